"""Turns an `ncu --set full` report into the per-kernel JSON bench.py reads (profiles/rNN_ncu_metrics.json) and a
readable table (profiles/rNN_ncu_summary.md). Run HERE on the .ncu-rep that came back in gpurun_out/:
    python tools/ncu_extract.py gpurun_out/prof.ncu-rep|gpurun_out/prof_raw.csv profiles/r02
Per kernel NAME the LAST captured launch is kept (the warmest one), plus the number of launches seen."""
import csv
import io
import json
import re
import subprocess
import sys

WANT = {
    "gpu__time_duration.sum": "duration_ns",
    "dram__bytes_read.sum": "dram_read",
    "dram__bytes_write.sum": "dram_write",
    "launch__registers_per_thread": "registers",
    "launch__shared_mem_per_block_static": "smem_static",
    "launch__shared_mem_per_block_dynamic": "smem_dynamic",
    "launch__grid_size": "grid",
    "launch__block_size": "block",
    "launch__occupancy_limit_registers": "occ_limit_regs",
    "launch__occupancy_limit_shared_mem": "occ_limit_smem",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "achieved_occupancy_pct",
    "smsp__inst_executed.sum": "inst_executed",
    "sm__issue_active.avg.pct_of_peak_sustained_elapsed": "issue_active_pct",
    "sm__inst_issued.avg.pct_of_peak_sustained_active": "inst_issued_pct_of_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active": "pipe_lsu_pct",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active": "pipe_xu_pct",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active": "pipe_fma_pct",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active": "pipe_alu_pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_throughput_pct",
    "gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed": "mem_throughput_pct",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed": "dram_throughput_pct",
    "l1tex__t_sector_hit_rate.pct": "l1_hit_pct",
    "lts__t_sector_hit_rate.pct": "l2_hit_pct",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum": "smem_wavefronts",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum": "smem_bank_conflicts",
    "lts__t_bytes.sum": "l2_bytes",
    "smsp__cycles_active.avg": "cycles_active",
}
STALL = re.compile(r"smsp__average_warps?_issue_stalled_(\w+)_per_issue_active\.ratio|"
                   r"smsp__average_warp_latency_issue_stalled_(\w+)\.ratio")
UNIT_SCALE = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "usecond": 1e3, "msecond": 1e6, "nsecond": 1.0,
              "second": 1e9, "us": 1e3, "ms": 1e6, "ns": 1.0, "s": 1e9}


def num(s):
    try:
        return float(s.replace(",", ""))
    except ValueError:
        return None


def main():
    rep, out = sys.argv[1], sys.argv[2]
    if rep.endswith(".csv"):       # the raw page already exported on the GPU box (tools/ncu_round.sh)
        raw = open(rep).read()
    else:
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    raw = raw[raw.index('"ID"'):]
    rows = list(csv.reader(io.StringIO(raw)))
    head, units, body = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(head)}
    kname = col["Kernel Name"]
    kernels, order = {}, []
    for r in body:
        name = r[kname].split("(")[0].replace("<unnamed>::", "").replace("void ", "").strip()
        name = re.sub(r"^gsradix::", "", name)
        base_name = re.sub(r"<.*", "", name)
        name = base_name if base_name.startswith(("raster_", "preprocess_")) else name      # keep template args of the radix passes
        ent = {"launches_captured": kernels.get(name, {}).get("launches_captured", 0) + 1, "full_name": r[kname][:160]}
        for m, key in WANT.items():
            if m in col:
                v = num(r[col[m]])
                if v is not None:
                    ent[key] = v * UNIT_SCALE.get(units[col[m]], 1.0) if key in ("duration_ns", "dram_read", "dram_write", "l2_bytes") else v
        stalls = {}
        for h, i in col.items():
            mm = STALL.match(h)
            if mm and "not_issued" not in h:
                v = num(r[i])
                if v:
                    stalls[mm.group(1) or mm.group(2)] = max(stalls.get(mm.group(1) or mm.group(2), 0.0), v)
        ent["top_stalls_warps_per_issue"] = dict(sorted(stalls.items(), key=lambda kv: -kv[1])[:5])
        if "dram_read" in ent and "dram_write" in ent:
            ent["dram_bytes"] = ent["dram_read"] + ent["dram_write"]
        if name not in kernels:
            order.append(name)
        kernels[name] = ent
    sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
    from bench import csrc_sha16        # the capture is of the tree this script runs in: do not edit csrc/ in between
    kernels["_source"] = {"csrc_sha16": csrc_sha16(), "report": rep, "command": "ncu --set full --clock-control none --import-source on (tools/ncu_round.sh)",
                          "note": "last captured launch per kernel name; times under ncu are serialised and cold-cache, "
                                  "bench.py measures times live"}
    json.dump(kernels, open(out + "_ncu_metrics.json", "w"), indent=1)
    with open(out + "_ncu_summary.md", "w") as f:
        f.write("| kernel | us | regs | smem B | occ % | issue % | DRAM MB | L2 hit % | top stalls (warps/issue) |\n|---|---|---|---|---|---|---|---|---|\n")
        for name in order:
            e = kernels[name]
            st = ", ".join(f"{k} {v:.2f}" for k, v in e["top_stalls_warps_per_issue"].items())
            f.write(f"| {name} | {e.get('duration_ns', 0) / 1e3:.1f} | {e.get('registers', 0):.0f} | "
                    f"{e.get('smem_static', 0) + e.get('smem_dynamic', 0):.0f} | {e.get('achieved_occupancy_pct', 0):.0f} | "
                    f"{e.get('issue_active_pct', 0):.0f} | {e.get('dram_bytes', 0) / 1e6:.2f} | {e.get('l2_hit_pct', 0):.0f} | {st} |\n")
    print(open(out + "_ncu_summary.md").read())


if __name__ == "__main__":
    main()
