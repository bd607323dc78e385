#!/bin/bash
# One GPU call of the build->measure loop: tests, smoke, bench (each under its own timeout); logs under gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.used --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -q --timeout=600 > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py --steps 50 --warmup 10 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
tail -5 gpurun_out/pytest.log; tail -3 gpurun_out/smoke.log; tail -c 600 gpurun_out/bench.json; tail -5 gpurun_out/bench.err
# A/B: binning with the separate scan + emission kernels (round-1 form)
GS_BIN_SPLIT_SCAN=1 timeout 300 python bench.py --steps 30 --warmup 10 --no-extra-configs > gpurun_out/bench_splitscan.json 2> gpurun_out/bench_splitscan.err
python - <<'PY'
import json
for f in ("bench.json", "bench_splitscan.json"):
    try:
        d = json.loads(open("gpurun_out/" + f).read().strip().splitlines()[-1])
        print(f, d["value"], d["e2e"]["value"], {k: v for k, v in d.get("stages", {}).items()})
    except Exception as e:
        print(f, "unreadable", e)
PY
