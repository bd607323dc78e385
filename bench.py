#!/usr/bin/env python
"""bench.py — views/sec, forward+backward, 1 M Gaussians, 512x512 (BASELINE.json `metric`).

One "step" = one training view through the host mirror of GaussCtrlModel.get_outputs
(gc_model.py:77-241, training branch) + loss.backward() down to the six leaf gradients
(gc_trainer.py:275): pre-activations, project_gaussians, spherical_harmonics (deg 3),
tile binning, forward compositing, backward compositing, SH backward, projection backward.
N > 1 (torchrun): parameters replicated, each rank takes a different view per step, one NCCL
allreduce of the parameter gradients per step; value = all ranks' views
/ max-over-ranks time.

    python bench.py [--gpus N --steps K --warmup W]            our CUDA path
    python bench.py --impl reference [...]                      the CPU oracle on the host cores

Prints ONE JSON line (rank 0).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_GAUSS = 1_000_000
SH_DEG = 3
N_CAMS = 40            # gc_datamanager.py:61-66: 4 subsets x 10 views
BG = (0.1, 0.2, 0.3)
METRIC = "views/sec fwd+bwd (1M Gaussians, 512^2)"


def workload_config(n_gauss, world):
    return {
        "workload": f"synthetic {n_gauss} Gaussians (seed 0, s0 per SURVEY 8d), SH deg {SH_DEG}, 512x512, "
                    f"{N_CAMS} cameras on a sphere r=1.5 (garden intrinsics), one view fwd+bwd per step per GPU, "
                    f"upstream grad randn/(H*W), bg {BG}",
        "n_gaussians": n_gauss, "image": [512, 512], "sh_degree": SH_DEG, "views_per_step": world,
        "parallelism": f"view-parallel dp{world}" + (" + grad exchange every step" if world > 1 else ""),
        "view_sampling": "next_train semantics (random permutation of the 40 cameras per epoch, seed 0), rank r takes "
                         "draw step * world + r (dist.ViewSampler)",
        "l2": "working set (236 MB params + 236 MB grads) exceeds the 126 MB L2; no explicit flush",
    }


# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) < 9:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def nvlink_kib(dev):
    """(tx KiB, rx KiB) moved over ALL NVLinks of `dev` since driver load: NVML field values
    NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX / _RX (138 / 139, scope = all links). None where NVML cannot say."""
    try:
        import pynvml
        import torch
        pynvml.nvmlInit()
        pr = torch.cuda.get_device_properties(dev)
        h = pynvml.nvmlDeviceGetHandleByPciBusId(f"{pr.pci_domain_id:08x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0".encode())
        vals = pynvml.nvmlDeviceGetFieldValues(h, [(138, 0xFFFFFFFF), (139, 0xFFFFFFFF)])
        out = []
        for v in vals:
            if v.nvmlReturn != 0:
                return None
            out.append(int(v.value.ullVal))
        return tuple(out)
    except Exception:
        return None


# ------------------------------------------------------------------------------------------------
class CpuReference:
    """The oracle (CPU restatement of the path, kind 'port') on the same workload, all host threads."""

    def __init__(self, n_gauss, threads=None):
        import numpy as np
        from gaussctrl_exp_b200 import scene
        from oracle import oracle as orc
        orc.build()
        # torchrun exports OMP_NUM_THREADS=1 to its workers: ask for every host core explicitly
        orc.set_num_threads(threads or os.cpu_count() or 1)
        self.orc = orc
        self.cores = orc.num_threads()
        self.p = {k: v.numpy() for k, v in scene.make_scene(n_gauss, 0).items()}
        self.cams = scene.make_cameras(N_CAMS, 0)
        from gaussctrl_exp_b200.dist import ViewSampler
        self.sampler = ViewSampler(N_CAMS, 0, 1, seed=0)      # the job's draw sequence: view k of the CPU arm = draw k
        self.v_out = (np.random.default_rng(1234).standard_normal((512, 512, 3)) / (512 * 512)).astype(np.float32)

    def view(self, i):
        out, ctx = self.orc.render_view(self.p, self.cams[self.sampler.draw(i)].as_dict(), SH_DEG, bg=BG, training=True)
        self.orc.render_view_backward(self.p, ctx, self.v_out)

    def run(self, views, first=0):
        t0 = time.perf_counter()
        for i in range(views):
            self.view(first + i)
        dt = time.perf_counter() - t0
        return views / dt, dt


def run_reference(args):
    """The reference arm: the CPU restatement of the path on ALL host cores, same config as our arm at this N
    (one step = `world` views, fwd + bwd each — the views the N ranks of our arm render in one step)."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    ref = CpuReference(args.n_gaussians)
    ref.run(min(max(args.warmup, 0), 2))
    # one step = `world` views; bounded sample: at most ~120 views in the whole run (~0.7 s each on 16 threads), i.e. a
    # step renders min(world, 120 // steps) views and its time is scaled to `world` views
    per_step = max(1, min(world, 120 // max(args.steps, 1)))
    _, dt = ref.run(args.steps * per_step, first=2)
    ms = 1e3 * dt / args.steps * (world / per_step)
    val = world * 1e3 / ms
    sample = (f"{per_step} of the {world} view(s) (fwd+bwd) of a step of the same {args.n_gaussians}-Gaussian/512^2 workload, "
              f"{args.steps} steps, time scaled to {world} view(s) per step; C oracle + OpenMP on {ref.cores} threads")
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "views/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": workload_config(args.n_gaussians, world),
        "cpu_baseline": {"value": val, "unit": "views/s", "cores": ref.cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "views/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "gsplat 0.1.2 (the reference's implementation of this path) is not in /root/reference and not "
                "installable offline; this arm times the CPU restatement (oracle/) on the host cores",
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
def timed_events(fn, n=20, warm=3):
    """Average device time of fn() in seconds: CUDA events on the current stream, sync on both sides."""
    import torch
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e-3


def stage_breakdown(model, cam, v_out, iters=20):
    """Per-stage device times (CUDA events on the current stream) for one view of `cam`, kernels called through the raw
    C ABI on device-resident inputs. Algorithmic bytes per SURVEY.md §8d (+ the 48-byte packed record where a kernel
    reads / writes it). Returns ({stage: (seconds, alg_bytes | None)}, intersections, visible Gaussians)."""
    import ctypes
    import torch
    from gaussctrl_exp_b200 import _lib, project_gaussians, render, spherical_harmonics
    from gaussctrl_exp_b200.model import camera_matrices
    dev = model.device
    H, W = cam.H, cam.W
    V, P, _, _, tb = camera_matrices(cam.c2w, cam.fx, cam.fy, W, H)
    Vd, PVd = V[:3].contiguous().to(dev), (P @ V).to(dev)
    N = model.num_points
    T = tb[0] * tb[1]
    L = _lib.lib()
    st = {}
    timed = lambda fn: timed_events(fn, iters)
    with torch.no_grad(), torch.cuda.device(dev):
        sp = _lib.stream_ptr()
        raw = [t.detach().float().contiguous() for t in (model.means, model.scales, model.quats, model.opacities,
                                                         model.features_dc, model.features_rest)]
        gc = render.gs_camera(cam)
        bg = torch.tensor(BG, device=dev)
        # the product forward once: workspace with records, tile lists, images
        ws, im, I, _ = render.forward_views(*raw, [gc], True, SH_DEG, bg)
        vis = int((ws.radii > 0).sum().item())
        rec, depths, radii, hit, ids_sorted, bins = ws.rec, ws.depths, ws.radii, ws.num_tiles_hit, ws.ids_sorted, ws.tile_bins
        cap = ws.cap
        # ---- fused stage 1: leaves -> packed records (+ depths, radii, tiles_hit) ----
        ws2 = render.ViewWorkspace(dev, N, 1, cap, T)
        scratch = torch.empty(L.gs_view_scratch_bytes(N, 1, cap, T), device=dev, dtype=torch.uint8)
        # gs_view_fwd = preprocess + binning + tile order + compositing; the pieces below are timed one by one
        o_xy, o_c, o_rgb = torch.empty(N, 2, device=dev), torch.empty(N, 3, device=dev), torch.empty(N, 3, device=dev)
        o_d, o_op = torch.empty(N, device=dev), torch.empty(N, device=dev)
        o_r, o_h = torch.empty(N, device=dev, dtype=torch.int32), torch.empty(N, device=dev, dtype=torch.int32)
        st["preprocess_fwd"] = (timed(lambda: L.gs_preprocess_fwd(
            N, raw[0].data_ptr(), raw[1].data_ptr(), raw[2].data_ptr(), raw[3].data_ptr(), raw[4].data_ptr(),
            raw[5].data_ptr(), 1, SH_DEG, ctypes.addressof(gc), o_xy.data_ptr(), o_d.data_ptr(), o_r.data_ptr(),
            o_c.data_ptr(), o_h.data_ptr(), o_rgb.data_ptr(), o_op.data_ptr(), sp)), 284 * N)
        # ---- binning through the raw ABI (no host read inside the timed loop) ----
        total = torch.zeros(1, device=dev, dtype=torch.int32)
        ids2, bins2 = torch.empty_like(ids_sorted), torch.empty_like(bins)
        st["bin_fast"] = (timed(lambda: L.gs_bin_fast(N, cap, o_xy.data_ptr(), o_d.data_ptr(), o_r.data_ptr(),
                                                       o_h.data_ptr(), tb[0], tb[1], ids2.data_ptr(), bins2.data_ptr(),
                                                       total.data_ptr(), scratch.data_ptr(), scratch.numel(), sp)),
                          8 * N + 20 * N + 12 * I + 24 * I + 8 * I + 8 * T)
        order_f = torch.empty(T, device=dev, dtype=torch.int32)
        order_b = torch.empty(T, device=dev, dtype=torch.int32)
        twork = torch.empty(T, device=dev, dtype=torch.int32)
        L.gs_tile_order(T, bins.data_ptr(), 1, order_f.data_ptr(), sp)
        st["tile_order"] = (timed(lambda: L.gs_tile_order(T, bins.data_ptr(), 1, order_f.data_ptr(), sp)), None)
        # ---- compositing over packed records (the product kernels) ----
        fwd = lambda: L.gs_raster_fwd_packed(1, H, W, tb[0], tb[1], ids_sorted.data_ptr(), bins.data_ptr(), rec.data_ptr(),
                                             bg.data_ptr(), im.out.data_ptr(), None, im.final_T.data_ptr(),
                                             im.final_idx.data_ptr(), order_f.data_ptr(), twork.data_ptr(),
                                             im.out_clamped.data_ptr(), im.alpha.data_ptr(), sp)
        st["raster_fwd"] = (timed(fwd), None)
        Np = (N + 3) // 4 * 4
        g = torch.zeros(9 * Np + 64, device=dev)
        gx, gc_, gr, go = g[0:2 * N], g[2 * Np:2 * Np + 3 * N], g[5 * Np:5 * Np + 3 * N], g[8 * Np:8 * Np + N]
        L.gs_tile_order(T, twork.data_ptr(), 0, order_b.data_ptr(), sp)
        bwd = lambda: L.gs_raster_bwd_packed(H, W, tb[0], tb[1], ids_sorted.data_ptr(), bins.data_ptr(), rec.data_ptr(),
                                             bg.data_ptr(), im.final_T.data_ptr(), im.final_idx.data_ptr(), v_out.data_ptr(),
                                             None, gx.data_ptr(), gc_.data_ptr(), gr.data_ptr(), go.data_ptr(),
                                             order_b.data_ptr(), None, sp)
        st["raster_bwd"] = (timed(bwd), None)
        # ---- the un-packed compositing kernels (round-1 design: register double buffer, five arrays) for A/B ----
        conics = ws.conics().contiguous()
        xys_c, rgbs_c, opac_c = ws.xys.contiguous(), ws.rgbs.contiguous(), ws.opac.contiguous()
        out_o = torch.empty((H, W, 3), device=dev)
        fT_o, fidx_o = torch.empty((H, W), device=dev), torch.empty((H, W), device=dev, dtype=torch.int32)
        st["raster_fwd_unpacked_r01"] = (timed(lambda: L.gs_raster_fwd_ordered(
            H, W, tb[0], tb[1], ids_sorted.data_ptr(), bins.data_ptr(), xys_c.data_ptr(), conics.data_ptr(),
            rgbs_c.data_ptr(), opac_c.data_ptr(), None, bg.data_ptr(), out_o.data_ptr(), None, fT_o.data_ptr(),
            fidx_o.data_ptr(), order_f.data_ptr(), twork.data_ptr(), sp)), None)
        g2 = [torch.zeros((N, k), device=dev) for k in (2, 3, 3, 1)]
        st["raster_bwd_unpacked_r01"] = (timed(lambda: L.gs_raster_bwd_ordered(
            H, W, tb[0], tb[1], ids_sorted.data_ptr(), bins.data_ptr(), xys_c.data_ptr(), conics.data_ptr(),
            rgbs_c.data_ptr(), opac_c.data_ptr(), bg.data_ptr(), fT_o.data_ptr(), fidx_o.data_ptr(), v_out.data_ptr(),
            None, g2[0].data_ptr(), g2[1].data_ptr(), g2[2].data_ptr(), g2[3].data_ptr(), order_b.data_ptr(), sp)), None)
        # ---- leaf backward ----
        gl = [torch.empty_like(t) for t in raw]
        st["preprocess_bwd"] = (timed(lambda: L.gs_preprocess_bwd(
            N, raw[0].data_ptr(), raw[1].data_ptr(), raw[2].data_ptr(), raw[3].data_ptr(), 1, SH_DEG,
            ctypes.addressof(gc), o_r.data_ptr(), o_rgb.data_ptr(), g2[0].data_ptr(), g2[1].data_ptr(), g2[2].data_ptr(),
            g2[3].data_ptr(), gl[0].data_ptr(), gl[1].data_ptr(), gl[2].data_ptr(), gl[3].data_ptr(), gl[4].data_ptr(),
            gl[5].data_ptr(), sp)), 332 * N)
        # ---- whole-view C entries (what the product calls): forward, backward ----
        pin = torch.zeros(1, dtype=torch.int32).pin_memory()
        im2 = render.ViewImages(dev, 1, H, W, False)
        view_fwd = lambda: L.gs_view_fwd(N, 1, raw[0].data_ptr(), raw[1].data_ptr(), raw[2].data_ptr(), raw[3].data_ptr(),
                                         raw[4].data_ptr(), raw[5].data_ptr(), None, 1, SH_DEG, ctypes.addressof(gc),
                                         bg.data_ptr(), cap, ws2.buf.data_ptr(), ws2.buf.numel(), scratch.data_ptr(),
                                         scratch.numel(), im2.out.data_ptr(), im2.out_clamped.data_ptr(),
                                         im2.alpha.data_ptr(), None, im2.final_T.data_ptr(), im2.final_idx.data_ptr(),
                                         None, None, sp)
        st["view_fwd_total"] = (timed(view_fwd), None)
        flat = torch.empty(59 * Np, device=dev)
        view_bwd = lambda: L.gs_view_bwd(N, raw[0].data_ptr(), raw[1].data_ptr(), raw[2].data_ptr(), raw[3].data_ptr(), 1,
                                         SH_DEG, ctypes.addressof(gc), bg.data_ptr(), ws2.buf.data_ptr(), ws2.buf.numel(),
                                         cap, im2.final_T.data_ptr(), im2.final_idx.data_ptr(), v_out.data_ptr(), None,
                                         im2.out.data_ptr(), g.data_ptr(), flat.data_ptr(), sp)
        st["view_bwd_total"] = (timed(view_bwd), None)
        st["preprocess_fwd_packed"] = (timed(lambda: L.gs_preprocess_fwd_packed(
            N, raw[0].data_ptr(), raw[1].data_ptr(), raw[2].data_ptr(), raw[3].data_ptr(), raw[4].data_ptr(),
            raw[5].data_ptr(), None, 1, SH_DEG, ctypes.addressof(gc), ws2.rec.data_ptr(), ws2.depths.data_ptr(),
            ws2.radii.data_ptr(), ws2.num_tiles_hit.data_ptr(), sp)), 296 * N)     # 236 in + 48 record + 12 ints/depth
        # ---- literal operators (what the drop-in API runs): projection, SH, their backward, packing ----
        means, scales = raw[0], raw[1].exp()
        quats = raw[2] / raw[2].norm(dim=-1, keepdim=True)
        coeffs = torch.cat((raw[4][:, None, :], raw[5]), 1).contiguous()
        dirs = means - cam.c2w[:3, 3].to(dev)
        dirs = dirs / dirs.norm(dim=-1, keepdim=True)
        st["project_fwd"] = (timed(lambda: project_gaussians(means, scales, 1, quats, Vd, PVd, cam.fx, cam.fy, cam.cx,
                                                             cam.cy, H, W, tb)), 96 * N)
        st["sh_fwd"] = (timed(lambda: spherical_harmonics(SH_DEG, dirs, coeffs)), 216 * N)
        vcoef = torch.empty((N, 16, 3), device=dev)
        st["sh_bwd"] = (timed(lambda: L.gs_sh_bwd(N, 16, SH_DEG, dirs.data_ptr(), g2[2].data_ptr(), vcoef.data_ptr(), sp)),
                        216 * N)
        vm, vs, vq = torch.empty_like(means), torch.empty_like(means), torch.empty_like(quats)
        vdep = torch.zeros(N, device=dev)
        st["project_bwd"] = (timed(lambda: L.gs_project_bwd(N, means.data_ptr(), scales.data_ptr(), 1.0, quats.data_ptr(),
                                                            Vd.data_ptr(), PVd.data_ptr(), cam.fx, cam.fy, H, W,
                                                            o_r.data_ptr(), g2[0].data_ptr(), vdep.data_ptr(),
                                                            g2[1].data_ptr(), vm.data_ptr(), vs.data_ptr(), vq.data_ptr(),
                                                            sp)), 144 * N)
        rec2 = torch.empty((N, 12), device=dev)
        st["pack_splats"] = (timed(lambda: L.gs_pack_splats(N, xys_c.data_ptr(), conics.data_ptr(), rgbs_c.data_ptr(),
                                                            opac_c.data_ptr(), None, rec2.data_ptr(), sp)), 84 * N)
    return st, I, vis


def csrc_sha16():
    """sha256 (first 16 hex digits) over the CUDA sources of the library, in name order: what an ncu capture is OF.
    tools/ncu_extract.py stores it in the metrics file; bench.py compares it with the tree it runs from."""
    import hashlib
    d = os.path.join(ROOT, "gaussctrl_exp_b200", "csrc")
    h = hashlib.sha256()
    for name in sorted(os.listdir(d)):
        if name.endswith((".cu", ".cuh")) or name == "Makefile":
            h.update(name.encode())
            h.update(open(os.path.join(d, name), "rb").read())
    return h.hexdigest()[:16]


def load_ncu_profile():
    """profiles/rNN_ncu_metrics.json (newest round): per-kernel metrics extracted (tools/ncu_extract.py) from the
    `ncu --set full` capture committed next to it. bench.py measures times live; instruction counts, issue utilisation
    and DRAM traffic can only come from a profiler run, so they are read from that file — and flagged as stale when
    the CUDA sources have changed since the capture (the file records the source hash it was taken from)."""
    prof_dir = os.path.join(ROOT, "profiles")
    try:
        names = sorted(n for n in os.listdir(prof_dir) if n.endswith("_ncu_metrics.json"))
        path = os.path.join(prof_dir, names[-1])
        prof = json.load(open(path))
    except Exception:
        return {}, None, None
    captured = (prof.get("_source") or {}).get("csrc_sha16")
    return prof, os.path.relpath(path, ROOT), {"captured_from_csrc_sha16": captured, "this_tree_csrc_sha16": csrc_sha16(),
                                               "current": captured == csrc_sha16()}


def time_model_steps(model, cams, v_out, steps, warmup, rank=0, world=1):
    import torch
    from gaussctrl_exp_b200.dist import ViewSampler
    params = [model.means, model.scales, model.quats, model.features_dc, model.features_rest, model.opacities]
    sampler = ViewSampler(len(cams), rank, world, seed=0)

    def once(i):
        for p in params:
            p.grad = None
        model(cams[sampler.view(i)])["rgb"].backward(v_out)

    def step(i):
        from gaussctrl_exp_b200 import render as grender
        try:
            once(i)
        except grender.CapacityOverflow:        # model.speculate only: redo the view whose tile lists overflowed
            once(i)
    for i in range(warmup):
        step(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        step(warmup + i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def config1_api_500k(dev, steps=20):
    """BASELINE.json configs[1]: 500 K Gaussians, SH deg 3, 512x512, fwd+bwd on one GPU THROUGH THE
    gsplat-0.1.2-COMPATIBLE API (the literal operator sequence of gc_model.py:172-222: torch pre-ops +
    project_gaussians + spherical_harmonics + rasterize_gaussians); the fused path on the same scene beside it."""
    import torch
    from gaussctrl_exp_b200 import scene
    from gaussctrl_exp_b200.model import GaussianModel
    p = scene.make_scene(500_000, 0)
    cams = scene.make_cameras(N_CAMS, 0)
    g = torch.Generator().manual_seed(1234)
    v = (torch.randn(512, 512, 3, generator=g) / (512 * 512)).to(dev)
    out = {"workload": "synthetic 500000 Gaussians, SH deg 3, 512x512, one view fwd+bwd per step, 1 GPU"}
    for name, fused in (("api", False), ("fused", True)):
        m = GaussianModel(p, sh_degree=SH_DEG, background_color=torch.tensor(BG), fused=fused).to(dev)
        m.train()
        ms = time_model_steps(m, cams, v, steps, 5)
        out[name + "_views_per_s"] = 1e3 / ms
        out[name + "_ms_per_step"] = ms
        del m
    torch.cuda.empty_cache()
    return out


def scene_large_splats(dev, v_out, hbm_peak):
    """Second workload (scene.LARGE_SPLATS): large anisotropic splats, tens of tiles each — where the per-warp 8x4 box
    cull and the direct red.global backward of the compositing kernels stop being cheap. 512^2 training view."""
    import torch
    from gaussctrl_exp_b200 import scene
    from gaussctrl_exp_b200.model import GaussianModel
    spec = scene.LARGE_SPLATS
    cams = scene.make_cameras(N_CAMS, 0)
    m = GaussianModel(scene.make_scene(spec["n"], 0, s0=spec["s0"], log_sigma=spec["log_sigma"]), sh_degree=SH_DEG,
                      background_color=torch.tensor(BG)).to(dev)
    m.train()
    ms = time_model_steps(m, cams, v_out, 30, 8)
    st, I, vis = stage_breakdown(m, cams[0], v_out, iters=5)
    out = {"workload": f"synthetic {spec['n']} Gaussians, s0 {spec['s0']}, log-scale sigma {spec['log_sigma']}, SH deg 3, 512x512",
           "views_per_s": 1e3 / ms, "ms_per_step": ms, "intersections_cam0": int(I), "visible_cam0": vis,
           "intersections_per_gaussian": I / spec["n"], "intersections_per_visible": I / max(vis, 1),
           "stages_us": {k: st[k][0] * 1e6 for k in ("preprocess_fwd_packed", "bin_fast", "raster_fwd", "raster_bwd", "preprocess_bwd") if k in st},
           "bin_fast_Gintersections_per_s": I / st["bin_fast"][0] / 1e9,
           "raster_fwd_Gintersections_per_s": I / st["raster_fwd"][0] / 1e9,
           "raster_bwd_Gintersections_per_s": I / st["raster_bwd"][0] / 1e9,
           "bin_fast_frac_hbm": st["bin_fast"][1] / st["bin_fast"][0] / 1e9 / hbm_peak if st["bin_fast"][1] else None}
    del m
    torch.cuda.empty_cache()
    return out


def config5_stress(dev, n_gauss=5_000_000, n_cams=64, batch=8):
    """BASELINE.json configs[4]: 5 M Gaussians at 1920x1080, a 64-camera batch: eval render (batched
    render_views), training step, and binning / compositing throughput in intersections/s."""
    import torch
    from gaussctrl_exp_b200 import scene
    from gaussctrl_exp_b200.model import GaussianModel
    cams = scene.make_cameras(n_cams, 0, intrinsics=scene.HD_INTRINSICS)
    H, W = cams[0].H, cams[0].W
    v = (torch.randn(H, W, 3, generator=torch.Generator().manual_seed(1234)) / (H * W)).to(dev)
    m = GaussianModel(scene.make_scene(n_gauss, 0), sh_degree=SH_DEG, background_color=torch.tensor(BG)).to(dev)
    m.eval()
    m.render_views(cams, batch=batch)                            # warm-up pass over all cameras: capacity hints, allocator
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    m.render_views(cams, batch=batch)
    e1.record()
    torch.cuda.synchronize()
    ms_eval = e0.elapsed_time(e1) / n_cams
    m.train()
    ms_train = time_model_steps(m, cams, v, 16, 3)
    st, I, vis = stage_breakdown(m, cams[0], v, iters=5)
    out = {"workload": f"stress: {n_gauss} Gaussians, {W}x{H}, {n_cams} cameras (batches of {batch}), SH deg 3",
           "intersections_cam0": int(I), "visible_cam0": vis, "eval_render_views_per_s": 1e3 / ms_eval,
           "train_fwd_bwd_views_per_s": 1e3 / ms_train,
           "bin_fast_us": st["bin_fast"][0] * 1e6, "bin_fast_Gintersections_per_s": I / st["bin_fast"][0] / 1e9,
           "raster_fwd_us": st["raster_fwd"][0] * 1e6, "raster_fwd_Gintersections_per_s": I / st["raster_fwd"][0] / 1e9,
           "raster_bwd_us": st["raster_bwd"][0] * 1e6, "peak_mem_GB": torch.cuda.max_memory_allocated() / 2**30}
    del m
    torch.cuda.empty_cache()
    # configs[4] as worded: binning/sort and compositing throughput VS GAUSSIAN COUNT (one 1080p training view, camera 0)
    point = lambda n, I_, vis_, st_, ms_: {
        "n_gaussians": n, "intersections": int(I_), "visible": vis_, "train_fwd_bwd_views_per_s": 1e3 / ms_,
        "bin_fast_us": st_["bin_fast"][0] * 1e6, "bin_fast_Gintersections_per_s": I_ / st_["bin_fast"][0] / 1e9,
        "raster_fwd_us": st_["raster_fwd"][0] * 1e6, "raster_fwd_Gintersections_per_s": I_ / st_["raster_fwd"][0] / 1e9,
        "raster_bwd_us": st_["raster_bwd"][0] * 1e6, "raster_bwd_Gintersections_per_s": I_ / st_["raster_bwd"][0] / 1e9}
    sweep = []
    for n in (1_000_000, 2_000_000):
        if n >= n_gauss:
            continue
        ms_ = GaussianModel(scene.make_scene(n, 0), sh_degree=SH_DEG, background_color=torch.tensor(BG)).to(dev)
        ms_.train()
        t_ = time_model_steps(ms_, cams, v, 16, 3)
        st_, I_, vis_ = stage_breakdown(ms_, cams[0], v, iters=5)
        sweep.append(point(n, I_, vis_, st_, t_))
        del ms_
        torch.cuda.empty_cache()
    sweep.append(point(n_gauss, I, vis, st, ms_train))
    out["sweep_vs_gaussian_count"] = sweep
    return out


def run_ours(args):
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this path has no CPU fallback (use --impl reference for the CPU arm)")
    from gaussctrl_exp_b200 import _lib, scene
    from gaussctrl_exp_b200 import dist as gdist
    from gaussctrl_exp_b200.model import GaussianModel
    import torch.distributed as dist
    _lib.lib()
    rank, local, world = gdist.init_from_env("nccl")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    H = W = 512
    model = GaussianModel(scene.make_scene(args.n_gaussians, 0), sh_degree=SH_DEG,
                          background_color=torch.tensor(BG), fused=(args.path == "fused")).to(dev)
    model.train()
    params = [model.means, model.scales, model.quats, model.features_dc, model.features_rest, model.opacities]
    cams = scene.make_cameras(N_CAMS, 0)
    # training views: the reference's next_train sampling (a random permutation of the cameras per epoch), the same on
    # every rank; rank r takes draw i * world + r of step i (gc_datamanager.py:213-235)
    sampler_v = gdist.ViewSampler(N_CAMS, rank, world, seed=0)
    g = torch.Generator().manual_seed(1234)
    v_host = (torch.randn(H, W, 3, generator=g) / (H * W)).pin_memory()
    v_out = v_host.to(dev)
    tgt_host = torch.rand(H, W, 3, generator=g).pin_memory()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # the 40 training cameras are known up front (gc_datamanager.py:61-66): their host-side matrices (gc_model.py:117-155)
    # are computed once and cached by value, as a fine-tune loop that cycles over the same views would find them
    from gaussctrl_exp_b200 import render as grender
    for c in cams:
        grender.gs_camera(c)
        model._caller_matrices(c)

    gather_mode = world > 1 and model.fused and args.reduce in ("gather", "peer")
    model.view_state.defer_leaf_grads = gather_mode
    peer = None
    if gather_mode and args.reduce == "peer":
        try:
            peer = gdist.PeerViewExchange(model, mode=args.exchange)   # NVLink symmetric memory
        except Exception as e:                         # (e.g. no P2P between the visible devices)
            if rank == 0:
                print(f"bench.py: peer-memory exchange unavailable ({type(e).__name__}: {e}); using NCCL all-gather",
                      file=sys.stderr)
            args.reduce = "gather"

    def reduce_grads():
        if world == 1:
            return
        if peer is not None:
            peer.reduce(model)                     # peers' view blocks read in place over NVLink inside the kernel
            return
        if gather_mode:
            gdist.reduce_view_grads(model)        # all-gather 36 B/Gaussian, replay the leaf backward per view
            return
        flat = model.view_state.flat_grads if model.fused else None
        if flat is not None:
            gdist.allreduce_flat(flat)
        else:
            gdist.allreduce_grads(params)

    steps_redone = [0]

    def step(i):
        # model.speculate (the timed loops below set it unless --forward strict): the forward reads nothing back, so
        # the host is never waiting for the binning when the backward has to be enqueued; a view whose tile lists
        # overflowed their capacity raises at the end of ITS backward, before any gradient is handed out, and the
        # step is redone — what a training loop on this mode has to do
        try:
            _step(i)
        except grender.CapacityOverflow:
            steps_redone[0] += 1
            _step(i)

    def _step(i):
        cam = cams[sampler_v.view(i)]
        for p in params:
            p.grad = None
        out = model(cam)
        # "a fixed upstream gradient, then a8 backward down to the leaf grads" — v_out IS the upstream
        # gradient of the rendered image, handed to autograd directly (no torch loss kernels inside the step; the
        # e2e loop below has a real L1 loss against a target image copied from the host)
        out["rgb"].backward(v_out)
        reduce_grads()

    # ---- N > 1: the exchange that the timed loop uses against a plain NCCL all-reduce of the leaf gradients ----
    grad_check = None
    allreduce_ms = None
    if world > 1 and model.fused:
        def grads_allreduce(i):
            model.view_state.defer_leaf_grads = False
            model.view_state.view_block_buffer = None
            for p in params:
                p.grad = None
            model(cams[sampler_v.view(i)])["rgb"].backward(v_out)
            gdist.allreduce_flat(model.view_state.flat_grads)
            flat = model.view_state.flat_grads
            model.view_state.defer_leaf_grads = gather_mode
            if peer is not None:
                peer.begin(model)
            return flat
        step(0)
        torch.cuda.synchronize()
        sliced = peer is not None and peer.mode == "slice"
        if sliced:
            # owner-sliced exchange: every rank holds the reduced gradients of ITS Gaussians [lo, hi) only
            lo, hi = model.view_state.shard
            mine = {k: v.clone() for k, v in model.view_state.sharded_grads.items()}
        else:
            a = model.view_state.flat_grads.clone()
        b = grads_allreduce(0)
        torch.cuda.synchronize()
        den = b.abs().max().clamp_min(1e-30)
        if sliced:
            full = dict(zip(gdist.GRAD_PARAMS, (lambda t: (t[0], t[1], t[2], t[4], t[5], t[3]))(
                grender.split_flat_grads(b, args.n_gaussians, tuple(model.opacities.shape)))))
            diff = torch.stack([(mine[k] - full[k][lo:hi]).abs().max() for k in gdist.GRAD_PARAMS]).max().div(den).reshape(1)
            covered = torch.tensor([hi - lo], device=dev)
            dist.all_reduce(covered)
            identical = bool(int(covered.item()) == args.n_gaussians)       # the slices partition the Gaussians
        else:
            diff = ((a - b).abs().max() / den).reshape(1)
            # bit pattern checksum of this rank's reduced gradients: identical on every rank?
            chk = a.view(torch.int32).to(torch.int64).sum().reshape(1)
            chks = [torch.zeros_like(chk) for _ in range(world)]
            dist.all_gather(chks, chk)
            identical = bool(all(int(c.item()) == int(chks[0].item()) for c in chks))
        dist.all_reduce(diff, op=dist.ReduceOp.MAX)
        grad_check = {"exchange": args.reduce + ("/" + args.exchange if args.reduce == "peer" else ""),
                      "against": "NCCL sum-allreduce of the 59-float leaf gradients, averaged" +
                                 (" (each rank: its owner slice)" if sliced else ""),
                      "max_rel_diff": float(diff.item()),
                      ("slices_partition_the_gaussians" if sliced else "ranks_identical"): identical,
                      "nonzero": bool(b.abs().max().item() > 0), "step": 0}
        # config 4 as BASELINE.json words it: fwd+bwd + NCCL allreduce of the parameter gradients
        for i in range(3):
            grads_allreduce(i)
        barrier()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record()
        for i in range(10):
            grads_allreduce(3 + i)
        a1.record()
        barrier()
        t = torch.tensor([a0.elapsed_time(a1) / 10], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        allreduce_ms = float(t.item())

    loss_pinned = [torch.zeros((), dtype=torch.float32).pin_memory() for _ in range(2)]
    loss_events = [torch.cuda.Event() for _ in range(2)]
    loss_pending = [False, False]
    losses = []

    # e2e inputs: the step's target image comes from pinned host memory EVERY step. The copy of step i+1's target is
    # issued on a copy stream while step i computes (double-buffered, the way a data loader prefetches), so the 3 MB
    # PCIe transfer (~57 us) does not sit on the compute stream's critical path; each step still waits for ITS copy.
    copy_stream = torch.cuda.Stream(device=dev)
    tgt_dev = [torch.empty(H, W, 3, device=dev) for _ in range(2)]
    tgt_ready = [torch.cuda.Event() for _ in range(2)]
    tgt_free = [torch.cuda.Event() for _ in range(2)]
    tgt_issued = [-1, -1]

    def prefetch_target(i):
        slot = i % 2
        if tgt_issued[slot] == i:
            return
        copy_stream.wait_event(tgt_free[slot])           # the step that last read this buffer has consumed it
        with torch.cuda.stream(copy_stream):
            tgt_dev[slot].copy_(tgt_host, non_blocking=True)
            tgt_ready[slot].record(copy_stream)
        tgt_issued[slot] = i

    e2e_retries = [0]

    def step_e2e(i):
        # the end-to-end loop runs the forward fully asynchronously (model.speculate: the host never waits for the
        # device inside a step) and does what a training loop has to do then: redo a step whose tile lists overflowed
        try:
            _step_e2e(i)
        except grender.CapacityOverflow:
            e2e_retries[0] += 1
            _step_e2e(i)

    def _step_e2e(i):
        cam = cams[sampler_v.view(i)]
        slot = i % 2
        prefetch_target(i)                                  # no-op except on the very first step
        cur = torch.cuda.current_stream(dev)
        tgt = tgt_dev[slot]
        prefetch_target(i + 1)                              # H2D of the NEXT step's target overlaps this step
        for p in params:
            p.grad = None
        out = model(cam)
        # the step waits for ITS target where it first needs it — at the loss, not in front of the render (a wait node
        # ahead of the preprocess kernel cost 23 us per step: tools/e2e_anatomy.py) — and hands the buffer back as soon
        # as the subtraction has read it
        cur.wait_event(tgt_ready[slot])
        loss = (out["rgb"] - tgt).abs().mean()
        tgt_free[slot].record(cur)
        loss.backward()
        reduce_grads()
        # D2H of the step's result: async copy into pinned memory, consumed one step later so the host
        # never idles the GPU (every step's loss is still read inside the timed region; see drain_losses)
        loss_pinned[slot].copy_(loss.detach(), non_blocking=True)
        loss_events[slot].record()
        prev = 1 - slot
        if loss_pending[prev]:
            loss_events[prev].synchronize()
            losses.append(float(loss_pinned[prev]))
        loss_pending[slot] = True

    def drain_losses():
        for slot in (0, 1):
            if loss_pending[slot]:
                loss_events[slot].synchronize()
                losses.append(float(loss_pinned[slot]))
                loss_pending[slot] = False

    host_ms = [0.0]

    def timed_loop(fn, warmup, steps, drain=None):
        for i in range(warmup):
            fn(i)
        if drain:
            drain()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = _lib.launch_count()
        e0.record()
        h0 = time.perf_counter()
        for i in range(steps):
            fn(warmup + i)
        host_ms[0] = 1e3 * (time.perf_counter() - h0) / steps     # host time per step (diagnostic)
        if drain:
            drain()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        t = torch.tensor([ms], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), _lib.launch_count() - l0

    # the same loop with the default, self-healing forward (capacity confirmed inside the forward: one host wait that
    # ends when the binning is done); diagnostic, shorter
    model.speculate = False
    ms_strict, _ = timed_loop(step, max(3, args.warmup // 2), max(10, args.steps // 2))
    ms_strict /= max(10, args.steps // 2)
    model.speculate = args.forward == "async"
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    nv0 = nvlink_kib(dev) if world > 1 else None
    ms_total, launches = timed_loop(step, args.warmup, args.steps)
    nv1 = nvlink_kib(dev) if world > 1 else None
    nvlink = None
    if world > 1 and peer is not None and getattr(peer, "mode", "") == "slice":
        # what the last timed step actually stored into the peers (counted from the pushed block). The hardware
        # counters are read as well, but this pool does not expose them (NVML NOT_SUPPORTED, `nvidia-smi nvlink -gt d`:
        # N/A; profiles/r02_nvlink_counters_unavailable.txt)
        sent, nnz = peer.pushed_bytes_last_step(model.num_points)
        nvlink = {"pushed_bytes_per_step_rank0": sent, "nonzero_gaussians_in_block": nnz,
                  "dense_allreduce_bytes_per_rank": 2 * (world - 1) / world * 59 * 4 * model.num_points,
                  "hw_counters": None,
                  "source": "counted from the compacted view block the last timed step pushed (run arithmetic of "
                            "vx_push_slices_kernel); dense figure = ring all-reduce of the 59-float leaf gradients"}
    if nv0 and nv1:        # rank 0's GPU; includes the warm-up steps of the same loop
        n_st = args.steps + args.warmup
        nvlink = dict(nvlink or {})
        nvlink["hw_counters"] = {"tx_bytes_per_step": 1024.0 * (nv1[0] - nv0[0]) / n_st,
                                 "rx_bytes_per_step": 1024.0 * (nv1[1] - nv0[1]) / n_st,
                                 "source": "NVML NVLINK_THROUGHPUT_DATA_TX/RX (all links of rank 0's GPU, KiB) around "
                                           "warm-up + timed steps; includes barrier flags and NCCL's timing all-reduce"}
    host_enqueue_ms = host_ms[0]
    clocks = sampler.stop() if rank == 0 else None

    # per-step distribution (median and p10/p90): the same steps, each bracketed by its own events
    # (events are read back afterwards). Diagnostic only: `value` is the loop above.
    n_dist = min(args.steps, 100)
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(n_dist + 1)]
    barrier()
    evs[0].record()
    for i in range(n_dist):
        step(args.warmup + i)
        evs[i + 1].record()
    barrier()
    per_step = sorted(evs[i].elapsed_time(evs[i + 1]) for i in range(n_dist))
    pct = lambda q: per_step[min(n_dist - 1, int(q * n_dist))]
    step_dist = {"p10_ms": pct(0.10), "median_ms": pct(0.50), "p90_ms": pct(0.90), "steps": n_dist,
                 "what": "per-step device time of this rank, CUDA events around each step"}
    ms_e2e, _ = timed_loop(step_e2e, max(3, args.warmup // 2), args.steps, drain=drain_losses)

    # config 4 with the optimiser step the trainer runs right after backward (gc_trainer.py:276-298; Adam groups of
    # gc_config.py:58-87): fwd + bwd + gradient exchange + Adam. Owner-sliced exchange: Adam over the rank's slice
    # (1/world of the state and of the update) + all-gather of the updated rows; otherwise a replicated fused Adam.
    # Parameters are restored afterwards so that the measurements below see the same scene.
    def finetune_with_optimizer(n_steps=10):
        lr = dict(means=1.6e-4, scales=5e-3, quats=1e-3, features_dc=2.5e-3, features_rest=1.25e-4, opacities=5e-2)
        named = {k: getattr(model, k) for k in gdist.GRAD_PARAMS}
        saved = {k: p.detach().clone() for k, p in named.items()}
        sliced = peer is not None and peer.mode == "slice"
        if sliced or world == 1:
            # world == 1: the slice is the whole scene (same fused Adam kernel, nothing to exchange)
            opt = gdist.ShardedAdam(named, rank, world, lr=lr, eps=1e-15, peer=True)

            def fstep(i):
                step(i)
                g = model.view_state.sharded_grads if world > 1 else {k: named[k].grad for k in named}
                opt.step(g)
                opt.sync()
        else:
            opt = torch.optim.Adam([dict(params=[named[k]], lr=lr[k]) for k in named], eps=1e-15, fused=True)

            def fstep(i):
                step(i)
                opt.step()
        for i in range(3):
            fstep(i)
        barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        for i in range(n_steps):
            fstep(3 + i)
        f1.record()
        barrier()
        t = torch.tensor([f0.elapsed_time(f1) / n_steps], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        with torch.no_grad():
            for k, p in named.items():
                p.copy_(saved[k])
                p.grad = None
        del opt, saved
        torch.cuda.empty_cache()
        return {"ms_per_step": float(t.item()), "views_per_s": world * 1e3 / float(t.item()), "n_gpus": world,
                "optimizer": ("gs_adam_slice over owner slices (1/%d of the rows per rank), updated rows stored into every peer's "
                              "parameters by the same kernel (NVLink symmetric memory) + device barrier" % world)
                if (sliced or world == 1) else "replicated torch.optim.Adam(fused=True)",
                "what": "fwd + bwd + gradient exchange + Adam step (gc_trainer.py:276-298, lr groups of gc_config.py:58-87)"}
    finetune_opt = finetune_with_optimizer() if model.fused else None
    model.speculate = False

    # config 3: all-view eval render, RGB + depth + alpha, 40 views sharded over the ranks, batched
    def render_all():
        return model.render_views(cams, rank, world, batch=args.render_batch)
    for _ in range(2):
        render_all()
    barrier()
    r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    r0.record()
    for _ in range(args.render_reps):
        render_all()
    r1.record()
    barrier()
    t_r = torch.tensor([r0.elapsed_time(r1)], device=dev)
    if world > 1:
        dist.all_reduce(t_r, op=dist.ReduceOp.MAX)
    render_vps = N_CAMS * args.render_reps / (float(t_r.item()) * 1e-3)

    line = None
    if rank == 0:
        ms_step = ms_total / args.steps
        value = world * 1e3 / ms_step
        line = {
            "metric": METRIC, "value": value, "unit": "views/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(args.n_gaussians, world),
            "clocks": clocks, "gpu_launches": int(launches), "path": args.path,
            "grad_exchange": ((args.reduce + ("/" + args.exchange if args.reduce == "peer" else "")) if world > 1 else None),
            "grad_exchange_check": grad_check,
            "nvlink": nvlink,
            "forward": ("asynchronous (model.speculate=True): nothing read back in the forward, an overflowed view raises "
                        "in its backward and the step is redone" if args.forward == "async" else
                        "strict: capacity confirmed inside the forward (host waits for the binning)"),
            "steps_redone": steps_redone[0],
            "strict_forward": {"views_per_s": world * 1e3 / ms_strict, "ms_per_step": ms_strict,
                               "what": "the same loop with model.speculate=False (self-healing forward, the model's default)"},
            "host_ms_per_step": host_enqueue_ms, "step_time_distribution": step_dist,
            "e2e": {"value": world * 1e3 / (ms_e2e / args.steps), "unit": "views/s",
                    "h2d_bytes_per_step": int(tgt_host.numel() * 4 + 12 * 4 + 6 * 4), "d2h_bytes_per_step": 4 + 4,
                    "forward": ("asynchronous (model.speculate=True), overflowed steps redone" if args.forward == "async"
                                else "strict"), "steps_redone": e2e_retries[0]},
            # BASELINE.json configs, one named key each (configs[0] is the CPU case = cpu_baseline / --impl reference)
            "config3_all_view_render": {
                "views_per_s": render_vps, "n_gpus": world,
                "what": f"all-view RGB+depth+alpha render of {N_CAMS} cameras, 1M Gaussians, 512^2 (gc_pipeline.py:222-228), "
                        f"view-sharded over {world} GPU(s), batches of {args.render_batch} cameras per call"},
        }
        line["config"]["forward_mode_of_value"] = "asynchronous (model.speculate=True)" if args.forward == "async" else "strict"
        line["config"]["strict_forward_views_per_s"] = world * 1e3 / ms_strict
        if grad_check is not None:
            line["config"]["grad_exchange_check_max_rel_diff"] = grad_check["max_rel_diff"]
            line["config"]["grad_exchange_check_ok"] = bool(
                grad_check["max_rel_diff"] < 1e-5 and grad_check["nonzero"]
                and grad_check.get("slices_partition_the_gaussians", grad_check.get("ranks_identical")))
        if finetune_opt is not None:
            line["config4_finetune_step_with_optimizer"] = finetune_opt
            line["config"]["finetune_step_with_optimizer_views_per_s"] = finetune_opt["views_per_s"]
            line["e2e"]["finetune_step_with_optimizer_views_per_s"] = finetune_opt["views_per_s"]
        if allreduce_ms is not None:
            line["config4_finetune_step_nccl_allreduce"] = {
                "views_per_s": world * 1e3 / allreduce_ms, "ms_per_step": allreduce_ms, "n_gpus": world,
                "what": "fwd+bwd + NCCL sum-allreduce of the 59-float leaf gradients (236 MB) every step — BASELINE.json "
                        "configs[3] as worded; `value` uses the compacted peer-memory exchange instead"}
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak, peak_src = (peaks.get("hbm_gbs"), "measured") if peaks.get("hbm_gbs") else (6650.0, "fallback")
        prof, prof_path, prof_state = load_ncu_profile()
        if not args.no_stages:
            st, I, vis = stage_breakdown(model, cams[0], v_out)
            stages = {}
            for k, (sec, byts) in st.items():
                stages[k] = {"us": sec * 1e6}
                if byts:
                    stages[k].update({"alg_bytes": int(byts), "GBps": byts / sec / 1e9, "frac_hbm": byts / sec / 1e9 / hbm_peak})
            line["stages"] = stages
            line["intersections"] = I
            line["visible"] = vis
            use_prof = args.n_gaussians == N_GAUSS          # the capture is of the headline scene
            f_clk = (clocks.get("sm_mhz") or 1965.0) * 1e6
            slots = 148 * 4 * f_clk                       # warp-instruction issue slots per second (4 schedulers per SM)
            step_us = ms_strict * 1e3

            def hbm_entry(stage, kname):
                pk = prof.get(kname, {}) if use_prof else {}
                e = stages[stage]
                return {"bound": "hbm", "kernel": kname, "us": e["us"], "share_of_step": e["us"] / step_us,
                        "achieved": e["GBps"], "peak": hbm_peak, "unit": "GB/s", "frac": e["frac_hbm"],
                        "alg_bytes_per_launch": e["alg_bytes"], "traffic": pk.get("dram_bytes")}

            def issue_entry(stage, kname):
                pk = prof.get(kname) if use_prof else None
                e = {"bound": "sm_issue", "kernel": kname, "us": stages[stage]["us"],
                     "share_of_step": stages[stage]["us"] / step_us, "peak": slots, "unit": "warp-inst/s"}
                if pk and pk.get("inst_executed"):
                    e.update({"achieved": pk["inst_executed"] / (e["us"] * 1e-6),
                              "frac": pk["inst_executed"] / (e["us"] * 1e-6) / slots,
                              "smsp__inst_executed.sum": pk["inst_executed"],
                              "sm__issue_active.avg.pct_of_peak_sustained_elapsed": pk.get("issue_active_pct"),
                              "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum": pk.get("smem_wavefronts"),
                              "traffic": pk.get("dram_bytes")})
                return e
            # `roofline` = the DOMINANT kernel of the step. It has no dense contraction and moves 77 MB: it is bound by
            # the SMs' instruction issue rate, so achieved / peak are warp-instructions per second (time measured live
            # with CUDA events, instruction count from the committed ncu capture of the same sources). The HBM-bound
            # kernels and the multi-kernel binning stage are nested under `stages` so that the record keeps them.
            dom = issue_entry("raster_bwd", "raster_bwd_direct_kernel")
            dom.update({"sm_mhz": f_clk / 1e6, "peak_source": "148 SMs x 4 schedulers x measured SM clock",
                        "hbm_peak": hbm_peak, "hbm_peak_source": peak_src + " (MEASURED_PEAKS.json hbm_gbs)",
                        "metrics_source": prof_path, "metrics_state": prof_state,
                        "step_us_strict_forward": step_us})
            bin_e = {"bound": "hbm", "kernel": "gs_bin_fast (multi-kernel stage)", "us": stages["bin_fast"]["us"],
                     "share_of_step": stages["bin_fast"]["us"] / step_us, "achieved": stages["bin_fast"]["GBps"],
                     "peak": hbm_peak, "unit": "GB/s", "frac": stages["bin_fast"]["frac_hbm"],
                     "alg_bytes_per_launch": stages["bin_fast"]["alg_bytes"],
                     "projection_plus_sort_frac": (stages["preprocess_fwd_packed"]["alg_bytes"] + stages["bin_fast"]["alg_bytes"])
                     / (stages["preprocess_fwd_packed"]["us"] + stages["bin_fast"]["us"]) / 1e3 / hbm_peak,
                     "note": "L2-resident, latency-bound at this size (DESIGN.md §4)"}
            dom["stages"] = {"raster_fwd_packed_kernel": issue_entry("raster_fwd", "raster_fwd_packed_kernel"),
                             "preprocess_bwd_kernel": hbm_entry("preprocess_bwd", "preprocess_bwd_kernel"),
                             "preprocess_fwd_kernel": hbm_entry("preprocess_fwd_packed", "preprocess_fwd_kernel"),
                             "gs_bin_fast": bin_e}
            if "frac" not in dom:       # no capture for this scene size: fall back to the largest HBM-bound kernel
                top = max(("preprocess_bwd_kernel", "preprocess_fwd_kernel"), key=lambda k: dom["stages"][k]["us"])
                dom = dict(dom["stages"][top], stages=dom["stages"], note="no ncu capture for this scene size: largest "
                           "HBM-bound kernel; the dominant kernel is raster_bwd_direct_kernel (SM-issue bound)")
            line["roofline"] = dom
            line["roofline_binning"] = bin_e              # round-2 key names, same objects
            line["roofline_raster"] = {"bwd": issue_entry("raster_bwd", "raster_bwd_direct_kernel"),
                                       "fwd": dom["stages"]["raster_fwd_packed_kernel"], "metrics_state": prof_state}
        if world == 1 and not args.no_cpu:
            ref = CpuReference(args.n_gaussians)
            out0, _ = ref.orc.render_view(ref.p, ref.cams[0].as_dict(), SH_DEG, bg=BG, training=True)
            vps, dt = ref.run(args.cpu_views, first=1)
            line["cpu_baseline"] = {"value": vps, "unit": "views/s", "cores": ref.cores, "kind": "port",
                                    "sample": f"{args.cpu_views} views of the same workload (fwd+bwd), C oracle + OpenMP, {dt:.1f} s"}
            if "roofline_raster" in line:
                line["roofline_raster"]["pairs_P_reference_loop"] = int(out0["pairs"])
            del ref
        if world == 1 and not args.no_extra_configs:
            # host cost of one training step: the same loop on a 3 000-Gaussian scene, where the device has nothing to do
            tiny = GaussianModel(scene.make_scene(3000, 0), sh_degree=SH_DEG, background_color=torch.tensor(BG)).to(dev)
            tiny.train()
            line["host_only_ms_per_step"] = time_model_steps(tiny, cams, v_out, 100, 20)
            tiny.speculate = True
            line["host_only_ms_per_step_async_forward"] = time_model_steps(tiny, cams, v_out, 100, 20)
            del tiny
            # a 200 K-Gaussian scene: small enough for the host's share of the step to show next to the device's
            small = GaussianModel(scene.make_scene(200_000, 0), sh_degree=SH_DEG, background_color=torch.tensor(BG)).to(dev)
            small.train()
            ms_strict_small = time_model_steps(small, cams, v_out, 60, 15)
            small.speculate = True
            ms_async_small = time_model_steps(small, cams, v_out, 60, 15)
            line["scene_200k"] = {"views_per_s": 1e3 / ms_async_small, "ms_per_step": ms_async_small,
                                  "strict_forward_ms_per_step": ms_strict_small,
                                  "what": "the training-view loop on 200 K Gaussians (512^2): where the host cost per step "
                                          "(host_only_ms_per_step) is of the order of the device time"}
            del small
            line["scene_large_splats"] = scene_large_splats(dev, v_out, hbm_peak)
            line["config1_api_500k"] = config1_api_500k(dev)
            del model
            torch.cuda.empty_cache()
            line["config5_stress"] = config5_stress(dev)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n-gaussians", type=int, default=N_GAUSS)
    ap.add_argument("--cpu-views", type=int, default=24)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--forward", choices=("async", "strict"), default="async",
                    help="async (default): model.speculate=True in the timed loops, overflowed steps redone; strict: the "
                         "model's default self-healing forward")
    ap.add_argument("--no-stages", action="store_true")
    ap.add_argument("--no-extra-configs", action="store_true",
                    help="skip BASELINE.json configs[1] (500 K through the operator API) and configs[4] (5 M / 1080p stress)")
    ap.add_argument("--render-batch", type=int, default=8)
    ap.add_argument("--render-reps", type=int, default=3)
    ap.add_argument("--reduce", default="peer", choices=["gather", "peer", "allreduce"],
                    help="N>1 gradient exchange: peer = each rank's multi-view leaf backward reads the peers' "
                         "compositor-level grads (36 B/Gaussian) in place over NVLink (symmetric memory, collective fused "
                         "into the kernel); gather = same payload through one NCCL all-gather; allreduce = NCCL "
                         "all-reduce of the 236 B/Gaussian leaf grads")
    ap.add_argument("--exchange", choices=("slice", "push", "dense", "sparse"), default="slice",
                    help="peer-memory exchange: compacted blocks pushed to the peers (default), dense pull, compacted pull")
    ap.add_argument("--path", default="fused", choices=["fused", "api"],
                    help="fused: one autograd node / two C calls per view (render.py); api: the literal "
                         "gsplat-0.1.2 three-operator sequence of gc_model.py")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
