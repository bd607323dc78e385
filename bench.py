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


# ------------------------------------------------------------------------------------------------
class CpuReference:
    """The oracle (CPU restatement of the path, kind 'port') on the same workload, all host threads."""

    def __init__(self, n_gauss, threads=None):
        import numpy as np
        from gaussctrl_exp_b200 import scene
        from oracle import oracle as orc
        orc.build()
        if threads:
            orc.set_num_threads(threads)
        self.orc = orc
        self.cores = orc.num_threads()
        self.p = {k: v.numpy() for k, v in scene.make_scene(n_gauss, 0).items()}
        self.cams = scene.make_cameras(N_CAMS, 0)
        self.v_out = (np.random.default_rng(1234).standard_normal((512, 512, 3)) / (512 * 512)).astype(np.float32)

    def view(self, i):
        out, ctx = self.orc.render_view(self.p, self.cams[i % N_CAMS].as_dict(), SH_DEG, bg=BG, training=True)
        self.orc.render_view_backward(self.p, ctx, self.v_out)

    def run(self, views, first=0):
        t0 = time.perf_counter()
        for i in range(views):
            self.view(first + i)
        dt = time.perf_counter() - t0
        return views / dt, dt


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    ref = CpuReference(args.n_gaussians)
    ref.run(min(max(args.warmup, 0), 2))
    _, dt = ref.run(args.steps, first=2)                  # one step = one view, fwd + bwd
    ms = 1e3 * dt / args.steps
    val = 1e3 / ms
    sample = "1 view (fwd+bwd) of the same 1M-Gaussian/512^2 workload per step, C oracle + OpenMP"
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "views/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": workload_config(args.n_gaussians, 1),
        "cpu_baseline": {"value": val, "unit": "views/s", "cores": ref.cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "views/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "gsplat 0.1.2 (the reference's implementation of this path) is not in /root/reference and not "
                "installable offline; this arm times the CPU restatement (oracle/) on the host cores",
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
def stage_breakdown(model, cams, v_out, iters=20):
    """Per-stage device times (CUDA events on the current stream) and rooflines for one view.
    Algorithmic bytes per SURVEY.md §8d."""
    import torch
    from gaussctrl_exp_b200 import project_gaussians, rasterize_gaussians, spherical_harmonics
    from gaussctrl_exp_b200 import utils as U
    from gaussctrl_exp_b200.model import camera_matrices
    dev = model.device
    cam = cams[0]
    H, W = cam.H, cam.W
    V, P, _, _, tb = camera_matrices(cam.c2w, cam.fx, cam.fy, W, H)
    Vd, PVd = V[:3].contiguous().to(dev), (P @ V).to(dev)
    N = model.num_points
    with torch.no_grad():
        means, scales = model.means.detach(), model.scales.detach().exp()
        quats = model.quats.detach() / model.quats.detach().norm(dim=-1, keepdim=True)
        coeffs = torch.cat((model.features_dc.detach()[:, None, :], model.features_rest.detach()), 1).contiguous()
        dirs = means - cam.c2w[:3, 3].to(dev)
        dirs = dirs / dirs.norm(dim=-1, keepdim=True)
        opac = torch.sigmoid(model.opacities.detach())
    bg = torch.tensor(BG, device=dev)

    def timed(fn, n=iters):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n * 1e-3

    from gaussctrl_exp_b200 import _lib
    L = _lib.lib()
    st = {}
    with torch.no_grad():
        xys, depths, radii, conics, hit, cov3d = project_gaussians(means, scales, 1, quats, Vd, PVd, cam.fx, cam.fy,
                                                                   cam.cx, cam.cy, H, W, tb)
        rgbs = torch.clamp(spherical_harmonics(SH_DEG, dirs, coeffs) + 0.5, min=0.0)
        ids_sorted, bins, I = U.bin_fast(xys, depths, radii, hit, tb)
        T = tb[0] * tb[1]
        st["project_fwd"] = (timed(lambda: project_gaussians(means, scales, 1, quats, Vd, PVd, cam.fx, cam.fy, cam.cx,
                                                             cam.cy, H, W, tb)), 96 * N)
        st["sh_fwd"] = (timed(lambda: spherical_harmonics(SH_DEG, dirs, coeffs)), 216 * N)
        # binning through the raw ABI (no host read inside the timed loop)
        cap = ids_sorted.numel()
        scratch = torch.empty(L.gs_bin_fast_scratch_bytes(N, cap, T), device=dev, dtype=torch.uint8)
        total = torch.zeros(1, device=dev, dtype=torch.int32)
        st["bin_fast"] = (timed(lambda: L.gs_bin_fast(N, cap, xys.data_ptr(), depths.data_ptr(), radii.data_ptr(),
                                                       hit.data_ptr(), tb[0], tb[1], ids_sorted.data_ptr(),
                                                       bins.data_ptr(), total.data_ptr(), scratch.data_ptr(),
                                                       scratch.numel(), _lib.stream_ptr())),
                          8 * N + 20 * N + 12 * I + 24 * I + 8 * I + 8 * T)
        scratch2 = torch.empty(L.gs_bin_tile_scratch_bytes(N, cap, T), device=dev, dtype=torch.uint8)
        st["bin_tile"] = (timed(lambda: L.gs_bin_tile(N, cap, xys.data_ptr(), depths.data_ptr(), radii.data_ptr(),
                                                       tb[0], tb[1], ids_sorted.data_ptr(), bins.data_ptr(),
                                                       total.data_ptr(), scratch2.data_ptr(), scratch2.numel(),
                                                       _lib.stream_ptr())),
                          8 * N + 20 * N + 12 * I + 24 * I + 8 * I + 8 * T)
        out = torch.empty((H, W, 3), device=dev)
        fT = torch.empty((H, W), device=dev)
        fidx = torch.empty((H, W), device=dev, dtype=torch.int32)
        opv = opac.view(-1).contiguous()
        order_f = torch.empty(T, device=dev, dtype=torch.int32)
        order_b = torch.empty(T, device=dev, dtype=torch.int32)
        twork = torch.empty(T, device=dev, dtype=torch.int32)
        L.gs_tile_order(T, bins.data_ptr(), 1, order_f.data_ptr(), _lib.stream_ptr())
        st["tile_order"] = (timed(lambda: L.gs_tile_order(T, bins.data_ptr(), 1, order_f.data_ptr(), _lib.stream_ptr())), None)
        fwd = lambda: L.gs_raster_fwd_ordered(H, W, tb[0], tb[1], ids_sorted.data_ptr(), bins.data_ptr(), xys.data_ptr(),
                                              conics.data_ptr(), rgbs.data_ptr(), opv.data_ptr(), None, bg.data_ptr(),
                                              out.data_ptr(), None, fT.data_ptr(), fidx.data_ptr(), order_f.data_ptr(),
                                              twork.data_ptr(), _lib.stream_ptr())
        st["raster_fwd"] = (timed(fwd), None)
        g = [torch.zeros((N, k), device=dev) for k in (2, 3, 3, 1)]
        L.gs_tile_order(T, twork.data_ptr(), 0, order_b.data_ptr(), _lib.stream_ptr())
        bwd = lambda: L.gs_raster_bwd_ordered(H, W, tb[0], tb[1], ids_sorted.data_ptr(), bins.data_ptr(), xys.data_ptr(),
                                              conics.data_ptr(), rgbs.data_ptr(), opv.data_ptr(), bg.data_ptr(),
                                              fT.data_ptr(), fidx.data_ptr(), v_out.data_ptr(), None, g[0].data_ptr(),
                                              g[1].data_ptr(), g[2].data_ptr(), g[3].data_ptr(), order_b.data_ptr(),
                                              _lib.stream_ptr())
        st["raster_bwd"] = (timed(bwd), None)
        vcoef = torch.empty((N, 16, 3), device=dev)
        st["sh_bwd"] = (timed(lambda: L.gs_sh_bwd(N, 16, SH_DEG, dirs.data_ptr(), g[2].data_ptr(), vcoef.data_ptr(),
                                                  _lib.stream_ptr())), 216 * N)
        # fused per-view kernels (render.py path)
        import ctypes
        from gaussctrl_exp_b200 import render
        gc = render.gs_camera(cam)
        raw = [t.detach().contiguous() for t in (model.means, model.scales, model.quats, model.opacities,
                                                 model.features_dc, model.features_rest)]
        o_xy, o_d, o_c, o_rgb, o_op = (torch.empty_like(t) for t in (xys, depths, conics, rgbs, opv))
        o_r, o_h = torch.empty_like(radii), torch.empty_like(hit)
        st["preprocess_fwd"] = (timed(lambda: L.gs_preprocess_fwd(
            N, raw[0].data_ptr(), raw[1].data_ptr(), raw[2].data_ptr(), raw[3].data_ptr(), raw[4].data_ptr(),
            raw[5].data_ptr(), 1, SH_DEG, ctypes.addressof(gc), o_xy.data_ptr(), o_d.data_ptr(), o_r.data_ptr(),
            o_c.data_ptr(), o_h.data_ptr(), o_rgb.data_ptr(), o_op.data_ptr(), _lib.stream_ptr())), 284 * N)
        gl = [torch.empty_like(t) for t in raw]
        st["preprocess_bwd"] = (timed(lambda: L.gs_preprocess_bwd(
            N, raw[0].data_ptr(), raw[1].data_ptr(), raw[2].data_ptr(), raw[3].data_ptr(), 1, SH_DEG,
            ctypes.addressof(gc), o_r.data_ptr(), o_rgb.data_ptr(), g[0].data_ptr(), g[1].data_ptr(), g[2].data_ptr(),
            g[3].data_ptr(), gl[0].data_ptr(), gl[1].data_ptr(), gl[2].data_ptr(), gl[3].data_ptr(), gl[4].data_ptr(),
            gl[5].data_ptr(), _lib.stream_ptr())), 332 * N)
        vm, vs, vq = torch.empty_like(means), torch.empty_like(means), torch.empty_like(quats)
        vdep = torch.zeros(N, device=dev)
        st["project_bwd"] = (timed(lambda: L.gs_project_bwd(N, means.data_ptr(), scales.data_ptr(), 1.0, quats.data_ptr(),
                                                            Vd.data_ptr(), PVd.data_ptr(), cam.fx, cam.fy, H, W,
                                                            radii.data_ptr(), g[0].data_ptr(), vdep.data_ptr(),
                                                            g[1].data_ptr(), vm.data_ptr(), vs.data_ptr(), vq.data_ptr(),
                                                            _lib.stream_ptr())), 144 * N)
    return st, I, int((radii > 0).sum().item())


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
    g = torch.Generator().manual_seed(1234)
    v_host = (torch.randn(H, W, 3, generator=g) / (H * W)).pin_memory()
    v_out = v_host.to(dev)
    tgt_host = torch.rand(H, W, 3, generator=g).pin_memory()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

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

    def step(i):
        cam = cams[(i * world + rank) % N_CAMS]
        for p in params:
            p.grad = None
        out = model(cam)
        # "a fixed upstream gradient, then a8 backward down to the leaf grads" — v_out IS the upstream
        # gradient of the rendered image, handed to autograd directly (no torch loss kernels inside the step; the
        # e2e loop below has a real L1 loss against a target image copied from the host)
        out["rgb"].backward(v_out)
        reduce_grads()

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

    def step_e2e(i):
        cam = cams[(i * world + rank) % N_CAMS]
        slot = i % 2
        prefetch_target(i)                                  # no-op except on the very first step
        cur = torch.cuda.current_stream(dev)
        cur.wait_event(tgt_ready[slot])
        tgt = tgt_dev[slot]
        prefetch_target(i + 1)                              # H2D of the NEXT step's target overlaps this step
        for p in params:
            p.grad = None
        out = model(cam)
        loss = (out["rgb"] - tgt).abs().mean()
        loss.backward()
        tgt_free[slot].record(cur)
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
        host_ms[0] = 1e3 * (time.perf_counter() - h0) / steps     # host enqueue time per step (diagnostic)
        if drain:
            drain()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        t = torch.tensor([ms], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), _lib.launch_count() - l0

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms_total, launches = timed_loop(step, args.warmup, args.steps)
    host_enqueue_ms = host_ms[0]
    clocks = sampler.stop() if rank == 0 else None

    # per-step distribution (median and p10/p90): the same steps, each bracketed by its own events
    # (no host sync in between; events are read back afterwards). Diagnostic only: `value` is the loop above.
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

    # config 3: all-view eval render, RGB + depth + alpha, 40 views sharded over the ranks
    model.eval_render = True
    def render_all():
        return model.render_views(cams, rank, world, num_streams=args.render_streams)
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
            "eval_render": {"views_per_s": render_vps, "what": f"all-view RGB+depth+alpha render of {N_CAMS} cameras "
                            f"(gc_pipeline.py:222-228), view-sharded over {world} GPU(s), {args.render_streams} streams/GPU"},
            "host_enqueue_ms_per_step": host_enqueue_ms, "step_time_distribution": step_dist,
            "e2e": {"value": world * 1e3 / (ms_e2e / args.steps), "unit": "views/s",
                    "h2d_bytes_per_step": int(tgt_host.numel() * 4 + 12 * 4 + 6 * 4), "d2h_bytes_per_step": 4},
        }
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak, peak_src = (peaks.get("hbm_gbs"), "measured") if peaks.get("hbm_gbs") else (6650.0, "fallback")
        if not args.no_stages:
            st, I, vis = stage_breakdown(model, cams, v_out)
            stages = {}
            for k, (sec, byts) in st.items():
                stages[k] = {"us": sec * 1e6}
                if byts:
                    stages[k].update({"alg_bytes": int(byts), "GBps": byts / sec / 1e9, "frac_hbm": byts / sec / 1e9 / hbm_peak})
            line["stages"] = stages
            line["intersections"] = I
            line["visible"] = vis
            # Kernels of the measured (fused) step: preprocess_fwd, binning, tile_order x2, raster_fwd, raster_bwd,
            # preprocess_bwd. The dominant one (raster_bwd3, ~35 % of the step) is bound by SM issue rate, not
            # HBM or tensor cores (no dense contraction anywhere on this path) -> `roofline_raster`.
            # `roofline` is the dominant HBM-bound kernel of the step, in the schema's units.
            step_hbm = {k: stages[k] for k in ("preprocess_fwd", "preprocess_bwd") if k in stages}
            top = max(step_hbm, key=lambda k: step_hbm[k]["us"])
            ncu_traffic = {"preprocess_bwd": 273.8e6, "preprocess_fwd": 268.6e6}   # dram rd+wr per launch (ncu --set full; profiles/README.md r01k)
            line["roofline"] = {"bound": "hbm", "kernel": top + "_kernel", "achieved": step_hbm[top]["GBps"],
                                "peak": hbm_peak, "peak_source": peak_src + " (MEASURED_PEAKS.json hbm_gbs)",
                                "unit": "GB/s", "frac": step_hbm[top]["frac_hbm"],
                                "alg_bytes_per_launch": step_hbm[top]["alg_bytes"],
                                "traffic": ncu_traffic.get(top) if args.n_gaussians == N_GAUSS else None,
                                "note": "dominant HBM-bound kernel of the step; the dominant kernel overall is "
                                        "SM-issue bound, see roofline_raster; multi-kernel binning stage: stages.bin_fast"}
            bin_key = "bin_tile" if os.environ.get("GS_BIN_IMPL", "lsd") == "tile" else "bin_fast"
            line["roofline_binning"] = {"bound": "hbm", "kernel": "gs_" + bin_key + " (multi-kernel stage)",
                                        "achieved": stages[bin_key]["GBps"], "peak": hbm_peak, "unit": "GB/s",
                                        "frac": stages[bin_key]["frac_hbm"],
                                        "note": "L2-resident, latency-bound at this size (see DESIGN.md §4)"}
        pairs = None
        if world == 1 and not args.no_cpu:
            ref = CpuReference(args.n_gaussians)
            out0, _ = ref.orc.render_view(ref.p, ref.cams[0].as_dict(), SH_DEG, bg=BG, training=True)
            pairs = int(out0["pairs"])                      # (pixel, Gaussian) pairs of the reference algorithm, view 0
            vps, dt = ref.run(args.cpu_views, first=1)
            line["cpu_baseline"] = {"value": vps, "unit": "views/s", "cores": ref.cores, "kind": "port",
                                    "sample": f"{args.cpu_views} views of the same workload (fwd+bwd), C oracle + OpenMP, {dt:.1f} s"}
        if not args.no_stages and pairs:
            # a front-to-back compositor must evaluate P pairs at ~16 FP32-pipe instructions each
            # (backward ~2.5x); ceiling = 148 SMs x 128 lanes x f_clk lane-instructions/s
            f_clk = (clocks.get("sm_mhz") or 1965.0) * 1e6
            peak_lane_ips = 148 * 128 * f_clk
            rf = {}
            for k, ipp in (("raster_fwd", 16.0), ("raster_bwd", 40.0)):
                t = stages[k]["us"] * 1e-6
                rf[k] = {"pairs_per_s": pairs / t, "lane_inst_per_s": pairs * ipp / t, "frac": pairs * ipp / t / peak_lane_ips,
                         "us": stages[k]["us"], "model_inst_per_pair": ipp}
            line["roofline_raster"] = {"bound": "sm_issue", "kernel": "raster_bwd3_kernel", "pairs_P": pairs,
                                       "peak_lane_inst_per_s": peak_lane_ips, "sm_mhz": f_clk / 1e6,
                                       "achieved": rf["raster_bwd"]["lane_inst_per_s"], "unit": "lane-inst/s",
                                       "frac": rf["raster_bwd"]["frac"], "fwd": rf["raster_fwd"], "bwd": rf["raster_bwd"],
                                       "note": "P = pairs the reference's per-pixel loop visits (oracle, camera 0); the "
                                               "kernels skip pairs whose alpha-extent box misses the warp, so frac is an "
                                               "algorithmic-work rate, not an issue-slot utilisation (ncu: profiles/)"}
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
    ap.add_argument("--no-stages", action="store_true")
    ap.add_argument("--render-streams", type=int, default=2)
    ap.add_argument("--render-reps", type=int, default=3)
    ap.add_argument("--reduce", default="peer", choices=["gather", "peer", "allreduce"],
                    help="N>1 gradient exchange: peer = each rank's multi-view leaf backward reads the peers' "
                         "compositor-level grads (36 B/Gaussian) in place over NVLink (symmetric memory, collective fused "
                         "into the kernel); gather = same payload through one NCCL all-gather; allreduce = NCCL "
                         "all-reduce of the 236 B/Gaussian leaf grads")
    ap.add_argument("--exchange", choices=("push", "dense", "sparse"), default="push",
                    help="peer-memory exchange: compacted blocks pushed to the peers (default), dense pull, compacted pull")
    ap.add_argument("--path", default="fused", choices=["fused", "api"],
                    help="fused: one autograd node / 5 kernels per view (render.py); api: the literal "
                         "gsplat-0.1.2 three-operator sequence of gc_model.py")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
