"""Diagnostic: the binning pipelines on the 1 M-Gaussian bench scene (512x512) and a 1080p view, timed warm with CUDA
events around the C call alone (no host read in the timed region), outputs compared with each other.
    python tools/binprof.py            # times "bucket" (default) and "lsd" (round 1)
Run under `ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none` for per-kernel durations."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussctrl_exp_b200 import _lib, project_gaussians, scene  # noqa: E402
from gaussctrl_exp_b200 import utils as U  # noqa: E402
from gaussctrl_exp_b200.model import camera_matrices  # noqa: E402

dev = torch.device("cuda:0")
REPS = int(os.environ.get("REPS", "30"))
L = _lib.lib()
SIZES = ((1_000_000, None), (1_000_000, scene.HD_INTRINSICS), (5_000_000, scene.HD_INTRINSICS))
for n, intr in SIZES[:1 if os.environ.get("ONESIZE") else 3]:
    p = scene.make_scene(n, 0, device=dev)
    cam = scene.make_cameras(1, 0, intrinsics=intr)[0]
    V, P, _, _, tb = camera_matrices(cam.c2w, cam.fx, cam.fy, cam.W, cam.H)
    with torch.no_grad():
        xys, depths, radii, conics, hit, _ = project_gaussians(
            p["means"], p["scales"].exp(), 1, p["quats"] / p["quats"].norm(dim=-1, keepdim=True), V[:3].to(dev),
            (P @ V).to(dev), cam.fx, cam.fy, cam.cx, cam.cy, cam.H, cam.W, tb)
    T = tb[0] * tb[1]
    ref = None
    for impl in [i for i in ("lsd", "lsd3", "bucket") if os.environ.get("ONLY", i) == i]:
        ids, bins, I = U.bin_fast(xys, depths, radii, hit, tb, impl=impl)
        if ref is None:
            ref = (ids[:I].clone(), bins.clone(), I)
        else:
            assert I == ref[2] and torch.equal(ids[:I], ref[0]) and torch.equal(bins, ref[1]), impl
        cap = I + 1024
        out = torch.empty(cap, device=dev, dtype=torch.int32)
        total = torch.empty(1, device=dev, dtype=torch.int32)
        scratch = torch.empty(L.gs_bin_fast_scratch_bytes(n, cap, T), device=dev, dtype=torch.uint8)
        prev = L.gs_set_bin_impl(0 if impl == "bucket" else 1); prevp = L.gs_set_depth_sort_passes(3 if impl == "lsd3" else 4)

        def call():
            rc = L.gs_bin_fast(n, cap, _lib.ptr(xys), _lib.ptr(depths), _lib.ptr(radii), _lib.ptr(hit), tb[0], tb[1],
                               _lib.ptr(out), _lib.ptr(bins), _lib.ptr(total), _lib.ptr(scratch), scratch.numel(),
                               _lib.stream_ptr())
            assert rc == 0
        for _ in range(5):
            call()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(REPS):
            call()
        b.record()
        torch.cuda.synchronize()
        L.gs_set_bin_impl(prev); L.gs_set_depth_sort_passes(prevp)
        us = a.elapsed_time(b) / REPS * 1e3
        print(f"N={n} {cam.W}x{cam.H} I={I} impl={impl}: {us:.1f} us  ({I / us / 1e3:.1f} G intersections/s)", flush=True)
    del p, xys, depths, radii, conics, hit
print("binprof ok")
