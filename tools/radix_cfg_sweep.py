"""Diagnostic: gs_bin_fast timed for every CTA shape of its radix passes (gs_set_radix_cfg: depth sort x partition) on
scenes of several sizes; outputs compared with the default's. CUDA events around the C call alone.
    python tools/radix_cfg_sweep.py"""
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
NAMES = {-1: "auto", 0: "256x16", 1: "512x8", 2: "384x8", 3: "256x12"}
for n, intr in ((200_000, None), (500_000, None), (1_000_000, None), (2_000_000, None), (1_000_000, scene.HD_INTRINSICS),
                (5_000_000, scene.HD_INTRINSICS)):
    p = scene.make_scene(n, 0, device=dev)
    cam = scene.make_cameras(1, 0, intrinsics=intr)[0]
    V, P, _, _, tb = camera_matrices(cam.c2w, cam.fx, cam.fy, cam.W, cam.H)
    with torch.no_grad():
        xys, depths, radii, conics, hit, _ = project_gaussians(
            p["means"], p["scales"].exp(), 1, p["quats"] / p["quats"].norm(dim=-1, keepdim=True), V[:3].to(dev),
            (P @ V).to(dev), cam.fx, cam.fy, cam.cx, cam.cy, cam.H, cam.W, tb)
    T = tb[0] * tb[1]
    ids0, bins0, I = U.bin_fast(xys, depths, radii, hit, tb, impl="lsd")
    ids0 = ids0[:I].clone()
    cap = I + 1024
    out = torch.empty(cap, device=dev, dtype=torch.int32)
    bins = torch.empty_like(bins0)
    total = torch.empty(1, device=dev, dtype=torch.int32)
    scratch = torch.empty(L.gs_bin_fast_scratch_bytes(n, cap, T), device=dev, dtype=torch.uint8)

    def call():
        rc = L.gs_bin_fast(n, cap, _lib.ptr(xys), _lib.ptr(depths), _lib.ptr(radii), _lib.ptr(hit), tb[0], tb[1],
                           _lib.ptr(out), _lib.ptr(bins), _lib.ptr(total), _lib.ptr(scratch), scratch.numel(),
                           _lib.stream_ptr())
        assert rc == 0, rc
    res = {}
    for cd in (-1, 0, 1, 2, 3):
        for cp in (-1, 0, 1, 2, 3):
            if (cd == -1) != (cp == -1):
                continue
            L.gs_set_radix_cfg(0, cd)
            L.gs_set_radix_cfg(1, cp)
            out.zero_()
            for _ in range(3):
                call()
            torch.cuda.synchronize()
            assert int(total.item()) == I and torch.equal(out[:I], ids0) and torch.equal(bins, bins0), (cd, cp)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(REPS):
                call()
            b.record()
            torch.cuda.synchronize()
            res[(cd, cp)] = a.elapsed_time(b) / REPS * 1e3
    L.gs_set_radix_cfg(0, -1)
    L.gs_set_radix_cfg(1, -1)
    print(f"N={n} {cam.W}x{cam.H} I={I}: auto {res[(-1, -1)]:.1f} us", flush=True)
    print("   depth \\ part " + " ".join(f"{NAMES[c]:>8s}" for c in range(4)))
    for cd in range(4):
        print(f"   {NAMES[cd]:>12s} " + " ".join(f"{res[(cd, cp)]:8.1f}" for cp in range(4)), flush=True)
    del p, xys, depths, radii, conics, hit, out, scratch
print("sweep ok")
