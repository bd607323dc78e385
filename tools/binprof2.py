"""Diagnostic: the default (LSD) binning pipeline on the 1 M-Gaussian bench scene, timed warm with CUDA events;
run under `ncu --cache-control none --clock-control none --metrics gpu__time_duration.sum` for warm
per-kernel durations (their sum vs the event time = launch gaps)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussctrl_exp_b200 import project_gaussians, scene  # noqa: E402
from gaussctrl_exp_b200 import utils as U  # noqa: E402
from gaussctrl_exp_b200.model import camera_matrices  # noqa: E402

dev = torch.device("cuda:0")
p = scene.make_scene(1_000_000, 0, device=dev)
cam = scene.make_cameras(1, 0)[0]
V, P, _, _, tb = camera_matrices(cam.c2w, cam.fx, cam.fy, cam.W, cam.H)
with torch.no_grad():
    xys, depths, radii, conics, hit, _ = project_gaussians(
        p["means"], p["scales"].exp(), 1, p["quats"] / p["quats"].norm(dim=-1, keepdim=True), V[:3].to(dev),
        (P @ V).to(dev), cam.fx, cam.fy, cam.cx, cam.cy, cam.H, cam.W, tb)
REPS = int(os.environ.get("REPS", "20"))
for rep in range(3):
    ids, bins, I = U.bin_fast(xys, depths, radii, hit, tb, impl="lsd")
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for rep in range(REPS):
    ids, bins, I = U.bin_fast(xys, depths, radii, hit, tb, impl="lsd")
b.record()
torch.cuda.synchronize()
print("I", I, "bin_fast (incl. host allocs)", a.elapsed_time(b) / REPS * 1e3, "us")
