"""Diagnostic: A/B of the packed backward compositing kernel (gs_set_raster_bwd_direct: reductions straight to global
memory vs shared-memory accumulators + flush) inside ONE process (alternating, so that clock / box differences cancel):
the kernel alone (CUDA events around the C call), the training step, at 512x512 / 1080p / a small scene / big splats."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from gaussctrl_exp_b200 import _lib, scene  # noqa: E402
from gaussctrl_exp_b200.model import GaussianModel  # noqa: E402

dev = torch.device("cuda:0")
L = _lib.lib()
for n, intr, s0, tag in ((1_000_000, None, None, "1M 512^2"), (1_000_000, scene.HD_INTRINSICS, None, "1M 1080p"),
                         (200_000, None, None, "200K 512^2"), (100_000, None, 0.03, "100K big splats 512^2")):
    m = GaussianModel(scene.make_scene(n, 0, s0=s0), sh_degree=3, background_color=torch.tensor([0.1, 0.2, 0.3])).to(dev)
    m.train()
    cams = scene.make_cameras(40, 0, intrinsics=intr)
    H, W = cams[0].H, cams[0].W
    v = (torch.randn(H, W, 3, generator=torch.Generator().manual_seed(1234)) / (H * W)).to(dev)
    modes = (0, 1)
    res = {k: [] for k in modes}
    for rep in range(3):
        for d in modes:
            L.gs_set_raster_bwd_direct(d)
            st, I, vis = bench.stage_breakdown(m, cams[0], v, iters=20)
            res[d].append((st["raster_bwd"][0] * 1e6, bench.time_model_steps(m, cams, v, 40, 10)))
    for d, r in res.items():
        print(f"{tag} I={I} direct={d}: raster_bwd {[round(a, 1) for a, _ in r]} us, step {[round(s, 4) for _, s in r]} ms",
              flush=True)
    del m
L.gs_set_raster_bwd_direct(1)
print("raster_ab ok")
