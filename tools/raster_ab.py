"""Diagnostic: A/B of the packed forward compositing kernel's unroll factor inside ONE process (alternating, so that
clock / box differences cancel): kernel alone (CUDA events around the C call) and inside gs_view_fwd."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from gaussctrl_exp_b200 import _lib, scene  # noqa: E402
from gaussctrl_exp_b200.model import GaussianModel  # noqa: E402

dev = torch.device("cuda:0")
L = _lib.lib()
m = GaussianModel(scene.make_scene(1_000_000, 0), sh_degree=3, background_color=torch.tensor([0.1, 0.2, 0.3])).to(dev)
m.train()
cams = scene.make_cameras(40, 0)
v = (torch.randn(512, 512, 3, generator=torch.Generator().manual_seed(1234)) / (512 * 512)).to(dev)
res = {u: [] for u in (1, 2, 3, 4)}
steps = {u: [] for u in (1, 2, 3, 4)}
for rep in range(3):
    for u in (1, 2, 3, 4):
        L.gs_set_raster_fwd_unroll(u)
        st, I, vis = bench.stage_breakdown(m, cams[0], v, iters=20)
        res[u].append((st["raster_fwd"][0] * 1e6, st["view_fwd_total"][0] * 1e6))
        bwd = st["raster_bwd"][0] * 1e6
        steps[u].append(bench.time_model_steps(m, cams, v, 40, 10))
for u in res:
    print(f"unroll {u}: raster_fwd {[round(a, 1) for a, _ in res[u]]} us, view_fwd_total {[round(b, 1) for _, b in res[u]]} us, "
          f"step {[round(s, 4) for s in steps[u]]} ms", flush=True)
print(f"raster_bwd (last measurement) {bwd:.1f} us")
