"""Diagnostic (torchrun, N GPUs): where the data-parallel step spends its time on each rank —
local fwd+bwd of the rank's view, wait at the exchange barrier (rank skew), multi-view leaf backward."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussctrl_exp_b200 import _lib, render, scene  # noqa: E402
from gaussctrl_exp_b200 import dist as gd  # noqa: E402
from gaussctrl_exp_b200.model import GaussianModel  # noqa: E402

rank, local, world = gd.init_from_env("nccl")
dev = torch.device("cuda", local)
m = GaussianModel(scene.make_scene(1_000_000, 0), background_color=torch.tensor([0.1, 0.2, 0.3])).to(dev)
m.train()
cams = scene.make_cameras(40, 0)
v = torch.randn(512, 512, 3, generator=torch.Generator().manual_seed(1)).to(dev) / (512 * 512)
MODE = os.environ.get("GS_EXCHANGE", "slice")
px = gd.PeerViewExchange(m, mode=MODE)
STEPS, WARM = 40, 10
ev = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(STEPS)]
for i in range(WARM + STEPS):
    k = i - WARM
    cam = cams[gd.ViewSampler(40, rank, world, seed=0).view(i)]
    for q in m.parameters():
        q.grad = None
    if k >= 0:
        ev[k][0].record()
    out = m(cam)
    (out["rgb"] * v).sum().backward()
    if k >= 0:
        ev[k][1].record()
    px.reduce(m, on_barrier=(lambda: ev[k][2].record()) if k >= 0 else None)
    if k >= 0:
        ev[k][3].record()
torch.cuda.synchronize()
loc = sum(e[0].elapsed_time(e[1]) for e in ev) / STEPS
bar = sum(e[1].elapsed_time(e[2]) for e in ev) / STEPS
mul = sum(e[2].elapsed_time(e[3]) for e in ev) / STEPS
tot = ev[0][0].elapsed_time(ev[-1][3]) / STEPS
locs = [e[0].elapsed_time(e[1]) for e in ev]
print(f"rank {rank}: step {tot:.3f} ms = local fwd+bwd {loc:.3f} (min {min(locs):.3f} max {max(locs):.3f}) + pack/push + barrier wait {bar:.3f} "
      f"+ multi-view leaf bwd {mul:.3f} [{MODE}]", flush=True)
dist.barrier()
dist.destroy_process_group()
