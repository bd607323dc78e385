"""Diagnostic: the 5 M-Gaussian / 1920x1080 stress configuration (BASELINE.json configs[4]) timed with CUDA events —
8-view eval batches and training views — for A/B runs by environment switch (e.g. GS_COLSCAN_SEG_MIN, GS_RADIX_CFG_*):
    python tools/stress_ab.py"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussctrl_exp_b200 import scene
from gaussctrl_exp_b200.model import GaussianModel
dev = torch.device("cuda", 0)
m = GaussianModel(scene.make_scene(5_000_000, 0), sh_degree=3, background_color=torch.tensor([0.1, 0.2, 0.3])).to(dev)
cams = scene.make_cameras(32, 0, intrinsics=scene.HD_INTRINSICS)
m.eval()
m.render_views(cams, batch=8); m.render_views(cams, batch=8)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(3): m.render_views(cams, batch=8)
b.record(); torch.cuda.synchronize()
ms = a.elapsed_time(b) / 3
print(f"eval 5M/1080p batch 8: {32e3/ms:.1f} views/s")
m.train()
v = (torch.randn(1080, 1920, 3, generator=torch.Generator().manual_seed(1)) / (1080 * 1920)).to(dev)
def step(i):
    for q in m.parameters(): q.grad = None
    m(cams[i % 32])["rgb"].backward(v)
for i in range(34): step(i)
torch.cuda.synchronize()
a.record()
for i in range(32): step(i)
b.record(); torch.cuda.synchronize()
print(f"train 5M/1080p: {32e3/a.elapsed_time(b):.1f} views/s")
