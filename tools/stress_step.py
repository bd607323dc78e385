"""Minimal workload for an ncu launch list of the stress configuration: one warm training view and one warm 8-view eval
batch at 5 M Gaussians / 1920x1080, bracketed by cudaProfilerStart/Stop (ncu --profile-from-start off):
    python tools/stress_step.py [n_gaussians]"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussctrl_exp_b200 import scene
from gaussctrl_exp_b200.model import GaussianModel
dev = torch.device("cuda", 0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 5_000_000
m = GaussianModel(scene.make_scene(n, 0), sh_degree=3, background_color=torch.tensor([0.1, 0.2, 0.3])).to(dev)
cams = scene.make_cameras(16, 0, intrinsics=scene.HD_INTRINSICS)
m.train()
v = (torch.randn(1080, 1920, 3, generator=torch.Generator().manual_seed(1)) / (1080 * 1920)).to(dev)
for i in range(3):
    if i == 2:
        torch.cuda.synchronize(); torch.cuda.profiler.start()
    for q in m.parameters(): q.grad = None
    m(cams[0])["rgb"].backward(v)
torch.cuda.synchronize()
m.eval()
torch.cuda.profiler.stop()
m.render_views(cams[:8], batch=8)
torch.cuda.synchronize()
torch.cuda.profiler.start()
m.render_views(cams[:8], batch=8)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok")
