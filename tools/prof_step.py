"""Minimal workload for ncu: a few training steps (fwd + bwd) and one eval view of the headline configuration
(1 M Gaussians, SH deg 3, 512x512) through the product path. Usage (under gpurun, see profiles/README.md):
    python tools/prof_step.py [steps] [n_gaussians]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussctrl_exp_b200 import scene  # noqa: E402
from gaussctrl_exp_b200.model import GaussianModel  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 4
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1_000_000
dev = torch.device("cuda", 0)
m = GaussianModel(scene.make_scene(n, 0), sh_degree=3, background_color=torch.tensor([0.1, 0.2, 0.3])).to(dev)
m.train()
cams = scene.make_cameras(40, 0)
v = (torch.randn(512, 512, 3, generator=torch.Generator().manual_seed(1234)) / (512 * 512)).to(dev)
for i in range(steps):
    if i == steps - 1:                     # ncu --profile-from-start off: only the last (warm) step and the eval view
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
    for q in m.parameters():
        q.grad = None
    m(cams[0])["rgb"].backward(v)          # the same camera every step: launch k of a kernel is comparable across runs
torch.cuda.synchronize()
m.get_outputs_for_camera(cams[0])
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("prof_step ok")
