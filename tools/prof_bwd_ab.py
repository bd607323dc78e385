"""ncu target: the backward compositing kernel in both reduction modes (gs_set_raster_bwd_direct 0 / 1) at 512x512 and
1080p, one warm training step each inside the profiler range. Launch order: 512 acc, 512 direct, 1080p acc, 1080p direct."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussctrl_exp_b200 import _lib, scene  # noqa: E402
from gaussctrl_exp_b200.model import GaussianModel  # noqa: E402

dev = torch.device("cuda", 0)
L = _lib.lib()
m = GaussianModel(scene.make_scene(1_000_000, 0), sh_degree=3, background_color=torch.tensor([0.1, 0.2, 0.3])).to(dev)
m.train()
for intr in (None, scene.HD_INTRINSICS):
    cam = scene.make_cameras(2, 0, intrinsics=intr)[0]
    v = (torch.randn(cam.H, cam.W, 3, generator=torch.Generator().manual_seed(1234)) / (cam.H * cam.W)).to(dev)
    for direct in (0, 1):
        L.gs_set_raster_bwd_direct(direct)
        for i in range(3):
            if i == 2:
                torch.cuda.synchronize()
                torch.cuda.profiler.start()
            for q in m.parameters():
                q.grad = None
            m(cam)["rgb"].backward(v)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
print("prof_bwd_ab ok")
