"""Diagnostic: all-view eval render throughput vs number of streams, with a host-side cProfile."""
import cProfile
import os
import pstats
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussctrl_exp_b200 import scene  # noqa: E402
from gaussctrl_exp_b200.model import GaussianModel  # noqa: E402

dev = torch.device("cuda:0")
m = GaussianModel(scene.make_scene(1000000, 0), background_color=torch.tensor([0.1, 0.2, 0.3])).to(dev)
cams = scene.make_cameras(40, 0)
for ns in (1, 2, 3, 4):
    for _ in range(2):
        m.render_views(cams, num_streams=ns)
    torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(3):
        m.render_views(cams, num_streams=ns)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t
    print(f"streams={ns}: {120 / dt:.1f} views/s ({dt / 120 * 1e3:.3f} ms/view)")
pr = cProfile.Profile()
pr.enable()
m.render_views(cams, num_streams=2)
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(18)
