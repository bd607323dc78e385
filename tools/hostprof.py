"""Diagnostic: where the HOST time of one training step goes. The step loop of bench.py on a 3 000-Gaussian scene
(the device has nothing to do, so the loop time is the host's), under cProfile; and the same loop at 1 M Gaussians
as host enqueue time vs wall time.   python tools/hostprof.py"""
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
cams = scene.make_cameras(40, 0)
v = torch.randn(512, 512, 3, device=dev) / (512 * 512)


def loop(m, steps, speculate=False):
    m.speculate = speculate
    params = list(m.parameters())

    def step(i):
        for p in params:
            p.grad = None
        m(cams[i % 40])["rgb"].backward(v)
    for i in range(50):
        step(i)
    torch.cuda.synchronize()
    t = time.perf_counter()
    for i in range(steps):
        step(i)
    h = time.perf_counter() - t
    torch.cuda.synchronize()
    w = time.perf_counter() - t
    return step, h / steps * 1e3, w / steps * 1e3


for n in (3000, 1_000_000):
    m = GaussianModel(scene.make_scene(n, 0), background_color=torch.tensor([0.1, 0.2, 0.3])).to(dev)
    m.train()
    for spec in (False, True):
        step, h, w = loop(m, 200, spec)
        print(f"N={n} speculate={spec}: host enqueue {h:.3f} ms/step, wall {w:.3f} ms/step", flush=True)
    if n == 3000:
        m.speculate = False
        pr = cProfile.Profile()
        pr.enable()
        for i in range(300):
            step(i)
        pr.disable()
        torch.cuda.synchronize()
        st = pstats.Stats(pr)
        st.sort_stats("tottime").print_stats(35)
