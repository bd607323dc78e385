"""Diagnostic: where the end-to-end step (bench.py's `e2e`) spends its time beyond the device-resident step (`value`).
The same loop with parts switched off, CUDA events around 50 steps each:
    python tools/e2e_anatomy.py"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussctrl_exp_b200 import scene  # noqa: E402
from gaussctrl_exp_b200 import render as grender  # noqa: E402
from gaussctrl_exp_b200.model import GaussianModel  # noqa: E402

dev = torch.device("cuda", 0)
H = W = 512
N = int(os.environ.get("N", "1000000"))
model = GaussianModel(scene.make_scene(N, 0), sh_degree=3, background_color=torch.tensor([0.1, 0.2, 0.3])).to(dev)
model.train()
model.speculate = True
params = [model.means, model.scales, model.quats, model.features_dc, model.features_rest, model.opacities]
cams = scene.make_cameras(40, 0)
for c in cams:
    grender.gs_camera(c)
    model._caller_matrices(c)
g = torch.Generator().manual_seed(1234)
v_out = (torch.randn(H, W, 3, generator=g) / (H * W)).to(dev)
tgt_host = torch.rand(H, W, 3, generator=g).pin_memory()
tgt_res = tgt_host.to(dev)
copy_stream = torch.cuda.Stream(device=dev)
tgt_dev = [torch.empty(H, W, 3, device=dev) for _ in range(2)]
tgt_ready = [torch.cuda.Event() for _ in range(2)]
tgt_free = [torch.cuda.Event() for _ in range(2)]
loss_pinned = [torch.zeros((), dtype=torch.float32).pin_memory() for _ in range(2)]
loss_events = [torch.cuda.Event() for _ in range(2)]


def run(name, h2d, loss_kind, readback, steps=50, warm=10, late_wait=False):
    issued = [-1, -1]
    pending = [False, False]

    def prefetch(i):
        slot = i % 2
        if issued[slot] == i:
            return
        copy_stream.wait_event(tgt_free[slot])
        with torch.cuda.stream(copy_stream):
            tgt_dev[slot].copy_(tgt_host, non_blocking=True)
            tgt_ready[slot].record(copy_stream)
        issued[slot] = i

    def step(i):
        cam = cams[i % 40]
        slot = i % 2
        cur = torch.cuda.current_stream(dev)
        if h2d:
            prefetch(i)
            if not late_wait:
                cur.wait_event(tgt_ready[slot])
            tgt = tgt_dev[slot]
            prefetch(i + 1)
        else:
            tgt = tgt_res
        for p in params:
            p.grad = None
        out = model(cam)
        if h2d and late_wait:
            cur.wait_event(tgt_ready[slot])          # the target is first needed by the loss, not by the render
        if loss_kind == "l1":
            loss = (out["rgb"] - tgt).abs().mean()
            if h2d and late_wait:
                tgt_free[slot].record(cur)           # and last read by the subtraction
            loss.backward()
        elif loss_kind == "l1_torch":
            loss = torch.nn.functional.l1_loss(out["rgb"], tgt)
            loss.backward()
        else:
            loss = None
            out["rgb"].backward(v_out)
        if h2d and not late_wait:
            tgt_free[slot].record(cur)
        if readback and loss is not None:
            loss_pinned[slot].copy_(loss.detach(), non_blocking=True)
            loss_events[slot].record()
            prev = 1 - slot
            if pending[prev]:
                loss_events[prev].synchronize()
                float(loss_pinned[prev])
            pending[slot] = True

    for i in range(warm):
        step(i)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    t0 = time.perf_counter()
    for i in range(steps):
        step(warm + i)
    host = (time.perf_counter() - t0) / steps * 1e3
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / steps
    print(f"{name:58s} {ms:.4f} ms/step  ({1e3 / ms:7.1f} views/s)  host enqueue {host:.3f} ms", flush=True)


run("device-resident, fixed upstream gradient (= value)", False, None, False)
run("+ L1 loss against a resident target (3 + 3 torch kernels)", False, "l1", False)
run("+ L1 via torch.nn.functional.l1_loss", False, "l1_torch", False)
run("+ loss read back (pipelined by one step)", False, "l1", True)
run("+ target from pinned host memory every step (= e2e)", True, "l1", True)
run("  e2e without the read-back", True, "l1", False)
run("  e2e, target awaited right before the loss", True, "l1", True, late_wait=True)
run("  e2e, late wait, without the read-back", True, "l1", False, late_wait=True)
run("device-resident again", False, None, False)
