"""One-GPU timing of the gradient-exchange kernels with V local view blocks (no NVLink): pack, dense
multi-view backward, sparse multi-view backward."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussctrl_exp_b200 import _lib, render, scene  # noqa: E402
from gaussctrl_exp_b200 import dist as gd  # noqa: E402
from gaussctrl_exp_b200.model import GaussianModel  # noqa: E402

dev = torch.device("cuda", 0)
N = 1_000_000
m = GaussianModel(scene.make_scene(N, 0), background_color=torch.tensor([0.1, 0.2, 0.3])).to(dev)
m.train()
cams = scene.make_cameras(16, 0)
v = torch.randn(512, 512, 3, generator=torch.Generator().manual_seed(1)).to(dev) / (512 * 512)
L = _lib.lib()
m.view_state.defer_leaf_grads = True
blocks = []
for c in range(8):
    for q in m.parameters():
        q.grad = None
    out = m(cams[c])
    (out["rgb"] * v).sum().backward()
    blocks.append(m.view_state.view_block.clone())
use_sh, deg, _ = m.view_state.view_meta
params = {n: getattr(m, n) for n in gd.GRAD_PARAMS}
xs = L.gs_view_xbuf_floats(N)
xbuf = torch.zeros(8 * xs, device=dev)
scratch = torch.empty(L.gs_view_pack_scratch_bytes(N), dtype=torch.uint8, device=dev)
ptrs = torch.tensor([xbuf.data_ptr() + k * xs * 4 for k in range(8)], dtype=torch.int64).to(dev)


def timeit(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps * 1e3


def pack(k):
    _lib.check(L.gs_pack_view_block(N, blocks[k].data_ptr(), xbuf.data_ptr() + k * xs * 4, scratch.data_ptr(),
                                    scratch.numel(), _lib.stream_ptr()), "pack")


for k in range(8):
    pack(k)
print(f"pack one view block: {timeit(lambda: pack(0)):.1f} us")
for V in (1, 2, 4, 8):
    dense = torch.cat(blocks[:V])
    td = timeit(lambda: render.leaf_grads_from_view_blocks(params, dense, V, bool(use_sh), deg, 1.0 / V))
    ts = timeit(lambda: render.leaf_grads_from_view_blocks(params, None, V, bool(use_sh), deg, 1.0 / V,
                                                           peer_ptrs_dev=ptrs.data_ptr(), peer_offset=0, sparse=True))
    print(f"V={V}: dense multi-view backward {td:.1f} us, sparse {ts:.1f} us (local memory)")
