"""2+ GPU check (torchrun): the NVLink peer-memory gradient exchange equals the NCCL all-gather exchange."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussctrl_exp_b200 import dist as gd  # noqa: E402
from gaussctrl_exp_b200 import scene  # noqa: E402
from gaussctrl_exp_b200.model import GaussianModel  # noqa: E402

rank, local, world = gd.init_from_env("nccl")
dev = torch.device("cuda", local)
m = GaussianModel(scene.make_scene(200000, 0, s0=0.006), background_color=torch.tensor([0.1, 0.2, 0.3])).to(dev)
m.train()
cams = scene.make_cameras(16, 0)
v = torch.randn(512, 512, 3, generator=torch.Generator().manual_seed(1)).to(dev) / (512 * 512)
m.view_state.defer_leaf_grads = True
res = {}
peer = None
for mode in ("gather", "peer", "peer"):
    if mode == "peer" and peer is None:
        peer = gd.PeerViewExchange(m)
    for q in m.parameters():
        q.grad = None
    out = m(cams[rank])
    (out["rgb"] * v).sum().backward()
    if mode == "gather":
        m.view_state.view_block_buffer = None
        gd.reduce_view_grads(m)
    else:
        peer.reduce(m)
    torch.cuda.synchronize()
    res[mode] = {k: getattr(m, k).grad.clone() for k in gd.GRAD_PARAMS}
# the compositor's float atomics make two runs differ at the 1e-7 relative level: compare with a tolerance
rel = max(((res["gather"][k] - res["peer"][k]).abs().max() / res["gather"][k].abs().max().clamp_min(1e-30)).item()
          for k in gd.GRAD_PARAMS)
print(f"rank {rank}: peer vs gather max rel diff {rel:.3e} -> {'OK' if rel < 1e-5 else 'MISMATCH'}", flush=True)
assert rel < 1e-5
dist.barrier()
dist.destroy_process_group()
