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
N_GAUSS = int(sys.argv[1]) if len(sys.argv) > 1 else 1000000
m = GaussianModel(scene.make_scene(N_GAUSS, 0), background_color=torch.tensor([0.1, 0.2, 0.3])).to(dev)
m.train()
cams = scene.make_cameras(16, 0)
v = torch.randn(512, 512, 3, generator=torch.Generator().manual_seed(1)).to(dev) / (512 * 512)
# independent yardstick: every rank's ORDINARY leaf gradients (single-view backward), summed by a plain NCCL all-reduce
for q in m.parameters():
    q.grad = None
(m(cams[rank])["rgb"] * v).sum().backward()
gd.allreduce_flat(m.view_state.flat_grads, average=True)
plain = {k: getattr(m, k).grad.clone() for k in gd.GRAD_PARAMS}
m.view_state.defer_leaf_grads = True
for q in m.parameters():
    q.grad = None
out = m(cams[rank])
(out["rgb"] * v).sum().backward()           # ONE backward: every exchange below consumes the same view block
block = m.view_state.view_block.clone()
res = {}
gd.reduce_view_grads(m)
res["gather"] = {k: getattr(m, k).grad.clone() for k in gd.GRAD_PARAMS}
rel_plain = {k: ((res["gather"][k] - plain[k]).abs().max() / plain[k].abs().max().clamp_min(1e-30)).item()
             for k in gd.GRAD_PARAMS}
print(f"rank {rank}: view-block exchange vs NCCL all-reduce of ordinary leaf gradients: max rel diff "
      f"{max(rel_plain.values()):.2e} {rel_plain}", flush=True)
assert max(rel_plain.values()) < 2e-5 and all(plain[k].abs().max() > 0 for k in gd.GRAD_PARAMS)
for mode in ("peer", "sparse", "push"):
    px = gd.PeerViewExchange(m, mode="dense" if mode == "peer" else mode)
    for rep in range(2):                     # twice: exercises the double buffering
        if mode == "peer":
            m.view_state.view_block_buffer.copy_(block)
            m.view_state.view_block = m.view_state.view_block_buffer
        else:
            m.view_state.view_block = block
        px.reduce(m)
        torch.cuda.synchronize()
    res[mode] = {k: getattr(m, k).grad.clone() for k in gd.GRAD_PARAMS}
    same = all(torch.equal(res["gather"][k], res[mode][k]) for k in gd.GRAD_PARAMS)
    rel = {k: ((res["gather"][k] - res[mode][k]).abs().max() / res["gather"][k].abs().max().clamp_min(1e-30)).item()
           for k in gd.GRAD_PARAMS}
    nbad = {k: int(((res["gather"][k] - res[mode][k]).abs() > 1e-6 * res["gather"][k].abs().max()).sum().item())
            for k in gd.GRAD_PARAMS}
    print(f"rank {rank}: {mode} vs NCCL all-gather on the SAME view blocks: bit-identical={same} "
          f"max rel diff={max(rel.values()):.2e} {rel} elements off by >1e-6: {nbad}", flush=True)
    # the dense kernel is the same code path -> bit-identical; the sparse kernel is a separate translation
    # unit (FMA contraction may differ) -> tolerance
    assert same if mode == "peer" else max(rel.values()) < 2e-6
# owner-sliced exchange: this rank's slice of the reduced gradients, against the same two yardsticks
px = gd.PeerViewExchange(m, mode="slice")
for rep in range(2):
    m.view_state.view_block = block
    px.reduce(m)
    torch.cuda.synchronize()
lo, hi = m.view_state.shard
assert (lo, hi) == gd.slice_bounds(m.num_points, rank, world) and all(getattr(m, k).grad is None for k in gd.GRAD_PARAMS)
sg = m.view_state.sharded_grads
rel_g = {k: ((sg[k] - res["gather"][k][lo:hi]).abs().max() / res["gather"][k].abs().max().clamp_min(1e-30)).item()
         for k in gd.GRAD_PARAMS}
rel_p = {k: ((sg[k] - plain[k][lo:hi]).abs().max() / plain[k].abs().max().clamp_min(1e-30)).item() for k in gd.GRAD_PARAMS}
print(f"rank {rank}: slice [{lo}, {hi}) vs NCCL all-gather exchange: max rel diff {max(rel_g.values()):.2e}; "
      f"vs NCCL all-reduce of ordinary leaf gradients: {max(rel_p.values()):.2e}", flush=True)
assert max(rel_g.values()) < 2e-6 and max(rel_p.values()) < 2e-5
# one optimiser step: sharded Adam on the slices + parameter all-gather == replicated Adam on the all-reduced gradients
lr = dict(means=1.6e-4, scales=5e-3, quats=1e-3, features_dc=2.5e-3, features_rest=1.25e-4, opacities=5e-2)
before = {k: getattr(m, k).detach().clone() for k in gd.GRAD_PARAMS}
ref_p = {k: torch.nn.Parameter(before[k].clone()) for k in gd.GRAD_PARAMS}
for k in gd.GRAD_PARAMS:
    ref_p[k].grad = plain[k].clone()
torch.optim.Adam([dict(params=[ref_p[k]], lr=lr[k]) for k in gd.GRAD_PARAMS], eps=1e-15).step()
for peer_mode in (False, True):        # NCCL broadcast of the updated rows / stored into the peers by the Adam kernel
    with torch.no_grad():
        for k in gd.GRAD_PARAMS:
            getattr(m, k).copy_(before[k])
    dist.barrier()
    opt = gd.ShardedAdam({k: getattr(m, k) for k in gd.GRAD_PARAMS}, rank, world, lr=lr, eps=1e-15, peer=peer_mode)
    opt.step(sg)
    opt.sync()
    torch.cuda.synchronize()
    far = max(((getattr(m, k).detach() - ref_p[k].detach()).abs().max() / lr[k]).item() for k in gd.GRAD_PARAMS)
    print(f"rank {rank}: sharded Adam (peer={peer_mode}) vs replicated Adam: max |delta| / lr {far:.2e}", flush=True)
# Adam's first step moves every touched weight by ~lr * sign(g): compare the updates, relative to lr
upd = {k: ((getattr(m, k).detach() - ref_p[k].detach()).abs().max() / lr[k]).item() for k in gd.GRAD_PARAMS}
# rows whose gradient is ~0 in one arithmetic and exactly 0 in the other flip sign(g): count instead of bounding those
bad = {k: int((((getattr(m, k).detach() - ref_p[k].detach()).abs() / lr[k]) > 1e-3).sum().item()) for k in gd.GRAD_PARAMS}
print(f"rank {rank}: sharded Adam step vs replicated Adam: max |delta| / lr {max(upd.values()):.2e}, "
      f"weights off by > 1e-3 lr: {bad}", flush=True)
assert sum(bad.values()) <= 1e-4 * sum(before[k].numel() for k in gd.GRAD_PARAMS)
with torch.no_grad():
    for k in gd.GRAD_PARAMS:
        getattr(m, k).copy_(before[k])
blk = m.view_state.view_block
Np = (m.num_points + 3) // 4 * 4
nz = (blk[:2 * Np].view(-1, 2).abs().sum(1)[:m.num_points] != 0) | (blk[8 * Np:9 * Np][:m.num_points] != 0) | \
     (blk[5 * Np:8 * Np].view(-1, 3).abs().sum(1)[:m.num_points] != 0)
print(f"rank {rank}: Gaussians with a non-zero view gradient: {nz.float().mean().item():.3f} "
      f"(visible {(m.view_state.radii > 0).float().mean().item():.3f})", flush=True)
# every rank ended with the same gradients (bit pattern checksum)
chk = torch.cat([res["push"][k].reshape(-1) for k in gd.GRAD_PARAMS]).view(torch.int32).to(torch.int64).sum().reshape(1)
chks = [torch.zeros_like(chk) for _ in range(world)]
dist.all_gather(chks, chk)
assert all(int(c.item()) == int(chks[0].item()) for c in chks), "ranks disagree on the reduced gradients"
print(f"rank {rank}: PEER_CHECK_OK", flush=True)
dist.barrier()
dist.destroy_process_group()
