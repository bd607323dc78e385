"""View-parallel multi-GPU plumbing: one process per GPU, Gaussian parameters
replicated, cameras sharded round-robin, and — for the edit fine-tune step
(/root/reference/gaussctrl/gc_trainer.py:258-301) — one sum-allreduce of the six parameter
gradients (59 floats per Gaussian; the optimiser groups of gc_config.py:58-87).

The all-view render (gc_pipeline.py:222-228) needs no collective; `gather_views` exists for a
consumer that wants every image on rank 0.
"""
from __future__ import annotations

import os
from typing import Dict, List, Sequence

import torch
import torch.distributed as dist

GRAD_PARAMS = ("means", "scales", "quats", "features_dc", "features_rest", "opacities")


def init_from_env(backend: str | None = None):
    """Initialise torch.distributed from torchrun's env (RANK/LOCAL_RANK/WORLD_SIZE/MASTER_*).
    Returns (rank, local_rank, world). world == 1 without torchrun: no process group is made."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
            dist.init_process_group(backend, device_id=torch.device("cuda", local))
        else:
            dist.init_process_group(backend)
    return rank, local, world


def shard_views(num_views: int, rank: int, world: int) -> List[int]:
    """Round-robin: rank r renders views {v : v mod world == r} (balances per-view cost variance
    better than contiguous blocks)."""
    return list(range(rank, num_views, world))


class ViewSampler:
    """The reference's training-view sampling (gc_datamanager.py:213-235 `next_train`: pop a random camera from the
    unseen list, refill the list when it is exhausted — i.e. one random permutation of the cameras per epoch), made
    rank-deterministic for data-parallel training: every rank derives the SAME permutation of epoch e
    from (seed, e), the k-th draw of the job is element k mod num_views of epoch k // num_views, and step i hands draw
    i * world + r to rank r — the views `world` successive next_train calls would have returned. No communication, no
    state: view(i) can be asked for any step in any order (a resumed run continues the sequence), the ranks of one step
    never render the same view twice within an epoch, and every camera visits every rank over time instead of the
    same few cameras staying on one rank (a static i * world + r mod num_views assignment pins the expensive views
    to the same ranks whenever world divides num_views)."""

    def __init__(self, num_views: int, rank: int = 0, world: int = 1, seed: int = 0):
        if num_views <= 0 or not (0 <= rank < world):
            raise ValueError("ViewSampler: num_views > 0 and 0 <= rank < world expected")
        self.num_views, self.rank, self.world, self.seed = int(num_views), int(rank), int(world), int(seed)
        self._epochs: Dict[int, List[int]] = {}

    def _epoch(self, e: int) -> List[int]:
        perm = self._epochs.get(e)
        if perm is None:
            import numpy as np
            if len(self._epochs) > 64:
                self._epochs.clear()
            perm = self._epochs[e] = np.random.default_rng([self.seed, e]).permutation(self.num_views).tolist()
        return perm

    def draw(self, k: int) -> int:
        """The k-th view of the whole job (k = 0, 1, ...), identical on every rank."""
        return self._epoch(k // self.num_views)[k % self.num_views]

    def view(self, step: int) -> int:
        """Index of the view this rank renders in `step`."""
        return self.draw(step * self.world + self.rank)

    def views_of_step(self, step: int) -> List[int]:
        """The views of all ranks in `step`, in rank order."""
        return [self.draw(step * self.world + r) for r in range(self.world)]


def owner_of_view(view: int, world: int) -> int:
    return view % world


def allreduce_flat(flat: torch.Tensor, average: bool = True) -> None:
    """The fused render path writes all six leaf gradients into ONE flat buffer (59 floats per
    Gaussian, render.ViewState.flat_grads) and the .grad tensors are views of it: one NCCL call."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return
    dist.all_reduce(flat, op=dist.ReduceOp.SUM)
    if average:
        flat.div_(float(dist.get_world_size()))


def allreduce_grads(params: Dict[str, torch.Tensor] | Sequence[torch.Tensor], average: bool = True) -> None:
    """Sum (or mean) the gradients of the replicated Gaussian parameters across ranks, in place.
    `average=True` keeps the reference's per-view learning-rate semantics."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return
    tensors = list(params.values()) if isinstance(params, dict) else list(params)
    grads = [p.grad for p in tensors if p.grad is not None]
    works = [dist.all_reduce(g, op=dist.ReduceOp.SUM, async_op=True) for g in grads]
    for w in works:
        w.wait()
    if average:
        torch._foreach_div_(grads, float(dist.get_world_size()))


def gather_views(local: Dict[int, torch.Tensor], num_views: int, dst: int = 0):
    """Collect per-view tensors (same shape/dtype everywhere) on rank `dst`; returns a list
    indexed by view on dst, None elsewhere."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return [local[v] for v in range(num_views)]
    world, rank = dist.get_world_size(), dist.get_rank()
    out = [None] * num_views if rank == dst else None
    sample = next(iter(local.values()))
    for v in range(num_views):
        src = owner_of_view(v, world)
        if src == dst:
            if rank == dst:
                out[v] = local[v]
        elif rank == src:
            dist.send(local[v].contiguous(), dst)
        elif rank == dst:
            buf = torch.empty_like(sample)
            dist.recv(buf, src)
            out[v] = buf
    return out


def slice_size(n: int, world: int) -> int:
    """Gaussians per owner: ceil(n / world) rounded up to 128 (= gs_view_slice_size; the leaf-backward CTA and the
    bitmap words of the exchange are aligned to it)."""
    per = (n + world - 1) // world
    return (per + 127) // 128 * 128


def slice_bounds(n: int, rank: int, world: int):
    """[lo, hi) of the Gaussians rank `rank` owns (possibly empty for the last ranks of a small scene)."""
    s = slice_size(n, world)
    lo = min(rank * s, n)
    return lo, min(lo + s, n)


class ShardedAdam:
    """Adam over this rank's slice of the replicated Gaussian parameters, the consumer of the owner-sliced exchange
    (PeerViewExchange mode="slice"): torch.optim.Adam's update rule (what gc_config.py:58-87 configures per parameter
    group and gc_trainer.py:276-298 steps after backward), with the moments kept for the owned rows only (1/world of
    the optimiser state and of the update work per rank). After `step` + `sync` every rank holds the same updated
    parameters as a replicated Adam step on all-reduced gradients would leave.

    CUDA parameters are updated by gs_adam_slice (one fused kernel per parameter tensor). peer=True (NCCL world > 1):
    the six parameters are moved into ONE symmetric-memory buffer and the Adam kernel stores every updated value into
    all peers' copies as it goes (NVLink, no NCCL call on the data path); `sync` is then a device-side barrier.
    peer=False: `sync` broadcasts every owner's rows (NCCL / gloo). CPU parameters (tests under gloo): torch ops."""

    def __init__(self, params: Dict[str, torch.Tensor], rank: int, world: int, lr: Dict[str, float] | float = 1e-3,
                 betas=(0.9, 0.999), eps: float = 1e-15, peer: bool = False):
        self.params, self.rank, self.world = params, rank, world
        first = next(iter(params.values()))
        n = first.shape[0]
        self.n = n
        self.lo, self.hi = slice_bounds(n, rank, world)
        self.lr = {k: (lr[k] if isinstance(lr, dict) else lr) for k in params}
        self.betas, self.eps, self.t = betas, eps, 0
        self.m = {k: torch.zeros_like(p.data[self.lo:self.hi]) for k, p in params.items()}
        self.v = {k: torch.zeros_like(p.data[self.lo:self.hi]) for k, p in params.items()}
        self.cuda = first.is_cuda
        self.peer = bool(peer) and self.cuda and world > 1 and dist.is_initialized()
        self.hdl, self.offset = None, {}
        if self.peer:
            import torch.distributed._symmetric_memory as symm_mem
            off = 0
            for k, p in params.items():
                self.offset[k] = off
                off += (p.numel() + 3) // 4 * 4
            self.buf = symm_mem.empty(off, dtype=torch.float32, device=first.device)
            with torch.no_grad():
                for k, p in params.items():
                    view = self.buf[self.offset[k]:self.offset[k] + p.numel()].view(p.shape)
                    view.copy_(p.data)
                    p.data = view                       # the parameter now lives in symmetric memory
            self.hdl = symm_mem.rendezvous(self.buf, dist.group.WORLD)

    @torch.no_grad()
    def load_reference_state(self, opt_state: Dict[str, dict]) -> None:
        """Resume from the per-group Adam states of a reference checkpoint (postops.load_splatfacto_checkpoint(...,
        with_optimizers=True); gc_trainer.py:158-168 saves them, :74 restores them): this rank keeps its slice of the
        moments, and the step count, learning rates, betas and eps of the file replace the constructor's."""
        steps = {int(s["step"]) for s in opt_state.values()}
        if len(steps) > 1:
            raise ValueError(f"parameter groups at different Adam steps: {sorted(steps)}")
        for k, s in opt_state.items():
            self.m[k].copy_(s["exp_avg"][self.lo:self.hi].to(self.m[k].device))
            self.v[k].copy_(s["exp_avg_sq"][self.lo:self.hi].to(self.v[k].device))
            self.lr[k] = float(s["lr"])
            self.betas, self.eps = tuple(s["betas"]), float(s["eps"])
        if steps:
            self.t = steps.pop()

    @torch.no_grad()
    def step(self, sharded_grads: Dict[str, torch.Tensor]) -> None:
        self.t += 1
        b1, b2 = self.betas
        c1, c2 = 1.0 - b1 ** self.t, 1.0 - b2 ** self.t
        if self.cuda:
            from . import _lib
            L = _lib.lib()
            for k, p in self.params.items():
                width = p.numel() // max(self.n, 1)
                count = (self.hi - self.lo) * width
                if count == 0:
                    continue
                g = sharded_grads[k]
                assert g.is_contiguous() and g.numel() == count and p.data.is_contiguous()
                with torch.cuda.device(p.device):
                    rc = L.gs_adam_slice(count, p.data.data_ptr() + 4 * self.lo * width, g.data_ptr(), self.m[k].data_ptr(),
                                         self.v[k].data_ptr(), self.lr[k] / c1, b1, b2, self.eps, 1.0 / c2 ** 0.5,
                                         self.world if self.peer else 0, self.rank,
                                         self.hdl.buffer_ptrs_dev if self.peer else None,
                                         (self.offset[k] + self.lo * width) if self.peer else 0, _lib.stream_ptr())
                _lib.check(rc, "gs_adam_slice")
            return
        names = list(self.params)
        g = [sharded_grads[k] for k in names]
        m, v = [self.m[k] for k in names], [self.v[k] for k in names]
        torch._foreach_lerp_(m, g, 1.0 - b1)                          # m = b1 m + (1 - b1) g
        torch._foreach_mul_(v, b2)
        torch._foreach_addcmul_(v, g, g, value=1.0 - b2)              # v = b2 v + (1 - b2) g^2
        den = torch._foreach_sqrt(v)
        torch._foreach_div_(den, c2 ** 0.5)
        torch._foreach_add_(den, self.eps)
        for k, mk, dk in zip(names, m, den):
            self.params[k].data[self.lo:self.hi].addcdiv_(mk, dk, value=-self.lr[k] / c1)

    def sync(self) -> None:
        """After this every rank sees every owner's updated rows."""
        if self.world == 1 or not dist.is_initialized():
            return
        if self.peer:
            self.hdl.barrier(channel=1)       # all pushes of this step have landed in every rank's parameters
            return
        self.allgather_params()

    def allgather_params(self) -> None:
        """Every owner broadcasts its updated rows (6 parameters x world owners)."""
        if self.world == 1 or not dist.is_initialized():
            return
        works = []
        for k, p in self.params.items():
            for r in range(self.world):
                lo, hi = slice_bounds(self.n, r, self.world)
                if hi > lo:
                    works.append(dist.broadcast(p.data[lo:hi], src=r, async_op=True))
        for w in works:
            w.wait()


def reduce_view_grads(model, average: bool = True) -> None:
    """Data-parallel gradient step for the fused path WITHOUT a 236 B/Gaussian all-reduce.

    Precondition: `model.view_state.defer_leaf_grads = True` during this step's backward, so each rank holds
    the compositor-level gradients of ITS view (36 B/Gaussian + camera) in `view_state.view_block`.
    The blocks are all-gathered (one NCCL call) and every rank replays the preprocess backward over all of
    them (gs_preprocess_bwd_multi): identical summed (or averaged) leaf gradients on every rank, written
    into one flat buffer and installed as the parameters' .grad. Works for world_size 1 too."""
    from . import render
    st = model.view_state
    block = st.view_block
    assert block is not None, "run backward with view_state.defer_leaf_grads = True first"
    use_sh, deg, _ = st.view_meta
    world = dist.get_world_size() if dist.is_initialized() else 1
    if world > 1:
        gathered = torch.empty(world * block.numel(), device=block.device, dtype=block.dtype)
        dist.all_gather_into_tensor(gathered, block)
    else:
        gathered = block
    params = {k: getattr(model, k) for k in GRAD_PARAMS}
    flat, grads = render.leaf_grads_from_view_blocks(params, gathered, world, bool(use_sh), deg,
                                                     1.0 / world if average else 1.0)
    for k in GRAD_PARAMS:
        getattr(model, k).grad = grads[k]
    st.flat_grads = flat


class PeerViewExchange:
    """reduce_view_grads with the collective fused into the kernels (NVLink peer memory, no NCCL on the data
    path). Three modes:

    "push" (default): the sender compacts its view block (bitmap + word prefix + packed values of the ~1/5 of
        Gaussians with a non-zero gradient) and the pack kernel stores it straight into EVERY peer's
        symmetric-memory slot (gs_pack_view_block_push: posted NVLink stores, ~8 MB per peer instead of 36 MB);
        after one device-side barrier each rank's multi-view backward reads its LOCAL slots only.
    "dense": every rank accumulates its dense block in symmetric memory; after the barrier the multi-view
        backward kernel PULLS all peers' blocks in place (NVLink-ingress bound: 36 MB x (world-1) per step).
    "sparse": compacted like push but pulled by the receivers (dependent remote loads: latency bound).
    "slice": owner-sliced. Rank d owns the Gaussians [d*S, (d+1)*S) (slice_bounds) for the leaf backward and for the
        optimiser step that follows it (ShardedAdam below). Every rank pushes each peer only the part of its compacted
        block that covers the peer's slice (gs_pack_view_block_push_slices: ~1/world of the block per peer) and
        runs the leaf backward over ITS slice and all views (gs_leaf_bwd_slice). NVLink bytes and leaf-backward work
        per rank are then independent of the number of ranks (push/dense/sparse replay ALL Gaussians x ALL views on
        every rank: an O(world) term, 209 us at 8 views against 61 us for one). Result: view_state.sharded_grads
        (slice-shaped) + view_state.shard = (lo, hi); the parameters' .grad are left None.
    Two buffers alternate so that a rank may start writing step s+1 while slower peers still read step s
    (the barrier of step s+1 orders the reuse at step s+2)."""

    def __init__(self, model, sparse: bool | None = None, mode: str | None = None):
        import torch.distributed._symmetric_memory as symm_mem
        from . import _lib
        assert dist.is_initialized()
        L = _lib.lib()
        if mode is None:
            mode = "push" if sparse is None else ("sparse" if sparse else "dense")
        assert mode in ("push", "dense", "sparse", "slice")
        self.mode = mode
        self.world = dist.get_world_size()
        self.rank = dist.get_rank()
        self.sparse = mode != "dense"
        N = model.num_points
        self.stride = L.gs_view_xbuf_floats(N) if self.sparse else L.gs_view_block_floats(N)
        self.slots = self.world if mode in ("push", "slice") else 1
        self.shard = slice_bounds(N, self.rank, self.world)
        self.buf = symm_mem.empty(2 * self.slots * self.stride, dtype=torch.float32, device=model.device)
        self.hdl = symm_mem.rendezvous(self.buf, dist.group.WORLD)
        self.scratch = (torch.empty(L.gs_view_pack_scratch_bytes(N), dtype=torch.uint8, device=model.device)
                        if self.sparse else None)
        if mode in ("push", "slice"):
            base = self.buf.data_ptr()
            self.local_ptrs = torch.tensor([base + v * self.stride * 4 for v in range(self.world)],
                                           dtype=torch.int64).to(model.device)
        self.step = 0
        self.begin(model)

    def pushed_bytes_last_step(self, n_points: int):
        """Bytes this rank stored into its peers' memory in the last `reduce` (mode "slice"), counted from the compacted
        block it pushed: the same float4 run arithmetic as vx_push_slices_kernel (meta words of the peer's slice, the
        camera slot, the run [k_lo, k_hi) of each packed array). A diagnostic (one device->host read of the meta words);
        it stands in for NVLink hardware counters where those are not exposed. Returns (total bytes, non-zero Gaussians)."""
        assert self.mode == "slice" and self.step > 0
        from . import _lib
        k = (self.step - 1) % 2
        off = (k * self.world + self.rank) * self.stride
        Wp = ((((n_points + 31) // 32) + 3) // 4) * 4
        meta = self.buf[off:off + 2 * Wp].view(torch.int32).reshape(Wp, 2).cpu().numpy()
        bitmap, prefix = meta[:, 0].view("uint32"), meta[:, 1]
        import numpy as np
        nnz = int(prefix[-1]) + bin(int(bitmap[-1])).count("1")
        S = _lib.lib().gs_view_slice_size(n_points, self.world)
        sw = S // 32
        total = 0
        for d in range(self.world):
            if d == self.rank:
                continue
            wlo = min(d * sw, Wp)
            whi = min(wlo + sw, Wp)
            if wlo >= whi:
                continue
            k_lo = int(prefix[wlo])
            k_hi = int(prefix[whi]) if whi < Wp else nnz
            quads = ((whi + 1) // 2 - wlo // 2) + 16
            quads += (2 * k_hi + 3) // 4 - (2 * k_lo) // 4
            quads += 2 * ((3 * k_hi + 3) // 4 - (3 * k_lo) // 4)
            quads += (k_hi + 3) // 4 - k_lo // 4
            total += 16 * quads
        return total, nnz

    def begin(self, model) -> None:
        """Point the next backward at this step's buffer (dense mode) / at a private dense block (sparse modes)."""
        k = self.step % 2
        model.view_state.defer_leaf_grads = True
        model.view_state.view_block_buffer = None if self.sparse else self.buf[k * self.stride:(k + 1) * self.stride]

    def reduce(self, model, average: bool = True, on_barrier=None) -> None:
        from . import _lib, render
        st = model.view_state
        assert st.view_block is not None
        use_sh, deg, _ = st.view_meta
        k = self.step % 2
        L = _lib.lib()
        if self.mode == "sparse":
            xbuf = self.buf[k * self.stride:(k + 1) * self.stride]
            with torch.cuda.device(xbuf.device):
                rc = L.gs_pack_view_block(model.num_points, st.view_block.data_ptr(), xbuf.data_ptr(),
                                          self.scratch.data_ptr(), self.scratch.numel(), _lib.stream_ptr())
            _lib.check(rc, "gs_pack_view_block")
        elif self.mode in ("push", "slice"):
            with torch.cuda.device(self.buf.device):
                off = (k * self.world + self.rank) * self.stride
                fn = L.gs_pack_view_block_push if self.mode == "push" else L.gs_pack_view_block_push_slices
                rc = fn(model.num_points, st.view_block.data_ptr(),
                                               self.buf.data_ptr() + 4 * off, self.world, self.rank,
                                               self.hdl.buffer_ptrs_dev, off, self.scratch.data_ptr(),
                                               self.scratch.numel(), _lib.stream_ptr())
            _lib.check(rc, "gs_pack_view_block_push")
        self.hdl.barrier(channel=0)                 # every rank's block of this step is complete and visible
        if on_barrier is not None:                  # profiling hook (tools/dp_profile.py)
            on_barrier()
        params = {n: getattr(model, n) for n in GRAD_PARAMS}
        if self.mode == "slice":
            lo, hi = self.shard
            flat, grads = render.leaf_grads_slice(params, lo, hi, self.world, bool(use_sh), deg,
                                                  1.0 / self.world if average else 1.0, self.local_ptrs.data_ptr(),
                                                  k * self.world * self.stride)
            for n in GRAD_PARAMS:
                getattr(model, n).grad = None
            st.flat_grads, st.sharded_grads, st.shard = flat, grads, (lo, hi)
            self.step += 1
            self.begin(model)
            return
        if self.mode == "push":
            ptrs, off = self.local_ptrs.data_ptr(), k * self.world * self.stride
        else:
            ptrs, off = self.hdl.buffer_ptrs_dev, k * self.stride
        flat, grads = render.leaf_grads_from_view_blocks(
            params, None, self.world, bool(use_sh), deg, 1.0 / self.world if average else 1.0,
            peer_ptrs_dev=ptrs, peer_offset=off, sparse=self.sparse)
        for n in GRAD_PARAMS:
            getattr(model, n).grad = grads[n]
        st.flat_grads = flat
        self.step += 1
        self.begin(model)
