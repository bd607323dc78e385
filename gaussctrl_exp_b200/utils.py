"""Tile-binning helpers with gsplat-0.1.2's `gsplat.utils` names,
plus `bin_fast`, the traffic-lean path `rasterize_gaussians` uses internally.

These are the steps gsplat's rasterize_gaussians performs before compositing for the calls at
/root/reference/gaussctrl/gc_model.py:208 and :225. All integer outputs are bit-exact against the
oracle (tests/test_parity_gpu.py, tests/test_parity_scale_gpu.py).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
from torch import Tensor

from . import _lib

_scratch_cache = {}


def _scratch(dev, nbytes: int, tag: str) -> Tensor:
    """Grow-only uint8 scratch per (device, stream, tag), owned by torch's caching allocator: two streams binning on one
    device never share radix ping-pong buffers, and a buffer outgrown on one stream is released on that stream."""
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    key = (idx, torch.cuda.current_stream(idx).cuda_stream, tag)
    buf = _scratch_cache.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(int(nbytes * 1.25) + 256, device=dev, dtype=torch.uint8)
        _scratch_cache[key] = buf
    return buf


def compute_cumulative_intersects(num_tiles_hit: Tensor) -> Tuple[int, Tensor]:
    """Inclusive int32 scan of num_tiles_hit; returns (num_intersects, cum_tiles_hit).
    One host read (the total), like the reference's `.item()`."""
    hit = num_tiles_hit.contiguous()
    assert hit.dtype == torch.int32 and hit.is_cuda
    N = hit.shape[0]
    dev = hit.device
    cum = torch.empty_like(hit)
    total = torch.zeros(1, device=dev, dtype=torch.int32)
    L = _lib.lib()
    nbytes = L.gs_scan_scratch_bytes(N)
    scratch = _scratch(dev, nbytes, "scan")
    with torch.cuda.device(dev):
        rc = L.gs_bin_scan(N, _lib.ptr(hit), _lib.ptr(cum), _lib.ptr(total), _lib.ptr(scratch),
                           scratch.numel(), _lib.stream_ptr())
    _lib.check(rc, "gs_bin_scan")
    return _lib.check_count(int(total.item()), "gs_bin_scan"), cum


def map_gaussian_to_intersects(num_points: int, num_intersects: int, xys: Tensor, depths: Tensor,
                               radii: Tensor, cum_tiles_hit: Tensor,
                               tile_bounds: Tuple[int, int, int]) -> Tuple[Tensor, Tensor]:
    """-> (isect_ids int64 [I] = tile_id << 32 | depth bits, gaussian_ids int32 [I])."""
    dev = xys.device
    keys = torch.empty((num_intersects,), device=dev, dtype=torch.int64)
    ids = torch.empty((num_intersects,), device=dev, dtype=torch.int32)
    if num_intersects == 0:
        return keys, ids
    xys, depths = xys.contiguous(), depths.contiguous()
    radii, cum = radii.contiguous(), cum_tiles_hit.contiguous()
    with torch.cuda.device(dev):
        rc = _lib.lib().gs_bin_emit(num_points, _lib.ptr(xys), _lib.ptr(depths), _lib.ptr(radii), _lib.ptr(cum),
                                    int(tile_bounds[0]), int(tile_bounds[1]), _lib.ptr(keys), _lib.ptr(ids),
                                    _lib.stream_ptr())
    _lib.check(rc, "gs_bin_emit")
    return keys, ids


def sort_intersects(isect_ids: Tensor, gaussian_ids: Tensor, num_tiles: int) -> Tuple[Tensor, Tensor]:
    """Stable ascending sort of the 64-bit keys (what torch.sort + gather do in gsplat)."""
    I = isect_ids.shape[0]
    dev = isect_ids.device
    ks = torch.empty_like(isect_ids)
    vs = torch.empty_like(gaussian_ids)
    if I == 0:
        return ks, vs
    L = _lib.lib()
    nbytes = L.gs_sort_scratch_bytes(I)
    scratch = _scratch(dev, nbytes, "sort")
    with torch.cuda.device(dev):
        rc = L.gs_bin_sort(I, int(num_tiles), _lib.ptr(isect_ids), _lib.ptr(gaussian_ids), _lib.ptr(ks),
                           _lib.ptr(vs), _lib.ptr(scratch), scratch.numel(), _lib.stream_ptr())
    _lib.check(rc, "gs_bin_sort")
    return ks, vs


def get_tile_bin_edges(num_intersects: int, isect_ids_sorted: Tensor,
                       tile_bounds: Optional[Tuple[int, int, int]] = None) -> Tensor:
    """-> tile_bins int32 (T,2) = [start,end) per tile, (0,0) for empty tiles. Without
    tile_bounds the table is sized like gsplat 0.1.2 sized it (num_intersects rows)."""
    dev = isect_ids_sorted.device
    T = int(tile_bounds[0]) * int(tile_bounds[1]) if tile_bounds is not None else max(int(num_intersects), 1)
    bins = torch.empty((T, 2), device=dev, dtype=torch.int32)
    with torch.cuda.device(dev):
        rc = _lib.lib().gs_bin_edges(int(num_intersects), T, _lib.ptr(isect_ids_sorted.contiguous()),
                                     _lib.ptr(bins), _lib.stream_ptr())
    _lib.check(rc, "gs_bin_edges")
    return bins


def bin_and_sort_gaussians(num_points: int, num_intersects: int, xys: Tensor, depths: Tensor, radii: Tensor,
                           cum_tiles_hit: Tensor, tile_bounds: Tuple[int, int, int]):
    """-> (isect_ids_unsorted, gaussian_ids_unsorted, isect_ids_sorted, gaussian_ids_sorted, tile_bins)."""
    keys, ids = map_gaussian_to_intersects(num_points, num_intersects, xys, depths, radii, cum_tiles_hit,
                                           tile_bounds)
    T = int(tile_bounds[0]) * int(tile_bounds[1])
    ks, vs = sort_intersects(keys, ids, T)
    bins = get_tile_bin_edges(num_intersects, ks, tile_bounds)
    return keys, ids, ks, vs, bins


# ---------------------------------------------------------------------------------------------
_cap_hint = {}


class PendingCount:
    """Intersection count of a binning call that is on its way to a pinned host word (`bin_fast(..., defer=True)`):
    `wait()` blocks until it has landed — i.e. until the BINNING is done, whatever was enqueued behind it — and returns
    (count, fits the capacity)."""
    __slots__ = ("pinned", "event", "cap", "key", "free")
    _free = {}          # device index -> recycled (pinned word, event): an event stays on the device that first recorded it

    def __init__(self, total: Tensor, cap: int, key):
        self.free = PendingCount._free.setdefault(total.device.index, [])
        if self.free:
            self.pinned, self.event = self.free.pop()
        else:
            self.pinned, self.event = torch.zeros(1, dtype=torch.int32).pin_memory(), torch.cuda.Event()
        self.cap, self.key = cap, key
        self.pinned.copy_(total, non_blocking=True)
        self.event.record(torch.cuda.current_stream(total.device))

    def wait(self):
        self.event.synchronize()
        I = int(self.pinned[0])
        self.free.append((self.pinned, self.event))
        self.pinned = self.event = None
        _lib.check_count(I, "gs_bin_fast")
        if I <= self.cap:
            _cap_hint[self.key] = max(_cap_hint.get(self.key, 0), int(I * 1.25) + 1024)
        else:
            _cap_hint[self.key] = int(I * 1.25) + 1024
        return I, I <= self.cap


def bin_fast(xys: Tensor, depths: Tensor, radii: Tensor, num_tiles_hit: Tensor,
             tile_bounds: Tuple[int, int, int], impl: Optional[str] = None, defer: bool = False):
    """Traffic-lean binning (gs_bin_fast): -> (ids_sorted int32 [>=I], tile_bins (T,2), I).

    The capacity of ids_sorted is speculative (1.25x the largest total seen for this N); the single host
    read of the true total doubles as the overflow check, and an overflow simply re-runs.
    defer=True: nothing is read here; the third value is a PendingCount (or the count itself for an empty set) that
    the caller waits on AFTER it has enqueued the work that consumes the lists — on overflow (wait() -> (I, False)) the
    lists are empty (every tile range is [0, 0)), nothing was written out of bounds, and the caller calls again."""
    dev = xys.device
    N = xys.shape[0]
    tx, ty = int(tile_bounds[0]), int(tile_bounds[1])
    T = tx * ty
    xys, depths = xys.contiguous(), depths.contiguous()
    radii, hit = radii.contiguous(), num_tiles_hit.contiguous()
    key = (dev.index, N, T)
    cap = _cap_hint.get(key, max(4 * N, 1024))
    L = _lib.lib()
    bins = torch.empty((T, 2), device=dev, dtype=torch.int32)
    total = torch.empty(1, device=dev, dtype=torch.int32)
    if impl is None:
        impl = "lsd"
    if impl not in ("bucket", "lsd", "lsd3"):
        raise ValueError(f"unknown binning implementation {impl!r}")
    while True:
        ids_sorted = torch.empty((cap,), device=dev, dtype=torch.int32)
        with torch.cuda.device(dev):
            # gs_bin_fast. "lsd" (default): global depth sort + scan/emission + stable partition by tile; "lsd3": the
            # same with three 11/11/10-bit depth passes; "bucket": per-tile segments filled unordered, then sorted per
            # tile in shared memory — identical output
            special = impl != "lsd"
            if special:
                prev = L.gs_set_depth_sort_passes(3 if impl == "lsd3" else 4) if impl != "bucket" else None
                prev_legacy = L.gs_set_bin_impl(0 if impl == "bucket" else 1)
            try:
                scratch = _scratch(dev, L.gs_bin_fast_scratch_bytes(N, cap, T), "binfast")
                rc = L.gs_bin_fast(N, cap, _lib.ptr(xys), _lib.ptr(depths), _lib.ptr(radii), _lib.ptr(hit), tx, ty,
                                   _lib.ptr(ids_sorted), _lib.ptr(bins), _lib.ptr(total), _lib.ptr(scratch),
                                   scratch.numel(), _lib.stream_ptr())
            finally:
                if special:
                    if prev is not None:
                        L.gs_set_depth_sort_passes(prev)
                    L.gs_set_bin_impl(prev_legacy)
        _lib.check(rc, "gs_bin_fast")
        if defer:
            return ids_sorted, bins, PendingCount(total, cap, key)
        I = _lib.check_count(int(total.item()), "gs_bin_fast")
        if I <= cap:
            _cap_hint[key] = max(_cap_hint.get(key, 0), int(I * 1.25) + 1024)
            return ids_sorted, bins, I
        cap = int(I * 1.25) + 1024
