"""`rasterize_gaussians` with the gsplat-0.1.2 signature, backed by the sm_100a kernels.

Replaces the calls at /root/reference/gaussctrl/gc_model.py:208-220 (RGB + alpha) and
:225-236 (depth as 3 replicated channels); import at :36. Semantics: SURVEY.md App. A.3-A.5.
Differentiable w.r.t. xys, conics, colors, opacity (exact derivative of the forward).
"""
from __future__ import annotations

import weakref
from typing import Optional

import torch
from torch import Tensor
from torch.autograd import Function

from . import _lib
from .utils import bin_fast

# The eval path composites twice per view on identical geometry (gc_model.py:208, :225) and
# gsplat re-bins and re-sorts for the second call. When the SAME xys/depths/radii tensor
# objects (unchanged versions) come back, the previous binning is reused.
BINNING_CACHE = True
_bin_cache = {}


def _binning(xys, depths, radii, num_tiles_hit, H, W):
    """-> ((ids_sorted, tile_bins, count | PendingCount), tile_bounds, cache slot | None). A PendingCount means the
    count has not been read yet: the caller enqueues the compositing first and confirms the capacity afterwards
    (`_confirm`), so the host never idles the device between binning and compositing."""
    if xys.dtype != torch.float32 or not xys.is_contiguous():
        xys = xys.float().contiguous()
    tile_bounds = ((W + 15) // 16, (H + 15) // 16, 1)
    slot = None
    if BINNING_CACHE:
        sig = (xys._version, depths._version, radii._version, H, W)
        hit = _bin_cache.get(xys.device.index)
        # weak references: the cache keeps neither the caller's tensors nor their autograd graph alive, and a
        # recycled allocation (same address, new tensor) can never hit
        if hit is not None and hit[0]() is xys and hit[1]() is depths and hit[2]() is radii and hit[3] == sig:
            return hit[4], tile_bounds, None
        slot = (weakref.ref(xys), weakref.ref(depths), weakref.ref(radii), sig)
    res = bin_fast(xys, depths, radii, num_tiles_hit, tile_bounds, defer=True)
    return res, tile_bounds, slot


def _confirm(res, slot, dev_index) -> bool:
    """Waits for a deferred count; True if the lists fitted their capacity (then they may be cached)."""
    ids_sorted, tile_bins, cnt = res
    if not isinstance(cnt, int):
        I, ok = cnt.wait()
        if not ok:
            return False
        res = (ids_sorted, tile_bins, I)
    if slot is not None:
        _bin_cache[dev_index] = slot + (res,)
    return True


def clear_binning_cache():
    _bin_cache.clear()


def rasterize_gaussians(
    xys: Tensor,
    depths: Tensor,
    radii: Tensor,
    conics: Tensor,
    num_tiles_hit: Tensor,
    colors: Tensor,
    opacity: Tensor,
    img_height: int,
    img_width: int,
    background: Optional[Tensor] = None,
    return_alpha: bool = False,
) -> Tensor:
    """-> out_img (H,W,C) [, out_alpha (H,W)]. Zero intersections is not an error: the image is
    the background and alpha is 0."""
    if colors.dtype == torch.uint8:
        colors = colors.float() / 255
    if background is not None:
        assert background.shape[0] == colors.shape[-1], \
            f"incorrect shape of background color tensor, expected shape {colors.shape[-1]}"
    else:
        background = torch.ones(colors.shape[-1], dtype=torch.float32, device=colors.device)
    if xys.ndimension() != 2 or xys.size(1) != 2:
        raise ValueError("xys must have dimensions (N, 2)")
    if colors.ndimension() != 2:
        raise ValueError("colors must have dimensions (N, D)")
    if not xys.is_cuda:
        raise ValueError("rasterize_gaussians needs CUDA tensors (no CPU fallback)")
    H, W = int(img_height), int(img_width)
    if xys.shape[0] == 0:
        # empty Gaussian set: nothing to bin or composite (gsplat returns the background when there are no
        # intersections; the caller's own early-out is gc_model.py:189-190). Kept on the autograd graph of the inputs.
        zero = xys.sum() + colors.sum()
        out_img = background.to(xys.device).float().expand(H, W, colors.shape[-1]) + zero
        if return_alpha:
            return out_img, torch.zeros((H, W), device=xys.device, dtype=torch.float32) + zero
        return out_img
    while True:
        with torch.no_grad():
            res, tb, slot = _binning(xys, depths, radii, num_tiles_hit, H, W)
        # the compositing is enqueued BEFORE the intersection count is read: the host's wait ends when the binning is
        # done, while the device goes on compositing (gsplat reads the count between the two and idles the device)
        out = _RasterizeGaussians.apply(xys, conics, colors, opacity, background, res[0], res[1], H, W, tb, return_alpha)
        if _confirm(res, slot, xys.device.index):
            return out
        # the tile lists did not fit their speculative capacity (nothing was written, the image above is the
        # background): the capacity hint has been raised — bin and composite again


class _RasterizeGaussians(Function):
    """Three channels (the RGB call of gc_model.py:208 and the replicated-depth call of :225) run on the packed
    compositing kernels: the operator's tensors are packed once per call into 48-byte splat records
    (gs_pack_splats, one pass over N), the tiles are launched heaviest-first, and forward / backward are
    gs_raster_fwd_packed / gs_raster_bwd_packed — the same kernels the fused path uses. Other channel counts go
    through the generic kernels (gs_raster_fwd / gs_raster_bwd)."""

    @staticmethod
    def forward(ctx, xys, conics, colors, opacity, background, ids_sorted, tile_bins, H, W, tb,
                return_alpha=False):
        N = xys.shape[0]
        if conics.shape != (N, 3) or colors.shape[0] != N or opacity.numel() != N:
            raise ValueError("conics (N,3), colors (N,C) and opacity (N,1) must match xys (N,2)")
        dev = xys.device
        C = colors.shape[1]
        xys_f, conics_f = xys.float().contiguous(), conics.float().contiguous()
        colors_f, opac_f = colors.float().contiguous(), opacity.float().contiguous().view(-1)
        bg = background.to(dev).float().contiguous()
        out_img = torch.empty((H, W, C), device=dev, dtype=torch.float32)
        final_T = torch.empty((H, W), device=dev, dtype=torch.float32)
        final_idx = torch.empty((H, W), device=dev, dtype=torch.int32)
        L = _lib.lib()
        T = tb[0] * tb[1]
        rec = order = work = None
        with torch.cuda.device(dev):
            st = _lib.stream_ptr()
            if C == 3:
                rec = torch.empty((N, 12), device=dev, dtype=torch.float32)
                order = torch.empty((T,), device=dev, dtype=torch.int32)
                work = torch.empty((T,), device=dev, dtype=torch.int32)
                _lib.check(L.gs_pack_splats(N, _lib.ptr(xys_f), _lib.ptr(conics_f), _lib.ptr(colors_f), _lib.ptr(opac_f),
                                            None, _lib.ptr(rec), st), "gs_pack_splats")
                _lib.check(L.gs_tile_order(T, _lib.ptr(tile_bins), 1, _lib.ptr(order), st), "gs_tile_order")
                rc = L.gs_raster_fwd_packed(1, H, W, tb[0], tb[1], _lib.ptr(ids_sorted), _lib.ptr(tile_bins), _lib.ptr(rec),
                                            _lib.ptr(bg), _lib.ptr(out_img), None, _lib.ptr(final_T), _lib.ptr(final_idx),
                                            _lib.ptr(order), _lib.ptr(work), None, None, st)
            else:
                rc = L.gs_raster_fwd(
                    C, H, W, tb[0], tb[1], _lib.ptr(ids_sorted), _lib.ptr(tile_bins), _lib.ptr(xys_f),
                    _lib.ptr(conics_f), _lib.ptr(colors_f), _lib.ptr(opac_f), _lib.ptr(bg), _lib.ptr(out_img),
                    _lib.ptr(final_T), _lib.ptr(final_idx), st)
        _lib.check(rc, "gs_raster_fwd")
        if C == 3:
            ctx.save_for_backward(ids_sorted, tile_bins, rec, work, bg, final_T, final_idx)
        else:
            ctx.save_for_backward(ids_sorted, tile_bins, xys_f, conics_f, colors_f, opac_f, bg, final_T, final_idx)
        ctx.meta = (H, W, C, tb, tuple(opacity.shape), N)
        if return_alpha:
            return out_img, 1.0 - final_T
        return out_img

    @staticmethod
    def backward(ctx, v_out_img, v_out_alpha=None):
        H, W, C, tb, opac_shape, N = ctx.meta
        dev = v_out_img.device
        v_out_img = v_out_img.float().contiguous()
        if v_out_alpha is not None:
            v_out_alpha = v_out_alpha.float().contiguous()
        # the four gradient arrays are accumulated into by atomics: ONE zero-filled allocation, carved up
        g = torch.zeros(((6 + C) * N,), device=dev, dtype=torch.float32)
        v_xys, v_conics = g[0:2 * N].view(N, 2), g[2 * N:5 * N].view(N, 3)
        v_colors, v_opac = g[5 * N:(5 + C) * N].view(N, C), g[(5 + C) * N:]
        L = _lib.lib()
        with torch.cuda.device(dev):
            st = _lib.stream_ptr()
            if C == 3:
                ids_sorted, tile_bins, rec, work, bg, final_T, final_idx = ctx.saved_tensors
                order = torch.empty_like(work)
                _lib.check(L.gs_tile_order(tb[0] * tb[1], _lib.ptr(work), 0, _lib.ptr(order), st), "gs_tile_order")
                rc = L.gs_raster_bwd_packed(H, W, tb[0], tb[1], _lib.ptr(ids_sorted), _lib.ptr(tile_bins), _lib.ptr(rec),
                                            _lib.ptr(bg), _lib.ptr(final_T), _lib.ptr(final_idx), _lib.ptr(v_out_img),
                                            _lib.ptr(v_out_alpha), _lib.ptr(v_xys), _lib.ptr(v_conics), _lib.ptr(v_colors),
                                            _lib.ptr(v_opac), _lib.ptr(order), None, st)
            else:
                ids_sorted, tile_bins, xys, conics, colors, opac, bg, final_T, final_idx = ctx.saved_tensors
                rc = L.gs_raster_bwd(
                    C, H, W, tb[0], tb[1], _lib.ptr(ids_sorted), _lib.ptr(tile_bins), _lib.ptr(xys), _lib.ptr(conics),
                    _lib.ptr(colors), _lib.ptr(opac), _lib.ptr(bg), _lib.ptr(final_T), _lib.ptr(final_idx),
                    _lib.ptr(v_out_img), _lib.ptr(v_out_alpha), _lib.ptr(v_xys), _lib.ptr(v_conics),
                    _lib.ptr(v_colors), _lib.ptr(v_opac), st)
        _lib.check(rc, "gs_raster_bwd")
        return (v_xys, v_conics, v_colors, v_opac.view(opac_shape), None, None, None, None, None, None, None)
