"""Fused per-view render: ONE autograd node from the six splatfacto leaf tensors to (rgb, alpha).

Same result as the three-operator sequence of /root/reference/gaussctrl/gc_model.py:172-222
(pre-activations -> project_gaussians -> spherical_harmonics -> rasterize_gaussians) and its
backward (gc_trainer.py:275), but as five kernel launches per training view:
    forward : gs_preprocess_fwd -> gs_bin_fast -> gs_raster_fwd
    backward: gs_raster_bwd -> gs_preprocess_bwd
with no device->host synchronisation other than one 4-byte read of the intersection count, which
is overlapped with the compositing kernel (the tile lists are allocated speculatively; the read
only confirms the capacity was enough, otherwise the binning is redone).
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import numpy as np
import torch
from torch import Tensor
from torch.autograd import Function

from . import _lib
from .scene import Camera

# Training views (render_view under autograd): do not read the intersection count in the forward; verify it
# at the end of backward instead (raises CapacityOverflow, before any gradient is returned, in the unlikely
# case the count grew by more than 50 % between consecutive views of the same shape).
SPECULATE = True

# "tile": count / scan / scatter / per-tile shared-memory sort (gs_bin_tile, 5 launches);
# "lsd": global depth sort + stable partition by tile (gs_bin_fast). Identical outputs.
import os as _os
BIN_IMPL = _os.environ.get("GS_BIN_IMPL", "lsd")

_cam_cache = {}
_cap_hint = {}
_scratch = {}
_pinned = {}


def gs_camera(camera: Camera, glob_scale: float = 1.0, clip: float = 0.01) -> _lib.GsCamera:
    """Host-side matrices of gc_model.py:117-155 packed into the C struct (cached per camera)."""
    key = id(camera)
    hit = _cam_cache.get(key)
    if hit is not None and hit[0] is camera:
        return hit[1]
    c2w = camera.c2w.detach().cpu().numpy().astype(np.float32)
    W, H = int(camera.W), int(camera.H)
    R = c2w[:3, :3] @ np.diag(np.array([1, -1, -1], np.float32))       # :129-130
    T = c2w[:3, 3:4]
    Rinv = R.T
    viewmat = np.eye(4, dtype=np.float32)
    viewmat[:3, :3] = Rinv
    viewmat[:3, 3:4] = -Rinv @ T
    fovx = 2 * math.atan(W / (2 * camera.fx))                           # :143-144
    fovy = 2 * math.atan(H / (2 * camera.fy))
    n, f = 0.001, 1000.0                                                # :147
    t = n * math.tan(0.5 * fovy)
    r = n * math.tan(0.5 * fovx)
    P = np.array([[2 * n / (2 * r), 0.0, 0.0, 0.0],
                  [0.0, 2 * n / (2 * t), 0.0, 0.0],
                  [0.0, 0.0, (f + n) / (f - n), -1.0 * f * n / (f - n)],
                  [0.0, 0.0, 1.0, 0.0]], dtype=np.float32)
    PV = (P @ viewmat).astype(np.float32)
    cam = _lib.GsCamera()
    cam.viewmat[:] = viewmat[:3].reshape(-1).tolist()
    cam.projmat[:] = PV.reshape(-1).tolist()
    cam.cam_pos[:] = c2w[:3, 3].tolist()
    cam.fx, cam.fy, cam.cx, cam.cy = float(camera.fx), float(camera.fy), float(camera.cx), float(camera.cy)
    cam.glob_scale, cam.clip = float(glob_scale), float(clip)
    cam.H, cam.W = H, W
    cam.tiles_x, cam.tiles_y = (W + 15) // 16, (H + 15) // 16
    if len(_cam_cache) > 4096:
        _cam_cache.clear()
    _cam_cache[key] = (camera, cam)
    return cam


def _get_scratch(dev, nbytes):
    """Grow-only scratch arena per (device, stream): views rendered concurrently on different streams must
    not share radix ping-pong buffers."""
    key = (dev.index, torch.cuda.current_stream(dev).cuda_stream)
    buf = _scratch.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(int(nbytes * 1.25) + 256, device=dev, dtype=torch.uint8)
        _scratch[key] = buf
    return buf


_side = {}
# Opt-in (GS_SPLIT_COLOR=1): run the SH colour half of the preprocess on a side stream underneath the binning chain.
# Measured SLOWER (0.787 vs 0.772 ms/step): the colour kernel's 7 800 CTAs flood the SMs and the short binning kernels
# queue behind them (no stream below the default priority exists to demote it).
SPLIT_COLOR = _os.environ.get("GS_SPLIT_COLOR") is not None


def _get_side_stream(dev, cur):
    """(side stream, geometry-done event, colour-done event) paired with the stream `cur`; the events come from a
    small ring because several views may be in flight on one stream."""
    key = (dev.index, cur.cuda_stream)
    ent = _side.get(key)
    if ent is None:
        ent = {"stream": torch.cuda.Stream(device=dev),
               "events": [(torch.cuda.Event(), torch.cuda.Event()) for _ in range(8)], "next": 0}
        _side[key] = ent
    a, b = ent["events"][ent["next"]]
    ent["next"] = (ent["next"] + 1) % len(ent["events"])
    return ent["stream"], a, b


_RING = 16


def _get_pinned(dev):
    """Next (pinned int32[1], event) slot of a small per-device ring: several views may be in flight
    before the host looks at their intersection counts."""
    ring = _pinned.get(dev.index)
    if ring is None:
        ring = {"slots": [(torch.zeros(1, dtype=torch.int32).pin_memory(), torch.cuda.Event()) for _ in range(_RING)],
                "next": 0}
        _pinned[dev.index] = ring
    slot = ring["slots"][ring["next"]]
    ring["next"] = (ring["next"] + 1) % _RING
    return slot


class CapacityOverflow(RuntimeError):
    """Raised by the deferred check of a speculative (non-strict) forward: the tile lists did not fit the
    speculative capacity, so that view's image/gradients are invalid. The capacity hint has already been
    raised: re-running the step succeeds."""


def _check_pending(pending):
    pinned, ev, cap, key = pending
    ev.synchronize()
    I = int(pinned[0])
    if I > cap:
        _cap_hint[key] = int(I * 1.5) + 1024
        raise CapacityOverflow(f"{I} intersections exceed the speculative capacity {cap}; re-run this view")
    _cap_hint[key] = max(_cap_hint.get(key, 0), int(I * 1.5) + 1024)
    return I


class ViewState:
    """Per-view by-products a caller may want (gc_model.py keeps self.xys / self.radii)."""
    __slots__ = ("xys", "radii", "depths", "num_intersects", "xys_grad", "flat_grads", "defer_leaf_grads",
                 "view_block", "view_meta", "view_block_buffer")

    def __init__(self):
        self.xys = self.radii = self.depths = self.xys_grad = self.flat_grads = None
        self.num_intersects = 0
        # data-parallel mode (dist.reduce_view_grads): backward stops at the compositor-level gradients,
        # leaves them (+ the camera) in view_block and returns no leaf gradients
        self.defer_leaf_grads = False
        self.view_block = None
        self.view_meta = None
        # optional caller-owned buffer (e.g. NVLink symmetric memory) the view block is accumulated into
        self.view_block_buffer = None


def _forward_impl(means, scales, quats, opacities, features_dc, features_rest, cam, use_sh, deg, bg, strict=True,
                  with_depth=False, post=None):
    """strict=True: the host confirms the tile-list capacity before returning (one 4-byte read, overlapped
    with the compositing kernel; overflow -> redo). strict=False: nothing is read; the caller must run
    _check_pending(result[-1]) before trusting the view (render_view does it at the end of backward).
    post: optional dict; the compositing kernel then also writes the caller's per-pixel post-ops
    (gc_model.py:221-222) and the dict receives "rgb_clamped" = clamp(out, max=1) and "alpha" = 1 - final_T."""
    L = _lib.lib()
    dev = means.device
    N = means.shape[0]
    H, W, tx, ty = cam.H, cam.W, cam.tiles_x, cam.tiles_y
    T = tx * ty
    Np = (N + 3) // 4 * 4
    fbuf = torch.empty(10 * Np, device=dev, dtype=torch.float32)
    ibuf = torch.empty(2 * Np, device=dev, dtype=torch.int32)
    xys = fbuf[0:2 * N].view(N, 2)
    depths = fbuf[2 * Np:2 * Np + N]
    conics = fbuf[3 * Np:3 * Np + 3 * N].view(N, 3)
    rgbs = fbuf[6 * Np:6 * Np + 3 * N].view(N, 3)
    opac = fbuf[9 * Np:9 * Np + N]
    radii, hit = ibuf[0:N], ibuf[Np:Np + N]
    st = _lib.stream_ptr()
    color_done = None
    pp_args = (N, means.data_ptr(), scales.data_ptr(), quats.data_ptr(), opacities.data_ptr(),
               features_dc.data_ptr(), features_rest.data_ptr(), int(use_sh), int(deg),
               _ctypes_addr(cam), xys.data_ptr(), depths.data_ptr(), radii.data_ptr(), conics.data_ptr(),
               hit.data_ptr(), rgbs.data_ptr(), opac.data_ptr())
    if SPLIT_COLOR:
        # geometry on this stream (the binning chain needs nothing else); the SH colour pass — 216 of the 284 bytes
        # per Gaussian — on a side stream, underneath the latency-bound binning kernels; joined before compositing.
        # fbuf outlives the join, and everything later on this stream is ordered after it: no record_stream needed.
        rc = L.gs_preprocess_fwd_part(1, *pp_args, st)
        _lib.check(rc, "gs_preprocess_fwd_part(geometry)")
        cur = torch.cuda.current_stream(dev)
        side, geom_done, color_done = _get_side_stream(dev, cur)
        geom_done.record(cur)
        side.wait_event(geom_done)
        rc = L.gs_preprocess_fwd_part(2, *pp_args, side.cuda_stream)
        _lib.check(rc, "gs_preprocess_fwd_part(colour)")
        color_done.record(side)
    else:
        rc = L.gs_preprocess_fwd(*pp_args, st)
        _lib.check(rc, "gs_preprocess_fwd")
    key = (dev.index, N, T)
    known = key in _cap_hint
    cap = _cap_hint.get(key, max(4 * N, 1024))
    strict = strict or not known           # never speculate blindly on the first view of a shape
    bins = torch.empty((T, 2), device=dev, dtype=torch.int32)
    total = torch.empty(1, device=dev, dtype=torch.int32)
    out = torch.empty((H, W, 3), device=dev, dtype=torch.float32)
    fT = torch.empty((H, W), device=dev, dtype=torch.float32)
    fidx = torch.empty((H, W), device=dev, dtype=torch.int32)
    out_depth = torch.empty((H, W), device=dev, dtype=torch.float32) if with_depth else None
    order_f = torch.empty((T,), device=dev, dtype=torch.int32)
    tile_work = torch.empty((T,), device=dev, dtype=torch.int32)
    out_c = alpha = None
    if post is not None:
        out_c = torch.empty((H, W, 3), device=dev, dtype=torch.float32)
        alpha = torch.empty((H, W), device=dev, dtype=torch.float32)
        post["rgb_clamped"], post["alpha"] = out_c, alpha
    while True:
        ids_sorted = torch.empty((cap,), device=dev, dtype=torch.int32)
        if BIN_IMPL == "tile":
            scratch = _get_scratch(dev, L.gs_bin_tile_scratch_bytes(N, cap, T))
            rc = L.gs_bin_tile(N, cap, xys.data_ptr(), depths.data_ptr(), radii.data_ptr(), tx, ty,
                               ids_sorted.data_ptr(), bins.data_ptr(), total.data_ptr(), scratch.data_ptr(),
                               scratch.numel(), st)
        else:
            scratch = _get_scratch(dev, L.gs_bin_fast_scratch_bytes(N, cap, T))
            rc = L.gs_bin_fast(N, cap, xys.data_ptr(), depths.data_ptr(), radii.data_ptr(), hit.data_ptr(), tx, ty,
                               ids_sorted.data_ptr(), bins.data_ptr(), total.data_ptr(), scratch.data_ptr(),
                               scratch.numel(), st)
        _lib.check(rc, "gs_bin_tile/fast")
        pinned, ev = _get_pinned(dev)
        pinned.copy_(total, non_blocking=True)
        ev.record()
        # compositing is enqueued BEFORE the host looks at the count: it only reads tile_bins
        # tiles are launched heaviest-first (list length as the forward's work proxy); the forward records
        # each tile's true depth of composition, which orders the backward
        rc = L.gs_tile_order(T, bins.data_ptr(), 1, order_f.data_ptr(), st)
        _lib.check(rc, "gs_tile_order")
        if color_done is not None:
            torch.cuda.current_stream(dev).wait_event(color_done)
            color_done = None
        rc = L.gs_raster_fwd_ordered_post(H, W, tx, ty, ids_sorted.data_ptr(), bins.data_ptr(), xys.data_ptr(),
                                          conics.data_ptr(), rgbs.data_ptr(), opac.data_ptr(),
                                          depths.data_ptr() if with_depth else None, bg.data_ptr(), out.data_ptr(),
                                          out_depth.data_ptr() if with_depth else None, fT.data_ptr(),
                                          fidx.data_ptr(), order_f.data_ptr(), tile_work.data_ptr(),
                                          _lib.ptr(out_c), _lib.ptr(alpha), st)
        _lib.check(rc, "gs_raster_fwd")
        if not strict:
            return (xys, depths, conics, rgbs, opac, radii, hit, ids_sorted, bins, out, fT, fidx, None,
                    (pinned, ev, cap, key), tile_work)
        ev.synchronize()
        I = int(pinned[0])
        if I <= cap:
            _cap_hint[key] = max(int(I * 1.5) + 1024, 1024)
            break
        cap = int(I * 1.5) + 1024      # capacity was too small: nothing was composited from a torn list; redo
    if with_depth:
        return (xys, depths, conics, rgbs, opac, radii, hit, ids_sorted, bins, out, fT, fidx, I, out_depth, tile_work)
    return (xys, depths, conics, rgbs, opac, radii, hit, ids_sorted, bins, out, fT, fidx, I, None, tile_work)


def _ctypes_addr(struct):
    import ctypes
    return ctypes.addressof(struct)


class _RenderView(Function):
    @staticmethod
    def forward(ctx, means, scales, quats, opacities, features_dc, features_rest, cam, use_sh, deg, bg, state,
                clamp_rgb=False):
        for t, shape in ((means, 3), (scales, 3), (quats, 4), (features_dc, 3)):
            if t.ndim != 2 or t.shape[1] != shape or t.shape[0] != means.shape[0]:
                raise ValueError("means/scales (N,3), quats (N,4), features_dc (N,3) expected")
        if not means.is_cuda:
            raise ValueError("render_view needs CUDA tensors (no CPU fallback)")
        if features_rest.shape[1:] != (15, 3):
            raise ValueError("features_rest must have dimensions (N, 15, 3)")
        args = [t.detach().float().contiguous() for t in (means, scales, quats, opacities, features_dc, features_rest)]
        bg = bg.detach().to(means.device).float().contiguous()
        post = {}
        with torch.cuda.device(means.device):
            (xys, depths, conics, rgbs, opac, radii, hit, ids_sorted, bins, out, fT, fidx, I, pending,
             tile_work) = _forward_impl(*args, cam, use_sh, deg, bg, strict=not SPECULATE, post=post)
        ctx.pending = pending
        # clamp_rgb: the returned colour is clamp(out, max=1) (written by the compositing kernel) and the raw
        # colour is kept so that the backward kernel can mask the incoming gradient like torch.clamp's backward
        ctx.save_for_backward(*args, bg, xys, conics, rgbs, opac, radii, ids_sorted, bins, fT, fidx, tile_work,
                              out if clamp_rgb else None)
        ctx.cam, ctx.use_sh, ctx.deg, ctx.state = cam, use_sh, deg, state
        if state is not None:
            state.xys, state.radii, state.depths, state.num_intersects = xys, radii, depths, I
        ctx.mark_non_differentiable(depths, radii, hit)
        ctx.set_materialize_grads(False)      # unused outputs arrive as None instead of freshly filled zeros
        return (post["rgb_clamped"] if clamp_rgb else out), post["alpha"], depths, radii, hit

    @staticmethod
    def backward(ctx, v_out, v_alpha, *_unused):
        (means, scales, quats, opacities, fdc, frest, bg, xys, conics, rgbs, opac, radii, ids_sorted, bins, fT,
         fidx, tile_work, clamp_ref) = ctx.saved_tensors
        cam = ctx.cam
        L = _lib.lib()
        dev = means.device
        N = means.shape[0]
        H, W, tx, ty = cam.H, cam.W, cam.tiles_x, cam.tiles_y
        if v_out is None:                       # only alpha was used downstream
            v_out = torch.zeros((H, W, 3), device=dev, dtype=torch.float32)
        v_out = v_out.float().contiguous()
        va_ptr = None
        if v_alpha is not None:
            v_alpha = v_alpha.float().contiguous()
            va_ptr = v_alpha.data_ptr()
        Np = (N + 3) // 4 * 4
        with torch.cuda.device(dev):
            st = _lib.stream_ptr()
            ext = ctx.state.view_block_buffer if (ctx.state is not None and ctx.state.defer_leaf_grads) else None
            if ext is not None:
                assert ext.numel() == 9 * Np + 64 and ext.is_contiguous() and ext.dtype == torch.float32
                g = ext
                g.zero_()
            else:
                g = torch.zeros(9 * Np + 64, device=dev, dtype=torch.float32)    # accumulated into by atomics
            v_xys = g[0:2 * N].view(N, 2)
            v_conics = g[2 * Np:2 * Np + 3 * N].view(N, 3)
            v_rgbs = g[5 * Np:5 * Np + 3 * N].view(N, 3)
            v_opac = g[8 * Np:8 * Np + N]
            order_b = torch.empty_like(tile_work)
            rc = L.gs_tile_order(tx * ty, tile_work.data_ptr(), 0, order_b.data_ptr(), st)
            _lib.check(rc, "gs_tile_order")
            rc = L.gs_raster_bwd_ordered_post(H, W, tx, ty, ids_sorted.data_ptr(), bins.data_ptr(), xys.data_ptr(),
                                              conics.data_ptr(), rgbs.data_ptr(), opac.data_ptr(), bg.data_ptr(),
                                              fT.data_ptr(), fidx.data_ptr(), v_out.data_ptr(), va_ptr,
                                              v_xys.data_ptr(), v_conics.data_ptr(), v_rgbs.data_ptr(),
                                              v_opac.data_ptr(), order_b.data_ptr(), _lib.ptr(clamp_ref), st)
            _lib.check(rc, "gs_raster_bwd")
            if ctx.state is not None and ctx.state.defer_leaf_grads:
                rc = L.gs_finish_view_block(N, int(ctx.use_sh), _ctypes_addr(cam), rgbs.data_ptr(), g.data_ptr(), st)
                _lib.check(rc, "gs_finish_view_block")
                if ctx.pending is not None:
                    ctx.state.num_intersects = _check_pending(ctx.pending)
                ctx.state.xys_grad = v_xys
                ctx.state.view_block = g
                ctx.state.view_meta = (int(ctx.use_sh), int(ctx.deg), tuple(opacities.shape))
                return (None,) * 12
            # the six leaf gradients live in ONE flat buffer (59 floats / Gaussian) so that the
            # data-parallel step can all-reduce them with a single NCCL call
            flat = torch.empty(59 * Np, device=dev, dtype=torch.float32)
            v_means = flat[0:3 * N].view(N, 3)
            v_scales = flat[3 * Np:3 * Np + 3 * N].view(N, 3)
            v_quats = flat[6 * Np:6 * Np + 4 * N].view(N, 4)
            v_opl = flat[10 * Np:10 * Np + N].view(opacities.shape)
            v_dc = flat[11 * Np:11 * Np + 3 * N].view(N, 3)
            v_rest = flat[14 * Np:14 * Np + 45 * N].view(N, 15, 3)
            rc = L.gs_preprocess_bwd(N, means.data_ptr(), scales.data_ptr(), quats.data_ptr(), opacities.data_ptr(),
                                     int(ctx.use_sh), int(ctx.deg), _ctypes_addr(cam), radii.data_ptr(), rgbs.data_ptr(),
                                     v_xys.data_ptr(), v_conics.data_ptr(), v_rgbs.data_ptr(), v_opac.data_ptr(),
                                     v_means.data_ptr(), v_scales.data_ptr(), v_quats.data_ptr(), v_opl.data_ptr(),
                                     v_dc.data_ptr(), v_rest.data_ptr(), st)
            _lib.check(rc, "gs_preprocess_bwd")
        if ctx.pending is not None:
            I = _check_pending(ctx.pending)      # long since complete: does not stall the device
            if ctx.state is not None:
                ctx.state.num_intersects = I
        if ctx.state is not None:
            ctx.state.xys_grad = v_xys       # what gc_model.py:193-194 keeps via xys.retain_grad()
            ctx.state.flat_grads = flat
        return v_means, v_scales, v_quats, v_opl, v_dc, v_rest, None, None, None, None, None, None


def render_view(means: Tensor, scales: Tensor, quats: Tensor, opacities: Tensor, features_dc: Tensor,
                features_rest: Tensor, camera: Camera, sh_degree_to_use: int, use_sh: bool = True,
                background: Optional[Tensor] = None, state: Optional[ViewState] = None, clamp_rgb: bool = False
                ) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor]:
    """RAW splatfacto leaves (log-scales, un-normalised wxyz quats, opacity logits (N,1),
    features_dc (N,3), features_rest (N,15,3)) + one camera -> (rgb (H,W,3) before the caller's
    clamp(max=1) — or after it with clamp_rgb=True: the clamp and its backward are then done inside the
    compositing kernels —, alpha (H,W), depths (N), radii (N) i32, num_tiles_hit (N) i32)."""
    if background is None:
        background = torch.ones(3, dtype=torch.float32, device=means.device)
    cam = gs_camera(camera)
    if means.shape[0] == 0:
        # empty model: background image, zero alpha, empty per-Gaussian outputs; stays on the autograd graph so that
        # backward hands (empty) gradients to the six leaves
        dev = means.device
        zero = (means.sum() + scales.sum() + quats.sum() + opacities.sum() + features_dc.sum() + features_rest.sum())
        rgb = background.to(dev).float().expand(cam.H, cam.W, 3) + zero
        if clamp_rgb:
            rgb = torch.clamp(rgb, max=1.0)
        if state is not None:
            state.xys = torch.zeros((0, 2), device=dev)
            state.radii = torch.zeros((0,), device=dev, dtype=torch.int32)
            state.depths = torch.zeros((0,), device=dev)
            state.num_intersects = 0
        return (rgb, torch.zeros((cam.H, cam.W), device=dev) + zero, torch.zeros((0,), device=dev),
                torch.zeros((0,), device=dev, dtype=torch.int32), torch.zeros((0,), device=dev, dtype=torch.int32))
    return _RenderView.apply(means, scales, quats, opacities, features_dc, features_rest, cam, bool(use_sh),
                             int(sh_degree_to_use), background, state, bool(clamp_rgb))


def composite_extra(state_tensors, colors: Tensor, camera: Camera, background: Tensor) -> Tensor:
    """Second compositing pass over the SAME binning (the eval depth pass of gc_model.py:225-236
    re-bins and re-sorts in gsplat; here the tile lists of the RGB pass are reused).
    state_tensors = (xys, conics, opac, ids_sorted, bins); colors (N,C) -> (H,W,C). No autograd."""
    xys, conics, opac, ids_sorted, bins = state_tensors
    cam = gs_camera(camera)
    dev = xys.device
    C = colors.shape[1]
    colors = colors.float().contiguous()
    bg = background.to(dev).float().contiguous()
    out = torch.empty((cam.H, cam.W, C), device=dev, dtype=torch.float32)
    fT = torch.empty((cam.H, cam.W), device=dev, dtype=torch.float32)
    fidx = torch.empty((cam.H, cam.W), device=dev, dtype=torch.int32)
    with torch.cuda.device(dev):
        rc = _lib.lib().gs_raster_fwd(C, cam.H, cam.W, cam.tiles_x, cam.tiles_y, ids_sorted.data_ptr(), bins.data_ptr(),
                                      xys.data_ptr(), conics.data_ptr(), colors.data_ptr(), opac.data_ptr(),
                                      bg.data_ptr(), out.data_ptr(), fT.data_ptr(), fidx.data_ptr(), _lib.stream_ptr())
    _lib.check(rc, "gs_raster_fwd")
    return out


def leaf_grads_from_view_blocks(params, blocks, num_views: int, use_sh: bool, deg: int, scale: float = 1.0,
                                peer_ptrs_dev: Optional[int] = None, peer_offset: int = 0, sparse: bool = False):
    """gs_preprocess_bwd_multi: `blocks` = num_views view blocks back to back (as all-gathered), or — with
    peer_ptrs_dev (device array of per-rank base pointers, symmetric memory) — the peers' blocks read in place
    over NVLink (gs_preprocess_bwd_multi_peer). Returns (flat, dict of the six leaf gradients as views of flat),
    gradients = scale * sum over the views."""
    means, scales, quats, opacities = params["means"], params["scales"], params["quats"], params["opacities"]
    dev = means.device
    N = means.shape[0]
    Np = (N + 3) // 4 * 4
    L = _lib.lib()
    stride = L.gs_view_block_floats(N)
    if peer_ptrs_dev is None:
        assert blocks.numel() == num_views * stride and blocks.is_contiguous()
    flat = torch.empty(59 * Np, device=dev, dtype=torch.float32)
    out = {"means": flat[0:3 * N].view(N, 3), "scales": flat[3 * Np:3 * Np + 3 * N].view(N, 3),
           "quats": flat[6 * Np:6 * Np + 4 * N].view(N, 4), "opacities": flat[10 * Np:10 * Np + N].view(opacities.shape),
           "features_dc": flat[11 * Np:11 * Np + 3 * N].view(N, 3),
           "features_rest": flat[14 * Np:14 * Np + 45 * N].view(N, 15, 3)}
    with torch.cuda.device(dev):
        if peer_ptrs_dev is not None:
            fn = L.gs_preprocess_bwd_multi_peer_sparse if sparse else L.gs_preprocess_bwd_multi_peer
            rc = fn(
                N, int(num_views), int(peer_ptrs_dev), int(peer_offset), float(scale), means.detach().data_ptr(),
                scales.detach().data_ptr(), quats.detach().data_ptr(), opacities.detach().data_ptr(), int(use_sh), int(deg),
                out["means"].data_ptr(), out["scales"].data_ptr(), out["quats"].data_ptr(), out["opacities"].data_ptr(),
                out["features_dc"].data_ptr(), out["features_rest"].data_ptr(), _lib.stream_ptr())
            _lib.check(rc, "gs_preprocess_bwd_multi_peer")
            return flat, out
        rc = L.gs_preprocess_bwd_multi(
            N, int(num_views), blocks.data_ptr(), stride, float(scale), means.detach().data_ptr(),
            scales.detach().data_ptr(), quats.detach().data_ptr(), opacities.detach().data_ptr(), int(use_sh), int(deg),
            out["means"].data_ptr(), out["scales"].data_ptr(), out["quats"].data_ptr(), out["opacities"].data_ptr(),
            out["features_dc"].data_ptr(), out["features_rest"].data_ptr(), _lib.stream_ptr())
    _lib.check(rc, "gs_preprocess_bwd_multi")
    return flat, out
