"""Fused per-view render: ONE autograd node from the six splatfacto leaf tensors to (rgb, alpha), and ONE C-ABI call
per direction.

Same result as the three-operator sequence of /root/reference/gaussctrl/gc_model.py:172-222
(pre-activations -> project_gaussians -> spherical_harmonics -> rasterize_gaussians) and its
backward (gc_trainer.py:275):
    forward : gs_view_fwd  = preprocess (-> packed splat records) -> binning -> tile order -> compositing
    backward: gs_view_bwd  = tile order -> compositing backward -> leaf backward
Each is a single ctypes call that enqueues the whole kernel chain from C (csrc/view.cu). Everything a view keeps
between forward and backward lives in ONE workspace tensor whose layout the library reports (gs_view_layout).

Tile-list capacity. ids_sorted is allocated for a capacity learnt from earlier views of the same shape; the true
intersection count is copied to a pinned host word right behind the binning. By default the host reads that word after
it has enqueued the compositing kernel (the wait ends when the BINNING is done, while compositing still runs, so the
device never idles) and redoes the view if the capacity was too small: a caller never sees a torn image.
`speculate=True` skips even that wait; an overflow then shows as a NaN image (the compositing kernel poisons it) and
as CapacityOverflow at the end of backward.
"""
from __future__ import annotations

import ctypes
import math
import os as _os
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch
from torch import Tensor
from torch.autograd import Function

from . import _lib
from .scene import Camera

# Fully asynchronous training forward (see module docstring). Off by default: correctness first.
# debug switch (GS_DETERMINISTIC=1 or render.DETERMINISTIC = True): training views run the bit-reproducible backward
DETERMINISTIC = _os.environ.get("GS_DETERMINISTIC", "") not in ("", "0")
SPECULATE = _os.environ.get("GS_SPECULATE") is not None

_cam_cache = {}
_cap_hint = {}
_scratch = {}
_pinned_free = {}
_layout_cache = {}
_scratch_bytes_cache = {}


def gs_camera(camera: Camera, glob_scale: float = 1.0, clip: float = 0.01) -> _lib.GsCamera:
    """Host-side matrices of gc_model.py:117-155 packed into the C struct. Cached by VALUE (pose + intrinsics):
    an in-place update of the camera (pose optimiser, rescale_output_resolution) can never hit a stale entry."""
    key = (camera.signature(), float(glob_scale), float(clip))
    hit = _cam_cache.get(key)
    if hit is not None:
        return hit
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
    _cam_cache[key] = cam
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


class _Pending:
    """One in-flight intersection count: its OWN pinned word + event (never shared between views)."""
    __slots__ = ("pinned", "event", "cap", "key", "dev")

    def __init__(self, dev):
        free = _pinned_free.setdefault(dev.index, [])
        if free:
            self.pinned, self.event = free.pop()
        else:
            self.pinned, self.event = torch.zeros(1, dtype=torch.int32).pin_memory(), torch.cuda.Event()
            self.event.record(torch.cuda.current_stream(dev))    # creates the cudaEvent_t the library records later
        self.dev, self.cap, self.key = dev, 0, None

    def wait(self) -> int:
        """Blocks until the count has landed; returns it and recycles the slot."""
        self.event.synchronize()
        I = ctypes.c_int32.from_address(self.pinned.data_ptr()).value
        _pinned_free[self.dev.index].append((self.pinned, self.event))
        self.pinned = self.event = None
        return _lib.check_count(I, "gs_view_fwd")


class CapacityOverflow(RuntimeError):
    """speculate=True only: the tile lists of a view did not fit the speculative capacity. The image the forward
    returned was NaN-poisoned; no gradient is handed out. The capacity hint has been raised: re-running succeeds."""


_CAP_STEP = 1 << 20


def _round_cap(c: int) -> int:
    """Capacities move on a 1 Mi-entry grid: the workspace size (and with it the caching allocator's block) changes
    rarely instead of with every new maximum."""
    return (int(c) + _CAP_STEP - 1) // _CAP_STEP * _CAP_STEP


def _raise_hint(key, I):
    _cap_hint[key] = max(_cap_hint.get(key, 0), _round_cap(int(I * 1.5) + 1024))      # monotone: never shrinks


class ViewWorkspace:
    """Everything one forward leaves behind for the backward and for the caller (gc_model.py keeps self.xys /
    self.radii): one uint8 tensor + typed views of it, following gs_view_layout."""

    def __init__(self, dev, n: int, views: int, cap: int, T_view: int):
        lay = _layout_cache.get((n, views, cap, T_view))
        if lay is None:
            lay = _lib.GsViewLayout()
            _lib.check(_lib.lib().gs_view_layout(n, views, cap, T_view, ctypes.byref(lay)), "gs_view_layout")
            if len(_layout_cache) > 256:
                _layout_cache.clear()
            _layout_cache[(n, views, cap, T_view)] = lay
        self.n, self.views, self.cap, self.T_view, self.lay = n, views, cap, T_view, lay
        self.buf = torch.empty(lay.bytes, device=dev, dtype=torch.uint8)

    def _view(self, off, count, dtype):
        nbytes = count * 4
        return self.buf[off:off + nbytes].view(dtype)

    @property
    def rec(self):                      # (views*n, 12) packed splat records
        return self._view(self.lay.rec, self.n * self.views * 12, torch.float32).view(-1, 12)

    @property
    def xys(self):                      # strided views into the records: no copies
        return self.rec[:, 0:2]

    @property
    def opac(self):
        return self.rec[:, 4]

    @property
    def rgbs(self):
        return self.rec[:, 8:11]

    @property
    def depths(self):
        return self._view(self.lay.depths, self.n * self.views, torch.float32)

    @property
    def radii(self):
        return self._view(self.lay.radii, self.n * self.views, torch.int32)

    @property
    def num_tiles_hit(self):
        return self._view(self.lay.tiles_hit, self.n * self.views, torch.int32)

    @property
    def ids_sorted(self):
        return self._view(self.lay.ids_sorted, self.cap, torch.int32)

    @property
    def tile_bins(self):
        return self._view(self.lay.tile_bins, self.T_view * self.views * 2, torch.int32).view(-1, 2)

    @property
    def tile_work(self):
        return self._view(self.lay.tile_work, self.T_view * self.views, torch.int32)

    @property
    def total(self):
        return self._view(self.lay.total, 1, torch.int32)

    def conics(self):
        """True conics (a, b, c) recovered from the records' pre-scaled form (diagnostics / tests)."""
        q = self.rec[:, 5:8]
        k = 1.0 / 1.4426950408889634
        return torch.stack((q[:, 0] * (-2.0 * k), q[:, 1] * (-1.0 * k), q[:, 2] * (-2.0 * k)), dim=1)


class ViewState:
    """Per-view by-products a caller may want (gc_model.py keeps self.xys / self.radii). The per-Gaussian and tile
    arrays are lazy views into the last view's workspace (no tensor is created unless somebody asks)."""
    __slots__ = ("num_intersects", "xys_grad", "flat_grads", "defer_leaf_grads", "view_block", "view_meta",
                 "view_block_buffer", "final_idx", "final_T", "workspace", "_empty", "sharded_grads", "shard", "verified")

    def __init__(self):
        self.xys_grad = self.flat_grads = None
        # owner-sliced data-parallel mode (dist.PeerViewExchange mode="slice"): gradients of this rank's Gaussians
        # [shard[0], shard[1]) summed over all ranks' views, slice-shaped
        self.sharded_grads = self.shard = None
        # False between a speculative forward (nothing read back) and the check at the end of its backward: until
        # then the image and everything derived from it are UNVERIFIED (an overflow NaN-poisons them)
        self.verified = True
        # integer by-products of the last view (what gsplat's rasterize keeps for its backward); the parity
        # tests compare them bit for bit with the oracle
        self.final_idx = self.final_T = self.workspace = self._empty = None
        self.num_intersects = 0
        # data-parallel mode (dist.reduce_view_grads): backward stops at the compositor-level gradients,
        # leaves them (+ the camera) in view_block and returns no leaf gradients
        self.defer_leaf_grads = False
        self.view_block = None
        self.view_meta = None
        # optional caller-owned buffer (e.g. NVLink symmetric memory) the view block is accumulated into
        self.view_block_buffer = None

    def _ws(self, name):
        if self.workspace is not None:
            return getattr(self.workspace, name)
        return None if self._empty is None else self._empty[name]

    xys = property(lambda self: self._ws("xys"))
    radii = property(lambda self: self._ws("radii"))
    depths = property(lambda self: self._ws("depths"))
    num_tiles_hit = property(lambda self: self._ws("num_tiles_hit"))
    ids_sorted = property(lambda self: self._ws("ids_sorted"))
    tile_bins = property(lambda self: self._ws("tile_bins"))


class ViewImages:
    """Image-shaped outputs of one forward, carved from ONE allocation."""

    def __init__(self, dev, views: int, H: int, W: int, with_depth: bool):
        hw = views * H * W
        self.buf = torch.empty((10 if with_depth else 9) * hw, device=dev, dtype=torch.float32)
        b = self.buf
        shape3 = (views, H, W, 3) if views > 1 else (H, W, 3)
        shape1 = (views, H, W) if views > 1 else (H, W)
        self.out = b[0:3 * hw].view(shape3)
        self.out_clamped = b[3 * hw:6 * hw].view(shape3)
        self.alpha = b[6 * hw:7 * hw].view(shape1)
        self.final_T = b[7 * hw:8 * hw].view(shape1)
        self.final_idx = b[8 * hw:9 * hw].view(torch.int32).view(shape1)
        self.depth = b[9 * hw:10 * hw].view(shape1) if with_depth else None


def forward_views(means, scales, quats, opacities, features_dc, features_rest, cams: Sequence[_lib.GsCamera],
                  use_sh: bool, deg: int, bg: Tensor, with_depth: bool = False, subset: Optional[Tensor] = None,
                  speculate: bool = False) -> Tuple[ViewWorkspace, ViewImages, Optional[int], Optional[_Pending]]:
    """gs_view_fwd for `len(cams)` cameras of one image size over the same Gaussians (detached, contiguous fp32
    CUDA tensors). Returns (workspace, images, intersection count | None, pending | None): with speculate=False the
    count has been verified (the view was redone with a larger capacity if necessary); with speculate=True nothing
    was read and the caller must call check_pending(pending) before trusting the view."""
    L = _lib.lib()
    dev = means.device
    n = int(subset.shape[0]) if subset is not None else means.shape[0]
    views = len(cams)
    c0 = cams[0]
    H, W, T_view = c0.H, c0.W, c0.tiles_x * c0.tiles_y
    cam_arr = (_lib.GsCamera * views)(*cams) if views > 1 else c0
    key = (dev.index, n, views, T_view)
    known = key in _cap_hint
    cap = _cap_hint.get(key, _round_cap(max(4 * n * views, 1024)))
    speculate = speculate and known           # never speculate blindly on the first view of a shape
    st = _lib.stream_ptr()
    while True:
        ws = ViewWorkspace(dev, n, views, cap, T_view)
        im = ViewImages(dev, views, H, W, with_depth)
        sb = _scratch_bytes_cache.get((n, views, cap, T_view))
        if sb is None:
            sb = _scratch_bytes_cache[(n, views, cap, T_view)] = L.gs_view_scratch_bytes(n, views, cap, T_view)
        scratch = _get_scratch(dev, sb)
        pend = _Pending(dev)
        pend.cap, pend.key = cap, key
        rc = L.gs_view_fwd(n, views, means.data_ptr(), scales.data_ptr(), quats.data_ptr(), opacities.data_ptr(),
                           features_dc.data_ptr(), features_rest.data_ptr(), _lib.ptr(subset), int(use_sh), int(deg),
                           ctypes.addressof(cam_arr), bg.data_ptr(), cap, ws.buf.data_ptr(), ws.buf.numel(),
                           scratch.data_ptr(), scratch.numel(), im.out.data_ptr(), im.out_clamped.data_ptr(),
                           im.alpha.data_ptr(), _lib.ptr(im.depth), im.final_T.data_ptr(), im.final_idx.data_ptr(),
                           pend.pinned.data_ptr(), pend.event.cuda_event, st)
        _lib.check(rc, "gs_view_fwd")
        # the library recorded the event right behind the count copy, BEFORE it enqueued the compositing kernel:
        # waiting here ends when the binning is done, while the compositing still runs
        if speculate:
            return ws, im, None, pend
        I = pend.wait()
        if I <= cap:
            _raise_hint(key, I)
            return ws, im, I, None
        cap = _round_cap(int(I * 1.5) + 1024)       # capacity was too small: the binning wrote nothing; redo with room


def check_pending(pend: _Pending) -> int:
    cap, key = pend.cap, pend.key
    I = pend.wait()
    _raise_hint(key, I)
    if I > cap:
        raise CapacityOverflow(f"{I} intersections exceed the speculative capacity {cap}; re-run this view")
    return I


def backward_view(means, scales, quats, opacities, cam: _lib.GsCamera, use_sh: bool, deg: int, bg: Tensor,
                  ws: ViewWorkspace, final_T: Tensor, final_idx: Tensor, v_out: Tensor, v_alpha: Optional[Tensor],
                  clamp_ref: Optional[Tensor], gbuf: Optional[Tensor] = None, leaf_grads: bool = True):
    """gs_view_bwd. Returns (gbuf = compositor-level gradient block, flat = 59*Np leaf-gradient buffer | None)."""
    L = _lib.lib()
    dev = means.device
    n = ws.n
    Np = (n + 3) // 4 * 4
    if gbuf is None:
        gbuf = torch.empty(9 * Np + 64, device=dev, dtype=torch.float32)      # zero-filled by the library
    flat = torch.empty(59 * Np, device=dev, dtype=torch.float32) if leaf_grads else None
    if DETERMINISTIC:       # debug switch: bit-reproducible gradients (gs_view_bwd_det), slower
        det = torch.empty(L.gs_view_bwd_det_scratch_bytes(n, ws.cap), device=dev, dtype=torch.uint8)
        rc = L.gs_view_bwd_det(n, means.data_ptr(), scales.data_ptr(), quats.data_ptr(), opacities.data_ptr(), int(use_sh),
                               int(deg), ctypes.addressof(cam), bg.data_ptr(), ws.buf.data_ptr(), ws.buf.numel(), ws.cap,
                               final_T.data_ptr(), final_idx.data_ptr(), v_out.data_ptr(), _lib.ptr(v_alpha),
                               _lib.ptr(clamp_ref), gbuf.data_ptr(), _lib.ptr(flat), det.data_ptr(), det.numel(),
                               _lib.stream_ptr())
        _lib.check(rc, "gs_view_bwd_det")
        return gbuf, flat
    rc = L.gs_view_bwd(n, means.data_ptr(), scales.data_ptr(), quats.data_ptr(), opacities.data_ptr(), int(use_sh),
                       int(deg), ctypes.addressof(cam), bg.data_ptr(), ws.buf.data_ptr(), ws.buf.numel(), ws.cap,
                       final_T.data_ptr(), final_idx.data_ptr(), v_out.data_ptr(), _lib.ptr(v_alpha),
                       _lib.ptr(clamp_ref), gbuf.data_ptr(), _lib.ptr(flat), _lib.stream_ptr())
    _lib.check(rc, "gs_view_bwd")
    return gbuf, flat


def split_flat_grads(flat: Tensor, n: int, opac_shape):
    """The six leaf gradients as views of the flat 59*Np buffer (layout of gs_view_bwd)."""
    Np = (n + 3) // 4 * 4
    return (flat[0:3 * n].view(n, 3), flat[3 * Np:3 * Np + 3 * n].view(n, 3), flat[6 * Np:6 * Np + 4 * n].view(n, 4),
            flat[10 * Np:10 * Np + n].view(opac_shape), flat[11 * Np:11 * Np + 3 * n].view(n, 3),
            flat[14 * Np:14 * Np + 45 * n].view(n, 15, 3))


class _RenderView(Function):
    @staticmethod
    def forward(ctx, means, scales, quats, opacities, features_dc, features_rest, cam, use_sh, deg, bg, state,
                clamp_rgb=False, speculate=False):
        for t, shape in ((means, 3), (scales, 3), (quats, 4), (features_dc, 3)):
            if t.ndim != 2 or t.shape[1] != shape or t.shape[0] != means.shape[0]:
                raise ValueError("means/scales (N,3), quats (N,4), features_dc (N,3) expected")
        if not means.is_cuda:
            raise ValueError("render_view needs CUDA tensors (no CPU fallback)")
        if features_rest.shape[1:] != (15, 3):
            raise ValueError("features_rest must have dimensions (N, 15, 3)")
        args = [t.detach() if (t.dtype is torch.float32 and t.is_contiguous()) else t.detach().float().contiguous()
                for t in (means, scales, quats, opacities, features_dc, features_rest)]
        bg = bg.detach().to(means.device).float().contiguous()
        with torch.cuda.device(means.device):
            ws, im, I, pend = forward_views(*args, [cam], use_sh, deg, bg, speculate=speculate)
        ctx.pending = pend
        ctx.ws = ws
        # clamp_rgb: the returned colour is clamp(out, max=1) (written by the compositing kernel) and the raw
        # colour is kept so that the backward kernel can mask the incoming gradient like torch.clamp's backward.
        # NOTHING that is (or holds) a returned tensor may hang off ctx directly: output -> grad_fn -> ctx -> output
        # is a reference cycle that only the cyclic GC breaks, i.e. ~100 MB of workspace per view would pile up.
        ctx.clamp_rgb = clamp_rgb
        ctx.save_for_backward(*args[:4], bg, im.final_T, im.final_idx, im.out if clamp_rgb else None)
        ctx.cam, ctx.use_sh, ctx.deg, ctx.state = cam, use_sh, deg, state
        ctx.opac_shape = tuple(opacities.shape)
        depths, radii, hit = ws.depths, ws.radii, ws.num_tiles_hit
        if state is not None:
            state.num_intersects, state.final_idx, state.final_T, state.workspace = I, im.final_idx, im.final_T, ws
            state.verified = pend is None
        ctx.mark_non_differentiable(depths, radii, hit)
        ctx.set_materialize_grads(False)      # unused outputs arrive as None instead of freshly filled zeros
        return (im.out_clamped if clamp_rgb else im.out), im.alpha, depths, radii, hit

    @staticmethod
    def backward(ctx, v_out, v_alpha, *_unused):
        means, scales, quats, opacities, bg, final_T, final_idx, clamp_ref = ctx.saved_tensors
        cam, ws = ctx.cam, ctx.ws
        dev = means.device
        n = means.shape[0]
        H, W = cam.H, cam.W
        if v_out is None:                       # only alpha was used downstream
            v_out = torch.zeros((H, W, 3), device=dev, dtype=torch.float32)
        v_out = v_out.float().contiguous()
        if v_alpha is not None:
            v_alpha = v_alpha.float().contiguous()
        defer = ctx.state is not None and ctx.state.defer_leaf_grads
        ext = ctx.state.view_block_buffer if defer else None
        Np = (n + 3) // 4 * 4
        if ext is not None:
            assert ext.numel() == 9 * Np + 64 and ext.is_contiguous() and ext.dtype == torch.float32
        with torch.cuda.device(dev):
            g, flat = backward_view(means, scales, quats, opacities, cam, ctx.use_sh, ctx.deg, bg, ws, final_T,
                                    final_idx, v_out, v_alpha, clamp_ref, gbuf=ext, leaf_grads=not defer)
        v_xys = g[0:2 * n].view(n, 2)
        if ctx.pending is not None:
            I = check_pending(ctx.pending)        # long since complete: does not stall the device
            ctx.pending = None
            if ctx.state is not None:
                ctx.state.num_intersects = I
                ctx.state.verified = True
        if ctx.state is not None:
            ctx.state.xys_grad = v_xys            # what gc_model.py:193-194 keeps via xys.retain_grad()
        if defer:
            ctx.state.view_block = g
            ctx.state.view_meta = (int(ctx.use_sh), int(ctx.deg), ctx.opac_shape)
            return (None,) * 13
        if ctx.state is not None:
            ctx.state.flat_grads = flat
        v_means, v_scales, v_quats, v_opl, v_dc, v_rest = split_flat_grads(flat, n, ctx.opac_shape)
        return v_means, v_scales, v_quats, v_opl, v_dc, v_rest, None, None, None, None, None, None, None


def render_view(means: Tensor, scales: Tensor, quats: Tensor, opacities: Tensor, features_dc: Tensor,
                features_rest: Tensor, camera: Camera, sh_degree_to_use: int, use_sh: bool = True,
                background: Optional[Tensor] = None, state: Optional[ViewState] = None, clamp_rgb: bool = False,
                speculate: Optional[bool] = None) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor]:
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
            state.workspace = None
            state._empty = {"xys": torch.zeros((0, 2), device=dev), "radii": torch.zeros((0,), device=dev, dtype=torch.int32),
                            "depths": torch.zeros((0,), device=dev), "num_tiles_hit": torch.zeros((0,), device=dev, dtype=torch.int32),
                            "ids_sorted": torch.zeros((0,), device=dev, dtype=torch.int32), "tile_bins": None}
            state.num_intersects = 0
        return (rgb, torch.zeros((cam.H, cam.W), device=dev) + zero, torch.zeros((0,), device=dev),
                torch.zeros((0,), device=dev, dtype=torch.int32), torch.zeros((0,), device=dev, dtype=torch.int32))
    return _RenderView.apply(means, scales, quats, opacities, features_dc, features_rest, cam, bool(use_sh),
                             int(sh_degree_to_use), background, state, bool(clamp_rgb),
                             SPECULATE if speculate is None else bool(speculate))


@torch.no_grad()
def render_eval(means, scales, quats, opacities, features_dc, features_rest, cameras: List[Camera],
                sh_degree_to_use: int, use_sh: bool, background: Tensor, with_depth: bool = True,
                subset: Optional[Tensor] = None):
    """No-grad render of a BATCH of cameras (one image size) in one call: RGB (clamped, gc_model.py:222),
    accumulation and — with_depth — the expected depth normalised by accumulation, 1000 where nothing was hit
    (gc_model.py:224-238), all from ONE binning and ONE compositing launch over cameras x tiles.
    Returns (workspace, rgb (V,H,W,3), accumulation (V,H,W,1), depth (V,H,W,1) | None) — V axis dropped for one camera."""
    dev = means.device
    cams = [gs_camera(c) for c in cameras]
    args = [t.detach().float().contiguous() for t in (means, scales, quats, opacities, features_dc, features_rest)]
    bg = background.detach().to(dev).float().contiguous()
    if subset is not None:
        subset = subset.to(device=dev, dtype=torch.int32).contiguous()
    with torch.cuda.device(dev):
        ws, im, I, _ = forward_views(*args, cams, use_sh, sh_degree_to_use, bg, with_depth=with_depth, subset=subset)
    ws.num_intersects = I                # host copy of the count the call already read (0 <=> every radius is 0)
    alpha = im.alpha[..., None]
    depth = None
    if with_depth:
        d = im.depth[..., None]
        depth = torch.where(alpha > 0, d / alpha, torch.full_like(d, 1000.0))
    return ws, im.out_clamped, alpha, depth


def composite_extra(state_tensors, colors: Tensor, camera: Camera, background: Tensor) -> Tensor:
    """Second compositing pass over the SAME binning (the eval depth pass of gc_model.py:225-236
    re-bins and re-sorts in gsplat; here the tile lists of the RGB pass are reused).
    state_tensors = (xys, conics, opac, ids_sorted, bins); colors (N,C) -> (H,W,C). No autograd."""
    xys, conics, opac, ids_sorted, bins = state_tensors
    cam = gs_camera(camera)
    dev = xys.device
    C = colors.shape[1]
    colors = colors.float().contiguous()
    xys, conics, opac = xys.float().contiguous(), conics.float().contiguous(), opac.float().contiguous()
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


def leaf_grads_slice(params, lo: int, hi: int, num_views: int, use_sh: bool, deg: int, scale: float, peer_ptrs_dev: int,
                     peer_offset: int):
    """gs_leaf_bwd_slice: leaf gradients of the Gaussians [lo, hi) summed over `num_views` compacted view blocks that sit
    in this rank's exchange slots. Returns (flat, dict of SLICE-shaped gradients: row 0 = Gaussian lo)."""
    means, scales, quats, opacities = params["means"], params["scales"], params["quats"], params["opacities"]
    dev = means.device
    N, n = means.shape[0], max(hi - lo, 0)
    npad = (n + 3) // 4 * 4
    L = _lib.lib()
    flat = torch.empty(59 * npad, device=dev, dtype=torch.float32)
    out = {"means": flat[0:3 * n].view(n, 3), "scales": flat[3 * npad:3 * npad + 3 * n].view(n, 3),
           "quats": flat[6 * npad:6 * npad + 4 * n].view(n, 4),
           "opacities": flat[10 * npad:10 * npad + n].view((n,) + tuple(opacities.shape[1:])),
           "features_dc": flat[11 * npad:11 * npad + 3 * n].view(n, 3),
           "features_rest": flat[14 * npad:14 * npad + 45 * n].view(n, 15, 3)}
    if n == 0:
        return flat, out
    with torch.cuda.device(dev):
        rc = L.gs_leaf_bwd_slice(
            N, int(lo), int(hi), int(num_views), int(peer_ptrs_dev), int(peer_offset), float(scale), means.detach().data_ptr(),
            scales.detach().data_ptr(), quats.detach().data_ptr(), opacities.detach().data_ptr(), int(use_sh), int(deg),
            out["means"].data_ptr(), out["scales"].data_ptr(), out["quats"].data_ptr(), out["opacities"].data_ptr(),
            out["features_dc"].data_ptr(), out["features_rest"].data_ptr(), _lib.stream_ptr())
    _lib.check(rc, "gs_leaf_bwd_slice")
    return flat, out


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
