"""`project_gaussians` with the gsplat-0.1.2 signature, backed by the sm_100a kernels.

Replaces the call at /root/reference/gaussctrl/gc_model.py:174-188 (import at :35).
Semantics: SURVEY.md App. A.1; backward = exact derivative (App. A.6 lists where the real
gsplat binary is believed to deviate).
"""
from __future__ import annotations

from typing import Tuple

import torch
from torch import Tensor
from torch.autograd import Function

from . import _lib


def _f32c(t: Tensor) -> Tensor:
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def project_gaussians(
    means3d: Tensor,
    scales: Tensor,
    glob_scale: float,
    quats: Tensor,
    viewmat: Tensor,
    projmat: Tensor,
    fx: float,
    fy: float,
    cx: float,
    cy: float,
    img_height: int,
    img_width: int,
    tile_bounds: Tuple[int, int, int],
    clip_thresh: float = 0.01,
) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor, Tensor]:
    """Returns (xys[N,2] f32, depths[N] f32, radii[N] i32, conics[N,3] f32,
    num_tiles_hit[N] i32, cov3d[N,6] f32). Differentiable w.r.t. means3d, scales, quats."""
    return _ProjectGaussians.apply(
        means3d, scales, glob_scale, quats, viewmat, projmat, fx, fy, cx, cy, img_height, img_width,
        tile_bounds, clip_thresh)


class _ProjectGaussians(Function):
    @staticmethod
    def forward(ctx, means3d, scales, glob_scale, quats, viewmat, projmat, fx, fy, cx, cy,
                img_height, img_width, tile_bounds, clip_thresh=0.01):
        if means3d.ndim != 2 or means3d.shape[1] != 3:
            raise ValueError("means3d must have dimensions (N, 3)")
        if scales.shape != means3d.shape:
            raise ValueError("scales must have dimensions (N, 3)")
        N = means3d.shape[0]
        if quats.ndim != 2 or quats.shape[0] != N or quats.shape[1] != 4:
            raise ValueError("quats must have dimensions (N, 4)")
        if not means3d.is_cuda:
            raise ValueError("project_gaussians needs CUDA tensors (no CPU fallback)")
        if viewmat.numel() < 12 or projmat.numel() != 16:
            raise ValueError("viewmat must hold >= 12 and projmat 16 elements")
        dev = means3d.device
        means3d, scales, quats = _f32c(means3d), _f32c(scales), _f32c(quats)
        V = _f32c(viewmat.to(dev)).reshape(-1)
        P = _f32c(projmat.to(dev)).reshape(-1)
        H, W = int(img_height), int(img_width)
        tx, ty = int(tile_bounds[0]), int(tile_bounds[1])
        # one allocation for the six outputs (12 float + 2 int32 words per Gaussian), carved into the operator's shapes
        buf = torch.empty((14 * N,), device=dev, dtype=torch.float32)
        xys, cov3d, conics = buf[0:2 * N].view(N, 2), buf[2 * N:8 * N].view(N, 6), buf[8 * N:11 * N].view(N, 3)   # xys: 8 B aligned
        depths = buf[11 * N:12 * N]
        radii, hit = buf[12 * N:13 * N].view(torch.int32), buf[13 * N:].view(torch.int32)
        with torch.cuda.device(dev):
            rc = _lib.lib().gs_project_fwd(
                N, _lib.ptr(means3d), _lib.ptr(scales), float(glob_scale), _lib.ptr(quats), _lib.ptr(V),
                _lib.ptr(P), float(fx), float(fy), float(cx), float(cy), H, W, tx, ty, float(clip_thresh),
                _lib.ptr(cov3d), _lib.ptr(xys), _lib.ptr(depths), _lib.ptr(radii), _lib.ptr(conics),
                _lib.ptr(hit), _lib.stream_ptr())
        _lib.check(rc, "gs_project_fwd")
        ctx.save_for_backward(means3d, scales, quats, V, P, radii)
        ctx.consts = (float(glob_scale), float(fx), float(fy), H, W)
        ctx.mark_non_differentiable(radii, hit)
        return xys, depths, radii, conics, hit, cov3d

    @staticmethod
    def backward(ctx, v_xys, v_depths, v_radii, v_conics, v_hit, v_cov3d):
        means3d, scales, quats, V, P, radii = ctx.saved_tensors
        glob, fx, fy, H, W = ctx.consts
        N = means3d.shape[0]
        dev = means3d.device
        v_xys = torch.zeros((N, 2), device=dev) if v_xys is None else _f32c(v_xys)
        v_depths = torch.zeros((N,), device=dev) if v_depths is None else _f32c(v_depths)
        v_conics = torch.zeros((N, 3), device=dev) if v_conics is None else _f32c(v_conics)
        g = torch.empty((10 * N,), device=dev, dtype=torch.float32)          # one allocation for the three gradients
        v_quats, v_means, v_scales = g[0:4 * N].view(N, 4), g[4 * N:7 * N].view(N, 3), g[7 * N:].view(N, 3)   # quats: 16 B aligned
        with torch.cuda.device(dev):
            rc = _lib.lib().gs_project_bwd(
                N, _lib.ptr(means3d), _lib.ptr(scales), glob, _lib.ptr(quats), _lib.ptr(V), _lib.ptr(P),
                fx, fy, H, W, _lib.ptr(radii), _lib.ptr(v_xys), _lib.ptr(v_depths), _lib.ptr(v_conics),
                _lib.ptr(v_means), _lib.ptr(v_scales), _lib.ptr(v_quats), _lib.stream_ptr())
        _lib.check(rc, "gs_project_bwd")
        return (v_means, v_scales, None, v_quats, None, None, None, None, None, None, None, None, None, None)
