"""`num_sh_bases` / `spherical_harmonics` with the gsplat-0.1.2 signatures.

Replaces the call at /root/reference/gaussctrl/gc_model.py:199-200 (import at :32).
Semantics: SURVEY.md App. A.2. Differentiable w.r.t. the coefficients only.
"""
from __future__ import annotations

import torch
from torch import Tensor
from torch.autograd import Function

from . import _lib


def num_sh_bases(degree: int) -> int:
    if degree == 0:
        return 1
    if degree == 1:
        return 4
    if degree == 2:
        return 9
    if degree == 3:
        return 16
    return 25


def spherical_harmonics(degrees_to_use: int, viewdirs: Tensor, coeffs: Tensor) -> Tensor:
    """viewdirs (N,3), coeffs (N,K,3) with K >= num_sh_bases(degrees_to_use) -> colors (N,3)."""
    assert coeffs.shape[-2] >= num_sh_bases(degrees_to_use)
    return _SphericalHarmonics.apply(degrees_to_use, viewdirs.contiguous(), coeffs.contiguous())


class _SphericalHarmonics(Function):
    @staticmethod
    def forward(ctx, degrees_to_use: int, viewdirs: Tensor, coeffs: Tensor):
        if viewdirs.ndim != 2 or viewdirs.shape[1] != 3:
            raise ValueError("viewdirs must have dimensions (N, 3)")
        if coeffs.ndim != 3 or coeffs.shape[2] != 3 or coeffs.shape[0] != viewdirs.shape[0]:
            raise ValueError("coeffs must have dimensions (N, K, 3)")
        if not coeffs.is_cuda:
            raise ValueError("spherical_harmonics needs CUDA tensors (no CPU fallback)")
        N, K = coeffs.shape[0], coeffs.shape[1]
        dev = coeffs.device
        viewdirs = viewdirs.float().contiguous()
        coeffs = coeffs.float().contiguous()
        colors = torch.empty((N, 3), device=dev, dtype=torch.float32)
        with torch.cuda.device(dev):
            rc = _lib.lib().gs_sh_fwd(N, K, int(degrees_to_use), _lib.ptr(viewdirs), _lib.ptr(coeffs),
                                      _lib.ptr(colors), _lib.stream_ptr())
        _lib.check(rc, "gs_sh_fwd")
        ctx.save_for_backward(viewdirs)
        ctx.meta = (int(degrees_to_use), K)
        return colors

    @staticmethod
    def backward(ctx, v_colors: Tensor):
        (viewdirs,) = ctx.saved_tensors
        deg, K = ctx.meta
        N = viewdirs.shape[0]
        v_colors = v_colors.float().contiguous()
        v_coeffs = torch.empty((N, K, 3), device=viewdirs.device, dtype=torch.float32)
        with torch.cuda.device(viewdirs.device):
            rc = _lib.lib().gs_sh_bwd(N, K, deg, _lib.ptr(viewdirs), _lib.ptr(v_colors), _lib.ptr(v_coeffs),
                                      _lib.stream_ptr())
        _lib.check(rc, "gs_sh_bwd")
        return None, None, v_coeffs
