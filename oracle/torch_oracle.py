"""Autograd restatement of the hot path in float64 (CPU ORACLE — test infrastructure only).

Purpose: an independent check of the C oracle's hand-derived backward (gs_oracle.c) — the
gradients here come from torch.autograd through a vectorised restatement of SURVEY.md App. A,
so "backward = exact derivative of the forward" is verified rather than asserted. Discrete
decisions (culling, tile lists, skip / termination tests) are taken from the C oracle's fp32
forward so both differentiate the same piecewise-smooth function.

Follows the operators called at /root/reference/gaussctrl/gc_model.py:174, :200, :208.
PARITY UNPINNED (see gs_oracle.h).
"""
from __future__ import annotations

import numpy as np
import torch

from . import oracle as orc

SH_C0 = 0.28209479177387814
SH_C1 = 0.4886025119029199
SH_C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
SH_C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154,
         -0.4570457994644658, 1.445305721320277, -0.5900435899266435]


def project(means, scales, glob, quats, V, P, fx, fy, cx, cy, H, W):
    """float64 tensors. Returns xys (N,2), depths (N), conics (N,3) (no culling applied)."""
    Vt = torch.as_tensor(np.asarray(V, np.float64))[:3]
    Pt = torch.as_tensor(np.asarray(P, np.float64)).reshape(4, 4)
    t = means @ Vt[:, :3].T + Vt[:, 3]
    q = quats / quats.norm(dim=-1, keepdim=True)
    w, x, y, z = q.unbind(-1)
    R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                     2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                     2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], -1).reshape(-1, 3, 3)
    M = R * (glob * scales)[:, None, :]
    S = M @ M.transpose(1, 2)
    tz = t[:, 2]
    limx, limy = 1.3 * (0.5 * W / fx), 1.3 * (0.5 * H / fy)
    tx = tz * torch.clamp(t[:, 0] / tz, -limx, limx)
    ty = tz * torch.clamp(t[:, 1] / tz, -limy, limy)
    zero = torch.zeros_like(tz)
    J = torch.stack([fx / tz, zero, -fx * tx / tz ** 2, zero, fy / tz, -fy * ty / tz ** 2], -1).reshape(-1, 2, 3)
    Tm = J @ Vt[:, :3]
    cov = Tm @ S @ Tm.transpose(1, 2)
    a = cov[:, 0, 0] + 0.3
    b = cov[:, 0, 1]
    c = cov[:, 1, 1] + 0.3
    det = a * c - b * b
    conics = torch.stack([c / det, -b / det, a / det], -1)
    h = means @ Pt[:, :3].T + Pt[:, 3]
    rw = 1.0 / (h[:, 3] + 1e-6)
    xys = torch.stack([0.5 * W * h[:, 0] * rw + cx - 0.5, 0.5 * H * h[:, 1] * rw + cy - 0.5], -1)
    return xys, tz, conics


def sh_colors(deg, dirs, coeffs):
    d = dirs / dirs.norm(dim=-1, keepdim=True)
    x, y, z = d.unbind(-1)
    Y = [torch.full_like(x, SH_C0)]
    if deg >= 1:
        Y += [-SH_C1 * y, SH_C1 * z, -SH_C1 * x]
    if deg >= 2:
        xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
        Y += [SH_C2[0] * xy, SH_C2[1] * yz, SH_C2[2] * (2 * zz - xx - yy), SH_C2[3] * xz, SH_C2[4] * (xx - yy)]
    if deg >= 3:
        Y += [SH_C3[0] * y * (3 * xx - yy), SH_C3[1] * xy * z, SH_C3[2] * y * (4 * zz - xx - yy),
              SH_C3[3] * z * (2 * zz - 3 * xx - 3 * yy), SH_C3[4] * x * (4 * zz - xx - yy),
              SH_C3[5] * z * (xx - yy), SH_C3[6] * x * (xx - 3 * yy)]
    Yt = torch.stack(Y, -1)
    return (Yt[:, :, None] * coeffs[:, : Yt.shape[1], :]).sum(1)


def composite(xys, conics, colors, opac, bg, H, W, b: orc.Binning, fidx, tile_bounds):
    """Differentiable compositing that replays the C oracle's discrete decisions: pixel p
    composites exactly the Gaussians k < final_idx[p] of its tile list that pass the skip
    tests (evaluated here in float64)."""
    C = colors.shape[1]
    out = torch.zeros(H, W, C, dtype=torch.float64)
    alpha_out = torch.zeros(H, W, dtype=torch.float64)
    tx_n, ty_n = tile_bounds[0], tile_bounds[1]
    fidx_t = torch.as_tensor(np.asarray(fidx, np.int64))
    bgt = torch.as_tensor(np.asarray(bg, np.float64))
    for ty in range(ty_n):
        for tx in range(tx_n):
            s, e = int(b.tile_bins[ty * tx_n + tx, 0]), int(b.tile_bins[ty * tx_n + tx, 1])
            y0, y1, x0, x1 = ty * 16, min(H, ty * 16 + 16), tx * 16, min(W, tx * 16 + 16)
            if y1 <= y0 or x1 <= x0:
                continue
            if e <= s:
                out[y0:y1, x0:x1] = bgt
                continue
            ids = torch.as_tensor(b.ids_sorted[s:e].astype(np.int64))
            py, px = torch.meshgrid(torch.arange(y0, y1, dtype=torch.float64),
                                    torch.arange(x0, x1, dtype=torch.float64), indexing="ij")
            dx = xys[ids, 0][None, None, :] - px[..., None]
            dy = xys[ids, 1][None, None, :] - py[..., None]
            cn = conics[ids]
            sigma = 0.5 * (cn[:, 0] * dx * dx + cn[:, 2] * dy * dy) + cn[:, 1] * dx * dy
            vis = torch.exp(-sigma)
            raw = opac[ids][None, None, :] * vis
            alpha = torch.clamp(raw, max=0.999)
            k = torch.arange(s, e)[None, None, :]
            live = (sigma >= 0) & (alpha >= 1.0 / 255.0) & (k < fidx_t[y0:y1, x0:x1, None])
            alpha = torch.where(live, alpha, torch.zeros_like(alpha))
            T = torch.cumprod(1 - alpha, dim=-1)
            Tprev = torch.cat([torch.ones_like(T[..., :1]), T[..., :-1]], -1)
            wgt = alpha * Tprev
            out[y0:y1, x0:x1] = wgt @ colors[ids] + T[..., -1:] * bgt
            alpha_out[y0:y1, x0:x1] = 1 - T[..., -1]
    return out, alpha_out
