"""Seeded synthetic scenes and cameras.

The reference ships no trained Gaussians, so benchmarks and tests use
synthetic splatfacto-shaped parameter sets: the six leaf tensors of
/root/reference/gaussctrl/gc_model.py:157-170 and cameras with the intrinsics of
/root/reference/data/garden/transforms.json:2-7.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, List

import torch

# fx, fy, cx, cy, W, H of data/garden/transforms.json:2-7
GARDEN_INTRINSICS = dict(fx=585.78, fy=585.51, cx=255.98, cy=253.71, W=512, H=512)
HD_INTRINSICS = dict(fx=1200.0, fy=1200.0, cx=960.0, cy=540.0, W=1920, H=1080)

PARAM_NAMES = ("means", "scales", "quats", "features_dc", "features_rest", "opacities")


def default_scale(n: int) -> float:
    """s0 per Gaussian count so that I/N stays at 3-5 at 512^2."""
    if n <= 50_000:
        return 0.005
    if n <= 500_000:
        return 0.004
    if n <= 1_000_000:
        return 0.003
    return 0.002


# Second synthetic workload: LARGE, anisotropic splats (the long tail of a trained splatfacto scene, scripts/garden.sh:3,
# which the small-splat headline scene of SURVEY.md 8d does not have): 4x the headline scene's median scale and a
# log-scale spread of 1.0 instead of 0.5, so a splat covers tens of tiles and axis ratios of 10+ are common.
LARGE_SPLATS = dict(n=300_000, s0=0.012, log_sigma=1.0)


def make_scene(n: int, seed: int = 0, s0: float | None = None, device="cpu", log_sigma: float = 0.5) -> Dict[str, torch.Tensor]:
    """Leaf parameters in splatfacto's raw (pre-activation) form. log-scales ~ N(log s0, log_sigma^2)."""
    g = torch.Generator().manual_seed(seed)
    s0 = default_scale(n) if s0 is None else s0
    p = {
        "means": torch.randn(n, 3, generator=g) * 0.3,
        "scales": math.log(s0) + log_sigma * torch.randn(n, 3, generator=g),
        "quats": torch.randn(n, 4, generator=g),
        "features_dc": torch.randn(n, 3, generator=g),
        "features_rest": 0.1 * torch.randn(n, 15, 3, generator=g),
        "opacities": 2.0 * torch.randn(n, 1, generator=g),
    }
    return {k: v.float().contiguous().to(device) for k, v in p.items()}


@dataclass
class Camera:
    """The fields of nerfstudio `Cameras` that gc_model.py:115-145 reads, for one camera."""
    c2w: torch.Tensor  # (3,4) OpenGL convention (-z forward, +y up), float32, CPU
    fx: float
    fy: float
    cx: float
    cy: float
    W: int
    H: int

    def as_dict(self):
        return dict(c2w=self.c2w.numpy(), fx=self.fx, fy=self.fy, cx=self.cx, cy=self.cy, W=self.W, H=self.H)

    def rescale_output_resolution(self, scaling_factor: float) -> None:
        """nerfstudio `Cameras.rescale_output_resolution` (called at gc_model.py:115 and :206), default
        "floor" rounding: intrinsics scale, width / height truncate. In place, like the original."""
        self.fx, self.fy = self.fx * scaling_factor, self.fy * scaling_factor
        self.cx, self.cy = self.cx * scaling_factor, self.cy * scaling_factor
        self.W, self.H = int(self.W * scaling_factor), int(self.H * scaling_factor)

    def signature(self):
        """Value key for caches: an in-place pose / intrinsics update (camera optimiser, rescale) must not
        hit a stale entry."""
        return (self.c2w.detach().cpu().numpy().tobytes(), float(self.fx), float(self.fy), float(self.cx),
                float(self.cy), int(self.W), int(self.H))


@dataclass
class OrientedBox:
    """nerfstudio `OrientedBox` as far as gc_model.py:108-113 / :244-251 use it (`crop_box.within(means)`):
    R (3,3) rotation, T (3,) centre, S (3,) full side lengths."""
    R: torch.Tensor
    T: torch.Tensor
    S: torch.Tensor

    def within(self, pts: torch.Tensor) -> torch.Tensor:
        R, T, S = self.R.to(pts), self.T.to(pts), self.S.to(pts)
        H = torch.eye(4, device=pts.device, dtype=pts.dtype)
        H[:3, :3] = R
        H[:3, 3] = T
        H_world2bbox = torch.inverse(H)
        ph = torch.cat((pts, torch.ones_like(pts[..., :1])), dim=-1)
        local = torch.matmul(H_world2bbox, ph.T).T[..., :3]
        return torch.all(torch.cat([local > -S / 2, local < S / 2], dim=-1), dim=-1)


def look_at(pos: torch.Tensor, target: torch.Tensor, up=(0.0, 0.0, 1.0)) -> torch.Tensor:
    """c2w (3,4) of a camera at `pos` looking at `target`; columns = right, up, back."""
    pos = pos.double()
    fwd = target.double() - pos
    fwd = fwd / fwd.norm()
    upv = torch.tensor(up, dtype=torch.float64)
    right = torch.linalg.cross(fwd, upv)
    if right.norm() < 1e-6:
        right = torch.linalg.cross(fwd, torch.tensor([0.0, 1.0, 0.0], dtype=torch.float64))
    right = right / right.norm()
    true_up = torch.linalg.cross(right, fwd)
    c2w = torch.stack([right, true_up, -fwd, pos], dim=1)
    return c2w.float()


def make_cameras(v: int, seed: int = 0, radius: float = 1.5, intrinsics=None) -> List[Camera]:
    """V poses on a sphere (Fibonacci lattice + seeded jitter) looking at the origin."""
    intr = dict(GARDEN_INTRINSICS if intrinsics is None else intrinsics)
    g = torch.Generator().manual_seed(1000 + seed)
    cams = []
    golden = math.pi * (3.0 - math.sqrt(5.0))
    for i in range(v):
        z = 1.0 - 2.0 * (i + 0.5) / v
        z = 0.8 * z  # keep away from the poles so the up vector stays well defined
        r = math.sqrt(max(0.0, 1.0 - z * z))
        th = golden * i
        d = torch.tensor([r * math.cos(th), r * math.sin(th), z], dtype=torch.float64)
        d = d + 0.02 * torch.randn(3, generator=g, dtype=torch.float64)
        pos = radius * d / d.norm()
        cams.append(Camera(look_at(pos, torch.zeros(3, dtype=torch.float64)), **intr))
    return cams
