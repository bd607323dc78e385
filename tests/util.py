"""Shared helpers for the parity tests: seeded inputs in numpy (oracle side) and torch (CUDA side)."""
import numpy as np
import torch

from gaussctrl_exp_b200 import scene as gscene


def small_cam(W=64, H=48, seed=0, f=None):
    f = f if f is not None else 1.1 * max(W, H)
    cam = gscene.make_cameras(1, seed, intrinsics=dict(fx=f, fy=f * 0.99, cx=W / 2 - 0.3, cy=H / 2 + 0.2, W=W, H=H))[0]
    return cam


def np_params(N, seed=0, s0=None):
    return {k: v.numpy() for k, v in gscene.make_scene(N, seed, s0=s0).items()}


def rel_linf(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    den = max(np.abs(b).max(), 1e-30)
    return float(np.abs(a - b).max() / den)


def to_dev(p, dev):
    return {k: torch.from_numpy(np.ascontiguousarray(v)).to(dev) for k, v in p.items()}


def projected(orc, N=2000, W=64, H=48, seed=0, s0=0.02):
    """Oracle-side projection of a seeded scene; returns dict with everything downstream stages need."""
    p = np_params(N, seed, s0)
    cam = small_cam(W, H, seed)
    V, PV, tb = orc.camera_matrices(cam.c2w.numpy(), cam.fx, cam.fy, cam.cx, cam.cy, W, H)
    scales, quats, opac, coeffs, dirs = orc.preactivate(p, cam.c2w.numpy(), 3)
    xys, depths, radii, conics, hit, cov3d = orc.project_fwd(p["means"], scales, 1.0, quats, V[:3], PV, cam.fx, cam.fy,
                                                             cam.cx, cam.cy, H, W, tb)
    return dict(p=p, cam=cam, V=V, PV=PV, tb=tb, scales=scales, quats=quats, opac=opac, coeffs=coeffs, dirs=dirs,
                xys=xys, depths=depths, radii=radii, conics=conics, hit=hit, cov3d=cov3d, W=W, H=H, N=N)
