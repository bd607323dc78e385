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


# ----------------------------------------------------------------------------------------------
# Parity at full size (tests/test_parity_scale_gpu.py, __graft_entry__.smoke()).
#
# The forward takes three discontinuous decisions per (pixel, Gaussian) pair:
# alpha >= 1/255, T(1 - alpha) <= 1e-4, and raw alpha vs the 0.999 clamp (derivative only). Two correct
# fp32 implementations (expf in the oracle, ex2.approx on the GPU, __expf in gsplat itself) differ by a
# few ulp in alpha, so they can decide differently ONLY where the quantity sits within rounding distance
# of its threshold. The oracle reports that distance per pixel (orc_raster_risk); the contract checked is
#   (1) every pixel that deviates by more than TOL is a pixel where a decision was within EPS_* of its
#       threshold ("at-risk"); all other pixels — ~99 % of the image — are within TOL, strictly;
#   (2) a deviating pixel is off by at most one alpha quantum;
#   (3) every gradient row of a Gaussian that influences no deviating / clamp-at-risk pixel is within TOL
#       (relative to the tensor's max magnitude).
TOL = 1e-4
EPS_ALPHA = 5e-5      # |raw alpha / (1/255) - 1|: conic / opacity rounding (~1e-6) x sigma (<= ~6) + ex2.approx
EPS_T = 5e-4          # |T(1-alpha) / 1e-4 - 1|: T is a product of up to hundreds of (1 - alpha) factors
EPS_CLAMP = 5e-5      # |raw alpha / 0.999 - 1|
QUANTUM = 1.5 / 255   # a flipped alpha >= 1/255 decision moves a pixel by at most ~one alpha quantum


def percentile(a, q):
    a = np.asarray(a, np.float64).reshape(-1)
    return float(np.quantile(a, q)) if a.size else 0.0


def forward_parity(orc, ctx, got_rgb, ref_rgb, got_acc, ref_acc, H, W, got_fidx=None, extra_risk=None):
    """-> (report dict, deviating-pixel mask F, gradient-exclusion pixel mask). Asserts (1) and (2).
    extra_risk: optional (H,W) bool mask of pixels excused for another stated reason (see tile_mask_of)."""
    ra, rt, rc = orc.raster_risk(ctx["b"], ctx["xys"], ctx["conics"], ctx["opac"], H, W, ctx["tb"])
    at_risk = (ra < EPS_ALPHA) | (rt < EPS_T)
    if extra_risk is not None:
        at_risk = at_risk | extra_risk
    err_rgb = np.abs(np.asarray(got_rgb, np.float64) - ref_rgb).max(-1)
    err_acc = np.abs(np.asarray(got_acc, np.float64).reshape(H, W) - ref_acc.reshape(H, W))
    F = (err_rgb > TOL) | (err_acc > TOL)
    rep = {
        "pixels": int(H * W), "at_risk_frac": float(at_risk.mean()), "deviating_pixels": int(F.sum()),
        "deviating_frac": float(F.mean()),
        "rgb_err_max": float(err_rgb.max()), "rgb_err_p9999": percentile(err_rgb, 0.9999),
        "rgb_err_max_not_at_risk": float(err_rgb[~at_risk].max()) if (~at_risk).any() else 0.0,
        "acc_err_max": float(err_acc.max()), "acc_err_p9999": percentile(err_acc, 0.9999),
        "acc_err_max_not_at_risk": float(err_acc[~at_risk].max()) if (~at_risk).any() else 0.0,
        "largest_margin_of_a_deviating_pixel": float(np.minimum(ra / EPS_ALPHA, rt / EPS_T)[F].max()) if F.any() else 0.0,
    }
    if got_fidx is not None:
        same = np.asarray(got_fidx).reshape(H, W) == ctx["fidx"].reshape(H, W)
        rep["final_idx_mismatch_pixels"] = int((~same).sum())
        assert same[~at_risk].all(), "final_idx differs on a pixel where no decision was near its threshold"
    assert not (F & ~at_risk).any(), f"{int((F & ~at_risk).sum())} pixels deviate > {TOL} without any decision near a threshold: {rep}"
    assert rep["rgb_err_max"] < QUANTUM and rep["acc_err_max"] < QUANTUM, rep
    excl = F | (rc < EPS_CLAMP)
    if got_fidx is not None:
        excl |= ~same
    return rep, F, excl


def tile_mask_of(ctx, gaussians, radii_a, radii_b, H, W):
    """(H,W) bool mask of the 16x16 tiles that the given Gaussians cover with either of two radii. Used for the
    operator-API path, whose pre-activations (torch.exp / norm on the GPU, the CALLER's ops, gc_model.py:176-178) can
    differ by one ulp from the oracle's: through ceil() that moves the integer radius of ~1e-6 of the Gaussians, i.e.
    changes which tiles list them. Those tiles are excused — and counted in the report."""
    mask = np.zeros((H, W), bool)
    tb = ctx["tb"]
    for g in gaussians:
        x, y = ctx["xys"][g]
        r = float(max(radii_a[g], radii_b[g]))
        x0, x1 = max(int((x - r) / 16.0), 0), min(int((x + r) / 16.0 + 1), tb[0])
        y0, y1 = max(int((y - r) / 16.0), 0), min(int((y + r) / 16.0 + 1), tb[1])
        mask[16 * y0:16 * y1, 16 * x0:16 * x1] = True
    return mask


def gradient_parity(orc, ctx, excl_pixels, got_grads, ref_grads, H, W):
    """(3): rel L-inf of every gradient tensor over the Gaussians that influence no excluded pixel.
    got/ref_grads: dict name -> array with leading dimension N. Returns report; asserts <= TOL."""
    N = ctx["xys"].shape[0]
    upto = np.full((H, W), np.iinfo(np.int32).max, np.int32)
    touched = orc.mark_contributors(ctx["b"], ctx["xys"], ctx["conics"], ctx["opac"], H, W, ctx["tb"],
                                    excl_pixels.astype(np.uint8), upto)
    rep = {"excluded_pixels": int(excl_pixels.sum()), "excluded_gaussians": int(touched.sum()),
           "excluded_gaussian_frac": float(touched.mean()), "rel_linf": {}, "rel_linf_all_rows": {}}
    keep = ~touched
    for name, ref in ref_grads.items():
        got = np.asarray(got_grads[name], np.float64).reshape(N, -1)
        ref = np.asarray(ref, np.float64).reshape(N, -1)
        den = max(np.abs(ref).max(), 1e-30)
        rep["rel_linf"][name] = float(np.abs(got[keep] - ref[keep]).max() / den)
        rep["rel_linf_all_rows"][name] = float(np.abs(got - ref).max() / den)
    for name, v in rep["rel_linf"].items():
        assert v <= TOL, f"gradient {name}: {v:.3e} > {TOL} on Gaussians that touch no deviating pixel ({rep})"
    return rep


def depth_parity(got_depth, ref_depth, ref_acc, F):
    """Eval depth = sum(w d) / alpha with alpha = 1 - T_final (gc_model.py:237): the subtraction cancels, so the
    fp32 rounding of T (2^-24 absolute) is amplified by 1/alpha — in the reference as much as here. Bound:
    relative error <= TOL + 4 * 2^-24 / alpha on every non-deviating pixel with alpha > 0."""
    H, W = F.shape
    acc = np.asarray(ref_acc, np.float64).reshape(H, W)
    ref = np.asarray(ref_depth, np.float64).reshape(H, W)
    got = np.asarray(got_depth, np.float64).reshape(H, W)
    ok = ~F & (acc > 0)
    rel = np.abs(got - ref) / np.maximum(np.abs(ref), 1e-30)
    bound = TOL + 4 * 2.0 ** -24 / np.maximum(acc, 1e-30)
    worst = float((rel[ok] / bound[ok]).max()) if ok.any() else 0.0
    assert worst <= 1.0, f"depth error is {worst:.2f}x the stated bound"
    solid = ok & (acc > 0.5)
    return {"rel_err_over_bound_max": worst, "rel_err_max_alpha_gt_0.5": float(rel[solid].max()) if solid.any() else 0.0}
