"""ctypes/numpy front-end of the CPU ORACLE (test infrastructure, NOT product code).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module. PARITY UNPINNED — see gs_oracle.h.

The functions mirror the gsplat-0.1.2 operators called at
/root/reference/gaussctrl/gc_model.py:174 (project), :200 (SH), :208/:225 (rasterize)
and `render_view` mirrors the caller itself (gc_model.py:117-241).
"""
from __future__ import annotations

import ctypes as C
import math
import os
import subprocess
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libgs_oracle.so")
_lib = None

f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "gs_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-B", "libgs_oracle.so"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        L.orc_scan.restype = C.c_int64
        L.orc_raster_fwd.restype = C.c_int64
        L.orc_num_threads.restype = C.c_int
        _lib = L
    return _lib


def _f(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def set_num_threads(n: int):
    lib().orc_set_num_threads(C.c_int(n))


def num_threads() -> int:
    return lib().orc_num_threads()


def project_fwd(means, scales, glob_scale, quats, viewmat, projmat, fx, fy, cx, cy, H, W,
                tile_bounds, clip=0.01):
    means, scales, quats = _f(means), _f(scales), _f(quats)
    V, P = _f(viewmat).reshape(-1), _f(projmat).reshape(-1)
    assert V.size >= 12 and P.size == 16
    N = means.shape[0]
    cov3d = np.empty((N, 6), np.float32)
    xys = np.empty((N, 2), np.float32)
    depths = np.empty((N,), np.float32)
    radii = np.empty((N,), np.int32)
    conics = np.empty((N, 3), np.float32)
    hit = np.empty((N,), np.int32)
    lib().orc_project_fwd(C.c_int(N), _p(means), _p(scales), C.c_float(glob_scale), _p(quats),
                          _p(V), _p(P), C.c_float(fx), C.c_float(fy), C.c_float(cx), C.c_float(cy),
                          C.c_int(H), C.c_int(W), C.c_int(tile_bounds[0]), C.c_int(tile_bounds[1]),
                          C.c_float(clip), _p(cov3d), _p(xys), _p(depths), _p(radii), _p(conics), _p(hit))
    return xys, depths, radii, conics, hit, cov3d


def project_bwd(means, scales, glob_scale, quats, viewmat, projmat, fx, fy, cx, cy, H, W, radii,
                v_xys, v_depths, v_conics):
    means, scales, quats = _f(means), _f(scales), _f(quats)
    V, P = _f(viewmat).reshape(-1), _f(projmat).reshape(-1)
    N = means.shape[0]
    radii = np.ascontiguousarray(radii, np.int32)
    v_xys, v_depths, v_conics = _f(v_xys), _f(v_depths), _f(v_conics)
    vm = np.empty((N, 3), np.float32)
    vs = np.empty((N, 3), np.float32)
    vq = np.empty((N, 4), np.float32)
    lib().orc_project_bwd(C.c_int(N), _p(means), _p(scales), C.c_float(glob_scale), _p(quats),
                          _p(V), _p(P), C.c_float(fx), C.c_float(fy), C.c_float(cx), C.c_float(cy),
                          C.c_int(H), C.c_int(W), _p(radii), _p(v_xys), _p(v_depths), _p(v_conics),
                          _p(vm), _p(vs), _p(vq))
    return vm, vs, vq


def sh_fwd(deg, dirs, coeffs):
    dirs, coeffs = _f(dirs), _f(coeffs)
    N, K = coeffs.shape[0], coeffs.shape[1]
    assert K >= (deg + 1) ** 2
    out = np.empty((N, 3), np.float32)
    lib().orc_sh_fwd(C.c_int(N), C.c_int(K), C.c_int(deg), _p(dirs), _p(coeffs), _p(out))
    return out


def sh_bwd(deg, K, dirs, v_colors):
    dirs, v_colors = _f(dirs), _f(v_colors)
    N = dirs.shape[0]
    out = np.empty((N, K, 3), np.float32)
    lib().orc_sh_bwd(C.c_int(N), C.c_int(K), C.c_int(deg), _p(dirs), _p(v_colors), _p(out))
    return out


@dataclass
class Binning:
    cum: np.ndarray
    keys: np.ndarray
    ids: np.ndarray
    keys_sorted: np.ndarray
    ids_sorted: np.ndarray
    tile_bins: np.ndarray
    I: int


def bin_and_sort(xys, depths, radii, tiles_hit, tile_bounds) -> Binning:
    xys, depths = _f(xys), _f(depths)
    radii = np.ascontiguousarray(radii, np.int32)
    hit = np.ascontiguousarray(tiles_hit, np.int32)
    N = hit.shape[0]
    tx, ty = int(tile_bounds[0]), int(tile_bounds[1])
    cum = np.empty((N,), np.int32)
    I = int(lib().orc_scan(C.c_int(N), _p(hit), _p(cum)))
    keys = np.empty((I,), np.int64)
    ids = np.empty((I,), np.int32)
    ks = np.empty((I,), np.int64)
    isd = np.empty((I,), np.int32)
    bins = np.zeros((tx * ty, 2), np.int32)
    if I > 0:
        lib().orc_emit(C.c_int(N), _p(xys), _p(depths), _p(radii), _p(cum), C.c_int(tx), C.c_int(ty),
                       _p(keys), _p(ids))
        lib().orc_sort(C.c_int64(I), _p(keys), _p(ids), _p(ks), _p(isd))
        lib().orc_tile_bins(C.c_int64(I), C.c_int(tx * ty), _p(ks), _p(bins))
    return Binning(cum, keys, ids, ks, isd, bins, I)


def raster_fwd(b: Binning, xys, conics, colors, opac, bg, H, W, tile_bounds):
    xys, conics, colors, opac, bg = _f(xys), _f(conics), _f(colors), _f(opac).reshape(-1), _f(bg)
    Cc = colors.shape[1]
    out = np.empty((H, W, Cc), np.float32)
    fT = np.empty((H, W), np.float32)
    fidx = np.empty((H, W), np.int32)
    pairs = lib().orc_raster_fwd(C.c_int(Cc), C.c_int(H), C.c_int(W), C.c_int(tile_bounds[0]),
                                 C.c_int(tile_bounds[1]), _p(b.ids_sorted), _p(b.tile_bins), _p(xys),
                                 _p(conics), _p(colors), _p(opac), _p(bg), _p(out), _p(fT), _p(fidx))
    return out, fT, fidx, int(pairs)


def raster_bwd(b: Binning, xys, conics, colors, opac, bg, H, W, tile_bounds, fT, fidx, v_out,
               v_out_alpha=None):
    xys, conics, colors, opac, bg = _f(xys), _f(conics), _f(colors), _f(opac).reshape(-1), _f(bg)
    v_out = _f(v_out)
    N, Cc = colors.shape
    va = _f(v_out_alpha) if v_out_alpha is not None else None
    v_xys = np.empty((N, 2), np.float32)
    v_con = np.empty((N, 3), np.float32)
    v_col = np.empty((N, Cc), np.float32)
    v_op = np.empty((N,), np.float32)
    lib().orc_raster_bwd(C.c_int(Cc), C.c_int(H), C.c_int(W), C.c_int(tile_bounds[0]),
                         C.c_int(tile_bounds[1]), C.c_int(N), _p(b.ids_sorted), _p(b.tile_bins),
                         _p(xys), _p(conics), _p(colors), _p(opac), _p(bg), _p(_f(fT)),
                         _p(np.ascontiguousarray(fidx, np.int32)), _p(v_out),
                         _p(va) if va is not None else None,
                         _p(v_xys), _p(v_con), _p(v_col), _p(v_op))
    return v_xys, v_con, v_col, v_op[:, None]


def raster_risk(b: Binning, xys, conics, opac, H, W, tile_bounds):
    """Per-pixel smallest relative distance of a forward decision to its threshold:
    -> (risk_alpha, risk_T, risk_clamp), each (H,W) float32 (1e30 where no decision was taken)."""
    xys, conics, opac = _f(xys), _f(conics), _f(opac).reshape(-1)
    ra = np.empty((H, W), np.float32)
    rt = np.empty((H, W), np.float32)
    rc = np.empty((H, W), np.float32)
    lib().orc_raster_risk(C.c_int(H), C.c_int(W), C.c_int(tile_bounds[0]), C.c_int(tile_bounds[1]),
                          _p(b.ids_sorted), _p(b.tile_bins), _p(xys), _p(conics), _p(opac), _p(ra), _p(rt), _p(rc))
    return ra, rt, rc


def mark_contributors(b: Binning, xys, conics, opac, H, W, tile_bounds, mask, upto):
    """-> touched (N,) bool: Gaussians that can influence a pixel with mask != 0 (list entries before upto[pix])."""
    xys, conics, opac = _f(xys), _f(conics), _f(opac).reshape(-1)
    mask = np.ascontiguousarray(mask, np.uint8)
    upto = np.ascontiguousarray(upto, np.int32)
    touched = np.zeros((xys.shape[0],), np.uint8)
    lib().orc_mark_contributors(C.c_int(H), C.c_int(W), C.c_int(tile_bounds[0]), C.c_int(tile_bounds[1]),
                                _p(b.ids_sorted), _p(b.tile_bins), _p(xys), _p(conics), _p(opac), _p(mask),
                                _p(upto), _p(touched))
    return touched.astype(bool)


# ----------------------------------------------------------------------------------------------
# Caller semantics: /root/reference/gaussctrl/gc_model.py:117-241 restated in numpy.
# ----------------------------------------------------------------------------------------------
def projection_matrix(znear, zfar, fovx, fovy):
    """nerfstudio's `projection_matrix`, as restated in-tree at ad_render.py:49-67."""
    t = znear * math.tan(0.5 * fovy)
    b = -t
    r = znear * math.tan(0.5 * fovx)
    l = -r
    n, f = znear, zfar
    return np.array([[2 * n / (r - l), 0.0, (r + l) / (r - l), 0.0],
                     [0.0, 2 * n / (t - b), (t + b) / (t - b), 0.0],
                     [0.0, 0.0, (f + n) / (f - n), -1.0 * f * n / (f - n)],
                     [0.0, 0.0, 1.0, 0.0]], dtype=np.float32)


def camera_matrices(c2w, fx, fy, cx, cy, W, H):
    """gc_model.py:117-155: flipped world->camera matrix, P@V, tile bounds."""
    c2w = np.asarray(c2w, np.float32)
    R = c2w[:3, :3] @ np.diag(np.array([1, -1, -1], np.float32))   # :129-130
    T = c2w[:3, 3:4]
    Rinv = R.T
    Tinv = -Rinv @ T
    viewmat = np.eye(4, dtype=np.float32)
    viewmat[:3, :3] = Rinv
    viewmat[:3, 3:4] = Tinv
    fovx = 2 * math.atan(W / (2 * fx))                              # :143-144
    fovy = 2 * math.atan(H / (2 * fy))
    projmat = projection_matrix(0.001, 1000, fovx, fovy)            # :147
    tile_bounds = ((W + 15) // 16, (H + 15) // 16, 1)               # :150-155
    return viewmat, (projmat @ viewmat).astype(np.float32), tile_bounds


def _sigmoid(x):
    return (1.0 / (1.0 + np.exp(-x.astype(np.float64)))).astype(np.float32)


def preactivate(params, c2w, sh_degree):
    """gc_model.py:172-203 pre-ops. params: dict of numpy leaf tensors."""
    # exp correctly rounded (double, rounded once) and the norm as ((w*w + x*x) + y*y) + z*z in fp32: radii /
    # num_tiles_hit are compared BIT-EXACTLY with the fused CUDA path, which evaluates the same expressions
    scales = np.exp(params["scales"].astype(np.float64)).astype(np.float32)
    q = params["quats"].astype(np.float32)
    nq = np.sqrt(((q[:, 0] * q[:, 0] + q[:, 1] * q[:, 1]) + q[:, 2] * q[:, 2]) + q[:, 3] * q[:, 3])
    quats = q / nq[:, None]
    opac = _sigmoid(params["opacities"]).reshape(-1, 1)
    coeffs = np.concatenate([params["features_dc"][:, None, :], params["features_rest"]], axis=1)
    d = params["means"].astype(np.float32) - np.asarray(c2w, np.float32)[:3, 3][None]
    dirs = d / np.linalg.norm(d, axis=-1, keepdims=True)
    return scales, quats, opac, coeffs.astype(np.float32), dirs.astype(np.float32)


def render_view(params, cam, sh_degree=3, bg=(0.1, 0.2, 0.3), training=True):
    """CPU mirror of GaussCtrlModel.get_outputs (gc_model.py:77-241) for one camera.

    cam: dict(c2w(3|4,4), fx, fy, cx, cy, W, H). Returns (outputs dict, ctx for backward)."""
    W, H = int(cam["W"]), int(cam["H"])
    V, PV, tb = camera_matrices(cam["c2w"], cam["fx"], cam["fy"], cam["cx"], cam["cy"], W, H)
    scales, quats, opac, coeffs, dirs = preactivate(params, cam["c2w"], sh_degree)
    bg = np.asarray(bg, np.float32)
    xys, depths, radii, conics, hit, cov3d = project_fwd(
        params["means"], scales, 1.0, quats, V[:3, :], PV, cam["fx"], cam["fy"], cam["cx"], cam["cy"], H, W, tb)
    if sh_degree > 0:
        sh = sh_fwd(sh_degree, dirs, coeffs)
        rgbs = np.maximum(sh + np.float32(0.5), np.float32(0.0))
    else:
        sh = None
        rgbs = _sigmoid(coeffs[:, 0, :])
    b = bin_and_sort(xys, depths, radii, hit, tb)
    img, fT, fidx, pairs = raster_fwd(b, xys, conics, rgbs, opac, bg, H, W, tb)
    alpha = (1.0 - fT)[..., None]
    out = {"rgb": np.minimum(img, np.float32(1.0)), "accumulation": alpha, "depth": None,
           "pairs": pairs, "I": b.I, "visible": int((radii > 0).sum())}
    if not training:                                                   # gc_model.py:224-238
        dcol = np.repeat(depths[:, None], 3, axis=1)
        dimg, _, _, _ = raster_fwd(b, xys, conics, dcol, opac, np.zeros(3, np.float32), H, W, tb)
        d = dimg[..., 0:1].copy()
        m = alpha > 0
        d[m] = d[m] / alpha[m]
        d[alpha == 0] = 1000.0
        out["depth"] = d
    ctx = dict(V=V, PV=PV, tb=tb, scales=scales, quats=quats, opac=opac, coeffs=coeffs, dirs=dirs,
               xys=xys, depths=depths, radii=radii, conics=conics, hit=hit, sh=sh, rgbs=rgbs, b=b,
               img=img, fT=fT, fidx=fidx, bg=bg, cam=cam, sh_degree=sh_degree)
    return out, ctx


def render_view_backward(params, ctx, v_rgb, v_alpha=None):
    """Backward of render_view w.r.t. the six leaf tensors (+ xys.grad), chaining the oracle's
    exact-derivative operators through the caller's pre-ops (gc_trainer.py:275)."""
    cam, tb = ctx["cam"], ctx["tb"]
    W, H = int(cam["W"]), int(cam["H"])
    v_img = np.where(ctx["img"] > 1.0, 0.0, v_rgb).astype(np.float32)   # clamp(max=1), :222
    va = None if v_alpha is None else np.asarray(v_alpha, np.float32).reshape(H, W)
    v_xys, v_con, v_col, v_op = raster_bwd(ctx["b"], ctx["xys"], ctx["conics"], ctx["rgbs"], ctx["opac"],
                                           ctx["bg"], H, W, tb, ctx["fT"], ctx["fidx"], v_img, va)
    g = {"xys": v_xys}
    K = ctx["coeffs"].shape[1]
    if ctx["sh_degree"] > 0:
        v_sh = np.where(ctx["sh"] + np.float32(0.5) > 0, v_col, 0.0).astype(np.float32)
        v_coef = sh_bwd(ctx["sh_degree"], K, ctx["dirs"], v_sh)
    else:
        s = ctx["rgbs"].astype(np.float64)
        v_coef = np.zeros((v_col.shape[0], K, 3), np.float32)
        v_coef[:, 0, :] = (v_col * s * (1 - s)).astype(np.float32)
    g["features_dc"] = v_coef[:, 0, :]
    g["features_rest"] = v_coef[:, 1:, :]
    o = ctx["opac"].astype(np.float64)
    g["opacities"] = (v_op * o * (1 - o)).astype(np.float32)
    vm, vs, vq = project_bwd(params["means"], ctx["scales"], 1.0, ctx["quats"], ctx["V"][:3, :], ctx["PV"],
                             cam["fx"], cam["fy"], cam["cx"], cam["cy"], H, W, ctx["radii"],
                             v_xys, np.zeros_like(ctx["depths"]), v_con)
    g["means"] = vm
    g["scales"] = (vs * ctx["scales"]).astype(np.float32)               # d exp
    q = params["quats"].astype(np.float64)
    nq = np.linalg.norm(q, axis=-1, keepdims=True)
    qn = q / nq
    vq64 = vq.astype(np.float64)
    g["quats"] = ((vq64 - qn * (qn * vq64).sum(-1, keepdims=True)) / nq).astype(np.float32)
    return g


def noise_mask(points, verts, tris, view, proj, depth, threshold):
    """orc_noise_mask: CPU restatement of ad_render.py:273-333 + shaders :384-479 (sequential draw order).
    points (P,3), verts (V,3), tris (F,3) int32, view / proj 4x4 row-major, depth (H,W) -> (H,W,4) uint8."""
    points, verts, depth = _f(points).reshape(-1, 3), _f(verts).reshape(-1, 3), _f(depth)
    tris = np.ascontiguousarray(tris, np.int32).reshape(-1, 3)
    Vm, Pm = _f(view).reshape(-1), _f(proj).reshape(-1)
    assert Vm.size == 16 and Pm.size == 16 and depth.ndim == 2
    H, W = depth.shape
    out = np.zeros((H, W, 4), np.uint8)
    lib().orc_noise_mask(C.c_int(points.shape[0]), _p(points), C.c_int(verts.shape[0]), _p(verts),
                         C.c_int(tris.shape[0]), _p(tris), _p(Vm), _p(Pm), C.c_int(H), C.c_int(W), _p(depth),
                         C.c_float(threshold), _p(out))
    return out
