"""Noise-mask renderer, CPU side: host helpers against the reference's own numpy functions
(executed unmodified by tests/golden/gen_golden.py) and the C oracle against an independent float64 ray-casting
restatement of the shaders (/root/reference/gaussctrl/ad_render.py:384-479). The GL draw itself is unavailable on this
box, so the rasterisation rules stay "parity unpinned"."""
import os

import numpy as np
import pytest

from gaussctrl_exp_b200 import noise_mask as nm
from gaussctrl_exp_b200 import scene
from gaussctrl_exp_b200.model import camera_matrices

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _camera(W, H, seed=0, idx=0, f=150.0):
    cam = scene.make_cameras(idx + 1, seed, intrinsics=dict(fx=f, fy=f * 0.99, cx=W / 2, cy=H / 2, W=W, H=H))[idx]
    projmat0 = camera_matrices(cam.c2w, cam.fx, cam.fy, W, H)[3].numpy()
    c2w = np.eye(4)
    c2w[:3, :4] = cam.c2w.numpy()[:3, :4]
    view = nm.compute_view_matrix_ogl(*nm.extract_camera_parameters(c2w))
    return cam, c2w, view, projmat0


def test_helpers_match_reference_goldens():
    g = np.load(os.path.join(GOLD, "reference_noise_helpers.npz"))
    v, i = nm.sphere_mesh(0.015, 8, 8)
    assert v.dtype == np.float32 and np.array_equal(v.reshape(-1), g["sphere_v"])
    assert np.array_equal(i.reshape(-1).astype(np.uint32), g["sphere_i"])
    v2, i2 = nm.sphere_mesh(0.3, 5, 7)
    assert np.array_equal(v2.reshape(-1), g["sphere2_v"]) and np.array_equal(i2.reshape(-1).astype(np.uint32), g["sphere2_i"])
    assert np.array_equal(nm.uniform_points_in_cube(2.0, 6), g["cube"])
    for c2w, view in zip(g["c2ws"], g["views"]):
        pos, up, fw = nm.extract_camera_parameters(c2w)
        assert np.abs(nm.compute_view_matrix_ogl(pos, up, fw) - view).max() < 1e-12
    assert nm.FRAG_DEPTH_THRESHOLD == pytest.approx(0.0161, abs=1e-7)


def _raycast_reference(points, verts, tris, view, proj, depth, thr):
    """Literal float64 restatement of the two shaders, by ray casting instead of edge functions: for every pixel the
    view ray is intersected with every triangle (Moeller-Trumbore), the fragment's view-space distance is compared
    with the depth texel, and the last primitive in draw order that survives gives the colour."""
    H, W = depth.shape
    Pinv = np.linalg.inv(proj.astype(np.float64))
    r, c = np.mgrid[0:H, 0:W]
    x, y = W - 1 - c, H - 1 - r                                    # GL window pixel of the returned pixel (r, c)
    ndc = np.stack([(x + 0.5) / W * 2 - 1, (y + 0.5) / H * 2 - 1, np.zeros_like(x, float), np.ones_like(x, float)], -1)
    q = ndc.reshape(-1, 4) @ Pinv.T
    dirs = q[:, :3] / q[:, 3:4]                                     # a view-space point on each pixel's ray
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    dtex = np.clip(depth.astype(np.float64), 0, 1000).reshape(-1)
    out = np.zeros((H * W, 4), np.uint8)
    V64 = view.astype(np.float64)
    for p in points.astype(np.float64):
        world = verts.astype(np.float64) + p
        vp = world @ V64[:3, :3].T + V64[:3, 3]
        col = (world + 1) / 2
        for ia, ib, ic in tris:
            a, e1, e2 = vp[ia], vp[ib] - vp[ia], vp[ic] - vp[ia]
            pv = np.cross(dirs, e2)
            det = pv @ e1
            ok = np.abs(det) > 1e-18
            inv = np.where(ok, 1.0 / np.where(ok, det, 1), 0)
            u = (pv @ (-a)) * inv
            qv = np.cross(-a, e1)
            v = (dirs @ qv) * inv
            t = (qv @ e2) * inv
            hit = ok & (u >= 0) & (v >= 0) & (u + v <= 1) & (t > 0)
            if not hit.any():
                continue
            keep = hit & (dtex - t >= 0) & (dtex - t <= thr)        # |t * dir| = t
            if not keep.any():
                continue
            k = np.nonzero(keep)[0]
            w = np.stack([1 - u[k] - v[k], u[k], v[k]], -1)
            rgb = np.floor(np.clip(w @ col[[ia, ib, ic]], 0, 1) * 255 + 0.5)
            out[k, :3] = rgb
            out[k, 3] = 255
    return out.reshape(H, W, 4)


def test_oracle_matches_float64_raycast(oracle):
    W, H = 96, 72
    cam, c2w, view, proj = _camera(W, H, seed=2)
    rng = np.random.default_rng(4)
    pts = (rng.random((70, 3)) * 0.5 - 0.25).astype(np.float32)
    verts, tris = nm.sphere_mesh(0.04, 8, 8)
    dist = np.linalg.norm(pts - c2w[:3, 3], axis=1)
    depth = np.full((H, W), float(np.median(dist)), np.float32) + rng.random((H, W)).astype(np.float32) * 0.05
    img = oracle.noise_mask(pts, verts, tris, view, proj, depth, 0.06)
    ref = _raycast_reference(pts, verts, tris, view, proj, depth, 0.06)
    cov, cov_ref = img[..., 3] > 0, ref[..., 3] > 0
    assert cov_ref.sum() > 150
    # coverage differs only where a pixel centre sits within the 1/256 px snapping distance of a triangle edge or a
    # fragment within float rounding of the shell limits
    assert (cov != cov_ref).sum() <= 0.01 * cov_ref.sum()
    both = cov & cov_ref
    diff = np.abs(img[both].astype(int) - ref[both].astype(int)).max(axis=1)
    assert (diff > 1).mean() < 0.01          # same primitive won: colours agree to one 8-bit step


def test_mask_is_aligned_with_the_gaussian_projection(oracle):
    """The point of the reference's flips (ad_render.py:110-111, :326-327): a world point lands on the same (row, col)
    in the noise mask as in the splatted image it is compared with."""
    W, H = 160, 128
    cam, c2w, view, proj = _camera(W, H, seed=1, idx=2)
    V, PV, tb = oracle.camera_matrices(cam.c2w.numpy(), cam.fx, cam.fy, cam.cx, cam.cy, W, H)
    verts, tris = nm.sphere_mesh(0.015, 8, 8)
    for p in ([0.2, 0.1, -0.15], [-0.3, 0.25, 0.1], [0.0, -0.35, 0.3]):
        p = np.array([p], np.float32)
        xys = oracle.project_fwd(p, np.full((1, 3), 0.01, np.float32), 1.0, np.array([[1, 0, 0, 0]], np.float32), V[:3], PV,
                                 cam.fx, cam.fy, cam.cx, cam.cy, H, W, tb)[0]
        d = np.linalg.norm(p[0] - c2w[:3, 3])
        img = oracle.noise_mask(p, verts, tris, view, proj, np.full((H, W), d + 0.001, np.float32), 0.05)
        rr, cc = np.nonzero(img[..., 3])
        assert len(rr) >= 4
        assert abs(cc.mean() - xys[0, 0]) < 0.75 and abs(rr.mean() - xys[0, 1]) < 0.75
        # colour = (world position + 1) / 2
        assert np.abs(img[rr, cc, :3].astype(float).mean(0) - (p[0] + 1) / 2 * 255).max() < 3


def test_oracle_edge_cases(oracle):
    W, H = 64, 48
    cam, c2w, view, proj = _camera(W, H)
    verts, tris = nm.sphere_mesh(0.05, 8, 8)
    depth = np.full((H, W), 1.5, np.float32)
    empty = oracle.noise_mask(np.zeros((0, 3), np.float32), verts, tris, view, proj, depth, 0.1)
    assert empty.shape == (H, W, 4) and not empty.any()
    p = np.array([[0.0, 0.0, 0.0]], np.float32)
    assert not oracle.noise_mask(p, verts, tris, view, proj, np.full((H, W), 1000.0, np.float32), 0.1).any()   # shell far behind
    assert not oracle.noise_mask(p, verts, tris, view, proj, np.full((H, W), 0.5, np.float32), 0.1).any()      # occluded
    behind = (c2w[:3, 3] * 1.5)[None].astype(np.float32)              # behind the camera: w <= 0, nothing drawn
    assert not oracle.noise_mask(behind, verts, tris, view, proj, depth, 1000.0).any()
    # draw order: of two overlapping instances the later one owns the shared pixels (no depth test)
    q = np.array([[0.0, 0.0, 0.0], [0.02, 0.0, 0.0]], np.float32)
    a = oracle.noise_mask(q, verts, tris, view, proj, np.full((H, W), 3.0, np.float32), 3.0)
    b = oracle.noise_mask(q[1:], verts, tris, view, proj, np.full((H, W), 3.0, np.float32), 3.0)
    own = b[..., 3] > 0
    assert own.sum() > 10 and np.array_equal(a[own], b[own])
    assert nm.MultiViewNoiseRenderer.binarize_mask(a).shape == (H, W, 1)
    assert nm.MultiViewNoiseRenderer.binarize_mask(a).sum() == (a[..., :3] != 0).any(-1).sum()


def test_renderer_refuses_cpu_and_missing_noise(tmp_path):
    r = nm.MultiViewNoiseRenderer(32, 32, device="cpu")
    r.resolution = 8
    with pytest.raises(FileNotFoundError):
        r.setup_noise(str(tmp_path / "absent.npy"))
    with pytest.raises(AssertionError):
        r.setup_noise(np.zeros((4, 4, 4), np.float32))
    with pytest.raises(ValueError):
        r.setup_noise(np.ones((8, 8, 8), np.float32))                  # CPU device: no fallback
    with pytest.raises(RuntimeError):
        nm.MultiViewNoiseRenderer(32, 32).render_noise_SS(np.zeros((32, 32)), None, np.eye(4), np.eye(4))
