"""Noise-mask renderer, CUDA side: gs_noise_mask through the C ABI against the C oracle, bit for bit
(integer coverage, fp32 arithmetic in a fixed order without FMA contraction on both sides)."""
import numpy as np
import pytest
import torch

from gaussctrl_exp_b200 import noise_mask as nm
from gaussctrl_exp_b200 import scene
from gaussctrl_exp_b200.model import GaussianModel
from test_noise_mask_cpu import _camera

pytestmark = pytest.mark.gpu


def _t(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


@pytest.mark.parametrize("W,H,R,thr,seed", [(128, 128, 40, nm.FRAG_DEPTH_THRESHOLD, 0), (160, 96, 30, 0.05, 1),
                                            (75, 50, 24, 5.0, 2)])
def test_noise_mask_bit_exact(oracle, cuda_dev, W, H, R, thr, seed):
    cam, c2w, view, proj = _camera(W, H, seed=seed)
    rng = np.random.default_rng(seed)
    noise = rng.random((R, R, R)).astype(np.float32)
    pts = nm.uniform_points_in_cube(2.0, R)[np.where(noise.reshape(-1) > 0.8)[0]].astype(np.float32)
    verts, tris = nm.sphere_mesh(0.015 if thr < 1 else 0.03, 8, 8)
    depth = (1.5 + 0.4 * np.sin(np.arange(W) / 9.0)[None, :] * np.cos(np.arange(H) / 7.0)[:, None]).astype(np.float32)
    ref = oracle.noise_mask(pts, verts, tris, view, proj, depth, thr)
    got = nm.rasterize_noise_mask(_t(pts, cuda_dev), _t(verts, cuda_dev), _t(tris, cuda_dev), view, proj,
                                  _t(depth, cuda_dev), thr).cpu().numpy()
    assert (ref[..., 3] > 0).sum() > 50
    assert np.array_equal(got, ref)


def test_renderer_on_rendered_depth_full_size(oracle, cuda_dev):
    """The reference's use (gc_pipeline.py:236-250): depth of a 3DGS eval render -> noise mask of the same camera;
    512x512, 100^3 noise voxels (about 200 K spheres, 25 M triangles)."""
    W = H = 512
    cam = scene.make_cameras(1, 0)[0]
    model = GaussianModel(scene.make_scene(200_000, 0, s0=0.01), sh_degree=3,
                          background_color=torch.tensor([0.1, 0.2, 0.3])).to(cuda_dev)
    out = model.get_outputs_for_camera(cam)
    depth = out["depth"]
    r = nm.MultiViewNoiseRenderer(W, H, device=cuda_dev)
    noise = np.random.default_rng(7).random((100, 100, 100)).astype(np.float32)
    r.setup_noise(noise)
    assert r.pts.shape[0] > 150_000
    img = r.render_noise_SS(depth, out["mat_view"], out["mat_proj"], out["mat_c2w"])
    assert img.shape == (H, W, 4) and img.dtype == np.uint8
    c2w = np.eye(4)
    c2w[:3, :4] = cam.c2w.numpy()[:3, :4]
    ref = oracle.noise_mask(r.pts, r.vertices, r.indices, r.c2w_to_viewmat(c2w), out["mat_proj"].numpy(),
                            depth.cpu().numpy()[..., 0], nm.FRAG_DEPTH_THRESHOLD)
    assert (ref[..., 3] > 0).mean() > 0.01
    assert np.array_equal(img, ref)
    # kept fragments lie in the shell in front of the rendered surface: colour -> world point -> distance check
    rr, cc = np.nonzero(img[..., 3])
    world = img[rr, cc, :3].astype(np.float64) / 255 * 2 - 1
    dist = np.linalg.norm(world - c2w[:3, 3], axis=1)
    d = depth.cpu().numpy()[rr, cc, 0]
    assert np.all(dist <= d + 0.02) and np.all(dist >= d - nm.FRAG_DEPTH_THRESHOLD - 0.02)
    assert r.binarize_mask(img).sum() == len(rr)


def test_noise_mask_edge_cases_and_errors(oracle, cuda_dev):
    from gaussctrl_exp_b200 import _lib
    W, H = 64, 48
    cam, c2w, view, proj = _camera(W, H)
    verts, tris = nm.sphere_mesh(0.05, 8, 8)
    depth = _t(np.full((H, W), 1.5, np.float32), cuda_dev)
    v, t = _t(verts, cuda_dev), _t(tris, cuda_dev)
    empty = nm.rasterize_noise_mask(torch.zeros((0, 3), device=cuda_dev), v, t, view, proj, depth, 0.1)
    assert empty.shape == (H, W, 4) and not empty.any()
    behind = _t((c2w[:3, 3] * 1.5)[None].astype(np.float32), cuda_dev)
    assert not nm.rasterize_noise_mask(behind, v, t, view, proj, depth, 1000.0).any()
    # overlapping instances: the later primitive wins, exactly like the sequential oracle
    q = np.array([[0.0, 0.0, 0.0], [0.02, 0.0, 0.0], [0.01, 0.01, 0.0]], np.float32)
    d3 = np.full((H, W), 3.0, np.float32)
    got = nm.rasterize_noise_mask(_t(q, cuda_dev), v, t, view, proj, _t(d3, cuda_dev), 3.0).cpu().numpy()
    assert np.array_equal(got, oracle.noise_mask(q, verts, tris, view, proj, d3, 3.0))
    with pytest.raises(ValueError):
        nm.rasterize_noise_mask(torch.zeros((1, 3)), v, t, view, proj, depth)
    big_v, big_t = nm.sphere_mesh(0.05, 16, 16)                        # 289 vertices > 128
    with pytest.raises(_lib.GsError):
        nm.rasterize_noise_mask(torch.zeros((1, 3), device=cuda_dev), _t(big_v, cuda_dev), _t(big_t, cuda_dev), view, proj, depth)
    r = nm.MultiViewNoiseRenderer(W, H, device=cuda_dev)
    r.resolution = 10
    r.setup_noise(np.ones((10, 10, 10), np.float32))
    with pytest.raises(ValueError):
        r.render_noise_SS(np.zeros((H + 1, W), np.float32), None, proj, c2w)
