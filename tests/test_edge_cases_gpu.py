"""Edge cases the reference's callers can produce (empty / fully culled Gaussian sets, ragged image sizes, uint8
colours, default background), through the gsplat-compatible operators and the fused model."""
import numpy as np
import pytest
import torch

from gaussctrl_exp_b200 import project_gaussians, rasterize_gaussians, scene, spherical_harmonics
from gaussctrl_exp_b200.model import GaussianModel, camera_matrices
from util import np_params, small_cam

pytestmark = pytest.mark.gpu


def _cam_mats(cam, dev):
    V, P, _, _, tb = camera_matrices(cam.c2w, cam.fx, cam.fy, cam.W, cam.H)
    return V[:3].contiguous().to(dev), (P @ V).to(dev), tb


def test_empty_gaussian_set_operators(cuda_dev):
    dev = cuda_dev
    cam = small_cam(50, 37, 0)
    Vd, PVd, tb = _cam_mats(cam, dev)
    e = lambda *s, dt=torch.float32: torch.zeros(s, device=dev, dtype=dt)
    means = e(0, 3).requires_grad_()
    xys, depths, radii, conics, hit, cov3d = project_gaussians(means, e(0, 3), 1, e(0, 4), Vd, PVd, cam.fx, cam.fy, cam.cx,
                                                               cam.cy, cam.H, cam.W, tb)
    assert xys.shape == (0, 2) and radii.dtype == torch.int32 and hit.shape == (0,) and cov3d.shape == (0, 6)
    cols = spherical_harmonics(3, e(0, 3), e(0, 16, 3))
    assert cols.shape == (0, 3)
    bg = torch.tensor([0.2, 0.4, 0.6], device=dev)
    out, alpha = rasterize_gaussians(xys, depths, radii, conics, hit, cols, e(0, 1), cam.H, cam.W, background=bg,
                                     return_alpha=True)
    assert out.shape == (cam.H, cam.W, 3) and torch.allclose(out, bg.expand(cam.H, cam.W, 3)) and (alpha == 0).all()
    out.sum().backward()
    assert means.grad is None or means.grad.shape == (0, 3)


@pytest.mark.parametrize("fused", [True, False])
def test_empty_and_fully_culled_model(cuda_dev, fused):
    dev = cuda_dev
    cam = small_cam(64, 48, 1)
    bg = torch.tensor([0.1, 0.2, 0.3])
    p = {k: v[:0] for k, v in scene.make_scene(8, 0).items()}
    if fused:                                     # literal path: the caller's own early-out covers the empty set
        m = GaussianModel(p, sh_degree=3, background_color=bg, fused=True).to(dev)
        m.train()
        out = m(cam)
        assert out["rgb"].shape == (48, 64, 3) and torch.allclose(out["rgb"], bg.to(dev).expand(48, 64, 3))
        assert (out["accumulation"] == 0).all()
        out["rgb"].sum().backward()
        ev = m.get_outputs_for_camera(cam)
        assert torch.allclose(ev["rgb"], bg.to(dev).expand(48, 64, 3)) and (ev["depth"] == 1000).all()
    # every Gaussian behind the camera: culled by the clip threshold -> gc_model.py:189-190 returns {'rgb': background}
    # and nothing else, on the literal AND the fused path (the fused one reads it off the intersection count)
    q = scene.make_scene(500, 2, s0=0.02)
    q["means"] = q["means"] * 0.05 + cam.c2w[:3, 3] * 2.0            # twice as far from the origin as the camera, behind it
    m = GaussianModel(q, sh_degree=3, background_color=bg, fused=fused).to(dev)
    m.train()
    out = m(cam)
    assert set(out) == {"rgb"} and out["rgb"].shape == (48, 64, 3)
    assert torch.allclose(out["rgb"], bg.to(dev).expand(48, 64, 3))
    ev = m.get_outputs_for_camera(cam)
    assert set(ev) == {"rgb"} and torch.allclose(ev["rgb"], bg.to(dev).expand(48, 64, 3))
    if fused:
        # speculative forward: the count is not on the host yet, so the view keeps its full dict (documented deviation):
        # background, zero accumulation, zero gradients
        m.speculate = True
        out = m(cam)
        assert torch.allclose(out["rgb"], bg.to(dev).expand(48, 64, 3)) and (out["accumulation"] == 0).all()
        out["rgb"].sum().backward()
        assert all(float(getattr(m, n).grad.abs().max()) == 0.0 for n in ("means", "scales", "quats", "opacities"))


@pytest.mark.parametrize("W,H", [(1, 1), (17, 9), (16, 16), (33, 16), (250, 3)])
def test_ragged_image_sizes(oracle, cuda_dev, W, H):
    """Image sizes that are not multiples of the 16-pixel tile, down to a single pixel."""
    N = 400
    p = np_params(N, seed=3, s0=0.08)
    cam = small_cam(W, H, 3, f=max(40.0, 1.2 * max(W, H)))
    from util import forward_parity
    ref, ctx = oracle.render_view(p, cam.as_dict(), 2, training=False)
    m = GaussianModel({k: torch.from_numpy(v) for k, v in p.items()}, sh_degree=2, background_color=torch.tensor([0.1, 0.2, 0.3]),
                      step=30000).to(cuda_dev)
    out = m.get_outputs_for_camera(cam)
    assert out["rgb"].shape == (H, W, 3)
    # deviations only where the oracle's own alpha / T decision sat on its threshold, then at most one alpha quantum
    forward_parity(oracle, ctx, out["rgb"].cpu().numpy(), ref["rgb"], out["accumulation"].cpu().numpy(),
                   ref["accumulation"], H, W)


def test_uint8_colors_and_default_background(oracle, cuda_dev):
    dev = cuda_dev
    from util import projected
    d = projected(oracle, N=3000, W=96, H=64, seed=4, s0=0.04)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    rng = np.random.default_rng(0)
    c8 = rng.integers(0, 256, (d["N"], 3), dtype=np.uint8)
    args = (t(d["xys"]), t(d["depths"]), t(d["radii"]), t(d["conics"]), t(d["hit"]))
    opac = t(d["opac"].reshape(-1, 1))
    a = rasterize_gaussians(*args, t(c8), opac, d["H"], d["W"])                       # uint8 -> /255, background ones
    b = rasterize_gaussians(*args, t(c8.astype(np.float32) / 255), opac, d["H"], d["W"], background=torch.ones(3, device=dev))
    assert torch.allclose(a, b, atol=2e-6, rtol=0)      # torch divides by 255 as a multiplication by 1/255 on the GPU
    bb = oracle.bin_and_sort(d["xys"], d["depths"], d["radii"], d["hit"], d["tb"])
    from util import forward_parity
    ref, fT, fidx, _ = oracle.raster_fwd(bb, d["xys"], d["conics"], c8.astype(np.float32) / 255, d["opac"].reshape(-1),
                                         np.ones(3, np.float32), d["H"], d["W"], d["tb"])
    ctx = {"b": bb, "xys": d["xys"], "conics": d["conics"], "opac": d["opac"], "tb": d["tb"], "fidx": fidx}
    a2, alpha = rasterize_gaussians(*args, t(c8), opac, d["H"], d["W"], return_alpha=True)
    forward_parity(oracle, ctx, a2.cpu().numpy(), ref, alpha.cpu().numpy(), 1 - fT, d["H"], d["W"])


def test_pair_count_beyond_int32_is_reported_not_wrapped(cuda_dev):
    """More than 2^31 - 1 (tile, Gaussian) pairs in one view (every splat covers the whole 1080p frame): the device
    prefix saturates instead of wrapping negative, nothing is written, and the host raises instead of returning a
    view built on a wrapped count (gsplat's int32 cumsum wraps silently here)."""
    from gaussctrl_exp_b200 import _lib, utils
    dev = cuda_dev
    W, H = 1920, 1080
    tx, ty = (W + 15) // 16, (H + 15) // 16
    T = tx * ty                                              # 8 160 tiles
    N = (2**31) // T + 40_000                                # ~303 K splats x 8 160 tiles = 2.47 G pairs
    xys = torch.tensor([W / 2, H / 2], device=dev).repeat(N, 1)
    depths = torch.linspace(1.0, 2.0, N, device=dev)
    radii = torch.full((N,), 4096, device=dev, dtype=torch.int32)
    hit = torch.full((N,), T, device=dev, dtype=torch.int32)
    assert N * T > 2**31
    with pytest.raises(_lib.GsError, match="int32 count saturated"):
        utils.bin_fast(xys, depths, radii, hit, (tx, ty, 1))
    with pytest.raises(_lib.GsError, match="int32 count saturated"):
        utils.compute_cumulative_intersects(hit)
    # just below the limit the count is exact (and merely does not fit the first capacity)
    n_ok = (2**31 - 2) // T
    L = _lib.lib()
    total = torch.zeros(1, device=dev, dtype=torch.int32)
    cum = torch.empty(n_ok, device=dev, dtype=torch.int32)
    _lib.check(L.gs_bin_scan_static(n_ok, _lib.ptr(hit[:n_ok].contiguous()), _lib.ptr(cum), _lib.ptr(total),
                                    _lib.stream_ptr()), "gs_bin_scan_static")
    assert int(total.item()) == n_ok * T == int(cum[-1].item())


def test_binning_on_two_streams_keeps_separate_scratch(oracle, cuda_dev):
    """The operator-API scratch is per (device, stream): two streams binning different views concurrently get the
    results of a serial run."""
    from gaussctrl_exp_b200 import utils
    dev = cuda_dev
    g = torch.Generator().manual_seed(3)
    views = []
    for k in range(2):
        N, W, H = 60_000 + 7_000 * k, 320, 240
        tx, ty = (W + 15) // 16, (H + 15) // 16
        xys = (torch.rand(N, 2, generator=g) * torch.tensor([W, H])).to(dev)
        depths = (torch.rand(N, generator=g) + 0.5).to(dev)
        radii = torch.randint(0, 30, (N,), generator=g, dtype=torch.int32).to(dev)
        tc, tr = xys / 16.0, radii.float() / 16.0             # gs_tile_bbox: truncation, fp32, in this order
        x0, x1 = (tc[:, 0] - tr).trunc().clamp(0, tx).int(), (tc[:, 0] + tr + 1.0).trunc().clamp(0, tx).int()
        y0, y1 = (tc[:, 1] - tr).trunc().clamp(0, ty).int(), (tc[:, 1] + tr + 1.0).trunc().clamp(0, ty).int()
        hit = ((x1 - x0) * (y1 - y0) * (radii > 0)).int()
        views.append((xys, depths, radii, hit, (tx, ty, 1)))
    want = [utils.bin_fast(*v) for v in views]
    torch.cuda.synchronize()
    streams = [torch.cuda.Stream(dev) for _ in views]
    for rep in range(3):
        got = []
        for v, s in zip(views, streams):
            with torch.cuda.stream(s):
                got.append(utils.bin_fast(*v, defer=True))
        for (ids, bins, pend), (wids, wbins, wI) in zip(got, want):
            I, ok = pend.wait() if not isinstance(pend, int) else (pend, True)
            assert ok and I == wI
            torch.cuda.synchronize()
            assert torch.equal(bins, wbins) and torch.equal(ids[:I], wids[:I])
