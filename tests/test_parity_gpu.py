"""GPU parity tests: every CUDA stage, called through the C ABI (ctypes wrappers), against the CPU
oracle on the same seeded inputs. Integer outputs (radii, tiles_hit, keys, ids, tile ranges,
final_idx) must be BIT-EXACT; floating point within 1e-4 relative L-inf (BASELINE.json north_star)."""
import numpy as np
import pytest
import torch

from util import np_params, projected, rel_linf, small_cam, to_dev

pytestmark = pytest.mark.gpu

TOL = 1e-4


def _t(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def _project_gpu(d, dev):
    from gaussctrl_exp_b200 import project_gaussians
    cam = d["cam"]
    means = _t(d["p"]["means"], dev).requires_grad_()
    scales = _t(d["scales"], dev).requires_grad_()
    quats = _t(d["quats"], dev).requires_grad_()
    out = project_gaussians(means, scales, 1.0, quats, _t(d["V"][:3], dev), _t(d["PV"], dev), cam.fx, cam.fy, cam.cx,
                            cam.cy, d["H"], d["W"], d["tb"])
    return (means, scales, quats), out


@pytest.mark.parametrize("N,W,H,seed,s0", [(5000, 96, 64, 1, 0.03), (20000, 512, 512, 2, 0.01), (777, 50, 37, 3, 0.2)])
def test_project_fwd(oracle, cuda_dev, N, W, H, seed, s0):
    d = projected(oracle, N=N, W=W, H=H, seed=seed, s0=s0)
    _, (xys, depths, radii, conics, hit, cov3d) = _project_gpu(d, cuda_dev)
    assert radii.dtype == torch.int32 and hit.dtype == torch.int32
    # integers: bit-exact
    assert np.array_equal(radii.cpu().numpy(), d["radii"])
    assert np.array_equal(hit.cpu().numpy(), d["hit"])
    # floats: the forward is compiled without FMA contraction and mirrors the oracle's op order
    assert np.array_equal(xys.detach().cpu().numpy(), d["xys"])
    assert np.array_equal(depths.detach().cpu().numpy(), d["depths"])
    assert rel_linf(conics.detach().cpu().numpy(), d["conics"]) < 1e-6
    assert rel_linf(cov3d.detach().cpu().numpy(), d["cov3d"]) < 1e-6


def test_project_bwd(oracle, cuda_dev):
    d = projected(oracle, N=6000, W=96, H=64, seed=4, s0=0.03)
    cam, N = d["cam"], d["N"]
    (means, scales, quats), (xys, depths, radii, conics, hit, cov3d) = _project_gpu(d, cuda_dev)
    rng = np.random.default_rng(0)
    v_xys = rng.standard_normal((N, 2)).astype(np.float32)
    v_dep = rng.standard_normal((N,)).astype(np.float32)
    v_con = rng.standard_normal((N, 3)).astype(np.float32)
    loss = (xys * _t(v_xys, cuda_dev)).sum() + (depths * _t(v_dep, cuda_dev)).sum() + (conics * _t(v_con, cuda_dev)).sum()
    loss.backward()
    vm, vs, vq = oracle.project_bwd(d["p"]["means"], d["scales"], 1.0, d["quats"], d["V"][:3], d["PV"], cam.fx, cam.fy,
                                    cam.cx, cam.cy, d["H"], d["W"], d["radii"], v_xys, v_dep, v_con)
    assert rel_linf(means.grad.cpu().numpy(), vm) < TOL
    assert rel_linf(scales.grad.cpu().numpy(), vs) < TOL
    assert rel_linf(quats.grad.cpu().numpy(), vq) < TOL


@pytest.mark.parametrize("K,deg", [(16, 0), (16, 1), (16, 2), (16, 3), (25, 4), (9, 2), (4, 1)])
def test_sh(oracle, cuda_dev, K, deg):
    from gaussctrl_exp_b200 import spherical_harmonics
    rng = np.random.default_rng(K * 10 + deg)
    N = 3001
    dirs = rng.standard_normal((N, 3)).astype(np.float32)
    coeffs = rng.standard_normal((N, K, 3)).astype(np.float32)
    vcol = rng.standard_normal((N, 3)).astype(np.float32)
    c = _t(coeffs, cuda_dev).requires_grad_()
    out = spherical_harmonics(deg, _t(dirs, cuda_dev), c)
    assert rel_linf(out.detach().cpu().numpy(), oracle.sh_fwd(deg, dirs, coeffs)) < 1e-5
    (out * _t(vcol, cuda_dev)).sum().backward()
    assert rel_linf(c.grad.cpu().numpy(), oracle.sh_bwd(deg, K, dirs, vcol)) < 1e-5


@pytest.mark.parametrize("N,W,H,seed,s0", [(8000, 128, 80, 1, 0.03), (50000, 512, 512, 0, 0.005), (300, 1920, 1080, 2, 0.3),
                                           (5, 33, 17, 3, 0.05), (120000, 1920, 1080, 4, 0.004), (4097, 256, 256, 6, 0.02),
                                           (30000, 1920, 1080, 7, 0.03)])      # last: 15.6 M pairs -> segmented column scan
def test_binning_bit_exact(oracle, cuda_dev, N, W, H, seed, s0):
    from gaussctrl_exp_b200 import utils as U
    d = projected(oracle, N=N, W=W, H=H, seed=seed, s0=s0)
    tb = d["tb"]
    T = tb[0] * tb[1]
    b = oracle.bin_and_sort(d["xys"], d["depths"], d["radii"], d["hit"], tb)
    xys, depths = _t(d["xys"], cuda_dev), _t(d["depths"], cuda_dev)
    radii, hit = _t(d["radii"], cuda_dev), _t(d["hit"], cuda_dev)
    I, cum = U.compute_cumulative_intersects(hit)
    assert I == b.I and np.array_equal(cum.cpu().numpy(), b.cum)
    keys, ids, ks, vs, bins = U.bin_and_sort_gaussians(N, I, xys, depths, radii, cum, tb)
    assert np.array_equal(keys.cpu().numpy(), b.keys) and np.array_equal(ids.cpu().numpy(), b.ids)
    assert np.array_equal(ks.cpu().numpy(), b.keys_sorted)
    assert np.array_equal(vs.cpu().numpy(), b.ids_sorted)
    assert np.array_equal(bins.cpu().numpy(), b.tile_bins)
    # product paths: identical ids_sorted and tile ranges without 64-bit keys
    for impl in ("bucket", "lsd", "lsd3"):
        ids2, bins2, I2 = U.bin_fast(xys, depths, radii, hit, tb, impl=impl)
        assert I2 == b.I, impl
        assert np.array_equal(ids2[:I2].cpu().numpy(), b.ids_sorted), impl
        assert np.array_equal(bins2.cpu().numpy(), b.tile_bins), impl
        assert bins2.shape == (T, 2)


@pytest.mark.parametrize("N,W,H,s0", [(70000, 512, 512, 0.004), (9000, 1920, 1080, 0.05), (300, 64, 48, 0.1),
                                     (20000, 640, 640, 0.01)])      # 10-bit, 7+6-bit, 4-bit and 11-bit tile digits
def test_radix_cta_shapes_identical(oracle, cuda_dev, N, W, H, s0):
    """gs_set_radix_cfg: every CTA shape of the radix passes (depth sort x partition) gives the oracle's lists."""
    from gaussctrl_exp_b200 import _lib, utils as U
    d = projected(oracle, N=N, W=W, H=H, seed=11, s0=s0)
    tb = d["tb"]
    b = oracle.bin_and_sort(d["xys"], d["depths"], d["radii"], d["hit"], tb)
    t = [_t(d[k], cuda_dev) for k in ("xys", "depths", "radii", "hit")]
    L = _lib.lib()
    try:
        for cd in range(4):
            for cp in range(4):
                L.gs_set_radix_cfg(0, cd)
                L.gs_set_radix_cfg(1, cp)
                ids, bins, I = U.bin_fast(*t, tb, impl="lsd")
                assert I == b.I, (cd, cp)
                assert np.array_equal(ids[:I].cpu().numpy(), b.ids_sorted), (cd, cp)
                assert np.array_equal(bins.cpu().numpy(), b.tile_bins), (cd, cp)
    finally:
        L.gs_set_radix_cfg(0, -1)
        L.gs_set_radix_cfg(1, -1)
    assert L.gs_set_radix_cfg(0, 99) == -1 and L.gs_set_radix_cfg(2, 0) != 0      # query only / bad selector


def test_binning_big_tiles_and_depth_ties(oracle, cuda_dev):
    """Very long tile lists (60 K entries per tile) and exact depth ties (must come out in Gaussian-id order)."""
    from gaussctrl_exp_b200 import utils as U
    rng = np.random.default_rng(5)
    N, W, H = 60000, 48, 32                      # 3 x 2 tiles, every splat covers all of them: 60 K per tile
    tb = (3, 2, 1)
    xys = (rng.random((N, 2)) * [W, H]).astype(np.float32)
    depths = (rng.random(N) * 3 + 0.1).astype(np.float32)
    depths[rng.integers(0, N, 5000)] = np.float32(1.25)      # many exact ties
    depths[::7] = depths[3]
    radii = np.full(N, 200, np.int32)
    radii[::11] = 0                              # culled ones in between
    hit = np.where(radii > 0, 6, 0).astype(np.int32)
    b = oracle.bin_and_sort(xys, depths, radii, hit, tb)
    assert b.I == int(hit.sum()) and (b.tile_bins[:, 1] - b.tile_bins[:, 0]).max() > 50000
    t = [_t(a, cuda_dev) for a in (xys, depths, radii, hit)]
    for impl in ("bucket", "lsd", "lsd3"):
        ids2, bins2, I2 = U.bin_fast(*t, tb, impl=impl)
        assert I2 == b.I
        assert np.array_equal(bins2.cpu().numpy(), b.tile_bins), impl
        assert np.array_equal(ids2[:I2].cpu().numpy(), b.ids_sorted), impl
    # medium tiles: ~9 K per tile
    sel = np.arange(N) < 10000
    r2 = np.where(sel, radii, 0).astype(np.int32)
    h2 = np.where(r2 > 0, 6, 0).astype(np.int32)
    b = oracle.bin_and_sort(xys, depths, r2, h2, tb)
    ids2, bins2, I2 = U.bin_fast(t[0], t[1], _t(r2, cuda_dev), _t(h2, cuda_dev), tb, impl="bucket")
    assert I2 == b.I and np.array_equal(ids2[:I2].cpu().numpy(), b.ids_sorted)


@pytest.mark.parametrize("case", ["same_depth", "wide_range", "narrow_range", "tiny", "one_visible", "all_culled",
                                  "two_values", "peaky_1m"])
def test_depth_sort_corner_cases(oracle, cuda_dev, case):
    """Depth distributions that stress a depth sort's digit handling: every key equal, keys spread over 2^30 bit
    patterns, keys that differ only in their lowest bits, fewer keys than one chunk, one / no visible Gaussian, two
    distinct values, half of a 1 M scene on one surface. Oracle for the small ones; at 1 M the three product pipelines
    against each other."""
    from gaussctrl_exp_b200 import utils as U
    rng = np.random.default_rng(17)
    N, W, H = 30000, 64, 48
    tb = (4, 3, 1)
    if case == "tiny":
        N = 37
    elif case == "peaky_1m":
        N = 1_000_000
    xys = (rng.random((N, 2)) * [W, H]).astype(np.float32)
    radii = rng.integers(1, 12, N).astype(np.int32)
    radii[rng.random(N) < 0.1] = 0
    depths = (rng.random(N) * 3 + 0.2).astype(np.float32)
    if case == "same_depth":
        depths[:] = np.float32(1.5)
    elif case == "wide_range":
        depths = np.exp(rng.uniform(np.log(0.011), np.log(1e30), N)).astype(np.float32)
    elif case == "narrow_range":
        depths = np.float32(2.0) + (rng.integers(0, 300, N) * np.float32(2.0 ** -22)).astype(np.float32)
    elif case == "one_visible":
        radii[:] = 0
        radii[N // 2] = 3
    elif case == "all_culled":
        radii[:] = 0
    elif case == "two_values":
        depths = np.where(rng.random(N) < 0.5, np.float32(0.7), np.float32(4.0e10)).astype(np.float32)
    elif case == "peaky_1m":
        depths[: N // 2] = (1.0 + 1e-4 * rng.standard_normal(N // 2)).astype(np.float32)     # half the scene on one surface
    # tile counts as the projection would give them
    f16, one = np.float32(16), np.float32(1)
    tr = radii.astype(np.float32) / f16
    tcx, tcy = xys[:, 0] / f16, xys[:, 1] / f16
    x0 = np.clip((tcx - tr).astype(np.int32), 0, tb[0])
    x1 = np.clip((tcx + tr + one).astype(np.int32), 0, tb[0])
    y0 = np.clip((tcy - tr).astype(np.int32), 0, tb[1])
    y1 = np.clip((tcy + tr + one).astype(np.int32), 0, tb[1])
    hit = np.where(radii > 0, (x1 - x0) * (y1 - y0), 0).astype(np.int32)
    radii = np.where(hit > 0, radii, 0).astype(np.int32)
    t = [_t(a, cuda_dev) for a in (xys, depths, radii, hit)]
    ids, bins, I = U.bin_fast(*t, tb, impl="lsd")
    assert I == int(hit.sum())
    if case == "peaky_1m":
        for impl in ("lsd3", "bucket"):
            ids4, bins4, I4 = U.bin_fast(*t, tb, impl=impl)
            assert I4 == I and torch.equal(bins, bins4) and torch.equal(ids[:I], ids4[:I]), impl
    else:
        b = oracle.bin_and_sort(xys, depths, radii, hit, tb)
        assert b.I == I
        assert np.array_equal(bins.cpu().numpy(), b.tile_bins)
        assert np.array_equal(ids[:I].cpu().numpy(), b.ids_sorted)
        for impl in ("lsd3", "bucket"):
            ids4, bins4, I4 = U.bin_fast(*t, tb, impl=impl)
            assert I4 == I and torch.equal(bins, bins4) and torch.equal(ids[:I], ids4[:I]), impl


def _raster_case(oracle, N, W, H, seed, s0, C=3):
    d = projected(oracle, N=N, W=W, H=H, seed=seed, s0=s0)
    rng = np.random.default_rng(seed + 100)
    colors = rng.random((N, C)).astype(np.float32)
    bg = rng.random(C).astype(np.float32)
    b = oracle.bin_and_sort(d["xys"], d["depths"], d["radii"], d["hit"], d["tb"])
    return d, colors, bg, b


def _rctx(d, b, fidx):
    """What tests/util.py's parity helpers need to know about a compositing case."""
    return {"b": b, "xys": d["xys"], "conics": d["conics"], "opac": d["opac"], "tb": d["tb"], "fidx": fidx}


@pytest.mark.parametrize("N,W,H,seed,s0,C", [(3000, 96, 64, 1, 0.04, 3), (50000, 512, 512, 0, 0.005, 3),
                                             (2000, 70, 45, 2, 0.05, 3), (1500, 64, 48, 3, 0.05, 1),
                                             (1500, 64, 48, 4, 0.05, 5), (800, 48, 32, 5, 0.06, 20)])
def test_raster_fwd_bwd(oracle, cuda_dev, N, W, H, seed, s0, C):
    """rasterize_gaussians (C = 3: pack + packed cp.async kernels; other C: generic kernels) against the oracle:
    pixels deviate only where the oracle's own alpha / T decision sat on its threshold, gradients of Gaussians that
    touch no such pixel are within 1e-4 (tests/util.py)."""
    from gaussctrl_exp_b200 import rasterize_gaussians
    from util import forward_parity, gradient_parity
    dev = cuda_dev
    d, colors, bg, b = _raster_case(oracle, N, W, H, seed, s0, C)
    img, fT, fidx, pairs = oracle.raster_fwd(b, d["xys"], d["conics"], colors, d["opac"], bg, H, W, d["tb"])
    xys = _t(d["xys"], dev).requires_grad_()
    conics = _t(d["conics"], dev).requires_grad_()
    col = _t(colors, dev).requires_grad_()
    opac = _t(d["opac"], dev).requires_grad_()
    out, alpha = rasterize_gaussians(xys, _t(d["depths"], dev), _t(d["radii"], dev), conics, _t(d["hit"], dev), col,
                                     opac, H, W, background=_t(bg, dev), return_alpha=True)
    assert out.shape == (H, W, C) and alpha.shape == (H, W)
    ctx = _rctx(d, b, fidx)
    _, F, excl = forward_parity(oracle, ctx, out.detach().cpu().numpy(), img, alpha.detach().cpu().numpy(), 1 - fT, H, W)
    rng = np.random.default_rng(seed)
    v_out = rng.standard_normal((H, W, C)).astype(np.float32)
    v_a = rng.standard_normal((H, W)).astype(np.float32)
    ((out * _t(v_out, dev)).sum() + (alpha * _t(v_a, dev)).sum()).backward()
    g_xy, g_cn, g_co, g_op = oracle.raster_bwd(b, d["xys"], d["conics"], colors, d["opac"], bg, H, W, d["tb"], fT, fidx,
                                               v_out, v_a)
    gradient_parity(oracle, ctx, excl, {"xys": xys.grad.cpu().numpy(), "conics": conics.grad.cpu().numpy(),
                                        "colors": col.grad.cpu().numpy(), "opac": opac.grad.cpu().numpy()},
                    {"xys": g_xy, "conics": g_cn, "colors": g_co, "opac": g_op}, H, W)


@pytest.mark.parametrize("packed", [False, True])
def test_raster_final_idx_exact_and_no_alpha_grad(oracle, cuda_dev, packed):
    """Raw ABI: final_idx (an index) is bit-exact wherever no decision sat on its threshold; v_out_alpha = NULL path.
    packed = the cp.async kernels over 48-byte records (gs_pack_splats + gs_raster_*_packed, LPT tile order),
    otherwise the un-packed kernels (gs_raster_fwd / gs_raster_bwd)."""
    from gaussctrl_exp_b200 import _lib
    from util import forward_parity, gradient_parity
    dev = cuda_dev
    d, colors, bg, b = _raster_case(oracle, 4000, 112, 80, 9, 0.03)
    H, W, tb = d["H"], d["W"], d["tb"]
    T = tb[0] * tb[1]
    img, fT, fidx, _ = oracle.raster_fwd(b, d["xys"], d["conics"], colors, d["opac"], bg, H, W, tb)
    t = {k: _t(v, dev) for k, v in dict(xys=d["xys"], conics=d["conics"], colors=colors, opac=d["opac"][:, 0], bg=bg,
                                        ids=b.ids_sorted, bins=b.tile_bins).items()}
    out = torch.empty((H, W, 3), device=dev)
    gT = torch.empty((H, W), device=dev)
    gidx = torch.empty((H, W), device=dev, dtype=torch.int32)
    L = _lib.lib()
    sp = _lib.stream_ptr()
    v_out = np.random.default_rng(3).standard_normal((H, W, 3)).astype(np.float32)
    g = [torch.zeros((d["N"], k), device=dev) for k in (2, 3, 3, 1)]
    tv = _t(v_out, dev)
    if packed:
        rec = torch.empty((d["N"], 12), device=dev)
        order, work = torch.empty(T, device=dev, dtype=torch.int32), torch.empty(T, device=dev, dtype=torch.int32)
        assert L.gs_pack_splats(d["N"], t["xys"].data_ptr(), t["conics"].data_ptr(), t["colors"].data_ptr(),
                                t["opac"].data_ptr(), None, rec.data_ptr(), sp) == 0
        assert L.gs_tile_order(T, t["bins"].data_ptr(), 1, order.data_ptr(), sp) == 0
        assert L.gs_raster_fwd_packed(1, H, W, tb[0], tb[1], t["ids"].data_ptr(), t["bins"].data_ptr(), rec.data_ptr(),
                                      t["bg"].data_ptr(), out.data_ptr(), None, gT.data_ptr(), gidx.data_ptr(),
                                      order.data_ptr(), work.data_ptr(), None, None, sp) == 0
        assert L.gs_tile_order(T, work.data_ptr(), 0, order.data_ptr(), sp) == 0
        assert L.gs_raster_bwd_packed(H, W, tb[0], tb[1], t["ids"].data_ptr(), t["bins"].data_ptr(), rec.data_ptr(),
                                      t["bg"].data_ptr(), gT.data_ptr(), gidx.data_ptr(), tv.data_ptr(), None,
                                      g[0].data_ptr(), g[1].data_ptr(), g[2].data_ptr(), g[3].data_ptr(), order.data_ptr(),
                                      None, sp) == 0
    else:
        assert L.gs_raster_fwd(3, H, W, tb[0], tb[1], t["ids"].data_ptr(), t["bins"].data_ptr(), t["xys"].data_ptr(),
                               t["conics"].data_ptr(), t["colors"].data_ptr(), t["opac"].data_ptr(), t["bg"].data_ptr(),
                               out.data_ptr(), gT.data_ptr(), gidx.data_ptr(), sp) == 0
        assert L.gs_raster_bwd(3, H, W, tb[0], tb[1], t["ids"].data_ptr(), t["bins"].data_ptr(), t["xys"].data_ptr(),
                               t["conics"].data_ptr(), t["colors"].data_ptr(), t["opac"].data_ptr(), t["bg"].data_ptr(),
                               gT.data_ptr(), gidx.data_ptr(), tv.data_ptr(), None, g[0].data_ptr(), g[1].data_ptr(),
                               g[2].data_ptr(), g[3].data_ptr(), sp) == 0
    ctx = _rctx(d, b, fidx)
    _, F, excl = forward_parity(oracle, ctx, out.cpu().numpy(), img, (1 - gT).cpu().numpy(), 1 - fT, H, W,
                                got_fidx=gidx.cpu().numpy())
    o_xy, o_cn, o_co, o_op = oracle.raster_bwd(b, d["xys"], d["conics"], colors, d["opac"], bg, H, W, tb, fT, fidx, v_out)
    gradient_parity(oracle, ctx, excl, {"xys": g[0].cpu().numpy(), "conics": g[1].cpu().numpy(),
                                        "colors": g[2].cpu().numpy(), "opac": g[3].cpu().numpy()},
                    {"xys": o_xy, "conics": o_cn, "colors": o_co, "opac": o_op}, H, W)


def test_zero_intersections_is_background(cuda_dev):
    from gaussctrl_exp_b200 import rasterize_gaussians
    dev = cuda_dev
    N, H, W = 10, 40, 50
    z = torch.zeros
    out, alpha = rasterize_gaussians(z((N, 2), device=dev), z(N, device=dev), z(N, device=dev, dtype=torch.int32),
                                     z((N, 3), device=dev), z(N, device=dev, dtype=torch.int32), z((N, 3), device=dev),
                                     z((N, 1), device=dev), H, W, background=torch.tensor([0.2, 0.4, 0.6], device=dev),
                                     return_alpha=True)
    assert torch.allclose(out, torch.tensor([0.2, 0.4, 0.6], device=dev).expand(H, W, 3)) and (alpha == 0).all()


def test_argument_errors(cuda_dev):
    from gaussctrl_exp_b200 import project_gaussians, rasterize_gaussians, spherical_harmonics
    dev = cuda_dev
    with pytest.raises(ValueError):
        rasterize_gaussians(torch.zeros((4, 3), device=dev), None, None, None, None, torch.zeros((4, 3), device=dev),
                            None, 16, 16)
    with pytest.raises(AssertionError):
        rasterize_gaussians(torch.zeros((4, 2), device=dev), None, None, None, None, torch.zeros((4, 3), device=dev),
                            None, 16, 16, background=torch.zeros(4, device=dev))
    with pytest.raises(AssertionError):
        spherical_harmonics(3, torch.zeros((4, 3), device=dev), torch.zeros((4, 9, 3), device=dev))
    with pytest.raises(ValueError):
        project_gaussians(torch.zeros((4, 2), device=dev), torch.zeros((4, 2), device=dev), 1, torch.zeros((4, 4), device=dev),
                          torch.eye(4, device=dev), torch.eye(4, device=dev), 1, 1, 0, 0, 16, 16, (1, 1, 1))
    with pytest.raises(ValueError):  # CPU tensors: no CPU fallback
        project_gaussians(torch.zeros((4, 3)), torch.zeros((4, 3)), 1, torch.zeros((4, 4)), torch.eye(4), torch.eye(4),
                          1, 1, 0, 0, 16, 16, (1, 1, 1))


@pytest.mark.parametrize("N,W,H,deg,use_sh", [(30000, 160, 128, 3, True), (5003, 96, 64, 1, True), (999, 64, 48, 0, False),
                                              (130, 64, 48, 2, True)])
def test_fused_preprocess_matches_operator_chain(oracle, cuda_dev, N, W, H, deg, use_sh):
    """gs_preprocess_fwd/bwd == the reference's op chain (gc_model.py:172-203, :215) evaluated with
    torch ops + the unfused operators, on the same raw leaves."""
    import ctypes
    from gaussctrl_exp_b200 import _lib, project_gaussians, render, spherical_harmonics
    dev = cuda_dev
    p = to_dev(np_params(N, seed=31, s0=0.03), dev)
    cam = small_cam(W, H, 31)
    gc = render.gs_camera(cam)
    leaves = {k: v.clone().requires_grad_() for k, v in p.items()}
    V = torch.tensor(list(gc.viewmat), device=dev).view(3, 4)
    PV = torch.tensor(list(gc.projmat), device=dev).view(4, 4)
    tb = (gc.tiles_x, gc.tiles_y, 1)
    xys, depths, radii, conics, hit, _ = project_gaussians(
        leaves["means"], torch.exp(leaves["scales"]), 1, leaves["quats"] / leaves["quats"].norm(dim=-1, keepdim=True),
        V, PV, cam.fx, cam.fy, cam.cx, cam.cy, H, W, tb)
    coeffs = torch.cat((leaves["features_dc"][:, None, :], leaves["features_rest"]), dim=1)
    if use_sh:
        d = leaves["means"].detach() - cam.c2w[:3, 3].to(dev)
        d = d / d.norm(dim=-1, keepdim=True)
        rgbs = torch.clamp(spherical_harmonics(deg, d, coeffs) + 0.5, min=0.0)
    else:
        rgbs = torch.sigmoid(coeffs[:, 0, :])
    opac = torch.sigmoid(leaves["opacities"])

    L = _lib.lib()
    o = dict(xys=torch.empty(N, 2, device=dev), depths=torch.empty(N, device=dev),
             radii=torch.empty(N, device=dev, dtype=torch.int32), conics=torch.empty(N, 3, device=dev),
             hit=torch.empty(N, device=dev, dtype=torch.int32), rgbs=torch.empty(N, 3, device=dev),
             opac=torch.empty(N, device=dev))
    rc = L.gs_preprocess_fwd(N, p["means"].data_ptr(), p["scales"].data_ptr(), p["quats"].data_ptr(),
                             p["opacities"].data_ptr(), p["features_dc"].data_ptr(), p["features_rest"].data_ptr(),
                             int(use_sh), deg, ctypes.addressof(gc), o["xys"].data_ptr(), o["depths"].data_ptr(),
                             o["radii"].data_ptr(), o["conics"].data_ptr(), o["hit"].data_ptr(), o["rgbs"].data_ptr(),
                             o["opac"].data_ptr(), _lib.stream_ptr())
    assert rc == 0
    vis = radii > 0
    # exp/sigmoid are evaluated by different libraries (torch vs CUDA libm): 1-ulp input differences may move a
    # radius across an integer for a vanishing fraction of Gaussians
    same = (o["radii"] == radii) & (o["hit"] == hit)
    assert (~same).float().mean().item() <= 1e-4
    m = (vis & same).cpu().numpy()
    for k, ref in (("xys", xys), ("depths", depths), ("conics", conics), ("opac", opac[:, 0])):
        assert rel_linf(o[k].cpu().numpy()[m], ref.detach().cpu().numpy()[m]) < 1e-5, k
    assert rel_linf(o["rgbs"].cpu().numpy()[m], rgbs.detach().cpu().numpy()[m]) < 2e-5
    assert (o["rgbs"][~(o["radii"] > 0)] == 0).all()
    # the two-part form of the same kernel (geometry, then colour from the stored radii) is bit-identical
    o2 = {k: torch.full_like(v, 7) for k, v in o.items()}
    for part in (1, 2):
        rc = L.gs_preprocess_fwd_part(part, N, p["means"].data_ptr(), p["scales"].data_ptr(), p["quats"].data_ptr(),
                                      p["opacities"].data_ptr(), p["features_dc"].data_ptr(),
                                      p["features_rest"].data_ptr(), int(use_sh), deg, ctypes.addressof(gc),
                                      o2["xys"].data_ptr(), o2["depths"].data_ptr(), o2["radii"].data_ptr(),
                                      o2["conics"].data_ptr(), o2["hit"].data_ptr(), o2["rgbs"].data_ptr(),
                                      o2["opac"].data_ptr(), _lib.stream_ptr())
        assert rc == 0
        if part == 1:
            assert (o2["rgbs"] == 7).all()
    for k in o:
        assert torch.equal(o[k], o2[k]), k

    g = torch.Generator().manual_seed(3)
    v_xys = torch.randn(N, 2, generator=g).to(dev) * vis[:, None]
    v_con = torch.randn(N, 3, generator=g).to(dev) * vis[:, None]
    v_rgb = torch.randn(N, 3, generator=g).to(dev) * vis[:, None]
    v_op = torch.randn(N, generator=g).to(dev) * vis
    ((xys * v_xys).sum() + (conics * v_con).sum() + (rgbs * v_rgb).sum() + (opac[:, 0] * v_op).sum()).backward()
    go = {k: torch.empty_like(v) for k, v in p.items()}
    rc = L.gs_preprocess_bwd(N, p["means"].data_ptr(), p["scales"].data_ptr(), p["quats"].data_ptr(),
                             p["opacities"].data_ptr(), int(use_sh), deg, ctypes.addressof(gc), o["radii"].data_ptr(),
                             o["rgbs"].data_ptr(), v_xys.data_ptr(), v_con.data_ptr(), v_rgb.data_ptr(), v_op.data_ptr(),
                             go["means"].data_ptr(), go["scales"].data_ptr(), go["quats"].data_ptr(),
                             go["opacities"].data_ptr(), go["features_dc"].data_ptr(), go["features_rest"].data_ptr(),
                             _lib.stream_ptr())
    assert rc == 0
    mm = same.cpu().numpy()
    for k in ("means", "scales", "quats", "opacities", "features_dc", "features_rest"):
        ref = leaves[k].grad.cpu().numpy()[mm]
        got = go[k].cpu().numpy()[mm]
        if np.abs(ref).max() == 0:
            assert np.abs(got).max() == 0, k
        else:
            assert rel_linf(got, ref) < 1e-4, k


@pytest.mark.parametrize("T", [1, 37, 1024, 2048, 2049, 8160, 70000])
def test_tile_order_is_a_heaviest_first_permutation(cuda_dev, T):
    """gs_tile_order: a permutation of the tiles, exact descending order (ties by tile id) up to 2 048 tiles, descending
    by 1 024 work levels above; both input forms (work array, tile ranges)."""
    from gaussctrl_exp_b200 import _lib
    L = _lib.lib()
    rng = np.random.default_rng(T)
    work = (rng.gamma(1.5, 900.0, T)).astype(np.int32)
    work[rng.random(T) < 0.2] = 0
    if T > 3:
        work[:3] = work.max()                                   # ties at the top
    starts = np.concatenate([[0], np.cumsum(work)[:-1]]).astype(np.int32)
    bins = np.stack([starts, starts + work], axis=1).astype(np.int32)
    for is_range, arr in ((0, work), (1, bins)):
        order = torch.full((T,), -1, device=cuda_dev, dtype=torch.int32)
        a = torch.from_numpy(np.ascontiguousarray(arr)).to(cuda_dev)
        assert L.gs_tile_order(T, a.data_ptr(), is_range, order.data_ptr(), _lib.stream_ptr()) == 0
        o = order.cpu().numpy()
        assert np.array_equal(np.sort(o), np.arange(T))         # every tile exactly once
        w = work[o].astype(np.int64)
        if T <= 2048:
            want = np.lexsort((np.arange(T), -work.astype(np.int64)))
            assert np.array_equal(o, want)
        else:
            wmax = max(int(work.max()), 1)
            lev = 1023 - (w * 1023) // wmax
            assert np.all(np.diff(lev) >= 0)                    # heaviest level first
