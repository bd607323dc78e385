"""CPU tests of the oracle itself (no GPU): the C restatement is checked against an independent
float64 autograd restatement (oracle/torch_oracle.py) and against the invariants of the domain.
The reference ships no golden vectors for this path — parity vs the real gsplat
binary is UNPINNED; what is pinned here is internal consistency: forward == float64 restatement,
backward == exact derivative of the forward."""
import numpy as np
import torch

from util import np_params, projected, rel_linf, small_cam


def test_project_fwd_matches_float64(oracle):
    from oracle import torch_oracle as to
    d = projected(oracle, N=3000, W=96, H=64, seed=1, s0=0.03)
    p, cam = d["p"], d["cam"]
    xys, tz, conics = to.project(torch.from_numpy(p["means"]).double(), torch.from_numpy(d["scales"]).double(), 1.0,
                                 torch.from_numpy(d["quats"]).double(), d["V"], d["PV"], cam.fx, cam.fy, cam.cx,
                                 cam.cy, d["H"], d["W"])
    vis = d["radii"] > 0
    assert vis.sum() > 500
    assert rel_linf(d["xys"][vis], xys.numpy()[vis]) < 1e-5
    assert rel_linf(d["depths"][vis], tz.numpy()[vis]) < 1e-6
    assert np.abs(d["conics"][vis] - conics.numpy()[vis]).max() / np.abs(conics.numpy()[vis]).max() < 1e-4
    # culled rows are all-zero
    assert (d["xys"][~vis] == 0).all() and (d["hit"][~vis] == 0).all() and (d["conics"][~vis] == 0).all()
    # tiles_hit equals the bbox area recomputed from (xys, radii)
    tc = d["xys"] / np.float32(16)
    tr = d["radii"].astype(np.float32) / np.float32(16)
    tx, ty = d["tb"][0], d["tb"][1]
    x0 = np.clip((tc[:, 0] - tr).astype(np.int32), 0, tx)
    x1 = np.clip((tc[:, 0] + tr + np.float32(1)).astype(np.int32), 0, tx)
    y0 = np.clip((tc[:, 1] - tr).astype(np.int32), 0, ty)
    y1 = np.clip((tc[:, 1] + tr + np.float32(1)).astype(np.int32), 0, ty)
    assert ((x1 - x0) * (y1 - y0))[vis].tolist() == d["hit"][vis].tolist()


def test_project_bwd_is_exact_derivative(oracle):
    from oracle import torch_oracle as to
    d = projected(oracle, N=1500, W=96, H=64, seed=2, s0=0.03)
    p, cam = d["p"], d["cam"]
    rng = np.random.default_rng(0)
    N = d["N"]
    v_xys = rng.standard_normal((N, 2)).astype(np.float32)
    v_dep = rng.standard_normal((N,)).astype(np.float32)
    v_con = rng.standard_normal((N, 3)).astype(np.float32)
    vm, vs, vq = oracle.project_bwd(p["means"], d["scales"], 1.0, d["quats"], d["V"][:3], d["PV"], cam.fx, cam.fy,
                                    cam.cx, cam.cy, d["H"], d["W"], d["radii"], v_xys, v_dep, v_con)
    m = torch.from_numpy(p["means"]).double().requires_grad_()
    s = torch.from_numpy(d["scales"]).double().requires_grad_()
    q = torch.from_numpy(d["quats"]).double().requires_grad_()
    xys, tz, conics = to.project(m, s, 1.0, q, d["V"], d["PV"], cam.fx, cam.fy, cam.cx, cam.cy, d["H"], d["W"])
    vis = torch.from_numpy(d["radii"] > 0)
    loss = ((xys * torch.from_numpy(v_xys).double())[vis].sum() + (tz * torch.from_numpy(v_dep).double())[vis].sum()
            + (conics * torch.from_numpy(v_con).double())[vis].sum())
    loss.backward()
    assert rel_linf(vm, m.grad.numpy()) < 1e-5
    assert rel_linf(vs, s.grad.numpy()) < 1e-5
    assert rel_linf(vq, q.grad.numpy()) < 1e-5
    assert (vm[~vis.numpy()] == 0).all()


def test_sh_fwd_bwd(oracle):
    from oracle import torch_oracle as to
    rng = np.random.default_rng(3)
    N = 500
    dirs = rng.standard_normal((N, 3)).astype(np.float32)
    coeffs = rng.standard_normal((N, 16, 3)).astype(np.float32)
    vcol = rng.standard_normal((N, 3)).astype(np.float32)
    for deg in range(4):
        out = oracle.sh_fwd(deg, dirs, coeffs)
        c = torch.from_numpy(coeffs).double().requires_grad_()
        ref = to.sh_colors(deg, torch.from_numpy(dirs).double(), c)
        assert rel_linf(out, ref.detach().numpy()) < 1e-6
        (ref * torch.from_numpy(vcol).double()).sum().backward()
        g = oracle.sh_bwd(deg, 16, dirs, vcol)
        assert rel_linf(g, c.grad.numpy()) < 1e-6
        assert (g[:, (deg + 1) ** 2:, :] == 0).all()
    # degree 4 (25 bases): linear in the coefficients, zero-degree term isolates C0
    c25 = np.zeros((N, 25, 3), np.float32)
    c25[:, 0, :] = 1.0
    assert np.allclose(oracle.sh_fwd(4, dirs, c25), 0.28209479177387814, atol=1e-7)


def test_binning_invariants(oracle):
    d = projected(oracle, N=4000, W=128, H=80, seed=4, s0=0.03)
    b = oracle.bin_and_sort(d["xys"], d["depths"], d["radii"], d["hit"], d["tb"])
    assert b.I == int(d["hit"].sum()) and b.I > 4000
    assert (np.diff(b.cum) >= 0).all() and b.cum[-1] == b.I
    # sorted, and stable: equal keys keep ascending Gaussian id (== emission order)
    assert (np.diff(b.keys_sorted) >= 0).all()
    same = np.diff(b.keys_sorted) == 0
    assert (np.diff(b.ids_sorted)[same] > 0).all()
    # same multiset
    order = np.lexsort((b.ids, b.keys))
    assert (b.keys[order] == b.keys_sorted).all() and (b.ids[order] == b.ids_sorted).all()
    # key layout: tile id in the high word, depth bits in the low word
    tiles = (b.keys_sorted >> 32).astype(np.int64)
    T = d["tb"][0] * d["tb"][1]
    assert tiles.min() >= 0 and tiles.max() < T
    assert ((b.keys_sorted & 0xFFFFFFFF).astype(np.uint32) == d["depths"][b.ids_sorted].view(np.uint32)).all()
    # ranges partition [0, I) in tile order; empty tiles are (0, 0)
    cnt = np.bincount(tiles, minlength=T)
    for t in range(T):
        s, e = b.tile_bins[t]
        if cnt[t] == 0:
            assert s == 0 and e == 0
        else:
            assert e - s == cnt[t] and (tiles[s:e] == t).all()


def test_raster_fwd_matches_float64_and_bwd_is_exact_derivative(oracle):
    from oracle import torch_oracle as to
    d = projected(oracle, N=1200, W=48, H=40, seed=5, s0=0.04)
    H, W, tb = d["H"], d["W"], d["tb"]
    rng = np.random.default_rng(7)
    colors = rng.random((d["N"], 3)).astype(np.float32)
    bg = np.array([0.1, 0.2, 0.3], np.float32)
    b = oracle.bin_and_sort(d["xys"], d["depths"], d["radii"], d["hit"], tb)
    img, fT, fidx, pairs = oracle.raster_fwd(b, d["xys"], d["conics"], colors, d["opac"], bg, H, W, tb)
    assert pairs > H * W
    xy = torch.from_numpy(d["xys"]).double().requires_grad_()
    cn = torch.from_numpy(d["conics"]).double().requires_grad_()
    co = torch.from_numpy(colors).double().requires_grad_()
    op = torch.from_numpy(d["opac"][:, 0]).double().requires_grad_()
    ref, ref_a = to.composite(xy, cn, co, op, bg, H, W, b, fidx, tb)
    assert np.abs(img - ref.detach().numpy()).max() < 2e-5
    assert np.abs((1 - fT) - ref_a.detach().numpy()).max() < 2e-5
    v_out = rng.standard_normal((H, W, 3)).astype(np.float32)
    v_a = rng.standard_normal((H, W)).astype(np.float32)
    ((ref * torch.from_numpy(v_out).double()).sum() + (ref_a * torch.from_numpy(v_a).double()).sum()).backward()
    g_xy, g_cn, g_co, g_op = oracle.raster_bwd(b, d["xys"], d["conics"], colors, d["opac"], bg, H, W, tb, fT, fidx,
                                               v_out, v_a)
    assert rel_linf(g_co, co.grad.numpy()) < 1e-4
    assert rel_linf(g_op[:, 0], op.grad.numpy()) < 1e-4
    assert rel_linf(g_xy, xy.grad.numpy()) < 1e-4
    assert rel_linf(g_cn, cn.grad.numpy()) < 1e-4


def test_raster_edge_cases(oracle):
    # no Gaussians in view -> background, alpha 0, final_idx == 0
    H, W, tb = 40, 50, (4, 3, 1)
    b = oracle.bin_and_sort(np.zeros((3, 2), np.float32), np.zeros(3, np.float32), np.zeros(3, np.int32),
                            np.zeros(3, np.int32), tb)
    assert b.I == 0 and (b.tile_bins == 0).all()
    bg = np.array([0.3, 0.6, 0.9], np.float32)
    img, fT, fidx, pairs = oracle.raster_fwd(b, np.zeros((3, 2), np.float32), np.zeros((3, 3), np.float32),
                                             np.zeros((3, 3), np.float32), np.zeros((3, 1), np.float32), bg, H, W, tb)
    assert pairs == 0 and (fT == 1).all() and (fidx == 0).all() and np.allclose(img, bg)
    # one opaque Gaussian at the image centre: its pixel saturates at alpha 0.999
    xy = np.array([[25.0, 20.0]], np.float32)
    con = np.array([[0.05, 0.0, 0.05]], np.float32)
    hit = np.array([tb[0] * tb[1]], np.int32)
    b = oracle.bin_and_sort(xy, np.array([1.0], np.float32), np.array([100], np.int32), hit, tb)
    assert b.I == 12
    img, fT, fidx, _ = oracle.raster_fwd(b, xy, con, np.ones((1, 3), np.float32), np.array([[5.0]], np.float32),
                                         np.zeros(3, np.float32), H, W, tb)
    assert abs(img[20, 25, 0] - 0.999) < 1e-6 and abs(fT[20, 25] - 0.001) < 1e-6


def test_render_view_runs_and_grads_flow(oracle):
    p = np_params(3000, seed=6, s0=0.03)
    cam = small_cam(80, 64, 6).as_dict()
    out, ctx = oracle.render_view(p, cam, 3, training=False)
    assert out["rgb"].shape == (64, 80, 3) and out["depth"].shape == (64, 80, 1)
    assert out["rgb"].max() <= 1.0 and (out["depth"][out["accumulation"] == 0] == 1000).all()
    v = np.random.default_rng(1).standard_normal((64, 80, 3)).astype(np.float32)
    g = oracle.render_view_backward(p, ctx, v)
    for k in ("means", "scales", "quats", "features_dc", "features_rest", "opacities"):
        assert g[k].shape == p[k].shape and np.isfinite(g[k]).all() and np.abs(g[k]).max() > 0
