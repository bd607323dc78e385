"""The reference's own caller as the yardstick. tests/golden/reference_caller.npz holds the outputs of
GaussCtrlModel.get_outputs / get_outputs_for_camera (/root/reference/gaussctrl/gc_model.py:77-255) extracted by AST and
executed UNMODIFIED (tests/golden/gen_golden.py::reference_caller_golden; nerfstudio's Cameras / SplatfactoModel
attributes stubbed, gsplat's three operators = the CPU oracle's). Checked against it:
  * CPU: the oracle's numpy mirror of the caller (oracle.render_view) and the host matrices of model.py;
  * CPU: the three import lines of gc_model.py:32-36 resolve through the gsplat shim;
  * GPU: GaussianModel, literal (through the shim's operators) and fused, for the eval view, the obb_box crop, the
    empty crop's early return and a training view at half resolution (coarse-to-fine downscale)."""
import os

import numpy as np
import pytest
import torch

from util import forward_parity, depth_parity

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_caller.npz")
GREY = (0.1490, 0.1647, 0.2157)          # splatfacto's resolved colour for background_color="random" in eval


def _inputs():
    from gaussctrl_exp_b200 import scene
    g = np.load(GOLD)
    N, seed, s0 = int(g["scene"][0]), int(g["scene"][1]), float(g["scene"][2])
    p = {k: v.numpy() for k, v in scene.make_scene(N, seed, s0=s0).items()}
    fx, fy, cx, cy, W, H = g["intr"]
    cam = scene.Camera(torch.from_numpy(g["c2w"]), float(fx), float(fy), float(cx), float(cy), int(W), int(H))
    box = scene.OrientedBox(torch.from_numpy(g["box_R"]), torch.from_numpy(g["box_T"]), torch.from_numpy(g["box_S"]))
    far = scene.OrientedBox(torch.eye(3), torch.from_numpy(g["far_T"]), torch.ones(3))
    return g, p, cam, box, far


def test_oracle_caller_mirror_matches_reference_caller(oracle):
    g, p, cam, box, far = _inputs()
    H, W = cam.H, cam.W
    ref, ctx = oracle.render_view(p, cam.as_dict(), 3, bg=GREY, training=False)
    # same operators underneath; the pre-activations are torch's in the reference and numpy's in the mirror (1 ulp)
    rep, F, _ = forward_parity(oracle, ctx, g["eval_rgb"], ref["rgb"], g["eval_accumulation"], ref["accumulation"], H, W)
    assert rep["rgb_err_max_not_at_risk"] < 1e-5
    depth_parity(g["eval_depth"], ref["depth"], ref["accumulation"], F)
    assert (g["eval_depth"][ref["accumulation"] == 0] == 1000).all()
    inside = box.within(torch.from_numpy(p["means"])).numpy()
    assert int(inside.sum()) == int(g["crop_count"])
    refc, ctxc = oracle.render_view({k: v[inside] for k, v in p.items()}, cam.as_dict(), 3, bg=GREY, training=False)
    rep, _, _ = forward_parity(oracle, ctxc, g["crop_rgb"], refc["rgb"], g["crop_accumulation"], refc["accumulation"], H, W)
    assert rep["rgb_err_max_not_at_risk"] < 1e-5
    assert list(g["empty_keys"]) == ["rgb"] and np.allclose(g["empty_rgb"], np.array(GREY, np.float32))
    from gaussctrl_exp_b200.model import camera_matrices
    _, _, V0, P0, _ = camera_matrices(cam.c2w, cam.fx, cam.fy, W, H)
    assert np.allclose(V0.numpy(), g["eval_mat_view"], atol=1e-6) and np.allclose(P0.numpy(), g["eval_mat_proj"], rtol=1e-6)
    assert np.array_equal(g["eval_mat_c2w"], cam.c2w.numpy())


def test_gsplat_shim_resolves_the_reference_imports():
    """gc_model.py:32-36 import gsplat.sh, gsplat.project_gaussians and gsplat.rasterize; nerfstudio 1.0.0's splatfacto
    additionally imports gsplat._torch_impl.quat_to_rotmat and gsplat.compute_cumulative_intersects."""
    import gaussctrl_exp_b200
    gaussctrl_exp_b200.install_gsplat_shim()
    env = {}
    exec("from gsplat.sh import num_sh_bases, spherical_harmonics\n"
         "from gsplat.project_gaussians import project_gaussians\n"
         "from gsplat.rasterize import rasterize_gaussians\n"
         "from gsplat._torch_impl import quat_to_rotmat\n"
         "from gsplat import compute_cumulative_intersects\n", env)
    assert env["project_gaussians"] is gaussctrl_exp_b200.project_gaussians
    assert env["rasterize_gaussians"] is gaussctrl_exp_b200.rasterize_gaussians
    assert env["spherical_harmonics"] is gaussctrl_exp_b200.spherical_harmonics
    assert [env["num_sh_bases"](d) for d in range(5)] == [1, 4, 9, 16, 25]
    R = env["quat_to_rotmat"](torch.tensor([[2.0, 0.0, 0.0, 0.0]]))
    assert torch.allclose(R[0], torch.eye(3))


@pytest.mark.gpu
@pytest.mark.parametrize("fused", [False, True])
def test_host_mirror_matches_reference_caller_on_gpu(oracle, cuda_dev, fused):
    from gaussctrl_exp_b200.model import GaussianModel
    g, p, cam, box, far = _inputs()
    H, W = cam.H, cam.W
    dev = cuda_dev
    pt = {k: torch.from_numpy(v).to(dev) for k, v in p.items()}
    m = GaussianModel(pt, sh_degree=3, background_color="random", fused=fused).to(dev)
    _, ctx = oracle.render_view(p, cam.as_dict(), 3, bg=GREY, training=False)
    out = m.get_outputs_for_camera(cam)
    assert sorted(k for k, v in out.items() if v is not None) == list(g["eval_keys"]) and m.training
    _, F, _ = forward_parity(oracle, ctx, out["rgb"].cpu().numpy(), g["eval_rgb"], out["accumulation"].cpu().numpy(),
                             g["eval_accumulation"], H, W)
    depth_parity(out["depth"].cpu().numpy(), g["eval_depth"], g["eval_accumulation"], F)
    assert torch.allclose(out["mat_view"], torch.from_numpy(g["eval_mat_view"]), atol=1e-6)
    assert torch.allclose(out["mat_proj"], torch.from_numpy(g["eval_mat_proj"]), rtol=1e-6)
    assert np.array_equal(out["mat_c2w"].cpu().numpy(), g["eval_mat_c2w"])
    # obb_box crop (gc_model.py:108-113, :157-170, :244-251)
    inside = box.within(torch.from_numpy(p["means"])).numpy()
    _, ctxc = oracle.render_view({k: v[inside] for k, v in p.items()}, cam.as_dict(), 3, bg=GREY, training=False)
    out = m.get_outputs_for_camera(cam, obb_box=box)
    assert m.xys.shape[0] == int(g["crop_count"])
    _, F, _ = forward_parity(oracle, ctxc, out["rgb"].cpu().numpy(), g["crop_rgb"], out["accumulation"].cpu().numpy(),
                             g["crop_accumulation"], H, W)
    depth_parity(out["depth"].cpu().numpy(), g["crop_depth"], g["crop_accumulation"], F)
    out = m.get_outputs_for_camera(cam, obb_box=far)         # empty crop: early return (gc_model.py:110-111)
    assert list(out) == ["rgb"] and np.allclose(out["rgb"].cpu().numpy(), g["empty_rgb"])
    # training view under the coarse-to-fine schedule: half resolution, white background, camera restored
    mw = GaussianModel(pt, sh_degree=3, background_color="white", fused=fused, num_downscales=1, step=3100,
                       resolution_schedule=10000).to(dev)
    mw.train()
    out = mw(cam)
    assert out["rgb"].shape == (H // 2, W // 2, 3) and out["depth"] is None and (cam.W, cam.H) == (W, H)
    from gaussctrl_exp_b200 import scene
    half = scene.Camera(cam.c2w, cam.fx / 2, cam.fy / 2, cam.cx / 2, cam.cy / 2, W // 2, H // 2)
    _, ctxh = oracle.render_view(p, half.as_dict(), 3, bg=(1.0, 1.0, 1.0), training=True)
    forward_parity(oracle, ctxh, out["rgb"].detach().cpu().numpy(), g["train_half_rgb"],
                   out["accumulation"].detach().cpu().numpy(), g["train_half_accumulation"], H // 2, W // 2)
    assert torch.allclose(out["mat_proj"], torch.from_numpy(g["train_half_mat_proj"]), rtol=1e-6)
