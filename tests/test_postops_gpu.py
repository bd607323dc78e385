"""SURVEY.md §8f rank 3 on the GPU: the depth post-ops kernels (csrc/postops.cu through the C ABI) against the
reference's OWN functions (gc_pipeline.py:286-314, :439-457, :503-536, executed unmodified by tests/golden/gen_golden.py
-> reference_postops.npz) and against the host mirror at the reference's working resolution; the checkpoint the
reference's own `save_checkpoint` wrote (reference_step-000029999.ckpt) loaded onto the GPU, rendered against the
oracle and resumed with the sliced Adam kernel."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def test_ray_directions_bit_identical_to_reference(cuda_dev):
    from gaussctrl_exp_b200 import postops
    g = np.load(os.path.join(GOLD, "reference_postops.npz"))
    h, w, _ = g["depth"].shape
    rays = postops.compute_ray_directions((h, w), device=cuda_dev).cpu().numpy()
    assert rays.shape == g["rays"].shape and np.array_equal(rays, g["rays"])
    rr = postops.compute_ray_directions((20, 12), device=cuda_dev).cpu().numpy()      # rectangular: (height=12, width=20)
    assert rr.shape == g["rays_rect"].shape == (12, 20, 3) and np.array_equal(rr, g["rays_rect"])
    # the reference's working size: against the reference's arithmetic restated with numpy's correctly rounded float32
    # operations (what gen_golden.py executed for the small shapes above), bit for bit; the torch host mirror to one ulp
    # (torch's CPU kernels differ between hosts: seen on one box of the pool)
    big = postops.compute_ray_directions((512, 512), device=cuda_dev)
    vh = 2 * np.tan(0.8880228256678113 / 2) * 0.001
    ax = (np.arange(512) + 0.5) / 512 * vh - vh / 2
    ref = np.zeros((512, 512, 3), np.float32)
    ref[..., 0], ref[..., 1], ref[..., 2] = ax[None, :], ax[:, None], -0.001
    ref /= np.linalg.norm(ref, axis=-1, keepdims=True)
    assert np.array_equal(big.cpu().numpy(), ref)
    host = postops.compute_ray_directions((512, 512))
    assert (big.cpu() - host).abs().max().item() <= 2.0 ** -23
    assert torch.allclose(big.norm(dim=-1), torch.ones(512, 512, device=cuda_dev), atol=1e-6)


def test_depth_to_world_matches_reference(cuda_dev):
    from gaussctrl_exp_b200 import _lib, postops
    g = np.load(os.path.join(GOLD, "reference_postops.npz"))
    depth = torch.from_numpy(g["depth"]).to(cuda_dev)
    pts = postops.depth_map_screen_to_world(depth, torch.from_numpy(g["mat_view"]))
    assert pts.is_cuda and pts.shape == g["pts_world"].shape
    assert np.abs(pts.cpu().numpy() - g["pts_world"]).max() < 1e-5
    # the reference's working size, a camera of the bench scene, depths as the eval path produces them (incl. the
    # far-plane fill of empty pixels)
    from gaussctrl_exp_b200 import scene
    from gaussctrl_exp_b200.model import camera_matrices
    cam = scene.make_cameras(3, 4)[2]
    gen = torch.Generator().manual_seed(2)
    d = torch.rand(512, 512, 1, generator=gen) * 4 + 0.1
    d[torch.rand(512, 512, 1, generator=gen) < 0.2] = 1000.0
    view0 = torch.linalg.inv(torch.cat([cam.c2w, torch.tensor([[0.0, 0.0, 0.0, 1.0]])], 0))
    want = postops.depth_map_screen_to_world(d, view0)
    got = postops.depth_map_screen_to_world(d.to(cuda_dev), view0)
    err = (got.cpu() - want).abs().max().item()
    assert err <= 1e-5 * want.abs().max().item(), err
    # points lie on their rays: distance from the camera centre equals the depth (unit rays)
    centre = cam.c2w[:3, 3].to(cuda_dev)
    dist = (got - centre).norm(dim=-1, keepdim=True)
    assert torch.allclose(dist, d.to(cuda_dev), rtol=2e-5, atol=1e-5)
    with pytest.raises(_lib.GsError):                           # the reference's broadcast fails for h != w
        postops.depth_map_screen_to_world(torch.ones(48, 64, 1, device=cuda_dev), view0)
    with pytest.raises(_lib.GsError):                           # singular view matrix
        postops.depth_map_screen_to_world(torch.ones(8, 8, 1, device=cuda_dev), torch.zeros(4, 4))


@pytest.mark.parametrize("shape", [(1, 48, 48), (1, 512, 512), (1, 1080, 1920), (2, 33, 7)])
def test_disparity_bit_identical_to_torch(cuda_dev, shape):
    from gaussctrl_exp_b200 import postops
    gen = torch.Generator().manual_seed(shape[1])
    d = torch.rand(shape, generator=gen) * 5 + 0.05
    d[torch.rand(shape, generator=gen) < 0.1] = 1000.0
    d.view(-1)[shape[2] + 3] = 0.0                               # an empty-depth pixel: disparity 1e5 becomes the maximum
    want = postops.depth2disparity(d)                            # host mirror == gc_pipeline.py:449-457 line by line
    got = postops.depth2disparity(d.to(cuda_dev))
    assert got.is_cuda and got.shape == want.shape == (1, 3 * shape[0]) + shape[1:]
    assert torch.equal(got.cpu(), want)
    assert got.max().item() == 1.0
    if shape == (1, 48, 48):
        g = np.load(os.path.join(GOLD, "reference_postops.npz"))
        dd = torch.from_numpy(g["depth"][:, :, 0][None]).to(cuda_dev)
        assert np.array_equal(postops.depth2disparity(dd).cpu().numpy(), g["disparity"])


def test_post_ops_follow_the_eval_render_without_host_sync(oracle, cuda_dev):
    """render -> disparity / world points on the same stream, fed by the model's own outputs (what
    gc_pipeline.py:222-236 does per camera), against the same post-ops applied to the oracle's depth."""
    from gaussctrl_exp_b200 import postops
    from gaussctrl_exp_b200.model import GaussianModel
    from util import np_params, small_cam, to_dev
    p = np_params(6000, seed=5, s0=0.04)
    cam = small_cam(96, 96, 5)
    ref, _ = oracle.render_view(p, cam.as_dict(), 3, training=False)
    m = GaussianModel(to_dev(p, cuda_dev), sh_degree=3, background_color=torch.tensor([0.1, 0.2, 0.3])).to(cuda_dev)
    out = m.get_outputs_for_camera(cam)
    disp = postops.depth2disparity(out["depth"].permute(2, 0, 1))
    pts = postops.depth_map_screen_to_world(out["depth"], out["mat_view"])
    rd = torch.from_numpy(ref["depth"])
    want_disp = postops.depth2disparity(rd.permute(2, 0, 1))
    want_pts = postops.depth_map_screen_to_world(rd, out["mat_view"].cpu())
    same = (out["depth"].cpu() - rd).abs() <= 1e-4 * rd.abs()    # pixels whose depth agrees (all but decision flips)
    assert same.float().mean().item() > 0.995
    assert ((disp.cpu() - want_disp).abs()[0, 0][same[..., 0]] <= 2e-4).all()
    assert ((pts.cpu() - want_pts).abs()[same[..., 0]] <= 2e-4 * want_pts.abs().max()).all()


def test_reference_written_checkpoint_renders_and_resumes_on_gpu(oracle, cuda_dev):
    from gaussctrl_exp_b200 import dist as gdist
    from gaussctrl_exp_b200 import postops
    from gaussctrl_exp_b200.model import GaussianModel
    from util import forward_parity, small_cam
    f = os.path.join(GOLD, "reference_step-000029999.ckpt")
    params, step, opt = postops.load_splatfacto_checkpoint(f, map_location=cuda_dev, with_optimizers=True)
    assert step == 29999 and all(v.is_cuda for v in params.values())
    cam = small_cam(64, 48, 3)
    p_np = {k: v.cpu().numpy() for k, v in params.items()}
    ref, ctx = oracle.render_view(p_np, cam.as_dict(), 3, training=False)
    m = GaussianModel(params, sh_degree=3, background_color=torch.tensor([0.1, 0.2, 0.3])).to(cuda_dev)
    out = m.get_outputs_for_camera(cam)
    forward_parity(oracle, ctx, out["rgb"].cpu().numpy(), ref["rgb"], out["accumulation"].cpu().numpy(),
                   ref["accumulation"], 48, 64)
    # next fine-tune step of the optimiser state in the file: gs_adam_slice vs torch.optim.Adam.load_state_dict
    raw = torch.load(f, weights_only=False)
    gen = torch.Generator().manual_seed(3)
    grads = {k: torch.randn(v.shape, generator=gen) * 1e-3 for k, v in params.items()}
    ours = {k: v.clone() for k, v in params.items()}
    adam = gdist.ShardedAdam(ours, rank=0, world=1)
    adam.load_reference_state(opt)
    adam.step({k: g.to(cuda_dev).contiguous() for k, g in grads.items()})
    for group, name in postops.OPTIMIZER_GROUPS.items():
        prm = torch.nn.Parameter(params[name].cpu().clone())
        ref_opt = torch.optim.Adam([prm], lr=1.0)
        ref_opt.load_state_dict(raw["optimizers"][group])
        prm.grad = grads[name]
        ref_opt.step()
        err = (ours[name].cpu() - prm.data).abs().max().item()
        assert err <= 1e-7 + 2e-6 * prm.data.abs().max().item(), (name, err)
