"""Generates tests/golden/*.npz. Run HERE (the container that has /root/reference); the GPU box
only sees the committed .npz files.

1. reference_matrices.npz — outputs of the reference's OWN functions, extracted by AST from
   /root/reference/gaussctrl/gc_model.py:39-57 (projection_matrix_ogl) and
   /root/reference/gaussctrl/ad_render.py:49-86 (projection_matrix_splatfacto = nerfstudio's
   projection_matrix, projection_matrix_ogl) and executed unmodified. These are the only pieces of
   the hot path whose code is in the reference tree; they pin the caller-side
   matrix conventions.
   (3. reference_postops.npz and 4. reference_noise_helpers.npz: see main().)
2. oracle_small_scene.npz — a small seeded scene pushed through the CPU oracle (NOT
   reference-derived: gsplat is absent). It freezes the oracle's behaviour so later edits to the
   oracle or to the scene generator cannot drift silently.
"""
import ast
import math
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
REF = "/root/reference/gaussctrl"


def extract(path, names, env):
    tree = ast.parse(open(path).read())
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in names:
            mod = ast.Module(body=[node], type_ignores=[])
            exec(compile(mod, path, "exec"), env)
    return env


def reference_caller_golden():
    """5. reference_caller.npz — the reference's OWN caller, GaussCtrlModel.get_outputs / get_outputs_for_camera
    (/root/reference/gaussctrl/gc_model.py:77-255), extracted by AST and executed UNMODIFIED on the CPU. What it needs
    from outside the file is stubbed: nerfstudio's `Cameras` / `SplatfactoModel` attributes it reads (a10 of SURVEY.md
    §8), `OrientedBox.within`, `renderers.BACKGROUND_COLOR_OVERRIDE`, nerfstudio's `projection_matrix` (pinned
    separately against ad_render.py:49-67 in reference_matrices.npz), and the three gsplat operators, which here are
    the CPU oracle's. The outputs pin every caller-side decision (background, crop, downscale, pre-activations,
    `+0.5`/clamp, depth normalisation, early returns, returned matrices) for the host mirrors in
    gaussctrl_exp_b200/model.py and oracle/oracle.py."""
    import torch
    from typing import Dict, List, Optional, Union
    from gaussctrl_exp_b200 import scene
    from gaussctrl_exp_b200.model import projection_matrix
    from oracle import oracle as orc

    def project_gaussians(means3d, scales, glob_scale, quats, viewmat, projmat, fx, fy, cx, cy, H, W, tile_bounds,
                          clip_thresh=0.01):
        r = orc.project_fwd(means3d.detach().numpy(), scales.detach().numpy(), float(glob_scale), quats.detach().numpy(),
                            viewmat.numpy(), projmat.numpy(), fx, fy, cx, cy, H, W, tile_bounds, clip_thresh)
        xys, depths, radii, conics, hit, cov3d = (torch.from_numpy(a) for a in r)
        return xys.requires_grad_(), depths, radii, conics, hit, cov3d

    def spherical_harmonics(n, viewdirs, coeffs):
        return torch.from_numpy(orc.sh_fwd(int(n), viewdirs.detach().numpy(), coeffs.detach().numpy()))

    def rasterize_gaussians(xys, depths, radii, conics, num_tiles_hit, colors, opacity, H, W, background=None,
                            return_alpha=False):
        tb = ((W + 15) // 16, (H + 15) // 16, 1)
        b = orc.bin_and_sort(xys.detach().numpy(), depths.numpy(), radii.numpy(), num_tiles_hit.numpy(), tb)
        bg = np.ones(colors.shape[-1], np.float32) if background is None else background.numpy()
        img, fT, fidx, _ = orc.raster_fwd(b, xys.detach().numpy(), conics.numpy(), colors.detach().numpy(),
                                          opacity.detach().numpy(), bg, H, W, tb)
        if return_alpha:
            return torch.from_numpy(img), torch.from_numpy(1.0 - fT)
        return torch.from_numpy(img)

    class Cameras:                              # the attributes gc_model.py reads
        def __init__(self, cam):
            self.camera_to_worlds = cam.c2w.clone()[None]
            self.fx, self.fy = torch.tensor([[cam.fx]]), torch.tensor([[cam.fy]])
            self.cx, self.cy = torch.tensor([[cam.cx]]), torch.tensor([[cam.cy]])
            self.width, self.height = torch.tensor([[cam.W]]), torch.tensor([[cam.H]])
            self.shape = (1,)

        def rescale_output_resolution(self, s):          # nerfstudio Cameras, "floor" rounding
            self.fx, self.fy, self.cx, self.cy = self.fx * s, self.fy * s, self.cx * s, self.cy * s
            self.width, self.height = (self.width * s).to(torch.int64), (self.height * s).to(torch.int64)

        def to(self, device):
            return self

    class Renderers:
        BACKGROUND_COLOR_OVERRIDE = None

    class Config:
        def __init__(self, background_color):
            self.background_color, self.sh_degree, self.sh_degree_interval = background_color, 3, 1000

    env = {"torch": torch, "math": math, "Dict": Dict, "Union": Union, "List": List, "Optional": Optional,
           "Cameras": Cameras, "renderers": Renderers, "OrientedBox": scene.OrientedBox,
           "projection_matrix": projection_matrix, "project_gaussians": project_gaussians,
           "spherical_harmonics": spherical_harmonics, "rasterize_gaussians": rasterize_gaussians}
    path = os.path.join(REF, "gc_model.py")
    tree = ast.parse(open(path).read())
    extract(path, {"projection_matrix_ogl"}, env)
    cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "GaussCtrlModel")
    methods = {}
    for node in cls.body:
        if isinstance(node, ast.FunctionDef) and node.name in ("get_outputs", "get_outputs_for_camera"):
            exec(compile(ast.Module(body=[node], type_ignores=[]), path, "exec"), env)
            methods[node.name] = env[node.name]

    class Stub:                                  # what GaussCtrlModel inherits from SplatfactoModel, as far as it is read
        device = torch.device("cpu")

        def __init__(self, p, background_color, num_downscales=0, step=30000):
            for k, v in p.items():
                setattr(self, k, torch.from_numpy(v))
            self.config = Config(background_color)
            self.background_color = (torch.tensor([0.1490, 0.1647, 0.2157]) if background_color == "random"
                                     else torch.tensor({"white": [1.0, 1.0, 1.0], "black": [0.0, 0.0, 0.0]}[background_color]))
            self.training, self.crop_box, self.step = True, None, step
            self.num_downscales, self.resolution_schedule = num_downscales, 250

        def set_crop(self, box):
            self.crop_box = box

        def _get_downscale_factor(self):          # nerfstudio splatfacto
            if self.training:
                return 2 ** max(self.num_downscales - self.step // self.resolution_schedule, 0)
            return 1
    Stub.get_outputs = methods["get_outputs"]
    Stub.get_outputs_for_camera = methods["get_outputs_for_camera"]

    N, W, H = 3000, 96, 64
    p = {k: v.numpy() for k, v in scene.make_scene(N, 31, s0=0.03).items()}
    cam = scene.make_cameras(3, 31, intrinsics=dict(fx=110.0, fy=109.0, cx=47.7, cy=32.2, W=W, H=H))[1]
    th = 0.3
    box = scene.OrientedBox(R=torch.tensor([[math.cos(th), -math.sin(th), 0.0], [math.sin(th), math.cos(th), 0.0],
                                            [0.0, 0.0, 1.0]]), T=torch.tensor([0.02, 0.03, -0.04]),
                            S=torch.tensor([0.45, 0.5, 0.4]))
    far = scene.OrientedBox(R=torch.eye(3), T=torch.tensor([50.0, 0.0, 0.0]), S=torch.ones(3))
    out = {"c2w": cam.c2w.numpy(), "intr": np.array([cam.fx, cam.fy, cam.cx, cam.cy, W, H], np.float64),
           "scene": np.array([N, 31, 0.03]), "box_R": box.R.numpy(), "box_T": box.T.numpy(), "box_S": box.S.numpy(),
           "far_T": far.T.numpy()}

    def put(tag, res):
        for k, v in res.items():
            if v is not None:
                out[f"{tag}_{k}"] = v.detach().numpy() if isinstance(v, torch.Tensor) else np.asarray(v)
        out[f"{tag}_keys"] = np.array(sorted(k for k, v in res.items() if v is not None))

    m = Stub(p, "random")
    put("eval", m.get_outputs_for_camera(Cameras(cam)))
    assert m.training and m.xys.shape == (N, 2)
    put("crop", m.get_outputs_for_camera(Cameras(cam), obb_box=box))
    out["crop_count"] = np.array(m.xys.shape[0])
    put("empty", m.get_outputs_for_camera(Cameras(cam), obb_box=far))
    mw = Stub(p, "white", num_downscales=1, step=3100)     # training: factor 2 ** max(1 - 3100 // 250, 0) = 1 ...
    assert mw._get_downscale_factor() == 1
    mw.step, mw.resolution_schedule = 3100, 10000           # ... and with a long schedule, 2
    assert mw._get_downscale_factor() == 2
    c = Cameras(cam)
    put("train_half", mw.get_outputs(c))
    assert int(c.width.item()) == W and int(c.height.item()) == H     # rescaled back (gc_model.py:206)
    np.savez_compressed(os.path.join(HERE, "reference_caller.npz"), **out)


def reference_checkpoint_golden():
    """6. reference_step-000029999.ckpt — written by the reference's OWN `GaussCtrlTrainer.save_checkpoint`
    (/root/reference/gaussctrl/gc_trainer.py:146-174), extracted by AST and executed unmodified. What it reads from
    nerfstudio is stubbed: `self.pipeline` is a torch Module whose `_model` carries the six splatfacto parameters under
    the names gc_model.py:157-170 reads (plus an `ip2p.` sub-module, which the reference filters out), `self.optimizers.
    optimizers` holds one real torch.optim.Adam per parameter group with the names, learning rates and eps of
    gc_config.py:58-87, stepped twice so that the moments are non-trivial, `self.grad_scaler` is torch's GradScaler.
    The file pins the checkpoint layout the loader (postops.load_splatfacto_checkpoint) has to read."""
    import pathlib
    import shutil
    import tempfile
    import torch
    from gaussctrl_exp_b200 import scene

    path = os.path.join(REF, "gc_trainer.py")
    tree = ast.parse(open(path).read())
    cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "GaussCtrlTrainer")
    node = next(n for n in cls.body if isinstance(n, ast.FunctionDef) and n.name == "save_checkpoint")
    env = {"torch": torch, "check_main_thread": lambda f: f}
    exec(compile(ast.Module(body=[node], type_ignores=[]), path, "exec"), env)

    N = 96
    p = scene.make_scene(N, 77, s0=0.05)

    class Model(torch.nn.Module):
        def __init__(self):
            super().__init__()
            for k, v in p.items():
                setattr(self, k, torch.nn.Parameter(v.clone()))

    class Pipeline(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self._model = Model()
            self.ip2p = torch.nn.Linear(2, 2)            # stands for the diffusion pipeline the reference drops

    pipe = Pipeline()
    groups = {"xyz": ("means", 1.6e-4), "features_dc": ("features_dc", 0.0025), "features_rest": ("features_rest", 0.0025 / 20),
              "opacity": ("opacities", 0.05), "scaling": ("scales", 0.005), "rotation": ("quats", 0.001)}
    opts = {g: torch.optim.Adam([getattr(pipe._model, name)], lr=lr, eps=1e-15) for g, (name, lr) in groups.items()}
    gen = torch.Generator().manual_seed(5)
    for _ in range(2):
        for g, (name, _) in groups.items():
            prm = getattr(pipe._model, name)
            prm.grad = torch.randn(prm.shape, generator=gen) * 1e-3
            opts[g].step()

    class Obj:
        pass
    tr = Obj()
    tmp = pathlib.Path(tempfile.mkdtemp())
    tr.checkpoint_dir = tmp / "nerfstudio_models"
    tr.pipeline = pipe
    tr.optimizers = Obj()
    tr.optimizers.optimizers = opts
    tr.grad_scaler = torch.amp.GradScaler("cpu", enabled=False)
    tr.config = Obj()
    tr.config.save_only_latest_checkpoint = True
    env["save_checkpoint"](tr, 29999)
    src = tr.checkpoint_dir / "step-000029999.ckpt"
    assert src.exists()
    shutil.copy(src, os.path.join(HERE, "reference_step-000029999.ckpt"))
    shutil.rmtree(tmp)


def main():
    import torch
    from typing import Union
    env_model = extract(os.path.join(REF, "gc_model.py"), {"projection_matrix_ogl"},
                        {"torch": torch, "math": math, "Union": Union})
    env_ad = extract(os.path.join(REF, "ad_render.py"), {"projection_matrix_splatfacto", "projection_matrix_ogl"},
                     {"np": np, "math": math})
    cases = []
    for W, H, fx, fy in [(512, 512, 585.78, 585.51), (1920, 1080, 1200.0, 1200.0), (64, 48, 70.4, 69.7)]:
        fovx = 2 * math.atan(W / (2 * fx))
        fovy = 2 * math.atan(H / (2 * fy))
        cases.append((W, H, fx, fy, fovx, fovy))
    out = {"cases": np.array(cases, np.float64)}
    for i, (W, H, fx, fy, fovx, fovy) in enumerate(cases):
        out[f"gc_model_ogl_{i}"] = env_model["projection_matrix_ogl"](0.001, 1000, fovx, fovy).numpy()
        out[f"ad_render_splatfacto_{i}"] = env_ad["projection_matrix_splatfacto"](fovx, fovy)
        out[f"ad_render_ogl_{i}"] = env_ad["projection_matrix_ogl"](fovx, fovy)
    np.savez(os.path.join(HERE, "reference_matrices.npz"), **out)

    # 3. reference_postops.npz — the reference's own depth post-ops (gc_pipeline.py:286-314, :449-457,
    #    :503-536), extracted by AST and executed unmodified on seeded inputs.
    from einops import rearrange
    env_p = {"np": np, "torch": torch, "rearrange": rearrange}
    extract(os.path.join(REF, "gc_pipeline.py"), {"compute_ray_directions"}, env_p)
    tree = ast.parse(open(os.path.join(REF, "gc_pipeline.py")).read())
    for node in ast.walk(tree):
        if isinstance(node, ast.FunctionDef) and node.name in ("depth_map_screen_to_world", "depth2disparity_torch"):
            exec(compile(ast.Module(body=[node], type_ignores=[]), "gc_pipeline.py", "exec"), env_p)
    g = torch.Generator().manual_seed(9)
    h = w = 48
    depth = torch.rand(h, w, 1, generator=g) * 3 + 0.2
    c2w = torch.eye(4)
    A = torch.linalg.qr(torch.randn(3, 3, generator=g))[0]
    c2w[:3, :3] = A
    c2w[:3, 3] = torch.randn(3, generator=g)
    mat_view = torch.linalg.inv(c2w)
    rays = env_p["compute_ray_directions"](resolution=(h, w))
    pts = env_p["depth_map_screen_to_world"](None, depth, mat_view)
    disp = env_p["depth2disparity_torch"](None, depth[:, :, 0][None])
    np.savez(os.path.join(HERE, "reference_postops.npz"), depth=depth.numpy(), mat_view=mat_view.numpy(), rays=rays,
             pts_world=pts.numpy(), disparity=disp.numpy(), rays_rect=env_p["compute_ray_directions"](resolution=(20, 12)))

    # 4. reference_noise_helpers.npz — the numpy-only helpers of the noise-mask renderer, executed unmodified:
    #    ad_noise.py:5-24 (create_sphere), :53-66 (generate_uniform_points_in_cube), ad_render.py:164-205
    #    (extract_camera_parameters, compute_view_matrix_ogl). The GL draw itself cannot run here.
    env_n = extract(os.path.join(REF, "ad_noise.py"), {"create_sphere", "generate_uniform_points_in_cube"}, {"np": np})
    env_v = extract(os.path.join(REF, "ad_render.py"), {"extract_camera_parameters", "compute_view_matrix_ogl"}, {"np": np})
    sv, si = env_n["create_sphere"](0.015, 8, 8)
    sv2, si2 = env_n["create_sphere"](0.3, 5, 7)
    cube = env_n["generate_uniform_points_in_cube"](2.0, 6)
    gq = torch.Generator().manual_seed(11)
    c2ws, views = [], []
    for _ in range(3):
        m = np.eye(4)
        m[:3, :3] = torch.linalg.qr(torch.randn(3, 3, generator=gq))[0].numpy()
        m[:3, 3] = torch.randn(3, generator=gq).numpy()
        pos, up, fw = env_v["extract_camera_parameters"](m)
        c2ws.append(m)
        views.append(env_v["compute_view_matrix_ogl"](pos, up, fw))
    np.savez(os.path.join(HERE, "reference_noise_helpers.npz"), sphere_v=sv, sphere_i=si, sphere2_v=sv2, sphere2_i=si2,
             cube=cube, c2ws=np.stack(c2ws), views=np.stack(views))

    reference_caller_golden()
    reference_checkpoint_golden()

    from oracle import oracle as orc
    from util import np_params, small_cam
    N, W, H = 600, 64, 48
    p = np_params(N, seed=21, s0=0.05)
    cam = small_cam(W, H, 21)
    res, ctx = orc.render_view(p, cam.as_dict(), 3, training=False)
    v = (np.random.default_rng(5).standard_normal((H, W, 3)) / (H * W)).astype(np.float32)
    g = orc.render_view_backward(p, ctx, v)
    b = ctx["b"]
    np.savez_compressed(
        os.path.join(HERE, "oracle_small_scene.npz"),
        c2w=cam.c2w.numpy(), intr=np.array([cam.fx, cam.fy, cam.cx, cam.cy, W, H], np.float64), v_out=v,
        xys=ctx["xys"], depths=ctx["depths"], radii=ctx["radii"], conics=ctx["conics"], hit=ctx["hit"],
        keys_sorted=b.keys_sorted, ids_sorted=b.ids_sorted, tile_bins=b.tile_bins,
        rgb=res["rgb"], depth=res["depth"], acc=res["accumulation"], fidx=ctx["fidx"],
        **{"g_" + k: val for k, val in g.items()}, **{"p_" + k: val for k, val in p.items()})
    print("wrote", os.listdir(HERE))


if __name__ == "__main__":
    main()
