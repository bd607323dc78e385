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
