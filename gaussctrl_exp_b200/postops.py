"""Callers either side of the render path: what the reference does with a
rendered view before ControlNet, and where the Gaussians come from.

* depth -> world points      /root/reference/gaussctrl/gc_pipeline.py:286-314 (+ ray builder :503-536)
* depth -> disparity         gc_pipeline.py:449-457
* splatfacto checkpoint I/O  gc_trainer.py:147-174 (torch.save dict; parameter names gc_model.py:157-170)

CUDA tensors go through the C ABI (csrc/postops.cu: gs_ray_directions, gs_depth_to_world, gs_depth_to_disparity —
one streaming kernel each, no host synchronisation; a missing library raises): that is the PRODUCT path, and a CUDA
tensor never falls back to the branch below. CPU tensors take the host mirror of the same arithmetic in torch ops (the
reference's per-pixel Python double loop vectorised) — the reference's own post-ops run on CPU numpy arrays
(gc_pipeline.py:230-236 calls `.cpu().numpy()` first), so a caller that still hands over CPU data gets the reference's
behaviour; it is kept for that caller and for the golden-vector tests, not as a fallback. Both are pinned
against the reference's own functions by tests/golden/reference_postops.npz; the checkpoint reader against a file
written by the reference's own `save_checkpoint` (tests/golden/reference_step-000029999.ckpt).
"""
from __future__ import annotations

import math
from typing import Dict, Tuple

import torch
from torch import Tensor

_ray_cache: Dict[Tuple, Tensor] = {}
_FOV_Y, _NEAR = 0.8880228256678113, 0.001          # compute_ray_directions' defaults, which its only caller uses

PARAM_KEYS = ("means", "scales", "quats", "features_dc", "features_rest", "opacities")


def compute_ray_directions(resolution: Tuple[int, int], fov_y_rad: float = 0.8880228256678113, near: float = 0.001,
                           far: float = 1000.0, device="cpu") -> Tensor:
    """gc_pipeline.py:503-536, vectorised. NOTE the reference unpacks `width, height = resolution` while its
    caller passes (h, w) (gc_pipeline.py:299): reproduced as is (identical for the square 512x512 views)."""
    width, height = resolution
    device = torch.device(device)
    key = (width, height, fov_y_rad, near, str(device))
    hit = _ray_cache.get(key)
    if hit is not None:
        return hit
    if device.type == "cuda":
        from . import _lib
        d = torch.empty((height, width, 3), dtype=torch.float32, device=device)
        with torch.cuda.device(device):
            _lib.check(_lib.lib().gs_ray_directions(width, height, fov_y_rad, near, d.data_ptr(), _lib.stream_ptr()),
                       "gs_ray_directions")
        _ray_cache[key] = d
        return d
    aspect_ratio = width / height
    viewport_height = 2 * math.tan(fov_y_rad / 2) * near
    viewport_width = viewport_height * aspect_ratio
    center_x, center_y = viewport_width / 2, viewport_height / 2
    xs = (torch.arange(width, dtype=torch.float64) + 0.5) / width * viewport_width - center_x
    ys = (torch.arange(height, dtype=torch.float64) + 0.5) / height * viewport_height - center_y
    d = torch.empty((height, width, 3), dtype=torch.float32)
    d[..., 0] = xs.float()[None, :]
    d[..., 1] = ys.float()[:, None]
    d[..., 2] = -near
    # np.linalg.norm's float32 order of operations, so that the host mirror is bit-identical to the reference too
    # (torch's vectorised float32 sqrt on the CPU is not correctly rounded: taken in double and rounded once)
    n = torch.sqrt(((d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]) + d[..., 2] * d[..., 2]).double()).float()
    d = d / n[..., None]
    d = d.to(device)
    _ray_cache[key] = d
    return d


def depth_map_screen_to_world(depth_map: Tensor, mats_view: Tensor) -> Tensor:
    """gc_pipeline.py:286-314: depth (h,w,1) along the per-pixel ray, un-projected with inverse(mat_view)
    (the un-flipped world->camera matrix returned as outputs['mat_view']). -> (h,w,3)."""
    h, w, _ = depth_map.shape
    if depth_map.is_cuda:
        from . import _lib
        depth = depth_map.detach().float().contiguous()
        view = mats_view.detach().to("cpu", torch.float32).contiguous()      # 64 bytes; host-resident in the model's outputs
        out = torch.empty((h, w, 3), dtype=torch.float32, device=depth.device)
        with torch.cuda.device(depth.device):
            _lib.check(_lib.lib().gs_depth_to_world(h, w, depth.data_ptr(), view.data_ptr(), _FOV_Y, _NEAR, out.data_ptr(),
                                                    _lib.stream_ptr()), "gs_depth_to_world")
        return out
    rays_d = compute_ray_directions((h, w), device=depth_map.device)
    pts_view = rays_d * depth_map
    pts_view = torch.cat((pts_view, torch.ones((h, w, 1), dtype=rays_d.dtype, device=rays_d.device)), dim=-1)
    inv_view = torch.linalg.inv(mats_view.to(depth_map.device).float())
    pts_world = pts_view.reshape(-1, 4).float() @ inv_view.T
    pts_world = pts_world / pts_world[:, 3:4]
    return pts_world.reshape(h, w, 4)[:, :, :3]


def depth2disparity(depth: Tensor) -> Tensor:
    """gc_pipeline.py:449-457: depth (1,h,w) -> normalised disparity (1,3,h,w) for the depth ControlNet."""
    if depth.is_cuda:
        from . import _lib
        d = depth.detach().float().contiguous()
        out = torch.empty((1, 3 * d.shape[0]) + tuple(d.shape[1:]), dtype=torch.float32, device=d.device)
        scratch = torch.empty(1, dtype=torch.int32, device=d.device)
        with torch.cuda.device(d.device):
            _lib.check(_lib.lib().gs_depth_to_disparity(d.numel(), d.data_ptr(), out.data_ptr(), scratch.data_ptr(),
                                                        _lib.stream_ptr()), "gs_depth_to_disparity")
        return out
    disparity = 1 / (depth + 1e-5)
    disparity_map = disparity / torch.max(disparity)
    return torch.cat([disparity_map, disparity_map, disparity_map], dim=0)[None]


# optimiser group of gc_config.py:58-87 -> the parameter it holds
OPTIMIZER_GROUPS = {"xyz": "means", "features_dc": "features_dc", "features_rest": "features_rest", "opacity": "opacities",
                    "scaling": "scales", "rotation": "quats"}


def load_splatfacto_checkpoint(path: str, map_location="cpu", with_optimizers: bool = False):
    """Reads the dict written at gc_trainer.py:158-168 ({'step', 'pipeline', 'optimizers', 'scalers'}) and
    returns (the six leaf tensors, step). Accepts nerfstudio 1.0.0 names (`_model.means`) and the later
    `_model.gauss_params.means` spelling.

    with_optimizers=True adds a third value: {parameter name: {"exp_avg", "exp_avg_sq", "step", "lr", "betas", "eps"}}
    read from the per-group torch.optim.Adam state dicts the reference saves (what `_load_checkpoint`,
    gc_trainer.py:74, hands to `optimizers.load_optimizers`), keyed by OUR parameter names; groups the file does not
    hold (e.g. a checkpoint exported without optimiser state) are absent from it. dist.ShardedAdam.load_reference_state
    consumes it."""
    ckpt = torch.load(path, map_location=map_location, weights_only=False)
    sd = ckpt["pipeline"] if "pipeline" in ckpt else ckpt
    out = {}
    for k in PARAM_KEYS:
        for name in (f"_model.{k}", f"_model.gauss_params.{k}", f"module._model.{k}", k):
            if name in sd:
                out[k] = sd[name].detach().float().contiguous()
                break
        else:
            raise KeyError(f"checkpoint has no parameter '{k}' (keys: {list(sd)[:8]} ...)")
    n = out["means"].shape[0]
    if out["features_dc"].ndim == 3:           # some exporters keep (N,1,3)
        out["features_dc"] = out["features_dc"].reshape(n, 3)
    out["opacities"] = out["opacities"].reshape(n, 1)
    step = int(ckpt.get("step", 0))
    if not with_optimizers:
        return out, step
    opt = {}
    for group, name in OPTIMIZER_GROUPS.items():
        osd = ckpt.get("optimizers", {}).get(group)
        if not osd or not osd.get("state"):
            continue
        st = next(iter(osd["state"].values()))           # one parameter per group
        pg = osd["param_groups"][0]
        opt[name] = {"exp_avg": st["exp_avg"].detach().float().reshape(out[name].shape).contiguous(),
                     "exp_avg_sq": st["exp_avg_sq"].detach().float().reshape(out[name].shape).contiguous(),
                     "step": int(st["step"]), "lr": float(pg["lr"]), "betas": tuple(pg["betas"]), "eps": float(pg["eps"])}
    return out, step, opt


def save_splatfacto_checkpoint(path: str, params: Dict[str, Tensor], step: int) -> None:
    """Writes the same layout (pipeline state dict with `_model.<name>` keys) so the reference's
    `_load_checkpoint` (gc_trainer.py:75) can read it back."""
    sd = {f"_model.{k}": params[k].detach().cpu() for k in PARAM_KEYS}
    torch.save({"step": int(step), "pipeline": sd, "optimizers": {}, "scalers": {}}, path)
