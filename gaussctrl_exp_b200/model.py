"""Host-side mirror of the reference's render host code for ONE camera.

`GaussianModel.get_outputs` follows /root/reference/gaussctrl/gc_model.py:77-241 step by step
(background, matrices, pre-activations, project -> SH -> rasterize RGB(+alpha) -> depth pass in
eval) and `get_outputs_for_camera` follows :243-255, with nerfstudio's `Cameras` replaced by the
plain `scene.Camera` record and `SplatfactoModel`'s parameters by a dict of the same six leaf
tensors (gc_model.py:157-170). The three operators are this package's gsplat-compatible
functions; nothing here touches a CPU fallback.
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Union

import torch
from torch import Tensor, nn

from . import render as _render
from .project_gaussians import project_gaussians
from .rasterize import rasterize_gaussians
from .scene import Camera
from .sh import num_sh_bases, spherical_harmonics


def projection_matrix(znear: float, zfar: float, fovx: float, fovy: float, device="cpu") -> Tensor:
    """nerfstudio's splatfacto `projection_matrix` (imported at gc_model.py:31; the same matrix is
    restated in-tree at /root/reference/gaussctrl/ad_render.py:49-67)."""
    t = znear * math.tan(0.5 * fovy)
    b = -t
    r = znear * math.tan(0.5 * fovx)
    l = -r
    n, f = znear, zfar
    return torch.tensor(
        [
            [2 * n / (r - l), 0.0, (r + l) / (r - l), 0.0],
            [0.0, 2 * n / (t - b), (t + b) / (t - b), 0.0],
            [0.0, 0.0, (f + n) / (f - n), -1.0 * f * n / (f - n)],
            [0.0, 0.0, 1.0, 0.0],
        ],
        device=device,
    )


def projection_matrix_ogl(znear: float, zfar: float, fovx: float, fovy: float, device="cpu") -> Tensor:
    """gc_model.py:39-57 — OpenGL clip-space matrix returned to the caller as `mat_proj`."""
    t = znear * math.tan(0.5 * fovy)
    b = -t
    r = znear * math.tan(0.5 * fovx)
    l = -r
    n, f = znear, zfar
    return torch.tensor(
        [
            [2 * n / (r - l), 0.0, (r + l) / (r - l), 0.0],
            [0.0, 2 * n / (t - b), (t + b) / (t - b), 0.0],
            [0.0, 0.0, -(f + n) / (f - n), -2.0 * f * n / (f - n)],
            [0.0, 0.0, -1.0, 0.0],
        ],
        device=device,
    )


def camera_matrices(c2w: Tensor, fx: float, fy: float, W: int, H: int):
    """gc_model.py:117-155 on the host (CPU tensors): returns
    (viewmat 4x4 with y/z flipped, projmat 4x4, viewmat0 un-flipped W2C, projmat0 OGL, tile_bounds)."""
    c2w = c2w.detach().float().cpu()
    R = c2w[:3, :3]
    T = c2w[:3, 3:4]
    R_inv0 = R.T                                  # :121-126 (returned as mat_view)
    viewmat0 = torch.eye(4, dtype=R.dtype)
    viewmat0[:3, :3] = R_inv0
    viewmat0[:3, 3:4] = -R_inv0 @ T
    R_edit = torch.diag(torch.tensor([1, -1, -1], dtype=R.dtype))   # :129-130
    R = R @ R_edit
    R_inv = R.T
    T_inv = -R_inv @ T
    viewmat = torch.eye(4, dtype=R.dtype)
    viewmat[:3, :3] = R_inv
    viewmat[:3, 3:4] = T_inv
    fovx = 2 * math.atan(W / (2 * fx))            # :143-144
    fovy = 2 * math.atan(H / (2 * fy))
    projmat = projection_matrix(0.001, 1000, fovx, fovy)             # :147
    projmat0 = projection_matrix_ogl(0.001, 1000, fovx, fovy)        # :148
    tile_bounds = ((W + 15) // 16, (H + 15) // 16, 1)                # :150-155
    return viewmat, projmat, viewmat0, projmat0, tile_bounds


# nerfstudio `renderers.BACKGROUND_COLOR_OVERRIDE` (read at gc_model.py:102-104): a module-level tensor that, when set,
# replaces the background of every eval render.
BACKGROUND_COLOR_OVERRIDE: Optional[Tensor] = None

# nerfstudio `get_color` names that splatfacto accepts for `background_color`
_NAMED_COLORS = {"white": (1.0, 1.0, 1.0), "black": (0.0, 0.0, 0.0), "red": (1.0, 0.0, 0.0), "green": (0.0, 1.0, 0.0),
                 "blue": (0.0, 0.0, 1.0)}


class GaussianModel(nn.Module):
    """The six splatfacto leaf tensors + the render host code of GaussCtrlModel."""

    def __init__(self, params: Dict[str, Tensor], sh_degree: int = 3, sh_degree_interval: int = 1000,
                 background_color: Union[str, Tensor] = "random", step: int = 30000, fused: bool = True,
                 num_downscales: int = 0, resolution_schedule: int = 250):
        super().__init__()
        self.means = nn.Parameter(params["means"].clone())
        self.scales = nn.Parameter(params["scales"].clone())
        self.quats = nn.Parameter(params["quats"].clone())
        self.features_dc = nn.Parameter(params["features_dc"].clone())
        self.features_rest = nn.Parameter(params["features_rest"].clone())
        self.opacities = nn.Parameter(params["opacities"].clone())
        self.sh_degree = sh_degree
        self.sh_degree_interval = sh_degree_interval
        # `config.background_color` (a name, or a tensor for a fixed colour) and splatfacto's resolved
        # `self.background_color` tensor (populate_modules): "random" resolves to the viewer grey
        # (0.1490, 0.1647, 0.2157) — random colours are drawn per training view only (gc_model.py:93-106)
        self.config_background_color = background_color
        if isinstance(background_color, Tensor):
            self.background_color = background_color.detach().float()
        elif background_color == "random":
            self.background_color = torch.tensor([0.1490, 0.1647, 0.2157])
        elif background_color in _NAMED_COLORS:
            self.background_color = torch.tensor(_NAMED_COLORS[background_color])
        else:
            raise ValueError(f"unknown background_color {background_color!r}")
        self.step = step                       # a 30k-step splatfacto checkpoint (scripts/garden.sh:3)
        # splatfacto's coarse-to-fine schedule (`_get_downscale_factor`, read at gc_model.py:114): off (factor 1) for a
        # fine-tune that starts from a 30k-step checkpoint, and by default
        self.num_downscales = num_downscales
        self.resolution_schedule = resolution_schedule
        self.crop_box = None                   # OrientedBox or None (`set_crop`; only read in eval, gc_model.py:108)
        # fused=True: one autograd node / two C calls per view (render.py); fused=False: the
        # reference's literal three-operator sequence through the gsplat-compatible functions.
        self.fused = fused
        # fused training views: False (default) = the forward confirms the tile-list capacity before it returns (one
        # host wait that ends when the BINNING is done; an overflow is healed by redoing the view); True = nothing is
        # read in the forward (the host runs ahead of the device), an overflow NaN-poisons the image and raises
        # render.CapacityOverflow at the end of backward — for training loops that catch it and redo the step
        self.speculate = False
        self.view_state = _render.ViewState()
        self._mat_cache = {}
        self._streams = []
        self.xys: Optional[Tensor] = None
        self.radii: Optional[Tensor] = None
        self.last_size = None

    @property
    def device(self):
        return self.means.device

    @property
    def num_points(self):
        return self.means.shape[0]

    def set_crop(self, crop_box) -> None:
        """nerfstudio `SplatfactoModel.set_crop` (gc_model.py:250)."""
        self.crop_box = crop_box

    def _get_downscale_factor(self) -> int:
        """nerfstudio splatfacto: 2 ** max(num_downscales - step // resolution_schedule, 0) while training, 1 in eval."""
        if self.training:
            return 2 ** max(self.num_downscales - self.step // self.resolution_schedule, 0)
        return 1

    def _background(self) -> Tensor:
        """gc_model.py:93-106."""
        dev = self.device
        if self.background_color.device != dev:      # move once: a pageable H2D copy per view would sync the stream
            self.background_color = self.background_color.to(dev)
        if self.training:
            bc = self.config_background_color
            if isinstance(bc, str):
                if bc == "random":
                    return torch.rand(3, device=dev)
                if bc == "white":
                    return torch.ones(3, device=dev)
                if bc == "black":
                    return torch.zeros(3, device=dev)
            return self.background_color
        if BACKGROUND_COLOR_OVERRIDE is not None:
            return BACKGROUND_COLOR_OVERRIDE.to(dev)
        return self.background_color

    def _caller_matrices(self, camera: Camera):
        """mat_view / mat_proj of gc_model.py:121-126, :148 (cached by the camera's VALUES)."""
        key = camera.signature()
        hit = self._mat_cache.get(key)
        if hit is None:
            _, _, viewmat0, projmat0, _ = camera_matrices(camera.c2w, camera.fx, camera.fy, int(camera.W), int(camera.H))
            if len(self._mat_cache) > 4096:
                self._mat_cache.clear()
            hit = (viewmat0, projmat0)
            self._mat_cache[key] = hit
        return hit

    def _crop_ids(self) -> Optional[Tensor]:
        """gc_model.py:108-113: boolean mask of the Gaussians inside the crop box (eval only)."""
        if self.crop_box is not None and not self.training:
            return self.crop_box.within(self.means.detach()).squeeze()
        return None

    def _get_outputs_fused(self, camera: Camera, background: Tensor, crop_ids: Optional[Tensor]):
        """Same outputs as the literal path below, through render.render_view / render.render_eval.
        `camera` already carries the downscaled resolution. (None, None, None) = no Gaussian of a non-empty set reaches
        the image (gc_model.py:189-190 `radii.sum() == 0`): a radius is positive exactly when the splat hits a tile, so
        the intersection count the call has already read is the same test without another device read."""
        W, H = int(camera.W), int(camera.H)
        dev = self.device
        use_sh = self.sh_degree > 0
        n = min(self.step // self.sh_degree_interval, self.sh_degree) if use_sh else 0
        if torch.is_grad_enabled() and self.training:
            rgb, alpha, depths, radii, hit = _render.render_view(
                self.means, self.scales, self.quats, self.opacities, self.features_dc, self.features_rest, camera,
                n, use_sh, background, self.view_state, clamp_rgb=True,       # gc_model.py:222 inside the kernels
                speculate=self.speculate)
            self.xys, self.radii = self.view_state.xys, radii
            if self.num_points > 0 and self.view_state.num_intersects == 0:   # strict forward only: a speculative one has no count yet
                return None, None, None
            return rgb, None, alpha[..., None]
        subset = None
        if crop_ids is not None:            # index list into the preprocess kernel instead of six gathered copies
            subset = torch.nonzero(crop_ids, as_tuple=False).view(-1).to(torch.int32)
        count = self.num_points if subset is None else int(subset.shape[0])
        if count == 0:               # empty model: background, zero accumulation, depth 1000 (gc_model.py:238)
            with torch.no_grad():
                rgb = torch.clamp(background.to(dev).float().expand(H, W, 3).clone(), max=1.0)
                alpha = torch.zeros((H, W, 1), device=dev)
                depth_im = None if self.training else torch.full((H, W, 1), 1000.0, device=dev)
            self.xys = torch.zeros((0, 2), device=dev)
            self.radii = torch.zeros((0,), device=dev, dtype=torch.int32)
            return rgb, depth_im, alpha
        ws, rgb, alpha, depth_im = _render.render_eval(
            self.means, self.scales, self.quats, self.opacities, self.features_dc, self.features_rest, [camera], n,
            use_sh, background, with_depth=not self.training, subset=subset)
        self.xys, self.radii = ws.xys, ws.radii
        if ws.num_intersects == 0:
            return None, None, None
        return rgb, depth_im, alpha

    def get_outputs(self, camera: Camera) -> Dict[str, Optional[Tensor]]:
        background = self._background()                               # gc_model.py:93-106
        crop_ids = self._crop_ids()                                   # :108-113
        if crop_ids is not None and crop_ids.sum() == 0:
            return {"rgb": background.repeat(int(camera.H), int(camera.W), 1)}
        camera_downscale = self._get_downscale_factor()               # :114-115
        camera.rescale_output_resolution(1 / camera_downscale)
        try:
            W, H = int(camera.W), int(camera.H)
            self.last_size = (H, W)
            viewmat0, projmat0 = self._caller_matrices(camera)
            if self.fused:
                rgb, depth_im, alpha = self._get_outputs_fused(camera, background, crop_ids)
                if rgb is None:                                       # :189-190 every Gaussian culled
                    return {"rgb": background.repeat(H, W, 1)}
                return {"rgb": rgb, "depth": depth_im, "accumulation": alpha,
                        "mat_view": viewmat0, "mat_proj": projmat0, "mat_c2w": camera.c2w}
            return self._get_outputs_literal(camera, background, crop_ids, viewmat0, projmat0)
        finally:
            camera.rescale_output_resolution(camera_downscale)        # :206 (back to the original dimensions)

    def _get_outputs_literal(self, camera: Camera, background: Tensor, crop_ids, viewmat0, projmat0):
        """gc_model.py:117-241 line by line through the gsplat-compatible operators."""
        W, H = int(camera.W), int(camera.H)
        viewmat, projmat, _, _, tile_bounds = camera_matrices(camera.c2w, camera.fx, camera.fy, W, H)
        dev = self.device
        viewmat_d = viewmat.to(dev)
        full_proj = (projmat @ viewmat).to(dev)                       # :180
        if crop_ids is not None:                                      # :157-170
            opacities_crop = self.opacities[crop_ids]
            means_crop = self.means[crop_ids]
            features_dc_crop = self.features_dc[crop_ids]
            features_rest_crop = self.features_rest[crop_ids]
            scales_crop = self.scales[crop_ids]
            quats_crop = self.quats[crop_ids]
        else:
            opacities_crop, means_crop = self.opacities, self.means
            features_dc_crop, features_rest_crop = self.features_dc, self.features_rest
            scales_crop, quats_crop = self.scales, self.quats
        colors_crop = torch.cat((features_dc_crop[:, None, :], features_rest_crop), dim=1)   # :172

        self.xys, depths, self.radii, conics, num_tiles_hit, cov3d = project_gaussians(
            means_crop,
            torch.exp(scales_crop),
            1,
            quats_crop / quats_crop.norm(dim=-1, keepdim=True),
            viewmat_d[:3, :],
            full_proj,
            camera.fx, camera.fy, camera.cx, camera.cy,
            H, W, tile_bounds,
        )                                                             # :174-188
        if (self.radii).sum() == 0:                                   # :189-190
            return {"rgb": background.repeat(H, W, 1)}
        if self.training and self.xys.requires_grad:
            self.xys.retain_grad()                                    # :193-194

        if self.sh_degree > 0:
            cam_pos = camera.c2w[:3, 3].to(dev)
            viewdirs = means_crop.detach() - cam_pos                  # :197
            viewdirs = viewdirs / viewdirs.norm(dim=-1, keepdim=True)
            n = min(self.step // self.sh_degree_interval, self.sh_degree)
            rgbs = spherical_harmonics(n, viewdirs, colors_crop)      # :199-200
            rgbs = torch.clamp(rgbs + 0.5, min=0.0)                   # :201
        else:
            rgbs = torch.sigmoid(colors_crop[:, 0, :])                # :203

        rgb, alpha = rasterize_gaussians(
            self.xys, depths, self.radii, conics, num_tiles_hit, rgbs, torch.sigmoid(opacities_crop),
            H, W, background=background, return_alpha=True,
        )                                                             # :208-220
        alpha = alpha[..., None]
        rgb = torch.clamp(rgb, max=1.0)                               # :222
        depth_im = None
        if not self.training:                                         # :224-238
            depth_im = rasterize_gaussians(
                self.xys, depths, self.radii, conics, num_tiles_hit, depths[:, None].repeat(1, 3),
                torch.sigmoid(opacities_crop), H, W, background=torch.zeros(3, device=dev),
            )[..., 0:1]
            depth_im = torch.where(alpha > 0, depth_im / alpha, torch.full_like(depth_im, 1000.0))
        return {"rgb": rgb, "depth": depth_im, "accumulation": alpha,
                "mat_view": viewmat0, "mat_proj": projmat0, "mat_c2w": camera.c2w}   # :240-241

    @torch.no_grad()
    def get_outputs_for_camera(self, camera: Camera, obb_box=None) -> Dict[str, Optional[Tensor]]:
        """gc_model.py:243-255: eval render (RGB + depth + accumulation) of one camera, cropped to `obb_box`.
        Deliberate deviation: the reference ends with `self.training = True` whatever the flag was (:253-255); here
        the previous flag is restored, so a model put in eval mode stays in eval mode."""
        assert camera is not None, "must provide camera to gaussian model"
        self.set_crop(obb_box)                                        # :250
        was_training = self.training
        self.training = False
        try:
            return self.get_outputs(camera)
        finally:
            self.training = was_training

    def forward(self, camera: Camera):
        return self.get_outputs(camera)

    @torch.no_grad()
    def render_views(self, cameras, rank: int = 0, world: int = 1, batch: int = 8, obb_box=None,
                     keep=("rgb", "depth", "accumulation")):
        """The all-view render loop of gc_pipeline.py:222-228 / gc_render.py:196-203, view-sharded (rank renders
        views v with v % world == rank) and BATCHED: `batch` cameras of one image size go through ONE call — one
        projection pass per camera into a common record array, ONE binning over cameras x Gaussians with the
        camera folded into the tile id, ONE compositing launch over cameras x tiles — with one host read (the
        intersection count) per batch instead of several per view. Returns {view index: {name: tensor}} for this
        rank's views; per-view results are identical to get_outputs_for_camera."""
        from .dist import shard_views
        dev = self.device
        mine = shard_views(len(cameras), rank, world)
        self.set_crop(obb_box)
        was_training = self.training
        self.training = False
        results = {}
        try:
            background = self._background()
            crop_ids = self._crop_ids()
            subset = None
            if crop_ids is not None:
                subset = torch.nonzero(crop_ids, as_tuple=False).view(-1).to(torch.int32)
            count = self.num_points if subset is None else int(subset.shape[0])
            use_sh = self.sh_degree > 0
            n = min(self.step // self.sh_degree_interval, self.sh_degree) if use_sh else 0
            # group by image size, keep the caller's order inside a group
            groups = {}
            for v in mine:
                groups.setdefault((int(cameras[v].H), int(cameras[v].W)), []).append(v)
            for (H, W), vs in groups.items():
                for b0 in range(0, len(vs), max(1, batch)):
                    chunk = vs[b0:b0 + max(1, batch)]
                    if count == 0:
                        for v in chunk:
                            results[v] = self.get_outputs(cameras[v])
                        continue
                    ws, rgb, alpha, depth = _render.render_eval(
                        self.means, self.scales, self.quats, self.opacities, self.features_dc, self.features_rest,
                        [cameras[v] for v in chunk], n, use_sh, background, with_depth=True, subset=subset)
                    if len(chunk) == 1:
                        rgb, alpha, depth = rgb[None], alpha[None], depth[None]
                    for k, v in enumerate(chunk):
                        full = {"rgb": rgb[k], "depth": depth[k], "accumulation": alpha[k]}
                        results[v] = {name: full[name] for name in keep}
        finally:
            self.training = was_training
        return results


__all__ = ["GaussianModel", "camera_matrices", "projection_matrix", "projection_matrix_ogl", "num_sh_bases"]
