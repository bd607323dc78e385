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


class GaussianModel(nn.Module):
    """The six splatfacto leaf tensors + the render host code of GaussCtrlModel."""

    def __init__(self, params: Dict[str, Tensor], sh_degree: int = 3, sh_degree_interval: int = 1000,
                 background_color: Union[str, Tensor] = "random", step: int = 30000, fused: bool = True):
        super().__init__()
        self.means = nn.Parameter(params["means"].clone())
        self.scales = nn.Parameter(params["scales"].clone())
        self.quats = nn.Parameter(params["quats"].clone())
        self.features_dc = nn.Parameter(params["features_dc"].clone())
        self.features_rest = nn.Parameter(params["features_rest"].clone())
        self.opacities = nn.Parameter(params["opacities"].clone())
        self.sh_degree = sh_degree
        self.sh_degree_interval = sh_degree_interval
        self.background_color = background_color
        self.step = step                       # a 30k-step splatfacto checkpoint (scripts/garden.sh:3)
        # fused=True: one autograd node / five kernels per view (render.py); fused=False: the
        # reference's literal three-operator sequence through the gsplat-compatible functions.
        self.fused = fused
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

    def _background(self) -> Tensor:
        bc = self.background_color                                    # gc_model.py:93-106
        if isinstance(bc, Tensor):
            if bc.device != self.device:      # move once: a pageable H2D copy per view would sync the stream
                self.background_color = bc = bc.to(self.device)
            return bc
        if self.training and bc == "random":
            return torch.rand(3, device=self.device)
        if bc == "white":
            return torch.ones(3, device=self.device)
        return torch.zeros(3, device=self.device)

    def _caller_matrices(self, camera: Camera):
        """mat_view / mat_proj of gc_model.py:121-126, :148 (cached per camera object)."""
        hit = self._mat_cache.get(id(camera))
        if hit is None or hit[0] is not camera:
            _, _, viewmat0, projmat0, _ = camera_matrices(camera.c2w, camera.fx, camera.fy, int(camera.W), int(camera.H))
            if len(self._mat_cache) > 4096:
                self._mat_cache.clear()
            hit = (camera, viewmat0, projmat0)
            self._mat_cache[id(camera)] = hit
        return hit[1], hit[2]

    def _get_outputs_fused(self, camera: Camera) -> Dict[str, Optional[Tensor]]:
        """Same outputs as the literal path below, through render.render_view."""
        background = self._background()
        W, H = int(camera.W), int(camera.H)
        self.last_size = (H, W)
        dev = self.device
        use_sh = self.sh_degree > 0
        n = min(self.step // self.sh_degree_interval, self.sh_degree) if use_sh else 0
        cam = _render.gs_camera(camera)
        if torch.is_grad_enabled() and self.training:
            rgb, alpha, depths, radii, hit = _render.render_view(
                self.means, self.scales, self.quats, self.opacities, self.features_dc, self.features_rest, camera,
                n, use_sh, background, self.view_state, clamp_rgb=True)       # gc_model.py:222 inside the kernels
            self.xys, self.radii = self.view_state.xys, radii
            viewmat0, projmat0 = self._caller_matrices(camera)
            return {"rgb": rgb, "depth": None, "accumulation": alpha[..., None],
                    "mat_view": viewmat0, "mat_proj": projmat0, "mat_c2w": camera.c2w}
        if self.num_points == 0:               # empty model: background, zero accumulation, depth 1000 (gc_model.py:238)
            with torch.no_grad():
                rgb = torch.clamp(background.to(dev).float().expand(H, W, 3).clone(), max=1.0)
                alpha = torch.zeros((H, W, 1), device=dev)
                depth_im = None if self.training else torch.full((H, W, 1), 1000.0, device=dev)
            self.xys = torch.zeros((0, 2), device=dev)
            self.radii = torch.zeros((0,), device=dev, dtype=torch.int32)
            viewmat0, projmat0 = self._caller_matrices(camera)
            return {"rgb": rgb, "depth": depth_im, "accumulation": alpha,
                    "mat_view": viewmat0, "mat_proj": projmat0, "mat_c2w": camera.c2w}
        with torch.no_grad(), torch.cuda.device(dev):
            bg = background.to(dev).float().contiguous()
            want_depth = not self.training
            post = {}
            res = _render._forward_impl(
                self.means.detach(), self.scales.detach(), self.quats.detach(), self.opacities.detach(),
                self.features_dc.detach(), self.features_rest.detach(), cam, use_sh, n, bg, strict=True,
                with_depth=want_depth, post=post)
            xys, radii, depth_acc = res[0], res[5], res[13]
            self.xys, self.radii = xys, radii
            alpha = post["alpha"][..., None]
            rgb = post["rgb_clamped"]                                 # gc_model.py:222, written by the kernel
            depth_im = None
            if want_depth:                                            # gc_model.py:224-238 in the same pass
                depth_im = depth_acc[..., None]
                depth_im = torch.where(alpha > 0, depth_im / alpha, torch.full_like(depth_im, 1000.0))
        viewmat0, projmat0 = self._caller_matrices(camera)
        return {"rgb": rgb, "depth": depth_im, "accumulation": alpha,
                "mat_view": viewmat0, "mat_proj": projmat0, "mat_c2w": camera.c2w}

    def get_outputs(self, camera: Camera) -> Dict[str, Optional[Tensor]]:
        if self.fused:
            return self._get_outputs_fused(camera)
        background = self._background()
        W, H = int(camera.W), int(camera.H)
        self.last_size = (H, W)
        viewmat, projmat, viewmat0, projmat0, tile_bounds = camera_matrices(camera.c2w, camera.fx, camera.fy, W, H)
        dev = self.device
        viewmat_d = viewmat.to(dev)
        full_proj = (projmat @ viewmat).to(dev)                       # :180
        colors_crop = torch.cat((self.features_dc[:, None, :], self.features_rest), dim=1)   # :172

        self.xys, depths, self.radii, conics, num_tiles_hit, cov3d = project_gaussians(
            self.means,
            torch.exp(self.scales),
            1,
            self.quats / self.quats.norm(dim=-1, keepdim=True),
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
            viewdirs = self.means.detach() - cam_pos                  # :197
            viewdirs = viewdirs / viewdirs.norm(dim=-1, keepdim=True)
            n = min(self.step // self.sh_degree_interval, self.sh_degree)
            rgbs = spherical_harmonics(n, viewdirs, colors_crop)      # :199-200
            rgbs = torch.clamp(rgbs + 0.5, min=0.0)                   # :201
        else:
            rgbs = torch.sigmoid(colors_crop[:, 0, :])                # :203

        rgb, alpha = rasterize_gaussians(
            self.xys, depths, self.radii, conics, num_tiles_hit, rgbs, torch.sigmoid(self.opacities),
            H, W, background=background, return_alpha=True,
        )                                                             # :208-220
        alpha = alpha[..., None]
        rgb = torch.clamp(rgb, max=1.0)                               # :222
        depth_im = None
        if not self.training:                                         # :224-238
            depth_im = rasterize_gaussians(
                self.xys, depths, self.radii, conics, num_tiles_hit, depths[:, None].repeat(1, 3),
                torch.sigmoid(self.opacities), H, W, background=torch.zeros(3, device=dev),
            )[..., 0:1]
            depth_im = torch.where(alpha > 0, depth_im / alpha, torch.full_like(depth_im, 1000.0))
        return {"rgb": rgb, "depth": depth_im, "accumulation": alpha,
                "mat_view": viewmat0, "mat_proj": projmat0, "mat_c2w": camera.c2w}   # :240-241

    @torch.no_grad()
    def get_outputs_for_camera(self, camera: Camera) -> Dict[str, Optional[Tensor]]:
        """gc_model.py:243-255: eval render (RGB + depth + accumulation) of one camera."""
        assert camera is not None, "must provide camera to gaussian model"
        was_training = self.training
        self.training = False
        try:
            return self.get_outputs(camera)
        finally:
            self.training = was_training

    def forward(self, camera: Camera):
        return self.get_outputs(camera)

    @torch.no_grad()
    def render_views(self, cameras, rank: int = 0, world: int = 1, num_streams: int = 2, keep=("rgb", "depth", "accumulation")):
        """The all-view render loop of gc_pipeline.py:222-228 / gc_render.py:196-203, view-sharded
        (rank renders views v with v % world == rank) and software-pipelined over `num_streams` CUDA
        streams: the latency-bound binning of one view overlaps the compositing of another.
        Returns {view index: {name: tensor}} for this rank's views."""
        from .dist import shard_views
        dev = self.device
        mine = shard_views(len(cameras), rank, world)
        # streams are created once and reused: the caching allocator keeps one memory pool per stream, so
        # fresh streams on every call would mean fresh cudaMallocs on every call
        while len(self._streams) < max(1, num_streams):
            self._streams.append(torch.cuda.Stream(device=dev))
        streams = self._streams[:max(1, num_streams)]
        cur = torch.cuda.current_stream(dev)
        for s_ in streams:
            s_.wait_stream(cur)
        results = {}
        for k, v in enumerate(mine):
            with torch.cuda.stream(streams[k % len(streams)]):
                out = self.get_outputs_for_camera(cameras[v])
                results[v] = {name: out[name] for name in keep}
                for t in results[v].values():
                    if t is not None and t.is_cuda:
                        t.record_stream(cur)        # handed to the caller's stream
        for s_ in streams:
            cur.wait_stream(s_)
        return results


__all__ = ["GaussianModel", "camera_matrices", "projection_matrix", "projection_matrix_ogl", "num_sh_bases"]
