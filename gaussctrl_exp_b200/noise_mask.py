"""Multi-view noise mask, headless.

Host-side mirror of the reference's `MultiVeiwNoiseRenderer`
(/root/reference/gaussctrl/ad_render.py:207-333, shaders :384-479; helpers :164-205 and
/root/reference/gaussctrl/ad_noise.py:5-24, :53-66): a cube of noise voxels is thresholded into a point set, every
point is drawn as a small instanced sphere, and a fragment survives only inside a thin shell just in front of the
rendered 3DGS depth — so the same world-space noise shows up consistently in every view.

The reference does this with pygame + GLUT + OpenGL shaders and `glReadPixels`; no graphics stack exists here
("unavailable on this box"), so the same arithmetic runs as a CUDA point-splat (`csrc/noise_mask.cu`,
`gs_noise_mask` in include/gs_b200.h) into a device buffer. Method names and argument meaning follow the
reference class so a caller (`gc_pipeline.py:242-250`, commented out upstream) can switch over; there is no CPU
fallback.
"""
from __future__ import annotations

import os
from typing import Optional, Tuple, Union

import numpy as np
import torch

from . import _lib

# the fragment shader ignores its `u_depth_threshold` uniform and hard-codes this sum (ad_render.py:423)
FRAG_DEPTH_THRESHOLD = float(np.float32(0.016) + np.float32(0.0001))


def sphere_mesh(radius: float, stacks: int, slices: int) -> Tuple[np.ndarray, np.ndarray]:
    """UV sphere in the vertex / index order of ad_noise.py:5-24 (`create_sphere`): vertices (V,3) float32 with
    V = (stacks+1)(slices+1), rows of constant polar angle from +y to -y; two triangles per quad,
    (first, second, first+1) and (second, second+1, first+1). Returned as (V,3) / (F,3) arrays (the reference
    keeps them flat)."""
    phi = np.arange(stacks + 1, dtype=np.float64) / stacks * np.pi
    theta = np.arange(slices + 1, dtype=np.float64) / slices * 2 * np.pi
    ph, th = np.meshgrid(phi, theta, indexing="ij")
    verts = np.stack([radius * np.sin(ph) * np.cos(th), radius * np.cos(ph), radius * np.sin(ph) * np.sin(th)], -1)
    st, sl = np.meshgrid(np.arange(stacks), np.arange(slices), indexing="ij")
    first = (st * (slices + 1) + sl).reshape(-1)
    second = first + slices + 1
    tris = np.stack([first, second, first + 1, second, second + 1, first + 1], -1).reshape(-1, 3)
    return verts.reshape(-1, 3).astype(np.float32), tris.astype(np.int32)


def uniform_points_in_cube(cube_size: float = 2.0, resolution: int = 50) -> np.ndarray:
    """Voxel centres of ad_noise.py:53-66 (`generate_uniform_points_in_cube`): np.meshgrid's default 'xy' indexing,
    i.e. point index = (iy * R + ix) * R + iz; (R^3, 3) float64."""
    a = np.linspace(-cube_size / 2, cube_size / 2, resolution)
    xx, yy, zz = np.meshgrid(a, a, a)
    return np.stack([xx.ravel(), yy.ravel(), zz.ravel()], axis=1)


def extract_camera_parameters(c2w) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """ad_render.py:164-175: (position, up = +y column, forward = -z column) of a camera-to-world matrix."""
    c2w = np.array(c2w)
    return c2w[:3, 3], c2w[:3, 1], -c2w[:3, 2]


def compute_view_matrix_ogl(cam_pos, up, forward) -> np.ndarray:
    """ad_render.py:177-205. Note right = up x forward (not forward x up): the x axis is mirrored with respect to a
    conventional look-at matrix, which the reference undoes with np.fliplr after read-back."""
    forward = forward / np.linalg.norm(forward)
    up = up / np.linalg.norm(up)
    right = np.cross(up, forward)
    right = right / np.linalg.norm(right)
    up = np.cross(forward, right)
    m = np.zeros((4, 4))
    m[0, :3], m[1, :3], m[2, :3] = right, up, -forward
    m[0, 3], m[1, 3], m[2, 3] = -np.dot(right, cam_pos), -np.dot(up, cam_pos), np.dot(forward, cam_pos)
    m[3, 3] = 1
    return m


class MultiViewNoiseRenderer:
    """Drop-in for `MultiVeiwNoiseRenderer` (ad_render.py:207): same constructor arguments, `setup_noise`,
    `render_noise_SS`, `c2w_to_viewmat`, `binarize_mask`; renders on `device` instead of a hidden GL window."""

    def __init__(self, width: int, height: int, z_near: float = 0.001, z_far: float = 1000.0,
                 device: Union[str, torch.device] = "cuda"):
        self.width, self.height = int(width), int(height)
        self.z_near, self.z_far = z_near, z_far
        self.cube_size = 2.0                 # ad_render.py:215-227
        self.resolution = 100
        self.noise_threshold = 0.8
        self.noise_unit_size = 0.015
        self.frag_depth_threshold = 0.016    # uniform the shader never reads; FRAG_DEPTH_THRESHOLD is what applies
        self.device = torch.device(device)
        self.pts: Optional[np.ndarray] = None
        self.noise_np: Optional[np.ndarray] = None
        self._pts_dev = self._verts_dev = self._tris_dev = None
        self.vertices = self.indices = None

    def setup_noise(self, noise: Union[str, np.ndarray]):
        """ad_render.py:239-250 (+ the mesh / buffer creation of :258-289). `noise` is the (R,R,R) volume or the
        path of its .npy; generating it needs the third-party `noise.pnoise3` (ad_noise.py:68-105), which is
        absent here, so a missing file is an error rather than a silently different volume."""
        R = self.resolution
        if isinstance(noise, str):
            if not os.path.exists(noise):
                raise FileNotFoundError(f"noise volume '{noise}' not found (Perlin generation needs the `noise` package)")
            noise = np.load(noise)
        noise = np.asarray(noise, dtype=np.float32)
        if noise.shape != (R, R, R):
            raise AssertionError("noise_np shape not match")            # ad_noise.py:121
        self.noise_np = noise
        keep = np.where(noise.reshape(-1) > self.noise_threshold)[0]
        self.pts = uniform_points_in_cube(self.cube_size, R)[keep].astype(np.float32)
        self.vertices, self.indices = sphere_mesh(self.noise_unit_size, 8, 8)
        if self.device.type != "cuda":
            raise ValueError("MultiViewNoiseRenderer needs a CUDA device (no CPU fallback)")
        self._pts_dev = torch.from_numpy(self.pts).to(self.device).contiguous()
        self._verts_dev = torch.from_numpy(self.vertices).to(self.device).contiguous()
        self._tris_dev = torch.from_numpy(self.indices).to(self.device).contiguous()

    def c2w_to_viewmat(self, mat_c2w) -> np.ndarray:
        """ad_render.py:346-350 (without the print)."""
        pos, up, fw = extract_camera_parameters(mat_c2w)
        return compute_view_matrix_ogl(pos, up, fw)

    def render_noise_SS(self, depth_np, mat_view, mat_proj, mat_c2w=None, return_tensor: bool = False):
        """ad_render.py:273-333. depth_np: (H,W) or (H,W,1) distance image of the SAME camera and size as this
        renderer (numpy, or a CUDA tensor to stay on the device); mat_proj: the OpenGL projection `get_outputs`
        returns as "mat_proj"; mat_view is ignored exactly like upstream (the view matrix is rebuilt from mat_c2w).
        Returns the (H,W,4) uint8 RGBA image (numpy, or a CUDA tensor with return_tensor=True)."""
        if self._pts_dev is None:
            raise RuntimeError("call setup_noise() first")
        if mat_c2w is None:
            raise ValueError("mat_c2w is required (the view matrix is derived from it)")
        to_np = lambda m: m.detach().cpu().numpy() if isinstance(m, torch.Tensor) else np.asarray(m)
        view = np.ascontiguousarray(self.c2w_to_viewmat(to_np(mat_c2w).astype(np.float64)), dtype=np.float32)
        proj = np.ascontiguousarray(to_np(mat_proj), dtype=np.float32)
        if proj.shape != (4, 4):
            raise ValueError("mat_proj must be 4x4")
        H, W = self.height, self.width
        depth = depth_np if isinstance(depth_np, torch.Tensor) else torch.from_numpy(np.asarray(depth_np, np.float32))
        depth = depth.to(self.device, torch.float32).reshape(depth.shape[0], depth.shape[1]).contiguous()
        if tuple(depth.shape) != (H, W):
            raise ValueError(f"depth image {tuple(depth.shape)} does not match the renderer size {(H, W)}")
        out = rasterize_noise_mask(self._pts_dev, self._verts_dev, self._tris_dev, view, proj, depth, FRAG_DEPTH_THRESHOLD)
        return out if return_tensor else out.cpu().numpy()

    @staticmethod
    def binarize_mask(rgb_mask) -> np.ndarray:
        """ad_render.py:352-360 (which reads an undefined name upstream): 1 where any channel is non-zero."""
        rgb_mask = np.asarray(rgb_mask)
        return np.any(rgb_mask != 0, axis=-1).astype(np.uint8)[:, :, None]


def rasterize_noise_mask(points: torch.Tensor, verts: torch.Tensor, tris: torch.Tensor, view: np.ndarray,
                         proj: np.ndarray, depth: torch.Tensor, threshold: float = FRAG_DEPTH_THRESHOLD) -> torch.Tensor:
    """gs_noise_mask: points (P,3) f32, verts (V,3) f32, tris (F,3) i32, depth (H,W) f32 on one CUDA device;
    view / proj row-major 4x4 float32 on the host -> (H,W,4) uint8 CUDA tensor."""
    for t, name in ((points, "points"), (verts, "verts"), (tris, "tris"), (depth, "depth")):
        if not t.is_cuda:
            raise ValueError(f"{name} must be a CUDA tensor (no CPU fallback)")
    if points.ndim != 2 or points.shape[1] != 3 or verts.ndim != 2 or verts.shape[1] != 3 or tris.ndim != 2 or tris.shape[1] != 3:
        raise ValueError("points (P,3), verts (V,3), tris (F,3) expected")
    dev = depth.device
    points, verts, depth = points.float().contiguous(), verts.float().contiguous(), depth.float().contiguous()
    tris = tris.to(torch.int32).contiguous()
    view = np.ascontiguousarray(view, dtype=np.float32).reshape(16)
    proj = np.ascontiguousarray(proj, dtype=np.float32).reshape(16)
    H, W = depth.shape
    L = _lib.lib()
    nbytes = L.gs_noise_mask_scratch_bytes(H, W)
    scratch = torch.empty(nbytes // 8, device=dev, dtype=torch.int64)
    out = torch.empty((H, W, 4), device=dev, dtype=torch.uint8)
    with torch.cuda.device(dev):
        rc = L.gs_noise_mask(points.shape[0], _lib.ptr(points), verts.shape[0], _lib.ptr(verts), tris.shape[0],
                             _lib.ptr(tris), view.ctypes.data, proj.ctypes.data, H, W, _lib.ptr(depth), float(threshold),
                             _lib.ptr(scratch), nbytes, _lib.ptr(out), _lib.stream_ptr())
    _lib.check(rc, "gs_noise_mask")
    return out


__all__ = ["MultiViewNoiseRenderer", "rasterize_noise_mask", "sphere_mesh", "uniform_points_in_cube",
           "extract_camera_parameters", "compute_view_matrix_ogl", "FRAG_DEPTH_THRESHOLD"]
