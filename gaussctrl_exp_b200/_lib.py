"""ctypes binding of the C ABI declared in include/gs_b200.h.

There is NO fallback: if libgs_b200.so is missing or a call fails, an exception is raised.
The library is built in-tree (gaussctrl_exp_b200/libgs_b200.so) by `build()` /
`__graft_entry__.build()` with nvcc for sm_100a.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.environ.get("GS_B200_LIB") or os.path.join(_HERE, "libgs_b200.so")   # GS_B200_LIB: A/B builds
CSRC = os.path.join(_HERE, "csrc")

_lib = None

_vp, _i, _f, _i64, _sz, _d = C.c_void_p, C.c_int, C.c_float, C.c_int64, C.c_size_t, C.c_double

# name -> (restype, argtypes); mirrors include/gs_b200.h line by line
SIGNATURES = {
    "gs_version": (_i, []),
    "gs_built_arch": (_i, []),
    "gs_launch_count": (_i64, []),
    "gs_project_fwd": (_i, [_i, _vp, _vp, _f, _vp, _vp, _vp, _f, _f, _f, _f, _i, _i, _i, _i, _f,
                            _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gs_project_bwd": (_i, [_i, _vp, _vp, _f, _vp, _vp, _vp, _f, _f, _i, _i, _vp, _vp, _vp, _vp,
                            _vp, _vp, _vp, _vp]),
    "gs_preprocess_fwd": (_i, [_i, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gs_preprocess_fwd_part": (_i, [_i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp,
                                    _vp, _vp]),
    "gs_preprocess_bwd": (_i, [_i, _vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp,
                               _vp, _vp, _vp, _vp, _vp]),
    "gs_view_block_floats": (_i, [_i]),
    "gs_finish_view_block": (_i, [_i, _i, _vp, _vp, _vp, _vp]),
    "gs_preprocess_bwd_multi": (_i, [_i, _i, _vp, _i64, _f, _vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp,
                                     _vp, _vp]),
    "gs_preprocess_bwd_multi_peer": (_i, [_i, _i, _vp, _i64, _f, _vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp,
                                          _vp, _vp]),
    "gs_view_xbuf_floats": (_i, [_i]),
    "gs_view_pack_scratch_bytes": (_sz, [_i]),
    "gs_pack_view_block": (_i, [_i, _vp, _vp, _vp, _sz, _vp]),
    "gs_pack_view_block_push": (_i, [_i, _vp, _vp, _i, _i, _vp, _i64, _vp, _sz, _vp]),
    "gs_adam_slice": (_i, [_i64, _vp, _vp, _vp, _vp, _f, _f, _f, _f, _f, _i, _i, _vp, _i64, _vp]),
    "gs_view_slice_size": (_i, [_i, _i]),
    "gs_pack_view_block_push_slices": (_i, [_i, _vp, _vp, _i, _i, _vp, _i64, _vp, _sz, _vp]),
    "gs_leaf_bwd_slice": (_i, [_i, _i, _i, _i, _vp, _i64, _f, _vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gs_preprocess_bwd_multi_peer_sparse": (_i, [_i, _i, _vp, _i64, _f, _vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp,
                                                 _vp, _vp, _vp]),
    "gs_sh_fwd": (_i, [_i, _i, _i, _vp, _vp, _vp, _vp]),
    "gs_sh_bwd": (_i, [_i, _i, _i, _vp, _vp, _vp, _vp]),
    "gs_scan_scratch_bytes": (_sz, [_i]),
    "gs_bin_scan": (_i, [_i, _vp, _vp, _vp, _vp, _sz, _vp]),
    "gs_bin_scan_static": (_i, [_i, _vp, _vp, _vp, _vp]),
    "gs_bin_emit": (_i, [_i, _vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp]),
    "gs_sort_scratch_bytes": (_sz, [_i64]),
    "gs_bin_sort": (_i, [_i64, _i, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "gs_bin_edges": (_i, [_i64, _i, _vp, _vp, _vp]),
    "gs_bin_fast_scratch_bytes": (_sz, [_i, _i64, _i]),
    "gs_bin_fast": (_i, [_i, _i64, _vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp, _sz, _vp]),
    "gs_set_depth_sort_passes": (_i, [_i]),
    "gs_set_radix_cfg": (_i, [_i, _i]),
    "gs_set_bin_impl": (_i, [_i]),
    "gs_set_raster_fwd_unroll": (_i, [_i]),
    "gs_set_raster_bwd_direct": (_i, [_i]),
    "gs_raster_fwd": (_i, [_i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gs_raster_fwd_rgbd": (_i, [_i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gs_tile_order": (_i, [_i, _vp, _i, _vp, _vp]),
    "gs_raster_fwd_ordered": (_i, [_i, _i, _i, _i] + [_vp] * 15),
    "gs_raster_bwd_ordered": (_i, [_i, _i, _i, _i] + [_vp] * 17),
    "gs_raster_fwd_ordered_post": (_i, [_i, _i, _i, _i] + [_vp] * 17),
    "gs_raster_bwd_ordered_post": (_i, [_i, _i, _i, _i] + [_vp] * 18),
    "gs_ray_directions": (_i, [_i, _i, _d, _d, _vp, _vp]),
    "gs_depth_to_world": (_i, [_i, _i, _vp, _vp, _d, _d, _vp, _vp]),
    "gs_depth_to_disparity": (_i, [_i64, _vp, _vp, _vp, _vp]),
    "gs_noise_mask_scratch_bytes": (_sz, [_i, _i]),
    "gs_noise_mask": (_i, [_i, _vp, _i, _vp, _i, _vp, _vp, _vp, _i, _i, _vp, _f, _vp, _sz, _vp, _vp]),
    "gs_raster_bwd": (_i, [_i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp,
                           _vp, _vp, _vp, _vp, _vp]),
    "gs_pack_splats": (_i, [_i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gs_preprocess_fwd_packed": (_i, [_i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gs_raster_fwd_packed": (_i, [_i, _i, _i, _i, _i] + [_vp] * 13),
    "gs_raster_bwd_packed": (_i, [_i, _i, _i, _i] + [_vp] * 15),
    "gs_view_layout": (_i, [_i, _i, _i64, _i, _vp]),
    "gs_view_scratch_bytes": (_sz, [_i, _i, _i64, _i]),
    "gs_view_fwd": (_i, [_i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _i64, _vp, _sz, _vp, _sz,
                         _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gs_view_bwd": (_i, [_i, _vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _sz, _i64, _vp, _vp, _vp, _vp, _vp, _vp,
                         _vp, _vp]),
    "gs_view_bwd_det_scratch_bytes": (_sz, [_i, _i64]),
    "gs_view_bwd_det": (_i, [_i, _vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _sz, _i64, _vp, _vp, _vp, _vp, _vp, _vp,
                             _vp, _vp, _sz, _vp]),
}



class GsCamera(C.Structure):
    """Mirror of `struct GsCamera` in include/gs_b200.h (host memory, passed by pointer)."""
    _fields_ = [("viewmat", C.c_float * 12), ("projmat", C.c_float * 16), ("cam_pos", C.c_float * 3),
                ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float),
                ("glob_scale", C.c_float), ("clip", C.c_float),
                ("H", C.c_int32), ("W", C.c_int32), ("tiles_x", C.c_int32), ("tiles_y", C.c_int32)]


class GsViewLayout(C.Structure):
    """Mirror of `struct GsViewLayout`: byte offsets of the arrays inside a view workspace."""
    _fields_ = [(n, C.c_int64) for n in ("rec", "depths", "radii", "tiles_hit", "ids_sorted", "tile_bins", "tile_work",
                                         "order_fwd", "order_bwd", "total", "bytes")]


_ERRORS = {-1: "GS_ERR_BAD_ARG", -2: "GS_ERR_SCRATCH_SMALL", -3: "GS_ERR_UNSUPPORTED", -4: "GS_ERR_MISALIGNED"}


class GsError(RuntimeError):
    pass


def build(verbose: bool = False) -> str:
    """Compile every CUDA source for sm_100a into the in-tree shared library."""
    cmd = ["make", "-C", CSRC, "-j8"]
    if not verbose:
        cmd.insert(1, "-s")
    subprocess.check_call(cmd)
    if not os.path.exists(SO_PATH):
        raise GsError(f"build did not produce {SO_PATH}")
    return SO_PATH


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise GsError(
                f"{SO_PATH} is missing: the CUDA extension was not built (run __graft_entry__.build()). "
                "There is no CPU or PyTorch fallback for this path.")
        L = C.CDLL(SO_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc: int, what: str):
    if rc == 0:
        return
    if rc < 0:
        raise GsError(f"{what}: {_ERRORS.get(rc, rc)}")
    raise GsError(f"{what}: CUDA error {rc}")


COUNT_SATURATED = 0x7FFFFFFF


def check_count(I: int, what: str) -> int:
    """Tile-intersection totals are int32 on the device (as gsplat's int32 cumsum is). The scans saturate at INT32_MAX
    instead of wrapping, and nothing is written for such a view: report it instead of treating it as a count."""
    if I < 0 or I >= COUNT_SATURATED:
        raise GsError(f"{what}: the view has more than 2^31 - 2 tile intersections (int32 count saturated): "
                      "GS_ERR_UNSUPPORTED")
    return I


def ptr(t):
    """Device pointer of a (contiguous) torch tensor, or NULL for None."""
    return None if t is None else t.data_ptr()


def stream_ptr():
    import torch
    return torch.cuda.current_stream().cuda_stream


def launch_count() -> int:
    return int(lib().gs_launch_count())
