"""B200-native 3D Gaussian splatting render path behind the gsplat-0.1.2 operator API.

Drop-in for the calls /root/reference/gaussctrl/gc_model.py makes at :174, :200, :208, :225
(imports at :32-36). Hand-written sm_100a CUDA kernels behind a C ABI (include/gs_b200.h);
no CPU fallback — every operator raises if libgs_b200.so is missing.
"""
from .project_gaussians import project_gaussians
from .rasterize import rasterize_gaussians
from .sh import num_sh_bases, spherical_harmonics
from .utils import (bin_and_sort_gaussians, compute_cumulative_intersects, get_tile_bin_edges,
                    map_gaussian_to_intersects)

__version__ = "0.1.0"


def install_gsplat_shim() -> None:
    """Make `import gsplat` (gsplat.sh / gsplat.project_gaussians / gsplat.rasterize / ...)
    resolve to this package, so gc_model.py:32-36 imports work unmodified."""
    import os
    import sys
    compat = os.path.join(os.path.dirname(os.path.abspath(__file__)), "compat")
    if compat not in sys.path:
        sys.path.insert(0, compat)


__all__ = ["project_gaussians", "rasterize_gaussians", "spherical_harmonics", "num_sh_bases",
           "compute_cumulative_intersects", "map_gaussian_to_intersects", "get_tile_bin_edges",
           "bin_and_sort_gaussians", "install_gsplat_shim"]
