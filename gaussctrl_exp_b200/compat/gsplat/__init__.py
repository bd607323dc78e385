"""`gsplat` import shim (gsplat 0.1.2 layout) over gaussctrl_exp_b200. Put
gaussctrl_exp_b200/compat on PYTHONPATH (or call gaussctrl_exp_b200.install_gsplat_shim())
and /root/reference/gaussctrl/gc_model.py:32-36 imports resolve here."""
from gaussctrl_exp_b200.project_gaussians import project_gaussians
from gaussctrl_exp_b200.rasterize import rasterize_gaussians
from gaussctrl_exp_b200.sh import num_sh_bases, spherical_harmonics
from gaussctrl_exp_b200.utils import (bin_and_sort_gaussians, compute_cumulative_intersects,
                                      get_tile_bin_edges, map_gaussian_to_intersects)

__version__ = "0.1.2+b200"
__all__ = ["project_gaussians", "rasterize_gaussians", "spherical_harmonics", "num_sh_bases",
           "bin_and_sort_gaussians", "compute_cumulative_intersects", "get_tile_bin_edges",
           "map_gaussian_to_intersects"]
