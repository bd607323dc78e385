from gaussctrl_exp_b200.utils import (bin_and_sort_gaussians, compute_cumulative_intersects,  # noqa: F401
                                      get_tile_bin_edges, map_gaussian_to_intersects)
