from gaussctrl_exp_b200.rasterize import rasterize_gaussians  # noqa: F401
