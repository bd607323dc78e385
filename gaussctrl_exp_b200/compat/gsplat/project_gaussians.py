from gaussctrl_exp_b200.project_gaussians import project_gaussians  # noqa: F401
