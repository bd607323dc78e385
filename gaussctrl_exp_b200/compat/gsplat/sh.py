from gaussctrl_exp_b200.sh import num_sh_bases, spherical_harmonics  # noqa: F401
