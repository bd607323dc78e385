"""Diagnostic: tile-local binning vs the LSD path on a few shapes (run under compute-sanitizer when debugging)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from gaussctrl_exp_b200 import utils as U  # noqa: E402
from oracle import oracle as orc  # noqa: E402
from util import projected  # noqa: E402

dev = torch.device("cuda:0")
cases = [(4000, 75, 50, 11, 0.05), (8000, 128, 80, 1, 0.03), (300, 1920, 1080, 2, 0.3), (5, 33, 17, 3, 0.05)]
if len(sys.argv) > 1:
    cases = cases[: int(sys.argv[1])]
for N, W, H, seed, s0 in cases:
    d = projected(orc, N=N, W=W, H=H, seed=seed, s0=s0)
    t = [torch.from_numpy(np.ascontiguousarray(d[k])).to(dev) for k in ("xys", "depths", "radii", "hit")]
    b = orc.bin_and_sort(d["xys"], d["depths"], d["radii"], d["hit"], d["tb"])
    for impl in ("lsd", "tile"):
        ids, bins, I = U.bin_fast(*t, d["tb"], impl=impl)
        torch.cuda.synchronize()
        ok = I == b.I and np.array_equal(ids[:I].cpu().numpy(), b.ids_sorted) and np.array_equal(bins.cpu().numpy(), b.tile_bins)
        print(N, W, H, impl, "I", I, "max tile", int((b.tile_bins[:, 1] - b.tile_bins[:, 0]).max()), "OK" if ok else "MISMATCH", flush=True)
