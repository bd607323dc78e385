"""One process for compute-sanitizer: smoke() and then the small-shape binning / compositing / fused-model / edge-case
GPU tests (both binning pipelines, packed and un-packed compositing kernels, the deterministic backward, empty and ragged
inputs). tools/sanitize_round.sh runs it under one sanitizer tool per GPU call."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pytest  # noqa: E402

import __graft_entry__ as ge  # noqa: E402

ge.smoke()
SMALL = ("binning_bit_exact and (8000 or 300-1920 or 5-33 or 4097)",
         "raster_fwd_bwd and (3000 or 2000 or 1500 or 800)", "raster_final_idx", "zero_intersections", "argument_errors",
         "fused_preprocess and (5003 or 999 or 130)", "empty", "ragged", "uint8", "deterministic_backward",
         "test_project_fwd and (5000 or 777)", "test_project_bwd", "test_sh")
expr = " or ".join(f"({e})" for e in SMALL)
rc = pytest.main([os.path.join(ROOT, "tests", "test_parity_gpu.py"), os.path.join(ROOT, "tests", "test_edge_cases_gpu.py"),
                  os.path.join(ROOT, "tests", "test_model_gpu.py"), "-m", "gpu", "-q", "-x", "--timeout=3000", "-k", expr,
                  "-p", "no:cacheprovider"])
print("sanitize_target pytest rc", int(rc))
sys.exit(int(rc))
