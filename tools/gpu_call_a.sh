#!/bin/bash
# one GPU call: binning parity, bench A/B of the depth-sort passes and the side-stream zero fill, then the full GPU suite
mkdir -p gpurun_out
( time timeout 200 python -m pytest tests/test_parity_gpu.py -x -q -k "binning" ) > gpurun_out/a_bin.log 2>&1
tail -3 gpurun_out/a_bin.log
timeout 150 python bench.py --no-cpu > gpurun_out/a_bench3.json 2> gpurun_out/a_bench3.err
GS_DEPTH_SORT_PASSES=4 timeout 120 python bench.py --no-cpu > gpurun_out/a_bench4.json 2> gpurun_out/a_bench4.err
GS_ZERO_SIDE=1 timeout 120 python bench.py --no-cpu --no-stages > gpurun_out/a_benchz.json 2> gpurun_out/a_benchz.err
python - <<'PY'
import json
for f in ("a_bench3","a_bench4","a_benchz"):
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, round(d["value"],1), round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"],1), d.get("step_time_distribution"), "bin_fast", d.get("stages",{}).get("bin_fast",{}).get("us"))
    except Exception as e:
        print(f, "FAILED", e)
PY
( time timeout 400 python -m pytest tests -x -q -m gpu ) > gpurun_out/a_pytest.log 2>&1
tail -5 gpurun_out/a_pytest.log
