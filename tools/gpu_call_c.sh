#!/bin/bash
mkdir -p gpurun_out
timeout 200 python bench.py --no-cpu > gpurun_out/c_bench.json 2> gpurun_out/c_bench.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/c_bench.json").read().strip().splitlines()[-1])
print(round(d["value"],1), round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"],1), "host", round(d["host_enqueue_ms_per_step"],3), "launches", d["gpu_launches"], "render", round(d["eval_render"]["views_per_s"]))
for k,v in d["stages"].items(): print(" ", k, round(v["us"],1))
PY
tail -3 gpurun_out/c_bench.err
( time timeout 400 python -m pytest tests -x -q -m gpu ) > gpurun_out/c_pytest.log 2>&1
tail -4 gpurun_out/c_pytest.log
