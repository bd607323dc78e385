#!/bin/bash
# Scaling on one box: tools/scale_round.sh N  (N = 2, 4 or 8 GPUs): the multi-GPU pytest, bench.py at 1, 2, .. N ranks,
# the step anatomy at N ranks (tools/dp_profile.py). Outputs under gpurun_out/.
mkdir -p gpurun_out
N=${1:-4}
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -q -x --timeout=500 > gpurun_out/pytest_mgpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_mgpu.log
tail -3 gpurun_out/pytest_mgpu.log
for n in 1 2 4 8; do
  [ $n -gt $N ] && break
  if [ $n -eq 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 30 --warmup 10 --no-extra-configs --no-stages > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 30 --warmup 10 --no-stages > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err
  fi
  echo "bench $n rc=$?"
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29520 tools/dp_profile.py > gpurun_out/dp_profile_${N}gpu.log 2>&1
grep "rank 0" gpurun_out/dp_profile_${N}gpu.log | head -2
python - <<'PY'
import json
base = None
for n in (1, 2, 4, 8):
    try:
        d = json.loads(open(f"gpurun_out/scale_{n}.json").read().strip().splitlines()[-1])
    except Exception as e:
        continue
    base = base or d["value"]
    o = d.get("config4_finetune_step_with_optimizer") or {}
    print(n, round(d["value"], 1), round(d["ms_per_step"], 4), "eff", round(d["value"] / (n * base), 3), "e2e", round(d["e2e"]["value"], 1),
          "opt", round(o.get("views_per_s", 0), 1), (d.get("grad_exchange_check") or {}).get("max_rel_diff"), d.get("config3_all_view_render", {}).get("views_per_s"))
PY
