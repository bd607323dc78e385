mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/smi2.txt
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -q -x --timeout=500 -s > gpurun_out/pytest_mgpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_mgpu.log
grep -h "slice\|sharded\|PEER_CHECK\|passed\|failed\|Error\|rc=" gpurun_out/pytest_mgpu.log | tail -20
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 30 --warmup 10 > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err; echo "bench2 rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 30 --warmup 10 --exchange push --no-stages > gpurun_out/bench_2gpu_push.json 2> gpurun_out/bench_2gpu_push.err; echo "bench2push rc=$?"
tail -3 gpurun_out/bench_2gpu.err
python - <<'PY'
import json
for f in ("bench_2gpu.json", "bench_2gpu_push.json"):
    try:
        d = json.loads(open("gpurun_out/" + f).read().strip().splitlines()[-1])
        print(f, d["value"], d["ms_per_step"], d["e2e"]["value"], d.get("grad_exchange_check"), d.get("config4_finetune_step_with_optimizer"), d.get("config4_finetune_step_nccl_allreduce"))
    except Exception as e:
        print(f, "unreadable", e)
PY
