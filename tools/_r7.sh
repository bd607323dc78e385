mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_parity_gpu.py tests/test_model_gpu.py tests/test_edge_cases_gpu.py -m gpu -q -x --timeout=200 > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
tail -4 gpurun_out/pytest.log
timeout 300 python tools/binprof.py > gpurun_out/binprof.log 2>&1
REPS=2 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none --csv --log-file gpurun_out/binprof_launches.csv python tools/binprof.py > gpurun_out/binprof_ncu.log 2>&1
grep impl gpurun_out/binprof.log
