#!/bin/bash
mkdir -p gpurun_out
REPS=20 python tools/binprof2.py > gpurun_out/b_binprof.log 2>&1
tail -2 gpurun_out/b_binprof.log
REPS=2 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none --csv --log-file gpurun_out/b_bin_launches.csv python tools/binprof2.py > gpurun_out/b_ncu.log 2>&1
tail -2 gpurun_out/b_ncu.log
