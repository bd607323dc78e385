#!/bin/bash
# One ncu pass per GPU call (B200_PROFILING.md), always after the same command exited 0 without ncu:
#   tools/ncu_round.sh list   launch list (gpu__time_duration.sum) of the whole program, warm-up included
#   tools/ncu_round.sh full   --set full capture of every kernel of the LAST training step + the eval view
#                             (prof_step.py brackets them with cudaProfilerStart/Stop)
# Outputs under gpurun_out/.
mkdir -p gpurun_out
CMD="python tools/prof_step.py 4"
if [ "${1:-full}" = "list" ]; then
  $CMD > gpurun_out/plain.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
  tail -3 gpurun_out/ncu_list.log
else
  $CMD > gpurun_out/plain.log 2>&1 &&
  ncu --set full --clock-control none --import-source on --profile-from-start off -f -o gpurun_out/prof $CMD > gpurun_out/ncu_full.log 2>&1
  tail -3 gpurun_out/ncu_full.log
fi
ls -la gpurun_out/
