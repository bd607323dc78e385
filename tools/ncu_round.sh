#!/bin/bash
# One ncu call (B200_PROFILING.md): plain run first, then the launch list and one --set full capture of the
# compositing kernels of a warm step. Outputs under gpurun_out/.
mkdir -p gpurun_out
CMD="python tools/prof_step.py 4"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"${NCU_KERNELS:-raster_(fwd|bwd)_packed}" -s ${NCU_SKIP:-4} -c ${NCU_COUNT:-2} -f -o gpurun_out/prof $CMD > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_list.log gpurun_out/ncu_full.log; ls -la gpurun_out/
