#!/bin/bash
# One ncu pass per GPU call (B200_PROFILING.md), always after the same command exited 0 without ncu:
#   tools/ncu_round.sh list     launch list (gpu__time_duration.sum) of the whole program, warm-up included
#   tools/ncu_round.sh full     --set full capture of every kernel of the LAST training step + the eval view
#                               (prof_step.py brackets them with cudaProfilerStart/Stop); raw page exported as CSV
#   tools/ncu_round.sh source   --set full + source of the compositing kernels only (small report, kept as .ncu-rep)
# Outputs under gpurun_out/ (at most 64 MiB travel back: a report above 40 MB is dropped after its CSV export).
mkdir -p gpurun_out
CMD="python tools/prof_step.py 4"
MODE="${1:-full}"
$CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
if [ "$MODE" = "list" ]; then
  # the bench command itself (short form), as B200_PROFILING.md asks: its per-launch times are serialised, so it is the
  # kernels' SHARE of the step that is compared with bench.py's live timing
  BCMD="python bench.py --steps 4 --warmup 3 --no-cpu --no-extra-configs --no-stages"
  $BCMD > gpurun_out/bench_plain.json 2> gpurun_out/bench_plain.err &&
  ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none --csv --log-file gpurun_out/launches.csv $BCMD > gpurun_out/ncu_list.log 2>&1
  tail -3 gpurun_out/ncu_list.log
elif [ "$MODE" = "source" ]; then
  ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"${NCU_KERNELS:-raster_(fwd_packed|bwd_direct)}" -f -o gpurun_out/prof_src $CMD > gpurun_out/ncu_src.log 2>&1
  tail -3 gpurun_out/ncu_src.log
  ncu -i gpurun_out/prof_src.ncu-rep --page source --csv > gpurun_out/prof_src_source.csv 2>/dev/null
  ncu -i gpurun_out/prof_src.ncu-rep --page raw --csv > gpurun_out/prof_src_raw.csv 2>/dev/null
else
  ncu --set full --clock-control none --profile-from-start off -f -o gpurun_out/prof $CMD > gpurun_out/ncu_full.log 2>&1
  tail -3 gpurun_out/ncu_full.log
  ncu -i gpurun_out/prof.ncu-rep --page raw --csv > gpurun_out/prof_raw.csv 2>/dev/null
fi
for f in gpurun_out/*.ncu-rep; do
  [ -f "$f" ] && [ $(stat -c %s "$f") -gt 40000000 ] && { echo "dropping $f ($(stat -c %s "$f") bytes; CSV export kept)"; rm -f "$f"; }
done
ls -la gpurun_out/
