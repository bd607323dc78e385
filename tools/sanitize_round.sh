#!/bin/bash
# compute-sanitizer on small shapes, ONE tool invocation per GPU call (B200_PROFILING.md), after the same program
# exited 0 plain. Usage: tools/sanitize_round.sh memcheck|racecheck|synccheck|initcheck. Logs under gpurun_out/.
TOOL="${1:-memcheck}"
mkdir -p gpurun_out
timeout 900 python tools/sanitize_target.py > gpurun_out/sanitize_plain.out 2>&1 || { echo "plain run failed"; tail -15 gpurun_out/sanitize_plain.out; exit 1; }
tail -2 gpurun_out/sanitize_plain.out
timeout 3000 compute-sanitizer --tool "$TOOL" --error-exitcode 7 --log-file gpurun_out/sanitizer_${TOOL}.log python tools/sanitize_target.py > gpurun_out/sanitizer_${TOOL}.out 2>&1
echo "under $TOOL rc=$?"
tail -4 gpurun_out/sanitizer_${TOOL}.log; tail -3 gpurun_out/sanitizer_${TOOL}.out
