#!/bin/bash
set -x
python scripts/micro/tcs_trace.py > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:tail_fused -s 10 -c 1 -o gpurun_out/prof_tail python scripts/micro/tcs_trace.py > gpurun_out/ncu_tail.log 2>&1
echo "ncu rc=$?"; tail -5 gpurun_out/ncu_tail.log
