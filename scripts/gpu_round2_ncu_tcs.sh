#!/bin/bash
set -x
python scripts/micro/tcs_trace.py > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:tcs_gemm -s 20 -c 2 -o gpurun_out/prof_tcs python scripts/micro/tcs_trace.py > gpurun_out/ncu_tcs.log 2>&1
echo "ncu rc=$?"; tail -5 gpurun_out/ncu_tcs.log
