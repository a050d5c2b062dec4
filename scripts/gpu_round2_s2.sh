#!/bin/bash
# ncu --set full of the three launches on the C2 step's critical path
set -x
python scripts/micro/tcs_trace.py > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"tcs_gemm|tail_fused" -s 30 -c 3 -o gpurun_out/prof_c2_final python scripts/micro/tcs_trace.py > gpurun_out/ncu_c2_final.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_c2_final.log
