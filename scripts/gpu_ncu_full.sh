#!/bin/bash
# ncu --set full captures of the top kernels (one GPU). Each only after the same command exited 0 without ncu.
mkdir -p gpurun_out
C4="python bench.py --workload c4 --steps 1 --warmup 3 --no-cpu-baseline"
C2="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
timeout 300 $C4 > gpurun_out/plain_c4.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_gemm_kernel -s 11 -c 4 -f -o gpurun_out/prof_c4 $C4 > gpurun_out/ncu_c4.log 2>&1
echo "ncu c4 rc=$?"
timeout 300 $C2 > gpurun_out/plain_c2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_f32_pipelined -s 16 -c 8 -f -o gpurun_out/prof_c2 $C2 > gpurun_out/ncu_c2.log 2>&1
echo "ncu c2 rc=$?"
ls -la gpurun_out/*.ncu-rep
