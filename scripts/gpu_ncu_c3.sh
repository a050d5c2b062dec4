#!/bin/bash
# ncu captures of the large-batch inference pass (C3, 262,144 rows): launch list, then --set full of the two
# launches of a pass (unit 0: fp32 rows -> bf16 planes in tensor memory -> tcgen05; unit 1 + head: the same kernel
# with the last unit in its epilogue).  Each only after the plain run exited 0.
mkdir -p gpurun_out
C3="python bench.py --workload c3 --steps 3 --warmup 3 --no-cpu-baseline"
timeout 300 $C3 > gpurun_out/c3_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/c3_plain.log; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv \
    --log-file gpurun_out/c3_launches.csv $C3 > gpurun_out/c3_ncu_list.log 2>&1
echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc3_fwd_x32t -s 4 -c 2 -f \
    -o gpurun_out/prof_c3_x32 $C3 > gpurun_out/c3_ncu_full.log 2>&1
echo "ncu full rc=$?"
ncu -i gpurun_out/prof_c3_x32.ncu-rep --page raw --csv > gpurun_out/c3_x32_raw.csv 2>/dev/null
ncu -i gpurun_out/prof_c3_x32.ncu-rep --page source --csv > gpurun_out/c3_x32_source.csv 2>/dev/null
ls -la gpurun_out/prof_c3_x32.ncu-rep gpurun_out/c3_x32_raw.csv gpurun_out/c3_launches.csv gpurun_out/c3_x32_source.csv
