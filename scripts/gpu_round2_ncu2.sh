#!/bin/bash
set -x
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
CMD="python bench.py --steps 3 --warmup 3 --no-secondary --no-class-api --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 1500 -c 400 --csv --log-file gpurun_out/launches_c2.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "list rc=$?"
python scripts/micro/tcs_trace.py > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:tcs_gemm -s 20 -c 2 -o gpurun_out/prof_tcs_v2 python scripts/micro/tcs_trace.py > gpurun_out/ncu_tcs.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_tcs.log
