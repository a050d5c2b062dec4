#!/bin/bash
# launch list of the default bench command's C2 step (ncu, per-launch durations; cold-cache and serialised)
set -x
CMD="python bench.py --steps 3 --warmup 3 --no-secondary --no-class-api --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 1200 -c 300 --csv --log-file gpurun_out/launches_c2.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "list rc=$?"; tail -2 gpurun_out/ncu_list.log; wc -l gpurun_out/launches_c2.csv
