#!/bin/bash
# BASELINE.json configs[2]: inference sweep on the 784-256-128-10 model, batch 1 .. 1M
mkdir -p gpurun_out; : > gpurun_out/c3_sweep.jsonl
for b in 1 8 64 512 4096 32768 262144 1048576; do
  steps=200; [ $b -ge 32768 ] && steps=30; [ $b -ge 1048576 ] && steps=10
  timeout 300 python bench.py --workload c3 --batch $b --steps $steps --warmup 5 --no-cpu-baseline 2>>gpurun_out/c3.err >> gpurun_out/c3_sweep.jsonl
done
timeout 300 python bench.py --workload c3 --batch 65536 --steps 30 --warmup 5 2>>gpurun_out/c3.err >> gpurun_out/c3_sweep.jsonl
wc -l gpurun_out/c3_sweep.jsonl; tail -2 gpurun_out/c3.err
