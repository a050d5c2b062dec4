#!/bin/bash
# Data-parallel scaling of C5 at N GPUs (N = $1); also the default (C2 replicas) line at N.
N=${1:-2}
mkdir -p gpurun_out
PORT=29517
if [ "$N" = "1" ]; then
  timeout 900 python bench.py --gpus 1 --workload c5 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/c5_n1.json 2> gpurun_out/c5_n1.err
else
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT \
     bench.py --gpus $N --workload c5 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/c5_n$N.json 2> gpurun_out/c5_n$N.err
fi
echo "c5 N=$N rc=$?"; tail -3 gpurun_out/c5_n$N.err; grep -o '"value": [0-9.]*, "unit": "samples/s", "n_gpus": [0-9]*.\{0,60\}' gpurun_out/c5_n$N.json
if [ "$N" != "1" ]; then
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((PORT+1)) \
     bench.py --gpus $N --steps 100 --warmup 10 --no-cpu-baseline > gpurun_out/c2_n$N.json 2> gpurun_out/c2_n$N.err
  echo "c2 N=$N rc=$?"; tail -2 gpurun_out/c2_n$N.err; cut -c1-200 gpurun_out/c2_n$N.json
fi
