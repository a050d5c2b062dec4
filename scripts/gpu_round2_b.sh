#!/bin/bash
# round 2: 2-GPU pass -- the NCCL parity test and the sharded c5 leg as the driver launches it
set -x
python -m pytest tests/test_gpu_dp.py -m gpu -x -q > gpurun_out/pytest_dp.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_dp.log; tail -3 gpurun_out/pytest_dp.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_c5_n2.json 2> gpurun_out/bench_c5_n2.err; echo "c5 n2 rc=$?"
python bench.py --gpus 1 --steps 5 --warmup 3 > gpurun_out/bench_n1_on_multigpu_box.json 2> gpurun_out/bench_n1_on_multigpu_box.err; echo "n1 rc=$?"
tail -c 1500 gpurun_out/bench_c5_n2.err
