#!/bin/bash
# two GPUs: the multi-GPU tests, then the N = 2 leg of the scaling bench as the driver launches it
set -x
timeout 1200 python -m pytest tests/test_gpu_dp.py tests/test_gpu_wide_class.py -m gpu -x -q > gpurun_out/pytest_dp.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_dp.log
tail -5 gpurun_out/pytest_dp.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_c5_n2.json 2> gpurun_out/bench_c5_n2.err; echo "c5 n2 rc=$?"
tail -c 600 gpurun_out/bench_c5_n2.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_c5_n2.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d.get('dp_check'), d['config'].get('workload'), d.get('exposed_comm_ms'))
PY
