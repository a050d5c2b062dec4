#!/bin/bash
set -x
timeout 1800 python -m pytest tests/test_gpu_model_api.py tests/test_gpu_ref_suite.py tests/test_gpu_net.py tests/test_gpu_forward_autograd.py tests/test_gpu_wide_class.py -m gpu -x -q > gpurun_out/pytest_api.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_api.log
tail -5 gpurun_out/pytest_api.log
timeout 600 python scripts/micro/api_train_rate.py 2>&1 | tail -5
timeout 600 python bench.py --steps 50 --warmup 10 --no-secondary --no-cpu-baseline > gpurun_out/bench_cls.json 2> gpurun_out/bench_cls.err; python -c "
import json
d=json.loads(open('gpurun_out/bench_cls.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['e2e']['class_api'])
"
