#!/bin/bash
set -x
timeout 1500 python -m pytest tests/test_gpu_tail_shapes.py tests/test_gpu_tail_dropout.py tests/test_gpu_parity.py tests/test_gpu_optin_paths.py tests/test_gpu_model_api.py tests/test_gpu_round2.py -m gpu -x -q > gpurun_out/pytest_ab.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_ab.log; tail -3 gpurun_out/pytest_ab.log
timeout 600 python bench.py --steps 200 --warmup 20 --no-secondary --no-cpu-baseline > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/bench_quick.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['e2e']['ms_per_step'], d['e2e']['class_api'], d.get('with_dropout_0.2')['ms_per_step'])
"
POZ_KERNEL_TRACE=1 timeout 300 python scripts/micro/tcs_trace.py 2>&1 | grep "tag 200\|tag 201" | tail -4
