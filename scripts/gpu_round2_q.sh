#!/bin/bash
set -x
timeout 1500 python -m pytest tests/test_gpu_tail_dropout.py tests/test_gpu_tail_shapes.py tests/test_gpu_wide.py tests/test_gpu_parity.py tests/test_gpu_round2.py tests/test_gpu_optin_paths.py -m gpu -x -q > gpurun_out/pytest_drop.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_drop.log
tail -6 gpurun_out/pytest_drop.log
timeout 600 python bench.py --steps 200 --warmup 20 --no-class-api > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/bench_quick.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['e2e']['ms_per_step'], d.get('with_dropout_0.2'))
"
