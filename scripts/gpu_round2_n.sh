#!/bin/bash
set -x
timeout 900 python -m pytest tests/test_gpu_infer_f16.py -m gpu -x -q > gpurun_out/pytest_f16.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_f16.log
tail -8 gpurun_out/pytest_f16.log
timeout 600 python bench.py --workload c3 --steps 20 --warmup 3 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench rc=$?"; tail -c 300 gpurun_out/bench_c3.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_c3.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['roofline'].get('kernels_ms'), d['roofline'].get('frac'))
PY
timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "config3" > gpurun_out/pytest_c3.log 2>&1; tail -3 gpurun_out/pytest_c3.log
