#!/bin/bash
set -x
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "host_buffer" > gpurun_out/pytest_host.log 2>&1; tail -2 gpurun_out/pytest_host.log
for i in 1 2; do
timeout 600 python bench.py --steps 200 --warmup 20 --no-class-api --no-secondary --no-cpu-baseline > gpurun_out/bench_e2e.json 2> gpurun_out/bench_e2e.err; echo "bench rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/bench_e2e.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['e2e']['ms_per_step'], d['e2e']['runs_ms_per_step'], d['e2e']['h2d_copy_us_per_step'])
"
done
