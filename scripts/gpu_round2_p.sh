#!/bin/bash
set -x
timeout 900 python -m pytest tests/test_gpu_optin_paths.py -m gpu -x -q > gpurun_out/pytest_optin.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_optin.log
tail -4 gpurun_out/pytest_optin.log
for pdl in 0 1; do
POZ_PDL=$pdl timeout 600 python bench.py --steps 200 --warmup 20 --no-class-api > gpurun_out/bench_pdl$pdl.json 2> gpurun_out/bench_pdl$pdl.err; echo "bench rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/bench_pdl$pdl.json').read().strip().splitlines()[-1])
print('pdl $pdl', d['ms_per_step'], d['value'], d['e2e']['ms_per_step'])
"
done
bash scripts/gpu_round2_o.sh
