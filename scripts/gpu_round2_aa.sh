#!/bin/bash
set -x
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_optin_paths.py tests/test_gpu_model_api.py -m gpu -x -q > gpurun_out/pytest_inf.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_inf.log; tail -3 gpurun_out/pytest_inf.log
for p in 1 0; do for b in 8 64 512 1500; do
POZ_PDL_INFER=$p timeout 300 python bench.py --workload c3 --batch $b --steps 300 --warmup 20 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('pdl_infer $p batch $b us/call', round(d['ms_per_step']*1e3,2))
"; done; done
