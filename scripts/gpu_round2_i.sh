#!/bin/bash
set -x
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_optin_paths.py tests/test_gpu_round2.py -m gpu -x -q > gpurun_out/pytest_tail.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_tail.log
tail -15 gpurun_out/pytest_tail.log
timeout 600 python bench.py --steps 200 --warmup 20 --no-class-api > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"
cat gpurun_out/bench_quick.json | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'], d['e2e']['ms_per_step'], d['roofline']['kernels_us'], d['roofline']['kernel'], d['roofline']['frac'])"
POZ_KERNEL_TRACE=1 timeout 300 python scripts/micro/tcs_trace.py 2> gpurun_out/trace_tail.txt; echo "trace rc=$?"
grep -n "tag 66\|tag=66\| 66 " gpurun_out/trace_tail.txt | head -20
