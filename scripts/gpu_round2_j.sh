#!/bin/bash
set -x
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_optin_paths.py -m gpu -x -q > gpurun_out/pytest_tail.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_tail.log
tail -5 gpurun_out/pytest_tail.log
timeout 600 python bench.py --steps 200 --warmup 20 --no-class-api > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"
cat gpurun_out/bench_quick.json | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'], d['e2e']['ms_per_step'], d['roofline']['kernels_us'], d['roofline']['kernel'], d['roofline']['frac'])"
POZ_KERNEL_TRACE=1 timeout 300 python scripts/micro/tcs_trace.py 2> gpurun_out/trace_tail.txt; echo "trace rc=$?"
grep -n "tag 66" gpurun_out/trace_tail.txt | head -4
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tail_fused -s 10 -c 1 -o gpurun_out/prof_tail python scripts/micro/tcs_trace.py > gpurun_out/ncu_tail.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_tail.log
