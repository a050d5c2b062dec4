#!/bin/bash
set -x
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_model_api.py tests/test_gpu_round2.py -m gpu -x -q -k "not sweep and not wide" > gpurun_out/pytest_tcs.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_tcs.log
tail -30 gpurun_out/pytest_tcs.log
timeout 300 python bench.py --steps 200 --warmup 20 --no-secondary --no-class-api --no-cpu-baseline > gpurun_out/bench_tcs.json 2> gpurun_out/bench_tcs.err; echo "bench rc=$?"
POZ_TCS_FUSE_UPDATE=0 timeout 300 python bench.py --steps 200 --warmup 20 --no-secondary --no-class-api --no-cpu-baseline > gpurun_out/bench_tcs_nofuse.json 2> gpurun_out/bench_tcs_nofuse.err; echo "bench rc=$?"
POZ_TCS=0 timeout 300 python bench.py --steps 200 --warmup 20 --no-secondary --no-class-api --no-cpu-baseline > gpurun_out/bench_notcs.json 2> gpurun_out/bench_notcs.err; echo "bench rc=$?"
tail -c 800 gpurun_out/bench_tcs.err
