#!/bin/bash
set -x
timeout 900 python -m pytest tests/test_gpu_wide_class.py tests/test_gpu_wide.py tests/test_gpu_tc_gemm.py tests/test_gpu_round2.py -m gpu -x -q -k "not philox and not replayed" > gpurun_out/pytest_wide.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_wide.log
tail -40 gpurun_out/pytest_wide.log
