#!/bin/bash
set -x
timeout 1800 python -m pytest tests/test_gpu_model_api.py tests/test_gpu_ref_suite.py tests/test_gpu_net.py tests/test_gpu_forward_autograd.py tests/test_gpu_wide_class.py -m gpu -x -q > gpurun_out/pytest_api.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_api.log
tail -4 gpurun_out/pytest_api.log
timeout 600 python scripts/micro/api_loop_cost.py 2>&1 | tail -6
