#!/bin/bash
set -x
python -m pytest tests/test_gpu_forward_autograd.py tests/test_gpu_ref_suite.py tests/test_gpu_model_api.py -m gpu -q > gpurun_out/pytest_boundary.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_boundary.log
tail -60 gpurun_out/pytest_boundary.log
