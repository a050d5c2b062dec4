#!/bin/bash
set -x
timeout 900 python -m pytest tests/test_gpu_infer_f16.py -m gpu -x -q > gpurun_out/pytest_f16.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_f16.log
tail -8 gpurun_out/pytest_f16.log
timeout 300 python scripts/micro/c3_pass.py > gpurun_out/c3_pass.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc3_fwd_x32t -s 9 -c 2 -o gpurun_out/prof_c3h python scripts/micro/c3_pass.py > gpurun_out/ncu_c3h.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_c3h.log
