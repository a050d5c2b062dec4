#!/bin/bash
# the GPU suite several times over: any flaky test shows up as a failure with its name
set -x
for i in 1 2 3 4; do
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_soak$i.log 2>&1; echo "soak $i rc=$?"; tail -2 gpurun_out/pytest_soak$i.log | head -1
done
