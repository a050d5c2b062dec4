#!/bin/bash
# GPU-box visit: parity tests + smoke only.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --tb=short --maxfail=40 -p no:cacheprovider "$@" > gpurun_out/pytest.log 2>&1
echo "pytest rc=$?" | tee -a gpurun_out/pytest.log
grep -E "^(FAILED|ERROR)" gpurun_out/pytest.log | head -20
tail -3 gpurun_out/pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
