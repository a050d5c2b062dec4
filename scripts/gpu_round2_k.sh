#!/bin/bash
set -x
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_full.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_full.log
tail -8 gpurun_out/pytest_full.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
tail -c 600 gpurun_out/bench_default.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_default.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['e2e']['ms_per_step'], d['e2e'].get('class_api'), d['roofline']['kernel'], d['roofline']['frac'], d['roofline']['kernels_us'])
for k,v in d.get('secondary',{}).items(): print(k, v.get('value'), v.get('ms_per_step'))
PY
