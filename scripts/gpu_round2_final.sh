#!/bin/bash
# end of round 2: the GPU suite twice and the default bench five times (supervised: reports attempts)
set -x
for i in 1 2; do
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_final$i.log 2>&1; echo "suite $i rc=$?"; tail -2 gpurun_out/pytest_final$i.log | head -1
done
for i in 1 2 3 4 5; do
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_final$i.json 2> gpurun_out/bench_final$i.err; echo "bench $i rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/bench_final$i.json').read().strip().splitlines()[-1])
print('attempts', d['attempts'], 'step us', round(d['ms_per_step']*1e3,2), 'e2e us', round(d['e2e']['ms_per_step']*1e3,2), 'class', d['e2e']['class_api']['us_per_epoch'], 'c3', round(d['secondary']['c3']['value']/1e6), 'c4', round(d['secondary']['c4']['value']/1e6,2))
"
done
