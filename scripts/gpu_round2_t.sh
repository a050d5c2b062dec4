#!/bin/bash
set -x
for f in 1 0; do
POZ_FORK_STEP_BEGIN=$f timeout 600 python bench.py --steps 200 --warmup 20 --no-class-api --no-secondary --no-cpu-baseline > gpurun_out/bench_sb$f.json 2> gpurun_out/bench_sb$f.err; echo "bench rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/bench_sb$f.json').read().strip().splitlines()[-1])
print('fork_step_begin $f', d['ms_per_step'], d['value'], d['e2e']['ms_per_step'], d['with_dropout_0.2']['ms_per_step'])
"
done
