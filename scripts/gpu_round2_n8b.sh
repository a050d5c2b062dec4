#!/bin/bash
# the 8-GPU leg of the scaling bench as the driver launches it (native arm only)
set -x
N=${1:-8}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_c5_n$N.json 2> gpurun_out/bench_c5_n$N.err; echo "c5 n$N rc=$?"
tail -c 400 gpurun_out/bench_c5_n$N.err
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_c5_n$N.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d.get('dp_check',{}).get('ok'), d['e2e']['value'], d.get('exposed_comm_ms'), d.get('host_numa'))
PY
