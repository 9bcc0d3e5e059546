#!/bin/bash
timeout -s KILL 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/r02_bench_n8.json 2> gpurun_out/r02_bench_n8.err; tail -c 400 gpurun_out/r02_bench_n8.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_bench_n8.json').read().strip().splitlines()[-1])
for k in ('value','ms_per_step','e2e','clocks'):
    print(k, json.dumps(d.get(k))[:300])
print('c3', json.dumps(d.get('c3_envs'))[:900])
print('c5', json.dumps(d.get('c5_domain'))[:1500])
P
