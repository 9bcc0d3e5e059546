#!/bin/bash
timeout -s KILL 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 4 --steps 20 --warmup 3 > gpurun_out/r02_bench_n4.json 2> gpurun_out/r02_bench_n4.err
tail -c 200 gpurun_out/r02_bench_n4.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_bench_n4.json').read().strip().splitlines()[-1])
print('value', d['value'], 'ms', d['ms_per_step'], 'n', d['n_gpus'], 'e2e', d['e2e']['value'])
c5=d['c5_domain']; print('c5', c5['value'], c5['ms_per_slice'], 'eff', c5.get('efficiency_vs_n1'), 'parity', c5.get('parity_selfcheck',{}).get('pass'))
c3=d['c3_envs']; print('c3 strong', c3['strong']['value'], 'weak', c3['weak']['value'])
P
