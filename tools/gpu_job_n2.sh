#!/bin/bash
timeout -s KILL 600 python -m pytest tests/test_gpu_domain.py -x -q 2>&1 | tail -4
timeout -s KILL 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r02_bench_n2.json 2> gpurun_out/r02_bench_n2.err
tail -c 400 gpurun_out/r02_bench_n2.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_bench_n2.json').read().strip().splitlines()[-1])
print('value', d['value'], 'ms', d['ms_per_step'], 'n', d['n_gpus'])
c5=d['c5_domain']; print('c5', c5['value'], c5['ms_per_slice'], 'eff', c5.get('efficiency_vs_n1'), 'parity', c5.get('parity_selfcheck',{}).get('pass'), c5.get('parity_selfcheck',{}).get('max_rel_err'), 'full', json.dumps(c5.get('parity_full'))[:300])
c3=d['c3_envs']; print('c3 strong', c3['strong']['value'], 'weak', c3['weak']['value'])
print('engine', json.dumps(d.get('e2e_engine'))[:300])
P
