#!/bin/bash
timeout -s KILL 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/r02_bench_n8.json 2> gpurun_out/r02_bench_n8.err
tail -c 300 gpurun_out/r02_bench_n8.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_bench_n8.json').read().strip().splitlines()[-1])
print('value', d['value'], 'ms', d['ms_per_step'], 'n', d['n_gpus'], 'e2e', d['e2e']['value'], 'engine', d['e2e_engine']['value'])
c5=d['c5_domain']; print('c5', c5['value'], c5['ms_per_slice'], 'eff', c5.get('efficiency_vs_n1'), 'n1', c5.get('n1_same_run'))
print('c5 timeline', c5.get('timeline_ms_per_slice_max_over_ranks'), 'parity', c5.get('parity_selfcheck',{}).get('pass'), c5.get('parity_selfcheck',{}).get('max_rel_err'), c5.get('parity_full',{}).get('max_rel_err_x_y_theta'))
c3=d['c3_envs']; print('c3 strong', c3['strong']['value'], 'weak', c3['weak']['value'])
P
