#!/bin/bash
# 8 GPUs: the driver's command line (main C4 line + c5_domain + c3_envs), then the 8 M-particle domain bench with and without the peer exchange
timeout -s KILL 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/r02_bench_n8.json 2> gpurun_out/r02_bench_n8.err
tail -c 300 gpurun_out/r02_bench_n8.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_bench_n8.json').read().strip().splitlines()[-1])
print('value', d['value'], 'ms', d['ms_per_step'], 'n', d['n_gpus'], 'e2e', d['e2e']['value'])
c5=d['c5_domain']; print('c5', c5['value'], c5['ms_per_slice'], 'eff', c5.get('efficiency_vs_n1'), 'n1', c5.get('n1_same_run'), 'transport', c5.get('transport'))
print('c5 timeline', c5.get('timeline_ms_per_slice_max_over_ranks'), 'parity', c5.get('parity_selfcheck',{}).get('pass'), c5.get('parity_selfcheck',{}).get('max_rel_err'), json.dumps(c5.get('parity_full'))[:300])
c3=d['c3_envs']; print('c3 strong', c3['strong']['value'], 'weak', c3['weak']['value'])
print('engine', json.dumps(d.get('e2e_engine'))[:200])
P
for peer in 1 0; do
SWARM_B200_DD_PEER=$peer timeout -s KILL 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 2953$peer tools/bench_domain.py --agents 8000000 --exchange-every 33 --skin 2.0 --margin 0.15 --slices 10 2>gpurun_out/n8_peer_$peer.err | tail -1 | python tools/dd_line.py "8M/8GPU peer=$peer"
done | tee gpurun_out/r02_domain_peer_ab_n8.txt
