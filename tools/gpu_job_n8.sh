#!/bin/bash
for cfg in "33 2.0 0.15" "50 2.0 0.15" "40 2.0 0.12" "25 2.0 0.15"; do set -- $cfg
timeout -s KILL 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 tools/bench_domain.py --agents 8000000 --exchange-every $1 --skin $2 --margin $3 --slices 5 2>/dev/null | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print('k', $1, 'skin', $2, 'margin', $3, 'ms/slice', round(d['ms_per_slice'], 3), 'value %.4g' % d['value'], 'redund', round(d.get('redundancy', 0), 4), 'rebuilds', d.get('rebuilds'), 'exch', d.get('exchanges'))"
done | tee gpurun_out/r02_domain_sweep_n8.txt
