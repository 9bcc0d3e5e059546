#!/bin/bash
for f in 0.9 0.95 1.0 0.9 0.95 1.0; do
SWARM_B200_TILE_SAFETY=$f timeout -s KILL 300 python bench.py --no-extra --no-cpu 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
t, te = d['roofline']['tiling'], d['e2e_engine']['tiling']
print('safety $f', 'value_ms %.4f' % d['ms_per_step'], 'launch_ms %.4f' % d['roofline']['launch_ms'], 'rollbacks', t['rollbacks'], 'cadence', t['cadence'], '| engine_ms %.4f' % d['e2e_engine']['ms_per_step'], 'rollbacks', te['rollbacks'], 'cadence', te['cadence'], 'rebuilds/slice', d['config'].get('rebuilds_per_slice'))
"
done | tee gpurun_out/r02_tile_safety_ab2.txt
