#!/bin/bash
timeout -s KILL 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -12 | tee gpurun_out/r02_full2_tests.txt
timeout -s KILL 900 python bench.py --no-extra > gpurun_out/r02_full2_bench.json 2> gpurun_out/r02_full2_bench.err; tail -c 300 gpurun_out/r02_full2_bench.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_full2_bench.json').read().strip().splitlines()[-1])
for k in ('value','ms_per_step','gpu_launches','e2e'):
    print(k, json.dumps(d.get(k))[:400])
print('launch_ms', d['roofline']['launch_ms'], 'frac', d['roofline']['frac'])
print('engine', d['e2e_engine']['ms_per_step'], d['e2e_engine'].get('gpu_launches_per_slice'))
P
timeout -s KILL 300 python tools/bench_envs.py 2>&1 | tail -2 | cut -c1-600
