#!/bin/bash
timeout -s KILL 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/r02_gputest2.txt
timeout -s KILL 900 python bench.py > gpurun_out/r02_bench1.json 2> gpurun_out/r02_bench1.err; tail -c 600 gpurun_out/r02_bench1.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_bench1.json').read().strip().splitlines()[-1])
for k in ('value','ms_per_step','gpu_launches','e2e','roofline','e2e_engine','clocks'):
    print(k, json.dumps(d.get(k))[:700])
print('c3', json.dumps(d.get('c3_envs',{}).get('strong'))[:300])
print('c5', d.get('c5_domain',{}).get('value'), d.get('c5_domain',{}).get('ms_per_slice'))
P
