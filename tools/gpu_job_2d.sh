#!/bin/bash
timeout -s KILL 600 python -m pytest tests/test_gpu_engine.py tests/test_gpu_rollout.py -x -q 2>&1 | tail -4
timeout -s KILL 900 python bench.py --no-extra --no-cpu > gpurun_out/r02_2d_bench.json 2> gpurun_out/r02_2d_bench.err; tail -c 200 gpurun_out/r02_2d_bench.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_2d_bench.json').read().strip().splitlines()[-1])
print('value', d['value'], d['ms_per_step'], 'e2e', json.dumps(d['e2e']))
P
