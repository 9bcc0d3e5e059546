#!/bin/bash
# full GPU test suite + the default bench line at HEAD
timeout -s KILL 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/r02_head_tests.txt
timeout -s KILL 900 python bench.py > gpurun_out/r02_head_bench_n1.json 2> gpurun_out/r02_head_bench_n1.err; tail -c 300 gpurun_out/r02_head_bench_n1.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_head_bench_n1.json').read().strip().splitlines()[-1])
for k in ('value','ms_per_step','gpu_launches','e2e','roofline','clocks'):
    print(k, json.dumps(d.get(k))[:700])
print('engine', d['e2e_engine']['ms_per_step'])
P
