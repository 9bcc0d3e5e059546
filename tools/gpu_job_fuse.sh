#!/bin/bash
timeout -s KILL 300 python tools/engine_phase_times.py 2>&1 | tail -12 | tee gpurun_out/r02_engine_phase_times.txt
timeout -s KILL 900 python bench.py --no-engine > gpurun_out/r02_bench_fuse.json 2> gpurun_out/r02_bench_fuse.err; tail -c 300 gpurun_out/r02_bench_fuse.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_bench_fuse.json').read().strip().splitlines()[-1])
for k in ('value','ms_per_step','gpu_launches','e2e','roofline','clocks'):
    print(k, json.dumps(d.get(k))[:500])
P
