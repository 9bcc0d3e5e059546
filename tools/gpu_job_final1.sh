#!/bin/bash
timeout -s KILL 900 python bench.py > gpurun_out/r02_bench_n1.json 2> gpurun_out/r02_bench_n1.err; tail -c 300 gpurun_out/r02_bench_n1.err
timeout -s KILL 900 python bench.py --impl reference > gpurun_out/r02_bench_reference_arm.json 2> gpurun_out/r02_bench_ref.err; tail -c 300 gpurun_out/r02_bench_ref.err
timeout -s KILL 300 python tools/engine_phase_times.py 2>&1 | tail -12 | tee gpurun_out/r02_engine_phase_times.txt
timeout -s KILL 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches_bench.csv python bench.py --steps 4 --warmup 3 --no-engine > gpurun_out/ncu_bench.log 2>&1
timeout -s KILL 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_engine.csv python tools/run_once_engine.py 12 > gpurun_out/ncu_engine.log 2>&1
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_bench_n1.json').read().strip().splitlines()[-1])
for k in ('value','ms_per_step','gpu_launches','e2e','roofline','e2e_engine','cpu_baseline','clocks'):
    print(k, json.dumps(d.get(k))[:900])
print('c3', json.dumps(d.get('c3_envs',{}).get('strong'))[:300])
print('c5', d.get('c5_domain',{}).get('value'), d.get('c5_domain',{}).get('ms_per_slice'))
print(open('gpurun_out/r02_bench_reference_arm.json').read()[:600])
P
