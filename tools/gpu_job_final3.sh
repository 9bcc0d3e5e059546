#!/bin/bash
timeout -s KILL 900 python bench.py > gpurun_out/r02_bench_n1.json 2> gpurun_out/r02_bench_n1.err; tail -c 200 gpurun_out/r02_bench_n1.err
timeout -s KILL 900 python bench.py --impl reference > gpurun_out/r02_bench_reference_arm.json 2> gpurun_out/r02_bench_ref.err; tail -c 200 gpurun_out/r02_bench_ref.err
timeout -s KILL 300 python tools/engine_phase_times.py 2>&1 | tail -12 > gpurun_out/r02_engine_phase_times.txt
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_bench_n1.json').read().strip().splitlines()[-1])
print('value', d['value'], d['ms_per_step'], 'launches', d['gpu_launches'])
print('e2e', json.dumps(d['e2e']))
r=d['roofline']; print('launch_ms', r['launch_ms'], 'frac', r['frac'], 'issue', r['issue_frac'], 'share', r['share_of_step'])
print('engine', d['e2e_engine']['ms_per_step'], d['e2e_engine']['value'])
print('cpu', d['cpu_baseline']['value'], 'c3', d['c3_envs']['strong']['value'], 'c5', d['c5_domain']['value'], d['c5_domain']['ms_per_slice'])
print(open('gpurun_out/r02_bench_reference_arm.json').read()[:160])
P
cat gpurun_out/r02_engine_phase_times.txt | head -6
