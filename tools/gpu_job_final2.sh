#!/bin/bash
# round-2 final N=1 set: representative ncu capture of the tiled kernel (a launch without a rollback), summary written where
# bench.py reads it, then the bench line, the reference arm, the engine phase times and the measurement table
timeout -s KILL 600 ncu --set full --clock-control none --import-source on -k regex:k_bd2d_tiled -s 3 -c 1 -o gpurun_out/r02_k_bd2d_tiled -f python tools/run_once_tiled.py 100 6 > gpurun_out/r02_ncu_tiled.log 2>&1; grep "^ok" gpurun_out/r02_ncu_tiled.log | cut -c1-200
python tools/ncu_summary.py gpurun_out/r02_k_bd2d_tiled.ncu-rep 100 > gpurun_out/r02_ncu_k_bd2d_tiled_summary.txt && cp gpurun_out/r02_ncu_k_bd2d_tiled_summary.txt profiles/r02_ncu_k_bd2d_tiled_summary.txt
grep "gpu__time_duration\|smsp__inst_executed.sum" profiles/r02_ncu_k_bd2d_tiled_summary.txt
timeout -s KILL 900 python bench.py > gpurun_out/r02_bench_n1.json 2> gpurun_out/r02_bench_n1.err; tail -c 300 gpurun_out/r02_bench_n1.err
timeout -s KILL 900 python bench.py --impl reference > gpurun_out/r02_bench_reference_arm.json 2> gpurun_out/r02_bench_ref.err; tail -c 300 gpurun_out/r02_bench_ref.err
timeout -s KILL 300 python tools/engine_phase_times.py 2>&1 | tail -12 | tee gpurun_out/r02_engine_phase_times.txt
timeout -s KILL 600 python tools/measure_table.py > gpurun_out/r02_measure_table.jsonl 2> gpurun_out/mt.err; tail -c 200 gpurun_out/mt.err; cut -c1-260 gpurun_out/r02_measure_table.jsonl
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_bench_n1.json').read().strip().splitlines()[-1])
for k in ('value','ms_per_step','gpu_launches','e2e','roofline','e2e_engine','cpu_baseline','clocks'):
    print(k, json.dumps(d.get(k))[:1000])
print('c3', json.dumps(d.get('c3_envs',{}).get('strong'))[:300])
print('c5', json.dumps({k: v for k, v in d.get('c5_domain',{}).items() if not isinstance(v, (dict, list))})[:900])
print(open('gpurun_out/r02_bench_reference_arm.json').read()[:700])
P
