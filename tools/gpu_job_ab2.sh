#!/bin/bash
timeout -s KILL 900 python -m pytest tests/test_gpu_policy.py tests/test_gpu_rollout.py tests/test_gpu_tiled.py tests/test_gpu_domain_engine.py -x -q 2>&1 | tail -5 | tee gpurun_out/r02_ab2_tests.txt
timeout -s KILL 300 python tools/engine_phase_times.py 2>&1 | tail -12 | tee gpurun_out/r02_engine_phase_times.txt
timeout -s KILL 120 python tools/pcie_probe.py 2>&1 | tee gpurun_out/r02_pcie_probe.txt
timeout -s KILL 900 python bench.py --no-extra > gpurun_out/r02_ab2_bench.json 2> gpurun_out/r02_ab2_bench.err; tail -c 300 gpurun_out/r02_ab2_bench.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_ab2_bench.json').read().strip().splitlines()[-1])
for k in ('value','ms_per_step','gpu_launches','e2e'):
    print(k, json.dumps(d.get(k))[:400])
print('launch_ms', d['roofline']['launch_ms'], 'frac', d['roofline']['frac'])
print('engine', d['e2e_engine']['ms_per_step'], d['e2e_engine'].get('gpu_launches_per_slice'))
P
