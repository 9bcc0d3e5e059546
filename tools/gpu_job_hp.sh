#!/bin/bash
timeout -s KILL 600 python -m pytest tests/test_gpu_obs.py tests/test_gpu_rollout.py tests/test_gpu_fullsize.py -x -q 2>&1 | tail -8
timeout -s KILL 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_launches_engine.csv python tools/run_once_engine.py 4 > gpurun_out/r02_ncu_engine.log 2>&1
python - <<'P'
import csv
rows = list(csv.reader(open('gpurun_out/r02_launches_engine.csv')))
hdr = next(i for i, r in enumerate(rows) if 'Kernel Name' in r)
h = rows[hdr]; ik, iv = h.index('Kernel Name'), h.index('Metric Value')
seq = [(r[ik].split('(')[0][:60], float(r[iv].replace(',', ''))) for r in rows[hdr + 1:] if len(r) > iv]
for name, v in seq[-12:]: print(f'{v/1000:9.1f} us  {name}')
P
timeout -s KILL 600 python tools/engine_first_vs_second.py 2>&1 | tail -4
