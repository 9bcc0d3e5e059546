#!/bin/bash
timeout -s KILL 300 python -m pytest tests/test_gpu_tiled.py -x -q 2>&1 | tail -15 | tee gpurun_out/r02_tiled_tests.txt
SWARM_B200_LIB=swarmrl_b200/libswarm_b200_timing.so timeout -s KILL 300 python tools/phase_timing.py 2>&1 | tail -4 | tee gpurun_out/r02_phase_timing.txt
timeout -s KILL 300 python tools/tiled_probe.py 2.0 4 2>&1 | tee gpurun_out/r02_tiled_probe.txt
for st in 100 400; do timeout -s KILL 120 python tools/run_once_tiled.py $st 8 2>&1 | tail -1; done | tee gpurun_out/r02_tiled_slope.txt
