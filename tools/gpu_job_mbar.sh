#!/bin/bash
timeout -s KILL 300 python -m pytest tests/test_gpu_tiled.py -x -q 2>&1 | tail -5 | tee gpurun_out/r02_tiled_tests.txt
timeout -s KILL 300 python tools/tiled_nosync_probe.py 2>&1 | cut -c1-200 | tee gpurun_out/r02_tiled_nosync.txt
timeout -s KILL 300 python tools/tiled_probe.py 2.0 4 2>&1 | grep -v '"mode": 0' | cut -c1-200 | tee gpurun_out/r02_tiled_probe.txt
SWARM_B200_LIB=swarmrl_b200/libswarm_b200_timing.so timeout -s KILL 300 python tools/phase_warps.py 2>&1 | tee gpurun_out/r02_phase_warps.txt
