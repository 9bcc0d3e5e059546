#!/bin/bash
# phase clocks (timing build), slope over launch length, ncu full capture of the tiled kernel
SWARM_B200_LIB=swarmrl_b200/libswarm_b200_timing.so timeout -s KILL 300 python tools/phase_timing.py 2>&1 | tee gpurun_out/r02_phase_timing.txt
for st in 50 100 200 400; do timeout -s KILL 120 python tools/run_once_tiled.py $st 8 2>&1 | tail -1; done | tee gpurun_out/r02_tiled_slope.txt
timeout -s KILL 600 ncu --set full --clock-control none --import-source on -k regex:k_bd2d_tiled -s 5 -c 1 -o gpurun_out/r02_k_bd2d_tiled -f python tools/run_once_tiled.py 100 8 > gpurun_out/r02_ncu_tiled.log 2>&1
tail -3 gpurun_out/r02_ncu_tiled.log
timeout -s KILL 300 python -m pytest tests/test_gpu_trajectory.py tests/test_gpu_tiled.py -x -q 2>&1 | tail -15 | tee gpurun_out/r02_traj_tests.txt
