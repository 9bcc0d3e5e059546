#!/bin/bash
timeout -s KILL 600 python -m pytest tests/test_gpu_tiled.py tests/test_gpu_engine.py -x -q 2>&1 | tail -4 | tee gpurun_out/r02_ab3_tests.txt
for rep in 1 2; do
for lib in swarmrl_b200/libswarm_b200.so swarmrl_b200/libswarm_b200_var_sw_tile_pair_serial.so; do
  echo "$lib $(SWARM_B200_LIB=$lib timeout -s KILL 120 python tools/run_once_tiled.py 100 12 2>&1 | tail -1 | cut -c1-150)"
done; done | tee gpurun_out/r02_pair4_ab.txt
