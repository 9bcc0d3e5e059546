#!/bin/bash
timeout -s KILL 600 python -m pytest tests/test_gpu_obs.py tests/test_gpu_fullsize.py tests/test_gpu_domain_engine.py -x -q 2>&1 | tail -5 | tee gpurun_out/r02_ab1_tests.txt
timeout -s KILL 300 python tools/bench_vision_queue.py 2>&1 | tail -4 | tee gpurun_out/r02_vision_queue_ab.txt
for rep in 1 2; do
for lib in swarmrl_b200/libswarm_b200.so swarmrl_b200/libswarm_b200_var_sw_tile_no_mbar.so; do
  echo "$lib $(SWARM_B200_LIB=$lib timeout -s KILL 120 python tools/run_once_tiled.py 100 10 2>&1 | tail -1 | cut -c1-220)"
done; done | tee gpurun_out/r02_mbar_ab.txt
