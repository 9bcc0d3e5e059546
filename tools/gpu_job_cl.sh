#!/bin/bash
timeout -s KILL 900 python -m pytest tests/test_gpu_engine.py tests/test_gpu_tiled.py tests/test_gpu_rollout.py tests/test_gpu_fullsize.py -x -q 2>&1 | tail -15 | tee gpurun_out/r02_cl_tests.txt
timeout -s KILL 600 python tools/measure_table.py 2> gpurun_out/mt.err | head -2 | cut -c1-330
SWARM_B200_NO_CLUSTER=1 timeout -s KILL 600 python tools/measure_table.py 2> gpurun_out/mt.err | head -2 | cut -c1-200
