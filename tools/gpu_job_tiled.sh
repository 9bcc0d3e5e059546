#!/bin/bash
timeout -s KILL 500 python -m pytest tests/test_gpu_tiled.py -x -q 2>&1 | tail -30 | tee gpurun_out/r02_tiled_tests.txt
timeout -s KILL 400 tools/ab_bench.sh "SWARM_B200_NO_TILED=1 SWARM_B200_RES_PIPE=0 X=1" 20 50 2>&1 | tee gpurun_out/r02_ab_tiled.txt
