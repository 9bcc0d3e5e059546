#!/bin/bash
timeout -s KILL 600 python -m pytest tests/test_zz_no_alloc.py -x -q 2>&1 | tail -8 | tee gpurun_out/r02_noalloc_alone.txt
timeout -s KILL 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/r02_gputest_all.txt
timeout -s KILL 600 ncu --set full --clock-control none --import-source on -k regex:k_policy_forward -s 3 -c 1 -o gpurun_out/r02_k_policy_forward -f python tools/run_once_engine.py 6 > gpurun_out/r02_ncu_policy.log 2>&1
timeout -s KILL 600 ncu --set full --clock-control none --import-source on -k regex:k_vis2_cones -s 3 -c 1 -o gpurun_out/r02_k_vis2_cones -f python tools/run_once_engine.py 6 > gpurun_out/r02_ncu_cones.log 2>&1
timeout -s KILL 600 ncu --set full --clock-control none --import-source on -k regex:k_bd2d_tiled -s 5 -c 1 -o gpurun_out/r02_k_bd2d_tiled_final -f python tools/run_once_tiled.py 100 8 > gpurun_out/r02_ncu_tiled.log 2>&1
tail -1 gpurun_out/r02_ncu_policy.log gpurun_out/r02_ncu_cones.log gpurun_out/r02_ncu_tiled.log
