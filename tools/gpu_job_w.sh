#!/bin/bash
SWARM_B200_TILE_KCAP=5 timeout -s KILL 600 ncu --set full --clock-control none --import-source on -k regex:k_bd2d_tiled -s 3 -c 1 -o gpurun_out/r02_k_bd2d_tiled_kcap5 -f python tools/run_once_tiled.py 50 6 > gpurun_out/r02_ncu_tiled_kcap5.log 2>&1
tail -2 gpurun_out/r02_ncu_tiled_kcap5.log
