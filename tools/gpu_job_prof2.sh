#!/bin/bash
# round-2 final profiles: ncu full captures of the tiled run kernel and the queued cone kernel, launch lists of the bench and the engine path
timeout -s KILL 120 python tools/run_once_tiled.py 100 8 2>&1 | tail -1 | cut -c1-200
timeout -s KILL 600 ncu --set full --clock-control none --import-source on -k regex:k_bd2d_tiled -s 5 -c 1 -o gpurun_out/r02_k_bd2d_tiled -f python tools/run_once_tiled.py 100 8 > gpurun_out/r02_ncu_tiled.log 2>&1; tail -2 gpurun_out/r02_ncu_tiled.log
timeout -s KILL 600 ncu --set full --clock-control none --import-source on -k regex:k_vis2_cones_q -s 3 -c 1 -o gpurun_out/r02_k_vis2_cones_q -f python tools/run_once_engine.py 6 > gpurun_out/r02_ncu_vis.log 2>&1; tail -2 gpurun_out/r02_ncu_vis.log
timeout -s KILL 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches_bench.csv python bench.py --steps 4 --warmup 3 --no-engine --no-extra --no-cpu > gpurun_out/ncu_bench.log 2>&1; tail -1 gpurun_out/ncu_bench.log | cut -c1-200
timeout -s KILL 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_engine.csv python tools/run_once_engine.py 12 > gpurun_out/ncu_engine.log 2>&1; tail -1 gpurun_out/ncu_engine.log | cut -c1-200
ls -la gpurun_out/*.ncu-rep
