#!/bin/bash
timeout -s KILL 600 python -m pytest tests/test_gpu_tiled.py -x -q 2>&1 | tail -8 | tee gpurun_out/r02_t7_tests.txt
timeout -s KILL 300 python __graft_entry__.py smoke 2>&1 | tail -2
timeout -s KILL 600 ncu --set full --clock-control none --import-source on -k regex:k_bd2d_env -s 3 -c 1 -o gpurun_out/r02_k_bd2d_env -f python tools/bench_envs.py --slices 6 > gpurun_out/r02_ncu_env.log 2>&1; tail -2 gpurun_out/r02_ncu_env.log | cut -c1-200
timeout -s KILL 600 ncu --set full --clock-control none --import-source on -k regex:k_bd2d_stream -s 4 -c 1 -o gpurun_out/r02_k_bd2d_stream -f python tools/bench_domain.py --agents 1000000 --exchange-every 33 --skin 2.0 --margin 0.15 --slices 2 --warmup 1 > gpurun_out/r02_ncu_stream.log 2>&1; tail -2 gpurun_out/r02_ncu_stream.log | cut -c1-200
ls -la gpurun_out/*.ncu-rep
