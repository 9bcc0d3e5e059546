#!/bin/bash
timeout -s KILL 600 python -m pytest tests/test_gpu_engine.py -x -q -k "cluster_kernel" 2>&1 | tail -5 | tee gpurun_out/r02_cl2_tests.txt
for skin in 0.4 1.0 2.0; do timeout -s KILL 120 python tools/run_once_cluster.py 1000 4 $skin 2>&1 | tail -1; done | tee gpurun_out/r02_cluster_skin.txt
timeout -s KILL 600 ncu --set full --clock-control none --import-source on -k regex:k_bd2d_cluster -s 2 -c 1 -o gpurun_out/r02_k_bd2d_cluster -f python tools/run_once_cluster.py 1000 4 > gpurun_out/r02_ncu_cluster.log 2>&1; tail -1 gpurun_out/r02_ncu_cluster.log
timeout -s KILL 600 python tools/measure_table.py > gpurun_out/r02_measure_table.jsonl 2> gpurun_out/mt.err; cut -c1-300 gpurun_out/r02_measure_table.jsonl | head -3
