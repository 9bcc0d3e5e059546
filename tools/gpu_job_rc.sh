#!/bin/bash
timeout -s KILL 300 python -m pytest tests/test_gpu_tiled.py -x -q 2>&1 | tail -3 | tee gpurun_out/r02_tiled_tests.txt
timeout -s KILL 300 python tools/tiled_rebuild_cost.py 2>&1 | tail -1 | tee gpurun_out/r02_tiled_rebuild_cost.txt
for sk in 1.5 2.0 2.5 3.0; do timeout -s KILL 200 python tools/tiled_probe.py $sk 3 2>&1 | grep '"mode": 1' | cut -c1-150; done | tee gpurun_out/r02_tiled_skins.txt
