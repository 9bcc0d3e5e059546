#!/bin/bash
timeout -s KILL 300 python tools/tiled_probe.py 2.0 4 2>&1 | tee gpurun_out/r02_tiled_probe.txt
timeout -s KILL 300 python -m pytest tests/test_gpu_tiled.py -q 2>&1 | tail -15 | tee gpurun_out/r02_tiled_tests.txt
timeout -s KILL 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/r02_gputest1.txt
