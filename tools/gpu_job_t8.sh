#!/bin/bash
timeout -s KILL 600 python -m pytest tests/test_gpu_tiled.py -x -q -k "shared_rotational or environment_kernel" 2>&1 | tail -60 | tee gpurun_out/r02_t8_tests.txt
