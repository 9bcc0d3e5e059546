#!/bin/bash
timeout -s KILL 600 python -m pytest tests/test_gpu_tiled.py tests/test_gpu_engine.py tests/test_gpu_fullsize.py -x -q 2>&1 | tail -4 | tee gpurun_out/r02_ab6_tests.txt
for rep in 1 2; do
  echo "wrapflag $(timeout -s KILL 120 python tools/run_once_tiled.py 100 12 2>&1 | tail -1 | cut -c1-150)"
done | tee gpurun_out/r02_wrapflag.txt
