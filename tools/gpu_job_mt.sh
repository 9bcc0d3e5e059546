#!/bin/bash
timeout -s KILL 600 python -m pytest tests/test_gpu_engine.py tests/test_gpu_obs.py tests/test_gpu_trajectory.py -x -q 2>&1 | tail -8
timeout -s KILL 900 python tools/measure_table.py 2> gpurun_out/r02_measure_table.err | tee gpurun_out/r02_measure_table.jsonl | cut -c1-400; tail -3 gpurun_out/r02_measure_table.err
