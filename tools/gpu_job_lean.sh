#!/bin/bash
timeout -s KILL 900 python -m pytest tests/test_gpu_domain.py tests/test_gpu_domain_engine.py -x -q 2>&1 | tail -4 | tee gpurun_out/r02_lean_tests.txt
timeout -s KILL 600 python tools/bench_domain.py --agents 1000000 --exchange-every 33 --skin 2.0 --margin 0.15 --slices 10 2>gpurun_out/dd1m.err | tail -1 | python tools/dd_line.py "1M/1GPU lean"
