#!/bin/bash
timeout -s KILL 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 | tee gpurun_out/r02_full3_tests.txt
timeout -s KILL 600 python tools/bench_domain.py --agents 8000000 --exchange-every 33 --skin 2.0 --margin 0.15 --slices 5 2>gpurun_out/dd8m.err | tail -1 | python tools/dd_line.py "8M/1GPU"
timeout -s KILL 600 python tools/bench_domain.py --agents 1000000 --exchange-every 33 --skin 2.0 --margin 0.15 --slices 10 2>>gpurun_out/dd8m.err | tail -1 | python tools/dd_line.py "1M/1GPU"
tail -c 300 gpurun_out/dd8m.err
