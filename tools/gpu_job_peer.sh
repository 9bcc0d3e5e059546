#!/bin/bash
timeout -s KILL 900 python -m pytest tests/test_gpu_domain.py -x -q 2>&1 | tail -15 | tee gpurun_out/r02_peer_tests.txt
timeout -s KILL 300 python -m pytest tests/test_gpu_policy.py -x -q 2>&1 | tail -3
timeout -s KILL 300 python tools/engine_phase_times.py 2>&1 | tail -12 | tee gpurun_out/r02_engine_phase_times.txt
