#!/bin/bash
timeout -s KILL 900 python -m pytest tests/test_gpu_domain_engine.py tests/test_gpu_rollout.py -x -q 2>&1 | tail -15 | tee gpurun_out/r02_lv_dd_tests.txt
