#!/bin/bash
timeout -s KILL 600 python -m pytest tests/test_gpu_obs.py tests/test_gpu_domain_engine.py -x -q 2>&1 | tail -3
timeout -s KILL 300 python tools/bench_vision_queue.py 2>&1 | tail -3
