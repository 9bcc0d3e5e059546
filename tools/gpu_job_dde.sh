#!/bin/bash
# domain-decomposed swarm behind SwarmEngine: the new tests, then the older domain / obs tests that share the kernels
timeout -s KILL 900 python -m pytest tests/test_gpu_domain_engine.py -x -q 2>&1 | tail -30 | tee gpurun_out/r02_dde_tests.txt
timeout -s KILL 600 python -m pytest tests/test_gpu_domain.py tests/test_gpu_obs.py tests/test_gpu_rollout.py -x -q 2>&1 | tail -5 | tee gpurun_out/r02_dde_tests2.txt
