#!/bin/bash
timeout -s KILL 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/r02_full4_tests.txt
timeout -s KILL 300 python __graft_entry__.py smoke 2>&1 | tail -1
