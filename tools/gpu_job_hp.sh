#!/bin/bash
timeout -s KILL 600 python tools/engine_phase_times.py 2>&1 | tail -8 | tee gpurun_out/r02_engine_phase_times.txt
