#!/bin/bash
timeout -s KILL 300 python tools/tiled_nosync_probe.py 2>&1 | tee gpurun_out/r02_tiled_nosync.txt
