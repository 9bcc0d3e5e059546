#!/bin/bash
for rep in 1; do
for lib in swarmrl_b200/libswarm_b200.so swarmrl_b200/libswarm_b200_var_sw_stream_keep.so _ab/libswarm_b200_232bf73.so; do
SWARM_B200_LIB=$lib timeout -s KILL 600 python tools/bench_domain.py --agents 8000000 --exchange-every 33 --skin 2.0 --margin 0.15 --slices 5 2>gpurun_out/dd8m.err | tail -1 | python tools/dd_line.py "8M/1GPU $lib"
done; done | tee gpurun_out/r02_stream_old_new_ab.txt
tail -c 200 gpurun_out/dd8m.err
