#!/bin/bash
# 2 GPUs: the 2M-particle domain bench with and without the peer exchange, alternating
for peer in 1 0 1 0; do
SWARM_B200_DD_PEER=$peer timeout -s KILL 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2951$peer tools/bench_domain.py --agents 2000000 --exchange-every 33 --skin 2.0 --margin 0.15 --slices 10 2>gpurun_out/n2_peer_$peer.err | tail -1 | python tools/dd_line.py "peer=$peer"
done | tee gpurun_out/r02_domain_peer_ab_n2.txt
tail -c 300 gpurun_out/n2_peer_1.err
