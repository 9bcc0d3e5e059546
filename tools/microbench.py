"""Device micro-benchmarks used to size the kernels (run under gpurun): Gaussian generation rate,
per-step cost of the persistent kernel vs Verlet skin."""
import sys, pathlib, json
import numpy as np, torch
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import bench

def ev():
    return torch.cuda.Event(enable_timing=True)

ro = bench.Rollout(0, 42, 0.4)
d = ro.d
n = 4_000_000
out = torch.empty((n, 4), dtype=torch.float64, device=d.dev)
import ctypes as C
from swarmrl_b200 import _lib
for _ in range(3):
    _lib.check(d.lib.swarm_gaussians(d.h, n, 1, 0, out.data_ptr(), d._stream()))
a, b = ev(), ev(); a.record()
for k in range(10):
    _lib.check(d.lib.swarm_gaussians(d.h, n, k, 0, out.data_ptr(), d._stream()))
b.record(); torch.cuda.synchronize()
ms = a.elapsed_time(b) / 10
print(json.dumps({"gaussians_4M_ms": ms, "ns_per_block_of_4": ms * 1e6 / n, "blocks_per_s": n / ms * 1e3}))
for skin in (0.4, 0.8, 1.2, 2.0, 3.0):
    ro = bench.Rollout(0, 42, skin)
    for k in range(3):
        ro.d.run(100, k * 100)
    r0 = ro.d.stats().rebuilds
    a, b = ev(), ev(); a.record()
    for k in range(10):
        ro.d.run(100, 300 + k * 100)
    b.record(); torch.cuda.synchronize()
    st = ro.d.stats()
    print(json.dumps({"skin": skin, "us_per_step": a.elapsed_time(b) / 1000 * 1e3, "rebuilds_per_100": (st.rebuilds - r0) / 10,
                      "max_nbr": st.max_neighbors_seen, "cap": st.capacity, "overflow": st.overflow}))
# grid-barrier cost: 1000 particles forced onto G CTAs (SWARM_B200_GRID), so a step is ~only the barrier
import os
for G in (74, 148, 296, 444):
    os.environ["SWARM_B200_GRID"] = str(G)
    from swarmrl_b200.device import DeviceSwarm
    pos, th = bench.initial_state(1000, 200.0, 1)
    dd = DeviceSwarm(2, (200.0,) * 3, True, 1e-3, 0.0, 1, bench.species_table(), bench.wca_table(), skin=5.0)
    dd.set_state(pos, np.zeros(1000, dtype=np.int32), theta=th)
    dd.run(1000, 0); torch.cuda.synchronize()
    a, b = ev(), ev(); a.record(); dd.run(4000, 1000); b.record(); torch.cuda.synchronize()
    print(json.dumps({"barrier_grid": G, "us_per_step": a.elapsed_time(b) / 4000 * 1e3, "rebuilds": dd.stats().rebuilds}))
os.environ.pop("SWARM_B200_GRID")
