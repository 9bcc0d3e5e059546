"""Smallest driver of k_bd2d_stream for ncu: 1M agents (one slab of BASELINE config 5), launches of `steps` steps."""
import sys, pathlib
import numpy as np, torch
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import bench
from swarmrl_b200.device import DeviceSwarm
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1_000_000
L = float(np.sqrt(n * np.pi / 0.10))
pos, th = bench.initial_state(n, L, 42)
d = DeviceSwarm(2, (L,) * 3, True, 1e-3, bench.KT, 42, bench.species_table(), bench.wca_table(), skin=2.0)
d.set_state(pos, np.zeros(n, dtype=np.int32), theta=th)
d.f_swim[:n] = 5.0 * bench.stokes(1.0)[0]
for k in range(4):
    d.run(steps, k * steps)
st = d.stats()
print("ok", st.steps_done, st.rebuilds, st.overflow)
