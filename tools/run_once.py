"""Smallest driver of the hot kernel for ncu: 100k agents, a few launches of `steps` Brownian steps."""
import sys, pathlib
import numpy as np, torch
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import bench
from swarmrl_b200.device import DeviceSwarm
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
kT = float(sys.argv[2]) if len(sys.argv) > 2 else bench.KT
pos, th = bench.initial_state(bench.N_AGENTS, bench.BOX, 42)
d = DeviceSwarm(2, (bench.BOX,) * 3, True, 1e-3, kT, 42, bench.species_table(), bench.wca_table(), skin=2.0)
d.set_state(pos, np.zeros(bench.N_AGENTS, dtype=np.int32), theta=th)
f, t = bench.action_sets(bench.N_AGENTS, 7, 1)[0]
d.set_actions_host(f, t)
for k in range(4):
    d.run(steps, k * steps)
st = d.stats()
print("ok", st.steps_done, st.rebuilds, st.overflow)
