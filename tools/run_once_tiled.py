"""Smallest driver of k_bd2d_tiled for ncu: 100k agents, a few launches of `steps` Brownian steps (the first sorts)."""
import sys, pathlib
import numpy as np, torch
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import bench
from swarmrl_b200.device import DeviceSwarm
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 100
launches = int(sys.argv[2]) if len(sys.argv) > 2 else 6
pos, th = bench.initial_state(bench.N_AGENTS, bench.BOX, 42)
d = DeviceSwarm(2, (bench.BOX,) * 3, True, 1e-3, bench.KT, 42, bench.species_table(), bench.wca_table(), skin=2.0)
d.set_state(pos, np.zeros(bench.N_AGENTS, dtype=np.int32), theta=th)
f, t = bench.action_sets(bench.N_AGENTS, 7, 1)[0]
d.set_actions_host(f, t)
ev = []
for k in range(launches):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); d.run(steps, k * steps); b.record(); ev.append((a, b))
st = d.stats()
print("ok", st.steps_done, st.rebuilds, st.overflow, [round(a.elapsed_time(b), 4) for a, b in ev], d.tiling_stats())
