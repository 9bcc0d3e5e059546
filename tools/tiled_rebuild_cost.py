"""Marginal cost of one synchronisation point of the tiled kernel: the same 1000 steps with the cadence capped at 5, 10, 20
and uncapped (SWARM_B200_TILE_KCAP); slope of time against rebuilds."""
import json, os, pathlib, sys
import numpy as np, torch
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import bench
from swarmrl_b200.device import DeviceSwarm
rows = []
for cap in (5, 10, 20, 512):
    os.environ["SWARM_B200_TILE_KCAP"] = str(cap)
    pos, th = bench.initial_state(bench.N_AGENTS, bench.BOX, 42)
    d = DeviceSwarm(2, (bench.BOX,) * 3, True, 1e-3, bench.KT, 42, bench.species_table(), bench.wca_table(), skin=2.0)
    d.set_state(pos, np.zeros(bench.N_AGENTS, dtype=np.int32), theta=th)
    f, t = bench.action_sets(bench.N_AGENTS, 7, 1)[0]
    d.set_actions_host(f, t)
    for k in range(3): d.run(100, k * 100)
    r0 = d.stats().rebuilds
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for k in range(3, 13): d.run(100, k * 100)
    b.record(); torch.cuda.synchronize()
    st = d.stats()
    rows.append((cap, a.elapsed_time(b) * 1e3, st.rebuilds - r0))
    print(json.dumps({"k_cap": cap, "us_per_1000_steps": rows[-1][1], "rebuilds": rows[-1][2], "tiling": d.tiling_stats()}), flush=True)
(c0, t0, r0), (c1, t1, r1) = rows[0], rows[-1]
print(json.dumps({"us_per_rebuild": (t0 - t1) / (r0 - r1), "us_per_step_without_rebuilds": (t1 - r1 * (t0 - t1) / (r0 - r1)) / 1000}))
