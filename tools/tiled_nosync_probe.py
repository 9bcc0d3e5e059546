"""How much of a tiled step is the synchronisation points: the same swarm with the physical temperature (a rebuild every
~45 steps) and with a tiny one (the noise code still runs, but nothing moves far: a rebuild every ~200 steps)."""
import json, pathlib, sys
import numpy as np, torch
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import bench
from swarmrl_b200.device import DeviceSwarm
for kT in (bench.KT, 1e-4):
    for mode in (0, 1):
        pos, th = bench.initial_state(bench.N_AGENTS, bench.BOX, 42)
        d = DeviceSwarm(2, (bench.BOX,) * 3, True, 1e-3, kT, 42, bench.species_table(), bench.wca_table(), skin=2.0)
        d.set_state(pos, np.zeros(bench.N_AGENTS, dtype=np.int32), theta=th)
        d.set_tiling(mode)
        f, t = bench.action_sets(bench.N_AGENTS, 7, 1)[0]
        d.set_actions_host(f, t)
        for k in range(3): d.run(100, k * 100)
        r0 = d.stats().rebuilds
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for k in range(3, 13): d.run(100, k * 100)
        b.record(); torch.cuda.synchronize()
        st = d.stats()
        print(json.dumps({"kT": kT, "tiled": mode, "us_per_step": a.elapsed_time(b), "rebuilds_per_100": (st.rebuilds - r0) / 10, "tiling": d.tiling_stats()}), flush=True)
