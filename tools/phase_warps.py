"""Per-warp clock64 breakdown of the tiled kernel (SW_TIMING build): which warp of which CTA spends how long where.
   SWARM_B200_LIB=swarmrl_b200/libswarm_b200_timing.so python tools/phase_warps.py"""
import ctypes as C, json, os, sys, pathlib
import numpy as np, torch
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import bench
from swarmrl_b200 import _lib
from swarmrl_b200.device import DeviceSwarm

NAMES = ["barrier_wait(deferred)", "pairs+update", "sync:links+vote", "sync:rest", "local|halo_load", "barrier", "sync:search", "sync:barrier+halo"]
pos, th = bench.initial_state(bench.N_AGENTS, bench.BOX, 42)
d = DeviceSwarm(2, (bench.BOX,) * 3, True, 1e-3, bench.KT, 42, bench.species_table(), bench.wca_table(), skin=2.0)
d.set_state(pos, np.zeros(bench.N_AGENTS, dtype=np.int32), theta=th)
f, t = bench.action_sets(bench.N_AGENTS, 7, 1)[0]
d.set_actions_host(f, t)
out = (C.c_int64 * 8)()
d.run(300, 0)
step = 300
for cta in (0, 74)[: int(os.environ.get('PW_CTAS', '2'))]:
    for warp in (0, 5, 12, 22):
        os.environ["SWARM_B200_PROF_CTA"], os.environ["SWARM_B200_PROF_WARP"] = str(cta), str(warp)
        _lib.check(d.lib.swarm_debug_timing(d.h, out))
        d.run(100, step); step += 100
        _lib.check(d.lib.swarm_debug_timing(d.h, out))
        cyc = (np.array(list(out)[:8], dtype=float) / 100).round(0)
        print(json.dumps({"cta": cta, "warp": warp, "cycles_per_step": {k: v for k, v in zip(NAMES, cyc.tolist()) if k != "-"}, "sum": float(cyc.sum())}), flush=True)
print(json.dumps(d.tiling_stats()))
