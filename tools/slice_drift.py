"""How the C4 slice time drifts as the active swarm evolves: per 10 slices, run-kernel ms, rebuilds, max neighbours."""
import json, pathlib, sys
import numpy as np, torch
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import bench

skin = float(sys.argv[1]) if len(sys.argv) > 1 else 2.0
ro = bench.Rollout(0, 42, skin)
d = ro.d
prev = d.stats().rebuilds
for blk in range(15):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    tot = 0.0
    for k in range(10):
        f, t = ro.act_dev[(blk * 10 + k) % 4]
        d.f_swim[: bench.N_AGENTS].copy_(f); d.torque[:, : bench.N_AGENTS].copy_(t)
        a.record(); d.run(100, ro.step_idx); b.record(); torch.cuda.synchronize()
        ro.step_idx += 100
        tot += a.elapsed_time(b)
    st = d.stats()
    cnt = d.lib and None
    print(json.dumps({"skin": skin, "slices": (blk + 1) * 10, "run_ms": tot / 10, "rebuilds_per_slice": (st.rebuilds - prev) / 10, "max_nbr_seen": st.max_neighbors_seen}))
    prev = st.rebuilds
