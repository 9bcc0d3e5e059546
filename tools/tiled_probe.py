"""Run-kernel time of the C4 swarm per block of 10 slices, resident (tiling off) against tiled (automatic), with the tiled
kernel's counters: python tools/tiled_probe.py [skin] [blocks]"""
import json, pathlib, sys
import numpy as np, torch
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import bench

skin = float(sys.argv[1]) if len(sys.argv) > 1 else 2.0
blocks = int(sys.argv[2]) if len(sys.argv) > 2 else 4
for mode in (0, 1):
    ro = bench.Rollout(0, 42, skin)
    d = ro.d
    d.set_tiling(mode)
    prev = d.stats().rebuilds
    for blk in range(blocks):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        tot = 0.0
        for k in range(10):
            f, t = ro.act_dev[(blk * 10 + k) % 4]
            d.f_swim[: bench.N_AGENTS].copy_(f); d.torque[:, : bench.N_AGENTS].copy_(t)
            ro.flush_l2()
            a.record(); d.run(100, ro.step_idx); b.record(); torch.cuda.synchronize()
            ro.step_idx += 100
            tot += a.elapsed_time(b)
        st = d.stats()
        print(json.dumps({"mode": mode, "skin": skin, "slices": (blk + 1) * 10, "run_ms": round(tot / 10, 4),
                          "rebuilds_per_slice": (st.rebuilds - prev) / 10, "max_nbr_seen": st.max_neighbors_seen,
                          "overflow": st.overflow, "tiling": d.tiling_stats()}), flush=True)
        prev = st.rebuilds
    chk = d.pos[:2, :bench.N_AGENTS].double().sum().item()
    print(json.dumps({"mode": mode, "checksum": chk}))
    d.close()
