"""Per-phase clock64 breakdown of the resident kernel (needs the SW_TIMING build):
   SWARM_B200_LIB=swarmrl_b200/libswarm_b200_timing.so python tools/phase_timing.py"""
import ctypes as C, json, sys, pathlib
import numpy as np, torch
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import bench
from swarmrl_b200 import _lib
from swarmrl_b200.device import DeviceSwarm

NAMES = ["head+rebuild", "pairs+update", "cta_sync_or", "-", "local work s+1 (noise, sincos)", "wait final sync", "ctrl: arrive", "ctrl: poll"]

def prof(d, steps, label):
    out = (C.c_int64 * 8)()
    d.run(steps, 0); _lib.check(d.lib.swarm_debug_timing(d.h, out))
    a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
    a.record(); d.run(steps, steps); b.record(); torch.cuda.synchronize()
    _lib.check(d.lib.swarm_debug_timing(d.h, out))
    cyc = np.array(list(out)[:8], dtype=float) / steps
    print(json.dumps({"case": label, "us_per_step": a.elapsed_time(b) / steps * 1e3, "cycles_per_step": dict(zip(NAMES, cyc.round(0).tolist())),
                      "sum_cycles": float(cyc.sum())}))

for kT in (0.0, bench.KT):
    pos, th = bench.initial_state(1000, 200.0, 1)
    d = DeviceSwarm(2, (200.0,) * 3, True, 1e-3, kT, 1, bench.species_table(), bench.wca_table(), skin=5.0)
    d.set_state(pos, np.zeros(1000, dtype=np.int32), theta=th)
    prof(d, 2000, f"n=1000 kT={kT:.2f}")
for mode in (0, 1):
    for kT in (0.0, bench.KT):
        pos, th = bench.initial_state(bench.N_AGENTS, bench.BOX, 42)
        d = DeviceSwarm(2, (bench.BOX,) * 3, True, 1e-3, kT, 42, bench.species_table(), bench.wca_table(), skin=2.0)
        d.set_state(pos, np.zeros(bench.N_AGENTS, dtype=np.int32), theta=th)
        d.set_tiling(mode)
        prof(d, 500, f"n=100k kT={kT:.2f} {'tiled' if mode else 'resident'}")
        if mode:
            print(json.dumps(d.tiling_stats()))
