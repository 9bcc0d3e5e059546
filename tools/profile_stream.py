"""Where a 1M-particle slab spends its time (run under gpurun): per-step cost of k_bd2d_stream with no
rebuild (huge skin budget, kT small), cost of one Verlet rebuild, and cost of one domain exchange."""
import json
import pathlib
import sys

import numpy as np
import torch

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import bench
from swarmrl_b200.device import DeviceSwarm
from swarmrl_b200.domain import DomainSwarm


def ev():
    return torch.cuda.Event(enable_timing=True)


n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
L = float(np.sqrt(n * np.pi / 0.10))
pos, th = bench.initial_state(n, L, 42)
gt = bench.stokes(1.0)[0]
for skin in (0.6, 1.0, 2.0):
    d = DeviceSwarm(2, (L,) * 3, True, 1e-3, bench.KT, 42, bench.species_table(), bench.wca_table(), skin=skin)
    d.set_state(pos, np.zeros(n, dtype=np.int32), theta=th)
    d.f_swim[:n] = 5.0 * gt
    d.run(100, 0)
    st0 = d.stats()
    a, b = ev(), ev()
    a.record()
    d.run(300, 100)
    b.record()
    torch.cuda.synchronize()
    st = d.stats()
    print(json.dumps({"n": n, "skin": skin, "us_per_step": a.elapsed_time(b) / 300 * 1e3, "rebuilds_per_100": (st.rebuilds - st0.rebuilds) / 3,
                      "max_nbr": st.max_neighbors_seen, "persistent": st.persistent}))
    # launches of 1 step with a forced rebuild each (set_max_neighbors invalidates the list)
    a, b = ev(), ev()
    a.record()
    for k in range(10):
        d.set_max_neighbors(st.capacity)
        d.run(1, 400 + k)
    b.record()
    torch.cuda.synchronize()
    print(json.dumps({"n": n, "skin": skin, "us_per_rebuild_plus_step_launch": a.elapsed_time(b) / 10 * 1e3}))
    d.close()

dd = DomainSwarm((L, L), 1e-3, bench.KT, 42, bench.species_table(), bench.wca_table(), exchange_every=20, skin=1.0)
dd.scatter(pos[:, :2], th)
dd.run(20, 0)
torch.cuda.synchronize()
a, b = ev(), ev()
a.record()
for _ in range(10):
    dd.exchange()
b.record()
torch.cuda.synchronize()
dd.sync()
print(json.dumps({"n": n, "us_per_exchange_only": a.elapsed_time(b) / 10 * 1e3, "n_local": dd.n}))

# ---- GPU-side timeline of one exchange period (events between the phases; one stream) ---------------
import ctypes as C
from swarmrl_b200 import _lib

names = ["pack", "transport", "merge", "bind", "run20"]
acc = {k: 0.0 for k in names}
reps = 10
torch.cuda.synchronize()
for r in range(reps):
    e = [ev() for _ in range(6)]
    b_, g = dd.bufs, dd.geom
    old, new = dd.sets[dd.cur], dd.sets[dd.cur ^ 1]
    c_old, c_new = dd.counts[dd.cur], dd.counts[dd.cur ^ 1]
    a_old, a_new = dd._dd(old), dd._dd(new)
    e[0].record()
    _lib.check(dd.lib.swarm_dd_pack(dd.h, C.byref(g), dd.cap, c_old.data_ptr(), C.byref(a_old), b_["send_l"].data_ptr(),
                                    b_["send_r"].data_ptr(), dd._stream()))
    e[1].record()
    dd.transport.exchange(b_["send_l"], b_["send_r"], b_["recv_l"], b_["recv_r"])
    e[2].record()
    _lib.check(dd.lib.swarm_dd_merge(dd.h, C.byref(g), dd.cap, c_old.data_ptr(), C.byref(a_old), b_["recv_l"].data_ptr(),
                                     b_["recv_r"].data_ptr(), dd.cap, C.byref(a_new), c_new.data_ptr(), dd._stream()))
    e[3].record()
    dd.cur ^= 1
    dd._bind()
    e[4].record()
    _lib.check(dd.lib.swarm_run(dd.h, 20, 1000 + 20 * r, dd._stream()))
    e[5].record()
    torch.cuda.synchronize()
    for k, nm in enumerate(names):
        acc[nm] += e[k].elapsed_time(e[k + 1]) * 1e3 / reps
dd.sync()
print(json.dumps({"n": n, "exchange_period_us": acc, "rebuilds_total": int(dd.sw.stats().rebuilds)}))
