"""Smallest driver of k_bd2d_cluster for ncu: BASELINE config 2 (1,000 colloids, L = 125), a few launches of `steps` steps."""
import sys, pathlib
import numpy as np, torch
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent)); sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent / "tests"))
from helpers import KT_300, jittered_lattice, species_for, wca_for
from swarmrl_b200.device import DeviceSwarm
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
launches = int(sys.argv[2]) if len(sys.argv) > 2 else 4
skin = float(sys.argv[3]) if len(sys.argv) > 3 else 0.4
rng = np.random.default_rng(1234)
n, L = 1000, 125.0
sp, wca = species_for([1.0]), wca_for([1.0])
d = DeviceSwarm(2, (L,) * 3, True, 1e-3, KT_300, 42, sp, wca, skin=skin)
d.set_state(jittered_lattice(n, L, 2, rng), np.zeros(n, dtype=np.int32), theta=rng.uniform(0, 2 * np.pi, n))
d.set_actions_host(np.full(n, 5.0 * sp[0]["gamma_t"][0]), np.zeros((n, 3)))
ev = []
for k in range(launches):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); d.run(steps, k * steps); b.record(); ev.append((a, b))
st = d.stats()
print("ok", st.steps_done, "rebuilds", st.rebuilds, "overflow", st.overflow, "us/step", [round(a.elapsed_time(b) / steps * 1e3, 3) for a, b in ev])
