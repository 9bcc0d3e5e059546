"""Device time of every backend call of one Engine-API slice and the idle gap in front of it (CUDA events around the
DeviceSwarm methods SwarmEngine.integrate reaches on the device path)."""
import sys, pathlib, types, collections
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import torch, bench
from swarmrl_b200.device import DeviceSwarm
log = []
def wrap(name):
    f = getattr(DeviceSwarm, name)
    def g(self, *a, **k):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); r = f(self, *a, **k); e1.record(); log.append((name, e0, e1)); return r
    setattr(DeviceSwarm, name, g)
for nm in ("obs_vision", "policy_forward", "run", "obs_field", "obs_field_mean", "reduce_mean"): wrap(nm)
rec = bench.engine_record(types.SimpleNamespace(skin=2.0), 0, 60, 3)
torch.cuda.synchronize()
print({k: rec[k] for k in ("ms_per_step", "gpu_launches_per_slice")})
# mean device time of the LAST 60 calls of every kind (the timed integrate(60)) and the idle gap in front of each
byname = collections.defaultdict(list)
for k, (name, e0, e1) in enumerate(log):
    byname[name].append((e0.elapsed_time(e1), log[k - 1][2].elapsed_time(e0) if k else 0.0))
for name, v in byname.items():
    v = v[-60:]
    print(f"{name:16s} device {sum(a for a, _ in v) / len(v) * 1e3:8.1f} us   idle gap before it {sum(b for _, b in v) / len(v) * 1e3:8.1f} us   ({len(v)} calls)")
# host enqueue rate of one integrate(K) call (until the engine's final counter check starts waiting for the device)
import time
import swarmrl_b200 as sw
marks = {}
_chk = sw.SwarmEngine._check_deferred_runs
def chk(self):
    marks["enqueued"] = time.perf_counter(); return _chk(self)
sw.SwarmEngine._check_deferred_runs = chk
for nm in ("obs_vision", "policy_forward", "run", "obs_field", "reduce_mean"):
    pass

import numpy as np
_int = sw.SwarmEngine.integrate
calls = []
def timed(self, n, ff=None):
    torch.cuda.synchronize(); t0 = time.perf_counter(); r = _int(self, n, ff); t1 = time.perf_counter()
    calls.append((n, marks.get("enqueued", t1) - t0, t1 - t0)); return r
sw.SwarmEngine.integrate = timed
log.clear()
rec = bench.engine_record(types.SimpleNamespace(skin=2.0), 0, 60, 3)
print({k: rec[k] for k in ("ms_per_step", "gpu_launches_per_slice")}, "flushed", rec["flushed_per_call"]["ms_per_step"])
for n, enq, tot in calls:
    if n >= 10: print(f"integrate({n}): host enqueue {enq / n * 1e3:.3f} ms/slice, until return {tot / n * 1e3:.3f} ms/slice")
