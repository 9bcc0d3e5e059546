"""Why is the first Engine-API record of a process slower than the second?  Host enqueue time and device time of both."""
import sys, pathlib, types, time
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import torch, bench
import swarmrl_b200 as sw
marks = {}
_chk = sw.SwarmEngine._check_deferred_runs
def chk(self):
    marks["enqueued"] = time.perf_counter(); return _chk(self)
sw.SwarmEngine._check_deferred_runs = chk
_int = sw.SwarmEngine.integrate
calls = []
def timed(self, n, ff=None):
    torch.cuda.synchronize(); t0 = time.perf_counter(); r = _int(self, n, ff); t1 = time.perf_counter()
    calls.append((n, marks.get("enqueued", t1) - t0, t1 - t0)); return r
sw.SwarmEngine.integrate = timed
for rep in range(3):
    calls.clear()
    rec = bench.engine_record(types.SimpleNamespace(skin=2.0), 0, 40, 3)
    print(rep, {k: round(rec[k], 4) for k in ("ms_per_step", "gpu_launches_per_slice")}, "tiling", rec["tiling"]["cadence"], rec["tiling"]["rollbacks"], rec["tiling"]["sorts"])
    for n, enq, tot in calls:
        if n >= 10: print(f"   integrate({n}): host enqueue {enq / n * 1e3:.3f} ms/slice, until return {tot / n * 1e3:.3f} ms/slice")
