"""Where the host time of SwarmEngine.integrate goes on the device path (cProfile over 200 slices of the C4 rollout)."""
import cProfile, pstats, sys, pathlib, types
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import bench, torch
args = types.SimpleNamespace(skin=2.0)
pr = cProfile.Profile()
orig = bench.engine_record
# run the record once with the profiler around its integrate calls
import swarmrl_b200 as sw
_int = sw.SwarmEngine.integrate
def wrapped(self, n, ff=None):
    if getattr(self, "_warm", 0) >= 1:
        pr.enable(); r = _int(self, n, ff); torch.cuda.synchronize(); pr.disable(); return r
    self._warm = getattr(self, "_warm", 0) + 1
    return _int(self, n, ff)
sw.SwarmEngine.integrate = wrapped
rec = bench.engine_record(args, 0, 100, 3)
print({k: rec[k] for k in ("ms_per_step", "gpu_launches_per_slice")})
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
