"""Smallest driver of the Engine-API slice (bench.engine_record) for an ncu launch list."""
import sys, pathlib, types
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import bench
rec = bench.engine_record(types.SimpleNamespace(skin=2.0), 0, int(sys.argv[1]) if len(sys.argv) > 1 else 4, 2)
print({k: rec[k] for k in ("ms_per_step", "gpu_launches_per_slice")})
