"""Vision-cone pipeline at C4 (100k observers, range 10, 5 cones): time per call vs lanes per observer."""
import json, os, pathlib, sys
import numpy as np, torch
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import bench

ro = bench.Rollout(0, 42, 2.0)
d = ro.d
ref = None
for gs in ("serial", "2", "4", "8", "16", "32"):
    os.environ.pop("SWARM_B200_VIS_SERIAL", None)
    if gs == "serial":
        os.environ["SWARM_B200_VIS_SERIAL"] = "1"
    else:
        os.environ["SWARM_B200_VIS_GROUP"] = gs
    f = lambda: d.obs_vision(ro.row_of, ro.rows, bench.VISION["range"], bench.VISION["half_angle"], bench.VISION["n_cones"], ro.radii, ro.type_col, 1)
    for _ in range(3):
        out = f()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        out = f()
    b.record(); torch.cuda.synchronize()
    o = out.cpu().numpy()
    if ref is None:
        ref = o
    print(json.dumps({"lanes_per_observer": gs, "us_per_call": a.elapsed_time(b) / 20 * 1e3, "max_abs_diff_vs_serial": float(np.max(np.abs(o - ref)))}))
