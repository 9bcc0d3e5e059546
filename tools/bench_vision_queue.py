"""Vision-cone pipeline at C4 (100k observers, range 10, 5 cones) after some clustering: k_vis2_cones (every iteration
evaluates) against k_vis2_cones_q (screen, queue, evaluate on full queues), and both against the serial kernel."""
import json, os, pathlib, sys
import numpy as np, torch
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import bench

ro = bench.Rollout(0, 42, 2.0)
d = ro.d
for k in range(int(sys.argv[1]) if len(sys.argv) > 1 else 5):   # a few slices so that the active swarm has started to cluster
    d.run(100, k * 100)
ref = None
for mode in ("serial", "noqueue", "queue"):
    for v in ("SWARM_B200_VIS_SERIAL", "SWARM_B200_VIS_NOQUEUE"):
        os.environ.pop(v, None)
    if mode == "serial":
        os.environ["SWARM_B200_VIS_SERIAL"] = "1"
    elif mode == "noqueue":
        os.environ["SWARM_B200_VIS_NOQUEUE"] = "1"
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
    print(json.dumps({"mode": mode, "us_per_call": a.elapsed_time(b) / 20 * 1e3, "max_abs_diff_vs_serial": float(np.max(np.abs(o - ref))),
                      "sum": float(o.sum())}))
