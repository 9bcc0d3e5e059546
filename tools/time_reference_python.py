"""Build-container only (/root/reference present): wall time of the reference's OWN Python observable / task
code, imported under the NumPy shim of oracle/ref_shim.py (real JAX is not installable offline), at sizes
where its O(N^2) Python loop finishes -- the per-slice host cost a SwarmRL user pays today.
    python tools/time_reference_python.py  ->  one JSON line per size (paste into BASELINE.md)"""
import json, pathlib, sys, time
import numpy as np
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
from oracle import ref_shim

VC = ref_shim.ref("observables.subdivided_vision_cones").SubdividedVisionCones
GS = ref_shim.ref("tasks.searching.gradient_sensing").GradientSensing
Colloid = ref_shim.ref("components.colloid").Colloid
for n in (250, 500, 1000, 2000):
    rng = np.random.default_rng(n)
    L = float(np.sqrt(n * np.pi / 0.10))
    pos = np.zeros((n, 3)); pos[:, :2] = rng.uniform(0, L, (n, 2))
    th = rng.uniform(0, 2 * np.pi, n)
    cols = [Colloid(pos=pos[i], director=np.array([np.cos(th[i]), np.sin(th[i]), 0.0]), id=i, velocity=np.zeros(3), type=0) for i in range(n)]
    vc = VC(vision_range=10.0, vision_half_angle=np.pi / 2, n_cones=5, radii=[1.0] * n, detected_types=[0], particle_type=0)
    gs = GS(source=np.array([L / 2, L / 2, 0]), decay_function=lambda d: 1 - d, box_length=np.array([L, L, L]), reward_scale_factor=10, particle_type=0)
    gs.initialize(cols)
    t0 = time.perf_counter(); vc.compute_observable(cols); t_vis = time.perf_counter() - t0
    t0 = time.perf_counter(); gs(cols); t_rew = time.perf_counter() - t0
    print(json.dumps({"n": n, "vision_cones_s": t_vis, "gradient_sensing_s": t_rew, "us_per_observer_pair": t_vis / n / n * 1e6,
                      "extrapolated_vision_s_at_100k": t_vis * (100000 / n) ** 2}))
