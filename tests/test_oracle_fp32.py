"""
float32 evaluation mode of the oracle: the reference computes its observables and
tasks with JAX in float32; the golden fixtures (tests/golden) were generated under an fp64 NumPy shim and the kernels
compute in fp64.  These tests replay the SAME fixtures with the reference's float32 rounding points
(oracle.{field_delta_f32, vision_cones_f32, pair_field_f32}) and state, per observable, how far an fp64 answer may
lie from what real JAX returns:

  ConcentrationField / GradientSensing   |fp64 - fp32| <= 8 eps32 * scale * (|c0| + |c1| d_max)   (cancellation of two O(1) decays)
  SubdividedVisionCones                  |fp64 - fp32| <= 1e-5 per cone sum, EXCEPT where a particle sits within fp32
                                         rounding of a rim or of the vision range -- such entries differ by that
                                         particle's whole weight and must be rare (< 0.5 % of the entries)
  ParticleSensing / SpeciesSearch        |fp64 - fp32| <= 1e-5 * scale * max(1, M |c0|) per field difference

eps32 = 2^-24.  The same bounds are asserted for the CUDA kernels in tests/test_gpu_obs.py::test_fp32_reference_bounds.
"""
import numpy as np
import pytest

from oracle import oracle

EPS32 = 2.0 ** -24


def test_field_delta_fp32_bound(golden):
    for name, clip in (("concentration_field.npz", False), ("gradient_sensing.npz", True)):
        g = golden(name)
        traj, box, src, scale = g["traj"], g["box"], g["src"], float(g["scale"])
        sel = g["types"] == int(g["particle_type"])
        ref = g["obs"][..., 0] if "obs" in g.files else g["rewards"]
        dmax = 2.0
        tol = 8 * EPS32 * scale * (1.0 + 1.0 * dmax)
        worst = 0.0
        for s in range(1, traj.shape[0]):
            v32 = oracle.field_delta_f32(traj[s][sel] / box, traj[s - 1][sel] / box, src / box, 1, 1.0, -1.0, scale, clip)
            assert v32.dtype == np.float32
            worst = max(worst, float(np.max(np.abs(v32.astype(np.float64) - ref[s - 1]))))
        assert worst <= tol, (name, worst, tol)
        assert worst > 0.0            # the two precisions do differ: the bound is not vacuous


@pytest.mark.parametrize("case", ["ref_unit", "rand64", "rand200_narrow", "rand150_wide"])
def test_vision_cones_fp32_bound(golden, case):
    g = golden("vision_cones.npz")
    out32 = oracle.vision_cones_f32(g[f"{case}/pos"], g[f"{case}/dirs"], g[f"{case}/types"], int(g[f"{case}/particle_type"]),
                                    float(g[f"{case}/vision_range"]), float(g[f"{case}/half_angle"]), int(g[f"{case}/n_cones"]),
                                    g[f"{case}/radii"], [int(t) for t in g[f"{case}/detected_types"]])
    ref = g[f"{case}/obs"]
    assert out32.shape == ref.shape and out32.dtype == np.float32
    diff = np.abs(out32.astype(np.float64) - ref)
    flipped = diff > 1e-5
    assert flipped.mean() < 0.005, (case, int(flipped.sum()), diff.max())
    assert np.all(diff[flipped] <= 1.0 + 1e-5)      # a flip moves one particle's weight (<= 1) between cones / in or out
    if case == "ref_unit":
        assert np.array_equal(out32[0], np.array([[1.0, 0.0], [0.8, 0.75], [0.0, 0.0]], dtype=np.float32))


PAIR_CASES = {"ps_same": (0, 0, 2, 1.0, 4.0, 100.0, 0), "ps_other": (0, 1, 3, 0.01, 0.05, 100.0, 0),
              "ss_same": (1, 1, 2, 1.0, 4.0, 100.0, 1), "ss_avoid": (1, 0, 3, 0.01, 0.05, 10.0, -1)}


@pytest.mark.parametrize("name", sorted(PAIR_CASES))
def test_pair_field_fp32_bound(golden, name):
    g = golden("pair_field.npz")
    pt, st, kind, c0, c1, scale, clip = PAIR_CASES[name]
    traj, types, box = g["traj"], g["types"], g["box"]
    M = int(np.sum(types == st))
    amp = c0 if kind == 2 else c0 / c1      # largest value one sensed particle contributes
    tol = 1e-5 * scale * max(1.0, M * amp)
    h = oracle.pair_field_f32(traj[0], types, pt, st, box, kind, c0, c1)
    for s in range(1, traj.shape[0]):
        f = oracle.pair_field_f32(traj[s], types, pt, st, box, kind, c0, c1)
        d = (f - h).astype(np.float32)
        h = f
        d = np.clip(d, 0, None) if clip > 0 else (np.clip(d, None, 0) if clip < 0 else d)
        assert np.max(np.abs(np.float32(scale) * d - g[name][s - 1])) <= tol, (name, s)
