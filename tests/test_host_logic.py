"""
Host logic of SwarmEngine on CPU (oracle-backed stand-in for the device, tests/oracle_backend.py):
the integrate loop's bookkeeping, manage_forces, setup errors -- the known answers of the
reference's own engine tests.

  loop counters            CI/espresso_tests/unit_tests/test_integration.py:107-160
  v = f e/gamma, x = x0+vt CI/espresso_tests/unit_tests/test_espresso.py:70-118,140-147
  2-D new_direction        CI/espresso_tests/unit_tests/test_espresso_2d.py:84-92
  placement / type clash   CI/espresso_tests/unit_tests/test_add_colloid_on_point.py:55-99
"""
import numpy as np
import pytest

import swarmrl_b200 as sw
from oracle_backend import OracleBackend
from swarmrl_b200.units import UnitRegistry


class CpuEngine(sw.SwarmEngine):
    backend_factory = staticmethod(lambda **kw: OracleBackend(**kw))


def _engine(tmp_path, ureg, n_dims=2, steps=(5, 9), box=10, temperature=293, seed=42, **kw):
    time_step = ureg.Quantity(0.1, "second")
    md = sw.MDParams(ureg=ureg, fluid_dyn_viscosity=ureg.Quantity(8.9e-4, "pascal * second"),
                     WCA_epsilon=ureg.Quantity(293, "kelvin") * ureg.boltzmann_constant,
                     temperature=ureg.Quantity(temperature, "kelvin"),
                     box_length=ureg.Quantity(3 * [box], "micrometer"), time_step=time_step,
                     time_slice=time_step * steps[0], write_interval=time_step * steps[1])
    return CpuEngine(md_params=md, n_dims=n_dims, seed=seed, out_folder=tmp_path, write_chunk_size=10, **kw)


COUNTER_CASES = {
    (5, 9): [(2, dict(step=10, slice=2, write=2, time=1.0, held=2)), (3, dict(step=25, slice=5, write=3, time=2.5, held=3))],
    (7, 3): [(4, dict(step=28, slice=4, write=10, time=2.8, held=0)), (2, dict(step=42, slice=6, write=14, time=4.2, held=4))],
    (2, 2): [(4, dict(step=8, slice=4, write=4, time=0.8, held=4)), (2, dict(step=12, slice=6, write=6, time=1.2, held=6))],
}


@pytest.mark.parametrize("steps", list(COUNTER_CASES))
def test_integrate_counters_match_reference_known_answers(tmp_path, steps):
    ureg = UnitRegistry()
    e = _engine(tmp_path, ureg, steps=steps)
    e.add_colloids(1, ureg.Quantity(0.2, "micrometer"), ureg.Quantity(np.array([5, 5, 0]), "micrometer"),
                   ureg.Quantity(1, "micrometer"), type_colloid=0)
    ff = sw.ForceFunction(agents={"0": sw.ConstForce(1)})
    assert e.system.time == 0
    assert (e.params.steps_per_slice, e.params.steps_per_write_interval) == steps
    for n_slices, want in COUNTER_CASES[steps]:
        e.integrate(n_slices, ff)
        assert e.step_idx == want["step"] and e.slice_idx == want["slice"] and e.write_idx == want["write"]
        assert abs(e.system.time - want["time"]) < 1e-12
        assert len(e.traj_holder["Times"]) == want["held"]
    # chunks never straddle a slice or a write boundary, and tile the run exactly
    runs = e._backend.runs
    assert sum(r[0] for r in runs) == e.step_idx
    assert [r[1] for r in runs] == list(np.cumsum([0] + [r[0] for r in runs[:-1]]))
    for n, s0 in runs:
        assert s0 // steps[0] == (s0 + n - 1) // steps[0] and s0 // steps[1] == (s0 + n - 1) // steps[1]
    e.finalize()
    written = sorted(tmp_path.glob("trajectory*"))
    assert written, "finalize() flushes the trajectory holder"


def test_reward_is_evaluated_after_every_chunk_and_kill_switch_breaks_before_manage_forces(tmp_path):
    ureg = UnitRegistry()
    e = _engine(tmp_path, ureg, steps=(7, 3))
    e.add_colloids(2, ureg.Quantity(0.2, "micrometer"), ureg.Quantity(np.array([5, 5, 0]), "micrometer"),
                   ureg.Quantity(1, "micrometer"), type_colloid=0)

    class Probe(sw.ConstForce):
        def __init__(self):
            super().__init__(1.0)
            self.n_action = self.n_reward = 0
            self.kill_after = None

        def calc_action(self, colloids):
            self.n_action += 1
            return super().calc_action(colloids)

        def calc_reward(self, colloids, external_reward=0.0):
            self.n_reward += 1
            assert all(hasattr(c, a) for c in colloids for a in ("id", "pos", "type", "director"))
            if self.kill_after is not None and self.n_reward >= self.kill_after:
                self.kill_switch = True

    agent = Probe()
    ff = sw.ForceFunction(agents={"0": agent})
    e.integrate(2, ff)  # 14 steps: chunks end at 3,6,7,9,12,14
    assert agent.n_action == 2 and agent.n_reward == 6
    agent.kill_after = agent.n_reward + 1
    e.integrate(3, ff)
    # the agent's switch flips after the chunk 14 -> 15, but the loop tests force_model.kill_switch
    # (espresso.py:1288), which ForceFunction refreshes only inside calc_action (force_fn.py:76,84):
    # the run continues to the slice boundary at 21, manage_forces latches the switch, one more
    # chunk (21 -> 24, the next write boundary) is integrated, and only then the loop breaks.
    assert ff.kill_switch is True
    assert e.step_idx == 24 and e.slice_idx == 4


def test_kinematics_3d_two_types_and_ext_force(tmp_path):
    """test_espresso.py:20-147 at kT = 0."""
    ureg = UnitRegistry()
    time_step = ureg.Quantity(0.01, "second")
    md = sw.MDParams(ureg=ureg, fluid_dyn_viscosity=ureg.Quantity(8.9e-4, "pascal * second"),
                     WCA_epsilon=ureg.Quantity(1e-20, "joule"), temperature=ureg.Quantity(0, "kelvin"),
                     box_length=ureg.Quantity(3 * [1000], "micrometer"), time_step=time_step,
                     time_slice=time_step * 10, write_interval=time_step * 10)
    e = CpuEngine(md_params=md, n_dims=3, seed=42, out_folder=tmp_path, write_chunk_size=1000)
    e.add_colloids(3, ureg.Quantity(1, "micrometer"), ureg.Quantity(np.array([500, 500, 500]), "micrometer"),
                   ureg.Quantity(100, "micrometer"), type_colloid=0)
    e.add_colloids(2, ureg.Quantity(2, "micrometer"), ureg.Quantity(np.array([500, 500, 500]), "micrometer"),
                   ureg.Quantity(100, "micrometer"), type_colloid=1)
    assert len(e.colloids) == 5
    part_data = e.get_particle_data()
    assert set(part_data) == {"Id", "Type", "Unwrapped_Positions", "Velocities", "Directors"}
    direc = np.array([1 / np.sqrt(3)] * 3)
    ff = sw.ForceFunction({"0": sw.ToConstDirection(direc), "1": sw.ToConstDirection(direc)})
    e.integrate(1, ff)
    assert np.allclose(e.get_particle_data()["Directors"], direc[None, :], atol=1e-12)
    f = 1.234
    ff = sw.ForceFunction({"0": sw.ConstForce(f), "1": sw.ConstForce(f)})
    e.integrate(1, ff)
    old = e.get_particle_data()
    e.integrate(10, ff)
    new = e.get_particle_data()
    gam = np.array([e.get_friction_coefficients(int(t))[0] for t in new["Type"]])
    v = f * new["Directors"] / gam[:, None]
    assert np.allclose(new["Velocities"], v, rtol=1e-12)
    t = 10 * 10 * 0.01
    assert np.allclose(new["Unwrapped_Positions"], old["Unwrapped_Positions"] + t * v, rtol=2e-6)
    # ext force on one type only (test_espresso.py:140-147)
    e.add_const_force_to_colloids(ureg.Quantity(np.array([0.5, 0, 0]), "sim_force"), 1)
    F = np.stack(e._backend._o.ext_force, axis=1)
    assert np.array_equal(F[new["Type"] == 1], np.tile([0.5, 0, 0], (2, 1))) and not F[new["Type"] == 0].any()
    with pytest.raises(ValueError):
        e.add_const_force_to_colloids(ureg.Quantity(np.array([0.5, 0, 0]), "sim_force"), 5)
    with pytest.raises(RuntimeError):
        e.add_colloids(1, ureg.Quantity(1, "micrometer"), ureg.Quantity(np.array([500, 500, 500]), "micrometer"),
                       ureg.Quantity(100, "micrometer"), type_colloid=0)


def test_2d_new_direction_and_planarity(tmp_path):
    ureg = UnitRegistry()
    e = _engine(tmp_path, ureg, n_dims=2, steps=(5, 5), box=100, temperature=300)
    e.add_colloids(10, ureg.Quantity(1, "micrometer"), ureg.Quantity(np.array([50, 50, 0]), "micrometer"),
                   ureg.Quantity(40, "micrometer"), type_colloid=0)
    d0 = e.get_particle_data()
    assert np.all(d0["Unwrapped_Positions"][:, 2] == 0) and np.all(d0["Directors"][:, 2] == 0)
    e.integrate(2, sw.ForceFunction({"0": sw.ConstForce(0.1)}))
    d1 = e.get_particle_data()
    assert np.all(d1["Unwrapped_Positions"][:, 2] == 0) and np.all(np.abs(d1["Directors"][:, 2]) == 0)
    assert not np.allclose(d0["Directors"], d1["Directors"])  # rotational noise acted
    new_dir = np.array([1 / np.sqrt(2), -1 / np.sqrt(2), 0.0])
    e.manage_forces(sw.ForceFunction({"0": sw.ToConstDirection(new_dir)}))
    assert np.allclose(e.get_particle_data()["Directors"], new_dir[None, :], atol=1e-12)


def test_add_colloid_on_point_errors_and_exact_placement(tmp_path):
    ureg = UnitRegistry()
    e = _engine(tmp_path, ureg, n_dims=2, steps=(2, 2), box=100, temperature=0)
    pts = np.array([[10.0, 10, 0], [20.0, 30, 0], [70.0, 15, 0]])
    for p in pts:
        e.add_colloid_on_point(ureg.Quantity(1, "micrometer"), ureg.Quantity(p, "micrometer"), type_colloid=3)
    with pytest.raises(ValueError):  # type 3 already registered with radius 1
        e.add_colloid_on_point(ureg.Quantity(2, "micrometer"), ureg.Quantity(pts[0], "micrometer"), type_colloid=3)
    with pytest.raises(ValueError):  # director out of plane in 2-D
        e.add_colloid_on_point(ureg.Quantity(1, "micrometer"), ureg.Quantity(pts[0], "micrometer"),
                               init_direction=np.array([0, 0, 1.0]), type_colloid=3)
    e.integrate(1, sw.ForceFunction({"3": sw.ConstForce(0.0)}))
    assert np.array_equal(e.get_particle_data()["Unwrapped_Positions"], pts)
    with pytest.raises(ValueError):
        sw.SwarmEngine(md_params=e.params, n_dims=4)


def test_non_integer_step_ratio_is_rejected(tmp_path):
    ureg = UnitRegistry()
    md = sw.MDParams(ureg=ureg, time_step=ureg.Quantity(0.1, "second"), time_slice=ureg.Quantity(0.25, "second"),
                     write_interval=ureg.Quantity(1, "second"))
    with pytest.raises(ValueError):
        CpuEngine(md_params=md, n_dims=2, out_folder=tmp_path)


def test_force_function_routing_matches_reference(golden):
    """force_fn.py:52-86, golden produced by the reference's own ForceFunction."""
    g = golden("force_fn_routing.npz")
    types = g["types"]

    class PerType(sw.ClassicalAgent):
        def __init__(self, t):
            self.t = t

        def calc_action(self, colloids):
            return [sw.Action(force=100.0 * self.t + i) for i, _ in enumerate([c for c in colloids if c.type == self.t])]

    cols = [sw.Colloid(pos=np.zeros(3), director=np.array([1.0, 0, 0]), id=i, type=int(t)) for i, t in enumerate(types)]
    acts = sw.ForceFunction({"0": PerType(0), "2": PerType(2)}).calc_action(cols)
    assert np.array_equal([a.force for a in acts], g["forces"])


def test_add_colloids_rng_draw_order(tmp_path):
    """espresso.py:94-97,521-533: per colloid r, angle, then orientation angle from default_rng(seed)."""
    ureg = UnitRegistry()
    e = _engine(tmp_path, ureg, n_dims=2, steps=(2, 2), box=1000, seed=42)
    e.add_colloids(3, ureg.Quantity(2.14, "micrometer"), ureg.Quantity(np.array([500, 500, 0]), "micrometer"),
                   ureg.Quantity(400, "micrometer"), type_colloid=0)
    rng = np.random.default_rng(42)
    for i in range(3):
        r = 400 * np.sqrt(rng.random()); th = 2 * np.pi * rng.random(); ph = 2 * np.pi * rng.random()
        d = e.get_particle_data()
        assert np.allclose(d["Unwrapped_Positions"][i], [500 + r * np.cos(th), 500 + r * np.sin(th), 0])
        assert np.allclose(d["Directors"][i], [np.cos(ph), np.sin(ph), 0], atol=1e-15)


def test_oracle_policy_forward_definitions():
    """The oracle's restatement of the reference's action path (flax_network.py:180-192, gumbel_distribution.py:37-40):
    greedy = argmax(logits), log_prob = log(softmax + 1e-8)[idx], value = critic head; sampling reproducible per call."""
    from oracle.oracle import policy_forward

    rng = np.random.default_rng(2)
    P = dict(w1=rng.normal(size=(3, 16)), b1=rng.normal(size=16), wa=rng.normal(size=(16, 4)), ba=rng.normal(size=4),
             wv=rng.normal(size=16), bv=rng.normal(size=1))
    obs = rng.normal(size=(50, 3))
    idx, logp, val, lg = policy_forward(obs, P)
    h = np.maximum(obs @ P["w1"] + P["b1"], 0)
    assert np.allclose(lg, h @ P["wa"] + P["ba"], atol=1e-4) and np.allclose(val, h @ P["wv"] + P["bv"], atol=1e-4)
    assert np.array_equal(idx, np.argmax(lg, axis=1))
    sm = np.exp(lg - lg.max(1, keepdims=True)); sm /= sm.sum(1, keepdims=True)
    assert np.allclose(logp, np.log(sm + 1e-8)[np.arange(50), idx], atol=1e-5)
    a = policy_forward(obs, P, seed=1, call=7, sample=True)[0]
    assert np.array_equal(a, policy_forward(obs, P, seed=1, call=7, sample=True)[0])
    assert (a != policy_forward(obs, P, seed=1, call=8, sample=True)[0]).any()


def test_langevin_thermostat_setup_and_terminal_velocity(tmp_path):
    """thermostat_type="langevin" (espresso.py:1187-1201 without a fluid): mass and rinertia are mandatory
    (:391-400); v is integrated state and relaxes to f e / gamma_t with time constant m / gamma_t."""
    ureg = UnitRegistry()
    time_step = ureg.Quantity(0.001, "second")
    kw = dict(ureg=ureg, fluid_dyn_viscosity=ureg.Quantity(8.9e-4, "pascal * second"),
              WCA_epsilon=ureg.Quantity(293, "kelvin") * ureg.boltzmann_constant,
              temperature=ureg.Quantity(0, "kelvin"), box_length=ureg.Quantity(3 * [1000], "micrometer"),
              time_step=time_step, time_slice=time_step * 100, write_interval=time_step * 100)
    with pytest.raises(ValueError):
        bad = CpuEngine(md_params=sw.MDParams(thermostat_type="nose-hoover", **kw), n_dims=2, out_folder=tmp_path)
        bad.add_colloids(1, ureg.Quantity(1, "micrometer"), ureg.Quantity(np.array([500, 500, 0]), "micrometer"),
                         ureg.Quantity(100, "micrometer"))
        bad.integrate(1, sw.ForceFunction({"0": sw.ConstForce(1.0)}))
    e = CpuEngine(md_params=sw.MDParams(thermostat_type="langevin", **kw), n_dims=2, seed=3, out_folder=tmp_path)
    place = (ureg.Quantity(1, "micrometer"), ureg.Quantity(np.array([500, 500, 0]), "micrometer"),
             ureg.Quantity(100, "micrometer"))
    with pytest.raises(ValueError, match="particle mass"):
        e.add_colloids(1, *place)
    gt, _ = e_gamma = sw.engine._calc_friction_coefficients(
        float(kw["fluid_dyn_viscosity"].m_as("sim_dyn_viscosity")), 1.0)
    mass = ureg.Quantity(gt * 0.001 / 0.02, "sim_mass")  # m / gamma = 50 time steps
    with pytest.raises(ValueError, match="rotational inertia"):
        e.add_colloids(1, *place, mass=mass)
    rin = ureg.Quantity(e_gamma[1] * 0.001 / 0.02, "sim_rinertia")
    e.add_colloids(4, *place, mass=mass, rinertia=rin)
    f = 2.0 * gt
    ff = sw.ForceFunction({"0": sw.ConstForce(f)})
    e.integrate(1, ff)  # 100 steps = 2 relaxation times
    d = e.get_particle_data()
    speed = np.sum(d["Velocities"] * d["Directors"], axis=1)
    assert np.all(np.abs(speed / 2.0 - (1 - np.exp(-2.0))) < 2e-2)
    e.integrate(20, ff)
    d = e.get_particle_data()
    assert np.allclose(d["Velocities"], 2.0 * d["Directors"], rtol=1e-9, atol=1e-12)
    assert e._backend.thermostat_type == "langevin"
