"""
BASELINE config 1 end to end on the device: SwarmEngine.integrate (espresso.py:1251-1308) with a chemotaxis agent
(ConcentrationField observable, GradientSensing task, 4 actions of test_rl_trainers.py:36-46) -- the drop-in
surface the reference's trainers use (engine.colloids, engine.integrate(n, force_model), engine.finalize()).
The device path (no host round trip) must reproduce the reference-shaped object path (lists of Colloid / Action)
exactly, and the fused policy kernel must drive a rollout.
"""
import numpy as np
import pytest

import swarmrl_b200 as sw
from swarmrl_b200 import decay
from swarmrl_b200.observables import ConcentrationField
from swarmrl_b200.tasks import GradientSensing
from swarmrl_b200.units import UnitRegistry

pytestmark = pytest.mark.gpu


def _engine(tmp_path, seed=42):
    ureg = UnitRegistry()
    dt = ureg.Quantity(0.1, "second")
    md = sw.MDParams(ureg=ureg, fluid_dyn_viscosity=ureg.Quantity(8.9e-4, "pascal * second"),
                     WCA_epsilon=ureg.Quantity(297.0, "kelvin") * ureg.boltzmann_constant,
                     temperature=ureg.Quantity(300.0, "kelvin"), box_length=ureg.Quantity(3 * [1000.0], "micrometer"),
                     time_step=dt, time_slice=dt * 10, write_interval=dt * 10)
    e = sw.SwarmEngine(md_params=md, n_dims=2, seed=seed, out_folder=tmp_path, write_chunk_size=100)
    e.add_colloids(10, ureg.Quantity(2.14, "micrometer"), ureg.Quantity(np.array([500.0, 500.0, 0.0]), "micrometer"),
                   ureg.Quantity(400.0, "micrometer"), type_colloid=0)
    return e


def _agent(policy):
    box = np.array([1000.0, 1000.0, 1000.0])
    src = np.array([500.0, 500.0, 0.0])
    actions = {"Translate": sw.Action(force=10.0), "RotateCW": sw.Action(torque=np.array([0.0, 0.0, 10.0])),
               "RotateCCW": sw.Action(torque=np.array([0.0, 0.0, -10.0])), "Nothing": sw.Action()}
    return sw.DeviceAgent(0, policy, GradientSensing(source=src, decay_function=decay.Linear(1.0, -1.0), box_length=box,
                                                     reward_scale_factor=10, particle_type=0),
                          ConcentrationField(source=src, decay_fn=decay.Linear(1.0, -1.0), box_length=box, scale_factor=10000,
                                             particle_type=0), actions)


class ObjectPathOnly(sw.ForceFunction):
    """forces the reference-shaped protocol: calc_action(list of Colloid) -> list of Action"""
    has_device_path = False


def _rollout(tmp_path, ff_cls, policy, slices=5):
    e = _engine(tmp_path)
    agent = _agent(policy)
    ff = ff_cls({"0": agent})
    agent.reset_agent(e.colloids)
    e.integrate(slices, ff)
    e.finalize()
    data = e.get_particle_data()
    return e, agent, data


def test_device_path_equals_object_path(tmp_path):
    e1, a1, d1 = _rollout(tmp_path / "dev", sw.ForceFunction, sw.CounterPolicy(4, seed=7))
    e2, a2, d2 = _rollout(tmp_path / "obj", ObjectPathOnly, sw.CounterPolicy(4, seed=7))
    assert (e1.step_idx, e1.slice_idx) == (e2.step_idx, e2.slice_idx) == (50, 5)
    for k in ("Unwrapped_Positions", "Directors", "Velocities"):
        assert np.allclose(d1[k], d2[k], rtol=0, atol=1e-12), k
    r1 = np.array([np.asarray(r.cpu() if hasattr(r, "cpu") else r) for r in a1.trajectory.rewards])
    r2 = np.array([np.asarray(r.cpu() if hasattr(r, "cpu") else r) for r in a2.trajectory.rewards])
    assert r1.shape == r2.shape == (5, 10) and np.allclose(r1, r2, atol=1e-12)
    assert np.all(r1 >= 0) and np.isfinite(r1).all()
    z = d1["Unwrapped_Positions"][:, 2]
    assert np.all(z == 0.0)                                   # 2-D: planar (test_espresso_2d.py:59-82)


def test_reward_mean_stays_on_the_device_in_one_launch(tmp_path):
    """DeviceAgent(reward_mean=True) with GradientSensing: reward and its mean come from one launch
    (swarm_obs_field_mean); the means equal the host mean of the per-agent rewards the trainer would take
    (trainers/trainer.py:92-95)."""
    e = _engine(tmp_path)
    agent = _agent(sw.CounterPolicy(4, seed=7))
    agent.reward_mean = True
    ff = sw.ForceFunction({"0": agent})
    agent.reset_agent(e.colloids)
    e.integrate(5, ff)
    e.finalize()
    assert len(agent.reward_means) == len(agent.trajectory.rewards) == 5
    assert agent.task.last_mean is not None          # the fused launch was taken
    for red, r in zip(agent.reward_means, agent.trajectory.rewards):
        red, r = red.cpu().numpy(), r.cpu().numpy()
        assert abs(red[0] - r.sum()) <= 1e-12 * max(1.0, abs(r.sum())) and abs(red[1] - r.mean()) <= 1e-12 * max(1.0, abs(r.mean()))


def test_fused_policy_drives_a_rollout(tmp_path):
    rng = np.random.default_rng(0)
    P = dict(w1=rng.normal(size=(1, 128)) * 0.5, b1=rng.normal(size=128) * 0.1, wa=rng.normal(size=(128, 4)) * 0.2,
             ba=np.zeros(4), wv=rng.normal(size=128) * 0.2, bv=np.zeros(1))
    e, agent, data = _rollout(tmp_path, sw.ForceFunction, sw.FusedActorCritic(P), slices=6)
    assert e.slice_idx == 6 and agent.policy.call == 6
    acts = np.array([a.cpu().numpy() for a in agent.trajectory.actions])
    logp = np.array([l.cpu().numpy() for l in agent.trajectory.log_probs])
    assert acts.shape == (6, 10) and acts.min() >= 0 and acts.max() <= 3
    assert np.all(logp <= 0) and np.isfinite(logp).all()
    assert np.isfinite(data["Unwrapped_Positions"]).all()
    # the last action written by the kernel is what the engine holds
    f = e._backend.f_swim[:10].cpu().numpy()
    assert np.array_equal(f != 0.0, acts[-1] == 0)


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_chemotaxis_reward_curve_device_engine_vs_oracle_engine(tmp_path, seed):
    """north_star's third parity check (reward curves with noise ON): the same rollout driven once by the CUDA engine and
    once by the CPU oracle behind the same SwarmEngine host code (tests/oracle_backend.py), same counter policy, same
    thermostat seed.  Because both draw the same Philox noise the curves must agree far inside 2 sigma: to 1e-7."""
    from oracle_backend import OracleBackend

    class CpuEngine(sw.SwarmEngine):
        backend_factory = staticmethod(lambda **kw: OracleBackend(**kw))

    class ListPath(ObjectPathOnly):
        """hands the agents plain lists of Colloid, as the reference does, so that observable and task kernels run on
        a scratch copy whichever backend integrates (the oracle backend has no device arrays to view)"""

        def calc_reward(self, colloids, external_reward=0.0):
            return super().calc_reward(colloids.as_colloids() if hasattr(colloids, "as_colloids") else colloids, external_reward)

    def rollout(cls, sub):
        ureg = UnitRegistry()
        dt = ureg.Quantity(0.1, "second")
        md = sw.MDParams(ureg=ureg, fluid_dyn_viscosity=ureg.Quantity(8.9e-4, "pascal * second"),
                         WCA_epsilon=ureg.Quantity(297.0, "kelvin") * ureg.boltzmann_constant,
                         temperature=ureg.Quantity(300.0, "kelvin"), box_length=ureg.Quantity(3 * [1000.0], "micrometer"),
                         time_step=dt, time_slice=dt * 10, write_interval=dt * 10)
        e = cls(md_params=md, n_dims=2, seed=seed, out_folder=tmp_path / sub, write_chunk_size=100)
        e.add_colloids(10, ureg.Quantity(2.14, "micrometer"), ureg.Quantity(np.array([500.0, 500.0, 0.0]), "micrometer"),
                       ureg.Quantity(400.0, "micrometer"), type_colloid=0)
        agent = _agent(sw.CounterPolicy(4, seed=7))
        ff = ListPath({"0": agent})
        agent.reset_agent(e.colloids.as_colloids())
        for _ in range(3):                       # 3 episodes x 5 slices, as the reference's trainer test loops
            e.integrate(5, ff)
        e.finalize()
        r = np.array([np.asarray(x.cpu() if hasattr(x, "cpu") else x) for x in agent.trajectory.rewards])
        return r, e.get_particle_data()["Unwrapped_Positions"]

    r_dev, p_dev = rollout(sw.SwarmEngine, "dev")
    r_cpu, p_cpu = rollout(CpuEngine, "cpu")
    assert r_dev.shape == r_cpu.shape == (15, 10)
    assert np.allclose(p_dev, p_cpu, rtol=1e-9, atol=1e-9)
    assert np.allclose(r_dev, r_cpu, rtol=1e-7, atol=1e-7)
    assert r_dev.sum() > 0
