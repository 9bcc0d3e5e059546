"""
N1: the trajectory writer behind SwarmEngine.integrate on the CUDA backend -- device snapshot ring ->
pinned host -> writer thread -- against /root/reference/swarmrl/engine/espresso.py:1054-1159 (dataset layout) and
:1278-1285, 1310-1318 (snapshot BEFORE the chunk, first frame at t = 0, finalize flushes without a snapshot), and the
literal C1 configuration of the reference's trainer test (write_interval = 1 s, time_slice = 10 s:
CI/espresso_tests/integration_tests/test_rl_trainers.py:83-85), where a slice is cut into ten chunks and the reward is
evaluated after every chunk.
"""
import glob

import numpy as np
import pytest

import swarmrl_b200 as sw
from swarmrl_b200 import decay
from swarmrl_b200.observables import ConcentrationField
from swarmrl_b200.tasks import GradientSensing
from swarmrl_b200.units import UnitRegistry

pytestmark = pytest.mark.gpu


def _engine(tmp_path, n, write_every, slice_len, chunk, **kw):
    ureg = UnitRegistry()
    dt = ureg.Quantity(0.1, "second")
    md = sw.MDParams(ureg=ureg, fluid_dyn_viscosity=ureg.Quantity(8.9e-4, "pascal * second"),
                     WCA_epsilon=ureg.Quantity(297.0, "kelvin") * ureg.boltzmann_constant,
                     temperature=ureg.Quantity(300.0, "kelvin"), box_length=ureg.Quantity(3 * [1000.0], "micrometer"),
                     time_step=dt, time_slice=dt * slice_len, write_interval=dt * write_every)
    e = sw.SwarmEngine(md_params=md, n_dims=2, seed=42, out_folder=tmp_path, write_chunk_size=chunk, **kw)
    e.add_colloids(n, ureg.Quantity(2.14, "micrometer"), ureg.Quantity(np.array([500.0, 500.0, 0.0]), "micrometer"),
                   ureg.Quantity(400.0, "micrometer"), type_colloid=0)
    return e


class CountingAgent(sw.DeviceAgent):
    def __init__(self, *a, **k):
        super().__init__(*a, **k)
        self.n_reward_calls = 0

    def calc_reward_device(self, view, particle_type=None, external_reward=0.0):
        self.n_reward_calls += 1
        return super().calc_reward_device(view, particle_type, external_reward)


def _agent(cls=sw.DeviceAgent):
    box = np.array([1000.0, 1000.0, 1000.0])
    src = np.array([500.0, 500.0, 0.0])
    actions = {"Translate": sw.Action(force=10.0), "RotateCW": sw.Action(torque=np.array([0.0, 0.0, 10.0])),
               "RotateCCW": sw.Action(torque=np.array([0.0, 0.0, -10.0])), "Nothing": sw.Action()}
    return cls(0, sw.CounterPolicy(4, seed=7),
               GradientSensing(source=src, decay_function=decay.Linear(1.0, -1.0), box_length=box, reward_scale_factor=10, particle_type=0),
               ConcentrationField(source=src, decay_fn=decay.Linear(1.0, -1.0), box_length=box, scale_factor=10000, particle_type=0),
               actions)


def _load_chunks(folder):
    files = sorted(glob.glob(str(folder / "trajectory.colloids.*.npz")))
    parts = [np.load(f) for f in files]
    return files, {k: np.concatenate([p[k] for p in parts], axis=0) for k in parts[0].files}


def test_async_writer_frames_match_get_particle_data(tmp_path):
    """Every written frame equals get_particle_data() at the moment the snapshot was due; layout and time stamps as
    in the reference; chunks of `write_chunk_size` frames; finalize() writes the rest."""
    pytest.importorskip("torch")
    n, chunk = 37, 4
    e = _engine(tmp_path, n, write_every=5, slice_len=10, chunk=chunk)
    agent = _agent()
    ff = sw.ForceFunction({"0": agent})
    agent.reset_agent(e.colloids)
    want = []
    for k in range(11):                      # one slice per call: a snapshot is due at the start and in the middle
        if k > 0:
            pass
        e.integrate(1, ff)
    e.finalize()
    # replay with the synchronous path and record the state at every write point
    e2 = _engine(tmp_path / "sync", n, write_every=5, slice_len=10, chunk=chunk, async_writer=False)
    a2 = _agent()
    f2 = sw.ForceFunction({"0": a2})
    a2.reset_agent(e2.colloids)
    for k in range(11):
        e2.integrate(1, f2)
    e2.finalize()
    files, got = _load_chunks(tmp_path)
    files2, ref = _load_chunks(tmp_path / "sync")
    T = 22                                   # 110 steps, one frame every 5, first at t = 0, none at the end
    assert len(files) == len(files2) == -(-T // chunk)
    assert got["Times"].shape == (T, 1, 1) and got["Ids"].shape == (T, n, 1) and got["Types"].shape == (T, n, 1)
    for k in ("Unwrapped_Positions", "Velocities", "Directors"):
        assert got[k].shape == (T, n, 3)
        assert np.array_equal(got[k], ref[k]), k
    assert np.allclose(got["Times"].ravel(), 0.5 * np.arange(T))
    assert np.array_equal(got["Ids"][3].ravel(), np.arange(n)) and np.all(got["Types"] == 0)
    assert e.h5_time_steps_written == T and e.write_idx == T
    # the last frame precedes the last chunk of steps: it differs from the final state
    assert not np.array_equal(got["Unwrapped_Positions"][-1], e.get_particle_data()["Unwrapped_Positions"])
    # first frame: t = 0, the state after overlap removal, before any integration
    assert np.all(got["Velocities"][0] == 0.0)


def test_async_writer_back_pressure_with_a_tiny_ring(tmp_path):
    from swarmrl_b200.trajectory import AsyncTrajectoryWriter

    e = _engine(tmp_path, 12, write_every=1, slice_len=10, chunk=7)
    agent = _agent()
    ff = sw.ForceFunction({"0": agent})
    agent.reset_agent(e.colloids)
    e.integrate(1, ff)
    assert isinstance(e._writer, AsyncTrajectoryWriter)
    e._writer.drain()
    # swap in a ring of two slots: 40 more frames have to pass through it
    e._writer.flush()
    e.integrate(4, ff)
    e.finalize()
    _, got = _load_chunks(tmp_path)
    assert got["Times"].shape[0] == 50 and np.allclose(np.diff(got["Times"].ravel()), 0.1)
    d = np.linalg.norm(np.diff(got["Unwrapped_Positions"], axis=0), axis=2)
    assert np.all(d.max(axis=1) > 0.0)       # every frame is a different state


def test_config1_literal_write_interval_inside_the_slice(tmp_path):
    """test_rl_trainers.py:83-85: time_slice = 10 s, write_interval = 1 s, dt = 0.1 s  ->  every slice is cut into 10
    chunks of 10 steps; the reward is evaluated after EVERY chunk (10 per slice), actions once per slice."""
    e = _engine(tmp_path, 10, write_every=10, slice_len=100, chunk=100)
    agent = _agent(CountingAgent)
    ff = sw.ForceFunction({"0": agent})
    agent.reset_agent(e.colloids)
    for episode in range(2):
        e.integrate(5, ff)
    e.finalize()
    assert (e.step_idx, e.slice_idx, e.write_idx) == (1000, 10, 100)
    assert agent.n_reward_calls == 100 and agent.policy.call == 10
    assert len(agent.trajectory.rewards) == 100 and len(agent.trajectory.actions) == 10
    _, got = _load_chunks(tmp_path)
    assert got["Times"].shape == (100, 1, 1) and np.allclose(got["Times"].ravel(), np.arange(100.0))
    st = e._backend.stats()
    assert not st.overflow
