"""
GPU parity of the Langevin thermostat + velocity Verlet path (MDParams.thermostat_type="langevin",
/root/reference/swarmrl/engine/espresso.py:1187-1201) against the CPU oracle (oracle/swarm_oracle.c
orc_lv_steps), through the C-ABI (swarm_set_integrator + swarm_run).  Bar: fp64 trajectories, velocities
and angular velocities within 1e-9 relative; same Philox streams under noise.
"""
import numpy as np
import pytest

from helpers import KT_300, jittered_lattice, make_pair, random_quats, rel_err, set_actions, stokes, wca_for

pytestmark = pytest.mark.gpu


def _species(radii=(1.0,), a=0.02, dt=1e-3, aniso=None):
    table = []
    for t, r in enumerate(radii):
        gt, gr = stokes(r)
        gt3 = (gt,) * 3 if aniso is None else tuple(aniso[0])
        gr3 = (gr,) * 3 if aniso is None else tuple(aniso[1])
        table.append(dict(type=t, gamma_t=gt3, gamma_r=gr3, mass=gt * dt / a, rinertia=(gr * dt / a,) * 3))
    return table


def _pair(n_dims, n, L, kT=0.0, radii=(1.0,), aniso=None, device_kw=None, seed=17):
    rng = np.random.default_rng(seed)
    sp, wca = _species(radii, aniso=aniso), wca_for(list(radii))
    pos = jittered_lattice(n, L, n_dims, rng)
    sidx = (np.arange(n) % len(radii)).astype(np.int32)
    kw = dict(theta=rng.uniform(0, 2 * np.pi, n)) if n_dims == 2 else dict(quat=random_quats(n, rng))
    o, d = make_pair(n_dims, (L,) * 3, True, 1e-3, kT, 42, sp, wca, pos, sidx, device_kw=device_kw, **kw)
    d.set_integrator("langevin")
    return o, d, sp


def _compare(o, d, tol=1e-9, tol_e=1e-9):
    n = o.n
    p, v, e = d.particle_data_host()
    w = d.omega[:, :n].cpu().numpy().T
    assert rel_err(p, o.pos) < tol and rel_err(e, o.directors, 1.0) < tol_e
    assert rel_err(v, o.velocities) < 10 * tol and rel_err(w, o.omegas, 1e-12) < 10 * tol


@pytest.mark.parametrize("persistent", [True, False])
def test_noise_free_langevin_trajectory_2d(persistent):
    o, d, sp = _pair(2, 1000, 125.0, radii=(1.0, 1.5), device_kw=dict(persistent=persistent))
    n = o.n
    gt, gr = stokes(1.0)
    set_actions(o, d, np.full(n, 5.0 * gt), np.tile([0, 0, gr * 4 * np.pi / 180], (n, 1)))
    steps = 600 if persistent else 150
    o.lv_steps(0, steps, mode=2)
    st = d.run_checked(steps, 0)
    assert st.steps_done == steps and not st.overflow and st.persistent == int(persistent)
    assert st.rebuilds >= 1
    _compare(o, d)
    p, v, e = d.particle_data_host()
    assert np.all(p[:, 2] == 0.0) and np.all(v[:, 2] == 0.0) and np.all(e[:, 2] == 0.0)
    assert np.abs(v).max() > 1.0  # the swimmers reached a finite speed: v is state, not a by-product


def test_noise_free_langevin_trajectory_3d_and_anisotropic_friction():
    o, d, _ = _pair(3, 512, 30.0)
    n = o.n
    gt, gr = stokes(1.0)
    rng = np.random.default_rng(5)
    set_actions(o, d, np.full(n, 5.0 * gt), rng.normal(size=(n, 3)) * gr)
    o.lv_steps(0, 300, mode=2)
    st = d.run_checked(300, 0)
    assert st.steps_done == 300
    _compare(o, d)
    for n_dims in (2, 3):
        o, d, _ = _pair(n_dims, 64, 40.0, aniso=((1.0, 2.0, 3.0), (0.5, 0.7, 1.1)))
        set_actions(o, d, np.full(64, 2.0), np.tile([0.3, -0.2, 1.0] if n_dims == 3 else [0, 0, 1.0], (64, 1)))
        o.lv_steps(0, 80, mode=0)
        d.run_checked(80, 0)
        _compare(o, d, tol=1e-10)


def test_noisy_langevin_same_streams_and_chunk_restarts():
    """Uniform thermostat noise from stream 4 (and 5 in 3-D) of the same Philox layout; the second chunk starts
    with a force recalculation on both sides (counter = step0 of the chunk)."""
    for n_dims, n, L in ((2, 1000, 125.0), (3, 216, 30.0)):
        o, d, _ = _pair(n_dims, n, L, kT=KT_300)
        set_actions(o, d, np.full(n, 5.0 * stokes(1.0)[0]), np.zeros((n, 3)))
        for s0, k in ((10, 40), (50, 25), (75, 1), (76, 0)):
            o.lv_steps(s0, k, mode=2)
            st = d.run_checked(k, s0)
            assert st.steps_done == k
        _compare(o, d, tol=1e-9, tol_e=1e-8)
    # thermal velocities are there
    _, v, _ = d.particle_data_host()
    assert np.std(v) > 0.0


def test_langevin_equipartition_on_the_device():
    """Size-independent property at a size the oracle is not asked to run: <v^2> = 2 kT/m, <omega^2> = kT/I."""
    n, L = 100_000, 1772.0
    rng = np.random.default_rng(0)
    a = 0.04
    sp = _species(a=a)
    from swarmrl_b200.device import DeviceSwarm

    pos = jittered_lattice(n, L, 2, rng)
    d = DeviceSwarm(2, (L,) * 3, True, 1e-3, KT_300, 5, sp, wca_for([1.0]))
    d.set_state(pos, np.zeros(n, dtype=np.int32), theta=rng.uniform(0, 2 * np.pi, n))
    d.set_integrator("langevin")
    st = d.run_checked(400, 0)
    assert st.steps_done == 400
    _, v, _ = d.particle_data_host()
    w = d.omega[2, :n].cpu().numpy()
    assert abs(np.mean(np.sum(v**2, axis=1)) / (2 * KT_300 / sp[0]["mass"]) - 1) < 0.02 + a
    assert abs(np.mean(w**2) / (KT_300 / sp[0]["rinertia"][0]) - 1) < 0.02 + a


def test_langevin_overflow_resumes_inside_the_step():
    """A Verlet overflow stops the chunk between the two half kicks of a step; the resumed run closes that step
    (no second force recalculation), so the result equals the uninterrupted oracle run."""
    rng = np.random.default_rng(2)
    n, L, steps = 400, 60.0, 600
    sp, wca = _species(), wca_for([1.0])
    pos = jittered_lattice(n, L, 2, rng)
    # the swimmers head for the centre: the neighbour count there outgrows the initial capacity mid-run
    th = np.arctan2(L / 2 - pos[:, 1], L / 2 - pos[:, 0])
    o, d = make_pair(2, (L,) * 3, True, 1e-3, 0.0, 1, sp, wca, pos, np.zeros(n, dtype=np.int32), theta=th,
                     device_kw=dict(max_neighbors=6))
    d.set_integrator("langevin")
    set_actions(o, d, np.full(n, 40.0 * stokes(1.0)[0]), np.zeros((n, 3)))
    o.lv_steps(0, steps, mode=2)
    done, interrupted = 0, []
    while True:
        d.run(steps - done, done)
        st = d.stats()
        done += int(st.steps_done)
        if not st.overflow:
            break
        interrupted.append(int(st.steps_done))
        d.set_max_neighbors(st.capacity + 2)
    assert done == steps and any(k > 0 for k in interrupted), interrupted
    _compare(o, d, tol=1e-7, tol_e=1e-7)


def test_langevin_with_walls_and_constant_force():
    """Non-periodic arena (espresso.py:667-711) and a constant external force (:834-851) under velocity Verlet:
    inertial particles bounce off the WCA walls but stay inside; the device follows the oracle."""
    rng = np.random.default_rng(4)
    n, L = 100, 40.0
    sp, wca = _species(), wca_for([1.0])
    walls = dict(sigma=[1.0 * 2 ** (-1 / 6)], eps=[1.0], cut=[1.0])
    pos = jittered_lattice(n, L, 2, rng)
    pos[:, :2] = 2.0 + pos[:, :2] * (L - 4.0) / L
    o, d = make_pair(2, (L,) * 3, False, 1e-3, 0.0, 1, sp, wca, pos, np.zeros(n, dtype=np.int32),
                     theta=rng.uniform(0, 2 * np.pi, n), walls=walls)
    d.set_integrator("langevin")
    set_actions(o, d, np.full(n, 4.0 * stokes(1.0)[0]), np.zeros((n, 3)))
    F = np.tile([0.5 * stokes(1.0)[0], 0.0, 0.0], (n, 1))
    for a in range(3):
        o.ext_force[a][:] = F[:, a]
    d.set_ext_force_host(F)
    o.lv_steps(0, 3000, mode=0)
    st = d.run_checked(3000, 0)
    assert st.steps_done == 3000
    _compare(o, d, tol=1e-7, tol_e=1e-7)
    p, _, _ = d.particle_data_host()
    assert p[:, :2].min() > 0.0 and p[:, :2].max() < L
    assert (p[:, :2].min() < 1.0) or (p[:, :2].max() > L - 1.0)


def test_langevin_setup_errors():
    from swarmrl_b200.device import DeviceSwarm

    gt, gr = stokes(1.0)
    bad_mass = [dict(type=0, gamma_t=(gt,) * 3, gamma_r=(gr,) * 3, mass=0.0, rinertia=(1.0,) * 3)]
    d = DeviceSwarm(2, (50.0,) * 3, True, 1e-3, 0.0, 1, bad_mass, wca_for([1.0]))
    with pytest.raises(ValueError):
        d.set_integrator("langevin")
    with pytest.raises(ValueError):
        d.set_integrator("nose-hoover")
    aniso_inertia = [dict(type=0, gamma_t=(gt,) * 3, gamma_r=(gr,) * 3, mass=1.0, rinertia=(1.0, 2.0, 3.0))]
    d3 = DeviceSwarm(3, (50.0,) * 3, True, 1e-3, 0.0, 1, aniso_inertia, wca_for([1.0]))
    with pytest.raises(ValueError):
        d3.set_integrator("langevin")
    d3.set_integrator("brownian")


def test_domain_exchange_refuses_the_langevin_integrator():
    """Halo records carry x, y, theta and aux channels only: velocities, state under velocity Verlet, do not
    travel, so the exchange must fail loudly rather than integrate ghosts with stale velocities."""
    from swarmrl_b200.domain import DomainSwarm

    rng = np.random.default_rng(1)
    n, L = 2000, 250.0
    pos = jittered_lattice(n, L, 2, rng)
    d = DomainSwarm((L, L), 1e-3, 0.0, 3, _species(), wca_for([1.0]), exchange_every=5)
    d.scatter(pos[:, :2], rng.uniform(0, 2 * np.pi, n))
    d.run(5, 0)
    d.sync()
    d.sw.set_integrator("langevin")
    with pytest.raises(RuntimeError, match="Brownian"):
        d.exchange()
