"""
GPU parity tests of the engine kernels (cell hash, Verlet lists, WCA forces, Brownian step, steepest
descent, Philox noise) against the CPU oracle, through the C-ABI of libswarm_b200.so.

Bars (BASELINE.json north_star): cell / neighbour lists bit-exact as integer sets; noise-free fp64
trajectories within 1e-9 relative per step; Philox words bit-exact.
"""
import numpy as np
import pytest

from helpers import KT_300, jittered_lattice, make_pair, random_quats, rel_err, set_actions, species_for, wca_for

pytestmark = pytest.mark.gpu


def _swarm2d(n=1000, L=125.0, kT=0.0, seed=1234, radii=(1.0,), device_kw=None, periodic=True, walls=None):
    rng = np.random.default_rng(seed)
    sp, wca = species_for(list(radii)), wca_for(list(radii))
    pos = jittered_lattice(n, L, 2, rng)
    th = rng.uniform(0, 2 * np.pi, n)
    sidx = (np.arange(n) % len(radii)).astype(np.int32)
    o, d = make_pair(2, (L, L, L), periodic, 1e-3, kT, 42, sp, wca, pos, sidx, theta=th, device_kw=device_kw,
                     walls=walls)
    return o, d, sp


def _swarm3d(n=512, L=40.0, kT=0.0, seed=99, device_kw=None, aniso=None):
    rng = np.random.default_rng(seed)
    sp, wca = species_for([1.0], aniso=aniso), wca_for([1.0])
    pos = jittered_lattice(n, L, 3, rng)
    q = random_quats(n, rng)
    o, d = make_pair(3, (L, L, L), True, 1e-3, kT, 42, sp, wca, pos, np.zeros(n, dtype=np.int32), quat=q,
                     device_kw=device_kw)
    return o, d, sp


def test_philox_gaussians_match_oracle():
    from oracle import oracle

    _, d, _ = _swarm2d(n=64)
    for step, stream in ((0, 0), (7, 3), (2**40 + 5, 1)):
        g = d.gaussians(64, step, stream).cpu().numpy()
        ref = np.array([oracle.gaussians(42, i, step, stream) for i in range(64)])
        # same Philox words; libm vs CUDA log/sincospi differ by a few ulp
        assert np.max(np.abs(g - ref)) < 1e-13
    big = d.gaussians(200000, 3, 0).cpu().numpy().ravel()
    assert abs(big.mean()) < 5e-3 and abs(big.std() - 1.0) < 5e-3


@pytest.mark.parametrize("nc", [(52, 52, 1), (7, 13, 1), (1, 1, 1)])
def test_cell_hash_bit_exact(nc):
    o, d, _ = _swarm2d(n=3000, L=125.0)
    # unfold some positions by whole boxes: the hash must wrap them
    o.x[::7] += 125.0; o.y[::5] -= 250.0
    d.pos[0, ::7] += 125.0; d.pos[1, ::5] -= 250.0
    assert np.array_equal(d.cell_index(nc), o.cell_index(nc))


@pytest.mark.parametrize("persistent", [True, False])
def test_neighbour_list_bit_exact_2d(persistent):
    o, d, _ = _swarm2d(n=2000, L=100.0, radii=(1.0, 1.5), device_kw=dict(persistent=persistent, max_neighbors=32))
    got = d.neighbors()
    ref = o.pairs_within(skin=0.4)
    assert got.shape == ref.shape and np.array_equal(got, ref)


def test_neighbour_list_bit_exact_3d_and_nonperiodic():
    o, d, _ = _swarm3d(n=1000, L=24.0)
    assert np.array_equal(d.neighbors(), o.pairs_within(skin=0.4))
    o, d, _ = _swarm2d(n=500, L=50.0, periodic=False)
    assert np.array_equal(d.neighbors(), o.pairs_within(skin=0.4))


def test_neighbour_list_edge_cases():
    # empty swarm, single particle, tiny box (1-2 cells per axis)
    from swarmrl_b200.device import DeviceSwarm

    sp, wca = species_for([1.0]), wca_for([1.0])
    d = DeviceSwarm(2, (10, 10, 10), True, 1e-3, 0.0, 1, sp, wca)
    d.set_state(np.zeros((0, 3)), np.zeros(0, dtype=np.int32), theta=np.zeros(0))
    assert d.neighbors().shape == (0, 2)
    st = d.run_checked(5, 0)
    assert st.steps_done == 0
    d.set_state(np.array([[1.0, 2.0, 0.0]]), np.zeros(1, dtype=np.int32), theta=np.zeros(1))
    assert d.neighbors().shape == (0, 2)
    for L in (2.5, 5.0, 7.3):
        rng = np.random.default_rng(3)
        pos = np.zeros((6, 3)); pos[:, :2] = rng.uniform(0, L, (6, 2))
        o, d = make_pair(2, (L, L, L), True, 1e-3, 0.0, 1, sp, wca, pos, np.zeros(6, dtype=np.int32),
                         theta=np.zeros(6), device_kw=dict(max_neighbors=16))
        assert np.array_equal(d.neighbors(), o.pairs_within(skin=0.4)), L


def test_forces_match_oracle():
    o, d, sp = _swarm2d(n=2000, L=80.0, radii=(1.0, 1.3))
    n = o.n
    set_actions(o, d, np.linspace(0, 3, n), np.zeros((n, 3)))
    F = d.forces().cpu().numpy().T
    ref = o.forces(mode=0)
    assert np.abs(ref).max() > 1.0  # overlapping pairs exist
    assert rel_err(F, ref) < 1e-12


@pytest.mark.parametrize("persistent", [True, False])
def test_noise_free_trajectory_2d_config2(persistent):
    """BASELINE config 2: 1,000 colloids, WCA, 1,000 steps per slice, kT=0, <= 1e-9 relative per step."""
    o, d, sp = _swarm2d(n=1000, L=125.0, device_kw=dict(persistent=persistent))
    n = o.n
    gt, gr = sp[0]["gamma_t"][0], sp[0]["gamma_r"][0]
    set_actions(o, d, np.full(n, 5.0 * gt), np.tile([0, 0, gr * 4 * np.pi / 180], (n, 1)))
    steps = 1000 if persistent else 200
    o.bd_steps(0, steps, mode=2)
    st = d.run_checked(steps, 0)
    assert st.steps_done == steps and not st.overflow and st.persistent == int(persistent)
    p, v, e = d.particle_data_host()
    assert rel_err(p, o.pos) < 1e-9 and rel_err(e, o.directors, 1.0) < 1e-9
    assert rel_err(v, o.velocities) < 1e-9
    assert np.all(p[:, 2] == 0.0) and np.all(e[:, 2] == 0.0)


def test_noise_free_trajectory_3d():
    o, d, sp = _swarm3d(n=512, L=30.0)
    n = o.n
    gt, gr = sp[0]["gamma_t"][0], sp[0]["gamma_r"][0]
    rng = np.random.default_rng(5)
    set_actions(o, d, np.full(n, 5.0 * gt), rng.normal(size=(n, 3)) * gr)
    o.bd_steps(0, 300, mode=2)
    st = d.run_checked(300, 0)
    assert st.steps_done == 300
    p, v, e = d.particle_data_host()
    assert rel_err(p, o.pos) < 1e-9 and rel_err(e, o.directors, 1.0) < 1e-9 and rel_err(v, o.velocities) < 1e-9
    w = d.omega[:, :n].cpu().numpy().T
    assert rel_err(w, o.omegas) < 1e-9


def test_anisotropic_friction_3d_and_2d():
    o, d, sp = _swarm3d(n=64, L=40.0, aniso=((1.0, 2.0, 3.0), (0.5, 0.7, 1.1)))
    n = o.n
    set_actions(o, d, np.full(n, 2.0), np.tile([0.3, -0.2, 1.0], (n, 1)))
    o.bd_steps(0, 50, mode=0)
    d.run_checked(50, 0)
    p, v, e = d.particle_data_host()
    assert rel_err(p, o.pos) < 1e-10 and rel_err(e, o.directors, 1.0) < 1e-10 and rel_err(v, o.velocities) < 1e-10


def test_noisy_trajectory_same_streams():
    """With identical Philox streams the noisy trajectory also agrees (short run: chaos amplifies ulps)."""
    o, d, sp = _swarm2d(n=1000, L=125.0, kT=KT_300)
    n = o.n
    set_actions(o, d, np.full(n, 5.0 * sp[0]["gamma_t"][0]), np.zeros((n, 3)))
    o.bd_steps(10, 100, mode=2)
    st = d.run_checked(100, 10)
    assert st.steps_done == 100
    p, v, e = d.particle_data_host()
    assert rel_err(p, o.pos) < 1e-9 and rel_err(e, o.directors, 1.0) < 1e-8
    assert rel_err(v, o.velocities) < 1e-7
    o3, d3, sp3 = _swarm3d(n=216, L=30.0, kT=KT_300)
    o3.bd_steps(0, 50, mode=2)
    d3.run_checked(50, 0)
    p, v, e = d3.particle_data_host()
    assert rel_err(p, o3.pos) < 1e-9 and rel_err(e, o3.directors, 1.0) < 1e-8


def test_chunked_runs_equal_one_run():
    """integrator.run(a); run(b) == run(a+b): the Philox counter is the global step index."""
    _, d1, sp = _swarm2d(n=500, L=90.0, kT=KT_300)
    _, d2, _ = _swarm2d(n=500, L=90.0, kT=KT_300)
    f = np.full(500, 5.0 * sp[0]["gamma_t"][0])
    set_actions(None, d1, f, np.zeros((500, 3)))
    set_actions(None, d2, f, np.zeros((500, 3)))
    d1.run_checked(60, 0)
    for s0, k in ((0, 7), (7, 20), (27, 33)):
        d2.run_checked(k, s0)
    p1, _, e1 = d1.particle_data_host()
    p2, _, e2 = d2.particle_data_host()
    assert np.array_equal(p1, p2) and np.array_equal(e1, e2)


def test_steepest_descent_relax_matches_oracle_and_removes_overlap():
    rng = np.random.default_rng(11)
    n, L = 400, 80.0  # phi = 0.2: random placement overlaps, and the overlaps can be relaxed away
    sp, wca = species_for([1.0]), wca_for([1.0])
    pos = np.zeros((n, 3)); pos[:, :2] = rng.uniform(0, L, (n, 2))
    o, d = make_pair(2, (L,) * 3, True, 1e-3, 0.0, 1, sp, wca, pos, np.zeros(n, dtype=np.int32),
                     theta=rng.uniform(0, 6.28, n), device_kw=dict(max_neighbors=32))
    o.sd_steps(50, 0.1, 0.1, mode=2)
    st = d.relax_checked(50)
    assert st.steps_done <= 50
    p, _, _ = d.particle_data_host()
    assert rel_err(p, o.pos) < 1e-8
    o.sd_steps(950, 0.1, 0.1, mode=2)
    d.relax_checked(950)
    p, _, _ = d.particle_data_host()
    dx = p[:, None, :2] - p[None, :, :2]
    dx -= L * np.round(dx / L)
    r = np.sqrt((dx**2).sum(-1)) + 10 * np.eye(n)
    assert r.min() > 1.9


def test_walls_keep_particles_inside():
    """test_confining_walls.py:40-51 -- driven particles stay in the box (inequalities only)."""
    walls = dict(sigma=[1.0 * 2 ** (-1 / 6)], eps=[1.0], cut=[1.0])
    o, d, sp = _swarm2d(n=100, L=40.0, periodic=False, walls=walls)
    n = o.n
    set_actions(o, d, np.full(n, 5.0 * sp[0]["gamma_t"][0]), np.zeros((n, 3)))
    o.bd_steps(0, 4000, mode=0)
    d.run_checked(4000, 0)
    p, _, _ = d.particle_data_host()
    assert rel_err(p, o.pos) < 1e-7
    for q in (p, o.pos):
        assert q[:, :2].min() > 0.0 and q[:, :2].max() < 40.0
    assert (p[:, :2].min() < 1.0) or (p[:, :2].max() > 39.0)  # somebody is actually leaning on a wall


def test_verlet_overflow_is_reported_and_recovered():
    rng = np.random.default_rng(2)
    n, L = 300, 12.0
    sp, wca = species_for([1.0]), wca_for([1.0])
    pos = np.zeros((n, 3)); pos[:, :2] = rng.uniform(0, L, (n, 2))
    _, d = make_pair(2, (L,) * 3, True, 1e-3, 0.0, 1, sp, wca, pos, np.zeros(n, dtype=np.int32), theta=np.zeros(n),
                     device_kw=dict(max_neighbors=2))
    d.relax(10)
    st = d.stats()
    assert st.overflow == 1 and st.steps_done == 0
    st = d.relax_checked(10)
    assert st.overflow == 0 and st.capacity > 2


def test_new_direction_2d_and_3d():
    """espresso.py:1236-1249; test_espresso_2d.py:84-92."""
    o, d, _ = _swarm2d(n=50)
    rng = np.random.default_rng(8)
    ang = rng.uniform(0, 2 * np.pi, 50)
    nd = np.stack([np.cos(ang), np.sin(ang), np.zeros(50)], 1)
    has = (np.arange(50) % 2).astype(np.int32)
    d.set_new_directions_host(nd, has)
    e0 = d.directors().cpu().numpy().T.copy()
    d.apply_new_direction()
    e1 = d.directors().cpu().numpy().T
    assert np.allclose(e1[has == 1], nd[has == 1], atol=1e-12) and np.array_equal(e1[has == 0], e0[has == 0])
    o3, d3, _ = _swarm3d(n=40)
    v = rng.normal(size=(40, 3)); v /= np.linalg.norm(v, axis=1, keepdims=True)
    d3.set_new_directions_host(v, np.ones(40, dtype=np.int32))
    d3.apply_new_direction()
    assert np.allclose(d3.directors().cpu().numpy().T, v, atol=1e-12)


def test_batched_environments_are_independent():
    """n_envs > 1: each block of n/n_envs particles is its own periodic box (BASELINE config 3)."""
    E, m, L = 8, 32, 60.0
    rng = np.random.default_rng(17)
    sp, wca = species_for([1.0]), wca_for([1.0])
    pos = np.concatenate([jittered_lattice(m, L, 2, rng) for _ in range(E)])
    th = rng.uniform(0, 2 * np.pi, E * m)
    from swarmrl_b200.device import DeviceSwarm
    from oracle.oracle import OracleSwarm

    d = DeviceSwarm(2, (L,) * 3, True, 1e-2, 0.0, 42, sp, wca, n_envs=E)
    d.set_state(pos, np.zeros(E * m, dtype=np.int32), theta=th)
    f = np.full(E * m, 5.0 * sp[0]["gamma_t"][0])
    d.set_actions_host(f, np.zeros((E * m, 3)))
    d.run_checked(400, 0)
    p, _, _ = d.particle_data_host()
    for e in range(E):
        o = OracleSwarm(2, (L,) * 3, True, 1e-2, 0.0, 42, sp, wca)
        sl = slice(e * m, (e + 1) * m)
        o.set_state(pos[sl], np.zeros(m, dtype=np.int32), theta=th[sl])
        o.f_swim[:] = f[sl]
        o.bd_steps(0, 400, mode=0)
        assert rel_err(p[sl], o.pos) < 1e-9, e


def test_ensemble_shards_reproduce_the_whole_ensemble_with_noise():
    """Ensemble sharding: two handles holding E/2 environments each, their Philox ids bound
    to the GLOBAL agent index (swarm_bind_ids), give bit-for-bit the trajectories of one handle holding all
    E environments -- noise on, so a single mis-keyed random number would show."""
    import torch

    from swarmrl_b200.device import DeviceSwarm

    E, m, L = 8, 32, 60.0
    rng = np.random.default_rng(23)
    sp, wca = species_for([1.0]), wca_for([1.0])
    pos = np.concatenate([jittered_lattice(m, L, 2, rng) for _ in range(E)])
    th = rng.uniform(0, 2 * np.pi, E * m)
    f = rng.uniform(0, 20.0, E * m) * sp[0]["gamma_t"][0]

    def run(lo, hi):
        sl = slice(lo * m, hi * m)
        d = DeviceSwarm(2, (L,) * 3, True, 1e-3, KT_300, 42, sp, wca, n_envs=hi - lo, skin=1.0)
        d.set_state(pos[sl], np.zeros((hi - lo) * m, dtype=np.int32), theta=th[sl])
        d.bind_ids(torch.arange(lo * m, hi * m, dtype=torch.int64, device=d.dev))
        d.set_actions_host(f[sl], np.zeros(((hi - lo) * m, 3)))
        d.run_checked(300, 0)
        p, v, e = d.particle_data_host()
        return p, v, e

    whole = run(0, E)
    for lo, hi in ((0, E // 2), (E // 2, E)):
        part = run(lo, hi)
        for a, b in zip(part, whole):
            assert np.array_equal(a, b[lo * m:hi * m])


def test_warp_per_environment_kernel_equals_the_list_kernels(monkeypatch):
    """k_bd2d_env (one warp per environment, all pairs by shuffle in ascending index, no lists) against the
    Verlet-list kernels on the same batched ensemble: same trajectories to 1e-12 (FMA contraction differs between
    the kernels), noise on, ragged environment size (29 < 32), two particle types; and the fp32 pre-filter of the
    pair loop changes NOTHING (bitwise) against the exact all-pairs loop."""
    import torch

    from swarmrl_b200.device import DeviceSwarm

    E, m, L = 64, 29, 60.0
    rng = np.random.default_rng(5)
    sp, wca = species_for([1.0, 1.5]), wca_for([1.0, 1.5])
    pos = np.concatenate([jittered_lattice(m, L, 2, rng, jitter=0.15) for _ in range(E)])
    th = rng.uniform(0, 2 * np.pi, E * m)
    sidx = rng.integers(0, 2, E * m).astype(np.int32)
    f = rng.uniform(0, 30.0, E * m) * sp[0]["gamma_t"][0]
    tq = np.zeros((E * m, 3)); tq[:, 2] = rng.uniform(-1, 1, E * m) * sp[0]["gamma_r"][0]

    def run(env_kernel, exact_pairs=False):
        monkeypatch.delenv("SWARM_B200_ENV_EXACT_PAIRS", raising=False)
        if exact_pairs:
            monkeypatch.setenv("SWARM_B200_ENV_EXACT_PAIRS", "1")
        if env_kernel:
            monkeypatch.delenv("SWARM_B200_NO_ENV_KERNEL", raising=False)
        else:
            monkeypatch.setenv("SWARM_B200_NO_ENV_KERNEL", "1")
        d = DeviceSwarm(2, (L,) * 3, True, 1e-3, KT_300, 7, sp, wca, n_envs=E, skin=1.0)
        d.set_state(pos, sidx, theta=th)
        d.set_actions_host(f, tq)
        st0 = d.run_checked(150, 0)
        st1 = d.run_checked(50, 150)
        assert st1.steps_done == 50
        return d.particle_data_host() + (int(st1.rebuilds),)

    a, b, c = run(True), run(False), run(True, exact_pairs=True)
    assert a[3] == 0 and b[3] > 0          # the environment kernel never builds a neighbour list
    for u, v, w in zip(a[:3], b[:3], c[:3]):
        assert rel_err(u, v) < 1e-12
        assert np.array_equal(u, w)


@pytest.mark.parametrize("scale", [1e3, 1e6, 3e9, 1e12])
def test_director_of_large_angles(scale):
    """The 2-D kernels reduce theta themselves (swm_sincos; library routine beyond 2^30 rad): wound-up angles,
    as they accumulate under a constant torque, still give the oracle's director and trajectory."""
    rng = np.random.default_rng(12)
    n, L = 600, 120.0
    sp, wca = species_for([1.0]), wca_for([1.0])
    pos = jittered_lattice(n, L, 2, rng)
    th = rng.uniform(-scale, scale, n)
    o, d = make_pair(2, (L, L, L), True, 1e-3, 0.0, 42, sp, wca, pos, np.zeros(n, dtype=np.int32), theta=th)
    set_actions(o, d, np.full(n, 5.0 * sp[0]["gamma_t"][0]), np.zeros((n, 3)))
    o.bd_steps(0, 50, mode=2)
    st = d.run_checked(50, 0)
    assert st.steps_done == 50
    p, v, e = d.particle_data_host()
    assert rel_err(p, o.pos) < 1e-12 and rel_err(v, o.velocities) < 1e-9


def test_overflow_at_the_capacity_ceiling_raises_instead_of_looping():
    """More than SWARM_MAX_NEIGHBORS particles within cut-off + skin of each other: the Verlet capacity cannot
    grow past its ceiling, so relax_checked / run_checked must raise (they used to retry for ever)."""
    rng = np.random.default_rng(4)
    n, L = 200, 40.0
    sp, wca = species_for([1.0]), wca_for([1.0])
    pos = np.zeros((n, 3)); pos[:, :2] = 20.0 + rng.uniform(-0.4, 0.4, (n, 2))
    _, d = make_pair(2, (L,) * 3, True, 1e-3, 0.0, 1, sp, wca, pos, np.zeros(n, dtype=np.int32), theta=np.zeros(n))
    with pytest.raises(OverflowError, match="128"):
        d.relax_checked(10)
    _, d2 = make_pair(2, (L,) * 3, True, 1e-3, 0.0, 1, sp, wca, pos, np.zeros(n, dtype=np.int32), theta=np.zeros(n))
    with pytest.raises(OverflowError):
        d2.run_checked(5, 0)


@pytest.mark.parametrize("persistent", [True, False])
def test_steps_done_counts_the_whole_run_across_sub_launches(persistent):
    """swarm_run splits long runs into sub-launches (8192 steps, 1024 without the persistent kernels); steps_done
    is the count over the WHOLE run, and a run resumed after it equals one uninterrupted run."""
    n_steps = 10000 if persistent else 2500
    o, d, sp = _swarm2d(n=256, L=64.0, device_kw=dict(persistent=persistent))
    set_actions(o, d, np.full(256, 5.0 * sp[0]["gamma_t"][0]), np.zeros((256, 3)))
    d.run(n_steps, 0)
    st = d.stats()
    assert st.steps_done == n_steps and st.steps_requested == n_steps and not st.overflow
    o.bd_steps(0, n_steps, mode=2)
    p, _, _ = d.particle_data_host()
    assert rel_err(p, o.pos) < 1e-8


def test_overflow_in_a_later_sub_launch_resumes_at_the_right_step():
    """A list overflow after the first sub-launch: steps_done must include the completed sub-launches so that
    run_checked resumes at step0 + steps_done (it used to replay the first 8192 steps' noise)."""
    rng = np.random.default_rng(31)
    n, L = 128, 44.0
    sp, wca = species_for([1.0]), wca_for([1.0])
    pos = jittered_lattice(n, L, 2, rng, jitter=0.1)
    # every particle swims towards the box centre: the swarm compacts and the 2-entry lists overflow late
    th = np.arctan2(L / 2 - pos[:, 1], L / 2 - pos[:, 0])
    o, d = make_pair(2, (L,) * 3, False, 1e-3, 0.0, 9, sp, wca, pos, np.zeros(n, dtype=np.int32), theta=th,
                     device_kw=dict(max_neighbors=2, skin=0.4))
    f = np.full(n, 0.15 * sp[0]["gamma_t"][0])   # first particle with 3 neighbours within 2.4 at step ~12 800 (oracle)
    set_actions(o, d, f, np.zeros((n, 3)))
    total = 20000
    d.run(total, 0)
    st = d.stats()
    assert st.overflow == 1 and 0 < st.steps_done < total, (st.overflow, st.steps_done)
    first = int(st.steps_done)
    assert first > 8192, "the overflow was meant to happen after the first sub-launch; re-tune the test swarm"
    o.bd_steps(0, first, mode=2)
    p, _, _ = d.particle_data_host()
    assert rel_err(p, o.pos) < 1e-8          # steps_done is the step the state actually stands at
    d.set_max_neighbors(16)
    st = d.run_checked(total - first, first)
    o.bd_steps(first, total - first, mode=2)
    p, _, _ = d.particle_data_host()
    assert rel_err(p, o.pos) < 1e-7


def test_antiparallel_new_direction_turns_by_pi_without_nan():
    """espresso.py:1240-1248 has a 0/0 axis here; the device kernel turns by +pi and never stores a NaN angle."""
    o, d, _ = _swarm2d(n=16)
    e0 = d.directors().cpu().numpy().T.copy()
    d.set_new_directions_host(-e0, np.ones(16, dtype=np.int32))
    d.apply_new_direction()
    e1 = d.directors().cpu().numpy().T
    assert np.all(np.isfinite(e1)) and np.allclose(e1, -e0, atol=1e-12)


def test_config2_noise_statistics_over_32_replicas():
    """BASELINE config 2 with noise ON (north_star's third check): 32 replicas of the 1,000-colloid
    WCA swarm in ONE batched launch (n_envs = 32; every replica draws its own noise because the Philox id is the global
    particle index).  (a) <e(0).e(t)> of every replica against exp(-D_r t) -- exact for the Euler-Maruyama angle update,
    whatever the forces do; (b) MSD(t) against the CPU oracle run on INDEPENDENT noise (other seeds), Welch's test on
    the replica means.  Bars: at least three quarters of the eight checks within 2 sigma of the standard error and all
    within 3 (the four time points of one quantity come from the same trajectories, so they are not independent draws:
    two of them beyond 2 sigma together is one fluctuation, not two)."""
    import torch

    from oracle.oracle import OracleSwarm
    from swarmrl_b200.device import DeviceSwarm

    E, m, L, dt = 32, 1000, 125.0, 1e-3
    rng = np.random.default_rng(1234)
    sp, wca = species_for([1.0]), wca_for([1.0])
    pos1 = jittered_lattice(m, L, 2, rng)
    th1 = rng.uniform(0, 2 * np.pi, m)
    pos, th = np.tile(pos1, (E, 1)), np.tile(th1, E)
    d = DeviceSwarm(2, (L,) * 3, True, dt, KT_300, 42, sp, wca, n_envs=E, skin=1.0)
    d.set_state(pos, np.zeros(E * m, dtype=np.int32), theta=th)
    gt, gr = sp[0]["gamma_t"][0], sp[0]["gamma_r"][0]
    Dr = KT_300 / gr
    times = [50, 100, 200, 300]
    msd_g, cor_g = [], []
    done = 0
    for t in times:
        st = d.run_checked(t - done, done)
        assert not st.overflow
        done = t
        p = d.pos[:2, : E * m].cpu().numpy().T
        a = d.theta[: E * m].cpu().numpy()
        msd_g.append(((p - pos[:, :2]) ** 2).sum(axis=1).reshape(E, m).mean(axis=1))
        cor_g.append(np.cos(a - th).reshape(E, m).mean(axis=1))
    msd_g, cor_g = np.array(msd_g), np.array(cor_g)          # (times, replicas)
    # (a) orientation autocorrelation against the analytic decay
    z = []
    for k, t in enumerate(times):
        se = cor_g[k].std(ddof=1) / np.sqrt(E)
        z.append(abs(cor_g[k].mean() - np.exp(-Dr * t * dt)) / se)
    # (b) MSD against oracle replicas on independent noise
    R = 8
    msd_o = np.zeros((len(times), R))
    for r in range(R):
        o = OracleSwarm(2, (L,) * 3, True, dt, KT_300, 1000 + r, sp, wca)
        o.set_state(pos1, np.zeros(m, dtype=np.int32), theta=th1)
        done = 0
        for k, t in enumerate(times):
            o.bd_steps(done, t - done, mode=2)
            done = t
            msd_o[k, r] = ((o.pos[:, :2] - pos1[:, :2]) ** 2).sum(axis=1).mean()
    for k in range(len(times)):
        se = np.sqrt(msd_g[k].var(ddof=1) / E + msd_o[k].var(ddof=1) / R)
        z.append(abs(msd_g[k].mean() - msd_o[k].mean()) / se)
    z = np.array(z)
    assert np.mean(z < 2.0) >= 0.75, z
    assert np.all(z < 3.0), z
    # free-diffusion scale (collisions slow the swarm down, the first contacts of the jittered lattice push it apart)
    D = KT_300 / gt
    assert 0.5 * 4 * D * times[-1] * dt < msd_g[-1].mean() < 1.5 * 4 * D * times[-1] * dt


def test_set_actions_host_2d_writes_what_the_three_vector_entry_writes():
    """swarm_set_actions_host_2d (f_swim, tau_z per particle: half the host->device bytes of a 2-D slice) against
    swarm_set_actions_host (espresso.py:1227-1235 writes a 3-vector torque); a 3-D swarm refuses it."""
    import ctypes as C

    from swarmrl_b200 import _lib
    from swarmrl_b200.device import DeviceSwarm

    rng = np.random.default_rng(2)
    n, L = 3000, 200.0
    sp, wca = species_for([1.0]), wca_for([1.0])
    pos = jittered_lattice(n, L, 2, rng)
    f = rng.uniform(0, 3, n)
    tq = np.zeros((n, 3)); tq[:, 2] = rng.uniform(-1, 1, n)
    got = []
    for two_d in (False, True):
        d = DeviceSwarm(2, (L,) * 3, True, 1e-3, 0.0, 1, sp, wca)
        d.set_state(pos, np.zeros(n, dtype=np.int32), theta=np.zeros(n))
        if two_d:
            tz = np.ascontiguousarray(tq[:, 2])
            _lib.check(d.lib.swarm_set_actions_host_2d(d.h, f.ctypes.data, tz.ctypes.data, d._stream()))
        else:
            _lib.check(d.lib.swarm_set_actions_host(d.h, f.ctypes.data, tq.ctypes.data, d._stream()))
        d.synchronize()
        d.run_checked(20, 0)
        got.append((d.f_swim[:n].cpu().numpy(), d.torque[2, :n].cpu().numpy(), d.particle_data_host()[0], d.theta[:n].cpu().numpy()))
    for a, b in zip(*got):
        assert np.array_equal(a, b)
    d3 = DeviceSwarm(3, (L,) * 3, True, 1e-3, 0.0, 1, sp, wca)
    p3 = jittered_lattice(200, L, 3, rng)
    q = np.zeros((200, 4)); q[:, 0] = 1.0
    d3.set_state(p3, np.zeros(200, dtype=np.int32), quat=q)
    assert d3.lib.swarm_set_actions_host_2d(d3.h, f.ctypes.data, f.ctypes.data, d3._stream()) != 0


@pytest.mark.parametrize("n,L,radii,walls", [(10, 1000.0, (1.0,), False), (100, 60.0, (1.0,), False), (1500, 150.0, (1.0, 1.5), False),
                                             (2048, 170.0, (1.0,), True)])
def test_cluster_kernel_is_the_resident_kernel_bit_for_bit(n, L, radii, walls):
    """k_bd2d_cluster (one 8-CTA thread-block cluster: positions in distributed shared memory, hardware cluster barrier,
    in-cluster rebuild -- all pairs up to 128 particles, a shared-memory cell list above) against k_bd2d_resident
    (SWARM_B200_NO_CLUSTER=1) on the same noisy swarm, launches cut at odd and even steps: the same bits.  A missed pair in
    the cluster kernel's Verlet lists would show as a different trajectory (two species, walls, 2 048 = the largest size)."""
    import os

    from swarmrl_b200.device import DeviceSwarm

    rng = np.random.default_rng(8)
    sp, wca = species_for(list(radii)), wca_for(list(radii))
    pos = jittered_lattice(n, L, 2, rng) if n > 10 else np.c_[rng.uniform(100, 900, (n, 2)), np.zeros(n)]
    th = rng.uniform(0, 2 * np.pi, n)
    sidx = (np.arange(n) % len(radii)).astype(np.int32)
    gt, gr = sp[0]["gamma_t"][0], sp[0]["gamma_r"][0]
    wl = None
    if walls:
        wl = dict(sigma=np.array([2.0 * 2 ** (-1 / 6)]), eps=np.array([1.0]), cut=np.array([2.0]))
    out = []
    for no_cluster in (False, True):
        if no_cluster:
            os.environ["SWARM_B200_NO_CLUSTER"] = "1"
        try:
            d = DeviceSwarm(2, (L, L, L), not walls, 1e-3, KT_300, 42, sp, wca, walls=wl, skin=0.4)
            d.set_state(pos, sidx, theta=th)
            d.set_actions_host(np.full(n, 5.0 * gt), np.tile([0, 0, gr * 4 * np.pi / 180], (n, 1)))
            step = 0
            for m in (37, 100, 1):
                st = d.run_checked(m, step)
                assert st.steps_done == m and not st.overflow
                step += m
            p, v, e = d.particle_data_host()
            out.append((p, v, e, int(st.rebuilds)))
        finally:
            os.environ.pop("SWARM_B200_NO_CLUSTER", None)
    (p0, v0, e0, rb0), (p1, v1, e1, rb1) = out
    assert np.array_equal(p0, p1) and np.array_equal(e0, e1) and np.array_equal(v0, v1)
    assert rb0 >= 3 and rb1 >= 1      # the cluster kernel rebuilds at the start of each of the three launches
