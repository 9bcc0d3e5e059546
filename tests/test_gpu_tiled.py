"""
GPU parity tests of the tiled 2-D Brownian kernel (csrc/swarm_tiled.cuh) and the spatial slot order behind it
(csrc/swarm_sort.cuh): the path BASELINE config 4 (100,000 agents) runs on.

What has to hold (BASELINE.json north_star): the Verlet list of the tiled path is the oracle's pair set bit for bit;
noise-free trajectories stay within 1e-9 of the oracle; and, because forces are summed in ascending particle id and
the noise is keyed by the id, the tiled kernel reproduces the per-step-barrier kernel (k_bd2d_resident) BIT FOR BIT
whatever the tiling, the permutation, the rebuild cadence or the number of rolled-back periods.
"""
import os

import numpy as np
import pytest

from helpers import KT_300, jittered_lattice, make_pair, rel_err, set_actions, species_for, wca_for

pytestmark = pytest.mark.gpu


def _pair(n, L, kT, radii=(1.0,), skin=0.4, seed=1234, periodic=True, walls=None, aniso=None, max_neighbors=0):
    rng = np.random.default_rng(seed)
    sp, wca = species_for(list(radii), aniso=aniso), wca_for(list(radii))
    pos = jittered_lattice(n, L, 2, rng)
    pos[::11, 0] += L; pos[::13, 1] -= 2 * L   # unfolded coordinates: the sort key and the cell hash must wrap them
    if not periodic:
        pos = jittered_lattice(n, L, 2, rng)
    th = rng.uniform(0, 2 * np.pi, n)
    sidx = (np.arange(n) % len(radii)).astype(np.int32)
    o, d = make_pair(2, (L, L, L), periodic, 1e-3, kT, 42, sp, wca, pos, sidx, theta=th, walls=walls,
                     device_kw=dict(skin=skin, max_neighbors=max_neighbors))
    gt, gr = sp[0]["gamma_t"][0], sp[0]["gamma_r"][0]
    set_actions(o, d, np.full(n, 5.0 * gt), np.tile([0, 0, gr * 4 * np.pi / 180], (n, 1)))
    return o, d


def _run(d, mode, launches, steps):
    d.set_tiling(mode)
    for k in range(launches):
        st = d.run_checked(steps, k * steps)
        assert st.steps_done == steps and not st.overflow
    return d.particle_data_host()


def test_tiled_list_is_the_oracles_pair_set():
    o, d = _pair(6000, 180.0, 0.0, radii=(1.0, 1.5), max_neighbors=32)
    d.set_tiling(2)
    d.run_checked(1, 0)               # first step of a launch: in-kernel rebuild at the initial positions
    ts = d.tiling_stats()
    assert ts["tiled"] == 1 and ts["sorts"] == 1 and ts["ctas"] >= 40
    got = d.tiled_neighbors()
    ref = o.pairs_within(skin=0.4)
    assert got.shape == ref.shape and np.array_equal(got, ref)


def test_tiled_noise_free_trajectory_config2():
    """BASELINE config 2 through the tiled kernel: 1,000 colloids, 1,000 steps, kT = 0, <= 1e-9 relative."""
    o, d = _pair(1000, 125.0, 0.0)
    o.bd_steps(0, 1000, mode=2)
    d.set_tiling(2)
    st = d.run_checked(1000, 0)
    assert st.steps_done == 1000 and d.tiling_stats()["tiled"] == 1
    p, v, e = d.particle_data_host()
    assert rel_err(p, o.pos) < 1e-9 and rel_err(e, o.directors, 1.0) < 1e-9 and rel_err(v, o.velocities) < 1e-9


@pytest.mark.parametrize("skin", [0.4, 2.0])
def test_tiled_equals_resident_bit_for_bit_with_noise(skin):
    res = []
    for mode in (0, 2):
        _, d = _pair(20000, 800.0, KT_300, skin=skin)
        res.append(_run(d, mode, launches=3, steps=100))
        ts = d.tiling_stats()
        assert ts["tiled"] == (1 if mode else 0)
        if mode:
            assert ts["cadence"] >= 1 and ts["ctas"] == 148 or ts["ctas"] > 100
    for a, b in zip(*res):
        assert np.array_equal(a, b)


def test_tiled_types_walls_extforce_aniso_equal_resident():
    L, n = 400.0, 12000
    walls = dict(sigma=np.array([1.0, 1.3]) * 2 ** (-1 / 6), eps=np.array([1.0, 1.0]), cut=np.array([1.0, 1.3]))
    gt, gr = 4.0, 6.0
    res = []
    for mode in (0, 2):
        o, d = _pair(n, L, KT_300, radii=(1.0, 1.3), skin=1.0, periodic=False, walls=walls,
                     aniso=((gt, 2 * gt, 3 * gt), (gr, gr, gr)), max_neighbors=32)
        rng = np.random.default_rng(9)
        d.set_ext_force_host(rng.normal(size=(n, 3)) * [1.0, 1.0, 0.0])
        res.append(_run(d, mode, launches=2, steps=150))
    for a, b in zip(*res):
        assert np.array_equal(a, b)


def test_tiled_rollback_and_resort_do_not_change_the_result():
    """A far too long first cadence forces rolled-back periods; a sort before every launch permutes the slots."""
    _, d0 = _pair(20000, 800.0, KT_300, skin=0.4)
    ref = _run(d0, 0, launches=4, steps=60)
    os.environ["SWARM_B200_TILE_KINIT"] = "64"
    try:
        _, d = _pair(20000, 800.0, KT_300, skin=0.4)
        d.set_tiling(2, resort_interval=1)
        got = _run(d, 2, launches=4, steps=60)
    finally:
        del os.environ["SWARM_B200_TILE_KINIT"]
    ts = d.tiling_stats()
    assert ts["rollbacks"] >= 1 and ts["sorts"] == 4
    for a, b in zip(ref, got):
        assert np.array_equal(a, b)


def test_tiled_positions_edited_between_launches_and_overflow_growth():
    """The caller may move particles between runs (zero-copy state): the start vote sees it and rebuilds; a list
    capacity that is too small halts the launch and run_checked grows it, as on the other kernels."""
    res = []
    for mode in (0, 2):
        _, d = _pair(20000, 450.0, KT_300, skin=2.0, max_neighbors=4)
        d.set_tiling(mode)
        st = d.run_checked(50, 0)
        assert st.capacity > 4
        d.pos[0, :100] += 3.0          # beyond skin / 2
        d.run_checked(50, 50)
        res.append(d.particle_data_host())
    for a, b in zip(*res):
        assert np.array_equal(a, b)


def test_small_swarms_keep_the_resident_kernel_by_default():
    _, d = _pair(1000, 125.0, 0.0)
    d.run_checked(10, 0)
    assert d.tiling_stats()["tiled"] == 0


@pytest.mark.parametrize("mode", [2, 0])
def test_shared_rotational_pair_does_not_depend_on_where_a_launch_starts(mode):
    """The rotational normal of an odd step is the cosine kept from the even step before it (swarm_math.cuh): a launch that
    STARTS on an odd step, or ends on an even one, has no kept value and must draw the even step's block again -- so cutting a
    run at odd and even steps must not change a single bit, in the tiled kernel (mode 2: the value lives in shared memory) and
    in the resident one (mode 0: drawn afresh every step)."""
    _, whole = _pair(20000, 800.0, KT_300, skin=1.0)
    whole.set_tiling(mode)
    st = whole.run_checked(50, 0)
    assert st.steps_done == 50 and not st.overflow
    p0, v0, e0 = whole.particle_data_host()
    _, cut = _pair(20000, 800.0, KT_300, skin=1.0)
    cut.set_tiling(mode)
    step = 0
    for n in (7, 10, 1, 2, 30):   # launches starting on steps 0, 7, 17, 18, 20
        st = cut.run_checked(n, step)
        assert st.steps_done == n and not st.overflow
        step += n
    p1, v1, e1 = cut.particle_data_host()
    assert np.array_equal(p0, p1) and np.array_equal(e0, e1)
    assert cut.tiling_stats()["tiled"] == (1 if mode == 2 else 0)


def test_environment_kernel_keeps_the_cosine_across_an_odd_cut():
    """k_bd2d_env (batched environments, one warp each) holds the kept normal in a register: same check."""
    from swarmrl_b200.device import DeviceSwarm

    rng = np.random.default_rng(3)
    E, m, L = 64, 32, 40.0
    sp, wca = species_for([1.0]), wca_for([1.0])
    pos = np.concatenate([jittered_lattice(m, L, 2, rng) for _ in range(E)])
    th = rng.uniform(0, 2 * np.pi, E * m)
    gt, gr = sp[0]["gamma_t"][0], sp[0]["gamma_r"][0]

    def make():
        d = DeviceSwarm(2, (L, L, L), True, 1e-3, KT_300, 42, sp, wca, skin=0.4, n_envs=E)
        d.set_state(pos, np.zeros(E * m, dtype=np.int32), theta=th)
        d.set_actions_host(np.full(E * m, 5.0 * gt), np.tile([0, 0, gr * 4 * np.pi / 180], (E * m, 1)))
        return d

    a = make()
    a.run_checked(40, 0)
    b = make()
    step = 0
    for n in (5, 8, 1, 26):
        b.run_checked(n, step)
        step += n
    pa, _, ea = a.particle_data_host()
    pb, _, eb = b.particle_data_host()
    assert np.array_equal(pa, pb) and np.array_equal(ea, eb)
