"""
Pins the engine half of the oracle (ESPResSo is not in /root/reference: SURVEY.md 8c) against
the deterministic known answers the reference's own ESPResSo tests assert, and against analytic
results where the reference has none ("parity unpinned" items of oracle/swarm_oracle.c).

  v = f e / gamma_t, x(t) = x0 + v t      CI/espresso_tests/unit_tests/test_espresso.py:102-111
  WCA cutoff = r_a + r_b                    test_espresso.py:114-118, test_rod.py:69-74
  anisotropic v, omega = tau / gamma_r      test_espresso_2d.py:165-183
  2-D planarity under noise                 test_espresso_2d.py:60-81
  exact placement survives the relax        test_add_colloid_on_point.py:93-99
"""
import numpy as np
import pytest

from helpers import KT_300, jittered_lattice, random_quats, species_for, stokes, wca_for
from oracle.oracle import OracleSwarm


def test_const_force_kinematics_3d():
    rng = np.random.default_rng(1)
    radii = [1.0, 2.0]
    sp, wca = species_for(radii), wca_for(radii)
    n = 5
    pos = jittered_lattice(n, 1000.0, 3, rng) 
    q = random_quats(n, rng)
    o = OracleSwarm(3, (1000.0,) * 3, True, 0.01, 0.0, 42, sp, wca)
    o.set_state(pos, np.array([0, 0, 0, 1, 1], dtype=np.int32), quat=q)
    o.f_swim[:] = 1.234
    x0, e0 = o.pos.copy(), o.directors.copy()
    o.bd_steps(0, 1000)
    gam = np.array([stokes(radii[t])[0] for t in (0, 0, 0, 1, 1)])
    v = 1.234 * e0 / gam[:, None]
    assert np.allclose(o.velocities, v, rtol=1e-13)
    assert np.allclose(o.pos, x0 + 10.0 * v, rtol=2e-6)  # the reference's own tolerance
    assert np.allclose(o.pos, x0 + 10.0 * v, rtol=1e-11)
    assert np.allclose(o.directors, e0, atol=1e-15)


def test_wca_cutoff_is_sum_of_radii():
    radii = [1.0, 2.5]
    wca = wca_for(radii)
    assert np.allclose(wca["cut"], np.add.outer(radii, radii), rtol=1e-15)
    sp = species_for(radii)
    for d, expect_force in [(3.5 - 1e-9, True), (3.5 + 1e-9, False)]:
        o = OracleSwarm(3, (100.0,) * 3, False, 1e-3, 0.0, 1, sp, wca)
        o.set_state(np.array([[10.0, 10, 10], [10.0 + d, 10, 10]]), np.array([0, 1], dtype=np.int32))
        F = o.forces()
        assert (np.abs(F[:, 0]) > 0).all() == expect_force
        assert np.allclose(F[0], -F[1])


def test_wca_force_formula():
    """F = 48 eps [(s/r)^12 - (s/r)^6 / 2] / r along the separation; repulsive."""
    radii = [1.0]
    wca, sp = wca_for(radii), species_for(radii)
    r = 1.7
    o = OracleSwarm(2, (50.0,) * 3, True, 1e-3, 0.0, 1, sp, wca)
    o.set_state(np.array([[49.5, 3.0, 0], [49.5 + r - 50.0, 3.0, 0]]), np.zeros(2, dtype=np.int32))  # across the boundary
    F = o.forces(with_swim=False)
    s, eps = wca["sigma"][0, 0], wca["eps"][0, 0]
    mag = 48 * eps * ((s / r) ** 12 - 0.5 * (s / r) ** 6) / r
    assert mag > 0
    assert np.allclose(F[0], [-mag, 0, 0], rtol=1e-12) and np.allclose(F[1], [mag, 0, 0], rtol=1e-12)


def test_anisotropic_friction_and_torque_2d():
    """test_espresso_2d.py:94-185: v = (1/g_eq, 1/g_ax, 0) for unit force (1,1,0) with the director
    along y; omega_z = 1/g_rot under unit torque."""
    g_ax, g_eq, g_rot = 0.7, 1.9, 2.3
    sp = [dict(type=0, gamma_t=(g_eq, g_eq, g_ax), gamma_r=(g_rot, g_rot, 9.9), mass=1.0, rinertia=(1.0,) * 3)]
    o = OracleSwarm(2, (100.0,) * 3, True, 0.01, 0.0, 1, sp, wca_for([1.0]))
    o.set_state(np.array([[50.0, 50, 0]]), np.zeros(1, dtype=np.int32), theta=np.array([np.pi / 2]))
    o.ext_force[0][:] = 1.0
    o.ext_force[1][:] = 1.0
    o.bd_steps(0, 1)
    assert np.allclose(o.velocities[0], [1 / g_eq, 1 / g_ax, 0.0], rtol=1e-14, atol=1e-16)
    o.ext_force[0][:] = 0.0
    o.ext_force[1][:] = 0.0
    o.torque[2][:] = 1.0
    th0 = o.theta.copy()
    o.bd_steps(1, 10)
    assert np.allclose(o.omegas[0], [0, 0, 1 / g_rot])
    assert np.allclose(o.theta - th0, 10 * 0.01 / g_rot)


def test_2d_stays_planar_under_noise():
    rng = np.random.default_rng(3)
    sp, wca = species_for([1.0]), wca_for([1.0])
    n = 64
    pos = jittered_lattice(n, 40.0, 2, rng)
    o = OracleSwarm(2, (40.0,) * 3, True, 1e-3, KT_300, 42, sp, wca)
    th0 = rng.uniform(0, 2 * np.pi, n)
    o.set_state(pos, np.zeros(n, dtype=np.int32), theta=th0)
    o.bd_steps(0, 50)
    assert np.all(o.z == 0.0) and np.all(o.directors[:, 2] == 0.0)
    assert np.all(o.theta != th0) and np.all(o.velocities[:, 2] == 0.0)


def test_cell_list_equals_brute_force_bitwise():
    rng = np.random.default_rng(5)
    for n_dims, n, L in [(2, 400, 45.0), (3, 300, 14.0)]:
        radii = [1.0, 0.6]
        sp, wca = species_for(radii), wca_for(radii)
        pos = np.zeros((n, 3))
        pos[:, :n_dims] = rng.uniform(-L, 2 * L, (n, n_dims))  # unfolded, overlapping on purpose
        spx = rng.integers(0, 2, n).astype(np.int32)
        a = OracleSwarm(n_dims, (L,) * 3, True, 1e-4, 0.0, 1, sp, wca)
        a.set_state(pos, spx, theta=rng.uniform(0, 6, n), quat=random_quats(n, rng))
        assert np.array_equal(a.forces(mode=0), a.forces(mode=1))
        assert np.array_equal(a.forces(mode=0), a.forces(mode=2))


def test_relax_keeps_non_overlapping_placement_and_separates_overlaps():
    sp, wca = species_for([1.0]), wca_for([1.0])
    o = OracleSwarm(2, (100.0,) * 3, True, 1e-3, 0.0, 1, sp, wca)
    pos = np.array([[10.0, 10, 0], [13.0, 10, 0], [50.0, 50, 0], [50.5, 50, 0]])
    o.set_state(pos, np.zeros(4, dtype=np.int32))
    o.sd_steps(1000, 0.1, 0.1)
    assert np.array_equal(o.pos[:2], pos[:2])  # test_add_colloid_on_point.py:93-99
    assert np.linalg.norm(o.pos[2] - o.pos[3]) >= 2.0 - 1e-6
    assert np.allclose(o.pos[2] + o.pos[3], pos[2] + pos[3])


@pytest.mark.parametrize("n_dims", [2, 3])
def test_free_diffusion_msd_and_orientation_decay(n_dims):
    """parity-unpinned item pinned analytically: MSD = 2 d D_t t, <e(0).e(t)> = exp(-(d-1) D_r t)."""
    n, steps, dt = 4000, 200, 5e-3
    sp = species_for([1.0])
    zero = dict(sigma=np.zeros((1, 1)), eps=np.zeros((1, 1)), cut=np.zeros((1, 1)))
    o = OracleSwarm(n_dims, (1e4,) * 3, True, dt, KT_300, 7, sp, zero)
    rng = np.random.default_rng(0)
    pos = np.zeros((n, 3))
    pos[:, :n_dims] = rng.uniform(0, 1e4, (n, n_dims))
    o.set_state(pos, np.zeros(n, dtype=np.int32), theta=rng.uniform(0, 2 * np.pi, n), quat=random_quats(n, rng))
    x0, e0 = o.pos.copy(), o.directors.copy()
    o.bd_steps(0, steps)
    t = steps * dt
    gt, gr = stokes(1.0)
    msd = np.mean(np.sum((o.pos - x0) ** 2, axis=1))
    want = 2 * n_dims * KT_300 / gt * t
    assert abs(msd / want - 1) < 4 * np.sqrt(2.0 / (n * n_dims))
    c = np.mean(np.sum(o.directors * e0, axis=1))
    want_c = np.exp(-(n_dims - 1) * KT_300 / gr * t)
    assert abs(c - want_c) < 4 / np.sqrt(n)
    # velocity noise of the last step: <v^2> = d kT/m (espresso.py:402-405)
    v2 = np.mean(np.sum(o.velocities**2, axis=1))
    assert abs(v2 / (n_dims * KT_300 / sp[0]["mass"]) - 1) < 0.1


def test_verlet_and_cell_trajectories_equal_brute_force_bitwise():
    """The three neighbour strategies of the oracle visit pairs in the same (ascending j) order, so
    whole noisy trajectories agree bit for bit -- the CUDA path is later held to 1e-9 of this."""
    from oracle import oracle

    rng = np.random.default_rng(9)
    n, L = 256, 36.0
    sp, wca = species_for([1.0]), wca_for([1.0])
    pos = jittered_lattice(n, L, 2, rng)
    th = rng.uniform(0, 2 * np.pi, n)
    res = []
    for mode in (0, 1, 2):
        o = OracleSwarm(2, (L,) * 3, True, 1e-3, KT_300, 3, sp, wca)
        o.set_state(pos, np.zeros(n, dtype=np.int32), theta=th)
        o.f_swim[:] = 40.0
        o.bd_steps(0, 300, mode=mode)
        res.append((o.pos.copy(), o.theta.copy(), o.velocities.copy()))
    assert 1 < oracle.last_verlet_builds() < 300
    for r in res[1:]:
        assert all(np.array_equal(a, b) for a, b in zip(res[0], r))


# ---- Langevin thermostat + velocity Verlet (espresso.py:1187-1201): parity-unpinned, pinned analytically ----
def _langevin_species(radius=1.0, a=0.02, dt=1e-3):
    """Stokes friction; mass / inertia chosen so that gamma dt / m = a (the reference demands explicit values,
    espresso.py:391-400: water-density colloids relax in 2.5e-7 s, far below any usable time step)."""
    gt, gr = stokes(radius)
    return [dict(type=0, gamma_t=(gt,) * 3, gamma_r=(gr,) * 3, mass=gt * dt / a, rinertia=(gr * dt / a,) * 3)]


@pytest.mark.parametrize("n_dims", [2, 3])
def test_langevin_terminal_velocity_and_relaxation_time(n_dims):
    """kT = 0: v(t) = (f/gamma)(1 - exp(-gamma t / m)) e, omega(t) likewise with tau/gamma_r and I/gamma_r."""
    dt, a = 1e-3, 0.01
    sp = _langevin_species(a=a, dt=dt)
    zero = dict(sigma=np.zeros((1, 1)), eps=np.zeros((1, 1)), cut=np.zeros((1, 1)))
    o = OracleSwarm(n_dims, (1e3,) * 3, True, dt, 0.0, 7, sp, zero)
    o.set_state(np.array([[500.0, 500.0, 0.0 if n_dims == 2 else 500.0]]), np.zeros(1, dtype=np.int32),
                theta=np.array([0.3]), quat=np.array([[1.0, 0, 0, 0]]))
    gt, gr = stokes(1.0)
    o.f_swim[:] = 2.0 * gt
    o.lv_steps(0, 100)  # t = m / gamma
    e = o.directors[0]
    assert abs(np.dot(o.velocities[0], e) / 2.0 - (1 - np.exp(-1.0))) < 5e-3
    o.lv_steps(100, 3000)  # 31 relaxation times
    assert np.allclose(o.velocities[0], 2.0 * e, rtol=1e-9, atol=1e-12)
    # rotation about lab z: in 2-D this is the only rotational degree of freedom; in 3-D the director is body z
    o.f_swim[:] = 0.0
    o.torque[2][:] = 0.5 * gr
    o.lv_steps(3100, 100)
    assert abs(o.omegas[0, 2] / 0.5 - (1 - np.exp(-1.0))) < 5e-3
    o.lv_steps(3200, 3000)
    assert abs(o.omegas[0, 2] - 0.5) < 1e-9
    if n_dims == 2:
        assert np.all(o.pos[:, 2] == 0.0) and np.all(o.omegas[:, :2] == 0.0)


@pytest.mark.parametrize("n_dims", [2, 3])
def test_langevin_equipartition_and_diffusion(n_dims):
    """<v^2> = d kT/m, <omega^2> = d_r kT/I, MSD = 2 d (kT/gamma)(t - m/gamma) for t >> m/gamma."""
    n, steps, dt, a = 4000, 1500, 1e-3, 0.04
    sp = _langevin_species(a=a, dt=dt)
    zero = dict(sigma=np.zeros((1, 1)), eps=np.zeros((1, 1)), cut=np.zeros((1, 1)))
    o = OracleSwarm(n_dims, (1e4,) * 3, True, dt, KT_300, 11, sp, zero)
    rng = np.random.default_rng(3)
    pos = np.zeros((n, 3))
    pos[:, :n_dims] = rng.uniform(0, 1e4, (n, n_dims))
    o.set_state(pos, np.zeros(n, dtype=np.int32), theta=rng.uniform(0, 2 * np.pi, n), quat=random_quats(n, rng))
    x0 = o.pos.copy()
    o.lv_steps(0, steps)
    gt, gr = stokes(1.0)
    m, I = sp[0]["mass"], sp[0]["rinertia"][0]
    tol = 4 * np.sqrt(2.0 / (n * n_dims)) + a  # statistics + O(gamma dt / m) discretisation bias
    v2 = np.mean(np.sum(o.velocities**2, axis=1))
    assert abs(v2 / (n_dims * KT_300 / m) - 1) < tol
    d_rot = 1 if n_dims == 2 else 3
    w2 = np.mean(np.sum(o.omegas**2, axis=1))
    assert abs(w2 / (d_rot * KT_300 / I) - 1) < 4 * np.sqrt(2.0 / (n * d_rot)) + a
    t = steps * dt
    msd = np.mean(np.sum((o.pos - x0) ** 2, axis=1))
    want = 2 * n_dims * KT_300 / gt * (t - m / gt)
    assert abs(msd / want - 1) < tol
    if n_dims == 2:
        assert np.all(o.pos[:, 2] == 0.0)


def test_langevin_chunks_restart_with_a_force_recalculation():
    """Every integrator.run() recomputes the forces at the current (x, v) (particle properties changed in
    manage_forces): with kT = 0 the recalculated force equals the one that closed the previous chunk except
    for the friction term, which now sees the full-step velocity -- a second-order difference."""
    dt = 1e-3
    sp = _langevin_species(a=0.02, dt=dt)
    rng = np.random.default_rng(5)
    pos = jittered_lattice(200, 60.0, 2, rng)
    th = rng.uniform(0, 2 * np.pi, 200)
    out = []
    for chunks in ((0, 60),), ((0, 25), (25, 35)):
        o = OracleSwarm(2, (60.0,) * 3, True, dt, 0.0, 1, sp, wca_for([1.0]))
        o.set_state(pos, np.zeros(200, dtype=np.int32), theta=th)
        o.f_swim[:] = 3.0 * stokes(1.0)[0]
        for s0, k in chunks:
            o.lv_steps(s0, k, mode=2)
        out.append(o.pos.copy())
    assert not np.array_equal(out[0], out[1])
    assert np.abs(out[0] - out[1]).max() < 1e-4


def test_2d_brownian_step_takes_its_normals_as_documented():
    """The noise rule of a 2-D Brownian step (swarm_math.cuh, orc_bd_steps): translation = pair 0 of the draw of step s
    (g[0], g[1] of orc_gaussians(seed, id, s, 0)); rotation = pair 1 of the draw of the EVEN step s & ~1 -- its sine
    (g[3]) on the even step, its cosine (g[2]) on the odd one.  Free particles (no force, no torque): every increment is
    sigma x one of those normals, so the rule is pinned value by value; and sine / cosine of one pair are uncorrelated."""
    from oracle import oracle

    sp, wca = species_for([1.0]), wca_for([1.0])
    n, dt, seed, step0 = 64, 1e-3, 42, 7      # an odd first step: its rotational normal comes from the draw of step 6
    rng = np.random.default_rng(0)
    pos = jittered_lattice(n, 2000.0, 2, rng)
    o = OracleSwarm(2, (2000.0,) * 3, True, dt, KT_300, seed, sp, wca)
    o.set_state(pos, np.zeros(n, dtype=np.int32), theta=np.zeros(n))
    gt, gr = sp[0]["gamma_t"][0], sp[0]["gamma_r"][0]
    sg_t, sg_r = np.sqrt(2 * KT_300 * dt / gt), np.sqrt(2 * KT_300 * dt / gr)
    for s in range(step0, step0 + 6):
        x0, th0 = o.pos.copy(), np.arctan2(o.directors[:, 1], o.directors[:, 0])
        o.bd_steps(s, 1)
        th1 = np.arctan2(o.directors[:, 1], o.directors[:, 0])
        g_own = np.array([oracle.gaussians(seed, i, s, 0) for i in range(n)])
        g_even = np.array([oracle.gaussians(seed, i, s & ~1, 0) for i in range(n)])
        assert np.allclose(o.pos[:, 0] - x0[:, 0], sg_t * g_own[:, 0], rtol=0, atol=1e-9 * sg_t)
        assert np.allclose(o.pos[:, 1] - x0[:, 1], sg_t * g_own[:, 1], rtol=0, atol=1e-9 * sg_t)
        want = sg_r * (g_even[:, 2] if s & 1 else g_even[:, 3])
        dth = (th1 - th0 + np.pi) % (2 * np.pi) - np.pi
        assert np.allclose(dth, want, rtol=0, atol=1e-9)
    g = np.array([oracle.gaussians(seed, i, 2 * k, 0) for i in range(200) for k in range(50)])
    c = np.corrcoef(g[:, 2], g[:, 3])[0, 1]
    assert abs(c) < 4.0 / np.sqrt(len(g))
    assert abs(g[:, 2].var() - 1) < 0.06 and abs(g[:, 3].var() - 1) < 0.06
