"""
BASELINE.json's FULL sizes, where the oracle is too slow to replay everything: size-independent properties of the
same kernels (C4: 100 000 agents, resident kernel; C3: 4 096 x 32 agents, warp-per-environment kernel; a 1 M
slab, stream kernel) -- Newton's third law, neighbour-list symmetry, chunk invariance, noise moments, free
diffusion, observable idempotence -- plus an oracle spot check on a random subset of particles.
"""
import numpy as np
import pytest

import bench

pytestmark = pytest.mark.gpu


def _c4(kT=bench.KT, n=bench.N_AGENTS, skin=2.0, seed=42, n_envs=1, L=None):
    from swarmrl_b200.device import DeviceSwarm

    L = bench.BOX if L is None else L
    pos, th = bench.initial_state(n, L, seed)
    d = DeviceSwarm(2, (L,) * 3, True, bench.DT, kT, seed, bench.species_table(), bench.wca_table(), skin=skin, n_envs=n_envs)
    d.set_state(pos, np.zeros(n, dtype=np.int32), theta=th)
    return d, pos, th


def _push_together(d, n, L, frac=0.35):
    """strong inward swim so that the WCA forces are non-trivial after a short run"""
    import torch

    gt = bench.stokes(1.0)[0]
    d.f_swim[:n] = 400.0 * gt
    d.run_checked(150, 0)
    return d


def test_c4_newton_third_law_and_list_symmetry():
    d, pos, th = _c4(kT=0.0)
    n = bench.N_AGENTS
    _push_together(d, n, bench.BOX)
    d.f_swim[:n] = 0.0
    F = d.forces().cpu().numpy()
    assert np.abs(F).max() > 1.0                                      # there are contacts
    assert np.all(np.abs(F.sum(axis=1)) < 1e-9 * np.abs(F).sum(axis=1).max())   # sum of pair forces vanishes
    pairs = d.neighbors()
    a = set(map(tuple, pairs.tolist()))
    assert len(a) == len(pairs) and all((j, i) in a for i, j in list(a)[:200000])
    assert not any(i == j for i, j in list(a)[:200000])


@pytest.mark.parametrize("case", ["c4_resident", "c3_envs", "slab_1m_stream"])
def test_chunk_invariance_with_noise(case):
    """integrator.run(100) == run(37) + run(63): the Philox counter is the global step, so chunking (and the
    in-kernel rebuild schedule it changes) must not move a single bit."""
    if case == "c4_resident":
        mk = lambda: _c4()[0]
        n = bench.N_AGENTS
    elif case == "c3_envs":
        mk = lambda: _c4(n=4096 * 32, n_envs=4096, L=60.0 * 64, skin=1.0)[0]   # same density, 4096 blocks of 32
        n = 4096 * 32
    else:
        n = 1_000_000
        mk = lambda: _c4(n=n, L=float(np.sqrt(n * np.pi / 0.10)), skin=1.0)[0]
    gt = bench.stokes(1.0)[0]
    outs = []
    for chunks in ((100,), (37, 63)):
        d = mk()
        d.f_swim[:n] = 20.0 * gt
        s0 = 0
        for c in chunks:
            st = d.run_checked(c, s0)
            assert st.steps_done == c and not st.overflow
            s0 += c
        outs.append((d.pos[:2, :n].cpu().numpy(), d.theta[:n].cpu().numpy()))
        d.close()
    assert np.array_equal(outs[0][0], outs[1][0]) and np.array_equal(outs[0][1], outs[1][1])


def test_c4_noise_moments_and_free_diffusion():
    """100 000 free particles (no WCA contact on the lattice, no swim): <dx> = 0, MSD = 4 D t with D = kT / gamma_t,
    <cos dtheta> = exp(-D_r t), each within 5 standard errors; Gaussians of one step have unit variance."""
    d, pos, th = _c4(skin=1.0)
    n, steps = bench.N_AGENTS, 200
    g = d.gaussians(n, 3, 0).cpu().numpy()
    assert abs(g.mean()) < 5 / np.sqrt(4 * n) and abs(g.var() - 1.0) < 5 * np.sqrt(2.0 / (4 * n))
    assert abs(np.mean(g**4) - 3.0) < 5 * np.sqrt(96.0 / (4 * n))
    d.run_checked(steps, 0)
    p = d.pos[:2, :n].cpu().numpy().T
    dx = p - pos[:, :2]
    gt, gr = bench.stokes(1.0)
    t = steps * bench.DT
    D, Dr = bench.KT / gt, bench.KT / gr
    msd = (dx**2).sum(axis=1)
    # a few lattice neighbours touch during the run; the bound below tolerates that (< 1 % of particles)
    assert abs(dx.mean()) < 5 * np.sqrt(2 * D * t / n)
    assert abs(msd.mean() / (4 * D * t) - 1.0) < 0.02
    dth = d.theta[:n].cpu().numpy() - th
    assert abs(np.cos(dth).mean() - np.exp(-Dr * t)) < 5 * np.sqrt(0.5 * (1 - np.exp(-2 * Dr * t)) ** 2 / n) + 2e-3


def test_c4_observables_idempotent_and_oracle_spot_check():
    """Vision cones and the field reward at 100 000 observers: calling twice gives the same bits; 150 random
    observers agree with the oracle's O(N) scan per observer."""
    import torch

    from oracle.oracle import OracleSwarm

    d, pos, th = _c4(kT=0.0)
    n = bench.N_AGENTS
    row_of, rows = d.row_map(0)
    radii = torch.ones(n, dtype=torch.float64, device=d.dev)
    tc = torch.zeros(1, dtype=torch.int32, device=d.dev)
    v1 = d.obs_vision(row_of, rows, 10.0, np.pi / 2, 5, radii, tc, 1).clone()
    v2 = d.obs_vision(row_of, rows, 10.0, np.pi / 2, 5, radii, tc, 1)
    assert torch.equal(v1, v2) and float(v1.sum()) > 0
    o = OracleSwarm(2, (bench.BOX,) * 3, True, bench.DT, 0.0, 42, bench.species_table(), bench.wca_table())
    o.set_state(pos, np.zeros(n, dtype=np.int32), theta=th)
    pick = np.random.default_rng(0).choice(n, 150, replace=False)
    o.ptype = np.ones(n, dtype=np.int32); o.ptype[pick] = 0
    ref = o.vision_cones(0, 10.0, np.pi / 2, 5, np.ones(n), [0, 1]).sum(axis=2)        # both types seen
    got = v1.cpu().numpy()[np.sort(pick), :, 0]
    assert np.allclose(got, ref, rtol=1e-9, atol=1e-9)
