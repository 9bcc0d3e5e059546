"""
GPU parity tests of the observable / reward kernels against the oracle and the golden vectors
generated from the reference's own Python (oracle/make_golden.py), through the C-ABI.
"""
import numpy as np
import pytest

from helpers import jittered_lattice, make_pair, species_for, wca_for

pytestmark = pytest.mark.gpu


def _pair(n=3000, L=300.0, radii=(1.0, 1.0), seed=4):
    rng = np.random.default_rng(seed)
    sp, wca = species_for(list(radii)), wca_for(list(radii))
    pos = jittered_lattice(n, L, 2, rng)
    pos[::9, 0] += L  # unfolded coordinates: the reference takes raw differences
    th = rng.uniform(0, 2 * np.pi, n)
    sidx = (rng.uniform(size=n) < 0.3).astype(np.int32)
    return make_pair(2, (L,) * 3, True, 1e-3, 0.0, 1, sp, wca, pos, sidx, theta=th), rng


SPECIES8 = [dict(type=t, gamma_t=(1.0,) * 3, gamma_r=(1.0,) * 3, mass=1.0, rinertia=(1.0,) * 3) for t in range(8)]
ZERO_WCA = dict(sigma=np.zeros((8, 8)), eps=np.zeros((8, 8)), cut=np.zeros((8, 8)))


def _golden_swarm(pos, types, dirs=None):
    from swarmrl_b200.device import DeviceSwarm

    d = DeviceSwarm(2, (1.0, 1.0, 1.0), False, 1.0, 0.0, 0, SPECIES8, ZERO_WCA)
    th = np.zeros(len(pos)) if dirs is None else np.arctan2(dirs[:, 1], dirs[:, 0])
    d.set_state(pos, np.asarray(types, dtype=np.int32), theta=th)
    return d


@pytest.mark.parametrize("case", ["ref_unit", "rand64", "rand200_narrow", "rand150_wide"])
def test_vision_cones_golden(golden, case):
    """Fixtures produced by the reference's own SubdividedVisionCones (ref_unit is the literal case of
    CI/unit_tests/observables/test_subdivided_vision_cone.py:17-53), replayed on the device."""
    import torch

    g = golden("vision_cones.npz")
    pos, dirs, types = g[f"{case}/pos"], g[f"{case}/dirs"], g[f"{case}/types"]
    d = _golden_swarm(pos, types, dirs)
    det = [int(t) for t in g[f"{case}/detected_types"]]
    tc = -np.ones(8, dtype=np.int32)
    for c, t in enumerate(det):
        tc[t] = c
    row_of, rows = d.row_map(int(g[f"{case}/particle_type"]))
    out = d.obs_vision(row_of, rows, float(g[f"{case}/vision_range"]), float(g[f"{case}/half_angle"]),
                       int(g[f"{case}/n_cones"]), torch.from_numpy(g[f"{case}/radii"].astype(np.float64)).to(d.dev),
                       torch.from_numpy(tc).to(d.dev), len(det)).cpu().numpy()
    ref = g[f"{case}/obs"]
    assert out.shape == ref.shape and np.max(np.abs(out - ref)) < 1e-12
    if case == "ref_unit":
        assert np.allclose(out[0], [[1.0, 0.0], [0.8, 0.75], [0.0, 0.0]], atol=1e-15)


@pytest.mark.parametrize("n_cones,half_angle,rng_", [(5, np.pi / 2, 10.0), (3, np.pi, 25.0), (1, 0.3, 40.0)])
def test_vision_cones_match_oracle(n_cones, half_angle, rng_):
    import torch

    (o, d), rng = _pair()
    n = o.n
    radii = rng.uniform(0.5, 2.0, n)
    for want, det in ((0, [0, 1]), (1, [1])):
        row_of, rows = d.row_map(want)
        tc = -np.ones(2, dtype=np.int32)
        for c, t in enumerate(det):
            tc[t] = c
        out = d.obs_vision(row_of, rows, rng_, half_angle, n_cones, torch.from_numpy(radii).to(d.dev),
                           torch.from_numpy(tc).to(d.dev), len(det)).cpu().numpy()
        ref = o.vision_cones(want, rng_, half_angle, n_cones, radii, det)
        assert out.shape == ref.shape and ref.sum() > 0
        # rim tests are strict inequalities on acos(): allow an ulp-level reassignment of one neighbour
        bad = np.abs(out - ref) > 1e-9
        assert bad.sum() <= 2, bad.sum()


def test_field_observable_and_gradient_reward_match_oracle():
    (o, d), rng = _pair(n=5000)
    n = o.n
    L = 300.0
    import torch

    src = (0.5, 0.4, 0.0)
    hist_o = np.ascontiguousarray((o.pos / L).T.copy())
    hist_d = torch.from_numpy(hist_o.copy()).to(d.dev)
    hist_before = hist_o.copy()
    # move, then evaluate
    shift = rng.normal(size=(n, 3)) * 3.0
    shift[:, 2] = 0.0
    o.x += shift[:, 0]; o.y += shift[:, 1]
    d.pos[:, :n] += torch.from_numpy(np.ascontiguousarray(shift.T)).to(d.dev)
    row_of, rows = d.row_map(0)
    dn, dp = o.field_distances(0, src, (L, L, L), hist_o)
    out, gn, gp = d.obs_field(row_of, rows, src, (L, L, L), hist_d, decay=1, c0=1.0, c1=-1.0, scale=100.0, clip=True)
    assert np.max(np.abs(gn.cpu().numpy() - dn)) < 1e-14 and np.max(np.abs(gp.cpu().numpy() - dp)) < 1e-14
    ref = np.clip(100.0 * ((1 - dn) - (1 - dp)), 0, None)
    assert np.max(np.abs(out.cpu().numpy() - ref)) < 1e-11
    assert np.array_equal(hist_d.cpu().numpy()[:, o.ptype == 0], hist_o[:, o.ptype == 0])
    # rows of the other type were left alone
    assert np.array_equal(hist_d.cpu().numpy()[:, o.ptype == 1], hist_before[:, o.ptype == 1])
    # unclipped observable (ConcentrationField) and the other decay kinds
    for kind, f in ((2, lambda x: 2.0 * np.exp(-0.5 * x)), (3, lambda x: 2.0 / (x + 0.5))):
        h2 = torch.from_numpy(hist_before.copy()).to(d.dev)
        out, gn, gp = d.obs_field(row_of, rows, src, (L, L, L), h2, decay=kind, c0=2.0, c1=0.5, scale=3.0, clip=False)
        ref = 3.0 * (f(gn.cpu().numpy()) - f(gp.cpu().numpy()))
        assert np.max(np.abs(out.cpu().numpy() - ref)) < 1e-12


def test_field_reward_and_mean_in_one_launch():
    """swarm_obs_field_mean == swarm_obs_field followed by the mean: rewards bit for bit, (sum, mean) over the rows of the
    selected type in a fixed order (twice the same bits), history advanced once."""
    import torch

    (o, d), rng = _pair(n=5000)
    n, L, src = o.n, 300.0, (0.5, 0.4, 0.0)
    hist0 = np.ascontiguousarray((o.pos / L).T.copy())
    shift = rng.normal(size=(n, 3)) * 3.0
    shift[:, 2] = 0.0
    d.pos[:, :n] += torch.from_numpy(np.ascontiguousarray(shift.T)).to(d.dev)
    row_of, rows = d.row_map(0)
    assert 0 < rows < n                       # a proper subset: rows of the other type must not enter the sum
    for kind, c0, c1 in ((1, 1.0, -1.0), (2, 2.0, 0.5), (3, 2.0, 0.5)):
        ha = torch.from_numpy(hist0.copy()).to(d.dev)
        hb, hc = ha.clone(), ha.clone()
        ref, _, _ = d.obs_field(row_of, rows, src, (L, L, L), ha, decay=kind, c0=c0, c1=c1, scale=100.0, clip=True, want_dist=False)
        out, red = d.obs_field_mean(row_of, rows, src, (L, L, L), hb, kind, c0, c1, scale=100.0, clip=True)
        out2, red2 = d.obs_field_mean(row_of, rows, src, (L, L, L), hc, kind, c0, c1, scale=100.0, clip=True)
        r = ref.cpu().numpy()
        assert np.array_equal(out.cpu().numpy(), r) and torch.equal(ha, hb)
        assert r.sum() > 0.0
        got = red.cpu().numpy()
        assert abs(got[0] - r.sum()) <= 1e-12 * r.sum() and abs(got[1] - r.mean()) <= 1e-12 * r.mean()
        assert np.array_equal(got, red2.cpu().numpy())        # the ticket re-arms; same bits


@pytest.mark.parametrize("name,clip", [("gradient_sensing.npz", True), ("concentration_field.npz", False)])
def test_field_sequences_golden(golden, name, clip):
    """Sequences produced by the reference's GradientSensing / ConcentrationField (decay 1 - d)."""
    import torch

    g = golden(name)
    traj, types, box, src = g["traj"], g["types"], g["box"], g["src"]
    n = traj.shape[1]
    d = _golden_swarm(traj[0], types)
    hist = torch.from_numpy(np.ascontiguousarray((traj[0] / box).T)).to(d.dev)
    row_of, rows = d.row_map(int(g["particle_type"]))
    outs = []
    for s in range(1, traj.shape[0]):
        d.pos[:, :n] = torch.from_numpy(np.ascontiguousarray(traj[s].T)).to(d.dev)
        out, _, _ = d.obs_field(row_of, rows, src / box, box, hist, decay=1, c0=1.0, c1=-1.0,
                                scale=float(g["scale"]), clip=clip, want_dist=False)
        outs.append(out.cpu().numpy().copy())
    ref = g["rewards"] if clip else g["obs"][..., 0]
    assert np.max(np.abs(np.array(outs) - ref)) < (1e-12 if clip else 1e-9)


def test_gradient_sensing_reference_unit(golden):
    """CI/unit_tests/tasks/test_gradient_sensing.py:70-113 known answers through the device kernel."""
    import torch

    g = golden("gradient_sensing_unit.npz")
    p0, p1 = g["p0"], g["p1"]
    d = _golden_swarm(p1, np.zeros(3, dtype=np.int32))
    hist = torch.from_numpy(np.ascontiguousarray(p0.T.copy())).to(d.dev)
    row_of, rows = d.row_map(0)
    out, _, _ = d.obs_field(row_of, rows, (0.5, 0.5, 0.0), (1.0, 1.0, 1.0), hist, decay=1, c0=1.0, c1=-1.0, scale=1.0,
                            clip=True)
    r = out.cpu().numpy()
    assert np.max(np.abs(r - g["rewards"])) < 1e-15 and r[1] == 0.0
    assert np.array_equal(hist.cpu().numpy().T, p1)


def test_position_golden(golden):
    g = golden("position_director.npz")
    d = _golden_swarm(g["pos"], g["types"], g["dirs"])
    row_of, rows = d.row_map(1)
    out = d.obs_position(row_of, rows, g["box"]).cpu().numpy()
    assert np.max(np.abs(out - g["position_obs_type1"])) < 1e-15
    e = d.directors().cpu().numpy().T
    assert np.max(np.abs(e[g["types"] == 2] - g["director_obs_type2"])) < 1e-12


def test_position_observable_and_mean_reduction():
    (o, d), rng = _pair(n=4097)
    row_of, rows = d.row_map(1)
    out = d.obs_position(row_of, rows, (300.0, 300.0, 300.0)).cpu().numpy()
    assert np.array_equal(out, o.pos[o.ptype == 1] / 300.0)
    import torch

    for m in (0, 1, 31, 100003):
        v = rng.normal(size=m)
        r = d.reduce_mean(torch.from_numpy(v).to(d.dev)).cpu().numpy()
        assert abs(r[0] - v.sum()) < 1e-9 and (m == 0 or abs(r[1] - v.mean()) < 1e-12)
    # deterministic: identical bits run to run
    v = torch.from_numpy(rng.normal(size=50000)).to(d.dev)
    assert np.array_equal(d.reduce_mean(v).cpu().numpy(), d.reduce_mean(v).cpu().numpy())


PAIR_CASES = {  # name: (particle_type, sensing_type, decay, scale, clip_mode)
    "ps_same": (0, 0, ("Exponential", 1.0, 4.0), 100.0, 0),
    "ps_other": (0, 1, ("Inverse", 0.01, 0.05), 100.0, 0),
    "ss_same": (1, 1, ("Exponential", 1.0, 4.0), 100.0, 1),
    "ss_avoid": (1, 0, ("Inverse", 0.01, 0.05), 10.0, -1),
}


class _C:
    def __init__(self, pos, i, t):
        self.pos, self.director, self.id, self.type, self.velocity = pos, np.array([1.0, 0, 0]), i, int(t), np.zeros(3)


@pytest.mark.parametrize("name", sorted(PAIR_CASES))
def test_pair_field_golden_through_the_host_classes(golden, name):
    """ParticleSensing / SpeciesSearch (same constructor arguments as the reference) on lists of colloids,
    replaying the sequences the reference produced -- k_pair_field underneath."""
    from swarmrl_b200 import decay
    from swarmrl_b200.observables import ParticleSensing
    from swarmrl_b200.tasks import SpeciesSearch

    g = golden("pair_field.npz")
    pt, st, (dk, c0, c1), scale, clip = PAIR_CASES[name]
    fn = getattr(decay, dk)(c0, c1)
    traj, types, box = g["traj"], g["types"], g["box"]
    cols = lambda s: [_C(traj[s, i], i, types[i]) for i in range(len(types))]
    if clip == 0:
        ob = ParticleSensing(decay_fn=fn, box_length=box, sensing_type=st, scale_factor=scale, particle_type=pt)
        call = lambda c: ob.compute_observable(c).reshape(-1)
    else:
        ob = SpeciesSearch(decay_fn=fn, box_length=box, sensing_type=st, avoid=clip < 0, scale_factor=scale, particle_type=pt)
        call = ob
    with pytest.raises(ValueError):
        call(cols(1))                      # not initialised: the reference raises ValueError
    ob.initialize(cols(0))
    for s in range(1, traj.shape[0]):
        assert np.allclose(call(cols(s)), g[name][s - 1], rtol=1e-9, atol=1e-10), s
    assert len(ob.historical_field) == int(np.sum(types == pt))


def test_pair_field_matches_oracle_at_scale_and_multi_classes():
    """6000 particles, three types: all-pairs kernel vs the oracle; MultiSensing / MultiTasking on top."""
    import torch

    from oracle.oracle import pair_field
    from swarmrl_b200 import decay
    from swarmrl_b200.device_view import SwarmDeviceView
    from swarmrl_b200.observables import MultiSensing, ParticleSensing
    from swarmrl_b200.tasks import MultiTasking, SpeciesSearch

    rng = np.random.default_rng(8)
    n, L = 6000, 500.0
    pos = np.zeros((n, 3)); pos[:, :2] = rng.uniform(0, L, (n, 2))
    pos[17] = pos[4000]                                  # a coincident pair
    types = rng.integers(0, 3, n)
    d = _golden_swarm(pos, types)
    box = np.array([L, L, L])
    view = SwarmDeviceView(type("E", (), {"_backend": d, "_types": types})())   # an engine-shaped holder
    ps = [ParticleSensing(decay.Exponential(1.0, 6.0), box, sensing_type=t, scale_factor=1.0, particle_type=0) for t in (0, 1)]
    ms = MultiSensing(ps)
    mt = MultiTasking(0, [SpeciesSearch(decay.Inverse(0.02, 0.1), box, sensing_type=2, scale_factor=3.0, particle_type=0),
                          SpeciesSearch(decay.Linear(1.0, -1.0), box, sensing_type=1, avoid=True, scale_factor=2.0, particle_type=0)])
    ms.initialize(view); mt.initialize(view)
    f0 = [pair_field(pos, types, 0, t, box, 2, 1.0, 6.0) for t in (0, 1)]
    g0 = [pair_field(pos, types, 0, 2, box, 3, 0.02, 0.1), pair_field(pos, types, 0, 1, box, 1, 1.0, -1.0)]
    pos2 = pos.copy(); pos2[:, :2] += rng.normal(0, 5.0, (n, 2))
    d.pos[:, :n] = torch.from_numpy(np.ascontiguousarray(pos2.T)).to(d.dev)
    obs = ms.compute_observable_device(view).cpu().numpy()
    rew = mt.call_device(view).cpu().numpy()
    f1 = [pair_field(pos2, types, 0, t, box, 2, 1.0, 6.0) for t in (0, 1)]
    g1 = [pair_field(pos2, types, 0, 2, box, 3, 0.02, 0.1), pair_field(pos2, types, 0, 1, box, 1, 1.0, -1.0)]
    assert obs.shape == (int(np.sum(types == 0)), 2)
    for k in range(2):
        assert np.allclose(obs[:, k], f1[k] - f0[k], rtol=1e-9, atol=1e-9)
    want = 3.0 * np.clip(g1[0] - g0[0], 0, None) + 2.0 * np.clip(g1[1] - g0[1], None, 0)
    assert np.allclose(rew, want, rtol=1e-9, atol=1e-9)


def test_fp32_reference_bounds(golden):
    """The reference evaluates these observables in JAX float32; the kernels compute in fp64.
    Kernel output against the oracle's float32 evaluation mode (oracle.{vision_cones_f32, field_delta_f32}), with the
    per-observable bounds stated in tests/test_oracle_fp32.py: vision cones 1e-5 per cone sum except entries where a
    particle sits within float32 rounding of a rim / of the range (< 0.5 % of the entries, off by <= one weight);
    field observables 8 eps32 * scale * (|c0| + |c1| d_max)."""
    import torch

    from oracle import oracle

    (o, d), rng = _pair(n=2500, L=260.0)
    n = o.n
    radii = rng.uniform(0.5, 1.5, n)
    row_of, rows = d.row_map(0)
    tc = torch.from_numpy(np.array([0, 1], dtype=np.int32)).to(d.dev)
    out = d.obs_vision(row_of, rows, 12.0, np.pi / 2, 5, torch.from_numpy(radii).to(d.dev), tc, 2).cpu().numpy()
    ref32 = oracle.vision_cones_f32(o.pos, o.directors, o.ptype, 0, 12.0, np.pi / 2, 5, radii, [0, 1]).astype(np.float64)
    diff = np.abs(out - ref32)
    flipped = diff > 1e-5
    assert out.shape == ref32.shape and flipped.mean() < 0.005 and np.all(diff[flipped] <= 1.0 + 1e-5)
    assert diff[~flipped].max() > 0.0          # the two precisions differ: the bound is not vacuous
    # field observable, scale 1e4 as in the reference's RL tests
    L, scale = 260.0, 10000.0
    src = np.array([0.5, 0.5, 0.0])
    prev = o.pos / L
    hist = torch.from_numpy(np.ascontiguousarray(prev.T)).to(d.dev)
    d.pos[:2, :n] += torch.from_numpy(rng.normal(size=(2, n)) * 0.5).to(d.dev)
    now = d.pos[:, :n].cpu().numpy().T / L
    got, _, _ = d.obs_field(row_of, rows, src, (L, L, L), hist, decay=1, c0=1.0, c1=-1.0, scale=scale, clip=False, want_dist=False)
    sel = o.ptype == 0
    ref = oracle.field_delta_f32(now[sel], prev[sel], src, 1, 1.0, -1.0, scale, False).astype(np.float64)
    tol = 8 * 2.0 ** -24 * scale * (1.0 + 1.0 * 3.0)
    assert np.max(np.abs(got.cpu().numpy() - ref)) <= tol


def test_vision_fixed_capacity_layout_second_call_moved_particles_and_overflow():
    """The vision grid (k_vis2_*: fixed capacity per cell) is laid over the bounding box of the PREVIOUS call -- the first
    call of a swarm measures its own.  Same answers as the oracle, call after call (bit-identical while nothing moves),
    also after particles have left the previous box by more than a range (they are clamped into edge cells) and with
    more particles in one cell than its capacity (overflow list)."""
    import torch

    (o, d), rng = _pair(n=3000, L=300.0)
    n = o.n
    radii = np.ones(n)
    rd = torch.from_numpy(radii).to(d.dev)
    tc = torch.from_numpy(np.array([0, 1], dtype=np.int32)).to(d.dev)
    row_of, rows = d.row_map(0)

    def both():
        got = d.obs_vision(row_of, rows, 12.0, np.pi / 2, 5, rd, tc, 2).cpu().numpy()
        ref = o.vision_cones(0, 12.0, np.pi / 2, 5, radii, [0, 1])
        assert np.max(np.abs(got - ref)) < 1e-11
        return got

    first = both()                    # box of the current positions
    second = both()                   # box of the previous call
    assert np.array_equal(first, second)
    assert d.stats().launches > 0
    # move: a block of particles far outside the previous box, a dense clump of 120 inside one cell
    shift = np.zeros((n, 3)); shift[:40, 0] = 80.0; shift[40:80, 1] = -95.0
    clump = np.arange(200, 320)
    newpos = o.pos + shift
    newpos[clump, :2] = np.array([150.0, 150.0]) + rng.uniform(-3.0, 3.0, (len(clump), 2))
    o.x[:], o.y[:] = newpos[:, 0], newpos[:, 1]
    d.pos[:2, :n] = torch.from_numpy(np.ascontiguousarray(newpos[:, :2].T)).to(d.dev)
    both()
    both()                            # and once more on the box of the moved swarm
