"""
Generate tests/golden/*.npz by running the REFERENCE's own Python
(/root/reference/swarmrl/{observables,tasks,utils,force_functions}) under the
NumPy shim of oracle/ref_shim.py on seeded inputs.

Run in the build container only:   python oracle/make_golden.py
The fixtures are committed; the GPU box (no /root/reference) only reads them.

Each fixture stores the inputs and the reference's outputs:
  vision_cones.npz        SubdividedVisionCones.compute_observable
                          (subdivided_vision_cones.py:240-258), incl. the reference's
                          own unit-test case (CI/unit_tests/observables/
                          test_subdivided_vision_cone.py:17-53)
  signed_angle.npz        calc_signed_angle_between_directors (utils.py:297-332) on the
                          9 cases of CI/unit_tests/utils/test_utils.py:79-119 + random
  concentration_field.npz ConcentrationField over a 4-call sequence
                          (concentration_field.py:66-138)
  gradient_sensing.npz    GradientSensing over a 4-call sequence
                          (gradient_sensing.py:61-150), incl. test_gradient_sensing.py:70-113
  position_director.npz   PositionObservable / Director (position.py:32-60, director.py:30-58)
  force_fn_routing.npz    ForceFunction.calc_action per-type routing (force_fn.py:52-86)
  pair_field.npz          ParticleSensing.compute_observable (particle_sensing.py:95-178) and
                          SpeciesSearch.__call__ (species_search.py:97-184) over call sequences:
                          observer among the sensed set, observer of another type (the
                          nonzero(size=n-1) quirk drops the LAST sensed particle), coincident
                          particles (padding re-reads sensed[0]), avoid=True
"""
from __future__ import annotations

import pathlib
import sys
import warnings

import numpy as np

ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle import ref_shim  # noqa: E402

OUT = ROOT / "tests" / "golden"


def _colloids(Colloid, pos, dirs, types):
    return [
        Colloid(pos=np.array(pos[i], dtype=float), director=np.array(dirs[i], dtype=float), id=i,
                velocity=np.zeros(3), type=int(types[i]))
        for i in range(len(pos))
    ]


def vision_cones():
    VC = ref_shim.ref("observables.subdivided_vision_cones").SubdividedVisionCones
    Colloid = ref_shim.ref("components.colloid").Colloid
    cases = {}
    # the reference's unit test
    pos = np.array([[0, 0, 0], [0, 5, 0], [0, 8, 0], [-7, 8, 0], [1, 1, 0]], dtype=float)
    dirs = np.array([[0, 1, 0], [1, 0, 0], [1, 0, 0], [0, 1, 0], [0, 1, 0]], dtype=float)
    types = np.array([0, 0, 1, 1, 0])
    radii = np.array([1, 2, 3, 4, 1], dtype=float)
    cases["ref_unit"] = dict(pos=pos, dirs=dirs, types=types, radii=radii, vision_range=10.0,
                             half_angle=np.pi / 2, n_cones=3, particle_type=0)
    rng = np.random.default_rng(20261104)
    for name, n, L, n_cones, half, rngv, ntypes, ptype in [
        ("rand64", 64, 40.0, 5, np.pi / 2, 10.0, 2, 0),
        ("rand200_narrow", 200, 60.0, 7, 0.6, 12.5, 3, 1),
        ("rand150_wide", 150, 50.0, 4, np.pi, 9.0, 1, 0),
    ]:
        pos = np.zeros((n, 3))
        pos[:, :2] = rng.uniform(-0.3 * L, 1.3 * L, size=(n, 2))  # unfolded positions may leave the box
        ang = rng.uniform(0, 2 * np.pi, n)
        dirs = np.stack([np.cos(ang), np.sin(ang), np.zeros(n)], axis=1)
        types = rng.integers(0, ntypes, n)
        types[0] = ptype
        radii = rng.uniform(0.5, 2.0, n)
        cases[name] = dict(pos=pos, dirs=dirs, types=types, radii=radii, vision_range=rngv,
                           half_angle=half, n_cones=n_cones, particle_type=ptype)
    out = {}
    for name, c in cases.items():
        vc = VC(vision_range=c["vision_range"], vision_half_angle=c["half_angle"], n_cones=c["n_cones"],
                radii=list(c["radii"]), particle_type=c["particle_type"])
        cols = _colloids(Colloid, c["pos"], c["dirs"], c["types"])
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            obs = np.array(vc.compute_observable(cols))
        for k, v in c.items():
            out[f"{name}/{k}"] = np.asarray(v)
        out[f"{name}/detected_types"] = np.asarray(vc.detected_types)
        out[f"{name}/obs"] = obs
    np.savez_compressed(OUT / "vision_cones.npz", **out)
    assert np.array_equal(out["ref_unit/obs"][0], np.array([[1.0, 0], [0.8, 0.75], [0, 0]]))


def signed_angle():
    fn = ref_shim.ref("utils.utils").calc_signed_angle_between_directors
    a1, a2 = np.array([1.0, 0, 0]), np.array([-1 / np.sqrt(2), -1 / np.sqrt(2), 0])
    o1, o2, o3 = np.array([1.0, 0, 0]), np.array([0, 1.0, 0]), np.array([0.5, np.sqrt(3) / 2, 0])
    mine = [a1, a1, a1, a1, a2, a2, a2, a1, a2 * 1.5]
    other = [o1, -o1, o2, o3, o1, o2, o3, o3 * 4.5, o3]
    rng = np.random.default_rng(7)
    for _ in range(64):
        t, u = rng.uniform(0, 2 * np.pi, 2)
        mine.append(np.array([np.cos(t), np.sin(t), 0.0]))
        other.append(np.array([np.cos(u), np.sin(u), 0.0]) * rng.uniform(0.5, 3))
    mine, other = np.array(mine), np.array(other)
    ang = np.array([fn(m.copy(), o.copy()) for m, o in zip(mine, other)])
    np.savez_compressed(OUT / "signed_angle.npz", mine=mine, other=other, angle=ang)
    assert ang[0] == 0 and ang[1] == np.pi


def field_sequences():
    CF = ref_shim.ref("observables.concentration_field").ConcentrationField
    GS = ref_shim.ref("tasks.searching.gradient_sensing").GradientSensing
    Colloid = ref_shim.ref("components.colloid").Colloid
    rng = np.random.default_rng(11)
    n, steps = 24, 4
    box = np.array([1000.0, 1000.0, 1000.0])
    src = np.array([500.0, 500.0, 0.0])
    types = rng.integers(0, 2, n)
    traj = np.zeros((steps + 1, n, 3))
    traj[0, :, :2] = rng.uniform(100, 900, (n, 2))
    for s in range(steps):
        traj[s + 1] = traj[s]
        traj[s + 1, :, :2] += rng.normal(0, 30, (n, 2))
    dirs = np.tile([1.0, 0, 0], (n, 1))

    def decay(d):
        return 1 - d

    cf = CF(source=src, decay_fn=decay, box_length=box, scale_factor=10000, particle_type=0)
    gs = GS(source=src, decay_function=decay, box_length=box, reward_scale_factor=10, particle_type=0)
    cf.initialize(_colloids(Colloid, traj[0], dirs, types))
    gs.initialize(_colloids(Colloid, traj[0], dirs, types))
    obs, rew = [], []
    for s in range(1, steps + 1):
        cols = _colloids(Colloid, traj[s], dirs, types)
        obs.append(np.asarray(cf.compute_observable(cols)))
        rew.append(np.asarray(gs(cols)))
    np.savez_compressed(OUT / "concentration_field.npz", traj=traj, types=types, box=box, src=src,
                        scale=10000.0, particle_type=0, obs=np.array(obs))
    np.savez_compressed(OUT / "gradient_sensing.npz", traj=traj, types=types, box=box, src=src,
                        scale=10.0, particle_type=0, rewards=np.array(rew))

    # the reference's own unit case (test_gradient_sensing.py:70-113)
    gs2 = GS(source=np.array([0.5, 0.5, 0.0]), decay_function=decay, box_length=np.array([1.0, 1.0, 1.0]),
             reward_scale_factor=1, particle_type=0)
    p0 = np.array([[0, 0, 0], [0, 0.6, 0], [1.0, 0, 0]], dtype=float)
    p1 = np.array([[0.2, 0.2, 0], [0, 1.0, 0], [0, 1.0, 0]], dtype=float)
    d3 = np.tile([0, 1.0, 0], (3, 1))
    gs2.initialize(_colloids(Colloid, p0, d3, [0, 0, 0]))
    r = np.asarray(gs2(_colloids(Colloid, p1, d3, [0, 0, 0])))
    np.savez_compressed(OUT / "gradient_sensing_unit.npz", p0=p0, p1=p1, rewards=r)
    assert r[1] == 0.0 and r[0] > 0


def position_director():
    PO = ref_shim.ref("observables.position").PositionObservable
    DI = ref_shim.ref("observables.director").Director
    Colloid = ref_shim.ref("components.colloid").Colloid
    rng = np.random.default_rng(3)
    n = 17
    pos = rng.uniform(0, 1000, (n, 3))
    ang = rng.uniform(0, 2 * np.pi, n)
    dirs = np.stack([np.cos(ang), np.sin(ang), np.zeros(n)], axis=1)
    types = rng.integers(0, 3, n)
    box = np.array([1000.0, 800.0, 500.0])
    cols = _colloids(Colloid, pos, dirs, types)
    po = np.array(PO(box_length=box, particle_type=1).compute_observable(cols))
    di = np.array(DI(particle_type=2).compute_observable(cols))
    np.savez_compressed(OUT / "position_director.npz", pos=pos, dirs=dirs, types=types, box=box,
                        position_obs_type1=po, director_obs_type2=di)


def force_fn_routing():
    FF = ref_shim.ref("force_functions.force_fn").ForceFunction
    dm = ref_shim.ref("agents.dummy_models")
    Colloid = ref_shim.ref("components.colloid").Colloid
    types = np.array([0, 1, 0, 2, 1, 0, 2])
    n = len(types)
    pos = np.arange(3 * n, dtype=float).reshape(n, 3)
    dirs = np.tile([1.0, 0, 0], (n, 1))

    class PerType(dm.ConstForce):
        """returns one action per colloid OF ITS TYPE, force = 100*type + running count"""

        def __init__(self, t):
            self.t = t

        def calc_action(self, colloids):
            Action = ref_shim.ref("actions.actions").Action
            k = [c for c in colloids if c.type == self.t]
            return [Action(force=100.0 * self.t + i) for i, _ in enumerate(k)]

    ff = FF({"0": PerType(0), "2": PerType(2)})
    acts = ff.calc_action(_colloids(Colloid, pos, dirs, types))
    forces = np.array([a.force for a in acts])
    np.savez_compressed(OUT / "force_fn_routing.npz", types=types, forces=forces)
    assert list(forces) == [0.0, 0.0, 1.0, 200.0, 0.0, 2.0, 201.0]


def pair_fields():
    PS = ref_shim.ref("observables.particle_sensing").ParticleSensing
    SS = ref_shim.ref("tasks.searching.species_search").SpeciesSearch
    Colloid = ref_shim.ref("components.colloid").Colloid
    rng = np.random.default_rng(31)
    n, steps = 40, 3
    box = np.array([1000.0, 1000.0, 1000.0])
    types = rng.integers(0, 2, n)
    traj = np.zeros((steps + 1, n, 3))
    traj[0, :, :2] = rng.uniform(100, 900, (n, 2))
    for s in range(steps):
        traj[s + 1] = traj[s]
        traj[s + 1, :, :2] += rng.normal(0, 40, (n, 2))
    # coincident particles: two type-0 particles share a position in the last frame
    zero = np.nonzero(types == 0)[0]
    traj[steps, zero[3]] = traj[steps, zero[1]]
    dirs = np.tile([1.0, 0, 0], (n, 1))

    def decay_exp(d):          # kind 2: c0 exp(-c1 d), c0 = 1, c1 = 4
        return np.exp(-4.0 * d)

    def decay_inv(d):          # kind 3: c0 / (d + c1), c0 = 0.01, c1 = 0.05
        return 0.01 / (d + 0.05)

    out = dict(traj=traj, types=types, box=box)
    cases = [("ps_same", PS, dict(decay_fn=decay_exp, box_length=box, sensing_type=0, scale_factor=100, particle_type=0)),
             ("ps_other", PS, dict(decay_fn=decay_inv, box_length=box, sensing_type=1, scale_factor=100, particle_type=0)),
             ("ss_same", SS, dict(decay_fn=decay_exp, box_length=box, sensing_type=1, avoid=False, scale_factor=100, particle_type=1)),
             ("ss_avoid", SS, dict(decay_fn=decay_inv, box_length=box, sensing_type=0, avoid=True, scale_factor=10, particle_type=1))]
    for name, cls, kw in cases:
        ob = cls(**kw)
        ob.initialize(_colloids(Colloid, traj[0], dirs, types))
        vals = []
        for s in range(1, steps + 1):
            cols = _colloids(Colloid, traj[s], dirs, types)
            r = ob.compute_observable(cols) if cls is PS else ob(cols)
            vals.append(np.asarray(r, dtype=float).reshape(-1))
        out[name] = np.array(vals)
    np.savez_compressed(OUT / "pair_field.npz", **out)

    # the reference's own unit case (CI/unit_tests/tasks/test_species_search.py:46-128): history values
    ss = SS(decay_fn=lambda x: 1 - x, box_length=np.array([1.0, 1.0, 1.0]), sensing_type=1, avoid=False, scale_factor=1,
            particle_type=0)
    p = np.array([[0.0, 0, 0], [1.0, 1, 0], [0.0, 1, 0], [1.0, 0, 0]])
    ss.initialize(_colloids(Colloid, p, np.tile([1.0, 0, 0], (4, 1)), [0, 1, 1, 1]))
    hist = np.array([ss.historical_field[str(i)] for i in [0]])
    np.savez_compressed(OUT / "species_search_unit.npz", pos=p, types=np.array([0, 1, 1, 1]), hist=hist)


if __name__ == "__main__":
    OUT.mkdir(parents=True, exist_ok=True)
    vision_cones()
    signed_angle()
    field_sequences()
    position_director()
    force_fn_routing()
    pair_fields()
    print("golden fixtures written to", OUT)
