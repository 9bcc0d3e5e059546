"""
ctypes front-end of the CPU oracle (oracle/swarm_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  Nothing under swarmrl_b200/
imports this module.
"""
from __future__ import annotations

import ctypes as C
import pathlib
import subprocess

import numpy as np

_HERE = pathlib.Path(__file__).resolve().parent
_LIB_PATH = _HERE / "_build" / "libswarm_oracle.so"

MAX_SPECIES = 64
MAX_TYPES = 16


def build(force: bool = False) -> pathlib.Path:
    """Compile the oracle with the committed Makefile (gcc, OpenMP)."""
    src = _HERE / "swarm_oracle.c"
    if force or not _LIB_PATH.exists() or _LIB_PATH.stat().st_mtime < src.stat().st_mtime:
        subprocess.run(["make", "-C", str(_HERE), "-B"], check=True, capture_output=True)
    return _LIB_PATH


class Params(C.Structure):
    _fields_ = [
        ("n_dims", C.c_int),
        ("periodic", C.c_int * 3),
        ("box", C.c_double * 3),
        ("dt", C.c_double),
        ("kT", C.c_double),
        ("seed", C.c_uint64),
        ("n_species", C.c_int),
        ("n_types", C.c_int),
        ("sp_type", C.c_int * MAX_SPECIES),
        ("sp_gamma_t", (C.c_double * 3) * MAX_SPECIES),
        ("sp_gamma_r", (C.c_double * 3) * MAX_SPECIES),
        ("sp_mass", C.c_double * MAX_SPECIES),
        ("sp_rinertia", (C.c_double * 3) * MAX_SPECIES),
        ("wca_sigma", (C.c_double * MAX_TYPES) * MAX_TYPES),
        ("wca_eps", (C.c_double * MAX_TYPES) * MAX_TYPES),
        ("wca_cut", (C.c_double * MAX_TYPES) * MAX_TYPES),
        ("walls", C.c_int),
        ("wall_sigma", C.c_double * MAX_TYPES),
        ("wall_eps", C.c_double * MAX_TYPES),
        ("wall_cut", C.c_double * MAX_TYPES),
    ]


_DP = C.POINTER(C.c_double)
_IP = C.POINTER(C.c_int32)


class State(C.Structure):
    _fields_ = [("n", C.c_int64)] + [
        (name, _DP)
        for name in (
            "x y z theta q0 q1 q2 q3 vx vy vz wx wy wz f_swim tx ty tz fx fy fz".split()
        )
    ] + [("species", _IP)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(str(_LIB_PATH))
        L.orc_philox4x64_10.argtypes = [C.POINTER(C.c_uint64)] * 3
        L.orc_gaussians.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, _DP]
        L.orc_philox4x32_10.argtypes = [C.POINTER(C.c_uint32)] * 3
        L.orc_noise_words.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64)]
        L.orc_forces.argtypes = [C.POINTER(Params), C.POINTER(State), C.c_int, C.c_int, _DP, _DP, _DP]
        L.orc_bd_steps.argtypes = [C.POINTER(Params), C.POINTER(State), C.c_uint64, C.c_int64, C.c_int]
        L.orc_lv_steps.argtypes = [C.POINTER(Params), C.POINTER(State), C.c_uint64, C.c_int64, C.c_int]
        L.orc_sd_steps.argtypes = [C.POINTER(Params), C.POINTER(State), C.c_int64, C.c_double, C.c_double, C.c_int]
        L.orc_pairs_within.argtypes = [C.POINTER(Params), C.POINTER(State), C.c_double, _IP, C.c_int64]
        L.orc_pairs_within.restype = C.c_int64
        L.orc_cell_index.argtypes = [C.POINTER(Params), C.POINTER(State), C.POINTER(C.c_int), _IP]
        L.orc_field_distances.argtypes = [C.POINTER(State), _IP, C.c_int, _DP, _DP, _DP, _DP, _DP]
        L.orc_field_distances.restype = C.c_int64
        L.orc_vision_cones.argtypes = [
            C.POINTER(State), _DP, _IP, C.c_int, C.c_double, C.c_double, C.c_int, _DP, _IP, C.c_int, C.c_int, _DP,
        ]
        L.orc_vision_cones.restype = C.c_int64
        L.orc_num_threads.restype = C.c_int
        L.orc_set_skin.argtypes = [C.c_double]
        L.orc_last_verlet_builds.restype = C.c_int64
        L.orc_set_num_threads.argtypes = [C.c_int]
        _lib = L
    return _lib


def _dp(a):
    return a.ctypes.data_as(_DP)


def _ip(a):
    return a.ctypes.data_as(_IP)


def philox4x64(ctr, key):
    c = (C.c_uint64 * 4)(*[int(v) for v in ctr])
    k = (C.c_uint64 * 2)(*[int(v) for v in key])
    o = (C.c_uint64 * 4)()
    lib().orc_philox4x64_10(c, k, o)
    return np.array(list(o), dtype=np.uint64)


def philox4x32(ctr, key):
    c = (C.c_uint32 * 4)(*[int(v) for v in ctr])
    k = (C.c_uint32 * 2)(*[int(v) for v in key])
    o = (C.c_uint32 * 4)()
    lib().orc_philox4x32_10(c, k, o)
    return np.array(list(o), dtype=np.uint32)


def noise_words(seed, pid, step, stream):
    """The thermostat's four 64-bit noise words (orc_noise_words)."""
    o = (C.c_uint64 * 4)()
    lib().orc_noise_words(int(seed), int(pid), int(step), int(stream), o)
    return np.array(list(o), dtype=np.uint64)


def gaussians(seed, pid, step, stream):
    g = np.zeros(4)
    lib().orc_gaussians(int(seed), int(pid), int(step), int(stream), _dp(g))
    return g


class OracleSwarm:
    """
    Host-side swarm advanced by the C oracle.  Mirrors the data model of the
    CUDA library (SoA fp64 + species table) so tests can drive both with the
    same inputs.
    """

    def __init__(self, n_dims, box, periodic, dt, kT, seed, species, wca, walls=None):
        """
        species: list of dicts {type, gamma_t(3), gamma_r(3), mass, rinertia(3)}
        wca: dict with 'sigma','eps','cut' (T,T) arrays
        walls: None or dict with 'sigma','eps','cut' (T,) arrays
        """
        p = Params()
        p.n_dims = int(n_dims)
        per = [bool(periodic)] * 3 if np.isscalar(periodic) else list(periodic)
        for a in range(3):
            p.periodic[a] = int(per[a])
            p.box[a] = float(box[a])
        p.dt, p.kT, p.seed = float(dt), float(kT), int(seed)
        p.n_species = len(species)
        T = np.asarray(wca["sigma"]).shape[0]
        p.n_types = T
        for s, sp in enumerate(species):
            p.sp_type[s] = int(sp["type"])
            p.sp_mass[s] = float(sp["mass"])
            for a in range(3):
                p.sp_gamma_t[s][a] = float(sp["gamma_t"][a])
                p.sp_gamma_r[s][a] = float(sp["gamma_r"][a])
                p.sp_rinertia[s][a] = float(sp["rinertia"][a])
        for a in range(T):
            for b in range(T):
                p.wca_sigma[a][b] = float(wca["sigma"][a][b])
                p.wca_eps[a][b] = float(wca["eps"][a][b])
                p.wca_cut[a][b] = float(wca["cut"][a][b])
        p.walls = 0
        if walls is not None:
            p.walls = 1
            for a in range(T):
                p.wall_sigma[a] = float(walls["sigma"][a])
                p.wall_eps[a] = float(walls["eps"][a])
                p.wall_cut[a] = float(walls["cut"][a])
        self.params = p
        self.n_dims = int(n_dims)
        self.species_table = species
        self.n = 0

    def set_state(self, pos, species, theta=None, quat=None):
        pos = np.asarray(pos, dtype=np.float64)
        n = pos.shape[0]
        self.n = n
        self.x, self.y, self.z = (np.ascontiguousarray(pos[:, a]).copy() for a in range(3))
        self.theta = np.zeros(n) if theta is None else np.array(theta, dtype=np.float64)
        q = np.tile([1.0, 0, 0, 0], (n, 1)) if quat is None else np.asarray(quat, dtype=np.float64)
        self.q = [np.ascontiguousarray(q[:, a]).copy() for a in range(4)]
        self.v = [np.zeros(n) for _ in range(3)]
        self.w = [np.zeros(n) for _ in range(3)]
        self.f_swim = np.zeros(n)
        self.torque = [np.zeros(n) for _ in range(3)]
        self.ext_force = [np.zeros(n) for _ in range(3)]
        self.species = np.ascontiguousarray(species, dtype=np.int32).copy()
        self.ptype = np.array([self.species_table[s]["type"] for s in self.species], dtype=np.int32)
        self._mk_state()

    def _mk_state(self):
        s = State()
        s.n = self.n
        s.x, s.y, s.z = _dp(self.x), _dp(self.y), _dp(self.z)
        s.theta = _dp(self.theta)
        s.q0, s.q1, s.q2, s.q3 = (_dp(a) for a in self.q)
        s.vx, s.vy, s.vz = (_dp(a) for a in self.v)
        s.wx, s.wy, s.wz = (_dp(a) for a in self.w)
        s.f_swim = _dp(self.f_swim)
        s.tx, s.ty, s.tz = (_dp(a) for a in self.torque)
        s.fx, s.fy, s.fz = (_dp(a) for a in self.ext_force)
        s.species = _ip(self.species)
        self.state = s

    # --- views -------------------------------------------------------
    @property
    def pos(self):
        return np.stack([self.x, self.y, self.z], axis=1)

    @property
    def directors(self):
        if self.n_dims == 2:
            return np.stack([np.cos(self.theta), np.sin(self.theta), np.zeros(self.n)], axis=1)
        a, b, c, d = self.q
        return np.stack([2 * (b * d + a * c), 2 * (c * d - a * b), a * a - b * b - c * c + d * d], axis=1)

    @property
    def velocities(self):
        return np.stack(self.v, axis=1)

    @property
    def omegas(self):
        return np.stack(self.w, axis=1)

    # --- arithmetic ----------------------------------------------------
    def forces(self, mode=0, with_swim=True):
        F = [np.zeros(self.n) for _ in range(3)]
        lib().orc_forces(C.byref(self.params), C.byref(self.state), mode, int(with_swim), *(_dp(f) for f in F))
        return np.stack(F, axis=1)

    def bd_steps(self, step0, nsteps, mode=0):
        lib().orc_bd_steps(C.byref(self.params), C.byref(self.state), int(step0), int(nsteps), mode)

    def lv_steps(self, step0, nsteps, mode=0):
        """Langevin thermostat + velocity Verlet (espresso.py:1187-1201); v and omega are state here."""
        lib().orc_lv_steps(C.byref(self.params), C.byref(self.state), int(step0), int(nsteps), mode)

    def sd_steps(self, nsteps, gamma=0.1, max_disp=0.1, mode=0):
        lib().orc_sd_steps(C.byref(self.params), C.byref(self.state), int(nsteps), gamma, max_disp, mode)

    def pairs_within(self, skin=0.0):
        cap = max(16, 64 * self.n)
        while True:
            buf = np.zeros((cap, 2), dtype=np.int32)
            cnt = lib().orc_pairs_within(C.byref(self.params), C.byref(self.state), float(skin), _ip(buf), cap)
            if cnt <= cap:
                return buf[:cnt].copy()
            cap = cnt

    def cell_index(self, nc):
        out = np.zeros(self.n, dtype=np.int32)
        ncv = (C.c_int * 3)(*[int(v) for v in nc])
        lib().orc_cell_index(C.byref(self.params), C.byref(self.state), ncv, _ip(out))
        return out

    def field_distances(self, want_type, src, box, hist):
        """hist: (3, n) array, updated in place. Returns (d_now, d_prev)."""
        rows = int(np.sum(self.ptype == want_type))
        d_now, d_prev = np.zeros(rows), np.zeros(rows)
        src = np.ascontiguousarray(src, dtype=np.float64)
        box = np.ascontiguousarray(box, dtype=np.float64)
        assert hist.shape == (3, self.n) and hist.flags.c_contiguous
        r = lib().orc_field_distances(
            C.byref(self.state), _ip(self.ptype), int(want_type), _dp(src), _dp(box), _dp(hist), _dp(d_now), _dp(d_prev)
        )
        assert r == rows
        return d_now, d_prev

    def vision_cones(self, want_type, vision_range, half_angle, n_cones, radii, detected_types):
        rows = int(np.sum(self.ptype == want_type))
        n_det = len(detected_types)
        tmax = int(max(int(self.ptype.max(initial=0)), max(detected_types, default=0))) + 1
        type_col = -np.ones(tmax, dtype=np.int32)
        for c, t in enumerate(detected_types):
            type_col[int(t)] = c
        out = np.zeros((rows, n_cones, n_det))
        dirs = np.ascontiguousarray(self.directors.T)
        radii = np.ascontiguousarray(radii, dtype=np.float64)
        lib().orc_vision_cones(
            C.byref(self.state), _dp(dirs), _ip(self.ptype), int(want_type), float(vision_range), float(half_angle),
            int(n_cones), _dp(radii), _ip(type_col), tmax, n_det, _dp(out),
        )
        return out


def position_observable(pos, ptype, particle_type, box_length):
    """PositionObservable (/root/reference/swarmrl/observables/position.py:32-60): pos / box_length of the colloids
    of `particle_type`, in colloid order."""
    sel = np.asarray(ptype) == int(particle_type)
    return np.asarray(pos, dtype=float)[sel] / np.asarray(box_length, dtype=float)


def director_observable(dirs, ptype, particle_type):
    """Director (/root/reference/swarmrl/observables/director.py:30-58): a copy of the director of each colloid of
    `particle_type`, in colloid order."""
    return np.asarray(dirs, dtype=float)[np.asarray(ptype) == int(particle_type)].copy()


def decay_eval(kind, c0, c1, d):
    """swarm_decay_t: 1 linear c0 + c1 d, 2 exponential c0 exp(-c1 d), 3 inverse c0 / (d + c1)."""
    d = np.asarray(d, dtype=np.float64)
    return {1: lambda: c0 + c1 * d, 2: lambda: c0 * np.exp(-c1 * d), 3: lambda: c0 / (d + c1)}[int(kind)]()


def pair_field(pos, ptype, particle_type, sensing_type, box_length, kind, c0, c1):
    """Field value of every particle of `particle_type` (index order) summed over the particles of
    `sensing_type`: ParticleSensing.compute_single_observable (/root/reference/swarmrl/observables/
    particle_sensing.py:95-131) == SpeciesSearch.compute_single_particle_task (tasks/searching/
    species_search.py:97-133):
        d = |(x_j - x_i) / box_length|;  idx = nonzero(d, size=M-1)  (first M-1 non-zero entries in
        index order, PADDED WITH INDEX 0 when fewer exist);  field = sum decay(d[idx]).
    So an observer outside the sensed set silently loses the LAST sensed particle, and coincident
    particles make sensed[0] count again."""
    pos = np.asarray(pos, dtype=np.float64).reshape(-1, 3)
    ptype = np.asarray(ptype)
    box = np.asarray(box_length, dtype=np.float64)
    sensed = pos[ptype == sensing_type]
    M = len(sensed)
    out = []
    for i in np.nonzero(ptype == particle_type)[0]:
        d = np.sqrt((((sensed - pos[i]) / box) ** 2).sum(axis=1))
        nz = np.nonzero(d)[0][: max(M - 1, 0)]
        idx = np.zeros(max(M - 1, 0), dtype=np.int64)
        idx[: len(nz)] = nz
        total = 0.0
        for v in decay_eval(kind, c0, c1, d[idx]) if M > 1 else []:   # ascending-index summation, as the kernel
            total += float(v)
        out.append(total)
    return np.array(out)


# ------------------------------------------------------------------------------------------------
# float32 evaluation mode: the reference evaluates its observables and tasks with
# JAX in its default float32 -- positions are normalised / packed in float64 numpy and rounded to float32 where they
# enter a jnp call.  The kernels (and the restatements above) compute in fp64; these functions repeat the reference's
# arithmetic with the reference's rounding points, so that tests can STATE how far an fp64 result may lie from what
# real JAX would return (tests/test_oracle_fp32.py on the CPU, tests/test_gpu_obs.py against the kernels).
# ------------------------------------------------------------------------------------------------
_F = np.float32


def decay_eval_f32(kind, c0, c1, d):
    d = np.asarray(d, dtype=_F)
    c0, c1 = _F(c0), _F(c1)
    return {1: lambda: c0 + c1 * d, 2: lambda: c0 * np.exp(-c1 * d), 3: lambda: c0 / (d + c1)}[int(kind)]().astype(_F)


def field_delta_f32(p_now, p_prev, src, kind, c0, c1, scale, clip):
    """ConcentrationField.compute_single_observable / GradientSensing.compute_colloid_reward
    (concentration_field.py:84-110, gradient_sensing.py:92-126) with JAX's float32: `source - position` is float64
    numpy, jnp.linalg.norm rounds it to float32 and everything after that is float32.
    p_now, p_prev: (n, 3) NORMALISED positions (fp64); src: normalised source.  Returns float32 (n,)."""
    dn = np.sqrt(((np.asarray(src, dtype=np.float64) - np.asarray(p_now, dtype=np.float64)).astype(_F) ** 2).sum(axis=1, dtype=_F))
    dp = np.sqrt(((np.asarray(src, dtype=np.float64) - np.asarray(p_prev, dtype=np.float64)).astype(_F) ** 2).sum(axis=1, dtype=_F))
    v = _F(scale) * (decay_eval_f32(kind, c0, c1, dn) - decay_eval_f32(kind, c0, c1, dp))
    return np.clip(v, _F(0), None) if clip else v


def vision_cones_f32(pos, dirs, ptype, want_type, vision_range, half_angle, n_cones, radii, detected_types):
    """SubdividedVisionCones (subdivided_vision_cones.py:83-258, utils.py:297-332) in float32: positions, directors and
    radii are rounded to float32 when they enter the jitted function; every later operation is float32.  The observer
    itself is visited (distance 0: 0/0 = NaN angle, no cone), as in the reference.  O(observers x particles) numpy."""
    pos = np.asarray(pos, dtype=np.float64).astype(_F)
    dirs = np.asarray(dirs, dtype=np.float64).astype(_F)
    rad = np.asarray(radii, dtype=np.float64).astype(_F)
    ptype = np.asarray(ptype)
    det = [int(t) for t in detected_types]
    rims = (_F(-half_angle) + np.arange(n_cones + 1, dtype=_F) * _F(half_angle) * _F(2) / _F(n_cones)).astype(_F)
    obs = np.nonzero(ptype == int(want_type))[0]
    out = np.zeros((len(obs), n_cones, len(det)), dtype=_F)
    with np.errstate(invalid="ignore", divide="ignore"):
        for r, i in enumerate(obs):
            dist = pos - pos[i]
            dn = np.sqrt((dist * dist).sum(axis=1, dtype=_F))
            w = np.minimum(_F(1), _F(2) * rad / dn).astype(_F) * (dn < _F(vision_range))
            e = dirs[i] / np.sqrt((dirs[i] * dirs[i]).sum(dtype=_F))
            o = dist / dn[:, None]
            o = o / np.sqrt((o * o).sum(axis=1, dtype=_F))[:, None]
            ang = np.arccos(np.clip((o * e).sum(axis=1, dtype=_F), _F(-1), _F(1))).astype(_F)
            orth = (o * np.array([-e[1], e[0], e[2]], dtype=_F)).sum(axis=1, dtype=_F)
            ang = ang * np.where(orth >= 0, _F(1), _F(-1))
            for k in range(n_cones):
                inc = (rims[k] < ang) & (rims[k + 1] > ang)       # NaN (the observer itself) is in no cone
                for c, t in enumerate(det):
                    sel = inc & (ptype == t)
                    out[r, k, c] = w[sel].sum(dtype=_F)
    return out


def pair_field_f32(pos, ptype, particle_type, sensing_type, box_length, kind, c0, c1):
    """pair_field() with the reference's float32 rounding points (particle_sensing.py:95-131): positions enter the
    jitted function as float32; division by the box, norm, decay and the sum are float32."""
    pos = np.asarray(pos, dtype=np.float64).reshape(-1, 3).astype(_F)
    box = np.asarray(box_length, dtype=np.float64).astype(_F)
    ptype = np.asarray(ptype)
    sensed = pos[ptype == sensing_type]
    M = len(sensed)
    out = []
    for i in np.nonzero(ptype == particle_type)[0]:
        dd = (sensed - pos[i]) / box
        d = np.sqrt((dd * dd).sum(axis=1, dtype=_F))
        nz = np.nonzero(d)[0][: max(M - 1, 0)]
        idx = np.zeros(max(M - 1, 0), dtype=np.int64)
        idx[: len(nz)] = nz
        out.append(decay_eval_f32(kind, c0, c1, d[idx]).sum(dtype=_F) if M > 1 else _F(0))
    return np.array(out, dtype=_F)


PHILOX_DOMAIN_POLICY = 0x706F6C696379  # counter word c2 of the policy's Gumbel noise (swarm_common.cuh)


def policy_forward(obs, params, seed=0, ids=None, call=0, sample=False):
    """fp32 restatement of the reference's action path for the network of its tests
    (CI/espresso_tests/integration_tests/test_rl_trainers.py:17-26):
        h = relu(obs @ w1 + b1); logits = h @ wa + ba; value = h @ wv + bv      (flax Dense: x @ kernel + bias)
        log_prob = log(softmax(logits) + 1e-8)[idx]                             (networks/flax_network.py:180-192)
        idx = argmax(logits - log(-log(u)))                                     (sampling_strategies/gumbel_distribution.py:37-40)
    u comes from the engine's documented Philox layout -- counter (call, 4 + 16 * (a // 4), "policy", 0), key (id, seed),
    u = ((r >> 11) + 0.5) / 2^53 for action a = word a % 4 -- not from JAX's threefry (unavailable offline)."""
    f = np.float32
    x = np.asarray(obs, dtype=np.float64).astype(f)
    w1, b1, wa, ba, wv, bv = (np.asarray(params[k], dtype=f) for k in ("w1", "b1", "wa", "ba", "wv", "bv"))
    h = np.maximum(x @ w1 + b1, f(0))
    logits = (h @ wa + ba).astype(f)
    value = (h @ wv.reshape(-1, 1) + bv.reshape(1, 1)).reshape(-1).astype(f)
    n, A = logits.shape
    if sample:
        ids = np.arange(n) if ids is None else np.asarray(ids)
        pert = np.empty_like(logits)
        for r in range(n):
            for a0 in range(0, A, 4):
                w = philox4x64((call, 4 + 16 * (a0 // 4), PHILOX_DOMAIN_POLICY, 0), (int(ids[r]), int(seed)))
                for u in range(min(4, A - a0)):
                    uu = (float(int(w[u]) >> 11) + 0.5) / 9007199254740992.0
                    pert[r, a0 + u] = logits[r, a0 + u] - f(np.log(-np.log(uu)))
        idx = np.argmax(pert, axis=1)
    else:
        idx = np.argmax(logits, axis=1)
    e = np.exp(logits - logits.max(axis=1, keepdims=True))
    sm = e / e.sum(axis=1, keepdims=True)
    logp = np.log(sm + f(1e-8))[np.arange(n), idx]
    return idx.astype(np.int32), logp.astype(f), value, logits


def set_skin(skin):
    """Verlet skin of mode 2 (cell_system.skin, espresso.py:269)."""
    lib().orc_set_skin(float(skin))


def last_verlet_builds():
    return int(lib().orc_last_verlet_builds())


def num_threads():
    return lib().orc_num_threads()


def set_num_threads(t):
    lib().orc_set_num_threads(int(t))
