"""
DeviceSwarm -- Python owner of one `swarm_t` handle (include/swarm_b200.h) and of the
structure-of-arrays fp64 state it is bound to.

The state lives in torch CUDA tensors (torch is plumbing here: device memory, streams, DLPack);
every computation goes through the C-ABI of libswarm_b200.so.  There is no CPU fallback: on a box
without the library or without a CUDA device the constructor raises.

This is the data model the reference obtains from `espressomd.System().part`
(/root/reference/swarmrl/engine/espresso.py:193,416-441): positions, orientation, velocity,
swim force, external torque / force, per-particle friction and type.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib


def _torch():
    import torch

    if not torch.cuda.is_available():
        raise _lib.SwarmLibraryError(
            "swarmrl_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback"
        )
    return torch


def director_to_quat(d):
    """ESPResSo's convert_director_to_quaternion (director = body z axis), vectorised, numpy."""
    d = np.atleast_2d(np.asarray(d, dtype=np.float64))
    dm = np.linalg.norm(d, axis=1)
    dxy = np.hypot(d[:, 0], d[:, 1])
    with np.errstate(invalid="ignore", divide="ignore"):
        theta2 = np.where(dxy == 0.0, np.where(d[:, 2] > 0, 0.0, 0.5 * np.pi), 0.5 * np.arccos(d[:, 2] / dm))
        phi2 = np.where(
            dxy == 0.0,
            0.0,
            0.5 * np.where(d[:, 1] > 0.0, 1.0, -1.0) * np.arccos(np.clip(d[:, 0] / np.where(dxy == 0, 1, dxy), -1, 1))
            - 0.25 * np.pi,
        )
    ct, st, cp, sp = np.cos(theta2), np.sin(theta2), np.cos(phi2), np.sin(phi2)
    return np.stack([ct * cp, -st * cp, -st * sp, ct * sp], axis=1)


class DeviceSwarm:
    """One swarm (or `n_envs` independent environments of equal size) resident in HBM."""

    def __init__(self, n_dims, box, periodic, dt, kT, seed, species, wca, walls=None, skin=0.4,
                 n_envs=1, max_neighbors=0, persistent=True, device=None):
        """
        species : list of dicts {type, gamma_t(3), gamma_r(3), mass, rinertia(3)}
        wca     : dict of (T,T) arrays 'sigma', 'eps', 'cut'
        walls   : None or dict of (T,) arrays 'sigma', 'eps', 'cut'
        """
        torch = _torch()
        self.torch = torch
        self.lib = _lib.load()
        self.device_index = torch.cuda.current_device() if device is None else int(device)
        self.dev = torch.device("cuda", self.device_index)
        self.n_dims = int(n_dims)
        self.n_envs = int(n_envs)
        cfg = _lib.Config()
        cfg.n_dims = self.n_dims
        per = [bool(periodic)] * 3 if np.isscalar(periodic) else [bool(p) for p in periodic]
        for a in range(3):
            cfg.periodic[a] = int(per[a])
            cfg.box[a] = float(box[a])
        cfg.dt, cfg.kT, cfg.skin, cfg.seed = float(dt), float(kT), float(skin), int(seed)
        cfg.device = self.device_index
        cfg.max_neighbors = int(max_neighbors)
        cfg.n_envs = self.n_envs
        cfg.flags = 0 if persistent else _lib.FLAG_NO_PERSISTENT
        self.box = np.array([float(b) for b in box])
        self.periodic = per
        self.dt, self.kT, self.seed = float(dt), float(kT), int(seed)
        h = C.c_void_p()
        _lib.check(self.lib.swarm_create(C.byref(cfg), C.byref(h)))
        self.h = h
        self.species_table = list(species)
        self._set_species(species)
        self._set_wca(wca)
        if walls is not None:
            self._set_walls(walls)
        self.n = 0
        self._hold = []
        self._row_maps, self._sensed = {}, {}

    # ------------------------------------------------------------------ tables
    def _set_species(self, species):
        S = len(species)
        typ = np.array([int(s["type"]) for s in species], dtype=np.int32)
        gt = np.array([np.broadcast_to(np.asarray(s["gamma_t"], dtype=float), (3,)) for s in species]).reshape(S, 3).copy()
        gr = np.array([np.broadcast_to(np.asarray(s["gamma_r"], dtype=float), (3,)) for s in species]).reshape(S, 3).copy()
        mass = np.array([float(s["mass"]) for s in species], dtype=np.float64)
        rin = np.array([np.broadcast_to(np.asarray(s["rinertia"], dtype=float), (3,)) for s in species]).reshape(S, 3).copy()
        _lib.check(self.lib.swarm_set_species(self.h, S, _lib.ip(typ), _lib.dp(gt), _lib.dp(gr), _lib.dp(mass), _lib.dp(rin)))
        self.type_of_species = typ

    def _set_wca(self, wca):
        sig = np.ascontiguousarray(wca["sigma"], dtype=np.float64)
        eps = np.ascontiguousarray(wca["eps"], dtype=np.float64)
        cut = np.ascontiguousarray(wca["cut"], dtype=np.float64)
        T = sig.shape[0]
        _lib.check(self.lib.swarm_set_wca(self.h, T, _lib.dp(sig), _lib.dp(eps), _lib.dp(cut)))
        self.n_types = T

    def _set_walls(self, walls):
        sig = np.zeros(_lib.MAX_TYPES); eps = np.zeros(_lib.MAX_TYPES); cut = np.zeros(_lib.MAX_TYPES)
        k = len(walls["sigma"])
        sig[:k], eps[:k], cut[:k] = walls["sigma"], walls["eps"], walls["cut"]
        _lib.check(self.lib.swarm_set_walls(self.h, 1, _lib.dp(sig), _lib.dp(eps), _lib.dp(cut)))

    def set_thermostat(self, kT, seed):
        _lib.check(self.lib.swarm_set_thermostat(self.h, float(kT), int(seed)))
        self.kT, self.seed = float(kT), int(seed)

    def set_integrator(self, thermostat_type):
        """MDParams.thermostat_type: "brownian" (default) or "langevin" (velocity Verlet; v, omega are state)."""
        kinds = {"brownian": _lib.INTEGRATOR_BROWNIAN, "langevin": _lib.INTEGRATOR_LANGEVIN}
        if thermostat_type not in kinds:
            raise ValueError(f"integrator_type must be one of {list(kinds)}")
        _lib.check(self.lib.swarm_set_integrator(self.h, kinds[thermostat_type]))
        self.thermostat_type = thermostat_type

    # ------------------------------------------------------------------- state
    def _stream(self):
        return C.c_void_p(self.torch.cuda.current_stream(self.dev).cuda_stream)

    def set_state(self, pos, species, theta=None, quat=None):
        """Upload positions (n,3), species indices (n,), 2-D angles or 3-D quaternions."""
        torch = self.torch
        pos = np.asarray(pos, dtype=np.float64).reshape(-1, 3)
        n = pos.shape[0]
        self.n = n
        self._row_maps, self._sensed = {}, {}
        f64 = dict(dtype=torch.float64, device=self.dev)
        m = max(n, 1)
        # one allocation per field group; rows are the SoA components
        self.pos = torch.zeros((3, m), **f64)
        self.pos[:, :n] = torch.from_numpy(np.ascontiguousarray(pos.T)).to(self.dev)
        self.theta = torch.zeros(m, **f64)
        self.quat = torch.zeros((4, m), **f64)
        self.quat[0] = 1.0
        if theta is not None:
            self.theta[:n] = torch.from_numpy(np.asarray(theta, dtype=np.float64)).to(self.dev)
        if quat is not None:
            self.quat[:, :n] = torch.from_numpy(np.ascontiguousarray(np.asarray(quat, dtype=np.float64).T)).to(self.dev)
        self.vel = torch.zeros((3, m), **f64)
        self.omega = torch.zeros((3, m), **f64)
        self.f_swim = torch.zeros(m, **f64)
        self.torque = torch.zeros((3, m), **f64)
        self.ext_force = torch.zeros((3, m), **f64)
        self.new_dir = torch.zeros((3, m), **f64)
        self.has_new_dir = torch.zeros(m, dtype=torch.int32, device=self.dev)
        sp = np.ascontiguousarray(species, dtype=np.int32)
        self.species = torch.zeros(m, dtype=torch.int32, device=self.dev)
        self.species[:n] = torch.from_numpy(sp).to(self.dev)
        self.species_host = sp.copy()
        self.ptype_host = self.type_of_species[sp] if n else np.zeros(0, dtype=np.int32)
        st = _lib.StateDev()
        st.x, st.y, st.z = (self.pos[a].data_ptr() for a in range(3))
        st.theta = self.theta.data_ptr()
        st.q0, st.q1, st.q2, st.q3 = (self.quat[a].data_ptr() for a in range(4))
        st.vx, st.vy, st.vz = (self.vel[a].data_ptr() for a in range(3))
        st.wx, st.wy, st.wz = (self.omega[a].data_ptr() for a in range(3))
        st.species = self.species.data_ptr()
        _lib.check(self.lib.swarm_bind_state(self.h, n, C.byref(st)))
        act = _lib.ActionsDev()
        act.f_swim = self.f_swim.data_ptr()
        act.tqx, act.tqy, act.tqz = (self.torque[a].data_ptr() for a in range(3))
        act.fex, act.fey, act.fez = (self.ext_force[a].data_ptr() for a in range(3))
        act.ndx, act.ndy, act.ndz = (self.new_dir[a].data_ptr() for a in range(3))
        act.has_nd = self.has_new_dir.data_ptr()
        _lib.check(self.lib.swarm_bind_actions(self.h, C.byref(act)))

    def bind_ids(self, gid):
        """Philox ids of the bound slots (int64 device tensor): the global agent index when this handle holds
        a shard of a larger ensemble / swarm, so results do not depend on the sharding."""
        self._gid = gid.contiguous()
        _lib.check(self.lib.swarm_bind_ids(self.h, self._gid.data_ptr()))

    # ------------------------------------------------------------ integration
    def run(self, n_steps, step0):
        """integrator.run(n_steps) -- enqueue only (espresso.py:1304-1306)."""
        _lib.check(self.lib.swarm_run(self.h, int(n_steps), int(step0), self._stream()))

    def relax(self, n_iter=1000, gamma=0.1, max_disp=0.1):
        """_remove_overlap's steepest descent (espresso.py:1163-1167)."""
        _lib.check(self.lib.swarm_relax(self.h, int(n_iter), float(gamma), float(max_disp), self._stream()))

    def stats(self):
        st = _lib.Stats()
        _lib.check(self.lib.swarm_sync_stats(self.h, self._stream(), C.byref(st)))
        return st

    def run_checked(self, n_steps, step0, grow=True):
        """run + sync; grows the Verlet capacity and resumes if a list overflowed."""
        done = 0
        while True:
            self.run(n_steps - done, step0 + done)
            st = self.stats()
            done += int(st.steps_done)
            if not st.overflow:
                return st
            if not grow:
                raise OverflowError("Verlet list capacity exceeded")
            self._grow_neighbors(st)

    def relax_checked(self, n_iter=1000, gamma=0.1, max_disp=0.1):
        done = 0
        while True:
            self.relax(n_iter - done, gamma, max_disp)
            st = self.stats()
            done += int(st.steps_done)
            if not st.overflow:
                return st
            self._grow_neighbors(st)

    def _grow_neighbors(self, st):
        """Double the Verlet capacity after an overflow; at the hard ceiling there is nothing left to grow, and
        retrying with the same capacity would overflow again at once (the run would never end)."""
        if st.capacity >= _lib.MAX_NEIGHBORS:
            raise OverflowError(
                f"Verlet list capacity exceeded: a particle has more than {_lib.MAX_NEIGHBORS} neighbours within "
                f"cut-off + skin (max_neighbors_seen >= {st.max_neighbors_seen}); lower the density or the skin")
        self.set_max_neighbors(min(_lib.MAX_NEIGHBORS, max(2 * st.capacity, st.max_neighbors_seen + 8)))

    def set_max_neighbors(self, cap):
        _lib.check(self.lib.swarm_set_max_neighbors(self.h, int(cap)))

    def set_tiling(self, mode, resort_interval=0):
        """Spatial tiling of the 2-D Brownian chunk (swarm_set_tiling): 0 off, 1 automatic, 2 whenever it fits."""
        _lib.check(self.lib.swarm_set_tiling(self.h, int(mode), int(resort_interval)))

    def tiling_stats(self):
        """dict of the tiled kernel's counters (swarm_tiling_stats); synchronises."""
        out = (C.c_int64 * 8)()
        _lib.check(self.lib.swarm_tiling_stats(self.h, out))
        keys = ("tiled", "cadence", "steps_on_list", "rollbacks", "rim_sum", "sorts", "ctas", "slots_per_cta")
        return dict(zip(keys, [int(v) for v in out]))

    def tiled_neighbors(self):
        """Sorted (i, j) id pairs of the tiled kernel's own Verlet list, as it stands."""
        cnt = C.c_int64(0)
        cap = max(64, 32 * self.n)
        while True:
            buf = np.zeros((cap, 2), dtype=np.int32)
            _lib.check(self.lib.swarm_tiled_neighbors_dump(self.h, _lib.ip(buf), cap, C.byref(cnt)))
            if cnt.value <= cap:
                return buf[: cnt.value].copy()
            cap = int(cnt.value)

    def apply_new_direction(self):
        _lib.check(self.lib.swarm_apply_new_direction(self.h, self._stream()))

    # ---------------------------------------------------------------- read-out
    def directors(self):
        """(3, n) device tensor of directors."""
        out = self.torch.empty((3, max(self.n, 1)), dtype=self.torch.float64, device=self.dev)
        _lib.check(self.lib.swarm_directors(self.h, out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(), self._stream()))
        return out[:, : self.n]

    def forces(self):
        out = self.torch.empty((3, max(self.n, 1)), dtype=self.torch.float64, device=self.dev)
        _lib.check(self.lib.swarm_forces(self.h, out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(), self._stream()))
        return out[:, : self.n]

    def particle_data_host(self):
        """get_particle_data through host buffers (espresso.py:1320-1336): three (n,3) arrays."""
        n = self.n
        pos, vel, dr = (np.zeros((n, 3)) for _ in range(3))
        if n:
            _lib.check(self.lib.swarm_get_particle_data_host(
                self.h, pos.ctypes.data, vel.ctypes.data, dr.ctypes.data, self._stream()))
        return pos, vel, dr

    def set_actions_host(self, f_swim, torque):
        """manage_forces' writes through host buffers (espresso.py:1227-1235)."""
        f = np.ascontiguousarray(f_swim, dtype=np.float64)
        t = np.ascontiguousarray(torque, dtype=np.float64).reshape(-1, 3)
        if self.n_dims == 2 and not t[:, :2].any():
            # in the plane only tau_z matters: half the bytes, no de-interleave kernel (swarm_set_actions_host_2d)
            tz = np.ascontiguousarray(t[:, 2])
            self._hold = [f, tz]
            _lib.check(self.lib.swarm_set_actions_host_2d(self.h, f.ctypes.data, tz.ctypes.data, self._stream()))
            self.torch.cuda.current_stream(self.dev).synchronize()
            return
        self._hold = [f, t]  # keep alive until the async copies have run
        _lib.check(self.lib.swarm_set_actions_host(self.h, f.ctypes.data, t.ctypes.data, self._stream()))
        self.torch.cuda.current_stream(self.dev).synchronize()

    def set_ext_force_host(self, F):
        """parts.ext_force = F (espresso.py:834-851); (n,3) host array."""
        F = np.ascontiguousarray(np.asarray(F, dtype=np.float64).reshape(-1, 3).T)
        self.ext_force[:, : self.n] = self.torch.from_numpy(F).to(self.dev)

    def set_new_directions_host(self, nd, has):
        """Action.new_direction per particle (espresso.py:1236-1249); applied by apply_new_direction."""
        nd = np.ascontiguousarray(np.asarray(nd, dtype=np.float64).reshape(-1, 3).T)
        self.new_dir[:, : self.n] = self.torch.from_numpy(nd).to(self.dev)
        self.has_new_dir[: self.n] = self.torch.from_numpy(np.ascontiguousarray(has, dtype=np.int32)).to(self.dev)

    def neighbors(self):
        """Sorted (i, j) pairs of the Verlet list rebuilt at the current positions."""
        cnt = C.c_int64(0)
        cap = max(64, 32 * self.n)
        while True:
            buf = np.zeros((cap, 2), dtype=np.int32)
            _lib.check(self.lib.swarm_neighbors_dump(self.h, _lib.ip(buf), cap, C.byref(cnt)))
            if cnt.value <= cap:
                return buf[: cnt.value].copy()
            cap = int(cnt.value)

    def cell_index(self, nc):
        out = np.zeros(max(self.n, 1), dtype=np.int32)
        ncv = (C.c_int32 * 3)(*[int(v) for v in nc])
        _lib.check(self.lib.swarm_cell_index(self.h, ncv, _lib.ip(out)))
        return out[: self.n]

    def gaussians(self, n, step, stream_id):
        out = self.torch.empty((max(n, 1), 4), dtype=self.torch.float64, device=self.dev)
        _lib.check(self.lib.swarm_gaussians(self.h, int(n), int(step), int(stream_id), out.data_ptr(), self._stream()))
        return out[:n]

    # ------------------------------------------------------------- observables
    def row_map(self, particle_type):
        """int32 device tensor: output row of every particle of `particle_type`, else -1.  Cached per type until the
        next set_state (the types of the bound particles do not change): no host work or copy on the slice path."""
        key = int(particle_type)
        hit = self._row_maps.get(key)
        if hit is None:
            rows = -np.ones(max(self.n, 1), dtype=np.int32)
            sel = np.nonzero(self.ptype_host == key)[0]
            rows[sel] = np.arange(len(sel), dtype=np.int32)
            hit = (self.torch.from_numpy(rows).to(self.dev), len(sel))
            self._row_maps[key] = hit
        return hit

    def obs_field(self, row_of, n_rows, src, box_length, hist, decay=_lib.DECAY_NONE, c0=0.0, c1=0.0,
                  scale=1.0, clip=False, want_dist=True):
        torch = self.torch
        f64 = dict(dtype=torch.float64, device=self.dev)
        # every row belongs to one particle and the kernel writes it: no fill needed
        out = torch.empty(max(n_rows, 1), **f64) if decay != _lib.DECAY_NONE else None
        dn = torch.empty(max(n_rows, 1), **f64) if want_dist else None
        dp = torch.empty(max(n_rows, 1), **f64) if want_dist else None
        s3 = (C.c_double * 3)(*[float(v) for v in src])
        b3 = (C.c_double * 3)(*[float(v) for v in box_length])
        _lib.check(self.lib.swarm_obs_field(
            self.h, row_of.data_ptr() if row_of is not None else None, s3, b3, hist.data_ptr(), int(decay),
            float(c0), float(c1), float(scale), int(bool(clip)),
            out.data_ptr() if out is not None else None,
            dn.data_ptr() if dn is not None else None, dp.data_ptr() if dp is not None else None, self._stream()))
        cut = (lambda t: None if t is None else t[:n_rows])
        return cut(out), cut(dn), cut(dp)

    def obs_field_mean(self, row_of, n_rows, src, box_length, hist, decay, c0, c1, scale=1.0, clip=False):
        """Rewards of a closed-form decay AND their (sum, mean) in one launch: (rewards[n_rows], red[2])."""
        torch = self.torch
        out = torch.empty(max(n_rows, 1), dtype=torch.float64, device=self.dev)
        red = torch.empty(2, dtype=torch.float64, device=self.dev)
        s3 = (C.c_double * 3)(*[float(v) for v in src])
        b3 = (C.c_double * 3)(*[float(v) for v in box_length])
        _lib.check(self.lib.swarm_obs_field_mean(
            self.h, row_of.data_ptr() if row_of is not None else None, int(n_rows), s3, b3, hist.data_ptr(), int(decay),
            float(c0), float(c1), float(scale), int(bool(clip)), out.data_ptr(), red.data_ptr(), self._stream()))
        return out[:n_rows], red

    def obs_vision(self, row_of, n_rows, vision_range, half_angle, n_cones, radii, type_col, n_cols, dirs=None):
        torch = self.torch
        out = torch.empty((max(n_rows, 1), n_cones, n_cols), dtype=torch.float64, device=self.dev)  # zeroed by the call
        _lib.check(self.lib.swarm_obs_vision(
            self.h, row_of.data_ptr(), int(n_rows), float(vision_range), float(half_angle), int(n_cones),
            radii.data_ptr(), type_col.data_ptr(), int(type_col.numel()), int(n_cols),
            dirs.data_ptr() if dirs is not None else None, out.data_ptr(), self._stream()))
        return out[:n_rows]

    def sensed_index(self, sensing_type):
        """int32 device tensor of the particles of `sensing_type`, ascending (the reference's `test_points`); cached."""
        key = int(sensing_type)
        hit = self._sensed.get(key)
        if hit is None:
            sel = np.nonzero(self.ptype_host == key)[0].astype(np.int32)
            hit = (self.torch.from_numpy(sel).to(self.dev), len(sel))
            self._sensed[key] = hit
        return hit

    def obs_pair_field(self, row_of, n_rows, sensed, n_sensed, box_length, hist_rows, decay, c0, c1, scale=1.0, clip_mode=0,
                       init_only=False):
        """ParticleSensing / SpeciesSearch field (swarm_obs_pair_field); hist_rows (n_rows,) is updated in place."""
        out = self.torch.zeros(max(n_rows, 1), dtype=self.torch.float64, device=self.dev)
        b3 = (C.c_double * 3)(*[float(v) for v in box_length])
        _lib.check(self.lib.swarm_obs_pair_field(
            self.h, row_of.data_ptr(), sensed.data_ptr() if n_sensed else None, int(n_sensed), b3, int(decay), float(c0),
            float(c1), float(scale), int(clip_mode), int(bool(init_only)), hist_rows.data_ptr(), out.data_ptr(), None,
            self._stream()))
        return out[:n_rows]

    def policy_forward(self, net, obs, row_of, n_rows, call, sample=True, act_force=None, act_torque=None, want_logits=False):
        """swarm_policy_forward: `net` is a dict of fp32 device tensors w1 [D,H], b1 [H], wa [H,A], ba [A], wv [H], bv [1].
        Returns (idx int32, log_prob fp32, value fp32[, logits]) device tensors of n_rows; when the action tables are
        given the chosen action is also written into f_swim / torque."""
        torch = self.torch
        D, H = net["w1"].shape
        A = net["wa"].shape[1]
        p = _lib.Policy(int(D), int(H), int(A), *[net[k].data_ptr() for k in ("w1", "b1", "wa", "ba", "wv", "bv")])
        m = max(n_rows, 1)
        idx = torch.empty(m, dtype=torch.int32, device=self.dev)     # every row belongs to one particle: the kernel writes it
        logp = torch.empty(m, dtype=torch.float32, device=self.dev)
        val = torch.empty(m, dtype=torch.float32, device=self.dev)
        lg = torch.zeros((m, A), dtype=torch.float32, device=self.dev) if want_logits else None
        obs = obs.contiguous()
        _lib.check(self.lib.swarm_policy_forward(
            self.h, C.byref(p), obs.data_ptr(), row_of.data_ptr() if row_of is not None else None, int(call), int(bool(sample)),
            act_force.data_ptr() if act_force is not None else None, act_torque.data_ptr() if act_torque is not None else None,
            idx.data_ptr(), logp.data_ptr(), val.data_ptr(), lg.data_ptr() if lg is not None else None, self._stream()))
        out = (idx[:n_rows], logp[:n_rows], val[:n_rows])
        return out + (lg[:n_rows],) if want_logits else out

    def obs_position(self, row_of, n_rows, box_length):
        torch = self.torch
        out = torch.zeros((max(n_rows, 1), 3), dtype=torch.float64, device=self.dev)
        b3 = (C.c_double * 3)(*[float(v) for v in box_length])
        _lib.check(self.lib.swarm_obs_position(self.h, row_of.data_ptr() if row_of is not None else None, b3,
                                               out.data_ptr(), self._stream()))
        return out[:n_rows]

    def reduce_mean(self, values):
        """(sum, mean) of a 1-D fp64 device tensor, deterministic order."""
        out = self.torch.empty(2, dtype=self.torch.float64, device=self.dev)
        v = values.contiguous()
        _lib.check(self.lib.swarm_reduce_mean(self.h, v.data_ptr(), int(v.numel()), out.data_ptr(), self._stream()))
        return out

    # ------------------------------------------------------------------ misc
    def synchronize(self):
        self.torch.cuda.current_stream(self.dev).synchronize()

    def close(self):
        if getattr(self, "h", None) is not None and self.h:
            self.lib.swarm_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
