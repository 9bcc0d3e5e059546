"""
One large periodic 2-D swarm, slab-decomposed along x over the GPUs of a node (BASELINE config 5,
SURVEY.md 8(e).2).  The reference has no counterpart -- SwarmRL drives a single-rank ESPResSo
(swarmrl/engine/espresso.py:193,1304-1306) -- so the contract is: same trajectories as the single-GPU
engine (and the oracle) on the same inputs, to the tolerance north_star states (1e-9 relative per
step, noise-free; identical noise because Philox is keyed by the global particle id).

Scheme: every rank keeps its slab plus two halos of width H and integrates the
ghosts redundantly with the same kernel; halos are re-shipped (and ownership migrates) only every
`exchange_every` = k steps, H = k * r_cut * (1 + margin).  Per exchange: k_dd_pack -> two NCCL
send/recv pairs over NVLink (torch.distributed) -> k_dd_merge -> re-bind.  No per-step collective.

`slab_plan` and `select_local` are pure host logic (tested on CPU, gloo world_size 2);
`DomainSwarm` needs the CUDA library and a device -- there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib


def slab_plan(box_x, world, rank, exchange_every, max_cut, margin=0.25, halo_extra=0.0):
    """Geometry of rank `rank`'s slab.  Raises ValueError when the slabs are narrower than two halos.

    halo_extra widens the halo beyond what k steps of integration need: a per-agent observable with a finite range
    (vision cones: `vision_range`) is exact for every particle within k * r_cut * (1 + margin) of the slab only if the
    rank also holds everything within `halo_extra` of THOSE particles ("per slice a wider halo")."""
    world, rank, k = int(world), int(rank), int(exchange_every)
    if world < 1 or not (0 <= rank < world):
        raise ValueError("rank must lie in [0, world)")
    if k < 1:
        raise ValueError("exchange_every must be >= 1")
    if not (max_cut > 0.0 and box_x > 0.0 and margin >= 0.0 and halo_extra >= 0.0):
        raise ValueError("box_x, max_cut must be positive and margin, halo_extra non-negative")
    w = float(box_x) / world
    h = k * float(max_cut) * (1.0 + float(margin)) + float(halo_extra)
    if w < 2.0 * h:
        raise ValueError(f"slab width {w:g} < 2 halos of {h:g}: lower exchange_every or the number of ranks")
    return {
        "world": world, "rank": rank, "k": k, "box_x": float(box_x), "slab_w": w, "halo": h, "halo_extra": float(halo_extra),
        "x0": rank * w,
        "left": (rank - 1) % world, "right": (rank + 1) % world,
        "image_shift_left": -1 if rank == 0 else 0,
        "image_shift_right": 1 if rank == world - 1 else 0,
        # local frame: x_global = x_local + x_offset + image * box_x
        "x_offset": rank * w - h,
    }


def select_local(x_global, plan):
    """Initial scatter: which particles (and which periodic image of each) rank `plan['rank']` keeps.

    Returns (index, x_local, image, owned) -- index into the global arrays, x in the local frame where
    the slab is [H, H + W), the periodic image count such that
    x_global = x_local + x_offset + image * box_x, and the ownership flag.
    """
    x = np.asarray(x_global, dtype=np.float64)
    L, w, h, x0 = plan["box_x"], plan["slab_w"], plan["halo"], plan["x0"]
    img = np.floor(x / L)
    xf = x - img * L
    idx, xl, im = [], [], []
    for m in (-1, 0, 1):
        cand = xf - x0 + h + m * L
        sel = np.nonzero((cand >= 0.0) & (cand < w + 2.0 * h))[0]
        idx.append(sel)
        xl.append(cand[sel])
        im.append(img[sel] - m)
    idx = np.concatenate(idx)
    xl = np.concatenate(xl)
    im = np.concatenate(im).astype(np.int32)
    order = np.argsort(idx, kind="stable")
    idx, xl, im = idx[order], xl[order], im[order]
    owned = ((xl >= h) & (xl < h + w)).astype(np.int32)
    return idx.astype(np.int64), xl, im, owned


def unfold_global(x_local, image, plan):
    """x_global of local-frame coordinates (inverse of select_local)."""
    return np.asarray(x_local) + plan["x_offset"] + np.asarray(image, dtype=np.float64) * plan["box_x"]


class _Transport:
    """Moves the two exchange buffers to the slab neighbours: NCCL device-to-device over NVLink when the
    process group is nccl, host-staged for gloo (single-GPU tests), a local swap when world == 1."""

    def __init__(self, plan, group=None):
        self.plan = plan
        self.group = group
        self.world = plan["world"]
        if self.world > 1:
            import torch.distributed as dist

            self.dist = dist
            self.backend = dist.get_backend(group)

    def exchange(self, send_l, send_r, recv_l, recv_r):
        if self.world == 1:
            recv_r.copy_(send_l)   # what goes out through my left face enters through my right face
            recv_l.copy_(send_r)
            return
        dist, p = self.dist, self.plan
        if self.backend == "nccl":
            # posting order matters when left == right (world 2): the peer's first send is its send_l,
            # which reaches me through my RIGHT face
            ops = [dist.P2POp(dist.isend, send_l, p["left"], self.group), dist.P2POp(dist.isend, send_r, p["right"], self.group),
                   dist.P2POp(dist.irecv, recv_r, p["right"], self.group), dist.P2POp(dist.irecv, recv_l, p["left"], self.group)]
            for r in dist.batch_isend_irecv(ops):
                r.wait()
            return
        sl, sr = send_l.cpu(), send_r.cpu()
        rl, rr = sl.new_empty(sl.shape), sr.new_empty(sr.shape)
        reqs = [dist.isend(sl, p["left"], group=self.group, tag=1), dist.isend(sr, p["right"], group=self.group, tag=2),
                dist.irecv(rr, p["right"], group=self.group, tag=1), dist.irecv(rl, p["left"], group=self.group, tag=2)]
        for r in reqs:
            r.wait()
        recv_l.copy_(rl)
        recv_r.copy_(rr)

    def all_reduce_sum(self, t):
        if self.world > 1:
            if self.backend == "nccl":
                self.dist.all_reduce(t, group=self.group)
            else:
                c = t.cpu()
                self.dist.all_reduce(c, group=self.group)
                t.copy_(c)
        return t


_F64 = ("x", "y", "theta", "f_swim", "tqz")


class DomainSwarm:
    """This rank's slab (+ halos) of a periodic 2-D swarm; see the module docstring."""

    def __init__(self, box, dt, kT, seed, species, wca, *, exchange_every=8, margin=0.25, skin=0.4, world=1, rank=0,
                 group=None, device=None, n_aux=2, max_neighbors=0, capacity_factor=1.3, persistent=True, halo_extra=0.0,
                 peer="auto", velocities=False):
        """velocities: ship vx, vy, omega_z with every halo record (three more aux channels) and cut the runs with
        swarm_run_part, so that the Langevin integrator (velocity Verlet: velocities are state) works on the decomposed swarm.

        peer: move the halo records with P2P stores into the neighbours' memory (swarm_dd_push / swarm_dd_wait, CUDA IPC)
        instead of NCCL send / recv.  "auto" = when the process group is nccl (one GPU per rank on one node); True forces
        it (ranks sharing one GPU over gloo: the tests); False keeps torch.distributed.  SWARM_B200_DD_PEER=0/1 overrides."""
        import os

        from .device import DeviceSwarm

        env = os.environ.get("SWARM_B200_DD_PEER")
        self._peer_req = peer if env is None else (env not in ("0", "", "false"))
        self.peer = None   # set by _alloc: dict(region, left, right) when the peer path is active

        self.max_cut = float(np.max(np.asarray(wca["cut"], dtype=np.float64)))
        self.plan = slab_plan(box[0], world, rank, exchange_every, self.max_cut, margin, halo_extra)
        self.velocities = bool(velocities)
        self._vel_base = int(n_aux)          # aux channels [vel_base, vel_base + 3) = vx, vy, omega_z
        n_aux = int(n_aux) + (3 if self.velocities else 0)
        if not (0 <= n_aux <= _lib.DD_MAX_AUX):
            raise ValueError(f"n_aux (+ 3 velocity channels) must lie in [0, {_lib.DD_MAX_AUX}]")
        p = self.plan
        self.box = (float(box[0]), float(box[1]))
        self.n_aux = int(n_aux)
        self.capacity_factor = float(capacity_factor)
        local_box = (p["slab_w"] + 2.0 * p["halo"], self.box[1], self.box[1])
        self.sw = DeviceSwarm(2, local_box, (False, True, False), dt, kT, seed, species, wca, skin=skin,
                              max_neighbors=max_neighbors, persistent=persistent, device=device)
        self.torch, self.dev, self.lib, self.h = self.sw.torch, self.sw.dev, self.sw.lib, self.sw.h
        self.transport = _Transport(p, group)
        self.n = 0
        self.n_owned = 0
        self.since_exchange = 0
        self.exchanges = 0
        self._deferred_sums = []   # per-slice reward sums of this rank awaiting reward_means()
        self.cur = 0
        # profile = True brackets every phase with CUDA events (asynchronous, no host sync): `timeline()` then tells
        # how a slice splits into integration, exchange (pack + transport + merge, i.e. including the wait for the
        # slab neighbours) and reward evaluation
        self.profile = False
        self._events = []

    # ---------------------------------------------------------------- storage
    def _alloc(self, cap, buf_cap):
        torch = self.torch
        f64 = dict(dtype=torch.float64, device=self.dev)
        self.cap = int(cap)
        self.sets = []
        for _ in range(2):   # ping-pong halves: merge reads one, writes the other
            s = {k: torch.zeros(self.cap, **f64) for k in _F64 if k != "tqz"}
            s["tq"] = torch.zeros((3, self.cap), **f64)   # rows x, y stay local (2-D: always 0); row z travels
            s["tqz"] = s["tq"][2]
            s["gid"] = torch.zeros(self.cap, dtype=torch.int64, device=self.dev)
            for k in ("species", "image", "owned"):
                s[k] = torch.zeros(self.cap, dtype=torch.int32, device=self.dev)
            s["aux"] = [torch.zeros(self.cap, **f64) for _ in range(self.n_aux)]
            self.sets.append(s)
        self.zeros = torch.zeros(self.cap, **f64)
        self.vel = torch.zeros((3, self.cap), **f64)
        self.omega = torch.zeros((3, self.cap), **f64)
        self.geom = _lib.DDGeom(self.plan["slab_w"], self.plan["halo"], self.n_aux, int(buf_cap),
                                self.plan["image_shift_left"], self.plan["image_shift_right"])
        nd = int(self.lib.swarm_dd_buffer_doubles(C.byref(self.geom)))
        self.bufs = {k: torch.zeros(nd, **f64) for k in ("send_l", "send_r", "recv_l", "recv_r")}
        # per half: [0] live particle count, [1] owned, [2] error bits of the last merge, [3] sticky error bits.
        # The count stays on the device: kernels read it there, so exchanges queue without a host sync.
        self.counts = [torch.zeros(8, dtype=torch.int32, device=self.dev) for _ in range(2)]
        self.counts_host = torch.zeros((2, 8), dtype=torch.int32).pin_memory()
        self.out = torch.zeros(self.cap, **f64)
        _lib.check(self.lib.swarm_reserve(self.h, self.cap))
        _lib.check(self.lib.swarm_dd_set_lean(self.h, 1))   # the two halves alternate: counters are cleared by the merge itself
        self._peer_setup()

    def _peer_setup(self):
        """Exchange region of this rank + the two neighbours' regions (CUDA IPC); see csrc/swarm_dd.cu."""
        t, p = self.transport, self.plan
        want = self._peer_req
        if want == "auto":
            want = t.world > 1 and t.backend == "nccl"
        if not want or t.world < 2:
            return
        # Every step is agreed on by ALL ranks (all_gather of the outcome): if CUDA IPC is not available to one of them (a
        # container without shared IPC, a device without peer access) every rank falls back to torch.distributed send / recv.
        region = C.c_void_p()
        handle = C.create_string_buffer(64)
        ok = self.lib.swarm_dd_peer_alloc(self.sw.device_index, C.byref(self.geom), C.byref(region), handle) == 0
        mine = (bytes(handle.raw) if ok else None, int(self.geom.cap), int(self.geom.n_aux))
        every = [None] * t.world
        t.dist.all_gather_object(every, mine, group=t.group)    # also the barrier between "zeroed" and "first push"
        if any((c, a) != mine[1:] for _, c, a in every):
            raise RuntimeError("the ranks of a domain-decomposed swarm must use the same exchange-buffer geometry")
        opened = {}
        if all(h is not None for h, _, _ in every):
            for nb in {p["left"], p["right"]}:
                r = C.c_void_p()
                if self.lib.swarm_dd_peer_open(self.sw.device_index, every[nb][0], C.byref(r)) != 0:
                    ok = False
                    break
                opened[nb] = r
        else:
            ok = False
        verdict = [None] * t.world
        t.dist.all_gather_object(verdict, bool(ok), group=t.group)
        if not all(verdict):
            for r in opened.values():
                self.lib.swarm_dd_peer_close(r)
            t.dist.barrier(group=t.group)
            if region:
                self.lib.swarm_dd_peer_free(region)
            if self._peer_req is True:
                raise RuntimeError("the peer-memory exchange was requested but CUDA IPC is not available: " + _lib.last_error())
            return
        self.peer = dict(region=region, left=opened[p["left"]], right=opened[p["right"]], opened=opened)

    def _dd(self, s):
        a = _lib.DDArrays()
        for k in _F64 + ("gid", "species", "image", "owned"):
            setattr(a, k, s[k].data_ptr())
        for c in range(self.n_aux):
            a.aux[c] = s["aux"][c].data_ptr()
        return a

    def _bind(self):
        s = self.sets[self.cur]
        st = _lib.StateDev()
        st.x, st.y, st.z, st.theta = s["x"].data_ptr(), s["y"].data_ptr(), self.zeros.data_ptr(), s["theta"].data_ptr()
        st.vx, st.vy, st.vz = (self.vel[a].data_ptr() for a in range(3))
        st.wx, st.wy, st.wz = (self.omega[a].data_ptr() for a in range(3))
        if self.velocities:   # state under velocity Verlet: they live in the exchange record's aux channels
            vb = self._vel_base
            st.vx, st.vy, st.wz = (s["aux"][vb + c].data_ptr() for c in range(3))
        st.species = s["species"].data_ptr()
        _lib.check(self.lib.swarm_bind_state(self.h, self.cap, C.byref(st)))   # cap = upper bound; live count below
        act = _lib.ActionsDev()
        act.f_swim, act.tqx, act.tqy, act.tqz = s["f_swim"].data_ptr(), s["tq"][0].data_ptr(), s["tq"][1].data_ptr(), s["tqz"].data_ptr()
        _lib.check(self.lib.swarm_bind_actions(self.h, C.byref(act)))
        _lib.check(self.lib.swarm_bind_ids(self.h, s["gid"].data_ptr()))
        _lib.check(self.lib.swarm_bind_count(self.h, self.counts[self.cur].data_ptr()))
        _lib.check(self.lib.swarm_bind_images(self.h, s["image"].data_ptr()))   # vision cones: same-image pairs only
        self.sw.n = self.cap

    @property
    def local(self):
        return self.sets[self.cur]

    # ------------------------------------------------------------------ setup
    def scatter(self, pos_global, theta_global, species_global=None, capacity=None, buffer_capacity=None):
        """Every rank passes the same global arrays (host, particle index == global id) and keeps its part."""
        torch = self.torch
        pos = np.asarray(pos_global, dtype=np.float64)
        n_glob = pos.shape[0]
        idx, xl, im, owned = select_local(pos[:, 0], self.plan)
        p = self.plan
        frac = (p["slab_w"] + 2.0 * p["halo"]) / p["box_x"]
        cap = capacity or int(max(len(idx), frac * n_glob) * self.capacity_factor) + 1024
        bcap = buffer_capacity or int(max(1.0, p["halo"] / p["box_x"] * n_glob) * 1.5) + 1024   # 1.5 x the mean face population
        self._alloc(cap, bcap)
        n = len(idx)
        s = self.sets[0]
        up = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a, dtype=dt)).to(self.dev)
        s["x"][:n] = up(xl, np.float64)
        s["y"][:n] = up(pos[idx, 1], np.float64)
        s["theta"][:n] = up(np.asarray(theta_global, dtype=np.float64)[idx], np.float64)
        s["gid"][:n] = up(idx, np.int64)
        sp = np.zeros(n_glob, dtype=np.int32) if species_global is None else np.asarray(species_global, dtype=np.int32)
        s["species"][:n] = up(sp[idx], np.int32)
        s["image"][:n] = up(im, np.int32)
        s["owned"][:n] = up(owned, np.int32)
        self.n_global = n_glob
        self.cur, self.n, self.n_owned = 0, n, int(owned.sum())
        self.counts[0][:2] = torch.tensor([self.n, self.n_owned], dtype=torch.int32, device=self.dev)
        self.since_exchange = 0
        self._bind()

    def set_actions_global(self, f_swim_dev, tqz_dev):
        """manage_forces (espresso.py:1227-1235) from device arrays indexed by global id; ghosts included,
        because they are integrated redundantly."""
        s = self.local   # whole capacity: slots past the live count hold stale but valid ids and are never read
        torch = self.torch
        torch.index_select(f_swim_dev, 0, s["gid"], out=s["f_swim"])
        torch.index_select(tqz_dev, 0, s["gid"], out=s["tqz"])

    def _stream(self):
        return self.sw._stream()

    # --------------------------------------------------------------- timeline
    def _timed(self, kind):
        swarm = self

        class _Span:
            def __enter__(self):
                if swarm.profile:
                    self.e0 = swarm.torch.cuda.Event(enable_timing=True)
                    self.e1 = swarm.torch.cuda.Event(enable_timing=True)
                    self.e0.record()

            def __exit__(self, *exc):
                if swarm.profile:
                    self.e1.record()
                    swarm._events.append((kind, self.e0, self.e1))

        return _Span()

    def timeline_reset(self):
        self._events = []

    def timeline(self):
        """Milliseconds per phase since `timeline_reset` (synchronises)."""
        self.torch.cuda.current_stream(self.dev).synchronize()
        out = {"run_ms": 0.0, "exchange_ms": 0.0, "obs_ms": 0.0}
        for kind, e0, e1 in self._events:
            out[kind + "_ms"] += e0.elapsed_time(e1)
        return out

    # --------------------------------------------------------------- exchange
    def exchange(self):
        """halo refresh + migration, enqueued only: the new particle count stays on the device (see `sync`)."""
        if getattr(self.sw, "thermostat_type", "brownian") != "brownian" and not self.velocities:
            raise RuntimeError("this decomposed swarm ships no velocities: Brownian integrator only (create it with velocities=True "
                               "for the Langevin integrator)")
        b, g = self.bufs, self.geom
        old, new = self.sets[self.cur], self.sets[self.cur ^ 1]
        c_old, c_new = self.counts[self.cur], self.counts[self.cur ^ 1]
        a_old, a_new = self._dd(old), self._dd(new)
        _lib.check(self.lib.swarm_dd_pack(self.h, C.byref(g), self.cap, c_old.data_ptr(), C.byref(a_old),
                                          b["send_l"].data_ptr(), b["send_r"].data_ptr(), self._stream()))
        if self.peer is not None:
            # records straight into the neighbours' memory (P2P stores + a flag); my own arrive the same way
            pr, epoch = self.peer, self.exchanges + 1
            _lib.check(self.lib.swarm_dd_push(self.h, C.byref(g), b["send_l"].data_ptr(), b["send_r"].data_ptr(), pr["left"], pr["right"],
                                              epoch, self._stream()))
            _lib.check(self.lib.swarm_dd_wait(self.h, pr["region"], epoch, c_new.data_ptr(), self._stream()))
            recv_l = self.lib.swarm_dd_peer_recv(pr["region"], C.byref(g), 0, epoch & 1)
            recv_r = self.lib.swarm_dd_peer_recv(pr["region"], C.byref(g), 1, epoch & 1)
        else:
            self.transport.exchange(b["send_l"], b["send_r"], b["recv_l"], b["recv_r"])
            recv_l, recv_r = b["recv_l"].data_ptr(), b["recv_r"].data_ptr()
        _lib.check(self.lib.swarm_dd_merge(self.h, C.byref(g), self.cap, c_old.data_ptr(), C.byref(a_old),
                                           recv_l, recv_r, self.cap, C.byref(a_new),
                                           c_new.data_ptr(), self._stream()))
        self.cur ^= 1
        self.since_exchange = 0
        self.exchanges += 1
        self._bind()

    def sync(self):
        """Wait for the queued work, refresh `n` / `n_owned`, and raise if any exchange since the last call
        overflowed a buffer or lost a particle (the device keeps sticky error bits)."""
        for h in range(2):
            self.counts_host[h].copy_(self.counts[h], non_blocking=True)
        self.torch.cuda.current_stream(self.dev).synchronize()
        err = int(self.counts_host[0, 3]) | int(self.counts_host[1, 3])
        if err:
            why = [m for bit, m in ((1, "exchange buffer overflow"), (2, "an owned particle outran the halo"),
                                    (4, "local capacity exceeded"), (8, "a slab neighbour never delivered its halo records")) if err & bit]
            raise OverflowError("domain exchange failed: " + ", ".join(why))
        self.n, self.n_owned = int(self.counts_host[self.cur, 0]), int(self.counts_host[self.cur, 1])
        # a Verlet-list overflow halts the integrator kernels of this rank (every later launch returns at once) while
        # the exchanges would keep shipping the frozen slab: surface it here instead of returning stale positions
        st = self.sw.stats()
        if st.overflow:
            raise OverflowError(
                f"Verlet list capacity exceeded on rank {self.plan['rank']} (capacity {st.capacity}, "
                f"max_neighbors_seen >= {st.max_neighbors_seen}): raise max_neighbors or lower the skin")
        return self.n

    # ------------------------------------------------------------ integration
    def run(self, n_steps, step0):
        """integrator.run(n_steps) (espresso.py:1304-1306) on the decomposed swarm."""
        done, k = 0, self.plan["k"]
        # velocity Verlet: the parts of one run between two exchanges neither recalculate nor close a step (swarm_run_part),
        # so the cut run is the uncut one; the exchange in between ships x(t_n) and v(t_n - dt/2) of every record
        lv = self.velocities and getattr(self.sw, "thermostat_type", "brownian") == "langevin"
        while done < n_steps:
            if self.since_exchange >= k:
                with self._timed("exchange"):
                    self.exchange()
            m = min(k - self.since_exchange, n_steps - done)
            with self._timed("run"):
                if lv:
                    flags = (_lib.RUN_CONTINUE if done > 0 else 0) | (_lib.RUN_LEAVE_OPEN if done + m < n_steps else 0)
                    _lib.check(self.lib.swarm_run_part(self.h, int(m), int(step0 + done), flags, self._stream()))
                else:
                    _lib.check(self.lib.swarm_run(self.h, int(m), int(step0 + done), self._stream()))
            self.since_exchange += m
            done += m

    def stats(self):
        """Counters of the local handle; raises like `sync` when an exchange or a Verlet list overflowed."""
        self.sync()
        return self.sw.stats()

    # --------------------------------------------------------------- read-out
    def owned_host(self):
        """(gid, x_global, y, theta) of the owned particles, host arrays sorted by gid."""
        s, n = self.local, self.sync()
        own = s["owned"][:n].bool()
        gid = s["gid"][:n][own].cpu().numpy()
        xg = unfold_global(s["x"][:n][own].cpu().numpy(), s["image"][:n][own].cpu().numpy(), self.plan)
        y, th = s["y"][:n][own].cpu().numpy(), s["theta"][:n][own].cpu().numpy()
        o = np.argsort(gid, kind="stable")
        return gid[o], xg[o], y[o], th[o]

    def field_reward(self, src, box_length, decay=_lib.DECAY_LINEAR, c0=1.0, c1=-1.0, scale=1.0, clip=True, init_only=False,
                     defer=False):
        """GradientSensing reward (tasks/searching/gradient_sensing.py:92-150) of the owned particles and
        its global mean (trainers/trainer.py:92-95): returns (per-slot rewards, device tensor [sum, mean]).

        defer=True keeps this rank's sum on the device and returns (per-slot rewards, None): the reference takes the
        mean of an EPISODE's rewards after the episode (trainer.py:92-95), so the ranks need not meet in an all-reduce
        after every slice -- `reward_means()` reduces all deferred sums in one collective."""
        torch = self.torch
        if self._vel_base < 2:
            raise ValueError("the field history needs n_aux >= 2")
        a = self._dd(self.local)
        s3 = (C.c_double * 3)(*[float(v) for v in src])
        b2 = (C.c_double * 2)(float(box_length[0]), float(box_length[1]))
        with self._timed("obs"):
            _lib.check(self.lib.swarm_dd_obs_field(self.h, self.cap, self.counts[self.cur].data_ptr(), C.byref(a), self.plan["x_offset"], self.plan["box_x"], s3, b2,
                                                   int(decay), float(c0), float(c1), float(scale), int(bool(clip)),
                                                   int(bool(init_only)), self.out.data_ptr(), self._stream()))
            red = self.sw.reduce_mean(self.out)      # slots past the live count were written as 0; a fresh 2-element tensor per call
            if defer:
                self._deferred_sums.append(red[:1])
                return self.out, None
            tot = red[:1].clone()
            self.transport.all_reduce_sum(tot)
        return self.out, torch.cat([tot, tot / float(self.n_global)])

    def field_eval(self, src, box_length, decay, c0, c1, scale=1.0, clip=False, init_only=False, aux_base=0, all_local=False,
                   out=None):
        """ConcentrationField / GradientSensing value per local slot with the history in aux channels `aux_base`,
        `aux_base + 1` (swarm_dd_obs_field_ch); all_local=True evaluates the ghosts too (an observable that feeds the
        policy: a ghost must take its owner's action), False the owned particles only (a reward).  Enqueue only."""
        if self._vel_base < aux_base + 2:
            raise ValueError("this history needs aux channels %d and %d" % (aux_base, aux_base + 1))
        out = self.torch.empty(self.cap, dtype=self.torch.float64, device=self.dev) if out is None else out
        a = self._dd(self.local)
        s3 = (C.c_double * 3)(*[float(v) for v in src])
        b2 = (C.c_double * 2)(float(box_length[0]), float(box_length[1]))
        with self._timed("obs"):
            _lib.check(self.lib.swarm_dd_obs_field_ch(self.h, self.cap, self.counts[self.cur].data_ptr(), C.byref(a), self.plan["x_offset"],
                                                      self.plan["box_x"], s3, b2, int(decay), float(c0), float(c1), float(scale),
                                                      int(bool(clip)), int(bool(init_only)), int(aux_base), int(bool(all_local)),
                                                      out.data_ptr(), self._stream()))
        return out

    def reward_means(self):
        """Global reward mean of every slice evaluated with defer=True since the last call: ONE all-reduce of the vector
        of per-slice sums.  Returns a device tensor (n_slices,)."""
        torch = self.torch
        if not self._deferred_sums:
            return torch.zeros(0, dtype=torch.float64, device=self.dev)
        v = torch.cat(self._deferred_sums)
        self._deferred_sums = []
        with self._timed("obs"):
            self.transport.all_reduce_sum(v)
        return v / float(self.n_global)

    def close(self):
        pr, self.peer = self.peer, None
        if pr is not None and getattr(self.sw, "h", None):
            # neighbours may still be reading / writing: meet first when the process group is alive; without one (interpreter
            # exit) only the imported mappings are dropped and the own region is left to the driver's process teardown
            self.torch.cuda.synchronize(self.dev)
            dist = self.transport.dist
            alive = dist.is_initialized()
            if alive:
                dist.barrier(group=self.transport.group)
            for r in pr["opened"].values():
                self.lib.swarm_dd_peer_close(r)
            if alive:
                dist.barrier(group=self.transport.group)
                self.lib.swarm_dd_peer_free(pr["region"])
        self.sw.close()
