"""
DomainBackend -- the slab-decomposed swarm (swarmrl_b200/domain.py) behind `SwarmEngine`.

`SwarmEngine(..., domain=dict(world=W, rank=r, group=g, exchange_every=k, ...))` creates this backend instead of a
`DeviceSwarm`; `integrate(n_slices, ForceFunction({...: DeviceAgent(...)}))` then runs the reference's loop
(swarmrl/engine/espresso.py:1251-1308) on every rank over its slab + halos, with the same call sequence as on one GPU.
The reference has no counterpart (single-rank ESPResSo, espresso.py:193): the contract is "same trajectories, same
observables and same rewards as the single-GPU engine on the same inputs".

How the agent pipeline stays exact on a decomposed swarm
  * every LOCAL particle (owned or ghost) gets an observable, a policy evaluation and an action: ghosts are integrated
    redundantly, so they must take the action their owner computes.  The policy's sampling noise is keyed by the global
    particle id (csrc/swarm_policy.cu), the observable is a function of positions only, so a ghost whose neighbourhood is
    complete reproduces its owner's action bit for bit;
  * an observable with a finite range R (SubdividedVisionCones: `vision_range`) needs the halo wider by R
    (`halo_extra`): the local set is refreshed right before every observation (`ensure_fresh`), so it is exact over the
    whole kept region, the observable is exact for everything within H = k r_cut (1 + margin) of the slab, and a wrong
    action further out travels inward by at most r_cut per step -- it cannot reach the slab before the next exchange;
  * per-particle histories (ConcentrationField, GradientSensing keep the previous normalised position) travel with the
    particle in the exchange record's aux channels (`new_history`);
  * vision is measured on unfolded positions without a minimum image, as in the reference: pairs are seen only within
    one periodic image (swarm_bind_images);
  * `thermostat_type="langevin"`: velocities are state under velocity Verlet, so vx, vy, omega_z travel as three more aux
    channels and the runs between two exchanges are cut with swarm_run_part (no recalculation, no closing half kick at a cut);
  * rewards are evaluated for owned particles; their sum stays on the device and `all_reduce_means` reduces a whole
    rollout's sums in ONE collective.

Not supported on the decomposed swarm (raises): 3-D, walls, constant external force, `new_direction` actions, policies without a fused `act` (their sampling is keyed by row, not by particle id).
"""
from __future__ import annotations

import numpy as np

from . import _lib
from .domain import DomainSwarm


class DomainHistory:
    """A pair of aux channels of the exchange record holding a previous normalised position per particle."""

    def __init__(self, base, all_local):
        self.base, self.all_local = int(base), bool(all_local)


class DomainBackend:
    is_domain = True

    def __init__(self, n_dims, box, periodic, dt, kT, seed, species, wca, walls=None, skin=0.4, persistent=True, *,
                 world=1, rank=0, group=None, exchange_every=8, margin=0.25, halo_extra=0.0, device=None,
                 max_neighbors=0, capacity_factor=1.3, velocities=False):
        if int(n_dims) != 2:
            raise NotImplementedError("the domain-decomposed swarm is 2-D (slabs along x)")
        if walls is not None:
            raise NotImplementedError("walls are not supported on the domain-decomposed swarm")
        if not bool(periodic):
            raise NotImplementedError("the domain-decomposed swarm is periodic")
        self.n_dims = 2
        self.box = np.array([float(b) for b in box])
        self.world, self.rank = int(world), int(rank)
        self.dd = DomainSwarm((self.box[0], self.box[1]), dt, kT, seed, species, wca, exchange_every=exchange_every,
                              margin=margin, skin=skin, world=world, rank=rank, group=group, device=device, n_aux=4,
                              max_neighbors=max_neighbors, capacity_factor=capacity_factor, persistent=persistent,
                              halo_extra=halo_extra, velocities=velocities)
        self.sw = self.dd.sw
        self.torch, self.dev, self.lib, self.h = self.dd.torch, self.dd.dev, self.dd.lib, self.dd.h
        self.type_of_species = self.sw.type_of_species
        self.n = 0
        self.ptype_host = np.zeros(0, dtype=np.int32)
        self._row_maps = {}
        self._hist_used = 0
        self._fresh = False
        self._radii_cache = None
        self.thermostat_type = "brownian"

    # ------------------------------------------------------------------ setup
    def set_state(self, pos, species, theta=None, quat=None):
        if theta is None:
            raise NotImplementedError("the domain-decomposed swarm is 2-D: pass theta")
        pos = np.asarray(pos, dtype=np.float64).reshape(-1, 3)
        sp = np.ascontiguousarray(species, dtype=np.int32)
        self.n = pos.shape[0]
        self.species_host = sp.copy()
        self.ptype_host = self.type_of_species[sp] if self.n else np.zeros(0, dtype=np.int32)
        self._type_lut = self.torch.from_numpy(np.ascontiguousarray(self.type_of_species, dtype=np.int64)).to(self.dev)
        self.dd.scatter(pos, theta, sp)
        self._arange = self.torch.arange(self.dd.cap, dtype=self.torch.int32, device=self.dev)
        self._row_maps = {}
        self._fresh = True

    def set_thermostat(self, kT, seed):
        self.sw.set_thermostat(kT, seed)

    def set_integrator(self, thermostat_type):
        if thermostat_type != "brownian" and not self.dd.velocities:
            raise NotImplementedError("this decomposed swarm ships no velocities: create it with velocities=True for Langevin")
        self.sw.set_integrator(thermostat_type)
        self.thermostat_type = thermostat_type

    def set_max_neighbors(self, cap):
        self.sw.set_max_neighbors(cap)

    def set_ext_force_host(self, F):
        raise NotImplementedError("constant external forces are not supported on the domain-decomposed swarm")

    def set_new_directions_host(self, nd, has):
        raise NotImplementedError("new_direction actions are not supported on the domain-decomposed swarm")

    # ------------------------------------------------------------ integration
    def ensure_fresh(self):
        """Refresh the halos unless nothing has moved since the last exchange: called before every observation."""
        if not self._fresh:
            self.dd.exchange()
            self._fresh = True

    def run(self, n_steps, step0):
        if int(n_steps) > 0:
            self.dd.run(int(n_steps), int(step0))
            self._fresh = False

    def stats(self):
        return self.dd.stats()

    def run_checked(self, n_steps, step0, grow=True):
        """run + sync.  A Verlet overflow cannot be replayed here (the exchanges in between are gone): it raises."""
        self.run(n_steps, step0)
        st = self.dd.stats()
        if int(st.steps_done) == 0 and int(n_steps) > 0:
            raise RuntimeError("the integrator of this rank made no progress")
        return st

    def relax_checked(self, n_iter=1000, gamma=0.1, max_disp=0.1):
        """_remove_overlap's steepest descent (espresso.py:1163-1167) with the halos refreshed every k iterations: a
        descent step moves a particle by at most max_disp, well inside what the halo margin allows per step."""
        dd, k, done = self.dd, self.dd.plan["k"], 0
        while done < int(n_iter):
            if dd.since_exchange >= k:
                dd.exchange()
            m = min(k - dd.since_exchange, int(n_iter) - done)
            _lib.check(self.lib.swarm_relax(self.h, int(m), float(gamma), float(max_disp), dd._stream()))
            dd.since_exchange += m
            done += m
        dd.exchange()
        self._fresh = True
        return dd.stats()

    # ---------------------------------------------------------------- actions
    def set_actions_host(self, f_swim, torque):
        """manage_forces through host arrays indexed by global id (object path)."""
        torch = self.torch
        f = torch.from_numpy(np.ascontiguousarray(f_swim, dtype=np.float64)).to(self.dev)
        t = np.ascontiguousarray(torque, dtype=np.float64).reshape(-1, 3)
        if np.any(t[:, :2] != 0.0):
            raise NotImplementedError("2-D swarm: only the z torque is supported")
        self.dd.set_actions_global(f, torch.from_numpy(np.ascontiguousarray(t[:, 2])).to(self.dev))

    @property
    def f_swim(self):
        return self.dd.local["f_swim"]

    @property
    def torque(self):
        return self.dd.local["tq"]

    # ------------------------------------------------------------- observables
    def row_map(self, particle_type):
        """(row_of, rows): rows are the local SLOTS (rows = capacity); row_of[i] = i for a live slot of `particle_type`,
        else -1.  Built on the device from the live count of the current exchange -- no host synchronisation.  Refreshes
        the halos first when particles have moved since the last exchange (the rows of a stale set would not be the rows
        of the set the observable is evaluated on); for a reward at the end of a slice that is the exchange the next
        slice's observation needs anyway."""
        self.ensure_fresh()
        dd = self.dd
        key = (int(particle_type), dd.exchanges, dd.cur)
        hit = self._row_maps.get(key)
        if hit is None:
            torch = self.torch
            s = dd.local
            live = self._arange < dd.counts[dd.cur][0]
            typ = self._type_lut[s["species"].long().clamp_(0, self._type_lut.numel() - 1)]
            row_of = torch.where(live & (typ == int(particle_type)), self._arange, torch.full_like(self._arange, -1))
            self._row_maps = {key: (row_of.contiguous(), dd.cap)}
            hit = self._row_maps[key]
        return hit

    def new_history(self, box_length, all_local):
        """Two aux channels seeded with the current normalised positions (the reference's `initialize`)."""
        if self._hist_used + 2 > self.dd._vel_base:
            raise ValueError("at most %d position histories fit the exchange record" % (self.dd._vel_base // 2))
        h = DomainHistory(self._hist_used, all_local)
        self._hist_used += 2
        self.dd.field_eval((0.0, 0.0, 0.0), (box_length[0], box_length[1]), _lib.DECAY_LINEAR, 0.0, 0.0, init_only=True,
                           aux_base=h.base, all_local=True)
        return h

    def _field(self, src, box_length, hist, decay, c0, c1, scale, clip):
        if not isinstance(hist, DomainHistory):
            raise TypeError("a domain-decomposed swarm keeps its histories in the exchange record (new_history)")
        if decay == _lib.DECAY_NONE:
            raise NotImplementedError("the domain-decomposed swarm needs a closed-form decay (swarmrl_b200.decay)")
        if hist.all_local:
            self.ensure_fresh()
        return self.dd.field_eval(src, box_length, decay, c0, c1, scale=scale, clip=clip, aux_base=hist.base, all_local=hist.all_local)

    def obs_field(self, row_of, n_rows, src, box_length, hist, decay=_lib.DECAY_NONE, c0=0.0, c1=0.0, scale=1.0, clip=False,
                  want_dist=True):
        return self._field(src, box_length, hist, decay, c0, c1, scale, clip), None, None

    def obs_field_mean(self, row_of, n_rows, src, box_length, hist, decay, c0, c1, scale=1.0, clip=False):
        """(rewards per slot, [this rank's sum, this rank's sum / n_global]): `all_reduce_means` finishes the mean."""
        out = self._field(src, box_length, hist, decay, c0, c1, scale, clip)
        red = self.sw.reduce_mean(out)
        return out, self.torch.stack([red[0], red[0] / float(max(self.n, 1))])

    def obs_vision(self, row_of, n_rows, vision_range, half_angle, n_cones, radii, type_col, n_cols, dirs=None):
        """SubdividedVisionCones per local slot; `radii` is indexed by global particle id."""
        dd, torch = self.dd, self.torch
        if float(vision_range) > dd.plan["halo_extra"] + 1e-12:
            raise ValueError(f"vision_range {vision_range:g} exceeds the halo's observable band ({dd.plan['halo_extra']:g}): "
                             "create the engine with domain=dict(..., halo_extra=vision_range)")
        self.ensure_fresh()
        key = (dd.exchanges, dd.cur, radii.data_ptr())
        if self._radii_cache is None or self._radii_cache[0] != key:
            gid = dd.local["gid"].clamp(0, max(self.n - 1, 0))
            self._radii_cache = (key, radii[gid].contiguous())
        out = torch.empty((dd.cap, int(n_cones), int(n_cols)), dtype=torch.float64, device=self.dev)
        _lib.check(self.lib.swarm_obs_vision(self.h, row_of.data_ptr(), int(dd.cap), float(vision_range), float(half_angle),
                                             int(n_cones), self._radii_cache[1].data_ptr(), type_col.data_ptr(),
                                             int(type_col.numel()), int(n_cols), None, out.data_ptr(), dd._stream()))
        return out

    def policy_forward(self, *a, **kw):
        return self.sw.policy_forward(*a, **kw)

    def reduce_mean(self, values):
        red = self.sw.reduce_mean(values)
        return self.torch.stack([red[0], red[0] / float(max(self.n, 1))])

    def all_reduce_means(self, means):
        """[sum, mean] over the WHOLE swarm of every reward evaluation of a rollout: `means` is the list of per-rank
        2-element tensors an agent collected (`DeviceAgent.reward_means`); one collective for all of them."""
        if not means:
            return self.torch.zeros((0, 2), dtype=self.torch.float64, device=self.dev)
        v = self.torch.stack(list(means)).contiguous()
        self.dd.transport.all_reduce_sum(v)
        return v

    # --------------------------------------------------------------- read-out
    def particle_data_host(self):
        """get_particle_data (espresso.py:1320-1336) of the WHOLE swarm on every rank -- a collective: three (n, 3) host
        arrays indexed by particle id, gathered from the owners."""
        dd = self.dd
        s, n = dd.local, dd.sync()
        own = s["owned"][:n].bool()
        gid = s["gid"][:n][own].cpu().numpy()
        from .domain import unfold_global

        xg = unfold_global(s["x"][:n][own].cpu().numpy(), s["image"][:n][own].cpu().numpy(), dd.plan)
        y, th = s["y"][:n][own].cpu().numpy(), s["theta"][:n][own].cpu().numpy()
        if dd.velocities:
            v = self.torch.stack([s["aux"][dd._vel_base + c][:n][own] for c in range(2)]).cpu().numpy()
        else:
            v = dd.vel[:, :n][:, own].cpu().numpy()
        part = (gid, xg, y, th, v)
        if self.world > 1:
            parts = [None] * self.world
            dd.transport.dist.all_gather_object(parts, part, group=dd.transport.group)
        else:
            parts = [part]
        pos, vel, dr = (np.zeros((self.n, 3)) for _ in range(3))
        seen = 0
        for g, x_, y_, t_, v_ in parts:
            pos[g, 0], pos[g, 1] = x_, y_
            dr[g, 0], dr[g, 1] = np.cos(t_), np.sin(t_)
            vel[g, 0], vel[g, 1] = v_[0], v_[1]
            seen += len(g)
        if seen != self.n:
            raise RuntimeError(f"{seen} owned particles for a swarm of {self.n}: ownership is broken")
        return pos, vel, dr

    def synchronize(self):
        self.torch.cuda.current_stream(self.dev).synchronize()

    def close(self):
        self.dd.close()
