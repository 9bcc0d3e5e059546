"""
Per-agent observables with the reference's class names and constructor signatures, evaluated by
the sm_100a kernels of libswarm_b200.so (swarm_obs_field / swarm_obs_vision / swarm_obs_position).

  ConcentrationField     /root/reference/swarmrl/observables/concentration_field.py:22-138
  PositionObservable     /root/reference/swarmrl/observables/position.py:15-60
  Director               /root/reference/swarmrl/observables/director.py:14-58
  SubdividedVisionCones  /root/reference/swarmrl/observables/subdivided_vision_cones.py:17-258

`compute_observable(colloids)` accepts either `engine.colloids` of a SwarmEngine (the state is
already in HBM: nothing is copied in) or any sequence of Colloid-like objects (uploaded once to a
scratch DeviceSwarm).  Results come back as numpy arrays shaped like the reference's; the
`*_device` variants return torch CUDA tensors and leave the data in HBM.  There is no CPU path.

Numerics: the reference evaluates these in JAX float32 after normalising positions in float64
(SURVEY.md App. B item 10); this build uses fp64 throughout.
"""
from __future__ import annotations

import numpy as np

from . import _lib


class Observable:
    """/root/reference/swarmrl/observables/observable.py:10-96."""

    def __init__(self, particle_type: int):
        self._shape = None
        self.particle_type: int = particle_type

    def initialize(self, colloids):
        pass

    def get_colloid_indices(self, colloids, p_type: int = None):
        if p_type is None:
            p_type = self.particle_type
        return [i for i, c in enumerate(colloids) if c.type == p_type]

    def compute_observable(self, colloids):
        raise NotImplementedError("Implemented in child class.")

    @property
    def observable_shape(self):
        return self._shape


# ------------------------------------------------------------------------------------------------
# resolving "colloids" to a DeviceSwarm
# ------------------------------------------------------------------------------------------------
_SCRATCH_SPECIES = [dict(type=t, gamma_t=(1.0,) * 3, gamma_r=(1.0,) * 3, mass=1.0, rinertia=(1.0,) * 3)
                    for t in range(_lib.MAX_TYPES)]


def _scratch_swarm(colloids):
    """Upload a sequence of Colloid-like objects (pos, director, type) to a scratch DeviceSwarm."""
    from .device import DeviceSwarm, director_to_quat

    n = len(colloids)
    pos = np.array([np.asarray(c.pos, dtype=float) for c in colloids]).reshape(n, 3)
    dirs = np.array([np.asarray(c.director, dtype=float) for c in colloids]).reshape(n, 3)
    types = np.array([int(c.type) for c in colloids], dtype=np.int32)
    if n and (types.min() < 0 or types.max() >= _lib.MAX_TYPES):
        raise ValueError("particle types must be in 0..7")
    T = _lib.MAX_TYPES
    z = np.zeros((T, T))
    sw = DeviceSwarm(n_dims=3, box=(1.0, 1.0, 1.0), periodic=False, dt=1.0, kT=0.0, seed=0,
                     species=_SCRATCH_SPECIES, wca=dict(sigma=z, eps=z, cut=z))
    sw.set_state(pos, types, quat=director_to_quat(dirs) if n else None)
    # exact directors (not re-derived from a quaternion): keep them beside the swarm
    sw._exact_dirs = sw.torch.from_numpy(np.ascontiguousarray(dirs.T)).to(sw.dev)
    return sw


def resolve_swarm(colloids):
    """(DeviceSwarm, ids) for an engine view, a device view or a plain list of colloids."""
    from .device_view import SwarmDeviceView
    from .engine import ColloidsView

    if isinstance(colloids, ColloidsView):
        e = colloids.engine
        if e._backend is None:
            e._create_backend()
        return e._backend, None
    if isinstance(colloids, SwarmDeviceView):
        return colloids.swarm, None
    return _scratch_swarm(colloids), [int(np.copy(c.id)) for c in colloids]


def _decay_spec(fn):
    """(kind, c0, c1) when `fn` is one of the closed forms in swarmrl_b200.decay, else None."""
    spec = getattr(fn, "device_spec", None)
    return spec() if callable(spec) else None


def _apply_decay(fn, d):
    """Apply a user callable to a device tensor of distances; element-wise numpy round trip only
    if the callable cannot operate on torch tensors (it is user code, not a kernel fallback)."""
    try:
        out = fn(d)
        if hasattr(out, "device"):
            return out
    except Exception:
        pass
    import torch

    return torch.from_numpy(np.asarray(fn(d.cpu().numpy()), dtype=np.float64)).to(d.device)


class _FieldHistory:
    """Device-resident previous normalised positions, indexed by particle id (the reference keeps
    a dict keyed by str(id): concentration_field.py:79-82, gradient_sensing.py:77-79)."""

    def __init__(self):
        self.hist = None      # (3, n) device tensor
        self.ids = None       # host ids for list inputs
        self.valid = None     # host bool (n,), which ids have been initialised

    def empty(self):
        return self.hist is None

    def init_from(self, swarm, box_length, mask=None, all_local=True):
        if getattr(swarm, "is_domain", False):
            # decomposed swarm: the history travels with the particle in the exchange record (domain_engine.py);
            # all_local = the ghosts are evaluated too (an observable), False = owned particles only (a reward)
            self.hist = swarm.new_history(box_length, all_local)
            self.valid = np.ones(swarm.n, dtype=bool) if mask is None else np.asarray(mask, dtype=bool)
            return
        torch = swarm.torch
        b = torch.tensor([float(v) for v in box_length], dtype=torch.float64, device=swarm.dev)
        self.hist = (swarm.pos[:, : swarm.n] / b[:, None]).contiguous()
        self.valid = np.ones(swarm.n, dtype=bool) if mask is None else np.asarray(mask, dtype=bool)


class ConcentrationField(Observable):
    def __init__(self, source: np.ndarray, decay_fn: callable, box_length: np.ndarray, scale_factor: int = 100,
                 particle_type: int = 0):
        super().__init__(particle_type=particle_type)
        self.box_length = np.asarray(box_length, dtype=float)
        self.source = np.asarray(source, dtype=float) / self.box_length
        self.decay_fn = decay_fn
        self.scale_factor = scale_factor
        self._observable_shape = (3,)
        self._history = _FieldHistory()

    @property
    def _historic_positions(self):
        """dict view {str(id): normalised position}, as the reference stores it."""
        if self._history.empty() or not hasattr(self._history.hist, "cpu"):
            return {}
        h = self._history.hist.cpu().numpy().T
        return {str(i): h[i] for i in range(h.shape[0]) if self._history.valid[i]}

    def initialize(self, colloids):
        swarm, _ = resolve_swarm(colloids)
        self._history.init_from(swarm, self.box_length)  # every colloid, any type (:76-82)

    def compute_observable_device(self, colloids):
        swarm, _ = resolve_swarm(colloids)
        if self._history.empty():
            raise ValueError(f"{type(self).__name__} requires initialization. Please set the "
                             "initialize attribute of the gym to true and try again.")
        row_of, rows = swarm.row_map(self.particle_type)
        spec = _decay_spec(self.decay_fn)
        if spec is not None:
            out, _, _ = swarm.obs_field(row_of, rows, self.source, self.box_length, self._history.hist, decay=spec[0],
                                        c0=spec[1], c1=spec[2], scale=self.scale_factor, want_dist=False)
            return out.reshape(-1, 1)
        _, dn, dp = swarm.obs_field(row_of, rows, self.source, self.box_length, self._history.hist)
        delta = _apply_decay(self.decay_fn, dn) - _apply_decay(self.decay_fn, dp)
        return (self.scale_factor * delta).reshape(-1, 1)

    def compute_observable(self, colloids):
        return self.compute_observable_device(colloids).cpu().numpy()


class PositionObservable(Observable):
    def __init__(self, box_length: np.ndarray, particle_type: int = 0):
        super().__init__(particle_type=particle_type)
        self.box_length = np.asarray(box_length, dtype=float)

    def compute_observable_device(self, colloids):
        swarm, _ = resolve_swarm(colloids)
        row_of, rows = swarm.row_map(self.particle_type)
        return swarm.obs_position(row_of, rows, self.box_length)

    def compute_observable(self, colloids):
        return self.compute_observable_device(colloids).cpu().numpy()


class Director(Observable):
    def __init__(self, particle_type: int = 0):
        super().__init__(particle_type=particle_type)

    def compute_observable_device(self, colloids):
        swarm, _ = resolve_swarm(colloids)
        d = getattr(swarm, "_exact_dirs", None)
        if d is None:
            d = swarm.directors()
        sel = np.nonzero(swarm.ptype_host == int(self.particle_type))[0]
        idx = swarm.torch.from_numpy(sel).to(swarm.dev)
        return d[:, idx].T.contiguous()

    def compute_observable(self, colloids):
        return self.compute_observable_device(colloids).cpu().numpy()


class SubdividedVisionCones(Observable):
    """Quirks kept: the observer itself is visited (distance 0 => no
    cone), strict rim inequalities, the weight uses the SEEN particle's radius, `detected_types`
    is auto-detected (sorted unique types) on first use."""

    def __init__(self, vision_range: float, vision_half_angle: float, n_cones: int, radii, detected_types=None,
                 particle_type: int = 0):
        super().__init__(particle_type=particle_type)
        self.vision_range = vision_range
        self.vision_half_angle = vision_half_angle
        self.n_cones = n_cones
        self.radii = radii
        self.detected_types = detected_types
        self._radii_dev = None

    def _detect_all_things_to_see(self, types):
        self.detected_types = np.array(np.sort(np.unique(types)))

    def compute_observable_device(self, colloids):
        swarm, _ = resolve_swarm(colloids)
        torch = swarm.torch
        if self.detected_types is None:
            self._detect_all_things_to_see(swarm.ptype_host)
        det = [int(t) for t in self.detected_types]
        ckey = (id(swarm), swarm.n, tuple(det))
        if getattr(self, "_type_col_key", None) != ckey:
            tmax = max([int(swarm.ptype_host.max(initial=0))] + det) + 1
            type_col = -np.ones(tmax, dtype=np.int32)
            for c, t in enumerate(det):
                type_col[t] = c
            self._type_col_d = torch.from_numpy(type_col).to(swarm.dev)
            self._type_col_key = ckey
        type_col_d = self._type_col_d
        if self._radii_dev is None or self._radii_dev.numel() != max(swarm.n, 1) or self._radii_dev.device != swarm.dev:
            r = np.zeros(max(swarm.n, 1))
            r[: swarm.n] = np.asarray(self.radii, dtype=float)[: swarm.n]
            self._radii_dev = torch.from_numpy(r).to(swarm.dev)
        row_of, rows = swarm.row_map(self.particle_type)
        exact = getattr(swarm, "_exact_dirs", None)
        return swarm.obs_vision(row_of, rows, self.vision_range, self.vision_half_angle, self.n_cones,
                                self._radii_dev, type_col_d, len(det), dirs=exact)

    def compute_observable(self, colloids):
        out = self.compute_observable_device(colloids).cpu().numpy()
        return [out[i] for i in range(out.shape[0])]


class _PairField:
    """Shared body of ParticleSensing (observable) and SpeciesSearch (task): a field summed over all
    particles of `sensing_type`, history per observer (the reference's dict keyed by str(id))."""

    def _pf_init(self, decay_fn, box_length, sensing_type, scale_factor):
        self.decay_fn = decay_fn
        self.box_length = np.asarray(box_length, dtype=float)
        self.sensing_type = sensing_type
        self.scale_factor = scale_factor
        self._hist = None

    @property
    def historical_field(self):
        if self._hist is None:
            return {}
        return {str(i): float(v) for i, v in zip(self._hist_ids, self._hist.cpu().numpy())}

    def _pf_call(self, colloids, clip_mode, init_only):
        swarm, ids = resolve_swarm(colloids)
        spec = _decay_spec(self.decay_fn)
        if spec is None:
            raise NotImplementedError(
                "the all-pairs field is summed in-kernel: pass one of swarmrl_b200.decay.{Linear, Exponential, Inverse} "
                "as decay_fn (an arbitrary Python callable would need the n x m distance matrix on the host)")
        row_of, rows = swarm.row_map(self.particle_type)
        sensed, m = swarm.sensed_index(self.sensing_type)
        if self._hist is None or init_only:
            self._hist = swarm.torch.zeros(max(rows, 1), dtype=swarm.torch.float64, device=swarm.dev)
            sel = np.nonzero(swarm.ptype_host == int(self.particle_type))[0]
            self._hist_ids = [int(ids[i]) for i in sel] if ids is not None else [int(i) for i in sel]
        return swarm.obs_pair_field(row_of, rows, sensed, m, self.box_length, self._hist, spec[0], spec[1], spec[2],
                                    scale=self.scale_factor, clip_mode=clip_mode, init_only=init_only)


class ParticleSensing(_PairField, Observable):
    """/root/reference/swarmrl/observables/particle_sensing.py:18-178."""

    def __init__(self, decay_fn: callable, box_length: np.ndarray, sensing_type: int = 0, scale_factor: int = 100,
                 particle_type: int = 0):
        Observable.__init__(self, particle_type=particle_type)
        self._pf_init(decay_fn, box_length, sensing_type, scale_factor)

    def initialize(self, colloids):
        self._pf_call(colloids, 0, True)

    def compute_observable_device(self, colloids):
        if self._hist is None:
            raise ValueError(f"{type(self).__name__} requires initialization. Please set the "
                             "initialize attribute of the gym to true and try again.")
        return self._pf_call(colloids, 0, False).reshape(-1, 1)

    def compute_observable(self, colloids):
        return self.compute_observable_device(colloids).cpu().numpy()


class MultiSensing(Observable):
    """/root/reference/swarmrl/observables/multi_sensing.py:18-88: the observables' outputs side by side,
    one row per colloid of the (common) particle type."""

    def __init__(self, observables):
        super().__init__(particle_type=observables[0].particle_type if observables else 0)
        self.observables = observables

    def initialize(self, colloids):
        for item in self.observables:
            item.initialize(colloids)

    def compute_observable_device(self, colloids):
        """(n_colloids, sum of feature sizes) device tensor: what a device policy consumes."""
        import torch

        outs = [item.compute_observable_device(colloids) for item in self.observables]
        return torch.cat([o.reshape(o.shape[0], -1) for o in outs], dim=1)

    def compute_observable(self, colloids):
        """Object array (n_colloids, n_observables) of the per-observable arrays, as the reference returns
        (multi_sensing.py:75-88)."""
        per_obs = [np.asarray(item.compute_observable(colloids)) for item in self.observables]
        n = len(per_obs[0])
        out = np.empty((n, len(per_obs)), dtype=object)
        for k, item in enumerate(per_obs):
            for j in range(n):
                out[j, k] = item[j]
        return out
