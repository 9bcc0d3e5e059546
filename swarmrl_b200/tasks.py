"""
Task rewards evaluated on the device.

  Task              /root/reference/swarmrl/tasks/task.py:15-118
  GradientSensing   /root/reference/swarmrl/tasks/searching/gradient_sensing.py:21-150
  SpeciesSearch     /root/reference/swarmrl/tasks/searching/species_search.py:17-184
  MultiTasking      /root/reference/swarmrl/tasks/multi_tasking.py:15-62
"""
from __future__ import annotations

import numpy as np

from .observables import _FieldHistory, _PairField, _apply_decay, _decay_spec, resolve_swarm


class Task:
    def __init__(self, particle_type: int = 0):
        self.particle_type = particle_type
        self._kill_switch = False

    def initialize(self, colloids):
        pass

    @property
    def kill_switch(self):
        return self._kill_switch

    @kill_switch.setter
    def kill_switch(self, value):
        self._kill_switch = value

    def get_colloid_indices(self, colloids, p_type: int = None):
        if p_type is None:
            p_type = self.particle_type
        return [i for i, c in enumerate(colloids) if c.type == p_type]

    def __call__(self, colloids):
        raise NotImplementedError("Implemented in child class.")


class GradientSensing(Task):
    """reward_i = clip(scale * (decay(|p_i - src|) - decay(|p_i_prev - src|)), 0, inf), history
    updated on every call (gradient_sensing.py:92-126)."""

    fused_mean = True   # call_device(view, want_mean=True) also leaves the reward mean on the device

    def __init__(self, source=np.array([0, 0, 0]), decay_function: callable = None,
                 box_length=np.array([1.0, 1.0, 0.0]), reward_scale_factor: int = 10, particle_type: int = 0):
        super().__init__(particle_type=particle_type)
        self.last_mean = None
        self.box_length = np.asarray(box_length, dtype=float)
        with np.errstate(divide="ignore", invalid="ignore"):
            self.source = np.asarray(source, dtype=float) / self.box_length
        self.decay_fn = decay_function
        self.reward_scale_factor = reward_scale_factor
        self._history = _FieldHistory()

    @property
    def _historic_positions(self):
        if self._history.empty() or not hasattr(self._history.hist, "cpu"):
            return {}
        h = self._history.hist.cpu().numpy().T
        return {str(i): h[i] for i in range(h.shape[0]) if self._history.valid[i]}

    def initialize(self, colloids):
        swarm, _ = resolve_swarm(colloids)
        self._history.init_from(swarm, self.box_length, mask=(swarm.ptype_host == int(self.particle_type)), all_local=False)

    def change_source(self, new_source):
        """Stores the source un-normalised, as the reference does (gradient_sensing.py:81-90)."""
        self.source = np.asarray(new_source, dtype=float)

    def call_device(self, colloids, want_mean=False):
        """want_mean=True: `self.last_mean` holds the device (sum, mean) of the returned rewards when the decay has a
        closed form (one launch for both), None otherwise."""
        swarm, _ = resolve_swarm(colloids)
        if self._history.empty():
            raise KeyError("GradientSensing was not initialized")
        row_of, rows = swarm.row_map(self.particle_type)
        spec = _decay_spec(self.decay_fn)
        self.last_mean = None
        if spec is not None and want_mean and rows > 0:
            out, self.last_mean = swarm.obs_field_mean(row_of, rows, self.source, self.box_length, self._history.hist, spec[0],
                                                       spec[1], spec[2], scale=self.reward_scale_factor, clip=True)
            return out
        if spec is not None:
            out, _, _ = swarm.obs_field(row_of, rows, self.source, self.box_length, self._history.hist, decay=spec[0],
                                        c0=spec[1], c1=spec[2], scale=self.reward_scale_factor, clip=True,
                                        want_dist=False)
            return out
        _, dn, dp = swarm.obs_field(row_of, rows, self.source, self.box_length, self._history.hist)
        delta = _apply_decay(self.decay_fn, dn) - _apply_decay(self.decay_fn, dp)
        return (self.reward_scale_factor * delta).clamp_min(0.0)

    def __call__(self, colloids):
        return self.call_device(colloids).cpu().numpy()


class SpeciesSearch(_PairField, Task):
    """reward_i = scale * clip(field_i - field_i_prev) with the field summed over `sensing_type`;
    avoid=True keeps the negative part (species_search.py:178-184)."""

    def __init__(self, decay_fn: callable, box_length: np.ndarray, sensing_type: int = 0, avoid: bool = False,
                 scale_factor: int = 100, particle_type: int = 0):
        Task.__init__(self, particle_type=particle_type)
        self._pf_init(decay_fn, box_length, sensing_type, scale_factor)
        self.avoid = avoid

    def initialize(self, colloids):
        self._pf_call(colloids, 0, True)

    def call_device(self, colloids):
        if self._hist is None:
            raise ValueError(f"{type(self).__name__} requires initialization. Please set the "
                             "initialize attribute of the gym to true and try again.")
        return self._pf_call(colloids, -1 if self.avoid else 1, False)

    def __call__(self, colloids):
        return self.call_device(colloids).cpu().numpy()


class MultiTasking(Task):
    """Sum of several tasks' rewards (multi_tasking.py:43-62)."""

    def __init__(self, particle_type: int = 0, tasks=None):
        super().__init__(particle_type)
        self.tasks = tasks or []

    def initialize(self, colloids):
        for item in self.tasks:
            item.initialize(colloids)

    def call_device(self, colloids):
        rewards = [item.call_device(colloids) for item in self.tasks]
        out = rewards[0].clone()
        for r in rewards[1:]:
            out += r
        return out

    def __call__(self, colloids):
        return self.call_device(colloids).cpu().numpy()
