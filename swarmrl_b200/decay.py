"""
Closed-form decay functions that the field kernels evaluate in-kernel (swarm_obs_field's `decay`
argument).  They are ordinary callables too, so they can be handed to the reference's own
ConcentrationField / GradientSensing unchanged.  Any other Python callable still works with the
device observables: the kernel then returns distances and the callable is applied to the tensor.
"""
from __future__ import annotations

import numpy as np

from . import _lib


class Linear:
    """c0 + c1 * d   (the reference's tests use 1 - d: test_rl_trainers.py:29-33)."""

    def __init__(self, c0=1.0, c1=-1.0):
        self.c0, self.c1 = float(c0), float(c1)

    def __call__(self, d):
        return self.c0 + self.c1 * d

    def device_spec(self):
        return _lib.DECAY_LINEAR, self.c0, self.c1


class Exponential:
    """c0 * exp(-c1 * d)"""

    def __init__(self, c0=1.0, c1=1.0):
        self.c0, self.c1 = float(c0), float(c1)

    def __call__(self, d):
        if hasattr(d, "exp"):
            return self.c0 * (-self.c1 * d).exp()
        return self.c0 * np.exp(-self.c1 * d)

    def device_spec(self):
        return _lib.DECAY_EXP, self.c0, self.c1


class Inverse:
    """c0 / (d + c1)"""

    def __init__(self, c0=1.0, c1=1.0):
        self.c0, self.c1 = float(c0), float(c1)

    def __call__(self, d):
        return self.c0 / (d + self.c1)

    def device_spec(self):
        return _lib.DECAY_INVERSE, self.c0, self.c1
