"""
SwarmDeviceView -- what device-path agents / observables / tasks receive instead of a list of
`Colloid`s: direct access to the SoA device tensors of a running SwarmEngine (zero copy; torch
tensors export DLPack, so JAX / CuPy consumers can alias them without a transfer).
"""
from __future__ import annotations


class SwarmDeviceView:
    def __init__(self, engine):
        self.engine = engine
        self.swarm = engine._backend

    @property
    def n(self):
        return self.swarm.n

    @property
    def types(self):
        """host int array of particle types (static after the first integrate)."""
        return self.engine._types

    # (3, n) / (n,) fp64 device tensors --------------------------------------------------------
    @property
    def pos(self):
        return self.swarm.pos[:, : self.n]

    @property
    def theta(self):
        return self.swarm.theta[: self.n]

    @property
    def vel(self):
        return self.swarm.vel[:, : self.n]

    @property
    def f_swim(self):
        return self.swarm.f_swim[: self.n]

    @property
    def torque(self):
        return self.swarm.torque[:, : self.n]

    def directors(self):
        return self.swarm.directors()

    def dlpack(self, name):
        """DLPack capsule of one state tensor ('pos', 'theta', 'vel', 'f_swim', 'torque')."""
        import torch.utils.dlpack as dl

        return dl.to_dlpack(getattr(self, name))
