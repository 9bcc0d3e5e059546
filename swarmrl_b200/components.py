"""
Data carriers of the agent <-> engine bridge, with the reference's names and fields:
`Colloid` (/root/reference/swarmrl/components/colloid.py:11-49) and `Action`
(/root/reference/swarmrl/actions/actions.py:10-19).  When the reference package itself is
importable its own classes can be used interchangeably: the engine only reads attributes.
"""
from __future__ import annotations

import dataclasses

import numpy as np


@dataclasses.dataclass(frozen=True)
class Colloid:
    pos: np.ndarray
    director: np.ndarray
    id: int
    velocity: np.ndarray = None
    type: int = 0

    def __eq__(self, other):
        return self.id == other.id

    def __hash__(self):
        return hash(int(self.id))


@dataclasses.dataclass
class Action:
    """force: scalar along the director; torque: lab-frame 3-vector; new_direction: unit 3-vector."""

    id = 0
    force: float = 0.0
    torque: np.ndarray = None
    new_direction: np.ndarray = None
