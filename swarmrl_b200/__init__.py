"""
swarmrl_b200 -- B200-native (sm_100a) engine for the SwarmRL rollout hot path.

Host-side mirror of the reference's interfaces for this path (`SwarmEngine` == the reference's
`EspressoMD`, `MDParams`, `Colloid`, `Action`, `ForceFunction`, observables, tasks, agents) over
the C-ABI of libswarm_b200.so (include/swarm_b200.h).  There is no CPU fallback: constructing a
`DeviceSwarm` without the library or without a CUDA device raises.
"""
from . import _lib, units  # noqa: F401
from .agents import (ActorCriticAgent, Agent, ClassicalAgent, ConstForce, ConstForceAndTorque,  # noqa: F401
                     ConstTorque, CounterPolicy, DeviceAgent, FusedActorCritic, ToConstDirection, TrajectoryInformation)
from .components import Action, Colloid  # noqa: F401
from .engine import Engine, EspressoMD, MDParams, SwarmEngine  # noqa: F401
from .force_functions import ForceFunction  # noqa: F401

__version__ = "0.1.0"
