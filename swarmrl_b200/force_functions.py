"""
ForceFunction -- bridge between agents and an engine, same contract as the reference's
(/root/reference/swarmrl/force_functions/force_fn.py:13-106): agents are keyed by
`str(particle_type)`; every agent is asked for its actions, which are handed out, in order, to
the colloids whose `str(type)` equals the agent's key; colloids without an agent keep the default
`Action()`; the kill switch is the OR over the agents' switches.

Additions for the device path (not in the reference): `calc_action_device` /
`calc_reward_device` are used by `SwarmEngine` when EVERY agent implements them, so that no
per-colloid Python object is created ("drop-in object API vs SoA speed").
"""
from __future__ import annotations

import typing

import numpy as np

from .components import Action


class ForceFunction:
    _kill_switch: bool = False

    def __init__(self, agents: dict):
        self.agents = agents
        self.particle_types = [type_ for type_ in self.agents]

    @property
    def kill_switch(self):
        return self._kill_switch

    @kill_switch.setter
    def kill_switch(self, value):
        self._kill_switch = value

    def calc_action(self, colloids) -> typing.List[Action]:
        actions = {int(np.copy(c.id)): Action() for c in colloids}
        switches = []
        for key, agent in self.agents.items():
            computed = agent.calc_action(colloids=colloids)
            switches.append(agent.kill_switch)
            count = 0
            for c in colloids:
                if str(c.type) == key:
                    actions[int(c.id)] = computed[count]
                    count += 1
        self.kill_switch = any(switches)
        return list(actions.values())

    def calc_reward(self, colloids, external_reward: float = 0.0) -> None:
        for agent in self.agents.values():
            agent.calc_reward(colloids=colloids, external_reward=external_reward)

    # ---- device path -----------------------------------------------------------------------
    @property
    def has_device_path(self) -> bool:
        return len(self.agents) > 0 and all(
            hasattr(a, "calc_action_device") and hasattr(a, "calc_reward_device") for a in self.agents.values()
        )

    def calc_action_device(self, swarm_view) -> None:
        """Each agent writes f_swim / torque of its own type straight into the device arrays."""
        switches = []
        for key, agent in self.agents.items():
            agent.calc_action_device(swarm_view, int(key))
            switches.append(agent.kill_switch)
        self.kill_switch = any(switches)

    def calc_reward_device(self, swarm_view, external_reward: float = 0.0) -> None:
        for key, agent in self.agents.items():
            agent.calc_reward_device(swarm_view, int(key), external_reward)
