"""
Agents -- the callers that sit between `ForceFunction` and the observable / task kernels.

Host mirror of the reference's agent classes (same names, arguments and behaviour):
  Agent                /root/reference/swarmrl/agents/agent.py:11-62
  ClassicalAgent       /root/reference/swarmrl/agents/classical_agent.py:15-61
  ConstForce, ConstTorque, ConstForceAndTorque, ToConstDirection
                       /root/reference/swarmrl/agents/dummy_models.py:9-38
  ActorCriticAgent     /root/reference/swarmrl/agents/actor_critic.py:20-216
  TrajectoryInformation /root/reference/swarmrl/utils/colloid_utils.py:15-26

One deliberate difference: the reference's `ClassicalAgent` inherits
`Agent.calc_reward`, which raises NotImplementedError, so every rule-based agent crashes inside
`integrate` at espresso.py:1307.  Here `ClassicalAgent.calc_reward` evaluates the task if one was
given and otherwise does nothing.

`DeviceAgent` is the zero-copy path (not in the reference): observable -> policy -> action table
-> swim force / torque, all on device tensors of a running SwarmEngine, no per-colloid objects.
"""
from __future__ import annotations

import dataclasses
import typing

import numpy as np

from .components import Action


@dataclasses.dataclass
class TrajectoryInformation:
    particle_type: int
    features: list = dataclasses.field(default_factory=list)
    actions: list = dataclasses.field(default_factory=list)
    log_probs: list = dataclasses.field(default_factory=list)
    rewards: list = dataclasses.field(default_factory=list)
    killed: bool = False


class Agent:
    _killed = False

    @property
    def kill_switch(self):
        return self._killed

    @kill_switch.setter
    def kill_switch(self, value):
        self._killed = value

    def calc_action(self, colloids) -> typing.List[Action]:
        raise NotImplementedError("Implemented in Child class.")

    def calc_reward(self, colloids, external_reward: float = 0.0) -> None:
        raise NotImplementedError("Implemented in Child class.")


class ClassicalAgent(Agent):
    def __init__(self, particle_type: int, actions: dict, task=None, observable=None):
        self.particle_type = particle_type
        self.task = task
        self.observable = observable
        self.actions = actions

    def calc_action(self, colloids) -> typing.List[Action]:
        raise NotImplementedError("Implement in subclass")

    def calc_reward(self, colloids, external_reward: float = 0.0):
        task = getattr(self, "task", None)
        if task is None:
            return None
        rewards = task(colloids) + external_reward
        self.kill_switch = task.kill_switch
        return rewards


class ConstForce(ClassicalAgent):
    def __init__(self, force: float):
        self.action = Action(force=force)

    def calc_action(self, colloids):
        return len(colloids) * [self.action]


class ConstTorque(ClassicalAgent):
    def __init__(self, torque: np.ndarray):
        self.action = Action(torque=torque)

    def calc_action(self, colloids):
        return len(colloids) * [self.action]


class ConstForceAndTorque(ClassicalAgent):
    def __init__(self, force: float, torque: np.ndarray):
        self.action = Action(force=force, torque=torque)

    def calc_action(self, colloids):
        return len(colloids) * [self.action]


class ToConstDirection(ClassicalAgent):
    def __init__(self, direction: np.ndarray):
        self.action = Action(new_direction=direction)

    def calc_action(self, colloids):
        return len(colloids) * [self.action]


class ActorCriticAgent(Agent):
    """`network` is any object with `compute_action(observables) -> (indices, log_probs)` (the
    reference's FlaxModel, flax_network.py:153-195, drops in); `loss`, when given, needs
    `compute_loss(network=, episode_data=)`."""

    def __init__(self, particle_type: int, network, task, observable, actions: dict, loss=None, train: bool = True,
                 intrinsic_reward=None):
        self.network = network
        self.particle_type = particle_type
        self.task = task
        self.observable = observable
        self.actions = actions
        self.train = train
        self.loss = loss
        self.intrinsic_reward = intrinsic_reward
        self.trajectory = TrajectoryInformation(particle_type=self.particle_type)

    def __name__(self) -> str:
        return "ActorCriticAgent"

    def update_agent(self) -> tuple:
        rewards = self.trajectory.rewards
        killed = self.trajectory.killed
        if self.loss is not None:
            self.loss.compute_loss(network=self.network, episode_data=self.trajectory)
        if self.intrinsic_reward:
            self.intrinsic_reward.update(self.trajectory)
        self.reset_trajectory()
        return rewards, killed

    def reset_agent(self, colloids):
        self.observable.initialize(colloids)
        self.task.initialize(colloids)

    def reset_trajectory(self):
        self.task.kill_switch = False
        self.trajectory = TrajectoryInformation(particle_type=self.particle_type)

    def calc_action(self, colloids) -> typing.List[Action]:
        state_description = self.observable.compute_observable(colloids)
        action_indices, log_probs = self.network.compute_action(observables=state_description)
        chosen_actions = np.take(list(self.actions.values()), action_indices, axis=-1)
        if self.train:
            self.trajectory.features.append(state_description)
            self.trajectory.actions.append(action_indices)
            self.trajectory.log_probs.append(log_probs)
            self.trajectory.killed = self.task.kill_switch
        return chosen_actions

    def calc_reward(self, colloids, external_reward: float = 0.0):
        rewards = self.task(colloids)
        if self.intrinsic_reward:
            rewards += self.intrinsic_reward.compute_reward(episode_data=self.trajectory)
        rewards += external_reward
        if self.train:
            self.trajectory.rewards.append(rewards)
        self.kill_switch = self.task.kill_switch
        return rewards


# -------------------------------------------------------------------------------------------------
# device path
# -------------------------------------------------------------------------------------------------
class DeviceAgent(Agent):
    """
    Same roles as ActorCriticAgent, but every array stays in HBM:

      observable.compute_observable_device(view) -> (n_type, ...) fp64 device tensor
      policy(obs) -> int64 device tensor of action indices (n_type,)  [+ optional log-probs]
      the action table (dict name -> Action, as in the reference) is flattened once into device
      tensors force[k], torque[k,3]; the chosen rows are scattered into the engine's f_swim / torque
      arrays of this agent's particle type.

    `policy` is a callable on torch tensors (a torch module, a DLPack-bridged JAX function, or one
    of the built-ins below).  `new_direction` actions are not supported on this path.
    """

    def __init__(self, particle_type: int, policy, task, observable, actions: dict, train: bool = True,
                 keep_trajectory: bool = True, reward_mean: bool = False):
        self.particle_type = particle_type
        self.policy = policy
        self.task = task
        self.observable = observable
        self.actions = actions
        self.train = train
        self.keep_trajectory = keep_trajectory
        self.trajectory = TrajectoryInformation(particle_type=particle_type)
        self.last_rewards = None
        self.last_observable = None
        # reward_mean=True: (sum, mean) of every reward evaluation stays on the device (trainer.py:92-95 takes the mean
        # on the host); `reward_means` is the list of those 2-element device tensors, one D2H for a whole rollout
        self.reward_mean = reward_mean
        self.reward_means = []
        self._table = None
        self._idx = None

    def reset_agent(self, colloids):
        self.observable.initialize(colloids)
        self.task.initialize(colloids)

    def reset_trajectory(self):
        self.task.kill_switch = False
        self.trajectory = TrajectoryInformation(particle_type=self.particle_type)

    def _tables(self, swarm):
        if self._table is None or self._table[0].device != swarm.dev:
            torch = swarm.torch
            acts = list(self.actions.values())
            if any(a.new_direction is not None for a in acts):
                raise NotImplementedError("new_direction actions need the object path")
            f = torch.tensor([float(a.force) for a in acts], dtype=torch.float64, device=swarm.dev)
            t = torch.tensor([[0.0, 0.0, 0.0] if a.torque is None else [float(v) for v in a.torque] for a in acts],
                             dtype=torch.float64, device=swarm.dev)
            self._table = (f, t)
            sel = np.nonzero(swarm.ptype_host == int(self.particle_type))[0]
            self._idx = torch.from_numpy(sel).to(swarm.dev)
            self._all = len(sel) == swarm.n
        return self._table

    def calc_action_device(self, view, particle_type=None):
        swarm = view.swarm
        f_tab, t_tab = self._tables(swarm)
        obs = self.observable.compute_observable_device(view)
        self.last_observable = obs
        if hasattr(self.policy, "act"):
            # fused path (FusedActorCritic): network, sampling and the action write in one kernel
            row_of, rows = swarm.row_map(self.particle_type)
            idx, logp = self.policy.act(swarm, obs.reshape(rows, -1), row_of, rows, f_tab, t_tab.reshape(-1))
            if self.train and self.keep_trajectory:
                self.trajectory.features.append(obs)
                self.trajectory.actions.append(idx)
                self.trajectory.log_probs.append(logp)
                self.trajectory.killed = self.task.kill_switch
            return
        if getattr(swarm, "is_domain", False):
            raise NotImplementedError("a domain-decomposed swarm needs a policy with a fused `act` (FusedActorCritic): its "
                                      "sampling is keyed by the particle id, so a ghost takes its owner's action")
        out = self.policy(obs)
        idx, logp = out if isinstance(out, tuple) else (out, None)
        if self._all:
            swarm.f_swim[: swarm.n] = f_tab[idx]
            swarm.torque[:, : swarm.n] = t_tab[idx].T
        else:
            swarm.f_swim[self._idx] = f_tab[idx]
            swarm.torque[:, self._idx] = t_tab[idx].T
        if self.train and self.keep_trajectory:
            self.trajectory.features.append(obs)
            self.trajectory.actions.append(idx)
            self.trajectory.log_probs.append(logp)
            self.trajectory.killed = self.task.kill_switch

    def calc_reward_device(self, view, particle_type=None, external_reward: float = 0.0):
        fused = self.reward_mean and not external_reward and getattr(self.task, "fused_mean", False)
        rewards = self.task.call_device(view, want_mean=True) if fused else self.task.call_device(view)
        if external_reward:
            rewards = rewards + external_reward
        self.last_rewards = rewards
        if self.reward_mean:
            red = self.task.last_mean if fused else None
            self.reward_means.append(red if red is not None else view.swarm.reduce_mean(rewards))
        if self.train and self.keep_trajectory:
            self.trajectory.rewards.append(rewards)
        self.kill_switch = self.task.kill_switch

    # the object protocol still works (slow path), so a DeviceAgent can sit in any ForceFunction
    def calc_action(self, colloids):
        obs = self.observable.compute_observable_device(colloids)
        out = self.policy(obs)
        idx = (out[0] if isinstance(out, tuple) else out).cpu().numpy()
        return np.take(list(self.actions.values()), idx, axis=-1)

    def calc_reward(self, colloids, external_reward: float = 0.0):
        rewards = self.task(colloids) + external_reward
        if self.train and self.keep_trajectory:
            self.trajectory.rewards.append(rewards)
        self.kill_switch = self.task.kill_switch
        return rewards


class CounterPolicy:
    """Random policy from a counter-based generator ('random-policy
    actions from a counter RNG so no network is needed'): action = hash(seed, call, agent) mod k.
    Pure integer torch ops on the device; independent of how agents are sharded over GPUs when
    `offset` is the global index of the first local agent."""

    def __init__(self, n_actions: int, seed: int = 7, offset: int = 0):
        self.k, self.seed, self.offset, self.call = int(n_actions), int(seed), int(offset), 0

    def __call__(self, obs):
        import torch

        n = obs.shape[0]
        i = torch.arange(self.offset, self.offset + n, dtype=torch.int64, device=obs.device)
        x = i * 0x9E3779B1 + (self.call * 0x85EBCA77 + self.seed * 0xC2B2AE3D)
        x = (x ^ (x >> 15)) * 0x2C1B3C6D & 0xFFFFFFFF
        x = (x ^ (x >> 12)) * 0x297A2D39 & 0xFFFFFFFF
        x = x ^ (x >> 15)
        self.call += 1
        return x % self.k


class FusedActorCritic:
    """Device policy for `DeviceAgent`: the network of the reference's tests (Dense(H)+ReLU trunk, actor and critic
    heads; test_rl_trainers.py:17-26) evaluated, sampled (Gumbel-max) and -- when the agent's action table is handed
    over -- written into the engine's action arrays by ONE kernel (swarm_policy_forward), so a slice needs no host
    round trip.  `params` mirrors a flax parameter tree: {"Dense_0": {"kernel", "bias"}, "Dense_1": critic,
    "Dense_2": actor} (creation order of the reference's module) or the flat keys w1, b1, wa, ba, wv, bv."""

    def __init__(self, params, swarm=None, sample=True):
        self.host = self._flatten(params)
        self.sample = sample
        self.call = 0
        self._dev = None
        self.last_value = None
        if swarm is not None:
            self._upload(swarm)

    @staticmethod
    def _flatten(p):
        if "w1" in p:
            src = p
        else:
            src = {"w1": p["Dense_0"]["kernel"], "b1": p["Dense_0"]["bias"], "wv": p["Dense_1"]["kernel"], "bv": p["Dense_1"]["bias"],
                   "wa": p["Dense_2"]["kernel"], "ba": p["Dense_2"]["bias"]}
        out = {k: np.ascontiguousarray(np.asarray(v, dtype=np.float32)) for k, v in src.items()}
        out["wv"] = out["wv"].reshape(-1)
        out["bv"] = out["bv"].reshape(1)
        return out

    def _upload(self, swarm):
        self._dev = {k: swarm.torch.from_numpy(v).to(swarm.dev) for k, v in self.host.items()}
        self._swarm = swarm

    def act(self, swarm, obs, row_of, n_rows, act_force=None, act_torque=None):
        if self._dev is None or self._swarm is not swarm:
            self._upload(swarm)
        idx, logp, val = swarm.policy_forward(self._dev, obs, row_of, n_rows, self.call, self.sample, act_force, act_torque)
        self.call += 1
        self.last_value = val
        return idx, logp
