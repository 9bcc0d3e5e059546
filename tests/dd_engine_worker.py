"""
Worker of tests/test_gpu_domain_engine.py: `python tests/dd_engine_worker.py <rank> <world> <port> <out> [backend]`.
Every rank builds the SAME SwarmEngine (same particles, same agent) with `domain=dict(world=..., rank=...)` and runs the
reference's rollout loop; world 0 means "no domain": the plain single-GPU engine, whose results are the expected ones.
"""
import pathlib
import sys

import numpy as np

ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

CASE = dict(n=4096, L=160.0, slices=3, steps_per_slice=12, k=5, vision_range=6.0, n_cones=5, relax_iters=1000)


def build_engine(tmp, domain, thermostat="brownian"):
    import swarmrl_b200 as sw
    from helpers import jittered_lattice
    from swarmrl_b200.units import UnitRegistry

    n, L = CASE["n"], CASE["L"]
    ureg = UnitRegistry()
    dt = ureg.Quantity(1e-3, "second")
    md = sw.MDParams(ureg=ureg, fluid_dyn_viscosity=ureg.Quantity(8.9e-4, "pascal * second"),
                     WCA_epsilon=ureg.Quantity(297.0, "kelvin") * ureg.boltzmann_constant,
                     temperature=ureg.Quantity(300.0, "kelvin"), box_length=ureg.Quantity(3 * [L], "micrometer"),
                     time_step=dt, time_slice=dt * CASE["steps_per_slice"], write_interval=dt * CASE["steps_per_slice"] * 2,
                     thermostat_type=thermostat)
    e = sw.SwarmEngine(md_params=md, n_dims=2, seed=11, out_folder=tmp, write_chunk_size=100, domain=domain)
    rng = np.random.default_rng(5)
    pos = jittered_lattice(n, L, 2, rng, jitter=0.3)
    th = rng.uniform(0, 2 * np.pi, n)
    dirs = np.stack([np.cos(th), np.sin(th), np.zeros(n)], axis=1)
    kw = {}
    if thermostat == "langevin":   # espresso.py:391-400: mass and rinertia are mandatory; velocity relaxation time = 50 steps
        eta = md.fluid_dyn_viscosity.m_as("sim_dyn_viscosity")
        gt, gr = 6 * np.pi * eta * 1.0, 8 * np.pi * eta * 1.0
        kw = dict(mass=ureg.Quantity(gt * 1e-3 / 0.02, "sim_mass"), rinertia=ureg.Quantity(gr * 1e-3 / 0.02, "sim_rinertia"))
    e.add_colloids_at(ureg.Quantity(pos, "micrometer"), dirs, radius_colloid=ureg.Quantity(1.0, "micrometer"), type_colloid=0, **kw)
    return e


def build_agent():
    import swarmrl_b200 as sw
    from swarmrl_b200 import decay
    from swarmrl_b200.observables import SubdividedVisionCones
    from swarmrl_b200.tasks import GradientSensing

    L, n = CASE["L"], CASE["n"]
    rng = np.random.default_rng(3)
    D = CASE["n_cones"]
    P = dict(w1=rng.normal(size=(D, 32)) * 0.8, b1=rng.normal(size=32) * 0.1, wa=rng.normal(size=(32, 4)) * 0.5,
             ba=np.zeros(4), wv=rng.normal(size=32) * 0.2, bv=np.zeros(1))
    actions = {"Translate": sw.Action(force=10.0), "RotateCW": sw.Action(torque=np.array([0.0, 0.0, 10.0])),
               "RotateCCW": sw.Action(torque=np.array([0.0, 0.0, -10.0])), "Nothing": sw.Action()}
    task = GradientSensing(source=np.array([0.5 * L, 0.5 * L, 0.0]), decay_function=decay.Linear(1.0, -1.0),
                           box_length=np.array([L, L, L]), reward_scale_factor=10, particle_type=0)
    obs = SubdividedVisionCones(CASE["vision_range"], np.pi / 2, CASE["n_cones"], np.ones(n), detected_types=[0], particle_type=0)
    return sw.DeviceAgent(0, sw.FusedActorCritic(P), task, obs, actions, reward_mean=True)


def rollout(tmp, domain, thermostat="brownian"):
    import swarmrl_b200 as sw

    e = build_engine(tmp, domain, thermostat)
    agent = build_agent()
    ff = sw.ForceFunction({"0": agent})
    agent.reset_agent(e.colloids)
    e.integrate(CASE["slices"], ff)
    e.finalize()
    data = e.get_particle_data()
    be = e._backend
    means = e.reward_means(agent).cpu().numpy()      # one collective on a decomposed swarm, a stack on one GPU
    extra = dict(exchanges=be.dd.exchanges, owned=be.dd.n_owned) if domain is not None else dict(exchanges=0, owned=CASE["n"])
    return dict(pos=data["Unwrapped_Positions"], dirs=data["Directors"], vel=data["Velocities"], means=means, step_idx=e.step_idx, **extra)


def main():
    rank, world, port, out = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
    backend = sys.argv[5] if len(sys.argv) > 5 else "gloo"
    thermostat = sys.argv[6] if len(sys.argv) > 6 else "brownian"
    import torch
    import torch.distributed as dist

    dev = rank if backend == "nccl" else 0
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group(backend, init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    domain = dict(world=world, rank=rank, exchange_every=CASE["k"], margin=0.25, halo_extra=CASE["vision_range"], device=dev)
    res = rollout(pathlib.Path(out).parent / f"traj{rank}", domain, thermostat)
    np.savez(f"{out}.{rank}.npz", **res)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
