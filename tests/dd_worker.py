"""
Worker of the multi-rank domain-decomposition tests: `python tests/dd_worker.py <rank> <world> <port> <out.npz> [backend]`.
Every rank builds the same seeded global swarm, keeps its slab, runs, and rank-local owned particles are
written to <out>.<rank>.npz.  With backend gloo all ranks may share cuda:0 (host-staged transport); with
nccl rank r uses cuda:r.
"""
import pathlib
import sys

import numpy as np

ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

CASE = dict(n=4096, L=160.0, dt=1e-3, seed=42, k=4, steps=61, slices=2)


def langevin_species():
    """mass / rinertia such that the velocity relaxation time is 50 time steps (tests/test_gpu_langevin.py)"""
    from helpers import stokes

    gt, gr = stokes(1.0)
    return [dict(type=0, gamma_t=(gt,) * 3, gamma_r=(gr,) * 3, mass=gt * CASE["dt"] / 0.02, rinertia=(gr * CASE["dt"] / 0.02,) * 3)]


def build_case():
    from helpers import jittered_lattice, species_for, stokes

    rng = np.random.default_rng(77)
    n, L = CASE["n"], CASE["L"]
    pos = jittered_lattice(n, L, 2, rng, jitter=0.1)
    pos[:, 0] += rng.integers(-1, 2, n) * L          # unfolded inputs: some particles start in other images
    theta = rng.uniform(0, 2 * np.pi, n)
    gt, gr = stokes(1.0)
    acts = [(rng.uniform(0, 40.0 * gt, n), rng.uniform(-1, 1, n) * gr * 2.0) for _ in range(CASE["slices"])]
    return pos, theta, acts, species_for([1.0])


def main():
    rank, world, port, out = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
    backend = sys.argv[5] if len(sys.argv) > 5 else "gloo"
    kT = float(sys.argv[6]) if len(sys.argv) > 6 else 300.0 / 293.0
    langevin = len(sys.argv) > 7 and sys.argv[7] == "langevin"
    import torch
    import torch.distributed as dist
    from helpers import wca_for
    from swarmrl_b200.domain import DomainSwarm

    dev = rank if backend == "nccl" else 0
    torch.cuda.set_device(dev)
    dist.init_process_group(backend, init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    pos, theta, acts, sp = build_case()
    L = CASE["L"]
    if langevin:
        sp = langevin_species()
    d = DomainSwarm((L, L), CASE["dt"], kT, CASE["seed"], sp, wca_for([1.0]), exchange_every=CASE["k"], world=world,
                    rank=rank, device=dev, skin=0.4, velocities=langevin)
    d.scatter(pos, theta)
    if langevin:
        d.sw.set_integrator("langevin")   # velocity Verlet: v, omega travel with the records, runs are cut with swarm_run_part
    d.field_reward((0.5, 0.5, 0.0), (L, L), init_only=True)
    step, means = 0, []
    for f, t in acts:
        d.set_actions_global(torch.from_numpy(f).to(d.dev), torch.from_numpy(t).to(d.dev))
        d.run(CASE["steps"], step)
        step += CASE["steps"]
        _, red = d.field_reward((0.5, 0.5, 0.0), (L, L), scale=10.0)
        means.append(red.cpu().numpy())
    st = d.stats()
    assert not st.overflow
    gid, x, y, th = d.owned_host()
    np.savez(f"{out}.{rank}.npz", gid=gid, x=x, y=y, theta=th, means=np.array(means), exchanges=d.exchanges,
             peer=int(d.peer is not None))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
