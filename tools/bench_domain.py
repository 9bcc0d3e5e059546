"""
BASELINE config 5: one periodic 2-D swarm, slab-decomposed over the GPUs of the node.

    python tools/bench_domain.py --agents 1000000                      # 1 GPU (self-exchange)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29511 tools/bench_domain.py --agents 8000000     # the 8M-particle swarm

Prints one JSON line (rank 0): agent-steps/s over the whole swarm, device-timed, max over ranks, with
the exchange count and the share of the time spent in exchanges.  --strong keeps n fixed as ranks grow
(default), --per-gpu N sets n = N * world (weak).
"""
import argparse
import json
import os
import pathlib
import sys

import numpy as np

ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--agents", dest="n", type=int, default=1_000_000)
    ap.add_argument("--per-gpu", type=int, default=0)
    ap.add_argument("--phi", type=float, default=0.10)
    ap.add_argument("--steps", type=int, default=100, help="integrator steps per action slice")
    ap.add_argument("--slices", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--exchange-every", dest="k", type=int, default=10, help="steps between halo exchanges")
    ap.add_argument("--skin", type=float, default=1.0)
    ap.add_argument("--margin", type=float, default=0.25, help="halo = k * r_cut * (1 + margin)")
    ap.add_argument("--capacity-factor", type=float, default=1.3, help="local array capacity / expected local particles")
    ap.add_argument("--kT", type=float, default=300.0 / 293.0)
    args = ap.parse_args()

    import torch
    import torch.distributed as dist

    from swarmrl_b200.domain import DomainSwarm

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n = args.per_gpu * world if args.per_gpu else args.n
    L = float(np.sqrt(n * np.pi / args.phi))
    eta = 0.2200
    gt, gr = 6 * np.pi * eta, 8 * np.pi * eta
    m = 1000.0 / 4.0453e-9 * 1e-18 * 4.0 / 3.0 * np.pi
    species = [dict(type=0, gamma_t=(gt,) * 3, gamma_r=(gr,) * 3, mass=m, rinertia=(0.4 * m,) * 3)]
    s = 2.0 * 2 ** (-1 / 6)
    wca = dict(sigma=np.array([[s]]), eps=np.array([[297.0 / 293.0]]), cut=np.array([[2.0]]))

    # same seeded global swarm on every rank (jittered lattice, no overlaps)
    rng = np.random.default_rng(42)
    side = int(np.ceil(np.sqrt(n)))
    a = L / side
    idx = np.arange(n)
    pos = np.zeros((n, 2))
    pos[:, 0] = (idx % side + 0.5) * a
    pos[:, 1] = (idx // side + 0.5) * a
    pos += rng.uniform(-0.25 * (a - 2.0), 0.25 * (a - 2.0), (n, 2))
    theta = rng.uniform(0, 2 * np.pi, n)

    d = DomainSwarm((L, L), 1e-3, args.kT, 42, species, wca, exchange_every=args.k, margin=args.margin, skin=args.skin, world=world, rank=rank,
                    device=local, capacity_factor=args.capacity_factor)
    d.scatter(pos, theta)
    del pos
    f_glob = torch.from_numpy(rng.uniform(0.0, 10.0 * gt, n)).to(d.dev)
    t_glob = torch.from_numpy(rng.uniform(-1.0, 1.0, n) * gr * 4 * np.pi / 180).to(d.dev)
    d.field_reward((0.5, 0.5, 0.0), (L, L), init_only=True)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def one_slice(step0, timed):
        d.set_actions_global(f_glob, t_glob)
        d.run(args.steps, step0)
        _, red = d.field_reward((0.5, 0.5, 0.0), (L, L), scale=10.0)
        return red

    step = 0
    for _ in range(args.warmup):
        one_slice(step, False)
        step += args.steps
    st0 = d.stats()
    ex0 = d.exchanges
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for _ in range(args.slices):
        red = one_slice(step, True)
        step += args.steps
    ev1.record()
    barrier()
    st = d.stats()
    assert not st.overflow, "Verlet overflow"
    ms = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=d.dev)
    cnt = torch.tensor([float(d.n), float(d.n_owned)], dtype=torch.float64, device=d.dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt)
    if rank == 0:
        total = n * args.steps * args.slices
        print(json.dumps({
            "metric": "agent_steps_per_sec_domain_decomposed", "value": total / (float(ms[0]) * 1e-3), "unit": "agent-steps/s",
            "n_gpus": world, "n_agents": n, "box": L, "steps_per_slice": args.steps, "slices": args.slices,
            "ms_per_slice": float(ms[0]) / args.slices, "exchange_every": args.k, "halo": d.plan["halo"],
            "exchanges": d.exchanges - ex0, "rebuilds": int(st.rebuilds - st0.rebuilds), "persistent": bool(st.persistent),
            "local_particles_sum": float(cnt[0]), "owned_sum": float(cnt[1]), "redundancy": float(cnt[0]) / n,
            "reward_mean": float(red[1]), "skin": args.skin, "dtype": "f64"}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
