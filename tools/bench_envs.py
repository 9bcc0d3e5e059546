"""
BASELINE config 3: E independent environments x 32 agents, batched in one handle per GPU and sharded over
the GPUs with no communication (ensemble parallel).  Per action slice and environment: ConcentrationField
observable -> random policy from a counter RNG (no network) -> 4-action table -> 100 Brownian steps with
WCA -> GradientSensing reward -> mean.  Prints env-steps/s (one env-step = one environment advanced by one
action slice) and agent-steps/s, device-timed, max over ranks.

    python tools/bench_envs.py [--envs 4096]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29512 tools/bench_envs.py --envs 4096 [--weak]
"""
import argparse
import json
import os
import pathlib
import sys

import numpy as np

ROOT = pathlib.Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def measure(world, rank, local, envs, weak=False, m=32, steps=100, slices=50, warmup=5, skin=1.0):
    """Device-timed batched rollouts on every rank of the default process group (none when world == 1); returns
    the record on every rank (time = max over ranks).  `envs` is the total number of environments (per GPU when
    `weak`): rank r holds the contiguous range [r E/world, (r+1) E/world) and keys its noise by the GLOBAL agent
    index, so the ensemble does not depend on the sharding."""
    import torch
    import torch.distributed as dist

    import bench
    from swarmrl_b200.agents import CounterPolicy
    from swarmrl_b200.device import DeviceSwarm

    E_tot = envs * world if weak else envs
    E = E_tot // world
    e0 = rank * E
    L = 60.0
    n = E * m
    # every environment: 32 agents on a jittered 6 x 6 lattice of its own periodic 60 um box, seeded by
    # its GLOBAL environment index so the ensemble does not depend on the sharding
    pos = np.zeros((n, 3))
    th = np.zeros(n)
    side = int(np.ceil(np.sqrt(m)))
    a = L / side
    for e in range(E):
        rng = np.random.default_rng([1234, e0 + e])
        idx = rng.permutation(side * side)[:m]
        pos[e * m:(e + 1) * m, 0] = (idx % side + 0.5) * a + rng.uniform(-0.3 * a, 0.3 * a, m)
        pos[e * m:(e + 1) * m, 1] = (idx // side + 0.5) * a + rng.uniform(-0.3 * a, 0.3 * a, m)
        th[e * m:(e + 1) * m] = rng.uniform(0, 2 * np.pi, m)
    d = DeviceSwarm(2, (L,) * 3, True, 1e-3, bench.KT, 42, bench.species_table(), bench.wca_table(), skin=skin,
                    n_envs=E, device=local)
    d.set_state(pos, np.zeros(n, dtype=np.int32), theta=th)
    d.bind_ids(torch.arange(e0 * m, e0 * m + n, dtype=torch.int64, device=d.dev))
    gt, gr = bench.stokes(1.0)
    f_tab = torch.tensor([10.0 * gt, 0.0, 0.0, 0.0], dtype=torch.float64, device=d.dev)       # test_rl_trainers.py:36-46
    t_tab = torch.tensor([0.0, 10.0 * gr, -10.0 * gr, 0.0], dtype=torch.float64, device=d.dev)
    policy = CounterPolicy(4, seed=7, offset=e0 * m)
    row_of, rows = d.row_map(0)
    hist_obs = (d.pos[:, :n] / L).contiguous()
    hist_task = hist_obs.clone()
    src, box = (0.5, 0.5, 0.0), (L, L, L)

    def one_slice(step0):
        obs, _, _ = d.obs_field(row_of, rows, src, box, hist_obs, decay=1, c0=1.0, c1=-1.0, scale=10000.0, clip=False,
                                want_dist=False)
        idx = policy(obs)
        d.f_swim[:n] = f_tab[idx]
        d.torque[2, :n] = t_tab[idx]
        d.run(steps, step0)
        rew, _, _ = d.obs_field(row_of, rows, src, box, hist_task, decay=1, c0=1.0, c1=-1.0, scale=10.0, clip=True,
                                want_dist=False)
        return d.reduce_mean(rew)

    step = 0
    for _ in range(warmup):
        one_slice(step)
        step += steps
    st0 = d.stats()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ev0.record()
    for _ in range(slices):
        red = one_slice(step)
        step += steps
    ev1.record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    st = d.stats()
    assert not st.overflow and st.steps_done == steps
    ms = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=d.dev)
    tot = red[:1].clone()
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot)
    sec = float(ms[0]) * 1e-3
    out = {
        "metric": "env_steps_per_sec_batched_rollouts", "value": E_tot * slices / sec, "unit": "env-steps/s",
        "agent_steps_per_s": E_tot * m * steps * slices / sec, "n_gpus": world, "envs": E_tot, "envs_per_gpu": E, "agents_per_env": m,
        "steps_per_slice": steps, "slices": slices, "ms_per_slice": float(ms[0]) / slices,
        "scaling": "weak" if weak else "strong", "rebuilds_per_slice": (st.rebuilds - st0.rebuilds) / slices,
        "launches_per_slice": (st.launches - st0.launches) / slices, "persistent": bool(st.persistent),
        "reward_mean_last": float(tot[0]) / (E_tot * m), "dtype": "f64"}
    d.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", type=int, default=4096, help="environments in total (per GPU with --weak)")
    ap.add_argument("--agents-per-env", dest="m", type=int, default=32)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--slices", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--weak", action="store_true")
    ap.add_argument("--skin", type=float, default=1.0)
    args = ap.parse_args()

    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    out = measure(world, rank, local, args.envs, weak=args.weak, m=args.m, steps=args.steps, slices=args.slices,
                  warmup=args.warmup, skin=args.skin)
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
