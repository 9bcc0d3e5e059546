"""
BASELINE config 5: one periodic 2-D swarm, slab-decomposed over the GPUs of the node.

    python tools/bench_domain.py --agents 1000000                      # 1 GPU (self-exchange)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29511 tools/bench_domain.py --agents 8000000     # the 8M-particle swarm

Prints one JSON line (rank 0): agent-steps/s over the whole swarm, device-timed, max over ranks, with
the exchange count and the share of the time spent in exchanges.  --strong keeps n fixed as ranks grow
(default), --per-gpu N sets n = N * world (weak).

`measure()` and `parity_selfcheck()` are also what `bench.py` calls for its `c5_domain` record, so the
driver's `bench.py --gpus N` line carries the same numbers.
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

ETA = 0.2200
SRC = (0.5, 0.5, 0.0)


def tables():
    gt, gr = 6 * np.pi * ETA, 8 * np.pi * ETA
    m = 1000.0 / 4.0453e-9 * 1e-18 * 4.0 / 3.0 * np.pi
    species = [dict(type=0, gamma_t=(gt,) * 3, gamma_r=(gr,) * 3, mass=m, rinertia=(0.4 * m,) * 3)]
    s = 2.0 * 2 ** (-1 / 6)
    wca = dict(sigma=np.array([[s]]), eps=np.array([[297.0 / 293.0]]), cut=np.array([[2.0]]))
    return species, wca, gt, gr


def build_swarm(n, phi, seed=42):
    """The same seeded global swarm on every rank: jittered square lattice (no overlaps), random directors, and
    one fixed set of per-particle actions.  Returns (pos (n,2), theta, f_swim, tq_z, L)."""
    _, _, gt, gr = tables()
    L = float(np.sqrt(n * np.pi / phi))
    rng = np.random.default_rng(seed)
    side = int(np.ceil(np.sqrt(n)))
    a = L / side
    idx = np.arange(n)
    pos = np.zeros((n, 2))
    pos[:, 0] = (idx % side + 0.5) * a
    pos[:, 1] = (idx // side + 0.5) * a
    pos += rng.uniform(-0.25 * (a - 2.0), 0.25 * (a - 2.0), (n, 2))
    theta = rng.uniform(0, 2 * np.pi, n)
    f = rng.uniform(0.0, 10.0 * gt, n)
    t = rng.uniform(-1.0, 1.0, n) * gr * 4 * np.pi / 180
    return pos, theta, f, t, L


def _make(world, rank, local, group, swarm, k, skin, margin, capf, kT, profile=False):
    import torch

    from swarmrl_b200.domain import DomainSwarm

    pos, theta, f, t, L = swarm
    species, wca, _, _ = tables()
    d = DomainSwarm((L, L), 1e-3, kT, 42, species, wca, exchange_every=k, margin=margin, skin=skin, world=world, rank=rank,
                    group=group, device=local, capacity_factor=capf)
    d.profile = profile
    d.scatter(pos, theta)
    f_glob = torch.from_numpy(f).to(d.dev)
    t_glob = torch.from_numpy(t).to(d.dev)
    d.field_reward(SRC, (L, L), init_only=True)
    return d, f_glob, t_glob


def _slice(d, f_glob, t_glob, L, steps, step0):
    d.set_actions_global(f_glob, t_glob)
    d.run(steps, step0)
    d.field_reward(SRC, (L, L), scale=10.0, defer=True)   # the reward MEAN is taken once per measurement (one all-reduce)


def measure(world, rank, local, swarm, *, steps=100, slices=5, warmup=2, k=33, skin=2.0, margin=0.15, capf=1.3,
            kT=300.0 / 293.0, group=None, keep=False):
    """Device-timed slices of the decomposed swarm on the ranks of `group` (None: the default group when
    world > 1, a lone self-exchanging rank when world == 1).  Returns a dict on every rank (values are the
    max / sum over the ranks); with keep=True also the DomainSwarm, left at step (warmup + slices) * steps."""
    import torch
    import torch.distributed as dist

    n, L = swarm[0].shape[0], swarm[4]
    d, f_glob, t_glob = _make(world, rank, local, group, swarm, k, skin, margin, capf, kT, profile=True)

    def barrier():
        if world > 1:
            dist.barrier(group=group)
        torch.cuda.synchronize()

    step = 0
    for _ in range(warmup):
        _slice(d, f_glob, t_glob, L, steps, step)
        step += steps
    st0 = d.stats()
    ex0 = d.exchanges
    d.timeline_reset()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    d.reward_means()
    for _ in range(slices):
        _slice(d, f_glob, t_glob, L, steps, step)
        step += steps
    means = d.reward_means()    # inside the timed region: the collective of the measurement
    red = torch.stack([means[-1] * float(d.n_global), means[-1]])
    ev1.record()
    barrier()
    st = d.stats()   # raises on a Verlet or exchange-buffer overflow
    tl = d.timeline()
    ms = torch.tensor([ev0.elapsed_time(ev1), tl["run_ms"], tl["exchange_ms"], tl["obs_ms"]], dtype=torch.float64, device=d.dev)
    cnt = torch.tensor([float(d.n), float(d.n_owned), float(st.rebuilds - st0.rebuilds)], dtype=torch.float64, device=d.dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX, group=group)
        dist.all_reduce(cnt, group=group)
    total = n * steps * slices
    bytes_per_exchange = 2 * int(d.bufs["send_l"].numel()) * 8   # both faces, buffer capacity (records are packed at the front)
    out = {
        "metric": "agent_steps_per_sec_domain_decomposed", "value": total / (float(ms[0]) * 1e-3), "unit": "agent-steps/s",
        "n_gpus": world, "n_agents": n, "box": L, "steps_per_slice": steps, "slices": slices,
        "ms_per_slice": float(ms[0]) / slices, "exchange_every": k, "halo": d.plan["halo"], "margin": margin,
        "exchanges": d.exchanges - ex0, "rebuilds": float(cnt[2]) / world, "persistent": bool(st.persistent),
        "local_particles_sum": float(cnt[0]), "owned_sum": float(cnt[1]), "redundancy": float(cnt[0]) / n,
        "bytes_per_exchange_capacity": bytes_per_exchange,
        "bytes_per_exchange_live": int(2 * (8 + d.n_aux) * 8 * d.plan["halo"] / d.plan["box_x"] * n),
        "transport": ("peer memory: P2P stores into the neighbours' CUDA-IPC regions + system-scope flag (swarm_dd_push / swarm_dd_wait)"
                      if d.peer is not None else ("nccl send/recv (torch.distributed)" if world > 1 else "self-exchange (device copy)")),
        "timeline_ms_per_slice_max_over_ranks": {"integrate": float(ms[1]) / slices, "exchange(pack+transport+merge)": float(ms[2]) / slices,
                                                 "reward": float(ms[3]) / slices},
        "reward_mean": float(red[1]), "skin": skin, "dtype": "f64", "scaling": "strong"}
    if keep:
        return out, d
    d.close()
    return out


def gather_owned(d, world, rank, group=None):
    """(x_global, y, theta) of ALL particles indexed by global id, on rank 0 (None elsewhere)."""
    import torch
    import torch.distributed as dist

    gid, x, y, th = d.owned_host()
    n = d.n_global
    if world == 1:
        full = np.zeros((3, n))
        full[:, gid] = np.stack([x, y, th])
        return full, len(gid)
    dev = d.dev
    cnt = torch.tensor([len(gid)], dtype=torch.int64, device=dev)
    cnts = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(cnts, cnt, group=group)
    cmax = int(max(int(c) for c in cnts))
    pack = torch.zeros((4, cmax), dtype=torch.float64, device=dev)
    pack[0, : len(gid)] = torch.from_numpy(gid.astype(np.float64)).to(dev)
    pack[1, : len(gid)] = torch.from_numpy(x).to(dev)
    pack[2, : len(gid)] = torch.from_numpy(y).to(dev)
    pack[3, : len(gid)] = torch.from_numpy(th).to(dev)
    parts = [torch.zeros_like(pack) for _ in range(world)] if rank == 0 else None
    dist.gather(pack, parts, dst=0, group=group)
    if rank != 0:
        return None, 0
    full = np.full((3, n), np.nan)
    owned_total = 0
    for c, p in zip(cnts, parts):
        c = int(c)
        q = p[:, :c].cpu().numpy()
        full[:, q[0].astype(np.int64)] = q[1:]
        owned_total += c
    return full, owned_total


def rel_err(a, b):
    scale = max(float(np.max(np.abs(b))), 1e-300)
    return float(np.max(np.abs(a - b)) / scale)


def parity_selfcheck(world, rank, local, *, n=20000, steps=61, slices=2, k=4, skin=0.4, kT=300.0 / 293.0, group=None):
    """The decomposed swarm over the live process group (NCCL when launched by torchrun) against the single-GPU
    engine (DeviceSwarm, no decomposition) on the same seeded swarm, noise ON: max relative error of x, y, theta
    over all particles, and whether every particle is owned exactly once.  Rank 0 returns the record."""
    import torch

    from swarmrl_b200.device import DeviceSwarm

    swarm = build_swarm(n, 0.10, seed=77)
    pos, theta, f, t, L = swarm
    pos = pos.copy()
    pos[:, 0] += np.random.default_rng(5).integers(-1, 2, n) * L   # unfolded inputs: some particles start in other images
    swarm = (pos, theta, f, t, L)
    d, f_glob, t_glob = _make(world, rank, local, group, swarm, k, skin, 0.25, 1.5, kT)
    step = 0
    means = []
    for _ in range(slices):
        _slice(d, f_glob, t_glob, L, steps, step)
        step += steps
    means = [float(v) for v in d.reward_means().cpu()]
    d.stats()
    full, owned_total = gather_owned(d, world, rank, group)
    ex = d.exchanges
    transport = "peer memory (swarm_dd_push / swarm_dd_wait)" if d.peer is not None else ("nccl" if world > 1 else "self-exchange")
    d.close()
    if rank != 0:
        return None
    species, wca, _, _ = tables()
    s = DeviceSwarm(2, (L, L, L), True, 1e-3, kT, 42, species, wca, skin=skin, device=local)
    p3 = np.zeros((n, 3)); p3[:, :2] = pos
    s.set_state(p3, np.zeros(n, dtype=np.int32), theta=theta)
    row_of, rows = s.row_map(0)
    hist = (s.pos[:, :n] / L).contiguous()
    tq = np.zeros((n, 3)); tq[:, 2] = t
    s.set_actions_host(f, tq)
    step, means_ref = 0, []
    for _ in range(slices):
        st = s.run_checked(steps, step)
        step += steps
        rew, _, _ = s.obs_field(row_of, rows, SRC, (L, L, L), hist, decay=1, c0=1.0, c1=-1.0, scale=10.0, clip=True, want_dist=False)
        means_ref.append(float(s.reduce_mean(rew)[1]))
    p, _, _ = s.particle_data_host()
    th = s.theta[:n].cpu().numpy()
    s.close()
    err = max(rel_err(full[0], p[:, 0]), rel_err(full[1], p[:, 1]), rel_err(full[2], th))
    return {"n_agents": n, "steps": steps * slices, "exchange_every": k, "exchanges": ex, "noise": kT > 0,
            "transport": transport, "against": "single-GPU engine (DeviceSwarm, no decomposition)",
            "max_rel_err": err, "tolerance": 1e-9, "every_particle_owned_once": bool(owned_total == n and not np.isnan(full).any()),
            "reward_mean_abs_err": float(np.max(np.abs(np.array(means) - np.array(means_ref)))), "pass": bool(err < 1e-9 and owned_total == n)}


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
    ap.add_argument("--selfcheck", action="store_true", help="also run the 20k-particle parity self-check")
    args = ap.parse_args()

    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n = args.per_gpu * world if args.per_gpu else args.n
    out = measure(world, rank, local, build_swarm(n, args.phi), steps=args.steps, slices=args.slices, warmup=args.warmup, k=args.k,
                  skin=args.skin, margin=args.margin, capf=args.capacity_factor, kT=args.kT)
    if args.selfcheck:
        chk = parity_selfcheck(world, rank, local)
        if rank == 0:
            out["parity_selfcheck"] = chk
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
