#!/usr/bin/env python
"""
bench.py -- headline benchmark of the SwarmRL rollout hot path on B200 (BASELINE.json metric:
agent-steps/sec at 100k agents).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (BASELINE config 4, SURVEY.md 8d "C4"): one swarm of 100,000 active colloids, r = 1 um,
periodic 2-D box L = 1772 um (phi = 0.10), dt = 1e-3 s, kT = 300 K, 100 integrator steps per action
slice.  One bench "step" = one action slice of `SwarmEngine.integrate`:
    manage_forces (write f_swim / ext torque of every agent)          espresso.py:1203-1249
    integrator.run(100)  Brownian dynamics + WCA + Verlet lists + Philox   espresso.py:1304-1306
    SubdividedVisionCones observable (range 10, half angle pi/2, 5 cones)  subdivided_vision_cones.py:83-258
    GradientSensing reward + its mean                                      gradient_sensing.py:92-150, trainer.py:92-95
value  = N * 100 * K / (device time of K slices), inputs resident in HBM.
e2e    = the same slice through the host-buffer C-ABI (swarm_set_actions_host_2d -> run -> observable ->
         D2H of the observable and the reward mean), actions coming from pinned host memory.
N > 1  : ensemble sharding -- every rank advances its own independent 100k-agent swarm (weak scaling,
         no data-path collective; torch.distributed only for the barrier and the max-over-ranks time).
Two more records ride on the same line, measured in the same invocation with the same timing rules, because
north_star names them and they are what exercises the N GPUs TOGETHER:
  c5_domain : BASELINE config 5, ONE 8,000,000-particle swarm slab-decomposed over the N GPUs (NCCL halo +
              migration exchange every k steps, swarmrl_b200/domain.py) -- strong scaling; rank 0 also times
              the same swarm on its GPU alone, so `efficiency_vs_n1` is measured inside one run on one box;
              plus `parity_selfcheck` (a 20k-particle swarm over the live NCCL group against the single-GPU
              engine, noise on) and, for N > 1, `parity_full` (all 8M particles against the one-GPU run).
  c3_envs   : BASELINE config 3, 4,096 environments x 32 agents sharded E/N per GPU (strong, as BASELINE.json
              states it) and 4,096 per GPU (weak); env-steps/s.
--impl reference : ESPResSo is not installable here, so the reference arm times the CPU oracle
         (oracle/swarm_oracle.c, OpenMP, all host threads) on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import pathlib
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = pathlib.Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

N_AGENTS = 100_000
BOX = 1772.0
DT = 1e-3
STEPS_PER_SLICE = 100
KT = 300.0 / 293.0
SEED = 42
ETA = 0.2200
VISION = dict(range=10.0, half_angle=np.pi / 2, n_cones=5)
BYTES_PER_AGENT_STEP = 68  # 2-D steady-state step, read 44 B + write 24 B
# dram__bytes_read.sum + dram__bytes_write.sum of one k_bd2d_resident launch (100 steps), from the ncu --set full
# capture profiles/r01_k_bd2d_resident.ncu-rep (10.70 MB + 0.12 MB): the state is register / L2 resident
ISSUE_SLOTS_PER_CLOCK = 148 * 4      # warp instructions a B200 can issue per SM clock (one per scheduler)


def ncu_profile(kernel):
    """Per-launch figures of the run kernel from the committed ncu --set full summary (tools/ncu_summary.py), scaled to
    the bench's 100-step launch: DRAM bytes (dram__bytes_read + write) and warp instructions.  Returns None when no
    summary of that kernel is committed -- nothing is hard-coded here."""
    for rnd in ("r02", "r01"):
        f = ROOT / "profiles" / f"{rnd}_ncu_{kernel}_summary.txt"
        if not f.exists():
            continue
        vals = {}
        for ln in f.read_text().splitlines():
            p = ln.split()
            if len(p) >= 2 and p[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum", "profiled_launch_steps",
                                        "gpu__time_duration.sum"):
                mult = {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0}.get(p[2] if len(p) > 2 else "", 1.0)
                vals[p[0]] = float(p[1]) * mult
        steps = vals.get("profiled_launch_steps", 100.0)   # round 1's summaries profiled 100-step launches
        if "dram__bytes_read.sum" not in vals:
            continue
        k = STEPS_PER_SLICE / steps
        return {"file": f"profiles/{f.name}", "dram_bytes_per_launch": (vals["dram__bytes_read.sum"] + vals.get("dram__bytes_write.sum", 0.0)) * k,
                "warp_instr_per_launch": vals.get("smsp__inst_executed.sum", 0.0) * k}
    return None


METRIC = "agent_steps_per_sec_100k_agents"
UNIT = "agent-steps/s"


def stokes(r):
    return 6 * np.pi * ETA * r, 8 * np.pi * ETA * r**3


def species_table():
    gt, gr = stokes(1.0)
    m = 1000.0 / 4.0453e-9 * 1e-18 * 4.0 / 3.0 * np.pi
    return [dict(type=0, gamma_t=(gt,) * 3, gamma_r=(gr,) * 3, mass=m, rinertia=(0.4 * m,) * 3)]


def wca_table():
    s = 2.0 * 2 ** (-1 / 6)
    return dict(sigma=np.array([[s]]), eps=np.array([[297.0 / 293.0]]), cut=np.array([[2.0]]))


def initial_state(n, L, seed):
    """Jittered square lattice (no overlaps at phi = 0.10), directors uniform on the circle."""
    rng = np.random.default_rng(seed)
    m = int(np.ceil(np.sqrt(n)))
    a = L / m
    idx = np.arange(n)
    pos = np.zeros((n, 3))
    pos[:, 0] = (idx % m + 0.5) * a
    pos[:, 1] = (idx // m + 0.5) * a
    pos[:, :2] += rng.uniform(-0.25 * a, 0.25 * a, (n, 2))
    return pos, rng.uniform(0, 2 * np.pi, n)


def action_sets(n, seed, n_sets=4):
    """Random-policy actions from the 4 actions of the reference's RL tests (test_rl_trainers.py:36-46):
    translate (f = gamma_t * 5 um/s), rotate CCW / CW (tau_z = +-gamma_r * 4 pi/180 rad/s ... scaled), none."""
    gt, gr = stokes(1.0)
    rng = np.random.default_rng(seed)
    table_f = np.array([5.0 * gt, 0.0, 0.0, 0.0])
    table_t = np.array([0.0, gr * 10.0, -gr * 10.0, 0.0])
    out = []
    for _ in range(n_sets):
        a = rng.integers(0, 4, n)
        tq = np.zeros((n, 3))
        tq[:, 2] = table_t[a]
        out.append((np.ascontiguousarray(table_f[a]), tq))
    return out


class ClockSampler(threading.Thread):
    """nvidia-smi SM clock / throttle reasons sampled while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self._halt = index, [], set(), None, threading.Event()

    def _q(self, fields):
        try:
            r = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={fields}", "--format=csv,noheader,nounits"],
                               capture_output=True, text=True, timeout=5)
            return [v.strip() for v in r.stdout.strip().split(",")]
        except Exception:
            return None

    def _run_nvml(self):
        """NVML in-process: a sample every 2 ms, so even a 15 ms timed region is covered (nvidia-smi takes ~30 ms
        per query).  Returns False when NVML is unusable and the nvidia-smi loop should run instead."""
        try:
            import pynvml as nv

            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
        except Exception:
            return False
        bits = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40,
                "hw_power_brake_slowdown": 0x80}
        while not self._halt.is_set():
            try:
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                r = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(h))
                for nme, b in bits.items():
                    if r & b:
                        self.reasons.add(nme)
            except Exception:
                pass
            self._halt.wait(0.002)
        return True

    def run(self):
        if self._run_nvml():
            return
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap", "hw_power_brake_slowdown"]
        q = "clocks.sm,clocks.max.sm," + ",".join("clocks_throttle_reasons." + n for n in names)
        while not self._halt.is_set():
            v = self._q(q)
            if v and len(v) >= 2 + len(names):
                try:
                    self.samples.append(float(v[0]))
                    self.max_mhz = float(v[1])
                except ValueError:
                    pass
                for nme, val in zip(names, v[2:]):
                    if val.lower().startswith("active"):
                        self.reasons.add(nme)
            self._halt.wait(0.05)

    def stop(self):
        self._halt.set()
        self.join(timeout=6)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


# --------------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference path on the host cores (ESPResSo itself is unavailable)
# --------------------------------------------------------------------------------------------------
def cpu_slice_rate(sample_steps=40, observers=2000, threads=None):
    """Times a bounded sample of the same workload with the CPU oracle and returns the agent-steps/s of
    one full action slice: t_slice = 100 * t_step + t_vision(N observers) + t_reward, where t_step comes
    from `sample_steps` BD steps of all 100k agents (Verlet lists, skin 0.4, Philox noise) and t_vision
    from `observers` observers against all 100k particles (the reference's O(N^2) scan) scaled to N."""
    from oracle import oracle
    from oracle.oracle import OracleSwarm

    # all host threads unless told otherwise: torchrun exports OMP_NUM_THREADS=1 to its workers, which would
    # silently turn the "all cores" baseline into a single-threaded one
    if not threads:
        try:
            threads = len(os.sched_getaffinity(0))
        except AttributeError:
            threads = os.cpu_count() or 1
    oracle.set_num_threads(threads)
    cores = oracle.num_threads()
    n = N_AGENTS
    pos, th = initial_state(n, BOX, SEED)
    o = OracleSwarm(2, (BOX,) * 3, True, DT, KT, SEED, species_table(), wca_table())
    o.set_state(pos, np.zeros(n, dtype=np.int32), theta=th)
    f, tq = action_sets(n, 7, 1)[0]
    o.f_swim[:] = f
    o.torque[2][:] = tq[:, 2]
    o.bd_steps(0, 2, mode=2)  # warm
    t0 = time.perf_counter()
    o.bd_steps(2, sample_steps, mode=2)
    t_step = (time.perf_counter() - t0) / sample_steps
    # vision: only `observers` particles carry the observer type; everybody is seen
    o.ptype = np.ones(n, dtype=np.int32)
    o.ptype[:: n // observers] = 0
    rows = int((o.ptype == 0).sum())
    t0 = time.perf_counter()
    o.vision_cones(0, VISION["range"], VISION["half_angle"], VISION["n_cones"], np.ones(n), [0, 1])
    t_vis = (time.perf_counter() - t0) * n / rows
    o.ptype = np.zeros(n, dtype=np.int32)
    hist = np.ascontiguousarray((o.pos / BOX).T)
    t0 = time.perf_counter()
    dn, dp = o.field_distances(0, (0.5, 0.5, 0.0), (BOX,) * 3, hist)
    float(np.clip(10.0 * ((1 - dn) - (1 - dp)), 0, None).mean())
    t_rew = time.perf_counter() - t0
    t_slice = STEPS_PER_SLICE * t_step + t_vis + t_rew
    return dict(value=n * STEPS_PER_SLICE / t_slice, unit=UNIT, cores=cores, kind="port",
                sample=f"{sample_steps} BD steps of all {n} agents + vision cones for {rows} observers (scaled to {n}) "
                       f"+ 1 reward pass; t_step={t_step*1e3:.2f} ms, t_vision(N)={t_vis:.2f} s",
                t_step_ms=t_step * 1e3, t_vision_s=t_vis, integrator_only=n / t_step)


def config_dict(n_gpus):
    return {"workload": "C4: 100,000-agent single swarm, 2-D periodic L=1772 (phi=0.10), WCA + Brownian dynamics "
                        "kT=300K dt=1e-3, 100 steps/slice, vision cones (range 10, 5 cones) + gradient-sensing reward mean",
            "n_agents": N_AGENTS, "steps_per_slice": STEPS_PER_SLICE,
            "parallelism": "single GPU" if n_gpus == 1 else f"ensemble-sharded: {n_gpus} independent 100k swarms, no collective",
            "l2": "working set (6.8 MB) is L2-resident; L2 flushed (256 MiB write) between timed slices, outside the events"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    vals = []
    for _ in range(max(1, args.warmup // 3)):
        cpu_slice_rate(sample_steps=5, observers=500)
    t_all = time.perf_counter()
    last = None
    for _ in range(args.steps):
        last = cpu_slice_rate(sample_steps=20, observers=1000)
        vals.append(last["value"])
    v = float(np.mean(vals))
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": N_AGENTS * STEPS_PER_SLICE / v * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config_dict(args.gpus),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": last["cores"], "kind": "port", "sample": last["sample"]},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "ESPResSo (the reference's physics backend) is not installable offline; this arm is the CPU oracle "
                    "port of the same path, OpenMP over all host threads", "wall_s": time.perf_counter() - t_all}
    print(json.dumps(line))


# --------------------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------------------
class Rollout:
    """One 100k-agent swarm on one GPU and the per-slice work of the rollout hot path."""

    def __init__(self, device_index, seed, skin):
        import torch
        from swarmrl_b200.device import DeviceSwarm

        self.torch = torch
        n = N_AGENTS
        pos, th = initial_state(n, BOX, seed)
        self.d = d = DeviceSwarm(2, (BOX,) * 3, True, DT, KT, seed, species_table(), wca_table(), skin=skin,
                                 device=device_index)
        d.set_state(pos, np.zeros(n, dtype=np.int32), theta=th)
        self.row_of, self.rows = d.row_map(0)
        f64 = dict(dtype=torch.float64, device=d.dev)
        self.radii = torch.ones(n, **f64)
        self.type_col = torch.zeros(1, dtype=torch.int32, device=d.dev)
        self.hist = (d.pos[:, :n] / BOX).contiguous()
        sets = action_sets(n, 7 + seed)
        # device-resident action sets (value) and pinned host copies (e2e)
        self.act_dev = [(torch.from_numpy(f).to(d.dev), torch.from_numpy(np.ascontiguousarray(t.T)).to(d.dev)) for f, t in sets]
        # a 2-D swarm: the host passes (f_swim, tau_z) per agent -- swarm_set_actions_host_2d (the x, y torque components are zero)
        assert all(not t[:, :2].any() for _, t in sets)
        self.act_pin = [(torch.from_numpy(f).pin_memory(), torch.from_numpy(np.ascontiguousarray(t[:, 2])).pin_memory()) for f, t in sets]
        self.obs_pin = torch.empty((n, VISION["n_cones"], 1), dtype=torch.float64).pin_memory()
        self.mean_pin = torch.empty(2, dtype=torch.float64).pin_memory()
        self.step_idx = 0
        self.flush_buf = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=d.dev)
        self.run_ms = []
        self.h2d = sum(f.numel() * 8 + t.numel() * 8 for f, t in self.act_pin[:1])
        self.d2h = self.obs_pin.numel() * 8 + 16

    def flush_l2(self):
        self.flush_buf.fill_(1)

    def _physics_and_obs(self, time_run):
        d, torch = self.d, self.torch
        if time_run:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        d.run(STEPS_PER_SLICE, self.step_idx)
        if time_run:
            e1.record()
            self.run_events.append((e0, e1))
        self.step_idx += STEPS_PER_SLICE
        obs = d.obs_vision(self.row_of, self.rows, VISION["range"], VISION["half_angle"], VISION["n_cones"], self.radii,
                           self.type_col, 1)
        # GradientSensing reward and its mean: one launch (swarm_obs_field_mean)
        rew, mean = d.obs_field_mean(self.row_of, self.rows, (0.5, 0.5, 0.0), (BOX,) * 3, self.hist, 1, 1.0, -1.0, scale=10.0, clip=True)
        return obs, mean

    def slice_resident(self, k, time_run=False):
        """manage_forces from device-resident actions -> run -> observable -> reward mean (all in HBM)."""
        f, t = self.act_dev[k % len(self.act_dev)]
        self.d.f_swim[: N_AGENTS].copy_(f)
        self.d.torque[:, : N_AGENTS].copy_(t)
        return self._physics_and_obs(time_run)

    def slice_e2e(self, k):
        """Same slice through the host-buffer C-ABI: pinned host actions in, observable + reward mean out."""
        import ctypes as C
        from swarmrl_b200 import _lib

        f, t = self.act_pin[k % len(self.act_pin)]
        d = self.d
        _lib.check(d.lib.swarm_set_actions_host_2d(d.h, C.c_void_p(f.data_ptr()), C.c_void_p(t.data_ptr()), d._stream()))
        obs, mean = self._physics_and_obs(False)
        self.obs_pin.copy_(obs, non_blocking=True)
        self.mean_pin.copy_(mean, non_blocking=True)
        self.torch.cuda.current_stream().synchronize()
        return float(self.mean_pin[1])

    run_events = []


def run_ours(args):
    import torch
    import torch.distributed as dist

    from swarmrl_b200 import _lib

    _lib.load()  # raises if the CUDA library is missing: there is no fallback
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (B200); there is no CPU fallback")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n_gpus = world

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    ro = Rollout(local, SEED + rank, args.skin)
    K, W = args.steps, args.warmup
    for k in range(W):
        ro.slice_resident(k)
    st = ro.d.stats()
    if st.overflow or st.steps_done != STEPS_PER_SLICE:
        raise SystemExit(f"warm-up run halted: steps_done={st.steps_done} overflow={st.overflow}")
    launches0, rebuilds0 = st.launches, st.rebuilds

    # ---- device-resident arm ---------------------------------------------------------------------
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    ro.run_events = []
    barrier()
    t_wall0 = time.perf_counter()
    for k in range(K):
        ro.flush_l2()
        ev[k][0].record()
        ro.slice_resident(W + k, time_run=True)
        ev[k][1].record()
    barrier()
    t_wall = time.perf_counter() - t_wall0
    st = ro.d.stats()
    if st.overflow or st.steps_done != STEPS_PER_SLICE:
        raise SystemExit(f"timed run halted: steps_done={st.steps_done} overflow={st.overflow}")
    dev_ms = sum(a.elapsed_time(b) for a, b in ev)
    run_ms = [a.elapsed_time(b) for a, b in ro.run_events]
    launches = int(st.launches - launches0) + 2 * K  # + the two D2D action copies per slice (torch)
    rebuilds = int(st.rebuilds - rebuilds0)

    # ---- end-to-end arm (host buffers through the C-ABI) ---------------------------------------------
    for k in range(min(W, 3)):
        ro.slice_e2e(k)
    ev2 = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    barrier()
    for k in range(K):
        ro.flush_l2()
        ev2[k][0].record()
        ro.slice_e2e(k)
        ev2[k][1].record()
    barrier()
    e2e_ms = sum(a.elapsed_time(b) for a, b in ev2)
    clocks = sampler.stop() if rank == 0 else None

    t = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms = float(t[0]), float(t[1])
    tiling = ro.d.tiling_stats()
    eng = None
    if not args.no_engine:
        eng = engine_record(args, local, K, W)
        if world > 1:
            te = torch.tensor([eng["ms_per_step"]], dtype=torch.float64, device="cuda")
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
            eng["ms_per_step"] = float(te[0])
            eng["value"] = N_AGENTS * STEPS_PER_SLICE * n_gpus / (eng["ms_per_step"] * 1e-3)
    extras = None
    if not args.no_extra:
        del ro.flush_buf
        ro.d.close()
        torch.cuda.empty_cache()
        extras = extra_records(args, world, rank, local)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    units = N_AGENTS * STEPS_PER_SLICE * K * n_gpus
    value = units / (dev_ms * 1e-3)
    e2e = units / (e2e_ms * 1e-3)
    peaks = {}
    try:
        peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6554.6))
    run_avg_ms = float(np.mean(run_ms))
    achieved = N_AGENTS * STEPS_PER_SLICE * BYTES_PER_AGENT_STEP / (run_avg_ms * 1e-3) / 1e9
    kname = "k_bd2d_tiled" if tiling["tiled"] else "k_bd2d_resident"
    prof = ncu_profile(kname)
    sm_hz = (clocks["sm_mhz"] or 1965.0) * 1e6 if clocks else 1965.0e6
    issue = None
    if prof and prof["warp_instr_per_launch"]:
        issue = {"warp_instr_per_launch": prof["warp_instr_per_launch"],
                 "warp_instr_per_agent_step": prof["warp_instr_per_launch"] * 32 / (N_AGENTS * STEPS_PER_SLICE),
                 "peak_warp_instr_per_s": ISSUE_SLOTS_PER_CLOCK * sm_hz,
                 "frac": prof["warp_instr_per_launch"] / (run_avg_ms * 1e-3) / (ISSUE_SLOTS_PER_CLOCK * sm_hz),
                 "source": prof["file"] + " (instruction count) / this run's launch time and SM clock"}
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n_gpus, "steps": K, "warmup": W,
            "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": dict(config_dict(n_gpus), skin=args.skin, rebuilds_per_slice=rebuilds / K,
                                               persistent=bool(st.persistent)),
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": ro.h2d, "d2h_bytes_per_step": ro.d2h,
                    "ms_per_step": e2e_ms / K},
            "gpu_launches": launches,
            "roofline": {"bound": "hbm", "kernel": kname + "<true> (one cooperative launch = 100 Brownian steps of 100k agents)",
                         "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": prof["dram_bytes_per_launch"] if prof else None,
                         "traffic_source": prof["file"] if prof else None,
                         "issue": issue, "issue_frac": issue["frac"] if issue else None,
                         "tiling": tiling,
                         "bytes_per_launch": N_AGENTS * STEPS_PER_SLICE * BYTES_PER_AGENT_STEP,
                         "launch_ms": run_avg_ms, "share_of_step": run_avg_ms * K / dev_ms,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback"},
            "clocks": clocks, "wall_s_timed": t_wall}
    if eng is not None:
        line["e2e_engine"] = eng
    if n_gpus == 1 and not args.no_cpu:
        cb = cpu_slice_rate(sample_steps=args.cpu_steps, observers=2000)
        line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
        # the reference's CI runs ESPResSo single-threaded (OMP_NUM_THREADS=1, .github/workflows/espresso.yml:41)
        cb1 = cpu_slice_rate(sample_steps=max(4, args.cpu_steps // 10), observers=250, threads=1)
        line["cpu_baseline"]["single_thread_value"] = cb1["value"]
        from oracle import oracle as _orc
        _orc.set_num_threads(cb["cores"])
    line.update(extras or {})
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def engine_record(args, local, K, W):
    """`e2e_engine`: the same C4 slice through the repo's own Engine API, the call a SwarmRL user makes --
    SwarmEngine.integrate(K, ForceFunction({'0': DeviceAgent(FusedActorCritic, GradientSensing, SubdividedVisionCones)}))
    (espresso.py:1251-1308, actor_critic.py:159-216) -- CUDA events around the WHOLE integrate() call, so the Python
    host loop counts; the only D2H is the K reward means.  Policy: the reference tests' network (Dense(128) + ReLU,
    actor 4 / critic 1, test_rl_trainers.py:17-26), random-init, evaluated and sampled on the device."""
    import tempfile

    import torch

    import swarmrl_b200 as sw
    from swarmrl_b200 import decay
    from swarmrl_b200.observables import SubdividedVisionCones
    from swarmrl_b200.tasks import GradientSensing
    from swarmrl_b200.units import UnitRegistry

    ureg = UnitRegistry()
    dt = ureg.Quantity(DT, "second")
    md = sw.MDParams(ureg=ureg, fluid_dyn_viscosity=ureg.Quantity(8.9e-4, "pascal * second"),
                     WCA_epsilon=ureg.Quantity(297.0, "kelvin") * ureg.boltzmann_constant,
                     temperature=ureg.Quantity(300.0, "kelvin"), box_length=ureg.Quantity(3 * [BOX], "micrometer"),
                     time_step=dt, time_slice=dt * STEPS_PER_SLICE, write_interval=dt * STEPS_PER_SLICE * 100000)
    tmp = tempfile.mkdtemp(prefix="swarm_bench_")
    e = sw.SwarmEngine(md_params=md, n_dims=2, seed=SEED, out_folder=tmp, skin=args.skin)
    pos, th = initial_state(N_AGENTS, BOX, SEED)
    dirs = np.stack([np.cos(th), np.sin(th), np.zeros_like(th)], axis=1)
    e.add_colloids_at(ureg.Quantity(pos, "micrometer"), dirs, ureg.Quantity(1.0, "micrometer"), type_colloid=0)
    rng = np.random.default_rng(0)
    D, H, A = VISION["n_cones"], 128, 4
    net = dict(w1=rng.normal(size=(D, H)) * 0.5, b1=rng.normal(size=H) * 0.1, wa=rng.normal(size=(H, A)) * 0.2,
               ba=np.zeros(A), wv=rng.normal(size=H) * 0.2, bv=np.zeros(1))
    gt, gr = stokes(1.0)
    actions = {"Translate": sw.Action(force=5.0 * gt), "RotateCCW": sw.Action(torque=np.array([0.0, 0.0, gr * 10.0])),
               "RotateCW": sw.Action(torque=np.array([0.0, 0.0, -gr * 10.0])), "Nothing": sw.Action()}
    box = np.array([BOX] * 3)
    src = np.array([0.5 * BOX, 0.5 * BOX, 0.0])
    agent = sw.DeviceAgent(
        0, sw.FusedActorCritic(net),
        GradientSensing(source=src, decay_function=decay.Linear(1.0, -1.0), box_length=box, reward_scale_factor=10, particle_type=0),
        SubdividedVisionCones(VISION["range"], VISION["half_angle"], VISION["n_cones"], np.ones(N_AGENTS), detected_types=[0],
                              particle_type=0),
        actions, train=True, keep_trajectory=False, reward_mean=True)
    ff = sw.ForceFunction({"0": agent})
    agent.reset_agent(e.colloids)
    e.integrate(max(W, 2), ff)          # first call: overlap removal (1000 steepest-descent steps), the checked chunk, warm-up
    # one untimed call of the timed shape: the host runs ahead of the device inside integrate(K), so the caching allocator
    # grows its pool (cudaMalloc, a device synchronisation each) the first time K slices are in flight
    e.integrate(K, ff)
    torch.stack(agent.reward_means)[:, 1].cpu()   # and of the read-out: torch loads the kernels of stack / index lazily (tens of ms, once per process)
    st0 = e._backend.stats()
    agent.reward_means.clear()
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=e._backend.dev)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    flush.fill_(1)
    torch.cuda.synchronize()
    a.record()
    e.integrate(K, ff)
    means = torch.stack(agent.reward_means)[:, 1].cpu()
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b)
    st1 = e._backend.stats()
    # the same, one integrate(1) call per slice with the L2 flushed between the calls (flush outside the events)
    agent.reward_means.clear()
    ms_f = 0.0
    for _ in range(K):
        flush.fill_(1)
        a.record()
        e.integrate(1, ff)
        m1 = agent.reward_means[-1].cpu()
        b.record()
        torch.cuda.synchronize()
        ms_f += a.elapsed_time(b)
    rec = {"api": "SwarmEngine.integrate(K, ForceFunction({'0': DeviceAgent(FusedActorCritic(5-128-4/1), GradientSensing, "
                  "SubdividedVisionCones)}))", "value": N_AGENTS * STEPS_PER_SLICE * K / (ms * 1e-3), "unit": UNIT,
           "ms_per_step": ms / K, "slices": K, "d2h_bytes_per_step": 8, "h2d_bytes_per_step": 0,
           "gpu_launches_per_slice": (st1.launches - st0.launches) / K, "reward_mean_last": float(means[-1]),
           "l2": "one integrate(K) call: no flush between its slices (the state is L2-resident in every arm)",
           "flushed_per_call": {"ms_per_step": ms_f / K, "note": "K calls of integrate(1), 256 MiB L2 flush between the calls, each "
                                "call ends with the engine's one synchronising counter check"},
           "tiling": e._backend.tiling_stats(), "reward_mean_check": float(m1[1])}
    e.finalize()
    del flush
    e._backend.close()
    torch.cuda.empty_cache()
    return rec


def extra_records(args, world, rank, local):
    """c5_domain and c3_envs (see the module docstring); every rank takes part, rank 0 gets the records."""
    import torch
    import torch.distributed as dist

    from tools import bench_domain, bench_envs

    out = {}
    # ---- C3: batched environments, ensemble-sharded --------------------------------------------------
    strong = bench_envs.measure(world, rank, local, args.c3_envs, weak=False, slices=args.c3_slices)
    weak = strong if world == 1 else bench_envs.measure(world, rank, local, args.c3_envs, weak=True, slices=args.c3_slices)
    keys = ("value", "unit", "agent_steps_per_s", "envs", "envs_per_gpu", "ms_per_slice", "slices", "launches_per_slice")
    out["c3_envs"] = {"workload": f"C3: {args.c3_envs} independent environments x 32 agents, 100 steps/slice, ConcentrationField "
                                  "observable -> counter-RNG policy -> 4-action table -> Brownian steps + WCA -> GradientSensing mean",
                      "metric": "env_steps_per_sec_batched_rollouts", "n_gpus": world,
                      "strong": {k: strong[k] for k in keys}, "weak": {k: weak[k] for k in keys},
                      "collective": "none on the data path (one all-reduce of the reward sum per measurement)"}
    # ---- C5: one large swarm, domain-decomposed ------------------------------------------------------
    swarm = bench_domain.build_swarm(args.c5_agents, 0.10)
    kw = dict(steps=STEPS_PER_SLICE, slices=args.c5_slices, warmup=2, k=args.c5_k, skin=args.c5_skin, margin=0.15)
    rec, dom = bench_domain.measure(world, rank, local, swarm, keep=True, **kw)
    rec["workload"] = (f"C5: {args.c5_agents:,}-particle single swarm, 2-D periodic phi=0.10, WCA + Brownian dynamics kT=300K dt=1e-3, "
                       f"{STEPS_PER_SLICE} steps/slice + GradientSensing reward mean; {world} slab(s), redundant-halo exchange every k steps")
    if world > 1:
        full_n, owned = bench_domain.gather_owned(dom, world, rank)
        dom.close()
        del dom
        torch.cuda.empty_cache()
        dist.barrier()
        if rank == 0:
            # the same swarm on this GPU alone, same parameters: the N = 1 point of the strong-scaling curve
            one, dom1 = bench_domain.measure(1, 0, local, swarm, keep=True, **kw)
            full_1, _ = bench_domain.gather_owned(dom1, 1, 0)
            dom1.close()
            rec["n1_same_run"] = {k: one[k] for k in ("value", "ms_per_slice", "rebuilds", "redundancy")}
            rec["efficiency_vs_n1"] = rec["value"] / (world * one["value"])
            err = max(bench_domain.rel_err(full_n[c], full_1[c]) for c in range(3))
            rec["parity_full"] = {"n_agents": args.c5_agents, "steps": (2 + args.c5_slices) * STEPS_PER_SLICE, "noise": True,
                                  "against": "the same swarm on one GPU (self-exchanging slab), same kernels",
                                  "max_rel_err_x_y_theta": err, "every_particle_owned_once": bool(owned == args.c5_agents),
                                  "note": "chaotic growth of rounding differences between local frames over 700 steps of contacts "
                                          "is expected above 1e-12; the per-step bound (1e-9) is what parity_selfcheck pins"}
        dist.barrier()
    else:
        dom.close()
        rec["efficiency_vs_n1"] = 1.0
    chk = bench_domain.parity_selfcheck(world, rank, local)
    if rank == 0:
        rec["parity_selfcheck"] = chk
        tl = rec["timeline_ms_per_slice_max_over_ranks"]
        rec["limiting_term"] = max(tl, key=tl.get)
    out["c5_domain"] = rec
    return out if rank == 0 else None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--skin", type=float, default=2.0, help="Verlet skin (performance only; espresso.py:269 uses 0.4)")
    ap.add_argument("--cpu-steps", type=int, default=100, help="BD steps of the cpu_baseline sample")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the c5_domain / c3_envs records")
    ap.add_argument("--no-engine", action="store_true", help="skip the e2e_engine record (SwarmEngine.integrate with a device agent)")
    ap.add_argument("--c5-agents", type=int, default=8_000_000)
    ap.add_argument("--c5-slices", type=int, default=5)
    ap.add_argument("--c5-k", type=int, default=33, help="steps between halo exchanges (halo = k * r_cut * 1.15)")
    ap.add_argument("--c5-skin", type=float, default=2.0)
    ap.add_argument("--c3-envs", type=int, default=4096)
    ap.add_argument("--c3-slices", type=int, default=50)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
