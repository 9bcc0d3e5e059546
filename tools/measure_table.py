"""Numbers of BASELINE.md section 5 that bench.py does not carry: C1 / C2 throughput (GPU and CPU oracle, 1 thread and all
threads), the fp64 peak (DFMA micro-benchmark), the cost of the trajectory writer at C4.  One JSON line per record.
    python tools/measure_table.py > profiles/r02_measure_table.jsonl"""
import json, os, sys, pathlib, tempfile, time, types
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent / "tests"))
import numpy as np, torch
import bench
from helpers import KT_300, jittered_lattice, species_for, wca_for
from swarmrl_b200.device import DeviceSwarm
from oracle import oracle
from oracle.oracle import OracleSwarm


def gpu_rate(n, L, steps, reps, skin=0.4, kT=KT_300):
    rng = np.random.default_rng(1234)
    sp, wca = species_for([1.0]), wca_for([1.0])
    pos = jittered_lattice(n, L, 2, rng) if n > 10 else np.c_[rng.uniform(100, 900, (n, 2)), np.zeros(n)]
    th = rng.uniform(0, 2 * np.pi, n)
    d = DeviceSwarm(2, (L,) * 3, True, 1e-3, kT, 42, sp, wca, skin=skin)
    d.set_state(pos, np.zeros(n, dtype=np.int32), theta=th)
    d.set_actions_host(np.full(n, 5.0 * sp[0]["gamma_t"][0]), np.zeros((n, 3)))
    d.run_checked(steps, 0)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for r in range(reps): d.run(steps, (r + 1) * steps)
    b.record(); torch.cuda.synchronize()
    st = d.stats()
    assert not st.overflow
    return n * steps * reps / (a.elapsed_time(b) * 1e-3), a.elapsed_time(b) / (steps * reps) * 1e3, (pos, th, sp, wca)


def cpu_rate(n, L, steps, threads, state):
    pos, th, sp, wca = state
    oracle.set_num_threads(threads)
    o = OracleSwarm(2, (L,) * 3, True, 1e-3, KT_300, 42, sp, wca)
    o.set_state(pos, np.zeros(n, dtype=np.int32), theta=th)
    o.f_swim[:] = 5.0 * sp[0]["gamma_t"][0]
    o.bd_steps(0, max(1, steps // 10), mode=2)
    t0 = time.perf_counter(); o.bd_steps(steps, steps, mode=2); t = time.perf_counter() - t0
    return n * steps / t


cores = len(os.sched_getaffinity(0))
for name, n, L, steps, reps, skin in (("C1 (10 colloids, box 1000, 100 steps/slice)", 10, 1000.0, 100, 200, 0.4),
                                      ("C2 (1,000 colloids, L=125, 1,000 steps/slice)", 1000, 125.0, 1000, 20, 0.4)):
    g, us, state = gpu_rate(n, L, steps, reps, skin)
    c1, ca = cpu_rate(n, L, steps * 5, 1, state), cpu_rate(n, L, steps * 5, cores, state)
    print(json.dumps({"config": name, "gpu_agent_steps_per_s": g, "gpu_us_per_step": us, "cpu_oracle_1_thread": c1,
                      f"cpu_oracle_{cores}_threads": ca, "integrator only (no observable)": True, "skin": skin}), flush=True)

# fp64 peak
import ctypes as C
from swarmrl_b200 import _lib
out = (C.c_double * 4)()
for ilp in (1, 4, 8):
    _lib.check(_lib.load().swarm_microbench_dfma(0, ilp, 2048, out))
    print(json.dumps({"dfma_microbench_ilp": ilp, "dfma_per_s": out[0], "fp64_tflops": 2 * out[0] / 1e12,
                      "dfma_per_clk_per_sm": out[0] / (out[2] * 1.965e9)}), flush=True)

# trajectory writer at C4: a frame every slice against no frames
args = types.SimpleNamespace(skin=2.0)
import swarmrl_b200 as sw
base = bench.engine_record(args, 0, 40, 3)
_MD = sw.MDParams
class EverySlice(_MD):
    def __init__(self, **kw):
        kw["write_interval"] = kw["time_slice"]
        super().__init__(**kw)
sw.MDParams = EverySlice
wr = bench.engine_record(args, 0, 40, 3)
sw.MDParams = _MD
print(json.dumps({"trajectory_writer_C4": {"ms_per_slice_no_frames": base["ms_per_step"], "ms_per_slice_frame_every_slice": wr["ms_per_step"],
                                           "overhead": wr["ms_per_step"] / base["ms_per_step"] - 1.0,
                                           "frame_bytes": 72 * bench.N_AGENTS, "note": "device ring -> pinned host on a copy stream -> writer thread; 40 timed slices after 43 untimed ones"}}))
