"""
Domain-decomposed swarm (swarmrl_b200/domain.py, csrc/swarm_dd.cu) against the single-GPU engine on the
same inputs: trajectories within 1e-9 relative (north_star's fp64 bound), with noise ON -- the Philox
key is the global particle id, so the decomposition must not change a single random number -- and the
all-reduced GradientSensing mean against swarm_obs_field + swarm_reduce_mean.
"""
import pathlib
import socket
import subprocess
import sys

import numpy as np
import pytest

from helpers import KT_300, rel_err, wca_for

pytestmark = pytest.mark.gpu
HERE = pathlib.Path(__file__).resolve().parent
TOL = 1e-9


def _single_gpu_reference(kT, langevin=False):
    import torch

    import dd_worker as W
    from swarmrl_b200.device import DeviceSwarm

    pos, theta, acts, sp = W.build_case()
    C = W.CASE
    n, L = C["n"], C["L"]
    if langevin:
        sp = W.langevin_species()
    d = DeviceSwarm(2, (L, L, L), True, C["dt"], kT, C["seed"], sp, wca_for([1.0]), skin=0.4)
    d.set_state(pos, np.zeros(n, dtype=np.int32), theta=theta)
    if langevin:
        d.set_integrator("langevin")
    row_of, rows = d.row_map(0)
    hist = (d.pos[:, :n] / L).contiguous()
    step, means = 0, []
    for f, t in acts:
        tq = np.zeros((n, 3)); tq[:, 2] = t
        d.set_actions_host(f, tq)
        st = d.run_checked(C["steps"], step)
        assert st.steps_done == C["steps"]
        step += C["steps"]
        rew, _, _ = d.obs_field(row_of, rows, (0.5, 0.5, 0.0), (L, L, L), hist, decay=1, c0=1.0, c1=-1.0, scale=10.0,
                                clip=True, want_dist=False)
        means.append(d.reduce_mean(rew).cpu().numpy())
    p, _, _ = d.particle_data_host()
    return p, d.theta[:n].cpu().numpy(), np.array(means)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run_ranks(world, tmp_path, kT, backend="gloo", peer=None, extra=()):
    import os

    port, out = _free_port(), str(tmp_path / "dd")
    env = dict(os.environ)
    if peer is not None:
        env["SWARM_B200_DD_PEER"] = "1" if peer else "0"
    procs = [subprocess.Popen([sys.executable, str(HERE / "dd_worker.py"), str(r), str(world), str(port), out, backend, repr(kT), *extra],
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env) for r in range(world)]
    logs = [p.communicate(timeout=600)[0] for p in procs]
    for p, lg in zip(procs, logs):
        assert p.returncode == 0, lg[-3000:]
    return [np.load(f"{out}.{r}.npz") for r in range(world)]


def _check(parts, ref, n):
    p_ref, th_ref, means_ref = ref
    gid = np.concatenate([q["gid"] for q in parts])
    assert np.array_equal(np.sort(gid), np.arange(n)), "every particle must be owned by exactly one rank"
    x, y, th = (np.concatenate([q[k] for q in parts]) for k in ("x", "y", "theta"))
    o = np.argsort(gid)
    assert rel_err(x[o], p_ref[:, 0]) < TOL
    assert rel_err(y[o], p_ref[:, 1]) < TOL
    assert rel_err(th[o], th_ref) < TOL
    for q in parts:
        assert int(q["exchanges"]) >= 28          # 2 x 61 steps, k = 4
        assert np.allclose(q["means"], means_ref, rtol=1e-9, atol=1e-12)


@pytest.mark.parametrize("kT", [0.0, KT_300])
def test_self_exchange_world1_matches_single_swarm(tmp_path, kT):
    """world 1: the slab is the whole box and its halos are its own periodic images."""
    import dd_worker as W

    parts = _run_ranks(1, tmp_path, kT)
    _check(parts, _single_gpu_reference(kT), W.CASE["n"])


@pytest.mark.parametrize("world", [2, 3])
def test_ranks_sharing_one_gpu_match_single_swarm(tmp_path, world):
    """world 2 / 3 on one device, gloo host-staged transport: same kernels, same protocol as NCCL."""
    import dd_worker as W

    parts = _run_ranks(world, tmp_path, KT_300)
    _check(parts, _single_gpu_reference(KT_300), W.CASE["n"])
    owned = [len(q["gid"]) for q in parts]
    assert min(owned) > 0.5 * W.CASE["n"] / world


@pytest.mark.parametrize("world", [2, 3])
def test_peer_memory_exchange_between_ranks_sharing_one_gpu(tmp_path, world):
    """The exchange over peer memory (swarm_dd_push / swarm_dd_wait: remote stores into the neighbour's CUDA-IPC region and
    a system-scope flag) instead of torch.distributed: same trajectories.  Two processes on ONE device map each other's
    memory exactly as two GPUs do over NVLink; the device time-slices between them, so this is slow but exact."""
    import dd_worker as W

    parts = _run_ranks(world, tmp_path, KT_300, peer=True)
    _check(parts, _single_gpu_reference(KT_300), W.CASE["n"])
    assert all(int(q["peer"]) == 1 for q in parts)


@pytest.mark.parametrize("world", [1, 2])
def test_langevin_integrator_on_the_decomposed_swarm(tmp_path, world):
    """thermostat_type="langevin" (velocity Verlet, velocities are state): vx, vy, omega_z travel as aux channels of the
    exchange record and the runs between two exchanges are cut with swarm_run_part -- no recalculation and no closing half
    kick at a cut -- so the decomposed run reproduces the single-GPU Langevin run, noise on."""
    import dd_worker as W

    parts = _run_ranks(world, tmp_path, KT_300, extra=("langevin",))
    _check(parts, _single_gpu_reference(KT_300, langevin=True), W.CASE["n"])


@pytest.mark.parametrize("peer", [True, False])
def test_nccl_two_gpus_match_single_swarm(tmp_path, peer):
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import dd_worker as W

    parts = _run_ranks(2, tmp_path, KT_300, backend="nccl", peer=peer)
    _check(parts, _single_gpu_reference(KT_300), W.CASE["n"])
    assert all(int(q["peer"]) == int(peer) for q in parts)


def test_halo_too_thin_is_rejected():
    from swarmrl_b200.domain import slab_plan

    with pytest.raises(ValueError):
        slab_plan(100.0, 4, 0, exchange_every=8, max_cut=2.0)
