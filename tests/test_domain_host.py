"""
Host logic of the slab domain decomposition (swarmrl_b200/domain.py) on CPU: slab geometry, the initial
scatter, and the neighbour transport + all-reduce over torch.distributed `gloo` at world_size 2 and 3
(the N > 1 path; on the GPU box the same code runs over NCCL, tests/test_gpu_domain.py).
"""
import socket

import numpy as np
import pytest

from swarmrl_b200.domain import select_local, slab_plan, unfold_global


def test_slab_plan_geometry_and_errors():
    p = slab_plan(160.0, 4, 3, exchange_every=2, max_cut=2.0, margin=0.25)
    assert p["slab_w"] == 40.0 and p["halo"] == 5.0 and p["x0"] == 120.0
    assert (p["left"], p["right"]) == (2, 0)
    assert (p["image_shift_left"], p["image_shift_right"]) == (0, 1)
    assert slab_plan(160.0, 4, 0, 2, 2.0)["image_shift_left"] == -1
    with pytest.raises(ValueError):
        slab_plan(160.0, 4, 0, exchange_every=9, max_cut=2.0)      # slab 40 < 2 x 22.5
    with pytest.raises(ValueError):
        slab_plan(160.0, 4, 4, 2, 2.0)
    with pytest.raises(ValueError):
        slab_plan(160.0, 4, 0, 0, 2.0)


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_scatter_owns_every_particle_once_and_round_trips(world):
    rng = np.random.default_rng(5)
    L, n = 400.0, 5000
    x = rng.uniform(-2 * L, 3 * L, n)            # unfolded inputs spread over five periodic images
    owners = np.zeros(n, dtype=int)
    for r in range(world):
        p = slab_plan(L, world, r, 4, 2.0)
        idx, xl, im, owned = select_local(x, p)
        assert np.all((xl >= 0.0) & (xl < p["slab_w"] + 2 * p["halo"]))
        # every kept copy is a periodic image of its particle: same x modulo L, image count consistent
        xg = unfold_global(xl, im, p)
        assert np.allclose(xg, x[idx], rtol=0, atol=1e-9)
        np.add.at(owners, idx[owned == 1], 1)
        # ghosts = exactly the particles within one halo outside the slab (modulo L)
        xf = np.mod(x - p["x0"], L)
        expect_ghost = ((xf >= p["slab_w"]) & (xf < p["slab_w"] + p["halo"])) | (xf >= L - p["halo"])
        if world == 1:
            expect_ghost = (xf < p["halo"]) | (xf >= L - p["halo"])
        got = np.zeros(n, dtype=int)
        np.add.at(got, idx[owned == 0], 1)
        assert np.array_equal(got > 0, expect_ghost)
    assert np.all(owners == 1)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _transport_rank(rank, world, port, q):
    import torch
    import torch.distributed as dist

    from swarmrl_b200.domain import _Transport

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        p = slab_plan(100.0 * world, world, rank, 2, 2.0)
        tr = _Transport(p)
        send_l = torch.full((9,), 10.0 * rank + 1.0, dtype=torch.float64)   # "through my left face"
        send_r = torch.full((9,), 10.0 * rank + 2.0, dtype=torch.float64)
        recv_l, recv_r = torch.zeros(9, dtype=torch.float64), torch.zeros(9, dtype=torch.float64)
        for _ in range(3):                                                  # repeated exchanges must not cross
            tr.exchange(send_l, send_r, recv_l, recv_r)
        tot = tr.all_reduce_sum(torch.tensor([float(rank + 1)], dtype=torch.float64))
        q.put((rank, float(recv_l[0]), float(recv_r[0]), float(tot[0])))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_gloo_neighbour_transport(world):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q, port = ctx.Queue(), _free_port()
    procs = [ctx.Process(target=_transport_rank, args=(r, world, port, q)) for r in range(world)]
    for pr in procs:
        pr.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for pr in procs:
        pr.join(timeout=60)
        assert pr.exitcode == 0
    for rank, from_left, from_right, tot in res:
        left, right = (rank - 1) % world, (rank + 1) % world
        assert from_left == 10.0 * left + 2.0      # my left neighbour's send_r
        assert from_right == 10.0 * right + 1.0    # my right neighbour's send_l
        assert tot == world * (world + 1) / 2


def test_slab_plan_widens_the_halo_by_the_observable_range():
    """"per slice a wider halo" -- a per-agent observable of range R (vision cones) is exact for the
    particles the integration needs only if the rank also holds everything within R of THEM."""
    from swarmrl_b200.domain import slab_plan

    base = slab_plan(1000.0, 4, 1, exchange_every=10, max_cut=2.0, margin=0.25)
    wide = slab_plan(1000.0, 4, 1, exchange_every=10, max_cut=2.0, margin=0.25, halo_extra=12.5)
    assert wide["halo"] == pytest.approx(base["halo"] + 12.5) and wide["halo_extra"] == 12.5 and base["halo_extra"] == 0.0
    assert wide["x_offset"] == pytest.approx(wide["x0"] - wide["halo"])
    with pytest.raises(ValueError):
        slab_plan(100.0, 4, 0, exchange_every=2, max_cut=2.0, halo_extra=10.0)     # 25-wide slabs cannot hold two 15-wide halos
    with pytest.raises(ValueError):
        slab_plan(1000.0, 4, 0, exchange_every=2, max_cut=2.0, halo_extra=-1.0)
