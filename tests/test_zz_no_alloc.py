"""
swarm_obs_field_mean_noalloc (include/swarm_b200.h): the gradient reward + its mean with the per-CTA sums in a static
device array instead of a cudaMalloc'd scratch buffer.  No counterpart in the reference (added
for callers that cannot allocate: CUDA graph capture, serving loops).  Checked through the C-ABI against
swarm_obs_field_mean on two calls in a row, and for its size limit.  Collected last on purpose (file name).
"""
import ctypes as C

import numpy as np
import pytest

from helpers import jittered_lattice, make_pair, species_for, wca_for

pytestmark = pytest.mark.gpu


def _swarm(n, L=300.0, seed=11):
    rng = np.random.default_rng(seed)
    sp, wca = species_for([1.0, 1.0]), wca_for([1.0, 1.0])
    pos = jittered_lattice(n, L, 2, rng)
    th = rng.uniform(0, 2 * np.pi, n)
    sidx = (rng.uniform(size=n) < 0.3).astype(np.int32)
    (_, d) = make_pair(2, (L,) * 3, True, 1e-3, 0.0, 1, sp, wca, pos, sidx, theta=th)
    return d, rng


def _call(d, fn, row_of, rows, src, box, hist, kind, c0, c1, out, red):
    from swarmrl_b200 import _lib

    s3 = (C.c_double * 3)(*src)
    b3 = (C.c_double * 3)(*box)
    _lib.check(fn(d.h, row_of.data_ptr(), int(rows), s3, b3, hist.data_ptr(), int(kind), float(c0), float(c1), 100.0, 1,
                  out.data_ptr(), red.data_ptr(), d._stream()))


def test_noalloc_field_mean_equals_the_allocating_entry_point_twice_in_a_row():
    import torch

    d, rng = _swarm(5000)
    n, L, src = d.n, 300.0, (0.5, 0.4, 0.0)
    row_of, rows = d.row_map(0)
    assert 0 < rows < n
    f64 = dict(dtype=torch.float64, device=d.dev)
    hist0 = (d.pos[:, :n] / L).clone().contiguous()
    for call, (kind, c0, c1) in enumerate(((1, 1.0, -1.0), (2, 2.0, 0.5))):   # two calls in a row on the shared scratch
        shift = rng.normal(size=(3, n)) * 3.0
        shift[2] = 0.0
        d.pos[:, :n] += torch.from_numpy(shift).to(d.dev)
        ha, hb = hist0.clone(), hist0.clone()
        # every output buffer exists BEFORE the call: the entry point itself allocates nothing
        out_a, out_b = torch.full((rows,), -1.0, **f64), torch.full((rows,), -2.0, **f64)
        red_a, red_b = torch.full((2,), -1.0, **f64), torch.full((2,), -2.0, **f64)
        _call(d, d.lib.swarm_obs_field_mean, row_of, rows, src, (L, L, L), ha, kind, c0, c1, out_a, red_a)
        free_before = torch.cuda.mem_get_info(d.dev)[0]
        _call(d, d.lib.swarm_obs_field_mean_noalloc, row_of, rows, src, (L, L, L), hb, kind, c0, c1, out_b, red_b)
        torch.cuda.synchronize(d.dev)
        assert torch.cuda.mem_get_info(d.dev)[0] == free_before, "the no-alloc entry point changed the free device memory"
        a, b = out_a.cpu().numpy(), out_b.cpu().numpy()
        assert a.sum() > 0.0, call
        assert np.array_equal(a, b), call                      # same kernel, same arithmetic: bit for bit
        assert torch.equal(ha, hb), call                       # history advanced identically
        ra, rb = red_a.cpu().numpy(), red_b.cpu().numpy()
        assert np.array_equal(ra, rb), (call, ra, rb)          # same CTA sums in the same order
        assert abs(rb[0] - a.sum()) <= 1e-12 * a.sum() and abs(rb[1] - a.mean()) <= 1e-12 * a.mean()


def test_noalloc_field_mean_rejects_bad_arguments():
    import torch
    from swarmrl_b200 import _lib

    d, _ = _swarm(300)
    row_of, rows = d.row_map(0)
    f64 = dict(dtype=torch.float64, device=d.dev)
    hist = torch.zeros((3, d.n), **f64)
    out, red = torch.zeros(max(rows, 1), **f64), torch.zeros(2, **f64)
    s3, b3 = (C.c_double * 3)(0.5, 0.5, 0.0), (C.c_double * 3)(300.0, 300.0, 300.0)
    fn = d.lib.swarm_obs_field_mean_noalloc
    # decay NONE has no reward to average
    rc = fn(d.h, row_of.data_ptr(), rows, s3, b3, hist.data_ptr(), int(_lib.DECAY_NONE), 1.0, -1.0, 1.0, 1, out.data_ptr(),
            red.data_ptr(), d._stream())
    assert rc != 0
    rc = fn(d.h, row_of.data_ptr(), rows, s3, b3, hist.data_ptr(), 1, 1.0, -1.0, 1.0, 1, out.data_ptr(), None, d._stream())
    assert rc != 0
    with pytest.raises(Exception):
        _lib.check(rc)


def test_noalloc_field_mean_refuses_more_particles_than_the_static_scratch_holds():
    """8 388 608 particles (32768 CTA sums) is the documented limit: one more is SWARM_ERR_INVALID with a clear message,
    the limit itself still matches the allocating entry point."""
    import torch
    from swarmrl_b200 import _lib
    from swarmrl_b200.device import DeviceSwarm

    sp = [dict(type=0, gamma_t=(1.0,) * 3, gamma_r=(1.0,) * 3, mass=1.0, rinertia=(1.0,) * 3)]
    zero = dict(sigma=np.zeros((1, 1)), eps=np.zeros((1, 1)), cut=np.zeros((1, 1)))
    rng = np.random.default_rng(3)
    limit = 32768 * 256
    s3, b3 = (C.c_double * 3)(0.5, 0.5, 0.0), (C.c_double * 3)(1.0, 1.0, 1.0)
    for n in (limit, limit + 1):
        d = DeviceSwarm(2, (1.0, 1.0, 1.0), False, 1.0, 0.0, 0, sp, zero)
        pos = np.zeros((n, 3))
        pos[:, :2] = rng.uniform(size=(n, 2))
        d.set_state(pos, np.zeros(n, dtype=np.int32), theta=np.zeros(n))
        row_of, rows = d.row_map(0)
        f64 = dict(dtype=torch.float64, device=d.dev)
        hist = torch.full((3, n), 0.5, **f64)
        out, red = torch.zeros(n, **f64), torch.zeros(2, **f64)
        rc = d.lib.swarm_obs_field_mean_noalloc(d.h, row_of.data_ptr(), rows, s3, b3, hist.data_ptr(), 1, 1.0, -1.0, 10.0, 1,
                                                out.data_ptr(), red.data_ptr(), d._stream())
        if n > limit:
            assert rc == _lib.ERR_INVALID and "8388608" in _lib.last_error()
        else:
            assert rc == 0
            hist2 = torch.full((3, n), 0.5, **f64)
            out2, red2 = torch.zeros(n, **f64), torch.zeros(2, **f64)
            _lib.check(d.lib.swarm_obs_field_mean(d.h, row_of.data_ptr(), rows, s3, b3, hist2.data_ptr(), 1, 1.0, -1.0, 10.0, 1,
                                                  out2.data_ptr(), red2.data_ptr(), d._stream()))
            assert torch.equal(out, out2) and torch.equal(red, red2) and float(red[0]) > 0.0
        d.close()
        del d, hist, out, red
