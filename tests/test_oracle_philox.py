"""Pins the oracle's Philox4x64-10 bit-exactly against numpy.random.Philox (Random123's
Philox4x64-10; numpy increments the counter BEFORE producing a block)."""
import numpy as np
import pytest

from oracle import oracle


@pytest.mark.parametrize("ctr,key", [
    ([0, 0, 0, 0], [0, 0]),
    ([5, 6, 7, 8], [11, 12]),
    ([2**64 - 2, 0, 0, 0], [2**64 - 1, 2**63 + 5]),
    ([123456789, 1, 0, 0], [99999, 42]),
])
def test_block_matches_numpy(ctr, key):
    raw = np.random.Philox(counter=np.array(ctr, dtype=np.uint64), key=np.array(key, dtype=np.uint64)).random_raw(4)
    bumped = list(ctr)
    bumped[0] += 1  # numpy pre-increments; none of the cases carries into word 1
    assert np.array_equal(oracle.philox4x64(bumped, key), raw)


def test_random123_known_answer():
    # Random123 kat_vectors: philox4x64 10, all-zero counter and key
    out = oracle.philox4x64([0, 0, 0, 0], [0, 0])
    assert [hex(int(v)) for v in out] == ["0x16554d9eca36314c", "0xdb20fe9d672d0fdc", "0xd7e772cee186176b",
                                          "0x7e68b68aec7ba23b"]


def test_gaussians_layout_and_moments():
    """counter = (step, stream, 0, 0), key = (id, seed); u = (r>>12 + 1/2) 2^-52; Box-Muller."""
    seed, pid, step, stream = 42, 7, 1000, 3
    r = oracle.philox4x64([step, stream, 0, 0], [pid, seed])
    u = ((r >> np.uint64(12)).astype(np.float64) + 0.5) / 2.0**52
    m0, m1 = np.sqrt(-2 * np.log(u[0])), np.sqrt(-2 * np.log(u[2]))
    want = [m0 * np.cos(2 * np.pi * u[1]), m0 * np.sin(2 * np.pi * u[1]), m1 * np.cos(2 * np.pi * u[3]),
            m1 * np.sin(2 * np.pi * u[3])]
    assert np.allclose(oracle.gaussians(seed, pid, step, stream), want, rtol=1e-14, atol=1e-15)
    g = np.array([oracle.gaussians(1, i, s, 0) for i in range(200) for s in range(25)]).ravel()
    assert abs(g.mean()) < 4 / np.sqrt(g.size)
    assert abs(g.var() - 1.0) < 4 * np.sqrt(2.0 / g.size)
