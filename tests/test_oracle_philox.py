"""Pins the oracle's Philox generators bit-exactly: Philox4x64-10 against numpy.random.Philox (Random123's
Philox4x64-10; numpy increments the counter BEFORE producing a block) and Philox4x32-10 -- the thermostat's --
against Random123's published known answers and an independent pure-Python restatement."""
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


def _philox4x32_py(ctr, key):
    M0, M1, W0, W1 = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85
    c, k = list(ctr), list(key)
    for _ in range(10):
        p0, p1 = M0 * c[0], M1 * c[2]
        c = [(p1 >> 32) ^ c[1] ^ k[0], p1 & 0xFFFFFFFF, (p0 >> 32) ^ c[3] ^ k[1], p0 & 0xFFFFFFFF]
        k = [(k[0] + W0) & 0xFFFFFFFF, (k[1] + W1) & 0xFFFFFFFF]
    return c


@pytest.mark.parametrize("ctr,key,want", [
    # Random123 kat_vectors, "philox4x32 10"
    ([0, 0, 0, 0], [0, 0], [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]),
    ([0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2, [0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD]),
    ([0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344], [0xA4093822, 0x299F31D0],
     [0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1]),
])
def test_philox4x32_random123_known_answers(ctr, key, want):
    assert [int(v) for v in oracle.philox4x32(ctr, key)] == want
    assert _philox4x32_py(ctr, key) == want


def test_noise_words_layout():
    """key = (seed lo, seed hi); counter = (step lo, step hi, id lo, id[32..55] << 8 | stream << 1 | block);
    words 2b, 2b + 1 = block b's (out[0] << 32 | out[1]), (out[2] << 32 | out[3]) -- one block per Box-Muller pair."""
    rng = np.random.default_rng(5)
    for _ in range(50):
        seed, step = (int(v) for v in rng.integers(0, 2**63, 2, dtype=np.uint64) * 2 + 1)
        pid = int(rng.integers(0, 2**56, dtype=np.uint64))
        stream = int(rng.integers(0, 6))
        key = [seed & 0xFFFFFFFF, seed >> 32]
        c3 = (((pid >> 32) & 0xFFFFFF) << 8) | (stream << 1)
        a = _philox4x32_py([step & 0xFFFFFFFF, step >> 32, pid & 0xFFFFFFFF, c3], key)
        b = _philox4x32_py([step & 0xFFFFFFFF, step >> 32, pid & 0xFFFFFFFF, c3 | 1], key)
        want = [(a[0] << 32) | a[1], (a[2] << 32) | a[3], (b[0] << 32) | b[1], (b[2] << 32) | b[3]]
        assert [int(v) for v in oracle.noise_words(seed, pid, step, stream)] == want
    # neighbouring ids / steps / streams never share a word
    seen = set()
    for pid in range(4):
        for step in range(4):
            for stream in range(6):
                seen.update(int(v) for v in oracle.noise_words(42, pid, step, stream))
    assert len(seen) == 4 * 4 * 6 * 4


def test_gaussians_layout_and_moments():
    """u = (word >> 12 + 1/2) 2^-52 of the four noise words; Box-Muller on (u0, u1), (u2, u3)."""
    seed, pid, step, stream = 42, 7, 1000, 3
    r = oracle.noise_words(seed, pid, step, stream)
    u = ((r >> np.uint64(12)).astype(np.float64) + 0.5) / 2.0**52
    m0, m1 = np.sqrt(-2 * np.log(u[0])), np.sqrt(-2 * np.log(u[2]))
    want = [m0 * np.cos(2 * np.pi * u[1]), m0 * np.sin(2 * np.pi * u[1]), m1 * np.cos(2 * np.pi * u[3]),
            m1 * np.sin(2 * np.pi * u[3])]
    assert np.allclose(oracle.gaussians(seed, pid, step, stream), want, rtol=1e-14, atol=1e-15)
    g = np.array([oracle.gaussians(1, i, s, 0) for i in range(200) for s in range(25)]).ravel()
    assert abs(g.mean()) < 4 / np.sqrt(g.size)
    assert abs(g.var() - 1.0) < 4 * np.sqrt(2.0 / g.size)
