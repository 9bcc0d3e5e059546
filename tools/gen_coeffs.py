"""Generates the fp64 polynomial coefficients of swarm_math.cuh (Taylor series, exact rationals rounded
to nearest double with mpmath) and checks the evaluation schemes in numpy against mpmath."""
import mpmath as mp
import numpy as np

mp.mp.prec = 200


def hexf(v):
    return float(v).hex()


def sin_coeffs(n):  # sin(2 pi t) = t * sum c_k t^(2k)
    return [(-1) ** k * (2 * mp.pi) ** (2 * k + 1) / mp.factorial(2 * k + 1) for k in range(n)]


def cos_coeffs(n):  # cos(2 pi t) = sum c_k t^(2k)
    return [(-1) ** k * (2 * mp.pi) ** (2 * k) / mp.factorial(2 * k) for k in range(n)]


def log_coeffs(n):  # log(m) = 2 s * sum s^(2k)/(2k+1), s = (m-1)/(m+1)
    return [mp.mpf(2) / (2 * k + 1) for k in range(n)]


def emit(name, cs):
    print(f"// {name}")
    print("{" + ", ".join(f"{float(c)!r}" for c in cs) + "}")


if __name__ == "__main__":
    S, Cc, Lg = sin_coeffs(9), cos_coeffs(10), log_coeffs(12)
    emit("SIN2PI", S); emit("COS2PI", Cc); emit("LOG", Lg)
    print("ln2_hi", repr(float(mp.log(2))), "ln2_lo", repr(float(mp.log(2) - mp.mpf(float(mp.log(2))))))
    # numpy check of the schemes
    rng = np.random.default_rng(0)
    t = rng.uniform(-0.125, 0.125, 200000)
    t2 = t * t
    s = np.zeros_like(t); c = np.zeros_like(t)
    for k in reversed(range(len(S))):
        s = s * t2 + float(S[k])
    s *= t
    for k in reversed(range(len(Cc))):
        c = c * t2 + float(Cc[k])
    es = max(abs(mp.sin(2 * mp.pi * mp.mpf(float(x))) - mp.mpf(float(y))) for x, y in zip(t[:4000], s[:4000]))
    ec = max(abs(mp.cos(2 * mp.pi * mp.mpf(float(x))) - mp.mpf(float(y))) for x, y in zip(t[:4000], c[:4000]))
    print("sin abs err", float(es), "cos abs err", float(ec))
    m = rng.uniform(2 ** -0.5, 2 ** 0.5, 200000)
    f = m - 1.0
    ss = f / (2.0 + f)
    s2 = ss * ss
    p = np.zeros_like(m)
    for k in reversed(range(len(Lg))):
        p = p * s2 + float(Lg[k])
    lg = ss * p
    el = max(abs(mp.log(mp.mpf(float(x))) - mp.mpf(float(y))) / max(abs(mp.log(mp.mpf(float(x)))), mp.mpf(1e-300)) for x, y in zip(m[:4000], lg[:4000]))
    print("log rel err", float(el))
