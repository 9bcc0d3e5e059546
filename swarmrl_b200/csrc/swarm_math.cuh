// swarm_math.cuh -- branch-free fp64 kernels for the Box-Muller transform of the thermostat noise.
//
// The inputs are Philox words, not arbitrary doubles: u = (k + 1/2) 2^-52 with k a 52-bit integer, so
// the range reduction of sin/cos(2 pi u) is done on the integer and log(u) never sees 0, inf, NaN or a
// denormal.  That removes every special-case branch of the CUDA math library versions and about half
// of their fp64 instructions (the step kernel is fp64-pipe / issue bound).  Accuracy:
// <= 2 ulp (checked against mpmath in tools/gen_coeffs.py and against the oracle's libm on the GPU in
// tests/test_gpu_engine.py::test_philox_gaussians_match_oracle).  Coefficients: tools/gen_coeffs.py.
#pragma once
#include "swarm_common.cuh"

// 64 x 64 -> 128 bit product from four 32 x 32 -> 64 partial products (IMAD.WIDE.U32 with a 64-bit
// addend); nvcc's separate mul.hi.u64 / mul.lo.u64 recompute the partial products.
__device__ __forceinline__ void swm_mulhilo(u64 a, u64 b, u64 &hi, u64 &lo) {
  const unsigned a0 = (unsigned)a, a1 = (unsigned)(a >> 32), b0 = (unsigned)b, b1 = (unsigned)(b >> 32);
  const u64 t = (u64)a0 * b0;
  const u64 u = (u64)a1 * b0 + (t >> 32);
  const u64 v = (u64)a0 * b1 + (unsigned)u;
  hi = (u64)a1 * b1 + (u >> 32) + (v >> 32);
  lo = (v << 32) | (unsigned)t;
}

__device__ __forceinline__ void swm_philox4x64_10(u64 c0, u64 c1, u64 c2, u64 c3, u64 k0, u64 k1, u64 out[4]) {
  const u64 M0 = 0xD2E7470EE14C6C93ULL, M1 = 0xCA5A826395121157ULL;
  const u64 W0 = 0x9E3779B97F4A7C15ULL, W1 = 0xBB67AE8584CAA73BULL;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    u64 hi0, lo0, hi1, lo1;
    swm_mulhilo(M0, c0, hi0, lo0);
    swm_mulhilo(M1, c2, hi1, lo1);
    const u64 n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += W0; k1 += W1;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// u = ((r >> 12) + 1/2) 2^-52, bit-identical to sw_u01 / the oracle's u01, without an int->double convert
__device__ __forceinline__ double swm_u01(u64 r) {
  const double d = __longlong_as_double((long long)((r >> 12) | 0x3FF0000000000000ULL));  // [1, 2)
  return d - 0.99999999999999988897769753748;  // 1 - 2^-53 (exact); the difference is exact
}

// log(u), u in (0, 1) normal
__device__ __forceinline__ double swm_log01(double u) {
  // u = m 2^e, m in [sqrt(1/2), sqrt(2))
  int hi = __double2hiint(u), lo = __double2loint(u);
  int e = (hi >> 20) - 1023;
  hi = (hi & 0x000FFFFF) | 0x3FF00000;   // m in [1, 2)
  if (hi >= 0x3FF6A09F) { hi -= 0x00100000; e += 1; }  // m >= ~sqrt(2): halve
  const double m = __hiloint2double(hi, lo);
  const double f = m - 1.0;
  const double s = f / (2.0 + f);
  const double z = s * s;
  double p = 0.08695652173913043;
  p = fma(p, z, 0.09523809523809523);
  p = fma(p, z, 0.10526315789473684);
  p = fma(p, z, 0.11764705882352941);
  p = fma(p, z, 0.13333333333333333);
  p = fma(p, z, 0.15384615384615385);
  p = fma(p, z, 0.18181818181818182);
  p = fma(p, z, 0.2222222222222222);
  p = fma(p, z, 0.2857142857142857);
  p = fma(p, z, 0.4);
  p = fma(p, z, 0.6666666666666666);
  // log m = 2 s + s z p ;  log u = e ln2 + log m
  const double ed = (double)e;
  const double r = fma(s * z, p, fma(ed, 2.3190468138462996e-17, 0.0));
  return fma(ed, 0.6931471805599453, fma(2.0, s, r));
}

// (sin, cos)(2 pi u) for the Philox word r, u = ((r >> 12) + 1/2) 2^-52
__device__ __forceinline__ void swm_sincos2pi(u64 r, double &sn, double &cs) {
  const u64 k = r >> 12;                               // 52 bits
  const unsigned q = (unsigned)((k + (1ULL << 49)) >> 50);  // rint(4 u) in 0..4
  // t = u - q/4 in [-1/8, 1/8], exact
  const double u = swm_u01(r);
  const double t = fma((double)q, -0.25, u);
  const double z = t * t;
  double ps = 0.10422916220813984;
  ps = fma(ps, z, -0.7181223017785006);
  ps = fma(ps, z, 3.819952584848282);
  ps = fma(ps, z, -15.09464257682299);
  ps = fma(ps, z, 42.058693944897655);
  ps = fma(ps, z, -76.70585975306139);
  ps = fma(ps, z, 81.60524927607506);
  ps = fma(ps, z, -41.34170224039976);
  ps = fma(ps, z, 6.283185307179586);
  ps *= t;
  double pc = -0.03638284114254567;
  pc = fma(pc, z, 0.28200596845579123);
  pc = fma(pc, z, -1.714390711088672);
  pc = fma(pc, z, 7.903536371318469);
  pc = fma(pc, z, -26.4262567833744);
  pc = fma(pc, z, 60.24464137187666);
  pc = fma(pc, z, -85.45681720669373);
  pc = fma(pc, z, 64.9393940226683);
  pc = fma(pc, z, -19.739208802178716);
  pc = fma(pc, z, 1.0);
  // quadrant: angle = q pi/2 + 2 pi t
  const bool swap = q & 1u;
  double s0 = swap ? pc : ps, c0 = swap ? ps : pc;
  if ((q + 1u) & 2u) c0 = -c0;   // q = 1, 2: cos negative
  if (q & 2u) s0 = -s0;          // q = 2, 3: sin negative
  sn = s0; cs = c0;
}

// three of the four normals of (seed, id, step, stream 0): g0, g1 (pair 0) and g3 (sine of pair 1) --
// what a 2-D Brownian step consumes (x, y, theta).  Same values as sw_gaussians()[0,1,3] to ~1e-15.
__device__ __forceinline__ void swm_gaussians_xyt(u64 seed, u64 id, u64 step, double &g0, double &g1, double &g3) {
  u64 r[4];
  swm_philox4x64_10(step, 0ULL, 0ULL, 0ULL, id, seed, r);
  const double m0 = sqrt(-2.0 * swm_log01(swm_u01(r[0])));
  const double m1 = sqrt(-2.0 * swm_log01(swm_u01(r[2])));
  double s0, c0, s1, c1;
  swm_sincos2pi(r[1], s0, c0);
  swm_sincos2pi(r[3], s1, c1);
  g0 = m0 * c0; g1 = m0 * s0; g3 = m1 * s1;
}

// all four normals with the fast kernels (parity hook / 3-D)
__device__ __forceinline__ void swm_gaussians(u64 seed, u64 id, u64 step, u64 stream, double g[4]) {
  u64 r[4];
  swm_philox4x64_10(step, stream, 0ULL, 0ULL, id, seed, r);
  const double m0 = sqrt(-2.0 * swm_log01(swm_u01(r[0])));
  const double m1 = sqrt(-2.0 * swm_log01(swm_u01(r[2])));
  double s0, c0, s1, c1;
  swm_sincos2pi(r[1], s0, c0);
  swm_sincos2pi(r[3], s1, c1);
  g[0] = m0 * c0; g[1] = m0 * s0; g[2] = m1 * c1; g[3] = m1 * s1;
}
