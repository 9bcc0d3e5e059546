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

// Polynomial coefficients live in constant memory: a DFMA takes a c[bank][offset] operand directly,
// whereas a 64-bit immediate costs two extra moves per coefficient (measured: 140 UMOV per step).
static __constant__ double SWM_LOG[11] = {0.6666666666666666, 0.4, 0.2857142857142857, 0.2222222222222222,
                                          0.18181818181818182, 0.15384615384615385, 0.13333333333333333,
                                          0.11764705882352941, 0.10526315789473684, 0.09523809523809523,
                                          0.08695652173913043};
static __constant__ double SWM_SIN[9] = {6.283185307179586, -41.34170224039976, 81.60524927607506, -76.70585975306139,
                                         42.058693944897655, -15.09464257682299, 3.819952584848282,
                                         -0.7181223017785006, 0.10422916220813984};
static __constant__ double SWM_COS[10] = {1.0, -19.739208802178716, 64.9393940226683, -85.45681720669373,
                                          60.24464137187666, -26.4262567833744, 7.903536371318469,
                                          -1.714390711088672, 0.28200596845579123, -0.03638284114254567};
static __constant__ double SWM_K[4] = {0.99999999999999988897769753748 /* 1 - 2^-53 */, 0.6931471805599453 /* ln2 hi */,
                                       2.3190468138462996e-17 /* ln2 lo */, 0.5};


// 1/d and sqrt(x) for normal positive arguments: MUFU seed + Newton steps, no special-case branches.
__device__ __forceinline__ double swm_rcp(double d) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
  double e = fma(-d, r, 1.0);   // seed: ~2^-20 relative
  r = fma(r, e, r);
  e = fma(-d, r, 1.0);
  r = fma(r, e, r);
  e = fma(-d, r, 1.0);
  return fma(r, e, r);
}
// n / d, correctly rounded except in rare double-rounding cases (<= 1 ulp always)
__device__ __forceinline__ double swm_div(double n, double d) {
  const double r = swm_rcp(d);
  double q = n * r;
  const double rem = fma(-d, q, n);
  return fma(rem, r, q);
}
// n * (1/d) with a residual correction, r = swm_rcp(d) precomputed
__device__ __forceinline__ double swm_div_r(double n, double d, double r) {
  double q = n * r;
  const double rem = fma(-d, q, n);
  return fma(rem, r, q);
}
__device__ __forceinline__ double swm_sqrt(double x) {
  double r;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  double g = x * r, h = 0.5 * r;
  double e = fma(-h, g, 0.5);
  g = fma(g, e, g); h = fma(h, e, h);
  e = fma(-h, g, 0.5);
  g = fma(g, e, g); h = fma(h, e, h);
  e = fma(-h, g, 0.5);
  g = fma(g, e, g); h = fma(h, e, h);
  const double d = fma(-g, g, x);
  return fma(d, h, g);
}

// u = ((r >> 12) + 1/2) 2^-52, bit-identical to sw_u01 / the oracle's u01, without an int->double convert
__device__ __forceinline__ double swm_u01(u64 r) {
  const double d = __longlong_as_double((long long)((r >> 12) | 0x3FF0000000000000ULL));  // [1, 2)
  return d - SWM_K[0];  // 1 - 2^-53 (exact); the difference is exact
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
  const double s = swm_div(f, 2.0 + f);
  const double z = s * s;
  double p = SWM_LOG[10];
#pragma unroll
  for (int k = 9; k >= 0; --k) p = fma(p, z, SWM_LOG[k]);
  // log m = 2 s + s z p ;  log u = e ln2 + log m
  const double ed = (double)e;
  const double r = fma(s * z, p, ed * SWM_K[2]);
  return fma(ed, SWM_K[1], fma(2.0, s, r));
}

// (sin, cos)(2 pi u) for the Philox word r, u = ((r >> 12) + 1/2) 2^-52
__device__ __forceinline__ void swm_sincos2pi(u64 r, double &sn, double &cs) {
  const u64 k = r >> 12;                               // 52 bits
  const unsigned q = (unsigned)((k + (1ULL << 49)) >> 50);  // rint(4 u) in 0..4
  // t = u - q/4 in [-1/8, 1/8], exact
  const double u = swm_u01(r);
  const double t = fma((double)q, -0.25, u);
  const double z = t * t;
  double ps = SWM_SIN[8];
#pragma unroll
  for (int k = 7; k >= 0; --k) ps = fma(ps, z, SWM_SIN[k]);
  ps *= t;
  double pc = SWM_COS[9];
#pragma unroll
  for (int k = 8; k >= 0; --k) pc = fma(pc, z, SWM_COS[k]);
  // quadrant: angle = q pi/2 + 2 pi t
  const bool swap = q & 1u;
  double s0 = swap ? pc : ps, c0 = swap ? ps : pc;
  if ((q + 1u) & 2u) c0 = -c0;   // q = 1, 2: cos negative
  if (q & 2u) s0 = -s0;          // q = 2, 3: sin negative
  sn = s0; cs = c0;
}

// (sin, cos)(theta) of the 2-D director.  theta / (2 pi) with a two-term reciprocal (product error recovered by one
// fma), quadrant by the round-to-nearest magic constant, then the polynomials of swm_sincos2pi on t in [-1/8, 1/8]
// turns: <= 1.2e-16 absolute against mpmath for |theta| <= 1e6 (tools/gen_coeffs.py), no slow-path call in the hot
// loop.  Beyond 2^30 rad the library routine takes over.
__device__ __forceinline__ void swm_sincos(double th, double &sn, double &cs) {
  if (!(fabs(th) < 1073741824.0)) { sincos(th, &sn, &cs); return; }
  const double c_hi = 0.15915494309189535, c_lo = -9.839338337591243e-18;  // 1 / (2 pi)
  const double turns = th * c_hi;
  const double rem = fma(th, c_hi, -turns) + th * c_lo;
  const double big = 6755399441055744.0;  // 1.5 * 2^52: (x + big) - big = rint(x)
  const double qs = fma(4.0, turns, big);
  const double qd = qs - big;
  const unsigned q = (unsigned)__double2loint(qs);
  const double t = fma(qd, -0.25, turns) + rem;
  const double z = t * t;
  double ps = SWM_SIN[8];
#pragma unroll
  for (int k = 7; k >= 0; --k) ps = fma(ps, z, SWM_SIN[k]);
  ps *= t;
  double pc = SWM_COS[9];
#pragma unroll
  for (int k = 8; k >= 0; --k) pc = fma(pc, z, SWM_COS[k]);
  const bool swap = q & 1u;
  double s0 = swap ? pc : ps, c0 = swap ? ps : pc;
  if ((q + 1u) & 2u) c0 = -c0;
  if (q & 2u) s0 = -s0;
  sn = s0; cs = c0;
}

// four uniforms in (-1/2, 1/2): the Langevin thermostat's noise (oracle: orc_uniforms)
__device__ __forceinline__ void swm_uniforms(u64 seed, u64 id, u64 step, u64 stream, double u[4]) {
  u64 r[4];
  sw_noise_words(seed, id, step, stream, r);
#pragma unroll
  for (int k = 0; k < 4; ++k) u[k] = swm_u01(r[k]) - 0.5;
}

// ---------------------------------------------------------------------------------------------
// The normals of a 2-D Brownian step (oracle: orc_bd2d_normals).  A step consumes three: two for the translation, one for
// the rotation -- but a Box-Muller pair yields two, so a draw of four would throw one away every step.  Instead
//   translation of step s : pair 0 (block 0) of the draw (seed, id, s, stream 0):       g0 = m0 cos, g1 = m0 sin
//   rotation of step s    : pair 1 (block 1) of the draw of the EVEN step s & ~1:       its sine on the even step,
//                                                                                       its cosine on the odd one
// (sine and cosine of one pair are independent standard normals), so two steps cost three blocks and three pairs
// instead of four and four: ~60 of a step's ~700 instructions.  A kernel that keeps the cosine from the even step
// (k_bd2d_tiled, k_bd2d_env) takes the short path on the odd step; one that does not simply draws block 1 of step s - 1
// again -- the values are the same.  The key schedule of the seed is precomputed (DevParams.rk, a constant-bank operand).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void swm_philox_xyt(const unsigned int *rk, u64 id, u64 step, u64 r[4]) {
  sw_noise_block_rk(rk, id, step, 0ULL, 0u, r[0], r[1]);
  sw_noise_block_rk(rk, id, step & ~1ULL, 0ULL, 1u, r[2], r[3]);
}
// the three normals from words drawn earlier (the resident kernel draws the words of step s+1 in the shadow of step s's
// neighbour loads and transforms them in the shadow of the grid barrier)
__device__ __forceinline__ void swm_boxmuller_xyt(const u64 r[4], bool odd, double &g0, double &g1, double &g3) {
  const double m0 = swm_sqrt(-2.0 * swm_log01(swm_u01(r[0])));
  const double m1 = swm_sqrt(-2.0 * swm_log01(swm_u01(r[2])));
  double s0, c0, s1, c1;
  swm_sincos2pi(r[1], s0, c0);
  swm_sincos2pi(r[3], s1, c1);
  g0 = m0 * c0; g1 = m0 * s0; g3 = m1 * (odd ? c1 : s1);
}
__device__ __forceinline__ void swm_gaussians_xyt(const unsigned int *rk, u64 id, u64 step, double &g0, double &g1, double &g3) {
  u64 r[4];
  swm_philox_xyt(rk, id, step, r);
  swm_boxmuller_xyt(r, (step & 1ULL) != 0ULL, g0, g1, g3);
}

// swm_gaussians_xyt for a kernel that keeps the even step's cosine normal in memory (k_bd2d_stream: one slot per particle):
// keep_ok = the slot was written by the step just before `step`.  Same values, bit for bit.
__device__ __forceinline__ void swm_gaussians_xyt_keep(const unsigned int *rk, u64 id, u64 step, bool keep_ok, double *slot,
                                                       double &g0, double &g1, double &g3) {
#ifdef SW_STREAM_NO_KEEP
  if (true) { swm_gaussians_xyt(rk, id, step, g0, g1, g3); return; }   // A/B: every step draws both blocks, no slot traffic
#endif
  if (keep_ok && (step & 1ULL)) {
    u64 w0, w1;
    sw_noise_block_rk(rk, id, step, 0ULL, 0u, w0, w1);
    const double m0 = swm_sqrt(-2.0 * swm_log01(swm_u01(w0)));
    double s0, c0;
    swm_sincos2pi(w1, s0, c0);
    g0 = m0 * c0; g1 = m0 * s0; g3 = *slot;
  } else {
    u64 r[4];
    swm_philox_xyt(rk, id, step, r);
    const double m0 = swm_sqrt(-2.0 * swm_log01(swm_u01(r[0])));
    const double m1 = swm_sqrt(-2.0 * swm_log01(swm_u01(r[2])));
    double s0, c0, s1, c1;
    swm_sincos2pi(r[1], s0, c0);
    swm_sincos2pi(r[3], s1, c1);
    const double gs = m1 * s1, gc = m1 * c1;
    g0 = m0 * c0; g1 = m0 * s0;
    if (step & 1ULL) g3 = gc;
    else { g3 = gs; *slot = gc; }
  }
}

// ---------------------------------------------------------------------------------------------
// Everything a 2-D Brownian step needs that depends on the particle alone, in ONE pass: the three normals of
// (key schedule, id, step) and the director (sin, cos)(theta).  Same operations, value by value, as
// swm_gaussians_xyt + swm_sincos (every product / fma is explicit, so the bits are the same) -- but the two logarithms
// and the three sine / cosine polynomials are evaluated side by side: a coefficient is fetched from the constant bank
// once for all chains instead of once per chain (70 constant loads per warp-step in the separate version, ncu
// profiles/r02_ncu_k_bd2d_tiled_flag_version_summary.txt: LDC + LDCU), and the independent chains interleave.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void swm_quadrant(unsigned q, double ps, double pc, double &sn, double &cs) {
  const bool swap = q & 1u;
  double s0 = swap ? pc : ps, c0 = swap ? ps : pc;
  if ((q + 1u) & 2u) c0 = -c0;
  if (q & 2u) s0 = -s0;
  sn = s0; cs = c0;
}

// SHORT = the odd step of a pair whose cosine normal `g3_keep` was kept from the even step: block 0 and ONE Box-Muller
// pair only.  Otherwise both blocks: g3 = the sine (even step) or the cosine (odd step) of pair 1 of the even step's draw,
// and g3_keep receives the cosine for the odd step that follows an even one.
template <bool SHORT>
__device__ __forceinline__ void swm_step_inputs(const unsigned int *rk, u64 id, u64 step, double th, double &g0, double &g1,
                                                double &g3, double &g3_keep, double &sn, double &cs) {
  constexpr int NP = SHORT ? 1 : 2;   // Box-Muller pairs evaluated
  u64 r[4] = {0ULL, 0ULL, 0ULL, 0ULL};
  sw_noise_block_rk(rk, id, step, 0ULL, 0u, r[0], r[1]);
  if constexpr (!SHORT) sw_noise_block_rk(rk, id, step & ~1ULL, 0ULL, 1u, r[2], r[3]);
  // ---- arguments of the (sin, cos) pairs: t in [-1/8, 1/8] turns, quadrant q; chain NP is the director ---------------
  double t[NP + 1];
  unsigned q[NP + 1];
#pragma unroll
  for (int i = 0; i < NP; ++i) {
    const u64 k = r[2 * i + 1] >> 12;
    q[i] = (unsigned)((k + (1ULL << 49)) >> 50);
    t[i] = fma((double)q[i], -0.25, swm_u01(r[2 * i + 1]));
  }
  const bool big = !(fabs(th) < 1073741824.0);   // beyond 2^30 rad the library routine takes over (swm_sincos)
  {
    const double c_hi = 0.15915494309189535, c_lo = -9.839338337591243e-18;
    const double turns = th * c_hi;
    const double rem = fma(th, c_hi, -turns) + th * c_lo;
    const double bigc = 6755399441055744.0;
    const double qs = fma(4.0, turns, bigc);
    const double qd = qs - bigc;
    q[NP] = (unsigned)__double2loint(qs);
    t[NP] = big ? 0.0 : fma(qd, -0.25, turns) + rem;
  }
  // ---- polynomials, coefficient by coefficient ---------------------------------------------------------------------
  double z[NP + 1], ps[NP + 1], pc[NP + 1];
#pragma unroll
  for (int i = 0; i <= NP; ++i) { z[i] = t[i] * t[i]; ps[i] = SWM_SIN[8]; pc[i] = SWM_COS[9]; }
#pragma unroll
  for (int k = 7; k >= 0; --k) {
    const double c = SWM_SIN[k];
#pragma unroll
    for (int i = 0; i <= NP; ++i) ps[i] = fma(ps[i], z[i], c);
  }
#pragma unroll
  for (int i = 0; i <= NP; ++i) ps[i] *= t[i];
#pragma unroll
  for (int k = 8; k >= 0; --k) {
    const double c = SWM_COS[k];
#pragma unroll
    for (int i = 0; i <= NP; ++i) pc[i] = fma(pc[i], z[i], c);
  }
  // ---- the logarithms (swm_log01), side by side ----------------------------------------------------------------------
  double lg[NP];
  {
    double s2[NP], zz[NP], p[NP], ed[NP];
#pragma unroll
    for (int i = 0; i < NP; ++i) {
      const double u = swm_u01(r[2 * i]);
      int hi = __double2hiint(u), lo = __double2loint(u);
      int e = (hi >> 20) - 1023;
      hi = (hi & 0x000FFFFF) | 0x3FF00000;
      if (hi >= 0x3FF6A09F) { hi -= 0x00100000; e += 1; }
      const double f = __hiloint2double(hi, lo) - 1.0;
      s2[i] = swm_div(f, 2.0 + f);
      zz[i] = s2[i] * s2[i];
      p[i] = SWM_LOG[10];
      ed[i] = (double)e;
    }
#pragma unroll
    for (int k = 9; k >= 0; --k) {
      const double c = SWM_LOG[k];
#pragma unroll
      for (int i = 0; i < NP; ++i) p[i] = fma(p[i], zz[i], c);
    }
#pragma unroll
    for (int i = 0; i < NP; ++i) {
      const double rr = fma(s2[i] * zz[i], p[i], ed[i] * SWM_K[2]);
      lg[i] = fma(ed[i], SWM_K[1], fma(2.0, s2[i], rr));
    }
  }
  const double m0 = swm_sqrt(-2.0 * lg[0]);
  double s0, c0;
  swm_quadrant(q[0], ps[0], pc[0], s0, c0);
  g0 = m0 * c0; g1 = m0 * s0;
  if constexpr (SHORT) {
    g3 = g3_keep;
  } else {
    const double m1 = swm_sqrt(-2.0 * lg[NP - 1]);
    double s1, c1;
    swm_quadrant(q[NP - 1], ps[NP - 1], pc[NP - 1], s1, c1);
    const double gs = m1 * s1, gc = m1 * c1;
    g3 = (step & 1ULL) ? gc : gs;
    g3_keep = gc;
  }
  if (big) sincos(th, &sn, &cs);
  else swm_quadrant(q[NP], ps[NP], pc[NP], sn, cs);
}

// all four normals with the fast kernels (parity hook / 3-D)
__device__ __forceinline__ void swm_gaussians(u64 seed, u64 id, u64 step, u64 stream, double g[4]) {
  u64 r[4];
  sw_noise_words(seed, id, step, stream, r);
  const double m0 = swm_sqrt(-2.0 * swm_log01(swm_u01(r[0])));
  const double m1 = swm_sqrt(-2.0 * swm_log01(swm_u01(r[2])));
  double s0, c0, s1, c1;
  swm_sincos2pi(r[1], s0, c0);
  swm_sincos2pi(r[3], s1, c1);
  g[0] = m0 * c0; g[1] = m0 * s0; g[2] = m1 * c1; g[3] = m1 * s1;
}
