// swarm_bd2d.cuh -- k_bd2d_resident, THE hot kernel of the 2-D rollout (BASELINE configs 2-4).
// Included by swarm_engine.cu after the rebuild phases and the grid barrier.
//
// Cooperative persistent chunk of Brownian steps, ONE PARTICLE PER THREAD, one CTA per SM.  A particle's
// state (x, y, theta, director, rebuild reference, swim force, torque, first Verlet entries) stays in
// registers for the whole chunk; per step a thread reads only its neighbours' positions (L2, ld.cg) and
// writes its own new position for them.  theta, v and omega go to memory once, at the end.
//
// Per step the only grid-wide dependency is "neighbours' positions after step s".  Everything else of
// step s+1 -- the Philox + Box-Muller noise (seed, i, s+1), the director sincos(theta_{s+1}), the wall
// and swim forces -- depends on the particle alone and is issued BETWEEN barrier arrive and barrier
// wait, so that work (about 3/4 of a step's instructions) hides the barrier's L2 round trips.  Warp 0
// of every CTA owns no particles: its lane 0 does the arrive / poll, so no particle warp sits behind it.
//
// Arithmetic per particle is identical to phase_step<2, NOISE> (the oracle's orc_bd_steps), except
// that F / gamma is computed as a reciprocal multiply with one residual correction (<= 1 ulp).
#pragma once

#ifdef SW_TIMING
#define SW_TICK(slot) do { if (prof) { long long t_ = clock64(); tp[slot] += t_ - tc; tc = t_; } } while (0)
#else
#define SW_TICK(slot) do { } while (0)
#endif

// sw_wca_factor with ONE reciprocal (MUFU seed + Newton, no slow-path branch) shared by its two divisions, each
// quotient corrected by its residual: <= 1 ulp from the IEEE divisions of the generic path and the oracle, ~20
// instead of ~85 instructions.  It matters because a warp runs this branch whenever ANY of its lanes is in contact,
// which in a clustering active swarm is almost always.
// Every product and sum below is spelled out (__dmul_rn / fma): nvcc otherwise decides per call site which
// multiply of an expression it fuses into an FMA, and the four kernels that inline these helpers (resident, tiled,
// stream, env) must produce the same bits for the same particle.
__device__ __forceinline__ double swm_wca_factor(double r2, double sigma, double eps) {
  const double r = swm_rcp(r2);
  const double f2 = swm_div_r(__dmul_rn(sigma, sigma), r2, r);
  const double f6 = __dmul_rn(__dmul_rn(f2, f2), f2);
  return swm_div_r(__dmul_rn(__dmul_rn(__dmul_rn(48.0, eps), f6), __dsub_rn(f6, 0.5)), r2, r);
}

// pair force of one Verlet entry on particle (xi, yi) of type ti
__device__ __forceinline__ void bd2d_pair(const DevParams &P, double xi, double yi, int ti, int tj, double xj, double yj,
                                          double &Fx, double &Fy) {
  double dx = sw_min_image(xi - xj, P.box[0], P.inv_box[0], P.periodic[0]);
  double dy = sw_min_image(yi - yj, P.box[1], P.inv_box[1], P.periodic[1]);
  double r2 = sw_r2(dx, dy, 0.0);
  double cut = P.wca_cut[ti][tj];
  if (r2 < __dmul_rn(cut, cut) && r2 > 0.0) {
    double fac = swm_wca_factor(r2, P.wca_sigma[ti][tj], P.wca_eps[ti][tj]);
    Fx = fma(fac, dx, Fx); Fy = fma(fac, dy, Fy);
  }
}

// bd2d_pair for an entry that knows whether its separation needs the minimum image (k_bd2d_tiled, SW_TILE_WRAP): for a zero
// image shift the minimum image returns d - L * 0 = d, so skipping it leaves every bit as it is and saves two
// multiplies, a rounding and a subtraction per dimension and entry.
__device__ __forceinline__ void bd2d_pair_w(const DevParams &P, double xi, double yi, int ti, int tj, bool wrap, double xj, double yj,
                                            double &Fx, double &Fy) {
  double dx = xi - xj, dy = yi - yj;
  if (wrap) {
    dx = sw_min_image(dx, P.box[0], P.inv_box[0], P.periodic[0]);
    dy = sw_min_image(dy, P.box[1], P.inv_box[1], P.periodic[1]);
  }
  double r2 = sw_r2(dx, dy, 0.0);
  double cut = P.wca_cut[ti][tj];
  if (r2 < __dmul_rn(cut, cut) && r2 > 0.0) {
    double fac = swm_wca_factor(r2, P.wca_sigma[ti][tj], P.wca_eps[ti][tj]);
    Fx = fma(fac, dx, Fx); Fy = fma(fac, dy, Fy);
  }
}

// -DSW_TILE_PAIR4 (A/B build, NOT the default; profiles/r02_pair4_ab.txt: 0.42-0.59 ms per launch against 0.385-0.494 ms
// with bd2d_pair -- the step is short of issue slots, and a warp with ONE contact now issues four chains).
// Four Verlet entries at once.  bd2d_pair evaluates entry after entry, each behind its own branch: a warp then runs the WCA
// chain (reciprocal + Newton steps + two corrected quotients, ~18 dependent fp64 operations, ~160 cycles) once for every
// entry index at which ANY of its lanes is in contact -- two to three times per step in a clustering swarm, one after the
// other, with only ~6 warps per scheduler to cover the latency.  Here the four cut-off tests come first, then ONE branch
// (any of my four entries in contact) guards four INDEPENDENT chains that the scheduler interleaves; an entry that is
// not in contact runs its chain on r2 = 1 and is not accumulated.  The accumulation keeps the entry order and touches
// the same values with the same operations as bd2d_pair, so the force is the same bit for bit.
__device__ __forceinline__ void bd2d_pair4(const DevParams &P, double xi, double yi, int ti, const int en[4], const double xj[4],
                                           const double yj[4], int nvalid, double &Fx, double &Fy) {
  double dx[4], dy[4], r2[4];
  bool in[4];
  int tj[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    tj[k] = en[k] >> SW_TYPE_SHIFT;
    dx[k] = sw_min_image(xi - xj[k], P.box[0], P.inv_box[0], P.periodic[0]);
    dy[k] = sw_min_image(yi - yj[k], P.box[1], P.inv_box[1], P.periodic[1]);
    r2[k] = sw_r2(dx[k], dy[k], 0.0);
    const double cut = P.wca_cut[ti][tj[k]];
    in[k] = (k < nvalid) && r2[k] < __dmul_rn(cut, cut) && r2[k] > 0.0;
  }
  if (in[0] || in[1] || in[2] || in[3]) {
    double fac[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) fac[k] = swm_wca_factor(in[k] ? r2[k] : 1.0, P.wca_sigma[ti][tj[k]], P.wca_eps[ti][tj[k]]);
#pragma unroll
    for (int k = 0; k < 4; ++k)
      if (in[k]) { Fx = fma(fac[k], dx[k], Fx); Fy = fma(fac[k], dy[k], Fy); }
  }
}

// forces that depend on the particle alone: walls + swim + external (added AFTER the pair forces,
// in the oracle's order: walls, swim, external)
struct Bd2dSelf { double wx, wy, sx, sy; };
__device__ __forceinline__ Bd2dSelf bd2d_self_forces(const DevParams &P, int ti, double xi, double yi, double ex, double ey,
                                                    double fsw) {
  Bd2dSelf o;
  o.wx = 0.0; o.wy = 0.0;
  if (P.walls) {
    const double cutw = P.wall_cut[ti];
    // two separate accumulations per axis, in the oracle's order (lower wall, upper wall)
    const double ws = P.wall_sigma[ti], we = P.wall_eps[ti];
    if (xi < cutw && xi > 0.0) o.wx = fma(swm_wca_factor(__dmul_rn(xi, xi), ws, we), xi, o.wx);
    double d2 = P.box[0] - xi;
    if (d2 < cutw && d2 > 0.0) o.wx = fma(-swm_wca_factor(__dmul_rn(d2, d2), ws, we), d2, o.wx);
    if (yi < cutw && yi > 0.0) o.wy = fma(swm_wca_factor(__dmul_rn(yi, yi), ws, we), yi, o.wy);
    d2 = P.box[1] - yi;
    if (d2 < cutw && d2 > 0.0) o.wy = fma(-swm_wca_factor(__dmul_rn(d2, d2), ws, we), d2, o.wy);
  }
  o.sx = __dmul_rn(fsw, ex); o.sy = __dmul_rn(fsw, ey);
  return o;
}

// The Brownian update of one particle from its total force (phase_step<2>, oracle: orc_bd_steps): velocity F / gamma
// (per body axis when the friction is anisotropic), Euler-Maruyama step of position and angle.
struct Bd2dConst { double gt1, gt2, r_gt1, r_gt2, sg_par, sg_per, sg_rot; };
template <bool NOISE>
__device__ __forceinline__ void bd2d_update(bool iso, const Bd2dConst &c, double dt, double Fx, double Fy, double ex, double ey,
                                            double w, double g0, double g1, double g3, double &x, double &y, double &th,
                                            double &vx, double &vy) {
  if (iso) {
    vx = swm_div_r(Fx, c.gt2, c.r_gt2); vy = swm_div_r(Fy, c.gt2, c.r_gt2);
    x = fma(vx, dt, x); y = fma(vy, dt, y);
    if (NOISE) { x = fma(c.sg_par, g0, x); y = fma(c.sg_par, g1, y); }
  } else {
    const double fpar = fma(Fx, ex, __dmul_rn(Fy, ey)), fper = fma(Fy, ex, -__dmul_rn(Fx, ey));
    const double vpar = swm_div_r(fpar, c.gt2, c.r_gt2), vper = swm_div_r(fper, c.gt1, c.r_gt1);
    vx = fma(vpar, ex, -__dmul_rn(vper, ey)); vy = fma(vpar, ey, __dmul_rn(vper, ex));
    x = fma(vx, dt, x); y = fma(vy, dt, y);
    if (NOISE) {
      const double spar = __dmul_rn(c.sg_par, g0), sper = __dmul_rn(c.sg_per, g1);
      x += fma(spar, ex, -__dmul_rn(sper, ey)); y += fma(spar, ey, __dmul_rn(sper, ex));
    }
  }
  th = fma(w, dt, th);
  if (NOISE) th = fma(c.sg_rot, g3, th);
}

// PIPE = 1 (default): the Philox words of step s+1 are drawn between the ISSUE of step s's neighbour loads and their
// first use, and turned into normals (Box-Muller) behind the grid barrier -- so both waits of a step, the L2 round trip
// of the neighbour positions and the barrier, are covered by work that depends on the particle alone.  PIPE = 0 keeps
// the whole draw behind the barrier (round 1's order; SWARM_B200_RES_PIPE=0 selects it for A/B runs).
template <bool NOISE, int PIPE>
__global__ void __launch_bounds__(SW_RES_MAXT, 1) k_bd2d_resident(const __grid_constant__ DevParams P, DevArrays A, RunArgs R) {
  Ctl *ctl = A.ctl;
  if (ldcg(&ctl->halted)) return;
  unsigned int epoch = 0;
  const int ptid = (int)threadIdx.x - 32;  // warp 0 is the barrier warp
  const int ppc = R.particles_per_cta;     // particles are spread evenly over the CTAs (one CTA per SM)
  const long long i = (long long)blockIdx.x * ppc + ptid;
  const bool act = ptid >= 0 && ptid < ppc && i < sw_live_n(A);
  const bool leader = act && i == 0;
  double xi = 0.0, yi = 0.0, th = 0.0, xr = 0.0, yr = 0.0, fsw = 0.0, Tz = 0.0, fxe = 0.0, fye = 0.0;
  int sp = 0, ti = 0;
  u64 pid = 0;
  if (act) {
    pid = sw_pid(A, i);
    xi = A.x[0][i]; yi = A.y[0][i]; th = A.theta[i];
    xr = A.xr[i]; yr = A.yr[i];
    fsw = A.f_swim[i]; Tz = A.tqz[i];
    if (A.fex) { fxe = A.fex[i]; fye = A.fey[i]; }
    sp = A.species[i]; ti = P.sp_type[sp];
  }
  const double lim2 = (0.5 * P.skin) * (0.5 * P.skin);
  const double dt = P.dt, kT = P.kT;
  const double gt1 = P.sp_gt[sp][1], gt2 = P.sp_gt[sp][2], gr0 = P.sp_gr[sp][0];
  const bool iso = (gt1 == gt2);
  const double r_gt1 = swm_rcp(gt1), r_gt2 = swm_rcp(gt2), r_gr0 = swm_rcp(gr0);
  const double sg_par = NOISE ? sqrt(2.0 * kT * dt / gt2) : 0.0;
  const double sg_per = NOISE ? sqrt(2.0 * kT * dt / gt1) : 0.0;
  const double sg_rot = NOISE ? sqrt(2.0 * kT * dt / gr0) : 0.0;
  const double w = swm_div_r(Tz, gr0, r_gr0);  // omega_z: the torque is constant over the chunk

  __shared__ int s_need;
  __shared__ volatile int s_moved;  // some particle of this CTA went beyond skin/2 in the current step
  if (threadIdx.x == 0) s_moved = 0;
  unsigned int movers_seen = 0;  // thread 0: high word of the barrier counter at the last poll
  bool need = true;
  if (!R.force_rebuild) {
    bool moved = false;
    if (act) {
      double ddx = xi - xr, ddy = yi - yr;
      moved = !(ddx * ddx + ddy * ddy <= lim2);
    }
    const int any = __syncthreads_or(moved);
    grid_arrive64(&ctl->bar64, epoch, any != 0);
    if (threadIdx.x == 0) {
      const unsigned int mv = (unsigned int)(grid_poll64(&ctl->bar64, epoch) >> 32);
      s_need = (mv != movers_seen);
      movers_seen = mv;
    }
    __syncthreads();
    need = s_need != 0;
  }
  // particle-local inputs of the first step
  double g0 = 0.0, g1 = 0.0, g3 = 0.0, ex = 1.0, ey = 0.0;
  Bd2dSelf sf = {0.0, 0.0, 0.0, 0.0};
  if (act) {
    swm_sincos(th, ey, ex);
    sf = bd2d_self_forces(P, ti, xi, yi, ex, ey, fsw);
    if (NOISE) swm_gaussians_xyt(P.rk, pid, R.step0, g0, g1, g3);
  }
  u64 rw[4] = {0, 0, 0, 0};   // PIPE: Philox words of the next step, carried from the load shadow to the barrier shadow
  int cnt = 0, nb[SW_RES_NBR];
  const int self_entry = act ? (int)i : 0;
#pragma unroll
  for (int k = 0; k < SW_RES_NBR; ++k) nb[k] = self_entry;
  if (act && !need) {
    cnt = A.nbr_cnt[i];
#pragma unroll
    for (int k = 0; k < SW_RES_NBR; ++k) nb[k] = (k < cnt) ? A.nbr[(long long)k * A.n_pad + i] : self_entry;
  }
  double vx_l = 0.0, vy_l = 0.0;  // velocity of the most recent step (materialised after the loop)
#ifdef SW_TIMING
  long long tp[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tc = clock64();
  const bool prof = (blockIdx.x == 0) && (threadIdx.x == 32 || threadIdx.x == 0);
#endif

  long long s = 0;
  for (; s < R.n_steps; ++s) {
    const int cur = (int)(s & 1), nxt = cur ^ 1;
    if (need) {
      if (act) phase_clear<2>(P, A, cur, i);
      grid_barrier64(&ctl->bar64, epoch);
      if (act) phase_insert(A, i);
      grid_barrier64(&ctl->bar64, epoch);
      if (act) {
        phase_search<2>(P, A, cur, i);
        xr = xi; yr = yi;
        cnt = A.nbr_cnt[i];
#pragma unroll
        for (int k = 0; k < SW_RES_NBR; ++k) nb[k] = (k < cnt) ? A.nbr[(long long)k * A.n_pad + i] : self_entry;
      }
      if (leader) ctl->rebuilds += 1;
      grid_barrier64(&ctl->bar64, epoch);
      if (ldcg(&ctl->overflow)) break;  // uniform: every CTA reads the same value after the barrier
    }
    SW_TICK(0);
    const bool last = (s == R.n_steps - 1);
    bool moved = false;
    if (act) {
      // ---- the part that needs the neighbours: WCA over the Verlet list (ascending j) ------------
      // register-resident entries are loaded unconditionally (unused slots point at the particle
      // itself) so the L2 round trips overlap
      double xj[SW_RES_NBR], yj[SW_RES_NBR];
#pragma unroll
      for (int k = 0; k < SW_RES_NBR; ++k) {
        const int j = nb[k] & SW_IDX_MASK;
        xj[k] = ldcg(&A.x[cur][j]); yj[k] = ldcg(&A.y[cur][j]);
      }
      // ---- shadow of those loads: the next step's Philox block (integer pipe, depends on (seed, id, step) only) ----
      if (NOISE && PIPE && !last) swm_philox_xyt(P.rk, pid, R.step0 + (u64)(s + 1), rw);
      double Fx = 0.0, Fy = 0.0;
#pragma unroll
      for (int k = 0; k < SW_RES_NBR; ++k)
        if (k < cnt) bd2d_pair(P, xi, yi, ti, nb[k] >> SW_TYPE_SHIFT, xj[k], yj[k], Fx, Fy);
      // list entries beyond the register-resident ones (clustered particles): four at a time, indices first and
      // then positions, so a particle with 8 neighbours costs two dependent L2 round trips instead of eight -- one
      // such particle used to hold its whole CTA, and through the barrier the whole grid (the slice slowed from
      // 0.58 to 0.70 ms as the active swarm clustered)
      for (int k0 = SW_RES_NBR; k0 < cnt; k0 += 4) {
        int en4[4];
        double xe[4], ye[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) en4[u] = (k0 + u < cnt) ? A.nbr[(long long)(k0 + u) * A.n_pad + i] : self_entry;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int j = en4[u] & SW_IDX_MASK;
          xe[u] = ldcg(&A.x[cur][j]); ye[u] = ldcg(&A.y[cur][j]);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (k0 + u < cnt) bd2d_pair(P, xi, yi, ti, en4[u] >> SW_TYPE_SHIFT, xe[u], ye[u], Fx, Fy);
      }
      Fx += sf.wx; Fy += sf.wy;
      Fx += sf.sx; Fy += sf.sy;
      if (A.fex) { Fx += fxe; Fy += fye; }
      // ---- Brownian update (phase_step<2>) ------------------------------------------------------
      const Bd2dConst cc = {gt1, gt2, r_gt1, r_gt2, sg_par, sg_per, sg_rot};
      bd2d_update<NOISE>(iso, cc, dt, Fx, Fy, ex, ey, w, g0, g1, g3, xi, yi, th, vx_l, vy_l);
      const double xn = xi, yn = yi;
      A.x[nxt][i] = xn; A.y[nxt][i] = yn;
      double ddx = xn - xr, ddy = yn - yr;
      moved = (ddx * ddx + ddy * ddy) > lim2;
    }
    SW_TICK(1);
    if (threadIdx.x < 32) {
      // ---- barrier warp: waits for the particle warps' stores (named barrier 1: they only ARRIVE, so a warp
      // that is done goes straight on to its local work instead of waiting for the slowest warp of the CTA),
      // then arrives at the grid barrier and polls; publishes the rebuild decision -------------------------
      asm volatile("barrier.cta.sync 1, %0;" ::"r"((int)blockDim.x) : "memory");
      SW_TICK(2);
      int any = 0;
      if (threadIdx.x == 0) { any = s_moved; s_moved = 0; }
      grid_arrive64(&ctl->bar64, epoch, any != 0);
      SW_TICK(6);
      if (threadIdx.x == 0) {
        const unsigned int mv = (unsigned int)(grid_poll64(&ctl->bar64, epoch) >> 32);
        s_need = (mv != movers_seen);
        movers_seen = mv;
      }
      SW_TICK(7);
    } else {
      if (moved) s_moved = 1;
      asm volatile("barrier.cta.arrive 1, %0;" ::"r"((int)blockDim.x) : "memory");
      SW_TICK(2);
      if (act && !last) {
        // ---- particle-local work of step s+1, overlapping the barrier ----------------------------
        swm_sincos(th, ey, ex);
        sf = bd2d_self_forces(P, ti, xi, yi, ex, ey, fsw);
        if (NOISE) {
          if (PIPE) swm_boxmuller_xyt(rw, ((R.step0 + (u64)(s + 1)) & 1ULL) != 0ULL, g0, g1, g3);
          else swm_gaussians_xyt(P.rk, pid, R.step0 + (u64)(s + 1), g0, g1, g3);
        }
      }
    }
    SW_TICK(4);
    __syncthreads();
    need = s_need != 0;
    SW_TICK(5);
  }
#ifdef SW_TIMING
  if (prof) {
    if (threadIdx.x == 0) { ctl->tprof[6] += tp[6]; ctl->tprof[7] += tp[7]; }
    else for (int k = 0; k < 6; ++k) ctl->tprof[k] += tp[k];
  }
#endif
  if (act) {
    A.x[0][i] = xi; A.y[0][i] = yi; A.theta[i] = th;
    if (s > 0) {
      // v and omega of the last completed step (stochastic parts: streams 2 and 3)
      double gv[4] = {0, 0, 0, 0}, gw[4] = {0, 0, 0, 0};
      double sv = 0.0, sw = 0.0;
      if (NOISE) {
        swm_gaussians(P.seed, pid, R.step0 + (u64)(s - 1), 2, gv);
        swm_gaussians(P.seed, pid, R.step0 + (u64)(s - 1), 3, gw);
        sv = sqrt(kT / P.sp_mass[sp]);
        sw = sqrt(kT / P.sp_rin[sp][0]);
      }
      A.vx[i] = vx_l + sv * gv[0]; A.vy[i] = vy_l + sv * gv[1]; A.vz[i] = 0.0;
      A.wx[i] = 0.0; A.wy[i] = 0.0; A.wz[i] = w + sw * gw[0];
    }
  }
  if (leader) {
    ctl->steps_done = R.steps_base + s;
    if (s < R.n_steps) ctl->halted = 1;
    ctl->rebuild_flag[0] = 0; ctl->rebuild_flag[1] = 0; ctl->rebuild_flag[2] = 0;
  }
}

// ---------------------------------------------------------------------------------------------
// k_bd2d_stream -- the 2-D Brownian chunk for swarms too large for one particle per thread (the
// 1M-particle slabs of the domain-decomposed swarm, BASELINE config 5).  Same loop as k_persist (one
// grid barrier per step, in-kernel Verlet rebuild), but a lean step: per-species constants hoisted to
// shared memory once per launch, the resident kernel's arithmetic (3-normal Philox draw, reciprocal
// multiplies), neighbour positions fetched four at a time so their L2 round trips overlap, and a
// register budget that keeps 3 CTAs (24 warps) per SM.
// ---------------------------------------------------------------------------------------------
struct Bd2dSpecies { double gt1, gt2, gr0, r_gt1, r_gt2, r_gr0, sg_par, sg_per, sg_rot, sv, sw; };

__device__ __forceinline__ void sw_prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

template <bool NOISE>
__device__ __forceinline__ bool bd2d_step_stream(const DevParams &P, const DevArrays &A, const Bd2dSpecies *S, int cur, u64 step,
                                                 bool last, long long i, long long i_next, double lim2, bool have_g,
                                                 double pg0, double pg1, double pg3, bool keep_ok) {
  const int nxt = cur ^ 1;
  // every per-particle input is requested up front: without __restrict__ the compiler keeps each load
  // behind the stores of the previous particle and next to its first use, which serialised four DRAM
  // round trips per particle (ncu: 44 % long-scoreboard stalls)
  const int sp = A.species[i];
  const int cnt = A.nbr_cnt[i];
  const double xi = ldcg(&A.x[cur][i]), yi = ldcg(&A.y[cur][i]);
  const double th = A.theta[i];
  const double fsw = A.f_swim[i], tqz = A.tqz[i];
  const double xr = A.xr[i], yr = A.yr[i];
  const u64 pid = sw_pid(A, i);
  double fxe = 0.0, fye = 0.0;
  if (A.fex) { fxe = A.fex[i]; fye = A.fey[i]; }
  if (i_next >= 0) {
    // pull the next particle of this thread towards L2 while this one computes
    sw_prefetch_l2(&A.x[cur][i_next]); sw_prefetch_l2(&A.y[cur][i_next]); sw_prefetch_l2(&A.theta[i_next]);
    sw_prefetch_l2(&A.f_swim[i_next]); sw_prefetch_l2(&A.tqz[i_next]); sw_prefetch_l2(&A.xr[i_next]); sw_prefetch_l2(&A.yr[i_next]);
    sw_prefetch_l2(&A.nbr_cnt[i_next]); sw_prefetch_l2(&A.species[i_next]);
    if (A.gid) sw_prefetch_l2(&A.gid[i_next]);
#pragma unroll
    for (int u = 0; u < 4; ++u) sw_prefetch_l2(&A.nbr[(long long)u * A.n_pad + i_next]);
  }
  const int ti = P.sp_type[sp];
  const Bd2dSpecies c = S[sp];
  double Fx = 0.0, Fy = 0.0;
  for (int k0 = 0; k0 < cnt; k0 += 4) {
    int en[4];
    double xj[4], yj[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      en[u] = (k0 + u < cnt) ? A.nbr[(long long)(k0 + u) * A.n_pad + i] : (int)i;
      const int j = en[u] & SW_IDX_MASK;
      xj[u] = ldcg(&A.x[cur][j]); yj[u] = ldcg(&A.y[cur][j]);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (k0 + u < cnt) bd2d_pair(P, xi, yi, ti, en[u] >> SW_TYPE_SHIFT, xj[u], yj[u], Fx, Fy);
  }
  double ex, ey;
  double g0 = pg0, g1 = pg1, g3 = pg3;  // drawn behind the previous step's barrier for the thread's first particle
#ifndef SW_STREAM_SEPARATE_INPUTS
  // normals and director in one pass (swm_step_inputs: the bits of swm_gaussians_xyt + swm_sincos, coefficient loads
  // shared between the chains): 8 M particles on one GPU 41.05 -> 39.31 ms per slice (profiles/r02_stream_old_new_ab.txt)
  if (NOISE && !have_g) {
#ifndef SW_STREAM_NO_KEEP
    // the cosine normal of the even step waits in the particle's slot of A.gkeep (8 M particles: 39.45 -> 37.34 ms per slice)
    double keep = 0.0;
    if (keep_ok && (step & 1ULL)) { keep = A.gkeep[i]; swm_step_inputs<true>(P.rk, pid, step, th, g0, g1, g3, keep, ey, ex); }
    else { swm_step_inputs<false>(P.rk, pid, step, th, g0, g1, g3, keep, ey, ex); if (!(step & 1ULL)) A.gkeep[i] = keep; }
#else
    double keep;
    swm_step_inputs<false>(P.rk, pid, step, th, g0, g1, g3, keep, ey, ex);
#endif
  }
  else swm_sincos(th, ey, ex);
#else
  swm_sincos(th, ey, ex);
  if (NOISE && !have_g) swm_gaussians_xyt_keep(P.rk, pid, step, keep_ok, &A.gkeep[i], g0, g1, g3);   // keep_ok: step - 1 ran in this launch
#endif
  const Bd2dSelf sf = bd2d_self_forces(P, ti, xi, yi, ex, ey, fsw);
  Fx += sf.wx; Fy += sf.wy;
  Fx += sf.sx; Fy += sf.sy;
  if (A.fex) { Fx += fxe; Fy += fye; }
  const double dt = P.dt;
  double xn = xi, yn = yi, thn = th, vx, vy;
  const double w = swm_div_r(tqz, c.gr0, c.r_gr0);
  const Bd2dConst cc = {c.gt1, c.gt2, c.r_gt1, c.r_gt2, c.sg_par, c.sg_per, c.sg_rot};
  bd2d_update<NOISE>(c.gt1 == c.gt2, cc, dt, Fx, Fy, ex, ey, w, g0, g1, g3, xn, yn, thn, vx, vy);
  A.theta[i] = thn;
  A.x[nxt][i] = xn; A.y[nxt][i] = yn;
  if (last) {
    double gv[4] = {0, 0, 0, 0}, gw[4] = {0, 0, 0, 0};
    if (NOISE) {
      swm_gaussians(P.seed, pid, step, 2, gv);
      swm_gaussians(P.seed, pid, step, 3, gw);
    }
    A.vx[i] = vx + c.sv * gv[0]; A.vy[i] = vy + c.sv * gv[1]; A.vz[i] = 0.0;
    A.wx[i] = 0.0; A.wy[i] = 0.0; A.wz[i] = w + c.sw * gw[0];
  }
  const double ddx = xn - xr, ddy = yn - yr;
  return (ddx * ddx + ddy * ddy) > lim2;
}

#define SW_STREAM_THREADS 256
#ifndef SW_STREAM_MINB
#define SW_STREAM_MINB 3
#endif
#ifndef SW_STREAM_PRE
#define SW_STREAM_PRE 1  // particles per thread whose next-step noise is drawn behind the grid barrier
#endif
template <bool NOISE>
__global__ void __launch_bounds__(SW_STREAM_THREADS, SW_STREAM_MINB) k_bd2d_stream(const __grid_constant__ DevParams P, DevArrays A, RunArgs R) {
  Ctl *ctl = A.ctl;
  if (ldcg(&ctl->halted)) return;
  __shared__ Bd2dSpecies S[SW_S];
  if (threadIdx.x < P.n_species) {
    const int sp = threadIdx.x;
    Bd2dSpecies c;
    c.gt1 = P.sp_gt[sp][1]; c.gt2 = P.sp_gt[sp][2]; c.gr0 = P.sp_gr[sp][0];
    c.r_gt1 = swm_rcp(c.gt1); c.r_gt2 = swm_rcp(c.gt2); c.r_gr0 = swm_rcp(c.gr0);
    c.sg_par = NOISE ? sqrt(2.0 * P.kT * P.dt / c.gt2) : 0.0;
    c.sg_per = NOISE ? sqrt(2.0 * P.kT * P.dt / c.gt1) : 0.0;
    c.sg_rot = NOISE ? sqrt(2.0 * P.kT * P.dt / c.gr0) : 0.0;
    c.sv = NOISE ? sqrt(P.kT / P.sp_mass[sp]) : 0.0;
    c.sw = NOISE ? sqrt(P.kT / P.sp_rin[sp][0]) : 0.0;
    S[sp] = c;
  }
  __syncthreads();
  unsigned int epoch = 0;
  const long long lo = (long long)blockIdx.x * blockDim.x, hi = sw_live_n(A);
  const long long gstride = (long long)gridDim.x * blockDim.x;
  const bool leader = (blockIdx.x == 0 && threadIdx.x == 0);
  const double lim2 = (0.5 * P.skin) * (0.5 * P.skin);

  // grid barrier = the resident kernel's: one acq_rel atomic per CTA on a 64-bit word whose high half counts
  // the CTAs that saw a particle beyond skin/2 (so the rebuild decision rides on the barrier), a separate
  // release word to poll, no fence.sc and no per-step flag traffic.
  __shared__ int s_need;
  unsigned int movers_seen = 0;
  // Between arrive and poll every thread draws the NEXT step's noise of its first particle (depends on
  // (seed, id, step) only): about one ninth of a step's instructions at 1M particles, which covers the
  // barrier's L2 round trips and the arrival skew of the other CTAs.
  const long long i0 = lo + threadIdx.x;
  __shared__ double s_g[SW_STREAM_PRE][3][SW_STREAM_THREADS];  // a thread only touches its own column
  bool have_g = false;
  auto step_barrier = [&](bool moved, bool draw, u64 next_step, bool keep_ok) -> bool {
    const int any = __syncthreads_or(moved);
    if (threadIdx.x == 0) grid_arrive64(&ctl->bar64, epoch, any != 0);
    have_g = false;
    if (NOISE && draw) {
#pragma unroll 1
      for (int k = 0; k < SW_STREAM_PRE; ++k) {
        const long long i = i0 + k * gstride;
        if (i >= hi) break;
        double a, b, c;
        swm_gaussians_xyt_keep(P.rk, sw_pid(A, i), next_step, keep_ok, &A.gkeep[i], a, b, c);
        s_g[k][0][threadIdx.x] = a; s_g[k][1][threadIdx.x] = b; s_g[k][2][threadIdx.x] = c;
      }
      have_g = true;
    }
    if (threadIdx.x == 0) {
      const unsigned int mv = (unsigned int)(grid_poll64(&ctl->bar64, epoch) >> 32);
      s_need = (mv != movers_seen);
      movers_seen = mv;
    }
    __syncthreads();
    return s_need != 0;
  };
  bool need = true;
  if (!R.force_rebuild) {
    bool moved = false;
    for (long long i = lo + threadIdx.x; i < hi; i += gstride) moved |= moved_since_build<2>(P, A, 0, i);
    need = step_barrier(moved, R.n_steps > 0, R.step0, false);
  }
  long long s = 0;
  for (; s < R.n_steps; ++s) {
    const int cur = (int)(s & 1);
    if (need) {
      for (long long i = lo + threadIdx.x; i < hi; i += gstride) phase_clear<2>(P, A, cur, i);
      grid_barrier64(&ctl->bar64, epoch);
      for (long long i = lo + threadIdx.x; i < hi; i += gstride) phase_insert(A, i);
      grid_barrier64(&ctl->bar64, epoch);
      for (long long i = lo + threadIdx.x; i < hi; i += gstride) phase_search<2>(P, A, cur, i);
      if (leader) ctl->rebuilds += 1;
      grid_barrier64(&ctl->bar64, epoch);
      if (ldcg(&ctl->overflow)) break;
    }
    const bool last = (s == R.n_steps - 1);
    bool moved = false;
    int k = 0;
    for (long long i = i0; i < hi; i += gstride, ++k) {
      const bool pre = NOISE && have_g && k < SW_STREAM_PRE;
      const int kk = pre ? k : 0;
      moved |= bd2d_step_stream<NOISE>(P, A, S, cur, R.step0 + (u64)s, last, i, (i + gstride < hi) ? i + gstride : -1, lim2,
                                       pre, s_g[kk][0][threadIdx.x], s_g[kk][1][threadIdx.x], s_g[kk][2][threadIdx.x], s > 0);
    }
    need = step_barrier(moved, !last, R.step0 + (u64)s + 1, true);
  }
  if (s & 1) {
    for (long long i = lo + threadIdx.x; i < hi; i += gstride) {
      A.x[0][i] = ldcg(&A.x[1][i]); A.y[0][i] = ldcg(&A.y[1][i]);
    }
  }
  if (leader) {
    ctl->steps_done = R.steps_base + s;
    if (s < R.n_steps) ctl->halted = 1;
    ctl->rebuild_flag[0] = 0; ctl->rebuild_flag[1] = 0; ctl->rebuild_flag[2] = 0;
  }
}

// ---------------------------------------------------------------------------------------------
// k_bd2d_env -- batched independent environments (BASELINE config 3): ONE WARP PER ENVIRONMENT, one agent per
// lane (env_size <= 32).  Environments never interact, so nothing here is grid-wide: no cell list, no barrier
// of any kind, and the Verlet list of an agent is a 32-bit mask in a register.  An agent's state stays in
// registers for the whole chunk of steps and its neighbours' positions arrive by warp shuffle, visited in
// ascending agent index, which is exactly the order of the (ascending) Verlet list the other kernels and the
// oracle sum in; pairs beyond the cut-off add nothing, so the result equals the list kernels' up to FMA
// contraction (1e-13) and equals the all-pairs loop (SWARM_B200_ENV_EXACT_PAIRS) bit for bit.
// ---------------------------------------------------------------------------------------------
#define SW_ENV_THREADS 256
// RunArgs.force_rebuild is meaningless for the list-free kernel: the launcher reuses it as "exact pairs only"
__device__ __forceinline__ bool getenv_no_prefilter(const RunArgs &R) { return R.force_rebuild == 2; }
template <bool NOISE>
__global__ void __launch_bounds__(SW_ENV_THREADS) k_bd2d_env(const __grid_constant__ DevParams P, DevArrays A, RunArgs R) {
  Ctl *ctl = A.ctl;
  if (ldcg(&ctl->halted)) return;
  const int lane = threadIdx.x & 31;
  const long long env = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (env >= P.n_envs) return;   // whole warp
  const int m = P.env_size;
  const bool act = lane < m;
  const long long i = env * m + (act ? lane : 0);
  double xi = A.x[0][i], yi = A.y[0][i], th = A.theta[i];
  const double fsw = A.f_swim[i], Tz = A.tqz[i];
  double fxe = 0.0, fye = 0.0;
  if (A.fex) { fxe = A.fex[i]; fye = A.fey[i]; }
  const int sp = A.species[i], ti = P.sp_type[sp];
  const u64 pid = sw_pid(A, i);
  const double dt = P.dt, kT = P.kT;
  const double gt1 = P.sp_gt[sp][1], gt2 = P.sp_gt[sp][2], gr0 = P.sp_gr[sp][0];
  const bool iso = (gt1 == gt2);
  const double r_gt1 = swm_rcp(gt1), r_gt2 = swm_rcp(gt2), r_gr0 = swm_rcp(gr0);
  const double sg_par = NOISE ? sqrt(2.0 * kT * dt / gt2) : 0.0;
  const double sg_per = NOISE ? sqrt(2.0 * kT * dt / gt1) : 0.0;
  const double sg_rot = NOISE ? sqrt(2.0 * kT * dt / gr0) : 0.0;
  const double w = swm_div_r(Tz, gr0, r_gr0);
  double vx_l = 0.0, vy_l = 0.0;
  const bool prefilter = P.periodic[0] && P.periodic[1] && !getenv_no_prefilter(R);
  const float bxf = (float)P.box[0], byf = (float)P.box[1], ibxf = (float)P.inv_box[0], ibyf = (float)P.inv_box[1];
  const float pr = (float)(P.max_cut + P.skin + 1e-4 * (1.0 + fmax(P.box[0], P.box[1])));
  const float pref2 = pr * pr;
  const double lim2 = (0.5 * P.skin) * (0.5 * P.skin);
  unsigned mask = 0u;
  double xr = xi, yr = yi;
  bool need = true;
  double g3_keep = 0.0;     // cosine normal of the last even step's rotational pair
  bool g3_have = false;
  for (long long s = 0; s < R.n_steps; ++s) {
    double ex, ey, g0 = 0.0, g1 = 0.0, g3 = 0.0;
    if (NOISE) {
      // normals and director in one pass (the bits of swm_gaussians_xyt + swm_sincos); the rotational normal of an odd
      // step is the cosine kept from the even step before it
      const u64 step = R.step0 + (u64)s;
      if ((step & 1ULL) && g3_have) swm_step_inputs<true>(P.rk, pid, step, th, g0, g1, g3, g3_keep, ey, ex);
      else swm_step_inputs<false>(P.rk, pid, step, th, g0, g1, g3, g3_keep, ey, ex);
      g3_have = !(step & 1ULL);
    } else {
      swm_sincos(th, ey, ex);
    }
    const Bd2dSelf sf = bd2d_self_forces(P, ti, xi, yi, ex, ey, fsw);
    double Fx = 0.0, Fy = 0.0;
    if (prefilter) {
      // A Verlet list in one register: bit j of `mask` = agent j was within cut-off + skin of this agent when the
      // environment was last screened.  The screen runs in fp32 on wrapped coordinates (all 32 x 31 pairs, one
      // shuffle per coordinate; its radius exceeds cut-off + skin by far more than the fp32 rounding of a wrapped
      // coordinate, so it can only admit extra pairs, never drop one) and is repeated when some agent of the
      // environment has moved more than skin / 2 since -- every 20-30 steps instead of every step.
      bool moved = false;
      if (act) {
        const double ddx = xi - xr, ddy = yi - yr;
        moved = !(ddx * ddx + ddy * ddy <= lim2);
      }
      if (need || __any_sync(0xffffffffu, moved)) {
        const float xf = (float)__dsub_rn(xi, __dmul_rn(P.box[0], rint(__dmul_rn(xi, P.inv_box[0]))));
        const float yf = (float)__dsub_rn(yi, __dmul_rn(P.box[1], rint(__dmul_rn(yi, P.inv_box[1]))));
        unsigned nm = 0u;
        for (int j = 0; j < m; ++j) {
          float dxf = xf - __shfl_sync(0xffffffffu, xf, j), dyf = yf - __shfl_sync(0xffffffffu, yf, j);
          dxf -= bxf * rintf(dxf * ibxf); dyf -= byf * rintf(dyf * ibyf);
          if (dxf * dxf + dyf * dyf < pref2) nm |= 1u << j;
        }
        mask = act ? (nm & ~(1u << lane)) : 0u;
        xr = xi; yr = yi;
        need = false;
      }
      // the listed partners in ascending index (the summation order of the other kernels and of the oracle); a lane
      // that has run out of partners shuffles with itself
      unsigned rem = mask;
      while (__any_sync(0xffffffffu, rem != 0u)) {
        const int j = rem ? (__ffs((int)rem) - 1) : lane;
        rem &= rem - 1u;
        const double xj = __shfl_sync(0xffffffffu, xi, j), yj = __shfl_sync(0xffffffffu, yi, j);
        const int tj = __shfl_sync(0xffffffffu, ti, j);
        if (j != lane) bd2d_pair(P, xi, yi, ti, tj, xj, yj, Fx, Fy);
      }
    } else {
      for (int j = 0; j < m; ++j) {
        const double xj = __shfl_sync(0xffffffffu, xi, j), yj = __shfl_sync(0xffffffffu, yi, j);
        const int tj = __shfl_sync(0xffffffffu, ti, j);
        if (j != lane) bd2d_pair(P, xi, yi, ti, tj, xj, yj, Fx, Fy);
      }
    }
    Fx += sf.wx; Fy += sf.wy;
    Fx += sf.sx; Fy += sf.sy;
    if (A.fex) { Fx += fxe; Fy += fye; }
    const Bd2dConst cc = {gt1, gt2, r_gt1, r_gt2, sg_par, sg_per, sg_rot};
    bd2d_update<NOISE>(iso, cc, dt, Fx, Fy, ex, ey, w, g0, g1, g3, xi, yi, th, vx_l, vy_l);
  }
  if (act) {
    A.x[0][i] = xi; A.y[0][i] = yi; A.theta[i] = th;
    if (R.n_steps > 0) {
      double gv[4] = {0, 0, 0, 0}, gw[4] = {0, 0, 0, 0};
      double sv = 0.0, sw = 0.0;
      if (NOISE) {
        swm_gaussians(P.seed, pid, R.step0 + (u64)(R.n_steps - 1), 2, gv);
        swm_gaussians(P.seed, pid, R.step0 + (u64)(R.n_steps - 1), 3, gw);
        sv = sqrt(kT / P.sp_mass[sp]);
        sw = sqrt(kT / P.sp_rin[sp][0]);
      }
      A.vx[i] = vx_l + sv * gv[0]; A.vy[i] = vy_l + sv * gv[1]; A.vz[i] = 0.0;
      A.wx[i] = 0.0; A.wy[i] = 0.0; A.wz[i] = w + sw * gw[0];
    }
  }
  if (env == 0 && lane == 0) ctl->steps_done = R.steps_base + R.n_steps;
}
