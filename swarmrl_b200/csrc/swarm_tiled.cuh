// swarm_tiled.cuh -- k_bd2d_tiled: the 2-D Brownian chunk on SPATIAL TILES (BASELINE config 4, 100k agents).
// Included by swarm_engine.cu after swarm_bd2d.cuh (same step arithmetic) and swarm_sort.cuh (the slot order).
//
// k_bd2d_resident pays, per step, one grid-wide barrier (148 arrivals on one L2 word, release, poll) and one L2 round
// trip for every neighbour position, because slot = particle id and a particle's neighbours live anywhere.  Here the
// slots are spatially sorted (swarm_sort.cuh), CTA b owns slots [b * ppc, (b + 1) * ppc) -- a compact tile -- and
//
//  * the positions of a tile live in SHARED MEMORY (double-buffered); a Verlet entry that points into the own tile is an
//    LDS, not an L2 load.  Only entries that point into another CTA's tile ("remote", the rim of the tile) go through L2;
//  * there is NO per-step grid barrier.  A CTA publishes the positions of its rim particles, then a step counter
//    (st.release.gpu flag[cta] = e + 1), and waits only for the flags of the CTAs its lists point into (its partners,
//    about 8 of them) -- a neighbourhood barrier whose latency hides behind the next step's particle-local work
//    (Philox + Box-Muller + director), exactly as the grid barrier's did, but with no global arrival chain and no skew
//    from far-away SMs;
//  * the Verlet rebuild cannot ride on a per-step barrier any more, so it runs on a CADENCE: every K steps all CTAs
//    meet at a grid barrier, VALIDATE that no position used for a force evaluation since the last rebuild had moved
//    more than skin / 2 (the criterion of ESPResSo's Verlet lists, espresso.py:269), and rebuild.  If the validation
//    fails, the period is ROLLED BACK to its checkpoint (positions at the last rebuild; the noise is counter-based, so
//    a replay draws the same numbers) and replayed with a shorter K; with K = 1 every step is validated, so this
//    terminates.  K adapts to the largest displacement seen.  Since forces are summed in ascending particle id
//    (DevArrays.okey) and a valid list contains every pair inside the cut-off, the trajectory does not depend on K, on
//    the tiling or on the permutation: it is bit-identical to k_bd2d_resident's.
//
// Memory-model notes: rim positions are plain stores followed by a CTA barrier and ONE st.release.gpu of the flag by
// the flag warp (cumulativity covers the other threads' stores); the reader's flag warp ld.acquire.gpu's the flag,
// then a CTA barrier, then the particle threads ld.cg (L2) the positions.  Two position buffers suffice: a CTA
// overwrites buffer p at epoch e + 2 only after all partners have published epoch e + 2, i.e. finished reading buffer p
// (the partner relation is made symmetric at every rebuild).
#pragma once

#define SW_TILE_MAXP (SW_RES_MAXT - 32)   // particles per CTA
#define SW_TILE_MAXCTA 256                // partner bitmask width
#define SW_TILE_KMAX 512
#define SW_TILE_COLS 11                     // columns of dynamic shared memory (see the kernel)

struct TileArgs {
  const int *perm;          // slot -> particle id
  unsigned int *flags;      // per CTA, 32 words apart: step epochs published
  unsigned int *pbits;      // per CTA, SW_TILE_MAXCTA / 32 words: CTAs its lists point into
  int ppc;                  // slots per CTA
  int k_init;               // first cadence when the handle has none yet
  unsigned int epoch0;      // first flag epoch of this launch: later than any epoch of the previous launches, so the
                            // flags are never reset (launch counter << 20; a launch runs far fewer than 2^20 epochs)
};

__device__ __forceinline__ unsigned int tile_ld_acquire(const unsigned int *p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void tile_st_release(unsigned int *p, unsigned int v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

template <bool NOISE>
__global__ void __launch_bounds__(SW_RES_MAXT, 1) k_bd2d_tiled(const __grid_constant__ DevParams P, DevArrays A, DevArrays T, TileArgs G,
                                                        RunArgs R) {
  Ctl *ctl = A.ctl;
  if (ldcg(&ctl->halted)) return;
  unsigned int epoch = 0;                  // grid-barrier arrivals expected so far
  const int ptid = (int)threadIdx.x - 32;  // warp 0 is the flag warp
  const int ppc = G.ppc;
  const int base = (int)blockIdx.x * ppc;
  const long long slot = (long long)base + ptid;
  const bool act = ptid >= 0 && ptid < ppc && slot < A.n;
  const bool leader = act && slot == 0;
  const int me = (int)blockIdx.x;

  // dynamic shared memory: SW_TILE_COLS columns of SW_TILE_MAXP doubles, one entry per particle thread
  extern __shared__ double dsm[];
  double (*sx)[SW_TILE_MAXP] = (double (*)[SW_TILE_MAXP])dsm;                          // [2]: positions of the tile, by step parity
  double (*sy)[SW_TILE_MAXP] = (double (*)[SW_TILE_MAXP])(dsm + 2 * SW_TILE_MAXP);      // [2]
  double *sxr = dsm + 4 * SW_TILE_MAXP, *syr = dsm + 5 * SW_TILE_MAXP, *sthr = dsm + 6 * SW_TILE_MAXP;  // checkpoint = state at the last rebuild
  double *sfx = dsm + 7 * SW_TILE_MAXP, *sfy = dsm + 8 * SW_TILE_MAXP;                  // constant external force
  double *swx = dsm + 9 * SW_TILE_MAXP, *swy = dsm + 10 * SW_TILE_MAXP;                 // wall force of the coming step
  __shared__ Bd2dSpecies S[SW_S];
  __shared__ unsigned int s_pbits[SW_TILE_MAXCTA / 32];
  __shared__ int s_partner[SW_TILE_MAXCTA];
  __shared__ int s_np;
  __shared__ int s_viol;                  // validation verdict (grid-wide), broadcast to the CTA
  __shared__ float s_dm;                  // grid-wide largest displacement^2 of the period
  __shared__ unsigned int s_wmax[SW_RES_MAXT / 32];

  // ---- particle state: caller arrays are indexed by id, registers hold it for the whole chunk --------------------
  int id = 0, sp = 0, ti = 0;
  u64 pid = 0;
  double xi = 0.0, yi = 0.0, th = 0.0, fsw = 0.0, w = 0.0;
  if (threadIdx.x < P.n_species) {
    // per-species constants: shared memory, not 20 registers of every thread
    const int k = threadIdx.x;
    Bd2dSpecies c;
    c.gt1 = P.sp_gt[k][1]; c.gt2 = P.sp_gt[k][2]; c.gr0 = P.sp_gr[k][0];
    c.r_gt1 = swm_rcp(c.gt1); c.r_gt2 = swm_rcp(c.gt2); c.r_gr0 = swm_rcp(c.gr0);
    c.sg_par = NOISE ? sqrt(2.0 * P.kT * P.dt / c.gt2) : 0.0;
    c.sg_per = NOISE ? sqrt(2.0 * P.kT * P.dt / c.gt1) : 0.0;
    c.sg_rot = NOISE ? sqrt(2.0 * P.kT * P.dt / c.gr0) : 0.0;
    c.sv = NOISE ? sqrt(P.kT / P.sp_mass[k]) : 0.0;
    c.sw = NOISE ? sqrt(P.kT / P.sp_rin[k][0]) : 0.0;
    S[k] = c;
  }
  __syncthreads();
  if (act) {
    id = G.perm[slot];
    pid = (u64)T.gid[slot];
    xi = A.x[0][id]; yi = A.y[0][id]; th = A.theta[id];
    fsw = A.f_swim[id];
    sp = A.species[id]; ti = P.sp_type[sp];
    w = swm_div_r(A.tqz[id], S[sp].gr0, S[sp].r_gr0);   // omega_z: the torque is constant over the chunk
    if (A.fex) { sfx[ptid] = A.fex[id]; sfy[ptid] = A.fey[id]; }
  }
  const bool iso = (S[sp].gt1 == S[sp].gt2);
  const double lim2 = (0.5 * P.skin) * (0.5 * P.skin);
  const double dt = P.dt;

  // ---- rebuild cadence carried over from the previous launch -----------------------------------------------------
  int K = ldcg(&ctl->tile_K);
  if (K < 1) K = G.k_init;
  int since = ldcg(&ctl->tile_since);
  unsigned int movers_seen = 0;   // thread 0: high word of the barrier counter at the last vote
  unsigned int vidx = 0;          // validation index (tile_dmax encoding)
  bool need = true;               // rebuild before the next step
  if (threadIdx.x == 0) { s_np = 0; s_viol = 0; }
  if (act) {
    sxr[ptid] = T.xr[slot]; syr[ptid] = T.yr[slot]; sthr[ptid] = th;
    sx[0][ptid] = xi; sy[0][ptid] = yi;
  }
  // grid-wide vote: `mine` of any thread set  ->  returns true everywhere; dm2 = grid-wide max of `d2`
  auto vote = [&](bool mine, double d2) -> bool {
    // largest displacement^2 of the CTA as float bits (non-negative floats order like unsigned integers)
    unsigned int fb = __float_as_uint(__double2float_ru(d2));
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) fb = max(fb, __shfl_xor_sync(0xffffffffu, fb, d));
    if ((threadIdx.x & 31) == 0) s_wmax[threadIdx.x >> 5] = fb;
    const int any = __syncthreads_or(mine);
    if (threadIdx.x < 32) {
      unsigned int m = (threadIdx.x < (blockDim.x >> 5)) ? s_wmax[threadIdx.x] : 0u;
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, d));
      if (threadIdx.x == 0) {
        ++vidx;
        atomicMax(&ctl->tile_dmax, ((unsigned long long)vidx << 32) | m);
        grid_arrive64(&ctl->bar64, epoch, any != 0);
        const unsigned int mv = (unsigned int)(grid_poll64(&ctl->bar64, epoch) >> 32);
        s_viol = (mv != movers_seen);
        movers_seen = mv;
        s_dm = __uint_as_float((unsigned int)ldcg((const long long *)&ctl->tile_dmax));
      }
    }
    __syncthreads();
    return s_viol != 0;
  };
  if (!R.force_rebuild) {
    // the list of the previous launch is still good if nobody has been moved beyond skin / 2 of its reference
    double d2 = 0.0;
    if (act) { const double ddx = xi - sxr[ptid], ddy = yi - syr[ptid]; d2 = ddx * ddx + ddy * ddy; }
    need = vote(act && !(d2 <= lim2), (d2 <= lim2) ? d2 : 0.0) || since >= K;
  } else {
    since = 0;
  }

  // ---- particle-local inputs of the first step ---------------------------------------------------------------------
  double g0 = 0.0, g1 = 0.0, g3 = 0.0, ex = 1.0, ey = 0.0;
  auto local_inputs = [&](u64 step) {
    swm_sincos(th, ey, ex);
    if (P.walls) {
      const Bd2dSelf sf = bd2d_self_forces(P, ti, xi, yi, ex, ey, fsw);
      swx[ptid] = sf.wx; swy[ptid] = sf.wy;
    }
    if (NOISE) swm_gaussians_xyt(P.seed, pid, step, g0, g1, g3);
  };
  if (act) local_inputs(R.step0);

  int cnt = 0, nb[SW_RES_NBR];
  const int self_entry = act ? (int)slot : 0;
  bool exported = false;   // some other CTA's list points at this particle (lists are symmetric: mine points into it)
  auto load_list = [&]() {
    cnt = T.nbr_cnt[slot];
    exported = false;
#pragma unroll
    for (int k = 0; k < SW_RES_NBR; ++k) nb[k] = (k < cnt) ? T.nbr[(long long)k * T.n_pad + slot] : self_entry;
    auto classify = [&](int en) {
      const int j = en & SW_IDX_MASK;
      if ((unsigned int)(j - base) >= (unsigned int)ppc) {
        exported = true;
        const int owner = j / ppc;
        atomicOr(&s_pbits[owner >> 5], 1u << (owner & 31));
      }
    };
#pragma unroll
    for (int k = 0; k < SW_RES_NBR; ++k) classify(nb[k]);   // unused entries point at the particle itself: local
    for (int k = SW_RES_NBR; k < cnt; ++k) classify(T.nbr[(long long)k * T.n_pad + slot]);
  };
#pragma unroll
  for (int k = 0; k < SW_RES_NBR; ++k) nb[k] = self_entry;
  // partner list of this CTA from the published bit rows (symmetric closure: c is a partner if my lists point into c
  // OR c's lists point into me); warp 0 only, followed by a CTA barrier at the call site
  auto build_partners = [&]() {
    const int G_ = (int)gridDim.x;
    int np = 0;
    for (int c0 = 0; c0 < G_; c0 += 32) {
      const int c = c0 + (int)threadIdx.x;
      bool p = false;
      if (c < G_ && c != me) {
        const unsigned int mine = s_pbits[c >> 5] >> (c & 31);
        const unsigned int theirs = __ldcg(&G.pbits[c * (SW_TILE_MAXCTA / 32) + (me >> 5)]) >> (me & 31);
        p = ((mine | theirs) & 1u) != 0u;
      }
      const unsigned int b = __ballot_sync(0xffffffffu, p);
      if (p) s_partner[np + __popc(b & ((1u << threadIdx.x) - 1u))] = c;
      np += __popc(b);
    }
    if (threadIdx.x == 0) { s_np = np; atomicAdd((unsigned long long *)&ctl->tile_partners, (unsigned long long)np); }
  };
  double vx_l = 0.0, vy_l = 0.0;
  bool viol = false;        // a position of this thread used for forces since the last rebuild was beyond skin / 2
  double dmax2 = 0.0;       // its largest displacement^2 in the period

#ifdef SW_TIMING
  // clock64 per phase of CTA 0: slots 0..5 from the first particle thread, 6..7 from the flag thread (swarm_debug_timing)
  long long tp[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tc = clock64();
  const bool prof = (blockIdx.x == 0) && (threadIdx.x == 32 || threadIdx.x == 0);
#endif
  long long s = 0;          // steps completed in this launch
  long long s_ck = 0;       // step index of the checkpoint
  bool ck_start = true;     // checkpoint = launch start (state still in the caller's arrays)
  unsigned int e = G.epoch0;  // step epochs executed, replays included (flag values, buffer parity)
  bool have_partners = false;
  bool stop = false;

  for (;;) {
    const bool at_end = (s == R.n_steps);
    if (need || since >= K || at_end || !have_partners) {
      // ================= synchronisation point: validate the period, then rebuild ================================
      if (s > s_ck) {
        double d2n = 0.0;
        if (act) { const double ddx = xi - sxr[ptid], ddy = yi - syr[ptid]; d2n = ddx * ddx + ddy * ddy; }
        const bool bad = vote(viol, fmax(dmax2, d2n));
        const double dm = (double)s_dm;
        const long long len = s - s_ck;
        // cadence from the largest displacement: linear extrapolation (exact for drift, conservative for diffusion)
        double ratio = (dm > 0.0) ? sqrt(lim2 / dm) : 2.0;
        if (ratio > 2.0) ratio = 2.0;
        long long Kn = (long long)((double)(since > 0 ? since : 1) * ratio * 0.8);
        if (bad) {
          if (Kn > K / 2) Kn = K / 2;
          K = (int)(Kn < 1 ? 1 : Kn);
          // ---- roll the period back to its checkpoint and replay it with the shorter cadence -------------------
          const int cur = (int)(e & 1u);
          if (ck_start) {
            if (act) { xi = A.x[0][id]; yi = A.y[0][id]; th = A.theta[id]; }
            since = 0; need = true;
          } else {
            if (act) { xi = sxr[ptid]; yi = syr[ptid]; th = sthr[ptid]; }
            since = 0; need = false;
          }
          s = s_ck;
          viol = false; dmax2 = 0.0;
          if (act) {
            sx[cur][ptid] = xi; sy[cur][ptid] = yi;
            T.x[cur][slot] = xi; T.y[cur][slot] = yi;
            local_inputs(R.step0 + (u64)s);
          }
          if (leader) ctl->tile_rollbacks += 1;
          grid_barrier64(&ctl->bar64, epoch);   // restored rim positions visible before anybody steps
          (void)len;
          continue;
        }
        if (Kn < 1) Kn = 1;
        if (Kn > SW_TILE_KMAX) Kn = SW_TILE_KMAX;
        if (since >= K || need) K = (int)Kn;   // a period that ended early (launch end) says nothing about a longer one
        else if (Kn < K) K = (int)Kn;
      }
      if (at_end) break;
      if (need || since >= K) {
        // ---- rebuild at the current positions (buffer e & 1) -----------------------------------------------------
        const int cur = (int)(e & 1u);
        if (threadIdx.x < SW_TILE_MAXCTA / 32) s_pbits[threadIdx.x] = 0u;
        if (act) {
          T.x[cur][slot] = xi; T.y[cur][slot] = yi;
          phase_clear<2>(P, T, cur, slot);
        }
        grid_barrier64(&ctl->bar64, epoch);
        if (act) phase_insert(T, slot);
        grid_barrier64(&ctl->bar64, epoch);
        if (act) {
          phase_search<2>(P, T, cur, slot);
          sxr[ptid] = xi; syr[ptid] = yi; sthr[ptid] = th;
          load_list();
        }
        if (leader) ctl->rebuilds += 1;
        __syncthreads();
        if (threadIdx.x < SW_TILE_MAXCTA / 32) G.pbits[me * (SW_TILE_MAXCTA / 32) + threadIdx.x] = s_pbits[threadIdx.x];
        grid_barrier64(&ctl->bar64, epoch);
        if (ldcg(&ctl->overflow)) { stop = true; break; }   // uniform: read after the barrier
        if (threadIdx.x < 32) build_partners();
        __syncthreads();
        have_partners = true;
        s_ck = s; ck_start = false;
        since = 0; need = false;
        viol = false; dmax2 = 0.0;
      } else {
        // list kept from the previous launch (validated by the start vote): classify it, publish, exchange partners
        const int cur = (int)(e & 1u);
        if (threadIdx.x < SW_TILE_MAXCTA / 32) s_pbits[threadIdx.x] = 0u;
        __syncthreads();
        if (act) {
          T.x[cur][slot] = xi; T.y[cur][slot] = yi;
          load_list();
        }
        __syncthreads();
        if (threadIdx.x < SW_TILE_MAXCTA / 32) G.pbits[me * (SW_TILE_MAXCTA / 32) + threadIdx.x] = s_pbits[threadIdx.x];
        grid_barrier64(&ctl->bar64, epoch);
        if (threadIdx.x < 32) build_partners();
        __syncthreads();
        have_partners = true;
      }
    }
    SW_TICK(0);   // synchronisation points (validation, rebuild)
    // ================= one Brownian step: read buffer cur, write buffer nxt ==========================================
    const int cur = (int)(e & 1u), nxt = cur ^ 1;
    const bool last = (s == R.n_steps - 1);
    if (act) {
      {
        // the position used for this force evaluation must be within skin / 2 of the list's reference
        const double ddx = xi - sxr[ptid], ddy = yi - syr[ptid];
        const double d2 = ddx * ddx + ddy * ddy;
        viol = viol || !(d2 <= lim2);
        dmax2 = fmax(dmax2, d2);
      }
      double xj[SW_RES_NBR], yj[SW_RES_NBR];
#pragma unroll
      for (int k = 0; k < SW_RES_NBR; ++k) {
        const int j = nb[k] & SW_IDX_MASK;
        const unsigned int l = (unsigned int)(j - base);
        if (l < (unsigned int)ppc) { xj[k] = sx[cur][l]; yj[k] = sy[cur][l]; }
        else { xj[k] = ldcg(&T.x[cur][j]); yj[k] = ldcg(&T.y[cur][j]); }
      }
      double Fx = 0.0, Fy = 0.0;
#pragma unroll
      for (int k = 0; k < SW_RES_NBR; ++k)
        if (k < cnt) bd2d_pair(P, xi, yi, ti, nb[k] >> SW_TYPE_SHIFT, xj[k], yj[k], Fx, Fy);
      for (int k0 = SW_RES_NBR; k0 < cnt; k0 += 4) {
        int en4[4];
        double xe[4], ye[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) en4[u] = (k0 + u < cnt) ? T.nbr[(long long)(k0 + u) * T.n_pad + slot] : self_entry;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int j = en4[u] & SW_IDX_MASK;
          const unsigned int l = (unsigned int)(j - base);
          if (l < (unsigned int)ppc) { xe[u] = sx[cur][l]; ye[u] = sy[cur][l]; }
          else { xe[u] = ldcg(&T.x[cur][j]); ye[u] = ldcg(&T.y[cur][j]); }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (k0 + u < cnt) bd2d_pair(P, xi, yi, ti, en4[u] >> SW_TYPE_SHIFT, xe[u], ye[u], Fx, Fy);
      }
      // walls, swim, external: added after the pair forces, in the oracle's order
      if (P.walls) { Fx += swx[ptid]; Fy += swy[ptid]; }
      Fx += __dmul_rn(fsw, ex); Fy += __dmul_rn(fsw, ey);   // product rounded on its own, as in the other kernels
      if (A.fex) { Fx += sfx[ptid]; Fy += sfy[ptid]; }
      const Bd2dSpecies &c = S[sp];
      const Bd2dConst cc = {c.gt1, c.gt2, c.r_gt1, c.r_gt2, c.sg_par, c.sg_per, c.sg_rot};
      bd2d_update<NOISE>(iso, cc, dt, Fx, Fy, ex, ey, w, g0, g1, g3, xi, yi, th, vx_l, vy_l);
      sx[nxt][ptid] = xi; sy[nxt][ptid] = yi;
      if (exported) { T.x[nxt][slot] = xi; T.y[nxt][slot] = yi; }
    }
    SW_TICK(1);   // pair forces + update
    // the next iteration is a grid-wide synchronisation point: its barrier replaces the neighbourhood barrier
    const bool next_sync = (since + 1 >= K) || last;
    if (threadIdx.x < 32) {
      // ---- flag warp: all rim stores of the CTA done -> publish the epoch -> wait for the partners' -----------------
      asm volatile("barrier.cta.sync 1, %0;" ::"r"((int)blockDim.x) : "memory");
      SW_TICK(2);   // flag warp: until every particle warp has stored
      if (threadIdx.x == 0) tile_st_release(&G.flags[me * 32], e + 1u);
      SW_TICK(6);   // flag thread: fence + release
      if (!next_sync) {
        const int np = s_np;
        for (int p = (int)threadIdx.x; p < np; p += 32) {
          const unsigned int *f = &G.flags[s_partner[p] * 32];
          while ((int)(tile_ld_acquire(f) - (e + 1u)) < 0) {
          }
        }
      }
      __syncwarp();
      SW_TICK(7);   // flag thread: partners' flags
    } else {
      asm volatile("barrier.cta.arrive 1, %0;" ::"r"((int)blockDim.x) : "memory");
      SW_TICK(2);
      if (act && !last) local_inputs(R.step0 + (u64)(s + 1));   // step s+1's particle-local work hides the flag wait
    }
    SW_TICK(4);   // particle-local work of the next step
    __syncthreads();
    SW_TICK(5);   // wait for the flag warp (partners)
    ++s; ++since; ++e;
  }
#ifdef SW_TIMING
  if (prof) {
    if (threadIdx.x == 0) { ctl->tprof[6] += tp[6]; ctl->tprof[7] += tp[7]; }
    else for (int k = 0; k < 6; ++k) ctl->tprof[k] += tp[k];
  }
#endif

  // ---- results back to the caller's arrays (indexed by id) --------------------------------------------------------
  if (act) {
    A.x[0][id] = xi; A.y[0][id] = yi; A.theta[id] = th;
    if (s > 0) {
      double gv[4] = {0, 0, 0, 0}, gw[4] = {0, 0, 0, 0};
      const double sv = S[sp].sv, sw = S[sp].sw;
      if (NOISE) {
        swm_gaussians(P.seed, pid, R.step0 + (u64)(s - 1), 2, gv);
        swm_gaussians(P.seed, pid, R.step0 + (u64)(s - 1), 3, gw);
      }
      A.vx[id] = vx_l + sv * gv[0]; A.vy[id] = vy_l + sv * gv[1]; A.vz[id] = 0.0;
      A.wx[id] = 0.0; A.wy[id] = 0.0; A.wz[id] = w + sw * gw[0];
    }
  }
  if (leader) {
    ctl->steps_done = R.steps_base + s;
    if (stop) ctl->halted = 1;
    ctl->tile_K = K;
    ctl->tile_since = since;
    ctl->rebuild_flag[0] = 0; ctl->rebuild_flag[1] = 0; ctl->rebuild_flag[2] = 0;
  }
}
