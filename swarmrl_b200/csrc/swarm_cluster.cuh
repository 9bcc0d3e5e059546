// swarm_cluster.cuh -- k_bd2d_cluster: SMALL 2-D swarms (N <= 8 x 256: BASELINE configs 1 and 2) on ONE thread-block cluster.
// Included by swarm_engine.cu after swarm_bd2d.cuh (same step arithmetic: bd2d_pair, bd2d_self_forces, bd2d_update,
// swm_step_inputs -- the trajectory is the resident kernel's bit for bit).
//
// Why: k_bd2d_resident pays one grid barrier per step -- store -> atomic -> release -> poll through L2, ~3 us whatever the
// swarm size -- so 1,000 particles ran at 8 us per step and 10 particles at 2.9 us (tools/measure_table.py): latency, not work.
// A cluster of 8 CTAs is co-scheduled on one GPC and owns two things the grid does not have:
//   * DISTRIBUTED SHARED MEMORY: CTA r keeps the positions of its particles in its own shared memory (double-buffered by step
//     parity); a neighbour that lives in another CTA is read with ld.shared::cluster through mapa -- an SM-to-SM access that
//     never touches L2;
//   * the HARDWARE CLUSTER BARRIER: barrier.cluster.arrive (release) right after a thread has written its new position,
//     barrier.cluster.wait (acquire) after the particle-local work of the next step (Philox, Box-Muller, director) -- no atomic,
//     no polling loop, no memory fence of its own.
// The "somebody moved beyond skin / 2" vote rides on the same barrier: every CTA leaves its verdict in a shared-memory flag
// of the step's parity before arriving, and reads the eight flags through DSMEM after the wait.
// The Verlet rebuild stays inside the cluster too: every CTA gathers all N positions into its shared memory (coalesced DSMEM
// reads), bins them by a counting sort on a grid of at most 64 x 64 cells in shared memory (histogram, exclusive scan, fill; all
// eight CTAs build the same table) and each thread tests the particles of its 3 x 3 stencil, keeping its list in ascending
// index -- the summation order of the forces in every kernel.  Up to 128 particles the scan simply visits all pairs.  The lists
// are written to A.nbr / A.nbr_cnt / A.xr / A.yr as well, in the common format, so the neighbour dumps, swarm_forces and the other
// kernels find what they expect.
#pragma once
#include <cooperative_groups.h>
namespace cg = cooperative_groups;

#define SW_CL_CTAS 8
#define SW_CL_THREADS 256
#define SW_CL_MAXN (SW_CL_CTAS * SW_CL_THREADS)
#define SW_CL_CAP 32                       // Verlet entries per particle held in shared memory
#define SW_CL_G 64                         // the rebuild's own cell grid: at most G x G cells of edge >= cut-off + skin
#define SW_CL_SCAN_N 128                   // swarms up to this size scan all pairs instead (10 colloids in a 1000 um box: C1)
// dynamic shared memory: doubles sx[2][T], sy[2][T], ax[MAXN], ay[MAXN]; ints atype[MAXN], snb[CAP][T], flags[8],
// cstart[G*G + 1], cptr[G*G], acell[MAXN], corder[MAXN]
#define SW_CL_DSM_BYTES (sizeof(double) * (4 * SW_CL_THREADS + 2 * SW_CL_MAXN) + \
                         sizeof(int) * (SW_CL_MAXN + SW_CL_CAP * SW_CL_THREADS + 8 + (SW_CL_G * SW_CL_G + 1) + SW_CL_G * SW_CL_G + 2 * SW_CL_MAXN + 16))

template <bool NOISE>
__global__ void __cluster_dims__(SW_CL_CTAS, 1, 1) __launch_bounds__(SW_CL_THREADS, 1)
k_bd2d_cluster(const __grid_constant__ DevParams P, DevArrays A, RunArgs R) {
  cg::cluster_group cluster = cg::this_cluster();
  Ctl *ctl = A.ctl;
  if (ldcg(&ctl->halted)) return;            // written only by the last kernel's leader: uniform over the cluster
  extern __shared__ double cl_dsm[];
  double *sx = cl_dsm, *sy = cl_dsm + 2 * SW_CL_THREADS;                       // [2][T] by step parity
  double *ax = cl_dsm + 4 * SW_CL_THREADS, *ay = ax + SW_CL_MAXN;             // all positions, gathered at a rebuild
  int *atype = reinterpret_cast<int *>(ay + SW_CL_MAXN);                       // type of every particle (does not change)
  int *snb = atype + SW_CL_MAXN;                                               // [CAP][T]: (type << 28) | (owner CTA << 16) | local index
  int *flags = snb + SW_CL_CAP * SW_CL_THREADS;                                // [0..1] moved-beyond-skin/2 by parity, [2] list overflow, [3] longest list
  int *cstart = flags + 8;                                                     // [G*G + 1] first slot of every cell in corder (exclusive scan)
  int *cptr = cstart + SW_CL_G * SW_CL_G + 1;                                  // [G*G] fill pointer of every cell
  int *acell = cptr + SW_CL_G * SW_CL_G;                                       // [MAXN] cell of every particle
  int *corder = acell + SW_CL_MAXN;                                            // [MAXN] particle indices, cell by cell
  int *wsum = corder + SW_CL_MAXN;                                             // [8] warp totals of the scan
  // the rebuild's grid: the engine's own cell counts capped at G per axis (coarser cells are still >= cut-off + skin)
  const int gx = P.nc[0] < SW_CL_G ? P.nc[0] : SW_CL_G, gy = P.nc[1] < SW_CL_G ? P.nc[1] : SW_CL_G;
  const int ncell = gx * gy;
  auto cl_coord = [&](double x, int a, int g) -> int {
    double u = x * P.inv_box[a];
    if (P.periodic[a]) u -= floor(u);
    int c = (int)floor(u * (double)g);
    return c < 0 ? 0 : (c >= g ? g - 1 : c);
  };
  const int rank = (int)cluster.block_rank();
  const int ppc = R.particles_per_cta;
  const int l = (int)threadIdx.x;
  const long long n = A.n;
  const long long i = (long long)rank * ppc + l;
  const bool act = l < ppc && i < n;
  const bool leader = (rank == 0 && l == 0);
  const int caps = A.cap < SW_CL_CAP ? A.cap : SW_CL_CAP;

  int sp = 0, ti = 0;
  u64 pid = 0;
  double xi = 0.0, yi = 0.0, th = 0.0, fsw = 0.0, w = 0.0, fxe = 0.0, fye = 0.0, xr = 0.0, yr = 0.0;
  if (act) {
    sp = A.species[i]; ti = P.sp_type[sp];
    pid = sw_pid(A, i);
    xi = A.x[0][i]; yi = A.y[0][i]; th = A.theta[i];
    fsw = A.f_swim[i];
    if (A.fex) { fxe = A.fex[i]; fye = A.fey[i]; }
  }
  const double gt1 = P.sp_gt[sp][1], gt2 = P.sp_gt[sp][2], gr0 = P.sp_gr[sp][0];
  const bool iso = (gt1 == gt2);
  const double r_gt1 = swm_rcp(gt1), r_gt2 = swm_rcp(gt2), r_gr0 = swm_rcp(gr0);
  const double kT = P.kT, dt = P.dt;
  const Bd2dConst cc = {gt1, gt2, r_gt1, r_gt2, NOISE ? sqrt(2.0 * kT * dt / gt2) : 0.0, NOISE ? sqrt(2.0 * kT * dt / gt1) : 0.0,
                        NOISE ? sqrt(2.0 * kT * dt / gr0) : 0.0};
  if (act) w = swm_div_r(A.tqz[i], gr0, r_gr0);
  const double lim2 = (0.5 * P.skin) * (0.5 * P.skin);
  for (long long j = l; j < n; j += blockDim.x) atype[j] = P.sp_type[A.species[j]];
  if (l < 4) flags[l] = 0;
  if (act) { sx[l] = xi; sy[l] = yi; }      // parity 0
  cluster.sync();

  // particle-local inputs of a step: normals + director in one pass; the cosine of the even step's rotational pair stays in
  // a register for the odd step (swarm_math.cuh)
  double g0 = 0.0, g1 = 0.0, g3 = 0.0, g3_keep = 0.0, ex = 1.0, ey = 0.0;
  bool keep_ok = false;
  auto local_inputs = [&](u64 step) {
    if (NOISE) {
      if (keep_ok && (step & 1ULL)) swm_step_inputs<true>(P.rk, pid, step, th, g0, g1, g3, g3_keep, ey, ex);
      else swm_step_inputs<false>(P.rk, pid, step, th, g0, g1, g3, g3_keep, ey, ex);
      keep_ok = !(step & 1ULL);
    } else {
      swm_sincos(th, ey, ex);
    }
  };
  if (act) local_inputs(R.step0);

  int cnt = 0;
  double vx_l = 0.0, vy_l = 0.0;
  bool need = true;        // the list is rebuilt at the start of every launch (a few microseconds at this size)
  bool stop = false;
  long long s = 0;
  for (; s < R.n_steps; ++s) {
    const int cur = (int)(s & 1), nxt = cur ^ 1;
    if (need) {
      // ---- Verlet rebuild: gather every position of the cluster, scan in ascending index ----------------------------------
      for (long long j = l; j < n; j += blockDim.x) {
        const int r = (int)(j / ppc), lj = (int)(j - (long long)r * ppc);
        const double *rx = cluster.map_shared_rank(sx + cur * SW_CL_THREADS, r);
        const double *ry = cluster.map_shared_rank(sy + cur * SW_CL_THREADS, r);
        ax[j] = rx[lj]; ay[j] = ry[lj];
      }
      __syncthreads();
      bool over = false;
      int lj_[SW_CL_CAP], lt_[SW_CL_CAP];    // this particle's list: neighbour index (kept ascending) and type
      cnt = 0;
      auto consider = [&](int j) {
        if (j == (int)i) return;
        const int tj = atype[j];
        const double ddx = sw_min_image(xi - ax[j], P.box[0], P.inv_box[0], P.periodic[0]);
        const double ddy = sw_min_image(yi - ay[j], P.box[1], P.inv_box[1], P.periodic[1]);
        const double rl = __dadd_rn(P.wca_cut[ti][tj], P.skin);
        if (sw_r2(ddx, ddy, 0.0) < __dmul_rn(rl, rl)) {
          if (cnt < caps) {
            int k = cnt++;
            while (k > 0 && lj_[k - 1] > j) { lj_[k] = lj_[k - 1]; lt_[k] = lt_[k - 1]; --k; }
            lj_[k] = j; lt_[k] = tj;
          } else {
            over = true;
          }
        }
      };
      if (n <= SW_CL_SCAN_N) {
        if (act) for (int j = 0; j < (int)n; ++j) consider(j);
      } else {
        // ---- counting sort of the N particles by cell, in this CTA's shared memory (every CTA builds the same table) -----------
        for (int c = l; c < ncell; c += blockDim.x) cptr[c] = 0;
        __syncthreads();
        for (long long j = l; j < n; j += blockDim.x) {
          const int c = cl_coord(ay[j], 1, gy) * gx + cl_coord(ax[j], 0, gx);
          acell[j] = c;
          atomicAdd(&cptr[c], 1);
        }
        __syncthreads();
        {
          // exclusive scan of cptr[0 .. ncell) -> cstart: 16 consecutive cells per thread, warp shuffles, warp totals
          const int per = (ncell + (int)blockDim.x - 1) / (int)blockDim.x;
          const int c0 = l * per;
          int sum = 0;
          for (int c = c0; c < c0 + per && c < ncell; ++c) sum += cptr[c];
          int inc = sum;
#pragma unroll
          for (int d = 1; d < 32; d <<= 1) { const int v = __shfl_up_sync(0xffffffffu, inc, d); if ((l & 31) >= d) inc += v; }
          if ((l & 31) == 31) wsum[l >> 5] = inc;
          __syncthreads();
          int base = inc - sum;
          for (int wq = 0; wq < (l >> 5); ++wq) base += wsum[wq];
          for (int c = c0; c < c0 + per && c < ncell; ++c) { const int m = cptr[c]; cstart[c] = base; base += m; }
          if (l == (int)blockDim.x - 1) cstart[ncell] = (int)n;
          __syncthreads();
          for (int c = l; c < ncell; c += blockDim.x) cptr[c] = cstart[c];
          __syncthreads();
        }
        for (long long j = l; j < n; j += blockDim.x) corder[atomicAdd(&cptr[acell[j]], 1)] = (int)j;
        __syncthreads();
        if (act) {
          const int cx = cl_coord(xi, 0, gx), cy = cl_coord(yi, 1, gy);
          // the stencil visits every cell once: one cell per axis in a 1-cell grid, both in a 2-cell grid, else -1, 0, +1
          const int nx = gx >= 3 ? 3 : gx, ny = gy >= 3 ? 3 : gy;
          for (int qy = 0; qy < ny; ++qy)
            for (int qx = 0; qx < nx; ++qx) {
              int xx = gx >= 3 ? cx + qx - 1 : qx, yy = gy >= 3 ? cy + qy - 1 : qy;
              if (gx >= 3) { if (xx < 0) { if (!P.periodic[0]) continue; xx += gx; } else if (xx >= gx) { if (!P.periodic[0]) continue; xx -= gx; } }
              if (gy >= 3) { if (yy < 0) { if (!P.periodic[1]) continue; yy += gy; } else if (yy >= gy) { if (!P.periodic[1]) continue; yy -= gy; } }
              const int c = yy * gx + xx;
              for (int q = cstart[c]; q < cstart[c + 1]; ++q) consider(corder[q]);
            }
        }
      }
      if (act) {
        for (int k = 0; k < cnt; ++k) {
          const int j = lj_[k], r = j / ppc;
          snb[k * SW_CL_THREADS + l] = (lt_[k] << SW_TYPE_SHIFT) | (r << 16) | (j - r * ppc);
          A.nbr[(long long)k * A.n_pad + i] = j | (lt_[k] << SW_TYPE_SHIFT);
        }
        A.nbr_cnt[i] = cnt;
        A.xr[i] = xi; A.yr[i] = yi;
        xr = xi; yr = yi;
        if (cnt > flags[3]) atomicMax(&flags[3], cnt);
      }
      if (__syncthreads_or(over)) flags[2] = 1;     // every thread of the CTA writes the same value
      cluster.sync();
      int any_over = 0;
      for (int r = 0; r < SW_CL_CTAS; ++r) any_over |= *cluster.map_shared_rank(&flags[2], r);
      if (l == 0 && flags[3] > ldcg(&ctl->max_nbr_seen)) atomicMax(&ctl->max_nbr_seen, flags[3]);
      if (leader) ctl->rebuilds += 1;
      if (any_over) {
        if (leader) atomicExch(&ctl->overflow, 1);
        stop = true;
        break;                                       // uniform over the cluster
      }
      need = false;
    }
    const bool last = (s == R.n_steps - 1);
    bool moved = false;
    if (act) {
      double Fx = 0.0, Fy = 0.0;
      for (int k0 = 0; k0 < cnt; k0 += 4) {
        int en[4];
        double xj[4], yj[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          en[u] = (k0 + u < cnt) ? snb[(k0 + u) * SW_CL_THREADS + l] : ((rank << 16) | l);
          const int r = (en[u] >> 16) & 0xFF, lj = en[u] & 0xFFFF;
          xj[u] = cluster.map_shared_rank(sx + cur * SW_CL_THREADS, r)[lj];
          yj[u] = cluster.map_shared_rank(sy + cur * SW_CL_THREADS, r)[lj];
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (k0 + u < cnt) bd2d_pair(P, xi, yi, ti, en[u] >> SW_TYPE_SHIFT, xj[u], yj[u], Fx, Fy);
      }
      const Bd2dSelf sf = bd2d_self_forces(P, ti, xi, yi, ex, ey, fsw);
      Fx += sf.wx; Fy += sf.wy;
      Fx += sf.sx; Fy += sf.sy;
      if (A.fex) { Fx += fxe; Fy += fye; }
      bd2d_update<NOISE>(iso, cc, dt, Fx, Fy, ex, ey, w, g0, g1, g3, xi, yi, th, vx_l, vy_l);
      sx[nxt * SW_CL_THREADS + l] = xi; sy[nxt * SW_CL_THREADS + l] = yi;
      const double ddx = xi - xr, ddy = yi - yr;
      moved = (ddx * ddx + ddy * ddy) > lim2;
    }
    const int any = __syncthreads_or(moved);
    if (l == 0) flags[nxt] = any;
    cluster.barrier_arrive();                        // release: the new positions and the flag
    if (act && !last) local_inputs(R.step0 + (u64)(s + 1));
    cluster.barrier_wait();                          // acquire
    int nd = 0;
    for (int r = 0; r < SW_CL_CTAS; ++r) nd |= *cluster.map_shared_rank(&flags[nxt], r);
    need = nd != 0;
  }
  if (act) {
    A.x[0][i] = xi; A.y[0][i] = yi; A.theta[i] = th;
    if (s > 0) {
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
    if (stop) ctl->halted = 1;
    ctl->rebuild_flag[0] = 0; ctl->rebuild_flag[1] = 0; ctl->rebuild_flag[2] = 0;
  }
  cluster.sync();                                    // nobody leaves while a neighbour may still read its shared memory
}
