// swarm_engine.cu -- Brownian-dynamics engine kernels (sm_100a) and their C-ABI.
//
// Replaces, for the SwarmRL rollout hot path, what the reference obtains from ESPResSo through
// `system.integrator.run` (swarmrl/engine/espresso.py:1304-1306), `_remove_overlap` (:1161-1201) and
// the per-particle property writes of `manage_forces` (:1203-1249).  See DESIGN.md for the layout
// and the roofline of each kernel.
//
// Design (B200-first, not a translation of ESPResSo's cell system):
//  * SoA fp64 state; the 100k..1M-particle working set is L2-resident, so the integrator loop is
//    bound by launch / grid-synchronisation latency and fp64+integer issue, not by HBM.  One
//    COOPERATIVE PERSISTENT kernel therefore advances a whole chunk of steps: one launch per
//    `integrator.run`, one grid barrier per step, positions double-buffered in L2.
//  * Neighbours: per-particle Verlet lists (ascending neighbour index => a deterministic force
//    summation order identical to the oracle's), rebuilt INSIDE the persistent kernel from a
//    linked-cell hash (atomicExch insertion, 3 barriers) whenever any particle has moved more than
//    skin/2 since the last build -- the criterion ESPResSo's Verlet lists use.
//  * Noise: one Philox4x64-10 block per particle-step (4 normals: x, y, (z), theta).
#include <cooperative_groups.h>
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <utility>

#include "swarm_common.cuh"
#include "swarm_math.cuh"

#define SW_MAXCAP 128         // hard ceiling of the Verlet capacity
#define SW_IDX_MASK 0x0FFFFFFF
#define SW_TYPE_SHIFT 28
#define SW_THREADS 256

enum { MODE_BD = 0, MODE_SD = 1, MODE_LV = 2 };

struct RunArgs {
  long long n_steps;
  u64 step0;
  int mode;
  int force_rebuild;
  double sd_gamma, sd_maxd;
  int particles_per_cta;  // resident kernel only
  long long steps_base;   // steps completed by the earlier sub-launches of this swarm_run (sw_run_common splits long runs)
  // MODE_LV (velocity Verlet): iteration 0 of the launch is the force recalculation at the start of a chunk
  // (no closing half kick); the last iteration only closes the final step (no drift)
  int lv_first, lv_close;
  int prof_cta, prof_warp;   // SW_TIMING builds: whose clock64 sums go to Ctl.tprof (SWARM_B200_PROF_CTA / _WARP)
};

// ---------------------------------------------------------------------------------------------
// error plumbing
// ---------------------------------------------------------------------------------------------
static thread_local std::string g_err;
void sw_set_error(const std::string &msg) { g_err = msg; }
int sw_cuda_fail(cudaError_t e, const char *what) {
  g_err = std::string("CUDA error: ") + cudaGetErrorString(e) + " in " + what;
  return SWARM_ERR_CUDA;
}
extern "C" const char *swarm_last_error(void) { return g_err.c_str(); }
extern "C" int swarm_abi_version(void) { return SWARM_ABI_VERSION; }

// ---------------------------------------------------------------------------------------------
// grid barrier (all CTAs co-resident: cooperative launch).  ctl->barrier is zeroed by a memset
// node before every launch; `epoch` counts the arrivals this CTA expects.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void grid_barrier(unsigned int *counter, unsigned int &epoch) {
  __syncthreads();
  if (threadIdx.x == 0) {
    epoch += gridDim.x;
    __threadfence();
    atomicAdd(counter, 1u);
    while (*((volatile unsigned int *)counter) < epoch) {
    }
    __threadfence();
  }
  __syncthreads();
}

// ---------------------------------------------------------------------------------------------
// cell hash of particle i (per-environment grid; global cell id includes the environment)
// ---------------------------------------------------------------------------------------------
template <int DIM>
__device__ __forceinline__ void cell_coords(const DevParams &P, double x, double y, double z, int c[3]) {
  c[0] = sw_cell_coord(x, P.inv_box[0], P.nc[0], P.periodic[0]);
  c[1] = sw_cell_coord(y, P.inv_box[1], P.nc[1], P.periodic[1]);
  c[2] = (DIM == 3) ? sw_cell_coord(z, P.inv_box[2], P.nc[2], P.periodic[2]) : 0;
}

__device__ __forceinline__ int cell_id(const DevParams &P, int env, const int c[3]) {
  return env * P.ncell_env + (c[2] * P.nc[1] + c[1]) * P.nc[0] + c[0];
}

// ---- rebuild phase 1: unlink the old and the new cell head ----------------------------------
template <int DIM>
__device__ __forceinline__ void phase_clear(const DevParams &P, const DevArrays &A, int cur, long long i) {
  int c[3];
  double z = (DIM == 3) ? ldcg(&A.z[cur][i]) : 0.0;
  cell_coords<DIM>(P, ldcg(&A.x[cur][i]), ldcg(&A.y[cur][i]), z, c);
  int env = (int)(i / P.env_size);
  int cid = cell_id(P, env, c);
  A.head[A.cellof[i]] = -1;
  A.head[cid] = -1;
  A.cellof[i] = cid;
}

// ---- rebuild phase 2: push onto the cell's chain --------------------------------------------
__device__ __forceinline__ void phase_insert(const DevArrays &A, long long i) {
  A.next[i] = atomicExch(&A.head[A.cellof[i]], (int)i);
}

// ---- rebuild phase 3: Verlet list of particle i, ascending neighbour index --------------------
template <int DIM>
__device__ __forceinline__ void phase_search(const DevParams &P, const DevArrays &A, int cur, long long i) {
  const double xi = ldcg(&A.x[cur][i]), yi = ldcg(&A.y[cur][i]);
  const double zi = (DIM == 3) ? ldcg(&A.z[cur][i]) : 0.0;
  const int ti = P.sp_type[A.species[i]];
  const int env = (int)(i / P.env_size);
  int c0[3];
  cell_coords<DIM>(P, xi, yi, zi, c0);
  int loc[SW_MAXCAP];
  int cnt = 0;
  bool over = false;
  // stencil offsets per axis: {-1,0,1}; {0,1} when only two cells; {0} when one
  int lo[3], hi[3];
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    if (a >= DIM || P.nc[a] == 1) { lo[a] = 0; hi[a] = 0; }
    else if (P.nc[a] == 2) { lo[a] = 0; hi[a] = 1; }
    else { lo[a] = -1; hi[a] = 1; }
  }
  for (int dz = lo[2]; dz <= hi[2]; ++dz)
    for (int dy = lo[1]; dy <= hi[1]; ++dy)
      for (int dx = lo[0]; dx <= hi[0]; ++dx) {
        int d[3] = {dx, dy, dz}, cc[3];
        bool ok = true;
#pragma unroll
        for (int a = 0; a < 3; ++a) {
          int v = c0[a] + d[a];
          if (v < 0 || v >= P.nc[a]) {
            if (P.periodic[a] && a < DIM) v = (v + P.nc[a]) % P.nc[a];
            else ok = false;
          }
          cc[a] = v;
        }
        if (!ok) continue;
        int j = ldcg(&A.head[cell_id(P, env, cc)]);
        while (j >= 0) {
          if (j != (int)i) {
            int tj = P.sp_type[A.species[j]];
            double ddx = sw_min_image(xi - ldcg(&A.x[cur][j]), P.box[0], P.inv_box[0], P.periodic[0]);
            double ddy = sw_min_image(yi - ldcg(&A.y[cur][j]), P.box[1], P.inv_box[1], P.periodic[1]);
            double ddz = (DIM == 3) ? sw_min_image(zi - ldcg(&A.z[cur][j]), P.box[2], P.inv_box[2], P.periodic[2]) : 0.0;
            double rl = __dadd_rn(P.wca_cut[ti][tj], P.skin);
            if (sw_r2(ddx, ddy, ddz) < __dmul_rn(rl, rl)) {
              if (cnt < A.cap) {
                // insertion sort by neighbour index (entries carry the type in the top bits)
                int e = j | (tj << SW_TYPE_SHIFT);
                int k = cnt++;
                while (k > 0 && (loc[k - 1] & SW_IDX_MASK) > j) { loc[k] = loc[k - 1]; --k; }
                loc[k] = e;
              } else {
                over = true;
              }
            }
          }
          j = ldcg(&A.next[j]);
        }
      }
  if (A.okey) {
    // spatially sorted slots: the summation order stays "ascending particle id", whatever slot a particle sits in
    // (keys re-read per comparison: the lists are short and a rebuild is rare; no second local array)
    for (int k = 1; k < cnt; ++k) {
      const int e = loc[k];
      const long long key = A.okey[e & SW_IDX_MASK];
      int m = k;
      while (m > 0 && A.okey[loc[m - 1] & SW_IDX_MASK] > key) { loc[m] = loc[m - 1]; --m; }
      loc[m] = e;
    }
  }
  for (int k = 0; k < cnt; ++k) A.nbr[(long long)k * A.n_pad + i] = loc[k];
  A.nbr_cnt[i] = cnt;
  A.xr[i] = xi; A.yr[i] = yi;
  if (DIM == 3) A.zr[i] = zi;
  if (over) atomicExch(&A.ctl->overflow, 1);
  if (cnt > ldcg(&A.ctl->max_nbr_seen)) atomicMax(&A.ctl->max_nbr_seen, cnt);
}

// ---------------------------------------------------------------------------------------------
// total force on particle i at buffer `cur`: WCA over the Verlet list (ascending j) + walls +
// swim + external  (oracle: orc_forces)
// ---------------------------------------------------------------------------------------------
template <int DIM>
__device__ __forceinline__ void dev_force(const DevParams &P, const DevArrays &A, int cur, long long i, double xi,
                                          double yi, double zi, int ti, const double e[3], bool with_swim,
                                          double F[3]) {
  double Fx = 0.0, Fy = 0.0, Fz = 0.0;
  const int cnt = A.nbr_cnt[i];
  for (int k = 0; k < cnt; ++k) {
    int en = A.nbr[(long long)k * A.n_pad + i];
    int j = en & SW_IDX_MASK, tj = en >> SW_TYPE_SHIFT;
    double dx = sw_min_image(xi - ldcg(&A.x[cur][j]), P.box[0], P.inv_box[0], P.periodic[0]);
    double dy = sw_min_image(yi - ldcg(&A.y[cur][j]), P.box[1], P.inv_box[1], P.periodic[1]);
    double dz = (DIM == 3) ? sw_min_image(zi - ldcg(&A.z[cur][j]), P.box[2], P.inv_box[2], P.periodic[2]) : 0.0;
    double r2 = sw_r2(dx, dy, dz);
    double cut = P.wca_cut[ti][tj];
    if (r2 < __dmul_rn(cut, cut) && r2 > 0.0) {
      double fac = sw_wca_factor(r2, P.wca_sigma[ti][tj], P.wca_eps[ti][tj]);
      Fx += fac * dx; Fy += fac * dy; Fz += fac * dz;
    }
  }
  if (P.walls) {
    const double pos[3] = {xi, yi, zi};
    double *Fp[3] = {&Fx, &Fy, &Fz};
    const double cut = P.wall_cut[ti];
#pragma unroll
    for (int a = 0; a < DIM; ++a) {
      double d = pos[a];
      if (d < cut && d > 0.0) *Fp[a] += sw_wca_factor(d * d, P.wall_sigma[ti], P.wall_eps[ti]) * d;
      double d2 = P.box[a] - pos[a];
      if (d2 < cut && d2 > 0.0) *Fp[a] -= sw_wca_factor(d2 * d2, P.wall_sigma[ti], P.wall_eps[ti]) * d2;
    }
  }
  if (with_swim) {
    double f = A.f_swim[i];
    Fx += f * e[0]; Fy += f * e[1]; Fz += f * e[2];
  }
  if (A.fex) { Fx += A.fex[i]; Fy += A.fey[i]; Fz += A.fez[i]; }
  F[0] = Fx; F[1] = Fy; F[2] = Fz;
}

__device__ __forceinline__ double clampd(double v, double lo, double hi) { return v < lo ? lo : (v > hi ? hi : v); }

// ---------------------------------------------------------------------------------------------
// one integrator step of particle i: read buffer cur, write buffer cur^1  (oracle: orc_bd_steps /
// orc_sd_steps).  Returns true when the particle has moved more than skin/2 since the last build.
// ---------------------------------------------------------------------------------------------
template <int DIM, bool NOISE>
__device__ __forceinline__ bool phase_step(const DevParams &P, const DevArrays &A, const RunArgs &R, int cur, u64 step,
                                           bool last, long long i) {
  const int nxt = cur ^ 1;
  const int sp = A.species[i];
  const u64 pid = sw_pid(A, i);
  const int ti = P.sp_type[sp];
  double xi = ldcg(&A.x[cur][i]), yi = ldcg(&A.y[cur][i]);
  double zi = (DIM == 3) ? ldcg(&A.z[cur][i]) : 0.0;
  const double dt = P.dt, kT = P.kT;
  double F[3];
  double xn, yn, zn = 0.0;

  if (DIM == 2) {
    double th = A.theta[i];
    double e[3];
    sincos(th, &e[1], &e[0]);
    e[2] = 0.0;
    dev_force<2>(P, A, cur, i, xi, yi, 0.0, ti, e, true, F);
    const double Tz = A.tqz[i];
    if (R.mode == MODE_SD) {
      xn = xi + clampd(R.sd_gamma * F[0], -R.sd_maxd, R.sd_maxd);
      yn = yi + clampd(R.sd_gamma * F[1], -R.sd_maxd, R.sd_maxd);
      double l = fabs(R.sd_gamma * Tz);
      if (l > 0.0) A.theta[i] = th + (Tz > 0 ? 1.0 : -1.0) * clampd(l, -R.sd_maxd, R.sd_maxd);
    } else if (R.mode == MODE_LV) {
      // Langevin thermostat + velocity Verlet (oracle: orc_lv_steps): this force evaluation closes the previous
      // step (second half kick) and opens the next one (first half kick + drift)
      const double gt1 = P.sp_gt[sp][1], gt2 = P.sp_gt[sp][2], gr0 = P.sp_gr[sp][0];
      const double m = P.sp_mass[sp], hk = 0.5 * dt;
      double u[4] = {0.0, 0.0, 0.0, 0.0};
      if (NOISE) swm_uniforms(P.seed, pid, step, 4, u);
      double vx = A.vx[i], vy = A.vy[i], w = A.wz[i];
      double fx = F[0], fy = F[1];
      if (gt1 == gt2) {
        fx -= gt2 * vx; fy -= gt2 * vy;
        if (NOISE) {
          double pref = sqrt(24.0 * kT * gt2 / dt);
          fx += pref * u[0]; fy += pref * u[1];
        }
      } else {
        double vpar = vx * e[0] + vy * e[1], vper = -vx * e[1] + vy * e[0];
        double fpar = -(gt2 * vpar), fper = -(gt1 * vper);
        if (NOISE) {
          fpar += sqrt(24.0 * kT * gt2 / dt) * u[0];
          fper += sqrt(24.0 * kT * gt1 / dt) * u[1];
        }
        fx += fpar * e[0] - fper * e[1]; fy += fpar * e[1] + fper * e[0];
      }
      double tq = Tz - gr0 * w;
      if (NOISE) tq += sqrt(24.0 * kT * gr0 / dt) * u[3];
      const double ax = fx / m, ay = fy / m, al = tq / P.sp_rin[sp][0];
      const bool first = R.lv_first && step == R.step0, close = R.lv_close && last;
      if (!first) { vx += hk * ax; vy += hk * ay; w += hk * al; }
      xn = xi; yn = yi;
      if (!close) {
        vx += hk * ax; vy += hk * ay; w += hk * al;
        xn = xi + dt * vx; yn = yi + dt * vy;
        A.theta[i] = th + dt * w;
      }
      A.vx[i] = vx; A.vy[i] = vy; A.vz[i] = 0.0;
      A.wx[i] = 0.0; A.wy[i] = 0.0; A.wz[i] = w;
    } else {
      const double gt1 = P.sp_gt[sp][1], gt2 = P.sp_gt[sp][2], gr0 = P.sp_gr[sp][0];
      double g[4] = {0.0, 0.0, 0.0, 0.0};
      if (NOISE) {
        swm_gaussians(P.seed, pid, step, 0, g);
        // the rotational normal of an odd step is the cosine of the even step's pair 1 (swm_gaussians_xyt, swarm_math.cuh)
        if (step & 1ULL) { double ge[4]; swm_gaussians(P.seed, pid, step - 1ULL, 0, ge); g[3] = ge[2]; }
      }
      double vx, vy;
      if (gt1 == gt2) {
        vx = F[0] / gt2; vy = F[1] / gt2;
        xn = xi + vx * dt; yn = yi + vy * dt;
        if (NOISE) {
          double sg = sqrt(2.0 * kT * dt / gt2);
          xn += sg * g[0]; yn += sg * g[1];
        }
      } else {
        double fpar = F[0] * e[0] + F[1] * e[1], fper = -F[0] * e[1] + F[1] * e[0];
        double vpar = fpar / gt2, vper = fper / gt1;
        vx = vpar * e[0] - vper * e[1]; vy = vpar * e[1] + vper * e[0];
        xn = xi + vx * dt; yn = yi + vy * dt;
        if (NOISE) {
          double spar = sqrt(2.0 * kT * dt / gt2) * g[0], sper = sqrt(2.0 * kT * dt / gt1) * g[1];
          xn += spar * e[0] - sper * e[1]; yn += spar * e[1] + sper * e[0];
        }
      }
      double w = Tz / gr0;
      double thn = th + w * dt;
      if (NOISE) thn += sqrt(2.0 * kT * dt / gr0) * g[3];
      A.theta[i] = thn;
      if (last) {
        double gv[4] = {0, 0, 0, 0}, gw[4] = {0, 0, 0, 0};
        double sv = 0.0, sw = 0.0;
        if (NOISE) {
          swm_gaussians(P.seed, pid, step, 2, gv);
          swm_gaussians(P.seed, pid, step, 3, gw);
          sv = sqrt(kT / P.sp_mass[sp]);
          sw = sqrt(kT / P.sp_rin[sp][0]);
        }
        A.vx[i] = vx + sv * gv[0]; A.vy[i] = vy + sv * gv[1]; A.vz[i] = 0.0;
        A.wx[i] = 0.0; A.wy[i] = 0.0; A.wz[i] = w + sw * gw[0];
      }
    }
  } else {
    double q[4] = {A.q0[i], A.q1[i], A.q2[i], A.q3[i]}, Rm[3][3];
    sw_quat_to_mat(q, Rm);
    double e[3] = {Rm[0][2], Rm[1][2], Rm[2][2]};
    dev_force<3>(P, A, cur, i, xi, yi, zi, ti, e, true, F);
    const double T[3] = {A.tqx[i], A.tqy[i], A.tqz[i]};
    if (R.mode == MODE_SD) {
      xn = xi + clampd(R.sd_gamma * F[0], -R.sd_maxd, R.sd_maxd);
      yn = yi + clampd(R.sd_gamma * F[1], -R.sd_maxd, R.sd_maxd);
      zn = zi + clampd(R.sd_gamma * F[2], -R.sd_maxd, R.sd_maxd);
      double dq[3];
#pragma unroll
      for (int b = 0; b < 3; ++b) dq[b] = R.sd_gamma * (Rm[0][b] * T[0] + Rm[1][b] * T[1] + Rm[2][b] * T[2]);
      double l = sqrt(dq[0] * dq[0] + dq[1] * dq[1] + dq[2] * dq[2]);
      if (l > 0.0) {
        double ax[3] = {dq[0] / l, dq[1] / l, dq[2] / l};
        sw_quat_rotate_body(q, ax, clampd(l, -R.sd_maxd, R.sd_maxd));
        A.q0[i] = q[0]; A.q1[i] = q[1]; A.q2[i] = q[2]; A.q3[i] = q[3];
      }
    } else if (R.mode == MODE_LV) {
      const double *gt = P.sp_gt[sp], *gr = P.sp_gr[sp];
      const double m = P.sp_mass[sp], hk = 0.5 * dt;
      double u[4] = {0.0, 0.0, 0.0, 0.0}, ur[4] = {0.0, 0.0, 0.0, 0.0};
      if (NOISE) { swm_uniforms(P.seed, pid, step, 4, u); swm_uniforms(P.seed, pid, step, 5, ur); }
      double v[3] = {A.vx[i], A.vy[i], A.vz[i]};
      const double wl[3] = {A.wx[i], A.wy[i], A.wz[i]};
      if (gt[0] == gt[1] && gt[1] == gt[2]) {
        const double pref = NOISE ? sqrt(24.0 * kT * gt[0] / dt) : 0.0;
#pragma unroll
        for (int a = 0; a < 3; ++a) {
          F[a] -= gt[0] * v[a];
          if (NOISE) F[a] += pref * u[a];
        }
      } else {
        double fb[3];
#pragma unroll
        for (int b = 0; b < 3; ++b) {
          double vb = Rm[0][b] * v[0] + Rm[1][b] * v[1] + Rm[2][b] * v[2];
          fb[b] = -(gt[b] * vb);
          if (NOISE) fb[b] += sqrt(24.0 * kT * gt[b] / dt) * u[b];
        }
#pragma unroll
        for (int a = 0; a < 3; ++a) F[a] += Rm[a][0] * fb[0] + Rm[a][1] * fb[1] + Rm[a][2] * fb[2];
      }
      double wb[3], alb[3];
#pragma unroll
      for (int b = 0; b < 3; ++b) {
        wb[b] = Rm[0][b] * wl[0] + Rm[1][b] * wl[1] + Rm[2][b] * wl[2];
        double tb = Rm[0][b] * T[0] + Rm[1][b] * T[1] + Rm[2][b] * T[2];
        tb -= gr[b] * wb[b];
        if (NOISE) tb += sqrt(24.0 * kT * gr[b] / dt) * ur[b];
        alb[b] = tb / P.sp_rin[sp][b];
      }
      const bool first = R.lv_first && step == R.step0, close = R.lv_close && last;
      if (!first) {
#pragma unroll
        for (int a = 0; a < 3; ++a) { v[a] += hk * (F[a] / m); wb[a] += hk * alb[a]; }
      }
      xn = xi; yn = yi; zn = zi;
      if (!close) {
#pragma unroll
        for (int a = 0; a < 3; ++a) { v[a] += hk * (F[a] / m); wb[a] += hk * alb[a]; }
        xn = xi + dt * v[0]; yn = yi + dt * v[1]; zn = zi + dt * v[2];
        double dphi[3] = {wb[0] * dt, wb[1] * dt, wb[2] * dt};
        double mg = sqrt(dphi[0] * dphi[0] + dphi[1] * dphi[1] + dphi[2] * dphi[2]);
        if (mg > 0.0) {
          double ax[3] = {dphi[0] / mg, dphi[1] / mg, dphi[2] / mg};
          sw_quat_rotate_body(q, ax, mg);
          sw_quat_to_mat(q, Rm);
        }
        A.q0[i] = q[0]; A.q1[i] = q[1]; A.q2[i] = q[2]; A.q3[i] = q[3];
      }
      A.vx[i] = v[0]; A.vy[i] = v[1]; A.vz[i] = v[2];
      A.wx[i] = Rm[0][0] * wb[0] + Rm[0][1] * wb[1] + Rm[0][2] * wb[2];
      A.wy[i] = Rm[1][0] * wb[0] + Rm[1][1] * wb[1] + Rm[1][2] * wb[2];
      A.wz[i] = Rm[2][0] * wb[0] + Rm[2][1] * wb[1] + Rm[2][2] * wb[2];
    } else {
      const double *gt = P.sp_gt[sp], *gr = P.sp_gr[sp];
      double g[4] = {0.0, 0.0, 0.0, 0.0};
      if (NOISE) swm_gaussians(P.seed, pid, step, 0, g);
      const bool aniso = (gt[0] != gt[1] || gt[1] != gt[2]);
      double v[3];
      if (!aniso) {
#pragma unroll
        for (int a = 0; a < 3; ++a) v[a] = F[a] / gt[0];
      } else {
        double vb[3];
#pragma unroll
        for (int b = 0; b < 3; ++b) vb[b] = (Rm[0][b] * F[0] + Rm[1][b] * F[1] + Rm[2][b] * F[2]) / gt[b];
#pragma unroll
        for (int a = 0; a < 3; ++a) v[a] = Rm[a][0] * vb[0] + Rm[a][1] * vb[1] + Rm[a][2] * vb[2];
      }
      double pn[3] = {xi + v[0] * dt, yi + v[1] * dt, zi + v[2] * dt};
      if (NOISE) {
        if (!aniso) {
          double sg = sqrt(2.0 * kT * dt / gt[0]);
#pragma unroll
          for (int a = 0; a < 3; ++a) pn[a] += sg * g[a];
        } else {
          double db[3];
#pragma unroll
          for (int b = 0; b < 3; ++b) db[b] = sqrt(2.0 * kT * dt / gt[b]) * g[b];
#pragma unroll
          for (int a = 0; a < 3; ++a) pn[a] += Rm[a][0] * db[0] + Rm[a][1] * db[1] + Rm[a][2] * db[2];
        }
      }
      xn = pn[0]; yn = pn[1]; zn = pn[2];
      double wb[3], dphi[3];
#pragma unroll
      for (int b = 0; b < 3; ++b) {
        wb[b] = (Rm[0][b] * T[0] + Rm[1][b] * T[1] + Rm[2][b] * T[2]) / gr[b];
        dphi[b] = wb[b] * dt;
      }
      double m = sqrt(dphi[0] * dphi[0] + dphi[1] * dphi[1] + dphi[2] * dphi[2]);
      if (m > 0.0) {
        double ax[3] = {dphi[0] / m, dphi[1] / m, dphi[2] / m};
        sw_quat_rotate_body(q, ax, m);
      }
      if (NOISE) {
        double g1[4];
        swm_gaussians(P.seed, pid, step, 1, g1);
        double dr[3];
#pragma unroll
        for (int b = 0; b < 3; ++b) dr[b] = sqrt(2.0 * kT * dt / gr[b]) * g1[b];
        double mr = sqrt(dr[0] * dr[0] + dr[1] * dr[1] + dr[2] * dr[2]);
        if (mr > 0.0) {
          double ax[3] = {dr[0] / mr, dr[1] / mr, dr[2] / mr};
          sw_quat_rotate_body(q, ax, mr);
        }
      }
      A.q0[i] = q[0]; A.q1[i] = q[1]; A.q2[i] = q[2]; A.q3[i] = q[3];
      if (last) {
        double gv[4] = {0, 0, 0, 0}, gw[4] = {0, 0, 0, 0};
        double sv = 0.0;
        if (NOISE) {
          swm_gaussians(P.seed, pid, step, 2, gv);
          swm_gaussians(P.seed, pid, step, 3, gw);
          sv = sqrt(kT / P.sp_mass[sp]);
        }
        A.vx[i] = v[0] + sv * gv[0]; A.vy[i] = v[1] + sv * gv[1]; A.vz[i] = v[2] + sv * gv[2];
        double wbn[3];
#pragma unroll
        for (int b = 0; b < 3; ++b) wbn[b] = wb[b] + (NOISE ? sqrt(kT / P.sp_rin[sp][b]) * gw[b] : 0.0);
        A.wx[i] = Rm[0][0] * wbn[0] + Rm[0][1] * wbn[1] + Rm[0][2] * wbn[2];
        A.wy[i] = Rm[1][0] * wbn[0] + Rm[1][1] * wbn[1] + Rm[1][2] * wbn[2];
        A.wz[i] = Rm[2][0] * wbn[0] + Rm[2][1] * wbn[1] + Rm[2][2] * wbn[2];
      }
    }
  }
  A.x[nxt][i] = xn; A.y[nxt][i] = yn;
  if (DIM == 3) A.z[nxt][i] = zn;
  double ddx = xn - A.xr[i], ddy = yn - A.yr[i], ddz = (DIM == 3) ? zn - A.zr[i] : 0.0;
  const double lim = 0.5 * P.skin;
  return (ddx * ddx + ddy * ddy + ddz * ddz) > lim * lim;
}

template <int DIM>
__device__ __forceinline__ bool moved_since_build(const DevParams &P, const DevArrays &A, int cur, long long i) {
  double ddx = ldcg(&A.x[cur][i]) - A.xr[i], ddy = ldcg(&A.y[cur][i]) - A.yr[i];
  double ddz = (DIM == 3) ? ldcg(&A.z[cur][i]) - A.zr[i] : 0.0;
  const double lim = 0.5 * P.skin;
  double d2 = ddx * ddx + ddy * ddy + ddz * ddz;
  return !(d2 <= lim * lim);  // NaN counts as moved
}

// ---------------------------------------------------------------------------------------------
// THE hot kernel: cooperative persistent chunk of n_steps integrator steps.
// ---------------------------------------------------------------------------------------------
template <int DIM, bool NOISE>
__global__ void __launch_bounds__(SW_THREADS) k_persist(const __grid_constant__ DevParams P, DevArrays A, RunArgs R) {
  Ctl *ctl = A.ctl;
  if (ldcg(&ctl->halted)) return;
  unsigned int epoch = 0;
  // grid-stride ownership: thread t owns particles t, t + T, t + 2T, ... (balanced to one particle)
  const long long lo = (long long)blockIdx.x * blockDim.x, hi = sw_live_n(A);
  const long long gstride = (long long)gridDim.x * blockDim.x;
  const bool leader = (blockIdx.x == 0 && threadIdx.x == 0);

  bool need = true;
  if (!R.force_rebuild) {
    bool moved = false;
    for (long long i = lo + threadIdx.x; i < hi; i += gstride) moved |= moved_since_build<DIM>(P, A, 0, i);
    if (moved) ctl->rebuild_flag[0] = 1;
    grid_barrier(&ctl->barrier, epoch);
    need = ldcg(&ctl->rebuild_flag[0]) != 0;
  }

  long long s = 0;
  for (; s < R.n_steps; ++s) {
    const int cur = (int)(s & 1);
    if (s > 0) need = ldcg(&ctl->rebuild_flag[s % 3]) != 0;
    if (need) {
      for (long long i = lo + threadIdx.x; i < hi; i += gstride) phase_clear<DIM>(P, A, cur, i);
      grid_barrier(&ctl->barrier, epoch);
      for (long long i = lo + threadIdx.x; i < hi; i += gstride) phase_insert(A, i);
      grid_barrier(&ctl->barrier, epoch);
      for (long long i = lo + threadIdx.x; i < hi; i += gstride) phase_search<DIM>(P, A, cur, i);
      if (leader) ctl->rebuilds += 1;
      grid_barrier(&ctl->barrier, epoch);
      if (ldcg(&ctl->overflow)) break;  // uniform: every CTA reads the same value after the barrier
    }
    const bool last = (s == R.n_steps - 1);
    bool moved = false;
    for (long long i = lo + threadIdx.x; i < hi; i += gstride)
      moved |= phase_step<DIM, NOISE>(P, A, R, cur, R.step0 + (u64)s, last, i);
    if (moved) ctl->rebuild_flag[(s + 1) % 3] = 1;
    if (leader) ctl->rebuild_flag[(s + 2) % 3] = 0;
    grid_barrier(&ctl->barrier, epoch);
  }
  // results live in buffer (s & 1); the caller's arrays are buffer 0
  if (s & 1) {
    for (long long i = lo + threadIdx.x; i < hi; i += gstride) {
      A.x[0][i] = ldcg(&A.x[1][i]); A.y[0][i] = ldcg(&A.y[1][i]);
      if (DIM == 3) A.z[0][i] = ldcg(&A.z[1][i]);
    }
  }
  if (leader) {
    ctl->steps_done = R.steps_base + s;
    if (s < R.n_steps) ctl->halted = 1;
    ctl->rebuild_flag[0] = 0; ctl->rebuild_flag[1] = 0; ctl->rebuild_flag[2] = 0;
  }
}

// ---------------------------------------------------------------------------------------------
// Split grid barrier: arrive, do independent work, wait.  The Brownian step of particle i at step
// s+1 needs its neighbours' positions after step s (a grid-wide dependency) but its NOISE depends on
// (seed, i, s+1) only -- so the Philox + Box-Muller work of the next step, more than half of a
// step's instructions, is issued between arrive and wait and hides the barrier's L2 round trips.
// ---------------------------------------------------------------------------------------------
// 64-bit counter: low word = arrivals, high word = number of CTA-steps that reported a particle beyond
// skin/2 (the Verlet rebuild criterion travels with the barrier instead of costing another L2 round trip).
// red.release / ld.acquire at gpu scope: no fence.sc (what __threadfence() is) on the critical path.
// Centralised barrier with a separate release word: arrivals are atomics on `bar64`; the LAST arriver
// publishes the final counter value to `rel64` (own 128-byte line) and everybody polls only that word,
// so the polling traffic never queues in front of the arrivals at the L2 slice.
__device__ __forceinline__ void grid_arrive64(unsigned long long *counter, unsigned int &epoch, bool cta_moved) {
  if (threadIdx.x == 0) {
    epoch += gridDim.x;
    const unsigned long long inc = 1ULL | (cta_moved ? (1ULL << 32) : 0ULL);
    unsigned long long old;
    asm volatile("atom.acq_rel.gpu.global.add.u64 %0, [%1], %2;" : "=l"(old) : "l"(counter), "l"(inc) : "memory");
    if ((unsigned int)old + 1u == epoch) {
      const unsigned long long fin = old + inc;
      asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(counter + 16), "l"(fin) : "memory");
    }
  }
}
__device__ __forceinline__ unsigned long long grid_poll64(unsigned long long *counter, unsigned int epoch) {
  unsigned long long v;
  for (;;) {
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(counter + 16) : "memory");
    if ((unsigned int)v >= epoch) break;
#ifdef SW_POLL_SLEEP
    __nanosleep(SW_POLL_SLEEP);
#endif
  }
  return v;
}
// full barrier on the same counter (rebuild phases)
__device__ __forceinline__ void grid_barrier64(unsigned long long *counter, unsigned int &epoch) {
  __syncthreads();
  grid_arrive64(counter, epoch, false);
  if (threadIdx.x == 0) grid_poll64(counter, epoch);
  __syncthreads();
}

#define SW_RES_MAXT 736  // resident kernel: at most 23 warps per SM (22 particle warps = 104k particles on 148 SMs): 88 registers each
#define SW_RES_MAXREG 88  // 65536 registers / 736 threads, rounded down to the allocation unit of 8
#define SW_RES_NBR 4  // neighbour entries of the Verlet list kept in registers by the resident kernel

#include "swarm_bd2d.cuh"  // k_bd2d_resident: the hot kernel of the 2-D rollout
#include "swarm_sort.cuh"  // spatial slot order (cell hash, counting sort, cell-start scan, permutation)
#include "swarm_tiled.cuh"
#include "swarm_cluster.cuh" // k_bd2d_tiled: the same chunk on spatial CTA tiles, neighbourhood barrier instead of the grid barrier

// ---------------------------------------------------------------------------------------------
// one-launch-per-phase variant (SWARM_FLAG_NO_PERSISTENT): same device functions, kernel
// boundaries instead of grid barriers.  Used as a cross-check of the persistent kernel.
// ---------------------------------------------------------------------------------------------
template <int DIM>
__global__ void k_np_check(const __grid_constant__ DevParams P, DevArrays A, int force, long long steps_base) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (ldcg(&A.ctl->halted)) return;
  if (i == 0) A.ctl->steps_done = steps_base;
  if (force) { if (i == 0) A.ctl->rebuild_flag[0] = 1; return; }
  if (i < A.n && moved_since_build<DIM>(P, A, 0, i)) A.ctl->rebuild_flag[0] = 1;
}
template <int DIM>
__global__ void k_np_clear(const __grid_constant__ DevParams P, DevArrays A, long long s) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (ldcg(&A.ctl->halted) || !ldcg(&A.ctl->rebuild_flag[s % 3]) || i >= A.n) return;
  phase_clear<DIM>(P, A, (int)(s & 1), i);
}
__global__ void k_np_insert(DevArrays A, long long s) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (ldcg(&A.ctl->halted) || !ldcg(&A.ctl->rebuild_flag[s % 3]) || i >= A.n) return;
  phase_insert(A, i);
}
template <int DIM>
__global__ void k_np_search(const __grid_constant__ DevParams P, DevArrays A, long long s) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (ldcg(&A.ctl->halted) || !ldcg(&A.ctl->rebuild_flag[s % 3])) return;
  if (i == 0) A.ctl->rebuilds += 1;
  if (i >= A.n) return;
  phase_search<DIM>(P, A, (int)(s & 1), i);
}
template <int DIM, bool NOISE>
__global__ void k_np_step(const __grid_constant__ DevParams P, DevArrays A, RunArgs R, long long s) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  Ctl *ctl = A.ctl;
  if (ldcg(&ctl->halted)) return;
  if (ldcg(&ctl->overflow)) return;  // k_np_halt (next launch) latches `halted`
  if (i == 0) { ctl->rebuild_flag[(s + 2) % 3] = 0; ctl->steps_done = R.steps_base + s + 1; }
  if (i >= A.n) return;
  if (phase_step<DIM, NOISE>(P, A, R, (int)(s & 1), R.step0 + (u64)s, s == R.n_steps - 1, i))
    ctl->rebuild_flag[(s + 1) % 3] = 1;
}
__global__ void k_np_halt(DevArrays A) {
  if (blockIdx.x == 0 && threadIdx.x == 0 && A.ctl->overflow) A.ctl->halted = 1;
}
template <int DIM>
__global__ void k_np_finish(DevArrays A, long long steps_base) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  long long s = ldcg(&A.ctl->steps_done) - steps_base;   // steps of THIS sub-launch: its parity says where the result lies
  if (i == 0) { A.ctl->rebuild_flag[0] = 0; A.ctl->rebuild_flag[1] = 0; A.ctl->rebuild_flag[2] = 0; }
  if (!(s & 1) || i >= A.n) return;
  A.x[0][i] = A.x[1][i]; A.y[0][i] = A.y[1][i];
  if (DIM == 3) A.z[0][i] = A.z[1][i];
}

// ---------------------------------------------------------------------------------------------
// small utility kernels
// ---------------------------------------------------------------------------------------------
template <int DIM>
__global__ void k_forces(const __grid_constant__ DevParams P, DevArrays A, double *fx, double *fy, double *fz) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= A.n) return;
  double e[3];
  if (DIM == 2) { sincos(A.theta[i], &e[1], &e[0]); e[2] = 0.0; }
  else {
    double q[4] = {A.q0[i], A.q1[i], A.q2[i], A.q3[i]}, Rm[3][3];
    sw_quat_to_mat(q, Rm);
    e[0] = Rm[0][2]; e[1] = Rm[1][2]; e[2] = Rm[2][2];
  }
  double F[3];
  dev_force<DIM>(P, A, 0, i, A.x[0][i], A.y[0][i], (DIM == 3) ? A.z[0][i] : 0.0, P.sp_type[A.species[i]], e, true, F);
  fx[i] = F[0]; fy[i] = F[1]; fz[i] = F[2];
}

__global__ void k_directors(int n_dims, long long n, const double *theta, const double *q0, const double *q1,
                            const double *q2, const double *q3, double *ex, double *ey, double *ez) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (n_dims == 2) {
    double s, c;
    sincos(theta[i], &s, &c);
    ex[i] = c; ey[i] = s; ez[i] = 0.0;
  } else {
    double a = q0[i], b = q1[i], c = q2[i], d = q3[i];
    ex[i] = 2 * (b * d + a * c); ey[i] = 2 * (c * d - a * b); ez[i] = a * a - b * b - c * c + d * d;
  }
}

// manage_forces new_direction branch (espresso.py:1236-1249)
__global__ void k_apply_new_direction(int n_dims, long long n, const int *has_nd, const double *ndx, const double *ndy,
                                      const double *ndz, double *theta, double *q0, double *q1, double *q2, double *q3) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || !has_nd[i]) return;
  double nx = ndx[i], ny = ndy[i], nz = ndz[i];
  if (n_dims == 2) {
    double s, c;
    sincos(theta[i], &s, &c);
    double dot = nx * c + ny * s;  // old director z = 0
    dot = fmin(1.0, fmax(-1.0, dot));   // rounding may leave |dot| = 1 + 1 ulp, whose acos is NaN (= "no turn" above)
    double ang = acos(dot);
    if (ang > 1e-6) {
      // axis = (0,0,round(cross_z/|cross|)); cross = old x new
      double cx = s * nz, cy = -c * nz, cz = c * ny - s * nx;
      double cn = sqrt(cx * cx + cy * cy + cz * cz);
      // exactly antiparallel: old x new = 0 and the reference's axis is 0/0 (its round(nan) raises, which the host
      // side of manage_forces reproduces); the device never writes a NaN angle -- it turns by +pi
      double az = (cn > 0.0) ? rint(cz / cn) : 1.0;
      theta[i] += az * ang;
    }
  } else {
    // director -> quaternion, ESPResSo convert_director_to_quaternion
    double dm = sqrt(nx * nx + ny * ny + nz * nz);
    double dxy = sqrt(nx * nx + ny * ny);
    double theta2, phi2;
    const double pi = 3.14159265358979323846;
    if (dxy == 0.0) { theta2 = (nz > 0.0) ? 0.0 : 0.5 * pi; phi2 = 0.0; }
    else {
      theta2 = 0.5 * acos(nz / dm);
      phi2 = 0.5 * ((ny > 0.0) ? 1.0 : -1.0) * acos(nx / dxy) - 0.25 * pi;
    }
    double st, ct, sp, cp;
    sincos(theta2, &st, &ct);
    sincos(phi2, &sp, &cp);
    q0[i] = ct * cp; q1[i] = -st * cp; q2[i] = -st * sp; q3[i] = ct * sp;
  }
}

__global__ void k_gaussians(u64 seed, long long n, u64 step, u64 stream, double *out) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double g[4];
  swm_gaussians(seed, (u64)i, step, stream, g);
  out[4 * i + 0] = g[0]; out[4 * i + 1] = g[1]; out[4 * i + 2] = g[2]; out[4 * i + 3] = g[3];
}

template <int DIM>
__global__ void k_cell_index(const __grid_constant__ DevParams P, DevArrays A, int nc0, int nc1, int nc2, int *out) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= A.n) return;
  int cx = sw_cell_coord(A.x[0][i], P.inv_box[0], nc0, P.periodic[0]);
  int cy = sw_cell_coord(A.y[0][i], P.inv_box[1], nc1, P.periodic[1]);
  int cz = (DIM == 3) ? sw_cell_coord(A.z[0][i], P.inv_box[2], nc2, P.periodic[2]) : 0;
  out[i] = (cz * nc1 + cy) * nc0 + cx;
}

__global__ void k_interleave3(long long n, const double *a, const double *b, const double *c, double *out) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  out[3 * i] = a[i]; out[3 * i + 1] = b[i]; out[3 * i + 2] = c[i];
}
__global__ void k_deinterleave3(long long n, const double *in, double *a, double *b, double *c) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  a[i] = in[3 * i]; b[i] = in[3 * i + 1]; c[i] = in[3 * i + 2];
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
static int sw_alloc(swarm *s, void **p, size_t bytes, int fill = 0) {
  SW_CUDA(cudaMalloc(p, bytes > 0 ? bytes : 8));
  SW_CUDA(cudaMemset(*p, fill, bytes > 0 ? bytes : 8));
  s->owned.push_back(*p);
  return 0;
}

static void sw_free_owned(swarm *s) {
  for (void *p : s->owned) cudaFree(p);
  s->owned.clear();
}

static void sw_grid_setup(swarm *s) {
  // Verlet-build cell grid: cell edge >= max_cut + skin; total cells bounded
  DevParams &P = s->P;
  double rl = P.max_cut + P.skin;
  if (!(rl > 0.0)) rl = 1.0;
  long long budget = (1LL << 26) / (P.n_envs > 0 ? P.n_envs : 1);
  if (budget < 1) budget = 1;
  for (;;) {
    long long tot = 1;
    for (int a = 0; a < 3; ++a) {
      int nc = 1;
      if (a < P.n_dims) {
        double f = floor(P.box[a] / rl);
        nc = (f < 1.0) ? 1 : (f > 4096.0 ? 4096 : (int)f);
      }
      P.nc[a] = nc;
      tot *= nc;
    }
    if (tot <= budget) { P.ncell_env = (int)tot; break; }
    rl *= 1.26;  // coarsen
  }
}

extern "C" int swarm_create(const swarm_config_t *cfg, swarm_t **out) {
  if (!cfg || !out) { sw_set_error("null argument"); return SWARM_ERR_INVALID; }
  if (cfg->n_dims != 2 && cfg->n_dims != 3) { sw_set_error("Only 2d and 3d are allowed"); return SWARM_ERR_INVALID; }
  for (int a = 0; a < 3; ++a)
    if (!(cfg->box[a] > 0.0)) { sw_set_error("box lengths must be positive"); return SWARM_ERR_INVALID; }
  if (!(cfg->dt > 0.0) || cfg->kT < 0.0 || cfg->skin < 0.0) { sw_set_error("dt > 0, kT >= 0, skin >= 0 required"); return SWARM_ERR_INVALID; }
  if (cfg->n_envs < 1) { sw_set_error("n_envs must be >= 1"); return SWARM_ERR_INVALID; }
  int ndev = 0;
  SW_CUDA(cudaGetDeviceCount(&ndev));
  if (cfg->device < 0 || cfg->device >= ndev) { sw_set_error("invalid CUDA device ordinal"); return SWARM_ERR_INVALID; }
  SW_CUDA(cudaSetDevice(cfg->device));
  swarm *s = new swarm();
  s->cfg = *cfg;
  memset(&s->P, 0, sizeof(DevParams));
  memset(&s->A, 0, sizeof(DevArrays));
  DevParams &P = s->P;
  P.n_dims = cfg->n_dims;
  for (int a = 0; a < 3; ++a) {
    P.periodic[a] = cfg->periodic[a] ? 1 : 0;
    P.box[a] = cfg->box[a];
    P.inv_box[a] = 1.0 / cfg->box[a];
  }
  P.dt = cfg->dt; P.kT = cfg->kT; P.seed = cfg->seed; P.skin = cfg->skin;
  sw_key_schedule(P.seed, P.rk);
  P.n_envs = cfg->n_envs;
  s->A.cap = cfg->max_neighbors > 0 ? cfg->max_neighbors : (cfg->n_dims == 2 ? 16 : 64);
  if (s->A.cap > SW_MAXCAP) s->A.cap = SW_MAXCAP;
  s->persistent = !(cfg->flags & SWARM_FLAG_NO_PERSISTENT);
  cudaDeviceProp prop;
  SW_CUDA(cudaGetDeviceProperties(&prop, cfg->device));
  s->num_sms = prop.multiProcessorCount;
  if (!prop.cooperativeLaunch) s->persistent = false;
  void *ctl = nullptr;
  int rc = sw_alloc(s, &ctl, sizeof(Ctl));
  if (rc) { delete s; return rc; }
  s->A.ctl = (Ctl *)ctl;
  cudaError_t e = cudaMallocHost((void **)&s->ctl_host, sizeof(Ctl));
  if (e != cudaSuccess) { sw_free_owned(s); delete s; return sw_cuda_fail(e, "cudaMallocHost"); }
  memset(s->ctl_host, 0, sizeof(Ctl));
  *out = s;
  return SWARM_OK;
}

extern "C" int swarm_destroy(swarm_t *s) {
  if (!s) return SWARM_OK;
  cudaSetDevice(s->cfg.device);
  sw_free_owned(s);
  if (s->ctl_host) cudaFreeHost(s->ctl_host);
  if (s->obs_scratch) cudaFree(s->obs_scratch);
  if (s->red_scratch) cudaFree(s->red_scratch);
  if (s->field_scratch) cudaFree(s->field_scratch);
  if (s->vis_buf) cudaFree(s->vis_buf);
  if (s->dd_tickets) cudaFree(s->dd_tickets);
  delete s;
  return SWARM_OK;
}

extern "C" int swarm_set_species(swarm_t *s, int32_t n_species, const int32_t *type, const double *gamma_t,
                                 const double *gamma_r, const double *mass, const double *rinertia) {
  if (!s || n_species < 1 || n_species > SW_S) { sw_set_error("n_species out of range (1..16)"); return SWARM_ERR_INVALID; }
  DevParams &P = s->P;
  for (int k = 0; k < n_species; ++k) {
    if (type[k] < 0 || type[k] >= SW_T) { sw_set_error("particle type out of range (0..7)"); return SWARM_ERR_INVALID; }
    P.sp_type[k] = type[k];
    P.sp_mass[k] = mass[k];
    for (int a = 0; a < 3; ++a) {
      if (!(gamma_t[3 * k + a] > 0.0) || !(gamma_r[3 * k + a] > 0.0)) { sw_set_error("friction coefficients must be positive"); return SWARM_ERR_INVALID; }
      P.sp_gt[k][a] = gamma_t[3 * k + a];
      P.sp_gr[k][a] = gamma_r[3 * k + a];
      P.sp_rin[k][a] = rinertia[3 * k + a];
    }
  }
  P.n_species = n_species;
  s->species_set = true;
  s->list_valid = false;
  return SWARM_OK;
}

extern "C" int swarm_set_wca(swarm_t *s, int32_t n_types, const double *sigma, const double *eps, const double *cut) {
  if (!s || n_types < 1 || n_types > SW_T) { sw_set_error("n_types out of range (1..8)"); return SWARM_ERR_INVALID; }
  DevParams &P = s->P;
  P.n_types = n_types;
  P.max_cut = 0.0;
  for (int a = 0; a < n_types; ++a)
    for (int b = 0; b < n_types; ++b) {
      P.wca_sigma[a][b] = sigma[a * n_types + b];
      P.wca_eps[a][b] = eps[a * n_types + b];
      P.wca_cut[a][b] = cut[a * n_types + b];
      if (P.wca_cut[a][b] > P.max_cut) P.max_cut = P.wca_cut[a][b];
    }
  s->wca_set = true;
  s->list_valid = false;
  return SWARM_OK;
}

extern "C" int swarm_set_walls(swarm_t *s, int32_t enabled, const double *sigma, const double *eps, const double *cut) {
  if (!s) return SWARM_ERR_INVALID;
  DevParams &P = s->P;
  if (enabled) {
    // the arrays hold one entry per particle type: their length is only known once the WCA table is
    if (!s->wca_set) { sw_set_error("call swarm_set_wca before swarm_set_walls (it fixes the number of types)"); return SWARM_ERR_STATE; }
    if (!sigma || !eps || !cut) { sw_set_error("wall tables must be given"); return SWARM_ERR_INVALID; }
    for (int a = 0; a < P.n_types; ++a) {
      P.wall_sigma[a] = sigma[a]; P.wall_eps[a] = eps[a]; P.wall_cut[a] = cut[a];
    }
  }
  P.walls = enabled ? 1 : 0;
  return SWARM_OK;
}

extern "C" int swarm_set_thermostat(swarm_t *s, double kT, uint64_t seed) {
  if (!s || kT < 0.0) { sw_set_error("kT must be >= 0"); return SWARM_ERR_INVALID; }
  s->P.kT = kT; s->P.seed = seed;
  sw_key_schedule(seed, s->P.rk);
  return SWARM_OK;
}

extern "C" int swarm_bind_state(swarm_t *s, int64_t n, const swarm_state_dev_t *st) {
  if (!s || !st || n < 0) { sw_set_error("bad arguments"); return SWARM_ERR_INVALID; }
  if (n >= (1LL << SW_TYPE_SHIFT)) { sw_set_error("n must be < 2^28 particles per handle"); return SWARM_ERR_INVALID; }
  if (n % s->P.n_envs != 0) { sw_set_error("n must be a multiple of n_envs"); return SWARM_ERR_INVALID; }
  if (!st->x || !st->y || !st->z || !st->species || !st->vx || !st->vy || !st->vz || !st->wx || !st->wy || !st->wz) {
    sw_set_error("x, y, z, v*, w*, species must be bound"); return SWARM_ERR_INVALID;
  }
  if (s->P.n_dims == 2 && !st->theta) { sw_set_error("theta must be bound in 2-D"); return SWARM_ERR_INVALID; }
  if (s->P.n_dims == 3 && (!st->q0 || !st->q1 || !st->q2 || !st->q3)) { sw_set_error("q0..q3 must be bound in 3-D"); return SWARM_ERR_INVALID; }
  SW_CUDA(cudaSetDevice(s->cfg.device));
  // (re)allocate library scratch -- unless a previous bind / swarm_reserve already holds enough of it
  long long need = n > s->n_reserved ? n : s->n_reserved;
  long long need_pad = ((need + 31) / 32) * 32;
  if (need_pad == 0) need_pad = 32;
  DevArrays &A = s->A;
  const bool reuse = s->state_bound && A.x[1] && A.n_pad == need_pad;
  if (!reuse) {
    Ctl *ctl = s->A.ctl;
    int cap = s->A.cap;
    for (void *p : s->owned) if (p != (void *)ctl) cudaFree(p);
    s->owned.clear();
    s->owned.push_back(ctl);
    memset(&A, 0, sizeof(A));
    A.ctl = ctl; A.cap = cap;
    A.n_pad = need_pad;
    const int tmode = s->tile.mode, tint = s->tile.resort_interval;
    s->tile = TileHost();   // its arrays were in `owned`
    s->tile.mode = tmode; s->tile.resort_interval = tint;
  }
  s->tile.sorted = false;   // other particles: the slot order has to be computed again
  s->vis_prev_n = -1;
  A.n = n;
  s->st = *st;
  A.x[0] = st->x; A.y[0] = st->y; A.z[0] = st->z;
  A.theta = st->theta; A.q0 = st->q0; A.q1 = st->q1; A.q2 = st->q2; A.q3 = st->q3;
  A.vx = st->vx; A.vy = st->vy; A.vz = st->vz; A.wx = st->wx; A.wy = st->wy; A.wz = st->wz;
  A.species = st->species;
  A.gid = nullptr;
  A.n_dev = nullptr;
  s->vis_image = nullptr;
  s->P.env_size = (int)(n / s->P.n_envs);
  if (s->P.env_size < 1) s->P.env_size = 1;
  if (!reuse) {
    size_t nb = sizeof(double) * (size_t)A.n_pad;
    int rc;
    if ((rc = sw_alloc(s, (void **)&A.x[1], nb))) return rc;
    if ((rc = sw_alloc(s, (void **)&A.y[1], nb))) return rc;
    if ((rc = sw_alloc(s, (void **)&A.z[1], nb))) return rc;
    if ((rc = sw_alloc(s, (void **)&A.xr, nb))) return rc;
    if ((rc = sw_alloc(s, (void **)&A.yr, nb))) return rc;
    if ((rc = sw_alloc(s, (void **)&A.zr, nb))) return rc;
    if ((rc = sw_alloc(s, (void **)&A.gkeep, nb))) return rc;
    if ((rc = sw_alloc(s, (void **)&s->zeros, nb))) return rc;
    if ((rc = sw_alloc(s, (void **)&A.next, sizeof(int) * (size_t)A.n_pad, 0xFF))) return rc;
    if ((rc = sw_alloc(s, (void **)&A.cellof, sizeof(int) * (size_t)A.n_pad))) return rc;
    if ((rc = sw_alloc(s, (void **)&A.nbr_cnt, sizeof(int) * (size_t)A.n_pad))) return rc;
    if ((rc = sw_alloc(s, (void **)&A.nbr, sizeof(int) * (size_t)A.n_pad * (size_t)A.cap))) return rc;
    A.head = nullptr;
    s->ncell_alloc = 0;
  } else {
    // same scratch, different particles: the linked cells of the previous occupants are stale; they are reset
    // by the next launch, on that launch's stream (sw_prepare), so the order holds on non-blocking streams too
    s->cells_dirty = true;
  }
  // default actions: zeros
  A.f_swim = s->zeros; A.tqx = s->zeros; A.tqy = s->zeros; A.tqz = s->zeros;
  A.fex = nullptr; A.fey = nullptr; A.fez = nullptr;
  s->actions_bound = false;
  s->state_bound = true;
  s->list_valid = false;
  return SWARM_OK;
}

extern "C" int swarm_bind_actions(swarm_t *s, const swarm_actions_dev_t *act) {
  if (!s || !act) { sw_set_error("bad arguments"); return SWARM_ERR_INVALID; }
  if (!s->state_bound) { sw_set_error("bind the state first"); return SWARM_ERR_STATE; }
  if (!act->f_swim || !act->tqx || !act->tqy || !act->tqz) { sw_set_error("f_swim and torque arrays must be bound"); return SWARM_ERR_INVALID; }
  if ((act->fex || act->fey || act->fez) && !(act->fex && act->fey && act->fez)) { sw_set_error("bind all of fex, fey, fez or none"); return SWARM_ERR_INVALID; }
  s->act = *act;
  DevArrays &A = s->A;
  A.f_swim = act->f_swim; A.tqx = act->tqx; A.tqy = act->tqy; A.tqz = act->tqz;
  A.fex = act->fex; A.fey = act->fey; A.fez = act->fez;
  s->actions_bound = true;
  return SWARM_OK;
}

extern "C" int swarm_reserve(swarm_t *s, int64_t n_max) {
  if (!s || n_max < 0) { sw_set_error("bad arguments"); return SWARM_ERR_INVALID; }
  if (n_max >= (1LL << SW_TYPE_SHIFT)) { sw_set_error("n_max must be < 2^28 particles per handle"); return SWARM_ERR_INVALID; }
  s->n_reserved = n_max;
  return SWARM_OK;
}

extern "C" int swarm_bind_count(swarm_t *s, const int32_t *n_dev) {
  if (!s) return SWARM_ERR_INVALID;
  if (!s->state_bound) { sw_set_error("bind the state first"); return SWARM_ERR_STATE; }
  if (n_dev && !s->persistent) { sw_set_error("a device-resident count needs the persistent kernels"); return SWARM_ERR_STATE; }
  s->A.n_dev = (const int *)n_dev;
  s->list_valid = false;
  return SWARM_OK;
}

extern "C" int swarm_bind_ids(swarm_t *s, const int64_t *gid_dev) {
  if (!s) return SWARM_ERR_INVALID;
  if (!s->state_bound) { sw_set_error("bind the state first"); return SWARM_ERR_STATE; }
  s->A.gid = (const long long *)gid_dev;
  s->tile.sorted = false;   // the sorted copy of the ids is stale
  return SWARM_OK;
}

extern "C" int swarm_set_max_neighbors(swarm_t *s, int32_t cap) {
  if (!s || cap < 1) { sw_set_error("max_neighbors must be >= 1"); return SWARM_ERR_INVALID; }
  if (cap > SW_MAXCAP) { sw_set_error("max_neighbors exceeds the hard ceiling of 128"); return SWARM_ERR_OVERFLOW; }
  SW_CUDA(cudaSetDevice(s->cfg.device));
  s->A.cap = cap;
  s->cfg.max_neighbors = cap;
  if (s->state_bound) {
    int *nbr = nullptr;
    SW_CUDA(cudaMalloc((void **)&nbr, sizeof(int) * (size_t)s->A.n_pad * (size_t)cap));
    for (auto &p : s->owned) if (p == (void *)s->A.nbr) { cudaFree(p); p = nbr; }
    s->A.nbr = nbr;
  }
  s->list_valid = false;
  // clear the halted / overflow latch
  SW_CUDA(cudaMemset(&s->A.ctl->halted, 0, sizeof(int) * 2));
  return SWARM_OK;
}

static int sw_prepare(swarm *s, cudaStream_t st) {
  if (!s->state_bound) { sw_set_error("no state bound"); return SWARM_ERR_STATE; }
  if (!s->species_set || !s->wca_set) { sw_set_error("species / WCA tables not set"); return SWARM_ERR_STATE; }
  SW_CUDA(cudaSetDevice(s->cfg.device));
  int old_nc[3] = {s->P.nc[0], s->P.nc[1], s->P.nc[2]};
  sw_grid_setup(s);
  long long ncell = (long long)s->P.ncell_env * s->P.n_envs;
  bool changed = old_nc[0] != s->P.nc[0] || old_nc[1] != s->P.nc[1] || old_nc[2] != s->P.nc[2];
  if (!s->A.head || ncell > s->ncell_alloc || changed) {
    if (ncell > s->ncell_alloc || !s->A.head) {
      int *h = nullptr;
      SW_CUDA(cudaMalloc((void **)&h, sizeof(int) * (size_t)ncell));
      if (s->A.head) {
        for (auto &p : s->owned) if (p == (void *)s->A.head) { cudaFree(p); p = h; }
      } else {
        s->owned.push_back(h);
      }
      s->A.head = h;
      s->ncell_alloc = ncell;
    }
    SW_CUDA(cudaMemsetAsync(s->A.head, 0xFF, sizeof(int) * (size_t)s->ncell_alloc, st));
    SW_CUDA(cudaMemsetAsync(s->A.cellof, 0, sizeof(int) * (size_t)s->A.n_pad, st));
    s->list_valid = false;
  } else if (s->cells_dirty) {
    SW_CUDA(cudaMemsetAsync(s->A.head, 0xFF, sizeof(int) * (size_t)s->ncell_alloc, st));
    SW_CUDA(cudaMemsetAsync(s->A.cellof, 0, sizeof(int) * (size_t)s->A.n_pad, st));
  }
  if (s->cells_dirty) {
    SW_CUDA(cudaMemsetAsync(s->A.next, 0xFF, sizeof(int) * (size_t)s->A.n_pad, st));
    s->cells_dirty = false;
  }
  return SWARM_OK;
}

template <int DIM, bool NOISE>
static int sw_launch_persist(swarm *s, const RunArgs &R, cudaStream_t st) {
  auto kern = k_persist<DIM, NOISE>;
  int bpsm = 0;
  SW_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bpsm, kern, SW_THREADS, 0));
  if (bpsm < 1) { sw_set_error("persistent kernel does not fit on an SM"); return SWARM_ERR_CUDA; }
  s->coop_blocks_per_sm = bpsm;
  long long want = (s->A.n + SW_THREADS - 1) / SW_THREADS;
  long long maxb = (long long)s->num_sms * bpsm;
  int grid;
  if (want >= maxb) grid = (int)maxb;
  else if (want > s->num_sms) grid = (int)(((want + s->num_sms - 1) / s->num_sms) * s->num_sms);  // whole waves of SMs
  else grid = (int)(want > 0 ? want : 1);
  if (grid > maxb) grid = (int)maxb;
  const char *env = getenv("SWARM_B200_GRID");
  if (env) { int g = atoi(env); if (g >= 1 && g <= maxb) grid = g; }
  SW_CUDA(cudaMemsetAsync(&s->A.ctl->barrier, 0, sizeof(unsigned int), st));
  DevParams P = s->P;
  DevArrays A = s->A;
  RunArgs Rr = R;
  void *args[] = {(void *)&P, (void *)&A, (void *)&Rr};
  SW_CUDA(cudaLaunchCooperativeKernel((void *)kern, dim3(grid), dim3(SW_THREADS), args, 0, st));
  s->launches += 1;
  return SWARM_OK;
}

// resident 2-D kernel: one particle per thread, every CTA co-resident
template <bool NOISE, int PIPE>
static int sw_launch_resident(swarm *s, const RunArgs &R, cudaStream_t st, bool *fits) {
  auto kern = k_bd2d_resident<NOISE, PIPE>;
  // one CTA per SM (or fewer, larger ones): the grid barrier sees <= num_sms arrivals and the
  // particles are spread over the SMs to within one warp
  // one CTA per SM, the particles spread evenly over them; + 32: warp 0 of every CTA is the barrier warp
  long long ctas = s->num_sms;
  if (s->A.n < ctas * 128) ctas = (s->A.n + 127) / 128;   // small swarms: at least 4 particle warps per CTA
  int ppc = (int)((s->A.n + ctas - 1) / ctas);
  int threads = 32 + ((ppc + 31) / 32) * 32;
  const char *envt = getenv("SWARM_B200_RES_THREADS");
  if (envt) { int t = atoi(envt); if (t >= 64 && t <= SW_RES_MAXT && t % 32 == 0) { threads = t; ppc = t - 32; } }
  *fits = threads <= SW_RES_MAXT;
  if (!*fits) return SWARM_OK;
  int bpsm = 0;
  SW_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bpsm, kern, threads, 0));
  const long long want = (s->A.n + ppc - 1) / ppc;
  *fits = bpsm >= 1 && want <= (long long)s->num_sms * bpsm;
  if (!*fits) return SWARM_OK;
  s->coop_blocks_per_sm = bpsm;
  SW_CUDA(cudaMemsetAsync(&s->A.ctl->bar64, 0, sizeof(unsigned long long) * 32, st));
  DevParams P = s->P;
  DevArrays A = s->A;
  RunArgs Rr = R;
  Rr.particles_per_cta = ppc;
  void *args[] = {(void *)&P, (void *)&A, (void *)&Rr};
  SW_CUDA(cudaLaunchCooperativeKernel((void *)kern, dim3((unsigned)want), dim3(threads), args, 0, st));
  s->launches += 1;
  return SWARM_OK;
}


// ---------------------------------------------------------------------------------------------
// cluster kernel (swarm_cluster.cuh): small 2-D swarms on one thread-block cluster
// ---------------------------------------------------------------------------------------------
static bool sw_cluster_eligible(const swarm *s, const RunArgs &R) {
  return s->persistent && s->P.n_dims == 2 && R.mode == MODE_BD && s->P.n_envs == 1 && s->A.n >= 1 && s->A.n <= SW_CL_MAXN &&
         s->A.cap <= SW_CL_CAP && !s->A.gid && !s->A.okey && !s->A.n_dev && s->cluster_ok && s->tile.mode != 2 /* tiled kernel forced */ &&
         !getenv("SWARM_B200_NO_CLUSTER");
}

template <bool NOISE>
static int sw_launch_cluster(swarm *s, const RunArgs &R, cudaStream_t st, bool *done) {
  auto kern = k_bd2d_cluster<NOISE>;
  *done = false;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SW_CL_DSM_BYTES) != cudaSuccess) {
    cudaGetLastError();
    s->cluster_ok = false;      // a device without clusters / the shared memory: the resident kernel takes over for good
    return SWARM_OK;
  }
  DevParams P = s->P;
  DevArrays A = s->A;
  RunArgs Rr = R;
  Rr.particles_per_cta = (int)((s->A.n + SW_CL_CTAS - 1) / SW_CL_CTAS);
  kern<<<SW_CL_CTAS, SW_CL_THREADS, SW_CL_DSM_BYTES, st>>>(P, A, Rr);
  if (cudaGetLastError() != cudaSuccess) { s->cluster_ok = false; return SWARM_OK; }
  s->launches += 1;
  s->list_space = 0;
  *done = true;
  return SWARM_OK;
}

// ---------------------------------------------------------------------------------------------
// tiled 2-D kernel (swarm_tiled.cuh): eligibility, spatial sort, launch
// ---------------------------------------------------------------------------------------------
#define SW_TILE_MIN_N 4096   // below this a swarm is a handful of CTAs and the grid barrier is cheap

static bool sw_wca_symmetric(const swarm *s) {
  const DevParams &P = s->P;
  for (int a = 0; a < P.n_types; ++a)
    for (int b = 0; b < a; ++b)
      if (P.wca_cut[a][b] != P.wca_cut[b][a]) return false;
  return true;
}

// CTAs and slots per CTA of the tiled layout: one particle per thread, the particles spread evenly over the SMs
static void sw_tile_shape(const swarm *s, long long *ctas_out, int *ppc_out, int *threads_out) {
  long long ctas = s->num_sms;
  if (s->A.n < ctas * 128) ctas = (s->A.n + 127) / 128;   // small swarms: at least 4 warps per CTA
  if (ctas < 1) ctas = 1;
  const int ppc = (int)((s->A.n + ctas - 1) / ctas);
  *ppc_out = ppc;
  *threads_out = ((ppc + 31) / 32) * 32 + 32;   // + the halo warp
  *ctas_out = (s->A.n + ppc - 1) / ppc;
}

static bool sw_tile_eligible(const swarm *s, const RunArgs &R) {
  if (!s->persistent || s->P.n_dims != 2 || R.mode != MODE_BD || s->P.n_envs != 1 || s->A.n_dev) return false;
  if (s->tile.mode == 0 || getenv("SWARM_B200_NO_TILED")) return false;
  long long min_n = SW_TILE_MIN_N;
  if (const char *e = getenv("SWARM_B200_TILED_MIN_N")) min_n = atoll(e);
  if (s->tile.mode == 1 && s->A.n < min_n) return false;
  if (s->A.n < 1 || !sw_wca_symmetric(s)) return false;
  long long ctas; int ppc, threads;
  sw_tile_shape(s, &ctas, &ppc, &threads);
  return threads <= SW_RES_MAXT && ctas <= SW_TILE_MAXCTA && ctas <= s->num_sms;
}

// (re)allocate the tile scratch for the bound particle count and, when due, recompute the slot order
static int sw_tile_prepare(swarm *s, cudaStream_t st, bool *resorted) {
  TileHost &t = s->tile;
  DevArrays &A = s->A;
  const long long n = A.n;
  *resorted = false;
  if (t.n != A.n_pad) {
    // sized for the padded capacity, so a rebind within the same scratch keeps them
    const size_t np = (size_t)A.n_pad;
    int rc;
    if ((rc = sw_alloc(s, (void **)&t.perm, sizeof(int) * np))) return rc;
    if ((rc = sw_alloc(s, (void **)&t.key, sizeof(int) * np))) return rc;
    if ((rc = sw_alloc(s, (void **)&t.species_s, sizeof(int) * np))) return rc;
    if ((rc = sw_alloc(s, (void **)&t.gid_s, sizeof(long long) * np))) return rc;
    for (int b = 0; b < 2; ++b) {
      if ((rc = sw_alloc(s, (void **)&t.xs[b], sizeof(double) * np))) return rc;
      if ((rc = sw_alloc(s, (void **)&t.ys[b], sizeof(double) * np))) return rc;
    }
    if ((rc = sw_alloc(s, &t.rim[0], sizeof(RimRec) * 2 * np))) return rc;
    if ((rc = sw_alloc(s, &t.crec, sizeof(CellRec) * np))) return rc;
    t.rim[1] = (RimRec *)t.rim[0] + np;
    t.n = A.n_pad;
    t.sorted = false;
    t.nbins_alloc = 0;
  }
  const bool due = !t.sorted || (t.resort_interval > 0 && t.launches_since_sort >= t.resort_interval);
  if (!due) return SWARM_OK;
  // strips of width ~ sqrt(particles per CTA / density), rows of ~8 particles
  long long ctas; int ppc, threads;
  sw_tile_shape(s, &ctas, &ppc, &threads);
  const double Lx = s->P.box[0], Ly = s->P.box[1];
  const double density = (double)n / (Lx * Ly);
  SortGeom G;
  double W = sqrt((double)ppc / density);
  G.n_strips = (int)floor(Lx / W + 0.5);
  if (G.n_strips < 1) G.n_strips = 1;
  if (G.n_strips > 4096) G.n_strips = 4096;
  W = Lx / G.n_strips;
  double h = 8.0 / (density * W);
  double nr = ceil(Ly / h);
  if (nr < 1.0) nr = 1.0;
  if (nr > 65536.0) nr = 65536.0;
  while ((double)G.n_strips * nr > 4194304.0) nr = floor(nr / 2.0);
  G.n_rows = (int)nr;
  G.inv_w = 1.0 / W;
  G.inv_h = (double)G.n_rows / Ly;
  const int nbins = G.n_strips * G.n_rows;
  if (nbins + 1 > t.nbins_alloc) {
    int rc;
    if ((rc = sw_alloc(s, (void **)&t.hist, sizeof(int) * (size_t)(nbins + 1)))) return rc;
    if ((rc = sw_alloc(s, (void **)&t.start, sizeof(int) * (size_t)(nbins + 1)))) return rc;
    if ((rc = sw_alloc(s, (void **)&t.cursor, sizeof(int) * (size_t)(nbins + 1)))) return rc;
    t.nbins_alloc = nbins + 1;
  }
  SW_CUDA(cudaMemsetAsync(t.hist, 0, sizeof(int) * (size_t)(nbins + 1), st));
  const int B = sw_blocks(n, 256);
  k_sort_key<<<B, 256, 0, st>>>(s->P, G, n, A.x[0], A.y[0], t.key, t.hist);
  k_sort_scan<<<1, 1024, 0, st>>>(nbins, t.hist, t.start, t.cursor);
  k_sort_place<<<B, 256, 0, st>>>(n, t.key, t.cursor, t.perm);
  k_sort_fix<<<sw_blocks(nbins, 128), 128, 0, st>>>(nbins, t.start, t.perm);
  if (!getenv("SWARM_B200_TILE_NO_RIMFIRST")) {
    const double rc = s->P.max_cut + s->P.skin;
    const int rows_margin = (int)ceil(rc * G.inv_h) + 1;
    k_sort_rimfirst<<<(unsigned)ctas, 1024, 0, st>>>(s->P, G, n, ppc, rc, rows_margin, A.x[0], t.key, t.perm);
    s->launches += 1;
  }
  k_sort_gather<<<B, 256, 0, st>>>(n, t.perm, A.species, A.gid, t.species_s, t.gid_s);
  s->launches += 5;
  SW_CUDA(cudaGetLastError());
  t.sorted = true;
  t.launches_since_sort = 0;
  t.sorts += 1;
  *resorted = true;
  return SWARM_OK;
}

template <bool NOISE>
static int sw_launch_tiled(swarm *s, const RunArgs &R, cudaStream_t st, bool *fits) {
  auto kern = k_bd2d_tiled<NOISE>;
  long long ctas; int ppc, threads;
  sw_tile_shape(s, &ctas, &ppc, &threads);
  int bpsm = 0;
  const size_t dsm = sizeof(double) * SW_TILE_DSM_DOUBLES;
  SW_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dsm));
  SW_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bpsm, kern, threads, dsm));
  *fits = bpsm >= 1 && ctas <= (long long)s->num_sms * bpsm;
  if (!*fits) return SWARM_OK;
  bool resorted = false;
  int rc = sw_tile_prepare(s, st, &resorted);
  if (rc) return rc;
  TileHost &t = s->tile;
  RunArgs Rr = R;
  Rr.particles_per_cta = ppc;
  Rr.prof_cta = getenv("SWARM_B200_PROF_CTA") ? atoi(getenv("SWARM_B200_PROF_CTA")) : 0;
  Rr.prof_warp = getenv("SWARM_B200_PROF_WARP") ? atoi(getenv("SWARM_B200_PROF_WARP")) : 0;
  if (resorted || s->list_space != 1) Rr.force_rebuild = 1;   // the list in A.nbr is in another index space
  s->list_space = 1;
  s->coop_blocks_per_sm = bpsm;
  SW_CUDA(cudaMemsetAsync(&s->A.ctl->bar64, 0, sizeof(unsigned long long) * 32, st));
  DevParams P = s->P;
  DevArrays A = s->A;
  DevArrays T = s->A;   // slot space: same scratch (cells, lists, references), sorted positions / species / ids
  T.x[0] = t.xs[0]; T.x[1] = t.xs[1]; T.y[0] = t.ys[0]; T.y[1] = t.ys[1];
  T.species = t.species_s;
  T.gid = t.gid_s;
  T.okey = t.gid_s;
  const long long ncell = (long long)s->P.ncell_env;
  if (ncell > t.head64_cells) {
    int rc2;
    if ((rc2 = sw_alloc(s, (void **)&t.head64, sizeof(unsigned long long) * (size_t)ncell))) return rc2;   // zero: tag 0 is never used
    t.head64_cells = ncell;
  }
  TileArgs G;
  G.head64 = t.head64;
  G.crec = (CellRec *)t.crec;
  G.tag0 = t.tag_next;
  t.tag_next += 4ULL * (unsigned long long)R.n_steps + 16ULL;   // more than the rebuilds (and rolled-back ones) of this launch
  G.perm = t.perm; G.rim = (RimRec *)t.rim[0]; G.rim_stride = s->A.n_pad; G.ppc = ppc; G.k_init = 8; G.k_cap = SW_TILE_KMAX;
  G.k_safety = 0.9;
  if (const char *e = getenv("SWARM_B200_TILE_SAFETY")) { const double f = atof(e); if (f > 0.1 && f <= 0.95) G.k_safety = f; }   // 1.0 predicts the limit itself: nearly every period is rolled back
  if (const char *e = getenv("SWARM_B200_TILE_KCAP")) { const int k = atoi(e); if (k >= 1 && k <= SW_TILE_KMAX) G.k_cap = k; }
  G.flag0 = (unsigned int)((t.launches_total + 1) << 15);
  if (const char *e = getenv("SWARM_B200_TILE_KINIT")) { const int k = atoi(e); if (k >= 1 && k <= SW_TILE_KMAX) G.k_init = k; }
  void *args[] = {(void *)&P, (void *)&A, (void *)&T, (void *)&G, (void *)&Rr};
  SW_CUDA(cudaLaunchCooperativeKernel((void *)kern, dim3((unsigned)ctas), dim3(threads), args, dsm, st));
  s->launches += 1;
  t.launches_since_sort += 1;
  t.launches_total += 1;
  return SWARM_OK;
}

// batched environments of <= 32 agents: one warp per environment, no lists, no barrier, plain launch
template <bool NOISE>
static int sw_launch_env(swarm *s, const RunArgs &R, cudaStream_t st) {
  const long long warps_per_block = SW_ENV_THREADS / 32;
  const long long blocks = (s->P.n_envs + warps_per_block - 1) / warps_per_block;
  RunArgs Rr = R;
  Rr.force_rebuild = getenv("SWARM_B200_ENV_EXACT_PAIRS") ? 2 : 0;   // 2 = skip the fp32 pre-filter (cross-check)
  k_bd2d_env<NOISE><<<(unsigned)blocks, SW_ENV_THREADS, 0, st>>>(s->P, s->A, Rr);
  s->launches += 1;
  SW_CUDA(cudaGetLastError());
  return SWARM_OK;
}

// large 2-D swarms: lean grid-stride chunk kernel, every CTA co-resident (3 per SM)
template <bool NOISE>
static int sw_launch_stream(swarm *s, const RunArgs &R, cudaStream_t st) {
  auto kern = k_bd2d_stream<NOISE>;
  int bpsm = 0;
  SW_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bpsm, kern, SW_STREAM_THREADS, 0));
  if (bpsm < 1) { sw_set_error("stream kernel does not fit on an SM"); return SWARM_ERR_CUDA; }
  s->coop_blocks_per_sm = bpsm;
  long long want = (s->A.n + SW_STREAM_THREADS - 1) / SW_STREAM_THREADS;
  long long maxb = (long long)s->num_sms * bpsm;
  int grid = (int)(want >= maxb ? maxb : (want > 0 ? want : 1));
  const char *env = getenv("SWARM_B200_GRID");
  if (env) { int g = atoi(env); if (g >= 1 && g <= maxb) grid = g; }
  SW_CUDA(cudaMemsetAsync(&s->A.ctl->bar64, 0, sizeof(unsigned long long) * 32, st));
  DevParams P = s->P;
  DevArrays A = s->A;
  RunArgs Rr = R;
  void *args[] = {(void *)&P, (void *)&A, (void *)&Rr};
  SW_CUDA(cudaLaunchCooperativeKernel((void *)kern, dim3(grid), dim3(SW_STREAM_THREADS), args, 0, st));
  s->launches += 1;
  return SWARM_OK;
}

template <int DIM, bool NOISE>
static int sw_launch_np(swarm *s, const RunArgs &R, cudaStream_t st) {
  const int T = SW_THREADS, B = sw_blocks(s->A.n > 0 ? s->A.n : 1, T);
  k_np_check<DIM><<<B, T, 0, st>>>(s->P, s->A, R.force_rebuild, R.steps_base);
  s->launches += 1;
  for (long long k = 0; k < R.n_steps; ++k) {
    k_np_clear<DIM><<<B, T, 0, st>>>(s->P, s->A, k);
    k_np_insert<<<B, T, 0, st>>>(s->A, k);
    k_np_search<DIM><<<B, T, 0, st>>>(s->P, s->A, k);
    k_np_halt<<<1, 32, 0, st>>>(s->A);
    k_np_step<DIM, NOISE><<<B, T, 0, st>>>(s->P, s->A, R, k);
    s->launches += 5;
  }
  k_np_finish<DIM><<<B, T, 0, st>>>(s->A, R.steps_base);
  s->launches += 1;
  SW_CUDA(cudaGetLastError());
  return SWARM_OK;
}

// velocity Verlet divides by mass and rotational inertia (espresso.py:391-400 makes both mandatory)
static int sw_check_inertial(const swarm *s) {
  for (int k = 0; k < s->P.n_species; ++k) {
    const double *ri = s->P.sp_rin[k];
    bool ok = s->P.sp_mass[k] > 0.0 && ri[0] > 0.0;
    if (s->P.n_dims == 3) ok = ok && ri[0] == ri[1] && ri[1] == ri[2];
    if (!ok) {
      sw_set_error("Langevin integrator: every species needs a positive mass and rotational inertia (isotropic in 3-D)");
      return SWARM_ERR_INVALID;
    }
  }
  return SWARM_OK;
}

static int sw_run_common(swarm *s, RunArgs R, cudaStream_t st) {
  int rc = sw_prepare(s, st);
  if (rc) return rc;
  if (R.n_steps < 0) { sw_set_error("n_steps must be >= 0"); return SWARM_ERR_INVALID; }
  if (R.mode == MODE_LV) {  // the species table may have been replaced since swarm_set_integrator
    rc = sw_check_inertial(s);
    if (rc) return rc;
  }
  s->steps_requested = R.n_steps;
  s->last_mode = R.mode;
  s->lv_was_resume = (R.mode == MODE_LV) && s->lv_resume;
  R.force_rebuild = s->list_valid ? 0 : 1;
  if (R.n_steps == 0 || s->A.n == 0) {
    SW_CUDA(cudaMemsetAsync(&s->A.ctl->steps_done, 0, sizeof(long long), st));
    return SWARM_OK;
  }
  const bool noise = (R.mode == MODE_BD || R.mode == MODE_LV) && (s->P.kT > 0.0);
  // velocity Verlet: n steps are n + 1 force evaluations (see RunArgs); the kernels count evaluations
  const bool lv = (R.mode == MODE_LV);
  // swarm_run_part: a chunk cut in parts (the domain exchange sits between them).  CONTINUE = the previous part left its last
  // step open: start by closing it, no recalculation.  LEAVE_OPEN = stop between the half kicks of the last step (n evaluations).
  const bool part_cont = lv && (s->run_flags & SWARM_RUN_CONTINUE), part_open = lv && (s->run_flags & SWARM_RUN_LEAVE_OPEN);
  s->run_flags = 0;
  if (lv && !part_open) R.n_steps += 1;
  const bool lv_resume = lv && (s->lv_resume || part_cont);
  s->lv_resume = false;
  // bound the length of one cooperative launch (32-bit barrier counter, watchdog friendliness)
  const long long max_chunk = s->persistent ? 8192 : 1024;
  long long done = 0;
  while (done < R.n_steps) {
    RunArgs Rc = R;
    Rc.n_steps = (R.n_steps - done < max_chunk) ? R.n_steps - done : max_chunk;
    Rc.step0 = R.step0 + (u64)done;
    Rc.steps_base = done;   // the kernels report steps_base + their own count: steps of the whole swarm_run
    if (done > 0) Rc.force_rebuild = 0;
    Rc.lv_first = (lv && done == 0 && !lv_resume) ? 1 : 0;
    Rc.lv_close = (lv && !part_open && done + Rc.n_steps == R.n_steps) ? 1 : 0;
    // only the last sub-launch of a chunk materialises v / omega on its final step: handled by
    // Rc.n_steps bookkeeping -- earlier sub-launches also write them, harmlessly.
    bool done_resident = false;
    if (s->persistent && s->P.n_dims == 2 && R.mode == MODE_BD && s->P.n_envs > 1 && s->P.env_size <= 32 && !s->A.n_dev &&
        !getenv("SWARM_B200_NO_ENV_KERNEL")) {
      rc = noise ? sw_launch_env<true>(s, Rc, st) : sw_launch_env<false>(s, Rc, st);
      if (rc) return rc;
      done += Rc.n_steps;
      continue;
    }
    if (sw_cluster_eligible(s, Rc)) {
      rc = noise ? sw_launch_cluster<true>(s, Rc, st, &done_resident) : sw_launch_cluster<false>(s, Rc, st, &done_resident);
      if (rc) return rc;
    }
    if (!done_resident && sw_tile_eligible(s, Rc)) {
      rc = noise ? sw_launch_tiled<true>(s, Rc, st, &done_resident) : sw_launch_tiled<false>(s, Rc, st, &done_resident);
      if (rc) return rc;
    }
    if (!done_resident && s->list_space != 0) {   // the list in A.nbr is in the tiled kernel's slot space
      Rc.force_rebuild = 1;
      s->list_space = 0;
    }
    if (done_resident) {
    } else if (s->persistent && s->P.n_dims == 2 && R.mode == MODE_BD && !getenv("SWARM_B200_NO_RESIDENT")) {
      static const bool pipe = !(getenv("SWARM_B200_RES_PIPE") && atoi(getenv("SWARM_B200_RES_PIPE")) == 0);
      rc = !noise ? sw_launch_resident<false, 0>(s, Rc, st, &done_resident)
                  : (pipe ? sw_launch_resident<true, 1>(s, Rc, st, &done_resident) : sw_launch_resident<true, 0>(s, Rc, st, &done_resident));
      if (rc) return rc;
    }
    if (done_resident) {
    } else if (s->persistent && s->P.n_dims == 2 && R.mode == MODE_BD && !getenv("SWARM_B200_NO_STREAM")) {
      rc = noise ? sw_launch_stream<true>(s, Rc, st) : sw_launch_stream<false>(s, Rc, st);
    } else if (s->persistent) {
      if (s->P.n_dims == 2) rc = noise ? sw_launch_persist<2, true>(s, Rc, st) : sw_launch_persist<2, false>(s, Rc, st);
      else rc = noise ? sw_launch_persist<3, true>(s, Rc, st) : sw_launch_persist<3, false>(s, Rc, st);
    } else {
      if (s->P.n_dims == 2) rc = noise ? sw_launch_np<2, true>(s, Rc, st) : sw_launch_np<2, false>(s, Rc, st);
      else rc = noise ? sw_launch_np<3, true>(s, Rc, st) : sw_launch_np<3, false>(s, Rc, st);
    }
    if (rc) return rc;
    done += Rc.n_steps;
  }
  s->list_valid = true;
  return SWARM_OK;
}

extern "C" int swarm_run(swarm_t *s, int64_t n_steps, uint64_t step0, void *stream) {
  if (!s) return SWARM_ERR_INVALID;
  RunArgs R{};
  R.n_steps = n_steps; R.step0 = step0; R.mode = (s->integrator == SWARM_INTEGRATOR_LANGEVIN) ? MODE_LV : MODE_BD;
  return sw_run_common(s, R, (cudaStream_t)stream);
}

extern "C" int swarm_run_part(swarm_t *s, int64_t n_steps, uint64_t step0, int32_t flags, void *stream) {
  if (!s) return SWARM_ERR_INVALID;
  if (flags & ~(SWARM_RUN_CONTINUE | SWARM_RUN_LEAVE_OPEN)) { sw_set_error("unknown swarm_run_part flags"); return SWARM_ERR_INVALID; }
  s->run_flags = flags;
  const int rc = swarm_run(s, n_steps, step0, stream);
  s->run_flags = 0;
  return rc;
}

extern "C" int swarm_set_integrator(swarm_t *s, int32_t kind) {
  if (!s) return SWARM_ERR_INVALID;
  if (kind != SWARM_INTEGRATOR_BROWNIAN && kind != SWARM_INTEGRATOR_LANGEVIN) {
    sw_set_error("integrator must be SWARM_INTEGRATOR_BROWNIAN or SWARM_INTEGRATOR_LANGEVIN");
    return SWARM_ERR_INVALID;
  }
  if (kind == SWARM_INTEGRATOR_LANGEVIN) {
    if (!s->species_set) { sw_set_error("set the species table before selecting the Langevin integrator"); return SWARM_ERR_STATE; }
    int rc = sw_check_inertial(s);
    if (rc) return rc;
  }
  s->integrator = kind;
  s->lv_resume = false;
  return SWARM_OK;
}

extern "C" int swarm_relax(swarm_t *s, int64_t n_iter, double gamma, double max_disp, void *stream) {
  if (!s) return SWARM_ERR_INVALID;
  RunArgs R{};
  R.n_steps = n_iter; R.step0 = 0; R.mode = MODE_SD; R.sd_gamma = gamma; R.sd_maxd = max_disp;
  return sw_run_common(s, R, (cudaStream_t)stream);
}

extern "C" int swarm_sync_stats(swarm_t *s, void *stream, swarm_stats_t *out) {
  if (!s || !out) return SWARM_ERR_INVALID;
  SW_CUDA(cudaSetDevice(s->cfg.device));
  cudaStream_t st = (cudaStream_t)stream;
  SW_CUDA(cudaMemcpyAsync(s->ctl_host, s->A.ctl, sizeof(Ctl), cudaMemcpyDeviceToHost, st));
  SW_CUDA(cudaStreamSynchronize(st));
  out->steps_done = s->ctl_host->steps_done;
  out->steps_requested = s->steps_requested;
  if (s->last_mode == MODE_LV) {
    // the kernels counted force evaluations; an interrupted chunk (Verlet overflow) stopped between the half kicks
    // of a step, and the resumed run closes that step instead of starting with a recalculation
    if (s->ctl_host->overflow && (out->steps_done > 0 || s->lv_was_resume)) s->lv_resume = true;
    if (out->steps_done > out->steps_requested) out->steps_done = out->steps_requested;
  }
  out->rebuilds = s->ctl_host->rebuilds;
  out->launches = s->launches;
  out->overflow = s->ctl_host->overflow;
  out->max_neighbors_seen = s->ctl_host->max_nbr_seen;
  out->capacity = s->A.cap;
  out->persistent = s->persistent ? 1 : 0;
  if (s->ctl_host->overflow) s->list_valid = false;
  return SWARM_OK;
}

extern "C" int swarm_debug_timing(swarm_t *s, int64_t out8[8]) {
  if (!s || !out8) return SWARM_ERR_INVALID;
  SW_CUDA(cudaSetDevice(s->cfg.device));
  SW_CUDA(cudaDeviceSynchronize());
  long long t[8];
  SW_CUDA(cudaMemcpy(t, s->A.ctl->tprof, sizeof(t), cudaMemcpyDeviceToHost));
  SW_CUDA(cudaMemset(s->A.ctl->tprof, 0, sizeof(t)));
  for (int k = 0; k < 8; ++k) out8[k] = t[k];
  return SWARM_OK;
}

extern "C" int swarm_apply_new_direction(swarm_t *s, void *stream) {
  if (!s || !s->state_bound) { sw_set_error("no state bound"); return SWARM_ERR_STATE; }
  if (!s->act.has_nd || !s->act.ndx || !s->act.ndy || !s->act.ndz) { sw_set_error("new_direction arrays not bound"); return SWARM_ERR_STATE; }
  SW_CUDA(cudaSetDevice(s->cfg.device));
  const DevArrays &A = s->A;
  if (A.n == 0) return SWARM_OK;
  k_apply_new_direction<<<sw_blocks(A.n, 256), 256, 0, (cudaStream_t)stream>>>(s->P.n_dims, A.n, s->act.has_nd, s->act.ndx, s->act.ndy,
                                                                              s->act.ndz, A.theta, A.q0, A.q1, A.q2, A.q3);
  s->launches += 1;
  SW_CUDA(cudaGetLastError());
  return SWARM_OK;
}

extern "C" int swarm_directors(swarm_t *s, double *ex, double *ey, double *ez, void *stream) {
  if (!s || !s->state_bound) { sw_set_error("no state bound"); return SWARM_ERR_STATE; }
  SW_CUDA(cudaSetDevice(s->cfg.device));
  const DevArrays &A = s->A;
  if (A.n == 0) return SWARM_OK;
  k_directors<<<sw_blocks(A.n, 256), 256, 0, (cudaStream_t)stream>>>(s->P.n_dims, A.n, A.theta, A.q0, A.q1, A.q2, A.q3, ex, ey, ez);
  s->launches += 1;
  SW_CUDA(cudaGetLastError());
  return SWARM_OK;
}

// rebuild the Verlet list at the current positions (0-step run with a forced rebuild)
static int sw_force_rebuild(swarm *s, cudaStream_t st) {
  int rc = sw_prepare(s, st);
  if (rc) return rc;
  if (s->A.n == 0) return SWARM_OK;
  const int T = SW_THREADS, B = sw_blocks(s->A.n, T);
  // phases as separate launches; flag slot 0 forced
  if (s->P.n_dims == 2) {
    k_np_check<2><<<B, T, 0, st>>>(s->P, s->A, 1, 0);
    k_np_clear<2><<<B, T, 0, st>>>(s->P, s->A, 0);
    k_np_insert<<<B, T, 0, st>>>(s->A, 0);
    k_np_search<2><<<B, T, 0, st>>>(s->P, s->A, 0);
    k_np_finish<2><<<B, T, 0, st>>>(s->A, 0);
  } else {
    k_np_check<3><<<B, T, 0, st>>>(s->P, s->A, 1, 0);
    k_np_clear<3><<<B, T, 0, st>>>(s->P, s->A, 0);
    k_np_insert<<<B, T, 0, st>>>(s->A, 0);
    k_np_search<3><<<B, T, 0, st>>>(s->P, s->A, 0);
    k_np_finish<3><<<B, T, 0, st>>>(s->A, 0);
  }
  s->launches += 5;
  SW_CUDA(cudaGetLastError());
  s->list_valid = true;
  s->list_space = 0;
  return SWARM_OK;
}

extern "C" int swarm_neighbors_dump(swarm_t *s, int32_t *pairs, int64_t cap, int64_t *count) {
  if (!s || !count) return SWARM_ERR_INVALID;
  cudaStream_t st = 0;
  int rc = sw_force_rebuild(s, st);
  if (rc) return rc;
  swarm_stats_t stats;
  if ((rc = swarm_sync_stats(s, st, &stats))) return rc;
  if (stats.overflow) { sw_set_error("Verlet list capacity exceeded; raise max_neighbors"); return SWARM_ERR_OVERFLOW; }
  long long n = s->A.n;
  std::vector<int> cnt((size_t)n), nbr((size_t)s->A.n_pad * (size_t)s->A.cap);
  if (n > 0) {
    SW_CUDA(cudaMemcpy(cnt.data(), s->A.nbr_cnt, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost));
    SW_CUDA(cudaMemcpy(nbr.data(), s->A.nbr, sizeof(int) * nbr.size(), cudaMemcpyDeviceToHost));
  }
  int64_t c = 0;
  for (long long i = 0; i < n; ++i)
    for (int k = 0; k < cnt[(size_t)i]; ++k) {
      if (c < cap && pairs) {
        pairs[2 * c] = (int32_t)i;
        pairs[2 * c + 1] = nbr[(size_t)k * (size_t)s->A.n_pad + (size_t)i] & SW_IDX_MASK;
      }
      ++c;
    }
  *count = c;
  return SWARM_OK;
}

extern "C" int swarm_set_tiling(swarm_t *s, int32_t mode, int32_t resort_interval) {
  if (!s || mode < 0 || mode > 2) { sw_set_error("tiling mode must be 0 (off), 1 (automatic) or 2 (always)"); return SWARM_ERR_INVALID; }
  s->tile.mode = mode;
  if (resort_interval > 0) s->tile.resort_interval = resort_interval;
  return SWARM_OK;
}

extern "C" int swarm_tiling_stats(swarm_t *s, int64_t out8[8]) {
  if (!s || !out8) return SWARM_ERR_INVALID;
  SW_CUDA(cudaSetDevice(s->cfg.device));
  SW_CUDA(cudaDeviceSynchronize());
  SW_CUDA(cudaMemcpy(s->ctl_host, s->A.ctl, sizeof(Ctl), cudaMemcpyDeviceToHost));
  long long ctas = 0; int ppc = 0, threads = 0;
  if (s->state_bound && s->A.n > 0) sw_tile_shape(s, &ctas, &ppc, &threads);
  out8[0] = s->list_space == 1; out8[1] = s->ctl_host->tile_K; out8[2] = s->ctl_host->tile_since;
  out8[3] = s->ctl_host->tile_rollbacks; out8[4] = s->ctl_host->tile_partners; out8[5] = s->tile.sorts;
  out8[6] = ctas; out8[7] = ppc;
  return SWARM_OK;
}

extern "C" int swarm_tiled_neighbors_dump(swarm_t *s, int32_t *pairs, int64_t cap, int64_t *count) {
  if (!s || !count) return SWARM_ERR_INVALID;
  if (!s->state_bound || s->list_space != 1 || !s->tile.sorted) { sw_set_error("the last run did not use the tiled kernel"); return SWARM_ERR_STATE; }
  SW_CUDA(cudaSetDevice(s->cfg.device));
  SW_CUDA(cudaDeviceSynchronize());
  const long long n = s->A.n;
  std::vector<int> cnt((size_t)n), perm((size_t)n), nbr((size_t)s->A.n_pad * (size_t)s->A.cap);
  SW_CUDA(cudaMemcpy(cnt.data(), s->A.nbr_cnt, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost));
  SW_CUDA(cudaMemcpy(perm.data(), s->tile.perm, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost));
  SW_CUDA(cudaMemcpy(nbr.data(), s->A.nbr, sizeof(int) * nbr.size(), cudaMemcpyDeviceToHost));
  std::vector<std::pair<int, int>> out;
  for (long long sl = 0; sl < n; ++sl)
    for (int k = 0; k < cnt[(size_t)sl]; ++k)
      out.emplace_back(perm[(size_t)sl], perm[(size_t)(nbr[(size_t)k * (size_t)s->A.n_pad + (size_t)sl] & SW_TILE_JMASK)]);
  std::sort(out.begin(), out.end());
  int64_t c = 0;
  for (const auto &pr : out) {
    if (c < cap && pairs) { pairs[2 * c] = pr.first; pairs[2 * c + 1] = pr.second; }
    ++c;
  }
  *count = c;
  return SWARM_OK;
}

extern "C" int swarm_cell_index(swarm_t *s, const int32_t nc[3], int32_t *cell_host) {
  if (!s || !s->state_bound || !nc || !cell_host) { sw_set_error("bad arguments"); return SWARM_ERR_INVALID; }
  SW_CUDA(cudaSetDevice(s->cfg.device));
  long long n = s->A.n;
  if (n == 0) return SWARM_OK;
  int *d = nullptr;
  SW_CUDA(cudaMalloc((void **)&d, sizeof(int) * (size_t)n));
  if (s->P.n_dims == 2) k_cell_index<2><<<sw_blocks(n, 256), 256>>>(s->P, s->A, nc[0], nc[1], nc[2], d);
  else k_cell_index<3><<<sw_blocks(n, 256), 256>>>(s->P, s->A, nc[0], nc[1], nc[2], d);
  s->launches += 1;
  cudaError_t e = cudaMemcpy(cell_host, d, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost);
  cudaFree(d);
  if (e != cudaSuccess) return sw_cuda_fail(e, "cudaMemcpy(cell index)");
  return SWARM_OK;
}

extern "C" int swarm_forces(swarm_t *s, double *fx, double *fy, double *fz, void *stream) {
  if (!s) return SWARM_ERR_INVALID;
  cudaStream_t st = (cudaStream_t)stream;
  int rc = sw_force_rebuild(s, st);
  if (rc) return rc;
  if (s->A.n == 0) return SWARM_OK;
  if (s->P.n_dims == 2) k_forces<2><<<sw_blocks(s->A.n, 256), 256, 0, st>>>(s->P, s->A, fx, fy, fz);
  else k_forces<3><<<sw_blocks(s->A.n, 256), 256, 0, st>>>(s->P, s->A, fx, fy, fz);
  s->launches += 1;
  SW_CUDA(cudaGetLastError());
  return SWARM_OK;
}

extern "C" int swarm_philox_host(const uint64_t ctr[4], const uint64_t key[2], uint64_t out[4]) {
  u64 o[4];
  sw_philox4x64_10(ctr[0], ctr[1], ctr[2], ctr[3], key[0], key[1], o);
  for (int k = 0; k < 4; ++k) out[k] = o[k];
  return SWARM_OK;
}

// fp64 peak as these kernels can see it: `ilp` independent DFMA chains per thread, `iters` x 64 DFMAs per chain, every
// SM full (the denominator of the issue / fp64-pipe fractions in DESIGN.md)
template <int ILP>
__global__ void __launch_bounds__(256) k_microbench_dfma(int iters, double a, double b, double *sink) {
  double v[ILP];
#pragma unroll
  for (int k = 0; k < ILP; ++k) v[k] = (double)(threadIdx.x + k);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 64; ++r)
#pragma unroll
      for (int k = 0; k < ILP; ++k) v[k] = fma(v[k], a, b);
  }
  double s = 0.0;
#pragma unroll
  for (int k = 0; k < ILP; ++k) s += v[k];
  if (s == 123.456) sink[0] = s;   // never true: keeps the chains alive
}

extern "C" int swarm_microbench_dfma(int32_t device, int32_t ilp, int32_t iters, double out4[4]) {
  if (!out4 || iters < 1) return SWARM_ERR_INVALID;
  SW_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  SW_CUDA(cudaGetDeviceProperties(&prop, device));
  double *sink = nullptr;
  SW_CUDA(cudaMalloc((void **)&sink, 8));
  const int blocks = prop.multiProcessorCount * 8, threads = 256;
  cudaEvent_t e0, e1;
  SW_CUDA(cudaEventCreate(&e0)); SW_CUDA(cudaEventCreate(&e1));
  const int I = (ilp >= 8) ? 8 : (ilp >= 4 ? 4 : 1);
  for (int rep = 0; rep < 2; ++rep) {   // the first repetition warms up
    SW_CUDA(cudaEventRecord(e0));
    if (I == 8) k_microbench_dfma<8><<<blocks, threads>>>(iters, 0.999999, 1e-9, sink);
    else if (I == 4) k_microbench_dfma<4><<<blocks, threads>>>(iters, 0.999999, 1e-9, sink);
    else k_microbench_dfma<1><<<blocks, threads>>>(iters, 0.999999, 1e-9, sink);
    SW_CUDA(cudaEventRecord(e1));
    SW_CUDA(cudaEventSynchronize(e1));
  }
  float ms = 0.0f;
  SW_CUDA(cudaEventElapsedTime(&ms, e0, e1));
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(sink);
  const double n = (double)blocks * threads * (double)iters * 64.0 * I;
  out4[0] = n / (ms * 1e-3); out4[1] = ms; out4[2] = prop.multiProcessorCount; out4[3] = I;
  return SWARM_OK;
}

extern "C" int swarm_noise_words_host(uint64_t seed, uint64_t id, uint64_t step, uint64_t stream_id, uint64_t out[4]) {
  if (!out) return SWARM_ERR_INVALID;
  u64 o[4];
  sw_noise_words(seed, id, step, stream_id, o);
  for (int k = 0; k < 4; ++k) out[k] = o[k];
  return SWARM_OK;
}

extern "C" int swarm_gaussians(swarm_t *s, int64_t n, uint64_t step, uint64_t stream_id, double *out, void *stream) {
  if (!s || n < 0 || !out) return SWARM_ERR_INVALID;
  SW_CUDA(cudaSetDevice(s->cfg.device));
  if (n == 0) return SWARM_OK;
  k_gaussians<<<sw_blocks(n, 256), 256, 0, (cudaStream_t)stream>>>(s->P.seed, n, step, stream_id, out);
  s->launches += 1;
  SW_CUDA(cudaGetLastError());
  return SWARM_OK;
}

extern "C" int swarm_set_actions_host(swarm_t *s, const double *f_swim_host, const double *torque3n_host, void *stream) {
  if (!s || !s->state_bound || !s->actions_bound) { sw_set_error("state and actions must be bound"); return SWARM_ERR_STATE; }
  SW_CUDA(cudaSetDevice(s->cfg.device));
  cudaStream_t st = (cudaStream_t)stream;
  long long n = s->A.n;
  if (n == 0) return SWARM_OK;
  if (f_swim_host) SW_CUDA(cudaMemcpyAsync(s->A.f_swim, f_swim_host, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, st));
  if (torque3n_host) {
    if (s->obs_scratch_bytes < sizeof(double) * 3 * (size_t)n) {
      if (s->obs_scratch) cudaFree(s->obs_scratch);
      s->obs_scratch = nullptr; s->obs_scratch_bytes = 0;
      SW_CUDA(cudaMalloc(&s->obs_scratch, sizeof(double) * 3 * (size_t)n));
      s->obs_scratch_bytes = sizeof(double) * 3 * (size_t)n;
    }
    SW_CUDA(cudaMemcpyAsync(s->obs_scratch, torque3n_host, sizeof(double) * 3 * (size_t)n, cudaMemcpyHostToDevice, st));
    k_deinterleave3<<<sw_blocks(n, 256), 256, 0, st>>>(n, (const double *)s->obs_scratch, s->A.tqx, s->A.tqy, s->A.tqz);
    s->launches += 1;
    SW_CUDA(cudaGetLastError());
  }
  return SWARM_OK;
}

// 2-D swarms: the torque of an action has one component that matters (espresso.py:1232 writes the 3-vector because ESPResSo
// is 3-D; in the plane only tau_z turns the director, and the 2-D kernels read nothing else).  Half the host -> device bytes
// of swarm_set_actions_host and no de-interleave kernel.
extern "C" int swarm_set_actions_host_2d(swarm_t *s, const double *f_swim_host, const double *torque_z_host, void *stream) {
  if (!s || !s->state_bound || !s->actions_bound) { sw_set_error("state and actions must be bound"); return SWARM_ERR_STATE; }
  if (s->P.n_dims != 2) { sw_set_error("swarm_set_actions_host_2d needs a 2-D swarm"); return SWARM_ERR_INVALID; }
  SW_CUDA(cudaSetDevice(s->cfg.device));
  cudaStream_t st = (cudaStream_t)stream;
  const long long n = s->A.n;
  if (n == 0) return SWARM_OK;
  if (f_swim_host) SW_CUDA(cudaMemcpyAsync(s->A.f_swim, f_swim_host, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, st));
  if (torque_z_host) SW_CUDA(cudaMemcpyAsync(s->A.tqz, torque_z_host, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, st));
  return SWARM_OK;
}

// One frame of the trajectory (espresso.py:1110-1130: Unwrapped_Positions, Velocities, Directors, each (N, 3)) written
// into DEVICE memory by one kernel on the caller's stream: the snapshot ring of the asynchronous trajectory writer
// (swarmrl_b200/trajectory.py) is filled without touching the host; the D2H copy runs on the writer's own stream.
__global__ void k_snapshot(int n_dims, long long n, const double *x, const double *y, const double *z, const double *vx, const double *vy,
                           const double *vz, const double *theta, const double *q0, const double *q1, const double *q2,
                           const double *q3, double *pos, double *vel, double *dir) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  pos[3 * i] = x[i]; pos[3 * i + 1] = y[i]; pos[3 * i + 2] = z[i];
  vel[3 * i] = vx[i]; vel[3 * i + 1] = vy[i]; vel[3 * i + 2] = vz[i];
  double ex, ey, ez;
  if (n_dims == 2) {
    double s, c;
    sincos(theta[i], &s, &c);
    ex = c; ey = s; ez = 0.0;
  } else {
    const double a = q0[i], b = q1[i], c = q2[i], d = q3[i];
    ex = 2 * (b * d + a * c); ey = 2 * (c * d - a * b); ez = a * a - b * b - c * c + d * d;
  }
  dir[3 * i] = ex; dir[3 * i + 1] = ey; dir[3 * i + 2] = ez;
}

extern "C" int swarm_snapshot_dev(swarm_t *s, double *pos_n3_dev, double *vel_n3_dev, double *dir_n3_dev, void *stream) {
  if (!s || !s->state_bound || !pos_n3_dev || !vel_n3_dev || !dir_n3_dev) { sw_set_error("bad arguments"); return SWARM_ERR_INVALID; }
  SW_CUDA(cudaSetDevice(s->cfg.device));
  const DevArrays &A = s->A;
  if (A.n == 0) return SWARM_OK;
  k_snapshot<<<sw_blocks(A.n, 256), 256, 0, (cudaStream_t)stream>>>(s->P.n_dims, A.n, A.x[0], A.y[0], A.z[0], A.vx, A.vy, A.vz, A.theta,
                                                                    A.q0, A.q1, A.q2, A.q3, pos_n3_dev, vel_n3_dev, dir_n3_dev);
  s->launches += 1;
  SW_CUDA(cudaGetLastError());
  return SWARM_OK;
}

extern "C" int swarm_get_particle_data_host(swarm_t *s, double *pos_n3, double *vel_n3, double *dir_n3, void *stream) {
  if (!s || !s->state_bound) { sw_set_error("no state bound"); return SWARM_ERR_STATE; }
  SW_CUDA(cudaSetDevice(s->cfg.device));
  cudaStream_t st = (cudaStream_t)stream;
  long long n = s->A.n;
  if (n == 0) return SWARM_OK;
  size_t need = sizeof(double) * 6 * (size_t)n;
  if (s->obs_scratch_bytes < need) {
    if (s->obs_scratch) cudaFree(s->obs_scratch);
    s->obs_scratch = nullptr; s->obs_scratch_bytes = 0;
    SW_CUDA(cudaMalloc(&s->obs_scratch, need));
    s->obs_scratch_bytes = need;
  }
  double *buf = (double *)s->obs_scratch, *tmp = buf + 3 * n;
  const int B = sw_blocks(n, 256);
  if (pos_n3) {
    k_interleave3<<<B, 256, 0, st>>>(n, s->A.x[0], s->A.y[0], s->A.z[0], buf);
    SW_CUDA(cudaMemcpyAsync(pos_n3, buf, sizeof(double) * 3 * (size_t)n, cudaMemcpyDeviceToHost, st));
    SW_CUDA(cudaStreamSynchronize(st));
    s->launches += 1;
  }
  if (vel_n3) {
    k_interleave3<<<B, 256, 0, st>>>(n, s->A.vx, s->A.vy, s->A.vz, buf);
    SW_CUDA(cudaMemcpyAsync(vel_n3, buf, sizeof(double) * 3 * (size_t)n, cudaMemcpyDeviceToHost, st));
    SW_CUDA(cudaStreamSynchronize(st));
    s->launches += 1;
  }
  if (dir_n3) {
    k_directors<<<B, 256, 0, st>>>(s->P.n_dims, n, s->A.theta, s->A.q0, s->A.q1, s->A.q2, s->A.q3, tmp, tmp + n, tmp + 2 * n);
    k_interleave3<<<B, 256, 0, st>>>(n, tmp, tmp + n, tmp + 2 * n, buf);
    SW_CUDA(cudaMemcpyAsync(dir_n3, buf, sizeof(double) * 3 * (size_t)n, cudaMemcpyDeviceToHost, st));
    SW_CUDA(cudaStreamSynchronize(st));
    s->launches += 2;
  }
  SW_CUDA(cudaGetLastError());
  return SWARM_OK;
}
