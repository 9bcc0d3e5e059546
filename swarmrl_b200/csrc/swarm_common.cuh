// swarm_common.cuh -- shared device structs, Philox4x64-10, geometry helpers (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <type_traits>
#include <vector>

#include "../../include/swarm_b200.h"

#define SW_S SWARM_MAX_SPECIES
#define SW_T SWARM_MAX_TYPES

typedef unsigned long long u64;

// ---------------------------------------------------------------------------------------------
// parameters passed by value to every kernel (< 4 KB)
// ---------------------------------------------------------------------------------------------
struct DevParams {
  int n_dims;
  int periodic[3];
  double box[3], inv_box[3];
  double dt, kT;
  u64 seed;
  unsigned int rk[20];  // Philox4x32-10 round keys of the seed (sw_key_schedule): the kernels read them as constant-bank operands
  int n_species, n_types;
  int sp_type[SW_S];
  double sp_gt[SW_S][3], sp_gr[SW_S][3], sp_mass[SW_S], sp_rin[SW_S][3];
  double wca_sigma[SW_T][SW_T], wca_eps[SW_T][SW_T], wca_cut[SW_T][SW_T];
  int walls;
  double wall_sigma[SW_T], wall_eps[SW_T], wall_cut[SW_T];
  double max_cut, skin;
  int nc[3];      // Verlet-build cell grid per environment
  int ncell_env;  // nc[0]*nc[1]*nc[2]
  int n_envs;
  int env_size;   // particles per environment
};

// device-side control block (one per swarm handle)
struct Ctl {
  unsigned int barrier;   // grid barrier arrival counter (zeroed before each cooperative launch)
  int halted;             // set when a run stopped early (Verlet overflow); later launches no-op
  int overflow;
  int rebuild_flag[3];    // rotating "some particle moved > skin/2" flags
  int max_nbr_seen;
  int pad;
  long long steps_done;   // steps completed by the last launch sequence
  long long rebuilds;
  long long tprof[8];         // SW_TIMING builds: clock64 sums of the resident kernel's phases (CTA 0, thread 0)
  unsigned long long bar64;
  unsigned long long tile_dmax;  // tiled kernel: (validation index << 32) | float bits of the largest displacement^2 of the period;
                                 // monotone within a launch, zeroed with bar64 (same memset, same line as the arrival counter)
  unsigned long long pad1[14];   // keep the release word on its own 128-byte line
  unsigned long long rel64;      // published by the last arriver: same encoding as bar64
  unsigned long long pad2[15];  // resident kernel's barrier: low 32 bits arrivals, high 32 bits CTAs with a mover
  // tiled kernel (swarm_tiled.cuh): rebuild cadence carried from launch to launch, and its statistics
  int tile_K;                    // steps between Verlet rebuilds chosen by the last launch (0: not chosen yet)
  int tile_since;                // steps integrated on the current list
  long long tile_rollbacks;      // periods replayed because a particle left skin/2 before the scheduled rebuild
  long long tile_partners;       // sum over rebuilds of the number of rim particles (lists pointing into another CTA's tile)
};

struct DevArrays {
  long long n;
  long long n_pad;
  double *x[2], *y[2], *z[2];  // [0] caller-bound, [1] library scratch (double buffer)
  double *theta, *q0, *q1, *q2, *q3;
  double *vx, *vy, *vz, *wx, *wy, *wz;
  double *f_swim, *tqx, *tqy, *tqz, *fex, *fey, *fez;
  const int *species;
  const int *n_dev;           // live particle count on the device (<= n), or NULL; see swarm_bind_count
  const long long *gid;       // Philox id of every slot (domain decomposition); NULL => the slot index
  const long long *okey;      // Verlet entries of a particle are kept in ascending okey[slot] (NULL => ascending slot): the force
                              // summation order when slots are spatially sorted and no longer the particle ids
  double *xr, *yr, *zr;        // positions at the last Verlet rebuild
  double *gkeep;               // k_bd2d_stream: cosine normal of the even step's rotational pair, per slot (swm_gaussians_xyt_keep)
  int *head, *next, *cellof;   // linked cells
  int *nbr, *nbr_cnt;          // Verlet list: nbr[k * n_pad + i], k < cap
  int cap;
  Ctl *ctl;
};

// live particle count: the device-resident one when bound (domain decomposition re-partitions without a host sync)
__device__ __forceinline__ long long sw_live_n(const DevArrays &A) {
  if (!A.n_dev) return A.n;
  const long long m = (long long)__ldg(A.n_dev);
  return m < A.n ? m : A.n;
}

// Philox key id of slot i: the global particle id when the slots hold a sub-domain of a larger swarm
__device__ __forceinline__ u64 sw_pid(const DevArrays &A, long long i) { return A.gid ? (u64)A.gid[i] : (u64)i; }

// ---------------------------------------------------------------------------------------------
// host handle
// ---------------------------------------------------------------------------------------------
// spatial slot order + tile scratch of the tiled 2-D kernel (swarm_sort.cuh, swarm_tiled.cuh); library-owned
struct TileHost {
  long long n = 0;             // particle count the arrays were sized / sorted for (0: nothing allocated)
  bool sorted = false;         // perm / species_s / gid_s describe the bound particles
  int launches_since_sort = 0;
  int *perm = nullptr, *key = nullptr, *hist = nullptr, *start = nullptr, *cursor = nullptr, *species_s = nullptr;
  long long *gid_s = nullptr;
  double *xs[2] = {nullptr, nullptr}, *ys[2] = {nullptr, nullptr};
  void *crec = nullptr;                    // CellRec[n_pad]: the rebuild's per-particle search records
  unsigned long long *head64 = nullptr;   // tagged cell heads of the tiled rebuild
  long long head64_cells = 0;
  unsigned long long tag_next = 1;         // first unused rebuild tag
  void *rim[2] = {nullptr, nullptr};   // RimRec[n_pad] per epoch parity (swarm_tiled.cuh)
  int nbins_alloc = 0;
  int mode = 1;                // swarm_set_tiling: 0 off, 1 automatic
  int resort_interval = 16;    // launches between two spatial sorts
  long long sorts = 0;
  long long launches_total = 0;  // tiled launches of this scratch: the flag epochs of a launch start above the previous launch's
};

struct swarm {
  swarm_config_t cfg;
  DevParams P;
  DevArrays A;
  bool species_set = false, wca_set = false, state_bound = false, actions_bound = false;
  swarm_actions_dev_t act{};
  swarm_state_dev_t st{};
  bool list_valid = false;
  int list_space = 0;        // index space of the Verlet list in A.nbr: 0 = particle ids, 1 = the tiled kernel's sorted slots
  TileHost tile;
  int integrator = 0;        // swarm_integrator_t
  int last_mode = 0;         // mode of the last run (steps_done bookkeeping)
  bool lv_resume = false;    // next Langevin run continues an interrupted chunk
  bool lv_was_resume = false;
  bool cells_dirty = false;  // scratch reused for new particles: linked cells reset on the next launch's stream
  long long launches = 0;
  long long steps_requested = 0;
  int num_sms = 0;
  int coop_blocks_per_sm = 0;
  bool persistent = true;
  long long ncell_alloc = 0;
  long long n_reserved = 0;   // scratch capacity in particles (swarm_reserve); bind_state reuses it
  const long long *gid = nullptr;
  // scratch owned by the library
  std::vector<void *> owned;
  double *zeros = nullptr;  // n zeros (for unbound optional arrays)
  // observables scratch
  void *obs_scratch = nullptr;
  size_t obs_scratch_bytes = 0;
  double *red_scratch = nullptr;
  double *field_scratch = nullptr;   // per-CTA reward sums + ticket of swarm_obs_field_mean
  int field_scratch_blocks = 0;
  void *vis_buf = nullptr;   // vision-cone cell structure + packed copies
  size_t vis_bytes = 0;
  long long vis_prev_n = -1;   // the bounding box of the previous vision call is usable for this n / range
  double vis_prev_range = 0.0;
  bool cluster_ok = true;               // k_bd2d_cluster can be launched on this device (cleared by a failed launch)
  bool dd_lean = false;                 // swarm_dd_set_lean: counters are cleared by the merge's last kernel, not by memsets
  int *dd_clean_counts = nullptr;       //   the counts8 buffer that kernel cleared (the next merge's output)
  double *dd_send_l = nullptr, *dd_send_r = nullptr;            //   the send buffers whose record counts it cleared
  double *dd_last_send_l = nullptr, *dd_last_send_r = nullptr;  //   the send buffers of the last pack
  int run_flags = 0;                    // swarm_run_part: SWARM_RUN_* of the run being enqueued
  unsigned int *dd_tickets = nullptr;   // swarm_dd_push: one ticket counter per face
  const int *vis_image = nullptr;   // swarm_bind_images: periodic image count per slot (domain decomposition), or NULL
  Ctl *ctl_host = nullptr;  // pinned
};

void sw_set_error(const std::string &msg);
int sw_cuda_fail(cudaError_t e, const char *what);
#define SW_CUDA(expr)                                             \
  do {                                                            \
    cudaError_t _e = (expr);                                      \
    if (_e != cudaSuccess) return sw_cuda_fail(_e, #expr);        \
  } while (0)

// ---------------------------------------------------------------------------------------------
// Philox (Salmon et al., SC'11).  The thermostat draws from Philox4x32-10 (sw_noise_words below); the policy's
// Gumbel noise from Philox4x64-10 with counter (call, 4 + 16 k, DOMAIN_POLICY, 0), key (particle id, seed) -- a
// different generator, so action sampling can never read the words of a thermostat kick.  Stream map: DESIGN.md section 3.
// ---------------------------------------------------------------------------------------------
#define SW_PHILOX_DOMAIN_POLICY 0x706f6c696379ULL /* "policy" */

__host__ __device__ __forceinline__ void sw_mulhilo(u64 a, u64 b, u64 &hi, u64 &lo) {
#ifdef __CUDA_ARCH__
  hi = __umul64hi(a, b);
  lo = a * b;
#else
  unsigned __int128 p = (unsigned __int128)a * b;
  hi = (u64)(p >> 64);
  lo = (u64)p;
#endif
}

__host__ __device__ __forceinline__ void sw_philox4x64_10(u64 c0, u64 c1, u64 c2, u64 c3, u64 k0, u64 k1, u64 out[4]) {
  const u64 M0 = 0xD2E7470EE14C6C93ULL, M1 = 0xCA5A826395121157ULL;
  const u64 W0 = 0x9E3779B97F4A7C15ULL, W1 = 0xBB67AE8584CAA73BULL;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    u64 hi0, lo0, hi1, lo1;
    sw_mulhilo(M0, c0, hi0, lo0);
    sw_mulhilo(M1, c2, hi1, lo1);
    u64 n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += W0; k1 += W1;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// Philox4x32-10 (Random123 philox4x32, 10 rounds).  A round is two 32 x 32 -> 64 multiplies (one IMAD.WIDE each on the
// GPU, both halves at once), two three-input XORs and the key bump.
__host__ __device__ __forceinline__ void sw_philox4x32_10(unsigned int c0, unsigned int c1, unsigned int c2, unsigned int c3,
                                                         unsigned int k0, unsigned int k1, unsigned int out[4]) {
  const unsigned int M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const u64 p0 = (u64)M0 * (u64)c0, p1 = (u64)M1 * (u64)c2;
    const unsigned int n0 = (unsigned int)(p1 >> 32) ^ c1 ^ k0, n2 = (unsigned int)(p0 >> 32) ^ c3 ^ k1;
    c0 = n0; c1 = (unsigned int)p1; c2 = n2; c3 = (unsigned int)p0;
    k0 += W0; k1 += W1;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// Round keys of Philox4x32-10 for key = (seed lo, seed hi): rk[2r], rk[2r+1] = the key of round r.
__host__ __device__ __forceinline__ void sw_key_schedule(u64 seed, unsigned int rk[20]) {
  unsigned int k0 = (unsigned int)seed, k1 = (unsigned int)(seed >> 32);
  for (int r = 0; r < 10; ++r) { rk[2 * r] = k0; rk[2 * r + 1] = k1; k0 += 0x9E3779B9u; k1 += 0xBB67AE85u; }
}

// sw_philox4x32_10 with the key schedule precomputed (DevParams.rk): a round is two IMAD.WIDE and two LOP3 with a
// constant-bank operand -- 81 instructions for the two blocks of a draw; the in-line key bumps, done by the uniform
// datapath, and the register moves around them cost another 60 (cuobjdump, sm_100a).
__device__ __forceinline__ void sw_philox4x32_10_rk(unsigned int c0, unsigned int c1, unsigned int c2, unsigned int c3,
                                                    const unsigned int *rk, unsigned int out[4]) {
  const unsigned int M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const unsigned int h0 = __umulhi(M0, c0), l0 = M0 * c0, h1 = __umulhi(M1, c2), l1 = M1 * c2;
    const unsigned int n0 = h1 ^ c1 ^ rk[2 * r], n2 = h0 ^ c3 ^ rk[2 * r + 1];
    c0 = n0; c1 = l1; c2 = n2; c3 = l0;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
// ONE block of a draw: words 2 blk, 2 blk + 1 of sw_noise_words (blk = 0, 1) -- a Box-Muller pair needs exactly one block
__device__ __forceinline__ void sw_noise_block_rk(const unsigned int *rk, u64 id, u64 step, u64 stream, unsigned int blk, u64 &w0, u64 &w1) {
  const unsigned int c0 = (unsigned int)step, c1 = (unsigned int)(step >> 32), c2 = (unsigned int)id;
  const unsigned int c3 = ((unsigned int)((id >> 32) & 0xFFFFFFULL) << 8) | ((unsigned int)(stream & 0x7FULL) << 1) | blk;
  unsigned int a[4];
  sw_philox4x32_10_rk(c0, c1, c2, c3, rk, a);
  w0 = ((u64)a[0] << 32) | (u64)a[1];
  w1 = ((u64)a[2] << 32) | (u64)a[3];
}
__device__ __forceinline__ void sw_noise_words_rk(const unsigned int *rk, u64 id, u64 step, u64 stream, u64 r[4]) {
  sw_noise_block_rk(rk, id, step, stream, 0u, r[0], r[1]);
  sw_noise_block_rk(rk, id, step, stream, 1u, r[2], r[3]);
}

// The thermostat's four 64-bit noise words of (seed, particle id, step, stream) -- oracle: orc_noise_words.  Two
// Philox4x32-10 blocks b = 0, 1: key = (seed lo, seed hi), counter = (step lo, step hi, id lo,
// id[32..55] << 8 | stream << 1 | b); words 2b, 2b + 1 = block b's (out[0] << 32 | out[1]), (out[2] << 32 | out[3]) -- each
// Box-Muller pair (words 0,1 and 2,3) comes from ONE block, so a consumer of one pair draws one block (the 2-D Brownian
// step shares its rotational pair between two steps: swm_gaussians_xyt).  The key is the SEED, uniform over the
// swarm, so its ten-round schedule costs nothing per particle; the 20 rounds are ~85 instructions where the
// Philox4x64-10 block of round 1 (64-bit multiplies out of 32-bit ones) took 248.  Stream map: DESIGN.md section 3.
__host__ __device__ __forceinline__ void sw_noise_words(u64 seed, u64 id, u64 step, u64 stream, u64 r[4]) {
  const unsigned int k0 = (unsigned int)seed, k1 = (unsigned int)(seed >> 32);
  const unsigned int c0 = (unsigned int)step, c1 = (unsigned int)(step >> 32), c2 = (unsigned int)id;
  const unsigned int c3 = ((unsigned int)((id >> 32) & 0xFFFFFFULL) << 8) | ((unsigned int)(stream & 0x7FULL) << 1);
  unsigned int a[4], b[4];
  sw_philox4x32_10(c0, c1, c2, c3, k0, k1, a);
  sw_philox4x32_10(c0, c1, c2, c3 | 1u, k0, k1, b);
  r[0] = ((u64)a[0] << 32) | (u64)a[1]; r[1] = ((u64)a[2] << 32) | (u64)a[3];
  r[2] = ((u64)b[0] << 32) | (u64)b[1]; r[3] = ((u64)b[2] << 32) | (u64)b[3];
}

__device__ __forceinline__ double sw_u01(u64 r) {
  return ((double)(r >> 12) + 0.5) * (1.0 / 4503599627370496.0);
}

// four standard normals for (seed, id, step, stream): Box-Muller on (u0,u1), (u2,u3)
__device__ __forceinline__ void sw_gaussians(u64 seed, u64 id, u64 step, u64 stream, double g[4]) {
  u64 r[4];
  sw_noise_words(seed, id, step, stream, r);
  double m0 = sqrt(-2.0 * log(sw_u01(r[0])));
  double m1 = sqrt(-2.0 * log(sw_u01(r[2])));
  double s0, c0, s1, c1;
  sincospi(2.0 * sw_u01(r[1]), &s0, &c0);
  sincospi(2.0 * sw_u01(r[3]), &s1, &c1);
  g[0] = m0 * c0; g[1] = m0 * s0; g[2] = m1 * c1; g[3] = m1 * s1;
}

// ---------------------------------------------------------------------------------------------
// geometry.  Everything that decides SET MEMBERSHIP (cell index, r < cut) uses the explicit
// round-to-nearest intrinsics so nvcc cannot contract it into FMAs: the oracle is compiled with
// -ffp-contract=off and the neighbour / cell lists must agree bit for bit.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ double sw_min_image(double d, double L, double invL, int periodic) {
  if (periodic) d = __dsub_rn(d, __dmul_rn(L, rint(__dmul_rn(d, invL))));
  return d;
}

__device__ __forceinline__ int sw_cell_coord(double x, double invL, int nc, int periodic) {
  double u = __dmul_rn(x, invL);
  if (periodic) u = __dsub_rn(u, floor(u));
  double cf = floor(__dmul_rn(u, (double)nc));
  int c;
  if (cf < 0.0) c = 0;
  else if (cf >= (double)nc) c = nc - 1;
  else c = (int)cf;
  return c;
}

__device__ __forceinline__ double sw_r2(double dx, double dy, double dz) {
  return __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
}

// WCA pair-force factor: F_i += fac * (x_i - x_j)   (oracle: wca_factor)
__device__ __forceinline__ double sw_wca_factor(double r2, double sigma, double eps) {
  double f2 = sigma * sigma / r2;
  double f6 = f2 * f2 * f2;
  return 48.0 * eps * f6 * (f6 - 0.5) / r2;
}

// loads of data that other CTAs rewrite inside the same (persistent) kernel: bypass L1
__device__ __forceinline__ double ldcg(const double *p) { return __ldcg(p); }
__device__ __forceinline__ int ldcg(const int *p) { return __ldcg(p); }
__device__ __forceinline__ long long ldcg(const long long *p) { return __ldcg(p); }

__device__ __forceinline__ void sw_quat_to_mat(const double q[4], double R[3][3]) {
  double a = q[0], b = q[1], c = q[2], d = q[3];
  R[0][0] = a * a + b * b - c * c - d * d; R[0][1] = 2 * (b * c - a * d); R[0][2] = 2 * (b * d + a * c);
  R[1][0] = 2 * (b * c + a * d); R[1][1] = a * a - b * b + c * c - d * d; R[1][2] = 2 * (c * d - a * b);
  R[2][0] = 2 * (b * d - a * c); R[2][1] = 2 * (c * d + a * b); R[2][2] = a * a - b * b - c * c + d * d;
}

__device__ __forceinline__ void sw_quat_rotate_body(double q[4], const double ax[3], double angle) {
  double s, c;
  sincos(0.5 * angle, &s, &c);
  double r0 = c, r1 = s * ax[0], r2 = s * ax[1], r3 = s * ax[2];
  double o0 = q[0] * r0 - q[1] * r1 - q[2] * r2 - q[3] * r3;
  double o1 = q[0] * r1 + q[1] * r0 + q[2] * r3 - q[3] * r2;
  double o2 = q[0] * r2 - q[1] * r3 + q[2] * r0 + q[3] * r1;
  double o3 = q[0] * r3 + q[1] * r2 - q[2] * r1 + q[3] * r0;
  double nrm = sqrt(o0 * o0 + o1 * o1 + o2 * o2 + o3 * o3);
  q[0] = o0 / nrm; q[1] = o1 / nrm; q[2] = o2 / nrm; q[3] = o3 / nrm;
}

// launch bookkeeping
static inline int sw_blocks(long long n, int threads) { return (int)((n + threads - 1) / threads); }
