// swarm_obs.cu -- per-agent observable and task-reward kernels (sm_100a) and their C-ABI.
//
// Replaces the per-colloid Python/JAX loops of
//   swarmrl/observables/concentration_field.py:84-138, swarmrl/tasks/searching/gradient_sensing.py:92-150,
//   swarmrl/observables/subdivided_vision_cones.py:83-258 (+ swarmrl/utils/utils.py:297-332),
//   swarmrl/observables/position.py:32-60 and the mean of swarmrl/trainers/trainer.py:92-95.
//
// Vision cones are O(N^2) in the reference (every observer visits every particle).  Here the
// particles are binned on a uniform grid of edge >= vision_range laid over the bounding box of the
// UNFOLDED positions (the reference takes raw differences, no minimum image):
//   cell hash -> counting sort on the cell key (a one-digit radix sort: histogram, exclusive scan =
//   cell-start table, scatter) -> per-cell index sort (makes the order deterministic and equal to
//   what a stable radix sort yields) -> packed cell-ordered copies of (x, y, z, radius, column) ->
//   one thread per observer walks the 3^d neighbouring cells, each row of which is ONE contiguous
//   range of the packed arrays.
#include <math.h>
#include <stdlib.h>
#include <mutex>

#include "swarm_common.cuh"

struct VisGrid {
  double lo[3];
  double cs;       // cell edge
  double inv_cs;
  int nc[3];
  int ncell;
  unsigned long long mn[3], mx[3];  // ordered-uint encodings for atomicMin / atomicMax
};

__device__ __forceinline__ unsigned long long enc_double(double v) {
  unsigned long long b = (unsigned long long)__double_as_longlong(v);
  return (b & 0x8000000000000000ULL) ? ~b : (b | 0x8000000000000000ULL);
}
__device__ __forceinline__ double dec_double(unsigned long long e) {
  unsigned long long b = (e & 0x8000000000000000ULL) ? (e & 0x7FFFFFFFFFFFFFFFULL) : ~e;
  return __longlong_as_double((long long)b);
}

__global__ void k_vis_init(VisGrid *G, int *count, long long ncell_max) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i == 0)
    for (int a = 0; a < 3; ++a) { G->mn[a] = ~0ULL; G->mx[a] = 0ULL; }
  for (; i <= ncell_max; i += (long long)gridDim.x * blockDim.x) count[i] = 0;
}

__global__ void k_vis_bbox(long long n, const double *x, const double *y, const double *z, VisGrid *G) {
  __shared__ unsigned long long smn[3][8], smx[3][8];
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  unsigned long long mn[3] = {~0ULL, ~0ULL, ~0ULL}, mx[3] = {0ULL, 0ULL, 0ULL};
  for (; i < n; i += (long long)gridDim.x * blockDim.x) {
    unsigned long long e[3] = {enc_double(x[i]), enc_double(y[i]), enc_double(z[i])};
#pragma unroll
    for (int a = 0; a < 3; ++a) { mn[a] = min(mn[a], e[a]); mx[a] = max(mx[a], e[a]); }
  }
#pragma unroll
  for (int a = 0; a < 3; ++a)
    for (int o = 16; o > 0; o >>= 1) {
      mn[a] = min(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], o));
      mx[a] = max(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], o));
    }
  int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0)
    for (int a = 0; a < 3; ++a) { smn[a][w] = mn[a]; smx[a][w] = mx[a]; }
  __syncthreads();
  if (threadIdx.x < 3) {
    int a = threadIdx.x;
    unsigned long long m0 = ~0ULL, m1 = 0ULL;
    for (int k = 0; k < (int)(blockDim.x >> 5); ++k) { m0 = min(m0, smn[a][k]); m1 = max(m1, smx[a][k]); }
    atomicMin(&G->mn[a], m0);
    atomicMax(&G->mx[a], m1);
  }
}

__global__ void k_vis_grid(VisGrid *G, double range, long long ncell_max) {
  if (blockIdx.x != 0 || threadIdx.x != 0) return;
  double ext[3];
  for (int a = 0; a < 3; ++a) {
    double lo = dec_double(G->mn[a]), hi = dec_double(G->mx[a]);
    if (!(hi >= lo)) { lo = 0.0; hi = 0.0; }
    G->lo[a] = lo;
    ext[a] = hi - lo;
  }
  double cs = range > 0.0 ? range : 1.0;
  for (;;) {
    long long tot = 1;
    for (int a = 0; a < 3; ++a) {
      double f = floor(ext[a] / cs) + 1.0;
      int nc = (f > 1048576.0) ? 1048576 : (int)f;
      if (nc < 1) nc = 1;
      G->nc[a] = nc;
      tot *= nc;
      if (tot > ncell_max) break;
    }
    if (tot <= ncell_max) { G->ncell = (int)tot; break; }
    cs *= 1.5;
  }
  G->cs = cs;
  G->inv_cs = 1.0 / cs;
}

__device__ __forceinline__ int vis_coord(double v, double lo, double inv_cs, int nc) {
  double cf = floor((v - lo) * inv_cs);
  if (!(cf >= 0.0)) return 0;
  if (cf >= (double)nc) return nc - 1;
  return (int)cf;
}

__global__ void k_vis_count(long long n, const double *x, const double *y, const double *z, const VisGrid *G,
                            int *cellid, int *count) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int cx = vis_coord(x[i], G->lo[0], G->inv_cs, G->nc[0]);
  int cy = vis_coord(y[i], G->lo[1], G->inv_cs, G->nc[1]);
  int cz = vis_coord(z[i], G->lo[2], G->inv_cs, G->nc[2]);
  int c = (cz * G->nc[1] + cy) * G->nc[0] + cx;
  cellid[i] = c;
  atomicAdd(&count[c], 1);
}

// ---- exclusive scan of count[0..m) (m = ncell_max + 1) in three launches ---------------------
#define SCAN_T 256
#define SCAN_E 4
#define SCAN_TILE (SCAN_T * SCAN_E)

__device__ __forceinline__ int block_excl_scan(int v, int *total) {
  __shared__ int wsum[SCAN_T / 32];
  int l = threadIdx.x & 31, w = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int t = __shfl_up_sync(0xffffffffu, inc, o);
    if (l >= o) inc += t;
  }
  if (l == 31) wsum[w] = inc;
  __syncthreads();
  if (w == 0) {
    int s = (l < SCAN_T / 32) ? wsum[l] : 0;
#pragma unroll
    for (int o = 1; o < SCAN_T / 32; o <<= 1) {
      int t = __shfl_up_sync(0xffffffffu, s, o);
      if (l >= o) s += t;
    }
    if (l < SCAN_T / 32) wsum[l] = s;
  }
  __syncthreads();
  int base = (w > 0) ? wsum[w - 1] : 0;
  *total = wsum[SCAN_T / 32 - 1];
  __syncthreads();
  return base + inc - v;
}

__global__ void k_scan_tiles(long long m, const int *in, int *out, int *tile_sum) {
  long long base = (long long)blockIdx.x * SCAN_TILE + (long long)threadIdx.x * SCAN_E;
  int v[SCAN_E], s = 0;
#pragma unroll
  for (int k = 0; k < SCAN_E; ++k) { v[k] = (base + k < m) ? in[base + k] : 0; s += v[k]; }
  int total;
  int ex = block_excl_scan(s, &total);
#pragma unroll
  for (int k = 0; k < SCAN_E; ++k) { if (base + k < m) out[base + k] = ex; ex += v[k]; }
  if (threadIdx.x == 0) tile_sum[blockIdx.x] = total;
}

__global__ void k_scan_sums(int ntiles, int *tile_sum) {
  // single block: sequential over chunks of SCAN_T
  __shared__ int carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int b = 0; b < ntiles; b += SCAN_T) {
    int idx = b + threadIdx.x;
    int v = (idx < ntiles) ? tile_sum[idx] : 0;
    int total;
    int ex = block_excl_scan(v, &total);
    int c = carry;
    if (idx < ntiles) tile_sum[idx] = ex + c;
    __syncthreads();
    if (threadIdx.x == 0) carry = c + total;
    __syncthreads();
  }
}

__global__ void k_scan_add(long long m, int *out, const int *tile_sum, int *fill) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  int v = out[i] + tile_sum[i / SCAN_TILE];
  out[i] = v;
  fill[i] = v;
}

__global__ void k_vis_fill(long long n, const int *cellid, int *fill, int *items) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int p = atomicAdd(&fill[cellid[i]], 1);
  items[p] = (int)i;
}

// per-cell insertion sort by particle index: deterministic, == stable radix sort order
__global__ void k_vis_sortcell(const VisGrid *G, const int *start, int *items) {
  long long c = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= G->ncell) return;
  int b = start[c], e = start[c + 1];
  for (int k = b + 1; k < e; ++k) {
    int v = items[k], j = k;
    while (j > b && items[j - 1] > v) { items[j] = items[j - 1]; --j; }
    items[j] = v;
  }
}

__global__ void k_vis_pack(long long n, const int *items, const double *x, const double *y, const double *z,
                           const double *radii, const int *species, DevParams P, const int *type_col, int n_type_col,
                           double *sx, double *sy, double *sz, double *sr, int *scol) {
  long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  int i = items[k];
  sx[k] = x[i]; sy[k] = y[i]; sz[k] = z[i]; sr[k] = radii[i];
  int t = P.sp_type[species[i]];
  scol[k] = (t >= 0 && t < n_type_col) ? type_col[t] : -1;
}


// ---------------------------------------------------------------------------------------------
// k_vis_build -- the whole grid construction above (bounding box, grid, cell count, cell-start scan, placement,
// per-cell order, packed copies) in ONE cooperative launch with five grid barriers instead of ten dependent launches
// of 3-8 us each (profiles/r01_launch_shares.txt: 45 us of a 690 us slice).  Same arithmetic, phase by phase, as the
// k_vis_* / k_scan_* kernels, which stay as the fallback for devices without cooperative launch.
// The memset before the launch zeroes VisBuild and count[]; the bounding box is kept as two maxima (the minimum as
// the maximum of the complement) so that zero is the neutral start of both.
// ---------------------------------------------------------------------------------------------
struct VisBuild {
  unsigned long long mnc[3], mx[3];   // max of ~enc(v), max of enc(v)
  unsigned int bar;
  unsigned int pad;
};
// Bounding box of the previous call (outside the memset).  A grid laid over ANY box is correct -- coordinates are clamped
// into the edge cells, which keeps cell indices of two particles within `range` of each other at most one apart -- so the
// box of the previous slice, widened by one range, saves the barrier between "bounding box" and "cell of every particle".
struct VisPrev { unsigned long long mnc[3], mx[3]; };

__device__ __forceinline__ void vis_grid_barrier(unsigned int *counter, unsigned int &epoch) {
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

#define VISB_T 512
__global__ void __launch_bounds__(VISB_T) k_vis_build(long long n, const double *x, const double *y, const double *z,
                                                      const double *radii, const int *species, const __grid_constant__ DevParams P,
                                                      const int *type_col, int n_type_col, double range, long long ncell_max,
                                                      VisBuild *B, VisPrev *prev, int use_prev, VisGrid *G, int *count, int *start, int *fill, int *cellid,
                                                      int *items, double *sx, double *sy, double *sz, double *sr, int *scol,
                                                      double *out, long long out_len) {
  unsigned int epoch = 0;
  const long long gtid = (long long)blockIdx.x * blockDim.x + threadIdx.x, gsz = (long long)gridDim.x * blockDim.x;
  __shared__ unsigned long long smn[3][VISB_T / 32], smx[3][VISB_T / 32];
  __shared__ VisGrid g;
  // ---- phase 0: zero the output, bounding box of the unfolded positions ------------------------------------------
  for (long long i = gtid; i < out_len; i += gsz) out[i] = 0.0;
  {
    unsigned long long mn[3] = {0ULL, 0ULL, 0ULL}, mx[3] = {0ULL, 0ULL, 0ULL};
    for (long long i = gtid; i < n; i += gsz) {
      const unsigned long long e[3] = {enc_double(x[i]), enc_double(y[i]), enc_double(z[i])};
#pragma unroll
      for (int a = 0; a < 3; ++a) { mn[a] = max(mn[a], ~e[a]); mx[a] = max(mx[a], e[a]); }
    }
#pragma unroll
    for (int a = 0; a < 3; ++a)
      for (int o = 16; o > 0; o >>= 1) {
        mn[a] = max(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], o));
        mx[a] = max(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], o));
      }
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0)
      for (int a = 0; a < 3; ++a) { smn[a][w] = mn[a]; smx[a][w] = mx[a]; }
    __syncthreads();
    if (threadIdx.x < 3) {
      const int a = threadIdx.x;
      unsigned long long m0 = 0ULL, m1 = 0ULL;
      for (int k = 0; k < VISB_T / 32; ++k) { m0 = max(m0, smn[a][k]); m1 = max(m1, smx[a][k]); }
      atomicMax(&B->mnc[a], m0);
      atomicMax(&B->mx[a], m1);
    }
  }
  if (!use_prev) vis_grid_barrier(&B->bar, epoch);
  // ---- phase 1: the grid (every CTA computes the same one), cell of every particle, histogram -----------------------
  if (threadIdx.x == 0) {
    double ext[3];
    for (int a = 0; a < 3; ++a) {
      double lo, hi;
      if (use_prev) {
        lo = dec_double(~prev->mnc[a]); hi = dec_double(prev->mx[a]);
        if (hi > lo || a < 2) { lo -= range; hi += range; }   // a flat axis (2-D: z) stays flat
      } else {
        lo = dec_double(~__ldcg(&B->mnc[a])); hi = dec_double(__ldcg(&B->mx[a]));
      }
      if (!(hi >= lo)) { lo = 0.0; hi = 0.0; }
      g.lo[a] = lo;
      ext[a] = hi - lo;
    }
    double cs = range > 0.0 ? range : 1.0;
    for (;;) {
      long long tot = 1;
      for (int a = 0; a < 3; ++a) {
        double f = floor(ext[a] / cs) + 1.0;
        int nc = (f > 1048576.0) ? 1048576 : (int)f;
        if (nc < 1) nc = 1;
        g.nc[a] = nc;
        tot *= nc;
        if (tot > ncell_max) break;
      }
      if (tot <= ncell_max) { g.ncell = (int)tot; break; }
      cs *= 1.5;
    }
    g.cs = cs;
    g.inv_cs = 1.0 / cs;
    if (blockIdx.x == 0) *G = g;   // for the cone kernel
  }
  __syncthreads();
  for (long long i = gtid; i < n; i += gsz) {
    const int cx = vis_coord(x[i], g.lo[0], g.inv_cs, g.nc[0]);
    const int cy = vis_coord(y[i], g.lo[1], g.inv_cs, g.nc[1]);
    const int cz = vis_coord(z[i], g.lo[2], g.inv_cs, g.nc[2]);
    const int c = (cz * g.nc[1] + cy) * g.nc[0] + cx;
    cellid[i] = c;
    atomicAdd(&count[c], 1);
  }
  vis_grid_barrier(&B->bar, epoch);
  // ---- phase 2: exclusive scan of the histogram = cell start (CTA 0; the grid has ~3 particles per cell) -------------
  if (blockIdx.x == 0) {
    __shared__ int wsum[VISB_T / 32];
    __shared__ int carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int m = g.ncell + 1;
    for (int b0 = 0; b0 < m; b0 += VISB_T * 8) {
      // eight consecutive cells per thread: 4096 per round
      const int b = b0 + threadIdx.x * 8;
      int v[8], tsum = 0;
#pragma unroll
      for (int q = 0; q < 8; ++q) { v[q] = (b + q < m) ? __ldcg(&count[b + q]) : 0; tsum += v[q]; }
      int inc = tsum;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += t;
      }
      if (lane == 31) wsum[warp] = inc;
      __syncthreads();
      if (warp == 0) {
        const int wv = (lane < VISB_T / 32) ? wsum[lane] : 0;
        int winc = wv;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const int t = __shfl_up_sync(0xffffffffu, winc, d);
          if (lane >= d) winc += t;
        }
        if (lane < VISB_T / 32) wsum[lane] = winc - wv;
      }
      __syncthreads();
      int excl = carry + wsum[warp] + inc - tsum;
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        if (b + q < m) { start[b + q] = excl; fill[b + q] = excl; }
        excl += v[q];
      }
      __syncthreads();
      if (threadIdx.x == VISB_T - 1) carry = excl;
      __syncthreads();
    }
  }
  vis_grid_barrier(&B->bar, epoch);
  // ---- phase 3: placement -----------------------------------------------------------------------------------------
  for (long long i = gtid; i < n; i += gsz) {
    const int p = atomicAdd(&fill[cellid[i]], 1);
    items[p] = (int)i;
  }
  vis_grid_barrier(&B->bar, epoch);
  // ---- phase 4: ids of a cell in ascending order (deterministic; == a stable radix sort) -----------------------------
  for (long long c = gtid; c < g.ncell; c += gsz) {
    const int b = __ldcg(&start[c]), e = __ldcg(&start[c + 1]);
    for (int k = b + 1; k < e; ++k) {
      const int v = __ldcg(&items[k]);
      int j = k;
      while (j > b && __ldcg(&items[j - 1]) > v) { items[j] = __ldcg(&items[j - 1]); --j; }
      items[j] = v;
    }
  }
  vis_grid_barrier(&B->bar, epoch);
  // ---- phase 5: packed, cell-ordered copies ---------------------------------------------------------------------------
  for (long long k = gtid; k < n; k += gsz) {
    const int i = __ldcg(&items[k]);
    sx[k] = x[i]; sy[k] = y[i]; sz[k] = z[i]; sr[k] = radii[i];
    const int t = P.sp_type[species[i]];
    scol[k] = (t >= 0 && t < n_type_col) ? type_col[t] : -1;
  }
  if (gtid < 3) { prev->mnc[gtid] = __ldcg(&B->mnc[gtid]); prev->mx[gtid] = __ldcg(&B->mx[gtid]); }   // complete since the first barrier passed
}

#define VIS_MAX_ACC 64

// one thread per cell-ordered slot; observers (row_of >= 0) accumulate their cones
__global__ void __launch_bounds__(128) k_vis_cones(long long n, const int *items, const int *row_of, const VisGrid *G,
                                                   const int *start, const double *sx, const double *sy,
                                                   const double *sz, const double *sr, const int *scol, int n_dims,
                                                   const double *theta, const double *q0, const double *q1,
                                                   const double *q2, const double *q3, const double *dirs,
                                                   double range, double half_angle, int n_cones, int n_cols,
                                                   double *out) {
  long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int i = items[k];
  const int row = row_of[i];
  if (row < 0) return;
  double acc[VIS_MAX_ACC];
  const int nacc = n_cones * n_cols;
  for (int a = 0; a < nacc; ++a) acc[a] = 0.0;
  const double xi = sx[k], yi = sy[k], zi = sz[k];
  double ex, ey, ez;
  if (dirs) { ex = dirs[i]; ey = dirs[n + i]; ez = dirs[2 * n + i]; }
  else if (n_dims == 2) { sincos(theta[i], &ey, &ex); ez = 0.0; }
  else {
    double a = q0[i], b = q1[i], c = q2[i], d = q3[i];
    ex = 2 * (b * d + a * c); ey = 2 * (c * d - a * b); ez = a * a - b * b - c * c + d * d;
  }
  {
    double en = sqrt(ex * ex + ey * ey + ez * ez);
    ex /= en; ey /= en; ez /= en;
  }
  const int cx = vis_coord(xi, G->lo[0], G->inv_cs, G->nc[0]);
  const int cy = vis_coord(yi, G->lo[1], G->inv_cs, G->nc[1]);
  const int cz = vis_coord(zi, G->lo[2], G->inv_cs, G->nc[2]);
  const int x0 = max(cx - 1, 0), x1 = min(cx + 1, G->nc[0] - 1);
  for (int zz = max(cz - 1, 0); zz <= min(cz + 1, G->nc[2] - 1); ++zz)
    for (int yy = max(cy - 1, 0); yy <= min(cy + 1, G->nc[1] - 1); ++yy) {
      const long long rowbase = ((long long)zz * G->nc[1] + yy) * G->nc[0];
      const int b = start[rowbase + x0], e = start[rowbase + x1 + 1];
      for (int m = b; m < e; ++m) {
        const int col = __ldg(&scol[m]);
        if (col < 0) continue;
        double dx = __ldg(&sx[m]) - xi, dy = __ldg(&sy[m]) - yi, dz = __ldg(&sz[m]) - zi;
        double d = sqrt(dx * dx + dy * dy + dz * dz);
        if (!(d < range) || d == 0.0) continue;
        double w = 2.0 * __ldg(&sr[m]) / d;
        if (w > 1.0) w = 1.0;
        double s0 = dx / d, s1 = dy / d, s2 = dz / d;
        double sn = sqrt(s0 * s0 + s1 * s1 + s2 * s2);
        s0 /= sn; s1 /= sn; s2 /= sn;
        double c = s0 * ex + s1 * ey + s2 * ez;
        c = fmin(1.0, fmax(-1.0, c));
        double ang = acos(c);
        double od = s0 * (-ey) + s1 * ex + s2 * ez;
        if (!(od >= 0.0)) ang = -ang;
        for (int kc = 0; kc < n_cones; ++kc) {
          double lo = -half_angle + (double)kc * half_angle * 2.0 / (double)n_cones;
          double hi = -half_angle + (double)(kc + 1) * half_angle * 2.0 / (double)n_cones;
          if (lo < ang && hi > ang) acc[kc * n_cols + col] += w;
        }
      }
    }
  double *o = out + (long long)row * nacc;
  for (int a = 0; a < nacc; ++a) o[a] = acc[a];
}

// One WARP per cell-ordered slot: the ~30 candidates of an observer (3 x 3 cells of edge `range`) are
// spread over the lanes, so the acos / divisions of one observer run 32-wide instead of in one thread's
// serial loop (the thread-per-observer kernel above was latency-bound: 128 us for 100k observers).  Each
// lane keeps private cone sums in shared memory (conflict-free [acc][thread] layout), combined at the end
// by a fixed shuffle tree: deterministic, though no longer in ascending particle order.
#define VISW_THREADS 128
#define VISW_MAX_ACC 32
// One candidate (position, radius, output column) seen by the observer at (xi, yi, zi) with unit director e: its weight goes
// into the lane's private cone sums.  Shared by the two cone kernels below (cell-start layout and fixed-capacity layout).
__device__ __forceinline__ void visw_candidate(double xm, double ym, double zm, double rm, int col, double xi, double yi, double zi,
                                               double ex, double ey, double ez, double range, double half_angle, double cone_w,
                                               int n_cones, int n_cols, const double *rim, const double *crim, double *sacc) {
        if (col < 0) return;
        double dx = xm - xi, dy = ym - yi, dz = zm - zi;
        double d = sqrt(dx * dx + dy * dy + dz * dz);
        if (!(d < range) || d == 0.0) return;
        double w = 2.0 * rm / d;
        if (w > 1.0) w = 1.0;
        {
          // fast cone decision: compare cos(angle) with the cosines of the rims instead of taking acos and
          // normalising the separation twice.  Whenever the cosine lies within 1e-9 of a rim's (or the side
          // test within 1e-9 of zero) the exact sequence below decides, so the chosen cone is always the one the
          // serial kernel / the oracle choose; everywhere else the margin exceeds the rounding by six decades.
          const double inv_d = 1.0 / d;
          const double cf = (dx * ex + dy * ey + dz * ez) * inv_d;
          const double of = (dx * (-ey) + dy * ex + dz * ez) * inv_d;
          bool near = fabs(of) < 1e-9;
          int tf = -1;
          for (int t = 0; t < n_cones; ++t) {
            const double r0 = rim[t], r1 = rim[t + 1], c0 = crim[t], c1 = crim[t + 1];
            near = near || fabs(cf - c0) < 1e-9 || fabs(cf - c1) < 1e-9;
            bool in;
            if (of >= 0.0) in = (r0 < 0.0 || cf < c0) && (r1 > 0.0 && cf > c1);      // r0 < acos(c) < r1
            else in = (r0 < 0.0 && cf > c0) && (r1 > 0.0 || cf < c1);               // r0 < -acos(c) < r1
            if (in) tf = t;
          }
          if (!near) {
            if (tf >= 0) sacc[(tf * n_cols + col) * VISW_THREADS + threadIdx.x] += w;
            return;
          }
        }
        double s0 = dx / d, s1 = dy / d, s2 = dz / d;
        double sn = sqrt(s0 * s0 + s1 * s1 + s2 * s2);
        s0 /= sn; s1 /= sn; s2 /= sn;
        double c = s0 * ex + s1 * ey + s2 * ez;
        c = fmin(1.0, fmax(-1.0, c));
        double ang = acos(c);
        double od = s0 * (-ey) + s1 * ex + s2 * ez;
        if (!(od >= 0.0)) ang = -ang;
        // the rims are evaluated exactly as in the serial kernel; only the candidate cone is guessed
        int kc = (int)floor((ang + half_angle) / cone_w);
        for (int t = kc - 1; t <= kc + 1; ++t) {
          if (t < 0 || t >= n_cones) continue;
          double lo = -half_angle + (double)t * half_angle * 2.0 / (double)n_cones;
          double hi = -half_angle + (double)(t + 1) * half_angle * 2.0 / (double)n_cones;
          if (lo < ang && hi > ang) sacc[(t * n_cols + col) * VISW_THREADS + threadIdx.x] += w;
        }
}

template <int GS>
__global__ void __launch_bounds__(VISW_THREADS) k_vis_cones_warp(long long n, const int *items, const int *row_of, const VisGrid *G,
                                                                  const int *start, const double *sx, const double *sy,
                                                                  const double *sz, const double *sr, const int *scol, int n_dims,
                                                                  const double *theta, const double *q0, const double *q1,
                                                                  const double *q2, const double *q3, const double *dirs,
                                                                  double range, double half_angle, int n_cones, int n_cols,
                                                                  double *out) {
  extern __shared__ double sacc[];   // [nacc][VISW_THREADS]
  const int lane = threadIdx.x & (GS - 1);   // GS lanes share one observer
  long long k = ((long long)blockIdx.x * blockDim.x + threadIdx.x) / GS;
  const bool in_range = k < n;
  if (!in_range) k = n - 1;          // no early exit: every lane takes part in the shuffles below
  const int i = items[k];
  const int row = in_range ? row_of[i] : -1;
  const int nacc = n_cones * n_cols;
  for (int a = 0; a < nacc; ++a) sacc[a * VISW_THREADS + threadIdx.x] = 0.0;
  const double xi = sx[k], yi = sy[k], zi = sz[k];
  double ex, ey, ez;
  if (dirs) { ex = dirs[i]; ey = dirs[n + i]; ez = dirs[2 * n + i]; }
  else if (n_dims == 2) { sincos(theta[i], &ey, &ex); ez = 0.0; }
  else {
    double a = q0[i], b = q1[i], c = q2[i], d = q3[i];
    ex = 2 * (b * d + a * c); ey = 2 * (c * d - a * b); ez = a * a - b * b - c * c + d * d;
  }
  {
    double en = sqrt(ex * ex + ey * ey + ez * ez);
    ex /= en; ey /= en; ez /= en;
  }
  const double cone_w = half_angle * 2.0 / (double)n_cones;
  // rims and their cosines, once per CTA (the rim expression is the serial kernel's, bit for bit)
  __shared__ double rim[VIS_MAX_ACC + 1], crim[VIS_MAX_ACC + 1];
  for (int t = threadIdx.x; t <= n_cones; t += blockDim.x) {
    rim[t] = -half_angle + (double)t * half_angle * 2.0 / (double)n_cones;
    crim[t] = cos(rim[t]);
  }
  __syncthreads();
  const int cx = vis_coord(xi, G->lo[0], G->inv_cs, G->nc[0]);
  const int cy = vis_coord(yi, G->lo[1], G->inv_cs, G->nc[1]);
  const int cz = vis_coord(zi, G->lo[2], G->inv_cs, G->nc[2]);
  const int x0 = max(cx - 1, 0), x1 = min(cx + 1, G->nc[0] - 1);
  for (int zz = max(cz - 1, 0); zz <= min(cz + 1, G->nc[2] - 1); ++zz)
    for (int yy = max(cy - 1, 0); yy <= min(cy + 1, G->nc[1] - 1); ++yy) {
      const long long rowbase = ((long long)zz * G->nc[1] + yy) * G->nc[0];
      const int b = start[rowbase + x0], e = (row >= 0) ? start[rowbase + x1 + 1] : b;
      for (int m = b + lane; m < e; m += GS) {
        visw_candidate(__ldg(&sx[m]), __ldg(&sy[m]), __ldg(&sz[m]), __ldg(&sr[m]), __ldg(&scol[m]), xi, yi, zi, ex, ey, ez, range, half_angle,
                       cone_w, n_cones, n_cols, rim, crim, sacc);
      }
    }
  __syncwarp();
  for (int a = 0; a < nacc; ++a) {
    double v = sacc[a * VISW_THREADS + threadIdx.x];
#pragma unroll
    for (int off = GS / 2; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    if (lane == 0 && row >= 0) out[(long long)row * nacc + a] = v;
  }
}


// ---------------------------------------------------------------------------------------------
// Vision-cone grid, second layout: FIXED CAPACITY per cell -- no histogram, no scan, no scatter.
//   k_vis2_fill : cell of every particle on the grid of the PREVIOUS call's bounding box (widened by one range; any
//                 box is correct, see VisPrev); slot = atomicAdd(count[cell]); the first VIS2_CAP ids of a cell go into
//                 its slot row, the rest into one overflow list; the bounding box of THIS call is reduced on the side.
//   k_vis2_pack : one warp per cell ranks the ids of its row (ascending id: a deterministic candidate order) and writes
//                 the packed records {x, y, z, radius} + column in rank order.
//   k_vis2_cones: as k_vis_cones_warp, walking the (up to) 27 slot rows, then the overflow list when it is not empty.
// Three dependent launches instead of a cooperative kernel with five grid barriers (k_vis_build: 66 us at C4 against
// ~20 us for these three; profiles/r02_launches_engine.csv).  The very first call has no previous box: it takes the
// cell-start path once, which also leaves the box behind.
// ---------------------------------------------------------------------------------------------
#define VIS2_CAP 32
struct Vis2Ovf { int n; int pad; };
__device__ __forceinline__ long long vis_live(long long n, const int *n_dev) {
  if (!n_dev) return n;
  const long long m = (long long)__ldg(n_dev);
  return m < n ? m : n;
}

// Grid over the previous call's box: a few dozen flops, so every CTA of k_vis2_fill derives it for itself (thread 0,
// shared memory) and CTA 0 leaves the copy k_vis2_pack / k_vis2_cones read -- no launch of its own.
__device__ void vis2_make_grid(const VisPrev *prev, double range, long long ncell_max, VisGrid *G) {
  double ext[3];
  for (int a = 0; a < 3; ++a) {
    double lo = dec_double(~prev->mnc[a]), hi = dec_double(prev->mx[a]);
    if (hi > lo || a < 2) { lo -= range; hi += range; }
    if (!(hi >= lo)) { lo = 0.0; hi = 0.0; }
    G->lo[a] = lo;
    ext[a] = hi - lo;
    G->mn[a] = 0ULL; G->mx[a] = 0ULL;
  }
  double cs = range > 0.0 ? range : 1.0;
  for (;;) {
    long long tot = 1;
    for (int a = 0; a < 3; ++a) {
      double f = floor(ext[a] / cs) + 1.0;
      int nc = (f > 1048576.0) ? 1048576 : (int)f;
      if (nc < 1) nc = 1;
      G->nc[a] = nc;
      tot *= nc;
      if (tot > ncell_max) break;
    }
    if (tot <= ncell_max) { G->ncell = (int)tot; break; }
    cs *= 1.5;
  }
  G->cs = cs;
  G->inv_cs = 1.0 / cs;
}

__global__ void __launch_bounds__(256) k_vis2_fill(long long n, const int *n_dev, const double *x, const double *y, const double *z,
                                                   const VisPrev *prev, double range, long long ncell_max, VisGrid *Gout,
                                                   int *count, int *ids, Vis2Ovf *ovf, int2 *ovf_items, VisBuild *B, double *out,
                                                   long long out_len) {
  __shared__ unsigned long long smn[3][8], smx[3][8];
  __shared__ VisGrid sG;
  if (threadIdx.x == 0) {
    vis2_make_grid(prev, range, ncell_max, &sG);
    if (blockIdx.x == 0) *Gout = sG;
  }
  __syncthreads();
  const VisGrid *G = &sG;
  const long long gtid = (long long)blockIdx.x * blockDim.x + threadIdx.x, gsz = (long long)gridDim.x * blockDim.x;
  for (long long i = gtid; i < out_len; i += gsz) out[i] = 0.0;
  unsigned long long mn[3] = {0ULL, 0ULL, 0ULL}, mx[3] = {0ULL, 0ULL, 0ULL};
  n = vis_live(n, n_dev);   // slots past the device-resident count (domain decomposition) hold stale particles
  for (long long i = gtid; i < n; i += gsz) {
    const double xi = x[i], yi = y[i], zi = z[i];
    const unsigned long long e[3] = {enc_double(xi), enc_double(yi), enc_double(zi)};
#pragma unroll
    for (int a = 0; a < 3; ++a) { mn[a] = max(mn[a], ~e[a]); mx[a] = max(mx[a], e[a]); }
    const int cx = vis_coord(xi, G->lo[0], G->inv_cs, G->nc[0]);
    const int cy = vis_coord(yi, G->lo[1], G->inv_cs, G->nc[1]);
    const int cz = vis_coord(zi, G->lo[2], G->inv_cs, G->nc[2]);
    const int c = (cz * G->nc[1] + cy) * G->nc[0] + cx;
    const int p = atomicAdd(&count[c], 1);
    if (p < VIS2_CAP) ids[(long long)c * VIS2_CAP + p] = (int)i;
    else { const int q = atomicAdd(&ovf->n, 1); ovf_items[q] = make_int2(c, (int)i); }
  }
#pragma unroll
  for (int a = 0; a < 3; ++a)
    for (int o = 16; o > 0; o >>= 1) {
      mn[a] = max(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], o));
      mx[a] = max(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], o));
    }
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0)
    for (int a = 0; a < 3; ++a) { smn[a][w] = mn[a]; smx[a][w] = mx[a]; }
  __syncthreads();
  if (threadIdx.x < 3) {
    const int a = threadIdx.x;
    unsigned long long m0 = 0ULL, m1 = 0ULL;
    for (int k = 0; k < 8; ++k) { m0 = max(m0, smn[a][k]); m1 = max(m1, smx[a][k]); }
    atomicMax(&B->mnc[a], m0);
    atomicMax(&B->mx[a], m1);
  }
}

// first call of a swarm: no previous box yet -- the box of the current positions
__global__ void __launch_bounds__(256) k_vis2_bbox(long long n, const int *n_dev, const double *x, const double *y, const double *z, VisPrev *prev) {
  __shared__ unsigned long long smn[3][8], smx[3][8];
  const long long gtid = (long long)blockIdx.x * blockDim.x + threadIdx.x, gsz = (long long)gridDim.x * blockDim.x;
  unsigned long long mn[3] = {0ULL, 0ULL, 0ULL}, mx[3] = {0ULL, 0ULL, 0ULL};
  n = vis_live(n, n_dev);
  for (long long i = gtid; i < n; i += gsz) {
    const unsigned long long e[3] = {enc_double(x[i]), enc_double(y[i]), enc_double(z[i])};
#pragma unroll
    for (int a = 0; a < 3; ++a) { mn[a] = max(mn[a], ~e[a]); mx[a] = max(mx[a], e[a]); }
  }
#pragma unroll
  for (int a = 0; a < 3; ++a)
    for (int o = 16; o > 0; o >>= 1) {
      mn[a] = max(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], o));
      mx[a] = max(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], o));
    }
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0)
    for (int a = 0; a < 3; ++a) { smn[a][w] = mn[a]; smx[a][w] = mx[a]; }
  __syncthreads();
  if (threadIdx.x < 3) {
    const int a = threadIdx.x;
    unsigned long long m0 = 0ULL, m1 = 0ULL;
    for (int k = 0; k < 8; ++k) { m0 = max(m0, smn[a][k]); m1 = max(m1, smx[a][k]); }
    atomicMax(&prev->mnc[a], m0);
    atomicMax(&prev->mx[a], m1);
  }
}

// one warp per cell
__global__ void __launch_bounds__(256) k_vis2_pack(const VisGrid *G, const int *count, const int *ids, const double *x, const double *y,
                                                   const double *z, const double *radii, const int *species, const __grid_constant__ DevParams P,
                                                   const int *type_col, int n_type_col, double4 *rec, int *rcol, const int *image,
                                                   int *rimg, const VisBuild *B, VisPrev *prev) {
  const long long gw = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (gw == 0 && lane < 3) { prev->mnc[lane] = B->mnc[lane]; prev->mx[lane] = B->mx[lane]; }   // the box for the next call
  if (gw >= G->ncell) return;
  int m = count[gw];
  if (m == 0) return;
  if (m > VIS2_CAP) m = VIS2_CAP;
  const int id = (lane < m) ? ids[gw * VIS2_CAP + lane] : 0x7fffffff;
  int rank = 0;
#pragma unroll
  for (int q = 0; q < 32; ++q) rank += (__shfl_sync(0xffffffffu, id, q) < id) ? 1 : 0;
  if (lane < m) {
    const long long o = gw * VIS2_CAP + rank;
    rec[o] = make_double4(x[id], y[id], z[id], radii[id]);
    const int t = P.sp_type[species[id]];
    rcol[o] = (t >= 0 && t < n_type_col) ? type_col[t] : -1;
    if (image) rimg[o] = image[id];
  }
}

// `image` (swarm_bind_images, domain decomposition): the periodic image count of every slot along the decomposed axis.  The
// reference measures vision on UNFOLDED positions without a minimum image (subdivided_vision_cones.py:113-157), so two
// slots that are close in a rank's local frame see each other only when they sit in the same image.
template <int GS>
__global__ void __launch_bounds__(VISW_THREADS) k_vis2_cones(long long n, const int *n_dev, const int *image, const int *rimg,
                                                             const int *row_of, const VisGrid *G, const int *count,
                                                             const double4 *rec, const int *rcol, const Vis2Ovf *ovf, const int2 *ovf_items,
                                                             const double *x, const double *y, const double *z, const double *radii,
                                                             const int *species, const __grid_constant__ DevParams P, const int *type_col,
                                                             int n_type_col, int n_dims, const double *theta, const double *q0,
                                                             const double *q1, const double *q2, const double *q3, const double *dirs,
                                                             double range, double half_angle, int n_cones, int n_cols, double *out) {
  extern __shared__ double sacc[];   // [nacc][VISW_THREADS]
  const int lane = threadIdx.x & (GS - 1);
  long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) / GS;
  const bool in_range = i < vis_live(n, n_dev);
  if (!in_range) i = n - 1;
  const int row = in_range ? row_of[i] : -1;
  const int nacc = n_cones * n_cols;
  for (int a = 0; a < nacc; ++a) sacc[a * VISW_THREADS + threadIdx.x] = 0.0;
  const double xi = x[i], yi = y[i], zi = z[i];
  const int img_i = image ? image[i] : 0;
  double ex, ey, ez;
  if (dirs) { ex = dirs[i]; ey = dirs[n + i]; ez = dirs[2 * n + i]; }
  else if (n_dims == 2) { sincos(theta[i], &ey, &ex); ez = 0.0; }
  else {
    double a = q0[i], b = q1[i], c = q2[i], d = q3[i];
    ex = 2 * (b * d + a * c); ey = 2 * (c * d - a * b); ez = a * a - b * b - c * c + d * d;
  }
  {
    double en = sqrt(ex * ex + ey * ey + ez * ez);
    ex /= en; ey /= en; ez /= en;
  }
  const double cone_w = half_angle * 2.0 / (double)n_cones;
  __shared__ double rim[VIS_MAX_ACC + 1], crim[VIS_MAX_ACC + 1];
  for (int t = threadIdx.x; t <= n_cones; t += blockDim.x) {
    rim[t] = -half_angle + (double)t * half_angle * 2.0 / (double)n_cones;
    crim[t] = cos(rim[t]);
  }
  __syncthreads();
  const int cx = vis_coord(xi, G->lo[0], G->inv_cs, G->nc[0]);
  const int cy = vis_coord(yi, G->lo[1], G->inv_cs, G->nc[1]);
  const int cz = vis_coord(zi, G->lo[2], G->inv_cs, G->nc[2]);
  if (row >= 0) {
    for (int zz = max(cz - 1, 0); zz <= min(cz + 1, G->nc[2] - 1); ++zz)
      for (int yy = max(cy - 1, 0); yy <= min(cy + 1, G->nc[1] - 1); ++yy)
        for (int xx = max(cx - 1, 0); xx <= min(cx + 1, G->nc[0] - 1); ++xx) {
          const long long c = ((long long)zz * G->nc[1] + yy) * G->nc[0] + xx;
          int m = __ldg(&count[c]);
          if (m > VIS2_CAP) m = VIS2_CAP;
          for (int k = lane; k < m; k += GS) {
            const long long o = c * VIS2_CAP + k;
            const double2 r01 = __ldg(reinterpret_cast<const double2 *>(&rec[o]));
            const double2 r23 = __ldg(reinterpret_cast<const double2 *>(&rec[o]) + 1);
            int col = __ldg(&rcol[o]);
            if (image && __ldg(&rimg[o]) != img_i) col = -1;
            visw_candidate(r01.x, r01.y, r23.x, r23.y, col, xi, yi, zi, ex, ey, ez, range, half_angle, cone_w, n_cones, n_cols,
                           rim, crim, sacc);
          }
        }
    // particles beyond the capacity of their cell (dense clusters): one list, scanned by every observer that has such a
    // cell in its neighbourhood -- empty in all but pathological packings
    const int no = ovf->n;
    for (int k = lane; k < no; k += GS) {
      const int2 it = ovf_items[k];
      const int ox = it.x % G->nc[0], oy = (it.x / G->nc[0]) % G->nc[1], oz = it.x / (G->nc[0] * G->nc[1]);
      if (abs(ox - cx) > 1 || abs(oy - cy) > 1 || abs(oz - cz) > 1) continue;
      const int j = it.y;
      const int t = P.sp_type[species[j]];
      int col = (t >= 0 && t < n_type_col) ? type_col[t] : -1;
      if (image && image[j] != img_i) col = -1;
      visw_candidate(x[j], y[j], z[j], radii[j], col, xi, yi, zi, ex, ey, ez, range, half_angle, cone_w, n_cones, n_cols, rim, crim, sacc);
    }
  }
  __syncwarp();
  for (int a = 0; a < nacc; ++a) {
    double v = sacc[a * VISW_THREADS + threadIdx.x];
#pragma unroll
    for (int off = GS / 2; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    if (lane == 0 && row >= 0) out[(long long)row * nacc + a] = v;
  }
}

// k_vis2_cones with a CANDIDATE QUEUE (round 2).  In k_vis2_cones a warp walks the stencil cells of its 32 / GS observers
// in lock step and runs the whole candidate evaluation (square root, two divisions, the cone search) in every iteration
// in which ANY of its lanes holds a particle inside the range -- nearly always, although only one candidate in three is
// (pi r^2 of the 9 r^2 the stencil covers) and a cell of ~3 particles fills 3 of an observer's 4 lanes.  Here the walk
// only SCREENS: squared distance against range^2 (1 + 1e-9) -- a superset of the exact test, which visw_candidate
// repeats -- and appends the survivors to the observer's queue in shared memory (position by warp ballot: the order,
// and with it the summation order, is fixed).  The expensive evaluation then runs on full queues: every pass gives each
// lane of a group one candidate, ~3 passes per warp instead of ~9.  Same cones and weights as k_vis2_cones, candidate by
// candidate; the sums associate differently (<= 1 ulp per term).
#define VISQ_CAP 32
template <int GS>
__global__ void __launch_bounds__(VISW_THREADS) k_vis2_cones_q(long long n, const int *n_dev, const int *image, const int *rimg,
                                                               const int *row_of, const VisGrid *G, const int *count,
                                                               const double4 *rec, const int *rcol, const Vis2Ovf *ovf, const int2 *ovf_items,
                                                               const double *x, const double *y, const double *z, const double *radii,
                                                               const int *species, const __grid_constant__ DevParams P, const int *type_col,
                                                               int n_type_col, int n_dims, const double *theta, const double *q0,
                                                               const double *q1, const double *q2, const double *q3, const double *dirs,
                                                               double range, double half_angle, int n_cones, int n_cols, double *out) {
  extern __shared__ double sacc[];   // [nacc][VISW_THREADS], then the queues [VISW_THREADS / GS][VISQ_CAP]
  const int nacc = n_cones * n_cols;
  int *qu = reinterpret_cast<int *>(sacc + (size_t)nacc * VISW_THREADS) + (threadIdx.x / GS) * VISQ_CAP;
  const int lane = threadIdx.x & (GS - 1);
  const int gshift = (threadIdx.x & 31) & ~(GS - 1);   // first lane of this observer's group within the warp
  long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) / GS;
  const bool in_range = i < vis_live(n, n_dev);
  if (!in_range) i = n - 1;
  const int row = in_range ? row_of[i] : -1;
  for (int a = 0; a < nacc; ++a) sacc[a * VISW_THREADS + threadIdx.x] = 0.0;
  const double xi = x[i], yi = y[i], zi = z[i];
  const int img_i = image ? image[i] : 0;
  double ex, ey, ez;
  if (dirs) { ex = dirs[i]; ey = dirs[n + i]; ez = dirs[2 * n + i]; }
  else if (n_dims == 2) { sincos(theta[i], &ey, &ex); ez = 0.0; }
  else {
    double a = q0[i], b = q1[i], c = q2[i], d = q3[i];
    ex = 2 * (b * d + a * c); ey = 2 * (c * d - a * b); ez = a * a - b * b - c * c + d * d;
  }
  {
    double en = sqrt(ex * ex + ey * ey + ez * ez);
    ex /= en; ey /= en; ez /= en;
  }
  const double cone_w = half_angle * 2.0 / (double)n_cones;
  __shared__ double rim[VIS_MAX_ACC + 1], crim[VIS_MAX_ACC + 1];
  for (int t = threadIdx.x; t <= n_cones; t += blockDim.x) {
    rim[t] = -half_angle + (double)t * half_angle * 2.0 / (double)n_cones;
    crim[t] = cos(rim[t]);
  }
  __syncthreads();
  const int ncx = G->nc[0], ncy = G->nc[1], ncz = G->nc[2];
  const int cx = vis_coord(xi, G->lo[0], G->inv_cs, ncx);
  const int cy = vis_coord(yi, G->lo[1], G->inv_cs, ncy);
  const int cz = vis_coord(zi, G->lo[2], G->inv_cs, ncz);
  const double r2max = range * range * (1.0 + 1e-9);
  int qn = 0;   // entries in this group's queue: the same number in its GS lanes
  // one pass: every group gives each of its lanes one queued candidate (fewer when it holds fewer)
  auto pass = [&]() {
    const int take = qn < GS ? qn : GS;
    if (lane < take) {
      const int en = qu[qn - take + lane];
      if (en >= 0) {
        const double2 r01 = __ldg(reinterpret_cast<const double2 *>(&rec[en]));
        const double2 r23 = __ldg(reinterpret_cast<const double2 *>(&rec[en]) + 1);
        visw_candidate(r01.x, r01.y, r23.x, r23.y, __ldg(&rcol[en]), xi, yi, zi, ex, ey, ez, range, half_angle, cone_w, n_cones, n_cols,
                       rim, crim, sacc);
      } else {
        const int j = ~en;
        const int t = P.sp_type[species[j]];
        const int col = (t >= 0 && t < n_type_col) ? type_col[t] : -1;
        visw_candidate(x[j], y[j], z[j], radii[j], col, xi, yi, zi, ex, ey, ez, range, half_angle, cone_w, n_cones, n_cols, rim, crim, sacc);
      }
    }
    qn -= take;
  };
  auto push = [&](bool ok, int en) {
    const unsigned int b = (__ballot_sync(0xffffffffu, ok) >> gshift) & ((1u << GS) - 1u);
    if (ok) qu[qn + __popc(b & ((1u << lane) - 1u))] = en;
    qn += __popc(b);
    __syncwarp();
    if (__any_sync(0xffffffffu, qn > VISQ_CAP - GS)) { pass(); __syncwarp(); }   // a queue could not take another round: one pass for all
  };
  // the stencil walk is warp-uniform (27 offsets, each cell's count padded to the warp's maximum): the ballot needs all lanes
  for (int dz = (ncz > 1 ? -1 : 0); dz <= (ncz > 1 ? 1 : 0); ++dz)
    for (int dy = -1; dy <= 1; ++dy)
      for (int dx = -1; dx <= 1; ++dx) {
        const int xx = cx + dx, yy = cy + dy, zz = cz + dz;
        const bool valid = row >= 0 && xx >= 0 && xx < ncx && yy >= 0 && yy < ncy && zz >= 0 && zz < ncz;
        const long long c = valid ? ((long long)zz * ncy + yy) * ncx + xx : 0;
        int m = valid ? __ldg(&count[c]) : 0;
        if (m > VIS2_CAP) m = VIS2_CAP;
        const int mw = (int)__reduce_max_sync(0xffffffffu, (unsigned int)m);
        for (int k0 = 0; k0 < mw; k0 += GS) {
          const int k = k0 + lane;
          bool ok = false;
          const long long o = c * VIS2_CAP + k;
          if (k < m) {
            const double2 r01 = __ldg(reinterpret_cast<const double2 *>(&rec[o]));
            const double rz = __ldg(reinterpret_cast<const double *>(&rec[o]) + 2);
            const double ddx = r01.x - xi, ddy = r01.y - yi, ddz = rz - zi;
            const double d2 = ddx * ddx + ddy * ddy + ddz * ddz;
            ok = d2 > 0.0 && d2 <= r2max && __ldg(&rcol[o]) >= 0;
            if (ok && image && __ldg(&rimg[o]) != img_i) ok = false;
          }
          push(ok, (int)o);
        }
      }
  // particles beyond the capacity of their cell (dense clusters): one list, empty in all but pathological packings
  const int no = ovf->n;
  for (int k0 = 0; k0 < no; k0 += GS) {
    const int k = k0 + lane;
    bool ok = false;
    int j = 0;
    if (k < no && row >= 0) {
      const int2 it = ovf_items[k];
      const int ox = it.x % ncx, oy = (it.x / ncx) % ncy, oz = it.x / (ncx * ncy);
      j = it.y;
      if (abs(ox - cx) <= 1 && abs(oy - cy) <= 1 && abs(oz - cz) <= 1 && !(image && image[j] != img_i)) {
        const double ddx = x[j] - xi, ddy = y[j] - yi, ddz = z[j] - zi;
        const double d2 = ddx * ddx + ddy * ddy + ddz * ddz;
        ok = d2 > 0.0 && d2 <= r2max;
      }
    }
    push(ok, ~j);
  }
  while (__any_sync(0xffffffffu, qn > 0)) pass();
  __syncwarp();
  for (int a = 0; a < nacc; ++a) {
    double v = sacc[a * VISW_THREADS + threadIdx.x];
#pragma unroll
    for (int off = GS / 2; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    if (lane == 0 && row >= 0) out[(long long)row * nacc + a] = v;
  }
}

// ---- concentration field / gradient sensing ---------------------------------------------------
__device__ __forceinline__ double decay_eval(int kind, double c0, double c1, double d) {
  switch (kind) {
    case SWARM_DECAY_LINEAR: return c0 + c1 * d;
    case SWARM_DECAY_EXP: return c0 * exp(-c1 * d);
    case SWARM_DECAY_INVERSE: return c0 / (d + c1);
    default: return d;
  }
}

// `partial` != nullptr: the launch also leaves (sum, mean over n_rows) of the written rewards in out2 -- every CTA adds
// its 256 values in a fixed tree, the CTA that draws the last ticket adds the per-CTA sums in a fixed order (the
// reference takes the mean on the host, trainers/trainer.py:92-95; here reward and mean are one launch).
__global__ void __launch_bounds__(256) k_obs_field(long long n, const double *x, const double *y, const double *z, const int *row_of,
                                                   double s0, double s1, double s2, double b0, double b1, double b2, double *hist,
                                                   int decay, double c0, double c1, double scale, int clip, double *out,
                                                   double *d_now, double *d_prev, double *partial, unsigned int *ticket,
                                                   double *out2, long long n_rows) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int row = (i < n) ? (row_of ? row_of[i] : (int)i) : -1;
  double v = 0.0;
  if (row >= 0) {
    double p0 = x[i] / b0, p1 = y[i] / b1, p2 = z[i] / b2;
    double h0 = hist[i], h1 = hist[n + i], h2 = hist[2 * n + i];
    double dn = sqrt((s0 - p0) * (s0 - p0) + (s1 - p1) * (s1 - p1) + (s2 - p2) * (s2 - p2));
    double dp = sqrt((s0 - h0) * (s0 - h0) + (s1 - h1) * (s1 - h1) + (s2 - h2) * (s2 - h2));
    hist[i] = p0; hist[n + i] = p1; hist[2 * n + i] = p2;
    if (d_now) d_now[row] = dn;
    if (d_prev) d_prev[row] = dp;
    if (out && decay != SWARM_DECAY_NONE) {
      v = scale * (decay_eval(decay, c0, c1, dn) - decay_eval(decay, c0, c1, dp));
      if (clip && !(v > 0.0)) v = 0.0;
      out[row] = v;
    }
  }
  if (!partial) return;
  __shared__ double ws[8];
  __shared__ bool last;
  double sum = v;
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = sum;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int k = 0; k < 8; ++k) t += ws[k];
    partial[blockIdx.x] = t;
    __threadfence();
    last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  double t = 0.0;
  for (long long k = threadIdx.x; k < (long long)gridDim.x; k += 256) t += __ldcg(&partial[k]);
  for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = t;
  __syncthreads();
  if (threadIdx.x == 0) {
    double tot = 0.0;
    for (int k = 0; k < 8; ++k) tot += ws[k];
    out2[0] = tot;
    out2[1] = (n_rows > 0) ? tot / (double)n_rows : 0.0;
    *ticket = 0u;
  }
}

__global__ void k_obs_position(long long n, const double *x, const double *y, const double *z, const int *row_of,
                               double b0, double b1, double b2, double *out) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int row = row_of ? row_of[i] : (int)i;
  if (row < 0) return;
  out[3LL * row] = x[i] / b0; out[3LL * row + 1] = y[i] / b1; out[3LL * row + 2] = z[i] / b2;
}

// ---- pairwise field: ParticleSensing / SpeciesSearch ------------------------------------------------
// field_i = sum of decay(|(x_j - x_i) / box|) over the first M-1 sensed particles at NON-ZERO distance, in
// ascending index, padded with sensed[0] when fewer exist -- the reference's nonzero(size=M-1)
// (observables/particle_sensing.py:124-131, tasks/searching/species_search.py:126-133).  All pairs, no
// cut-off (the decay function has none): sensed positions are staged through shared memory tile by tile,
// one thread per particle, fixed summation order.
#define PF_TILE 256
__global__ void __launch_bounds__(PF_TILE) k_pair_field(long long n, const double *x, const double *y, const double *z,
                                                        const int *row_of, const int *sensed, long long M, double ib0, double ib1,
                                                        double ib2, int decay, double c0, double c1, double scale, int clip_mode,
                                                        int init_only, double *hist, double *out, double *field) {
  __shared__ double sx[PF_TILE], sy[PF_TILE], sz[PF_TILE];
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int row = (i < n) ? (row_of ? row_of[i] : (int)i) : -1;
  const bool act = row >= 0;
  double xi = 0.0, yi = 0.0, zi = 0.0;
  if (act) { xi = x[i]; yi = y[i]; zi = z[i]; }
  double total = 0.0, pend = 0.0, v0 = 0.0;
  long long nnz = 0;
  bool have = false;
  for (long long t0 = 0; t0 < M; t0 += PF_TILE) {
    const long long j = t0 + threadIdx.x;
    __syncthreads();
    if (j < M) { const int p = sensed[j]; sx[threadIdx.x] = x[p]; sy[threadIdx.x] = y[p]; sz[threadIdx.x] = z[p]; }
    __syncthreads();
    if (!act) continue;
    const int m = (int)((M - t0 < PF_TILE) ? (M - t0) : PF_TILE);
    for (int k = 0; k < m; ++k) {
      const double dx = (sx[k] - xi) * ib0, dy = (sy[k] - yi) * ib1, dz = (sz[k] - zi) * ib2;
      const double d = sqrt(dx * dx + dy * dy + dz * dz);
      if (t0 == 0 && k == 0) v0 = decay_eval(decay, c0, c1, d);
      if (d != 0.0) {
        if (have) total += pend;
        pend = decay_eval(decay, c0, c1, d);
        have = true;
        ++nnz;
      }
    }
  }
  if (!act) return;
  if (M <= 1) total = 0.0;
  else {
    if (have && nnz <= M - 1) total += pend;          // nnz == M: the last sensed particle is dropped
    for (long long q = nnz; q < M - 1; ++q) total += v0;  // padding re-reads sensed[0]
  }
  const double h = hist[row];
  hist[row] = total;
  if (field) field[row] = total;
  if (init_only || !out) return;
  double v = total - h;
  if (clip_mode > 0 && !(v > 0.0)) v = 0.0;
  if (clip_mode < 0 && !(v < 0.0)) v = 0.0;
  out[row] = scale * v;
}

// ---- deterministic two-stage sum -------------------------------------------------------------
#define RED_BLOCKS 296
// One launch: every CTA writes its slice's sum, takes a ticket, and the CTA that draws the last ticket adds the
// partials in a fixed order (what k_reduce_final did in a second launch) and re-arms the ticket.
__global__ void k_reduce_mean(long long n, const double *in, double *partial, unsigned int *ticket, double *out2) {
  __shared__ double ws[8];
  __shared__ bool last;
  long long per = (n + RED_BLOCKS - 1) / RED_BLOCKS;
  long long lo = (long long)blockIdx.x * per, hi = min(n, lo + per);
  double s = 0.0;
  for (long long i = lo + threadIdx.x; i < hi; i += blockDim.x) s += in[i];
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int k = 0; k < (int)(blockDim.x >> 5); ++k) t += ws[k];
    partial[blockIdx.x] = t;
    __threadfence();
    last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!last || threadIdx.x >= 32) return;
  __threadfence();
  double t = 0.0;
  for (int k = threadIdx.x; k < RED_BLOCKS; k += 32) t += __ldcg(&partial[k]);
  for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  if (threadIdx.x == 0) {
    out2[0] = t;
    out2[1] = (n > 0) ? t / (double)n : 0.0;
    *ticket = 0u;
  }
}

__global__ void k_reduce_partial(long long n, const double *in, double *partial) {
  __shared__ double ws[8];
  // fixed, grid-independent assignment: block b owns a contiguous slice
  long long per = (n + RED_BLOCKS - 1) / RED_BLOCKS;
  long long lo = (long long)blockIdx.x * per, hi = min(n, lo + per);
  double s = 0.0;
  for (long long i = lo + threadIdx.x; i < hi; i += blockDim.x) s += in[i];
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int k = 0; k < (int)(blockDim.x >> 5); ++k) t += ws[k];
    partial[blockIdx.x] = t;
  }
}
__global__ void k_reduce_final(long long n, const double *partial, double *out2) {
  // one warp, fixed order: lane l adds partial[l], partial[l + 32], ... then a fixed shuffle tree (deterministic;
  // the former single-thread loop over 296 dependent DADDs took 10 us)
  if (blockIdx.x != 0 || threadIdx.x >= 32) return;
  double t = 0.0;
  for (int k = threadIdx.x; k < RED_BLOCKS; k += 32) t += partial[k];
  for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  if (threadIdx.x == 0) {
    out2[0] = t;
    out2[1] = (n > 0) ? t / (double)n : 0.0;
  }
}

// ---------------------------------------------------------------------------------------------
// C-ABI
// ---------------------------------------------------------------------------------------------
static inline size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

static int sw_obs_field_launch(swarm_t *s, const int32_t *row_of, const double src[3], const double box_length[3], double *hist,
                               int32_t decay, double c0, double c1, double scale, int32_t clip, double *out, double *d_now,
                               double *d_prev, double *out2, int64_t n_rows, void *stream) {
  if (!s || !s->state_bound || !src || !box_length || !hist) { sw_set_error("bad arguments"); return SWARM_ERR_INVALID; }
  if (decay < SWARM_DECAY_NONE || decay > SWARM_DECAY_INVERSE) { sw_set_error("unknown decay kind"); return SWARM_ERR_INVALID; }
  SW_CUDA(cudaSetDevice(s->cfg.device));
  const DevArrays &A = s->A;
  if (A.n == 0) return SWARM_OK;
  const int blocks = sw_blocks(A.n, 256);
  double *partial = nullptr;
  unsigned int *ticket = nullptr;
  if (out2) {
    if (!out || decay == SWARM_DECAY_NONE) { sw_set_error("the mean needs a closed-form decay and a reward array"); return SWARM_ERR_INVALID; }
    if (s->field_scratch_blocks < blocks) {
      if (s->field_scratch) cudaFree(s->field_scratch);
      s->field_scratch = nullptr; s->field_scratch_blocks = 0;
      SW_CUDA(cudaMalloc((void **)&s->field_scratch, sizeof(double) * ((size_t)blocks + 1)));
      SW_CUDA(cudaMemset(s->field_scratch, 0, sizeof(double) * ((size_t)blocks + 1)));   // [blocks]: the ticket
      s->field_scratch_blocks = blocks;
    }
    partial = s->field_scratch;
    ticket = (unsigned int *)(s->field_scratch + s->field_scratch_blocks);
  }
  k_obs_field<<<blocks, 256, 0, (cudaStream_t)stream>>>(A.n, A.x[0], A.y[0], A.z[0], row_of, src[0], src[1], src[2], box_length[0],
                                                        box_length[1], box_length[2], hist, decay, c0, c1, scale, clip, out, d_now,
                                                        d_prev, partial, ticket, out2, (long long)n_rows);
  s->launches += 1;
  SW_CUDA(cudaGetLastError());
  return SWARM_OK;
}

extern "C" int swarm_obs_field(swarm_t *s, const int32_t *row_of, const double src[3], const double box_length[3],
                               double *hist, int32_t decay, double c0, double c1, double scale, int32_t clip,
                               double *out, double *d_now, double *d_prev, void *stream) {
  return sw_obs_field_launch(s, row_of, src, box_length, hist, decay, c0, c1, scale, clip, out, d_now, d_prev, nullptr, 0, stream);
}

extern "C" int swarm_obs_field_mean(swarm_t *s, const int32_t *row_of, int64_t n_rows, const double src[3],
                                    const double box_length[3], double *hist, int32_t decay, double c0, double c1, double scale,
                                    int32_t clip, double *out, double *out2, void *stream) {
  if (!out2 || n_rows < 0) { sw_set_error("bad arguments"); return SWARM_ERR_INVALID; }
  return sw_obs_field_launch(s, row_of, src, box_length, hist, decay, c0, c1, scale, clip, out, nullptr, nullptr, out2, n_rows, stream);
}

// ---- swarm_obs_field_mean without any allocation ----------------------------------------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops).  The per-CTA reward sums and the ticket of k_obs_field live in a file-scope
// device array instead of swarm_t::field_scratch: SW_NOALLOC_MAX_BLOCKS CTAs of 256 particles = 8 388 608 particles,
// 256 KiB + one 256-byte slot for the ticket, held by every GPU that loads the library until the process exits.
// Host code must not pass the symbol itself to a kernel (in host code it is a stand-in, not a device address): its
// address is resolved once per device with cudaGetSymbolAddress, that device current.  All calls share the array -- two
// calls whose kernels overlap on one GPU (different streams) race on the sums and on the ticket.
#define SW_NOALLOC_MAX_BLOCKS 32768
#define SW_NOALLOC_MAX_DEVICES 64
static __device__ __align__(256) double g_noalloc_field_scratch[SW_NOALLOC_MAX_BLOCKS + 32];   // [MAX_BLOCKS]: the ticket
static double *g_noalloc_field_ptr[SW_NOALLOC_MAX_DEVICES];   // device address of the array, per device (nullptr: not resolved yet)
static std::mutex g_noalloc_mutex;

extern "C" int swarm_obs_field_mean_noalloc(swarm_t *s, const int32_t *row_of, int64_t n_rows, const double src[3],
                                            const double box_length[3], double *hist, int32_t decay, double c0, double c1,
                                            double scale, int32_t clip, double *out, double *out2, void *stream) {
  if (!s || !s->state_bound || !src || !box_length || !hist || !out || !out2 || n_rows < 0) { sw_set_error("bad arguments"); return SWARM_ERR_INVALID; }
  if (decay <= SWARM_DECAY_NONE || decay > SWARM_DECAY_INVERSE) { sw_set_error("the mean needs a closed-form decay"); return SWARM_ERR_INVALID; }
  const DevArrays &A = s->A;
  if (A.n > (long long)SW_NOALLOC_MAX_BLOCKS * 256) {
    sw_set_error("swarm_obs_field_mean_noalloc: more than 8388608 particles (the static scratch holds 32768 CTA sums); use swarm_obs_field_mean");
    return SWARM_ERR_INVALID;
  }
  const int dev = s->cfg.device;
  if (dev < 0 || dev >= SW_NOALLOC_MAX_DEVICES) { sw_set_error("swarm_obs_field_mean_noalloc: device ordinal out of range"); return SWARM_ERR_INVALID; }
  SW_CUDA(cudaSetDevice(dev));
  double *scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_field_ptr[dev]) {
      double *p = nullptr;
      SW_CUDA(cudaGetSymbolAddress((void **)&p, g_noalloc_field_scratch));   // the address on the CURRENT device
      g_noalloc_field_ptr[dev] = p;
    }
    scratch = g_noalloc_field_ptr[dev];
  }
  if (A.n == 0) return SWARM_OK;
  const int blocks = sw_blocks(A.n, 256);
  unsigned int *ticket = (unsigned int *)(scratch + SW_NOALLOC_MAX_BLOCKS);
  // the ticket must start at zero whatever an earlier (possibly failed) call left; the sums are overwritten before they are read
  SW_CUDA(cudaMemsetAsync(ticket, 0, sizeof(double), (cudaStream_t)stream));
  k_obs_field<<<blocks, 256, 0, (cudaStream_t)stream>>>(A.n, A.x[0], A.y[0], A.z[0], row_of, src[0], src[1], src[2], box_length[0],
                                                        box_length[1], box_length[2], hist, decay, c0, c1, scale, clip, out, nullptr,
                                                        nullptr, scratch, ticket, out2, (long long)n_rows);
  s->launches += 1;
  SW_CUDA(cudaGetLastError());
  return SWARM_OK;
}

extern "C" int swarm_obs_pair_field(swarm_t *s, const int32_t *row_of, const int32_t *sensed_idx, int64_t n_sensed,
                                    const double box_length[3], int32_t decay, double c0, double c1, double scale,
                                    int32_t clip_mode, int32_t init_only, double *hist_rows, double *out, double *field,
                                    void *stream) {
  if (!s || !s->state_bound || !box_length || !hist_rows || n_sensed < 0 || (n_sensed > 0 && !sensed_idx)) {
    sw_set_error("bad arguments"); return SWARM_ERR_INVALID;
  }
  if (decay < SWARM_DECAY_LINEAR || decay > SWARM_DECAY_INVERSE) { sw_set_error("unknown decay kind"); return SWARM_ERR_INVALID; }
  SW_CUDA(cudaSetDevice(s->cfg.device));
  const DevArrays &A = s->A;
  if (A.n == 0) return SWARM_OK;
  k_pair_field<<<sw_blocks(A.n, PF_TILE), PF_TILE, 0, (cudaStream_t)stream>>>(
      A.n, A.x[0], A.y[0], A.z[0], row_of, sensed_idx, n_sensed, 1.0 / box_length[0], 1.0 / box_length[1], 1.0 / box_length[2],
      decay, c0, c1, scale, clip_mode, init_only, hist_rows, out, field);
  s->launches += 1;
  SW_CUDA(cudaGetLastError());
  return SWARM_OK;
}

extern "C" int swarm_obs_position(swarm_t *s, const int32_t *row_of, const double box_length[3], double *out, void *stream) {
  if (!s || !s->state_bound || !box_length || !out) { sw_set_error("bad arguments"); return SWARM_ERR_INVALID; }
  SW_CUDA(cudaSetDevice(s->cfg.device));
  const DevArrays &A = s->A;
  if (A.n == 0) return SWARM_OK;
  k_obs_position<<<sw_blocks(A.n, 256), 256, 0, (cudaStream_t)stream>>>(A.n, A.x[0], A.y[0], A.z[0], row_of, box_length[0],
                                                                        box_length[1], box_length[2], out);
  s->launches += 1;
  SW_CUDA(cudaGetLastError());
  return SWARM_OK;
}

extern "C" int swarm_reduce_mean(swarm_t *s, const double *in, int64_t n, double *out2, void *stream) {
  if (!s || n < 0 || !out2 || (n > 0 && !in)) { sw_set_error("bad arguments"); return SWARM_ERR_INVALID; }
  SW_CUDA(cudaSetDevice(s->cfg.device));
  if (!s->red_scratch) {
    SW_CUDA(cudaMalloc((void **)&s->red_scratch, sizeof(double) * (RED_BLOCKS + 1)));
    SW_CUDA(cudaMemset(s->red_scratch, 0, sizeof(double) * (RED_BLOCKS + 1)));   // [RED_BLOCKS]: the ticket
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (getenv("SWARM_B200_REDUCE_TWO_PASS")) {
    k_reduce_partial<<<RED_BLOCKS, 256, 0, st>>>(n, in, s->red_scratch);
    k_reduce_final<<<1, 32, 0, st>>>(n, s->red_scratch, out2);
    s->launches += 2;
  } else {
    k_reduce_mean<<<RED_BLOCKS, 256, 0, st>>>(n, in, s->red_scratch, (unsigned int *)(s->red_scratch + RED_BLOCKS), out2);
    s->launches += 1;
  }
  SW_CUDA(cudaGetLastError());
  return SWARM_OK;
}

extern "C" int swarm_bind_images(swarm_t *s, const int32_t *image_dev) {
  if (!s) { sw_set_error("bad arguments"); return SWARM_ERR_INVALID; }
  s->vis_image = (const int *)image_dev;
  return SWARM_OK;
}

extern "C" int swarm_obs_vision(swarm_t *s, const int32_t *row_of, int64_t rows, double vision_range, double half_angle,
                                int32_t n_cones, const double *radii, const int32_t *type_col, int32_t n_type_col,
                                int32_t n_cols, const double *dirs, double *out, void *stream) {
  if (!s || !s->state_bound || !row_of || !radii || !type_col || !out) { sw_set_error("bad arguments"); return SWARM_ERR_INVALID; }
  if (n_cones < 1 || n_cols < 1 || n_cones * n_cols > VIS_MAX_ACC) {
    sw_set_error("n_cones * n_detected_types must be in 1..64"); return SWARM_ERR_INVALID;
  }
  if (!(vision_range > 0.0)) { sw_set_error("vision_range must be positive"); return SWARM_ERR_INVALID; }
  SW_CUDA(cudaSetDevice(s->cfg.device));
  const DevArrays &A = s->A;
  const long long n = A.n;
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0 || rows == 0) return SWARM_OK;
  long long ncell_max = 4 * n;
  if (ncell_max < 4096) ncell_max = 4096;
  if (ncell_max > (1LL << 24)) ncell_max = (1LL << 24);
  const long long m = ncell_max + 1;
  const int ntiles = (int)((m + SCAN_TILE - 1) / SCAN_TILE);
  // carve the scratch buffer
  size_t off = 0;
  auto carve = [&](size_t bytes) { size_t o = off; off += align256(bytes); return o; };
  size_t o_grid = carve(sizeof(VisGrid));
  size_t o_prev = carve(sizeof(VisPrev));
  size_t o_build = carve(sizeof(VisBuild));   // directly in front of count[]: one memset zeroes both
  size_t o_count = carve(sizeof(int) * (size_t)m), o_start = carve(sizeof(int) * (size_t)m), o_fill = carve(sizeof(int) * (size_t)m);
  size_t o_tiles = carve(sizeof(int) * (size_t)ntiles);
  size_t o_cell = carve(sizeof(int) * (size_t)n), o_items = carve(sizeof(int) * (size_t)n), o_col = carve(sizeof(int) * (size_t)n);
  size_t o_sx = carve(sizeof(double) * (size_t)n), o_sy = carve(sizeof(double) * (size_t)n);
  size_t o_sz = carve(sizeof(double) * (size_t)n), o_sr = carve(sizeof(double) * (size_t)n);
  // fixed-capacity layout (k_vis2_*): at most n / 2 cells of VIS2_CAP slots
  long long ncell2 = n / 2;
  if (ncell2 < 4096) ncell2 = 4096;
  if (ncell2 > (1LL << 20)) ncell2 = (1LL << 20);
  size_t o_ids2 = carve(sizeof(int) * (size_t)ncell2 * VIS2_CAP), o_rcol2 = carve(sizeof(int) * (size_t)ncell2 * VIS2_CAP);
  size_t o_rimg2 = carve(s->vis_image ? sizeof(int) * (size_t)ncell2 * VIS2_CAP : 0);
  size_t o_rec2 = carve(sizeof(double4) * (size_t)ncell2 * VIS2_CAP), o_ovf = carve(sizeof(Vis2Ovf)), o_ovfi = carve(sizeof(int2) * (size_t)n);
  if (s->vis_bytes < off) {
    if (s->vis_buf) cudaFree(s->vis_buf);
    s->vis_buf = nullptr; s->vis_bytes = 0;
    SW_CUDA(cudaMalloc(&s->vis_buf, off));
    s->vis_bytes = off;
    s->vis_prev_n = -1;
  }
  char *base = (char *)s->vis_buf;
  VisGrid *G = (VisGrid *)(base + o_grid);
  int *count = (int *)(base + o_count), *start = (int *)(base + o_start), *fill = (int *)(base + o_fill);
  int *tiles = (int *)(base + o_tiles), *cellid = (int *)(base + o_cell), *items = (int *)(base + o_items);
  int *scol = (int *)(base + o_col);
  double *sx = (double *)(base + o_sx), *sy = (double *)(base + o_sy), *sz = (double *)(base + o_sz), *sr = (double *)(base + o_sr);
  const int T = 256, B = sw_blocks(n, T);
  const int gb = B < 4 * s->num_sms ? B : 4 * s->num_sms;
  const long long out_len = (long long)rows * n_cones * n_cols;
  const bool dd_view = A.n_dev || s->vis_image;   // a device-resident count / image counts: only the fixed-capacity kernels honour them
  if (dd_view && n_cones * n_cols > VISW_MAX_ACC) {
    sw_set_error("vision cones on a domain-decomposed swarm: n_cones * n_detected_types must be <= 32"); return SWARM_ERR_INVALID;
  }
  if (n_cones * n_cols <= VISW_MAX_ACC && (dd_view || (!getenv("SWARM_B200_VIS_V1") && !getenv("SWARM_B200_VIS_SERIAL")))) {
    // ---- fixed-capacity layout on the previous call's bounding box: memset, fill, pack, cones -------------------
    VisBuild *VB = (VisBuild *)(base + o_build);
    VisPrev *VP = (VisPrev *)(base + o_prev);
    if (!(s->vis_prev_n == n && s->vis_prev_range == vision_range)) {
      SW_CUDA(cudaMemsetAsync(VP, 0, sizeof(VisPrev), st));
      k_vis2_bbox<<<gb > 0 ? gb : 1, T, 0, st>>>(n, A.n_dev, A.x[0], A.y[0], A.z[0], VP);
      s->launches += 1;
      s->vis_prev_n = n; s->vis_prev_range = vision_range;
    }
    int *ids2 = (int *)(base + o_ids2), *rcol2 = (int *)(base + o_rcol2), *rimg2 = s->vis_image ? (int *)(base + o_rimg2) : nullptr;
    double4 *rec2 = (double4 *)(base + o_rec2);
    Vis2Ovf *ovf = (Vis2Ovf *)(base + o_build + 64);   // inside VisBuild's 256-byte slot
    static_assert(sizeof(VisBuild) <= 64, "overflow counter sits behind VisBuild");
    int2 *ovfi = (int2 *)(base + o_ovfi);
    // one memset: this call's box accumulator, the overflow counter (both in the slot in front of count[]) and the cell counts
    SW_CUDA(cudaMemsetAsync(VB, 0, (size_t)((char *)(count + ncell2 + 1) - (char *)VB), st));
    k_vis2_fill<<<gb > 0 ? gb : 1, T, 0, st>>>(n, A.n_dev, A.x[0], A.y[0], A.z[0], VP, vision_range, ncell2, G, count, ids2, ovf, ovfi, VB, out, out_len);
    k_vis2_pack<<<sw_blocks(ncell2 * 32, T), T, 0, st>>>(G, count, ids2, A.x[0], A.y[0], A.z[0], radii, A.species, s->P, type_col, n_type_col,
                                                        rec2, rcol2, s->vis_image, rimg2, VB, VP);
    const size_t smem = sizeof(double) * (size_t)(n_cones * n_cols) * VISW_THREADS;
    if (getenv("SWARM_B200_VIS_NOQUEUE"))
      k_vis2_cones<4><<<sw_blocks(n * 4, VISW_THREADS), VISW_THREADS, smem, st>>>(
          n, A.n_dev, s->vis_image, rimg2, row_of, G, count, rec2, rcol2, ovf, ovfi, A.x[0], A.y[0], A.z[0], radii, A.species, s->P, type_col, n_type_col, s->P.n_dims,
          A.theta, A.q0, A.q1, A.q2, A.q3, dirs, vision_range, half_angle, n_cones, n_cols, out);
    else
      k_vis2_cones_q<4><<<sw_blocks(n * 4, VISW_THREADS), VISW_THREADS, smem + sizeof(int) * (VISW_THREADS / 4) * VISQ_CAP, st>>>(
          n, A.n_dev, s->vis_image, rimg2, row_of, G, count, rec2, rcol2, ovf, ovfi, A.x[0], A.y[0], A.z[0], radii, A.species, s->P, type_col, n_type_col, s->P.n_dims,
          A.theta, A.q0, A.q1, A.q2, A.q3, dirs, vision_range, half_angle, n_cones, n_cols, out);
    s->launches += 3;
    SW_CUDA(cudaGetLastError());
    return SWARM_OK;
  }
  bool fused = s->persistent && !getenv("SWARM_B200_VIS_UNFUSED");
  int bpsm = 0;
  if (fused) {
    SW_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bpsm, k_vis_build, VISB_T, 0));
    fused = bpsm >= 1;
  }
  if (fused) {
    VisBuild *VB = (VisBuild *)(base + o_build);
    VisPrev *VP = (VisPrev *)(base + o_prev);
    int use_prev = (s->vis_prev_n == n && s->vis_prev_range == vision_range && !getenv("SWARM_B200_VIS_NO_PREV")) ? 1 : 0;
    s->vis_prev_n = n; s->vis_prev_range = vision_range;
    SW_CUDA(cudaMemsetAsync(VB, 0, (size_t)((char *)(count + m) - (char *)VB), st));
    long long want = (n + VISB_T - 1) / VISB_T;
    if (bpsm > 2) bpsm = 2;
    const long long maxb = (long long)s->num_sms * bpsm;
    int grid = (int)(want < maxb ? (want > 0 ? want : 1) : maxb);
    DevParams P = s->P;
    long long nn = n, ncm = ncell_max, ol = out_len;
    const double *px = A.x[0], *py = A.y[0], *pz = A.z[0];
    const int *spc = A.species;
    double rng_ = vision_range;
    int ntc = n_type_col;
    void *args[] = {&nn, &px, &py, &pz, &radii, &spc, &P, &type_col, &ntc, &rng_, &ncm, &VB, &VP, &use_prev, &G, &count, &start, &fill,
                    &cellid, &items, &sx, &sy, &sz, &sr, &scol, &out, &ol};
    SW_CUDA(cudaLaunchCooperativeKernel((void *)k_vis_build, dim3(grid), dim3(VISB_T), args, 0, st));
    s->launches += 1;
  } else {
  if (rows > 0) SW_CUDA(cudaMemsetAsync(out, 0, sizeof(double) * (size_t)rows * (size_t)n_cones * (size_t)n_cols, st));
  k_vis_init<<<gb > 0 ? gb : 1, T, 0, st>>>(G, count, ncell_max);
  k_vis_bbox<<<gb > 0 ? gb : 1, T, 0, st>>>(n, A.x[0], A.y[0], A.z[0], G);
  k_vis_grid<<<1, 32, 0, st>>>(G, vision_range, ncell_max);
  k_vis_count<<<B, T, 0, st>>>(n, A.x[0], A.y[0], A.z[0], G, cellid, count);
  k_scan_tiles<<<ntiles, SCAN_T, 0, st>>>(m, count, start, tiles);
  k_scan_sums<<<1, SCAN_T, 0, st>>>(ntiles, tiles);
  k_scan_add<<<sw_blocks(m, T), T, 0, st>>>(m, start, tiles, fill);
  k_vis_fill<<<B, T, 0, st>>>(n, cellid, fill, items);
  k_vis_sortcell<<<sw_blocks(ncell_max, T), T, 0, st>>>(G, start, items);
  k_vis_pack<<<B, T, 0, st>>>(n, items, A.x[0], A.y[0], A.z[0], radii, A.species, s->P, type_col, n_type_col, sx, sy, sz, sr, scol);
  s->launches += 10;
  }
  if (n_cones * n_cols <= VISW_MAX_ACC && !getenv("SWARM_B200_VIS_SERIAL")) {
    const size_t smem = sizeof(double) * (size_t)(n_cones * n_cols) * VISW_THREADS;
    const char *eg = getenv("SWARM_B200_VIS_GROUP");
    const int gs = eg ? atoi(eg) : 4;   // measured at C4 (tools/bench_vision.py): 4 lanes per observer is the fastest
#define VISW_LAUNCH(GS_)                                                                                                         \
  k_vis_cones_warp<GS_><<<sw_blocks(n * GS_, VISW_THREADS), VISW_THREADS, smem, st>>>(                                           \
      n, items, row_of, G, start, sx, sy, sz, sr, scol, s->P.n_dims, A.theta, A.q0, A.q1, A.q2, A.q3, dirs, vision_range,        \
      half_angle, n_cones, n_cols, out)
    if (gs == 32) VISW_LAUNCH(32);
    else if (gs == 16) VISW_LAUNCH(16);
    else if (gs == 8) VISW_LAUNCH(8);
    else if (gs == 2) VISW_LAUNCH(2);
    else VISW_LAUNCH(4);
#undef VISW_LAUNCH
  } else {
    k_vis_cones<<<sw_blocks(n, 128), 128, 0, st>>>(n, items, row_of, G, start, sx, sy, sz, sr, scol, s->P.n_dims, A.theta, A.q0,
                                                   A.q1, A.q2, A.q3, dirs, vision_range, half_angle, n_cones, n_cols, out);
  }
  s->launches += 1;
  SW_CUDA(cudaGetLastError());
  return SWARM_OK;
}
