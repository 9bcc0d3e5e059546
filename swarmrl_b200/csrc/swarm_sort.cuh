// swarm_sort.cuh -- spatial ordering of the particle slots (cell hash -> counting sort -> cell-start scan -> permutation).
// Included by swarm_engine.cu.
//
// ESPResSo keeps its particles in the cells of a regular decomposition (selected at
// swarmrl/engine/espresso.py:267-270); the kernels here keep them in caller-owned SoA arrays indexed by particle id.
// For the tiled kernel (swarm_tiled.cuh) the library additionally orders the particles in SPACE: slot -> id = perm[slot],
// such that consecutive slots are consecutive along a strip-wise traversal of the box --
//
//     key(i) = strip(x_i) * n_rows + row(y_i)        strips of width ~ sqrt(particles per CTA / density) along x,
//                                                     rows of a few particles each along y
//
// so that the `ppc` consecutive slots a CTA owns form a compact, roughly square tile with the SAME particle count in
// every CTA (what a geometric grid cannot give once an active swarm has clustered).  One radix pass with the whole key as
// the digit: histogram (atomicAdd), exclusive scan (= cell start), placement, and a per-bin ordering by id that makes
// the permutation independent of the atomics' arrival order.
//
// Nothing here decides a RESULT: forces are summed in ascending particle id (DevArrays.okey) and the noise is keyed by
// the id, so trajectories are bit-identical for any permutation; the sort only decides how many Verlet entries of a CTA
// point into another CTA.
#pragma once

struct SortGeom {
  int n_strips, n_rows;   // bins = n_strips * n_rows
  double inv_w, inv_h;    // 1 / strip width, 1 / row height (on coordinates wrapped into the box)
};

__device__ __forceinline__ int sw_sort_key(const DevParams &P, const SortGeom &G, double x, double y) {
  // wrapped (or clamped, without periodicity) coordinate in [0, L)
  double ux = x * P.inv_box[0], uy = y * P.inv_box[1];
  ux -= floor(ux); uy -= floor(uy);
  if (!P.periodic[0]) ux = fmin(fmax(x * P.inv_box[0], 0.0), 1.0);
  if (!P.periodic[1]) uy = fmin(fmax(y * P.inv_box[1], 0.0), 1.0);
  int sx = (int)(ux * P.box[0] * G.inv_w), ry = (int)(uy * P.box[1] * G.inv_h);
  sx = sx < 0 ? 0 : (sx >= G.n_strips ? G.n_strips - 1 : sx);
  ry = ry < 0 ? 0 : (ry >= G.n_rows ? G.n_rows - 1 : ry);
  // boustrophedon: odd strips run downwards, so the end of one strip is next to the start of the following one
  if (sx & 1) ry = G.n_rows - 1 - ry;
  return sx * G.n_rows + ry;
}

// K1: cell hash + histogram
__global__ void k_sort_key(const __grid_constant__ DevParams P, SortGeom G, long long n, const double *x, const double *y,
                           int *key, int *hist) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int k = sw_sort_key(P, G, x[i], y[i]);
  key[i] = k;
  atomicAdd(&hist[k], 1);
}

// K2: exclusive scan of the histogram (cell start), one CTA; also the placement cursors
__global__ void __launch_bounds__(1024) k_sort_scan(int nbins, const int *hist, int *start, int *cursor) {
  __shared__ int wsum[32];
  __shared__ int carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int b0 = 0; b0 < nbins; b0 += 1024) {
    const int b = b0 + threadIdx.x;
    const int v = (b < nbins) ? hist[b] : 0;
    int inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane >= d) inc += t;
    }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      int w = wsum[lane], winc = w;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, winc, d);
        if (lane >= d) winc += t;
      }
      wsum[lane] = winc - w;  // exclusive prefix of the warp sums
    }
    __syncthreads();
    const int excl = carry + wsum[warp] + inc - v;
    if (b < nbins) { start[b] = excl; cursor[b] = excl; }
    __syncthreads();
    if (threadIdx.x == 1023) carry = excl + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) start[nbins] = carry;
}

// K3: placement (slot within the bin in arrival order; K4 makes it deterministic)
__global__ void k_sort_place(long long n, const int *key, int *cursor, int *perm) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int slot = atomicAdd(&cursor[key[i]], 1);
  perm[slot] = (int)i;
}

// K4: ids of a bin in ascending order (bins hold a handful of particles; a clustered bin a few hundred)
__global__ void k_sort_fix(int nbins, const int *start, int *perm) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nbins) return;
  const int lo = start[b], hi = start[b + 1];
  for (int k = lo + 1; k < hi; ++k) {
    const int v = perm[k];
    int m = k;
    while (m > lo && perm[m - 1] > v) { perm[m] = perm[m - 1]; --m; }
    perm[m] = v;
  }
}

// K4b: within every tile (the `ppc` consecutive slots of one CTA) the particles that can have a neighbour in another tile
// come FIRST: those within `rc` = cut-off + skin of a strip edge, and those in the first / last rows of the tile.  The
// rim of a tile then occupies its first few warps and the remaining warps never execute (or wait for) a remote read.
// Stable partition, one CTA per tile.
__global__ void __launch_bounds__(1024) k_sort_rimfirst(const __grid_constant__ DevParams P, SortGeom G, long long n, int ppc, double rc,
                                                        int rows_margin, const double *x, const int *key, int *perm) {
  __shared__ int s_id[1024];
  __shared__ int wsum[32];
  __shared__ int s_total;
  const long long base = (long long)blockIdx.x * ppc;
  const int cnt = (int)((n - base < ppc) ? (n - base) : ppc);
  const int t = threadIdx.x;
  int id = -1, flag = 0;
  if (t < cnt) {
    id = perm[base + t];
    const int kf = key[perm[base]], kl = key[perm[base + cnt - 1]], k = key[id];
    double ux = x[id] * P.inv_box[0];
    ux -= floor(ux);
    if (!P.periodic[0]) ux = fmin(fmax(x[id] * P.inv_box[0], 0.0), 1.0);
    const double xs = ux * P.box[0] * G.inv_w;          // in strip widths
    const double fr = xs - floor(xs);
    const double w = 1.0 / G.inv_w;
    const bool edge = fr * w < rc || (1.0 - fr) * w < rc;
    flag = (edge || k - kf <= rows_margin || kl - k <= rows_margin) ? 1 : 0;
  }
  // exclusive scan of the flags over the CTA
  const int lane = t & 31, warp = t >> 5;
  const unsigned int b = __ballot_sync(0xffffffffu, flag);
  if (lane == 0) wsum[warp] = __popc(b);
  __syncthreads();
  if (warp == 0) {
    const int v = wsum[lane];
    int inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane >= d) inc += u;
    }
    wsum[lane] = inc - v;
    if (lane == 31) s_total = inc;
  }
  __syncthreads();
  const int before = wsum[warp] + __popc(b & ((1u << lane) - 1u));   // flagged particles in front of this one
  if (t < cnt) {
    const int dst = flag ? before : s_total + (t - before);
    s_id[dst] = id;
  }
  __syncthreads();
  if (t < cnt) perm[base + t] = s_id[t];
}

// K5: the per-slot constants the rebuild phases gather by slot (species, ordering key = Philox id)
__global__ void k_sort_gather(long long n, const int *perm, const int *species, const long long *gid, int *species_s,
                              long long *gid_s) {
  const long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  const int id = perm[s];
  species_s[s] = species[id];
  gid_s[s] = gid ? gid[id] : (long long)id;
}
