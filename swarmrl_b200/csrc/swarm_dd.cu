// swarm_dd.cu -- slab domain decomposition of ONE large periodic 2-D swarm over several GPUs
// (BASELINE config 5).  The reference has no counterpart: SwarmRL drives a
// single-rank ESPResSo (swarmrl/engine/espresso.py:193,1304-1306).
//
// Communication-avoiding scheme: rank g owns the slab x in [g W, (g+1) W) and
// keeps, in a LOCAL frame where the slab is [H, H + W), every particle with 0 <= x_loc < W + 2H.  The
// ghosts in the two halos are integrated REDUNDANTLY by the same Brownian kernel as the owned
// particles -- their noise is keyed by the global particle id (swarm_bind_ids), so a ghost follows
// bit-for-bit the random path its owner computes -- and the halos are re-shipped only every k steps,
// with H >= k (r_cut + margin): an error entering at the outer halo edge travels at most r_cut + 2
// max-displacements per step and never reaches the slab between two exchanges.
//
// One exchange = k_dd_pack (owned particles inside the two send bands -> fixed-capacity SoA record
// buffers with the record count in their first 8 bytes), NCCL send/recv of the two buffers (host side,
// torch.distributed), k_dd_merge (old owned + the two received buffers -> the other half of the
// ping-pong state, owned flag recomputed).  Migration needs nothing extra: a particle that left the
// slab lies in a send band, the receiver finds it inside its own slab and owns it from then on.
#include "swarm_common.cuh"
#include <cstring>

#define DD_FIXED 8  // x, y, theta, f_swim, tqz, gid, species, image

struct DdArrays {
  double *x, *y, *theta, *f_swim, *tqz;
  long long *gid;
  int *species, *image, *owned;
  double *aux[SWARM_DD_MAX_AUX];
};

static DdArrays dd_arrays(const swarm_dd_arrays_t *a) {
  DdArrays d;
  d.x = a->x; d.y = a->y; d.theta = a->theta; d.f_swim = a->f_swim; d.tqz = a->tqz;
  d.gid = (long long *)a->gid; d.species = a->species; d.image = a->image; d.owned = a->owned;
  for (int c = 0; c < SWARM_DD_MAX_AUX; ++c) d.aux[c] = a->aux[c];
  return d;
}

// live count: device-resident when given (no host sync between exchanges), bounded by the host's n
__device__ __forceinline__ long long dd_live(long long n, const int *n_dev) {
  if (!n_dev) return n;
  const long long m = (long long)__ldg(n_dev);   // L1-cached: a .cg load from a million threads hammers one L2 sector
  return m < n ? m : n;
}

// CTA-aggregated append: one atomicAdd per CTA (same-address atomics serialise in L2 at ~2.5 ns each, so
// one per warp made the 1M-particle merge atomic-bound).  Every thread of the CTA must call it; `slot`
// selects one of two independent appends per kernel.  Returns this thread's slot, or -1 when !want.
#define DD_THREADS 512
__device__ __forceinline__ long long dd_append(unsigned long long *counter, bool want, int slot) {
  __shared__ unsigned int wcnt[2][DD_THREADS / 32];
  __shared__ unsigned long long base[2];
  const unsigned m = __ballot_sync(0xFFFFFFFFu, want);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) wcnt[slot][warp] = __popc(m);
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int tot = 0;
    for (int w = 0; w < DD_THREADS / 32; ++w) { const unsigned int c = wcnt[slot][w]; wcnt[slot][w] = tot; tot += c; }
    base[slot] = tot ? atomicAdd(counter, (unsigned long long)tot) : 0ULL;
  }
  __syncthreads();
  return want ? (long long)(base[slot] + wcnt[slot][warp] + __popc(m & ((1u << lane) - 1u))) : -1;
}

__device__ __forceinline__ void dd_store(double *buf, long long cap, int n_aux, long long j, const DdArrays &L, long long i,
                                         double x_new, int image_new) {
  double *r = buf + 1;
  r[0 * cap + j] = x_new;
  r[1 * cap + j] = L.y[i];
  r[2 * cap + j] = L.theta[i];
  r[3 * cap + j] = L.f_swim[i];
  r[4 * cap + j] = L.tqz[i];
  r[5 * cap + j] = (double)L.gid[i];
  r[6 * cap + j] = (double)L.species[i];
  r[7 * cap + j] = (double)image_new;
#pragma unroll
  for (int c = 0; c < SWARM_DD_MAX_AUX; ++c)
    if (c < n_aux) r[(DD_FIXED + c) * cap + j] = L.aux[c][i];
}

__global__ void k_dd_pack(swarm_dd_geom_t G, long long n, const int *n_dev, DdArrays L, double *send_left, double *send_right) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const bool act = i < dd_live(n, n_dev) && L.owned[i] != 0;
  const double x = act ? L.x[i] : 0.0;
  const bool to_l = act && x < 2.0 * G.halo;   // within H of the left face, or already beyond it
  const bool to_r = act && x >= G.slab_w;      // within H of the right face (H + W - H), or beyond
  long long j = dd_append((unsigned long long *)send_left, to_l, 0);
  if (to_l && j < G.cap) dd_store(send_left, G.cap, G.n_aux, j, L, i, x + G.slab_w, L.image[i] + G.image_shift_left);
  j = dd_append((unsigned long long *)send_right, to_r, 1);
  if (to_r && j < G.cap) dd_store(send_right, G.cap, G.n_aux, j, L, i, x - G.slab_w, L.image[i] + G.image_shift_right);
}

// t < n_old: old local slot t (kept when owned); then the records of recv_left, then of recv_right
__global__ void k_dd_merge(swarm_dd_geom_t G, long long n_bound, const int *n_dev, DdArrays O, const double *recv_left, const double *recv_right,
                           long long cap_new, DdArrays N, int *counts) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long n_old = dd_live(n_bound, n_dev);
  const unsigned long long cl = *(const unsigned long long *)recv_left, cr = *(const unsigned long long *)recv_right;
  if (t == 0 && (cl > (unsigned long long)G.cap || cr > (unsigned long long)G.cap)) atomicOr(&counts[2], 1);
  const long long nl = cl > (unsigned long long)G.cap ? G.cap : (long long)cl;
  const long long nr = cr > (unsigned long long)G.cap ? G.cap : (long long)cr;
  const double *rec = nullptr;
  long long j = -1;
  bool cand = false;
  double x = 0.0;
  if (t < n_old) {
    cand = O.owned[t] != 0;
    if (cand) x = O.x[t];
  } else if (t < n_old + nl) {
    rec = recv_left + 1; j = t - n_old; cand = true; x = rec[j];
  } else if (t < n_old + nl + nr) {
    rec = recv_right + 1; j = t - n_old - nl; cand = true; x = rec[j];
  }
  const bool keep = cand && x >= 0.0 && x < G.slab_w + 2.0 * G.halo;
  if (cand && !keep && !rec) atomicOr(&counts[2], 2);  // an owned particle outran the halo: k or H too small
  const long long o = dd_append((unsigned long long *)(counts + 4), keep, 0);
  const bool fits = keep && o < cap_new;
  const int own = (fits && x >= G.halo && x < G.halo + G.slab_w) ? 1 : 0;
  const int own_cta = __syncthreads_count(own);
  if (threadIdx.x == 0 && own_cta) atomicAdd(&counts[1], own_cta);
  if (!keep) return;
  if (!fits) { atomicOr(&counts[2], 4); return; }
  N.x[o] = x;
  N.owned[o] = own;
  if (!rec) {
    N.y[o] = O.y[t]; N.theta[o] = O.theta[t]; N.f_swim[o] = O.f_swim[t]; N.tqz[o] = O.tqz[t];
    N.gid[o] = O.gid[t]; N.species[o] = O.species[t]; N.image[o] = O.image[t];
#pragma unroll
    for (int c = 0; c < SWARM_DD_MAX_AUX; ++c)
      if (c < G.n_aux) N.aux[c][o] = O.aux[c][t];
  } else {
    const long long cap = G.cap;
    N.y[o] = rec[1 * cap + j]; N.theta[o] = rec[2 * cap + j]; N.f_swim[o] = rec[3 * cap + j]; N.tqz[o] = rec[4 * cap + j];
    N.gid[o] = (long long)rec[5 * cap + j]; N.species[o] = (int)rec[6 * cap + j]; N.image[o] = (int)rec[7 * cap + j];
#pragma unroll
    for (int c = 0; c < SWARM_DD_MAX_AUX; ++c)
      if (c < G.n_aux) N.aux[c][o] = rec[(DD_FIXED + c) * cap + j];
  }
}

// lean mode (swarm_dd_set_lean): the counters the NEXT exchange starts from are cleared here, behind the merge that was their
// last reader, instead of by four memset launches in front of the next pack and merge
__global__ void k_dd_finish(int *counts, int *old_counts, double *send_left, double *send_right) {
  const unsigned long long n = *(unsigned long long *)(counts + 4);
  counts[0] = (int)n;
  counts[3] |= counts[2];   // sticky error bits: survive the next merge's reset of counts[2]
  if (old_counts) {
    old_counts[0] = 0; old_counts[1] = 0; old_counts[2] = 0;
    old_counts[4] = 0; old_counts[5] = 0; old_counts[6] = 0; old_counts[7] = 0;
    *(unsigned long long *)send_left = 0ULL;
    *(unsigned long long *)send_right = 0ULL;
  }
}

static int dd_check(const swarm_dd_geom_t *g) {
  if (!g || g->n_aux < 0 || g->n_aux > SWARM_DD_MAX_AUX || g->cap < 1 || !(g->slab_w > 0.0) || !(g->halo >= 0.0)) {
    sw_set_error("bad domain-decomposition geometry"); return SWARM_ERR_INVALID;
  }
  if (g->slab_w < 2.0 * g->halo) { sw_set_error("slab narrower than two halos: use fewer ranks or a smaller halo"); return SWARM_ERR_INVALID; }
  return SWARM_OK;
}

extern "C" int64_t swarm_dd_buffer_doubles(const swarm_dd_geom_t *g) {
  if (!g) return -1;
  return 1 + (int64_t)(DD_FIXED + g->n_aux) * g->cap;
}

extern "C" int swarm_dd_pack(swarm_t *s, const swarm_dd_geom_t *g, int64_t n, const int32_t *n_dev, const swarm_dd_arrays_t *loc,
                             double *send_left_dev, double *send_right_dev, void *stream) {
  if (!s || !loc || !send_left_dev || !send_right_dev || n < 0) { sw_set_error("bad arguments"); return SWARM_ERR_INVALID; }
  // a halo record carries x, y, theta and the aux channels.  Under the Langevin integrator v and omega are state: the caller
  // binds vx, vy, wz to three aux channels of the record (DomainSwarm(velocities=True)) and cuts its runs with swarm_run_part
  int rc = dd_check(g);
  if (rc) return rc;
  SW_CUDA(cudaSetDevice(s->cfg.device));
  cudaStream_t st = (cudaStream_t)stream;
  if (!(s->dd_lean && s->dd_send_l == send_left_dev && s->dd_send_r == send_right_dev)) {   // lean: cleared by the last merge
    SW_CUDA(cudaMemsetAsync(send_left_dev, 0, 8, st));
    SW_CUDA(cudaMemsetAsync(send_right_dev, 0, 8, st));
  }
  s->dd_send_l = nullptr; s->dd_send_r = nullptr;
  s->dd_last_send_l = send_left_dev; s->dd_last_send_r = send_right_dev;
  if (n > 0) {
    k_dd_pack<<<(unsigned)((n + DD_THREADS - 1) / DD_THREADS), DD_THREADS, 0, st>>>(*g, n, (const int *)n_dev, dd_arrays(loc), send_left_dev, send_right_dev);
    s->launches += 1;
  }
  SW_CUDA(cudaGetLastError());
  return SWARM_OK;
}

extern "C" int swarm_dd_merge(swarm_t *s, const swarm_dd_geom_t *g, int64_t n_old, const int32_t *n_old_dev, const swarm_dd_arrays_t *old_loc,
                              const double *recv_left_dev, const double *recv_right_dev, int64_t cap_new,
                              const swarm_dd_arrays_t *new_loc, int32_t *counts8_dev, void *stream) {
  if (!s || !old_loc || !new_loc || !recv_left_dev || !recv_right_dev || !counts8_dev || n_old < 0 || cap_new < 1) {
    sw_set_error("bad arguments"); return SWARM_ERR_INVALID;
  }
  int rc = dd_check(g);
  if (rc) return rc;
  SW_CUDA(cudaSetDevice(s->cfg.device));
  cudaStream_t st = (cudaStream_t)stream;
  // counts[3] (sticky error bits) survives; everything else restarts
  if (!(s->dd_lean && s->dd_clean_counts == counts8_dev)) {   // lean: cleared by the previous merge
    SW_CUDA(cudaMemsetAsync(counts8_dev, 0, 3 * sizeof(int), st));
    SW_CUDA(cudaMemsetAsync(counts8_dev + 4, 0, 4 * sizeof(int), st));
  }
  s->dd_clean_counts = nullptr;
  const long long total = n_old + 2LL * g->cap;
  k_dd_merge<<<(unsigned)((total + DD_THREADS - 1) / DD_THREADS), DD_THREADS, 0, st>>>(*g, n_old, (const int *)n_old_dev, dd_arrays(old_loc), recv_left_dev, recv_right_dev, cap_new,
                                                             dd_arrays(new_loc), counts8_dev);
  const bool lean = s->dd_lean && n_old_dev && s->dd_last_send_l && s->dd_last_send_r;
  k_dd_finish<<<1, 1, 0, st>>>(counts8_dev, lean ? (int *)n_old_dev : nullptr, s->dd_last_send_l, s->dd_last_send_r);
  if (lean) { s->dd_clean_counts = (int *)n_old_dev; s->dd_send_l = s->dd_last_send_l; s->dd_send_r = s->dd_last_send_r; }
  s->launches += 2;
  SW_CUDA(cudaGetLastError());
  return SWARM_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// Exchange over PEER MEMORY (NVLink / NVSwitch P2P stores) instead of NCCL send/recv.
//
// Every rank owns one exchange region, allocated here with cudaMalloc and exported with cudaIpcGetMemHandle:
//     [ 4 flags, 256 bytes apart: left[0], left[1], right[0], right[1] ] [ recv_left[0] ][ recv_left[1] ][ recv_right[0] ][ recv_right[1] ]
// (two parities of each receive buffer).  Exchange e (1, 2, 3, ...) uses parity e & 1:
//   sender   k_dd_push: the LIVE records of its packed send buffers (count x fields doubles, not the buffer's capacity)
//            go straight into the neighbours' receive buffers with coalesced remote stores; every CTA fences at system
//            scope and takes a ticket, the last one stores flag = e with release semantics at system scope;
//   receiver k_dd_wait: ONE thread polls its own two flags (local memory) with acquire loads at system scope; the merge
//            kernel behind it in the stream then reads the records from local memory.
// No collective, no proxy thread, no rendezvous: the data path of an exchange is two kernels on the sender's stream and a
// one-thread kernel on the receiver's.  Two parities suffice: a sender overwrites parity p at exchange e + 2 after its own
// merge of e + 1, which waited for the neighbour's push of e + 1, which the neighbour issued behind its merge of e.
// A flag that does not arrive within SW_DD_WAIT_NS sets error bit 8 instead of hanging the GPU.
// ---------------------------------------------------------------------------------------------------------------------
#define SW_DD_FLAG_STRIDE 256
#define SW_DD_WAIT_NS 20000000000ULL   // 20 s

__device__ __forceinline__ unsigned long long dd_globaltimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// blockIdx.y = face (0: my send_left -> the left neighbour's recv_right, 1: my send_right -> the right neighbour's recv_left)
__global__ void __launch_bounds__(256) k_dd_push(const double *send_left, const double *send_right, double *peer_of_left, double *peer_of_right,
                                                  unsigned long long *flag_of_left, unsigned long long *flag_of_right, unsigned long long epoch,
                                                  long long cap, int nfields, unsigned int *tickets) {
  const int face = blockIdx.y;
  const double *src = face ? send_right : send_left;
  double *dst = face ? peer_of_right : peer_of_left;
  const unsigned long long raw = *(const unsigned long long *)src;
  const long long cnt = raw > (unsigned long long)cap ? cap : (long long)raw;
  const long long total = cnt * nfields;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const long long f = t / cnt, j = t - f * cnt;
    dst[1 + f * cap + j] = src[1 + f * cap + j];
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) *(unsigned long long *)dst = raw;   // the record count (beyond cap = overflow, seen by the merge)
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int t = atomicAdd(&tickets[face], 1u);
    if (t == gridDim.x - 1) {
      tickets[face] = 0u;
      __threadfence_system();
      unsigned long long *flag = face ? flag_of_right : flag_of_left;
      asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(flag), "l"(epoch) : "memory");
    }
  }
}

__global__ void k_dd_wait(const unsigned long long *flag_left, const unsigned long long *flag_right, unsigned long long epoch, int *counts) {
  const unsigned long long t0 = dd_globaltimer();
  for (int side = 0; side < 2; ++side) {
    const unsigned long long *f = side ? flag_right : flag_left;
    for (;;) {
      unsigned long long v;
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(f) : "memory");
      if (v == epoch) break;
      if (dd_globaltimer() - t0 > SW_DD_WAIT_NS) { atomicOr(&counts[3], 8); return; }   // sticky: the neighbour never delivered
      __nanosleep(200);
    }
  }
}

extern "C" int swarm_dd_set_lean(swarm_t *s, int32_t on) {
  if (!s) return SWARM_ERR_INVALID;
  s->dd_lean = on != 0;
  s->dd_clean_counts = nullptr; s->dd_send_l = nullptr; s->dd_send_r = nullptr; s->dd_last_send_l = nullptr; s->dd_last_send_r = nullptr;
  return SWARM_OK;
}

extern "C" int64_t swarm_dd_peer_region_bytes(const swarm_dd_geom_t *g) {
  if (!g) return -1;
  return 4 * SW_DD_FLAG_STRIDE + 4 * (int64_t)sizeof(double) * swarm_dd_buffer_doubles(g);
}

extern "C" int swarm_dd_peer_alloc(int32_t device, const swarm_dd_geom_t *g, void **region_out, unsigned char handle64_out[64]) {
  if (!g || !region_out || !handle64_out) { sw_set_error("bad arguments"); return SWARM_ERR_INVALID; }
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "the IPC handle travels as 64 bytes");
  SW_CUDA(cudaSetDevice(device));
  const size_t bytes = (size_t)swarm_dd_peer_region_bytes(g);
  void *p = nullptr;
  SW_CUDA(cudaMalloc(&p, bytes));
  SW_CUDA(cudaMemset(p, 0, bytes));
  SW_CUDA(cudaDeviceSynchronize());
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) { cudaFree(p); sw_set_error(cudaGetErrorString(e)); return SWARM_ERR_CUDA; }
  memcpy(handle64_out, &h, 64);
  *region_out = p;
  return SWARM_OK;
}

extern "C" int swarm_dd_peer_open(int32_t device, const unsigned char handle64[64], void **region_out) {
  if (!handle64 || !region_out) { sw_set_error("bad arguments"); return SWARM_ERR_INVALID; }
  SW_CUDA(cudaSetDevice(device));
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  SW_CUDA(cudaIpcOpenMemHandle(region_out, h, cudaIpcMemLazyEnablePeerAccess));
  return SWARM_OK;
}

extern "C" int swarm_dd_peer_close(void *region) {
  if (region) SW_CUDA(cudaIpcCloseMemHandle(region));
  return SWARM_OK;
}

extern "C" int swarm_dd_peer_free(void *region) {
  if (region) SW_CUDA(cudaFree(region));
  return SWARM_OK;
}

static inline unsigned long long *dd_flag(void *region, int side, int parity) {
  return (unsigned long long *)((char *)region + (size_t)(2 * side + parity) * SW_DD_FLAG_STRIDE);
}
static inline double *dd_recv(void *region, const swarm_dd_geom_t *g, int side, int parity) {
  return (double *)((char *)region + 4 * SW_DD_FLAG_STRIDE) + (size_t)(2 * side + parity) * (size_t)swarm_dd_buffer_doubles(g);
}

extern "C" double *swarm_dd_peer_recv(void *region, const swarm_dd_geom_t *g, int32_t side, int32_t parity) {
  if (!region || !g) return nullptr;
  return dd_recv(region, g, side ? 1 : 0, parity & 1);
}

extern "C" int swarm_dd_push(swarm_t *s, const swarm_dd_geom_t *g, const double *send_left_dev, const double *send_right_dev,
                             void *left_neighbour_region, void *right_neighbour_region, uint64_t epoch, void *stream) {
  if (!s || !send_left_dev || !send_right_dev || !left_neighbour_region || !right_neighbour_region || epoch == 0) {
    sw_set_error("bad arguments"); return SWARM_ERR_INVALID;
  }
  int rc = dd_check(g);
  if (rc) return rc;
  SW_CUDA(cudaSetDevice(s->cfg.device));
  if (!s->dd_tickets) {
    SW_CUDA(cudaMalloc((void **)&s->dd_tickets, 256));
    SW_CUDA(cudaMemset(s->dd_tickets, 0, 256));
  }
  const int par = (int)(epoch & 1ULL);
  // my left face enters the left neighbour through ITS right face, and the other way round
  k_dd_push<<<dim3(32, 2), 256, 0, (cudaStream_t)stream>>>(send_left_dev, send_right_dev, dd_recv(left_neighbour_region, g, 1, par),
                                                           dd_recv(right_neighbour_region, g, 0, par), dd_flag(left_neighbour_region, 1, par),
                                                           dd_flag(right_neighbour_region, 0, par), (unsigned long long)epoch, (long long)g->cap,
                                                           DD_FIXED + g->n_aux, s->dd_tickets);
  s->launches += 1;
  SW_CUDA(cudaGetLastError());
  return SWARM_OK;
}

extern "C" int swarm_dd_wait(swarm_t *s, void *my_region, uint64_t epoch, int32_t *counts8_dev, void *stream) {
  if (!s || !my_region || !counts8_dev || epoch == 0) { sw_set_error("bad arguments"); return SWARM_ERR_INVALID; }
  SW_CUDA(cudaSetDevice(s->cfg.device));
  const int par = (int)(epoch & 1ULL);
  k_dd_wait<<<1, 1, 0, (cudaStream_t)stream>>>(dd_flag(my_region, 0, par), dd_flag(my_region, 1, par), (unsigned long long)epoch, counts8_dev);
  s->launches += 1;
  SW_CUDA(cudaGetLastError());
  return SWARM_OK;
}

// ConcentrationField / GradientSensing of the owned particles on their GLOBAL unfolded coordinates
// (observables/concentration_field.py:84-138, tasks/searching/gradient_sensing.py:92-150).  The history
// p_prev = (x, y)/box_length travels with the particle in aux channels 0 and 1.  out[i] = 0 for ghosts.
__device__ __forceinline__ double dd_decay(int kind, double c0, double c1, double d) {
  switch (kind) {
    case SWARM_DECAY_LINEAR: return c0 + c1 * d;
    case SWARM_DECAY_EXP: return c0 * exp(-c1 * d);
    case SWARM_DECAY_INVERSE: return c0 / (d + c1);
    default: return d;
  }
}

__global__ void k_dd_obs_field(long long n, const int *n_dev, DdArrays L, double x_off, double box_x, double s0, double s1, double s2, double b0,
                               double b1, int decay, double c0, double c1, double scale, int clip, int init_only, double *out,
                               double *ha, double *hb, int all_local) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (i >= dd_live(n, n_dev) || !(all_local || L.owned[i])) { out[i] = 0.0; return; }
  const double xg = L.x[i] + x_off + (double)L.image[i] * box_x;
  const double p0 = xg / b0, p1 = L.y[i] / b1;
  const double h0 = ha[i], h1 = hb[i];
  ha[i] = p0; hb[i] = p1;
  if (init_only) { out[i] = 0.0; return; }
  const double dn = sqrt((s0 - p0) * (s0 - p0) + (s1 - p1) * (s1 - p1) + s2 * s2);
  const double dp = sqrt((s0 - h0) * (s0 - h0) + (s1 - h1) * (s1 - h1) + s2 * s2);
  double v = scale * (dd_decay(decay, c0, c1, dn) - dd_decay(decay, c0, c1, dp));
  if (clip && !(v > 0.0)) v = 0.0;
  out[i] = v;
}

static int dd_obs_field(swarm_t *s, int64_t n, const int32_t *n_dev, const swarm_dd_arrays_t *loc, double x_offset, double box_x,
                        const double src[3], const double box_length[2], int32_t decay, double c0, double c1, double scale,
                        int32_t clip_nonneg, int32_t init_only, int32_t aux_base, int32_t all_local, double *out_dev, void *stream) {
  if (!s || !loc || !src || !box_length || !out_dev || n < 0 || aux_base < 0 || aux_base + 2 > SWARM_DD_MAX_AUX ||
      !loc->aux[aux_base] || !loc->aux[aux_base + 1]) {
    sw_set_error("bad arguments (the history needs two aux channels)"); return SWARM_ERR_INVALID;
  }
  if (decay < SWARM_DECAY_NONE || decay > SWARM_DECAY_INVERSE) { sw_set_error("unknown decay kind"); return SWARM_ERR_INVALID; }
  SW_CUDA(cudaSetDevice(s->cfg.device));
  if (n == 0) return SWARM_OK;
  k_dd_obs_field<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(n, (const int *)n_dev, dd_arrays(loc), x_offset, box_x, src[0], src[1],
                                                                               src[2], box_length[0], box_length[1], decay, c0, c1,
                                                                               scale, clip_nonneg, init_only, out_dev, loc->aux[aux_base],
                                                                               loc->aux[aux_base + 1], all_local);
  s->launches += 1;
  SW_CUDA(cudaGetLastError());
  return SWARM_OK;
}

extern "C" int swarm_dd_obs_field(swarm_t *s, int64_t n, const int32_t *n_dev, const swarm_dd_arrays_t *loc, double x_offset, double box_x,
                                  const double src[3], const double box_length[2], int32_t decay, double c0, double c1,
                                  double scale, int32_t clip_nonneg, int32_t init_only, double *out_dev, void *stream) {
  return dd_obs_field(s, n, n_dev, loc, x_offset, box_x, src, box_length, decay, c0, c1, scale, clip_nonneg, init_only, 0, 0, out_dev, stream);
}

extern "C" int swarm_dd_obs_field_ch(swarm_t *s, int64_t n, const int32_t *n_dev, const swarm_dd_arrays_t *loc, double x_offset,
                                     double box_x, const double src[3], const double box_length[2], int32_t decay, double c0, double c1,
                                     double scale, int32_t clip_nonneg, int32_t init_only, int32_t aux_base, int32_t all_local,
                                     double *out_dev, void *stream) {
  return dd_obs_field(s, n, n_dev, loc, x_offset, box_x, src, box_length, decay, c0, c1, scale, clip_nonneg, init_only, aux_base, all_local,
                      out_dev, stream);
}
