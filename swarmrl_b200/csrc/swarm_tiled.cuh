// swarm_tiled.cuh -- k_bd2d_tiled: the 2-D Brownian chunk on SPATIAL TILES (BASELINE config 4, 100k agents).
// Included by swarm_engine.cu after swarm_bd2d.cuh (same step arithmetic) and swarm_sort.cuh (the slot order).
//
// k_bd2d_resident pays, per step, one grid-wide barrier (148 arrivals on one L2 word, release, poll) and one L2 round
// trip for every neighbour position, because slot = particle id and a particle's neighbours live anywhere.  Here the
// slots are spatially sorted (swarm_sort.cuh), CTA b owns slots [b * ppc, (b + 1) * ppc) -- a compact tile -- and
//
//  * the positions of a tile live in SHARED MEMORY (double-buffered); a Verlet entry that points into the own tile is an
//    LDS, not an L2 load.  Only entries that point into another CTA's tile ("remote", the rim of the tile) go through L2;
//  * there is NO per-step barrier beyond the CTA's own, and NO memory fence on the step path.  A particle whose list
//    points into another tile ("rim" particle; lists are symmetric, so it is also read from there) publishes, after every
//    step, a 32-byte RECORD of four 8-byte words {epoch flag, 32 bits of x or y} into the slot of the epoch's parity; the
//    reader loads the record with relaxed loads and accepts it only when all four flags are the epoch it needs (an
//    aligned 8-byte store is atomic, so a word with the right flag carries the right data -- NCCL's LL protocol); an
//    incomplete or old record is simply read again.  A record is its own acknowledgement: no release / acquire pair -- the fence of a release (MEMBAR.GPU, 3,700 cycles measured on
//    the flag-based first version, profiles/r02_phase_timing_flag_version.txt) was the longest link of the step.  The
//    data arrive one L2 round trip after the neighbour's step, i.e. long before they are needed: the reader's own
//    particle-local work of the next step (Philox, Box-Muller, director) lies in between.  Two record slots suffice:
//    thread i overwrites slot p at epoch e + 2 only after it has read every partner j's record of epoch e + 2, which j
//    wrote after reading i's record of epoch e + 1 from slot p (same thread, program order);
//  * the records a tile needs (the rim of its neighbours, ~150 of them) are fetched by ONE dedicated HALO WARP into a
//    halo region behind the tile's own positions in shared memory, while the 22 particle warps do the next step's
//    particle-local work: every Verlet entry of the step -- own tile or halo -- is then a plain LDS, no particle warp
//    ever waits for L2 (with the particle threads polling their own remote entries one after the other, the rim warps
//    spent 5,000-8,000 cycles per step on serialised L2 round trips and the whole CTA waited for them at the step
//    barrier: profiles/r02_phase_warps_direct_read.txt).  Which remote slots a tile needs is collected at every rebuild
//    in a shared-memory hash table; an entry that does not fit keeps its raw slot and is read directly (correct for any
//    slot order, merely slow);
//  * the Verlet rebuild cannot ride on a per-step barrier any more, so it runs on a CADENCE: every K steps all CTAs
//    meet at a grid barrier, VALIDATE that no position used for a force evaluation since the last rebuild had moved
//    more than skin / 2 (the criterion of ESPResSo's Verlet lists, espresso.py:269), and rebuild.  If the validation
//    fails, the period is ROLLED BACK to its checkpoint (positions at the last rebuild; the noise is counter-based, so
//    a replay draws the same numbers) and replayed with a shorter K; with K = 1 every step is validated, so this
//    terminates.  K adapts to the largest displacement seen.  Since forces are summed in ascending particle id
//    (DevArrays.okey) and a valid list contains every pair inside the cut-off, the trajectory does not depend on K, on
//    the tiling or on the permutation: it is bit-identical to k_bd2d_resident's.
#pragma once

#define SW_TILE_MAXP (SW_RES_MAXT - 32)   // particles per CTA = threads per CTA (22 warps: 88 registers each)
#define SW_TILE_MAXCTA 256
#define SW_TILE_KMAX 512
#define SW_TILE_SNBR 8                      // Verlet entries SW_RES_NBR .. SW_RES_NBR + 7 of a particle are cached in shared memory
#define SW_TILE_HT 1024                     // halo hash table: remote slots this tile's lists point at (load factor <= 1/2)
#define SW_TILE_W (SW_TILE_MAXP + SW_TILE_HT)   // positions per buffer: the tile, then its halo
#define SW_TILE_RAW 0x08000000              // entry flag: remote slot that did not fit the halo table (read directly)
#define SW_TILE_WRAP 0x04000000             // entry flag: the pair's separation needs the minimum image (set at the rebuild)
#define SW_TILE_JMASK 0x03FFFFFF            // slot / buffer index of an entry (below the two flags)
// dynamic shared memory in doubles: x, y double-buffered over tile + halo; 8 per-thread columns; the list cache and the
// hash table / compact list as ints
#define SW_TILE_DSM_DOUBLES (4 * SW_TILE_W + 8 * SW_TILE_MAXP + (SW_TILE_SNBR * SW_TILE_MAXP + 2 * SW_TILE_HT) / 2)

// A rim record is four 8-byte words, each {flag (low half), 32 bits of data (high half)}: x lo, x hi, y lo, y hi -- the
// layout of NCCL's LL protocol.  An aligned 8-byte store is single-copy atomic, so a word whose flag matches carries
// data of that epoch; the record is complete when all four flags match.  32 bytes = one L2 sector, two 16-byte accesses.
struct RimRec { unsigned long long w[4]; };

// One 32-byte record per particle for the rebuild's search: position, chain link, species and ordering key travel in ONE
// L2 sector (the search is bound by L2 sectors: five separate arrays cost five sectors per candidate, 113 MB per rebuild).
struct CellRec { double x, y; int next, species; long long key; };

struct TileArgs {
  CellRec *crec;            // [slot]
  const int *perm;          // slot -> particle id
  RimRec *rim;              // [epoch parity][rim_stride]: one buffer, so that the parity is an offset, not an index into
                            // a kernel-parameter array (a dynamically indexed parameter array is copied to local memory)
  long long rim_stride;
  unsigned long long *head64;  // linked-cell heads of the tiled rebuild: (rebuild tag << 28) | slot; an entry with another tag is empty
  unsigned long long tag0;  // first rebuild tag of this launch (host: grows by more than a launch can use)
  int ppc;                  // slots per CTA
  int k_init;               // first cadence when the handle has none yet
  int k_cap;                // upper bound of the cadence (SW_TILE_KMAX; lower for measurements: SWARM_B200_TILE_KCAP)
  double k_safety;          // the next period = (steps the list lasted, extrapolated) x this factor (SWARM_B200_TILE_SAFETY)
  unsigned int flag0;       // flag of this launch's epoch 0: EVEN (flag parity = epoch parity = record slot) and above
                            // every flag of the previous launch (launch counter << 15; a launch runs < 2^15 epochs)
};

// -DSW_TILE_MBAR (A/B build, NOT the default): split CTA barrier of the step loop.  Measured on the same box against the
// plain __syncthreads at the end of the iteration (profiles/r02_mbar_ab.txt): 0.455-0.482 ms per 100-step launch with the
// mbarrier against 0.445-0.459 ms without -- the polling loop of 23 warps costs more issue slots than the overlap of a
// slow warp's pair phase with the others' local work wins, so __syncthreads stays.
// Split CTA barrier (mbarrier in shared memory): a warp ARRIVES as soon as its new positions are in
// shared memory and WAITS only after the particle-local work of the next step.  With __syncthreads at the end of the
// iteration every warp waited for the slowest warp's pair phase AND local work; now a warp waits only if another warp
// has not finished its pair phase by the time the waiter has finished its own local work, and the latency-bound pair
// phase of a slow warp overlaps the issue-bound local work of the fast ones.  Lane 0 arrives for its warp after
// __syncwarp (orders the lanes' stores before the releasing arrive); every lane polls (acquire).
__device__ __forceinline__ void tile_mbar_init(unsigned long long *b, unsigned int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((unsigned int)__cvta_generic_to_shared(b)), "r"(count) : "memory");
}
__device__ __forceinline__ void tile_mbar_arrive_warp(unsigned long long *b) {
  __syncwarp();
  if ((threadIdx.x & 31) == 0)
    asm volatile("mbarrier.arrive.release.cta.shared::cta.b64 _, [%0];" ::"r"((unsigned int)__cvta_generic_to_shared(b)) : "memory");
}
__device__ __forceinline__ void tile_mbar_wait(unsigned long long *b, unsigned int parity) {
  const unsigned int a = (unsigned int)__cvta_generic_to_shared(b);
  unsigned int ok;
  do {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.acquire.cta.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(a), "r"(parity) : "memory");
  } while (!ok);
}

__device__ __forceinline__ void rim_write(RimRec *r, unsigned int flag, double x, double y) {
  const unsigned long long f = flag;
  const unsigned long long w0 = ((unsigned long long)(unsigned int)__double2loint(x) << 32) | f;
  const unsigned long long w1 = ((unsigned long long)(unsigned int)__double2hiint(x) << 32) | f;
  const unsigned long long w2 = ((unsigned long long)(unsigned int)__double2loint(y) << 32) | f;
  const unsigned long long w3 = ((unsigned long long)(unsigned int)__double2hiint(y) << 32) | f;
  asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(r), "l"(w0), "l"(w1) : "memory");
  asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(&r->w[2]), "l"(w2), "l"(w3) : "memory");
}
// the neighbour's position at the epoch of `flag`: spins until its record of that epoch is complete (it normally has
// been for thousands of cycles: the reader's own particle-local work lies between the neighbour's store and this load)
__device__ __forceinline__ void rim_read(const RimRec *r, unsigned int flag, double &x, double &y) {
  unsigned long long w0, w1, w2, w3;
  for (;;) {
    asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(w0), "=l"(w1) : "l"(r) : "memory");
    asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(w2), "=l"(w3) : "l"(&r->w[2]) : "memory");
    const unsigned int bad = ((unsigned int)w0 ^ flag) | ((unsigned int)w1 ^ flag) | ((unsigned int)w2 ^ flag) | ((unsigned int)w3 ^ flag);
    if (bad == 0u) break;
  }
  x = __hiloint2double((int)(w1 >> 32), (int)(w0 >> 32));
  y = __hiloint2double((int)(w3 >> 32), (int)(w2 >> 32));
}

// ---------------------------------------------------------------------------------------------------------------------
// Rebuild of the tiled kernel: linked cells WITHOUT a clear phase and a search whose L2 round trips overlap.
//  * a cell head is 64 bits, (rebuild tag << 28) | slot.  A head whose tag is not the current rebuild's is an empty cell,
//    so nothing has to be unlinked first: phase_clear and its grid barrier are gone (insert -> barrier -> search);
//  * the search loads the heads of all stencil cells at once, then the first particle of three cells at a time, and
//    takes the ordering key (particle id) of a candidate together with its position -- phase_search walks cell after
//    cell and re-reads the keys while sorting: ~20 dependent L2 round trips per particle against ~5 here.
// Same pair test, bit for bit, as phase_search (the lists are the oracle's pair sets: tests/test_gpu_tiled.py).
// ---------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void tile_insert(const DevParams &P, const DevArrays &T, CellRec *crec, unsigned long long *head64,
                                            unsigned long long tag, long long slot, double xi, double yi, int sp, long long key) {
  int c[3];
  cell_coords<2>(P, xi, yi, 0.0, c);
  const unsigned long long mine = (tag << 28) | (unsigned long long)slot;
  const unsigned long long old = atomicExch(&head64[cell_id(P, 0, c)], mine);
  const int nx = ((old >> 28) == tag) ? (int)(old & 0x0FFFFFFFULL) : -1;
  // two 16-byte stores; read (after a grid barrier) with .cg loads
  double2 a = make_double2(xi, yi);
  longlong2 b = make_longlong2((long long)(((unsigned long long)(unsigned int)sp << 32) | (unsigned int)nx), key);
  *reinterpret_cast<double2 *>(&crec[slot].x) = a;
  *reinterpret_cast<longlong2 *>(&crec[slot].next) = b;
}

// Verlet list of `slot` at (xi, yi): entries (type << 28 | slot of the neighbour) in ascending okey into loc[], count returned
__device__ __forceinline__ int tile_search(const DevParams &P, const DevArrays &T, const CellRec *crec, const unsigned long long *head64,
                                           unsigned long long tag, long long slot, double xi, double yi, int ti, int *loc, long long *lkey, bool &over) {
  int c0[3];
  cell_coords<2>(P, xi, yi, 0.0, c0);
  int lo[2], hi[2];
#pragma unroll
  for (int a = 0; a < 2; ++a) {
    if (P.nc[a] == 1) { lo[a] = 0; hi[a] = 0; }
    else if (P.nc[a] == 2) { lo[a] = 0; hi[a] = 1; }
    else { lo[a] = -1; hi[a] = 1; }
  }
  // heads of the (up to) nine cells, all loads in flight together
  int hd[9];
#pragma unroll
  for (int q = 0; q < 9; ++q) {
    const int dx = q % 3 - 1, dy = q / 3 - 1;
    bool ok = dx >= lo[0] && dx <= hi[0] && dy >= lo[1] && dy <= hi[1];
    int cc[3] = {c0[0] + dx, c0[1] + dy, 0};
#pragma unroll
    for (int a = 0; a < 2; ++a) {
      if (cc[a] < 0 || cc[a] >= P.nc[a]) {
        if (P.periodic[a]) cc[a] = (cc[a] + P.nc[a]) % P.nc[a];
        else ok = false;
      }
    }
    hd[q] = -1;
    if (ok) {
      const unsigned long long v = __ldcg(&head64[cell_id(P, 0, cc)]);
      if ((v >> 28) == tag) hd[q] = (int)(v & 0x0FFFFFFFULL);
    }
  }
  int cnt = 0;
  over = false;
  auto consider = [&](int j, double xj, double yj, int spj, long long kj) {
    if (j == (int)slot) return;
    const int tj = P.sp_type[spj];
    const double rx = xi - xj, ry = yi - yj;
    const double ddx = sw_min_image(rx, P.box[0], P.inv_box[0], P.periodic[0]);
    const double ddy = sw_min_image(ry, P.box[1], P.inv_box[1], P.periodic[1]);
    const double rl = __dadd_rn(P.wca_cut[ti][tj], P.skin);
    if (sw_r2(ddx, ddy, 0.0) < __dmul_rn(rl, rl)) {
      // Does this pair need the minimum image at all?  Its image shift rint(d / L) is constant until the next rebuild (the
      // separation changes by less than the skin, far below L / 2 in a box of at least four list radii), and it is zero
      // for all but the pairs across a periodic face or between particles with different winding numbers: the step then
      // takes d = x_i - x_j as it is -- the value the minimum image returns for a zero shift, bit for bit.
      const double reach = 4.0 * (rl + P.skin);
      const bool wrap = ddx != rx || ddy != ry || (P.periodic[0] && P.box[0] < reach) || (P.periodic[1] && P.box[1] < reach);
      if (cnt < T.cap) {
        int k = cnt++;
        while (k > 0 && lkey[k - 1] > kj) { loc[k] = loc[k - 1]; lkey[k] = lkey[k - 1]; --k; }
        loc[k] = j | (tj << SW_TYPE_SHIFT) | (wrap ? SW_TILE_WRAP : 0);
        lkey[k] = kj;
      } else {
        over = true;
      }
    }
  };
#pragma unroll
  for (int g = 0; g < 3; ++g) {
    // first particle of the three cells of one stencil row: independent loads, one record (one sector) each
    double2 XY[3];
    longlong2 NK[3];
#pragma unroll
    for (int u = 0; u < 3; ++u) {
      const int j = hd[3 * g + u];
      XY[u] = make_double2(0.0, 0.0); NK[u] = make_longlong2(-1LL, 0LL);
      if (j >= 0) {
        XY[u] = __ldcg(reinterpret_cast<const double2 *>(&crec[j].x));
        NK[u] = __ldcg(reinterpret_cast<const longlong2 *>(&crec[j].next));
      }
    }
#pragma unroll
    for (int u = 0; u < 3; ++u) {
      int j = hd[3 * g + u];
      if (j < 0) continue;
      consider(j, XY[u].x, XY[u].y, (int)((unsigned long long)NK[u].x >> 32), NK[u].y);
      j = (int)(unsigned int)NK[u].x;
      while (j >= 0) {   // the rest of the chain (half a particle per cell on average)
        const double2 xy = __ldcg(reinterpret_cast<const double2 *>(&crec[j].x));
        const longlong2 nk = __ldcg(reinterpret_cast<const longlong2 *>(&crec[j].next));
        consider(j, xy.x, xy.y, (int)((unsigned long long)nk.x >> 32), nk.y);
        j = (int)(unsigned int)nk.x;
      }
    }
  }
  for (int k = 0; k < cnt; ++k) T.nbr[(long long)k * T.n_pad + slot] = loc[k];
  T.nbr_cnt[slot] = cnt;
  T.xr[slot] = xi; T.yr[slot] = yi;
  return cnt;   // overflow / largest count: reduced per CTA by the caller (100k threads polling one L2 word is a hot spot)
}

// Halo warp: the records of epoch flag `flag` of every remote slot of the tile -> halo region (sxh, syh) of one position
// buffer.  Four records per lane are in flight at a time; an incomplete one is polled again.  Not inlined: its registers
// (16 64-bit words in flight) must not count against the particle warps' allocation.
__device__ __noinline__ void tile_halo_load(const RimRec *rim, unsigned int flag, const int *htab, const int *hlist, int nh, int hlane,
                                            double *sxh, double *syh) {
  for (int c0 = hlane; c0 < nh; c0 += 128) {
    unsigned long long w[4][4];
    int pos[4];
    bool pend[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int c = c0 + 32 * u;
      pend[u] = c < nh;
      pos[u] = pend[u] ? hlist[c] : 0;
    }
    bool any = true;
    while (any) {
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (pend[u]) {
          const RimRec *r = &rim[htab[pos[u]]];
          asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(w[u][0]), "=l"(w[u][1]) : "l"(r) : "memory");
          asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(w[u][2]), "=l"(w[u][3]) : "l"(&r->w[2]) : "memory");
        }
      any = false;
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (pend[u]) {
          const unsigned int bad = ((unsigned int)w[u][0] ^ flag) | ((unsigned int)w[u][1] ^ flag) | ((unsigned int)w[u][2] ^ flag) |
                                   ((unsigned int)w[u][3] ^ flag);
          if (bad == 0u) {
            sxh[pos[u]] = __hiloint2double((int)(w[u][1] >> 32), (int)(w[u][0] >> 32));
            syh[pos[u]] = __hiloint2double((int)(w[u][3] >> 32), (int)(w[u][2] >> 32));
            pend[u] = false;
          } else {
            any = true;
          }
        }
      if (any) __nanosleep(100);
    }
  }
}

template <bool NOISE>
__global__ void __launch_bounds__(SW_RES_MAXT, 1) k_bd2d_tiled(const __grid_constant__ DevParams P, DevArrays A, DevArrays T, TileArgs G,
                                                        RunArgs R) {
  Ctl *ctl = A.ctl;
  if (ldcg(&ctl->halted)) return;
  unsigned int epoch = 0;                  // grid-barrier arrivals expected so far
  const int ptid = (int)threadIdx.x;       // one particle per thread; the LAST warp of the CTA is the halo warp
  const bool hw = threadIdx.x >= blockDim.x - 32;
  const int hlane = (int)threadIdx.x - ((int)blockDim.x - 32);
  const int ppc = G.ppc;
  const int base = (int)blockIdx.x * ppc;
  const long long slot = (long long)base + ptid;
  const bool act = !hw && ptid < ppc && slot < A.n;
  const bool leader = act && slot == 0;

  // dynamic shared memory (SW_TILE_DSM_DOUBLES)
  extern __shared__ double dsm[];
  double (*sx)[SW_TILE_W] = (double (*)[SW_TILE_W])dsm;                        // [2][tile + halo]: positions by step parity
  double (*sy)[SW_TILE_W] = (double (*)[SW_TILE_W])(dsm + 2 * SW_TILE_W);      // [2][tile + halo]
  double *col = dsm + 4 * SW_TILE_W;                                           // per-thread columns
  double *sxr = col, *syr = col + SW_TILE_MAXP, *sthr = col + 2 * SW_TILE_MAXP;  // checkpoint = state at the last rebuild
  double *sfx = col + 3 * SW_TILE_MAXP, *sfy = col + 4 * SW_TILE_MAXP;           // constant external force
  double *swx = col + 5 * SW_TILE_MAXP, *swy = col + 6 * SW_TILE_MAXP;           // wall force of the coming step
  double *sgk = col + 7 * SW_TILE_MAXP;                                          // cosine normal kept from the even step (local_inputs)
  int *snb = (int *)(col + 8 * SW_TILE_MAXP);    // [SW_TILE_SNBR][SW_TILE_MAXP]: list entries beyond the register-resident ones
  int *htab = snb + SW_TILE_SNBR * SW_TILE_MAXP; // [SW_TILE_HT]: remote slot at each halo position, -1 = free
  int *hlist = htab + SW_TILE_HT;                // [s_nh]: the occupied halo positions, compact
  __shared__ int s_nh;                           // halo records of this tile
  __shared__ Bd2dSpecies S[SW_S];
  __shared__ int s_viol;                  // validation verdict (grid-wide), broadcast to the CTA
  __shared__ float s_dm;                  // grid-wide largest displacement^2 of the period
  __shared__ unsigned int s_wmax[SW_RES_MAXT / 32];
  __shared__ int s_rim;                   // rim particles of this CTA at the last rebuild (statistics)
  __shared__ int s_cmax;                  // longest Verlet list of this CTA

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
  __shared__ unsigned long long s_stepbar;
  if (threadIdx.x == 0) tile_mbar_init(&s_stepbar, blockDim.x >> 5);   // one arrival per warp
  unsigned int stepbar_parity = 0u;
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
  if (threadIdx.x == 0) { s_viol = 0; s_rim = 0; s_nh = 0; s_cmax = 0; }
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
  bool keep_ok = false;   // sgk[ptid] holds the cosine normal of the even step just before `step` (uniform over the grid)
  auto local_inputs = [&](u64 step) {
    if (NOISE) {
      // an odd step right after its even partner takes the short path (one Philox block, one Box-Muller pair); anything
      // else -- even steps, a launch that starts on an odd step, the first step after a rollback -- draws both.  The kept
      // normal lives in this thread's shared-memory slot, not in a register pair the pair phase would have to spill.
      double keep = 0.0;
      if (keep_ok && (step & 1ULL)) { keep = sgk[ptid]; swm_step_inputs<true>(P.rk, pid, step, th, g0, g1, g3, keep, ey, ex); }
      else { swm_step_inputs<false>(P.rk, pid, step, th, g0, g1, g3, keep, ey, ex); sgk[ptid] = keep; }
      keep_ok = !(step & 1ULL);
    }
    else swm_sincos(th, ey, ex);
    if (P.walls) {
      const Bd2dSelf sf = bd2d_self_forces(P, ti, xi, yi, ex, ey, fsw);
      swx[ptid] = sf.wx; swy[ptid] = sf.wy;
    }
  };
  if (act) local_inputs(R.step0);

  int cnt = 0, nb[SW_RES_NBR];
  bool exported = false;   // rim particle: its list points into another CTA's tile (and, lists being symmetric, is pointed at from there)
  // A list entry of the step loop: type in the top bits (as in the list), below them an index into the position
  // buffers -- own tile [0, ppc), halo [SW_TILE_MAXP, SW_TILE_W) -- or, flagged SW_TILE_RAW, a remote slot that did not
  // fit the halo table.
  auto translate = [&](int en) -> int {
    const int j = en & SW_TILE_JMASK;
    const int keep = en & (~SW_IDX_MASK | SW_TILE_WRAP);   // type and the minimum-image flag travel with the entry
    const unsigned int l = (unsigned int)(j - base);
    if (l < (unsigned int)ppc) return keep | (int)l;
    exported = true;
    unsigned int h = ((unsigned int)j * 2654435761u) >> 22;   // 10 bits
#pragma unroll 1
    for (int probe = 0; probe < 64; ++probe) {
      const int old = atomicCAS(&htab[h], -1, j);
      if (old == -1 || old == j) return keep | (SW_TILE_MAXP + (int)h);
      h = (h + 1u) & (SW_TILE_HT - 1);
    }
    return keep | SW_TILE_RAW | j;
  };
  // the same mapping for a list entry read from memory inside the step loop (beyond the cached ones): lookup only
  auto translate_ro = [&](int en) -> int {
    const int j = en & SW_TILE_JMASK;
    const int keep = en & (~SW_IDX_MASK | SW_TILE_WRAP);
    const unsigned int l = (unsigned int)(j - base);
    if (l < (unsigned int)ppc) return keep | (int)l;
    unsigned int h = ((unsigned int)j * 2654435761u) >> 22;
#pragma unroll 1
    for (int probe = 0; probe < 64; ++probe) {
      const int v = htab[h];
      if (v == j) return keep | (SW_TILE_MAXP + (int)h);
      if (v == -1) break;
      h = (h + 1u) & (SW_TILE_HT - 1);
    }
    return keep | SW_TILE_RAW | j;
  };
  auto load_list = [&]() {
    cnt = T.nbr_cnt[slot];
    exported = false;
#pragma unroll
    for (int k = 0; k < SW_RES_NBR; ++k) nb[k] = (k < cnt) ? translate(T.nbr[(long long)k * T.n_pad + slot]) : (int)ptid;
    for (int k = SW_RES_NBR; k < cnt; ++k) {
      const int en = translate(T.nbr[(long long)k * T.n_pad + slot]);
      if (k < SW_RES_NBR + SW_TILE_SNBR) snb[(k - SW_RES_NBR) * SW_TILE_MAXP + ptid] = en;
    }
  };
  // halo warp: compact list of the occupied table positions (after every thread's load_list, behind a CTA barrier)
  auto halo_compact = [&]() {
    int n = 0;
    for (int h0 = 0; h0 < SW_TILE_HT; h0 += 32) {
      const bool occ = htab[h0 + hlane] >= 0;
      const unsigned int b = __ballot_sync(0xffffffffu, occ);
      if (occ) hlist[n + __popc(b & ((1u << hlane) - 1u))] = h0 + hlane;
      n += __popc(b);
    }
    if (hlane == 0) s_nh = n;
    __syncwarp();
  };
#pragma unroll
  for (int k = 0; k < SW_RES_NBR; ++k) nb[k] = act ? ptid : 0;
  double vx_l = 0.0, vy_l = 0.0;
  bool viol = false;        // a position of this thread used for forces since the last rebuild was beyond skin / 2
  double dmax2 = 0.0;       // its largest displacement^2 in the period

#ifdef SW_TIMING
  // clock64 per phase of CTA 0: slots 0..5 from the first particle thread, 6..7 from the flag thread (swarm_debug_timing)
  long long tp[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tc = clock64();
  const bool prof = ((int)blockIdx.x == R.prof_cta) && ((int)threadIdx.x == 32 * R.prof_warp);   // warp 22 = the halo warp
#endif
  long long s = 0;          // steps completed in this launch
  long long s_ck = 0;       // step index of the checkpoint
  bool ck_start = true;     // checkpoint = launch start (state still in the caller's arrays)
  unsigned int e = 0;               // step epochs executed in this launch, replays included: buffer parity, record flag G.flag0 + e
  bool have_partners = false;       // the rim records of the current epoch have been published
  unsigned int nrb = 0;             // rebuilds started in this launch (tag of the linked cells)
  bool stop = false;

  // The loop below is instantiated TWICE: for the particle warps and for the halo warp (same source, hence the same
  // sequence of CTA and grid barriers and the same uniform decisions).  In the halo warp's copy everything that
  // belongs to a particle is compiled out, and in the particle warps' copy the halo loader is -- so neither role's
  // registers are held, spilled or saved around a call for the sake of the other (with one shared loop and the loader
  // as a called function, the director, the noise and theta of every particle lived in local memory).
  auto run = [&](auto role) {
  constexpr bool HALO = decltype(role)::value;
  const bool actv = HALO ? false : act;
  const bool leadv = HALO ? false : leader;
  for (;;) {
    const bool at_end = (s == R.n_steps);
    SW_TICK(0);   // the wait of the step barrier lands here (BAR.SYNC defers its blocking to the next use)
    if (need || since >= K || at_end || !have_partners) {
      // ================= synchronisation point: validate the period, then rebuild ================================
      const int cur = (int)(e & 1u);
      const bool do_vote = (s > s_ck);                                        // steps ran on the current list
      const bool maybe = !at_end && (do_vote || need || since >= K);          // a rebuild may follow (the vote may extend the period instead)
      const unsigned long long tag = G.tag0 + (unsigned long long)nrb;
      if (maybe) {
        // publish every position and link the particles into their cells now: the barrier of the vote (or a plain one)
        // then separates the links from the search.  If the rebuild does not happen after all (vote failed, or the
        // period was extended) the links stay behind with a tag no later rebuild uses.
        ++nrb;
        if (actv) tile_insert(P, T, G.crec, G.head64, tag, slot, xi, yi, sp, (long long)pid);
      }
      if (do_vote) {
        double d2n = 0.0;
        if (actv) { const double ddx = xi - sxr[ptid], ddy = yi - syr[ptid]; d2n = ddx * ddx + ddy * ddy; }
        const bool bad = vote(viol, fmax(dmax2, d2n));
        const double dm = (double)s_dm;
        // How long does this list last?  Linear extrapolation of the largest displacement to 0.9 * skin / 2 (exact for
        // drift, conservative for diffusion, whose largest displacement grows like the square root of the time).
        double ratio = (dm > 0.0) ? sqrt(lim2 / dm) : 4.0;
        if (ratio > 4.0) ratio = 4.0;
        long long Kn = (long long)((double)(since > 0 ? since : 1) * ratio * G.k_safety);
        if (bad) {
          if (Kn > K / 2) Kn = K / 2;
          K = (int)(Kn < 1 ? 1 : Kn);
          // ---- roll the period back to its checkpoint and replay it with the shorter cadence -------------------
          if (ck_start) {
            if (actv) { xi = A.x[0][id]; yi = A.y[0][id]; th = A.theta[id]; }
            since = 0; need = true;
          } else {
            if (actv) { xi = sxr[ptid]; yi = syr[ptid]; th = sthr[ptid]; }
            since = 0; need = false;
          }
          s = s_ck;
          keep_ok = false;   // the kept normal belongs to the step pair that was abandoned
          viol = false; dmax2 = 0.0;
          if (actv) {
            sx[cur][ptid] = xi; sy[cur][ptid] = yi;
            if (exported) { rim_write(&(G.rim + cur * G.rim_stride)[slot], G.flag0 + e, xi, yi); rim_write(&(G.rim + (cur ^ 1) * G.rim_stride)[slot], G.flag0 + e, xi, yi); }
            local_inputs(R.step0 + (u64)s);
          }
          if (leadv) ctl->tile_rollbacks += 1;
          grid_barrier64(&ctl->bar64, epoch);   // nobody steps before every CTA has restored (and republished) its state
          if constexpr (HALO) tile_halo_load(G.rim + cur * G.rim_stride, G.flag0 + e, htab, hlist, s_nh, hlane, &sx[cur][SW_TILE_MAXP], &sy[cur][SW_TILE_MAXP]);
          __syncthreads();   // the halo of the current epoch is in shared memory
          continue;
        }
        if (Kn > G.k_cap) Kn = G.k_cap;
        if (since >= K) {
          // scheduled point: extend the period when the list is good for at least a quarter more (one more vote, no
          // rebuild); otherwise rebuild now and give the next period this one's length (a little more if there was slack)
          const long long room = since / 4 > 4 ? since / 4 : 4;
          if (Kn >= since + room && since + room <= G.k_cap) K = (int)(Kn < 2LL * since ? Kn : 2LL * since);
          else {
            const long long up = (long long)since + since / 8 + 1;
            Kn = Kn < up ? Kn : up;
            K = (int)(Kn < 1 ? 1 : Kn);
            need = true;
          }
        } else if (Kn < K) {
          K = (int)(Kn < 1 ? 1 : Kn);   // launch end inside a period: only ever shorten
        }
      } else if (maybe) {
        grid_barrier64(&ctl->bar64, epoch);
      }
      SW_TICK(2);   // links + vote (grid barrier)
      if (at_end) break;
      if (need || since >= K) {
        // ---- the lists at the current positions ------------------------------------------------------------------
        for (int h = threadIdx.x; h < SW_TILE_HT; h += blockDim.x) htab[h] = -1;
        __syncthreads();
        bool over = false;
        if (actv) {
          int loc[SW_MAXCAP];
          long long lkey[SW_MAXCAP];
          const int c = tile_search(P, T, G.crec, G.head64, tag, slot, xi, yi, ti, loc, lkey, over);
          if (c > s_cmax) atomicMax(&s_cmax, c);
          sxr[ptid] = xi; syr[ptid] = yi; sthr[ptid] = th;
          cnt = c;
          exported = false;
#pragma unroll
          for (int k = 0; k < SW_RES_NBR; ++k) nb[k] = (k < c) ? translate(loc[k]) : (int)ptid;
          for (int k = SW_RES_NBR; k < c; ++k) {
            const int en = translate(loc[k]);
            if (k < SW_RES_NBR + SW_TILE_SNBR) snb[(k - SW_RES_NBR) * SW_TILE_MAXP + ptid] = en;
          }
          // the rim may have changed: everybody on it republishes -- and marks the record of the OTHER parity with this
          // epoch's flag, which no reader of that slot ever waits for (a slot left over from an earlier launch could
          // otherwise carry, after a wrap of the 32-bit flag, exactly the flag of the next epoch)
          if (exported) { rim_write(&(G.rim + cur * G.rim_stride)[slot], G.flag0 + e, xi, yi); rim_write(&(G.rim + (cur ^ 1) * G.rim_stride)[slot], G.flag0 + e, xi, yi); }
        }
        SW_TICK(6);   // search + translation + rim records
        const int any_over = __syncthreads_or(over);
        const int nrim = __syncthreads_count(actv && exported);
        if (threadIdx.x == 0) {
          s_rim = nrim;
          atomicAdd((unsigned long long *)&ctl->tile_partners, (unsigned long long)nrim);
          if (any_over) atomicExch(&ctl->overflow, 1);
          if (s_cmax > ldcg(&ctl->max_nbr_seen)) atomicMax(&ctl->max_nbr_seen, s_cmax);
        }
        if (leadv) ctl->rebuilds += 1;
        if constexpr (HALO) halo_compact();
        grid_barrier64(&ctl->bar64, epoch);
        if (threadIdx.x == 0) s_viol = ldcg(&ctl->overflow);   // one reader per CTA, broadcast through shared memory
        __syncthreads();
        if (s_viol) { stop = true; break; }   // uniform
        if constexpr (HALO) tile_halo_load(G.rim + cur * G.rim_stride, G.flag0 + e, htab, hlist, s_nh, hlane, &sx[cur][SW_TILE_MAXP], &sy[cur][SW_TILE_MAXP]);
        __syncthreads();   // the halo of the current epoch is in shared memory
        SW_TICK(7);   // CTA barrier, halo list, grid barrier, halo load
        have_partners = true;
        s_ck = s; ck_start = false;
        since = 0; need = false;
        viol = false; dmax2 = 0.0;
      } else if (have_partners) {
        // the vote extended the period (K grew): lists, rim and records are as they were -- only the halo of this epoch
        // was not fetched, because the previous step expected a synchronisation point
        if constexpr (HALO) tile_halo_load(G.rim + cur * G.rim_stride, G.flag0 + e, htab, hlist, s_nh, hlane, &sx[cur][SW_TILE_MAXP], &sy[cur][SW_TILE_MAXP]);
        __syncthreads();
      } else {
        // list kept from the previous launch (validated by the start vote): classify it and publish the rim
        for (int h = threadIdx.x; h < SW_TILE_HT; h += blockDim.x) htab[h] = -1;
        __syncthreads();
        if (actv) {
          load_list();
          if (exported) { rim_write(&(G.rim + cur * G.rim_stride)[slot], G.flag0 + e, xi, yi); rim_write(&(G.rim + (cur ^ 1) * G.rim_stride)[slot], G.flag0 + e, xi, yi); }
        }
        __syncthreads();
        if constexpr (HALO) halo_compact();
        grid_barrier64(&ctl->bar64, epoch);
        if constexpr (HALO) tile_halo_load(G.rim + cur * G.rim_stride, G.flag0 + e, htab, hlist, s_nh, hlane, &sx[cur][SW_TILE_MAXP], &sy[cur][SW_TILE_MAXP]);
        __syncthreads();   // the halo of the current epoch is in shared memory
        have_partners = true;
      }
    }
    SW_TICK(3);   // synchronisation points (validation, rebuild, halo refresh)
    // ================= one Brownian step: read buffer cur, write buffer nxt ==========================================
    const int cur = (int)(e & 1u), nxt = cur ^ 1;
    const bool last = (s == R.n_steps - 1);
    if (actv) {
      {
        // the position used for this force evaluation must be within skin / 2 of the list's reference
        const double ddx = xi - sxr[ptid], ddy = yi - syr[ptid];
        const double d2 = ddx * ddx + ddy * ddy;
        viol = viol || !(d2 <= lim2);
        dmax2 = fmax(dmax2, d2);
      }
      // every entry is an index into the position buffer (own tile, then halo): plain LDS.  A raw remote entry (halo
      // table full) is the exception and reads its record itself.
      auto fetch = [&](int en, double &x, double &y) {
        if (en & SW_TILE_RAW) rim_read(&(G.rim + cur * G.rim_stride)[en & SW_TILE_JMASK], G.flag0 + e, x, y);
        else { const int l = en & SW_TILE_JMASK; x = sx[cur][l]; y = sy[cur][l]; }
      };
      double xj[SW_RES_NBR], yj[SW_RES_NBR];
#pragma unroll
      for (int k = 0; k < SW_RES_NBR; ++k) fetch(nb[k], xj[k], yj[k]);
      double Fx = 0.0, Fy = 0.0;
#ifndef SW_TILE_PAIR4
#pragma unroll
      for (int k = 0; k < SW_RES_NBR; ++k)
        if (k < cnt) bd2d_pair_w(P, xi, yi, ti, nb[k] >> SW_TYPE_SHIFT, (nb[k] & SW_TILE_WRAP) != 0, xj[k], yj[k], Fx, Fy);
#else
      static_assert(SW_RES_NBR == 4, "bd2d_pair4 takes the four register-resident entries");
      bd2d_pair4(P, xi, yi, ti, nb, xj, yj, cnt, Fx, Fy);
#endif
      for (int k0 = SW_RES_NBR; k0 < cnt; k0 += 4) {
        int en4[4];
        double xe[4], ye[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          // a clustered particle's further neighbours: shared memory too (the first SW_TILE_SNBR of them)
          const int k = k0 + u;
          en4[u] = (k >= cnt) ? ptid
                              : (k < SW_RES_NBR + SW_TILE_SNBR ? snb[(k - SW_RES_NBR) * SW_TILE_MAXP + ptid]
                                                               : translate_ro(T.nbr[(long long)k * T.n_pad + slot]));
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) fetch(en4[u], xe[u], ye[u]);
#ifndef SW_TILE_PAIR4
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (k0 + u < cnt) bd2d_pair_w(P, xi, yi, ti, en4[u] >> SW_TYPE_SHIFT, (en4[u] & SW_TILE_WRAP) != 0, xe[u], ye[u], Fx, Fy);
#else
        bd2d_pair4(P, xi, yi, ti, en4, xe, ye, cnt - k0, Fx, Fy);
#endif
      }
      // walls, swim, external: added after the pair forces, in the oracle's order
      if (P.walls) { Fx += swx[ptid]; Fy += swy[ptid]; }
      Fx += __dmul_rn(fsw, ex); Fy += __dmul_rn(fsw, ey);   // product rounded on its own, as in the other kernels
      if (A.fex) { Fx += sfx[ptid]; Fy += sfy[ptid]; }
      const Bd2dSpecies &c = S[sp];
      const Bd2dConst cc = {c.gt1, c.gt2, c.r_gt1, c.r_gt2, c.sg_par, c.sg_per, c.sg_rot};
      bd2d_update<NOISE>(iso, cc, dt, Fx, Fy, ex, ey, w, g0, g1, g3, xi, yi, th, vx_l, vy_l);
      sx[nxt][ptid] = xi; sy[nxt][ptid] = yi;
      if (exported) rim_write(&(G.rim + nxt * G.rim_stride)[slot], G.flag0 + e + 1u, xi, yi);
    }
    SW_TICK(1);   // pair forces + update
#ifdef SW_TILE_MBAR
    if constexpr (!HALO) tile_mbar_arrive_warp(&s_stepbar);   // this warp's positions of the next step are in shared memory
#endif
    if (actv && !last) local_inputs(R.step0 + (u64)(s + 1));   // step s+1's particle-local work: the rim records travel meanwhile
    // the next iteration is a synchronisation point: it republishes and reloads the halo behind its grid barrier
    const bool next_sync = (since + 1 >= K) || last;
    if constexpr (HALO) if (!next_sync) tile_halo_load((G.rim + nxt * G.rim_stride), G.flag0 + e + 1u, htab, hlist, s_nh, hlane, &sx[nxt][SW_TILE_MAXP], &sy[nxt][SW_TILE_MAXP]);
#ifdef SW_TILE_MBAR
    if constexpr (HALO) tile_mbar_arrive_warp(&s_stepbar);    // the halo of the next step is in shared memory
#endif
    SW_TICK(4);   // particle-local work of the next step / halo warp: the neighbours' records
#ifndef SW_TILE_MBAR
    __syncthreads();   // the positions of the next step -- tile and halo -- are in shared memory
#else
    tile_mbar_wait(&s_stepbar, stepbar_parity);   // the positions of the next step -- tile and halo -- are in shared memory
    stepbar_parity ^= 1u;
#endif
    SW_TICK(5);
    ++s; ++since; ++e;
  }
  };
  if (hw) run(std::true_type()); else run(std::false_type());
#ifdef SW_TIMING
  if (prof) {
    for (int k = 0; k < 8; ++k) ctl->tprof[k] += tp[k];
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
