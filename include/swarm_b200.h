/*
 * swarm_b200.h -- C-ABI of libswarm_b200.so, the B200-native (sm_100a) engine for
 * the SwarmRL rollout hot path.
 *
 * The reference reaches its physics through the Python/Cython handles of
 * `espressomd` (a third-party C++ package); every entry point below cites the
 * reference call site it replaces (paths relative to /root/reference).  All
 * pointers named *_dev are CUDA device pointers owned by the caller (torch
 * tensors on the Python side, exchanged zero-copy); *_host are host pointers.
 * No call throws; every call returns 0 on success or a negative swarm_status_t
 * and leaves a message for swarm_last_error().  Calls that take a `stream`
 * only enqueue work on that cudaStream_t (passed as void*) and do not
 * synchronise unless stated.
 *
 * State layout (HBM, structure of arrays, fp64, one element per particle,
 * particle index == particle id == creation order, as in ESPResSo):
 *   x,y,z            unfolded positions           (espresso.py:1122,1333)
 *   theta            2-D orientation angle; director = (cos, sin, 0)
 *   q0..q3           3-D orientation quaternion; director = body z axis
 *   vx..vz, wx..wz   velocity / lab angular velocity, materialised at the end
 *                    of every swarm_run chunk
 *   f_swim           Action.force  (espresso.py:1228)
 *   tqx..tqz         Action.torque (espresso.py:1229-1235), lab frame
 *   fex..fez         ext_force     (espresso.py:834-851); may be NULL
 *   species          int32 index into the species table (see swarm_set_species)
 */
#ifndef SWARM_B200_H
#define SWARM_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SWARM_MAX_SPECIES 16
#define SWARM_MAX_TYPES 8
#define SWARM_MAX_NEIGHBORS 128 /* hard ceiling of the Verlet-list capacity (swarm_set_max_neighbors) */
#define SWARM_ABI_VERSION 1

typedef enum {
  SWARM_OK = 0,
  SWARM_ERR_INVALID = -1,   /* bad argument (the reference raises ValueError) */
  SWARM_ERR_STATE = -2,     /* call order violated (the reference raises RuntimeError) */
  SWARM_ERR_CUDA = -3,      /* CUDA runtime error; message holds cudaGetErrorString */
  SWARM_ERR_OVERFLOW = -4,  /* a fixed-capacity device structure overflowed */
  SWARM_ERR_NOMEM = -5
} swarm_status_t;

typedef struct swarm swarm_t;
typedef enum { SWARM_DECAY_NONE = 0, SWARM_DECAY_LINEAR = 1, SWARM_DECAY_EXP = 2, SWARM_DECAY_INVERSE = 3 } swarm_decay_t;

/* espresso.py:143-152 (n_dims, seed), :267-270 (box, time_step, skin, periodicity),
 * :1170-1186 (kT, thermostat seed). */
typedef struct {
  int32_t n_dims;        /* 2 or 3 */
  int32_t periodic[3];
  double box[3];
  double dt;
  double kT;
  double skin;           /* Verlet skin; affects performance only (espresso.py:268-269) */
  uint64_t seed;
  int32_t device;        /* CUDA device ordinal */
  int32_t max_neighbors; /* Verlet-list capacity per particle; 0 = default (16 in 2-D, 64 in 3-D) */
  int32_t n_envs;        /* 1 = one swarm; >1 = that many independent environments of n/n_envs
                            consecutive particles each (ensemble batching, BASELINE config 3) */
  int32_t flags;         /* SWARM_FLAG_* */
} swarm_config_t;

#define SWARM_FLAG_NO_PERSISTENT 1 /* one launch per step instead of the cooperative kernel */

typedef struct {
  double *x, *y, *z;
  double *theta;
  double *q0, *q1, *q2, *q3;
  double *vx, *vy, *vz;
  double *wx, *wy, *wz;
  int32_t *species;
} swarm_state_dev_t;

typedef struct {
  double *f_swim;
  double *tqx, *tqy, *tqz;
  double *fex, *fey, *fez;     /* may be NULL */
  double *ndx, *ndy, *ndz;     /* Action.new_direction; may be NULL */
  int32_t *has_nd;             /* 1 where new_direction is set; may be NULL */
} swarm_actions_dev_t;

typedef struct {
  int64_t steps_done;      /* integrator steps completed by the last swarm_run / swarm_relax */
  int64_t steps_requested;
  int64_t rebuilds;        /* neighbour-list rebuilds since creation */
  int64_t launches;        /* kernels launched by this handle since creation */
  int32_t overflow;        /* 1 if the Verlet list capacity was exceeded (run halted early) */
  int32_t max_neighbors_seen;
  int32_t capacity;        /* current Verlet capacity */
  int32_t persistent;      /* 1 if the cooperative persistent kernel is in use */
} swarm_stats_t;

const char *swarm_last_error(void);
int swarm_abi_version(void);

/* espressomd.System(...) + _init_system  (espresso.py:193, 236-288) */
int swarm_create(const swarm_config_t *cfg, swarm_t **out);
int swarm_destroy(swarm_t *s);

/* system.part.add(gamma=, gamma_rot=, mass=, rinertia=, type=) (espresso.py:416-441):
 * one species per distinct property set.  gamma_t / gamma_r / rinertia are body-frame
 * diagonals, 3 per species. */
int swarm_set_species(swarm_t *s, int32_t n_species, const int32_t *type, const double *gamma_t,
                      const double *gamma_r, const double *mass, const double *rinertia);

/* non_bonded_inter[a,b].wca.set_params(sigma, epsilon)  (espresso.py:810-819); row-major T x T.
 * cut = sigma * 2^(1/6). */
int swarm_set_wca(swarm_t *s, int32_t n_types, const double *sigma, const double *eps, const double *cut);

/* constraints.add(ShapeBasedConstraint(Wall...)) on the box faces (espresso.py:667-711).  The three arrays hold
 * n_types entries (wall interaction of every particle type): call swarm_set_wca FIRST, it fixes n_types --
 * otherwise SWARM_ERR_STATE. */
int swarm_set_walls(swarm_t *s, int32_t enabled, const double *sigma, const double *eps, const double *cut);

/* thermostat.set_brownian(kT, seed) (espresso.py:1180-1186) -- may be changed between runs */
int swarm_set_thermostat(swarm_t *s, double kT, uint64_t seed);

/* MDParams.thermostat_type (espresso.py:62, 1173-1201): "brownian" = thermostat.set_brownian +
 * integrator.set_brownian_dynamics (the default, :1179-1186); "langevin" = thermostat.set_langevin +
 * integrator.set_vv (:1187-1201, without a lattice-Boltzmann fluid).  Under LANGEVIN swarm_run advances
 * velocity Verlet steps: v and omega of the bound state are INPUTS as well as outputs, every species
 * needs mass > 0 and rinertia > 0 (espresso.py:391-400; isotropic rinertia in 3-D), every run starts
 * with a force recalculation at the current (x, v) as ESPResSo does after particle properties change.
 * On a decomposed swarm the caller ships vx, vy, wz as aux channels of the exchange record and cuts runs with swarm_run_part. */
typedef enum { SWARM_INTEGRATOR_BROWNIAN = 0, SWARM_INTEGRATOR_LANGEVIN = 1 } swarm_integrator_t;
int swarm_set_integrator(swarm_t *s, int32_t kind);

/* Bind caller-owned device arrays (no copy).  n must be a multiple of n_envs. */
int swarm_bind_state(swarm_t *s, int64_t n, const swarm_state_dev_t *st);
int swarm_bind_actions(swarm_t *s, const swarm_actions_dev_t *act);

/* manage_forces' new_direction branch (espresso.py:1236-1249): 3-D overwrite the director,
 * 2-D rotate in plane by acos(new . old) when > 1e-6, about sign((old x new)_z). */
int swarm_apply_new_direction(swarm_t *s, void *stream);

/* _remove_overlap: integrator.set_steepest_descent(f_max=0, gamma, max_displacement); run(n_iter)
 * (espresso.py:1163-1167).  Simulation time is not advanced. */
int swarm_relax(swarm_t *s, int64_t n_iter, double gamma, double max_disp, void *stream);

/* integrator.run(n_steps)  (espresso.py:1304-1306) with the Brownian integrator: WCA + swim +
 * external force/torque + Philox noise, fused.  step0 = global index of the first step (the
 * Philox counter).  Enqueues only; call swarm_sync_stats to learn whether it ran to completion. */
int swarm_run(swarm_t *s, int64_t n_steps, uint64_t step0, void *stream);

/* One integrator.run cut in parts, for the Langevin integrator on a decomposed swarm (no reference counterpart: ESPResSo
 * runs single-rank; the domain exchange sits between the parts).  A velocity-Verlet chunk of n steps is n + 1 force
 * evaluations, the first a recalculation and the last closing the final step; SWARM_RUN_LEAVE_OPEN stops a part between
 * the two half kicks of its last step (positions at t_n, velocities at t_n - dt/2: the state the exchange ships) and
 * SWARM_RUN_CONTINUE starts the next part by closing that step, so the parts together are the uncut run, bit for bit.
 * Flags are ignored by the Brownian integrator, where swarm_run_part == swarm_run. */
#define SWARM_RUN_CONTINUE 1
#define SWARM_RUN_LEAVE_OPEN 2
int swarm_run_part(swarm_t *s, int64_t n_steps, uint64_t step0, int32_t flags, void *stream);

/* Synchronises `stream` and reports counters of the last run. */
int swarm_sync_stats(swarm_t *s, void *stream, swarm_stats_t *out);

/* Profiling hook: clock64 sums of the resident kernel's phases since the last call (all zero unless the
 * library was built with -DSW_TIMING, `python -m swarmrl_b200.build --timing`); synchronises the device. */
int swarm_debug_timing(swarm_t *s, int64_t out8[8]);

/* Spatial tiling of the 2-D Brownian chunk (csrc/swarm_tiled.cuh, csrc/swarm_sort.cuh).  ESPResSo stores its
 * particles by cell of the regular decomposition selected at espresso.py:267-270; here the library keeps a spatially
 * sorted slot order (cell hash, counting sort, cell-start scan) beside the caller's id-indexed arrays and gives every
 * CTA a compact tile.  mode: 0 = off, 1 = automatic (swarms of >= 4096 particles; the default), 2 = whenever the swarm
 * fits one particle per thread.  resort_interval: launches between two sorts (<= 0 keeps the current value).
 * Trajectories do not depend on any of this (forces are summed in ascending particle id, noise is keyed by id). */
int swarm_set_tiling(swarm_t *s, int32_t mode, int32_t resort_interval);
/* out8: [0] 1 if the Verlet list currently lives in the tiled kernel's slot space (the last Brownian run was tiled),
 * [1] rebuild cadence K, [2] steps integrated on the current list, [3] periods rolled back so far, [4] sum of
 * rim particles (particles whose list points into another CTA's tile) over all rebuilds, [5] spatial sorts so far, [6] CTAs, [7] slots per CTA.  Synchronises the device. */
int swarm_tiling_stats(swarm_t *s, int64_t out8[8]);
/* The tiled kernel's OWN Verlet list as it stands (built at the last in-kernel rebuild, slots mapped back to particle
 * ids), every ordered pair sorted by (i,j) -- for the bit-exact list check of that path.  SWARM_ERR_STATE unless the
 * last run was tiled. */
int swarm_tiled_neighbors_dump(swarm_t *s, int32_t *pairs_host, int64_t cap, int64_t *count);

/* Grow the Verlet capacity after an overflow (invalidates the current list). */
int swarm_set_max_neighbors(swarm_t *s, int32_t max_neighbors);

/* director of every particle (particle.director), 3 arrays of n */
int swarm_directors(swarm_t *s, double *ex_dev, double *ey_dev, double *ez_dev, void *stream);

/* Neighbour-structure introspection for the bit-exact list checks (synchronous, host out).
 * Rebuilds the Verlet list at the current positions, then writes every ordered pair (i,j)
 * with r_ij < cut(type_i,type_j) + skin, sorted by (i,j).  *count = number of pairs. */
int swarm_neighbors_dump(swarm_t *s, int32_t *pairs_host, int64_t cap, int64_t *count);
/* cell index of every particle on a grid of nc[] cells over the box (the cell hash) */
int swarm_cell_index(swarm_t *s, const int32_t nc[3], int32_t *cell_host);
/* total force on every particle at the current state (debug / parity), device out */
int swarm_forces(swarm_t *s, double *fx_dev, double *fy_dev, double *fz_dev, void *stream);

/* Parity hooks of the noise (ESPResSo: thermostat RNG seeded at espresso.py:1184).  swarm_philox_host: one
 * Philox4x64-10 block (the policy's Gumbel noise).  swarm_noise_words_host: the thermostat's four 64-bit words of
 * (seed, id, step, stream) -- two Philox4x32-10 blocks, layout in csrc/swarm_common.cuh -- pure host code.
 * swarm_gaussians: the 4 normals Box-Muller makes of those words, per id 0..n-1, on the device. */
int swarm_philox_host(const uint64_t ctr[4], const uint64_t key[2], uint64_t out[4]);
int swarm_noise_words_host(uint64_t seed, uint64_t id, uint64_t step, uint64_t stream_id, uint64_t out[4]);
int swarm_gaussians(swarm_t *s, int64_t n, uint64_t step, uint64_t stream_id, double *out4n_dev, void *stream);

/* Scratch capacity: later swarm_bind_state calls with n <= n_max keep the library's scratch (Verlet
 * lists, cell links) instead of reallocating it -- the domain-decomposed swarm re-binds after every
 * halo exchange.  Call before swarm_bind_state. */
int swarm_reserve(swarm_t *s, int64_t n_max);

/* Philox key id of every slot (int64 [n], device).  The reference's thermostat keys its noise by the
 * particle id (espresso.py:1180-1186); slot index == id unless a sub-domain of a larger swarm is bound.
 * Reset to NULL (slot index) by swarm_bind_state. */
int swarm_bind_ids(swarm_t *s, const int64_t *gid_dev);

/* Device-resident particle count (int32, device): the kernels of swarm_run then advance
 * min(*n_dev, n) particles, n being the bound passed to swarm_bind_state.  Persistent kernels only.
 * Reset to NULL by swarm_bind_state. */
int swarm_bind_count(swarm_t *s, const int32_t *n_dev);

/* ---- domain decomposition (no reference counterpart, ESPResSo runs single-rank) ---
 * Slab decomposition along x of a 2-D periodic swarm.  Local frame of a rank: owned slab
 * [halo, halo + slab_w), kept region [0, slab_w + 2 halo); the local swarm_t is created with
 * box = (slab_w + 2 halo, L_y, .), periodic = (0, 1, 0).  Exchange buffers are SoA records of
 * (x, y, theta, f_swim, tqz, gid, species, image, aux...) doubles behind an 8-byte record count. */
#define SWARM_DD_MAX_AUX 8
typedef struct {
  double slab_w;
  double halo;
  int32_t n_aux;             /* channels that travel with a particle: observable histories (two per history), and vx, vy,
                              * omega_z when the Langevin integrator runs on the decomposed swarm */
  int32_t cap;               /* record capacity of one exchange buffer */
  int32_t image_shift_left;  /* added to the periodic image count of records sent left (-1 on rank 0) */
  int32_t image_shift_right; /* ... sent right (+1 on the last rank) */
} swarm_dd_geom_t;

typedef struct {
  double *x, *y, *theta, *f_swim, *tqz;
  int64_t *gid;
  int32_t *species, *image, *owned;
  double *aux[SWARM_DD_MAX_AUX];
} swarm_dd_arrays_t;

/* doubles per exchange buffer for this geometry */
int64_t swarm_dd_buffer_doubles(const swarm_dd_geom_t *g);
/* owned particles within `halo` of a slab face (or beyond it: migrants) -> send buffers, x already in
 * the receiver's frame */
int swarm_dd_pack(swarm_t *s, const swarm_dd_geom_t *g, int64_t n, const int32_t *n_dev, const swarm_dd_arrays_t *loc,
                  double *send_left_dev, double *send_right_dev, void *stream);
/* new local set = old owned + both received buffers, restricted to the kept region; owned recomputed.
 * counts8_dev: [0] new n, [1] owned, [2] error bits (1 buffer overflow, 2 a particle outran the halo,
 * 4 local capacity exceeded), [3] the same bits, sticky across calls (never reset by the library).
 * Every `n` of the dd calls is an upper bound when the matching *_dev count pointer is non-NULL: the live
 * count is then read on the device (counts8_dev[0] of the previous merge), so a rank can queue
 * exchange after exchange without a host synchronisation. */
int swarm_dd_merge(swarm_t *s, const swarm_dd_geom_t *g, int64_t n_old, const int32_t *n_old_dev, const swarm_dd_arrays_t *old_loc,
                   const double *recv_left_dev, const double *recv_right_dev, int64_t cap_new,
                   const swarm_dd_arrays_t *new_loc, int32_t *counts8_dev, void *stream);

/* ---- the exchange over peer memory (NVLink / NVSwitch P2P stores) instead of NCCL send / recv ---------------------
 * No reference counterpart (see above).  Every rank allocates ONE exchange region (4 flags + two parities of its two
 * receive buffers) and publishes its 64-byte CUDA IPC handle to its slab neighbours, which open it once.  Per exchange
 * `epoch` = 1, 2, 3, ...: swarm_dd_pack (local send buffers) -> swarm_dd_push (the live records go into the neighbours'
 * receive buffers of parity epoch & 1 with remote stores, then a system-scope release of flag = epoch) ; the receiver
 * enqueues swarm_dd_wait (one thread polls its own two flags) in front of swarm_dd_merge, which reads
 * swarm_dd_peer_recv(my_region, g, side, epoch & 1).  A flag that does not arrive within 20 s sets the sticky error bit 8
 * of counts8_dev[3] instead of hanging.  All ranks must use the same geometry (buffer capacity and n_aux). */
int64_t swarm_dd_peer_region_bytes(const swarm_dd_geom_t *g);
int swarm_dd_peer_alloc(int32_t device, const swarm_dd_geom_t *g, void **region_out, unsigned char handle64_out[64]);
int swarm_dd_peer_open(int32_t device, const unsigned char handle64[64], void **region_out);
int swarm_dd_peer_close(void *region);
int swarm_dd_peer_free(void *region);
/* side 0: the buffer filled through my left face, 1: through my right face */
double *swarm_dd_peer_recv(void *region, const swarm_dd_geom_t *g, int32_t side, int32_t parity);
int swarm_dd_push(swarm_t *s, const swarm_dd_geom_t *g, const double *send_left_dev, const double *send_right_dev,
                  void *left_neighbour_region, void *right_neighbour_region, uint64_t epoch, void *stream);
int swarm_dd_wait(swarm_t *s, void *my_region, uint64_t epoch, int32_t *counts8_dev, void *stream);

/* Lean exchange bookkeeping: with on = 1 the last kernel of swarm_dd_merge also clears what the NEXT exchange starts from -- the
 * record counts of the send buffers of the last swarm_dd_pack and the counts8 buffer passed as n_old_dev (which must then be
 * the counts8 buffer of the old half, the next merge's output) -- so the next pack / merge skip their four memset launches
 * when they are handed exactly those buffers (any other buffer is cleared as before).  For callers that alternate
 * pack -> transport -> merge on two fixed halves, as DomainSwarm does. */
int swarm_dd_set_lean(swarm_t *s, int32_t on);

/* swarm_obs_field for the owned particles of a sub-domain, on global unfolded coordinates
 * x_global = x_local + x_offset + image * box_x; history (p_prev) in aux channels 0, 1; out_dev[n],
 * 0 for ghosts.  init_only = 1 just seeds the history. */
int swarm_dd_obs_field(swarm_t *s, int64_t n, const int32_t *n_dev, const swarm_dd_arrays_t *loc, double x_offset, double box_x,
                       const double src[3], const double box_length[2], int32_t decay, double c0, double c1,
                       double scale, int32_t clip_nonneg, int32_t init_only, double *out_dev, void *stream);

/* The same with the history in aux channels aux_base, aux_base + 1 and, with all_local = 1, for the ghosts as well (a
 * ConcentrationField OBSERVABLE feeds the policy of every local particle: a ghost must take the action its owner
 * computes; a task reward is owned-only).  observables/concentration_field.py:84-138. */
int swarm_dd_obs_field_ch(swarm_t *s, int64_t n, const int32_t *n_dev, const swarm_dd_arrays_t *loc, double x_offset, double box_x,
                          const double src[3], const double box_length[2], int32_t decay, double c0, double c1, double scale,
                          int32_t clip_nonneg, int32_t init_only, int32_t aux_base, int32_t all_local, double *out_dev,
                          void *stream);

/* Periodic image count of every slot along the decomposed axis (int32, device), or NULL to unbind.  swarm_obs_vision then
 * lets two slots see each other only when they sit in the same image: the reference evaluates vision cones on unfolded
 * positions without a minimum image (observables/subdivided_vision_cones.py:113-157), and a rank's local frame places the
 * ghosts from across the periodic seam next to its own particles.  With a device-resident count bound (swarm_bind_count)
 * swarm_obs_vision also skips the slots past it.  Reset to NULL by swarm_bind_state. */
int swarm_bind_images(swarm_t *s, const int32_t *image_dev);

/* ---- observables / tasks -------------------------------------------------------------- */

/* ConcentrationField.compute_observable (observables/concentration_field.py:84-138) and
 * GradientSensing.__call__ (tasks/searching/gradient_sensing.py:92-150):
 *   p = pos/box_length; d_now = |src - p|; d_prev = |src - hist|; hist <- p
 *   out = scale * (decay(d_now) - decay(d_prev)), clipped to >= 0 when clip_nonneg.
 * row_of_dev[i] = output row of particle i, or -1 to skip it (type filter).
 * decay: NONE writes only d_now/d_prev; LINEAR c0 + c1 d; EXP c0 exp(-c1 d); INVERSE c0/(d + c1).
 * hist_dev = [3][n].  out_dev / d_now_dev / d_prev_dev may be NULL. */
int swarm_obs_field(swarm_t *s, const int32_t *row_of_dev, const double src[3], const double box_length[3],
                    double *hist_dev, int32_t decay, double c0, double c1, double scale, int32_t clip_nonneg,
                    double *out_dev, double *d_now_dev, double *d_prev_dev, void *stream);

/* SubdividedVisionCones.compute_observable (observables/subdivided_vision_cones.py:83-258):
 * out[row, cone, col] over every particle j (raw unfolded differences, observer included),
 * col = type_col_dev[type_j] (or -1 = undetected), radius of the SEEN particle.
 * dirs_dev = [3][n] observer directors, or NULL to use the bound orientation state.
 * out_dev = [rows][n_cones][n_cols], overwritten. */
int swarm_obs_vision(swarm_t *s, const int32_t *row_of_dev, int64_t rows, double vision_range, double half_angle,
                     int32_t n_cones, const double *radii_dev, const int32_t *type_col_dev, int32_t n_type_col,
                     int32_t n_cols, const double *dirs_dev, double *out_dev, void *stream);

/* ParticleSensing.compute_observable (observables/particle_sensing.py:95-178) and SpeciesSearch.__call__
 * (tasks/searching/species_search.py:97-184): field of the n_sensed particles listed (ascending) in
 * sensed_idx_dev at every particle with row_of >= 0, with the reference's nonzero(size=n_sensed-1) rule
 * (zero distances skipped, at most n_sensed-1 terms, padding re-reads sensed[0]).
 *   out[row] = scale * clip(field - hist[row]); hist[row] <- field.
 * clip_mode 0 none (observable), 1 keep >= 0 (task), -1 keep <= 0 (task, avoid=True).  init_only = 1 only
 * seeds hist (the classes' initialize()).  out_dev / field_dev may be NULL. */
int swarm_obs_pair_field(swarm_t *s, const int32_t *row_of_dev, const int32_t *sensed_idx_dev, int64_t n_sensed,
                         const double box_length[3], int32_t decay, double c0, double c1, double scale, int32_t clip_mode,
                         int32_t init_only, double *hist_rows_dev, double *out_dev, double *field_dev, void *stream);

/* PositionObservable (observables/position.py:32-60): out[row] = pos / box_length, [rows][3] */
int swarm_obs_position(swarm_t *s, const int32_t *row_of_dev, const double box_length[3], double *out_dev, void *stream);

/* Trainer.update_rl mean reward (trainers/trainer.py:92-95): out_dev[0] = sum, out_dev[1] = mean.
 * Deterministic two-stage reduction. */
int swarm_reduce_mean(swarm_t *s, const double *in_dev, int64_t n, double *out2_dev, void *stream);

/* swarm_obs_field (closed-form decay, rewards into out_dev) and swarm_reduce_mean over its n_rows rewards in ONE launch:
 * GradientSensing.__call__ (tasks/searching/gradient_sensing.py:140-183) followed by the trainer's np.mean
 * (trainers/trainer.py:92-95).  out2_dev[0] = sum, out2_dev[1] = sum / n_rows; fixed summation order. */
int swarm_obs_field_mean(swarm_t *s, const int32_t *row_of_dev, int64_t n_rows, const double src[3],
                         const double box_length[3], double *hist_dev, int32_t decay, double c0, double c1, double scale,
                         int32_t clip, double *out_dev, double *out2_dev, void *stream);

/* swarm_obs_field_mean that allocates NOTHING.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments, same results (bit for bit) as swarm_obs_field_mean; the per-CTA sums live in a static device array of
 * 32768 CTA slots (256 KiB per GPU that loads the library, held until exit), so a swarm of more than 8 388 608 particles
 * is refused with SWARM_ERR_INVALID.  Every call shares that array: calls that overlap on one GPU race. */
int swarm_obs_field_mean_noalloc(swarm_t *s, const int32_t *row_of_dev, int64_t n_rows, const double src[3],
                                 const double box_length[3], double *hist_dev, int32_t decay, double c0, double c1,
                                 double scale, int32_t clip, double *out_dev, double *out2_dev, void *stream);

/* ---- fused policy forward ----------------------------------------------------
 * ActorCriticAgent.calc_action (agents/actor_critic.py:159-184) -> FlaxModel.compute_action
 * (networks/flax_network.py:153-195) -> GumbelDistribution (sampling_strategies/gumbel_distribution.py:
 * 19-42) -> manage_forces' writes (engine/espresso.py:1227-1235), for the network of the reference's
 * tests (test_rl_trainers.py:17-26): Dense(hidden)+ReLU, actor Dense(n_actions), critic Dense(1).
 * Parameters are fp32 device arrays laid out like flax Dense kernels, [in][out] row-major. */
typedef struct {
  int32_t obs_dim, hidden, n_actions;
  const float *w1, *b1;   /* [obs_dim][hidden], [hidden] */
  const float *wa, *ba;   /* [hidden][n_actions], [n_actions] */
  const float *wv, *bv;   /* [hidden], [1] */
} swarm_policy_t;

/* obs_dev [rows][obs_dim] fp64 (what the observable kernels write); row_of as elsewhere (NULL: every
 * particle).  sample = 1: Gumbel-max with Philox uniforms keyed (particle id, seed; call); 0: greedy.
 * Outputs per row (each may be NULL): idx, log(softmax + 1e-8)[idx], value, logits [rows][n_actions].
 * act_force_dev [n_actions] / act_torque_dev [n_actions][3]: when given, the chosen action is written
 * into the bound f_swim / torque arrays of every selected particle (no host round trip). */
int swarm_policy_forward(swarm_t *s, const swarm_policy_t *net, const double *obs_dev, const int32_t *row_of_dev, uint64_t call,
                         int32_t sample, const double *act_force_dev, const double *act_torque_dev, int32_t *idx_out_dev,
                         float *logp_out_dev, float *value_out_dev, float *logits_out_dev, void *stream);

/* ---- host-buffer entry points (what a cgo/ctypes binding of the reference would call) --- */
/* manage_forces: write f_swim and ext torque of every particle (espresso.py:1227-1235) */
int swarm_set_actions_host(swarm_t *s, const double *f_swim_host, const double *torque3n_host, void *stream);
/* the same for a 2-D swarm: only the z component of the torque turns a director in the plane (the x, y components the
 * reference writes at espresso.py:1232 are zero there and never read by the 2-D kernels), so the host passes tau_z [n]:
 * half the host -> device bytes, no de-interleave kernel */
int swarm_set_actions_host_2d(swarm_t *s, const double *f_swim_host, const double *torque_z_host, void *stream);
/* Measurement hook (no reference counterpart): DFMA throughput of the device with `ilp` (1, 4 or 8) independent chains per
 * thread -- the fp64 peak the roofline notes of DESIGN.md / BASELINE.md divide by.  out4: DFMA/s, ms, SM count, ilp used. */
int swarm_microbench_dfma(int32_t device, int32_t ilp, int32_t iters, double out4[4]);

/* One trajectory frame (espresso.py:1110-1130: Unwrapped_Positions, Velocities, Directors) into three [n][3]
 * row-major DEVICE buffers, one kernel on `stream`, no host synchronisation: the snapshot ring of the asynchronous
 * trajectory writer (espresso.py:1278-1285 takes the snapshot before the chunk is integrated). */
int swarm_snapshot_dev(swarm_t *s, double *pos_n3_dev, double *vel_n3_dev, double *dir_n3_dev, void *stream);
/* get_particle_data (espresso.py:1320-1336): [n][3] row-major outputs; synchronises */
int swarm_get_particle_data_host(swarm_t *s, double *pos_n3, double *vel_n3, double *dir_n3, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* SWARM_B200_H */
