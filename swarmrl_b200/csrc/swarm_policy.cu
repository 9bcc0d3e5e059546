// swarm_policy.cu -- fused batched policy forward: the step a rollout takes
// right after the observable, kept on the device so a slice needs no host round trip.
//
// Replaces, for the architecture of the reference's own tests (CI/espresso_tests/integration_tests/
// test_rl_trainers.py:17-26: Dense(H) + ReLU trunk, Dense(A) actor head, Dense(1) critic head):
//   ActorCriticAgent.calc_action            swarmrl/agents/actor_critic.py:159-184
//   FlaxModel.compute_action                swarmrl/networks/flax_network.py:153-195
//     logits, value = net(obs); idx = sampling_strategy(logits);
//     log_prob = log(softmax(logits) + 1e-8)[idx]
//   GumbelDistribution.__call__             swarmrl/sampling_strategies/gumbel_distribution.py:19-42
//     idx = argmax(logits - log(-log(u))),  u ~ U(0, 1)
//   manage_forces' property writes          swarmrl/engine/espresso.py:1227-1235
//
// fp32 arithmetic like JAX's default; one thread per agent, the hidden layer is never materialised (each
// hidden unit is consumed by both heads as soon as it is computed), weights staged in shared memory.
// With obs_dim <= 16 and 4 actions this is ~2.5 kFLOP per agent: CUDA cores, not tensor cores -- a
// K = 5 GEMM cannot feed tcgen05.  The uniforms come from the engine's Philox
// (key = (agent id, seed), counter = (call, 4)): JAX's threefry stream cannot be reproduced without JAX,
// so sampling parity is statistical, the deterministic outputs (logits, value, log-prob) are exact.
#include "swarm_common.cuh"

#define POL_THREADS 128
#define POL_MAX_OBS 32
#define POL_MAX_ACT 16

struct PolNet {
  int obs_dim, hidden, n_actions;
  const float *w1, *b1, *wa, *ba, *wv, *bv;
};

// DMAX / AMAX bound the unrolled loops: the network of the reference's tests (obs_dim <= 8, 4 actions) runs the <8, 4>
// instance.  The parameters of one hidden unit sit in ONE shared-memory record, zero-padded to the instance's widths,
//     rec[j] = { w1[0..DMAX) , wa[0..AMAX) , b1 , wv , 0 , 0 }        (DMAX + AMAX + 4 floats, 16-byte aligned)
// so a unit costs (DMAX + AMAX + 4) / 4 broadcast LDS.128 and DMAX + AMAX + 1 unpredicated FFMAs; a padded term is
// fma(0, 0, h) = h, so the sums are the ones of the unpadded network bit for bit.  The earlier form read every weight
// with its own predicated LDS.32 (58 instructions per unit in SASS, 41 us at C4: profiles/r02_launches_engine.csv).
template <int DMAX, int AMAX>
__global__ void __launch_bounds__(POL_THREADS) k_policy_forward(long long n, PolNet N, const double *obs, const int *row_of,
                                                                const long long *gid, u64 seed, u64 call, int sample,
                                                                const double *act_force, const double *act_torque,
                                                                double *f_swim, double *tqx, double *tqy, double *tqz, int *idx_out,
                                                                float *logp_out, float *value_out, float *logits_out) {
  extern __shared__ float4 sm4[];
  constexpr int REC = DMAX + AMAX + 4;   // floats per hidden unit
  float *sm = reinterpret_cast<float *>(sm4);
  const int D = N.obs_dim, H = N.hidden, A = N.n_actions;
  for (int t = threadIdx.x; t < H * REC; t += blockDim.x) {
    const int j = t / REC, k = t - j * REC;
    float v = 0.0f;
    if (k < DMAX) { if (k < D) v = N.w1[k * H + j]; }
    else if (k < DMAX + AMAX) { if (k - DMAX < A) v = N.wa[j * A + (k - DMAX)]; }
    else if (k == DMAX + AMAX) v = N.b1[j];
    else if (k == DMAX + AMAX + 1) v = N.wv[j];
    sm[t] = v;
  }
  __syncthreads();
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int row = row_of ? row_of[i] : (int)i;
  if (row < 0) return;
  float x[DMAX], lg[AMAX];
#pragma unroll
  for (int d = 0; d < DMAX; ++d) x[d] = (d < D) ? (float)obs[(long long)row * D + d] : 0.0f;
#pragma unroll
  for (int a = 0; a < AMAX; ++a) lg[a] = (a < A) ? N.ba[a] : -INFINITY;
  float val = N.bv[0];
#pragma unroll 4
  for (int j = 0; j < H; ++j) {
    float r[REC];
#pragma unroll
    for (int q = 0; q < REC / 4; ++q) {
      const float4 v = sm4[j * (REC / 4) + q];
      r[4 * q] = v.x; r[4 * q + 1] = v.y; r[4 * q + 2] = v.z; r[4 * q + 3] = v.w;
    }
    float h = r[DMAX + AMAX];
#pragma unroll
    for (int d = 0; d < DMAX; ++d) h = fmaf(x[d], r[d], h);
    h = fmaxf(h, 0.0f);
#pragma unroll
    for (int a = 0; a < AMAX; ++a) lg[a] = fmaf(h, r[DMAX + a], lg[a]);
    val = fmaf(h, r[DMAX + AMAX + 1], val);
  }
  // log(softmax + 1e-8), max-shifted as jax.nn.softmax does
  float mx = -INFINITY;
#pragma unroll
  for (int a = 0; a < AMAX; ++a) if (a < A) mx = fmaxf(mx, lg[a]);
  float den = 0.0f;
#pragma unroll
  for (int a = 0; a < AMAX; ++a) if (a < A) den += expf(lg[a] - mx);
  int best = 0;
  if (sample) {
    const u64 pid = gid ? (u64)gid[i] : (u64)i;
    float bv = -INFINITY;
    for (int a0 = 0; a0 < A; a0 += 4) {
      u64 r[4];
      sw_philox4x64_10(call, 4ULL + (u64)(a0 >> 2) * 16ULL, SW_PHILOX_DOMAIN_POLICY, 0ULL, pid, seed, r);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int a = a0 + u;
        if (a >= A) break;
        const double uu = ((double)(r[u] >> 11) + 0.5) * (1.0 / 9007199254740992.0);
        float v = -INFINITY;
#pragma unroll
        for (int q = 0; q < AMAX; ++q) if (q == a) v = lg[q];
        v -= (float)log(-log(uu));
        if (v > bv) { bv = v; best = a; }
      }
    }
  } else {
    float bv = -INFINITY;
#pragma unroll
    for (int a = 0; a < AMAX; ++a) if (a < A && lg[a] > bv) { bv = lg[a]; best = a; }
  }
  float lb = 0.0f;
#pragma unroll
  for (int a = 0; a < AMAX; ++a) if (a == best) lb = lg[a];
  if (idx_out) idx_out[row] = best;
  if (logp_out) logp_out[row] = logf(expf(lb - mx) / den + 1e-8f);
  if (value_out) value_out[row] = val;
  if (logits_out)
    for (int a = 0; a < A; ++a) {
      float v = 0.0f;
#pragma unroll
      for (int q = 0; q < AMAX; ++q) if (q == a) v = lg[q];
      logits_out[(long long)row * A + a] = v;
    }
  if (act_force && f_swim) {   // manage_forces: the chosen action straight into the engine's arrays
    f_swim[i] = act_force[best];
    tqx[i] = act_torque[3 * best]; tqy[i] = act_torque[3 * best + 1]; tqz[i] = act_torque[3 * best + 2];
  }
}

extern "C" int swarm_policy_forward(swarm_t *s, const swarm_policy_t *net, const double *obs_dev, const int32_t *row_of_dev,
                                    uint64_t call, int32_t sample, const double *act_force_dev, const double *act_torque_dev,
                                    int32_t *idx_out_dev, float *logp_out_dev, float *value_out_dev, float *logits_out_dev,
                                    void *stream) {
  if (!s || !s->state_bound || !net || !obs_dev) { sw_set_error("bad arguments"); return SWARM_ERR_INVALID; }
  if (net->obs_dim < 1 || net->obs_dim > POL_MAX_OBS || net->n_actions < 1 || net->n_actions > POL_MAX_ACT || net->hidden < 1 ||
      !net->w1 || !net->b1 || !net->wa || !net->ba || !net->wv || !net->bv) {
    sw_set_error("policy: obs_dim must be 1..32, n_actions 1..16, all six parameter arrays bound"); return SWARM_ERR_INVALID;
  }
  if ((act_force_dev == nullptr) != (act_torque_dev == nullptr)) { sw_set_error("bind both action tables or none"); return SWARM_ERR_INVALID; }
  if (act_force_dev && !s->actions_bound) { sw_set_error("bind the action arrays before asking the policy to write them"); return SWARM_ERR_STATE; }
  const int dmax = (net->obs_dim <= 8 && net->n_actions <= 4) ? 8 : ((net->obs_dim <= 16 && net->n_actions <= 8) ? 16 : POL_MAX_OBS);
  const int amax = dmax == 8 ? 4 : (dmax == 16 ? 8 : POL_MAX_ACT);
  const size_t smem = sizeof(float) * (size_t)net->hidden * (size_t)(dmax + amax + 4);
  if (smem > 200 * 1024) { sw_set_error("policy: parameters do not fit in shared memory"); return SWARM_ERR_INVALID; }
  SW_CUDA(cudaSetDevice(s->cfg.device));
  const DevArrays &A = s->A;
  if (A.n == 0) return SWARM_OK;
  PolNet N{net->obs_dim, net->hidden, net->n_actions, net->w1, net->b1, net->wa, net->ba, net->wv, net->bv};
#define POL_LAUNCH(DM, AM)                                                                                                        \
  do {                                                                                                                            \
    if (smem > 48 * 1024) SW_CUDA(cudaFuncSetAttribute(k_policy_forward<DM, AM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    k_policy_forward<DM, AM><<<sw_blocks(A.n, POL_THREADS), POL_THREADS, smem, (cudaStream_t)stream>>>(                           \
        A.n, N, obs_dev, row_of_dev, A.gid, s->P.seed, call, sample, act_force_dev, act_torque_dev, A.f_swim, A.tqx, A.tqy,      \
        A.tqz, idx_out_dev, logp_out_dev, value_out_dev, logits_out_dev);                                                          \
  } while (0)
  if (net->obs_dim <= 8 && net->n_actions <= 4) POL_LAUNCH(8, 4);
  else if (net->obs_dim <= 16 && net->n_actions <= 8) POL_LAUNCH(16, 8);
  else POL_LAUNCH(POL_MAX_OBS, POL_MAX_ACT);
#undef POL_LAUNCH
  s->launches += 1;
  SW_CUDA(cudaGetLastError());
  return SWARM_OK;
}
