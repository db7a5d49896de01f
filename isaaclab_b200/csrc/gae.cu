// gae.cu -- RolloutStorage.compute_returns (RSL/storage/rollout_storage.py:127-149), add_transitions
// (rollout_storage.py:77-103 + PPO.process_env_step time-out bootstrap, RSL/algorithms/ppo.py:161-164) and the
// minibatch gather of mini_batch_generator (rollout_storage.py:188-197).
//
// HBM-bound byte work: one thread per env scans the T steps of its column (coalesced across the warp for
// every t), loads for a chunk of steps are issued before the dependent recurrence consumes them.
// Compiled with -fmad=false: the recurrence keeps torch's eager rounding order.
#include "common.cuh"

namespace b200lab {

constexpr int kGaeBlock = 64;
constexpr int kGaeChunk = 8;

__global__ void __launch_bounds__(kGaeBlock)
gae_scan_kernel(const float* __restrict__ rewards, const float* __restrict__ values,
                const uint8_t* __restrict__ dones, const float* __restrict__ last_values,
                float* __restrict__ returns, float* __restrict__ advantages, int T, int N, float gamma,
                float lam, double* __restrict__ partials) {
  int n = blockIdx.x * kGaeBlock + threadIdx.x;
  double s1 = 0.0, s2 = 0.0;
  if (n < N) {
    float adv = 0.0f;
    float next_v = last_values[n];
    for (int t0 = T; t0 > 0; t0 -= kGaeChunk) {
      float r[kGaeChunk], v[kGaeChunk];
      uint8_t d[kGaeChunk];
#pragma unroll
      for (int k = 0; k < kGaeChunk; ++k) {
        int t = t0 - 1 - k;
        if (t >= 0) {
          size_t o = (size_t)t * N + n;
          r[k] = __ldg(rewards + o);
          v[k] = __ldg(values + o);
          d[k] = __ldg(dones + o);
        }
      }
#pragma unroll
      for (int k = 0; k < kGaeChunk; ++k) {
        int t = t0 - 1 - k;
        if (t >= 0) {
          float nnt = 1.0f - (float)d[k];
          float ng = nnt * gamma;
          float delta = (r[k] + ng * next_v) - v[k];
          adv = delta + (ng * lam) * adv;
          float ret = adv + v[k];
          float a = ret - v[k];  // rollout_storage.py:146: advantages = returns - values
          size_t o = (size_t)t * N + n;
          returns[o] = ret;
          advantages[o] = a;
          s1 += (double)a;
          s2 += (double)a * (double)a;
          next_v = v[k];
        }
      }
    }
  }
  // block reduction of (sum, sumsq) in double, fixed order
  __shared__ double sh1[kGaeBlock / 32], sh2[kGaeBlock / 32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s1 += __shfl_xor_sync(0xffffffffu, s1, o);
    s2 += __shfl_xor_sync(0xffffffffu, s2, o);
  }
  int w = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0) { sh1[w] = s1; sh2[w] = s2; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0.0, b = 0.0;
    for (int i = 0; i < kGaeBlock / 32; ++i) { a += sh1[i]; b += sh2[i]; }
    partials[2 * blockIdx.x] = a;
    partials[2 * blockIdx.x + 1] = b;
  }
}

// (adv - mean) / (std + 1e-8), std unbiased over all T*N elements (torch.Tensor.std default).
__global__ void __launch_bounds__(256)
gae_normalize_kernel(float* __restrict__ advantages, int64_t total, const double* __restrict__ partials,
                     int num_partials) {
  __shared__ float s_mean, s_inv;
  if (threadIdx.x == 0) {
    double a = 0.0, b = 0.0;
    for (int i = 0; i < num_partials; ++i) { a += partials[2 * i]; b += partials[2 * i + 1]; }
    double mean = a / (double)total;
    double var = (b - a * mean) / (double)(total - 1);
    if (var < 0.0) var = 0.0;
    s_mean = (float)mean;
    s_inv = (float)sqrt(var) + 1e-8f;
  }
  __syncthreads();
  float mean = s_mean, den = s_inv;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < total; i += stride) advantages[i] = (advantages[i] - mean) / den;
}

// ---- add_transitions ----
__global__ void __launch_bounds__(256)
add_transition_kernel(B200LabTransition tr) {
  int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int N = tr.N;
  int64_t total_obs = (int64_t)N * tr.obs_dim;
  int64_t total_act = (int64_t)N * tr.act_dim;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  size_t t = (size_t)tr.t;
  if (tr.obs)   // NULL: the caller already wrote the observation row of step t (PPO.act stores it before env.step)
    for (int64_t i = n; i < total_obs; i += stride) tr.s_obs[t * total_obs + i] = tr.obs[i];
  for (int64_t i = n; i < total_act; i += stride) {
    tr.s_actions[t * total_act + i] = tr.actions[i];
    tr.s_mu[t * total_act + i] = tr.mu[i];
    tr.s_sigma[t * total_act + i] = tr.sigma[i];
  }
  for (int64_t i = n; i < N; i += stride) {
    float r = tr.rewards[i];
    float v = tr.values[i];
    if (tr.time_outs) r = r + tr.gamma * (v * (float)tr.time_outs[i]);  // ppo.py:161-164
    tr.s_rewards[t * N + i] = r;
    tr.s_dones[t * N + i] = tr.dones[i];
    tr.s_values[t * N + i] = v;
    tr.s_log_prob[t * N + i] = tr.log_prob[i];
  }
}

// ---- finished-episode bookkeeping of OnPolicyRunner.learn (RSL/runners/on_policy_runner.py:113-132) ----
// The reference does `(dones > 0).nonzero()` + `.cpu().tolist()` into two deque(maxlen=100) every env step (a host
// sync per step).  Here the deques are device rings with a device push counter; the host reads them once per iteration.
// One CTA: thread i owns a contiguous span of envs so pushes keep nonzero()'s ascending env order.
constexpr int kEpBlock = 1024;

template <typename DoneT>
__global__ void __launch_bounds__(kEpBlock)
episode_stats_kernel(const float* __restrict__ rewards, const DoneT* __restrict__ dones, int N,
                     float* __restrict__ cur_sum, float* __restrict__ cur_len, float* __restrict__ ring_rew,
                     float* __restrict__ ring_len, int ring, long long* __restrict__ pushed) {
  __shared__ int warp_tot[kEpBlock / 32];
  int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  int span = (N + kEpBlock - 1) / kEpBlock;
  int lo = min(tid * span, N), hi = min(lo + span, N);
  int mine = 0;
  for (int n = lo; n < hi; ++n) mine += dones[n] > 0;
  int incl = mine;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    int v = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += v;
  }
  if (lane == 31) warp_tot[wid] = incl;
  __syncthreads();
  int before = 0, total = 0;
  for (int w = 0; w < kEpBlock / 32; ++w) {
    int t = warp_tot[w];
    if (w < wid) before += t;
    total += t;
  }
  long long base = *pushed;
  int rank = before + incl - mine;  // exclusive rank of this thread's first finished episode within the step
  for (int n = lo; n < hi; ++n) {
    float s = cur_sum[n] + rewards[n];
    float l = cur_len[n] + 1.0f;
    if (dones[n] > 0) {
      if (total - rank <= ring) {  // older pushes of the same step would be evicted by deque(maxlen) anyway
        int slot = (int)((base + rank) % ring);
        ring_rew[slot] = s;
        ring_len[slot] = l;
      }
      ++rank;
      s = 0.0f;
      l = 0.0f;
    }
    cur_sum[n] = s;
    cur_len[n] = l;
  }
  __syncthreads();
  if (tid == 0) *pushed = base + total;
}

// ---- minibatch gather: one warp per sample row, lanes stride the columns (coalesced) ----
__global__ void __launch_bounds__(256)
gather_minibatch_kernel(B200LabGather g) {
  int lane = threadIdx.x & 31;
  int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t m = warp; m < g.M; m += nwarps) {
    int64_t src = __ldg(g.idx + m);
    const float* so = g.obs + src * g.obs_dim;
    float* dobs = g.o_obs + m * (g.o_obs_ld > 0 ? g.o_obs_ld : (int64_t)g.obs_dim);
    for (int c = lane; c < g.obs_dim; c += 32) dobs[c] = __ldg(so + c);
    for (int c = lane; c < g.act_dim; c += 32) {
      g.o_actions[m * g.act_dim + c] = __ldg(g.actions + src * g.act_dim + c);
      g.o_mu[m * g.act_dim + c] = __ldg(g.mu + src * g.act_dim + c);
      g.o_sigma[m * g.act_dim + c] = __ldg(g.sigma + src * g.act_dim + c);
    }
    if (lane == 0) g.o_values[m] = __ldg(g.values + src);
    if (lane == 1) g.o_returns[m] = __ldg(g.returns + src);
    if (lane == 2) g.o_log_prob[m] = __ldg(g.log_prob + src);
    if (lane == 3) g.o_advantages[m] = __ldg(g.advantages + src);
  }
}

// ---- library-owned scratch for b200lab_gae_static (no allocation anywhere on that path) ----
// File-scope device array: every GPU that loads the library holds these 64 KB until the process exits.  The name is a
// HOST-SIDE STAND-IN in host code, not a device address -- the address a kernel may be given comes from
// cudaGetSymbolAddress, resolved once per device with that device current (gae_static_scratch below).
constexpr int kGaeStaticMaxEnvs = B200LAB_GAE_STATIC_MAX_ENVS;
constexpr int kGaeStaticDoubles = 2 * (kGaeStaticMaxEnvs / kGaeBlock) + 2;
static_assert(kGaeStaticMaxEnvs % kGaeBlock == 0, "B200LAB_GAE_STATIC_MAX_ENVS must be a multiple of the GAE block");
static __device__ __align__(256) double g_gae_static_scratch[kGaeStaticDoubles];

constexpr int kGaeStaticMaxDevices = 64;
static double* g_gae_static_ptr[kGaeStaticMaxDevices];   // per-device address of the array, resolved on first use

static int gae_static_scratch(double** out) {
  int dev = -1;
  cudaError_t err = cudaGetDevice(&dev);
  if (err != cudaSuccess) { set_error("b200lab_gae_static: cudaGetDevice: %s", cudaGetErrorString(err)); return B200LAB_ERR_CUDA; }
  if (dev < 0 || dev >= kGaeStaticMaxDevices) { set_error("b200lab_gae_static: device index %d out of range", dev); return B200LAB_ERR_ARG; }
  if (!g_gae_static_ptr[dev]) {
    double* p = nullptr;
    err = cudaGetSymbolAddress((void**)&p, g_gae_static_scratch);
    if (err != cudaSuccess || !p) {
      set_error("b200lab_gae_static: cudaGetSymbolAddress: %s", cudaGetErrorString(err));
      return B200LAB_ERR_CUDA;
    }
    g_gae_static_ptr[dev] = p;
  }
  *out = g_gae_static_ptr[dev];
  return B200LAB_OK;
}

}  // namespace b200lab

using namespace b200lab;

extern "C" {

int64_t b200lab_gae_scratch_bytes(int32_t T, int32_t N) {
  (void)T;
  return (int64_t)(2 * ceil_div(N, kGaeBlock) + 2) * (int64_t)sizeof(double);
}

int b200lab_gae(const float* rewards, const float* values, const uint8_t* dones, const float* last_values,
                float* returns, float* advantages, int32_t T, int32_t N, float gamma, float lam,
                int32_t normalize_advantage, void* scratch, void* stream) {
  B200LAB_CHECK_ARG(T > 0 && N > 0, "T and N must be positive");
  B200LAB_CHECK_ARG(rewards && values && dones && last_values && returns && advantages && scratch, "null pointer");
  cudaStream_t s = (cudaStream_t)stream;
  int blocks = (int)ceil_div(N, kGaeBlock);
  gae_scan_kernel<<<blocks, kGaeBlock, 0, s>>>(rewards, values, dones, last_values, returns, advantages, T, N,
                                               gamma, lam, (double*)scratch);
  count_launch();
  int rc = check_launch("gae_scan");
  if (rc) return rc;
  if (normalize_advantage) {
    int64_t total = (int64_t)T * N;
    int nb = (int)min64(ceil_div(total, 256 * 4), 4 * kNumSMs);
    gae_normalize_kernel<<<nb, 256, 0, s>>>(advantages, total, (const double*)scratch, blocks);
    count_launch();
    rc = check_launch("gae_normalize");
  }
  return rc;
}

int b200lab_gae_static(const float* rewards, const float* values, const uint8_t* dones, const float* last_values,
                       float* returns, float* advantages, int32_t T, int32_t N, float gamma, float lam,
                       int32_t normalize_advantage, void* stream) {
  B200LAB_CHECK_ARG(T > 0 && N > 0, "T and N must be positive");
  if (N > kGaeStaticMaxEnvs) {
    set_error("b200lab_gae_static: N = %d exceeds B200LAB_GAE_STATIC_MAX_ENVS = %d (the library-owned scratch is fixed-size; "
              "use b200lab_gae with caller scratch)", (int)N, kGaeStaticMaxEnvs);
    return B200LAB_ERR_ARG;
  }
  B200LAB_CHECK_ARG(rewards && values && dones && last_values && returns && advantages, "null pointer");
  double* scratch = nullptr;
  int rc = gae_static_scratch(&scratch);
  if (rc) return rc;
  // Same two kernels, same launch geometry as b200lab_gae => bitwise the same results.  Nothing in the scratch has to
  // start at zero: gae_scan_kernel overwrites partial slots [0, 2*blocks) and gae_normalize_kernel reads exactly those.
  cudaStream_t s = (cudaStream_t)stream;
  int blocks = (int)ceil_div(N, kGaeBlock);
  gae_scan_kernel<<<blocks, kGaeBlock, 0, s>>>(rewards, values, dones, last_values, returns, advantages, T, N,
                                               gamma, lam, scratch);
  count_launch();
  rc = check_launch("gae_scan (static scratch)");
  if (rc) return rc;
  if (normalize_advantage) {
    int64_t total = (int64_t)T * N;
    int nb = (int)min64(ceil_div(total, 256 * 4), 4 * kNumSMs);
    gae_normalize_kernel<<<nb, 256, 0, s>>>(advantages, total, (const double*)scratch, blocks);
    count_launch();
    rc = check_launch("gae_normalize (static scratch)");
  }
  return rc;
}

int b200lab_add_transition(const B200LabTransition* tr, void* stream) {
  B200LAB_CHECK_ARG(tr && tr->t >= 0 && tr->t < tr->T, "bad transition index (rollout buffer overflow)");
  int64_t work = (int64_t)tr->N * tr->obs_dim;
  int nb = (int)min64(ceil_div(work, 256), 8 * kNumSMs);
  add_transition_kernel<<<nb, 256, 0, (cudaStream_t)stream>>>(*tr);
  count_launch();
  return check_launch("add_transition");
}

int b200lab_episode_stats(const float* rewards, const void* dones, int32_t dones_elem_size, int32_t N,
                          float* cur_reward_sum, float* cur_episode_length, float* ring_reward, float* ring_length,
                          int32_t ring_size, int64_t* pushed, void* stream) {
  B200LAB_CHECK_ARG(rewards && dones && cur_reward_sum && cur_episode_length && ring_reward && ring_length && pushed,
                    "episode_stats: null pointer");
  B200LAB_CHECK_ARG(N > 0 && ring_size > 0, "episode_stats: bad sizes");
  B200LAB_CHECK_ARG(dones_elem_size == 1 || dones_elem_size == 8, "episode_stats: dones must be u8/bool or int64");
  cudaStream_t s = (cudaStream_t)stream;
  if (dones_elem_size == 1)
    episode_stats_kernel<uint8_t><<<1, kEpBlock, 0, s>>>(rewards, (const uint8_t*)dones, N, cur_reward_sum,
                                                         cur_episode_length, ring_reward, ring_length, ring_size,
                                                         (long long*)pushed);
  else
    episode_stats_kernel<long long><<<1, kEpBlock, 0, s>>>(rewards, (const long long*)dones, N, cur_reward_sum,
                                                           cur_episode_length, ring_reward, ring_length, ring_size,
                                                           (long long*)pushed);
  count_launch();
  return check_launch("episode_stats");
}

int b200lab_gather_minibatch(const B200LabGather* g, void* stream) {
  B200LAB_CHECK_ARG(g && g->M > 0, "bad gather args");
  int nb = (int)min64(ceil_div(g->M * 32, 256), 16 * kNumSMs);
  gather_minibatch_kernel<<<nb, 256, 0, (cudaStream_t)stream>>>(*g);
  count_launch();
  return check_launch("gather_minibatch");
}
}
