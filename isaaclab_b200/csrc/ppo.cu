// ppo.cu -- the elementwise half of rsl_rl PPO (RSL/algorithms/ppo.py:129-140, 249-315, 380-381):
// action sampling + log-prob, the fused loss forward/backward, KL-adaptive learning rate, grad-norm clip + Adam.
//
// Everything here is HBM-bound streaming work over [M, A] minibatch rows; each kernel reads its inputs once,
// writes the gradients w.r.t. the network outputs once and reduces the scalar statistics with warp shuffles +
// one atomic per CTA, so no intermediate tensor of the reference's ~100 eager kernels is materialised.
#include "common.cuh"

namespace b200lab {

constexpr float kHalfLog2Pi = 0.91893853320467274178f;  // 0.5*log(2*pi)

// ActorCritic.act: Normal(mu, std).sample() = mu + std * eps (torch.normal semantic); get_actions_log_prob:
// sum_a [-(a-mu)^2/(2 var) - log(std) - 0.5 log(2 pi)]  (torch.distributions.Normal.log_prob).
__global__ void __launch_bounds__(256)
sample_actions_kernel(const float* __restrict__ mu, const float* __restrict__ std, const float* __restrict__ eps,
                      float* __restrict__ actions, float* __restrict__ log_prob, float* __restrict__ sigma_out,
                      int64_t M, int A) {
  int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= M) return;
  float lp = 0.0f;
  for (int a = 0; a < A; ++a) {
    float s = __ldg(std + a);
    float mean = mu[m * A + a];
    float act = mean + s * eps[m * A + a];
    actions[m * A + a] = act;
    if (sigma_out) sigma_out[m * A + a] = s;
    float d = act - mean;
    float var = s * s;
    lp += (-(d * d) / (2.0f * var) - logf(s)) - kHalfLog2Pi;
  }
  log_prob[m] = lp;
}

__device__ __forceinline__ void block_atomic_add(float v, float* dst, float* sh) {
  v = warp_sum(v);
  int w = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0) sh[w] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.0f;
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) t += sh[i];
    atomicAdd(dst, t);
  }
  __syncthreads();
}

constexpr int kMaxA = 64;

// One thread per minibatch sample.  Gradients are those autograd produces for
//   loss = mean(max(-A r, -A clamp(r))) + c_v mean(max((v-R)^2, (vc-R)^2)) - c_e mean(entropy)
// including torch.max's 0.5/0.5 split on ties and clamp's pass-through inside [lo, hi].
__global__ void __launch_bounds__(128)
ppo_loss_kernel(B200LabPpoLoss p) {
  __shared__ float sh[4];
  __shared__ float sh_dstd[kMaxA];
  const int A = p.A;
  for (int a = threadIdx.x; a < A; a += blockDim.x) sh_dstd[a] = 0.0f;
  __syncthreads();
  int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool ok = m < p.M;
  const float invM = 1.0f / (float)p.M;
  float surr_loss = 0.f, v_loss = 0.f, ent = 0.f, kl = 0.f;
  float d_lp = 0.f;
  if (ok) {
    // log prob of the stored actions under the current policy, entropy, KL to the old policy
    float lp = 0.f;
    for (int a = 0; a < A; ++a) {
      float s = __ldg(p.std + a), mu = p.mu[m * A + a], act = p.actions[m * A + a];
      float d = act - mu;
      lp += (-(d * d) / (2.0f * (s * s)) - logf(s)) - kHalfLog2Pi;
      ent += 0.5f + kHalfLog2Pi + logf(s);
      float os = p.old_sigma[m * A + a], om = p.old_mu[m * A + a];
      float dm = om - mu;
      kl += (logf(s / os + 1.0e-5f) + (os * os + dm * dm) / (2.0f * (s * s))) - 0.5f;
    }
    float adv = p.advantages[m];
    float ratio = expf(lp - p.old_log_prob[m]);
    float lo = 1.0f - p.clip_param, hi = 1.0f + p.clip_param;
    float rc = fminf(fmaxf(ratio, lo), hi);
    float s1 = -adv * ratio, s2 = -adv * rc;
    surr_loss = fmaxf(s1, s2);
    // d loss / d ratio
    float g1 = s1 > s2 ? 1.0f : (s1 == s2 ? 0.5f : 0.0f);
    float g2 = 1.0f - g1;
    float in_range = (ratio >= lo && ratio <= hi) ? 1.0f : 0.0f;
    float d_ratio = (-adv) * (g1 + g2 * in_range) * invM;
    d_lp = d_ratio * ratio;
    // value loss
    float v = p.value[m], R = p.returns[m];
    float dv;
    if (p.use_clipped_value_loss) {
      float tv = p.target_values[m];
      float diff = v - tv;
      float vc = tv + fminf(fmaxf(diff, -p.clip_param), p.clip_param);
      float l1 = (v - R) * (v - R), l2 = (vc - R) * (vc - R);
      v_loss = fmaxf(l1, l2);
      float h1 = l1 > l2 ? 1.0f : (l1 == l2 ? 0.5f : 0.0f);
      float h2 = 1.0f - h1;
      float inr = (diff >= -p.clip_param && diff <= p.clip_param) ? 1.0f : 0.0f;
      dv = h1 * 2.0f * (v - R) + h2 * 2.0f * (vc - R) * inr;
    } else {
      v_loss = (R - v) * (R - v);
      dv = 2.0f * (v - R);
    }
    p.d_value[m] = dv * p.value_loss_coef * invM;
  }
  // per-action gradients
  for (int a = 0; a < A; ++a) {
    float dstd = 0.f;
    if (ok) {
      float s = __ldg(p.std + a), mu = p.mu[m * A + a], act = p.actions[m * A + a];
      float d = act - mu;
      float var = s * s;
      p.d_mu[m * A + a] = d_lp * (d / var);
      // d lp / d std = d^2/s^3 - 1/s ; entropy term: -c_e * mean(sum log s) -> -c_e/(M s)
      dstd = d_lp * ((d * d) / (var * s) - 1.0f / s) - p.entropy_coef * invM / s;
    }
    dstd = warp_sum(dstd);
    if ((threadIdx.x & 31) == 0) atomicAdd(&sh_dstd[a], dstd);
  }
  __syncthreads();
  for (int a = threadIdx.x; a < A; a += blockDim.x) atomicAdd(p.d_std + a, sh_dstd[a]);
  block_atomic_add(surr_loss * invM, p.stats + 0, sh);
  block_atomic_add(v_loss * invM, p.stats + 1, sh);
  block_atomic_add(ent * invM, p.stats + 2, sh);
  block_atomic_add(kl * invM, p.stats + 3, sh);
}

// ppo.py:276-294 on the device: lr stays in HBM so the schedule needs no host round trip.  stats[3] is the
// (all-reduced SUM of the) mean KL; inv_world_size turns the sum into the mean over ranks (ppo.py:272-274).
__global__ void adapt_lr_kernel(float* lr, const float* stats, float desired_kl, float inv_world_size) {
  float kl = stats[3] * inv_world_size;
  float v = *lr;
  if (kl > desired_kl * 2.0f) v = fmaxf(1e-5f, v / 1.5f);
  else if (kl < desired_kl / 2.0f && kl > 0.0f) v = fminf(1e-2f, v * 1.5f);
  *lr = v;
}

__global__ void __launch_bounds__(256)
grad_sumsq_kernel(const float* __restrict__ g, int64_t n, float* __restrict__ out) {
  __shared__ float sh[8];
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  float s = 0.f;
  for (; i < n; i += stride) { float x = __ldg(g + i); s = fmaf(x, x, s); }
  block_atomic_add(s, out, sh);
}

// nn.utils.clip_grad_norm_ (clip_coef = max_norm / (total_norm + 1e-6), clamped to 1) followed by torch.optim.Adam
// (no amsgrad, no weight decay): identical update formula to torch/optim/adam.py _single_tensor_adam.
__global__ void __launch_bounds__(256)
adam_clip_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                 int64_t n, const float* __restrict__ lr_dev, const float* __restrict__ sumsq, float grad_scale,
                 float max_norm, float b1, float b2, float eps, float bc1, float bc2_sqrt) {
  float lr = *lr_dev;
  float coef = 1.0f;
  if (max_norm > 0.0f) {
    float total = sqrtf(*sumsq) * grad_scale;
    coef = fminf(max_norm / (total + 1e-6f), 1.0f);
  }
  coef *= grad_scale;
  float step_size = lr / bc1;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    float gr = g[i] * coef;
    float mi = m[i] + (gr - m[i]) * (1.0f - b1);          // exp_avg.lerp_(grad, 1 - beta1)
    float vi = v[i] * b2 + (1.0f - b2) * gr * gr;         // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1 - beta2)
    m[i] = mi;
    v[i] = vi;
    float denom = sqrtf(vi) / bc2_sqrt + eps;
    p[i] = p[i] - step_size * (mi / denom);
  }
}

// ---- EmpiricalNormalization (RSL/networks/normalization.py:14-72) ----
// update: batch mean / biased variance per column (two passes over the [M,D] batch, as torch.var does), then
//   mean += rate*(mean_x - mean);  var += rate*(var_x - var + delta*(mean_x - mean_new));  std = sqrt(var)
// with rate = M / count_new.  One CTA per 32 columns: 8 row-lanes x 32 column-lanes, coalesced along the columns.
__global__ void __launch_bounds__(256)
empirical_norm_update_kernel(const float* __restrict__ x, int64_t M, int D, float rate, float* __restrict__ mean,
                             float* __restrict__ var, float* __restrict__ stdv) {
  __shared__ float sh[8][33];
  __shared__ float mean_x_s[32];
  const int cl = threadIdx.x & 31, rl = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + cl;
  float s = 0.0f;
  if (c < D)
    for (int64_t r = rl; r < M; r += 8) s += __ldg(x + r * D + c);
  sh[rl][cl] = s;
  __syncthreads();
  if (rl == 0) {
    float t = 0.0f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += sh[k][cl];
    mean_x_s[cl] = t / (float)M;
  }
  __syncthreads();
  const float mx = mean_x_s[cl];
  float q = 0.0f;
  if (c < D)
    for (int64_t r = rl; r < M; r += 8) {
      float d = __ldg(x + r * D + c) - mx;
      q += d * d;
    }
  __syncthreads();
  sh[rl][cl] = q;
  __syncthreads();
  if (rl == 0 && c < D) {
    float t = 0.0f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += sh[k][cl];
    const float var_x = t / (float)M;
    const float m_old = mean[c];
    const float delta = mx - m_old;
    const float m_new = m_old + rate * delta;
    const float v_new = var[c] + rate * (var_x - var[c] + delta * (mx - m_new));
    mean[c] = m_new;
    var[c] = v_new;
    stdv[c] = sqrtf(v_new);
  }
}

// Global statistics across ranks (north_star: NCCL all-reduce of the observation-normaliser statistics): every rank
// reduces its batch to per-column (sum, sum of squares) in fp64, the [2D+1] buffers (last slot: the rank's sample count)
// are summed over the ranks by ONE all-reduce, and the merge applies the reference's running update
// (normalization.py:50-66) with the batch mean / biased variance of the UNION of the ranks' batches.
__global__ void __launch_bounds__(256)
empirical_norm_partial_kernel(const float* __restrict__ x, int64_t M, int D, double* __restrict__ sums) {
  __shared__ double sh1[8][33], sh2[8][33];
  const int cl = threadIdx.x & 31, rl = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + cl;
  double s = 0.0, q = 0.0;
  if (c < D)
    for (int64_t r = rl; r < M; r += 8) {
      const double v = (double)__ldg(x + r * D + c);
      s += v;
      q += v * v;
    }
  sh1[rl][cl] = s;
  sh2[rl][cl] = q;
  __syncthreads();
  if (rl == 0 && c < D) {
    double a = 0.0, b = 0.0;
#pragma unroll
    for (int k = 0; k < 8; ++k) { a += sh1[k][cl]; b += sh2[k][cl]; }
    sums[c] = a;
    sums[D + c] = b;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) sums[2 * D] = (double)M;
}

__global__ void __launch_bounds__(256)
empirical_norm_merge_kernel(const double* __restrict__ sums, int D, int64_t count_old, float* __restrict__ mean,
                            float* __restrict__ var, float* __restrict__ stdv) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= D) return;
  const double Mtot = sums[2 * D];
  const float rate = (float)(Mtot / ((double)count_old + Mtot));
  const double mxd = sums[c] / Mtot;
  double vxd = sums[D + c] / Mtot - mxd * mxd;
  if (vxd < 0.0) vxd = 0.0;
  const float mx = (float)mxd, var_x = (float)vxd;
  const float m_old = mean[c];
  const float delta = mx - m_old;
  const float m_new = m_old + rate * delta;
  const float v_new = var[c] + rate * (var_x - var[c] + delta * (mx - m_new));
  mean[c] = m_new;
  var[c] = v_new;
  stdv[c] = sqrtf(v_new);
}

// forward: (x - mean) / (std + eps)
__global__ void __launch_bounds__(256)
empirical_norm_apply_kernel(const float* __restrict__ x, const float* __restrict__ mean, const float* __restrict__ stdv,
                            float eps, float* __restrict__ out, int64_t M, int D) {
  const int64_t total = M * D;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int c = (int)(i % D);
    out[i] = (__ldg(x + i) - __ldg(mean + c)) / (__ldg(stdv + c) + eps);
  }
}

}  // namespace b200lab

using namespace b200lab;

extern "C" {

int b200lab_sample_actions(const float* mu, const float* std, const float* eps, float* actions, float* log_prob,
                           float* sigma_out, int64_t M, int32_t A, void* stream) {
  B200LAB_CHECK_ARG(mu && std && eps && actions && log_prob && M > 0 && A > 0, "bad argument");
  sample_actions_kernel<<<(unsigned)ceil_div(M, 256), 256, 0, (cudaStream_t)stream>>>(mu, std, eps, actions, log_prob,
                                                                                       sigma_out, M, A);
  count_launch();
  return check_launch("sample_actions");
}

int b200lab_ppo_loss(const B200LabPpoLoss* a, void* stream) {
  B200LAB_CHECK_ARG(a && a->M > 0 && a->A > 0 && a->A <= kMaxA, "bad argument (A must be <= 64)");
  B200LAB_CHECK_ARG(a->mu && a->std && a->value && a->actions && a->old_log_prob && a->old_mu && a->old_sigma &&
                        a->advantages && a->returns && a->target_values && a->d_mu && a->d_value && a->d_std && a->stats,
                    "null pointer");
  ppo_loss_kernel<<<(unsigned)ceil_div(a->M, 128), 128, 0, (cudaStream_t)stream>>>(*a);
  count_launch();
  return check_launch("ppo_loss");
}

int b200lab_adapt_lr(float* lr_dev, const float* stats, float desired_kl, float inv_world_size, void* stream) {
  B200LAB_CHECK_ARG(lr_dev && stats, "null pointer");
  adapt_lr_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(lr_dev, stats, desired_kl, inv_world_size);
  count_launch();
  return check_launch("adapt_lr");
}

int b200lab_grad_sumsq(const float* grad, int64_t n, float* out_sumsq, void* stream) {
  B200LAB_CHECK_ARG(grad && out_sumsq && n > 0, "bad argument");
  int nb = (int)min64(ceil_div(n, 256 * 4), 2 * kNumSMs);
  grad_sumsq_kernel<<<nb, 256, 0, (cudaStream_t)stream>>>(grad, n, out_sumsq);
  count_launch();
  return check_launch("grad_sumsq");
}

int b200lab_adam_clip_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n,
                           const float* lr_dev, const float* grad_sumsq, float grad_scale, float max_grad_norm,
                           float beta1, float beta2, float eps, int64_t step, void* stream) {
  B200LAB_CHECK_ARG(param && grad && exp_avg && exp_avg_sq && lr_dev && n > 0 && step > 0, "bad argument");
  B200LAB_CHECK_ARG(max_grad_norm <= 0.0f || grad_sumsq, "grad_sumsq required when clipping");
  float bc1 = 1.0f - (float)pow((double)beta1, (double)step);
  float bc2s = (float)sqrt(1.0 - pow((double)beta2, (double)step));
  int nb = (int)min64(ceil_div(n, 256 * 4), 4 * kNumSMs);
  adam_clip_kernel<<<nb, 256, 0, (cudaStream_t)stream>>>(param, grad, exp_avg, exp_avg_sq, n, lr_dev, grad_sumsq,
                                                         grad_scale, max_grad_norm, beta1, beta2, eps, bc1, bc2s);
  count_launch();
  return check_launch("adam_clip_step");
}
int b200lab_empirical_norm_update(const float* x, int64_t M, int32_t D, int64_t count_new, float* mean, float* var,
                                  float* std, void* stream) {
  B200LAB_CHECK_ARG(x && mean && var && std && M > 0 && D > 0 && count_new >= M, "bad argument");
  const float rate = (float)((double)M / (double)count_new);
  empirical_norm_update_kernel<<<(unsigned)ceil_div(D, 32), 256, 0, (cudaStream_t)stream>>>(x, M, D, rate, mean, var, std);
  count_launch();
  return check_launch("empirical_norm_update");
}

int b200lab_empirical_norm_partial(const float* x, int64_t M, int32_t D, double* sums, void* stream) {
  B200LAB_CHECK_ARG(x && sums && M > 0 && D > 0, "bad argument");
  empirical_norm_partial_kernel<<<(unsigned)ceil_div(D, 32), 256, 0, (cudaStream_t)stream>>>(x, M, D, sums);
  count_launch();
  return check_launch("empirical_norm_partial");
}

int b200lab_empirical_norm_merge(const double* sums, int32_t D, int64_t count_old, float* mean, float* var, float* std,
                                 void* stream) {
  B200LAB_CHECK_ARG(sums && mean && var && std && D > 0 && count_old >= 0, "bad argument");
  empirical_norm_merge_kernel<<<(unsigned)ceil_div(D, 256), 256, 0, (cudaStream_t)stream>>>(sums, D, count_old, mean, var, std);
  count_launch();
  return check_launch("empirical_norm_merge");
}

int b200lab_empirical_norm_apply(const float* x, const float* mean, const float* std, float eps, float* out, int64_t M,
                                 int32_t D, void* stream) {
  B200LAB_CHECK_ARG(x && mean && std && out && M > 0 && D > 0, "bad argument");
  int nb = (int)min64(ceil_div(M * D, 256), 16 * kNumSMs);
  empirical_norm_apply_kernel<<<nb, 256, 0, (cudaStream_t)stream>>>(x, mean, std, eps, out, M, D);
  count_launch();
  return check_launch("empirical_norm_apply");
}
}
