// common.cuh -- shared device helpers for libb200lab (sm_100a only).
//
// The math helpers restate the exact fp32 formulas of LAB/utils/math.py (operation order preserved so that
// float-derived masks are bit-exact; the MDP translation units are compiled with -fmad=false).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>

#include "../../include/b200lab.h"

namespace b200lab {

void set_error(const char* fmt, ...);
void count_launch(int n = 1);
int check_launch(const char* what);

#define B200LAB_CHECK_ARG(cond, msg)        \
  do {                                      \
    if (!(cond)) {                          \
      b200lab::set_error("%s: %s", __func__, msg); \
      return B200LAB_ERR_ARG;               \
    }                                       \
  } while (0)

constexpr int kNumSMs = 148;  // B200

__host__ __device__ inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
__host__ __device__ inline int64_t min64(int64_t a, int64_t b) { return a < b ? a : b; }
__device__ __forceinline__ float elu_f(float x) { return x > 0.0f ? x : expm1f(x); }   // nn.ELU, alpha = 1

struct f3 {
  float x, y, z;
};
struct f4 {
  float w, x, y, z;
};

__device__ __forceinline__ f3 cross3(f3 a, f3 b) {
  return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}

// LAB/utils/math.py:606-625  quat_rotate_inverse: a - b + c
__device__ __forceinline__ f3 quat_rotate_inverse(f4 q, f3 v) {
  float s = 2.0f * (q.w * q.w) - 1.0f;
  f3 qv{q.x, q.y, q.z};
  f3 cr = cross3(qv, v);
  float d = qv.x * v.x + qv.y * v.y + qv.z * v.z;
  f3 r;
  r.x = (v.x * s - cr.x * q.w * 2.0f) + qv.x * d * 2.0f;
  r.y = (v.y * s - cr.y * q.w * 2.0f) + qv.y * d * 2.0f;
  r.z = (v.z * s - cr.z * q.w * 2.0f) + qv.z * d * 2.0f;
  return r;
}

// LAB/utils/math.py:546-564  quat_apply: vec + w*t + xyz x t, t = 2 (xyz x vec)
__device__ __forceinline__ f3 quat_apply(f4 q, f3 v) {
  f3 qv{q.x, q.y, q.z};
  f3 t = cross3(qv, v);
  t.x *= 2.0f; t.y *= 2.0f; t.z *= 2.0f;
  f3 c = cross3(qv, t);
  return {(v.x + q.w * t.x) + c.x, (v.y + q.w * t.y) + c.y, (v.z + q.w * t.z) + c.z};
}

// LAB/utils/math.py:522-542  yaw_quat (+ normalize, math.py:82)
__device__ __forceinline__ f4 yaw_quat(f4 q) {
  float yaw = atan2f(2.0f * (q.w * q.z + q.x * q.y), 1.0f - 2.0f * (q.y * q.y + q.z * q.z));
  float sz = sinf(yaw / 2.0f);
  float cw = cosf(yaw / 2.0f);
  float n = sqrtf(cw * cw + 0.0f + 0.0f + sz * sz);
  n = fmaxf(n, 1e-9f);
  return {cw / n, 0.0f / n, 0.0f / n, sz / n};
}

// LAB/utils/math.py:96-117  wrap_to_pi
__device__ __forceinline__ float wrap_to_pi(float a) {
  const float pi = 3.14159265358979323846f;
  const float two_pi = 6.28318530717958647692f;
  float x = a + pi;
  float m = fmodf(x, two_pi);
  if (m != 0.0f && (m < 0.0f)) m += two_pi;  // torch.remainder: result takes the sign of the divisor
  return (m == 0.0f && a > 0.0f) ? pi : m - pi;
}

// LAB/utils/math.py:252-280 quat_from_euler_xyz
__device__ __forceinline__ f4 quat_from_euler_xyz(float roll, float pitch, float yaw) {
  float cy = cosf(yaw * 0.5f), sy = sinf(yaw * 0.5f);
  float cr = cosf(roll * 0.5f), sr = sinf(roll * 0.5f);
  float cp = cosf(pitch * 0.5f), sp = sinf(pitch * 0.5f);
  f4 q;
  q.w = cy * cr * cp + sy * sr * sp;
  q.x = cy * sr * cp - sy * cr * sp;
  q.y = cy * cr * sp + sy * sr * cp;
  q.z = sy * cr * cp - cy * sr * sp;
  return q;
}

// LAB/utils/math.py:464-500 quat_mul
__device__ __forceinline__ f4 quat_mul(f4 a, f4 b) {
  float ww = (a.z + a.x) * (b.x + b.y);
  float yy = (a.w - a.y) * (b.w + b.z);
  float zz = (a.w + a.y) * (b.w - b.z);
  float xx = ww + yy + zz;
  float qq = 0.5f * (xx + (a.z - a.x) * (b.x - b.y));
  f4 r;
  r.w = qq - ww + (a.z - a.y) * (b.y - b.z);
  r.x = qq - xx + (a.x + a.w) * (b.x + b.w);
  r.y = qq - yy + (a.w - a.x) * (b.y + b.z);
  r.z = qq - zz + (a.z + a.y) * (b.w - b.x);
  return r;
}

__device__ __forceinline__ float norm3(float x, float y, float z) { return sqrtf(x * x + y * y + z * z); }

__device__ __forceinline__ float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }

// ---- Philox4x32-10 counter RNG (perf mode; parity mode reads pre-drawn uniforms) ----
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += W0;
    key.y += W1;
  }
  return ctr;
}

// uniform in [0,1) for (seed, step, env, slot); `stream_id` separates u_reset / u_step / u_obs
__device__ __forceinline__ float philox_uniform(uint64_t seed, int64_t step, uint32_t stream_id, uint32_t env,
                                                uint32_t slot) {
  uint4 ctr = make_uint4(env, slot >> 2, (uint32_t)step, ((uint32_t)((uint64_t)step >> 32)) ^ (stream_id << 24));
  uint2 key = make_uint2((uint32_t)seed, (uint32_t)(seed >> 32));
  uint4 r = philox4x32_10(ctr, key);
  uint32_t v = (slot & 3) == 0 ? r.x : (slot & 3) == 1 ? r.y : (slot & 3) == 2 ? r.z : r.w;
  return (float)(v >> 8) * (1.0f / 16777216.0f);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ int warp_sum_int(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace b200lab
