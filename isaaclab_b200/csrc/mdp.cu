// mdp.cu -- the fused non-physics step of ManagerBasedRLEnv.step (LAB/envs/manager_based_rl_env.py:153-242).
//
// Kernels (one launch each, all on the caller's stream, no host sync anywhere):
//   process_action_kernel : ActionManager.process_action + JointAction.process_actions
//   substep_pre_kernel    : apply_action + Articulation._apply_actuator_model (Implicit/IdealPD/DCMotor/NetLSTM/NetMLP)
//   substep_post_kernel   : ArticulationData.update (joint_acc) + ContactSensor update + RayCaster timestamps
//   post_step_kernel      : terminations, rewards, masked _reset_idx, command compute, interval events (warp per env)
//   compute_obs_kernel    : ObservationManager.compute incl. the height-scan ray cast (mdp_obs section below)
//
// Memory: everything is env-major fp32 as the reference keeps it; each warp / thread group touches contiguous
// per-env rows so every sector read is fully used.  Compiled with -fmad=false (bit-exact masks); the tiny
// actuator nets use explicit fmaf().
#include <climits>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "common.cuh"
#include "raycast.cuh"

namespace b200lab {

struct EnvHandle {
  B200LabEnvState host;   // passed BY VALUE to the kernels whose descriptor reads are warp-uniform (__grid_constant__,
                          // 12 KB of the 32 KB parameter space): constant-bank operands instead of global round trips
  B200LabEnvState* dev;   // device copy (kept for callers that want the descriptor on the device)
  float* derived;      // [N, kDerived] per-env derived state for compute_obs
  bool rays_vertical;  // all ray directions are (0,0,dz), dz != 0, and the sensor is yaw-only
  int vert_batch;      // candidates per round trip of the vertical ray query (B200LAB_VERT_BATCH: 2 default, 3, 4)
  bool obs_plain;      // no observation term keeps history or carries modifiers
  int obs_impl;        // 0 (default): per-ray kernel; 1: staged height scan (compute_obs_patch_kernel), opt-in via B200LAB_OBS_IMPL
  struct ObsCol* cols; // [obs_dim] per-column program of the observation group (compute_obs_kernel)
  MeshAccel mesh;      // host copy of the ray-caster mesh descriptor (passed by value to compute_obs)
  bool has_mesh;
  // host copy of the ActuatorNetLSTM weights when their shape has a constant-bank instantiation (H=8, L=2: the ANYdrive
  // net) -- handed to the kernels as a launch parameter
  cudaStream_t side;       // second stream + fork/join events of b200lab_step_tail
  cudaEvent_t ev_fork, ev_join;
  // speculative height scan (b200lab_height_scan_prefetch): hit heights [N,R] and the sensor pose [N,8] they were cast
  // from, produced on a third stream while the decimation / post_step kernels run
  cudaStream_t side2;
  cudaEvent_t ev_spec_fork, ev_spec_done;
  float* spec_z;
  float* spec_pose;
  int64_t spec_cap;        // N*R the buffers were sized for
  bool spec_pending;       // a prefetch has been enqueued and not yet joined
  // post_step leaves the derived per-env quantities of its final state in `derived`; they stay valid -- and the
  // observation kernels use them instead of running derive_obs_kernel -- until the next entry point that advances the
  // env (process_action, decimation_fused, substep_pre/post, env_update)
  const float* derived_root;   // root_state_w pointer of the view post_step derived from (or NULL)
  bool lstm_const;
  bool mlp_const;          // the ActuatorNetMLP has the 6-32-32-32-1 shape of NetMlp63 (compile-time-shaped kernels)
  const float* lstm_src;   // device pointer the copy was taken from
  float lstm_w[2560];   // packed NetLstm<8,2> image (see NetLstm) or the raw NetMlp<6,32,3> weights
};

typedef B200LabEnvState Env;
typedef B200LabSimView Sim;

// Programmatic dependent launch (PDL): a kernel launched with the programmatic-stream-serialization attribute may start
// while its predecessor in the stream is still running; it must not touch anything the predecessor writes before
// pdl_wait() (griddepcontrol.wait: predecessor complete, its memory visible).  pdl_launch_dependents() in the predecessor
// lets the scheduler place the successor's blocks as soon as SM resources free up, so the successor's launch latency,
// parameter loads and predecessor-independent loads overlap the predecessor's tail.  Both are no-ops without the attribute.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

template <class... KArgs, class... Args>
static cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(std::forward<Args>(args))...);
}

// ------------------------------------------------------------------------------------------------
// process_action
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
process_action_kernel(const __grid_constant__ Env env_param, const float* __restrict__ in) {
  const Env* __restrict__ e = &env_param;

  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int A = e->A;
  int64_t total = (int64_t)e->N * A;
  if (i == 0) {  // first kernel of the step sequence
    if (e->step_counter_dev) *e->step_counter_dev += 1;
    if (e->any_reset_dev) *e->any_reset_dev = 0;
  }
  if (i >= total) return;
  int a = (int)(i % A);
  float x = in[i];
  e->prev_action[i] = e->action[i];  // action_manager.py:331
  e->action[i] = x;                  // :332
  e->raw_actions[i] = x;             // joint_actions.py:132
  float p = x * e->action_scale[a] + e->action_offset[a];  // joint_actions.py:134
  if (e->action_clip) p = fminf(fmaxf(p, e->action_clip[2 * a]), e->action_clip[2 * a + 1]);
  e->processed_actions[i] = p;
}

// ------------------------------------------------------------------------------------------------
// substep_pre: actuator models.  One thread per (env, joint).
// ------------------------------------------------------------------------------------------------
constexpr int kActBlock = 128;
constexpr int kMaxNetFloats = 8192;  // staged net weights (32 KB)

// LSTM gate activations on the SFU path (ex2.approx + fast divide): |error| ~ 1e-6 absolute on the gate values, far below
// the 2e-5 torque tolerance of the parity tests; the IEEE expf/tanhf/divide sequences were half of the actuator kernel's
// instructions (profiles/r1_notes.md).
__device__ __forceinline__ float ex2_ftz(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float rcp_ftz(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
// 4 instructions (FMUL, MUFU.EX2, FADD, MUFU.RCP); the non-ftz __expf/__fdividef forms add denormal range scaling
// around each MUFU (~2.5x the instructions) for inputs the gates never see
__device__ __forceinline__ float sigmoidf_(float x) { return rcp_ftz(1.0f + ex2_ftz(-1.4426950408889634f * x)); }
__device__ __forceinline__ float tanhf_(float x) {
  float e = ex2_ftz(2.8853900817779268f * fabsf(x));   // exp(2|x|); overflow -> inf -> rcp -> 0 -> t = 1
  float t = fmaf(-2.0f, rcp_ftz(e + 1.0f), 1.0f);
  return copysignf(t, x);
}

// Two gate activations behind ONE reciprocal: 1/a and 1/b from r = 1/(a*b) (r*b, r*a).  The cell is MUFU-pipe bound
// (ex2 + rcp per activation); pairing (input, forget), (output, tanh(cell gate)) and the tanh(c) of adjacent hidden units
// takes a cell from 10 to 7.5 MUFU per unit.  Exponents are clamped at 2^60 so the product stays finite (the clamp moves
// a sigmoid below 1e-18, far under the torque tolerance).
#ifndef B200LAB_GATE_PAIR
#define B200LAB_GATE_PAIR 1
#endif
__device__ __forceinline__ float gate_den(float x_scaled) { return 1.0f + ex2_ftz(fminf(x_scaled, 60.0f)); }
__device__ __forceinline__ void sigmoid_pair(float x, float y, float& sx, float& sy) {
  const float a = gate_den(-1.4426950408889634f * x), b = gate_den(-1.4426950408889634f * y);
  const float r = rcp_ftz(a * b);
  sx = r * b; sy = r * a;
}
__device__ __forceinline__ void sigmoid_tanh_pair(float x, float y, float& sx, float& ty) {
  const float a = gate_den(-1.4426950408889634f * x), b = gate_den(2.8853900817779268f * fabsf(y));
  const float r = rcp_ftz(a * b);
  sx = r * b;
  ty = copysignf(fmaf(-2.0f, r * a, 1.0f), y);
}
__device__ __forceinline__ void tanh_pair(float x, float y, float& tx, float& ty) {
  const float a = gate_den(2.8853900817779268f * fabsf(x)), b = gate_den(2.8853900817779268f * fabsf(y));
  const float r = rcp_ftz(a * b);
  tx = copysignf(fmaf(-2.0f, r * b, 1.0f), x);
  ty = copysignf(fmaf(-2.0f, r * a, 1.0f), y);
}

__device__ __forceinline__ float dc_motor_clip(float tau, float qd, float sat, float vlim, float elim) {
  // actuator_pd.py:276-286
  float max_e = sat * (1.0f - qd / vlim);
  max_e = fminf(fmaxf(max_e, 0.0f), elim);
  float min_e = sat * (-1.0f - qd / vlim);
  min_e = fminf(fmaxf(min_e, -elim), 0.0f);
  return fminf(fmaxf(tau, min_e), max_e);
}

// LSTM cell stack, hidden size H compile-time; weights in shared memory (broadcast reads).
template <int H>
__device__ __forceinline__ float lstm_forward(const float* __restrict__ w, int L, float x0, float x1,
                                              float* __restrict__ hs, float* __restrict__ cs, int64_t row,
                                              int64_t rows) {
  float xin[H];
  xin[0] = x0;
  xin[1] = x1;
  int in_dim = 2;
  for (int l = 0; l < L; ++l) {
    const float* w_ih = w;
    const float* w_hh = w_ih + 4 * H * in_dim;
    const float* b_ih = w_hh + 4 * H * H;
    const float* b_hh = b_ih + 4 * H;
    w = b_hh + 4 * H;
    float h[H], c[H];
    float* hp = hs + ((int64_t)l * rows + row) * H;
    float* cp = cs + ((int64_t)l * rows + row) * H;
#pragma unroll
    for (int k = 0; k < H; k += 4) {
      float4 a = *reinterpret_cast<const float4*>(hp + k);
      float4 b = *reinterpret_cast<const float4*>(cp + k);
      h[k] = a.x; h[k + 1] = a.y; h[k + 2] = a.z; h[k + 3] = a.w;
      c[k] = b.x; c[k + 1] = b.y; c[k + 2] = b.z; c[k + 3] = b.w;
    }
    float hn[H];
#pragma unroll
    for (int u = 0; u < H; ++u) {
      float g4[4];
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        int r = g * H + u;
        float acc = b_ih[r] + b_hh[r];
        if (l == 0) {
          acc = fmaf(w_ih[r * 2], xin[0], acc);
          acc = fmaf(w_ih[r * 2 + 1], xin[1], acc);
        } else {
#pragma unroll
          for (int k = 0; k < H; ++k) acc = fmaf(w_ih[r * H + k], xin[k], acc);
        }
#pragma unroll
        for (int k = 0; k < H; ++k) acc = fmaf(w_hh[r * H + k], h[k], acc);
        g4[g] = acc;
      }
      float ig = sigmoidf_(g4[0]), fg = sigmoidf_(g4[1]), gg = tanhf_(g4[2]), og = sigmoidf_(g4[3]);
      float cn = fmaf(fg, c[u], ig * gg);
      c[u] = cn;
      hn[u] = og * tanhf_(cn);
    }
#pragma unroll
    for (int k = 0; k < H; k += 4) {
      *reinterpret_cast<float4*>(hp + k) = make_float4(hn[k], hn[k + 1], hn[k + 2], hn[k + 3]);
      *reinterpret_cast<float4*>(cp + k) = make_float4(c[k], c[k + 1], c[k + 2], c[k + 3]);
    }
#pragma unroll
    for (int k = 0; k < H; ++k) xin[k] = hn[k];
    in_dim = H;
  }
  // head: Linear(H -> 1)
  float out = w[H];
#pragma unroll
  for (int k = 0; k < H; ++k) out = fmaf(w[k], xin[k], out);
  return out;
}

// ---- actuator-net weights as a KERNEL PARAMETER (constant bank, c[0][imm] operands) ---------------------------------
// Every lane of a warp reads the same weight.  From shared memory that is one LDS per FMA, and the single LSU port per SM
// (vs. four FMA pipes) makes the cell LDS-bound; as immediate-offset constant-bank operands the weights ride on the FFMA
// itself and no load instruction is issued.  All indices below are compile-time after unrolling (H, L template params).
struct NetNone { static constexpr bool kLstm = false, kMlp = false; };

// Packed layout for the two-wide FMA (fma.rn.f32x2 -> FFMA2, one FMA-pipe slot for two IEEE fp32 FMAs; the cell is
// FMA-pipe bound).  Two gate rows of the same hidden unit share their input vector, so the rows are paired --
// (input, forget) and (cell, output) -- and their weights interleaved: per layer l, per unit u, per pair:
//   {bias_a, bias_b} with bias = b_ih + b_hh (summed in fp32 on the host: the same IEEE add the kernel did), then
//   in_l x {w_ih[a][k], w_ih[b][k]}, then H x {w_hh[a][k], w_hh[b][k]}.
// Every half of an FFMA2 is the scalar fmaf of the unpacked code, in the same order: results are bitwise unchanged.
template <int H, int L>
struct NetLstm {
  static constexpr bool kLstm = true, kMlp = false;
  static constexpr int kH = H, kL = L;
  static constexpr int kRawCount = (4 * H * 2 + 4 * H * H + 8 * H) + (L - 1) * (8 * H * H + 8 * H) + H + 1;
  static constexpr int in_dim(int l) { return l == 0 ? 2 : H; }
  static constexpr int pair_len(int l) { return 1 + in_dim(l) + H; }                 // float2 per (unit, pair)
  static constexpr int layer_off(int l) { return l == 0 ? 0 : 2 * H * pair_len(0) + (l - 1) * 2 * H * pair_len(1); }
  static constexpr int kPairs = layer_off(L);
  float2 w2[kPairs];
  float head[H + 1];
};

__device__ __forceinline__ uint64_t pack2(float a, float b) { uint64_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void unpack2(uint64_t v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}

template <class Net>
__device__ __forceinline__ float lstm_cells_const(const Net& net, float x0, float x1, float (&h)[Net::kL][Net::kH],
                                                  float (&c)[Net::kL][Net::kH]) {
  constexpr int H = Net::kH, L = Net::kL;
  float xin[H];
  xin[0] = x0;
  xin[1] = x1;
#pragma unroll
  for (int l = 0; l < L; ++l) {
    const int in_dim = Net::in_dim(l), plen = Net::pair_len(l);
    float hn[H];
#pragma unroll
    for (int u = 0; u < H; ++u) {
      float g4[4];
#pragma unroll
      for (int pr = 0; pr < 2; ++pr) {
        const int o = Net::layer_off(l) + (u * 2 + pr) * plen;
        uint64_t acc = pack2(net.w2[o].x, net.w2[o].y);
#pragma unroll
        for (int k = 0; k < in_dim; ++k) acc = fma2(pack2(net.w2[o + 1 + k].x, net.w2[o + 1 + k].y), pack2(xin[k], xin[k]), acc);
#pragma unroll
        for (int k = 0; k < H; ++k)
          acc = fma2(pack2(net.w2[o + 1 + in_dim + k].x, net.w2[o + 1 + in_dim + k].y), pack2(h[l][k], h[l][k]), acc);
        unpack2(acc, g4[2 * pr], g4[2 * pr + 1]);
      }
#if B200LAB_GATE_PAIR
      float ig, fg, gg, og;
      sigmoid_pair(g4[0], g4[1], ig, fg);
      sigmoid_tanh_pair(g4[3], g4[2], og, gg);
      float cn = fmaf(fg, c[l][u], ig * gg);
      c[l][u] = cn;
      hn[u] = og;   // multiplied by tanh(c) below, two units per reciprocal
    }
#pragma unroll
    for (int u = 0; u < H; u += 2) {
      float t0, t1;
      tanh_pair(c[l][u], c[l][u + 1], t0, t1);
      hn[u] *= t0; hn[u + 1] *= t1;
    }
#else
      float ig = sigmoidf_(g4[0]), fg = sigmoidf_(g4[1]), gg = tanhf_(g4[2]), og = sigmoidf_(g4[3]);
      float cn = fmaf(fg, c[l][u], ig * gg);
      c[l][u] = cn;
      hn[u] = og * tanhf_(cn);
    }
#endif
#pragma unroll
    for (int k = 0; k < H; ++k) { h[l][k] = hn[k]; xin[k] = hn[k]; }
  }
  float out = net.head[H];
#pragma unroll
  for (int k = 0; k < H; ++k) out = fmaf(net.head[k], xin[k], out);
  return out;
}

template <class Net>
__device__ __forceinline__ void lstm_state_load(const B200LabActuatorGroup& g, int64_t row, int64_t rows,
                                                float (&h)[Net::kL][Net::kH], float (&c)[Net::kL][Net::kH]) {
#pragma unroll
  for (int l = 0; l < Net::kL; ++l)
#pragma unroll
    for (int k = 0; k < Net::kH; k += 4) {
      float4 a = *reinterpret_cast<const float4*>(g.lstm_h + ((int64_t)l * rows + row) * Net::kH + k);
      float4 b = *reinterpret_cast<const float4*>(g.lstm_c + ((int64_t)l * rows + row) * Net::kH + k);
      h[l][k] = a.x; h[l][k + 1] = a.y; h[l][k + 2] = a.z; h[l][k + 3] = a.w;
      c[l][k] = b.x; c[l][k + 1] = b.y; c[l][k + 2] = b.z; c[l][k + 3] = b.w;
    }
}

template <class Net>
__device__ __forceinline__ void lstm_state_store(const B200LabActuatorGroup& g, int64_t row, int64_t rows,
                                                 const float (&h)[Net::kL][Net::kH], const float (&c)[Net::kL][Net::kH]) {
#pragma unroll
  for (int l = 0; l < Net::kL; ++l)
#pragma unroll
    for (int k = 0; k < Net::kH; k += 4) {
      *reinterpret_cast<float4*>(g.lstm_h + ((int64_t)l * rows + row) * Net::kH + k) =
          make_float4(h[l][k], h[l][k + 1], h[l][k + 2], h[l][k + 3]);
      *reinterpret_cast<float4*>(g.lstm_c + ((int64_t)l * rows + row) * Net::kH + k) =
          make_float4(c[l][k], c[l][k + 1], c[l][k + 2], c[l][k + 3]);
    }
}

__device__ __forceinline__ float mlp_act(float x, int kind) {
  switch (kind) {
    case 0: return x / (1.0f + fabsf(x));            // softsign
    case 1: return x > 0.0f ? x : expm1f(x);         // elu
    case 2: return fmaxf(x, 0.0f);                   // relu
    default: return tanhf(x);
  }
}

// MLP with all widths <= 32; activations in registers, weights in shared memory.
__device__ __forceinline__ float mlp_forward32(const float* __restrict__ w, const B200LabActuatorGroup& g,
                                               const float* in) {
  float a[32], b[32];
  int d0 = g.mlp_dims[0];
#pragma unroll
  for (int i = 0; i < 32; ++i) a[i] = i < d0 ? in[i] : 0.0f;
  for (int l = 0; l < g.mlp_layers; ++l) {
    int din = g.mlp_dims[l], dout = g.mlp_dims[l + 1];
    const float* W = w;
    const float* bias = W + din * dout;
    w = bias + dout;
    bool last = (l == g.mlp_layers - 1);
#pragma unroll
    for (int o = 0; o < 32; ++o) {
      float acc = 0.0f;
      if (o < dout) {
        acc = bias[o];
#pragma unroll
        for (int i = 0; i < 32; ++i)
          if (i < din) acc = fmaf(W[o * din + i], a[i], acc);
        if (!last) acc = mlp_act(acc, g.mlp_act);
      }
      b[o] = acc;
    }
#pragma unroll
    for (int i = 0; i < 32; ++i) a[i] = b[i];
  }
  return a[0];
}

// ---- ActuatorNetMLP with compile-time shape, weights as a kernel parameter (constant bank) -------------------------------
// IN inputs -> NH hidden layers of width H -> 1 (the walk-these-ways actuator net of the Unitree configs: 6-32-32-32-1).
// Raw layout as the host packs it (b200lab.h): W0[H,IN] b0[H] W1[H,H] b1[H] ... Wout[1,H] bout[1].  The accumulation order
// is the one of mlp_forward32 (acc = bias; acc = fmaf(W[o][i], a[i], acc) for i ascending) => bitwise the same torques;
// what changes is that no load instruction is issued per FMA and no predicated padding up to 32 x 32 is executed.
template <int IN, int H, int NH>
struct NetMlp {
  static constexpr bool kLstm = false, kMlp = true;
  static constexpr int kIn = IN, kH = H, kNH = NH;
  static constexpr int kCount = IN * H + H + (NH - 1) * (H * H + H) + H + 1;
  // shape tag only: 9.5 KB of weights as constant-bank operands thrash the constant cache (measured: 461 us per fused
  // launch at 16384 envs), so the weights are staged in shared memory and the H-wide rows are read with 128-bit loads
};

// shared-memory weight loads as volatile asm: the weights are invariant across the sub-step loop, and the compiler would
// otherwise hoist all 2369 of them out of it and spill them to local memory (9.5 KB stack frame)
__device__ __forceinline__ float lds1(const float* p) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"((unsigned)__cvta_generic_to_shared(p)));
  return v;
}
__device__ __forceinline__ float4 lds4(const float* p) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"((unsigned)__cvta_generic_to_shared(p)));
  return v;
}

// softsign through the SFU reciprocal (1 ulp-level error; torque tolerance of the parity tests is 2e-5, as for the
// LSTM's SFU gate activations): the IEEE division was a third of the net's instructions
__device__ __forceinline__ float mlp_act_fast(float x, int kind) {
  if (kind == 0) return x * rcp_ftz(1.0f + fabsf(x));
  return mlp_act(x, kind);
}

// `sb`: this thread's column of a [H][stride] shared-memory scratch (element o at sb[o * stride]; thread-contiguous =>
// conflict-free).  The loop over output neurons is ROLLED -- fully unrolled the net is ~70 KB of code and the kernel
// stalls on instruction fetch (ncu: no_instruction 5.8 per issue, issue-active 29 %) -- so a layer's outputs are parked in
// `sb` and re-read into registers for the next layer's unrolled 128-bit-weight inner products.
#ifndef B200LAB_MLP_NEURONS
#define B200LAB_MLP_NEURONS 4
#endif
constexpr int kMlpNeurons = B200LAB_MLP_NEURONS;

template <class Net>
__device__ __forceinline__ float mlp_const(const float* w, float* sb, int stride, const float (&in)[Net::kIn], int act) {
  constexpr int IN = Net::kIn, H = Net::kH, NH = Net::kNH;
  static_assert(H % 4 == 0 && (IN * H + H) % 4 == 0, "128-bit weight rows");
  static_assert(H % kMlpNeurons == 0, "neurons per trip must divide the layer width");
  float a[H];
#pragma unroll 1
  for (int o = 0; o < H; ++o) {
    float acc = lds1(w + IN * H + o);
#pragma unroll
    for (int i = 0; i < IN; ++i) acc = fmaf(lds1(w + o * IN + i), in[i], acc);
    sb[o * stride] = mlp_act_fast(acc, act);
  }
#pragma unroll 1
  for (int l = 1; l < NH; ++l) {
    const float* wl = w + (IN * H + H) + (l - 1) * (H * H + H);
#pragma unroll
    for (int i = 0; i < H; ++i) a[i] = sb[i * stride];
    // kMlpNeurons output neurons per trip of the rolled loop: their inner products are independent FMA chains (each in
    // the reference order, i ascending), so the 4-cycle dependent-issue latency of one chain is covered by the others
#pragma unroll 1
    for (int o = 0; o < H; o += kMlpNeurons) {
      float acc[kMlpNeurons];
#pragma unroll
      for (int u = 0; u < kMlpNeurons; ++u) acc[u] = lds1(wl + H * H + o + u);
#pragma unroll
      for (int i = 0; i < H; i += 4) {
        float4 w4[kMlpNeurons];
#pragma unroll
        for (int u = 0; u < kMlpNeurons; ++u) w4[u] = lds4(wl + (o + u) * H + i);
#pragma unroll
        for (int u = 0; u < kMlpNeurons; ++u) acc[u] = fmaf(w4[u].x, a[i], acc[u]);
#pragma unroll
        for (int u = 0; u < kMlpNeurons; ++u) acc[u] = fmaf(w4[u].y, a[i + 1], acc[u]);
#pragma unroll
        for (int u = 0; u < kMlpNeurons; ++u) acc[u] = fmaf(w4[u].z, a[i + 2], acc[u]);
#pragma unroll
        for (int u = 0; u < kMlpNeurons; ++u) acc[u] = fmaf(w4[u].w, a[i + 3], acc[u]);
      }
#pragma unroll
      for (int u = 0; u < kMlpNeurons; ++u) sb[(o + u) * stride] = mlp_act_fast(acc[u], act);
    }
  }
#pragma unroll
  for (int i = 0; i < H; ++i) a[i] = sb[i * stride];
  const float* wo = w + (IN * H + H) + (NH - 1) * (H * H + H);
  float out = lds1(wo + H);
#pragma unroll
  for (int i = 0; i < H; i += 4) {
    const float4 w4 = lds4(wo + i);
    out = fmaf(w4.x, a[i], out); out = fmaf(w4.y, a[i + 1], out); out = fmaf(w4.z, a[i + 2], out); out = fmaf(w4.w, a[i + 3], out);
  }
  return out;
}

constexpr int kMlpMaxHist = 4;
__device__ __forceinline__ float hist_at(const float (&h)[kMlpMaxHist], int idx) {
  return idx == 0 ? h[0] : idx == 1 ? h[1] : idx == 2 ? h[2] : h[3];
}

// one ActuatorNetMLP evaluation on register-resident histories (actuator_net.py:147-189): roll, insert, select, scale, net
template <class Net>
__device__ __forceinline__ float mlp_actuator_step(const float* w, float* sb, int stride, const B200LabActuatorGroup& g, float err_p, float qd,
                                                   float (&ph)[kMlpMaxHist], float (&vh)[kMlpMaxHist]) {
#pragma unroll
  for (int k = kMlpMaxHist - 1; k > 0; --k) { ph[k] = ph[k - 1]; vh[k] = vh[k - 1]; }   // slots >= hist are never read
  ph[0] = err_p; vh[0] = qd;
  constexpr int NI = Net::kIn / 2;
  float in[Net::kIn];
#pragma unroll
  for (int k = 0; k < NI; ++k) {
    const float pv = hist_at(ph, g.mlp_input_idx[k]) * g.mlp_pos_scale;
    const float vv = hist_at(vh, g.mlp_input_idx[k]) * g.mlp_vel_scale;
    if (g.mlp_vel_pos_order) { in[k] = vv; in[NI + k] = pv; } else { in[k] = pv; in[NI + k] = vv; }
  }
  return mlp_const<Net>(w, sb, stride, in, g.mlp_act) * g.mlp_torque_scale;
}

template <int HIST_MAX = kMlpMaxHist>
__device__ __forceinline__ void mlp_hist_load(const B200LabActuatorGroup& g, int n, int jl, float (&ph)[HIST_MAX], float (&vh)[HIST_MAX]) {
  const int Jg = g.group_joints, Hh = g.mlp_hist;
#pragma unroll
  for (int k = 0; k < HIST_MAX; ++k) {
    ph[k] = vh[k] = 0.f;
    if (k < Hh) {
      ph[k] = g.mlp_pos_hist[((int64_t)n * Hh + k) * Jg + jl];
      vh[k] = g.mlp_vel_hist[((int64_t)n * Hh + k) * Jg + jl];
    }
  }
}
template <int HIST_MAX = kMlpMaxHist>
__device__ __forceinline__ void mlp_hist_store(const B200LabActuatorGroup& g, int n, int jl, const float (&ph)[HIST_MAX], const float (&vh)[HIST_MAX]) {
  const int Jg = g.group_joints, Hh = g.mlp_hist;
#pragma unroll
  for (int k = 0; k < HIST_MAX; ++k)
    if (k < Hh) {
      g.mlp_pos_hist[((int64_t)n * Hh + k) * Jg + jl] = ph[k];
      g.mlp_vel_hist[((int64_t)n * Hh + k) * Jg + jl] = vh[k];
    }
}

// stage the weights of the (single) net-based actuator group, if any, in shared memory (block-cooperative)
__device__ __forceinline__ void stage_net_weights(const Env* __restrict__ e, float* s_w) {
  int net_group = -1;
  for (int g = 0; g < e->num_act_groups; ++g)
    if (e->act[g].kind == B200LAB_ACT_NET_LSTM || e->act[g].kind == B200LAB_ACT_NET_MLP) net_group = g;
  if (net_group >= 0) {
    const B200LabActuatorGroup& g = e->act[net_group];
    int nw = 0;
    if (g.kind == B200LAB_ACT_NET_LSTM) {
      int H = g.lstm_hidden;
      for (int l = 0; l < g.lstm_layers; ++l) nw += 4 * H * (l == 0 ? 2 : H) + 4 * H * H + 8 * H;
      nw += H + 1;
    } else {
      for (int l = 0; l < g.mlp_layers; ++l) nw += g.mlp_dims[l] * g.mlp_dims[l + 1] + g.mlp_dims[l + 1];
    }
    for (int i = threadIdx.x; i < nw; i += blockDim.x) s_w[i] = g.net_weights[i];
  }
  __syncthreads();
}

// One actuator evaluation for joint j of env n (Articulation._apply_actuator_model, articulation.py:1432-1469).
// q, qd: joint state BEFORE the physics step; q_t, qd_t, ff: the joint targets.
__device__ __noinline__ void actuator_eval(const Env* __restrict__ e, const float* s_w, int n, int j, float q, float qd,
                                              float q_t, float qd_t, float ff, float& computed, float& applied) {
  const int N = e->N;
  const B200LabActuatorGroup& g = e->act[e->joint_group[j]];
  if (g.kind == B200LAB_ACT_DELAYED_PD || g.kind == B200LAB_ACT_REMOTIZED_PD) {
    // DelayBuffer.compute on the three set-points (actuator_pd.py:336-344, delay_buffer.py:150-176): push, then read the
    // row pushed `lag` steps ago (capped at what has been pushed since the reset; the first push fills the ring)
    const int L = g.max_delay + 1;
    const int64_t row = (int64_t)n * g.group_joints + e->joint_local[j];
    float* ring = g.delay_ring + row * L * 3;
    int cnt = g.delay_count[row];
    const int slot = cnt % L;
    if (cnt == 0) {
      for (int k = 0; k < L; ++k) { ring[k * 3] = q_t; ring[k * 3 + 1] = qd_t; ring[k * 3 + 2] = ff; }
    } else {
      ring[slot * 3] = q_t; ring[slot * 3 + 1] = qd_t; ring[slot * 3 + 2] = ff;
    }
    const int lag = min(g.delay_lag[n], cnt);
    const int rd = (slot - lag + L) % L;
    q_t = ring[rd * 3]; qd_t = ring[rd * 3 + 1]; ff = ring[rd * 3 + 2];
    cnt += 1;
    if (cnt >= 2 * L) cnt -= L;   // keeps cnt mod L and cnt >= L >= lag: no overflow however long the episode
    g.delay_count[row] = cnt;
  }
  float err_p = q_t - q, err_v = qd_t - qd;
  float elim = g.effort_limit[j];
  if (g.kind == B200LAB_ACT_IMPLICIT || g.kind == B200LAB_ACT_IDEAL_PD || g.kind == B200LAB_ACT_DC_MOTOR ||
      g.kind == B200LAB_ACT_DELAYED_PD || g.kind == B200LAB_ACT_REMOTIZED_PD) {
    computed = (g.stiffness[j] * err_p + g.damping[j] * err_v) + ff;  // actuator_pd.py:136 / :190
    if (g.kind == B200LAB_ACT_DC_MOTOR)
      applied = dc_motor_clip(computed, qd, g.saturation_effort, g.velocity_limit[j], elim);
    else
      applied = fminf(fmaxf(computed, -elim), elim);  // actuator_base.py:309-318
    if (g.kind == B200LAB_ACT_REMOTIZED_PD) {
      // LinearInterpolation.compute(joint_pos): samples below q -> bracket -> weight, zero-order hold outside the table
      int ns = 0;
      for (int i = 0; i < g.lookup_n; ++i) ns += g.lookup_x[i] < q ? 1 : 0;
      const int lo = max(ns - 1, 0), hi = min(ns, g.lookup_n - 1);
      const float xl = g.lookup_x[lo], xh = g.lookup_x[hi], yl = g.lookup_y[lo], yh = g.lookup_y[hi];
      const float w = hi == lo ? 0.0f : (q - xl) / (xh - xl);
      const float lim = yl + w * (yh - yl);
      applied = fminf(fmaxf(applied, -lim), lim);   // actuator_pd.py:411-414
    }
  } else if (g.kind == B200LAB_ACT_NET_LSTM) {
    int64_t rows = (int64_t)N * g.group_joints;
    int64_t row = (int64_t)n * g.group_joints + e->joint_local[j];
    float t;
    if (g.lstm_hidden == 8) t = lstm_forward<8>(s_w, g.lstm_layers, err_p, qd, g.lstm_h, g.lstm_c, row, rows);
    else if (g.lstm_hidden == 16) t = lstm_forward<16>(s_w, g.lstm_layers, err_p, qd, g.lstm_h, g.lstm_c, row, rows);
    else t = lstm_forward<4>(s_w, g.lstm_layers, err_p, qd, g.lstm_h, g.lstm_c, row, rows);
    computed = t;  // actuator_net.py:90
    applied = dc_motor_clip(computed, qd, g.saturation_effort, g.velocity_limit[j], elim);
  } else {  // B200LAB_ACT_NET_MLP, actuator_net.py:147-189
    int Jg = g.group_joints, Hh = g.mlp_hist, jl = e->joint_local[j];
    float* ph = g.mlp_pos_hist + (int64_t)n * Hh * Jg + jl;
    float* vh = g.mlp_vel_hist + (int64_t)n * Hh * Jg + jl;
    // roll(1, dim=1): slot k <- slot k-1 ; slot 0 <- newest
    for (int k = Hh - 1; k > 0; --k) {
      ph[(int64_t)k * Jg] = ph[(int64_t)(k - 1) * Jg];
      vh[(int64_t)k * Jg] = vh[(int64_t)(k - 1) * Jg];
    }
    ph[0] = err_p;
    vh[0] = qd;
    float in[32];
    int ni = g.mlp_num_idx;
    for (int k = 0; k < ni; ++k) {
      float pv = ph[(int64_t)g.mlp_input_idx[k] * Jg] * g.mlp_pos_scale;
      float vv = vh[(int64_t)g.mlp_input_idx[k] * Jg] * g.mlp_vel_scale;
      if (g.mlp_vel_pos_order) { in[k] = vv; in[ni + k] = pv; }
      else { in[k] = pv; in[ni + k] = vv; }
    }
    computed = mlp_forward32(s_w, g, in) * g.mlp_torque_scale;
    applied = dc_motor_clip(computed, qd, g.saturation_effort, g.velocity_limit[j], elim);
  }
}

template <class Net>
__global__ void __launch_bounds__(kActBlock)
substep_pre_kernel(const __grid_constant__ Env env_param, Sim sim, bool apply_action, bool only_if_any_reset,
                   const __grid_constant__ Net net) {
  const Env* __restrict__ e = &env_param;
  extern __shared__ float s_w[];
  const int J = e->J, N = e->N;
  if (only_if_any_reset && *e->any_reset_dev == 0) return;   // uniform over the grid
  if constexpr (!Net::kLstm) stage_net_weights(e, s_w);
  int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= (int64_t)N * J) return;
  int n = (int)(gid / J), j = (int)(gid % J);
  // action_manager.apply_action -> set_joint_{position,effort}_target (joint_actions.py:158-160, 223-225)
  int col = e->joint_action_col[j];
  if (apply_action && col >= 0) {
    float p = e->processed_actions[(int64_t)n * e->A + col];
    if (e->action_kind == 0) e->joint_pos_target[gid] = p;
    else e->joint_effort_target[gid] = p;
  }
  int gi = e->joint_group[j];
  if (gi < 0) return;
  const B200LabActuatorGroup& g = e->act[gi];
  float q_t = e->joint_pos_target[gid], qd_t = e->joint_vel_target[gid], ff = e->joint_effort_target[gid];
  float computed, applied;
  bool done = false;
  if constexpr (Net::kLstm) {
    if (g.kind == B200LAB_ACT_NET_LSTM) {
      float h[Net::kL][Net::kH], c[Net::kL][Net::kH];
      const int64_t rows = (int64_t)N * g.group_joints, row = (int64_t)n * g.group_joints + e->joint_local[j];
      lstm_state_load<Net>(g, row, rows, h, c);
      const float qd = sim.joint_vel[gid];
      computed = lstm_cells_const(net, q_t - sim.joint_pos[gid], qd, h, c);   // actuator_net.py:76-90
      applied = dc_motor_clip(computed, qd, g.saturation_effort, g.velocity_limit[j], g.effort_limit[j]);
      lstm_state_store<Net>(g, row, rows, h, c);
      done = true;
    }
  }
  if constexpr (Net::kMlp) {
    if (g.kind == B200LAB_ACT_NET_MLP) {
      float ph[kMlpMaxHist], vh[kMlpMaxHist];
      const int jl = e->joint_local[j];
      mlp_hist_load(g, n, jl, ph, vh);
      const float qd = sim.joint_vel[gid];
      computed = mlp_actuator_step<Net>(s_w, s_w + ((Net::kCount + 3) & ~3) + threadIdx.x, (int)blockDim.x, g, q_t - sim.joint_pos[gid], qd, ph, vh);
      applied = dc_motor_clip(computed, qd, g.saturation_effort, g.velocity_limit[j], g.effort_limit[j]);
      mlp_hist_store(g, n, jl, ph, vh);
      done = true;
    }
  }
  if (!done) actuator_eval(e, s_w, n, j, sim.joint_pos[gid], sim.joint_vel[gid], q_t, qd_t, ff, computed, applied);
  e->computed_torque[gid] = computed;
  e->applied_torque[gid] = applied;
  if (g.kind == B200LAB_ACT_IMPLICIT) {
    e->joint_pos_target_sim[gid] = q_t;      // control_action returned unchanged (actuator_pd.py:140)
    e->joint_effort_target_sim[gid] = ff;
  } else {
    e->joint_effort_target_sim[gid] = applied;  // articulation.py:1460
  }
}


// ------------------------------------------------------------------------------------------------
// substep_post: scene.update(dt) -- ArticulationData.update (joint_acc, articulation_data.py:78-83,547-556),
// SensorBase.update (sensor_base.py:197-205) for the contact sensor (recomputed every sub-step because
// history_length > 0) and the ray caster (only marked outdated; lazily recomputed by compute_obs).
// Thread layout: blocks own whole envs; W = max(J, B) lanes per env.
// ------------------------------------------------------------------------------------------------
constexpr int kPostBlock = 256;

__global__ void __launch_bounds__(kPostBlock)
substep_post_kernel(const __grid_constant__ Env env_param, Sim sim, int W, int envs_per_block) {
  const Env* __restrict__ e = &env_param;
  const int N = e->N, J = e->J, B = e->B;
  int el = threadIdx.x / W, i = threadIdx.x % W;
  int n = blockIdx.x * envs_per_block + el;
  bool active = el < envs_per_block && n < N;
  float dt = e->physics_dt;
  float ts = 0.f, tl = 0.f;
  bool outdated = false;
  if (active) {
    if (i < J) {
      // joint_acc = (joint_vel - previous_joint_vel) / time_elapsed ; previous <- joint_vel
      int64_t o = (int64_t)n * J + i;
      float qd = sim.joint_vel[o];
      e->joint_acc[o] = (qd - e->prev_joint_vel[o]) / dt;
      e->prev_joint_vel[o] = qd;
    }
    if (e->cs_timestamp) {
      ts = e->cs_timestamp[n] + dt;                       // sensor_base.py:199
      tl = e->cs_timestamp_last_update[n];
      outdated = e->cs_is_outdated[n] || ((ts - tl) + 1e-6f >= e->cs_update_period);  // :200
    }
  }
  __syncthreads();  // every lane of the env has read the env-level scalars
  if (!active) return;
  if (e->cs_timestamp && i < B) {
    if (outdated) {
      int64_t ob = (int64_t)n * B + i;
      float fx = 0.f, fy = 0.f, fz = 0.f;
      if (sim.net_forces_w) { fx = sim.net_forces_w[ob * 3]; fy = sim.net_forces_w[ob * 3 + 1]; fz = sim.net_forces_w[ob * 3 + 2]; }
      e->cs_net_forces_w[ob * 3] = fx; e->cs_net_forces_w[ob * 3 + 1] = fy; e->cs_net_forces_w[ob * 3 + 2] = fz;
      // history shift (contact_sensor.py:333-335)
      int Hh = e->hist;
      float* hp = e->cs_forces_hist + ((int64_t)n * Hh * B + i) * 3;
      for (int k = Hh - 1; k > 0; --k) {
        float* dst = hp + (int64_t)k * B * 3;
        const float* src = hp + (int64_t)(k - 1) * B * 3;
        dst[0] = src[0]; dst[1] = src[1]; dst[2] = src[2];
      }
      hp[0] = fx; hp[1] = fy; hp[2] = fz;
      if (e->cs_current_air_time) {
        float elapsed = ts - tl;  // contact_sensor.py:353
        bool is_contact = norm3(fx, fy, fz) > e->cs_force_threshold;
        float ca = e->cs_current_air_time[ob], cc = e->cs_current_contact_time[ob];
        bool first_contact = (ca > 0.f) && is_contact;
        bool first_detached = (cc > 0.f) && !is_contact;
        if (first_contact) e->cs_last_air_time[ob] = ca + elapsed;
        e->cs_current_air_time[ob] = !is_contact ? ca + elapsed : 0.f;
        if (first_detached) e->cs_last_contact_time[ob] = cc + elapsed;
        e->cs_current_contact_time[ob] = is_contact ? cc + elapsed : 0.f;
      }
    }
  }
  if (i == 0) {
    if (e->cs_timestamp) {
      e->cs_timestamp[n] = ts;
      if (outdated) e->cs_timestamp_last_update[n] = ts;  // sensor_base.py:294
      e->cs_is_outdated[n] = 0;
    }
    if (e->rc_timestamp) {
      float rts = e->rc_timestamp[n] + dt;
      e->rc_timestamp[n] = rts;
      if ((rts - e->rc_timestamp_last_update[n]) + 1e-6f >= e->rc_update_period) e->rc_is_outdated[n] = 1;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// decimation_fused: the whole physics loop of one env step in ONE launch, for replayed simulator state
// (recorded or synthetic: all S post-physics states are known up front, manager_based_rl_env.py:173-196):
//   process_action, then for s in 0..S-1: apply_action + actuator model on state s-1, scene.update on state s.
// Same arithmetic and the same per-sub-step order as process_action / substep_pre / substep_post above; what changes
// is data movement: action processing, joint targets, previous joint velocity, contact history and air-time state stay
// in registers across the S sub-steps and are written once.  Thread layout as substep_post (W = max(J,B) lanes / env).
// ------------------------------------------------------------------------------------------------
constexpr int kMaxDecimation = 8;
struct SimViews {
  Sim prev;                  // state before the first physics step
  Sim v[kMaxDecimation];     // state after each physics step
  int S;
};

#ifndef B200LAB_FUSED_THREADS
#define B200LAB_FUSED_THREADS 128
#endif
constexpr int kFusedBlock = B200LAB_FUSED_THREADS;   // 512 threads / SM at <= 128 registers: the body-role blocks are pure memory latency

template <class Net>
__global__ void __launch_bounds__(kFusedBlock, 512 / kFusedBlock)
decimation_fused_kernel(const __grid_constant__ Env env_param, const float* __restrict__ action_in,
                        const __grid_constant__ SimViews sv, int joint_blocks, int W, int envs_per_block,
                        const __grid_constant__ Net net) {
  const Env* __restrict__ e = &env_param;
  pdl_launch_dependents();   // post_step may start its prologue under this kernel's tail (it waits before consuming)
  extern __shared__ float s_w[];
  const int N = e->N, J = e->J, B = e->B, A = e->A, S = sv.S;
  const float dt = e->physics_dt;
  // Blocks [0, joint_blocks): one thread per (env, joint), dense.  Blocks [joint_blocks, ...): whole envs, W = B lanes per
  // env for the contact sensor + the env-level sensor clocks (they need a block barrier; see the end of the kernel).
  int n, i;
  const bool joint_role = (int)blockIdx.x < joint_blocks;
  if (joint_role) {
    if constexpr (!Net::kLstm) stage_net_weights(e, s_w);
    int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= (int64_t)N * J) return;
    n = (int)(gid / J); i = (int)(gid % J);
  } else {
    int el = threadIdx.x / W;
    i = threadIdx.x % W;
    n = ((int)blockIdx.x - joint_blocks) * envs_per_block + el;
    if (blockIdx.x == joint_blocks && threadIdx.x == 0) {
      if (e->step_counter_dev) *e->step_counter_dev += 1;
      if (e->any_reset_dev) *e->any_reset_dev = 0;
    }
    if (el >= envs_per_block || n >= N) return;
  }
  // ---- joint role
  if (joint_role) {
    const int j = i;
    const int64_t o = (int64_t)n * J + j;
    // ActionManager.process_action + JointAction.process_actions for the action column driving this joint
    int col = e->joint_action_col[j];
    if (col >= 0) {
      int64_t ao = (int64_t)n * A + col;
      float x = action_in[ao];
      e->prev_action[ao] = e->action[ao];
      e->action[ao] = x;
      e->raw_actions[ao] = x;
      float p = x * e->action_scale[col] + e->action_offset[col];
      if (e->action_clip) p = fminf(fmaxf(p, e->action_clip[2 * col]), e->action_clip[2 * col + 1]);
      e->processed_actions[ao] = p;
      if (e->action_kind == 0) e->joint_pos_target[o] = p; else e->joint_effort_target[o] = p;
    }
    const int gi = e->joint_group[j];
    float q_t = e->joint_pos_target[o], qd_t = e->joint_vel_target[o], ff = e->joint_effort_target[o];
    float prev_qd = e->prev_joint_vel[o], acc = e->joint_acc[o];
    float computed = 0.f, applied = 0.f;
    float q = sv.prev.joint_pos[o], qd = sv.prev.joint_vel[o];
    bool done = false;
    if constexpr (Net::kLstm) {
      if (gi >= 0 && e->act[gi].kind == B200LAB_ACT_NET_LSTM) {
        // recurrent state stays in registers across the S evaluations (actuator_net.py:76-99)
        const B200LabActuatorGroup& g = e->act[gi];
        float h[Net::kL][Net::kH], c[Net::kL][Net::kH];
        const int64_t rows = (int64_t)N * g.group_joints, row = (int64_t)n * g.group_joints + e->joint_local[j];
        lstm_state_load<Net>(g, row, rows, h, c);
        const float sat = g.saturation_effort, vlim = g.velocity_limit[j], elim = g.effort_limit[j];
#pragma unroll 1
        for (int s = 0; s < S; ++s) {
          // next state requested before the cell math: its latency hides behind ~1k FFMAs
          const float qn = sv.v[s].joint_pos[o], qdn = sv.v[s].joint_vel[o];
          computed = lstm_cells_const(net, q_t - q, qd, h, c);
          applied = dc_motor_clip(computed, qd, sat, vlim, elim);
          acc = (qdn - prev_qd) / dt;
          prev_qd = qdn;
          q = qn;
          qd = qdn;
        }
        lstm_state_store<Net>(g, row, rows, h, c);
        done = true;
      }
    }
    if constexpr (Net::kMlp) {
      if (gi >= 0 && e->act[gi].kind == B200LAB_ACT_NET_MLP) {
        // the position-error / velocity histories stay in registers across the S evaluations (actuator_net.py:147-189)
        const B200LabActuatorGroup& g = e->act[gi];
        float ph[kMlpMaxHist], vh[kMlpMaxHist];
        const int jl = e->joint_local[j];
        mlp_hist_load(g, n, jl, ph, vh);
        const float sat = g.saturation_effort, vlim = g.velocity_limit[j], elim = g.effort_limit[j];
#pragma unroll 1
        for (int s = 0; s < S; ++s) {
          const float qn = sv.v[s].joint_pos[o], qdn = sv.v[s].joint_vel[o];
          computed = mlp_actuator_step<Net>(s_w, s_w + ((Net::kCount + 3) & ~3) + threadIdx.x, (int)blockDim.x, g, q_t - q, qd, ph, vh);
          applied = dc_motor_clip(computed, qd, sat, vlim, elim);
          acc = (qdn - prev_qd) / dt;
          prev_qd = qdn;
          q = qn;
          qd = qdn;
        }
        mlp_hist_store(g, n, jl, ph, vh);
        done = true;
      }
    }
    if (!done) {
#pragma unroll 1
      for (int s = 0; s < S; ++s) {
        const float qn = sv.v[s].joint_pos[o], qdn = sv.v[s].joint_vel[o];
        if (gi >= 0) actuator_eval(e, s_w, n, j, q, qd, q_t, qd_t, ff, computed, applied);
        acc = (qdn - prev_qd) / dt;  // ArticulationData.update (articulation_data.py:547-556)
        prev_qd = qdn;
        q = qn;
        qd = qdn;
      }
    }
    e->prev_joint_vel[o] = prev_qd;
    e->joint_acc[o] = acc;
    if (gi >= 0 && S > 0) {
      e->computed_torque[o] = computed;
      e->applied_torque[o] = applied;
      if (e->act[gi].kind == B200LAB_ACT_IMPLICIT) { e->joint_pos_target_sim[o] = q_t; e->joint_effort_target_sim[o] = ff; }
      else e->joint_effort_target_sim[o] = applied;
    }
  }
  // ---- body role: ContactSensor (contact_sensor.py:320-379) under SensorBase.update (sensor_base.py:197-205)
  if (joint_role) return;

  if (e->cs_timestamp && i < B) {
    const int64_t ob = (int64_t)n * B + i;
    const int Hh = e->hist;
    float ts = e->cs_timestamp[n], tl = e->cs_timestamp_last_update[n];
    bool outdated = e->cs_is_outdated[n] != 0;
    const bool track = e->cs_current_air_time != nullptr;
    float la = 0.f, ca = 0.f, lc = 0.f, cc = 0.f;
    if (track) { la = e->cs_last_air_time[ob]; ca = e->cs_current_air_time[ob]; lc = e->cs_last_contact_time[ob]; cc = e->cs_current_contact_time[ob]; }
    float* hp = e->cs_forces_hist + ((int64_t)n * Hh * B + i) * 3;
    float hx[4], hy[4], hz[4];  // history depth <= 4 kept in registers
    const bool hreg = Hh <= 4;
    if (hreg) {
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (k < Hh) { hx[k] = hp[(int64_t)k * B * 3]; hy[k] = hp[(int64_t)k * B * 3 + 1]; hz[k] = hp[(int64_t)k * B * 3 + 2]; }
    }
    float fx = e->cs_net_forces_w[ob * 3], fy = e->cs_net_forces_w[ob * 3 + 1], fz = e->cs_net_forces_w[ob * 3 + 2];
    // the S force samples are requested together (independent loads) instead of one dependent round trip per sub-step
    float pfx[kMaxDecimation], pfy[kMaxDecimation], pfz[kMaxDecimation];
#pragma unroll
    for (int s = 0; s < kMaxDecimation; ++s) {
      pfx[s] = pfy[s] = pfz[s] = 0.f;
      if (s < S) {
        const float* nf = sv.v[s].net_forces_w;
        if (nf) { pfx[s] = nf[ob * 3]; pfy[s] = nf[ob * 3 + 1]; pfz[s] = nf[ob * 3 + 2]; }
      }
    }
#pragma unroll
    for (int s = 0; s < kMaxDecimation; ++s) {
      if (s >= S) break;
      ts = ts + dt;
      outdated = outdated || ((ts - tl) + 1e-6f >= e->cs_update_period);
      if (outdated) {
        fx = pfx[s]; fy = pfy[s]; fz = pfz[s];
        if (hreg) {
#pragma unroll
          for (int k = 3; k > 0; --k) if (k < Hh) { hx[k] = hx[k - 1]; hy[k] = hy[k - 1]; hz[k] = hz[k - 1]; }
          hx[0] = fx; hy[0] = fy; hz[0] = fz;
        } else {
          for (int k = Hh - 1; k > 0; --k) {
            float* dst = hp + (int64_t)k * B * 3; const float* src = hp + (int64_t)(k - 1) * B * 3;
            dst[0] = src[0]; dst[1] = src[1]; dst[2] = src[2];
          }
          hp[0] = fx; hp[1] = fy; hp[2] = fz;
        }
        if (track) {
          float elapsed = ts - tl;
          bool is_contact = norm3(fx, fy, fz) > e->cs_force_threshold;
          bool first_contact = (ca > 0.f) && is_contact;
          bool first_detached = (cc > 0.f) && !is_contact;
          if (first_contact) la = ca + elapsed;
          ca = !is_contact ? ca + elapsed : 0.f;
          if (first_detached) lc = cc + elapsed;
          cc = is_contact ? cc + elapsed : 0.f;
        }
        tl = ts;
        outdated = false;
      }
    }
    e->cs_net_forces_w[ob * 3] = fx; e->cs_net_forces_w[ob * 3 + 1] = fy; e->cs_net_forces_w[ob * 3 + 2] = fz;
    if (hreg) {
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (k < Hh) { hp[(int64_t)k * B * 3] = hx[k]; hp[(int64_t)k * B * 3 + 1] = hy[k]; hp[(int64_t)k * B * 3 + 2] = hz[k]; }
    }
    if (track) { e->cs_last_air_time[ob] = la; e->cs_current_air_time[ob] = ca; e->cs_last_contact_time[ob] = lc; e->cs_current_contact_time[ob] = cc; }
  }
  // ---- env scalars (one lane).  The sensor timestamps are re-derived with the same fp32 recurrence; they are
  // written after every lane of this env has read them (the block barrier below).
  float cts = 0.f, ctl = 0.f; bool cout_ = false; float rts = 0.f; bool rc_mark = false;
  if (i == 0) {
    if (e->cs_timestamp) {
      cts = e->cs_timestamp[n]; ctl = e->cs_timestamp_last_update[n]; cout_ = e->cs_is_outdated[n] != 0;
      for (int s = 0; s < S; ++s) {
        cts = cts + dt;
        cout_ = cout_ || ((cts - ctl) + 1e-6f >= e->cs_update_period);
        if (cout_) { ctl = cts; cout_ = false; }
      }
    }
    if (e->rc_timestamp) {
      rts = e->rc_timestamp[n];
      float rtl = e->rc_timestamp_last_update[n];
      for (int s = 0; s < S; ++s) { rts = rts + dt; if ((rts - rtl) + 1e-6f >= e->rc_update_period) rc_mark = true; }
    }
  }
  __syncthreads();
  if (i == 0) {
    if (e->cs_timestamp) { e->cs_timestamp[n] = cts; e->cs_timestamp_last_update[n] = ctl; e->cs_is_outdated[n] = cout_ ? 1 : 0; }
    if (e->rc_timestamp) { e->rc_timestamp[n] = rts; if (rc_mark) e->rc_is_outdated[n] = 1; }
  }
}

// ------------------------------------------------------------------------------------------------
// post_step: one warp per env.
// ------------------------------------------------------------------------------------------------
// Per-env working set of one warp.  Everything the terms read is loaded ONCE, coalesced, at kernel entry -- lane l holds
// joint l (and l+32), action column l (and l+32), body l (and l+32) -- so every term is register arithmetic plus warp
// reductions; the reference re-reads (and re-derives) these tensors in every term function.
struct EnvCtx {
  int n, lane;
  f4 q;
  f3 pos, lin_w, ang_w, lin_b, ang_b, grav_b;
  f3 cmd;
  bool terminated;
  // lane-distributed rows (slot 0: index lane, slot 1: index lane + 32)
  float jq[2], jqd[2], jacc[2], jtau[2], jdef[2], jlo[2], jhi[2];
  float act[2], pact[2];
  float bnorm[2];            // max over the contact history of |F| for the body (net_forces_w_history.norm(-1).max(1))
  float b_cc[2], b_ca[2], b_la[2];  // current contact time, current air time, last air time
};

__device__ __forceinline__ float u_reset_at(const Env* e, int n, int slot, int64_t step) {
  return e->u_reset ? e->u_reset[(int64_t)n * e->n_u_reset + slot] : philox_uniform(e->seed, step, 1u, (uint32_t)n, (uint32_t)slot);
}
__device__ __forceinline__ float u_step_at(const Env* e, int n, int slot, int64_t step) {
  return e->u_step ? e->u_step[(int64_t)n * e->n_u_step + slot] : philox_uniform(e->seed, step, 2u, (uint32_t)n, (uint32_t)slot);
}

// The uniforms one env consumes from a stream in this step, drawn ONCE, lane-parallel (slot = lane, lane + 32, ...), into a
// per-warp shared-memory table.  post_step used to inline philox_uniform at every use (~30 copies of the 10-round generator,
// 2850 instructions = 45 KB that only the resetting warps -- the kernel's critical path -- ever fetched).  Same counters,
// same values.
constexpr int kDrawCapReset = 384, kDrawCapStep = 64;
__device__ __noinline__ void fill_draws(float* dst, const float* __restrict__ recorded, int width, uint64_t seed, int64_t step,
                                        uint32_t stream_id, int n, int lane) {
  for (int s = lane; s < width; s += 32)
    dst[s] = recorded ? recorded[(int64_t)n * width + s] : philox_uniform(seed, step, stream_id, (uint32_t)n, (uint32_t)s);
  __syncwarp();
}

__device__ __forceinline__ void derive_state(EnvCtx& c, const float* rs) {
  c.pos = {rs[0], rs[1], rs[2]};
  c.q = {rs[3], rs[4], rs[5], rs[6]};
  c.lin_w = {rs[7], rs[8], rs[9]};
  c.ang_w = {rs[10], rs[11], rs[12]};
  c.lin_b = quat_rotate_inverse(c.q, c.lin_w);   // articulation_data.py:604-610
  c.ang_b = quat_rotate_inverse(c.q, c.ang_w);   // :613-619
  c.grav_b = quat_rotate_inverse(c.q, f3{0.f, 0.f, -1.f});  // :513-515
}

__device__ __forceinline__ void load_rows(EnvCtx& c, const Env* __restrict__ e, const Sim& sim) {
  const int n = c.n, lane = c.lane, J = e->J, A = e->A, B = e->B;
#pragma unroll
  for (int sl = 0; sl < 2; ++sl) {
    int j = lane + 32 * sl;
    c.jq[sl] = c.jqd[sl] = c.jacc[sl] = c.jtau[sl] = c.jdef[sl] = c.jlo[sl] = c.jhi[sl] = 0.f;
    c.act[sl] = c.pact[sl] = 0.f;
    c.bnorm[sl] = -INFINITY; c.b_cc[sl] = c.b_ca[sl] = c.b_la[sl] = 0.f;
    if (j < J) {
      int64_t o = (int64_t)n * J + j;
      c.jq[sl] = sim.joint_pos[o]; c.jqd[sl] = sim.joint_vel[o];
      c.jacc[sl] = e->joint_acc[o]; c.jtau[sl] = e->applied_torque[o];
      c.jdef[sl] = __ldg(e->default_joint_pos + j);
      c.jlo[sl] = __ldg(e->soft_joint_pos_limits + 2 * j); c.jhi[sl] = __ldg(e->soft_joint_pos_limits + 2 * j + 1);
    }
    if (j < A) { c.act[sl] = e->action[(int64_t)n * A + j]; c.pact[sl] = e->prev_action[(int64_t)n * A + j]; }
    if (j < B && e->cs_forces_hist) {
      const float* hp = e->cs_forces_hist + ((int64_t)n * e->hist * B + j) * 3;
      float m = -INFINITY;
      for (int k = 0; k < e->hist; ++k) {
        const float* f = hp + (int64_t)k * B * 3;
        m = fmaxf(m, norm3(f[0], f[1], f[2]));
      }
      c.bnorm[sl] = m;
      if (e->cs_current_air_time) {
        c.b_cc[sl] = e->cs_current_contact_time[(int64_t)n * B + j];
        c.b_ca[sl] = e->cs_current_air_time[(int64_t)n * B + j];
        c.b_la[sl] = e->cs_last_air_time[(int64_t)n * B + j];
      }
    }
  }
}

__device__ __forceinline__ bool sel(uint64_t mask, int lane, int sl) { return (mask >> (lane + 32 * sl)) & 1ull; }

// value held by the lane owning index `idx` (idx uniform across the warp)
__device__ __forceinline__ float row_at(const float (&row)[2], int idx) {
  float a = __shfl_sync(0xffffffffu, row[0], idx & 31);
  float b = __shfl_sync(0xffffffffu, row[1], idx & 31);
  return idx < 32 ? a : b;
}

__device__ float eval_termination(const Env* e, const Sim& sim, const B200LabTerm& t, const EnvCtx& c, int64_t ep_len) {
  const int lane = c.lane;
  switch (t.op) {
    case B200LAB_TERM_TIME_OUT:
      return ep_len >= e->max_episode_length ? 1.f : 0.f;
    case B200LAB_TERM_ILLEGAL_CONTACT: {
      int any = (sel(t.body_mask, lane, 0) && c.bnorm[0] > t.p[0]) || (sel(t.body_mask, lane, 1) && c.bnorm[1] > t.p[0]);
      return __any_sync(0xffffffffu, any) ? 1.f : 0.f;
    }
    case B200LAB_TERM_JOINT_POS_OUT_OF_MANUAL_LIMIT: {
      int any = 0;
#pragma unroll
      for (int sl = 0; sl < 2; ++sl)
        if (sel(t.joint_mask, lane, sl) && lane + 32 * sl < e->J) any |= (c.jq[sl] > t.p[1]) || (c.jq[sl] < t.p[0]);
      return __any_sync(0xffffffffu, any) ? 1.f : 0.f;
    }
    case B200LAB_TERM_ROOT_HEIGHT_BELOW_MINIMUM:
      return c.pos.z < t.p[0] ? 1.f : 0.f;
    case B200LAB_TERM_BAD_ORIENTATION:
      return fabsf(acosf(-c.grav_b.z)) > t.p[0] ? 1.f : 0.f;
    case B200LAB_TERM_EXTERNAL:
      return e->ext_values[(int64_t)t.i0 * e->N + c.n] != 0.f ? 1.f : 0.f;
  }
  return 0.f;
}

__device__ float eval_reward(const Env* e, const Sim& sim, const B200LabTerm& t, const EnvCtx& c) {
  const int n = c.n, lane = c.lane, J = e->J, B = e->B;
  switch (t.op) {
    case B200LAB_REW_IS_ALIVE: return c.terminated ? 0.f : 1.f;
    case B200LAB_REW_IS_TERMINATED: return c.terminated ? 1.f : 0.f;
    case B200LAB_REW_LIN_VEL_Z_L2: return c.lin_b.z * c.lin_b.z;
    case B200LAB_REW_ANG_VEL_XY_L2: return c.ang_b.x * c.ang_b.x + c.ang_b.y * c.ang_b.y;
    case B200LAB_REW_FLAT_ORIENTATION_L2: return c.grav_b.x * c.grav_b.x + c.grav_b.y * c.grav_b.y;
    case B200LAB_REW_JOINT_TORQUES_L2:
    case B200LAB_REW_JOINT_VEL_L1:
    case B200LAB_REW_JOINT_VEL_L2:
    case B200LAB_REW_JOINT_ACC_L2:
    case B200LAB_REW_JOINT_DEVIATION_L1:
    case B200LAB_REW_JOINT_POS_LIMITS:
    case B200LAB_REW_JOINT_POS_TARGET_L2: {
      float acc = 0.f;
#pragma unroll
      for (int sl = 0; sl < 2; ++sl) {
        if (lane + 32 * sl >= J || !sel(t.joint_mask, lane, sl)) continue;
        float v;
        switch (t.op) {
          case B200LAB_REW_JOINT_TORQUES_L2: v = c.jtau[sl] * c.jtau[sl]; break;
          case B200LAB_REW_JOINT_VEL_L1: v = fabsf(c.jqd[sl]); break;
          case B200LAB_REW_JOINT_VEL_L2: v = c.jqd[sl] * c.jqd[sl]; break;
          case B200LAB_REW_JOINT_ACC_L2: v = c.jacc[sl] * c.jacc[sl]; break;
          case B200LAB_REW_JOINT_DEVIATION_L1: v = fabsf(c.jq[sl] - c.jdef[sl]); break;
          case B200LAB_REW_JOINT_POS_LIMITS: {
            float lo = c.jq[sl] - c.jlo[sl], hi = c.jq[sl] - c.jhi[sl];
            v = -fminf(lo, 0.f) + fmaxf(hi, 0.f);   // rewards.py:190-196
            break;
          }
          default: {  // JOINT_POS_TARGET_L2
            float d = wrap_to_pi(c.jq[sl]) - t.p[0];
            v = d * d;
          }
        }
        acc += v;
      }
      return warp_sum(acc);
    }
    case B200LAB_REW_ACTION_RATE_L2:
    case B200LAB_REW_ACTION_L2: {
      float acc = 0.f;
#pragma unroll
      for (int sl = 0; sl < 2; ++sl) {
        if (lane + 32 * sl >= e->A) continue;
        float x = c.act[sl];
        if (t.op == B200LAB_REW_ACTION_RATE_L2) x = x - c.pact[sl];
        acc += x * x;
      }
      return warp_sum(acc);
    }
    case B200LAB_REW_UNDESIRED_CONTACTS:
    case B200LAB_REW_CONTACT_FORCES: {
      float acc = 0.f;
#pragma unroll
      for (int sl = 0; sl < 2; ++sl)
        if (lane + 32 * sl < B && sel(t.body_mask, lane, sl)) {
          float m = c.bnorm[sl];
          if (t.op == B200LAB_REW_UNDESIRED_CONTACTS) acc += (m > t.p[0]) ? 1.f : 0.f;
          else acc += fmaxf(m - t.p[0], 0.f);
        }
      return warp_sum(acc);
    }
    case B200LAB_REW_TRACK_LIN_VEL_XY_EXP: {
      float dx = c.cmd.x - c.lin_b.x, dy = c.cmd.y - c.lin_b.y;
      return expf(-(dx * dx + dy * dy) / t.p[0]);
    }
    case B200LAB_REW_TRACK_ANG_VEL_Z_EXP: {
      float d = c.cmd.z - c.ang_b.z;
      return expf(-(d * d) / t.p[0]);
    }
    case B200LAB_REW_TRACK_LIN_VEL_XY_YAW_FRAME_EXP: {
      f3 v = quat_rotate_inverse(yaw_quat(c.q), c.lin_w);
      float dx = c.cmd.x - v.x, dy = c.cmd.y - v.y;
      return expf(-(dx * dx + dy * dy) / t.p[0]);
    }
    case B200LAB_REW_TRACK_ANG_VEL_Z_WORLD_EXP: {
      float d = c.cmd.z - c.ang_w.z;
      return expf(-(d * d) / t.p[0]);
    }
    case B200LAB_REW_FEET_AIR_TIME: {
      float acc = 0.f;
#pragma unroll
      for (int sl = 0; sl < 2; ++sl)
        if (lane + 32 * sl < B && sel(t.body_mask, lane, sl)) {
          float cc = c.b_cc[sl];
          bool first = (cc > 0.f) && (cc < t.p[1]);   // contact_sensor.py:208-210
          acc += (c.b_la[sl] - t.p[0]) * (first ? 1.f : 0.f);
        }
      float r = warp_sum(acc);
      return r * ((sqrtf(c.cmd.x * c.cmd.x + c.cmd.y * c.cmd.y) > 0.1f) ? 1.f : 0.f);
    }
    case B200LAB_REW_FEET_AIR_TIME_POSITIVE_BIPED: {
      int cnt = 0;
      float mn = INFINITY;
      for (int b = 0; b < B; ++b)
        if ((t.body_mask >> b) & 1ull) cnt += row_at(c.b_cc, b) > 0.f;
      for (int b = 0; b < B; ++b)
        if ((t.body_mask >> b) & 1ull) {
          float ct = row_at(c.b_cc, b), at = row_at(c.b_ca, b);
          float mode = ct > 0.f ? ct : at;
          mn = fminf(mn, cnt == 1 ? mode : 0.f);
        }
      float r = fminf(mn, t.p[0]);
      return r * ((sqrtf(c.cmd.x * c.cmd.x + c.cmd.y * c.cmd.y) > 0.1f) ? 1.f : 0.f);
    }
    case B200LAB_REW_FEET_SLIDE: {
      float acc = 0.f;
      uint64_t m1 = t.body_mask, m2 = t.body_mask2;
      while (m1 && m2) {
        int b1 = __ffsll((long long)m1) - 1, b2 = __ffsll((long long)m2) - 1;
        m1 &= m1 - 1; m2 &= m2 - 1;
        bool contact = row_at(c.bnorm, b1) > 1.0f;
        const float* v = sim.body_lin_vel_w + ((int64_t)n * B + b2) * 3;
        acc += sqrtf(v[0] * v[0] + v[1] * v[1]) * (contact ? 1.f : 0.f);
      }
      return acc;
    }
    case B200LAB_REW_EXTERNAL:
      return e->ext_values[(int64_t)t.i0 * e->N + n];
  }
  return 0.f;
}

// UniformVelocityCommand._resample_command + CommandTerm._resample (velocity_command.py:122-138,
// command_manager.py:176-187).  u[0..6] = time_left, x, y, z, heading, is_heading, is_standing.
__device__ __forceinline__ void resample_command(const Env* e, int n, const float* u, f3& cmd) {
  e->cmd_time_left[n] = u[0] * e->cmd_resample_span + e->cmd_resample_lo;
  e->cmd_counter[n] += 1;
  const float* r = e->cmd_range;
  cmd.x = u[1] * e->cmd_span[0] + r[0];
  cmd.y = u[2] * e->cmd_span[1] + r[2];
  cmd.z = u[3] * e->cmd_span[2] + r[4];
  if (e->cmd_heading_command) {
    e->cmd_heading_target[n] = u[4] * e->cmd_span[3] + r[6];
    e->cmd_is_heading[n] = u[5] <= e->cmd_rel_heading;
  }
  e->cmd_is_standing[n] = u[6] <= e->cmd_rel_standing;
}

// Per-env derived quantities shared by all obs columns / rays of an env: [N, kDerived] floats in a scratch buffer owned
// by the env handle, filled by derive_env -- at the end of post_step (lane 0, from the root state it holds in registers)
// or by derive_obs_kernel when compute_obs is called on its own.
//   0-2 lin_b, 3-5 ang_b, 6-8 grav_b, 9-11 sensor pos (root pos + drift), 12-15 sensor-frame quaternion (yaw-only when
//   attach_yaw_only), 16 root z, 17 ray-caster outdated flag, 18 speculative scan valid
constexpr int kDerived = 20;
constexpr int kSpecPose = 8;   // sensor position (root + drift) xyz, sensor-frame quaternion wxyz, pad

// `rc_flag`, `drift`, `rc_ts`: the ray caster's is_outdated flag, drift and timestamp of this env (callers that hold them
// in registers avoid two dependent round trips at the end of their serial chain)
__device__ __forceinline__ void derive_env(const Env* __restrict__ e, const float (&rs)[13], int n, float* __restrict__ derived,
                                           const float* __restrict__ spec_pose, bool rc_flag, const float (&drift)[3],
                                           float rc_ts) {
  float* o = derived + (int64_t)n * kDerived;
  f4 q{rs[3], rs[4], rs[5], rs[6]};
  f3 lb = quat_rotate_inverse(q, f3{rs[7], rs[8], rs[9]});       // articulation_data.py:604-610
  f3 ab = quat_rotate_inverse(q, f3{rs[10], rs[11], rs[12]});    // :613-619
  f3 gb = quat_rotate_inverse(q, f3{0.f, 0.f, -1.f});            // :513-515
  o[0] = lb.x; o[1] = lb.y; o[2] = lb.z;
  o[3] = ab.x; o[4] = ab.y; o[5] = ab.z;
  o[6] = gb.x; o[7] = gb.y; o[8] = gb.z;
  o[16] = rs[2];
  float rc_out = 0.f, spec_ok = 0.f;
  if (e->rc_is_outdated) {
    rc_out = rc_flag ? 1.f : 0.f;
    if (rc_out != 0.f) {
      // RayCaster._update_buffers_impl (ray_caster.py:225-241): pose (+ drift) of the outdated sensors
      float px = rs[0] + drift[0], py = rs[1] + drift[1], pz = rs[2] + drift[2];
      o[9] = px; o[10] = py; o[11] = pz;
      f4 qy = e->rc_attach_yaw_only ? yaw_quat(q) : q;
      o[12] = qy.w; o[13] = qy.x; o[14] = qy.y; o[15] = qy.z;
      e->rc_pos_w[n * 3] = px; e->rc_pos_w[n * 3 + 1] = py; e->rc_pos_w[n * 3 + 2] = pz;
      e->rc_quat_w[n * 4] = q.w; e->rc_quat_w[n * 4 + 1] = q.x; e->rc_quat_w[n * 4 + 2] = q.y; e->rc_quat_w[n * 4 + 3] = q.z;
      e->rc_timestamp_last_update[n] = rc_ts;   // sensor_base.py:294-296
      e->rc_is_outdated[n] = 0;
      // the speculative scan is valid for this env iff it was cast from bitwise the same sensor pose
      if (spec_pose) {
        const float* sp = spec_pose + (int64_t)n * kSpecPose;
        const float pose[7] = {px, py, pz, qy.w, qy.x, qy.y, qy.z};
        bool same = true;
#pragma unroll
        for (int k = 0; k < 7; ++k) same = same && __float_as_uint(sp[k]) == __float_as_uint(pose[k]);
        spec_ok = same ? 1.f : 0.f;
      }
    } else {
      o[11] = e->rc_pos_w[n * 3 + 2];
    }
  }
  o[17] = rc_out;
  o[18] = spec_ok;
}

#ifndef B200LAB_STEP_WARPS
#define B200LAB_STEP_WARPS 4
#endif
constexpr int kStepWarps = B200LAB_STEP_WARPS;
#ifndef B200LAB_POST_BLOCKS
#define B200LAB_POST_BLOCKS 6
#endif

// kMode 0: the whole of post_step in one kernel.  The split form runs the same body as two kernels so that the code a
// step executes for (almost) every env stays small enough for the instruction cache:
//   kMode 1 "hot" : counters, terminations, rewards for every env; command countdown / interval events / derive for the
//                   envs that were NOT reset.  No reset code is compiled in.
//   kMode 2 "cold": one warp per env again, exits at once unless the hot kernel set reset_buf; reloads the env's context,
//                   runs _reset_idx and then the same tail.  The reference has the same structure: work on
//                   `reset_env_ids` only (manager_based_rl_env.py:215-220).
template <int kMode>
__global__ void __launch_bounds__(kStepWarps * 32, B200LAB_POST_BLOCKS * 4 / kStepWarps)
post_step_kernel(const __grid_constant__ Env env_param, Sim sim, int64_t step, const Env* __restrict__ env_dev,
                 float* __restrict__ derived) {
  const Env* __restrict__ e = &env_param;   // descriptor in the constant bank: warp-uniform reads, no global round trips
  __shared__ float s_draw_reset[kStepWarps][kDrawCapReset], s_draw_step[kStepWarps][kDrawCapStep];
  // The termination / reward term tables are staged from the device copy of the descriptor into shared memory: read
  // through the constant bank they push the kernel's constant footprint past ~4 KB, where every launch pays several
  // microseconds of serialised constant-cache misses (tools/imc_probe.cu).
  extern __shared__ uint2 s_tab[];   // the tables sit at 8-byte-aligned offsets of the descriptor
  if constexpr (kMode != 2) {
    const int n8_term = e->num_term * (int)(sizeof(B200LabTerm) / 8), n8_rew = e->num_rew * (int)(sizeof(B200LabTerm) / 8);
    const uint2* g_term = reinterpret_cast<const uint2*>(env_dev->term);
    const uint2* g_rew = reinterpret_cast<const uint2*>(env_dev->rew);
    for (int i = threadIdx.x; i < n8_term; i += blockDim.x) s_tab[i] = __ldg(g_term + i);
    for (int i = threadIdx.x; i < n8_rew; i += blockDim.x) s_tab[n8_term + i] = __ldg(g_rew + i);
  }
  __syncthreads();
  const B200LabTerm* __restrict__ s_term = reinterpret_cast<const B200LabTerm*>(s_tab);
  const B200LabTerm* __restrict__ s_rew = s_term + e->num_term;
  pdl_launch_dependents();   // compute_obs may be placed under this kernel's tail
  const int lane = threadIdx.x & 31;
  const int n = blockIdx.x * kStepWarps + (threadIdx.x >> 5);
  float* const ur = s_draw_reset[threadIdx.x >> 5];   // this env's reset-stream / step-stream uniforms (fill_draws)
  float* const us = s_draw_step[threadIdx.x >> 5];
  const int N = e->N, J = e->J, B = e->B, A = e->A;
  if (n >= N) { pdl_wait(); return; }
  if constexpr (kMode == 2) {
    if (!e->reset_buf[n]) return;   // the hot kernel finished this env
  }
  EnvCtx c;
  c.n = n; c.lane = lane;
  // ---- loads of LIBRARY-OWNED state that the preceding library kernel (decimation / sub-step) does not write: issued
  // before the dependency wait.  The simulator's tensors (root state, joints, forces) belong to the caller, whose own
  // kernel may be the predecessor on this stream: they are read after the wait.
  c.cmd = {0.f, 0.f, 0.f};
  if (e->has_command) c.cmd = {e->cmd_vel_b[n * 3], e->cmd_vel_b[n * 3 + 1], e->cmd_vel_b[n * 3 + 2]};
  // Every per-env scalar the tail of the step reads is requested HERE, together with the rows below: the stores that
  // follow (no-alias cannot be proven) would otherwise serialise each of these loads behind an L2 round trip.
  float pre_mxy = 0.f, pre_myaw = 0.f, pre_tl = 0.f, pre_ht = 0.f;
  bool pre_ih = false, pre_is = false;
  if (e->has_command) {
    pre_mxy = e->cmd_metric_xy[n]; pre_myaw = e->cmd_metric_yaw[n]; pre_tl = e->cmd_time_left[n];
    pre_is = e->cmd_is_standing[n] != 0;
    if (e->cmd_heading_command) { pre_ih = e->cmd_is_heading[n] != 0; pre_ht = e->cmd_heading_target[n]; }
  }
  float pre_itl[B200LAB_MAX_INTERVAL_EVENTS];
#pragma unroll
  for (int k = 0; k < B200LAB_MAX_INTERVAL_EVENTS; ++k) pre_itl[k] = (k < e->num_interval && e->ev_interval_time_left[k]) ? e->ev_interval_time_left[k][n] : 0.f;
  const float pre_esum = lane < e->num_rew ? e->episode_sums[(int64_t)lane * N + n] : 0.f;
  const int64_t pre_ep_len = e->episode_length_buf[n];
  bool pre_rc_flag = false;
  float pre_drift[3] = {0.f, 0.f, 0.f}, pre_rc_ts = 0.f;
  if (derived && e->rc_is_outdated) {
#pragma unroll
    for (int k = 0; k < 3; ++k) pre_drift[k] = e->rc_drift[n * 3 + k];
  }
  // ---- everything below reads what the preceding kernel wrote
  pdl_wait();
  float rs[13];
#pragma unroll
  for (int k = 0; k < 13; ++k) rs[k] = sim.root_state_w[(int64_t)n * 13 + k];
  derive_state(c, rs);
  if (e->step_counter_dev) step = *e->step_counter_dev;
  if constexpr (kMode != 2) load_rows(c, e, sim);
  if (derived && e->rc_is_outdated) {
    pre_rc_flag = e->rc_is_outdated[n] != 0;
    pre_rc_ts = e->rc_timestamp[n];
  }
  // -- episode counter (manager_based_rl_env.py:200)
  int64_t ep_len = pre_ep_len + 1;
  // -- terminations (termination_manager.py:151-174)
  bool terminated = false, truncated = false;
  if constexpr (kMode != 2)
  for (int t = 0; t < e->num_term; ++t) {
    bool v = eval_termination(e, sim, s_term[t], c, ep_len) != 0.f;
    if (s_term[t].time_out) truncated |= v; else terminated |= v;
    if (lane == 0) e->term_dones[(int64_t)t * N + n] = v;
  }
  c.terminated = terminated;
  const bool reset = kMode == 2 ? true : (terminated || truncated);
  // -- rewards (reward_manager.py:128-157)
  if constexpr (kMode != 2) {
    float total = 0.f;
    const float dt = e->step_dt;
    float mine = 0.f;        // lane r keeps term r's value: the per-term buffers are then updated by all lanes at once
    bool mine_on = false;
    for (int r = 0; r < e->num_rew; ++r) {
      const B200LabTerm& t = s_rew[r];
      if (t.weight == 0.0f) continue;
      float value = eval_reward(e, sim, t, c) * t.weight * dt;
      total += value;          // reward_buf += value, in term order (reward_manager.py:150)
      if (lane == r) { mine = value; mine_on = true; }
    }
    if (mine_on) {
      e->episode_sums[(int64_t)lane * N + n] = pre_esum + mine;
      e->step_reward[(int64_t)n * e->num_rew + lane] = mine / dt;
    }
    if (lane == 0) {
      e->reward_buf[n] = total;
      e->terminated[n] = terminated;
      e->truncated[n] = truncated;
      e->reset_buf[n] = reset;
    }
  }
  __syncwarp();
  if constexpr (kMode == 1) {
    if (reset) {   // the cold kernel takes this env from here
      if (lane == 0 && e->any_reset_dev) *e->any_reset_dev = 1;
      return;
    }
  }
  // -- _reset_idx (manager_based_rl_env.py:347-392), masked per env
  if constexpr (kMode != 1)
  if (reset) {
    if (lane == 0 && e->any_reset_dev) *e->any_reset_dev = 1;
    fill_draws(ur, e->u_reset, e->n_u_reset, e->seed, step, 1u, n, lane);   // every reset-stream uniform of this env
    // curriculum: terrain_levels_vel (curriculums.py:26-55) + update_env_origins (terrain_importer.py:307-322)
    if (e->has_terrain_curriculum) {
      float ox = e->env_origins[n * 3], oy = e->env_origins[n * 3 + 1];
      float dx = c.pos.x - ox, dy = c.pos.y - oy;
      float dist = sqrtf(dx * dx + dy * dy);
      bool up = dist > e->terrain_size_x / 2.0f;
      bool down = dist < sqrtf(c.cmd.x * c.cmd.x + c.cmd.y * c.cmd.y) * e->max_episode_length_s * 0.5f;
      down = down && !up;
      int64_t lvl = e->terrain_levels[n] + (up ? 1 : 0) - (down ? 1 : 0);
      if (lvl >= e->terrain_rows) {
        float u = ur[B200LAB_U_RESET_CURRICULUM];
        lvl = min((int64_t)(u * (float)e->terrain_rows), (int64_t)e->terrain_rows - 1);
      } else if (lvl < 0) lvl = 0;
      const float* to = e->terrain_origins + ((int64_t)lvl * e->terrain_cols + e->terrain_types[n]) * 3;
      __syncwarp();
      if (lane == 0) {
        e->terrain_levels[n] = lvl;
        e->env_origins[n * 3] = to[0]; e->env_origins[n * 3 + 1] = to[1]; e->env_origins[n * 3 + 2] = to[2];
      }
      __syncwarp();
    }
    // scene.reset: actuators (actuator_net.py:70-74,142-145)
    for (int g = 0; g < e->num_act_groups; ++g) {
      const B200LabActuatorGroup& ag = e->act[g];
      if (ag.kind == B200LAB_ACT_NET_LSTM) {
        int per = ag.group_joints * ag.lstm_hidden;
        int64_t rows = (int64_t)N * per;
        for (int l = 0; l < ag.lstm_layers; ++l)
          for (int k = lane; k < per; k += 32) {
            ag.lstm_h[l * rows + (int64_t)n * per + k] = 0.f;
            ag.lstm_c[l * rows + (int64_t)n * per + k] = 0.f;
          }
      } else if (ag.kind == B200LAB_ACT_NET_MLP) {
        int per = ag.mlp_hist * ag.group_joints;
        for (int k = lane; k < per; k += 32) {
          ag.mlp_pos_hist[(int64_t)n * per + k] = 0.f;
          ag.mlp_vel_hist[(int64_t)n * per + k] = 0.f;
        }
      } else if (ag.kind == B200LAB_ACT_DELAYED_PD || ag.kind == B200LAB_ACT_REMOTIZED_PD) {   // DelayedPDActuator.reset (actuator_pd.py:311-334)
        for (int k = lane; k < ag.group_joints; k += 32) ag.delay_count[(int64_t)n * ag.group_joints + k] = 0;
        if (lane == 0) {
          const int span = ag.max_delay - ag.min_delay + 1;
          ag.delay_lag[n] = ag.min_delay + min((int)(ur[ag.delay_u_slot] * (float)span), span - 1);
        }
      }
    }
    // scene.reset: contact sensor (contact_sensor.py:143-162, sensor_base.py:182-195)
    if (e->cs_timestamp) {
      for (int k = lane; k < B * 3; k += 32) e->cs_net_forces_w[(int64_t)n * B * 3 + k] = 0.f;
      for (int k = lane; k < e->hist * B * 3; k += 32) e->cs_forces_hist[(int64_t)n * e->hist * B * 3 + k] = 0.f;
      if (e->cs_current_air_time)
        for (int k = lane; k < B; k += 32) {
          e->cs_current_air_time[(int64_t)n * B + k] = 0.f;
          e->cs_last_air_time[(int64_t)n * B + k] = 0.f;
          e->cs_current_contact_time[(int64_t)n * B + k] = 0.f;
          e->cs_last_contact_time[(int64_t)n * B + k] = 0.f;
        }
      if (lane == 0) { e->cs_timestamp[n] = 0.f; e->cs_timestamp_last_update[n] = 0.f; e->cs_is_outdated[n] = 1; }
    }
    // scene.reset: ray caster (ray_caster.py:107-114)
    if (e->rc_timestamp && lane == 0) {
      e->rc_timestamp[n] = 0.f; e->rc_timestamp_last_update[n] = 0.f; e->rc_is_outdated[n] = 1;
      pre_rc_ts = 0.f; pre_rc_flag = true;
      for (int k = 0; k < 3; ++k) {
        pre_drift[k] = ur[B200LAB_U_RESET_DRIFT + k] * e->rc_drift_span + e->rc_drift_lo;
        e->rc_drift[n * 3 + k] = pre_drift[k];
      }
    }
    // EventManager.apply(mode="reset") trigger gating (event_manager.py:232-270)
    const int64_t gstep = step + 1;   // env._sim_step_counter // decimation at the time of _reset_idx
    auto gate = [&](int32_t min_steps, int32_t* last_step, uint8_t* once) -> bool {
      if (min_steps <= 0 || !last_step || !once) return true;
      const int32_t last = last_step[n];
      const bool valid = (gstep - (int64_t)last >= (int64_t)min_steps) || (last == 0 && once[n] == 0);
      __syncwarp();
      if (valid && lane == 0) { last_step[n] = (int32_t)gstep; once[n] = 1; }
      __syncwarp();
      return valid;
    };
    // reset events: reset_root_state_uniform (events.py:823-865)
    if (e->ev_reset_root_state && gate(e->ev_root_min_steps, e->ev_root_last_step, e->ev_root_once)) {
      float up[6], uv[6];
#pragma unroll
      for (int k = 0; k < 6; ++k) {
        float a = ur[B200LAB_U_RESET_POSE + k], b = ur[B200LAB_U_RESET_VEL + k];
        up[k] = a * e->ev_pose_range[2 * k + 1] + e->ev_pose_range[2 * k];
        uv[k] = b * e->ev_vel_range[2 * k + 1] + e->ev_vel_range[2 * k];
      }
      const float* d = e->default_root_state;
      const float* org = e->env_origins;
      rs[0] = (d[0] + (org ? org[n * 3] : 0.f)) + up[0];
      rs[1] = (d[1] + (org ? org[n * 3 + 1] : 0.f)) + up[1];
      rs[2] = (d[2] + (org ? org[n * 3 + 2] : 0.f)) + up[2];
      f4 dq = quat_from_euler_xyz(up[3], up[4], up[5]);
      f4 qn = quat_mul(f4{d[3], d[4], d[5], d[6]}, dq);
      rs[3] = qn.w; rs[4] = qn.x; rs[5] = qn.y; rs[6] = qn.z;
#pragma unroll
      for (int k = 0; k < 6; ++k) rs[7 + k] = d[7 + k] + uv[k];
      if (lane < 13) sim.root_state_w[(int64_t)n * 13 + lane] = rs[lane];
      derive_state(c, rs);
    }
    // reset events: reset_joints_by_scale / by_offset (events.py:987-1051)
    for (int k = 0; k < e->ev_num_reset_joints; ++k) {
      if (!gate(e->ev_reset_joints_min_steps[k], e->ev_reset_joints_last_step[k], e->ev_reset_joints_once[k])) continue;
      const int base = B200LAB_U_RESET_JOINT + 2 * J * k;
      for (int j = lane; j < J; j += 32) {
        float u0 = ur[base + j], u1 = ur[base + J + j];
        float sp = u0 * e->ev_joint_pos_range[k][1] + e->ev_joint_pos_range[k][0];
        float sv = u1 * e->ev_joint_vel_range[k][1] + e->ev_joint_vel_range[k][0];
        float qp = e->default_joint_pos[j], qv = e->default_joint_vel[j];
        if (e->ev_reset_joints_mode[k] == 1) { qp = qp * sp; qv = qv * sv; } else { qp = qp + sp; qv = qv + sv; }
        qp = fminf(fmaxf(qp, e->soft_joint_pos_limits[2 * j]), e->soft_joint_pos_limits[2 * j + 1]);
        float vl = e->soft_joint_vel_limits[j];
        qv = fminf(fmaxf(qv, -vl), vl);
        int64_t o = (int64_t)n * J + j;
        sim.joint_pos[o] = qp;                 // articulation.py:514
        sim.joint_vel[o] = qv;                 // :551-553
        e->prev_joint_vel[o] = qv;
        e->joint_acc[o] = 0.f;
      }
    }
    // manager resets, in the order of manager_based_rl_env.py:364-389
    // action manager (action_manager.py:296-316)
    for (int a = lane; a < A; a += 32) {
      e->prev_action[(int64_t)n * A + a] = 0.f;
      e->action[(int64_t)n * A + a] = 0.f;
      e->raw_actions[(int64_t)n * A + a] = 0.f;
    }
    // reward manager: episodic sums -> log, zero (reward_manager.py:100-126)
    if (lane == 0 && e->log_accum) atomicAdd(e->log_accum, 1.0);
    for (int r = lane; r < e->num_rew; r += 32) {
      if (e->log_accum) atomicAdd(e->log_accum + 1 + r, (double)e->episode_sums[(int64_t)r * N + n]);
      e->episode_sums[(int64_t)r * N + n] = 0.f;
    }
    // termination manager log (termination_manager.py:127-149)
    if (e->log_accum)
      for (int t = lane; t < e->num_term; t += 32)
        if (e->term_dones[(int64_t)t * N + n]) atomicAdd(e->log_accum + 1 + e->num_rew + t, 1.0);
    // command manager reset (command_manager.py:120-149)
    if (e->has_command) {
      if (lane == 0) {
        if (e->log_accum) {
          atomicAdd(e->log_accum + 1 + e->num_rew + e->num_term, (double)e->cmd_metric_xy[n]);
          atomicAdd(e->log_accum + 2 + e->num_rew + e->num_term, (double)e->cmd_metric_yaw[n]);
        }
        e->cmd_metric_xy[n] = 0.f;
        e->cmd_metric_yaw[n] = 0.f;
        e->cmd_counter[n] = 0;
      }
      float u[7];
#pragma unroll
      for (int k = 0; k < 7; ++k) u[k] = ur[B200LAB_U_RESET_CMD + k];
      __syncwarp();
      if (lane == 0) resample_command(e, n, u, c.cmd);
      c.cmd.x = __shfl_sync(0xffffffffu, c.cmd.x, 0);
      c.cmd.y = __shfl_sync(0xffffffffu, c.cmd.y, 0);
      c.cmd.z = __shfl_sync(0xffffffffu, c.cmd.z, 0);
    }
    ep_len = 0;  // manager_based_rl_env.py:392
    if (e->has_command) {   // the command reset above rewrote these (same lane: program order)
      pre_mxy = 0.f; pre_myaw = 0.f; pre_tl = e->cmd_time_left[n];
      pre_is = e->cmd_is_standing[n] != 0;
      if (e->cmd_heading_command) { pre_ih = e->cmd_is_heading[n] != 0; pre_ht = e->cmd_heading_target[n]; }
    }
  }
  if (lane == 0) e->episode_length_buf[n] = ep_len;
  __syncwarp();
  // step-stream uniforms (command resampling, interval pushes): drawn only by the envs that consume some this step; the
  // predicates repeat, warp-uniformly, the tests lane 0 makes below
  {
    bool need_us = e->has_command && (__shfl_sync(0xffffffffu, pre_tl, 0) - e->step_dt) <= 0.0f;
#pragma unroll
    for (int k = 0; k < B200LAB_MAX_INTERVAL_EVENTS; ++k) {
      if (k >= e->num_interval) break;
      const bool global_time = e->ev_interval_global[k] != 0 && e->ev_interval_global_time[k] != nullptr;
      const float tl = (global_time ? e->ev_interval_global_time[k][step & 1] : pre_itl[k]) - e->step_dt;
      need_us = need_us || tl < 1e-6f;
    }
    if (need_us) fill_draws(us, e->u_step, e->n_u_step, e->seed, step, 2u, n, lane);
  }
  // -- command_manager.compute (command_manager.py:151-166)
  if (e->has_command) {
    if (lane == 0) {
      // _update_metrics (velocity_command.py:111-120)
      float ex = c.cmd.x - c.lin_b.x, ey = c.cmd.y - c.lin_b.y;
      e->cmd_metric_xy[n] = pre_mxy + sqrtf(ex * ex + ey * ey) / e->cmd_max_command_step;
      e->cmd_metric_yaw[n] = pre_myaw + fabsf(c.cmd.z - c.ang_b.z) / e->cmd_max_command_step;
      float tl = pre_tl - e->step_dt;
      e->cmd_time_left[n] = tl;
      if (tl <= 0.0f) {
        float u[7];
#pragma unroll
        for (int k = 0; k < 7; ++k) u[k] = us[B200LAB_U_STEP_CMD + k];
        resample_command(e, n, u, c.cmd);
        pre_is = e->cmd_is_standing[n] != 0;
        if (e->cmd_heading_command) { pre_ih = e->cmd_is_heading[n] != 0; pre_ht = e->cmd_heading_target[n]; }
      }
      // _update_command (velocity_command.py:140-160)
      if (e->cmd_heading_command && pre_ih) {
        f3 fw = quat_apply(c.q, f3{1.f, 0.f, 0.f});            // heading_w, articulation_data.py:518-526
        float heading = atan2f(fw.y, fw.x);
        float err = wrap_to_pi(pre_ht - heading);
        c.cmd.z = fminf(fmaxf(e->cmd_heading_stiffness * err, e->cmd_range[4]), e->cmd_range[5]);
      }
      if (pre_is) c.cmd = {0.f, 0.f, 0.f};
      e->cmd_vel_b[n * 3] = c.cmd.x; e->cmd_vel_b[n * 3 + 1] = c.cmd.y; e->cmd_vel_b[n * 3 + 2] = c.cmd.z;
    }
  }
  // -- interval events: push_by_setting_velocity (event_manager.py:204-229, events.py:795-820)
  if (lane == 0) {
    if (e->obs_hist_reset_step) {   // observation history: reset envs refill their rings in this step's compute; one append per step
      if (reset) e->obs_hist_reset_step[n] = step;
      if (n == 0) *e->obs_hist_count = (*e->obs_hist_count + 1) & 0x3fffffff;
    }
#pragma unroll
    for (int k = 0; k < B200LAB_MAX_INTERVAL_EVENTS; ++k) {
      if (k >= e->num_interval) break;
      const bool global_time = e->ev_interval_global[k] != 0 && e->ev_interval_global_time[k] != nullptr;
      float tl = (global_time ? e->ev_interval_global_time[k][step & 1] : pre_itl[k]) - e->step_dt;
      if (tl < 1e-6f) {
        int base = B200LAB_U_STEP_INTERVAL + 7 * k;
        // is_global_time: one draw for the new interval (env 0's slot), every env is pushed (event_manager.py:213-220)
        tl = (global_time ? u_step_at(e, 0, base, step) : us[base]) * e->ev_interval_range[k][1] + e->ev_interval_range[k][0];
        for (int a = 0; a < 6; ++a) {
          float lo = e->ev_push_range[k][2 * a], span = e->ev_push_range[k][2 * a + 1];
          rs[7 + a] = rs[7 + a] + (us[base + 1 + a] * span + lo);
          sim.root_state_w[(int64_t)n * 13 + 7 + a] = rs[7 + a];
        }
      }
      if (global_time) { if (n == 0) e->ev_interval_global_time[k][(step + 1) & 1] = tl; }
      else e->ev_interval_time_left[k][n] = tl;
    }
    // the quantities compute_obs shares across the columns of this env, from the final root state of the step (lane 0
    // holds it: resets were applied by every lane, pushes by lane 0) -- saves the derive launch on the critical path
    if (derived) derive_env(e, rs, n, derived, nullptr, pre_rc_flag, pre_drift, pre_rc_ts);
  }
}

// Per-column program of the concatenated group (48 B, three coalesced 128-bit loads per thread), built at env_create:
//   a = {op | src << 8 | noise_kind << 24 | clip_on << 30 | scale_on << 31, noise_lo, noise_span, scale}
//   b = {clip_lo, clip_hi, p[0], EXTERNAL column}
//   c = {ray local start xyz, ray direction z}   (height-scan columns)
//   d = {history: H | row << 8 | modifier kinds << 16 (3 x 2 bits), ring offset of (term, element) in the env's obs_hist
//        row, term width, column of the NEWEST row's element (where the element's noise draw is indexed)}   as bit patterns
//   m = modifier parameters
struct ObsCol { float4 a, b, c, d, m; };


__global__ void __launch_bounds__(128)
derive_obs_kernel(const __grid_constant__ Env env_param, Sim sim, float* __restrict__ derived,
                  const float* __restrict__ spec_pose, int env_begin, int env_end) {
  const Env* __restrict__ e = &env_param;
  int n = env_begin + blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= env_end) return;
  float rs[13];
#pragma unroll
  for (int k = 0; k < 13; ++k) rs[k] = sim.root_state_w[(int64_t)n * 13 + k];
  bool rc_flag = false;
  float drift[3] = {0.f, 0.f, 0.f}, rc_ts = 0.f;
  if (e->rc_is_outdated) {
    rc_flag = e->rc_is_outdated[n] != 0;
    rc_ts = e->rc_timestamp[n];
#pragma unroll
    for (int k = 0; k < 3; ++k) drift[k] = e->rc_drift[n * 3 + k];
  }
  derive_env(e, rs, n, derived, spec_pose, rc_flag, drift, rc_ts);
}

// ------------------------------------------------------------------------------------------------
// Speculative height scan (vertical rays, yaw-only sensor).  The sensor pose after the last physics sub-step is known
// before the actuator / contact / termination / reward kernels of the step have run, and only a reset (which rewrites
// the root pose and the drift) changes it afterwards.  One block per env: thread 0 derives the sensor pose exactly as
// derive_obs_kernel does and records it; every thread casts its rays and stores the hit height.  compute_obs consumes
// the heights of the envs whose recorded pose is bitwise the pose it derives itself and casts the others in place.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
height_scan_spec_kernel(const __grid_constant__ Env env_param, Sim sim, const ObsCol* __restrict__ ray_cols,
                        const __grid_constant__ MeshAccel mesh, float* __restrict__ spec_z, float* __restrict__ spec_pose) {
  const Env* __restrict__ e = &env_param;
  __shared__ float s_pose[kSpecPose];
  const int n = blockIdx.x;
  if (threadIdx.x == 0) {
    const float* rs = sim.root_state_w + (int64_t)n * 13;
    f4 q{rs[3], rs[4], rs[5], rs[6]};
    float px = rs[0] + e->rc_drift[n * 3], py = rs[1] + e->rc_drift[n * 3 + 1], pz = rs[2] + e->rc_drift[n * 3 + 2];
    f4 qy = yaw_quat(q);
    float* sp = spec_pose + (int64_t)n * kSpecPose;
    s_pose[0] = px; s_pose[1] = py; s_pose[2] = pz; s_pose[3] = qy.w; s_pose[4] = qy.x; s_pose[5] = qy.y; s_pose[6] = qy.z;
#pragma unroll
    for (int k = 0; k < 7; ++k) sp[k] = s_pose[k];
  }
  __syncthreads();
  const f4 qy{s_pose[3], s_pose[4], s_pose[5], s_pose[6]};
  for (int r = threadIdx.x; r < e->R; r += blockDim.x) {
    const float4 cc = __ldg(&ray_cols[r].c);
    f3 st = quat_apply(qy, f3{cc.x, cc.y, cc.z});
    st.x += s_pose[0]; st.y += s_pose[1]; st.z += s_pose[2];
    float tt; int face;
    float z = INFINITY;
    if (raycast_vertical(mesh, st, cc.w, e->rc_max_distance, tt, face)) z = st.z + tt * cc.w;
    spec_z[(int64_t)n * e->R + r] = z;
  }
}

// ------------------------------------------------------------------------------------------------
// compute_obs: ObservationManager.compute_group (observation_manager.py:260-335) for one concatenated group.
// One thread per (env, obs column), no barriers: the term of a column comes from the per-column table built at
// env_create, the env-level quantities from the derived scratch.  Height-scan columns cast their ray.
// ------------------------------------------------------------------------------------------------

__device__ __forceinline__ float obs_finish(const Env* __restrict__ e, float v, float4 ca, float4 cb, unsigned w0, int kind,
                                            int64_t step, unsigned n, unsigned k, unsigned gid);
__device__ __forceinline__ float obs_apply(float v, float4 ca, float4 cb, unsigned w0, int kind, float u1, float u2, bool recorded);

#ifndef B200LAB_OBS_BLOCKS
#define B200LAB_OBS_BLOCKS 4
#endif
#ifndef B200LAB_OBS_THREADS
#define B200LAB_OBS_THREADS 64
#endif
// kBatch: candidates of the ray's cell requested per round trip (raycast_vertical); kBatch = 4 trades the 64-register cap
// (and with it a quarter of the resident warps) for half the dependent round trips of a 4-candidate cell
// kPlain: no term keeps history or carries modifiers -- that code (and its column-program word) is compiled out, which is
// what the 64-register budget of the headline configs was tuned for
template <bool kVerticalRays, int kBatch = kVertBatch, bool kPlain = true>
__global__ void __launch_bounds__(B200LAB_OBS_THREADS, kVerticalRays ? (kBatch > 2 ? 3 : B200LAB_OBS_BLOCKS) * (256 / B200LAB_OBS_THREADS) : 2)
compute_obs_kernel(const __grid_constant__ Env env_param, Sim sim, int64_t step, const ObsCol* __restrict__ cols,
                   const float* __restrict__ derived, const __grid_constant__ MeshAccel mesh, int has_mesh,
                   const float* __restrict__ spec_z, unsigned env_begin, unsigned env_count) {
  const Env* __restrict__ e = &env_param;
  const unsigned N = (unsigned)e->N, D = (unsigned)e->obs_dim;
  const unsigned lid = blockIdx.x * blockDim.x + threadIdx.x;   // host guarantees N * D < 2^31
  if (lid >= env_count * D) return;
  const unsigned n = env_begin + lid / D, k = lid - (lid / D) * D;
  const unsigned gid = n * D + k;
  const float* o = derived + (size_t)n * kDerived;
  const ObsCol* col = cols + k;
  const float4 ca = __ldg(&col->a), cb = __ldg(&col->b);   // static program: read before the dependency wait
  const float4 cd = kPlain ? make_float4(0.f, 0.f, 0.f, 0.f) : __ldg(&col->d);
  pdl_wait();
  const unsigned w0 = __float_as_uint(ca.x);
  const int op = w0 & 0xffu, li = (w0 >> 8) & 0xffffu, kind = (w0 >> 24) & 0x3fu;
  // history (observation_manager.py:319-325): column = (row h of H, element); rows oldest first, the newest row is the
  // one computed now.  Slot of row h in the ring: (appends so far + 1 + h) mod H.
  const unsigned hw = kPlain ? 0u : __float_as_uint(cd.x), H = hw & 0xffu, hrow = (hw >> 8) & 0xffu, modk = hw >> 16;
  float* ring = nullptr;
  if (H) {
    if (e->step_counter_dev) step = *e->step_counter_dev;
    const unsigned cnt = (unsigned)*e->obs_hist_count;
    ring = e->obs_hist + (size_t)n * e->obs_hist_total + __float_as_uint(cd.y) + ((cnt + 1u + hrow) % H) * __float_as_uint(cd.z);
    if (hrow != H - 1 && e->obs_hist_reset_step[n] != step) {   // an older row: already noised / clipped / scaled when it was new
      e->obs_buf[gid] = *ring;
      return;
    }
  }
  float v = 0.f;
  switch (op) {
    case B200LAB_OBS_BASE_LIN_VEL: v = o[li]; break;
    case B200LAB_OBS_BASE_ANG_VEL: v = o[3 + li]; break;
    case B200LAB_OBS_PROJECTED_GRAVITY: v = o[6 + li]; break;
    case B200LAB_OBS_GENERATED_COMMANDS: v = e->cmd_vel_b[n * 3 + li]; break;
    case B200LAB_OBS_JOINT_POS_REL: v = sim.joint_pos[(size_t)n * e->J + li] - e->default_joint_pos[li]; break;
    case B200LAB_OBS_JOINT_VEL_REL: v = sim.joint_vel[(size_t)n * e->J + li] - e->default_joint_vel[li]; break;
    case B200LAB_OBS_JOINT_POS: v = sim.joint_pos[(size_t)n * e->J + li]; break;
    case B200LAB_OBS_JOINT_VEL: v = sim.joint_vel[(size_t)n * e->J + li]; break;
    case B200LAB_OBS_LAST_ACTION: v = e->action[(size_t)n * e->A + li]; break;
    case B200LAB_OBS_BASE_POS_Z: v = o[16]; break;
    case B200LAB_OBS_EXTERNAL: v = e->ext_values[((size_t)__float_as_int(cb.w) + li) * N + n]; break;
    case B200LAB_OBS_HEIGHT_SCAN: {
      float pz = o[11], hz;
      float* hp = e->rc_ray_hits_w ? e->rc_ray_hits_w + ((size_t)n * e->R + li) * 3 : nullptr;
      if (o[17] != 0.f) {
        // ray start = quat_apply(_yaw)(q, local start) + pos ; directions unrotated when attach_yaw_only (ray_caster.py:243-253)
        const float4 cc = __ldg(&col->c);
        f4 qy{o[12], o[13], o[14], o[15]};
        f3 st = quat_apply(qy, f3{cc.x, cc.y, cc.z});
        st.x += o[9]; st.y += o[10]; st.z += o[11];
        float tt; int face;
        f3 hit{INFINITY, INFINITY, INFINITY};
        // kVerticalRays: every pattern direction is (0, 0, dz) and the sensor is yaw-only (the height scanner) ->
        // single-cell query, no DDA code in the kernel (fewer registers, higher occupancy)
        if (kVerticalRays) {
          if (o[18] != 0.f) {   // speculative scan cast from this very pose (height_scan_spec_kernel)
            float z = spec_z[(size_t)n * e->R + li];
            if (z != INFINITY) hit = {st.x, st.y, z};
          } else if (has_mesh && raycast_vertical<kBatch>(mesh, st, cc.w, e->rc_max_distance, tt, face)) {
            hit = {st.x, st.y, st.z + tt * cc.w};
          }
        } else {
          f3 ld{__ldg(e->rc_ray_dirs + li * 3), __ldg(e->rc_ray_dirs + li * 3 + 1), __ldg(e->rc_ray_dirs + li * 3 + 2)};
          f3 dr = e->rc_attach_yaw_only ? ld : quat_apply(qy, ld);
          if (has_mesh && raycast_mesh_dev(mesh, st, dr, e->rc_max_distance, tt, face))
            hit = {st.x + tt * dr.x, st.y + tt * dr.y, st.z + tt * dr.z};
        }
        if (hp) { hp[0] = hit.x; hp[1] = hit.y; hp[2] = hit.z; }
        hz = hit.z;
      } else {
        hz = hp ? hp[2] : INFINITY;   // the host drops the hit buffer only for sensors that are outdated at every compute
      }
      v = (pz - hz) - cb.z;  // observations.py:173
      break;
    }
  }
  if (modk) {   // function modifiers, before the noise (observation_manager.py:310-312)
    const float4 cm = __ldg(&col->m);
    float p0 = cm.x, p1 = cm.y, p2 = cm.z, p3 = cm.w;   // parameters are consumed front to back
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const unsigned mk = (modk >> (2 * i)) & 3u;
      if (mk == 3u) { v = fminf(fmaxf(v, p0), p1); p0 = p2; p1 = p3; }
      else if (mk != 0u) { v = mk == 1u ? v * p0 : v + p0; p0 = p1; p1 = p2; p2 = p3; }
    }
  }
  // the element's noise draw is indexed by the column of its newest row (one draw per element and step)
  const unsigned kn = H ? __float_as_uint(cd.w) : k;
  v = obs_finish(e, v, ca, cb, w0, kind, step, n, kn, n * D + kn);
  if (H) *ring = v;   // the newest row's slot -- or, right after a reset, this row's slot: the ring is filled with the fresh row
  e->obs_buf[gid] = v;
}

// ------------------------------------------------------------------------------------------------
// compute_obs, Q columns per thread (vertical rays, no history / modifiers).
//
// The one-column kernel above is a chain of dependent L2 round trips per ray (cell range -> entries -> records) run by
// ~27 warps per SM; ncu shows it latency-bound with nothing saturated, and half of its time is the per-thread prologue
// (column program, derived pose, Philox) paid by N * obs_dim threads (profiles/r2_notes.md).  Here a thread owns Q
// consecutive columns of one env: the pose and the Philox block are fetched once per thread, and the Q ray queries advance
// in lockstep -- the entry of candidate i+1 of every ray is requested together with the records of candidate i, so each trip
// of the loop is ONE round trip with 4Q independent loads in flight per thread.  The loop body is call-free: a borderline
// barycentric (|.| < 2e-5) only flags its ray, which is then redone by raycast_vertical (the fp64 decision) -- same bits as
// the one-column kernel for every ray.
// ------------------------------------------------------------------------------------------------
template <int Q>
__global__ void __launch_bounds__(64, Q >= 4 ? 8 : 12)
compute_obs_quad_kernel(const __grid_constant__ Env env_param, Sim sim, int64_t step, const ObsCol* __restrict__ cols,
                        const float* __restrict__ derived, const __grid_constant__ MeshAccel mesh, int has_mesh,
                        const float* __restrict__ spec_z, unsigned env_begin, unsigned env_count) {
  const Env* __restrict__ e = &env_param;
  const unsigned N = (unsigned)e->N, D = (unsigned)e->obs_dim, QD = (D + Q - 1) / Q;
  const unsigned lid = blockIdx.x * blockDim.x + threadIdx.x;
  if (lid >= env_count * QD) return;
  const unsigned n = env_begin + lid / QD, k0 = (lid - (lid / QD) * QD) * Q;
  float4 ca[Q], cb[Q];
  bool has_hs = false;
#pragma unroll
  for (int j = 0; j < Q; ++j) {   // static program: read before the dependency wait; the tail of a row repeats its last column
    const ObsCol* col = cols + min(k0 + j, D - 1);
    ca[j] = __ldg(&col->a); cb[j] = __ldg(&col->b);
    has_hs |= (__float_as_uint(ca[j].x) & 0xffu) == B200LAB_OBS_HEIGHT_SCAN;
  }
  pdl_wait();
  const float* o = derived + (size_t)n * kDerived;
  float px = 0.f, py = 0.f, pz = 0.f;
  f4 qy{1.f, 0.f, 0.f, 0.f};
  bool upd = false, spec = false;
  if (has_hs) {
    px = o[9]; py = o[10]; pz = o[11];
    qy = f4{o[12], o[13], o[14], o[15]};
    upd = o[17] != 0.f; spec = o[18] != 0.f;
  }
  float v[Q], sx[Q], sy[Q], sz[Q], inv_dz[Q], dzv[Q], tb[Q];
  int s0[Q], cnt[Q];
  bool hs[Q], hit[Q], border[Q];
#pragma unroll
  for (int j = 0; j < Q; ++j) {
    const unsigned w0 = __float_as_uint(ca[j].x);
    const int op = w0 & 0xffu, li = (w0 >> 8) & 0xffffu;
    v[j] = 0.f; cnt[j] = 0; s0[j] = 0; hs[j] = false; hit[j] = false; border[j] = false;
    sx[j] = sy[j] = sz[j] = 0.f; inv_dz[j] = dzv[j] = -1.f; tb[j] = e->rc_max_distance;
    switch (op) {
      case B200LAB_OBS_BASE_LIN_VEL: v[j] = o[li]; break;
      case B200LAB_OBS_BASE_ANG_VEL: v[j] = o[3 + li]; break;
      case B200LAB_OBS_PROJECTED_GRAVITY: v[j] = o[6 + li]; break;
      case B200LAB_OBS_GENERATED_COMMANDS: v[j] = e->cmd_vel_b[n * 3 + li]; break;
      case B200LAB_OBS_JOINT_POS_REL: v[j] = sim.joint_pos[(size_t)n * e->J + li] - e->default_joint_pos[li]; break;
      case B200LAB_OBS_JOINT_VEL_REL: v[j] = sim.joint_vel[(size_t)n * e->J + li] - e->default_joint_vel[li]; break;
      case B200LAB_OBS_JOINT_POS: v[j] = sim.joint_pos[(size_t)n * e->J + li]; break;
      case B200LAB_OBS_JOINT_VEL: v[j] = sim.joint_vel[(size_t)n * e->J + li]; break;
      case B200LAB_OBS_LAST_ACTION: v[j] = e->action[(size_t)n * e->A + li]; break;
      case B200LAB_OBS_BASE_POS_Z: v[j] = o[16]; break;
      case B200LAB_OBS_EXTERNAL: v[j] = e->ext_values[((size_t)__float_as_int(cb[j].w) + li) * N + n]; break;
      case B200LAB_OBS_HEIGHT_SCAN: {
        hs[j] = true;
        if (upd) {   // ray start = quat_apply_yaw(q, local start) + pos (ray_caster.py:243-253)
          const float4 cc = __ldg(&cols[min(k0 + j, D - 1)].c);
          f3 st = quat_apply(qy, f3{cc.x, cc.y, cc.z});
          sx[j] = st.x + px; sy[j] = st.y + py; sz[j] = st.z + pz;
          dzv[j] = cc.w; inv_dz[j] = 1.0f / cc.w;
          if (!spec && has_mesh) {
            const int cx = (int)floorf((sx[j] - mesh.ox) * mesh.inv_cell), cy = (int)floorf((sy[j] - mesh.oy) * mesh.inv_cell);
            if (cx >= 0 && cy >= 0 && cx < mesh.nx && cy < mesh.ny) {
              const int c = cy * mesh.nx + cx;
              s0[j] = __ldg(mesh.cell_start + c);
              cnt[j] = __ldg(mesh.cell_start + c + 1) - s0[j];
            }
          }
        }
        break;
      }
    }
  }
  // ---- the Q ray queries in lockstep
  int maxc = 0;
  int fn[Q];
#pragma unroll
  for (int j = 0; j < Q; ++j) {
    maxc = max(maxc, cnt[j]);
    fn[j] = cnt[j] > 0 ? __ldg(mesh.cell_tris + s0[j]) : -1;
  }
  for (int it = 0; it < maxc; ++it) {
    float4 ra[Q], rb[Q], rc[Q];
    bool on[Q];
#pragma unroll
    for (int j = 0; j < Q; ++j) {
      on[j] = fn[j] >= 0;
      const float4* r = mesh.rec + (size_t)(on[j] ? (fn[j] & kEntryIdMask) : 0) * 3;
      if (on[j]) { ra[j] = __ldg(r); rb[j] = __ldg(r + 1); rc[j] = __ldg(r + 2); }
    }
#pragma unroll
    for (int j = 0; j < Q; ++j) fn[j] = it + 1 < cnt[j] ? __ldg(mesh.cell_tris + s0[j] + it + 1) : -1;
#pragma unroll
    for (int j = 0; j < Q; ++j) {
      if (on[j]) {
        const float t = ray_tri_vertical_fast(ra[j], rb[j], rc[j], sx[j], sy[j], sz[j], inv_dz[j], border[j]);
        if (t > 0.0f && t <= tb[j]) { tb[j] = t; hit[j] = true; }
      }
    }
  }
  // ---- finish: hit heights, noise (one Philox block for the thread's columns), clip, scale
  const bool recorded = e->u_obs != nullptr;
  bool need3 = false, need4 = false;
#pragma unroll
  for (int j = 0; j < Q; ++j) {
    const int kind = (__float_as_uint(ca[j].x) >> 24) & 0x3fu;
    const bool rnd = (kind >= B200LAB_NOISE_UNIFORM_ADD && kind <= B200LAB_NOISE_UNIFORM_ABS) || kind > B200LAB_NOISE_CONSTANT_ABS;
    need3 |= rnd; need4 |= kind > B200LAB_NOISE_CONSTANT_ABS;
  }
  uint4 r3 = make_uint4(0, 0, 0, 0), r4 = r3;
  if (!recorded && need3) {   // Q divides 4: the thread's columns share one counter block (slot >> 2), as philox_uniform forms it
    if (e->step_counter_dev) step = *e->step_counter_dev;
    const uint2 key = make_uint2((uint32_t)e->seed, (uint32_t)(e->seed >> 32));
    const uint32_t hi = (uint32_t)((uint64_t)step >> 32);
    r3 = philox4x32_10(make_uint4(n, k0 >> 2, (uint32_t)step, hi ^ (3u << 24)), key);
    if (need4) r4 = philox4x32_10(make_uint4(n, k0 >> 2, (uint32_t)step, hi ^ (4u << 24)), key);
  }
#pragma unroll
  for (int j = 0; j < Q; ++j) {
    const unsigned k = k0 + j;
    if (k >= D) break;
    const unsigned w0 = __float_as_uint(ca[j].x), gid = n * D + k;
    const int li = (w0 >> 8) & 0xffffu, kind = (w0 >> 24) & 0x3fu;
    if (hs[j]) {
      float* hp = e->rc_ray_hits_w ? e->rc_ray_hits_w + ((size_t)n * e->R + li) * 3 : nullptr;
      float hz;
      if (upd) {
        f3 h3{INFINITY, INFINITY, INFINITY};
        if (spec) {   // speculative scan cast from this very pose (height_scan_spec_kernel)
          const float z = spec_z[(size_t)n * e->R + li];
          if (z != INFINITY) h3 = {sx[j], sy[j], z};
        } else if (border[j]) {   // some candidate sat on an edge: the fp64 path decides the whole ray
          float tt; int face;
          if (raycast_vertical(mesh, f3{sx[j], sy[j], sz[j]}, dzv[j], e->rc_max_distance, tt, face)) h3 = {sx[j], sy[j], sz[j] + tt * dzv[j]};
        } else if (hit[j]) {
          h3 = {sx[j], sy[j], sz[j] + tb[j] * dzv[j]};
        }
        if (hp) { hp[0] = h3.x; hp[1] = h3.y; hp[2] = h3.z; }
        hz = h3.z;
      } else {
        hz = hp ? hp[2] : INFINITY;
      }
      v[j] = (pz - hz) - cb[j].z;   // observations.py:173
    }
    float u1 = 0.f, u2 = 0.f;
    if (recorded) u1 = e->u_obs[gid];
    else {
      const unsigned lane = k & 3u;
      const uint32_t a3 = lane == 0 ? r3.x : lane == 1 ? r3.y : lane == 2 ? r3.z : r3.w;
      const uint32_t a4 = lane == 0 ? r4.x : lane == 1 ? r4.y : lane == 2 ? r4.z : r4.w;
      u1 = (float)(a3 >> 8) * (1.0f / 16777216.0f);
      u2 = (float)(a4 >> 8) * (1.0f / 16777216.0f);
    }
    e->obs_buf[gid] = obs_apply(v[j], ca[j], cb[j], w0, kind, u1, u2, recorded);
  }
}

// noise -> clip -> scale of one observation element (observation_manager.py:313-318).  u1: the element's uniform draw (or,
// in parity mode, the recorded draw of whatever distribution the term uses); u2: second uniform of a Gaussian's Box-Muller.
__device__ __forceinline__ float obs_apply(float v, float4 ca, float4 cb, unsigned w0, int kind, float u1, float u2, bool recorded) {
  if (kind != B200LAB_NOISE_NONE) {
    const float noise_lo = ca.y, noise_span = ca.z;
    if (kind == B200LAB_NOISE_CONSTANT_ADD) v = v + noise_lo;             // noise_model.py:33-38
    else if (kind == B200LAB_NOISE_CONSTANT_SCALE) v = v * noise_lo;
    else if (kind == B200LAB_NOISE_CONSTANT_ABS) v = 0.0f + noise_lo;
    else if (kind <= B200LAB_NOISE_UNIFORM_ABS) {
      float u = u1;
      float nz = u * noise_span + noise_lo;
      if (kind == B200LAB_NOISE_UNIFORM_ADD) v = (v + u * noise_span) + noise_lo;  // noise_model.py:64
      else if (kind == B200LAB_NOISE_UNIFORM_SCALE) v = v * nz;
      else v = nz;
    } else {
      // parity mode: the recorded randn_like draw; else Box-Muller on two Philox uniforms of this element
      float z = recorded ? u1 : sqrtf(-2.0f * logf(1.0f - u1)) * cospif(2.0f * u2);
      float nz = noise_lo + noise_span * z;
      if (kind == B200LAB_NOISE_GAUSSIAN_ADD) v = (v + noise_lo) + noise_span * z;   // noise_model.py:91
      else if (kind == B200LAB_NOISE_GAUSSIAN_SCALE) v = v * nz;
      else v = nz;
    }
  }
  if (w0 & (1u << 30)) v = fminf(fmaxf(v, cb.x), cb.y);
  if (w0 & (1u << 31)) v = v * ca.w;
  return v;
}

__device__ __forceinline__ float obs_finish(const Env* __restrict__ e, float v, float4 ca, float4 cb, unsigned w0, int kind,
                                            int64_t step, unsigned n, unsigned k, unsigned gid) {
  float u1 = 0.f, u2 = 0.f;
  if (kind >= B200LAB_NOISE_UNIFORM_ADD && kind != B200LAB_NOISE_CONSTANT_ADD && kind != B200LAB_NOISE_CONSTANT_SCALE &&
      kind != B200LAB_NOISE_CONSTANT_ABS) {
    if (e->u_obs) u1 = e->u_obs[gid];
    else {
      if (e->step_counter_dev) step = *e->step_counter_dev;
      u1 = philox_uniform(e->seed, step, 3u, n, k);
      if (kind > B200LAB_NOISE_CONSTANT_ABS) u2 = philox_uniform(e->seed, step, 4u, n, k);
    }
  }
  return obs_apply(v, ca, cb, w0, kind, u1, u2, e->u_obs != nullptr);
}

// ------------------------------------------------------------------------------------------------
// compute_obs, STAGED variant: the height scan of one env runs out of shared memory.
//
// One CTA per env.  The 187 rays of an env land in a ~1.6 x 1.0 m yaw-rotated footprint, i.e. a few dozen rows of the
// mesh's cell grid.  Instead of every ray walking  cell range -> triangle ids -> records  through three dependent,
// fully divergent L1/L2 requests (the per-ray kernel above), the CTA
//   A. computes its rays' cells and, per grid row, the span of cells they touch (shared-memory atomics),
//   B. stages, with coalesced loads, (1) the `cell_start` values of those spans, (2) the entry lists of those cells,
//      translated on the fly into shared-memory record slots, and (3) the packed records of every small triangle whose
//      home cell can overlap the spans -- triangles are sorted by home cell at build time (raycast.cuh), so this is one
//      contiguous run per grid row;
//   C. runs the unchanged per-ray test (same candidates, same arithmetic => bitwise the per-ray kernel's hits) against
//      shared memory.  Entries that were not staged (large triangles, capacity overflow) fall back to the global record.
// The proprioceptive columns ride in the same CTA; the observation row is written coalesced.
// ------------------------------------------------------------------------------------------------
constexpr int kPatchThreads = 256;
#ifndef B200LAB_PATCH_REC
#define B200LAB_PATCH_REC 704
#endif
#ifndef B200LAB_PATCH_BLOCKS
#define B200LAB_PATCH_BLOCKS 4
#endif
constexpr int kPatchRec = B200LAB_PATCH_REC;   // staged records (48 B each)
constexpr int kPatchIds = 2816;                // staged cell entries
constexpr int kPatchCs = 2048;                 // staged cell_start values
constexpr int kPatchRows = 56;                 // grid rows under one footprint (<= 64: two per lane of one warp)
constexpr int kPatchHRows = kPatchRows + kHomeMargin;
constexpr int kPatchMaxCols = 2;               // observation columns per thread (obs_dim <= 512)

struct PatchSmem {
  float4 rec[kPatchRec * 3];
  int ids[kPatchIds];
  int cs[kPatchCs];
  int row_xmin[kPatchRows], row_xmax[kPatchRows];
  int row_cs_off[kPatchRows];     // offset in cs[] of the row's span; -1: row not staged
  int row_cs_end[kPatchRows];     // end offset (== start for a row that is not staged)
  int row_ids_off[kPatchRows];    // offset in ids[] of the row's first entry
  int row_ids_end[kPatchRows];
  int row_ids_base[kPatchRows];   // index in mesh.cell_tris of that entry
  int hrow_h0[kPatchHRows], hrow_h1[kPatchHRows], hrow_off[kPatchHRows];   // staged id range / first slot per home row
  int ymin, ymax;
  int tot_rec, tot_cs, tot_ids;   // staged item counts (flat staging loops)
};

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}

__device__ __forceinline__ int warp_excl_scan(int v, int lane, int& total) {
  int incl = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    int t = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += t;
  }
  total = __shfl_sync(0xffffffffu, incl, 31);
  return incl - v;
}

template <int kCols>   // observation columns per thread: 1 for obs_dim <= 256 (half the live registers), else 2
__global__ void __launch_bounds__(kPatchThreads, B200LAB_PATCH_BLOCKS)
compute_obs_patch_kernel(const __grid_constant__ Env env_param, Sim sim, int64_t step, const ObsCol* __restrict__ cols,
                         const float* __restrict__ derived, const __grid_constant__ MeshAccel mesh, unsigned env_begin) {
  extern __shared__ __align__(16) unsigned char patch_smem_raw[];
  PatchSmem& S = *reinterpret_cast<PatchSmem*>(patch_smem_raw);
  const Env* __restrict__ e = &env_param;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const unsigned n = env_begin + blockIdx.x;
  const unsigned N = (unsigned)e->N, D = (unsigned)e->obs_dim;
  const float* o = derived + (size_t)n * kDerived;
  // static per-column programs: read before the dependency wait
  float4 ca[kCols], cb[kCols], cc[kCols];
#pragma unroll
  for (int s = 0; s < kCols; ++s) {
    const unsigned k = tid + s * kPatchThreads;
    ca[s] = cb[s] = cc[s] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (k < D) { ca[s] = __ldg(&cols[k].a); cb[s] = __ldg(&cols[k].b); cc[s] = __ldg(&cols[k].c); }
  }
  if (tid == 0) { S.ymin = INT_MAX; S.ymax = INT_MIN; S.tot_rec = S.tot_cs = S.tot_ids = 0; }
  if (tid < kPatchRows) { S.row_xmin[tid] = INT_MAX; S.row_xmax[tid] = INT_MIN; S.row_cs_off[tid] = -1; }
  __syncthreads();
  pdl_wait();
  // ---- A: ray starts and their cells
  const bool outdated = o[17] != 0.f;
  const float pos_x = o[9], pos_y = o[10], pos_z = o[11];
  const f4 qy{o[12], o[13], o[14], o[15]};
  f3 st[kCols];
  int cx[kCols], cy[kCols];
  bool is_ray[kCols], in_grid[kCols];
#pragma unroll
  for (int s = 0; s < kCols; ++s) {
    const unsigned k = tid + s * kPatchThreads;
    const int op = __float_as_uint(ca[s].x) & 0xffu;
    is_ray[s] = k < D && op == B200LAB_OBS_HEIGHT_SCAN && outdated;
    in_grid[s] = false;
    cx[s] = cy[s] = 0;
    st[s] = f3{0.f, 0.f, 0.f};
    if (is_ray[s]) {
      // ray start = quat_apply_yaw(q, local start) + pos (ray_caster.py:243-253); same expressions as the per-ray kernel
      f3 p = quat_apply(qy, f3{cc[s].x, cc[s].y, cc[s].z});
      p.x += pos_x; p.y += pos_y; p.z += pos_z;
      st[s] = p;
      cx[s] = (int)floorf((p.x - mesh.ox) * mesh.inv_cell);
      cy[s] = (int)floorf((p.y - mesh.oy) * mesh.inv_cell);
      in_grid[s] = cx[s] >= 0 && cy[s] >= 0 && cx[s] < mesh.nx && cy[s] < mesh.ny;
    }
  }
  {
    int lo = INT_MAX, hi = INT_MIN;
#pragma unroll
    for (int s = 0; s < kCols; ++s)
      if (in_grid[s]) { lo = min(lo, cy[s]); hi = max(hi, cy[s]); }
    lo = __reduce_min_sync(0xffffffffu, lo);
    hi = __reduce_max_sync(0xffffffffu, hi);
    if (lane == 0 && lo <= hi) { atomicMin(&S.ymin, lo); atomicMax(&S.ymax, hi); }
  }
  __syncthreads();
  const int ry0 = S.ymin;
  const int nrows = S.ymax >= ry0 ? S.ymax - ry0 + 1 : 0;
  const bool staged = nrows > 0 && nrows <= kPatchRows;
  if (staged) {
#pragma unroll
    for (int s = 0; s < kCols; ++s)
      if (in_grid[s]) { atomicMin(&S.row_xmin[cy[s] - ry0], cx[s]); atomicMax(&S.row_xmax[cy[s] - ry0], cx[s]); }
  }
  __syncthreads();
  // ---- B1: per-row spans -> global ranges and shared-memory offsets (warp 0: cell rows, warp 1: home rows).
  // Offsets are the raw exclusive scans; a row that would overflow the staging area keeps count 0 (its rays take the
  // global path), which leaves a hole in the item index space that the flat loops below skip.
  if (staged && warp == 0) {
    int carry_ids = 0, carry_cs = 0;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const int r = lane + 32 * half;
      int nids = 0, ncs = 0, a = 0;
      if (r < nrows) {
        const int x0 = S.row_xmin[r], x1 = S.row_xmax[r];
        if (x0 <= x1) {
          const int base = (ry0 + r) * mesh.nx;
          a = __ldg(mesh.cell_start + base + x0);
          nids = __ldg(mesh.cell_start + base + x1 + 1) - a;
          ncs = x1 - x0 + 2;
        }
      }
      int tot_ids, tot_cs;
      const int off_ids = carry_ids + warp_excl_scan(nids, lane, tot_ids);
      const int off_cs = carry_cs + warp_excl_scan(ncs, lane, tot_cs);
      carry_ids += tot_ids; carry_cs += tot_cs;
      if (r < kPatchRows) {
        const bool ok = r < nrows && ncs > 0 && off_ids + nids <= kPatchIds && off_cs + ncs <= kPatchCs;
        S.row_cs_off[r] = ok ? off_cs : -1;
        S.row_cs_end[r] = off_cs + (ok ? ncs : 0);
        S.row_ids_off[r] = off_ids;
        S.row_ids_end[r] = off_ids + (ok ? nids : 0);
        S.row_ids_base[r] = a;
      }
    }
    if (lane == 0) { S.tot_ids = min(carry_ids, kPatchIds); S.tot_cs = min(carry_cs, kPatchCs); }
  } else if (staged && warp == 1) {
    int carry = 0;
    const int nh = nrows + kHomeMargin;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const int t = lane + 32 * half;     // home row ry0 - kHomeMargin + t serves cell rows t-kHomeMargin .. t
      int h0 = 0, cnt = 0;
      if (t < nh) {
        int hx0 = INT_MAX, hx1 = INT_MIN;
        for (int r = max(t - kHomeMargin, 0); r <= min(t, nrows - 1); ++r) {
          const int x0 = S.row_xmin[r], x1 = S.row_xmax[r];
          if (x0 <= x1) { hx0 = min(hx0, x0 - kHomeMargin); hx1 = max(hx1, x1); }
        }
        const int hr = ry0 - kHomeMargin + t;
        if (hx0 <= hx1 && hr >= 0 && hr < mesh.ny) {
          hx0 = max(hx0, 0);
          const int base = hr * mesh.nx;
          h0 = __ldg(mesh.home_start + base + hx0);
          cnt = __ldg(mesh.home_start + base + hx1 + 1) - h0;
        }
      }
      int tot;
      const int off = carry + warp_excl_scan(cnt, lane, tot);
      carry += tot;
      if (t < kPatchHRows) {
        if (t >= nh || off + cnt > kPatchRec) cnt = 0;   // capacity: the row's triangles stay in global memory
        S.hrow_h0[t] = h0; S.hrow_h1[t] = h0 + cnt; S.hrow_off[t] = off;
      }
    }
    if (lane == 0) S.tot_rec = min(carry, kPatchRec);
  }
  __syncthreads();
  // ---- B2: stage with cp.async (global -> shared without a register round trip): a warp owns rows round-robin and
  // issues the copies of ALL its rows back to back -- nothing here waits on a load, so every structure costs the CTA one
  // L2 round trip.  Cell entries are staged raw; they are translated to record slots where they are used (stage C).
  if (staged) {
    const int nh = nrows + kHomeMargin;
    for (int t = warp; t < nh; t += kPatchThreads / 32) {
      const int h0 = S.hrow_h0[t], n3 = (S.hrow_h1[t] - h0) * 3;
      const float4* src = mesh.rec + (size_t)h0 * 3;
      float4* dst = S.rec + S.hrow_off[t] * 3;
      for (int i = lane; i < n3; i += 32) cp_async16(dst + i, src + i);
    }
    for (int r = warp; r < nrows; r += kPatchThreads / 32) {
      const int cs_off = S.row_cs_off[r];
      if (cs_off < 0) continue;
      const int ncs = S.row_cs_end[r] - cs_off;
      const int* src = mesh.cell_start + (ry0 + r) * mesh.nx + S.row_xmin[r];
      for (int i = lane; i < ncs; i += 32) cp_async4(S.cs + cs_off + i, src + i);
      const int ids_off = S.row_ids_off[r], nids = S.row_ids_end[r] - ids_off;
      const int* isrc = mesh.cell_tris + S.row_ids_base[r];
      for (int i = lane; i < nids; i += 32) cp_async4(S.ids + ids_off + i, isrc + i);
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
  }
  __syncthreads();
  // ---- C + D: per-column values
#pragma unroll
  for (int s = 0; s < kCols; ++s) {
    const unsigned k = tid + s * kPatchThreads;
    if (k >= D) continue;
    const unsigned gid = n * D + k;
    const unsigned w0 = __float_as_uint(ca[s].x);
    const int op = w0 & 0xffu, li = (w0 >> 8) & 0xffffu, kind = (w0 >> 24) & 0x3fu;
    float v = 0.f;
    switch (op) {
      case B200LAB_OBS_BASE_LIN_VEL: v = o[li]; break;
      case B200LAB_OBS_BASE_ANG_VEL: v = o[3 + li]; break;
      case B200LAB_OBS_PROJECTED_GRAVITY: v = o[6 + li]; break;
      case B200LAB_OBS_GENERATED_COMMANDS: v = e->cmd_vel_b[n * 3 + li]; break;
      case B200LAB_OBS_JOINT_POS_REL: v = sim.joint_pos[(size_t)n * e->J + li] - e->default_joint_pos[li]; break;
      case B200LAB_OBS_JOINT_VEL_REL: v = sim.joint_vel[(size_t)n * e->J + li] - e->default_joint_vel[li]; break;
      case B200LAB_OBS_JOINT_POS: v = sim.joint_pos[(size_t)n * e->J + li]; break;
      case B200LAB_OBS_JOINT_VEL: v = sim.joint_vel[(size_t)n * e->J + li]; break;
      case B200LAB_OBS_LAST_ACTION: v = e->action[(size_t)n * e->A + li]; break;
      case B200LAB_OBS_BASE_POS_Z: v = o[16]; break;
      case B200LAB_OBS_EXTERNAL: v = e->ext_values[((size_t)__float_as_int(cb[s].w) + li) * N + n]; break;
      case B200LAB_OBS_HEIGHT_SCAN: {
        float* hp = e->rc_ray_hits_w ? e->rc_ray_hits_w + ((size_t)n * e->R + li) * 3 : nullptr;
        float hz;
        if (is_ray[s]) {
          const float dz = cc[s].w;
          float t_best = e->rc_max_distance;
          int face_best = -1;
          const int r = cy[s] - ry0;
          if (in_grid[s] && staged && S.row_cs_off[r] >= 0) {
            const int ci = S.row_cs_off[r] + (cx[s] - S.row_xmin[r]);
            const int k0 = S.cs[ci], k1 = S.cs[ci + 1];
            const int ib = S.row_ids_off[r] - S.row_ids_base[r];
            const float inv_dz = 1.0f / dz;
            for (int kk = k0; kk < k1; ++kk) {
              const int ent = S.ids[ib + kk];
              const int f = ent & kEntryIdMask;
              const int t = r - (ent >> kEntryShift) + kHomeMargin;   // home row of a small triangle, relative to the patch
              float4 ra, rb, rc;
              int h0;
              if (f < mesh.first_large && t >= 0 && f >= (h0 = S.hrow_h0[t]) && f < S.hrow_h1[t]) {
                const int slot = S.hrow_off[t] + (f - h0);
                ra = S.rec[slot * 3]; rb = S.rec[slot * 3 + 1]; rc = S.rec[slot * 3 + 2];
              } else {   // large triangle, or a home row that did not fit the staging area
                const float4* rp = mesh.rec + (size_t)f * 3;
                ra = __ldg(rp); rb = __ldg(rp + 1); rc = __ldg(rp + 2);
              }
              const float th = ray_tri_vertical_rec(mesh, f, ra, rb, rc, st[s], dz, inv_dz);
              if (th > 0.0f && th <= t_best) { t_best = th; face_best = f; }
            }
          } else if (in_grid[s]) {
            float tt; int face;
            if (raycast_vertical(mesh, st[s], dz, e->rc_max_distance, tt, face)) { t_best = tt; face_best = face; }
          }
          f3 hit{INFINITY, INFINITY, INFINITY};
          if (face_best >= 0) hit = {st[s].x, st[s].y, st[s].z + t_best * dz};
          if (hp) { hp[0] = hit.x; hp[1] = hit.y; hp[2] = hit.z; }
          hz = hit.z;
        } else {
          hz = hp ? hp[2] : INFINITY;
        }
        v = (pos_z - hz) - cb[s].z;   // observations.py:173
        break;
      }
    }
    e->obs_buf[gid] = obs_finish(e, v, ca[s], cb[s], w0, kind, step, n, k, gid);
  }
}


// extras["log"] fold (see b200lab.h): one block; runs on the side stream of step_tail, off the critical path
__global__ void __launch_bounds__(256)
log_fold_kernel(const __grid_constant__ Env env_param) {
  const Env* __restrict__ e = &env_param;
  __shared__ double s_red[256];
  const int tid = threadIdx.x, nr = e->num_rew, nt = e->num_term;
  const double cnt = e->log_accum[0];
  double lv = 0.0;
  if (cnt > 0.0 && e->has_terrain_curriculum)
    for (int n = tid; n < e->N; n += blockDim.x) lv += (double)e->terrain_levels[n];
  s_red[tid] = lv;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (tid < o) s_red[tid] += s_red[tid + o];
    __syncthreads();
  }
  const double mean_level = s_red[0] / (double)e->N;
  const int slots = 4 + nr + nt;
  double* run = e->log_running;
  double* last = run + B200LAB_LOG_SLOTS;
  for (int i = tid; i < slots; i += blockDim.x) {
    double v = 0.0;
    if (cnt > 0.0) {
      const double a = e->log_accum[i];
      if (i == 0) v = 1.0;
      else if (i < 1 + nr) v = a / cnt / (double)e->max_episode_length_s;
      else if (i < 1 + nr + nt) v = a;
      else if (i < 3 + nr + nt) v = a / cnt;
      else v = mean_level;
      run[i] += v;
    }
    last[i] = i == 0 ? cnt : v;
  }
  __syncthreads();
  for (int i = tid; i < slots; i += blockDim.x) e->log_accum[i] = 0.0;
}

static int launch_geometry(int width, int* wpad, int* epb) {
  int w = width < 1 ? 1 : width;
  if (w > 256) return -1;
  *wpad = w;
  *epb = 256 / w;
  return 0;
}
}  // namespace b200lab

using namespace b200lab;

// per-column observation program, rebuilt whenever the descriptor changes
static int upload_obs_tables(EnvHandle* h) {
  const B200LabEnvState& s = h->host;
  h->rays_vertical = false;
  h->has_mesh = false;
  std::vector<float> starts, dirs;
  if (s.R > 0 && s.rc_ray_dirs && s.rc_ray_starts) {
    starts.resize((size_t)s.R * 3);
    dirs.resize((size_t)s.R * 3);
    cudaError_t err = cudaMemcpy(dirs.data(), s.rc_ray_dirs, dirs.size() * sizeof(float), cudaMemcpyDeviceToHost);
    if (err == cudaSuccess) err = cudaMemcpy(starts.data(), s.rc_ray_starts, starts.size() * sizeof(float), cudaMemcpyDeviceToHost);
    if (err != cudaSuccess) { set_error("ray pattern: %s", cudaGetErrorString(err)); return B200LAB_ERR_CUDA; }
    if (s.rc_attach_yaw_only) {
      bool ok = true;
      for (int r = 0; r < s.R; ++r) ok = ok && dirs[3 * r] == 0.f && dirs[3 * r + 1] == 0.f && dirs[3 * r + 2] != 0.f;
      h->rays_vertical = ok;
    }
  }
  if (s.rc_mesh) {
    cudaError_t err = cudaMemcpy(&h->mesh, s.rc_mesh, sizeof(MeshAccel), cudaMemcpyDeviceToHost);
    if (err != cudaSuccess) { set_error("ray-caster mesh descriptor: %s", cudaGetErrorString(err)); return B200LAB_ERR_CUDA; }
    h->has_mesh = getenv("B200LAB_OBS_NOMESH") == nullptr;   // measurement knob (tools/obs_probe.py): every ray misses
  }
  // speculative-scan buffers; any descriptor change invalidates the recorded poses (NaN never compares equal to a pose)
  if (h->side2) cudaStreamSynchronize(h->side2);
  h->spec_pending = false;
  h->derived_root = nullptr;
  const int64_t need = h->rays_vertical && h->has_mesh ? (int64_t)s.N * s.R : 0;
  if (need > h->spec_cap) {
    if (h->spec_z) cudaFree(h->spec_z);
    if (h->spec_pose) cudaFree(h->spec_pose);
    h->spec_z = h->spec_pose = nullptr;
    h->spec_cap = 0;
    cudaError_t err = cudaMalloc((void**)&h->spec_z, (size_t)need * sizeof(float));
    if (err == cudaSuccess) err = cudaMalloc((void**)&h->spec_pose, (size_t)s.N * kSpecPose * sizeof(float));
    if (err != cudaSuccess) { set_error("speculative scan buffers: %s", cudaGetErrorString(err)); return B200LAB_ERR_CUDA; }
    h->spec_cap = need;
  }
  if (h->spec_pose) cudaMemset(h->spec_pose, 0xFF, (size_t)s.N * kSpecPose * sizeof(float));
  int D = s.obs_dim;
  if (D <= 0) return B200LAB_OK;
  if ((int64_t)s.N * D >= (int64_t)1 << 31) { set_error("N * obs_dim must be < 2^31"); return B200LAB_ERR_ARG; }
  std::vector<ObsCol> cols((size_t)D);
  h->obs_plain = true;   // no history, no modifiers: the staged kernel (which knows neither) may be used
  auto as_f = [](uint32_t u) { float f; memcpy(&f, &u, 4); return f; };
  int k = 0;
  for (int t = 0; t < s.num_obs; ++t) {
    const B200LabTerm& tm = s.obs[t];
    bool joint = tm.op == B200LAB_OBS_JOINT_POS_REL || tm.op == B200LAB_OBS_JOINT_VEL_REL ||
                 tm.op == B200LAB_OBS_JOINT_POS || tm.op == B200LAB_OBS_JOINT_VEL;
    if (tm.op < 0 || tm.op > 255 || tm.noise_kind < 0 || tm.noise_kind > 63) { set_error("observation term %d: bad op / noise kind", t); return B200LAB_ERR_ARG; }
    const int Hn = tm.history_length > 0 ? tm.history_length : 0, rows = Hn > 0 ? Hn : 1;
    if (Hn > 255 || tm.num_mod < 0 || tm.num_mod > 3) { set_error("observation term %d: history_length > 255 or > 3 modifiers", t); return B200LAB_ERR_ARG; }
    if (Hn > 0 && (!s.obs_hist || !s.obs_hist_count || !s.obs_hist_reset_step || tm.hist_offset < 0 ||
                   tm.hist_offset + Hn * tm.dim > s.obs_hist_total)) {
      set_error("observation term %d: history buffers missing or too small", t); return B200LAB_ERR_ARG;
    }
    uint32_t modk = 0;
    for (int q = 0; q < tm.num_mod; ++q) {
      if (tm.mod_kind[q] < 1 || tm.mod_kind[q] > 3) { set_error("observation term %d: bad modifier kind", t); return B200LAB_ERR_ARG; }
      modk |= (uint32_t)tm.mod_kind[q] << (2 * q);
    }
    if (Hn > 0 || modk) h->obs_plain = false;
    const int k_term = k;
    for (int row = 0; row < rows; ++row) {
    uint64_t m = tm.joint_mask;
    for (int li = 0; li < tm.dim; ++li) {
      if (k >= D) { set_error("observation term dims exceed obs_dim"); return B200LAB_ERR_ARG; }
      int src = li;
      if (joint) {
        if (!m) { set_error("observation term %d: dim exceeds the number of selected joints", t); return B200LAB_ERR_ARG; }
        src = __builtin_ctzll(m);
        m &= m - 1;
      }
      if (src > 0xffff) { set_error("observation term %d: more than 65535 columns", t); return B200LAB_ERR_ARG; }
      uint32_t kind = s.obs_enable_corruption ? (uint32_t)tm.noise_kind : (uint32_t)B200LAB_NOISE_NONE;
      uint32_t w0 = (uint32_t)tm.op | ((uint32_t)src << 8) | (kind << 24) | (tm.clip_on ? 1u << 30 : 0u) | (tm.scale_on ? 1u << 31 : 0u);
      ObsCol& c = cols[k];
      c.a = make_float4(as_f(w0), tm.noise_lo, tm.noise_span, tm.scale);
      c.b = make_float4(tm.clip_lo, tm.clip_hi, tm.p[0], as_f((uint32_t)tm.i0));
      c.c = make_float4(0.f, 0.f, 0.f, 0.f);
      if (tm.op == B200LAB_OBS_HEIGHT_SCAN) {
        if (li >= s.R || starts.empty()) { set_error("height_scan term wider than the ray pattern"); return B200LAB_ERR_ARG; }
        c.c = make_float4(starts[3 * li], starts[3 * li + 1], starts[3 * li + 2], dirs[3 * li + 2]);
      }
      c.d = make_float4(as_f((uint32_t)Hn | ((uint32_t)row << 8) | (modk << 16)), as_f((uint32_t)(tm.hist_offset + li)),
                        as_f((uint32_t)tm.dim), as_f((uint32_t)(k_term + (rows - 1) * tm.dim + li)));
      c.m = make_float4(tm.mod_p[0], tm.mod_p[1], tm.mod_p[2], tm.mod_p[3]);
      ++k;
    }
    }
  }
  if (k != D) { set_error("observation term dims (%d) do not add up to obs_dim (%d)", k, D); return B200LAB_ERR_ARG; }
  if (h->cols) { cudaFree(h->cols); h->cols = nullptr; }
  cudaError_t err = cudaMalloc((void**)&h->cols, (size_t)D * sizeof(ObsCol));
  if (err == cudaSuccess) err = cudaMemcpy(h->cols, cols.data(), (size_t)D * sizeof(ObsCol), cudaMemcpyHostToDevice);
  if (err != cudaSuccess) { set_error("obs tables: %s", cudaGetErrorString(err)); return B200LAB_ERR_CUDA; }
  return B200LAB_OK;
}

typedef NetLstm<8, 2> NetLstm82;
typedef NetMlp<6, 32, 3> NetMlp63;
static_assert(sizeof(NetLstm82) <= 2560 * sizeof(float), "EnvHandle::lstm_w too small");

// dynamic shared memory of the NetMlp63 kernels: raw weights (16-byte padded) + one [H][threads] activation scratch
static size_t mlp63_smem(int threads) {
  return (size_t)(((NetMlp63::kCount + 3) & ~3) + NetMlp63::kH * threads) * sizeof(float);
}

static int stage_const_net(EnvHandle* h) {
  const B200LabEnvState& s = h->host;
  const B200LabActuatorGroup* net = nullptr;
  for (int g = 0; g < s.num_act_groups; ++g)
    if (s.act[g].kind == B200LAB_ACT_NET_LSTM) net = &s.act[g];
  // ActuatorNetMLP of the shape that has a constant-bank instantiation (6-32-32-32-1, history <= 4)
  const B200LabActuatorGroup* mlp = nullptr;
  for (int g = 0; g < s.num_act_groups; ++g)
    if (s.act[g].kind == B200LAB_ACT_NET_MLP) mlp = &s.act[g];
  h->mlp_const = false;
  if (!net && mlp && mlp->net_weights && mlp->mlp_layers == NetMlp63::kNH + 1 && mlp->mlp_dims[0] == NetMlp63::kIn &&
      mlp->mlp_dims[1] == NetMlp63::kH && mlp->mlp_dims[2] == NetMlp63::kH && mlp->mlp_dims[3] == NetMlp63::kH &&
      mlp->mlp_dims[4] == 1 && mlp->mlp_hist <= kMlpMaxHist && 2 * mlp->mlp_num_idx == NetMlp63::kIn) {
    h->mlp_const = true;
    h->lstm_const = false;
    h->lstm_src = nullptr;
    return B200LAB_OK;
  }
  if (!net || net->lstm_hidden != NetLstm82::kH || net->lstm_layers != NetLstm82::kL || !net->net_weights) {
    h->lstm_const = false;
    h->lstm_src = nullptr;
    return B200LAB_OK;
  }
  if (h->lstm_const && h->lstm_src == net->net_weights) return B200LAB_OK;
  // raw layout (b200lab.h): per layer W_ih[4H,in] W_hh[4H,H] b_ih[4H] b_hh[4H]; then head W[H] b[1] -> packed image
  std::vector<float> raw(NetLstm82::kRawCount);
  cudaError_t err = cudaMemcpy(raw.data(), net->net_weights, raw.size() * sizeof(float), cudaMemcpyDeviceToHost);
  if (err != cudaSuccess) { set_error("actuator net weights: %s", cudaGetErrorString(err)); return B200LAB_ERR_CUDA; }
  {
    constexpr int H = NetLstm82::kH, L = NetLstm82::kL;
    NetLstm82* pk = reinterpret_cast<NetLstm82*>(h->lstm_w);
    const float* w = raw.data();
    for (int l = 0; l < L; ++l) {
      const int in = NetLstm82::in_dim(l), plen = NetLstm82::pair_len(l);
      const float* w_ih = w;
      const float* w_hh = w_ih + 4 * H * in;
      const float* b_ih = w_hh + 4 * H * H;
      const float* b_hh = b_ih + 4 * H;
      w = b_hh + 4 * H;
      for (int u = 0; u < H; ++u)
        for (int pr = 0; pr < 2; ++pr) {
          const int ra = (2 * pr) * H + u, rb = (2 * pr + 1) * H + u;   // gate order i, f, g, o (torch.nn.LSTM)
          float2* o = pk->w2 + NetLstm82::layer_off(l) + (u * 2 + pr) * plen;
          volatile float ba = b_ih[ra] + b_hh[ra], bb = b_ih[rb] + b_hh[rb];   // fp32 adds, as the kernel performed them
          o[0] = make_float2(ba, bb);
          for (int k = 0; k < in; ++k) o[1 + k] = make_float2(w_ih[ra * in + k], w_ih[rb * in + k]);
          for (int k = 0; k < H; ++k) o[1 + in + k] = make_float2(w_hh[ra * H + k], w_hh[rb * H + k]);
        }
    }
    for (int k = 0; k <= H; ++k) pk->head[k] = w[k];
  }
  h->lstm_const = true;
  h->lstm_src = net->net_weights;
  return B200LAB_OK;
}

static void launch_substep_pre(EnvHandle* h, const B200LabSimView* sim, bool apply_action, bool only_if_any_reset,
                               size_t smem, cudaStream_t stream) {
  int64_t total = (int64_t)h->host.N * h->host.J;
  unsigned blocks = (unsigned)ceil_div(total, kActBlock);
  if (h->lstm_const)
    substep_pre_kernel<NetLstm82><<<blocks, kActBlock, 0, stream>>>(h->host, *sim, apply_action, only_if_any_reset,
                                                                    *reinterpret_cast<const NetLstm82*>(h->lstm_w));
  else if (h->mlp_const)
    substep_pre_kernel<NetMlp63><<<blocks, kActBlock, mlp63_smem(kActBlock), stream>>>(h->host, *sim, apply_action, only_if_any_reset, NetMlp63{});
  else
    substep_pre_kernel<NetNone><<<blocks, kActBlock, smem, stream>>>(h->host, *sim, apply_action, only_if_any_reset, NetNone{});
}

extern "C" {

void* b200lab_env_create(const B200LabEnvState* st) {
  if (!st) { set_error("b200lab_env_create: null state"); return nullptr; }
  EnvHandle* h = (EnvHandle*)malloc(sizeof(EnvHandle));
  if (!h) return nullptr;
  h->host = *st;
  h->cols = nullptr;
  h->has_mesh = false;
  {
    const char* impl = getenv("B200LAB_OBS_IMPL");
    const char* vb = getenv("B200LAB_VERT_BATCH");
    h->vert_batch = vb ? atoi(vb) : kVertBatch;
    h->obs_impl = impl ? atoi(impl) : 0;   // measured on B200 (profiles/r2_notes.md): per-ray 45 us, staged 90 us at 4096 envs
  }
  h->side2 = nullptr;
  h->ev_spec_fork = h->ev_spec_done = nullptr;
  h->spec_z = h->spec_pose = nullptr;
  h->spec_cap = 0;
  h->spec_pending = false;
  h->derived = nullptr;
  h->lstm_const = false;
  h->mlp_const = false;
  h->lstm_src = nullptr;
  h->dev = nullptr;
  h->side = nullptr;
  h->ev_fork = h->ev_join = nullptr;
  cudaError_t err = cudaMalloc((void**)&h->derived, (size_t)st->N * 20 * sizeof(float));
  if (err == cudaSuccess) err = cudaStreamCreateWithFlags(&h->side, cudaStreamNonBlocking);
  if (err == cudaSuccess) err = cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming);
  if (err == cudaSuccess) err = cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming);
  if (err == cudaSuccess) err = cudaStreamCreateWithFlags(&h->side2, cudaStreamNonBlocking);
  if (err == cudaSuccess) err = cudaEventCreateWithFlags(&h->ev_spec_fork, cudaEventDisableTiming);
  if (err == cudaSuccess) err = cudaEventCreateWithFlags(&h->ev_spec_done, cudaEventDisableTiming);
  if (err == cudaSuccess) err = cudaMalloc((void**)&h->dev, sizeof(B200LabEnvState));
  if (err == cudaSuccess) err = cudaMemcpy(h->dev, st, sizeof(B200LabEnvState), cudaMemcpyHostToDevice);
  if (err == cudaSuccess && upload_obs_tables(h) != 0) err = cudaErrorUnknown;
  if (err == cudaSuccess && stage_const_net(h) != 0) err = cudaErrorUnknown;
  if (err != cudaSuccess) {
    if (err != cudaErrorUnknown) set_error("b200lab_env_create: %s", cudaGetErrorString(err));
    if (h->derived) cudaFree(h->derived);
    if (h->dev) cudaFree(h->dev);
    free(h);
    return nullptr;
  }
  return h;
}

int b200lab_env_update(void* env, const B200LabEnvState* st) {
  B200LAB_CHECK_ARG(env && st, "null argument");
  EnvHandle* h = (EnvHandle*)env;
  h->host = *st;
  cudaError_t err = cudaMemcpy(h->dev, st, sizeof(B200LabEnvState), cudaMemcpyHostToDevice);
  if (err != cudaSuccess) { set_error("b200lab_env_update: %s", cudaGetErrorString(err)); return B200LAB_ERR_CUDA; }
  int rc = stage_const_net(h);
  if (rc) return rc;
  return upload_obs_tables(h);
}

void b200lab_env_destroy(void* env) {
  if (!env) return;
  EnvHandle* h = (EnvHandle*)env;
  if (h->dev) cudaFree(h->dev);
  if (h->side) cudaStreamDestroy(h->side);
  if (h->ev_fork) cudaEventDestroy(h->ev_fork);
  if (h->ev_join) cudaEventDestroy(h->ev_join);
  if (h->cols) cudaFree(h->cols);
  if (h->side2) { cudaStreamSynchronize(h->side2); cudaStreamDestroy(h->side2); }
  if (h->ev_spec_fork) cudaEventDestroy(h->ev_spec_fork);
  if (h->ev_spec_done) cudaEventDestroy(h->ev_spec_done);
  if (h->spec_z) cudaFree(h->spec_z);
  if (h->spec_pose) cudaFree(h->spec_pose);
  if (h->derived) cudaFree(h->derived);
  free(h);
}

int b200lab_process_action(void* env, const float* action_in, void* stream) {
  B200LAB_CHECK_ARG(env && action_in, "null argument");
  EnvHandle* h = (EnvHandle*)env;
  h->derived_root = nullptr;
  int64_t total = (int64_t)h->host.N * h->host.A;
  process_action_kernel<<<(unsigned)ceil_div(total, 256), 256, 0, (cudaStream_t)stream>>>(h->host, action_in);
  count_launch();
  return check_launch("process_action");
}

static int net_smem_bytes(const B200LabEnvState& s, size_t* out) {
  size_t smem = 0;
  int nets = 0;
  for (int g = 0; g < s.num_act_groups; ++g) {
    const B200LabActuatorGroup& a = s.act[g];
    size_t nw = 0;
    if (a.kind == B200LAB_ACT_NET_LSTM) {
      int H = a.lstm_hidden;
      B200LAB_CHECK_ARG(H == 4 || H == 8 || H == 16, "ActuatorNetLSTM hidden size must be 4, 8 or 16");
      for (int l = 0; l < a.lstm_layers; ++l) nw += 4 * H * (l == 0 ? 2 : H) + 4 * H * H + 8 * H;
      nw += H + 1;
      ++nets;
    } else if (a.kind == B200LAB_ACT_NET_MLP) {
      B200LAB_CHECK_ARG(a.mlp_layers >= 1 && a.mlp_layers <= 5, "ActuatorNetMLP supports 1..5 layers");
      for (int l = 0; l <= a.mlp_layers; ++l) B200LAB_CHECK_ARG(a.mlp_dims[l] <= 32, "ActuatorNetMLP width must be <= 32");
      B200LAB_CHECK_ARG(2 * a.mlp_num_idx == a.mlp_dims[0], "ActuatorNetMLP input dim must be 2*len(input_idx)");
      for (int l = 0; l < a.mlp_layers; ++l) nw += a.mlp_dims[l] * a.mlp_dims[l + 1] + a.mlp_dims[l + 1];
      ++nets;
    }
    if (nw * sizeof(float) > smem) smem = nw * sizeof(float);
  }
  B200LAB_CHECK_ARG(nets <= 1, "at most one network-based actuator group per articulation is supported");
  B200LAB_CHECK_ARG(smem <= kMaxNetFloats * sizeof(float), "actuator network too large for shared-memory staging");
  *out = smem;
  return B200LAB_OK;
}

int b200lab_substep_pre(void* env, const B200LabSimView* sim, void* stream) {
  B200LAB_CHECK_ARG(env && sim && sim->joint_pos && sim->joint_vel, "null argument");
  EnvHandle* h = (EnvHandle*)env;
  h->derived_root = nullptr;
  const B200LabEnvState& s = h->host;
  size_t smem = 0;
  int rc = net_smem_bytes(s, &smem);
  if (rc) return rc;
  int64_t total = (int64_t)s.N * s.J;
  launch_substep_pre(h, sim, true, false, smem, (cudaStream_t)stream);
  count_launch();
  return check_launch("substep_pre");
}

int b200lab_post_reset_write(void* env, const B200LabSimView* sim, void* stream) {
  B200LAB_CHECK_ARG(env && sim && sim->joint_pos && sim->joint_vel, "null argument");
  EnvHandle* h = (EnvHandle*)env;
  const B200LabEnvState& s = h->host;
  B200LAB_CHECK_ARG(s.any_reset_dev != nullptr, "any_reset_dev is required by post_reset_write");
  size_t smem = 0;
  int rc = net_smem_bytes(s, &smem);
  if (rc) return rc;
  int64_t total = (int64_t)s.N * s.J;
  launch_substep_pre(h, sim, false, true, smem, (cudaStream_t)stream);
  count_launch();
  return check_launch("post_reset_write");
}

int b200lab_substep_post(void* env, const B200LabSimView* sim, void* stream) {
  B200LAB_CHECK_ARG(env && sim && sim->joint_vel, "null argument");
  EnvHandle* h = (EnvHandle*)env;
  h->derived_root = nullptr;
  const B200LabEnvState& s = h->host;
  int W, epb;
  int width = s.J > s.B ? s.J : s.B;
  B200LAB_CHECK_ARG(launch_geometry(width, &W, &epb) == 0, "max(J,B) must be <= 256");
  substep_post_kernel<<<(unsigned)ceil_div(s.N, epb), kPostBlock, 0, (cudaStream_t)stream>>>(h->host, *sim, W, epb);
  count_launch();
  return check_launch("substep_post");
}

int b200lab_height_scan_prefetch(void* env, const B200LabSimView* sim, void* stream) {
  B200LAB_CHECK_ARG(env && sim && sim->root_state_w, "null argument");
  EnvHandle* h = (EnvHandle*)env;
  const B200LabEnvState& s = h->host;
  // only the height-scanner configuration (vertical pattern, yaw-only attachment) is speculated; otherwise a no-op
  if (!h->rays_vertical || !h->has_mesh || !h->spec_z || s.R <= 0 || !s.rc_drift) return B200LAB_OK;
  cudaStream_t st_ = (cudaStream_t)stream;
  cudaError_t err = cudaEventRecord(h->ev_spec_fork, st_);
  if (err == cudaSuccess) err = cudaStreamWaitEvent(h->side2, h->ev_spec_fork, 0);
  if (err != cudaSuccess) { set_error("b200lab_height_scan_prefetch: %s", cudaGetErrorString(err)); return B200LAB_ERR_CUDA; }
  int threads = ((s.R + 31) / 32) * 32;
  if (threads > 256) threads = 256;
  // the ray columns of the observation program carry the local ray starts; a sensor that is not observed has none
  const ObsCol* ray_cols = nullptr;
  int k = 0;
  for (int t = 0; t < s.num_obs && !ray_cols; ++t) {
    if (s.obs[t].op == B200LAB_OBS_HEIGHT_SCAN && s.obs[t].dim == s.R) ray_cols = h->cols + k;
    k += s.obs[t].dim;
  }
  if (!ray_cols) return B200LAB_OK;
  height_scan_spec_kernel<<<(unsigned)s.N, threads, 0, h->side2>>>(h->host, *sim, ray_cols, h->mesh, h->spec_z, h->spec_pose);
  count_launch();
  int rc = check_launch("height_scan_prefetch");
  if (rc) return rc;
  err = cudaEventRecord(h->ev_spec_done, h->side2);
  if (err != cudaSuccess) { set_error("b200lab_height_scan_prefetch: %s", cudaGetErrorString(err)); return B200LAB_ERR_CUDA; }
  h->spec_pending = true;
  return B200LAB_OK;
}

int b200lab_decimation_fused(void* env, const float* action_in, const B200LabSimView* prev,
                             const B200LabSimView* views, int32_t num_substeps, void* stream) {
  B200LAB_CHECK_ARG(env && action_in && prev && views, "null argument");
  B200LAB_CHECK_ARG(num_substeps >= 1 && num_substeps <= kMaxDecimation, "decimation must be in 1..8");
  EnvHandle* h = (EnvHandle*)env;
  h->derived_root = nullptr;
  const B200LabEnvState& s = h->host;
  size_t smem = 0;
  int rc = net_smem_bytes(s, &smem);
  if (rc) return rc;
  SimViews sv;
  sv.prev = *prev;
  sv.S = num_substeps;
  for (int k = 0; k < num_substeps; ++k) {
    B200LAB_CHECK_ARG(views[k].joint_pos && views[k].joint_vel, "null simulator view");
    sv.v[k] = views[k];
  }
  for (int k = num_substeps; k < kMaxDecimation; ++k) sv.v[k] = views[num_substeps - 1];
  int W, epb;
  B200LAB_CHECK_ARG(s.B <= kFusedBlock, "B must be <= 128");
  W = s.B > 0 ? s.B : 1;
  epb = kFusedBlock / W;
  int joint_blocks = (int)ceil_div((int64_t)s.N * s.J, kFusedBlock);
  int body_blocks = (int)ceil_div(s.N, epb);
  if (h->lstm_const) {
    decimation_fused_kernel<NetLstm82><<<(unsigned)(joint_blocks + body_blocks), kFusedBlock, 0, (cudaStream_t)stream>>>(
        h->host, action_in, sv, joint_blocks, W, epb, *reinterpret_cast<const NetLstm82*>(h->lstm_w));
  } else if (h->mlp_const) {
    decimation_fused_kernel<NetMlp63><<<(unsigned)(joint_blocks + body_blocks), kFusedBlock, mlp63_smem(kFusedBlock), (cudaStream_t)stream>>>(
        h->host, action_in, sv, joint_blocks, W, epb, NetMlp63{});
  } else {
    decimation_fused_kernel<NetNone><<<(unsigned)(joint_blocks + body_blocks), kFusedBlock, smem, (cudaStream_t)stream>>>(
        h->host, action_in, sv, joint_blocks, W, epb, NetNone{});
  }
  count_launch();
  return check_launch("decimation_fused");
}

int b200lab_post_step(void* env, const B200LabSimView* sim, int64_t step_counter, void* stream) {
  B200LAB_CHECK_ARG(env && sim && sim->root_state_w && sim->joint_pos && sim->joint_vel, "null argument");
  EnvHandle* h = (EnvHandle*)env;
  const B200LabEnvState& s = h->host;
  B200LAB_CHECK_ARG(s.J <= 64 && s.B <= 64, "joint/body masks support up to 64 joints/bodies");
  B200LAB_CHECK_ARG(s.n_u_reset <= kDrawCapReset && s.n_u_step <= kDrawCapStep, "more random draws per env than the draw tables hold");
  const size_t tab_bytes = (size_t)(s.num_term + s.num_rew) * sizeof(B200LabTerm);
  // a pending speculative scan is validated by the stand-alone derive kernel (it must be joined first)
  const bool fuse_derive = !h->spec_pending && s.obs_dim > 0;
  {
    // measured on B200 at 4096 envs (profiles/r2_notes.md): combined kernel 36.9 us, hot + cold 43.5 us -> default combined
    static const int split = [] { const char* v = getenv("B200LAB_POST_SPLIT"); return v ? atoi(v) : 0; }();
    const dim3 grid((unsigned)ceil_div(s.N, kStepWarps)), block(kStepWarps * 32);
    float* dptr = fuse_derive ? h->derived : (float*)nullptr;
    cudaError_t err;
    if (split) {
      err = launch_pdl(post_step_kernel<1>, grid, block, tab_bytes, (cudaStream_t)stream, h->host, *sim, step_counter,
                       (const Env*)h->dev, dptr);
      if (err == cudaSuccess) {   // full dependency on the hot kernel (it reads reset_buf); compute_obs may overlap ITS tail
        post_step_kernel<2><<<grid, block, 0, (cudaStream_t)stream>>>(h->host, *sim, step_counter, (const Env*)h->dev, dptr);
        err = cudaGetLastError();
        count_launch();
      }
    } else {
      err = launch_pdl(post_step_kernel<0>, grid, block, tab_bytes, (cudaStream_t)stream, h->host, *sim, step_counter,
                       (const Env*)h->dev, dptr);
    }
    if (err != cudaSuccess) { set_error("post_step: %s", cudaGetErrorString(err)); return B200LAB_ERR_CUDA; }
  }
  h->derived_root = fuse_derive ? sim->root_state_w : nullptr;
  count_launch();
  return check_launch("post_step");
}

int b200lab_compute_obs_range(void* env, const B200LabSimView* sim, int64_t step_counter, int32_t env_begin,
                              int32_t env_end, void* stream) {
  B200LAB_CHECK_ARG(env && sim && sim->root_state_w, "null argument");
  EnvHandle* h = (EnvHandle*)env;
  const B200LabEnvState& s = h->host;
  B200LAB_CHECK_ARG(env_begin >= 0 && env_begin <= env_end && env_end <= s.N, "env range out of bounds");
  cudaStream_t st_ = (cudaStream_t)stream;
  bool joined_spec = false;
  if (h->spec_pending) {   // join the speculative scan
    cudaError_t err = cudaStreamWaitEvent(st_, h->ev_spec_done, 0);
    if (err != cudaSuccess) { set_error("b200lab_compute_obs: %s", cudaGetErrorString(err)); return B200LAB_ERR_CUDA; }
    h->spec_pending = false;
    joined_spec = true;
  }
  const int cnt = env_end - env_begin;
  if (cnt == 0) return B200LAB_OK;
  if (!(h->derived_root && h->derived_root == sim->root_state_w)) {
    derive_obs_kernel<<<(unsigned)ceil_div(cnt, 128), 128, 0, st_>>>(h->host, *sim, h->derived,
                                                                   h->rays_vertical && h->has_mesh ? h->spec_pose : nullptr,
                                                                   env_begin, env_end);
    count_launch();
  }
  int64_t total = (int64_t)cnt * s.obs_dim;
  if (total <= 0) return check_launch("compute_obs");
  if (h->obs_impl == 1 && h->obs_plain && h->rays_vertical && h->has_mesh && !joined_spec && s.obs_dim <= kPatchMaxCols * kPatchThreads) {
    // staged height scan: one CTA per env (a speculated scan, when one was joined, is consumed by the per-ray kernel)
    static bool attr_set = false;
    if (!attr_set) {
      cudaError_t aerr = cudaFuncSetAttribute(compute_obs_patch_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              (int)sizeof(PatchSmem));
      if (aerr == cudaSuccess)
        aerr = cudaFuncSetAttribute(compute_obs_patch_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(PatchSmem));
      if (aerr != cudaSuccess) { set_error("compute_obs: %s", cudaGetErrorString(aerr)); return B200LAB_ERR_CUDA; }
      attr_set = true;
    }
    auto pk = s.obs_dim <= kPatchThreads ? compute_obs_patch_kernel<1> : compute_obs_patch_kernel<2>;
    cudaError_t err = launch_pdl(pk, dim3((unsigned)cnt), dim3(kPatchThreads), sizeof(PatchSmem), st_,
                                 h->host, *sim, step_counter, (const ObsCol*)h->cols, (const float*)h->derived, h->mesh,
                                 (unsigned)env_begin);
    if (err != cudaSuccess) { set_error("compute_obs: %s", cudaGetErrorString(err)); return B200LAB_ERR_CUDA; }
    count_launch();
    return check_launch("compute_obs");
  }
  if (h->obs_impl >= 2 && h->obs_plain && h->rays_vertical) {
    // Q columns per thread; obs_impl 2 -> Q = 4, 3 -> Q = 2
    const int Q = h->obs_impl == 3 ? 2 : 4;
    auto kq = Q == 4 ? compute_obs_quad_kernel<4> : compute_obs_quad_kernel<2>;
    const int64_t threads = (int64_t)cnt * ((s.obs_dim + Q - 1) / Q);
    cudaError_t err = launch_pdl(kq, dim3((unsigned)ceil_div(threads, 64)), dim3(64), 0, st_, h->host, *sim, step_counter,
                                 (const ObsCol*)h->cols, (const float*)h->derived, h->mesh, h->has_mesh ? 1 : 0,
                                 (const float*)h->spec_z, (unsigned)env_begin, (unsigned)cnt);
    if (err != cudaSuccess) { set_error("compute_obs: %s", cudaGetErrorString(err)); return B200LAB_ERR_CUDA; }
    count_launch();
    return check_launch("compute_obs");
  }
  {
    auto kern = !h->obs_plain ? (h->rays_vertical ? compute_obs_kernel<true, kVertBatch, false> : compute_obs_kernel<false, kVertBatch, false>)
              : !h->rays_vertical ? compute_obs_kernel<false> : h->vert_batch == 4 ? compute_obs_kernel<true, 4>
              : h->vert_batch == 3 ? compute_obs_kernel<true, 3> : compute_obs_kernel<true>;
    cudaError_t err = launch_pdl(kern, dim3((unsigned)ceil_div(total, B200LAB_OBS_THREADS)), dim3(B200LAB_OBS_THREADS), 0, st_, h->host, *sim, step_counter,
                                 (const ObsCol*)h->cols, (const float*)h->derived, h->mesh, h->has_mesh ? 1 : 0,
                                 (const float*)h->spec_z, (unsigned)env_begin, (unsigned)cnt);
    if (err != cudaSuccess) { set_error("compute_obs: %s", cudaGetErrorString(err)); return B200LAB_ERR_CUDA; }
  }
  count_launch();
  return check_launch("compute_obs");
}

int b200lab_log_fold(void* env, void* stream) {
  B200LAB_CHECK_ARG(env, "null argument");
  EnvHandle* h = (EnvHandle*)env;
  if (!h->host.log_accum || !h->host.log_running) return B200LAB_OK;
  B200LAB_CHECK_ARG(4 + h->host.num_rew + h->host.num_term <= B200LAB_LOG_SLOTS, "too many log slots");
  log_fold_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(h->host);
  count_launch();
  return check_launch("log_fold");
}

int b200lab_compute_obs(void* env, const B200LabSimView* sim, int64_t step_counter, void* stream) {
  B200LAB_CHECK_ARG(env, "null argument");
  return b200lab_compute_obs_range(env, sim, step_counter, 0, ((EnvHandle*)env)->host.N, stream);
}

int b200lab_step_tail_begin(void* env, const B200LabSimView* sim, void* stream) {
  B200LAB_CHECK_ARG(env && sim && sim->root_state_w && sim->joint_pos && sim->joint_vel, "null argument");
  EnvHandle* h = (EnvHandle*)env;
  cudaStream_t st_ = (cudaStream_t)stream;
  // fork: the actuator pass and the observation kernels touch disjoint buffers (none of the library's observation terms
  // reads actuator outputs) and are both latency-bound, so they share the SMs
  cudaError_t err = cudaEventRecord(h->ev_fork, st_);
  if (err == cudaSuccess) err = cudaStreamWaitEvent(h->side, h->ev_fork, 0);
  if (err != cudaSuccess) { set_error("b200lab_step_tail: %s", cudaGetErrorString(err)); return B200LAB_ERR_CUDA; }
  int rc = b200lab_post_reset_write(env, sim, h->side);
  if (rc) return rc;
  rc = b200lab_log_fold(env, h->side);
  if (rc) return rc;
  err = cudaEventRecord(h->ev_join, h->side);
  if (err != cudaSuccess) { set_error("b200lab_step_tail: %s", cudaGetErrorString(err)); return B200LAB_ERR_CUDA; }
  return B200LAB_OK;
}

int b200lab_step_tail_end(void* env, void* stream) {
  B200LAB_CHECK_ARG(env, "null argument");
  EnvHandle* h = (EnvHandle*)env;
  cudaError_t err = cudaStreamWaitEvent((cudaStream_t)stream, h->ev_join, 0);
  if (err != cudaSuccess) { set_error("b200lab_step_tail: %s", cudaGetErrorString(err)); return B200LAB_ERR_CUDA; }
  return B200LAB_OK;
}

int b200lab_step_tail(void* env, const B200LabSimView* sim, int64_t step_counter, void* stream) {
  int rc = b200lab_step_tail_begin(env, sim, stream);
  if (rc) return rc;
  rc = b200lab_compute_obs(env, sim, step_counter, stream);
  if (rc) return rc;
  return b200lab_step_tail_end(env, stream);
}
}
