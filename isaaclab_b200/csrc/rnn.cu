// rnn.cu -- pointwise halves of the recurrent policy's cells (RSL/networks/memory.py:21-24 builds nn.LSTM / nn.GRU;
// RSL/modules/actor_critic_recurrent.py:141-160 runs them in inference mode over one step and in batch mode over padded
// trajectories).  The two matrix products of a step -- x W_ih^T + b_ih for ALL time steps at once, h W_hh^T + b_hh per
// step -- go through b200lab_linear_fwd (tcgen05 for the wide shapes); these kernels do the gate arithmetic of
// torch.nn.LSTM / torch.nn.GRU (gate order i,f,g,o and r,z,n) and its exact derivative for back-propagation through time,
// one thread per (trajectory, hidden unit), every access coalesced along the hidden dimension.
#include "common.cuh"

namespace b200lab {

__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }

__global__ void __launch_bounds__(256)
lstm_cell_fwd_kernel(const float* __restrict__ gi, const float* __restrict__ gh, const float* __restrict__ c_prev,
                     float* __restrict__ h, float* __restrict__ c, float* __restrict__ gates, int64_t B, int H) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * H) return;
  const int64_t b = idx / H;
  const int j = (int)(idx - b * H);
  const int64_t g0 = b * 4 * H + j;
  const float i_ = sigmoidf_(gi[g0] + gh[g0]);
  const float f_ = sigmoidf_(gi[g0 + H] + gh[g0 + H]);
  const float g_ = tanhf(gi[g0 + 2 * H] + gh[g0 + 2 * H]);
  const float o_ = sigmoidf_(gi[g0 + 3 * H] + gh[g0 + 3 * H]);
  const float cn = f_ * c_prev[idx] + i_ * g_;
  c[idx] = cn;
  h[idx] = o_ * tanhf(cn);
  if (gates) { gates[g0] = i_; gates[g0 + H] = f_; gates[g0 + 2 * H] = g_; gates[g0 + 3 * H] = o_; }
}

// dc: in = dLoss/dc_t carried from step t+1, out = dLoss/dc_{t-1}.  dgates = dLoss/d(pre-activation gates) of step t.
__global__ void __launch_bounds__(256)
lstm_cell_bwd_kernel(const float* __restrict__ gates, const float* __restrict__ c_prev, const float* __restrict__ c,
                     const float* __restrict__ dh_out, const float* __restrict__ dh_rec, float* __restrict__ dc,
                     float* __restrict__ dgates, int64_t B, int H) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * H) return;
  const int64_t b = idx / H;
  const int j = (int)(idx - b * H);
  const int64_t g0 = b * 4 * H + j;
  const float i_ = gates[g0], f_ = gates[g0 + H], g_ = gates[g0 + 2 * H], o_ = gates[g0 + 3 * H];
  const float dh = (dh_out ? dh_out[idx] : 0.f) + (dh_rec ? dh_rec[idx] : 0.f);
  const float tc = tanhf(c[idx]);
  const float dct = dc[idx] + dh * o_ * (1.0f - tc * tc);
  dgates[g0] = dct * g_ * i_ * (1.0f - i_);
  dgates[g0 + H] = dct * c_prev[idx] * f_ * (1.0f - f_);
  dgates[g0 + 2 * H] = dct * i_ * (1.0f - g_ * g_);
  dgates[g0 + 3 * H] = dh * tc * o_ * (1.0f - o_);
  dc[idx] = dct * f_;
}

// saved [B, 4H] = r | z | n | (W_hn h + b_hn)
__global__ void __launch_bounds__(256)
gru_cell_fwd_kernel(const float* __restrict__ gi, const float* __restrict__ gh, const float* __restrict__ h_prev,
                    float* __restrict__ h, float* __restrict__ saved, int64_t B, int H) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * H) return;
  const int64_t b = idx / H;
  const int j = (int)(idx - b * H);
  const int64_t g0 = b * 3 * H + j;
  const float r = sigmoidf_(gi[g0] + gh[g0]);
  const float z = sigmoidf_(gi[g0 + H] + gh[g0 + H]);
  const float hn = gh[g0 + 2 * H];
  const float n = tanhf(gi[g0 + 2 * H] + r * hn);
  h[idx] = (1.0f - z) * n + z * h_prev[idx];
  if (saved) {
    const int64_t s0 = b * 4 * H + j;
    saved[s0] = r; saved[s0 + H] = z; saved[s0 + 2 * H] = n; saved[s0 + 3 * H] = hn;
  }
}

// dh_direct = dLoss/dh_{t-1} through the (z * h_{t-1}) term; the W_hh path is added by the caller (dgh . W_hh).
__global__ void __launch_bounds__(256)
gru_cell_bwd_kernel(const float* __restrict__ saved, const float* __restrict__ h_prev, const float* __restrict__ dh_out,
                    const float* __restrict__ dh_rec, const float* dh_dir_in, float* dh_dir_out,   // may be the same buffer
                    float* __restrict__ dgi, float* __restrict__ dgh, int64_t B, int H) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * H) return;
  const int64_t b = idx / H;
  const int j = (int)(idx - b * H);
  const int64_t s0 = b * 4 * H + j, g0 = b * 3 * H + j;
  const float r = saved[s0], z = saved[s0 + H], n = saved[s0 + 2 * H], hn = saved[s0 + 3 * H];
  const float dh = (dh_out ? dh_out[idx] : 0.f) + (dh_rec ? dh_rec[idx] : 0.f) + (dh_dir_in ? dh_dir_in[idx] : 0.f);
  const float dn_pre = dh * (1.0f - z) * (1.0f - n * n);
  const float dz_pre = dh * (h_prev[idx] - n) * z * (1.0f - z);
  const float dr_pre = dn_pre * hn * r * (1.0f - r);
  dgi[g0] = dr_pre; dgi[g0 + H] = dz_pre; dgi[g0 + 2 * H] = dn_pre;
  dgh[g0] = dr_pre; dgh[g0 + H] = dz_pre; dgh[g0 + 2 * H] = dn_pre * r;
  dh_dir_out[idx] = dh * z;
}

}  // namespace b200lab

using namespace b200lab;

extern "C" {

int b200lab_lstm_cell_fwd(const float* gi, const float* gh, const float* c_prev, float* h, float* c, float* gates,
                          int64_t B, int32_t H, void* stream) {
  B200LAB_CHECK_ARG(gi && gh && c_prev && h && c && B > 0 && H > 0, "bad argument");
  lstm_cell_fwd_kernel<<<(unsigned)ceil_div(B * H, 256), 256, 0, (cudaStream_t)stream>>>(gi, gh, c_prev, h, c, gates, B, H);
  count_launch();
  return check_launch("lstm_cell_fwd");
}

int b200lab_lstm_cell_bwd(const float* gates, const float* c_prev, const float* c, const float* dh_out,
                          const float* dh_rec, float* dc, float* dgates, int64_t B, int32_t H, void* stream) {
  B200LAB_CHECK_ARG(gates && c_prev && c && dc && dgates && B > 0 && H > 0, "bad argument");
  lstm_cell_bwd_kernel<<<(unsigned)ceil_div(B * H, 256), 256, 0, (cudaStream_t)stream>>>(gates, c_prev, c, dh_out, dh_rec,
                                                                                        dc, dgates, B, H);
  count_launch();
  return check_launch("lstm_cell_bwd");
}

int b200lab_gru_cell_fwd(const float* gi, const float* gh, const float* h_prev, float* h, float* saved, int64_t B,
                         int32_t H, void* stream) {
  B200LAB_CHECK_ARG(gi && gh && h_prev && h && B > 0 && H > 0, "bad argument");
  gru_cell_fwd_kernel<<<(unsigned)ceil_div(B * H, 256), 256, 0, (cudaStream_t)stream>>>(gi, gh, h_prev, h, saved, B, H);
  count_launch();
  return check_launch("gru_cell_fwd");
}

int b200lab_gru_cell_bwd(const float* saved, const float* h_prev, const float* dh_out, const float* dh_rec,
                         const float* dh_dir_in, float* dh_dir_out, float* dgi, float* dgh, int64_t B, int32_t H,
                         void* stream) {
  B200LAB_CHECK_ARG(saved && h_prev && dh_dir_out && dgi && dgh && B > 0 && H > 0, "bad argument");
  gru_cell_bwd_kernel<<<(unsigned)ceil_div(B * H, 256), 256, 0, (cudaStream_t)stream>>>(saved, h_prev, dh_out, dh_rec,
                                                                                       dh_dir_in, dh_dir_out, dgi, dgh, B, H);
  count_launch();
  return check_launch("gru_cell_bwd");
}

}  // extern "C"
