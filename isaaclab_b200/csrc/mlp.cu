// mlp.cu -- actor-critic MLP layers (RSL/networks/mlp.py:60-80: Linear+ELU ... Linear), fp32 SIMT path.
//
// precision 0: fp32 FFMA register-tiled GEMM (the exact-order parity path; the reference's own CPU path is fp32).
// precision 1/2: tcgen05 kind::tf32 paths live in mlp_tc.cu (dispatch below).
//
// Three GEMM shapes per layer (row-major everywhere, nn.Linear layout W[Nout,K]):
//   fwd : Y[M,Nout]  = act(X[M,K] . W^T + b)
//   dX  : dX[M,K]    = dZ[M,Nout] . W[Nout,K]
//   dW  : dW[Nout,K] = dZ^T . X   (reduction over M: split across CTAs, deterministic two-pass reduce)
// Tiles are staged through shared memory with 128-bit loads when the contiguous dimension allows it;
// M is large (4096 .. 98304 rows), K/Nout <= 512, so grids are M/64 x Nout/64 CTAs -- several waves of 148 SMs.
#include "common.cuh"

namespace b200lab {

// tcgen05 path (mlp_tc.cu): same operand contract as sgemm_kernel below
int gemm_tc(const float* A, int64_t lda_i, int64_t lda_k, const float* B, int64_t ldb_k, int64_t ldb_j, float* C, int64_t ldc,
            int64_t M, int N, int64_t K, int64_t k_per_split, int splits, const float* bias, int epi, int precision,
            cudaStream_t s);
// a layer goes to the tensor path when its input OR output is at least this wide: the 12-wide action head and the 1-wide
// value head ride on zero-filled tiles (masked copies) -- mostly padding, but their cost is the [M,128] activation read,
// and the FFMA kernel's 64 x 64 tiles were slower on them; layers narrow on both sides (cartpole's 4 -> 32) stay FFMA
constexpr int kMinTcWidth = 32;

constexpr int BM = 64, BN = 64, BK = 16;
constexpr int kGemmThreads = 256;

// Generic strided tile GEMM: C[i,j] = sum_k A(i,k) * B(k,j), A(i,k) = A[i*lda_i + k*lda_k], B(k,j) = B[k*ldb_k + j*ldb_j].
// grid.z splits the k range (used for the dW reduction over M); each z-slice writes its own C slab.
// EPI: 0 = store, 1 = +bias, 2 = +bias then ELU.
template <int EPI>
__global__ void __launch_bounds__(kGemmThreads)
sgemm_kernel(const float* __restrict__ A, int64_t lda_i, int64_t lda_k, const float* __restrict__ B, int64_t ldb_k,
             int64_t ldb_j, float* __restrict__ C, int64_t ldc, int64_t M, int N, int64_t K, int64_t k_per_split,
             const float* __restrict__ bias) {
  __shared__ float sA[BK][BM + 4];
  __shared__ float sB[BK][BN + 4];
  const int tid = threadIdx.x;
  const int tx = tid % 16, ty = tid / 16;  // 16 x 16 threads, 4 x 4 outputs each
  const int64_t i0 = (int64_t)blockIdx.x * BM;
  const int j0 = blockIdx.y * BN;
  const int64_t k_begin = (int64_t)blockIdx.z * k_per_split;
  const int64_t k_end = min64(K, k_begin + k_per_split);
  float acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b] = 0.0f;

  // loader mapping: choose the thread->element map so that the unit-stride dimension is the fastest
  const bool a_k_contig = (lda_k == 1);
  const bool b_k_contig = (ldb_k == 1);
  for (int64_t k0 = k_begin; k0 < k_end; k0 += BK) {
#pragma unroll
    for (int r = 0; r < (BM * BK) / kGemmThreads; ++r) {
      int e = tid + r * kGemmThreads;
      int kk, ii;
      if (a_k_contig) { kk = e % BK; ii = e / BK; } else { ii = e % BM; kk = e / BM; }
      int64_t gi = i0 + ii, gk = k0 + kk;
      sA[kk][ii] = (gi < M && gk < k_end) ? __ldg(A + gi * lda_i + gk * lda_k) : 0.0f;
    }
#pragma unroll
    for (int r = 0; r < (BN * BK) / kGemmThreads; ++r) {
      int e = tid + r * kGemmThreads;
      int kk, jj;
      if (b_k_contig) { kk = e % BK; jj = e / BK; } else { jj = e % BN; kk = e / BN; }
      int64_t gk = k0 + kk;
      int gj = j0 + jj;
      sB[kk][jj] = (gj < N && gk < k_end) ? __ldg(B + gk * ldb_k + (int64_t)gj * ldb_j) : 0.0f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float4 a4 = *reinterpret_cast<const float4*>(&sA[kk][ty * 4]);
      float4 b4 = *reinterpret_cast<const float4*>(&sB[kk][tx * 4]);
      float av[4] = {a4.x, a4.y, a4.z, a4.w}, bv[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = fmaf(av[a], bv[b], acc[a][b]);
    }
    __syncthreads();
  }
  float* Cz = C + (int64_t)blockIdx.z * M * ldc;
#pragma unroll
  for (int a = 0; a < 4; ++a) {
    int64_t gi = i0 + ty * 4 + a;
    if (gi >= M) continue;
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      int gj = j0 + tx * 4 + b;
      if (gj >= N) continue;
      float v = acc[a][b];
      if (EPI >= 1) v += __ldg(bias + gj);
      if (EPI == 2) v = elu_f(v);
      Cz[gi * ldc + gj] = v;
    }
  }
}

// ELU'(pre-activation) is recovered from the saved post-activation Y: Y > 0 -> 1, else Y + 1 (alpha = 1).

// column sums of dZ[M,N] (bias gradient), partials per row-slab then a fixed-order reduce.
__global__ void __launch_bounds__(256)
colsum_partial_kernel(const float* __restrict__ dZ, int64_t M, int N, int64_t rows_per_block, float* __restrict__ part) {
  // block handles rows [r0, r1), threads stride columns; 256 threads = 8 row-lanes x 32 col-lanes
  __shared__ float sh[8][33];
  int cl = threadIdx.x % 32, rl = threadIdx.x / 32;
  int64_t r0 = (int64_t)blockIdx.x * rows_per_block, r1 = min64(M, r0 + rows_per_block);
  for (int c0 = 0; c0 < N; c0 += 32) {
    int c = c0 + cl;
    float s = 0.0f;
    if (c < N)
      for (int64_t r = r0 + rl; r < r1; r += 8) s += __ldg(dZ + r * N + c);
    sh[rl][cl] = s;
    __syncthreads();
    if (rl == 0 && c < N) {
      float t = 0.0f;
#pragma unroll
      for (int k = 0; k < 8; ++k) t += sh[k][cl];
      part[(int64_t)blockIdx.x * N + c] = t;
    }
    __syncthreads();
  }
}

// dZ = dY * ELU'(.) AND the per-slab column sums of dZ in one pass over the layer's [M,N] activations (the bias gradient
// used to re-read dZ in a second kernel).  A block owns `rows_per_block` rows; warp w takes rows w, w+8, ...; lanes take 4
// consecutive columns (128-bit loads/stores, 512 B per warp instruction, 8 rows in flight); columns in chunks of 128.
// VEC = false: scalar columns for N % 4 != 0.
template <bool VEC>
__global__ void __launch_bounds__(256)
elu_bwd_colsum_kernel(const float* __restrict__ Y, const float* dY, float* dZ, int64_t M, int N,   // dZ may be dY (in place)
                      int64_t rows_per_block, float* __restrict__ part) {
  constexpr int kCols = VEC ? 128 : 32;
  __shared__ float sh[8][kCols + 4];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_block, r1 = min64(M, r0 + rows_per_block);
  for (int c0 = 0; c0 < N; c0 += kCols) {
    if (VEC) {
      const int c = c0 + lane * 4;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      if (c < N) {
#pragma unroll 8
        for (int64_t r = r0 + warp; r < r1; r += 8) {
          const float4 y = __ldg(reinterpret_cast<const float4*>(Y + r * N + c));
          float4 g = *reinterpret_cast<const float4*>(dY + r * N + c);
          g.x *= y.x > 0.0f ? 1.0f : y.x + 1.0f;
          g.y *= y.y > 0.0f ? 1.0f : y.y + 1.0f;
          g.z *= y.z > 0.0f ? 1.0f : y.z + 1.0f;
          g.w *= y.w > 0.0f ? 1.0f : y.w + 1.0f;
          *reinterpret_cast<float4*>(dZ + r * N + c) = g;
          acc.x += g.x; acc.y += g.y; acc.z += g.z; acc.w += g.w;
        }
      }
      sh[warp][lane * 4 + 0] = acc.x; sh[warp][lane * 4 + 1] = acc.y;
      sh[warp][lane * 4 + 2] = acc.z; sh[warp][lane * 4 + 3] = acc.w;
    } else {
      const int c = c0 + lane;
      float acc = 0.0f;
      if (c < N)
        for (int64_t r = r0 + warp; r < r1; r += 8) {
          const float y = __ldg(Y + r * N + c);
          const float g = dY[r * N + c] * (y > 0.0f ? 1.0f : y + 1.0f);
          dZ[r * N + c] = g;
          acc += g;
        }
      sh[warp][lane] = acc;
    }
    __syncthreads();
    if (threadIdx.x < kCols && c0 + (int)threadIdx.x < N) {
      float t = 0.0f;
#pragma unroll
      for (int k = 0; k < 8; ++k) t += sh[k][threadIdx.x];
      part[(int64_t)blockIdx.x * N + c0 + threadIdx.x] = t;
    }
    __syncthreads();
  }
}

// rows per block of the bias-gradient column sum: 64 rows -> M/64 blocks (384 at M = 24576: 2.6 waves of 148 SMs; at 512
// rows per block the grid was 48 blocks and this memory-bound pass took longer than the GEMMs)
constexpr int kColsumRows = 64;

// K-splits of the dW = dZ^T X reduction over the M samples.  SIMT path: ~4 CTAs per SM.  Tensor-core path (128 x 128|64
// tiles, one CTA per SM): as many splits as still fit ONE wave -- 152 CTAs on 148 SMs would run as two.
static int dw_splits(int64_t M, int K, int Nout, bool use_tc) {
  int64_t tiles = use_tc ? ceil_div(Nout, 128) * ceil_div(K, K > 64 ? 128 : 64) : ceil_div(Nout, BM) * ceil_div(K, BN);
  int64_t want = use_tc ? kNumSMs / tiles : ceil_div(4 * kNumSMs, tiles);
  int64_t max_by_m = ceil_div(M, 256);
  int64_t s = want < max_by_m ? want : max_by_m;
  return (int)(s < 1 ? 1 : s);
}

// The tail of a layer's backward in ONE launch: dW = sum of the K-split slabs, and up to two bias gradients (column sums
// of row-block partials).  Blocks [0, slab_blocks) take the slabs, the rest one warp per column (fixed order: deterministic).
struct ColTask {
  const float* part;
  int rows, n;
  float* out;
};
__global__ void __launch_bounds__(256)
reduce_tail_kernel(const float* __restrict__ slabs, int splits, int64_t n, float* __restrict__ out, int slab_blocks, ColTask c0,
                   ColTask c1) {
  if ((int)blockIdx.x < slab_blocks) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float s = 0.0f;
    for (int z = 0; z < splits; ++z) s += __ldg(slabs + (int64_t)z * n + i);
    out[i] = s;
    return;
  }
  int b = (int)blockIdx.x - slab_blocks;
  const int c0_blocks = (c0.n + 7) / 8;
  const ColTask t = b < c0_blocks ? c0 : c1;
  if (b >= c0_blocks) b -= c0_blocks;
  const int lane = threadIdx.x & 31, c = b * 8 + (threadIdx.x >> 5);
  if (c >= t.n) return;
  float s = 0.0f;
  for (int z = lane; z < t.rows; z += 32) s += __ldg(t.part + (int64_t)z * t.n + c);
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) s += __shfl_down_sync(0xffffffffu, s, d);
  if (lane == 0) t.out[c] = s;
}

static int linear_bwd_impl(const float* X, const float* W, const float* Y, const float* dY, float* dZ, float* dX, float* dW,
                           float* db, float* db_in, int64_t M, int32_t K, int32_t Nout, int32_t flags, int32_t precision,
                           void* workspace, cudaStream_t s) {
  const bool elu = flags & B200LAB_BWD_ELU_OUT, elu_in = flags & B200LAB_BWD_ELU_IN, dz_given = flags & B200LAB_BWD_DZ_GIVEN;
  B200LAB_CHECK_ARG(X && W && dY && dW && db && workspace && M > 0 && K > 0 && Nout > 0, "bad argument");
  B200LAB_CHECK_ARG(!elu || (Y && dZ), "ELU backward needs the saved output Y and the dZ scratch");
  B200LAB_CHECK_ARG(precision >= 0 && precision <= 2, "precision must be 0 (fp32 FFMA), 1 (tf32) or 2 (tf32x3)");
  const bool use_tc = precision != 0 && (Nout >= kMinTcWidth || K >= kMinTcWidth);
  const float* dz = dY;
  float* ws = (float*)workspace;
  float* colsum_part = ws + (int64_t)dw_splits(M, K, Nout, use_tc) * Nout * K;
  float* in_part = colsum_part + ceil_div(M, kColsumRows) * Nout;
  const int colsum_blocks = (int)ceil_div(M, kColsumRows);
  ColTask own{nullptr, 0, 0, nullptr}, in{nullptr, 0, 0, nullptr};
  if (!dz_given) {
    if (elu) {   // dZ and the bias-gradient partial sums in one pass
      const bool vec = Nout % 4 == 0 && ((uintptr_t)Y | (uintptr_t)dY | (uintptr_t)dZ) % 16 == 0;
      if (vec) elu_bwd_colsum_kernel<true><<<colsum_blocks, 256, 0, s>>>(Y, dY, dZ, M, Nout, kColsumRows, colsum_part);
      else     elu_bwd_colsum_kernel<false><<<colsum_blocks, 256, 0, s>>>(Y, dY, dZ, M, Nout, kColsumRows, colsum_part);
      count_launch();
      dz = dZ;
    } else {
      colsum_partial_kernel<<<colsum_blocks, 256, 0, s>>>(dz, M, Nout, kColsumRows, colsum_part);
      count_launch();
    }
    own = ColTask{colsum_part, colsum_blocks, Nout, db};
  }
  if (dX) {  // dX[M,K] = dZ[M,Nout] . W[Nout,K]
    if (use_tc) {
      int rc = gemm_tc(dz, Nout, 1, W, K, 1, dX, K, M, K, Nout, Nout, 1, nullptr, 0, precision, s);
      if (rc) return rc;
    } else {
      dim3 grid((unsigned)ceil_div(M, BM), (unsigned)ceil_div(K, BN), 1);
      sgemm_kernel<0><<<grid, kGemmThreads, 0, s>>>(dz, Nout, 1, W, K, 1, dX, K, M, K, Nout, Nout, nullptr);
      count_launch();
    }
    if (elu_in) {
      // X is the ELU output this gradient belongs to: dX becomes the input-side layer's dZ, in place, with the partial sums
      // of its bias gradient.  (Doing this in the epilogue of the dX product was measured: four epilogue warps per SM cannot
      // keep enough loads of X in flight, the product ran 3x longer -- profiles/r2_notes.md.)
      const bool vec = K % 4 == 0 && ((uintptr_t)X | (uintptr_t)dX) % 16 == 0;
      if (vec) elu_bwd_colsum_kernel<true><<<colsum_blocks, 256, 0, s>>>(X, dX, dX, M, K, kColsumRows, in_part);
      else     elu_bwd_colsum_kernel<false><<<colsum_blocks, 256, 0, s>>>(X, dX, dX, M, K, kColsumRows, in_part);
      count_launch();
      in = ColTask{in_part, colsum_blocks, K, db_in};
    }
  }
  // dW[Nout,K] = sum_m dZ[m,n] X[m,k]: A(i=n,k=m) = dZ[m*Nout + n], B(k=m,j=k) = X[m*K + k]
  int splits = dw_splits(M, K, Nout, use_tc);
  int64_t k_per = ceil_div(ceil_div(M, splits), 32) * 32;   // multiple of both paths' K tile
  splits = (int)ceil_div(M, k_per);
  float* dst = splits == 1 ? dW : ws;
  if (use_tc) {
    int rc = gemm_tc(dz, 1, Nout, X, K, 1, dst, K, Nout, K, M, k_per, splits, nullptr, 0, precision, s);
    if (rc) return rc;
  } else {
    dim3 grid((unsigned)ceil_div(Nout, BM), (unsigned)ceil_div(K, BN), (unsigned)splits);
    sgemm_kernel<0><<<grid, kGemmThreads, 0, s>>>(dz, 1, Nout, X, K, 1, dst, K, Nout, K, M, k_per, nullptr);
    count_launch();
  }
  {
    const int64_t n = (int64_t)Nout * K;
    const int slab_blocks = splits > 1 ? (int)ceil_div(n, 256) : 0;
    const int col_blocks = (own.n + 7) / 8 + (in.n + 7) / 8;
    if (slab_blocks + col_blocks > 0) {
      reduce_tail_kernel<<<(unsigned)(slab_blocks + col_blocks), 256, 0, s>>>(ws, splits, n, dW, slab_blocks, own, in);
      count_launch();
    }
  }
  return check_launch("linear_bwd");
}

}  // namespace b200lab

using namespace b200lab;

extern "C" {

int64_t b200lab_linear_bwd_workspace_bytes(int64_t M, int32_t K, int32_t Nout) {
  int s_simt = dw_splits(M, K, Nout, false), s_tc = dw_splits(M, K, Nout, true);
  int splits = s_simt > s_tc ? s_simt : s_tc;   // the caller does not say which path will run
  int64_t colsum_blocks = ceil_div(M, kColsumRows);
  // + the partials of the input side's bias gradient (b200lab_linear_bwd_chain)
  return ((int64_t)splits * Nout * K + colsum_blocks * Nout + colsum_blocks * K) * (int64_t)sizeof(float);
}

int b200lab_linear_fwd(const float* X, const float* W, const float* b, float* Y, int64_t M, int32_t K, int32_t Nout,
                       int32_t elu, int32_t precision, void* stream) {
  B200LAB_CHECK_ARG(X && W && b && Y && M > 0 && K > 0 && Nout > 0, "bad argument");
  cudaStream_t s = (cudaStream_t)stream;
  if (precision != 0 && (Nout >= kMinTcWidth || K >= kMinTcWidth)) {
    int rc = gemm_tc(X, K, 1, W, 1, K, Y, Nout, M, Nout, K, K, 1, b, elu ? 2 : 1, precision, s);
    return rc ? rc : check_launch("linear_fwd (tcgen05)");
  }
  B200LAB_CHECK_ARG(precision >= 0 && precision <= 2, "precision must be 0 (fp32 FFMA), 1 (tf32) or 2 (tf32x3)");
  dim3 grid((unsigned)ceil_div(M, BM), (unsigned)ceil_div(Nout, BN), 1);
  if (elu)
    sgemm_kernel<2><<<grid, kGemmThreads, 0, s>>>(X, K, 1, W, 1, K, Y, Nout, M, Nout, K, K, b);
  else
    sgemm_kernel<1><<<grid, kGemmThreads, 0, s>>>(X, K, 1, W, 1, K, Y, Nout, M, Nout, K, K, b);
  count_launch();
  return check_launch("linear_fwd");
}

int b200lab_linear_bwd(const float* X, const float* W, const float* Y, const float* dY, float* dZ, float* dX,
                       float* dW, float* db, int64_t M, int32_t K, int32_t Nout, int32_t elu, int32_t precision,
                       void* workspace, void* stream) {
  return linear_bwd_impl(X, W, Y, dY, dZ, dX, dW, db, nullptr, M, K, Nout, elu ? B200LAB_BWD_ELU_OUT : 0, precision, workspace,
                         (cudaStream_t)stream);
}

int b200lab_linear_bwd_chain(const float* X, const float* W, const float* Y, const float* dY, float* dZ, float* dX,
                             float* dW, float* db, float* db_in, int64_t M, int32_t K, int32_t Nout, int32_t flags,
                             int32_t precision, void* workspace, void* stream) {
  B200LAB_CHECK_ARG((flags & ~(B200LAB_BWD_ELU_OUT | B200LAB_BWD_ELU_IN | B200LAB_BWD_DZ_GIVEN)) == 0, "unknown flag");
  B200LAB_CHECK_ARG(!(flags & B200LAB_BWD_ELU_IN) || (dX && db_in), "B200LAB_BWD_ELU_IN needs dX and db_in");
  B200LAB_CHECK_ARG(!((flags & B200LAB_BWD_ELU_OUT) && (flags & B200LAB_BWD_DZ_GIVEN)), "ELU_OUT and DZ_GIVEN exclude each other");
  return linear_bwd_impl(X, W, Y, dY, dZ, dX, dW, db, db_in, M, K, Nout, flags, precision, workspace, (cudaStream_t)stream);
}

}  // extern "C"
