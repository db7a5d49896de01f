// mlp_tc.cu -- tcgen05 (5th-gen tensor core) GEMM for the actor-critic MLP layers (RSL/networks/mlp.py:15-120 forward,
// autograd backward of RSL/algorithms/ppo.py:178-422), sm_100a only.
//
// C[M,N] (+bias, +ELU) = A[M,K] . B[N,K]^T with arbitrary element strides for A and B (so the same kernel serves
// Y = X W^T, dX = dZ W and dW = dZ^T X without transposed copies) and an optional split of K over grid.z.
//
// Arithmetic: the reference computes these layers in fp32 (torch's default: TF32 matmul off).  Tensor cores offer
// kind::tf32 (10-bit mantissa), so every operand is split while it is staged into shared memory,
//     x = hi + lo,  hi = x with the low 13 mantissa bits cleared (exactly representable in TF32),  lo = x - hi (exact),
// and each K step issues THREE MMAs into the same fp32 TMEM accumulator:  Ahi.Bhi + Ahi.Blo + Alo.Bhi  ("3xTF32").
// The dropped Alo.Blo term and the TF32 rounding of lo are O(2^-22) relative -- the result agrees with an fp32 FFMA
// GEMM to a few ulp.  precision 1 = single TF32 MMA (fast, reduced precision, opt-in only), precision 2 = 3xTF32.
//
// Structure (one CTA = one 128 x BN tile, 256 threads, 1 CTA / SM):
//   all 8 warps : global -> registers -> hi/lo split -> st.shared into the canonical K-major SWIZZLE_NONE UMMA layout
//                 (core matrix = 8 rows x 16 B; LBO padded by 16 B so the 128-bit stores are bank-conflict free)
//   thread 0    : tcgen05.mma.cta_group::1.kind::tf32 over the BK/8 K steps of the stage, tcgen05.commit -> mbarrier
//   2 stages    : the MMAs of k-tile t run while every warp stages k-tile t+1
//   epilogue    : tcgen05.ld (32 lanes x 32 bit, 16 columns at a time) -> +bias -> ELU -> global
// Operands are split by the threads, so they cannot come through TMA; the staging traffic (L2 -> SM) is what bounds the
// kernel, not the tensor pipe.
#include "common.cuh"

namespace b200lab {

namespace tc {

constexpr int BM = 128;        // UMMA M (cta_group::1)
constexpr int BK = 32;         // K elements per stage = 8 chunks of 4 tf32 (16 B) = 4 MMA K-steps of 8
constexpr int kThreads = 256;
constexpr int kStages = 1;     // one smem stage per CTA, 2-3 CTAs per SM: the CTAs' load latencies, stores and MMAs interleave

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred P1;\n\t"
      "LAB_WAIT:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
      "@P1 bra DONE;\n\t"
      "bra LAB_WAIT;\n\t"
      "DONE:\n\t"
      "}" ::"r"(bar), "r"(parity)
      : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}

// D[tmem] (+)= A[smem desc] . B[smem desc]^T, kind::tf32, issued by ONE thread
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// mbarrier arrive once every tcgen05.mma issued so far by this thread has completed
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// Shared-memory matrix descriptor, K-major, SWIZZLE_NONE (cute::UMMA::SmemDescriptor bit layout, version 1):
//   [0,14) start address >> 4 | [16,30) leading-dimension byte offset >> 4 (between the 16-byte K chunks)
//   [32,46) stride-dimension byte offset >> 4 (between 8-row groups) | [46,48) version = 1 | [61,64) layout type = 0
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr >> 4) & 0x3FFFu) | ((uint64_t)((lbo >> 4) & 0x3FFFu) << 16) |
         ((uint64_t)((sbo >> 4) & 0x3FFFu) << 32) | (1ull << 46);
}

// Instruction descriptor (cute::UMMA::InstrDescriptor): c_format F32 = 1 @[4,6), a/b_format TF32 = 2 @[7,10)/[10,13),
// a/b_major K = 0 @15/16, n_dim = N >> 3 @[17,23), m_dim = M >> 4 @[24,29)
__host__ __device__ constexpr uint32_t make_idesc(int m, int n) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

// Operand tile of ROWS x BK tf32 values in the canonical K-major SWIZZLE_NONE layout: core matrix = 8 rows x 16 B (4 k),
//   offset(r, k) = (k/4)*kLboK + r*16 + (k%4)*4          i.e. [K chunks][rows][4]
// with the chunk stride padded by 16 B so that the lanes of a store instruction fall into different banks.
template <int ROWS>
struct Tile {
  static constexpr uint32_t kLboK = ROWS * 16 + 16, kSboK = 128;
  static constexpr uint32_t kBytes = (BK / 4) * kLboK;
};

// how an operand is brought in: scalar loads | float4 along k | float4 along the rows, transposed while it is stored
enum StageMode { kScalarK = 0, kVecK = 1, kVecMN = 2 };

template <int BN>
struct Smem {
  static constexpr uint32_t kA = Tile<BM>::kBytes, kB = Tile<BN>::kBytes;
  static constexpr uint32_t kStage = 2 * kA + 2 * kB;         // A hi, A lo, B hi, B lo
  static constexpr uint32_t kTotal = kStages * kStage + 64;   // + barriers / tmem pointer
};

__device__ __forceinline__ float exp2f_fast(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

__device__ __forceinline__ void split_tf32(float x, float& hi, float& lo) {
  hi = __uint_as_float(__float_as_uint(x) & 0xFFFFE000u);
  lo = x - hi;
}

// Staging of ROWS x BK elements of a strided operand (element (r, k) at src + r*ld_r + k*ld_k; rows < rows_valid,
// k in [k0, k_end)) in two phases, so that ALL of a thread's loads are in flight together and -- issued right after the
// MMAs of the previous k-tile -- complete under them:  load_operand (global -> registers, raw; out-of-range slots read
// a safe address, no branches around the loads)  and  store_operand (zero out-of-range, hi/lo split, st.shared).
//   kVecK   : ld_k == 1, 16-byte aligned rows      -> float4 along k
//   kVecMN  : ld_r == 1, 16-byte aligned k-columns -> float4 along the rows, stored as 4 scalars (transposed into the
//             K-major layout); lanes = 2 adjacent row groups x 16 k, which keeps both the 32-byte sectors of the loads
//             full and the 32 scalar stores of a warp in 32 different banks
//   kScalarK: anything else (e.g. the 235-wide observation rows) -> scalar loads, K-major layout
template <int ROWS>
struct Frag { float f[ROWS * BK / kThreads]; };

// Per-thread addressing, fixed for the whole K loop: slot `it` lives at base + it*step (+ k0*ld_k per k-tile).
template <int ROWS>
struct Cursor {
  const float* base;
  int64_t step;
  uint32_t row_ok;       // bit `it`: the slot's row(s) are inside the matrix
  int k_local, k_step;   // k of slot `it` inside the k-tile = k_local + it * k_step
  int r_local, r_step;   // row of slot `it` inside the tile = r_local + it * r_step
};

template <int ROWS>
__device__ __forceinline__ Cursor<ROWS> make_cursor(const float* src, int64_t ld_r, int64_t ld_k, int64_t r0, int64_t rows_valid,
                                                    int mode, int tid) {
  constexpr int kElems = ROWS * BK / kThreads;
  Cursor<ROWS> c;
  c.row_ok = 0;
  if (mode == kVecK) {            // slot it: k = 4*(tid%8), r = tid/8 + 32*it
    c.k_local = (tid % (BK / 4)) * 4; c.k_step = 0;
    c.r_local = tid / (BK / 4);       c.r_step = kThreads / (BK / 4);
  } else if (mode == kVecMN) {    // slot it: k = 16*(warp%2) + lane/2, r = 4*(2*(warp/2 + 4*it) + lane%2)
    const int warp = tid >> 5, lane = tid & 31;
    c.k_local = (warp & 1) * 16 + (lane >> 1); c.k_step = 0;
    c.r_local = ((warp >> 1) * 2 + (lane & 1)) * 4; c.r_step = 32;
  } else if ((ld_k == 1) || (ld_k < ld_r)) {   // scalar, k fastest: k = tid%32, r = tid/32 + 8*it
    c.k_local = tid % BK; c.k_step = 0;
    c.r_local = tid / BK; c.r_step = kThreads / BK;
  } else {                        // scalar, rows fastest: r = tid%ROWS, k = tid/ROWS + (256/ROWS)*it
    c.r_local = tid % ROWS; c.r_step = 0;
    c.k_local = tid / ROWS; c.k_step = kThreads / ROWS;
  }
  c.base = src + (r0 + c.r_local) * ld_r + c.k_local * ld_k;
  c.step = c.r_step * ld_r + c.k_step * ld_k;
  const int slots = (mode == kScalarK) ? kElems : kElems / 4;
#pragma unroll
  for (int it = 0; it < kElems; ++it)
    if (it < slots && r0 + c.r_local + it * c.r_step < rows_valid) c.row_ok |= 1u << it;   // kVecMN: rows_valid % 4 == 0
  return c;
}

template <int ROWS>
__device__ __forceinline__ void load_operand(Frag<ROWS>& fr, const Cursor<ROWS>& c, const float* __restrict__ safe,
                                             int64_t ld_k, int64_t k0, int64_t k_end, int mode) {
  constexpr int kElems = ROWS * BK / kThreads;
  const float* p = c.base + k0 * ld_k;
  // raw values only: nothing may consume a load here, or the thread stalls on it instead of letting it fly under the
  // MMAs -- out-of-range slots are zeroed in store_operand
  if (mode != kScalarK) {
#pragma unroll
    for (int it = 0; it < kElems / 4; ++it) {
      const bool ok = ((c.row_ok >> it) & 1u) && (k0 + c.k_local + it * c.k_step < k_end);
      float4 v = __ldg(reinterpret_cast<const float4*>(ok ? p + it * c.step : safe));
      fr.f[it * 4 + 0] = v.x; fr.f[it * 4 + 1] = v.y; fr.f[it * 4 + 2] = v.z; fr.f[it * 4 + 3] = v.w;
    }
  } else {
#pragma unroll
    for (int it = 0; it < kElems; ++it) {
      const bool ok = ((c.row_ok >> it) & 1u) && (k0 + c.k_local + it * c.k_step < k_end);
      fr.f[it] = __ldg(ok ? p + it * c.step : safe);
    }
  }
}

template <int ROWS, bool SPLIT>
__device__ __forceinline__ void store_operand(const Frag<ROWS>& fr, const Cursor<ROWS>& c, int64_t k0, int64_t k_end, int mode,
                                              uint8_t* s_hi, uint8_t* s_lo) {
  constexpr int kElems = ROWS * BK / kThreads;
  if (mode != kScalarK) {
#pragma unroll
    for (int it = 0; it < kElems / 4; ++it) {
      const int k = c.k_local + it * c.k_step, r = c.r_local + it * c.r_step;
      const int64_t gk = k0 + k;
      const bool ok = ((c.row_ok >> it) & 1u) != 0;
      const bool along_k = mode == kVecK;   // the 4 lanes of the float4 step along k (ragged K tail) or along rows
      float x0 = (ok && gk < k_end) ? fr.f[it * 4 + 0] : 0.f;
      float x1 = (ok && gk + (along_k ? 1 : 0) < k_end) ? fr.f[it * 4 + 1] : 0.f;
      float x2 = (ok && gk + (along_k ? 2 : 0) < k_end) ? fr.f[it * 4 + 2] : 0.f;
      float x3 = (ok && gk + (along_k ? 3 : 0) < k_end) ? fr.f[it * 4 + 3] : 0.f;
      float4 h, l;
      split_tf32(x0, h.x, l.x); split_tf32(x1, h.y, l.y); split_tf32(x2, h.z, l.z); split_tf32(x3, h.w, l.w);
      const uint32_t off = (uint32_t)(k >> 2) * Tile<ROWS>::kLboK + (uint32_t)r * 16 + (uint32_t)(k & 3) * 4;
      if (along_k) {
        *reinterpret_cast<float4*>(s_hi + off) = h;
        if (SPLIT) *reinterpret_cast<float4*>(s_lo + off) = l;
      } else {   // the float4 runs along 4 consecutive rows of one k
        *reinterpret_cast<float*>(s_hi + off) = h.x; *reinterpret_cast<float*>(s_hi + off + 16) = h.y;
        *reinterpret_cast<float*>(s_hi + off + 32) = h.z; *reinterpret_cast<float*>(s_hi + off + 48) = h.w;
        if (SPLIT) {
          *reinterpret_cast<float*>(s_lo + off) = l.x; *reinterpret_cast<float*>(s_lo + off + 16) = l.y;
          *reinterpret_cast<float*>(s_lo + off + 32) = l.z; *reinterpret_cast<float*>(s_lo + off + 48) = l.w;
        }
      }
    }
  } else {
#pragma unroll
    for (int it = 0; it < kElems; ++it) {
      const int k = c.k_local + it * c.k_step, r = c.r_local + it * c.r_step;
      const bool ok = ((c.row_ok >> it) & 1u) && (k0 + k < k_end);
      float h, l;
      split_tf32(ok ? fr.f[it] : 0.f, h, l);
      const uint32_t off = (uint32_t)(k >> 2) * Tile<ROWS>::kLboK + (uint32_t)r * 16 + (uint32_t)(k & 3) * 4;
      *reinterpret_cast<float*>(s_hi + off) = h;
      if (SPLIT) *reinterpret_cast<float*>(s_lo + off) = l;
    }
  }
}

// EPI: 0 = store, 1 = +bias, 2 = +bias then ELU.  SPLIT: 3xTF32 (true) or single TF32 (false).
template <int BN, int EPI, bool SPLIT>
__global__ void __launch_bounds__(kThreads, 2)
gemm_tf32_kernel(const float* __restrict__ A, int64_t lda_i, int64_t lda_k, const float* __restrict__ B, int64_t ldb_k,
                 int64_t ldb_j, float* __restrict__ C, int64_t ldc, int64_t M, int N, int64_t K, int64_t k_per_split,
                 const float* __restrict__ bias, int a_mode, int b_mode) {
  extern __shared__ __align__(1024) uint8_t smem[];
  using S = Smem<BN>;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t i0 = (int64_t)blockIdx.x * BM;
  const int j0 = blockIdx.y * BN;
  const int64_t k_begin = (int64_t)blockIdx.z * k_per_split;
  const int64_t k_end = min64(K, k_begin + k_per_split);
  const int num_kt = (int)((k_end - k_begin + BK - 1) / BK);

  uint8_t* ctl = smem + kStages * S::kStage;
  const uint32_t bar_stage0 = smem_u32(ctl), bar_done = smem_u32(ctl + 8 * kStages);
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(ctl + 8 * kStages + 8);

  if (tid == 0) {
    for (int s = 0; s < kStages; ++s) mbar_init(bar_stage0 + 8 * s, 1);
    mbar_init(bar_done, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) tmem_alloc(smem_u32(tmem_ptr_smem), BN);   // BN in {32,64,128,256}: power of two >= 32 columns
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = *tmem_ptr_smem;

  constexpr uint32_t idesc = make_idesc(BM, BN);   // both operands K-major in shared memory
  uint32_t phase_bits = 0;   // bit s = parity to wait for on stage s
  Frag<BM> fa;
  Frag<BN> fb;
  const Cursor<BM> ca = make_cursor<BM>(A, lda_i, lda_k, i0, M, a_mode, tid);
  const Cursor<BN> cb = make_cursor<BN>(B, ldb_j, ldb_k, j0, N, b_mode, tid);
  if (num_kt > 0) {
    load_operand<BM>(fa, ca, A, lda_k, k_begin, k_end, a_mode);
    load_operand<BN>(fb, cb, B, ldb_k, k_begin, k_end, b_mode);
  }
  for (int kt = 0; kt < num_kt; ++kt) {
    const int s = kt % kStages;
    uint8_t* st = smem + s * S::kStage;
    uint8_t *a_hi = st, *a_lo = st + S::kA, *b_hi = st + 2 * S::kA, *b_lo = st + 2 * S::kA + S::kB;
    if (kt >= kStages) {   // the MMAs that read this stage (k-tile kt - kStages) must have drained
      mbar_wait(bar_stage0 + 8 * s, (phase_bits >> s) & 1u);
      phase_bits ^= 1u << s;
    }
    const int64_t k_cur = k_begin + (int64_t)kt * BK;
    store_operand<BM, SPLIT>(fa, ca, k_cur, k_end, a_mode, a_hi, a_lo);
    store_operand<BN, SPLIT>(fb, cb, k_cur, k_end, b_mode, b_hi, b_lo);
    fence_proxy_async();   // generic-proxy smem writes -> visible to the tensor core (async proxy)
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      const uint32_t sa_hi = smem_u32(a_hi), sa_lo = smem_u32(a_lo), sb_hi = smem_u32(b_hi), sb_lo = smem_u32(b_lo);
#pragma unroll
#pragma unroll
      for (int ks = 0; ks < BK / 8; ++ks) {   // one MMA K-step = 8 tf32 = two 16-byte chunks
        const uint32_t oa = ks * 2 * Tile<BM>::kLboK, ob = ks * 2 * Tile<BN>::kLboK;
        const uint64_t da_hi = make_desc(sa_hi + oa, Tile<BM>::kLboK, Tile<BM>::kSboK);
        const uint64_t db_hi = make_desc(sb_hi + ob, Tile<BN>::kLboK, Tile<BN>::kSboK);
        umma_tf32(tmem_d, da_hi, db_hi, idesc, (kt | ks) != 0);
        if (SPLIT) {
          const uint64_t da_lo = make_desc(sa_lo + oa, Tile<BM>::kLboK, Tile<BM>::kSboK);
          const uint64_t db_lo = make_desc(sb_lo + ob, Tile<BN>::kLboK, Tile<BN>::kSboK);
          umma_tf32(tmem_d, da_hi, db_lo, idesc, 1u);
          umma_tf32(tmem_d, da_lo, db_hi, idesc, 1u);
        }
      }
      umma_commit(bar_stage0 + 8 * s);            // frees the stage for k-tile kt + kStages
      if (kt == num_kt - 1) umma_commit(bar_done);  // accumulator complete
    }
    if (kt + 1 < num_kt) {   // next k-tile's loads fly under this k-tile's MMAs (issued AFTER the fence above: the
                             // membar it lowers to would otherwise wait for them)
      const int64_t k0 = k_begin + (int64_t)(kt + 1) * BK;
      load_operand<BM>(fa, ca, A, lda_k, k0, k_end, a_mode);
      load_operand<BN>(fb, cb, B, ldb_k, k0, k_end, b_mode);
    }
  }
  // ---- epilogue: TMEM -> registers -> (32 x 32 transpose through shared memory) -> global.  Warp w may read TMEM lanes
  // 32*(w%4)..+31 (= tile rows, one per lane); the two warps that share a lane quarter split the columns.  Written straight
  // from the TMEM layout every store instruction would touch 32 rows (32 sectors); after the transpose it writes 128
  // contiguous bytes of one row.
  if (num_kt > 0) mbar_wait(bar_done, 0);
  tc_fence_after();
  {
    const int q = warp & 3, half = warp >> 2;
    float* Cz = C + (int64_t)blockIdx.z * M * ldc;
    float* tile = reinterpret_cast<float*>(smem) + warp * (32 * 33);   // stage buffers are free: every MMA has completed
    constexpr int kColsPerHalf = BN / 2;
#pragma unroll 1
    for (int c0 = half * kColsPerHalf; c0 < (half + 1) * kColsPerHalf; c0 += 32) {
      if (j0 + c0 >= N) break;   // warp-uniform
      float v[32];
      if (num_kt > 0) {
        float lo16[16], hi16[16];
        tmem_ld16(tmem_d + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, lo16);
        tmem_ld16(tmem_d + ((uint32_t)(q * 32) << 16) + (uint32_t)(c0 + 16), hi16);
#pragma unroll
        for (int i = 0; i < 16; ++i) { v[i] = lo16[i]; v[16 + i] = hi16[i]; }
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = 0.f;
      }
#pragma unroll
      for (int i = 0; i < 32; ++i) tile[lane * 33 + i] = v[i];
      __syncwarp();
      const int gj = j0 + c0 + lane;
      const bool col_ok = gj < N;
      float bv = 0.f;
      if (EPI >= 1 && col_ok) bv = __ldg(bias + gj);
#pragma unroll 8
      for (int rr = 0; rr < 32; ++rr) {
        const int64_t gi = i0 + q * 32 + rr;
        if (gi >= M) break;      // warp-uniform
        float x = tile[rr * 33 + lane];
        if (EPI >= 1) x += bv;
        if (EPI == 2) x = x > 0.0f ? x : exp2f_fast(1.4426950408889634f * x) - 1.0f;   // |abs err| ~1e-7
        if (col_ok) Cz[gi * ldc + gj] = x;
      }
      __syncwarp();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_d, BN);
}

template <int BN, int EPI, bool SPLIT>
static int launch(const float* A, int64_t lda_i, int64_t lda_k, const float* B, int64_t ldb_k, int64_t ldb_j, float* C,
                  int64_t ldc, int64_t M, int N, int64_t K, int64_t k_per_split, int splits, const float* bias,
                  cudaStream_t s) {
  auto kern = gemm_tf32_kernel<BN, EPI, SPLIT>;
  static bool configured = false;
  if (!configured) {
    cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Smem<BN>::kTotal);
    if (err != cudaSuccess) { set_error("tcgen05 gemm: %s", cudaGetErrorString(err)); return B200LAB_ERR_CUDA; }
    configured = true;
  }
  // 128-bit loads along whichever dimension is contiguous; both need 16-byte aligned vectors, kVecK split boundaries on
  // multiples of 4, kVecMN whole groups of 4 rows
  auto mode_of = [&](const float* p, int64_t ld_r, int64_t ld_k, int64_t rows) -> int {
    if ((uintptr_t)p % 16 != 0) return kScalarK;
    if (ld_k == 1 && ld_r % 4 == 0 && k_per_split % 4 == 0) return kVecK;
    if (ld_r == 1 && ld_k % 4 == 0 && rows % 4 == 0) return kVecMN;
    return kScalarK;
  };
  dim3 grid((unsigned)ceil_div(M, BM), (unsigned)ceil_div(N, BN), (unsigned)splits);
  kern<<<grid, kThreads, Smem<BN>::kTotal, s>>>(A, lda_i, lda_k, B, ldb_k, ldb_j, C, ldc, M, N, K, k_per_split, bias,
                                               mode_of(A, lda_i, lda_k, M), mode_of(B, ldb_j, ldb_k, N));
  count_launch();
  return B200LAB_OK;
}

template <int EPI, bool SPLIT>
static int launch_bn(const float* A, int64_t lda_i, int64_t lda_k, const float* B, int64_t ldb_k, int64_t ldb_j, float* C,
                     int64_t ldc, int64_t M, int N, int64_t K, int64_t k_per_split, int splits, const float* bias,
                     cudaStream_t s) {
  if (N > 64) return launch<128, EPI, SPLIT>(A, lda_i, lda_k, B, ldb_k, ldb_j, C, ldc, M, N, K, k_per_split, splits, bias, s);
  return launch<64, EPI, SPLIT>(A, lda_i, lda_k, B, ldb_k, ldb_j, C, ldc, M, N, K, k_per_split, splits, bias, s);
}

}  // namespace tc

// Same contract as the SIMT sgemm_kernel launch in mlp.cu.  epi: 0 store, 1 +bias, 2 +bias+ELU.
int gemm_tc(const float* A, int64_t lda_i, int64_t lda_k, const float* B, int64_t ldb_k, int64_t ldb_j, float* C, int64_t ldc,
            int64_t M, int N, int64_t K, int64_t k_per_split, int splits, const float* bias, int epi, int precision,
            cudaStream_t s) {
  if (precision != 1 && precision != 2) {
    set_error("gemm: precision %d unknown (0 fp32 FFMA, 1 tf32, 2 tf32x3)", precision);
    return B200LAB_ERR_UNSUPPORTED;
  }
  const bool split = precision == 2;
#define B200LAB_TC_DISPATCH(E)                                                                                          \
  return split ? tc::launch_bn<E, true>(A, lda_i, lda_k, B, ldb_k, ldb_j, C, ldc, M, N, K, k_per_split, splits, bias, s) \
               : tc::launch_bn<E, false>(A, lda_i, lda_k, B, ldb_k, ldb_j, C, ldc, M, N, K, k_per_split, splits, bias, s)
  if (epi == 2) { B200LAB_TC_DISPATCH(2); }
  if (epi == 1) { B200LAB_TC_DISPATCH(1); }
  B200LAB_TC_DISPATCH(0);
#undef B200LAB_TC_DISPATCH
}

}  // namespace b200lab
