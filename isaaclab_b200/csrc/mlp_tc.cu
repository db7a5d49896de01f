// mlp_tc.cu -- tcgen05 (5th-gen tensor core) GEMM for the actor-critic MLP layers (RSL/networks/mlp.py:15-120 forward,
// autograd backward of RSL/algorithms/ppo.py:178-422), sm_100a only.
//
// C[M,N] (+bias, +ELU) = A[M,K] . B[N,K]^T with arbitrary element strides for A and B (so the same kernel serves
// Y = X W^T, dX = dZ W and dW = dZ^T X without transposed copies) and an optional split of K over grid.z.
//
// Arithmetic: the reference computes these layers in fp32 (torch's default: TF32 matmul off).  Tensor cores offer
// kind::tf32, which reads the top 19 bits of each 32-bit operand word and IGNORES the low 13 mantissa bits (measured:
// tools/umma_probe.cu -- truncation, not rounding).  So the raw fp32 tile in shared memory already is the "hi" operand,
//     x = hi + lo,  hi = x with the low 13 mantissa bits cleared,  lo = x - hi (exact in fp32),
// only `lo` has to be computed, and each K step issues THREE MMAs into the same fp32 TMEM accumulator:
//     Araw.Braw + Araw.Blo + Alo.Braw  ("3xTF32").  The dropped Alo.Blo term and the truncation of lo are O(2^-22)
// relative: the result agrees with an fp32 FFMA GEMM to a few ulp.  precision 1 = the first MMA only (opt-in).
//
// Structure (persistent: one CTA per SM walks the 128 x BN output tiles b, b + gridDim.x, ...; 16 producer warps + 1 MMA
// warp + 4 epilogue warps; ring of kStages raw stages + 2 lo buffers; TWO accumulators in TMEM):
//   producer warps : TMA boxes (K-contiguous, aligned operands: one thread issues, completion by mbarrier transaction
//                    bytes) or cp.async global -> shared (16-byte copies wherever the operand is contiguous and aligned
//                    along M/N, 4-byte copies otherwise; zero-fill for out-of-range rows / K tail), then a linear sweep
//                    writes lo = x - trunc(x) of the stage to a lo buffer, fence.proxy.async, mbarrier.arrive(full[s]).
//                    The ring does not drain between tiles: the k-tiles of the CTA's next output tile are requested
//                    while the last MMAs of the current one run.
//   MMA warp       : one lane waits full[s], issues the BK/8 x {1,3} tcgen05.mma.cta_group::1.kind::tf32 into accumulator
//                    n & 1, and tcgen05.commit -> empty[s] (releases the stage to the producers); after the tile's last
//                    k-tile, commit -> acc_full[n & 1].
//   epilogue warps : wait acc_full, tcgen05.ld the accumulator (lane = row, one warp per lane quarter), arrive acc_empty
//                    as soon as the last columns are in registers, transpose 32 x 32 through shared memory, +bias, ELU,
//                    128-bit stores of 4 rows x 128 contiguous bytes -- all under the MMAs of the next tile.
//   Measured (profiles/r2_notes.md): the one-tile-per-CTA form of this kernel paid ~40 % of a short-K tile's time in
//   prologue (barrier init, TMEM allocation, ring fill) and epilogue; cfg 3 update 22.5 -> 20.6 ms.
// Shared-memory operand layouts (word maps read back from the hardware with tools/umma_probe.cu):
//   K-contiguous operand  : K-major SWIZZLE_128B (layout type 2): offset(r,k) = (r/8)*1024 + (r%8)*128 +
//                           (((k/4) ^ (r%8))*16) + (k%4)*4 -- a row's 32 tf32 are one 128-byte line, so the copies of a
//                           warp land as full 128-byte wavefronts (the un-swizzled core-matrix layout scatters a row's
//                           16-byte chunks and cost 9x the shared-memory wavefronts: profiles/r1_notes.md)
//   M/N-contiguous operand: MN-major SWIZZLE_128B_BASE32B (layout type 1, the only MN-major layout 32-bit operands
//                           have): offset(mn,k) = swz((mn/32)*LBO + (k/4)*SBO + (k%4)*128 + (mn%32)*4), LBO = 4096,
//                           SBO = 512, swz: byte-address bits [5,7) ^= bits [7,9)
// so dX (B = W read along its rows) and dW (both operands read along the sample dimension) need no transposition.
#include <cuda.h>   // CUtensorMap and the cuTensorMapEncodeTiled prototype; the entry point is resolved at run time

#include "common.cuh"

namespace b200lab {

namespace tc {

constexpr int BM = 128;               // UMMA M (cta_group::1)
constexpr int BK = 32;                // K elements per stage = 4 MMA K-steps of 8 tf32
constexpr int kProducers = 512;       // 16 warps stage operands (4 per scheduler hide the ALU latency of the lo sweep)
constexpr int kEpiWarps = 4;          // one per TMEM lane quarter: drain accumulator n while the MMAs fill accumulator n+1
constexpr int kThreads = kProducers + 32 + kEpiWarps * 32 + 32;   // producers | the MMA warp | epilogue warps | the TMA warp
constexpr int kMaxStages = 8;
constexpr int kEpiLd = 36;            // floats per row of an epilogue warp's 32 x 32 transpose tile

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
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
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// TMA: one 2-D box (c0 = element index along K, c1 = row) global -> shared, completion counted in bytes on `bar`
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
               ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(bar) : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async4(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

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
// same with the A operand in tensor memory (lane = row, one 32-bit column per tf32 element; tools/umma_tmem_a_probe.cu)
__device__ __forceinline__ void umma_tf32_ta(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t"
      "}" ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// 8 consecutive columns of this thread's TMEM lane (warp w owns lanes 32*(w%4)..+31)
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const float (&v)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr),
               "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])), "r"(__float_as_uint(v[3])),
               "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])), "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7]))
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

// 32 consecutive columns with ONE wait (two x16 loads in flight together)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, "
      "%19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]),
        "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]),
        "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

// Shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout, version 1):
//   [0,14) start address >> 4 | [16,30) leading-dimension byte offset >> 4 | [32,46) stride-dimension byte offset >> 4
//   [46,48) version = 1 | [61,64) layout type (0 = SWIZZLE_NONE, 1 = SWIZZLE_128B_BASE32B, 2 = SWIZZLE_128B)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
  return (uint64_t)((saddr >> 4) & 0x3FFFu) | ((uint64_t)((lbo >> 4) & 0x3FFFu) << 16) |
         ((uint64_t)((sbo >> 4) & 0x3FFFu) << 32) | (1ull << 46) | ((uint64_t)layout << 61);
}

// Instruction descriptor (cute::UMMA::InstrDescriptor): c_format F32 = 1 @[4,6), a/b_format TF32 = 2 @[7,10)/[10,13),
// a/b_major (0 = K, 1 = MN) @15/16, n_dim = N >> 3 @[17,23), m_dim = M >> 4 @[24,29)
__host__ __device__ constexpr uint32_t make_idesc(int m, int n) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

// How an operand tile of ROWS x BK elements (element (r, k) at src + r*ld_r + k*ld_k) is brought in and laid out.
enum OperandMode {
  kKVec = 0,      // ld_k == 1, rows 16-byte aligned      -> 16-byte copies along k,    K-major layout
  kKScalar = 1,   // k fastest but unaligned (235-wide)   ->  4-byte copies,            K-major layout
  kMnVec = 2,     // ld_r == 1, k-columns 16-byte aligned -> 16-byte copies along rows, MN-major layout
  kMnScalar = 3,  // rows fastest but unaligned           ->  4-byte copies,            MN-major layout
  kKTma = 4,      // as kKVec, moved by TMA: ONE thread issues a ROWS x BK box per k-tile (SWIZZLE_128B tensor map writes
                  // exactly the K-major layout; rows / K beyond the matrix arrive as zeros), completion on an mbarrier
  kMnTma = 5,     // as kMnVec, moved by TMA: ROWS/32 boxes of 32 rows x BK per k-tile (SWIZZLE_128B_ATOM_32B tensor map = the
                  // MN-major layout: a box is one 4 KB group of BK/4 k-atoms)
};

template <int ROWS>
struct Tile {
  // K-major SWIZZLE_128B: a row's BK = 32 tf32 are one 128-byte line, 8-row groups 1024 B apart (SBO); LBO unused
  static constexpr uint32_t kSboK = 1024;
  // MN-major SWIZZLE_128B_BASE32B: 32 rows x 4 k atoms of 512 B; k-atoms adjacent (SBO), 32-row groups LBO apart
  static constexpr uint32_t kLboMn = (BK / 4) * 512, kSboMn = 512;
  static constexpr uint32_t kBytes = ROWS * BK * 4;   // either layout, no padding; buffers are 1024-byte aligned
};

// K-major SWIZZLE_128B: 16-byte chunk index ^= row % 8 (byte-address bits [4,7) ^= bits [7,10))
__device__ __forceinline__ uint32_t k_offset(int r, int k) {
  return ((uint32_t)(r >> 3) << 10) + ((uint32_t)(r & 7) << 7) + ((((uint32_t)(k >> 2) ^ (uint32_t)r) & 7u) << 4) + ((uint32_t)(k & 3) << 2);
}
__device__ __forceinline__ uint32_t swz_mn(uint32_t off) { return off ^ (((off >> 7) & 3u) << 5); }

// Per-thread, per-operand addressing, fixed for the whole K loop (MODE is a template parameter: no runtime dispatch in
// the loop).  Slot `it` of a thread copies the element(s) at row r_local + it*r_step, k = k_local + it*k_step (first
// element of the 16-byte vector in the vec modes) from  cur + it*step  to shared offset  off0 + it*off_step;  `cur`
// advances by BK*ld_k per k-tile.  One of r_step / k_step is always 0.
template <int ROWS, int MODE>
struct Cursor {
  static constexpr bool kVec = MODE == kKVec || MODE == kMnVec;
  static constexpr int kSlots = kVec ? ROWS * BK / 4 / kProducers : ROWS * BK / kProducers;
  static constexpr int kRowsPerPass = MODE == kKVec ? kProducers / 8 : kProducers / BK;              // K-major modes
  static constexpr int kKPerPass = MODE == kMnVec ? kProducers / (ROWS / 4) : kProducers / ROWS;     // MN-major modes
  static constexpr bool kKMajor = MODE == kKVec || MODE == kKScalar;
  static constexpr uint32_t kOffStep = kKMajor ? (uint32_t)(kRowsPerPass / 8) * Tile<ROWS>::kSboK : (uint32_t)(kKPerPass / 4) * Tile<ROWS>::kSboMn;
  static_assert(!kKMajor || kRowsPerPass % 8 == 0, "K-major slots must keep their swizzle phase");
  static_assert(kSlots >= 1, "tile too small for the producer count");
  static_assert(kKMajor || kKPerPass % 4 == 0, "MN-major slots must keep their swizzle phase");
  const float* cur;     // source of slot 0 for the NEXT k-tile to be issued
  int64_t step, tile_step;
  uint32_t off0;
  uint32_t row_ok;      // K-major: bit it = slot's row is inside the matrix.  MN-major: valid rows of the vector / element (0..4)
  int k_left;           // k elements left from slot 0's k for the next k-tile to be issued (may be <= 0)

  __device__ __forceinline__ void init(const float* src, int64_t ld_r, int64_t ld_k, int64_t r0, int64_t rows_valid,
                                       int64_t k_begin, int64_t k_end, int tid) {
    int r_local, k_local;
    if (MODE == kKVec)         { k_local = (tid & 7) * 4; r_local = tid >> 3; }
    else if (MODE == kKScalar) { k_local = tid & (BK - 1); r_local = tid >> 5; }
    else if (MODE == kMnVec)   { r_local = (tid % (ROWS / 4)) * 4; k_local = tid / (ROWS / 4); }
    else                       { r_local = tid % ROWS; k_local = tid / ROWS; }
    if (kKMajor) {
      off0 = k_offset(r_local, k_local);
      step = (int64_t)kRowsPerPass * ld_r;
      row_ok = 0;
#pragma unroll
      for (int it = 0; it < kSlots; ++it)
        if (r0 + r_local + it * kRowsPerPass < rows_valid) row_ok |= 1u << it;
    } else {
      off0 = swz_mn((uint32_t)(r_local >> 5) * Tile<ROWS>::kLboMn + (uint32_t)(k_local >> 2) * Tile<ROWS>::kSboMn +
                    (uint32_t)(k_local & 3) * 128 + (uint32_t)(r_local & 31) * 4);
      step = (int64_t)kKPerPass * ld_k;
      const int64_t left = rows_valid - (r0 + r_local);
      row_ok = (uint32_t)(left < 0 ? 0 : (left > 4 ? 4 : left));
    }
    cur = src + (r0 + r_local) * ld_r + (k_begin + k_local) * ld_k;
    tile_step = (int64_t)BK * ld_k;
    k_left = (int)min64(k_end - k_begin - k_local, 1 << 28);
  }

  // cp.async of this thread's slots of the next k-tile; rows outside the matrix and the K tail are zero-filled by the
  // copy (src-size < cp-size reads only src-size bytes -- the address of a fully masked copy is never dereferenced)
  __device__ __forceinline__ void issue(uint32_t s_raw) {
    const float* p = cur;
#pragma unroll
    for (int it = 0; it < kSlots; ++it) {
      uint32_t bytes;
      if (MODE == kKVec) {
        const int kl = k_left < 0 ? 0 : (k_left > 4 ? 4 : k_left);
        bytes = ((row_ok >> it) & 1u) ? (uint32_t)kl * 4u : 0u;
      } else if (MODE == kKScalar) {
        bytes = (((row_ok >> it) & 1u) && k_left > 0) ? 4u : 0u;
      } else if (MODE == kMnVec) {
        bytes = (k_left - it * kKPerPass > 0) ? row_ok * 4u : 0u;
      } else {
        bytes = (k_left - it * kKPerPass > 0 && row_ok) ? 4u : 0u;
      }
      if (kVec) cp_async16(s_raw + off0 + (uint32_t)it * kOffStep, p, bytes);
      else      cp_async4(s_raw + off0 + (uint32_t)it * kOffStep, p, bytes);
      p += step;
    }
    cur += tile_step;
    k_left -= BK;
  }
};

template <int ROWS>
struct Cursor<ROWS, kKTma> {
  const CUtensorMap* map;
  int r0, k_next;
  bool leader;
  __device__ __forceinline__ void init(const CUtensorMap* m, int64_t row0, int64_t k_begin, int tid) {
    map = m; r0 = (int)row0; k_next = (int)k_begin; leader = tid == 0;
  }
  __device__ __forceinline__ void issue(uint32_t s_raw, uint32_t bar) {
    if (leader) tma_load_2d(s_raw, map, k_next, r0, bar);
    k_next += BK;
  }
};

template <int ROWS>
struct Cursor<ROWS, kMnTma> {
  const CUtensorMap* map;
  int r0, k_next;
  bool leader;
  __device__ __forceinline__ void init(const CUtensorMap* m, int64_t row0, int64_t k_begin, int tid) {
    map = m; r0 = (int)row0; k_next = (int)k_begin; leader = tid == 0;
  }
  __device__ __forceinline__ void issue(uint32_t s_raw, uint32_t bar) {
    if (leader) {
#pragma unroll
      for (int grp = 0; grp < ROWS / 32; ++grp) tma_load_2d(s_raw + (uint32_t)grp * Tile<ROWS>::kLboMn, map, r0 + grp * 32, k_next, bar);
    }
    k_next += BK;
  }
};

// Shared memory: a ring of kStages RAW stages (A | B, filled by cp.async) and TWO lo buffers (A lo | B lo).  Only the
// k-tile being multiplied needs its lo part, so the bytes in flight from L2 are not halved by it: kStages-1 raw tiles
// are outstanding while one tile is split and another is multiplied.
template <int BN>
struct Smem {
  static constexpr uint32_t kA = Tile<BM>::kBytes, kB = Tile<BN>::kBytes;
  static constexpr uint32_t kStage = kA + kB;                      // one k-tile: A | B (raw, or their lo parts)
  static constexpr int kSlots = BN > 64 ? 6 : 8;                   // k-tile slots of 32 KB / 24 KB
  // all-TMA pipeline: TF32 -> kSlots raw stages; 3xTF32 -> kNsSplit raw stages + kLoRing lo buffers of the B part only (the
  // A operand of the split goes to the tensor core through tensor memory: 64 columns per stage next to the accumulators)
  // kLoRing: lo buffers of B (shared memory) and raw | lo copies of A (tensor memory) live from a k-tile's sweep to the end
  // of its MMAs only, so a ring of 2 serves any number of raw stages (the sweep of k-tile g waits for the MMAs of g - 2)
  static constexpr int kLoRing = 2;
  static constexpr int kNsSplit = (kSlots * kStage - kLoRing * kB) / kStage;
  static_assert(2 * BN + 64 * kLoRing <= 512, "accumulators + A copies exceed tensor memory");
  // cp.async pipeline: kSlots - 2 raw stages + 2 lo buffers
  static constexpr int kStages = kSlots - 2;
  static constexpr uint32_t kLo = kStages * kStage;
  static constexpr uint32_t kEpi = kSlots * kStage;                // one 32 x 32 transpose tile per epilogue warp
  static constexpr uint32_t kEpiBytes = kEpiWarps * 32 * kEpiLd * 4;
  static constexpr uint32_t kCtl = kEpi + kEpiBytes;               // full / empty / tma [kMaxStages], acc_full[2], acc_empty[2], tmem pointer
  static constexpr uint32_t kTotal = kCtl + 24 * kMaxStages + 32 + 16 + 1024;   // + slack to align the base to 1024
  static_assert(kSlots <= kMaxStages, "barrier arrays too small");
};

__device__ __forceinline__ float exp2f_fast(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

// lo = x - trunc_tf32(x) over a whole raw stage (A | B), written to the lo buffer in the SAME layout -- so the pass is a
// linear sweep, independent of the operand layouts.  (Each thread must only touch bytes whose copies it can see: the
// sweep runs after a CTA-wide hand-over of "my copies have landed", see the producer loop.)
template <uint32_t BYTES>
__device__ __forceinline__ void lo_sweep(const uint8_t* __restrict__ raw, uint8_t* __restrict__ lo, int tid) {
  constexpr int kChunks = BYTES / 16;
#pragma unroll
  for (int c0 = 0; c0 < kChunks; c0 += kProducers) {
    const int c = c0 + tid;
    if (c0 + kProducers <= kChunks || c < kChunks) {
      const float4 v = *reinterpret_cast<const float4*>(raw + (size_t)c * 16);
      float4 l;
      l.x = v.x - __uint_as_float(__float_as_uint(v.x) & 0xFFFFE000u);
      l.y = v.y - __uint_as_float(__float_as_uint(v.y) & 0xFFFFE000u);
      l.z = v.z - __uint_as_float(__float_as_uint(v.z) & 0xFFFFE000u);
      l.w = v.w - __uint_as_float(__float_as_uint(v.w) & 0xFFFFE000u);
      *reinterpret_cast<float4*>(lo + (size_t)c * 16) = l;
    }
  }
}

template <int ROWS>
__device__ __forceinline__ uint64_t operand_desc(uint32_t base, int ks, bool mn_major) {
  // one MMA K-step = 8 tf32 = 32 bytes along the swizzled 128-byte row (K-major; the hardware applies the XOR to the
  // final address) or two 4-row k-atoms (MN-major)
  return mn_major ? make_desc(base + (uint32_t)ks * 2u * Tile<ROWS>::kSboMn, Tile<ROWS>::kLboMn, Tile<ROWS>::kSboMn, 1u)
                  : make_desc(base + (uint32_t)ks * 32u, 16u, Tile<ROWS>::kSboK, 2u);
}

// One unit of work of the persistent loop: output tile (i0, j0) of K range [k_begin, k_end) (a split of K writes slab z).
struct Unit {
  int64_t i0, k_begin, k_end;
  int j0, z, num_kt;
};
__device__ __forceinline__ Unit unit_of(int64_t u, int n_i, int n_j, int64_t K, int64_t k_per_split) {
  // j fastest: CTAs that run side by side share the A panel (read from HBM once, then from L2)
  const int64_t per_z = (int64_t)n_i * n_j;
  Unit w;
  w.z = (int)(u / per_z);
  const int64_t r = u - (int64_t)w.z * per_z;
  w.i0 = (r / n_j) * BM;
  w.j0 = (int)(r % n_j);
  w.k_begin = (int64_t)w.z * k_per_split;
  w.k_end = min64(K, w.k_begin + k_per_split);
  w.num_kt = (int)((w.k_end - w.k_begin + BK - 1) / BK);
  return w;
}

// epi: 0 = store, 1 = +bias, 2 = +bias then ELU.  split: 3xTF32 (1) or single TF32 (0).
// Persistent: CTA b works through units b, b + gridDim.x, ... .  The operand ring runs across unit boundaries (the loads
// of the next tile are in flight under the last MMAs of this one) and the accumulator is double-buffered in TMEM, so the
// epilogue of unit n (TMEM -> registers -> transpose -> global, by its own 4 warps) runs under the MMAs of unit n+1.
template <int BN, int AMODE, int BMODE>
__global__ void __launch_bounds__(kThreads, 1)
gemm_tf32_kernel(const float* __restrict__ A, int64_t lda_i, int64_t lda_k, const float* __restrict__ B, int64_t ldb_k,
                 int64_t ldb_j, float* __restrict__ C, int64_t ldc, int64_t M, int N, int64_t K, int64_t k_per_split,
                 int splits, const float* __restrict__ bias, int epi, int split, const __grid_constant__ CUtensorMap map_a,
                 const __grid_constant__ CUtensorMap map_b) {
  extern __shared__ uint8_t smem_dyn[];
  constexpr bool kTmaA = AMODE == kKTma || AMODE == kMnTma, kTmaB = BMODE == kKTma || BMODE == kMnTma, kAnyTma = kTmaA || kTmaB, kAllTma = kTmaA && kTmaB;
  constexpr uint32_t kTmaBytes = (kTmaA ? Tile<BM>::kBytes : 0u) + (kTmaB ? Tile<BN>::kBytes : 0u);
  using S = Smem<BN>;
  constexpr int kStages = S::kStages;
  constexpr int kMmaWarp = kProducers / 32;
  uint8_t* smem = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);   // MN-major swizzle atoms: 1024-byte aligned
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n_i = (int)((M + BM - 1) / BM), n_j = (N + BN - 1) / BN;
  const int64_t num_units = (int64_t)n_i * n_j * splits;

  uint8_t* ctl = smem + S::kCtl;
  const uint32_t bar_full = smem_u32(ctl), bar_empty = bar_full + 8 * kMaxStages, bar_tma = bar_full + 16 * kMaxStages;
  const uint32_t bar_acc_full = bar_full + 24 * kMaxStages, bar_acc_empty = bar_acc_full + 16;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(ctl + 24 * kMaxStages + 32);
  // all-TMA pipeline: ns raw stages (+ ns lo buffers behind them with the 3xTF32 split); cp.async pipeline: kStages + 2 lo
  const int ns = kAllTma ? (split ? S::kNsSplit : S::kSlots) : kStages;
  constexpr uint32_t kTmemCols = kAllTma ? 512u : 2u * BN;   // two accumulators (+ a ring of 2 x (A raw | A lo), 32 + 32 columns each)

  if (tid == 0) {
    if (kTmaA) tma_prefetch_desc(&map_a);
    if (kTmaB) tma_prefetch_desc(&map_b);
    for (int s = 0; s < kMaxStages; ++s) {
      mbar_init(bar_full + 8 * s, kProducers / 32);   // one arrival per producer warp
      mbar_init(bar_empty + 8 * s, 1);
      mbar_init(bar_tma + 8 * s, 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(bar_acc_full + 8 * b, 1);
      mbar_init(bar_acc_empty + 8 * b, kEpiWarps);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kMmaWarp) tmem_alloc(smem_u32(tmem_ptr_smem), kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = *tmem_ptr_smem;
  const uint32_t tmem_a = tmem_d + 2u * BN;   // k-tile g: raw A in columns [64 (g&1), +32), its lo part in the next 32

  if (warp < kMmaWarp) {
    if constexpr (kAllTma) {
      // ---------------- sweepers (3xTF32 only).  Per k-tile the TMA warp brings in:  B -> lo = x - trunc(x) to the stage's
      // lo buffer (linear sweep);  A -> each thread reads 8 k of its row from the raw stage and stores them and their lo
      // parts to tensor memory (lane = row: warp w owns rows 32*(w%4)..+31 and the k octet w/4), from where the MMAs take
      // the A operand -- no lo copy of A in shared memory and no tensor-core reads of A from it.  They wait for the stage's
      // TMA barrier and for the MMAs of k-tile g - 2 (the previous user of lo ring slot g & 1), so they run at most two
      // k-tiles ahead of the tensor core while the TMA warp keeps all the other raw stages in flight.
      if (split) {
        int total_g = 0;   // k-tiles of all my units
        for (int64_t u = blockIdx.x; u < num_units; u += gridDim.x) total_g += unit_of(u, n_i, n_j, K, k_per_split).num_kt;
        const int row = (warp & 3) * 32 + lane, kq = warp >> 2;
        const uint32_t a_lane = ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)(kq * 8);
        uint32_t a_off[2];   // K-major: the two 16-byte chunks of k = 8 kq .. 8 kq + 7;  MN-major: offset of k = 8 kq (see below)
        if constexpr (AMODE == kKTma) { a_off[0] = k_offset(row, kq * 8); a_off[1] = k_offset(row, kq * 8 + 4); }
        else { a_off[0] = (uint32_t)(row >> 5) * Tile<BM>::kLboMn + (uint32_t)(kq * 2) * Tile<BM>::kSboMn + (uint32_t)(row & 31) * 4; a_off[1] = 0; }
        int s = 0, s2 = 0;   // stage of k-tile g, of k-tile g - 2
        uint32_t ph = 0, ph2 = 0;
#pragma unroll 1
        for (int g = 0; g < total_g; ++g) {
#ifndef B200LAB_GEMM_NOLOAD
          mbar_wait(bar_tma + 8 * s, ph);
#endif
          if (g >= 2) {   // lo ring slot g & 1 (B lo in shared memory, A raw | lo in tensor memory): the MMAs of k-tile g - 2 are done
            mbar_wait(bar_empty + 8 * s2, ph2);
            if (++s2 == ns) { s2 = 0; ph2 ^= 1u; }
          }
          tc_fence_after();
#ifndef B200LAB_GEMM_NOSWEEP
          const uint8_t* raw = smem + s * S::kStage;
          float v[8], l[8];
          if constexpr (AMODE == kKTma) {
            const float4 x0 = *reinterpret_cast<const float4*>(raw + a_off[0]), x1 = *reinterpret_cast<const float4*>(raw + a_off[1]);
            v[0] = x0.x; v[1] = x0.y; v[2] = x0.z; v[3] = x0.w; v[4] = x1.x; v[5] = x1.y; v[6] = x1.z; v[7] = x1.w;
          } else {
#pragma unroll
            for (int k = 0; k < 8; ++k)   // k-atom (k/4), line k%4 of it; 32-byte chunks of a line XORed with the line index
              v[k] = *reinterpret_cast<const float*>(raw + swz_mn(a_off[0] + (uint32_t)(k >> 2) * Tile<BM>::kSboMn + (uint32_t)(k & 3) * 128));
          }
#pragma unroll
          for (int k = 0; k < 8; ++k) l[k] = v[k] - __uint_as_float(__float_as_uint(v[k]) & 0xFFFFE000u);
          tmem_st8(tmem_a + (uint32_t)((g & 1) * 64) + a_lane, v);
          tmem_st8(tmem_a + (uint32_t)((g & 1) * 64 + 32) + a_lane, l);
          lo_sweep<S::kB>(raw + S::kA, smem + ns * S::kStage + (g & 1) * S::kB, tid);
          asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
#endif
          tc_fence_before();
          fence_proxy_async();            // st.shared (generic proxy) -> visible to the tensor core (async proxy)
          __syncwarp();                   // one arrival per warp
          if (lane == 0) mbar_arrive(bar_full + 8 * s);
          if (++s == ns) { s = 0; ph ^= 1u; }
        }
      }
    } else {
      // ---------------- producers (an operand that cannot go by TMA: unaligned, or TMA switched off) ----------------
      int total_g = 0;   // k-tiles of all my units
      for (int64_t u = blockIdx.x; u < num_units; u += gridDim.x) total_g += unit_of(u, n_i, n_j, K, k_per_split).num_kt;
      Cursor<BM, AMODE> ca;
      Cursor<BN, BMODE> cb;
      int64_t u_issue = blockIdx.x;
      int kt_issue = 0, nkt_issue = 0;
      auto open_unit = [&]() {
        const Unit w = unit_of(u_issue, n_i, n_j, K, k_per_split);
        nkt_issue = w.num_kt;
        kt_issue = 0;
        if constexpr (kTmaA) ca.init(&map_a, w.i0, w.k_begin, tid); else ca.init(A, lda_i, lda_k, w.i0, M, w.k_begin, w.k_end, tid);
        if constexpr (kTmaB) cb.init(&map_b, (int64_t)w.j0 * BN, w.k_begin, tid);
        else cb.init(B, ldb_j, ldb_k, (int64_t)w.j0 * BN, N, w.k_begin, w.k_end, tid);
      };
      if (u_issue < num_units) open_unit();
      auto issue = [&](int g_new) {   // loads of my g_new-th k-tile (callers keep g_new < total_g)
        if (kt_issue == nkt_issue) { u_issue += gridDim.x; open_unit(); }
        ++kt_issue;
        const int sn = g_new % kStages;
        const uint32_t st = smem_u32(smem + sn * S::kStage);
#ifdef B200LAB_GEMM_NOLOAD   // timing experiment: no operand traffic (results are garbage)
        return;
#endif
        if constexpr (kAnyTma) { if (tid == 0) mbar_arrive_expect_tx(bar_tma + 8 * sn, kTmaBytes); }
        if constexpr (kTmaA) ca.issue(st, bar_tma + 8 * sn); else ca.issue(st);
        if constexpr (kTmaB) cb.issue(st + S::kA, bar_tma + 8 * sn); else cb.issue(st + S::kA);
      };
#pragma unroll
      for (int p = 0; p < kStages - 1; ++p) {
        if (p < total_g) issue(p);
        if constexpr (!kAllTma) cp_async_commit();
      }
#pragma unroll 1
      for (int g = 0; g < total_g; ++g) {
        const int s = g % kStages;
        if constexpr (!kAllTma) cp_async_wait<kStages - 2>();   // this thread's copies of k-tile g have landed ...
        // ... the TMA boxes of the tile too (the mbarrier completes when all their bytes are in shared memory, for every waiter)
#ifndef B200LAB_GEMM_NOLOAD
        if constexpr (kAnyTma) mbar_wait(bar_tma + 8 * s, (uint32_t)((g / kStages) & 1));
#endif
        if (split) {
          // ... and, after the producers' barrier, everybody's cp.async copies: the lo sweep is linear over the stage, so a
          // thread reads bytes other threads copied (no barrier when both operands came by TMA).  Lo buffer g&1 was last
          // read by the MMAs of k-tile g-2 (waited for below, one iteration ago).
          if constexpr (!kAllTma) asm volatile("bar.sync 1, %0;" ::"n"(kProducers) : "memory");
#ifndef B200LAB_GEMM_NOSWEEP   // timing experiment: the MMAs read stale lo buffers
          lo_sweep<S::kStage>(smem + s * S::kStage, smem + S::kLo + (g & 1) * S::kStage, tid);
#endif
        }
        fence_proxy_async();            // cp.async / st.shared (generic proxy) -> visible to the tensor core (async proxy)
        __syncwarp();                   // one arrival per warp: 512 arrivals on one mbarrier word serialise in shared memory
        if (lane == 0) mbar_arrive(bar_full + 8 * s);
        // k-tile g-1's MMAs done: its raw stage can be refilled (with k-tile g+kStages-1) and, one iteration from now,
        // its lo buffer rewritten -- so this wait is unconditional
        if (g >= 1) mbar_wait(bar_empty + 8 * ((g - 1) % kStages), (uint32_t)(((g - 1) / kStages) & 1));
        if (g + kStages - 1 < total_g) issue(g + kStages - 1);
        if constexpr (!kAllTma) cp_async_commit();
      }
    }
  } else if (warp == kMmaWarp) {
    if (lane == 0) {
      // ---------------- MMA issuer ----------------
      constexpr bool a_mn = AMODE == kMnVec || AMODE == kMnScalar || AMODE == kMnTma, b_mn = BMODE == kMnVec || BMODE == kMnScalar || BMODE == kMnTma;
      constexpr uint32_t idesc = make_idesc(BM, BN) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16);
      int g = 0, n = 0, s = 0;
      uint32_t ph = 0;   // parity of stage s's current use
#pragma unroll 1
      for (int64_t u = blockIdx.x; u < num_units; u += gridDim.x, ++n) {
        const int num_kt = unit_of(u, n_i, n_j, K, k_per_split).num_kt;
        const int buf = n & 1, use = n >> 1;
        if (use >= 1) {   // the epilogue has read this accumulator's previous tile out of TMEM
          mbar_wait(bar_acc_empty + 8 * buf, (uint32_t)((use - 1) & 1));
          tc_fence_after();
        }
        const uint32_t acc = tmem_d + (uint32_t)(buf * BN);
#pragma unroll 1
        for (int kt = 0; kt < num_kt; ++kt, ++g) {
          const uint32_t sa = smem_u32(smem + s * S::kStage), sb = sa + S::kA;
          // lo part: behind the raw ring, one buffer per stage (all-TMA) or two alternating ones (cp.async pipeline)
          const uint32_t sa_lo = smem_u32(smem + (kStages + (g & 1)) * S::kStage);            // cp.async pipeline only
          const uint32_t sb_lo = kAllTma ? smem_u32(smem + ns * S::kStage + (g & 1) * S::kB) : sa_lo + S::kA;
          const uint32_t ta = tmem_a + (uint32_t)((g & 1) * 64);                                // all-TMA split only
#ifndef B200LAB_GEMM_NOLOAD
          if constexpr (kAllTma) mbar_wait(bar_tma + 8 * s, ph);     // raw operands: straight from the TMA's barrier
#endif
          if (!kAllTma || split) mbar_wait(bar_full + 8 * s, ph);     // cp.async copies landed / lo part written
          tc_fence_after();
#ifndef B200LAB_GEMM_NOMMA   // timing experiment: no tensor-core work (results are garbage)
#pragma unroll
#else
          if (g < 0)
#endif
          for (int ks = 0; ks < BK / 8; ++ks) {
            const uint64_t da = operand_desc<BM>(sa, ks, a_mn), db = operand_desc<BN>(sb, ks, b_mn);
            if (kAllTma && split) {   // A from tensor memory (always "K-major" there)
              constexpr uint32_t idesc_ta = make_idesc(BM, BN) | ((uint32_t)b_mn << 16);
              umma_tf32_ta(acc, ta + (uint32_t)(ks * 8), db, idesc_ta, (kt | ks) != 0);
              umma_tf32_ta(acc, ta + (uint32_t)(ks * 8), operand_desc<BN>(sb_lo, ks, b_mn), idesc_ta, 1u);
              umma_tf32_ta(acc, ta + 32u + (uint32_t)(ks * 8), db, idesc_ta, 1u);
            } else {
              umma_tf32(acc, da, db, idesc, (kt | ks) != 0);
              if (split) {
                umma_tf32(acc, da, operand_desc<BN>(sb_lo, ks, b_mn), idesc, 1u);
                umma_tf32(acc, operand_desc<BM>(sa_lo, ks, a_mn), db, idesc, 1u);
              }
            }
          }
          umma_commit(bar_empty + 8 * s);   // releases the stage (and its lo buffer) once these MMAs have read it
          if (++s == ns) { s = 0; ph ^= 1u; }
        }
        umma_commit(bar_acc_full + 8 * buf);   // accumulator complete
      }
    }
  } else if (warp == kMmaWarp + kEpiWarps + 1) {
    // ---------------- TMA warp (all-TMA pipeline): refills stage s as soon as the MMAs of its previous k-tile are done --
    // nobody else is in that loop, so ns - 1 k-tiles stay in flight whatever the sweepers and the MMA warp are doing
    if constexpr (kAllTma) {
      if (lane == 0) {
        Cursor<BM, AMODE> ca;
        Cursor<BN, BMODE> cb;
        int s = 0;
        uint32_t ph = 0;
        bool wrapped = false;
#pragma unroll 1
        for (int64_t u = blockIdx.x; u < num_units; u += gridDim.x) {
          const Unit w = unit_of(u, n_i, n_j, K, k_per_split);
          ca.init(&map_a, w.i0, w.k_begin, 0);
          cb.init(&map_b, (int64_t)w.j0 * BN, w.k_begin, 0);
#pragma unroll 1
          for (int kt = 0; kt < w.num_kt; ++kt) {
            if (wrapped) mbar_wait(bar_empty + 8 * s, ph ^ 1u);   // the stage's previous use has been multiplied
#ifndef B200LAB_GEMM_NOLOAD
            const uint32_t st = smem_u32(smem + s * S::kStage);
            mbar_arrive_expect_tx(bar_tma + 8 * s, kTmaBytes);
            ca.issue(st, bar_tma + 8 * s);
            cb.issue(st + S::kA, bar_tma + 8 * s);
#endif
            if (++s == ns) { s = 0; ph ^= 1u; wrapped = true; }
          }
        }
      }
    }
  } else if (warp > kMmaWarp) {
    // ---------------- epilogue: TMEM -> registers -> (32 x 32 transpose through shared memory) -> global.  A warp may read
    // the TMEM lanes 32*(warp%4)..+31 (= tile rows, one per lane); it walks the BN columns in groups of 32.  Written
    // straight from the TMEM layout every store instruction would touch 32 rows (32 sectors); after the transpose it
    // writes 128 contiguous bytes of one row.
    const int q = warp & 3;
    float* tile = reinterpret_cast<float*>(smem + S::kEpi) + (warp - kMmaWarp - 1) * (32 * kEpiLd);
    const bool elu = epi == 2;
    // 128-bit path: a lane re-reads 4 columns of one row, so a store instruction writes 4 rows x 128 contiguous bytes
    const bool vec = ldc % 4 == 0 && N % 4 == 0 && (reinterpret_cast<uintptr_t>(C) & 15u) == 0;
    int n = 0;
#pragma unroll 1
    for (int64_t u = blockIdx.x; u < num_units; u += gridDim.x, ++n) {
      const Unit w = unit_of(u, n_i, n_j, K, k_per_split);
      const int buf = n & 1, use = n >> 1;
      const int j0 = w.j0 * BN;
      float* Cz = C + (int64_t)w.z * M * ldc;
      float4 bvs[BN / 32];   // the tile's bias values, fetched while the MMAs are still running
#pragma unroll
      for (int c = 0; c < BN / 32; ++c) {
        const int gj = j0 + c * 32 + (lane & 7) * 4;
        bvs[c] = (vec && epi >= 1 && gj < N) ? make_float4(__ldg(bias + gj), __ldg(bias + gj + 1), __ldg(bias + gj + 2), __ldg(bias + gj + 3))
                                             : make_float4(0.f, 0.f, 0.f, 0.f);
      }
      mbar_wait(bar_acc_full + 8 * buf, (uint32_t)(use & 1));
      tc_fence_after();
      const uint32_t acc = tmem_d + (uint32_t)(buf * BN) + ((uint32_t)(q * 32) << 16);
#pragma unroll
      for (int c0 = 0; c0 < BN; c0 += 32) {
        const bool last = c0 + 32 == BN;
        if (j0 + c0 < N) {   // warp-uniform
          float v32[32];
          tmem_ld32(acc + (uint32_t)c0, v32);
          if (last) {   // every column of this warp's quarter is in registers: hand the accumulator back
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_acc_empty + 8 * buf);
          }
          float4* trow = reinterpret_cast<float4*>(tile + lane * kEpiLd);   // row stride 144 B: conflict-free 128-bit accesses
#pragma unroll
          for (int i = 0; i < 8; ++i) trow[i] = make_float4(v32[4 * i], v32[4 * i + 1], v32[4 * i + 2], v32[4 * i + 3]);
          __syncwarp();
          if (vec) {
            const int cg = (lane & 7) * 4, gj = j0 + c0 + cg;
            const bool col_ok = gj < N;   // N % 4 == 0: the whole vector is inside
            const float4 bv = bvs[c0 / 32];
#pragma unroll
            for (int rr = 0; rr < 8; ++rr) {
              const int r = rr * 4 + (lane >> 3);
              const int64_t gi = w.i0 + q * 32 + r;
              float4 x = *reinterpret_cast<const float4*>(tile + r * kEpiLd + cg);
              x.x += bv.x; x.y += bv.y; x.z += bv.z; x.w += bv.w;
              if (elu) {   // |abs err| ~1e-7
                x.x = x.x > 0.0f ? x.x : exp2f_fast(1.4426950408889634f * x.x) - 1.0f;
                x.y = x.y > 0.0f ? x.y : exp2f_fast(1.4426950408889634f * x.y) - 1.0f;
                x.z = x.z > 0.0f ? x.z : exp2f_fast(1.4426950408889634f * x.z) - 1.0f;
                x.w = x.w > 0.0f ? x.w : exp2f_fast(1.4426950408889634f * x.w) - 1.0f;
              }
#ifndef B200LAB_GEMM_NOSTORE
              if (col_ok && gi < M) *reinterpret_cast<float4*>(Cz + gi * ldc + gj) = x;
#endif
            }
          } else {
            const int gj = j0 + c0 + lane;
            const bool col_ok = gj < N;
            float bv = 0.f;
            if (epi >= 1 && col_ok) bv = __ldg(bias + gj);
#pragma unroll 8
            for (int rr = 0; rr < 32; ++rr) {
              const int64_t gi = w.i0 + q * 32 + rr;
              if (gi >= M) break;      // warp-uniform
              float x = tile[rr * kEpiLd + lane] + bv;
              if (elu) x = x > 0.0f ? x : exp2f_fast(1.4426950408889634f * x) - 1.0f;
              if (col_ok) Cz[gi * ldc + gj] = x;
            }
          }
          __syncwarp();   // the tile is rewritten by the next column group
        } else if (last) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_acc_empty + 8 * buf);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kMmaWarp) tmem_dealloc(tmem_d, kTmemCols);
}

// cuTensorMapEncodeTiled, resolved through the runtime (no link-time dependency on libcuda)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
    (void)cudaGetLastError();
  }
  return fn;
}

// Tensor map of a K-contiguous fp32 operand [rows, K] with row stride ld_r (elements): boxes of box_rows x BK, written to
// shared memory in the K-major SWIZZLE_128B layout the MMA descriptors expect; out-of-range elements arrive as zeros.
static bool make_k_major_map(CUtensorMap* map, const float* base, int64_t rows, int64_t K, int64_t ld_r, int box_rows) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return false;
  const cuuint64_t gdim[2] = {(cuuint64_t)K, (cuuint64_t)rows};
  const cuuint64_t gstride[1] = {(cuuint64_t)ld_r * 4u};
  const cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1u, 1u};
  return fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)base, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// Tensor map of an M/N-contiguous fp32 operand (element (r, k) at base + r + k*ld_k): boxes of 32 rows x BK, written in the
// MN-major SWIZZLE_128B_BASE32B layout (32-byte chunks of a 128-byte line XORed with the line index mod 4).
static bool make_mn_major_map(CUtensorMap* map, const float* base, int64_t rows, int64_t K, int64_t ld_k) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return false;
  const cuuint64_t gdim[2] = {(cuuint64_t)rows, (cuuint64_t)K};
  const cuuint64_t gstride[1] = {(cuuint64_t)ld_k * 4u};
  const cuuint32_t box[2] = {32u, (cuuint32_t)BK};
  const cuuint32_t estr[2] = {1u, 1u};
  return fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)base, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <int BN, int AMODE, int BMODE>
static int launch(const float* A, int64_t lda_i, int64_t lda_k, const float* B, int64_t ldb_k, int64_t ldb_j, float* C,
                  int64_t ldc, int64_t M, int N, int64_t K, int64_t k_per_split, int splits, const float* bias, int epi,
                  int split, cudaStream_t s) {
  auto kern = gemm_tf32_kernel<BN, AMODE, BMODE>;
  alignas(64) CUtensorMap map_a, map_b;
  memset(&map_a, 0, sizeof(map_a));
  memset(&map_b, 0, sizeof(map_b));
  if (AMODE == kKTma && !make_k_major_map(&map_a, A, M, K, lda_i, BM)) { set_error("tcgen05 gemm: cuTensorMapEncodeTiled (A) failed"); return B200LAB_ERR_CUDA; }
  if (BMODE == kKTma && !make_k_major_map(&map_b, B, N, K, ldb_j, BN)) { set_error("tcgen05 gemm: cuTensorMapEncodeTiled (B) failed"); return B200LAB_ERR_CUDA; }
  if (AMODE == kMnTma && !make_mn_major_map(&map_a, A, M, K, lda_k)) { set_error("tcgen05 gemm: cuTensorMapEncodeTiled (A, MN-major) failed"); return B200LAB_ERR_CUDA; }
  if (BMODE == kMnTma && !make_mn_major_map(&map_b, B, N, K, ldb_k)) { set_error("tcgen05 gemm: cuTensorMapEncodeTiled (B, MN-major) failed"); return B200LAB_ERR_CUDA; }
  static bool configured = false;
  if (!configured) {
    cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Smem<BN>::kTotal);
    if (err != cudaSuccess) { set_error("tcgen05 gemm: %s", cudaGetErrorString(err)); return B200LAB_ERR_CUDA; }
    configured = true;
  }
  static const int sm_count = [] {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = kNumSMs;
    return n;
  }();
  const int64_t units = ceil_div(M, BM) * ceil_div(N, BN) * splits;
  const unsigned grid = (unsigned)(units < sm_count ? units : sm_count);   // one persistent CTA per SM
  kern<<<grid, kThreads, Smem<BN>::kTotal, s>>>(A, lda_i, lda_k, B, ldb_k, ldb_j, C, ldc, M, N, K, k_per_split, splits, bias, epi,
                                                split, map_a, map_b);
  count_launch();
  return B200LAB_OK;
}

template <int BN, int AMODE>
static int launch_b(int b_mode, const float* A, int64_t lda_i, int64_t lda_k, const float* B, int64_t ldb_k, int64_t ldb_j,
                    float* C, int64_t ldc, int64_t M, int N, int64_t K, int64_t k_per_split, int splits, const float* bias,
                    int epi, int split, cudaStream_t s) {
#define B200LAB_TC_B(MODE) \
  case MODE: return launch<BN, AMODE, MODE>(A, lda_i, lda_k, B, ldb_k, ldb_j, C, ldc, M, N, K, k_per_split, splits, bias, epi, split, s)
  switch (b_mode) {
    B200LAB_TC_B(kKVec); B200LAB_TC_B(kKScalar); B200LAB_TC_B(kMnVec); B200LAB_TC_B(kKTma); B200LAB_TC_B(kMnTma);
    default: return launch<BN, AMODE, kMnScalar>(A, lda_i, lda_k, B, ldb_k, ldb_j, C, ldc, M, N, K, k_per_split, splits, bias, epi, split, s);
  }
#undef B200LAB_TC_B
}

template <int BN>
static int launch_a(int a_mode, int b_mode, const float* A, int64_t lda_i, int64_t lda_k, const float* B, int64_t ldb_k,
                    int64_t ldb_j, float* C, int64_t ldc, int64_t M, int N, int64_t K, int64_t k_per_split, int splits,
                    const float* bias, int epi, int split, cudaStream_t s) {
#define B200LAB_TC_A(MODE) \
  case MODE: return launch_b<BN, MODE>(b_mode, A, lda_i, lda_k, B, ldb_k, ldb_j, C, ldc, M, N, K, k_per_split, splits, bias, epi, split, s)
  switch (a_mode) {
    B200LAB_TC_A(kKVec); B200LAB_TC_A(kKScalar); B200LAB_TC_A(kMnVec); B200LAB_TC_A(kKTma); B200LAB_TC_A(kMnTma);
    default: return launch_b<BN, kMnScalar>(b_mode, A, lda_i, lda_k, B, ldb_k, ldb_j, C, ldc, M, N, K, k_per_split, splits, bias, epi, split, s);
  }
#undef B200LAB_TC_A
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
  // 16-byte copies along whichever dimension is contiguous, when every vector is 16-byte aligned: base pointer, leading
  // dimension and (K-major) the split boundaries multiples of 4 elements; tile origins are multiples of 64 already
  auto mode_of = [&](const float* p, int64_t ld_r, int64_t ld_k) -> int {
    const bool base_ok = (uintptr_t)p % 16 == 0;
    if (ld_k == 1) return (base_ok && ld_r % 4 == 0 && (splits == 1 || k_per_split % 4 == 0)) ? tc::kKVec : tc::kKScalar;
    if (ld_r == 1) return (base_ok && ld_k % 4 == 0) ? tc::kMnVec : tc::kMnScalar;
    return ld_k < ld_r ? tc::kKScalar : tc::kMnScalar;
  };
  int a_mode = mode_of(A, lda_i, lda_k), b_mode = mode_of(B, ldb_j, ldb_k);
  // K-contiguous, 16-byte aligned operands go by TMA (B200LAB_GEMM_TMA=0 keeps the cp.async path, for A/B measurements);
  // a K split must fall on k-tile boundaries, since a box reads BK columns whatever the split's end
  static const bool use_tma = [] { const char* v = getenv("B200LAB_GEMM_TMA"); return !(v && v[0] == '0'); }() && tc::encode_tiled_fn() != nullptr;
  if (use_tma && (splits == 1 || k_per_split % tc::BK == 0) && K < (1ll << 31) && M < (1ll << 31)) {
    if (a_mode == tc::kKVec) a_mode = tc::kKTma;
    if (b_mode == tc::kKVec) b_mode = tc::kKTma;
    static const bool mn_tma = [] { const char* v = getenv("B200LAB_GEMM_MN_TMA"); return !(v && v[0] == '0'); }();
    if (mn_tma && a_mode == tc::kMnVec) a_mode = tc::kMnTma;
    if (mn_tma && b_mode == tc::kMnVec && N < (1 << 30)) b_mode = tc::kMnTma;
  }
  const int split = precision == 2 ? 1 : 0;
  if (N > 64) return tc::launch_a<128>(a_mode, b_mode, A, lda_i, lda_k, B, ldb_k, ldb_j, C, ldc, M, N, K, k_per_split, splits, bias, epi, split, s);
  return tc::launch_a<64>(a_mode, b_mode, A, lda_i, lda_k, B, ldb_k, ldb_j, C, ldc, M, N, K, k_per_split, splits, bias, epi, split, s);
}

}  // namespace b200lab
