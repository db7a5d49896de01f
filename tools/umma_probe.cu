// umma_probe.cu -- reads out how tcgen05.mma (kind::tf32, cta_group::1, M=128) maps shared memory to operand elements for
// a given shared-memory descriptor, and how it treats the low 13 mantissa bits of fp32 inputs.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o tools/_bin/umma_probe tools/umma_probe.cu && tools/_bin/umma_probe
// Method: smem word w of the probed operand holds the value (float)w; the other operand is an identity selector in the
// known-good K-major SWIZZLE_NONE layout, so D[m][n] = probed(m, k = n) = index of the word the hardware read.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>
#include <vector>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
  return (uint64_t)((saddr >> 4) & 0x3FFFu) | ((uint64_t)((lbo >> 4) & 0x3FFFu) << 16) |
         ((uint64_t)((sbo >> 4) & 0x3FFFu) << 32) | (1ull << 46) | ((uint64_t)layout << 61);
}

struct Probe {
  int which;        // 0: probe A (B = selector), 1: probe B (A = selector)
  int major;        // major bit of the probed operand in the instruction descriptor (0 K-major, 1 MN-major)
  uint32_t lbo, sbo, layout;
  int pre;          // 1: first run a known-good K-major MMA (D = selector pattern), then the probed MMA with accumulate=1
  int raw_bits;     // 1: fill the probed operand with bit patterns for the truncation test instead of word indices
};

// D is 128 x 64 fp32; N = 64, one MMA K-step (8 tf32).
__global__ void __launch_bounds__(128) probe_kernel(Probe p, float* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  float* probed = reinterpret_cast<float*>(smem);              // 32 KB region, word w = (float)w
  float* sel = reinterpret_cast<float*>(smem + 32768);         // selector operand, K-major SWIZZLE_NONE
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 32768 + 8192);
  uint32_t* tptr = reinterpret_cast<uint32_t*>(smem + 32768 + 8192 + 16);
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int w = tid; w < 8192; w += 128) {
    float v = (float)(w + 1);
    if (p.raw_bits) v = __uint_as_float(0x3F800000u | (uint32_t)(w * 2654435761u >> 9));   // [1,2) with random mantissa
    probed[w] = v;
  }
  // selector: rows r (128 for A-role, 64 for B-role), k in 0..7, value (r == k); layout off = (k/4)*LBOk + r*16 + (k%4)*4
  const int sel_rows = p.which == 0 ? 64 : 128;
  const uint32_t sel_lbo = sel_rows * 16;
  for (int i = tid; i < 2048; i += 128) sel[i] = 0.f;
  __syncthreads();
  for (int i = tid; i < sel_rows * 8; i += 128) {
    int r = i / 8, k = i % 8;
    sel[((k >> 2) * sel_lbo + r * 16 + (k & 3) * 4) / 4] = (r == k) ? 1.0f : 0.0f;
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(smem_u32(tptr)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *tptr;
  if (tid == 0) {
    uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    uint64_t d_probe = make_desc(smem_u32(probed), p.lbo, p.sbo, p.layout);
    uint64_t d_sel = make_desc(smem_u32(sel), sel_lbo, 128, 0);
    uint64_t da, db;
    if (p.which == 0) { da = d_probe; db = d_sel; idesc |= (uint32_t)p.major << 15; }
    else { da = d_sel; db = d_probe; idesc |= (uint32_t)p.major << 16; }
    if (p.pre) {   // D = 1000 * [m == n] via selector x selector, K-major
      uint32_t idk = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
      uint64_t dk = make_desc(smem_u32(probed), 128 * 16 + 16, 128, 0);
      asm volatile(
          "{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %4, 0;\n\t"
          "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, q;\n\t}" ::"r"(tmem), "l"(dk), "l"(d_sel), "r"(idk), "r"(0)
          : "memory");
    }
    asm volatile(
        "{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, q;\n\t}" ::"r"(tmem), "l"(da), "l"(db), "r"(idesc), "r"(p.pre)
        : "memory");
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
  }
  asm volatile(
      "{\n\t.reg .pred P1;\n\tW:\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], 0;\n\t@P1 bra D;\n\tbra W;\n\tD:\n\t}" ::"r"(
          smem_u32(bar))
      : "memory");
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  // lane = row (warp w reads TMEM lanes 32w..32w+31), 64 columns
  for (int c0 = 0; c0 < 64; c0 += 16) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int i = 0; i < 16; ++i) out[tid * 64 + c0 + i] = __uint_as_float(r[i]);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" ::"r"(tmem) : "memory");
}

static std::vector<float> run(const Probe& p, float* d_out) {
  cudaMemset(d_out, 0xFF, 128 * 64 * 4);
  probe_kernel<<<1, 128, 32768 + 8192 + 64>>>(p, d_out);
  cudaError_t e = cudaDeviceSynchronize();
  std::vector<float> h(128 * 64);
  if (e != cudaSuccess) { printf("  CUDA error: %s\n", cudaGetErrorString(e)); exit(1); }
  cudaMemcpy(h.data(), d_out, h.size() * 4, cudaMemcpyDeviceToHost);
  return h;
}

// A-probe: D[m][n<8] = word index of A(m, k=n).  B-probe: D[m<8][n] = word index of B(n, k=m).
static void describe(const Probe& p, const std::vector<float>& h) {
  auto at = [&](int mn, int k) { return p.which == 0 ? h[mn * 64 + k] : h[k * 64 + mn]; };
  const int rows = p.which == 0 ? 128 : 64;
  printf("probe %s major=%d lbo=%u sbo=%u layout=%u:\n", p.which == 0 ? "A" : "B", p.major, p.lbo, p.sbo, p.layout);
  printf("  word(mn=0..9, k=0): ");
  for (int m = 0; m < 10; ++m) printf("%g ", at(m, 0));
  printf("\n  word(mn=0, k=0..7): ");
  for (int k = 0; k < 8; ++k) printf("%g ", at(0, k));
  printf("\n  word(mn=8,16,32,64; k=0): %g %g %g %g", at(8, 0), at(16, 0), at(32, 0), rows > 64 ? at(64, 0) : -1.f);
  printf("\n  word(mn=5, k=0..7): ");
  for (int k = 0; k < 8; ++k) printf("%g ", at(5, k));
  printf("\n  raw D[0][0..11]: ");
  for (int n = 0; n < 12; ++n) printf("%g ", h[n]);
  printf("\n  raw D[1][0..11]: ");
  for (int n = 0; n < 12; ++n) printf("%g ", h[64 + n]);
  printf("\n");
}

int main() {
  cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768 + 8192 + 64);
  float* d_out;
  cudaMalloc(&d_out, 128 * 64 * 4);
  // 1. sanity: K-major SWIZZLE_NONE, the layout in production use (expect word = (k/4)*LBO/4 + mn*4 + k%4)
  Probe p{0, 0, 128 * 16 + 16, 128, 0, 0, 0};
  describe(p, run(p, d_out));
  // 2. MN-major SWIZZLE_NONE with distinguishable LBO / SBO
  for (int which = 0; which < 2; ++which) {
    const uint32_t combos[][2] = {{2048, 128}, {128, 2048}, {4096, 256}, {256, 4096}, {1024, 512}, {512, 1024}};
    for (auto& c : combos) {
      Probe q{which, 1, c[0], c[1], 0, which, 0};
      describe(q, run(q, d_out));
    }
  }
  // 3. K-major SWIZZLE_128B (layout type 2), rows of 128 B, 8-row groups 1024 B apart
  for (int which = 0; which < 2; ++which) {
    Probe q{which, 0, 16, 1024, 2, 0, 0};
    auto h = run(q, d_out);
    describe(q, h);
    printf("  full 8x8 (mn, k) word map:\n");
    for (int m = 0; m < 8; ++m) {
      printf("   ");
      for (int k = 0; k < 8; ++k) printf("%5g ", which == 0 ? h[m * 64 + k] : h[k * 64 + m]);
      printf("\n");
    }
  }
  // 4. MN-major SWIZZLE_128B
  for (int which = 0; which < 2; ++which) {
    const uint32_t combos[][2] = {{4096, 1024}, {1024, 4096}};
    for (auto& c : combos) {
      Probe q{which, 1, c[0], c[1], 2, 0, 0};
      auto h = run(q, d_out);
      describe(q, h);
    }
  }
  // 4b. MN-major SWIZZLE_128B_BASE32B (layout type 1): the only MN-major layout for 32-bit operands
  for (int which = 0; which < 2; ++which) {
    const uint32_t combos[][2] = {{4096, 512}, {512, 4096}, {8192, 1024}, {1024, 8192}};
    for (auto& c : combos) {
      Probe q{which, 1, c[0], c[1], 1, 0, 0};
      auto h = run(q, d_out);
      describe(q, h);
      printf("  (mn 0..39 step 1, k 0..7) word map (minus 1):\n");
      for (int k = 0; k < 8; ++k) {
        printf("   k=%d:", k);
        for (int m = 0; m < 40; ++m) printf(" %g", (which == 0 ? h[m * 64 + k] : h[k * 64 + m]) - 1);
        printf("\n");
      }
    }
  }
  // 5. low mantissa bits: A(m,k) raw fp32 in [1,2) with random low bits, B selects -> D = what the tensor core saw
  {
    Probe q{0, 0, 128 * 16 + 16, 128, 0, 0, 1};
    auto h = run(q, d_out);
    int trunc = 0, rn = 0, other = 0;
    for (int m = 0; m < 128; ++m)
      for (int k = 0; k < 8; ++k) {
        uint32_t w = (uint32_t)((k >> 2) * (128 * 16 + 16) / 4 + m * 4 + (k & 3));
        uint32_t bits = 0x3F800000u | (uint32_t)(w * 2654435761u >> 9);
        uint32_t t = bits & 0xFFFFE000u, r = (bits + 0x1000u) & 0xFFFFE000u;
        uint32_t got;
        float g = h[m * 64 + k];
        memcpy(&got, &g, 4);
        if (got == t && t != r) ++trunc; else if (got == r && t != r) ++rn; else if (got != t) ++other;
      }
    printf("low-bit handling of raw fp32 operands: truncated=%d rounded_nearest=%d other=%d\n", trunc, rn, other);
  }
  return 0;
}
