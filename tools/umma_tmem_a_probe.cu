// umma_tmem_a_probe.cu -- does tcgen05.mma (kind::tf32, cta_group::1, M=128) take its A operand from tensor memory as
// "lane = row m, column = k (one 32-bit column per tf32 element)"?  A(m, k) = +-(32 (m % 64) + k + 1) is written with tcgen05.st by the
// thread that owns lane m; B is an identity selector in shared memory (K-major SWIZZLE_NONE, as tools/umma_probe.cu), so
// D[m][n < 8] must read back A(m, n).  Also times nothing: a layout probe only.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o tools/_bin/umma_tmem_a_probe tools/umma_tmem_a_probe.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <vector>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
  return (uint64_t)((saddr >> 4) & 0x3FFFu) | ((uint64_t)((lbo >> 4) & 0x3FFFu) << 16) |
         ((uint64_t)((sbo >> 4) & 0x3FFFu) << 32) | (1ull << 46) | ((uint64_t)layout << 61);
}

// ksteps MMAs of K = 8; A columns [64, 64 + 8*ksteps); D = 128 x 64 in columns [0, 64)
__global__ void __launch_bounds__(128) probe_kernel(int ksteps, float* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  float* sel = reinterpret_cast<float*>(smem);   // per k-step: 64 rows x 8 k, K-major SWIZZLE_NONE, 2048 B each
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 16384);
  uint32_t* tptr = reinterpret_cast<uint32_t*>(smem + 16384 + 16);
  const int tid = threadIdx.x, warp = tid >> 5;
  const uint32_t sel_lbo = 64 * 16;
  for (int i = tid; i < 4096; i += 128) sel[i] = 0.f;
  __syncthreads();
  // k-step ks selects n = 8*ks + k:  B_ks(n, k) = (n == 8*ks + k)
  for (int ks = 0; ks < ksteps; ++ks)
    for (int i = tid; i < 64 * 8; i += 128) {
      int r = i / 8, k = i % 8;
      sel[ks * 512 + ((k >> 2) * sel_lbo + r * 16 + (k & 3) * 4) / 4] = (r == 8 * ks + k) ? 1.0f : 0.0f;
    }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 128;" ::"r"(smem_u32(tptr)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *tptr;
  // A(m = tid, k) for k in 0 .. 8*ksteps-1: 8 columns per store
  for (int ks = 0; ks < ksteps; ++ks) {
    uint32_t v[8];
    for (int k = 0; k < 8; ++k) v[k] = __float_as_uint((tid >= 64 ? -1.f : 1.f) * (float)((tid % 64) * 32 + ks * 8 + k + 1));   // exact in tf32
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(
                     tmem + ((uint32_t)(warp * 32) << 16) + 64u + (uint32_t)(ks * 8)),
                 "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
                 : "memory");
  }
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  if (tid == 0) {
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    for (int ks = 0; ks < ksteps; ++ks) {
      const uint64_t db = make_desc(smem_u32(sel) + ks * 2048, sel_lbo, 128, 0);
      asm volatile(
          "{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %4, 0;\n\t"
          "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, q;\n\t}" ::"r"(tmem), "r"(tmem + 64u + (uint32_t)(ks * 8)), "l"(db),
          "r"(idesc), "r"(ks)
          : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
  }
  asm volatile(
      "{\n\t.reg .pred P1;\n\tW:\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], 0;\n\t@P1 bra D;\n\tbra W;\n\tD:\n\t}" ::"r"(
          smem_u32(bar))
      : "memory");
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
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
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 128;" ::"r"(tmem) : "memory");
}

int main() {
  cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384 + 64);
  float* d_out;
  cudaMalloc(&d_out, 128 * 64 * 4);
  for (int ksteps = 1; ksteps <= 4; ksteps += 3) {
    cudaMemset(d_out, 0xFF, 128 * 64 * 4);
    probe_kernel<<<1, 128, 16384 + 64>>>(ksteps, d_out);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
    std::vector<float> h(128 * 64);
    cudaMemcpy(h.data(), d_out, h.size() * 4, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int m = 0; m < 128; ++m)
      for (int n = 0; n < 64; ++n) {
        const float want = n < 8 * ksteps ? (m >= 64 ? -1.f : 1.f) * (float)((m % 64) * 32 + n + 1) : 0.f;
        if (h[m * 64 + n] != want) ++bad;
      }
    printf("ksteps=%d: %d mismatches of %d\n", ksteps, bad, 128 * 64);
    for (int m : {0, 1, 5, 33, 127}) {
      printf("  D[%d][0..11]:", m);
      for (int n = 0; n < 12; ++n) printf(" %g", h[m * 64 + n]);
      printf("\n");
    }
  }
  return 0;
}
