// Micro-benchmark: issue rate of tcgen05.mma kind::f16 (M=128, N, K=16, bf16 -> fp32) with shared-memory operands in
// different layouts - is the DAC conv kernel limited by how the tensor core fetches its no-swizzle "panel" operands?
//   ./umma_rate            prints cycles per MMA and TFLOP/s (bf16) for every configuration
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o umma_rate umma_rate.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void umma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(d),
      "l"(a), "l"(b), "r"(idesc), "r"(acc)
      : "memory");
}

struct Cfg {
  int layout;      // 0 = no swizzle (panels), 2 = SWIZZLE_128B
  int lbo_a;       // bytes (no swizzle): stride between the two K halves
  int lbo_b;
  int shift_rows;  // start-address shift of A in rows
  int N;
  int mmas_per_batch;  // distinct descriptors cycled through
  int a_reuse;     // 1: 3-product pattern (am.bh, ah.bm, ah.bh), 0: all the same
};

__global__ void __launch_bounds__(128, 1) rate_kernel(Cfg c, int reps, unsigned long long* out) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t slot;
  const uint32_t base = (smem_u32(smem) + 1023u) & ~1023u;
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(&slot)), "r"(256)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
  }
  if (threadIdx.x == 32) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_u32(&bar)), "r"(1) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(smem + (base - smem_u32(smem)))[i] = 0;
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
  const uint32_t tmem = slot;
  const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(c.N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  if (threadIdx.x == 0) {
    // A region: 64 KB at base (hi plane at 0, mid plane at 32 KB); B region: 64 KB at base + 64 KB
    const uint32_t a0 = base + (c.layout == 2 ? c.shift_rows * 128 : c.shift_rows * 16), a1 = a0 + 32768;
    const uint32_t b0 = base + 65536, b1 = b0 + 32768;
    auto desc = [&](uint32_t addr, int lbo) -> uint64_t {
      if (c.layout == 2) {
        const uint64_t bo = (addr >> 7) & 7;  // base offset: start not aligned to the 1024-byte swizzle atom
        return (uint64_t)((addr >> 4) & 0x3FFF) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (bo << 49) | (2ull << 61);
      }
      return (uint64_t)((addr >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) | (8ull << 32) | (1ull << 46);
    };
    const long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
#pragma unroll
      for (int ks = 0; ks < 2; ++ks) {
        const uint32_t ka = c.layout == 2 ? ks * 32 : 2 * ks * c.lbo_a, kb = c.layout == 2 ? ks * 32 : 2 * ks * c.lbo_b;
        const uint64_t ah = desc(a0 + ka, c.lbo_a), am = desc(a1 + ka, c.lbo_a);
        const uint64_t bh = desc(b0 + kb, c.lbo_b), bm = desc(b1 + kb, c.lbo_b);
        if (c.a_reuse) {
          umma(tmem, am, bh, idesc, 1);
          umma(tmem, ah, bm, idesc, 1);
          umma(tmem, ah, bh, idesc, 1);
        } else {
          umma(tmem, ah, bh, idesc, 1);
          umma(tmem, ah, bh, idesc, 1);
          umma(tmem, ah, bh, idesc, 1);
        }
      }
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(smem_u32(&bar))
                 : "memory");
    uint32_t done = 0;
    for (uint32_t spin = 0; !done && spin < 400000000u; ++spin)
      asm volatile(
          "{\n\t.reg .pred P1;\n\tmbarrier.test_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.b32 %0, 1, 0, P1;\n\t}\n"
          : "=r"(done)
          : "r"(smem_u32(&bar)), "r"(0)
          : "memory");
    const long long t1 = clock64();
    out[blockIdx.x] = (unsigned long long)(t1 - t0);
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 32)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem), "r"(256) : "memory");
}

int main() {
  unsigned long long* d;
  cudaMalloc(&d, 148 * 8);
  cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const Cfg cfgs[] = {
      {0, 2912, 2048, 0, 128, 6, 1}, {0, 2912, 2048, 3, 128, 6, 1}, {0, 2048, 2048, 0, 128, 6, 1},
      {0, 2048, 2048, 0, 128, 6, 0}, {2, 0, 0, 0, 128, 6, 1},       {2, 0, 0, 3, 128, 6, 1},
      {2, 0, 0, 0, 128, 6, 0},       {0, 2912, 4096, 0, 256, 6, 1}, {2, 0, 0, 0, 256, 6, 1},
      {2, 0, 0, 3, 256, 6, 1},       {0, 2912, 1536, 0, 96, 6, 1},  {2, 0, 0, 0, 96, 6, 1},
      {0, 2912, 3072, 0, 192, 6, 1}, {2, 0, 0, 3, 192, 6, 1},
  };
  const int reps = 2000;
  for (const Cfg& c : cfgs) {
    for (int grid : {1, 148}) {
      rate_kernel<<<grid, 128, 200 * 1024>>>(c, reps, d);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) {
        printf("error: %s\n", cudaGetErrorString(e));
        return 1;
      }
      unsigned long long h[148];
      cudaMemcpy(h, d, grid * 8, cudaMemcpyDeviceToHost);
      unsigned long long mx = 0;
      for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
      const double cyc = (double)mx / (reps * 6.0);
      printf("layout %s lboA %4d lboB %4d shift %d N %3d pattern %s grid %3d : %7.1f cycles / MMA  (%6.1f flop/cycle/SM)\n",
             c.layout == 2 ? "sw128" : "none ", c.lbo_a, c.lbo_b, c.shift_rows, c.N, c.a_reuse ? "3-product" : "same     ",
             grid, cyc, 2.0 * 128 * c.N * 16 / cyc);
    }
  }
  return 0;
}
