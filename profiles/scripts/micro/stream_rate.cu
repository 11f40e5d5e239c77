// Micro-benchmark: what read bandwidth can a flash-decode-shaped grid reach at all?  1024 CTAs x 128 threads, each CTA
// streams two 60 KB regions ("K" then "V" of one (b, head)) with 16-byte ld.global.cg loads, U loads in flight per
// thread, and only XORs the data.  24 distinct 124 MB caches are cycled so nothing is served from L2.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o stream_rate stream_rate.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

template <int U, bool TWO_STREAMS>
__global__ void __launch_bounds__(128) stream_kernel(const uint4* __restrict__ k, const uint4* __restrict__ v, int rows,
                                                     int c_stride, uint32_t* sink) {
  const int bh = blockIdx.x;
  const uint4* kb = k + (size_t)bh * c_stride * 8;  // 8 uint4 per 128-byte row
  const uint4* vb = v + (size_t)bh * c_stride * 8;
  const int n = rows * 8;
  uint32_t x = 0;
  if (TWO_STREAMS) {  // K and V interleaved (one-pass shape)
    for (int i = threadIdx.x; i < n; i += 128 * U) {
      uint4 a[U], b[U];
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (i + u * 128 < n) {
          asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(a[u].x), "=r"(a[u].y), "=r"(a[u].z), "=r"(a[u].w) : "l"(kb + i + u * 128));
          asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(b[u].x), "=r"(b[u].y), "=r"(b[u].z), "=r"(b[u].w) : "l"(vb + i + u * 128));
        }
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (i + u * 128 < n) x ^= a[u].x ^ a[u].w ^ b[u].y ^ b[u].z;
    }
  } else {  // K pass then V pass (three-pass shape, no barrier in between)
    for (int pass = 0; pass < 2; ++pass) {
      const uint4* p = pass ? vb : kb;
      for (int i = threadIdx.x; i < n; i += 128 * U) {
        uint4 a[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
          if (i + u * 128 < n)
            asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(a[u].x), "=r"(a[u].y), "=r"(a[u].z), "=r"(a[u].w) : "l"(p + i + u * 128));
#pragma unroll
        for (int u = 0; u < U; ++u)
          if (i + u * 128 < n) x ^= a[u].x ^ a[u].w;
      }
    }
  }
  if (x == 0x12345678u) sink[0] = x;
}

template <int U, bool TWO>
void run(const char* name, const uint4* k, const uint4* v, int rows, int c_stride, uint32_t* sink) {
  const size_t layer = (size_t)1024 * c_stride * 8;  // uint4 per layer
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  for (int rep = 0; rep < 2; ++rep) {
    cudaEventRecord(e0);
    for (int l = 0; l < 24; ++l) stream_kernel<U, TWO><<<1024, 128>>>(k + l * layer, v + l * layer, rows, c_stride, sink);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
  }
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  const double bytes = 24.0 * 1024 * rows * 128 * 2;
  printf("%-34s U=%2d : %7.2f us / launch  %7.0f GB/s\n", name, U, ms * 1e3 / 24, bytes / (ms * 1e-3) / 1e9);
}

int main() {
  const int rows = 472, c_stride = 512;
  const size_t n = (size_t)24 * 1024 * c_stride * 8;
  uint4 *k, *v;
  uint32_t* sink;
  cudaMalloc(&k, n * sizeof(uint4));
  cudaMalloc(&v, n * sizeof(uint4));
  cudaMalloc(&sink, 4);
  cudaMemset(k, 1, n * sizeof(uint4));
  cudaMemset(v, 2, n * sizeof(uint4));
  {  // steady state: ONE launch over all 24 caches (24576 CTAs): no per-launch ramp / tail
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(e0);
      stream_kernel<4, false><<<24 * 1024, 128>>>(k, v, rows, c_stride, sink);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
    }
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double bytes = 24.0 * 1024 * rows * 128 * 2;
    printf("one launch over 24 caches, U=4      : %7.2f us / cache   %7.0f GB/s\n", ms * 1e3 / 24, bytes / (ms * 1e-3) / 1e9);
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(e0);
      stream_kernel<8, true><<<24 * 1024, 128>>>(k, v, rows, c_stride, sink);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
    }
    cudaEventElapsedTime(&ms, e0, e1);
    printf("one launch over 24 caches, K+V U=8  : %7.2f us / cache   %7.0f GB/s\n", ms * 1e3 / 24, bytes / (ms * 1e-3) / 1e9);
  }
  run<4, false>("K pass then V pass", k, v, rows, c_stride, sink);
  run<8, false>("K pass then V pass", k, v, rows, c_stride, sink);
  run<16, false>("K pass then V pass", k, v, rows, c_stride, sink);
  run<2, true>("K and V together", k, v, rows, c_stride, sink);
  run<4, true>("K and V together", k, v, rows, c_stride, sink);
  run<8, true>("K and V together", k, v, rows, c_stride, sink);
  cudaError_t e = cudaDeviceSynchronize();
  printf("%s\n", cudaGetErrorString(e));
  return 0;
}
