// Compiled into libparler_b200_diag.so only (make DIAG=1).
// Diagnostics (not on the product path): dependent-launch chains used by profiles/ scripts to measure the per-kernel
// latency floor of a CUDA-graph replay on the box (empty kernels, one-load kernels, with / without PDL).
#include <stdlib.h>

#include "../../include/parler_b200_diag.h"
#include "gemm_common.cuh"

namespace ptts {
__global__ void diag_empty_kernel(float* buf) {
  pdl_wait();
  pdl_launch();
  if (buf == nullptr && threadIdx.x == 9999) printf("x");
}
__global__ void diag_touch_kernel(float* buf, int n) {
  pdl_wait();
  pdl_launch();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) buf[i] = __ldcg(buf + i) + 1.0f;
}
}  // namespace ptts

namespace ptts {
static unsigned long long* g_tl_dev = nullptr;
static int g_tl_used = 0;
static int g_tl_kind[4096];
constexpr int kTlSlots = 4096;
__global__ void tl_reset_kernel(unsigned long long* p, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    p[4 * i] = ~0ull;
    p[4 * i + 1] = ~0ull;
    p[4 * i + 2] = 0ull;
    p[4 * i + 3] = 0ull;
  }
}
TL tl_next(int kind) {
  static int enabled = -1;
  if (enabled < 0) {
    enabled = PTTS_KNOB("PTTS_TIMELINE", 0) ? 1 : 0;
    if (enabled) {
      cudaMalloc(&g_tl_dev, sizeof(unsigned long long) * 4 * kTlSlots);
      tl_reset_kernel<<<(kTlSlots + 255) / 256, 256>>>(g_tl_dev, kTlSlots);
      cudaDeviceSynchronize();
    }
  }
  TL t;
  t.p = nullptr;
  if (enabled && g_tl_used < kTlSlots) {
    g_tl_kind[g_tl_used] = kind;
    t.p = g_tl_dev + 4 * g_tl_used++;
  }
  return t;
}
}  // namespace ptts

using namespace ptts;

// Timeline diagnostics (PTTS_TIMELINE=1).  cmd 0: reset all slots (device sync); cmd 1: copy 4 uint64 per slot to
// `host` and the launch-site tags to `kinds`, return the number of slots handed out so far.
extern "C" int ptts_diag_timeline(int32_t cmd, uint64_t* host, int32_t* kinds, int32_t max_slots) {
  cudaDeviceSynchronize();
  if (!g_tl_dev) return 0;
  if (cmd == 0) {
    tl_reset_kernel<<<(kTlSlots + 255) / 256, 256>>>(g_tl_dev, kTlSlots);
    cudaDeviceSynchronize();
    return g_tl_used;
  }
  const int n = g_tl_used < max_slots ? g_tl_used : max_slots;
  cudaMemcpy(host, g_tl_dev, sizeof(unsigned long long) * 4 * n, cudaMemcpyDeviceToHost);
  for (int i = 0; i < n; ++i) kinds[i] = g_tl_kind[i];
  return n;
}

// mode 0: n empty kernels (grid ctas x 128); mode 1: n kernels doing one dependent load+store per thread
extern "C" int ptts_diag_chain(int32_t n, int32_t mode, int32_t ctas, float* buf, ptts_stream_t s) {
  PTTS_CHECK_ARG(n >= 0 && ctas > 0 && buf, "diag_chain: bad arguments");
  for (int i = 0; i < n; ++i) {
    cudaError_t e;
    if (mode == 0)
      e = launch_k(diag_empty_kernel, dim3(ctas), dim3(128), 0, as_stream(s), buf);
    else
      e = launch_k(diag_touch_kernel, dim3(ctas), dim3(128), 0, as_stream(s), buf, ctas * 128);
    if (e != cudaSuccess) {
      set_error("diag_chain: %s", cudaGetErrorString(e));
      return PTTS_E_LAUNCH;
    }
    PTTS_CHECK_LAUNCH("diag_chain");
  }
  return PTTS_OK;
}

// PTTS_XB_DBG=1: per-phase SM-clock sums of the fused cross-attention block (fast_xblock.cu); reset != 0 clears them
namespace ptts { int xblock_read_stamps(unsigned long long* host, int reset); }
extern "C" int ptts_diag_xblock(uint64_t* host, int32_t reset) {
  cudaDeviceSynchronize();
  return xblock_read_stamps((unsigned long long*)host, reset);
}

// Copies the globaltimer stamps recorded by the tcgen05 GEMM under PTTS_GEMM_DBG=1 (8 x uint64 per launch, ring of 256)
// to host memory; returns the number of launches recorded so far.  Synchronises the device (diagnostic only).
extern "C" int ptts_diag_gemm_stamps(uint64_t* host, int32_t n_launches) {
  cudaDeviceSynchronize();
  return gemm::umma_read_stamps((unsigned long long*)host, n_launches);
}
