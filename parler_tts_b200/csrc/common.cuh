// Shared helpers for libparler_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <float.h>

#include "../../include/parler_b200.h"

namespace ptts {

void set_error(const char* fmt, ...);
void count_launch();  // host-side tally of kernel launches enqueued by this library (ptts_launch_count)

#define PTTS_CHECK_ARG(cond, ...)                    \
  do {                                               \
    if (!(cond)) {                                   \
      ::ptts::set_error(__VA_ARGS__);                \
      return PTTS_E_ARG;                             \
    }                                                \
  } while (0)

#define PTTS_CHECK_LAUNCH(what)                                              \
  do {                                                                       \
    ::ptts::count_launch();                                                  \
    cudaError_t e__ = cudaPeekAtLastError();                                 \
    if (e__ != cudaSuccess) {                                                \
      ::ptts::set_error("%s: %s", what, cudaGetErrorString(e__));            \
      (void)cudaGetLastError();                                              \
      return PTTS_E_LAUNCH;                                                  \
    }                                                                        \
  } while (0)

#define PTTS_TRY(expr)                 \
  do {                                 \
    int rc__ = (expr);                 \
    if (rc__ != PTTS_OK) return rc__;  \
  } while (0)

// ---- knobs.  The product library is built WITHOUT PTTS_DIAG: every PTTS_KNOB is its compile-time default, nothing
// reads the environment and the ablation / timeline code is unreachable.  The diagnostic library (make DIAG=1 ->
// libparler_b200_diag.so, used by profiles/scripts only) reads the named environment variable instead.
#ifdef PTTS_DIAG
int knob_env(const char* name, int dflt);  // atoi(getenv(name)) or dflt; cached per call site by the caller
#define PTTS_KNOB(name, dflt)                              \
  ([]() -> int {                                           \
    static const int v__ = ::ptts::knob_env(name, dflt);   \
    return v__;                                            \
  }())
#else
#define PTTS_KNOB(name, dflt) (dflt)
#endif

inline bool pdl_enabled() { return PTTS_KNOB("PTTS_PDL", 1) != 0; }  // programmatic dependent launch

// Launch helper of the fast family: cudaLaunchKernelEx, with the programmatic-stream-serialization attribute when
// PDL is enabled (works eagerly and under stream capture into a CUDA graph).
template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                            Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  int n = 0;
  if (pdl_enabled()) {
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    n = 1;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

inline cudaStream_t as_stream(ptts_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }
inline size_t dtype_size(int dt) { return dt == PTTS_BF16 ? 2 : 4; }

template <typename T> struct DT;
template <> struct DT<float> { static constexpr int id = PTTS_F32; };
template <> struct DT<__nv_bfloat16> { static constexpr int id = PTTS_BF16; };

__device__ __forceinline__ float to_f(float v) { return v; }
__device__ __forceinline__ float to_f(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

// finfo(dtype).min as a float: what the reference's additive attention masks hold (modeling_parler_tts.py:875-879)
template <typename T> __device__ __forceinline__ float finfo_min();
template <> __device__ __forceinline__ float finfo_min<float>() { return -FLT_MAX; }
template <> __device__ __forceinline__ float finfo_min<__nv_bfloat16>() { return -3.3895313892515355e38f; }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__device__ __forceinline__ float apply_act(float x, int act) {
  switch (act) {
    case PTTS_ACT_GELU: return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f));
    case PTTS_ACT_RELU: return fmaxf(x, 0.0f);
    case PTTS_ACT_SILU: return x / (1.0f + expf(-x));
    case PTTS_ACT_GELU_TANH: {
      float u = 0.79788456080286535588f * (x + 0.044715f * x * x * x);
      return 0.5f * x * (1.0f + tanhf(u));
    }
    default: return x;
  }
}

// ---- programmatic dependent launch (PDL): every kernel of the fast family calls pdl_wait() before it touches data
// written by an earlier kernel, and pdl_launch() as early as possible so that the next kernel's prologue (weight
// prefetch, smem carve-up) overlaps this kernel's tail.  Both are no-ops for a launch without the attribute.
// Election of ONE lane of a converged warp (elect.sync).  Code guarded by it directly - `if (elect_one()) { ... }` -
// is known to ptxas to run in a single lane, so cp.async.bulk.tensor / tcgen05.mma / tcgen05.commit are issued from
// uniform registers back to back; guarded by `lane == 0` instead, every such instruction becomes an
// ELECT ... R2UR ... BRA.U.ANY loop over the (one) active lane: ~10 SASS instructions per MMA on the issuing thread.
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.b32 %0, 1, 0, p;\n\t}\n" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// ---- timeline diagnostics (PTTS_TIMELINE=1): every fast-family launch gets 4 uint64 slots {first CTA start, first /
// last CTA past griddepcontrol.wait, last CTA end} filled with globaltimer by atomics; null pointer = disabled.
struct TL {
  unsigned long long* p;
};
#ifdef PTTS_DIAG
TL tl_next(int kind);  // host: next timeline slot (null unless PTTS_TIMELINE=1); kind = tag of the launch site
#else
inline TL tl_next(int) { return TL{nullptr}; }
#endif
__device__ __forceinline__ unsigned long long gtimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(t));
  return t;
}
__device__ __forceinline__ void tl_start(TL t) {
  if (t.p && threadIdx.x == 0) atomicMin(t.p, gtimer_ns());
}
__device__ __forceinline__ void tl_dep(TL t) {
  if (t.p && threadIdx.x == 0) {
    const unsigned long long n = gtimer_ns();
    atomicMin(t.p + 1, n);
    atomicMax(t.p + 2, n);
  }
}
__device__ __forceinline__ void tl_end(TL t) {
  if (t.p && threadIdx.x == 0) atomicMax(t.p + 3, gtimer_ns());
}

// Asynchronous L2 prefetch of `bytes` (multiple of 16, 16-byte aligned) of global memory: fire and forget, holds no
// registers and no shared memory; used to keep HBM busy while the SMs wait on a dependency.
__device__ __forceinline__ void l2_prefetch_bulk(const void* p, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;\n" ::"l"(p), "r"(bytes) : "memory");
}
// the calling CTA prefetches [p, p + bytes) in 8 KB pieces, one per thread
__device__ __forceinline__ void l2_prefetch_range(const void* p, int64_t bytes) {
  constexpr int kPiece = 8192;
  for (int64_t o = (int64_t)threadIdx.x * kPiece; o < bytes; o += (int64_t)blockDim.x * kPiece) {
    const int64_t left = bytes - o;
    l2_prefetch_bulk(reinterpret_cast<const char*>(p) + o, (uint32_t)(left < kPiece ? left : kPiece));
  }
}

// same with one prefetch.global.L2 per 128-byte line (classic, LSU-issued)
__device__ __forceinline__ void l2_prefetch_lines(const void* p, int64_t bytes) {
  for (int64_t o = (int64_t)threadIdx.x * 128; o < bytes; o += (int64_t)blockDim.x * 128)
    asm volatile("prefetch.global.L2 [%0];\n" ::"l"(reinterpret_cast<const char*>(p) + o) : "memory");
}

// ---- weight prefetch hints: the decode step is a chain of short dependent kernels, and a GEMM's weight slab only
// starts to travel from HBM when its CTAs become resident (~4 us before their dependency resolves, 6 us of stream at
// 128 KB per SM).  The orchestrator (decoder.cu) therefore hands the matrix a LATER launch will stream to a kernel a
// few launches ahead of it (set_l2_hint, consumed by the next launch site that supports hints); that grid queues
// cp.async.bulk.prefetch.L2 over the matrix, one contiguous share per CTA, so the later slab loads hit L2.
struct L2Hint {
  const void* ptr;
  long long bytes;
};
void set_l2_hint(const void* ptr, long long bytes);  // host; PTTS_WPF bit mask selects which sites call it
L2Hint take_l2_hint();                               // host: the pending hint (or {null, 0}); clears it
// bit mask of hint sites; measured on B200 (B=64, 860 frames): every site costs 0-17 us per decode step -> 0
inline int wpf_mask() { return PTTS_KNOB("PTTS_WPF", 0); }
__device__ __forceinline__ void l2_hint_prefetch(const L2Hint& hint, int cta, int n_ctas) {
  if (!hint.ptr) return;
  constexpr long long kPiece = 8192;
  const long long share = ((hint.bytes + n_ctas - 1) / n_ctas + 127) / 128 * 128;
  const long long lo = (long long)cta * share;
  const long long hi = lo + share < hint.bytes ? lo + share : hint.bytes;
  for (long long o = lo + (long long)threadIdx.x * kPiece; o < hi; o += (long long)blockDim.x * kPiece) {
    const long long left = hi - o;
    l2_prefetch_bulk(reinterpret_cast<const char*>(hint.ptr) + o, (uint32_t)(left < kPiece ? left : kPiece));
  }
}

// value of a dynamic scalar: base + *dyn (dyn may be null)
__device__ __forceinline__ int dyn_value(int base, const int* dyn) { return base + (dyn ? *dyn : 0); }

}  // namespace ptts
