// ptts_attention_noalloc: the flash-decode attention of ptts_attention (fast family, one query token per utterance) for
// callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc).  No counterpart in the reference.
//
// ptts_attention needs caller-provided scratch for its split-KV partial results (small batches: the context of every
// (utterance, head) pair is split over several CTAs, each writes {running max, sum of exp, unnormalised output[64]}, a
// combine kernel merges them).  Here that scratch is a file-scope static __device__ array of kNoAllocScratchFloats
// floats (1 MiB): every GPU that loads the library holds it until the process exits.
//   * host code must not use the array's NAME as a pointer (it is a host-side stand-in): the device address comes from
//     cudaGetSymbolAddress, once per device, with that device current, return code checked;
//   * every call shares the array and it keeps the last call's contents.  Nothing has to start at zero: the kernel
//     writes every slot the combine kernel reads within the same call.  Calls that OVERLAP on the GPU (two streams, two
//     host threads) race on it - serialise them (one stream) or use ptts_attention with per-call scratch;
//   * no cudaMalloc / cudaFree / allocating library routine is called here or below (attention_fast only launches).
#include <mutex>

#include "ops.cuh"

namespace ptts {
namespace {

constexpr int64_t kNoAllocScratchFloats = 262144;  // 1 MiB: attn_scratch_floats(B, heads) fits for B * heads <= 3363
constexpr int kMaxDevices = 64;

static __device__ __align__(256) float g_noalloc_scratch[kNoAllocScratchFloats];

// device address of g_noalloc_scratch on the CURRENT device (looked up once per device); null + error text on failure
float* noalloc_scratch_ptr() {
  static std::mutex mu;
  static float* ptr[kMaxDevices] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    set_error("attention_noalloc: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
    return nullptr;
  }
  std::lock_guard<std::mutex> lock(mu);
  if (!ptr[dev]) {
    float* p = nullptr;
    e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_scratch);
    if (e != cudaSuccess || !p) {
      set_error("attention_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
      (void)cudaGetLastError();
      return nullptr;
    }
    ptr[dev] = p;
  }
  return ptr[dev];
}

}  // namespace
}  // namespace ptts

using namespace ptts;

extern "C" int64_t ptts_attention_noalloc_max_floats(void) { return kNoAllocScratchFloats; }

extern "C" int ptts_attention_noalloc(const void* q, const void* cache_k, const void* cache_v, int32_t dtype, int32_t B,
                                      int32_t heads, int32_t d, int32_t c_stride, int32_t ctx_base,
                                      const int32_t* ctx_dyn, const int32_t* key_mask, int32_t mask_len, void* out,
                                      ptts_stream_t s) {
  PTTS_CHECK_ARG(q && cache_k && cache_v && out, "attention_noalloc: null pointer");
  PTTS_CHECK_ARG(B > 0 && heads > 0 && c_stride > 0, "attention_noalloc: bad dims");
  PTTS_CHECK_ARG(fast_attention_supported(dtype, d, 1), "attention_noalloc: needs fp32 / bf16 and head_dim 64");
  const int64_t need = attn_scratch_floats(B, heads);
  if (need > kNoAllocScratchFloats) {
    set_error("attention_noalloc: B * heads = %d needs %lld floats of split-KV scratch, the static array holds %lld "
              "(B * heads <= 3363); use ptts_attention with caller scratch",
              B * heads, (long long)need, (long long)kNoAllocScratchFloats);
    return PTTS_E_ARG;
  }
  float* scratch = noalloc_scratch_ptr();
  if (!scratch) return PTTS_E_LAUNCH;
  return attention_fast(q, cache_k, cache_v, dtype, B, heads, d, 1, c_stride, ctx_base, ctx_dyn, 0, key_mask, mask_len,
                        out, scratch, as_stream(s));
}
