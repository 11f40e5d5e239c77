// Delay-pattern codec-token layout kernels (int64, bit-exact vs the oracle).
//   build : /root/reference/parler_tts/modeling_parler_tts.py:77-139
//   apply : modeling_parler_tts.py:68-74
//   de-delay : modeling_parler_tts.py:2633-2648        frame filter : modeling_parler_tts.py:2668-2670
#include "common.cuh"

namespace ptts {

__global__ void build_delay_mask_kernel(const int64_t* __restrict__ start, int B, int K, int T0, int M, int64_t bos,
                                        int64_t pad, int64_t* __restrict__ mask, int* __restrict__ first_start) {
  const bool degenerate = M < 2 * K - 1;
  const int64_t total = (int64_t)B * K * M;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int j = (int)(i % M);
    const int r = (int)(i / M);
    const int k = r % K;
    int64_t v = -1;
    if (!degenerate) {
      if (j <= k) v = bos;
      else if (j >= M - K + 1 + k) v = pad;
      else if (j >= k && j < T0 + k) v = start[(int64_t)r * T0 + (j - k)];
    }
    mask[i] = v;
    // first column of codebook 0 that is still -1, minimum over the batch
    if (!degenerate && k == 0 && v == -1) atomicMin(first_start, j);
  }
}

__global__ void apply_delay_mask_kernel(const int64_t* __restrict__ ids, int64_t ld_ids,
                                        const int64_t* __restrict__ mask, int64_t ld_mask, int rows, int len,
                                        int64_t* __restrict__ out, int64_t ld_out) {
  const int64_t total = (int64_t)rows * len;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int j = (int)(i % len);
    const int64_t r = i / len;
    const int64_t m = mask[r * ld_mask + j];
    out[r * ld_out + j] = (m == -1) ? ids[r * ld_ids + j] : m;
  }
}

__global__ void dedelay_kernel(const int64_t* __restrict__ ids, int64_t ld_ids, const int64_t* __restrict__ mask,
                               int64_t ld_mask, int B, int K, int F, int shift, int64_t* __restrict__ codes) {
  const int64_t total = (int64_t)B * K * F;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int f = (int)(i % F);
    const int64_t r = i / F;
    const int k = (int)(r % K);
    const int j = f + (shift ? k + 1 : 0);
    const int64_t m = mask[r * ld_mask + j];
    codes[i] = (m == -1) ? ids[r * ld_ids + j] : m;
  }
}

// one block per utterance: ordered compaction of the frames whose K codes are all < codebook_size
__global__ void filter_frames_kernel(const int64_t* __restrict__ codes, int K, int F, int64_t cbs,
                                     int64_t* __restrict__ packed, int* __restrict__ counts) {
  const int b = blockIdx.x;
  const int64_t* c = codes + (int64_t)b * K * F;
  int64_t* p = packed + (int64_t)b * K * F;
  __shared__ int warp_tot[32];
  __shared__ int base;
  if (threadIdx.x == 0) base = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  for (int f0 = 0; f0 < F; f0 += blockDim.x) {
    const int f = f0 + threadIdx.x;
    bool ok = f < F;
    if (ok)
      for (int k = 0; k < K; ++k) ok = ok && (c[(int64_t)k * F + f] < cbs);
    const unsigned bal = __ballot_sync(0xffffffffu, ok);
    if (lane == 0) warp_tot[warp] = __popc(bal);
    __syncthreads();
    int off = base;
    for (int w = 0; w < warp; ++w) off += warp_tot[w];
    off += __popc(bal & ((1u << lane) - 1u));
    if (ok)
      for (int k = 0; k < K; ++k) p[(int64_t)k * F + off] = c[(int64_t)k * F + f];
    __syncthreads();
    if (threadIdx.x == 0) {
      int t = 0;
      for (int w = 0; w < nwarp; ++w) t += warp_tot[w];
      base += t;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) counts[b] = base;
}

__global__ void set_int_kernel(int* p, int v) { *p = v; }

static inline int grid_for(int64_t total, int threads) {
  int64_t g = (total + threads - 1) / threads;
  if (g < 1) g = 1;
  if (g > 148 * 8) g = 148 * 8;
  return (int)g;
}

}  // namespace ptts

using namespace ptts;

extern "C" int ptts_build_delay_mask(const int64_t* start_ids, int32_t B, int32_t K, int32_t T0, int32_t max_len,
                                     int64_t bos, int64_t pad, int64_t* mask_out, int32_t* first_start,
                                     ptts_stream_t s) {
  PTTS_CHECK_ARG(start_ids && mask_out && first_start, "build_delay_mask: null pointer");
  PTTS_CHECK_ARG(B > 0 && K > 0 && K <= PTTS_MAX_CODEBOOKS && T0 > 0 && max_len > 0, "build_delay_mask: bad dims");
  PTTS_CHECK_ARG(max_len < 2 * K - 1 || T0 + K - 1 <= max_len, "build_delay_mask: prompt longer than max_length");
  cudaStream_t st = as_stream(s);
  set_int_kernel<<<1, 1, 0, st>>>(first_start, T0);  // "no -1 left" -> seq_len (modeling_parler_tts.py:134)
  const int64_t total = (int64_t)B * K * max_len;
  build_delay_mask_kernel<<<grid_for(total, 256), 256, 0, st>>>(start_ids, B, K, T0, max_len, bos, pad, mask_out,
                                                                first_start);
  PTTS_CHECK_LAUNCH("build_delay_mask");
  return PTTS_OK;
}

extern "C" int ptts_apply_delay_mask(const int64_t* ids, int64_t ld_ids, const int64_t* mask, int64_t ld_mask,
                                     int32_t rows, int32_t len, int64_t* out, int64_t ld_out, ptts_stream_t s) {
  PTTS_CHECK_ARG(ids && mask && out, "apply_delay_mask: null pointer");
  PTTS_CHECK_ARG(rows >= 0 && len >= 0 && ld_ids >= len && ld_mask >= len && ld_out >= len, "apply_delay_mask: bad dims");
  if (rows == 0 || len == 0) return PTTS_OK;
  apply_delay_mask_kernel<<<grid_for((int64_t)rows * len, 256), 256, 0, as_stream(s)>>>(ids, ld_ids, mask, ld_mask,
                                                                                       rows, len, out, ld_out);
  PTTS_CHECK_LAUNCH("apply_delay_mask");
  return PTTS_OK;
}

extern "C" int ptts_dedelay(const int64_t* ids, int64_t ld_ids, const int64_t* mask, int64_t ld_mask, int32_t B,
                            int32_t K, int32_t l_out, int64_t* codes, ptts_stream_t s) {
  PTTS_CHECK_ARG(B > 0 && K > 0 && l_out >= 1 && ld_ids >= l_out && ld_mask >= l_out, "dedelay: bad dims");
  // an output shorter than 2K-1 rebuilds an all -1 mask (modeling_parler_tts.py:105-106), so every column survives
  const bool degenerate = l_out < 2 * K - 1;
  const int F = degenerate ? l_out : l_out - K;
  if (F == 0) return PTTS_OK;  // nothing survives the de-delay: empty (B, K, 0) result
  PTTS_CHECK_ARG(ids && mask && codes, "dedelay: null pointer");
  dedelay_kernel<<<grid_for((int64_t)B * K * F, 256), 256, 0, as_stream(s)>>>(ids, ld_ids, mask, ld_mask, B, K, F,
                                                                             degenerate ? 0 : 1, codes);
  PTTS_CHECK_LAUNCH("dedelay");
  return PTTS_OK;
}

extern "C" int ptts_filter_frames(const int64_t* codes, int32_t B, int32_t K, int32_t F, int64_t codebook_size,
                                  int64_t* packed, int32_t* counts, ptts_stream_t s) {
  PTTS_CHECK_ARG(codes && packed && counts, "filter_frames: null pointer");
  PTTS_CHECK_ARG(B > 0 && K > 0 && F >= 0, "filter_frames: bad dims");
  filter_frames_kernel<<<B, 256, 0, as_stream(s)>>>(codes, K, F, codebook_size, packed, counts);
  PTTS_CHECK_LAUNCH("filter_frames");
  return PTTS_OK;
}
