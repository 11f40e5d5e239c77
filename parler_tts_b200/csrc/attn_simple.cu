// Simple attention over a K/V cache: one block per (head, batch, query).  Replaces bmm + additive mask + softmax + bmm
// of /root/reference/parler_tts/modeling_parler_tts.py:326-365.  Masked keys take the score finfo(dtype).min exactly
// like the reference's additive mask (so a fully-masked row degenerates to a uniform average over ALL keys, as there).
#include "ops.cuh"

namespace ptts {

template <typename T>
__global__ void __launch_bounds__(128)
attention_simple_kernel(const T* __restrict__ q, const T* __restrict__ ck, const T* __restrict__ cv, int heads, int d,
                        int Tq, int c_stride, int ctx_base, const int* __restrict__ ctx_dyn, int causal,
                        const int* __restrict__ key_mask, int mask_len, T* __restrict__ out) {
  extern __shared__ float smem[];
  const int hd = blockIdx.x, b = blockIdx.y, t = blockIdx.z;
  const int H = heads * d;
  const int ctx = dyn_value(ctx_base, ctx_dyn);
  const int limit = causal ? (ctx - Tq + t + 1) : ctx;  // keys j < limit are causally visible
  float* sc = smem;              // [ctx] scores -> probabilities
  float* qs = smem + c_stride;   // [d]
  float* red = qs + d;           // [8]
  float* acc2 = red + 8;         // [4][d] partial outputs per warp
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const T* qrow = q + ((int64_t)b * Tq + t) * H + hd * d;
  for (int i = threadIdx.x; i < d; i += blockDim.x) qs[i] = to_f(qrow[i]);
  __syncthreads();
  const T* kb = ck + ((int64_t)b * heads + hd) * (int64_t)c_stride * d;
  const T* vb = cv + ((int64_t)b * heads + hd) * (int64_t)c_stride * d;
  const float neg = finfo_min<T>();
  // pass 1: scores
  float lmax = -INFINITY;
  for (int j = warp; j < ctx; j += 4) {
    float part = 0.f;
    for (int i = lane; i < d; i += 32) part = fmaf(qs[i], to_f(kb[(int64_t)j * d + i]), part);
    part = warp_sum(part);
    bool keep = j < limit;
    if (keep && key_mask && j < mask_len) keep = key_mask[(int64_t)b * mask_len + j] != 0;
    const float s = keep ? part : neg;
    if (lane == 0) sc[j] = s;
    lmax = fmaxf(lmax, s);
  }
  if (lane == 0) red[warp] = lmax;
  __syncthreads();
  const float mx = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
  __syncthreads();
  // pass 2: exp and sum
  float lsum = 0.f;
  for (int j = threadIdx.x; j < ctx; j += blockDim.x) {
    const float e = expf(sc[j] - mx);
    sc[j] = e;
    lsum += e;
  }
  lsum = warp_sum(lsum);
  if (lane == 0) red[4 + warp] = lsum;
  __syncthreads();
  const float inv = 1.0f / (red[4] + red[5] + red[6] + red[7]);
  // pass 3: out = sum_j p_j v_j ; warp w takes keys j = w mod 4, lanes take dims
  for (int i0 = 0; i0 < d; i0 += 32) {
    const int i = i0 + lane;
    float o = 0.f;
    if (i < d)
      for (int j = warp; j < ctx; j += 4) o = fmaf(sc[j] * inv, to_f(vb[(int64_t)j * d + i]), o);
    if (i < d) acc2[warp * d + i] = o;
  }
  __syncthreads();
  T* orow = out + ((int64_t)b * Tq + t) * H + hd * d;
  for (int i = threadIdx.x; i < d; i += blockDim.x)
    orow[i] = from_f<T>(acc2[i] + acc2[d + i] + acc2[2 * d + i] + acc2[3 * d + i]);
}

int attention_simple(const void* q, const void* ck, const void* cv, int dtype, int B, int heads, int d, int Tq,
                     int c_stride, int ctx_base, const int* ctx_dyn, int causal, const int* key_mask, int mask_len,
                     void* out, cudaStream_t st) {
  if (B == 0 || Tq == 0) return PTTS_OK;
  const size_t smem = sizeof(float) * ((size_t)c_stride + d + 8 + 4 * d);
  dim3 grid(heads, B, Tq);
  if (dtype == PTTS_F32) {
    if (smem > 48 * 1024)
      cudaFuncSetAttribute(attention_simple_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    attention_simple_kernel<float><<<grid, 128, smem, st>>>((const float*)q, (const float*)ck, (const float*)cv, heads,
                                                           d, Tq, c_stride, ctx_base, ctx_dyn, causal, key_mask,
                                                           mask_len, (float*)out);
  } else {
    if (smem > 48 * 1024)
      cudaFuncSetAttribute(attention_simple_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int)smem);
    attention_simple_kernel<__nv_bfloat16><<<grid, 128, smem, st>>>(
        (const __nv_bfloat16*)q, (const __nv_bfloat16*)ck, (const __nv_bfloat16*)cv, heads, d, Tq, c_stride, ctx_base,
        ctx_dyn, causal, key_mask, mask_len, (__nv_bfloat16*)out);
  }
  PTTS_CHECK_LAUNCH("attention_simple");
  return PTTS_OK;
}

}  // namespace ptts
