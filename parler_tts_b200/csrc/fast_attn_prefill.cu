// Attention of the prompt prefill (Tq > 1 queries per utterance over a short context; self-attention with the causal
// mask over P + T0 positions, cross-attention over the S cached T5 keys): replaces bmm + additive mask + softmax + bmm of
// /root/reference/parler_tts/modeling_parler_tts.py:326-365 on the un-cached branch.
//
// The simple family runs one CTA per (head, utterance, QUERY) and re-reads the K / V rows of the pair for every query
// (33 792 CTAs and 280 MB of L2 reads for B=64, 33 positions: 201 us per launch, 41 % of the prefill).  Here one CTA owns
// a (head, utterance) pair: its K rows are staged in shared memory TRANSPOSED ([d][ctx + 1]: lanes = keys read
// consecutive elements), its V rows row-major ([ctx][d]: lanes = dims), once; warp w then takes queries w, w + NW, ...
// with the scores of a query in the warp's own strip.  fp32 accumulation, masked keys take finfo(dtype).min like the
// reference's additive mask.  Once per generate() call, not on the per-step path.
#include <stdlib.h>

#include "ops.cuh"

namespace ptts {

namespace {

constexpr int PD = 64;  // head dim
constexpr int PNW = 8;  // warps per CTA

template <typename T>
__global__ void __launch_bounds__(PNW * 32)
attn_prefill_kernel(const T* __restrict__ q, const T* __restrict__ ck, const T* __restrict__ cv, int heads, int Tq,
                    int c_stride, int ctx_base, const int* __restrict__ ctx_dyn, int causal,
                    const int* __restrict__ key_mask, int mask_len, T* __restrict__ out) {
  extern __shared__ __align__(16) unsigned char psm[];
  const int hd = blockIdx.x, b = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int H = heads * PD;
  const int ctx = dyn_value(ctx_base, ctx_dyn);
  const int CP = ctx + 1;  // pitch of the transposed K rows (odd pitches keep the staging stores off one bank)
  T* kt = reinterpret_cast<T*>(psm);                  // [PD][CP]
  T* vs = kt + (size_t)PD * CP;                       // [ctx][PD]
  float* strip = reinterpret_cast<float*>(psm + (((size_t)PD * CP + (size_t)ctx * PD) * sizeof(T) + 15) / 16 * 16);
  float* qs = strip + (size_t)warp * (PD + ctx);      // per warp: [PD] query, [ctx] scores -> probabilities
  float* sc = qs + PD;
  const T* kb = ck + ((int64_t)b * heads + hd) * (int64_t)c_stride * PD;
  const T* vb = cv + ((int64_t)b * heads + hd) * (int64_t)c_stride * PD;
  for (int e0 = threadIdx.x; e0 < ctx * PD; e0 += PNW * 32 * 4) {  // eight loads in flight per thread
    T kk[4], vv[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int e = e0 + u * PNW * 32;
      if (e < ctx * PD) {
        kk[u] = kb[e];
        vv[u] = vb[e];
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int e = e0 + u * PNW * 32;
      if (e < ctx * PD) {
        const int j = e / PD, d = e - j * PD;
        kt[d * CP + j] = kk[u];
        vs[e] = vv[u];
      }
    }
  }
  __syncthreads();
  const float neg = finfo_min<T>();
  for (int t = warp; t < Tq; t += PNW) {
    const T* qrow = q + ((int64_t)b * Tq + t) * H + hd * PD;
    qs[lane] = to_f(qrow[lane]);
    qs[lane + 32] = to_f(qrow[lane + 32]);
    __syncwarp();
    const int limit = causal ? (ctx - Tq + t + 1) : ctx;  // keys j < limit are causally visible
    float lmax = -INFINITY;
    for (int j = lane; j < ctx; j += 32) {
      float dot = 0.f;
#pragma unroll 16
      for (int d = 0; d < PD; ++d) dot = fmaf(qs[d], to_f(kt[d * CP + j]), dot);
      bool keep = j < limit;
      if (keep && key_mask && j < mask_len) keep = key_mask[(int64_t)b * mask_len + j] != 0;
      const float sv = keep ? dot : neg;
      sc[j] = sv;
      lmax = fmaxf(lmax, sv);
    }
    lmax = warp_max(lmax);
    float lsum = 0.f;
    for (int j = lane; j < ctx; j += 32) {
      const float e = expf(sc[j] - lmax);
      sc[j] = e;
      lsum += e;
    }
    lsum = warp_sum(lsum);
    __syncwarp();
    const float inv = 1.0f / lsum;
    float o0 = 0.f, o1 = 0.f;  // dims lane and lane + 32
    for (int j = 0; j < ctx; ++j) {
      const float pj = sc[j] * inv;
      o0 = fmaf(pj, to_f(vs[j * PD + lane]), o0);
      o1 = fmaf(pj, to_f(vs[j * PD + lane + 32]), o1);
    }
    T* orow = out + ((int64_t)b * Tq + t) * H + hd * PD;
    orow[lane] = from_f<T>(o0);
    orow[lane + 32] = from_f<T>(o1);
    __syncwarp();  // the next query overwrites qs / sc
  }
}

size_t prefill_smem(int dtype, int ctx_max) {
  const size_t es = dtype_size(dtype);
  return (((size_t)PD * (ctx_max + 1) + (size_t)ctx_max * PD) * es + 15) / 16 * 16 + (size_t)PNW * (PD + ctx_max) * sizeof(float);
}

}  // namespace

// ctx_max: upper bound of the context this launch can see (host-side: ctx_base + what ctx_dyn may add)
bool attention_prefill_supported(int dtype, int d, int Tq, int ctx_max) {
  return (dtype == PTTS_F32 || dtype == PTTS_BF16) && d == PD && Tq > 1 && ctx_max > 0 &&
         prefill_smem(dtype, ctx_max) <= 200 * 1024;
}

int attention_prefill(const void* q, const void* ck, const void* cv, int dtype, int B, int heads, int Tq, int c_stride,
                      int ctx_base, const int* ctx_dyn, int ctx_max, int causal, const int* key_mask, int mask_len,
                      void* out, cudaStream_t st) {
  if (B == 0 || Tq == 0) return PTTS_OK;
  const size_t smem = prefill_smem(dtype, ctx_max);
  const dim3 grid(heads, B);
  if (dtype == PTTS_F32) {
    if (smem > 48 * 1024)
      cudaFuncSetAttribute(attn_prefill_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    attn_prefill_kernel<float><<<grid, PNW * 32, smem, st>>>((const float*)q, (const float*)ck, (const float*)cv, heads, Tq,
                                                             c_stride, ctx_base, ctx_dyn, causal, key_mask, mask_len,
                                                             (float*)out);
  } else {
    if (smem > 48 * 1024)
      cudaFuncSetAttribute(attn_prefill_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    attn_prefill_kernel<__nv_bfloat16><<<grid, PNW * 32, smem, st>>>(
        (const __nv_bfloat16*)q, (const __nv_bfloat16*)ck, (const __nv_bfloat16*)cv, heads, Tq, c_stride, ctx_base,
        ctx_dyn, causal, key_mask, mask_len, (__nv_bfloat16*)out);
  }
  PTTS_CHECK_LAUNCH("attention_prefill");
  return PTTS_OK;
}

}  // namespace ptts
