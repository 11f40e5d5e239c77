// Flash-decoding self-attention over an FP8 (e4m3) static K/V cache [B][heads][c_max][64] with one dequantisation scale
// per (layer, K|V, head): the opt-in `kv_dtype = fp8` configuration of the decode step.  Same role as
// attn_decode_kernel (fast_attn.cu): replaces bmm + additive mask + softmax + bmm and the torch.cat of the cache of
// /root/reference/parler_tts/modeling_parler_tts.py:299-304, 326-365 on the cached branch.
//
// Why: at B=64 / 860 frames the K/V stream is 82 % of the bytes a decode step must move; one byte
// per element halves it.  The kernel keeps the proven shape of the bf16 one (one CTA per (head, utterance[, split]),
// 128 threads, 16-byte ld.global.cg, two 8 KB batches in flight per CTA, K pass -> softmax -> V pass with the scores in
// shared memory), but a 64-byte row is covered by 4 threads instead of 8, so the arithmetic per loaded byte doubles:
//   K pass: e4m3x2 -> f16x2 (one cvt per two elements), HFMA2 against q / max|q| held as f16x2 (products <= 448, four per
//           f16 lane before the fp32 reduction), the K scale and max|q| are applied once per score in fp32;
//   V pass: e4m3x2 -> f16x2 -> fp32 FMA with the fp32 probability (fp32 accumulation over the whole context), the V scale
//           is applied once per output element.
// The fused prologue folds the split-K partial sums of the QKV GEMM, scales q and APPENDS the new K / V row quantised
// with the layer's scales.  Scales come from kv_quantize_fp8 (below): after the prefill has run on a bf16 staging
// cache, amax over the prompt positions of every (layer, K|V, head), times a head-room factor of 4 for later tokens,
// divided by 448; the staged rows are converted with cvt.rn.satfinite (later outliers saturate instead of wrapping).
#include <cuda_fp16.h>

#include "ops.cuh"

namespace ptts {

namespace {

constexpr int NT = 128;
constexpr int D = 64;
constexpr int TPK = 4;           // threads per 64-byte cache row
constexpr int KPB = NT / TPK;    // rows per CTA pass (32)
constexpr int U = 4;             // independent 16-byte loads per thread per batch
constexpr int BATCH = KPB * U;   // 128 rows = 8 KB per batch, two batches in flight
constexpr float kE4M3Max = 448.0f;
constexpr float kHeadRoom = 4.0f;

__device__ __forceinline__ uint4 ld_cg16(const void* p) {
  uint4 u;
  asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];\n" : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w) : "l"(p));
  return u;
}

// two e4m3 bytes (low 16 bits of `v`) -> f16x2, element 0 in the low half
__device__ __forceinline__ __half2 e4m3x2_to_half2(uint32_t v) {
  uint32_t r;
  const unsigned short s = (unsigned short)v;
  asm("cvt.rn.f16x2.e4m3x2 %0, %1;\n" : "=r"(r) : "h"(s));
  return *reinterpret_cast<__half2*>(&r);
}
// 16 bytes -> 8 x f16x2 in element order
__device__ __forceinline__ void unpack16(const uint4& u, __half2 (&h)[8]) {
  h[0] = e4m3x2_to_half2(u.x); h[1] = e4m3x2_to_half2(u.x >> 16);
  h[2] = e4m3x2_to_half2(u.y); h[3] = e4m3x2_to_half2(u.y >> 16);
  h[4] = e4m3x2_to_half2(u.z); h[5] = e4m3x2_to_half2(u.z >> 16);
  h[6] = e4m3x2_to_half2(u.w); h[7] = e4m3x2_to_half2(u.w >> 16);
}
__device__ __forceinline__ unsigned char f32_to_e4m3(float x) {
  unsigned short r;
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;\n" : "=h"(r) : "f"(0.0f), "f"(x));  // low byte = second operand
  return (unsigned char)(r & 0xff);
}

struct Fp8Fuse {
  const float* part;  // split-K partial sums [S][rows][3H] of the QKV GEMM, or null (q read from the q buffer)
  int S;
  int64_t stride;
  int ncols;
  float scale;        // 1 / sqrt(d)
  int slot_base;
  const int* slot_dyn;
};

// part layout (splits > 1): [(b*heads + h)*splits + s][D + 2] = { running max, sum of exp, unnormalised output[D] }
// PLAIN: no key mask and no KV split (the B=64 decode step): both are compiled out
template <bool FUSED, bool PLAIN>
__global__ void __launch_bounds__(NT, 7)
attn_fp8_kernel(const Fp8Fuse qf, const __nv_bfloat16* __restrict__ q, unsigned char* __restrict__ ck,
                unsigned char* __restrict__ cv, const float* __restrict__ k_scale, const float* __restrict__ v_scale,
                int heads, int c_stride, int ctx_base, const int* __restrict__ ctx_dyn,
                const int* __restrict__ key_mask_in, int mask_len, __nv_bfloat16* __restrict__ out,
                float* __restrict__ part, int splits_in, int sc_cap) {
  const int* key_mask = PLAIN ? nullptr : key_mask_in;
  const int splits = PLAIN ? 1 : splits_in;
  __shared__ float qsf[D];
  __shared__ float red[12];
  extern __shared__ float sm[];
  float* sc = sm;                // [sc_cap] scores -> probabilities
  float* ored = sm + sc_cap;     // [KPB][D] output partials
  const int hd = blockIdx.x, b = blockIdx.y, sp = blockIdx.z;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = tid / TPK, c = tid % TPK;
  const int H = heads * D;
  const unsigned char* kb = ck + ((int64_t)b * heads + hd) * (int64_t)c_stride * D + c * 16;
  const unsigned char* vb = cv + ((int64_t)b * heads + hd) * (int64_t)c_stride * D + c * 16;
  // Decode-step contract (as in fast_attn.cu): rows below the slot this launch appends and the position counter are at
  // least two launches old, so they may be read BEFORE griddepcontrol.wait, while the QKV GEMM still runs.
  const int ctx = dyn_value(ctx_base, ctx_dyn);
  const int chunk = (ctx + splits - 1) / splits;
  const int j0 = sp * chunk;
  const int j1 = min(ctx, j0 + chunk);
  const int n = max(j1 - j0, 0);
  auto load_batch = [&](uint4 (&raw)[U], const unsigned char* base_ptr, int base) {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int j = base + u * KPB + g;
      if (j < j1) raw[u] = ld_cg16(base_ptr + (int64_t)j * D);
    }
  };
  uint4 cur[U], nxt[U];
  int pre = 0;  // K batches requested before the dependency resolved
  if constexpr (FUSED) {
    const int old_end = dyn_value(qf.slot_base, qf.slot_dyn);
    if (j0 < j1 && min(j1, j0 + BATCH) <= old_end) {
      load_batch(cur, kb, j0);
      pre = 1;
      if (j0 + BATCH < j1 && min(j1, j0 + 2 * BATCH) <= old_end) {
        load_batch(nxt, kb, j0 + BATCH);
        pre = 2;
      }
    }
  }
  pdl_wait();
  pdl_launch();
  const float ks = k_scale[hd], vs = v_scale[hd];
  // ---- prologue: q (fp32) into shared memory; fused: fold the partial sums, append the new K / V row
  if constexpr (FUSED) {
    if (tid < D) {
      const float* p0 = qf.part + (int64_t)b * qf.ncols + hd * D + tid;
      float aq = 0.f, ak = 0.f, av = 0.f;
      for (int s0 = 0; s0 < qf.S; s0 += 4) {  // 4 slices (12 loads) in flight, added in slice order
        float tq[4], tk[4], tv[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (s0 + u < qf.S) {
            const float* ps = p0 + (s0 + u) * qf.stride;
            tq[u] = __ldcg(ps);
            tk[u] = __ldcg(ps + H);
            tv[u] = __ldcg(ps + 2 * H);
          }
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (s0 + u < qf.S) {
            aq += tq[u];
            ak += tk[u];
            av += tv[u];
          }
      }
      qsf[tid] = aq * qf.scale;
      {  // every KV split of (b, head) writes the same bytes: whichever split reads the slot sees its own write
        const int slot = dyn_value(qf.slot_base, qf.slot_dyn);
        const int64_t o = (((int64_t)b * heads + hd) * c_stride + slot) * D + tid;
        ck[o] = f32_to_e4m3(ak / ks);
        cv[o] = f32_to_e4m3(av / vs);
      }
    }
  } else {
    if (tid < D) qsf[tid] = __bfloat162float(q[(int64_t)b * H + hd * D + tid]);
  }
  __syncthreads();
  // q / max|q| as f16x2 (16 dims of this thread): products with e4m3 values stay below 448, four of them per f16 lane
  float qmax = 0.f;
  {
    float m = fmaxf(fabsf(qsf[lane]), fabsf(qsf[lane + 32]));
    qmax = warp_max(m);
  }
  const float qinv = qmax > 0.f ? 1.0f / qmax : 0.f;
  __half2 q2[8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
    q2[i] = __floats2half2_rn(qsf[c * 16 + 2 * i] * qinv, qsf[c * 16 + 2 * i + 1] * qinv);
  const float sfac = qmax * ks;
  const float neg = -3.3895313892515355e38f;  // finfo(bf16).min: the reference's additive mask value in bf16 runs
  // ---- pass 1: scores
  float lmax = -INFINITY;
  if (pre == 0) load_batch(cur, kb, j0);
  for (int base = j0; base < j1; base += BATCH) {
    if (base + BATCH < j1 && !(pre == 2 && base == j0)) load_batch(nxt, kb, base + BATCH);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int j = base + u * KPB + g;
      float dot = 0.f;
      if (j < j1) {
        __half2 kf[8];
        unpack16(cur[u], kf);
        __half2 a0 = __hmul2(q2[0], kf[0]), a1 = __hmul2(q2[1], kf[1]);
#pragma unroll
        for (int i = 2; i < 8; i += 2) {
          a0 = __hfma2(q2[i], kf[i], a0);
          a1 = __hfma2(q2[i + 1], kf[i + 1], a1);
        }
        const float2 f0 = __half22float2(a0), f1 = __half22float2(a1);
        dot = (f0.x + f0.y) + (f1.x + f1.y);
      }
      dot += __shfl_xor_sync(0xffffffffu, dot, 1);
      dot += __shfl_xor_sync(0xffffffffu, dot, 2);
      if (c == 0 && j < j1) {
        bool keep = true;
        if (key_mask && j < mask_len) keep = key_mask[(int64_t)b * mask_len + j] != 0;
        const float s = keep ? dot * sfac : neg;
        sc[j - j0] = s;
        lmax = fmaxf(lmax, s);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) cur[u] = nxt[u];
  }
  load_batch(cur, vb, j0);  // the first V rows travel while the block reduces max / sum
  lmax = warp_max(lmax);
  if (lane == 0) red[warp] = lmax;
  __syncthreads();
  const float mx = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
  float lsum = 0.f;
#pragma unroll 4
  for (int i = tid; i < n; i += NT) {
    const float e = expf(sc[i] - mx);
    sc[i] = e;
    lsum += e;
  }
  lsum = warp_sum(lsum);
  if (lane == 0) red[4 + warp] = lsum;
  __syncthreads();
  const float total = red[4] + red[5] + red[6] + red[7];
  const float inv = ((splits == 1 && n > 0) ? 1.0f / total : 1.0f) * vs;
  // ---- pass 3: out = sum_j p_j V_j
  float acc[16];
#pragma unroll
  for (int e = 0; e < 16; ++e) acc[e] = 0.f;
  for (int base = j0; base < j1; base += BATCH) {
    if (base + BATCH < j1) load_batch(nxt, vb, base + BATCH);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int j = base + u * KPB + g;
      if (j < j1) {
        __half2 vf[8];
        unpack16(cur[u], vf);
        const float pj = sc[j - j0];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float2 f = __half22float2(vf[i]);
          acc[2 * i] = fmaf(pj, f.x, acc[2 * i]);
          acc[2 * i + 1] = fmaf(pj, f.y, acc[2 * i + 1]);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) cur[u] = nxt[u];
  }
#pragma unroll
  for (int e = 0; e < 16; e += 4)
    *reinterpret_cast<float4*>(ored + g * D + c * 16 + e) = make_float4(acc[e], acc[e + 1], acc[e + 2], acc[e + 3]);
  __syncthreads();
  if (tid < D) {
    float o = 0.f;
#pragma unroll
    for (int gg = 0; gg < KPB; ++gg) o += ored[gg * D + tid];
    o *= inv;
    if (splits == 1) {
      out[(int64_t)b * H + hd * D + tid] = __float2bfloat16_rn(o);
    } else {
      float* pp = part + (((int64_t)b * heads + hd) * splits + sp) * (D + 2);
      pp[2 + tid] = o;
      if (tid == 0) {
        pp[0] = n > 0 ? mx : -INFINITY;
        pp[1] = n > 0 ? total : 0.f;
      }
    }
  }
}

__global__ void __launch_bounds__(D)
attn_fp8_combine_kernel(const float* __restrict__ part, int heads, int splits, __nv_bfloat16* __restrict__ out) {
  const int hd = blockIdx.x, b = blockIdx.y, t = threadIdx.x;
  pdl_wait();
  pdl_launch();
  const float* pp = part + ((int64_t)b * heads + hd) * splits * (D + 2);
  float mx = -INFINITY;
  for (int s = 0; s < splits; ++s) mx = fmaxf(mx, pp[s * (D + 2)]);
  float l = 0.f, o = 0.f;
  for (int s = 0; s < splits; ++s) {
    const float m = pp[s * (D + 2)];
    if (m == -INFINITY) continue;
    const float w = expf(m - mx);
    l = fmaf(pp[s * (D + 2) + 1], w, l);
    o = fmaf(pp[s * (D + 2) + 2 + t], w, o);
  }
  out[((int64_t)b * heads + hd) * D + t] = __float2bfloat16_rn(o / l);
}

int plan_splits_fp8(int B, int heads, int c_stride) {
  const int pairs = B * heads;
  if (pairs >= 512) return 1;
  int s = (592 + pairs - 1) / pairs;
  const int cap = c_stride / 128 > 0 ? c_stride / 128 : 1;  // at least one 128-row batch per split
  if (s > cap) s = cap;
  if (s > 16) s = 16;
  return s < 1 ? 1 : s;
}

int launch_fp8(const Fp8Fuse& qf, const void* q, void* ck, void* cv, const float* k_scale, const float* v_scale, int B,
               int heads, int c_stride, int ctx_base, const int* ctx_dyn, const int* key_mask, int mask_len, void* out,
               float* scratch, cudaStream_t st) {
  const int splits = plan_splits_fp8(B, heads, c_stride);
  const int cap = ((c_stride + splits - 1) / splits + 1 + 63) / 64 * 64 + 64;
  PTTS_CHECK_ARG(splits == 1 || scratch, "attention_fp8: split-KV needs scratch");
  const size_t smem = sizeof(float) * ((size_t)cap + (size_t)KPB * D);
  PTTS_CHECK_ARG(smem <= 200 * 1024, "attention_fp8: cache too long for the score buffer");
  if (smem > 48 * 1024) {
    static size_t configured = 0;  // high-water mark of the opt-in dynamic shared memory (host state, see parler_b200.h)
    if (smem > configured) {
      cudaFuncSetAttribute(attn_fp8_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      cudaFuncSetAttribute(attn_fp8_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      cudaFuncSetAttribute(attn_fp8_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      cudaFuncSetAttribute(attn_fp8_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      configured = smem;
    }
  }
  const bool plain = !key_mask && splits == 1;
  const dim3 grid(heads, B, splits), block(NT);
  cudaError_t e;
#define PTTS_FP8_LAUNCH(F, P)                                                                                         \
  launch_k(attn_fp8_kernel<F, P>, grid, block, smem, st, qf, (const __nv_bfloat16*)q, (unsigned char*)ck,            \
           (unsigned char*)cv, k_scale, v_scale, heads, c_stride, ctx_base, ctx_dyn, key_mask, mask_len,              \
           (__nv_bfloat16*)out, scratch, splits, cap)
  if (qf.part)
    e = plain ? PTTS_FP8_LAUNCH(true, true) : PTTS_FP8_LAUNCH(true, false);
  else
    e = plain ? PTTS_FP8_LAUNCH(false, true) : PTTS_FP8_LAUNCH(false, false);
#undef PTTS_FP8_LAUNCH
  if (e != cudaSuccess) {
    set_error("attn_fp8 launch: %s", cudaGetErrorString(e));
    return PTTS_E_LAUNCH;
  }
  PTTS_CHECK_LAUNCH("attn_fp8");
  if (splits > 1) {
    e = launch_k(attn_fp8_combine_kernel, dim3(heads, B), dim3(D), 0, st, (const float*)scratch, heads, splits,
                 (__nv_bfloat16*)out);
    if (e != cudaSuccess) {
      set_error("attn_fp8_combine launch: %s", cudaGetErrorString(e));
      return PTTS_E_LAUNCH;
    }
    PTTS_CHECK_LAUNCH("attn_fp8_combine");
  }
  return PTTS_OK;
}

// ---- quantisation of the staged prefill cache
// amax[(l*2 + kv)*heads + h] = max |x| over the B x slots x 64 staged elements (float bits ordered as unsigned)
__global__ void __launch_bounds__(256)
kv_amax_kernel(const __nv_bfloat16* __restrict__ stage, int heads, int slots, unsigned int* __restrict__ amax) {
  const int hd = blockIdx.x, b = blockIdx.y, lk = blockIdx.z;
  const int B = gridDim.y;
  const __nv_bfloat16* src = stage + (((int64_t)lk * B + b) * heads + hd) * (int64_t)slots * D;
  float m = 0.f;
  for (int i = threadIdx.x; i < slots * D; i += blockDim.x) m = fmaxf(m, fabsf(__bfloat162float(src[i])));
  m = warp_max(m);
  __shared__ float r[8];
  if ((threadIdx.x & 31) == 0) r[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = r[0];
    for (int i = 1; i < 8; ++i) t = fmaxf(t, r[i]);
    atomicMax(amax + (int64_t)lk * heads + hd, __float_as_uint(t));
  }
}

__global__ void __launch_bounds__(256)
kv_quant_kernel(const __nv_bfloat16* __restrict__ stage, int heads, int slots, int c_max,
                const unsigned int* __restrict__ amax, float* __restrict__ scale, unsigned char* __restrict__ cache) {
  const int hd = blockIdx.x, b = blockIdx.y, lk = blockIdx.z;
  const int B = gridDim.y;
  const float am = __uint_as_float(amax[(int64_t)lk * heads + hd]);
  const float sc = fmaxf(am * kHeadRoom, 1e-12f) / kE4M3Max;
  if (b == 0 && threadIdx.x == 0) scale[(int64_t)lk * heads + hd] = sc;
  const float inv = 1.0f / sc;
  const __nv_bfloat16* src = stage + (((int64_t)lk * B + b) * heads + hd) * (int64_t)slots * D;
  unsigned char* dst = cache + (((int64_t)lk * B + b) * heads + hd) * (int64_t)c_max * D;
  for (int i = threadIdx.x; i < slots * D; i += blockDim.x) dst[i] = f32_to_e4m3(__bfloat162float(src[i]) * inv);
}

}  // namespace

bool attention_fp8_supported(int d) { return d == D; }

int attention_fp8_fused(const float* qkv_part, int S, int64_t part_stride, int ncols, float scale, int slot_base,
                        const int* slot_dyn, void* ck, void* cv, const float* k_scale, const float* v_scale, int B,
                        int heads, int c_stride, int ctx_base, const int* ctx_dyn, const int* key_mask, int mask_len,
                        void* out, float* scratch, cudaStream_t st) {
  Fp8Fuse qf = {};
  qf.part = qkv_part; qf.S = S; qf.stride = part_stride; qf.ncols = ncols; qf.scale = scale;
  qf.slot_base = slot_base; qf.slot_dyn = slot_dyn;
  return launch_fp8(qf, nullptr, ck, cv, k_scale, v_scale, B, heads, c_stride, ctx_base, ctx_dyn, key_mask, mask_len,
                    out, scratch, st);
}

int attention_fp8(const void* q, const void* ck, const void* cv, const float* k_scale, const float* v_scale, int B,
                  int heads, int c_stride, int ctx_base, const int* ctx_dyn, const int* key_mask, int mask_len,
                  void* out, float* scratch, cudaStream_t st) {
  Fp8Fuse qf = {};
  return launch_fp8(qf, q, const_cast<void*>(ck), const_cast<void*>(cv), k_scale, v_scale, B, heads, c_stride,
                    ctx_base, ctx_dyn, key_mask, mask_len, out, scratch, st);
}

// stage [n_lk][B][heads][slots][64] bf16 -> cache [n_lk][B][heads][c_max][64] e4m3 rows [0, slots), scale [n_lk][heads];
// amax_ws: n_lk * heads uint32 of workspace
int kv_quantize_fp8(const void* stage, int n_lk, int B, int heads, int slots, int c_max, unsigned int* amax_ws,
                    float* scale, void* cache, cudaStream_t st) {
  PTTS_CHECK_ARG(stage && amax_ws && scale && cache && n_lk > 0 && B > 0 && heads > 0 && slots > 0 && slots <= c_max,
                 "kv_quantize_fp8: bad arguments");
  cudaError_t e = cudaMemsetAsync(amax_ws, 0, sizeof(unsigned int) * (size_t)n_lk * heads, st);
  if (e != cudaSuccess) {
    set_error("kv_quantize_fp8 memset: %s", cudaGetErrorString(e));
    return PTTS_E_LAUNCH;
  }
  const dim3 grid(heads, B, n_lk);
  kv_amax_kernel<<<grid, 256, 0, st>>>((const __nv_bfloat16*)stage, heads, slots, amax_ws);
  PTTS_CHECK_LAUNCH("kv_amax");
  kv_quant_kernel<<<grid, 256, 0, st>>>((const __nv_bfloat16*)stage, heads, slots, c_max, amax_ws, scale,
                                        (unsigned char*)cache);
  PTTS_CHECK_LAUNCH("kv_quant");
  return PTTS_OK;
}

}  // namespace ptts
