// T5 text encoder (the description encoder of generate(), run once per call): replaces the third-party
// transformers T5EncoderModel that /root/reference/parler_tts/modeling_parler_tts.py:2197-2241
// (_prepare_text_encoder_kwargs_for_generation -> encoder(**kwargs), line 2229) calls.  Arithmetic restated from
// site-packages/transformers/models/t5/modeling_t5.py: T5LayerNorm 46-68 (RMS norm, no mean, no bias), T5Attention
// 189-345 (no 1/sqrt(d) scaling, bucketed relative-position bias of block 0 shared by every block, additive
// finfo.min mask, fp32 softmax), T5DenseGatedActDense 106-132 / T5DenseActDense, T5Stack final_layer_norm.
// The linears run on the same tcgen05 / simple GEMM family as the decoder; the residual stream stays fp32.
#include "ops.cuh"

namespace ptts {

namespace {

__global__ void t5_embed_kernel(const int64_t* __restrict__ ids, const float* __restrict__ table, int vocab, int D,
                                float* __restrict__ h) {
  const int row = blockIdx.x;
  int64_t id = ids[row];
  id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
  const float4* src = reinterpret_cast<const float4*>(table + id * D);
  float4* dst = reinterpret_cast<float4*>(h + (int64_t)row * D);
  for (int i = threadIdx.x; i < D / 4; i += blockDim.x) dst[i] = src[i];
}

// y = w * T(h * rsqrt(mean(h^2) + eps))   (modeling_t5.py:61-68; the cast to the weight dtype happens before the scale)
template <typename T>
__global__ void __launch_bounds__(256) t5_rmsnorm_kernel(const float* __restrict__ h, const float* __restrict__ w, int D,
                                                         float eps, T* __restrict__ y) {
  __shared__ float red[8];
  const int row = blockIdx.x, tid = threadIdx.x;
  const float* x = h + (int64_t)row * D;
  float ss = 0.f;
  for (int i = tid; i < D; i += 256) ss = fmaf(x[i], x[i], ss);
  ss = warp_sum(ss);
  if ((tid & 31) == 0) red[tid >> 5] = ss;
  __syncthreads();
  float tot = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) tot += red[i];
  const float r = rsqrtf(tot / (float)D + eps);
  for (int i = tid; i < D; i += 256) y[(int64_t)row * D + i] = from_f<T>(w[i] * to_f(from_f<T>(x[i] * r)));
}

// One warp per query row i of (b, head): scores_j = q_i . k_j + bias[bucket(j - i)][head] (+ finfo.min on masked keys),
// softmax in fp32, probabilities rounded to T like `.type_as(scores)`, out = P V.  qkv rows are [q | k | v] of width
// 3 * inner with head h at columns h * dkv.
template <typename T, int QW>
__global__ void __launch_bounds__(QW * 32)
t5_attention_kernel(const T* __restrict__ qkv, const int32_t* __restrict__ mask, const float* __restrict__ rel_bias,
                    const int32_t* __restrict__ bucket_of, int max_rel, int S, int heads, int dkv,
                    T* __restrict__ out) {
  extern __shared__ float sm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int i = blockIdx.x * QW + warp, hd = blockIdx.y, b = blockIdx.z;
  const int inner = heads * dkv;
  float* qs = sm + warp * (dkv + S);  // [dkv] query, then [S] scores
  float* sc = qs + dkv;
  if (i >= S) return;  // whole warp
  const T* base = qkv + (int64_t)b * S * 3 * inner + hd * dkv;
  for (int d = lane; d < dkv; d += 32) qs[d] = to_f(base[(int64_t)i * 3 * inner + d]);
  __syncwarp();
  const float neg = finfo_min<T>();
  float lmax = -INFINITY;
  for (int j = lane; j < S; j += 32) {
    const T* kr = base + (int64_t)j * 3 * inner + inner;
    float dot = 0.f;
    for (int d = 0; d < dkv; ++d) dot = fmaf(qs[d], to_f(kr[d]), dot);
    int rel = j - i;
    rel = rel < -max_rel ? -max_rel : (rel > max_rel ? max_rel : rel);
    float pb = to_f(from_f<T>(rel_bias[bucket_of[rel + max_rel] * heads + hd]));
    if (mask && mask[(int64_t)b * S + j] == 0) pb = to_f(from_f<T>(pb + neg));
    const float s = to_f(from_f<T>(to_f(from_f<T>(dot)) + pb));
    sc[j] = s;
    lmax = fmaxf(lmax, s);
  }
  lmax = warp_max(lmax);
  float lsum = 0.f;
  for (int j = lane; j < S; j += 32) {
    const float e = expf(sc[j] - lmax);
    sc[j] = e;
    lsum += e;
  }
  lsum = warp_sum(lsum);
  __syncwarp();
  const float inv = 1.0f / lsum;
  for (int d = lane; d < dkv; d += 32) {
    float acc = 0.f;
    const T* vr = base + 2 * inner + d;
    for (int j = 0; j < S; ++j) acc = fmaf(to_f(from_f<T>(sc[j] * inv)), to_f(vr[(int64_t)j * 3 * inner]), acc);
    out[((int64_t)b * S + i) * inner + hd * dkv + d] = from_f<T>(acc);
  }
}

// The same arithmetic with the K and V rows of (b, head) staged in shared memory once per CTA: K transposed [dkv][S + 1]
// (lanes = keys read consecutive words), V row-major [S][dkv] (lanes = dims), both already converted to float.  One CTA
// per (head, b), warp w takes queries w, w + NW, ...; every dot product and every P.V sum runs in the order of the kernel
// above, so the results are bit-identical - the 2-byte strided global loads (455 us per launch at B=64, S=48, 58 % of
// the encoder) become shared-memory reads.  Used when the staged rows fit (S <= ~380 at dkv = 64).
template <typename T, int NW>
__global__ void __launch_bounds__(NW * 32)
t5_attention_staged_kernel(const T* __restrict__ qkv, const int32_t* __restrict__ mask, const float* __restrict__ rel_bias,
                           const int32_t* __restrict__ bucket_of, int max_rel, int S, int heads, int dkv,
                           T* __restrict__ out) {
  extern __shared__ float sm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int hd = blockIdx.x, b = blockIdx.y;
  const int inner = heads * dkv, SP = S + 1;
  float* kt = sm;                       // [dkv][SP]
  float* vs = kt + (size_t)dkv * SP;    // [S][dkv]
  float* qs = vs + (size_t)S * dkv + warp * (dkv + S);  // per warp: [dkv] query, [S] scores
  float* sc = qs + dkv;
  const T* base = qkv + (int64_t)b * S * 3 * inner + hd * dkv;
  for (int e = threadIdx.x; e < S * dkv; e += NW * 32) {
    const int j = e / dkv, d = e - j * dkv;
    const T* row = base + (int64_t)j * 3 * inner;
    kt[d * SP + j] = to_f(row[inner + d]);
    vs[e] = to_f(row[2 * inner + d]);
  }
  __syncthreads();
  const float neg = finfo_min<T>();
  for (int i = warp; i < S; i += NW) {
    for (int d = lane; d < dkv; d += 32) qs[d] = to_f(base[(int64_t)i * 3 * inner + d]);
    __syncwarp();
    float lmax = -INFINITY;
    for (int j = lane; j < S; j += 32) {
      float dot = 0.f;
      for (int d = 0; d < dkv; ++d) dot = fmaf(qs[d], kt[d * SP + j], dot);
      int rel = j - i;
      rel = rel < -max_rel ? -max_rel : (rel > max_rel ? max_rel : rel);
      float pb = to_f(from_f<T>(rel_bias[bucket_of[rel + max_rel] * heads + hd]));
      if (mask && mask[(int64_t)b * S + j] == 0) pb = to_f(from_f<T>(pb + neg));
      const float sv = to_f(from_f<T>(to_f(from_f<T>(dot)) + pb));
      sc[j] = sv;
      lmax = fmaxf(lmax, sv);
    }
    lmax = warp_max(lmax);
    float lsum = 0.f;
    for (int j = lane; j < S; j += 32) {
      const float e = expf(sc[j] - lmax);
      sc[j] = e;
      lsum += e;
    }
    lsum = warp_sum(lsum);
    __syncwarp();
    const float inv = 1.0f / lsum;
    for (int d = lane; d < dkv; d += 32) {
      float acc = 0.f;
      for (int j = 0; j < S; ++j) acc = fmaf(to_f(from_f<T>(sc[j] * inv)), vs[j * dkv + d], acc);
      out[((int64_t)b * S + i) * inner + hd * dkv + d] = from_f<T>(acc);
    }
    __syncwarp();  // the next query overwrites qs / sc
  }
}

// g = T(act(a)) * b over rows of [a | b] (wi_0 | wi_1), modeling_t5.py:116-118
template <typename T>
__global__ void t5_gated_act_kernel(const T* __restrict__ ab, int64_t rows, int F, int act, T* __restrict__ g) {
  const int64_t n = rows * F;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = e / F;
    const int c = (int)(e - r * F);
    const float a = to_f(ab[r * 2 * F + c]), bb = to_f(ab[r * 2 * F + F + c]);
    g[e] = from_f<T>(to_f(from_f<T>(apply_act(a, act))) * bb);
  }
}

int t5_linear(int impl, const void* x, const void* w, int dt, int M, int N, int K, int act, const float* residual,
              void* y, int ydtype, cudaStream_t st) {
  if (impl == PTTS_IMPL_FAST && dt == PTTS_BF16 && gemm_bf16_supported(M, N, K))
    return gemm_bf16_store(x, w, M, N, K, 1.0f, act, nullptr, residual, nullptr, y, ydtype, st);
  return linear_simple(x, dt, w, dt, nullptr, M, N, K, 1.0f, act, residual, nullptr, y, ydtype, st);
}

inline size_t align256(size_t v) { return (v + 255) / 256 * 256; }

struct Workspace {
  float* h;
  char *xn, *qkv, *attn, *ff, *g;
  size_t total;
};

Workspace carve(const PttsT5* m, int64_t rows, void* ws) {
  const size_t es = dtype_size(m->wdtype);
  const int inner = m->n_heads * m->d_kv;
  const int wi = (m->gated ? 2 : 1) * m->d_ff;
  Workspace w;
  size_t o = 0;
  char* p = reinterpret_cast<char*>(ws);
  w.h = reinterpret_cast<float*>(p + o);  o += align256((size_t)rows * m->d_model * sizeof(float));
  w.xn = p + o;                            o += align256((size_t)rows * m->d_model * es);
  w.qkv = p + o;                           o += align256((size_t)rows * 3 * inner * es);
  w.attn = p + o;                          o += align256((size_t)rows * inner * es);
  w.ff = p + o;                            o += align256((size_t)rows * wi * es);
  w.g = p + o;                             o += align256((size_t)rows * m->d_ff * es);
  w.total = o;
  return w;
}

template <typename T>
int t5_encode_t(const PttsT5* m, const int64_t* ids, const int32_t* mask, int B, int S, const Workspace& w, void* out,
                int impl, cudaStream_t st) {
  constexpr int QW = 4;
  const int D = m->d_model, heads = m->n_heads, dkv = m->d_kv, F = m->d_ff, dt = m->wdtype;
  const int inner = heads * dkv, rows = B * S;
  const size_t att_smem = (size_t)QW * (dkv + S) * sizeof(float);
  PTTS_CHECK_ARG(att_smem <= 200 * 1024, "t5_encode: sequence too long for the score buffer");
  if (att_smem > 48 * 1024)
    cudaFuncSetAttribute(t5_attention_kernel<T, QW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)att_smem);
  // K / V of a (b, head) staged in shared memory when they fit (fast family; the simple family keeps the plain kernel)
  constexpr int SNW = 8;
  const size_t staged_smem = ((size_t)dkv * (S + 1) + (size_t)S * dkv + (size_t)SNW * (dkv + S)) * sizeof(float);
  const bool staged = impl == PTTS_IMPL_FAST && staged_smem <= 200 * 1024;
  if (staged && staged_smem > 48 * 1024)
    cudaFuncSetAttribute(t5_attention_staged_kernel<T, SNW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)staged_smem);
  t5_embed_kernel<<<rows, 192, 0, st>>>(ids, m->embed, m->vocab, D, w.h);
  PTTS_CHECK_LAUNCH("t5_embed");
  for (int l = 0; l < m->n_layers; ++l) {
    const PttsT5Layer& L = m->layers[l];
    t5_rmsnorm_kernel<T><<<rows, 256, 0, st>>>(w.h, L.ln0, D, m->eps, reinterpret_cast<T*>(w.xn));
    PTTS_CHECK_LAUNCH("t5_rmsnorm");
    PTTS_TRY(t5_linear(impl, w.xn, L.w_qkv, dt, rows, 3 * inner, D, PTTS_ACT_NONE, nullptr, w.qkv, dt, st));
    if (staged)
      t5_attention_staged_kernel<T, SNW><<<dim3(heads, B), SNW * 32, staged_smem, st>>>(
          reinterpret_cast<const T*>(w.qkv), mask, m->rel_bias, m->bucket_of, m->max_rel, S, heads, dkv,
          reinterpret_cast<T*>(w.attn));
    else
      t5_attention_kernel<T, QW><<<dim3((S + QW - 1) / QW, heads, B), QW * 32, att_smem, st>>>(
          reinterpret_cast<const T*>(w.qkv), mask, m->rel_bias, m->bucket_of, m->max_rel, S, heads, dkv,
          reinterpret_cast<T*>(w.attn));
    PTTS_CHECK_LAUNCH("t5_attention");
    PTTS_TRY(t5_linear(impl, w.attn, L.w_o, dt, rows, D, inner, PTTS_ACT_NONE, w.h, w.h, PTTS_F32, st));
    t5_rmsnorm_kernel<T><<<rows, 256, 0, st>>>(w.h, L.ln1, D, m->eps, reinterpret_cast<T*>(w.xn));
    PTTS_CHECK_LAUNCH("t5_rmsnorm");
    const void* ffn_in = w.ff;
    if (m->gated) {
      PTTS_TRY(t5_linear(impl, w.xn, L.w_wi, dt, rows, 2 * F, D, PTTS_ACT_NONE, nullptr, w.ff, dt, st));
      const int64_t n = (int64_t)rows * F;
      const int blocks = (int)((n + 255) / 256 < 148 * 16 ? (n + 255) / 256 : 148 * 16);
      t5_gated_act_kernel<T><<<blocks, 256, 0, st>>>(reinterpret_cast<const T*>(w.ff), rows, F, m->act_fn,
                                                      reinterpret_cast<T*>(w.g));
      PTTS_CHECK_LAUNCH("t5_gated_act");
      ffn_in = w.g;
    } else {
      PTTS_TRY(t5_linear(impl, w.xn, L.w_wi, dt, rows, F, D, m->act_fn, nullptr, w.ff, dt, st));
    }
    PTTS_TRY(t5_linear(impl, ffn_in, L.w_wo, dt, rows, D, F, PTTS_ACT_NONE, w.h, w.h, PTTS_F32, st));
  }
  t5_rmsnorm_kernel<T><<<rows, 256, 0, st>>>(w.h, m->ln_f, D, m->eps, reinterpret_cast<T*>(out));
  PTTS_CHECK_LAUNCH("t5_rmsnorm");
  return PTTS_OK;
}

}  // namespace

}  // namespace ptts

using namespace ptts;

extern "C" int64_t ptts_t5_workspace_bytes(const PttsT5* m, int32_t rows) {
  if (!m || rows <= 0) return 0;
  return (int64_t)carve(m, rows, nullptr).total;
}

extern "C" int ptts_t5_encode(const PttsT5* m, const int64_t* ids, const int32_t* mask, int32_t B, int32_t S,
                              void* workspace, void* out, int32_t impl, ptts_stream_t s) {
  PTTS_CHECK_ARG(m && m->layers && m->embed && m->rel_bias && m->bucket_of && m->ln_f, "t5_encode: null model tensor");
  PTTS_CHECK_ARG(m->n_layers > 0 && m->d_model > 0 && m->d_model % 4 == 0 && m->n_heads > 0 && m->d_kv > 0 &&
                     m->d_ff > 0 && m->max_rel > 0,
                 "t5_encode: bad model dims");
  PTTS_CHECK_ARG(m->wdtype == PTTS_F32 || m->wdtype == PTTS_BF16, "t5_encode: bad wdtype");
  PTTS_CHECK_ARG(ids && workspace && out && B > 0 && S > 0, "t5_encode: bad arguments");
  const Workspace w = carve(m, (int64_t)B * S, workspace);
  if (m->wdtype == PTTS_F32) return t5_encode_t<float>(m, ids, mask, B, S, w, out, impl, as_stream(s));
  return t5_encode_t<__nv_bfloat16>(m, ids, mask, B, S, w, out, impl, as_stream(s));
}
