// "Simple" kernel family: straightforward, dtype-generic (fp32 / bf16 storage, fp32 accumulate) kernels.
// They serve the fp32 parity configuration (token-exact greedy vs the oracle) and validate the fast sm_100a kernels.
// Reference ops replaced: see include/parler_b200.h next to each entry point.
#include "common.cuh"
#include "ops.cuh"

namespace ptts {

// ------------------------------------------------------------------------------------------------ embeddings
// modeling_parler_tts.py:836 (sum over codebooks, in codebook order), 839-840 (prompt prefix), 884-886 (positions)
__global__ void embed_prefill_kernel(const float* __restrict__ embed, const float* __restrict__ pos, int K, int V1,
                                     int H, const int64_t* __restrict__ ids, int64_t ld_ids,
                                     const int64_t* __restrict__ mask, int64_t ld_mask, int ids_batch, int T0,
                                     const float* __restrict__ prompt_table, const int64_t* __restrict__ prompt_ids,
                                     const float* __restrict__ prompt_hidden, int P, int pos0, float* __restrict__ h) {
  const int T = P + T0;
  const int b = blockIdx.x / T, t = blockIdx.x % T;
  const int bi = b % ids_batch;  // classifier-free guidance: the unconditional half reuses the same ids
  float* out = h + (int64_t)blockIdx.x * H;
  const float* prow = pos + (int64_t)(pos0 + t) * H;
  if (t < P) {
    const float* src = prompt_hidden ? prompt_hidden + ((int64_t)b * P + t) * H
                                     : prompt_table + prompt_ids[(int64_t)b * P + t] * (int64_t)H;
    for (int c = threadIdx.x; c < H; c += blockDim.x) out[c] = src[c] + prow[c];
  } else {
    for (int c = threadIdx.x; c < H; c += blockDim.x) {
      float acc = 0.f;
      for (int k = 0; k < K; ++k) {
        int64_t id = ids[((int64_t)bi * K + k) * ld_ids + (t - P)];
        if (mask) {  // apply_delay_pattern_mask, modeling_parler_tts.py:68-74
          const int64_t mv = mask[((int64_t)bi * K + k) * ld_mask + (t - P)];
          if (mv != -1) id = mv;
        }
        acc += embed[((int64_t)k * V1 + id) * H + c];
      }
      out[c] = acc + prow[c];
    }
  }
}

int embed_prefill(const PttsModel* m, const int64_t* ids, int64_t ld_ids, const int64_t* mask, int64_t ld_mask, int B,
                  int ids_batch, int T0, const float* prompt_table,
                  const int64_t* prompt_ids, const float* prompt_hidden, int P, float* h, cudaStream_t st, int pos0) {
  embed_prefill_kernel<<<B * (P + T0), 256, 0, st>>>(m->embed, m->pos, m->n_codebooks, m->vocab + 1, m->hidden, ids,
                                                     ld_ids, mask, ld_mask, ids_batch, T0, prompt_table, prompt_ids,
                                                     prompt_hidden, P, pos0, h);
  PTTS_CHECK_LAUNCH("embed_prefill");
  return PTTS_OK;
}

// ------------------------------------------------------------------------------------------------ LayerNorm
template <typename TY>
__global__ void layernorm_kernel(const float* __restrict__ x, const float* __restrict__ g,
                                 const float* __restrict__ bta, int H, float eps, int64_t in_stride,
                                 int64_t in_offset, TY* __restrict__ y) {
  const float* xr = x + (int64_t)blockIdx.x * in_stride + in_offset;
  TY* yr = y + (int64_t)blockIdx.x * H;
  __shared__ float red[32];
  __shared__ float stat[2];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  float s = 0.f;
  for (int c = threadIdx.x; c < H; c += blockDim.x) s += xr[c];
  s = warp_sum(s);
  if (lane == 0) red[warp] = s;
  __syncthreads();
  if (warp == 0) {
    float v = lane < nw ? red[lane] : 0.f;
    v = warp_sum(v);
    if (lane == 0) stat[0] = v / (float)H;
  }
  __syncthreads();
  const float mean = stat[0];
  float q = 0.f;
  for (int c = threadIdx.x; c < H; c += blockDim.x) {
    const float d = xr[c] - mean;
    q += d * d;
  }
  q = warp_sum(q);
  if (lane == 0) red[warp] = q;
  __syncthreads();
  if (warp == 0) {
    float v = lane < nw ? red[lane] : 0.f;
    v = warp_sum(v);
    if (lane == 0) stat[1] = rsqrtf(v / (float)H + eps);
  }
  __syncthreads();
  const float rstd = stat[1];
  for (int c = threadIdx.x; c < H; c += blockDim.x) yr[c] = from_f<TY>((xr[c] - mean) * rstd * g[c] + bta[c]);
}

int layernorm(const float* x, const float* g, const float* b, int rows, int H, float eps, int64_t in_stride,
              int64_t in_offset, void* y, int ydtype, cudaStream_t st) {
  if (rows == 0) return PTTS_OK;
  if (ydtype == PTTS_F32)
    layernorm_kernel<float><<<rows, 256, 0, st>>>(x, g, b, H, eps, in_stride, in_offset, (float*)y);
  else
    layernorm_kernel<__nv_bfloat16><<<rows, 256, 0, st>>>(x, g, b, H, eps, in_stride, in_offset, (__nv_bfloat16*)y);
  PTTS_CHECK_LAUNCH("layernorm");
  return PTTS_OK;
}

// ------------------------------------------------------------------------------------------------ Linear (SIMT)
// Y[M][N] = epi(X[M][K] . W[N][K]^T); 64x64 block tile, 16-deep K slices, 4x4 register tile per thread.
template <typename TX, typename TW, typename TY>
__global__ void __launch_bounds__(256)
linear_simt_kernel(const TX* __restrict__ X, const TW* __restrict__ W, const float* __restrict__ bias, int M, int N,
                   int K, float alpha, int act, const float* residual, const int* __restrict__ row_mask, TY* Y) {
  constexpr int BM = 64, BN = 64, BK = 16;
  __shared__ float Xs[BK][BM + 4];
  __shared__ float Ws[BK][BN + 4];
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int lr = threadIdx.x >> 2, lk = (threadIdx.x & 3) * 4;  // loader: row, first k of 4
  float acc[4][4] = {};
  for (int k0 = 0; k0 < K; k0 += BK) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int kk = k0 + lk + i;
      const int m = m0 + lr, n = n0 + lr;
      Xs[lk + i][lr] = (m < M && kk < K) ? to_f(X[(int64_t)m * K + kk]) : 0.f;
      Ws[lk + i][lr] = (n < N && kk < K) ? to_f(W[(int64_t)n * K + kk]) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = Xs[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = Ws[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m >= M) continue;
    const float rm = row_mask ? (float)row_mask[m] : 1.0f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= N) continue;
      float v = acc[i][j] * alpha;
      if (bias) v += bias[n];
      v = apply_act(v, act);
      if (row_mask) v *= rm;
      if (residual) v += residual[(int64_t)m * N + n];
      Y[(int64_t)m * N + n] = from_f<TY>(v);
    }
  }
}

template <typename TX, typename TW>
static void linear_simt_y(const void* x, const void* w, const float* bias, int M, int N, int K, float alpha, int act,
                          const float* residual, const int* row_mask, void* y, int ydtype, cudaStream_t st) {
  dim3 grid((N + 63) / 64, (M + 63) / 64);
  if (ydtype == PTTS_F32)
    linear_simt_kernel<TX, TW, float><<<grid, 256, 0, st>>>((const TX*)x, (const TW*)w, bias, M, N, K, alpha, act,
                                                            residual, row_mask, (float*)y);
  else
    linear_simt_kernel<TX, TW, __nv_bfloat16><<<grid, 256, 0, st>>>((const TX*)x, (const TW*)w, bias, M, N, K, alpha,
                                                                    act, residual, row_mask, (__nv_bfloat16*)y);
}

int linear_simple(const void* x, int xdtype, const void* w, int wdtype, const float* bias, int M, int N, int K,
                  float alpha, int act, const float* residual, const int* row_mask, void* y, int ydtype,
                  cudaStream_t st) {
  if (M == 0 || N == 0) return PTTS_OK;
  if (xdtype == PTTS_F32 && wdtype == PTTS_F32)
    linear_simt_y<float, float>(x, w, bias, M, N, K, alpha, act, residual, row_mask, y, ydtype, st);
  else if (xdtype == PTTS_F32 && wdtype == PTTS_BF16)
    linear_simt_y<float, __nv_bfloat16>(x, w, bias, M, N, K, alpha, act, residual, row_mask, y, ydtype, st);
  else if (xdtype == PTTS_BF16 && wdtype == PTTS_F32)
    linear_simt_y<__nv_bfloat16, float>(x, w, bias, M, N, K, alpha, act, residual, row_mask, y, ydtype, st);
  else
    linear_simt_y<__nv_bfloat16, __nv_bfloat16>(x, w, bias, M, N, K, alpha, act, residual, row_mask, y, ydtype, st);
  PTTS_CHECK_LAUNCH("linear_simple");
  return PTTS_OK;
}

// ------------------------------------------------------------------------------------------------ QKV scatter
// qkv [rows][3H] -> q [rows][H] * scale; k, v -> cache [B][heads][c_stride][d] at slot pos0 + t (+ *dyn).
// Replaces the torch.cat cache growth of modeling_parler_tts.py:301-304 with an indexed write.
template <typename T>
__global__ void qkv_scatter_kernel(const T* __restrict__ qkv, int H, int heads, int d, int Tq, int c_stride,
                                   int pos_base, const int* __restrict__ pos_dyn, float scale, T* __restrict__ q,
                                   T* __restrict__ ck, T* __restrict__ cv) {
  const int row = blockIdx.x;
  const int b = row / Tq, t = row % Tq;
  const int slot = dyn_value(pos_base, pos_dyn) + t;
  const T* src = qkv + (int64_t)row * 3 * H;
  for (int c = threadIdx.x; c < H; c += blockDim.x) {
    const int hd = c / d, dd = c % d;
    q[(int64_t)row * H + c] = from_f<T>(to_f(src[c]) * scale);
    const int64_t o = (((int64_t)b * heads + hd) * c_stride + slot) * d + dd;
    ck[o] = src[H + c];
    cv[o] = src[2 * H + c];
  }
}

// kv [rows=B*S][2H] -> cross cache [B][heads][S][d]
template <typename T>
__global__ void kv_scatter_kernel(const T* __restrict__ kv, int H, int heads, int d, int S, T* __restrict__ ck,
                                  T* __restrict__ cv) {
  const int row = blockIdx.x;
  const int b = row / S, s = row % S;
  const T* src = kv + (int64_t)row * 2 * H;
  for (int c = threadIdx.x; c < H; c += blockDim.x) {
    const int hd = c / d, dd = c % d;
    const int64_t o = (((int64_t)b * heads + hd) * S + s) * d + dd;
    ck[o] = src[c];
    cv[o] = src[H + c];
  }
}

template <typename T>
__global__ void scale_copy_kernel(const T* __restrict__ x, int64_t n, float scale, T* __restrict__ y) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    y[i] = from_f<T>(to_f(x[i]) * scale);
}

int qkv_scatter(const void* qkv, int dtype, int rows, int H, int heads, int d, int Tq, int c_stride, int pos_base,
                const int* pos_dyn, float scale, void* q, void* ck, void* cv, cudaStream_t st) {
  if (rows == 0) return PTTS_OK;
  if (dtype == PTTS_F32)
    qkv_scatter_kernel<float><<<rows, 256, 0, st>>>((const float*)qkv, H, heads, d, Tq, c_stride, pos_base, pos_dyn,
                                                    scale, (float*)q, (float*)ck, (float*)cv);
  else
    qkv_scatter_kernel<__nv_bfloat16><<<rows, 256, 0, st>>>((const __nv_bfloat16*)qkv, H, heads, d, Tq, c_stride,
                                                            pos_base, pos_dyn, scale, (__nv_bfloat16*)q,
                                                            (__nv_bfloat16*)ck, (__nv_bfloat16*)cv);
  PTTS_CHECK_LAUNCH("qkv_scatter");
  return PTTS_OK;
}

int kv_scatter(const void* kv, int dtype, int rows, int H, int heads, int d, int S, void* ck, void* cv,
               cudaStream_t st) {
  if (rows == 0) return PTTS_OK;
  if (dtype == PTTS_F32)
    kv_scatter_kernel<float><<<rows, 256, 0, st>>>((const float*)kv, H, heads, d, S, (float*)ck, (float*)cv);
  else
    kv_scatter_kernel<__nv_bfloat16><<<rows, 256, 0, st>>>((const __nv_bfloat16*)kv, H, heads, d, S,
                                                           (__nv_bfloat16*)ck, (__nv_bfloat16*)cv);
  PTTS_CHECK_LAUNCH("kv_scatter");
  return PTTS_OK;
}

}  // namespace ptts
