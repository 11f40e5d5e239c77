// DAC encode for audio-prompt continuation: waveform -> codes.  Replaces
// /root/reference/parler_tts/dac_wrapper/modeling_dac.py:25-96 (DACModel.encode -> descript DAC.encode), called from
// /root/reference/parler_tts/modeling_parler_tts.py:2247-2292.  Two new pieces next to the conv kernels the decoder
// already has (the encoder's residual units reuse ptts_dac_conv1d):
//   * a strided / arbitrary-kernel conv1d (first conv k7, down-sampling convs k = 2s stride s, last conv k3);
//   * residual vector quantisation: per codebook in_proj (latent -> 8), nearest l2-normalised codebook row, subtract
//     out_proj(codebook row) from the residual (transformers models/dac/modeling_dac.py:122-170, 281-343).
// Prompts are seconds long and encoded once per call: these are straightforward fp32 CUDA-core kernels.
#include "ops.cuh"

namespace ptts {

namespace {

__device__ __forceinline__ float snake_e(float v, float a) {
  const float s = sinf(a * v);
  return v + (1.0f / (a + 1e-9f)) * s * s;
}

// y[b][co][to] = bias[co] + sum_ci sum_k w[co][ci][k] * act(x)[b][ci][to*stride - pad + k]
// Block tile: 64 output channels x 64 output steps, 256 threads x (4 co x 4 t), 8 input channels per stage.
__global__ void __launch_bounds__(256)
dac_conv1d_strided_kernel(const float* __restrict__ x, int Cin, int T, const float* __restrict__ w,
                          const float* __restrict__ bias, int Cout, int ks, int stride, int pad,
                          const float* __restrict__ alpha, int Tout, float* __restrict__ y) {
  constexpr int CO = 64, TT = 64, CI = 8;
  extern __shared__ float sm[];
  const int XW = (TT - 1) * stride + ks;
  float* xs = sm;             // [CI][XW]
  float* ws = sm + CI * XW;   // [CI][ks][CO]
  const int b = blockIdx.z, co0 = blockIdx.y * CO, t0 = blockIdx.x * TT;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const float* xb = x + (int64_t)b * Cin * T;
  float acc[4][4] = {};
  for (int ci0 = 0; ci0 < Cin; ci0 += CI) {
    for (int i = threadIdx.x; i < CI * XW; i += 256) {
      const int ci = i / XW, tt = i % XW;
      const int t = t0 * stride - pad + tt, c = ci0 + ci;
      float v = 0.f;
      if (c < Cin && t >= 0 && t < T) {
        v = xb[(int64_t)c * T + t];
        if (alpha) v = snake_e(v, alpha[c]);
      }
      xs[i] = v;
    }
    for (int i = threadIdx.x; i < CI * ks * CO; i += 256) {
      const int co = i % CO, k = (i / CO) % ks, ci = i / (CO * ks);
      const int c = ci0 + ci, o = co0 + co;
      ws[i] = (c < Cin && o < Cout) ? w[((int64_t)o * Cin + c) * ks + k] : 0.f;
    }
    __syncthreads();
    for (int ci = 0; ci < CI; ++ci) {
      for (int k = 0; k < ks; ++k) {
        float wv[4], xv[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) wv[i] = ws[(ci * ks + k) * CO + ty * 4 + i];
#pragma unroll
        for (int j = 0; j < 4; ++j) xv[j] = xs[ci * XW + (tx * 4 + j) * stride + k];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(wv[i], xv[j], acc[i][j]);
      }
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int o = co0 + ty * 4 + i;
    if (o >= Cout) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int t = t0 + tx * 4 + j;
      if (t < Tout) y[((int64_t)b * Cout + o) * Tout + t] = acc[i][j] + (bias ? bias[o] : 0.f);
    }
  }
}

// Residual VQ over a tile of TF = 32 consecutive frames of one utterance; the residual tile [C][32] lives in shared
// memory across all codebooks.  256 threads.
//   e      = in_proj(residual)                               [D][32]
//   code   = argmax_j -(|e^|^2 - 2 e^.c^_j) + |c^_j|^2        (e^, c^ l2-normalised; first index on ties)
//   resid -= out_proj(e + (codebook[code] - e))               (the straight-through form the reference evaluates)
constexpr int VQ_TF = 32, VQ_MAXD = 16;
__global__ void __launch_bounds__(256)
dac_vq_encode_kernel(const float* __restrict__ z, int C, int F, int K, int D, int n_codes,
                     const float* __restrict__ w_in,    // [K][D][C]
                     const float* __restrict__ b_in,    // [K][D]
                     const float* __restrict__ cb,      // [K][n_codes][D] raw codebook
                     const float* __restrict__ cbn,     // [K][n_codes][D] l2-normalised rows
                     const float* __restrict__ w_out,   // [K][C][D]
                     const float* __restrict__ b_out,   // [K][C]
                     int64_t* __restrict__ codes) {     // [B][K][F]
  extern __shared__ float sm[];
  float* res = sm;                          // [C][TF]
  float* e = res + (size_t)C * VQ_TF;       // [D][TF] projected latents
  float* en = e + VQ_MAXD * VQ_TF;          // [D][TF] normalised
  float* zq = en + VQ_MAXD * VQ_TF;         // [D][TF] straight-through quantised latents
  float* bval = zq + VQ_MAXD * VQ_TF;       // [8][TF]
  int* bidx = reinterpret_cast<int*>(bval + 8 * VQ_TF);  // [8][TF]
  __shared__ float l2s[VQ_TF];
  const int b = blockIdx.y, f0 = blockIdx.x * VQ_TF, tid = threadIdx.x;
  const int fl = tid & 31, grp = tid >> 5;  // frame lane, group 0..7
  const float* zb = z + (int64_t)b * C * F;
  for (int i = tid; i < C * VQ_TF; i += 256) {
    const int c = i >> 5, f = f0 + (i & 31);
    res[i] = f < F ? zb[(int64_t)c * F + f] : 0.f;
  }
  __syncthreads();
  for (int q = 0; q < K; ++q) {
    // ---- in_proj: thread (d = grp, grp + 8, ..; frame fl)
    for (int d = grp; d < D; d += 8) {
      const float* wr = w_in + ((int64_t)q * D + d) * C;
      float a = 0.f;
      for (int c = 0; c < C; ++c) a = fmaf(wr[c], res[c * VQ_TF + fl], a);
      e[d * VQ_TF + fl] = a + b_in[q * D + d];
    }
    __syncthreads();
    if (tid < VQ_TF) {  // F.normalize: x / max(||x||, 1e-12)
      float ss = 0.f;
      for (int d = 0; d < D; ++d) ss = fmaf(e[d * VQ_TF + tid], e[d * VQ_TF + tid], ss);
      const float inv = 1.0f / fmaxf(sqrtf(ss), 1e-12f);
      float l2 = 0.f;
      for (int d = 0; d < D; ++d) {
        const float v = e[d * VQ_TF + tid] * inv;
        en[d * VQ_TF + tid] = v;
        l2 = fmaf(v, v, l2);
      }
      l2s[tid] = l2;
    }
    __syncthreads();
    // ---- nearest codebook row: group g scans rows [g * n/8, (g+1) * n/8) in ascending order
    {
      float ev[VQ_MAXD];
      for (int d = 0; d < D; ++d) ev[d] = en[d * VQ_TF + fl];
      const float l2 = l2s[fl];
      const int per = (n_codes + 7) / 8;
      const int j0 = grp * per, j1 = min(n_codes, j0 + per);
      float best = -INFINITY;
      int bj = j0;
      for (int j = j0; j < j1; ++j) {
        const float* cr = cbn + ((int64_t)q * n_codes + j) * D;
        float dot = 0.f, c2 = 0.f;
        for (int d = 0; d < D; ++d) {
          const float cv = cr[d];
          dot = fmaf(ev[d], cv, dot);
          c2 = fmaf(cv, cv, c2);
        }
        const float dist = -(l2 - 2.0f * dot) + c2;
        if (dist > best) {
          best = dist;
          bj = j;
        }
      }
      bval[grp * VQ_TF + fl] = best;
      bidx[grp * VQ_TF + fl] = bj;
    }
    __syncthreads();
    if (tid < VQ_TF) {
      float best = bval[tid];
      int bj = bidx[tid];
      for (int g = 1; g < 8; ++g)
        if (bval[g * VQ_TF + tid] > best) {  // strict: the lowest index wins a tie, like torch.max
          best = bval[g * VQ_TF + tid];
          bj = bidx[g * VQ_TF + tid];
        }
      if (f0 + tid < F) codes[((int64_t)b * K + q) * F + f0 + tid] = bj;
      const float* cr = cb + ((int64_t)q * n_codes + bj) * D;
      for (int d = 0; d < D; ++d) {
        const float p = e[d * VQ_TF + tid];
        zq[d * VQ_TF + tid] = p + (cr[d] - p);
      }
    }
    __syncthreads();
    // ---- residual -= out_proj(zq)
    if (q + 1 < K) {
      float zv[VQ_MAXD];
      for (int d = 0; d < D; ++d) zv[d] = zq[d * VQ_TF + fl];
      for (int c = grp; c < C; c += 8) {
        const float* wr = w_out + ((int64_t)q * C + c) * D;
        float a = 0.f;
        for (int d = 0; d < D; ++d) a = fmaf(wr[d], zv[d], a);
        res[c * VQ_TF + fl] -= a + b_out[q * C + c];
      }
      __syncthreads();
    }
  }
}

}  // namespace

}  // namespace ptts

using namespace ptts;

extern "C" int ptts_dac_conv1d_strided(const float* x, int32_t B, int32_t Cin, int32_t T, const float* w,
                                       const float* bias, int32_t Cout, int32_t ksize, int32_t stride, int32_t pad,
                                       const float* snake_alpha, float* y, ptts_stream_t s) {
  PTTS_CHECK_ARG(x && w && y, "dac_conv1d_strided: null pointer");
  PTTS_CHECK_ARG(B > 0 && Cin > 0 && Cout > 0 && T > 0 && ksize > 0 && ksize <= 32 && stride > 0 && stride <= 16 &&
                     pad >= 0,
                 "dac_conv1d_strided: bad dims");
  const int Tout = (T + 2 * pad - ksize) / stride + 1;
  PTTS_CHECK_ARG(Tout > 0, "dac_conv1d_strided: input shorter than the kernel");
  const size_t smem = sizeof(float) * (size_t)(8 * ((64 - 1) * stride + ksize) + 8 * ksize * 64);
  PTTS_CHECK_ARG(smem <= 200 * 1024, "dac_conv1d_strided: tile does not fit shared memory");
  if (smem > 48 * 1024)
    cudaFuncSetAttribute(dac_conv1d_strided_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  dim3 grid((Tout + 63) / 64, (Cout + 63) / 64, B);
  dac_conv1d_strided_kernel<<<grid, 256, smem, as_stream(s)>>>(x, Cin, T, w, bias, Cout, ksize, stride, pad,
                                                               snake_alpha, Tout, y);
  PTTS_CHECK_LAUNCH("dac_conv1d_strided");
  return PTTS_OK;
}

extern "C" int ptts_dac_vq_encode(const float* z, int32_t B, int32_t C, int32_t F, int32_t K, int32_t D,
                                  int32_t codebook_size, const float* w_in, const float* b_in, const float* codebook,
                                  const float* codebook_normalized, const float* w_out, const float* b_out,
                                  int64_t* codes, ptts_stream_t s) {
  PTTS_CHECK_ARG(z && w_in && b_in && codebook && codebook_normalized && w_out && b_out && codes,
                 "dac_vq_encode: null pointer");
  PTTS_CHECK_ARG(B > 0 && C > 0 && F > 0 && K > 0 && D > 0 && D <= VQ_MAXD && codebook_size > 0,
                 "dac_vq_encode: bad dims");
  const size_t smem = sizeof(float) * ((size_t)C * VQ_TF + 3 * VQ_MAXD * VQ_TF + 16 * VQ_TF);
  PTTS_CHECK_ARG(smem <= 200 * 1024, "dac_vq_encode: latent dimension too large for the residual tile");
  if (smem > 48 * 1024)
    cudaFuncSetAttribute(dac_vq_encode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  dim3 grid((F + VQ_TF - 1) / VQ_TF, B);
  dac_vq_encode_kernel<<<grid, 256, smem, as_stream(s)>>>(z, C, F, K, D, codebook_size, w_in, b_in, codebook,
                                                          codebook_normalized, w_out, b_out, codes);
  PTTS_CHECK_LAUNCH("dac_vq_encode");
  return PTTS_OK;
}
