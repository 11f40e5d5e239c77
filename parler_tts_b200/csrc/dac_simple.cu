// DAC decoder, "simple" fp32 kernel family (CUDA cores): codebook dequant, dilated conv1d with fused Snake prologue,
// bias / residual / tanh epilogue, and transposed conv.  Arithmetic of descript-audio-codec as called from
// /root/reference/parler_tts/dac_wrapper/modeling_dac.py:130-131 (see oracle/dac_oracle.py for the restatement).
// Activations are channel-major [B][C][T] with time contiguous, so every global access is coalesced along time.
#include "ops.cuh"

namespace ptts {

__device__ __forceinline__ float snake_f(float x, float alpha, float inv_alpha) {
  const float s = sinf(alpha * x);
  return x + inv_alpha * s * s;
}

// z[b][c][f] = sum_i tables[i][codes[b][i][f]][c]   (tables already contain out_proj_i . codebook_i + bias_i)
// block = 32 frames x 32 channels tile, transposed through shared memory so both sides stay coalesced
__global__ void dac_dequant_kernel(const int64_t* __restrict__ codes, int K, int F, int64_t ld_b, int64_t ld_k,
                                   const float* __restrict__ tables, const float* __restrict__ bias_sum, int cbs,
                                   int latent, float* __restrict__ z) {
  __shared__ float tile[32][33];
  __shared__ int64_t cd[PTTS_MAX_CODEBOOKS][32];
  const int b = blockIdx.z, f0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int tx = threadIdx.x, ty = threadIdx.y;  // 32 x 8
  for (int i = ty; i < K; i += 8) {
    const int f = f0 + tx;
    int64_t v = (f < F) ? codes[(int64_t)b * ld_b + (int64_t)i * ld_k + f] : 0;
    if (v < 0 || v >= cbs) v = 0;  // caller filters invalid frames (modeling_parler_tts.py:2668); stay in bounds
    cd[i][tx] = v;
  }
  __syncthreads();
  for (int fr = ty; fr < 32; fr += 8) {  // thread (tx = channel, fr = frame)
    float acc = 0.f;
    const int c = c0 + tx;
    if (c < latent)
      for (int i = 0; i < K; ++i) acc += tables[((int64_t)i * cbs + cd[i][fr]) * latent + c];
    if (bias_sum && c < latent) acc += bias_sum[c];
    tile[fr][tx] = acc;
  }
  __syncthreads();
  for (int cr = ty; cr < 32; cr += 8) {  // thread (tx = frame, cr = channel)
    const int f = f0 + tx, c = c0 + cr;
    if (f < F && c < latent) z[((int64_t)b * latent + c) * F + f] = tile[tx][cr];
  }
}

// Direct conv1d, "same" padding.  Block tile: 64 output channels x 64 time steps, 256 threads x (4 co x 4 t).
template <int KS>
__global__ void __launch_bounds__(256)
dac_conv1d_kernel(const float* __restrict__ x, int Cin, int T, const float* __restrict__ w,
                  const float* __restrict__ bias, int Cout, int dil, const float* __restrict__ alpha,
                  const float* residual, int tanh_out, float* y) {
  constexpr int CO = 64, TT = 64, CI = 8;
  extern __shared__ float sm[];
  const int halo = dil * (KS - 1) / 2;
  const int XW = TT + 2 * halo;
  float* xs = sm;                 // [CI][XW]
  float* ws = sm + CI * XW;       // [CI][KS][CO]
  const int b = blockIdx.z, co0 = blockIdx.y * CO, t0 = blockIdx.x * TT;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;  // tx -> time quad, ty -> cout quad
  const float* xb = x + (int64_t)b * Cin * T;
  float acc[4][4] = {};
  for (int ci0 = 0; ci0 < Cin; ci0 += CI) {
    for (int i = threadIdx.x; i < CI * XW; i += 256) {
      const int ci = i / XW, tt = i % XW;
      const int t = t0 - halo + tt, c = ci0 + ci;
      float v = 0.f;
      if (c < Cin && t >= 0 && t < T) {
        v = xb[(int64_t)c * T + t];
        if (alpha) {
          const float a = alpha[c];
          v = snake_f(v, a, 1.0f / (a + 1e-9f));
        }
      }
      xs[i] = v;
    }
    for (int i = threadIdx.x; i < CI * KS * CO; i += 256) {
      const int co = i % CO, k = (i / CO) % KS, ci = i / (CO * KS);
      const int c = ci0 + ci, o = co0 + co;
      ws[i] = (c < Cin && o < Cout) ? w[((int64_t)o * Cin + c) * KS + k] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int ci = 0; ci < CI; ++ci) {
#pragma unroll
      for (int k = 0; k < KS; ++k) {
        float wv[4], xv[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) wv[i] = ws[(ci * KS + k) * CO + ty * 4 + i];
#pragma unroll
        for (int j = 0; j < 4; ++j) xv[j] = xs[ci * XW + tx * 4 + j + k * dil];
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
      if (t >= T) continue;
      const int64_t idx = ((int64_t)b * Cout + o) * T + t;
      float v = acc[i][j] + (bias ? bias[o] : 0.f);
      if (residual) v += residual[idx];
      if (tanh_out) v = tanhf(v);
      y[idx] = v;
    }
  }
}

// Transposed conv, kernel 2*stride, padding ceil(stride/2): every output sample has exactly two taps per input channel.
// w layout [Cin][Cout][2*stride].  Block tile 64 co x 64 t_out.
__global__ void __launch_bounds__(256)
dac_convtr1d_kernel(const float* __restrict__ x, int Cin, int T, const float* __restrict__ w,
                    const float* __restrict__ bias, int Cout, int stride, const float* __restrict__ alpha,
                    float* __restrict__ y) {
  constexpr int CO = 64, TT = 64, CI = 8;
  extern __shared__ float sm[];
  const int KS = 2 * stride, pad = (stride + 1) / 2, Tout = T * stride;
  const int b = blockIdx.z, co0 = blockIdx.y * CO, t0 = blockIdx.x * TT;
  // input window that can reach outputs [t0, t0+TT): q in [(t0+pad)/stride - 1, (t0+TT-1+pad)/stride]
  const int q_lo = (t0 + pad) / stride - 1;
  const int XW = TT / stride + 3;
  float* xs = sm;              // [CI][XW]
  float* ws = sm + CI * XW;    // [CI][KS][CO]
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const float* xb = x + (int64_t)b * Cin * T;
  float acc[4][4] = {};
  for (int ci0 = 0; ci0 < Cin; ci0 += CI) {
    for (int i = threadIdx.x; i < CI * XW; i += 256) {
      const int ci = i / XW, qq = i % XW;
      const int q = q_lo + qq, c = ci0 + ci;
      float v = 0.f;
      if (c < Cin && q >= 0 && q < T) {
        v = xb[(int64_t)c * T + q];
        if (alpha) {
          const float a = alpha[c];
          v = snake_f(v, a, 1.0f / (a + 1e-9f));
        }
      }
      xs[i] = v;
    }
    for (int i = threadIdx.x; i < CI * KS * CO; i += 256) {
      const int co = i % CO, k = (i / CO) % KS, ci = i / (CO * KS);
      const int c = ci0 + ci, o = co0 + co;
      ws[i] = (c < Cin && o < Cout) ? w[((int64_t)c * Cout + o) * KS + k] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int t = t0 + tx * 4 + j;
      const int q = (t + pad) / stride, ph = (t + pad) % stride;
      const int qi = q - q_lo;
#pragma unroll
      for (int ci = 0; ci < CI; ++ci) {
        const float x0 = xs[ci * XW + qi], x1 = xs[ci * XW + qi - 1];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          acc[i][j] = fmaf(x0, ws[(ci * KS + ph) * CO + ty * 4 + i], acc[i][j]);
          acc[i][j] = fmaf(x1, ws[(ci * KS + ph + stride) * CO + ty * 4 + i], acc[i][j]);
        }
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
      if (t >= Tout) continue;
      y[((int64_t)b * Cout + o) * Tout + t] = acc[i][j] + (bias ? bias[o] : 0.f);
    }
  }
}

}  // namespace ptts

using namespace ptts;

extern "C" int ptts_dac_dequant(const int64_t* codes, int32_t B, int32_t K, int32_t F, int64_t ld_b, int64_t ld_k,
                                const float* tables, const float* bias_sum, int32_t codebook_size, int32_t latent,
                                float* z, ptts_stream_t s) {
  PTTS_CHECK_ARG(codes && tables && z, "dac_dequant: null pointer");
  PTTS_CHECK_ARG(B > 0 && K > 0 && K <= PTTS_MAX_CODEBOOKS && F > 0 && latent > 0, "dac_dequant: bad dims");
  dim3 grid((F + 31) / 32, (latent + 31) / 32, B), block(32, 8);
  dac_dequant_kernel<<<grid, block, 0, as_stream(s)>>>(codes, K, F, ld_b, ld_k, tables, bias_sum, codebook_size,
                                                       latent, z);
  PTTS_CHECK_LAUNCH("dac_dequant");
  return PTTS_OK;
}

extern "C" int ptts_dac_conv1d(const float* x, int32_t B, int32_t Cin, int32_t T, const float* w, const float* bias,
                               int32_t Cout, int32_t ksize, int32_t dilation, const float* snake_alpha,
                               const float* residual, int32_t tanh_out, float* y, int32_t impl, ptts_stream_t s) {
  PTTS_CHECK_ARG(x && w && y, "dac_conv1d: null pointer");
  PTTS_CHECK_ARG(B > 0 && Cin > 0 && Cout > 0 && T > 0 && dilation > 0, "dac_conv1d: bad dims");
  PTTS_CHECK_ARG(ksize == 1 || ksize == 7, "dac_conv1d: kernel size must be 1 or 7");
  (void)impl;  // the NCW entry points are the simple family; the fast family is ptts_dac_conv1d_px (channels-last)
  const int halo = dilation * (ksize - 1) / 2;
  const size_t smem = sizeof(float) * (8 * (64 + 2 * halo) + 8 * ksize * 64);
  dim3 grid((T + 63) / 64, (Cout + 63) / 64, B);
  if (ksize == 1)
    dac_conv1d_kernel<1><<<grid, 256, smem, as_stream(s)>>>(x, Cin, T, w, bias, Cout, dilation, snake_alpha, residual,
                                                            tanh_out, y);
  else
    dac_conv1d_kernel<7><<<grid, 256, smem, as_stream(s)>>>(x, Cin, T, w, bias, Cout, dilation, snake_alpha, residual,
                                                            tanh_out, y);
  PTTS_CHECK_LAUNCH("dac_conv1d");
  return PTTS_OK;
}

extern "C" int ptts_dac_convtr1d(const float* x, int32_t B, int32_t Cin, int32_t T, const float* w, const float* bias,
                                 int32_t Cout, int32_t stride, const float* snake_alpha, float* y, int32_t impl,
                                 ptts_stream_t s) {
  PTTS_CHECK_ARG(x && w && y, "dac_convtr1d: null pointer");
  PTTS_CHECK_ARG(B > 0 && Cin > 0 && Cout > 0 && T > 0, "dac_convtr1d: bad dims");
  // (16 would need 66 KB of dynamic shared memory, above the 48 KB a launch gets without opting in: rejected here
  // instead of failing at the launch)
  PTTS_CHECK_ARG(stride == 2 || stride == 4 || stride == 8, "dac_convtr1d: stride must be 2, 4 or 8");
  (void)impl;
  const int Tout = T * stride;
  const size_t smem = sizeof(float) * (8 * (64 / stride + 3) + 8 * 2 * stride * 64);
  dim3 grid((Tout + 63) / 64, (Cout + 63) / 64, B);
  dac_convtr1d_kernel<<<grid, 256, smem, as_stream(s)>>>(x, Cin, T, w, bias, Cout, stride, snake_alpha, y);
  PTTS_CHECK_LAUNCH("dac_convtr1d");
  return PTTS_OK;
}
