// DAC decoder, fast family: C-ABI entry points of the channels-last ("NWC", [B][T][C] fp32) path and its two small
// non-GEMM kernels (codebook lookup, Cout = 1 output conv).  Arithmetic of descript-audio-codec as called from
// /root/reference/parler_tts/dac_wrapper/modeling_dac.py:130-131; restated in oracle/dac_oracle.py:
//   conv1d k=7 dilated / k=1 : y[t][co] = sum_tap sum_ci snake(x)[t + (tap - c) * dil][ci] * W[tap][co][ci]
//   conv_transpose1d k=2s    : y[q*s + ph - pad][co] = sum_ci snake(x)[q][ci] W[ci][co][ph] + snake(x)[q-1][ci] W[ci][co][ph+s]
//                              = the same kernel with 2 taps (row offsets 0, -1) and N = s * Cout columns (ph, co).
// Every convolution of the stack is ONE implicit-GEMM kernel: the warp-specialised persistent tcgen05 kernel of
// fast_dac3.cu (M = time rows, N = output columns, K = taps x Cin; fp32 semantics through hi/mid bf16 planes).
#include "gemm_common.cuh"

namespace ptts {

using namespace gemm;

namespace {

constexpr int CIK = 32;       // input channels per chunk
constexpr int MAX_SPAN = 54;  // max (hi - lo) tap span in rows: k = 7, dilation 9

// Output conv (Cout = 1, k = 7) + tanh over channels-last fp32 input: y[t] = tanh(b + sum_k sum_c snake(x)[t+k-3][c] w[k][c]).
// HBM-bound (one read of x).  Phase 1: z[r][k] = sum_c snake(x[r][c]) w[k][c] for the 134 input rows of a 128-row tile:
// 8 threads per row (fully coalesced 128-byte reads), every element is activated once and used for all 7 taps; the 8
// partial vectors are combined by a 7-shuffle butterfly that leaves tap k in lane k.  Phase 2: y[t] = sum_k z[t+k][k].
constexpr int CO_ROWS = 128, CO_HALO = 6, CO_MAXC = 256;
__device__ __forceinline__ float snake_fast(float x, float a, float inv) {  // range-reduced SFU sine
  const float ax = a * x;
  const float k = rintf(ax * 0.15915494309189535f);
  float r = fmaf(k, -6.2831854820251465f, ax);
  r = fmaf(k, 1.7484556000744883e-7f, r);
  const float sn = __sinf(r);
  return fmaf(inv * sn, sn, x);
}
__global__ void __launch_bounds__(256)
dac_conv_out_kernel(const float* __restrict__ x, int T, int Cin, const float* __restrict__ w /*[7][Cin]*/,
                    const float* __restrict__ bias, const float* __restrict__ alpha, float* __restrict__ y) {
  __shared__ __align__(16) float ws[8 * CO_MAXC];                    // [tap 0..6, zero tap 7][Cin]
  __shared__ float zs[(CO_ROWS + CO_HALO) * 8];                      // [row][tap]
  const int b = blockIdx.y, t0 = blockIdx.x * CO_ROWS, tid = threadIdx.x, lane = tid & 31;
  const int sub = tid & 7, rg = tid >> 3;                            // 8 threads per row, 32 rows per pass
  const float* xb = x + (int64_t)b * T * Cin;
  for (int i = tid; i < 8 * Cin; i += 256) ws[i] = i < 7 * Cin ? w[i] : 0.f;
  const int nj = Cin / 32;                                           // float4 chunks per thread and row (<= 8)
  __syncthreads();
  for (int r0 = 0; r0 < CO_ROWS + CO_HALO; r0 += 32) {  // uniform trip count: the shuffles below need whole warps
    const int r = r0 + rg;
    const int t = r < CO_ROWS + CO_HALO ? t0 - 3 + r : -1;
    float acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0.f;
    if (t >= 0 && t < T) {
      const float* xr = xb + (int64_t)t * Cin + sub * 4;
      for (int j = 0; j < nj; ++j) {
        float4 v = __ldg(reinterpret_cast<const float4*>(xr + j * 32));
        if (alpha) {
          const float4 al = __ldg(reinterpret_cast<const float4*>(alpha + j * 32 + sub * 4));
          v.x = snake_fast(v.x, al.x, __frcp_rn(al.x + 1e-9f));
          v.y = snake_fast(v.y, al.y, __frcp_rn(al.y + 1e-9f));
          v.z = snake_fast(v.z, al.z, __frcp_rn(al.z + 1e-9f));
          v.w = snake_fast(v.w, al.w, __frcp_rn(al.w + 1e-9f));
        }
#pragma unroll
        for (int k = 0; k < 7; ++k) {
          const float4 wk = *reinterpret_cast<const float4*>(ws + k * Cin + j * 32 + sub * 4);
          acc[k] = fmaf(v.x, wk.x, fmaf(v.y, wk.y, fmaf(v.z, wk.z, fmaf(v.w, wk.w, acc[k]))));
        }
      }
    }
    // butterfly over the 8 lanes of the row: after step s every lane keeps the half of its vector selected by its
    // own bit; lane `sub` ends with the full sum of tap `sub`
    {
      const bool up = sub & 4;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float send = up ? acc[k] : acc[k + 4];
        const float keep = up ? acc[k + 4] : acc[k];
        acc[k] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
      }
    }
    {
      const bool up = sub & 2;
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const float send = up ? acc[k] : acc[k + 2];
        const float keep = up ? acc[k + 2] : acc[k];
        acc[k] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
      }
    }
    {
      const bool up = sub & 1;
      const float send = up ? acc[0] : acc[1];
      const float keep = up ? acc[1] : acc[0];
      acc[0] = keep + __shfl_xor_sync(0xffffffffu, send, 1);
    }
    if (r < CO_ROWS + CO_HALO) zs[r * 8 + sub] = acc[0];
  }
  (void)lane;
  __syncthreads();
  if (tid < CO_ROWS) {
    const int t = t0 + tid;
    if (t < T) {
      float o = bias ? bias[0] : 0.f;
#pragma unroll
      for (int k = 0; k < 7; ++k) o += zs[(tid + k) * 8 + k];
      y[(int64_t)b * T + t] = tanhf(o);
    }
  }
}

// z[b][f][c] = sum_i tables[i][codes[b][i][f]][c]   (tables = codebook_i x out_proj_i + bias_i, folded at load)
__global__ void __launch_bounds__(256)
dac_dequant_nwc_kernel(const int64_t* __restrict__ codes, int K, int F, int64_t ld_b, int64_t ld_k,
                       const float* __restrict__ tables, int cbs, int latent, float* __restrict__ z) {
  __shared__ int64_t cd[PTTS_MAX_CODEBOOKS];
  const int f = blockIdx.x, b = blockIdx.y;
  if (threadIdx.x < K) {
    int64_t v = codes[(int64_t)b * ld_b + (int64_t)threadIdx.x * ld_k + f];
    if (v < 0 || v >= cbs) v = 0;  // the caller filters invalid frames (modeling_parler_tts.py:2668); stay in bounds
    cd[threadIdx.x] = v;
  }
  __syncthreads();
  for (int c = threadIdx.x * 4; c < latent; c += 256 * 4) {
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int i = 0; i < K; ++i) {
      const float4 t = __ldg(reinterpret_cast<const float4*>(tables + ((int64_t)i * cbs + cd[i]) * latent + c));
      a.x += t.x; a.y += t.y; a.z += t.z; a.w += t.w;
    }
    *reinterpret_cast<float4*>(z + ((int64_t)b * F + f) * latent + c) = a;
  }
}

}  // namespace
}  // namespace ptts

using namespace ptts;

extern "C" PTTS_API int ptts_dac_tile_n(int32_t N) { return dac_tile_n(N); }

extern "C" PTTS_API int ptts_dac_set_cta_limit(int32_t n) { return dac_set_cta_limit(n); }

extern "C" PTTS_API int ptts_dac_dequant_nwc(const int64_t* codes, int32_t B, int32_t K, int32_t F, int64_t ld_b, int64_t ld_k,
                                    const float* tables, int32_t codebook_size, int32_t latent, float* z,
                                    ptts_stream_t s) {
  PTTS_CHECK_ARG(codes && tables && z, "dac_dequant_nwc: null pointer");
  PTTS_CHECK_ARG(B > 0 && K > 0 && K <= PTTS_MAX_CODEBOOKS && F > 0 && latent > 0 && latent % 4 == 0,
                 "dac_dequant_nwc: bad dims");
  dac_dequant_nwc_kernel<<<dim3(F, B), 256, 0, as_stream(s)>>>(codes, K, F, ld_b, ld_k, tables, codebook_size, latent,
                                                              z);
  PTTS_CHECK_LAUNCH("dac_dequant_nwc");
  return PTTS_OK;
}

// conv1d / transposed conv over pre-tiled weights with optional "planes" input and output (include/parler_b200.h)
extern "C" PTTS_API int ptts_dac_conv1d_px(const float* x, const void* x_planes, int32_t B, int32_t T, int32_t Cin,
                                           const void* w_tiled, const float* bias, int32_t Cout, int32_t ksize,
                                           int32_t dilation, const float* snake_alpha, const float* residual, float* y,
                                           void* y_planes, const float* out_alpha, ptts_stream_t s) {
  PTTS_CHECK_ARG((x != nullptr) != (x_planes != nullptr) && w_tiled && (y || y_planes), "dac_conv1d_px: null pointer");
  PTTS_CHECK_ARG(!(x_planes && snake_alpha), "dac_conv1d_px: planes are already activated");
  PTTS_CHECK_ARG(B > 0 && T > 0 && Cin > 0 && Cin % CIK == 0 && Cout > 0 && Cout % 16 == 0,
                 "dac_conv1d_px: Cin must be a multiple of 32 and Cout of 16");
  PTTS_CHECK_ARG((ksize == 1 || ksize == 7) && dilation >= 1 && (ksize - 1) * dilation <= MAX_SPAN,
                 "dac_conv1d_px: kernel size 1 or 7, (k-1)*dilation <= 54");
  PTTS_CHECK_ARG(!x_planes || ((ksize - 1) / 2) * dilation <= PTTS_DAC_PLANE_PAD_LO, "dac_conv1d_px: planes halo");
  DacConv a = {};
  a.x = x; a.x_planes = (const __nv_bfloat16*)x_planes; a.w = (const __nv_bfloat16*)w_tiled; a.bias = bias;
  a.alpha = snake_alpha; a.residual = residual; a.y = y; a.y_planes = (__nv_bfloat16*)y_planes; a.out_alpha = out_alpha;
  a.T = T; a.Cin = Cin; a.N = Cout; a.Cout = Cout; a.taps = ksize; a.tap_center = (ksize - 1) / 2;
  a.tap_step = dilation; a.lo = -((ksize - 1) / 2) * dilation; a.span = (ksize - 1) * dilation;
  a.Mrows = T; a.out_stride = 1; a.out_pad = 0; a.Tout = T;
  const int rc = dac_conv_ws(a, B, as_stream(s));
  if (rc == PTTS_E_UNSUPPORTED) set_error("dac_conv1d_px: unsupported shape");
  return rc;
}

extern "C" PTTS_API int ptts_dac_convtr1d_px(const float* x, const void* x_planes, int32_t B, int32_t T, int32_t Cin,
                                             const void* w_tiled, const float* bias, int32_t Cout, int32_t stride,
                                             const float* snake_alpha, float* y, ptts_stream_t s) {
  PTTS_CHECK_ARG((x != nullptr) != (x_planes != nullptr) && w_tiled && y, "dac_convtr1d_px: null pointer");
  PTTS_CHECK_ARG(!(x_planes && snake_alpha), "dac_convtr1d_px: planes are already activated");
  PTTS_CHECK_ARG(B > 0 && T > 0 && Cin > 0 && Cin % CIK == 0 && Cout > 0 && Cout % 16 == 0 && stride >= 2,
                 "dac_convtr1d_px: Cin must be a multiple of 32, Cout of 16, stride >= 2");
  DacConv a = {};
  a.x = x; a.x_planes = (const __nv_bfloat16*)x_planes; a.w = (const __nv_bfloat16*)w_tiled; a.bias = bias;
  a.alpha = snake_alpha; a.y = y;
  a.T = T; a.Cin = Cin; a.N = stride * Cout; a.Cout = Cout; a.taps = 2; a.tap_center = 0; a.tap_step = -1;
  a.lo = -1; a.span = 1;
  a.Mrows = T + 1; a.out_stride = stride; a.out_pad = (stride + 1) / 2; a.Tout = T * stride;
  const int rc = dac_conv_ws(a, B, as_stream(s));
  if (rc == PTTS_E_UNSUPPORTED) set_error("dac_convtr1d_px: unsupported shape");
  return rc;
}

extern "C" PTTS_API int ptts_dac_conv_out_nwc(const float* x, int32_t B, int32_t T, int32_t Cin, const float* w,
                                     const float* bias, const float* snake_alpha, float* y, ptts_stream_t s) {
  PTTS_CHECK_ARG(x && w && y, "dac_conv_out_nwc: null pointer");
  PTTS_CHECK_ARG(B > 0 && T > 0 && Cin > 0 && Cin % 32 == 0 && Cin <= CO_MAXC,
                 "dac_conv_out_nwc: Cin must be a multiple of 32, <= 256");
  dac_conv_out_kernel<<<dim3((T + CO_ROWS - 1) / CO_ROWS, B), 256, 0, as_stream(s)>>>(x, T, Cin, w, bias, snake_alpha, y);
  PTTS_CHECK_LAUNCH("dac_conv_out_nwc");
  return PTTS_OK;
}
