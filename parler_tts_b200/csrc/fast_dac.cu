// DAC decoder, fast family: channels-last ("NWC", [B][T][C] fp32) activations and ONE tensor-core implicit-GEMM kernel
// for every convolution of the stack (arithmetic of descript-audio-codec as called from
// /root/reference/parler_tts/dac_wrapper/modeling_dac.py:130-131; restated in oracle/dac_oracle.py):
//   conv1d k=7 dilated / k=1 : y[t][co] = sum_tap sum_ci snake(x)[t + (tap - c) * dil][ci] * W[tap][co][ci]
//   conv_transpose1d k=2s    : y[q*s + ph - pad][co] = sum_ci snake(x)[q][ci] W[ci][co][ph] + snake(x)[q-1][ci] W[ci][co][ph+s]
//                              = the same kernel with 2 taps (row offsets 0, -1) and N = s * Cout columns (ph, co).
// GEMM view: M = time rows (128 per CTA), N = output columns (96 / 128 per CTA), K = taps x Cin.
// Precision: the waveform must match the fp32 oracle to 1e-3 max-abs through ~30 layers, which rules out single bf16 /
// tf32 products; every fp32 operand is split x = hi + mid (two bf16, 16 mantissa bits) and each MAC is
// hi*hi + hi*mid + mid*hi on mma.sync.m16n8k16 with fp32 accumulation (error ~2^-16 per product).
// Data movement: activations are read once per CTA as fp32 (coalesced float4 along channels), Snake + split happen in
// registers, the bf16 planes go to shared memory [time][32 ch] with an 80-byte pitch (conflict-free ldmatrix; a time
// shift of a tap is just a row offset); the pre-split weights stream through a 3-stage cp.async ring.
#include "gemm_common.cuh"

namespace ptts {

using namespace gemm;

namespace {

constexpr int TM = 128;    // time rows per CTA
constexpr int CIK = 32;    // input channels per chunk
constexpr int ROWB = 80;   // shared-memory row pitch in bytes (64 B of data + 16 B pad)
constexpr int NT2 = 256;   // threads
constexpr int WST = 3;     // weight ring stages
constexpr int MAX_SPAN = 54;  // max (hi - lo) tap span in rows: k = 7, dilation 9
constexpr int XU = ((TM + MAX_SPAN) * 8 + NT2 - 1) / NT2;  // float4 units per thread per x chunk

struct ConvParams {
  const float* x;              // [B][T][Cin]
  const __nv_bfloat16* w_hi;   // [taps][N][Cin]
  const __nv_bfloat16* w_mid;
  const float* bias;           // [Cout] or null
  const float* alpha;          // [Cin] Snake alpha, or null (no activation)
  const float* residual;       // [B][Tout][Cout] or null
  float* y;                    // [B][Tout][Cout]
  int T, Cin, N, Cout, taps, tap_center, tap_step, lo, span;
  int Mrows, out_stride, out_pad, Tout;
};

__device__ __forceinline__ float snake_f(float x, float a, float inv) {
  const float s = sinf(a * x);
  return x + inv * s * s;
}

__device__ __forceinline__ void split_bf16(float v, __nv_bfloat16& hi, __nv_bfloat16& mid) {
  hi = __float2bfloat16_rn(v);
  mid = __float2bfloat16_rn(v - __bfloat162float(hi));
}

template <int BN>
__global__ void __launch_bounds__(NT2, 1) dac_conv_mma_kernel(const ConvParams p) {
  extern __shared__ __align__(128) unsigned char smem[];
  constexpr int NP = BN / 32;            // pairs of n8 tiles per warp (warp covers BN/2 columns)
  constexpr int WPL = BN * ROWB;         // bytes of one weight plane per stage
  const int xrows = TM + p.span;
  const int XPL = xrows * ROWB;          // bytes of one activation plane
  // smem: xs[2 buffers][2 planes][XPL] | ws[WST][2 planes][WPL]
  const uint32_t sx = smem_u32(smem);
  const uint32_t sw = sx + 4 * XPL;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wm = warp & 3, wn = warp >> 2;
  const int m0 = blockIdx.x * TM, n0 = blockIdx.y * BN, b = blockIdx.z;
  const int nchunks = p.Cin / CIK;
  const int nit = nchunks * p.taps;
  const float* xb = p.x + (int64_t)b * p.T * p.Cin;

  float4 pre[XU];
  auto prefetch_x = [&](int c) {
#pragma unroll
    for (int j = 0; j < XU; ++j) {
      const int u = tid + NT2 * j;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (u < xrows * 8) {
        const int row = u >> 3, c4 = u & 7;
        const int t = m0 + p.lo + row;
        if (t >= 0 && t < p.T) v = __ldg(reinterpret_cast<const float4*>(xb + (int64_t)t * p.Cin + c * CIK + c4 * 4));
      }
      pre[j] = v;
    }
  };
  auto store_x = [&](int c, int buf) {
    unsigned char* base = smem + (size_t)buf * 2 * XPL;
#pragma unroll
    for (int j = 0; j < XU; ++j) {
      const int u = tid + NT2 * j;
      if (u < xrows * 8) {
        const int row = u >> 3, c4 = u & 7;
        float4 v = pre[j];
        if (p.alpha) {
          const float4 a = __ldg(reinterpret_cast<const float4*>(p.alpha + c * CIK + c4 * 4));
          v.x = snake_f(v.x, a.x, 1.0f / (a.x + 1e-9f));
          v.y = snake_f(v.y, a.y, 1.0f / (a.y + 1e-9f));
          v.z = snake_f(v.z, a.z, 1.0f / (a.z + 1e-9f));
          v.w = snake_f(v.w, a.w, 1.0f / (a.w + 1e-9f));
        }
        __nv_bfloat16 h[4], m[4];
        split_bf16(v.x, h[0], m[0]);
        split_bf16(v.y, h[1], m[1]);
        split_bf16(v.z, h[2], m[2]);
        split_bf16(v.w, h[3], m[3]);
        *reinterpret_cast<uint2*>(base + row * ROWB + c4 * 8) = *reinterpret_cast<uint2*>(h);
        *reinterpret_cast<uint2*>(base + XPL + row * ROWB + c4 * 8) = *reinterpret_cast<uint2*>(m);
      }
    }
  };
  auto load_w = [&](int it) {
    if (it < nit) {
      const int c = it / p.taps, tap = it % p.taps;
      const uint32_t dst0 = sw + (it % WST) * 2 * WPL;
#pragma unroll
      for (int j = 0; j < (BN * 8 + NT2 - 1) / NT2; ++j) {
        const int q = tid + NT2 * j;
        if (q < BN * 8) {
          const int plane = q / (BN * 4), rem = q % (BN * 4);
          const int n = rem >> 2, ch = rem & 3;
          const bool ok = n0 + n < p.N;
          const __nv_bfloat16* wsrc = plane ? p.w_mid : p.w_hi;
          const __nv_bfloat16* src = ok ? wsrc + ((int64_t)tap * p.N + n0 + n) * p.Cin + c * CIK + ch * 8 : wsrc;
          cp_async16(dst0 + plane * WPL + n * ROWB + ch * 16, src, ok);
        }
      }
    }
    cp_async_commit();
  };

  float acc[2][BN / 16][4];
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < BN / 16; ++j)
#pragma unroll
      for (int k = 0; k < 4; ++k) acc[i][j][k] = 0.f;

  load_w(0);
  load_w(1);
  prefetch_x(0);
  store_x(0, 0);
  if (nchunks > 1) prefetch_x(1);

  for (int it = 0; it < nit; ++it) {
    const int c = it / p.taps, tap = it % p.taps;
    cp_async_wait<WST - 2>();
    __syncthreads();
    load_w(it + WST - 1);
    const uint32_t xh = sx + (c & 1) * 2 * XPL, xm = xh + XPL;
    const uint32_t wh = sw + (it % WST) * 2 * WPL, wmid = wh + WPL;
    const int shift = (tap - p.tap_center) * p.tap_step - p.lo;
    const int arow = wm * 32 + (lane & 7) + ((lane >> 3) & 1) * 8 + shift;
#pragma unroll
    for (int ks = 0; ks < CIK / 16; ++ks) {
      uint32_t ah[2][4], am[2][4];
#pragma unroll
      for (int mt = 0; mt < 2; ++mt) {
        const uint32_t off = (arow + mt * 16) * ROWB + (ks * 2 + (lane >> 4)) * 16;
        ldmatrix_x4(ah[mt], xh + off);
        ldmatrix_x4(am[mt], xm + off);
      }
#pragma unroll
      for (int np = 0; np < NP; ++np) {
        uint32_t bh[4], bm[4];
        const uint32_t off =
            (wn * (BN / 2) + np * 16 + (lane & 7) + (lane >> 4) * 8) * ROWB + (ks * 2 + ((lane >> 3) & 1)) * 16;
        ldmatrix_x4(bh, wh + off);
        ldmatrix_x4(bm, wmid + off);
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            mma_bf16(acc[mt][np * 2 + h], am[mt], bh[2 * h], bh[2 * h + 1]);
            mma_bf16(acc[mt][np * 2 + h], ah[mt], bm[2 * h], bm[2 * h + 1]);
            mma_bf16(acc[mt][np * 2 + h], ah[mt], bh[2 * h], bh[2 * h + 1]);
          }
      }
    }
    if (tap == p.taps - 1 && c + 1 < nchunks) {  // hand the prefetched chunk over, start fetching the one after
      store_x(c + 1, (c + 1) & 1);
      if (c + 2 < nchunks) prefetch_x(c + 2);
    }
  }
  cp_async_wait<0>();

  // ---- epilogue: bias, residual, (ph, co) -> output row mapping of the transposed conv
  const float* resb = p.residual ? p.residual + (int64_t)b * p.Tout * p.Cout : nullptr;
  float* yb = p.y + (int64_t)b * p.Tout * p.Cout;
#pragma unroll
  for (int mt = 0; mt < 2; ++mt)
#pragma unroll
    for (int nt = 0; nt < BN / 16; ++nt) {
      const int n = n0 + wn * (BN / 2) + nt * 8 + (lane & 3) * 2;
      if (n >= p.N) continue;
      const int ph = n / p.Cout, co = n - ph * p.Cout;
      const float b0 = p.bias ? p.bias[co] : 0.f, b1 = p.bias ? p.bias[co + 1] : 0.f;
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        const int m = m0 + wm * 32 + mt * 16 + (lane >> 2) + hh * 8;
        if (m >= p.Mrows) continue;
        const int t = m * p.out_stride + ph - p.out_pad;
        if (t < 0 || t >= p.Tout) continue;
        const int64_t o = (int64_t)t * p.Cout + co;
        float v0 = acc[mt][nt][hh * 2] + b0, v1 = acc[mt][nt][hh * 2 + 1] + b1;
        if (resb) {
          const float2 r = *reinterpret_cast<const float2*>(resb + o);
          v0 += r.x;
          v1 += r.y;
        }
        *reinterpret_cast<float2*>(yb + o) = make_float2(v0, v1);
      }
    }
}

// Output conv (Cout = 1, k = 7) + tanh over channels-last fp32 input: y[t] = tanh(b + sum_k sum_c snake(x)[t+k-3][c] w[k][c]).
// HBM-bound (one read of x).  Phase 1: z[r][k] = sum_c snake(x[r][c]) w[k][c] for the 134 input rows of a 128-row tile:
// 8 threads per row (fully coalesced 128-byte reads), every element is activated once and used for all 7 taps; the 8
// partial vectors are combined by a 7-shuffle butterfly that leaves tap k in lane k.  Phase 2: y[t] = sum_k z[t+k][k].
constexpr int CO_ROWS = 128, CO_HALO = 6, CO_MAXC = 256;
__device__ __forceinline__ float snake_fast(float x, float a, float inv) {  // range-reduced SFU sine (see fast_dac2.cu)
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

template <int BN>
int launch_conv(const ConvParams& p, int B, cudaStream_t st) {
  const size_t smem = (size_t)4 * (TM + p.span) * ROWB + (size_t)WST * 2 * BN * ROWB;
  static size_t configured = 0;
  if (smem > configured) {
    cudaFuncSetAttribute(dac_conv_mma_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    configured = smem;
  }
  dim3 grid((p.Mrows + TM - 1) / TM, (p.N + BN - 1) / BN, B);
  dac_conv_mma_kernel<BN><<<grid, NT2, smem, st>>>(p);
  PTTS_CHECK_LAUNCH("dac_conv_mma");
  return PTTS_OK;
}

int launch_conv_any(const ConvParams& p, int B, cudaStream_t st) {
  if (!p.w_mid) {  // pre-tiled weights: the warp-specialised persistent tcgen05 kernel (fast_dac3.cu)
    DacConv a = {};
    a.x = p.x; a.w = p.w_hi; a.bias = p.bias; a.alpha = p.alpha; a.residual = p.residual; a.y = p.y;
    a.T = p.T; a.Cin = p.Cin; a.N = p.N; a.Cout = p.Cout; a.taps = p.taps; a.tap_center = p.tap_center;
    a.tap_step = p.tap_step; a.lo = p.lo; a.span = p.span; a.Mrows = p.Mrows; a.out_stride = p.out_stride;
    a.out_pad = p.out_pad; a.Tout = p.Tout;
    const int rc = dac_conv_ws(a, B, st);
    if (rc == PTTS_E_UNSUPPORTED) set_error("dac conv: pre-tiled weights need Cin %% 32 == 0 and Cout %% 16 == 0");
    return rc;
  }
  if (dac_umma_enabled()) {  // tcgen05 kernel (fast_dac2.cu); shapes it does not take fall through to mma.sync
    const int rc = dac_conv_umma(&p, B, st);
    if (rc != PTTS_E_UNSUPPORTED) return rc;
  }
  if (p.N % 128 != 0 && p.N % 96 == 0) return launch_conv<96>(p, B, st);
  return launch_conv<128>(p, B, st);
}

}  // namespace
}  // namespace ptts

using namespace ptts;

extern "C" PTTS_API int ptts_dac_tile_n(int32_t N) { return dac_tile_n(N); }

// 1: the host should pass pre-tiled weights (default); 0: PTTS_DAC=umma|mma selected a two-plane kernel
extern "C" PTTS_API int ptts_dac_pretiled(void) {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("PTTS_DAC");
    v = (e && (e[0] == 'm' || e[0] == 'u')) ? 0 : 1;
  }
  return v;
}

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

extern "C" PTTS_API int ptts_dac_conv1d_nwc(const float* x, int32_t B, int32_t T, int32_t Cin, const void* w_hi,
                                   const void* w_mid, const float* bias, int32_t Cout, int32_t ksize, int32_t dilation,
                                   const float* snake_alpha, const float* residual, float* y, ptts_stream_t s) {
  PTTS_CHECK_ARG(x && w_hi && y, "dac_conv1d_nwc: null pointer");
  PTTS_CHECK_ARG(B > 0 && T > 0 && Cin > 0 && Cin % CIK == 0 && Cout > 0 && Cout % 2 == 0,
                 "dac_conv1d_nwc: Cin must be a multiple of 32 and Cout even");
  PTTS_CHECK_ARG((ksize == 1 || ksize == 7) && dilation >= 1 && (ksize - 1) * dilation <= MAX_SPAN,
                 "dac_conv1d_nwc: kernel size 1 or 7, (k-1)*dilation <= 54");
  ConvParams p = {};
  p.x = x; p.w_hi = (const __nv_bfloat16*)w_hi; p.w_mid = (const __nv_bfloat16*)w_mid; p.bias = bias;
  p.alpha = snake_alpha; p.residual = residual; p.y = y;
  p.T = T; p.Cin = Cin; p.N = Cout; p.Cout = Cout; p.taps = ksize; p.tap_center = (ksize - 1) / 2;
  p.tap_step = dilation; p.lo = -((ksize - 1) / 2) * dilation; p.span = (ksize - 1) * dilation;
  p.Mrows = T; p.out_stride = 1; p.out_pad = 0; p.Tout = T;
  return launch_conv_any(p, B, as_stream(s));
}

extern "C" PTTS_API int ptts_dac_convtr1d_nwc(const float* x, int32_t B, int32_t T, int32_t Cin, const void* w_hi,
                                     const void* w_mid, const float* bias, int32_t Cout, int32_t stride,
                                     const float* snake_alpha, float* y, ptts_stream_t s) {
  PTTS_CHECK_ARG(x && w_hi && y, "dac_convtr1d_nwc: null pointer");
  PTTS_CHECK_ARG(B > 0 && T > 0 && Cin > 0 && Cin % CIK == 0 && Cout > 0 && Cout % 2 == 0 && stride >= 2,
                 "dac_convtr1d_nwc: Cin must be a multiple of 32, Cout even, stride >= 2");
  ConvParams p = {};
  p.x = x; p.w_hi = (const __nv_bfloat16*)w_hi; p.w_mid = (const __nv_bfloat16*)w_mid; p.bias = bias;
  p.alpha = snake_alpha; p.residual = nullptr; p.y = y;
  p.T = T; p.Cin = Cin; p.N = stride * Cout; p.Cout = Cout; p.taps = 2; p.tap_center = 0; p.tap_step = -1;
  p.lo = -1; p.span = 1;
  p.Mrows = T + 1; p.out_stride = stride; p.out_pad = (stride + 1) / 2; p.Tout = T * stride;
  return launch_conv_any(p, B, as_stream(s));
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
