// Weight-streaming "skinny" GEMM of the decode step: Y[M][N] = epi(X[M][K] . W[N][K]^T), bf16 operands, fp32
// accumulate.  At decode time M = batch (<= a few hundred rows) so every linear of the decoder layer
// (/root/reference/parler_tts/modeling_parler_tts.py:282,301-302,380,488-490,1143) is bound by streaming W from HBM
// once; the kernel is organised around that stream:
//   * grid = (N / BN) x (M / 64) x split_k CTAs, chosen so that >= ~1 CTA per SM pulls weights concurrently;
//   * W and X tiles travel global -> shared with 16-byte cp.async in a 3-4 stage ring (XOR-swizzled 256-byte rows, so
//     ldmatrix is bank-conflict free); the W tiles of the first stages are requested BEFORE griddepcontrol.wait, i.e.
//     while the previous kernel of the step is still draining (programmatic dependent launch);
//   * math: mma.sync.m16n8k16 bf16 -> fp32 (at M = 64 the tensor pipe is < 10 % busy: HBM is the bound, the roofline
//     in DESIGN.md section 4 is bytes(W) / time);
//   * epilogues are fused: STORE (alpha, bias, activation, row mask, fp32 residual, bf16|fp32 out), SCATTER (q scaled to
//     the q buffer, k / v written in place into the static KV cache - replaces the torch.cat cache growth of
//     modeling_parler_tts.py:303-304), PARTIAL (split-K partial sums, reduced deterministically by the LayerNorm
//     kernel that consumes the residual stream next).
#include <stdlib.h>

#include "gemm_common.cuh"

namespace ptts {

using namespace gemm;

namespace {

template <int BN, int EPI, int STAGES>
__global__ void __launch_bounds__(NTH) skinny_gemm_kernel(const GemmParams p) {
  extern __shared__ __align__(128) unsigned char smem[];
  constexpr int XB = BM * 256, WB = BN * 256, SB = XB + WB;
  const int n0 = blockIdx.x * BN, m0 = blockIdx.y * BM, z = blockIdx.z;
  const int kbeg = z * p.k_per_split, KT = p.k_per_split / BK;
  const uint32_t s0 = smem_u32(smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // weights do not depend on the previous kernel: request them before the grid dependency resolves
#pragma unroll
  for (int s = 0; s < STAGES - 1; ++s)
    if (s < KT) load_tile<BN>(s0 + s * SB + XB, p.W, p.K, n0, p.N, kbeg + s * BK);
  pdl_wait();
  pdl_launch();
#pragma unroll
  for (int s = 0; s < STAGES - 1; ++s) {
    if (s < KT) load_tile<BM>(s0 + s * SB, p.X, p.K, m0, p.M, kbeg + s * BK);
    cp_async_commit();
  }

  float acc[BN / 8][4];
#pragma unroll
  for (int i = 0; i < BN / 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int kt = 0; kt < KT; ++kt) {
    cp_async_wait<STAGES - 2>();
    __syncthreads();
    {
      const int nk = kt + STAGES - 1;
      if (nk < KT) {
        const uint32_t sb = s0 + (nk % STAGES) * SB;
        load_tile<BN>(sb + XB, p.W, p.K, n0, p.N, kbeg + nk * BK);
        load_tile<BM>(sb, p.X, p.K, m0, p.M, kbeg + nk * BK);
      }
      cp_async_commit();
    }
    const uint32_t xs = s0 + (kt % STAGES) * SB, ws = xs + XB;
#pragma unroll
    for (int ks = 0; ks < BK / 16; ++ks) {
      uint32_t a[4];
      {
        const int r = warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
        const int c = ks * 2 + (lane >> 4);
        ldmatrix_x4(a, xs + r * 256 + ((c ^ (r & 7)) << 4));
      }
      if constexpr (BN == 8) {
        uint32_t b[2];
        const int r = lane & 7;
        const int c = ks * 2 + ((lane >> 3) & 1);
        ldmatrix_x2(b, ws + r * 256 + ((c ^ (r & 7)) << 4));
        mma_bf16(acc[0], a, b[0], b[1]);
      } else {
#pragma unroll
        for (int np = 0; np < BN / 16; ++np) {
          uint32_t b[4];
          const int r = np * 16 + (lane & 7) + (lane >> 4) * 8;
          const int c = ks * 2 + ((lane >> 3) & 1);
          ldmatrix_x4(b, ws + r * 256 + ((c ^ (r & 7)) << 4));
          mma_bf16(acc[2 * np], a, b[0], b[1]);
          mma_bf16(acc[2 * np + 1], a, b[2], b[3]);
        }
      }
    }
  }
  cp_async_wait<0>();

  // ---- epilogue: thread holds (row, col..col+1) and (row+8, col..col+1) of every 8-wide n tile
  const int rbase = m0 + warp * 16 + (lane >> 2);
#pragma unroll
  for (int nt = 0; nt < BN / 8; ++nt) {
    const int col = n0 + nt * 8 + (lane & 3) * 2;
    if (col >= p.N) continue;
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
      const int row = rbase + hh * 8;
      if (row >= p.M) continue;
      float v0 = acc[nt][hh * 2], v1 = acc[nt][hh * 2 + 1];
      gemm::epilogue_pair<EPI>(p, row, col, z, v0, v1);
    }
  }
}

template <int BN> constexpr int stages_for() { return BN >= 64 ? 3 : 4; }

template <int BN, int EPI>
int launch_one(const GemmParams& p, int splits, cudaStream_t st) {
  constexpr int ST = stages_for<BN>();
  constexpr size_t smem = (size_t)ST * (BM * 256 + BN * 256);
  static bool configured = false;
  if (!configured) {
    cudaFuncSetAttribute(skinny_gemm_kernel<BN, EPI, ST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    configured = true;
  }
  dim3 grid((p.N + BN - 1) / BN, (p.M + BM - 1) / BM, splits);
  cudaError_t e = launch_k(skinny_gemm_kernel<BN, EPI, ST>, grid, dim3(NTH), smem, st, p);
  if (e != cudaSuccess) {
    set_error("skinny_gemm launch: %s", cudaGetErrorString(e));
    return PTTS_E_LAUNCH;
  }
  PTTS_CHECK_LAUNCH("skinny_gemm");
  return PTTS_OK;
}

template <int EPI>
int launch_bn(int bn, const GemmParams& p, int splits, cudaStream_t st) {
  switch (bn) {
    case 8: return launch_one<8, EPI>(p, splits, st);
    case 16: return launch_one<16, EPI>(p, splits, st);
    case 32: return launch_one<32, EPI>(p, splits, st);
    default: return launch_one<64, EPI>(p, splits, st);
  }
}

constexpr int kTargetCtas = 120;

int pick_bn(int M, int N) {
  const int mt = (M + BM - 1) / BM;
  for (int bn = 64; bn >= 8; bn >>= 1)
    if (((N + bn - 1) / bn) * mt >= kTargetCtas) return bn;
  return 8;
}

// 2: the tcgen05 swap-AB kernel (fast_gemm3.cu) wherever its shape constraints hold, this file's mma.sync pipelined
// kernel for the rest (K % 128 != 0, odd tile counts); diagnostic build, PTTS_GEMM_RING=1: always the latter (A/B runs)
int gemm_mode() { return PTTS_KNOB("PTTS_GEMM_RING", 0) ? 0 : 2; }

}  // namespace

bool gemm_bf16_supported(int M, int N, int K) { return M > 0 && N > 0 && (N % 2) == 0 && K > 0 && (K % BK) == 0; }

int gemm_bf16_store(const void* x, const void* w, int M, int N, int K, float alpha, int act, const float* bias,
                    const float* residual, const int* row_mask, void* y, int ydtype, cudaStream_t st) {
  GemmParams p = {};
  p.X = (const __nv_bfloat16*)x;
  p.W = (const __nv_bfloat16*)w;
  p.M = M; p.N = N; p.K = K; p.k_per_split = K;
  p.alpha = alpha; p.act = act; p.bias = bias; p.residual = residual; p.row_mask = row_mask; p.y = y;
  p.ydtype = ydtype;
#ifndef PTTS_GEMM_BIG
#define PTTS_GEMM_BIG 1
#endif
  if (PTTS_GEMM_BIG && gemm_mode() == 2 && M >= 256) {  // prefill / T5 / cross-KV sizes: the plain tile kernel
    const int rc = big_gemm_store(p, st);
    if (rc != PTTS_E_UNSUPPORTED) return rc;
  }
  if (gemm_mode() == 2) {
    const int rc = umma_gemm_store(p, st);
    if (rc != PTTS_E_UNSUPPORTED) return rc;
  }
  return launch_bn<EPI_STORE>(pick_bn(M, N), p, 1, st);
}

int gemm_bf16_scatter(const void* x, const void* w, int M, int N, int K, int has_q, float q_scale, void* q, void* ck,
                      void* cv, int H, int heads, int d, int Tq, int c_stride, int pos_base, const int* pos_dyn,
                      cudaStream_t st) {
  GemmParams p = {};
  p.X = (const __nv_bfloat16*)x;
  p.W = (const __nv_bfloat16*)w;
  p.M = M; p.N = N; p.K = K; p.k_per_split = K;
  p.q = (__nv_bfloat16*)q; p.ck = (__nv_bfloat16*)ck; p.cv = (__nv_bfloat16*)cv;
  p.H = H; p.heads = heads; p.d = d; p.Tq = Tq; p.c_stride = c_stride; p.pos_base = pos_base; p.pos_dyn = pos_dyn;
  p.has_q = has_q; p.q_scale = q_scale;
  if (PTTS_GEMM_BIG && gemm_mode() == 2 && M >= 256) {
    const int rc = big_gemm_scatter(p, st);
    if (rc != PTTS_E_UNSUPPORTED) return rc;
  }
  if (gemm_mode() == 2) {
    const int rc = umma_gemm_scatter(p, st);
    if (rc != PTTS_E_UNSUPPORTED) return rc;
  }
  return launch_bn<EPI_SCATTER>(pick_bn(M, N), p, 1, st);
}

// Split-K plan for a residual GEMM: the partial sums [S][M][N] are added to the residual stream by the next LayerNorm.
// Returns S (1 = the caller should use the fused STORE epilogue instead).
int gemm_bf16_plan_splits(int M, int N, int K, int* bn_out) {
  if (gemm_mode() == 2) {  // tcgen05 kernel: bn_out = 0 marks the plan as "umma"
    const int s = umma_plan_splits(M, N, K);
    if (s > 1) {
      if (bn_out) *bn_out = 0;
      return s;
    }
  }
  const int mt = (M + BM - 1) / BM;
  int best_bn = 64, best_s = 1;
  for (int bn = 64; bn >= 16; bn >>= 1) {
    const int nt = ((N + bn - 1) / bn) * mt;
    int s = (kTargetCtas + nt - 1) / nt;
    const int smax = K / 256 > 0 ? K / 256 : 1;
    if (s > smax) s = smax;
    if (s > kMaxSplitK) s = kMaxSplitK;
    while (s > 1 && ((K / BK) % s) != 0) --s;  // equal K slices, multiples of BK
    best_bn = bn;
    best_s = s;
    if (nt * s >= kTargetCtas) break;
  }
  if (bn_out) *bn_out = best_bn;
  return best_s;
}

int gemm_bf16_partial(const void* x, const void* w, int M, int N, int K, int splits, int bn, float* part,
                      cudaStream_t st) {
  GemmParams p = {};
  p.X = (const __nv_bfloat16*)x;
  p.W = (const __nv_bfloat16*)w;
  p.M = M; p.N = N; p.K = K; p.k_per_split = K / splits;
  p.part = part;
  if (bn == 0) return umma_gemm_partial(p, splits, st);
  return launch_bn<EPI_PARTIAL>(bn, p, splits, st);
}

}  // namespace ptts
