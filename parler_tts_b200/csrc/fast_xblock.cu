// The whole cross-attention block of one decoder layer at decode time (one query token per utterance) in ONE launch:
//   h'   = h + sum_s part_in[s]                       (folds the split-K partial sums of the self-attention out_proj)
//   xn   = LayerNorm(h') * gamma + beta               (encoder_attn_layer_norm, modeling_parler_tts.py:467)
//   q    = (xn . Wq^T) * d^-0.5                       (encoder_attn.q_proj, modeling_parler_tts.py:282)
//   a    = softmax(q . K^T + mask) . V                (cached T5 K/V, modeling_parler_tts.py:292-294, 326-365)
//   part_out[head] = a_head . Wo[:, head]^T           (encoder_attn.out_proj as split-K-by-head partial sums, :380)
// The residual add of modeling_parler_tts.py:483 is done by the NEXT LayerNorm launch, which folds part_in and
// part_out (stored back to back) into the residual stream in slice order; this kernel never writes h.
//
// Replaces four launches of the chain (LayerNorm, q GEMM, cross flash-decode, out_proj GEMM: ~19.5 us per layer at
// B=64 although they move only ~17 MB) - the block is pure dependency latency, so the fix is to keep it on chip:
// one CTA per (head, group of 8 utterances).  Per head the q projection needs only the 64 rows of Wq of that head
// (128 KB at H=1024, requested into shared memory BEFORE griddepcontrol.wait so that it streams while the producer
// still runs) and the out projection only the 64 columns of Wo of that head (L2-prefetched before the wait, staged
// into the same shared memory while the attention phase runs).  The 8 rows of the group ride in the lower half of
// an m16n8k16 mma.sync tile (the upper half is fed zeros): at 8 x 64 x 1024 MACs per GEMM the tensor pipe is idle
// either way, and tcgen05 would add a TMEM round trip to a 1 us phase.
#include <stdlib.h>

#include "gemm_common.cuh"

namespace ptts {

namespace {

using gemm::cp_async16;
using gemm::cp_async_commit;
using gemm::cp_async_wait;
using gemm::ldmatrix_x4;
using gemm::mma_bf16;
using gemm::smem_u32;

constexpr int XG = 8;     // utterances per CTA = warps per CTA (one warp owns one utterance in the per-row phases)
constexpr int XT = XG * 32;
constexpr int XD = 64;    // head dim

struct XBlockParams {
  const float* h;              // [B][H] residual stream (read only)
  const float* part_in;        // [S_in][B][H] pending partial sums of the previous residual GEMM (or S_in = 0)
  int S_in;
  int64_t stride;              // B * H
  const float* gamma;
  const float* beta;
  float eps;
  const __nv_bfloat16* wq;     // [H][H]
  const __nv_bfloat16* wo;     // [H][H]
  const __nv_bfloat16* kc;     // [B][heads][S][64]
  const __nv_bfloat16* vc;
  const int* key_mask;         // [B][S] or null
  int B, heads, S;
  float scale;
  float* part_out;             // [heads][B][H]
  // L2 prefetch of the next self-attention launch's K rows (see fast_attn.cu)
  const __nv_bfloat16* pf_next;
  int pf_stride, pf_rows_base, pf_pct;
  const int* pf_rows_dyn;
  L2Hint hint;                 // matrix of a later GEMM (fc1) to pull into L2 while this block runs
};

// 16-byte chunk c of row r lives at chunk position c ^ (r & 7): conflict-free ldmatrix over 8 consecutive rows
__device__ __forceinline__ uint32_t swz(int r, int c, int row_bytes) {
  return (uint32_t)(r * row_bytes + ((c ^ (r & 7)) << 4));
}

// PTTS_XB_DBG=1: sums over all CTAs (and launches) of SM-clock cycles since the CTA's griddepcontrol.wait returned:
// [0] LayerNorm done, [1] Wq slab landed (barrier), [2] q done (barrier), [3] attention done, [4] Wo slab landed
// (barrier), [5] end, [6] CTA count, [7] max end; read by ptts_diag_xblock
__device__ unsigned long long g_xb_agg[8];

template <int HT>  // H = 128 * HT
__global__ void __launch_bounds__(XT, 1) xblock_kernel(const XBlockParams p, int sc_cap, int dbg, TL tl) {
  constexpr int H = 128 * HT;
  constexpr int ROWB = 2 * H;  // bytes per bf16 row of H
  extern __shared__ __align__(128) unsigned char smraw[];
  unsigned char* wbuf = smraw;                         // Wq head slab [64][H] bf16, later Wo head slab [H][64] bf16
  unsigned char* xn = wbuf + XD * ROWB;                // [8][H] bf16 (swizzled)
  float* qs = reinterpret_cast<float*>(xn + XG * ROWB);  // [8][64] scaled q (rounded to bf16 like the q buffer)
  unsigned char* at = reinterpret_cast<unsigned char*>(qs + XG * XD);  // [8][64] bf16 attention output (swizzled)
  float* sc = reinterpret_cast<float*>(at + XG * XD * 2);              // [8][sc_cap] scores -> probabilities
  const int hd = blockIdx.x, b0 = blockIdx.y * XG;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nrows = min(XG, p.B - b0);
  const uint32_t wbuf_s = smem_u32(wbuf), xn_s = smem_u32(xn), at_s = smem_u32(at);
  tl_start(tl);

  // ---- before the dependency: everything below only touches immutable data (weights, the cross cache, cache rows of
  // the next self-attention launch that are older than this step)
  {
    const __nv_bfloat16* src = p.wq + (int64_t)hd * XD * H;
    constexpr int CPR = ROWB / 16;  // chunks per row
    for (int i = tid; i < XD * CPR; i += XT) {
      const int r = i / CPR, c = i % CPR;
      cp_async16(wbuf_s + swz(r, c, ROWB), src + (int64_t)r * H + c * 8, true);
    }
    cp_async_commit();
  }
  pdl_launch();
  l2_hint_prefetch(p.hint, blockIdx.y * gridDim.x + blockIdx.x, gridDim.x * gridDim.y);
  for (int n = tid; n < H; n += XT)  // Wo[:, head]: one 128-byte line per output feature
    asm volatile("prefetch.global.L2 [%0];\n" ::"l"(p.wo + (int64_t)n * H + hd * XD) : "memory");
  {
    const int64_t kv_bytes = (int64_t)p.S * XD * 2;
    for (int i = tid; i < nrows * 2; i += XT) {  // K and V of the group's utterances for this head
      const int b = b0 + (i >> 1);
      const __nv_bfloat16* base = ((i & 1) ? p.vc : p.kc) + ((int64_t)b * p.heads + hd) * p.S * XD;
      for (int64_t o = 0; o < kv_bytes; o += 8192)
        l2_prefetch_bulk(reinterpret_cast<const char*>(base) + o, (uint32_t)min((int64_t)8192, kv_bytes - o));
    }
    if (p.pf_next && p.pf_pct > 0) {
      const int rows = dyn_value(p.pf_rows_base, p.pf_rows_dyn) * min(p.pf_pct, 100) / 100;
      const int64_t bytes = (int64_t)rows * XD * 2;
      const int pair = tid >> 5;  // warp w pulls the rows of utterance b0 + w
      if (pair < nrows) {
        const char* src = reinterpret_cast<const char*>(p.pf_next + ((int64_t)(b0 + pair) * p.heads + hd) * p.pf_stride * XD);
        for (int64_t o = (int64_t)lane * 8192; o < bytes; o += 32 * 8192)
          l2_prefetch_bulk(src + o, (uint32_t)min((int64_t)8192, bytes - o));
      }
    }
  }
  pdl_wait();
  tl_dep(tl);
  const long long c_dep = dbg ? clock64() : 0;
  long long c_st[5] = {0, 0, 0, 0, 0};

  // ---- phase 1: fold + LayerNorm of row b0 + warp -> xn (bf16, shared memory)
  {
    const int b = b0 + warp;
    float4 v[HT];
    if (warp < nrows) {
      const float* hr = p.h + (int64_t)b * H;
      constexpr int SB = HT <= 8 ? 4 : 2;  // partial slices in flight next to the residual row (register budget)
#pragma unroll
      for (int i = 0; i < HT; ++i) v[i] = __ldcg(reinterpret_cast<const float4*>(hr) + i * 32 + lane);
      for (int s0 = 0; s0 < p.S_in; s0 += SB) {  // SB slices (SB x HT loads) in flight, added in slice order
        float4 t[SB][HT];
#pragma unroll
        for (int u = 0; u < SB; ++u)
          if (s0 + u < p.S_in) {
            const float4* ps = reinterpret_cast<const float4*>(p.part_in + (int64_t)(s0 + u) * p.stride + (int64_t)b * H);
#pragma unroll
            for (int i = 0; i < HT; ++i) t[u][i] = __ldcg(ps + i * 32 + lane);
          }
#pragma unroll
        for (int u = 0; u < SB; ++u)
          if (s0 + u < p.S_in) {
#pragma unroll
            for (int i = 0; i < HT; ++i) {
              v[i].x += t[u][i].x; v[i].y += t[u][i].y; v[i].z += t[u][i].z; v[i].w += t[u][i].w;
            }
          }
      }
      float s = 0.f;
#pragma unroll
      for (int i = 0; i < HT; ++i) s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
      const float mean = warp_sum(s) / (float)H;
      float qq = 0.f;
#pragma unroll
      for (int i = 0; i < HT; ++i) {
        const float a = v[i].x - mean, bb = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
        qq += (a * a + bb * bb) + (c * c + d * d);
      }
      const float rstd = rsqrtf(warp_sum(qq) / (float)H + p.eps);
#pragma unroll
      for (int i = 0; i < HT; ++i) {
        const int col = (i * 32 + lane) * 4;
        const float4 g = __ldg(reinterpret_cast<const float4*>(p.gamma + col));
        const float4 be = __ldg(reinterpret_cast<const float4*>(p.beta + col));
        const __nv_bfloat162 lo = __floats2bfloat162_rn((v[i].x - mean) * rstd * g.x + be.x, (v[i].y - mean) * rstd * g.y + be.y);
        const __nv_bfloat162 hi = __floats2bfloat162_rn((v[i].z - mean) * rstd * g.z + be.z, (v[i].w - mean) * rstd * g.w + be.w);
        uint2 u;
        u.x = *reinterpret_cast<const uint32_t*>(&lo);
        u.y = *reinterpret_cast<const uint32_t*>(&hi);
        *reinterpret_cast<uint2*>(xn + swz(warp, col >> 3, ROWB) + (col & 7) * 2) = u;
      }
    } else {  // rows past the batch: zeros (their results are never stored)
#pragma unroll
      for (int i = 0; i < HT; ++i) {
        const int col = (i * 32 + lane) * 4;
        *reinterpret_cast<uint2*>(xn + swz(warp, col >> 3, ROWB) + (col & 7) * 2) = make_uint2(0u, 0u);
      }
    }
  }
  if (dbg) c_st[0] = clock64();
  cp_async_wait<0>();
  __syncthreads();  // xn complete, the Wq slab has landed
  if (dbg) c_st[1] = clock64();

  // ---- phase 2: q[8][64] = xn . Wq_head^T; warp w owns output features [8w, 8w + 8)
  {
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    const int mi = lane >> 3, ri = lane & 7;  // ldmatrix: lane supplies row ri of 8x8 matrix mi (= k chunk kc + mi)
    const uint32_t a_base = xn_s + ri * ROWB, b_base = wbuf_s + (warp * 8 + ri) * ROWB;
#pragma unroll 4
    for (int kc = 0; kc < ROWB / 16; kc += 4) {  // 32 k per iteration = two k16 steps
      uint32_t a[4], bq[4];
      const uint32_t off = (uint32_t)(((kc + mi) ^ ri) << 4);
      ldmatrix_x4(a, a_base + off);
      ldmatrix_x4(bq, b_base + off);
      const uint32_t a0[4] = {a[0], 0u, a[1], 0u}, a1[4] = {a[2], 0u, a[3], 0u};
      mma_bf16(acc[(kc >> 2) & 1 ? 2 : 0], a0, bq[0], bq[1]);
      mma_bf16(acc[(kc >> 2) & 1 ? 3 : 1], a1, bq[2], bq[3]);
    }
    const int g = lane >> 2, t = lane & 3;
    const float q0 = (acc[0][0] + acc[1][0]) + (acc[2][0] + acc[3][0]);
    const float q1 = (acc[0][1] + acc[1][1]) + (acc[2][1] + acc[3][1]);
    float* qd = qs + g * XD + warp * 8 + 2 * t;
    qd[0] = __bfloat162float(__float2bfloat16_rn(q0 * p.scale));
    qd[1] = __bfloat162float(__float2bfloat16_rn(q1 * p.scale));
  }
  __syncthreads();  // q visible; every warp is done with the Wq slab
  if (dbg) c_st[2] = clock64();

  // ---- the Wo head slab (L2-resident by now) replaces the Wq slab while the attention phase runs
  {
    const __nv_bfloat16* src = p.wo + hd * XD;
    for (int i = tid; i < H * 8; i += XT) {
      const int r = i >> 3, c = i & 7;
      cp_async16(wbuf_s + swz(r, c, 128), src + (int64_t)r * H + c * 8, true);
    }
    cp_async_commit();
  }

  // ---- phase 3: cross-attention of utterance b0 + warp over its S cached keys (4 lanes per key row, 8 rows per pass)
  if (warp < nrows) {
    const int b = b0 + warp;
    const int kg = lane >> 2, c4 = lane & 3;
    const __nv_bfloat16* kb = p.kc + ((int64_t)b * p.heads + hd) * p.S * XD + c4 * 16;
    const __nv_bfloat16* vb = p.vc + ((int64_t)b * p.heads + hd) * p.S * XD + c4 * 16;
    float* scw = sc + warp * sc_cap;
    float qv[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) qv[e] = qs[warp * XD + c4 * 16 + e];
    const float neg = finfo_min<__nv_bfloat16>();
    constexpr int UP = 6;  // passes (of 8 keys) in flight per batch
    float lmax = -INFINITY;
    for (int j0 = 0; j0 < p.S; j0 += 8 * UP) {
      uint4 raw[UP][2];
#pragma unroll
      for (int u = 0; u < UP; ++u) {
        const int j = j0 + u * 8 + kg;
        if (j < p.S) {
          raw[u][0] = __ldcg(reinterpret_cast<const uint4*>(kb + (int64_t)j * XD));
          raw[u][1] = __ldcg(reinterpret_cast<const uint4*>(kb + (int64_t)j * XD + 8));
        }
      }
#pragma unroll
      for (int u = 0; u < UP; ++u) {
        const int j = j0 + u * 8 + kg;
        float dot = 0.f;
        if (j < p.S) {
#pragma unroll
          for (int hlf = 0; hlf < 2; ++hlf) {
            const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&raw[u][hlf]);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float2 f = __bfloat1622float2(h2[e]);
              dot = fmaf(qv[hlf * 8 + 2 * e], f.x, dot);
              dot = fmaf(qv[hlf * 8 + 2 * e + 1], f.y, dot);
            }
          }
        }
        dot += __shfl_xor_sync(0xffffffffu, dot, 1);
        dot += __shfl_xor_sync(0xffffffffu, dot, 2);
        if (j < p.S && c4 == 0) {
          bool keep = true;
          if (p.key_mask) keep = p.key_mask[(int64_t)b * p.S + j] != 0;
          const float s = keep ? dot : neg;
          scw[j] = s;
          lmax = fmaxf(lmax, s);
        }
      }
    }
    const float mx = warp_max(lmax);
    __syncwarp();
    float lsum = 0.f;
    for (int j = lane; j < p.S; j += 32) {
      const float e = expf(scw[j] - mx);
      scw[j] = e;
      lsum += e;
    }
    const float inv = 1.0f / warp_sum(lsum);
    __syncwarp();
    float acc[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) acc[e] = 0.f;
    for (int j0 = 0; j0 < p.S; j0 += 8 * UP) {
      uint4 raw[UP][2];
#pragma unroll
      for (int u = 0; u < UP; ++u) {
        const int j = j0 + u * 8 + kg;
        if (j < p.S) {
          raw[u][0] = __ldcg(reinterpret_cast<const uint4*>(vb + (int64_t)j * XD));
          raw[u][1] = __ldcg(reinterpret_cast<const uint4*>(vb + (int64_t)j * XD + 8));
        }
      }
#pragma unroll
      for (int u = 0; u < UP; ++u) {
        const int j = j0 + u * 8 + kg;
        if (j < p.S) {
          const float pj = scw[j] * inv;
#pragma unroll
          for (int hlf = 0; hlf < 2; ++hlf) {
            const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&raw[u][hlf]);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float2 f = __bfloat1622float2(h2[e]);
              acc[hlf * 8 + 2 * e] = fmaf(pj, f.x, acc[hlf * 8 + 2 * e]);
              acc[hlf * 8 + 2 * e + 1] = fmaf(pj, f.y, acc[hlf * 8 + 2 * e + 1]);
            }
          }
        }
      }
    }
#pragma unroll
    for (int e = 0; e < 16; ++e) {  // sum over the 8 key groups of the warp (fixed order)
      acc[e] += __shfl_xor_sync(0xffffffffu, acc[e], 4);
      acc[e] += __shfl_xor_sync(0xffffffffu, acc[e], 8);
      acc[e] += __shfl_xor_sync(0xffffffffu, acc[e], 16);
    }
    if (kg == 0) {
#pragma unroll
      for (int hlf = 0; hlf < 2; ++hlf) {
        uint4 u;
        uint32_t* w32 = reinterpret_cast<uint32_t*>(&u);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const __nv_bfloat162 t2 = __floats2bfloat162_rn(acc[hlf * 8 + 2 * e], acc[hlf * 8 + 2 * e + 1]);
          w32[e] = *reinterpret_cast<const uint32_t*>(&t2);
        }
        *reinterpret_cast<uint4*>(at + swz(warp, c4 * 2 + hlf, 128)) = u;
      }
    }
  } else if (lane < 8) {
    *reinterpret_cast<uint4*>(at + swz(warp, lane, 128)) = make_uint4(0u, 0u, 0u, 0u);
  }
  if (dbg) c_st[3] = clock64();
  cp_async_wait<0>();
  __syncthreads();  // attention tile complete, the Wo slab has landed
  if (dbg) c_st[4] = clock64();

  // ---- phase 4: part_out[head][b][n] = a[b][0:64] . Wo[n][head*64 : head*64+64]; warp w owns n in [w*H/8, (w+1)*H/8)
  {
    const int mi = lane >> 3, ri = lane & 7;
    uint32_t a_lo[4], a_hi[4];  // k 0..31 and k 32..63 of the 8 x 64 tile
    ldmatrix_x4(a_lo, at_s + ri * 128 + ((mi ^ ri) << 4));
    ldmatrix_x4(a_hi, at_s + ri * 128 + (((4 + mi) ^ ri) << 4));
    const uint32_t f0[4] = {a_lo[0], 0u, a_lo[1], 0u}, f1[4] = {a_lo[2], 0u, a_lo[3], 0u};
    const uint32_t f2[4] = {a_hi[0], 0u, a_hi[1], 0u}, f3[4] = {a_hi[2], 0u, a_hi[3], 0u};
    const int g = lane >> 2, t = lane & 3;
    float* orow = p.part_out + ((int64_t)hd * p.B + b0 + g) * H + 2 * t;
    constexpr int NTW = H / 8 / 8;  // n-tiles per warp
#pragma unroll 4
    for (int nt = 0; nt < NTW; ++nt) {
      const int n0 = warp * (H / 8) + nt * 8;
      uint32_t w_lo[4], w_hi[4];
      const uint32_t rb = wbuf_s + (n0 + ri) * 128;
      ldmatrix_x4(w_lo, rb + ((mi ^ ri) << 4));
      ldmatrix_x4(w_hi, rb + (((4 + mi) ^ ri) << 4));
      float c[4] = {0.f, 0.f, 0.f, 0.f}, d[4] = {0.f, 0.f, 0.f, 0.f};
      mma_bf16(c, f0, w_lo[0], w_lo[1]);
      mma_bf16(d, f1, w_lo[2], w_lo[3]);
      mma_bf16(c, f2, w_hi[0], w_hi[1]);
      mma_bf16(d, f3, w_hi[2], w_hi[3]);
      if (g < nrows) *reinterpret_cast<float2*>(orow + n0) = make_float2(c[0] + d[0], c[1] + d[1]);
    }
  }
  if (dbg && tid == 0) {
    const long long c_end = clock64();
#pragma unroll
    for (int i = 0; i < 5; ++i) atomicAdd(g_xb_agg + i, (unsigned long long)(c_st[i] - c_dep));
    atomicAdd(g_xb_agg + 5, (unsigned long long)(c_end - c_dep));
    atomicAdd(g_xb_agg + 6, 1ull);
    atomicMax(g_xb_agg + 7, (unsigned long long)(c_end - c_dep));
  }
  tl_end(tl);
}

}  // namespace

int xblock_read_stamps(unsigned long long* host, int reset) {
  if (reset) {
    static unsigned long long zero[8];
    cudaMemcpyToSymbol(g_xb_agg, zero, sizeof(zero));
    return 0;
  }
  cudaMemcpyFromSymbol(host, g_xb_agg, sizeof(unsigned long long) * 8);
  return 8;
}

// PTTS_XBLOCK=0 keeps the four separate launches (LayerNorm, q GEMM, cross flash-decode, out_proj GEMM)
bool xblock_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("PTTS_XBLOCK");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v == 1;
}

static size_t xblock_smem(int H, int S) {
  const int sc_cap = (S + 7) / 8 * 8;
  return (size_t)XD * 2 * H + (size_t)XG * 2 * H + sizeof(float) * XG * XD + (size_t)XG * XD * 2 +
         sizeof(float) * XG * sc_cap;
}

bool xblock_supported(int H, int heads, int d, int S) {
  if (!xblock_enabled()) return false;
  if (d != XD || heads * d != H || (H != 1024 && H != 1536) || S < 1) return false;
  return xblock_smem(H, S) <= 227 * 1024;
}

int xblock_fast(const float* h, const float* part_in, int S_in, int64_t stride, const float* gamma, const float* beta,
                float eps, const void* wq, const void* wo, const void* kc, const void* vc, const int* key_mask, int B,
                int heads, int H, int S, float scale, float* part_out, const void* pf_next, int pf_stride,
                int pf_rows_base, const int* pf_rows_dyn, cudaStream_t st) {
  PTTS_CHECK_ARG(xblock_supported(H, heads, H / heads, S), "xblock_fast: unsupported shape");
  XBlockParams p = {};
  p.h = h; p.part_in = part_in; p.S_in = S_in; p.stride = stride; p.gamma = gamma; p.beta = beta; p.eps = eps;
  p.wq = (const __nv_bfloat16*)wq; p.wo = (const __nv_bfloat16*)wo;
  p.kc = (const __nv_bfloat16*)kc; p.vc = (const __nv_bfloat16*)vc; p.key_mask = key_mask;
  p.B = B; p.heads = heads; p.S = S; p.scale = scale; p.part_out = part_out;
  static int pf_pct = -1;  // same knob as the cross flash-decode kernel (fast_attn.cu)
  if (pf_pct < 0) {
    const char* e = getenv("PTTS_L2PF_PCT");
    pf_pct = e ? atoi(e) : 50;
    const char* m = getenv("PTTS_L2PF");
    if (m && !(atoi(m) & 2)) pf_pct = 0;
  }
  p.pf_next = (const __nv_bfloat16*)pf_next; p.pf_stride = pf_stride; p.pf_rows_base = pf_rows_base;
  p.pf_rows_dyn = pf_rows_dyn; p.pf_pct = pf_pct;
  p.hint = take_l2_hint();
  const size_t smem = xblock_smem(H, S);
  const int sc_cap = (S + 7) / 8 * 8;
  const dim3 grid(heads, (B + XG - 1) / XG);
  static int dbg = -1;
  if (dbg < 0) {
    const char* ev = getenv("PTTS_XB_DBG");
    dbg = ev ? atoi(ev) : 0;
  }
  cudaError_t e;
  if (H == 1024) {
    static size_t configured = 0;
    if (smem > configured) {
      cudaFuncSetAttribute(xblock_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      configured = smem;
    }
    e = launch_k(xblock_kernel<8>, grid, dim3(XT), smem, st, p, sc_cap, dbg, tl_next(5));
  } else {
    static size_t configured = 0;
    if (smem > configured) {
      cudaFuncSetAttribute(xblock_kernel<12>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      configured = smem;
    }
    e = launch_k(xblock_kernel<12>, grid, dim3(XT), smem, st, p, sc_cap, dbg, tl_next(5));
  }
  if (e != cudaSuccess) {
    set_error("xblock_fast launch: %s", cudaGetErrorString(e));
    return PTTS_E_LAUNCH;
  }
  PTTS_CHECK_LAUNCH("xblock_fast");
  return PTTS_OK;
}

}  // namespace ptts
