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
  float* part_out;             // [heads][B][H], or (out_by_row) [B][heads][H]
  int out_by_row;
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

__device__ __forceinline__ uint32_t cluster_rank_x() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;\n" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_arrive() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
}
__device__ __forceinline__ void cluster_wait() {
  asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
__device__ __forceinline__ void st_dsmem_v2(uint32_t local_addr, uint32_t rank, uint32_t x, uint32_t y) {
  uint32_t ra;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;\n" : "=r"(ra) : "r"(local_addr), "r"(rank));
  asm volatile("st.shared::cluster.v2.b32 [%0], {%1, %2};\n" ::"r"(ra), "r"(x), "r"(y) : "memory");
}

// grid (heads, ceil(B / 8)), cluster (CS, 1, 1), CS in {1, 2, 4}: the CTAs of a cluster are CS heads of the SAME group
// of 8 utterances and share the fold + LayerNorm of the group's rows: cluster rank r normalises rows [r*8/CS, (r+1)*8/CS)
// (32*CS threads per row) and pushes them into the xn tile of all CS CTAs through distributed shared memory.  Every
// head needs every row; computed per CTA (CS = 1) the fold costs 2.3 us of L2 bandwidth (128 CTAs x 160 KB), while
// CS = 8 (one row per CTA) cannot be placed next to the neighbouring kernels' CTAs: measured 12 us between the first
// and the last CTA passing the dependency wait.
__device__ __forceinline__ void xb_tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void xb_mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  for (uint32_t spin = 0; !done; ++spin) {
    asm volatile(
        "{\n\t.reg .pred P1;\n\tmbarrier.test_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.b32 %0, 1, 0, P1;\n\t}\n"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (spin > 200000000u) __trap();  // a protocol bug must fail loudly, not hang the box
  }
}

// Shared-memory image of both weight slabs = what a SWIZZLE_128B tensor-map copy writes: blocks of [rows][128 bytes]
// whose 16-byte chunk c of row r sits at chunk position c ^ (r & 7) (base 1024-byte aligned).
//   Wq head slab: H/64 blocks (one per 64-element k block) of 64 rows   -> block kb at kb * 8 KB
//   Wo head slab: one block of H rows (the head's 64 input columns)
// LEAN: S <= 48 keys (all in registers), no hint, no next-layer prefetch, no diagnostics - the optional paths are
// compiled out of the instantiation the benchmark shape launches (one-shot kernel: instruction fetch is on its
// critical path, see fast_gemm3.cu).
// PLAIN (lean launches without a key mask and with B % 8 == 0): no mask tests, no ragged-group predicates
template <int HT, int CS, bool LEAN, bool PLAIN = false>  // H = 128 * HT
__global__ void __launch_bounds__(XT, 1)
xblock_kernel(const XBlockParams p, int sc_cap, int dbg_in, const __grid_constant__ CUtensorMap tm_wq,
              const __grid_constant__ CUtensorMap tm_wo, TL tl_in) {
  const int dbg = LEAN ? (dbg_in & 2) : dbg_in;
  TL tl = tl_in;
  if constexpr (LEAN) tl.p = nullptr;
  constexpr int H = 128 * HT;
  constexpr int ROWB = 2 * H;  // bytes per bf16 row of H
  constexpr int UP = 6;        // key passes (8 keys each) whose K and V rows a warp keeps in registers: 48 keys
  extern __shared__ __align__(128) unsigned char smraw0[];
  unsigned char* smraw = smraw0 + ((1024u - (smem_u32(smraw0) & 1023u)) & 1023u);  // swizzle atoms: 1024-byte aligned
  unsigned char* wbuf = smraw;                         // Wq head slab [64][H] bf16, later Wo head slab [H][64] bf16
  unsigned char* xn = wbuf + XD * ROWB;                // [8][H] bf16 (swizzled); later the q split-K reduction tile
  float* qs = reinterpret_cast<float*>(xn + XG * ROWB);  // [8][64] scaled q (rounded to bf16 like the q buffer)
  unsigned char* at = reinterpret_cast<unsigned char*>(qs + XG * XD);  // [8][64] bf16 attention output (swizzled)
  float* sc = reinterpret_cast<float*>(at + XG * XD * 2);              // [8][sc_cap] scores -> probabilities
  float* gb = sc + XG * sc_cap;                        // [2][H] gamma | beta (only when H <= 1024: shared-memory budget)
  constexpr bool kStageGB = HT <= 8;
  __shared__ float red[16];
  __shared__ __align__(8) uint64_t wq_bar, wo_bar;
  const int hd = blockIdx.x, b0 = blockIdx.y * XG;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nrows = PLAIN ? XG : min(XG, p.B - b0);
  const int* key_mask = PLAIN ? nullptr : p.key_mask;
  const uint32_t wbuf_s = smem_u32(wbuf), xn_s = smem_u32(xn), at_s = smem_u32(at);
  const uint32_t crank = cluster_rank_x();
  tl_start(tl);
  if constexpr (CS > 1) cluster_arrive();  // "my shared memory is live": peers wait for this before they push rows

  // ---- before the dependency: everything below only touches immutable data (weights, the cross cache, cache rows of
  // the next self-attention launch that are older than this step)
  if (warp == 0 && elect_one()) {  // single-lane issue from uniform registers (common.cuh)
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_u32(&wq_bar)), "r"(1) : "memory");
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_u32(&wo_bar)), "r"(1) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
    // the head's 64 rows of Wq as H/64 box copies (64 rows x 128 bytes each) - one thread, the TMA engine does the rest
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(smem_u32(&wq_bar)), "r"(XD * ROWB) : "memory");
    for (int kb = 0; kb < H / 64; ++kb) xb_tma_load_2d(wbuf_s + kb * 8192, &tm_wq, smem_u32(&wq_bar), kb * 64, hd * XD);
  }
  if constexpr (kStageGB) {  // gamma / beta are immutable: stage them now, the normalise step reads shared memory
    for (int i = tid; i < H / 4; i += XT) {
      cp_async16(smem_u32(gb) + i * 16, p.gamma + i * 4, true);
      cp_async16(smem_u32(gb + H) + i * 16, p.beta + i * 4, true);
    }
    cp_async_commit();
  }
  if (!(dbg & 2)) pdl_launch();  // PTTS_XB_EARLY=0 (dbg bit 1): release the next LayerNorm only after the wait
  // K and V rows of utterance b0 + warp for this head travel into registers now (4 lanes per 128-byte row, 8 rows per
  // pass): by the time q exists they have long landed, and the attention phase is pure arithmetic
  const int kg = lane >> 2, c4 = lane & 3;
  const int bw = b0 + warp;
  const bool live = warp < nrows;
  const __nv_bfloat16* kb = p.kc + ((int64_t)(live ? bw : b0) * p.heads + hd) * p.S * XD + c4 * 16;
  const __nv_bfloat16* vb = p.vc + ((int64_t)(live ? bw : b0) * p.heads + hd) * p.S * XD + c4 * 16;
  uint4 rk[UP][2], rv[UP][2];
  if (live) {
#pragma unroll
    for (int u = 0; u < UP; ++u) {
      const int j = u * 8 + kg;
      if (j < p.S) {
        rk[u][0] = __ldcg(reinterpret_cast<const uint4*>(kb + (int64_t)j * XD));
        rk[u][1] = __ldcg(reinterpret_cast<const uint4*>(kb + (int64_t)j * XD + 8));
      }
    }
#pragma unroll
    for (int u = 0; u < UP; ++u) {
      const int j = u * 8 + kg;
      if (j < p.S) {
        rv[u][0] = __ldcg(reinterpret_cast<const uint4*>(vb + (int64_t)j * XD));
        rv[u][1] = __ldcg(reinterpret_cast<const uint4*>(vb + (int64_t)j * XD + 8));
      }
    }
  }
  // the key mask of the description is immutable too: the lanes that will own the scores of the register-held keys
  // (kg == 0: keys u*8 + 2*c4 + e) fetch their 12 mask words now - in the attention phase a mask load would be one more
  // L2 round trip on the chain (masked batches ran 1.259 against 1.203 ms per decode step)
  uint32_t mbits = 0xffffffffu;
  if (key_mask && live && kg == 0) {
    mbits = 0u;
#pragma unroll
    for (int u = 0; u < UP; ++u)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int j = u * 8 + 2 * c4 + e;
        if (j < p.S && __ldg(key_mask + (int64_t)bw * p.S + j) != 0) mbits |= 1u << (2 * u + e);
      }
  }
  if constexpr (!LEAN) l2_hint_prefetch(p.hint, blockIdx.y * gridDim.x + blockIdx.x, gridDim.x * gridDim.y);
  for (int n = tid; n < H; n += XT)  // Wo[:, head]: one 128-byte line per output feature
    asm volatile("prefetch.global.L2 [%0];\n" ::"l"(p.wo + (int64_t)n * H + hd * XD) : "memory");
  if (!LEAN && p.S > 8 * UP && live && lane < 2) {  // keys beyond the register-held ones: L2 prefetch of K (lane 0) / V (lane 1)
    const char* base = reinterpret_cast<const char*>((lane ? p.vc : p.kc) + (((int64_t)bw * p.heads + hd) * p.S + 8 * UP) * XD);
    const int64_t bytes = (int64_t)(p.S - 8 * UP) * XD * 2;
    for (int64_t o = 0; o < bytes; o += 8192) l2_prefetch_bulk(base + o, (uint32_t)min((int64_t)8192, bytes - o));
  }
  if (!LEAN && p.pf_next && p.pf_pct > 0 && live) {
    const int rows = dyn_value(p.pf_rows_base, p.pf_rows_dyn) * min(p.pf_pct, 100) / 100;
    const int64_t bytes = (int64_t)rows * XD * 2;
    const char* src = reinterpret_cast<const char*>(p.pf_next + ((int64_t)bw * p.heads + hd) * p.pf_stride * XD);
    for (int64_t o = (int64_t)lane * 8192; o < bytes; o += 32 * 8192)
      l2_prefetch_bulk(src + o, (uint32_t)min((int64_t)8192, bytes - o));
  }
  if constexpr (kStageGB) cp_async_wait<0>();
  __syncthreads();  // gamma / beta staged; the barrier initialisations are visible to every thread
  pdl_wait();
  if (dbg & 2) pdl_launch();
  tl_dep(tl);
  const long long c_dep = (dbg & 1) ? clock64() : 0;
  long long c_st[5] = {0, 0, 0, 0, 0};

  // ---- phase 1: fold + LayerNorm; team `team` (TS = 32*CS threads) of this CTA owns row crank*RP + team of the group
  {
    constexpr int RP = XG / CS;                 // rows normalised by this CTA
    constexpr int TS = XT / RP;                 // threads per row
    constexpr int V4 = (H / 4 + TS - 1) / TS;   // float4 per thread
    const int team = tid / TS, tt = tid % TS;
    const int row = (int)crank * RP + team;     // row of the group
    const bool mine = row < nrows;
    const int b = b0 + row;
    float4 v[V4];
#pragma unroll
    for (int i = 0; i < V4; ++i) v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (mine) {
      const float4* hr = reinterpret_cast<const float4*>(p.h + (int64_t)b * H);
#pragma unroll
      for (int i = 0; i < V4; ++i)
        if (tt + i * TS < H / 4) v[i] = __ldcg(hr + tt + i * TS);
      constexpr int SB = V4 <= 4 ? 4 : (V4 <= 8 ? 2 : 1);  // slices in flight (register budget: K/V rows are held too)
      for (int s0 = 0; s0 < p.S_in; s0 += SB) {  // added in slice order (deterministic)
        float4 t[SB][V4];
#pragma unroll
        for (int u = 0; u < SB; ++u)
          if (s0 + u < p.S_in) {
            const float4* ps = reinterpret_cast<const float4*>(p.part_in + (int64_t)(s0 + u) * p.stride + (int64_t)b * H);
#pragma unroll
            for (int i = 0; i < V4; ++i)
              if (tt + i * TS < H / 4) t[u][i] = __ldcg(ps + tt + i * TS);
          }
#pragma unroll
        for (int u = 0; u < SB; ++u)
          if (s0 + u < p.S_in) {
#pragma unroll
            for (int i = 0; i < V4; ++i)
              if (tt + i * TS < H / 4) {
                v[i].x += t[u][i].x; v[i].y += t[u][i].y; v[i].z += t[u][i].z; v[i].w += t[u][i].w;
              }
          }
      }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < V4; ++i) s += (v[i].x + v[i].y) + (v[i].z + v[i].w);  // threads past H / 4 hold zeros
    s = warp_sum(s);
    if constexpr (CS > 1) {  // a row spans CS warps
      if (lane == 0) red[warp] = s;
      __syncthreads();
      s = 0.f;
#pragma unroll
      for (int w = 0; w < CS; ++w) s += red[team * CS + w];
    }
    const float mean = s / (float)H;
    float qq = 0.f;
#pragma unroll
    for (int i = 0; i < V4; ++i)
      if (tt + i * TS < H / 4) {
        const float a = v[i].x - mean, bb = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
        qq += (a * a + bb * bb) + (c * c + d * d);
      }
    qq = warp_sum(qq);
    if constexpr (CS > 1) {
      if (lane == 0) red[8 + warp] = qq;
      __syncthreads();
      qq = 0.f;
#pragma unroll
      for (int w = 0; w < CS; ++w) qq += red[8 + team * CS + w];
    }
    const float rstd = rsqrtf(qq / (float)H + p.eps);
    if constexpr (CS > 1) cluster_wait();  // every CTA of the cluster is running: its shared memory may be written
#pragma unroll
    for (int i = 0; i < V4; ++i) {
      const int c4i = tt + i * TS;
      if (c4i < H / 4) {
        const int col = c4i * 4;
        uint32_t lo = 0u, hi = 0u;
        if (mine) {
          const float4 g = kStageGB ? *reinterpret_cast<const float4*>(gb + col) : __ldg(reinterpret_cast<const float4*>(p.gamma + col));
          const float4 be = kStageGB ? *reinterpret_cast<const float4*>(gb + H + col) : __ldg(reinterpret_cast<const float4*>(p.beta + col));
          const __nv_bfloat162 l2 = __floats2bfloat162_rn((v[i].x - mean) * rstd * g.x + be.x, (v[i].y - mean) * rstd * g.y + be.y);
          const __nv_bfloat162 h2 = __floats2bfloat162_rn((v[i].z - mean) * rstd * g.z + be.z, (v[i].w - mean) * rstd * g.w + be.w);
          lo = *reinterpret_cast<const uint32_t*>(&l2);
          hi = *reinterpret_cast<const uint32_t*>(&h2);
        }
        const uint32_t off = swz(row, col >> 3, ROWB) + (col & 7) * 2;
        if constexpr (CS > 1) {
#pragma unroll
          for (uint32_t r = 0; r < CS; ++r) st_dsmem_v2(xn_s + off, r, lo, hi);
        } else {
          *reinterpret_cast<uint2*>(xn + off) = make_uint2(lo, hi);
        }
      }
    }
  }
  if (dbg & 1) c_st[0] = clock64();
  xb_mbar_wait(smem_u32(&wq_bar), 0);  // the Wq slab has landed (async proxy -> visible after the barrier wait)
  if constexpr (CS > 1) {
    cluster_arrive();
    cluster_wait();  // all 8 rows of xn are here (pushed by the CS CTAs), and this CTA's Wq slab has landed
  } else {
    __syncthreads();
  }
  if (dbg & 1) c_st[1] = clock64();

  // ---- phase 2: q[8][64] = xn . Wq_head^T; warp w owns output features [8w, 8w + 8)  (splitting K over the warps
  // instead reads every operand byte once but measured slower: 1.73 us against 1.24 us with its reduction)
  {
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    const int mi = lane >> 3, ri = lane & 7;  // ldmatrix: lane supplies row ri of 8x8 matrix mi (= k chunk kc + mi)
    const uint32_t a_base = xn_s + ri * ROWB, b_base = wbuf_s + (warp * 8 + ri) * 128;
#pragma unroll 4
    for (int kc = 0; kc < ROWB / 16; kc += 8) {  // 64 k per iteration (= one 8 KB block of the slab): four k16 steps
      uint32_t a[4], bq[4], a2[4], bq2[4];
      const uint32_t off = (uint32_t)(((kc + mi) ^ ri) << 4), off2 = (uint32_t)(((kc + 4 + mi) ^ ri) << 4);
      const uint32_t bb = b_base + (kc >> 3) * 8192;
      ldmatrix_x4(a, a_base + off);
      ldmatrix_x4(bq, bb + ((mi ^ ri) << 4));
      ldmatrix_x4(a2, a_base + off2);
      ldmatrix_x4(bq2, bb + (((4 + mi) ^ ri) << 4));
      const uint32_t f0[4] = {a[0], 0u, a[1], 0u}, f1[4] = {a[2], 0u, a[3], 0u};
      const uint32_t f2[4] = {a2[0], 0u, a2[1], 0u}, f3[4] = {a2[2], 0u, a2[3], 0u};
      mma_bf16(acc[0], f0, bq[0], bq[1]);
      mma_bf16(acc[1], f1, bq[2], bq[3]);
      mma_bf16(acc[2], f2, bq2[0], bq2[1]);
      mma_bf16(acc[3], f3, bq2[2], bq2[3]);
    }
    const int g = lane >> 2, t = lane & 3;
    const float q0 = (acc[0][0] + acc[1][0]) + (acc[2][0] + acc[3][0]);
    const float q1 = (acc[0][1] + acc[1][1]) + (acc[2][1] + acc[3][1]);
    float* qd = qs + g * XD + warp * 8 + 2 * t;
    qd[0] = __bfloat162float(__float2bfloat16_rn(q0 * p.scale));
    qd[1] = __bfloat162float(__float2bfloat16_rn(q1 * p.scale));
  }
  __syncthreads();  // q visible; every warp is done with the Wq slab
  if (dbg & 1) c_st[2] = clock64();

  // ---- the Wo head slab (L2-resident by now) replaces the Wq slab while the attention phase runs: H/256 box copies
  if (warp == 0 && elect_one()) {
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");  // generic-proxy reads of the Wq slab are done
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(smem_u32(&wo_bar)), "r"(H * 128) : "memory");
    for (int nb = 0; nb < H / 256; ++nb) xb_tma_load_2d(wbuf_s + nb * 32768, &tm_wo, smem_u32(&wo_bar), hd * XD, nb * 256);
  }

  // ---- phase 3: cross-attention of utterance b0 + warp over its S cached keys (4 lanes per key row, 8 rows per pass)
  if (live) {
    float* scw = sc + warp * sc_cap;
    float qv[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) qv[e] = qs[warp * XD + c4 * 16 + e];
    const float neg = finfo_min<__nv_bfloat16>();
    float lmax = -INFINITY;
    auto score_batch = [&](const uint4 (&raw)[UP][2], int j0) {
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
          if (key_mask) keep = key_mask[(int64_t)bw * p.S + j] != 0;
          const float sj = keep ? dot : neg;
          scw[j] = sj;
          lmax = fmaxf(lmax, sj);
        }
      }
    };
    auto load_batch = [&](uint4 (&raw)[UP][2], const __nv_bfloat16* base, int j0) {
#pragma unroll
      for (int u = 0; u < UP; ++u) {
        const int j = j0 + u * 8 + kg;
        if (j < p.S) {
          raw[u][0] = __ldcg(reinterpret_cast<const uint4*>(base + (int64_t)j * XD));
          raw[u][1] = __ldcg(reinterpret_cast<const uint4*>(base + (int64_t)j * XD + 8));
        }
      }
    };
    {
      // the first 48 keys on the tensor core: scores[key] = sum_d q[d] K[key][d] as m16n8k16 tiles whose row 0 is the
      // query and whose 8 columns are the 8 key rows of a pass.  The lane (kg, c4) already holds K[u*8 + kg][c4*16 ..
      // c4*16 + 15] - exactly a B fragment if the k slots of step kk are mapped to head dims c4*16 + 4kk + {0,1} (b0)
      // and c4*16 + 4kk + 2 + {0,1} (b1); the A fragment uses the same mapping, so no data moves.
      uint32_t qa[4][2];
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        const __nv_bfloat162 lo2 = __floats2bfloat162_rn(qv[4 * kk], qv[4 * kk + 1]);      // exact: q is bf16-rounded
        const __nv_bfloat162 hi2 = __floats2bfloat162_rn(qv[4 * kk + 2], qv[4 * kk + 3]);
        qa[kk][0] = kg == 0 ? *reinterpret_cast<const uint32_t*>(&lo2) : 0u;
        qa[kk][1] = kg == 0 ? *reinterpret_cast<const uint32_t*>(&hi2) : 0u;
      }
#pragma unroll
      for (int u = 0; u < UP; ++u) {
        float c[4] = {0.f, 0.f, 0.f, 0.f};
        const uint32_t kb0[4] = {rk[u][0].x, rk[u][0].z, rk[u][1].x, rk[u][1].z};
        const uint32_t kb1[4] = {rk[u][0].y, rk[u][0].w, rk[u][1].y, rk[u][1].w};
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          const uint32_t a[4] = {qa[kk][0], 0u, qa[kk][1], 0u};
          mma_bf16(c, a, kb0[kk], kb1[kk]);
        }
        if (kg == 0) {  // row 0 of the tile: lane c4 holds the scores of keys u*8 + 2*c4 and + 1
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const int j = u * 8 + 2 * c4 + e;
            if (j < p.S) {
              const bool keep = (mbits >> (2 * u + e)) & 1u;
              const float sj = keep ? c[e] : neg;
              scw[j] = sj;
              lmax = fmaxf(lmax, sj);
            }
          }
        }
      }
    }
    for (int j0 = 8 * UP; !LEAN && j0 < p.S; j0 += 8 * UP) {  // descriptions longer than 48 tokens
      uint4 raw[UP][2];
      load_batch(raw, kb, j0);
      score_batch(raw, j0);
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
    auto accum_batch = [&](const uint4 (&raw)[UP][2], int j0) {
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
    };
    accum_batch(rv, 0);
    for (int j0 = 8 * UP; !LEAN && j0 < p.S; j0 += 8 * UP) {
      uint4 raw[UP][2];
      load_batch(raw, vb, j0);
      accum_batch(raw, j0);
    }
    // sum over the 8 key groups of the warp (lane bits 2..4) as a butterfly that halves the values each lane keeps:
    // 8 + 4 + 2 shuffles instead of 3 x 16, fixed order; the lane ends up with two neighbouring head dims
    const bool b2 = lane & 4, b3 = lane & 8, b4 = lane & 16;
    float r8[8], r4[4], r2[2];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float send = b2 ? acc[i] : acc[8 + i], keep = b2 ? acc[8 + i] : acc[i];
      r8[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float send = b3 ? r8[i] : r8[4 + i], keep = b3 ? r8[4 + i] : r8[i];
      r4[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const float send = b4 ? r4[i] : r4[2 + i], keep = b4 ? r4[2 + i] : r4[i];
      r2[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
    {
      const int d0 = c4 * 16 + (b2 ? 8 : 0) + (b3 ? 4 : 0) + (b4 ? 2 : 0);
      const __nv_bfloat162 t2 = __floats2bfloat162_rn(r2[0], r2[1]);
      *reinterpret_cast<uint32_t*>(at + swz(warp, d0 >> 3, 128) + (d0 & 7) * 2) = *reinterpret_cast<const uint32_t*>(&t2);
    }
  } else if (lane < 8) {
    *reinterpret_cast<uint4*>(at + swz(warp, lane, 128)) = make_uint4(0u, 0u, 0u, 0u);
  }
  if (dbg & 1) c_st[3] = clock64();
  xb_mbar_wait(smem_u32(&wo_bar), 0);
  __syncthreads();  // attention tile complete, the Wo slab has landed
  if (dbg & 1) c_st[4] = clock64();

  // ---- phase 4: part_out[head][b][n] = a[b][0:64] . Wo[n][head*64 : head*64+64]; warp w owns n in [w*H/8, (w+1)*H/8)
  {
    const int mi = lane >> 3, ri = lane & 7;
    uint32_t a_lo[4], a_hi[4];  // k 0..31 and k 32..63 of the 8 x 64 tile
    ldmatrix_x4(a_lo, at_s + ri * 128 + ((mi ^ ri) << 4));
    ldmatrix_x4(a_hi, at_s + ri * 128 + (((4 + mi) ^ ri) << 4));
    const uint32_t f0[4] = {a_lo[0], 0u, a_lo[1], 0u}, f1[4] = {a_lo[2], 0u, a_lo[3], 0u};
    const uint32_t f2[4] = {a_hi[0], 0u, a_hi[1], 0u}, f3[4] = {a_hi[2], 0u, a_hi[3], 0u};
    const int g = lane >> 2, t = lane & 3;
    float* orow = p.part_out + (p.out_by_row ? ((int64_t)(b0 + g) * p.heads + hd) : ((int64_t)hd * p.B + b0 + g)) * H + 2 * t;
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
  if ((dbg & 1) && tid == 0) {
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

// diagnostic build, PTTS_XBLOCK=0: keeps the four separate launches (LayerNorm, q GEMM, cross flash-decode, out_proj GEMM)
bool xblock_enabled() { return PTTS_KNOB("PTTS_XBLOCK", 1) != 0; }

static size_t xblock_smem(int H, int S) {
  const int sc_cap = (S + 7) / 8 * 8;
  return 1024 + (size_t)XD * 2 * H + (size_t)XG * 2 * H + sizeof(float) * XG * XD + (size_t)XG * XD * 2 +
         sizeof(float) * XG * sc_cap + (H <= 1024 ? sizeof(float) * 2 * H : 0);
}

bool xblock_supported(int H, int heads, int d, int S) {
  if (!xblock_enabled()) return false;
  if (d != XD || heads * d != H || (H != 1024 && H != 1536) || heads % 4 != 0 || S < 1) return false;
  return xblock_smem(H, S) <= 227 * 1024 && gemm::tma_available();  // else the 12-launch chain runs
}

int xblock_fast(const float* h, const float* part_in, int S_in, int64_t stride, const float* gamma, const float* beta,
                float eps, const void* wq, const void* wo, const void* kc, const void* vc, const int* key_mask, int B,
                int heads, int H, int S, float scale, float* part_out, int out_by_row, const void* pf_next,
                int pf_stride, int pf_rows_base, const int* pf_rows_dyn, cudaStream_t st) {
  PTTS_CHECK_ARG(xblock_supported(H, heads, H / heads, S), "xblock_fast: unsupported shape");
  XBlockParams p = {};
  p.h = h; p.part_in = part_in; p.S_in = S_in; p.stride = stride; p.gamma = gamma; p.beta = beta; p.eps = eps;
  p.wq = (const __nv_bfloat16*)wq; p.wo = (const __nv_bfloat16*)wo;
  p.kc = (const __nv_bfloat16*)kc; p.vc = (const __nv_bfloat16*)vc; p.key_mask = key_mask;
  p.B = B; p.heads = heads; p.S = S; p.scale = scale; p.part_out = part_out; p.out_by_row = out_by_row;
  // share of the next self-attention launch's K rows to pull into L2 from here.  Measured on B200 with this kernel
  // (B=64, 860 frames): 50 % 1.694, 25 % 1.674, 0 % 1.652 ms per decode step - the prefetch queues in front of the
  // block's own slab loads: 0
  const int pf_pct = PTTS_KNOB("PTTS_XB_L2PF_PCT", 0);
  p.pf_next = (const __nv_bfloat16*)pf_next; p.pf_stride = pf_stride; p.pf_rows_base = pf_rows_base;
  p.pf_rows_dyn = pf_rows_dyn; p.pf_pct = pf_pct;
  p.hint = take_l2_hint();
  const size_t smem = xblock_smem(H, S);
  const int sc_cap = (S + 7) / 8 * 8;
  const dim3 grid(heads, (B + XG - 1) / XG);
  // bit 0: per-phase stamps (diagnostic build, PTTS_XB_DBG=1); bit 1: release the next LayerNorm AFTER the own wait:
  // measured 1.496 (before) against 1.451 ms per decode step - released early, the LayerNorm releases fc1 in turn,
  // whose TMA slab burst lands on this kernel
  const int dbg = (PTTS_KNOB("PTTS_XB_DBG", 0) & 1) | (PTTS_KNOB("PTTS_XB_EARLY", 0) ? 0 : 2);
  // CTAs (heads) that share the fold + LayerNorm of a group.  Measured (B=64, 860 frames, TMA slabs, late releases):
  // 1 -> 1.421, 2 -> 1.436 ms per decode step; 4 was 0.13 ms behind 2 with the cp.async slabs, 8 cannot be placed
  // (12 us dep-spread).  The cluster instantiations exist in the diagnostic build only (PTTS_XB_CLUSTER=2|4).
  int cs = PTTS_KNOB("PTTS_XB_CLUSTER", 1);
  if (cs != 1 && cs != 2 && cs != 4) cs = 1;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(XT);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (cs > 1) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = cs;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = 1;
    ++na;
  }
  if (pdl_enabled()) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  CUtensorMap tm_wq, tm_wo;
  if (!gemm::tensor_map_2d(wq, H, H, XD, &tm_wq) || !gemm::tensor_map_2d(wo, H, H, 256, &tm_wo)) {
    set_error("xblock_fast: no tensor map (driver entry point missing or misaligned weights)");
    return PTTS_E_UNSUPPORTED;
  }
  using KernelFn = void (*)(const XBlockParams, int, int, const CUtensorMap, const CUtensorMap, TL);
  KernelFn fn = nullptr;
  const TL tl = tl_next(5);
  const int lean_ok = PTTS_KNOB("PTTS_XB_LEAN", 1);  // diagnostic build, 0: always the full instantiation
  const bool lean = lean_ok && S <= 48 && !p.hint.ptr && !(p.pf_next && p.pf_pct > 0) && !(dbg & 1) && !tl.p;
#ifdef PTTS_DIAG
#define PTTS_XB_PICK(HT, LL) (cs == 1 ? xblock_kernel<HT, 1, LL> : (cs == 2 ? xblock_kernel<HT, 2, LL> : xblock_kernel<HT, 4, LL>))
#else
#define PTTS_XB_PICK(HT, LL) xblock_kernel<HT, 1, LL>
#endif
  if (H == 1024) {
    if (lean && cs == 1 && !key_mask && B % XG == 0) fn = xblock_kernel<8, 1, true, true>;
    else if (lean) fn = PTTS_XB_PICK(8, true);
    else fn = PTTS_XB_PICK(8, false);
  } else {
    if (lean && cs == 1 && !key_mask && B % XG == 0) fn = xblock_kernel<12, 1, true, true>;
    else if (lean) fn = PTTS_XB_PICK(12, true);
    else fn = PTTS_XB_PICK(12, false);
  }
#undef PTTS_XB_PICK
  const bool plain = lean && cs == 1 && !key_mask && B % XG == 0;
  static size_t configured[3][2][3] = {};
  size_t& conf = configured[plain ? 2 : (lean ? 1 : 0)][H == 1024 ? 0 : 1][cs == 1 ? 0 : (cs == 2 ? 1 : 2)];
  if (smem > conf) {
    cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    conf = smem;
  }
  cudaError_t e = cudaLaunchKernelEx(&cfg, fn, p, sc_cap, dbg, tm_wq, tm_wo, tl);
  if (e != cudaSuccess) {
    set_error("xblock_fast launch: %s", cudaGetErrorString(e));
    return PTTS_E_LAUNCH;
  }
  PTTS_CHECK_LAUNCH("xblock_fast");
  return PTTS_OK;
}

}  // namespace ptts
