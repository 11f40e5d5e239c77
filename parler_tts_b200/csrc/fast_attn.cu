// Flash-decoding attention for one query token over a static K/V cache (self-attention over [B][heads][c_max][64] and
// cross-attention over the cached T5 K/V): replaces bmm + additive mask + softmax + bmm of
// /root/reference/parler_tts/modeling_parler_tts.py:326-365 on the cached decode branch (299-304 / 292-294).
//
// HBM-bound by construction (every cached K and V element is read exactly once per step, 2 FLOP per byte), so the
// kernel is built around the stream: one CTA per (head, utterance[, KV split]); 128 threads read 16-byte chunks so that
// a warp covers 4 consecutive 128-byte key rows per instruction (fully coalesced), 4 independent loads in flight per
// thread; scores live in shared memory between the two passes (K pass -> softmax -> V pass).  With fewer than ~600
// (head, utterance) pairs the context is split across CTAs (split-KV) and merged by a tiny combine kernel.
// Masked keys take finfo(dtype).min like the reference's additive mask (a fully masked row averages all keys).
#include <stdlib.h>

#include "ops.cuh"

namespace ptts {

namespace {

constexpr int NT = 128;

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <typename T> struct Chunk;
template <> struct Chunk<__nv_bfloat16> {
  static constexpr int E = 8;
  static __device__ __forceinline__ void load(const __nv_bfloat16* p, float (&f)[8]) {
    const uint4 u = __ldcg(reinterpret_cast<const uint4*>(p));  // L2-coherent: written by the previous kernel
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 t = __bfloat1622float2(h[i]);
      f[2 * i] = t.x;
      f[2 * i + 1] = t.y;
    }
  }
  // streaming (read-once) load of cache rows: L2 only (.cg), no L1 allocation, coherent with the K/V write of the
  // preceding kernel
  static __device__ __forceinline__ uint4 load_raw(const __nv_bfloat16* p) {
    uint4 u;
    asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];\n"
                 : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w)
                 : "l"(p));
    return u;
  }
  static __device__ __forceinline__ void unpack(const uint4& u, float (&f)[8]) {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 t = __bfloat1622float2(h[i]);
      f[2 * i] = t.x;
      f[2 * i + 1] = t.y;
    }
  }
};
template <> struct Chunk<float> {
  static constexpr int E = 4;
  static __device__ __forceinline__ void load(const float* p, float (&f)[4]) {
    const float4 u = __ldcg(reinterpret_cast<const float4*>(p));
    f[0] = u.x; f[1] = u.y; f[2] = u.z; f[3] = u.w;
  }
  static __device__ __forceinline__ uint4 load_raw(const float* p) {
    uint4 u;
    asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];\n"
                 : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w)
                 : "l"(p));
    return u;
  }
  static __device__ __forceinline__ void unpack(const uint4& u, float (&f)[4]) {
    f[0] = __uint_as_float(u.x); f[1] = __uint_as_float(u.y); f[2] = __uint_as_float(u.z);
    f[3] = __uint_as_float(u.w);
  }
};

// (32-byte chunks - ld.global.v8 / LDG.256, new with sm_100: 4 threads per bf16 key row, two 32-byte loads in flight per
// thread and batch instead of four of 16 - were measured in this kernel on B200 at B=64, 860 frames: 1.2832 / 1.2845
// against 1.2203 / 1.2212 ms per decode step, same box; removed.  The DAC conv epilogue keeps its 32-byte row accesses:
// 138.9 against 147 - 151 ms for 32 x 860 frames, profiles/r02_wide_access.txt.)
// Optional fused prologue: the q (and, for self-attention, the new k / v) projections arrive as split-K partial sums
// [S][rows][ncols] of the tcgen05 GEMM; the CTA of (b, head) adds them up in slice order (deterministic), scales q,
// writes the new K / V row into its static-cache slot (replaces torch.cat, modeling_parler_tts.py:303-304) and goes on -
// no cluster reduction and no scatter epilogue in the GEMM, no extra launch.
struct QFuse {
  const float* part;   // null: q is read from the q buffer
  int S;
  int64_t stride;      // rows * ncols
  int ncols;           // 3H (q | k | v) or H (q only)
  int has_kv;
  float scale;
  int slot_base;
  const int* slot_dyn;
  int pf_own_v;        // prefetch this CTA's V rows into L2 while its K pass runs
  const void* pf_next; // cache of the NEXT attention launch [B][heads][pf_stride][D] whose (b, head) rows [0, pf_rows)
  int pf_stride;       //   this CTA pulls into L2 (immutable until then), or null
  int pf_rows_base;
  const int* pf_rows_dyn;
  int pf_pct;
  int pf_late;         // queue that prefetch after griddepcontrol.wait instead of before it
  int64_t pf_v_off;    // elements from pf_next to the V cache of the same launch (pf_pct > 100 also pulls V rows)
  L2Hint tail;         // matrix of the GEMM that consumes this launch (out_proj): each CTA pulls its share into L2 when
                       // its own stream is done - issued earlier it would be evicted by the K/V stream
};

// part layout: [(b*heads + h)*splits + s][D + 2] = { running max, sum of exp, unnormalised output[D] }
// LEAN (FUSED only): no K staging, no L2 prefetches, no tail hint, no timeline - the optional paths are compiled out of
// the instantiation the decode step launches by default (see the note on one-shot code size in fast_gemm3.cu).
// PLAIN (lean launches without a key mask and without a KV split): the mask test leaves the K pass and the split
// bookkeeping leaves the epilogue
template <typename T, int D, bool FUSED, bool LEAN = false, bool PLAIN = false>
__global__ void __launch_bounds__(NT, FUSED ? 7 : 8)  // 7 x 148 = 1036 CTA slots >= 64 x 16 (b, head) pairs
attn_decode_kernel(const QFuse qf, const T* __restrict__ q, const T* __restrict__ ck, const T* __restrict__ cv, int heads, int c_stride,
                   int ctx_base, const int* __restrict__ ctx_dyn, const int* __restrict__ key_mask_in, int mask_len,
                   T* __restrict__ out, float* __restrict__ part, int splits_in, int sc_cap, int stage_cap_in, TL tl_in) {
  const int* key_mask = PLAIN ? nullptr : key_mask_in;
  const int splits = PLAIN ? 1 : splits_in;
  const int stage_cap = LEAN ? 0 : stage_cap_in;
  TL tl = tl_in;
  if constexpr (LEAN) tl.p = nullptr;
  constexpr int E = Chunk<T>::E;     // elements per 16-byte chunk
  constexpr int TPK = D / E;         // threads per key row
  constexpr int KPB = NT / TPK;      // key rows per CTA pass
  constexpr int U = 4;               // independent loads in flight per thread
  constexpr int ROWB = D * (int)sizeof(T);  // bytes per cache row
  __shared__ __align__(16) T qs[D];
  extern __shared__ float sm[];
  float* sc = sm;                    // [sc_cap] scores -> probabilities
  float* red = sm + sc_cap;          // [8] block reductions
  float* ored = red + 8;             // [KPB][D] output partials
  unsigned char* ksm = reinterpret_cast<unsigned char*>(ored + KPB * D);  // [stage_cap] K rows staged before the wait
  const int hd = blockIdx.x, b = blockIdx.y, sp = blockIdx.z;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = tid / TPK, c = tid % TPK;
  const int H = heads * D;
  tl_start(tl);
  const T* kb = ck + ((int64_t)b * heads + hd) * (int64_t)c_stride * D + c * E;
  const T* vb = cv + ((int64_t)b * heads + hd) * (int64_t)c_stride * D + c * E;
  int ctx = 0, j0 = 0, j1 = 0, n = 0;
  auto plan = [&]() {
    ctx = dyn_value(ctx_base, ctx_dyn);
    const int chunk = (ctx + splits - 1) / splits;
    j0 = sp * chunk;
    j1 = min(ctx, j0 + chunk);
    n = max(j1 - j0, 0);
  };
  // Loads are software-pipelined: the next batch of U rows per thread is requested before the current one is
  // consumed, and the first V batch is requested before the softmax reductions, so HBM never idles behind a barrier.
  auto load_batch = [&](uint4 (&raw)[U], const T* base_ptr, int base) {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int j = base + u * KPB + g;
      if (j < j1) raw[u] = Chunk<T>::load_raw(base_ptr + (int64_t)j * D);
    }
  };
  uint4 cur[U], nxt[U];
  bool pre_k = false, pre_v = false;
  int staged = 0;
  auto prefetch_next = [&]() {
    const int all = dyn_value(qf.pf_rows_base, qf.pf_rows_dyn);
    const int rows = all * (qf.pf_pct > 100 ? 100 : qf.pf_pct) / 100;
    const T* src = reinterpret_cast<const T*>(qf.pf_next) + ((int64_t)b * heads + hd) * qf.pf_stride * D;
    if (qf.pf_pct < 0)
      l2_prefetch_lines(src, (int64_t)(rows * -1) * D * sizeof(T));
    else
      l2_prefetch_range(src, (int64_t)rows * D * sizeof(T));
    if (qf.pf_pct > 100 && qf.pf_v_off > 0)  // and the first (pct - 100) % of the V rows of that launch
      l2_prefetch_range(src + qf.pf_v_off, (int64_t)(all * (qf.pf_pct - 100) / 100) * D * sizeof(T));
  };
  if constexpr (FUSED) {
    // Decode-step contract of the fused entry point: cache rows below the slot this launch appends (self-attention) or
    // the whole cross-attention cache, and the position counter (written by the previous step's epilogue), are at
    // least two launches old - they may be read BEFORE griddepcontrol.wait, while the producer GEMM still runs.
    plan();
    const int old_end = qf.has_kv ? dyn_value(qf.slot_base, qf.slot_dyn) : ctx;
    // Self-attention: while the QKV GEMM still runs, the first `staged` (immutable) K rows of this CTA travel into
    // shared memory with cp.async - that share of the stream no longer sits behind the dependency.
    if (stage_cap > 0 && qf.has_kv) {
      staged = min(stage_cap, max(min(j1, old_end) - j0, 0) / (KPB * U) * (KPB * U));
      const uint32_t dst0 = smem_addr(ksm) + c * 16;
      for (int r = g; r < staged; r += KPB)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(dst0 + r * ROWB), "l"(kb + (int64_t)(j0 + r) * D)
                     : "memory");
      asm volatile("cp.async.commit_group;\n" ::: "memory");
    }
    const int jg = j0 + staged;  // first row of the register-pipelined (global) part of the K pass
    pre_k = jg < j1 && min(j1, jg + KPB * U) <= old_end;
    if (pre_k) load_batch(cur, kb, jg);
    // (Requesting the SECOND K batch here too - 2 x 8 KB per CTA in flight while the prologue folds q / k / v - was
    // measured on B200 at B=64, 860 frames: 1.2321 against 1.2247 ms per decode step; the 72-register cap of 7 CTAs per
    // SM makes the extra live batch spill.)
    pre_v = pre_k && !qf.has_kv && n <= KPB * U;
    if (pre_v) load_batch(nxt, vb, j0);
    // HBM would idle between the K pass and the V pass of the wave, and during the short kernels between two
    // attention launches: queue asynchronous L2 prefetches for what is read next
    if (!LEAN && qf.pf_own_v && n > 0)
      l2_prefetch_range(cv + (((int64_t)b * heads + hd) * c_stride + j0) * D, (int64_t)min(n, old_end - j0) * D * sizeof(T));
    if (!LEAN && qf.pf_next && sp == 0 && !qf.pf_late) prefetch_next();
  }
  pdl_wait();
  pdl_launch();  // (release after the K pass: 1.2205, only at exit: 1.2367 against 1.2194 / 1.2205 ms per decode step)
  tl_dep(tl);
  // pf_late (opt-in, measured slower): prefetch only after the producer GEMM (cross q projection) has finished
  if constexpr (FUSED)
    if (!LEAN && qf.pf_next && sp == 0 && qf.pf_late) prefetch_next();
  if constexpr (!FUSED) plan();
  float qv[E];
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
            if (qf.has_kv) {
              tk[u] = __ldcg(ps + H);
              tv[u] = __ldcg(ps + 2 * H);
            }
          }
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (s0 + u < qf.S) {
            aq += tq[u];
            if (qf.has_kv) {
              ak += tk[u];
              av += tv[u];
            }
          }
      }
      qs[tid] = from_f<T>(aq * qf.scale);
      if (qf.has_kv) {
        const int slot = dyn_value(qf.slot_base, qf.slot_dyn);
        const int64_t o = (((int64_t)b * heads + hd) * c_stride + slot) * D + tid;
        const_cast<T*>(ck)[o] = from_f<T>(ak);
        const_cast<T*>(cv)[o] = from_f<T>(av);
      }
    }
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");  // my share of the staged K rows has landed
    __syncthreads();  // q in shared memory, the new cache row and the staged K rows visible to the whole CTA
    Chunk<T>::unpack(*reinterpret_cast<const uint4*>(reinterpret_cast<const unsigned char*>(qs) + c * 16), qv);
  } else {
    Chunk<T>::load(q + (int64_t)b * H + hd * D + c * E, qv);
  }
  const float neg = finfo_min<T>();
  // ---- pass 1: scores = q . K^T (masked)
  float lmax = -INFINITY;
  if (!pre_k) load_batch(cur, kb, j0 + staged);
  if constexpr (FUSED) {
    for (int r0 = 0; r0 < staged; r0 += KPB) {  // staged rows: all below j1 and below the append slot
      const int r = r0 + g;
      float kf[E];
      Chunk<T>::unpack(*reinterpret_cast<const uint4*>(ksm + r * ROWB + c * 16), kf);
      float dot = 0.f;
#pragma unroll
      for (int e = 0; e < E; ++e) dot = fmaf(qv[e], kf[e], dot);
#pragma unroll
      for (int o = TPK / 2; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
      if (c == 0) {
        const int j = j0 + r;
        bool keep = true;
        if (key_mask && j < mask_len) keep = key_mask[(int64_t)b * mask_len + j] != 0;
        const float s = keep ? dot : neg;
        sc[r] = s;
        lmax = fmaxf(lmax, s);
      }
    }
  }
  for (int base = j0 + staged; base < j1; base += KPB * U) {
    if (base + KPB * U < j1) load_batch(nxt, kb, base + KPB * U);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int j = base + u * KPB + g;
      float dot = 0.f;
      if (j < j1) {
        float kf[E];
        Chunk<T>::unpack(cur[u], kf);
#pragma unroll
        for (int e = 0; e < E; ++e) dot = fmaf(qv[e], kf[e], dot);
      }
#pragma unroll
      for (int o = TPK / 2; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
      if (c == 0 && j < j1) {
        bool keep = true;
        if (key_mask && j < mask_len) keep = key_mask[(int64_t)b * mask_len + j] != 0;
        const float s = keep ? dot : neg;
        sc[j - j0] = s;
        lmax = fmaxf(lmax, s);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) cur[u] = nxt[u];
  }
  // (Requesting these first V rows one step earlier - under the LAST K batch, when nxt is free - was measured on B200 at
  // B=64, 860 frames: 1.2149 / 1.2159 against 1.2115 / 1.2095 ms per decode step, fp8 cache 1.0302 against 1.0257 / 1.0271.)
  if (!pre_v) load_batch(cur, vb, j0);  // first V rows travel while the block reduces max / sum (pre_v: cur = nxt)
  lmax = warp_max(lmax);
  if (lane == 0) red[warp] = lmax;
  __syncthreads();
  const float mx = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
  // ---- pass 2: exp and sum
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
  const float inv = (splits == 1 && n > 0) ? 1.0f / total : 1.0f;
  // ---- pass 3: out = sum_j p_j V_j
  float acc[E];
#pragma unroll
  for (int e = 0; e < E; ++e) acc[e] = 0.f;
  for (int base = j0; base < j1; base += KPB * U) {
    if (base + KPB * U < j1) load_batch(nxt, vb, base + KPB * U);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int j = base + u * KPB + g;
      if (j < j1) {
        float vf[E];
        Chunk<T>::unpack(cur[u], vf);
        const float pj = sc[j - j0] * inv;
#pragma unroll
        for (int e = 0; e < E; ++e) acc[e] = fmaf(pj, vf[e], acc[e]);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) cur[u] = nxt[u];
  }
#pragma unroll
  for (int e = 0; e < E; ++e) ored[g * D + c * E + e] = acc[e];
  __syncthreads();
  if (tid < D) {
    float o = 0.f;
#pragma unroll
    for (int gg = 0; gg < KPB; ++gg) o += ored[gg * D + tid];
    if (splits == 1) {
      out[(int64_t)b * H + hd * D + tid] = from_f<T>(o);
    } else {
      float* pp = part + (((int64_t)b * heads + hd) * splits + sp) * (D + 2);
      pp[2 + tid] = o;
      if (tid == 0) {
        pp[0] = n > 0 ? mx : -INFINITY;
        pp[1] = n > 0 ? total : 0.f;
      }
    }
  }
  if constexpr (FUSED && !LEAN)
    l2_hint_prefetch(qf.tail, (blockIdx.y * gridDim.x + blockIdx.x) * gridDim.z + blockIdx.z,
                     gridDim.x * gridDim.y * gridDim.z);
  tl_end(tl);
}

template <typename T, int D>
__global__ void __launch_bounds__(D)
attn_combine_kernel(const float* __restrict__ part, int heads, int splits, T* __restrict__ out) {
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
  out[((int64_t)b * heads + hd) * D + t] = from_f<T>(o / l);
}

// (Two KV splits per pair when the pairs already fill the GPU - 2048 CTAs in two waves, so that the K / softmax / V
// phases of co-resident CTAs de-synchronise and the tail shortens - were measured on B200 at B=64, 860 frames with the
// combine kernel: 27.7 against 23.4 us per launch, 1.374 against 1.243 ms per decode step.)
int plan_splits(int B, int heads, int c_stride) {
  const int pairs = B * heads;
  if (pairs >= 512) return 1;
  int s = (592 + pairs - 1) / pairs;
  const int cap = c_stride / 64 > 0 ? c_stride / 64 : 1;  // at least ~64 keys per split
  if (s > cap) s = cap;
  if (s > 16) s = 16;
  return s < 1 ? 1 : s;
}

template <typename T>
int launch_decode(const QFuse& qf, const void* q, const void* ck, const void* cv, int B, int heads, int c_stride, int ctx_base,
                  const int* ctx_dyn, const int* key_mask, int mask_len, void* out, float* scratch, cudaStream_t st) {
  constexpr int D = 64;
  constexpr int KPB = NT / (D / Chunk<T>::E);
  const int splits = plan_splits(B, heads, c_stride);
  const int cap = ((c_stride + splits - 1) / splits + 1 + 63) / 64 * 64 + 64;
  // A cp.async.bulk (TMA engine) ring variant of this kernel (4 stages x 8 KB per CTA, one issuing thread) was measured
  // on B200 at B=64, ctx 466: 28.3 us = 4.33 TB/s against 23.2 us = 5.26 TB/s for the software-pipelined ld.global.cg
  // below (80 % of the measured copy peak) and removed.
  PTTS_CHECK_ARG(splits == 1 || scratch, "attention_fast: split-KV needs scratch");
  cudaError_t e;
  // Two one-pass (online softmax) variants were tried and removed: a scalar one (33.1 us per launch at B=64, ctx 472
  // against 24.7 us for this three-pass kernel) and a tensor-core one (mma.sync for q.K^T and P.V, 26.8 us).  Even on an
  // L2-resident cache all three top out at 6..7 TB/s (profiles/scripts/attn_l2.py), i.e. the limit is the 16-byte
  // load path per SM, not the arithmetic and not the pass structure.
  {
    size_t smem = sizeof(float) * ((size_t)cap + 8 + (size_t)KPB * D);
    // K rows staged in shared memory before the dependency wait (fused self-attention only): what fits next to the
    // score buffer with 7 CTAs per SM, in units of one K-pass batch (64 rows = 8 KB).  PTTS_ATTN_STAGE=<rows>.
    // Measured on B200 (B=64, 860 frames): 0 rows 1.686, 64 rows 1.698, 128 rows 1.707 ms per decode step - the rows
    // are L2-resident anyway and the extra shared memory / barrier costs more than the earlier start saves: OFF.
    const int stage_req = PTTS_KNOB("PTTS_ATTN_STAGE", 0);
    int stage_cap = 0;
    if (qf.part && qf.has_kv && sizeof(T) == 2) {
      const int unit = KPB * 4, budget = 31 * 1024;
      stage_cap = smem < (size_t)budget ? (int)((budget - smem) / (D * sizeof(T))) / unit * unit : 0;
      if (stage_cap > stage_req / unit * unit) stage_cap = stage_req / unit * unit;
      smem += (size_t)stage_cap * D * sizeof(T);
    }
    if (smem > 48 * 1024) {
      static size_t configured = 0;
      if (smem > configured) {
        cudaFuncSetAttribute(attn_decode_kernel<T, D, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        cudaFuncSetAttribute(attn_decode_kernel<T, D, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        cudaFuncSetAttribute(attn_decode_kernel<T, D, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        cudaFuncSetAttribute(attn_decode_kernel<T, D, true, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        configured = smem;
      }
    }
    PTTS_CHECK_ARG(smem <= 200 * 1024, "attention_fast: cache too long for the score buffer");
    if (qf.part) {
      const TL tl = tl_next(c_stride > 64 ? 2 : 3);
      const int lean_ok = PTTS_KNOB("PTTS_ATTN_LEAN", 1);  // diagnostic build, 0: always the full instantiation
      const bool lean = lean_ok && stage_cap == 0 && !qf.pf_own_v && !qf.pf_next && !qf.tail.ptr && !tl.p;
      if (lean && !key_mask && splits == 1)
        e = launch_k(attn_decode_kernel<T, D, true, true, true>, dim3(heads, B, splits), dim3(NT), smem, st, qf, (const T*)q,
                     (const T*)ck, (const T*)cv, heads, c_stride, ctx_base, ctx_dyn, key_mask, mask_len, (T*)out, scratch,
                     splits, cap, 0, tl);
      else if (lean)
        e = launch_k(attn_decode_kernel<T, D, true, true>, dim3(heads, B, splits), dim3(NT), smem, st, qf, (const T*)q,
                     (const T*)ck, (const T*)cv, heads, c_stride, ctx_base, ctx_dyn, key_mask, mask_len, (T*)out, scratch,
                     splits, cap, 0, tl);
      else
        e = launch_k(attn_decode_kernel<T, D, true, false>, dim3(heads, B, splits), dim3(NT), smem, st, qf, (const T*)q,
                     (const T*)ck, (const T*)cv, heads, c_stride, ctx_base, ctx_dyn, key_mask, mask_len, (T*)out, scratch,
                     splits, cap, stage_cap, tl);
    }
    else
      e = launch_k(attn_decode_kernel<T, D, false>, dim3(heads, B, splits), dim3(NT), smem, st, qf, (const T*)q,
                   (const T*)ck, (const T*)cv, heads, c_stride, ctx_base, ctx_dyn, key_mask, mask_len, (T*)out, scratch,
                   splits, cap, 0, tl_next(c_stride > 64 ? 2 : 3));
  }
  if (e != cudaSuccess) {
    set_error("attn_decode launch: %s", cudaGetErrorString(e));
    return PTTS_E_LAUNCH;
  }
  PTTS_CHECK_LAUNCH("attn_decode");
  if (splits > 1) {
    e = launch_k(attn_combine_kernel<T, D>, dim3(heads, B), dim3(D), 0, st, (const float*)scratch, heads, splits,
                 (T*)out);
    if (e != cudaSuccess) {
      set_error("attn_combine launch: %s", cudaGetErrorString(e));
      return PTTS_E_LAUNCH;
    }
    PTTS_CHECK_LAUNCH("attn_combine");
  }
  return PTTS_OK;
}

}  // namespace

bool fast_attention_supported(int dtype, int d, int Tq) {
  return (dtype == PTTS_F32 || dtype == PTTS_BF16) && d == 64 && Tq == 1;
}

int64_t attn_scratch_floats(int B, int heads) { return (((int64_t)B * heads + 608) * 66 + 255) / 256 * 256; }

int attention_fast(const void* q, const void* ck, const void* cv, int dtype, int B, int heads, int d, int Tq,
                   int c_stride, int ctx_base, const int* ctx_dyn, int causal, const int* key_mask, int mask_len,
                   void* out, float* scratch, cudaStream_t st) {
  (void)causal;  // a single query at the end of the context sees every cached key
  (void)d;
  (void)Tq;
  QFuse qf = {};
  if (dtype == PTTS_F32)
    return launch_decode<float>(qf, q, ck, cv, B, heads, c_stride, ctx_base, ctx_dyn, key_mask, mask_len, out, scratch,
                                st);
  return launch_decode<__nv_bfloat16>(qf, q, ck, cv, B, heads, c_stride, ctx_base, ctx_dyn, key_mask, mask_len, out,
                                      scratch, st);
}

// bf16 decode attention whose q (and new k / v) come as split-K partial sums of the preceding tcgen05 GEMM
int attention_fast_fused(const float* qkv_part, int S, int64_t part_stride, int ncols, int has_kv, float scale,
                         int slot_base, const int* slot_dyn, void* ck, void* cv, int B, int heads, int c_stride,
                         int ctx_base, const int* ctx_dyn, const int* key_mask, int mask_len, void* out, float* scratch,
                         const void* pf_next, int pf_stride, int pf_rows_base, const int* pf_rows_dyn,
                         cudaStream_t st) {
  QFuse qf = {};
  qf.part = qkv_part; qf.S = S; qf.stride = part_stride; qf.ncols = ncols; qf.has_kv = has_kv; qf.scale = scale;
  qf.slot_base = slot_base; qf.slot_dyn = slot_dyn;
  qf.tail = take_l2_hint();
  // PTTS_L2PF: bit 0 = own V rows (self-attention), bit 1 = next launch's K rows.  Measured on B200 (B=64, 860
  // frames): 3 -> 1.748 ms per decode step, 2 -> 1.708 ms (the own-V prefetch competes with the K pass) - default 2.
  const int pf = PTTS_KNOB("PTTS_L2PF", 2);
  // share of the next launch's K rows to pull (L2 capacity is finite).  Measured (early release on, own-V prefetch
  // off): 0 % 1.677, 50 % 1.673, 100 % 1.690, 160 % 1.711, 200 % (K and V) 1.724 ms per decode step
  const int pf_pct = PTTS_KNOB("PTTS_L2PF_PCT", 50);
  // PTTS_L2PF_LATE=1 queues the next-launch prefetch after the dependency wait, so that the producer GEMM's epilogue
  // stores do not queue behind it.  Measured on B200 (B=64, 860 frames): 1.838 ms per decode step against 1.745 ms
  // with the early prefetch - the ~5 us longer prefetch window is worth more than the GEMM epilogue it delays - OFF.
  qf.pf_late = PTTS_KNOB("PTTS_L2PF_LATE", 0);
  qf.pf_own_v = has_kv && (pf & 1);
  if (pf & 2) {
    qf.pf_next = pf_next; qf.pf_stride = pf_stride; qf.pf_rows_base = pf_rows_base; qf.pf_rows_dyn = pf_rows_dyn;
    qf.pf_pct = pf_pct;
    qf.pf_v_off = (int64_t)B * heads * pf_stride * 64;  // [L][2][B][heads][c_max][64]: V follows K
  }
  return launch_decode<__nv_bfloat16>(qf, nullptr, ck, cv, B, heads, c_stride, ctx_base, ctx_dyn, key_mask, mask_len,
                                      out, scratch, st);
}

}  // namespace ptts
