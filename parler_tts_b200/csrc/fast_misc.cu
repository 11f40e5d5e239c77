// LayerNorm of the residual stream with the split-K reduction of the preceding residual GEMM folded in:
//   h[row] += sum_s part[s][row]   (out_proj / fc2 partial sums, fixed order => deterministic)
//   y[row]  = LN(h[row]) * gamma + beta   (nn.LayerNorm eps 1e-5, /root/reference/parler_tts/modeling_parler_tts.py:446,467,487,959)
// One CTA per row, the row stays in shared memory between the mean / variance / normalise passes (two-pass variance
// like the simple family, so fp32 results agree to rounding).
#include <stdlib.h>

#include "ops.cuh"

namespace ptts {

namespace {

// rowblk (optional): n_blk further slices stored ROW-contiguously, [row][n_blk][H] (what the fused cross-attention block
// writes): the 64 KB of a row arrive as ONE bulk copy (cp.async.bulk, TMA engine) into shared memory instead of 4 096
// 16-byte LSU requests, and are folded from there after the strided slices, in slice order.
// U = strided slice loads a thread keeps in flight per round (2 / 8 / 20, picked from S on the host: predicated-off
// loads and adds of a larger U are still fetched and issued)
// BLK: the row-block path exists; LEAN: no hint, no timeline (one-shot kernel: what is not needed is compiled out, see
// fast_gemm3.cu)
template <typename TY, int U, bool BLK, bool LEAN>
__global__ void __launch_bounds__(256)
ln_reduce_kernel(float* __restrict__ h, const float* __restrict__ part, int S, int64_t part_stride,
                 const float* __restrict__ rowblk_in, int n_blk, const float* __restrict__ g,
                 const float* __restrict__ bta, int H, float eps, int row_mul, int row_add, TY* __restrict__ y,
                 const L2Hint hint, int early, TL tl_in) {
  const float* rowblk = BLK ? rowblk_in : nullptr;
  TL tl = tl_in;
  if constexpr (LEAN) tl.p = nullptr;
  extern __shared__ __align__(128) float buf[];  // [H] row | [n_blk][H] bulk-copied slices
  __shared__ float red[8], red2[8];
  __shared__ __align__(8) uint64_t blk_bar;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t rin = (int64_t)blockIdx.x * row_mul + row_add;
  float* hr = h + rin * H;
  tl_start(tl);
  // Trigger BEFORE waiting: the consumer of a LayerNorm is always a GEMM whose pre-wait section only prefetches
  // immutable weights, so it may become resident (and pull its W slab from HBM) while the kernel before this one is
  // still running; its own griddepcontrol.wait still orders it after this grid.  (early = 0: release after the wait -
  // with TMA slab loads the consumer needs ~1.5 us of lead, and its burst no longer collides with this kernel's loads.)
  if (early) pdl_launch();
  float4 gg4[4], bb4[4];  // gamma / beta are immutable too: fetch them while the producer finishes
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = tid * 4 + i * 1024;
    if (c < H) {
      gg4[i] = __ldg(reinterpret_cast<const float4*>(g + c));
      bb4[i] = __ldg(reinterpret_cast<const float4*>(bta + c));
    }
  }
  if constexpr (!LEAN) l2_hint_prefetch(hint, blockIdx.x, gridDim.x);
  const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&blk_bar);
  if (rowblk && tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(bar), "r"(1) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  if (rowblk) __syncthreads();
  pdl_wait();
  if (!early) pdl_launch();
  tl_dep(tl);
  float* blk = buf + H;
  if (rowblk && tid == 0) {
    const uint32_t bytes = (uint32_t)n_blk * (uint32_t)H * 4u;
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(
                     (uint32_t)__cvta_generic_to_shared(blk)),
                 "l"(rowblk + rin * (int64_t)n_blk * H), "r"(bytes), "r"(bar)
                 : "memory");
  }
  float s = 0.f;
#pragma unroll 1
  for (int c = tid * 4; c < H; c += 256 * 4) {
    float4 v = __ldcg(reinterpret_cast<const float4*>(hr + c));
    // up to U independent loads in flight (one L2 round trip), summed in split order (deterministic)
    for (int i0 = 0; i0 < S; i0 += U) {
      float4 p[U];
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (i0 + u < S) p[u] = __ldcg(reinterpret_cast<const float4*>(part + (i0 + u) * part_stride + rin * H + c));
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (i0 + u < S) { v.x += p[u].x; v.y += p[u].y; v.z += p[u].z; v.w += p[u].w; }
    }
    if (rowblk) {
      uint32_t done = 0;
      for (uint32_t spin = 0; !done; ++spin) {
        asm volatile(
            "{\n\t.reg .pred P1;\n\tmbarrier.test_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.b32 %0, 1, 0, P1;\n\t}\n"
            : "=r"(done)
            : "r"(bar), "r"(0)
            : "memory");
        if (spin > 200000000u) __trap();
      }
      // eight shared-memory loads in flight before the first add (rolled, every add waited for its own load); the adds
      // keep the slice order
#pragma unroll 1
      for (int k0 = 0; k0 < n_blk; k0 += 8) {
        float4 t[8];
#pragma unroll
        for (int u = 0; u < 8; ++u)
          if (k0 + u < n_blk) t[u] = *reinterpret_cast<const float4*>(blk + (size_t)(k0 + u) * H + c);
#pragma unroll
        for (int u = 0; u < 8; ++u)
          if (k0 + u < n_blk) { v.x += t[u].x; v.y += t[u].y; v.z += t[u].z; v.w += t[u].w; }
      }
    }
    if (S > 0 || rowblk) *reinterpret_cast<float4*>(hr + c) = v;
    *reinterpret_cast<float4*>(buf + c) = v;
    s += (v.x + v.y) + (v.z + v.w);
  }
  // block reductions: one barrier each - every thread adds the 8 warp sums itself, in fixed order (deterministic)
  s = warp_sum(s);
  if (lane == 0) red[warp] = s;
  __syncthreads();
  const float mean = (((red[0] + red[1]) + (red[2] + red[3])) + ((red[4] + red[5]) + (red[6] + red[7]))) / (float)H;
  float qq = 0.f;
#pragma unroll 1
  for (int c = tid * 4; c < H; c += 256 * 4) {
    const float4 v = *reinterpret_cast<const float4*>(buf + c);
    const float a = v.x - mean, b = v.y - mean, cc = v.z - mean, d = v.w - mean;
    qq += (a * a + b * b) + (cc * cc + d * d);
  }
  qq = warp_sum(qq);
  if (lane == 0) red2[warp] = qq;
  __syncthreads();
  const float rstd =
      rsqrtf((((red2[0] + red2[1]) + (red2[2] + red2[3])) + ((red2[4] + red2[5]) + (red2[6] + red2[7]))) / (float)H + eps);
  TY* yr = y + (int64_t)blockIdx.x * H;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = tid * 4 + i * 1024;
    if (c >= H) break;
    const float4 v = *reinterpret_cast<const float4*>(buf + c);
    const float4 gg = gg4[i], bb = bb4[i];
    const float o0 = (v.x - mean) * rstd * gg.x + bb.x, o1 = (v.y - mean) * rstd * gg.y + bb.y;
    const float o2 = (v.z - mean) * rstd * gg.z + bb.z, o3 = (v.w - mean) * rstd * gg.w + bb.w;
    if constexpr (sizeof(TY) == 4) {
      *reinterpret_cast<float4*>(yr + c) = make_float4(o0, o1, o2, o3);
    } else {
      __nv_bfloat162 lo = __floats2bfloat162_rn(o0, o1), hi = __floats2bfloat162_rn(o2, o3);
      uint2 u;
      u.x = *reinterpret_cast<uint32_t*>(&lo);
      u.y = *reinterpret_cast<uint32_t*>(&hi);
      *reinterpret_cast<uint2*>(yr + c) = u;
    }
  }
  tl_end(tl);
}

}  // namespace

bool fast_layernorm_supported(int H) { return H % 4 == 0 && H <= 4096; }  // gamma / beta live in registers

// rows output rows; output row r reads input row r*row_mul + row_add (of h and of every partial)
int layernorm_fast(float* h, const float* part, int S, int64_t part_stride, const float* g, const float* b, int rows,
                   int H, float eps, int row_mul, int row_add, void* y, int ydtype, cudaStream_t st,
                   const float* rowblk, int n_blk) {
  if (rows == 0) return PTTS_OK;
  const size_t smem = sizeof(float) * (size_t)H * (size_t)(1 + (rowblk ? n_blk : 0));
  PTTS_CHECK_ARG(smem <= 200 * 1024, "layernorm_fast: row block too large for shared memory");
  if (smem > 48 * 1024) {
    static size_t configured = 0;
    if (smem > configured) {
#define PTTS_LN_ATTR(TY, UU)                                                                                          \
  cudaFuncSetAttribute(ln_reduce_kernel<TY, UU, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);  \
  cudaFuncSetAttribute(ln_reduce_kernel<TY, UU, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
  cudaFuncSetAttribute(ln_reduce_kernel<TY, UU, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
  cudaFuncSetAttribute(ln_reduce_kernel<TY, UU, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)
      PTTS_LN_ATTR(float, 2);
      PTTS_LN_ATTR(float, 8);
      PTTS_LN_ATTR(float, 20);
      PTTS_LN_ATTR(__nv_bfloat16, 2);
      PTTS_LN_ATTR(__nv_bfloat16, 8);
      PTTS_LN_ATTR(__nv_bfloat16, 20);
#undef PTTS_LN_ATTR
      configured = smem;
    }
  }
  const L2Hint hint = take_l2_hint();
  // measured on B200 (B=64, 860 frames, TMA GEMMs): early 1.496, late 1.481 ms per decode step; together with the late
  // cross-attention block 1.427 -> late (diagnostic build, PTTS_LN_EARLY=1: before the wait)
    const int early = PTTS_KNOB("PTTS_LN_EARLY", 0);  // re-measured on the round-2 kernels: 1.2338 against 1.2206 ms
  cudaError_t e;
  const TL tl = tl_next(1);
  const bool lean = !hint.ptr && !tl.p;
#define PTTS_LN_LAUNCH2(TY, UU, BB, LL)                                                                               \
  e = launch_k(ln_reduce_kernel<TY, UU, BB, LL>, dim3(rows), dim3(256), smem, st, h, part, S, part_stride, rowblk, n_blk, \
               g, b, H, eps, row_mul, row_add, (TY*)y, hint, early, tl)
#define PTTS_LN_LAUNCH(TY, UU)                                      \
  do {                                                              \
    if (rowblk) {                                                   \
      if (lean) PTTS_LN_LAUNCH2(TY, UU, true, true);                \
      else PTTS_LN_LAUNCH2(TY, UU, true, false);                    \
    } else {                                                        \
      if (lean) PTTS_LN_LAUNCH2(TY, UU, false, true);               \
      else PTTS_LN_LAUNCH2(TY, UU, false, false);                   \
    }                                                               \
  } while (0)
  if (ydtype == PTTS_F32) {
    if (S <= 2) PTTS_LN_LAUNCH(float, 2);
    else if (S <= 8) PTTS_LN_LAUNCH(float, 8);
    else PTTS_LN_LAUNCH(float, 20);
  } else {
    if (S <= 2) PTTS_LN_LAUNCH(__nv_bfloat16, 2);
    else if (S <= 8) PTTS_LN_LAUNCH(__nv_bfloat16, 8);
    else PTTS_LN_LAUNCH(__nv_bfloat16, 20);
  }
#undef PTTS_LN_LAUNCH2
#undef PTTS_LN_LAUNCH
  if (e != cudaSuccess) {
    set_error("layernorm_fast launch: %s", cudaGetErrorString(e));
    return PTTS_E_LAUNCH;
  }
  PTTS_CHECK_LAUNCH("layernorm_fast");
  return PTTS_OK;
}

// y[r][c] = act(sum_s part[s][r][c]) as bf16: finishes the split-K fc1 of the decode step (partial sums are added in
// slice order, deterministic).  One float4 column group per thread; all S loads of a thread are in flight together.
// ACT >= 0: the activation is a compile-time constant (the GELU instantiation is a third of the generic code)
template <int ACT, bool LEAN>
__global__ void __launch_bounds__(256)
partial_act_kernel(const float* __restrict__ part, int S, int64_t part_stride, int64_t n4, int act_in,
                   __nv_bfloat16* __restrict__ y, const L2Hint hint, int early, TL tl_in) {
  TL tl = tl_in;
  if constexpr (LEAN) tl.p = nullptr;
  tl_start(tl);
  if (early) pdl_launch();  // before the wait: the consumer (fc2 GEMM) only prefetches immutable weights ahead of its own wait
  if constexpr (!LEAN) l2_hint_prefetch(hint, blockIdx.x, gridDim.x);
  pdl_wait();
  if (!early) pdl_launch();
  tl_dep(tl);
  const int act = ACT >= 0 ? ACT : act_in;
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i < n4) {
    const float4* p0 = reinterpret_cast<const float4*>(part) + i;
    const int64_t st4 = part_stride / 4;
    float4 v = __ldcg(p0);
    for (int s0 = 1; s0 < S; s0 += 4) {
      float4 t[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (s0 + u < S) t[u] = __ldcg(p0 + (s0 + u) * st4);
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (s0 + u < S) { v.x += t[u].x; v.y += t[u].y; v.z += t[u].z; v.w += t[u].w; }
    }
    __nv_bfloat162 lo = __floats2bfloat162_rn(apply_act(v.x, act), apply_act(v.y, act));
    __nv_bfloat162 hi = __floats2bfloat162_rn(apply_act(v.z, act), apply_act(v.w, act));
    uint2 u;
    u.x = *reinterpret_cast<uint32_t*>(&lo);
    u.y = *reinterpret_cast<uint32_t*>(&hi);
    reinterpret_cast<uint2*>(y)[i] = u;
  }
  tl_end(tl);
}

int partial_act_fast(const float* part, int S, int64_t part_stride, int rows, int N, int act, void* y,
                     cudaStream_t st) {
  PTTS_CHECK_ARG(part && y && S >= 1 && N % 4 == 0 && part_stride % 4 == 0, "partial_act_fast: bad arguments");
  const int64_t n4 = (int64_t)rows * N / 4;
  if (n4 == 0) return PTTS_OK;
  const int early = PTTS_KNOB("PTTS_ACT_EARLY", 1);  // measured: early 1.496, late 1.522 ms per decode step -> early
  const L2Hint hint = take_l2_hint();
  const TL tl = tl_next(4);
  const dim3 grid((unsigned)((n4 + 255) / 256));
  cudaError_t e;
  if (act == PTTS_ACT_GELU && !hint.ptr && !tl.p)
    e = launch_k(partial_act_kernel<PTTS_ACT_GELU, true>, grid, dim3(256), 0, st, part, S, part_stride, n4, act,
                 (__nv_bfloat16*)y, hint, early, tl);
  else
    e = launch_k(partial_act_kernel<-1, false>, grid, dim3(256), 0, st, part, S, part_stride, n4, act,
                 (__nv_bfloat16*)y, hint, early, tl);
  if (e != cudaSuccess) {
    set_error("partial_act_fast launch: %s", cudaGetErrorString(e));
    return PTTS_E_LAUNCH;
  }
  PTTS_CHECK_LAUNCH("partial_act_fast");
  return PTTS_OK;
}


}  // namespace ptts
