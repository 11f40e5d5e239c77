// DAC decoder convolutions, warp-specialised persistent tcgen05 kernel (the default conv path of the fast family).
//   D[t][n] = sum_tap sum_ci snake(x)[t + off(tap)][ci] * W[tap][n][ci]        one tile = 256 time rows x BN columns
//
// Why a third kernel (measured on B200, profiles/scripts/micro/umma_rate.cu and dac_layers.py): a 128 x 128 x 16 bf16
// tcgen05.mma takes 74.5 cycles whatever the operand layout (no swizzle / 128-byte swizzle / shifted start address),
// so the ~2500 cycles per (tap, 32-channel chunk) of fast_dac2.cu are not the tensor core: they are the per-iteration
// cp.async + __syncthreads hand-shake of all 256 threads, the activation staging that stalls the issuing thread, and
// the weight stream (16 KB per 6 MMAs per SM = ~10 TB/s of L2 reads at full MMA rate, above what L2 delivers).
//
// This kernel removes all three:
//  * roles: warp 0 streams weight stages with ONE cp.async.bulk each (weights are pre-tiled on the host into the
//    exact shared-memory image of a stage, ptts_dac_tile_n / dac.py), warp 1 issues the MMAs, warps 2..9 stage the
//    activations (fp32 -> Snake -> hi/mid bf16 panels) one chunk ahead, warps 10..17 (plus the stager warps when the
//    input arrives as planes and they have nothing to stage) run the epilogue of tile i
//    while the tensor core already works on tile i+1 (double-buffered TMEM accumulators when 4 x BN <= 512
//    columns); their residual rows are requested before the accumulator is ready.  All hand-offs are mbarriers
//    (tcgen05.commit on the consumer side); there is no __syncthreads in the steady state.
//  * "planes": a conv can WRITE its output already Snake-activated and split into hi/mid bf16 panels
//    ([B][2][C/8][T + pad][8], same bytes as fp32) and the next conv READS such planes with plain bulk copies - the
//    Snake of a k=1 conv or of a transposed conv (whose 6..48 column tiles would each redo it) then costs nothing:
//    it rides in the epilogue warps of the producing k=7 conv, which are idle most of the time.
//  * a tile is 256 rows (two 128-row accumulators share every weight stage): half the L2 weight traffic per FLOP.
//  * persistent CTAs (one per SM) loop over tiles, so TMEM allocation, barrier set-up and pipeline fill are paid once.
// Operand layout and the three-product fp32 emulation (mid*hi + hi*mid + hi*hi) are those of fast_dac2.cu.
#include <stdlib.h>

#include "gemm_common.cuh"

namespace ptts {

using namespace gemm;

namespace {

constexpr int TMW = 256, CIK3 = 32, NTHR = 576, NSTG = 256, WST3 = 4, MAX_SPAN3 = 54, MAXXB = 4;
constexpr int XU3 = ((TMW + MAX_SPAN3) * 8 + NSTG - 1) / NSTG;
constexpr int PAD_LO = PTTS_DAC_PLANE_PAD_LO, PAD_HI = PTTS_DAC_PLANE_PAD_HI;

using ConvParams3 = DacConv;  // ops.cuh

__device__ __forceinline__ float snake3(float x, float a, float inv) {  // see fast_dac2.cu: range-reduced SFU sine
  const float ax = a * x;
  const float k = rintf(ax * 0.15915494309189535f);
  float r = fmaf(k, -6.2831854820251465f, ax);
  r = fmaf(k, 1.7484556000744883e-7f, r);
  const float s = __sinf(r);
  return fmaf(inv * s, s, x);
}
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<const uint32_t*>(&h);
}
__device__ __forceinline__ uint64_t desc_np(uint32_t saddr, uint32_t lbo_bytes) {  // K-major, no swizzle, SBO = 128 B
  return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) | (8ull << 32) | (1ull << 46);
}
__device__ __forceinline__ void mma3(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void commit3(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void bar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void bar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void bar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory");
}
// bounded wait: a protocol bug must trap (the launch fails loudly), never hang the GPU
__device__ __forceinline__ void bar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  long long t0 = 0;
  for (uint32_t spin = 0; !done; ++spin) {
    asm volatile(
        "{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.b32 %0, 1, 0, P1;\n\t}\n"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if ((spin & 1023u) == 1023u) {  // ~2 s of SM clock without progress
      if (t0 == 0) t0 = clock64();
      else if (clock64() - t0 > 4000000000ll) __trap();
    }
  }
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(dst),
      "l"(src), "r"(bytes), "r"(bar)
      : "memory");
}
__device__ __forceinline__ void tmem_ld16b(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, "
      "[%16];\n"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}

// 32-byte accesses (LDG.256 / STG.256, sm_100): a lane moves a whole 32-byte sector of its row per instruction - the
// fp32 rows of the epilogue are 64 bytes per (time step, 16-channel group), i.e. two instructions instead of four
#ifndef PTTS_DAC_WIDE
#define PTTS_DAC_WIDE 1
#endif
__device__ __forceinline__ void ld_f8(const float* p, float4& a, float4& b) {
  asm volatile("ld.global.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];\n"
               : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
               : "l"(p));
}
__device__ __forceinline__ void st_f8(float* p, const float4& a, const float4& b) {
  asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};\n" ::"l"(p), "f"(a.x), "f"(a.y), "f"(a.z), "f"(a.w),
               "f"(b.x), "f"(b.y), "f"(b.z), "f"(b.w)
               : "memory");
}

struct Tile {
  int b, m0, n0, nt;
};

template <int BN, bool DBL>
__global__ void __launch_bounds__(NTHR, 1) dac_conv_ws_kernel(const ConvParams3 p, int MT, int NTn, int ntiles, int NXB) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ __align__(8) uint64_t bars[2 * WST3 + 2 * MAXXB + 4];
  __shared__ uint32_t tmem_slot;
  constexpr uint32_t IDESC = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  constexpr int WPANEL = BN * 16;     // one 8-channel weight panel
  constexpr int WPLANE = 4 * WPANEL;  // hi or mid plane of a (tap, chunk) stage
  constexpr int WSTAGE = 2 * WPLANE;
  const int xrows = TMW + p.span;
  const int XPANEL = xrows * 16;
  const int XPLANE = 4 * XPANEL;
  const int XBUF = 2 * XPLANE;
  const uint32_t sx = smem_u32(smem);
  const uint32_t sw = sx + NXB * XBUF;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nchunks = p.Cin / CIK3;
  const int nit = nchunks * p.taps;
  const bool in_planes = p.x_planes != nullptr;
  const uint32_t w_full = smem_u32(&bars[0]), w_empty = smem_u32(&bars[WST3]);
  const uint32_t x_full = smem_u32(&bars[2 * WST3]), x_empty = smem_u32(&bars[2 * WST3 + MAXXB]);
  const uint32_t acc_full = smem_u32(&bars[2 * WST3 + 2 * MAXXB]), acc_empty = smem_u32(&bars[2 * WST3 + 2 * MAXXB + 2]);

  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(&tmem_slot)),
                 "r"(512)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
  }
  if (tid == 32) {
    for (int i = 0; i < WST3; ++i) {
      bar_init(w_full + 8 * i, 1);
      bar_init(w_empty + 8 * i, 1);
    }
    for (int i = 0; i < MAXXB; ++i) {
      bar_init(x_full + 8 * i, in_planes ? 1 : NSTG);  // planes: the producer's expect_tx; fp32: every stager thread
      bar_init(x_empty + 8 * i, 1);
    }
    for (int i = 0; i < 2; ++i) {
      bar_init(acc_full + 8 * i, 1);
      bar_init(acc_empty + 8 * i, in_planes ? 512 : 256);  // every epilogue thread
    }
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  // Snake parameters of the consumer (planes output): alpha and 1 / (alpha + 1e-9) per output channel, once per CTA
  float* s_al = reinterpret_cast<float*>(smem + (size_t)NXB * XBUF + (size_t)WST3 * WSTAGE);
  float* s_ia = s_al + p.Cout;
  if (p.y_planes)
    for (int i = tid; i < p.Cout; i += NTHR) {
      const float a = p.out_alpha ? __ldg(p.out_alpha + i) : 0.f;
      s_al[i] = a;
      s_ia[i] = 1.0f / (a + 1e-9f);
    }
  asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
  const uint32_t tmem_d = tmem_slot;

  auto tile_of = [&](int tile) {
    Tile t;
    t.nt = tile % NTn;
    const int r = tile / NTn;
    t.m0 = (r % MT) * TMW;
    t.b = r / MT;
    t.n0 = t.nt * BN;
    return t;
  };

  // ---------------------------------------------------------------------- epilogue: a warp owns TMEM lanes
  // 32 (warp % 4) .. +31, i.e. its 32 time rows of both 128-row accumulators; the EW warps of a lane quarter split
  // the 2 * BN / 16 column groups
  auto run_epilogue = [&](const int e_idx, const int EW) {
    const int q = warp & 3;
    constexpr int NG = BN / 16;  // column groups per accumulator
    const int Tpo = p.Tout + PAD_LO + PAD_HI;
    uint32_t tt = 0;
    const int g_first = e_idx, g_step = EW;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++tt) {
      const Tile t = tile_of(tile);
      const uint32_t a = DBL ? (tt & 1) : 0, ak = DBL ? (tt >> 1) : tt;
      const float* resb = p.residual ? p.residual + (int64_t)t.b * p.Tout * p.Cout : nullptr;
      float* yb = p.y ? p.y + (int64_t)t.b * p.Tout * p.Cout : nullptr;
      // group gi = (sub, cc): output position of this thread's row and its first channel
      auto locate = [&](int gi, int& to, int& co, bool& ok) {
        const int sub = gi / NG, cc = (gi % NG) * 16;
        const int row = t.m0 + sub * 128 + q * 32 + lane;
        const int n = t.n0 + cc;
        const int ph = n / p.Cout;  // a 16-column group never straddles a phase (Cout % 16 == 0 checked on the host)
        co = n - ph * p.Cout;
        to = row * p.out_stride + ph - p.out_pad;
        ok = row < p.Mrows && n < p.N && to >= 0 && to < p.Tout;
      };
      float4 res[4];
      auto fetch_res = [&](int gi) {
        int to, co;
        bool ok;
        locate(gi, to, co, ok);
#pragma unroll
        for (int i = 0; i < 4; ++i) res[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (resb && ok) {
          const float* rp = resb + (int64_t)to * p.Cout + co;
          if (PTTS_DAC_WIDE) {
            ld_f8(rp, res[0], res[1]);
            ld_f8(rp + 8, res[2], res[3]);
          } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) res[i] = *reinterpret_cast<const float4*>(rp + 4 * i);
          }
        }
      };
      fetch_res(g_first);  // the residual rows do not depend on the accumulator: request them before waiting for it
      bar_wait(acc_full + 8 * a, ak & 1);
      asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
#pragma unroll 1
      for (int gi = g_first; gi < 2 * NG; gi += g_step) {
        const int sub = gi / NG, cc = (gi % NG) * 16;
        uint32_t v[16];
        tmem_ld16b(tmem_d + ((uint32_t)(q * 32) << 16) + a * 2 * BN + sub * BN + cc, v);
        float4 r[4] = {res[0], res[1], res[2], res[3]};
        if (gi + g_step < 2 * NG) fetch_res(gi + g_step);
        asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
        int to, co;
        bool ok;
        locate(gi, to, co, ok);
        if (ok) {
          float o[16];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            float4 bb = make_float4(0.f, 0.f, 0.f, 0.f);
            if (p.bias) bb = __ldg(reinterpret_cast<const float4*>(p.bias + co + 4 * i));
            o[4 * i + 0] = __uint_as_float(v[4 * i + 0]) + bb.x + r[i].x;
            o[4 * i + 1] = __uint_as_float(v[4 * i + 1]) + bb.y + r[i].y;
            o[4 * i + 2] = __uint_as_float(v[4 * i + 2]) + bb.z + r[i].z;
            o[4 * i + 3] = __uint_as_float(v[4 * i + 3]) + bb.w + r[i].w;
          }
          if (yb) {
            float* yp = yb + (int64_t)to * p.Cout + co;
            if (PTTS_DAC_WIDE) {
              st_f8(yp, make_float4(o[0], o[1], o[2], o[3]), make_float4(o[4], o[5], o[6], o[7]));
              st_f8(yp + 8, make_float4(o[8], o[9], o[10], o[11]), make_float4(o[12], o[13], o[14], o[15]));
            } else {
#pragma unroll
              for (int i = 0; i < 4; ++i)
                *reinterpret_cast<float4*>(yp + 4 * i) = make_float4(o[4 * i], o[4 * i + 1], o[4 * i + 2], o[4 * i + 3]);
            }
          }
          if (p.y_planes) {
            // the consumer's Snake and hi/mid split, written as two 8-channel panels per plane: lanes are consecutive
            // time rows, so every store instruction of the warp covers 512 contiguous bytes
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              uint4 hi, mid;
              uint32_t* hp = reinterpret_cast<uint32_t*>(&hi);
              uint32_t* mp = reinterpret_cast<uint32_t*>(&mid);
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                float v0 = o[8 * h + 2 * e], v1 = o[8 * h + 2 * e + 1];
                if (p.out_alpha) {
                  const float2 aa = *reinterpret_cast<const float2*>(s_al + co + 8 * h + 2 * e);
                  const float2 ii = *reinterpret_cast<const float2*>(s_ia + co + 8 * h + 2 * e);
                  v0 = snake3(v0, aa.x, ii.x);
                  v1 = snake3(v1, aa.y, ii.y);
                }
                const __nv_bfloat162 hh = __floats2bfloat162_rn(v0, v1);
                const float2 ff = __bfloat1622float2(hh);
                hp[e] = *reinterpret_cast<const uint32_t*>(&hh);
                mp[e] = pack2(v0 - ff.x, v1 - ff.y);
              }
              const int64_t panel = (co >> 3) + h;
              __nv_bfloat16* dst_hi = p.y_planes + ((((int64_t)t.b * 2 + 0) * (p.Cout / 8) + panel) * Tpo + PAD_LO + to) * 8;
              __nv_bfloat16* dst_mid = p.y_planes + ((((int64_t)t.b * 2 + 1) * (p.Cout / 8) + panel) * Tpo + PAD_LO + to) * 8;
              *reinterpret_cast<uint4*>(dst_hi) = hi;
              *reinterpret_cast<uint4*>(dst_mid) = mid;
            }
          }
        }
      }
      asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
      bar_arrive(acc_empty + 8 * a);
    }
  };

  if (warp == 0) {
    // ------------------------------------------------------------------ producer: bulk copies of weight stages and,
    // for "planes" input, of the activation panels (8 copies of xrows x 16 B per 32-channel chunk)
    if (elect_one()) {
      uint32_t k = 0, g = 0;
      const int Tp = p.T + PAD_LO + PAD_HI;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const Tile t = tile_of(tile);
        const unsigned char* src = reinterpret_cast<const unsigned char*>(p.w) + (size_t)t.nt * nit * WSTAGE;
        for (int c = 0; c < nchunks; ++c, ++g) {
          if (in_planes) {
            const uint32_t buf = g % NXB;
            bar_wait(x_empty + 8 * buf, ((g / NXB) & 1) ^ 1);
            bar_expect_tx(x_full + 8 * buf, 8 * XPANEL);
#pragma unroll
            for (int pl = 0; pl < 2; ++pl)
#pragma unroll
              for (int kc = 0; kc < 4; ++kc) {
                const __nv_bfloat16* xs =
                    p.x_planes + ((((int64_t)t.b * 2 + pl) * (p.Cin / 8) + c * 4 + kc) * Tp + PAD_LO + t.m0 + p.lo) * 8;
                bulk_g2s(sx + buf * XBUF + pl * XPLANE + kc * XPANEL, xs, XPANEL, x_full + 8 * buf);
              }
          }
          for (int tap = 0; tap < p.taps; ++tap, ++k) {
            const uint32_t s = k % WST3;
            bar_wait(w_empty + 8 * s, ((k / WST3) & 1) ^ 1);
            bar_expect_tx(w_full + 8 * s, WSTAGE);
            bulk_g2s(sw + s * WSTAGE, src + (size_t)(c * p.taps + tap) * WSTAGE, WSTAGE, w_full + 8 * s);
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (elect_one()) {  // single-lane issue from uniform registers (common.cuh)
      uint32_t k = 0, g = 0, tt = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++tt) {
        const uint32_t a = DBL ? (tt & 1) : 0, ak = DBL ? (tt >> 1) : tt;
        bar_wait(acc_empty + 8 * a, (ak & 1) ^ 1);  // the epilogue has drained this accumulator pair
        asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
        const uint32_t d0 = tmem_d + a * 2 * BN, d1 = d0 + BN;
        for (int c = 0; c < nchunks; ++c, ++g) {
          const uint32_t buf = g % NXB;
          bar_wait(x_full + 8 * buf, (g / NXB) & 1);
          asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
          const uint32_t xh = sx + buf * XBUF, xm = xh + XPLANE;
          for (int tap = 0; tap < p.taps; ++tap, ++k) {
            const uint32_t s = k % WST3;
            bar_wait(w_full + 8 * s, (k / WST3) & 1);
            asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
            const uint32_t wh = sw + s * WSTAGE, wm = wh + WPLANE;
            const uint32_t shift = (uint32_t)((tap - p.tap_center) * p.tap_step - p.lo) * 16u;
#pragma unroll
            for (int ks = 0; ks < 2; ++ks) {
              const uint64_t bh = desc_np(wh + 2 * ks * WPANEL, WPANEL);
              const uint64_t bm = desc_np(wm + 2 * ks * WPANEL, WPANEL);
#pragma unroll
              for (int sub = 0; sub < 2; ++sub) {
                const uint32_t ro = shift + sub * 128 * 16;
                const uint64_t ah = desc_np(xh + 2 * ks * XPANEL + ro, XPANEL);
                const uint64_t am = desc_np(xm + 2 * ks * XPANEL + ro, XPANEL);
                const uint32_t d = sub ? d1 : d0;
                mma3(d, am, bh, IDESC, (c | tap | ks) ? 1u : 0u);
                mma3(d, ah, bm, IDESC, 1u);
                mma3(d, ah, bh, IDESC, 1u);
              }
            }
            commit3(w_empty + 8 * s);  // the stage may be refilled once these MMAs have read it
          }
          commit3(x_empty + 8 * buf);
        }
        commit3(acc_full + 8 * a);
      }
    }
    __syncwarp();
  } else if (warp < 10) {
    // ------------------------------------------------------------------ activation stagers (fp32 input); with planes
    // input there is nothing to stage and these warps help with the epilogue
    if (in_planes) {
      run_epilogue(2 + ((warp - 2) >> 2), 4);
    } else {
      const int st = tid - 64;
      const int c4 = st & 7;  // a thread always handles the same 4 channels of a chunk
      float4 pre[XU3];
      auto prefetch_x = [&](const Tile& t, int c) {
        const float* xb = p.x + (int64_t)t.b * p.T * p.Cin + c * CIK3 + c4 * 4;
#pragma unroll
        for (int j = 0; j < XU3; ++j) {
          const int u = st + NSTG * j;
          float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
          if (u < xrows * 8) {
            const int tt = t.m0 + p.lo + (u >> 3);
            if (tt >= 0 && tt < p.T) v = __ldg(reinterpret_cast<const float4*>(xb + (int64_t)tt * p.Cin));
          }
          pre[j] = v;
        }
      };
      auto store_x = [&](int c, uint32_t buf) {
        unsigned char* base = smem + (size_t)buf * XBUF;
        float4 al = make_float4(1.f, 1.f, 1.f, 1.f), ia = al;
        if (p.alpha) {
          al = __ldg(reinterpret_cast<const float4*>(p.alpha + c * CIK3 + c4 * 4));
          ia = make_float4(1.0f / (al.x + 1e-9f), 1.0f / (al.y + 1e-9f), 1.0f / (al.z + 1e-9f), 1.0f / (al.w + 1e-9f));
        }
#pragma unroll
        for (int j = 0; j < XU3; ++j) {
          const int u = st + NSTG * j;
          if (u < xrows * 8) {
            const int row = u >> 3;
            float4 v = pre[j];
            if (p.alpha) {
              v.x = snake3(v.x, al.x, ia.x);
              v.y = snake3(v.y, al.y, ia.y);
              v.z = snake3(v.z, al.z, ia.z);
              v.w = snake3(v.w, al.w, ia.w);
            }
            const __nv_bfloat162 h0 = __floats2bfloat162_rn(v.x, v.y), h1 = __floats2bfloat162_rn(v.z, v.w);
            const float2 f0 = __bfloat1622float2(h0), f1 = __bfloat1622float2(h1);
            uint2 hi, mid;
            hi.x = *reinterpret_cast<const uint32_t*>(&h0);
            hi.y = *reinterpret_cast<const uint32_t*>(&h1);
            mid.x = pack2(v.x - f0.x, v.y - f0.y);
            mid.y = pack2(v.z - f1.x, v.w - f1.y);
            const int off = (c4 >> 1) * XPANEL + row * 16 + (c4 & 1) * 8;
            *reinterpret_cast<uint2*>(base + off) = hi;
            *reinterpret_cast<uint2*>(base + XPLANE + off) = mid;
          }
        }
      };
      uint32_t g = 0;
      if ((int)blockIdx.x < ntiles) prefetch_x(tile_of(blockIdx.x), 0);
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const Tile t = tile_of(tile);
        for (int c = 0; c < nchunks; ++c, ++g) {
          const uint32_t buf = g % NXB;
          bar_wait(x_empty + 8 * buf, ((g / NXB) & 1) ^ 1);  // the MMAs that read this buffer NXB chunks ago are done
          store_x(c, buf);
          asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");  // generic-proxy writes -> async proxy (UMMA)
          bar_arrive(x_full + 8 * buf);
          if (c + 1 < nchunks)
            prefetch_x(t, c + 1);
          else if (tile + (int)gridDim.x < ntiles)
            prefetch_x(tile_of(tile + gridDim.x), 0);
        }
      }
    }
  } else {
    run_epilogue((warp - 10) >> 2, in_planes ? 4 : 2);
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
  __syncthreads();
  if (warp == 0)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_d), "r"(512) : "memory");
}

int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

// CTAs a conv launch may occupy (0 = one per SM): the streamed chunk decode runs next to the token loop and leaves
// the other SMs to the loop's kernels (ptts_dac_set_cta_limit)
int g_cta_limit = 0;

template <int BN, bool DBL>
int launch_ws(const ConvParams3& p, int B, cudaStream_t st) {
  const size_t xbuf = (size_t)2 * 4 * (TMW + p.span) * 16, wring = (size_t)WST3 * 2 * 4 * BN * 16;
  // activation buffers: 2 when the stager warps fill them, up to 4 (whatever fits) when they arrive by bulk copy -
  // the planes path is HBM-bound and wants more bytes in flight
  int nxb = 2;
  if (p.x_planes)
    while (nxb < MAXXB && (nxb + 1) * xbuf + wring <= 188 * 1024) ++nxb;
  const size_t smem = nxb * xbuf + wring + (p.y_planes ? (size_t)2 * p.Cout * sizeof(float) : 0);
  static size_t configured = 0;
  if (smem > configured) {
    cudaFuncSetAttribute(dac_conv_ws_kernel<BN, DBL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    configured = smem;
  }
  const int MT = (p.Mrows + TMW - 1) / TMW, NTn = (p.N + BN - 1) / BN;
  const int ntiles = MT * NTn * B;
  int grid = ntiles < sm_count() ? ntiles : sm_count();
  if (g_cta_limit > 0 && grid > g_cta_limit) grid = g_cta_limit;  // persistent CTAs: any grid walks all tiles
  dac_conv_ws_kernel<BN, DBL><<<grid, NTHR, smem, st>>>(p, MT, NTn, ntiles, nxb);
  PTTS_CHECK_LAUNCH("dac_conv_ws");
  return PTTS_OK;
}

}  // namespace

int dac_set_cta_limit(int n) {
  const int prev = g_cta_limit;
  g_cta_limit = n > 0 ? n : 0;
  return prev;
}

// column-tile width of the pre-tiled weight layout for a conv with N output columns (ptts_dac_tile_n)
int dac_tile_n(int N) {
  if (N % 128 == 0) return 128;
  if (N % 192 == 0) return 192;
  if (N % 96 == 0) return 96;
  return 128;
}

// p.w = weights pre-tiled as [n_tile][chunk][tap][plane hi|mid][4 panels][BN][8] bf16, BN = dac_tile_n(N)
int dac_conv_ws(const DacConv& p, int B, cudaStream_t st) {
  if (p.Cout % 16 != 0 || p.span > MAX_SPAN3 || p.Cin % CIK3 != 0) return PTTS_E_UNSUPPORTED;
  if (p.y_planes && p.out_stride != 1) return PTTS_E_UNSUPPORTED;
  switch (dac_tile_n(p.N)) {
    case 192: return launch_ws<192, false>(p, B, st);
    case 96: return launch_ws<96, true>(p, B, st);
    default: return launch_ws<128, true>(p, B, st);
  }
}

}  // namespace ptts
