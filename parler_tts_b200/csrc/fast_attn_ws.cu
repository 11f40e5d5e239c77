// Warp-specialised, TMA-staged flash-decoding self-attention for the decode step (one query token per utterance over the
// static K/V cache [B][heads][c_max][64], bf16 or fp8 e4m3): replaces bmm + additive mask + softmax + bmm and the
// torch.cat of the cache of /root/reference/parler_tts/modeling_parler_tts.py:299-304, 326-365 on the cached branch.
//
// One PERSISTENT CTA per SM walks its (utterance, head) items; per item the old cache rows stream through a deep
// shared-memory ring as 128-row tiles (cp.async.bulk.tensor, SWIZZLE_128B / SWIZZLE_64B for one-byte elements), the
// two products run on the 5th-generation tensor core with the tile as the A operand straight from the ring, and 128
// softmax threads (one per key of a tile = one TMEM lane) only touch scores and probabilities:
//   warp 0      producer: one lane issues the tile copies of K(item), V(item), K(next item), ... - everything streamed
//               is immutable during the step (rows below the slot this step appends), so the producer runs ahead of the
//               dependency on the QKV GEMM and of the softmax of earlier items; up to ~190 KB in flight per SM;
//   warp 1      score-MMA issuer (one lane): S[128 keys x 16] = Ktile . q^T (tcgen05.mma M=128 N=16, A K-major) into a
//               ring of 8 accumulators in tensor memory, completion by tcgen05.commit;
//   warp 2      P.V-MMA issuer (one lane): O[64 dims x 8] += Vtile^T . P^T (tcgen05.mma M=64 N=8, A MN-major = the very
//               [key][dim] tile the copy wrote); a thread of its own, so S(i+1) never queues behind the wait for P(i);
//   warp 3      helper: folds the split-K partial sums of the QKV GEMM for the items ahead (q scaled, new K / V row),
//               appends the new row to the cache, builds the q operand tile, scores the new row in fp32;
//   warps 4-11  two softmax groups of 128 threads taking alternate items: tcgen05.ld of the scores, mask, max,
//               probabilities as the B operand of the second product, final normalisation and the bf16 output row.
// Narrow operands keep fp32-class accuracy by splitting the small operand over the idle N columns: q = q_hi + q_lo and
// p = p_hi + p_lo (bf16), or three e4m3 digits q1 + q2/16 + q3/256 (fp8); the big operand (the cache) is used as stored.
// The new row (slot ctx-1) never travels through the ring: its score and its p.v term come from registers.
#include <cuda.h>
#include <cuda_fp16.h>

#include "gemm_common.cuh"

namespace ptts {

using gemm::smem_u32;

namespace {

constexpr int WS_THREADS = 384;
constexpr int TILE = 128;  // keys per tile = TMEM lanes of the score accumulator
constexpr int WD = 64;     // head dim
constexpr uint32_t kTmemCols = 512;  // [0,480): 30 score accumulators (N = 16), [480,496): one output accumulator (N = 8) per softmax group
constexpr uint32_t kSpinLimit = 1u << 22;  // ~1 s of try_wait polls
constexpr int kMaxRing = 24;

struct WsParams {
  const float* part;  // split-K partial sums [S][B][3H] of the QKV GEMM
  int S;
  long long stride;
  int ncols;          // 3H
  float scale;        // 1 / sqrt(d)
  int slot_base;      // rows already cached = slot_base + *slot_dyn; the new row is written to that slot
  const int* slot_dyn;
  void* ck;           // [B][heads][c_stride][64] of this layer
  void* cv;
  const float* k_scale;  // fp8: [heads]
  const float* v_scale;
  int heads, c_stride, n_items;
  const int* key_mask;   // [B][mask_len] or null
  int mask_len;
  __nv_bfloat16* out;    // [B][H]
  int ring;              // ring slots
  int sc_cap;            // floats of the score buffer
};

// ---- PTX wrappers (bounded waits: a protocol bug must trap, not hang the box)
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  for (uint32_t spin = 0; !done; ++spin) {
    asm volatile(
        "{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.b32 %0, 1, 0, P1;\n\t}\n"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (spin > kSpinLimit) __trap();
  }
}
// the single issuing lanes spin with test_wait: a try_wait that finds the phase incomplete suspends the thread for a
// while (measured in fast_gemm3.cu: eight sequential waits cost ~4000 cycles that way)
__device__ __forceinline__ void mbar_spin(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  for (uint32_t spin = 0; !done; ++spin) {
    asm volatile(
        "{\n\t.reg .pred P1;\n\tmbarrier.test_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.b32 %0, 1, 0, P1;\n\t}\n"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (spin > kSpinLimit) __trap();
  }
}
__device__ __forceinline__ void tma_tile_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }
__device__ __forceinline__ void bar_named(int id, int count) {
  asm volatile("bar.sync %0, %1;\n" ::"r"(id), "r"(count) : "memory");
}
template <bool FP8>
__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  if constexpr (FP8)
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
  else
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_x8(uint32_t taddr, uint32_t (&v)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];\n"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "r"(taddr)
               : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
}

// Shared-memory matrix descriptors (cute::UMMA::SmemDescriptor): start >> 4 | LBO >> 4 << 16 | SBO >> 4 << 32 |
// version 1 << 46 | layout << 61 (SWIZZLE_128B = 2, SWIZZLE_64B = 4).  SBO = distance between groups of 8 rows along
// the non-contiguous direction (8 x 128 B or 8 x 64 B), for K-major and MN-major operands alike; LBO (distance between
// swizzle atoms along the contiguous direction) is not exercised: every operand here is one atom wide.
__device__ __forceinline__ uint64_t desc_sw128(uint32_t saddr) {
  return (uint64_t)((saddr >> 4) & 0x3FFF) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ uint64_t desc_sw64(uint32_t saddr) {
  return (uint64_t)((saddr >> 4) & 0x3FFF) | (1ull << 16) | (32ull << 32) | (1ull << 46) | (4ull << 61);
}
// Instruction descriptors: D fp32 (bit 4) | A fmt << 7 | B fmt << 10 | A MN-major << 15 | N >> 3 << 17 | M >> 4 << 24
// (cta_group::1 shapes: M = 64 takes N % 8 == 0, M = 128 takes N % 16 == 0)
constexpr uint32_t idesc(int fmt, int a_mn, int M, int N) {
  return (1u << 4) | ((uint32_t)fmt << 7) | ((uint32_t)fmt << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)(N >> 3) << 17) |
         ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ unsigned char to_e4m3(float x) {
  unsigned short r;
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;\n" : "=h"(r) : "f"(0.0f), "f"(x));
  return (unsigned char)(r & 0xff);
}
__device__ __forceinline__ float from_e4m3(unsigned char b) {
  uint32_t r;
  const unsigned short s = b;
  asm("cvt.rn.f16x2.e4m3x2 %0, %1;\n" : "=r"(r) : "h"(s));
  return __half2float(__ushort_as_half((unsigned short)(r & 0xffff)));
}
// x (|x| <= 256) as three e4m3 digits: x ~ d0 + d1 / 16 + d2 / 256
__device__ __forceinline__ void split3_e4m3(float x, unsigned char (&d)[3], float& rec) {
  d[0] = to_e4m3(x);
  const float f0 = from_e4m3(d[0]);
  const float r1 = (x - f0) * 16.0f;
  d[1] = to_e4m3(r1);
  const float f1 = from_e4m3(d[1]);
  const float r2 = (r1 - f1) * 16.0f;
  d[2] = to_e4m3(r2);
  rec = f0 + f1 * 0.0625f + from_e4m3(d[2]) * 0.00390625f;
}

// byte offset of element (row r < 8, element e) of an 8-row operand tile: 128-byte rows with SWIZZLE_128B (16-byte
// chunk c at c ^ r; a second 64-element atom follows at +1024 B), or 64-byte rows with SWIZZLE_64B (c ^ (r >> 1))
__device__ __forceinline__ int off_sw128(int r, int byte_in_row) {
  const int atom = byte_in_row >> 7, b = byte_in_row & 127;
  return atom * 1024 + r * 128 + ((((b >> 4) ^ r) & 7) << 4) + (b & 15);
}
__device__ __forceinline__ int off_sw64(int r, int byte_in_row) {
  return r * 64 + ((((byte_in_row >> 4) ^ (r >> 1)) & 3) << 4) + (byte_in_row & 15);
}

}  // namespace

namespace {

// Score accumulators in tensor memory (16 columns each).  mbarrier waits are by phase PARITY, so a waiter that is a whole
// phase ahead of its barrier gets a false positive: a softmax group's first wait of an item is only ordered behind the
// buffer's previous use if that use belongs to an item the group has seen complete, which needs NS > tiles per item
// (profiles/scripts/attn_ws_protocol_sim.py replays the protocol with parity semantics and random schedules).
constexpr int NS = 30;

// v2 roles (384 threads): warp 0 TMA producer | warp 1 score-MMA issuer | warp 2 P.V-MMA issuer | warp 3 helper (q operand,
// new K / V row) | warps 4-7 softmax group 0 | warps 8-11 softmax group 1 (groups take alternate items; warp % 4 = the TMEM
// lane quadrant a warp may read).  The two MMA issuers are independent threads, so S(i+1) is issued while P.V(i) still
// waits for its probabilities: consecutive items of a CTA overlap instead of paying their latency chains back to back.
template <bool FP8>
__global__ void __launch_bounds__(WS_THREADS, 1)
attn_ws_kernel(const WsParams p, const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_v) {
  constexpr int ES = FP8 ? 1 : 2;
  constexpr int ROWB = WD * ES;                // bytes per cache row
  constexpr uint32_t TB = TILE * ROWB;         // bytes per ring slot (16 KB | 8 KB)
  constexpr int KS_QK = FP8 ? 2 : 4;           // 64 dims / UMMA_K (32 | 16)
  constexpr int KS_PV = FP8 ? 4 : 8;           // 128 keys / UMMA_K
  constexpr uint32_t IDESC_QK = idesc(FP8 ? 0 : 1, 0, 128, 16);  // E4M3 = 0, BF16 = 1
  constexpr uint32_t IDESC_PV = idesc(FP8 ? 0 : 1, 1, 64, 8);
  constexpr uint32_t O_COL = NS * 16;          // output accumulators (8 columns per softmax group) behind the scores
  extern __shared__ unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t bar_full[kMaxRing], bar_empty[kMaxRing];
  __shared__ __align__(8) uint64_t bar_qfull[4], bar_done[4], bar_sfull[NS], bar_sempty[NS];
  __shared__ __align__(8) uint64_t bar_pfull[2][2], bar_pempty[2][2], bar_ofull[2], bar_oempty[2];
  __shared__ uint32_t tmem_slot;
  __shared__ float vnew[4][WD], snew[4], sfac_s[4], red_max[2][4], red_sum[2][4];

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;
  unsigned char* smem = smem_raw + (sbase - smem_u32(smem_raw));
  const int ring = p.ring;
  const uint32_t s_ring = sbase;
  const uint32_t s_ptile = s_ring + (uint32_t)ring * TB;  // [group][buffer] x 2 KB
  const uint32_t s_qtile = s_ptile + 4 * 2048;            // [item & 3] x 2 KB (16 rows: N = 16 of the first product)
  unsigned char* g_ptile = smem + (size_t)ring * TB;
  unsigned char* g_qtile = g_ptile + 4 * 2048;
  float* sc_all = reinterpret_cast<float*>(g_qtile + 4 * 2048);  // [group][sc_cap]

  // rows already in the cache; immutable for the whole step, and the counter itself is two launches old (decode-step
  // contract, see fast_attn.cu): everything below may run before griddepcontrol.wait
  const int n_old = dyn_value(p.slot_base, p.slot_dyn);
  const int nt = (n_old + TILE - 1) / TILE;
  const int H = p.heads * WD;
  const int n_it = ((int)p.n_items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;  // items of this CTA

  // ---- one-time setup
  for (int i = tid; i < (4 * 2048 + 4 * 2048) / 16; i += WS_THREADS)
    reinterpret_cast<uint4*>(g_ptile)[i] = make_uint4(0u, 0u, 0u, 0u);  // unused operand rows stay zero
  if (tid == 0) {
    for (int i = 0; i < ring; ++i) {
      mbar_init(smem_u32(&bar_full[i]), 1);
      mbar_init(smem_u32(&bar_empty[i]), 1);
    }
    for (int i = 0; i < 4; ++i) {
      mbar_init(smem_u32(&bar_qfull[i]), 32);
      mbar_init(smem_u32(&bar_done[i]), 128);
    }
    for (int i = 0; i < NS; ++i) {
      mbar_init(smem_u32(&bar_sfull[i]), 1);
      mbar_init(smem_u32(&bar_sempty[i]), 128);
    }
    for (int g = 0; g < 2; ++g) {
      for (int i = 0; i < 2; ++i) {
        mbar_init(smem_u32(&bar_pfull[g][i]), 128);
        mbar_init(smem_u32(&bar_pempty[g][i]), 1);
      }
      mbar_init(smem_u32(&bar_ofull[g]), 1);
      mbar_init(smem_u32(&bar_oempty[g]), 128);
    }
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(&tmem_slot)),
                 "r"(kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
  }
  fence_async_smem();  // the zeroed operand tiles are read through the async proxy
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;

  if (warp == 0) {
    // ================================================================ producer: K(0) V(0) K(1) V(1) ... in ring order
    if (elect_one()) {  // single-lane issue from uniform registers (common.cuh)
      uint32_t cnt = 0;
      for (int it = 0; it < n_it; ++it) {
        const int row0 = ((int)blockIdx.x + it * (int)gridDim.x) * p.c_stride;
        for (int pass = 0; pass < 2; ++pass)
          for (int t = 0; t < nt; ++t, ++cnt) {
            const uint32_t slot = cnt % (uint32_t)ring, use = cnt / (uint32_t)ring;
            if (use > 0) mbar_spin(smem_u32(&bar_empty[slot]), (use & 1u) ^ 1u);
            mbar_expect_tx(smem_u32(&bar_full[slot]), TB);
            tma_tile_2d(s_ring + slot * TB, pass ? &tm_v : &tm_k, smem_u32(&bar_full[slot]), 0, row0 + t * TILE);
          }
      }
    }
  } else if (warp == 1) {
    // ================================================================ score-MMA issuer: S = Ktile . q^T
    if (elect_one()) {  // single-lane issue from uniform registers (common.cuh)
      for (int it = 0; it < n_it; ++it) {
        const int par = it & 3;
        mbar_spin(smem_u32(&bar_qfull[par]), (uint32_t)(it >> 2) & 1u);
        for (int t = 0; t < nt; ++t) {
          const uint32_t cnt = (uint32_t)(it * 2 * nt + t), slot = cnt % (uint32_t)ring;
          const uint32_t n = (uint32_t)(it * nt + t), buf = n % NS;
          // the slot's previous use (possibly a V tile the other issuer consumes) must be CONSUMED before this thread
          // may interpret the parity of `full`: otherwise a wait one phase ahead returns at once (needs ring >= nt + 2)
          const uint32_t use = cnt / (uint32_t)ring;
          if (use > 0) mbar_spin(smem_u32(&bar_empty[slot]), (use & 1u) ^ 1u);
          mbar_spin(smem_u32(&bar_full[slot]), use & 1u);
          if (n >= NS) mbar_spin(smem_u32(&bar_sempty[buf]), ((n / NS) & 1u) ^ 1u);
          tc_fence_after();
          const uint32_t a0 = s_ring + slot * TB, b0 = s_qtile + par * 2048;
#pragma unroll
          for (int ks = 0; ks < KS_QK; ++ks)
            umma<FP8>(tmem + buf * 16, FP8 ? desc_sw64(a0 + ks * 32) : desc_sw128(a0 + ks * 32),
                      FP8 ? desc_sw64(b0 + ks * 32) : desc_sw128(b0 + ks * 32), IDESC_QK, ks > 0);
          umma_commit(smem_u32(&bar_sfull[buf]));
          umma_commit(smem_u32(&bar_empty[slot]));
        }
      }
    }
  } else if (warp == 2) {
    // ================================================================ P.V-MMA issuer: O[g] += Vtile^T . P^T
    if (elect_one()) {  // single-lane issue from uniform registers (common.cuh)
      for (int it = 0; it < n_it; ++it) {
        const int g = it & 1, my = it >> 1;
        for (int t = 0; t < nt; ++t) {
          const uint32_t cnt = (uint32_t)(it * 2 * nt + nt + t), slot = cnt % (uint32_t)ring;
          const uint32_t pb_cnt = (uint32_t)(my * nt + t), pb = pb_cnt & 1u;
          // probabilities first: P(i, t) ready means S(i) is complete, hence every earlier tile of this slot has been
          // consumed and the parity of `full` is unambiguous (this thread waits on `full` from kernel start otherwise)
          mbar_spin(smem_u32(&bar_pfull[g][pb]), (pb_cnt >> 1) & 1u);
          mbar_spin(smem_u32(&bar_full[slot]), (cnt / (uint32_t)ring) & 1u);
          if (t == 0 && my > 0) mbar_spin(smem_u32(&bar_oempty[g]), (uint32_t)(my - 1) & 1u);
          tc_fence_after();
          const uint32_t a0 = s_ring + slot * TB, b0 = s_ptile + (uint32_t)(g * 2 + (int)pb) * 2048;
#pragma unroll
          for (int ks = 0; ks < KS_PV; ++ks) {
            const uint32_t bk = FP8 ? b0 + ks * 32 : b0 + (ks >> 2) * 1024 + (ks & 3) * 32;
            umma<FP8>(tmem + O_COL + g * 8, FP8 ? desc_sw64(a0 + ks * 2048) : desc_sw128(a0 + ks * 2048), desc_sw128(bk),
                      IDESC_PV, (t > 0 || ks > 0) ? 1u : 0u);
          }
          umma_commit(smem_u32(&bar_empty[slot]));
          umma_commit(smem_u32(&bar_pempty[g][pb]));
          if (t == nt - 1) umma_commit(smem_u32(&bar_ofull[g]));
        }
      }
    }
  } else if (warp == 3) {
    // ================================================================ helper: q operand, new K / V row (lane = dims lane, lane + 32)
    pdl_wait();    // the QKV partial sums are the only producer-dependent input of this kernel
    pdl_launch();
    for (int it = 0; it < n_it; ++it) {
      const int par = it & 3;
      const int item = (int)blockIdx.x + it * (int)gridDim.x;
      const int b = item / p.heads, hd = item - b * p.heads;
      if (it >= 4) mbar_wait(smem_u32(&bar_done[par]), (uint32_t)((it >> 2) - 1) & 1u);
      const float* p0 = p.part + (long long)b * p.ncols + hd * WD + lane;
      float aq[2] = {0.f, 0.f}, ak[2] = {0.f, 0.f}, av[2] = {0.f, 0.f};
      for (int s0 = 0; s0 < p.S; s0 += 2) {  // 2 slices (12 loads) in flight, added in slice order
        float tq[2][2], tk[2][2], tv[2][2];
#pragma unroll
        for (int u = 0; u < 2; ++u)
          if (s0 + u < p.S) {
            const float* ps = p0 + (long long)(s0 + u) * p.stride;
#pragma unroll
            for (int e = 0; e < 2; ++e) {
              tq[u][e] = __ldcg(ps + e * 32);
              tk[u][e] = __ldcg(ps + H + e * 32);
              tv[u][e] = __ldcg(ps + 2 * H + e * 32);
            }
          }
#pragma unroll
        for (int u = 0; u < 2; ++u)
          if (s0 + u < p.S) {
#pragma unroll
            for (int e = 0; e < 2; ++e) {
              aq[e] += tq[u][e];
              ak[e] += tk[u][e];
              av[e] += tv[u][e];
            }
          }
      }
      const long long o = ((long long)item * p.c_stride + n_old) * WD + lane;
      unsigned char* qt = g_qtile + par * 2048;
      float dot = 0.f, sf = 1.0f;
      if constexpr (FP8) {
        const float q0 = aq[0] * p.scale, q1 = aq[1] * p.scale;
        const float qmax = warp_max(fmaxf(fabsf(q0), fabsf(q1)));
        const float qs = qmax > 0.f ? 256.0f / qmax : 0.f;
        const float ks = p.k_scale[hd], vs = p.v_scale[hd];
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          unsigned char dgt[3];
          float qrec;
          split3_e4m3((e ? q1 : q0) * qs, dgt, qrec);
#pragma unroll
          for (int r = 0; r < 3; ++r) qt[off_sw64(r, lane + e * 32)] = dgt[r];
          const unsigned char k8 = to_e4m3(ak[e] / ks), v8 = to_e4m3(av[e] / vs);
          reinterpret_cast<unsigned char*>(p.ck)[o + e * 32] = k8;
          reinterpret_cast<unsigned char*>(p.cv)[o + e * 32] = v8;
          dot += qrec * from_e4m3(k8);
          vnew[par][lane + e * 32] = from_e4m3(v8);
        }
        sf = qmax * (1.0f / 256.0f) * ks;
      } else {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const float q = aq[e] * p.scale;
          const __nv_bfloat16 qh = __float2bfloat16_rn(q);
          const __nv_bfloat16 ql = __float2bfloat16_rn(q - __bfloat162float(qh));
          *reinterpret_cast<__nv_bfloat16*>(qt + off_sw128(0, (lane + e * 32) * 2)) = qh;
          *reinterpret_cast<__nv_bfloat16*>(qt + off_sw128(1, (lane + e * 32) * 2)) = ql;
          const __nv_bfloat16 kr = __float2bfloat16_rn(ak[e]), vr = __float2bfloat16_rn(av[e]);
          reinterpret_cast<__nv_bfloat16*>(p.ck)[o + e * 32] = kr;
          reinterpret_cast<__nv_bfloat16*>(p.cv)[o + e * 32] = vr;
          dot += (__bfloat162float(qh) + __bfloat162float(ql)) * __bfloat162float(kr);
          vnew[par][lane + e * 32] = __bfloat162float(vr);
        }
      }
      dot = warp_sum(dot);  // score of the new row, fp32, from the values the cache now holds
      if (lane == 0) {
        snew[par] = dot * sf;
        sfac_s[par] = sf;
      }
      fence_async_smem();  // the q tile is read by the tensor core (async proxy)
      mbar_arrive(smem_u32(&bar_qfull[par]));
    }
  } else {
    // ================================================================ softmax groups: 128 threads, thread = key of a tile
    const int g = (warp - 4) >> 2;      // group 0: warps 4-7, group 1: warps 8-11
    const int quad = warp & 3;          // TMEM lane quadrant this warp may read
    const int lt = quad * 32 + lane;    // key within the tile
    const uint32_t t_lane = tmem + ((uint32_t)(quad * 32) << 16);
    const float neg = -3.3895313892515355e38f;  // finfo(bf16).min, the reference's additive mask value
    float* sc = sc_all + (size_t)g * p.sc_cap;
    int my = 0;
    for (int it = g; it < n_it; it += 2, ++my) {
      const int par = it & 3;
      const int item = (int)blockIdx.x + it * (int)gridDim.x;
      const int b = item / p.heads, hd = item - b * p.heads;
      mbar_wait(smem_u32(&bar_qfull[par]), (uint32_t)(it >> 2) & 1u);
      const float sf = sfac_s[par], s_new = snew[par];
      float lmax = s_new;
      for (int t = 0; t < nt; ++t) {
        const uint32_t n = (uint32_t)(it * nt + t), buf = n % NS;
        mbar_wait(smem_u32(&bar_sfull[buf]), (n / NS) & 1u);
        tc_fence_after();
        uint32_t v[8];
        tmem_ld_x8(t_lane + buf * 16, v);
        tc_fence_before();
        mbar_arrive(smem_u32(&bar_sempty[buf]));
        const int j = t * TILE + lt;
        float s = FP8 ? (__uint_as_float(v[0]) + __uint_as_float(v[1]) * 0.0625f + __uint_as_float(v[2]) * 0.00390625f)
                      : (__uint_as_float(v[0]) + __uint_as_float(v[1]));
        s *= sf;
        if (j >= n_old)
          s = -INFINITY;  // rows past the context (the tile is padded by the neighbouring cache rows)
        else if (p.key_mask && j < p.mask_len && p.key_mask[(long long)b * p.mask_len + j] == 0)
          s = neg;
        sc[j] = s;
        lmax = fmaxf(lmax, s);
      }
      lmax = warp_max(lmax);
      if (lane == 0) red_max[g][quad] = lmax;
      bar_named(1 + g, 128);
      const float mx = fmaxf(fmaxf(red_max[g][0], red_max[g][1]), fmaxf(red_max[g][2], red_max[g][3]));
      float lsum = 0.f;
      for (int t = 0; t < nt; ++t) {
        const uint32_t pb_cnt = (uint32_t)(my * nt + t), pb = pb_cnt & 1u;
        if (pb_cnt >= 2) mbar_wait(smem_u32(&bar_pempty[g][pb]), ((pb_cnt >> 1) & 1u) ^ 1u);
        const float pj = __expf(sc[t * TILE + lt] - mx);
        lsum += pj;
        unsigned char* pt = g_ptile + (g * 2 + (int)pb) * 2048;
        if constexpr (FP8) {
          unsigned char dgt[3];
          float rec;
          split3_e4m3(pj * 256.0f, dgt, rec);
#pragma unroll
          for (int r = 0; r < 3; ++r) pt[off_sw128(r, lt)] = dgt[r];
        } else {
          const __nv_bfloat16 ph = __float2bfloat16_rn(pj);
          const __nv_bfloat16 pl = __float2bfloat16_rn(pj - __bfloat162float(ph));
          *reinterpret_cast<__nv_bfloat16*>(pt + off_sw128(0, lt * 2)) = ph;
          *reinterpret_cast<__nv_bfloat16*>(pt + off_sw128(1, lt * 2)) = pl;
        }
        fence_async_smem();
        mbar_arrive(smem_u32(&bar_pfull[g][pb]));
      }
      const float p_new = __expf(s_new - mx);
      lsum = warp_sum(lsum);
      if (lane == 0) red_sum[g][quad] = lsum;
      bar_named(1 + g, 128);
      const float total = ((red_sum[g][0] + red_sum[g][1]) + (red_sum[g][2] + red_sum[g][3])) + p_new;
      float o = 0.f;
      if (nt > 0) {
        mbar_wait(smem_u32(&bar_ofull[g]), (uint32_t)my & 1u);
        tc_fence_after();
        uint32_t v[8];
        tmem_ld_x8(t_lane + O_COL + g * 8, v);  // M = 64: output row m sits in lane (m % 16) + 32 * (m / 16)
        tc_fence_before();
        mbar_arrive(smem_u32(&bar_oempty[g]));
        o = FP8 ? (__uint_as_float(v[0]) + __uint_as_float(v[1]) * 0.0625f + __uint_as_float(v[2]) * 0.00390625f) *
                      0.00390625f
                : (__uint_as_float(v[0]) + __uint_as_float(v[1]));
      }
      if (lane < 16) {
        const int dim = quad * 16 + lane;
        o = (o + p_new * vnew[par][dim]) / total;
        if constexpr (FP8) o *= p.v_scale[hd];
        p.out[(long long)b * H + hd * WD + dim] = __float2bfloat16_rn(o);
      }
      mbar_arrive(smem_u32(&bar_done[par]));
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem), "r"(kTmemCols) : "memory");
}

}  // namespace

// ---- host side
bool attention_ws_supported(int B, int heads, int d, int c_stride) {
  if (d != WD || !gemm::tma_available()) return false;
  static int sms = 0;  // host state: SM count of the device (queried once)
  if (sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  }
  // a persistent CTA per SM pays off once every SM has several items; smaller batches keep the split-KV kernels
  const int nt = (c_stride + TILE - 1) / TILE;
  // protocol constraints (see NS): fewer tiles per pass than score accumulators, and a ring of 16 KB slots (bf16, the
  // larger case) that holds more than one pass next to the operand tiles and the two score buffers
  const long long fixed = 1024 + 4 * 2048 + 4 * 2048 + 2LL * 4 * nt * TILE;
  const long long ring = (225LL * 1024 - fixed) / (TILE * WD * 2);
  return B * heads >= 2 * sms && nt < NS && ring >= nt + 2;
}

template <bool FP8>
static int launch_ws(WsParams p, int B, cudaStream_t st) {
  constexpr int ES = FP8 ? 1 : 2;
  constexpr size_t TB = (size_t)TILE * WD * ES;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  p.n_items = B * p.heads;
  p.sc_cap = (p.c_stride + TILE - 1) / TILE * TILE;
  const size_t fixed = 1024 + 4 * 2048 + 4 * 2048 + 2 * sizeof(float) * (size_t)p.sc_cap;
  const size_t budget = 225 * 1024;
  PTTS_CHECK_ARG(fixed + 4 * TB <= budget, "attention_ws: cache too long for the score buffer");
  int ring = (int)((budget - fixed) / TB);
  if (ring > kMaxRing) ring = kMaxRing;
  if (ring < p.sc_cap / TILE + 2) {  // protocol constraint (see NS): the ring must hold more than one pass of an item
    set_error("attention_ws: cache too long for the tile ring (%d slots, %d tiles per pass)", ring, p.sc_cap / TILE);
    return PTTS_E_UNSUPPORTED;
  }
  p.ring = ring;
  const size_t smem = fixed + (size_t)ring * TB;
  CUtensorMap tm_k, tm_v;
  const int rows = p.n_items * p.c_stride;
  if (!gemm::tensor_map_2d_dt(p.ck, rows, WD, TILE, FP8 ? PTTS_FP8 : PTTS_BF16, &tm_k) ||
      !gemm::tensor_map_2d_dt(p.cv, rows, WD, TILE, FP8 ? PTTS_FP8 : PTTS_BF16, &tm_v)) {
    set_error("attention_ws: no tensor map (driver entry point missing or misaligned cache)");
    return PTTS_E_UNSUPPORTED;
  }
  static size_t configured[2] = {0, 0};  // host state: opt-in dynamic shared memory high-water mark per instantiation
  if (smem > configured[FP8 ? 1 : 0]) {
    cudaFuncSetAttribute(attn_ws_kernel<FP8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    configured[FP8 ? 1 : 0] = smem;
  }
  const int grid = p.n_items < sms ? p.n_items : sms;
  const cudaError_t e = launch_k(attn_ws_kernel<FP8>, dim3(grid), dim3(WS_THREADS), smem, st, p, tm_k, tm_v);
  if (e != cudaSuccess) {
    set_error("attn_ws launch: %s", cudaGetErrorString(e));
    (void)cudaGetLastError();
    return PTTS_E_LAUNCH;
  }
  PTTS_CHECK_LAUNCH("attn_ws");
  return PTTS_OK;
}

// Fused decode-step self-attention: q / new k / new v arrive as split-K partial sums [S][B][3H] of the QKV GEMM.
// kv_dtype PTTS_BF16 or PTTS_FP8 (then k_scale / v_scale [heads]).  slot = rows already cached = index of the new row.
int attention_ws_fused(const float* qkv_part, int S, int64_t part_stride, int ncols, float scale, int slot_base,
                       const int* slot_dyn, void* ck, void* cv, int kv_dtype, const float* k_scale, const float* v_scale,
                       int B, int heads, int c_stride, const int* key_mask, int mask_len, void* out, cudaStream_t st) {
  WsParams p = {};
  p.part = qkv_part; p.S = S; p.stride = part_stride; p.ncols = ncols; p.scale = scale;
  p.slot_base = slot_base; p.slot_dyn = slot_dyn; p.ck = ck; p.cv = cv; p.k_scale = k_scale; p.v_scale = v_scale;
  p.heads = heads; p.c_stride = c_stride; p.key_mask = key_mask; p.mask_len = mask_len;
  p.out = (__nv_bfloat16*)out;
  if (kv_dtype == PTTS_FP8) return launch_ws<true>(p, B, st);
  return launch_ws<false>(p, B, st);
}

}  // namespace ptts
