// tcgen05 (UMMA) split-K GEMM for the decode step, "swap-AB": the weight tile is the 128-row A operand, the batch
// rows are the 64-wide N operand, the fp32 accumulator [128 weight rows x 64 batch rows] lives in tensor memory.
//   D[n][m] = sum_k W[n0 + n][k] * X[m][k]       (both operands K-major, SWIZZLE_128B shared-memory layout)
// Why tcgen05 here although the step is HBM-bound: at M = 64 the decode step still needs ~46 GFLOP per 0.6 ms of
// weight streaming (~75 TFLOP/s sustained, bursts of 400+ while a tile is resident); measured on B200 the legacy
// mma.sync path with 4 warps per SM made every linear compute/latency-bound (5-17 us per launch, <= 1.2 TB/s of
// weights, profiles/r01_gemm_sweep.txt).  One elected thread issues K/16 tcgen05.mma instructions per CTA instead.
// Structure per CTA (128 threads): cp.async the whole W slab [128 x K/CS] (before griddepcontrol.wait: weights do
// not depend on the previous kernel) and X slab [64 x K/CS] into 128-byte-swizzled rows -> fence.proxy.async ->
// tcgen05.mma burst -> tcgen05.commit -> mbarrier -> tcgen05.ld (32x32b) -> transposed fp32 tile in shared memory ->
// cluster barrier -> batched DSMEM loads of the CS K-slices, summed in rank order (deterministic) -> arrive at the
// exit barrier -> fused epilogue (gemm_common.cuh) on 64/CS rows per CTA -> wait on the exit barrier.
// (In-kernel stamps, profiles/scripts/gemm_stamps.py: pushing partials with 4-byte st.shared::cluster costs 7.9 us per
// launch; unbatched DSMEM loads + a blocking exit barrier cost 3.2 us; hence batched pulls and a split barrier.)
#include <cuda.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <unordered_map>

#include "gemm_common.cuh"

namespace ptts {

using namespace gemm;

namespace {

constexpr int UM = 128;            // weight rows per CTA (UMMA M)
constexpr int UN = 64;             // batch rows per CTA (UMMA N)
constexpr int KB64 = 64;           // K elements per 128-byte swizzled row block
constexpr int PITCH3 = UM + 4;     // fp32 pitch of the transposed reduction tile [m][n]
constexpr int WBLK = UM * 128;     // bytes of one W k-block (128 rows x 128 B)
constexpr int XBLK = UN * 128;     // bytes of one X k-block
constexpr uint32_t TMEM_COLS = 64;
constexpr int kMaxKB64 = 8;  // <= 512 K elements per CTA: (128 + 64) rows x 1 KB = 192 KB of operand slabs

__device__ __forceinline__ uint32_t cluster_rank3() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;\n" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync3() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
__device__ __forceinline__ float2 ld_dsmem_f2_3(uint32_t local_addr, uint32_t rank) {
  uint32_t ra;
  float2 v;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;\n" : "=r"(ra) : "r"(local_addr), "r"(rank));
  asm volatile("ld.shared::cluster.v2.f32 {%0,%1}, [%2];\n" : "=f"(v.x), "=f"(v.y) : "r"(ra) : "memory");
  return v;
}

// ROWS x 64 bf16 block: 128-byte rows, 16-byte chunk c of row r at chunk position c ^ (r & 7)  (= SWIZZLE_128B)
template <int ROWS>
__device__ __forceinline__ void load_block_sw128(uint32_t sbase, const __nv_bfloat16* g, int ld, int row0,
                                                 int rows_total, int k0) {
#pragma unroll
  for (int i = threadIdx.x; i < ROWS * 8; i += NTH) {
    const int r = i >> 3, c = i & 7;
    const int gr = row0 + r;
    const bool ok = gr < rows_total;
    const __nv_bfloat16* src = ok ? g + (int64_t)gr * ld + k0 + c * 8 : g;
    cp_async16(sbase + r * 128 + ((c ^ (r & 7)) << 4), src, ok);
  }
}

// K-major, SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor): start >> 4 | LBO (unused for
// swizzled K-major, 1) << 16 | SBO = 1024 B (8 rows x 128 B) >> 4 << 32 | version 1 << 46 | layout SWIZZLE_128B (2) << 61
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
  return (uint64_t)((saddr >> 4) & 0x3FFF) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// kind::f16 instruction descriptor: D fp32, A/B bf16, both K-major, N = 64, M = 128
constexpr uint32_t kIdesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(UN >> 3) << 17) | ((uint32_t)(UM >> 4) << 24);

__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(kIdesc), "r"(accumulate)
      : "memory");
}

__device__ __forceinline__ void tmem_ld_32x32b_x32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
               : "r"(taddr)
               : "memory");
}
// PTTS_GEMM_DBG=1: CTA (0,0,0) of every launch records SM-clock / globaltimer stamps (read back by
// ptts_diag_gemm_stamps; profiles/scripts/gemm_stamps.py) - where inside the kernel does the time go?
__device__ unsigned long long g_stamps[256 * 8];
__device__ unsigned int g_stamp_idx;
// PTTS_GEMM_DBG=3: per launch slot (dbg >> 8), sums over ALL CTAs (and replays) of SM-clock cycles since the CTA's
// griddepcontrol.wait returned: [0] W slab landed, [1] X slab landed, [2] MMA done, [3] CTA end, [4] CTA count,
// [5] max W, [6] max end, [7] max X
__device__ unsigned long long g_agg[256 * 8];
int g_agg_seq = 0;  // host: launch slot of the next launch
__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(t));
  return t;
}
#define STAMP(i)                                                    \
  do {                                                              \
    if (dbg_slot >= 0) g_stamps[dbg_slot * 8 + (i)] = (dbg & 3) == 2 ? (unsigned long long)clock64() : gtime(); \
  } while (0)

// ---- TMA operand loads (dbg bit 32, PTTS_GEMM_TMA=0 to disable): the W slab [128 x KB*64] and the X slab [64 x KB*64]
// arrive as KB box copies each (cp.async.bulk.tensor.2d, box = 64 elements x 128 / 64 rows, SWIZZLE_128B = the very
// shared-memory image load_block_sw128 builds), issued by ONE thread and tracked by an mbarrier.  With 16-byte
// cp.async the 8192 requests of a 128 KB slab trickle through the LSU (measured: slab landed 1.6 us AFTER the
// dependency resolved although it was requested 4+ us before); the TMA engine keeps the whole slab in flight.
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void mbar_expect_tx3(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait3(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  for (uint32_t spin = 0; !done; ++spin) {
    asm volatile(
        "{\n\t.reg .pred P1;\n\tmbarrier.test_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.b32 %0, 1, 0, P1;\n\t}\n"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    // test_wait, not try_wait: a try_wait that finds the phase incomplete suspends the thread for a while - eight
    // sequential k-block waits cost 4000 cycles that way (measured: X "landed" 5172 instead of 1550 cycles after the
    // dependency)
    if (spin > 200000000u) __trap();  // a protocol bug must fail loudly, not hang the box
  }
}

// 3-D box copy: tensor map dims {64 elements, rows, k-blocks}, box {64, box_rows, box_kb} = box_kb consecutive
// SWIZZLE_128B k-blocks of box_rows rows, laid out back to back exactly like box_kb 2-D copies
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];\n" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
// tcgen05.mma from the 32-bit low words of two K-major SWIZZLE_128B descriptors (high word: SBO 1024 B, version 1,
// SWIZZLE_128B = 0x40004040, see umma_desc_sw128)
__device__ __forceinline__ void umma_f16_lo(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "mov.b64 da, {%1, %5};\n\tmov.b64 db, {%2, %5};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %3, p;\n\t}\n" ::"r"(tmem_d),
      "r"(a_lo), "r"(b_lo), "r"(kIdesc), "r"(accumulate), "r"(0x40004040u)
      : "memory");
}

// grid = (N/128, M/64, CS), cluster (1,1,CS); KB = 64-element k-blocks per CTA (<= 8)
// LEAN = the instantiation the decode step launches when nothing optional is asked for (TMA loads, single X barrier,
// no diagnostics, no hint, no finisher): every optional path is compiled out.  It matters: the code of an unused
// finisher in this one-shot kernel cost 0.24 us per launch (1.457 against 1.434 ms per decode step, same box) - the
// instruction fetch of a kernel that runs each line once is part of its critical path.
// FULLM (lean PARTIAL launches with M % 64 == 0): the ragged-tile copy of the store loop is compiled out too
template <int EPI, bool LEAN, bool FULLM = false>
__global__ void __launch_bounds__(NTH)
umma_gemm_kernel(const GemmParams p, int KB, int CS, int dbg_in, const L2Hint hint, const __grid_constant__ CUtensorMap tm_w,
                 const __grid_constant__ CUtensorMap tm_x, TL tl_in) {
  const int dbg = LEAN ? ((dbg_in & (16 | 128)) | 32 | 64) : dbg_in;
  TL tl = tl_in;
  if constexpr (LEAN) tl.p = nullptr;
  extern __shared__ unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t mma_bar;
  __shared__ __align__(8) uint64_t w_bar, x_bar[kMaxKB64];  // one barrier per X k-block: MMAs start as blocks land
  const bool tma = (dbg & 32) != 0;
  __shared__ uint32_t tmem_base_slot;
  const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SWIZZLE_128B atoms need 1024-byte alignment
  unsigned char* smem = smem_raw + (sbase - smem_u32(smem_raw));
  const uint32_t sw = sbase, sx = sbase + KB * WBLK;
  const int n0 = blockIdx.x * UM, m0 = blockIdx.y * UN;
  const uint32_t rank = (EPI == EPI_PARTIAL) ? blockIdx.z : cluster_rank3();  // K slice of this CTA
  const int kbeg = (int)rank * KB * KB64;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int dbg_slot = -1;
  if ((dbg & 3) && (dbg & 3) != 3 && threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0 && rank == 0)
    dbg_slot = (int)(atomicAdd(&g_stamp_idx, 1u) & 255u);
  STAMP(0);
  tl_start(tl);

  if (!tma)
    for (int kb = 0; kb < KB; ++kb) load_block_sw128<UM>(sw + kb * WBLK, p.W, p.K, n0, p.N, kbeg + kb * KB64);
  if (warp == 0) {  // tensor-memory allocation (one warp), barrier init (one thread)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(&tmem_base_slot)),
                 "r"(TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
  }
  if (warp == 1 && elect_one()) {  // single-lane issue from uniform registers (common.cuh)
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_u32(&mma_bar)), "r"(1) : "memory");
    if (tma) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_u32(&w_bar)), "r"(1) : "memory");
      for (int kb = 0; kb < KB; ++kb)
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_u32(&x_bar[kb])), "r"(1) : "memory");
    }
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    if (tma) {  // the whole W slab of this CTA, before the dependency resolves (weights are immutable)
      asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
      mbar_expect_tx3(smem_u32(&w_bar), (uint32_t)(KB * WBLK));
      for (int kb = 0; kb < KB; ++kb) tma_load_2d(sw + kb * WBLK, &tm_w, smem_u32(&w_bar), kbeg + kb * KB64, n0);
    }
  }
  if constexpr (!LEAN) {
    cp_async_commit();  // group 0: the W slab
    l2_hint_prefetch(hint, (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x,
                     gridDim.x * gridDim.y * gridDim.z);
  }
  // dbg bit 4 (split-K partial GEMMs, PTTS_PDL_EARLY=0 to disable): release the dependent kernel BEFORE waiting on
  // the own producer, so that its CTAs are resident and its pre-wait prologue (gamma / beta, first K batch, L2
  // prefetches) has run by the time this grid completes.  Safe because every consumer touches producer data only
  // after its own griddepcontrol.wait (which waits for THIS grid to complete and flush), and what the consumers read
  // before their wait is immutable for the whole step: weights, cache rows below the append slot, and the position
  // counter, which only the step epilogue writes - the epilogue never triggers early and a graph's first node has a
  // full dependency on the previous replay, so no early-resident kernel can see a stale counter.
  if (dbg & 16) pdl_launch();
  pdl_wait();
  if (!(dbg & (16 | 128))) pdl_launch();  // bit 7: release the dependent only when the MMA burst is done (see below)
  STAMP(1);
  tl_dep(tl);
  const long long c_dep = (dbg & 3) == 3 ? clock64() : 0;
  long long c_w = 0, c_x = 0, c_mma = 0;
  if (tma) {
    if (warp == 1 && elect_one()) {
      if (dbg & 64) {  // PTTS_GEMM_XBAR=1: one barrier for the whole X slab
        mbar_expect_tx3(smem_u32(&x_bar[0]), (uint32_t)(KB * XBLK));
        for (int kb = 0; kb < KB; ++kb) tma_load_2d(sx + kb * XBLK, &tm_x, smem_u32(&x_bar[0]), kbeg + kb * KB64, m0);
      } else {
        // all expectations first, then the copies back to back (an arrive between two copies delays the second)
        for (int kb = 0; kb < KB; ++kb) mbar_expect_tx3(smem_u32(&x_bar[kb]), (uint32_t)XBLK);
        for (int kb = 0; kb < KB; ++kb) tma_load_2d(sx + kb * XBLK, &tm_x, smem_u32(&x_bar[kb]), kbeg + kb * KB64, m0);
      }
    }
  } else {
    for (int kb = 0; kb < KB; ++kb) load_block_sw128<UN>(sx + kb * XBLK, p.X, p.K, m0, p.M, kbeg + kb * KB64);
    cp_async_commit();  // group 1: the X slab
    if ((dbg & 3) == 3) {
      cp_async_wait<1>();
      c_w = clock64();
    }
    cp_async_wait<0>();
    if ((dbg & 3) == 3) c_x = clock64();
  }
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");  // cp.async (generic proxy) -> UMMA (async proxy)
  asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
  const uint32_t tmem_d = tmem_base_slot;
  STAMP(2);

  if (warp == 0 && elect_one()) {
    if (tma) {  // barrier inits are visible since the __syncthreads above; the slabs arrive through the async proxy
      mbar_wait3(smem_u32(&w_bar), 0);
      if ((dbg & 3) == 3) c_w = clock64();
    }
    for (int kb = 0; kb < KB; ++kb) {
      if (tma && (kb == 0 || !(dbg & 64))) {
        mbar_wait3(smem_u32(&x_bar[kb]), 0);
        if ((dbg & 3) == 3 && (kb == KB - 1 || (dbg & 64))) c_x = clock64();
        asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
      }
      const uint64_t ad = umma_desc_sw128(sw + kb * WBLK), bd = umma_desc_sw128(sx + kb * XBLK);
#pragma unroll
      for (int k = 0; k < KB64 / 16; ++k)  // +32 bytes (2 x 16 B) per UMMA_K = 16 inside the swizzle atom
        umma_f16(tmem_d, ad + 2 * k, bd + 2 * k, (kb | k) ? 1u : 0u);
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(
                     smem_u32(&mma_bar))
                 : "memory");
  }
  {  // every thread waits for the MMA burst (bounded spin: a descriptor bug must trap, not hang the box)
    uint32_t done = 0;
    for (uint32_t spin = 0; !done; ++spin) {
      asm volatile(
          "{\n\t.reg .pred P1;\n\tmbarrier.test_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.b32 %0, 1, 0, P1;\n\t}\n"
          : "=r"(done)
          : "r"(smem_u32(&mma_bar)), "r"(0)
          : "memory");
      if (spin > 200000000u) __trap();
    }
  }
  asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
  // the out_proj in front of the fused cross-attention block releases it here: released at the dependency wait, the
  // block's pre-wait burst (128 KB of Wq per SM) lands on this kernel's MMA / epilogue (measured 5.4 us against 3.3 us
  // for the other GEMMs)
  if ((dbg & 128) && !(dbg & 16)) pdl_launch();
  STAMP(3);
  if ((dbg & 3) == 3) c_mma = clock64();
  auto agg = [&]() {
    if ((dbg & 3) == 3 && threadIdx.x == 0) {
      unsigned long long* a = g_agg + ((dbg >> 8) & 255) * 8;
      const long long c_end = clock64();
      atomicAdd(a + 0, (unsigned long long)(c_w - c_dep));
      atomicAdd(a + 1, (unsigned long long)(c_x - c_dep));
      atomicAdd(a + 2, (unsigned long long)(c_mma - c_dep));
      atomicAdd(a + 3, (unsigned long long)(c_end - c_dep));
      atomicAdd(a + 4, 1ull);
      atomicMax(a + 5, (unsigned long long)(c_w - c_dep));
      atomicMax(a + 6, (unsigned long long)(c_end - c_dep));
      atomicMax(a + 7, (unsigned long long)(c_x - c_dep));
    }
  };

  if constexpr (EPI == EPI_PARTIAL) {
    // split-K partial sums straight to global memory [slice][m][n] (coalesced along n); the LayerNorm that consumes
    // the residual stream next adds them up in slice order (deterministic) - no cluster, no on-chip reduction
    const int n = n0 + warp * 32 + lane;
    STAMP(4);
    // one running pointer per thread and no per-element index arithmetic: this loop is the kernel's critical path
    const size_t ldn = (size_t)p.N;
    float* dst = p.part + ((size_t)rank * p.M + m0) * ldn + n;
    const int rows = min(UN, p.M - m0);
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      uint32_t v[32];
      tmem_ld_32x32b_x32(tmem_d + ((uint32_t)(warp * 32) << 16) + half * 32, v);
      asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
      if (n < p.N) {
        if (FULLM || rows == UN) {
#pragma unroll
          for (int j = 0; j < 32; ++j) dst[(size_t)j * ldn] = __uint_as_float(v[j]);
        } else if constexpr (!FULLM) {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (half * 32 + j < rows) dst[(size_t)j * ldn] = __uint_as_float(v[j]);
        }
      }
      dst += 32 * ldn;
    }
    STAMP(5);
    asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
    __syncthreads();
    if (warp == 0) {
      asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_d), "r"(TMEM_COLS) : "memory");
    }
    STAMP(6);
    agg();
    tl_end(tl);
    return;
  }
  // accumulator -> registers -> transposed fp32 tile red[m][n] in my own shared memory (the operand slabs are dead):
  // thread (warp w, lane l) owns weight row n = 32 w + l and the 64 batch-row columns
  const int rows_per = UN / CS;
  float* red = reinterpret_cast<float*>(smem);
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    uint32_t v[32];
    tmem_ld_32x32b_x32(tmem_d + ((uint32_t)(warp * 32) << 16) + half * 32, v);
    asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
#pragma unroll
    for (int j = 0; j < 32; ++j) red[(half * 32 + j) * PITCH3 + warp * 32 + lane] = __uint_as_float(v[j]);
  }
  // residual pairs of my first 4 epilogue elements depend on earlier kernels only: fetch them now, the L2 round trip
  // hides behind the cluster barrier
  const int pairs = rows_per * (UM / 2);
  float2 res[4];
  auto fetch_res = [&](int e0) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      res[u] = make_float2(0.f, 0.f);
      if constexpr (EPI == EPI_STORE) {
        const int e = e0 + u * NTH;
        const int row = m0 + (int)rank * rows_per + e / (UM / 2), col = n0 + (e % (UM / 2)) * 2;
        if (p.residual && e < pairs && row < p.M && col < p.N)
          res[u] = __ldcg(reinterpret_cast<const float2*>(p.residual + (int64_t)row * p.N + col));
      }
    }
  };
  fetch_res(threadIdx.x);
  asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
  cluster_sync3();  // all CS partial tiles are visible cluster-wide
  STAMP(4);
  if (warp == 0) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_d), "r"(TMEM_COLS) : "memory");
  }

  // deterministic reduction over the CS K-slices (rank order): all DSMEM loads of up to 4 elements are issued
  // back to back; after the last batch the CTA ARRIVES at the exit barrier (peers may release my tile as soon as they
  // have read it) and only waits on it after its own global stores
  for (int e0 = threadIdx.x; e0 < pairs; e0 += NTH * 4) {
    if (e0 != (int)threadIdx.x) fetch_res(e0);
    float2 part[4][8];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int e = e0 + u * NTH;
      const int lr = (int)rank * rows_per + e / (UM / 2), lc = (e % (UM / 2)) * 2;
      const uint32_t la = smem_u32(red + lr * PITCH3 + lc);
#pragma unroll
      for (int r = 0; r < 8; ++r)
        if (r < CS && e < pairs) part[u][r] = ld_dsmem_f2_3(la, (uint32_t)r);
    }
    if (e0 + NTH * 4 >= pairs + (int)threadIdx.x)  // uniform: the last batch of this CTA
      asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int e = e0 + u * NTH;
      if (e >= pairs) break;
      float v0 = 0.f, v1 = 0.f;
#pragma unroll
      for (int r = 0; r < 8; ++r)
        if (r < CS) {
          v0 += part[u][r].x;
          v1 += part[u][r].y;
        }
      const int row = m0 + (int)rank * rows_per + e / (UM / 2), col = n0 + (e % (UM / 2)) * 2;
      if (row < p.M && col < p.N) epilogue_pair<EPI>(p, row, col, 0, v0, v1, &res[u]);
    }
  }
  STAMP(5);
  asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");  // nobody exits while a peer reads its tile
  STAMP(6);
  agg();
  tl_end(tl);
}


// ---- lean split-K partial kernel, second generation (round 2).  Same math and the same output as
// umma_gemm_kernel<EPI_PARTIAL, true, true>, rebuilt around what its SASS showed: the single issuing thread was the
// critical path.  Issued from a lane of a divergent branch, every cp.async.bulk.tensor / tcgen05.mma became an
// ELECT ... BRA.U.ANY loop with R2UR moves and a 64-bit descriptor rebuild in front of it (~45 instructions per
// k-block: ~2500 cycles from "X landed" to "MMA done" for a burst the tensor core finishes in ~1100, and ~100 instructions
// for the eight X box copies after the dependency wait).  Here
//   * one warp-uniform branch + elect.sync, so the uniform datapath issues without the per-lane loops;
//   * ONE 3-D box copy for the whole X slab (tensor map viewed as [k-block][row][64]), two for the W slab;
//   * the KB*4 MMAs fully unrolled, descriptors = a 32-bit low word + compile-time offset (the high word is constant);
//   * barrier init / TMEM allocation / the only __syncthreads in front of the dependency wait.
// X slab as XP box copies with a barrier each: the MMAs of a part run while the later parts land.  Measured on B200
// (B=64, 860 frames, same box): 1 part 1.2514, 2 parts 1.2403 ms per decode step
#ifndef PTTS_GEMM_XPARTS
#define PTTS_GEMM_XPARTS 2
#endif
template <int KB>
__global__ void __launch_bounds__(NTH)
umma_partial_lean_kernel(float* __restrict__ part, int M, int N, int late, const __grid_constant__ CUtensorMap tm_w,
                         const __grid_constant__ CUtensorMap tm_x) {
  extern __shared__ unsigned char smem_raw[];
  constexpr int XP = (KB % PTTS_GEMM_XPARTS == 0) ? PTTS_GEMM_XPARTS : 1;
  __shared__ __align__(8) uint64_t bars[2 + XP];  // W slab, MMA burst, X parts
  __shared__ uint32_t tmem_base_slot;
  const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SWIZZLE_128B atoms need 1024-byte alignment
  const uint32_t sw = sbase, sx = sbase + KB * WBLK;
  const int n0 = blockIdx.x * UM, m0 = blockIdx.y * UN, rank = blockIdx.z;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t w_bar = smem_u32(&bars[0]), mma_bar = smem_u32(&bars[1]), x_bar = smem_u32(&bars[2]);
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(&tmem_base_slot)),
                 "r"(TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
  } else if (warp == 1) {
    if (elect_one()) {
#pragma unroll
      for (int i = 0; i < 2 + XP; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(w_bar + 8 * i) : "memory");
      asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
      asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
      // the whole W slab of this CTA, before the dependency resolves (weights are immutable)
      mbar_expect_tx3(w_bar, (uint32_t)(KB * WBLK));
      tma_load_3d(sw, &tm_w, w_bar, 0, n0, rank * KB);
      tma_load_3d(sw + (KB / 2) * WBLK, &tm_w, w_bar, 0, n0, rank * KB + KB / 2);
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
  __syncthreads();  // barrier inits and the TMEM base address are visible to everybody - still in front of the wait
  asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
  const uint32_t tmem_d = tmem_base_slot;
  if (late == 2) pdl_launch();  // before the own wait (diagnostic / variant builds only)
  pdl_wait();
  if (late == 0) pdl_launch();
  if (warp == 1 && elect_one()) {
    const uint32_t a0 = ((sw >> 4) & 0x3FFFu) | (1u << 16), b0 = ((sx >> 4) & 0x3FFFu) | (1u << 16);
    // all expectations first, then the copies back to back (an arrive between two copies delays the second)
#pragma unroll
    for (int i = 0; i < XP; ++i) mbar_expect_tx3(x_bar + 8 * i, (uint32_t)(KB / XP * XBLK));
#pragma unroll
    for (int i = 0; i < XP; ++i)
      tma_load_3d(sx + i * (KB / XP) * XBLK, &tm_x, x_bar + 8 * i, 0, m0, rank * KB + i * (KB / XP));
    mbar_wait3(w_bar, 0);
#pragma unroll
    for (int i = 0; i < XP; ++i) {
      mbar_wait3(x_bar + 8 * i, 0);
      asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
      // k-block loop rolled (fully unrolled, ptxas hoists all 2 * 4 * KB descriptor pairs and spills uniform registers)
#pragma unroll 1
      for (int kb = i * (KB / XP); kb < (i + 1) * (KB / XP); ++kb)
#pragma unroll
        for (int k = 0; k < KB64 / 16; ++k)  // +32 bytes (2 x 16 B) per UMMA_K = 16 inside the swizzle atom
          umma_f16_lo(tmem_d, a0 + kb * (WBLK >> 4) + 2 * k, b0 + kb * (XBLK >> 4) + 2 * k, (kb | k) ? 1u : 0u);
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(mma_bar) : "memory");
  }
  __syncwarp();
  mbar_wait3(mma_bar, 0);
  asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
  if (late == 1) pdl_launch();
  // split-K partial sums straight to global memory [slice][m][n] (coalesced along n); the consumer folds them in slice
  // order.  Thread (warp w, lane l) owns weight row n0 + 32 w + l and its 64 batch-row columns.
  const int n = n0 + warp * 32 + lane;
  const size_t ldn = (size_t)N;
  float* dst = part + ((size_t)rank * M + m0) * ldn + n;
  uint32_t v0[32], v1[32];
  tmem_ld_32x32b_x32(tmem_d + ((uint32_t)(warp * 32) << 16), v0);
  tmem_ld_32x32b_x32(tmem_d + ((uint32_t)(warp * 32) << 16) + 32, v1);
  asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
  if (n < N) {
#pragma unroll
    for (int j = 0; j < 32; ++j) dst[(size_t)j * ldn] = __uint_as_float(v0[j]);
    dst += 32 * ldn;
#pragma unroll
    for (int j = 0; j < 32; ++j) dst[(size_t)j * ldn] = __uint_as_float(v1[j]);
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
  __syncthreads();
  if (warp == 0)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_d), "r"(TMEM_COLS) : "memory");
}

int plan_cluster3(int M, int N, int K) {
  if (K % 128 != 0) return 0;
  const int kblocks = K / KB64;
  const int nt = ((N + UM - 1) / UM) * ((M + UN - 1) / UN);
#ifdef PTTS_DIAG
  const int forced = PTTS_KNOB("PTTS_CS", 0);
  if (forced > 0 && kblocks % forced == 0 && kblocks / forced <= kMaxKB64) return forced;
#endif
  for (int cs = 1; cs <= 8; cs <<= 1) {
    if (kblocks % cs) continue;
    const int kb = kblocks / cs;
    if (kb > kMaxKB64) continue;
    const bool last = cs == 8 || (kblocks % (cs * 2)) != 0;
    if ((nt * cs >= 120 && (kb <= 4 || nt * cs <= 148)) || last) return cs;
  }
  return 0;
}

thread_local int g_late_once = 0;

// ---- tensor maps (host): one per (matrix, shape, box rows), created on first use through the driver entry point (no
// link-time dependency on libcuda) and cached; the map is passed to the kernel by value (__grid_constant__)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
    (void)cudaGetLastError();
  }
  return fn;
}
// Everything a descriptor encodes is part of the key (base, extents, row pitch, box, element type): a buffer that is
// freed and re-allocated at the same address with another shape or pitch gets its own entry, never a stale map.
struct MapKey {
  const void* ptr;
  int rows, cols, box_rows, dtype;
  long long pitch_bytes;
  int box_kb;  // 0: 2-D map; > 0: 3-D map [k-block][row][64] whose box spans box_kb k-blocks
  bool operator==(const MapKey& o) const {
    return ptr == o.ptr && rows == o.rows && cols == o.cols && box_rows == o.box_rows && dtype == o.dtype &&
           pitch_bytes == o.pitch_bytes && box_kb == o.box_kb;
  }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    return std::hash<const void*>()(k.ptr) ^ (size_t)k.rows * 0x9E3779B97F4A7C15ull ^ (size_t)k.cols * 0xC2B2AE3D27D4EB4Full ^
           (size_t)k.box_rows ^ ((size_t)k.dtype << 20) ^ (size_t)k.pitch_bytes * 0x165667B19E3779F9ull ^
           ((size_t)k.box_kb << 40);
  }
};
// row-major [rows][cols] matrix, box = 64 columns x box_rows rows, out-of-range rows read as 0:
//   PTTS_BF16: 128-byte box rows, SWIZZLE_128B;  PTTS_FP8 (bytes): 64-byte box rows, SWIZZLE_64B
//   box_kb > 0 (bf16 only): the same matrix viewed as [cols/64][rows][64] with box {64, box_rows, box_kb} - ONE copy
//   brings box_kb consecutive SWIZZLE_128B k-blocks, laid out back to back like box_kb 2-D copies
bool tensor_map_2d_any(const void* ptr, int rows, int cols, int box_rows, int dtype, CUtensorMap* out, int box_kb = 0) {
  static std::mutex mu;
  static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
  EncodeTiledFn fn = encode_tiled_fn();
  const int es = dtype == PTTS_FP8 ? 1 : 2;
  if (!fn || (reinterpret_cast<uintptr_t>(ptr) & 15) || (cols * es) % 16) return false;
  std::lock_guard<std::mutex> lock(mu);
  if (box_kb > 0 && (dtype != PTTS_BF16 || cols % KB64)) return false;
  const MapKey key = {ptr, rows, cols, box_rows, dtype, (long long)cols * es, box_kb};
  auto it = cache.find(key);
  if (it != cache.end()) {
    *out = it->second;
    return true;
  }
  CUtensorMap m;
  const cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t gstr[1] = {(cuuint64_t)cols * es};
  const cuuint32_t box[2] = {(cuuint32_t)KB64, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  CUresult r;
  if (box_kb > 0) {
    const cuuint64_t gdim3[3] = {(cuuint64_t)KB64, (cuuint64_t)rows, (cuuint64_t)(cols / KB64)};
    const cuuint64_t gstr3[2] = {(cuuint64_t)cols * es, (cuuint64_t)KB64 * es};
    const cuuint32_t box3[3] = {(cuuint32_t)KB64, (cuuint32_t)box_rows, (cuuint32_t)box_kb};
    const cuuint32_t estr3[3] = {1, 1, 1};
    r = fn(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), gdim3, gstr3, box3, estr3,
           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  } else {
    r = fn(&m, dtype == PTTS_FP8 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
           const_cast<void*>(ptr), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
           dtype == PTTS_FP8 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
           CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  if (r != CUDA_SUCCESS) return false;
  if (cache.size() > 4096) cache.clear();
  cache.emplace(key, m);
  *out = m;
  return true;
}
bool tensor_map_2d_impl(const void* ptr, int rows, int cols, int box_rows, CUtensorMap* out) {
  return tensor_map_2d_any(ptr, rows, cols, box_rows, PTTS_BF16, out);
}

template <int EPI>
int launch_umma(const GemmParams& p, cudaStream_t st, int splits = 0) {
  const int cs = EPI == EPI_PARTIAL ? splits : plan_cluster3(p.M, p.N, p.K);
  if (cs == 0) return PTTS_E_UNSUPPORTED;
  const int kb = p.K / KB64 / cs;
  size_t smem = (size_t)kb * (WBLK + XBLK);  // operand slabs, reused as the fp32 reduction tile
  const size_t red = (size_t)UN * PITCH3 * sizeof(float);
  if (smem < red) smem = red;
  smem += 1024;  // alignment slack
  static size_t configured = 0;
  if (smem > configured) {
    cudaFuncSetAttribute(umma_gemm_kernel<EPI, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(umma_gemm_kernel<EPI, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (EPI == EPI_PARTIAL) {
      cudaFuncSetAttribute(umma_gemm_kernel<EPI_PARTIAL, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      cudaFuncSetAttribute(umma_partial_lean_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      cudaFuncSetAttribute(umma_partial_lean_kernel<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      cudaFuncSetAttribute(umma_partial_lean_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    }
    configured = smem;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((p.N + UM - 1) / UM, (p.M + UN - 1) / UN, cs);
  cfg.blockDim = dim3(NTH);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  int n = 0;
  attr[n].id = cudaLaunchAttributeClusterDimension;
  attr[n].val.clusterDim.x = 1;
  attr[n].val.clusterDim.y = 1;
  attr[n].val.clusterDim.z = EPI == EPI_PARTIAL ? 1 : cs;
  ++n;
  if (pdl_enabled()) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  // diagnostic build: 1 globaltimer stamps (256 ns ticks), 2 SM clock cycles (intra-kernel phases), 3 per-CTA sums
  const int dbg = PTTS_KNOB("PTTS_GEMM_DBG", 0);
  // with cp.async slabs the early release won (1.727 against 1.748 ms per decode step); with TMA slabs the slab needs
  // ~1.5 us of lead only and the late release wins (1.486 against 1.496 ms) -> late
    // (re-measured on the round-2 kernels: early 1.3299 against 1.2206 ms; release at MMA done 1.2664)
  const int early = PTTS_KNOB("PTTS_PDL_EARLY", 0) ? 16 : 0;
  int early_now = early;
  if (EPI == EPI_PARTIAL && g_late_once) {
    early_now = g_late_once == 2 ? 128 : 0;
    g_late_once = 0;
  }
  const int dbg_arg = (dbg == 3 ? (3 | ((g_agg_seq++ & 255) << 8)) : dbg) | (EPI == EPI_PARTIAL ? early_now : 0);
  const L2Hint hint = EPI == EPI_PARTIAL ? take_l2_hint() : L2Hint{nullptr, 0};
  const int use_tma = PTTS_KNOB("PTTS_GEMM_TMA", 1);
  CUtensorMap tm_w, tm_x;
  memset(&tm_w, 0, sizeof(tm_w));
  memset(&tm_x, 0, sizeof(tm_x));
  const bool tma = use_tma && tensor_map_2d_impl(p.W, p.N, p.K, UM, &tm_w) && tensor_map_2d_impl(p.X, p.M, p.K, UN, &tm_x);
  // one barrier for the whole X slab, not one per k-block.  Measured on B200: per k-block the last X block "lands"
  // 5300 cycles after the dependency instead of 1500 (each arrive.expect_tx between two bulk-tensor issues costs ~500
  // cycles) - 1.566 against 1.485 ms per decode step
  const int xbar = PTTS_KNOB("PTTS_GEMM_XBAR", 1) ? 64 : 0;
  const TL tl = tl_next(10 + EPI);
  const int lean_ok = PTTS_KNOB("PTTS_GEMM_LEAN", 1);  // diagnostic build, 0: always the full instantiation
  const bool lean = lean_ok && tma && xbar && !(dbg_arg & ~(16 | 128)) && !hint.ptr && !tl.p;
  // second-generation lean kernel (3-D box copies, unrolled issue): k-block counts the decode step uses
  // (a ragged last weight tile is fine: the box copy zero-fills rows >= N and the store loop skips them)
  if (lean && EPI == EPI_PARTIAL && p.M % UN == 0 && (kb == 8 || kb == 6 || kb == 4) &&
      PTTS_KNOB("PTTS_GEMM_GEN2", 1)) {
    CUtensorMap t3w, t3x;
    if (tensor_map_2d_any(p.W, p.N, p.K, UM, PTTS_BF16, &t3w, kb / 2) &&
        tensor_map_2d_any(p.X, p.M, p.K, UN, PTTS_BF16, &t3x, kb % PTTS_GEMM_XPARTS == 0 ? kb / PTTS_GEMM_XPARTS : kb)) {
            const int late = (dbg_arg & 16) ? 2 : (dbg_arg & 128) ? 1 : 0;
      cudaError_t e2 = kb == 8   ? cudaLaunchKernelEx(&cfg, umma_partial_lean_kernel<8>, p.part, p.M, p.N, late, t3w, t3x)
                       : kb == 6 ? cudaLaunchKernelEx(&cfg, umma_partial_lean_kernel<6>, p.part, p.M, p.N, late, t3w, t3x)
                                 : cudaLaunchKernelEx(&cfg, umma_partial_lean_kernel<4>, p.part, p.M, p.N, late, t3w, t3x);
      if (e2 != cudaSuccess) {
        set_error("umma_partial_lean launch: %s", cudaGetErrorString(e2));
        (void)cudaGetLastError();
        return PTTS_E_LAUNCH;
      }
      PTTS_CHECK_LAUNCH("umma_partial_lean");
      return PTTS_OK;
    }
  }
  cudaError_t e =
      (lean && EPI == EPI_PARTIAL && p.M % UN == 0)
          ? cudaLaunchKernelEx(&cfg, umma_gemm_kernel<EPI_PARTIAL, true, true>, p, kb, cs, dbg_arg, hint, tm_w, tm_x, tl)
      : lean ? cudaLaunchKernelEx(&cfg, umma_gemm_kernel<EPI, true>, p, kb, cs, dbg_arg, hint, tm_w, tm_x, tl)
           : cudaLaunchKernelEx(&cfg, umma_gemm_kernel<EPI, false>, p, kb, cs, dbg_arg | (tma ? 32 : 0) | xbar, hint, tm_w,
                                tm_x, tl);
  if (e != cudaSuccess) {
    set_error("umma_gemm launch: %s", cudaGetErrorString(e));
    (void)cudaGetLastError();
    return PTTS_E_LAUNCH;
  }
  PTTS_CHECK_LAUNCH("umma_gemm");
  return PTTS_OK;
}

}  // namespace

void gemm_next_trigger_late(int at_mma_done) { g_late_once = at_mma_done ? 2 : 1; }

namespace gemm {
bool tma_available() { return encode_tiled_fn() != nullptr; }
bool tensor_map_2d(const void* ptr, int rows, int cols, int box_rows, CUtensorMap* out) {
  return tensor_map_2d_impl(ptr, rows, cols, box_rows, out);
}
bool tensor_map_3d(const void* ptr, int rows, int cols, int box_rows, int box_kb, CUtensorMap* out) {
  return tensor_map_2d_any(ptr, rows, cols, box_rows, PTTS_BF16, out, box_kb);
}
bool tensor_map_2d_dt(const void* ptr, int rows, int cols, int box_rows, int dtype, CUtensorMap* out) {
  return tensor_map_2d_any(ptr, rows, cols, box_rows, dtype, out);
}
int umma_read_stamps(unsigned long long* host, int n_launches) {
  if (n_launches == -2) {  // reset the aggregate statistics
    static unsigned long long zero[256 * 8];
    cudaMemcpyToSymbol(g_agg, zero, sizeof(zero));
    return 0;
  }
  if (n_launches < 0) {  // aggregate statistics of PTTS_GEMM_DBG=3 (256 x 8)
    cudaMemcpyFromSymbol(host, g_agg, sizeof(unsigned long long) * 8 * 256);
    return 256;
  }
  unsigned int idx = 0;
  cudaMemcpyFromSymbol(&idx, g_stamp_idx, sizeof(idx));
  cudaMemcpyFromSymbol(host, g_stamps, sizeof(unsigned long long) * 8 * (n_launches > 256 ? 256 : n_launches));
  return (int)idx;
}
// split-K plan of the PARTIAL variant: smallest S in {2,4,8,16} that gives >= ~30 CTAs (else the largest that fits)
int umma_plan_splits(int M, int N, int K) {
  if (K % 128 != 0) return 0;
  const int kblocks = K / KB64;
  const int nt = ((N + UM - 1) / UM) * ((M + UN - 1) / UN);
  // CTAs wanted per residual GEMM (fewer slices = cheaper LayerNorm fold): with TMA slab loads fewer, larger CTAs
  // win, 30 -> 16 measured 1.654 -> 1.559 ms per step
    const int target = PTTS_KNOB("PTTS_SPLIT_TARGET", 16);  // 32 re-measured in round 2: 1.2811 against 1.2206 ms
  // 64-element k-blocks per CTA (8 = 128 KB of W per CTA; 4 = 64 KB measured slower, 1.630 against 1.497 ms)
  int kbmax = PTTS_KNOB("PTTS_KBMAX", kMaxKB64);
  if (kbmax < 1 || kbmax > kMaxKB64) kbmax = kMaxKB64;
  int best = 0;
  for (int s = 2; s <= kMaxSplitK; s <<= 1) {
    if (kblocks % s || kblocks / s > kbmax) continue;
    best = s;
    if (nt * s >= target) break;
  }
  return best;
}
int umma_gemm_partial(const GemmParams& p, int splits, cudaStream_t st) {
  return launch_umma<EPI_PARTIAL>(p, st, splits);
}
int umma_gemm_store(const GemmParams& p, cudaStream_t st) { return launch_umma<EPI_STORE>(p, st); }
int umma_gemm_scatter(const GemmParams& p, cudaStream_t st) { return launch_umma<EPI_SCATTER>(p, st); }
}  // namespace gemm

}  // namespace ptts
