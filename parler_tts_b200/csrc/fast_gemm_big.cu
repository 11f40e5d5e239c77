// Large-M bf16 GEMM of the once-per-call phases (prompt prefill, cross K/V projection, T5 encoder): y = epilogue(x . W^T)
// for M >= 256 token rows (nn.Linear of /root/reference/parler_tts/modeling_parler_tts.py:282,297-302,380,488-490 on the
// un-cached branch).
//
// The decode-step kernels of fast_gemm3.cu are weight-streaming designs (128 weight rows x 64 token rows per CTA, K split
// across a cluster): at M = 2112 (B=64 x 33 prefill positions) every weight tile is re-read 33 times and every token tile
// 24 times from L2 - 312 MB of L2 traffic for a 13 GFLOP GEMM, 99 us per launch, 134 TFLOP/s
// (profiles/r02_prefill_t5_kernels_v14.txt).  This is the plain tile kernel those shapes want:
//   * CTA tile 128 token rows (tcgen05.mma M) x 128 weight rows (N), K in 64-element blocks through a 5-stage ring of
//     32 KB (A 16 KB + B 16 KB, cp.async.bulk.tensor 2-D box copies, SWIZZLE_128B);
//   * persistent CTAs (one per SM), token tiles fastest so that the CTAs of a wave share weight tiles in L2;
//   * warp 0 = TMA producer, warp 1 = MMA issuer (both a single elected lane), warps 2..5 = epilogue; two 128-column
//     TMEM accumulators so that the epilogue of tile i runs under the MMAs of tile i + 1;
//   * the fused epilogues of gemm_common.cuh (STORE: alpha / bias / activation / row mask / residual; SCATTER: q | k | v
//     into their buffers and cache slots), row by row: a TMEM lane is a token row, so a warp that stored straight from
//     its registers wrote 32 different rows per instruction (one 4-byte pair per 32-byte sector; with an fp32 residual
//     that cost 27 us per launch at M = 2112: 48 against 21 us for the same shape without one).  Each epilogue warp turns its
//     32 x 64 block through a strip of shared memory and then walks the rows with lanes = consecutive column pairs:
//     256-byte residual reads and 128-byte (bf16) / 256-byte (fp32) stores per instruction.
//   Known limit (profiles/r02_prefill_t5_launches_v19.txt): the operand stages arrive at ~13 bytes per clock per SM and the
//   tensor pipe is busy 11 - 21 % of the time; ~200 - 300 TFLOP/s on the prefill shapes.  Once-per-call work (~1 % of a call).
#include <cuda.h>
#include <string.h>

#include "gemm_common.cuh"

namespace ptts {

using namespace gemm;

namespace {

constexpr int GM = 128, GN = 128, GK = 64;
constexpr int GSTAGES = 5;
constexpr int GSTRIP_PITCH = 64 + 4;                    // floats per strip row (the pad keeps float4 row writes at 4 wavefronts)
constexpr int GSTRIP_BYTES = 32 * GSTRIP_PITCH * 4;     // one epilogue warp's 32 x 64 strip
constexpr int GA_BYTES = GM * 128, GB_BYTES = GN * 128, GSTAGE_BYTES = GA_BYTES + GB_BYTES;
constexpr int GTHREADS = 192;
constexpr uint32_t GTMEM_COLS = 256;
// kind::f16 instruction descriptor: D fp32, A/B bf16, both K-major, N = 128, M = 128
constexpr uint32_t kIdescBig = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(GN >> 3) << 17) | ((uint32_t)(GM >> 4) << 24);

__device__ __forceinline__ void big_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void big_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(bar) : "memory");
}
// single issuing lanes spin (test_wait); the 128 epilogue threads use try_wait
template <bool SPIN>
__device__ __forceinline__ void big_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  for (uint32_t spin = 0; !done; ++spin) {
    if constexpr (SPIN)
      asm volatile(
          "{\n\t.reg .pred P1;\n\tmbarrier.test_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.b32 %0, 1, 0, P1;\n\t}\n"
          : "=r"(done)
          : "r"(bar), "r"(parity)
          : "memory");
    else
      asm volatile(
          "{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.b32 %0, 1, 0, P1;\n\t}\n"
          : "=r"(done)
          : "r"(bar), "r"(parity)
          : "memory");
    if (spin > (SPIN ? 400000000u : 8000000u)) __trap();  // a protocol bug must fail loudly, not hang the box
  }
}
__device__ __forceinline__ void big_tma_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
// tcgen05.mma from the low words of two K-major SWIZZLE_128B descriptors (high word 0x40004040, see fast_gemm3.cu)
__device__ __forceinline__ void big_umma(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "mov.b64 da, {%1, %5};\n\tmov.b64 db, {%2, %5};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %3, p;\n\t}\n" ::"r"(tmem_d),
      "r"(a_lo), "r"(b_lo), "r"(kIdescBig), "r"(accumulate), "r"(0x40004040u)
      : "memory");
}
__device__ __forceinline__ void big_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void big_tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
               : "r"(taddr)
               : "memory");
}

// tiles are numbered token-tile fastest: tile = nt * MT + mt
template <int EPI>
__global__ void __launch_bounds__(GTHREADS, 1)
gemm_big_kernel(const GemmParams p, int MT, int NT, const __grid_constant__ CUtensorMap tm_x,
                const __grid_constant__ CUtensorMap tm_w) {
  extern __shared__ unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t bars[2 * GSTAGES + 4];  // full[6] | empty[6] | acc_full[2] | acc_empty[2]
  __shared__ uint32_t tmem_base_slot;
  const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SWIZZLE_128B atoms need 1024-byte alignment
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t full = smem_u32(&bars[0]), empty = smem_u32(&bars[GSTAGES]);
  const uint32_t acc_full = smem_u32(&bars[2 * GSTAGES]), acc_empty = smem_u32(&bars[2 * GSTAGES + 2]);
  const int ntiles = MT * NT, KB = p.K / GK;
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(&tmem_base_slot)),
                 "r"(GTMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
  } else if (warp == 0) {
    if (elect_one()) {
      for (int i = 0; i < GSTAGES; ++i) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(full + 8 * i) : "memory");
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(empty + 8 * i) : "memory");
      }
      for (int i = 0; i < 2; ++i) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(acc_full + 8 * i) : "memory");
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 128;\n" ::"r"(acc_empty + 8 * i) : "memory");
      }
      asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
      asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
  const uint32_t tmem_d = tmem_base_slot;
  pdl_wait();    // x (and a residual) come from the previous kernel of the chain
  pdl_launch();

  if (warp == 0) {
    // ------------------------------------------------------------------ producer
    if (elect_one()) {
      uint32_t it = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int m0 = (tile % MT) * GM, n0 = (tile / MT) * GN;
        for (int kb = 0; kb < KB; ++kb, ++it) {
          const uint32_t s = it % GSTAGES, use = it / GSTAGES;
          if (use > 0) big_wait<true>(empty + 8 * s, (use & 1u) ^ 1u);
          big_expect_tx(full + 8 * s, GSTAGE_BYTES);
          big_tma_2d(sbase + s * GSTAGE_BYTES, &tm_x, full + 8 * s, kb * GK, m0);
          big_tma_2d(sbase + s * GSTAGE_BYTES + GA_BYTES, &tm_w, full + 8 * s, kb * GK, n0);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (elect_one()) {
      uint32_t it = 0, tt = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++tt) {
        const uint32_t buf = tt & 1u, use_acc = tt >> 1;
        if (use_acc > 0) big_wait<true>(acc_empty + 8 * buf, (use_acc & 1u) ^ 1u);  // the epilogue has drained this buffer
        asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
        const uint32_t d = tmem_d + buf * GN;
        for (int kb = 0; kb < KB; ++kb, ++it) {
          const uint32_t s = it % GSTAGES, use = it / GSTAGES;
          big_wait<true>(full + 8 * s, use & 1u);
          asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
          const uint32_t a0 = (((sbase + s * GSTAGE_BYTES) >> 4) & 0x3FFFu) | (1u << 16);
          const uint32_t b0 = (((sbase + s * GSTAGE_BYTES + GA_BYTES) >> 4) & 0x3FFFu) | (1u << 16);
#pragma unroll
          for (int k = 0; k < GK / 16; ++k) big_umma(d, a0 + 2 * k, b0 + 2 * k, (kb | k) ? 1u : 0u);
          big_commit(empty + 8 * s);  // the stage may be refilled once these MMAs have read it
        }
        big_commit(acc_full + 8 * buf);
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------------ epilogue: warp w owns TMEM lanes 32 (w % 4) ..
    const int q = warp & 3;
    uint32_t tt = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++tt) {
      const int m0 = (tile % MT) * GM, n0 = (tile / MT) * GN;
      const uint32_t buf = tt & 1u, use_acc = tt >> 1;
      big_wait<false>(acc_full + 8 * buf, use_acc & 1u);
      asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
      float* strip = reinterpret_cast<float*>(smem_raw + (sbase - smem_u32(smem_raw)) + GSTAGES * GSTAGE_BYTES) +
                     (size_t)(warp - 2) * (GSTRIP_BYTES / 4);
#pragma unroll 1
      for (int c0 = 0; c0 < GN; c0 += 64) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {  // this lane's row, 2 x 32 columns -> strip[lane][..]
          uint32_t v[32];
          big_tmem_ld32(tmem_d + ((uint32_t)(q * 32) << 16) + buf * GN + c0 + 32 * h, v);
          asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
#pragma unroll
          for (int j = 0; j < 32; j += 4)
            *reinterpret_cast<uint4*>(strip + lane * GSTRIP_PITCH + 32 * h + j) = make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]);
        }
        __syncwarp();
        const int col = n0 + c0 + 2 * lane;  // lanes = consecutive column pairs of one row
        if (col < p.N) {
          bool has_res = false;
          if constexpr (EPI == EPI_STORE) has_res = p.residual != nullptr;
#pragma unroll 1
          for (int r0 = 0; r0 < 32; r0 += 16) {
            // the residual pairs of 16 rows are requested together (one L2 round trip instead of one per row)
            float2 res[16];
            if (has_res) {
#pragma unroll
              for (int u = 0; u < 16; ++u) {
                const int row = m0 + q * 32 + r0 + u;
                if (row < p.M) res[u] = __ldcg(reinterpret_cast<const float2*>(p.residual + (int64_t)row * p.N + col));
              }
            }
#pragma unroll
            for (int u = 0; u < 16; ++u) {
              const int row = m0 + q * 32 + r0 + u;
              if (row < p.M) {
                const float2 a = *reinterpret_cast<const float2*>(strip + (r0 + u) * GSTRIP_PITCH + 2 * lane);
                epilogue_pair<EPI>(p, row, col, 0, a.x, a.y, has_res ? &res[u] : nullptr);
              }
            }
          }
        }
        __syncwarp();  // the strip is rewritten by the next 64 columns
      }
      asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
      big_arrive(acc_empty + 8 * buf);
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
  __syncthreads();
  if (warp == 1)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_d), "r"(GTMEM_COLS) : "memory");
}

int g_sm_count = 0;

template <int EPI>
int launch_big(const GemmParams& p, cudaStream_t st) {
  if (p.M < 256 || p.K % GK != 0 || !tma_available()) return PTTS_E_UNSUPPORTED;
  CUtensorMap tm_x, tm_w;
  if (!tensor_map_2d(p.X, p.M, p.K, GM, &tm_x) || !tensor_map_2d(p.W, p.N, p.K, GN, &tm_w)) return PTTS_E_UNSUPPORTED;
  if (g_sm_count == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_sm_count, cudaDevAttrMultiProcessorCount, dev);
    if (g_sm_count <= 0) g_sm_count = 148;
  }
  const int MT = (p.M + GM - 1) / GM, NT = (p.N + GN - 1) / GN;
  const size_t smem = (size_t)GSTAGES * GSTAGE_BYTES + 4 * GSTRIP_BYTES + 1024;
  static bool configured = false;
  if (!configured) {
    cudaFuncSetAttribute(gemm_big_kernel<EPI_STORE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(gemm_big_kernel<EPI_SCATTER>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    configured = true;
  }
  const int grid = MT * NT < g_sm_count ? MT * NT : g_sm_count;
  cudaError_t e = launch_k(gemm_big_kernel<EPI>, dim3(grid), dim3(GTHREADS), smem, st, p, MT, NT, tm_x, tm_w);
  if (e != cudaSuccess) {
    set_error("gemm_big launch: %s", cudaGetErrorString(e));
    (void)cudaGetLastError();
    return PTTS_E_LAUNCH;
  }
  PTTS_CHECK_LAUNCH("gemm_big");
  return PTTS_OK;
}

}  // namespace

namespace gemm {
int big_gemm_store(const GemmParams& p, cudaStream_t st) { return launch_big<EPI_STORE>(p, st); }
int big_gemm_scatter(const GemmParams& p, cudaStream_t st) { return launch_big<EPI_SCATTER>(p, st); }
}  // namespace gemm

}  // namespace ptts
