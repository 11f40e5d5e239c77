// fc1 of the decode step with the activation in the GEMM's own epilogue (round 2): y = act(x . W^T) as bf16, ONE launch
// instead of split-K partial sums + a reduce / activation kernel (nn.Linear + ACT2FN["gelu"],
// /root/reference/parler_tts/modeling_parler_tts.py:488).
//
// Why a kernel of its own: the activation needs the FULL K sum of an output element, and the split-K kernel of
// fast_gemm3.cu splits K exactly because a 128-row weight tile over the whole K (256 KB at K = 1024) does not fit next to
// the activations.  Here a CTA owns 64 weight rows (tcgen05.mma M = 64, N = 64 batch rows) over the whole K:
//   * W slab [64 x K] = K/64 k-blocks of 8 KB (128 KB at K = 1024), two 3-D box copies BEFORE griddepcontrol.wait;
//   * X [64 x K] (another 128 KB) does not fit beside it: it streams through a ring of 12 k-block slots in groups of 4
//     (one 3-D box copy + one mbarrier per group); the group that re-uses the first four slots is requested as soon as the
//     MMAs that read them have completed (tcgen05.commit -> mbarrier);
//   * accumulator D[weight row][batch row] in 64 TMEM columns; M = 64 puts row r in lane 32 * (r / 16) + r % 16
//     (cute's "half subpartitions" atom for UMMA_1SM M_MMA = 64), so every warp reads 16 live lanes;
//   * epilogue on 8 warps: tcgen05.ld, the idle half of each warp takes half of the values by shuffle, 16 activations per
//     thread, bf16 tile transposed through shared memory, 128-byte row stores.
// With more than 64 batch rows a second kernel takes 32 batch rows per CTA (N = 32): the X slab of a CTA is then 64 KB and
// sits beside the W slab without a ring; the CTAs of one weight tile read the same W slab (L2 hits after the first).
// Measured on one box (profiles/r02_fc1_act.txt, call 45): at B = 128 the 32-row kernel gives 2.024 ms per decode step
// against 2.149 ms for two 64-row tiles; at B = 64 it is the slower one (1.2097 against 1.2024 ms), so 64 rows keep the ring.
// (The "pair finisher" of round 1 - the second-arriving K-slice CTA folds both slices - lost because 32 CTAs did the
// activation of a whole launch; here all 64 CTAs do 4096 each.)
#include <cuda.h>
#include <string.h>

#include "gemm_common.cuh"

namespace ptts {

using namespace gemm;

namespace {

constexpr int AM = 64;             // weight rows per CTA (UMMA M)
constexpr int AN = 64;             // batch rows per CTA (UMMA N), ring kernel
constexpr int AN32 = 32;           // batch rows per CTA, ring-less kernel
constexpr int AKB = 64;            // K elements per SWIZZLE_128B k-block
constexpr int ABLK = 64 * 128;     // bytes of one k-block of either operand (64 rows x 128 B)
constexpr int AXBLK32 = AN32 * 128;  // bytes of one X k-block of the ring-less kernel
constexpr int AGRP = 4;            // k-blocks per X box copy / barrier
constexpr int ASLOTS = 12;         // X ring slots (k-blocks)
constexpr int AMAXKB = 16;         // K <= 1024
constexpr int ATH = 256;
constexpr uint32_t ATMEM_COLS = 64;
constexpr uint32_t ATMEM_COLS32 = 32;
// kind::f16 instruction descriptor: D fp32, A/B bf16, both K-major, N = 64, M = 64
constexpr uint32_t kIdescAct = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(AN >> 3) << 17) | ((uint32_t)(AM >> 4) << 24);
constexpr uint32_t kIdescAct32 = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(AN32 >> 3) << 17) | ((uint32_t)(AM >> 4) << 24);

__device__ __forceinline__ void act_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void act_wait(uint32_t bar, uint32_t parity) {
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
__device__ __forceinline__ void act_tma_3d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];\n" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
// tcgen05.mma from the low words of two K-major SWIZZLE_128B descriptors (high word 0x40004040: SBO 1024 B, version 1,
// SWIZZLE_128B - see umma_desc_sw128 in fast_gemm3.cu)
template <uint32_t IDESC = kIdescAct>
__device__ __forceinline__ void act_umma(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "mov.b64 da, {%1, %5};\n\tmov.b64 db, {%2, %5};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %3, p;\n\t}\n" ::"r"(tmem_d),
      "r"(a_lo), "r"(b_lo), "r"(IDESC), "r"(accumulate), "r"(0x40004040u)
      : "memory");
}
__device__ __forceinline__ void act_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void act_tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
               : "r"(taddr)
               : "memory");
}

// grid = (N / 64, M / 64); KBT = K / 64 k-blocks, a multiple of AGRP, <= AMAXKB
// ACT >= 0: the activation is a compile-time constant (one-shot kernel: what is not needed is compiled out)
template <int ACT>
__global__ void __launch_bounds__(ATH)
umma_act_kernel(__nv_bfloat16* __restrict__ y, int N, int KBT, int act_in, const __grid_constant__ CUtensorMap tm_w,
                const __grid_constant__ CUtensorMap tm_x) {
  extern __shared__ unsigned char smem_raw[];
  // [0] W slab, [1] MMA burst, [2] first X group consumed (its slots may be refilled), [3..6] X groups
  __shared__ __align__(8) uint64_t bars[3 + AMAXKB / AGRP];
  __shared__ uint32_t tmem_base_slot;
  const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SWIZZLE_128B atoms need 1024-byte alignment
  unsigned char* smem = smem_raw + (sbase - smem_u32(smem_raw));
  const uint32_t sw = sbase, sx = sbase + KBT * ABLK;
  const int n0 = blockIdx.x * AM, m0 = blockIdx.y * AN;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t w_bar = smem_u32(&bars[0]), mma_bar = smem_u32(&bars[1]), free_bar = smem_u32(&bars[2]);
  const uint32_t x_bar = smem_u32(&bars[3]);
  const int groups = KBT / AGRP;                       // X box copies
  const int ring_groups = ASLOTS / AGRP;               // groups the ring holds
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(&tmem_base_slot)),
                 "r"(ATMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
  } else if (warp == 1) {
    if (elect_one()) {
#pragma unroll
      for (int i = 0; i < 3 + AMAXKB / AGRP; ++i)
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(w_bar + 8 * i) : "memory");
      asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
      asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
      // the whole W slab of this CTA, before the dependency resolves (weights are immutable); tm_w's box = KBT / 2 k-blocks
      act_expect_tx(w_bar, (uint32_t)(KBT * ABLK));
      act_tma_3d(sw, &tm_w, w_bar, 0, n0, 0);
      act_tma_3d(sw + (KBT / 2) * ABLK, &tm_w, w_bar, 0, n0, KBT / 2);
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
  __syncthreads();  // barrier inits and the TMEM base address are visible to everybody - still in front of the wait
  asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
  const uint32_t tmem_d = tmem_base_slot;
  pdl_wait();
  pdl_launch();  // (releasing fc2 only when the MMAs are done was measured: 1.2317 against 1.2194 ms per decode step)
  if (warp == 1 && elect_one()) {
    const int first = groups < ring_groups ? groups : ring_groups;
    for (int g = 0; g < first; ++g) act_expect_tx(x_bar + 8 * g, (uint32_t)(AGRP * ABLK));
    for (int g = 0; g < first; ++g) act_tma_3d(sx + g * AGRP * ABLK, &tm_x, x_bar + 8 * g, 0, m0, g * AGRP);
    act_wait(w_bar, 0);
    const uint32_t a0 = ((sw >> 4) & 0x3FFFu) | (1u << 16), b0 = ((sx >> 4) & 0x3FFFu) | (1u << 16);
    for (int g = 0; g < groups; ++g) {
      act_wait(x_bar + 8 * g, 0);
      asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
      const int slot0 = (g % ring_groups) * AGRP;
#pragma unroll
      for (int j = 0; j < AGRP; ++j)
#pragma unroll
        for (int k = 0; k < AKB / 16; ++k)  // +32 bytes (2 x 16 B) per UMMA_K = 16 inside the swizzle atom
          act_umma(tmem_d, a0 + (g * AGRP + j) * (ABLK >> 4) + 2 * k, b0 + (slot0 + j) * (ABLK >> 4) + 2 * k,
                   (g | j | k) ? 1u : 0u);
      if (g == 0 && groups > ring_groups) {
        // the last group re-uses the first group's slots: request it the moment these MMAs have read them
        act_commit(free_bar);
        act_wait(free_bar, 0);
        act_expect_tx(x_bar + 8 * ring_groups, (uint32_t)(AGRP * ABLK));
        act_tma_3d(sx, &tm_x, x_bar + 8 * ring_groups, 0, m0, ring_groups * AGRP);
      }
    }
    act_commit(mma_bar);
  }
  __syncwarp();
  act_wait(mma_bar, 0);
  asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
  // ---- epilogue.  Warp w reads TMEM lanes [32 q, 32 q + 32), q = w % 4, of which lanes 0..15 hold weight rows
  // 16 q .. 16 q + 15; warps 0-3 take batch columns 0..31, warps 4-7 columns 32..63.
  const int q = warp & 3, c0 = (warp >> 2) * 32;
  uint32_t v[32];
  act_tmem_ld32(tmem_d + ((uint32_t)(q * 32) << 16) + c0, v);
  asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
  // lanes 16..31 hold nothing: each takes the upper 16 columns of lane - 16
  float mine[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    const uint32_t hi = __shfl_sync(0xffffffffu, v[16 + j], lane & 15);
    mine[j] = __uint_as_float(lane < 16 ? v[j] : hi);
  }
  const int act = ACT >= 0 ? ACT : act_in;
  // bf16 tile [batch row][weight row] in shared memory (the operand slabs are dead: every MMA has completed)
  __nv_bfloat16* tile = reinterpret_cast<__nv_bfloat16*>(smem);
  const int nrow = q * 16 + (lane & 15), mbase = c0 + (lane < 16 ? 0 : 16);
#pragma unroll
  for (int j = 0; j < 16; ++j) tile[(mbase + j) * AM + nrow] = __float2bfloat16_rn(apply_act(mine[j], act));
  asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
  __syncthreads();
  if (warp == 0)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_d), "r"(ATMEM_COLS) : "memory");
  // 64 rows x 128 bytes: thread t stores 32 bytes of row t / 4
  const int r = threadIdx.x >> 2, part = threadIdx.x & 3;
  const uint4* src = reinterpret_cast<const uint4*>(tile + r * AM + part * 16);
  uint4* dst = reinterpret_cast<uint4*>(y + (size_t)(m0 + r) * N + n0 + part * 16);
  dst[0] = src[0];
  dst[1] = src[1];
}

__device__ __forceinline__ void act_tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
        "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}

// The ring-less kernel: 32 batch rows per CTA.  grid = (N / 64, M / 32); KBT = K / 64 k-blocks, a multiple of AGRP, <= AMAXKB
// ACT >= 0: the activation is a compile-time constant (one-shot kernel: what is not needed is compiled out)
template <int ACT>
__global__ void __launch_bounds__(ATH)
umma_act32_kernel(__nv_bfloat16* __restrict__ y, int N, int KBT, int act_in, const __grid_constant__ CUtensorMap tm_w,
                const __grid_constant__ CUtensorMap tm_x) {
  extern __shared__ unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t bars[2 + AMAXKB / AGRP];  // [0] W slab, [1] MMA burst, [2..5] X groups
  __shared__ uint32_t tmem_base_slot;
  const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SWIZZLE_128B atoms need 1024-byte alignment
  unsigned char* smem = smem_raw + (sbase - smem_u32(smem_raw));
  const uint32_t sw = sbase, sx = sbase + KBT * ABLK;
  const int n0 = blockIdx.x * AM, m0 = blockIdx.y * AN32;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t w_bar = smem_u32(&bars[0]), mma_bar = smem_u32(&bars[1]), x_bar = smem_u32(&bars[2]);
  const int groups = KBT / AGRP;  // X box copies
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(&tmem_base_slot)),
                 "r"(ATMEM_COLS32)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
  } else if (warp == 1) {
    if (elect_one()) {
#pragma unroll
      for (int i = 0; i < 2 + AMAXKB / AGRP; ++i)
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(w_bar + 8 * i) : "memory");
      asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
      asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
      // the whole W slab of this CTA, before the dependency resolves (weights are immutable); tm_w's box = KBT / 2 k-blocks
      act_expect_tx(w_bar, (uint32_t)(KBT * ABLK));
      act_tma_3d(sw, &tm_w, w_bar, 0, n0, 0);
      act_tma_3d(sw + (KBT / 2) * ABLK, &tm_w, w_bar, 0, n0, KBT / 2);
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
  __syncthreads();  // barrier inits and the TMEM base address are visible to everybody - still in front of the wait
  asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
  const uint32_t tmem_d = tmem_base_slot;
  pdl_wait();
  pdl_launch();
  if (warp == 1 && elect_one()) {
    // all expectations first, then the copies back to back (an arrive between two copies delays the second)
    for (int g = 0; g < groups; ++g) act_expect_tx(x_bar + 8 * g, (uint32_t)(AGRP * AXBLK32));
    for (int g = 0; g < groups; ++g) act_tma_3d(sx + g * AGRP * AXBLK32, &tm_x, x_bar + 8 * g, 0, m0, g * AGRP);
    act_wait(w_bar, 0);
    const uint32_t a0 = ((sw >> 4) & 0x3FFFu) | (1u << 16), b0 = ((sx >> 4) & 0x3FFFu) | (1u << 16);
    for (int g = 0; g < groups; ++g) {
      act_wait(x_bar + 8 * g, 0);
      asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
#pragma unroll
      for (int j = 0; j < AGRP; ++j)
#pragma unroll
        for (int k = 0; k < AKB / 16; ++k)  // +32 bytes (2 x 16 B) per UMMA_K = 16 inside the swizzle atom
          act_umma<kIdescAct32>(tmem_d, a0 + (g * AGRP + j) * (ABLK >> 4) + 2 * k, b0 + (g * AGRP + j) * (AXBLK32 >> 4) + 2 * k,
                   (g | j | k) ? 1u : 0u);
    }
    act_commit(mma_bar);
  }
  __syncwarp();
  act_wait(mma_bar, 0);
  asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
  // ---- epilogue.  Warp w reads TMEM lanes [32 q, 32 q + 32), q = w % 4, of which lanes 0..15 hold weight rows
  // 16 q .. 16 q + 15; warps 0-3 take batch columns 0..15, warps 4-7 columns 16..31.
  const int q = warp & 3, c0 = (warp >> 2) * 16;
  uint32_t v[16];
  act_tmem_ld16(tmem_d + ((uint32_t)(q * 32) << 16) + c0, v);
  asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
  // lanes 16..31 hold nothing: each takes the upper 8 columns of lane - 16
  float mine[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const uint32_t hi = __shfl_sync(0xffffffffu, v[8 + j], lane & 15);
    mine[j] = __uint_as_float(lane < 16 ? v[j] : hi);
  }
  const int act = ACT >= 0 ? ACT : act_in;
  // bf16 tile [batch row][weight row] in shared memory (the operand slabs are dead: every MMA has completed)
  __nv_bfloat16* tile = reinterpret_cast<__nv_bfloat16*>(smem);
  const int nrow = q * 16 + (lane & 15), mbase = c0 + (lane < 16 ? 0 : 8);
#pragma unroll
  for (int j = 0; j < 8; ++j) tile[(mbase + j) * AM + nrow] = __float2bfloat16_rn(apply_act(mine[j], act));
  asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
  __syncthreads();
  if (warp == 0)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_d), "r"(ATMEM_COLS32) : "memory");
  // 32 rows x 128 bytes: thread t stores 16 bytes of row t / 8
  const int r = threadIdx.x >> 3, part = threadIdx.x & 7;
  *reinterpret_cast<uint4*>(y + (size_t)(m0 + r) * N + n0 + part * 8) = *reinterpret_cast<const uint4*>(tile + r * AM + part * 8);
}

}  // namespace

namespace {
// 64 batch rows: one ring CTA per weight tile; anything else in 32-row tiles
bool act_use_ring(int M, int kbt) { return M == AN && (kbt <= ASLOTS || kbt == ASLOTS + AGRP); }
}  // namespace

bool gemm_bf16_act_supported(int M, int N, int K) {
  const int kbt = K / AKB;
  return M > 0 && M % AN32 == 0 && N % AM == 0 && K % AKB == 0 && kbt % AGRP == 0 && kbt <= AMAXKB && gemm::tma_available();
}

// y[M][N] (bf16) = act(x[M][K] . w[N][K]^T)
int gemm_bf16_act(const void* x, const void* w, int M, int N, int K, int act, void* y, cudaStream_t st) {
  if (!gemm_bf16_act_supported(M, N, K)) return PTTS_E_UNSUPPORTED;
  const int kbt = K / AKB;
  const bool ring = act_use_ring(M, kbt);
  const int rows = ring ? AN : AN32;
  CUtensorMap tm_w, tm_x;
  if (!gemm::tensor_map_3d(w, N, K, AM, kbt / 2, &tm_w) || !gemm::tensor_map_3d(x, M, K, rows, AGRP, &tm_x))
    return PTTS_E_UNSUPPORTED;
  const size_t smem = ring ? (size_t)(kbt + (kbt < ASLOTS ? kbt : ASLOTS)) * ABLK + 1024 : (size_t)kbt * (ABLK + AXBLK32) + 1024;
  static size_t configured[2] = {0, 0};
  if (smem > configured[ring]) {
    if (ring) {
      cudaFuncSetAttribute(umma_act_kernel<PTTS_ACT_GELU>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      cudaFuncSetAttribute(umma_act_kernel<-1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    } else {
      cudaFuncSetAttribute(umma_act32_kernel<PTTS_ACT_GELU>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      cudaFuncSetAttribute(umma_act32_kernel<-1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    }
    configured[ring] = smem;
  }
  const dim3 grid(N / AM, M / rows);
  __nv_bfloat16* yo = (__nv_bfloat16*)y;
  cudaError_t e;
  if (ring)
    e = act == PTTS_ACT_GELU ? launch_k(umma_act_kernel<PTTS_ACT_GELU>, grid, dim3(ATH), smem, st, yo, N, kbt, act, tm_w, tm_x)
                             : launch_k(umma_act_kernel<-1>, grid, dim3(ATH), smem, st, yo, N, kbt, act, tm_w, tm_x);
  else
    e = act == PTTS_ACT_GELU ? launch_k(umma_act32_kernel<PTTS_ACT_GELU>, grid, dim3(ATH), smem, st, yo, N, kbt, act, tm_w, tm_x)
                             : launch_k(umma_act32_kernel<-1>, grid, dim3(ATH), smem, st, yo, N, kbt, act, tm_w, tm_x);
  if (e != cudaSuccess) {
    set_error("umma_act launch: %s", cudaGetErrorString(e));
    (void)cudaGetLastError();
    return PTTS_E_LAUNCH;
  }
  PTTS_CHECK_LAUNCH("umma_act");
  return PTTS_OK;
}

}  // namespace ptts
