// Cluster split-K GEMM for the decode step (M = batch rows <= a few hundred): Y = epi(X . W^T), bf16 in, fp32 acc.
//
// Why a second kernel: with one launch per linear and ~270 dependent launches per step, the step time is the SUM of
// per-kernel latencies, so every kernel must be "one memory round trip" long.  Here each CTA owns a 64 x 64 output
// tile over ONE K slice (K / CS columns, CS = cluster size along K, 1..8) and requests its whole W slab and X slab
// at once (no ring): the W slab (HBM) is requested before griddepcontrol.wait, i.e. while the previous kernel of the
// step is still running; the X slab (L2 hits) right after.  The CS partial tiles are then reduced through
// distributed shared memory (mapa + ld.shared::cluster) in fixed rank order (deterministic), each CTA finishing
// 64/CS rows of the tile and running the fused epilogue (STORE / SCATTER of gemm_common.cuh) on them - no partial
// sums in global memory, no separate reduction kernel, small per-CTA activation ingest (64 x K/CS instead of 64 x K).
#include <stdlib.h>

#include "gemm_common.cuh"

namespace ptts {

using namespace gemm;

namespace {

constexpr int BN2 = 64;
constexpr int PITCH = BN2 + 4;  // fp32 pitch of the reduction tile (keeps 16-byte alignment, spreads banks)

__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;\n" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
__device__ __forceinline__ float2 ld_dsmem_f2(uint32_t local_addr, uint32_t rank) {
  uint32_t ra;
  float2 v;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;\n" : "=r"(ra) : "r"(local_addr), "r"(rank));
  asm volatile("ld.shared::cluster.v2.f32 {%0,%1}, [%2];\n" : "=f"(v.x), "=f"(v.y) : "r"(ra) : "memory");
  return v;
}

__device__ __forceinline__ void wait_groups_pending(int n) {  // cp.async.wait_group needs an immediate
  switch (n) {
    case 0: cp_async_wait<0>(); break;
    case 1: cp_async_wait<1>(); break;
    case 2: cp_async_wait<2>(); break;
    case 3: cp_async_wait<3>(); break;
    case 4: cp_async_wait<4>(); break;
    default: cp_async_wait<5>(); break;
  }
}

// grid = (N/64, M/64, CS), cluster = (1, 1, CS); KB = k blocks of 128 per CTA (<= 6)
template <int EPI>
__global__ void __launch_bounds__(NTH) cluster_gemm_kernel(const GemmParams p, int KB, int CS) {
  extern __shared__ __align__(128) unsigned char smem[];
  constexpr int XB = BM * 256, WB = BN2 * 256;
  const int n0 = blockIdx.x * BN2, m0 = blockIdx.y * BM;
  const uint32_t rank = cluster_rank();
  const int kbeg = (int)rank * KB * BK;
  const uint32_t sx = smem_u32(smem), sw = sx + KB * XB;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  for (int kb = 0; kb < KB; ++kb) load_tile<BN2>(sw + kb * WB, p.W, p.K, n0, p.N, kbeg + kb * BK);
  pdl_wait();
  pdl_launch();
  for (int kb = 0; kb < KB; ++kb) {
    load_tile<BM>(sx + kb * XB, p.X, p.K, m0, p.M, kbeg + kb * BK);
    cp_async_commit();
  }

  float acc[BN2 / 8][4];
#pragma unroll
  for (int i = 0; i < BN2 / 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int kb = 0; kb < KB; ++kb) {
    wait_groups_pending(KB - 1 - kb);
    __syncthreads();
    const uint32_t xs = sx + kb * XB, ws = sw + kb * WB;
#pragma unroll
    for (int ks = 0; ks < BK / 16; ++ks) {
      uint32_t a[4];
      {
        const int r = warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
        const int c = ks * 2 + (lane >> 4);
        ldmatrix_x4(a, xs + r * 256 + ((c ^ (r & 7)) << 4));
      }
#pragma unroll
      for (int np = 0; np < BN2 / 16; ++np) {
        uint32_t b[4];
        const int r = np * 16 + (lane & 7) + (lane >> 4) * 8;
        const int c = ks * 2 + ((lane >> 3) & 1);
        ldmatrix_x4(b, ws + r * 256 + ((c ^ (r & 7)) << 4));
        mma_bf16(acc[2 * np], a, b[0], b[1]);
        mma_bf16(acc[2 * np + 1], a, b[2], b[3]);
      }
    }
  }
  __syncthreads();  // every warp is done with the operand slabs: reuse them as the fp32 reduction tile

  float* red = reinterpret_cast<float*>(smem);
  {
    const int r0 = warp * 16 + (lane >> 2);
#pragma unroll
    for (int nt = 0; nt < BN2 / 8; ++nt) {
      const int c = nt * 8 + (lane & 3) * 2;
      *reinterpret_cast<float2*>(red + r0 * PITCH + c) = make_float2(acc[nt][0], acc[nt][1]);
      *reinterpret_cast<float2*>(red + (r0 + 8) * PITCH + c) = make_float2(acc[nt][2], acc[nt][3]);
    }
  }
  cluster_sync_all();  // all CS partial tiles are visible cluster-wide

  const int rows_per = BM / CS;
  const int pairs = rows_per * (BN2 / 2);
  for (int e = threadIdx.x; e < pairs; e += NTH) {
    const int lr = (int)rank * rows_per + e / (BN2 / 2);
    const int lc = (e % (BN2 / 2)) * 2;
    const uint32_t la = smem_u32(red + lr * PITCH + lc);
    float2 v[8];
#pragma unroll
    for (int r = 0; r < 8; ++r)
      if (r < CS) v[r] = ld_dsmem_f2(la, (uint32_t)r);
    float v0 = 0.f, v1 = 0.f;
#pragma unroll
    for (int r = 0; r < 8; ++r)
      if (r < CS) {
        v0 += v[r].x;
        v1 += v[r].y;
      }
    const int row = m0 + lr, col = n0 + lc;
    if (row < p.M && col < p.N) epilogue_pair<EPI>(p, row, col, 0, v0, v1);
  }
  cluster_sync_all();  // nobody exits while a peer may still read its tile
}

constexpr int kTarget = 120;
constexpr int kMaxKB = 6;

// cluster size along K for an (M, N, K) problem, or 0 if this kernel does not apply
int plan_cluster(int M, int N, int K) {
  if (K % BK != 0) return 0;
  const int kblocks = K / BK;
  const int nt = ((N + BN2 - 1) / BN2) * ((M + BM - 1) / BM);
  static int forced = -1;  // PTTS_CS=<1|2|4|8>: force the cluster size where it fits (sweeps in profiles/scripts)
  if (forced < 0) {
    const char* e = getenv("PTTS_CS");
    forced = e ? atoi(e) : 0;
  }
  if (forced > 0 && kblocks % forced == 0 && kblocks / forced <= kMaxKB) return forced;
  for (int cs = 1; cs <= 8; cs <<= 1) {
    if (kblocks % cs) continue;
    const int kb = kblocks / cs;
    if (kb > kMaxKB) continue;
    const bool last = cs == 8 || (kblocks % (cs * 2)) != 0;  // no larger cluster divides K
    // enough CTAs to pull weights on every SM; slabs > 64 KB (1 CTA per SM) only if that is still a single wave
    if ((nt * cs >= kTarget && (kb <= 2 || nt * cs <= 148)) || last) return cs;
  }
  return 0;
}

template <int EPI>
int launch_cluster(const GemmParams& p, cudaStream_t st) {
  const int cs = plan_cluster(p.M, p.N, p.K);
  if (cs == 0) return PTTS_E_UNSUPPORTED;
  const int kb = p.K / BK / cs;
  size_t smem = (size_t)kb * (BM * 256 + BN2 * 256);
  const size_t red = (size_t)BM * PITCH * sizeof(float);
  if (smem < red) smem = red;
  static size_t configured = 0;
  if (smem > configured) {
    cudaFuncSetAttribute(cluster_gemm_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    configured = smem;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((p.N + BN2 - 1) / BN2, (p.M + BM - 1) / BM, cs);
  cfg.blockDim = dim3(NTH);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  int n = 0;
  attr[n].id = cudaLaunchAttributeClusterDimension;
  attr[n].val.clusterDim.x = 1;
  attr[n].val.clusterDim.y = 1;
  attr[n].val.clusterDim.z = cs;
  ++n;
  if (pdl_enabled()) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  cudaError_t e = cudaLaunchKernelEx(&cfg, cluster_gemm_kernel<EPI>, p, kb, cs);
  if (e != cudaSuccess) {
    set_error("cluster_gemm launch: %s", cudaGetErrorString(e));
    (void)cudaGetLastError();
    return PTTS_E_LAUNCH;
  }
  PTTS_CHECK_LAUNCH("cluster_gemm");
  return PTTS_OK;
}

}  // namespace

namespace gemm {
int cluster_gemm_store(const GemmParams& p, cudaStream_t st) { return launch_cluster<EPI_STORE>(p, st); }
int cluster_gemm_scatter(const GemmParams& p, cudaStream_t st) { return launch_cluster<EPI_SCATTER>(p, st); }
}  // namespace gemm

}  // namespace ptts
