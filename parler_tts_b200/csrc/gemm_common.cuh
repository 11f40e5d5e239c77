// Shared pieces of the bf16 GEMM kernels of the fast family (fast_gemm.cu: pipelined mma.sync kernel for the shapes the
// tcgen05 kernel of fast_gemm3.cu does not take): parameter block, cp.async / ldmatrix / mma.sync wrappers, swizzled tile loader, fused epilogues.
#pragma once
#include <cuda.h>

#include "ops.cuh"

namespace ptts {
namespace gemm {

constexpr int BM = 64, BK = 128, NTH = 128;
enum { EPI_PARTIAL = 0, EPI_STORE = 1, EPI_SCATTER = 2 };

struct GemmParams {
  const __nv_bfloat16* X;
  const __nv_bfloat16* W;
  int M, N, K, k_per_split;
  // STORE
  float alpha;
  int act;
  const float* bias;
  const float* residual;
  const int* row_mask;
  void* y;
  int ydtype;
  // SCATTER
  __nv_bfloat16 *q, *ck, *cv;
  int H, heads, d, Tq, c_stride, pos_base, has_q;
  const int* pos_dyn;
  float q_scale;
  // PARTIAL
  float* part;
  // (Two in-kernel finishers of the PARTIAL variant were measured and removed: a grid-wide arrival counter after which
  // every CTA folds + LayerNorms its share of the rows - 1.99 against 1.78 ms per decode step - and a "pair finisher" in
  // which the second-arriving K-slice CTA of an fc1 tile folds both slices and applies the activation - 1.539 against
  // 1.457 ms.  Both lose to the separate small kernels that spread the same work over all SMs.)
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
  const int sz = valid ? 16 : 0;  // src-size 0: the 16 destination bytes are zero-filled, nothing is read
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory");
}

__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x2(uint32_t (&r)[2], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.shared.b16 {%0,%1}, [%2];\n" : "=r"(r[0]), "=r"(r[1]) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// ROWS x 128 bf16 tile (256-byte rows, 16-byte chunk c of row r stored at chunk position c ^ (r & 7))
template <int ROWS>
__device__ __forceinline__ void load_tile(uint32_t sbase, const __nv_bfloat16* g, int ld, int row0, int rows_total,
                                          int k0) {
#pragma unroll
  for (int i = threadIdx.x; i < ROWS * 16; i += NTH) {
    const int r = i >> 4, c = i & 15;
    const int gr = row0 + r;
    const bool ok = gr < rows_total;
    const __nv_bfloat16* src = ok ? g + (int64_t)gr * ld + k0 + c * 8 : g;
    cp_async16(sbase + r * 256 + ((c ^ (r & 7)) << 4), src, ok);
  }
}

// Cached tensor map (fast_gemm3.cu) of a bf16 row-major matrix [rows][cols]: box = 64 columns (128 bytes, SWIZZLE_128B)
// x box_rows rows, out-of-range rows read as 0.  false: no driver entry point / misaligned matrix.
bool tensor_map_2d(const void* ptr, int rows, int cols, int box_rows, CUtensorMap* out);
// same for PTTS_BF16 (128-byte box rows, SWIZZLE_128B) or PTTS_FP8 bytes (64-byte box rows, SWIZZLE_64B)
bool tensor_map_2d_dt(const void* ptr, int rows, int cols, int box_rows, int dtype, CUtensorMap* out);
// bf16 [rows][cols] viewed as [cols/64][rows][64]: box {64, box_rows, box_kb} = box_kb consecutive SWIZZLE_128B k-blocks
bool tensor_map_3d(const void* ptr, int rows, int cols, int box_rows, int box_kb, CUtensorMap* out);
bool tma_available();  // the driver entry point for tensor-map creation was found

// plain 128 x 128 tile kernel for M >= 256 (fast_gemm_big.cu: prefill, cross K/V, T5)
int big_gemm_store(const GemmParams& p, cudaStream_t st);
int big_gemm_scatter(const GemmParams& p, cudaStream_t st);
// tcgen05 swap-AB cluster split-K kernel (fast_gemm3.cu)
int umma_gemm_store(const GemmParams& p, cudaStream_t st);
int umma_gemm_scatter(const GemmParams& p, cudaStream_t st);
int umma_plan_splits(int M, int N, int K);
int umma_gemm_partial(const GemmParams& p, int splits, cudaStream_t st);
int umma_read_stamps(unsigned long long* host, int n_launches);

// Fused epilogue for the element pair (row, col..col+1); z = split index (PARTIAL only).
// `pre` (optional): the residual pair already loaded by the caller (hides the L2 round trip behind other work).
template <int EPI>
__device__ __forceinline__ void epilogue_pair(const GemmParams& p, int row, int col, int z, float v0, float v1,
                                              const float2* pre = nullptr) {
  if constexpr (EPI == EPI_PARTIAL) {
    *reinterpret_cast<float2*>(p.part + ((int64_t)z * p.M + row) * p.N + col) = make_float2(v0, v1);
  } else if constexpr (EPI == EPI_STORE) {
    v0 *= p.alpha;
    v1 *= p.alpha;
    if (p.bias) {
      v0 += p.bias[col];
      v1 += p.bias[col + 1];
    }
    v0 = apply_act(v0, p.act);
    v1 = apply_act(v1, p.act);
    if (p.row_mask) {
      const float rm = (float)p.row_mask[row];
      v0 *= rm;
      v1 *= rm;
    }
    const int64_t o = (int64_t)row * p.N + col;
    if (p.residual) {
      const float2 r = pre ? *pre : __ldcg(reinterpret_cast<const float2*>(p.residual + o));
      v0 += r.x;
      v1 += r.y;
    }
    if (p.ydtype == PTTS_F32)
      *reinterpret_cast<float2*>((float*)p.y + o) = make_float2(v0, v1);
    else
      *reinterpret_cast<__nv_bfloat162*>((__nv_bfloat16*)p.y + o) = __floats2bfloat162_rn(v0, v1);
  } else {  // EPI_SCATTER: columns are [q | k | v] (has_q) or [k | v], each H wide
    const int which = col / p.H + (p.has_q ? 0 : 1);
    const int c = col % p.H;
    if (which == 0) {
      *reinterpret_cast<__nv_bfloat162*>(p.q + (int64_t)row * p.H + c) =
          __floats2bfloat162_rn(v0 * p.q_scale, v1 * p.q_scale);
    } else {
      const int b = row / p.Tq, t = row % p.Tq;
      const int slot = dyn_value(p.pos_base, p.pos_dyn) + t;
      const int hd = c / p.d, dd = c % p.d;
      __nv_bfloat16* dst = which == 1 ? p.ck : p.cv;
      *reinterpret_cast<__nv_bfloat162*>(dst + (((int64_t)b * p.heads + hd) * p.c_stride + slot) * p.d + dd) =
          __floats2bfloat162_rn(v0, v1);
    }
  }
}

}  // namespace gemm
}  // namespace ptts
