// Internal (non-ABI) declarations shared by the translation units of libparler_b200.so.
#pragma once
#include "common.cuh"

namespace ptts {

int embed_prefill(const PttsModel* m, const int64_t* ids, int64_t ld_ids, const int64_t* mask, int64_t ld_mask, int B,
                  int ids_batch, int T0, const float* prompt_table,
                  const int64_t* prompt_ids, const float* prompt_hidden, int P, float* h, cudaStream_t st, int pos0 = 0);
int layernorm(const float* x, const float* g, const float* b, int rows, int H, float eps, int64_t in_stride,
              int64_t in_offset, void* y, int ydtype, cudaStream_t st);
int linear_simple(const void* x, int xdtype, const void* w, int wdtype, const float* bias, int M, int N, int K,
                  float alpha, int act, const float* residual, const int* row_mask, void* y, int ydtype,
                  cudaStream_t st);
int qkv_scatter(const void* qkv, int dtype, int rows, int H, int heads, int d, int Tq, int c_stride, int pos_base,
                const int* pos_dyn, float scale, void* q, void* ck, void* cv, cudaStream_t st);
int kv_scatter(const void* kv, int dtype, int rows, int H, int heads, int d, int S, void* ck, void* cv,
               cudaStream_t st);
int attention_simple(const void* q, const void* ck, const void* cv, int dtype, int B, int heads, int d, int Tq,
                     int c_stride, int ctx_base, const int* ctx_dyn, int causal, const int* key_mask, int mask_len,
                     void* out, cudaStream_t st);
// logit_parts (optional): split-K partial sums [nslice][B*K][V] of the heads GEMM instead of s->logits
int step_epilogue(const PttsModel* m, const PttsState* s, cudaStream_t st, const float* logit_parts = nullptr,
                  int nslice = 1, int64_t slice_stride = 0);

// fast family -- fast_*.cu
constexpr int kMaxSplitK = 16;
bool gemm_bf16_supported(int M, int N, int K);
int gemm_bf16_store(const void* x, const void* w, int M, int N, int K, float alpha, int act, const float* bias,
                    const float* residual, const int* row_mask, void* y, int ydtype, cudaStream_t st);
int gemm_bf16_scatter(const void* x, const void* w, int M, int N, int K, int has_q, float q_scale, void* q, void* ck,
                      void* cv, int H, int heads, int d, int Tq, int c_stride, int pos_base, const int* pos_dyn,
                      cudaStream_t st);
// y (bf16) = act(x . w^T) over the whole K in one launch, activation in the GEMM epilogue (fast_gemm_act.cu; decode sizes)
bool gemm_bf16_act_supported(int M, int N, int K);
int gemm_bf16_act(const void* x, const void* w, int M, int N, int K, int act, void* y, cudaStream_t st);
int gemm_bf16_plan_splits(int M, int N, int K, int* bn_out);
int gemm_bf16_partial(const void* x, const void* w, int M, int N, int K, int splits, int bn, float* part,
                      cudaStream_t st);
// the next gemm_bf16_partial launch releases its dependent kernel AFTER its own dependency wait whatever PTTS_PDL_EARLY
// says, or (at_mma_done) only when its MMA burst has completed
void gemm_next_trigger_late(int at_mma_done);
// prefill attention (Tq > 1) with the pair's K / V staged in shared memory (fast_attn_prefill.cu)
bool attention_prefill_supported(int dtype, int d, int Tq, int ctx_max);
int attention_prefill(const void* q, const void* ck, const void* cv, int dtype, int B, int heads, int Tq, int c_stride,
                      int ctx_base, const int* ctx_dyn, int ctx_max, int causal, const int* key_mask, int mask_len,
                      void* out, cudaStream_t st);
bool fast_layernorm_supported(int H);
// y = act(sum of S split-K partial slices) as bf16 (fast_misc.cu)
int partial_act_fast(const float* part, int S, int64_t part_stride, int rows, int N, int act, void* y,
                     cudaStream_t st);
int layernorm_fast(float* h, const float* part, int S, int64_t part_stride, const float* g, const float* b, int rows,
                   int H, float eps, int row_mul, int row_add, void* y, int ydtype, cudaStream_t st,
                   const float* rowblk = nullptr, int n_blk = 0);
int64_t attn_scratch_floats(int B, int heads);

bool fast_attention_supported(int dtype, int d, int Tq);
int attention_fast(const void* q, const void* ck, const void* cv, int dtype, int B, int heads, int d, int Tq,
                   int c_stride, int ctx_base, const int* ctx_dyn, int causal, const int* key_mask, int mask_len,
                   void* out, float* scratch, cudaStream_t st);
// warp-specialised persistent tcgen05 conv over pre-tiled weights (fast_dac3.cu)
struct DacConv {
  const float* x;                  // [B][T][Cin] fp32 (Snake `alpha` applied while staging), or null
  const __nv_bfloat16* x_planes;   // or: activated hi/mid planes [B][2][Cin/8][T + pads][8]
  const __nv_bfloat16* w;          // pre-tiled stages
  const float* bias;
  const float* alpha;
  const float* residual;           // [B][Tout][Cout] or null
  float* y;                        // [B][Tout][Cout] fp32, or null
  __nv_bfloat16* y_planes;         // or: Snake(out_alpha) + hi/mid planes [B][2][Cout/8][Tout + pads][8]
  const float* out_alpha;
  int T, Cin, N, Cout, taps, tap_center, tap_step, lo, span;
  int Mrows, out_stride, out_pad, Tout;
};
int dac_tile_n(int N);
int dac_set_cta_limit(int n);  // persistent conv grids; returns the previous limit (0 = one CTA per SM)
int dac_conv_ws(const DacConv& conv, int B, cudaStream_t st);
int attention_fast_fused(const float* qkv_part, int S, int64_t part_stride, int ncols, int has_kv, float scale,
                         int slot_base, const int* slot_dyn, void* ck, void* cv, int B, int heads, int c_stride,
                         int ctx_base, const int* ctx_dyn, const int* key_mask, int mask_len, void* out, float* scratch,
                         const void* pf_next, int pf_stride, int pf_rows_base, const int* pf_rows_dyn,
                         cudaStream_t st);
// warp-specialised, TMA-staged persistent flash-decode on tcgen05 (fast_attn_ws.cu): bf16 or fp8 cache
bool attention_ws_supported(int B, int heads, int d, int c_stride);
int attention_ws_fused(const float* qkv_part, int S, int64_t part_stride, int ncols, float scale, int slot_base,
                       const int* slot_dyn, void* ck, void* cv, int kv_dtype, const float* k_scale, const float* v_scale,
                       int B, int heads, int c_stride, const int* key_mask, int mask_len, void* out, cudaStream_t st);
// the cross-attention block of a decode step in one launch (fast_xblock.cu)
bool xblock_supported(int H, int heads, int d, int S);
int xblock_fast(const float* h, const float* part_in, int S_in, int64_t stride, const float* gamma, const float* beta,
                float eps, const void* wq, const void* wo, const void* kc, const void* vc, const int* key_mask, int B,
                int heads, int H, int S, float scale, float* part_out, int out_by_row, const void* pf_next,
                int pf_stride, int pf_rows_base, const int* pf_rows_dyn, cudaStream_t st);
// fp8 (e4m3) self-attention cache (fast_attn_fp8.cu)
bool attention_fp8_supported(int d);
int attention_fp8_fused(const float* qkv_part, int S, int64_t part_stride, int ncols, float scale, int slot_base,
                        const int* slot_dyn, void* ck, void* cv, const float* k_scale, const float* v_scale, int B,
                        int heads, int c_stride, int ctx_base, const int* ctx_dyn, const int* key_mask, int mask_len,
                        void* out, float* scratch, cudaStream_t st);
int attention_fp8(const void* q, const void* ck, const void* cv, const float* k_scale, const float* v_scale, int B,
                  int heads, int c_stride, int ctx_base, const int* ctx_dyn, const int* key_mask, int mask_len,
                  void* out, float* scratch, cudaStream_t st);
int kv_quantize_fp8(const void* stage, int n_lk, int B, int heads, int slots, int c_max, unsigned int* amax_ws,
                    float* scale, void* cache, cudaStream_t st);

}  // namespace ptts
