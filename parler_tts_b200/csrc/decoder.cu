// Orchestration of the decoder forward passes (prefill, one cached decode step) and the exported building blocks.
// Every function only enqueues kernels on the caller's stream: no allocation, no synchronisation, graph-capturable.
// Replaces /root/reference/parler_tts/modeling_parler_tts.py:794-978 (ParlerTTSDecoder.forward), 415-502 (layer),
// 264-382 (attention), 1096-1185 (lm heads) for the generation path.
#include <stdlib.h>

#include "gemm_common.cuh"
#include "ops.cuh"

namespace ptts {

static int linear_dispatch(int impl, const void* x, int xdtype, const void* w, int wdtype, const float* bias, int M,
                           int N, int K, float alpha, int act, const float* residual, const int* row_mask, void* y,
                           int ydtype, cudaStream_t st) {
  if (impl == PTTS_IMPL_FAST && xdtype == PTTS_BF16 && wdtype == PTTS_BF16 && gemm_bf16_supported(M, N, K)) {
    // decode-sized activation GEMM (fc1): full-K tiles with the activation in the epilogue (fast_gemm_act.cu)
    if (act != PTTS_ACT_NONE && !bias && !residual && !row_mask && alpha == 1.0f && ydtype == PTTS_BF16 && M <= 128 &&
        gemm_bf16_act_supported(M, N, K)) {
      const int rc = gemm_bf16_act(x, w, M, N, K, act, y, st);
      if (rc != PTTS_E_UNSUPPORTED) return rc;
    }
    return gemm_bf16_store(x, w, M, N, K, alpha, act, bias, residual, row_mask, y, ydtype, st);
  }
  return linear_simple(x, xdtype, w, wdtype, bias, M, N, K, alpha, act, residual, row_mask, y, ydtype, st);
}

static int attention_dispatch(int impl, const void* q, const void* ck, const void* cv, int dtype, int B, int heads,
                              int d, int Tq, int c_stride, int ctx_base, const int* ctx_dyn, int causal,
                              const int* key_mask, int mask_len, void* out, float* scratch, cudaStream_t st) {
  if (impl == PTTS_IMPL_FAST && fast_attention_supported(dtype, d, Tq))
    return attention_fast(q, ck, cv, dtype, B, heads, d, Tq, c_stride, ctx_base, ctx_dyn, causal, key_mask, mask_len,
                          out, scratch, st);
  // prefill (Tq > 1): K / V of a (head, utterance) pair staged in shared memory once for all its queries
  const int ctx_max = ctx_dyn ? c_stride : ctx_base;
  if (impl == PTTS_IMPL_FAST && attention_prefill_supported(dtype, d, Tq, ctx_max))
    return attention_prefill(q, ck, cv, dtype, B, heads, Tq, c_stride, ctx_base, ctx_dyn, ctx_max, causal, key_mask,
                             mask_len, out, st);
  return attention_simple(q, ck, cv, dtype, B, heads, d, Tq, c_stride, ctx_base, ctx_dyn, causal, key_mask, mask_len,
                          out, st);
}

static int check_model(const PttsModel* m) {
  PTTS_CHECK_ARG(m && m->layers, "null model descriptor");
  PTTS_CHECK_ARG(m->n_layers > 0 && m->hidden > 0 && m->n_heads > 0 && m->head_dim * m->n_heads == m->hidden,
                 "bad model dims");
  PTTS_CHECK_ARG(m->n_codebooks > 0 && m->n_codebooks <= PTTS_MAX_CODEBOOKS, "bad codebook count");
  PTTS_CHECK_ARG(m->wdtype == PTTS_F32 || m->wdtype == PTTS_BF16, "bad wdtype");
  PTTS_CHECK_ARG(m->embed && m->pos && m->w_heads && m->lnf_g && m->lnf_b, "null model tensor");
  return PTTS_OK;
}

static int check_state(const PttsModel* m, const PttsState* s) {
  PTTS_CHECK_ARG(s, "null state descriptor");
  PTTS_CHECK_ARG(s->B > 0 && s->P >= 0 && s->T0 > 0 && s->S >= 0, "bad state dims");
  PTTS_CHECK_ARG(s->c_max >= s->P + s->max_len, "c_max smaller than P + max_len");
  PTTS_CHECK_ARG(s->P + s->max_len <= m->max_pos, "P + max_len exceeds max_position_embeddings");
  PTTS_CHECK_ARG(s->ctrl && s->kv_self && s->h && s->xn && s->q && s->attn && s->ffn && s->logits && s->scratch &&
                     s->ids && s->mask && s->unfinished,
                 "null state tensor");
  PTTS_CHECK_ARG(s->S == 0 || s->kv_cross, "null cross cache");
  if (s->kv_dtype == PTTS_FP8) {
    PTTS_CHECK_ARG(m->wdtype == PTTS_BF16 && s->impl == PTTS_IMPL_FAST && attention_fp8_supported(m->head_dim),
                   "fp8 KV cache needs bf16 weights, the fast kernel family and head_dim 64");
    PTTS_CHECK_ARG(s->kv_scale && s->kv_stage, "fp8 KV cache: null kv_scale / kv_stage");
  } else {
    PTTS_CHECK_ARG(s->kv_dtype == m->wdtype, "kv_dtype must be the weight dtype or PTTS_FP8");
  }
  return PTTS_OK;
}

static inline size_t kv_layer_elems(const PttsModel* m, int B, int slots) {
  return (size_t)B * m->n_heads * slots * m->head_dim;
}

// Self-attention cache of layer l as the kernels of this call see it.  With an fp8 cache the PREFILL (Tq > 1) runs on
// the bf16 staging cache of P + T0 slots (kv_stage) and ptts_prefill quantises it afterwards; decode steps see the
// e4m3 cache and the per-(layer, K|V, head) scales.
struct KvView {
  void* k;
  void* v;
  const float* k_scale;
  const float* v_scale;
  int c_stride;
  int dtype;
};
static KvView self_kv(const PttsModel* m, const PttsState* s, int l, int Tq) {
  KvView kv = {};
  if (s->kv_dtype == PTTS_FP8 && Tq == 1) {
    const size_t le = kv_layer_elems(m, s->B, s->c_max);
    kv.k = (char*)s->kv_self + (size_t)(2 * l) * le;
    kv.v = (char*)s->kv_self + (size_t)(2 * l + 1) * le;
    kv.k_scale = s->kv_scale + (size_t)(2 * l) * m->n_heads;
    kv.v_scale = s->kv_scale + (size_t)(2 * l + 1) * m->n_heads;
    kv.c_stride = s->c_max;
    kv.dtype = PTTS_FP8;
    return kv;
  }
  const bool staged = s->kv_dtype == PTTS_FP8;
  const int slots = staged ? s->P + s->T0 : s->c_max;
  const size_t es = dtype_size(m->wdtype), le = kv_layer_elems(m, s->B, slots);
  char* base = (char*)(staged ? s->kv_stage : s->kv_self);
  kv.k = base + (size_t)(2 * l) * le * es;
  kv.v = base + (size_t)(2 * l + 1) * le * es;
  kv.c_stride = slots;
  kv.dtype = m->wdtype;
  return kv;
}

// Scratch layout: [0, 64) grid-barrier counters of the fused GEMM finisher (zero between launches; the caller hands
// over zero-initialised scratch) | split-KV attention partials | work area (qkv staging of the simple family,
// all-position logits staging, split-K partial sums of the residual GEMMs).
constexpr int kScratchHeader = 64;
static inline float* attn_area(const PttsState* s) { return s->scratch + kScratchHeader; }
static inline float* work_area(const PttsModel* m, const PttsState* s) {
  return attn_area(s) + attn_scratch_floats(s->B, m->n_heads);
}

// floats the caller-provided work area is guaranteed to hold for `rows` token rows (see ptts_scratch_floats)
static inline int64_t work_area_floats(const PttsModel* m, int rows) {
  return (int64_t)kMaxSplitK * (rows < 256 ? rows : 256) * 3 * m->hidden;
}

// Split-K partial sums of a residual GEMM that the next LayerNorm still has to fold into the residual stream.
struct Pending {
  const float* part = nullptr;
  int S = 0;
  int64_t stride = 0;
  const float* rowblk = nullptr;   // further slices stored row-contiguously [rows][n_blk][H] (cross-attention block)
  int n_blk = 0;
};

struct Path {
  bool gemm;  // bf16 tensor-core weight-streaming GEMMs with fused epilogues
  bool ln;    // LayerNorm with the split-K reduction folded in
};

static Path pick_path(const PttsModel* m, const PttsState* s) {
  Path p;
  const bool fast = s->impl == PTTS_IMPL_FAST;
  p.gemm = fast && m->wdtype == PTTS_BF16 && gemm_bf16_supported(1, m->hidden, m->hidden) &&
           gemm_bf16_supported(1, m->hidden, m->ffn);
  p.ln = fast && fast_layernorm_supported(m->hidden);
  return p;
}

// y = LN(h [+ pending partial sums]) for rows_out rows; output row r reads residual row r*row_mul + row_add
static int ln_step(const PttsModel* m, const PttsState* s, const Path& path, Pending& pend, const float* g,
                   const float* b, int rows_out, int row_mul, int row_add, cudaStream_t st) {
  const int H = m->hidden;
  if (path.ln) {
    PTTS_CHECK_ARG(!pend.rowblk || (row_mul == 1 && row_add == 0), "internal: row block with a row map");
    PTTS_TRY(layernorm_fast(s->h, pend.part, pend.S, pend.stride, g, b, rows_out, H, 1e-5f, row_mul, row_add, s->xn,
                            m->wdtype, st, pend.rowblk, pend.n_blk));
    pend = Pending();
    return PTTS_OK;
  }
  PTTS_CHECK_ARG(pend.S == 0, "internal: pending split-K partials without the fast LayerNorm");
  return layernorm(s->h, g, b, rows_out, H, 1e-5f, (int64_t)row_mul * H, (int64_t)row_add * H, s->xn, m->wdtype, st);
}

// Residual GEMMs (out_proj, fc2) write split-K partial sums that the next LayerNorm folds into the residual stream in
// slice order (deterministic, no on-chip reduction); the diagnostic build can fall back to the STORE epilogue with
// PTTS_SPLITK_PARTIALS=0.
static bool split_k_partials_enabled() { return PTTS_KNOB("PTTS_SPLITK_PARTIALS", 1) != 0; }

// diagnostic build, PTTS_QKV_FUSE=0: the decode-step QKV / cross-q projections keep the scatter epilogue
static bool qkv_fusion_enabled() { return PTTS_KNOB("PTTS_QKV_FUSE", 1) != 0; }

// diagnostic build, PTTS_FC1_PARTIALS=0: the decode-step fc1 keeps the STORE epilogue with the fused activation
static bool fc1_partials_enabled() { return PTTS_KNOB("PTTS_FC1_PARTIALS", 1) != 0; }

// PTTS_SKIP=<bitmask>: timing ablation ONLY (results are garbage): 1 LayerNorm, 2 QKV GEMM, 4 self-attention,
// 8 residual GEMMs, 16 cross q GEMM, 32 cross-attention, 64 fc1, 128 heads, 256 the one-launch cross-attention block.  Used by profiles/scripts/ablate.sh to get
// the marginal cost of each kernel class inside the PDL-overlapped graph (ncu serialises launches and cannot).
// Compiled into the DIAGNOSTIC library only (make DIAG=1): the product library always runs every kernel.
#ifdef PTTS_DIAG
static int skip_mask() {
  static const int v = PTTS_KNOB("PTTS_SKIP", 0);
  return v;
}
#define PTTS_UNLESS_SKIPPED(bit, expr) \
  do {                                  \
    if (!(skip_mask() & (bit))) PTTS_TRY(expr); \
  } while (0)
#else
static constexpr int skip_mask() { return 0; }
#define PTTS_UNLESS_SKIPPED(bit, expr) PTTS_TRY(expr)
#endif

// h += x . w^T   (out_proj / fc2 with the residual add of modeling_parler_tts.py:460,483,492)
static int residual_linear(const PttsModel* m, const PttsState* s, const Path& path, Pending& pend, const void* x,
                           const void* w, int rows, int K, cudaStream_t st) {
  const int H = m->hidden, dt = m->wdtype;
  if (path.gemm && path.ln && rows <= 256 && split_k_partials_enabled()) {
    int bn = 64;
    const int S = gemm_bf16_plan_splits(rows, H, K, &bn);
    if (S > 1) {
      float* part = work_area(m, s);
      PTTS_TRY(gemm_bf16_partial(x, w, rows, H, K, S, bn, part, st));
      pend.part = part;
      pend.S = S;
      pend.stride = (int64_t)rows * H;
      return PTTS_OK;
    }
  }
  return linear_dispatch(s->impl, x, dt, w, dt, nullptr, rows, H, K, 1.0f, PTTS_ACT_NONE, s->h, nullptr, s->h,
                         PTTS_F32, st);
}

// one decoder layer over `rows` = B*Tq tokens.  ctx_base/ctx_dyn give the self-attention context length AFTER the
// new K/V have been written; the write slot of token t is ctx - Tq + t.
static int run_layer(const PttsModel* m, const PttsState* s, const Path& path, Pending& pend, int l, int Tq,
                     int ctx_base, const int* ctx_dyn, cudaStream_t st) {
  const PttsLayer& L = m->layers[l];
  const int H = m->hidden, F = m->ffn, heads = m->n_heads, d = m->head_dim, B = s->B, dt = m->wdtype;
  const int rows = B * Tq;
  const KvView kv = self_kv(m, s, l, Tq);
  const float scale = 1.0f / sqrtf((float)d);
  // ---- self attention block
  PTTS_UNLESS_SKIPPED(1, ln_step(m, s, path, pend, L.ln1_g, L.ln1_b, rows, 1, 0, st));
  int bn_q = 64;
  const bool fuse = path.gemm && Tq == 1 && d == 64 && qkv_fusion_enabled();
  const int Sq = fuse ? gemm_bf16_plan_splits(rows, 3 * H, H, &bn_q) : 0;
  const bool fused_self = fuse && Sq > 1 && bn_q == 0;
  PTTS_CHECK_ARG(kv.dtype != PTTS_FP8 || fused_self || Tq > 1,
                 "fp8 KV cache: the decode step needs the fused tcgen05 path (bf16 weights, head_dim 64)");
  if (fused_self) {
    // decode: QKV as split-K partial sums; the attention CTA of (b, head) folds them, scales q and writes the new
    // K / V row into the cache itself
    float* qpart = work_area(m, s);
    // (The QKV GEMM releasing the self-attention kernel BEFORE its own wait - attention CTAs resident, first K batch in
    // flight while LayerNorm / QKV still run - was measured: 1.2200 / 1.2202 against 1.2209 / 1.2219 ms, noise level.)
    PTTS_UNLESS_SKIPPED(2, gemm_bf16_partial(s->xn, L.w_qkv, rows, 3 * H, H, Sq, 0, qpart, st));
    if (wpf_mask() & 2) set_l2_hint(L.w_o, (long long)H * H * dtype_size(dt));
    // self-attention kernel choice (PttsState.self_attn_kernel): one CTA per (utterance, head) item (split-KV for small
    // batches), or - opt-in, 2 - the persistent warp-specialised TMA + tcgen05 kernel.  Measured on B200 (B=64, 860
    // frames, profiles/r02_attn_ws.txt): per item 1.285 (bf16 cache) / 1.097 ms (fp8) per decode step, persistent 1.475 /
    // 1.372 ms - its exact two-pass softmax parks the V tiles of an item in the shared-memory ring until the item's
    // probabilities exist, the ring holds ~1.5 items, and the HBM latency is exposed once per item; automatic (0)
    // therefore stays with the per-item kernels
    const bool ws = s->self_attn_kernel == 2 && dt == PTTS_BF16 && attention_ws_supported(B, heads, d, kv.c_stride);
    if (ws)
      PTTS_UNLESS_SKIPPED(4, attention_ws_fused(qpart, Sq, (int64_t)rows * 3 * H, 3 * H, scale, ctx_base - 1, ctx_dyn,
                                                kv.k, kv.v, kv.dtype, kv.k_scale, kv.v_scale, B, heads, kv.c_stride,
                                                s->prompt_mask, s->P, s->attn, st));
    else if (kv.dtype == PTTS_FP8)
      PTTS_UNLESS_SKIPPED(4, attention_fp8_fused(qpart, Sq, (int64_t)rows * 3 * H, 3 * H, scale, ctx_base - 1, ctx_dyn,
                                                 kv.k, kv.v, kv.k_scale, kv.v_scale, B, heads, kv.c_stride, ctx_base,
                                                 ctx_dyn, s->prompt_mask, s->P, s->attn, attn_area(s), st));
    else
      PTTS_UNLESS_SKIPPED(4, attention_fast_fused(qpart, Sq, (int64_t)rows * 3 * H, 3 * H, 1, scale, ctx_base - 1,
                                                  ctx_dyn, kv.k, kv.v, B, heads, kv.c_stride, ctx_base, ctx_dyn,
                                                  s->prompt_mask, s->P, s->attn, attn_area(s), nullptr, 0, 0, nullptr,
                                                  st));
  } else {
    if (path.gemm) {  // q (scaled) and the new K/V rows go straight to their buffers / cache slots
      PTTS_UNLESS_SKIPPED(2, gemm_bf16_scatter(s->xn, L.w_qkv, rows, 3 * H, H, 1, scale, s->q, kv.k, kv.v, H, heads, d,
                                               Tq, kv.c_stride, ctx_base - Tq, ctx_dyn, st));
    } else {
      void* qkv = (void*)work_area(m, s);
      PTTS_TRY(linear_dispatch(s->impl, s->xn, dt, L.w_qkv, dt, nullptr, rows, 3 * H, H, 1.0f, PTTS_ACT_NONE, nullptr,
                               nullptr, qkv, dt, st));
      PTTS_TRY(qkv_scatter(qkv, dt, rows, H, heads, d, Tq, kv.c_stride, ctx_base - Tq, ctx_dyn, scale, s->q, kv.k, kv.v,
                           st));
    }
    PTTS_UNLESS_SKIPPED(4, attention_dispatch(s->impl, s->q, kv.k, kv.v, dt, B, heads, d, Tq, kv.c_stride, ctx_base,
                                              ctx_dyn, Tq > 1, s->prompt_mask, s->P, s->attn, attn_area(s), st));
  }
  const bool cross = s->S > 0;
  const bool use_xblock = cross && fuse && path.ln && dt == PTTS_BF16 && xblock_supported(H, heads, d, s->S);
  // In front of the fused cross-attention block the out_proj GEMM releases its dependent only AFTER its own wait:
  // released early, the block's CTAs become resident under the self-attention kernel and their pre-wait traffic
  // (128 KB of Wq per SM, cross K/V) competes with the K/V stream (diagnostic build: PTTS_XB_LATE=0 early, 2 only when
  // the out_proj's MMA burst is done)
  if (use_xblock) {
    const int xb_late = PTTS_KNOB("PTTS_XB_LATE", 1);
    if (xb_late) gemm_next_trigger_late(xb_late == 2);
  }
  // (An L2 prefetch of the NEXT layer's K/V rows queued from this out_proj, right after the self-attention kernel has
  // finished its stream, was measured on B200 at B=64 / 860 frames: fp8 cache 1.097 -> 1.477 (K) / 1.485 ms (K+V) per
  // decode step, bf16 cache 1.286 -> 2.024 / 2.050 ms - the prefetch costs its full HBM time in front of the chain's
  // weight loads and the later attention launch does not find the rows in L2.  Removed; profiles/r02_l2_prefetch.txt.)
  PTTS_UNLESS_SKIPPED(8, residual_linear(m, s, path, pend, s->attn, L.w_o, rows, H, st));
  // ---- cross attention block (modeling_parler_tts.py:465-483)
  if (cross) {
    const size_t es = dtype_size(dt);
    char* kc = (char*)s->kv_cross + (size_t)(2 * l) * kv_layer_elems(m, B, s->S) * es;
    char* vc = (char*)s->kv_cross + (size_t)(2 * l + 1) * kv_layer_elems(m, B, s->S) * es;
    // decode: the whole block (fold + LayerNorm, q projection, attention over the cached T5 K/V, out projection as
    // per-head partial sums) is ONE launch (fast_xblock.cu); the next LayerNorm folds the self-attention out_proj
    // slices and the per-head slices, stored back to back, into the residual stream
    if (use_xblock && pend.part && pend.part == work_area(m, s) && pend.stride == (int64_t)rows * H &&
        (int64_t)(pend.S + heads) * rows * H <= work_area_floats(m, rows) && !(skip_mask() & (1 | 16 | 32))) {
      float* xout = work_area(m, s) + (int64_t)pend.S * pend.stride;
      if (wpf_mask() & 4) set_l2_hint(L.w_fc1, (long long)F * H * es);  // fc1 is two launches ahead
      // per-head slices stored [B][heads][H]: the next LayerNorm pulls a row's slices with one bulk copy (diagnostic
      // build, PTTS_XB_ROWOUT=0: [heads][B][H] folded with strided loads)
      const int by_row = PTTS_KNOB("PTTS_XB_ROWOUT", 1) && (int64_t)heads * H * 4 <= 160 * 1024;
      PTTS_UNLESS_SKIPPED(256, xblock_fast(s->h, pend.part, pend.S, pend.stride, L.ln2_g, L.ln2_b, 1e-5f, L.w_qc, L.w_oc,
                                           kc, vc, s->enc_mask, B, heads, H, s->S, scale, xout, by_row, nullptr, 0, 0,
                                           nullptr, st));
      if (by_row) {
        pend.rowblk = xout;
        pend.n_blk = heads;
      } else {
        pend.S += heads;
      }
    } else {
      PTTS_UNLESS_SKIPPED(1, ln_step(m, s, path, pend, L.ln2_g, L.ln2_b, rows, 1, 0, st));
      int bn_c = 64;
      const int Sc = fuse ? gemm_bf16_plan_splits(rows, H, H, &bn_c) : 0;
      if (fuse && Sc > 1 && bn_c == 0) {
        float* qpart = work_area(m, s);
        PTTS_UNLESS_SKIPPED(16, gemm_bf16_partial(s->xn, L.w_qc, rows, H, H, Sc, 0, qpart, st));
        PTTS_UNLESS_SKIPPED(32, attention_fast_fused(qpart, Sc, (int64_t)rows * H, H, 0, scale, 0, nullptr, kc, vc, B,
                                                     heads, s->S, s->S, nullptr, s->enc_mask, s->S, s->attn,
                                                     attn_area(s), nullptr, 0, 0, nullptr, st));
      } else {
        PTTS_UNLESS_SKIPPED(16, linear_dispatch(s->impl, s->xn, dt, L.w_qc, dt, nullptr, rows, H, H, scale,
                                                PTTS_ACT_NONE, nullptr, nullptr, s->q, dt, st));
        PTTS_UNLESS_SKIPPED(32, attention_dispatch(s->impl, s->q, kc, vc, dt, B, heads, d, Tq, s->S, s->S, nullptr, 0,
                                                   s->enc_mask, s->S, s->attn, attn_area(s), st));
      }
      PTTS_UNLESS_SKIPPED(8, residual_linear(m, s, path, pend, s->attn, L.w_oc, rows, H, st));
    }
  }
  // ---- feed forward block
  if (fuse && (wpf_mask() & 8)) set_l2_hint(L.w_fc2, (long long)F * H * dtype_size(dt));  // fc2 is three launches ahead
  PTTS_UNLESS_SKIPPED(1, ln_step(m, s, path, pend, L.ln3_g, L.ln3_b, rows, 1, 0, st));
  int bn_f = 64;
  const int Sf = (fuse && fc1_partials_enabled()) ? gemm_bf16_plan_splits(rows, F, H, &bn_f) : 0;
#ifndef PTTS_FC1_ACT
#define PTTS_FC1_ACT 1
#endif
#ifndef PTTS_FC1_ACT_MAXROWS
#define PTTS_FC1_ACT_MAXROWS 128
#endif
  // decode: fc1 over the whole K with the activation in its own epilogue - one launch instead of split-K partial sums +
  // reduce / activation.  Measured on B200 (B=64, 860 frames, same box): see profiles/r02_fc1_act.txt
  if (PTTS_FC1_ACT && fuse && rows <= PTTS_FC1_ACT_MAXROWS && gemm_bf16_act_supported(rows, F, H) && !(skip_mask() & 64)) {
    PTTS_TRY(gemm_bf16_act(s->xn, L.w_fc1, rows, F, H, m->act_fn, s->ffn, st));
  } else if (Sf > 1 && bn_f == 0 && (int64_t)Sf * rows * F <= work_area_floats(m, rows)) {
    // decode: fc1 as split-K partial sums + one small reduce / activation kernel spread over all SMs.  (Folding the
    // activation into the second-arriving K-slice CTA of each output tile was measured slower: 1.539 against 1.457 ms
    // per decode step, 32 finishing CTAs do what the separate kernel spreads over 148 SMs.)
    float* fpart = work_area(m, s);
    PTTS_UNLESS_SKIPPED(64, gemm_bf16_partial(s->xn, L.w_fc1, rows, F, H, Sf, 0, fpart, st));
    if (wpf_mask() & 1) {  // the next layer's QKV matrix (or the LM heads) is three launches ahead
      if (l + 1 < m->n_layers)
        set_l2_hint(m->layers[l + 1].w_qkv, (long long)3 * H * H * dtype_size(dt));
      else
        set_l2_hint(m->w_heads, (long long)m->n_codebooks * m->vocab * H * dtype_size(dt));
    }
    PTTS_UNLESS_SKIPPED(64, partial_act_fast(fpart, Sf, (int64_t)rows * F, rows, F, m->act_fn, s->ffn, st));
  } else {
    PTTS_UNLESS_SKIPPED(64, linear_dispatch(s->impl, s->xn, dt, L.w_fc1, dt, nullptr, rows, F, H, 1.0f, m->act_fn,
                                            nullptr, nullptr, s->ffn, dt, st));
  }
  PTTS_UNLESS_SKIPPED(8, residual_linear(m, s, path, pend, s->ffn, L.w_fc2, rows, F, st));
  return PTTS_OK;
}

// [B][T][K][V] -> [B][K][T][V]
__global__ void logits_bktv_kernel(const float* __restrict__ in, int T, int K, int V, float* __restrict__ out) {
  const int row = blockIdx.x;  // (b, t, k)
  const int k = row % K, t = (row / K) % T, b = row / (K * T);
  const float* src = in + (int64_t)row * V;
  float* dst = out + (((int64_t)b * K + k) * T + t) * V;
  for (int i = threadIdx.x; i < V; i += blockDim.x) dst[i] = src[i];
}

}  // namespace ptts

using namespace ptts;

extern "C" int64_t ptts_scratch_floats(const PttsModel* m, int32_t rows) {
  if (!m || rows <= 0) return 0;
  int64_t w = (int64_t)rows * 3 * m->hidden;                                  // fused qkv / cross kv staging
  const int64_t b = (int64_t)rows * m->n_codebooks * m->vocab;                // all-position logits staging
  const int64_t c = (int64_t)kMaxSplitK * (rows < 256 ? rows : 256) * 3 * m->hidden;  // split-K partial sums (qkv)
  if (b > w) w = b;
  if (c > w) w = c;
  return kScratchHeader + attn_scratch_floats(rows, m->n_heads) + w + 4096;  // rows >= B bounds the split-KV partial area
}

extern "C" int ptts_embed_prefill(const PttsModel* m, const int64_t* ids, int64_t ld_ids, int32_t B, int32_t T0,
                                  const float* prompt_table, const int64_t* prompt_ids, const float* prompt_hidden,
                                  int32_t P, float* h_out, ptts_stream_t s) {
  PTTS_TRY(check_model(m));
  PTTS_CHECK_ARG(ids && h_out && B > 0 && T0 > 0 && P >= 0, "embed_prefill: bad arguments");
  PTTS_CHECK_ARG(P == 0 || prompt_hidden || (prompt_table && prompt_ids), "embed_prefill: no prompt source");
  return embed_prefill(m, ids, ld_ids, nullptr, 0, B, B, T0, prompt_table, prompt_ids, prompt_hidden, P, h_out,
                       as_stream(s));
}

extern "C" int ptts_embed_tokens(const PttsModel* m, const int64_t* ids, int64_t ld_ids, int32_t B, int32_t T,
                                 int32_t pos0, float* h_out, ptts_stream_t s) {
  PTTS_TRY(check_model(m));
  PTTS_CHECK_ARG(ids && h_out && B > 0 && T > 0 && pos0 >= 0 && pos0 + T <= m->max_pos, "embed_tokens: bad arguments");
  return embed_prefill(m, ids, ld_ids, nullptr, 0, B, B, T, nullptr, nullptr, nullptr, 0, h_out, as_stream(s), pos0);
}

extern "C" int ptts_layernorm(const float* x, const float* g, const float* b, int32_t rows, int32_t H, float eps,
                              void* y, int32_t ydtype, ptts_stream_t s) {
  PTTS_CHECK_ARG(x && g && b && y && rows >= 0 && H > 0, "layernorm: bad arguments");
  if (fast_layernorm_supported(H))  // no pending split-K partials: the residual rows are only read
    return layernorm_fast(const_cast<float*>(x), nullptr, 0, 0, g, b, rows, H, eps, 1, 0, y, ydtype, as_stream(s));
  return layernorm(x, g, b, rows, H, eps, H, 0, y, ydtype, as_stream(s));
}

extern "C" int ptts_linear(const void* x, int32_t xdtype, const void* w, int32_t wdtype, const float* bias, int32_t M,
                           int32_t N, int32_t K, float alpha, int32_t act, const float* residual,
                           const int32_t* row_mask, void* y, int32_t ydtype, int32_t impl, ptts_stream_t s) {
  PTTS_CHECK_ARG(x && w && y && M >= 0 && N > 0 && K > 0, "linear: bad arguments");
  return linear_dispatch(impl, x, xdtype, w, wdtype, bias, M, N, K, alpha, act, residual, row_mask, y, ydtype,
                         as_stream(s));
}

extern "C" int ptts_attention(const void* q, const void* cache_k, const void* cache_v, int32_t dtype, int32_t B,
                              int32_t heads, int32_t d, int32_t Tq, int32_t c_stride, int32_t ctx_base,
                              const int32_t* ctx_dyn, int32_t causal, const int32_t* key_mask, int32_t mask_len,
                              void* out, float* scratch, int32_t impl, ptts_stream_t s) {
  PTTS_CHECK_ARG(q && cache_k && cache_v && out, "attention: null pointer");
  PTTS_CHECK_ARG(B > 0 && heads > 0 && d > 0 && Tq > 0 && c_stride > 0, "attention: bad dims");
  return attention_dispatch(impl, q, cache_k, cache_v, dtype, B, heads, d, Tq, c_stride, ctx_base, ctx_dyn, causal,
                            key_mask, mask_len, out, scratch, as_stream(s));
}

extern "C" int ptts_attention_fp8(const void* q, const void* cache_k, const void* cache_v, const float* k_scale,
                                  const float* v_scale, int32_t B, int32_t heads, int32_t d, int32_t c_stride,
                                  int32_t ctx_base, const int32_t* ctx_dyn, const int32_t* key_mask, int32_t mask_len,
                                  void* out, float* scratch, ptts_stream_t s) {
  PTTS_CHECK_ARG(q && cache_k && cache_v && k_scale && v_scale && out, "attention_fp8: null pointer");
  PTTS_CHECK_ARG(B > 0 && heads > 0 && c_stride > 0 && attention_fp8_supported(d), "attention_fp8: bad dims");
  return attention_fp8(q, cache_k, cache_v, k_scale, v_scale, B, heads, c_stride, ctx_base, ctx_dyn, key_mask,
                       mask_len, out, scratch, as_stream(s));
}

extern "C" int ptts_attention_ws(const float* part, int32_t S, int64_t part_stride, int32_t ncols, float scale,
                                 int32_t slot_base, const int32_t* slot_dyn, void* cache_k, void* cache_v,
                                 int32_t kv_dtype, const float* k_scale, const float* v_scale, int32_t B, int32_t heads,
                                 int32_t d, int32_t c_stride, const int32_t* key_mask, int32_t mask_len, void* out,
                                 ptts_stream_t s) {
  PTTS_CHECK_ARG(part && cache_k && cache_v && out && S > 0 && B > 0 && heads > 0 && c_stride > 0,
                 "attention_ws: bad arguments");
  PTTS_CHECK_ARG(kv_dtype == PTTS_BF16 || (kv_dtype == PTTS_FP8 && k_scale && v_scale), "attention_ws: bad kv_dtype");
  if (!attention_ws_supported(B, heads, d, c_stride)) {
    set_error("attention_ws: unsupported shape (head_dim 64, at least 2 items per SM)");
    return PTTS_E_UNSUPPORTED;
  }
  return attention_ws_fused(part, S, part_stride, ncols, scale, slot_base, slot_dyn, cache_k, cache_v, kv_dtype, k_scale,
                            v_scale, B, heads, c_stride, key_mask, mask_len, out, as_stream(s));
}

extern "C" int ptts_kv_quantize_fp8(const void* stage, int32_t n_lk, int32_t B, int32_t heads, int32_t d, int32_t slots,
                                    int32_t c_max, uint32_t* amax_ws, float* scale, void* cache, ptts_stream_t s) {
  PTTS_CHECK_ARG(attention_fp8_supported(d), "kv_quantize_fp8: head_dim must be 64");
  return kv_quantize_fp8(stage, n_lk, B, heads, slots, c_max, amax_ws, scale, cache, as_stream(s));
}

extern "C" int ptts_cross_kv(const PttsModel* m, const PttsState* s, const void* enc, ptts_stream_t stream) {
  PTTS_TRY(check_model(m));
  PTTS_TRY(check_state(m, s));
  PTTS_CHECK_ARG(enc && s->S > 0, "cross_kv: no encoder states");
  cudaStream_t st = as_stream(stream);
  const int H = m->hidden, dt = m->wdtype, rows = s->B * s->S;
  const size_t es = dtype_size(dt);
  const Path path = pick_path(m, s);
  for (int l = 0; l < m->n_layers; ++l) {
    char* kc = (char*)s->kv_cross + (size_t)(2 * l) * kv_layer_elems(m, s->B, s->S) * es;
    char* vc = (char*)s->kv_cross + (size_t)(2 * l + 1) * kv_layer_elems(m, s->B, s->S) * es;
    if (path.gemm) {
      PTTS_TRY(gemm_bf16_scatter(enc, m->layers[l].w_kvc, rows, 2 * H, H, 0, 1.0f, nullptr, kc, vc, H, m->n_heads,
                                 m->head_dim, s->S, s->S, 0, nullptr, st));
    } else {
      float* stage = work_area(m, s);
      PTTS_TRY(linear_dispatch(s->impl, enc, dt, m->layers[l].w_kvc, dt, nullptr, rows, 2 * H, H, 1.0f, PTTS_ACT_NONE,
                               nullptr, nullptr, stage, dt, st));
      PTTS_TRY(kv_scatter(stage, dt, rows, H, m->n_heads, m->head_dim, s->S, kc, vc, st));
    }
  }
  return PTTS_OK;
}

extern "C" int ptts_prefill(const PttsModel* m, const PttsState* s, const float* prompt_table,
                            const int64_t* prompt_ids, const float* prompt_hidden, float* all_logits,
                            int32_t run_epilogue, ptts_stream_t stream) {
  PTTS_TRY(check_model(m));
  PTTS_TRY(check_state(m, s));
  PTTS_CHECK_ARG(s->P == 0 || prompt_hidden || (prompt_table && prompt_ids), "prefill: no prompt source");
  cudaStream_t st = as_stream(stream);
  const int T = s->P + s->T0, B = s->B, H = m->hidden, K = m->n_codebooks, V = m->vocab, dt = m->wdtype;
  const int cfg = s->guidance_scale > 1.0f;
  PTTS_CHECK_ARG(!cfg || B % 2 == 0, "prefill: guidance needs an even model batch");
  // with classifier-free guidance rows [B/2, B) replay the ids of rows [0, B/2) (modeling_parler_tts.py:2109-2118)
  PTTS_TRY(embed_prefill(m, s->ids, s->max_len, s->mask, s->max_len, B, cfg ? B / 2 : B, s->T0, prompt_table, prompt_ids, prompt_hidden,
                         s->P, s->h, st));
  const Path path = pick_path(m, s);
  Pending pend;
  for (int l = 0; l < m->n_layers; ++l) PTTS_TRY(run_layer(m, s, path, pend, l, T, T, nullptr, st));
  if (all_logits) {
    float* stage = work_area(m, s);
    PTTS_TRY(ln_step(m, s, path, pend, m->lnf_g, m->lnf_b, B * T, 1, 0, st));
    PTTS_TRY(linear_dispatch(s->impl, s->xn, dt, m->w_heads, dt, nullptr, B * T, K * V, H, 1.0f, PTTS_ACT_NONE,
                             nullptr, nullptr, stage, PTTS_F32, st));
    logits_bktv_kernel<<<B * T * K, 256, 0, st>>>(stage, T, K, V, all_logits);
    PTTS_CHECK_LAUNCH("logits_bktv");
  }
  // logits of the last position only (transformers takes outputs.logits[:, -1, :])
  PTTS_TRY(ln_step(m, s, path, pend, m->lnf_g, m->lnf_b, B, T, T - 1, st));
  PTTS_TRY(linear_dispatch(s->impl, s->xn, dt, m->w_heads, dt, nullptr, B, K * V, H, 1.0f, PTTS_ACT_NONE, nullptr,
                           nullptr, s->logits, PTTS_F32, st));
  if (run_epilogue) PTTS_TRY(step_epilogue(m, s, st));
  if (s->kv_dtype == PTTS_FP8)  // scales from the staged prompt positions, then the e4m3 rows [0, P + T0) of every layer
    PTTS_TRY(kv_quantize_fp8(s->kv_stage, 2 * m->n_layers, B, m->n_heads, T, s->c_max,
                             reinterpret_cast<unsigned int*>(work_area(m, s)), s->kv_scale, s->kv_self, st));
  (void)take_l2_hint();  // a hint never outlives the call that set it (its matrix may be freed afterwards)
  return PTTS_OK;
}

extern "C" int ptts_decode_step(const PttsModel* m, const PttsState* s, int32_t run_epilogue, ptts_stream_t stream) {
  PTTS_TRY(check_model(m));
  PTTS_TRY(check_state(m, s));
  cudaStream_t st = as_stream(stream);
  const int B = s->B, H = m->hidden, K = m->n_codebooks, V = m->vocab, dt = m->wdtype;
  // context after this step's K/V write = P + len   (len = columns already in ids; input column = len-1)
  const Path path = pick_path(m, s);
  Pending pend;
  for (int l = 0; l < m->n_layers; ++l) PTTS_TRY(run_layer(m, s, path, pend, l, 1, s->P, &s->ctrl->len, st));
  PTTS_TRY(ln_step(m, s, path, pend, m->lnf_g, m->lnf_b, B, 1, 0, st));
#ifndef PTTS_HEADS_PARTIAL
#define PTTS_HEADS_PARTIAL 1
#endif
  // The fused step (epilogue in the same call): the 9 lm_heads as split-K partial sums of the lean tcgen05 kernel - 4
  // K slices, 308 CTAs of 96 KB, two per SM - that the epilogue kernel adds up while it reads them; the logits matrix is
  // only materialised when somebody else reads it (host-side processors, forward()).  Measured: profiles/r02_heads_partial.txt
  constexpr int kHeadSlices = 4;  // <= 4: the epilogue kernel folds at most four slices
  if (PTTS_HEADS_PARTIAL && run_epilogue && path.gemm && path.ln && B % 64 == 0 && H % (64 * kHeadSlices) == 0 &&
      H / 64 / kHeadSlices <= 8 && (int64_t)kHeadSlices * B * K * V <= work_area_floats(m, B)) {
    float* hpart = work_area(m, s);
    PTTS_TRY(gemm_bf16_partial(s->xn, m->w_heads, B, K * V, H, kHeadSlices, 0, hpart, st));
    PTTS_TRY(step_epilogue(m, s, st, hpart, kHeadSlices, (int64_t)B * K * V));
    (void)take_l2_hint();
    return PTTS_OK;
  }
  PTTS_TRY(linear_dispatch(s->impl, s->xn, dt, m->w_heads, dt, nullptr, B, K * V, H, 1.0f, PTTS_ACT_NONE, nullptr,
                           nullptr, s->logits, PTTS_F32, st));
  if (run_epilogue) PTTS_TRY(step_epilogue(m, s, st));
  (void)take_l2_hint();  // a hint never outlives the call that set it (its matrix may be freed afterwards)
  return PTTS_OK;
}
