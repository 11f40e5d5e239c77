/*
 * parler_b200.h - C ABI of libparler_b200.so: the B200 (sm_100a) codec-token generation path of Parler-TTS.
 *
 * The reference (danlyth/parler-tts, /root/reference) is pure Python and has NO FFI; every entry point below
 * names the reference Python function (file:line under /root/reference/parler_tts/) whose device work it
 * replaces.  INTEGRATION.md shows the ctypes binding a maintainer of the reference would add.
 *
 * Conventions
 *   - extern "C", plain pointers and sizes.  All tensor pointers are DEVICE pointers owned by the caller (PyTorch
 *     allocates; the library never allocates, frees or keeps a pointer after the call returns).
 *   - ptts_stream_t is a cudaStream_t (torch.cuda.current_stream().cuda_stream).  No call synchronises the host and
 *     no call changes state that a later call's RESULT depends on, so every call is CUDA-graph capturable and
 *     replayable.  Host state the library does keep (process-wide, guarded by a mutex where it is shared):
 *       * a cache of CUtensorMap descriptors keyed by (base pointer, rows, cols, row pitch in bytes, box rows, element
 *         type, k-blocks per box: 0 for a 2-D map, > 0 for the 3-D [k-block][row][64] view whose one copy brings several
 *         swizzled k-blocks) for the TMA operand loads of the tcgen05 GEMMs / cross-attention block.  A descriptor only encodes that
 *         key, so a buffer freed and re-allocated at the same address with another shape gets another entry; the
 *         cache is cleared when it exceeds 4096 entries;
 *       * per-kernel high-water marks of cudaFuncAttributeMaxDynamicSharedMemorySize (raised on first use);
 *       * the launch tally behind ptts_launch_count() and the thread-local text behind ptts_last_error_string().
 *     Environment variables are NOT read by this library (the diagnostic build libparler_b200_diag.so, `make DIAG=1`,
 *     is the one with PTTS_* knobs and the ptts_diag_* exports of include/parler_b200_diag.h).
 *   - return value: 0 = ok; non-zero = PTTS_E_* (argument or launch error), text in ptts_last_error_string().
 *   - "T" tensors have the element type named by a PTTS_F32 / PTTS_BF16 enum; accumulation is always fp32.
 *   - dynamic scalars that change between graph replays (current sequence length) live in the device-resident
 *     control block PttsCtrl.
 */
#ifndef PARLER_B200_H
#define PARLER_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* ptts_stream_t;

#if defined(__GNUC__)
#define PTTS_API __attribute__((visibility("default")))
#else
#define PTTS_API
#endif

enum { PTTS_F32 = 0, PTTS_BF16 = 1, PTTS_FP8 = 2 /* e4m3, self-attention KV cache only */ };
enum { PTTS_ACT_NONE = 0, PTTS_ACT_GELU = 1, PTTS_ACT_RELU = 2, PTTS_ACT_SILU = 3, PTTS_ACT_GELU_TANH = 4 };
enum { PTTS_OK = 0, PTTS_E_ARG = 1, PTTS_E_LAUNCH = 2, PTTS_E_UNSUPPORTED = 3 };
enum { PTTS_IMPL_SIMPLE = 0, PTTS_IMPL_FAST = 1 };

#define PTTS_ABI_VERSION 2
#define PTTS_MAX_CODEBOOKS 16

/* Device-resident control block (int32 words), one per generation state.  Written only by kernels. */
typedef struct {
  int32_t len;          /* columns filled in `ids` (the next token goes to column `len`) */
  int32_t done;         /* 1 once every row has finished or len reached max_len */
  int32_t stop_len;     /* value of len at which the reference loop would have returned */
  int32_t n_unfinished; /* rows (of B*K) still unfinished (transformers greedy_search/sample bookkeeping) */
  int32_t ticket;       /* internal: block arrival counter of the epilogue kernel */
  int32_t steps;        /* decode-step epilogues executed (diagnostic) */
  int32_t reserved[2];  /* Philox key (seed lo, hi) of the sampling epilogue; written by the host at reset */
} PttsCtrl;

/* One decoder layer, weights packed row-major [out_features][in_features], element type = PttsModel.wdtype.
 * Reference keys: model.decoder.layers.{i}.* (modeling_parler_tts.py:386-413). */
typedef struct {
  const void* w_qkv;   /* [3H][H]   self_attn q_proj | k_proj | v_proj stacked */
  const void* w_o;     /* [H][H]    self_attn.out_proj */
  const void* w_qc;    /* [H][H]    encoder_attn.q_proj */
  const void* w_kvc;   /* [2H][H]   encoder_attn k_proj | v_proj stacked (prefill only) */
  const void* w_oc;    /* [H][H]    encoder_attn.out_proj */
  const void* w_fc1;   /* [F][H] */
  const void* w_fc2;   /* [H][F] */
  const float *ln1_g, *ln1_b; /* self_attn_layer_norm */
  const float *ln2_g, *ln2_b; /* encoder_attn_layer_norm */
  const float *ln3_g, *ln3_b; /* final_layer_norm */
} PttsLayer;

typedef struct {
  int32_t n_layers, hidden, n_heads, head_dim, ffn, n_codebooks, vocab, max_pos;
  int32_t wdtype;      /* PTTS_F32 | PTTS_BF16: matrices, KV caches and GEMM-input activations */
  int32_t act_fn;      /* PTTS_ACT_* */
  int32_t pad_id, eos_id, bos_id, reserved0;
  const float* embed;  /* [K][vocab+1][H] fp32  model.decoder.embed_tokens.{k}.weight */
  const float* pos;    /* [max_pos][H] fp32     model.decoder.embed_positions.weights */
  const float *lnf_g, *lnf_b; /* model.decoder.layer_norm */
  const void* w_heads; /* [K*vocab][H]          lm_heads.{k}.weight stacked */
  const PttsLayer* layers; /* HOST array of n_layers entries */
} PttsModel;

/* Generation state for a batch of B utterances (all device pointers; buffers sized for `rows` = max rows any call
 * processes: B for decode, B*(P+T0) for prefill). */
typedef struct {
  int32_t B, S, P, T0;     /* batch, encoder length, prompt length, start columns (1 without audio prompt) */
  int32_t c_max;           /* self-attention cache slots per sequence (>= P + max_len) */
  int32_t max_len;         /* M: generation max_length (columns of ids / mask) */
  int32_t impl;            /* PTTS_IMPL_* kernel family */
  int32_t do_sample, top_k; float temperature; int32_t suppress_from; /* epilogue: sampling + optional vocab cut */
  float guidance_scale;    /* >1: rows [B/2, B) are the unconditional half (classifier-free guidance) */
  uint64_t seed; int64_t row_offset; /* seed: informational (the kernels read the key from PttsCtrl.reserved);
                                        row_offset: global row index of local row 0 (batch sharding) */
  PttsCtrl* ctrl;
  void* kv_self;           /* [L][2][B][heads][c_max][d]  kv_dtype */
  void* kv_cross;          /* [L][2][B][heads][S][d]      T */
  const int32_t* enc_mask; /* [B][S] or NULL */
  const int32_t* prompt_mask; /* [B][P] or NULL */
  float* h;                /* [rows][H] fp32 residual stream */
  void* xn;                /* [rows][H] T   LayerNorm output */
  void* q;                 /* [rows][H] T */
  void* attn;              /* [rows][H] T */
  void* ffn;               /* [rows][F] T */
  float* logits;           /* [B*K][vocab] fp32 */
  float* scratch;          /* fp32 scratch, >= ptts_scratch_floats(...) */
  int64_t* ids;            /* [B*K][max_len] raw generated ids */
  const int64_t* mask;     /* [B*K][max_len] delay-pattern mask */
  int32_t* unfinished;     /* [B*K] */
  /* ---- ABI 2 */
  int32_t kv_dtype;        /* element type of kv_self: the weight dtype (PTTS_F32 / PTTS_BF16), or PTTS_FP8: e4m3 rows with
                              one dequantisation scale per (layer, K|V, head); needs bf16 weights + PTTS_IMPL_FAST */
  int32_t min_length;      /* epilogue: EOS is suppressed while `ids` holds fewer than this many columns = max(min_length,
                              T0 + min_new_tokens) of transformers' MinLength / MinNewTokensLength processors; 0 = off */
  float top_p;             /* epilogue, sampling only: nucleus filter (TopPLogitsWarper, min_tokens_to_keep 1); >= 1 = off */
  float repetition_penalty;/* epilogue: RepetitionPenaltyLogitsProcessor over the row's own history; 1 = off */
  int32_t logits_premixed; /* guidance_scale > 1 only: rows [0, B/2*K) of `logits` already hold the guidance-mixed scores
                              (host-side logits processors ran); the epilogue skips its own mix */
  int32_t self_attn_kernel;/* decode-step self-attention kernel: 0 = automatic (today: one CTA per (utterance, head) item,
                              split-KV for small batches), 1 = always one CTA per item, 2 = the warp-specialised persistent
                              TMA + tcgen05 kernel whenever its shape constraints hold (bf16 weights, d = 64, >= 2 items / SM) */
  float* kv_scale;         /* [L][2][heads] fp32, PTTS_FP8 only: written by ptts_prefill, read by ptts_decode_step */
  void* kv_stage;          /* [L][2][B][heads][P+T0][d] T, PTTS_FP8 only: staging cache of the prefill */
  int64_t pad_id, eos_id;  /* generation-config ids used by the epilogue (may differ from the decoder config's) */
} PttsState;

PTTS_API int32_t ptts_abi_version(void);
PTTS_API const char* ptts_last_error_string(void);
PTTS_API int64_t ptts_scratch_floats(const PttsModel* m, int32_t rows);
/* Cumulative number of kernel launches this library has enqueued from the calling process (one per PTTS launch
 * site executed on the host; a CUDA-graph replay re-runs the launches recorded at capture time). */
PTTS_API int64_t ptts_launch_count(void);

/* ---- integer kernels -------------------------------------------------------------------------------------- */
/* build_delay_pattern_mask, modeling_parler_tts.py:77-139.  first_start (device int32[1]) receives the number of
 * leading columns that are already determined (the reference's `first_start_id`). */
PTTS_API int ptts_build_delay_mask(const int64_t* start_ids, int32_t B, int32_t K, int32_t T0, int32_t max_len, int64_t bos,
                          int64_t pad, int64_t* mask_out, int32_t* first_start, ptts_stream_t s);
/* apply_delay_pattern_mask, modeling_parler_tts.py:68-74 */
PTTS_API int ptts_apply_delay_mask(const int64_t* ids, int64_t ld_ids, const int64_t* mask, int64_t ld_mask, int32_t rows,
                          int32_t len, int64_t* out, int64_t ld_out, ptts_stream_t s);
/* de-delay of generate(), modeling_parler_tts.py:2633-2648 (no-prompt geometry): codes[b][k][f] = applied[b*K+k][f+k+1],
 * f < l_out-K; when l_out < 2K-1 the rebuilt mask is all -1 and codes[b][k][f] = applied[b*K+k][f], f < l_out. */
PTTS_API int ptts_dedelay(const int64_t* ids, int64_t ld_ids, const int64_t* mask, int64_t ld_mask, int32_t B, int32_t K,
                 int32_t l_out, int64_t* codes, ptts_stream_t s);
/* frame filter of the ragged DAC path, modeling_parler_tts.py:2668: valid[b][f] = all_k codes[b][k][f] < codebook_size;
 * counts[b] = number of valid frames; packed[b][k][0..counts[b]) = surviving frames in order. */
PTTS_API int ptts_filter_frames(const int64_t* codes, int32_t B, int32_t K, int32_t F, int64_t codebook_size, int64_t* packed,
                       int32_t* counts, ptts_stream_t s);

/* ---- decoder building blocks (each replaces eager torch ops of modeling_parler_tts.py) ---------------------- */
/* Sum_k embed_tokens[k](ids) + positions, modeling_parler_tts.py:836,884-886; prompt rows are gathered from
 * embed_prompts (modeling_parler_tts.py:2243-2245) and prepended (839-840).  ids: applied ids [B*K][ld_ids], columns
 * [0,T0).  h_out [B][P+T0][H]. */
PTTS_API int ptts_embed_prefill(const PttsModel* m, const int64_t* ids, int64_t ld_ids, int32_t B, int32_t T0,
                       const float* prompt_table, const int64_t* prompt_ids, const float* prompt_hidden, int32_t P,
                       float* h_out, ptts_stream_t s);
/* The same for T token columns without prompt: h_out[b][t] = sum_k embed_tokens[k](ids[b*K+k][t]) + positions[pos0 + t]
 * (the decode-time branch of modeling_parler_tts.py:836,884-886 with past_key_values_length = pos0); used by
 * ParlerTTSForCausalLM.forward(past_key_values=...). */
PTTS_API int ptts_embed_tokens(const PttsModel* m, const int64_t* ids, int64_t ld_ids, int32_t B, int32_t T,
                               int32_t pos0, float* h_out, ptts_stream_t s);
/* nn.LayerNorm (eps 1e-5), modeling_parler_tts.py:446,467,487,959.  x fp32 [rows][H] -> y T */
PTTS_API int ptts_layernorm(const float* x, const float* g, const float* b, int32_t rows, int32_t H, float eps, void* y,
                   int32_t ydtype, ptts_stream_t s);
/* nn.Linear: Y = act(alpha * X W^T + bias) [* row_mask] [+ residual], modeling_parler_tts.py:282,301-302,380,488-490,
 * 1143,2014-2017.  X [M][K] xdtype, W [N][K] wdtype, Y [M][N] ydtype (ld = N), residual fp32 [M][N] may alias Y. */
PTTS_API int ptts_linear(const void* x, int32_t xdtype, const void* w, int32_t wdtype, const float* bias, int32_t M, int32_t N,
                int32_t K, float alpha, int32_t act, const float* residual, const int32_t* row_mask, void* y,
                int32_t ydtype, int32_t impl, ptts_stream_t s);
/* Attention over a cache, modeling_parler_tts.py:326-378.  q [B][Tq][H] (already scaled), cache_k/v
 * [B][heads][c_stride][d]; query t of batch b sees keys j < ctx_base + (*ctx_dyn) and, if causal, j <= ctx - Tq + t;
 * key_mask [B][mask_len] (1 = keep) covers keys [0, mask_len), later keys are unmasked.  Masked scores are the
 * dtype's finfo.min exactly as the reference's additive mask produces. */
PTTS_API int ptts_attention(const void* q, const void* cache_k, const void* cache_v, int32_t dtype, int32_t B, int32_t heads,
                   int32_t d, int32_t Tq, int32_t c_stride, int32_t ctx_base, const int32_t* ctx_dyn, int32_t causal,
                   const int32_t* key_mask, int32_t mask_len, void* out, float* scratch, int32_t impl,
                   ptts_stream_t s);

/* No counterpart in the reference.  ptts_attention (fast family, Tq = 1,
 * d = 64, T = PTTS_F32 | PTTS_BF16) for callers that cannot allocate device memory: the split-KV scratch is a static
 * __device__ array inside the library (ptts_attention_noalloc_max_floats() = 262144 floats = 1 MiB per GPU, held until
 * the process exits) instead of a caller buffer; the call allocates nothing.  Limit: B * heads <= 3363 (the scratch
 * ptts_attention would need must fit), larger sizes return PTTS_E_ARG with a message.  All calls share the array, which
 * keeps the last call's contents (nothing needs clearing: every slot read is written within the same call); calls that
 * overlap on the GPU (different streams / threads) race on it - serialise them or use ptts_attention. */
PTTS_API int64_t ptts_attention_noalloc_max_floats(void);
PTTS_API int ptts_attention_noalloc(const void* q, const void* cache_k, const void* cache_v, int32_t dtype, int32_t B,
                                    int32_t heads, int32_t d, int32_t c_stride, int32_t ctx_base, const int32_t* ctx_dyn,
                                    const int32_t* key_mask, int32_t mask_len, void* out, ptts_stream_t s);
/* The same over an e4m3 cache [B][heads][c_stride][64] (PttsState.kv_dtype = PTTS_FP8): q bf16 [B][H] already
 * scaled, k_scale / v_scale [heads] fp32 dequantisation multipliers, out bf16 [B][H].  Tq = 1, d = 64. */
PTTS_API int ptts_attention_fp8(const void* q, const void* cache_k, const void* cache_v, const float* k_scale,
                                const float* v_scale, int32_t B, int32_t heads, int32_t d, int32_t c_stride,
                                int32_t ctx_base, const int32_t* ctx_dyn, const int32_t* key_mask, int32_t mask_len,
                                void* out, float* scratch, ptts_stream_t s);
/* The decode-step self-attention of ptts_decode_step alone, on the warp-specialised persistent kernel (TMA-staged
 * 128-row tiles, both products on tcgen05): q | k | v of the new token arrive as S split-K partial slices
 * part[s][b][3H] (fp32, summed in slice order), q is scaled by `scale`, the new K / V row is appended at slot
 * slot_base + *slot_dyn (= rows already cached) and the bf16 output row out[b][H] attends over rows [0, slot].
 * kv_dtype PTTS_BF16 or PTTS_FP8 (k_scale / v_scale [heads]).  Returns PTTS_E_UNSUPPORTED for shapes it does not take
 * (d != 64, fewer than 2 items per SM). */
PTTS_API int ptts_attention_ws(const float* part, int32_t S, int64_t part_stride, int32_t ncols, float scale,
                               int32_t slot_base, const int32_t* slot_dyn, void* cache_k, void* cache_v, int32_t kv_dtype,
                               const float* k_scale, const float* v_scale, int32_t B, int32_t heads, int32_t d,
                               int32_t c_stride, const int32_t* key_mask, int32_t mask_len, void* out, ptts_stream_t s);
/* Quantises staged bf16 cache rows [n_lk][B][heads][slots][64] (n_lk = layers x {K, V}) into the e4m3 cache
 * [n_lk][B][heads][c_max][64]: scale[n_lk][heads] = 4 * amax / 448 over the staged rows (head-room for later tokens),
 * rows converted with saturation.  amax_ws: n_lk * heads uint32 of workspace.  ptts_prefill calls this itself when
 * kv_dtype = PTTS_FP8; exported for tests. */
PTTS_API int ptts_kv_quantize_fp8(const void* stage, int32_t n_lk, int32_t B, int32_t heads, int32_t d, int32_t slots,
                                  int32_t c_max, uint32_t* amax_ws, float* scale, void* cache, ptts_stream_t s);

/* ---- the hot path -------------------------------------------------------------------------------------------- */
/* Cross-attention K/V of all layers from the projected encoder states (modeling_parler_tts.py:297-298), once per
 * utterance.  enc [B][S][H] T. */
PTTS_API int ptts_cross_kv(const PttsModel* m, const PttsState* st, const void* enc, ptts_stream_t s);
/* First forward of generate(): prompt + start columns (modeling_parler_tts.py:794-978 with past=None), logits of the
 * last position, then the step epilogue.  If all_logits != NULL, also writes logits of every position
 * [B*K][P+T0][vocab] (ParlerTTSForCausalLM.forward, modeling_parler_tts.py:1096-1185). */
PTTS_API int ptts_prefill(const PttsModel* m, const PttsState* st, const float* prompt_table, const int64_t* prompt_ids,
                 const float* prompt_hidden, float* all_logits, int32_t run_epilogue, ptts_stream_t s);
/* One cached decode step: 24 x (LN, QKV, self-attn, out_proj, LN, q, cross-attn, out_proj, LN, fc1+GELU, fc2), final
 * LN, 9 lm_heads, epilogue.  Replaces prepare_inputs_for_generation + forward + greedy_search/sample body
 * (modeling_parler_tts.py:2081-2148, 1927-2079, call sites 2582-2620).  Graph-capturable.  run_epilogue = 0 stops
 * after the logits so that user logits_processors can run before ptts_step_epilogue. */
PTTS_API int ptts_decode_step(const PttsModel* m, const PttsState* st, int32_t run_epilogue, ptts_stream_t s);
/* The fused per-step epilogue alone: logits processing + argmax / top-k sample + EOS/PAD freeze + token write +
 * delay-mask apply + next-step embedding gather-sum + position add. */
PTTS_API int ptts_step_epilogue(const PttsModel* m, const PttsState* st, ptts_stream_t s);

/* ---- DAC decode (dac_wrapper/modeling_dac.py:98-134; arithmetic of descript-audio-codec) ----------------------- */
/* z[b][c][f] = sum_i (out_proj_i . codebook_i[codes[b][i][f]])[c] + bias; tables[i] = folded [codebook_size][latent]
 * (codebook x weight-normalised out_proj, pre-multiplied at load).  z fp32 [B][latent][F]. */
PTTS_API int ptts_dac_dequant(const int64_t* codes, int32_t B, int32_t K, int32_t F, int64_t ld_b, int64_t ld_k,
                     const float* tables, const float* bias_sum, int32_t codebook_size, int32_t latent, float* z,
                     ptts_stream_t s);
/* y = [tanh]( conv1d( snake_alpha(x) or x, w, bias, dilation, same padding ) ) [+ residual]; w folded [Cout][Cin][k]. */
PTTS_API int ptts_dac_conv1d(const float* x, int32_t B, int32_t Cin, int32_t T, const float* w, const float* bias,
                    int32_t Cout, int32_t ksize, int32_t dilation, const float* snake_alpha, const float* residual,
                    int32_t tanh_out, float* y, int32_t impl, ptts_stream_t s);
/* y = conv_transpose1d(snake(x), w [Cin][Cout][2*stride], stride, padding ceil(stride/2)); stride in {2, 4, 8} (the
 * 44.1 kHz codec of the reference uses 8, 8, 4, 2; an odd stride such as the 5 of descript's 16 / 24 kHz models is rejected
 * with an argument error, not approximated) */
PTTS_API int ptts_dac_convtr1d(const float* x, int32_t B, int32_t Cin, int32_t T, const float* w, const float* bias,
                      int32_t Cout, int32_t stride, const float* snake_alpha, float* y, int32_t impl,
                      ptts_stream_t s);


/* ---- DAC decode, fast family: channels-last activations [B][T][C] fp32, tcgen05 implicit GEMM ------------------- */
/* fp32 semantics on the bf16 tensor core: every operand is split x = hi + mid (two bf16) and each product is three
 * MMAs.  Weights are pre-split and PRE-TILED on the host (parler_tts_b200/dac.py) into ONE buffer
 * [n_tile][Cin/32][taps][plane hi|mid][4][bn][8] bf16 with bn = ptts_dac_tile_n(N), N = Cout (conv1d) or stride*Cout
 * (convtr1d), rows n >= N zero-filled: the shared-memory image of one pipeline stage, moved by one bulk copy. */
PTTS_API int ptts_dac_tile_n(int32_t N);
/* CTAs the persistent conv kernels (ptts_dac_conv1d_px / ptts_dac_convtr1d_px) launched from now on may occupy; 0 = one
 * per SM (default).  Returns the previous value.  Host-side launch setting of the calling process (one process per GPU),
 * not synchronised: the chunked decode that runs beside the token loop (streaming.py) sets it around its launches so that
 * the loop's kernels keep SMs of their own.  The results do not depend on it. */
PTTS_API int ptts_dac_set_cta_limit(int32_t n);
/* "Planes": an activation tensor stored the way the tensor core wants it - Snake already applied, split into two
 * bf16 planes (x = hi + mid), 8-channel panels: [B][2 (hi, mid)][C/8][PAD_LO + T + PAD_HI][8] bf16 (4 bytes per
 * element, like fp32).  Row PAD_LO - 1 and row PAD_LO + T must be zero (the caller zeroes the pads once), the other pad
 * rows only need to exist.  A conv writes planes for its consumer (out_alpha = the consumer's Snake alpha, NULL = no
 * activation) from its otherwise idle epilogue warps, and the consumer then streams them with bulk copies instead of
 * activating its input again for every column tile (a k=1 conv, or a transposed conv with 6..48 column tiles).
 * Exactly one of x / x_planes; y and/or y_planes; snake_alpha only with x. */
#define PTTS_DAC_PLANE_PAD_LO 8
#define PTTS_DAC_PLANE_PAD_HI 264
PTTS_API int ptts_dac_conv1d_px(const float* x, const void* x_planes, int32_t B, int32_t T, int32_t Cin,
                                const void* w_tiled, const float* bias, int32_t Cout, int32_t ksize, int32_t dilation,
                                const float* snake_alpha, const float* residual, float* y, void* y_planes,
                                const float* out_alpha, ptts_stream_t s);
PTTS_API int ptts_dac_convtr1d_px(const float* x, const void* x_planes, int32_t B, int32_t T, int32_t Cin,
                                  const void* w_tiled, const float* bias, int32_t Cout, int32_t stride,
                                  const float* snake_alpha, float* y, ptts_stream_t s);
PTTS_API int ptts_dac_dequant_nwc(const int64_t* codes, int32_t B, int32_t K, int32_t F, int64_t ld_b, int64_t ld_k,
                         const float* tables, int32_t codebook_size, int32_t latent, float* z, ptts_stream_t s);
/* final conv (Cout = 1, k = 7) + tanh: w [7][Cin] fp32, y [B][T] */
PTTS_API int ptts_dac_conv_out_nwc(const float* x, int32_t B, int32_t T, int32_t Cin, const float* w, const float* bias,
                          const float* snake_alpha, float* y, ptts_stream_t s);

/* ---- DAC encode for audio-prompt continuation ------------------------------------------------ */
/* dac_wrapper/modeling_dac.py:25-96 (DACModel.encode -> descript DAC.encode), called by modeling_parler_tts.py:2247-2292.
 * The encoder's residual units run on ptts_dac_conv1d; these are the two remaining pieces. */
/* y[b][co][to] = bias + sum_ci,k w[co][ci][k] * snake_alpha(x)[b][ci][to*stride - pad + k];  x [B][Cin][T] fp32,
 * w folded [Cout][Cin][ksize], y [B][Cout][(T + 2 pad - ksize)/stride + 1]. */
PTTS_API int ptts_dac_conv1d_strided(const float* x, int32_t B, int32_t Cin, int32_t T, const float* w, const float* bias,
                                     int32_t Cout, int32_t ksize, int32_t stride, int32_t pad, const float* snake_alpha,
                                     float* y, ptts_stream_t s);
/* Residual vector quantisation of latents z [B][C][F]: for codebook i: e = in_proj_i(residual); code = nearest
 * l2-normalised codebook row (first index on ties); residual -= out_proj_i(codebook_i[code]).  w_in [K][D][C],
 * b_in [K][D], codebook / codebook_normalized [K][codebook_size][D], w_out [K][C][D], b_out [K][C] (weight norm folded);
 * codes [B][K][F] int64. */
PTTS_API int ptts_dac_vq_encode(const float* z, int32_t B, int32_t C, int32_t F, int32_t K, int32_t D, int32_t codebook_size,
                                const float* w_in, const float* b_in, const float* codebook,
                                const float* codebook_normalized, const float* w_out, const float* b_out, int64_t* codes,
                                ptts_stream_t s);

/* ---- T5 text encoder: the description encoder generate() runs once per call --------------- */
/* Replaces the third-party transformers T5EncoderModel called at modeling_parler_tts.py:2229
 * (_prepare_text_encoder_kwargs_for_generation, 2197-2241).  Weight layout = HF T5 state-dict keys
 * text_encoder.encoder.block.{i}.layer.0.SelfAttention.{q,k,v,o}.weight, ...layer.0.layer_norm.weight,
 * ...layer.1.DenseReluDense.{wi_0,wi_1,wo | wi,wo}.weight, ...layer.1.layer_norm.weight,
 * text_encoder.encoder.final_layer_norm.weight, text_encoder.shared.weight and block 0's
 * relative_attention_bias.weight. */
typedef struct {
  const void* w_qkv;   /* [3*inner][D]  q | k | v stacked, inner = n_heads * d_kv */
  const void* w_o;     /* [D][inner] */
  const void* w_wi;    /* gated: [2*d_ff][D] wi_0 | wi_1 stacked; else [d_ff][D] */
  const void* w_wo;    /* [D][d_ff] */
  const float *ln0, *ln1; /* RMS-norm weights in front of attention / feed-forward */
} PttsT5Layer;

typedef struct {
  int32_t n_layers, d_model, n_heads, d_kv, d_ff, gated, act_fn, wdtype;
  int32_t vocab, n_buckets, max_rel, reserved0;
  float eps, reserved1;
  const float* embed;       /* [vocab][D] fp32 */
  const float* rel_bias;    /* [n_buckets][n_heads] fp32 */
  const int32_t* bucket_of; /* [2*max_rel+1]: bucket of relative position (key - query) + max_rel, clamped outside */
  const float* ln_f;        /* final_layer_norm */
  const PttsT5Layer* layers; /* HOST array of n_layers entries */
} PttsT5;

PTTS_API int64_t ptts_t5_workspace_bytes(const PttsT5* m, int32_t rows);
/* ids [B][S] int64, mask [B][S] int32 (1 = keep) or NULL, out [B][S][D] T (last_hidden_state); workspace device
 * bytes >= ptts_t5_workspace_bytes(m, B*S). */
PTTS_API int ptts_t5_encode(const PttsT5* m, const int64_t* ids, const int32_t* mask, int32_t B, int32_t S, void* workspace,
                            void* out, int32_t impl, ptts_stream_t s);

#ifdef __cplusplus
}
#endif
#endif /* PARLER_B200_H */
