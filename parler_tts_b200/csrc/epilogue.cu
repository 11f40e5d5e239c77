// The fused per-step epilogue: ONE kernel per decode step that replaces, for every row r = b*K + k,
//   logits processors: repetition penalty, classifier-free guidance,           transformers logits_process
//     vocabulary cut, minimum length (EOS masked)                              (_get_logits_processor, :2561-2567)
//   greedy argmax | temperature + top-k + top-p + softmax + draw                transformers greedy_search / sample
//   next = next*unfinished + pad*(1-unfinished); append; unfinished &= next!=eos     (same)
//   apply_delay_pattern_mask on the new column                                  modeling_parler_tts.py:68-74, 2107
//   sum_k embed_tokens[k](input) + position row                                 modeling_parler_tts.py:836, 884-886
// (call sites /root/reference/parler_tts/modeling_parler_tts.py:2582-2620).  One block per utterance, one warp per
// codebook row; all sequence-length state lives in the device control block so the launch is graph-replayable.
#include "ops.cuh"

namespace ptts {

__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                              uint32_t k1, uint32_t out[4]) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

__device__ __forceinline__ uint32_t order_key(float f) {
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

// warp-wide log-sum-exp helpers over a row in shared memory
__device__ __forceinline__ float row_max(const float* s, int V, int lane) {
  float m = -INFINITY;
  for (int i = lane; i < V; i += 32) m = fmaxf(m, s[i]);
  return warp_max(m);
}

__global__ void __launch_bounds__(32 * PTTS_MAX_CODEBOOKS)
step_epilogue_kernel(const float* __restrict__ embed, const float* __restrict__ pos, int K, int V, int H, int Bu,
                     int cfg, float guidance, int P, int T0, int M, int do_sample, int top_k, float temperature,
                     int suppress_from, int min_length, float top_p, float rep_penalty, int premixed, int64_t row_offset,
                     int64_t pad_id, int64_t eos_id,
                     PttsCtrl* __restrict__ ctrl, const float* __restrict__ logits, int64_t* __restrict__ ids,
                     const int64_t* __restrict__ mask, int* __restrict__ unfinished, float* __restrict__ h,
                     int h_rows_per_batch, int nslice, int64_t slice_stride) {
  extern __shared__ float smem[];
  __shared__ int64_t next_in[PTTS_MAX_CODEBOOKS];
  __shared__ int transitions;
  // Launched with the programmatic-serialization attribute: the CTAs are resident when the lm_heads GEMM finishes.  The
  // control block was last written by the PREVIOUS step's epilogue (this kernel never releases its dependent early, so the
  // next step's first kernel starts only when this grid has completed): safe to read before the wait.
  if (ctrl->done) return;
  const int len = ctrl->len;
  if (len >= M) return;
  pdl_wait();
  const int b = blockIdx.x;
  const int k = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) transitions = 0;
  __syncthreads();
  float* sc = smem + (size_t)k * V;        // this warp's score row
  // compacted kept indices (sampling only); with top-p the warp's region is 2 V words: indices | probabilities
  int* kept = (int*)(smem + (size_t)K * V) + (size_t)k * V * (top_p < 1.0f ? 2 : 1);
  const int r = b * K + k;

  // ---- logits processors, in the order transformers applies them (defaults first, then the caller's list with the
  // guidance processor last, modeling_parler_tts.py:2556-2567)
  // nslice > 1: the logits arrive as split-K partial sums [nslice][B*K][V] of the heads GEMM (decode step of the captured
  // graph) and are added up here in slice order (deterministic); nslice == 1: the plain logits matrix
  const float* lc = logits + (int64_t)r * V;
  // (all slice loads of an element are issued before the first add: with a rolled loop every load waits for the one
  // before it - measured 17 us per step)
  auto raw = [&](const float* row, int i) {
    float t[4];
#pragma unroll
    for (int sl = 0; sl < 4; ++sl) t[sl] = sl < nslice ? __ldcg(row + (int64_t)sl * slice_stride + i) : 0.f;
    return ((t[0] + t[1]) + t[2]) + t[3];
  };
  const int mix = cfg && !premixed;  // premixed: rows [0, Bu*K) already hold the guidance-mixed scores (host processors)
  if ((V & 3) == 0 && (slice_stride & 3) == 0) {
    // 16-byte loads, three column groups x four slices in flight per lane: the 1088-wide row folds in three L2 round trips
    const float4* l4 = reinterpret_cast<const float4*>(lc);
    const int64_t st4 = slice_stride >> 2;
    float4* sc4 = reinterpret_cast<float4*>(sc);
#pragma unroll 3
    for (int i = lane; i < (V >> 2); i += 32) {
      float4 t[4];
#pragma unroll
      for (int sl = 0; sl < 4; ++sl) t[sl] = sl < nslice ? __ldcg(l4 + sl * st4 + i) : make_float4(0.f, 0.f, 0.f, 0.f);
      sc4[i] = make_float4(((t[0].x + t[1].x) + t[2].x) + t[3].x, ((t[0].y + t[1].y) + t[2].y) + t[3].y,
                           ((t[0].z + t[1].z) + t[2].z) + t[3].z, ((t[0].w + t[1].w) + t[2].w) + t[3].w);
    }
  } else {
#pragma unroll 4
    for (int i = lane; i < V; i += 32) sc[i] = raw(lc, i);
  }
  __syncwarp();
  if (rep_penalty != 1.0f) {
    // RepetitionPenaltyLogitsProcessor over this row's own raw history (with guidance: the conditional rows only, which
    // is what gather/scatter over (B*K, len) ids does to the (2*B*K, V) scores).  Every occurrence of a token writes
    // the same value computed from the UNmodified logit, so duplicates are penalised once.
    const int64_t* hist = ids + (int64_t)r * M;
    for (int i = lane; i < len; i += 32) {
      const int64_t t = hist[i];
      if (t >= 0 && t < V) {
        const float v = raw(lc, (int)t);
        sc[t] = v < 0.f ? v * rep_penalty : v / rep_penalty;
      }
    }
    __syncwarp();
  }
  if (mix) {
    const float* lu = logits + ((int64_t)(b + Bu) * K + k) * V;
    float mc = -INFINITY, mu = -INFINITY;
    for (int i = lane; i < V; i += 32) { mc = fmaxf(mc, sc[i]); mu = fmaxf(mu, raw(lu, i)); }
    mc = warp_max(mc); mu = warp_max(mu);
    float sc_ = 0.f, su = 0.f;
    for (int i = lane; i < V; i += 32) { sc_ += expf(sc[i] - mc); su += expf(raw(lu, i) - mu); }
    sc_ = warp_sum(sc_); su = warp_sum(su);
    const float lzc = mc + logf(sc_), lzu = mu + logf(su);
    for (int i = lane; i < V; i += 32) {
      const float c = sc[i] - lzc, u = raw(lu, i) - lzu;
      sc[i] = u + guidance * (c - u);
    }
  }
  if (suppress_from >= 0)
    for (int i = suppress_from + lane; i < V; i += 32) sc[i] = -INFINITY;
  // MinLengthLogitsProcessor / MinNewTokensLengthLogitsProcessor: no EOS while the sequence is shorter than min_length
  // columns.  (Applied after the guidance mix: the reference masks both halves first and its mix then yields NaN.)
  if (lane == 0 && len < min_length && eos_id >= 0 && eos_id < V) sc[eos_id] = -INFINITY;
  __syncwarp();

  // ---- token selection
  int64_t tok;
  if (!do_sample) {
    float bv = -INFINITY; int bi = 0x7fffffff;
#pragma unroll 8
    for (int i = lane; i < V; i += 32) {
      const float v = sc[i];
      if (v > bv || (v == bv && i < bi)) { bv = v; bi = i; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
    }
    tok = bi;
  } else {
    if (temperature != 1.0f) {
      for (int i = lane; i < V; i += 32) sc[i] = sc[i] / temperature;
      __syncwarp();
    }
    // k-th largest ordered key by MSB-first search: keep every score >= it (ties kept, TopKLogitsWarper)
    uint32_t thr = 0u;
    const int kk = top_k > 0 ? min(top_k, V) : V;
    if (top_k > 0) {
      constexpr int KREG = 36;  // ordered keys a lane keeps in registers for the 32 search steps: rows of up to 1152 logits
      if (V <= KREG * 32) {
        uint32_t key[KREG];
#pragma unroll
        for (int t = 0; t < KREG; ++t) {
          const int i = lane + 32 * t;
          key[t] = i < V ? order_key(sc[i]) : 0u;  // 0 is below every candidate: slots past the row never count
        }
#pragma unroll 1
        for (int bit = 31; bit >= 0; --bit) {
          const uint32_t cand = thr | (1u << bit);
          int cnt = 0;
#pragma unroll
          for (int t = 0; t < KREG; ++t) cnt += key[t] >= cand;
          cnt = __reduce_add_sync(0xffffffffu, cnt);
          if (cnt >= kk) thr = cand;
        }
      } else {
        for (int bit = 31; bit >= 0; --bit) {
          const uint32_t cand = thr | (1u << bit);
          int cnt = 0;
          for (int i = lane; i < V; i += 32) cnt += order_key(sc[i]) >= cand;
          cnt = __reduce_add_sync(0xffffffffu, cnt);
          if (cnt >= kk) thr = cand;
        }
      }
    }
    // ordered compaction of the kept set
    int nk = 0;
    for (int i0 = 0; i0 < V; i0 += 32) {
      const int i = i0 + lane;
      const bool keep = i < V && order_key(sc[i]) >= thr;
      const unsigned bal = __ballot_sync(0xffffffffu, keep);
      if (keep) kept[nk + __popc(bal & ((1u << lane) - 1u))] = i;
      nk += __popc(bal);
    }
    __syncwarp();
    const float mx = row_max(sc, V, lane);
    if (top_p < 1.0f) {
      // TopPLogitsWarper (min_tokens_to_keep 1): with the candidates sorted ascending, drop those whose cumulative
      // probability stays <= 1 - top_p.  The dropped set is a down-set in score order, so it is found as the largest
      // ordered-key threshold T with sum{p_y : key_y < T} <= 1 - top_p (32-step bitwise search), keep key >= T.
      float* prob = reinterpret_cast<float*>(kept) + V;  // third smem row of this warp (sampling with top_p only)
      float tot = 0.f;
      for (int j = lane; j < nk; j += 32) {
        const float e = expf(sc[kept[j]] - mx);
        prob[j] = e;
        tot += e;
      }
      tot = warp_sum(tot);
      __syncwarp();
      const float budget = (1.0f - top_p) * tot;
      uint32_t T = 0u;
      for (int bit = 31; bit >= 0; --bit) {
        const uint32_t cand = T | (1u << bit);
        float below = 0.f;
        for (int j = lane; j < nk; j += 32)
          if (order_key(sc[kept[j]]) < cand) below += prob[j];
        below = warp_sum(below);
        if (below <= budget) T = cand;
      }
      int nk2 = 0;
      for (int j0 = 0; j0 < nk; j0 += 32) {  // in-place ordered compaction (reads run ahead of writes)
        const int j = j0 + lane;
        const int idx = j < nk ? kept[j] : 0;
        const bool keep = j < nk && order_key(sc[idx]) >= T;
        const unsigned bal = __ballot_sync(0xffffffffu, keep);
        __syncwarp();
        if (keep) kept[nk2 + __popc(bal & ((1u << lane) - 1u))] = idx;
        nk2 += __popc(bal);
      }
      nk = nk2;
      __syncwarp();
    }
    int choice = 0;
    uint32_t rnd[4];
    const uint64_t grow = (uint64_t)(row_offset + r);
    // the Philox key lives in the control block so that a captured graph can be replayed with a new seed
    philox4x32_10((uint32_t)grow, (uint32_t)(len - T0), 0u, 0u, (uint32_t)ctrl->reserved[0], (uint32_t)ctrl->reserved[1],
                  rnd);
    const float u = (float)(((double)rnd[0] + 0.5) * 2.3283064365386963e-10);
    if (nk <= 64 && V >= 64) {
      // the usual top-k case.  The sums keep the fixed ascending-index fp32 order (the draw must not depend on the launch
      // geometry), but the exponentials are taken by all lanes at once and the serial part is a chain of adds over a
      // contiguous strip: every lane runs it (broadcast reads) and notes the prefix sums of its own two candidates
      const int k0 = lane < nk ? kept[lane] : 0, k1 = lane + 32 < nk ? kept[lane + 32] : 0;
      const float e0 = lane < nk ? expf(sc[k0] - mx) : 0.f, e1 = lane + 32 < nk ? expf(sc[k1] - mx) : 0.f;
      __syncwarp();  // every score has been read: the head of the row becomes the strip
      sc[lane] = e0;
      sc[lane + 32] = e1;
      __syncwarp();
      float acc = 0.f, p0 = 0.f, p1 = 0.f;
      for (int j = 0; j < nk; ++j) {
        acc += sc[j];
        if (j == lane) p0 = acc;
        if (j == lane + 32) p1 = acc;
      }
      const float thresh = u * acc;
      const unsigned b0 = __ballot_sync(0xffffffffu, lane < nk && p0 > thresh);
      const unsigned b1 = __ballot_sync(0xffffffffu, lane + 32 < nk && p1 > thresh);
      const int pick = b0 ? __ffs(b0) - 1 : (b1 ? 32 + __ffs(b1) - 1 : nk - 1);
      choice = pick < 32 ? __shfl_sync(0xffffffffu, k0, pick & 31) : __shfl_sync(0xffffffffu, k1, pick & 31);
    } else {
      if (lane == 0) {
        float total = 0.f;
        for (int j = 0; j < nk; ++j) total += expf(sc[kept[j]] - mx);  // fixed ascending-index fp32 order
        const float thresh = u * total;
        float acc = 0.f;
        choice = kept[nk - 1];
        for (int j = 0; j < nk; ++j) {
          acc += expf(sc[kept[j]] - mx);
          if (acc > thresh) { choice = kept[j]; break; }
        }
      }
      choice = __shfl_sync(0xffffffffu, choice, 0);
    }
    tok = choice;
  }

  // ---- EOS / PAD bookkeeping, token write, delay mask for the next input
  if (lane == 0) {
    const int unf = unfinished[r];
    tok = unf ? tok : pad_id;
    ids[(int64_t)r * M + len] = tok;
    const int unf2 = unf && (tok != eos_id);
    unfinished[r] = unf2;
    if (unf && !unf2) atomicAdd(&transitions, 1);
    const int64_t mv = (len < M) ? mask[(int64_t)r * M + len] : -1;
    next_in[k] = (mv == -1) ? tok : mv;
  }
  __syncthreads();

  // ---- next-step hidden state: sum_k embed[k][input_k] + pos[P + len]   (written for both CFG halves)
  if (len + 1 < M) {
    const float* prow = pos + (int64_t)(P + len) * H;
    if ((H & 3) == 0) {
      // all K embedding rows of a column group are requested before the first add (with a rolled loop every load waited
      // for the one before it: up to K dependent HBM round trips per column - the epilogue took ~20 us of the step);
      // the adds keep the order k = 0 .. K-1, then the position row
      for (int c4 = threadIdx.x; c4 < H / 4; c4 += blockDim.x) {
        float4 t[PTTS_MAX_CODEBOOKS];
#pragma unroll
        for (int kk2 = 0; kk2 < PTTS_MAX_CODEBOOKS; ++kk2)
          if (kk2 < K)
            t[kk2] = __ldg(reinterpret_cast<const float4*>(embed + ((int64_t)kk2 * (V + 1) + next_in[kk2]) * H) + c4);
        const float4 pr = __ldg(reinterpret_cast<const float4*>(prow) + c4);
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int kk2 = 0; kk2 < PTTS_MAX_CODEBOOKS; ++kk2)
          if (kk2 < K) {
            acc.x += t[kk2].x; acc.y += t[kk2].y; acc.z += t[kk2].z; acc.w += t[kk2].w;
          }
        const float4 v = make_float4(acc.x + pr.x, acc.y + pr.y, acc.z + pr.z, acc.w + pr.w);
        reinterpret_cast<float4*>(h + (int64_t)b * h_rows_per_batch * H)[c4] = v;
        if (cfg) reinterpret_cast<float4*>(h + (int64_t)(b + Bu) * h_rows_per_batch * H)[c4] = v;
      }
    } else {
      for (int c = threadIdx.x; c < H; c += blockDim.x) {
        float acc = 0.f;
        for (int kk2 = 0; kk2 < K; ++kk2) acc += embed[((int64_t)kk2 * (V + 1) + next_in[kk2]) * H + c];
        const float v = acc + prow[c];
        h[(int64_t)b * h_rows_per_batch * H + c] = v;
        if (cfg) h[(int64_t)(b + Bu) * h_rows_per_batch * H + c] = v;
      }
    }
  }

  // ---- global bookkeeping: the last block to arrive advances len and decides termination
  if (threadIdx.x == 0) {
    if (transitions) atomicSub(&ctrl->n_unfinished, transitions);
    __threadfence();
    const int t = atomicAdd(&ctrl->ticket, 1);
    if (t == (int)gridDim.x - 1) {
      __threadfence();
      const int nl = len + 1;
      const int nu = atomicAdd(&ctrl->n_unfinished, 0);
      ctrl->ticket = 0;
      ctrl->steps += 1;
      if (nu == 0 || nl >= M) { ctrl->stop_len = nl; ctrl->done = 1; }
      ctrl->len = nl;
    }
  }
}

int step_epilogue(const PttsModel* m, const PttsState* s, cudaStream_t st, const float* logit_parts, int nslice,
                  int64_t slice_stride) {
  const int K = m->n_codebooks, V = m->vocab;
  const int cfg = s->guidance_scale > 1.0f;
  PTTS_CHECK_ARG(!cfg || (s->B % 2 == 0), "step_epilogue: guidance needs an even model batch");
  const int Bu = cfg ? s->B / 2 : s->B;
  const bool nucleus = s->do_sample && s->top_p > 0.0f && s->top_p < 1.0f;
  // per codebook row: scores | kept indices + (top-p) candidate probabilities
  const size_t smem = (size_t)K * V * sizeof(float) * (s->do_sample ? (nucleus ? 3 : 2) : 1);
  if (smem > 48 * 1024) {
    static size_t configured = 0;
    if (smem > configured) {
      cudaFuncSetAttribute(step_epilogue_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      configured = smem;
    }
  }
  PTTS_CHECK_ARG(smem <= 200 * 1024, "step_epilogue: K*vocab too large");
  // the hidden buffer holds one row per utterance during decode (rows_per_batch = 1)
  const cudaError_t e = launch_k(step_epilogue_kernel, dim3(Bu), dim3(32 * K), smem, st, m->embed, m->pos, K, V, m->hidden,
                                 Bu, cfg, s->guidance_scale, s->P, s->T0, s->max_len, s->do_sample, s->top_k,
                                 s->temperature, s->suppress_from, s->min_length, nucleus ? s->top_p : 1.0f,
                                 s->repetition_penalty > 0.0f ? s->repetition_penalty : 1.0f, s->logits_premixed,
                                 s->row_offset, s->pad_id, s->eos_id, s->ctrl, logit_parts ? logit_parts : s->logits,
                                 s->ids, s->mask, s->unfinished, s->h, 1, logit_parts ? nslice : 1,
                                 logit_parts ? slice_stride : (int64_t)0);
  if (e != cudaSuccess) {
    set_error("step_epilogue launch: %s", cudaGetErrorString(e));
    (void)cudaGetLastError();
    return PTTS_E_LAUNCH;
  }
  PTTS_CHECK_LAUNCH("step_epilogue");
  return PTTS_OK;
}

}  // namespace ptts

extern "C" int ptts_step_epilogue(const PttsModel* m, const PttsState* st, ptts_stream_t s) {
  PTTS_CHECK_ARG(m && st, "step_epilogue: null descriptor");
  PTTS_CHECK_ARG(m->n_codebooks <= PTTS_MAX_CODEBOOKS, "step_epilogue: too many codebooks");
  return ptts::step_epilogue(m, st, ptts::as_stream(s));
}
