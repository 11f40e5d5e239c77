/*
 * parler_b200_diag.h - exports that exist ONLY in the diagnostic build libparler_b200_diag.so (`make -C
 * parler_tts_b200/csrc DIAG=1`).  The product library libparler_b200.so has none of them, reads no environment variable
 * and compiles the ablation switches out; tests/, bench.py and __graft_entry__.py never load the diagnostic build.
 * Users: profiles/scripts/*.py (timelines, per-phase stamps, launch-latency floor) with
 * PTTS_LIBRARY=parler_tts_b200/libparler_b200_diag.so.
 */
#ifndef PARLER_B200_DIAG_H
#define PARLER_B200_DIAG_H

#include "parler_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Enqueues a chain of n dependent trivial kernels (mode 0: empty, mode 1: one load + store per thread over
 * buf[ctas*128]) so that the per-launch latency floor of a graph replay can be measured. */
PTTS_API int ptts_diag_chain(int32_t n, int32_t mode, int32_t ctas, float* buf, ptts_stream_t s);
/* globaltimer stamps of the tcgen05 GEMM recorded under PTTS_GEMM_DBG=1 (8 x uint64 per launch). */
PTTS_API int ptts_diag_gemm_stamps(uint64_t* host, int32_t n_launches);
/* PTTS_XB_DBG=1: 8 x uint64 per-phase SM-clock sums of the fused cross-attention block; reset != 0 clears them. */
PTTS_API int ptts_diag_xblock(uint64_t* host, int32_t reset);
/* PTTS_TIMELINE=1: per-launch globaltimer slots {first start, first/last dependency resolved, last end} of the
 * fast-family kernels.  cmd 0 = reset, cmd 1 = read (4 x uint64 per slot + launch-site tag); returns #slots. */
PTTS_API int ptts_diag_timeline(int32_t cmd, uint64_t* host, int32_t* kinds, int32_t max_slots);

#ifdef __cplusplus
}
#endif
#endif /* PARLER_B200_DIAG_H */
