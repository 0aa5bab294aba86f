/* libb200sam - test / measurement hooks.  NOT part of the drop-in surface of include/b200sam.h: nothing the
 * reference's callers need is declared here, INTEGRATION.md does not use it, and a deployment may strip these symbols.
 * Used by tests/ (per-block parity taps, SIMT bisecting) and scripts/ (kernel A/B measurements). */
#ifndef B200SAM_DEBUG_H_
#define B200SAM_DEBUG_H_

#include "b200sam.h"

#ifdef __cplusplus
extern "C" {
#endif

/* 1 = evaluate every GEMM of this handle with the one-thread-per-row SIMT kernel instead of tcgen05 (bisecting only). */
int b200sam_set_debug_simt_gemm(b200sam_t* h, int on);

/* Make b200sam_encode stop after `nblocks` transformer blocks (-1 = run everything, the default) and copy the fp32
 * residual stream (B,64,64,D) out; used by the per-block parity tests against the reference's forward-hook taps. */
int b200sam_debug_stop_after_block(b200sam_t* h, int nblocks);
int b200sam_debug_read_tokens(b200sam_t* h, float* out, int B, void* stream);

/* Process-wide kernel variant switches for same-box A/B measurements (scripts/bench_attn.py).  which: 0 = windowed
 * attention kernel (0 shipped, 9 round-1 kernel, 4 shipped + L2 prefetch, 2 / 3 phase trace), 1 = global attention
 * kernel (0 shipped, 1 round-1 softmax arithmetic, 2 phase trace), 2 = big-GEMM kernel (0 single CTA, 1 CTA pair = default), 3 = the encoder's two
 * convolutions (0 = A tiles gathered by TMA from the activation = default, 1 = im2col copy + plain GEMM), 4 = preprocess kernel (0 = shared-memory staged
 * second generation = default, 1 = the round-1 kernel), 5 = LayerNorm (0 = folded into the qkv / lin1 GEMMs = default,
 * 1 = stand-alone kernel), 6 = programmatic dependent launch of the encoder kernels (0 = off = default, 1 = on),
 * 7 = window-pad rows of the qkv buffer (0 = written by the idle warp of the previous block's lin2 GEMM = default,
 * 1 = by the fill kernel behind every windowed qkv GEMM).
 * An unknown
 * variant is an error at launch time, never a fallback. */
int b200sam_debug_set_variant(int which, int value);

/* Device buffer (>= 16 * 16 * 10 uint64, zeroed by the caller) for the tracing kernel variants: CTA 0 records
 * clock64() at the phase boundaries of its first 16 work items, indexed [warp][item][tag].  NULL switches tracing off. */
int b200sam_debug_set_trace(void* dev_buffer);

/* Time `iters` launches of the attention op (same inputs as b200sam_op_attention) with CUDA events on `stream`,
 * tables packed once outside the timed region; *host_ms_out = average milliseconds per launch (S = 64: the G GEMM +
 * the attention kernel). */
int b200sam_debug_attention_bench(const void* qkv_bf16, int nbatch, int S, int heads, int hd, const float* rel_h,
                                  const float* rel_w, void* out_bf16, int iters, float* host_ms_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200SAM_DEBUG_H_ */
