// ViT attention with decomposed relative-position bias (reference: Attention.forward image_encoder.py:235-260,
// get_rel_pos :321-351, add_decomposed_rel_pos :354-392): dispatch to the two tcgen05 kernels.
//   S == 14  windowed blocks   -> attn_tc.cu         (one (window, head) problem at a time, scores stay in TMEM)
//   S == 64  global blocks     -> attn_global_tc.cu  (flash-style over 128-key tiles) fed by the batched tcgen05 GEMM
//                                  G = Q [Rh; Rw]^T (fp16) for the bias projections
#include "common.cuh"
#include "gemm_tc.cuh"
#include "kernels.cuh"

namespace bsam {

cudaError_t attention_init() { return cudaSuccess; }

namespace {
std::atomic<int> g_variant[VK_COUNT];   // zero-initialised: variant 0 everywhere, except the GEMM pair switch below
std::atomic<bool> g_pair_set{false};
std::atomic<void*> g_trace{nullptr};
}  // namespace

void* debug_trace_buffer() { return g_trace.load(); }
void set_debug_trace_buffer(void* p) { g_trace.store(p); }

int kernel_variant(int which) {
  if (which < 0 || which >= VK_COUNT) return 0;
  if (which == VK_GEMM_PAIR && !g_pair_set.load()) return 1;   // CTA-pair kernel is the default for the big GEMMs
  return g_variant[which].load(std::memory_order_relaxed);
}
void set_kernel_variant(int which, int value) {
  if (which < 0 || which >= VK_COUNT) return;
  if (which == VK_GEMM_PAIR) g_pair_set.store(true);
  g_variant[which].store(value, std::memory_order_relaxed);
}

cudaError_t launch_attention(const __nv_bfloat16* qkv, int nbatch, int S, int heads, int hd,
                             const __nv_bfloat16* relh, const __nv_bfloat16* relw, const __nv_bfloat16* relg,
                             void* gbuf, __nv_bfloat16* out, cudaStream_t s, bool window_token_out,
                             int* f16_overflow) {
  if (nbatch <= 0) return cudaSuccess;
  if (hd != 64 && hd != 80) return cudaErrorInvalidValue;
  if (S == kWin) {
    // the two padded tables must be contiguous: [relh (32 rows); relw (32 rows)]
    if (relw != relh + 32 * hd) return cudaErrorInvalidValue;
    return launch_attention_window_tc(qkv, nbatch, heads, hd, relh, out, s, window_token_out);
  }
  if (S == kGrid) {
    if (!relg || !gbuf) return cudaErrorInvalidValue;
    // G[row, head, 0:256] = q[row, head, :] . [Rh ; Rw]^T for every head in one batched launch: the A tile of head h
    // is the 128-column slice starting at h*hd of the q part of the row; relg is zero beyond column hd.
    GemmEpi e;
    e.out = gbuf;
    e.ld_out = heads * 256;
    e.batch = heads;
    e.a_batch_off = hd;
    e.out_batch_off = 256;
    // w part stored as the 64 biases of the row's own qx, in units of the softmax scale (x sqrt(hd)): the attention
    // kernel adds them to q.k on the tensor core (S += Gw E) before scaling by hd^-1/2
    e.rowmap = RM_RELW_SHIFT;
    e.relw_scale = sqrtf((float)hd);
    e.f16_overflow = f16_overflow;   // fp16 range check of the projections (see GemmEpi)
    cudaError_t err = gemm_tc(false, GM_F16, qkv, 3 * heads * hd, relg, 128, nbatch * kTokens, 256, 128, e, s);
    if (err != cudaSuccess) return err;
    return launch_attention_global_tc(qkv, nbatch, heads, hd, reinterpret_cast<const __half*>(gbuf), out, s);
  }
  return cudaErrorInvalidValue;
}

}  // namespace bsam
