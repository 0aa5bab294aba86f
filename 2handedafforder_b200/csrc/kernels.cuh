// Launchers of the non-GEMM kernels of the SAM branch (elementwise.cu, attn.cu, decoder.cu, postprocess.cu).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace bsam {

enum DType : int { DT_F32 = 0, DT_BF16 = 1, DT_F16 = 2 };

// ---- elementwise.cu
cudaError_t launch_layernorm_bf16(const float* x, long rows, int D, const float* g, const float* b, float eps,
                                  __nv_bfloat16* out, bool windowed, cudaStream_t s);
cudaError_t launch_im2col_patch(const void* frames, int dtype, int B, __nv_bfloat16* out, cudaStream_t s);
cudaError_t launch_im2col_3x3(const __nv_bfloat16* z, int B, __nv_bfloat16* out, cudaStream_t s);
cudaError_t launch_fold_ln_weights(const float* Wt, const float* gamma, const float* beta, const float* bias, int N, int K,
                                   __nv_bfloat16* Wf, float* c, float* d, cudaStream_t s);
cudaError_t launch_tokens_to_nchw(const float* tok, int B, void* out, int dtype, cudaStream_t s);
cudaError_t launch_nchw_to_tokens(const void* in, int dtype, int B, float* tok, cudaStream_t s);
// window-pad rows of the packed qkv buffer (B*4900, 3D) := bias (qkv of a zero row, image_encoder.py:179-183 pads after norm1)
cudaError_t launch_fill_pad_rows(__nv_bfloat16* qkv, const float* bias, const int* pad_rows, int n_pad, int B,
                                 int cols, cudaStream_t s);
cudaError_t launch_cast_f32_bf16(const float* in, __nv_bfloat16* out, long n, cudaStream_t s);
cudaError_t launch_to_f32(const void* in, int dtype, float* out, long n, cudaStream_t s);
cudaError_t launch_dense_pe(const float* gauss, float* pe_tok, cudaStream_t s);
cudaError_t launch_keys_init(const float* emb_tok, const float* no_mask, const int* prompt_frame, int P, float* keys,
                             cudaStream_t s);
// rel-pos table (2S-1, hd) f32 -> bf16 padded to RP rows (32 for S=14, 128 for S=64), extra rows zero
cudaError_t launch_relpos_pack(const float* rel, int S, int hd, __nv_bfloat16* out, cudaStream_t s);
cudaError_t launch_relpos_pack_g(const float* rel_h, const float* rel_w, int hd, __nv_bfloat16* out, cudaStream_t s);

// ---- attn.cu
// qkv: (nbatch*S*S, 3*heads*hd) bf16 packed [q | k | v] per row, heads-major inside each; out (rows, heads*hd).
// relh/relw: bf16 tables from launch_relpos_pack.
// S == 64 additionally needs relg = (256, 128) bf16 [rel_pos_h (127 rows) ; 0 ; rel_pos_w REVERSED (127 rows) ; 0], columns
// >= hd zero (launch_relpos_pack_g), and gbuf = (nbatch*4096, heads*256) fp16 scratch for the bias projections.
cudaError_t launch_attention(const __nv_bfloat16* qkv, int nbatch, int S, int heads, int hd,
                             const __nv_bfloat16* relh, const __nv_bfloat16* relw, const __nv_bfloat16* relg,
                             void* gbuf, __nv_bfloat16* out, cudaStream_t s, bool window_token_out = false,
                             int* f16_overflow = nullptr);
cudaError_t attention_init();
// ---- attn_tc.cu (tcgen05): windowed attention; rel = (64, hd) bf16 [rel_pos_h padded to 32 rows; rel_pos_w padded]
cudaError_t launch_attention_global_tc(const __nv_bfloat16* qkv, int nbatch, int heads, int hd, const __half* G,
                                       __nv_bfloat16* out, cudaStream_t s);
// token_out: rows are written in token order (b*4096 + y*64 + x, window-pad rows dropped) instead of window order
cudaError_t launch_attention_window_tc(const __nv_bfloat16* qkv, int nwin, int heads, int hd, const __nv_bfloat16* rel,
                                       __nv_bfloat16* out, cudaStream_t s, bool token_out = false);

// Kernel variant switches for same-box A/B measurements (include/b200sam_debug.h); 0 = the shipped default.
enum VariantKind : int { VK_ATTN_WINDOW = 0, VK_ATTN_GLOBAL = 1, VK_GEMM_PAIR = 2, VK_IMPLICIT_CONV = 3, VK_PREPROCESS = 4, VK_LN_FOLD = 5, VK_PDL = 6, VK_PAD_FILL = 7, VK_COUNT = 8 };
int kernel_variant(int which);
void set_kernel_variant(int which, int value);
// device buffer of at least 16 * 16 * 10 uint64 that the tracing kernel variants fill (null: tracing variants refuse)
void* debug_trace_buffer();

// Launch with the programmatic-dependent-launch attribute (common.cuh griddep_*) when variant 6 is on.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = kernel_variant(VK_PDL) == 1 ? 1 : 0;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
void set_debug_trace_buffer(void* dev_ptr);

// ---- decoder.cu
struct DecSideWeights;   // device pointers of one mask decoder (api.cu owns it)
struct DecoderWorkspace;

// ---- postprocess.cu
// one side: any of logits (n,H,W) f32 / mask (n,H,W) u8 / bits (n,H,ceil(W/8)) packed (numpy.packbits order) may be null
cudaError_t launch_postprocess(const float* lowres, int n, int in_h, int in_w, int out_h, int out_w, float* logits,
                               uint8_t* mask, uint8_t* bits, float thr, cudaStream_t s);
// both sides of n prompts in one pass -> packed bits; with gt != null also the taxonomy-gated evaluation counts
// (n,3) {intersection, union, area} (zeroed here) and the packed gated prediction
cudaError_t launch_postprocess_pair(const float* low_left, const float* low_right, int n, int in_h, int in_w, int out_h,
                                    int out_w, uint8_t* bits_left, uint8_t* bits_right, float thr,
                                    const float* taxonomy, const uint8_t* gt, const int* gt_index, int64_t* counts,
                                    uint8_t* pred_bits, cudaStream_t s);
// out[i] = position-sensitive 64-bit checksum of data[i*stride, i*stride + bytes) (zeroed here)
cudaError_t launch_checksum64(const uint8_t* data, int n_items, long bytes, long stride, uint64_t* out,
                              cudaStream_t s);

}  // namespace bsam
