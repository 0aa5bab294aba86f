// Interface of the tcgen05 GEMM family (gemm_tc.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace bsam {

// Epilogue selector.  C[m, n] = sum_k A[m, k] * W[n, k]  (both operands K-major, i.e. nn.Linear layout),
// then per mode:
enum GemmMode : int {
  GM_BIAS_BF16 = 0,       // out_bf16[m, n] = acc + bias[n]
  GM_BIAS_GELU_BF16 = 1,  // out_bf16[m, n] = gelu_erf(acc + bias[n])
  GM_F32 = 2,             // out_f32[orow(m), n] = acc (+ bias[n]) (+ addend[arow, n]); row maps below
  GM_LN = 3,              // N == 256: y = LayerNorm_n(acc (+bias) (+addend[m])) * g + b -> out_f32 and/or out2_bf16
  GM_UPS1 = 4,            // N == 256 = 4 x 64: per 64-group LN (eps) -> GELU -> pixel-shuffle store (ConvT k2 s2)
  GM_UPS2 = 5,            // N == 128 = 4 x 32: GELU -> dot with hyper[p, 0:32] -> pixel-shuffle store of the mask logit
  GM_F16 = 6,             // out_f16[m, n] = acc (batched over column slices of A: rel-pos projections of every head)
  GM_F32_FOLD = 7,        // GM_F32 + the producer side of the folded LayerNorm (bf16 copy, partial row sums); bf16 operands only
};

enum RowMap : int {
  RM_IDENT = 0,
  RM_WINDOW_TO_TOKEN = 1,  // input rows are in 14x14-window order (25 windows x 196 per frame, pad rows skipped)
  RM_RELW_SHIFT = 3,       // fp16 mode, global rel-pos projections: columns >= 128 of token row r hold q.Rw[126-j];
                           // they are stored shifted by qx = r % 64 so that out[128 + kx] = q.Rw[qx - kx + 63], kx < 64
  RM_TOKEN_TO_PAD66 = 4,   // GM_LN: the bf16 copy (out2) of token row b*4096 + y*64 + x goes to row b*4356 + (y+1)*66 + x+1
                           // (the zero-ringed layout AM_CONV3 reads); the fp32 output, if any, stays in token order
  RM_TOKEN_TO_WINDOW = 2,  // bf16 modes: input rows in token order (B*4096), output rows scattered into 14x14-window order (B*4900)
};

// How the A operand is addressed.  The two convolutions of the encoder are GEMMs whose A tiles are gathered by the TMA
// engine from the activation itself (no im2col copy): multi-dimensional tensor maps turn a (patch, k) / (token, tap, c)
// tile into one or four box loads.
enum AMode : int {
  AM_PLAIN = 0,    // A is a row-major (M, K) matrix
  AM_PATCH16 = 1,  // PatchEmbed Conv2d(k16, s16) (image_encoder.py:395-426): A = bf16 frames (B, 3, 1024, 1024), row =
                   // b*4096 + py*64 + px, k = c*256 + ky*16 + kx; M % 256 == 0, K == 768; CTA-pair kernel only
  AM_CONV3 = 2,    // neck Conv2d 3x3 pad 1 (image_encoder.py:100-106): A = bf16 (B, 66, 66, 256) zero-ringed activation,
                   // row = b*4096 + y*64 + x, k = (dy*3 + dx)*256 + c; M % 128 == 0, K == 2304; single-CTA kernel only
};

struct GemmEpi {
  const float* bias = nullptr;    // [N] (for UPS1/UPS2 already expanded over the 4 (ky,kx) groups)
  const float* addend = nullptr;  // fp32 [*, ld_add]
  int ld_add = 0;
  int add_mod = 0;                // >0: addend row = orow % add_mod (broadcast tables), else addend row = orow
  void* out = nullptr;            // bf16 or fp32 depending on mode
  void* out2 = nullptr;           // GM_LN: optional bf16 copy
  int ld_out = 0;
  int rowmap = RM_IDENT;
  int a_mode = AM_PLAIN;
  float relw_scale = 1.f;      // RM_RELW_SHIFT: factor applied (in fp32) to the shifted w columns before the fp16 store
  int* f16_overflow = nullptr; // GM_F16: set to 1 (host-mapped flag) when a value had to be saturated to +-65504
  const float* ln_g = nullptr;
  const float* ln_b = nullptr;
  float ln_eps = 0.f;
  const float* hyper = nullptr;   // GM_UPS2: (P, 32) mask-0 hypernetwork vector per prompt
  // GM_F16 only: `batch` independent products sharing W; batch b reads A columns [b*a_batch_off, b*a_batch_off + K)
  // (the A tensor map then spans the full lda-wide rows, so W must be zero in any K padding) and writes output
  // columns [b*out_batch_off, b*out_batch_off + N).
  int batch = 1;
  int a_batch_off = 0;
  int out_batch_off = 0;
  // GM_F32 only, > 1: the K range is cut into `ksplit` pieces computed by different CTAs and combined with fp32 atomicAdd
  // into `out`, which the caller has zeroed (no bias / addend / rowmap / batch); every piece must be non-empty.  For
  // products with a huge K and a small output (weight gradients: K = all rows of the batch).
  int ksplit = 1;
  // ---- LayerNorm folded into the GEMM that consumes it (image_encoder.py:179,191: norm1 -> qkv, norm2 -> lin1):
  //   LN(x) W^T + b = rstd * (x (gamma . W)^T) - rstd * mean * c + d,   c[n] = sum_k gamma[k] W[n,k],  d = W beta + b
  // Producer side (GM_F32, the residual-stream epilogues of patch / proj / lin2): besides the fp32 rows the epilogue
  // writes a bf16 copy (the A operand of the next GEMM) and, per row and 128-column half tile, the partial sums
  // {sum x, sum x^2} into stats_out[(row * stats_slots + slot) * 2].
  void* xb_out = nullptr;          // bf16 (M, ld_out)
  float* stats_out = nullptr;
  int stats_slots = 0;             // = ceil(N / BN) * 2 of the PRODUCING GEMM
  // Consumer side (GM_BIAS_BF16 / GM_BIAS_GELU_BF16): W already holds bf16(gamma . W), `bias` holds d, fold_c holds c;
  // mean / rstd of input row m come from stats_in (the producer's slots) over ln_dim = K columns.
  const float* stats_in = nullptr;
  const float* fold_c = nullptr;
  float fold_eps = 0.f;
  int stats_dim = 0;               // columns the producer's sums run over (= K of the consuming GEMM)
  // Side job for the otherwise idle fourth warp of every CTA of the pair kernel (only there: gemm_tc_is_pair()): write
  // bf16(pad_bias) into the pad_n window-pad rows (positions pad_rows[], relative to a frame's 4900-row window layout) of
  // pad_frames frames of pad_dst (pad_cols bf16 per row).  The lin2 launch of block i does this for the qkv buffer of a
  // windowed block i + 1 (image_encoder.py:179-183: the padded tokens are zero after norm1, so their qkv is the bias).
  void* pad_dst = nullptr;
  const float* pad_bias = nullptr;
  const int* pad_rows = nullptr;
  int pad_n = 0, pad_frames = 0, pad_cols = 0;
};

// A: (M, K) row-major with leading dimension lda (elements); W: (N, K) row-major, ldw.
// tf32 == false: A, W are bf16.  tf32 == true: A, W are fp32, multiplied as TF32.
// Requirements: K % 8 == 0 (tf32) / K % 8 == 0 (bf16) and 16-byte aligned rows; N % 32 == 0.
// Returns cudaSuccess or an error; never falls back to another implementation.
cudaError_t gemm_tc(bool tf32, int mode, const void* A, int lda, const void* W, int ldw, int M, int N, int K,
                    const GemmEpi& epi, cudaStream_t stream, bool debug_simt = false);

// Whether gemm_tc() would run this launch on the CTA-pair kernel (the only one that takes GemmEpi::pad_dst).
bool gemm_tc_is_pair(bool tf32, int mode, int M, int N, const GemmEpi& epi, bool debug_simt = false);

// 2-D tiled tensor map over a row-major (rows, cols) matrix with leading dimension ld (elements).  elem_bytes 2 or 4,
// swizzle_bytes in {32, 64, 128} with box_cols * elem_bytes == swizzle_bytes.  Needs gemm_tc_init.
cudaError_t make_tmap_2d(void* tensor_map_out /*CUtensorMap*/, const void* ptr, int elem_bytes, long rows, long cols,
                         long ld, int box_cols, int box_rows, int swizzle_bytes);
int gemm_num_sms();

// Must be called once per process before gemm_tc (resolves cuTensorMapEncodeTiled, sets smem attributes).
cudaError_t gemm_tc_init(int device);

}  // namespace bsam
