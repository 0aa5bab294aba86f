/* libb200sam — C ABI of the B200-native SAM segmentation branch of 2HandedAfforder.
 *
 * Drop-in boundary for the reference's `visual_model` submodule (2Haff/model/segment_anything/modeling/sam.py:18-55)
 * as driven by 2Haff/model/LISA.py:157-168 (get_visual_embs), :284-325 (model_forward) and :487-532 (evaluate).
 * Plain pointers and sizes only; no torch types.  All device work is enqueued on the caller's stream; no hidden
 * synchronisation on the hot path.  Every function returns 0 on success or a negative b200sam_status; the message is
 * available from b200sam_last_error().  There is no CPU fallback: a missing GPU / wrong architecture is an error.
 *
 * Pointers are DEVICE pointers unless the parameter name starts with `host_`.
 */
#ifndef B200SAM_H_
#define B200SAM_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200sam b200sam_t;

enum b200sam_status {
  B200SAM_OK = 0,
  B200SAM_ERR_INVALID = -1,      /* bad argument (shape, dtype, batch > max, unknown weight name) */
  B200SAM_ERR_CUDA = -2,         /* a CUDA call failed; see last_error */
  B200SAM_ERR_STATE = -3,        /* weights missing / not finalised */
  B200SAM_ERR_UNSUPPORTED = -4   /* device is not sm_100 */
};

enum b200sam_dtype { B200SAM_F32 = 0, B200SAM_BF16 = 1, B200SAM_F16 = 2 };

/* build_sam.py:15-48 hyper-parameter table. */
enum b200sam_arch { B200SAM_VIT_B = 0, B200SAM_VIT_L = 1, B200SAM_VIT_H = 2 };

/* Replaces build_sam_vit_{b,l,h}() (build_sam.py:15-48) + Sam.__init__ (sam.py:24-55).
 * Workspace for `max_batch` frames and `max_prompts` [SEG] prompts (summed over the batch) is allocated here, so
 * the hot path never allocates. */
int b200sam_create(b200sam_t** out, int arch, int device, int max_batch, int max_prompts);

/* Same, with explicit ViT hyper-parameters (_build_sam, build_sam.py:59-121); used for small test shapes.
 * dim % 32 == 0, dim / heads in {64, 80}. */
int b200sam_create_custom(b200sam_t** out, int dim, int depth, int heads, const int* host_global_idx, int n_global,
                          int device, int max_batch, int max_prompts);

void b200sam_destroy(b200sam_t* h);
const char* b200sam_last_error(const b200sam_t* h);

/* Replaces Sam.load_state_dict for one tensor.  `name` is the reference state_dict key relative to visual_model
 * (e.g. "image_encoder.blocks.0.attn.qkv.weight"); `ptr` may be a host or a device pointer.
 * The tensor is copied; the caller keeps ownership. */
int b200sam_load_weight(b200sam_t* h, const char* name, const void* ptr, int dtype, const int64_t* host_shape,
                        int ndim);
/* Checks that every tensor of the manifest was loaded, repacks into kernel layouts (bf16 GEMM operands, permuted
 * conv weights, positional tables).  Must be called after the last load_weight and before any compute call. */
int b200sam_finalize_weights(b200sam_t* h);

/* ImageEncoderViT.forward (image_encoder.py:110-125) for B frames at once (LISA.get_visual_embs loops B=1,
 * LISA.py:157-168).  frames: (B,3,1024,1024) NCHW, normalised and zero padded, dtype f32/bf16.
 * emb_out: (B,256,64,64) NCHW in `emb_dtype` (f32 or bf16); may be NULL, the fp32 embedding also stays in the
 * handle for b200sam_decode(..., emb = NULL). */
int b200sam_encode(b200sam_t* h, const void* frames, int frames_dtype, int B, void* emb_out, int emb_dtype,
                   void* stream);

/* PromptEncoder text path + both MaskDecoders (prompt_encoder.py:140-186, mask_decoder.py:79-178) for all prompts
 * of a batch.  emb: (B,256,64,64) NCHW in emb_dtype, or NULL to use the embedding left by the last encode.
 * seg_embeds: (P,256) f32, P = host_prompt_offsets[B]; prompts of frame i are rows [off[i], off[i+1]).
 * Outputs (all f32, any may be NULL): lowres_left/right (P,256,256) = mask slot 0 (multimask_output=False,
 * mask_decoder.py:110-115); iou_left/right (P,4) full IoU head; taxonomy (P,4) softmax (left decoder only).
 * A decoder whose outputs are all NULL is not run (the facade's mask_decoder_left / _right call one side each). */
int b200sam_decode(b200sam_t* h, const void* emb, int emb_dtype, int B, const float* seg_embeds,
                   const int* host_prompt_offsets, float* lowres_left, float* lowres_right, float* iou_left,
                   float* iou_right, float* taxonomy, void* stream);

/* Sam.postprocess_masks (sam.py:155-189) + the callers' threshold (chat.py:231, train_ds.py:662).
 * lowres: (n,256,256) f32.  Every mask uses the same input_size (h,w) and original_size (H,W).
 * logits_out: (n,H,W) f32 or NULL; mask_out: (n,H,W) u8 {0,1} = logit > threshold, or NULL. */
int b200sam_postprocess(b200sam_t* h, const float* lowres, int n_masks, int in_h, int in_w, int out_h, int out_w,
                        float* logits_out, uint8_t* mask_out, float threshold, void* stream);

/* Same resampling + threshold, written in the 1-bit wire format (1.56 MB instead of 12.4 MB per
 * 1080p frame through NVLink and PCIe): bits_out (n, H, ceil(W/8)) u8, bit 7 of byte X/8 = pixel 8*(X/8) - the layout
 * of numpy.packbits(mask, axis=-1); the pad bits of a row's last byte are 0.  bits_out may be any device address, e.g.
 * a slice of the send buffer of the gather (clip.segment_clip writes there directly, no staging copy). */
int b200sam_postprocess_bits(b200sam_t* h, const float* lowres, int n_masks, int in_h, int in_w, int out_h,
                             int out_w, uint8_t* bits_out, float threshold, void* stream);

/* Both decoders' masks of P prompts in ONE pass over the low-res logits: postprocess + threshold of the left and the
 * right mask -> bits_left / bits_right (packed as above, either may be NULL) and, when gt != NULL, the evaluation of
 * train_ds.py:662-709,761-796 fused behind the threshold - the taxonomy-gated prediction (argmax 0: left, 1: right,
 * 2/3: left|right) against gt[gt_index[p]] ((G,H,W) u8, non-zero = set; gt_index NULL: gt[p]) reduced to
 * counts_out (P,3) int64 {intersection, union, predicted area}, plus the packed gated prediction pred_bits_out
 * (nullable).  No full-resolution u8 plane is written or re-read. */
int b200sam_postprocess_pair(b200sam_t* h, const float* lowres_left, const float* lowres_right, int P, int in_h,
                             int in_w, int out_h, int out_w, uint8_t* bits_left, uint8_t* bits_right, float threshold,
                             const float* taxonomy, const uint8_t* gt, const int* gt_index, int64_t* counts_out,
                             uint8_t* pred_bits_out, void* stream);

/* out[i] = 64-bit position-sensitive checksum of the byte string data[i*stride, i*stride + bytes): the sum over its
 * 8-byte little-endian words w_k (tail zero padded) of splitmix64_final(w_k ^ (k+1)*0x9E3779B97F4A7C15) mod 2^64.
 * Shipped beside the packed masks by clip.segment_clip so that the gathering rank can verify what arrived. */
int b200sam_checksum64(b200sam_t* h, const void* data, int n_items, int64_t bytes, int64_t stride, uint64_t* out,
                       void* stream);

/* ---- the callers either side of the branch ---- */

/* ResizeLongestSide(1024).apply_image (segment_anything/utils/transforms.py:27-34: torchvision resize of a PIL image
 * = Pillow's antialiased 8-bit bilinear resample) + preprocess (inference.py:91-105 = chat.py:49-63: (x - mean) / std,
 * zero pad bottom/right to 1024^2) + the callers' cast to the model dtype (inference.py:251-256), bit-exact.
 * frames_rgb: (B,H,W,3) uint8 RGB; out: (B,3,1024,1024) in out_dtype (f32 / bf16 / f16).
 * host_resize_hw (host, 2 ints, may be NULL) receives get_preprocess_shape(H, W) (transforms.py:103-113), the
 * `resize_list` entry the callers pass on to postprocess_masks. */
int b200sam_preprocess(b200sam_t* h, const uint8_t* frames_rgb, int B, int H, int W, void* out, int out_dtype,
                       int* host_resize_hw, void* stream);

/* text_hidden_fcs[0] (LISA.py:95-101: Linear(h,h) - ReLU - Linear(h,256)) evaluated on the [SEG] rows only, with the
 * row selection of LISAForCausalLM.evaluate (LISA.py:457-485): mask = output_ids[:,1:] == seg_token_idx shifted by the
 * 255 extra image-token positions.  hidden: (b,L,hdim) last hidden states in hidden_dtype, L >= 255 + T - 1;
 * output_ids: (b,T) int64; w1 (hdim,hdim), b1 (hdim), w2 (256,hdim), b2 (256) all in w_dtype (f32 or bf16).
 * emb_out: (max_rows,256) f32, rows [0,total) written in row-major [SEG] order; counts_out: device (b+1) ints =
 * prompts per image, then the total (if the total exceeds max_rows only max_rows rows are written: the caller checks). */
int b200sam_seg_project(b200sam_t* h, const void* hidden, int hidden_dtype, int b, int L, int hdim,
                        const int64_t* output_ids, int T, int64_t seg_token_idx, const void* w1, const void* b1,
                        const void* w2, const void* b2, int w_dtype, float* emb_out, int max_rows, int* counts_out,
                        void* stream);

/* b200sam_seg_project without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same arguments, same results as b200sam_seg_project; its scratch ([SEG] row indices, counts, the
 * ReLU(Linear1) rows) is a file-scope device array of fixed size instead of handle-owned buffers, so
 *   - sizes are limited to max_rows <= 64, hdim <= 5120, b <= 1024 (larger: B200SAM_ERR_INVALID with the limits in
 *     b200sam_last_error; use b200sam_seg_project);
 *   - every GPU that loads the library holds the array (~1.26 MB) until exit;
 *   - all calls on a device share it: calls that overlap on the GPU (different streams / host threads) race.
 * Calls no cudaMalloc / cudaFree / allocating routine; the handle `h` is created beforehand at set-up. */
int b200sam_seg_project_noalloc(b200sam_t* h, const void* hidden, int hidden_dtype, int b, int L, int hdim,
                                const int64_t* output_ids, int T, int64_t seg_token_idx, const void* w1,
                                const void* b1, const void* w2, const void* b2, int w_dtype, float* emb_out,
                                int max_rows, int* counts_out, void* stream);

/* Evaluation of train_ds.py::validate on the device (train_ds.py:662-709,761-796).  masks_left / masks_right:
 * (P,H,W) u8 thresholded masks (b200sam_postprocess mask_out); taxonomy (P,4) f32; gt: (G,H,W) u8 ground-truth masks,
 * prompt p is compared with gt[gt_index[p]] (gt_index == NULL: gt[p]).  The predicted mask is left (taxonomy argmax
 * 0), right (1) or left|right (2,3).  counts_out: (P,3) int64 = {intersection, union, predicted area} in pixels, from
 * which IoU = I / U and the reference's IoCM = I / (255 * area) follow; pred_out (nullable): (P,H,W) u8 gated mask
 * with values {0,255}. */
int b200sam_mask_metrics(b200sam_t* h, const uint8_t* masks_left, const uint8_t* masks_right, const float* taxonomy,
                         const uint8_t* gt, const int* gt_index, int P, int H, int W, int64_t* counts_out,
                         uint8_t* pred_out, void* stream);

/* PromptEncoder.get_dense_pe (prompt_encoder.py:67-76): writes (1,256,64,64) f32. */
int b200sam_dense_pe(b200sam_t* h, float* out, void* stream);

/* Number of kernels this library has launched on behalf of handle h since creation (bench.py's gpu_launches). */
int64_t b200sam_launch_count(const b200sam_t* h);

/* Kernel classes of the per-launch profile. */
enum b200sam_profile_class {
  B200SAM_PC_ELEMENTWISE = 0,     /* im2col of the patches, layout transposes */
  B200SAM_PC_GEMM_PATCH = 1,
  B200SAM_PC_LAYERNORM = 2,
  B200SAM_PC_GEMM_QKV = 3,
  B200SAM_PC_ATTN_WINDOW = 4,
  B200SAM_PC_ATTN_GLOBAL = 5,
  B200SAM_PC_GEMM_PROJ = 6,
  B200SAM_PC_GEMM_LIN1 = 7,
  B200SAM_PC_GEMM_LIN2 = 8,
  B200SAM_PC_NECK = 9,
  B200SAM_PC_DEC_ELEMENTWISE = 10,
  B200SAM_PC_DEC_TOKEN = 11,
  B200SAM_PC_DEC_GEMM = 12,
  B200SAM_PC_DEC_UPSCALE = 13,
  B200SAM_PC_POSTPROCESS = 14,
  B200SAM_PC_PREPROCESS = 15,
  B200SAM_PC_SEG_PROJECT = 16,
  B200SAM_PC_METRICS = 17,
  B200SAM_PC_COUNT = 18
};
/* Measurement aid for bench.py: while enabled, every kernel launch of this handle is bracketed by a pair of CUDA
 * events on the launching stream.  b200sam_profile_read synchronises on them, writes the summed milliseconds and
 * the launch count of every class into the two HOST arrays (length n_classes >= B200SAM_PC_COUNT) and resets. */
int b200sam_profile_enable(b200sam_t* h, int on);
int b200sam_profile_read(b200sam_t* h, float* host_ms, int64_t* host_count, int n_classes);

/* ---- operator-level entry points (used by the per-op parity tests; same kernels the path uses) ---- */

/* C = A(M,K) * W(N,K)^T with fused epilogue.  mode: 0 bias->bf16, 1 bias+GELU->bf16, 2 (+bias)(+addend)->f32,
 * 3 (+bias)(+addend)->LayerNorm(N=256)->f32.  tf32: 0 = bf16 operands, 1 = fp32 operands multiplied as TF32. */
int b200sam_op_gemm(int device, int tf32, int mode, const void* A, int lda, const void* W, int ldw, int M, int N,
                    int K, const float* bias, const float* addend, int ld_add, int add_mod, const float* ln_g,
                    const float* ln_b, float ln_eps, void* out, int ld_out, int rowmap, int debug_simt, void* stream);

/* The residual-stream GEMM of the encoder with the producer half of the folded LayerNorm (proj / lin2 / patch embed;
 * image_encoder.py:187-191 with norm1 / norm2 of the NEXT consumer folded), bf16 operands:
 * out(M,N) f32 = A W^T + bias + addend[row % add_mod or row] (addend may alias out), xb_out(M,N) = bf16(out) with the
 * same leading dimension, stats_out(M, 2*ceil(N/256)) float2 = per row and 128-column slice {sum, sum of squares} of
 * out (two slots per 256-wide tile; a last tile of 128 columns splits them 64 + 64; consumers add all slots of a row).
 * addend is required; add_mod must be 0 or a multiple of 32. */
int b200sam_op_gemm_fold_producer(int device, const void* A, int lda, const void* W, int ldw, int M, int N, int K,
                                  const float* bias, const float* addend, int ld_add, int add_mod, float* out,
                                  int ld_out, void* xb_out, float* stats_out, void* stream);

/* The consuming half (qkv / lin1; image_encoder.py:179,191 + common.py:25-26): out(M,N) bf16 =
 * [GELU](LayerNorm_{gamma,beta,eps}(x) W^T + bias), evaluated as the path evaluates it.  W (N,K) f32 is folded by the
 * weight-finalisation kernel into caller scratch - wf_scratch (N,K) bf16 = bf16(gamma . W), cd_scratch (2N) f32 =
 * {c[n] = sum_k wf[n,k], d = W beta + bias} - then out = rstd * (xb wf^T) - rstd * mean * c + d with xb (M,K) the
 * bf16 copy of x and mean / rstd from stats (M, 2*ceil(K/256)) float2, the producer's per-slice {sum, sum of squares}. */
int b200sam_op_gemm_fold_consumer(int device, int gelu, const void* xb, int lda, const float* W, const float* gamma,
                                  const float* beta, const float* bias, float eps, const float* stats, int M, int N,
                                  int K, void* wf_scratch, float* cd_scratch, void* out, int ld_out, void* stream);

/* LayerNorm over the last dim of x (rows, D) f32 -> bf16; windowed != 0 scatters rows into 14x14-window order
 * (rows must then be B*4096; out has B*4900 rows whose pad rows are left untouched). */
int b200sam_op_layernorm(const float* x, int rows, int D, const float* g, const float* b, float eps, void* out_bf16,
                         int windowed, void* stream);

/* Attention with decomposed rel-pos (image_encoder.py:235-260,354-392) on a packed qkv buffer
 * (rows, 3*heads*hd) bf16; S = 14 (rows = nbatch*196) or 64 (rows = nbatch*4096); rel_h / rel_w: (2S-1, hd) f32.
 * out: (rows, heads*hd) bf16. */
int b200sam_op_attention(const void* qkv_bf16, int nbatch, int S, int heads, int hd, const float* rel_h,
                         const float* rel_w, void* out_bf16, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200SAM_H_ */
