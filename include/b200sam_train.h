/* libb200sam - training forward / backward of the two SAM mask decoders.
 *
 * What it replaces: the autograd graph PyTorch records when LISA.model_forward runs mask_decoder_left /
 * mask_decoder_right with requires_grad parameters (2Haff/model/LISA.py:84-90, 284-325), postprocess_masks under
 * autograd (segment_anything/modeling/sam.py:155-189) and the two mask losses (LISA.py:16-59, 346-420).
 * Every node of that graph, forward and backward, is one of the fp32 primitives below; the host side
 * (2handedafforder_b200/train.py) calls them in the order of mask_decoder.py:122-170 and transformer.py:61-240.
 *
 * Conventions: all tensors are device pointers to contiguous fp32 unless strides are given; `stream` is a
 * cudaStream_t; `h` is any handle created on the device that owns the buffers (error text via b200sam_last_error,
 * launches counted by b200sam_launch_count).  Return 0 or a negative b200sam_status.  No function synchronises.
 */
#ifndef B200SAM_TRAIN_H_
#define B200SAM_TRAIN_H_

#include <stdint.h>

#include "b200sam.h"

#ifdef __cplusplus
extern "C" {
#endif

/* C[b1][b2] = alpha * A[b1][b2] (M x K) * B[b1][b2] (K x N) + beta * C[b1][b2] (+ bias) (then ReLU if relu != 0).
 * sa / sb / sc: four element strides each = {outer batch, inner batch, row, column} (rows of B are the K index).
 * bias_mode 0 none, 1 bias[n], 2 bias[m].  Replaces: nn.Linear forward / dX / dW (torch.addmm), the per-head
 * q @ k^T and attn @ v of transformer.py:222-233 and their adjoints, hyper_in @ upscaled_embedding
 * (mask_decoder.py:160-163) and, with a pixel shuffle, ConvTranspose2d(k=2, s=2) (mask_decoder.py:52-61).
 * Long reductions into few output tiles are split over K and combined with fp32 atomics. */
int b200sam_train_gemm(b200sam_t* h, const float* A, const int64_t* sa, const float* B, const int64_t* sb, float* C,
                       const int64_t* sc, const float* bias, int bias_mode, int nb1, int nb2, int M, int N, int K,
                       float alpha, float beta, int relu, void* stream);

/* out (M x N, row pitch ldo) = A (M x K, pitch lda) * W^T (W: N x K, pitch ldw) (+ bias[n]) on the tensor cores
 * (tcgen05, TF32 operands, fp32 accumulation in TMEM; the TMA-fed kernel of the inference decoder).  N % 32 == 0,
 * K % 8 == 0, 16-byte aligned rows.  The image-side Linear layers (rows = P * 4096) and both ConvTranspose2d use it
 * for the forward, the input gradient (W^T supplied transposed) and - with split_k != 0: K cut into one range per
 * CTA, partial products combined with fp32 atomics, no bias - the weight gradient dW = dY^T X on transposed copies. */
int b200sam_train_gemm_tc(b200sam_t* h, const float* A, int64_t lda, const float* W, int64_t ldw, float* out,
                          int64_t ldo, const float* bias, int M, int N, int K, int split_k, void* stream);

/* y (cols x rows) = x (rows x cols)^T. */
int b200sam_train_transpose(b200sam_t* h, const float* x, int64_t rows, int cols, float* y, void* stream);

/* nn.LayerNorm over the last dimension (transformer.py:97-107; LayerNorm2d of common.py:31-43 on channels-last
 * rows).  C % 32 == 0, C <= 1024.  Saves mean / rstd per row for the backward.  dgamma / dbeta are overwritten. */
int b200sam_train_layernorm_fwd(b200sam_t* h, const float* x, const float* gamma, const float* beta, int64_t rows,
                                int C, float eps, float* y, float* mean, float* rstd, void* stream);
int b200sam_train_layernorm_bwd(b200sam_t* h, const float* dy, const float* x, const float* gamma, const float* mean,
                                const float* rstd, int64_t rows, int C, float* dx, float* dgamma, float* dbeta,
                                void* stream);

/* torch.softmax(x, dim=-1) over rows of length L (transformer.py:229, mask_decoder.py:177) and its adjoint
 * dx = y * (dy - sum(dy * y)).  In place allowed. */
int b200sam_train_softmax_fwd(b200sam_t* h, const float* x, int64_t rows, int L, float* y, void* stream);
int b200sam_train_softmax_bwd(b200sam_t* h, const float* dy, const float* y, int64_t rows, int L, float* dx,
                              void* stream);

/* kind 0: ReLU (MLP blocks, transformer.py / mask_decoder.py:200-204); kind 1: exact-erf GELU (nn.GELU,
 * mask_decoder.py:56,60).  Backward: `ref` is the OUTPUT for ReLU and the INPUT for GELU. */
int b200sam_train_act_fwd(b200sam_t* h, int kind, const float* x, int64_t n, float* y, void* stream);
int b200sam_train_act_bwd(b200sam_t* h, int kind, const float* dy, const float* ref, int64_t n, float* dx,
                          void* stream);

/* out[r][c] = a[r][c] + b[r % mod][c]: residual adds (mod = rows) and the positional adds `keys + key_pe`
 * (transformer.py:170,186,203) where key_pe is shared by every prompt (mod = 4096). */
int b200sam_train_add_bcast(b200sam_t* h, const float* a, const float* b, int64_t rows, int C, int64_t mod, float* out,
                            void* stream);

/* out[c] = sum_r x[r][c] (bias gradients). */
int b200sam_train_colsum(b200sam_t* h, const float* x, int64_t rows, int C, float* out, void* stream);

/* (n, hgt, wid, co*4), column = co*4 + ky*2 + kx  ->  (n, 2 hgt, 2 wid, co) (inverse != 0: the other way). */
int b200sam_train_pixel_shuffle(b200sam_t* h, const float* x, int64_t n_img, int hgt, int wid, int co, int inverse,
                                float* y, void* stream);

/* src[p] = image_embeddings[frame_of_prompt[p]] + dense, NCHW (n_frames,256,64,64) -> token-major (P,4096,256) fp32
 * (mask_decoder.py:141-143 + the flatten/permute of transformer.py:83-85).  emb_dtype: B200SAM_F32/BF16/F16. */
int b200sam_train_src_tokens(b200sam_t* h, const void* emb_nchw, int emb_dtype, int n_frames,
                             const int32_t* frame_of_prompt, const float* dense_vec, int P, float* out, void* stream);

/* F.interpolate(mode="bilinear", align_corners=False) of n planes and its adjoint (sam.py:174-188 under autograd).
 * The input plane is a pitched view: in_h x in_w live pixels, row pitch `pitch`, plane stride `plane` (elements), so
 * the crop masks[..., :h, :w] costs nothing.  resize_bwd overwrites the whole dx (n * plane elements). */
int b200sam_train_resize_fwd(b200sam_t* h, const float* x, int64_t n, int in_h, int in_w, int64_t pitch, int64_t plane,
                             int out_h, int out_w, float* y, void* stream);
int b200sam_train_resize_bwd(b200sam_t* h, const float* dy, int64_t n, int in_h, int in_w, int64_t pitch,
                             int64_t plane, int out_h, int out_w, float* dx, void* stream);

/* sigmoid_ce_loss + dice_loss (LISA.py:16-59) per mask on z = weight[p] * logits[p] (weight = the taxonomy gate of
 * LISA.py:357-358, NULL = 1).  Forward: sums[p] = {sum bce_with_logits, sum sigmoid, sum sigmoid * gt, sum gt} in
 * double.  Backward: dlogits = d(g_bce[p] * bce_mean[p] + g_dice[p] * dice[p]) / dlogits with dice's scale / eps. */
int b200sam_train_mask_loss_fwd(b200sam_t* h, const float* logits, const float* gt, const float* weight, int P,
                                int64_t hw, double* sums, void* stream);
int b200sam_train_mask_loss_bwd(b200sam_t* h, const float* logits, const float* gt, const float* weight,
                                const double* sums, const float* g_bce, const float* g_dice, int P, int64_t hw,
                                float scale, float eps, float* dlogits, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200SAM_TRAIN_H_ */
