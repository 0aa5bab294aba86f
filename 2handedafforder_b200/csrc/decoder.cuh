// Token-side kernels of the two-way mask decoder (decoder.cu).
#pragma once
#include <cuda_runtime.h>

namespace bsam {

struct AttnPtrs {
  const float *qw, *qb, *kw, *kb, *vw, *vb, *ow, *ob;
};

struct TokStageA {
  int layer;                      // 0: build tokens + replace queries; 1: residual form
  const float* iou_token;         // (1,256)
  const float* mask_tokens;       // (4,256)
  const float* seg;               // (P,256) projected [SEG] embeddings
  AttnPtrs sa;                    // self attention (internal 256)
  const float *n1g, *n1b;
  const float *t2i_qw, *t2i_qb;   // cross_attn_token_to_image.q_proj (128,256)
  float* tok_state;               // (P,6,256) in/out
  float* tok_pe;                  // (P,6,256) out (layer 0) / in
  float* tok_q;                   // (P,6,128) out
};

struct TokStageB {
  const float *t2i_ow, *t2i_ob;   // cross_attn_token_to_image.out_proj (256,128)
  const float *n2g, *n2b, *n3g, *n3b;
  const float *mlp1w, *mlp1b, *mlp2w, *mlp2b;
  const float *i2t_kw, *i2t_kb, *i2t_vw, *i2t_vb;   // cross_attn_image_to_token k/v proj (128,256)
  const float* tok_o;             // (P,6,128)
  const float* tok_pe;
  float* tok_state;
  float* tok_kv;                  // (P,6,256) out: [k | v]
};

struct TokHeads {
  const float *ow, *ob, *nfg, *nfb;   // final_attn_token_to_image.out_proj, norm_final_attn
  const float* hyp_w[3];
  const float* hyp_b[3];
  const float* iou_w[3];
  const float* iou_b[3];
  const float* tax_w[3];              // nullptr on the right decoder
  const float* tax_b[3];
  const float* tok_state;
  const float* tok_o;
  float* hyper;                       // (P,32)
  float* iou_out;                     // (P,4) or nullptr
  float* tax_out;                     // (P,4) or nullptr
};

cudaError_t decoder_init();
cudaError_t launch_dec_stage_a(const TokStageA& a, int P, cudaStream_t s);
cudaError_t launch_dec_stage_q(const float* tok_state, const float* tok_pe, const float* qw, const float* qb,
                               float* tok_q, int P, cudaStream_t s);
cudaError_t launch_dec_t2i(const float* tok_q, const float* kv, float* tok_o, int P, cudaStream_t s);
cudaError_t launch_dec_stage_b(const TokStageB& a, int P, cudaStream_t s);
cudaError_t launch_dec_i2t(const float* qimg, const float* tok_kv, float* oimg, int P, cudaStream_t s);
cudaError_t launch_dec_heads(const TokHeads& a, int P, cudaStream_t s);

}  // namespace bsam
