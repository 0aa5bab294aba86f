// Handle of libb200sam: hyper-parameters, packed weights and the pre-allocated workspace.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <map>
#include <utility>
#include <string>
#include <vector>

#include "aux.cuh"
#include "gemm_tc.cuh"
#include "kernels.cuh"

namespace bsam {

struct Tensor {
  float* d = nullptr;               // fp32 master copy on the device
  std::vector<int64_t> shape;
  long numel = 0;
};

struct EncBlock {
  bool global = false;
  const float *ln1g, *ln1b, *ln2g, *ln2b;
  const float *qkv_b, *proj_b, *lin1_b, *lin2_b;
  __nv_bfloat16 *qkv_w, *proj_w, *lin1_w, *lin2_w;
  // LayerNorm folded into qkv / lin1 (launch_fold_ln_weights): bf16(gamma . W), c = its row sums, d = W beta + b
  __nv_bfloat16 *qkv_wf = nullptr, *lin1_wf = nullptr;
  float *qkv_c = nullptr, *qkv_d = nullptr, *lin1_c = nullptr, *lin1_d = nullptr;
  __nv_bfloat16 *relh, *relw;       // padded bf16 rel-pos tables
  __nv_bfloat16* relg = nullptr;    // global blocks: (256, 128) stacked + zero padded tables (B operand of the G GEMM)
};

// transformer.Attention weights (transformer.py:185-218), fp32 device pointers into the masters
struct AttnW {
  const float *qw, *qb, *kw, *kb, *vw, *vb, *ow, *ob;
  int internal;                      // 256 (self) or 128 (cross)
};

struct DecLayer {
  AttnW self_attn, t2i, i2t;
  const float *n1g, *n1b, *n2g, *n2b, *n3g, *n3b, *n4g, *n4b;
  const float *mlp1w, *mlp1b, *mlp2w, *mlp2b;
  // derived (finalize): stacked K|V projection of the image side and its positional table, i2t query table
  float* kv_w = nullptr;     // (256, 256): rows 0..127 = k_proj, 128..255 = v_proj        [t2i]
  float* kv_tab = nullptr;   // (4096, 256): [pe @ Wk^T + bk | bv]
  float* q_tab = nullptr;    // (4096, 128): pe @ Wq^T + bq                                 [i2t]
};

struct DecSide {
  bool has_taxonomy = false;
  const float *iou_token, *mask_tokens;
  DecLayer layer[2];
  AttnW final_attn;
  const float *nfg, *nfb;
  float* final_kv_w = nullptr;
  float* final_kv_tab = nullptr;
  const float *up1_lng, *up1_lnb;
  float *up0_w = nullptr, *up0_b = nullptr;   // (256 = (ky,kx,c64), 256), expanded bias (256)
  float *up3_w = nullptr, *up3_b = nullptr;   // (128 = (ky,kx,c32), 64), expanded bias (128)
  const float* hyp_w[4][3];
  const float* hyp_b[4][3];
  const float* iou_w[3];
  const float* iou_b[3];
  const float* tax_w[3];
  const float* tax_b[3];
};

// decoder workspace of ONE side, sized for max_prompts
struct DecWorkspace {
  float* keys = nullptr;             // (P*4096, 256)
  float* kv = nullptr;               // (P*4096, 256)  [k | v] of the image side
  float* qimg = nullptr;             // (P*4096, 128)
  float* oimg = nullptr;             // (P*4096, 128)
  float* up1 = nullptr;              // (P*16384, 64)
  float* tok_state = nullptr;        // (P, 6, 256) queries
  float* tok_pe = nullptr;           // (P, 6, 256) initial tokens (= query PE)
  float* tok_q = nullptr;            // (P, 6, 128) projected queries for t2i
  float* tok_o = nullptr;            // (P, 6, 128) t2i attention output before out_proj
  float* tok_kv = nullptr;           // (P, 6, 256) [k | v] of the tokens for i2t
  float* hyper = nullptr;            // (P, 32) mask-0 hypernetwork output
};

// encode_body captured for one batch size (api.cu)
struct EncGraph {
  cudaGraphExec_t exec = nullptr;
  int sig = 0;                       // kernel-variant signature the capture was made with
  int calls = 0;
  int64_t n_launches = 0;            // kernels inside the graph (b200sam_launch_count keeps counting them)
};

}  // namespace bsam

struct b200sam {
  int dim = 0, depth = 0, heads = 0, hd = 0;
  std::vector<int> global_idx;
  int device = 0, max_batch = 0, max_prompts = 0;
  std::string err;
  int64_t launches = 0;
  bool debug_simt = false;
  // per-kernel-class CUDA-event profile (b200sam_profile_enable / _read)
  bool prof_on = false;
  int pc = 0;
  cudaStream_t prof_stream = nullptr;
  size_t prof_used = 0;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> prof_ev;
  std::vector<int> prof_cls;
  int debug_stop_block = -1;         // >= 0: encode() stops after that many blocks (bisecting only)
  bool finalized = false;
  bool use_graphs = true;
  std::map<int, bsam::EncGraph> enc_graphs;   // by batch size
  cudaStream_t graph_stream = nullptr;        // where the encoder graphs are captured and replayed
  cudaEvent_t ev_g0 = nullptr, ev_g1 = nullptr;
  bool have_emb = false;
  int emb_batch = 0;

  std::map<std::string, bsam::Tensor> w;
  std::vector<void*> owned;          // workspace + grow-on-demand buffers (freed in destroy)
  // buffers derived from the weights by finalize_weights (bf16 packs, rel-pos / positional tables, permuted conv
  // weights).  They are freed and rebuilt when a tensor of their group was re-loaded, so repeated
  // load_weight + finalize cycles (validation between training epochs) keep the footprint flat.
  std::vector<void*> derived_enc, derived_dec;
  std::vector<void*>* alloc_into = &owned;
  bool dirty_enc = true, dirty_dec = true;

  // encoder
  __nv_bfloat16* patch_w = nullptr;
  const float* patch_b = nullptr;
  const float* pos = nullptr;
  std::vector<bsam::EncBlock> blk;
  __nv_bfloat16 *neck0_w = nullptr, *neck2_w = nullptr;
  const float *neck1g, *neck1b, *neck3g, *neck3b;
  // prompt encoder
  float* pe_tok = nullptr;           // (4096, 256) dense positional encoding, token-major
  const float* no_mask = nullptr;    // (256)
  // decoders
  bsam::DecSide side[2];

  // workspace (sized for max_batch / max_prompts)
  float* X = nullptr;                // (B*4096, D) fp32 residual stream
  __nv_bfloat16* ln_plain = nullptr; // (B*4096, D)
  __nv_bfloat16* qkv = nullptr;      // (B*4900, 3D)
  int* pad_rows = nullptr;           // (n_pad_rows = 804) window-pad positions inside one frame's 4900-row window layout
  int n_pad_rows = 0;
  __nv_bfloat16* attn_o = nullptr;   // (B*4900, D)
  void* gbuf = nullptr;              // (B*4096, heads*256) fp16 rel-pos projections of the global blocks
  // host-mapped flag the G GEMM epilogue raises when a projection left the fp16 range and was saturated: checked at
  // the next API call, which then fails (an out-of-range checkpoint must not produce quietly wrong masks)
  volatile int* ovf_host = nullptr;
  int* ovf_dev = nullptr;
  float* ln_stats = nullptr;         // (B*4096, 2*ceil(D/256), 2) partial row sums of the residual stream (folded LayerNorm)
  int ln_slots = 0;
  __nv_bfloat16* zpad = nullptr;     // (B, 66, 66, 256) neck.1 output with a zero ring (3x3 convolution operand)
  __nv_bfloat16* hid = nullptr;      // (B*4096, 4D) MLP hidden; also im2col scratch
  float* emb_tok = nullptr;          // (B*4096, 256) fp32 image embedding, token-major
  // decoder workspace
  int* prompt_frame = nullptr;       // (P) device
  bsam::DecWorkspace dws[2];         // one per mask decoder (left / right run concurrently)
  cudaStream_t side_stream = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  float* emb_in = nullptr;           // (B*4096, 256) staging when decode() is given an external embedding

  // callers either side of the branch (aux.cu); grown on first use / new input size, never on a repeated call
  int n_sm = 148;
  std::map<std::pair<int, int>, bsam::ResampleTables> resample;   // (H, W) of the raw frame -> Pillow tables
  int* seg_rows = nullptr;           // (seg_cap_rows) flat row index of every [SEG] token
  int* seg_counts = nullptr;         // (1025) prompts per image, then the total
  float* seg_hidden = nullptr;       // (seg_cap_rows, seg_cap_h) ReLU(fc1)
  int seg_cap_rows = 0, seg_cap_h = 0;
};
