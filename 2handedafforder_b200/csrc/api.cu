// libb200sam C ABI (include/b200sam.h): handle, weight ingestion/repacking, and the launch sequences of
// ImageEncoderViT.forward (image_encoder.py:110-125), PromptEncoder text path + both MaskDecoders
// (prompt_encoder.py:140-186, mask_decoder.py:79-178) and Sam.postprocess_masks (sam.py:155-189).
//
// No work is done on the host per element and nothing is allocated on the hot path: the workspace is sized at
// create() and every compute entry point only enqueues kernels on the caller's stream.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <string>
#include <vector>

#include <nvtx3/nvToolsExt.h>

#include "../../include/b200sam.h"
#include "../../include/b200sam_debug.h"
#include "common.cuh"
#include "decoder.cuh"
#include "gemm_tc.cuh"
#include "kernels.cuh"
#include "model.cuh"

using namespace bsam;

namespace {

int fail(b200sam_t* h, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  if (h) h->err = buf;
  return code;
}

// NVTX range per stage of the path ("Tracing / profiling"): visible to nsys timelines and usable as
// an ncu filter (`ncu --nvtx --nvtx-include "b200sam_decode/" ...`).  Header-only NVTX3: a no-op costing a few
// nanoseconds when no tool is attached.
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
  NvtxRange(const NvtxRange&) = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
};

#define CU(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess)                                                                         \
      return fail(h, B200SAM_ERR_CUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
  } while (0)

// kernel launch through one of the launch_* helpers: counts it and, in profiling mode, brackets it with CUDA
// events on the launching stream tagged with the current kernel class (h->pc)
#define KL(call)                                                                                   \
  do {                                                                                             \
    prof_begin(h);                                                                                 \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess)                                                                         \
      return fail(h, B200SAM_ERR_CUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
    prof_end(h);                                                                                   \
    ++h->launches;                                                                                 \
  } while (0)

void prof_begin(b200sam_t* h) {
  if (!h->prof_on) return;
  if (h->prof_used == h->prof_ev.size()) {
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    h->prof_ev.push_back({a, b});
    h->prof_cls.push_back(0);
  }
  cudaEventRecord(h->prof_ev[h->prof_used].first, h->prof_stream);
}
void prof_end(b200sam_t* h) {
  if (!h->prof_on) return;
  cudaEventRecord(h->prof_ev[h->prof_used].second, h->prof_stream);
  h->prof_cls[h->prof_used] = h->pc;
  ++h->prof_used;
}

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    cudaGetDevice(&prev);
    if (prev != dev) cudaSetDevice(dev);
    else prev = -1;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

template <typename T>
cudaError_t dalloc(b200sam_t* h, T** p, size_t count, bool zero = false) {
  void* q = nullptr;
  cudaError_t e = cudaMalloc(&q, count * sizeof(T) + 256);
  if (e != cudaSuccess) return e;
  if (zero) {
    e = cudaMemset(q, 0, count * sizeof(T));
    if (e != cudaSuccess) return e;
  }
  h->alloc_into->push_back(q);
  *p = reinterpret_cast<T*>(q);
  return cudaSuccess;
}

void free_list(std::vector<void*>& v) {
  for (void* p : v) cudaFree(p);
  v.clear();
}

// routes dalloc() into one of the handle's derived-buffer lists for the lifetime of the scope
struct AllocScope {
  b200sam_t* h;
  AllocScope(b200sam_t* h_, std::vector<void*>* list) : h(h_) { h->alloc_into = list; }
  ~AllocScope() { h->alloc_into = &h->owned; }
};

struct ManifestEntry {
  std::string name;
  std::vector<int64_t> shape;
  bool required;
};

void add(std::vector<ManifestEntry>& m, const std::string& n, std::vector<int64_t> s, bool req = true) {
  m.push_back({n, std::move(s), req});
}
void add_lin(std::vector<ManifestEntry>& m, const std::string& p, int64_t out, int64_t in) {
  add(m, p + ".weight", {out, in});
  add(m, p + ".bias", {out});
}
void add_ln(std::vector<ManifestEntry>& m, const std::string& p, int64_t n, bool req = true) {
  add(m, p + ".weight", {n}, req);
  add(m, p + ".bias", {n}, req);
}
void add_attn(std::vector<ManifestEntry>& m, const std::string& p, int64_t internal) {
  add_lin(m, p + ".q_proj", internal, 256);
  add_lin(m, p + ".k_proj", internal, 256);
  add_lin(m, p + ".v_proj", internal, 256);
  add_lin(m, p + ".out_proj", 256, internal);
}

// every tensor of visual_model.state_dict().  `required == false` marks tensors that must be
// accepted by load_weight but are not on the text-prompt path (point / box / mask prompt embeddings).
std::vector<ManifestEntry> build_manifest(const b200sam_t* h) {
  std::vector<ManifestEntry> m;
  const int64_t D = h->dim, hd = h->hd;
  const std::string e = "image_encoder.";
  add(m, e + "pos_embed", {1, kGrid, kGrid, D});
  add(m, e + "patch_embed.proj.weight", {D, 3, kPatch, kPatch});
  add(m, e + "patch_embed.proj.bias", {D});
  for (int i = 0; i < h->depth; ++i) {
    const std::string b = e + "blocks." + std::to_string(i) + ".";
    bool glob = false;
    for (int g : h->global_idx) glob |= (g == i);
    const int64_t S = glob ? kGrid : kWin;
    add_ln(m, b + "norm1", D);
    add(m, b + "attn.rel_pos_h", {2 * S - 1, hd});
    add(m, b + "attn.rel_pos_w", {2 * S - 1, hd});
    add_lin(m, b + "attn.qkv", 3 * D, D);
    add_lin(m, b + "attn.proj", D, D);
    add_ln(m, b + "norm2", D);
    add_lin(m, b + "mlp.lin1", 4 * D, D);
    add_lin(m, b + "mlp.lin2", D, 4 * D);
  }
  add(m, e + "neck.0.weight", {256, D, 1, 1});
  add_ln(m, e + "neck.1", 256);
  add(m, e + "neck.2.weight", {256, 256, 3, 3});
  add_ln(m, e + "neck.3", 256);
  const std::string p = "prompt_encoder.";
  add(m, p + "pe_layer.positional_encoding_gaussian_matrix", {2, 128});
  add(m, p + "no_mask_embed.weight", {1, 256});
  for (int k = 0; k < 4; ++k) add(m, p + "point_embeddings." + std::to_string(k) + ".weight", {1, 256}, false);
  add(m, p + "not_a_point_embed.weight", {1, 256}, false);
  add(m, p + "mask_downscaling.0.weight", {4, 1, 2, 2}, false);
  add(m, p + "mask_downscaling.0.bias", {4}, false);
  add_ln(m, p + "mask_downscaling.1", 4, false);
  add(m, p + "mask_downscaling.3.weight", {16, 4, 2, 2}, false);
  add(m, p + "mask_downscaling.3.bias", {16}, false);
  add_ln(m, p + "mask_downscaling.4", 16, false);
  add(m, p + "mask_downscaling.6.weight", {256, 16, 1, 1}, false);
  add(m, p + "mask_downscaling.6.bias", {256}, false);
  for (int s = 0; s < 2; ++s) {
    const std::string d = s == 0 ? "mask_decoder_left." : "mask_decoder_right.";
    add(m, d + "iou_token.weight", {1, 256});
    add(m, d + "mask_tokens.weight", {4, 256});
    for (int l = 0; l < 2; ++l) {
      const std::string t = d + "transformer.layers." + std::to_string(l) + ".";
      add_attn(m, t + "self_attn", 256);
      add_ln(m, t + "norm1", 256);
      add_attn(m, t + "cross_attn_token_to_image", 128);
      add_ln(m, t + "norm2", 256);
      add_lin(m, t + "mlp.lin1", 2048, 256);
      add_lin(m, t + "mlp.lin2", 256, 2048);
      add_ln(m, t + "norm3", 256);
      add_ln(m, t + "norm4", 256);
      add_attn(m, t + "cross_attn_image_to_token", 128);
    }
    add_attn(m, d + "transformer.final_attn_token_to_image", 128);
    add_ln(m, d + "transformer.norm_final_attn", 256);
    add(m, d + "output_upscaling.0.weight", {256, 64, 2, 2});
    add(m, d + "output_upscaling.0.bias", {64});
    add_ln(m, d + "output_upscaling.1", 64);
    add(m, d + "output_upscaling.3.weight", {64, 32, 2, 2});
    add(m, d + "output_upscaling.3.bias", {32});
    for (int k = 0; k < 4; ++k) {
      const std::string q = d + "output_hypernetworks_mlps." + std::to_string(k) + ".layers.";
      add_lin(m, q + "0", 256, 256);
      add_lin(m, q + "1", 256, 256);
      add_lin(m, q + "2", 32, 256);
    }
    const std::string q = d + "iou_prediction_head.layers.";
    add_lin(m, q + "0", 256, 256);
    add_lin(m, q + "1", 256, 256);
    add_lin(m, q + "2", 4, 256);
    if (s == 0) {  // build_sam.py:103,116: taxonomy head only on the left decoder
      const std::string x = d + "taxonomy_embed.layers.";
      add_lin(m, x + "0", 1024, 1024);
      add_lin(m, x + "1", 1024, 1024);
      add_lin(m, x + "2", 4, 1024);
    }
  }
  return m;
}

const float* W(b200sam_t* h, const std::string& name) { return h->w.at(name).d; }

AttnW attn_w(b200sam_t* h, const std::string& p, int internal) {
  AttnW a;
  a.qw = W(h, p + ".q_proj.weight");
  a.qb = W(h, p + ".q_proj.bias");
  a.kw = W(h, p + ".k_proj.weight");
  a.kb = W(h, p + ".k_proj.bias");
  a.vw = W(h, p + ".v_proj.weight");
  a.vb = W(h, p + ".v_proj.bias");
  a.ow = W(h, p + ".out_proj.weight");
  a.ob = W(h, p + ".out_proj.bias");
  a.internal = internal;
  return a;
}

AttnPtrs attn_ptrs(const AttnW& a) { return AttnPtrs{a.qw, a.qb, a.kw, a.kb, a.vw, a.vb, a.ow, a.ob}; }

int make_bf16(b200sam_t* h, const float* src, long n, __nv_bfloat16** out) {
  CU(dalloc(h, out, (size_t)n));
  KL(launch_cast_f32_bf16(src, *out, n, 0));
  return 0;
}

// Stacked [k_proj ; v_proj] weight of a token->image attention and its positional table:
//   k = (keys + pe) Wk^T + bk = keys Wk^T + (pe Wk^T + bk),  v = keys Wv^T + bv      (transformer.py:163-165,220-225)
// so one GEMM over the image side yields [k | v] with the constant part added from a (4096, 256) table.
int make_kv_tables(b200sam_t* h, const AttnW& a, float** kv_w, float** kv_tab) {
  CU(dalloc(h, kv_w, 256 * 256));
  CU(cudaMemcpy(*kv_w, a.kw, 128 * 256 * 4, cudaMemcpyDeviceToDevice));
  CU(cudaMemcpy(*kv_w + 128 * 256, a.vw, 128 * 256 * 4, cudaMemcpyDeviceToDevice));
  CU(dalloc(h, kv_tab, (size_t)kTokens * 256));
  // exact fp32 evaluation (SIMT) of pe @ Wk^T + bk: one-time cost at finalisation
  GemmEpi e;
  e.bias = a.kb;
  e.out = *kv_tab;
  e.ld_out = 256;
  KL(gemm_tc(true, GM_F32, h->pe_tok, 256, a.kw, 256, kTokens, 128, 256, e, 0, /*debug_simt=*/true));
  std::vector<float> bv(128);
  CU(cudaMemcpy(bv.data(), a.vb, 128 * 4, cudaMemcpyDeviceToHost));
  std::vector<float> rows((size_t)kTokens * 128);
  for (int t = 0; t < kTokens; ++t) memcpy(&rows[(size_t)t * 128], bv.data(), 128 * 4);
  CU(cudaMemcpy2D(*kv_tab + 128, 256 * 4, rows.data(), 128 * 4, 128 * 4, kTokens, cudaMemcpyHostToDevice));
  return 0;
}

int finalize_side(b200sam_t* h, int s) {
  DecSide& d = h->side[s];
  const std::string p = s == 0 ? "mask_decoder_left." : "mask_decoder_right.";
  d.has_taxonomy = (s == 0);
  d.iou_token = W(h, p + "iou_token.weight");
  d.mask_tokens = W(h, p + "mask_tokens.weight");
  for (int l = 0; l < 2; ++l) {
    DecLayer& L = d.layer[l];
    const std::string t = p + "transformer.layers." + std::to_string(l) + ".";
    L.self_attn = attn_w(h, t + "self_attn", 256);
    L.t2i = attn_w(h, t + "cross_attn_token_to_image", 128);
    L.i2t = attn_w(h, t + "cross_attn_image_to_token", 128);
    L.n1g = W(h, t + "norm1.weight"); L.n1b = W(h, t + "norm1.bias");
    L.n2g = W(h, t + "norm2.weight"); L.n2b = W(h, t + "norm2.bias");
    L.n3g = W(h, t + "norm3.weight"); L.n3b = W(h, t + "norm3.bias");
    L.n4g = W(h, t + "norm4.weight"); L.n4b = W(h, t + "norm4.bias");
    L.mlp1w = W(h, t + "mlp.lin1.weight"); L.mlp1b = W(h, t + "mlp.lin1.bias");
    L.mlp2w = W(h, t + "mlp.lin2.weight"); L.mlp2b = W(h, t + "mlp.lin2.bias");
    if (int r = make_kv_tables(h, L.t2i, &L.kv_w, &L.kv_tab)) return r;
    // image->token query table: q = (keys + pe) Wq^T + bq                                  (transformer.py:174-176)
    CU(dalloc(h, &L.q_tab, (size_t)kTokens * 128));
    GemmEpi e;
    e.bias = L.i2t.qb;
    e.out = L.q_tab;
    e.ld_out = 128;
    KL(gemm_tc(true, GM_F32, h->pe_tok, 256, L.i2t.qw, 256, kTokens, 128, 256, e, 0, true));
  }
  d.final_attn = attn_w(h, p + "transformer.final_attn_token_to_image", 128);
  d.nfg = W(h, p + "transformer.norm_final_attn.weight");
  d.nfb = W(h, p + "transformer.norm_final_attn.bias");
  if (int r = make_kv_tables(h, d.final_attn, &d.final_kv_w, &d.final_kv_tab)) return r;

  // output_upscaling (mask_decoder.py:54-64): ConvTranspose2d(k2,s2) == GEMM with N = (ky, kx, out) + pixel shuffle.
  // ConvT weight layout is (in, out, ky, kx).
  {
    std::vector<float> w0(256 * 64 * 4), b0(64), w0p(256 * 256), b0p(256);
    CU(cudaMemcpy(w0.data(), W(h, p + "output_upscaling.0.weight"), w0.size() * 4, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(b0.data(), W(h, p + "output_upscaling.0.bias"), b0.size() * 4, cudaMemcpyDeviceToHost));
    for (int ky = 0; ky < 2; ++ky)
      for (int kx = 0; kx < 2; ++kx)
        for (int o = 0; o < 64; ++o) {
          const int n = (ky * 2 + kx) * 64 + o;
          b0p[n] = b0[o];
          for (int c = 0; c < 256; ++c) w0p[(size_t)n * 256 + c] = w0[((size_t)c * 64 + o) * 4 + ky * 2 + kx];
        }
    CU(dalloc(h, &d.up0_w, w0p.size()));
    CU(dalloc(h, &d.up0_b, b0p.size()));
    CU(cudaMemcpy(d.up0_w, w0p.data(), w0p.size() * 4, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d.up0_b, b0p.data(), b0p.size() * 4, cudaMemcpyHostToDevice));
    std::vector<float> w3(64 * 32 * 4), b3(32), w3p(128 * 64), b3p(128);
    CU(cudaMemcpy(w3.data(), W(h, p + "output_upscaling.3.weight"), w3.size() * 4, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(b3.data(), W(h, p + "output_upscaling.3.bias"), b3.size() * 4, cudaMemcpyDeviceToHost));
    for (int ky = 0; ky < 2; ++ky)
      for (int kx = 0; kx < 2; ++kx)
        for (int o = 0; o < 32; ++o) {
          const int n = (ky * 2 + kx) * 32 + o;
          b3p[n] = b3[o];
          for (int c = 0; c < 64; ++c) w3p[(size_t)n * 64 + c] = w3[((size_t)c * 32 + o) * 4 + ky * 2 + kx];
        }
    CU(dalloc(h, &d.up3_w, w3p.size()));
    CU(dalloc(h, &d.up3_b, b3p.size()));
    CU(cudaMemcpy(d.up3_w, w3p.data(), w3p.size() * 4, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d.up3_b, b3p.data(), b3p.size() * 4, cudaMemcpyHostToDevice));
  }
  d.up1_lng = W(h, p + "output_upscaling.1.weight");
  d.up1_lnb = W(h, p + "output_upscaling.1.bias");
  for (int k = 0; k < 4; ++k)
    for (int j = 0; j < 3; ++j) {
      const std::string q = p + "output_hypernetworks_mlps." + std::to_string(k) + ".layers." + std::to_string(j);
      d.hyp_w[k][j] = W(h, q + ".weight");
      d.hyp_b[k][j] = W(h, q + ".bias");
    }
  for (int j = 0; j < 3; ++j) {
    const std::string q = p + "iou_prediction_head.layers." + std::to_string(j);
    d.iou_w[j] = W(h, q + ".weight");
    d.iou_b[j] = W(h, q + ".bias");
    if (s == 0) {
      const std::string x = p + "taxonomy_embed.layers." + std::to_string(j);
      d.tax_w[j] = W(h, x + ".weight");
      d.tax_b[j] = W(h, x + ".bias");
    } else {
      d.tax_w[j] = nullptr;
      d.tax_b[j] = nullptr;
    }
  }
  return 0;
}

// bf16 GEMM operand packs, padded rel-pos tables and the permuted neck conv weight of the image encoder
int finalize_encoder(b200sam_t* h) {
  NvtxRange nvtx_("finalize_encoder");
  free_list(h->derived_enc);
  AllocScope scope(h, &h->derived_enc);
  const long D = h->dim;
  const std::string e = "image_encoder.";
  if (int r = make_bf16(h, W(h, e + "patch_embed.proj.weight"), D * 768, &h->patch_w)) return r;
  h->patch_b = W(h, e + "patch_embed.proj.bias");
  h->pos = W(h, e + "pos_embed");
  h->blk.assign(h->depth, EncBlock());
  for (int i = 0; i < h->depth; ++i) {
    EncBlock& b = h->blk[i];
    const std::string p = e + "blocks." + std::to_string(i) + ".";
    for (int gi : h->global_idx) b.global |= (gi == i);
    b.ln1g = W(h, p + "norm1.weight"); b.ln1b = W(h, p + "norm1.bias");
    b.ln2g = W(h, p + "norm2.weight"); b.ln2b = W(h, p + "norm2.bias");
    b.qkv_b = W(h, p + "attn.qkv.bias"); b.proj_b = W(h, p + "attn.proj.bias");
    b.lin1_b = W(h, p + "mlp.lin1.bias"); b.lin2_b = W(h, p + "mlp.lin2.bias");
    if (int r = make_bf16(h, W(h, p + "attn.qkv.weight"), 3 * D * D, &b.qkv_w)) return r;
    if (int r = make_bf16(h, W(h, p + "attn.proj.weight"), D * D, &b.proj_w)) return r;
    if (int r = make_bf16(h, W(h, p + "mlp.lin1.weight"), 4 * D * D, &b.lin1_w)) return r;
    if (int r = make_bf16(h, W(h, p + "mlp.lin2.weight"), 4 * D * D, &b.lin2_w)) return r;
    CU(dalloc(h, &b.qkv_wf, (size_t)3 * D * D));
    CU(dalloc(h, &b.qkv_c, (size_t)3 * D));
    CU(dalloc(h, &b.qkv_d, (size_t)3 * D));
    KL(launch_fold_ln_weights(W(h, p + "attn.qkv.weight"), b.ln1g, b.ln1b, b.qkv_b, (int)(3 * D), (int)D, b.qkv_wf, b.qkv_c,
                              b.qkv_d, 0));
    CU(dalloc(h, &b.lin1_wf, (size_t)4 * D * D));
    CU(dalloc(h, &b.lin1_c, (size_t)4 * D));
    CU(dalloc(h, &b.lin1_d, (size_t)4 * D));
    KL(launch_fold_ln_weights(W(h, p + "mlp.lin1.weight"), b.ln2g, b.ln2b, b.lin1_b, (int)(4 * D), (int)D, b.lin1_wf,
                              b.lin1_c, b.lin1_d, 0));
    const int S = b.global ? kGrid : kWin;
    const int RP = b.global ? 128 : 32;
    CU(dalloc(h, &b.relh, (size_t)2 * RP * h->hd));   // [relh ; relw] contiguous: one tensor map covers both
    b.relw = b.relh + (size_t)RP * h->hd;
    KL(launch_relpos_pack(W(h, p + "attn.rel_pos_h"), S, h->hd, b.relh, 0));
    KL(launch_relpos_pack(W(h, p + "attn.rel_pos_w"), S, h->hd, b.relw, 0));
    if (b.global) {
      CU(dalloc(h, &b.relg, (size_t)256 * 128));
      KL(launch_relpos_pack_g(W(h, p + "attn.rel_pos_h"), W(h, p + "attn.rel_pos_w"), h->hd, b.relg, 0));
    }
  }
  if (int r = make_bf16(h, W(h, e + "neck.0.weight"), 256 * D, &h->neck0_w)) return r;
  {
    // neck.2 Conv2d 3x3 weight (out, in, dy, dx) -> (out, dy, dx, in): the K order of launch_im2col_3x3
    std::vector<float> w((size_t)256 * 256 * 9), wp(w.size());
    CU(cudaMemcpy(w.data(), W(h, e + "neck.2.weight"), w.size() * 4, cudaMemcpyDeviceToHost));
    for (int o = 0; o < 256; ++o)
      for (int c = 0; c < 256; ++c)
        for (int t = 0; t < 9; ++t) wp[((size_t)o * 9 + t) * 256 + c] = w[((size_t)o * 256 + c) * 9 + t];
    float* tmp = nullptr;
    CU(dalloc(h, &tmp, wp.size()));
    CU(cudaMemcpy(tmp, wp.data(), wp.size() * 4, cudaMemcpyHostToDevice));
    if (int r = make_bf16(h, tmp, (long)wp.size(), &h->neck2_w)) return r;
  }
  h->neck1g = W(h, e + "neck.1.weight"); h->neck1b = W(h, e + "neck.1.bias");
  h->neck3g = W(h, e + "neck.3.weight"); h->neck3b = W(h, e + "neck.3.bias");
  return 0;
}

// prompt-encoder constants + both mask decoders
int finalize_decoders(b200sam_t* h) {
  NvtxRange nvtx_("finalize_decoders");
  free_list(h->derived_dec);
  AllocScope scope(h, &h->derived_dec);
  // prompt encoder constants (prompt_encoder.py:67-76,182-184): dense PE once, in fp32
  CU(dalloc(h, &h->pe_tok, (size_t)kTokens * 256));
  KL(launch_dense_pe(W(h, "prompt_encoder.pe_layer.positional_encoding_gaussian_matrix"), h->pe_tok, 0));
  h->no_mask = W(h, "prompt_encoder.no_mask_embed.weight");
  for (int s = 0; s < 2; ++s)
    if (int r = finalize_side(h, s)) return r;
  return 0;
}

int alloc_workspace(b200sam_t* h) {
  const size_t B = h->max_batch, P = h->max_prompts, D = h->dim;
  CU(dalloc(h, &h->X, B * kTokens * D));
  CU(dalloc(h, &h->ln_plain, B * kTokens * D));
  CU(dalloc(h, &h->qkv, B * kWinRowsPerFrame * 3 * D));
  CU(dalloc(h, &h->attn_o, B * kWinRowsPerFrame * D));
  {
    std::vector<int> pads;
    for (int rem = 0; rem < kWinRowsPerFrame; ++rem) {
      const int win = rem / kWinTokens, t = rem - win * kWinTokens;
      const int y = (win / kWinPerSide) * kWin + t / kWin, x = (win % kWinPerSide) * kWin + t % kWin;
      if (y >= kGrid || x >= kGrid) pads.push_back(rem);
    }
    h->n_pad_rows = (int)pads.size();
    CU(dalloc(h, &h->pad_rows, pads.size()));
    CU(cudaMemcpy(h->pad_rows, pads.data(), pads.size() * sizeof(int), cudaMemcpyHostToDevice));
  }
  {
    __half* g = nullptr;
    CU(dalloc(h, &g, B * kTokens * (size_t)h->heads * 256));
    h->gbuf = g;
  }
  {
    void* flag = nullptr;
    CU(cudaHostAlloc(&flag, sizeof(int), cudaHostAllocMapped));
    h->ovf_host = reinterpret_cast<volatile int*>(flag);
    *h->ovf_host = 0;
    void* dflag = nullptr;
    CU(cudaHostGetDevicePointer(&dflag, flag, 0));
    h->ovf_dev = reinterpret_cast<int*>(dflag);
  }
  const size_t hid_cols = 4 * D > 9 * 256 ? 4 * D : 9 * 256;
  CU(dalloc(h, &h->hid, B * kTokens * hid_cols));
  // neck.1 output in the zero-ringed (B, 66, 66, 256) layout the 3x3 convolution reads through a 4-D tensor map; the
  // ring is zeroed here once and never written again
  h->ln_slots = 2 * (int)((D + 255) / 256);
  CU(dalloc(h, &h->ln_stats, B * kTokens * (size_t)h->ln_slots * 2));
  CU(dalloc(h, &h->zpad, B * 66 * 66 * 256, /*zero=*/true));
  CU(cudaDeviceSynchronize());
  CU(dalloc(h, &h->emb_tok, B * kTokens * 256));
  CU(dalloc(h, &h->emb_in, B * kTokens * 256));
  if (P > 0) {
    CU(dalloc(h, &h->prompt_frame, P));
    for (int i = 0; i < 2; ++i) {
      DecWorkspace& w = h->dws[i];
      CU(dalloc(h, &w.keys, P * kTokens * 256));
      CU(dalloc(h, &w.kv, P * kTokens * 256));
      CU(dalloc(h, &w.qimg, P * kTokens * 128));
      CU(dalloc(h, &w.oimg, P * kTokens * 128));
      CU(dalloc(h, &w.up1, P * 16384 * 64));
      CU(dalloc(h, &w.tok_state, P * 6 * 256));
      CU(dalloc(h, &w.tok_pe, P * 6 * 256));
      CU(dalloc(h, &w.tok_q, P * 6 * 128));
      CU(dalloc(h, &w.tok_o, P * 6 * 128));
      CU(dalloc(h, &w.tok_kv, P * 6 * 256));
      CU(dalloc(h, &w.hyper, P * 32));
    }
    CU(cudaStreamCreateWithFlags(&h->side_stream, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming));
  }
  return 0;
}

int create_common(b200sam_t** out, int dim, int depth, int heads, const int* gidx, int ng, int device, int max_batch,
                  int max_prompts) {
  if (!out) return B200SAM_ERR_INVALID;
  *out = nullptr;
  b200sam_t* h = new b200sam();
  *out = h;   // returned even on failure so the caller can read last_error, then destroy
  if (dim <= 0 || depth <= 0 || heads <= 0 || dim % heads != 0 || dim % 64 != 0 || dim > 1280 || max_batch <= 0 ||
      max_prompts < 0)
    return fail(h, B200SAM_ERR_INVALID, "bad hyper-parameters dim=%d depth=%d heads=%d max_batch=%d max_prompts=%d", dim,
                depth, heads, max_batch, max_prompts);
  h->dim = dim;
  h->depth = depth;
  h->heads = heads;
  h->hd = dim / heads;
  if (h->hd != 64 && h->hd != 80) return fail(h, B200SAM_ERR_INVALID, "head dim %d not in {64, 80}", h->hd);
  for (int i = 0; i < ng; ++i) {
    if (gidx[i] < 0 || gidx[i] >= depth) return fail(h, B200SAM_ERR_INVALID, "global index %d out of range", gidx[i]);
    h->global_idx.push_back(gidx[i]);
  }
  h->device = device;
  h->max_batch = max_batch;
  h->max_prompts = max_prompts;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    return fail(h, B200SAM_ERR_UNSUPPORTED, "no CUDA device: libb200sam has no CPU fallback");
  if (device < 0 || device >= ndev) return fail(h, B200SAM_ERR_INVALID, "device %d of %d", device, ndev);
  cudaDeviceProp prop;
  CU(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10)
    return fail(h, B200SAM_ERR_UNSUPPORTED, "device %d is sm_%d%d; libb200sam is built for sm_100a only", device,
                prop.major, prop.minor);
  h->n_sm = prop.multiProcessorCount;
  DeviceGuard g(device);
  CU(gemm_tc_init(device));
  CU(attention_init());
  CU(decoder_init());
  return alloc_workspace(h);
}

int check_ready(b200sam_t* h) {
  if (!h) return B200SAM_ERR_INVALID;
  if (!h->finalized) return fail(h, B200SAM_ERR_STATE, "weights not finalised: call b200sam_finalize_weights first");
  if (h->ovf_host && *h->ovf_host) {
    *h->ovf_host = 0;
    return fail(h, B200SAM_ERR_STATE,
                "an earlier b200sam_encode saturated the fp16 rel-pos projections q.Rh / q.Rw (|value| > 65504) in a "
                "global-attention block: its embeddings are invalid (checkpoint outside the supported range)");
  }
  return 0;
}

int run_gemm(b200sam_t* h, bool tf32, int mode, const void* A, int lda, const void* Wt, int ldw, long M, int N, int K,
             const GemmEpi& e, cudaStream_t s) {
  if (M > 0x7fffffffL) return fail(h, B200SAM_ERR_INVALID, "GEMM M too large");
  KL(gemm_tc(tf32, mode, A, lda, Wt, ldw, (int)M, N, K, e, s, h->debug_simt));
  return 0;
}

}  // namespace

// ------------------------------------------------------------------------------------------------ lifecycle

extern "C" int b200sam_create(b200sam_t** out, int arch, int device, int max_batch, int max_prompts) {
  // build_sam.py:15-48
  static const int gb[4] = {2, 5, 8, 11}, gl[4] = {5, 11, 17, 23}, gh[4] = {7, 15, 23, 31};
  switch (arch) {
    case B200SAM_VIT_B: return create_common(out, 768, 12, 12, gb, 4, device, max_batch, max_prompts);
    case B200SAM_VIT_L: return create_common(out, 1024, 24, 16, gl, 4, device, max_batch, max_prompts);
    case B200SAM_VIT_H: return create_common(out, 1280, 32, 16, gh, 4, device, max_batch, max_prompts);
    default:
      if (out) {
        *out = new b200sam();
        (*out)->err = "unknown arch id";
      }
      return B200SAM_ERR_INVALID;
  }
}

extern "C" int b200sam_create_custom(b200sam_t** out, int dim, int depth, int heads, const int* host_global_idx,
                                     int n_global, int device, int max_batch, int max_prompts) {
  return create_common(out, dim, depth, heads, host_global_idx, n_global, device, max_batch, max_prompts);
}

extern "C" void b200sam_destroy(b200sam_t* h) {
  if (!h) return;
  {
    DeviceGuard g(h->device);
    free_list(h->owned);
    free_list(h->derived_enc);
    free_list(h->derived_dec);
    for (auto& kv : h->enc_graphs)
      if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
    if (h->graph_stream) cudaStreamDestroy(h->graph_stream);
    if (h->ev_g0) cudaEventDestroy(h->ev_g0);
    if (h->ev_g1) cudaEventDestroy(h->ev_g1);
    if (h->ovf_host) cudaFreeHost(const_cast<int*>(h->ovf_host));
    if (h->side_stream) cudaStreamDestroy(h->side_stream);
    if (h->ev_fork) cudaEventDestroy(h->ev_fork);
    if (h->ev_join) cudaEventDestroy(h->ev_join);
    for (auto& ev : h->prof_ev) {
      cudaEventDestroy(ev.first);
      cudaEventDestroy(ev.second);
    }
    for (auto& kv : h->w) cudaFree(kv.second.d);
  }
  delete h;
}

extern "C" const char* b200sam_last_error(const b200sam_t* h) { return h ? h->err.c_str() : "null handle"; }

extern "C" int64_t b200sam_launch_count(const b200sam_t* h) { return h ? h->launches : 0; }

extern "C" int b200sam_set_debug_simt_gemm(b200sam_t* h, int on) {
  if (!h) return B200SAM_ERR_INVALID;
  h->debug_simt = on != 0;
  return 0;
}

extern "C" int b200sam_profile_enable(b200sam_t* h, int on) {
  if (!h) return B200SAM_ERR_INVALID;
  h->prof_on = on != 0;
  h->prof_used = 0;
  return 0;
}

extern "C" int b200sam_profile_read(b200sam_t* h, float* host_ms, int64_t* host_count, int n_classes) {
  if (!h || !host_ms || !host_count || n_classes < B200SAM_PC_COUNT) return B200SAM_ERR_INVALID;
  DeviceGuard g(h->device);
  for (int i = 0; i < n_classes; ++i) {
    host_ms[i] = 0.f;
    host_count[i] = 0;
  }
  for (size_t i = 0; i < h->prof_used; ++i) {
    CU(cudaEventSynchronize(h->prof_ev[i].second));
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, h->prof_ev[i].first, h->prof_ev[i].second));
    host_ms[h->prof_cls[i]] += ms;
    host_count[h->prof_cls[i]] += 1;
  }
  h->prof_used = 0;
  return 0;
}

extern "C" int b200sam_debug_stop_after_block(b200sam_t* h, int nblocks) {
  if (!h) return B200SAM_ERR_INVALID;
  h->debug_stop_block = nblocks;
  return 0;
}

extern "C" int b200sam_debug_read_tokens(b200sam_t* h, float* out, int B, void* stream) {
  if (!h || !out || B < 0 || B > h->max_batch) return B200SAM_ERR_INVALID;
  DeviceGuard g(h->device);
  CU(cudaMemcpyAsync(out, h->X, (size_t)B * kTokens * h->dim * sizeof(float), cudaMemcpyDeviceToDevice,
                     reinterpret_cast<cudaStream_t>(stream)));
  return 0;
}

// ------------------------------------------------------------------------------------------------ weights

extern "C" int b200sam_load_weight(b200sam_t* h, const char* name, const void* ptr, int dtype,
                                   const int64_t* host_shape, int ndim) {
  if (!h) return B200SAM_ERR_INVALID;
  if (!name || !ptr || !host_shape || ndim < 0 || ndim > 8) return fail(h, B200SAM_ERR_INVALID, "bad load_weight args");
  DeviceGuard g(h->device);
  std::vector<ManifestEntry> man = build_manifest(h);
  const ManifestEntry* me = nullptr;
  for (const auto& m : man)
    if (m.name == name) me = &m;
  if (!me) return fail(h, B200SAM_ERR_INVALID, "unknown weight name '%s'", name);
  std::vector<int64_t> shape(host_shape, host_shape + ndim);
  if (shape != me->shape) {
    std::string got, want;
    for (auto v : shape) got += std::to_string(v) + ",";
    for (auto v : me->shape) want += std::to_string(v) + ",";
    return fail(h, B200SAM_ERR_INVALID, "shape mismatch for '%s': got (%s) want (%s)", name, got.c_str(), want.c_str());
  }
  long n = 1;
  for (auto v : shape) n *= v;
  const int esz = dtype == B200SAM_F32 ? 4 : (dtype == B200SAM_BF16 || dtype == B200SAM_F16 ? 2 : 0);
  if (!esz) return fail(h, B200SAM_ERR_INVALID, "bad dtype %d for '%s'", dtype, name);
  // Not a hot-path call: the masters may still be read by kernels of an earlier compute call (the decoders use the
  // fp32 masters directly) and the copies below run on the legacy stream, so quiesce the device when the first tensor
  // of a re-load arrives.  The caller synchronises the stream that produced `ptr` (sam.py::_sync_weights does).
  if (h->finalized) CU(cudaDeviceSynchronize());
  if (strncmp(name, "image_encoder.", 14) == 0) h->dirty_enc = true;
  else h->dirty_dec = true;
  Tensor& t = h->w[name];
  if (!t.d) {
    void* q = nullptr;
    CU(cudaMalloc(&q, (size_t)n * 4 + 256));
    t.d = reinterpret_cast<float*>(q);
  }
  t.shape = shape;
  t.numel = n;
  if (dtype == B200SAM_F32) {
    CU(cudaMemcpy(t.d, ptr, (size_t)n * 4, cudaMemcpyDefault));
  } else {
    void* tmp = nullptr;
    CU(cudaMalloc(&tmp, (size_t)n * 2 + 256));
    cudaError_t e = cudaMemcpy(tmp, ptr, (size_t)n * 2, cudaMemcpyDefault);
    if (e == cudaSuccess) e = launch_to_f32(tmp, dtype, t.d, n, 0);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    cudaFree(tmp);
    CU(e);
    ++h->launches;
  }
  h->finalized = false;
  return 0;
}

extern "C" int b200sam_finalize_weights(b200sam_t* h) {
  if (!h) return B200SAM_ERR_INVALID;
  DeviceGuard g(h->device);
  const bool prof_was = h->prof_on;
  h->prof_on = false;   // one-time packing kernels are not part of any profile
  for (const auto& m : build_manifest(h))
    if (m.required && !h->w.count(m.name)) return fail(h, B200SAM_ERR_STATE, "missing weight '%s'", m.name.c_str());
  // Buffers derived by a previous finalize are freed and rebuilt per group (encoder / prompt encoder + decoders),
  // and only for a group one of whose tensors was re-loaded: sizes never change, the footprint stays flat.
  CU(cudaDeviceSynchronize());
  if (h->dirty_enc) {
    if (int r = finalize_encoder(h)) return r;
    h->dirty_enc = false;
  }
  if (h->dirty_dec) {
    if (int r = finalize_decoders(h)) return r;
    h->dirty_dec = false;
  }
  CU(cudaDeviceSynchronize());
  for (auto& kv : h->enc_graphs)       // captured launches hold pointers into the buffers rebuilt above
    if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
  h->enc_graphs.clear();
  h->prof_on = prof_was;
  h->finalized = true;
  h->have_emb = false;
  return 0;
}

// ------------------------------------------------------------------------------------------------ encode

namespace {

// PatchEmbed (image_encoder.py:111-113, 395-426) + absolute position embedding (:115-116) as one GEMM with bias and
// pos_embed in the epilogue.  bf16 frames are read by the GEMM's own TMA loads through a 5-D tensor map (AM_PATCH16,
// no im2col copy); fp32 / fp16 frames, widths that are not a multiple of 256 and the SIMT bisect path go through the
// im2col index map, which also casts to bf16.  Reads the caller's frames, so it stays outside the CUDA graph.
// LayerNorm folded into the consuming GEMMs (gemm_tc.cuh GemmEpi::stats_in): the shipped path.  Debug variant 5 = 1 and
// the SIMT bisect path use the stand-alone LayerNorm kernel.
bool ln_fold_on(const b200sam_t* h) { return kernel_variant(VK_LN_FOLD) == 0 && !h->debug_simt; }

int encode_patch(b200sam_t* h, const void* frames, int frames_dtype, int B, cudaStream_t s) {
  NvtxRange nvtx_("patch_embed");
  const int D = h->dim;
  const long T = (long)B * kTokens;
  GemmEpi e;
  e.bias = h->patch_b;
  e.addend = h->pos;
  e.ld_add = D;
  e.add_mod = kTokens;
  e.out = h->X;
  e.ld_out = D;
  const bool fold = ln_fold_on(h);
  if (fold) {          // block 0's norm1 is folded into its qkv GEMM: emit the bf16 rows and the row sums here
    e.xb_out = h->ln_plain;
    e.stats_out = h->ln_stats;
    e.stats_slots = h->ln_slots;
  }
  const int mode = fold ? GM_F32_FOLD : GM_F32;
  const bool implicit = frames_dtype == DT_BF16 && D % 256 == 0 && !h->debug_simt &&
                        kernel_variant(VK_IMPLICIT_CONV) == 0 && (reinterpret_cast<uintptr_t>(frames) & 15) == 0;
  if (implicit) {
    e.a_mode = AM_PATCH16;
    h->pc = B200SAM_PC_GEMM_PATCH;
    return run_gemm(h, false, mode, frames, 768, h->patch_w, 768, T, D, 768, e, s);
  }
  h->pc = B200SAM_PC_ELEMENTWISE;
  KL(launch_im2col_patch(frames, frames_dtype, B, h->hid, s));
  h->pc = B200SAM_PC_GEMM_PATCH;
  return run_gemm(h, false, mode, h->hid, 768, h->patch_w, 768, T, D, 768, e, s);
}

// Everything of ImageEncoderViT.forward that reads and writes only the handle's own buffers: the
// transformer blocks, the neck.  No caller pointer appears in any launch, which is what makes the sequence
// capturable into a CUDA graph (b200sam_encode).  Returns 1 when a debug stop cut the sequence short.
int encode_body(b200sam_t* h, int B, cudaStream_t s) {
  const int D = h->dim, heads = h->heads, hd = h->hd;
  const long T = (long)B * kTokens;
  const int nblocks = (h->debug_stop_block >= 0 && h->debug_stop_block < h->depth) ? h->debug_stop_block : h->depth;
  const bool fold = ln_fold_on(h);
  // window-pad rows of a windowed block's qkv buffer: written by the idle warp of the PREVIOUS block's lin2 launch
  // (GemmEpi::pad_dst) when that launch runs on the CTA-pair kernel, else by the small fill kernel behind the qkv GEMM
  bool pads_written = false;
  for (int i = 0; i < nblocks; ++i) {
    const EncBlock& b = h->blk[i];
    NvtxRange nvtx_blk(b.global ? "block_global" : "block_windowed");
    // norm1 in token order.  The windowed blocks never materialise the zero-padded window tensor: the qkv epilogue
    // scatters the 4096 real rows of a frame into the 4900-row window layout (window_partition, image_encoder.py:263-288),
    // the 804 pad rows (zero after norm1 => qkv = bias, :179-183) are filled by a small kernel, and the attention
    // kernel stores straight back into token order (window_unpartition), so qkv and proj run on 4096 rows per frame.
    h->pc = B200SAM_PC_LAYERNORM;
    if (!fold) KL(launch_layernorm_bf16(h->X, T, D, b.ln1g, b.ln1b, 1e-6f, h->ln_plain, false, s));
    {
      GemmEpi e;
      e.bias = fold ? b.qkv_d : b.qkv_b;
      if (fold) {
        e.stats_in = h->ln_stats;
        e.stats_slots = h->ln_slots;
        e.stats_dim = D;
        e.fold_c = b.qkv_c;
        e.fold_eps = 1e-6f;
      }
      e.out = h->qkv;
      e.ld_out = 3 * D;
      e.rowmap = b.global ? RM_IDENT : RM_TOKEN_TO_WINDOW;
      h->pc = B200SAM_PC_GEMM_QKV;
      if (int r = run_gemm(h, false, GM_BIAS_BF16, h->ln_plain, D, fold ? b.qkv_wf : b.qkv_w, D, T, 3 * D, D, e, s)) return r;
      if (!b.global && !pads_written) KL(launch_fill_pad_rows(h->qkv, b.qkv_b, h->pad_rows, h->n_pad_rows, B, 3 * D, s));
      pads_written = false;
    }
    h->pc = b.global ? B200SAM_PC_ATTN_GLOBAL : B200SAM_PC_ATTN_WINDOW;
    KL(launch_attention(h->qkv, b.global ? B : B * kWinPerSide * kWinPerSide, b.global ? kGrid : kWin, heads, hd, b.relh,
                        b.relw, b.relg, h->gbuf, h->attn_o, s, /*window_token_out=*/true, h->ovf_dev));
    if (b.global) ++h->launches;   // the batched bias-projection GEMM in front of the global kernel
    {
      // proj + residual: x = shortcut + attn(x)                                    (image_encoder.py:187-190)
      GemmEpi e;
      e.bias = b.proj_b;
      e.addend = h->X;
      e.ld_add = D;
      e.out = h->X;
      e.ld_out = D;
      if (fold) {
        e.xb_out = h->ln_plain;
        e.stats_out = h->ln_stats;
        e.stats_slots = h->ln_slots;
      }
      h->pc = B200SAM_PC_GEMM_PROJ;
      if (int r = run_gemm(h, false, fold ? GM_F32_FOLD : GM_F32, h->attn_o, D, b.proj_w, D, T, D, D, e, s)) return r;
    }
    // x = x + lin2(GELU(lin1(norm2(x))))                                          (image_encoder.py:191, common.py:25-26)
    h->pc = B200SAM_PC_LAYERNORM;
    if (!fold) KL(launch_layernorm_bf16(h->X, T, D, b.ln2g, b.ln2b, 1e-6f, h->ln_plain, false, s));
    {
      GemmEpi e;
      e.bias = fold ? b.lin1_d : b.lin1_b;
      if (fold) {
        e.stats_in = h->ln_stats;
        e.stats_slots = h->ln_slots;
        e.stats_dim = D;
        e.fold_c = b.lin1_c;
        e.fold_eps = 1e-6f;
      }
      e.out = h->hid;
      e.ld_out = 4 * D;
      h->pc = B200SAM_PC_GEMM_LIN1;
      if (int r = run_gemm(h, false, GM_BIAS_GELU_BF16, h->ln_plain, D, fold ? b.lin1_wf : b.lin1_w, D, T, 4 * D, D, e, s)) return r;
    }
    {
      GemmEpi e;
      e.bias = b.lin2_b;
      e.addend = h->X;
      e.ld_add = D;
      e.out = h->X;
      e.ld_out = D;
      if (fold) {          // the bf16 rows feed the next block's qkv (or the neck), the sums its norm1
        e.xb_out = h->ln_plain;
        e.stats_out = h->ln_stats;
        e.stats_slots = h->ln_slots;
      }
      h->pc = B200SAM_PC_GEMM_LIN2;
      const int lin2_mode = fold ? GM_F32_FOLD : GM_F32;
      if (i + 1 < nblocks && !h->blk[i + 1].global && kernel_variant(VK_PAD_FILL) == 0 && !h->debug_simt &&
          gemm_tc_is_pair(false, lin2_mode, (int)T, D, e, false)) {
        // attention(i) is done with the qkv buffer; nothing reads it before block i + 1's attention
        e.pad_dst = h->qkv;
        e.pad_bias = h->blk[i + 1].qkv_b;
        e.pad_rows = h->pad_rows;
        e.pad_n = h->n_pad_rows;
        e.pad_frames = B;
        e.pad_cols = 3 * D;
        pads_written = true;
      }
      if (int r = run_gemm(h, false, lin2_mode, h->hid, 4 * D, b.lin2_w, 4 * D, T, D, 4 * D, e, s)) return r;
    }
  }
  if (nblocks != h->depth) return 1;   // debug: residual stream after `nblocks` blocks is left in X
  // neck (image_encoder.py:92-108): conv1x1 -> LayerNorm2d -> conv3x3 -> LayerNorm2d, token-major
  NvtxRange nvtx_neck("neck");
  h->pc = B200SAM_PC_NECK;
  if (!fold) KL(launch_cast_f32_bf16(h->X, h->ln_plain, T * D, s));     // folded: the last lin2 epilogue wrote the bf16 rows
  // neck.2 Conv2d 3x3: the GEMM gathers its A tiles from the zero-ringed neck.1 output (AM_CONV3); the SIMT bisect
  // path and variant 1 materialise the (T, 9*256) im2col rows instead
  const bool implicit3 = !h->debug_simt && kernel_variant(VK_IMPLICIT_CONV) == 0;
  __nv_bfloat16* z = implicit3 ? h->zpad : h->attn_o;   // (B, 66, 66, 256) / (T, 256) bf16
  {
    GemmEpi e;
    e.rowmap = implicit3 ? RM_TOKEN_TO_PAD66 : RM_IDENT;
    e.out2 = z;
    e.ld_out = 256;
    e.ln_g = h->neck1g;
    e.ln_b = h->neck1b;
    e.ln_eps = 1e-6f;
    if (int r = run_gemm(h, false, GM_LN, h->ln_plain, D, h->neck0_w, D, T, 256, D, e, s)) return r;
  }
  if (!implicit3) KL(launch_im2col_3x3(z, B, h->hid, s));
  {
    GemmEpi e;
    e.a_mode = implicit3 ? AM_CONV3 : AM_PLAIN;
    e.out = h->emb_tok;
    e.ld_out = 256;
    e.ln_g = h->neck3g;
    e.ln_b = h->neck3b;
    e.ln_eps = 1e-6f;
    if (int r = run_gemm(h, false, GM_LN, implicit3 ? z : h->hid, 9 * 256, h->neck2_w, 9 * 256, T, 256, 9 * 256, e, s))
      return r;
  }
  return 0;
}

// The ~300 launches of encode_body replayed as one CUDA graph per batch size: at B = 1 - the
// reference's real call pattern, LISA.py:160-166 - the kernels are short enough for launch overhead and the two
// cuTensorMapEncodeTiled calls per GEMM to show.  The first call of a batch size runs eagerly (it also sets the
// kernels' shared-memory attributes), the second is captured on the caller's stream, later ones replay.  Profiling,
// debug switches and a caller who is capturing the stream himself use the plain launch sequence.
int encode_body_graphed(b200sam_t* h, int B, cudaStream_t s) {
  const int sig = kernel_variant(VK_ATTN_WINDOW) * 10000 + kernel_variant(VK_ATTN_GLOBAL) * 100 +
                  kernel_variant(VK_GEMM_PAIR) * 10 + kernel_variant(VK_IMPLICIT_CONV) + kernel_variant(VK_LN_FOLD) * 1000000 +
                  kernel_variant(VK_PDL) * 10000000 + kernel_variant(VK_PAD_FILL) * 100000000;
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  const bool plain = h->prof_on || h->debug_simt || h->debug_stop_block >= 0 || !h->use_graphs ||
                     cudaStreamIsCapturing(s, &cap) != cudaSuccess || cap != cudaStreamCaptureStatusNone;
  if (plain) return encode_body(h, B, s);
  EncGraph& g = h->enc_graphs[B];
  if (g.exec && g.sig != sig) {
    cudaGraphExecDestroy(g.exec);
    g = EncGraph();
  }
  if (!g.exec && g.calls++ == 0) return encode_body(h, B, s);
  // The caller's stream may be the legacy default stream (PyTorch's default), which can be neither captured nor
  // graph-launched into without serialising the device: capture and replay happen on a stream of the handle, ordered
  // behind / in front of the caller's stream by two events.
  if (!h->graph_stream) {
    CU(cudaStreamCreateWithFlags(&h->graph_stream, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&h->ev_g0, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&h->ev_g1, cudaEventDisableTiming));
  }
  cudaStream_t gs = h->graph_stream;
  CU(cudaEventRecord(h->ev_g0, s));
  CU(cudaStreamWaitEvent(gs, h->ev_g0, 0));
  if (!g.exec) {
    const int64_t before = h->launches;
    CU(cudaStreamBeginCapture(gs, cudaStreamCaptureModeThreadLocal));
    const int r = encode_body(h, B, gs);
    cudaGraph_t graph = nullptr;
    cudaError_t ce = cudaStreamEndCapture(gs, &graph);
    if (r == 0 && ce == cudaSuccess && graph) ce = cudaGraphInstantiate(&g.exec, graph, 0);
    if (graph) cudaGraphDestroy(graph);
    const int64_t captured = h->launches - before;
    h->launches = before;
    if (r != 0 || ce != cudaSuccess || !g.exec) {
      g.exec = nullptr;
      cudaGetLastError();
      h->use_graphs = false;          // same kernels, launched one by one from now on
      return r < 0 ? r : encode_body(h, B, s);
    }
    g.sig = sig;
    g.n_launches = captured;
  }
  CU(cudaGraphLaunch(g.exec, gs));
  h->launches += g.n_launches;
  CU(cudaEventRecord(h->ev_g1, gs));
  CU(cudaStreamWaitEvent(s, h->ev_g1, 0));
  return 0;
}

}  // namespace

extern "C" int b200sam_encode(b200sam_t* h, const void* frames, int frames_dtype, int B, void* emb_out, int emb_dtype,
                              void* stream) {
  NvtxRange nvtx_("b200sam_encode");
  if (int r = check_ready(h)) return r;
  if (B < 0 || B > h->max_batch) return fail(h, B200SAM_ERR_INVALID, "batch %d exceeds max_batch %d", B, h->max_batch);
  if (B == 0) return 0;
  if (!frames) return fail(h, B200SAM_ERR_INVALID, "frames is NULL");
  if (frames_dtype < 0 || frames_dtype > 2 || (emb_out && (emb_dtype < 0 || emb_dtype > 2)))
    return fail(h, B200SAM_ERR_INVALID, "unsupported dtype");
  DeviceGuard g(h->device);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  h->prof_stream = s;

  if (int r = encode_patch(h, frames, frames_dtype, B, s)) return r;
  {
    const int r = encode_body_graphed(h, B, s);
    if (r < 0) return r;
    if (r > 0) return 0;               // debug stop: no embedding
  }
  h->pc = B200SAM_PC_ELEMENTWISE;
  if (emb_out) KL(launch_tokens_to_nchw(h->emb_tok, B, emb_out, emb_dtype, s));
  h->have_emb = true;
  h->emb_batch = B;
  return 0;
}

// ------------------------------------------------------------------------------------------------ decode

namespace {

int decode_side(b200sam_t* h, int sidx, const float* emb_tok, const float* seg, int P, float* lowres, float* iou,
                float* tax, cudaStream_t s) {
  NvtxRange nvtx_(sidx == 0 ? "mask_decoder_left" : "mask_decoder_right");
  const DecSide& d = h->side[sidx];
  const DecWorkspace& w = h->dws[sidx];   // one workspace per side: the two decoders run concurrently
  const long rows = (long)P * kTokens;
  // src = image_embeddings (repeat_interleave per prompt) + dense_prompt_embeddings      (mask_decoder.py:143-144)
  h->pc = B200SAM_PC_DEC_ELEMENTWISE;
  KL(launch_keys_init(emb_tok, h->no_mask, h->prompt_frame, P, w.keys, s));
  for (int l = 0; l < 2; ++l) {
    const DecLayer& L = d.layer[l];
    TokStageA a;
    a.layer = l;
    a.iou_token = d.iou_token;
    a.mask_tokens = d.mask_tokens;
    a.seg = seg;
    a.sa = attn_ptrs(L.self_attn);
    a.n1g = L.n1g;
    a.n1b = L.n1b;
    a.t2i_qw = L.t2i.qw;
    a.t2i_qb = L.t2i.qb;
    a.tok_state = w.tok_state;
    a.tok_pe = w.tok_pe;
    a.tok_q = w.tok_q;
    h->pc = B200SAM_PC_DEC_TOKEN;
    KL(launch_dec_stage_a(a, P, s));
    {
      GemmEpi e;   // [k | v] of the image side for token->image attention
      e.addend = L.kv_tab;
      e.ld_add = 256;
      e.add_mod = kTokens;
      e.out = w.kv;
      e.ld_out = 256;
      h->pc = B200SAM_PC_DEC_GEMM;
      if (int r = run_gemm(h, true, GM_F32, w.keys, 256, L.kv_w, 256, rows, 256, 256, e, s)) return r;
    }
    h->pc = B200SAM_PC_DEC_TOKEN;
    KL(launch_dec_t2i(w.tok_q, w.kv, w.tok_o, P, s));
    TokStageB b;
    b.t2i_ow = L.t2i.ow; b.t2i_ob = L.t2i.ob;
    b.n2g = L.n2g; b.n2b = L.n2b; b.n3g = L.n3g; b.n3b = L.n3b;
    b.mlp1w = L.mlp1w; b.mlp1b = L.mlp1b; b.mlp2w = L.mlp2w; b.mlp2b = L.mlp2b;
    b.i2t_kw = L.i2t.kw; b.i2t_kb = L.i2t.kb; b.i2t_vw = L.i2t.vw; b.i2t_vb = L.i2t.vb;
    b.tok_o = w.tok_o;
    b.tok_pe = w.tok_pe;
    b.tok_state = w.tok_state;
    b.tok_kv = w.tok_kv;
    KL(launch_dec_stage_b(b, P, s));
    {
      GemmEpi e;   // image->token queries
      e.addend = L.q_tab;
      e.ld_add = 128;
      e.add_mod = kTokens;
      e.out = w.qimg;
      e.ld_out = 128;
      h->pc = B200SAM_PC_DEC_GEMM;
      if (int r = run_gemm(h, true, GM_F32, w.keys, 256, L.i2t.qw, 256, rows, 128, 256, e, s)) return r;
    }
    h->pc = B200SAM_PC_DEC_TOKEN;
    KL(launch_dec_i2t(w.qimg, w.tok_kv, w.oimg, P, s));
    {
      GemmEpi e;   // keys = norm4(keys + out_proj(attn))                                  (transformer.py:176-180)
      e.bias = L.i2t.ob;
      e.addend = w.keys;
      e.ld_add = 256;
      e.out = w.keys;
      e.ld_out = 256;
      e.ln_g = L.n4g;
      e.ln_b = L.n4b;
      e.ln_eps = 1e-5f;
      h->pc = B200SAM_PC_DEC_GEMM;
      if (int r = run_gemm(h, true, GM_LN, w.oimg, 128, L.i2t.ow, 128, rows, 256, 128, e, s)) return r;
    }
  }
  // final token->image attention (transformer.py:100-104)
  h->pc = B200SAM_PC_DEC_TOKEN;
  KL(launch_dec_stage_q(w.tok_state, w.tok_pe, d.final_attn.qw, d.final_attn.qb, w.tok_q, P, s));
  {
    GemmEpi e;
    e.addend = d.final_kv_tab;
    e.ld_add = 256;
    e.add_mod = kTokens;
    e.out = w.kv;
    e.ld_out = 256;
    h->pc = B200SAM_PC_DEC_GEMM;
    if (int r = run_gemm(h, true, GM_F32, w.keys, 256, d.final_kv_w, 256, rows, 256, 256, e, s)) return r;
  }
  h->pc = B200SAM_PC_DEC_TOKEN;
  KL(launch_dec_t2i(w.tok_q, w.kv, w.tok_o, P, s));
  TokHeads th;
  th.ow = d.final_attn.ow; th.ob = d.final_attn.ob; th.nfg = d.nfg; th.nfb = d.nfb;
  for (int j = 0; j < 3; ++j) {
    th.hyp_w[j] = d.hyp_w[0][j]; th.hyp_b[j] = d.hyp_b[0][j];
    th.iou_w[j] = d.iou_w[j]; th.iou_b[j] = d.iou_b[j];
    th.tax_w[j] = d.tax_w[j]; th.tax_b[j] = d.tax_b[j];
  }
  th.tok_state = w.tok_state;
  th.tok_o = w.tok_o;
  th.hyper = w.hyper;
  th.iou_out = iou;
  th.tax_out = d.has_taxonomy ? tax : nullptr;
  KL(launch_dec_heads(th, P, s));
  if (lowres) {
    // output_upscaling + hypernetwork product, mask slot 0 only (multimask_output=False, mask_decoder.py:110-115)
    {
      GemmEpi e;
      e.bias = d.up0_b;
      e.ln_g = d.up1_lng;
      e.ln_b = d.up1_lnb;
      e.ln_eps = 1e-6f;
      e.out = w.up1;
      h->pc = B200SAM_PC_DEC_UPSCALE;
      if (int r = run_gemm(h, true, GM_UPS1, w.keys, 256, d.up0_w, 256, rows, 256, 256, e, s)) return r;
    }
    {
      GemmEpi e;
      e.bias = d.up3_b;
      e.hyper = w.hyper;
      e.out = lowres;
      if (int r = run_gemm(h, true, GM_UPS2, w.up1, 64, d.up3_w, 64, (long)P * 16384, 128, 64, e, s)) return r;
    }
  }
  return 0;
}

}  // namespace

extern "C" int b200sam_decode(b200sam_t* h, const void* emb, int emb_dtype, int B, const float* seg_embeds,
                              const int* host_prompt_offsets, float* lowres_left, float* lowres_right, float* iou_left,
                              float* iou_right, float* taxonomy, void* stream) {
  NvtxRange nvtx_("b200sam_decode");
  if (int r = check_ready(h)) return r;
  if (B < 0 || B > h->max_batch) return fail(h, B200SAM_ERR_INVALID, "batch %d exceeds max_batch %d", B, h->max_batch);
  if (!host_prompt_offsets) return fail(h, B200SAM_ERR_INVALID, "host_prompt_offsets is NULL");
  if (host_prompt_offsets[0] != 0) return fail(h, B200SAM_ERR_INVALID, "prompt offsets must start at 0");
  for (int i = 0; i < B; ++i)
    if (host_prompt_offsets[i + 1] < host_prompt_offsets[i])
      return fail(h, B200SAM_ERR_INVALID, "prompt offsets must be non-decreasing");   // LISA.py:193 assert
  const int P = host_prompt_offsets[B];
  if (P > h->max_prompts) return fail(h, B200SAM_ERR_INVALID, "%d prompts exceed max_prompts %d", P, h->max_prompts);
  if (P == 0) return 0;   // no [SEG] token generated: empty outputs (chat.py:227)
  if (!seg_embeds) return fail(h, B200SAM_ERR_INVALID, "seg_embeds is NULL");
  DeviceGuard g(h->device);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  h->prof_stream = s;
  const float* emb_tok = h->emb_tok;
  if (emb) {
    if (emb_dtype < 0 || emb_dtype > 2) return fail(h, B200SAM_ERR_INVALID, "unsupported emb dtype");
    h->pc = B200SAM_PC_DEC_ELEMENTWISE;
    KL(launch_nchw_to_tokens(emb, emb_dtype, B, h->emb_in, s));
    emb_tok = h->emb_in;
  } else if (!h->have_emb || h->emb_batch < B) {
    return fail(h, B200SAM_ERR_STATE, "decode(emb = NULL) needs a preceding encode of at least %d frames", B);
  }
  std::vector<int> pf(P);
  for (int i = 0; i < B; ++i)
    for (int p = host_prompt_offsets[i]; p < host_prompt_offsets[i + 1]; ++p) pf[p] = i;
  // small pageable H2D copy, stream ordered: returns once the data is staged, so `pf` may go out of scope
  CU(cudaMemcpyAsync(h->prompt_frame, pf.data(), (size_t)P * sizeof(int), cudaMemcpyHostToDevice, s));
  const bool left = lowres_left || iou_left || taxonomy, right = lowres_right || iou_right;
  // The two mask decoders are independent (mask_decoder_left / _right, LISA.py:505-530) and individually too small
  // to fill the GPU: the right one is forked onto the handle's side stream and joined back into the caller's
  // stream (event fork/join: no host sync, graph-capturable).  The per-launch profile needs one stream.
  const bool fork = left && right && !h->prof_on && h->side_stream;
  if (fork) {
    CU(cudaEventRecord(h->ev_fork, s));
    CU(cudaStreamWaitEvent(h->side_stream, h->ev_fork, 0));
  }
  if (left)
    if (int r = decode_side(h, 0, emb_tok, seg_embeds, P, lowres_left, iou_left, taxonomy, s)) return r;
  if (right)
    if (int r = decode_side(h, 1, emb_tok, seg_embeds, P, lowres_right, iou_right, nullptr, fork ? h->side_stream : s))
      return r;
  if (fork) {
    CU(cudaEventRecord(h->ev_join, h->side_stream));
    CU(cudaStreamWaitEvent(s, h->ev_join, 0));
  }
  return 0;
}

// ------------------------------------------------------------------------------------------------ postprocess / PE

extern "C" int b200sam_postprocess(b200sam_t* h, const float* lowres, int n_masks, int in_h, int in_w, int out_h,
                                   int out_w, float* logits_out, uint8_t* mask_out, float threshold, void* stream) {
  NvtxRange nvtx_("b200sam_postprocess");
  if (!h) return B200SAM_ERR_INVALID;
  if (n_masks < 0 || in_h < 1 || in_w < 1 || in_h > kImg || in_w > kImg || out_h < 1 || out_w < 1)
    return fail(h, B200SAM_ERR_INVALID, "bad postprocess sizes n=%d in=(%d,%d) out=(%d,%d)", n_masks, in_h, in_w, out_h,
                out_w);
  if (n_masks == 0) return 0;
  if (!lowres) return fail(h, B200SAM_ERR_INVALID, "lowres is NULL");
  DeviceGuard g(h->device);
  h->prof_stream = reinterpret_cast<cudaStream_t>(stream);
  h->pc = B200SAM_PC_POSTPROCESS;
  KL(launch_postprocess(lowres, n_masks, in_h, in_w, out_h, out_w, logits_out, mask_out, nullptr, threshold,
                        reinterpret_cast<cudaStream_t>(stream)));
  return 0;
}

extern "C" int b200sam_postprocess_bits(b200sam_t* h, const float* lowres, int n_masks, int in_h, int in_w, int out_h,
                                        int out_w, uint8_t* bits_out, float threshold, void* stream) {
  NvtxRange nvtx_("b200sam_postprocess");
  if (!h) return B200SAM_ERR_INVALID;
  if (n_masks < 0 || in_h < 1 || in_w < 1 || in_h > kImg || in_w > kImg || out_h < 1 || out_w < 1)
    return fail(h, B200SAM_ERR_INVALID, "bad postprocess sizes n=%d in=(%d,%d) out=(%d,%d)", n_masks, in_h, in_w, out_h,
                out_w);
  if (n_masks == 0) return 0;
  if (!lowres || !bits_out) return fail(h, B200SAM_ERR_INVALID, "lowres / bits_out is NULL");
  DeviceGuard g(h->device);
  h->prof_stream = reinterpret_cast<cudaStream_t>(stream);
  h->pc = B200SAM_PC_POSTPROCESS;
  KL(launch_postprocess(lowres, n_masks, in_h, in_w, out_h, out_w, nullptr, nullptr, bits_out, threshold,
                        reinterpret_cast<cudaStream_t>(stream)));
  return 0;
}

extern "C" int b200sam_postprocess_pair(b200sam_t* h, const float* lowres_left, const float* lowres_right, int P,
                                        int in_h, int in_w, int out_h, int out_w, uint8_t* bits_left,
                                        uint8_t* bits_right, float threshold, const float* taxonomy,
                                        const uint8_t* gt, const int* gt_index, int64_t* counts_out,
                                        uint8_t* pred_bits_out, void* stream) {
  NvtxRange nvtx_("b200sam_postprocess");
  if (!h) return B200SAM_ERR_INVALID;
  if (P < 0 || in_h < 1 || in_w < 1 || in_h > kImg || in_w > kImg || out_h < 1 || out_w < 1)
    return fail(h, B200SAM_ERR_INVALID, "bad postprocess sizes P=%d in=(%d,%d) out=(%d,%d)", P, in_h, in_w, out_h, out_w);
  if (P == 0) return 0;
  if (!lowres_left || !lowres_right) return fail(h, B200SAM_ERR_INVALID, "lowres is NULL");
  if (gt && (!taxonomy || !counts_out))
    return fail(h, B200SAM_ERR_INVALID, "evaluation against gt needs taxonomy and counts_out");
  if (!gt && !bits_left && !bits_right) return fail(h, B200SAM_ERR_INVALID, "nothing to compute: no output given");
  DeviceGuard g(h->device);
  h->prof_stream = reinterpret_cast<cudaStream_t>(stream);
  h->pc = B200SAM_PC_POSTPROCESS;
  KL(launch_postprocess_pair(lowres_left, lowres_right, P, in_h, in_w, out_h, out_w, bits_left, bits_right, threshold,
                             taxonomy, gt, gt_index, counts_out, pred_bits_out, reinterpret_cast<cudaStream_t>(stream)));
  return 0;
}

extern "C" int b200sam_checksum64(b200sam_t* h, const void* data, int n_items, int64_t bytes, int64_t stride,
                                  uint64_t* out, void* stream) {
  NvtxRange nvtx_("b200sam_checksum64");
  if (!h) return B200SAM_ERR_INVALID;
  if (n_items < 0 || n_items > 65535 || bytes < 0 || stride < bytes)
    return fail(h, B200SAM_ERR_INVALID, "bad checksum sizes n=%d bytes=%lld stride=%lld", n_items, (long long)bytes,
                (long long)stride);
  if (n_items == 0) return 0;
  if (!data || !out) return fail(h, B200SAM_ERR_INVALID, "NULL argument");
  DeviceGuard g(h->device);
  h->prof_stream = reinterpret_cast<cudaStream_t>(stream);
  h->pc = B200SAM_PC_METRICS;
  KL(launch_checksum64(reinterpret_cast<const uint8_t*>(data), n_items, bytes, stride, out,
                       reinterpret_cast<cudaStream_t>(stream)));
  return 0;
}

extern "C" int b200sam_dense_pe(b200sam_t* h, float* out, void* stream) {
  if (int r = check_ready(h)) return r;
  if (!out) return fail(h, B200SAM_ERR_INVALID, "out is NULL");
  DeviceGuard g(h->device);
  h->prof_stream = reinterpret_cast<cudaStream_t>(stream);
  h->pc = B200SAM_PC_ELEMENTWISE;
  KL(launch_tokens_to_nchw(h->pe_tok, 1, out, DT_F32, reinterpret_cast<cudaStream_t>(stream)));
  return 0;
}

// ------------------------------------------------------------------------------------------------ callers either side

namespace {

int upload_ints(b200sam_t* h, const std::vector<int>& v, int** out) {
  CU(dalloc(h, out, v.size() ? v.size() : 1));
  if (!v.empty()) CU(cudaMemcpy(*out, v.data(), v.size() * sizeof(int), cudaMemcpyHostToDevice));
  return 0;
}

// transforms.py:103-113
void preprocess_shape(int oldh, int oldw, int* newh, int* neww) {
  const double scale = kImg * 1.0 / (oldh > oldw ? oldh : oldw);
  *newh = (int)(oldh * scale + 0.5);
  *neww = (int)(oldw * scale + 0.5);
}

// Pillow coefficient tables of one raw frame size, built and uploaded the first time that size is seen.
int resample_tables(b200sam_t* h, int H, int W, const ResampleTables** out) {
  auto it = h->resample.find({H, W});
  if (it == h->resample.end()) {
    ResampleTables t;
    t.H = H;
    t.W = W;
    preprocess_shape(H, W, &t.new_h, &t.new_w);
    if (t.new_h < 1 || t.new_w < 1 || t.new_h > kImg || t.new_w > kImg)
      return fail(h, B200SAM_ERR_INVALID, "frame %dx%d resizes to %dx%d", H, W, t.new_h, t.new_w);
    ResampleAxis ax, ay;
    build_resample_axis(W, t.new_w, &ax);
    build_resample_axis(H, t.new_h, &ay);
    t.kx = ax.ksize;
    t.ky = ay.ksize;
    t.ident_x = ax.identity;
    t.ident_y = ay.identity;
    t.tile_rows = preprocess_tile_rows(ay, t.new_h);
    preprocess2_extents(ax, ay, t.new_w, t.new_h, &t.span_px128, &t.rows16);
    if ((size_t)t.tile_rows * 128 * 3 > 200 * 1024)
      return fail(h, B200SAM_ERR_INVALID, "frame height %d needs a %d-row tile: downscale factor not supported", H,
                  t.tile_rows);
    {
      float lut[768];
      preprocess_lut(lut);
      std::vector<int> bits(768);
      memcpy(bits.data(), lut, sizeof lut);
      int* d = nullptr;
      if (int r = upload_ints(h, bits, &d)) return r;
      t.d_lut = reinterpret_cast<float*>(d);
    }
    if (int r = upload_ints(h, ax.lo, &t.d_xlo)) return r;
    if (int r = upload_ints(h, ax.cnt, &t.d_xcnt)) return r;
    if (int r = upload_ints(h, ax.kk, &t.d_xkk)) return r;
    if (int r = upload_ints(h, ay.lo, &t.d_ylo)) return r;
    if (int r = upload_ints(h, ay.cnt, &t.d_ycnt)) return r;
    if (int r = upload_ints(h, ay.kk, &t.d_ykk)) return r;
    it = h->resample.emplace(std::make_pair(H, W), t).first;
  }
  *out = &it->second;
  return 0;
}

}  // namespace

extern "C" int b200sam_preprocess(b200sam_t* h, const uint8_t* frames_rgb, int B, int H, int W, void* out,
                                  int out_dtype, int* host_resize_hw, void* stream) {
  NvtxRange nvtx_("b200sam_preprocess");
  if (!h) return B200SAM_ERR_INVALID;
  if (B < 0 || H < 1 || W < 1 || H > 16384 || W > 16384)
    return fail(h, B200SAM_ERR_INVALID, "bad preprocess sizes B=%d H=%d W=%d", B, H, W);
  if (out_dtype != B200SAM_F32 && out_dtype != B200SAM_BF16 && out_dtype != B200SAM_F16)
    return fail(h, B200SAM_ERR_INVALID, "bad output dtype %d", out_dtype);
  DeviceGuard g(h->device);
  const ResampleTables* t = nullptr;
  if (int r = resample_tables(h, H, W, &t)) return r;
  if (host_resize_hw) host_resize_hw[0] = t->new_h, host_resize_hw[1] = t->new_w;
  if (B == 0) return 0;
  if (!frames_rgb || !out) return fail(h, B200SAM_ERR_INVALID, "frames / out is NULL");
  if (reinterpret_cast<uintptr_t>(out) & 15) return fail(h, B200SAM_ERR_INVALID, "out must be 16-byte aligned");
  h->prof_stream = reinterpret_cast<cudaStream_t>(stream);
  h->pc = B200SAM_PC_PREPROCESS;
  KL(launch_preprocess(frames_rgb, B, H, W, *t, out, out_dtype, reinterpret_cast<cudaStream_t>(stream)));
  return 0;
}

extern "C" int b200sam_seg_project(b200sam_t* h, const void* hidden, int hidden_dtype, int b, int L, int hdim,
                                   const int64_t* output_ids, int T, int64_t seg_token_idx, const void* w1,
                                   const void* b1, const void* w2, const void* b2, int w_dtype, float* emb_out,
                                   int max_rows, int* counts_out, void* stream) {
  NvtxRange nvtx_("b200sam_seg_project");
  if (!h) return B200SAM_ERR_INVALID;
  if (b < 1 || b > 1024 || T < 1 || hdim < 8 || hdim % 8 != 0 || max_rows < 1)
    return fail(h, B200SAM_ERR_INVALID, "bad seg_project sizes b=%d T=%d hdim=%d max_rows=%d", b, T, hdim, max_rows);
  if (L < 255 + T - 1)   // LISA.py:461-466: the hidden states carry 255 more positions than output_ids[:, 1:]
    return fail(h, B200SAM_ERR_INVALID, "hidden length %d < 255 + T - 1 = %d", L, 255 + T - 1);
  if ((size_t)hdim * 8 * sizeof(float) > 227 * 1024)
    return fail(h, B200SAM_ERR_INVALID, "hidden size %d does not fit the 8-row shared-memory stage", hdim);
  if (hidden_dtype < 0 || hidden_dtype > 2 || (w_dtype != B200SAM_F32 && w_dtype != B200SAM_BF16))
    return fail(h, B200SAM_ERR_INVALID, "bad dtypes hidden=%d weights=%d", hidden_dtype, w_dtype);
  if (!hidden || !output_ids || !w1 || !b1 || !w2 || !b2 || !emb_out || !counts_out)
    return fail(h, B200SAM_ERR_INVALID, "NULL argument");
  DeviceGuard g(h->device);
  if (!h->seg_counts) CU(dalloc(h, &h->seg_counts, 1025));
  if (max_rows > h->seg_cap_rows || hdim > h->seg_cap_h) {   // first use / larger request: grow (not on the hot path)
    const int rows = max_rows > h->seg_cap_rows ? max_rows : h->seg_cap_rows;
    const int hd = hdim > h->seg_cap_h ? hdim : h->seg_cap_h;
    CU(dalloc(h, &h->seg_rows, rows));
    CU(dalloc(h, &h->seg_hidden, (size_t)rows * hd));
    h->seg_cap_rows = rows;
    h->seg_cap_h = hd;
  }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  h->prof_stream = s;
  h->pc = B200SAM_PC_SEG_PROJECT;
  KL(launch_seg_scan(output_ids, b, T, L, seg_token_idx, h->seg_rows, max_rows, h->seg_counts, s));
  KL(launch_rows_fc(hidden, hidden_dtype, hdim, h->seg_rows, h->seg_counts + b, max_rows, w1, b1, w_dtype, hdim, hdim,
                    true, h->seg_hidden, h->n_sm, s));
  KL(launch_rows_fc(h->seg_hidden, DT_F32, hdim, nullptr, h->seg_counts + b, max_rows, w2, b2, w_dtype, hdim, kOutCh,
                    false, emb_out, h->n_sm, s));
  CU(cudaMemcpyAsync(counts_out, h->seg_counts, sizeof(int) * (b + 1), cudaMemcpyDeviceToDevice, s));
  return 0;
}

extern "C" int b200sam_mask_metrics(b200sam_t* h, const uint8_t* masks_left, const uint8_t* masks_right,
                                    const float* taxonomy, const uint8_t* gt, const int* gt_index, int P, int H,
                                    int W, int64_t* counts_out, uint8_t* pred_out, void* stream) {
  NvtxRange nvtx_("b200sam_mask_metrics");
  if (!h) return B200SAM_ERR_INVALID;
  if (P < 0 || P > 65535 || H < 1 || W < 1)
    return fail(h, B200SAM_ERR_INVALID, "bad mask_metrics sizes P=%d H=%d W=%d", P, H, W);
  if (P == 0) return 0;
  if (!masks_left || !masks_right || !taxonomy || !gt || !counts_out) return fail(h, B200SAM_ERR_INVALID, "NULL argument");
  DeviceGuard g(h->device);
  h->prof_stream = reinterpret_cast<cudaStream_t>(stream);
  h->pc = B200SAM_PC_METRICS;
  KL(launch_mask_metrics(masks_left, masks_right, taxonomy, gt, gt_index, P, H, W, counts_out, pred_out, h->n_sm,
                         reinterpret_cast<cudaStream_t>(stream)));
  return 0;
}

// ------------------------------------------------------------------------------------------------ operator entry points

extern "C" int b200sam_op_gemm(int device, int tf32, int mode, const void* A, int lda, const void* Wt, int ldw, int M,
                               int N, int K, const float* bias, const float* addend, int ld_add, int add_mod,
                               const float* ln_g, const float* ln_b, float ln_eps, void* out, int ld_out, int rowmap,
                               int debug_simt, void* stream) {
  DeviceGuard g(device);
  if (gemm_tc_init(device) != cudaSuccess) return B200SAM_ERR_CUDA;
  GemmEpi e;
  e.bias = bias;
  e.addend = addend;
  e.ld_add = ld_add;
  e.add_mod = add_mod;
  e.out = out;
  e.ld_out = ld_out;
  e.rowmap = rowmap;
  e.ln_g = ln_g;
  e.ln_b = ln_b;
  e.ln_eps = ln_eps;
  if (mode < GM_BIAS_BF16 || mode > GM_LN) return B200SAM_ERR_INVALID;
  cudaError_t err = gemm_tc(tf32 != 0, mode, A, lda, Wt, ldw, M, N, K, e, reinterpret_cast<cudaStream_t>(stream),
                            debug_simt != 0);
  if (err == cudaErrorInvalidValue) return B200SAM_ERR_INVALID;
  return err == cudaSuccess ? 0 : B200SAM_ERR_CUDA;
}

extern "C" int b200sam_op_gemm_fold_producer(int device, const void* A, int lda, const void* Wt, int ldw, int M, int N,
                                             int K, const float* bias, const float* addend, int ld_add, int add_mod,
                                             float* out, int ld_out, void* xb_out, float* stats_out, void* stream) {
  DeviceGuard g(device);
  if (gemm_tc_init(device) != cudaSuccess) return B200SAM_ERR_CUDA;
  GemmEpi e;
  e.bias = bias;
  e.addend = addend;
  e.ld_add = ld_add;
  e.add_mod = add_mod;
  e.out = out;
  e.ld_out = ld_out;
  e.xb_out = xb_out;
  e.stats_out = stats_out;
  e.stats_slots = 2 * ((N + 255) / 256);
  cudaError_t err = gemm_tc(false, GM_F32_FOLD, A, lda, Wt, ldw, M, N, K, e, reinterpret_cast<cudaStream_t>(stream), false);
  if (err == cudaErrorInvalidValue) return B200SAM_ERR_INVALID;
  return err == cudaSuccess ? 0 : B200SAM_ERR_CUDA;
}

extern "C" int b200sam_op_gemm_fold_consumer(int device, int gelu, const void* xb, int lda, const float* W,
                                             const float* gamma, const float* beta, const float* bias, float eps,
                                             const float* stats, int M, int N, int K, void* wf_scratch,
                                             float* cd_scratch, void* out, int ld_out, void* stream) {
  DeviceGuard g(device);
  if (gemm_tc_init(device) != cudaSuccess) return B200SAM_ERR_CUDA;
  if (!xb || !W || !gamma || !beta || !bias || !stats || !wf_scratch || !cd_scratch || !out || M < 1 || N < 1 || K < 1)
    return B200SAM_ERR_INVALID;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  __nv_bfloat16* wf = reinterpret_cast<__nv_bfloat16*>(wf_scratch);
  if (launch_fold_ln_weights(W, gamma, beta, bias, N, K, wf, cd_scratch, cd_scratch + N, s) != cudaSuccess)
    return B200SAM_ERR_CUDA;
  GemmEpi e;
  e.bias = cd_scratch + N;       // d
  e.fold_c = cd_scratch;         // c
  e.fold_eps = eps;
  e.stats_in = stats;
  e.stats_slots = 2 * ((K + 255) / 256);
  e.stats_dim = K;
  e.out = out;
  e.ld_out = ld_out;
  cudaError_t err = gemm_tc(false, gelu ? GM_BIAS_GELU_BF16 : GM_BIAS_BF16, xb, lda, wf, K, M, N, K, e, s, false);
  if (err == cudaErrorInvalidValue) return B200SAM_ERR_INVALID;
  return err == cudaSuccess ? 0 : B200SAM_ERR_CUDA;
}

extern "C" int b200sam_op_layernorm(const float* x, int rows, int D, const float* g, const float* b, float eps,
                                    void* out_bf16, int windowed, void* stream) {
  cudaError_t err = launch_layernorm_bf16(x, rows, D, g, b, eps, reinterpret_cast<__nv_bfloat16*>(out_bf16),
                                          windowed != 0, reinterpret_cast<cudaStream_t>(stream));
  if (err == cudaErrorInvalidValue) return B200SAM_ERR_INVALID;
  return err == cudaSuccess ? 0 : B200SAM_ERR_CUDA;
}

extern "C" int b200sam_op_attention(const void* qkv_bf16, int nbatch, int S, int heads, int hd, const float* rel_h,
                                    const float* rel_w, void* out_bf16, void* stream) {
  if ((S != kWin && S != kGrid) || (hd != 64 && hd != 80) || heads <= 0 || nbatch < 0) return B200SAM_ERR_INVALID;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || gemm_tc_init(dev) != cudaSuccess) return B200SAM_ERR_CUDA;
  const int RP = S == kWin ? 32 : 128;
  __nv_bfloat16* tabs = nullptr;
  if (cudaMalloc(reinterpret_cast<void**>(&tabs), (size_t)2 * RP * hd * 2) != cudaSuccess) return B200SAM_ERR_CUDA;
  __nv_bfloat16* relg = nullptr;
  void* gbuf = nullptr;
  cudaError_t err = launch_relpos_pack(rel_h, S, hd, tabs, s);
  if (err == cudaSuccess) err = launch_relpos_pack(rel_w, S, hd, tabs + RP * hd, s);
  if (err == cudaSuccess && S == kGrid) {
    err = cudaMalloc(reinterpret_cast<void**>(&relg), 256 * 128 * 2);
    if (err == cudaSuccess) err = cudaMalloc(&gbuf, (size_t)nbatch * kTokens * heads * 256 * 2 + 256);
    if (err == cudaSuccess) err = launch_relpos_pack_g(rel_h, rel_w, hd, relg, s);
  }
  if (err == cudaSuccess)
    err = launch_attention(reinterpret_cast<const __nv_bfloat16*>(qkv_bf16), nbatch, S, heads, hd, tabs, tabs + RP * hd,
                           relg, gbuf, reinterpret_cast<__nv_bfloat16*>(out_bf16), s);
  if (err == cudaSuccess) err = cudaStreamSynchronize(s);
  cudaFree(tabs);
  cudaFree(relg);
  cudaFree(gbuf);
  if (err != cudaSuccess) fprintf(stderr, "b200sam_op_attention: %s\n", cudaGetErrorString(err));
  return err == cudaSuccess ? 0 : B200SAM_ERR_CUDA;
}

// ------------------------------------------------------------------------------------------------ measurement hooks

extern "C" int b200sam_debug_set_variant(int which, int value) {
  if (which < 0 || which >= VK_COUNT) return B200SAM_ERR_INVALID;
  set_kernel_variant(which, value);
  return 0;
}

extern "C" int b200sam_debug_set_trace(void* dev_buffer) {
  set_debug_trace_buffer(dev_buffer);
  return 0;
}

extern "C" int b200sam_debug_attention_bench(const void* qkv_bf16, int nbatch, int S, int heads, int hd,
                                             const float* rel_h, const float* rel_w, void* out_bf16, int iters,
                                             float* host_ms_out, void* stream) {
  if ((S != kWin && S != kGrid) || (hd != 64 && hd != 80) || heads <= 0 || nbatch <= 0 || iters <= 0 || !host_ms_out)
    return B200SAM_ERR_INVALID;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || gemm_tc_init(dev) != cudaSuccess) return B200SAM_ERR_CUDA;
  const int RP = S == kWin ? 32 : 128;
  __nv_bfloat16 *tabs = nullptr, *relg = nullptr;
  void* gbuf = nullptr;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  cudaError_t err = cudaMalloc(reinterpret_cast<void**>(&tabs), (size_t)2 * RP * hd * 2);
  if (err == cudaSuccess) err = launch_relpos_pack(rel_h, S, hd, tabs, s);
  if (err == cudaSuccess) err = launch_relpos_pack(rel_w, S, hd, tabs + RP * hd, s);
  if (err == cudaSuccess && S == kGrid) {
    err = cudaMalloc(reinterpret_cast<void**>(&relg), 256 * 128 * 2);
    if (err == cudaSuccess) err = cudaMalloc(&gbuf, (size_t)nbatch * kTokens * heads * 256 * 2 + 256);
    if (err == cudaSuccess) err = launch_relpos_pack_g(rel_h, rel_w, hd, relg, s);
  }
  if (err == cudaSuccess) err = cudaEventCreate(&e0);
  if (err == cudaSuccess) err = cudaEventCreate(&e1);
  for (int it = -3; it < iters && err == cudaSuccess; ++it) {   // 3 warm-up launches
    if (it == 0) err = cudaEventRecord(e0, s);
    if (err == cudaSuccess)
      err = launch_attention(reinterpret_cast<const __nv_bfloat16*>(qkv_bf16), nbatch, S, heads, hd, tabs,
                             tabs + RP * hd, relg, gbuf, reinterpret_cast<__nv_bfloat16*>(out_bf16), s);
  }
  if (err == cudaSuccess) err = cudaEventRecord(e1, s);
  if (err == cudaSuccess) err = cudaEventSynchronize(e1);
  float ms = 0.f;
  if (err == cudaSuccess) err = cudaEventElapsedTime(&ms, e0, e1);
  *host_ms_out = ms / iters;
  if (e0) cudaEventDestroy(e0);
  if (e1) cudaEventDestroy(e1);
  cudaFree(tabs);
  cudaFree(relg);
  cudaFree(gbuf);
  if (err != cudaSuccess) fprintf(stderr, "b200sam_debug_attention_bench: %s\n", cudaGetErrorString(err));
  return err == cudaSuccess ? 0 : B200SAM_ERR_CUDA;
}
