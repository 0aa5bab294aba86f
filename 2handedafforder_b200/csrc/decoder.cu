// Token-side and attention kernels of the two-way mask decoder (fp32 SIMT; the image-side projections run on the
// tcgen05 TF32 GEMM).  Reference: TwoWayTransformer / TwoWayAttentionBlock / Attention (transformer.py:62-242),
// MaskDecoder.predict_masks / predict_taxonomy / MLP (mask_decoder.py:122-205).
//
// One CTA per prompt keeps the 6 tokens (iou token, 4 mask tokens, the projected [SEG] embedding) in shared
// memory; weights stream from L2 with 16-byte loads, one warp per output feature and warp-shuffle reductions.
#include "decoder.cuh"

#include <cooperative_groups.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace bsam {

namespace {

constexpr int kTok = 6;
constexpr int kC = 256;
constexpr int kThreadsTok = 256;

// y[r][o] = act(sum_i x[r][i] * W[o][i] + b[o]) for r < R; x, y in shared memory; IN % 128 == 0.
template <int R>
__device__ void block_linear(const float* __restrict__ xs, int IN, const float* __restrict__ W,
                             const float* __restrict__ b, int OUT, float* __restrict__ ys, int ldy, bool relu) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  for (int o = warp; o < OUT; o += nw) {
    const float4* wr = reinterpret_cast<const float4*>(W + (long)o * IN);
    float acc[R];
#pragma unroll
    for (int r = 0; r < R; ++r) acc[r] = 0.f;
    for (int i = lane; i < IN / 4; i += 32) {
      const float4 wv = __ldg(wr + i);
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const float4 xv = *reinterpret_cast<const float4*>(xs + r * IN + 4 * i);
        acc[r] += wv.x * xv.x + wv.y * xv.y + wv.z * xv.z + wv.w * xv.w;
      }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) acc[r] = warp_sum(acc[r]);
    if (lane == 0) {
      const float bb = b ? __ldg(b + o) : 0.f;
#pragma unroll
      for (int r = 0; r < R; ++r) {
        float v = acc[r] + bb;
        ys[r * ldy + o] = relu ? fmaxf(v, 0.f) : v;
      }
    }
  }
}

// rows of 256: y = LayerNorm(x (+ res)) * g + b, eps 1e-5 (nn.LayerNorm default, transformer.py:134-144). In place on xs.
__device__ void block_add_ln(float* xs, const float* res, const float* __restrict__ g, const float* __restrict__ b,
                             int rows) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  for (int r = warp; r < rows; r += nw) {
    float v[8];
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      v[j] = xs[r * kC + lane + 32 * j] + (res ? res[r * kC + lane + 32 * j] : 0.f);
      s += v[j];
    }
    const float mean = warp_sum(s) * (1.f / kC);
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) q += (v[j] - mean) * (v[j] - mean);
    const float rstd = rsqrtf(warp_sum(q) * (1.f / kC) + 1e-5f);
#pragma unroll
    for (int j = 0; j < 8; ++j)
      xs[r * kC + lane + 32 * j] = (v[j] - mean) * rstd * __ldg(g + lane + 32 * j) + __ldg(b + lane + 32 * j);
  }
}

// ------------------------------------------------------------------------------------------------ cluster helpers
// The token side is 6 rows per prompt: far too little work for one CTA per prompt to hide the latency of streaming
// ~6 MB of fp32 weights.  Each prompt is therefore handled by a thread-block cluster of 8 CTAs: every CTA keeps the
// full token state in its own shared memory, computes 1/8 of the output features of each Linear and all-gathers
// its slice into the shared memory of all 8 CTAs through distributed shared memory; cheap steps (6x6 attention,
// LayerNorm) are recomputed redundantly instead of synchronised.
constexpr int kCl = 8;

// Features [rank*OUT/8, (rank+1)*OUT/8) of y = act(x W^T + b) for R rows; x in local smem; the slice is written to
// `ys` (same smem offset) in every CTA of the cluster.  Caller must cluster.sync() before anyone reads ys.
// Each warp works on kLinU output features at a time: kLinU independent 16-byte weight loads are in flight per lane
// (the loop used to expose one L2 latency per feature) and every activation vector read from shared memory is used
// for kLinU dot products.  The sum of every single feature is formed in the same order as before.
constexpr int kLinU = 4;

template <int R, class Store>
__device__ __forceinline__ void cl_linear_core(const float* __restrict__ xs, int IN, const float* __restrict__ W, int o_lo,
                                               int o_hi, const Store& store) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  for (int o0 = o_lo + warp; o0 < o_hi; o0 += nw * kLinU) {
    const float4* wr[kLinU];
    float acc[kLinU][R];
#pragma unroll
    for (int u = 0; u < kLinU; ++u) {
      const int o = o0 + u * nw;
      wr[u] = reinterpret_cast<const float4*>(W + (long)(o < o_hi ? o : o0) * IN);
#pragma unroll
      for (int r = 0; r < R; ++r) acc[u][r] = 0.f;
    }
    for (int i = lane; i < IN / 4; i += 32) {
      float4 wv[kLinU];
#pragma unroll
      for (int u = 0; u < kLinU; ++u) wv[u] = __ldg(wr[u] + i);
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const float4 xv = *reinterpret_cast<const float4*>(xs + r * IN + 4 * i);
#pragma unroll
        for (int u = 0; u < kLinU; ++u) acc[u][r] += wv[u].x * xv.x + wv[u].y * xv.y + wv[u].z * xv.z + wv[u].w * xv.w;
      }
    }
#pragma unroll
    for (int u = 0; u < kLinU; ++u) {
      const int o = o0 + u * nw;
      if (o < o_hi) {                      // warp-uniform
#pragma unroll
        for (int r = 0; r < R; ++r) acc[u][r] = warp_sum(acc[u][r]);
        store(o, acc[u]);
      }
    }
  }
}

template <int R>
__device__ void cl_linear(cg::cluster_group& cl, const float* __restrict__ xs, int IN, const float* __restrict__ W,
                          const float* __restrict__ b, int OUT, float* ys, int ldy, bool relu) {
  const int lane = threadIdx.x & 31;
  const int per = OUT / kCl, o_lo = (int)cl.block_rank() * per;
  float* dst = cl.map_shared_rank(ys, lane & (kCl - 1));
  cl_linear_core<R>(xs, IN, W, o_lo, o_lo + per, [&](int o, const float (&acc)[R]) {
    if (lane < kCl) {
      const float bb = b ? __ldg(b + o) : 0.f;
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const float v = acc[r] + bb;
        dst[r * ldy + o] = relu ? fmaxf(v, 0.f) : v;
      }
    }
  });
}

// Same split, result straight to global memory (row stride ldy).
template <int R>
__device__ void cl_linear_global(cg::cluster_group& cl, const float* __restrict__ xs, int IN,
                                 const float* __restrict__ W, const float* __restrict__ b, int OUT,
                                 float* __restrict__ yg, int ldy) {
  const int lane = threadIdx.x & 31;
  const int per = (OUT + kCl - 1) / kCl, o_lo = (int)cl.block_rank() * per;
  const int o_hi = min(OUT, o_lo + per);
  cl_linear_core<R>(xs, IN, W, o_lo, o_hi, [&](int o, const float (&acc)[R]) {
    if (lane == 0) {
      const float bb = b ? __ldg(b + o) : 0.f;
#pragma unroll
      for (int r = 0; r < R; ++r) yg[r * ldy + o] = acc[r] + bb;
    }
  });
}

// ------------------------------------------------------------------------------------------------ stage A
// Builds the tokens (layer 0), token self-attention + norm1, and the token->image query projection.
// transformer.py:153-167 (layer 0 REPLACES the queries and uses no PE; later layers add), mask_decoder.py:131-138.
__global__ void __cluster_dims__(kCl, 1, 1) __launch_bounds__(kThreadsTok)
dec_stage_a_kernel(TokStageA a) {
  cg::cluster_group cl = cg::this_cluster();
  extern __shared__ __align__(16) float sm[];
  float* sQ = sm;                  // (6,256) queries
  float* sT = sQ + kTok * kC;      // (6,256) token PE (initial tokens)
  float* sX = sT + kTok * kC;      // (6,256) q = k input
  float* sq = sX + kTok * kC;      // (6,256) projected q
  float* sk = sq + kTok * kC;
  float* sv = sk + kTok * kC;
  float* sA = sv + kTok * kC;      // (6,256) attention output
  float* sY = sA + kTok * kC;      // (6,256) out_proj result
  const int p = blockIdx.x / kCl, tid = threadIdx.x;
  const bool lead = cl.block_rank() == 0;
  float* gQ = a.tok_state + (long)p * kTok * kC;
  float* gT = a.tok_pe + (long)p * kTok * kC;

  if (a.layer == 0) {
    for (int i = tid; i < kTok * kC; i += kThreadsTok) {
      const int r = i / kC, c = i % kC;
      float v = r == 0 ? a.iou_token[c] : (r < 5 ? a.mask_tokens[(r - 1) * kC + c] : a.seg[(long)p * kC + c]);
      sT[i] = v;
      sQ[i] = v;
      sX[i] = v;
      if (lead) gT[i] = v;
    }
  } else {
    for (int i = tid; i < kTok * kC; i += kThreadsTok) {
      sT[i] = gT[i];
      sQ[i] = gQ[i];
      sX[i] = sQ[i] + sT[i];
    }
  }
  cl.sync();                       // every CTA of the cluster is running; local tiles are complete
  cl_linear<kTok>(cl, sX, kC, a.sa.qw, a.sa.qb, kC, sq, kC, false);
  cl_linear<kTok>(cl, sX, kC, a.sa.kw, a.sa.kb, kC, sk, kC, false);
  cl_linear<kTok>(cl, sQ, kC, a.sa.vw, a.sa.vb, kC, sv, kC, false);   // v = queries (no PE), transformer.py:160
  cl.sync();
  // 8 heads x 32, 6x6 scores: one thread per (head, query); recomputed by every CTA
  if (tid < 8 * kTok) {
    const int h = tid / kTok, qi = tid % kTok;
    float s[kTok], mx = -INFINITY;
    for (int j = 0; j < kTok; ++j) {
      float d = 0.f;
      for (int c = 0; c < 32; ++c) d += sq[qi * kC + h * 32 + c] * sk[j * kC + h * 32 + c];
      s[j] = d * 0.17677669529663687f;   // 1/sqrt(32)
      mx = fmaxf(mx, s[j]);
    }
    float l = 0.f;
    for (int j = 0; j < kTok; ++j) {
      s[j] = __expf(s[j] - mx);
      l += s[j];
    }
    for (int c = 0; c < 32; ++c) {
      float o = 0.f;
      for (int j = 0; j < kTok; ++j) o += s[j] * sv[j * kC + h * 32 + c];
      sA[qi * kC + h * 32 + c] = o / l;
    }
  }
  __syncthreads();
  cl_linear<kTok>(cl, sA, kC, a.sa.ow, a.sa.ob, kC, sY, kC, false);
  cl.sync();
  block_add_ln(sY, a.layer == 0 ? nullptr : sQ, a.n1g, a.n1b, kTok);   // queries = norm1(attn_out (+ queries))
  __syncthreads();
  for (int i = tid; i < kTok * kC; i += kThreadsTok) {
    if (lead) gQ[i] = sY[i];
    sX[i] = sY[i] + sT[i];                                             // q = queries + query_pe
  }
  __syncthreads();
  cl_linear_global<kTok>(cl, sX, kC, a.t2i_qw, a.t2i_qb, 128, a.tok_q + (long)p * kTok * 128, 128);
}

// Only the q projection (final_attn_token_to_image, transformer.py:100-102).
__global__ void __launch_bounds__(kThreadsTok)
dec_stage_q_kernel(const float* __restrict__ tok_state, const float* __restrict__ tok_pe, const float* qw,
                   const float* qb, float* __restrict__ tok_q) {
  __shared__ __align__(16) float sQ[kTok * kC];
  const int p = blockIdx.x;
  for (int i = threadIdx.x; i < kTok * kC; i += kThreadsTok)
    sQ[i] = tok_state[(long)p * kTok * kC + i] + tok_pe[(long)p * kTok * kC + i];
  __syncthreads();
  block_linear<kTok>(sQ, kC, qw, qb, 128, tok_q + (long)p * kTok * 128, 128, false);
}

// ------------------------------------------------------------------------------------------------ token -> image attention
// 6 queries x 4096 keys, 8 heads x 16 (cross attention, downsample_rate 2; transformer.py:136-147,199,220-242).
// grid (P, 8): one CTA per (prompt, head); keys/values stream once from the [k | v] buffer; online softmax per
// thread, merged with warp shuffles and a final shared-memory pass.
__global__ void __launch_bounds__(256)
dec_t2i_attn_kernel(const float* __restrict__ tok_q, const float* __restrict__ kv, float* __restrict__ tok_o) {
  const int p = blockIdx.x, h = blockIdx.y, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  __shared__ float sq[kTok][16];
  __shared__ float red[8][kTok][18];
  if (tid < kTok * 16) sq[tid / 16][tid % 16] = tok_q[((long)p * kTok + tid / 16) * 128 + h * 16 + tid % 16] * 0.25f;
  __syncthreads();
  float m[kTok], l[kTok], acc[kTok][16];
#pragma unroll
  for (int i = 0; i < kTok; ++i) {
    m[i] = -INFINITY;
    l[i] = 0.f;
#pragma unroll
    for (int d = 0; d < 16; ++d) acc[i][d] = 0.f;
  }
  const float* base = kv + (long)p * kTokens * 256 + h * 16;
  for (int j = tid; j < kTokens; j += 256) {
    float k[16], v[16];
    const float4* kp = reinterpret_cast<const float4*>(base + (long)j * 256);
    const float4* vp = reinterpret_cast<const float4*>(base + (long)j * 256 + 128);
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      float4 a = __ldg(kp + c), b = __ldg(vp + c);
      k[4 * c] = a.x; k[4 * c + 1] = a.y; k[4 * c + 2] = a.z; k[4 * c + 3] = a.w;
      v[4 * c] = b.x; v[4 * c + 1] = b.y; v[4 * c + 2] = b.z; v[4 * c + 3] = b.w;
    }
#pragma unroll
    for (int i = 0; i < kTok; ++i) {
      float s = 0.f;
#pragma unroll
      for (int d = 0; d < 16; ++d) s += sq[i][d] * k[d];
      const float mn = fmaxf(m[i], s);
      const float al = __expf(m[i] - mn), pe = __expf(s - mn);
      m[i] = mn;
      l[i] = l[i] * al + pe;
#pragma unroll
      for (int d = 0; d < 16; ++d) acc[i][d] = acc[i][d] * al + pe * v[d];
    }
  }
#pragma unroll
  for (int i = 0; i < kTok; ++i) {
    const float mw = warp_max(m[i]);
    const float sc = __expf(m[i] - mw);
    const float lw = warp_sum(l[i] * sc);
    if (lane == 0) {
      red[warp][i][16] = mw;
      red[warp][i][17] = lw;
    }
#pragma unroll
    for (int d = 0; d < 16; ++d) {
      const float a = warp_sum(acc[i][d] * sc);
      if (lane == 0) red[warp][i][d] = a;
    }
  }
  __syncthreads();
  if (tid < kTok * 16) {
    const int i = tid / 16, d = tid % 16;
    float mx = -INFINITY;
    for (int w = 0; w < 8; ++w) mx = fmaxf(mx, red[w][i][16]);
    float L = 0.f, A = 0.f;
    for (int w = 0; w < 8; ++w) {
      const float sc = __expf(red[w][i][16] - mx);
      L += red[w][i][17] * sc;
      A += red[w][i][d] * sc;
    }
    tok_o[((long)p * kTok + i) * 128 + h * 16 + d] = A / L;
  }
}

// ------------------------------------------------------------------------------------------------ stage B
// out_proj of the token->image attention + norm2, MLP (ReLU) + norm3, and the token k/v projections for the
// image->token attention (transformer.py:163-181).  Cluster of 8 CTAs per prompt (see cl_linear).
__global__ void __cluster_dims__(kCl, 1, 1) __launch_bounds__(kThreadsTok)
dec_stage_b_kernel(TokStageB a) {
  cg::cluster_group cl = cg::this_cluster();
  extern __shared__ __align__(16) float sm[];
  float* sQ = sm;                    // (6,256) queries, later queries + pe
  float* sT = sQ + kTok * kC;        // (6,256) token pe
  float* sO = sT + kTok * kC;        // (6,128) attention output
  float* sX = sO + kTok * 128;       // (6,256) out_proj -> norm2 output
  float* sY = sX + kTok * kC;        // (6,256) mlp output -> norm3 output
  float* sH = sY + kTok * kC;        // (6,2048) MLP hidden
  const int p = blockIdx.x / kCl, tid = threadIdx.x;
  const bool lead = cl.block_rank() == 0;
  float* gQ = a.tok_state + (long)p * kTok * kC;
  for (int i = tid; i < kTok * kC; i += kThreadsTok) {
    sQ[i] = gQ[i];
    sT[i] = a.tok_pe[(long)p * kTok * kC + i];
  }
  for (int i = tid; i < kTok * 128; i += kThreadsTok) sO[i] = a.tok_o[(long)p * kTok * 128 + i];
  cl.sync();
  cl_linear<kTok>(cl, sO, 128, a.t2i_ow, a.t2i_ob, kC, sX, kC, false);
  cl.sync();
  block_add_ln(sX, sQ, a.n2g, a.n2b, kTok);          // queries = norm2(queries + attn_out)
  __syncthreads();
  cl_linear<kTok>(cl, sX, kC, a.mlp1w, a.mlp1b, 2048, sH, 2048, true);
  cl.sync();
  cl_linear<kTok>(cl, sH, 2048, a.mlp2w, a.mlp2b, kC, sY, kC, false);
  cl.sync();
  block_add_ln(sY, sX, a.n3g, a.n3b, kTok);          // queries = norm3(queries + mlp_out)
  __syncthreads();
  for (int i = tid; i < kTok * kC; i += kThreadsTok) {
    if (lead) gQ[i] = sY[i];
    sQ[i] = sY[i] + sT[i];                           // k input = queries + query_pe
  }
  __syncthreads();
  float* gkv = a.tok_kv + (long)p * kTok * 256;
  cl_linear_global<kTok>(cl, sQ, kC, a.i2t_kw, a.i2t_kb, 128, gkv, 256);
  cl_linear_global<kTok>(cl, sY, kC, a.i2t_vw, a.i2t_vb, 128, gkv + 128, 256);
}

// ------------------------------------------------------------------------------------------------ image -> token attention
// Every image token attends to the 6 tokens (transformer.py:174-180): thread per (token row, head).
__global__ void __launch_bounds__(256)
dec_i2t_attn_kernel(const float* __restrict__ qimg, const float* __restrict__ tok_kv, float* __restrict__ oimg) {
  const int p = blockIdx.y;
  __shared__ float skv[kTok][256];
  for (int i = threadIdx.x; i < kTok * 256; i += 256) skv[i / 256][i % 256] = tok_kv[(long)p * kTok * 256 + i];
  __syncthreads();
  const int h = threadIdx.x & 7;
  const long row = (long)p * kTokens + blockIdx.x * 32 + (threadIdx.x >> 3);
  const float4* qp = reinterpret_cast<const float4*>(qimg + row * 128 + h * 16);
  float q[16];
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    float4 a = __ldg(qp + c);
    q[4 * c] = a.x; q[4 * c + 1] = a.y; q[4 * c + 2] = a.z; q[4 * c + 3] = a.w;
  }
  float s[kTok], mx = -INFINITY;
#pragma unroll
  for (int j = 0; j < kTok; ++j) {
    float d = 0.f;
#pragma unroll
    for (int c = 0; c < 16; ++c) d += q[c] * skv[j][h * 16 + c];
    s[j] = d * 0.25f;
    mx = fmaxf(mx, s[j]);
  }
  float l = 0.f;
#pragma unroll
  for (int j = 0; j < kTok; ++j) {
    s[j] = __expf(s[j] - mx);
    l += s[j];
  }
  const float inv = 1.f / l;
  float4* op = reinterpret_cast<float4*>(oimg + row * 128 + h * 16);
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    float o[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      float acc = 0.f;
#pragma unroll
      for (int j = 0; j < kTok; ++j) acc += s[j] * skv[j][128 + h * 16 + 4 * c + e];
      o[e] = acc * inv;
    }
    op[c] = make_float4(o[0], o[1], o[2], o[3]);
  }
}

// ------------------------------------------------------------------------------------------------ heads
// Final attention out_proj + norm_final_attn, hypernetwork MLP of mask token 0, IoU head, taxonomy head
// (transformer.py:100-104, mask_decoder.py:150-178).  Cluster of 8 CTAs per prompt; the three MLPs advance layer by
// layer together so there is one cluster barrier per layer.
__global__ void __cluster_dims__(kCl, 1, 1) __launch_bounds__(kThreadsTok)
dec_heads_kernel(TokHeads a) {
  cg::cluster_group cl = cg::this_cluster();
  extern __shared__ __align__(16) float sm[];
  float* sQ = sm;                  // (6,256)
  float* sO = sQ + kTok * kC;      // (6,128)
  float* sX = sO + kTok * 128;     // (6,256)
  float* sA = sX + kTok * kC;      // 256 hyper | 256 iou | 1024 taxonomy
  float* sB = sA + 1536;           // same split
  float* sC = sB + 1536;           // 4 taxonomy logits
  const int p = blockIdx.x / kCl, tid = threadIdx.x;
  const bool tax = a.tax_w[0] != nullptr && a.tax_out != nullptr;
  for (int i = tid; i < kTok * kC; i += kThreadsTok) sQ[i] = a.tok_state[(long)p * kTok * kC + i];
  for (int i = tid; i < kTok * 128; i += kThreadsTok) sO[i] = a.tok_o[(long)p * kTok * 128 + i];
  cl.sync();
  cl_linear<kTok>(cl, sO, 128, a.ow, a.ob, kC, sX, kC, false);
  cl.sync();
  block_add_ln(sX, sQ, a.nfg, a.nfb, kTok);          // hs = norm_final_attn(queries + attn_out)
  __syncthreads();
  // layer 1: hypernetwork MLP 0 on mask token 0 (row 1), IoU head on the iou token (row 0), taxonomy on
  // mask_tokens_out.flatten(1) (rows 1..4 are contiguous = 1024)
  cl_linear<1>(cl, sX + 1 * kC, kC, a.hyp_w[0], a.hyp_b[0], 256, sA, 256, true);
  cl_linear<1>(cl, sX, kC, a.iou_w[0], a.iou_b[0], 256, sA + 256, 256, true);
  if (tax) cl_linear<1>(cl, sX + 1 * kC, 1024, a.tax_w[0], a.tax_b[0], 1024, sA + 512, 1024, true);
  cl.sync();
  cl_linear<1>(cl, sA, kC, a.hyp_w[1], a.hyp_b[1], 256, sB, 256, true);
  cl_linear<1>(cl, sA + 256, kC, a.iou_w[1], a.iou_b[1], 256, sB + 256, 256, true);
  if (tax) cl_linear<1>(cl, sA + 512, 1024, a.tax_w[1], a.tax_b[1], 1024, sB + 512, 1024, true);
  cl.sync();
  cl_linear_global<1>(cl, sB, kC, a.hyp_w[2], a.hyp_b[2], 32, a.hyper + (long)p * 32, 32);
  if (a.iou_out) cl_linear_global<1>(cl, sB + 256, kC, a.iou_w[2], a.iou_b[2], 4, a.iou_out + (long)p * 4, 4);
  if (tax && cl.block_rank() == kCl - 1) {
    block_linear<1>(sB + 512, 1024, a.tax_w[2], a.tax_b[2], 4, sC, 4, false);
    __syncthreads();
    if (tid == 0) {
      float mx = fmaxf(fmaxf(sC[0], sC[1]), fmaxf(sC[2], sC[3]));
      float e0 = expf(sC[0] - mx), e1 = expf(sC[1] - mx), e2 = expf(sC[2] - mx), e3 = expf(sC[3] - mx);
      float inv = 1.f / (e0 + e1 + e2 + e3);
      float* o = a.tax_out + (long)p * 4;
      o[0] = e0 * inv; o[1] = e1 * inv; o[2] = e2 * inv; o[3] = e3 * inv;
    }
  }
}

constexpr int kSmemA = 8 * kTok * kC * 4;
constexpr int kSmemB = (4 * kTok * kC + kTok * 128 + kTok * 2048) * 4;
constexpr int kSmemH = (2 * kTok * kC + kTok * 128 + 2 * 1536 + 16) * 4;

}  // namespace

cudaError_t decoder_init() {
  cudaError_t e = cudaFuncSetAttribute(dec_stage_a_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemA);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(dec_stage_b_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemB);
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(dec_heads_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemH);
}

cudaError_t launch_dec_stage_a(const TokStageA& a, int P, cudaStream_t s) {
  if (P <= 0) return cudaSuccess;
  dec_stage_a_kernel<<<P * kCl, kThreadsTok, kSmemA, s>>>(a);
  return cudaGetLastError();
}
cudaError_t launch_dec_stage_q(const float* tok_state, const float* tok_pe, const float* qw, const float* qb,
                               float* tok_q, int P, cudaStream_t s) {
  if (P <= 0) return cudaSuccess;
  dec_stage_q_kernel<<<P, kThreadsTok, 0, s>>>(tok_state, tok_pe, qw, qb, tok_q);
  return cudaGetLastError();
}
cudaError_t launch_dec_t2i(const float* tok_q, const float* kv, float* tok_o, int P, cudaStream_t s) {
  if (P <= 0) return cudaSuccess;
  dec_t2i_attn_kernel<<<dim3(P, 8), 256, 0, s>>>(tok_q, kv, tok_o);
  return cudaGetLastError();
}
cudaError_t launch_dec_stage_b(const TokStageB& a, int P, cudaStream_t s) {
  if (P <= 0) return cudaSuccess;
  dec_stage_b_kernel<<<P * kCl, kThreadsTok, kSmemB, s>>>(a);
  return cudaGetLastError();
}
cudaError_t launch_dec_i2t(const float* qimg, const float* tok_kv, float* oimg, int P, cudaStream_t s) {
  if (P <= 0) return cudaSuccess;
  dec_i2t_attn_kernel<<<dim3(kTokens / 32, P), 256, 0, s>>>(qimg, tok_kv, oimg);
  return cudaGetLastError();
}
cudaError_t launch_dec_heads(const TokHeads& a, int P, cudaStream_t s) {
  if (P <= 0) return cudaSuccess;
  dec_heads_kernel<<<P * kCl, kThreadsTok, kSmemH, s>>>(a);
  return cudaGetLastError();
}

}  // namespace bsam
