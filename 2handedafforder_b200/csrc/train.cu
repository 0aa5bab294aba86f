// Training forward / backward of the two mask decoders (LISA.py:284-430, losses :16-59).
//
// The reference trains mask_decoder_left / mask_decoder_right (LISA.py:84-90) through PyTorch autograd.  Here every
// arithmetic step of that graph, forward and backward, is one of the fp32 kernels below; 2handedafforder_b200/train.py
// composes them in the order of mask_decoder.py / transformer.py and lets torch.autograd do nothing but remember the
// order (each primitive is an autograd.Function whose forward and backward are calls into this file).
//
//   gemm_strided_kernel   C = alpha * A B (+ beta C) (+ bias) (ReLU), two batch dimensions, arbitrary element strides:
//                         every Linear (forward, input gradient, weight gradient), the per-head QK^T / PV products and
//                         their four adjoints, the hypernetwork x upscaled-embedding product and the two ConvTranspose2d
//                         (k = s = 2 => a GEMM followed by a pixel shuffle).  Split-K with fp32 atomics for the weight
//                         gradients (64 K-deep reductions over P * 4096 rows into a 256 x 256 result).
//   layernorm_*           LayerNorm / LayerNorm2d (channels-last rows), saves mean and rstd
//   softmax_*             row softmax and its adjoint
//   act_*                 ReLU / exact-erf GELU
//   add_bcast, colsum     residual / positional adds with a broadcast period, bias gradients
//   pixel_shuffle         (n, h, w, co*4) <-> (n, 2h, 2w, co)
//   src_tokens            repeat_interleave(image_embeddings) + dense, NCHW -> token-major (mask_decoder.py:141-143)
//   resize_*              F.interpolate(bilinear, align_corners=False) and its adjoint (sam.py:155-189 under autograd)
//   mask_loss_*           sigmoid_ce_loss + dice_loss (LISA.py:16-59) per mask, fused forward sums and fused backward
//
// Layout: activations are token-major / channels-last fp32 throughout, so heads, transposes and the NCHW views of the
// reference are stride patterns of the GEMM, never copies.
#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>

#include "b200sam.h"
#include "b200sam_train.h"
#include "common.cuh"
#include "kernels.cuh"
#include "model.cuh"

namespace bsam {
namespace {

// ------------------------------------------------------------------------------------------------ strided GEMM
struct GemmStrides {
  long b1, b2, r, c;   // element strides: outer batch, inner batch, row, column
};

constexpr int kGT = 64;    // tile edge
constexpr int kGK = 16;    // K step

// 256 threads; each owns a TM x 4 block of the BM x 64 output tile (BM = 64: TM = 4; BM = 16: TM = 1, for the thin
// products of the decoder's attention - 6 tokens against 4096 - where a 64-row tile would be 90 % padding).  Tiles of A
// (BM x 16) and B (16 x 64) are staged through shared memory with the thread -> element map chosen so that the
// unit-stride dimension of each operand is the fastest-varying one across a warp (coalesced whatever the transposition).
template <int BM>
__global__ void __launch_bounds__(256) gemm_strided_kernel(const float* __restrict__ A, GemmStrides sa,
                                                           const float* __restrict__ B, GemmStrides sb,
                                                           float* __restrict__ Cm, GemmStrides sc,
                                                           const float* __restrict__ bias, int bias_mode, int nb2, int M,
                                                           int N, int K, float alpha, float beta, int relu, int ksplit) {
  constexpr int TM = BM / 16;
  constexpr int kALoads = BM * kGK / 256;      // 4 or 1 elements of the A tile per thread
  __shared__ __align__(16) float As[kGK][BM + 4];
  __shared__ __align__(16) float Bs[kGK][kGT + 4];
  const int bz = blockIdx.z / ksplit, ks = blockIdx.z - bz * ksplit;
  const int b1 = bz / nb2, b2 = bz - b1 * nb2;
  A += b1 * sa.b1 + b2 * sa.b2;
  B += b1 * sb.b1 + b2 * sb.b2;
  Cm += b1 * sc.b1 + b2 * sc.b2;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * kGT;
  const int t = threadIdx.x;
  const int tm = (t >> 4) * TM, tn = (t & 15) * 4;
  // K range of this split (multiples of kGK)
  const int ktiles = (K + kGK - 1) / kGK;
  const int per = (ktiles + ksplit - 1) / ksplit;
  const int kt0 = ks * per, kt1 = min(ktiles, kt0 + per);
  float acc[TM][4];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  const bool a_kfast = sa.c == 1 || sa.r != 1;   // A(m, k): k contiguous (or neither): lanes walk k
  const bool b_nfast = sb.c == 1 || sb.r != 1;   // B(k, n): n contiguous (or neither): lanes walk n
  for (int kt = kt0; kt < kt1; ++kt) {
    const int k0 = kt * kGK;
#pragma unroll
    for (int i = 0; i < kALoads; ++i) {
      const int e = t + i * 256;
      int m, k;
      if (a_kfast) {
        m = e >> 4;
        k = e & 15;
      } else {
        m = e % BM;
        k = e / BM;
      }
      float v = 0.f;
      if (m0 + m < M && k0 + k < K) v = A[(long)(m0 + m) * sa.r + (long)(k0 + k) * sa.c];
      As[k][m] = v;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int e = t + i * 256;
      int n, k;
      if (b_nfast) {
        n = e & 63;
        k = e >> 6;
      } else {
        n = e >> 4;
        k = e & 15;
      }
      float v = 0.f;
      if (n0 + n < N && k0 + k < K) v = B[(long)(k0 + k) * sb.r + (long)(n0 + n) * sb.c];
      Bs[k][n] = v;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kGK; ++k) {
      float av[TM];
      if constexpr (TM == 4) {
        const float4 a = *reinterpret_cast<const float4*>(&As[k][tm]);
        av[0] = a.x, av[1] = a.y, av[2] = a.z, av[3] = a.w;
      } else {
#pragma unroll
        for (int i = 0; i < TM; ++i) av[i] = As[k][tm + i];
      }
      const float4 b = *reinterpret_cast<const float4*>(&Bs[k][tn]);
      const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const int m = m0 + tm + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tn + j;
      if (n >= N) continue;
      float* c = Cm + (long)m * sc.r + (long)n * sc.c;
      float v = alpha * acc[i][j];
      if (ksplit > 1) {          // C was prepared (beta applied, bias added) by the launcher's first pass
        atomicAdd(c, v);
      } else {
        if (bias_mode == 1) v += bias[n];
        else if (bias_mode == 2) v += bias[m];
        if (beta != 0.f) v += beta * *c;
        if (relu) v = fmaxf(v, 0.f);
        *c = v;
      }
    }
  }
}

// ---- thin products: two of (M, N, K) are <= 32 and one is long (the decoder's attention between 6 tokens and 4096
// image positions with 16-wide heads; the hypernetwork product and its adjoints over 65536 pixels).  The tiled kernel
// wastes 75-90 % of its tile on them; these two stream the long dimension instead.
//
// thin_expand: N long, M <= MT, K <= 32.  One thread per column n: A (M x K) sits in shared memory, the thread reads its K
// values of B and produces M outputs.
template <int MT>
__global__ void __launch_bounds__(256) thin_expand_kernel(const float* __restrict__ A, GemmStrides sa,
                                                          const float* __restrict__ B, GemmStrides sb,
                                                          float* __restrict__ Cm, GemmStrides sc,
                                                          const float* __restrict__ bias, int bias_mode, int nb2, int M,
                                                          int N, int K, float alpha, float beta) {
  __shared__ float As[32][MT + 1];       // [k][m]
  const int bz = blockIdx.y;
  const int b1 = bz / nb2, b2 = bz - b1 * nb2;
  A += b1 * sa.b1 + b2 * sa.b2;
  B += b1 * sb.b1 + b2 * sb.b2;
  Cm += b1 * sc.b1 + b2 * sc.b2;
  for (int i = threadIdx.x; i < 32 * MT; i += 256) {
    const int k = i / MT, m = i - k * MT;
    As[k][m] = (m < M && k < K) ? A[(long)m * sa.r + (long)k * sa.c] : 0.f;
  }
  __syncthreads();
  const int n = blockIdx.x * 256 + threadIdx.x;
  if (n >= N) return;
  float acc[MT];
#pragma unroll
  for (int m = 0; m < MT; ++m) acc[m] = 0.f;
  const float* bp = B + (long)n * sb.c;
  if (sb.r == 1 && (K & 3) == 0 && ((reinterpret_cast<uintptr_t>(bp) & 15) == 0)) {   // K values contiguous per column
    for (int k = 0; k < K; k += 4) {
      const float4 v = *reinterpret_cast<const float4*>(bp + k);
      const float bv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int m = 0; m < MT; ++m) acc[m] = fmaf(As[k + j][m], bv[j], acc[m]);
    }
  } else {
    for (int k = 0; k < K; ++k) {
      const float bv = bp[(long)k * sb.r];
#pragma unroll
      for (int m = 0; m < MT; ++m) acc[m] = fmaf(As[k][m], bv, acc[m]);
    }
  }
#pragma unroll
  for (int m = 0; m < MT; ++m) {
    if (m < M) {
      float* c = Cm + (long)m * sc.r + (long)n * sc.c;
      float v = alpha * acc[m];
      if (bias_mode == 1) v += bias[n];
      else if (bias_mode == 2) v += bias[m];
      if (beta != 0.f) v += beta * *c;
      *c = v;
    }
  }
}

// thin_reduce: K long, M <= MT, N <= NT (MT * NT <= 128).  A block of 64 threads owns one K chunk of one batch: each
// thread accumulates the whole M x N result over its k values in registers, the block sums the 64 partial results
// through shared memory and adds them to C with one atomic per element (C prepared by gemm_prepare_kernel).
template <int MT, int NT>
__global__ void __launch_bounds__(64) thin_reduce_kernel(const float* __restrict__ A, GemmStrides sa,
                                                         const float* __restrict__ B, GemmStrides sb,
                                                         float* __restrict__ Cm, GemmStrides sc, int nb2, int M, int N,
                                                         int K, int kchunk, float alpha) {
  __shared__ float red[MT * NT][65];
  const int bz = blockIdx.y;
  const int b1 = bz / nb2, b2 = bz - b1 * nb2;
  A += b1 * sa.b1 + b2 * sa.b2;
  B += b1 * sb.b1 + b2 * sb.b2;
  Cm += b1 * sc.b1 + b2 * sc.b2;
  float acc[MT][NT];
#pragma unroll
  for (int m = 0; m < MT; ++m)
#pragma unroll
    for (int n = 0; n < NT; ++n) acc[m][n] = 0.f;
  const int k0 = blockIdx.x * kchunk, k1 = min(K, k0 + kchunk);
  const bool bvec = sb.c == 1 && N == NT && (sb.r & 3) == 0 && (reinterpret_cast<uintptr_t>(B) & 15) == 0;
  for (int k = k0 + threadIdx.x; k < k1; k += 64) {
    float av[MT], bv[NT];
#pragma unroll
    for (int m = 0; m < MT; ++m) av[m] = m < M ? A[(long)m * sa.r + (long)k * sa.c] : 0.f;
    const float* bp = B + (long)k * sb.r;
    if (bvec) {
#pragma unroll
      for (int n = 0; n < NT; n += 4) {
        const float4 v = *reinterpret_cast<const float4*>(bp + n);
        bv[n] = v.x, bv[n + 1] = v.y, bv[n + 2] = v.z, bv[n + 3] = v.w;
      }
    } else {
#pragma unroll
      for (int n = 0; n < NT; ++n) bv[n] = n < N ? bp[(long)n * sb.c] : 0.f;
    }
#pragma unroll
    for (int m = 0; m < MT; ++m)
#pragma unroll
      for (int n = 0; n < NT; ++n) acc[m][n] = fmaf(av[m], bv[n], acc[m][n]);
  }
#pragma unroll
  for (int m = 0; m < MT; ++m)
#pragma unroll
    for (int n = 0; n < NT; ++n) red[m * NT + n][threadIdx.x] = acc[m][n];
  __syncthreads();
  for (int e = threadIdx.x; e < MT * NT; e += 64) {
    const int m = e / NT, n = e - m * NT;
    if (m >= M || n >= N) continue;
    float sum = 0.f;
#pragma unroll 8
    for (int t = 0; t < 64; ++t) sum += red[e][t];     // row pitch 65: lanes hit distinct banks
    atomicAdd(Cm + (long)m * sc.r + (long)n * sc.c, alpha * sum);
  }
}

// first pass of a split-K product: C = beta * C + bias
__global__ void __launch_bounds__(256) gemm_prepare_kernel(float* __restrict__ Cm, GemmStrides sc,
                                                           const float* __restrict__ bias, int bias_mode, int nb2, int M,
                                                           int N, float beta) {
  const int bz = blockIdx.z;
  const int b1 = bz / nb2, b2 = bz - b1 * nb2;
  Cm += b1 * sc.b1 + b2 * sc.b2;
  const long i = (long)blockIdx.x * 256 + threadIdx.x;
  if (i >= (long)M * N) return;
  int m, n;
  if (sc.c == 1 || sc.r != 1) {
    m = (int)(i / N);
    n = (int)(i - (long)m * N);
  } else {
    n = (int)(i / M);
    m = (int)(i - (long)n * M);
  }
  float* c = Cm + (long)m * sc.r + (long)n * sc.c;
  float v = beta != 0.f ? beta * *c : 0.f;
  if (bias_mode == 1) v += bias[n];
  else if (bias_mode == 2) v += bias[m];
  *c = v;
}


// ------------------------------------------------------------------------------------------------ LayerNorm
// One warp per row of C channels (C % 32 == 0, C <= 1024).  Two-pass statistics in registers (biased variance, as
// nn.LayerNorm and LayerNorm2d, common.py:31-43).
template <int kMaxPerLane>
__global__ void __launch_bounds__(256) layernorm_fwd_kernel(const float* __restrict__ x, const float* __restrict__ g,
                                                            const float* __restrict__ b, long rows, int C, float eps,
                                                            float* __restrict__ y, float* __restrict__ mean_out,
                                                            float* __restrict__ rstd_out) {
  const int lane = threadIdx.x & 31;
  const long row = (long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int per = C >> 5;
  float v[kMaxPerLane];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < kMaxPerLane; ++i)
    if (i < per) {
      v[i] = x[row * C + i * 32 + lane];
      s += v[i];
    }
#pragma unroll
  for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const float mean = s / C;
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < kMaxPerLane; ++i)
    if (i < per) {
      const float d = v[i] - mean;
      q += d * d;
    }
#pragma unroll
  for (int o = 16; o; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
  const float rstd = rsqrtf(q / C + eps);
#pragma unroll
  for (int i = 0; i < kMaxPerLane; ++i)
    if (i < per) {
      const int c = i * 32 + lane;
      y[row * C + c] = (v[i] - mean) * rstd * g[c] + b[c];
    }
  if (lane == 0) {
    mean_out[row] = mean;
    rstd_out[row] = rstd;
  }
}

// dx = rstd * (dy*g - mean_c(dy*g) - xhat * mean_c(dy*g*xhat));  dgamma += sum_r dy*xhat;  dbeta += sum_r dy.
// Each warp walks a strided set of rows keeping its dgamma / dbeta partials in registers; one shared-memory reduction
// per block, then one atomicAdd per channel per block into the (zeroed) outputs.
template <int kMaxPerLane>
__global__ void __launch_bounds__(256) layernorm_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ x,
                                                            const float* __restrict__ g, const float* __restrict__ mean,
                                                            const float* __restrict__ rstd, long rows, int C,
                                                            float* __restrict__ dx, float* __restrict__ dgamma,
                                                            float* __restrict__ dbeta) {
  __shared__ float red[2][1024];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int per = C >> 5;
  float pg[kMaxPerLane], pb[kMaxPerLane], gg[kMaxPerLane];
#pragma unroll
  for (int i = 0; i < kMaxPerLane; ++i) {
    pg[i] = pb[i] = 0.f;
    gg[i] = i < per ? g[i * 32 + lane] : 0.f;
  }
  for (int c = threadIdx.x; c < C; c += 256) red[0][c] = red[1][c] = 0.f;
  __syncthreads();
  for (long row = (long)blockIdx.x * 8 + warp; row < rows; row += (long)gridDim.x * 8) {
    const float mu = mean[row], rs = rstd[row];
    float xh[kMaxPerLane], dg[kMaxPerLane];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < kMaxPerLane; ++i)
      if (i < per) {
        const long o = row * C + i * 32 + lane;
        const float d = dy[o];
        xh[i] = (x[o] - mu) * rs;
        dg[i] = d * gg[i];
        s1 += dg[i];
        s2 += dg[i] * xh[i];
        pg[i] += d * xh[i];
        pb[i] += d;
      }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
      s1 += __shfl_xor_sync(0xffffffffu, s1, o);
      s2 += __shfl_xor_sync(0xffffffffu, s2, o);
    }
    const float m1 = s1 / C, m2 = s2 / C;
#pragma unroll
    for (int i = 0; i < kMaxPerLane; ++i)
      if (i < per) dx[row * C + i * 32 + lane] = rs * (dg[i] - m1 - xh[i] * m2);
  }
#pragma unroll
  for (int i = 0; i < kMaxPerLane; ++i)
    if (i < per) {
      atomicAdd(&red[0][i * 32 + lane], pg[i]);
      atomicAdd(&red[1][i * 32 + lane], pb[i]);
    }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += 256) {
    atomicAdd(dgamma + c, red[0][c]);
    atomicAdd(dbeta + c, red[1][c]);
  }
}

// ------------------------------------------------------------------------------------------------ softmax
// One warp per row of length L (any L); in place allowed.
__global__ void __launch_bounds__(256) softmax_fwd_kernel(const float* __restrict__ x, long rows, int L,
                                                          float* __restrict__ y) {
  const int lane = threadIdx.x & 31;
  const long row = (long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const float* xr = x + row * L;
  float* yr = y + row * L;
  float m = -INFINITY;
  for (int i = lane; i < L; i += 32) m = fmaxf(m, xr[i]);
#pragma unroll
  for (int o = 16; o; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  float s = 0.f;
  for (int i = lane; i < L; i += 32) {
    const float e = expf(xr[i] - m);
    yr[i] = e;
    s += e;
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const float inv = 1.f / s;
  for (int i = lane; i < L; i += 32) yr[i] *= inv;
}

// dx = y * (dy - sum(dy * y)); in place on dy allowed
__global__ void __launch_bounds__(256) softmax_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ y,
                                                          long rows, int L, float* __restrict__ dx) {
  const int lane = threadIdx.x & 31;
  const long row = (long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const float* dr = dy + row * L;
  const float* yr = y + row * L;
  float s = 0.f;
  for (int i = lane; i < L; i += 32) s += dr[i] * yr[i];
#pragma unroll
  for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  for (int i = lane; i < L; i += 32) dx[row * L + i] = yr[i] * (dr[i] - s);
}

// ------------------------------------------------------------------------------------------------ elementwise
__device__ __forceinline__ float gelu_exact(float x) { return 0.5f * x * (1.f + erff(x * 0.70710678118654752f)); }
__device__ __forceinline__ float gelu_exact_grad(float x) {
  const float cdf = 0.5f * (1.f + erff(x * 0.70710678118654752f));
  const float pdf = 0.39894228040143268f * expf(-0.5f * x * x);
  return cdf + x * pdf;
}

// kind 0: ReLU, 1: GELU (erf form, nn.GELU default)
__global__ void __launch_bounds__(256) act_fwd_kernel(int kind, const float* __restrict__ x, long n,
                                                      float* __restrict__ y) {
  for (long i = (long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long)gridDim.x * 256) {
    const float v = x[i];
    y[i] = kind == 0 ? fmaxf(v, 0.f) : gelu_exact(v);
  }
}
// ReLU: `ref` is the OUTPUT (y > 0); GELU: `ref` is the INPUT
__global__ void __launch_bounds__(256) act_bwd_kernel(int kind, const float* __restrict__ dy,
                                                      const float* __restrict__ ref, long n, float* __restrict__ dx) {
  for (long i = (long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long)gridDim.x * 256) {
    const float r = ref[i];
    dx[i] = kind == 0 ? (r > 0.f ? dy[i] : 0.f) : dy[i] * gelu_exact_grad(r);
  }
}

// out[r][c] = a[r][c] + b[r % mod][c]
__global__ void __launch_bounds__(256) add_bcast_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                                        long rows, int C, long mod, float* __restrict__ out) {
  const long n = rows * C;
  for (long i = (long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long)gridDim.x * 256) {
    const long r = i / C;
    const int c = (int)(i - r * C);
    out[i] = a[i] + b[(r % mod) * C + c];
  }
}

// out[c] += sum over rows of x[r][c]  (out zeroed by the launcher).  Block = 256 threads = (256 / C') row lanes.
__global__ void __launch_bounds__(256) colsum_kernel(const float* __restrict__ x, long rows, int C,
                                                     float* __restrict__ out) {
  // thread t handles column c = t % C (C <= 256 per pass) for rows t / C, stepping by rows-per-block
  for (int c0 = 0; c0 < C; c0 += 256) {
    const int cw = min(256, C - c0);
    const int c = threadIdx.x % cw, rl = threadIdx.x / cw, rpb = 256 / cw;
    if (rl >= rpb) continue;
    float s = 0.f;
    for (long r = (long)blockIdx.x * rpb + rl; r < rows; r += (long)gridDim.x * rpb) s += x[r * C + c0 + c];
    atomicAdd(out + c0 + c, s);
  }
}

// (n, h, w, co*4) with column = co*4 + ky*2 + kx  <->  (n, 2h, 2w, co): the scatter of ConvTranspose2d(k=2, s=2)
__global__ void __launch_bounds__(256) pixel_shuffle_kernel(const float* __restrict__ x, long n_img, int hgt, int wid,
                                                            int co, int inverse, float* __restrict__ y) {
  const long total = n_img * hgt * wid * co * 4;
  for (long i = (long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long)gridDim.x * 256) {
    // i enumerates the (n, 2h, 2w, co) side
    const int c = (int)(i % co);
    long r = i / co;
    const int X = (int)(r % (2 * wid));
    r /= 2 * wid;
    const int Y = (int)(r % (2 * hgt));
    const long n = r / (2 * hgt);
    const long j = (((n * hgt + (Y >> 1)) * wid + (X >> 1)) * co + c) * 4 + (Y & 1) * 2 + (X & 1);
    if (inverse) y[j] = x[i];
    else y[i] = x[j];
  }
}

// src[p][t][c] = emb[frame[p]][c][t] + dense[c]   (emb NCHW fp32 / bf16 / f16, t = y*64 + x); 32 x 32 transposing tiles
template <typename T>
__global__ void __launch_bounds__(256) src_tokens_kernel(const T* __restrict__ emb, const int* __restrict__ frame,
                                                         const float* __restrict__ dense, float* __restrict__ out) {
  __shared__ float tile[32][33];
  const long p = blockIdx.z;
  const int f = frame ? frame[p] : 0;
  const int t0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int j = ty; j < 32; j += 8) tile[j][tx] = static_cast<float>(emb[((long)f * kOutCh + c0 + j) * kTokens + t0 + tx]);
  __syncthreads();
  for (int j = ty; j < 32; j += 8)
    out[(p * kTokens + t0 + j) * kOutCh + c0 + tx] = tile[tx][j] + dense[c0 + tx];
}


// ------------------------------------------------------------------------------------------------ bilinear resize
// F.interpolate(mode="bilinear", align_corners=False) index arithmetic (ATen area_pixel_compute_source_index with
// scale = in / out in fp32): src = scale * (dst + 0.5) - 0.5 clamped at 0, i1 = min(i0 + 1, in - 1).
__device__ __forceinline__ void bilinear_src(int dst, float scale, int in, int* i0, int* i1, float* l1) {
  float s = scale * (dst + 0.5f) - 0.5f;
  if (s < 0.f) s = 0.f;
  int a = (int)s;
  if (a > in - 1) a = in - 1;
  *i0 = a;
  *i1 = a + (a < in - 1 ? 1 : 0);
  *l1 = s - a;
}

// x: n planes of (in_h, in_w) live pixels inside a pitched plane (the crop masks[..., :h, :w] of sam.py:181 is a view)
__global__ void __launch_bounds__(256) resize_fwd_kernel(const float* __restrict__ x, long n, int in_h, int in_w,
                                                         long pitch, long plane, int out_h, int out_w,
                                                         float* __restrict__ y) {
  const float sh = (float)in_h / out_h, sw = (float)in_w / out_w;
  const long total = n * out_h * out_w;
  for (long i = (long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long)gridDim.x * 256) {
    const int X = (int)(i % out_w);
    long r = i / out_w;
    const int Y = (int)(r % out_h);
    const long p = r / out_h;
    int y0, y1, x0, x1;
    float ly, lx;
    bilinear_src(Y, sh, in_h, &y0, &y1, &ly);
    bilinear_src(X, sw, in_w, &x0, &x1, &lx);
    const float* s = x + p * plane;
    const float hy = 1.f - ly, hx = 1.f - lx;
    y[i] = hy * (hx * s[y0 * pitch + x0] + lx * s[y0 * pitch + x1]) +
           ly * (hx * s[y1 * pitch + x0] + lx * s[y1 * pitch + x1]);
  }
}

// adjoint: dx (zeroed by the caller, same pitched layout) += the four weights of every output pixel
__global__ void __launch_bounds__(256) resize_bwd_kernel(const float* __restrict__ dy, long n, int in_h, int in_w,
                                                         long pitch, long plane, int out_h, int out_w,
                                                         float* __restrict__ dx) {
  const float sh = (float)in_h / out_h, sw = (float)in_w / out_w;
  const long total = n * out_h * out_w;
  for (long i = (long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long)gridDim.x * 256) {
    const int X = (int)(i % out_w);
    long r = i / out_w;
    const int Y = (int)(r % out_h);
    const long p = r / out_h;
    int y0, y1, x0, x1;
    float ly, lx;
    bilinear_src(Y, sh, in_h, &y0, &y1, &ly);
    bilinear_src(X, sw, in_w, &x0, &x1, &lx);
    float* d = dx + p * plane;
    const float g = dy[i], hy = 1.f - ly, hx = 1.f - lx;
    atomicAdd(d + y0 * pitch + x0, g * hy * hx);
    atomicAdd(d + y0 * pitch + x1, g * hy * lx);
    atomicAdd(d + y1 * pitch + x0, g * ly * hx);
    atomicAdd(d + y1 * pitch + x1, g * ly * lx);
  }
}

// ------------------------------------------------------------------------------------------------ mask losses
// Per mask p with z = w[p] * logit (the taxonomy gate of LISA.py:357-358 applied before both losses):
//   sums[p] = { sum bce_with_logits(z, t), sum sigmoid(z), sum sigmoid(z) * t, sum t }   (double accumulators)
// sigmoid_ce_loss (LISA.py:42-59) = sums[0] / HW per mask; dice_loss (:16-39) = 1 - (2*sums[2]/scale + eps) /
// (sums[1]/scale + sums[3]/scale + eps) per mask.  grid (blocks, P).
__global__ void __launch_bounds__(256) mask_loss_fwd_kernel(const float* __restrict__ logits,
                                                            const float* __restrict__ gt, const float* __restrict__ w,
                                                            long hw, double* __restrict__ sums) {
  __shared__ double red[8][4];
  const long p = blockIdx.y;
  const float wp = w ? w[p] : 1.f;
  const float* x = logits + p * hw;
  const float* t = gt + p * hw;
  double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
  for (long i0 = (long)blockIdx.x * 256 * 8; i0 < hw; i0 += (long)gridDim.x * 256 * 8) {
    float f0 = 0.f, f1 = 0.f, f2 = 0.f, f3 = 0.f;   // fp32 over 8 elements, then promoted
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const long i = i0 + j * 256 + threadIdx.x;
      if (i < hw) {
        const float z = wp * x[i], tt = t[i];
        const float sg = 1.f / (1.f + expf(-z));
        f0 += fmaxf(z, 0.f) - z * tt + log1pf(expf(-fabsf(z)));
        f1 += sg;
        f2 += sg * tt;
        f3 += tt;
      }
    }
    a0 += f0, a1 += f1, a2 += f2, a3 += f3;
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    a0 += __shfl_xor_sync(0xffffffffu, a0, o);
    a1 += __shfl_xor_sync(0xffffffffu, a1, o);
    a2 += __shfl_xor_sync(0xffffffffu, a2, o);
    a3 += __shfl_xor_sync(0xffffffffu, a3, o);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) red[warp][0] = a0, red[warp][1] = a1, red[warp][2] = a2, red[warp][3] = a3;
  __syncthreads();
  if (threadIdx.x < 4) {
    double s = 0;
    for (int k = 0; k < 8; ++k) s += red[k][threadIdx.x];
    atomicAdd(sums + p * 4 + threadIdx.x, s);
  }
}

// dlogit = w * ( g_bce/HW * (s - t) + g_dice * ddice/ds * s (1 - s) ),
// ddice/ds_i = -( (2 t_i / scale) * (den + eps) - (num + eps) / scale ) / (den + eps)^2
__global__ void __launch_bounds__(256) mask_loss_bwd_kernel(const float* __restrict__ logits,
                                                            const float* __restrict__ gt, const float* __restrict__ w,
                                                            const double* __restrict__ sums,
                                                            const float* __restrict__ g_bce,
                                                            const float* __restrict__ g_dice, long hw, float scale,
                                                            float eps, float* __restrict__ dlogits) {
  const long p = blockIdx.y;
  const float wp = w ? w[p] : 1.f;
  const double num = 2.0 * sums[p * 4 + 2] / scale + eps;
  const double den = sums[p * 4 + 1] / scale + sums[p * 4 + 3] / scale + eps;
  const float ca = (float)(-2.0 / scale / den);            // coefficient of t_i
  const float cb = (float)(num / scale / (den * den));     // constant term
  const float gb = g_bce[p] / (float)hw, gd = g_dice[p];
  const float* x = logits + p * hw;
  const float* t = gt + p * hw;
  float* d = dlogits + p * hw;
  for (long i = (long)blockIdx.x * 256 + threadIdx.x; i < hw; i += (long)gridDim.x * 256) {
    const float z = wp * x[i], tt = t[i];
    const float sg = 1.f / (1.f + expf(-z));
    d[i] = wp * (gb * (sg - tt) + gd * (ca * tt + cb) * sg * (1.f - sg));
  }
}

// y (cols, rows) = x (rows, cols)^T, 32 x 32 tiles through shared memory (both sides coalesced)
__global__ void __launch_bounds__(256) transpose_kernel(const float* __restrict__ x, long rows, int cols,
                                                        float* __restrict__ y) {
  __shared__ float tile[32][33];
  const long r0 = (long)blockIdx.x * 32;
  const int c0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int j = ty; j < 32; j += 8)
    if (r0 + j < rows && c0 + tx < cols) tile[j][tx] = x[(r0 + j) * cols + c0 + tx];
  __syncthreads();
  for (int j = ty; j < 32; j += 8)
    if (c0 + j < cols && r0 + tx < rows) y[(long)(c0 + j) * rows + r0 + tx] = tile[tx][j];
}

int tfail(b200sam_t* h, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  if (h) h->err = buf;
  return code;
}

int grid_for(long n, int per_block = 256) {
  long g = (n + per_block - 1) / per_block;
  const long cap = (long)gemm_num_sms() * 16;
  if (g > cap) g = cap;
  return g < 1 ? 1 : (int)g;
}

struct TGuard {
  int prev = -1;
  explicit TGuard(int dev) {
    cudaGetDevice(&prev);
    if (prev != dev) cudaSetDevice(dev);
    else prev = -1;
  }
  ~TGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

#define T_ENTER()                                                    \
  if (!h) return B200SAM_ERR_INVALID;                                \
  TGuard guard_(h->device);                                          \
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);           \
  (void)s
#define T_DONE(nl)                                                                                            \
  do {                                                                                                        \
    cudaError_t e_ = cudaGetLastError();                                                                      \
    if (e_ != cudaSuccess) return tfail(h, B200SAM_ERR_CUDA, "%s: %s", __func__, cudaGetErrorString(e_));     \
    h->launches += (nl);                                                                                      \
    return 0;                                                                                                 \
  } while (0)

}  // namespace
}  // namespace bsam

using namespace bsam;

extern "C" int b200sam_train_gemm(b200sam_t* h, const float* A, const int64_t* sa, const float* B, const int64_t* sb,
                                  float* Cm, const int64_t* sc, const float* bias, int bias_mode, int nb1, int nb2,
                                  int M, int N, int K, float alpha, float beta, int relu, void* stream) {
  T_ENTER();
  if (!A || !B || !Cm || !sa || !sb || !sc || nb1 < 0 || nb2 < 0 || M < 0 || N < 0 || K < 0 || bias_mode < 0 ||
      bias_mode > 2 || (bias_mode && !bias))
    return tfail(h, B200SAM_ERR_INVALID, "train_gemm: bad arguments");
  const long nb = (long)nb1 * nb2;
  if (nb == 0 || M == 0 || N == 0) return 0;
  if (nb > 65535L) return tfail(h, B200SAM_ERR_INVALID, "train_gemm: too many batches");
  GemmStrides a{sa[0], sa[1], sa[2], sa[3]}, b{sb[0], sb[1], sb[2], sb[3]}, c{sc[0], sc[1], sc[2], sc[3]};
  // thin products: put the small dimension on the rows (C^T = B^T A^T is the same kernel with the operands and the
  // row / column strides exchanged) and use the 16-row tile
  if (N <= 16 && M > N) {
    const float* tp = A;
    A = B;
    B = tp;
    GemmStrides na{b.b1, b.b2, b.c, b.r}, nb_{a.b1, a.b2, a.c, a.r};
    a = na;
    b = nb_;
    const long t2 = c.r;
    c.r = c.c;
    c.c = t2;
    const int t3 = M;
    M = N;
    N = t3;
    if (bias_mode) bias_mode = 3 - bias_mode;
  }
  // thin products (see thin_expand_kernel / thin_reduce_kernel)
  if (!relu && M <= 32 && K <= 32 && N >= 1024 && nb <= 65535) {
    dim3 grid((unsigned)((N + 255) / 256), (unsigned)nb);
    if (M <= 8) thin_expand_kernel<8><<<grid, 256, 0, s>>>(A, a, B, b, Cm, c, bias, bias_mode, nb2, M, N, K, alpha, beta);
    else if (M <= 16) thin_expand_kernel<16><<<grid, 256, 0, s>>>(A, a, B, b, Cm, c, bias, bias_mode, nb2, M, N, K, alpha, beta);
    else thin_expand_kernel<32><<<grid, 256, 0, s>>>(A, a, B, b, Cm, c, bias, bias_mode, nb2, M, N, K, alpha, beta);
    T_DONE(1);
  }
  if (!relu && K >= 2048 && nb <= 65535 && ((M <= 8 && N <= 16) || (M <= 4 && N <= 32))) {
    const int kchunk = K >= 32768 ? 4096 : 1024;
    dim3 g0((unsigned)(((long)M * N + 255) / 256), 1, (unsigned)nb);
    gemm_prepare_kernel<<<g0, 256, 0, s>>>(Cm, c, bias, bias_mode, nb2, M, N, beta);
    dim3 grid((unsigned)((K + kchunk - 1) / kchunk), (unsigned)nb);
    if (M <= 4 && N <= 16) thin_reduce_kernel<4, 16><<<grid, 64, 0, s>>>(A, a, B, b, Cm, c, nb2, M, N, K, kchunk, alpha);
    else if (M <= 4) thin_reduce_kernel<4, 32><<<grid, 64, 0, s>>>(A, a, B, b, Cm, c, nb2, M, N, K, kchunk, alpha);
    else if (M <= 6) thin_reduce_kernel<6, 16><<<grid, 64, 0, s>>>(A, a, B, b, Cm, c, nb2, M, N, K, kchunk, alpha);
    else thin_reduce_kernel<8, 16><<<grid, 64, 0, s>>>(A, a, B, b, Cm, c, nb2, M, N, K, kchunk, alpha);
    T_DONE(2);
  }
  const int bm = M <= 16 ? 16 : kGT;
  const int tx = (N + kGT - 1) / kGT, ty = (M + bm - 1) / bm;
  const long tiles = (long)tx * ty * nb;
  const int ktiles = (K + kGK - 1) / kGK;
  // split K when the output has too few tiles to fill the machine and the reduction is long (weight gradients)
  int ksplit = 1;
  if (!relu && tiles < 2L * gemm_num_sms() && ktiles >= 64) {
    long want = (4L * gemm_num_sms() + tiles - 1) / tiles;
    if (want > ktiles / 8) want = ktiles / 8;
    if (nb * want > 65535) want = 65535 / nb;
    if (want > 1) ksplit = (int)want;
  }
  if (ty > 65535) return tfail(h, B200SAM_ERR_INVALID, "train_gemm: too many row tiles");
  int nl = 1;
  if (ksplit > 1) {
    dim3 g((unsigned)(((long)M * N + 255) / 256), 1, (unsigned)nb);
    gemm_prepare_kernel<<<g, 256, 0, s>>>(Cm, c, bias, bias_mode, nb2, M, N, beta);
    ++nl;
  }
  dim3 grid(tx, ty, (unsigned)(nb * ksplit));
  if (bm == 16)
    gemm_strided_kernel<16><<<grid, 256, 0, s>>>(A, a, B, b, Cm, c, bias, bias_mode, nb2, M, N, K, alpha, beta, relu, ksplit);
  else
    gemm_strided_kernel<64><<<grid, 256, 0, s>>>(A, a, B, b, Cm, c, bias, bias_mode, nb2, M, N, K, alpha, beta, relu, ksplit);
  T_DONE(nl);
}

extern "C" int b200sam_train_layernorm_fwd(b200sam_t* h, const float* x, const float* gamma, const float* beta,
                                           int64_t rows, int C, float eps, float* y, float* mean, float* rstd,
                                           void* stream) {
  T_ENTER();
  if (!x || !gamma || !beta || !y || !mean || !rstd || rows < 0 || C <= 0 || C % 32 || C > 1024)
    return tfail(h, B200SAM_ERR_INVALID, "train_layernorm_fwd: bad arguments (C %% 32 == 0, C <= 1024)");
  if (rows == 0) return 0;
  const unsigned g = (unsigned)((rows + 7) / 8);
  if (C <= 256) layernorm_fwd_kernel<8><<<g, 256, 0, s>>>(x, gamma, beta, rows, C, eps, y, mean, rstd);
  else layernorm_fwd_kernel<32><<<g, 256, 0, s>>>(x, gamma, beta, rows, C, eps, y, mean, rstd);
  T_DONE(1);
}

extern "C" int b200sam_train_layernorm_bwd(b200sam_t* h, const float* dy, const float* x, const float* gamma,
                                           const float* mean, const float* rstd, int64_t rows, int C, float* dx,
                                           float* dgamma, float* dbeta, void* stream) {
  T_ENTER();
  if (!dy || !x || !gamma || !mean || !rstd || !dx || !dgamma || !dbeta || rows < 0 || C <= 0 || C % 32 || C > 1024)
    return tfail(h, B200SAM_ERR_INVALID, "train_layernorm_bwd: bad arguments");
  if (cudaMemsetAsync(dgamma, 0, C * sizeof(float), s) != cudaSuccess ||
      cudaMemsetAsync(dbeta, 0, C * sizeof(float), s) != cudaSuccess)
    return tfail(h, B200SAM_ERR_CUDA, "train_layernorm_bwd: memset failed");
  if (rows == 0) return 0;
  long g = (rows + 7) / 8;
  const long cap = (long)gemm_num_sms() * 4;
  if (g > cap) g = cap;
  if (C <= 256) layernorm_bwd_kernel<8><<<(unsigned)g, 256, 0, s>>>(dy, x, gamma, mean, rstd, rows, C, dx, dgamma, dbeta);
  else layernorm_bwd_kernel<32><<<(unsigned)g, 256, 0, s>>>(dy, x, gamma, mean, rstd, rows, C, dx, dgamma, dbeta);
  T_DONE(1);
}

extern "C" int b200sam_train_softmax_fwd(b200sam_t* h, const float* x, int64_t rows, int L, float* y, void* stream) {
  T_ENTER();
  if (!x || !y || rows < 0 || L <= 0) return tfail(h, B200SAM_ERR_INVALID, "train_softmax_fwd: bad arguments");
  if (rows == 0) return 0;
  softmax_fwd_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, s>>>(x, rows, L, y);
  T_DONE(1);
}

extern "C" int b200sam_train_softmax_bwd(b200sam_t* h, const float* dy, const float* y, int64_t rows, int L, float* dx,
                                         void* stream) {
  T_ENTER();
  if (!dy || !y || !dx || rows < 0 || L <= 0) return tfail(h, B200SAM_ERR_INVALID, "train_softmax_bwd: bad arguments");
  if (rows == 0) return 0;
  softmax_bwd_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, s>>>(dy, y, rows, L, dx);
  T_DONE(1);
}

extern "C" int b200sam_train_act_fwd(b200sam_t* h, int kind, const float* x, int64_t n, float* y, void* stream) {
  T_ENTER();
  if (!x || !y || n < 0 || kind < 0 || kind > 1) return tfail(h, B200SAM_ERR_INVALID, "train_act_fwd: bad arguments");
  if (n == 0) return 0;
  act_fwd_kernel<<<grid_for(n), 256, 0, s>>>(kind, x, n, y);
  T_DONE(1);
}

extern "C" int b200sam_train_act_bwd(b200sam_t* h, int kind, const float* dy, const float* ref, int64_t n, float* dx,
                                     void* stream) {
  T_ENTER();
  if (!dy || !ref || !dx || n < 0 || kind < 0 || kind > 1)
    return tfail(h, B200SAM_ERR_INVALID, "train_act_bwd: bad arguments");
  if (n == 0) return 0;
  act_bwd_kernel<<<grid_for(n), 256, 0, s>>>(kind, dy, ref, n, dx);
  T_DONE(1);
}

extern "C" int b200sam_train_add_bcast(b200sam_t* h, const float* a, const float* b, int64_t rows, int C, int64_t mod,
                                       float* out, void* stream) {
  T_ENTER();
  if (!a || !b || !out || rows < 0 || C <= 0 || mod <= 0)
    return tfail(h, B200SAM_ERR_INVALID, "train_add_bcast: bad arguments");
  if (rows == 0) return 0;
  add_bcast_kernel<<<grid_for(rows * C), 256, 0, s>>>(a, b, rows, C, mod, out);
  T_DONE(1);
}

extern "C" int b200sam_train_colsum(b200sam_t* h, const float* x, int64_t rows, int C, float* out, void* stream) {
  T_ENTER();
  if (!x || !out || rows < 0 || C <= 0) return tfail(h, B200SAM_ERR_INVALID, "train_colsum: bad arguments");
  if (cudaMemsetAsync(out, 0, C * sizeof(float), s) != cudaSuccess)
    return tfail(h, B200SAM_ERR_CUDA, "train_colsum: memset failed");
  if (rows == 0) return 0;
  const int cw = C < 256 ? C : 256, rpb = 256 / cw;
  long g = (rows + rpb - 1) / rpb;
  const long cap = (long)gemm_num_sms() * 4;
  if (g > cap) g = cap;
  colsum_kernel<<<(unsigned)g, 256, 0, s>>>(x, rows, C, out);
  T_DONE(1);
}

extern "C" int b200sam_train_pixel_shuffle(b200sam_t* h, const float* x, int64_t n_img, int hgt, int wid, int co,
                                           int inverse, float* y, void* stream) {
  T_ENTER();
  if (!x || !y || n_img < 0 || hgt <= 0 || wid <= 0 || co <= 0)
    return tfail(h, B200SAM_ERR_INVALID, "train_pixel_shuffle: bad arguments");
  if (n_img == 0) return 0;
  pixel_shuffle_kernel<<<grid_for(n_img * hgt * wid * co * 4), 256, 0, s>>>(x, n_img, hgt, wid, co, inverse, y);
  T_DONE(1);
}

extern "C" int b200sam_train_src_tokens(b200sam_t* h, const void* emb_nchw, int emb_dtype, int n_frames,
                                        const int32_t* frame_of_prompt, const float* dense_vec, int P, float* out,
                                        void* stream) {
  T_ENTER();
  if (!emb_nchw || !dense_vec || !out || P < 0 || n_frames <= 0 || emb_dtype < 0 || emb_dtype > 2)
    return tfail(h, B200SAM_ERR_INVALID, "train_src_tokens: bad arguments");
  if (P == 0) return 0;
  if (!frame_of_prompt && n_frames != 1)
    return tfail(h, B200SAM_ERR_INVALID, "train_src_tokens: frame_of_prompt is required for more than one frame");
  dim3 grid(kTokens / 32, kOutCh / 32, P);
  if (emb_dtype == DT_F32)
    src_tokens_kernel<float><<<grid, 256, 0, s>>>(reinterpret_cast<const float*>(emb_nchw), frame_of_prompt, dense_vec, out);
  else if (emb_dtype == DT_BF16)
    src_tokens_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>(reinterpret_cast<const __nv_bfloat16*>(emb_nchw), frame_of_prompt, dense_vec, out);
  else
    src_tokens_kernel<__half><<<grid, 256, 0, s>>>(reinterpret_cast<const __half*>(emb_nchw), frame_of_prompt, dense_vec, out);
  T_DONE(1);
}

extern "C" int b200sam_train_resize_fwd(b200sam_t* h, const float* x, int64_t n, int in_h, int in_w, int64_t pitch,
                                        int64_t plane, int out_h, int out_w, float* y, void* stream) {
  T_ENTER();
  if (!x || !y || n < 0 || in_h <= 0 || in_w <= 0 || out_h <= 0 || out_w <= 0 || pitch < in_w || plane < pitch * in_h)
    return tfail(h, B200SAM_ERR_INVALID, "train_resize_fwd: bad arguments");
  if (n == 0) return 0;
  resize_fwd_kernel<<<grid_for(n * out_h * out_w), 256, 0, s>>>(x, n, in_h, in_w, pitch, plane, out_h, out_w, y);
  T_DONE(1);
}

extern "C" int b200sam_train_resize_bwd(b200sam_t* h, const float* dy, int64_t n, int in_h, int in_w, int64_t pitch,
                                        int64_t plane, int out_h, int out_w, float* dx, void* stream) {
  T_ENTER();
  if (!dy || !dx || n < 0 || in_h <= 0 || in_w <= 0 || out_h <= 0 || out_w <= 0 || pitch < in_w || plane < pitch * in_h)
    return tfail(h, B200SAM_ERR_INVALID, "train_resize_bwd: bad arguments");
  if (n == 0) return 0;
  if (cudaMemsetAsync(dx, 0, (size_t)n * plane * sizeof(float), s) != cudaSuccess)
    return tfail(h, B200SAM_ERR_CUDA, "train_resize_bwd: memset failed");
  resize_bwd_kernel<<<grid_for(n * out_h * out_w), 256, 0, s>>>(dy, n, in_h, in_w, pitch, plane, out_h, out_w, dx);
  T_DONE(1);
}

extern "C" int b200sam_train_mask_loss_fwd(b200sam_t* h, const float* logits, const float* gt, const float* weight,
                                           int P, int64_t hw, double* sums, void* stream) {
  T_ENTER();
  if (!logits || !gt || !sums || P < 0 || P > 65535 || hw <= 0)
    return tfail(h, B200SAM_ERR_INVALID, "train_mask_loss_fwd: bad arguments");
  if (P == 0) return 0;
  if (cudaMemsetAsync(sums, 0, (size_t)P * 4 * sizeof(double), s) != cudaSuccess)
    return tfail(h, B200SAM_ERR_CUDA, "train_mask_loss_fwd: memset failed");
  long gx = (hw + 2047) / 2048;
  const long cap = ((long)gemm_num_sms() * 8 + P - 1) / P;
  if (gx > cap) gx = cap;
  dim3 grid((unsigned)gx, P);
  mask_loss_fwd_kernel<<<grid, 256, 0, s>>>(logits, gt, weight, hw, sums);
  T_DONE(1);
}

extern "C" int b200sam_train_mask_loss_bwd(b200sam_t* h, const float* logits, const float* gt, const float* weight,
                                           const double* sums, const float* g_bce, const float* g_dice, int P,
                                           int64_t hw, float scale, float eps, float* dlogits, void* stream) {
  T_ENTER();
  if (!logits || !gt || !sums || !g_bce || !g_dice || !dlogits || P < 0 || P > 65535 || hw <= 0 || scale <= 0.f)
    return tfail(h, B200SAM_ERR_INVALID, "train_mask_loss_bwd: bad arguments");
  if (P == 0) return 0;
  long gx = (hw + 255) / 256;
  const long cap = ((long)gemm_num_sms() * 16 + P - 1) / P;
  if (gx > cap) gx = cap;
  dim3 grid((unsigned)gx, P);
  mask_loss_bwd_kernel<<<grid, 256, 0, s>>>(logits, gt, weight, sums, g_bce, g_dice, hw, scale, eps, dlogits);
  T_DONE(1);
}

extern "C" int b200sam_train_transpose(b200sam_t* h, const float* x, int64_t rows, int cols, float* y, void* stream) {
  T_ENTER();
  if (!x || !y || rows < 0 || cols <= 0) return tfail(h, B200SAM_ERR_INVALID, "train_transpose: bad arguments");
  if (rows == 0) return 0;
  const long gx = (rows + 31) / 32;
  if (gx > 0x7fffffffL || (cols + 31) / 32 > 65535) return tfail(h, B200SAM_ERR_INVALID, "train_transpose: too large");
  dim3 grid((unsigned)gx, (unsigned)((cols + 31) / 32));
  transpose_kernel<<<grid, 256, 0, s>>>(x, rows, cols, y);
  T_DONE(1);
}

extern "C" int b200sam_train_gemm_tc(b200sam_t* h, const float* A, int64_t lda, const float* W, int64_t ldw, float* out,
                                     int64_t ldo, const float* bias, int M, int N, int K, int split_k, void* stream) {
  T_ENTER();
  if (!A || !W || !out || M < 0 || N <= 0 || K <= 0 || N % 32 || K % 8 || lda < K || ldw < K || ldo < N ||
      lda > 0x7fffffffL || ldw > 0x7fffffffL || ldo > 0x7fffffffL)
    return tfail(h, B200SAM_ERR_INVALID, "train_gemm_tc: bad arguments (N %% 32 == 0, K %% 8 == 0)");
  if (M == 0) return 0;
  GemmEpi e;
  e.out = out;
  e.ld_out = (int)ldo;
  int nl = 1;
  if (split_k) {
    // long reduction into a small output: one K range per CTA, about two CTAs per SM, combined with atomics
    if (bias) return tfail(h, B200SAM_ERR_INVALID, "train_gemm_tc: split_k has no bias");
    const int bn = N <= 128 ? 128 : 256;
    const long tiles = (long)((M + 127) / 128) * ((N + bn - 1) / bn);
    const int kblocks = (K + 31) / 32;
    long want = (2L * gemm_num_sms() + tiles - 1) / tiles;
    if (want > kblocks / 4) want = kblocks / 4;
    if (want > 1) {
      const int kper = (int)((kblocks + want - 1) / want);
      e.ksplit = (kblocks + kper - 1) / kper;
      if (cudaMemset2DAsync(out, (size_t)ldo * sizeof(float), 0, (size_t)N * sizeof(float), M, s) != cudaSuccess)
        return tfail(h, B200SAM_ERR_CUDA, "train_gemm_tc: memset failed");
      ++nl;
    }
  }
  e.bias = bias;
  cudaError_t err = gemm_tc(true, GM_F32, A, (int)lda, W, (int)ldw, M, N, K, e, s, false);
  if (err != cudaSuccess) return tfail(h, B200SAM_ERR_CUDA, "train_gemm_tc: %s", cudaGetErrorString(err));
  T_DONE(nl);
}
