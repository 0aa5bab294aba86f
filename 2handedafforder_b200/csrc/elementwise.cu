// Memory-bound kernels of the encoder / prompt encoder: LayerNorm (+ window partition), im2col index maps,
// layout transposes, casts, the dense positional encoding.  All are coalesced, 8/16-byte vectorised HBM traffic
// with warp-shuffle reductions.
#include "common.cuh"
#include "kernels.cuh"

namespace bsam {

namespace {

// ------------------------------------------------------------------------------------------------ LayerNorm
// Block.norm1 / norm2 (image_encoder.py:160,170,179,191; eps 1e-6 from build_sam.py:77), fp32 in, bf16 out.
// One warp per token row; the row stays in registers between the statistics and the normalise pass.
// windowed: fuse window_partition (image_encoder.py:263-288) into the store; pad rows of `out` are never written
// (they were zeroed once at workspace creation, matching F.pad after the norm).
template <int NJ>
__global__ void __launch_bounds__(256) layernorm_bf16_kernel(const float* __restrict__ x, long rows, int D,
                                                             const float* __restrict__ g,
                                                             const float* __restrict__ b, float eps,
                                                             __nv_bfloat16* __restrict__ out, int windowed) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // rows are walked from the END of the residual stream: the GEMM epilogue that produced X wrote it front to back,
  // so its last ~100 MB are still in the 126 MB L2 when this kernel starts - reading those first (before this
  // kernel's own traffic evicts them) turns most of the fp32 read into L2 hits
  const long row = rows - 1 - ((long)blockIdx.x * 8 + warp);
  if (row < 0) return;
  const int nvec = D >> 2;
  const float4* xr = reinterpret_cast<const float4*>(x + row * D);
  float4 v[NJ];
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < NJ; ++j) {
    int i = lane + 32 * j;
    if (i < nvec) {
      v[j] = xr[i];
      s += v[j].x + v[j].y + v[j].z + v[j].w;
    }
  }
  const float mean = warp_sum(s) / D;
  float q = 0.f;
#pragma unroll
  for (int j = 0; j < NJ; ++j) {
    int i = lane + 32 * j;
    if (i < nvec) {
      float a = v[j].x - mean, c = v[j].y - mean, d = v[j].z - mean, e = v[j].w - mean;
      q += a * a + c * c + d * d + e * e;
    }
  }
  const float rstd = rsqrtf(warp_sum(q) / D + eps);
  long orow = row;
  if (windowed) {
    long bi = row / kTokens;
    int t = (int)(row - bi * kTokens);
    int y = t / kGrid, xx = t % kGrid;
    orow = bi * kWinRowsPerFrame + ((y / kWin) * kWinPerSide + xx / kWin) * kWinTokens + (y % kWin) * kWin + xx % kWin;
  }
  uint2* o = reinterpret_cast<uint2*>(out + orow * D);
  const float4* g4 = reinterpret_cast<const float4*>(g);
  const float4* b4 = reinterpret_cast<const float4*>(b);
#pragma unroll
  for (int j = 0; j < NJ; ++j) {
    int i = lane + 32 * j;
    if (i < nvec) {
      float4 gg = __ldg(g4 + i), bb = __ldg(b4 + i);
      uint2 w;
      w.x = pack_bf16((v[j].x - mean) * rstd * gg.x + bb.x, (v[j].y - mean) * rstd * gg.y + bb.y);
      w.y = pack_bf16((v[j].z - mean) * rstd * gg.z + bb.z, (v[j].w - mean) * rstd * gg.w + bb.w);
      o[i] = w;
    }
  }
}

// ------------------------------------------------------------------------------------------------ im2col maps
// PatchEmbed Conv2d(k16,s16) (image_encoder.py:395-426): non-overlapping patches => im2col is a pure index map.
// out[(b, py, px)][c*256 + ky*16 + kx] = frame[b][c][py*16+ky][px*16+kx], cast to bf16.
// One thread moves 8 consecutive kx (reads are fully coalesced along the image row).
template <typename T>
__global__ void __launch_bounds__(256) im2col_patch_kernel(const T* __restrict__ frames, long total,
                                                           __nv_bfloat16* __restrict__ out) {
  long idx = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  // idx enumerates 8-pixel chunks of the NCHW frame in memory order
  const int xc = (int)(idx % (kImg / 8));
  long r = idx / (kImg / 8);
  const int y = (int)(r % kImg);
  r /= kImg;
  const int c = (int)(r % 3);
  const long b = r / 3;
  const T* src = frames + ((b * 3 + c) * kImg + y) * (long)kImg + xc * 8;
  float f[8];
  if constexpr (sizeof(T) == 4) {
    float4 a = reinterpret_cast<const float4*>(src)[0], d = reinterpret_cast<const float4*>(src)[1];
    f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = d.x; f[5] = d.y; f[6] = d.z; f[7] = d.w;
  } else {
    uint4 a = reinterpret_cast<const uint4*>(src)[0];
    const T* h = reinterpret_cast<const T*>(&a);
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] = static_cast<float>(h[j]);
  }
  const int py = y / kPatch, ky = y % kPatch, px = (xc * 8) / kPatch, kx = (xc * 8) % kPatch;
  uint4 w;
  w.x = pack_bf16(f[0], f[1]); w.y = pack_bf16(f[2], f[3]); w.z = pack_bf16(f[4], f[5]); w.w = pack_bf16(f[6], f[7]);
  __nv_bfloat16* dst = out + ((b * kTokens + py * kGrid + px) * (long)(3 * kPatch * kPatch)) + c * 256 + ky * 16 + kx;
  *reinterpret_cast<uint4*>(dst) = w;
}

// neck Conv2d 3x3 pad 1 (image_encoder.py:100-106) as a 9-tap implicit GEMM: K order (tap = dy*3+dx, c).
// z: (B,64,64,256) bf16 token-major.  One thread moves 8 channels of one tap.
__global__ void __launch_bounds__(256) im2col_3x3_kernel(const __nv_bfloat16* __restrict__ z, long total,
                                                         __nv_bfloat16* __restrict__ out) {
  long idx = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int c8 = (int)(idx % 32);
  long r = idx / 32;
  const int tap = (int)(r % 9);
  r /= 9;                                    // r = b*4096 + y*64 + x
  const int t = (int)(r % kTokens);
  const long b = r / kTokens;
  const int y = t / kGrid + tap / 3 - 1, x = t % kGrid + tap % 3 - 1;
  uint4 v = make_uint4(0, 0, 0, 0);
  if (y >= 0 && y < kGrid && x >= 0 && x < kGrid)
    v = *reinterpret_cast<const uint4*>(z + ((b * kTokens + y * kGrid + x) * (long)kOutCh) + c8 * 8);
  *reinterpret_cast<uint4*>(out + r * (9L * kOutCh) + tap * kOutCh + c8 * 8) = v;
}

// ------------------------------------------------------------------------------------------------ layout
// (B, 4096, 256) fp32 token-major  <->  (B, 256, 64, 64) NCHW in f32 / bf16 / f16 (image_encoder.py:124 permute).
template <typename T, bool kToNchw>
__global__ void __launch_bounds__(256) emb_transpose_kernel(const void* __restrict__ in_, void* __restrict__ out_) {
  __shared__ float tile[32][33];
  const long b = blockIdx.z;
  const int t0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8
  if constexpr (kToNchw) {
    const float* in = reinterpret_cast<const float*>(in_) + b * kTokens * kOutCh;
    T* out = reinterpret_cast<T*>(out_) + b * kTokens * kOutCh;
#pragma unroll
    for (int j = 0; j < 4; ++j) tile[ty + 8 * j][tx] = in[(long)(t0 + ty + 8 * j) * kOutCh + c0 + tx];
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 4; ++j) out[(long)(c0 + ty + 8 * j) * kTokens + t0 + tx] = static_cast<T>(tile[tx][ty + 8 * j]);
  } else {
    const T* in = reinterpret_cast<const T*>(in_) + b * kTokens * kOutCh;
    float* out = reinterpret_cast<float*>(out_) + b * kTokens * kOutCh;
#pragma unroll
    for (int j = 0; j < 4; ++j) tile[ty + 8 * j][tx] = static_cast<float>(in[(long)(c0 + ty + 8 * j) * kTokens + t0 + tx]);
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 4; ++j) out[(long)(t0 + ty + 8 * j) * kOutCh + c0 + tx] = tile[tx][ty + 8 * j];
  }
}

__global__ void cast_f32_bf16_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, long n) {
  long i = ((long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i + 3 < n) {
    float4 v = *reinterpret_cast<const float4*>(in + i);
    uint2 w;
    w.x = pack_bf16(v.x, v.y);
    w.y = pack_bf16(v.z, v.w);
    *reinterpret_cast<uint2*>(out + i) = w;
  } else {
    for (; i < n; ++i) out[i] = __float2bfloat16(in[i]);
  }
}

template <typename T>
__global__ void to_f32_kernel(const T* __restrict__ in, float* __restrict__ out, long n) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = static_cast<float>(in[i]);
}

// PositionEmbeddingRandom._pe_encoding + forward (prompt_encoder.py:203-229), token-major (4096, 256):
// [sin(2*pi*(c @ G)), cos(...)], c = 2*((x+.5)/64, (y+.5)/64) - 1.  Computed once at weight finalisation in fp32.
__global__ void dense_pe_kernel(const float* __restrict__ gauss, float* __restrict__ pe) {
  const int t = blockIdx.x, c = threadIdx.x;   // 4096 x 128
  const int y = t / kGrid, x = t % kGrid;
  const float cx = 2.f * ((x + 0.5f) / kGrid) - 1.f, cy = 2.f * ((y + 0.5f) / kGrid) - 1.f;
  const float ang = 6.283185307179586f * (cx * gauss[c] + cy * gauss[128 + c]);
  pe[(long)t * kOutCh + c] = sinf(ang);
  pe[(long)t * kOutCh + 128 + c] = cosf(ang);
}

// MaskDecoder.predict_masks src = image_embeddings (repeat per prompt) + dense_prompt_embeddings
// (mask_decoder.py:143-144), where dense = no_mask_embed broadcast (prompt_encoder.py:182-184).
__global__ void __launch_bounds__(256) keys_init_kernel(const float* __restrict__ emb_tok,
                                                        const float* __restrict__ no_mask,
                                                        const int* __restrict__ prompt_frame, float* __restrict__ keys) {
  const int p = blockIdx.y;
  const long i = ((long)blockIdx.x * blockDim.x + threadIdx.x) * 4;   // over 4096*256
  const float4 e = *reinterpret_cast<const float4*>(emb_tok + (long)prompt_frame[p] * kTokens * kOutCh + i);
  const float4 m = *reinterpret_cast<const float4*>(no_mask + (i % kOutCh));
  *reinterpret_cast<float4*>(keys + (long)p * kTokens * kOutCh + i) = make_float4(e.x + m.x, e.y + m.y, e.z + m.z, e.w + m.w);
}

__global__ void relpos_pack_kernel(const float* __restrict__ rel, int nrows, int hd, int RP,
                                   __nv_bfloat16* __restrict__ out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= RP * hd) return;
  int r = i / hd;
  out[i] = __float2bfloat16(r < nrows ? rel[i] : 0.f);
}

// (256, 128) bf16: rows 0..126 = rel_pos_h, 128..254 = rel_pos_w, everything else (incl. columns >= hd) zero
__global__ void relpos_pack_g_kernel(const float* __restrict__ rel_h, const float* __restrict__ rel_w, int hd,
                                     __nv_bfloat16* __restrict__ out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 256 * 128) return;
  int r = i >> 7, c = i & 127;
  float v = 0.f;
  if (c < hd) {
    if (r < 127) v = rel_h[r * hd + c];
    else if (r >= 128 && r < 255) v = rel_w[(254 - r) * hd + c];   // w table reversed: row 128 + j = Rw[126 - j]
  }
  out[i] = __float2bfloat16(v);
}

}  // namespace

cudaError_t launch_layernorm_bf16(const float* x, long rows, int D, const float* g, const float* b, float eps,
                                  __nv_bfloat16* out, bool windowed, cudaStream_t s) {
  if (rows <= 0) return cudaSuccess;
  if (D % 4 != 0 || D > 1280) return cudaErrorInvalidValue;
  const int grid = (int)((rows + 7) / 8);
  const int nj = (D / 4 + 31) / 32;
  if (nj <= 2) layernorm_bf16_kernel<2><<<grid, 256, 0, s>>>(x, rows, D, g, b, eps, out, windowed);
  else if (nj <= 3) layernorm_bf16_kernel<3><<<grid, 256, 0, s>>>(x, rows, D, g, b, eps, out, windowed);
  else if (nj <= 6) layernorm_bf16_kernel<6><<<grid, 256, 0, s>>>(x, rows, D, g, b, eps, out, windowed);
  else if (nj <= 8) layernorm_bf16_kernel<8><<<grid, 256, 0, s>>>(x, rows, D, g, b, eps, out, windowed);
  else layernorm_bf16_kernel<10><<<grid, 256, 0, s>>>(x, rows, D, g, b, eps, out, windowed);
  return cudaGetLastError();
}

// One block per (window-pad row, frame): pad_rows lists the 804 pad positions of the 4900-row window layout.
__global__ void __launch_bounds__(128) fill_pad_rows_kernel(__nv_bfloat16* __restrict__ qkv,
                                                            const float* __restrict__ bias,
                                                            const int* __restrict__ pad_rows, int cols) {
  griddep_launch_dependents();
  griddep_wait();     // PDL: keeps the chain qkv GEMM -> fill -> attention transitive (the attention waits only on this grid)
  const int rem = pad_rows[blockIdx.x];
  uint2* o = reinterpret_cast<uint2*>(qkv + ((long)blockIdx.y * kWinRowsPerFrame + rem) * cols);
  const float4* b4 = reinterpret_cast<const float4*>(bias);
  for (int i = threadIdx.x; i < cols / 4; i += 128) {
    const float4 v = __ldg(b4 + i);
    uint2 w;
    w.x = pack_bf16(v.x, v.y);
    w.y = pack_bf16(v.z, v.w);
    o[i] = w;
  }
}

cudaError_t launch_fill_pad_rows(__nv_bfloat16* qkv, const float* bias, const int* pad_rows, int n_pad, int B,
                                 int cols, cudaStream_t s) {
  if (B <= 0 || n_pad <= 0) return cudaSuccess;
  if (cols % 4 != 0) return cudaErrorInvalidValue;
  return launch_pdl(fill_pad_rows_kernel, dim3(n_pad, B), dim3(128), 0, s, qkv, bias, pad_rows, cols);
}

cudaError_t launch_im2col_patch(const void* frames, int dtype, int B, __nv_bfloat16* out, cudaStream_t s) {
  if (B <= 0) return cudaSuccess;
  const long total = (long)B * 3 * kImg * (kImg / 8);
  const int grid = (int)((total + 255) / 256);
  if (dtype == DT_F32) im2col_patch_kernel<float><<<grid, 256, 0, s>>>(reinterpret_cast<const float*>(frames), total, out);
  else if (dtype == DT_BF16) im2col_patch_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>(reinterpret_cast<const __nv_bfloat16*>(frames), total, out);
  else if (dtype == DT_F16) im2col_patch_kernel<__half><<<grid, 256, 0, s>>>(reinterpret_cast<const __half*>(frames), total, out);
  else return cudaErrorInvalidValue;
  return cudaGetLastError();
}

cudaError_t launch_im2col_3x3(const __nv_bfloat16* z, int B, __nv_bfloat16* out, cudaStream_t s) {
  if (B <= 0) return cudaSuccess;
  const long total = (long)B * kTokens * 9 * 32;
  im2col_3x3_kernel<<<(int)((total + 255) / 256), 256, 0, s>>>(z, total, out);
  return cudaGetLastError();
}

cudaError_t launch_tokens_to_nchw(const float* tok, int B, void* out, int dtype, cudaStream_t s) {
  if (B <= 0) return cudaSuccess;
  dim3 grid(kTokens / 32, kOutCh / 32, B);
  if (dtype == DT_F32) emb_transpose_kernel<float, true><<<grid, 256, 0, s>>>(tok, out);
  else if (dtype == DT_BF16) emb_transpose_kernel<__nv_bfloat16, true><<<grid, 256, 0, s>>>(tok, out);
  else if (dtype == DT_F16) emb_transpose_kernel<__half, true><<<grid, 256, 0, s>>>(tok, out);
  else return cudaErrorInvalidValue;
  return cudaGetLastError();
}

cudaError_t launch_nchw_to_tokens(const void* in, int dtype, int B, float* tok, cudaStream_t s) {
  if (B <= 0) return cudaSuccess;
  dim3 grid(kTokens / 32, kOutCh / 32, B);
  if (dtype == DT_F32) emb_transpose_kernel<float, false><<<grid, 256, 0, s>>>(in, tok);
  else if (dtype == DT_BF16) emb_transpose_kernel<__nv_bfloat16, false><<<grid, 256, 0, s>>>(in, tok);
  else if (dtype == DT_F16) emb_transpose_kernel<__half, false><<<grid, 256, 0, s>>>(in, tok);
  else return cudaErrorInvalidValue;
  return cudaGetLastError();
}

// LayerNorm folded into the consuming Linear (gemm_tc.cuh, GemmEpi::stats_in):  LN(x) W^T + b =
// rstd * (x (gamma . W)^T) - rstd * mean * c + d.  One block per output feature n: Wf[n, :] = bf16(gamma . W[n, :]),
// c[n] = sum_k float(Wf[n, k]) (the ROUNDED weights, so that a constant row still cancels exactly), d[n] = W[n, :] . beta + b[n].
static __global__ void __launch_bounds__(128) fold_ln_weights_kernel(const float* __restrict__ Wt, const float* __restrict__ gamma,
                                                              const float* __restrict__ beta, const float* __restrict__ bias,
                                                              int K, __nv_bfloat16* __restrict__ Wf, float* __restrict__ c,
                                                              float* __restrict__ d) {
  __shared__ float red[2][4];
  const long n = blockIdx.x;
  float sc = 0.f, sd = 0.f;
  for (int k = threadIdx.x; k < K; k += 128) {
    const float w = Wt[n * K + k];
    const __nv_bfloat16 wf = __float2bfloat16_rn(w * gamma[k]);
    Wf[n * K + k] = wf;
    sc += __bfloat162float(wf);
    sd += w * beta[k];
  }
  sc = warp_sum(sc);
  sd = warp_sum(sd);
  if ((threadIdx.x & 31) == 0) red[0][threadIdx.x >> 5] = sc, red[1][threadIdx.x >> 5] = sd;
  __syncthreads();
  if (threadIdx.x == 0) {
    c[n] = (red[0][0] + red[0][1]) + (red[0][2] + red[0][3]);
    d[n] = (red[1][0] + red[1][1]) + (red[1][2] + red[1][3]) + bias[n];
  }
}

cudaError_t launch_fold_ln_weights(const float* Wt, const float* gamma, const float* beta, const float* bias, int N, int K,
                                   __nv_bfloat16* Wf, float* c, float* d, cudaStream_t s) {
  fold_ln_weights_kernel<<<N, 128, 0, s>>>(Wt, gamma, beta, bias, K, Wf, c, d);
  return cudaGetLastError();
}

cudaError_t launch_cast_f32_bf16(const float* in, __nv_bfloat16* out, long n, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  const long nthreads = (n + 3) / 4;
  cast_f32_bf16_kernel<<<(int)((nthreads + 255) / 256), 256, 0, s>>>(in, out, n);
  return cudaGetLastError();
}

cudaError_t launch_to_f32(const void* in, int dtype, float* out, long n, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  const int grid = (int)((n + 255) / 256);
  if (dtype == DT_F32) return cudaMemcpyAsync(out, in, n * 4, cudaMemcpyDeviceToDevice, s);
  if (dtype == DT_BF16) to_f32_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>(reinterpret_cast<const __nv_bfloat16*>(in), out, n);
  else if (dtype == DT_F16) to_f32_kernel<__half><<<grid, 256, 0, s>>>(reinterpret_cast<const __half*>(in), out, n);
  else return cudaErrorInvalidValue;
  return cudaGetLastError();
}

cudaError_t launch_dense_pe(const float* gauss, float* pe_tok, cudaStream_t s) {
  dense_pe_kernel<<<kTokens, 128, 0, s>>>(gauss, pe_tok);
  return cudaGetLastError();
}

cudaError_t launch_keys_init(const float* emb_tok, const float* no_mask, const int* prompt_frame, int P, float* keys,
                             cudaStream_t s) {
  if (P <= 0) return cudaSuccess;
  dim3 grid(kTokens * kOutCh / 4 / 256, P);
  keys_init_kernel<<<grid, 256, 0, s>>>(emb_tok, no_mask, prompt_frame, keys);
  return cudaGetLastError();
}

cudaError_t launch_relpos_pack(const float* rel, int S, int hd, __nv_bfloat16* out, cudaStream_t s) {
  const int RP = (S == 14) ? 32 : 128;
  relpos_pack_kernel<<<(RP * hd + 255) / 256, 256, 0, s>>>(rel, 2 * S - 1, hd, RP, out);
  return cudaGetLastError();
}

cudaError_t launch_relpos_pack_g(const float* rel_h, const float* rel_w, int hd, __nv_bfloat16* out, cudaStream_t s) {
  relpos_pack_g_kernel<<<(256 * 128 + 255) / 256, 256, 0, s>>>(rel_h, rel_w, hd, out);
  return cudaGetLastError();
}

}  // namespace bsam
