// Sam.postprocess_masks (sam.py:155-189) + the callers' threshold (chat.py:231, train_ds.py:662-663,
// inference.py:283-297 via logit(th)) + the evaluation counts of train_ds.py:662-709,761-796, as ONE HBM-bound kernel.
//
// The reference resamples twice with a crop in between and materialises the 1024^2 intermediate:
//   a = bilinear(m 256^2 -> 1024^2) ; a = a[:h, :w] ; out = bilinear(a -> H x W)          (both align_corners=False)
// A single direct 256^2 -> (H, W) resample is NOT equivalent, so the two resamplings are
// composed exactly, but separably and in shared memory - no 1024^2 tensor touches HBM:
//   R[q]  = oy.l0 * (y0.l0 m[y0.i0][q] + y0.l1 m[y0.i1][q]) + oy.l1 * (y1.l0 m[y1.i0][q] + y1.l1 m[y1.i1][q])
//           the (at most four) low-res rows an output row touches, collapsed along y            (256 floats / row)
//   A[x]  = tx.l0 R[tx.i0] + tx.l1 R[tx.i1]      the 1024-grid row, columns [0, w) of the crop      (w floats / row)
//   out[X] = ox.l0 A[ox.i0] + ox.l1 A[ox.i1]
// One block owns kRows consecutive output rows of one mask (both sides of a prompt in the fused-metrics mode); the
// column taps are computed once per thread and reused for every row.  Outputs, each optional: fp32 logits (what
// postprocess_masks returns), u8 {0,1} masks, and the 1-bit wire format - bit 7 of byte X/8 is pixel 8*(X/8), i.e.
// numpy.packbits(mask, axis=-1), rows padded to whole bytes: 8x fewer bytes through NVLink / PCIe than u8.
//
// ATen index rule (UpSampleBilinear2d): src = max(0, scale * (dst + 0.5) - 0.5), scale = float(in) / out,
// i0 = int(src), i1 = i0 + (i0 < in - 1), lambda = src - i0.
#include "common.cuh"
#include "kernels.cuh"

namespace bsam {

namespace {

struct Tap {
  int i0, i1;
  float l0, l1;
};

__device__ __forceinline__ Tap make_tap(int dst, float scale, int n_in) {
  float src = fmaxf(scale * (dst + 0.5f) - 0.5f, 0.f);
  int i0 = min((int)src, n_in - 1);
  Tap t;
  t.i0 = i0;
  t.i1 = i0 + (i0 < n_in - 1 ? 1 : 0);
  t.l1 = src - (float)i0;
  t.l0 = 1.f - t.l1;
  return t;
}

constexpr int kRows = 4;          // output rows per block
constexpr int kPostThreads = 256;
constexpr int kPix = 8;           // consecutive output pixels per thread = one byte of the bit format

struct PostArgs {
  const float* low[2];            // (n, 256, 256) low-res logits of side 0 / 1 (side 1 only in the 2-side modes)
  float* logits[2];               // (n, H, W) f32 or null
  uint8_t* mask[2];               // (n, H, W) u8 {0,1} or null
  uint8_t* bits[2];               // (n, H, ceil(W/8)) packed or null
  int n, h, w, H, W;
  float sy, sx, thr;
  // fused evaluation (2 sides): taxonomy-gated prediction vs ground truth
  const float* taxonomy;          // (n, 4)
  const uint8_t* gt;              // (G, H, W) u8, non-zero = set
  const int* gt_index;            // (n) or null
  unsigned long long* counts;     // (n, 3) {intersection, union, predicted area}
  uint8_t* pred_bits;             // (n, H, ceil(W/8)) gated prediction, packed, or null
};

template <int SIDES, bool METRICS>
__global__ void __launch_bounds__(kPostThreads)
postprocess_kernel(const PostArgs a) {
  __shared__ float R[SIDES][kRows][256];
  __shared__ float A[SIDES][kRows][kImg];
  const long n = blockIdx.y;
  const int Y0 = blockIdx.x * kRows;
  const int tid = threadIdx.x;
  constexpr float kUp = 256.f / 1024.f;
  // ---- rows: y-collapse of the low-res mask (one column per thread)
#pragma unroll
  for (int r = 0; r < kRows; ++r) {
    const int Y = min(Y0 + r, a.H - 1);
    const Tap oy = make_tap(Y, a.sy, a.h);
    const Tap y0 = make_tap(oy.i0, kUp, 256), y1 = make_tap(oy.i1, kUp, 256);
#pragma unroll
    for (int s = 0; s < SIDES; ++s) {
      const float* m = a.low[s] + n * 65536 + tid;
      const float u0 = fmaf(y0.l0, __ldg(m + y0.i0 * 256), y0.l1 * __ldg(m + y0.i1 * 256));
      const float u1 = fmaf(y1.l0, __ldg(m + y1.i0 * 256), y1.l1 * __ldg(m + y1.i1 * 256));
      R[s][r][tid] = fmaf(oy.l0, u0, oy.l1 * u1);
    }
  }
  __syncthreads();
  // ---- the cropped 1024-grid rows
  for (int x = tid; x < a.w; x += kPostThreads) {
    const Tap tx = make_tap(x, kUp, 256);
#pragma unroll
    for (int s = 0; s < SIDES; ++s)
#pragma unroll
      for (int r = 0; r < kRows; ++r) A[s][r][x] = fmaf(tx.l0, R[s][r][tx.i0], tx.l1 * R[s][r][tx.i1]);
  }
  __syncthreads();
  // ---- output pixels: 8 per thread and row
  int use_l = 1, use_r = 1;
  const uint8_t* G = nullptr;
  if constexpr (METRICS) {
    int k = 0;
    float best = a.taxonomy[n * 4];
#pragma unroll
    for (int i = 1; i < 4; ++i)
      if (a.taxonomy[n * 4 + i] > best) best = a.taxonomy[n * 4 + i], k = i;
    use_l = k != 1;   // train_ds.py:699-706: 0 = left, 1 = right, 2 / 3 = left | right
    use_r = k != 0;
    G = a.gt + (long)(a.gt_index ? a.gt_index[n] : n) * a.H * a.W;
  }
  unsigned inter = 0, uni = 0, area = 0;
  const int wb = (a.W + 7) >> 3;
  for (int X0 = tid * kPix; X0 < a.W; X0 += kPostThreads * kPix) {
    int i0[kPix], i1[kPix];
    float l1[kPix];
#pragma unroll
    for (int j = 0; j < kPix; ++j) {
      const Tap ox = make_tap(min(X0 + j, a.W - 1), a.sx, a.w);
      i0[j] = ox.i0, i1[j] = ox.i1, l1[j] = ox.l1;
    }
    const bool full = X0 + kPix <= a.W;
#pragma unroll
    for (int r = 0; r < kRows; ++r) {
      const int Y = Y0 + r;
      if (Y >= a.H) break;
      const long base = (n * a.H + Y) * (long)a.W + X0;
      uint32_t by[2] = {0u, 0u};
#pragma unroll
      for (int s = 0; s < SIDES; ++s) {
        float v[kPix];
#pragma unroll
        for (int j = 0; j < kPix; ++j) v[j] = fmaf(1.f - l1[j], A[s][r][i0[j]], l1[j] * A[s][r][i1[j]]);
        uint32_t b = 0;
#pragma unroll
        for (int j = 0; j < kPix; ++j) b |= (v[j] > a.thr ? 1u : 0u) << (7 - j);
        if (!full) b &= 0xFFu << (kPix - (a.W - X0));   // pixels beyond the row end stay 0 in the padded byte
        by[s] = b;
        if (a.logits[s]) {
          float* o = a.logits[s] + base;
          if (full && ((reinterpret_cast<uintptr_t>(o) & 15) == 0)) {
            reinterpret_cast<float4*>(o)[0] = make_float4(v[0], v[1], v[2], v[3]);
            reinterpret_cast<float4*>(o)[1] = make_float4(v[4], v[5], v[6], v[7]);
          } else {
#pragma unroll
            for (int j = 0; j < kPix; ++j)
              if (X0 + j < a.W) o[j] = v[j];
          }
        }
        if (a.mask[s]) {
          uint8_t* o = a.mask[s] + base;
          if (full && ((reinterpret_cast<uintptr_t>(o) & 7) == 0)) {
            uint2 t;
            t.x = ((b >> 7) & 1u) | (((b >> 6) & 1u) << 8) | (((b >> 5) & 1u) << 16) | (((b >> 4) & 1u) << 24);
            t.y = ((b >> 3) & 1u) | (((b >> 2) & 1u) << 8) | (((b >> 1) & 1u) << 16) | ((b & 1u) << 24);
            *reinterpret_cast<uint2*>(o) = t;
          } else {
#pragma unroll
            for (int j = 0; j < kPix; ++j)
              if (X0 + j < a.W) o[j] = (b >> (7 - j)) & 1u;
          }
        }
        if (a.bits[s]) a.bits[s][(n * a.H + Y) * (long)wb + (X0 >> 3)] = (uint8_t)b;
      }
      if constexpr (METRICS) {
        const uint32_t pm = (use_l ? by[0] : 0u) | (use_r ? by[1] : 0u);
        uint32_t gm = 0;
        const uint8_t* g = G + (long)Y * a.W + X0;
        if (full && ((reinterpret_cast<uintptr_t>(g) & 7) == 0)) {
          const uint2 t = __ldg(reinterpret_cast<const uint2*>(g));
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            gm |= (((t.x >> (8 * j)) & 0xFFu) ? 1u : 0u) << (7 - j);
            gm |= (((t.y >> (8 * j)) & 0xFFu) ? 1u : 0u) << (3 - j);
          }
        } else {
#pragma unroll
          for (int j = 0; j < kPix; ++j)
            if (X0 + j < a.W) gm |= (__ldg(g + j) ? 1u : 0u) << (7 - j);
        }
        inter += __popc(pm & gm);
        uni += __popc(pm | gm);
        area += __popc(pm);
        if (a.pred_bits) a.pred_bits[(n * a.H + Y) * (long)wb + (X0 >> 3)] = (uint8_t)pm;
      }
    }
  }
  if constexpr (METRICS) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      inter += __shfl_xor_sync(0xffffffffu, inter, o);
      uni += __shfl_xor_sync(0xffffffffu, uni, o);
      area += __shfl_xor_sync(0xffffffffu, area, o);
    }
    if ((tid & 31) == 0) {
      if (inter) atomicAdd(a.counts + n * 3 + 0, (unsigned long long)inter);
      if (uni) atomicAdd(a.counts + n * 3 + 1, (unsigned long long)uni);
      if (area) atomicAdd(a.counts + n * 3 + 2, (unsigned long long)area);
    }
  }
}

// 64-bit position-sensitive checksum of `n_items` byte strings of `bytes` each (stride `stride` bytes):
//   sum over 8-byte little-endian words w_i (tail zero padded) of mix(w_i ^ (i + 1) * K),  mod 2^64
// with the splitmix64 finaliser as mix.  A sum, so any number of threads may contribute in any order and the
// producing rank and the gathering rank get the same value for the same bytes (clip.segment_clip ships it beside the
// packed masks; bench.py --gpus N compares).
__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

__global__ void __launch_bounds__(256)
checksum64_kernel(const uint8_t* __restrict__ data, long bytes, long stride, unsigned long long* __restrict__ out) {
  const uint8_t* p = data + (long)blockIdx.y * stride;
  const long nwords = (bytes + 7) >> 3, nfull = bytes >> 3;
  const bool aligned = (reinterpret_cast<uintptr_t>(p) & 7) == 0;
  unsigned long long acc = 0;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nwords; i += (long)gridDim.x * blockDim.x) {
    unsigned long long w = 0;
    if (aligned && i < nfull) {
      w = __ldg(reinterpret_cast<const unsigned long long*>(p) + i);
    } else {
      for (int b = 0; b < 8; ++b)
        if (i * 8 + b < bytes) w |= (unsigned long long)p[i * 8 + b] << (8 * b);
    }
    acc += mix64(w ^ ((unsigned long long)(i + 1) * 0x9E3779B97F4A7C15ull));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0 && acc) atomicAdd(out + blockIdx.y, acc);
}

template <int SIDES, bool METRICS>
cudaError_t launch_post(const PostArgs& a, cudaStream_t s) {
  dim3 grid((a.H + kRows - 1) / kRows, a.n);
  postprocess_kernel<SIDES, METRICS><<<grid, kPostThreads, 0, s>>>(a);
  return cudaGetLastError();
}

bool sizes_ok(int n, int in_h, int in_w, int out_h, int out_w) {
  return in_h >= 1 && in_w >= 1 && in_h <= kImg && in_w <= kImg && n <= 65535 && out_h <= 65535 * kRows;
}

}  // namespace

cudaError_t launch_postprocess(const float* lowres, int n, int in_h, int in_w, int out_h, int out_w, float* logits,
                               uint8_t* mask, uint8_t* bits, float thr, cudaStream_t s) {
  if (n <= 0 || out_h <= 0 || out_w <= 0) return cudaSuccess;
  if (!sizes_ok(n, in_h, in_w, out_h, out_w)) return cudaErrorInvalidValue;
  PostArgs a = {};
  a.low[0] = lowres;
  a.logits[0] = logits;
  a.mask[0] = mask;
  a.bits[0] = bits;
  a.n = n, a.h = in_h, a.w = in_w, a.H = out_h, a.W = out_w;
  a.sy = (float)in_h / (float)out_h, a.sx = (float)in_w / (float)out_w, a.thr = thr;
  return launch_post<1, false>(a, s);
}

cudaError_t launch_postprocess_pair(const float* low_left, const float* low_right, int n, int in_h, int in_w, int out_h,
                                    int out_w, uint8_t* bits_left, uint8_t* bits_right, float thr,
                                    const float* taxonomy, const uint8_t* gt, const int* gt_index, int64_t* counts,
                                    uint8_t* pred_bits, cudaStream_t s) {
  if (n <= 0 || out_h <= 0 || out_w <= 0) return cudaSuccess;
  if (!sizes_ok(n, in_h, in_w, out_h, out_w) || !low_left || !low_right) return cudaErrorInvalidValue;
  PostArgs a = {};
  a.low[0] = low_left, a.low[1] = low_right;
  a.bits[0] = bits_left, a.bits[1] = bits_right;
  a.n = n, a.h = in_h, a.w = in_w, a.H = out_h, a.W = out_w;
  a.sy = (float)in_h / (float)out_h, a.sx = (float)in_w / (float)out_w, a.thr = thr;
  if (!gt) return launch_post<2, false>(a, s);
  if (!taxonomy || !counts) return cudaErrorInvalidValue;
  a.taxonomy = taxonomy, a.gt = gt, a.gt_index = gt_index;
  a.counts = reinterpret_cast<unsigned long long*>(counts);
  a.pred_bits = pred_bits;
  cudaError_t e = cudaMemsetAsync(counts, 0, (size_t)n * 3 * sizeof(int64_t), s);
  if (e != cudaSuccess) return e;
  return launch_post<2, true>(a, s);
}

cudaError_t launch_checksum64(const uint8_t* data, int n_items, long bytes, long stride, uint64_t* out,
                              cudaStream_t s) {
  if (n_items <= 0) return cudaSuccess;
  if (bytes < 0 || stride < bytes || n_items > 65535) return cudaErrorInvalidValue;
  cudaError_t e = cudaMemsetAsync(out, 0, (size_t)n_items * sizeof(uint64_t), s);
  if (e != cudaSuccess || bytes == 0) return e;
  const long nwords = (bytes + 7) >> 3;
  long bx = (nwords + 256 * 8 - 1) / (256 * 8);
  if (bx > 64) bx = 64;
  checksum64_kernel<<<dim3((unsigned)bx, n_items), 256, 0, s>>>(data, bytes, stride,
                                                                 reinterpret_cast<unsigned long long*>(out));
  return cudaGetLastError();
}

}  // namespace bsam
