// Sam.postprocess_masks (sam.py:155-189) + the callers' threshold (chat.py:231, train_ds.py:662-663,
// inference.py:283-297 via logit(th)), as ONE HBM-bound kernel.
//
// The reference resamples twice with a crop in between and materialises the 1024^2 intermediate:
//   a = bilinear(m 256^2 -> 1024^2) ; a = a[:h, :w] ; out = bilinear(a -> H x W)          (both align_corners=False)
// A single direct 256^2 -> (H, W) resample is NOT equivalent, so every output pixel
// evaluates the four 1024-grid taps it needs, each from four low-res taps, in registers: 16 cached low-res reads
// (a 256 KB mask stays in L1/L2) and one coalesced 16-byte store per 4 output pixels.  No 1024^2 tensor touches HBM.
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

// value of the virtual 1024^2 intermediate at (y, x): bilinear from the 256^2 low-res mask
__device__ __forceinline__ float mid_value(const float* __restrict__ m, const Tap& ty, const Tap& tx) {
  const float* r0 = m + ty.i0 * 256;
  const float* r1 = m + ty.i1 * 256;
  const float top = tx.l0 * __ldg(r0 + tx.i0) + tx.l1 * __ldg(r0 + tx.i1);
  const float bot = tx.l0 * __ldg(r1 + tx.i0) + tx.l1 * __ldg(r1 + tx.i1);
  return ty.l0 * top + ty.l1 * bot;
}

__device__ __forceinline__ float out_value(const float* __restrict__ m, int Y, int X, float sy, float sx, int h,
                                           int w) {
  const Tap oy = make_tap(Y, sy, h), ox = make_tap(X, sx, w);
  constexpr float kUp = 256.f / 1024.f;
  const Tap y0 = make_tap(oy.i0, kUp, 256), y1 = make_tap(oy.i1, kUp, 256);
  const Tap x0 = make_tap(ox.i0, kUp, 256), x1 = make_tap(ox.i1, kUp, 256);
  const float top = ox.l0 * mid_value(m, y0, x0) + ox.l1 * mid_value(m, y0, x1);
  const float bot = ox.l0 * mid_value(m, y1, x0) + ox.l1 * mid_value(m, y1, x1);
  return oy.l0 * top + oy.l1 * bot;
}

// grid: (ceil(W/4/128), H, n).  One thread = 4 consecutive output pixels of one row.
__global__ void __launch_bounds__(128)
postprocess_kernel(const float* __restrict__ lowres, int h, int w, int H, int W, float sy, float sx,
                   float* __restrict__ logits, uint8_t* __restrict__ mask, float thr) {
  const int X0 = (blockIdx.x * 128 + threadIdx.x) * 4;
  if (X0 >= W) return;
  const int Y = blockIdx.y;
  const long n = blockIdx.z;
  const float* m = lowres + n * 65536;
  float v[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) v[j] = (X0 + j < W) ? out_value(m, Y, X0 + j, sy, sx, h, w) : 0.f;
  const long base = (n * H + Y) * (long)W + X0;
  const bool full = X0 + 3 < W;
  if (logits) {
    float* o = logits + base;
    if (full && ((reinterpret_cast<uintptr_t>(o) & 15) == 0)) {
      *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (X0 + j < W) o[j] = v[j];
    }
  }
  if (mask) {
    uint8_t* o = mask + base;
    if (full && ((reinterpret_cast<uintptr_t>(o) & 3) == 0)) {
      uchar4 b = make_uchar4(v[0] > thr, v[1] > thr, v[2] > thr, v[3] > thr);
      *reinterpret_cast<uchar4*>(o) = b;
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (X0 + j < W) o[j] = v[j] > thr;
    }
  }
}

}  // namespace

cudaError_t launch_postprocess(const float* lowres, int n, int in_h, int in_w, int out_h, int out_w, float* logits,
                               uint8_t* mask, float thr, cudaStream_t s) {
  if (n <= 0 || out_h <= 0 || out_w <= 0) return cudaSuccess;
  if (in_h < 1 || in_w < 1 || in_h > kImg || in_w > kImg || out_h > 65535 || n > 65535) return cudaErrorInvalidValue;
  dim3 grid((out_w + 511) / 512, out_h, n);
  postprocess_kernel<<<grid, 128, 0, s>>>(lowres, in_h, in_w, out_h, out_w, (float)in_h / (float)out_h,
                                           (float)in_w / (float)out_w, logits, mask, thr);
  return cudaGetLastError();
}

}  // namespace bsam
