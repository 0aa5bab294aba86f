// Sam.postprocess_masks (sam.py:155-189) + the callers' threshold (chat.py:231, train_ds.py:662-663,
// inference.py:283-297 via logit(th)), as ONE HBM-bound kernel.
//
// The reference resamples twice with a crop in between and materialises the 1024^2 intermediate:
//   a = bilinear(m 256^2 -> 1024^2) ; a = a[:h, :w] ; out = bilinear(a -> H x W)          (both align_corners=False)
// A single direct 256^2 -> (H, W) resample is NOT equivalent, so every output pixel
// composes the two resamplings exactly (four 1024-grid taps, each from four low-res taps); the row direction is
// collapsed once per output row in shared memory.  One coalesced 16-byte store per 4 output pixels; no 1024^2
// tensor touches HBM.
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

// Both resamplings are separable, so the row direction is collapsed first: for one output row Y the (at most four)
// low-res rows it touches are blended into ONE 256-float row in shared memory,
//   R[q] = sum_a oy.l[a] * sum_p ty_a.l[p] * m[ty_a.i[p]][q],
// and every output pixel is then 4 shared-memory reads: out = sum_b ox.l[b] * sum_q tx_b.l[q] * R[tx_b.i[q]].
// grid: (ceil(W/4/128), H, n).  One thread = 4 consecutive output pixels of one row.
__global__ void __launch_bounds__(128)
postprocess_kernel(const float* __restrict__ lowres, int h, int w, int H, int W, float sy, float sx,
                   float* __restrict__ logits, uint8_t* __restrict__ mask, float thr) {
  __shared__ float R[256];
  const int Y = blockIdx.y;
  const long n = blockIdx.z;
  const float* m = lowres + n * 65536;
  constexpr float kUp = 256.f / 1024.f;
  {
    const Tap oy = make_tap(Y, sy, h);
    const Tap y0 = make_tap(oy.i0, kUp, 256), y1 = make_tap(oy.i1, kUp, 256);
    const float* r00 = m + y0.i0 * 256;
    const float* r01 = m + y0.i1 * 256;
    const float* r10 = m + y1.i0 * 256;
    const float* r11 = m + y1.i1 * 256;
#pragma unroll
    for (int q = threadIdx.x; q < 256; q += 128)
      R[q] = oy.l0 * (y0.l0 * __ldg(r00 + q) + y0.l1 * __ldg(r01 + q)) +
             oy.l1 * (y1.l0 * __ldg(r10 + q) + y1.l1 * __ldg(r11 + q));
  }
  __syncthreads();
  const int X0 = (blockIdx.x * 128 + threadIdx.x) * 4;
  if (X0 >= W) return;
  float v[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    v[j] = 0.f;
    if (X0 + j < W) {
      const Tap ox = make_tap(X0 + j, sx, w);
      const Tap x0 = make_tap(ox.i0, kUp, 256), x1 = make_tap(ox.i1, kUp, 256);
      v[j] = ox.l0 * (x0.l0 * R[x0.i0] + x0.l1 * R[x0.i1]) + ox.l1 * (x1.l0 * R[x1.i0] + x1.l1 * R[x1.i1]);
    }
  }
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
