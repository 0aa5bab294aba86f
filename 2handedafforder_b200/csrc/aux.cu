// The callers either side of the SAM branch, as HBM-bound integer / byte kernels:
//
//   preprocess_kernel   raw uint8 RGB frame -> ResizeLongestSide.apply_image (transforms.py:27-34, = Pillow's 8-bit
//                       two-pass antialiased bilinear resample) -> (x - mean) / std -> zero pad to 1024^2 -> model
//                       dtype (inference.py:91-105,251-256).  Bit-exact against the reference: the resample is
//                       integer fixed point with the coefficient tables built on the host in double precision exactly
//                       as Pillow builds them; the normalisation is one IEEE fp32 subtract and one IEEE fp32 divide.
//   seg_scan_kernel /   `[SEG]` rows of the LLM output (LISA.py:457-476) and text_hidden_fcs = Linear-ReLU-Linear
//   rows_fc_kernel      (LISA.py:95-101) evaluated ONLY on those rows (the reference projects every row, then gathers).
//   mask_metrics_kernel threshold-mask evaluation of train_ds.py:662-709,761-796: taxonomy-gated left / right / OR mask
//                       and the intersection / union / area pixel counts behind IoU and IoCM, so validate() never
//                       copies full-resolution masks to the host.
#include <math.h>

#include <algorithm>
#include <vector>

#include "aux.cuh"
#include "common.cuh"

namespace bsam {

// ------------------------------------------------------------------------------------------------ resample tables

// Pillow precompute_coeffs + normalize_coeffs_8bpc for the bilinear (triangle, support 1) filter over the whole
// axis.  Double precision, same operation order, so the 22-bit fixed-point weights are identical.
void build_resample_axis(int in_size, int out_size, ResampleAxis* ax) {
  const double scale = (double)in_size / out_size;
  const double filterscale = scale < 1.0 ? 1.0 : scale;
  const double support = 1.0 * filterscale;
  const int ksize = (int)ceil(support) * 2 + 1;
  ax->in_size = in_size;
  ax->out_size = out_size;
  ax->ksize = ksize;
  ax->identity = in_size == out_size;
  ax->lo.assign(out_size, 0);
  ax->cnt.assign(out_size, 0);
  ax->kk.assign((size_t)out_size * ksize, 0);
  std::vector<double> w(ksize);
  const double ss = 1.0 / filterscale;
  for (int xx = 0; xx < out_size; ++xx) {
    const double center = (xx + 0.5) * scale;
    int lo = (int)(center - support + 0.5);
    if (lo < 0) lo = 0;
    int hi = (int)(center + support + 0.5);
    if (hi > in_size) hi = in_size;
    const int n = hi - lo;
    double ww = 0.0;
    for (int x = 0; x < n; ++x) {
      double t = (x + lo - center + 0.5) * ss;
      if (t < 0.0) t = -t;
      w[x] = t < 1.0 ? 1.0 - t : 0.0;
      ww += w[x];
    }
    for (int x = 0; x < n; ++x)
      if (ww != 0.0) w[x] /= ww;
    for (int x = 0; x < n; ++x) {
      const double v = w[x] * (double)(1 << kResamplePrecisionBits);
      ax->kk[(size_t)xx * ksize + x] = w[x] < 0 ? (int)(-0.5 + v) : (int)(0.5 + v);
    }
    ax->lo[xx] = lo;
    ax->cnt[xx] = n;
  }
}

namespace {

constexpr int kTileW = 128, kTileH = 8, kPreThreads = 256;

__device__ __forceinline__ uint8_t clip8(int acc) {
  const int v = acc >> kResamplePrecisionBits;      // arithmetic shift, as Pillow's clip8 lookup index
  return (uint8_t)min(max(v, 0), 255);
}

template <typename T>
__device__ __forceinline__ void store4(T* p, float a, float b, float c, float d);
template <>
__device__ __forceinline__ void store4<float>(float* p, float a, float b, float c, float d) {
  *reinterpret_cast<float4*>(p) = make_float4(a, b, c, d);
}
template <>
__device__ __forceinline__ void store4<__nv_bfloat16>(__nv_bfloat16* p, float a, float b, float c, float d) {
  uint2 v;
  v.x = (uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn(a)) | ((uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn(b)) << 16);
  v.y = (uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn(c)) | ((uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn(d)) << 16);
  *reinterpret_cast<uint2*>(p) = v;
}
template <>
__device__ __forceinline__ void store4<__half>(__half* p, float a, float b, float c, float d) {
  uint2 v;
  v.x = (uint32_t)__half_as_ushort(__float2half_rn(a)) | ((uint32_t)__half_as_ushort(__float2half_rn(b)) << 16);
  v.y = (uint32_t)__half_as_ushort(__float2half_rn(c)) | ((uint32_t)__half_as_ushort(__float2half_rn(d)) << 16);
  *reinterpret_cast<uint2*>(p) = v;
}

// Horizontal pass of one (output column, channel) down `nrows` input rows with an N-tap window held in registers.
template <int N>
__device__ __forceinline__ void hpass_col(const uint8_t* __restrict__ col, const int* __restrict__ k,
                                          uint8_t* __restrict__ dst, int nrows, int W) {
  int kk[N];
#pragma unroll
  for (int j = 0; j < N; ++j) kk[j] = k[j];
  const long pitch = (long)W * 3;
  int r = 0;
  for (; r + 1 < nrows; r += 2) {                  // two rows per trip: 2N independent loads in flight
    const uint8_t* r0p = col + r * pitch;
    const uint8_t* r1p = r0p + pitch;
    int a0 = 1 << (kResamplePrecisionBits - 1), a1 = a0;
#pragma unroll
    for (int j = 0; j < N; ++j) {
      a0 += (int)r0p[j * 3] * kk[j];
      a1 += (int)r1p[j * 3] * kk[j];
    }
    dst[r * kTileW * 3] = clip8(a0);
    dst[(r + 1) * kTileW * 3] = clip8(a1);
  }
  if (r < nrows) {
    const uint8_t* r0p = col + r * pitch;
    int a0 = 1 << (kResamplePrecisionBits - 1);
#pragma unroll
    for (int j = 0; j < N; ++j) a0 += (int)r0p[j * 3] * kk[j];
    dst[r * kTileW * 3] = clip8(a0);
  }
}

struct PreArgs {
  const uint8_t* in;     // (B, H, W, 3)
  int H, W, new_h, new_w;
  const int *xlo, *xcnt, *xkk;   // horizontal tables (new_w entries), ksize kx
  const int *ylo, *ycnt, *ykk;   // vertical tables (new_h entries), ksize ky
  int kx, ky, ident_x, ident_y;
  float mean[3], stdv[3];
};

// grid (1024/128, 1024/8, B), 256 threads.  One block = one 128 x 8 tile of the padded 1024^2 output, 3 channels.
// Stage 1: horizontal pass of the input rows this tile's vertical windows touch -> uint8 tile in shared memory
// (Pillow rounds to uint8 between the passes).  Stage 2: vertical pass, normalise, 4 pixels per thread, one 8/16-byte
// store each; tiles in the pad region store zeros.
template <typename T>
__global__ void __launch_bounds__(kPreThreads) preprocess_kernel(PreArgs a, T* __restrict__ out) {
  extern __shared__ uint8_t tile[];                 // (rows, kTileW, 3)
  const int x0 = blockIdx.x * kTileW, y0 = blockIdx.y * kTileH;
  const long b = blockIdx.z;
  const int tw = min(kTileW, a.new_w - x0), th = min(kTileH, a.new_h - y0);   // live part of the tile (may be <= 0)
  int r0 = 0, nrows = 0;
  if (tw > 0 && th > 0) {
    if (a.ident_y) {
      r0 = y0;
      nrows = th;
    } else {
      r0 = a.ylo[y0];
      nrows = a.ylo[y0 + th - 1] + a.ycnt[y0 + th - 1] - r0;
    }
    const uint8_t* src = a.in + b * (long)a.H * a.W * 3;
    // one thread = one (output column, channel) of the tile; its window and fixed-point weights stay in registers
    // while it walks down the input rows (adjacent threads read adjacent bytes of the interleaved RGB row)
    for (int i = threadIdx.x; i < tw * 3; i += kPreThreads) {
      const int x = i / 3, c = i - x * 3;
      uint8_t* dst = tile + x * 3 + c;
      if (a.ident_x) {
        const uint8_t* col = src + (long)r0 * a.W * 3 + (x0 + x) * 3 + c;
        for (int r = 0; r < nrows; ++r) dst[r * kTileW * 3] = col[(long)r * a.W * 3];
      } else {
        const int lo = a.xlo[x0 + x], n = a.xcnt[x0 + x];
        const int* k = a.xkk + (long)(x0 + x) * a.kx;
        const uint8_t* col = src + (long)r0 * a.W * 3 + lo * 3 + c;
        if (n <= 8) {
          switch (n) {
            case 1: hpass_col<1>(col, k, dst, nrows, a.W); break;
            case 2: hpass_col<2>(col, k, dst, nrows, a.W); break;
            case 3: hpass_col<3>(col, k, dst, nrows, a.W); break;
            case 4: hpass_col<4>(col, k, dst, nrows, a.W); break;
            case 5: hpass_col<5>(col, k, dst, nrows, a.W); break;
            case 6: hpass_col<6>(col, k, dst, nrows, a.W); break;
            case 7: hpass_col<7>(col, k, dst, nrows, a.W); break;
            default: hpass_col<8>(col, k, dst, nrows, a.W); break;
          }
        } else {
          for (int r = 0; r < nrows; ++r) {
            const uint8_t* row = col + (long)r * a.W * 3;
            int acc = 1 << (kResamplePrecisionBits - 1);
            for (int j = 0; j < n; ++j) acc += (int)row[j * 3] * k[j];
            dst[r * kTileW * 3] = clip8(acc);
          }
        }
      }
    }
  }
  __syncthreads();
  // stage 2: items = 3 channels x 8 rows x 32 groups of 4 pixels
  for (int i = threadIdx.x; i < 3 * kTileH * (kTileW / 4); i += kPreThreads) {
    const int g = i % (kTileW / 4), y = (i / (kTileW / 4)) % kTileH, c = i / (kTileH * (kTileW / 4));
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    if (y < th) {
      const int Y = y0 + y;
      int lo = Y - r0, n = 1;
      const int* k = nullptr;
      if (!a.ident_y) {
        lo = a.ylo[Y] - r0;
        n = a.ycnt[Y];
        k = a.ykk + (long)Y * a.ky;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int x = g * 4 + j;
        if (x < tw) {
          int u;
          if (a.ident_y) {
            u = tile[(lo * kTileW + x) * 3 + c];
          } else {
            int acc = 1 << (kResamplePrecisionBits - 1);
            for (int t = 0; t < n; ++t) acc += (int)tile[((lo + t) * kTileW + x) * 3 + c] * k[t];
            u = clip8(acc);
          }
          v[j] = __fdiv_rn(__fsub_rn((float)u, a.mean[c]), a.stdv[c]);   // inference.py:99
        }
      }
    }
    T* o = out + ((b * 3 + c) * (long)kImg + (y0 + y)) * kImg + x0 + g * 4;
    store4<T>(o, v[0], v[1], v[2], v[3]);
  }
}


// ---- second-generation kernel (the shipped one for horizontal windows of up to 10 taps, i.e. frames up to 4608 px) ----
// Same arithmetic, restructured around what bounded the first kernel (one-byte global loads walked row by row at
// 11 % of HBM): a block = one 128 x 16 tile of the 1024^2 output, 3 channels, in three phases.
//   A  the input rows the tile's vertical windows touch, restricted to the byte span its horizontal windows touch, are
//      copied to shared memory with 16-byte loads (each row keeps its own 0..15-byte alignment phase);
//   B  horizontal pass from shared memory: one thread = one output column, all three channels; its window (3 * taps
//      interleaved RGB bytes) is fetched as NW 32-bit words, re-aligned with funnel shifts and unpacked in registers;
//      the window start and the fixed-point weights stay in registers while the thread walks down the rows.  Output:
//      channel-planar uint8 tile (Pillow rounds to uint8 between the passes);
//   C  vertical pass: one thread = 8 consecutive pixels of one channel and output row (two 32-bit loads per tap), then
//      (x - mean) / std through a 3 x 256-entry table built per block with the same IEEE subtract / divide / convert as
//      before (a uint8 has 256 values: identical bits, no division per pixel), one 16-byte (bf16 / f16) or two (f32) stores.
constexpr int kP2W = 128, kP2H = 16;

// Pillow's clip8 without the clamp: the triangle filter's fixed-point weights are >= 0 and sum to 2^22 +- (taps / 2)
// (normalize_coeffs_8bpc rounds each to nearest), so 2^21 <= acc <= 255 * (2^22 + taps) + 2^21 < 256 * 2^22 and the
// shifted value is already in [0, 255]  (255 * taps < 2^21 for any window this kernel accepts).
__device__ __forceinline__ uint32_t shr8(int acc) { return (uint32_t)acc >> kResamplePrecisionBits; }

struct Pre2Args {
  PreArgs p;
  int rows_max;      // most input rows any 16-row output tile touches
  int raw_pitch;     // bytes per staged input row (multiple of 16; covers phase + span + window over-read)
  long in_bytes;     // B * H * W * 3
  const float* lut;  // [3][256] (u - mean[c]) / std[c] in fp32, built once on the host (preprocess_lut)
};

template <typename T> __device__ __forceinline__ T cvt_out(float v);
template <> __device__ __forceinline__ float cvt_out<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 cvt_out<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ __half cvt_out<__half>(float v) { return __float2half_rn(v); }

template <typename T>
__device__ __forceinline__ void store8(T* o, const T (&v)[8]) {
  if constexpr (sizeof(T) == 4) {
    reinterpret_cast<float4*>(o)[0] = make_float4(v[0], v[1], v[2], v[3]);
    reinterpret_cast<float4*>(o)[1] = make_float4(v[4], v[5], v[6], v[7]);
  } else {
    *reinterpret_cast<uint4*>(o) = *reinterpret_cast<const uint4*>(v);
  }
}

// shared-memory layout: [lut 3*256 T][rshift rows_max int, padded to 16 B][raw rows_max * raw_pitch][hb 3 * rows_max * 128]
// [ytab 16 * (2 + ky) int]
__device__ __forceinline__ uint8_t* hb_end(uint8_t* sm, const Pre2Args& q, size_t esz) {
  return sm + 3 * 256 * esz + ((q.rows_max * 4 + 15) & ~15) + (size_t)q.rows_max * q.raw_pitch +
         (size_t)3 * q.rows_max * kP2W;
}

template <typename T, int NPX>     // NPX: taps of the horizontal window the instantiation can hold (3, 5, 8 or 10)
__global__ void __launch_bounds__(kPreThreads) preprocess2_kernel(Pre2Args q, T* __restrict__ out) {
  constexpr int NV = (3 * NPX + 3) / 4;     // aligned 32-bit words holding the 3 * NPX window bytes
  extern __shared__ __align__(16) uint8_t sm2[];
  const PreArgs& a = q.p;
  const int x0 = blockIdx.x * kP2W, y0 = blockIdx.y * kP2H;
  const long b = blockIdx.z;
  const int tw = min(kP2W, a.new_w - x0), th = min(kP2H, a.new_h - y0);
  const int tid = threadIdx.x;
  if (tw <= 0 || th <= 0) {                 // tile entirely in the zero pad
    __align__(16) T z[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) z[j] = cvt_out<T>(0.f);
    for (int i = tid; i < 3 * kP2H * (kP2W / 8); i += kPreThreads) {
      const int g = i & 15, y = (i >> 4) & 15, c = i >> 8;
      store8<T>(out + ((b * 3 + c) * (long)kImg + (y0 + y)) * kImg + x0 + g * 8, z);
    }
    return;
  }
  T* lut = reinterpret_cast<T*>(sm2);                               // [3][256]
  int* rshift = reinterpret_cast<int*>(sm2 + 3 * 256 * sizeof(T));  // [rows_max] alignment phase of each staged row
  uint8_t* raw = sm2 + 3 * 256 * sizeof(T) + ((q.rows_max * 4 + 15) & ~15);
  uint8_t* hb = raw + (size_t)q.rows_max * q.raw_pitch;             // [3][rows_max][128]
  for (int i = tid; i < 3 * 256; i += kPreThreads) lut[i] = cvt_out<T>(__ldg(q.lut + i));   // inference.py:99
  // horizontal window of this thread's output column: independent of the staged rows, so the table loads overlap phase A
  const int bx = tid & (kP2W - 1);
  int bn = 0, bob = 0;
  int bk[NPX];
#pragma unroll
  for (int j = 0; j < NPX; ++j) bk[j] = 0;
  if (bx < tw) {
    const int X = x0 + bx;
    bn = a.xcnt[X];
    bob = a.xlo[X];
#pragma unroll
    for (int j = 0; j < NPX; ++j) bk[j] = j < bn ? __ldg(a.xkk + (long)X * a.kx + j) : 0;
  }
  // vertical windows of the tile's 16 output rows -> shared memory (phase C reads them 3 * 16 times each)
  int* ytab = reinterpret_cast<int*>(hb_end(sm2, q, sizeof(T)));     // [16][2 + ky]: lo - r0, n, weights
  const int r0 = a.ylo[y0];
  for (int i = tid; i < kP2H * (2 + a.ky); i += kPreThreads) {
    const int y = i / (2 + a.ky), f = i - y * (2 + a.ky);
    int v = 0;
    if (y < th) v = f == 0 ? a.ylo[y0 + y] - r0 : f == 1 ? a.ycnt[y0 + y] : __ldg(a.ykk + (long)(y0 + y) * a.ky + (f - 2));
    ytab[i] = v;
  }
  const int nrows = a.ylo[y0 + th - 1] + a.ycnt[y0 + th - 1] - r0;
  const int px0 = a.xlo[x0];
  const int nbytes = (a.xlo[x0 + tw - 1] + a.xcnt[x0 + tw - 1] - px0) * 3;
  // ---- phase A: one warp per staged row, one 16-byte vector per lane and trip
  {
    const uint8_t* frame = a.in + b * (long)a.H * a.W * 3 + (long)px0 * 3;
    const uint8_t* lo_ok = a.in;
    const uint8_t* hi_ok = a.in + q.in_bytes;
    const int warp = tid >> 5, lane = tid & 31;
    for (int r = warp; r < nrows; r += kPreThreads / 32) {
      const uint8_t* src = frame + (long)(r0 + r) * a.W * 3;
      const int sh = (int)(reinterpret_cast<uintptr_t>(src) & 15);
      if (lane == 0) rshift[r] = sh;
      const uint8_t* base = src - sh;
      uint8_t* dst = raw + (size_t)r * q.raw_pitch;
      const int nvec = (sh + nbytes + 15) >> 4;
      if (base >= lo_ok && base + nvec * 16 <= hi_ok) {      // whole staged row inside the caller's buffer (warp-uniform)
        for (int v = lane; v < nvec; v += 32)
          *reinterpret_cast<uint4*>(dst + v * 16) = __ldg(reinterpret_cast<const uint4*>(base + v * 16));
      } else {                                               // first / last row of the buffer: byte by byte at the ends
        for (int v = lane; v < nvec; v += 32) {
          const uint8_t* p = base + v * 16;
          __align__(16) uint8_t tmp[16];
#pragma unroll
          for (int jj = 0; jj < 16; ++jj) tmp[jj] = (p + jj >= lo_ok && p + jj < hi_ok) ? p[jj] : (uint8_t)0;
          *reinterpret_cast<uint4*>(dst + v * 16) = *reinterpret_cast<const uint4*>(tmp);
        }
      }
    }
  }
  __syncthreads();
  // ---- phase B
  {
    const int x = bx, par = tid >> 7;
    if (x < tw) {
      const int ob = (bob - px0) * 3;
      const int (&k)[NPX] = bk;
      const uint8_t* rowp = raw + (size_t)par * q.raw_pitch;
      uint8_t* hp = hb + par * kP2W + x;
      const int rstep = (kPreThreads / kP2W) * q.raw_pitch, plane = q.rows_max * kP2W;
      for (int r = par; r < nrows; r += kPreThreads / kP2W, rowp += rstep, hp += (kPreThreads / kP2W) * kP2W) {
        const int o = ob + rshift[r];
        const uint32_t* wp = reinterpret_cast<const uint32_t*>(rowp + (o & ~3));
        const uint32_t sh = (uint32_t)(o & 3) * 8;
        uint32_t w[NV + 1], v[NV];
#pragma unroll
        for (int j = 0; j <= NV; ++j) w[j] = wp[j];
#pragma unroll
        for (int j = 0; j < NV; ++j) v[j] = __funnelshift_r(w[j], w[j + 1], sh);
        int acc[3] = {1 << (kResamplePrecisionBits - 1), 1 << (kResamplePrecisionBits - 1),
                      1 << (kResamplePrecisionBits - 1)};
#pragma unroll
        for (int j = 0; j < 3 * NPX; ++j) acc[j % 3] += (int)__byte_perm(v[j >> 2], 0u, 0x4440u + (j & 3)) * k[j / 3];
        hp[0] = shr8(acc[0]);
        hp[plane] = shr8(acc[1]);
        hp[2 * plane] = shr8(acc[2]);
      }
    }
  }
  __syncthreads();
  // ---- phase C
  for (int i = tid; i < 3 * kP2H * (kP2W / 8); i += kPreThreads) {
    const int g = i & 15, y = (i >> 4) & 15, c = i >> 8;
    __align__(16) T o8[8];
    if (y < th && g * 8 < tw) {
      const int* yt = ytab + y * (2 + a.ky);
      const int lo = yt[0], n = yt[1];
      int acc[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = 1 << (kResamplePrecisionBits - 1);
      const uint8_t* col = hb + ((size_t)c * q.rows_max + lo) * kP2W + g * 8;
      for (int t = 0; t < n; ++t) {
        const int kt = yt[2 + t];
        const uint2 wv = *reinterpret_cast<const uint2*>(col + (size_t)t * kP2W);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          acc[j] += (int)__byte_perm(wv.x, 0u, 0x4440u + j) * kt;
          acc[4 + j] += (int)__byte_perm(wv.y, 0u, 0x4440u + j) * kt;
        }
      }
      const T* lc = lut + c * 256;
#pragma unroll
      for (int j = 0; j < 8; ++j) o8[j] = lc[shr8(acc[j])];
      if (g * 8 + 8 > tw) {                       // the one group that straddles new_w: columns right of it are zero pad
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (g * 8 + j >= tw) o8[j] = cvt_out<T>(0.f);
      }
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) o8[j] = cvt_out<T>(0.f);
    }
    store8<T>(out + ((b * 3 + c) * (long)kImg + (y0 + y)) * kImg + x0 + g * 8, o8);
  }
}

}  // namespace

// (u - mean[c]) / std[c] for every uint8 value and channel: one IEEE fp32 subtract and one IEEE fp32 divide, exactly what
// the per-pixel form computes (inference.py:93-99), so a table lookup gives identical bits
void preprocess_lut(float* out768) {
  const float mean[3] = {123.675f, 116.28f, 103.53f}, stdv[3] = {58.395f, 57.12f, 57.375f};
  for (int c = 0; c < 3; ++c)
    for (int u = 0; u < 256; ++u) {
      volatile float d = (float)u - mean[c];
      volatile float v = d / stdv[c];
      out768[c * 256 + u] = v;
    }
}

void preprocess2_extents(const ResampleAxis& ax, const ResampleAxis& ay, int new_w, int new_h, int* span_px,
                         int* rows) {
  int ws = 0, wr = 0;
  for (int x0 = 0; x0 < new_w; x0 += kP2W) {
    const int x1 = (x0 + kP2W < new_w ? x0 + kP2W : new_w) - 1;
    ws = std::max(ws, ax.lo[x1] + ax.cnt[x1] - ax.lo[x0]);
  }
  for (int y0 = 0; y0 < new_h; y0 += kP2H) {
    const int y1 = (y0 + kP2H < new_h ? y0 + kP2H : new_h) - 1;
    wr = std::max(wr, ay.lo[y1] + ay.cnt[y1] - ay.lo[y0]);
  }
  *span_px = ws;
  *rows = wr;
}

int preprocess_tile_rows(const ResampleAxis& ay, int new_h) {
  int worst = 0;
  for (int y0 = 0; y0 < new_h; y0 += kTileH) {
    const int y1 = (y0 + kTileH < new_h ? y0 + kTileH : new_h) - 1;
    const int rows = ay.identity ? (y1 - y0 + 1) : (ay.lo[y1] + ay.cnt[y1] - ay.lo[y0]);
    if (rows > worst) worst = rows;
  }
  return worst;
}

cudaError_t launch_preprocess(const uint8_t* frames, int B, int H, int W, const ResampleTables& t, void* out,
                              int out_dtype, cudaStream_t s) {
  if (B <= 0) return cudaSuccess;
  PreArgs a;
  a.in = frames;
  a.H = H;
  a.W = W;
  a.new_h = t.new_h;
  a.new_w = t.new_w;
  a.xlo = t.d_xlo;
  a.xcnt = t.d_xcnt;
  a.xkk = t.d_xkk;
  a.ylo = t.d_ylo;
  a.ycnt = t.d_ycnt;
  a.ykk = t.d_ykk;
  a.kx = t.kx;
  a.ky = t.ky;
  a.ident_x = t.ident_x;
  a.ident_y = t.ident_y;
  const float mean[3] = {123.675f, 116.28f, 103.53f}, stdv[3] = {58.395f, 57.12f, 57.375f};   // inference.py:93-94
  for (int c = 0; c < 3; ++c) a.mean[c] = mean[c], a.stdv[c] = stdv[c];
  // second-generation kernel whenever its window registers and shared-memory stage can hold the resample
  if (kernel_variant(VK_PREPROCESS) == 0 && t.kx <= 10 && t.span_px128 > 0 && t.rows16 > 0 && t.d_lut) {
    Pre2Args q;
    q.p = a;
    q.rows_max = t.rows16;
    q.raw_pitch = (15 + t.span_px128 * 3 + 15) & ~15;
    q.in_bytes = (long)B * H * W * 3;
    const size_t esz = out_dtype == DT_F32 ? 4 : 2;
    const size_t smem2 = 3 * 256 * esz + (((size_t)q.rows_max * 4 + 15) & ~(size_t)15) +
                         (size_t)q.rows_max * q.raw_pitch + (size_t)3 * q.rows_max * kP2W +
                         (size_t)kP2H * (2 + t.ky) * sizeof(int) + 64;
    q.lut = t.d_lut;
    if (smem2 <= 200 * 1024) {
      dim3 grid2(kImg / kP2W, kImg / kP2H, B);
      cudaError_t e2 = cudaSuccess;
#define BSAM_PRE2(T, NPX)                                                                                          \
  do {                                                                                                             \
    e2 = cudaFuncSetAttribute(preprocess2_kernel<T, NPX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2); \
    if (e2 != cudaSuccess) return e2;                                                                              \
    preprocess2_kernel<T, NPX><<<grid2, kPreThreads, smem2, s>>>(q, reinterpret_cast<T*>(out));                     \
  } while (0)
#define BSAM_PRE2_T(T)                 \
  do {                                 \
    if (t.kx <= 3) BSAM_PRE2(T, 3);    \
    else if (t.kx <= 5) BSAM_PRE2(T, 5); \
    else if (t.kx <= 8) BSAM_PRE2(T, 8); \
    else BSAM_PRE2(T, 10);             \
  } while (0)
      if (out_dtype == DT_F32) BSAM_PRE2_T(float);
      else if (out_dtype == DT_BF16) BSAM_PRE2_T(__nv_bfloat16);
      else if (out_dtype == DT_F16) BSAM_PRE2_T(__half);
      else return cudaErrorInvalidValue;
#undef BSAM_PRE2_T
#undef BSAM_PRE2
      return cudaGetLastError();
    }
  }
  const size_t smem = (size_t)t.tile_rows * kTileW * 3;
  dim3 grid(kImg / kTileW, kImg / kTileH, B);
  cudaError_t e = cudaSuccess;
  if (out_dtype == DT_F32) {
    if (smem > 48 * 1024) e = cudaFuncSetAttribute(preprocess_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    preprocess_kernel<float><<<grid, kPreThreads, smem, s>>>(a, reinterpret_cast<float*>(out));
  } else if (out_dtype == DT_BF16) {
    if (smem > 48 * 1024) e = cudaFuncSetAttribute(preprocess_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    preprocess_kernel<__nv_bfloat16><<<grid, kPreThreads, smem, s>>>(a, reinterpret_cast<__nv_bfloat16*>(out));
  } else if (out_dtype == DT_F16) {
    if (smem > 48 * 1024) e = cudaFuncSetAttribute(preprocess_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    preprocess_kernel<__half><<<grid, kPreThreads, smem, s>>>(a, reinterpret_cast<__half*>(out));
  } else {
    return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ [SEG] rows

namespace {

// One block.  LISA.py:457-476: mask = output_ids[:, 1:] == seg_idx, shifted right by 255 columns (the single image
// token stands for 256 CLIP patches); rows are collected in row-major order.  counts: (b + 1) ints = prompts per
// image, then the total; rows: flat row index (bi * L + 255 + t - 1) of every [SEG] token, at most max_rows
// (a larger total is reported through counts[b] and the caller raises).
__global__ void __launch_bounds__(1024) seg_scan_kernel(const int64_t* __restrict__ ids, int b, int T, int L,
                                                         int64_t seg_idx, int* __restrict__ rows, int max_rows,
                                                         int* __restrict__ counts) {
  __shared__ int s_cnt[1024];
  __shared__ int s_off[1025];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarp = blockDim.x >> 5;
  for (int bi = warp; bi < b; bi += nwarp) {      // pass 1: count per image
    int c = 0;
    for (int t0 = 1; t0 < T; t0 += 32) {
      const int t = t0 + lane;
      c += __popc(__ballot_sync(0xffffffffu, t < T && ids[(long)bi * T + t] == seg_idx));
    }
    if (lane == 0) s_cnt[bi] = c;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int acc = 0;
    for (int bi = 0; bi < b; ++bi) {
      s_off[bi] = acc;
      acc += s_cnt[bi];
      counts[bi] = s_cnt[bi];
    }
    s_off[b] = acc;
    counts[b] = acc;
  }
  __syncthreads();
  for (int bi = warp; bi < b; bi += nwarp) {      // pass 2: ordered compaction
    int base = s_off[bi];
    for (int t0 = 1; t0 < T; t0 += 32) {
      const int t = t0 + lane;
      const bool hit = t < T && ids[(long)bi * T + t] == seg_idx;
      const unsigned m = __ballot_sync(0xffffffffu, hit);
      const int pos = base + __popc(m & ((1u << lane) - 1u));
      if (hit && pos < max_rows) rows[pos] = bi * L + 255 + (t - 1);
      base += __popc(m);
    }
  }
}

constexpr int kFcRows = 8, kFcWarps = 16;

__device__ __forceinline__ void load8(const float* p, float (&w)[8]) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  w[0] = a.x, w[1] = a.y, w[2] = a.z, w[3] = a.w, w[4] = b.x, w[5] = b.y, w[6] = b.z, w[7] = b.w;
}
__device__ __forceinline__ void load8(const __nv_bfloat16* p, float (&w)[8]) {
  const uint4 v = __ldg(reinterpret_cast<const uint4*>(p));
  const uint32_t u[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    w[2 * i] = __uint_as_float(u[i] << 16);
    w[2 * i + 1] = __uint_as_float(u[i] & 0xffff0000u);
  }
}
__device__ __forceinline__ float ld_scalar(const float* p, long i) { return __ldg(p + i); }
__device__ __forceinline__ float ld_scalar(const __nv_bfloat16* p, long i) { return __bfloat162float(p[i]); }

// y[p, j] = act(bias[j] + sum_k x[row(p), k] * Wt[j, k]) for the (device-side) n = *n_ptr gathered rows.
// grid (ctas, ceil(max_rows / 8)); block = 16 warps; the 8 input rows of the chunk are staged once in shared memory as
// fp32 (8 * K * 4 bytes), each warp then streams whole weight rows with 16-byte loads (the kernel is bound by the
// one read of W: K * N * sizeof(W) bytes) and reduces 8 dot products by warp shuffle.
template <typename WT>
__global__ void __launch_bounds__(kFcWarps * 32)
rows_fc_kernel(const void* __restrict__ x, int x_dtype, long x_ld, const int* __restrict__ rows,
               const int* __restrict__ n_ptr, int max_rows, const WT* __restrict__ Wt, const WT* __restrict__ bias,
               int K, int N, int relu, float* __restrict__ y) {
  extern __shared__ float xs[];                    // (8, K)
  const int n = min(*n_ptr, max_rows);
  const int p0 = blockIdx.y * kFcRows;
  if (p0 >= n) return;
  const int np = min(kFcRows, n - p0);
  for (int i = threadIdx.x; i < kFcRows * K; i += blockDim.x) {
    const int r = i / K, k = i - r * K;
    float v = 0.f;
    if (r < np) {
      const long row = rows ? rows[p0 + r] : (p0 + r);
      const long off = row * x_ld + k;
      if (x_dtype == DT_F32) v = reinterpret_cast<const float*>(x)[off];
      else if (x_dtype == DT_BF16) v = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(x)[off]);
      else v = __half2float(reinterpret_cast<const __half*>(x)[off]);
    }
    xs[i] = v;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int j = blockIdx.x * kFcWarps + warp; j < N; j += gridDim.x * kFcWarps) {
    float acc[kFcRows];
#pragma unroll
    for (int r = 0; r < kFcRows; ++r) acc[r] = 0.f;
    const WT* wrow = Wt + (long)j * K;
#pragma unroll 4                                   // 4 x 16-byte weight loads in flight per lane
    for (int k = lane * 8; k < K; k += 256) {
      float w[8];
      load8(wrow + k, w);
#pragma unroll
      for (int r = 0; r < kFcRows; ++r) {
        const float4 a = *reinterpret_cast<const float4*>(xs + r * K + k);
        const float4 c = *reinterpret_cast<const float4*>(xs + r * K + k + 4);
        acc[r] += w[0] * a.x + w[1] * a.y + w[2] * a.z + w[3] * a.w + w[4] * c.x + w[5] * c.y + w[6] * c.z + w[7] * c.w;
      }
    }
    const float bj = ld_scalar(bias, j);
#pragma unroll
    for (int r = 0; r < kFcRows; ++r) {
      float v = warp_sum(acc[r]);
      if (lane == 0 && r < np) {
        v += bj;
        if (relu) v = fmaxf(v, 0.f);
        y[(long)(p0 + r) * N + j] = v;
      }
    }
  }
}

}  // namespace

cudaError_t launch_seg_scan(const int64_t* ids, int b, int T, int L, int64_t seg_idx, int* rows, int max_rows,
                            int* counts, cudaStream_t s) {
  if (b <= 0 || b > 1024) return cudaErrorInvalidValue;
  seg_scan_kernel<<<1, 1024, 0, s>>>(ids, b, T, L, seg_idx, rows, max_rows, counts);
  return cudaGetLastError();
}

cudaError_t launch_rows_fc(const void* x, int x_dtype, long x_ld, const int* rows, const int* n_ptr, int max_rows,
                           const void* Wt, const void* bias, int w_dtype, int K, int N, bool relu, float* y,
                           int n_sm, cudaStream_t s) {
  if (K % 8 != 0 || max_rows <= 0) return cudaErrorInvalidValue;
  const size_t smem = (size_t)kFcRows * K * sizeof(float);
  if (smem > 227 * 1024) return cudaErrorInvalidValue;
  const int ctas = (N + kFcWarps - 1) / kFcWarps < n_sm ? (N + kFcWarps - 1) / kFcWarps : n_sm;
  dim3 grid(ctas, (max_rows + kFcRows - 1) / kFcRows);
  cudaError_t e;
  if (w_dtype == DT_F32) {
    e = cudaFuncSetAttribute(rows_fc_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    rows_fc_kernel<float><<<grid, kFcWarps * 32, smem, s>>>(x, x_dtype, x_ld, rows, n_ptr, max_rows,
                                                            reinterpret_cast<const float*>(Wt),
                                                            reinterpret_cast<const float*>(bias), K, N, relu, y);
  } else if (w_dtype == DT_BF16) {
    e = cudaFuncSetAttribute(rows_fc_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    rows_fc_kernel<__nv_bfloat16><<<grid, kFcWarps * 32, smem, s>>>(
        x, x_dtype, x_ld, rows, n_ptr, max_rows, reinterpret_cast<const __nv_bfloat16*>(Wt),
        reinterpret_cast<const __nv_bfloat16*>(bias), K, N, relu, y);
  } else {
    return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ evaluation metrics

namespace {

// bit j*8 of the result = byte j of v is non-zero
__device__ __forceinline__ uint32_t nz_bytes(uint32_t v) { return __vcmpne4(v, 0u) & 0x01010101u; }

// grid (ctas_per_prompt, P).  Per prompt p: k = argmax(taxonomy[p]) (first maximum, like torch.argmax);
// pred = left (k == 0) | right (k == 1) | left OR right (otherwise)   [train_ds.py:690-706];
// counts[p] = {|pred & gt|, |pred | gt|, |pred|}                       [train_ds.py:761-796];
// pred_out (optional) = the gated mask as {0, 255} bytes.  16 bytes of each mask per thread per iteration.
__global__ void __launch_bounds__(256)
mask_metrics_kernel(const uint8_t* __restrict__ left, const uint8_t* __restrict__ right,
                    const float* __restrict__ taxonomy, const uint8_t* __restrict__ gt,
                    const int* __restrict__ gt_index, long HW, unsigned long long* __restrict__ counts,
                    uint8_t* __restrict__ pred_out, int vec_ok) {
  const long p = blockIdx.y;
  int k = 0;
  {
    float best = taxonomy[p * 4];
#pragma unroll
    for (int i = 1; i < 4; ++i)
      if (taxonomy[p * 4 + i] > best) best = taxonomy[p * 4 + i], k = i;
  }
  const uint32_t use_l = k != 1 ? 0xffffffffu : 0u, use_r = k != 0 ? 0xffffffffu : 0u;
  const uint8_t* L = left + p * HW;
  const uint8_t* R = right + p * HW;
  const uint8_t* G = gt + (long)(gt_index ? gt_index[p] : p) * HW;
  uint8_t* O = pred_out ? pred_out + p * HW : nullptr;
  unsigned inter = 0, uni = 0, area = 0;
  const long tid = (long)blockIdx.x * blockDim.x + threadIdx.x, nth = (long)gridDim.x * blockDim.x;
  if (vec_ok) {
    const long nvec = HW / 16;
    for (long i = tid; i < nvec; i += nth) {
      const uint4 l = __ldg(reinterpret_cast<const uint4*>(L) + i), r = __ldg(reinterpret_cast<const uint4*>(R) + i);
      const uint4 g = __ldg(reinterpret_cast<const uint4*>(G) + i);
      const uint32_t lw[4] = {l.x, l.y, l.z, l.w}, rw[4] = {r.x, r.y, r.z, r.w}, gw[4] = {g.x, g.y, g.z, g.w};
      uint32_t ow[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const uint32_t pm = (nz_bytes(lw[j]) & use_l) | (nz_bytes(rw[j]) & use_r), gm = nz_bytes(gw[j]);
        inter += __popc(pm & gm);
        uni += __popc(pm | gm);
        area += __popc(pm);
        ow[j] = pm * 255u;
      }
      if (O) reinterpret_cast<uint4*>(O)[i] = make_uint4(ow[0], ow[1], ow[2], ow[3]);
    }
    for (long i = nvec * 16 + tid; i < HW; i += nth) {      // tail
      const bool pm = (L[i] != 0 && use_l) || (R[i] != 0 && use_r), gm = G[i] != 0;
      inter += pm && gm, uni += pm || gm, area += pm;
      if (O) O[i] = pm ? 255 : 0;
    }
  } else {
    for (long i = tid; i < HW; i += nth) {
      const bool pm = (L[i] != 0 && use_l) || (R[i] != 0 && use_r), gm = G[i] != 0;
      inter += pm && gm, uni += pm || gm, area += pm;
      if (O) O[i] = pm ? 255 : 0;
    }
  }
  // block reduction: warp shuffles, then one atomic per warp and counter
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    inter += __shfl_xor_sync(0xffffffffu, inter, o);
    uni += __shfl_xor_sync(0xffffffffu, uni, o);
    area += __shfl_xor_sync(0xffffffffu, area, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (inter) atomicAdd(counts + p * 3 + 0, (unsigned long long)inter);
    if (uni) atomicAdd(counts + p * 3 + 1, (unsigned long long)uni);
    if (area) atomicAdd(counts + p * 3 + 2, (unsigned long long)area);
  }
}

}  // namespace

cudaError_t launch_mask_metrics(const uint8_t* left, const uint8_t* right, const float* taxonomy, const uint8_t* gt,
                                const int* gt_index, int P, int H, int W, int64_t* counts, uint8_t* pred_out,
                                int n_sm, cudaStream_t s) {
  if (P <= 0) return cudaSuccess;
  if (P > 65535) return cudaErrorInvalidValue;
  const long HW = (long)H * W;
  cudaError_t e = cudaMemsetAsync(counts, 0, sizeof(int64_t) * 3 * P, s);
  if (e != cudaSuccess) return e;
  auto al = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  const int vec_ok = HW % 16 == 0 && al(left) && al(right) && al(gt) && (!pred_out || al(pred_out));
  long per = (HW / 16 + 255) / 256;                      // blocks that would give one vector per thread
  long want = (8L * n_sm + P - 1) / P;                   // ~8 resident blocks per SM over the whole grid
  int ctas = (int)(per < want ? (per > 0 ? per : 1) : want);
  dim3 grid(ctas, P);
  mask_metrics_kernel<<<grid, 256, 0, s>>>(left, right, taxonomy, gt, gt_index, HW,
                                           reinterpret_cast<unsigned long long*>(counts), pred_out, vec_ok);
  return cudaGetLastError();
}

}  // namespace bsam

// ------------------------------------------------------------------------------------------------ no-allocation variant
// b200sam_seg_project_noalloc: b200sam_seg_project for callers that cannot call cudaMalloc (code under CUDA graph
// capture, tight serving loops, harnesses that forbid it).  No counterpart in the reference.
// The scratch b200sam_seg_project keeps in handle-owned cudaMalloc buffers ([SEG] row indices,
// per-image counts, the ReLU(Linear1) rows) lives in ONE file-scope device array of fixed size instead: every GPU that
// loads the library holds its ~1.26 MB until exit.
//
// In host code the array's NAME is a host-side stand-in, not a device address: the device address is resolved once per
// device (with that device current) through cudaGetSymbolAddress and only that pointer is handed to the kernels.
//
// Every call on a device shares the array, which keeps the previous call's contents.  Nothing in it has to start at
// zero (the scan overwrites all b + 1 counts and every row index it later reads; the first FC overwrites every hidden
// row the second one reads), but the counts are cleared anyway so that a failed launch cannot hand the previous call's
// totals to the caller.  Two calls that overlap on the GPU (different streams, or concurrent host threads) RACE on it.
#include <atomic>

#include "b200sam.h"
#include "model.cuh"

namespace bsam {
namespace {

constexpr int kNoAllocMaxRows = 64;        // [SEG] prompts per call
constexpr int kNoAllocMaxHidden = 5120;    // LLM hidden size (LLaVA-7B 4096, LLaVA-13B 5120)
constexpr int kNoAllocMaxImages = 1024;    // images per call (same limit as b200sam_seg_project)
constexpr size_t kNoAllocCountsOff = 0;                                                   // (1024 + 1) ints
constexpr size_t kNoAllocRowsOff = ((kNoAllocMaxImages + 1) * sizeof(int) + 255) / 256 * 256;
constexpr size_t kNoAllocHiddenOff = kNoAllocRowsOff + (kNoAllocMaxRows * sizeof(int) + 255) / 256 * 256;
constexpr size_t kNoAllocBytes = kNoAllocHiddenOff + (size_t)kNoAllocMaxRows * kNoAllocMaxHidden * sizeof(float);

static __device__ __align__(256) unsigned char g_seg_noalloc_scratch[kNoAllocBytes];

constexpr int kNoAllocMaxDevices = 64;
std::atomic<unsigned char*> g_seg_noalloc_ptr[kNoAllocMaxDevices];   // device address of the array, per device ordinal

int na_fail(b200sam_t* h, int code, const char* msg) {
  if (h) h->err = msg;
  return code;
}

}  // namespace
}  // namespace bsam

extern "C" int b200sam_seg_project_noalloc(b200sam_t* h, const void* hidden, int hidden_dtype, int b, int L, int hdim,
                                           const int64_t* output_ids, int T, int64_t seg_token_idx, const void* w1,
                                           const void* b1, const void* w2, const void* b2, int w_dtype,
                                           float* emb_out, int max_rows, int* counts_out, void* stream) {
  using namespace bsam;
  if (!h) return B200SAM_ERR_INVALID;
  if (max_rows > kNoAllocMaxRows || hdim > kNoAllocMaxHidden || b > kNoAllocMaxImages) {
    char buf[256];
    snprintf(buf, sizeof buf,
             "seg_project_noalloc: max_rows=%d / hdim=%d / b=%d exceed the static scratch (max_rows <= %d, hdim <= %d, "
             "b <= %d); use b200sam_seg_project",
             max_rows, hdim, b, kNoAllocMaxRows, kNoAllocMaxHidden, kNoAllocMaxImages);
    return na_fail(h, B200SAM_ERR_INVALID, buf);
  }
  if (b < 1 || T < 1 || hdim < 8 || hdim % 8 != 0 || max_rows < 1)
    return na_fail(h, B200SAM_ERR_INVALID, "seg_project_noalloc: bad sizes");
  if (L < 255 + T - 1) return na_fail(h, B200SAM_ERR_INVALID, "seg_project_noalloc: hidden length < 255 + T - 1");
  if (hidden_dtype < 0 || hidden_dtype > 2 || (w_dtype != B200SAM_F32 && w_dtype != B200SAM_BF16))
    return na_fail(h, B200SAM_ERR_INVALID, "seg_project_noalloc: bad dtypes");
  if (!hidden || !output_ids || !w1 || !b1 || !w2 || !b2 || !emb_out || !counts_out)
    return na_fail(h, B200SAM_ERR_INVALID, "seg_project_noalloc: NULL argument");
  const int dev = h->device;
  if (dev < 0 || dev >= kNoAllocMaxDevices) return na_fail(h, B200SAM_ERR_INVALID, "seg_project_noalloc: device ordinal");
  int prev = -1;
  if (cudaGetDevice(&prev) != cudaSuccess) return na_fail(h, B200SAM_ERR_CUDA, "seg_project_noalloc: cudaGetDevice");
  if (prev != dev && cudaSetDevice(dev) != cudaSuccess)
    return na_fail(h, B200SAM_ERR_CUDA, "seg_project_noalloc: cudaSetDevice");
  int rc = 0;
  do {
    unsigned char* base = g_seg_noalloc_ptr[dev].load(std::memory_order_acquire);
    if (!base) {   // set-up, once per device, with that device current: resolve the array's DEVICE address
      void* p = nullptr;
      const cudaError_t e = cudaGetSymbolAddress(&p, g_seg_noalloc_scratch);
      if (e != cudaSuccess || !p) {
        rc = na_fail(h, B200SAM_ERR_CUDA, cudaGetErrorString(e));
        break;
      }
      base = static_cast<unsigned char*>(p);
      g_seg_noalloc_ptr[dev].store(base, std::memory_order_release);
    }
    int* counts = reinterpret_cast<int*>(base + kNoAllocCountsOff);
    int* rows = reinterpret_cast<int*>(base + kNoAllocRowsOff);
    float* hid = reinterpret_cast<float*>(base + kNoAllocHiddenOff);
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    cudaError_t e = cudaMemsetAsync(counts, 0, sizeof(int) * (b + 1), s);
    if (e == cudaSuccess) e = launch_seg_scan(output_ids, b, T, L, seg_token_idx, rows, max_rows, counts, s);
    if (e == cudaSuccess)
      e = launch_rows_fc(hidden, hidden_dtype, hdim, rows, counts + b, max_rows, w1, b1, w_dtype, hdim, hdim, true, hid,
                         h->n_sm, s);
    if (e == cudaSuccess)
      e = launch_rows_fc(hid, DT_F32, hdim, nullptr, counts + b, max_rows, w2, b2, w_dtype, hdim, kOutCh, false, emb_out,
                         h->n_sm, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(counts_out, counts, sizeof(int) * (b + 1), cudaMemcpyDeviceToDevice, s);
    if (e != cudaSuccess) {
      rc = na_fail(h, B200SAM_ERR_CUDA, cudaGetErrorString(e));
      break;
    }
    h->launches += 3;
  } while (false);
  if (prev != dev) cudaSetDevice(prev);
  return rc;
}
