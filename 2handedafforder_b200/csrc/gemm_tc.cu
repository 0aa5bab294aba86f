// tcgen05 GEMM family for the SAM branch (B200 / sm_100a).
//
// One persistent, warp-specialised kernel: warp 0 = TMA producer (cp.async.bulk.tensor, SWIZZLE_128B tiles
// into a 4-6 stage shared-memory ring), warp 1 = single-thread tcgen05.mma issuer (M=128 x N=BN x K=16/8,
// fp32 accumulators in TMEM, two accumulator buffers so the epilogue of tile i overlaps the MMAs of tile
// i+1), warp 2 = TMEM allocator, warps 4-7 = epilogue (tcgen05.ld -> registers -> fused epilogue -> global).
//
// The epilogues are the fusions the reference executes as separate ATen ops:
//   GM_BIAS_BF16       qkv Linear                      image_encoder.py:238
//   GM_BIAS_GELU_BF16  MLPBlock.lin1 + nn.GELU (erf)   common.py:25-26
//   GM_F32             patch-embed bias + pos_embed    image_encoder.py:111-113,
//                      attn.proj + window_unpartition + residual  image_encoder.py:187-190,258,
//                      MLPBlock.lin2 + residual        image_encoder.py:191,
//                      decoder k/v/q projections with the constant positional part folded into a table
//                                                      transformer.py:220-225
//   GM_LN              neck conv + LayerNorm2d         image_encoder.py:92-108, common.py:38-43,
//                      cross_attn_image_to_token out_proj + residual + norm4  transformer.py:176-180
//   GM_UPS1 / GM_UPS2  output_upscaling ConvT/LN2d/GELU/ConvT/GELU and the hypernetwork mask product
//                                                      mask_decoder.py:54-64,154-165
#include "gemm_tc.cuh"

#include <cudaTypedefs.h>
#include <stdio.h>
#include <stdlib.h>

#include <type_traits>

#include <atomic>
#include <mutex>

#include "common.cuh"
#include "kernels.cuh"

namespace bsam {

namespace {

constexpr int kBM = 128;
constexpr int kThreads = 384;   // warps 0-3: TMA / MMA / TMEM alloc / spare; warps 4-11: epilogue

template <int BN>
struct TileCfg {
  static constexpr int kStageBytes = kBM * 128 + BN * 128;
  static constexpr int kStages = (BN == 256) ? 4 : 6;
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/ + 32768 /*epilogue tiles*/;
};

// token row (b*4096 + y*64 + x) -> row of the 14x14-window layout (b*4900 + window*196 + in-window index);
// window_partition of image_encoder.py:263-288 as an index map
__device__ __forceinline__ long token_to_window_row(long row) {
  const long b = row / kTokens;
  const int t = (int)(row - b * kTokens);
  const int y = t / kGrid, x = t % kGrid;
  return b * kWinRowsPerFrame + ((y / kWin) * kWinPerSide + x / kWin) * kWinTokens + (y % kWin) * kWin + x % kWin;
}

struct TmemSrc {
  uint32_t taddr;  // lane offset | first column of this accumulator buffer
  __device__ __forceinline__ void load(int chunk, float (&v)[32]) const {
    uint32_t r[32];
    tmem_ld_32x32(taddr + chunk * 32, r);
    tmem_wait_ld();
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
  }
};

template <bool kTf32>
struct SimtSrc {
  const void* A;
  const void* W;
  int lda, ldw, K, N, n0;
  long row;
  bool row_ok;
  __device__ void load(int chunk, float (&v)[32]) const {
#pragma unroll 1
    for (int j = 0; j < 32; ++j) {
      int col = n0 + chunk * 32 + j;
      float acc = 0.f;
      if (row_ok && col < N) {
        if constexpr (kTf32) {
          const float* a = reinterpret_cast<const float*>(A) + row * lda;
          const float* w = reinterpret_cast<const float*>(W) + (long)col * ldw;
          for (int k = 0; k < K; ++k) acc += a[k] * w[k];
        } else {
          const __nv_bfloat16* a = reinterpret_cast<const __nv_bfloat16*>(A) + row * lda;
          const __nv_bfloat16* w = reinterpret_cast<const __nv_bfloat16*>(W) + (long)col * ldw;
          for (int k = 0; k < K; ++k) acc += __bfloat162float(a[k]) * __bfloat162float(w[k]);
        }
      }
      v[j] = acc;
    }
  }
};

__device__ __forceinline__ void add_vec32(float (&v)[32], const float* __restrict__ p) {
  const float4* p4 = reinterpret_cast<const float4*>(p);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    float4 t = __ldg(p4 + j);
    v[4 * j + 0] += t.x;
    v[4 * j + 1] += t.y;
    v[4 * j + 2] += t.z;
    v[4 * j + 3] += t.w;
  }
}
// plain (coherent) loads: the addend may alias the output (in-place residual update)
__device__ __forceinline__ void add_vec32_plain(float (&v)[32], const float* p) {
  const float4* p4 = reinterpret_cast<const float4*>(p);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    float4 t = p4[j];
    v[4 * j + 0] += t.x;
    v[4 * j + 1] += t.y;
    v[4 * j + 2] += t.z;
    v[4 * j + 3] += t.w;
  }
}
__device__ __forceinline__ void store_f32x32(float* o, const float (&v)[32]) {
  float4* o4 = reinterpret_cast<float4*>(o);
#pragma unroll
  for (int j = 0; j < 8; ++j) o4[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
}
__device__ __forceinline__ void store_bf16x32(__nv_bfloat16* o, const float (&v)[32]) {
  uint4* o4 = reinterpret_cast<uint4*>(o);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint4 t;
    t.x = pack_bf16(v[8 * j + 0], v[8 * j + 1]);
    t.y = pack_bf16(v[8 * j + 2], v[8 * j + 3]);
    t.z = pack_bf16(v[8 * j + 4], v[8 * j + 5]);
    t.w = pack_bf16(v[8 * j + 6], v[8 * j + 7]);
    o4[j] = t;
  }
}

// fp16 stores of the rel-pos projections saturate instead of producing inf (an inf bias would turn a whole softmax row
// into NaN); the caller is told through a host-mapped flag, so an out-of-range checkpoint fails loudly, not silently.
constexpr float kF16Max = 65504.f;
__device__ __forceinline__ float sat_f16(float x, bool& hit) {
  hit |= fabsf(x) > kF16Max;
  return fminf(fmaxf(x, -kF16Max), kF16Max);
}

// One output row (row of the 128-row tile owned by this thread), columns [n0, n0 + BN).
// `src.load` is warp-collective for the TMEM source: every lane must call it the same number of times.
struct NoWait {
  __device__ __forceinline__ void operator()() const {}
};

// `half`/`nhalves`: the 32-column chunks of the tile are split between `nhalves` warps that share a TMEM lane
// quadrant (modes with a row-wise reduction over the tile keep nhalves == 1).  `wait_acc()` blocks until the
// accumulator is complete; it is called after the first global prefetch so that load overlaps the main loop.
template <int MODE, int BN, class Src, class Wait>
__device__ __forceinline__ void epilogue_row(const GemmEpi& e, const Src& src, long row, int M, int N, int n0,
                                             long out_off, int half, int nhalves, const Wait& wait_acc) {
  const bool row_in = row < M;
  const int ncols = min(BN, N - n0);
  const int nchunks = ncols / 32;
  const int c_lo = (nchunks * half) / nhalves, c_hi = (nchunks * (half + 1)) / nhalves;
  float v[32];

  if constexpr (MODE == GM_BIAS_BF16 || MODE == GM_BIAS_GELU_BF16) {
    const long orow_b = e.rowmap == RM_TOKEN_TO_WINDOW ? token_to_window_row(row_in ? row : 0) : row;
    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(e.out) + orow_b * (long)e.ld_out + n0;
    wait_acc();
    for (int c = c_lo; c < c_hi; ++c) {
      src.load(c, v);
      if (row_in) {
        add_vec32(v, e.bias + n0 + c * 32);
        if constexpr (MODE == GM_BIAS_GELU_BF16) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = gelu_erf_fast(v[j]);
        }
        store_bf16x32(o + c * 32, v);
      }
    }
  } else if constexpr (MODE == GM_F16) {
    __half* o = reinterpret_cast<__half*>(e.out) + row * (long)e.ld_out + n0 + out_off;
    wait_acc();
    for (int c = c_lo; c < c_hi; ++c) {
      src.load(c, v);
      if (row_in) {
        bool hit = false;
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = sat_f16(v[j], hit);
        if (hit && e.f16_overflow) *e.f16_overflow = 1;
        uint4* o4 = reinterpret_cast<uint4*>(o + c * 32);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint4 t;
          __half2 h0 = __floats2half2_rn(v[8 * j + 0], v[8 * j + 1]), h1 = __floats2half2_rn(v[8 * j + 2], v[8 * j + 3]);
          __half2 h2 = __floats2half2_rn(v[8 * j + 4], v[8 * j + 5]), h3 = __floats2half2_rn(v[8 * j + 6], v[8 * j + 7]);
          t.x = *reinterpret_cast<uint32_t*>(&h0);
          t.y = *reinterpret_cast<uint32_t*>(&h1);
          t.z = *reinterpret_cast<uint32_t*>(&h2);
          t.w = *reinterpret_cast<uint32_t*>(&h3);
          o4[j] = t;
        }
      }
    }
  } else if constexpr (MODE == GM_F32) {
    long orow = row;
    bool ok = row_in;
    if (e.rowmap == RM_WINDOW_TO_TOKEN) {
      long b = row / kWinRowsPerFrame;
      int rem = (int)(row - b * kWinRowsPerFrame);
      int win = rem / kWinTokens, t = rem - win * kWinTokens;
      int y = (win / kWinPerSide) * kWin + t / kWin;
      int x = (win % kWinPerSide) * kWin + t % kWin;
      ok = ok && (y < kGrid) && (x < kGrid);
      orow = b * kTokens + y * kGrid + x;
    }
    float* o = reinterpret_cast<float*>(e.out) + orow * (long)e.ld_out + n0;
    const float* a = nullptr;
    if (e.addend) {
      long arow = e.add_mod > 0 ? (orow % e.add_mod) : orow;
      a = e.addend + arow * (long)e.ld_add + n0;
    }
    // the addend (residual stream) chunk is fetched one step ahead: its HBM/L2 latency overlaps the accumulator
    // wait and the previous chunk's TMEM load.  The addend may alias the output; chunks are disjoint columns.
    const bool pf = ok && a != nullptr;
    float4 nxt[8];
    if (pf && c_lo < c_hi) {
#pragma unroll
      for (int j = 0; j < 8; ++j) nxt[j] = reinterpret_cast<const float4*>(a + c_lo * 32)[j];
    }
    wait_acc();
    for (int c = c_lo; c < c_hi; ++c) {
      float4 cur[8];
      if (pf) {
#pragma unroll
        for (int j = 0; j < 8; ++j) cur[j] = nxt[j];
        if (c + 1 < c_hi) {
#pragma unroll
          for (int j = 0; j < 8; ++j) nxt[j] = reinterpret_cast<const float4*>(a + (c + 1) * 32)[j];
        }
      }
      src.load(c, v);
      if (ok) {
        if (e.bias) add_vec32(v, e.bias + n0 + c * 32);
        if (pf) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            v[4 * j + 0] += cur[j].x;
            v[4 * j + 1] += cur[j].y;
            v[4 * j + 2] += cur[j].z;
            v[4 * j + 3] += cur[j].w;
          }
        }
        store_f32x32(o + c * 32, v);
      }
    }
  } else if constexpr (MODE == GM_LN) {
    // N == BN == 256; shifted single-pass statistics, then a second pass over TMEM to normalise.
    const float* a = e.addend ? e.addend + row * (long)e.ld_add : nullptr;
    float shift = 0.f, s1 = 0.f, s2 = 0.f;
    wait_acc();
    for (int c = 0; c < 8; ++c) {
      src.load(c, v);
      if (row_in) {
        if (e.bias) add_vec32(v, e.bias + c * 32);
        if (a) add_vec32_plain(v, a + c * 32);
        if (c == 0) shift = v[0];
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          float d = v[j] - shift;
          s1 += d;
          s2 += d * d;
        }
      }
    }
    const float md = s1 * (1.f / 256.f);
    const float var = fmaxf(s2 * (1.f / 256.f) - md * md, 0.f);
    const float mean = shift + md;
    const float rstd = rsqrtf(var + e.ln_eps);
    for (int c = 0; c < 8; ++c) {
      src.load(c, v);
      if (row_in) {
        if (e.bias) add_vec32(v, e.bias + c * 32);
        if (a) add_vec32_plain(v, a + c * 32);
#pragma unroll
        for (int j = 0; j < 32; ++j)
          v[j] = (v[j] - mean) * rstd * __ldg(e.ln_g + c * 32 + j) + __ldg(e.ln_b + c * 32 + j);
        if (e.out) store_f32x32(reinterpret_cast<float*>(e.out) + row * (long)e.ld_out + c * 32, v);
        if (e.out2) {
          long r2 = row;
          if (e.rowmap == RM_TOKEN_TO_PAD66) {
            const long b = row / kTokens;
            const int t = (int)(row - b * kTokens);
            r2 = b * (66 * 66) + (t / kGrid + 1) * 66 + (t % kGrid) + 1;
          }
          store_bf16x32(reinterpret_cast<__nv_bfloat16*>(e.out2) + r2 * (long)e.ld_out + c * 32, v);
        }
      }
    }
  } else if constexpr (MODE == GM_UPS1) {
    // row = p*4096 + y*64 + x; columns (ky, kx, c64).  Output (P,128,128,64) token-major fp32.
    long p = row / kTokens;
    int t = (int)(row - p * kTokens);
    int y = t / kGrid, x = t % kGrid;
    float w[32];
    wait_acc();
    for (int g = 0; g < 4; ++g) {
      src.load(2 * g, v);
      src.load(2 * g + 1, w);
      if (row_in) {
        add_vec32(v, e.bias + g * 64);
        add_vec32(w, e.bias + g * 64 + 32);
        float s = 0.f;
#pragma unroll
        for (int j = 0; j < 32; ++j) s += v[j] + w[j];
        const float mean = s * (1.f / 64.f);
        float q = 0.f;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          float d0 = v[j] - mean, d1 = w[j] - mean;
          q += d0 * d0 + d1 * d1;
        }
        const float rstd = rsqrtf(q * (1.f / 64.f) + e.ln_eps);
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          v[j] = gelu_erf_poly((v[j] - mean) * rstd * __ldg(e.ln_g + j) + __ldg(e.ln_b + j));
          w[j] = gelu_erf_poly((w[j] - mean) * rstd * __ldg(e.ln_g + 32 + j) + __ldg(e.ln_b + 32 + j));
        }
        int ky = g >> 1, kx = g & 1;
        float* o = reinterpret_cast<float*>(e.out) + ((p * 128 + 2 * y + ky) * 128 + 2 * x + kx) * 64;
        store_f32x32(o, v);
        store_f32x32(o + 32, w);
      }
    }
  } else if constexpr (MODE == GM_UPS2) {
    // row = (p*128 + y)*128 + x; columns (ky, kx, c32).  Output mask logits (P,256,256) fp32.
    long p = row / (128 * 128);
    int t = (int)(row - p * 128 * 128);
    int y = t / 128, x = t % 128;
    const float* h = e.hyper + p * 32;
    wait_acc();
    for (int g = 0; g < 4; ++g) {
      src.load(g, v);
      if (row_in) {
        add_vec32(v, e.bias + g * 32);
        float dot = 0.f;
#pragma unroll
        for (int j = 0; j < 32; ++j) dot += gelu_erf_poly(v[j]) * __ldg(h + j);
        int ky = g >> 1, kx = g & 1;
        reinterpret_cast<float*>(e.out)[p * 65536 + (2 * y + ky) * 256 + 2 * x + kx] = dot;
      }
    }
  }
}

// Warp-collective epilogue of the splittable modes (bias->bf16, bias+GELU->bf16, fp16, bias+addend->fp32) with
// coalesced global traffic.  tcgen05.ld hands every lane one accumulator ROW (32 columns); storing that directly makes
// each 16-byte store instruction touch 32 different 128-byte lines.  Instead each 32x32 chunk is transposed through a
// private 4 KB shared-memory tile (XOR-swizzled float4 slots, conflict free both ways) so that 8 consecutive lanes
// cover 128 contiguous bytes of one row and a load/store instruction touches 4 lines instead of 32.
constexpr int kEpiStageBytes = 8 * 32 * 32 * 4;   // one 32x32 fp32 tile per epilogue warp

// GM_F32_FOLD (residual-stream epilogue + producer side of the folded LayerNorm).  gemm_tc() guarantees the identity
// row map, an addend, add_mod % 32 == 0 and no K split, so the eight row slices of a lane sit at a constant stride from
// one base pointer and the row loop has no per-slice branch: the compiler interleaves the slices (the general loop
// below serialises them behind its row-map / modulo / split-K branches, and with two epilogue warps per scheduler
// that chain of dependent instructions - not HBM - was what bounded the proj GEMM).  The {sum, sum of squares} of a
// row are kept per lane over the whole half tile and combined once at the end by a transposing butterfly
// (7 shuffles per quantity instead of 24 per chunk); lane (sub, cg) ends up with the pair of row cg * 4 + sub.
template <int BN, int LD, class Wait>
__device__ __forceinline__ void epilogue_warp_fold(const GemmEpi& e, const TmemSrc& src, long row0, int lane, int M,
                                                   int N, int n0, int half, int nhalves, const Wait& wait_acc,
                                                   float4* stage) {
  const int ncols = min(BN, N - n0);
  const int nchunks = ncols / 32;
  const int c_lo = (nchunks * half) / nhalves, c_hi = (nchunks * (half + 1)) / nhalves;
  const int sub = lane >> 3, cg = lane & 7;     // rows row0 + i*4 + sub (i < 8), columns cg*4 .. cg*4+3 of a chunk
  const long r0 = row0 + sub;
  const int left = (int)min(32L, max(0L, (long)M - r0));   // slice i holds a row of the matrix iff 4 i < left
  const bool all_rows = __all_sync(0xffffffffu, left >= 29);
  // LD > 0: the leading dimension of out / xb_out / addend is a compile-time constant, so the eight slices of a lane
  // are immediate offsets from three running pointers (advanced by one chunk per iteration) - no address arithmetic
  const int ld_o = LD > 0 ? LD : e.ld_out, ld_a = LD > 0 ? LD : e.ld_add;
  const long arow0 = (e.add_mod > 0 ? row0 % e.add_mod : row0) + sub;
  const float* ap = e.addend + arow0 * (long)ld_a + n0 + cg * 4 + c_lo * 32;
  const long obase = r0 * (long)ld_o + n0 + cg * 4 + c_lo * 32;
  float* op = reinterpret_cast<float*>(e.out) + obase;
  __nv_bfloat16* xp = reinterpret_cast<__nv_bfloat16*>(e.xb_out) + obase;
  const int astride = 4 * ld_a, ostride = 4 * ld_o;

  float4 nxt[8];
  float ps[8], pq[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    nxt[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    ps[i] = pq[i] = 0.f;
  }
  if (c_lo < c_hi) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      if (4 * i < left) nxt[i] = *reinterpret_cast<const float4*>(ap + i * astride);
  }
  // bias of the whole half tile (at most 4 chunks = 128 columns: nhalves = 2, BN <= 256) in ONE load per lane, lane L
  // holding columns 4 L .. 4 L + 3; a chunk takes its four values by shuffle.  A per-chunk bias load inside the loop
  // shares a scoreboard with the residual prefetches issued after it, so consuming it at the top of the next chunk made
  // the warp wait for those prefetches too (19 % of the kernel's stall samples, profiles/r6_prof_gemm_summary.csv)
  float4 b_own = make_float4(0.f, 0.f, 0.f, 0.f);
  if (e.bias && c_lo * 32 + lane * 4 < c_hi * 32)
    b_own = __ldg(reinterpret_cast<const float4*>(e.bias + n0 + c_lo * 32 + lane * 4));
  wait_acc();
  const uint32_t stage_s = smem_u32(stage);
  uint32_t v[32];
  if (c_lo < c_hi) tmem_ld_32x32(src.taddr + c_lo * 32, v);
  for (int c = c_lo; c < c_hi; ++c) {
    tmem_wait_ld();
#pragma unroll
    for (int g = 0; g < 8; ++g)
      sts128(stage_s + (lane * 8 + (g ^ (lane & 7))) * 16, __uint_as_float(v[4 * g]), __uint_as_float(v[4 * g + 1]),
             __uint_as_float(v[4 * g + 2]), __uint_as_float(v[4 * g + 3]));
    __syncwarp();
    float4 b4;
    {
      const int src_lane = (c - c_lo) * 8 + cg;
      b4.x = __shfl_sync(0xffffffffu, b_own.x, src_lane), b4.y = __shfl_sync(0xffffffffu, b_own.y, src_lane);
      b4.z = __shfl_sync(0xffffffffu, b_own.z, src_lane), b4.w = __shfl_sync(0xffffffffu, b_own.w, src_lane);
    }
    const bool more = c + 1 < c_hi;
    if (more) tmem_ld_32x32(src.taddr + (c + 1) * 32, v);
    auto rows = [&](auto all_tag) {
      constexpr bool kAll = decltype(all_tag)::value;      // the 32 rows of this warp all exist: no store predicates
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int r = i * 4 + sub;
        float4 x = lds128(stage_s + (r * 8 + (cg ^ (r & 7))) * 16);
        add4(x, b4);
        add4(x, nxt[i]);
        const bool st = kAll || 4 * i < left;
        // single-buffered prefetch: the register of slice i is refilled with the next chunk as soon as it is consumed
        if (more && st) nxt[i] = *reinterpret_cast<const float4*>(ap + i * astride + 32);
        if (st) {
          *reinterpret_cast<float4*>(op + i * ostride) = x;
          uint2 t;
          t.x = pack_bf16(x.x, x.y);
          t.y = pack_bf16(x.z, x.w);
          *reinterpret_cast<uint2*>(xp + i * ostride) = t;
        }
        ps[i] += (x.x + x.y) + (x.z + x.w);
        pq[i] = fmaf(x.x, x.x, fmaf(x.y, x.y, fmaf(x.z, x.z, fmaf(x.w, x.w, pq[i]))));
      }
    };
    if constexpr (LD > 0) {
      if (all_rows) rows(std::true_type{});
      else rows(std::false_type{});
    } else {
      rows(std::false_type{});      // one instantiation only: two of them spill with run-time strides
    }
    ap += 32, op += 32, xp += 32;
    __syncwarp();
  }
  // transposing butterfly over the eight column lanes of a row group: after the rounds (xor 4, 2, 1) lane cg holds the
  // total of slice cg
  {
    const bool h4 = (cg & 4) != 0, h2 = (cg & 2) != 0, h1 = (cg & 1) != 0;
    float s4[4], q4[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float ks = h4 ? ps[j + 4] : ps[j], ss = h4 ? ps[j] : ps[j + 4];
      const float kq = h4 ? pq[j + 4] : pq[j], sq = h4 ? pq[j] : pq[j + 4];
      s4[j] = ks + __shfl_xor_sync(0xffffffffu, ss, 4);
      q4[j] = kq + __shfl_xor_sync(0xffffffffu, sq, 4);
    }
    float s2[2], q2[2];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const float ks = h2 ? s4[j + 2] : s4[j], ss = h2 ? s4[j] : s4[j + 2];
      const float kq = h2 ? q4[j + 2] : q4[j], sq = h2 ? q4[j] : q4[j + 2];
      s2[j] = ks + __shfl_xor_sync(0xffffffffu, ss, 2);
      q2[j] = kq + __shfl_xor_sync(0xffffffffu, sq, 2);
    }
    const float st_s = (h1 ? s2[1] : s2[0]) + __shfl_xor_sync(0xffffffffu, h1 ? s2[0] : s2[1], 1);
    const float st_q = (h1 ? q2[1] : q2[0]) + __shfl_xor_sync(0xffffffffu, h1 ? q2[0] : q2[1], 1);
    // also when this half owns no chunk: the slot is then written as zeros
    const long row = row0 + cg * 4 + sub;
    if (row < M)
      reinterpret_cast<float2*>(e.stats_out)[row * e.stats_slots + (n0 / BN) * nhalves + half] = make_float2(st_s, st_q);
  }
}

// LD: compile-time leading dimension of the GM_F32_FOLD operands (0 = run time), see epilogue_warp_fold
template <int MODE, int BN, int LD, class Wait>
__device__ __forceinline__ void epilogue_warp(const GemmEpi& e, const TmemSrc& src, long row0, int lane, int M, int N,
                                              int n0, long out_off, int half, int nhalves, const Wait& wait_acc,
                                              float4* stage) {
  if constexpr (MODE == GM_F32_FOLD) {
    epilogue_warp_fold<BN, LD>(e, src, row0, lane, M, N, n0, half, nhalves, wait_acc, stage);
    return;
  }
  constexpr bool kF32 = MODE == GM_F32;
  const int ncols = min(BN, N - n0);
  const int nchunks = ncols / 32;
  const int c_lo = (nchunks * half) / nhalves, c_hi = (nchunks * (half + 1)) / nhalves;
  // output row of the accumulator row this lane owns (-1: nothing to store)
  int orow_own;
  {
    const long row = row0 + lane;
    bool ok = row < M;
    long orow = row;
    if (kF32 && e.rowmap == RM_WINDOW_TO_TOKEN) {
      long b = row / kWinRowsPerFrame;
      int rem = (int)(row - b * kWinRowsPerFrame);
      int win = rem / kWinTokens, t = rem - win * kWinTokens;
      int y = (win / kWinPerSide) * kWin + t / kWin;
      int x = (win % kWinPerSide) * kWin + t % kWin;
      ok = ok && (y < kGrid) && (x < kGrid);
      orow = b * kTokens + y * kGrid + x;
    }
    if ((MODE == GM_BIAS_BF16 || MODE == GM_BIAS_GELU_BF16) && e.rowmap == RM_TOKEN_TO_WINDOW && ok)
      orow = token_to_window_row(row);
    orow_own = ok ? (int)orow : -1;
  }
  const bool all_rows = __all_sync(0xffffffffu, orow_own >= 0);
  const int sub = lane >> 3, cg = lane & 7;     // transposed role: rows i*4 + sub (i < 8), columns cg*4 .. cg*4+3
  int orow[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) orow[i] = __shfl_sync(0xffffffffu, orow_own, i * 4 + sub);

  const bool has_add = kF32 && e.addend != nullptr;
  float4 nxt[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) nxt[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  auto fetch_row = [&](int c, int i) {
    if (orow[i] >= 0) {
      const long arow = e.add_mod > 0 ? (orow[i] % e.add_mod) : orow[i];
      nxt[i] = *reinterpret_cast<const float4*>(e.addend + arow * (long)e.ld_add + n0 + c * 32 + cg * 4);
    }
  };
  auto fetch = [&](int c) {
#pragma unroll
    for (int i = 0; i < 8; ++i) fetch_row(c, i);
  };
  auto bias4 = [&](int c) {
    return (MODE != GM_F16 && e.bias) ? __ldg(reinterpret_cast<const float4*>(e.bias + n0 + c * 32 + cg * 4))
                                      : make_float4(0.f, 0.f, 0.f, 0.f);
  };
  // folded LayerNorm, consumer side: lane L derives rstd and -rstd * mean of accumulator row row0 + L from the
  // producer's partial sums; the row loop below fetches the pair of row i*4 + sub by shuffle
  constexpr bool kFoldIn = MODE == GM_BIAS_BF16 || MODE == GM_BIAS_GELU_BF16;
  const bool fold_in = kFoldIn && e.stats_in != nullptr;
  float rs_own = 1.f, nm_own = 0.f;
  if constexpr (kFoldIn) if (fold_in) {
    const long row = row0 + lane;
    if (row < M) {
      const float2* sp = reinterpret_cast<const float2*>(e.stats_in) + row * e.stats_slots;
      float s1 = 0.f, s2 = 0.f;
      for (int j = 0; j < e.stats_slots; ++j) {
        const float2 v2 = __ldg(sp + j);
        s1 += v2.x;
        s2 += v2.y;
      }
      const float inv = 1.f / (float)e.stats_dim;
      const float mu = s1 * inv;
      rs_own = rsqrtf(fmaxf(s2 * inv - mu * mu, 0.f) + e.fold_eps);
      nm_own = -rs_own * mu;
    }
  }
  // the pairs of this lane's eight row slices (row i*4 + sub), fetched once per tile
  float rs_i[kFoldIn ? 8 : 1], nm_i[kFoldIn ? 8 : 1];
  if constexpr (kFoldIn) {
    if (fold_in) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        rs_i[i] = __shfl_sync(0xffffffffu, rs_own, i * 4 + sub);
        nm_i[i] = __shfl_sync(0xffffffffu, nm_own, i * 4 + sub);
      }
    }
  }
  auto foldc4 = [&](int c) {
    return fold_in ? __ldg(reinterpret_cast<const float4*>(e.fold_c + n0 + c * 32 + cg * 4))
                   : make_float4(0.f, 0.f, 0.f, 0.f);
  };
  if (has_add && c_lo < c_hi) fetch(c_lo);
  float4 b_nxt = c_lo < c_hi ? bias4(c_lo) : make_float4(0.f, 0.f, 0.f, 0.f);
  float4 c_nxt = make_float4(0.f, 0.f, 0.f, 0.f);
  if constexpr (kFoldIn) {
    if (c_lo < c_hi) c_nxt = foldc4(c_lo);
  }
  wait_acc();
  // Software pipeline, one chunk deep: the accumulator registers are dead as soon as they sit in the transposition
  // tile, so the tcgen05.ld of chunk c+1 (and its bias / addend loads) is in flight while chunk c is finished.
  const uint32_t stage_s = smem_u32(stage);
  uint32_t v[32];
  if (c_lo < c_hi) tmem_ld_32x32(src.taddr + c_lo * 32, v);
  for (int c = c_lo; c < c_hi; ++c) {
    tmem_wait_ld();
#pragma unroll
    for (int g = 0; g < 8; ++g)
      sts128(stage_s + (lane * 8 + (g ^ (lane & 7))) * 16, __uint_as_float(v[4 * g]), __uint_as_float(v[4 * g + 1]),
             __uint_as_float(v[4 * g + 2]), __uint_as_float(v[4 * g + 3]));
    __syncwarp();
    const float4 b4 = b_nxt, c4 = c_nxt;
    if (c + 1 < c_hi) {
      tmem_ld_32x32(src.taddr + (c + 1) * 32, v);
      b_nxt = bias4(c + 1);
      if constexpr (kFoldIn) c_nxt = foldc4(c + 1);
    }
    // the row loop, instantiated with and without the folded LayerNorm: the (uniform) choice is made once per chunk,
    // so the eight row slices stay one branch-free block that the compiler interleaves
    auto rows = [&](auto fold_tag, auto all_tag) {
    constexpr bool kFold = decltype(fold_tag)::value;
    constexpr bool kAll = decltype(all_tag)::value;      // every row of this warp is stored: no store predicate
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int r = i * 4 + sub;
      float4 x = lds128(stage_s + (r * 8 + (cg ^ (r & 7))) * 16);
      if constexpr (kFoldIn && kFold) {
        // rstd * x + (-rstd * mean * c + bias) on packed fp32x2: four FMA-pipe instructions per float4
        const uint64_t rs2 = f32x2(rs_i[i], rs_i[i]), nm2 = f32x2(nm_i[i], nm_i[i]);
        const uint64_t y01 = fma2(rs2, f32x2(x.x, x.y), fma2(nm2, f32x2(c4.x, c4.y), f32x2(b4.x, b4.y)));
        const uint64_t y23 = fma2(rs2, f32x2(x.z, x.w), fma2(nm2, f32x2(c4.z, c4.w), f32x2(b4.z, b4.w)));
        unpack2(y01, x.x, x.y);
        unpack2(y23, x.z, x.w);
      } else {
        add4(x, b4);
      }
      const bool st = kAll || orow[i] >= 0;
      const long col = n0 + c * 32 + cg * 4;
      if constexpr (kF32) {
        if (has_add) {
          // single-buffered prefetch of the residual rows: as soon as row slice i of this chunk is consumed its
          // register is refilled with the same rows of the next chunk (one chunk of latency cover, 32 registers)
          add4(x, nxt[i]);
          if (c + 1 < c_hi) fetch_row(c + 1, i);
        }
        float* op = reinterpret_cast<float*>(e.out) + orow[i] * (long)e.ld_out + col;
        if (e.ksplit > 1) {          // partial product of one K range: combine in place
          if (st) {
            atomicAdd(op, x.x);
            atomicAdd(op + 1, x.y);
            atomicAdd(op + 2, x.z);
            atomicAdd(op + 3, x.w);
          }
        } else if (st) {
          *reinterpret_cast<float4*>(op) = x;
        }
      } else if constexpr (MODE == GM_F16) {
        const float ws = (e.rowmap == RM_RELW_SHIFT && col >= 128) ? e.relw_scale : 1.f;
        bool hit = false;
        x.x = sat_f16(x.x * ws, hit), x.y = sat_f16(x.y * ws, hit);
        x.z = sat_f16(x.z * ws, hit), x.w = sat_f16(x.w * ws, hit);
        if (hit && st && e.f16_overflow) *e.f16_overflow = 1;
        __half2 h0 = __floats2half2_rn(x.x, x.y), h1 = __floats2half2_rn(x.z, x.w);
        uint2 t;
        t.x = *reinterpret_cast<uint32_t*>(&h0);
        t.y = *reinterpret_cast<uint32_t*>(&h1);
        __half* orp = reinterpret_cast<__half*>(e.out) + orow[i] * (long)e.ld_out + out_off;
        if (e.rowmap == RM_RELW_SHIFT && col >= 128) {
          // per-row shift: source column 128 + j goes to 128 + (j - 63 + qx) when that lands in [0, 64)
          const int kx0 = (int)col - 191 + (orow[i] & 63);
          const __half hv[4] = {__float2half_rn(x.x), __float2half_rn(x.y), __float2half_rn(x.z),
                                __float2half_rn(x.w)};   // already scaled by relw_scale and saturated above
#pragma unroll
          for (int q = 0; q < 4; ++q)
            if (st && (unsigned)(kx0 + q) < 64u) orp[128 + kx0 + q] = hv[q];
        } else if (st) {
          *reinterpret_cast<uint2*>(orp + col) = t;
        }
      } else {
        if constexpr (MODE == GM_BIAS_GELU_BF16) {
          gelu_erf_poly2(x.x, x.y);
          gelu_erf_poly2(x.z, x.w);
        }
        uint2 t;
        t.x = pack_bf16(x.x, x.y);
        t.y = pack_bf16(x.z, x.w);
        if (st) *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(e.out) + orow[i] * (long)e.ld_out + col) = t;
      }
    }
    };
    if constexpr (kFoldIn) {
      if (fold_in) {
        if (all_rows) rows(std::true_type{}, std::true_type{});
        else rows(std::true_type{}, std::false_type{});
      } else {
        if (all_rows) rows(std::false_type{}, std::true_type{});
        else rows(std::false_type{}, std::false_type{});
      }
    } else {
      if (all_rows) rows(std::false_type{}, std::true_type{});
      else rows(std::false_type{}, std::false_type{});
    }
    __syncwarp();
  }
}

template <bool kTf32, int MODE, int BN>
__global__ void __launch_bounds__(kThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, int M, int N, int K,
               GemmEpi e) {
  griddep_launch_dependents();
  using Cfg = TileCfg<BN>;
  constexpr int kStages = Cfg::kStages;
  constexpr int kBK = kTf32 ? 32 : 64;                 // elements per 128-byte row
  constexpr uint32_t kIdesc = (1u << 4) /*D=f32*/ | ((kTf32 ? 2u : 1u) << 7) | ((kTf32 ? 2u : 1u) << 10) |
                              (uint32_t(BN >> 3) << 17) | (uint32_t(kBM >> 4) << 24);

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * Cfg::kStageBytes);
  uint64_t* full = bars;
  uint64_t* empty = bars + kStages;
  uint64_t* tmem_full = bars + 2 * kStages;
  uint64_t* tmem_empty = bars + 2 * kStages + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int m_tiles = (M + kBM - 1) / kBM;
  const int n_tiles = (N + BN - 1) / BN;
  const int tiles_per_batch = m_tiles * n_tiles;
  const int kblocks = (K + kBK - 1) / kBK;
  // split-K (GM_F32, weight gradients of the training path): the "batch" index enumerates K ranges of kper blocks whose
  // partial products are combined with fp32 atomics by the epilogue (the launcher guarantees no range is empty)
  const bool ksplit = e.ksplit > 1;
  const int kper = ksplit ? (kblocks + e.ksplit - 1) / e.ksplit : kblocks;
  const int num_tiles = tiles_per_batch * (ksplit ? e.ksplit : e.batch);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);
      mbar_init(&tmem_empty[s], 256);
    }
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 2 * BN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  griddep_wait();     // PDL: everything above overlapped the previous kernel's tail

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int bidx = tile / tiles_per_batch, trem = tile - bidx * tiles_per_batch;
        const int m_blk = trem / n_tiles, n_blk = trem % n_tiles;
        const int a_col0 = ksplit ? 0 : bidx * e.a_batch_off;
        const int kb0 = ksplit ? bidx * kper : 0, kb1 = ksplit ? min(kblocks, kb0 + kper) : kblocks;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          uint8_t* sA = smem + stage * Cfg::kStageBytes;
          uint8_t* sB = sA + kBM * 128;
          mbar_expect_tx(&full[stage], Cfg::kStageBytes);
          if (!kTf32 && e.a_mode == AM_CONV3) {
            // 128 tokens = image rows y0, y0+1 of frame b; k block kb = tap (dy, dx), channels c0 .. c0+63.  The box
            // (64 c, 64 x, 2 y) of the zero-ringed activation lands as 128 rows x 128 bytes, i.e. the K-major SW128 tile
            const int tap = kb >> 2, c0 = (kb & 3) * 64;
            tma_load_4d(sA, &tmA, &full[stage], c0, tap % 3, (m_blk & 31) * 2 + tap / 3, m_blk >> 5);
          } else {
            tma_load_2d(sA, &tmA, &full[stage], a_col0 + kb * kBK, m_blk * kBM);
          }
          tma_load_2d(sB, &tmB, &full[stage], kb * kBK, n_blk * BN);
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
        const int as = it & 1;
        const uint32_t aph = (it >> 1) & 1;
        mbar_wait(&tmem_empty[as], aph ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        const int ks = tile / tiles_per_batch;
        const int kb0 = ksplit ? ks * kper : 0, kb1 = ksplit ? min(kblocks, kb0 + kper) : kblocks;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t a0 = smem_u32(smem + stage * Cfg::kStageBytes);
          const uint32_t b0 = a0 + kBM * 128;
#pragma unroll
          for (int k = 0; k < 4; ++k) {   // 4 MMAs of K=16 (bf16) / K=8 (tf32): 32 bytes of K each
            umma_ss<kTf32>(d_tmem, umma_desc_sw128(a0 + k * 32), umma_desc_sw128(b0 + k * 32), kIdesc,
                           (kb != kb0 || k != 0) ? 1u : 0u);
          }
          umma_commit(&empty[stage]);     // frees the smem slot when these MMAs retire
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit(&tmem_full[as]);      // accumulator complete -> epilogue
      }
    }
  } else if (warp >= 4) {
    const int ew = warp & 3;              // TMEM lane quadrant this warp may access
    const int eh = (warp - 4) >> 2;       // which half of the tile's column chunks
    constexpr bool kSplit = MODE == GM_BIAS_BF16 || MODE == GM_BIAS_GELU_BF16 || MODE == GM_F32 || MODE == GM_F16 ||
                            MODE == GM_F32_FOLD;
    int it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int bidx = tile / tiles_per_batch, trem = tile - bidx * tiles_per_batch;
      const int m_blk = trem / n_tiles, n_blk = trem % n_tiles;
      const int as = it & 1;
      const uint32_t aph = (it >> 1) & 1;
      auto wait_acc = [&]() {
        mbar_wait(&tmem_full[as], aph);
        tc_fence_after();
      };
      TmemSrc src{tmem_base + (uint32_t(ew * 32) << 16) + uint32_t(as * BN)};
      const long row = (long)m_blk * kBM + ew * 32 + lane;
      if constexpr (kSplit)
        epilogue_warp<MODE, BN, 0>(e, src, (long)m_blk * kBM + ew * 32, lane, M, N, n_blk * BN,
                                ksplit ? 0L : (long)bidx * e.out_batch_off, eh, 2, wait_acc,
                                reinterpret_cast<float4*>(smem + kStages * Cfg::kStageBytes + 256) + (warp - 4) * 256);
      else if (eh == 0)
        epilogue_row<MODE, BN>(e, src, row, M, N, n_blk * BN, (long)bidx * e.out_batch_off, 0, 1, wait_acc);
      else
        wait_acc();
      tc_fence_before();
      mbar_arrive(&tmem_empty[as]);
    }
  }

  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 2) tmem_dealloc(tmem_base, 2 * BN);
}

// ------------------------------------------------------------------------------------------------ CTA-pair kernel
// Same contract for the four big encoder GEMMs (bf16 operands), computed by CTA pairs (cluster of 2, cta_group::2):
// one 256 x 256 output tile per pair, each CTA stages its own 128 rows of A and 128 of the 256 rows of W (so every
// operand byte is read from L2 once per pair instead of once per CTA, and the 32 KB stages allow a 6-deep ring), the
// leader issues tcgen05.mma with M = 256 and each CTA drains its own 128 accumulator lanes.
constexpr int kStages2 = 6;
constexpr int kStageBytes2 = kBM * 128 + 128 * 128;   // A half + W half
constexpr int kSmemBytes2 = kStages2 * kStageBytes2 + 1024 + 256 + 32768 /*epilogue transposition tiles*/;

template <int MODE, int LD = 0>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
gemm_tc2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, int M, int N, int K,
                GemmEpi e) {
  griddep_launch_dependents();
  constexpr int BN = 256;
  constexpr uint32_t kIdesc = umma_idesc_bf16(256, BN, false, false);
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages2 * kStageBytes2);
  uint64_t* full = bars;
  uint64_t* empty = bars + kStages2;
  uint64_t* tmem_full = bars + 2 * kStages2;
  uint64_t* tmem_empty = bars + 2 * kStages2 + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages2 + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;
  const int m_tiles = (M + 255) / 256;
  const int n_tiles = (N + BN - 1) / BN;
  const int num_tiles = m_tiles * n_tiles;
  const int kblocks = (K + 63) / 64;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages2; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);
      mbar_init(&tmem_empty[s], 512);     // the epilogue threads of both CTAs arrive on the leader's barrier
    }
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc_2cta(tmem_slot, 2 * BN);
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  griddep_wait();     // PDL: everything above overlapped the previous kernel's tail

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = pair; tile < num_tiles; tile += npairs) {
        const int m_blk = tile / n_tiles, n_blk = tile % n_tiles;
        for (int kb = 0; kb < kblocks; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          uint8_t* sA = smem + stage * kStageBytes2;
          uint8_t* sB = sA + kBM * 128;
          if (rank == 0) mbar_expect_tx(&full[stage], 2 * kStageBytes2);
          if (e.a_mode == AM_PATCH16) {
            // this CTA's 128 patches = patch rows py0, py0+1 of frame b; k block kb = channel c, kernel rows ky0..ky0+3.
            // One box per kernel row: (16 kx, 64 px, 1 ky, 2 py) = 128 rows x 32 bytes, the K-major SW32 tile of one
            // K = 16 MMA step (4 KB each, 4 per stage).
            const int p0 = m_blk * 256 + (int)rank * 128;
            const int bc = (p0 >> 12) * 3 + (kb >> 2), py0 = (p0 & 4095) >> 6, ky0 = (kb & 3) * 4;
#pragma unroll
            for (int j = 0; j < 4; ++j) tma_load_5d_2cta(sA + j * 4096, &tmA, &full[stage], 0, 0, ky0 + j, py0, bc);
          } else {
            tma_load_2d_2cta(sA, &tmA, &full[stage], kb * 64, m_blk * 256 + (int)rank * 128);
          }
          tma_load_2d_2cta(sB, &tmB, &full[stage], kb * 64, n_blk * BN + (int)rank * 128);
          if (++stage == kStages2) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && rank == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int tile = pair; tile < num_tiles; tile += npairs, ++it) {
        const int as = it & 1;
        const uint32_t aph = (it >> 1) & 1;
        mbar_wait(&tmem_empty[as], aph ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = 0; kb < kblocks; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t a0 = smem_u32(smem + stage * kStageBytes2);
          const uint32_t b0 = a0 + kBM * 128;
          if (e.a_mode == AM_PATCH16) {
#pragma unroll
            for (int k = 0; k < 4; ++k)    // A: one 128 x 32 B SW32 sub-tile per K step (8-row groups 256 B apart)
              umma_ss_2cta(d_tmem, umma_desc(a0 + k * 4096, 256, 0, 6), umma_desc_sw128(b0 + k * 32), kIdesc,
                           (kb | k) != 0 ? 1u : 0u);
          } else {
#pragma unroll
            for (int k = 0; k < 4; ++k)
              umma_ss_2cta(d_tmem, umma_desc_sw128(a0 + k * 32), umma_desc_sw128(b0 + k * 32), kIdesc,
                           (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit_2cta(&empty[stage], 3);      // frees the slot in both CTAs
          if (++stage == kStages2) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit_2cta(&tmem_full[as], 3);
      }
    }
  } else if (warp == 3) {
    // GemmEpi::pad_dst: the idle warp writes the window-pad rows of the next block's qkv buffer (rows strided over CTAs)
    if (e.pad_dst) {
      const int rows = e.pad_n * e.pad_frames, cols4 = e.pad_cols >> 2;
      const float4* b4 = reinterpret_cast<const float4*>(e.pad_bias);
      for (int r = blockIdx.x; r < rows; r += gridDim.x) {
        const int frame = r / e.pad_n, rem = __ldg(e.pad_rows + (r - frame * e.pad_n));
        uint2* o = reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(e.pad_dst) +
                                            ((long)frame * kWinRowsPerFrame + rem) * e.pad_cols);
        for (int i = lane; i < cols4; i += 32) {
          const float4 v = __ldg(b4 + i);
          uint2 w;
          w.x = pack_bf16(v.x, v.y);
          w.y = pack_bf16(v.z, v.w);
          o[i] = w;
        }
      }
    }
  } else if (warp >= 4) {
    const int ew = warp & 3;
    const int eh = (warp - 4) >> 2;
    int it = 0;
    for (int tile = pair; tile < num_tiles; tile += npairs, ++it) {
      const int m_blk = tile / n_tiles, n_blk = tile % n_tiles;
      const int as = it & 1;
      const uint32_t aph = (it >> 1) & 1;
      auto wait_acc = [&]() {
        mbar_wait(&tmem_full[as], aph);
        tc_fence_after();
      };
      TmemSrc src{tmem_base + (uint32_t(ew * 32) << 16) + uint32_t(as * BN)};
      const long row = (long)m_blk * 256 + rank * 128 + ew * 32 + lane;
      (void)row;
      epilogue_warp<MODE, BN, LD>(e, src, (long)m_blk * 256 + rank * 128 + ew * 32, lane, M, N, n_blk * BN, 0, eh, 2,
                              wait_acc, reinterpret_cast<float4*>(smem + kStages2 * kStageBytes2 + 256) + (warp - 4) * 256);
      tc_fence_before();
      mbar_arrive_cluster(mapa_shared(smem_u32(&tmem_empty[as]), 0));
    }
  }

  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  if (warp == 2) tmem_dealloc_2cta(tmem_base, 2 * BN);
}

// Debug-only SIMT evaluation of the same contract (one thread per output row); selected explicitly by the
// caller to bisect tcgen05/TMA problems.  Never used unless requested.
template <bool kTf32, int MODE, int BN>
__global__ void gemm_simt_debug_kernel(const void* A, int lda, const void* W, int ldw, int M, int N, int K,
                                       GemmEpi e) {
  const long row = (long)blockIdx.x * kBM + threadIdx.x;
  const int n0 = blockIdx.y * BN;
  SimtSrc<kTf32> src{A, W, lda, ldw, K, N, n0, row, row < M};
  epilogue_row<MODE, BN>(e, src, row, M, N, n0, 0, 0, 1, NoWait());
}

// Process-wide state is limited to what really is process-wide: the driver entry point and a build-time A/B switch.
// SM counts are kept per device ordinal (a second handle on another GPU must not inherit the first one's count).
std::atomic<PFN_cuTensorMapEncodeTiled_v12000> g_encode{nullptr};
constexpr int kMaxDevices = 64;
std::atomic<int> g_num_sms[kMaxDevices];

cudaError_t make_tmap(CUtensorMap* tm, const void* ptr, bool f32, long rows, long cols, long ld, int box_rows) {
  const int esz = f32 ? 4 : 2;
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * esz};
  cuuint32_t box[2] = {(cuuint32_t)(128 / esz), (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) || (strides[0] & 15)) return cudaErrorInvalidValue;
  CUresult r = g_encode.load()(tm, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                        const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? cudaSuccess : cudaErrorInvalidValue;
}

// bf16 frames (B, 3, 1024, 1024) as (kx, px, ky, py, b*3+c): strides ascending, box = one kernel row of two patch rows
cudaError_t make_tmap_patch16(CUtensorMap* tm, const void* frames, long B) {
  cuuint64_t dims[5] = {16, 64, 16, 64, (cuuint64_t)(3 * B)};
  cuuint64_t strides[4] = {32, 2048, 32768, 2097152};
  cuuint32_t box[5] = {16, 64, 1, 2, 1};
  cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  if (reinterpret_cast<uintptr_t>(frames) & 15) return cudaErrorInvalidValue;
  CUresult r = g_encode.load()(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(frames), dims, strides, box,
                               estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_32B,
                               CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? cudaSuccess : cudaErrorInvalidValue;
}

// bf16 zero-ringed activation (B, 66, 66, 256) as (c, x, y, b); box = 64 channels of 64 x 2 positions
cudaError_t make_tmap_conv3(CUtensorMap* tm, const void* z, long B) {
  cuuint64_t dims[4] = {256, 66, 66, (cuuint64_t)B};
  cuuint64_t strides[3] = {512, 512 * 66, 512 * 66 * 66};
  cuuint32_t box[4] = {64, 64, 2, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  if (reinterpret_cast<uintptr_t>(z) & 15) return cudaErrorInvalidValue;
  CUresult r = g_encode.load()(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(z), dims, strides, box, estr,
                               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                               CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? cudaSuccess : cudaErrorInvalidValue;
}

template <bool kTf32, int MODE, int BN>
cudaError_t launch(const void* A, int lda, const void* W, int ldw, int M, int N, int K, const GemmEpi& e,
                   cudaStream_t stream, bool debug_simt) {
  if (M <= 0) return cudaSuccess;
  if (debug_simt) {
    dim3 grid((M + kBM - 1) / kBM, (N + BN - 1) / BN);
    gemm_simt_debug_kernel<kTf32, MODE, BN><<<grid, kBM, 0, stream>>>(A, lda, W, ldw, M, N, K, e);
    return cudaGetLastError();
  }
  static std::atomic<uint64_t> attr_set{0};
  cudaError_t err = smem_opt_in(gemm_tc_kernel<kTf32, MODE, BN>, TileCfg<BN>::kSmemBytes, attr_set);
  if (err != cudaSuccess) return err;
  CUtensorMap tmA, tmB;
  if (e.a_mode == AM_CONV3) {
    if (kTf32 || M % (2 * kGrid) != 0 || M % kTokens != 0 || K != 9 * 256 || e.batch != 1) return cudaErrorInvalidValue;
    err = make_tmap_conv3(&tmA, A, M / kTokens);
  } else if (e.a_mode != AM_PLAIN) {
    return cudaErrorInvalidValue;
  } else {
    err = make_tmap(&tmA, A, kTf32, M, e.batch > 1 ? lda : K, lda, kBM);
  }
  if (err != cudaSuccess) return err;
  err = make_tmap(&tmB, W, kTf32, N, K, ldw, BN);
  if (err != cudaSuccess) return err;
  const int tiles = ((M + kBM - 1) / kBM) * ((N + BN - 1) / BN) * (e.ksplit > 1 ? e.ksplit : e.batch);
  const int n_sm = gemm_num_sms();
  const int grid = tiles < n_sm ? tiles : n_sm;
  return launch_pdl(gemm_tc_kernel<kTf32, MODE, BN>, dim3(grid), dim3(kThreads), TileCfg<BN>::kSmemBytes, stream, tmA, tmB, M, N,
                    K, e);
}

template <int MODE, int LD = 0>
cudaError_t launch2(const void* A, int lda, const void* W, int ldw, int M, int N, int K, const GemmEpi& e,
                    cudaStream_t stream) {
  static std::atomic<uint64_t> attr_set{0};
  cudaError_t err = smem_opt_in(gemm_tc2_kernel<MODE, LD>, kSmemBytes2, attr_set);
  if (err != cudaSuccess) return err;
  CUtensorMap tmA, tmB;
  if (e.a_mode == AM_PATCH16) {
    if (M % kTokens != 0 || K != 3 * kPatch * kPatch) return cudaErrorInvalidValue;
    err = make_tmap_patch16(&tmA, A, M / kTokens);
  } else if (e.a_mode != AM_PLAIN) {
    return cudaErrorInvalidValue;
  } else {
    err = make_tmap(&tmA, A, false, M, K, lda, 128);
  }
  if (err != cudaSuccess) return err;
  err = make_tmap(&tmB, W, false, N, K, ldw, 128);
  if (err != cudaSuccess) return err;
  const int tiles = ((M + 255) / 256) * ((N + 255) / 256);
  int pairs = gemm_num_sms() / 2;
  if (tiles < pairs) pairs = tiles;
  return launch_pdl(gemm_tc2_kernel<MODE, LD>, dim3(2 * pairs), dim3(kThreads), kSmemBytes2, stream, tmA, tmB, M, N, K, e);
}

}  // namespace

cudaError_t gemm_tc_init(int device) {
  if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
  cudaError_t err;
  if (!g_encode.load(std::memory_order_acquire)) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    err = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (err != cudaSuccess) return err;
    if (qres != cudaDriverEntryPointSuccess || !fn) return cudaErrorNotSupported;
    g_encode.store(reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn), std::memory_order_release);
  }
  if (g_num_sms[device].load(std::memory_order_acquire) == 0) {
    int sms = 0;
    err = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    if (err != cudaSuccess) return err;
    g_num_sms[device].store(sms, std::memory_order_release);
  }
  return cudaSuccess;
}

cudaError_t make_tmap_2d(void* out, const void* ptr, int esz, long rows, long cols, long ld, int box_cols, int box_rows,
                         int swizzle_bytes) {
  if (!g_encode.load()) return cudaErrorInitializationError;
  if (box_cols * esz != swizzle_bytes || box_rows > 256) return cudaErrorInvalidValue;
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * esz};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) || (strides[0] & 15)) return cudaErrorInvalidValue;
  const CUtensorMapSwizzle sw = swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                                : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B;
  CUresult r = g_encode.load()(reinterpret_cast<CUtensorMap*>(out),
                        esz == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                        const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? cudaSuccess : cudaErrorInvalidValue;
}

// SM count of the CURRENT device (every launcher runs under the handle's DeviceGuard)
int gemm_num_sms() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return 148;
  const int n = g_num_sms[dev].load(std::memory_order_acquire);
  return n > 0 ? n : 148;
}

bool gemm_tc_is_pair(bool tf32, int mode, int M, int N, const GemmEpi& e, bool dbg) {
  return !tf32 && !dbg && e.batch == 1 && e.ksplit <= 1 && M >= 1024 && N % 256 == 0 && e.a_mode != AM_CONV3 &&
         (kernel_variant(VK_GEMM_PAIR) != 0 || e.a_mode == AM_PATCH16) &&
         (mode == GM_BIAS_BF16 || mode == GM_BIAS_GELU_BF16 || mode == GM_F32 || mode == GM_F32_FOLD);
}

cudaError_t gemm_tc(bool tf32, int mode, const void* A, int lda, const void* W, int ldw, int M, int N, int K,
                    const GemmEpi& e, cudaStream_t s, bool dbg) {
  if (!g_encode.load() && !dbg) return cudaErrorInitializationError;
  if (N % 32 != 0 || K % 8 != 0) return cudaErrorInvalidValue;
  if (e.batch != 1 && (mode != GM_F16 || tf32 || dbg || e.batch < 1)) return cudaErrorInvalidValue;
  if ((mode == GM_LN || mode == GM_UPS1) && N != 256) return cudaErrorInvalidValue;
  if (e.a_mode != AM_PLAIN && (dbg || tf32)) return cudaErrorInvalidValue;   // the SIMT bisect path reads a plain A
  if (mode == GM_F32_FOLD &&
      (tf32 || dbg || e.rowmap != RM_IDENT || e.ksplit > 1 || e.batch != 1 || !e.xb_out || !e.stats_out ||
       e.stats_slots != 2 * ((N + 255) / 256) || !e.addend || (e.add_mod > 0 && e.add_mod % 32 != 0)))
    return cudaErrorInvalidValue;
  if (e.stats_out || e.xb_out) {
    if (mode != GM_F32_FOLD) return cudaErrorInvalidValue;
  }
  if (e.stats_in) {
    if ((mode != GM_BIAS_BF16 && mode != GM_BIAS_GELU_BF16) || tf32 || dbg || !e.fold_c || !e.bias || e.stats_slots < 1 ||
        e.stats_dim < 1)
      return cudaErrorInvalidValue;
  }
  if (e.pad_dst && (!gemm_tc_is_pair(tf32, mode, M, N, e, dbg) || !e.pad_bias || !e.pad_rows || e.pad_n < 1 ||
                    e.pad_frames < 1 || e.pad_cols < 4 || e.pad_cols % 4 != 0))
    return cudaErrorInvalidValue;
  if (e.ksplit > 1) {
    const int kblocks = (K + (tf32 ? 32 : 64) - 1) / (tf32 ? 32 : 64), kper = (kblocks + e.ksplit - 1) / e.ksplit;
    if (mode != GM_F32 || dbg || e.batch != 1 || e.bias || e.addend || e.a_mode != AM_PLAIN || e.rowmap != RM_IDENT ||
        (long)(e.ksplit - 1) * kper >= kblocks)      // every K range must hold at least one block
      return cudaErrorInvalidValue;
  }
  if (e.a_mode == AM_PATCH16 && !(e.batch == 1 && M >= 1024 && N % 256 == 0 && (mode == GM_F32 || mode == GM_F32_FOLD)))
    return cudaErrorInvalidValue;
  if (mode == GM_UPS2 && N != 128) return cudaErrorInvalidValue;
  if (gemm_tc_is_pair(tf32, mode, M, N, e, dbg)) {
    switch (mode) {
      case GM_BIAS_BF16: return launch2<GM_BIAS_BF16>(A, lda, W, ldw, M, N, K, e, s);
      case GM_BIAS_GELU_BF16: return launch2<GM_BIAS_GELU_BF16>(A, lda, W, ldw, M, N, K, e, s);
      case GM_F32: return launch2<GM_F32>(A, lda, W, ldw, M, N, K, e, s);
      case GM_F32_FOLD:      // ViT-H / ViT-L / ViT-B widths get compile-time strides in the epilogue
        if (e.ld_out == e.ld_add && e.ld_out == 1280) return launch2<GM_F32_FOLD, 1280>(A, lda, W, ldw, M, N, K, e, s);
        if (e.ld_out == e.ld_add && e.ld_out == 1024) return launch2<GM_F32_FOLD, 1024>(A, lda, W, ldw, M, N, K, e, s);
        if (e.ld_out == e.ld_add && e.ld_out == 768) return launch2<GM_F32_FOLD, 768>(A, lda, W, ldw, M, N, K, e, s);
        return launch2<GM_F32_FOLD>(A, lda, W, ldw, M, N, K, e, s);
      default: break;
    }
  }
  if (!tf32) {
    switch (mode) {
      case GM_BIAS_BF16: return launch<false, GM_BIAS_BF16, 256>(A, lda, W, ldw, M, N, K, e, s, dbg);
      case GM_BIAS_GELU_BF16: return launch<false, GM_BIAS_GELU_BF16, 256>(A, lda, W, ldw, M, N, K, e, s, dbg);
      case GM_F32: return launch<false, GM_F32, 256>(A, lda, W, ldw, M, N, K, e, s, dbg);
      case GM_F32_FOLD: return launch<false, GM_F32_FOLD, 256>(A, lda, W, ldw, M, N, K, e, s, dbg);
      case GM_LN: return launch<false, GM_LN, 256>(A, lda, W, ldw, M, N, K, e, s, dbg);
      case GM_F16: return launch<false, GM_F16, 256>(A, lda, W, ldw, M, N, K, e, s, dbg);
      default: return cudaErrorInvalidValue;
    }
  }
  switch (mode) {
    case GM_F32:
      if (N <= 128) return launch<true, GM_F32, 128>(A, lda, W, ldw, M, N, K, e, s, dbg);
      return launch<true, GM_F32, 256>(A, lda, W, ldw, M, N, K, e, s, dbg);
    case GM_LN: return launch<true, GM_LN, 256>(A, lda, W, ldw, M, N, K, e, s, dbg);
    case GM_UPS1: return launch<true, GM_UPS1, 256>(A, lda, W, ldw, M, N, K, e, s, dbg);
    case GM_UPS2: return launch<true, GM_UPS2, 128>(A, lda, W, ldw, M, N, K, e, s, dbg);
    default: return cudaErrorInvalidValue;
  }
}

}  // namespace bsam
