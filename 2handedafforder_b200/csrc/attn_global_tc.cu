// Global (64x64 tokens) attention with decomposed relative-position bias on tcgen05 tensor cores
// (reference: Attention.forward image_encoder.py:235-260, add_decomposed_rel_pos :354-392).
//
//   A[q,k] = (q . k) * hd^-1/2 + q . Rh[qy - ky + 63] + q . Rw[qx - kx + 63]      (bias uses the UNSCALED q)
//   O = softmax_k(A) V
//
// Flash-style, the 4096 x 4096 score matrix never exists in HBM (the reference materialises it, 1 GB fp32 per frame,
// and touches it ~5x).  A persistent CTA owns 256 query rows (two M = 128 tiles that ping-pong on the tensor core)
// of one (frame, head) and streams 128-key tiles:
//   warp 0      TMA producer: Q tiles once per work item; K and V tiles through two 3-stage rings
//               (head dim 80: Q / K = one SWIZZLE_128B box of 64 columns + one SWIZZLE_32B box of 16;
//               V = two SWIZZLE_128B boxes, columns [0,64) and [16,80), consumed by one N = 128 MMA)
//   warp 1      tcgen05.mma issuer: S_t = Q_t K_j^T into TMEM, then O_t += P_t V_j with P read from TMEM and V as an
//               MN-major B operand; S_t(j+1) is issued right behind it so the tensor core works while the other
//               tile's softmax runs
//   warps 4-7 / 8-11   softmax of tile 0 / tile 1, one thread per query row: a 128-key tile is two image rows, so
//               the h-bias is two scalars per tile; the w-bias is added on the tensor core (S_t += Gw_t E with a
//               one-hot E, see GlbCfg); lazy rescaling (O and the row sum are
//               only rescaled when the running maximum grows by more than 2^8); P (bf16) goes back into the S columns
// The bias projections G = Q [Rh; Rw]^T come from a batched tcgen05 GEMM (gemm_tc GM_F16) as fp16.
#include <cudaTypedefs.h>

#include "common.cuh"
#include "gemm_tc.cuh"
#include "kernels.cuh"

namespace bsam {

namespace {

constexpr int kGT = kTokens;          // 4096
constexpr int kBKeys = 128;           // keys per tile = two image rows
constexpr int kNIter = kGT / kBKeys;  // 32
constexpr int kQBlocks = kGT / 256;   // 16 work items per (frame, head)
constexpr int kStagesK = 2;   // K tiles in flight (20 KB each)
constexpr int kStagesV = 3;   // V tiles in flight (32 KB each)
constexpr int kThreadsG = 384;
constexpr int kOColG = 256;           // O_t at kOColG + t * 128

template <int HD>
struct GlbCfg {
  static constexpr int kRem = HD - 64;
  static constexpr int kMain = 128 * 128;                 // one 128-row tile, 64 columns: 16 KB
  static constexpr int kRemB = kRem ? 128 * 32 : 0;       // 16 columns: 4 KB
  static constexpr int kTile = kMain + kRemB;
  // V's second box is 64 columns wide too (head columns [HD-64, HD)): with the two boxes one LBO apart a single
  // N = 128 MN-major MMA per 16 keys covers the whole head (columns 64..HD-1 land in O columns 128-kRem..127) instead
  // of an N = 64 and an N = 16 instruction - the single issuing thread is the bottleneck of the P.V step
  static constexpr int kVRemB = kRem ? kMain : 0;
  static constexpr int kVTile = kMain + kVRemB;
  // w-bias through the tensor core: S_t += Gw_t E, Gw_t = the 128 x 64 fp16 tile of the row's w-biases (stored by the
  // G GEMM as q . Rw[qx - kx + 63] * sqrt(hd), kx < 64) and E the constant 64 x 128 one-hot matrix E[kx][n] = (n % 64
  // == kx): exact in fp32, and the softmax threads keep no per-column bias registers.  The two halves of E are equal,
  // so only 64 rows are kept and each half of the score tile gets its own N = 64 MMA on the same operand.
  static constexpr int oQm = 0;                           // 2 tiles
  static constexpr int oGw = 2 * kMain;                   // 2 tiles (fp16)
  static constexpr int oE = oGw + 2 * kMain;              // 64 rows (n % 64) x 64 (kx) fp16, K-major, swizzle-128: 8 KB
  static constexpr int oKm = oE + kMain;                  // kStagesK tiles
  static constexpr int oVm = oKm + kStagesK * kMain;      // kStagesV tiles
  static constexpr int oVr = oVm + kStagesV * kMain;      // kStagesV second boxes
  static constexpr int oQr = oVr + kStagesV * kVRemB;
  static constexpr int oKr = oQr + 2 * kRemB;
  static constexpr int oBar = oKr + kStagesK * kRemB;
  static constexpr int kSmem = oBar + 512 + 1024;
};

__device__ __forceinline__ float ex2_fast(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// POLY: fp32x2 softmax arithmetic with two key pairs of every five on the FMA-pipe exp2 polynomial (ex2_poly2)
// instead of MUFU.EX2 - the special-function unit (8 pipe cycles per warp instruction) is the co-limiter of this kernel.
// TRACE: CTA 0 records clock64() at the phase boundaries of the first 16 key tiles of its first item
// ([warp][tile j][tag], include/b200sam_debug.h); the shipped instantiation has TRACE = false.
template <int HD, bool POLY, bool TRACE>
__global__ void __launch_bounds__(kThreadsG, 1)
attn_global_tc_kernel(const __grid_constant__ CUtensorMap tm128, const __grid_constant__ CUtensorMap tm32,
                      const __grid_constant__ CUtensorMap tmG,
                      const __half* __restrict__ G, __nv_bfloat16* __restrict__ out, int nbatch, int heads,
                      float scale, unsigned long long* __restrict__ trace) {
  griddep_launch_dependents();
  auto mark = [&](int w, int item_it, int j, int tag) {
    if constexpr (TRACE) {
      if (blockIdx.x == 0 && item_it == 1 && j < 16) trace[(w * 16 + j) * 10 + tag] = clock64();
    }
  };
  using C = GlbCfg<HD>;
  constexpr bool kHasRem = C::kRem > 0;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::oBar);
  uint64_t* q_full = bars + 0;
  uint64_t* q_empty = bars + 1;
  uint64_t* k_full = bars + 2;                    // [kStagesK]
  uint64_t* k_empty = k_full + kStagesK;
  uint64_t* v_full = k_empty + kStagesK;          // [kStagesV]
  uint64_t* v_empty = v_full + kStagesV;
  uint64_t* s_full = v_empty + kStagesV;          // [2]
  uint64_t* p_full = s_full + 2;                  // [2]
  uint64_t* o_full = p_full + 2;                  // [2]
  uint64_t* p_half = o_full + 2;                  // [2] POLY variant: P of the first 64 keys of the tile is stored
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(p_half + 2);

  const int warp = (int)uniform_u32(threadIdx.x >> 5), lane = threadIdx.x & 31;
  const int nitems = nbatch * heads * kQBlocks;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm128);
    if (kHasRem) tma_prefetch_desc(&tm32);
    tma_prefetch_desc(&tmG);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(q_full, 1);
    mbar_init(q_empty, 1);
    for (int i = 0; i < kStagesK; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&k_empty[i], 1);
    }
    for (int i = 0; i < kStagesV; ++i) {
      mbar_init(&v_full[i], 1);
      mbar_init(&v_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s_full[i], 1);
      mbar_init(&p_full[i], 4);
      mbar_init(&p_half[i], 4);
      mbar_init(&o_full[i], 1);
    }
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 512);
  // one-hot B operand of the w-bias MMA: row n (key inside the tile) holds fp16 1.0 at column n % 64
  // (K-major, 128-byte rows, 16-byte chunk index XOR row % 8 = the SWIZZLE_128B pattern the descriptor expects)
  for (int i = threadIdx.x; i < C::kMain / 32; i += blockDim.x)
    reinterpret_cast<uint4*>(smem + C::oE)[i] = make_uint4(0u, 0u, 0u, 0u);
  __syncthreads();
  if (threadIdx.x < 64) {
    const int n = threadIdx.x, kx = n;
    *reinterpret_cast<uint16_t*>(smem + C::oE + n * 128 + (((kx >> 3) ^ (n & 7)) << 4) + (kx & 7) * 2) = 0x3C00;
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  griddep_wait();     // PDL: barrier / TMEM / one-hot operand set-up overlapped the tail of the rel-pos GEMM
  // register re-allocation between the warpgroups: the softmax threads keep a 128-key score row + 64 bias values
  // (setmaxnreg at the top of every role branch)

  if (warp == 0) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
    if (lane == 0) {
      int it = 0, kc = 0;
      for (int item = blockIdx.x; item < nitems; item += gridDim.x, ++it) {
        const int ir = nitems - 1 - item;   // back to front: the qkv GEMM's last tiles are still in L2
        const int bh = ir / kQBlocks, qb = ir - bh * kQBlocks;
        const int b = bh / heads, head = bh - b * heads;
        const int row0 = b * kGT;
        const int cq = head * HD, ck = heads * HD + head * HD, cv = 2 * heads * HD + head * HD;
        mbar_wait(q_empty, (it & 1) ^ 1);
        mbar_expect_tx(q_full, 2 * C::kTile + 2 * C::kMain);
        for (int t = 0; t < 2; ++t) {
          tma_load_2d(smem + C::oGw + t * C::kMain, &tmG, q_full, head * 256 + 128, row0 + qb * 256 + t * 128);
          tma_load_2d(smem + C::oQm + t * C::kMain, &tm128, q_full, cq, row0 + qb * 256 + t * 128);
          if (kHasRem) tma_load_2d(smem + C::oQr + t * C::kRemB, &tm32, q_full, cq + 64, row0 + qb * 256 + t * 128);
        }
        for (int j = 0; j < kNIter; ++j, ++kc) {
          const int sk = kc % kStagesK;
          mbar_wait(&k_empty[sk], ((kc / kStagesK) & 1) ^ 1);
          mbar_expect_tx(&k_full[sk], C::kTile);
          tma_load_2d(smem + C::oKm + sk * C::kMain, &tm128, &k_full[sk], ck, row0 + j * kBKeys);
          if (kHasRem) tma_load_2d(smem + C::oKr + sk * C::kRemB, &tm32, &k_full[sk], ck + 64, row0 + j * kBKeys);
          const int s = kc % kStagesV;
          mbar_wait(&v_empty[s], ((kc / kStagesV) & 1) ^ 1);
          mbar_expect_tx(&v_full[s], C::kVTile);
          tma_load_2d(smem + C::oVm + s * C::kMain, &tm128, &v_full[s], cv, row0 + j * kBKeys);
          if (kHasRem) tma_load_2d(smem + C::oVr + s * C::kVRemB, &tm128, &v_full[s], cv + HD - 64, row0 + j * kBKeys);
        }
      }
    }
  } else if (warp == 1) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
    // All 32 lanes run the loops (warp-uniform control flow, uniform-register descriptors); one elected lane issues
    // the tcgen05 instructions.
    const bool leader = elect_one();
    const uint32_t tm = uniform_u32(tmem);
    constexpr uint32_t idS = umma_idesc_bf16(128, kBKeys, false, false);
    constexpr uint32_t idO64 = umma_idesc_bf16(128, 64, false, true);
    constexpr uint32_t idO128 = umma_idesc_bf16(128, 128, false, true);
    // fp16 x fp16 -> fp32, both K-major, M = 128, N = 64 (a_format = b_format = 0)
    constexpr uint32_t idB = (1u << 4) | (uint32_t(64 >> 3) << 17) | (uint32_t(128 >> 4) << 24);
    const uint32_t sGw = smem_u32(smem + C::oGw), sE = smem_u32(smem + C::oE);
    const uint32_t sQm = smem_u32(smem + C::oQm), sQr = smem_u32(smem + C::oQr);
    const uint32_t sKm = smem_u32(smem + C::oKm), sKr = smem_u32(smem + C::oKr);
    const uint32_t sVm = smem_u32(smem + C::oVm), sVr = smem_u32(smem + C::oVr);
    const uint32_t lbo = kHasRem ? (sVr - sVm) : 1024;   // V main box -> second box of the same stage
    auto issue_S = [&](int t, int ks) {
      const uint32_t aM = sQm + t * C::kMain, aR = sQr + t * C::kRemB;
      const uint32_t bM = sKm + ks * C::kMain, bR = sKr + ks * C::kRemB;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint64_t da = umma_desc(aM + k * 32, 1024, 0, 2), db = umma_desc(bM + k * 32, 1024, 0, 2);
        if (leader) umma_ss<false>(tm + t * 128, da, db, idS, k != 0);
      }
      if (kHasRem) {
        const uint64_t da = umma_desc(aR, 256, 0, 6), db = umma_desc(bR, 256, 0, 6);
        if (leader) umma_ss<false>(tm + t * 128, da, db, idS, 1);
      }
      // + w-bias: S_t += Gw_t E
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint64_t da = umma_desc(sGw + t * C::kMain + k * 32, 1024, 0, 2), db = umma_desc(sE + k * 32, 1024, 0, 2);
        if (leader) {
          umma_ss<false>(tm + t * 128, da, db, idB, 1);
          umma_ss<false>(tm + t * 128 + 64, da, db, idB, 1);
        }
      }
    };
    int it = 0, kc = 0;          // kc: index of the K/V tile of iteration 0 of this item in the rings
    uint32_t np0 = 0, np1 = 0;   // P tiles consumed per M tile (parity of p_full)
    for (int item = blockIdx.x; item < nitems; item += gridDim.x, ++it, kc += kNIter) {
      mbar_wait(q_full, it & 1);
      {
        const int s = kc % kStagesK;
        mbar_wait(&k_full[s], (kc / kStagesK) & 1);
        tc_fence_after();
        issue_S(0, s);
        if (leader) umma_commit(&s_full[0]);
        issue_S(1, s);
        if (leader) {
          umma_commit(&s_full[1]);
          umma_commit(&k_empty[s]);
        }
      }
#pragma unroll 1
      for (int j = 0; j < kNIter; ++j) {
        const int sv = (kc + j) % kStagesV;
        const int sk = (kc + j + 1) % kStagesK;
        const bool more = j + 1 < kNIter;
        if (more) mbar_wait(&k_full[sk], ((kc + j + 1) / kStagesK) & 1);
        mbar_wait(&v_full[sv], ((kc + j) / kStagesV) & 1);
#pragma unroll
        for (int t = 0; t < 2; ++t) {
          const uint32_t dO = tm + kOColG + t * 128, aP = tm + t * 128;
          const uint32_t bM = sVm + sv * C::kMain;
          if (lane == 0) mark(1, it, j, 4 * t);
          if constexpr (POLY) {
            // P.V in two steps (O does not alias the scores here): the first four key groups start as soon as their P
            // is stored, while the softmax warps still work on the second half of the tile
            mbar_wait(&p_half[t], (t ? np1 : np0) & 1);
            tc_fence_after();
#pragma unroll
            for (int ks = 0; ks < kBKeys / 32; ++ks) {
              const uint64_t db = umma_desc(bM + ks * 2048, 1024, lbo, 2);
              if (leader) umma_ts_bf16(dO, aP + ks * 8, db, kHasRem ? idO128 : idO64, (j | ks) != 0);
            }
          }
          mbar_wait(&p_full[t], (t ? np1 : np0) & 1);
          if (t) ++np1; else ++np0;
          tc_fence_after();
          if (lane == 0) mark(1, it, j, 4 * t + 1);
#pragma unroll
          for (int ks = POLY ? kBKeys / 32 : 0; ks < kBKeys / 16; ++ks) {
            const uint64_t db = umma_desc(bM + ks * 2048, 1024, lbo, 2);
            if (leader) umma_ts_bf16(dO, aP + ks * 8, db, kHasRem ? idO128 : idO64, (j | ks) != 0);
          }
          if (t == 1 && leader) umma_commit(&v_empty[sv]);
          if (more) {
            issue_S(t, sk);
            if (leader) {
              umma_commit(&s_full[t]);
              if (t == 1) {
                umma_commit(&k_empty[sk]);
                if (j + 2 == kNIter) umma_commit(q_empty);   // last score MMAs of this item: Q tiles are dead
              }
            }
          } else {
            if (leader) umma_commit(&o_full[t]);
          }
          if (lane == 0) mark(1, it, j, 4 * t + 2);
        }
      }
    }
  } else if (warp < 4) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
    const int t = (warp - 4) >> 2, q = warp & 3;
    const int rl = q * 32 + lane;                       // row inside the M tile
    const uint32_t tl = tmem + (uint32_t(q * 32) << 16);
    const uint32_t tS = tl + t * 128, tO = tl + kOColG + t * 128;
    constexpr float kLog2e = 1.4426950408889634f;
    const float cs = scale * kLog2e;
    const long ldg = (long)heads * 256, ldo = (long)heads * HD;
    uint32_t ns = 0;                                     // score tiles consumed (parity of s_full[t])
    int it = 0;
    for (int item = blockIdx.x; item < nitems; item += gridDim.x, ++it) {
      const int ir = nitems - 1 - item;
      const int bhh = ir / kQBlocks, qb = ir - bhh * kQBlocks;
      const int b = bhh / heads, head = bhh - b * heads;
      const int qrow = qb * 256 + t * 128 + rl;          // token index inside the frame
      const int qy = qrow >> 6;
      const __half* g = G + ((long)b * kGT + qrow) * ldg + head * 256;
      const __half* gh = g + qy + 63;                    // h-bias of key row ky: gh[-ky]
      float bh0 = __half2float(__ldg(gh)) * kLog2e, bh1 = __half2float(__ldg(gh - 1)) * kLog2e;
      float m_used = -INFINITY, l = 0.f;
#pragma unroll 1
      for (int j = 0; j < kNIter; ++j) {
        // prefetch the next tile's two h-bias values
        float nb0 = 0.f, nb1 = 0.f;
        if (j + 1 < kNIter) {
          nb0 = __half2float(__ldg(gh - 2 * (j + 1)));
          nb1 = __half2float(__ldg(gh - 2 * (j + 1) - 1));
        }
        mbar_wait(&s_full[t], ns & 1);
        ++ns;
        tc_fence_after();
        if (lane == 0) mark(warp, it, j, 0);
        // ---- the whole 128-key score row of this thread in registers (one TMEM round trip per tile)
        uint32_t v[128];
#pragma unroll
        for (int c = 0; c < 4; ++c) tmem_ld_32x32(tS + c * 32, *reinterpret_cast<uint32_t(*)[32]>(&v[32 * c]));
        tmem_wait_ld();
        if (lane == 0) mark(warp, it, j, 1);
        // maximum of the raw scores -> upper bound of the biased tile maximum
        float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
        for (int i = 0; i < 128; i += 4) {
          mx0 = fmaxf(mx0, __uint_as_float(v[i]));
          mx1 = fmaxf(mx1, __uint_as_float(v[i + 1]));
          mx2 = fmaxf(mx2, __uint_as_float(v[i + 2]));
          mx3 = fmaxf(mx3, __uint_as_float(v[i + 3]));
        }
        const float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
        const float Mt = mx * cs + fmaxf(bh0, bh1);   // the scores already hold the w-bias (in units of 1/cs)
        const bool grow = Mt > m_used + 8.f;
        if (__any_sync(0xffffffffu, grow)) {
          const float m_new = grow ? Mt : m_used;
          const float f = ex2_fast(m_used - m_new);     // 0 on the first tile (m_used = -inf), 1 for unchanged rows
          l *= f;
          m_used = m_new;
          if (j > 0) {
            uint32_t o[32];
#pragma unroll
            for (int c = 0; c < 2; ++c) {
              tmem_ld_32x32(tO + c * 32, o);
              tmem_wait_ld();
              uint32_t a[16], bq[16];
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                a[i] = __float_as_uint(__uint_as_float(o[i]) * f);
                bq[i] = __float_as_uint(__uint_as_float(o[16 + i]) * f);
              }
              tmem_st_32x32_16(tO + c * 32, a);
              tmem_st_32x32_16(tO + c * 32 + 16, bq);
            }
            if (kHasRem) {
              uint32_t w[16];
              tmem_ld_32x32_16(tO + 128 - C::kRem, w);
              tmem_wait_ld();
#pragma unroll
              for (int i = 0; i < 16; ++i) w[i] = __float_as_uint(__uint_as_float(w[i]) * f);
              tmem_st_32x32_16(tO + 128 - C::kRem, w);
            }
          }
        }
        const float c0 = bh0 - m_used, c1 = bh1 - m_used;
        if (lane == 0) mark(warp, it, j, 2);
        // ---- p = 2^(s*cs + bh - m) (s holds q.k + w-bias), row sum (4 partial chains), P (bf16 pairs) into the
        // first 64 columns of S_t
        float sum;
        if constexpr (POLY) {
          uint64_t acc2 = f32x2(0.f, 0.f);
          const uint64_t cs2 = f32x2(cs, cs);
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const float cc = c < 2 ? c0 : c1;
            const uint64_t cc2 = f32x2(cc, cc);
            uint32_t u[16];
#pragma unroll
            for (int i = 0; i < 32; i += 2) {
              const uint64_t tt = fma2(f32x2(__uint_as_float(v[32 * c + i]), __uint_as_float(v[32 * c + i + 1])), cs2, cc2);
              const int pi = ((32 * c + i) >> 1) % 5;
              const uint64_t e = (pi == 1 || pi == 3) ? ex2_poly2(tt) : ex2_mufu2(tt);
              acc2 = add2(acc2, e);
              float e0, e1;
              unpack2(e, e0, e1);
              u[i >> 1] = pack_bf16(e0, e1);
            }
            tmem_st_32x32_16(tS + c * 16, u);
            if (c == 1) {                        // P of keys 0..63 is complete
              tmem_wait_st();
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive(&p_half[t]);
            }
          }
          float a0, a1;
          unpack2(acc2, a0, a1);
          sum = a0 + a1;
        } else {
        float sum0 = 0.f, sum1 = 0.f, sum2 = 0.f, sum3 = 0.f;
  #pragma unroll
          for (int c = 0; c < 4; ++c) {
            const float cc = c < 2 ? c0 : c1;
            uint32_t u[16];
  #pragma unroll
            for (int i = 0; i < 32; i += 4) {
              const float p0 = ex2_fast(fmaf(__uint_as_float(v[32 * c + i]), cs, cc));
              const float p1 = ex2_fast(fmaf(__uint_as_float(v[32 * c + i + 1]), cs, cc));
              const float p2 = ex2_fast(fmaf(__uint_as_float(v[32 * c + i + 2]), cs, cc));
              const float p3 = ex2_fast(fmaf(__uint_as_float(v[32 * c + i + 3]), cs, cc));
              sum0 += p0;
              sum1 += p1;
              sum2 += p2;
              sum3 += p3;
              u[i >> 1] = pack_bf16(p0, p1);
              u[(i >> 1) + 1] = pack_bf16(p2, p3);
            }
            tmem_st_32x32_16(tS + c * 16, u);
          }
          sum = (sum0 + sum1) + (sum2 + sum3);
        }
        l += sum;
        if (lane == 0) mark(warp, it, j, 3);
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_full[t]);
        if (lane == 0) mark(warp, it, j, 4);
        bh0 = nb0 * kLog2e;
        bh1 = nb1 * kLog2e;
      }
      // ---- epilogue: O / l -> bf16
      mbar_wait(&o_full[t], it & 1);
      tc_fence_after();
      const float inv = 1.f / l;
      __nv_bfloat16* go = out + ((long)b * kGT + qrow) * ldo + head * HD;
      {
        uint32_t v[32];
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          tmem_ld_32x32(tO + c * 32, v);
          tmem_wait_ld();
          uint4* o4 = reinterpret_cast<uint4*>(go + c * 32);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            uint4 x;
            x.x = pack_bf16(__uint_as_float(v[8 * i + 0]) * inv, __uint_as_float(v[8 * i + 1]) * inv);
            x.y = pack_bf16(__uint_as_float(v[8 * i + 2]) * inv, __uint_as_float(v[8 * i + 3]) * inv);
            x.z = pack_bf16(__uint_as_float(v[8 * i + 4]) * inv, __uint_as_float(v[8 * i + 5]) * inv);
            x.w = pack_bf16(__uint_as_float(v[8 * i + 6]) * inv, __uint_as_float(v[8 * i + 7]) * inv);
            o4[i] = x;
          }
        }
        if (kHasRem) {
          uint32_t w[16];
          tmem_ld_32x32_16(tO + 128 - C::kRem, w);
          tmem_wait_ld();
          uint4* o4 = reinterpret_cast<uint4*>(go + 64);
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            uint4 x;
            x.x = pack_bf16(__uint_as_float(w[8 * i + 0]) * inv, __uint_as_float(w[8 * i + 1]) * inv);
            x.y = pack_bf16(__uint_as_float(w[8 * i + 2]) * inv, __uint_as_float(w[8 * i + 3]) * inv);
            x.z = pack_bf16(__uint_as_float(w[8 * i + 4]) * inv, __uint_as_float(w[8 * i + 5]) * inv);
            x.w = pack_bf16(__uint_as_float(w[8 * i + 6]) * inv, __uint_as_float(w[8 * i + 7]) * inv);
            o4[i] = x;
          }
        }
      }
      tc_fence_before();   // orders these TMEM reads before the next item's p_full arrive -> PV(0) overwrite
    }
  }

  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 2) tmem_dealloc(tmem, 512);
}

template <int HD>
cudaError_t launch_global(const __nv_bfloat16* qkv, int nbatch, int heads, const __half* G, __nv_bfloat16* out,
                          cudaStream_t s) {
  using C = GlbCfg<HD>;
  // 0 = shipped (fp32x2 + polynomial-exp2 softmax, P.V in two steps), 1 = the round-1 arithmetic (MUFU.EX2 only, A/B),
  // 2 = shipped kernel with the phase trace (include/b200sam_debug.h)
  const int variant = kernel_variant(VK_ATTN_GLOBAL);
  if (variant < 0 || variant > 2 || (variant == 2 && !debug_trace_buffer())) return cudaErrorInvalidValue;
  const long rows = (long)nbatch * kGT, cols = 3L * heads * HD;
  CUtensorMap tm128, tm32;
  cudaError_t e = make_tmap_2d(&tm128, qkv, 2, rows, cols, cols, 64, 128, 128);
  if (e != cudaSuccess) return e;
  if (C::kRem) {
    e = make_tmap_2d(&tm32, qkv, 2, rows, cols, cols, 16, 128, 32);
    if (e != cudaSuccess) return e;
  } else {
    tm32 = tm128;
  }
  CUtensorMap tmG;   // fp16 bias projections (rows, heads * 256): the 64 w-biases of a row start at column 128
  e = make_tmap_2d(&tmG, G, 2, rows, (long)heads * 256, (long)heads * 256, 64, 128, 128);
  if (e != cudaSuccess) return e;
  const int nitems = nbatch * heads * kQBlocks;
  const int grid = nitems < gemm_num_sms() ? nitems : gemm_num_sms();
  const float scale = 1.0f / sqrtf((float)HD);
  unsigned long long* tr = reinterpret_cast<unsigned long long*>(debug_trace_buffer());
#define BSAM_LAUNCH_GLOBAL(POLY, TRACE, TRACEPTR)                                                                  \
  do {                                                                                                             \
    static std::atomic<uint64_t> attr{0};                                                                          \
    e = smem_opt_in(attn_global_tc_kernel<HD, POLY, TRACE>, C::kSmem, attr);                                       \
    if (e != cudaSuccess) return e;                                                                                \
    e = launch_pdl(attn_global_tc_kernel<HD, POLY, TRACE>, dim3(grid), dim3(kThreadsG), C::kSmem, s, tm128, tm32, tmG, G, \
                   out, nbatch, heads, scale, (unsigned long long*)(TRACEPTR));                                     \
    if (e != cudaSuccess) return e;                                                                                \
  } while (0)
  if (variant == 2) BSAM_LAUNCH_GLOBAL(true, true, tr);
  else if (variant == 1) BSAM_LAUNCH_GLOBAL(false, false, nullptr);
  else BSAM_LAUNCH_GLOBAL(true, false, nullptr);
#undef BSAM_LAUNCH_GLOBAL
  return cudaGetLastError();
}

}  // namespace

// qkv: (nbatch*4096, 3*heads*hd) bf16; G: (nbatch*4096, heads*256) fp16 = per head [q.Rh[0..126], 0, q.Rw[0..126], 0].
cudaError_t launch_attention_global_tc(const __nv_bfloat16* qkv, int nbatch, int heads, int hd, const __half* G,
                                       __nv_bfloat16* out, cudaStream_t s) {
  if (nbatch <= 0) return cudaSuccess;
  if (hd == 80) return launch_global<80>(qkv, nbatch, heads, G, out, s);
  if (hd == 64) return launch_global<64>(qkv, nbatch, heads, G, out, s);
  return cudaErrorInvalidValue;
}

}  // namespace bsam
