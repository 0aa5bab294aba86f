// 14x14 windowed attention with decomposed relative-position bias on tcgen05 tensor cores
// (reference: Attention.forward image_encoder.py:235-260, add_decomposed_rel_pos :354-392, window_partition :263-288).
//
//   A[q,k] = (q . k) * hd^-1/2 + q . Rh[qy - ky + 13] + q . Rw[qx - kx + 13]      (bias uses the UNSCALED q)
//   O = softmax_k(A) V                     (window-pad tokens are real keys; only the MMA tile padding is masked)
//
// One persistent CTA per SM walks the (window, head) problems.  Per problem (T = 196 tokens, two M = 128 tiles):
//   warp 0      TMA producer: Q, K (single buffers: dead once the score MMAs retire) and V (double buffered) tiles
//               straight out of the packed qkv rows; head dim 80 = one SWIZZLE_128B box of 64 + one SWIZZLE_32B box of 16
//   warp 1      tcgen05.mma issuer: G_t = Q_t [Rh;Rw]^T (rel-pos projections, N = 64), S_t = Q_t K^T (N = 208) into
//               TMEM; later O_t = P_t V with P read from TMEM (A operand) and V as an MN-major B operand
//   warps 4-11  softmax / epilogue, one thread per query row: bias gather from G, row max, exp2, P (bf16) written back
//               into the S columns of TMEM, then O * (1 / rowsum) -> bf16 -> global
// The 196x196 score matrix never leaves TMEM.
#include <cudaTypedefs.h>

#include "common.cuh"
#include "gemm_tc.cuh"
#include "kernels.cuh"

namespace bsam {

namespace {

constexpr int kT = kWinTokens;        // 196
constexpr int kRowsPad = 208;         // 13 x 16: MMA N / K granularity
constexpr int kThreadsWin = 384;
constexpr int kThreadsWin1 = 512;     // variant 1: + an epilogue warpgroup (warps 12-15)
// variant 1: 4 transposition tiles + the 1/rowsum hand-over + pad to 1024 + the second Q tile-0 buffer's 16-column
// part (head dim 80) or main part (head dim 64: no spare room in the V regions) at a 1024-byte aligned offset
constexpr int kWin1Tiles = 4 * 2048 + 2 * 128 * 4;            // 9216
template <int HD>
struct Win1Extra {
  static constexpr int value = kWin1Tiles + 1024 + (HD > 64 ? 128 * 32 : 128 * 128);
};
constexpr int kSCols = 224;           // TMEM columns reserved per M tile (S: 208 fp32; P aliases 0..103, O 112..191)
constexpr int kOCol = 112;
constexpr int kGCol = 448;            // rel-pos projections: 64 columns

template <int HD>
struct WinCfg {
  static constexpr int kRem = HD - 64;                      // 16 (ViT-H/L) or 0 (ViT-B)
  static constexpr int kMain = kRowsPad * 128;              // 26,624 = 26 x 1024
  static constexpr int kRemB = kRem ? kRowsPad * 32 : 0;    // 6,656 = 26 x 256
  static constexpr int kOperand = kMain + kRemB;
  static constexpr int kRMain = 64 * 128, kRRem = kRem ? 64 * 32 : 0;
  // shared-memory map (offsets from a 1024-byte aligned base): swizzle-128 parts first, then swizzle-32 parts
  // V's second box is a full 64-column swizzle-128 box over head columns [HD-32, HD+32): one LBO after the first it
  // lets a single N = 96 MN-major MMA per 16 keys produce the whole head (columns 64..HD-1 land in O columns
  // 96-kRem..95) instead of an N = 64 plus an N = 16 instruction - MMA issue is what bounds the P.V step
  static constexpr int kVRemB = kRem ? kMain : 0;
  static constexpr int kVOperand = kMain + kVRemB;
  static constexpr int oQm = 0, oKm = kMain, oV0m = 2 * kMain, oV1m = 3 * kMain, oV0r = 4 * kMain,
                       oV1r = oV0r + kVRemB, oRm = oV1r + kVRemB;
  static constexpr int oQr = oRm + kRMain, oKr = oQr + kRemB, oRr = oKr + kRemB;
  static constexpr int oStage = oRr + kRRem;                // 256 threads x 27 floats
  static constexpr int oBar = oStage + 256 * 27 * 4;
  static constexpr int kSmem = oBar + 256 + 1024;
};

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <int HD>
__global__ void __launch_bounds__(kThreadsWin, 1)
attn_window_tc_kernel(const __grid_constant__ CUtensorMap tm128, const __grid_constant__ CUtensorMap tm32,
                      const __grid_constant__ CUtensorMap tmR128, const __grid_constant__ CUtensorMap tmR32,
                      __nv_bfloat16* __restrict__ out, int nwin, int heads, float scale, int token_out) {
  using C = WinCfg<HD>;
  constexpr bool kHasRem = C::kRem > 0;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::oBar);
  uint64_t* qk_full = bars + 0;
  uint64_t* qk_empty = bars + 1;
  uint64_t* v_full = bars + 2;     // [2]
  uint64_t* v_empty = bars + 4;    // [2]
  uint64_t* r_full = bars + 6;
  uint64_t* s_full = bars + 7;     // [2]
  uint64_t* g_free = bars + 9;
  uint64_t* p_full = bars + 10;    // [2]
  uint64_t* o_full = bars + 12;    // [2]
  uint64_t* o_free = bars + 14;    // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nprob = nwin * heads;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm128);
    tma_prefetch_desc(&tmR128);
    if (kHasRem) {
      tma_prefetch_desc(&tm32);
      tma_prefetch_desc(&tmR32);
    }
  }
  if (warp == 1 && lane == 0) {
    mbar_init(qk_full, 1);
    mbar_init(qk_empty, 1);
    mbar_init(r_full, 1);
    mbar_init(g_free, 4);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&v_full[i], 1);
      mbar_init(&v_empty[i], 1);
      mbar_init(&s_full[i], 1);
      mbar_init(&p_full[i], 4);
      mbar_init(&o_full[i], 1);
      mbar_init(&o_free[i], 4);
    }
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(r_full, C::kRMain + C::kRRem);
      tma_load_2d(smem + C::oRm, &tmR128, r_full, 0, 0);
      if (kHasRem) tma_load_2d(smem + C::oRr, &tmR32, r_full, 64, 0);
      int it = 0;
      for (int p = blockIdx.x; p < nprob; p += gridDim.x, ++it) {
        const int pr = nprob - 1 - p;      // back to front: the qkv GEMM's last tiles are still in L2
        const int win = pr / heads, head = pr - win * heads;
        const int row0 = win * kT;
        const int cq = head * HD, ck = heads * HD + head * HD, cv = 2 * heads * HD + head * HD;
        mbar_wait(qk_empty, (it & 1) ^ 1);
        mbar_expect_tx(qk_full, 2 * C::kOperand);
        tma_load_2d(smem + C::oQm, &tm128, qk_full, cq, row0);
        tma_load_2d(smem + C::oKm, &tm128, qk_full, ck, row0);
        if (kHasRem) {
          tma_load_2d(smem + C::oQr, &tm32, qk_full, cq + 64, row0);
          tma_load_2d(smem + C::oKr, &tm32, qk_full, ck + 64, row0);
        }
        const int s = it & 1;
        mbar_wait(&v_empty[s], ((it >> 1) & 1) ^ 1);
        mbar_expect_tx(&v_full[s], C::kVOperand);
        tma_load_2d(smem + (s ? C::oV1m : C::oV0m), &tm128, &v_full[s], cv, row0);
        if (kHasRem) tma_load_2d(smem + (s ? C::oV1r : C::oV0r), &tm128, &v_full[s], cv + HD - 32, row0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idS = umma_idesc_bf16(128, kRowsPad, false, false);
      constexpr uint32_t idG = umma_idesc_bf16(128, 64, false, false);
      constexpr uint32_t idO64 = umma_idesc_bf16(128, 64, false, true);
      constexpr uint32_t idO96 = umma_idesc_bf16(128, 96, false, true);
      const uint32_t sQm = smem_u32(smem + C::oQm), sKm = smem_u32(smem + C::oKm), sRm = smem_u32(smem + C::oRm);
      const uint32_t sQr = smem_u32(smem + C::oQr), sKr = smem_u32(smem + C::oKr), sRr = smem_u32(smem + C::oRr);
      mbar_wait(r_full, 0);
      int it = 0;
      for (int p = blockIdx.x; p < nprob; p += gridDim.x, ++it) {
        const uint32_t ph = it & 1;
        mbar_wait(qk_full, ph);
#pragma unroll 1
        for (int t = 0; t < 2; ++t) {
          // tile t's TMEM columns (S/P/O of the previous problem) and the shared G columns must be drained
          mbar_wait(&o_free[t], ph ^ 1);
          if (t == 1) mbar_wait(g_free, ph);
          tc_fence_after();
          const uint32_t aM = sQm + t * 128 * 128, aR = sQr + t * 128 * 32;
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_ss<false>(tmem + kGCol, umma_desc(aM + k * 32, 1024, 0, 2), umma_desc(sRm + k * 32, 1024, 0, 2), idG,
                           k != 0);
          if (kHasRem) umma_ss<false>(tmem + kGCol, umma_desc(aR, 256, 0, 6), umma_desc(sRr, 256, 0, 6), idG, 1);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_ss<false>(tmem + t * kSCols, umma_desc(aM + k * 32, 1024, 0, 2), umma_desc(sKm + k * 32, 1024, 0, 2),
                           idS, k != 0);
          if (kHasRem)
            umma_ss<false>(tmem + t * kSCols, umma_desc(aR, 256, 0, 6), umma_desc(sKr, 256, 0, 6), idS, 1);
          umma_commit(&s_full[t]);
        }
        umma_commit(qk_empty);                       // Q and K tiles are dead once these MMAs retire
        const int s = it & 1;
        const uint32_t sVm = smem_u32(smem + (s ? C::oV1m : C::oV0m)), sVr = smem_u32(smem + (s ? C::oV1r : C::oV0r));
        mbar_wait(&v_full[s], (it >> 1) & 1);
#pragma unroll 1
        for (int t = 0; t < 2; ++t) {
          mbar_wait(&p_full[t], ph);
          tc_fence_after();
          const uint32_t dO = tmem + t * kSCols + kOCol, aP = tmem + t * kSCols;
#pragma unroll
          for (int ks = 0; ks < kRowsPad / 16; ++ks)
            umma_ts_bf16(dO, aP + ks * 8, umma_desc(sVm + ks * 2048, 1024, kHasRem ? sVr - sVm : 1024, 2),
                         kHasRem ? idO96 : idO64, ks != 0);
          umma_commit(&o_full[t]);
        }
        umma_commit(&v_empty[s]);
      }
    }
  } else if (warp >= 4) {
    const int t = (warp - 4) >> 2, q = warp & 3;
    const int r = t * 128 + q * 32 + lane;            // query row inside the window
    const bool valid = r < kT;
    const int rr = valid ? r : kT - 1;
    const int qy = rr / kWin, qx = rr - qy * kWin;
    const uint32_t tl = tmem + (uint32_t(q * 32) << 16);
    float* slot = reinterpret_cast<float*>(smem + C::oStage) + ((warp - 4) * 32 + lane) * 27;
    constexpr float kLog2e = 1.4426950408889634f;
    const float cs = scale * kLog2e;
    const long ldo = (long)heads * HD;
    int it = 0;
    for (int p = blockIdx.x; p < nprob; p += gridDim.x, ++it) {
      const uint32_t ph = it & 1;
      const int pr = nprob - 1 - p;      // back to front: the qkv GEMM's last tiles are still in L2
        const int win = pr / heads, head = pr - win * heads;
      mbar_wait(&s_full[t], ph);
      tc_fence_after();
      // ---- bias tables: bh[ky] = G[qy - ky + 13], bw[kx] = G[32 + qx - kx + 13] (pre-multiplied by log2 e)
      float bh[kWin], bw[kWin];
      {
        uint32_t g[32];
        tmem_ld_32x32(tl + kGCol, g);
        tmem_wait_ld();
#pragma unroll
        for (int j = 0; j < 27; ++j) slot[j] = __uint_as_float(g[j]);
#pragma unroll
        for (int k = 0; k < kWin; ++k) bh[k] = slot[qy + 13 - k] * kLog2e;
        tmem_ld_32x32(tl + kGCol + 32, g);
        tmem_wait_ld();
#pragma unroll
        for (int j = 0; j < 27; ++j) slot[j] = __uint_as_float(g[j]);
#pragma unroll
        for (int k = 0; k < kWin; ++k) bw[k] = slot[qx + 13 - k] * kLog2e;
      }
      if (t == 0) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(g_free);
      }
      // ---- pass 1: row maximum of the raw scores -> a safe upper bound of the biased maximum
      float mx = -INFINITY;
      {
        uint32_t v[32];
#pragma unroll
        for (int c = 0; c < 6; ++c) {
          tmem_ld_32x32(tl + t * kSCols + c * 32, v);
          tmem_wait_ld();
#pragma unroll
          for (int j = 0; j < 32; ++j) mx = fmaxf(mx, __uint_as_float(v[j]));
        }
        uint32_t w[16];
        tmem_ld_32x32_16(tl + t * kSCols + 192, w);
        tmem_wait_ld();
#pragma unroll
        for (int j = 0; j < 4; ++j) mx = fmaxf(mx, __uint_as_float(w[j]));
      }
      float bhx = -INFINITY, bwx = -INFINITY;   // scale > 0: max_k(s * cs) = cs * max_k(s)
#pragma unroll
      for (int k = 0; k < kWin; ++k) {
        bhx = fmaxf(bhx, bh[k]);
        bwx = fmaxf(bwx, bw[k]);
      }
      const float M = mx * cs + bhx + bwx;
#pragma unroll
      for (int k = 0; k < kWin; ++k) bh[k] -= M;
      // ---- pass 2: p = 2^(s*cs + bh + bw - M), row sum, P (bf16 pairs) back into the S columns
      float sum = 0.f;
      {
        uint32_t v[32];
#pragma unroll
        for (int c = 0; c < 6; ++c) {
          tmem_ld_32x32(tl + t * kSCols + c * 32, v);
          tmem_wait_ld();
          uint32_t u[16];
#pragma unroll
          for (int j = 0; j < 32; j += 2) {
            const int k0 = c * 32 + j, k1 = k0 + 1;
            const float p0 = ex2_approx(fmaf(__uint_as_float(v[j]), cs, bh[k0 / kWin]) + bw[k0 % kWin]);
            const float p1 = ex2_approx(fmaf(__uint_as_float(v[j + 1]), cs, bh[k1 / kWin]) + bw[k1 % kWin]);
            sum += p0 + p1;
            u[j >> 1] = pack_bf16(p0, p1);
          }
          tmem_st_32x32_16(tl + t * kSCols + c * 16, u);
        }
        uint32_t w[16];
        tmem_ld_32x32_16(tl + t * kSCols + 192, w);
        tmem_wait_ld();
        uint32_t u[8];
#pragma unroll
        for (int j = 0; j < 4; j += 2) {
          const int k0 = 192 + j, k1 = k0 + 1;
          const float p0 = ex2_approx(fmaf(__uint_as_float(w[j]), cs, bh[k0 / kWin]) + bw[k0 % kWin]);
          const float p1 = ex2_approx(fmaf(__uint_as_float(w[j + 1]), cs, bh[k1 / kWin]) + bw[k1 % kWin]);
          sum += p0 + p1;
          u[j >> 1] = pack_bf16(p0, p1);
        }
#pragma unroll
        for (int j = 2; j < 8; ++j) u[j] = 0u;   // keys 196..207 are MMA tile padding: P = 0
        tmem_st_32x32_8(tl + t * kSCols + 96, u);
      }
      tmem_wait_st();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_full[t]);
      // ---- epilogue: O / rowsum -> bf16 -> out[(win*196 + r), head*HD ...]
      mbar_wait(&o_full[t], ph);
      tc_fence_after();
      const float inv = 1.f / sum;
      // token_out: window_unpartition (image_encoder.py:291-318) folded into the store - rows go to b*4096 + y*64 + x and
      // the window-pad rows are dropped, so the projection GEMM runs on 4096 instead of 4900 rows per frame
      long orow = (long)win * kT + rr;
      bool vst = valid;
      if (token_out) {
        const int bfr = win / (kWinPerSide * kWinPerSide), wi = win - bfr * (kWinPerSide * kWinPerSide);
        const int y = (wi / kWinPerSide) * kWin + qy, x = (wi % kWinPerSide) * kWin + qx;
        vst = valid && y < kGrid && x < kGrid;
        orow = (long)bfr * kTokens + y * kGrid + x;
      }
      __nv_bfloat16* go = out + orow * ldo + head * HD;
      {
        uint32_t v[32];
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          tmem_ld_32x32(tl + t * kSCols + kOCol + c * 32, v);
          tmem_wait_ld();
          if (vst) {
            uint4* o4 = reinterpret_cast<uint4*>(go + c * 32);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              uint4 x;
              x.x = pack_bf16(__uint_as_float(v[8 * j + 0]) * inv, __uint_as_float(v[8 * j + 1]) * inv);
              x.y = pack_bf16(__uint_as_float(v[8 * j + 2]) * inv, __uint_as_float(v[8 * j + 3]) * inv);
              x.z = pack_bf16(__uint_as_float(v[8 * j + 4]) * inv, __uint_as_float(v[8 * j + 5]) * inv);
              x.w = pack_bf16(__uint_as_float(v[8 * j + 6]) * inv, __uint_as_float(v[8 * j + 7]) * inv);
              o4[j] = x;
            }
          }
        }
        if (kHasRem) {
          uint32_t w[16];
          tmem_ld_32x32_16(tl + t * kSCols + kOCol + 96 - C::kRem, w);
          tmem_wait_ld();
          if (vst) {
            uint4* o4 = reinterpret_cast<uint4*>(go + 64);
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              uint4 x;
              x.x = pack_bf16(__uint_as_float(w[8 * j + 0]) * inv, __uint_as_float(w[8 * j + 1]) * inv);
              x.y = pack_bf16(__uint_as_float(w[8 * j + 2]) * inv, __uint_as_float(w[8 * j + 3]) * inv);
              x.z = pack_bf16(__uint_as_float(w[8 * j + 4]) * inv, __uint_as_float(w[8 * j + 5]) * inv);
              x.w = pack_bf16(__uint_as_float(w[8 * j + 6]) * inv, __uint_as_float(w[8 * j + 7]) * inv);
              o4[j] = x;
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&o_free[t]);
    }
  }

  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 2) tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------------------------------------ shipped kernel
// (the kernel above is the round-1 version, kept as variant 9 for same-box A/B)
// Same data path and TMEM / shared-memory layout; what changes is the scheduling of the two M tiles ("units"):
//  * one MMA-issuing warp per unit, each blocking on its own unit's barriers, so tile 0 of the NEXT problem starts as
//    soon as its own TMEM columns are drained instead of queueing behind tile 1's P.V in program order (a single
//    polling issuer was measured first: its reaction time of ~1300 cycles per hand-over ate the gain);
//  * the softmax warps software-pipeline their TMEM reads (the load of chunk c+1 is in flight while chunk c is
//    processed; tcgen05.wait::ld carries the register dependency), read the bias projections with one wait and through
//    explicit shared-space stores / loads, and drain O with a single wait;
//  * the warp whose 32 rows are all tile padding (rows 224..255) does not run at all;
//  * O leaves through its own warpgroup (warps 12-15): it drains the unit's O columns into registers as soon as the
//    P.V MMAs retire, frees the unit for the next problem's scores, and only then normalises and stores (coalesced
//    through a shared-memory transposition).  Measured with the phase trace (scripts/trace_attn.py): the store phase
//    was ~2000 of the ~9000 cycles of a unit's chain when the softmax warps did it themselves.
// TRACE: CTA 0 records clock64() at the phase boundaries of its first kTraceIts problems into `trace`
// ([warp][problem][tag]); measurement builds only (include/b200sam_debug.h), the shipped instantiation has TRACE = false.
constexpr int kTraceIts = 16, kTraceTags = 10;
template <int HD, bool TRACE>
__global__ void __launch_bounds__(kThreadsWin1, 1)
attn_window_tc_kernel_v1(const __grid_constant__ CUtensorMap tm128, const __grid_constant__ CUtensorMap tm32,
                         const __grid_constant__ CUtensorMap tmR128, const __grid_constant__ CUtensorMap tmR32,
                         const __grid_constant__ CUtensorMap tmQa128, const __grid_constant__ CUtensorMap tmQa32,
                         const __grid_constant__ CUtensorMap tmQb128, const __grid_constant__ CUtensorMap tmQb32,
                         __nv_bfloat16* __restrict__ out, int nwin, int heads, float scale, int token_out,
                         unsigned long long* __restrict__ trace) {
  griddep_launch_dependents();
  auto mark = [&](int w, int it, int tag) {
    if constexpr (TRACE) {
      if (blockIdx.x == 0 && it < kTraceIts) trace[(w * kTraceIts + it) * kTraceTags + tag] = clock64();
    }
  };
  using C = WinCfg<HD>;
  constexpr bool kHasRem = C::kRem > 0;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::oBar);
  // Operand buffers of variant 1 (same regions as WinCfg lays out, used differently): Q is one buffer but each M
  // tile has its own full / empty pair (tile 0 is re-loaded as soon as ITS MMAs retire), K is double buffered (it is
  // shared by both units and freed by the later one: with a single buffer its re-load sat on the chain of the
  // earlier unit), V is a single buffer (it is needed ~3500 cycles after it can be re-loaded).
  // Tile 0 of Q is double buffered as well: it heads the chain of the leading unit (its bias projections are issued as
  // early as the shared G columns allow).
  //   K[0] = oKm / oKr, K[1] = oV1m / oV1r, V = oV0m / oV0r, Q tile 0 [1] = oV1r + 7168 / tail of the dynamic smem.
  uint64_t* q_full = bars + 0;     // [2]: tile 1 is [1]; tile 0 uses q0_full / q0_empty below
  uint64_t* q_empty = bars + 2;    // [2]
  uint64_t* v_full = bars + 4;
  uint64_t* v_empty = bars + 5;
  uint64_t* r_full = bars + 6;
  uint64_t* s_full = bars + 7;     // [2]
  uint64_t* g_read = bars + 9;     // [2] the unit's softmax warps hold the bias projections in registers
  uint64_t* p_full = bars + 11;    // [2]
  uint64_t* o_full = bars + 13;    // [2]
  uint64_t* o_free = bars + 15;    // [2]
  uint64_t* g_full = bars + 17;    // [2] bias projections of the unit are in the shared G columns
  uint64_t* k_full = bars + 19;    // [2]
  uint64_t* k_empty = bars + 21;   // [2]
  uint64_t* q0_full = bars + 23;   // [2]
  uint64_t* q0_empty = bars + 25;  // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 27);
  static_assert(C::kSmem + Win1Extra<HD>::value <= 227 * 1024, "variant 1 does not fit the shared memory of an SM");
  uint8_t* const tail = reinterpret_cast<uint8_t*>(
      (reinterpret_cast<uintptr_t>(smem + C::oBar + 256 + kWin1Tiles) + 1023) & ~uintptr_t(1023));
  uint8_t* const q0m_b = kHasRem ? smem + C::oV1r + 7168 : tail;             // 16 KB, 1024-aligned
  uint8_t* const q0r_b = tail;                                                // 4 KB (head dim 80 only)

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nprob = nwin * heads;
  const int nmine = blockIdx.x < nprob ? (nprob - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  // softmax warps with at least one real query row: 4 of tile 0, 3 of tile 1 (rows 224..255 are padding only)
  constexpr int kWarpsOfUnit1 = (kT - 128 + 31) / 32;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm128);
    tma_prefetch_desc(&tmR128);
    tma_prefetch_desc(&tmQa128);
    tma_prefetch_desc(&tmQb128);
    if (kHasRem) {
      tma_prefetch_desc(&tm32);
      tma_prefetch_desc(&tmR32);
      tma_prefetch_desc(&tmQa32);
      tma_prefetch_desc(&tmQb32);
    }
  }
  if (warp == 1 && lane == 0) {
    mbar_init(v_full, 1);
    mbar_init(v_empty, 2);
    mbar_init(r_full, 1);
    for (int i = 0; i < 2; ++i) {
      const int nw = i == 0 ? 4 : kWarpsOfUnit1;
      mbar_init(&q_full[i], 1);
      mbar_init(&q_empty[i], 1);
      mbar_init(&k_full[i], 1);
      mbar_init(&k_empty[i], 2);
      mbar_init(&q0_full[i], 1);
      mbar_init(&q0_empty[i], 1);
      mbar_init(&s_full[i], 1);
      mbar_init(&g_full[i], 1);
      mbar_init(&g_read[i], nw);
      mbar_init(&p_full[i], nw);
      mbar_init(&o_full[i], 1);
      mbar_init(&o_free[i], nw);      // the epilogue warps that own real rows of the unit
    }
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  griddep_wait();     // PDL: the set-up above overlapped the tail of the qkv GEMM / pad-row fill

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(r_full, C::kRMain + C::kRRem);
      tma_load_2d(smem + C::oRm, &tmR128, r_full, 0, 0);
      if (kHasRem) tma_load_2d(smem + C::oRr, &tmR32, r_full, 64, 0);
      // The TMA unit works through its queue in order at about one 128-byte row per 2-3 cycles (measured: a problem's
      // 1248 rows are ~3700 cycles), so the ISSUE ORDER is a priority order.  Per period: V of this problem (free once
      // the previous P.V's retired, needed ~3000 cycles later), then the next problem's Q tile 0 and K (double
      // buffered, needed first by the leading unit), then its Q tile 1 (free once this problem's tile-1 MMAs retired).
      auto coords = [&](int it, int& row0, int& cq, int& ck, int& cv) {
        const int pr = nprob - 1 - ((int)blockIdx.x + it * (int)gridDim.x);   // back to front: the qkv GEMM's last
        const int win = pr / heads, head = pr - win * heads;                  // tiles are still in L2
        row0 = win * kT;
        cq = head * HD, ck = heads * HD + head * HD, cv = 2 * heads * HD + head * HD;
      };
      auto load_q0k = [&](int it) {
        int row0, cq, ck, cv;
        coords(it, row0, cq, ck, cv);
        const int kb = it & 1;
        mbar_wait(&q0_empty[kb], ((it >> 1) & 1) ^ 1);
        mark(0, it, 0);
        mbar_expect_tx(&q0_full[kb], 128 * 128 + (kHasRem ? 128 * 32 : 0));
        tma_load_2d(kb ? q0m_b : smem + C::oQm, &tmQa128, &q0_full[kb], cq, row0);
        if (kHasRem) tma_load_2d(kb ? q0r_b : smem + C::oQr, &tmQa32, &q0_full[kb], cq + 64, row0);
        mbar_wait(&k_empty[kb], ((it >> 1) & 1) ^ 1);
        mbar_expect_tx(&k_full[kb], C::kOperand);
        tma_load_2d(smem + (kb ? C::oV1m : C::oKm), &tm128, &k_full[kb], ck, row0);
        if (kHasRem) tma_load_2d(smem + (kb ? C::oV1r : C::oKr), &tm32, &k_full[kb], ck + 64, row0);
      };
      auto load_q1 = [&](int it) {
        int row0, cq, ck, cv;
        coords(it, row0, cq, ck, cv);
        mbar_wait(&q_empty[1], (it & 1) ^ 1);
        mbar_expect_tx(&q_full[1], (kRowsPad - 128) * 128 + (kHasRem ? (kRowsPad - 128) * 32 : 0));
        tma_load_2d(smem + C::oQm + 128 * 128, &tmQb128, &q_full[1], cq, row0 + 128);
        if (kHasRem) tma_load_2d(smem + C::oQr + 128 * 32, &tmQb32, &q_full[1], cq + 64, row0 + 128);
      };
      if (nmine > 0) {
        load_q0k(0);
        load_q1(0);
      }
      for (int it = 0; it < nmine; ++it) {
        int row0, cq, ck, cv;
        coords(it, row0, cq, ck, cv);
        mbar_wait(v_empty, (it & 1) ^ 1);
        mark(0, it, 1);
        mbar_expect_tx(v_full, C::kVOperand);
        tma_load_2d(smem + C::oV0m, &tm128, v_full, cv, row0);
        if (kHasRem) tma_load_2d(smem + C::oV0r, &tm128, v_full, cv + HD - 32, row0);
        if (it + 1 < nmine) {
          load_q0k(it + 1);
          load_q1(it + 1);
          // the single-buffered operands are re-loaded behind MMAs of the problem before: pull the boxes of problem
          // it + 2 into L2 behind the demand loads (optional, A/B switch)
          if ((token_out & 4) && it + 2 < nmine) {
            int rown, cqn, ckn, cvn;
            coords(it + 2, rown, cqn, ckn, cvn);
            tma_prefetch_2d(&tmQb128, cqn, rown + 128);
            if (kHasRem) tma_prefetch_2d(&tmQb32, cqn + 64, rown + 128);
            tma_prefetch_2d(&tm128, cvn, rown);
            if (kHasRem) tma_prefetch_2d(&tm128, cvn + HD - 32, rown);
          }
        }
      }
    }
  } else if (warp == 1 || warp == 3) {
    // One MMA-issuing warp PER UNIT (warp 1: tile 0, warp 3: tile 1): each runs its own in-order stream
    // S(u,it) -> P.V(u,it) -> S(u,it+1) and blocks on its own unit's barriers, so neither unit queues behind the
    // other.  What the units share is ordered by barriers: the single-buffered G columns (g_read of the other unit),
    // K (k_empty counts both units' score MMAs) and the V tile (v_empty counts both P.V's).
    // All 32 lanes run the loop in warp-uniform code and one elected lane issues the tcgen05 instructions: descriptors
    // and TMEM addresses then live in uniform registers and the MMAs issue back to back.  Inside an `if (lane == 0)`
    // region every operand goes through an ELECT / R2UR.BROADCAST loop (~100 cycles per MMA, 23 MMAs per problem and
    // unit: more than the tensor time of those MMAs).
    {
      const bool leader = elect_one();
      const int u = (int)uniform_u32(warp == 1 ? 0u : 1u);
      const uint32_t tm = uniform_u32(tmem);
      constexpr uint32_t idS = umma_idesc_bf16(128, kRowsPad, false, false);
      constexpr uint32_t idG = umma_idesc_bf16(128, 64, false, false);
      constexpr uint32_t idO64 = umma_idesc_bf16(128, 64, false, true);
      constexpr uint32_t idO96 = umma_idesc_bf16(128, 96, false, true);
      const uint32_t sQm = smem_u32(smem + C::oQm), sRm = smem_u32(smem + C::oRm);
      const uint32_t sQr = smem_u32(smem + C::oQr), sRr = smem_u32(smem + C::oRr);
      const uint32_t dS = tm + u * kSCols, dO = dS + kOCol;
      // A operand: tile 1 lives behind tile 0 in the single Q buffer; tile 0 alternates between two buffers
      auto q_main = [&](int it) { return u ? sQm + 128 * 128 : ((it & 1) ? smem_u32(q0m_b) : sQm); };
      auto q_rem = [&](int it) { return u ? sQr + 128 * 32 : ((it & 1) ? smem_u32(q0r_b) : sQr); };
      // Bias projections G(u, it).  They only need Q and the shared G columns - not the unit's own TMEM columns - so
      // they are issued one step ahead, between the score MMAs and the P.V MMAs of the PREVIOUS problem, and the softmax
      // warps gather their bias tables while P.V, the O drain and the score MMAs of this problem run.
      auto issue_G = [&](int it) {
        const uint32_t ph = it & 1;
        const uint32_t aM = q_main(it), aR = q_rem(it);
        if (u == 0) mbar_wait(&q0_full[it & 1], (it >> 1) & 1);
        else mbar_wait(&q_full[1], ph);
        if (lane == 0) mark(warp, it, 2);
        if (u == 0) {
          if (it > 0) mbar_wait(&g_read[1], ph ^ 1);     // tile 1 of the previous problem holds its projections
        } else {
          mbar_wait(&g_read[0], ph);                     // tile 0 of this problem holds its projections
        }
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const uint64_t da = umma_desc(aM + k * 32, 1024, 0, 2), db = umma_desc(sRm + k * 32, 1024, 0, 2);
          if (leader) umma_ss<false>(tm + kGCol, da, db, idG, k != 0);
        }
        if (kHasRem) {
          const uint64_t da = umma_desc(aR, 256, 0, 6), db = umma_desc(sRr, 256, 0, 6);
          if (leader) umma_ss<false>(tm + kGCol, da, db, idG, 1);
        }
        if (leader) umma_commit(&g_full[u]);
        if (lane == 0) mark(warp, it, 4);
      };
      mbar_wait(r_full, 0);
      if (nmine > 0) issue_G(0);
#pragma unroll 1
      for (int it = 0; it < nmine; ++it) {
        const uint32_t ph = it & 1;
        const int kb = it & 1;
        const uint32_t sKm = smem_u32(smem + (kb ? C::oV1m : C::oKm)), sKr = smem_u32(smem + (kb ? C::oV1r : C::oKr));
        const uint32_t aM = q_main(it), aR = q_rem(it);
        mbar_wait(&k_full[kb], (it >> 1) & 1);
        mbar_wait(&o_free[u], ph ^ 1);                   // the unit's S / P / O columns are drained
        tc_fence_after();
        if (lane == 0) mark(warp, it, 3);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const uint64_t da = umma_desc(aM + k * 32, 1024, 0, 2), db = umma_desc(sKm + k * 32, 1024, 0, 2);
          if (leader) umma_ss<false>(dS, da, db, idS, k != 0);
        }
        if (kHasRem) {
          const uint64_t da = umma_desc(aR, 256, 0, 6), db = umma_desc(sKr, 256, 0, 6);
          if (leader) umma_ss<false>(dS, da, db, idS, 1);
        }
        if (leader) {
          umma_commit(&s_full[u]);
          umma_commit(u ? &q_empty[1] : &q0_empty[kb]);  // this tile of Q is dead once these MMAs retire
          umma_commit(&k_empty[kb]);                     // count 2: K is dead once both units' score MMAs retire
        }
        if (lane == 0) mark(warp, it, 0);
        if (it + 1 < nmine) issue_G(it + 1);
        const uint32_t sVm = smem_u32(smem + C::oV0m), sVr = smem_u32(smem + C::oV0r);
        const uint32_t lbo = kHasRem ? sVr - sVm : 1024;
        mbar_wait(v_full, ph);
        // (P.V cannot start on a first part of P while the softmax warps still read scores: O aliases the score
        // columns 112..207 of the unit)
        mbar_wait(&p_full[u], ph);
        tc_fence_after();
        if (lane == 0) mark(warp, it, 5);
#pragma unroll
        for (int ks = 0; ks < kRowsPad / 16; ++ks) {
          const uint64_t db = umma_desc(sVm + ks * 2048, 1024, lbo, 2);
          if (leader) umma_ts_bf16(dO, dS + ks * 8, db, kHasRem ? idO96 : idO64, ks != 0);
        }
        if (leader) {
          umma_commit(&o_full[u]);
          umma_commit(v_empty);                          // count 2
        }
        if (lane == 0) mark(warp, it, 1);
      }
    }
  } else if (TRACE && warp == 2) {
    if (lane == 0) {   // observer: when do the TMA loads land, when do the score MMAs retire?
      for (int it = 0; it < nmine && it < kTraceIts; ++it) {
        mbar_wait(&q0_full[it & 1], (it >> 1) & 1);
        mark(2, it, 0);
        mbar_wait(&s_full[0], it & 1);
        mark(2, it, 2);
        mbar_wait(&s_full[1], it & 1);
        mark(2, it, 3);
      }
    } else if (lane == 1) {
      for (int it = 0; it < nmine && it < kTraceIts; ++it) {
        mbar_wait(v_full, it & 1);
        mark(2, it, 1);
      }
    } else if (lane == 2) {
      for (int it = 0; it < nmine && it < kTraceIts; ++it) {
        mbar_wait(&o_full[0], it & 1);
        mark(2, it, 4);
        mbar_wait(&o_full[1], it & 1);
        mark(2, it, 5);
      }
    }
  } else if (warp >= 4 && (warp < 8 || warp - 8 < kWarpsOfUnit1)) {
    const int t = (warp - 4) >> 2, q = warp & 3;
    const int r = t * 128 + q * 32 + lane;            // query row inside the window
    const bool valid = r < kT;
    const int rr = valid ? r : kT - 1;
    const int qy = rr / kWin, qx = rr - qy * kWin;
    const uint32_t tl = tmem + (uint32_t(q * 32) << 16);
    const uint32_t tS = tl + t * kSCols;
    const uint32_t slot = smem_u32(smem + C::oStage) + ((warp - 4) * 32 + lane) * 27 * 4;
    float* inv_s = reinterpret_cast<float*>(smem + C::oBar + 256 + 4 * 2048);
    constexpr float kLog2e = 1.4426950408889634f;
    const float cs = scale * kLog2e;
    for (int it = 0; it < nmine; ++it) {
      const uint32_t ph = it & 1;
      // ---- bias projections -> registers (two loads, one wait), then the unit's G columns are free again
      uint32_t va[32], vb[32];
      mbar_wait(&g_full[t], ph);
      tc_fence_after();
      if (lane == 0) mark(warp, it, 0);
      tmem_ld_32x32(tl + kGCol, va);
      tmem_ld_32x32(tl + kGCol + 32, vb);
      tmem_wait_ld_dep(va);
      tmem_wait_ld_dep(vb);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&g_read[t]);
      // bias tables: bh[ky] = G[qy - ky + 13], bw[kx] = G[32 + qx - kx + 13] (pre-multiplied by log2 e), gathered
      // through a private shared-memory slot (the row's shift is a run-time index)
      float bh[kWin], bw[kWin];
#pragma unroll
      for (int j = 0; j < 27; ++j) sts32(slot + 4 * j, va[j]);
#pragma unroll
      for (int k = 0; k < kWin; ++k) bh[k] = lds32(slot + 4 * (qy + 13 - k)) * kLog2e;
#pragma unroll
      for (int j = 0; j < 27; ++j) sts32(slot + 4 * j, vb[j]);
#pragma unroll
      for (int k = 0; k < kWin; ++k) bw[k] = lds32(slot + 4 * (qx + 13 - k)) * kLog2e;
      if (lane == 0) mark(warp, it, 1);
      mbar_wait(&s_full[t], ph);
      tc_fence_after();
      if (lane == 0) mark(warp, it, 2);
      // ---- pass 1: row maximum of the raw scores -> a safe upper bound of the biased maximum.  Software pipeline:
      // the load of chunk c+1 is in flight while chunk c is reduced.
      float mx = -INFINITY;
      uint32_t w[16];
      tmem_ld_32x32(tS, va);
#pragma unroll
      for (int c = 0; c < 6; ++c) {
        uint32_t(&cur)[32] = (c & 1) ? vb : va;
        uint32_t(&nxt)[32] = (c & 1) ? va : vb;
        tmem_wait_ld_dep(cur);
        if (c < 5) tmem_ld_32x32(tS + (c + 1) * 32, nxt);
        else tmem_ld_32x32_16(tS + 192, w);
#pragma unroll
        for (int j = 0; j < 32; j += 2)
          mx = fmaxf(mx, fmaxf(__uint_as_float(cur[j]), __uint_as_float(cur[j + 1])));
      }
      tmem_wait_ld_dep16(w);
#pragma unroll
      for (int j = 0; j < 4; ++j) mx = fmaxf(mx, __uint_as_float(w[j]));
      tmem_ld_32x32(tS, va);                 // first chunk of pass 2
      float bhx = -INFINITY, bwx = -INFINITY;   // scale > 0: max_k(s * cs) = cs * max_k(s)
#pragma unroll
      for (int k = 0; k < kWin; ++k) {
        bhx = fmaxf(bhx, bh[k]);
        bwx = fmaxf(bwx, bw[k]);
      }
      const float M = mx * cs + bhx + bwx;
      // Packed operands of pass 2.  Keys are handled in pairs (k, k+1) with k even; a key row holds 14 keys, so a pair
      // never straddles two key rows: one (bh, bh) pair per key row, one (bw[2i], bw[2i+1]) pair per column pair.
      uint64_t bhp[kWin], bwp[kWin / 2];
#pragma unroll
      for (int k = 0; k < kWin; ++k) bhp[k] = f32x2(bh[k] - M, bh[k] - M);
#pragma unroll
      for (int k = 0; k < kWin / 2; ++k) bwp[k] = f32x2(bw[2 * k], bw[2 * k + 1]);
      const uint64_t cs2 = f32x2(cs, cs);
      if (lane == 0) mark(warp, it, 3);
      // ---- pass 2: p = 2^(s*cs + bh + bw - M), row sum, P (bf16 pairs) back into the S columns.  fp32x2 arithmetic
      // (half the issue slots), and two key pairs of every five take the FMA-pipe polynomial instead of MUFU.EX2
      // (kPolyOf5): the special-function unit is the contended resource of this phase.
      uint64_t sum2 = f32x2(0.f, 0.f);
      auto exp_pair = [&](uint32_t s0, uint32_t s1, int k0) -> uint32_t {
        uint64_t tt = fma2(f32x2(__uint_as_float(s0), __uint_as_float(s1)), cs2, bhp[k0 / kWin]);
        tt = add2(tt, bwp[(k0 % kWin) >> 1]);
        const int pi = (k0 >> 1) % 5;
        const uint64_t e = (pi == 1 || pi == 3) ? ex2_poly2(tt) : ex2_mufu2(tt);
        sum2 = add2(sum2, e);
        float e0, e1;
        unpack2(e, e0, e1);
        return pack_bf16(e0, e1);
      };
#pragma unroll
      for (int c = 0; c < 6; ++c) {
        uint32_t(&cur)[32] = (c & 1) ? vb : va;
        uint32_t(&nxt)[32] = (c & 1) ? va : vb;
        tmem_wait_ld_dep(cur);
        if (c < 5) tmem_ld_32x32(tS + (c + 1) * 32, nxt);
        else tmem_ld_32x32_16(tS + 192, w);
        uint32_t u[16];
#pragma unroll
        for (int j = 0; j < 32; j += 2) u[j >> 1] = exp_pair(cur[j], cur[j + 1], c * 32 + j);
        tmem_st_32x32_16(tS + c * 16, u);      // columns [16c, 16c+16) hold scores of chunk c/2: already consumed
      }
      {
        tmem_wait_ld_dep16(w);
        uint32_t u[8];
#pragma unroll
        for (int j = 0; j < 4; j += 2) u[j >> 1] = exp_pair(w[j], w[j + 1], 192 + j);
#pragma unroll
        for (int j = 2; j < 8; ++j) u[j] = 0u;   // keys 196..207 are MMA tile padding: P = 0
        tmem_st_32x32_8(tS + 96, u);
      }
      float sum0, sum1;
      unpack2(sum2, sum0, sum1);
      if (lane == 0) mark(warp, it, 4);
      inv_s[t * 128 + q * 32 + lane] = 1.f / (sum0 + sum1);   // read by the epilogue warps behind p_full
      tmem_wait_st();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_full[t]);
      if (lane == 0) mark(warp, it, 5);
    }
  } else if (warp >= 12) {
    // ---- epilogue warpgroup: O / rowsum -> bf16 -> out.  token_out: window_unpartition (image_encoder.py:291-318)
    // folded into the store - rows go to b*4096 + y*64 + x and the window-pad rows are dropped
    const int q = warp & 3;
    const uint32_t tl = tmem + (uint32_t(q * 32) << 16);
    const uint32_t wtile = smem_u32(smem + C::oBar + 256) + q * 2048;
    const float* inv_s = reinterpret_cast<const float*>(smem + C::oBar + 256 + 4 * 2048);
    const long ldo = (long)heads * HD;
    const int sub = lane >> 2, cg = lane & 3;         // transposed role: rows i*8 + sub, 16-byte unit cg of the chunk
    int it = 0;
    for (int p = blockIdx.x; p < nprob; p += gridDim.x, ++it) {
      const uint32_t ph = it & 1;
      const int pr = nprob - 1 - p;
      const int win = pr / heads, head = pr - win * heads;
#pragma unroll 1
      for (int t = 0; t < 2; ++t) {
        if (t == 1 && q >= kWarpsOfUnit1) break;      // rows 224..255 do not exist
        const int r = t * 128 + q * 32 + lane;
        const bool valid = r < kT;
        const int rr = valid ? r : kT - 1;
        const int qy = rr / kWin, qx = rr - qy * kWin;
        long orow = (long)win * kT + rr;
        bool vst = valid;
        if (token_out & 1) {
          const int bfr = win / (kWinPerSide * kWinPerSide), wi = win - bfr * (kWinPerSide * kWinPerSide);
          const int y = (wi / kWinPerSide) * kWin + qy, x = (wi % kWinPerSide) * kWin + qx;
          vst = valid && y < kGrid && x < kGrid;
          orow = (long)bfr * kTokens + y * kGrid + x;
        }
        const int orow_i = vst ? (int)orow : -1;
        const uint32_t tO = tl + t * kSCols + kOCol;
        uint32_t va[32], vb[32], w[16];
        mbar_wait(&p_full[t], ph);                    // orders the read of 1/rowsum behind the softmax warps' writes
        const float inv = inv_s[t * 128 + q * 32 + lane];
        mbar_wait(&o_full[t], ph);
        tc_fence_after();
        if (lane == 0) mark(warp, it, 2 * t);
        tmem_ld_32x32(tO, va);
        tmem_ld_32x32(tO + 32, vb);
        if (kHasRem) tmem_ld_32x32_16(tO + 96 - C::kRem, w);
        tmem_wait_ld_dep(va);
        tmem_wait_ld_dep(vb);
        if (kHasRem) tmem_wait_ld_dep16(w);
        // all of O sits in registers: the unit's TMEM columns can take the next problem's scores right away
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&o_free[t]);
        if (lane == 0) mark(warp, it, 2 * t + 1);
        // A thread owns one query ROW; stored directly, every 16-byte store instruction of a warp would touch 32
        // different 128-byte lines.  Each 32-column chunk is transposed through a 2 KB tile (16-byte units
        // XOR-swizzled by the row, conflict free both ways): four lanes cover 64 contiguous bytes of a row.
        auto store_chunk = [&](const uint32_t* v, int ncols, int col0) {
          __syncwarp();
          const int nunits = ncols / 8;               // 16-byte units per row in this chunk (4 or 2)
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (j < nunits) {
              const uint32_t x0 = pack_bf16(__uint_as_float(v[8 * j + 0]) * inv, __uint_as_float(v[8 * j + 1]) * inv);
              const uint32_t x1 = pack_bf16(__uint_as_float(v[8 * j + 2]) * inv, __uint_as_float(v[8 * j + 3]) * inv);
              const uint32_t x2 = pack_bf16(__uint_as_float(v[8 * j + 4]) * inv, __uint_as_float(v[8 * j + 5]) * inv);
              const uint32_t x3 = pack_bf16(__uint_as_float(v[8 * j + 6]) * inv, __uint_as_float(v[8 * j + 7]) * inv);
              sts128u(wtile + lane * 64 + ((j ^ ((lane >> 1) & 3)) << 4), x0, x1, x2, x3);
            }
          }
          __syncwarp();
          if (ncols == 32) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const int rw = i * 8 + sub;
              const uint4 x = lds128u(wtile + rw * 64 + ((cg ^ ((rw >> 1) & 3)) << 4));
              const int orw = __shfl_sync(0xffffffffu, orow_i, rw);
              if (orw >= 0 && !(TRACE && (token_out & 2)))
                *reinterpret_cast<uint4*>(out + (long)orw * ldo + head * HD + col0 + cg * 8) = x;
            }
          } else {
#pragma unroll
            for (int i = 0; i < 2; ++i) {
              const int rw = i * 16 + (lane >> 1), c2 = lane & 1;
              const uint4 x = lds128u(wtile + rw * 64 + ((c2 ^ ((rw >> 1) & 3)) << 4));
              const int orw = __shfl_sync(0xffffffffu, orow_i, rw);
              if (orw >= 0 && !(TRACE && (token_out & 2)))
                *reinterpret_cast<uint4*>(out + (long)orw * ldo + head * HD + col0 + c2 * 8) = x;
            }
          }
        };
        store_chunk(va, 32, 0);
        store_chunk(vb, 32, 32);
        if (kHasRem) store_chunk(w, 16, 64);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 2) tmem_dealloc(tmem, 512);
}

template <int HD>
cudaError_t launch_window(const __nv_bfloat16* qkv, int nwin, int heads, const __nv_bfloat16* rel, __nv_bfloat16* out,
                          bool token_out, cudaStream_t s) {
  using C = WinCfg<HD>;
  static std::atomic<uint64_t> attr{0};
  cudaError_t e = smem_opt_in(attn_window_tc_kernel<HD>, C::kSmem, attr);
  if (e != cudaSuccess) return e;
  const long rows = (long)nwin * kT;
  const long cols = 3L * heads * HD;
  CUtensorMap tm128, tm32, tmR128, tmR32;
  e = make_tmap_2d(&tm128, qkv, 2, rows, cols, cols, 64, kRowsPad, 128);
  if (e != cudaSuccess) return e;
  e = make_tmap_2d(&tmR128, rel, 2, 64, HD, HD, 64, 64, 128);
  if (e != cudaSuccess) return e;
  if (C::kRem) {
    e = make_tmap_2d(&tm32, qkv, 2, rows, cols, cols, 16, kRowsPad, 32);
    if (e != cudaSuccess) return e;
    e = make_tmap_2d(&tmR32, rel, 2, 64, HD, HD, 16, 64, 32);
    if (e != cudaSuccess) return e;
  } else {
    tm32 = tm128;
    tmR32 = tmR128;
  }
  // variant 1 loads the two M tiles of Q separately: 128-row and 80-row boxes
  CUtensorMap tmQa128, tmQa32, tmQb128, tmQb32;
  e = make_tmap_2d(&tmQa128, qkv, 2, rows, cols, cols, 64, 128, 128);
  if (e == cudaSuccess) e = make_tmap_2d(&tmQb128, qkv, 2, rows, cols, cols, 64, kRowsPad - 128, 128);
  if (e != cudaSuccess) return e;
  if (C::kRem) {
    e = make_tmap_2d(&tmQa32, qkv, 2, rows, cols, cols, 16, 128, 32);
    if (e == cudaSuccess) e = make_tmap_2d(&tmQb32, qkv, 2, rows, cols, cols, 16, kRowsPad - 128, 32);
    if (e != cudaSuccess) return e;
  } else {
    tmQa32 = tmQa128;
    tmQb32 = tmQb128;
  }
  const int nprob = nwin * heads;
  const int grid = nprob < gemm_num_sms() ? nprob : gemm_num_sms();
  const float scale = 1.0f / sqrtf((float)HD);
  // Variants (include/b200sam_debug.h): 0 = shipped kernel (attn_window_tc_kernel_v1), 4 = same with the L2 prefetch of
  // the single-buffered operands, 2 / 3 = with the phase trace (3: without the global stores), 9 = the round-1 kernel,
  // kept for same-box A/B (scripts/bench_attn.py).
  const int variant = kernel_variant(VK_ATTN_WINDOW);
  switch (variant) {
    case 9:
      attn_window_tc_kernel<HD><<<grid, kThreadsWin, C::kSmem, s>>>(tm128, tm32, tmR128, tmR32, out, nwin, heads, scale,
                                                                  token_out ? 1 : 0);
      break;
    case 4:
    case 0: {
      static std::atomic<uint64_t> attr1{0};
      e = smem_opt_in(attn_window_tc_kernel_v1<HD, false>, C::kSmem + Win1Extra<HD>::value, attr1);
      if (e != cudaSuccess) return e;
      e = launch_pdl(attn_window_tc_kernel_v1<HD, false>, dim3(grid), dim3(kThreadsWin1), C::kSmem + Win1Extra<HD>::value, s,
                     tm128, tm32, tmR128, tmR32, tmQa128, tmQa32, tmQb128, tmQb32, out, nwin, heads, scale,
                     (token_out ? 1 : 0) | (variant == 4 ? 4 : 0), (unsigned long long*)nullptr);
      if (e != cudaSuccess) return e;
      break;
    }
    case 3:     // variant 2 without the global stores of O (measurement only: how much of the chain is the store?)
    case 2: {   // the shipped kernel with the phase trace of CTA 0 (b200sam_debug_set_trace)
      static std::atomic<uint64_t> attr2{0};
      if (!debug_trace_buffer()) return cudaErrorInvalidValue;
      e = smem_opt_in(attn_window_tc_kernel_v1<HD, true>, C::kSmem + Win1Extra<HD>::value, attr2);
      if (e != cudaSuccess) return e;
      attn_window_tc_kernel_v1<HD, true><<<grid, kThreadsWin1, C::kSmem + Win1Extra<HD>::value, s>>>(
          tm128, tm32, tmR128, tmR32, tmQa128, tmQa32, tmQb128, tmQb32, out, nwin, heads, scale,
          (token_out ? 1 : 0) | (variant == 3 ? 2 : 0),
          reinterpret_cast<unsigned long long*>(debug_trace_buffer()));
      break;
    }
    default: return cudaErrorInvalidValue;   // unknown variant: an error, never a fallback
  }
  return cudaGetLastError();
}

}  // namespace

// qkv: (nwin*196, 3*heads*hd) bf16; rel: (64, hd) bf16 = [rel_pos_h (27 rows, zero padded to 32); rel_pos_w (same)].
cudaError_t launch_attention_window_tc(const __nv_bfloat16* qkv, int nwin, int heads, int hd, const __nv_bfloat16* rel,
                                       __nv_bfloat16* out, cudaStream_t s, bool token_out) {
  if (nwin <= 0) return cudaSuccess;
  if (hd == 80) return launch_window<80>(qkv, nwin, heads, rel, out, token_out, s);
  if (hd == 64) return launch_window<64>(qkv, nwin, heads, rel, out, token_out, s);
  return cudaErrorInvalidValue;
}

}  // namespace bsam
