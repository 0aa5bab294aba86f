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
  const int nprob = nwin * heads;
  const int grid = nprob < gemm_num_sms() ? nprob : gemm_num_sms();
  attn_window_tc_kernel<HD><<<grid, kThreadsWin, C::kSmem, s>>>(tm128, tm32, tmR128, tmR32, out, nwin, heads,
                                                              1.0f / sqrtf((float)HD), token_out ? 1 : 0);
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
