// ViT attention with decomposed relative-position bias folded into the score tile
// (reference: Attention.forward image_encoder.py:235-260, get_rel_pos :321-351, add_decomposed_rel_pos :354-392).
//
//   A[q,k] = (q . k) * hd^-1/2 + q . Rh[qy - ky + S-1] + q . Rw[qx - kx + S-1]     (bias uses the UNSCALED q)
//   O = softmax_k(A) V                                                               (pad tokens are real keys)
//
// Flash-style: one CTA owns NW*16 query rows of one (batch item, head); K/V stream through a double-buffered
// cp.async ring in 64-key tiles; the TxT score matrix never exists in HBM (the reference materialises it and
// touches it ~5x).  The prologue computes G = Q Rh^T and Q Rw^T with tensor-core MMAs and scatters them into
// per-row bias tables; for S=64 a key tile is exactly one image row, so the h-bias is one scalar per (row, tile)
// and the w-bias is the same register fragment for every tile.
// Round-1 tensor path: mma.sync m16n8k16 bf16 (fp32 accumulate, fp32 softmax); a tcgen05 version is planned.
#include "common.cuh"
#include "kernels.cuh"

namespace bsam {

namespace {

template <int S, int HD, int NW>
struct AttnCfg {
  static constexpr int T = S * S;
  static constexpr int BQ = NW * 16;
  static constexpr int KT = 64;
  static constexpr int HDP = HD + 8;                       // padded row: odd number of 16-byte chunks
  static constexpr int RP = (S == 14) ? 32 : 128;          // padded rel-pos table rows
  static constexpr int kQBytes = BQ * HDP * 2;
  static constexpr int kKVBytes = 2 * 2 * KT * HDP * 2;    // 2 stages x (K,V)
  static constexpr int kRelHBytes = BQ * S * 4;
  static constexpr int kRelWBytes = (S == 14) ? BQ * S * 4 : 0;     // S=64: aliased onto the K/V ring
  static constexpr int kRBytes = (S == 14) ? RP * HDP * 2 : 0;      // S=64: aliased onto the K/V ring
  static constexpr int kSmem = kQBytes + kKVBytes + kRelHBytes + kRelWBytes + kRBytes;
  static_assert(S == 14 || (RP * HDP * 2 <= kKVBytes && BQ * S * 4 <= kKVBytes), "alias region too small");
};

template <int S, int HD, int NW>
__global__ void __launch_bounds__(NW * 32, 1)
attn_relpos_kernel(const __nv_bfloat16* __restrict__ qkv, const __nv_bfloat16* __restrict__ relh,
                   const __nv_bfloat16* __restrict__ relw, __nv_bfloat16* __restrict__ out, int heads, float scale) {
  using C = AttnCfg<S, HD, NW>;
  constexpr int T = C::T, BQ = C::BQ, KT = C::KT, HDP = C::HDP, RP = C::RP;
  constexpr int NKS = HD / 16;     // k-steps over the head dim
  constexpr int NDT = HD / 8;      // n-tiles of the output
  constexpr int NKT = (T + KT - 1) / KT;
  constexpr int CPR = HD / 8;      // 16-byte chunks per row

  extern __shared__ __align__(16) uint8_t smem[];
  __nv_bfloat16* sQ = reinterpret_cast<__nv_bfloat16*>(smem);
  __nv_bfloat16* sKV = reinterpret_cast<__nv_bfloat16*>(smem + C::kQBytes);
  float* sRelH = reinterpret_cast<float*>(smem + C::kQBytes + C::kKVBytes);
  float* sRelW = (S == 14) ? reinterpret_cast<float*>(smem + C::kQBytes + C::kKVBytes + C::kRelHBytes)
                           : reinterpret_cast<float*>(sKV);
  __nv_bfloat16* sR = (S == 14) ? reinterpret_cast<__nv_bfloat16*>(smem + C::kQBytes + C::kKVBytes + C::kRelHBytes + C::kRelWBytes)
                                : sKV;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, t4 = lane & 3;
  const int q0 = blockIdx.x * BQ;
  const int head = blockIdx.y;
  const long bi = blockIdx.z;
  const int ld = 3 * heads * HD;
  const __nv_bfloat16* base = qkv + bi * T * (long)ld;
  const __nv_bfloat16* gQ = base + head * HD;
  const __nv_bfloat16* gK = base + heads * HD + head * HD;
  const __nv_bfloat16* gV = base + 2 * heads * HD + head * HD;

  auto load_rows = [&](__nv_bfloat16* dst, const __nv_bfloat16* src, long src_ld, int row0, int nrows, int limit) {
    for (int i = tid; i < nrows * CPR; i += NW * 32) {
      int r = i / CPR, c = i % CPR;
      bool ok = (row0 + r) < limit;
      const __nv_bfloat16* p = src + (long)(ok ? row0 + r : 0) * src_ld + c * 8;
      cp_async16(dst + r * HDP + c * 8, p, ok);
    }
  };
  auto load_kv = [&](int stage, int kt) {
    __nv_bfloat16* dK = sKV + stage * 2 * KT * HDP;
    load_rows(dK, gK, ld, kt * KT, KT, T);
    load_rows(dK + KT * HDP, gV, ld, kt * KT, KT, T);
  };

  // ---------------- prologue: Q tile + rel-pos tables
  load_rows(sQ, gQ, ld, q0, BQ, T);
  load_rows(sR, relh, HD, 0, RP, RP);
  cp_async_commit();
  cp_async_wait<0>();
  __syncthreads();

  // this thread's two query rows
  const int rl0 = warp * 16 + g, rl1 = rl0 + 8;
  const int qi0 = min(q0 + rl0, T - 1), qi1 = min(q0 + rl1, T - 1);
  const int qy0 = qi0 / S, qx0 = qi0 % S, qy1 = qi1 / S, qx1 = qi1 % S;

  uint32_t qf[NKS][4];
  {
    const int r = warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
#pragma unroll
    for (int ks = 0; ks < NKS; ++ks)
      ldmatrix_x4(smem_u32(sQ + r * HDP + ks * 16 + (lane >> 4) * 8), qf[ks][0], qf[ks][1], qf[ks][2], qf[ks][3]);
  }

  auto rel_gemm_scatter = [&](float* table, int pos0, int pos1) {
    // G[row, r] = q[row] . R[r]; table[row][pos + S-1 - r] = G[row, r]
#pragma unroll
    for (int nt2 = 0; nt2 < RP / 16; ++nt2) {
      float acc0[4] = {0.f, 0.f, 0.f, 0.f}, acc1[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int ks = 0; ks < NKS; ++ks) {
        uint32_t b0, b1, b2, b3;
        const int rr = nt2 * 16 + (lane & 7) + ((lane >> 4) << 3);
        ldmatrix_x4(smem_u32(sR + rr * HDP + ks * 16 + ((lane >> 3) & 1) * 8), b0, b1, b2, b3);
        mma_bf16_16816(acc0, qf[ks], b0, b1);
        mma_bf16_16816(acc1, qf[ks], b2, b3);
      }
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const float* a = h ? acc1 : acc0;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int r = nt2 * 16 + h * 8 + 2 * t4 + j;
          const int k0 = pos0 + (S - 1) - r, k1 = pos1 + (S - 1) - r;
          if (r < 2 * S - 1 && k0 >= 0 && k0 < S) table[rl0 * S + k0] = a[j];
          if (r < 2 * S - 1 && k1 >= 0 && k1 < S) table[rl1 * S + k1] = a[2 + j];
        }
      }
    }
  };

  rel_gemm_scatter(sRelH, qy0, qy1);
  __syncthreads();                                   // everyone is done reading sR (Rh)
  load_rows(sR, relw, HD, 0, RP, RP);
  cp_async_commit();
  cp_async_wait<0>();
  __syncthreads();

  float bw[8][4];                                    // S=64 only: w-bias fragment, identical for every key tile
  if constexpr (S == 14) {
    rel_gemm_scatter(sRelW, qx0, qx1);
    __syncthreads();
  } else {
    // G in registers first (sR aliases sRelW), then scatter after a barrier.
    float gacc[RP / 8][4];
#pragma unroll
    for (int nt2 = 0; nt2 < RP / 16; ++nt2) {
#pragma unroll
      for (int j = 0; j < 4; ++j) gacc[2 * nt2][j] = gacc[2 * nt2 + 1][j] = 0.f;
#pragma unroll
      for (int ks = 0; ks < NKS; ++ks) {
        uint32_t b0, b1, b2, b3;
        const int rr = nt2 * 16 + (lane & 7) + ((lane >> 4) << 3);
        ldmatrix_x4(smem_u32(sR + rr * HDP + ks * 16 + ((lane >> 3) & 1) * 8), b0, b1, b2, b3);
        mma_bf16_16816(gacc[2 * nt2], qf[ks], b0, b1);
        mma_bf16_16816(gacc[2 * nt2 + 1], qf[ks], b2, b3);
      }
    }
    __syncthreads();                                 // all warps finished reading sR
#pragma unroll
    for (int nt = 0; nt < RP / 8; ++nt) {
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int r = nt * 8 + 2 * t4 + j;
        const int k0 = qx0 + (S - 1) - r, k1 = qx1 + (S - 1) - r;
        if (r < 2 * S - 1 && k0 >= 0 && k0 < S) sRelW[rl0 * S + k0] = gacc[nt][j];
        if (r < 2 * S - 1 && k1 >= 0 && k1 < S) sRelW[rl1 * S + k1] = gacc[nt][2 + j];
      }
    }
    __syncwarp();                                    // rows rl0/rl1 are written and read by the same warp only
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        bw[nt][j] = sRelW[rl0 * S + nt * 8 + 2 * t4 + j];
        bw[nt][2 + j] = sRelW[rl1 * S + nt * 8 + 2 * t4 + j];
      }
    }
    __syncthreads();                                 // alias region free for the K/V ring
  }

  // ---------------- main loop
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
  float o[NDT][4];
#pragma unroll
  for (int i = 0; i < NDT; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
  constexpr float kLog2e = 1.4426950408889634f;

  load_kv(0, 0);
  cp_async_commit();
#pragma unroll 1
  for (int kt = 0; kt < NKT; ++kt) {
    const int stage = kt & 1;
    if (kt + 1 < NKT) load_kv(stage ^ 1, kt + 1);
    cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();
    const __nv_bfloat16* sK = sKV + stage * 2 * KT * HDP;
    const __nv_bfloat16* sV = sK + KT * HDP;

    float sc[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i) sc[i][0] = sc[i][1] = sc[i][2] = sc[i][3] = 0.f;
#pragma unroll
    for (int ks = 0; ks < NKS; ++ks) {
#pragma unroll
      for (int np = 0; np < 4; ++np) {
        uint32_t b0, b1, b2, b3;
        const int key = np * 16 + (lane & 7) + ((lane >> 4) << 3);
        ldmatrix_x4(smem_u32(sK + key * HDP + ks * 16 + ((lane >> 3) & 1) * 8), b0, b1, b2, b3);
        mma_bf16_16816(sc[2 * np], qf[ks], b0, b1);
        mma_bf16_16816(sc[2 * np + 1], qf[ks], b2, b3);
      }
    }
    // scale + bias (+ mask of keys beyond T in the last tile)
    float mx0 = -INFINITY, mx1 = -INFINITY;
    if constexpr (S == 64) {
      const float bh0 = sRelH[rl0 * S + kt], bh1 = sRelH[rl1 * S + kt];
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          sc[nt][j] = sc[nt][j] * scale + bh0 + bw[nt][j];
          sc[nt][2 + j] = sc[nt][2 + j] * scale + bh1 + bw[nt][2 + j];
          mx0 = fmaxf(mx0, sc[nt][j]);
          mx1 = fmaxf(mx1, sc[nt][2 + j]);
        }
      }
    } else {
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int key = kt * KT + nt * 8 + 2 * t4 + j;
          if (key < T) {
            const int ky = key / S, kx = key - ky * S;
            sc[nt][j] = sc[nt][j] * scale + sRelH[rl0 * S + ky] + sRelW[rl0 * S + kx];
            sc[nt][2 + j] = sc[nt][2 + j] * scale + sRelH[rl1 * S + ky] + sRelW[rl1 * S + kx];
          } else {
            sc[nt][j] = -INFINITY;
            sc[nt][2 + j] = -INFINITY;
          }
          mx0 = fmaxf(mx0, sc[nt][j]);
          mx1 = fmaxf(mx1, sc[nt][2 + j]);
        }
      }
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);   // finite: every tile holds at least one valid key
    const float al0 = exp2f((m0 - mn0) * kLog2e), al1 = exp2f((m1 - mn1) * kLog2e);
    m0 = mn0;
    m1 = mn1;
    float rs0 = 0.f, rs1 = 0.f;
    uint32_t pf[4][4];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      const float p0 = exp2f((sc[nt][0] - mn0) * kLog2e), p1 = exp2f((sc[nt][1] - mn0) * kLog2e);
      const float p2 = exp2f((sc[nt][2] - mn1) * kLog2e), p3 = exp2f((sc[nt][3] - mn1) * kLog2e);
      rs0 += p0 + p1;
      rs1 += p2 + p3;
      pf[nt >> 1][(nt & 1) * 2 + 0] = pack_bf16(p0, p1);
      pf[nt >> 1][(nt & 1) * 2 + 1] = pack_bf16(p2, p3);
    }
    l0 = l0 * al0 + rs0;
    l1 = l1 * al1 + rs1;
#pragma unroll
    for (int i = 0; i < NDT; ++i) {
      o[i][0] *= al0;
      o[i][1] *= al0;
      o[i][2] *= al1;
      o[i][3] *= al1;
    }
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {          // 16 keys per step
#pragma unroll
      for (int dp = 0; dp < NDT / 2; ++dp) {
        uint32_t b0, b1, b2, b3;
        const int key = kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
        ldmatrix_x4_trans(smem_u32(sV + key * HDP + dp * 16 + (lane >> 4) * 8), b0, b1, b2, b3);
        mma_bf16_16816(o[2 * dp], pf[kk], b0, b1);
        mma_bf16_16816(o[2 * dp + 1], pf[kk], b2, b3);
      }
    }
    __syncthreads();                          // stage may be overwritten by the next prefetch
  }

  l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
  l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  const float inv0 = 1.f / l0, inv1 = 1.f / l1;
  const int ldo = heads * HD;
  __nv_bfloat16* go = out + bi * T * (long)ldo + head * HD;
  if (q0 + rl0 < T) {
    uint32_t* p = reinterpret_cast<uint32_t*>(go + (long)(q0 + rl0) * ldo);
#pragma unroll
    for (int i = 0; i < NDT; ++i) p[i * 4 + t4] = pack_bf16(o[i][0] * inv0, o[i][1] * inv0);
  }
  if (q0 + rl1 < T) {
    uint32_t* p = reinterpret_cast<uint32_t*>(go + (long)(q0 + rl1) * ldo);
#pragma unroll
    for (int i = 0; i < NDT; ++i) p[i * 4 + t4] = pack_bf16(o[i][2] * inv1, o[i][3] * inv1);
  }
}

template <int S, int HD, int NW>
cudaError_t launch_one(const __nv_bfloat16* qkv, int nbatch, int heads, const __nv_bfloat16* relh,
                       const __nv_bfloat16* relw, __nv_bfloat16* out, cudaStream_t s) {
  using C = AttnCfg<S, HD, NW>;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(attn_relpos_kernel<S, HD, NW>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmem);
    if (e != cudaSuccess) return e;
    attr = true;
  }
  dim3 grid((C::T + C::BQ - 1) / C::BQ, heads, nbatch);
  attn_relpos_kernel<S, HD, NW><<<grid, NW * 32, C::kSmem, s>>>(qkv, relh, relw, out, heads, 1.0f / sqrtf((float)HD));
  return cudaGetLastError();
}

}  // namespace

cudaError_t attention_init() { return cudaSuccess; }

cudaError_t launch_attention(const __nv_bfloat16* qkv, int nbatch, int S, int heads, int hd,
                             const __nv_bfloat16* relh, const __nv_bfloat16* relw, __nv_bfloat16* out,
                             cudaStream_t s) {
  if (nbatch <= 0) return cudaSuccess;
  if (S == 14 && (hd == 80 || hd == 64)) {
    // tcgen05 path (attn_tc.cu); the two padded tables must be contiguous: [relh (32 rows); relw (32 rows)]
    if (relw != relh + 32 * hd) return cudaErrorInvalidValue;
    return launch_attention_window_tc(qkv, nbatch, heads, hd, relh, out, s);
  }
  if (S == 64 && hd == 80) return launch_one<64, 80, 8>(qkv, nbatch, heads, relh, relw, out, s);
  if (S == 64 && hd == 64) return launch_one<64, 64, 8>(qkv, nbatch, heads, relh, relw, out, s);
  return cudaErrorInvalidValue;
}

}  // namespace bsam
