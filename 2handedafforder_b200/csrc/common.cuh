// Shared device helpers for the B200 (sm_100a) SAM-branch kernels: PTX wrappers for
// mbarrier, TMA (cp.async.bulk.tensor), tcgen05 (MMA / TMEM alloc / TMEM load / commit),
// cp.async, ldmatrix and mma.sync, plus small math helpers.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>

namespace bsam {

// Opt a kernel in to more than 48 KB of dynamic shared memory ONCE PER DEVICE: the attribute belongs to the
// (function, device context) pair, so a process-wide flag would leave the second GPU of a process without it.
// `done` is a per-kernel bitmask indexed by the current device ordinal.
template <class Kernel>
inline cudaError_t smem_opt_in(Kernel kernel, int bytes, std::atomic<uint64_t>& done) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  const uint64_t bit = 1ull << (dev & 63);
  if (done.load(std::memory_order_acquire) & bit) return cudaSuccess;
  e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess) done.fetch_or(bit, std::memory_order_release);
  return e;
}

// Fixed geometry of the path: 1024^2 frames, 16x16 patches, 14x14 windows
// (reference: 2Haff/model/segment_anything/build_sam.py:67-84).
constexpr int kImg = 1024;
constexpr int kPatch = 16;
constexpr int kGrid = 64;            // tokens per side
constexpr int kTokens = kGrid * kGrid;
constexpr int kWin = 14;
constexpr int kWinPerSide = 5;       // ceil(64/14)
constexpr int kWinTokens = kWin * kWin;                       // 196
constexpr int kWinRowsPerFrame = kWinPerSide * kWinPerSide * kWinTokens;  // 4900
constexpr int kOutCh = 256;

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// ---------------------------------------------------------------- mbarrier
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}

// non-blocking probe of a phase (event-driven issue loops poll several barriers instead of blocking on one)
__device__ __forceinline__ bool mbar_test(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n.reg .pred p;\n"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}

// ---------------------------------------------------------------- TMA
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* tm) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(tm)) : "memory");
}
// Programmatic dependent launch (PDL): a kernel launched with cudaLaunchAttributeProgrammaticStreamSerialization may
// start while its predecessor in the stream still runs.  launch_dependents lets the successor's CTAs be scheduled as
// soon as every CTA of this grid is resident; wait blocks until the predecessor grid has completed and its memory is
// visible.  Everything before the wait (barrier init, TMEM allocation, descriptor prefetch) overlaps the predecessor's
// tail; no global memory that another kernel writes or reads may be touched before it.  Both are no-ops for a kernel
// launched without the attribute.
__device__ __forceinline__ void griddep_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// 2-D tiled load: coordinates are (inner = element column, outer = row).
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* tm, uint64_t* bar, int c_inner,
                                            int c_outer) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar)), "r"(c_inner),
      "r"(c_outer)
      : "memory");
}

// 4-D tiled load (coordinates innermost first): the 3x3 neck convolution reads its A tiles straight from the
// zero-padded (B, 66, 66, 256) activation, one (64 channels, 64 x, 2 y) box per tap (gemm_tc.cu, AM_CONV3).
__device__ __forceinline__ void tma_load_4d(void* smem_dst, const CUtensorMap* tm, uint64_t* bar, int c0, int c1,
                                            int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2),
      "r"(c3)
      : "memory");
}

// L2 prefetch of the box a later tma_load_2d will fetch (no shared memory, no barrier): turns the demand load of a
// single-buffered operand into an L2 hit.
__device__ __forceinline__ void tma_prefetch_2d(const CUtensorMap* tm, int c_inner, int c_outer) {
  asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global.tile [%0, {%1, %2}];" ::"l"(reinterpret_cast<uint64_t>(tm)),
               "r"(c_inner), "r"(c_outer)
               : "memory");
}

// ---------------------------------------------------------------- tcgen05 / TMEM
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)),
               "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// Same wait, with the destination registers of the outstanding tcgen05.ld tied to it as read-write operands: in a
// software-pipelined loop (load chunk c+1, process chunk c, wait, process chunk c+1) nothing but this dependency keeps
// the compiler from scheduling a use of the asynchronously written registers above the wait.
__device__ __forceinline__ void tmem_wait_ld_dep(uint32_t (&r)[32]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                 "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]),
                 "+r"(r[16]), "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]),
                 "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31])
               :
               : "memory");
}
__device__ __forceinline__ void tmem_wait_ld_dep16(uint32_t (&r)[16]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                 "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])
               :
               : "memory");
}

// D[tmem] (+)= A[smem desc] * B[smem desc]; kind::f16 covers bf16/fp16 operands, kind::tf32 fp32 operands.
template <bool kTf32>
__device__ __forceinline__ void umma_ss(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                        uint32_t accumulate) {
  if constexpr (kTf32) {
    asm volatile(
        "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
  } else {
    asm volatile(
        "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
  }
}
// Arrive on an mbarrier once all previously issued tcgen05.mma of this thread have completed.
// One lane of a converged warp (elect.sync).  The MMA warps run their loops with all 32 lanes and wrap only the
// tcgen05 instructions in `if (leader)`: descriptors and TMEM addresses computed in warp-uniform code stay in uniform
// registers, whereas inside an `if (lane == 0)` region every operand goes through an ELECT / R2UR.BROADCAST waterfall
// (about a dozen dependent instructions per MMA, which bounds kernels that issue many small MMAs).
__device__ __forceinline__ bool elect_one() {
  uint32_t p;
  asm volatile("{\n.reg .pred e;\nelect.sync _|e, 0xffffffff;\nselp.u32 %0, 1, 0, e;\n}\n" : "=r"(p));
  return p != 0;
}
// warp-uniform copy of a value (lane 0's): lets the compiler prove uniformity
__device__ __forceinline__ uint32_t uniform_u32(uint32_t v) { return __shfl_sync(0xffffffffu, v, 0); }
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
// 32 lanes x 32 consecutive fp32 columns: thread i of the warp gets lane (base_lane + i), columns c..c+31.
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}

// ---------------------------------------------------------------- CTA pairs (cta_group::2, cluster of 2)
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `p` (a shared::cta address of this CTA) in the CTA of rank `rank`
__device__ __forceinline__ uint32_t mapa_shared(uint32_t p, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(p), "r"(rank));
  return r;
}
// Remote arrive on a peer CTA's barrier.  Default semantics (release at CTA scope): the consumers only need the TMEM
// reads ordered (tcgen05.fence does that); a cluster-scope release would drain every outstanding global store first.
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void tmem_alloc_2cta(uint32_t* smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)),
               "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_2cta(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// 2-D tiled load issued by either CTA of a pair; the transaction bytes are credited to the mbarrier of the
// pair's leader (cluster address with the peer bit cleared).
__device__ __forceinline__ void tma_load_2d_2cta(void* smem_dst, const CUtensorMap* tm, uint64_t* bar, int c_inner,
                                                 int c_outer) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], "
      "[%2];" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar) & 0xFEFFFFFFu), "r"(c_inner), "r"(c_outer)
      : "memory");
}
// 5-D tiled load issued by either CTA of a pair (transaction bytes credited to the leader's mbarrier): the patch
// embedding reads (16 kx, 64 px, 1 ky, 2 py, 1 channel) boxes of the caller's NCHW frames (gemm_tc.cu, AM_PATCH16).
__device__ __forceinline__ void tma_load_5d_2cta(void* smem_dst, const CUtensorMap* tm, uint64_t* bar, int c0, int c1,
                                                 int c2, int c3, int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], "
      "[%1, {%3, %4, %5, %6, %7}], [%2];" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar) & 0xFEFFFFFFu), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
      : "memory");
}
// D[tmem, both CTAs] (+)= A * B with M = 256 split over the pair (bf16 operands); issued by the leader only.
__device__ __forceinline__ void umma_ss_2cta(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Arrive on the mbarrier at this shared-memory offset in every CTA of `cta_mask` once all MMAs issued so far retire.
__device__ __forceinline__ void umma_commit_2cta(uint64_t* bar, uint16_t cta_mask) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
          smem_u32(bar)),
      "h"(cta_mask)
      : "memory");
}

// 32 lanes x 16 consecutive 32-bit columns.
__device__ __forceinline__ void tmem_ld_32x32_16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_32x32_8(uint32_t taddr, uint32_t (&r)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
}
// registers -> TMEM: thread i of the warp writes lane (base_lane + i), 16 consecutive 32-bit columns.
__device__ __forceinline__ void tmem_st_32x32_16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st_32x32_8(uint32_t taddr, const uint32_t (&r)[8]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
      : "memory");
}
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// D[tmem] (+)= A[tmem] * B[smem desc], bf16 operands: A is read from tensor memory (lane = row, two bf16 per
// 32-bit column along K), as written by tcgen05.st.
__device__ __forceinline__ void umma_ts_bf16(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n}\n" ::"r"(tmem_d),
      "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// Generic shared-memory matrix descriptor (Blackwell version 1).  layout: 0 none, 2 SWIZZLE_128B, 4 SWIZZLE_64B,
// 6 SWIZZLE_32B.  K-major operand: sbo = byte stride between 8-row groups (8 x swizzle span), lbo unused.
// MN-major operand (rows of the tile are K indices, the MN extent is contiguous): sbo = byte stride between groups
// of 8 K-rows, lbo = byte stride between swizzle-span-wide atoms along MN.
__device__ __forceinline__ uint64_t umma_desc(uint32_t smem_addr, uint32_t sbo_bytes, uint32_t lbo_bytes,
                                              uint32_t layout) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFF) >> 4);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(layout & 7) << 61;
  return d;
}
// Instruction descriptor for kind::f16 with bf16 operands and fp32 accumulation.
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int M, int N, bool a_mn_major, bool b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (a_mn_major ? (1u << 15) : 0u) | (b_mn_major ? (1u << 16) : 0u) |
         (uint32_t(N >> 3) << 17) | (uint32_t(M >> 4) << 24);
}

// K-major operand tile in shared memory, rows of exactly 128 bytes, TMA SWIZZLE_128B layout:
// 8-row groups are 1024 B apart (SBO), LBO unused, descriptor version 1 (Blackwell), layout type 2.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFF) >> 4);          // start address, bits [0,14)
  d |= static_cast<uint64_t>(0) << 16;                             // leading byte offset (unused)
  d |= static_cast<uint64_t>(1024 >> 4) << 32;                     // stride byte offset, bits [32,46)
  d |= static_cast<uint64_t>(1) << 46;                             // version = 1
  d |= static_cast<uint64_t>(2) << 61;                             // SWIZZLE_128B
  return d;
}

// ---------------------------------------------------------------- legacy tensor path (attention)
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src, bool valid) {
  uint32_t bytes = valid ? 16u : 0u;   // src-size 0 => zero fill
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(smem_dst)), "l"(gmem_src), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2,
                                                  uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}
// D(16x8,f32) += A(16x16,bf16,row) * B(16x8,bf16,col)
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// ---------------------------------------------------------------- math
__device__ __forceinline__ float gelu_erf(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f)); }
// GELU (erf form) with erf from Abramowitz & Stegun 7.1.26 (|erf error| <= 1.5e-7): two MUFU ops (rcp, ex2) and
// ~12 FMA-pipe ops instead of the branchy erff; used where the result is rounded to bf16 anyway.
//   gelu(x) = relu(x) - 0.5 |x| erfc(|x| / sqrt 2),  erfc(z) ~= t (a1 + t (a2 + t (a3 + t (a4 + t a5)))) exp(-z^2)
__device__ __forceinline__ float gelu_erf_fast(float x) {
  const float z = fabsf(x) * 0.70710678118654752440f;
  float t;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(z, 0.3275911f, 1.0f)));
  float p = fmaf(t, 1.061405429f, -1.453152027f);
  p = fmaf(p, t, 1.421413741f);
  p = fmaf(p, t, -0.284496736f);
  p = fmaf(p, t, 0.254829592f);
  float ex;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(ex) : "f"(z * z * -1.4426950408889634f));
  const float erfc_z = p * t * ex;
  return fmaf(-0.5f * fabsf(x), erfc_z, fmaxf(x, 0.f));
}
// GELU (erf form) without special-function-unit work: erf(x / sqrt 2) = xc * Q(t), xc = clamp(x, +-4.75),
// t = xc^2 * (2 / 4.75^2) - 1 in [-1, 1], Q a degree-10 least-squares fit on Chebyshev nodes (|erf error| <= 4e-6 over
// the real line, |gelu error| <= 7e-6: far below the bf16 rounding of the result).  15 FMA-pipe + 2 ALU ops.
__device__ __forceinline__ float gelu_erf_poly(float x) {
  const float xc = fminf(fmaxf(x, -4.75f), 4.75f);
  const float t = fmaf(xc * xc, 8.864265928e-02f, -1.0f);
  float q = 3.041163087e-03f;
  q = fmaf(q, t, -7.997395471e-03f);
  q = fmaf(q, t, 9.152978659e-03f);
  q = fmaf(q, t, -1.351587381e-02f);
  q = fmaf(q, t, 2.733652107e-02f);
  q = fmaf(q, t, -4.377972335e-02f);
  q = fmaf(q, t, 6.069134176e-02f);
  q = fmaf(q, t, -8.103607595e-02f);
  q = fmaf(q, t, 1.064774096e-01f);
  q = fmaf(q, t, -1.473404765e-01f);
  q = fmaf(q, t, 2.974965572e-01f);
  const float h = 0.5f * x;
  return fmaf(h, xc * q, h);
}
// Two GELUs at once on packed fp32x2 registers (FFMA2 / FMUL2): same arithmetic as gelu_erf_poly, half the issue slots.
__device__ __forceinline__ uint64_t f32x2(float a, float b) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ void unpack2(uint64_t v, float& a, float& b) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
}
// 2^x for two values at once WITHOUT the special-function unit: round-to-nearest range reduction with the 1.5 * 2^23
// trick, a degree-3 minimax polynomial for 2^f on [-0.5, 0.5] (relative error 7.5e-5, far below the bf16 rounding of the
// softmax probabilities it feeds) on packed fp32x2 FMAs, and the integer part added into the exponent field.  x <= 0 is
// expected (softmax arguments after the maximum is subtracted); x is clamped at -125.  MUFU.EX2 costs 8 pipe cycles per
// warp instruction on B200 (measured, scripts/ubench/mufu.cu) and is THE contended unit of the attention softmax:
// evaluating a fraction of the exponentials here lets the FMA pipe share that load.
__device__ __forceinline__ uint64_t ex2_poly2(uint64_t x2) {
  float x0, x1;
  unpack2(x2, x0, x1);
  const uint64_t x = f32x2(fmaxf(x0, -125.f), fmaxf(x1, -125.f));
  const uint64_t y = add2(x, f32x2(12582912.f, 12582912.f));
  const uint64_t r = add2(y, f32x2(-12582912.f, -12582912.f));
  const uint64_t f = fma2(r, f32x2(-1.f, -1.f), x);
  uint64_t p = fma2(f, f32x2(0.05517153f, 0.05517153f), f32x2(0.24261111f, 0.24261111f));
  p = fma2(p, f, f32x2(0.69326101f, 0.69326101f));
  p = fma2(p, f, f32x2(0.99992807f, 0.99992807f));
  float p0, p1, y0, y1;
  unpack2(p, p0, p1);
  unpack2(y, y0, y1);
  return f32x2(__int_as_float(__float_as_int(p0) + (__float_as_int(y0) << 23)),
               __int_as_float(__float_as_int(p1) + (__float_as_int(y1) << 23)));
}
__device__ __forceinline__ uint64_t ex2_mufu2(uint64_t x2) {
  float x0, x1, e0, e1;
  unpack2(x2, x0, x1);
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e0) : "f"(x0));
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e1) : "f"(x1));
  return f32x2(e0, e1);
}

// float4 += float4 as two packed adds (bit-identical to four scalar adds)
__device__ __forceinline__ void add4(float4& x, const float4& y) {
  const uint64_t lo = add2(f32x2(x.x, x.y), f32x2(y.x, y.y)), hi = add2(f32x2(x.z, x.w), f32x2(y.z, y.w));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(x.x), "=f"(x.y) : "l"(lo));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(x.z), "=f"(x.w) : "l"(hi));
}
// explicit shared-space 16-byte accesses (a generic pointer into dynamic smem compiles to generic LD/ST)
__device__ __forceinline__ void sts128(uint32_t addr, float a, float b, float c, float d) {
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ void sts128u(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ uint4 lds128u(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void sts32(uint32_t addr, uint32_t v) {
  asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ float lds32(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ float4 lds128(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void gelu_erf_poly2(float& a, float& b) {
  const uint64_t x = f32x2(a, b);
  const uint64_t xc = f32x2(fminf(fmaxf(a, -4.75f), 4.75f), fminf(fmaxf(b, -4.75f), 4.75f));
  const uint64_t t = fma2(mul2(xc, xc), f32x2(8.864265928e-02f, 8.864265928e-02f), f32x2(-1.0f, -1.0f));
  uint64_t q = f32x2(3.041163087e-03f, 3.041163087e-03f);
  q = fma2(q, t, f32x2(-7.997395471e-03f, -7.997395471e-03f));
  q = fma2(q, t, f32x2(9.152978659e-03f, 9.152978659e-03f));
  q = fma2(q, t, f32x2(-1.351587381e-02f, -1.351587381e-02f));
  q = fma2(q, t, f32x2(2.733652107e-02f, 2.733652107e-02f));
  q = fma2(q, t, f32x2(-4.377972335e-02f, -4.377972335e-02f));
  q = fma2(q, t, f32x2(6.069134176e-02f, 6.069134176e-02f));
  q = fma2(q, t, f32x2(-8.103607595e-02f, -8.103607595e-02f));
  q = fma2(q, t, f32x2(1.064774096e-01f, 1.064774096e-01f));
  q = fma2(q, t, f32x2(-1.473404765e-01f, -1.473404765e-01f));
  q = fma2(q, t, f32x2(2.974965572e-01f, 2.974965572e-01f));
  const uint64_t h = mul2(x, f32x2(0.5f, 0.5f));
  const uint64_t r = fma2(h, mul2(xc, q), h);
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(r));
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

}  // namespace bsam
