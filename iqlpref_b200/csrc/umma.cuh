// sm_100a primitives used by the bf16 path: mbarrier, bulk async copies (TMA engine, 1-D),
// tensor memory (TMEM) management, tcgen05.mma issue/commit and TMEM loads.
//
// Operand "images".  Every bf16 GEMM operand, in global memory and in shared memory alike, is
// stored in ONE physical layout, the canonical 128-byte-swizzled UMMA layout:
//   a matrix [R rows][C cols] (C padded to a multiple of 64) is a sequence of column blocks of
//   64 bf16; block cb holds R rows of 128 bytes; inside a row the 16-byte chunk c (8 bf16) sits
//   at chunk position c ^ (row % 8):
//       byte(r, c) = cb*R*128 + r*128 + ((((c % 64) / 8) ^ (r % 8)) * 16) + (c % 8) * 2
// Because global and shared images are byte-identical, tiles move with plain 1-D bulk copies
// (cp.async.bulk, no tensor maps), and the same image serves as
//   * a K-major operand  (rows = M/N index, cols = K):  SBO = 1024 B, K advances 32 B per
//     UMMA_K=16 inside a block and R*128 B between blocks;
//   * an MN-major operand (cols = M/N index, rows = K):  LBO = R*128 B (block stride),
//     SBO = 1024 B, K advances 2048 B (16 rows) per UMMA_K=16.
// (cute/atom/mma_traits_sm100.hpp documents the canonical layouts these descriptors encode.)
#pragma once
#include <cuda_bf16.h>
#include <stdint.h>

namespace iqlb {
namespace umma {

__host__ __device__ __forceinline__ int64_t image_bytes(int rows, int cols) {
  return (int64_t)((cols + 63) / 64) * rows * 128;
}
// byte offset of element (r, c) in the image of an [rows x *] matrix
__host__ __device__ __forceinline__ int64_t image_off(int rows, int r, int c) {
  return (int64_t)(c >> 6) * rows * 128 + (int64_t)r * 128 + ((((c & 63) >> 3) ^ (r & 7)) << 4) + ((c & 7) << 1);
}

#ifdef __CUDACC__
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- single-thread issue --------------------------------------------------------------------------
// tcgen05.mma / tcgen05.commit / cp.async.bulk take their operands from UNIFORM registers.  Guarded by a per-thread
// condition (`threadIdx.x == 0`, `lane == 0`) the compiler cannot prove the operands warp-uniform and wraps EVERY such
// instruction in a loop over the active threads (SASS: VOTEU / ELECT / R2UR.BROADCAST / UTCHMMA / BRA.U.ANY), ~120 clocks
// per MMA whatever its shape (profiles/r01_mma_issue_rate.md).  Guarded by `elect.sync` inside a warp-uniform branch the
// operands stay in uniform registers and consecutive MMAs issue back to back (profiles/r02_mma_issue_rate.md).
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.b32 %0, 1, 0, P;\n\t}" : "=r"(pred));
  return pred != 0;
}
// warp index as a value the compiler knows to be the same in every lane
__device__ __forceinline__ int warp_uniform() { return __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0); }

// ---- mbarrier -----------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}

// ---- proxies / fences -----------------------------------------------------------------------------
// generic-proxy writes (st.shared) -> visible to the async proxy (tcgen05.mma / bulk copies)
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// ---- 1-D bulk copies (TMA engine, SASS UBLKCP) -----------------------------------------------------
__device__ __forceinline__ void bulk_g2s(uint32_t dst_smem, const void* src_gmem, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_smem),
               "l"(src_gmem), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* dst_gmem, uint32_t src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem), "r"(src_smem), "r"(bytes)
               : "memory");
}
// HBM -> L2 only (no shared-memory destination): hides DRAM latency of a later bulk copy or ld.global
__device__ __forceinline__ void bulk_prefetch_l2(const void* p, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// ---- TMEM ------------------------------------------------------------------------------------------
// one full warp; writes the base address (lane 0, column c) to *slot in shared memory
__device__ __forceinline__ void tmem_alloc(uint32_t slot_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot_smem), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// 32 consecutive fp32 columns of this thread's TMEM lane (warp w of the CTA owns lanes 32*(w%4)..+31)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// Same wait, declared to the compiler as the producer of v: a tcgen05.ld issued earlier for v may still be in flight
// while other registers are being worked on (software-pipelined epilogues), and nothing may read v before this point.
__device__ __forceinline__ void tmem_ld_wait_on(float (&v)[32]) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.wait::ld.sync.aligned;"
      : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]),
        "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]), "+r"(r[16]),
        "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]), "+r"(r[24]),
        "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31])
      :
      : "memory");
}

// ---- descriptors ---------------------------------------------------------------------------------
// shared-memory matrix descriptor, SWIZZLE_128B, Blackwell version field = 1
//   bits [0,14) start>>4 | [16,30) LBO>>4 | [32,46) SBO>>4 | [46,48) version=1 | [61,64) layout=2
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3FFFF) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// instruction descriptor, kind::f16, bf16 x bf16 -> fp32, M = 128
//   [4,6) c=F32(1) | [7,10) a=BF16(1) | [10,13) b=BF16(1) | 15 a_major | 16 b_major | [17,23) N>>3 | [24,29) M>>4
__device__ __forceinline__ uint32_t instr_desc_bf16(int n, int a_mn_major, int b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn_major << 15) | ((uint32_t)b_mn_major << 16) |
         ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}
// D[tmem] (+)= A[smem] * B[smem]; one thread issues
__device__ __forceinline__ void mma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Same, with the A operand read from tensor memory: lane = row (M = 128), each 32-bit column holds two
// consecutive bf16 of K (even k in the low half), so one UMMA_K = 16 step spans 8 columns.
__device__ __forceinline__ void mma_bf16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// 16 consecutive 32-bit columns of this thread's TMEM lane <- registers
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
        "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// arrive on an mbarrier once all previously issued tcgen05.mma of this thread have completed
__device__ __forceinline__ void mma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

// Operand view of an image resident in shared memory.
struct Operand {
  uint32_t base;        // shared address of the image (1024-byte aligned) + row/block offset of the tile
  uint32_t lbo, sbo;    // descriptor strides in bytes
  uint32_t kstep;       // bytes per UMMA_K=16 step inside a group of 4 steps
  uint32_t kblock;      // bytes per group of 4 steps
  int mn_major;
  __device__ __forceinline__ uint64_t desc(int ks) const {
    return smem_desc(base + (uint32_t)(ks >> 2) * kblock + (uint32_t)(ks & 3) * kstep, lbo, sbo);
  }
};
// rows of the image are the M/N index, columns are K.  `rows_total` = rows of the whole image.
__device__ __forceinline__ Operand operand_k_major(uint32_t img, int rows_total, int row0) {
  Operand o;
  o.base = img + (uint32_t)row0 * 128;
  o.lbo = 16; o.sbo = 1024; o.kstep = 32; o.kblock = (uint32_t)rows_total * 128; o.mn_major = 0;
  return o;
}
// columns of the image are the M/N index (starting at col0, multiple of 64), rows are K (from row0)
__device__ __forceinline__ Operand operand_mn_major(uint32_t img, int rows_total, int col0, int row0) {
  Operand o;
  o.base = img + (uint32_t)(col0 >> 6) * rows_total * 128 + (uint32_t)row0 * 128;
  o.lbo = (uint32_t)rows_total * 128; o.sbo = 1024; o.kstep = 2048; o.kblock = 8192; o.mn_major = 1;
  return o;
}
// issue ksteps MMAs of shape 128 x n x 16 (one thread)
__device__ __forceinline__ void gemm_issue(uint32_t d_tmem, const Operand& a, const Operand& b, int n, int ksteps,
                                           bool accumulate_first) {
  const uint32_t idesc = instr_desc_bf16(n, a.mn_major, b.mn_major);
  for (int ks = 0; ks < ksteps; ++ks) mma_bf16(d_tmem, a.desc(ks), b.desc(ks), idesc, (ks > 0 || accumulate_first) ? 1u : 0u);
}

// pack two floats to a bf16x2 word (lo = a)
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
#endif  // __CUDACC__

}  // namespace umma
}  // namespace iqlb
