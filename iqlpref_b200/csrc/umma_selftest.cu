// Self-test of the tcgen05 building blocks (descriptors, operand majorness, TMEM read-back):
// one CTA computes D[128 x n] = A * B from two operand images with every operand arrangement the
// IQL step uses.  Exercised by tests/test_gpu_umma.py against numpy.
#include "common.cuh"
#include "umma.cuh"

namespace iqlb {
using namespace umma;

__global__ void __launch_bounds__(128, 1) umma_selftest_kernel(const uint8_t* __restrict__ a_img, uint32_t a_bytes,
                                                               const uint8_t* __restrict__ b_img, uint32_t b_bytes,
                                                               int a_mn, int a_rows_total, int a_off, int b_mn,
                                                               int b_rows_total, int b_off, int n, int ksteps,
                                                               float* __restrict__ d_out) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[2];
  __shared__ uint32_t tmem_slot;
  // dynamic shared memory is only guaranteed 16-byte aligned: round up to the 1024 B swizzle atom
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t a_s = base, b_s = base + ((a_bytes + 1023u) & ~1023u);
  const uint32_t bar_ld = smem_u32(&bars[0]), bar_mma = smem_u32(&bars[1]);
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    mbar_init(bar_ld, 1);
    mbar_init(bar_mma, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(smem_u32(&tmem_slot), 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (threadIdx.x == 0) {
    mbar_expect_tx(bar_ld, a_bytes + b_bytes);
    bulk_g2s(a_s, a_img, a_bytes, bar_ld);
    bulk_g2s(b_s, b_img, b_bytes, bar_ld);
    mbar_wait(bar_ld, 0);
    tc_fence_after();
    Operand a = a_mn ? operand_mn_major(a_s, a_rows_total, a_off, 0) : operand_k_major(a_s, a_rows_total, a_off);
    Operand b = b_mn ? operand_mn_major(b_s, b_rows_total, b_off, 0) : operand_k_major(b_s, b_rows_total, b_off);
    gemm_issue(tmem, a, b, n, ksteps, false);
    mma_commit(bar_mma);
  }
  mbar_wait(bar_mma, 0);
  tc_fence_after();
  const int row = threadIdx.x;
  const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
  for (int c0 = 0; c0 < n; c0 += 16) {
    float v[16];
    tmem_ld16(lane_base + (uint32_t)c0, v);
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 16; ++j) d_out[(int64_t)row * n + c0 + j] = v[j];
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

// Issue-rate probe: `count` back-to-back MMAs (128 x n x 16, operands uninitialised) into one accumulator,
// A from shared memory (mode 0) or tensor memory (mode 1); clocks from the first issue to the commit arrival.
__global__ void __launch_bounds__(128, 1) umma_rate_kernel(int mode, int n, int count, unsigned long long* out) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[1];
  __shared__ uint32_t tmem_slot;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t a_s = base, b_s = base + 16384u;
  const uint32_t bar = smem_u32(&bars[0]);
  const int warp = threadIdx.x >> 5;
  for (uint32_t i = threadIdx.x; i < (16384u + 32768u) / 4; i += blockDim.x)
    reinterpret_cast<uint32_t*>(smem_raw + (base - smem_u32(smem_raw)))[i] = 0;
  fence_async_smem();
  if (threadIdx.x == 0) { mbar_init(bar, (mode & 2) ? 2 : 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(smem_u32(&tmem_slot), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  // mode bit 0: A from tensor memory; bit 1: two issuing warps (accumulators 128 columns apart, n <= 128)
  const int issuers = (mode & 2) ? 2 : 1;
  if (mode & 4) {  // issue guarded by elect.sync inside a warp-uniform branch (umma.cuh: elect_one)
    if (warp_uniform() < issuers) {
      if (elect_one()) {
        const int wu = (int)(threadIdx.x >> 5);
        const uint32_t idesc = instr_desc_bf16(n, 0, 0);
        const Operand a = operand_k_major(a_s, 128, 0), b = operand_k_major(b_s, n, 0);
        const uint32_t acc = tmem + 128u * (uint32_t)wu, at = tmem + 256u + 64u * (uint32_t)wu;
        const long long t0 = clock64();
        for (int i = 0; i < count; ++i) {
          if ((mode & 1) == 0) mma_bf16(acc, a.desc(i & 3), b.desc(i & 3), idesc, i > 0);
          else mma_bf16_ts(acc, at + 8u * (uint32_t)(i & 7), b.desc(i & 3), idesc, i > 0);
        }
        const long long t1 = clock64();
        mma_commit(bar);
        mbar_wait(bar, 0);
        const long long t2 = clock64();
        if (wu == 0) {
          out[0] = (unsigned long long)(t1 - t0);
          out[1] = (unsigned long long)(t2 - t0);
        }
      }
    }
  } else if ((threadIdx.x & 31) == 0 && warp < issuers) {
    const uint32_t idesc = instr_desc_bf16(n, 0, 0);
    const Operand a = operand_k_major(a_s, 128, 0), b = operand_k_major(b_s, n, 0);
    const uint32_t acc = tmem + 128u * (uint32_t)warp, at = tmem + 256u + 64u * (uint32_t)warp;
    const long long t0 = clock64();
    for (int i = 0; i < count; ++i) {
      if ((mode & 1) == 0) mma_bf16(acc, a.desc(i & 3), b.desc(i & 3), idesc, i > 0);
      else mma_bf16_ts(acc, at + 8u * (uint32_t)(i & 7), b.desc(i & 3), idesc, i > 0);
    }
    const long long t1 = clock64();
    mma_commit(bar);
    mbar_wait(bar, 0);
    const long long t2 = clock64();
    if (warp == 0) {
      out[0] = (unsigned long long)(t1 - t0);
      out[1] = (unsigned long long)(t2 - t0);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

}  // namespace iqlb

using namespace iqlb;

extern "C" int iqlb_selftest_mma_rate(int32_t a_from_tmem, int32_t n, int32_t count, unsigned long long* clocks_out,
                                      void* stream) {
  IQLB_REQUIRE(clocks_out != nullptr, "null pointer");
  IQLB_REQUIRE(n >= 16 && n <= 256 && n % 16 == 0 && count >= 1, "n multiple of 16 in [16,256]");
  IQLB_REQUIRE(a_from_tmem >= 0 && a_from_tmem <= 7 && (!(a_from_tmem & 2) || n <= 128), "mode");
  const size_t smem = 16384 + 32768 + 1024;
  IQLB_CUDA(cudaFuncSetAttribute(umma_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  umma_rate_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(a_from_tmem, n, count, clocks_out);
  IQLB_LAUNCH_CHECK();
  return IQLB_OK;
}

extern "C" int iqlb_selftest_umma(const void* a_img, int64_t a_bytes, const void* b_img, int64_t b_bytes, int32_t a_mn,
                                  int32_t a_rows_total, int32_t a_off, int32_t b_mn, int32_t b_rows_total,
                                  int32_t b_off, int32_t n, int32_t ksteps, float* d_out, void* stream) {
  IQLB_REQUIRE(a_img && b_img && d_out, "null pointer");
  IQLB_REQUIRE(a_bytes > 0 && b_bytes > 0 && a_bytes % 16 == 0 && b_bytes % 16 == 0, "image sizes");
  IQLB_REQUIRE(n >= 16 && n <= 256 && n % 16 == 0 && ksteps >= 1, "n multiple of 16 in [16,256]");
  const size_t smem = (size_t)((a_bytes + 1023) / 1024 * 1024 + (b_bytes + 1023) / 1024 * 1024 + 1024);
  IQLB_REQUIRE(smem <= 227 * 1024, "operands exceed shared memory");
  IQLB_CUDA(cudaFuncSetAttribute(umma_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  umma_selftest_kernel<<<1, 128, smem, (cudaStream_t)stream>>>((const uint8_t*)a_img, (uint32_t)a_bytes,
                                                              (const uint8_t*)b_img, (uint32_t)b_bytes, a_mn,
                                                              a_rows_total, a_off, b_mn, b_rows_total, b_off, n, ksteps,
                                                              d_out);
  IQLB_LAUNCH_CHECK();
  return IQLB_OK;
}
