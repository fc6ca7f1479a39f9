// Shared helpers for libiqlb: error plumbing, Philox4x32-10, warp/block reductions.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/iqlb.h"

namespace iqlb {

extern thread_local char g_err[512];

inline int fail(int code, const char* fmt, const char* a = "", long long b = 0) {
  snprintf(g_err, sizeof(g_err), fmt, a, b);
  return code;
}

#define IQLB_CUDA(expr)                                                                     \
  do {                                                                                      \
    cudaError_t e__ = (expr);                                                               \
    if (e__ != cudaSuccess) {                                                               \
      snprintf(iqlb::g_err, sizeof(iqlb::g_err), "%.300s failed: %s (%s:%d)", #expr,            \
               cudaGetErrorString(e__), __FILE__, __LINE__);                                \
      return IQLB_ERR_CUDA;                                                                 \
    }                                                                                       \
  } while (0)

#define IQLB_LAUNCH_CHECK() IQLB_CUDA(cudaGetLastError())

#define IQLB_REQUIRE(cond, msg)                                                             \
  do {                                                                                      \
    if (!(cond)) {                                                                          \
      snprintf(iqlb::g_err, sizeof(iqlb::g_err), "invalid argument: %s (%s)", msg, #cond);  \
      return IQLB_ERR_INVALID;                                                              \
    }                                                                                       \
  } while (0)

// RewardMLP activations (optbnn/bnn/nets/mlp.py:45-48), ids of include/iqlb.h
#ifdef __CUDACC__
__device__ __forceinline__ float apply_act(int act, float x) {
  switch (act) {
    case IQLB_ACT_RELU: return x > 0.f ? x : 0.f;
    case IQLB_ACT_TANH: return tanhf(x);
    case IQLB_ACT_COS: return cosf(x);
    case IQLB_ACT_SIN: return sinf(x);
    case IQLB_ACT_LINEAR: return x;
    case IQLB_ACT_RBF: return expf(-x * x);
    case IQLB_ACT_SWISH: return x * (1.0f / (1.0f + expf(-x)));
    case IQLB_ACT_SOFTPLUS: return x > 20.f ? x : log1pf(expf(x));
    case IQLB_ACT_LEAKY_RELU: return x > 0.f ? x : 0.01f * x;
    default: return x;
  }
}
// d act / dz at pre-activation z (backward pass of the reward-MLP trainer)
__device__ __forceinline__ float act_grad(int act, float z) {
  switch (act) {
    case IQLB_ACT_RELU: return z > 0.f ? 1.f : 0.f;
    case IQLB_ACT_TANH: { const float t = tanhf(z); return 1.f - t * t; }
    case IQLB_ACT_COS: return -sinf(z);
    case IQLB_ACT_SIN: return cosf(z);
    case IQLB_ACT_LINEAR: return 1.f;
    case IQLB_ACT_RBF: return -2.f * z * expf(-z * z);
    case IQLB_ACT_SWISH: { const float s = 1.0f / (1.0f + expf(-z)); return s * (1.f + z * (1.f - s)); }
    case IQLB_ACT_SOFTPLUS: return z > 20.f ? 1.f : 1.0f / (1.0f + expf(-z));
    case IQLB_ACT_LEAKY_RELU: return z > 0.f ? 1.f : 0.01f;
    default: return 1.f;
  }
}
#endif

#ifdef __CUDACC__
// GEMM epilogues evaluate an activation per output element, 32-64 elements per thread, fully unrolled: with the switch
// inlined at every element (cosf / sinf / tanhf bring their slow paths along) a kernel grows to 20-80 thousand
// instructions and every element's path through it misses the instruction cache - the tensor-core GEMM spent 60 of its 75
// thousand clocks per tile there.  The two activations the hot paths use stay inline; the rest is ONE out-of-line copy.
static __device__ __noinline__ float apply_act_call(int act, float x) { return apply_act(act, x); }
static __device__ __noinline__ float act_grad_call(int act, float z) { return act_grad(act, z); }
__device__ __forceinline__ float apply_act_lean(int act, float x) {
  if (act == IQLB_ACT_RELU) return x > 0.f ? x : 0.f;
  if (act == IQLB_ACT_LINEAR) return x;
  return apply_act_call(act, x);
}
__device__ __forceinline__ float act_grad_lean(int act, float z) {
  if (act == IQLB_ACT_RELU) return z > 0.f ? 1.f : 0.f;
  if (act == IQLB_ACT_LINEAR) return 1.f;
  return act_grad_call(act, z);
}
#endif

// ---- Philox4x32-10 exactly as cuRAND's curand_Philox4x32_10 (curand_philox4x32_x.h) ----------
__host__ __device__ __forceinline__ void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
#ifdef __CUDA_ARCH__
  uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
  uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
#else
  uint64_t p0 = (uint64_t)0xD2511F53u * c[0], p1 = (uint64_t)0xCD9E8D57u * c[2];
  uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0, hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
#endif
  uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
  c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
}

// First curand4() of curand_init(seed, subsequence, offset) with offset % 4 == 0.
__host__ __device__ __forceinline__ void philox4(uint64_t seed, uint64_t subsequence, uint64_t offset,
                                                 uint32_t (&out)[4]) {
  uint64_t blk = offset >> 2;
  uint32_t c[4] = {(uint32_t)blk, (uint32_t)(blk >> 32), (uint32_t)subsequence, (uint32_t)(subsequence >> 32)};
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    philox_round(c, k0, k1);
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  out[0] = c[0]; out[1] = c[1]; out[2] = c[2]; out[3] = c[3];
}

// torch.randint on CUDA (DistributionTemplates.h:319-345): 32-bit path below 2^28.
__host__ __device__ __forceinline__ int64_t randint_from_philox(const uint32_t (&r)[4], uint64_t range) {
  if (range >= (1ull << 28)) {
    uint64_t v = ((uint64_t)r[0] << 32) | r[1];
    return (int64_t)(v % range);
  }
  return (int64_t)(r[0] % (uint32_t)range);
}

// ---- L2 eviction hints: optimizer state is touched once per step and is larger than L2, so it streams with
// evict_first and leaves the operand images / activations of the tensor-core kernels resident ----------------
#ifdef __CUDACC__
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ float4 ld_f4_hint(const float* ptr, uint64_t pol) {
  float4 v;
  asm volatile("ld.global.L2::cache_hint.v4.f32 {%0, %1, %2, %3}, [%4], %5;"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(ptr), "l"(pol));
  return v;
}
__device__ __forceinline__ void st_f4_hint(float* ptr, const float4& v, uint64_t pol) {
  asm volatile("st.global.L2::cache_hint.v4.f32 [%0], {%1, %2, %3, %4}, %5;" ::"l"(ptr), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w), "l"(pol)
               : "memory");
}
__device__ __forceinline__ void st_u2_hint(void* ptr, uint32_t a, uint32_t b, uint64_t pol) {
  asm volatile("st.global.L2::cache_hint.v2.b32 [%0], {%1, %2}, %3;" ::"l"(ptr), "r"(a), "r"(b), "l"(pol) : "memory");
}
#endif

// ---- programmatic dependent launch (PDL) ----------------------------------------------------------------
// Every kernel of the bf16 step is launched with programmatic stream serialization: its CTAs may start while
// the previous kernel's last wave is still running, do the part of their prologue that does not touch the
// previous kernel's outputs (barrier / TMEM set-up, weight loads, bookkeeping), then pdl_wait() until the
// previous grid has completed and flushed.  pdl_launch() comes AFTER the wait in every kernel, so at most
// two adjacent kernels overlap and anything older than the immediate predecessor is complete.
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, bool pdl,
                              Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
// IQLB_PDL: bit mask of kernels launched with the attribute (1 gather, 2 forward, 4 loss, 8 data gradient,
// 16 weight gradient, 32 Adam)
inline bool pdl_enabled(int bit) {
  static const int mask = getenv("IQLB_PDL") ? atoi(getenv("IQLB_PDL")) : 27;  // measured on B200: loss and Adam lose with it
  return (mask & bit) != 0;
}
#endif

// ---- reductions -----------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Deterministic block sum (fixed order); blockDim.x multiple of 32, <= 1024.  All threads get it.
__device__ __forceinline__ float block_sum(float v, float* smem32) {
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) smem32[w] = v;
  __syncthreads();
  float t = (lane < nw) ? smem32[lane] : 0.f;
  t = warp_sum(t);
  return t;
}

}  // namespace iqlb
