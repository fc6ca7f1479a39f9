// Generic fp32-in / fp32-out GEMM on the tensor cores (tcgen05, TMEM accumulator), "split-bf16" arithmetic:
// every product a * b is evaluated as a_hi*b_hi + a_lo*b_hi + a_hi*b_lo with hi = bf16(v), lo = bf16(v - hi) and fp32
// accumulation, i.e. ~2^-17 relative per product instead of bf16's 2^-9 - the same formulation as relabel_mlp_tc.cu,
// here for operands that live in global memory as plain fp32 matrices.  Serves the Linear layers of the preference
// transformer (pref_trans.py:82-87, 119-125, 174-183; relabel_pt.cu) and is selected by launch_gemm (gemm_f32.cu) for
// tasks that set GemmTask::allow_tc.
//
//   LINEAR   C[m,n] = act(sum_k A[m,k] * W[n,k] + bias[n]) (+ residual)      W stored [N][K] (nn.Linear)
//            with t.pre:  C[m,n] = scale * acc * act'(pre[m,n])               (data gradient through an activation)
//   FWD_KN   C[m,n] = act(scale * sum_k A[m,k] * W[k,n] + bias[n])            W stored [K][N]; pre-activation kept in t.pre
//   BWD_WEIGHT C[m,n] = sum_k A[k,m] * X[k,n]                                 dW = (layer input)^T dZ over the rows k
//            (both operands read MN-major; the caller splits the rows over tasks and sums the partial tiles)
//
// One CTA = one 128 x 128 output tile.  K is walked in blocks of 64: all 256 threads load the fp32 tiles (128-bit loads,
// the next block's loads are issued before the current block is converted), split them into hi / lo bf16 and write the
// four operand images of a stage (A_hi, A_lo, B_hi, B_lo; 16 KB each, 128-byte-swizzled UMMA layout, umma.cuh) into a
// 3-stage shared-memory ring; one thread issues the 12 MMAs of a stage (4 k-steps x 3 products, M = 128, N = 128) and
// commits them to the stage's "empty" barrier.  Epilogue: 8 warps drain the 128 x 128 fp32 accumulator from TMEM.
#include "iql_step.cuh"
#include "umma.cuh"

namespace iqlb {
using namespace umma;

namespace gtc {
constexpr int BM = 128, BN = 128, BK = 64, kStages = 3;
constexpr uint32_t kImg = 128 * 128;             // bytes of one operand image of a stage (128 rows x 64 bf16)
constexpr uint32_t kStageBytes = 4 * kImg;       // A_hi | A_lo | B_hi | B_lo
constexpr uint32_t oCtrl = kStages * kStageBytes;
constexpr uint32_t kSmem = oCtrl + 128;
}  // namespace gtc

__device__ __forceinline__ void split8(const float (&v)[8], uint4& hi, uint4& lo) {
  uint32_t h[4], l[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const __nv_bfloat162 hh = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
    const float2 hf = __bfloat1622float2(hh);
    const __nv_bfloat162 ll = __floats2bfloat162_rn(v[2 * j] - hf.x, v[2 * j + 1] - hf.y);
    h[j] = *reinterpret_cast<const uint32_t*>(&hh);
    l[j] = *reinterpret_cast<const uint32_t*>(&ll);
  }
  hi = make_uint4(h[0], h[1], h[2], h[3]);
  lo = make_uint4(l[0], l[1], l[2], l[3]);
}

// 8 consecutive floats p[0..8) of a row, zero beyond `valid` (<= 8); fast: 16-byte aligned
__device__ __forceinline__ void load8(const float* __restrict__ p, int valid, bool fast, float (&v)[8]) {
  if (valid >= 8 && fast) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  } else {
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = j < valid ? __ldg(p + j) : 0.f;
  }
}

template <int MODE>
__device__ __forceinline__ void gemm_tc_body(const GemmTask& t, const int tile) {
  using namespace gtc;
  constexpr bool B_KN = (MODE == GEMM_FWD_KN || MODE == GEMM_BWD_WEIGHT);
  constexpr bool A_MN = (MODE == GEMM_BWD_WEIGHT);
  extern __shared__ __align__(1024) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int warp_u = warp_uniform();
  const uint32_t sbase = smem_u32(smem);
  const uint32_t bar_empty = sbase + oCtrl, bar_done = bar_empty + 8 * kStages;
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + oCtrl + 8 * kStages + 8);
  const int tiles_n = (t.N + BN - 1) / BN;
  if (tile >= ((t.M + BM - 1) / BM) * tiles_n) return;  // grouped launches size the grid for the largest task
  const int tm = tile / tiles_n, tn = tile % tiles_n;
  const int m0 = tm * BM, n0 = tn * BN;
  const int nkb = (t.K + BK - 1) / BK;
  const bool dbg = t.dbg != nullptr && tile == (int)gridDim.x / 2 && tid == 64;
#define GTC_TS(i) do { if (dbg) t.dbg[(i)] = (unsigned long long)clock64(); } while (0)
  GTC_TS(0);

  if (tid == 0) {
    if (sbase & 1023u) __trap();
    for (int i = 0; i < kStages; ++i) mbar_init(bar_empty + 8 * i, 1);
    mbar_init(bar_done, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(sbase + oCtrl + 8 * kStages + 8, 128);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  const bool fastA = ((reinterpret_cast<uintptr_t>(t.A) & 15) == 0) && (t.lda % 4 == 0);
  const bool fastB = ((reinterpret_cast<uintptr_t>(t.Bm) & 15) == 0) && (t.ldb % 4 == 0);
  // chunk = 8 consecutive elements along the contiguous dimension; 1024 chunks per operand and stage, 4 per thread
  float ra[4][8], rb[4][8];
  auto load_stage = [&](int kb) {
    const int k0 = kb * BK;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int q = tid + 256 * i;
      if (!A_MN) {  // A[m][k], k contiguous: row r, chunk c
        const int r = q >> 3, c = q & 7;
        const int gm = m0 + r, gk = k0 + 8 * c;
        const int valid = gm < t.M ? min(8, t.K - gk) : 0;
        load8(t.A + (int64_t)gm * t.lda + gk, valid, fastA, ra[i]);
      } else {      // A[k][m], m contiguous
        const int kk = q >> 4, cm = q & 15;
        const int gk = k0 + kk, gm = m0 + 8 * cm;
        const int valid = gk < t.K ? min(8, t.M - gm) : 0;
        load8(t.A + (int64_t)gk * t.lda + gm, valid, fastA, ra[i]);
      }
      if (!B_KN) {  // W[n][k], k contiguous
        const int r = q >> 3, c = q & 7;
        const int gn = n0 + r, gk = k0 + 8 * c;
        const int valid = gn < t.N ? min(8, t.K - gk) : 0;
        load8(t.Bm + (int64_t)gn * t.ldb + gk, valid, fastB, rb[i]);
      } else {      // W[k][n], n contiguous: k row kk, chunk cn of 16
        const int kk = q >> 4, cn = q & 15;
        const int gk = k0 + kk, gn = n0 + 8 * cn;
        const int valid = gk < t.K ? min(8, t.N - gn) : 0;
        load8(t.Bm + (int64_t)gk * t.ldb + gn, valid, fastB, rb[i]);
      }
    }
  };
  auto store_stage = [&](int slot) {
    uint8_t* st = smem + (uint32_t)slot * kStageBytes;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int q = tid + 256 * i;
      uint4 hi, lo;
      {
        uint32_t off;
        if (!A_MN) {
          const int r = q >> 3, c = q & 7;
          off = (uint32_t)r * 128u + (uint32_t)((c ^ (r & 7)) << 4);
        } else {
          const int kk = q >> 4, cm = q & 15;
          off = (uint32_t)(cm >> 3) * (64u * 128u) + (uint32_t)kk * 128u + (uint32_t)(((cm & 7) ^ (kk & 7)) << 4);
        }
        split8(ra[i], hi, lo);
        *reinterpret_cast<uint4*>(st + off) = hi;
        *reinterpret_cast<uint4*>(st + kImg + off) = lo;
      }
      split8(rb[i], hi, lo);
      uint32_t off;
      if (!B_KN) {
        const int r = q >> 3, c = q & 7;
        off = (uint32_t)r * 128u + (uint32_t)((c ^ (r & 7)) << 4);
      } else {  // MN-major image: 64-column (n) blocks of 64 rows (k) x 128 bytes
        const int kk = q >> 4, cn = q & 15;
        off = (uint32_t)(cn >> 3) * (64u * 128u) + (uint32_t)kk * 128u + (uint32_t)(((cn & 7) ^ (kk & 7)) << 4);
      }
      *reinterpret_cast<uint4*>(st + 2 * kImg + off) = hi;
      *reinterpret_cast<uint4*>(st + 3 * kImg + off) = lo;
    }
  };

  GTC_TS(1);
  load_stage(0);
  for (int kb = 0; kb < nkb; ++kb) {
    const int slot = kb % kStages, use = kb / kStages;
    if (use > 0) mbar_wait(bar_empty + 8 * slot, (uint32_t)((use - 1) & 1));  // the MMAs that read this slot are done
    if (kb < 4) GTC_TS(2 + 3 * kb);
    store_stage(slot);
    if (kb < 4) GTC_TS(3 + 3 * kb);
    if (kb + 1 < nkb) load_stage(kb + 1);  // in flight while this stage is multiplied
    fence_async_smem();
    __syncthreads();
    if (kb < 4) GTC_TS(4 + 3 * kb);
    if (warp_u == 0 && elect_one()) {  // one issuing thread, chosen so that the operands stay in uniform registers (umma.cuh)
      tc_fence_after();
      const uint32_t st = sbase + (uint32_t)slot * kStageBytes;
      const Operand a_hi = A_MN ? operand_mn_major(st, 64, 0, 0) : operand_k_major(st, 128, 0);
      const Operand a_lo = A_MN ? operand_mn_major(st + kImg, 64, 0, 0) : operand_k_major(st + kImg, 128, 0);
      const Operand b_hi = B_KN ? operand_mn_major(st + 2 * kImg, 64, 0, 0) : operand_k_major(st + 2 * kImg, 128, 0);
      const Operand b_lo = B_KN ? operand_mn_major(st + 3 * kImg, 64, 0, 0) : operand_k_major(st + 3 * kImg, 128, 0);
      const int ks = min(BK, t.K - kb * BK);
      const int ksteps = (ks + 15) / 16;  // the tail of K is zero-filled up to a multiple of 16
      gemm_issue(tmem, a_lo, b_hi, BN, ksteps, kb > 0);  // small terms first
      gemm_issue(tmem, a_hi, b_lo, BN, ksteps, true);
      gemm_issue(tmem, a_hi, b_hi, BN, ksteps, true);
      mma_commit(bar_empty + 8 * slot);
      if (kb + 1 == nkb) mma_commit(bar_done);
    }
  }
  GTC_TS(14);
  mbar_wait(bar_done, 0);
  tc_fence_after();
  GTC_TS(15);

  // ---- epilogue, two passes through shared memory (every stage of the ring is free once the last MMA has completed):
  //  1. warp w drains TMEM lane quarter w % 4 (32 rows), column half w / 4 into a padded fp32 tile (row stride 132
  //     floats: the 128-bit stores of 8 lanes - 8 different rows - cover the 32 banks exactly once);
  //  2. warp w walks rows w, w + 8, ...: lane l owns columns 4l .. 4l+3 of every row, so bias is loaded once per thread
  //     and the residual / pre-activation reads and the output stores are 512 contiguous bytes per warp instruction.
  //     (One thread per output ROW wrote 32 scattered 16-byte pieces per instruction: 11 us of a 14 us tile.)
  constexpr int kLd = 132;
  float* ctile = reinterpret_cast<float*>(smem);
  {
    const int quarter = warp & 3, chalf = warp >> 2;
    const int r = quarter * 32 + lane;
    const uint32_t lane_base = tmem + ((uint32_t)(quarter * 32) << 16);
#pragma unroll
    for (int cc = 0; cc < 2; ++cc) {
      const int c0 = chalf * 64 + cc * 32;
      float v[32];
      tmem_ld32(lane_base + (uint32_t)c0, v);
      tmem_ld_wait();
#pragma unroll
      for (int j4 = 0; j4 < 8; ++j4)
        *reinterpret_cast<float4*>(ctile + r * kLd + c0 + 4 * j4) = make_float4(v[4 * j4], v[4 * j4 + 1], v[4 * j4 + 2], v[4 * j4 + 3]);
    }
  }
  tc_fence_before();
  __syncthreads();
  {
    const int nb = n0 + 4 * lane;
    const bool cvec = (t.ldc % 4 == 0) && ((reinterpret_cast<uintptr_t>(t.C) & 15) == 0) && nb + 4 <= t.N;
    float bias4[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) bias4[j] = (MODE != GEMM_BWD_WEIGHT && t.bias != nullptr && nb + j < t.N) ? __ldg(t.bias + nb + j) : 0.f;
    const int act = t.act;
    for (int r = warp; r < BM; r += 8) {
      const int m = m0 + r;
      if (m >= t.M) break;
      const float4 a4 = *reinterpret_cast<const float4*>(ctile + r * kLd + 4 * lane);
      float x[4] = {a4.x, a4.y, a4.z, a4.w};
      const int64_t o = (int64_t)m * t.ldc + nb;
      float side[4] = {0.f, 0.f, 0.f, 0.f};  // residual (LINEAR) or pre-activation (LINEAR with t.pre)
      const float* sp = (MODE == GEMM_LINEAR) ? (t.pre != nullptr ? t.pre : t.mask) : nullptr;
      if (sp != nullptr) {
        if (cvec) { const float4 s4 = *reinterpret_cast<const float4*>(sp + o); side[0] = s4.x; side[1] = s4.y; side[2] = s4.z; side[3] = s4.w; }
        else {
#pragma unroll
          for (int j = 0; j < 4; ++j) if (nb + j < t.N) side[j] = sp[o + j];
        }
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (MODE == GEMM_BWD_WEIGHT) {
          // plain partial sum
        } else if (MODE == GEMM_FWD_KN) {
          x[j] = fmaf(t.scale, x[j], bias4[j]);
        } else if (t.pre != nullptr) {
          x[j] = t.scale * x[j] * act_grad_lean(act, side[j]);
        } else {
          x[j] = apply_act_lean(act, x[j] + bias4[j]) + side[j];
        }
      }
      if (MODE == GEMM_FWD_KN) {
        if (t.pre != nullptr) {
          if (cvec) *reinterpret_cast<float4*>(t.pre + o) = make_float4(x[0], x[1], x[2], x[3]);
          else {
#pragma unroll
            for (int j = 0; j < 4; ++j) if (nb + j < t.N) t.pre[o + j] = x[j];
          }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) x[j] = apply_act_lean(act, x[j]);
      }
      if (cvec) *reinterpret_cast<float4*>(t.C + o) = make_float4(x[0], x[1], x[2], x[3]);
      else {
#pragma unroll
        for (int j = 0; j < 4; ++j) if (nb + j < t.N) t.C[o + j] = x[j];
      }
    }
  }
  GTC_TS(16);
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 128);
  GTC_TS(17);
}

template <int MODE>
__global__ void __launch_bounds__(256, 1) gemm_tc_kernel(const GemmTask t) {
  gemm_tc_body<MODE>(t, (int)blockIdx.x);
}

// one launch over a device-resident task array (blockIdx.y = task)
template <int MODE>
__global__ void __launch_bounds__(256, 1) gemm_tc_grouped_kernel(const GemmTask* __restrict__ tasks) {
  __shared__ GemmTask t;
  if (threadIdx.x == 0) t = tasks[blockIdx.y];
  __syncthreads();
  gemm_tc_body<MODE>(t, (int)blockIdx.x);
}

// Shapes worth a 128 x 128 tensor-core tile; anything else stays on the FFMA kernel.
bool gemm_tc_eligible(GemmMode mode, const GemmTask& t) {
  static const bool off = getenv("IQLB_GEMM_TC") && atoi(getenv("IQLB_GEMM_TC")) == 0;
  if (off || !t.allow_tc) return false;
  if (mode != GEMM_LINEAR && mode != GEMM_FWD_KN) return false;
  if (t.drop_thresh != 0) return false;
  return t.M >= 128 && t.N >= 64 && t.K >= 16;
}

int launch_gemm_tc(GemmMode mode, const GemmTask& t, cudaStream_t stream) {
  static bool attr = false;
  if (!attr) {
    IQLB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<GEMM_LINEAR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gtc::kSmem));
    IQLB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<GEMM_FWD_KN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gtc::kSmem));
    attr = true;
  }
  const unsigned grid = (unsigned)(((t.M + gtc::BM - 1) / gtc::BM) * ((t.N + gtc::BN - 1) / gtc::BN));
  if (mode == GEMM_LINEAR) gemm_tc_kernel<GEMM_LINEAR><<<grid, 256, gtc::kSmem, stream>>>(t);
  else gemm_tc_kernel<GEMM_FWD_KN><<<grid, 256, gtc::kSmem, stream>>>(t);
  IQLB_LAUNCH_CHECK();
  return IQLB_OK;
}

// Weight-gradient partials of the reward-model trainers: tasks[i] is one row chunk, every task has the same M x N.
int launch_grouped_gemm_tc_bwd_weight(const GemmTask* d_tasks, int n_tasks, int M, int N, cudaStream_t stream) {
  static bool attr = false;
  if (!attr) {
    IQLB_CUDA(cudaFuncSetAttribute(gemm_tc_grouped_kernel<GEMM_BWD_WEIGHT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)gtc::kSmem));
    attr = true;
  }
  const dim3 grid((unsigned)(((M + gtc::BM - 1) / gtc::BM) * ((N + gtc::BN - 1) / gtc::BN)), (unsigned)n_tasks);
  gemm_tc_grouped_kernel<GEMM_BWD_WEIGHT><<<grid, 256, gtc::kSmem, stream>>>(d_tasks);
  IQLB_LAUNCH_CHECK();
  return IQLB_OK;
}

// Grouped launch of any arrangement over a device-resident task array (every task at most maxM x maxN): the batched small
// GEMMs of the preference transformer's attention (Q K^T, P V and their gradients per (sequence, head), pt_train.cu).
int launch_grouped_gemm_tc(GemmMode mode, const GemmTask* d_tasks, int n_tasks, int maxM, int maxN, cudaStream_t stream) {
  if (mode == GEMM_BWD_WEIGHT) return launch_grouped_gemm_tc_bwd_weight(d_tasks, n_tasks, maxM, maxN, stream);
  static bool attr = false;
  if (!attr) {
    IQLB_CUDA(cudaFuncSetAttribute(gemm_tc_grouped_kernel<GEMM_LINEAR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gtc::kSmem));
    IQLB_CUDA(cudaFuncSetAttribute(gemm_tc_grouped_kernel<GEMM_FWD_KN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gtc::kSmem));
    attr = true;
  }
  const dim3 grid((unsigned)(((maxM + gtc::BM - 1) / gtc::BM) * ((maxN + gtc::BN - 1) / gtc::BN)), (unsigned)n_tasks);
  if (mode == GEMM_LINEAR) gemm_tc_grouped_kernel<GEMM_LINEAR><<<grid, 256, gtc::kSmem, stream>>>(d_tasks);
  else if (mode == GEMM_FWD_KN) gemm_tc_grouped_kernel<GEMM_FWD_KN><<<grid, 256, gtc::kSmem, stream>>>(d_tasks);
  else return fail(IQLB_ERR_UNSUPPORTED, "launch_grouped_gemm_tc: unsupported mode%s", "");
  IQLB_LAUNCH_CHECK();
  return IQLB_OK;
}

}  // namespace iqlb

using namespace iqlb;

// Diagnostic entry: one GEMM through launch_gemm (tests/test_gpu_gemm_tc.py pins the tensor-core kernel against numpy
// for both operand arrangements, ragged shapes and unaligned leading dimensions).
extern "C" int iqlb_selftest_gemm(int32_t kn_layout, const float* a, int32_t lda, const float* w, int32_t ldw,
                                  const float* bias, const float* residual, float* c, int32_t ldc, int32_t m, int32_t n,
                                  int32_t k, int32_t activation, float scale, int32_t allow_tc, void* stream) {
  static unsigned long long* dbg_buf = nullptr;  // IQLB_GEMM_TL=1: stamps of one CTA, printed after the call (synchronises)
  static const bool tl = getenv("IQLB_GEMM_TL") && atoi(getenv("IQLB_GEMM_TL")) != 0;
  if (tl && !dbg_buf) cudaMalloc((void**)&dbg_buf, 32 * sizeof(unsigned long long));
  IQLB_REQUIRE(a && w && c, "null pointer");
  IQLB_REQUIRE(m > 0 && n > 0 && k > 0, "shape");
  IQLB_REQUIRE(!kn_layout || bias != nullptr, "the [K][N] arrangement needs a bias vector");
  GemmTask t{};
  t.A = a; t.lda = lda; t.Bm = w; t.ldb = ldw; t.bias = bias; t.mask = kn_layout ? nullptr : residual;
  t.C = c; t.ldc = ldc; t.M = m; t.N = n; t.K = k; t.act = activation; t.scale = scale; t.allow_tc = allow_tc;
  t.dbg = tl ? dbg_buf : nullptr;
  const int rc = launch_gemm(kn_layout ? GEMM_FWD_KN : GEMM_LINEAR, t, (cudaStream_t)stream);
  if (tl && allow_tc && rc == 0) {
    unsigned long long h[32];
    cudaStreamSynchronize((cudaStream_t)stream);
    cudaMemcpy(h, dbg_buf, sizeof(h), cudaMemcpyDeviceToHost);
    fprintf(stderr, "[gemm_tc timeline m=%d n=%d k=%d] setup %llu |", m, n, k, h[1] - h[0]);
    for (int kb = 0; kb < 4 && kb * 64 < k; ++kb)
      fprintf(stderr, " kb%d: wait@%llu store %llu load+sync %llu |", kb, h[2 + 3 * kb] - h[0], h[3 + 3 * kb] - h[2 + 3 * kb],
              h[4 + 3 * kb] - h[3 + 3 * kb]);
    fprintf(stderr, " loop end@%llu mma done +%llu epilogue %llu dealloc %llu\n", h[14] - h[0], h[15] - h[14], h[16] - h[15],
            h[17] - h[16]);
  }
  return rc;
}
