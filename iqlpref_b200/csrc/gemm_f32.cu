// Grouped fp32 FFMA GEMM (parity mode).  One launch covers one layer of every (agent, network)
// pair: blockIdx.y = task, blockIdx.x = output tile.  Three operand arrangements:
//   FWD        C[m,n] = relu(sum_k A[m,k] * W[n,k] + bias[n])        (nn.Linear forward)
//   BWD_DATA   C[m,n] = mask[m,n] > 0 ? sum_k A[m,k] * W[k,n] : 0     (dX = dY W, through ReLU)
//   BWD_WEIGHT C[m,n] = sum_k A[k,m] * X[k,n];  bias_grad[m] = sum_k A[k,m]   (dW = dY^T X, db)
//   FWD_KN     C[m,n] = act(scale * sum_k A[m,k] * W[k,n] + bias[n])   (RewardMLP scaled Linear; pre-activation kept in t.pre)
//   LINEAR     C[m,n] = act(sum_k A[m,k] * W[n,k] + bias[n]) (+ residual); with t.pre: scale * acc * act'(pre[m,n])
// 128 x BN tile, BK = 8, 256 threads, 8 x TN outputs per thread, register-prefetched.
#include "iql_step.cuh"

namespace iqlb {

constexpr int BM = 128;
constexpr int BK = 8;
constexpr int TM = 8;

// element (r, k) of an operand tile: KC = k contiguous (ptr[r*ld+k]), RC = r contiguous (ptr[k*ld+r])
template <bool KC, int ROWS>
__device__ __forceinline__ void load_tile(const float* __restrict__ p, int ld, int r0, int rmax, int k0, int kmax,
                                          float (&reg)[ROWS * BK / 256]) {
  constexpr int PER = ROWS * BK / 256;
#pragma unroll
  for (int i = 0; i < PER; ++i) {
    const int e = threadIdx.x + i * 256;
    int r, k;
    if (KC) { r = e / BK; k = e % BK; } else { k = e / ROWS; r = e % ROWS; }
    const int gr = r0 + r, gk = k0 + k;
    float v = 0.f;
    if (gr < rmax && gk < kmax) v = KC ? __ldg(p + (int64_t)gr * ld + gk) : __ldg(p + (int64_t)gk * ld + gr);
    reg[i] = v;
  }
}

template <bool KC, int ROWS>
__device__ __forceinline__ void store_tile(float (*s)[ROWS + 4], const float (&reg)[ROWS * BK / 256]) {
  constexpr int PER = ROWS * BK / 256;
#pragma unroll
  for (int i = 0; i < PER; ++i) {
    const int e = threadIdx.x + i * 256;
    int r, k;
    if (KC) { r = e / BK; k = e % BK; } else { k = e / ROWS; r = e % ROWS; }
    s[k][r] = reg[i];
  }
}

template <int MODE, int BN>
__device__ __forceinline__ void gemm_body(const GemmTask& t, int tile) {
  constexpr int TN = BN / 16;
  constexpr bool A_KC = (MODE != GEMM_BWD_WEIGHT);
  constexpr bool B_KC = (MODE == GEMM_FWD || MODE == GEMM_LINEAR);
  const int tiles_n = (t.N + BN - 1) / BN;
  const int tiles_m = (t.M + BM - 1) / BM;
  if (tile >= tiles_m * tiles_n) return;
  const int tm = tile / tiles_n, tn = tile % tiles_n;
  const int m0 = tm * BM, n0 = tn * BN;

  // two stages: the tile of step k+1 is written while step k is being read, one barrier per step
  __shared__ __align__(16) float As2[2][BK][BM + 4];
  __shared__ __align__(16) float Bs2[2][BK][BN + 4];

  const int ty = threadIdx.x / 16, tx = threadIdx.x % 16;
  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;
  float asum[TM];
#pragma unroll
  for (int i = 0; i < TM; ++i) asum[i] = 0.f;

  float ra[BM * BK / 256], rb[BN * BK / 256];
  load_tile<A_KC, BM>(t.A, t.lda, m0, t.M, 0, t.K, ra);
  load_tile<B_KC, BN>(t.Bm, t.ldb, n0, t.N, 0, t.K, rb);
  const bool want_bsum = (MODE == GEMM_BWD_WEIGHT) && tn == 0 && t.bias_grad != nullptr;

  store_tile<A_KC, BM>(As2[0], ra);
  store_tile<B_KC, BN>(Bs2[0], rb);
  __syncthreads();
  int stage = 0;
  for (int k0 = 0; k0 < t.K; k0 += BK, stage ^= 1) {
    float (*As)[BM + 4] = As2[stage];
    float (*Bs)[BN + 4] = Bs2[stage];
    const bool more = k0 + BK < t.K;
    if (more) {
      load_tile<A_KC, BM>(t.A, t.lda, m0, t.M, k0 + BK, t.K, ra);
      load_tile<B_KC, BN>(t.Bm, t.ldb, n0, t.N, k0 + BK, t.K, rb);
    }
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      float a[TM], b[TN];
      // rows ty*4..+3 and 64+ty*4..+3: two conflict-free 128-bit reads
      *(float4*)&a[0] = *(const float4*)&As[k][ty * 4];
      *(float4*)&a[4] = *(const float4*)&As[k][BM / 2 + ty * 4];
      if constexpr (TN == 8) {
        *(float4*)&b[0] = *(const float4*)&Bs[k][tx * 4];
        *(float4*)&b[4] = *(const float4*)&Bs[k][BN / 2 + tx * 4];
      } else if constexpr (TN == 4) {
        *(float4*)&b[0] = *(const float4*)&Bs[k][tx * 4];
      } else {
#pragma unroll
        for (int j = 0; j < TN; ++j) b[j] = Bs[k][tx * TN + j];
      }
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
      if constexpr (MODE == GEMM_BWD_WEIGHT) {
#pragma unroll
        for (int i = 0; i < TM; ++i) asum[i] += a[i];
      }
    }
    if (more) {  // the other stage was last read before the previous barrier
      store_tile<A_KC, BM>(As2[stage ^ 1], ra);
      store_tile<B_KC, BN>(Bs2[stage ^ 1], rb);
    }
    __syncthreads();
  }

  // full-width tiles of the row-major outputs: 128-bit stores (a warp writes two 256-byte row segments per instruction)
  if constexpr (TN == 8 && (MODE == GEMM_FWD_KN || MODE == GEMM_LINEAR)) {
    auto al16 = [](const void* p) { return p == nullptr || (reinterpret_cast<uintptr_t>(p) & 15u) == 0; };
    if (n0 + BN <= t.N && (t.ldc & 3) == 0 && al16(t.C) && al16(t.pre) && al16(t.mask) && al16(t.bias)) {
#pragma unroll
      for (int i = 0; i < TM; ++i) {
        const int m = m0 + (i < 4 ? ty * 4 + i : BM / 2 + ty * 4 + (i - 4));
        if (m >= t.M) continue;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int n = n0 + h * (BN / 2) + tx * 4;
          const int64_t o = (int64_t)m * t.ldc + n;
          float v[4] = {acc[i][h * 4], acc[i][h * 4 + 1], acc[i][h * 4 + 2], acc[i][h * 4 + 3]};
          float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
          if (t.bias != nullptr) b4 = __ldg(reinterpret_cast<const float4*>(t.bias + n));
          const float b[4] = {b4.x, b4.y, b4.z, b4.w};
          if (MODE == GEMM_FWD_KN) {
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] = fmaf(t.scale, v[j], b[j]);
            if (t.pre != nullptr) *reinterpret_cast<float4*>(t.pre + o) = make_float4(v[0], v[1], v[2], v[3]);
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] = apply_act_lean(t.act, v[j]);
          } else if (t.pre != nullptr) {  // data gradient through the activation below: scale * acc * act'(z)
            const float4 z4 = *reinterpret_cast<const float4*>(t.pre + o);
            const float z[4] = {z4.x, z4.y, z4.z, z4.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] = t.scale * v[j] * act_grad_lean(t.act, z[j]);
          } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] = apply_act_lean(t.act, v[j] + b[j]);
            if (t.mask != nullptr) {
              const float4 r4 = *reinterpret_cast<const float4*>(t.mask + o);
              v[0] += r4.x; v[1] += r4.y; v[2] += r4.z; v[3] += r4.w;
            }
          }
          *reinterpret_cast<float4*>(t.C + o) = make_float4(v[0], v[1], v[2], v[3]);
        }
      }
      return;
    }
  }
  uint64_t drop_seed = 0, drop_ctr = 0;
  if (MODE == GEMM_FWD && t.drop_thresh != 0) { drop_seed = *t.drop_rng; drop_ctr = (uint64_t)(*t.drop_it) * 4ull; }
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const int m = m0 + (i < 4 ? ty * 4 + i : BM / 2 + ty * 4 + (i - 4));
    if (m >= t.M) continue;
    uint32_t dr[4] = {0, 0, 0, 0};
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      int n;
      if (TN == 8) n = n0 + (j < 4 ? tx * 4 + j : BN / 2 + tx * 4 + (j - 4));
      else n = n0 + tx * TN + j;
      if (n >= t.N) continue;
      float v = acc[i][j];
      if (MODE == GEMM_FWD) {
        v += __ldg(t.bias + n);
        v = v > 0.f ? v : 0.f;
        if (t.drop_thresh != 0) {
          const int64_t e = (int64_t)m * t.N + n;
          if ((j & 3) == 0 || (n & 3) == 0 || TN != 8)  // columns come in aligned groups of four: one Philox call per group
            philox4(drop_seed, ((uint64_t)(t.drop_layer + 1) << 32) | (uint64_t)(e >> 2), drop_ctr, dr);
          v = dr[n & 3] >= t.drop_thresh ? v * t.drop_scale : 0.f;
        }
      } else if (MODE == GEMM_BWD_DATA) {
        v = (__ldg(t.mask + (int64_t)m * t.ldc + n) > 0.f) ? v * t.scale : 0.f;
      } else if (MODE == GEMM_FWD_KN) {
        v = fmaf(t.scale, v, __ldg(t.bias + n));
        if (t.pre != nullptr) t.pre[(int64_t)m * t.ldc + n] = v;
        v = apply_act_lean(t.act, v);
      } else if (MODE == GEMM_LINEAR && t.pre != nullptr) {  // scale * (A W^T) * act'(pre): back through an activation
        v = t.scale * v * act_grad_lean(t.act, t.pre[(int64_t)m * t.ldc + n]);
      } else if (MODE == GEMM_LINEAR) {  // nn.Linear: act(A W^T + bias) (+ residual, carried in t.mask)
        v = apply_act_lean(t.act, v + (t.bias ? __ldg(t.bias + n) : 0.f));
        if (t.mask != nullptr) v += t.mask[(int64_t)m * t.ldc + n];
      }
      t.C[(int64_t)m * t.ldc + n] = v;
    }
    if (want_bsum && tx == 0) t.bias_grad[m] = asum[i];
  }
}

template <int MODE, int BN>
__global__ void __launch_bounds__(256, 2) grouped_gemm_kernel(const GemmTask* __restrict__ tasks) {
  const GemmTask t = tasks[blockIdx.y];
  gemm_body<MODE, BN>(t, (int)blockIdx.x);
}

template <int MODE, int BN>
__global__ void __launch_bounds__(256, 2) single_gemm_kernel(const GemmTask t) {
  gemm_body<MODE, BN>(t, (int)blockIdx.x);
}

// One GEMM with the task passed by value (no device-side task array).
int launch_gemm(GemmMode mode, const GemmTask& t, cudaStream_t stream) {
  if (t.M <= 0 || t.N <= 0) return IQLB_OK;
  if (gemm_tc_eligible(mode, t)) return launch_gemm_tc(mode, t, stream);
  const int tiles_m = (t.M + BM - 1) / BM;
  const bool narrow = t.N <= 32;
  const unsigned grid = (unsigned)(tiles_m * (narrow ? (t.N + 31) / 32 : (t.N + 127) / 128));
  switch (mode) {
    case GEMM_LINEAR:
      if (narrow) single_gemm_kernel<GEMM_LINEAR, 32><<<grid, 256, 0, stream>>>(t);
      else single_gemm_kernel<GEMM_LINEAR, 128><<<grid, 256, 0, stream>>>(t);
      break;
    case GEMM_FWD_KN:
      if (narrow) single_gemm_kernel<GEMM_FWD_KN, 32><<<grid, 256, 0, stream>>>(t);
      else single_gemm_kernel<GEMM_FWD_KN, 128><<<grid, 256, 0, stream>>>(t);
      break;
    default:
      return fail(IQLB_ERR_UNSUPPORTED, "launch_gemm: unsupported mode%s", "");
  }
  IQLB_LAUNCH_CHECK();
  return IQLB_OK;
}

template <int MODE>
static int launch_mode(const GemmTask* d_tasks, int n_tasks, int maxM, int maxN, bool narrow_n, cudaStream_t stream) {
  const int tiles_m = (maxM + BM - 1) / BM;
  if (narrow_n) {
    dim3 grid((unsigned)(tiles_m * ((maxN + 31) / 32)), (unsigned)n_tasks);
    grouped_gemm_kernel<MODE, 32><<<grid, 256, 0, stream>>>(d_tasks);
  } else {
    dim3 grid((unsigned)(tiles_m * ((maxN + 127) / 128)), (unsigned)n_tasks);
    grouped_gemm_kernel<MODE, 128><<<grid, 256, 0, stream>>>(d_tasks);
  }
  IQLB_LAUNCH_CHECK();
  return IQLB_OK;
}

int launch_grouped_gemm(GemmMode mode, const GemmTask* d_tasks, int n_tasks, int maxM, int maxN, bool narrow_n,
                        cudaStream_t stream) {
  switch (mode) {
    case GEMM_FWD: return launch_mode<GEMM_FWD>(d_tasks, n_tasks, maxM, maxN, narrow_n, stream);
    case GEMM_BWD_DATA: return launch_mode<GEMM_BWD_DATA>(d_tasks, n_tasks, maxM, maxN, narrow_n, stream);
    case GEMM_FWD_KN: return launch_mode<GEMM_FWD_KN>(d_tasks, n_tasks, maxM, maxN, narrow_n, stream);
    case GEMM_LINEAR: return launch_mode<GEMM_LINEAR>(d_tasks, n_tasks, maxM, maxN, narrow_n, stream);
    default: return launch_mode<GEMM_BWD_WEIGHT>(d_tasks, n_tasks, maxM, maxN, narrow_n, stream);
  }
}

}  // namespace iqlb
