// Dataset preparation on the device (the step between the relabel and the replay buffer):
//   return_reward_range  (/root/reference/algorithms/offline/iql.py:344-360)  episode returns -> (min, max), trj_lens
//   modify_reward        (iql.py:363-401)                                     the normalize_reward variants 0..7
//   compute_mean_std / normalize_states (iql.py:132-139)
// The reference walks the dataset in a Python loop (minutes per run, once per normalize_reward agent).  Here:
//   * episode boundaries = terminal, or episode length == max_episode_steps.  Between two terminals ("run") the
//     boundaries are periodic, so a max-scan of the last terminal index gives every row its position in its run and
//     hence its position in its episode;
//   * the thread that sits on an episode's LAST row walks the episode once, first row to last, adding the rewards
//     in double in exactly the reference's order (ep_ret += float(r)): min / max of the returns are bit-identical to
//     the reference's, and the same walk writes the episode length into trj_lens;
//   * modify_reward replays numpy's in-place op sequence with its dtypes (fp32 array, weak Python scalars, fp64
//     temporaries for min_ret / trj_lens).
// HBM-bound (5 B read + 4-8 B written per row), a handful of launches, no host synchronisation.
#include <cub/block/block_reduce.cuh>
#include <cub/block/block_scan.cuh>

#include "common.cuh"

namespace iqlb {

constexpr int kScanBlock = 1024;

struct MaxOp {
  __device__ __forceinline__ long long operator()(long long a, long long b) const { return a > b ? a : b; }
};

// per block: index + 1 of the last terminal inside the block (0: none)
__global__ void __launch_bounds__(kScanBlock) last_terminal_kernel(const uint8_t* __restrict__ terminals, int64_t n,
                                                                   long long* __restrict__ blk_last) {
  const int64_t j = (int64_t)blockIdx.x * kScanBlock + threadIdx.x;
  long long v = (j < n && terminals[j]) ? (long long)j + 1 : 0;
  typedef cub::BlockReduce<long long, kScanBlock> Reduce;
  __shared__ typename Reduce::TempStorage tmp;
  const long long m = Reduce(tmp).Reduce(v, MaxOp());
  if (threadIdx.x == 0) blk_last[blockIdx.x] = m;
}

// one block: carry[b] = index + 1 of the last terminal before block b
__global__ void __launch_bounds__(1024) carry_scan_kernel(const long long* __restrict__ blk_last, int n_blocks,
                                                          long long* __restrict__ carry) {
  typedef cub::BlockScan<long long, 1024> Scan;
  __shared__ typename Scan::TempStorage tmp;
  __shared__ long long running;
  if (threadIdx.x == 0) running = 0;
  __syncthreads();
  for (int b0 = 0; b0 < n_blocks; b0 += 1024) {
    const int b = b0 + threadIdx.x;
    const long long v = b < n_blocks ? blk_last[b] : 0;
    long long ex, total;
    Scan(tmp).ExclusiveScan(v, ex, 0ll, MaxOp(), total);
    const long long base = running;
    if (b < n_blocks) carry[b] = ex > base ? ex : base;
    __syncthreads();
    if (threadIdx.x == 0) running = total > base ? total : base;
    __syncthreads();
  }
}

__device__ __forceinline__ unsigned long long double_key(double d) {  // order-preserving
  const unsigned long long u = (unsigned long long)__double_as_longlong(d);
  return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
}
__device__ __forceinline__ double key_double(unsigned long long k) {
  return __longlong_as_double((long long)((k >> 63) ? (k & 0x7fffffffffffffffull) : ~k));
}

// range[0] = min key, range[1] = max key (initialised to ~0 / 0), range[2] = number of completed episodes
__global__ void __launch_bounds__(kScanBlock) episode_walk_kernel(const float* __restrict__ rewards,
                                                                  const uint8_t* __restrict__ terminals, int64_t n,
                                                                  int max_steps, const long long* __restrict__ carry,
                                                                  unsigned long long* __restrict__ range,
                                                                  float* __restrict__ trj_lens) {
  const int64_t j = (int64_t)blockIdx.x * kScanBlock + threadIdx.x;
  const bool term = j < n && terminals[j] != 0;
  // run start of row j = index + 1 of the last terminal BEFORE j: exclusive max-scan of (terminal ? j + 1 : 0)
  typedef cub::BlockScan<long long, kScanBlock> Scan;
  __shared__ typename Scan::TempStorage tmp;
  long long ex;
  Scan(tmp).ExclusiveScan(term ? (long long)j + 1 : 0ll, ex, 0ll, MaxOp());
  if (j >= n) return;
  const long long c = carry[blockIdx.x];
  const int64_t run_start = ex > c ? ex : c;
  const int64_t pos = j - run_start;             // position inside the run
  const int len = (int)(pos % max_steps) + 1;    // episode length up to and including row j
  const bool ends = term || len == max_steps;
  if (!ends && j != n - 1) return;
  // this row closes an episode (or is the trailing, unfinished one): walk it in the reference's order
  const int64_t start = j - len + 1;
  double ret = 0.0;
  for (int64_t k = start; k <= j; ++k) {
    ret += (double)rewards[k];
    if (trj_lens) trj_lens[k] = (float)len;
  }
  if (ends) {
    const unsigned long long key = double_key(ret);
    atomicMin(range + 0, key);
    atomicMax(range + 1, key);
    atomicAdd(range + 2, 1ull);
  }
}

__global__ void range_init_kernel(unsigned long long* __restrict__ range) {
  range[0] = ~0ull; range[1] = 0ull; range[2] = 0ull;
}

__global__ void range_finish_kernel(const unsigned long long* __restrict__ range, double* __restrict__ min_max) {
  min_max[0] = key_double(range[0]);
  min_max[1] = key_double(range[1]);
}

// numpy's in-place sequence of modify_reward (iql.py:363-401).  kind: 0 locomotion (halfcheetah / hopper / walker2d),
// 1 antmaze; other environments are left untouched by the reference.
__global__ void __launch_bounds__(256) modify_reward_kernel(float* __restrict__ rewards, int64_t n, int kind, int nr,
                                                            int max_steps, const double* __restrict__ min_max,
                                                            const float* __restrict__ trj_lens) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float r = rewards[i];
  const bool ranged = kind == 0 || nr != 1;
  const double mn = ranged ? min_max[0] : 0.0, mx = ranged ? min_max[1] : 1.0;
  const float span = (float)(mx - mn), steps = (float)max_steps;
  // (explicit _rn intrinsics: numpy rounds after every step, so the multiply and the "- 1" must not fuse into an FMA)
  if (kind == 0) {
    r = __fmul_rn(__fdiv_rn(r, span), steps);
  } else if (nr == 1) {
    r = __fsub_rn(r, 1.0f);
  } else {
    if (nr == 4 || nr == 5) r = __fsub_rn(r, (float)mn);
    else if (nr != 2 && nr != 3) r = (float)__dsub_rn((double)r, __ddiv_rn(mn, (double)trj_lens[i]));  // fp64 temporary
    r = __fmul_rn(__fdiv_rn(r, span), steps);
    if (nr == 3 || nr == 5 || nr >= 7 || nr < 1) r = __fsub_rn(r, 1.0f);
  }
  rewards[i] = r;
}

// column sums in double: grid (row chunks), block 256 threads = 256 / dim_pad rows x dim_pad columns
__global__ void __launch_bounds__(256) column_moments_kernel(const float* __restrict__ x, int64_t n, int dim, int ld,
                                                             const double* __restrict__ mean, double* __restrict__ acc) {
  // pass 0 (mean == nullptr): acc[c] += sum x;  pass 1: acc[c] += sum (x - mean)^2
  const int c = threadIdx.x % 32, lane_rows = blockDim.x / 32, r0 = threadIdx.x / 32;
  for (int c0 = 0; c0 < dim; c0 += 32) {
    const int col = c0 + c;
    double s = 0.0;
    if (col < dim) {
      const double m = mean ? mean[col] : 0.0;
      for (int64_t r = (int64_t)blockIdx.x * lane_rows + r0; r < n; r += (int64_t)gridDim.x * lane_rows) {
        const double v = (double)x[r * ld + col];
        s += mean ? (v - m) * (v - m) : v;
      }
    }
    __shared__ double red[8][32];
    red[r0][c] = s;
    __syncthreads();
    if (r0 == 0 && col < dim) {
      double t = 0.0;
      for (int k = 0; k < lane_rows; ++k) t += red[k][c];
      atomicAdd(acc + col, t);
    }
    __syncthreads();
  }
}

__global__ void mean_std_finish_kernel(double* __restrict__ acc, int64_t n, int dim, int pass, float eps,
                                       double* __restrict__ mean_d, float* __restrict__ mean, float* __restrict__ std) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= dim) return;
  if (pass == 0) {
    mean_d[c] = acc[c] / (double)n;
    mean[c] = (float)mean_d[c];
    acc[c] = 0.0;
  } else {
    std[c] = (float)sqrt(acc[c] / (double)n) + eps;
  }
}

__global__ void __launch_bounds__(256) normalize_states_kernel(float* __restrict__ x, int64_t n, int dim, int ld,
                                                               const float* __restrict__ mean,
                                                               const float* __restrict__ std) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n * dim) return;
  const int64_t r = i / dim;
  const int c = (int)(i - r * dim);
  x[r * ld + c] = (x[r * ld + c] - mean[c]) / std[c];
}

}  // namespace iqlb

using namespace iqlb;

extern "C" int iqlb_reward_range(const float* rewards, const uint8_t* terminals, int64_t n, int32_t max_episode_steps,
                                 double* min_max, float* trj_lens, void* stream) {
  IQLB_REQUIRE(rewards && terminals && min_max, "null pointer");
  IQLB_REQUIRE(n >= 1 && max_episode_steps >= 1, "empty dataset / episode cap");
  cudaStream_t st = (cudaStream_t)stream;
  const int nb = (int)((n + kScanBlock - 1) / kScanBlock);
  long long* scratch = nullptr;  // blk_last[nb] | carry[nb] | range[3]
  IQLB_CUDA(cudaMallocAsync((void**)&scratch, (size_t)(2 * nb + 3) * sizeof(long long), st));
  long long *blk_last = scratch, *carry = scratch + nb;
  unsigned long long* range = reinterpret_cast<unsigned long long*>(scratch + 2 * nb);
  cudaError_t e = cudaSuccess;
  {
    range_init_kernel<<<1, 1, 0, st>>>(range);
    last_terminal_kernel<<<nb, kScanBlock, 0, st>>>(terminals, n, blk_last);
    carry_scan_kernel<<<1, 1024, 0, st>>>(blk_last, nb, carry);
    episode_walk_kernel<<<nb, kScanBlock, 0, st>>>(rewards, terminals, n, max_episode_steps, carry, range, trj_lens);
    range_finish_kernel<<<1, 1, 0, st>>>(range, min_max);
    e = cudaGetLastError();
  }
  cudaFreeAsync(scratch, st);
  if (e != cudaSuccess) return fail(IQLB_ERR_CUDA, "iqlb_reward_range: %s", cudaGetErrorString(e));
  return IQLB_OK;
}

// ---- allocation-free variant (no counterpart in the reference) --------------
// Same launches as iqlb_reward_range, with the scratch (blk_last[nb] | carry[nb] | range[3]) in a file-scope device array
// instead of cudaMallocAsync: for callers that may not allocate (CUDA graph capture, tight serving loops).  256 KB per GPU
// that loads the library, held until exit.  Every call on a device shares it: calls that overlap on the GPU race.
static constexpr int64_t kStaticMaxRows = IQLB_REWARD_RANGE_STATIC_MAX_ROWS;
static constexpr int kStaticMaxBlocks = (int)((kStaticMaxRows + kScanBlock - 1) / kScanBlock);
static __device__ __align__(256) long long g_range_scratch[2 * kStaticMaxBlocks + 3];

extern "C" int iqlb_reward_range_static(const float* rewards, const uint8_t* terminals, int64_t n, int32_t max_episode_steps,
                                        double* min_max, float* trj_lens, void* stream) {
  IQLB_REQUIRE(rewards && terminals && min_max, "null pointer");
  IQLB_REQUIRE(n >= 1 && max_episode_steps >= 1, "empty dataset / episode cap");
  if (n > kStaticMaxRows)
    return fail(IQLB_ERR_INVALID, "invalid argument: iqlb_reward_range_static handles at most %s%lld rows (use iqlb_reward_range)",
                "", (long long)kStaticMaxRows);
  // The array's name is a host-side stand-in in host code: its device address is looked up once per device, with that
  // device current (the caller's), and cached.
  constexpr int kMaxDev = 64;
  static long long* dev_ptr[kMaxDev] = {};
  int dev = 0;
  IQLB_CUDA(cudaGetDevice(&dev));
  IQLB_REQUIRE(dev >= 0 && dev < kMaxDev, "device ordinal");
  if (dev_ptr[dev] == nullptr) {
    long long* p = nullptr;
    IQLB_CUDA(cudaGetSymbolAddress((void**)&p, g_range_scratch));
    dev_ptr[dev] = p;
  }
  cudaStream_t st = (cudaStream_t)stream;
  const int nb = (int)((n + kScanBlock - 1) / kScanBlock);
  long long *blk_last = dev_ptr[dev], *carry = blk_last + nb;
  unsigned long long* range = reinterpret_cast<unsigned long long*>(blk_last + 2 * nb);
  // nothing of the previous call survives: range_init rewrites range[3], the two scans write all nb entries they read
  range_init_kernel<<<1, 1, 0, st>>>(range);
  last_terminal_kernel<<<nb, kScanBlock, 0, st>>>(terminals, n, blk_last);
  carry_scan_kernel<<<1, 1024, 0, st>>>(blk_last, nb, carry);
  episode_walk_kernel<<<nb, kScanBlock, 0, st>>>(rewards, terminals, n, max_episode_steps, carry, range, trj_lens);
  range_finish_kernel<<<1, 1, 0, st>>>(range, min_max);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(IQLB_ERR_CUDA, "iqlb_reward_range_static: %s", cudaGetErrorString(e));
  return IQLB_OK;
}

extern "C" int iqlb_modify_reward(float* rewards, const uint8_t* terminals, int64_t n, int32_t env_kind,
                                  int32_t normalize_reward, int32_t max_episode_steps, void* stream) {
  IQLB_REQUIRE(rewards && terminals, "null pointer");
  IQLB_REQUIRE(n >= 1 && max_episode_steps >= 1, "empty dataset / episode cap");
  IQLB_REQUIRE(env_kind == 0 || env_kind == 1, "env_kind: 0 locomotion, 1 antmaze");
  cudaStream_t st = (cudaStream_t)stream;
  const bool ranged = env_kind == 0 || normalize_reward != 1;
  const bool per_len = env_kind == 1 && normalize_reward != 1 && !(normalize_reward >= 2 && normalize_reward <= 5);
  double* min_max = nullptr;
  float* trj = nullptr;
  if (ranged) {
    IQLB_CUDA(cudaMallocAsync((void**)&min_max, 2 * sizeof(double), st));
    if (per_len) IQLB_CUDA(cudaMallocAsync((void**)&trj, (size_t)n * sizeof(float), st));
    const int rc = iqlb_reward_range(rewards, terminals, n, max_episode_steps, min_max, trj, stream);
    if (rc) { cudaFreeAsync(min_max, st); if (trj) cudaFreeAsync(trj, st); return rc; }
  }
  modify_reward_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(rewards, n, env_kind, normalize_reward,
                                                                    max_episode_steps, min_max, trj);
  const cudaError_t e = cudaGetLastError();
  if (min_max) cudaFreeAsync(min_max, st);
  if (trj) cudaFreeAsync(trj, st);
  if (e != cudaSuccess) return fail(IQLB_ERR_CUDA, "iqlb_modify_reward: %s", cudaGetErrorString(e));
  return IQLB_OK;
}

extern "C" int iqlb_state_mean_std(const float* states, int64_t n, int32_t dim, int32_t ld, float eps, float* mean,
                                   float* std, void* stream) {
  IQLB_REQUIRE(states && mean && std, "null pointer");
  IQLB_REQUIRE(n >= 1 && dim >= 1 && ld >= dim, "shape");
  cudaStream_t st = (cudaStream_t)stream;
  double* acc = nullptr;  // acc[dim] | mean_d[dim]
  IQLB_CUDA(cudaMallocAsync((void**)&acc, (size_t)2 * dim * sizeof(double), st));
  cudaError_t e = cudaMemsetAsync(acc, 0, (size_t)2 * dim * sizeof(double), st);
  if (e == cudaSuccess) {
    int64_t blocks = (n + 7) / 8;
    if (blocks > 148 * 8) blocks = 148 * 8;
    column_moments_kernel<<<(unsigned)blocks, 256, 0, st>>>(states, n, dim, ld, nullptr, acc);
    mean_std_finish_kernel<<<(dim + 127) / 128, 128, 0, st>>>(acc, n, dim, 0, eps, acc + dim, mean, std);
    column_moments_kernel<<<(unsigned)blocks, 256, 0, st>>>(states, n, dim, ld, acc + dim, acc);
    mean_std_finish_kernel<<<(dim + 127) / 128, 128, 0, st>>>(acc, n, dim, 1, eps, acc + dim, mean, std);
    e = cudaGetLastError();
  }
  cudaFreeAsync(acc, st);
  if (e != cudaSuccess) return fail(IQLB_ERR_CUDA, "iqlb_state_mean_std: %s", cudaGetErrorString(e));
  return IQLB_OK;
}

extern "C" int iqlb_normalize_states(float* states, int64_t n, int32_t dim, int32_t ld, const float* mean,
                                     const float* std, void* stream) {
  IQLB_REQUIRE(states && mean && std, "null pointer");
  IQLB_REQUIRE(n >= 0 && dim >= 1 && ld >= dim, "shape");
  if (n == 0) return IQLB_OK;
  normalize_states_kernel<<<(unsigned)((n * dim + 255) / 256), 256, 0, (cudaStream_t)stream>>>(states, n, dim, ld, mean, std);
  IQLB_LAUNCH_CHECK();
  return IQLB_OK;
}
