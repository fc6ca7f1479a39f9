// Replay sampling: Philox-indexed row gather with 128-bit loads.
// Replaces ReplayBuffer.sample (/root/reference/algorithms/offline/iql.py:211-221).
#include "common.cuh"

namespace iqlb {

thread_local char g_err[512] = "";

__global__ void philox_randint_kernel(uint64_t seed, uint64_t offset, uint64_t n_rows, int batch,
                                      int64_t* __restrict__ idx_out) {
  int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= batch) return;
  uint32_t r[4];
  philox4(seed, (uint64_t)b, offset, r);
  idx_out[b] = randint_from_philox(r, n_rows);
}

// One thread per (row, 16-byte vector).  Threads of a row are adjacent, so a warp reads whole
// rows as consecutive LDG.128s and writes the packed batch fully coalesced.  HBM-bound: random
// row reads (row_stride*4 bytes, 32 B sectors) + sequential writes.
template <bool kSample>
__global__ void __launch_bounds__(256) replay_gather_kernel(
    const float4* __restrict__ table, uint64_t n_rows, int vec_per_row, int reward_col,
    const float* __restrict__ reward_cols, const int32_t* __restrict__ reward_variant,
    const uint64_t* __restrict__ rng, const int64_t* __restrict__ idx_in, int batch,
    int64_t* __restrict__ idx_out, float4* __restrict__ batch_out) {
  const int agent = blockIdx.y;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int b = t / vec_per_row;
  const int v = t - b * vec_per_row;
  if (b >= batch) return;
  int64_t row;
  if (kSample) {
    uint32_t r[4];
    philox4(rng[2 * agent], (uint64_t)b, rng[2 * agent + 1], r);
    row = randint_from_philox(r, n_rows);
  } else {
    row = idx_in[(int64_t)agent * batch + b];
  }
  if (idx_out != nullptr && v == 0) idx_out[(int64_t)agent * batch + b] = row;
  float4 val = __ldg(table + row * vec_per_row + v);
  if (reward_cols != nullptr && (reward_col >> 2) == v) {
    const int variant = reward_variant ? reward_variant[agent] : 0;
    float rw = __ldg(reward_cols + (int64_t)variant * (int64_t)n_rows + row);
    switch (reward_col & 3) {
      case 0: val.x = rw; break;
      case 1: val.y = rw; break;
      case 2: val.z = rw; break;
      default: val.w = rw; break;
    }
  }
  batch_out[((int64_t)agent * batch + b) * vec_per_row + v] = val;
}

int launch_gather(const float* table, int64_t n_rows, int row_stride, int reward_col,
                  const float* reward_cols, const int32_t* reward_variant, const uint64_t* rng,
                  const int64_t* idx_in, int n_agents, int batch, int64_t* idx_out, float* batch_out,
                  cudaStream_t stream) {
  IQLB_REQUIRE(table && batch_out, "null table/batch_out");
  IQLB_REQUIRE(n_rows > 0 && batch > 0 && n_agents > 0, "empty sample");
  IQLB_REQUIRE(row_stride > 0 && row_stride % 4 == 0, "row_stride must be a multiple of 4 floats");
  IQLB_REQUIRE(((uintptr_t)table & 15) == 0 && ((uintptr_t)batch_out & 15) == 0, "16-byte alignment");
  IQLB_REQUIRE(reward_col >= 0 && reward_col < row_stride, "reward_col");
  const int vpr = row_stride / 4;
  dim3 grid((unsigned)(((int64_t)batch * vpr + 255) / 256), (unsigned)n_agents);
  if (rng != nullptr) {
    replay_gather_kernel<true><<<grid, 256, 0, stream>>>((const float4*)table, (uint64_t)n_rows, vpr, reward_col,
                                                         reward_cols, reward_variant, rng, nullptr, batch, idx_out,
                                                         (float4*)batch_out);
  } else {
    IQLB_REQUIRE(idx_in != nullptr, "null idx");
    replay_gather_kernel<false><<<grid, 256, 0, stream>>>((const float4*)table, (uint64_t)n_rows, vpr, reward_col,
                                                          reward_cols, reward_variant, nullptr, idx_in, batch,
                                                          idx_out, (float4*)batch_out);
  }
  IQLB_LAUNCH_CHECK();
  return IQLB_OK;
}

__global__ void polyak_kernel(float4* __restrict__ target, const float4* __restrict__ source, int64_t n4,
                              float* __restrict__ target_tail, const float* __restrict__ source_tail, int tail,
                              float tau) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n4; i += stride) {
    float4 t = target[i];
    const float4 s = __ldg(source + i);
    // Tensor.lerp_(end, w) for w < 0.5: self + w * (end - self)
    t.x = t.x + tau * (s.x - t.x);
    t.y = t.y + tau * (s.y - t.y);
    t.z = t.z + tau * (s.z - t.z);
    t.w = t.w + tau * (s.w - t.w);
    target[i] = t;
  }
  if (blockIdx.x == 0 && (int)threadIdx.x < tail) {
    float t = target_tail[threadIdx.x];
    target_tail[threadIdx.x] = t + tau * (source_tail[threadIdx.x] - t);
  }
}

}  // namespace iqlb

using namespace iqlb;

extern "C" int iqlb_version(void) { return IQLB_VERSION; }
extern "C" const char* iqlb_last_error(void) { return g_err; }

extern "C" int iqlb_philox_randint(uint64_t seed, uint64_t offset, int64_t n_rows, int32_t batch, int64_t* idx_out,
                                   void* stream) {
  IQLB_REQUIRE(idx_out != nullptr, "null idx_out");
  IQLB_REQUIRE(n_rows > 0 && batch > 0, "empty range");
  IQLB_REQUIRE((offset & 3) == 0, "philox offset must be a multiple of 4 (torch generator offsets are)");
  philox_randint_kernel<<<(batch + 255) / 256, 256, 0, (cudaStream_t)stream>>>(seed, offset, (uint64_t)n_rows, batch,
                                                                              idx_out);
  IQLB_LAUNCH_CHECK();
  return IQLB_OK;
}

extern "C" int iqlb_replay_sample(const float* table, int64_t n_rows, int32_t row_stride, int32_t reward_col,
                                  const float* reward_cols, const int32_t* reward_variant, const uint64_t* rng,
                                  int32_t n_agents, int32_t batch, int64_t* idx_out, float* batch_out, void* stream) {
  IQLB_REQUIRE(rng != nullptr, "null rng");
  return launch_gather(table, n_rows, row_stride, reward_col, reward_cols, reward_variant, rng, nullptr, n_agents,
                       batch, idx_out, batch_out, (cudaStream_t)stream);
}

extern "C" int iqlb_replay_gather(const float* table, int64_t n_rows, int32_t row_stride, int32_t reward_col,
                                  const float* reward_cols, const int32_t* reward_variant, const int64_t* idx,
                                  int32_t n_agents, int32_t batch, float* batch_out, void* stream) {
  IQLB_REQUIRE(idx != nullptr, "null idx");
  return launch_gather(table, n_rows, row_stride, reward_col, reward_cols, reward_variant, nullptr, idx, n_agents,
                       batch, nullptr, batch_out, (cudaStream_t)stream);
}

extern "C" int iqlb_polyak(float* target, const float* source, int64_t n, float tau, void* stream) {
  IQLB_REQUIRE(target && source, "null pointer");
  IQLB_REQUIRE(n >= 0, "negative n");
  IQLB_REQUIRE(((uintptr_t)target & 15) == 0 && ((uintptr_t)source & 15) == 0, "16-byte alignment");
  if (n == 0) return IQLB_OK;
  const int64_t n4 = n / 4;
  const int tail = (int)(n - n4 * 4);
  int64_t blocks = (n4 + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (blocks < 1) blocks = 1;
  polyak_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((float4*)target, (const float4*)source, n4,
                                                                    target + n4 * 4, source + n4 * 4, tail, tau);
  IQLB_LAUNCH_CHECK();
  return IQLB_OK;
}
