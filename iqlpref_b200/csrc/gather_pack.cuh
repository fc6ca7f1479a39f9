// ReplayBuffer.sample (iql.py:211-221) fused with the packing of the tensor-core input images: the body of
// replay_gather_pack_kernel (bf16_step.cu).
#pragma once
#include "iql_step.cuh"
#include "umma.cuh"

namespace iqlb {

struct GatherArgs {
  const float4* table;
  uint64_t n_rows;
  int vec_per_row, rows_per_cta, reward_col;
  const float* reward_cols;
  const int32_t* reward_variant;
  const uint64_t* rng;
  int64_t* idx_out;
  float4* batch_out;
  int blocks_x;  // gather CTAs per agent
};

// Block (bx, agent index by) of 256 threads: Philox indices, TMA staging of the rows (one cp.async.bulk per 16-byte aligned
// row of obs | action | next_obs | reward | done into shared memory, completion on an mbarrier), fp32 batch rows (the loss kernel
// reads reward / done / action from them) and both bf16 input images in one pass.  srow: rows_per_cta * R floats.
// call: number of earlier calls by this CTA (persistent callers loop over blocks; the mbarrier is initialised by call 0 and
// waited on with the call's phase parity; the caller separates calls with __syncthreads()).
// agent = first agent of the launch + by (the caller passes it: writing P.agent0 in a kernel would copy the whole parameter
// struct to local memory).
__device__ __forceinline__ void gather_pack_body(const BfParams& P, const GatherArgs& G, int bx, int agent, float* srow, int call = 0) {
  __shared__ __align__(8) uint64_t gather_bar;
  __shared__ int64_t gather_row[32];
  const Layout& L = P.L;
  const int t = threadIdx.x;
  const int vec_per_row = G.vec_per_row, rows_per_cta = G.rows_per_cta;
  const int b0 = bx * rows_per_cta;
  const int n_here = L.B - b0 < rows_per_cta ? L.B - b0 : rows_per_cta;
  const uint32_t bar = umma::smem_u32(&gather_bar);
  const bool tma = rows_per_cta <= 32;  // one issuing lane per row (rows of >= 32 floats: every reference shape)
  if (tma) {
    if (call == 0) {
      if (t == 0) { umma::mbar_init(bar, 1); umma::fence_barrier_init(); }
      __syncthreads();
    }
    if (t < 32) {
      if (t < n_here) {
        uint32_t r[4];
        philox4(G.rng[2 * agent], (uint64_t)(b0 + t), G.rng[2 * agent + 1], r);
        const int64_t row = randint_from_philox(r, G.n_rows);
        gather_row[t] = row;
        if (G.idx_out != nullptr) G.idx_out[(int64_t)agent * L.B + b0 + t] = row;
        umma::bulk_g2s(umma::smem_u32(srow + t * L.R), G.table + row * vec_per_row, (uint32_t)L.R * 4u, bar);
      }
      if (t == 0) umma::mbar_expect_tx(bar, (uint32_t)n_here * (uint32_t)L.R * 4u);
    }
    __syncthreads();  // gather_row visible
    umma::mbar_wait(bar, (uint32_t)(call & 1));
  }
  const int lb = t / vec_per_row, v = t - lb * vec_per_row;
  const int b = b0 + lb;
  if (lb < rows_per_cta && b < L.B) {
    int64_t row;
    float4 val;
    if (tma) {
      row = gather_row[lb];
      val = *reinterpret_cast<const float4*>(srow + lb * L.R + 4 * v);
    } else {
      uint32_t r[4];
      philox4(G.rng[2 * agent], (uint64_t)b, G.rng[2 * agent + 1], r);
      row = randint_from_philox(r, G.n_rows);
      if (G.idx_out != nullptr && v == 0) G.idx_out[(int64_t)agent * L.B + b] = row;
      val = __ldg(G.table + row * vec_per_row + v);
    }
    if (G.reward_cols != nullptr && (G.reward_col >> 2) == v) {
      const int variant = G.reward_variant ? G.reward_variant[agent] : 0;
      const float rw = __ldg(G.reward_cols + (int64_t)variant * (int64_t)G.n_rows + row);
      switch (G.reward_col & 3) {
        case 0: val.x = rw; break;
        case 1: val.y = rw; break;
        case 2: val.z = rw; break;
        default: val.w = rw; break;
      }
    }
    G.batch_out[((int64_t)agent * L.B + b) * vec_per_row + v] = val;
    *reinterpret_cast<float4*>(srow + lb * L.R + 4 * v) = val;
  }
  __syncthreads();
  const int cpr = P.KXB * 8;
  for (int i = t; i < rows_per_cta * cpr; i += 256) {
    const int rb = i / cpr, ch = i - rb * cpr;
    const int bb = bx * rows_per_cta + rb;
    if (bb >= L.B) break;
    const float* row = srow + rb * L.R;
    uint32_t sa[4], nx[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int c = ch * 8 + 2 * q;
      const float a0 = c < L.Dq ? row[c] : 0.f, a1 = c + 1 < L.Dq ? row[c + 1] : 0.f;
      const float n0 = c < L.S ? row[L.Dq + c] : 0.f, n1 = c + 1 < L.S ? row[L.Dq + c + 1] : 0.f;
      sa[q] = umma::pack_bf16(a0, a1);
      nx[q] = umma::pack_bf16(n0, n1);
    }
    const int64_t off = umma::image_off(L.B, bb, ch * 8);
    *reinterpret_cast<uint4*>(P.xsa + (int64_t)agent * P.x_bytes + off) = make_uint4(sa[0], sa[1], sa[2], sa[3]);
    *reinterpret_cast<uint4*>(P.xnext + (int64_t)agent * P.x_bytes + off) = make_uint4(nx[0], nx[1], nx[2], nx[3]);
  }
}

}  // namespace iqlb
