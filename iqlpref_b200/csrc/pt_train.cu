// Preference-transformer training step (the PT half):
//   PTTrainer.train / .evaluation   gp_reward-priors/optbnn/training/training.py:308-416
//   PT.forward in training mode     gp_reward-priors/optbnn/bnn/nets/pref_trans.py:269-325 (GPT2Model :186-227,
//                                   GPT2Block :147-183, GPT2SelfAttention :91-144, GPT2MLP :69-88)
//   bt_pool_logit                   gp_reward-priors/optbnn/utils/util.py:356-364
//   optimizer                       torch.optim.Adam(net.parameters(), lr) (scripts_pt/run_pt_training.py:137)
// One call = one forward over the 2B sequences of both segments (activations kept), Bradley-Terry pooling of
// `weighted_sum`, soft-label cross entropy, the backward pass through the preference-attention head, the final
// LayerNorm, every GPT2 block (MLP, causal attention, two LayerNorms, residual dropouts), the embedding LayerNorm and the
// three embeddings, and Adam - ordered on the caller's stream, no host synchronisation, deterministic (every reduction
// over rows is a fixed-order sum of per-chunk partials).
//
// fp32 throughout (the reference trains in fp32).  Every Linear - forward, data gradient, weight gradient - is a GEMM
// through launch_gemm / the grouped weight-gradient launchers: fp32 FFMA, or with precision 1 split-bf16 on the tensor
// cores (gemm_tc.cu).  So are the attention products (Q K^T, P V and their four gradients per (sequence, head), the preference
// head's per sequence): grouped launches over task tables built once per handle.  LayerNorm / softmax / dropout are
// warp-per-row kernels.  The attention probabilities of the forward pass are kept ([n, H, 2T, 2T]).
//
// Dropout (embd / attention / the two residual sites of a block): the reference draws its masks from torch's CUDA generator,
// which cannot be mirrored; element e (row-major index in the tensor the mask multiplies) of site s at Adam step k is KEPT iff
//   word (e & 3) of Philox4x32-10(key = seed, counter = {k, 0, lo32(e >> 2), (s << 8) | hi(e >> 2)}) >= p * 2^32
// and kept values are scaled by 1 / (1 - p)   (include/iqlb.h; parity by injecting these masks into the oracle).
#include <vector>

#include "pt.cuh"

namespace iqlb {

struct PtDrop {
  uint64_t seed, ctr;  // ctr = 4 * adam_step (philox4's offset argument)
  uint32_t thresh;     // 0: no dropout at this site
  float scale;
  int site;
};

__device__ __forceinline__ float drop_factor(const PtDrop& d, int64_t e) {
  if (d.thresh == 0) return 1.f;
  uint32_t r[4];
  philox4(d.seed, ((uint64_t)d.site << 40) | (uint64_t)(e >> 2), d.ctr, r);
  return r[e & 3] >= d.thresh ? d.scale : 0.f;
}

// ------------------------------------------------------------------------------------------------
// LayerNorm forward, one warp per row: y = LN(x (+ temb[ts])) * g + b (* dropout); keeps xhat and 1/std for the backward pass
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ptt_ln_fwd_kernel(int E, float eps, const float* __restrict__ g,
                                                         const float* __restrict__ b, const float* x, int64_t in_stride,
                                                         int64_t n_rows, const int32_t* __restrict__ ts_idx,
                                                         const float* __restrict__ temb, float* y, float* __restrict__ xhat,
                                                         float* __restrict__ rstd_out, PtDrop drop) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t r = (int64_t)blockIdx.x * 8 + warp;
  if (r >= n_rows) return;
  const int per = E / 32;
  float v[kPtMaxPerLane];
  const float* xr = x + r * in_stride;
  const float* te = temb ? temb + (int64_t)ts_idx[r >> 1] * E : nullptr;
#pragma unroll
  for (int c = 0; c < kPtMaxPerLane; ++c)
    if (c < per) v[c] = xr[lane + 32 * c] + (te ? __ldg(te + lane + 32 * c) : 0.f);
  float s = 0.f;
#pragma unroll
  for (int c = 0; c < kPtMaxPerLane; ++c) if (c < per) s += v[c];
  const float mean = warp_sum(s) / (float)E;
  float q = 0.f;
#pragma unroll
  for (int c = 0; c < kPtMaxPerLane; ++c) if (c < per) { const float d = v[c] - mean; q += d * d; }
  const float rstd = rsqrtf(warp_sum(q) / (float)E + eps);
#pragma unroll
  for (int c = 0; c < kPtMaxPerLane; ++c) {
    if (c < per) {
      const int col = lane + 32 * c;
      const float xh = (v[c] - mean) * rstd;
      float yv = xh * __ldg(g + col) + __ldg(b + col);
      yv *= drop_factor(drop, r * E + col);
      y[r * E + col] = yv;
      if (xhat) xhat[r * E + col] = xh;
    }
  }
  if (rstd_out && lane == 0) rstd_out[r] = rstd;
}

// LayerNorm backward over a block of rows.  dy (optionally through a dropout mask) -> dx (written or accumulated);
// partial[block][0:E] = sum_rows dy * xhat (d gain), partial[block][E:2E] = sum_rows dy (d bias).
// (PM = upper bound of E / 32 the register arrays are sized for: at PM = 16 the kernel needs 162 registers)
template <int PM>
__global__ void __launch_bounds__(256) ptt_ln_bwd_kernel(int E, const float* __restrict__ g, const float* dy,
                                                         int64_t dy_stride, PtDrop drop, const float* __restrict__ xhat,
                                                         const float* __restrict__ rstd, int64_t n_rows, int rows_per_block,
                                                         float* dx, int64_t dx_stride, int accumulate,
                                                         float* __restrict__ partial) {
  extern __shared__ float red[];  // [8][2E]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int per = E / 32;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
  const int64_t r1 = r0 + rows_per_block < n_rows ? r0 + rows_per_block : n_rows;
  float ag[PM], ab[PM], gv[PM];
#pragma unroll
  for (int c = 0; c < PM; ++c) { ag[c] = 0.f; ab[c] = 0.f; gv[c] = c < per ? __ldg(g + lane + 32 * c) : 0.f; }
  const float invE = 1.0f / (float)E;
  for (int64_t r = r0 + warp; r < r1; r += 8) {
    float d[PM], xh[PM];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int c = 0; c < PM; ++c) {
      if (c < per) {
        const int col = lane + 32 * c;
        d[c] = dy[r * dy_stride + col] * drop_factor(drop, r * E + col);
        xh[c] = xhat[r * E + col];
        ag[c] = fmaf(d[c], xh[c], ag[c]);
        ab[c] += d[c];
        d[c] *= gv[c];
        s1 += d[c];
        s2 = fmaf(d[c], xh[c], s2);
      }
    }
    s1 = warp_sum(s1) * invE;
    s2 = warp_sum(s2) * invE;
    const float rs = rstd[r];
#pragma unroll
    for (int c = 0; c < PM; ++c) {
      if (c < per) {
        const int col = lane + 32 * c;
        const float o = rs * (d[c] - s1 - xh[c] * s2);
        float* dst = dx + r * dx_stride + col;
        *dst = accumulate ? *dst + o : o;
      }
    }
  }
#pragma unroll
  for (int c = 0; c < PM; ++c) {
    if (c < per) {
      red[warp * 2 * E + lane + 32 * c] = ag[c];
      red[warp * 2 * E + E + lane + 32 * c] = ab[c];
    }
  }
  __syncthreads();
  for (int j = threadIdx.x; j < 2 * E; j += blockDim.x) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += red[w * 2 * E + j];
    partial[(int64_t)blockIdx.x * 2 * E + j] = s;
  }
}

// ------------------------------------------------------------------------------------------------
// Causal self-attention (pref_trans.py:118-140).  The products are batched GEMMs over the (sequence, head) pairs
// (S = Q K^T, O = Pd V; backward dPd = dO V^T, dQ = dS K, dK = dS^T Q, dV = Pd^T dO - grouped launches, on the tensor cores
// in split-bf16 mode); the kernels below are the row-wise softmax between them.  Score buffers are [n][H][2T][2T].
// Forward: P[i][j] = softmax_{j <= i}(S[i][j] / sqrt(dh) + (1 - m[j]) * -10000), 0 above the diagonal (kept for the
// backward pass); Pd = P * dropout mask (a second buffer when dropout is on, else Pd aliases P).  One warp per row.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ptt_softmax_fwd_kernel(int T2, int H, float scale, float* S, float* Pd,
                                                              const float* __restrict__ maskbuf, int64_t n_rows, PtDrop drop) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t r = (int64_t)blockIdx.x * 8 + warp;
  if (r >= n_rows) return;
  const int i = (int)(r % T2);
  const int64_t w = r / T2 / H;
  const float* mk = maskbuf + w * (T2 / 2);
  float* srow = S + r * T2;
  const int nchunk = (T2 + 31) / 32;
  float s[kPtMaxKeyChunks];
  float mx = -INFINITY;
#pragma unroll
  for (int c = 0; c < kPtMaxKeyChunks; ++c) {
    if (c < nchunk) {
      const int j = lane + 32 * c;
      float v = -INFINITY;
      if (j <= i) v = srow[j] * scale + (1.0f - mk[j >> 1]) * -10000.0f;
      s[c] = v;
      mx = fmaxf(mx, v);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  float sum = 0.f;
#pragma unroll
  for (int c = 0; c < kPtMaxKeyChunks; ++c) {
    if (c < nchunk) {
      const float e = (s[c] == -INFINITY) ? 0.f : expf(s[c] - mx);
      s[c] = e;
      sum += e;
    }
  }
  const float inv = 1.0f / warp_sum(sum);
#pragma unroll
  for (int c = 0; c < kPtMaxKeyChunks; ++c) {
    const int j = lane + 32 * c;
    if (c < nchunk && j < T2) {
      const float pv = s[c] * inv;  // 0 above the diagonal
      srow[j] = pv;
      if (Pd != S) Pd[r * T2 + j] = pv * drop_factor(drop, r * T2 + j);
    }
  }
}

// Backward: dS (in: dPd = dO V^T) <- scale * P * (dP - sum_j P dP) with dP = dPd * dropout mask; 0 above the diagonal.
__global__ void __launch_bounds__(256) ptt_softmax_bwd_kernel(int T2, float scale, const float* __restrict__ P, float* dS,
                                                              int64_t n_rows, PtDrop drop) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t r = (int64_t)blockIdx.x * 8 + warp;
  if (r >= n_rows) return;
  const int i = (int)(r % T2);
  const int nchunk = (T2 + 31) / 32;
  float dp[kPtMaxKeyChunks], pv[kPtMaxKeyChunks];
  float rowsum = 0.f;
#pragma unroll
  for (int c = 0; c < kPtMaxKeyChunks; ++c) {
    dp[c] = 0.f; pv[c] = 0.f;
    const int j = lane + 32 * c;
    if (c < nchunk && j <= i) {
      const int64_t e = r * T2 + j;
      pv[c] = P[e];
      dp[c] = dS[e] * drop_factor(drop, e);
      rowsum = fmaf(pv[c], dp[c], rowsum);
    }
  }
  rowsum = warp_sum(rowsum);
#pragma unroll
  for (int c = 0; c < kPtMaxKeyChunks; ++c) {
    const int j = lane + 32 * c;
    if (c < nchunk && j < T2) dS[r * T2 + j] = j <= i ? scale * pv[c] * (dp[c] - rowsum) : 0.f;
  }
}

// ------------------------------------------------------------------------------------------------
// Preference-attention head (pref_trans.py:303-320): y rows [q(P) | k(P) | v], full attention, key-padding bias only; the
// products Q K^T, dS K and dS^T Q are grouped GEMMs per sequence.  Forward (one warp per query): W[t][j] = softmax_j(S / sqrt(P)
// + (1 - m[j]) * -10000) (in place, kept), ws[t] = sum_j W[t][j] v[j].  Backward (one block per sequence), with g = dL/dws:
// dS[t][j] = scale * W[t][j] g[t] (v[j] - ws[t]);  dv[j] = sum_t W[t][j] g[t] (written into dy).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ptt_pref_softmax_fwd_kernel(int T, int Pd, float scale, float* S,
                                                                   const float* __restrict__ y,
                                                                   const float* __restrict__ maskbuf, int64_t n_rows,
                                                                   float* __restrict__ ws) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t r = (int64_t)blockIdx.x * 8 + warp;
  if (r >= n_rows) return;
  const int64_t n = r / T;
  const int ld = 2 * Pd + 1;
  const float* mk = maskbuf + n * T;
  const float* vb = y + n * T * ld + 2 * Pd;
  float* srow = S + r * T;
  const int nchunk = (T + 31) / 32;
  float s[kPtMaxKeyChunks];
  float mx = -INFINITY;
#pragma unroll
  for (int c = 0; c < kPtMaxKeyChunks; ++c) {
    const int j = lane + 32 * c;
    if (c < nchunk) {
      s[c] = j < T ? srow[j] * scale + (1.0f - mk[j]) * -10000.0f : -INFINITY;
      mx = fmaxf(mx, s[c]);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  float sum = 0.f, wsv = 0.f;
#pragma unroll
  for (int c = 0; c < kPtMaxKeyChunks; ++c) {
    const int j = lane + 32 * c;
    if (c < nchunk && j < T) {
      const float e = expf(s[c] - mx);
      s[c] = e;
      sum += e;
      wsv = fmaf(e, vb[(int64_t)j * ld], wsv);
    }
  }
  sum = warp_sum(sum);
  wsv = warp_sum(wsv);
  const float inv = 1.0f / sum;
#pragma unroll
  for (int c = 0; c < kPtMaxKeyChunks; ++c) {
    const int j = lane + 32 * c;
    if (c < nchunk && j < T) srow[j] = s[c] * inv;
  }
  if (lane == 0) ws[r] = wsv * inv;
}

__global__ void __launch_bounds__(256) ptt_pref_softmax_bwd_kernel(int T, int Pd, float scale, const float* __restrict__ W,
                                                                   const float* __restrict__ y, const float* __restrict__ ws,
                                                                   const float* __restrict__ g, float* __restrict__ dS,
                                                                   float* __restrict__ dy) {
  const int n = blockIdx.x;
  const int ld = 2 * Pd + 1;
  const float* wb = W + (int64_t)n * T * T;
  float* db = dS + (int64_t)n * T * T;
  const float* vb = y + (int64_t)n * T * ld + 2 * Pd;
  const float* gn = g + (int64_t)n * T;
  const float* wsn = ws + (int64_t)n * T;
  for (int e = threadIdx.x; e < T * T; e += blockDim.x) {
    const int t = e / T, j = e - t * T;
    db[e] = scale * wb[e] * gn[t] * (vb[(int64_t)j * ld] - wsn[t]);
  }
  for (int j = threadIdx.x; j < T; j += blockDim.x) {  // column sums, coalesced over j
    float dv = 0.f;
    for (int t = 0; t < T; ++t) dv = fmaf(wb[(int64_t)t * T + j], gn[t], dv);
    dy[((int64_t)n * T + j) * ld + 2 * Pd] = dv;
  }
}

// x[e] += a[e] * dropout factor (the residual connections of a block, pref_trans.py:86, 142, 176-182)
__global__ void ptt_resid_kernel(float* __restrict__ x, const float* __restrict__ a, int64_t total, PtDrop drop) {
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x)
    x[e] = fmaf(a[e], drop_factor(drop, e), x[e]);
}

__global__ void ptt_mask_kernel(const float* __restrict__ src, float* __restrict__ dst, int64_t total, PtDrop drop) {
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x)
    dst[e] = src[e] * drop_factor(drop, e);
}

__global__ void ptt_relu_mask_kernel(float* __restrict__ d, const float* __restrict__ z, int64_t total) {
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x)
    if (!(z[e] > 0.f)) d[e] = 0.f;
}

// Gradient of timestep_embed (pref_trans.py:275-278: both tokens of a step add row ts of the table).  Deterministic
// scatter-add in two levels: block (v, c) sums du[2 i] + du[2 i + 1] over the positions i of chunk c with ts[i] == v, in index
// order, into partial[c][v]; the Adam kernel adds the chunks in a fixed order.  (One block per table row walking all
// positions took 1.2 ms: the padded positions all carry ts = 0, thousands of dependent row loads in one block.)
constexpr int kTembChunk = 1024;
__global__ void __launch_bounds__(256) ptt_temb_grad_kernel(int E, const int32_t* __restrict__ ts, int64_t n_pos,
                                                            const float* __restrict__ du, float* __restrict__ partial,
                                                            int64_t plen) {
  __shared__ int list[kTembChunk];
  __shared__ int wcount[8];
  __shared__ int n_list;
  const int v = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t p0 = (int64_t)blockIdx.y * kTembChunk;
  const int64_t p1 = p0 + kTembChunk < n_pos ? p0 + kTembChunk : n_pos;
  if (threadIdx.x == 0) n_list = 0;
  __syncthreads();
  for (int64_t base = p0; base < p1; base += 256) {  // ordered compaction of the matching positions
    const int64_t idx = base + threadIdx.x;
    const bool match = idx < p1 && ts[idx] == v;
    const unsigned bal = __ballot_sync(0xffffffffu, match);
    if (lane == 0) wcount[warp] = __popc(bal);
    __syncthreads();
    int off = n_list;
    for (int w = 0; w < warp; ++w) off += wcount[w];
    if (match) list[off + __popc(bal & ((1u << lane) - 1u))] = (int)(idx - p0);
    __syncthreads();
    if (threadIdx.x == 0) { int t = 0; for (int w = 0; w < 8; ++w) t += wcount[w]; n_list += t; }
    __syncthreads();
  }
  const int nl = n_list;
  float acc[2] = {0.f, 0.f};  // E <= 512: columns threadIdx.x, threadIdx.x + 256
  for (int k = 0; k < nl; ++k) {
    const float* r = du + (p0 + list[k]) * 2 * E;
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const int col = threadIdx.x + 256 * c;
      if (col < E) acc[c] += r[col] + r[E + col];
    }
  }
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    const int col = threadIdx.x + 256 * c;
    if (col < E) partial[(int64_t)blockIdx.y * plen + (int64_t)v * E + col] = acc[c];
  }
}

}  // namespace iqlb

using namespace iqlb;

struct iqlb_pt_trainer {
  iqlb_pt_desc d{};
  PtOff P{};
  int pool_mean = 1, tc = 0;
  int n_max = 0;  // sequences (2 * pairs) the buffers hold
  float p_embd = 0.f, p_attn = 0.f, p_ra = 0.f, p_rm = 0.f;
  uint64_t seed = 0;
  // inputs of both segments, concatenated
  float *St = nullptr, *Ac = nullptr, *Mk = nullptr;
  int32_t* Ts = nullptr;
  // forward state
  float *x = nullptr, *xhat_s = nullptr, *rstd_s = nullptr;
  struct Layer { float *h0, *xhat0, *rstd0, *qkv, *Pb, *Pd, *o, *h1, *xhat1, *rstd1, *z; };  // Pb: softmax, Pd: after dropout
  std::vector<Layer> layers;
  float *hid = nullptr, *xhat_f = nullptr, *rstd_f = nullptr, *y = nullptr, *ws = nullptr, *val = nullptr;
  // backward scratch
  float *rowgrad = nullptr, *Sp = nullptr, *dy = nullptr, *dx = nullptr, *tmpA = nullptr, *tmpB = nullptr, *dbig = nullptr;
  float *dS = nullptr, *dSp = nullptr, *partial = nullptr, *zero_bias = nullptr;
  float *out2 = nullptr, *pair_out = nullptr;
  unsigned int* counter = nullptr;
  GemmTask* d_tasks = nullptr;  // [n_groups][max_chunks] weight-gradient tasks
  // Batched attention GEMMs, built once for n_max sequences (a launch over the first n * H tasks serves any batch size).
  // Per layer kAttnArrays arrays of n_max * H tasks: 0 S = Q K^T | 1 O = Pd V | 2 O = P V (no attention dropout) |
  // 3 dPd = dO V^T | 4 dQ = dS K | 5 dK = dS^T Q | 6 dV = Pd^T dO | 7 dV = P^T dO; then three arrays of n_max tasks for the
  // preference head: S = Q K^T | dQ = dS K | dK = dS^T Q.
  GemmTask* d_attn = nullptr;
  static constexpr int kAttnArrays = 8;
  GemmTask* attn_tasks(int layer, int which) const { return d_attn + ((size_t)layer * kAttnArrays + which) * n_max * P.H; }
  GemmTask* pref_tasks(int which) const { return d_attn + (size_t)P.L * kAttnArrays * n_max * P.H + (size_t)which * n_max; }
  int n_groups = 0, max_chunks = 0;
  int64_t max_plen = 0;
  int tasks_n = -1;  // sequence count the task table was built for
  std::vector<void*> allocs;
};

static void ptt_free(iqlb_pt_trainer* h) {
  if (!h) return;
  for (void* p : h->allocs) cudaFree(p);
  delete h;
}

static constexpr int kPttMaxChunks = 296;
static constexpr int kPttChunkRows = 512;
static constexpr int kLnRowsPerBlock = 32;

static inline int ptt_chunks(int64_t rows, int& chunk_rows) {
  int n_chunks = (int)((rows + kPttChunkRows - 1) / kPttChunkRows);
  if (n_chunks > kPttMaxChunks) n_chunks = kPttMaxChunks;
  if (n_chunks < 1) n_chunks = 1;
  chunk_rows = (int)((rows + n_chunks - 1) / n_chunks);
  chunk_rows = (chunk_rows + 7) / 8 * 8;
  return (int)((rows + chunk_rows - 1) / chunk_rows);
}

extern "C" int iqlb_pt_train_param_count(const iqlb_pt_desc* d, int64_t* n_floats) {
  int rc = pt_check(d);
  if (rc) return rc;
  IQLB_REQUIRE(n_floats != nullptr, "null output");
  IQLB_REQUIRE(d->pref_dim >= 1, "pref_dim");
  PtOff P;
  P.init(*d, true);
  *n_floats = P.total;
  return IQLB_OK;
}

extern "C" int iqlb_pt_trainer_create(const iqlb_pt_desc* d, int32_t pool_mean, int32_t max_pairs, iqlb_pt_trainer** out) {
  IQLB_REQUIRE(out != nullptr, "null out");
  int rc = pt_check(d);
  if (rc) return rc;
  IQLB_REQUIRE(d->pref_dim >= 1, "pref_dim");
  IQLB_REQUIRE(max_pairs >= 1, "max_pairs");
  auto* h = new iqlb_pt_trainer();
  h->d = *d;
  h->P.init(*d, true);
  h->pool_mean = pool_mean ? 1 : 0;
  h->n_max = 2 * max_pairs;
  const PtOff& P = h->P;
  const int E = P.E, T = P.T, T2 = 2 * T, I = P.I, H = P.H, L = P.L, ldy = 2 * P.Pd + 1;
  const int64_t n = h->n_max, M = n * T2, nT = n * T;
  const int64_t bigc = 3 * E > I ? 3 * E : I;
  bool ok = true;
  auto alloc = [&](int64_t floats) -> float* {
    void* p = nullptr;
    if (!ok) return nullptr;
    if (cudaMalloc(&p, (size_t)(floats > 0 ? floats : 1) * sizeof(float)) != cudaSuccess) { cudaGetLastError(); ok = false; return nullptr; }
    h->allocs.push_back(p);
    return (float*)p;
  };
  h->St = alloc(nT * P.S); h->Ac = alloc(nT * P.A); h->Mk = alloc(nT); h->Ts = (int32_t*)alloc(nT);
  h->x = alloc(M * E); h->xhat_s = alloc(M * E); h->rstd_s = alloc(M);
  h->layers.resize(L);
  for (int l = 0; l < L; ++l) {
    auto& y = h->layers[l];
    y.h0 = alloc(M * E); y.xhat0 = alloc(M * E); y.rstd0 = alloc(M); y.qkv = alloc(M * 3 * E);
    y.Pb = alloc(n * H * T2 * T2); y.Pd = alloc(n * H * T2 * T2); y.o = alloc(M * E); y.h1 = alloc(M * E); y.xhat1 = alloc(M * E); y.rstd1 = alloc(M);
    y.z = alloc(M * I);
  }
  h->hid = alloc(nT * E); h->xhat_f = alloc(nT * E); h->rstd_f = alloc(nT); h->y = alloc(nT * ldy); h->ws = alloc(nT);
  h->val = alloc(nT);
  h->rowgrad = alloc(nT); h->Sp = alloc(nT * T); h->dy = alloc(nT * ldy); h->dx = alloc(M * E); h->tmpA = alloc(M * E);
  h->tmpB = alloc(M * E); h->dbig = alloc(M * bigc); h->dS = alloc(n * H * T2 * T2); h->dSp = alloc(nT * T);
  h->zero_bias = alloc(bigc > ldy ? bigc : ldy);
  h->out2 = alloc(2); h->pair_out = alloc(2 * n); h->counter = (unsigned int*)alloc(1);
  // weight-gradient partials: the largest Linear (+ bias), or the LayerNorm partials [M / 64][2E]
  int64_t plen = (int64_t)I * E + I;
  const int64_t p_in = (int64_t)3 * E * E + 3 * E, p_pref = (int64_t)ldy * E + ldy;
  if (p_in > plen) plen = p_in;
  if (p_pref > plen) plen = p_pref;
  h->max_plen = plen;
  h->max_chunks = kPttMaxChunks;
  int cr = 0;
  const int64_t need_w = (int64_t)ptt_chunks(M, cr) * plen;
  const int64_t need_ln = ((M + kLnRowsPerBlock - 1) / kLnRowsPerBlock) * 2 * E;
  const int64_t need_te = ((nT + kTembChunk - 1) / kTembChunk) * P.n_ts * E;
  int64_t need = need_w > need_ln ? need_w : need_ln;
  if (need_te > need) need = need_te;
  h->partial = alloc(need);
  h->n_groups = 3 + 4 * L;
  if (ok) {
    void* p = nullptr;
    if (cudaMalloc(&p, (size_t)h->n_groups * h->max_chunks * sizeof(GemmTask)) != cudaSuccess) { cudaGetLastError(); ok = false; }
    else { h->allocs.push_back(p); h->d_tasks = (GemmTask*)p; }
  }
  const size_t n_attn = (size_t)L * iqlb_pt_trainer::kAttnArrays * n * H + 3 * (size_t)n;
  if (ok) {
    void* p = nullptr;
    if (cudaMalloc(&p, n_attn * sizeof(GemmTask)) != cudaSuccess) { cudaGetLastError(); ok = false; }
    else { h->allocs.push_back(p); h->d_attn = (GemmTask*)p; }
  }
  if (!ok) {
    ptt_free(h);
    return fail(IQLB_ERR_NOMEM, "iqlb_pt_trainer_create: device allocation failed%s", "");
  }
  {
    std::vector<GemmTask> tk(n_attn);
    const int dh = E / H;
    auto task = [](const float* A, int lda, const float* B, int ldb, float* C, int ldc, int Mm, int Nn, int Kk, const float* bias) {
      GemmTask t{};
      t.A = A; t.lda = lda; t.Bm = B; t.ldb = ldb; t.C = C; t.ldc = ldc; t.M = Mm; t.N = Nn; t.K = Kk; t.bias = bias;
      t.scale = 1.f; t.act = IQLB_ACT_LINEAR; t.allow_tc = 1;
      return t;
    };
    for (int l = 0; l < L; ++l) {
      auto& y = h->layers[l];
      for (int64_t w = 0; w < n; ++w) {
        for (int hh = 0; hh < H; ++hh) {
          const size_t i = (size_t)w * H + hh;
          const float* q = y.qkv + w * T2 * 3 * E + hh * dh;
          const float *k = q + E, *v = q + 2 * E;
          float* Pb = y.Pb + (int64_t)i * T2 * T2;
          float* Pdr = y.Pd + (int64_t)i * T2 * T2;
          float* dS = h->dS + (int64_t)i * T2 * T2;
          float* o = y.o + w * T2 * E + hh * dh;
          const float* dO = h->tmpB + w * T2 * E + hh * dh;
          float* dq = h->dbig + w * T2 * 3 * E + hh * dh;
          const size_t base = (size_t)l * iqlb_pt_trainer::kAttnArrays * n * H;
          auto at = [&](int which) -> GemmTask& { return tk[base + (size_t)which * n * H + i]; };
          at(0) = task(q, 3 * E, k, 3 * E, Pb, T2, T2, T2, dh, nullptr);             // LINEAR: S = Q K^T
          at(1) = task(Pdr, T2, v, 3 * E, o, E, T2, dh, T2, h->zero_bias);           // FWD_KN: O = Pd V
          at(2) = task(Pb, T2, v, 3 * E, o, E, T2, dh, T2, h->zero_bias);            //         O = P V
          at(3) = task(dO, E, v, 3 * E, dS, T2, T2, T2, dh, nullptr);                // LINEAR: dPd = dO V^T
          at(4) = task(dS, T2, k, 3 * E, dq, 3 * E, T2, dh, T2, h->zero_bias);       // FWD_KN: dQ = dS K
          at(5) = task(dS, T2, q, 3 * E, dq + E, 3 * E, T2, dh, T2, nullptr);        // BWD_WEIGHT: dK = dS^T Q
          at(6) = task(Pdr, T2, dO, E, dq + 2 * E, 3 * E, T2, dh, T2, nullptr);      //             dV = Pd^T dO
          at(7) = task(Pb, T2, dO, E, dq + 2 * E, 3 * E, T2, dh, T2, nullptr);       //             dV = P^T dO
        }
      }
    }
    const size_t pbase = (size_t)L * iqlb_pt_trainer::kAttnArrays * n * H;
    for (int64_t w = 0; w < n; ++w) {
      const float* yq = h->y + w * T * ldy;
      float* dyq = h->dy + w * T * ldy;
      tk[pbase + w] = task(yq, ldy, yq + P.Pd, ldy, h->Sp + w * T * T, T, T, T, P.Pd, nullptr);                    // S = Q K^T
      tk[pbase + n + w] = task(h->dSp + w * T * T, T, yq + P.Pd, ldy, dyq, ldy, T, P.Pd, T, h->zero_bias);        // dQ = dS K
      tk[pbase + 2 * n + w] = task(h->dSp + w * T * T, T, yq, ldy, dyq + P.Pd, ldy, T, P.Pd, T, nullptr);         // dK = dS^T Q
    }
    if (cudaMemcpy(h->d_attn, tk.data(), n_attn * sizeof(GemmTask), cudaMemcpyHostToDevice) != cudaSuccess) {
      cudaGetLastError(); ptt_free(h);
      return fail(IQLB_ERR_CUDA, "iqlb_pt_trainer_create: task table upload failed%s", "");
    }
  }
  cudaMemset(h->counter, 0, sizeof(unsigned int));
  cudaMemset(h->zero_bias, 0, (size_t)(bigc > ldy ? bigc : ldy) * sizeof(float));
  *out = h;
  return IQLB_OK;
}

extern "C" int iqlb_pt_trainer_destroy(iqlb_pt_trainer* h) {
  if (h) cudaDeviceSynchronize();
  ptt_free(h);
  return IQLB_OK;
}

extern "C" int iqlb_pt_trainer_set_precision(iqlb_pt_trainer* h, int32_t precision) {
  IQLB_REQUIRE(h != nullptr, "null handle");
  IQLB_REQUIRE(precision == 0 || precision == 1, "precision must be 0 (fp32 FFMA) or 1 (split-bf16 tcgen05)");
  h->tc = precision;
  return IQLB_OK;
}

extern "C" int iqlb_pt_trainer_set_dropout(iqlb_pt_trainer* h, float p_embd, float p_attn, float p_resid_attn,
                                           float p_resid_mlp, uint64_t seed) {
  IQLB_REQUIRE(h != nullptr, "null handle");
  auto okp = [](float p) { return p >= 0.f && p < 1.f; };
  IQLB_REQUIRE(okp(p_embd) && okp(p_attn) && okp(p_resid_attn) && okp(p_resid_mlp), "dropout probabilities in [0, 1)");
  h->p_embd = p_embd; h->p_attn = p_attn; h->p_ra = p_resid_attn; h->p_rm = p_resid_mlp; h->seed = seed;
  return IQLB_OK;
}

static PtDrop make_drop(const iqlb_pt_trainer* h, float p, int site, int64_t adam_step, bool training) {
  PtDrop d{};
  d.seed = h->seed; d.ctr = (uint64_t)adam_step * 4ull; d.site = site; d.thresh = 0; d.scale = 1.f;
  if (training && p > 0.f) {
    const double t = (double)p * 4294967296.0;
    d.thresh = t >= 4294967295.0 ? 4294967295u : (uint32_t)t;
    if (d.thresh == 0) d.thresh = 1;
    d.scale = 1.0f / (1.0f - p);
  }
  return d;
}

static int ptt_check_batch(const iqlb_pt_trainer* h, const void* params, const void* s1, const void* a1, const void* t1,
                           const void* m1, const void* s2, const void* a2, const void* t2, const void* m2,
                           const void* labels, int B) {
  IQLB_REQUIRE(h != nullptr, "null trainer");
  IQLB_REQUIRE(params && s1 && a1 && t1 && m1 && s2 && a2 && t2 && m2 && labels, "null pointer");
  IQLB_REQUIRE(B > 0 && 2 * B <= h->n_max, "pairs per batch must be in [1, max_pairs]");
  return IQLB_OK;
}

static int ptt_pack(iqlb_pt_trainer* h, const float* s1, const float* a1, const int32_t* t1, const float* m1, const float* s2,
                    const float* a2, const int32_t* t2, const float* m2, int B, cudaStream_t st) {
  const PtOff& P = h->P;
  const int64_t BT = (int64_t)B * P.T;
  const auto D = cudaMemcpyDeviceToDevice;
  IQLB_CUDA(cudaMemcpyAsync(h->St, s1, (size_t)BT * P.S * 4, D, st));
  IQLB_CUDA(cudaMemcpyAsync(h->St + BT * P.S, s2, (size_t)BT * P.S * 4, D, st));
  IQLB_CUDA(cudaMemcpyAsync(h->Ac, a1, (size_t)BT * P.A * 4, D, st));
  IQLB_CUDA(cudaMemcpyAsync(h->Ac + BT * P.A, a2, (size_t)BT * P.A * 4, D, st));
  IQLB_CUDA(cudaMemcpyAsync(h->Ts, t1, (size_t)BT * 4, D, st));
  IQLB_CUDA(cudaMemcpyAsync(h->Ts + BT, t2, (size_t)BT * 4, D, st));
  IQLB_CUDA(cudaMemcpyAsync(h->Mk, m1, (size_t)BT * 4, D, st));
  IQLB_CUDA(cudaMemcpyAsync(h->Mk + BT, m2, (size_t)BT * 4, D, st));
  return IQLB_OK;
}

// nn.Linear forward: C = act(A W^T + b)
static int ptt_linear(const iqlb_pt_trainer* h, const float* A, int lda, const float* W, const float* bias, float* C, int ldc,
                      int64_t M, int N, int K, bool relu, cudaStream_t st) {
  GemmTask t{};
  t.A = A; t.lda = lda; t.Bm = W; t.ldb = K; t.bias = bias; t.C = C; t.ldc = ldc;
  t.M = (int)M; t.N = N; t.K = K; t.act = relu ? IQLB_ACT_RELU : IQLB_ACT_LINEAR; t.scale = 1.f; t.allow_tc = h->tc;
  return launch_gemm(GEMM_LINEAR, t, st);
}

// data gradient of nn.Linear: dX[M, in] = dY[M, out] W[out][in]
static int ptt_linear_bwd_data(const iqlb_pt_trainer* h, const float* dY, int ldy, const float* W, float* dX, int64_t M,
                               int n_in, int n_out, cudaStream_t st) {
  GemmTask t{};
  t.A = dY; t.lda = ldy; t.Bm = W; t.ldb = n_in; t.bias = h->zero_bias; t.C = dX; t.ldc = n_in;
  t.M = (int)M; t.N = n_in; t.K = n_out; t.act = IQLB_ACT_LINEAR; t.scale = 1.f; t.allow_tc = h->tc;
  return launch_gemm(GEMM_FWD_KN, t, st);
}

// one grouped launch of an attention product (tasks built at create): tensor cores in split-bf16 mode when a 128 x 128 tile
// is worth it, else the grouped FFMA kernel
static int ptt_batched(const iqlb_pt_trainer* h, GemmMode mode, const GemmTask* tasks, int n_tasks, int M, int N, int K,
                       cudaStream_t st) {
  if (h->tc && M >= 64 && N >= 64 && K >= 16) return launch_grouped_gemm_tc(mode, tasks, n_tasks, M, N, st);
  return launch_grouped_gemm(mode, tasks, n_tasks, M, N, N <= 32, st);
}

static inline unsigned ew_grid(int64_t total) {
  const int64_t b = (total + 255) / 256;
  return (unsigned)(b < 148 * 16 ? (b > 0 ? b : 1) : 148 * 16);
}

static int ptt_forward_train(iqlb_pt_trainer* h, const float* w, int n, int64_t adam_step, bool training, cudaStream_t st) {
  const PtOff& P = h->P;
  const int E = P.E, T = P.T, T2 = 2 * T, I = P.I, H = P.H, dh = E / H, ldy = 2 * P.Pd + 1;
  const int64_t M = (int64_t)n * T2, nT = (int64_t)n * T;
  int rc;
  // embeddings: x[2t] = state_linear(s), x[2t+1] = action_linear(a); the LayerNorm kernel adds the timestep rows
  if ((rc = ptt_linear(h, h->St, P.S, w + P.sw, w + P.sb, h->x, 2 * E, nT, E, P.S, false, st))) return rc;
  if ((rc = ptt_linear(h, h->Ac, P.A, w + P.aw, w + P.ab, h->x + E, 2 * E, nT, E, P.A, false, st))) return rc;
  const unsigned ln_grid = (unsigned)((M + 7) / 8);
  ptt_ln_fwd_kernel<<<ln_grid, 256, 0, st>>>(E, P.eps, w + P.slg, w + P.slb, h->x, E, M, h->Ts, w + P.temb, h->x, h->xhat_s,
                                             h->rstd_s, make_drop(h, h->p_embd, 0, adam_step, training));
  IQLB_LAUNCH_CHECK();
  const PtDrop none = make_drop(h, 0.f, 0, adam_step, false);
  const int nh = n * H;
  const int64_t srows = (int64_t)nh * T2;
  const bool adrop = training && h->p_attn > 0.f;
  for (int l = 0; l < P.L; ++l) {
    const float* lw = w + P.layer0 + (int64_t)l * P.layer_stride;
    auto& y = h->layers[l];
    ptt_ln_fwd_kernel<<<ln_grid, 256, 0, st>>>(E, P.eps, lw + P.ln0g, lw + P.ln0b, h->x, E, M, nullptr, nullptr, y.h0, y.xhat0,
                                               y.rstd0, none);
    IQLB_LAUNCH_CHECK();
    if ((rc = ptt_linear(h, y.h0, E, lw + P.inw, lw + P.inb, y.qkv, 3 * E, M, 3 * E, E, false, st))) return rc;
    if ((rc = ptt_batched(h, GEMM_LINEAR, h->attn_tasks(l, 0), nh, T2, T2, dh, st))) return rc;
    ptt_softmax_fwd_kernel<<<(unsigned)((srows + 7) / 8), 256, 0, st>>>(T2, H, rsqrtf((float)dh), y.Pb, adrop ? y.Pd : y.Pb, h->Mk,
                                                                       srows, make_drop(h, h->p_attn, 1 + 4 * l, adam_step, training));
    IQLB_LAUNCH_CHECK();
    if ((rc = ptt_batched(h, GEMM_FWD_KN, h->attn_tasks(l, adrop ? 1 : 2), nh, T2, dh, T2, st))) return rc;
    if ((rc = ptt_linear(h, y.o, E, lw + P.outw, lw + P.outb, h->tmpA, E, M, E, E, false, st))) return rc;
    ptt_resid_kernel<<<ew_grid(M * E), 256, 0, st>>>(h->x, h->tmpA, M * E, make_drop(h, h->p_ra, 2 + 4 * l, adam_step, training));
    IQLB_LAUNCH_CHECK();
    ptt_ln_fwd_kernel<<<ln_grid, 256, 0, st>>>(E, P.eps, lw + P.ln1g, lw + P.ln1b, h->x, E, M, nullptr, nullptr, y.h1, y.xhat1,
                                               y.rstd1, none);
    IQLB_LAUNCH_CHECK();
    if ((rc = ptt_linear(h, y.h1, E, lw + P.m1w, lw + P.m1b, y.z, I, M, I, E, true, st))) return rc;
    if ((rc = ptt_linear(h, y.z, I, lw + P.m2w, lw + P.m2b, h->tmpA, E, M, E, I, false, st))) return rc;
    ptt_resid_kernel<<<ew_grid(M * E), 256, 0, st>>>(h->x, h->tmpA, M * E, make_drop(h, h->p_rm, 3 + 4 * l, adam_step, training));
    IQLB_LAUNCH_CHECK();
  }
  // final LayerNorm of the action tokens (pref_trans.py:298-301) -> pref_linear -> preference attention
  ptt_ln_fwd_kernel<<<(unsigned)((nT + 7) / 8), 256, 0, st>>>(E, P.eps, w + P.fg, w + P.fb, h->x + E, 2 * E, nT, nullptr, nullptr,
                                                              h->hid, h->xhat_f, h->rstd_f, none);
  IQLB_LAUNCH_CHECK();
  if ((rc = ptt_linear(h, h->hid, E, w + P.pw, w + P.pb, h->y, ldy, nT, ldy, E, false, st))) return rc;
  if ((rc = ptt_batched(h, GEMM_LINEAR, h->pref_tasks(0), n, T, T, P.Pd, st))) return rc;
  ptt_pref_softmax_fwd_kernel<<<(unsigned)((nT + 7) / 8), 256, 0, st>>>(T, P.Pd, rsqrtf((float)P.Pd), h->Sp, h->y, h->Mk, nT, h->ws);
  IQLB_LAUNCH_CHECK();
  return IQLB_OK;
}

extern "C" int iqlb_pt_eval(iqlb_pt_trainer* h, const float* params, const float* states1, const float* actions1,
                            const int32_t* timesteps1, const float* mask1, const float* states2, const float* actions2,
                            const int32_t* timesteps2, const float* mask2, const float* labels, int32_t B, float* out2,
                            float* wsum_out, void* stream) {
  int rc = ptt_check_batch(h, params, states1, actions1, timesteps1, mask1, states2, actions2, timesteps2, mask2, labels, B);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  if ((rc = ptt_pack(h, states1, actions1, timesteps1, mask1, states2, actions2, timesteps2, mask2, B, st))) return rc;
  const int n = 2 * B;
  if ((rc = ptt_forward_train(h, params, n, 1, false, st))) return rc;
  if ((rc = launch_bt_loss(h->ws, h->Mk, h->Mk + (int64_t)B * h->P.T, labels, B, h->P.T, h->pool_mean, nullptr, h->pair_out,
                           h->counter, out2 ? out2 : h->out2, st)))
    return rc;
  if (wsum_out) IQLB_CUDA(cudaMemcpyAsync(wsum_out, h->ws, (size_t)n * h->P.T * sizeof(float), cudaMemcpyDeviceToDevice, st));
  return IQLB_OK;
}

extern "C" int iqlb_pt_train_step(iqlb_pt_trainer* h, float* params, float* exp_avg, float* exp_avg_sq, const float* states1,
                                  const float* actions1, const int32_t* timesteps1, const float* mask1, const float* states2,
                                  const float* actions2, const int32_t* timesteps2, const float* mask2, const float* labels,
                                  int32_t B, float lr, float beta1, float beta2, float eps, int64_t adam_step,
                                  int32_t training, float* out2, float* grad_out, void* stream) {
  int rc = ptt_check_batch(h, params, states1, actions1, timesteps1, mask1, states2, actions2, timesteps2, mask2, labels, B);
  if (rc) return rc;
  IQLB_REQUIRE((exp_avg != nullptr && exp_avg_sq != nullptr) || lr == 0.f, "null optimizer state");
  IQLB_REQUIRE(adam_step >= 1, "adam_step counts from 1 (the step being taken)");
  cudaStream_t st = (cudaStream_t)stream;
  const PtOff& P = h->P;
  const int E = P.E, T = P.T, T2 = 2 * T, I = P.I, H = P.H, dh = E / H, L = P.L, ldy = 2 * P.Pd + 1;
  const int n = 2 * B;
  const int64_t M = (int64_t)n * T2, nT = (int64_t)n * T;
  const bool tr = training != 0;
  float* w = params;

  if ((rc = ptt_pack(h, states1, actions1, timesteps1, mask1, states2, actions2, timesteps2, mask2, B, st))) return rc;
  if ((rc = ptt_forward_train(h, w, n, adam_step, tr, st))) return rc;
  if ((rc = launch_bt_loss(h->ws, h->Mk, h->Mk + (int64_t)B * T, labels, B, T, h->pool_mean, h->rowgrad, h->pair_out, h->counter,
                           out2 ? out2 : h->out2, st)))
    return rc;

  // ---- weight-gradient task table: group g = one Linear, its tasks = row chunks ---------------------------------------
  int cr_tok = 0, cr_pos = 0;
  const int nc_tok = ptt_chunks(M, cr_tok), nc_pos = ptt_chunks(nT, cr_pos);
  struct Group { const float* A; int lda; const float* X; int ldx; int n_out, n_in; int64_t rows; int chunk_rows, n_chunks; int64_t off; };
  std::vector<Group> groups;
  groups.push_back({h->dx, 2 * E, h->St, P.S, E, P.S, nT, cr_pos, nc_pos, P.sw});       // 0 state_linear
  groups.push_back({h->dx + E, 2 * E, h->Ac, P.A, E, P.A, nT, cr_pos, nc_pos, P.aw});   // 1 action_linear
  for (int l = 0; l < L; ++l) {
    auto& y = h->layers[l];
    const int64_t lo = P.layer0 + (int64_t)l * P.layer_stride;
    groups.push_back({h->dbig, 3 * E, y.h0, E, 3 * E, E, M, cr_tok, nc_tok, lo + P.inw});                        // 2+4l attention in
    groups.push_back({h->p_ra > 0.f && tr ? h->tmpA : h->dx, E, y.o, E, E, E, M, cr_tok, nc_tok, lo + P.outw});  // 3+4l attention out
    groups.push_back({h->dbig, I, y.h1, E, I, E, M, cr_tok, nc_tok, lo + P.m1w});                                // 4+4l mlp in
    groups.push_back({h->p_rm > 0.f && tr ? h->tmpA : h->dx, E, y.z, I, E, I, M, cr_tok, nc_tok, lo + P.m2w});   // 5+4l mlp out
  }
  groups.push_back({h->dy, ldy, h->hid, E, ldy, E, nT, cr_pos, nc_pos, P.pw});          // last: pref_linear
  const int task_key = n * 2 + (tr ? 1 : 0);
  if (h->tasks_n != task_key) {
    std::vector<GemmTask> tasks((size_t)h->n_groups * h->max_chunks);
    for (size_t g = 0; g < groups.size(); ++g) {
      const Group& G = groups[g];
      const int64_t plen = (int64_t)G.n_out * G.n_in + G.n_out;
      for (int c = 0; c < G.n_chunks; ++c) {
        const int64_t r0 = (int64_t)c * G.chunk_rows;
        GemmTask t{};
        t.A = G.A + r0 * G.lda; t.lda = G.lda;  // A[k, m] : rows of dY
        t.Bm = G.X + r0 * G.ldx; t.ldb = G.ldx; // X[k, n] : rows of the layer input
        t.C = h->partial + (int64_t)c * plen; t.ldc = G.n_in;
        t.M = G.n_out; t.N = G.n_in; t.K = (int)((G.rows - r0 < G.chunk_rows) ? (G.rows - r0) : G.chunk_rows);
        tasks[g * h->max_chunks + c] = t;
      }
    }
    IQLB_CUDA(cudaStreamSynchronize(st));  // an earlier step on this stream may still read the old table
    IQLB_CUDA(cudaMemcpy(h->d_tasks, tasks.data(), tasks.size() * sizeof(GemmTask), cudaMemcpyHostToDevice));
    h->tasks_n = task_key;
  }

  const double bc1 = 1.0 - pow((double)beta1, (double)adam_step);
  const double bc2 = 1.0 - pow((double)beta2, (double)adam_step);
  const float step_size = (float)((double)lr / bc1), bc2_sqrt = (float)sqrt(bc2);
  const bool apply = exp_avg != nullptr;
  auto reduce_adam = [&](int64_t off, int64_t plen, int n_chunks, const float* partial) -> int {
    return launch_reduce_adam(partial, n_chunks, plen, 0, 1.f, apply ? params + off : nullptr, apply ? exp_avg + off : nullptr,
                              apply ? exp_avg_sq + off : nullptr, grad_out ? grad_out + off : nullptr, beta1, beta2, eps,
                              step_size, bc2_sqrt, st);
  };
  // dW = dY^T X (+ db = column sums of dY) of group g, then Adam on [W | b]
  auto weight_grad = [&](int g) -> int {
    const Group& G = groups[g];
    const int64_t plen = (int64_t)G.n_out * G.n_in + G.n_out;
    int r;
    if (h->tc && G.n_in >= 64 && G.n_out >= 64)
      r = launch_grouped_gemm_tc_bwd_weight(h->d_tasks + (size_t)g * h->max_chunks, G.n_chunks, G.n_out, G.n_in, st);
    else
      r = launch_grouped_gemm(GEMM_BWD_WEIGHT, h->d_tasks + (size_t)g * h->max_chunks, G.n_chunks, G.n_out, G.n_in, G.n_in <= 32, st);
    if (r) return r;
    if ((r = launch_colsum(G.A, G.lda, G.n_out, nullptr, 0, G.rows, G.chunk_rows, G.n_chunks, h->partial + (int64_t)G.n_out * G.n_in,
                           plen, st)))
      return r;
    return reduce_adam(G.off, plen, G.n_chunks, h->partial);
  };
  auto ln_bwd = [&](const float* g_ptr, int64_t g_off, const float* dyp, int64_t dy_stride, const PtDrop& drop, const float* xhat,
                    const float* rstd, int64_t rows, float* dxp, int64_t dx_stride, int accumulate) -> int {
    const int blocks = (int)((rows + kLnRowsPerBlock - 1) / kLnRowsPerBlock);
    if (E <= 256)
      ptt_ln_bwd_kernel<8><<<blocks, 256, (size_t)16 * E * sizeof(float), st>>>(E, g_ptr, dyp, dy_stride, drop, xhat, rstd, rows,
                                                                               kLnRowsPerBlock, dxp, dx_stride, accumulate, h->partial);
    else
      ptt_ln_bwd_kernel<kPtMaxPerLane><<<blocks, 256, (size_t)16 * E * sizeof(float), st>>>(
          E, g_ptr, dyp, dy_stride, drop, xhat, rstd, rows, kLnRowsPerBlock, dxp, dx_stride, accumulate, h->partial);
    IQLB_LAUNCH_CHECK();
    return reduce_adam(g_off, 2 * E, blocks, h->partial);
  };
  const PtDrop none = make_drop(h, 0.f, 0, adam_step, false);

  // ---- head: preference attention -> pref_linear -> final LayerNorm ----------------------------------------------------
  ptt_pref_softmax_bwd_kernel<<<n, 256, 0, st>>>(T, P.Pd, rsqrtf((float)P.Pd), h->Sp, h->y, h->ws, h->rowgrad, h->dSp, h->dy);
  IQLB_LAUNCH_CHECK();
  if ((rc = ptt_batched(h, GEMM_FWD_KN, h->pref_tasks(1), n, T, P.Pd, T, st))) return rc;      // dq = dS k
  if ((rc = ptt_batched(h, GEMM_BWD_WEIGHT, h->pref_tasks(2), n, T, P.Pd, T, st))) return rc;  // dk = dS^T q
  if ((rc = ptt_linear_bwd_data(h, h->dy, ldy, w + P.pw, h->tmpB, nT, E, ldy, st))) return rc;
  if ((rc = weight_grad((int)groups.size() - 1))) return rc;
  IQLB_CUDA(cudaMemsetAsync(h->dx, 0, (size_t)M * E * sizeof(float), st));  // state tokens of the last layer feed nothing
  if ((rc = ln_bwd(w + P.fg, P.fg, h->tmpB, E, none, h->xhat_f, h->rstd_f, nT, h->dx + E, 2 * E, 0))) return rc;

  const int nh = n * H;
  const int64_t srows = (int64_t)nh * T2;
  const bool adrop = tr && h->p_attn > 0.f;

  for (int l = L - 1; l >= 0; --l) {
    const int64_t lo = P.layer0 + (int64_t)l * P.layer_stride;
    const float* lw = w + lo;
    auto& y = h->layers[l];
    // MLP: x2 = x1 + drop(out_linear(relu(in_linear(LN1(x1)))))
    const float* dm = h->dx;
    if (h->p_rm > 0.f && tr) {
      ptt_mask_kernel<<<ew_grid(M * E), 256, 0, st>>>(h->dx, h->tmpA, M * E, make_drop(h, h->p_rm, 3 + 4 * l, adam_step, tr));
      IQLB_LAUNCH_CHECK();
      dm = h->tmpA;
    }
    if ((rc = ptt_linear_bwd_data(h, dm, E, lw + P.m2w, h->dbig, M, I, E, st))) return rc;
    ptt_relu_mask_kernel<<<ew_grid(M * I), 256, 0, st>>>(h->dbig, y.z, M * I);
    IQLB_LAUNCH_CHECK();
    if ((rc = weight_grad(5 + 4 * l))) return rc;
    if ((rc = ptt_linear_bwd_data(h, h->dbig, I, lw + P.m1w, h->tmpB, M, E, I, st))) return rc;
    if ((rc = weight_grad(4 + 4 * l))) return rc;
    if ((rc = ln_bwd(lw + P.ln1g, lo + P.ln1g, h->tmpB, E, none, y.xhat1, y.rstd1, M, h->dx, E, 1))) return rc;
    // attention: x1 = x0 + drop(out_linear(attn(in_linear(LN0(x0)))))
    const float* da = h->dx;
    if (h->p_ra > 0.f && tr) {
      ptt_mask_kernel<<<ew_grid(M * E), 256, 0, st>>>(h->dx, h->tmpA, M * E, make_drop(h, h->p_ra, 2 + 4 * l, adam_step, tr));
      IQLB_LAUNCH_CHECK();
      da = h->tmpA;
    }
    if ((rc = ptt_linear_bwd_data(h, da, E, lw + P.outw, h->tmpB, M, E, E, st))) return rc;
    if ((rc = weight_grad(3 + 4 * l))) return rc;
    const PtDrop ad = make_drop(h, h->p_attn, 1 + 4 * l, adam_step, tr);
    if ((rc = ptt_batched(h, GEMM_LINEAR, h->attn_tasks(l, 3), nh, T2, T2, dh, st))) return rc;                      // dPd = dO V^T
    if ((rc = ptt_batched(h, GEMM_BWD_WEIGHT, h->attn_tasks(l, adrop ? 6 : 7), nh, T2, dh, T2, st))) return rc;      // dV = Pd^T dO
    ptt_softmax_bwd_kernel<<<(unsigned)((srows + 7) / 8), 256, 0, st>>>(T2, rsqrtf((float)dh), y.Pb, h->dS, srows, ad);
    IQLB_LAUNCH_CHECK();
    if ((rc = ptt_batched(h, GEMM_FWD_KN, h->attn_tasks(l, 4), nh, T2, dh, T2, st))) return rc;                      // dQ = dS K
    if ((rc = ptt_batched(h, GEMM_BWD_WEIGHT, h->attn_tasks(l, 5), nh, T2, dh, T2, st))) return rc;                  // dK = dS^T Q
    if ((rc = ptt_linear_bwd_data(h, h->dbig, 3 * E, lw + P.inw, h->tmpB, M, E, 3 * E, st))) return rc;
    if ((rc = weight_grad(2 + 4 * l))) return rc;
    if ((rc = ln_bwd(lw + P.ln0g, lo + P.ln0g, h->tmpB, E, none, y.xhat0, y.rstd0, M, h->dx, E, 1))) return rc;
  }

  // ---- embeddings: dropout -> stacked LayerNorm -> state / action Linear + timestep table -------------------------------
  if ((rc = ln_bwd(w + P.slg, P.slg, h->dx, E, make_drop(h, h->p_embd, 0, adam_step, tr), h->xhat_s, h->rstd_s, M, h->dx, E, 0)))
    return rc;
  const int te_chunks = (int)((nT + kTembChunk - 1) / kTembChunk);
  ptt_temb_grad_kernel<<<dim3(P.n_ts, te_chunks), 256, 0, st>>>(E, h->Ts, nT, h->dx, h->partial, (int64_t)P.n_ts * E);
  IQLB_LAUNCH_CHECK();
  if ((rc = reduce_adam(P.temb, (int64_t)P.n_ts * E, te_chunks, h->partial))) return rc;
  if ((rc = weight_grad(0))) return rc;
  if ((rc = weight_grad(1))) return rc;
  return IQLB_OK;
}
