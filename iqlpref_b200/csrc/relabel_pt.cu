// Preference-transformer relabel.  Replaces the forward passes of qlearning_dataset_pt
// (/root/reference/algorithms/offline/iql.py:1259-1301) and RewardPT.forward
// (gp_reward-priors/optbnn/bnn/nets/pref_trans.py:269-325, GPT2 block :118-183), eval mode.
//
// fp32 throughout (the reference runs this pass in fp32, no autocast).  Windows are processed in
// chunks; inside a chunk every Linear is one fp32 GEMM over all tokens (gemm_f32.cu), LayerNorm /
// embedding / attention are warp-per-row kernels.  The relabel only reads value[:, 0, -1, 0], so in
// the LAST transformer layer only the final token's query, output projection, MLP and LayerNorm are
// evaluated (K and V still cover all 2T tokens): for num_layers == 1 that removes ~85 % of the flops.
//
// Packed weights (fp32), in the order reward_models/pref_transformer.py:pack_pt_weights emits them:
//   state_linear W[E,S] b[E] | action_linear W[E,A] b[E] | timestep_embed [n_ts,E] | stacked LN g,b
//   per layer: LN0 g,b | attn in W[3E,E] b[3E] | attn out W[E,E] b[E] | LN1 g,b | mlp in W[I,E] b[I] | mlp out W[E,I] b[E]
//   final LN g,b | pref_linear.weight[2P] (value row) [E] | its bias [1] | pref_linear W[2P+1,E] b[2P+1] (the whole head:
//   query / key / value of the preference attention, pref_trans.py:303-320, needed for `weighted_sum` only)
#include "pt.cuh"

namespace iqlb {

// ------------------------------------------------------------------------------------------------
// Embedding: one warp per (window, position).  x[2t] = LN(state_linear(s) + temb[ts]),
// x[2t+1] = LN(action_linear(a) + temb[ts])  (pref_trans.py:272-286).  Gather mode builds the window
// the way iql.py:1264-1292 does: rows win_start.. left-padded with zeros, timesteps 0..seq_len-1.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) pt_embed_kernel(PtOff P, const float* __restrict__ w, int gather,
                                                       const float* __restrict__ obs, const float* __restrict__ act,
                                                       const int32_t* __restrict__ win_start, const int32_t* __restrict__ seq_len,
                                                       const int32_t* __restrict__ timesteps, const float* __restrict__ attn_mask,
                                                       int64_t win0, int n_win, float* __restrict__ x, float* __restrict__ maskbuf) {
  // Persistent CTAs; the two embedding matrices live in shared memory TRANSPOSED ([in][E]), so lane `l` reads output
  // feature l + 32 c of input i at consecutive addresses (the row-major weights read through __ldg cost 32 cache lines
  // per warp instruction: 3.5 ms of an 8.6 ms relabel of 4096 windows).
  extern __shared__ float sh[];  // WsT [S][E] | WaT [A][E] | per warp: S + A floats
  float* WsT = sh;
  float* WaT = WsT + (size_t)P.S * P.E;
  float* stage = WaT + (size_t)P.A * P.E;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < P.S * P.E; i += blockDim.x) {
    const int e = i / P.S, k = i - e * P.S;  // read row-major [E][S] coalesced, write transposed
    WsT[(size_t)k * P.E + e] = __ldg(w + P.sw + i);
  }
  for (int i = threadIdx.x; i < P.A * P.E; i += blockDim.x) {
    const int e = i / P.A, k = i - e * P.A;
    WaT[(size_t)k * P.E + e] = __ldg(w + P.aw + i);
  }
  __syncthreads();
  float* ss = stage + warp * (P.S + P.A);
  float* sa = ss + P.S;
  const int per = P.E / 32;
  const int64_t n_pairs = (int64_t)n_win * P.T;
  for (int64_t tp = (int64_t)blockIdx.x * 8 + warp; tp < n_pairs; tp += (int64_t)gridDim.x * 8) {
    const int wl = (int)(tp / P.T), t = (int)(tp % P.T);
    const int64_t wg = win0 + wl;
    const float* sp = nullptr;
    const float* ap = nullptr;
    int ts = 0;
    float m = 0.f;
    if (gather) {
      const int sl = seq_len[wg], pad = P.T - sl;
      if (t >= pad) {
        const int64_t row = (int64_t)win_start[wg] + (t - pad);
        sp = obs + row * P.S; ap = act + row * P.A; ts = t - pad; m = 1.f;
      }
    } else {
      sp = obs + (wg * P.T + t) * P.S; ap = act + (wg * P.T + t) * P.A;
      ts = timesteps[wg * P.T + t]; m = attn_mask[wg * P.T + t];
    }
    __syncwarp();
    for (int i = lane; i < P.S; i += 32) ss[i] = sp ? sp[i] : 0.f;
    for (int i = lane; i < P.A; i += 32) sa[i] = ap ? ap[i] : 0.f;
    __syncwarp();
    const float* te = w + P.temb + (int64_t)ts * P.E;
    float es[kPtMaxPerLane], ea[kPtMaxPerLane];
#pragma unroll
    for (int c = 0; c < kPtMaxPerLane; ++c) { es[c] = 0.f; ea[c] = 0.f; }
    for (int i = 0; i < P.S; ++i) {
      const float sv = ss[i];
      const float* wr = WsT + (size_t)i * P.E + lane;
#pragma unroll
      for (int c = 0; c < kPtMaxPerLane; ++c) if (c < per) es[c] = fmaf(sv, wr[32 * c], es[c]);
    }
    for (int i = 0; i < P.A; ++i) {
      const float av = sa[i];
      const float* wr = WaT + (size_t)i * P.E + lane;
#pragma unroll
      for (int c = 0; c < kPtMaxPerLane; ++c) if (c < per) ea[c] = fmaf(av, wr[32 * c], ea[c]);
    }
#pragma unroll
    for (int c = 0; c < kPtMaxPerLane; ++c) {
      if (c < per) {
        const int e = lane + 32 * c;
        const float tv = __ldg(te + e);
        es[c] = es[c] + __ldg(w + P.sb + e) + tv;
        ea[c] = ea[c] + __ldg(w + P.ab + e) + tv;
      }
    }
    warp_layer_norm(es, per, P.E, P.eps, w + P.slg, w + P.slb, lane);
    warp_layer_norm(ea, per, P.E, P.eps, w + P.slg, w + P.slb, lane);
    float* xs = x + ((int64_t)wl * 2 * P.T + 2 * t) * P.E;
#pragma unroll
    for (int c = 0; c < kPtMaxPerLane; ++c)
      if (c < per) { xs[lane + 32 * c] = es[c]; xs[P.E + lane + 32 * c] = ea[c]; }
    if (lane == 0) maskbuf[(int64_t)wl * P.T + t] = m;
  }
}

// y[r] = LN(x[r * in_stride .. + E))  (rows may be strided: the last token of every window)
__global__ void __launch_bounds__(256) pt_ln_kernel(int E, float eps, const float* __restrict__ g, const float* __restrict__ b,
                                                    const float* __restrict__ x, int64_t in_stride, int64_t n_rows,
                                                    float* __restrict__ y) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t r = (int64_t)blockIdx.x * 8 + warp;
  if (r >= n_rows) return;
  const int per = E / 32;
  float v[kPtMaxPerLane];
  const float* xr = x + r * in_stride;
#pragma unroll
  for (int c = 0; c < kPtMaxPerLane; ++c) if (c < per) v[c] = xr[lane + 32 * c];
  warp_layer_norm(v, per, E, eps, g, b, lane);
#pragma unroll
  for (int c = 0; c < kPtMaxPerLane; ++c) if (c < per) y[r * E + lane + 32 * c] = v[c];
}

__global__ void __launch_bounds__(256) pt_copy_rows_kernel(int E, const float* __restrict__ x, int64_t in_stride,
                                                           int64_t n_rows, float* __restrict__ y) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_rows * E) return;
  y[i] = x[(i / E) * in_stride + (i % E)];
}

// ------------------------------------------------------------------------------------------------
// Causal self-attention over the 2T tokens of one window (pref_trans.py:118-140): additive
// key-padding bias (1 - m) * -10000 (:35-40), -inf above the diagonal, softmax, scale 1/sqrt(dh).
// grid (windows, heads), 8 warps; K and V of the head live in shared memory; a warp owns a query.
// qkv rows are [q(E) | k(E) | v(E)]; head h is the slice [h*dh, (h+1)*dh) of each.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) pt_attn_full_kernel(int T2, int E, int dh, const float* __restrict__ qkv,
                                                           const float* __restrict__ maskbuf, float* __restrict__ out) {
  extern __shared__ float sh[];
  const int w = blockIdx.x, h = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ldk = dh + 1;
  float* Ks = sh;                       // [T2][dh+1]
  float* Vs = Ks + (size_t)T2 * ldk;    // [T2][dh+1]
  float* qs = Vs + (size_t)T2 * ldk;    // [8][dh]
  float* ps = qs + 8 * dh;              // [8][T2]
  const float* base = qkv + (int64_t)w * T2 * 3 * E + h * dh;
  for (int i = threadIdx.x; i < T2 * dh; i += blockDim.x) {
    const int j = i / dh, d = i % dh;
    Ks[j * ldk + d] = base[(int64_t)j * 3 * E + E + d];
    Vs[j * ldk + d] = base[(int64_t)j * 3 * E + 2 * E + d];
  }
  __syncthreads();
  const float scale = rsqrtf((float)dh);
  const float* mk = maskbuf + (int64_t)w * (T2 / 2);
  const int nchunk = (T2 + 31) / 32;
  float* q = qs + warp * dh;
  float* p = ps + warp * T2;
  for (int i = warp; i < T2; i += 8) {
    for (int d = lane; d < dh; d += 32) q[d] = base[(int64_t)i * 3 * E + d];
    __syncwarp();
    float s[kPtMaxKeyChunks];
    float mx = -INFINITY;
#pragma unroll
    for (int c = 0; c < kPtMaxKeyChunks; ++c) {
      if (c < nchunk) {
        const int j = lane + 32 * c;
        float v = -INFINITY;
        if (j < T2 && j <= i) {
          float acc = 0.f;
          for (int d = 0; d < dh; ++d) acc = fmaf(q[d], Ks[j * ldk + d], acc);
          v = acc * scale + (1.0f - mk[j >> 1]) * -10000.0f;
        }
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
    sum = warp_sum(sum);
    const float inv = 1.0f / sum;
#pragma unroll
    for (int c = 0; c < kPtMaxKeyChunks; ++c)
      if (c < nchunk && lane + 32 * c < T2) p[lane + 32 * c] = s[c] * inv;
    __syncwarp();
    for (int d = lane; d < dh; d += 32) {
      float acc = 0.f;
      for (int j = 0; j <= i; ++j) acc = fmaf(p[j], Vs[j * ldk + d], acc);
      out[((int64_t)w * T2 + i) * E + h * dh + d] = acc;
    }
    __syncwarp();
  }
}

// Attention of the LAST token only (it sees every key).  grid (windows), one warp per head.
// q: [windows][E]; kv rows: [k(E) | v(E)] for all 2T tokens.
__global__ void pt_attn_last_kernel(int T2, int E, int dh, const float* __restrict__ q, const float* __restrict__ kv,
                                    const float* __restrict__ maskbuf, float* __restrict__ out) {
  extern __shared__ float sh[];  // per head: p[T2]
  const int w = blockIdx.x, h = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* p = sh + h * T2;
  const float* qh = q + (int64_t)w * E + h * dh;
  const float* kb = kv + (int64_t)w * T2 * 2 * E + h * dh;
  const float* mk = maskbuf + (int64_t)w * (T2 / 2);
  const float scale = rsqrtf((float)dh);
  const int nchunk = (T2 + 31) / 32;
  float s[kPtMaxKeyChunks];
  float mx = -INFINITY;
#pragma unroll
  for (int c = 0; c < kPtMaxKeyChunks; ++c) {
    if (c < nchunk) {
      const int j = lane + 32 * c;
      float v = -INFINITY;
      if (j < T2) {
        const float* kj = kb + (int64_t)j * 2 * E;
        float acc = 0.f;
        for (int d = 0; d < dh; ++d) acc = fmaf(__ldg(qh + d), kj[d], acc);
        v = acc * scale + (1.0f - mk[j >> 1]) * -10000.0f;
      }
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
  sum = warp_sum(sum);
  const float inv = 1.0f / sum;
#pragma unroll
  for (int c = 0; c < kPtMaxKeyChunks; ++c)
    if (c < nchunk && lane + 32 * c < T2) p[lane + 32 * c] = s[c] * inv;
  __syncwarp();
  for (int d = lane; d < dh; d += 32) {
    float acc = 0.f;
    for (int j = 0; j < T2; ++j) acc = fmaf(p[j], kb[(int64_t)j * 2 * E + E + d], acc);
    out[(int64_t)w * E + h * dh + d] = acc;
  }
}

// value = LN_f(x_row) . wv + bv   (pref_trans.py:223, 303-312: the last column of pref_linear)
// rows: row r of x at x + r*in_stride; writes out[r * out_stride]
__global__ void __launch_bounds__(256) pt_value_kernel(int E, float eps, const float* __restrict__ g, const float* __restrict__ b,
                                                       const float* __restrict__ wv, const float* __restrict__ bv,
                                                       const float* __restrict__ x, int64_t in_stride, int64_t n_rows,
                                                       float* __restrict__ out) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t r = (int64_t)blockIdx.x * 8 + warp;
  if (r >= n_rows) return;
  const int per = E / 32;
  float v[kPtMaxPerLane];
  const float* xr = x + r * in_stride;
#pragma unroll
  for (int c = 0; c < kPtMaxPerLane; ++c) if (c < per) v[c] = xr[lane + 32 * c];
  warp_layer_norm(v, per, E, eps, g, b, lane);
  float acc = 0.f;
#pragma unroll
  for (int c = 0; c < kPtMaxPerLane; ++c) if (c < per) acc = fmaf(v[c], __ldg(wv + lane + 32 * c), acc);
  acc = warp_sum(acc);
  if (lane == 0) out[r] = acc + __ldg(bv);
}

// Preference attention head (pref_trans.py:303-320): y[n][t] = pref_linear(hidden action token t) = [q(P) | k(P) | v(1)];
// weighted_sum[n][t] = sum_j softmax_j(q_t . k_j / sqrt(P) + (1 - m_j) * -10000) v_j - full (non-causal) attention, one
// head, only the key-padding mask.  grid (sequences), one warp per query.
__global__ void __launch_bounds__(256) pt_pref_attn_kernel(int T, int Pd, const float* __restrict__ y,
                                                           const float* __restrict__ maskbuf, float* __restrict__ out) {
  extern __shared__ float sh[];  // per warp: scores [T]
  const int n = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ld = 2 * Pd + 1;
  const float* yb = y + (int64_t)n * T * ld;
  const float* mk = maskbuf + (int64_t)n * T;
  float* sc = sh + warp * T;
  const float scale = rsqrtf((float)Pd);
  for (int t = warp; t < T; t += 8) {
    const float* q = yb + (int64_t)t * ld;
    float mx = -INFINITY;
    for (int j = 0; j < T; ++j) {
      const float* k = yb + (int64_t)j * ld + Pd;
      float acc = 0.f;
      for (int d = lane; d < Pd; d += 32) acc = fmaf(q[d], k[d], acc);
      acc = warp_sum(acc);
      const float v = acc * scale + (1.0f - mk[j]) * -10000.0f;
      if (lane == 0) sc[j] = v;
      mx = fmaxf(mx, v);
    }
    __syncwarp();
    float sum = 0.f, ws = 0.f;
    for (int j = lane; j < T; j += 32) {
      const float e = expf(sc[j] - mx);
      sum += e;
      ws = fmaf(e, yb[(int64_t)j * ld + 2 * Pd], ws);
    }
    sum = warp_sum(sum);
    ws = warp_sum(ws);
    if (lane == 0) out[(int64_t)n * T + t] = ws / sum;
    __syncwarp();
  }
}

int launch_pt_pref_attn(int T, int Pd, const float* y, const float* maskbuf, float* out, int n, cudaStream_t st) {
  pt_pref_attn_kernel<<<n, 256, (size_t)8 * T * sizeof(float), st>>>(T, Pd, y, maskbuf, out);
  IQLB_LAUNCH_CHECK();
  return IQLB_OK;
}

// ------------------------------------------------------------------------------------------------
// host orchestration
// ------------------------------------------------------------------------------------------------
static int pt_linear(const float* A, int64_t lda, const float* W, const float* bias, float* C, const float* residual,
                     int M, int N, int K, bool relu, cudaStream_t st) {
  GemmTask t{};
  t.A = A; t.lda = (int)lda; t.Bm = W; t.ldb = K; t.bias = bias; t.C = C; t.ldc = N; t.mask = residual;
  t.M = M; t.N = N; t.K = K; t.act = relu ? IQLB_ACT_RELU : IQLB_ACT_LINEAR; t.scale = 1.f;
  t.allow_tc = 1;  // split-bf16 on the tensor cores for the token-level GEMMs (~1e-5 relative; the relabel bar is 1e-3)
  return launch_gemm(GEMM_LINEAR, t, st);
}

int pt_check(const iqlb_pt_desc* d) {
  IQLB_REQUIRE(d != nullptr, "null descriptor");
  IQLB_REQUIRE(d->state_dim > 0 && d->action_dim > 0 && d->num_layers >= 1 && d->intermediate_dim > 0, "dims");
  IQLB_REQUIRE(d->embd_dim >= 32 && d->embd_dim % 32 == 0 && d->embd_dim <= 32 * kPtMaxPerLane, "embd_dim multiple of 32, <= 512");
  IQLB_REQUIRE(d->num_heads >= 1 && d->num_heads <= 32 && d->embd_dim % d->num_heads == 0, "num_heads");
  IQLB_REQUIRE(d->query_len >= 1 && 2 * d->query_len <= 32 * kPtMaxKeyChunks, "query_len <= 256");
  IQLB_REQUIRE(d->n_timesteps >= d->query_len, "timestep table shorter than query_len");
  return IQLB_OK;
}

// all_tokens: value for every position (RewardPT.forward); else only the last token (relabel)
int pt_run(const iqlb_pt_desc* d, const float* w, int gather, const float* obs, const float* act,
           const int32_t* win_start, const int32_t* seq_len, const int32_t* timesteps, const float* attn_mask,
           int64_t n, bool all_tokens, float* out, float* wsum_out, cudaStream_t st, bool train_layout) {
  int rc = pt_check(d);
  if (rc) return rc;
  if (n == 0) return IQLB_OK;
  PtOff P;
  P.init(*d, train_layout);
  const int E = P.E, T2 = 2 * P.T, I = P.I, dh = E / P.H;
  const int full_layers = all_tokens ? P.L : P.L - 1;
  int64_t big_cols = full_layers > 0 ? (3 * E > I ? 3 * E : I) : 2 * E;
  if (wsum_out != nullptr && big_cols * 2 < 2 * P.Pd + 1) big_cols = (2 * P.Pd + 2) / 2;  // [n T, 2P+1] lives in `big` ([n 2T, big_cols])
  const int64_t per_tok = (int64_t)(2 * E + big_cols) * sizeof(float);
  int64_t chunk = (int64_t)(768ll << 20) / (per_tok * T2);
  if (chunk < 1) chunk = 1;
  if (chunk > 2048) chunk = 2048;
  if (chunk > n) chunk = n;
  const int64_t Mt = chunk * T2;
  float *x = nullptr, *hbuf = nullptr, *big = nullptr, *maskbuf = nullptr, *xl = nullptr, *hl = nullptr, *ql = nullptr, *midl = nullptr;
  // stream-ordered scratch from the device's default memory pool: after the first call the pool keeps the blocks, so a
  // relabel call neither synchronises nor pays cudaMalloc (the release threshold is raised once)
  static bool pool_set = false;
  if (!pool_set) {
    int dev = 0;
    cudaMemPool_t pool = nullptr;
    if (cudaGetDevice(&dev) == cudaSuccess && cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
      uint64_t keep = ~0ull;
      cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    }
    pool_set = true;
  }
  auto cleanup = [&]() {
    float* bufs[8] = {x, hbuf, big, maskbuf, xl, hl, ql, midl};
    for (float* b : bufs) if (b) cudaFreeAsync(b, st);
  };
#define PT_CUDA(e) do { cudaError_t e__ = (e); if (e__ != cudaSuccess) { snprintf(g_err, sizeof(g_err), "%s: %s", #e, cudaGetErrorString(e__)); cleanup(); return IQLB_ERR_CUDA; } } while (0)
#define PT_RC(e) do { int r__ = (e); if (r__) { cleanup(); return r__; } } while (0)
  PT_CUDA(cudaMallocAsync((void**)&x, (size_t)Mt * E * sizeof(float), st));
  PT_CUDA(cudaMallocAsync((void**)&hbuf, (size_t)Mt * E * sizeof(float), st));
  PT_CUDA(cudaMallocAsync((void**)&big, (size_t)Mt * big_cols * sizeof(float), st));
  PT_CUDA(cudaMallocAsync((void**)&maskbuf, (size_t)chunk * P.T * sizeof(float), st));
  if (!all_tokens) {
    PT_CUDA(cudaMallocAsync((void**)&xl, (size_t)chunk * E * sizeof(float), st));
    PT_CUDA(cudaMallocAsync((void**)&hl, (size_t)chunk * E * sizeof(float), st));
    PT_CUDA(cudaMallocAsync((void**)&ql, (size_t)chunk * E * sizeof(float), st));
    PT_CUDA(cudaMallocAsync((void**)&midl, (size_t)chunk * I * sizeof(float), st));
  }
  const size_t embed_smem = ((size_t)(P.S + P.A) * E + 8 * (P.S + P.A)) * sizeof(float);
  int n_sm = 148;
  {
    int dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
  }
  PT_CUDA(cudaFuncSetAttribute(pt_embed_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)embed_smem));
  const size_t attn_smem = ((size_t)2 * T2 * (dh + 1) + 8 * dh + 8 * T2) * sizeof(float);
  PT_CUDA(cudaFuncSetAttribute(pt_attn_full_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)attn_smem));
  for (int64_t w0 = 0; w0 < n; w0 += chunk) {
    const int nw = (int)(n - w0 < chunk ? n - w0 : chunk);
    const int64_t mt = (int64_t)nw * T2;
    {
      const int64_t want = ((int64_t)nw * P.T + 7) / 8;
      const unsigned grid = (unsigned)(want < 4 * n_sm ? want : 4 * n_sm);  // persistent: the weights are staged once per CTA
      pt_embed_kernel<<<grid, 256, embed_smem, st>>>(P, w, gather, obs, act, win_start, seq_len, timesteps, attn_mask, w0,
                                                     nw, x, maskbuf);
    }
    PT_CUDA(cudaGetLastError());
    for (int l = 0; l < P.L; ++l) {
      const float* lw = w + P.layer0 + (int64_t)l * P.layer_stride;
      const bool full = l < full_layers;
      pt_ln_kernel<<<(unsigned)((mt + 7) / 8), 256, 0, st>>>(E, P.eps, lw + P.ln0g, lw + P.ln0b, x, E, mt, hbuf);
      PT_CUDA(cudaGetLastError());
      if (full) {
        PT_RC(pt_linear(hbuf, E, lw + P.inw, lw + P.inb, big, nullptr, (int)mt, 3 * E, E, false, st));
        pt_attn_full_kernel<<<dim3(nw, P.H), 256, attn_smem, st>>>(T2, E, dh, big, maskbuf, hbuf);
        PT_CUDA(cudaGetLastError());
        PT_RC(pt_linear(hbuf, E, lw + P.outw, lw + P.outb, x, x, (int)mt, E, E, false, st));
        pt_ln_kernel<<<(unsigned)((mt + 7) / 8), 256, 0, st>>>(E, P.eps, lw + P.ln1g, lw + P.ln1b, x, E, mt, hbuf);
        PT_CUDA(cudaGetLastError());
        PT_RC(pt_linear(hbuf, E, lw + P.m1w, lw + P.m1b, big, nullptr, (int)mt, I, E, true, st));
        PT_RC(pt_linear(big, I, lw + P.m2w, lw + P.m2b, x, x, (int)mt, E, I, false, st));
      } else {
        // K, V for every token; Q, projection, MLP for the last token of each window only
        PT_RC(pt_linear(hbuf, E, lw + P.inw + (int64_t)E * E, lw + P.inb + E, big, nullptr, (int)mt, 2 * E, E, false, st));
        PT_RC(pt_linear(hbuf + (int64_t)(T2 - 1) * E, (int64_t)T2 * E, lw + P.inw, lw + P.inb, ql, nullptr, nw, E, E, false, st));
        pt_attn_last_kernel<<<nw, 32 * P.H, (size_t)P.H * T2 * sizeof(float), st>>>(T2, E, dh, ql, big, maskbuf, hl);
        PT_CUDA(cudaGetLastError());
        pt_copy_rows_kernel<<<(unsigned)(((int64_t)nw * E + 255) / 256), 256, 0, st>>>(E, x + (int64_t)(T2 - 1) * E,
                                                                                   (int64_t)T2 * E, nw, xl);
        PT_CUDA(cudaGetLastError());
        PT_RC(pt_linear(hl, E, lw + P.outw, lw + P.outb, xl, xl, nw, E, E, false, st));
        pt_ln_kernel<<<(unsigned)((nw + 7) / 8), 256, 0, st>>>(E, P.eps, lw + P.ln1g, lw + P.ln1b, xl, E, nw, hl);
        PT_CUDA(cudaGetLastError());
        PT_RC(pt_linear(hl, E, lw + P.m1w, lw + P.m1b, midl, nullptr, nw, I, E, true, st));
        PT_RC(pt_linear(midl, I, lw + P.m2w, lw + P.m2b, xl, xl, nw, E, I, false, st));
      }
    }
    if (all_tokens) {
      pt_value_kernel<<<(unsigned)(((int64_t)nw * P.T + 7) / 8), 256, 0, st>>>(E, P.eps, w + P.fg, w + P.fb, w + P.vw, w + P.vb,
                                                                           x + E, 2 * E, (int64_t)nw * P.T, out + w0 * P.T);
      if (wsum_out != nullptr) {
        // final LayerNorm of the action tokens -> pref_linear (q | k | v) -> preference attention
        const int64_t nt = (int64_t)nw * P.T;
        const int ldy = 2 * P.Pd + 1;
        pt_ln_kernel<<<(unsigned)((nt + 7) / 8), 256, 0, st>>>(E, P.eps, w + P.fg, w + P.fb, x + E, 2 * E, nt, hbuf);
        PT_CUDA(cudaGetLastError());
        PT_RC(pt_linear(hbuf, E, w + P.pw, w + P.pb, big, nullptr, (int)nt, ldy, E, false, st));
        pt_pref_attn_kernel<<<nw, 256, (size_t)8 * P.T * sizeof(float), st>>>(P.T, P.Pd, big, maskbuf, wsum_out + w0 * P.T);
        PT_CUDA(cudaGetLastError());
      }
    } else {
      pt_value_kernel<<<(unsigned)((nw + 7) / 8), 256, 0, st>>>(E, P.eps, w + P.fg, w + P.fb, w + P.vw, w + P.vb, xl, E, nw,
                                                            out + w0);
    }
    PT_CUDA(cudaGetLastError());
  }
  cleanup();
#undef PT_CUDA
#undef PT_RC
  return IQLB_OK;
}

}  // namespace iqlb

using namespace iqlb;

extern "C" int iqlb_pt_param_count(const iqlb_pt_desc* d, int64_t* n_floats) {
  int rc = pt_check(d);
  if (rc) return rc;
  IQLB_REQUIRE(n_floats != nullptr, "null output");
  PtOff P;
  P.init(*d);
  *n_floats = P.total;
  return IQLB_OK;
}

extern "C" int iqlb_relabel_pt(const iqlb_pt_desc* d, const float* weights, const float* obs, const float* act,
                               int64_t n_data, const int32_t* win_start, const int32_t* seq_len, int64_t n, float* out,
                               void* stream) {
  IQLB_REQUIRE(weights && obs && act && (n == 0 || (win_start && seq_len && out)), "null pointer");
  IQLB_REQUIRE(n >= 0 && n_data > 0, "sizes");
  return pt_run(d, weights, 1, obs, act, win_start, seq_len, nullptr, nullptr, n, false, out, nullptr, (cudaStream_t)stream);
}

extern "C" int iqlb_pt_forward(const iqlb_pt_desc* d, const float* weights, const float* states, const float* actions,
                               const int32_t* timesteps, const float* attn_mask, int64_t n, float* value_out,
                               float* weighted_sum_out, void* stream) {
  IQLB_REQUIRE(weights && (n == 0 || (states && actions && timesteps && attn_mask && value_out)), "null pointer");
  IQLB_REQUIRE(n >= 0, "sizes");
  IQLB_REQUIRE(d != nullptr && d->pref_dim >= 1, "pref_dim");
  return pt_run(d, weights, 0, states, actions, nullptr, nullptr, timesteps, attn_mask, n, true, value_out, weighted_sum_out,
                (cudaStream_t)stream);
}
