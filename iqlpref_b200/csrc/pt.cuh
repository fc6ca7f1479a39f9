// Preference-transformer pieces shared by the relabel pass (relabel_pt.cu) and the training step (pt_train.cu).
#pragma once
#include "iql_step.cuh"

namespace iqlb {

struct PtOff {
  int S, A, E, H, I, L, T, n_ts, Pd;
  float eps;
  int64_t sw, sb, aw, ab, temb, slg, slb, layer0, layer_stride, fg, fb, vw, vb, pw, pb, total;
  // inside a layer
  int64_t ln0g, ln0b, inw, inb, outw, outb, ln1g, ln1b, m1w, m1b, m2w, m2b;
  // train_layout: the parameters exactly in `net.parameters()` order (what torch.optim.Adam indexes): the relabel layout's
  // copy of the value row of pref_linear (vw, vb) is dropped and aliases the last row of pref_linear itself
  __host__ void init(const iqlb_pt_desc& d, bool train_layout = false) {
    S = d.state_dim; A = d.action_dim; E = d.embd_dim; H = d.num_heads; I = d.intermediate_dim; L = d.num_layers;
    T = d.query_len; n_ts = d.n_timesteps; eps = d.eps; Pd = d.pref_dim;
    int64_t o = 0;
    sw = o; o += (int64_t)E * S; sb = o; o += E;
    aw = o; o += (int64_t)E * A; ab = o; o += E;
    temb = o; o += (int64_t)n_ts * E;
    slg = o; o += E; slb = o; o += E;
    layer0 = o;
    int64_t q = 0;
    ln0g = q; q += E; ln0b = q; q += E;
    inw = q; q += (int64_t)3 * E * E; inb = q; q += 3 * E;
    outw = q; q += (int64_t)E * E; outb = q; q += E;
    ln1g = q; q += E; ln1b = q; q += E;
    m1w = q; q += (int64_t)I * E; m1b = q; q += I;
    m2w = q; q += (int64_t)E * I; m2b = q; q += E;
    layer_stride = q;
    o += (int64_t)L * layer_stride;
    fg = o; o += E; fb = o; o += E;
    if (!train_layout) { vw = o; o += E; vb = o; o += 1; }
    pw = o; o += (int64_t)(2 * Pd + 1) * E; pb = o; o += 2 * Pd + 1;
    if (train_layout) { vw = pw + (int64_t)2 * Pd * E; vb = pb + 2 * Pd; }
    total = o;
  }
};

constexpr int kPtMaxPerLane = 16;  // E <= 512

// LayerNorm of a row held as vals[c] = row[lane + 32 c]; two-pass (mean, then centred variance)
__device__ __forceinline__ void warp_layer_norm(float (&vals)[kPtMaxPerLane], int per, int E, float eps,
                                                const float* __restrict__ g, const float* __restrict__ b, int lane) {
  float s = 0.f;
#pragma unroll
  for (int c = 0; c < kPtMaxPerLane; ++c) if (c < per) s += vals[c];
  const float mean = warp_sum(s) / (float)E;
  float q = 0.f;
#pragma unroll
  for (int c = 0; c < kPtMaxPerLane; ++c) if (c < per) { const float d = vals[c] - mean; q += d * d; }
  const float rstd = rsqrtf(warp_sum(q) / (float)E + eps);
#pragma unroll
  for (int c = 0; c < kPtMaxPerLane; ++c)
    if (c < per) vals[c] = (vals[c] - mean) * rstd * __ldg(g + lane + 32 * c) + __ldg(b + lane + 32 * c);
}

constexpr int kPtMaxKeyChunks = 16;  // 2T <= 512

// relabel_pt.cu: the eval-mode forward (all_tokens: value / weighted_sum of every position; else the last token only)
int pt_check(const iqlb_pt_desc* d);
int pt_run(const iqlb_pt_desc* d, const float* w, int gather, const float* obs, const float* act,
           const int32_t* win_start, const int32_t* seq_len, const int32_t* timesteps, const float* attn_mask,
           int64_t n, bool all_tokens, float* out, float* wsum_out, cudaStream_t st, bool train_layout = false);

// preference-attention head (pref_trans.py:303-320): y [n T][2P+1] -> weighted_sum [n][T]
int launch_pt_pref_attn(int T, int Pd, const float* y, const float* maskbuf, float* out, int n, cudaStream_t st);

// mr_train.cu: pieces of the Bradley-Terry step both reward-model trainers use
int launch_bt_loss(const float* pred, const float* m1, const float* m2, const float* labels, int B, int T, int pool_mean,
                   float* rowgrad, float* pair_out, unsigned int* counter, float* out2, cudaStream_t st);
int launch_colsum(const float* X, int ld, int ncols, const float* rs, int rs_sum, int64_t M, int chunk_rows, int n_chunks,
                  float* dst, int64_t plen, cudaStream_t st);
int launch_reduce_adam(const float* partial, int n_chunks, int64_t plen, int64_t n_w, float wscale, float* p, float* m,
                       float* v, float* grad_out, float beta1, float beta2, float eps, float step_size, float bc2_sqrt,
                       cudaStream_t st);

}  // namespace iqlb
