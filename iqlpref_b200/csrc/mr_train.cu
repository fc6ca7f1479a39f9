// Bradley-Terry preference-model training step of the MR reward model:
//   MRTrainer.train / .evaluation   gp_reward-priors/optbnn/training/training.py:27-72, 74-114
//   bt_pool_logit                   gp_reward-priors/optbnn/utils/util.py:356-364
//   RewardMLP forward               gp_reward-priors/optbnn/bnn/nets/mlp.py:73-111, layers/linear.py:34-47
//   optimizer                       torch.optim.Adam(net.parameters(), lr) (scripts_mr/run_mr_training.py:105)
// One call = forward over the 2*B*T rows of both segments, masked pooling, soft-label cross entropy, backward
// and Adam, ordered on the caller's stream (weight-gradient partials fork to a side stream and join back), no host
// synchronisation.  fp32 FFMA GEMMs (the reference trains in fp32,
// no autocast).  Weight gradients are split over row chunks (one CTA per chunk and tile) and summed in a fixed
// order by the kernel that also applies Adam, so a step is deterministic.
#include <cstdlib>
#include <vector>

#include "pt.cuh"

namespace iqlb {

// X[(seg*B + b)*T + t] = cat(states_seg[b,t,:], actions_seg[b,t,:])   (training.py:47-54)
__global__ void mr_pack_kernel(const float* __restrict__ s1, const float* __restrict__ a1,
                               const float* __restrict__ s2, const float* __restrict__ a2, int64_t BT, int S, int A,
                               float* __restrict__ X) {
  const int D = S + A;
  const int64_t total = 2 * BT * D;
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = e / D;
    const int c = (int)(e % D);
    const int seg = row >= BT;
    const int64_t r = row - (seg ? BT : 0);
    const float* s = seg ? s2 : s1;
    const float* a = seg ? a2 : a1;
    X[e] = c < S ? __ldg(s + r * S + c) : __ldg(a + r * A + (c - S));
  }
}

// pred[m] = scale * <H[m,:], w_out> + b_out, one warp per row
__global__ void mr_head_kernel(const float* __restrict__ H, int ld, int w, const float* __restrict__ w_out,
                               const float* __restrict__ b_out, float scale, int64_t M, float* __restrict__ pred) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  if (row >= M) return;
  float acc = 0.f;
  for (int k = lane; k < w; k += 32) acc = fmaf(__ldg(H + row * ld + k), __ldg(w_out + k), acc);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) pred[row] = fmaf(scale, acc, __ldg(b_out));
}

// Masked pooling, soft-label cross entropy (mean over pairs), accuracy, and the gradient of the loss with
// respect to every per-row prediction.  One 64-thread block per pair: warp `seg` pools its segment (coalesced over
// t, warp-shuffle sums), thread 0 forms the pair's loss, the two warps write
//   rowgrad[m] = dL/dpred[m] = dlogit[b, seg] * mask[b, t] / n[b, seg]   (0 where pred*mask is NaN: nansum)
// and the last block to finish sums the per-pair terms in a fixed order (deterministic mean).
__global__ void __launch_bounds__(64) mr_loss_kernel(const float* __restrict__ pred, const float* __restrict__ m1,
                                                     const float* __restrict__ m2, const float* __restrict__ labels,
                                                     int B, int T, int pool_mean, float* __restrict__ rowgrad,
                                                     float* __restrict__ pair_out, unsigned int* __restrict__ counter,
                                                     float* __restrict__ out2) {
  __shared__ float s_logit[2], s_ninv[2], s_dl[2];
  __shared__ bool s_last;
  const int b = blockIdx.x, seg = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t BT = (int64_t)B * T;
  const float* mk = (seg ? m2 : m1) + (int64_t)b * T;
  const float* pr = pred + seg * BT + (int64_t)b * T;
  float s = 0.f, n = 0.f;
  for (int t = lane; t < T; t += 32) {
    const float m = __ldg(mk + t), v = pr[t] * m;
    if (!isnan(v)) s += v;
    if (!isnan(m)) n += m;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { s += __shfl_xor_sync(0xffffffffu, s, o); n += __shfl_xor_sync(0xffffffffu, n, o); }
  if (lane == 0) {
    n = pool_mean ? fmaxf(n, 1.0f) : 1.0f;
    s_logit[seg] = s / n;
    s_ninv[seg] = 1.0f / n;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const float l0 = s_logit[0], l1 = s_logit[1];
    const float y0 = __ldg(labels + 2 * b), y1 = __ldg(labels + 2 * b + 1);
    const float mx = fmaxf(l0, l1);
    const float e0 = expf(l0 - mx), e1 = expf(l1 - mx);
    const float lse = mx + logf(e0 + e1);
    // torch.argmax returns the first maximal index
    const int pc = l1 > l0 ? 1 : 0, tc = y1 > y0 ? 1 : 0;
    pair_out[2 * b] = -(y0 * (l0 - lse) + y1 * (l1 - lse));
    pair_out[2 * b + 1] = (pc == tc) ? 1.f : 0.f;
    const float ys = y0 + y1, inv = 1.0f / (e0 + e1), invB = 1.0f / (float)B;
    s_dl[0] = (e0 * inv * ys - y0) * invB;
    s_dl[1] = (e1 * inv * ys - y1) * invB;
    __threadfence();
    s_last = atomicAdd(counter, 1u) == (unsigned)(B - 1);
  }
  __syncthreads();
  if (rowgrad != nullptr) {
    float* rg = rowgrad + seg * BT + (int64_t)b * T;
    const float c = s_dl[seg] * s_ninv[seg];
    for (int t = lane; t < T; t += 32) {
      const float m = __ldg(mk + t);
      rg[t] = isnan(pr[t] * m) ? 0.f : c * m;
    }
  }
  if (!s_last) return;
  __threadfence();
  __shared__ float r_loss[64], r_acc[64];
  float loss = 0.f, acc = 0.f;
  for (int i = threadIdx.x; i < B; i += 64) { loss += __ldcg(pair_out + 2 * i); acc += __ldcg(pair_out + 2 * i + 1); }
  r_loss[threadIdx.x] = loss; r_acc[threadIdx.x] = acc;
  __syncthreads();
  for (int o = 32; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) { r_loss[threadIdx.x] += r_loss[threadIdx.x + o]; r_acc[threadIdx.x] += r_acc[threadIdx.x + o]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) { out2[0] = r_loss[0] / (float)B; out2[1] = r_acc[0] / (float)B; *counter = 0u; }
}

// dZ[m,k] = rowgrad[m] * w_out[k] * scale * act'(Z[m,k])   (back through the output layer and the last activation)
__global__ void mr_head_bwd_kernel(const float* __restrict__ rowgrad, const float* __restrict__ w_out, float scale,
                                   const float* __restrict__ Z, int ld, int w, int64_t M, int act,
                                   float* __restrict__ dZ) {
  const int64_t total = M * w;
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t m = e / w;
    const int k = (int)(e % w);
    dZ[m * ld + k] = rowgrad[m] * __ldg(w_out + k) * scale * act_grad(act, Z[m * ld + k]);
  }
}

// Column sums of one row chunk per block: dst[c*plen + j] = sum_{m in chunk c} rs[m] * X[m, j]  (rs == null: 1);
// with rs_sum the extra column j == ncols receives sum_m rs[m].
__global__ void __launch_bounds__(128) mr_colsum_kernel(const float* __restrict__ X, int ld, int ncols,
                                                        const float* __restrict__ rs, int rs_sum, int64_t M,
                                                        int chunk_rows, float* __restrict__ dst, int64_t plen) {
  const int64_t r0 = (int64_t)blockIdx.x * chunk_rows;
  const int64_t r1 = r0 + chunk_rows < M ? r0 + chunk_rows : M;
  float* d = dst + (int64_t)blockIdx.x * plen;
  // (gridDim.y column tiles: the preference transformer's 768 / 1024-wide layers would leave 25-50 blocks on 148 SMs)
  for (int j = blockIdx.y * blockDim.x + threadIdx.x; j < ncols + (rs_sum ? 1 : 0); j += blockDim.x * gridDim.y) {
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    if (j < ncols) {
      int64_t m = r0;
      for (; m + 7 < r1; m += 8) {  // eight rows in flight per thread
        float x[8], r[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) { x[u] = X[(m + u) * ld + j]; r[u] = rs ? rs[m + u] : 1.f; }
        a0 = fmaf(r[0], x[0], a0); a1 = fmaf(r[1], x[1], a1); a2 = fmaf(r[2], x[2], a2); a3 = fmaf(r[3], x[3], a3);
        a0 = fmaf(r[4], x[4], a0); a1 = fmaf(r[5], x[5], a1); a2 = fmaf(r[6], x[6], a2); a3 = fmaf(r[7], x[7], a3);
      }
      for (; m < r1; ++m) a0 = fmaf(rs ? rs[m] : 1.f, X[m * ld + j], a0);
    } else {
      for (int64_t m = r0; m < r1; ++m) a0 += rs[m];
    }
    d[j] = (a0 + a1) + (a2 + a3);
  }
}

// g[e] = (e < n_w ? wscale : 1) * sum_c partial[c*plen + e], then torch.optim.Adam on element e
// (optim/adam.py: exp_avg.lerp_(g, 1-b1); exp_avg_sq = b2*v + (1-b2) g^2; p -= step_size * m / (sqrt(v)/sqrt(bc2) + eps)).
__global__ void __launch_bounds__(256) mr_reduce_adam_kernel(const float* __restrict__ partial, int n_chunks,
                                                             int64_t plen, int64_t n_w, float wscale,
                                                             float* __restrict__ p, float* __restrict__ m,
                                                             float* __restrict__ v, float* __restrict__ grad_out,
                                                             float beta1, float beta2, float eps, float step_size,
                                                             float bc2_sqrt) {
  // 64 elements x 4 chunk groups per block: group q sums chunks q, q+4, ... (eight loads in flight), then the four group sums
  // are added in a fixed order
  __shared__ float s_part[4][64];
  const int ex = threadIdx.x & 63, q = threadIdx.x >> 6;
  const int64_t e = blockIdx.x * 64ll + ex;
  float g0 = 0.f, g1 = 0.f;
  if (e < plen) {
    int c = q;
    for (; c + 28 < n_chunks; c += 32) {  // eight loads in flight per thread
      float t[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) t[u] = partial[(int64_t)(c + 4 * u) * plen + e];
      g0 += (t[0] + t[2]) + (t[4] + t[6]);
      g1 += (t[1] + t[3]) + (t[5] + t[7]);
    }
    for (; c < n_chunks; c += 4) g0 += partial[(int64_t)c * plen + e];
  }
  s_part[q][ex] = g0 + g1;
  __syncthreads();
  if (q != 0 || e >= plen) return;
  float g = (s_part[0][ex] + s_part[1][ex]) + (s_part[2][ex] + s_part[3][ex]);
  if (e < n_w) g *= wscale;
  if (grad_out != nullptr) grad_out[e] = g;
  if (p == nullptr) return;
  float mm = m[e], vv = v[e];
  mm = mm + (g - mm) * (1.0f - beta1);
  vv = beta2 * vv + (1.0f - beta2) * g * g;
  m[e] = mm; v[e] = vv;
  const float denom = sqrtf(vv) / bc2_sqrt + eps;
  p[e] = p[e] - step_size * (mm / denom);
}

// host launchers shared with the preference-transformer trainer (pt_train.cu; declared in pt.cuh)
// ------------------------------------------------------------------------------------------------
// Functional GP prior of an LCF model at the measurement points (MRTrainerF; gp/models/model.py:262-325, 422-449,
// 451-513): K = Phi Sigma Phi^T has rank <= d, so alpha = (K + eps I)^-1 (f - Phi p_mean) comes from the d x d system
//   eps = jitter * max(trace(Sigma Phi^T Phi) / n, 1),  A = Sigma^-1 + Phi^T Phi / eps = L L^T,
//   alpha = (r - Phi A^-1 Phi^T r / eps) / eps,  log|K + eps I| = log|A| + log|Sigma| + n log eps,
// all in fp64 as the reference does.  One CTA (d <= 64, a few hundred points: the algebra is ~1 MFLOP).
//   rowgrad != null : rowgrad[i] = float(alpha_i) * cot_scale   (the cotangent of f(X_M) in the backward pass: the step
//                     adds J^T alpha / N to the likelihood gradient, training.py:233-238)
//   prior_out != null : prior_out[0] = log p_GP(f) = -r.alpha / 2 - log|K + eps I| / 2 - n log(2 pi) / 2, prior_out[1] = eps
// ------------------------------------------------------------------------------------------------
constexpr int kFpMaxD = 64;
__global__ void __launch_bounds__(256) mr_fprior_kernel(const float* __restrict__ f, const double* __restrict__ phi, int n,
                                                        int d, const double* __restrict__ p_mean,
                                                        const double* __restrict__ sigma,
                                                        const double* __restrict__ sigma_inv, double logdet_sigma,
                                                        double jitter, float cot_scale, double* __restrict__ r_scratch,
                                                        float* __restrict__ rowgrad, double* __restrict__ prior_out) {
  __shared__ double A[kFpMaxD * kFpMaxD];
  __shared__ double bvec[kFpMaxD], red[256];
  __shared__ double s_eps, s_logdet;
  const int tid = threadIdx.x;
  auto block_sum = [&](double v) -> double {  // fixed-order tree: the result does not depend on scheduling
    red[tid] = v;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
      if (tid < o) red[tid] += red[tid + o];
      __syncthreads();
    }
    const double t = red[0];
    __syncthreads();
    return t;
  };
  for (int i = tid; i < n; i += 256) {  // r = f - Phi p_mean
    double m = 0.0;
    for (int j = 0; j < d; ++j) m += phi[(int64_t)i * d + j] * p_mean[j];
    r_scratch[i] = (double)f[i] - m;
  }
  for (int idx = tid; idx < d * d; idx += 256) {  // Phi^T Phi
    const int j = idx / d, k = idx % d;
    double acc = 0.0;
    for (int i = 0; i < n; ++i) acc += phi[(int64_t)i * d + j] * phi[(int64_t)i * d + k];
    A[idx] = acc;
  }
  __syncthreads();
  double part = 0.0;
  for (int idx = tid; idx < d * d; idx += 256) part += sigma[idx] * A[idx];
  const double kdm = block_sum(part) / (double)n;
  const double eps = jitter * (kdm > 1.0 ? kdm : 1.0);
  for (int idx = tid; idx < d * d; idx += 256) A[idx] = sigma_inv[idx] + A[idx] / eps;
  if (tid < d) {  // b = Phi^T r
    double acc = 0.0;
    for (int i = 0; i < n; ++i) acc += phi[(int64_t)i * d + tid] * r_scratch[i];
    bvec[tid] = acc;
  }
  __syncthreads();
  for (int k = 0; k < d; ++k) {  // Cholesky, lower triangle in place (a non-positive pivot gives NaN, where torch raises)
    if (tid == 0) A[k * d + k] = sqrt(A[k * d + k]);
    __syncthreads();
    for (int i = k + 1 + tid; i < d; i += 256) A[i * d + k] /= A[k * d + k];
    __syncthreads();
    const int rem = d - k - 1;
    for (int idx = tid; idx < rem * rem; idx += 256) {
      const int i = k + 1 + idx / rem, j = k + 1 + idx % rem;
      if (j <= i) A[i * d + j] -= A[i * d + k] * A[j * d + k];
    }
    __syncthreads();
  }
  if (tid == 0) {  // z = A^-1 b (two triangular solves), log|A|
    double ld = 0.0;
    for (int i = 0; i < d; ++i) {
      double v = bvec[i];
      for (int j = 0; j < i; ++j) v -= A[i * d + j] * bvec[j];
      bvec[i] = v / A[i * d + i];
      ld += log(A[i * d + i]);
    }
    for (int i = d - 1; i >= 0; --i) {
      double v = bvec[i];
      for (int j = i + 1; j < d; ++j) v -= A[j * d + i] * bvec[j];
      bvec[i] = v / A[i * d + i];
    }
    s_logdet = 2.0 * ld;
    s_eps = eps;
  }
  __syncthreads();
  double quad = 0.0;
  for (int i = tid; i < n; i += 256) {
    double pz = 0.0;
    for (int j = 0; j < d; ++j) pz += phi[(int64_t)i * d + j] * bvec[j];
    const double r = r_scratch[i], alpha = (r - pz / eps) / eps;
    if (rowgrad) rowgrad[i] = (float)alpha * cot_scale;
    quad += r * alpha;
  }
  quad = block_sum(quad);
  if (tid == 0 && prior_out) {
    const double logdet = s_logdet + logdet_sigma + (double)n * log(s_eps);
    prior_out[0] = -0.5 * quad - 0.5 * logdet - 0.5 * (double)n * 1.8378770664093453;  // log(2 pi)
    prior_out[1] = s_eps;
  }
}

int launch_bt_loss(const float* pred, const float* m1, const float* m2, const float* labels, int B, int T, int pool_mean,
                   float* rowgrad, float* pair_out, unsigned int* counter, float* out2, cudaStream_t st) {
  mr_loss_kernel<<<B, 64, 0, st>>>(pred, m1, m2, labels, B, T, pool_mean, rowgrad, pair_out, counter, out2);
  IQLB_LAUNCH_CHECK();
  return IQLB_OK;
}

int launch_colsum(const float* X, int ld, int ncols, const float* rs, int rs_sum, int64_t M, int chunk_rows, int n_chunks,
                  float* dst, int64_t plen, cudaStream_t st) {
  const int col_tiles = (ncols + (rs_sum ? 1 : 0) + 127) / 128;
  mr_colsum_kernel<<<dim3(n_chunks, col_tiles), 128, 0, st>>>(X, ld, ncols, rs, rs_sum, M, chunk_rows, dst, plen);
  IQLB_LAUNCH_CHECK();
  return IQLB_OK;
}

int launch_reduce_adam(const float* partial, int n_chunks, int64_t plen, int64_t n_w, float wscale, float* p, float* m,
                       float* v, float* grad_out, float beta1, float beta2, float eps, float step_size, float bc2_sqrt,
                       cudaStream_t st) {
  mr_reduce_adam_kernel<<<(unsigned)((plen + 63) / 64), 256, 0, st>>>(partial, n_chunks, plen, n_w, wscale, p, m, v, grad_out,
                                                                   beta1, beta2, eps, step_size, bc2_sqrt);
  IQLB_LAUNCH_CHECK();
  return IQLB_OK;
}

}  // namespace iqlb

using namespace iqlb;

struct iqlb_mr_trainer {
  int d_in = 0, depth = 0, act = 0, pool_mean = 1, maxw = 0;
  std::vector<int> widths;
  std::vector<int64_t> offW, offB;  // depth + 1 entries (last = output layer)
  int64_t pm = 0, max_rows = 0, max_plen = 0;
  int max_chunks = 0;
  float *X = nullptr, *pred = nullptr, *rowgrad = nullptr, *partial = nullptr, *G0 = nullptr, *G1 = nullptr;
  float* out2 = nullptr;
  float* pair_out = nullptr;        // [max_rows] per-pair loss / accuracy terms
  unsigned int* counter = nullptr;  // blocks of the loss kernel that have finished (reset by the last one)
  std::vector<float*> Z, H;
  GemmTask* d_tasks = nullptr;  // [depth][max_chunks] weight-gradient tasks
  int64_t tasks_rows = -1;      // row count the device task array was built for
  // the weight-gradient partials of a layer and its data-gradient GEMM both read dZ and nothing of each other: the
  // partials run on a side stream (fork / join by events, capturable) and fill the idle tail of the GEMM's last wave
  cudaStream_t side = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  bool overlap = true;  // IQLB_MR_OVERLAP=0: everything on the caller's stream
  int tc = 0;           // iqlb_mr_set_precision: 1 = the seven GEMMs of a step run on the tensor cores in split-bf16
  double* fp_r = nullptr;  // [max_rows] residual f - Phi p_mean of the functional-prior kernel (first iqlb_mr_train_step_f)
};

struct MrFPrior {  // arguments of iqlb_mr_train_step_f beyond those of iqlb_mr_train_step
  const float* x_meas; const double* phi; int n_meas, d;
  const double *p_mean, *sigma, *sigma_inv;
  double logdet_sigma, jitter, num_datapoints;
  double* prior_out;
};

static void mr_free(iqlb_mr_trainer* h) {
  if (!h) return;
  cudaFree(h->X); cudaFree(h->pred); cudaFree(h->rowgrad); cudaFree(h->partial); cudaFree(h->G0); cudaFree(h->G1);
  cudaFree(h->out2); cudaFree(h->d_tasks); cudaFree(h->pair_out); cudaFree(h->counter); cudaFree(h->fp_r);
  for (float* p : h->Z) cudaFree(p);
  for (float* p : h->H) cudaFree(p);
  if (h->side) cudaStreamDestroy(h->side);
  if (h->ev_fork) cudaEventDestroy(h->ev_fork);
  if (h->ev_join) cudaEventDestroy(h->ev_join);
  delete h;
}

static constexpr int kMrMaxChunks = 296;  // two one-CTA waves of weight-gradient tiles on 148 SMs

extern "C" int iqlb_mr_create(int32_t d_in, const int32_t* host_widths, int32_t depth, int32_t activation,
                              int32_t pool_mean, int64_t max_rows, iqlb_mr_trainer** out) {
  IQLB_REQUIRE(out != nullptr, "null out");
  IQLB_REQUIRE(d_in > 0 && depth >= 1 && depth <= 16 && host_widths != nullptr, "network shape");
  IQLB_REQUIRE(activation >= IQLB_ACT_RELU && activation <= IQLB_ACT_LEAKY_RELU, "activation id");
  IQLB_REQUIRE(max_rows > 0 && max_rows % 2 == 0, "max_rows = 2 * pairs * segment length");
  auto* h = new iqlb_mr_trainer();
  h->d_in = d_in; h->depth = depth; h->act = activation; h->pool_mean = pool_mean ? 1 : 0; h->max_rows = max_rows;
  int prev = d_in;
  for (int l = 0; l < depth; ++l) {
    if (host_widths[l] <= 0) { delete h; return fail(IQLB_ERR_INVALID, "invalid argument: hidden width%s", ""); }
    h->widths.push_back(host_widths[l]);
    h->offW.push_back(h->pm); h->pm += (int64_t)prev * host_widths[l];
    h->offB.push_back(h->pm); h->pm += host_widths[l];
    const int64_t plen = (int64_t)prev * host_widths[l] + host_widths[l];
    if (plen > h->max_plen) h->max_plen = plen;
    prev = host_widths[l];
    if (prev > h->maxw) h->maxw = prev;
  }
  h->offW.push_back(h->pm); h->pm += prev;
  h->offB.push_back(h->pm); h->pm += 1;
  h->max_chunks = kMrMaxChunks;
#define MR_ALLOC(ptr, n)                                                                          \
  do {                                                                                            \
    if (cudaMalloc((void**)&(ptr), (size_t)(n) * sizeof(float)) != cudaSuccess) {                 \
      cudaGetLastError(); mr_free(h);                                                             \
      return fail(IQLB_ERR_NOMEM, "iqlb_mr_create: device allocation failed%s", "");              \
    }                                                                                             \
  } while (0)
  MR_ALLOC(h->X, max_rows * d_in);
  MR_ALLOC(h->pred, max_rows);
  MR_ALLOC(h->rowgrad, max_rows);
  MR_ALLOC(h->partial, (int64_t)h->max_chunks * h->max_plen);
  MR_ALLOC(h->G0, max_rows * h->maxw);
  MR_ALLOC(h->G1, max_rows * h->maxw);
  MR_ALLOC(h->out2, 2);
  MR_ALLOC(h->pair_out, max_rows);
  MR_ALLOC(h->counter, 1);
  cudaMemset(h->counter, 0, sizeof(unsigned int));
  h->Z.assign(depth, nullptr); h->H.assign(depth, nullptr);
  for (int l = 0; l < depth; ++l) { MR_ALLOC(h->Z[l], max_rows * h->widths[l]); MR_ALLOC(h->H[l], max_rows * h->widths[l]); }
#undef MR_ALLOC
  if (cudaMalloc((void**)&h->d_tasks, (size_t)depth * h->max_chunks * sizeof(GemmTask)) != cudaSuccess) {
    cudaGetLastError(); mr_free(h);
    return fail(IQLB_ERR_NOMEM, "iqlb_mr_create: device allocation failed%s", "");
  }
  const char* ov = getenv("IQLB_MR_OVERLAP");
  h->overlap = !(ov && ov[0] == '0');
  if (h->overlap && (cudaStreamCreateWithFlags(&h->side, cudaStreamNonBlocking) != cudaSuccess ||
                     cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming) != cudaSuccess ||
                     cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming) != cudaSuccess)) {
    cudaGetLastError(); mr_free(h);
    return fail(IQLB_ERR_CUDA, "iqlb_mr_create: stream / event creation failed%s", "");
  }
  *out = h;
  return IQLB_OK;
}

extern "C" int iqlb_mr_destroy(iqlb_mr_trainer* h) {
  if (h) cudaDeviceSynchronize();
  mr_free(h);
  return IQLB_OK;
}

extern "C" int iqlb_mr_set_precision(iqlb_mr_trainer* h, int32_t precision) {
  IQLB_REQUIRE(h != nullptr, "null handle");
  IQLB_REQUIRE(precision == 0 || precision == 1, "precision must be 0 (fp32 FFMA) or 1 (split-bf16 tcgen05)");
  h->tc = precision;
  return IQLB_OK;
}

extern "C" int iqlb_mr_param_count(const iqlb_mr_trainer* h, int64_t* n_params) {
  IQLB_REQUIRE(h != nullptr && n_params != nullptr, "null argument");
  *n_params = h->pm;
  return IQLB_OK;
}

// Forward over the 2*B*T rows of both segments (B == 0: none) followed by n_extra plain rows x_extra [n_extra][d_in]
// (the measurement points of MRTrainerF): pred[0 .. 2BT + n_extra).
static int mr_forward(iqlb_mr_trainer* h, const float* params, const float* s1, const float* a1, const float* s2,
                      const float* a2, int B, int T, int S, int A, bool keep_pre, cudaStream_t st,
                      const float* x_extra = nullptr, int n_extra = 0) {
  const int64_t BT = (int64_t)B * T, M = 2 * BT + n_extra;
  if (BT > 0) {
    const int64_t total = 2 * BT * h->d_in;
    const int blocks = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
    mr_pack_kernel<<<blocks, 256, 0, st>>>(s1, a1, s2, a2, BT, S, A, h->X);
    IQLB_LAUNCH_CHECK();
  }
  if (n_extra > 0)
    IQLB_CUDA(cudaMemcpyAsync(h->X + 2 * BT * h->d_in, x_extra, (size_t)n_extra * h->d_in * sizeof(float),
                              cudaMemcpyDeviceToDevice, st));
  int prev = h->d_in;
  const float* src = h->X;
  for (int l = 0; l < h->depth; ++l) {
    const int w = h->widths[l];
    GemmTask t{};
    t.A = src; t.lda = prev;
    t.Bm = params + h->offW[l]; t.ldb = w;
    t.bias = params + h->offB[l];
    t.C = h->H[l]; t.ldc = w;
    t.pre = keep_pre ? h->Z[l] : nullptr;
    t.M = (int)M; t.N = w; t.K = prev;
    t.scale = 1.0f / sqrtf((float)prev);
    t.act = h->act;
    t.allow_tc = h->tc;
    const int rc = launch_gemm(GEMM_FWD_KN, t, st);
    if (rc) return rc;
    src = h->H[l]; prev = w;
  }
  mr_head_kernel<<<(unsigned)((M * 32 + 255) / 256), 256, 0, st>>>(src, prev, prev, params + h->offW[h->depth],
                                                                  params + h->offB[h->depth],
                                                                  1.0f / sqrtf((float)prev), M, h->pred);
  IQLB_LAUNCH_CHECK();
  return IQLB_OK;
}

static int mr_check_batch(const iqlb_mr_trainer* h, const void* params, const void* s1, const void* a1,
                          const void* m1, const void* s2, const void* a2, const void* m2, const void* labels,
                          int B, int T, int S, int A) {
  IQLB_REQUIRE(h != nullptr, "null trainer");
  IQLB_REQUIRE(params && s1 && m1 && s2 && m2 && labels, "null pointer");
  IQLB_REQUIRE((a1 && a2) || A == 0, "null actions with action_dim > 0");
  IQLB_REQUIRE(B > 0 && T > 0 && S > 0 && A >= 0, "batch shape");
  IQLB_REQUIRE(S + A == h->d_in, "state_dim + action_dim must equal the network input width");
  IQLB_REQUIRE(2ll * B * T <= h->max_rows, "batch larger than max_rows given to iqlb_mr_create");
  return IQLB_OK;
}

extern "C" int iqlb_mr_eval(iqlb_mr_trainer* h, const float* params, const float* states1, const float* actions1,
                            const float* mask1, const float* states2, const float* actions2, const float* mask2,
                            const float* labels, int32_t B, int32_t T, int32_t S, int32_t A, float* out2,
                            float* pred_out, void* stream) {
  int rc = mr_check_batch(h, params, states1, actions1, mask1, states2, actions2, mask2, labels, B, T, S, A);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  rc = mr_forward(h, params, states1, actions1, states2, actions2, B, T, S, A, false, st);
  if (rc) return rc;
  mr_loss_kernel<<<B, 64, 0, st>>>(h->pred, mask1, mask2, labels, B, T, h->pool_mean, nullptr, h->pair_out, h->counter,
                                   out2 ? out2 : h->out2);
  IQLB_LAUNCH_CHECK();
  if (pred_out) IQLB_CUDA(cudaMemcpyAsync(pred_out, h->pred, (size_t)2 * B * T * sizeof(float), cudaMemcpyDeviceToDevice, st));
  return IQLB_OK;
}

static int mr_train_impl(iqlb_mr_trainer* h, float* params, float* exp_avg, float* exp_avg_sq, const float* states1,
                         const float* actions1, const float* mask1, const float* states2, const float* actions2,
                         const float* mask2, const float* labels, int32_t B, int32_t T, int32_t S, int32_t A, float lr,
                         float beta1, float beta2, float eps, int64_t adam_step, float* out2, float* grad_out,
                         const MrFPrior* fp, void* stream) {
  int rc = mr_check_batch(h, params, states1, actions1, mask1, states2, actions2, mask2, labels, B, T, S, A);
  if (rc) return rc;
  IQLB_REQUIRE((exp_avg != nullptr && exp_avg_sq != nullptr) || lr == 0.f, "null optimizer state");
  IQLB_REQUIRE(adam_step >= 1, "adam_step counts from 1 (the step being taken)");
  cudaStream_t st = (cudaStream_t)stream;
  const int n_meas = fp ? fp->n_meas : 0;
  if (fp) {
    IQLB_REQUIRE(fp->x_meas && fp->phi && fp->p_mean && fp->sigma && fp->sigma_inv, "null prior pointer");
    IQLB_REQUIRE(n_meas > 0 && fp->d > 0 && fp->d <= kFpMaxD, "1 <= n_concepts <= 64 and at least one measurement point");
    IQLB_REQUIRE(2ll * B * T + n_meas <= h->max_rows, "2*B*T + n_meas larger than max_rows given to iqlb_mr_create");
    IQLB_REQUIRE(fp->num_datapoints > 0 && fp->jitter > 0, "num_datapoints and jitter must be positive");
    if (!h->fp_r) IQLB_CUDA(cudaMalloc((void**)&h->fp_r, (size_t)h->max_rows * sizeof(double)));
  }
  const int64_t M = 2ll * B * T + n_meas;  // the measurement points ride through the same GEMMs as extra rows
  const int L = h->depth;
  rc = mr_forward(h, params, states1, actions1, states2, actions2, B, T, S, A, true, st, fp ? fp->x_meas : nullptr, n_meas);
  if (rc) return rc;
  mr_loss_kernel<<<B, 64, 0, st>>>(h->pred, mask1, mask2, labels, B, T, h->pool_mean, h->rowgrad, h->pair_out,
                                   h->counter, out2 ? out2 : h->out2);
  IQLB_LAUNCH_CHECK();
  if (fp) {  // dU/df(X_M) = alpha / N (training.py:233-238: param.grad += J^T alpha / N)
    mr_fprior_kernel<<<1, 256, 0, st>>>(h->pred + 2ll * B * T, fp->phi, n_meas, fp->d, fp->p_mean, fp->sigma, fp->sigma_inv,
                                        fp->logdet_sigma, fp->jitter, (float)(1.0 / fp->num_datapoints), h->fp_r,
                                        h->rowgrad + 2ll * B * T, nullptr);
    IQLB_LAUNCH_CHECK();
  }

  // row chunks of the split weight-gradient reductions
  int n_chunks = (int)((M + 127) / 128);
  if (n_chunks > h->max_chunks) n_chunks = h->max_chunks;
  int chunk_rows = (int)((M + n_chunks - 1) / n_chunks);
  chunk_rows = (chunk_rows + 7) / 8 * 8;
  n_chunks = (int)((M + chunk_rows - 1) / chunk_rows);

  if (h->tasks_rows != M) {
    std::vector<GemmTask> tasks((size_t)L * h->max_chunks);
    int prev = h->d_in;
    for (int l = 0; l < L; ++l) {
      const int w = h->widths[l];
      const float* below = l == 0 ? h->X : h->H[l - 1];
      const float* dz = ((L - 1 - l) % 2 == 0) ? h->G0 : h->G1;
      const int64_t plen = (int64_t)prev * w + w;
      for (int c = 0; c < n_chunks; ++c) {
        const int64_t r0 = (int64_t)c * chunk_rows;
        GemmTask t{};
        t.A = below + r0 * prev; t.lda = prev;      // A[k, m'] : rows of the layer input
        t.Bm = dz + r0 * w; t.ldb = w;              // X[k, n'] : rows of dZ
        t.C = h->partial + (int64_t)c * plen; t.ldc = w;
        t.M = prev; t.N = w; t.K = (int)((M - r0 < chunk_rows) ? (M - r0) : chunk_rows);
        tasks[(size_t)l * h->max_chunks + c] = t;
      }
      prev = w;
    }
    // the previous task array may still be in use by launches of an earlier step on this stream
    IQLB_CUDA(cudaStreamSynchronize(st));
    IQLB_CUDA(cudaMemcpy(h->d_tasks, tasks.data(), tasks.size() * sizeof(GemmTask), cudaMemcpyHostToDevice));
    h->tasks_rows = M;
  }

  const double bc1 = 1.0 - pow((double)beta1, (double)adam_step);
  const double bc2 = 1.0 - pow((double)beta2, (double)adam_step);
  const float step_size = (float)((double)lr / bc1), bc2_sqrt = (float)sqrt(bc2);
  const bool apply = lr != 0.f || exp_avg != nullptr;
  auto reduce_adam = [&](int64_t off, int64_t plen, int64_t n_w, float wscale) -> int {
    mr_reduce_adam_kernel<<<(unsigned)((plen + 63) / 64), 256, 0, st>>>(
        h->partial, n_chunks, plen, n_w, wscale, apply ? params + off : nullptr, apply ? exp_avg + off : nullptr,
        apply ? exp_avg_sq + off : nullptr, grad_out ? grad_out + off : nullptr, beta1, beta2, eps, step_size,
        bc2_sqrt);
    IQLB_LAUNCH_CHECK();
    return IQLB_OK;
  };
  const int ew_blocks = 148 * 8;

  // output layer: dZ of the last hidden layer first (it reads w_out before Adam rewrites it)
  const int wl = h->widths[L - 1];
  const float out_scale = 1.0f / sqrtf((float)wl);
  mr_head_bwd_kernel<<<ew_blocks, 256, 0, st>>>(h->rowgrad, params + h->offW[L], out_scale, h->Z[L - 1], wl, wl, M,
                                                h->act, h->G0);
  IQLB_LAUNCH_CHECK();
  mr_colsum_kernel<<<n_chunks, 128, 0, st>>>(h->H[L - 1], wl, wl, h->rowgrad, 1, M, chunk_rows, h->partial, wl + 1);
  IQLB_LAUNCH_CHECK();
  rc = reduce_adam(h->offW[L], wl + 1, wl, out_scale);
  if (rc) return rc;

  for (int l = L - 1; l >= 0; --l) {
    const int w = h->widths[l];
    const int prev = l == 0 ? h->d_in : h->widths[l - 1];
    float* dz = ((L - 1 - l) % 2 == 0) ? h->G0 : h->G1;
    float* other = ((L - 1 - l) % 2 == 0) ? h->G1 : h->G0;
    const int64_t plen = (int64_t)prev * w + w;
    // the partial buffer is free again (the previous reduce on `st` has been enqueued) and dZ of this layer is ready
    const bool fork = h->overlap && l > 0;
    cudaStream_t ws = fork ? h->side : st;
    if (fork) {
      IQLB_CUDA(cudaEventRecord(h->ev_fork, st));
      IQLB_CUDA(cudaStreamWaitEvent(h->side, h->ev_fork, 0));
    }
    if (h->tc && w >= 64)
      rc = launch_grouped_gemm_tc_bwd_weight(h->d_tasks + (size_t)l * h->max_chunks, n_chunks, prev, w, ws);
    else
      rc = launch_grouped_gemm(GEMM_BWD_WEIGHT, h->d_tasks + (size_t)l * h->max_chunks, n_chunks, prev, w, w <= 32, ws);
    if (rc) return rc;
    mr_colsum_kernel<<<n_chunks, 128, 0, ws>>>(dz, w, w, nullptr, 0, M, chunk_rows, h->partial + (int64_t)prev * w,
                                               plen);
    IQLB_LAUNCH_CHECK();
    if (fork) IQLB_CUDA(cudaEventRecord(h->ev_join, h->side));
    if (l > 0) {
      GemmTask t{};  // dH[m, k_in] = sum_n dZ[m, n] W[k_in, n], through the activation of the layer below
      t.A = dz; t.lda = w;
      t.Bm = params + h->offW[l]; t.ldb = w;
      t.C = other; t.ldc = prev;
      t.M = (int)M; t.N = prev; t.K = w;
      t.pre = h->Z[l - 1];                    // epilogue: dZ of the layer below = scale * dH * act'(Z)
      t.scale = 1.0f / sqrtf((float)prev);
      t.act = h->act;
      t.allow_tc = h->tc;
      rc = launch_gemm(GEMM_LINEAR, t, st);
      if (rc) return rc;
    }
    if (fork) IQLB_CUDA(cudaStreamWaitEvent(st, h->ev_join, 0));
    rc = reduce_adam(h->offW[l], plen, (int64_t)prev * w, 1.0f / sqrtf((float)prev));
    if (rc) return rc;
  }
  if (fp && fp->prior_out) {  // log p_GP(f(X_M)) with the updated weights (training.py:242-250)
    rc = mr_forward(h, params, nullptr, nullptr, nullptr, nullptr, 0, 0, S, A, false, st, fp->x_meas, n_meas);
    if (rc) return rc;
    mr_fprior_kernel<<<1, 256, 0, st>>>(h->pred, fp->phi, n_meas, fp->d, fp->p_mean, fp->sigma, fp->sigma_inv,
                                        fp->logdet_sigma, fp->jitter, 0.f, h->fp_r, nullptr, fp->prior_out);
    IQLB_LAUNCH_CHECK();
  }
  return IQLB_OK;
}

extern "C" int iqlb_mr_train_step(iqlb_mr_trainer* h, float* params, float* exp_avg, float* exp_avg_sq,
                                  const float* states1, const float* actions1, const float* mask1,
                                  const float* states2, const float* actions2, const float* mask2,
                                  const float* labels, int32_t B, int32_t T, int32_t S, int32_t A, float lr,
                                  float beta1, float beta2, float eps, int64_t adam_step, float* out2,
                                  float* grad_out, void* stream) {
  return mr_train_impl(h, params, exp_avg, exp_avg_sq, states1, actions1, mask1, states2, actions2, mask2, labels, B, T, S,
                       A, lr, beta1, beta2, eps, adam_step, out2, grad_out, nullptr, stream);
}

extern "C" int iqlb_mr_train_step_f(iqlb_mr_trainer* h, float* params, float* exp_avg, float* exp_avg_sq,
                                    const float* states1, const float* actions1, const float* mask1,
                                    const float* states2, const float* actions2, const float* mask2,
                                    const float* labels, int32_t B, int32_t T, int32_t S, int32_t A, float lr,
                                    float beta1, float beta2, float eps, int64_t adam_step, const float* x_meas,
                                    const double* phi, int32_t n_meas, int32_t n_concepts, const double* p_mean,
                                    const double* sigma, const double* sigma_inv, double logdet_sigma, double jitter,
                                    double num_datapoints, float* out2, double* prior_out, float* grad_out,
                                    void* stream) {
  MrFPrior fp{x_meas, phi, n_meas, n_concepts, p_mean, sigma, sigma_inv, logdet_sigma, jitter, num_datapoints, prior_out};
  return mr_train_impl(h, params, exp_avg, exp_avg_sq, states1, actions1, mask1, states2, actions2, mask2, labels, B, T, S,
                       A, lr, beta1, beta2, eps, adam_step, out2, grad_out, &fp, stream);
}
