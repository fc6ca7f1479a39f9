// Ensemble reward-MLP relabel with mean / CVaR reduction over members.
// Replaces the forward + np.partition blocks of qlearning_dataset_mr / _bnn / _mr_ensemble
// (/root/reference/algorithms/offline/iql.py:719-724, 982-1011, 1162-1187) and RewardMLP.forward
// (gp_reward-priors/optbnn/bnn/nets/mlp.py:73-93, layers/linear.py:43-47).
//
// Row chunks stream through: hidden layers as grouped fp32 GEMMs (one task per member), the
// width->1 output layer as a warp-per-row dot, then a per-row radix select of the n_tail-th
// smallest member prediction (bit-exact selection; ties broken by member index).
#include "iql_step.cuh"

namespace iqlb {

struct RelabelTc;  // relabel_mlp_tc.cu
bool relabel_mlp_tc_supported(int d_in, const int32_t* widths, int depth);
int relabel_tc_begin(RelabelTc** ctx, int ldx, int d_in, const float* weights, int S, const int32_t* widths, int depth,
                     int activation, cudaStream_t st);
int relabel_tc_forward(RelabelTc* c, const float* x, int64_t rows, float* preds, int64_t pred_stride, cudaStream_t st);
void relabel_tc_end(RelabelTc* c, cudaStream_t st);

// preds[s, i] = scale * H[s, i, :] . Wout[s, :] + bout[s]     (one warp per (row, member))
__global__ void __launch_bounds__(256) mlp_out_kernel(const float* __restrict__ Hbuf, int64_t h_member_stride, int ldh,
                                                      int width, const float* __restrict__ weights,
                                                      int64_t w_member_stride, int64_t off_wout, float scale,
                                                      int rows, float* __restrict__ preds, int64_t pred_stride) {
  const int s = blockIdx.y;
  const int lane = threadIdx.x & 31;
  const int i = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (i >= rows) return;
  const float* h = Hbuf + (int64_t)s * h_member_stride + (int64_t)i * ldh;
  const float* w = weights + (int64_t)s * w_member_stride + off_wout;
  float acc = 0.f;
  for (int j = lane; j < width; j += 32) acc = fmaf(h[j], __ldg(w + j), acc);
  acc = warp_sum(acc);
  if (lane == 0) preds[(int64_t)s * pred_stride + i] = fmaf(scale, acc, __ldg(w + width));
}

__device__ __forceinline__ uint32_t float_key(float f) {
  uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);  // ascending order-preserving
}

// One thread per row: mean of the n_tail smallest of preds[0..S-1, i] (iql.py:1009-1011).
// Radix select on order-preserving keys: 32 counting passes over the column (coalesced across
// the rows of a warp, L2-resident), then one pass that sums everything below the threshold.
// mean_out (nullable): the plain mean over ALL members of the same row, from the same pass (the reference prints
// "ensemble mean +- std" next to the CVaR reward, iql.py:1013-1021, 1196-1204: no second ensemble pass, no [S, N] matrix).
__global__ void __launch_bounds__(256) cvar_select_kernel(const float* __restrict__ preds, int64_t pred_stride, int S,
                                                          int n_tail, int rows, float* __restrict__ out,
                                                          uint8_t* __restrict__ sel, int64_t sel_stride,
                                                          float* __restrict__ mean_out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows) return;
  const float* col = preds + i;
  if (n_tail >= S) {
    double acc = 0.0;
    for (int s = 0; s < S; ++s) acc += (double)col[(int64_t)s * pred_stride];
    out[i] = (float)(acc / (double)S);
    if (mean_out) mean_out[i] = (float)(acc / (double)S);
    if (sel) for (int s = 0; s < S; ++s) sel[(int64_t)s * sel_stride + i] = 1;
    return;
  }
  uint32_t prefix = 0, mask = 0;
  int k = n_tail;  // rank (1-based) still to locate inside the current prefix class
  for (int bit = 31; bit >= 0; --bit) {
    const uint32_t bm = 1u << bit;
    int c0 = 0;
    for (int s = 0; s < S; ++s) {
      const uint32_t key = float_key(col[(int64_t)s * pred_stride]);
      c0 += ((key & mask) == prefix && (key & bm) == 0) ? 1 : 0;
    }
    if (k > c0) { k -= c0; prefix |= bm; }
    mask |= bm;
  }
  // prefix == key of the n_tail-th smallest; k == how many ties at the threshold are inside the tail
  double acc = 0.0, tot = 0.0;
  int ties_left = k;
  for (int s = 0; s < S; ++s) {
    const float v = col[(int64_t)s * pred_stride];
    const uint32_t key = float_key(v);
    bool in = key < prefix;
    if (key == prefix && ties_left > 0) { in = true; --ties_left; }
    if (in) acc += (double)v;
    tot += (double)v;
    if (sel) sel[(int64_t)s * sel_stride + i] = in ? 1 : 0;
  }
  out[i] = (float)(acc / (double)n_tail);
  if (mean_out) mean_out[i] = (float)(tot / (double)S);
}

}  // namespace iqlb

using namespace iqlb;

// Tensor-core path (relabel_mlp_tc.cu): member predictions by split-bf16 tcgen05 MMAs, then the same selection.
static int relabel_mlp_tensor(const float* x, int64_t n_rows, int ldx, int d_in, const float* weights, int S,
                              const int32_t* widths, int depth, int activation, int n_tail, float* out,
                              float* preds_out, uint8_t* sel_out, float* mean_out, cudaStream_t st) {
  RelabelTc* ctx = nullptr;
  int rc = relabel_tc_begin(&ctx, ldx, d_in, weights, S, widths, depth, activation, st);
  if (rc) return rc;
  const bool direct = S == 1 && !sel_out && !preds_out && !mean_out;  // a single model: its prediction is the result
  int64_t chunk = n_rows < (1ll << 30) ? n_rows : (1ll << 30);
  float* scratch = nullptr;
  if (!direct && !preds_out) {
    int64_t c = (int64_t)(512ll << 20) / ((int64_t)S * (int64_t)sizeof(float));
    c = c / 256 * 256;
    if (c < 256) c = 256;
    if (c < chunk) chunk = c;
    cudaError_t e = cudaMallocAsync((void**)&scratch, (size_t)S * (size_t)chunk * sizeof(float), st);
    if (e != cudaSuccess) {
      relabel_tc_end(ctx, st);
      return fail(IQLB_ERR_NOMEM, "relabel scratch: %s", cudaGetErrorString(e));
    }
  }
  for (int64_t r0 = 0; r0 < n_rows && rc == IQLB_OK; r0 += chunk) {
    const int64_t rows = n_rows - r0 < chunk ? n_rows - r0 : chunk;
    float* preds = preds_out ? preds_out + r0 : (direct ? out + r0 : scratch);
    const int64_t pstride = (preds_out || direct) ? n_rows : chunk;
    rc = relabel_tc_forward(ctx, x + r0 * ldx, rows, preds, pstride, st);
    if (rc == IQLB_OK && !direct) {
      cvar_select_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, st>>>(preds, pstride, S, n_tail, (int)rows, out + r0,
                                                                        sel_out ? sel_out + r0 : nullptr, n_rows,
                                                                        mean_out ? mean_out + r0 : nullptr);
      if (cudaGetLastError() != cudaSuccess) rc = fail(IQLB_ERR_CUDA, "cvar_select_kernel launch failed%s", "");
    }
  }
  if (scratch) cudaFreeAsync(scratch, st);
  relabel_tc_end(ctx, st);
  return rc;
}

extern "C" int iqlb_relabel_mlp_p(const float* x, int64_t n_rows, int32_t ldx, int32_t d_in, const float* weights,
                                  int32_t n_members, const int32_t* host_widths, int32_t depth, int32_t activation,
                                  int32_t n_tail, int32_t precision, float* out, float* preds_out, uint8_t* sel_out,
                                  float* mean_out, void* stream) {
  IQLB_REQUIRE(x && weights && host_widths && out, "null pointer");
  IQLB_REQUIRE(n_rows >= 0 && d_in > 0 && ldx >= d_in, "row shape");
  IQLB_REQUIRE(n_members > 0 && depth > 0 && depth <= 16, "ensemble shape");
  IQLB_REQUIRE(n_tail >= 1 && n_tail <= n_members, "n_tail must be in [1, S]");
  IQLB_REQUIRE(activation >= IQLB_ACT_RELU && activation <= IQLB_ACT_LEAKY_RELU, "activation id");
  IQLB_REQUIRE(precision >= IQLB_RELABEL_AUTO && precision <= IQLB_RELABEL_SPLIT_BF16, "precision id");
  if (n_rows == 0) return IQLB_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int S = n_members;
  const bool tc_ok = relabel_mlp_tc_supported(d_in, host_widths, depth);
  if (precision == IQLB_RELABEL_SPLIT_BF16 && !tc_ok) {
    snprintf(g_err, sizeof(g_err), "tensor-core relabel covers d_in <= 128, equal hidden widths of 64 or 128, depth <= 3");
    return IQLB_ERR_UNSUPPORTED;
  }
  if (precision != IQLB_RELABEL_FP32 && tc_ok)
    return relabel_mlp_tensor(x, n_rows, ldx, d_in, weights, S, host_widths, depth, activation, n_tail, out, preds_out,
                              sel_out, mean_out, st);
  // member parameter offsets
  std::vector<int64_t> offW(depth + 1), offB(depth + 1);
  int64_t pm = 0;
  int maxw = 0;
  int prev = d_in;
  for (int l = 0; l < depth; ++l) {
    IQLB_REQUIRE(host_widths[l] > 0, "hidden width");
    offW[l] = pm; pm += (int64_t)prev * host_widths[l];
    offB[l] = pm; pm += host_widths[l];
    prev = host_widths[l];
    if (prev > maxw) maxw = prev;
  }
  offW[depth] = pm; pm += prev;  // Wout[w_last][1]
  offB[depth] = pm; pm += 1;
  // chunk rows so that the two ping-pong activation buffers stay <= ~1 GiB in total
  int64_t chunk = (int64_t)(512ll << 20) / ((int64_t)S * maxw * (int64_t)sizeof(float));
  chunk = chunk / 128 * 128;
  if (chunk < 128) chunk = 128;
  if (chunk > 65536) chunk = 65536;
  if (chunk > n_rows) chunk = (n_rows + 127) / 128 * 128;
  float *actA = nullptr, *actB = nullptr, *preds_scratch = nullptr;
  GemmTask* d_tasks = nullptr;
  const int64_t act_elems = (int64_t)S * chunk * maxw;
  int rc = IQLB_OK;
  auto cleanup = [&]() {
    cudaStreamSynchronize(st);
    cudaFree(actA); cudaFree(actB); cudaFree(preds_scratch); cudaFree(d_tasks);
  };
#define RL_CUDA(e) do { cudaError_t e__ = (e); if (e__ != cudaSuccess) { snprintf(g_err, sizeof(g_err), "%s: %s", #e, cudaGetErrorString(e__)); cleanup(); return IQLB_ERR_CUDA; } } while (0)
  RL_CUDA(cudaMalloc((void**)&actA, (size_t)act_elems * sizeof(float)));
  if (depth > 1) RL_CUDA(cudaMalloc((void**)&actB, (size_t)act_elems * sizeof(float)));
  if (!preds_out) RL_CUDA(cudaMalloc((void**)&preds_scratch, (size_t)S * chunk * sizeof(float)));
  RL_CUDA(cudaMalloc((void**)&d_tasks, (size_t)S * depth * sizeof(GemmTask)));
  std::vector<GemmTask> tasks((size_t)S * depth);
  for (int64_t r0 = 0; r0 < n_rows; r0 += chunk) {
    const int rows = (int)((n_rows - r0 < chunk) ? (n_rows - r0) : chunk);
    prev = d_in;
    for (int l = 0; l < depth; ++l) {
      const int w = host_widths[l];
      float* dst = (l % 2 == 0) ? actA : actB;
      const float* src = (l % 2 == 0) ? actB : actA;
      for (int s = 0; s < S; ++s) {
        GemmTask t{};
        if (l == 0) { t.A = x + r0 * ldx; t.lda = ldx; }
        else { t.A = src + (int64_t)s * chunk * maxw; t.lda = maxw; }
        t.Bm = weights + (int64_t)s * pm + offW[l]; t.ldb = w;
        t.bias = weights + (int64_t)s * pm + offB[l];
        t.C = dst + (int64_t)s * chunk * maxw; t.ldc = maxw;
        t.M = rows; t.N = w; t.K = prev;
        t.scale = 1.0f / sqrtf((float)prev);
        t.act = activation;
        tasks[(size_t)l * S + s] = t;
      }
      prev = w;
    }
    RL_CUDA(cudaMemcpyAsync(d_tasks, tasks.data(), tasks.size() * sizeof(GemmTask), cudaMemcpyHostToDevice, st));
    for (int l = 0; l < depth; ++l) {
      rc = launch_grouped_gemm(GEMM_FWD_KN, d_tasks + (size_t)l * S, S, rows, host_widths[l], host_widths[l] <= 32, st);
      if (rc) { cleanup(); return rc; }
    }
    const float* last = ((depth - 1) % 2 == 0) ? actA : actB;
    float* preds = preds_out ? preds_out + r0 : preds_scratch;
    const int64_t pstride = preds_out ? n_rows : chunk;
    mlp_out_kernel<<<dim3((rows + 7) / 8, S), 256, 0, st>>>(last, chunk * maxw, maxw, prev, weights, pm, offW[depth],
                                                           1.0f / sqrtf((float)prev), rows, preds, pstride);
    RL_CUDA(cudaGetLastError());
    cvar_select_kernel<<<(rows + 255) / 256, 256, 0, st>>>(preds, pstride, S, n_tail, rows, out + r0,
                                                           sel_out ? sel_out + r0 : nullptr, n_rows,
                                                           mean_out ? mean_out + r0 : nullptr);
    RL_CUDA(cudaGetLastError());
    // the task array is re-written next chunk: wait for this chunk's launches to consume it
    if (r0 + chunk < n_rows) RL_CUDA(cudaStreamSynchronize(st));
  }
  cleanup();
#undef RL_CUDA
  return IQLB_OK;
}

extern "C" int iqlb_relabel_mlp(const float* x, int64_t n_rows, int32_t ldx, int32_t d_in, const float* weights,
                                int32_t n_members, const int32_t* host_widths, int32_t depth, int32_t activation,
                                int32_t n_tail, float* out, float* preds_out, uint8_t* sel_out, void* stream) {
  return iqlb_relabel_mlp_p(x, n_rows, ldx, d_in, weights, n_members, host_widths, depth, activation, n_tail,
                            IQLB_RELABEL_AUTO, out, preds_out, sel_out, nullptr, stream);
}

extern "C" int iqlb_keep_mask_scan(const uint8_t* terminals, const uint8_t* timeouts, int64_t n,
                                   int32_t max_episode_steps, int32_t terminate_on_end, uint8_t* keep,
                                   int32_t* ep_steps) {
  IQLB_REQUIRE(terminals != nullptr, "null terminals");
  IQLB_REQUIRE(n >= 1, "empty dataset");
  int32_t ep = 0;
  for (int64_t i = 0; i < n - 1; ++i) {
    if (ep_steps) ep_steps[i] = ep;
    const bool done = terminals[i] != 0;
    const bool final = timeouts ? (timeouts[i] != 0) : (ep == max_episode_steps - 1);
    if (!terminate_on_end && final) {
      if (keep) keep[i] = 0;
      ep = 0;
      continue;
    }
    if (keep) keep[i] = 1;
    if (done || final) ep = 0;
    ep += 1;
  }
  return IQLB_OK;
}
