// Internal declarations shared by the IQL step translation units.
#pragma once
#include <vector>

#include "common.cuh"
#include "umma.cuh"

namespace iqlb {

// Stream order of the 7 MLP forwards of one update:
//   0 V(next_obs)  1 V(obs)  2 Qt1(sa)  3 Qt2(sa)  4 Q1(sa)  5 Q2(sa)  6 actor(obs)
// Trainable streams (backward + Adam): 1 (V), 4 (Q1), 5 (Q2), 6 (actor).
constexpr int kStreams = 7;
constexpr int kTrain = 4;
__host__ __device__ inline int train_stream(int t) { return t == 0 ? 1 : (t == 1 ? 4 : (t == 2 ? 5 : 6)); }

struct Layout {
  int S, A, H, Dq, R, B;
  int deterministic;
  int64_t qnet, vnet, anet;       // floats per network
  int64_t off_q1, off_q2, off_v, off_a, off_logstd;
  int64_t P, PQ, Pstride, PQstride;
  __host__ __device__ static int64_t net_size(int in, int H, int out) {
    return (int64_t)H * in + H + (int64_t)H * H + H + (int64_t)out * H + out;
  }
  __host__ __device__ static int64_t al8(int64_t n) { return (n + 7) / 8 * 8; }
  __host__ __device__ void init(int S_, int A_, int H_, int det, int R_, int B_) {
    S = S_; A = A_; H = H_; Dq = S_ + A_; R = R_; B = B_; deterministic = det;
    qnet = net_size(Dq, H, 1);
    vnet = net_size(S, H, 1);
    anet = net_size(S, H, A);
    // every network starts on a 32-byte boundary, so rows of W1 are 128-bit addressable in all four
    off_q1 = 0; off_q2 = al8(qnet); off_v = off_q2 + al8(qnet); off_a = off_v + al8(vnet); off_logstd = off_a + al8(anet);
    PQ = off_v;
    P = off_logstd + (det ? 0 : A);
    Pstride = al8(P);
    PQstride = al8(PQ);
  }
  // offsets inside one network
  __host__ __device__ int64_t w0() const { return 0; }
  __host__ __device__ int64_t b0(int in) const { return (int64_t)H * in; }
  __host__ __device__ int64_t w1(int in) const { return b0(in) + H; }
  __host__ __device__ int64_t b1(int in) const { return w1(in) + (int64_t)H * H; }
  __host__ __device__ int64_t w2(int in) const { return b1(in) + H; }
  __host__ __device__ int64_t b2(int in, int out) const { return w2(in) + (int64_t)out * H; }
  // per stream: input width, output width, offset of the net inside params (or target for 2,3)
  __host__ __device__ int in_dim(int s) const { return (s >= 2 && s <= 5) ? Dq : S; }
  __host__ __device__ int out_dim(int s) const { return s == 6 ? A : 1; }
  __host__ __device__ int64_t net_off(int s) const {
    switch (s) {
      case 0: case 1: return off_v;
      case 2: case 4: return off_q1;
      case 3: case 5: return off_q2;
      default: return off_a;
    }
  }
  __host__ __device__ bool uses_target(int s) const { return s == 2 || s == 3; }
  // head_out rows: 0 next_v, 1 v, 2 tq1, 3 tq2, 4 q1, 5 q2, 6.. mu[a]
  __host__ __device__ int head_rows() const { return 6 + A; }
  // d_out rows: 0 dv, 1 dq1, 2 dq2, 3.. dz[a]
  __host__ __device__ int dout_rows() const { return 3 + A; }
};

struct GemmTask {
  const float* A;
  const float* Bm;
  float* C;
  const float* bias;    // FWD: bias[n]
  const float* mask;    // BWD_DATA: relu mask source (same shape/ld as C)
  float* bias_grad;     // BWD_WEIGHT: column sums of A
  int M, N, K, lda, ldb, ldc;
  float scale;          // FWD_KN: act(scale * acc + bias); BWD_DATA: gradient scale (1, or 1/(1-p) through a dropout)
  int act;              // FWD_KN: IQLB_ACT_* id
  // FWD with actor dropout (iql.py:437-438, nn.Dropout after every hidden ReLU of the actor): element (m, n) is kept
  // when word (n & 3) of Philox4x32-10(key = rng seed of the agent, subsequence = (layer+1) << 32 | (m*N + n) / 4,
  // counter = Adam step of the agent) is >= drop_thresh; kept values are scaled by drop_scale = 1/(1-p)
  uint32_t drop_thresh;  // 0: no dropout
  float drop_scale;
  int drop_layer;
  const uint64_t* drop_rng;  // &rng[2 * agent] (seed)
  const int64_t* drop_it;    // &it[agent]
  float* pre;                // FWD_KN (nullable): pre-activation scale * acc + bias, same shape/ld as C (kept for the backward pass);
                             // LINEAR (nullable): read instead - C = scale * acc * act'(pre) (data gradient through an activation)
  unsigned long long* dbg;   // gemm_tc.cu: CTA 0 records clock64 stamps here (diagnostics; nullptr otherwise)
  int allow_tc;              // 1: launch_gemm may run this task on the tensor cores in split-bf16 arithmetic (gemm_tc.cu,
                             // ~1e-5 relative) when the shape is worth a 128 x 128 tile; 0: fp32 FFMA always
};

enum GemmMode { GEMM_FWD = 0, GEMM_BWD_DATA = 1, GEMM_BWD_WEIGHT = 2, GEMM_FWD_KN = 3, GEMM_LINEAR = 4 };

struct StepScalars {  // per agent, written by the loss kernel, read by Adam
  float ss_q, ss_v, ss_a, bc2_sqrt, tau;
};

int launch_gemm(GemmMode mode, const GemmTask& t, cudaStream_t stream);
bool gemm_tc_eligible(GemmMode mode, const GemmTask& t);
int launch_gemm_tc(GemmMode mode, const GemmTask& t, cudaStream_t stream);
int launch_grouped_gemm_tc_bwd_weight(const GemmTask* d_tasks, int n_tasks, int M, int N, cudaStream_t stream);
int launch_grouped_gemm_tc(GemmMode mode, const GemmTask* d_tasks, int n_tasks, int maxM, int maxN, cudaStream_t stream);
int launch_grouped_gemm(GemmMode mode, const GemmTask* d_tasks, int n_tasks, int maxM, int maxN, bool narrow_n,
                        cudaStream_t stream);

constexpr int kOutMax = 24;  // widest head handled in registers (actor action_dim)

struct BfParams {
  Layout L;
  int KXB;                // 64-column blocks of the input images (ceil((S+A)/64))
  int agent0;             // first agent of this launch (agent groups are pipelined, iql_step.cu)
  int rev;                // 1: this launch walks its agents from the last to the first (L2 zig-zag, iql_step.cu)
  int64_t w0_bytes;       // W0 image bytes  = KXB * H * 128
  int64_t net_bytes;      // W0 + W1 images of one network
  int64_t agent_w_bytes;  // 6 networks: Q1 Q2 V actor Qt1 Qt2
  int64_t x_bytes;        // one input image [B x 64*KXB]
  int64_t h_bytes;        // one activation image [B x H]
  int64_t z_bytes;        // one dz image [B x 64]
  uint8_t *wimg, *xsa, *xnext, *h1, *h2, *dh2, *dh1, *dz;
  uint32_t *mask1, *mask2;  // [agent][train net][B][H/32]
  __nv_bfloat16* gradh;     // [agent][Pstride] gradients of the step in bf16 (what the reference's autocast GEMMs emit)
  // actor dropout (iql.py:436-437; include/iqlb.h iqlb_agents_set_actor_dropout): 0 = off
  uint32_t drop_thresh;   // keep an element when its Philox word >= drop_thresh (= p * 2^32)
  float drop_scale;       // 1 / (1 - p); 1 when off
  const uint64_t* drop_rng;  // rng[2 * agent] = seed (Philox key)
  const int64_t* drop_it;    // it[agent] = Adam steps done so far (Philox counter)
  int opt_lpt;            // data-gradient kernel: actor tiles (the longest) first (IQLB_LPT)
  int dbg_z;              // timeline: blockIdx.z of the recording CTA
  int opt_issuers;        // MMA-issuing threads of the gradient kernels (IQLB_ISSUERS, 1 or 2)
  unsigned long long* dbg;  // IQLB_TIMELINE=1: clock64 stamps of CTA 0 (kernel timelines, debug tensor 8)
};

struct Bf16State {
  BfParams P;
  bool dirty = true;
};

__host__ __device__ inline int net_of_stream(int s) {  // index into the weight images
  switch (s) {
    case 0: case 1: return 2;  // V
    case 2: return 4;          // Q target 1
    case 3: return 5;          // Q target 2
    case 4: return 0;
    case 5: return 1;
    default: return 3;         // actor
  }
}
__host__ __device__ inline int train_index(int s) { return s == 1 ? 0 : (s == 4 ? 1 : (s == 5 ? 2 : (s == 6 ? 3 : -1))); }

#ifdef __CUDACC__
// Actor dropout on 32 consecutive hidden units [c0, c0 + 32) of batch row `brow` of hidden layer `layer` (0 or 1):
// element (m, n) is kept when word (n & 3) of Philox4x32-10(key = seed, subsequence = (layer + 1) << 32 | (m * H + n) / 4,
// counter = step) is >= thresh, and kept values are scaled by 1 / (1 - p) - the mask documented in include/iqlb.h, the
// same one the fp32 path draws (gemm_f32.cu) and tests/test_gpu_iql_step.py rebuilds with the numpy Philox.
__device__ __forceinline__ void dropout32(float (&v)[32], uint64_t seed, uint64_t step, int layer, int brow, int H, int c0,
                                          uint32_t thresh, float scale) {
#pragma unroll
  for (int j4 = 0; j4 < 8; ++j4) {
    uint32_t r[4];
    philox4(seed, ((uint64_t)(layer + 1) << 32) | (uint64_t)(((int64_t)brow * H + c0 + 4 * j4) >> 2), step * 4ull, r);
#pragma unroll
    for (int k = 0; k < 4; ++k) v[4 * j4 + k] = r[k] >= thresh ? v[4 * j4 + k] * scale : 0.f;
  }
}
#endif

// image location of parameter element e (offset inside one agent's packed parameters)
__device__ __forceinline__ bool shadow_loc(const Layout& L, int KXB, int64_t e, int& net, int64_t& off) {
  int n, in;
  int64_t r;
  if (e < L.off_q2) { n = 0; r = e; in = L.Dq; }
  else if (e < L.off_v) { n = 1; r = e - L.off_q2; in = L.Dq; }
  else if (e < L.off_a) { n = 2; r = e - L.off_v; in = L.S; }
  else if (e < L.off_logstd) { n = 3; r = e - L.off_a; in = L.S; }
  else return false;
  const int64_t w0n = (int64_t)L.H * in;
  if (r < w0n) {
    off = umma::image_off(L.H, (int)(r / in), (int)(r % in));
  } else {
    r -= w0n + L.H;
    if (r < 0 || r >= (int64_t)L.H * L.H) return false;
    off = (int64_t)KXB * L.H * 128 + umma::image_off(L.H, (int)(r / L.H), (int)(r % L.H));
  }
  net = n;
  return true;
}


}  // namespace iqlb

struct iqlb_agents {
  iqlb_agents_desc desc;
  iqlb::Layout L;
  // caller-owned
  float *params, *exp_avg, *exp_avg_sq, *target;
  const float* hparams;
  int64_t* it;
  double* actor_lr;
  uint64_t* rng;
  float *losses, *loss_sum;
  // replay
  const float* table = nullptr;
  int64_t n_rows = 0;
  const float* reward_cols = nullptr;
  const int32_t* reward_variant = nullptr;
  int64_t* idx_out = nullptr;
  // lib-owned scratch
  float *batch = nullptr, *H1 = nullptr, *H2 = nullptr, *head_out = nullptr, *d_out = nullptr, *dH2 = nullptr,
        *dH1 = nullptr, *grads = nullptr;
  iqlb::StepScalars* scal = nullptr;
  iqlb::GemmTask *t_fwd0 = nullptr, *t_fwd1 = nullptr, *t_bw1 = nullptr, *t_bd1 = nullptr, *t_bw0 = nullptr;
  void* bf16_state = nullptr;  // precision==1 extension (tcgen05 path)
  float drop_p = 0.f;          // actor dropout probability (precision 0 only)
  // per-kernel CUDA-event timing (iqlb_agents_profile): ev[i], ev[i+1] bracket launch i of a step
  static constexpr int kMaxProf = 24;
  bool prof = false;
  cudaEvent_t ev[kMaxProf + 1] = {};
  const char* prof_name[kMaxProf] = {};
  double prof_ms[kMaxProf] = {};
  int prof_n = 0, prof_steps = 0;
  int launches_last = 0;
  // agent-group pipelining (precision 1): Adam of group g runs on side_stream while the tensor-core
  // chain of group g+1 runs on the caller's stream
  static constexpr int kMaxGroups = 8;
  cudaStream_t side_stream = nullptr;
  cudaEvent_t ev_chain[kMaxGroups] = {};
  cudaEvent_t ev_adam[kMaxGroups] = {};     // optimizer of group g done (side stream)
  bool adam_pending[kMaxGroups] = {};       // ... and not yet waited for by the caller's stream
  int n_groups = 1;
  int overlap = 0;           // 1: the join with the side stream is deferred to the next step / iqlb_agents_sync
  int n_sm = 0;
  int adam_ctas_per_sm = 2;  // residency bound of the optimizer grid when it runs under tensor-core kernels
  // pipelined host feed (iqlb_agents_step_host_pipelined): H2D of step k+1 on copy_stream under the update of step k
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_h2d[2] = {}, ev_done[2] = {};
  cudaStream_t d2h_stream = nullptr;  // loss read-back
  float* loss_snap[2] = {};           // per-parity device copy of the step's losses, written by the loss kernel
  float* loss_snap_cur = nullptr;
  cudaEvent_t ev_step[2] = {};  // step k enqueued on the caller's stream (the loss read-back waits for it)
  float* stage[2] = {};
  float* host_loss[2] = {};  // pinned
  long long pipe_k = 0;
  // L2 zig-zag: consecutive kernels walk the agents in opposite directions, so a consumer starts with what its
  // producer wrote last (still in L2); the direction flips every step (IQLB_ZIGZAG=0 turns it off)
  const float* batch_src = nullptr;  // explicit-batch step reads its rows from here instead of `batch` (pipelined host feed, precision 1)
  int zig = 0;
  // look-ahead sampling (precision 1): the Philox gather + packing of step t+1 runs on pre_stream under the
  // HBM-bound optimizer kernel of step t (its consumers of step t - loss, forward, weight gradients - are done and
  // the RNG offset has been advanced by then).  pre_valid: h->batch / the X images already hold the next batch.
  cudaStream_t pre_stream = nullptr;
  cudaEvent_t ev_pre_fork = nullptr, ev_pre_join = nullptr;
  bool pre_valid = false;
  int presample = 0;
};

namespace iqlb {
// Launch bookkeeping: when profiling is on, an event is recorded after every launch of the step.
struct Prof {
  iqlb_agents* h; cudaStream_t st; int i = 0;
  Prof(iqlb_agents* h_, cudaStream_t st_) : h(h_), st(st_) {
    if (h->prof) cudaEventRecord(h->ev[0], st);
  }
  void mark(const char* name) {
    if (h->prof && i < iqlb_agents::kMaxProf) {
      h->prof_name[i] = name;
      cudaEventRecord(h->ev[i + 1], st);
    }
    ++i;
  }
  void done() {
    if (h->prof) h->prof_n = i < iqlb_agents::kMaxProf ? i : iqlb_agents::kMaxProf;
    h->launches_last = i;
  }
};
}  // namespace iqlb
