/*
 * iqlb.h - C ABI of libiqlb.so, the B200 (sm_100a) implementation of the IQL hot path of
 * ml4ai/iqlpref: replay sampling, the fused IQL update over N independent agents, the Polyak
 * target update and the reward-model relabelling pass.
 *
 * The reference is pure Python/PyTorch and has no native boundary; each entry point below
 * names the reference interface (file:line under /root/reference) it replaces.  INTEGRATION.md
 * shows the ctypes binding a maintainer of the reference would add.
 *
 * Conventions: every function returns 0 on success and a negative iqlb_status on failure and
 * never throws; iqlb_last_error() gives the thread-local message.  All pointers are DEVICE
 * pointers unless the parameter name starts with `host_`.  `stream` is a cudaStream_t passed as
 * void* (NULL = legacy default stream).  Nothing synchronises the host unless stated.  Buffers
 * are borrowed for the duration of the call (or, for iqlb_agents_create, until destroy).
 */
#ifndef IQLB_H_
#define IQLB_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define IQLB_VERSION 100

typedef enum {
  IQLB_OK = 0,
  IQLB_ERR_INVALID = -1,   /* bad argument (the reference raises ValueError/RuntimeError) */
  IQLB_ERR_CUDA = -2,      /* CUDA runtime error, see iqlb_last_error() */
  IQLB_ERR_UNSUPPORTED = -3,
  IQLB_ERR_NOMEM = -4
} iqlb_status;

int iqlb_version(void);
const char* iqlb_last_error(void);

/* ---------------------------------------------------------------------------------------------
 * (1) Replay buffer.  Replaces ReplayBuffer.sample (algorithms/offline/iql.py:211-221):
 *     torch.randint (Philox4x32-10, ATen DistributionTemplates.h:66-90,319-345) + 5 gathers.
 *
 * The table is ONE packed fp32 matrix [n_rows, row_stride]; a row is
 *     [ obs(S) | act(A) | next_obs(S) | reward | done | zero pad ]   (row_stride % 4 == 0)
 * so that [obs|act] is the contiguous twin-Q input and a row moves as 128-bit vectors.
 * --------------------------------------------------------------------------------------------- */

/* idx_out[b] = Philox(seed, subsequence=b, offset).x % n_rows  (bit-exact with
 * torch.randint(0, n_rows, (batch,), device="cuda") for a generator at (seed, offset)). */
int iqlb_philox_randint(uint64_t seed, uint64_t offset, int64_t n_rows, int32_t batch,
                        int64_t* idx_out, void* stream);

/* For each agent a: draw `batch` indices from (rng[a][0]=seed, rng[a][1]=offset), copy the rows
 * to batch_out[a, b, :], optionally replacing the reward column by
 * reward_cols[reward_variant[a], idx] (sweeps whose agents differ only in the reward labels).
 * idx_out (nullable) receives the indices [n_agents, batch].  Does NOT advance rng. */
int iqlb_replay_sample(const float* table, int64_t n_rows, int32_t row_stride, int32_t reward_col,
                       const float* reward_cols, const int32_t* reward_variant,
                       const uint64_t* rng, int32_t n_agents, int32_t batch,
                       int64_t* idx_out, float* batch_out, void* stream);

/* Same gather with caller-supplied indices idx[n_agents, batch]. */
int iqlb_replay_gather(const float* table, int64_t n_rows, int32_t row_stride, int32_t reward_col,
                       const float* reward_cols, const int32_t* reward_variant,
                       const int64_t* idx, int32_t n_agents, int32_t batch,
                       float* batch_out, void* stream);

/* ---------------------------------------------------------------------------------------------
 * (2)-(4) The IQL update for N independent agents.  Replaces ImplicitQLearning.train
 * (iql.py:639-662): _update_v (:581-593), _update_q (:595-613) incl. soft_update (:127-129),
 * _update_policy (:615-637), three torch.optim.Adam steps and CosineAnnealingLR.step.
 *
 * Parameter packing (fp32, per agent, nn.Linear layout W[out,in] row-major then b[out]):
 *   Q1: W0[H,S+A] b0[H] W1[H,H] b1[H] W2[1,H] b2[1] | Q2: same | V: W0[H,S] ... W2[1,H] b2[1]
 *   | actor: W0[H,S] b0 W1 b1 W2[A,H] b2[A] | log_std[A] (Gaussian policy only)
 * Each network segment is zero-padded to a multiple of 8 floats (every network starts 32-byte
 * aligned).  iqlb_param_count() gives P (all, padding included) and PQ (Q1+Q2 segments = target size).
 * --------------------------------------------------------------------------------------------- */
typedef struct {
  int32_t n_agents;
  int32_t state_dim;
  int32_t action_dim;
  int32_t hidden_dim;    /* 2 hidden layers of this width (iql.py:458-459,519,537) */
  int32_t batch_size;
  int32_t deterministic; /* 1: DeterministicPolicy (iql.py:485-504), 0: GaussianPolicy (:452-474) */
  int32_t row_stride;    /* floats per packed replay row */
  int32_t precision;     /* 0: fp32 FFMA (parity mode, 1e-3 vs fp32 oracle); 1: bf16 tcgen05 */
} iqlb_agents_desc;

/* per-agent hyper-parameters, float32[n_agents][8]:
 *   0 discount  1 tau(polyak)  2 beta  3 iql_tau  4 vf_lr  5 qf_lr  6 actor_lr(base)  7 max_steps */
#define IQLB_HP_STRIDE 8

typedef struct iqlb_agents iqlb_agents;

int iqlb_param_count(int32_t state_dim, int32_t action_dim, int32_t hidden_dim, int32_t deterministic,
                     int64_t* p_all, int64_t* p_q);

/* All arrays are caller-owned device memory that must outlive the handle:
 *   params, exp_avg, exp_avg_sq : float[n_agents][P8]     target : float[n_agents][PQ8]
 *                                 (row strides P8 / PQ8 = P / PQ rounded up to a multiple of 8, padding zero)
 *   hparams : float[n_agents][8]
 *   it      : int64[n_agents]   (ImplicitQLearning.total_it == every Adam `step`)
 *   actor_lr: double[n_agents]  (current CosineAnnealingLR value, recursive form)
 *   rng     : uint64[n_agents][2] (seed, philox offset); offset advances by 4 per sampled step
 *   losses  : float[n_agents][3] (value_loss, q_loss, actor_loss of the last step)
 *   loss_sum: float[n_agents][3] (running sum; the host divides by log_freq and zeroes it) */
int iqlb_agents_create(const iqlb_agents_desc* desc, float* params, float* exp_avg, float* exp_avg_sq,
                       float* target, const float* hparams, int64_t* it, double* actor_lr,
                       uint64_t* rng, float* losses, float* loss_sum, iqlb_agents** out);
int iqlb_agents_destroy(iqlb_agents* h);

/* Attach the replay table the on-device sampler reads (see (1)).  idx_out nullable. */
int iqlb_agents_set_replay(iqlb_agents* h, const float* table, int64_t n_rows,
                           const float* reward_cols, const int32_t* reward_variant, int64_t* idx_out);

/* One update of every agent.  batch == NULL: sample on device from the attached table and
 * advance rng; otherwise use the packed rows batch[n_agents][batch_size][row_stride]. */
int iqlb_agents_step(iqlb_agents* h, const float* batch, void* stream);

/* Same, with the batch in (pinned) HOST memory: H2D copy of the rows, the update, D2H copy of
 * losses[n_agents][3] into host_losses, then a stream synchronise (the reference's .item()). */
int iqlb_agents_step_host(iqlb_agents* h, const float* host_batch, float* host_losses, void* stream);

/* Pipelined host feed for training loops that log with one step of lag (the reference averages losses over
 * log_freq = 250 steps, iql.py:1537-1544).  Call k copies host_batch (pinned) to the device on an internal copy
 * stream, enqueues update k on `stream`, and returns the losses[n_agents][3] of update k-1 (zeros for k = 0)
 * after waiting for that update only - so the H2D copy of step k+1 runs under update k.  host_batch of call k
 * must stay untouched until call k+1 has returned.  host_batch == NULL: no new update, returns the losses of
 * the last one (flush). */
int iqlb_agents_step_host_pipelined(iqlb_agents* h, const float* host_batch, float* host_losses_prev, void* stream);

/* Actor dropout (nn.Dropout(p) after every hidden ReLU of the policy MLP, iql.py:437-438, 459, 492; the adroit /
 * pen configs use p = 0.1).  Both precisions.  The keep mask of element (row m, unit n) of hidden layer l at Adam
 * step t of agent a is word (n & 3) of Philox4x32-10(key = rng[a].seed, subsequence = (l+1) << 32 | (m*H + n) / 4,
 * counter = t) >= p * 2^32; kept activations are scaled by 1/(1-p).  The reference draws its masks from torch's CUDA
 * generator, which cannot be mirrored: parity is tested by injecting this mask into the oracle. */
int iqlb_agents_set_actor_dropout(iqlb_agents* h, float p);

/* Batched policy inference for evaluation (GaussianPolicy.act / DeterministicPolicy.act in eval mode, iql.py:476-482,
 * 506-514, called from eval_actor :304-319): actions[a][i][:] = clamp(max_action * tanh(actor_a(states_a[i])), +-max_action)
 * for every agent, fp32.  states: device, agent a reads n_rows x state_dim floats at states + a * agent_stride
 * (agent_stride 0: all agents evaluate the same rows); actions: device [n_agents][n_rows][action_dim]. */
int iqlb_agents_act(iqlb_agents* h, const float* states, int64_t agent_stride, int32_t n_rows, float max_action,
                    float* actions, void* stream);

/* Tell the library that the caller wrote params/target/it/rng directly (adopting modules, loading a
 * checkpoint, reseeding): precision 1 rebuilds the bf16 operand images of the weights and drops the
 * look-ahead batch; precision 0: no-op.
 * Look-ahead sampling (precision 1, on-device sampling, no idx_out; opt-in with IQLB_PRESAMPLE=1): the
 * gather of step t+1 is enqueued on an internal stream under the optimizer kernel of step t, after the
 * RNG offset of step t has been advanced, so the index stream is unchanged; an explicit-batch step,
 * iqlb_agents_set_replay or this call discards it. */
int iqlb_agents_refresh(iqlb_agents* h, void* stream);

/* Optimizer overlap (precision 1, opt-in with IQLB_GROUPS=2; with the default single agent group both calls are no-ops -
 * on B200 the overlapped schedule measured slower than the serial one).  The HBM-bound Adam +
 * Polyak kernel of agent group g runs on an internal low-priority stream, as a residency-bounded persistent grid, UNDER the
 * latency-bound tensor-core kernels of the next group on the caller's stream: agents are independent (the reference runs
 * one OS process per agent, tr_sweeps/launch.sh:82-92), so group g's update t+1 needs only group g's optimizer of update t.
 * Default (overlap 0): every step call ends with the caller's stream waiting for the internal one - ordinary stream
 * semantics.  iqlb_agents_set_overlap(h, 1, stream) defers that join to the next step, so the last group's optimizer also
 * overlaps the first group's next update; the caller must then call iqlb_agents_sync(h, stream) before reading or writing
 * params / exp_avg / exp_avg_sq / target on `stream` (losses, loss_sum, it, actor_lr and rng are written on the caller's
 * stream and need no sync; refresh, act, debug_copy and destroy join by themselves).  Arithmetic is unchanged. */
int iqlb_agents_set_overlap(iqlb_agents* h, int32_t enable, void* stream);
int iqlb_agents_sync(iqlb_agents* h, void* stream);

/* Number of kernels one iqlb_agents_step enqueues (for launch accounting). */
int iqlb_agents_launches_per_step(const iqlb_agents* h, int32_t sampled);

/* Per-kernel timing of the step with CUDA events on the launching stream (bench.py's roofline leg).
 * profile(h,1) zeroes the totals and makes every later step record one event per launch;
 * profile_collect(h) waits for the last recorded step and adds its per-launch times to the totals
 * (call it after each profiled step); profile_read returns kernel i's name, total ms and the number
 * of collected steps (error once i >= number of launches in a step). */
int iqlb_agents_profile(iqlb_agents* h, int32_t enable);
int iqlb_agents_profile_collect(iqlb_agents* h);
int iqlb_agents_profile_read(iqlb_agents* h, int32_t i, const char** host_name, double* host_ms_total,
                             int32_t* host_steps);

/* Debug/parity switch.  option 0: precision 1 runs Adam + Polyak inside the weight-gradient kernel
 * (value 1, default) or materialises fp32 gradients and runs the stand-alone Adam kernel (value 0,
 * so that iqlb_agents_debug_copy(which = 6) can be compared with the fp32 path). */
int iqlb_agents_debug_option(iqlb_agents* h, int32_t option, int32_t value);

/* Debug/parity hook: copy an internal scratch tensor to dst (device).  which: 0 H1, 1 H2,
 * 2 head_out, 3 d_out, 4 dH2, 5 dH1, 6 grads, 7 batch.  Returns bytes via *nbytes if dst NULL. */
int iqlb_agents_debug_copy(iqlb_agents* h, int32_t which, void* dst, int64_t* nbytes, void* stream);

/* (4) Stand-alone soft_update (iql.py:127-129): target += tau * (source - target), n floats. */
int iqlb_polyak(float* target, const float* source, int64_t n, float tau, void* stream);

/* ---------------------------------------------------------------------------------------------
 * (5) Relabelling.  Replaces the reward-model forward + reduction inside
 * qlearning_dataset_mr (iql.py:719-724), qlearning_dataset_bnn (:982-1011),
 * qlearning_dataset_mr_ensemble (:1162-1187) and qlearning_dataset_pt (:1259-1301).
 * --------------------------------------------------------------------------------------------- */

/* activation ids of RewardMLP (optbnn/bnn/nets/mlp.py:45-48) */
enum { IQLB_ACT_RELU = 0, IQLB_ACT_TANH, IQLB_ACT_COS, IQLB_ACT_SIN, IQLB_ACT_LINEAR, IQLB_ACT_RBF,
       IQLB_ACT_SWISH, IQLB_ACT_SOFTPLUS, IQLB_ACT_LEAKY_RELU };

/* Ensemble of S reward MLPs (scaled-variance Linear, optbnn/bnn/layers/linear.py:43-47) over the
 * rows x[n_rows][d_in] (row stride ldx floats), reduced per row to the mean of the n_tail
 * smallest member predictions (n_tail == S: plain mean; S == 1: the single model).
 *   weights : float[S][PM], member layout W0[d_in][w0] b0[w0] W1[w0][w1] b1 ... Wout[w_last][1] bout[1]
 *             (W stored (n_in, n_out) exactly as the reference's state_dict / BNN tuples)
 *   widths  : host int32[depth] hidden widths
 *   out     : float[n_rows]
 *   preds_out (nullable): float[S][n_rows] all member predictions (the reference's all_preds)
 *   sel_out (nullable)  : uint8[S][n_rows] 1 where member s is in the tail set of row i */
int iqlb_relabel_mlp(const float* x, int64_t n_rows, int32_t ldx, int32_t d_in,
                     const float* weights, int32_t n_members, const int32_t* host_widths, int32_t depth,
                     int32_t activation, int32_t n_tail, float* out, float* preds_out, uint8_t* sel_out,
                     void* stream);

/* Same with an explicit arithmetic mode for the hidden layers:
 *   IQLB_RELABEL_FP32       fp32 FFMA GEMMs (any shape; the 1e-6 parity mode)
 *   IQLB_RELABEL_SPLIT_BF16 tcgen05 tensor cores, every product as x_hi*W_hi + x_lo*W_hi + x_hi*W_lo with fp32
 *                           accumulation (~1e-5 relative per layer); d_in <= 128, equal hidden widths of 64 or
 *                           128, depth <= 3 - IQLB_ERR_UNSUPPORTED otherwise
 *   IQLB_RELABEL_AUTO       split-bf16 where it applies, fp32 elsewhere (what iqlb_relabel_mlp does)
 * The member selection (CVaR tail) is exact given the predictions in every mode.
 * mean_out (nullable): float[n_rows], the plain mean over all S members of every row, from the same pass (the
 * "ensemble mean" diagnostic the reference prints next to the CVaR reward, iql.py:1013-1021, 1196-1204).  With
 * preds_out == NULL the member predictions only ever exist as a row-chunked scratch of at most 512 MB. */
enum { IQLB_RELABEL_AUTO = -1, IQLB_RELABEL_FP32 = 0, IQLB_RELABEL_SPLIT_BF16 = 1 };
int iqlb_relabel_mlp_p(const float* x, int64_t n_rows, int32_t ldx, int32_t d_in,
                       const float* weights, int32_t n_members, const int32_t* host_widths, int32_t depth,
                       int32_t activation, int32_t n_tail, int32_t precision, float* out, float* preds_out,
                       uint8_t* sel_out, float* mean_out, void* stream);

/* ---------------------------------------------------------------------------------------------
 * (6) Dataset preparation on the device (between the relabel and the replay buffer).
 * --------------------------------------------------------------------------------------------- */

/* return_reward_range (iql.py:344-360): episodes end at a terminal or after max_episode_steps rows; returns are
 * summed in double in the reference's order.  min_max (device, 2 doubles) = min / max over the COMPLETED episodes;
 * trj_lens (device, nullable, float[n]) = length of the episode each row belongs to (the trailing unfinished
 * episode reports its current length, as the reference does).  All pointers are device pointers. */
int iqlb_reward_range(const float* rewards, const uint8_t* terminals, int64_t n, int32_t max_episode_steps,
                      double* min_max, float* trj_lens, void* stream);

/* Allocation-free variant of iqlb_reward_range.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same arguments, same result bit for bit; the scan scratch lives in a static device array of the library
 * (256 KB on every GPU that loads it, held until exit) instead of the stream-ordered pool, so the call makes no
 * cudaMalloc / cudaFree of any kind.  n > IQLB_REWARD_RANGE_STATIC_MAX_ROWS is rejected with IQLB_ERR_INVALID.  All calls on a
 * device share the array: calls that overlap on the GPU (different streams) race - serialise them. */
#define IQLB_REWARD_RANGE_STATIC_MAX_ROWS 16777216
int iqlb_reward_range_static(const float* rewards, const uint8_t* terminals, int64_t n, int32_t max_episode_steps,
                             double* min_max, float* trj_lens, void* stream);

/* modify_reward (iql.py:363-401), in place on rewards[n] with numpy's op sequence and dtypes.
 * env_kind 0: halfcheetah / hopper / walker2d; 1: antmaze with normalize_reward 0..7 (other environments are not
 * touched by the reference: do not call). */
int iqlb_modify_reward(float* rewards, const uint8_t* terminals, int64_t n, int32_t env_kind,
                       int32_t normalize_reward, int32_t max_episode_steps, void* stream);

/* compute_mean_std (iql.py:132-135): mean[dim], std[dim] = population std + eps over states[n][dim] (row stride
 * ld floats), accumulated in double.  normalize_states (iql.py:138-139): in place (x - mean) / std. */
int iqlb_state_mean_std(const float* states, int64_t n, int32_t dim, int32_t ld, float eps, float* mean, float* std,
                        void* stream);
int iqlb_normalize_states(float* states, int64_t n, int32_t dim, int32_t ld, const float* mean, const float* std,
                          void* stream);

/* Debug: CTA 0 of the following tensor-core relabel launches records clock64 stamps into buf (device memory,
 * >= 5*512 uint64; scripts/relabel_timeline.py decodes them).  NULL switches the recording off. */
int iqlb_relabel_debug_timeline(unsigned long long* buf);

/* Preference transformer (optbnn/bnn/nets/pref_trans.py:230-325) evaluated on the windows that
 * qlearning_dataset_pt builds (iql.py:1264-1292) and reduced to value[:,0,-1,0].
 *   obs[n_data][S], act[n_data][A]      : the dataset
 *   win_start[n], seq_len[n] (int32)    : window i covers rows win_start[i] .. +seq_len[i]-1,
 *                                         left-padded to query_len; timesteps 0..seq_len-1
 *   weights: packed fp32, layout given by iqlb_pt_param_count / DESIGN.md
 *   out[n] */
typedef struct {
  int32_t state_dim, action_dim, embd_dim, num_heads, intermediate_dim, num_layers;
  int32_t pref_dim;        /* pref_attn_embd_dim P: pref_linear is E -> 2P+1 = [query | key | value] */
  int32_t n_timesteps;     /* rows of timestep_embed = max_episode_steps + 1 */
  int32_t query_len;
  float eps;               /* LayerNorm eps (model_eps) */
} iqlb_pt_desc;

int iqlb_pt_param_count(const iqlb_pt_desc* d, int64_t* n_floats);
int iqlb_relabel_pt(const iqlb_pt_desc* d, const float* weights, const float* obs, const float* act,
                    int64_t n_data, const int32_t* win_start, const int32_t* seq_len, int64_t n,
                    float* out, void* stream);

/* RewardPT.forward on explicit windows (pref_trans.py:269-325; what the relabel loop calls at
 * iql.py:1294-1299): states[n][T][S], actions[n][T][A], timesteps int32[n][T], attn_mask[n][T]
 * -> value_out[n][T] (= result["value"][:, 0, :, 0]) and, when weighted_sum_out is not NULL, weighted_sum_out[n][T]
 * (= result["weighted_sum"][:, :, 0]: the preference-attention head, pref_trans.py:303-325 - what PTTrainer's
 * Bradley-Terry loss pools, training.py:348); T = d->query_len. */
int iqlb_pt_forward(const iqlb_pt_desc* d, const float* weights, const float* states, const float* actions,
                    const int32_t* timesteps, const float* attn_mask, int64_t n, float* value_out,
                    float* weighted_sum_out, void* stream);

/* ---------------------------------------------------------------------------------------------
 * MR preference-model training step: replaces MRTrainer.train / .evaluation
 * (gp_reward-priors/optbnn/training/training.py:27-72, 74-114) with bt_pool_logit
 * (optbnn/utils/util.py:356-364), RewardMLP (optbnn/bnn/nets/mlp.py:73-111) and
 * torch.optim.Adam(net.parameters(), lr) (scripts_mr/run_mr_training.py:105).
 *   params / exp_avg / exp_avg_sq : device float[PM] in the member layout of iqlb_relabel_mlp
 *       (W0[d_in][w0] b0[w0] ... Wout[w_last][1] bout[1]); owned by the caller, updated in place
 *   states{1,2} [B][T][S], actions{1,2} [B][T][A], mask{1,2} [B][T] (0/1), labels [B][2] : device fp32
 *   out2 (device float[2], nullable): mean soft-label cross entropy, accuracy (argmax of pooled
 *       logits vs argmax of labels)
 *   grad_out (nullable): device float[PM], the gradient the step used
 *   adam_step : 1 for the first step (the value of Adam's `step` after this update)
 *   pool_mean : 1 = bt_pool "mean" (masked mean), 0 = "sum"
 * A call with lr == 0 and exp_avg == NULL computes loss and gradients only.  No host sync, except on
 * the first call for a new row count (the weight-gradient task table is rebuilt).
 * --------------------------------------------------------------------------------------------- */
typedef struct iqlb_mr_trainer iqlb_mr_trainer;
int iqlb_mr_create(int32_t d_in, const int32_t* host_widths, int32_t depth, int32_t activation,
                   int32_t pool_mean, int64_t max_rows /* 2*B*T */, iqlb_mr_trainer** out);
int iqlb_mr_destroy(iqlb_mr_trainer* h);
/* Arithmetic of the step's GEMMs (forward, data gradients, weight-gradient partials): 0 = fp32 FFMA (default), 1 = the
 * tensor cores in split-bf16 (csrc/gemm_tc.cu: x_hi*w_hi + x_lo*w_hi + x_hi*w_lo, fp32 accumulation, ~1e-5 relative). */
int iqlb_mr_set_precision(iqlb_mr_trainer* h, int32_t precision);
int iqlb_mr_param_count(const iqlb_mr_trainer* h, int64_t* n_params);
int iqlb_mr_train_step(iqlb_mr_trainer* h, float* params, float* exp_avg, float* exp_avg_sq,
                       const float* states1, const float* actions1, const float* mask1,
                       const float* states2, const float* actions2, const float* mask2,
                       const float* labels, int32_t B, int32_t T, int32_t S, int32_t A, float lr,
                       float beta1, float beta2, float eps, int64_t adam_step, float* out2,
                       float* grad_out, void* stream);
/* MRTrainerF.train (training.py:166-262): the step above with the functional GP prior of an LCF model
 * (optbnn/gp/models/model.py:202-513: f = Phi(x)^T w, w ~ N(p_mean, Sigma)) in place of a weight prior.  The gradient the
 * step descends is dCE/dw + J_w(X_M)^T alpha / N with alpha = (Phi Sigma Phi^T + eps I)^-1 (f(X_M) - Phi p_mean), solved
 * through the d x d Woodbury system in fp64 (model.py:262-325: eps = jitter * max(trace(Sigma Phi^T Phi) / n_meas, 1)).
 *   x_meas [n_meas][d_in] device fp32 : the measurement points; they go through the step's GEMMs as extra rows, so
 *       iqlb_mr_create's max_rows must cover 2*B*T + n_meas
 *   phi [n_meas][n_concepts] device fp64 : Phi(X_M), evaluated by the caller (the reference's `function_vect` is user code)
 *   p_mean [d], sigma [d][d], sigma_inv [d][d] device fp64, logdet_sigma = log|Sigma| : the prior's constants; d <= 64
 *   num_datapoints : N of training.py:238;  jitter : relative nugget (1e-6 in the reference)
 *   prior_out (nullable) device double[2] : log p_GP(f(X_M)) evaluated with the UPDATED weights (training.py:242-250; the
 *       reference logs training_loss = out2[0] - prior_out[0] / N) and eps
 * A covariance that is not positive definite gives NaN where the reference's torch.linalg.cholesky raises. */
int iqlb_mr_train_step_f(iqlb_mr_trainer* h, float* params, float* exp_avg, float* exp_avg_sq,
                         const float* states1, const float* actions1, const float* mask1,
                         const float* states2, const float* actions2, const float* mask2,
                         const float* labels, int32_t B, int32_t T, int32_t S, int32_t A, float lr,
                         float beta1, float beta2, float eps, int64_t adam_step, const float* x_meas,
                         const double* phi, int32_t n_meas, int32_t n_concepts, const double* p_mean,
                         const double* sigma, const double* sigma_inv, double logdet_sigma, double jitter,
                         double num_datapoints, float* out2, double* prior_out, float* grad_out, void* stream);
/* MRTrainer.evaluation: forward + loss/accuracy only; pred_out (nullable) float[2*B*T] per-row rewards */
int iqlb_mr_eval(iqlb_mr_trainer* h, const float* params, const float* states1, const float* actions1,
                 const float* mask1, const float* states2, const float* actions2, const float* mask2,
                 const float* labels, int32_t B, int32_t T, int32_t S, int32_t A, float* out2,
                 float* pred_out, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Preference-transformer training step (the PT half): replaces PTTrainer.train / .evaluation
 * (gp_reward-priors/optbnn/training/training.py:308-416) with PT.forward in training mode
 * (optbnn/bnn/nets/pref_trans.py:269-325 and the GPT2 blocks :69-227), bt_pool_logit on `weighted_sum`
 * (optbnn/utils/util.py:356-364) and torch.optim.Adam(net.parameters(), lr) (scripts_pt/run_pt_training.py:137).
 *   params / exp_avg / exp_avg_sq : device float[iqlb_pt_train_param_count], the parameters in net.parameters() order
 *       (state_linear W b | action_linear W b | timestep_embed | stacked LN g b | per block: LN0 g b, attention in W b,
 *       attention out W b, LN1 g b, mlp in W b, mlp out W b | final LN g b | pref_linear W b); caller-owned, updated in place
 *   states{1,2} [B][T][S], actions{1,2} [B][T][A], timesteps{1,2} int32 [B][T] (< d->n_timesteps), mask{1,2} [B][T],
 *   labels [B][2] : device; T = d->query_len
 *   out2 (device float[2], nullable): mean soft-label cross entropy, accuracy;  grad_out (nullable): the gradient used
 *   adam_step : 1 for the first step.  lr == 0 with exp_avg == NULL computes loss and gradients only.
 *   training != 0 applies the dropouts set by iqlb_pt_trainer_set_dropout.  The reference draws its dropout masks from
 *   torch's generator, which cannot be mirrored; here element e (row-major index in the tensor the mask multiplies:
 *   [2B][2T][E] for the embedding / residual sites, [2B][H][2T][2T] for the attention probabilities) of site s
 *   (0 = embd_dropout; for block l: 1 + 4l = attention probabilities, 2 + 4l = after attention.out_linear,
 *   3 + 4l = after mlp.out_linear) at Adam step k is KEPT iff word (e & 3) of
 *   Philox4x32-10(key = seed, counter = {lo32(k), hi32(k), lo32(e >> 2), (s << 8) | hi32(e >> 2)}) >= floor(p * 2^32),
 *   kept values scaled by 1 / (1 - p).
 * No host synchronisation, except on the first call for a new batch size (the weight-gradient task table is rebuilt).
 * --------------------------------------------------------------------------------------------- */
typedef struct iqlb_pt_trainer iqlb_pt_trainer;
int iqlb_pt_train_param_count(const iqlb_pt_desc* d, int64_t* n_floats);
int iqlb_pt_trainer_create(const iqlb_pt_desc* d, int32_t pool_mean, int32_t max_pairs, iqlb_pt_trainer** out);
int iqlb_pt_trainer_destroy(iqlb_pt_trainer* h);
/* 0 = fp32 FFMA GEMMs (default), 1 = every Linear (forward, data and weight gradients) on the tensor cores in split-bf16 */
int iqlb_pt_trainer_set_precision(iqlb_pt_trainer* h, int32_t precision);
int iqlb_pt_trainer_set_dropout(iqlb_pt_trainer* h, float p_embd, float p_attn, float p_resid_attn, float p_resid_mlp,
                                uint64_t seed);
int iqlb_pt_train_step(iqlb_pt_trainer* h, float* params, float* exp_avg, float* exp_avg_sq, const float* states1,
                       const float* actions1, const int32_t* timesteps1, const float* mask1, const float* states2,
                       const float* actions2, const int32_t* timesteps2, const float* mask2, const float* labels,
                       int32_t B, float lr, float beta1, float beta2, float eps, int64_t adam_step, int32_t training,
                       float* out2, float* grad_out, void* stream);
/* PTTrainer.evaluation: eval-mode forward + loss / accuracy; wsum_out (nullable) float[2B][T] = weighted_sum */
int iqlb_pt_eval(iqlb_pt_trainer* h, const float* params, const float* states1, const float* actions1,
                 const int32_t* timesteps1, const float* mask1, const float* states2, const float* actions2,
                 const int32_t* timesteps2, const float* mask2, const float* labels, int32_t B, float* out2,
                 float* wsum_out, void* stream);

/* Diagnostic: one CTA computes D[128][n] (fp32, row-major) = A * B with tcgen05.mma from two bf16
 * operand images (128-byte-swizzled column blocks of 64, see csrc/umma.cuh).  a_mn / b_mn select
 * MN-major (1: image columns are the M/N index, starting at column a_off) or K-major (0: image rows
 * are the M/N index, starting at row a_off); *_rows_total = rows of the whole image.  Used by
 * tests/test_gpu_umma.py to pin the descriptors against numpy for every arrangement the step uses. */
int iqlb_selftest_umma(const void* a_img, int64_t a_bytes, const void* b_img, int64_t b_bytes, int32_t a_mn,
                       int32_t a_rows_total, int32_t a_off, int32_t b_mn, int32_t b_rows_total, int32_t b_off,
                       int32_t n, int32_t ksteps, float* d_out, void* stream);

/* Diagnostic: one fp32 GEMM through the library's dispatcher.  kn_layout 0: C = act(A W^T + bias) (+ residual), W [n][k]
 * (nn.Linear); 1: C = act(scale * A W + bias), W [k][n] (RewardMLP's scaled Linear).  allow_tc 1 lets shapes of at least
 * 128 x 64 x 16 run on the tensor cores in split-bf16 arithmetic (csrc/gemm_tc.cu; ~1e-5 relative), 0 forces fp32 FFMA. */
int iqlb_selftest_gemm(int32_t kn_layout, const float* a, int32_t lda, const float* w, int32_t ldw, const float* bias,
                       const float* residual, float* c, int32_t ldc, int32_t m, int32_t n, int32_t k, int32_t activation,
                       float scale, int32_t allow_tc, void* stream);

/* Tensor-pipe issue-rate probe: `count` back-to-back 128 x n x 16 bf16 MMAs into one accumulator, A from shared
 * memory (a_from_tmem = 0) or from tensor memory (1).  clocks_out (device, 2 x uint64): SM clocks spent issuing,
 * and from the first issue to the completion of the last MMA. */
int iqlb_selftest_mma_rate(int32_t a_from_tmem, int32_t n, int32_t count, unsigned long long* clocks_out, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Host helper (plain C loop, no GPU): the keep-mask / episode-step scan every qlearning_dataset_*
 * starts with (iql.py:701-716, 1236-1253).  terminals/timeouts are host uint8[n]; timeouts may be
 * NULL (then `final` is ep == max_episode_steps - 1).  Writes host_keep[n-1], host_ep_steps[n-1]
 * (either may be NULL).
 * --------------------------------------------------------------------------------------------- */
int iqlb_keep_mask_scan(const uint8_t* host_terminals, const uint8_t* host_timeouts, int64_t n,
                        int32_t max_episode_steps, int32_t terminate_on_end, uint8_t* host_keep,
                        int32_t* host_ep_steps);

#ifdef __cplusplus
}
#endif
#endif /* IQLB_H_ */
