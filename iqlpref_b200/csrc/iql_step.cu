// The IQL update for N independent agents (fp32 parity path + orchestration).
// Replaces ImplicitQLearning.train (/root/reference/algorithms/offline/iql.py:639-662).
//
// Launch sequence of one step (every launch covers ALL agents):
//   gather -> fwd L0 (7 streams) -> fwd L1 -> heads fwd -> losses + output grads + bookkeeping
//   -> heads bwd -> bwd L1 weights -> bwd L1 data -> bwd L0 weights -> Adam + Polyak
// Data dependences follow SURVEY.md 3.2: every forward reads start-of-step parameters; only the
// Polyak update must see the post-Adam Q parameters, so it is fused into the Adam kernel.
#include <new>
#include <stdlib.h>

#include "iql_step.cuh"

namespace iqlb {

int launch_gather(const float* table, int64_t n_rows, int row_stride, int reward_col, const float* reward_cols,
                  const int32_t* reward_variant, const uint64_t* rng, const int64_t* idx_in, int n_agents, int batch,
                  int64_t* idx_out, float* batch_out, cudaStream_t stream);
int bf16_create(iqlb_agents* h);
void bf16_destroy(iqlb_agents* h);
int bf16_step_gemms_fwd(iqlb_agents* h, int a0, int n, Prof& pf, cudaStream_t stream, bool packed);
int bf16_gather_pack(iqlb_agents* h, int a0, int n, Prof& pf, cudaStream_t stream);
int bf16_step_gemms_bwd(iqlb_agents* h, int a0, int n, Prof& pf, cudaStream_t stream);
const BfParams* bf16_params(const iqlb_agents* h);
void bf16_mark_dirty(iqlb_agents* h);
int bf16_refresh(iqlb_agents* h, cudaStream_t stream);

// ------------------------------------------------------------------------------------------
// Heads forward: out[o] = H2[b,:] . W2[o,:] + b2[o]   (N = 1 for V/Q, N = A (+tanh) for actor)
// One warp per (stream, batch row); lanes stride the hidden dimension, warp-shuffle reduce.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) head_fwd_kernel(Layout L, const float* __restrict__ params,
                                                       const float* __restrict__ target,
                                                       const float* __restrict__ H2, float* __restrict__ head_out) {
  const int agent = blockIdx.z, s = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x * 8 + warp;
  if (b >= L.B) return;
  const int in = L.in_dim(s), nout = L.out_dim(s);
  const float* base = L.uses_target(s) ? target + (int64_t)agent * L.PQstride + L.net_off(s)
                                       : params + (int64_t)agent * L.Pstride + L.net_off(s);
  const float* W2 = base + L.w2(in);
  const float* b2 = base + L.b2(in, nout);
  const float* h = H2 + (((int64_t)agent * kStreams + s) * L.B + b) * L.H;
  float* ho = head_out + (int64_t)agent * L.head_rows() * L.B;
  for (int o = 0; o < nout; ++o) {
    float acc = 0.f;
    for (int j = lane; j < L.H; j += 32) acc = fmaf(h[j], __ldg(W2 + (int64_t)o * L.H + j), acc);
    acc = warp_sum(acc);
    if (lane == 0) {
      float v = acc + __ldg(b2 + o);
      if (s == 6) ho[(int64_t)(6 + o) * L.B + b] = tanhf(v);
      else ho[(int64_t)s * L.B + b] = v;
    }
  }
}

// ------------------------------------------------------------------------------------------
// Losses, output gradients and per-agent bookkeeping.  One CTA per agent, one thread per row.
//   value_loss (iql.py:404-405,587-588), q_loss (:604-606), actor_loss (:622-632)
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ double cosine_next(double lr, double base_lr, long long epoch, long long t_max) {
  // cospi(x) = cos(pi x) without the rounding of pi * x; 32-bit remainder when the counters fit (they do: max_steps
  // and the step counter are < 2^31 in every reference config), 64-bit otherwise
  bool restart;
  if (epoch < 0x7fffffffll && t_max < 0x3fffffffll && epoch - 1 - t_max >= 0)
    restart = (uint32_t)(epoch - 1 - t_max) % (2u * (uint32_t)t_max) == 0u;
  else
    restart = (epoch - 1 - t_max) % (2 * t_max) == 0;
  const double inv_t = 1.0 / (double)t_max;
  if (restart) return lr + base_lr * (1.0 - cospi(inv_t)) / 2.0;
  return (1.0 + cospi((double)epoch * inv_t)) / (1.0 + cospi((double)(epoch - 1) * inv_t)) * lr;
}

template <int kMaxThreads>
__global__ void __launch_bounds__(kMaxThreads) loss_kernel(Layout L, const float* __restrict__ params, const float* __restrict__ hparams,
                            const float* __restrict__ batch, const float* __restrict__ head_out,
                            float* __restrict__ d_out, float* __restrict__ grads, float* __restrict__ losses,
                            float* __restrict__ loss_sum, int64_t* __restrict__ it, double* __restrict__ actor_lr,
                            uint64_t* __restrict__ rng, StepScalars* __restrict__ scal, int advance_rng,
                            uint8_t* __restrict__ dzimg, int64_t z_bytes, int agent0, unsigned long long* dbg,
                            __nv_bfloat16* __restrict__ gradh, float* __restrict__ losses_snap) {
#ifdef IQLB_NO_TIMELINE
#define IQLB_TSL(slot) do { } while (0)
#else
#define IQLB_TSL(slot) do { if (dbg && blockIdx.x == 0) dbg[3072 + (slot)] = (unsigned long long)clock64(); } while (0)
#endif
  if (threadIdx.x == 0) IQLB_TSL(0);
  // per-warp partial sums of every reduced quantity: 0 vl | 1 ql | 2 al | 3..5 head bias grads (V, Q1, Q2)
  // | 6.. actor head bias grads [A] | 6+A.. log_std grads [A]; one block-wide pass at the end
  constexpr int kNV = 6 + 2 * kOutMax;
  __shared__ float red[32][kNV];
  const int agent = agent0 + blockIdx.x, b = threadIdx.x;
  const int lane = b & 31, warp = b >> 5, nw = (blockDim.x + 31) >> 5;
  const bool live = b < L.B;
  const bool bf = dzimg != nullptr;
  const float* hp = hparams + agent * IQLB_HP_STRIDE;
  const float discount = hp[0], beta = hp[2], iql_tau = hp[3];
  // per-agent bookkeeping (fp64 pow / cos: a long dependent chain).  When the launch carries an extra warp
  // beyond the batch rows, its first lane does this at once, under the loss arithmetic of the other warps.
  const bool extra_warp = (int)blockDim.x >= ((L.B + 31) / 32) * 32 + 32;
  auto bookkeeping = [&]() {
    IQLB_TSL(5);
    // Adam step scalars of THIS step (torch optim/adam.py:529-547) and scheduler/rng advance.
    const long long t = it[agent] + 1;
    // beta^t as exp(t ln beta): one fp64 exp instead of a pow (a third of the dependent chain); the result differs
    // from Python's beta ** t by ~1e-15 relative, far below the fp32 step sizes it ends up in
    const double bc1 = 1.0 - exp((double)t * -0.10536051565782628234);   // ln 0.9
    const double bc2 = 1.0 - exp((double)t * -0.00100050033358353350);   // ln 0.999
    const double lr_a = actor_lr[agent];
    StepScalars sc;
    sc.ss_v = (float)((double)hp[4] / bc1);
    sc.ss_q = (float)((double)hp[5] / bc1);
    sc.ss_a = (float)(lr_a / bc1);
    sc.bc2_sqrt = (float)sqrt(bc2);
    sc.tau = hp[1];
    scal[agent] = sc;
    it[agent] = t;
    actor_lr[agent] = cosine_next(lr_a, (double)hp[6], t, (long long)hp[7]);
    if (advance_rng & 1) rng[2 * agent + 1] += 4;  // one torch.randint of <= 256*grid draws (ATen :50-63)
    IQLB_TSL(6);
  };
  if (extra_warp && b == (int)blockDim.x - 32) bookkeeping();
  const float* ho = head_out + (int64_t)agent * L.head_rows() * L.B;
  float* dout = d_out + (int64_t)agent * L.dout_rows() * L.B;
  const int bb = live ? b : 0;
  const float* row = batch + ((int64_t)agent * L.B + bb) * L.R;
  const float invB = 1.0f / (float)L.B;
  const float* log_std = params + (int64_t)agent * L.Pstride + L.off_logstd;

  // every global load of this thread is issued before the first use; what the forward kernel (the PDL
  // predecessor) wrote - head_out - is loaded after the wait, batch rows / log_std / bookkeeping before it
  const float rew = __ldg(row + 2 * L.S + L.A), done = __ldg(row + 2 * L.S + L.A + 1);
  float mu[kOutMax], act[kOutMax], lsr[kOutMax];
  // (unconditional loads on clamped indices, through volatile asm: the compiler would otherwise sink each
  //  predicated load to its first use inside the action loop and pay one memory latency per action)
  auto ldg = [](const float* p) { float x; asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(x) : "l"(p)); return x; };
  auto ldv = [](const float* p) { float x; asm volatile("ld.global.f32 %0, [%1];" : "=f"(x) : "l"(p)); return x; };
  const int amax = L.A - 1;
#pragma unroll
  for (int a = 0; a < kOutMax; ++a) {
    if (a < 8 || a < L.A) {  // uniform; the first eight always (antmaze / halfcheetah never branch)
      act[a] = ldg(row + L.S + (a < amax ? a : amax));
      lsr[a] = L.deterministic ? 0.f : ldg(log_std + (a < amax ? a : amax));
    } else {
      act[a] = 0.f; lsr[a] = 0.f;
    }
  }
  pdl_wait();
  pdl_launch();
  const float nv = ldv(ho + 0 * L.B + bb), v = ldv(ho + 1 * L.B + bb), tq1 = ldv(ho + 2 * L.B + bb), tq2 = ldv(ho + 3 * L.B + bb);
  const float q1 = ldv(ho + 4 * L.B + bb), q2 = ldv(ho + 5 * L.B + bb);
#pragma unroll
  for (int a = 0; a < kOutMax; ++a) {
    if (a < 8 || a < L.A) mu[a] = ldv(ho + (int64_t)(6 + (a < amax ? a : amax)) * L.B + bb);
    else mu[a] = 0.f;
  }

  const float tq = fminf(tq1, tq2);
  const float adv = tq - v;
  if (threadIdx.x == 0) IQLB_TSL(1);
  const float wt = adv < 0.f ? fabsf(iql_tau - 1.0f) : fabsf(iql_tau);
  const float y = rew + (1.0f - done) * discount * nv;
  const float e1 = q1 - y, e2 = q2 - y;
  const float w = fminf(expf(beta * adv), 100.0f);
  const float d0 = live ? -2.0f * wt * adv * invB : 0.f;
  const float d1 = live ? e1 * invB : 0.f, d2 = live ? e2 * invB : 0.f;
  float vl = live ? wt * (adv * adv) : 0.f;
  float ql = live ? e1 * e1 + e2 * e2 : 0.f;
  if (live) {
    dout[0 * L.B + b] = d0;
    dout[1 * L.B + b] = d1;
    dout[2 * L.B + b] = d2;
  }
  // bf16 path: the dz image (dz columns + a column of ones) feeds the weight-gradient MMAs, and the
  // head bias gradients are reduced here (the fp32 path does both in head_bwd_kernel)
  if (bf && live) {
    for (int t = 0; t < 3; ++t)  // columns 0 (dz) and 1 (ones) share one 32-bit word of the image
      *reinterpret_cast<uint32_t*>(dzimg + ((int64_t)agent * kTrain + t) * z_bytes + umma::image_off(L.B, b, 0)) =
          umma::pack_bf16(t == 0 ? d0 : (t == 1 ? d1 : d2), 1.0f);
  }
  float bc = 0.f;
  // phase A: per-action terms in registers (no cross-lane traffic); phase B: warp sums of eight actions at a
  // time, level by level, so the 16 shuffle chains of a group overlap instead of running one after another
  float dzs[kOutMax], gss[kOutMax];
#pragma unroll
  for (int a = 0; a < kOutMax; ++a) {
    float gls = 0.f, dza = 0.f;
    if (live && a < L.A) {
      const float diff = mu[a] - act[a];
      float dmu;
      if (L.deterministic) {
        bc += diff * diff;
        dmu = 2.0f * w * diff * invB;
      } else {
        const float ls = fminf(fmaxf(lsr[a], -20.0f), 2.0f);
        const float sigma = expf(ls);
        const float inv_var = 1.0f / (sigma * sigma);
        bc += 0.5f * diff * diff * inv_var + ls + 0.91893853320467274178f;
        dmu = w * diff * inv_var * invB;
        gls = (lsr[a] >= -20.0f && lsr[a] <= 2.0f) ? w * (1.0f - diff * diff * inv_var) * invB : 0.f;
      }
      dza = dmu * (1.0f - mu[a] * mu[a]);
      dout[(int64_t)(3 + a) * L.B + b] = dza;
    }
    dzs[a] = dza; gss[a] = gls;
  }
  if (bf && live) {
    // the actor's dz row (columns 0..A-1) and its column of ones (A): eight bf16 = one 16-byte chunk of the image per store
    // (one 2-byte store per action cost a 32-line LSU pass each)
    uint8_t* zrow = dzimg + ((int64_t)agent * kTrain + 3) * z_bytes;
#pragma unroll
    for (int a0 = 0; a0 <= kOutMax; a0 += 8) {
      if (a0 <= L.A) {  // uniform
        float c[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) c[k] = (a0 + k < kOutMax && a0 + k < L.A) ? dzs[a0 + k < kOutMax ? a0 + k : 0] : (a0 + k == L.A ? 1.0f : 0.f);
        *reinterpret_cast<uint4*>(zrow + umma::image_off(L.B, b, a0)) =
            make_uint4(umma::pack_bf16(c[0], c[1]), umma::pack_bf16(c[2], c[3]), umma::pack_bf16(c[4], c[5]), umma::pack_bf16(c[6], c[7]));
      }
    }
  }
#pragma unroll
  for (int a0 = 0; a0 < kOutMax; a0 += 8) {
    if (a0 < L.A) {  // uniform
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          dzs[a0 + k] += __shfl_xor_sync(0xffffffffu, dzs[a0 + k], o);
          gss[a0 + k] += __shfl_xor_sync(0xffffffffu, gss[a0 + k], o);
        }
      }
      if (lane == 0) {
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (a0 + k < L.A) { red[warp][6 + a0 + k] = dzs[a0 + k]; red[warp][6 + L.A + a0 + k] = gss[a0 + k]; }
      }
    }
  }
  const float al = live ? w * bc : 0.f;
  {
    const float s0 = warp_sum(vl), s1 = warp_sum(ql), s2 = warp_sum(al);
    const float s3 = warp_sum(d0), s4 = warp_sum(d1), s5 = warp_sum(d2);
    if (lane == 0) {
      red[warp][0] = s0; red[warp][1] = s1; red[warp][2] = s2; red[warp][3] = s3; red[warp][4] = s4; red[warp][5] = s5;
    }
  }
  if (threadIdx.x == 0) IQLB_TSL(2);
  __syncthreads();
  if (threadIdx.x == 0) IQLB_TSL(3);
  const int nvals = 6 + 2 * L.A;
  for (int j = b; j < nvals; j += blockDim.x) {  // thread j adds the per-warp partials of quantity j in warp order
    float tot = 0.f;
    for (int wi = 0; wi < nw; ++wi) tot += red[wi][j];
    float* g = grads + (int64_t)agent * L.Pstride;
    if (j == 0) {
      const float value_loss = tot * invB;
      losses[agent * 3 + 0] = value_loss; if (losses_snap) losses_snap[agent * 3 + 0] = value_loss; atomicAdd(loss_sum + agent * 3 + 0, value_loss);  // one writer: a fire-and-forget add
    } else if (j == 1) {
      const float q_loss = tot * invB * 0.5f;
      losses[agent * 3 + 1] = q_loss; if (losses_snap) losses_snap[agent * 3 + 1] = q_loss; atomicAdd(loss_sum + agent * 3 + 1, q_loss);  // one writer: a fire-and-forget add
    } else if (j == 2) {
      const float actor_loss = tot * invB;
      losses[agent * 3 + 2] = actor_loss; if (losses_snap) losses_snap[agent * 3 + 2] = actor_loss; atomicAdd(loss_sum + agent * 3 + 2, actor_loss);  // one writer: a fire-and-forget add
    } else {
      int64_t e = -1;
      if (j < 6) { if (bf) e = (j == 3 ? L.off_v + L.b2(L.S, 1) : (j == 4 ? L.off_q1 : L.off_q2) + L.b2(L.Dq, 1)); }
      else if (j < 6 + L.A) { if (bf) e = L.off_a + L.b2(L.S, L.A) + (j - 6); }
      else if (!L.deterministic) e = L.off_logstd + (j - 6 - L.A);
      if (e >= 0) {
        g[e] = tot;
        if (gradh != nullptr) gradh[(int64_t)agent * L.Pstride + e] = __float2bfloat16_rn(tot);
      }
    }
  }
  if (threadIdx.x == 0) IQLB_TSL(4);
  if (!extra_warp && (b == 32 || (blockDim.x <= 32 && b == 0))) bookkeeping();
}

// ------------------------------------------------------------------------------------------
// Heads backward for the 4 trainable streams:  dW2 = dOut^T H2, db2 = sum_b dOut,
// dH2pre[b,j] = (H2[b,j] > 0) * sum_o dOut[o,b] * W2[o,j].   One CTA per (stream, agent).
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) head_bwd_kernel(Layout L, const float* __restrict__ params,
                                                       const float* __restrict__ H2,
                                                       const float* __restrict__ d_out, float* __restrict__ dH2,
                                                       float* __restrict__ grads, float actor_drop_scale) {
  const int agent = blockIdx.y, t = blockIdx.x, s = train_stream(t);
  const int in = L.in_dim(s), nout = L.out_dim(s);
  const int64_t noff = L.net_off(s);
  const float* W2 = params + (int64_t)agent * L.Pstride + noff + L.w2(in);
  float* gW2 = grads + (int64_t)agent * L.Pstride + noff + L.w2(in);
  float* gb2 = grads + (int64_t)agent * L.Pstride + noff + L.b2(in, nout);
  const float* h2 = H2 + ((int64_t)agent * kStreams + s) * L.B * L.H;
  float* dh2 = dH2 + ((int64_t)agent * kTrain + t) * L.B * L.H;
  const float* dout = d_out + (int64_t)agent * L.dout_rows() * L.B;
  const int drow0 = (t == 3) ? 3 : t;  // V->0, Q1->1, Q2->2, actor->3..
  for (int j = threadIdx.x; j < L.H; j += blockDim.x) {
    for (int o0 = 0; o0 < nout; o0 += 8) {
      const int no = min(8, nout - o0);
      float w2[8], acc[8];
#pragma unroll
      for (int oo = 0; oo < 8; ++oo) {
        w2[oo] = oo < no ? W2[(int64_t)(o0 + oo) * L.H + j] : 0.f;
        acc[oo] = 0.f;
      }
      const bool first = (o0 == 0), last = (o0 + 8 >= nout);
#pragma unroll 4
      for (int b = 0; b < L.B; ++b) {
        const float h = h2[(int64_t)b * L.H + j];
        float g = first ? 0.f : dh2[(int64_t)b * L.H + j];
#pragma unroll
        for (int oo = 0; oo < 8; ++oo) {
          if (oo < no) {
            const float d = __ldg(dout + (int64_t)(drow0 + o0 + oo) * L.B + b);
            acc[oo] = fmaf(d, h, acc[oo]);
            g = fmaf(d, w2[oo], g);
          }
        }
        dh2[(int64_t)b * L.H + j] = last ? (h > 0.f ? (t == 3 ? g * actor_drop_scale : g) : 0.f) : g;
      }
#pragma unroll
      for (int oo = 0; oo < 8; ++oo)
        if (oo < no) gW2[(int64_t)(o0 + oo) * L.H + j] = acc[oo];
    }
  }
  for (int o = threadIdx.x; o < nout; o += blockDim.x) {
    float sacc = 0.f;
    for (int b = 0; b < L.B; ++b) sacc += dout[(int64_t)(drow0 + o) * L.B + b];
    gb2[o] = sacc;
  }
}

// ------------------------------------------------------------------------------------------
// Adam (torch _fused_adam math, betas (0.9, 0.999), eps 1e-8) for the three optimizers of every
// agent, with the Polyak target update (iql.py:127-129,613) fused for the Q segment.
// HBM-bound: 28 B read + 16 B written per parameter (+8 B per Q parameter for the target).
// ------------------------------------------------------------------------------------------
#ifndef IQLB_ADAM_SLOTS
#define IQLB_ADAM_SLOTS 1
#endif
// One CTA owns `256 * IQLB_ADAM_SLOTS` 16-byte words (a "chunk") of one agent's parameter vector at a time and walks
// chunks blockIdx.x, blockIdx.x + gridDim.x, ... of the launch (agent-major order).  Two launch shapes:
//   * one chunk per CTA (gridDim.x == n_chunks): the stand-alone optimizer pass, every SM at full occupancy;
//   * a RESIDENCY-BOUNDED persistent grid (gridDim.x == k * SMs, k small): the optimizer of agent group g runs on a side
//     stream UNDER the tensor-core kernels of group g+1 (iql_step.cu step_device).  Those kernels keep one CTA per SM
//     with ~39 K registers and all of shared memory; a 256-thread CTA of this kernel needs no shared memory and
//     <= 48 registers per thread (launch bound: 5 CTAs per SM), so two of them fit beside it on every SM and stream HBM while the GEMM CTA waits on
//     its own latencies.  An unbounded grid would instead fill every SM's thread slots and lock the GEMM kernels out.
// (Round 2, measured and removed: the next step's gather as extra CTAs of this launch, and a one-shot persistent grid of 5
//  CTAs per SM.  With the gather code merely COMPILED INTO this kernel - shared memory, an mbarrier, one __syncthreads, none
//  of it executed by the optimizer CTAs - the optimizer pass went from 121 to 133 us; same box, same step otherwise:
//  241.6 us with this plain kernel and a separate gather launch, 257.3 with the gather CTAs riding along, 252.6 persistent.)
template <bool kPersist>
__global__ void __launch_bounds__(256, kPersist ? 5 : 1) adam_polyak_kernel(Layout L, float* __restrict__ params,
                                                          float* __restrict__ exp_avg, float* __restrict__ exp_avg_sq,
                                                          float* __restrict__ target, const float* __restrict__ grads,
                                                          const StepScalars* __restrict__ scal, BfParams BP, int shadow,
                                                          int agent0, int n_agents_launch, int chunks_per_agent) {
  // Thread t handles the 16-byte words t (and t + 256, ...) of the chunk, so every warp access is 512 contiguous bytes per
  // array.  Network segments start on multiples of 8 floats: a word never straddles two optimizers, and inside W1 it
  // is half of a 16-byte chunk of the image.
  const bool hint = (shadow & 2) != 0;  // stream the optimizer state through L2 (evict_first), keep the images (evict_last)
  const uint64_t pol_s = l2_policy_evict_first(), pol_k = l2_policy_evict_last();
  constexpr int kS = IQLB_ADAM_SLOTS;  // 16-byte words per thread and array
  auto body = [&](const int chunk, const bool first) {
  const int aidx = chunk / chunks_per_agent, cblk = chunk - aidx * chunks_per_agent;
  const int agent = agent0 + ((shadow & 4) ? n_agents_launch - 1 - aidx : aidx);
  const StepScalars sc = scal[agent];
  float4 p[kS], m[kS], v[kS], g[kS], tg[kS];
  int64_t e[kS];
  bool live[kS], has_t[kS];
#pragma unroll
  for (int h = 0; h < kS; ++h) {  // optimizer state: not written by the weight-gradient kernel (the PDL predecessor)
    e[h] = ((int64_t)cblk * (256 * kS) + threadIdx.x + 256 * h) * 4;
    live[h] = e[h] < L.Pstride;
    has_t[h] = e[h] < L.PQ;
    if (live[h]) {
      const int64_t g4 = (int64_t)agent * L.Pstride + e[h];
      if (hint) {
        p[h] = ld_f4_hint(params + g4, pol_s);
        m[h] = ld_f4_hint(exp_avg + g4, pol_s);
        v[h] = ld_f4_hint(exp_avg_sq + g4, pol_s);
        if (has_t[h]) tg[h] = ld_f4_hint(target + (int64_t)agent * L.PQstride + e[h], pol_s);
      } else {
        p[h] = *reinterpret_cast<const float4*>(params + g4);
        m[h] = *reinterpret_cast<const float4*>(exp_avg + g4);
        v[h] = *reinterpret_cast<const float4*>(exp_avg_sq + g4);
        if (has_t[h]) tg[h] = *reinterpret_cast<const float4*>(target + (int64_t)agent * L.PQstride + e[h]);
      }
    }
  }
  if (first) {
    pdl_wait();
    pdl_launch();
  }
#pragma unroll
  for (int h = 0; h < kS; ++h) {  // gradients of this step
    if (!live[h]) continue;
    const int64_t g4 = (int64_t)agent * L.Pstride + e[h];
    if (shadow & 1) {  // bf16 path: gradients arrive in bf16
      const uint2 gh = *reinterpret_cast<const uint2*>(BP.gradh + g4);
      g[h] = make_float4(__uint_as_float(gh.x << 16), __uint_as_float(gh.x & 0xffff0000u), __uint_as_float(gh.y << 16),
                         __uint_as_float(gh.y & 0xffff0000u));
    } else {
      g[h] = *reinterpret_cast<const float4*>(grads + g4);
    }
  }
#pragma unroll
  for (int h = 0; h < kS; ++h) {
    if (!live[h]) continue;
    const float ss = e[h] < L.off_v ? sc.ss_q : (e[h] < L.off_a ? sc.ss_v : sc.ss_a);
    float* pp = &p[h].x; float* mp = &m[h].x; float* vp = &v[h].x; const float* gp = &g[h].x; float* tp = &tg[h].x;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const float gg = gp[c];
      const float mm = mp[c] + (1.0f - 0.9f) * (gg - mp[c]);
      const float vv = 0.999f * vp[c] + (1.0f - 0.999f) * gg * gg;
      const float denom = sqrtf(vv) / sc.bc2_sqrt + 1e-8f;
      pp[c] = pp[c] - ss * (mm / denom);
      mp[c] = mm;
      vp[c] = vv;
      if (has_t[h]) tp[c] = tp[c] + sc.tau * (pp[c] - tp[c]);  // Tensor.lerp_ (weight < 0.5), with the post-Adam qf
    }
    const int64_t g4 = (int64_t)agent * L.Pstride + e[h];
    if (hint) {
      st_f4_hint(params + g4, p[h], pol_s);
      st_f4_hint(exp_avg + g4, m[h], pol_s);
      st_f4_hint(exp_avg_sq + g4, v[h], pol_s);
      if (has_t[h]) st_f4_hint(target + (int64_t)agent * L.PQstride + e[h], tg[h], pol_s);
    } else {
      *reinterpret_cast<float4*>(params + g4) = p[h];
      *reinterpret_cast<float4*>(exp_avg + g4) = m[h];
      *reinterpret_cast<float4*>(exp_avg_sq + g4) = v[h];
      if (has_t[h]) *reinterpret_cast<float4*>(target + (int64_t)agent * L.PQstride + e[h]) = tg[h];
    }
    if (shadow & 1) {  // refresh the bf16 operand images of the weights just updated
      uint8_t* wimg = BP.wimg + (int64_t)agent * BP.agent_w_bytes;
      int net;
      int64_t off;
      if (shadow_loc(L, BP.KXB, e[h], net, off) && off >= BP.w0_bytes) {  // W1: four columns of one image chunk
        if (hint) {
          st_u2_hint(wimg + (int64_t)net * BP.net_bytes + off, umma::pack_bf16(pp[0], pp[1]), umma::pack_bf16(pp[2], pp[3]), pol_k);
          if (has_t[h])
            st_u2_hint(wimg + (int64_t)(net + 4) * BP.net_bytes + off, umma::pack_bf16(tp[0], tp[1]),
                       umma::pack_bf16(tp[2], tp[3]), pol_k);
        } else {
          *reinterpret_cast<uint2*>(wimg + (int64_t)net * BP.net_bytes + off) =
              make_uint2(umma::pack_bf16(pp[0], pp[1]), umma::pack_bf16(pp[2], pp[3]));
          if (has_t[h])
            *reinterpret_cast<uint2*>(wimg + (int64_t)(net + 4) * BP.net_bytes + off) =
                make_uint2(umma::pack_bf16(tp[0], tp[1]), umma::pack_bf16(tp[2], tp[3]));
        }
      } else {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          if (shadow_loc(L, BP.KXB, e[h] + c, net, off)) {
            *reinterpret_cast<__nv_bfloat16*>(wimg + (int64_t)net * BP.net_bytes + off) = __float2bfloat16_rn(pp[c]);
            if (has_t[h])
              *reinterpret_cast<__nv_bfloat16*>(wimg + (int64_t)(net + 4) * BP.net_bytes + off) = __float2bfloat16_rn(tp[c]);
          }
        }
      }
    }
  }
  };
  if (kPersist) {
    const int n_chunks = n_agents_launch * chunks_per_agent;
    bool first = true;
    for (int chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
      body(chunk, first);
      first = false;
    }
  } else {
    body((int)blockIdx.x, true);
  }
}

__global__ void bf16_to_f32_kernel(const __nv_bfloat16* __restrict__ src, float* __restrict__ dst, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = __bfloat162float(src[i]);
}

// ------------------------------------------------------------------------------------------
static int build_tasks(iqlb_agents* h) {
  const Layout& L = h->L;
  const int N = h->desc.n_agents;
  cudaFree(h->t_fwd0); cudaFree(h->t_fwd1); cudaFree(h->t_bw1); cudaFree(h->t_bd1); cudaFree(h->t_bw0);
  h->t_fwd0 = h->t_fwd1 = h->t_bw1 = h->t_bd1 = h->t_bw0 = nullptr;
  const bool drop = h->drop_p > 0.f;
  const uint32_t drop_thresh = drop ? (uint32_t)((double)h->drop_p * 4294967296.0) : 0u;
  const float drop_scale = drop ? 1.0f / (1.0f - h->drop_p) : 1.0f;
  std::vector<GemmTask> f0((size_t)N * kStreams), f1((size_t)N * kStreams), bw1((size_t)N * kTrain),
      bd1((size_t)N * kTrain), bw0((size_t)N * kTrain);
  for (int a = 0; a < N; ++a) {
    float* P = h->params + (int64_t)a * L.Pstride;
    float* T = h->target + (int64_t)a * L.PQstride;
    float* G = h->grads + (int64_t)a * L.Pstride;
    float* X = h->batch + (int64_t)a * L.B * L.R;
    for (int s = 0; s < kStreams; ++s) {
      const int in = L.in_dim(s);
      const float* net = (L.uses_target(s) ? T : P) + L.net_off(s);
      float* h1 = h->H1 + ((int64_t)a * kStreams + s) * L.B * L.H;
      float* h2 = h->H2 + ((int64_t)a * kStreams + s) * L.B * L.H;
      GemmTask t{};
      t.A = X + (s == 0 ? L.S + L.A : 0); t.lda = L.R;
      t.Bm = net + L.w0(); t.ldb = in; t.bias = net + L.b0(in);
      t.C = h1; t.ldc = L.H; t.M = L.B; t.N = L.H; t.K = in;
      if (drop && s == 6) { t.drop_thresh = drop_thresh; t.drop_scale = drop_scale; t.drop_layer = 0; t.drop_rng = h->rng + 2 * a; t.drop_it = h->it + a; }
      f0[(size_t)a * kStreams + s] = t;
      GemmTask u{};
      u.A = h1; u.lda = L.H; u.Bm = net + L.w1(in); u.ldb = L.H; u.bias = net + L.b1(in);
      u.C = h2; u.ldc = L.H; u.M = L.B; u.N = L.H; u.K = L.H;
      if (drop && s == 6) { u.drop_thresh = drop_thresh; u.drop_scale = drop_scale; u.drop_layer = 1; u.drop_rng = h->rng + 2 * a; u.drop_it = h->it + a; }
      f1[(size_t)a * kStreams + s] = u;
    }
    for (int tt = 0; tt < kTrain; ++tt) {
      const int s = train_stream(tt), in = L.in_dim(s);
      const float* net = P + L.net_off(s);
      float* gnet = G + L.net_off(s);
      const float* h1 = h->H1 + ((int64_t)a * kStreams + s) * L.B * L.H;
      float* dh2 = h->dH2 + ((int64_t)a * kTrain + tt) * L.B * L.H;
      float* dh1 = h->dH1 + ((int64_t)a * kTrain + tt) * L.B * L.H;
      GemmTask w1{};  // dW1[H,H] = dH2^T H1 ; db1
      w1.A = dh2; w1.lda = L.H; w1.Bm = h1; w1.ldb = L.H; w1.C = gnet + L.w1(in); w1.ldc = L.H;
      w1.bias_grad = gnet + L.b1(in); w1.M = L.H; w1.N = L.H; w1.K = L.B;
      bw1[(size_t)a * kTrain + tt] = w1;
      GemmTask d1{};  // dH1 = (dH2 W1) * relu'(H1)
      d1.A = dh2; d1.lda = L.H; d1.Bm = net + L.w1(in); d1.ldb = L.H; d1.C = dh1; d1.ldc = L.H;
      d1.mask = h1; d1.M = L.B; d1.N = L.H; d1.K = L.H;
      d1.scale = (drop && tt == 3) ? drop_scale : 1.0f;
      bd1[(size_t)a * kTrain + tt] = d1;
      GemmTask w0{};  // dW0[H,in] = dH1^T X ; db0
      w0.A = dh1; w0.lda = L.H; w0.Bm = X; w0.ldb = L.R; w0.C = gnet + L.w0(); w0.ldc = in;
      w0.bias_grad = gnet + L.b0(in); w0.M = L.H; w0.N = in; w0.K = L.B;
      bw0[(size_t)a * kTrain + tt] = w0;
    }
  }
  auto up = [](GemmTask** dst, const std::vector<GemmTask>& v) -> int {
    IQLB_CUDA(cudaMalloc((void**)dst, v.size() * sizeof(GemmTask)));
    IQLB_CUDA(cudaMemcpy(*dst, v.data(), v.size() * sizeof(GemmTask), cudaMemcpyHostToDevice));
    return IQLB_OK;
  };
  int rc;
  if ((rc = up(&h->t_fwd0, f0))) return rc;
  if ((rc = up(&h->t_fwd1, f1))) return rc;
  if ((rc = up(&h->t_bw1, bw1))) return rc;
  if ((rc = up(&h->t_bd1, bd1))) return rc;
  if ((rc = up(&h->t_bw0, bw0))) return rc;
  return IQLB_OK;
}

// The launch chain of one update for agents [a0, a0+n) up to (not including) the optimizer.
static int step_chain(iqlb_agents* h, int a0, int n, bool sampled, Prof& pf, cudaStream_t st) {
  const Layout& L = h->L;
  const bool bf = h->desc.precision == 1;
  const BfParams* bp = bf16_params(h);
  int rc;
  if (sampled && bf && h->pre_valid) {
    // gathered and packed during the previous step's optimizer kernel (look-ahead sampling)
    IQLB_CUDA(cudaStreamWaitEvent(st, h->ev_pre_join, 0));
    h->pre_valid = false;
  } else if (sampled && bf) {
    if ((rc = bf16_gather_pack(h, a0, n, pf, st))) return rc;
  } else if (sampled) {
    IQLB_REQUIRE(h->table != nullptr, "no replay table attached (iqlb_agents_set_replay)");
    rc = launch_gather(h->table, h->n_rows, L.R, 2 * L.S + L.A, h->reward_cols,
                       h->reward_variant ? h->reward_variant + a0 : nullptr, h->rng + 2 * a0, nullptr, n, L.B,
                       h->idx_out ? h->idx_out + (int64_t)a0 * L.B : nullptr, h->batch + (int64_t)a0 * L.B * L.R, st);
    if (rc) return rc;
    pf.mark("replay_gather");
  }
  if (bf) {
    if ((rc = bf16_step_gemms_fwd(h, a0, n, pf, st, sampled))) return rc;
  } else {
    const int N = h->desc.n_agents;  // the fp32 parity path always runs all agents in one group
    if ((rc = launch_grouped_gemm(GEMM_FWD, h->t_fwd0, N * kStreams, L.B, L.H, false, st))) return rc;
    pf.mark("gemm_fwd_l0");
    if ((rc = launch_grouped_gemm(GEMM_FWD, h->t_fwd1, N * kStreams, L.B, L.H, false, st))) return rc;
    pf.mark("gemm_fwd_l1");
    head_fwd_kernel<<<dim3((L.B + 7) / 8, kStreams, N), 256, 0, st>>>(L, h->params, h->target, h->H2, h->head_out);
    IQLB_LAUNCH_CHECK();
    pf.mark("head_fwd");
  }
  int lb = (L.B + 31) / 32 * 32;
  if (lb + 32 <= 1024) lb += 32;  // bookkeeping warp
  IQLB_CUDA(launch_pdl(lb <= 320 ? loss_kernel<320> : loss_kernel<1024>, dim3(n), dim3(lb), 0, st, bf && pdl_enabled(4) && !h->prof, L, (const float*)h->params,
                       h->hparams, (const float*)((bf && h->batch_src) ? h->batch_src : h->batch), (const float*)h->head_out, h->d_out, h->grads, h->losses,
                       h->loss_sum, h->it, h->actor_lr, h->rng, h->scal, sampled ? 1 : 0, bf ? bp->dz : (uint8_t*)nullptr,
                       bf ? bp->z_bytes : (int64_t)0, a0, bf ? bp->dbg : (unsigned long long*)nullptr,
                       bf ? bp->gradh : (__nv_bfloat16*)nullptr, h->loss_snap_cur));
  pf.mark("loss");
  if (bf) {
    if ((rc = bf16_step_gemms_bwd(h, a0, n, pf, st))) return rc;
  } else {
    const int N = h->desc.n_agents;
    head_bwd_kernel<<<dim3(kTrain, N), 256, 0, st>>>(L, h->params, h->H2, h->d_out, h->dH2, h->grads,
                                                     h->drop_p > 0.f ? 1.0f / (1.0f - h->drop_p) : 1.0f);
    IQLB_LAUNCH_CHECK();
    pf.mark("head_bwd");
    if ((rc = launch_grouped_gemm(GEMM_BWD_WEIGHT, h->t_bw1, N * kTrain, L.H, L.H, false, st))) return rc;
    pf.mark("gemm_bwd_w1");
    if ((rc = launch_grouped_gemm(GEMM_BWD_DATA, h->t_bd1, N * kTrain, L.B, L.H, false, st))) return rc;
    pf.mark("gemm_bwd_d1");
    if ((rc = launch_grouped_gemm(GEMM_BWD_WEIGHT, h->t_bw0, N * kTrain, L.H, L.Dq, true, st))) return rc;
    pf.mark("gemm_bwd_w0");
  }
  return IQLB_OK;
}

static int step_adam(iqlb_agents* h, int a0, int n, Prof& pf, cudaStream_t st, bool bounded) {
  const Layout& L = h->L;
  const bool bf = h->desc.precision == 1;
  const BfParams* bp = bf16_params(h);
  static const int l2hint_env = getenv("IQLB_L2HINT") ? atoi(getenv("IQLB_L2HINT")) : -1;
  // bounded (under the next group's tensor-core kernels): the optimizer state streams through L2 with evict_first so that
  // it does not push the other group's activations (written by one kernel, read by the next) out to HBM
  const bool l2hint = l2hint_env >= 0 ? l2hint_env != 0 : bounded;
  const int cpa = (int)((L.Pstride + 1024 * IQLB_ADAM_SLOTS - 1) / (1024 * IQLB_ADAM_SLOTS));
  const int n_chunks = cpa * n;
  int grid = n_chunks;
  if (bounded && h->n_sm > 0 && h->adam_ctas_per_sm > 0 && grid > h->n_sm * h->adam_ctas_per_sm) grid = h->n_sm * h->adam_ctas_per_sm;
  IQLB_CUDA(launch_pdl(grid < n_chunks ? adam_polyak_kernel<true> : adam_polyak_kernel<false>, dim3((unsigned)grid), dim3(256), 0, st,
                       bf && !bounded && pdl_enabled(32) && !h->prof, L, h->params, h->exp_avg, h->exp_avg_sq, h->target,
                       (const float*)h->grads, (const StepScalars*)h->scal, bf ? *bp : BfParams{},
                       (bf ? 1 : 0) | (l2hint ? 2 : 0) | ((bf && h->zig == 1) ? 4 : 0), a0, n, cpa));
  pf.mark(bf ? "adam_polyak_shadow" : "adam_polyak");
  return IQLB_OK;
}

// The caller's stream waits for every optimizer launch still running on the side stream.
static int join_groups(iqlb_agents* h, cudaStream_t st) {
  for (int g = 0; g < iqlb_agents::kMaxGroups; ++g) {
    if (h->adam_pending[g]) {
      IQLB_CUDA(cudaStreamWaitEvent(st, h->ev_adam[g], 0));
      h->adam_pending[g] = false;
    }
  }
  return IQLB_OK;
}

// One update of every agent.
// precision 1 with >= 2 agent groups (opt-in, IQLB_GROUPS; one group by default): the HBM-bound optimizer of group g runs on
// an internal low-priority stream as a residency-bounded persistent grid (adam_polyak_kernel) UNDER the latency-bound
// tensor-core chain of the next group on the caller's stream - agents are independent, so group g's chain of step t+1
// needs only group g's optimizer of step t.  Default: the caller's stream joins the side stream at the end of every call
// (stream semantics unchanged: chain(A), [Adam(A) || chain(B)], Adam(B)).  With iqlb_agents_set_overlap(h, 1) the join
// is deferred, so Adam(B) of step t also overlaps chain(A) of step t+1; the caller must then call iqlb_agents_sync
// before touching params / exp_avg / exp_avg_sq / target (losses, it, rng and actor_lr are written on the caller's
// stream and need no sync).  Profiling runs one group.
static int step_device(iqlb_agents* h, bool sampled, cudaStream_t st) {
  const int N = h->desc.n_agents;
  int rc;
  Prof pf(h, st);
  int G = (h->desc.precision == 1 && !h->prof) ? h->n_groups : 1;
  if (h->pre_valid && !sampled) {  // an explicit batch follows a look-ahead gather: let it finish, then overwrite it
    IQLB_CUDA(cudaStreamWaitEvent(st, h->ev_pre_join, 0));
    h->pre_valid = false;
  }
  if (G <= 1) {
    if ((rc = join_groups(h, st))) return rc;  // (a profiled or fused-Adam step after overlapped ones)
    if ((rc = step_chain(h, 0, N, sampled, pf, st))) return rc;
    const bool lookahead = sampled && h->desc.precision == 1 && !h->prof && h->idx_out == nullptr;
    if (lookahead && h->presample == 1) {
      if (!h->pre_stream) {
        int prio_lo = 0, prio_hi = 0;  // highest priority: the small gather must not queue behind the optimizer's CTAs
        IQLB_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
        IQLB_CUDA(cudaStreamCreateWithPriority(&h->pre_stream, cudaStreamNonBlocking, prio_hi));
        IQLB_CUDA(cudaEventCreateWithFlags(&h->ev_pre_fork, cudaEventDisableTiming));
        IQLB_CUDA(cudaEventCreateWithFlags(&h->ev_pre_join, cudaEventDisableTiming));
      }
      IQLB_CUDA(cudaEventRecord(h->ev_pre_fork, st));
      IQLB_CUDA(cudaStreamWaitEvent(h->pre_stream, h->ev_pre_fork, 0));
      Prof side(h, h->pre_stream);  // not profiling here: only counts
      if ((rc = bf16_gather_pack(h, 0, N, side, h->pre_stream))) return rc;
      IQLB_CUDA(cudaEventRecord(h->ev_pre_join, h->pre_stream));
      h->pre_valid = true;
      pf.mark("replay_gather_pack(next step)");
    }
    if ((rc = step_adam(h, 0, N, pf, st, false))) return rc;
    pf.done();
    if (h->zig >= 0 && h->desc.precision == 1) h->zig ^= 1;
    return IQLB_OK;
  }
  const int per = (N + G - 1) / G;
  static const bool tl = getenv("IQLB_GROUP_TL") && atoi(getenv("IQLB_GROUP_TL")) != 0;  // debug: per-step group timeline
  static cudaEvent_t tl_start = nullptr;
  if (tl) {
    if (!tl_start) cudaEventCreate(&tl_start);
    cudaEventRecord(tl_start, st);
  }
  for (int g = 0, a0 = 0; a0 < N; ++g, a0 += per) {
    const int n = a0 + per <= N ? per : N - a0;
    if (h->adam_pending[g]) {  // this group's optimizer of the previous step (deferred join)
      IQLB_CUDA(cudaStreamWaitEvent(st, h->ev_adam[g], 0));
      h->adam_pending[g] = false;
    }
    if ((rc = step_chain(h, a0, n, sampled, pf, st))) return rc;
    IQLB_CUDA(cudaEventRecord(h->ev_chain[g], st));
    IQLB_CUDA(cudaStreamWaitEvent(h->side_stream, h->ev_chain[g], 0));
    if ((rc = step_adam(h, a0, n, pf, h->side_stream, true))) return rc;
    IQLB_CUDA(cudaEventRecord(h->ev_adam[g], h->side_stream));
    h->adam_pending[g] = true;
  }
  if (tl) {  // synchronises: shows the schedule inside ONE step (no cross-step overlap while this is on)
    static int n_print = 0;
    cudaStreamSynchronize(h->side_stream);
    cudaStreamSynchronize(st);
    if (n_print++ % 20 == 15) {
      for (int g = 0; g < G; ++g) {
        float tc = 0.f, ta = 0.f;
        cudaEventElapsedTime(&tc, tl_start, h->ev_chain[g]);
        cudaEventElapsedTime(&ta, tl_start, h->ev_adam[g]);
        fprintf(stderr, "[group timeline] group %d: chain done at %.1f us, optimizer done at %.1f us\n", g, tc * 1e3, ta * 1e3);
      }
    }
  }
  if (!h->overlap && (rc = join_groups(h, st))) return rc;
  pf.done();
  if (h->zig >= 0) h->zig ^= 1;
  return IQLB_OK;
}

}  // namespace iqlb

using namespace iqlb;

extern "C" int iqlb_param_count(int32_t state_dim, int32_t action_dim, int32_t hidden_dim, int32_t deterministic,
                                int64_t* p_all, int64_t* p_q) {
  IQLB_REQUIRE(state_dim > 0 && action_dim > 0 && hidden_dim > 0, "dims must be positive");
  Layout L;
  L.init(state_dim, action_dim, hidden_dim, deterministic, 4, 1);
  if (p_all) *p_all = L.P;
  if (p_q) *p_q = L.PQ;
  return IQLB_OK;
}

extern "C" int iqlb_agents_create(const iqlb_agents_desc* d, float* params, float* exp_avg, float* exp_avg_sq,
                                  float* target, const float* hparams, int64_t* it, double* actor_lr, uint64_t* rng,
                                  float* losses, float* loss_sum, iqlb_agents** out) {
  IQLB_REQUIRE(d && out, "null desc/out");
  IQLB_REQUIRE(params && exp_avg && exp_avg_sq && target && hparams && it && actor_lr && rng && losses && loss_sum,
               "null state buffer");
  IQLB_REQUIRE(d->n_agents > 0 && d->state_dim > 0 && d->action_dim > 0 && d->hidden_dim > 0, "dims");
  IQLB_REQUIRE(d->batch_size > 0 && d->batch_size <= 1024, "batch_size in [1, 1024]");
  IQLB_REQUIRE(d->action_dim <= kOutMax, "action_dim <= 24 (widest D4RL action space in the reference configs)");
  IQLB_REQUIRE(d->row_stride % 4 == 0 && d->row_stride >= 2 * d->state_dim + d->action_dim + 2,
               "row_stride must hold obs|act|next_obs|rew|done and be a multiple of 4");
  IQLB_REQUIRE(d->precision == 0 || d->precision == 1, "precision must be 0 (fp32) or 1 (bf16 tcgen05)");
  IQLB_REQUIRE(((uintptr_t)params & 15) == 0 && ((uintptr_t)exp_avg & 15) == 0 && ((uintptr_t)exp_avg_sq & 15) == 0 &&
                   ((uintptr_t)target & 15) == 0,
               "state buffers must be 16-byte aligned");
  iqlb_agents* h = new (std::nothrow) iqlb_agents();
  if (!h) return fail(IQLB_ERR_NOMEM, "out of host memory%s", "");
  h->desc = *d;
  h->L.init(d->state_dim, d->action_dim, d->hidden_dim, d->deterministic, d->row_stride, d->batch_size);
  h->params = params; h->exp_avg = exp_avg; h->exp_avg_sq = exp_avg_sq; h->target = target;
  h->hparams = hparams; h->it = it; h->actor_lr = actor_lr; h->rng = rng; h->losses = losses; h->loss_sum = loss_sum;
  const Layout& L = h->L;
  const int64_t N = d->n_agents, BH = (int64_t)L.B * L.H;
  auto alloc = [&](float** p, int64_t n) -> int {
    IQLB_CUDA(cudaMalloc((void**)p, (size_t)n * sizeof(float)));
    IQLB_CUDA(cudaMemset(*p, 0, (size_t)n * sizeof(float)));
    return IQLB_OK;
  };
  int rc = IQLB_OK;
  if (!rc) rc = alloc(&h->batch, N * L.B * L.R);
  if (!rc) rc = alloc(&h->H1, N * kStreams * BH);
  if (!rc) rc = alloc(&h->H2, N * kStreams * BH);
  if (!rc) rc = alloc(&h->head_out, N * L.head_rows() * L.B);
  if (!rc) rc = alloc(&h->d_out, N * L.dout_rows() * L.B);
  if (!rc) rc = alloc(&h->dH2, N * kTrain * BH);
  if (!rc) rc = alloc(&h->dH1, N * kTrain * BH);
  if (!rc) rc = alloc(&h->grads, N * L.Pstride);
  if (!rc) rc = alloc((float**)&h->scal, N * (int64_t)(sizeof(StepScalars) / sizeof(float)));
  if (!rc) rc = build_tasks(h);
  if (!rc && d->precision == 1) rc = bf16_create(h);
  h->zig = (getenv("IQLB_ZIGZAG") && atoi(getenv("IQLB_ZIGZAG")) == 0) ? -1 : 0;
  // measured on B200 (64 agents): 268.7 vs 269.1 us per step - the forward kernel loses its programmatic overlap
  // with the optimizer's tail to the cross-stream join, which eats what the hidden gather saves.  Off by default.
  // look-ahead sampling: the next step's gather on a side stream under the optimizer kernel.  Measured at the end of round 2:
  // slower with 64 agents (228.3 vs 225.2 us per step: the optimizer is HBM-bound and the forward kernel loses its programmatic
  // overlap with the optimizer's tail to the cross-stream join), faster with 8 (58.1 vs 59.6 us: the optimizer's 34 MB sit in
  // L2 and every kernel is a partial wave).  Opt-in (IQLB_PRESAMPLE=1) at any size: the look-ahead batch is gathered one
  // call early, so a caller that rewrites rows of the attached table in place between two steps would train on the old
  // rows - iqlb_agents_set_replay and every state-changing entry point drop it, an in-place write cannot be seen.
  h->presample = (getenv("IQLB_PRESAMPLE") && atoi(getenv("IQLB_PRESAMPLE")) != 0) ? 1 : 0;
  if (!rc && d->precision == 1) {
    const char* env = getenv("IQLB_GROUPS");
    // Opt-in (IQLB_GROUPS=2): measured on B200 with 64 agents, two groups are SLOWER than one (313-370 vs 260 us per
    // step, profiles/r02_overlap_ab.jsonl).  The group timeline shows why: a 32-agent chain alone takes 96 us and its
    // optimizer 65 us, but run together they take 150 us and 113 us - the latency-bound tensor-core kernels see every
    // operand load queue behind the optimizer's saturated HBM stream, and two half-size chains (2 x 96 us) cost 53 us
    // more than one full-size chain (139 us) to begin with.
    int G = env ? atoi(env) : 1;
    if (G < 1) G = 1;
    if (G > iqlb_agents::kMaxGroups) G = iqlb_agents::kMaxGroups;
    if (G > d->n_agents) G = d->n_agents;
    h->n_groups = G;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&h->n_sm, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
      rc = fail(IQLB_ERR_CUDA, "cudaDeviceGetAttribute%s", "");
    h->adam_ctas_per_sm = getenv("IQLB_ADAM_CTAS") ? atoi(getenv("IQLB_ADAM_CTAS")) : 2;
    h->overlap = (getenv("IQLB_OVERLAP") && atoi(getenv("IQLB_OVERLAP")) != 0) ? 1 : 0;
    if (G > 1 && !rc) {
      int prio_lo = 0, prio_hi = 0;  // lowest priority: tensor-core CTAs of the caller's stream are placed first
      if (cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi) != cudaSuccess ||
          cudaStreamCreateWithPriority(&h->side_stream, cudaStreamNonBlocking, prio_lo) != cudaSuccess)
        rc = fail(IQLB_ERR_CUDA, "cudaStreamCreate%s", "");
      const unsigned evf = (getenv("IQLB_GROUP_TL") && atoi(getenv("IQLB_GROUP_TL")) != 0) ? cudaEventDefault : cudaEventDisableTiming;
      for (int g = 0; g < G && !rc; ++g)
        if (cudaEventCreateWithFlags(&h->ev_chain[g], evf) != cudaSuccess ||
            cudaEventCreateWithFlags(&h->ev_adam[g], evf) != cudaSuccess)
          rc = fail(IQLB_ERR_CUDA, "cudaEventCreate%s", "");
    }
  }
  if (rc) { iqlb_agents_destroy(h); return rc; }
  *out = h;
  return IQLB_OK;
}

extern "C" int iqlb_agents_destroy(iqlb_agents* h) {
  if (!h) return IQLB_OK;
  bf16_destroy(h);
  if (h->side_stream) { cudaStreamSynchronize(h->side_stream); cudaStreamDestroy(h->side_stream); }
  if (h->pre_stream) {
    cudaStreamSynchronize(h->pre_stream);
    cudaStreamDestroy(h->pre_stream);
    cudaEventDestroy(h->ev_pre_fork);
    cudaEventDestroy(h->ev_pre_join);
  }
  for (int g = 0; g < iqlb_agents::kMaxGroups; ++g) {
    if (h->ev_chain[g]) cudaEventDestroy(h->ev_chain[g]);
    if (h->ev_adam[g]) cudaEventDestroy(h->ev_adam[g]);
  }
  if (h->copy_stream) {
    cudaStreamSynchronize(h->copy_stream);
    cudaStreamDestroy(h->copy_stream);
    if (h->d2h_stream) { cudaStreamSynchronize(h->d2h_stream); cudaStreamDestroy(h->d2h_stream); }
    for (int i = 0; i < 2; ++i) {
      if (h->ev_h2d[i]) cudaEventDestroy(h->ev_h2d[i]);
      if (h->ev_done[i]) cudaEventDestroy(h->ev_done[i]);
      if (h->ev_step[i]) cudaEventDestroy(h->ev_step[i]);
      cudaFree(h->stage[i]);
      cudaFree(h->loss_snap[i]);
      if (h->host_loss[i]) cudaFreeHost(h->host_loss[i]);
    }
  }
  cudaFree(h->batch); cudaFree(h->H1); cudaFree(h->H2); cudaFree(h->head_out); cudaFree(h->d_out);
  cudaFree(h->dH2); cudaFree(h->dH1); cudaFree(h->grads); cudaFree(h->scal);
  for (int i = 0; i <= iqlb_agents::kMaxProf; ++i) if (h->ev[i]) cudaEventDestroy(h->ev[i]);
  cudaFree(h->t_fwd0); cudaFree(h->t_fwd1); cudaFree(h->t_bw1); cudaFree(h->t_bd1); cudaFree(h->t_bw0);
  delete h;
  return IQLB_OK;
}

extern "C" int iqlb_agents_set_replay(iqlb_agents* h, const float* table, int64_t n_rows, const float* reward_cols,
                                      const int32_t* reward_variant, int64_t* idx_out) {
  IQLB_REQUIRE(h && table, "null handle/table");
  IQLB_REQUIRE(n_rows > 0, "empty table");
  IQLB_REQUIRE(((uintptr_t)table & 15) == 0, "table must be 16-byte aligned");
  if (h->pre_valid) {  // a look-ahead gather of the old table is in flight (or done): drop it
    IQLB_CUDA(cudaStreamSynchronize(h->pre_stream));
    h->pre_valid = false;
  }
  h->table = table; h->n_rows = n_rows; h->reward_cols = reward_cols; h->reward_variant = reward_variant;
  h->idx_out = idx_out;
  return IQLB_OK;
}

extern "C" int iqlb_agents_step(iqlb_agents* h, const float* batch, void* stream) {
  IQLB_REQUIRE(h != nullptr, "null handle");
  cudaStream_t st = (cudaStream_t)stream;
  if (batch != nullptr && batch != h->batch) {
    const Layout& L = h->L;
    IQLB_CUDA(cudaMemcpyAsync(h->batch, batch, (size_t)h->desc.n_agents * L.B * L.R * sizeof(float),
                              cudaMemcpyDeviceToDevice, st));
  }
  return step_device(h, batch == nullptr, st);
}

extern "C" int iqlb_agents_step_host(iqlb_agents* h, const float* host_batch, float* host_losses, void* stream) {
  IQLB_REQUIRE(h && host_batch && host_losses, "null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  const Layout& L = h->L;
  IQLB_CUDA(cudaMemcpyAsync(h->batch, host_batch, (size_t)h->desc.n_agents * L.B * L.R * sizeof(float),
                            cudaMemcpyHostToDevice, st));
  int rc = step_device(h, false, st);
  if (rc) return rc;
  IQLB_CUDA(cudaMemcpyAsync(host_losses, h->losses, (size_t)h->desc.n_agents * 3 * sizeof(float),
                            cudaMemcpyDeviceToHost, st));
  IQLB_CUDA(cudaStreamSynchronize(st));
  return IQLB_OK;
}

extern "C" int iqlb_agents_step_host_pipelined(iqlb_agents* h, const float* host_batch, float* host_losses_prev,
                                               void* stream) {
  IQLB_REQUIRE(h && host_losses_prev, "null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  const Layout& L = h->L;
  const size_t bytes = (size_t)h->desc.n_agents * L.B * L.R * sizeof(float);
  const size_t lbytes = (size_t)h->desc.n_agents * 3 * sizeof(float);
  if (!h->copy_stream) {
    IQLB_CUDA(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    IQLB_CUDA(cudaStreamCreateWithFlags(&h->d2h_stream, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) {
      IQLB_CUDA(cudaEventCreateWithFlags(&h->ev_h2d[i], cudaEventDisableTiming));
      IQLB_CUDA(cudaEventCreateWithFlags(&h->ev_done[i], cudaEventDisableTiming));
      IQLB_CUDA(cudaEventCreateWithFlags(&h->ev_step[i], cudaEventDisableTiming));
      IQLB_CUDA(cudaMalloc((void**)&h->stage[i], bytes));
      IQLB_CUDA(cudaMalloc((void**)&h->loss_snap[i], lbytes));
      IQLB_CUDA(cudaMallocHost((void**)&h->host_loss[i], lbytes));
    }
  }
  const int s = (int)(h->pipe_k & 1);
  if (host_batch != nullptr) {
    // stage[s] was last read by step k-2, whose completion the previous call waited for
    IQLB_CUDA(cudaMemcpyAsync(h->stage[s], host_batch, bytes, cudaMemcpyHostToDevice, h->copy_stream));
    IQLB_CUDA(cudaEventRecord(h->ev_h2d[s], h->copy_stream));
    IQLB_CUDA(cudaStreamWaitEvent(st, h->ev_h2d[s], 0));
    // precision 1 reads the staged rows in place (the packing and the loss kernel take the pointer); the fp32 path's
    // GEMM tasks are bound to h->batch
    if (h->desc.precision == 1) h->batch_src = h->stage[s];
    else IQLB_CUDA(cudaMemcpyAsync(h->batch, h->stage[s], bytes, cudaMemcpyDeviceToDevice, st));
    h->loss_snap_cur = h->loss_snap[s];
    int rc = step_device(h, false, st);
    h->batch_src = nullptr;
    h->loss_snap_cur = nullptr;
    if (rc) return rc;
    // the losses leave on the copy stream, so that the caller's stream carries kernels only (a copy node between the
    // optimizer kernel and the next step's first kernel would serialise their hand-over)
    // (read from the per-parity snapshot the loss kernel wrote: `losses` itself is rewritten by the next step)
    IQLB_CUDA(cudaEventRecord(h->ev_step[s], st));
    IQLB_CUDA(cudaStreamWaitEvent(h->d2h_stream, h->ev_step[s], 0));
    IQLB_CUDA(cudaMemcpyAsync(h->host_loss[s], h->loss_snap[s], lbytes, cudaMemcpyDeviceToHost, h->d2h_stream));
    IQLB_CUDA(cudaEventRecord(h->ev_done[s], h->d2h_stream));
  }
  const bool flush = host_batch == nullptr;
  const long long prev = flush ? h->pipe_k - 1 : h->pipe_k - 1;  // the step whose losses are returned
  if (prev >= 0) {
    const int ps = (int)(prev & 1);
    IQLB_CUDA(cudaEventSynchronize(h->ev_done[ps]));
    memcpy(host_losses_prev, h->host_loss[ps], lbytes);
  } else {
    memset(host_losses_prev, 0, lbytes);
  }
  if (!flush) h->pipe_k += 1;
  return IQLB_OK;
}

extern "C" int iqlb_agents_sync(iqlb_agents* h, void* stream) {
  IQLB_REQUIRE(h != nullptr, "null handle");
  return join_groups(h, (cudaStream_t)stream);
}

extern "C" int iqlb_agents_set_overlap(iqlb_agents* h, int32_t enable, void* stream) {
  IQLB_REQUIRE(h != nullptr, "null handle");
  h->overlap = enable ? 1 : 0;
  return enable ? IQLB_OK : join_groups(h, (cudaStream_t)stream);
}

extern "C" int iqlb_agents_launches_per_step(const iqlb_agents* h, int32_t sampled) {
  if (!h) return 0;
  if (h->launches_last > 0) return h->launches_last;  // counted by the last step
  // gather + fwd0 + fwd1 + head_fwd + loss + head_bwd + bw1 + bd1 + bw0 + adam
  return (sampled ? 1 : 0) + 5 + 4;
}

extern "C" int iqlb_agents_refresh(iqlb_agents* h, void* stream) {
  IQLB_REQUIRE(h != nullptr, "null handle");
  { int rcj = join_groups(h, (cudaStream_t)stream); if (rcj) return rcj; }
  if (h->pre_valid) {  // caller-owned state (parameters, RNG offsets) changed: the look-ahead batch is stale
    IQLB_CUDA(cudaStreamWaitEvent((cudaStream_t)stream, h->ev_pre_join, 0));
    h->pre_valid = false;
  }
  if (h->desc.precision == 1) return bf16_refresh(h, (cudaStream_t)stream);
  return IQLB_OK;
}

// max_action * tanh-head output, clamped (GaussianPolicy.act / DeterministicPolicy.act, iql.py:476-482, 506-514)
__global__ void __launch_bounds__(256) act_scale_clamp_kernel(float* __restrict__ a, int64_t n, float max_action) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) a[i] = fminf(fmaxf(max_action * a[i], -max_action), max_action);
}

// Batched deterministic policy inference, the device half of eval_actor (iql.py:304-319): every agent's actor in eval
// mode (no dropout) on its own block of states (or on one shared block), fp32 like the reference's evaluation.
extern "C" int iqlb_agents_act(iqlb_agents* h, const float* states, int64_t agent_stride, int32_t n_rows,
                               float max_action, float* actions, void* stream) {
  IQLB_REQUIRE(h && states && actions, "null pointer");
  IQLB_REQUIRE(n_rows >= 0 && agent_stride >= 0, "shape");
  if (n_rows == 0) return IQLB_OK;
  cudaStream_t st = (cudaStream_t)stream;
  { int rcj = join_groups(h, st); if (rcj) return rcj; }
  const Layout& L = h->L;
  const int N = h->desc.n_agents;
  float* hid = nullptr;       // H1 | H2, [N][n_rows][H] each
  GemmTask* d_tasks = nullptr;
  const int64_t hn = (int64_t)N * n_rows * L.H;
  IQLB_CUDA(cudaMallocAsync((void**)&hid, (size_t)(2 * hn) * sizeof(float), st));
  cudaError_t e = cudaMallocAsync((void**)&d_tasks, (size_t)(3 * N) * sizeof(GemmTask), st);
  if (e != cudaSuccess) { cudaFreeAsync(hid, st); return fail(IQLB_ERR_NOMEM, "iqlb_agents_act: %s", cudaGetErrorString(e)); }
  std::vector<GemmTask> tasks((size_t)3 * N);
  for (int a = 0; a < N; ++a) {
    const float* net = h->params + (int64_t)a * L.Pstride + L.off_a;
    float* h1 = hid + (int64_t)a * n_rows * L.H;
    float* h2 = hid + hn + (int64_t)a * n_rows * L.H;
    GemmTask t{};
    t.M = n_rows; t.scale = 1.f;
    t.A = states + (int64_t)a * agent_stride; t.lda = L.S; t.Bm = net + L.w0(); t.ldb = L.S; t.bias = net + L.b0(L.S);
    t.C = h1; t.ldc = L.H; t.N = L.H; t.K = L.S; t.act = IQLB_ACT_RELU;
    tasks[a] = t;
    t.A = h1; t.lda = L.H; t.Bm = net + L.w1(L.S); t.ldb = L.H; t.bias = net + L.b1(L.S); t.C = h2; t.K = L.H;
    tasks[N + a] = t;
    t.A = h2; t.Bm = net + L.w2(L.S); t.bias = net + L.b2(L.S, L.A); t.C = actions + (int64_t)a * n_rows * L.A;
    t.ldc = L.A; t.N = L.A; t.act = IQLB_ACT_TANH;
    tasks[2 * N + a] = t;
  }
  // (synchronous copy of a few KB: an evaluation call, not the training path)
  e = cudaMemcpy(d_tasks, tasks.data(), tasks.size() * sizeof(GemmTask), cudaMemcpyHostToDevice);
  int rc = e == cudaSuccess ? IQLB_OK : fail(IQLB_ERR_CUDA, "iqlb_agents_act: %s", cudaGetErrorString(e));
  if (!rc) rc = launch_grouped_gemm(GEMM_LINEAR, d_tasks, N, n_rows, L.H, false, st);
  if (!rc) rc = launch_grouped_gemm(GEMM_LINEAR, d_tasks + N, N, n_rows, L.H, false, st);
  if (!rc) rc = launch_grouped_gemm(GEMM_LINEAR, d_tasks + 2 * N, N, n_rows, L.A, true, st);
  if (!rc) {
    const int64_t tot = (int64_t)N * n_rows * L.A;
    act_scale_clamp_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(actions, tot, max_action);
    if (cudaGetLastError() != cudaSuccess) rc = fail(IQLB_ERR_CUDA, "act_scale_clamp_kernel launch failed%s", "");
  }
  cudaFreeAsync(hid, st);
  cudaFreeAsync(d_tasks, st);
  return rc;
}

extern "C" int iqlb_agents_set_actor_dropout(iqlb_agents* h, float p) {
  IQLB_REQUIRE(h != nullptr, "null handle");
  IQLB_REQUIRE(p >= 0.f && p < 1.f, "dropout probability must be in [0, 1)");
  h->drop_p = p;  // precision 1 reads it at every launch (bf16_step.cu bf16_set_dropout)
  return build_tasks(h);
}

extern "C" int iqlb_agents_debug_option(iqlb_agents* h, int32_t option, int32_t value) {
  IQLB_REQUIRE(h != nullptr, "null handle");
  IQLB_REQUIRE(option == 0, "unknown option");
  // option 0 selected the optimizer fused into the weight-gradient kernel (round 1; measured slower, removed in round 2):
  // only 0 - gradients materialised for the stand-alone optimizer kernel, the one schedule left - is accepted
  IQLB_REQUIRE(value == 0, "option 0 (fused optimizer) has been removed: value must be 0");
  return IQLB_OK;
}

extern "C" int iqlb_agents_profile(iqlb_agents* h, int32_t enable) {
  IQLB_REQUIRE(h != nullptr, "null handle");
  if (enable && !h->ev[0]) {
    for (int i = 0; i <= iqlb_agents::kMaxProf; ++i) IQLB_CUDA(cudaEventCreate(&h->ev[i]));
  }
  h->prof = enable != 0;
  if (enable) {
    for (int i = 0; i < iqlb_agents::kMaxProf; ++i) h->prof_ms[i] = 0.0;
    h->prof_steps = 0;
    h->prof_n = 0;
  }
  return IQLB_OK;
}

extern "C" int iqlb_agents_profile_collect(iqlb_agents* h) {
  IQLB_REQUIRE(h != nullptr && h->prof, "profiling is off");
  if (h->prof_n == 0) return IQLB_OK;
  IQLB_CUDA(cudaEventSynchronize(h->ev[h->prof_n]));
  for (int i = 0; i < h->prof_n; ++i) {
    float ms = 0.f;
    IQLB_CUDA(cudaEventElapsedTime(&ms, h->ev[i], h->ev[i + 1]));
    h->prof_ms[i] += ms;
  }
  h->prof_steps += 1;
  return IQLB_OK;
}

extern "C" int iqlb_agents_profile_read(iqlb_agents* h, int32_t i, const char** host_name, double* host_ms_total,
                                        int32_t* host_steps) {
  IQLB_REQUIRE(h != nullptr, "null handle");
  IQLB_REQUIRE(i >= 0 && i < h->prof_n, "kernel index out of range");
  if (host_name) *host_name = h->prof_name[i];
  if (host_ms_total) *host_ms_total = h->prof_ms[i];
  if (host_steps) *host_steps = h->prof_steps;
  return IQLB_OK;
}

extern "C" int iqlb_agents_debug_copy(iqlb_agents* h, int32_t which, void* dst, int64_t* nbytes, void* stream) {
  IQLB_REQUIRE(h != nullptr, "null handle");
  if (dst) { int rcj = join_groups(h, (cudaStream_t)stream); if (rcj) return rcj; }
  const Layout& L = h->L;
  const int64_t N = h->desc.n_agents, BH = (int64_t)L.B * L.H;
  const float* src = nullptr;
  int64_t n = 0;
  switch (which) {
    case 0: src = h->H1; n = N * kStreams * BH; break;
    case 1: src = h->H2; n = N * kStreams * BH; break;
    case 2: src = h->head_out; n = N * L.head_rows() * L.B; break;
    case 3: src = h->d_out; n = N * L.dout_rows() * L.B; break;
    case 4: src = h->dH2; n = N * kTrain * BH; break;
    case 5: src = h->dH1; n = N * kTrain * BH; break;
    case 6:
      n = N * L.Pstride;
      if (h->desc.precision == 1 && bf16_params(h)) {
        if (nbytes) *nbytes = n * (int64_t)sizeof(float);
        if (dst) {
          bf16_to_f32_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(bf16_params(h)->gradh, (float*)dst, n);
          IQLB_LAUNCH_CHECK();
        }
        return IQLB_OK;
      }
      src = h->grads;
      break;
    case 7: src = h->batch; n = N * L.B * L.R; break;
    case 8: src = bf16_params(h) ? (const float*)bf16_params(h)->dbg : nullptr; n = src ? 16384 : 0; break;
    default: return fail(IQLB_ERR_INVALID, "unknown debug tensor%s", "");
  }
  if (nbytes) *nbytes = n * (int64_t)sizeof(float);
  if (dst) IQLB_CUDA(cudaMemcpyAsync(dst, src, (size_t)n * sizeof(float), cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return IQLB_OK;
}
