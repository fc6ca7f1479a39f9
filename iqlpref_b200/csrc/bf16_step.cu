// bf16 tcgen05/TMEM path of the IQL step (precision == 1).
// Replaces the 7 MLP forwards and 4 MLP backwards of ImplicitQLearning.train
// (/root/reference/algorithms/offline/iql.py:581-637) with three tensor-core kernels:
//
//   bf16_fwd_persist_kernel (bf16_fwd.cu; tile = half-batch, stream, agent): X -> relu(X W0^T + b0) -> relu(H1 W1^T + b1)
//                           -> head (W2, b2, tanh for the actor); keeps H1, H2 (bf16 images) and the
//                           ReLU masks of the four trainable streams for the backward pass.
//   bf16_bwd_data_kernel    (half-batch, net, agent): dH2 = (dz W2) * relu'(H2) on CUDA cores,
//                           dH1 = (dH2 W1) * relu'(H1) on the tensor cores (W1 read MN-major).
//   bf16_bwd_weight_kernel  (128 output features, net, agent): dW1 = dH2^T H1, dW0 = dH1^T X,
//                           dW2 = dz^T H2, db1, db0 - all as MN-major x MN-major MMAs over the
//                           batch dimension (bias gradients ride on a ones column of the dz image).
//
// All operands are bf16 "images" (umma.cuh): identical bytes in HBM and in shared memory, moved by
// 1-D bulk copies.  Accumulation is fp32 in TMEM; parameters, Adam state and the Polyak target stay
// fp32 (iql_step.cu), and the Adam kernel refreshes the bf16 weight images it just updated.
#include "iql_step.cuh"
#include <stdlib.h>

#include "umma.cuh"
#include "gather_pack.cuh"

namespace iqlb {
using namespace umma;

// ------------------------------------------------------------------------------------------------
// params/target fp32 -> bf16 weight images (after adopt / load_state_dict; Adam keeps them fresh)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) shadow_refresh_kernel(BfParams P, const float* __restrict__ params,
                                                             const float* __restrict__ target) {
  const Layout& L = P.L;
  const int agent = blockIdx.y;
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= L.off_logstd) return;
  int net;
  int64_t off;
  if (!shadow_loc(L, P.KXB, e, net, off)) return;
  uint8_t* w = P.wimg + (int64_t)agent * P.agent_w_bytes;
  *reinterpret_cast<__nv_bfloat16*>(w + (int64_t)net * P.net_bytes + off) = __float2bfloat16_rn(params[(int64_t)agent * L.Pstride + e]);
  if (e < L.PQ)
    *reinterpret_cast<__nv_bfloat16*>(w + (int64_t)(net + 4) * P.net_bytes + off) =
        __float2bfloat16_rn(target[(int64_t)agent * L.PQstride + e]);
}

// ------------------------------------------------------------------------------------------------
// packed fp32 batch rows -> bf16 input images  X_sa = [obs|act|0..], X_next = [next_obs|0..]
// one thread per 16-byte chunk (8 columns)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) pack_x_kernel(BfParams P, const float* __restrict__ batch) {
  const Layout& L = P.L;
  const int agent = P.agent0 + blockIdx.y;
  const int chunks_per_row = P.KXB * 8;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  // Programmatic launch: nothing here depends on the predecessor (the previous step's Adam kernel) - the input images'
  // readers finished before it started - so the packing runs under its tail; the wait sits at the END, so that this
  // grid's completion still implies the predecessor's for the forward kernel (which reads the weight images).
  pdl_launch();
  if (i >= L.B * chunks_per_row) { pdl_wait(); return; }
  const int b = i / chunks_per_row, ch = i % chunks_per_row;
  const float* row = batch + ((int64_t)agent * L.B + b) * L.R;
  uint32_t sa[4], nx[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int c = ch * 8 + 2 * q;
    const float a0 = c < L.Dq ? row[c] : 0.f, a1 = c + 1 < L.Dq ? row[c + 1] : 0.f;
    const float n0 = c < L.S ? row[L.Dq + c] : 0.f, n1 = c + 1 < L.S ? row[L.Dq + c + 1] : 0.f;
    sa[q] = pack_bf16(a0, a1);
    nx[q] = pack_bf16(n0, n1);
  }
  const int64_t off = image_off(L.B, b, ch * 8);
  *reinterpret_cast<uint4*>(P.xsa + (int64_t)agent * P.x_bytes + off) = make_uint4(sa[0], sa[1], sa[2], sa[3]);
  *reinterpret_cast<uint4*>(P.xnext + (int64_t)agent * P.x_bytes + off) = make_uint4(nx[0], nx[1], nx[2], nx[3]);
  pdl_wait();
}

// ------------------------------------------------------------------------------------------------
// ReplayBuffer.sample (iql.py:211-221) fused with the image packing: Philox index, 128-bit row gather,
// fp32 batch rows (the loss kernel reads reward / done / action from them) and both bf16 input images
// in one pass.  A CTA stages 256 / vec_per_row rows in shared memory.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) replay_gather_pack_kernel(BfParams P, GatherArgs G) {
  extern __shared__ __align__(16) float srow[];  // [rows_per_cta][R]
  pdl_wait();  // first kernel of a step: the previous step's Adam (launched before it) must be complete
  pdl_launch();
  gather_pack_body(P, G, (int)blockIdx.x, P.agent0 + (int)blockIdx.y, srow);
}

__device__ __forceinline__ void bulk_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }

// store 8 consecutive columns (one 16-byte chunk) of tile row `row` into a tile image with 128 rows
__device__ __forceinline__ void st_chunk(uint8_t* tile, int row, int col0, const float* v) {
  const int off = (col0 >> 6) * 16384 + row * 128 + ((((col0 & 63) >> 3) ^ (row & 7)) << 4);
  *reinterpret_cast<uint4*>(tile + off) =
      make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
}

// ------------------------------------------------------------------------------------------------
// Forward: bf16_fwd_persist_kernel (bf16_fwd.cu) - persistent, warp-specialised, one CTA per SM.  (Round 1's one-tile-per-CTA
// kernel lived here; the persistent kernel now covers every shape bf16_create accepts - pen-human forward 118 -> 75 us,
// profiles/r02f_fwd_wide_ab.jsonl - so it was removed.)
// ------------------------------------------------------------------------------------------------
// ------------------------------------------------------------------------------------------------
// Backward, data gradients.  grid (B/128, 4 trainable nets, agents), 256 threads.
// Shared memory: [dH2 tile 64K | W1 image 128K | 2 x 16K output staging (first: W2 fp32) | ctrl].
// dH2[b,j] = relu'(H2[b,j]) * sum_o dz[b,o] W2[o,j] on CUDA cores (thread = row x column half), then
// dH1[:, 64*ib..] = dH2 * W1[:, 64*ib..] as four N=64 MMAs (W1 image read MN-major: N = in, K = out),
// each drained (mask, bf16) while the next one runs.  Both tiles go to HBM as images for the weight kernel.
// ------------------------------------------------------------------------------------------------
// 16 worker warps (four per TMEM lane quarter) + one issuing warp.  With eight workers the CUDA-core dH2 = (dz W2) * relu'
// tile took 4.3 k of a CTA's 12 k clocks at two warps per scheduler.  Round 2: the tile is produced K-block by K-block (64 of
// its 256 columns at a time, every worker 16 of them) and the issuing thread multiplies block kb by W1 (N = 256, 4 k-steps)
// while the workers compute block kb + 1, so the tensor-core product hides under the CUDA-core one instead of following it.
constexpr int kBdWorkerWarps = 16, kBdWorkerThreads = kBdWorkerWarps * 32, kBdThreads = kBdWorkerThreads + 32;
__device__ __forceinline__ void bwd_data_body(const BfParams& P, const float* __restrict__ params,
                                              const float* __restrict__ d_out, const int cta_lin, const int n_ag,
                                              const int bx, const int by, const int bz) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const Layout& L = P.L;
  // Longest tiles first: the actor's dH2 = dz W2 has out_dim columns of dz per row (8-24) against one for V / Q, and
  // its CTAs run ~14.5 us against ~8 us - launched in (agent, net) order they end up as the kernel's tail
  // (per-SM CTA timeline, scripts/timeline.py).  CTAs are scheduled in linear block order: actor tiles take the front.
  int half, t, aidx;
  if (P.opt_lpt) {
    if (cta_lin < 2 * n_ag) { half = cta_lin & 1; t = 3; aidx = cta_lin >> 1; }
    else { const int r = cta_lin - 2 * n_ag; half = r & 1; t = (r >> 1) % 3; aidx = (r >> 1) / 3; }
  } else {
    half = bx; t = by; aidx = bz;
  }
  const int agent = P.agent0 + (P.rev ? n_ag - 1 - aidx : aidx);
  const int s = train_stream(t), net = net_of_stream(s);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, wu = warp_uniform();  // wu: single-thread roles only (elect_one(), umma.cuh)
  const int quarter = warp & 3, cgrp = (warp >> 2) & 3;  // column group: 64 columns of the dH2 tile, 16 of every dH1 block
  const uint32_t sbase = smem_u32(smem);
  const uint32_t oA = 0, oW = 65536u, oStage = oW + 131072u, oCtrl = oStage + 32768u;
  const uint32_t bar_w = sbase + oCtrl, bar_mma = bar_w + 32, bar_k = bar_w + 72;  // (tmem slot at + 64)
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + oCtrl + 64);
  const int in = L.in_dim(s), nout = L.out_dim(s);
  const uint8_t* w1img = P.wimg + (int64_t)agent * P.agent_w_bytes + (int64_t)net * P.net_bytes + P.w0_bytes;
  const float* fnet = params + (int64_t)agent * L.Pstride + L.net_off(s);
  const int64_t at = (int64_t)agent * kTrain + t;

  // actor dropout: both hidden layers' gradients pass through a kept-and-scaled mask (the stored bit masks already are
  // relu' AND keep); exactly 1 when dropout is off
  const float gscale = (t == 3 && P.drop_thresh != 0) ? P.drop_scale : 1.0f;
  [[maybe_unused]] const bool dbgd = P.dbg != nullptr && half == 0 && t == 1 && aidx == P.dbg_z && tid == 64;
#ifdef IQLB_NO_TIMELINE
#define IQLB_TSD(slot) do { } while (0)
#else
#define IQLB_TSD(slot) do { if (dbgd) P.dbg[2560 + (slot)] = (unsigned long long)clock64(); } while (0)
#endif
  IQLB_TSD(0);
  if (P.dbg != nullptr && tid == 0 && cta_lin < 1024) {  // (SM, start) of every CTA: scripts/timeline.py --waves
    unsigned smid; unsigned long long gt;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
    P.dbg[4096 + cta_lin * 3] = smid; P.dbg[4096 + cta_lin * 3 + 1] = gt;
  }
  if (wu == 0 && elect_one()) {
    if (sbase & 1023u) __trap();
    for (int i = 0; i < 4; ++i) { mbar_init(bar_w + 8 * i, 1); mbar_init(bar_mma + 8 * i, 1); mbar_init(bar_k + 8 * i, kBdWorkerWarps); }
    fence_barrier_init();
    for (int ib = 0; ib < 4; ++ib) {
      mbar_expect_tx(bar_w + 8 * ib, 32768u);
      bulk_g2s(sbase + oW + (uint32_t)ib * 32768u, w1img + (int64_t)ib * 32768, 32768u, bar_w + 8 * ib);
    }
  }
  if (warp == 1) tmem_alloc(sbase + oCtrl + 64, 256);
  // the worker warps compute and drain; the first lanes of the two extra warps issue the dH2 store and the MMAs, so the
  // drain of block ib overlaps the issue of block ib+1 instead of waiting for the issuing thread's warp
  const bool worker = warp < kBdWorkerWarps;
  float* sw2 = reinterpret_cast<float*>(smem + oStage);
  if (worker)
    for (int i = tid; i < nout * 256; i += kBdWorkerThreads) sw2[i] = __ldg(fnet + L.w2(in) + i);
  const int row = quarter * 32 + lane, brow = half * 128 + row;
  // this thread's 16 columns of K-block kb of the dH2 tile: bits (cgrp & 1) * 16 .. + 15 of mask word 2 kb + cgrp / 2
  const uint32_t* m2p = P.mask2 + (at * L.B + brow) * (L.H / 32);
  uint64_t m2h = 0;
#pragma unroll
  for (int kb = 0; kb < 4; ++kb)
    m2h |= (uint64_t)(worker ? (m2p[kb * 2 + (cgrp >> 1)] >> ((cgrp & 1) * 16)) & 0xffffu : 0u) << (16 * kb);
  const uint32_t* m1p = P.mask1 + (at * L.B + brow) * (L.H / 32);
  uint32_t m1w[4];  // this thread's 16 columns of block ib: bits (cgrp & 1) * 16 .. + 15 of word 2 ib + cgrp / 2
#pragma unroll
  for (int ib = 0; ib < 4; ++ib) m1w[ib] = worker ? (m1p[ib * 2 + (cgrp >> 1)] >> ((cgrp & 1) * 16)) & 0xffffu : 0u;
  // everything above reads the forward kernel's and Adam's outputs only; dz comes from the loss kernel
  pdl_wait();
  pdl_launch();
  float dz[kOutMax];
  {
    const float* dout = d_out + (int64_t)agent * L.dout_rows() * L.B;
    const int drow0 = (t == 3) ? 3 : t;
#pragma unroll
    for (int o = 0; o < kOutMax; ++o) dz[o] = (worker && o < nout) ? dout[(int64_t)(drow0 + o) * L.B + brow] : 0.f;
  }
  IQLB_TSD(1);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  IQLB_TSD(2);
  const uint32_t tmem = *tmem_slot;
  if (wu < kBdWorkerWarps) {  // dH2 tile, one 64-column K-block at a time: this thread's row, columns [kb * 64 + cgrp * 16, + 16)
    // (not unrolled: copies of the 24-output body would not fit the instruction cache)
#pragma unroll 1
    for (int kb = 0; kb < 4; ++kb) {
      const uint32_t mh = (uint32_t)(m2h >> (16 * kb)) & 0xffffu;
#pragma unroll 1
      for (int hh = 0; hh < 2; ++hh) {
        const int c0 = kb * 64 + cgrp * 16 + hh * 8;
        float g[8];
        if (nout == 1) {
          const float4 w0 = *reinterpret_cast<const float4*>(sw2 + c0), w1 = *reinterpret_cast<const float4*>(sw2 + c0 + 4);
          g[0] = dz[0] * w0.x; g[1] = dz[0] * w0.y; g[2] = dz[0] * w0.z; g[3] = dz[0] * w0.w;
          g[4] = dz[0] * w1.x; g[5] = dz[0] * w1.y; g[6] = dz[0] * w1.z; g[7] = dz[0] * w1.w;
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) g[j] = 0.f;
          // groups of eight outputs behind a uniform branch: an 8-action actor issues one group, not three
          // (predicated-off instructions still take issue slots)
#pragma unroll
          for (int ob = 0; ob < kOutMax; ob += 8) {
            if (ob < nout) {
#pragma unroll
              for (int oo = 0; oo < 8; ++oo) {
                const int o = ob + oo;
                if (oo == 0 || o < nout) {
                  const float4 w0 = *reinterpret_cast<const float4*>(sw2 + o * 256 + c0),
                               w1 = *reinterpret_cast<const float4*>(sw2 + o * 256 + c0 + 4);
                  g[0] = fmaf(dz[o], w0.x, g[0]); g[1] = fmaf(dz[o], w0.y, g[1]); g[2] = fmaf(dz[o], w0.z, g[2]);
                  g[3] = fmaf(dz[o], w0.w, g[3]); g[4] = fmaf(dz[o], w1.x, g[4]); g[5] = fmaf(dz[o], w1.y, g[5]);
                  g[6] = fmaf(dz[o], w1.z, g[6]); g[7] = fmaf(dz[o], w1.w, g[7]);
                }
              }
            }
          }
        }
        const uint32_t m = mh >> (8 * hh);
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] = ((m >> j) & 1u) ? g[j] : 0.f;
        if (gscale != 1.0f) {
#pragma unroll
          for (int j = 0; j < 8; ++j) g[j] *= gscale;
        }
        st_chunk(smem + oA, row, c0, g);
      }
      fence_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_k + 8 * kb);  // K-block kb of this warp's rows is in shared memory
    }
    IQLB_TSD(3);
  } else if (elect_one()) {  // the issuing thread
    for (int ib = 0; ib < 4; ++ib) mbar_wait(bar_w + 8 * ib, 0);
    for (int kb = 0; kb < 4; ++kb) {
      mbar_wait(bar_k + 8 * kb, 0);
      tc_fence_after();
      gemm_issue(tmem, operand_k_major(sbase + oA + (uint32_t)kb * 16384u, 128, 0), operand_mn_major(sbase + oW, L.H, 0, 64 * kb), 256, 4,
                 kb > 0);
    }
    mma_commit(bar_mma);
    uint8_t* dh2 = P.dh2 + at * P.h_bytes;  // the finished dH2 tile also goes to HBM for the weight-gradient kernel
    for (int cb = 0; cb < 4; ++cb)
      bulk_s2g(dh2 + (int64_t)cb * L.B * 128 + (int64_t)half * 16384, sbase + oA + (uint32_t)cb * 16384u, 16384u);
    bulk_commit();
    bulk_wait_read0();
  }
  __syncwarp();
  if (worker) {
    mbar_wait(bar_mma, 0);
    tc_fence_after();
    IQLB_TSD(4);
  }
  const uint32_t lane_base = tmem + ((uint32_t)(quarter * 32) << 16);
  uint8_t* dh1 = P.dh1 + at * P.h_bytes;
#pragma unroll
  for (int ib = 0; worker && ib < 4; ++ib) {
    IQLB_TSD(5 + ib);
    uint8_t* stage = smem + oW + ib * 32768;  // W1 block ib has been consumed by MMA ib: its space stages dH1 block ib
    float v[16];
    tmem_ld16(lane_base + 64u * ib + 16u * cgrp, v);
    tmem_ld_wait();
    const uint32_t m = m1w[ib];
#pragma unroll
    for (int j = 0; j < 16; ++j) v[j] = ((m >> j) & 1u) ? v[j] : 0.f;
    if (gscale != 1.0f) {
#pragma unroll
      for (int j = 0; j < 16; ++j) v[j] *= gscale;
    }
#pragma unroll
    for (int q = 0; q < 2; ++q) st_chunk(stage, row, 16 * cgrp + 8 * q, v + 8 * q);
    fence_async_smem();
    asm volatile("bar.sync 1, %0;" ::"n"(kBdWorkerThreads) : "memory");
    if (wu == 0 && elect_one()) {
      bulk_s2g(dh1 + (int64_t)ib * L.B * 128 + (int64_t)half * 16384, sbase + oW + (uint32_t)ib * 32768u, 16384u);
      bulk_commit();
    }
  }
  IQLB_TSD(9);
  if (wu == 0 && elect_one()) bulk_wait_read0();
  if (P.dbg != nullptr && tid == 0 && cta_lin < 1024) {
    unsigned long long gt;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
    P.dbg[4096 + cta_lin * 3 + 2] = gt;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 256);
}

__global__ void __launch_bounds__(kBdThreads, 1) bf16_bwd_data_kernel(BfParams P, const float* __restrict__ params,
                                                               const float* __restrict__ d_out) {
  bwd_data_body(P, params, d_out, (int)blockIdx.x + 2 * ((int)blockIdx.y + kTrain * (int)blockIdx.z), (int)gridDim.z,
                (int)blockIdx.x, (int)blockIdx.y, (int)blockIdx.z);
}

// ------------------------------------------------------------------------------------------------
// Backward, weight gradients.  Tiles = (H/128 output-feature tiles, 4 nets, agents), one persistent CTA of 544 threads per SM.
// Every product reduces over the batch (K = B), with both operands read MN-major from images whose
// rows are batch rows.  blk = B*128 bytes = one 64-column block of a batch-row image.
// Shared memory: [bufA 2 blk | bufB 4 blk | dz image 1 blk | ctrl]
//   phase 1: bufA = dH2[:, tile], bufB = H1          -> dW1 tile (TMEM 0..255), db1 (col 448 or 480 + out: clear of a 128-column dW0 tile)
//   phase 2: bufA = dH1[:, tile], bufB[0..] = X_sa   -> dW0 tile (TMEM 256..), db0 (col 384+out)
//            H2[:, tile] (2 blk of bufB)             -> dW2^T tile (TMEM 416 + o)
// The dz image holds dz[b, 0..out) and a column of ones at `out`, so bias gradients are MMA columns.
// Gradients leave as bf16 (what the reference's autocast weight-gradient GEMMs emit) for the stand-alone optimizer kernel.
//   warps 0-15  drain: 4 warps per TMEM lane quarter (16 of every 64 columns each) - the round-1 kernel (8 warps that also
//               issued the MMAs) spent 5 of its 10 us per CTA moving the 128 x 256 dW1 tile TMEM -> shared memory -> global
//               at 12 % warp occupancy;
//   warp 16     one elected thread does every bulk copy and every MMA issue, so the phase-2 products (dW0, db0, and with
//               one input block also dW2) are loaded and ISSUED under the dW1 drain instead of after it.
// With one input block (state_dim + action_dim <= 64) the H2 tile is requested right after phase 1 into the two H1 blocks
// behind X and the dW1 drain stages through one 32 KB buffer (H1 block 3); with two input blocks H2 follows the drain.
// Measured (scripts/gpu_bw_ab.sh, same box): CTA 9.9 -> 8.0 us, kernel 53.7 -> 45.9 us.  Not kept: the dW1 product as four
// 64-column MMAs with a barrier each, so that the drain starts after the first (the products are shared-memory bound, the
// A tile is re-read per block and the phase-2 operands arrive later: CTA 8.4 us); Adam / Polyak fused into this kernel's
// epilogue (round 1: with one CTA per SM its HBM streaming is serialised behind operand loads and MMAs, 276 vs 55 + 119 us).
// ------------------------------------------------------------------------------------------------
constexpr int kBwDrainThreads = 512, kBwThreads = kBwDrainThreads + 32;

__device__ __forceinline__ void bw_bar(int id) { asm volatile("bar.sync %0, %1;" ::"r"(id), "n"(kBwDrainThreads) : "memory"); }

// the 16 drain warps of bf16_bwd_weight_kernel (ph = parity of the tile count, d1col = TMEM column of this tile's db1,
// bar_d2 = "final drain done")
__device__ __forceinline__ void bwd_weight_drain_p(const BfParams& P, uint8_t* smem, const uint32_t tmem, const uint32_t bar_p1,
                                                   const uint32_t bar_p2, const uint32_t bar_d1, const uint32_t bar_d2,
                                                   const uint32_t ph, const uint32_t d1col, const int ot, const int t,
                                                   const int agent, const int in, const int nout, const int n0, const int warp,
                                                   const int lane, const int quarter, const int cgrp, const int tid,
                                                   const int tile) {
  const Layout& L = P.L;
  const int s = train_stream(t);
  const uint32_t blk = (uint32_t)L.B * 128u;
  const uint32_t oB = 2 * blk;
  const int trow = quarter * 32 + lane;  // tile row (= output feature ot * 128 + trow) this thread drains
  const uint32_t lane_base = tmem + ((uint32_t)(quarter * 32) << 16);
  const int64_t noff = L.net_off(s);
  __nv_bfloat16* gagent = P.gradh + (int64_t)agent * L.Pstride;
  const bool persist = bar_d2 != 0;
  if (P.dbg != nullptr && tid == 0 && tile < 512) {  // (SM, start) of every tile: scripts/timeline.py
    unsigned smid; unsigned long long gt;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
    P.dbg[4096 + 1536 + tile * 4] = smid; P.dbg[4096 + 1536 + tile * 4 + 1] = gt; P.dbg[4096 + 1536 + tile * 4 + 3] = (unsigned)t;
  }
#ifdef IQLB_NO_TIMELINE
#define IQLB_TSWD(k) do { } while (0)
#else
#define IQLB_TSWD(k) do { if (P.dbg != nullptr && blockIdx.x == 0 && tid == 0 && tile / (int)gridDim.x < 8) P.dbg[3200 + (tile / (int)gridDim.x) * 16 + (k)] = (unsigned long long)clock64(); } while (0)
#endif
  mbar_wait(bar_p1, ph);
  IQLB_TSWD(9);
  tc_fence_after();
  {
    // ---- dW1: four 64-column chunks through one [128 rows][64] fp32 buffer (float4 index ^ (row & 15)); bf16 rows leave
    //      as 128-bit stores, 8 lanes per 128-byte row segment
    float* sb = reinterpret_cast<float*>(smem + oB + 3 * blk);
    const int rr = lane >> 3, l8 = lane & 7;
#pragma unroll 1
    for (int cc = 0; cc < 4; ++cc) {
      float v[16];
      tmem_ld16(lane_base + (uint32_t)(cc * 64 + cgrp * 16), v);
      tmem_ld_wait();
      if (cc > 0 || persist) bw_bar(2);  // every warp has read chunk cc - 1 (persistent: the previous tile's staged gradients) out of the buffer
#pragma unroll
      for (int q = 0; q < 4; ++q)
        *reinterpret_cast<float4*>(sb + trow * 64 + (((cgrp * 4 + q) ^ (trow & 15)) << 2)) =
            make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
      bw_bar(1);  // chunk cc staged
      __nv_bfloat16* gw1 = gagent + noff + L.w1(in) + (int64_t)(ot * 128) * L.H + cc * 64 + 8 * l8;
#pragma unroll
      for (int it = 0; it < 2; ++it) {  // bf16, as the reference's autocast weight-gradient GEMM emits them
        const int r = warp * 8 + it * 4 + rr;
        const float4 g0 = *reinterpret_cast<const float4*>(sb + r * 64 + (((2 * l8) ^ (r & 15)) << 2));
        const float4 g1 = *reinterpret_cast<const float4*>(sb + r * 64 + (((2 * l8 + 1) ^ (r & 15)) << 2));
        *reinterpret_cast<uint4*>(gw1 + (int64_t)r * L.H) =
            make_uint4(pack_bf16(g0.x, g0.y), pack_bf16(g0.z, g0.w), pack_bf16(g1.x, g1.y), pack_bf16(g1.z, g1.w));
      }
    }
  }
  if (P.KXB != 1 || persist) {  // the staging buffer (and, for the next tile, TMEM columns 0..255) may be reused
    fence_async_smem();
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_d1);
  }
  IQLB_TSWD(10);
  mbar_wait(bar_p2, ph);
  IQLB_TSWD(12);
  tc_fence_after();
  // ---- everything else through shared memory in packed-parameter order (the operands of phase 2 are dead):
  //      G0 = dW0 tile [128][in] | G1 = db1[128] | G2 = db0[128] | G3 = dW2[nout][128]
  // each a contiguous, 16-byte aligned span of the parameter vector -> one flat 128-bit pass
  float* stg = reinterpret_cast<float*>(smem + oB + 2 * blk);
  const int g1 = 128 * in, g2 = g1 + 128, g3 = g2 + 128, gtot = g3 + 128 * nout;
  const int64_t e_w0 = noff + L.w0() + (int64_t)ot * 128 * in, e_b1 = noff + L.b1(in) + ot * 128,
                e_b0 = noff + L.b0(in) + ot * 128, e_w2 = noff + L.w2(in) + ot * 128;
  auto flat_elem = [&](int i) -> int64_t {
    if (i < g1) return e_w0 + i;
    if (i < g2) return e_b1 + (i - g1);
    if (i < g3) return e_b0 + (i - g2);
    return e_w2 + (int64_t)((i - g3) >> 7) * L.H + ((i - g3) & 127);
  };
  if (cgrp < 2) {
    for (int c0 = cgrp * 16; c0 < n0; c0 += 32) {
      float v[16];
      tmem_ld16(lane_base + 256u + (uint32_t)c0, v);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 16; ++j)
        if (c0 + j < in) stg[trow * in + c0 + j] = v[j];
    }
  } else if (cgrp == 2) {  // the bias gradients ride on the ones column (index nout) of the dz image
    float v[32];
    tmem_ld32(lane_base + 384u, v);
    tmem_ld_wait();
    float db = 0.f;
#pragma unroll
    for (int j = 0; j < 32; ++j) db = (j == nout) ? v[j] : db;
    stg[g2 + trow] = db;
    tmem_ld32(lane_base + d1col, v);
    tmem_ld_wait();
    db = 0.f;
#pragma unroll
    for (int j = 0; j < 32; ++j) db = (j == nout) ? v[j] : db;
    stg[g1 + trow] = db;
  } else {
    float v[32];
    tmem_ld32(lane_base + 416u, v);
    tmem_ld_wait();
#pragma unroll
    for (int o = 0; o < kOutMax; ++o)
      if (o < nout) stg[g3 + o * 128 + trow] = v[o];
  }
  bw_bar(1);
  for (int i = tid * 4; i < gtot; i += 4 * kBwDrainThreads) {
    const float4 g = *reinterpret_cast<const float4*>(stg + i);
    *reinterpret_cast<uint2*>(gagent + flat_elem(i)) = make_uint2(pack_bf16(g.x, g.y), pack_bf16(g.z, g.w));
  }
  if (persist) {  // staging area and TMEM columns 256.. are free for the next tile
    fence_async_smem();
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_d2);
    if (P.dbg != nullptr && tid == 0 && tile < 512) {
      unsigned long long gt;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
      P.dbg[4096 + 1536 + tile * 4 + 2] = gt;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// The kernel is persistent: one CTA per SM walks the tiles (ot, net, agent) round-robin (253.2 -> 251.4 us per step against
// one CTA per tile).  What the loop adds:
//   * the next tile's phase-1 operands are requested as soon as the current tile's last product has completed (dH2, dz and
//     the first two H1 blocks at once, H1 blocks 2-3 when the final drain has left the staging area), i.e. under the final
//     drain instead of behind a CTA exit / launch (1.2 us) + barrier / TMEM set-up;
//   * db1 alternates between TMEM columns 480 and 448, so phase 1 of tile i+1 never waits for the final drain of tile i.
// Every barrier completes once per tile: parity = tile count & 1.
// ------------------------------------------------------------------------------------------------
struct BwTile {
  int ot, t, agent, s, in, nout, n0;
  int64_t at;
};
__device__ __forceinline__ BwTile bw_tile(const BfParams& P, int tile, int n_ag) {
  BwTile T;
  T.ot = tile & 1;
  T.t = (tile >> 1) & 3;
  const int aidx = tile >> 3;
  T.agent = P.agent0 + (P.rev ? n_ag - 1 - aidx : aidx);
  T.s = train_stream(T.t);
  T.in = P.L.in_dim(T.s);
  T.nout = P.L.out_dim(T.s);
  T.n0 = (T.in + 15) / 16 * 16;
  T.at = (int64_t)T.agent * kTrain + T.t;
  return T;
}

__global__ void __launch_bounds__(kBwThreads, 1) bf16_bwd_weight_kernel(BfParams P, int n_tiles) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const Layout& L = P.L;
  const int n_ag = n_tiles >> 3;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, wu = warp_uniform();
  const int quarter = warp & 3, cgrp = (warp >> 2) & 3;
  const uint32_t blk = (uint32_t)L.B * 128u;
  const uint32_t sbase = smem_u32(smem);
  const uint32_t oA = 0, oB = 2 * blk, oZ = 6 * blk, oCtrl = 7 * blk;
  const uint32_t bar_ld1 = sbase + oCtrl, bar_ld2 = bar_ld1 + 8, bar_ld3 = bar_ld1 + 16, bar_p1 = bar_ld1 + 24,
                 bar_p2 = bar_ld1 + 32, bar_d1 = bar_ld1 + 40, bar_d2 = bar_ld1 + 48;
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + oCtrl + 56);
  const int ksteps = L.B / 16;
  const bool preload_h2 = P.KXB == 1;
  const uint32_t oH2 = preload_h2 ? oB + blk : oB + 2 * blk;
  const bool issuer_warp = wu == kBwDrainThreads / 32;

  if (issuer_warp && elect_one()) {
    if (sbase & 1023u) __trap();
    for (int i = 0; i < 5; ++i) mbar_init(bar_ld1 + 8 * i, 1);
    mbar_init(bar_d1, kBwDrainThreads / 32);
    mbar_init(bar_d2, kBwDrainThreads / 32);
    fence_barrier_init();
    if ((int)blockIdx.x < n_tiles) {
      const BwTile T = bw_tile(P, blockIdx.x, n_ag);
      mbar_expect_tx(bar_ld1, 7 * blk);
      bulk_g2s(sbase + oB, P.h1 + T.at * P.h_bytes, 4 * blk, bar_ld1);  // forward output: before the PDL wait
      bulk_g2s(sbase + oZ, P.dz + T.at * P.z_bytes, blk, bar_ld1);      // loss output, two kernels back
      pdl_wait();  // dH2 / dH1 come from the data-gradient kernel
      bulk_g2s(sbase + oA, P.dh2 + T.at * P.h_bytes + (int64_t)(2 * T.ot) * blk, 2 * blk, bar_ld1);
    }
  }
  if (warp == 1) tmem_alloc(sbase + oCtrl + 56, 512);
  pdl_wait();
  pdl_launch();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
#ifdef IQLB_NO_TIMELINE
#define IQLB_TSW(slot) do { } while (0)
#else  // clock64 stamps of CTA 0, 16 slots per tile (issuing thread 0..7, drain thread 0 8..11): scripts/timeline.py
#define IQLB_TSW(slot) do { if (P.dbg != nullptr && blockIdx.x == 0 && (slot) < 128) P.dbg[3200 + (slot)] = (unsigned long long)clock64(); } while (0)
#endif

  if (issuer_warp) {
    if (elect_one()) {
      const Operand z = operand_mn_major(sbase + oZ, L.B, 0, 0);
      const Operand a = operand_mn_major(sbase + oA, L.B, 0, 0);
      int it = 0;
      for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
        const uint32_t ph = (uint32_t)(it & 1);
        const BwTile T = bw_tile(P, tile, n_ag);
        const uint32_t d1col = (it & 1) ? 448u : 480u;
        // phase 1: dW1 tile = dH2[:, ot]^T H1 (128 x 256), db1 = dH2[:, ot]^T [dz | 1]
        IQLB_TSW(it * 16 + 0);
        mbar_wait(bar_ld1, ph);
        IQLB_TSW(it * 16 + 1);
        if (it > 0) mbar_wait(bar_d1, ph ^ 1u);  // the previous tile's dW1 has left TMEM columns 0..255
        IQLB_TSW(it * 16 + 2);
        tc_fence_after();
        gemm_issue(tmem, a, operand_mn_major(sbase + oB, L.B, 0, 0), 256, ksteps, false);
        gemm_issue(tmem + d1col, a, z, 32, ksteps, false);
        mma_commit(bar_p1);
        mbar_wait(bar_p1, ph);  // dH2 and H1 consumed: their space takes the phase-2 operands
        IQLB_TSW(it * 16 + 3);
        mbar_expect_tx(bar_ld2, 2 * blk + (uint32_t)P.KXB * blk);
        bulk_g2s(sbase + oA, P.dh1 + T.at * P.h_bytes + (int64_t)(2 * T.ot) * blk, 2 * blk, bar_ld2);
        bulk_g2s(sbase + oB, P.xsa + (int64_t)T.agent * P.x_bytes, (uint32_t)P.KXB * blk, bar_ld2);
        if (preload_h2) {
          mbar_expect_tx(bar_ld3, 2 * blk);
          bulk_g2s(sbase + oH2, P.h2 + T.at * P.h_bytes + (int64_t)(2 * T.ot) * blk, 2 * blk, bar_ld3);
        }
        // phase 2 (under the dW1 drain): dW0 tile = dH1[:, ot]^T X, db0 = dH1[:, ot]^T [dz | 1], dW2 = H2[:, ot]^T [dz | 1]
        mbar_wait(bar_ld2, ph);
        IQLB_TSW(it * 16 + 4);
        if (it > 0) mbar_wait(bar_d2, ph ^ 1u);  // the previous tile's final drain has read TMEM columns 256..447
        tc_fence_after();
        gemm_issue(tmem + 256, a, operand_mn_major(sbase + oB, L.B, 0, 0), T.n0, ksteps, false);
        gemm_issue(tmem + 384, a, z, 32, ksteps, false);
        if (!preload_h2) {  // the H2 tile overlaps the drain's staging buffer: wait for the drain
          mbar_wait(bar_d1, ph);
          mbar_expect_tx(bar_ld3, 2 * blk);
          bulk_g2s(sbase + oH2, P.h2 + T.at * P.h_bytes + (int64_t)(2 * T.ot) * blk, 2 * blk, bar_ld3);
        }
        mbar_wait(bar_ld3, ph);
        IQLB_TSW(it * 16 + 5);
        tc_fence_after();
        gemm_issue(tmem + 416, operand_mn_major(sbase + oH2, L.B, 0, 0), z, 32, ksteps, false);
        mma_commit(bar_p2);
        // the next tile's phase-1 operands, under this tile's final drain
        const int next = tile + (int)gridDim.x;
        if (next < n_tiles) {
          const BwTile Nx = bw_tile(P, next, n_ag);
          mbar_wait(bar_p2, ph);  // every operand of this tile is dead
          IQLB_TSW(it * 16 + 6);
          mbar_expect_tx(bar_ld1, 7 * blk);
          bulk_g2s(sbase + oA, P.dh2 + Nx.at * P.h_bytes + (int64_t)(2 * Nx.ot) * blk, 2 * blk, bar_ld1);
          bulk_g2s(sbase + oZ, P.dz + Nx.at * P.z_bytes, blk, bar_ld1);
          bulk_g2s(sbase + oB, P.h1 + Nx.at * P.h_bytes, 2 * blk, bar_ld1);
          mbar_wait(bar_d2, ph);  // the final drain has read its staging area (H1 blocks 2-3)
          IQLB_TSW(it * 16 + 7);
          bulk_g2s(sbase + oB + 2 * blk, P.h1 + Nx.at * P.h_bytes + 2 * (int64_t)blk, 2 * blk, bar_ld1);
        }
      }
    }
  } else {
    int it = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
      const BwTile T = bw_tile(P, tile, n_ag);
      if (tid == 0) IQLB_TSW(it * 16 + 8);
      bwd_weight_drain_p(P, smem, tmem, bar_p1, bar_p2, bar_d1, bar_d2, (uint32_t)(it & 1), (it & 1) ? 448u : 480u, T.ot, T.t,
                         T.agent, T.in, T.nout, T.n0, warp, lane, quarter, cgrp, tid, tile);
      if (tid == 0) IQLB_TSW(it * 16 + 11);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static size_t bwd_data_smem() { return 65536u + 131072u + 32768u + 128u; }
static size_t bwd_weight_smem(int B) { return 7u * (size_t)B * 128u + 64u; }

int bf16_fwd_persist_launch(const BfParams& P, const float* params, const float* target, float* head_out, int n_agents,
                            cudaStream_t stream, bool pdl);

int bf16_bwd_data_persist_launch(const BfParams& P, const float* params, const float* d_out, int n_agents, int n_sm,
                                 cudaStream_t stream, bool pdl);

int bf16_create(iqlb_agents* h) {
  const Layout& L = h->L;
  IQLB_REQUIRE(L.H == 256, "the tcgen05 path is built for hidden_dim == 256");
  IQLB_REQUIRE(L.B == 256, "the tcgen05 path is built for batch_size 256 (every reference config); use precision fp32 otherwise");
  // (125, not 128: the weight-gradient kernel stages dW0 | db1 | db0 | dW2 of a 128-row tile, (128 in + 256 + 128 out) floats,
  //  in the 64 KB behind the X image; Ant-v2, 111 + 8, is the widest D4RL shape this covers)
  IQLB_REQUIRE(L.A <= kOutMax && L.Dq <= 125, "action_dim <= 24 and state_dim + action_dim <= 125");
  IQLB_REQUIRE(1 + L.A <= 16 * ((L.Dq + 63) / 64), "action_dim too wide for the head staging area");
  Bf16State* st = new (std::nothrow) Bf16State();
  if (!st) return fail(IQLB_ERR_NOMEM, "out of host memory%s", "");
  h->bf16_state = st;
  BfParams& P = st->P;
  P.L = L;
  P.agent0 = 0;
  P.KXB = (L.Dq + 63) / 64;
  P.w0_bytes = (int64_t)P.KXB * L.H * 128;
  P.net_bytes = P.w0_bytes + (int64_t)(L.H / 64) * L.H * 128;
  P.agent_w_bytes = 6 * P.net_bytes;
  P.x_bytes = (int64_t)P.KXB * L.B * 128;
  P.h_bytes = (int64_t)(L.H / 64) * L.B * 128;
  P.z_bytes = (int64_t)L.B * 128;
  const int64_t N = h->desc.n_agents;
  auto alloc = [&](void** p, int64_t bytes) -> int {
    IQLB_CUDA(cudaMalloc(p, (size_t)bytes));
    IQLB_CUDA(cudaMemset(*p, 0, (size_t)bytes));
    return IQLB_OK;
  };
  int rc = IQLB_OK;
  if (!rc) rc = alloc((void**)&P.wimg, N * P.agent_w_bytes);
  if (!rc) rc = alloc((void**)&P.xsa, N * P.x_bytes);
  if (!rc) rc = alloc((void**)&P.xnext, N * P.x_bytes);
  if (!rc) rc = alloc((void**)&P.h1, N * kTrain * P.h_bytes);
  if (!rc) rc = alloc((void**)&P.h2, N * kTrain * P.h_bytes);
  if (!rc) rc = alloc((void**)&P.dh2, N * kTrain * P.h_bytes);
  if (!rc) rc = alloc((void**)&P.dh1, N * kTrain * P.h_bytes);
  if (!rc) rc = alloc((void**)&P.dz, N * kTrain * P.z_bytes);
  if (!rc) rc = alloc((void**)&P.mask1, N * kTrain * L.B * (L.H / 32) * 4);
  if (!rc) rc = alloc((void**)&P.mask2, N * kTrain * L.B * (L.H / 32) * 4);
  if (!rc) rc = alloc((void**)&P.gradh, N * L.Pstride * 2);
  P.dbg = nullptr;
  P.rev = 0;
  auto envi = [](const char* k, int d) { const char* v = getenv(k); return v ? atoi(v) : d; };
  P.opt_issuers = envi("IQLB_ISSUERS", 2);
  P.opt_lpt = envi("IQLB_LPT", 1);
  P.dbg_z = envi("IQLB_DBG_Z", 0);  // which agent's CTA records the clock64 timeline
  if (!rc && getenv("IQLB_TIMELINE") && atoi(getenv("IQLB_TIMELINE"))) rc = alloc((void**)&P.dbg, 8192 * 8);
  if (rc) return rc;
  P.drop_thresh = 0; P.drop_scale = 1.0f; P.drop_rng = h->rng; P.drop_it = h->it;
  IQLB_CUDA(cudaFuncSetAttribute(bf16_bwd_data_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bwd_data_smem()));
  IQLB_CUDA(cudaFuncSetAttribute(bf16_bwd_weight_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)bwd_weight_smem(L.B)));
  st->dirty = true;
  return IQLB_OK;
}

void bf16_destroy(iqlb_agents* h) {
  Bf16State* st = (Bf16State*)h->bf16_state;
  if (!st) return;
  BfParams& P = st->P;
  cudaFree(P.wimg); cudaFree(P.xsa); cudaFree(P.xnext); cudaFree(P.h1); cudaFree(P.h2); cudaFree(P.dh2);
  cudaFree(P.dh1); cudaFree(P.dz); cudaFree(P.mask1); cudaFree(P.mask2); cudaFree(P.gradh); cudaFree(P.dbg);
  delete st;
  h->bf16_state = nullptr;
}

const BfParams* bf16_params(const iqlb_agents* h) {
  const Bf16State* st = (const Bf16State*)h->bf16_state;
  return st ? &st->P : nullptr;
}

void bf16_mark_dirty(iqlb_agents* h) {
  if (h->bf16_state) ((Bf16State*)h->bf16_state)->dirty = true;
}

int bf16_refresh(iqlb_agents* h, cudaStream_t stream) {
  Bf16State* st = (Bf16State*)h->bf16_state;
  IQLB_REQUIRE(st != nullptr, "not a bf16 handle");
  const Layout& L = h->L;
  shadow_refresh_kernel<<<dim3((unsigned)((L.off_logstd + 255) / 256), h->desc.n_agents), 256, 0, stream>>>(
      st->P, h->params, h->target);
  IQLB_LAUNCH_CHECK();
  st->dirty = false;
  return IQLB_OK;
}

GatherArgs bf16_gather_args(const iqlb_agents* h) {
  const Layout& L = h->L;
  GatherArgs G{};
  G.table = (const float4*)h->table; G.n_rows = (uint64_t)h->n_rows; G.vec_per_row = L.R / 4;
  G.rows_per_cta = 256 / G.vec_per_row; G.reward_col = 2 * L.S + L.A; G.reward_cols = h->reward_cols;
  G.reward_variant = h->reward_variant; G.rng = (const uint64_t*)h->rng; G.idx_out = h->idx_out; G.batch_out = (float4*)h->batch;
  G.blocks_x = (L.B + G.rows_per_cta - 1) / G.rows_per_cta;
  return G;
}

// sampled steps: Philox gather + image packing in one launch (replaces replay_gather + pack_x)
int bf16_gather_pack(iqlb_agents* h, int a0, int n, Prof& pf, cudaStream_t stream) {
  Bf16State* st = (Bf16State*)h->bf16_state;
  IQLB_REQUIRE(st != nullptr, "not a bf16 handle");
  const Layout& L = h->L;
  IQLB_REQUIRE(h->table != nullptr, "no replay table attached (iqlb_agents_set_replay)");
  BfParams P = st->P;
  P.agent0 = a0;
  const int vpr = L.R / 4;
  IQLB_REQUIRE(vpr <= 256, "row_stride too wide");
  const int rpc = 256 / vpr;
  const GatherArgs G = bf16_gather_args(h);
  IQLB_CUDA(launch_pdl(replay_gather_pack_kernel, dim3((unsigned)G.blocks_x, n), dim3(256), (size_t)rpc * L.R * 4, stream,
                       pdl_enabled(1) && !h->prof, P, G));
  pf.mark("replay_gather_pack");
  return IQLB_OK;
}

// actor dropout parameters of this launch (iqlb_agents_set_actor_dropout)
static void bf16_set_dropout(const iqlb_agents* h, BfParams& P) {
  const bool on = h->drop_p > 0.f;
  P.drop_thresh = on ? (uint32_t)((double)h->drop_p * 4294967296.0) : 0u;
  P.drop_scale = on ? 1.0f / (1.0f - h->drop_p) : 1.0f;
  P.drop_rng = h->rng;
  P.drop_it = h->it;
}

int bf16_step_gemms_fwd(iqlb_agents* h, int a0, int n, Prof& pf, cudaStream_t stream, bool packed) {
  Bf16State* st = (Bf16State*)h->bf16_state;
  IQLB_REQUIRE(st != nullptr, "not a bf16 handle");
  const Layout& L = h->L;
  const int N = n;
  BfParams P = st->P;
  P.agent0 = a0;
  P.rev = h->zig == 1;
  int rc;
  if (st->dirty) {
    if ((rc = bf16_refresh(h, stream))) return rc;
    pf.mark("shadow_refresh");
  }
  if (!packed) {
    IQLB_CUDA(launch_pdl(pack_x_kernel, dim3((unsigned)((L.B * P.KXB * 8 + 255) / 256), N), dim3(256), 0, stream,
                         pdl_enabled(1) && !h->prof, P, h->batch_src ? h->batch_src : (const float*)h->batch));
    pf.mark("pack_x");
  }
  bf16_set_dropout(h, P);
  if ((rc = bf16_fwd_persist_launch(P, h->params, h->target, h->head_out, N, stream, pdl_enabled(2) && !h->prof))) return rc;
  pf.mark("bf16_fwd");
  return IQLB_OK;
}

int bf16_step_gemms_bwd(iqlb_agents* h, int a0, int n, Prof& pf, cudaStream_t stream) {
  Bf16State* st = (Bf16State*)h->bf16_state;
  IQLB_REQUIRE(st != nullptr, "not a bf16 handle");
  const Layout& L = h->L;
  const int N = n;
  BfParams P = st->P;
  P.agent0 = a0;
  bf16_set_dropout(h, P);
  P.rev = h->zig == 1;  // data gradients in the forward's direction, weight gradients against it
  // More half-tiles than SMs: the persistent kernel (bf16_bwd_data.cu; shared W1 loads, loads a tile ahead, packed schedule:
  // 40.0 -> 31.7 us at 64 agents).  Otherwise every half-tile gets its own SM either way and the one-tile-per-CTA kernel
  // below is the one that measured faster inside the PDL chain (8 agents: 59.7 vs 61.0 us per step, 16: 81.1 vs 83.2,
  // although the persistent kernel alone is 1.5 us shorter; profiles/r02f_bwd_data_persist_ab.jsonl).
  static const int bd_env = getenv("IQLB_BD_PERSIST") ? atoi(getenv("IQLB_BD_PERSIST")) : -1;
  const int n_sm_bd = h->n_sm > 0 ? h->n_sm : 148;
  const bool bd_persist = bd_env >= 0 ? bd_env != 0 : N * kTrain * (L.B / 128) > n_sm_bd;
  if (bd_persist) {
    int rc = bf16_bwd_data_persist_launch(P, h->params, h->d_out, N, n_sm_bd, stream,
                                          pdl_enabled(8) && !h->prof);
    if (rc) return rc;
  } else {
    IQLB_CUDA(launch_pdl(bf16_bwd_data_kernel, dim3(L.B / 128, kTrain, N), dim3(kBdThreads), bwd_data_smem(), stream,
                         pdl_enabled(8) && !h->prof, P,
                         (const float*)h->params, (const float*)h->d_out));
  }
  pf.mark("bf16_bwd_data");
  if (h->zig >= 0) P.rev ^= 1;
  const int n_tiles = (L.H / 128) * kTrain * N;
  const int n_sm = h->n_sm > 0 ? h->n_sm : 148;
  IQLB_CUDA(launch_pdl(bf16_bwd_weight_kernel, dim3((unsigned)(n_tiles < n_sm ? n_tiles : n_sm)), dim3(kBwThreads),
                       bwd_weight_smem(L.B), stream, pdl_enabled(16) && !h->prof, P, n_tiles));
  pf.mark("bf16_bwd_weight");
  return IQLB_OK;
}

}  // namespace iqlb
