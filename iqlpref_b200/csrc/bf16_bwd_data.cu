// Persistent data-gradient kernel of the bf16 IQL step (replaces round 1's one-tile-per-CTA bf16_bwd_data_kernel).
//   dH2[b, j] = relu'(H2[b, j]) * sum_o dz[b, o] W2[o, j]      CUDA cores, fp32 W2 (the reference's head backward)
//   dH1[b, i] = relu'(H1[b, i]) * sum_j dH2[b, j] W1[j, i]     tcgen05, W1 image read MN-major (N = in, K = out)
// Both tiles go to HBM as bf16 images for the weight-gradient kernel (ImplicitQLearning.train's loss.backward(),
// /root/reference/algorithms/offline/iql.py:581-637).
//
// One CTA per SM walks a host-built list of half-tiles (bd_schedule below); the two 128-row half batches of a UNIT =
// (agent, trainable net) share one W1 image and are adjacent in a CTA's list whenever the schedule can keep them together:
//   * W1 (128 KB) is loaded once per unit instead of once per half batch, and the next unit's W1 is requested as soon as
//     the last MMA of the current one has completed - under the drain and the next unit's dH2 production - instead of at
//     the head of a new CTA (where it cost ~3 us of a ~9 us CTA, plus the 1.2 us exit / launch gap);
//   * what a half-tile reads from global memory (ReLU masks, dz, W2) is loaded into registers one half-tile / one unit
//     ahead: measured on the first version of this kernel, those loads were 3.3 k exposed clocks per half-tile and the W2
//     load 3.6 k per unit, of a 10 k half-tile;
//   * the dH1 tile is staged in the A-tile region (free once the MMAs and the dH2 store have read it) and leaves as bulk
//     stores; the next dH2 K-block is written as soon as the store of the dH1 block in its place has been read.  (Stores
//     straight from registers - 32 lanes x 16 bytes to 32 different lines per instruction - cost 5.6 k clocks per half-tile.)
//   * half-tiles are dealt by cost (the actor's dH2 has out_dim = 8-24 columns of dz per row against one for V / Q),
//     heaviest first, each to the least-loaded CTA.
//   warps 0-15  workers: 4 per TMEM lane quarter; drain of the previous half, then dH2 K-block by K-block
//   warp 16     one elected thread: W1 loads, dH2 / dH1 bulk stores and every MMA issue
#include <vector>

#include "iql_step.cuh"
#include "umma.cuh"

namespace iqlb {
using namespace umma;

namespace bdp {
constexpr int kWorkerWarps = 16, kWorkerThreads = kWorkerWarps * 32, kThreads = kWorkerThreads + 32;
// shared memory: A tile (dH2 K-blocks, then the staged dH1 blocks; 4 x 16 KB) | W1 image (4 x 32 KB column blocks)
//                | W2 fp32 [out][256] | control
constexpr uint32_t oA = 0, oW = 65536u, oW2 = oW + 131072u, kW2Bytes = kOutMax * 1024u, oCtrl = oW2 + kW2Bytes,
                   kSmem = oCtrl + 256u;
static_assert(kSmem <= 232448u, "shared memory budget of one CTA (227 KB)");
constexpr int kPre = 8;  // dz columns / W2 values per thread that are prefetched in registers (wider actors load the rest late)
// barriers (8 bytes each)
enum { BAR_W = 0, BAR_K = 1 /* ..4: dH2 K-block written */, BAR_MMA = 5, BAR_A_FREE = 6, BAR_DRAINED = 7 /* ..10: dH1 block staged */,
       BAR_OUT = 11 /* ..14: the store of dH1 block ib has been read */, N_BARS = 15 };
}  // namespace bdp

__device__ __forceinline__ uint32_t bd_chunk_off(int row, int col0) {
  return (uint32_t)((col0 >> 6) * 16384 + row * 128 + ((((col0 & 63) >> 3) ^ (row & 7)) << 4));
}

struct BdUnit {
  int agent, t, s, in, nout;
  int64_t at;
};

// work-list entry: (agent index within the launch) << 3 | trainable net << 1 | half batch
__device__ __forceinline__ BdUnit bd_unit(const BfParams& P, int e, int n_ag) {
  const int r = e >> 3, t = (e >> 1) & 3;
  BdUnit U;
  U.t = t;
  U.agent = P.agent0 + (P.rev ? n_ag - 1 - r : r);
  U.s = train_stream(t);
  U.in = P.L.in_dim(U.s);
  U.nout = P.L.out_dim(U.s);
  U.at = (int64_t)U.agent * kTrain + t;
  return U;
}

__global__ void __launch_bounds__(bdp::kThreads, 1) bf16_bwd_data_persist_kernel(BfParams P, const float* __restrict__ params,
                                                                                 const float* __restrict__ d_out, int n_ag,
                                                                                 const int32_t* __restrict__ wl_off,
                                                                                 const int32_t* __restrict__ wl) {
  using namespace bdp;
  extern __shared__ __align__(1024) uint8_t smem[];
  const Layout& L = P.L;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, wu = warp_uniform();  // wu: role branches only (umma.cuh)
  const int quarter = warp & 3, cgrp = (warp >> 2) & 3;
  const uint32_t sbase = smem_u32(smem);
  auto bar = [&](int i) -> uint32_t { return sbase + oCtrl + 8u * (uint32_t)i; };
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + oCtrl + 8 * N_BARS);
  // this CTA's half-tiles (host-built schedule, bf16_bwd_data_persist_launch): entries of one unit are adjacent
  const int wl_beg = __ldg(wl_off + blockIdx.x), cnt = __ldg(wl_off + blockIdx.x + 1) - wl_beg;
  auto entry = [&](int k) -> int { return __ldg(wl + wl_beg + k); };
  auto next_run = [&](int k) -> int {  // first entry after k that belongs to another unit, or -1
    const int e = entry(k);
    for (int k2 = k + 1; k2 < cnt; ++k2)
      if ((entry(k2) >> 1) != (e >> 1)) return entry(k2);
    return -1;
  };

  if (tid == 0) {
    if (sbase & 1023u) __trap();
    mbar_init(bar(BAR_W), 1); mbar_init(bar(BAR_MMA), 1); mbar_init(bar(BAR_A_FREE), 1);
    for (int i = 0; i < 4; ++i) {
      mbar_init(bar(BAR_K + i), kWorkerWarps); mbar_init(bar(BAR_DRAINED + i), kWorkerWarps); mbar_init(bar(BAR_OUT + i), 1);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(sbase + oCtrl + 8 * N_BARS, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
#ifdef IQLB_NO_TIMELINE
#define IQLB_TSB(slot) do { } while (0)
#else  // clock64 stamps of CTA 0 (worker thread 0 and the issuing thread), 16 slots per half-tile: scripts/timeline.py
#define IQLB_TSB(slot) do { if (P.dbg != nullptr && blockIdx.x == 0 && (slot) < 512) P.dbg[2560 + (slot)] = (unsigned long long)clock64(); } while (0)
#endif

  auto w1_of = [&](const BdUnit& U) -> const uint8_t* {
    return P.wimg + (int64_t)U.agent * P.agent_w_bytes + (int64_t)net_of_stream(U.s) * P.net_bytes + P.w0_bytes;
  };

  if (wu == kWorkerWarps) {
    // ------------------------------------------------------------------ issuing thread
    if (elect_one()) {
      // the first W1 image is Adam's output of the previous step: older than the PDL predecessor (the loss kernel)
      auto load_w1 = [&](int e) {
        const BdUnit U = bd_unit(P, e, n_ag);
        mbar_expect_tx(bar(BAR_W), 131072u);
        for (int ib = 0; ib < 4; ++ib)
          bulk_g2s(sbase + oW + (uint32_t)ib * 32768u, w1_of(U) + (int64_t)ib * 32768, 32768u, bar(BAR_W));
      };
      if (cnt > 0) load_w1(entry(0));
      pdl_wait();
      pdl_launch();
      auto store_dh1 = [&](uint8_t* dh1, uint32_t dph) {  // the staged dH1 blocks of the half-tile drained last
        for (int ib = 0; ib < 4; ++ib) {
          mbar_wait(bar(BAR_DRAINED + ib), dph);
          bulk_s2g(dh1 + (int64_t)ib * L.B * 128, sbase + oA + (uint32_t)ib * 16384u, 16384u);
          bulk_commit();
        }
        asm volatile("cp.async.bulk.wait_group.read 3;" ::: "memory"); mbar_arrive(bar(BAR_OUT + 0));
        asm volatile("cp.async.bulk.wait_group.read 2;" ::: "memory"); mbar_arrive(bar(BAR_OUT + 1));
        asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); mbar_arrive(bar(BAR_OUT + 2));
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); mbar_arrive(bar(BAR_OUT + 3));
      };
      int irun = -1;  // units (runs of entries that share a W1 image) started so far - 1
      uint8_t* prev_dh1 = nullptr;
      for (int n = 0; n < cnt; ++n) {
        {
          const int e = entry(n), h = e & 1;
          const BdUnit U = bd_unit(P, e, n_ag);
          const bool new_run = n == 0 || (entry(n - 1) >> 1) != (e >> 1);
          const bool last_of_run = n + 1 == cnt || (entry(n + 1) >> 1) != (e >> 1);
          const uint32_t nph = (uint32_t)(n & 1);
          if (n > 0) store_dh1(prev_dh1, nph ^ 1u);
          if (new_run) {
            ++irun;
            const int en = next_run(n);  // the next unit's W1 starts its trip HBM -> L2 one unit ahead
            if (en >= 0) bulk_prefetch_l2(w1_of(bd_unit(P, en, n_ag)), 131072u);
            mbar_wait(bar(BAR_W), (uint32_t)(irun & 1));
          }
          IQLB_TSB(n * 16 + 8);
          for (int kb = 0; kb < 4; ++kb) {
            mbar_wait(bar(BAR_K + kb), nph);
            tc_fence_after();
            gemm_issue(tmem, operand_k_major(sbase + oA + (uint32_t)kb * 16384u, 128, 0),
                       operand_mn_major(sbase + oW, L.H, 0, 64 * kb), 256, 4, kb > 0);
          }
          mma_commit(bar(BAR_MMA));
          IQLB_TSB(n * 16 + 9);
          // the dH2 tile also goes to HBM for the weight-gradient kernel (one group behind the MMAs: stored K-block by
          // K-block as each is produced, the reads competed with the MMAs' operand fetches and the kernel was 1 us slower)
          uint8_t* dh2 = P.dh2 + U.at * P.h_bytes + (int64_t)h * 16384;
          for (int cb = 0; cb < 4; ++cb) bulk_s2g(dh2 + (int64_t)cb * L.B * 128, sbase + oA + (uint32_t)cb * 16384u, 16384u);
          bulk_commit();
          mbar_wait(bar(BAR_MMA), nph);
          IQLB_TSB(n * 16 + 10);
          // W1 consumed by every half of this unit: the next unit's image, under this unit's last drain
          if (last_of_run && n + 1 < cnt) load_w1(entry(n + 1));
          bulk_wait_read0();
          IQLB_TSB(n * 16 + 11);
          mbar_arrive(bar(BAR_A_FREE));  // accumulator complete, A tile read by the MMAs and by the store: drain, then produce
          prev_dh1 = P.dh1 + U.at * P.h_bytes + (int64_t)h * 16384;
        }
      }
      if (cnt > 0) store_dh1(prev_dh1, (uint32_t)((cnt - 1) & 1));
      // (every store has been READ out of shared memory: the CTA may exit, the writes complete with the grid - waiting for
      //  them here, cp.async.bulk.wait_group 0, kept every CTA ~1 us longer and cost 1.2 us per step at 8 agents)
    }
  } else {
    // ------------------------------------------------------------------ workers
    const int row = quarter * 32 + lane;
    const uint32_t lane_base = tmem + ((uint32_t)(quarter * 32) << 16);
    float* sw2 = reinterpret_cast<float*>(smem + oW2);
    // registers loaded one half-tile ahead: mask words of this thread's 16 columns per block, and the first kPre dz columns
    uint32_t nm2[4], nm1[4];
    float ndz[kPre], nw2[kPre];
    auto issue_half = [&](const BdUnit& U, int h) {
      const int brow = h * 128 + row;
      const uint32_t* m2p = P.mask2 + (U.at * L.B + brow) * (L.H / 32) + (cgrp >> 1);
      const uint32_t* m1p = P.mask1 + (U.at * L.B + brow) * (L.H / 32) + (cgrp >> 1);
#pragma unroll
      for (int kb = 0; kb < 4; ++kb) { nm2[kb] = m2p[kb * 2]; nm1[kb] = m1p[kb * 2]; }
    };
    auto issue_dz = [&](const BdUnit& U, int h) {  // (dz is the loss kernel's output: after the PDL wait)
      const float* dout = d_out + (int64_t)U.agent * L.dout_rows() * L.B + (int64_t)(U.t == 3 ? 3 : U.t) * L.B + h * 128 + row;
#pragma unroll
      for (int o = 0; o < kPre; ++o) ndz[o] = o < U.nout ? dout[(int64_t)o * L.B] : 0.f;
    };
    auto issue_w2 = [&](const BdUnit& U) {  // this thread's share of the unit's W2 (out_dim <= kPre: 256 * out / 512 values)
      const float* w2 = params + (int64_t)U.agent * L.Pstride + L.net_off(U.s) + L.w2(U.in);
#pragma unroll
      for (int k = 0; k < kPre / 2; ++k) {
        const int i = tid + k * kWorkerThreads;
        nw2[k] = (U.nout <= kPre && i < U.nout * 256) ? __ldg(w2 + i) : 0.f;
      }
    };
    if (cnt > 0) {
      const BdUnit U0 = bd_unit(P, entry(0), n_ag);
      issue_half(U0, entry(0) & 1);
      issue_w2(U0);
    }
    pdl_wait();
    pdl_launch();
    if (cnt > 0) issue_dz(bd_unit(P, entry(0), n_ag), entry(0) & 1);
    // the half-tile whose accumulator is still to be drained
    bool pend = false;
    uint32_t pm1[4] = {0u, 0u, 0u, 0u};
    float pend_gs = 1.0f;
    int n = 0;  // half-tiles produced so far by this CTA
    auto drain = [&]() {  // accumulator of half-tile n - 1 -> masked bf16 dH1 blocks staged in the A tile
      mbar_wait(bar(BAR_A_FREE), (uint32_t)((n - 1) & 1));
      tc_fence_after();
#pragma unroll
      for (int ib = 0; ib < 4; ++ib) {
        float v[16];
        tmem_ld16(lane_base + 64u * ib + 16u * cgrp, v);
        tmem_ld_wait();
        const uint32_t m = pm1[ib];
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = ((m >> j) & 1u) ? v[j] : 0.f;
        if (pend_gs != 1.0f) {
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] *= pend_gs;
        }
#pragma unroll
        for (int q = 0; q < 2; ++q)
          *reinterpret_cast<uint4*>(smem + oA + ib * 16384 + bd_chunk_off(row, 16 * cgrp + 8 * q)) =
              make_uint4(pack_bf16(v[8 * q], v[8 * q + 1]), pack_bf16(v[8 * q + 2], v[8 * q + 3]),
                         pack_bf16(v[8 * q + 4], v[8 * q + 5]), pack_bf16(v[8 * q + 6], v[8 * q + 7]));
        fence_async_smem();
        if (ib == 3) tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(BAR_DRAINED + ib));
      }
    };
    for (; n < cnt; ++n) {
      const int e = entry(n), h = e & 1;
      const BdUnit U = bd_unit(P, e, n_ag);
      const float gscale = (U.t == 3 && P.drop_thresh != 0) ? P.drop_scale : 1.0f;
      const float* dout = d_out + (int64_t)U.agent * L.dout_rows() * L.B + (int64_t)(U.t == 3 ? 3 : U.t) * L.B;
      if (n == 0 || (entry(n - 1) >> 1) != (e >> 1)) {
      if (tid == 0) IQLB_TSB(n * 16 + 0);
      // W2 of this unit (every worker has finished the previous unit's production, which read the buffer)
      asm volatile("bar.sync 1, %0;" ::"n"(kWorkerThreads) : "memory");
      if (U.nout <= kPre) {
#pragma unroll
        for (int k = 0; k < kPre / 2; ++k) {
          const int i = tid + k * kWorkerThreads;
          if (i < U.nout * 256) sw2[i] = nw2[k];
        }
      } else {
        const float* w2 = params + (int64_t)U.agent * L.Pstride + L.net_off(U.s) + L.w2(U.in);
        for (int i = tid; i < U.nout * 256; i += kWorkerThreads) sw2[i] = __ldg(w2 + i);
      }
      asm volatile("bar.sync 1, %0;" ::"n"(kWorkerThreads) : "memory");
      const int en = next_run(n);
      if (en >= 0) issue_w2(bd_unit(P, en, n_ag));  // the next unit's, a whole unit ahead
      }
      {
        if (tid == 0) IQLB_TSB(n * 16 + 1);
        const int brow = h * 128 + row;
        // this half's registers, and the loads of the next half-tile
        uint32_t m2w[4], m1w[4];
        float dz[kOutMax];
#pragma unroll
        for (int kb = 0; kb < 4; ++kb) {
          m2w[kb] = (nm2[kb] >> ((cgrp & 1) * 16)) & 0xffffu;
          m1w[kb] = (nm1[kb] >> ((cgrp & 1) * 16)) & 0xffffu;
        }
#pragma unroll
        for (int o = 0; o < kOutMax; ++o) dz[o] = o < kPre ? ndz[o] : 0.f;
        if (U.nout > kPre) {
#pragma unroll
          for (int o = kPre; o < kOutMax; ++o) dz[o] = o < U.nout ? dout[(int64_t)o * L.B + brow] : 0.f;
        }
        if (n + 1 < cnt) { const int e1 = entry(n + 1); const BdUnit Nx = bd_unit(P, e1, n_ag); issue_half(Nx, e1 & 1); issue_dz(Nx, e1 & 1); }
        if (tid == 0) IQLB_TSB(n * 16 + 2);
        const bool had = pend;
        if (pend) drain();
        if (tid == 0) IQLB_TSB(n * 16 + 3);
        const uint32_t oph = (uint32_t)((n - 1) & 1);
        if (U.nout == 1 && gscale == 1.0f) {  // V / Q: one dz column, the whole tile unrolled
#pragma unroll
          for (int kb = 0; kb < 4; ++kb) {
            if (had) mbar_wait(bar(BAR_OUT + kb), oph);  // the dH1 block staged here has left
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
              const int c0 = kb * 64 + cgrp * 16 + hh * 8;
              const float4 w0 = *reinterpret_cast<const float4*>(sw2 + c0), w1 = *reinterpret_cast<const float4*>(sw2 + c0 + 4);
              const uint32_t m = m2w[kb] >> (8 * hh);
              float g[8] = {dz[0] * w0.x, dz[0] * w0.y, dz[0] * w0.z, dz[0] * w0.w, dz[0] * w1.x, dz[0] * w1.y, dz[0] * w1.z, dz[0] * w1.w};
#pragma unroll
              for (int j = 0; j < 8; ++j) g[j] = ((m >> j) & 1u) ? g[j] : 0.f;
              *reinterpret_cast<uint4*>(smem + oA + bd_chunk_off(row, c0)) =
                  make_uint4(pack_bf16(g[0], g[1]), pack_bf16(g[2], g[3]), pack_bf16(g[4], g[5]), pack_bf16(g[6], g[7]));
            }
            fence_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(BAR_K + kb));  // K-block kb of this warp's rows is in shared memory
          }
        } else {
          // (not unrolled: copies of the 24-output body would not fit the instruction cache)
#pragma unroll 1
          for (int kb = 0; kb < 4; ++kb) {
            if (had) mbar_wait(bar(BAR_OUT + kb), oph);
            const uint32_t mh = m2w[kb];
#pragma unroll 1
            for (int hh = 0; hh < 2; ++hh) {
              const int c0 = kb * 64 + cgrp * 16 + hh * 8;
              float g[8];
#pragma unroll
              for (int j = 0; j < 8; ++j) g[j] = 0.f;
              // groups of eight outputs behind a uniform branch: an 8-action actor issues one group, not three
#pragma unroll
              for (int ob = 0; ob < kOutMax; ob += 8) {
                if (ob < U.nout) {
#pragma unroll
                  for (int oo = 0; oo < 8; ++oo) {
                    const int o = ob + oo;
                    if (oo == 0 || o < U.nout) {
                      const float4 w0 = *reinterpret_cast<const float4*>(sw2 + o * 256 + c0),
                                   w1 = *reinterpret_cast<const float4*>(sw2 + o * 256 + c0 + 4);
                      g[0] = fmaf(dz[o], w0.x, g[0]); g[1] = fmaf(dz[o], w0.y, g[1]); g[2] = fmaf(dz[o], w0.z, g[2]);
                      g[3] = fmaf(dz[o], w0.w, g[3]); g[4] = fmaf(dz[o], w1.x, g[4]); g[5] = fmaf(dz[o], w1.y, g[5]);
                      g[6] = fmaf(dz[o], w1.z, g[6]); g[7] = fmaf(dz[o], w1.w, g[7]);
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
              *reinterpret_cast<uint4*>(smem + oA + bd_chunk_off(row, c0)) =
                  make_uint4(pack_bf16(g[0], g[1]), pack_bf16(g[2], g[3]), pack_bf16(g[4], g[5]), pack_bf16(g[6], g[7]));
            }
            fence_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(BAR_K + kb));
          }
        }
        if (tid == 0) IQLB_TSB(n * 16 + 4);
        pend = true;
        pend_gs = gscale;
#pragma unroll
        for (int ib = 0; ib < 4; ++ib) pm1[ib] = m1w[ib];
      }
    }
    if (pend) drain();
    if (tid == 0) IQLB_TSB(n * 16 + 5);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 256);
}

// ------------------------------------------------------------------------------------------------
// Host-built schedule.  Half-tiles cost ~8 k clocks for V / Q and ~12-20 k for the actor (its dH2 has out_dim columns of dz
// per row); 512 of them on 148 SMs do not divide evenly, so instead of dealing units round-robin the host packs them:
// heaviest units first, each half to the least-loaded CTA - except that the second half of a unit joins the first (one W1
// load for both) as long as that CTA stays within 6 % of the mean load.  With 64 antmaze agents: 48 CTAs get (A, A, L),
// 16 (A, A), 84 four light halves; the makespan is 31.8 against 40 cost units for whole units dealt round-robin.
// ------------------------------------------------------------------------------------------------
struct BdSchedule {
  int dev = -1, n_agents = 0, n_sm = 0, A = 0, grid = 0;
  int32_t *d_off = nullptr, *d_list = nullptr;
};

static int bd_schedule(int n_agents, int n_sm, int A, const BdSchedule** out) {
  static std::vector<BdSchedule> cache;  // (a handful of shapes per process; entries live until exit)
  int dev = 0;
  IQLB_CUDA(cudaGetDevice(&dev));
  for (const BdSchedule& c : cache)
    if (c.dev == dev && c.n_agents == n_agents && c.n_sm == n_sm && c.A == A) { *out = &c; return IQLB_OK; }
  const int n_halves = n_agents * kTrain * 2;
  const int G = n_halves < n_sm ? n_halves : n_sm;
  const double cost_light = 8.0, cost_actor = 8.0 + 4.0 * ((A + 7) / 8);
  const double mean = (n_agents * 2 * (3 * cost_light + cost_actor)) / G;
  std::vector<double> load(G, 0.0);
  std::vector<std::vector<int32_t>> lists(G);
  auto least = [&]() { int b = 0; for (int i = 1; i < G; ++i) if (load[i] < load[b]) b = i; return b; };
  for (int pass = 0; pass < 2; ++pass) {  // actor units, then V / Q1 / Q2
    for (int r = 0; r < n_agents; ++r) {
      for (int t = (pass == 0 ? 3 : 0); t < (pass == 0 ? 4 : 3); ++t) {
        const double c = t == 3 ? cost_actor : cost_light;
        const int c0 = least();
        lists[c0].push_back((int32_t)(r << 3 | t << 1 | 0));
        load[c0] += c;
        const int c1 = (load[c0] + c <= 1.06 * mean) ? c0 : least();
        lists[c1].push_back((int32_t)(r << 3 | t << 1 | 1));
        load[c1] += c;
      }
    }
  }
  std::vector<int32_t> off(G + 1, 0), flat;
  for (int i = 0; i < G; ++i) {
    off[i + 1] = off[i] + (int32_t)lists[i].size();
    flat.insert(flat.end(), lists[i].begin(), lists[i].end());
  }
  BdSchedule sc;
  sc.dev = dev; sc.n_agents = n_agents; sc.n_sm = n_sm; sc.A = A; sc.grid = G;
  IQLB_CUDA(cudaMalloc((void**)&sc.d_off, off.size() * sizeof(int32_t)));
  IQLB_CUDA(cudaMalloc((void**)&sc.d_list, flat.size() * sizeof(int32_t)));
  IQLB_CUDA(cudaMemcpy(sc.d_off, off.data(), off.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
  IQLB_CUDA(cudaMemcpy(sc.d_list, flat.data(), flat.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
  cache.push_back(sc);
  *out = &cache.back();
  return IQLB_OK;
}

int bf16_bwd_data_persist_launch(const BfParams& P, const float* params, const float* d_out, int n_agents, int n_sm,
                                 cudaStream_t stream, bool pdl) {
  static bool attr_set = false;
  if (!attr_set) {
    IQLB_CUDA(cudaFuncSetAttribute(bf16_bwd_data_persist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bdp::kSmem));
    attr_set = true;
  }
  IQLB_REQUIRE(P.L.B == 256 && P.L.H == 256, "data-gradient kernel: batch 256, hidden 256");
  const BdSchedule* sc = nullptr;
  int rc = bd_schedule(n_agents, n_sm, P.L.A, &sc);
  if (rc) return rc;
  IQLB_CUDA(launch_pdl(bf16_bwd_data_persist_kernel, dim3(sc->grid), dim3(bdp::kThreads), bdp::kSmem, stream, pdl, P, params, d_out,
                       n_agents, (const int32_t*)sc->d_off, (const int32_t*)sc->d_list));
  return IQLB_OK;
}

}  // namespace iqlb
