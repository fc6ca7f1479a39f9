// Persistent, warp-specialised forward kernel of the bf16 IQL step (one CTA per SM, tiles =
// (agent, stream, half batch) taken round-robin): X -> relu(X W0^T + b0) -> relu(H1 W1^T + b1) -> head (W2, b2 in fp32,
// tanh for the actor); keeps H1, H2 (bf16 images) and the ReLU masks of the four trainable streams for the backward pass.
//
//   warp 0  producer : bulk copies (X tile, b0 | W0 and the four W1 K-blocks through a 4-slot ring); b1, W2 by the epilogue
//   warp 1  MMA      : tcgen05.mma issue; layer 1 starts on K-block kb as soon as the epilogue has
//                      written H1 block kb, so the tensor core trails the epilogue by one block
//   warps 2-17       : epilogue (4 warps per TMEM lane quarter: 32 of every 64 columns of the even or the odd blocks each;
//                      16 warps = 4 per scheduler: with 8 the TMEM-load / convert / store chain of a block ran at 36 % issue
//                      utilisation; epilogue 0 3.5 k -> 2.3 k clocks, forward 52.4 -> 50.0 us at 64 agents)
//   warp 18          : stores: H1 / H2 tiles staged in shared memory leave as bulk stores (one commit group per
//                      64-column block, so H2 block b moves in as soon as H1 block b has left), b1 | W2 | b2 loads
//
// While the epilogue of tile i drains layer 1, the producer already holds tile i+1's operands and the
// MMA warp has issued its layer 0, so loads, both GEMMs, TMEM allocation and barrier set-up disappear
// from the critical path: a tile costs roughly its two epilogues.
// Tiles come from a host-built work list (fwd_schedule below).  Two instantiations: narrow (the input image is one 64-column block: state_dim + action_dim <= 64, action_dim <= 8, no actor
// dropout - the D4RL locomotion / antmaze shapes) and wide (two input blocks, action_dim <= 24, actor dropout: pen / adroit).
// The wide one trades a ring slot for the second X block and the 24-row W2; W0 takes one ring entry per input block.
#include <vector>

#include "iql_step.cuh"
#include "umma.cuh"

namespace iqlb {
using namespace umma;

namespace fwdp {
// shared memory: H tile | ring of 32 KB operand blocks (the W0 K-blocks, then the four W1 K-blocks of a tile)
//                | X tile | b0 (x2) | b1 + W2 + b2 | barriers
constexpr uint32_t oH = 0, oRing = 65536, kSlotBytes = 32768;
#ifndef IQLB_FWD_EPI_WARPS
#define IQLB_FWD_EPI_WARPS 16
#endif
constexpr int kEpiWarps = IQLB_FWD_EPI_WARPS, kEpiThreads = kEpiWarps * 32, kAuxWarp = 2 + kEpiWarps, kThreads = 64 + kEpiThreads + 32;
constexpr int kGroups = kEpiWarps / 4, kPar = kGroups / 2;  // column-half x block-parity groups of epilogue warps
template <bool kWide>
struct Lay {
  static constexpr int kSlots = kWide ? 3 : 4, kXB = kWide ? 2 : 1, kOut = kWide ? kOutMax : 8;
  static constexpr uint32_t oX = oRing + kSlots * kSlotBytes, oB0 = oX + kXB * 16384, oBW = oB0 + 2048,
                            kBWBytes = kWide ? 1024 + kOutMax * 1024 + 128 : 9216 + 32, oCtrl = oBW + kBWBytes,
                            kSmem = oCtrl + 256;
  static_assert(kSmem <= 232448, "shared memory budget of one CTA (227 KB)");
  static_assert(kSlots <= 4, "barrier indices reserve four ring slots");
  static_assert(kOut <= 32, "partial head sums are parked in 32 accumulator columns");
};
// barrier indices (8 bytes each) inside the control block
enum { FULL_L0 = 0, MMA0 = 1, MMA1 = 2, FULL_R = 3, EMPTY_R = 7, HBLK = 11, ACC0_FREE = 15, ACC1_FREE = 16, FULL_BW = 17,
       EPI0_DONE = 18, EPI1_DONE = 19, H1_OUT = 20 /* ..23: H1 block b has left the H tile */, H2_OUT = 24 /* ..27 */, N_BARS = 28 };
}  // namespace fwdp

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

__device__ __forceinline__ void st_chunk128(uint8_t* tile, int row, int col0, const float* v) {
  const int off = (col0 >> 6) * 16384 + row * 128 + ((((col0 & 63) >> 3) ^ (row & 7)) << 4);
  *reinterpret_cast<uint4*>(tile + off) =
      make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
}

// v = relu(v + bias[0..32)); returns the 32-bit ReLU mask (bit j = v[j] > 0) when kMask
template <bool kMask>
__device__ __forceinline__ uint32_t bias_relu32(float (&v)[32], const float* __restrict__ sb) {
  uint32_t m = 0;
#pragma unroll
  for (int j4 = 0; j4 < 8; ++j4) {
    const float4 bb = *reinterpret_cast<const float4*>(sb + 4 * j4);
    v[4 * j4] = fmaxf(v[4 * j4] + bb.x, 0.f);
    v[4 * j4 + 1] = fmaxf(v[4 * j4 + 1] + bb.y, 0.f);
    v[4 * j4 + 2] = fmaxf(v[4 * j4 + 2] + bb.z, 0.f);
    v[4 * j4 + 3] = fmaxf(v[4 * j4 + 3] + bb.w, 0.f);
  }
  if (kMask) {
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (v[j] > 0.f) m |= 1u << j;
  }
  return m;
}

template <bool kWide>
__global__ void __launch_bounds__(fwdp::kThreads, 1) bf16_fwd_persist_kernel(BfParams P, const float* __restrict__ params,
                                                                             const float* __restrict__ target,
                                                                             float* __restrict__ head_out, int n_agents,
                                                                             const int32_t* __restrict__ wl_off,
                                                                             const int32_t* __restrict__ wl) {
  using namespace fwdp;
  using Ly = Lay<kWide>;
  constexpr int kSlots = Ly::kSlots, kOut = Ly::kOut;
  constexpr uint32_t oX = Ly::oX, oB0 = Ly::oB0, oBW = Ly::oBW, oCtrl = Ly::oCtrl;
  extern __shared__ __align__(1024) uint8_t smem[];
  const Layout& L = P.L;
  // single-thread roles (producer, MMA issuer, store thread) are chosen by elect.sync inside warp-uniform branches, so the
  // operands of their bulk copies / MMAs / commits stay in uniform registers (umma.cuh: elect_one)
  const int tid = threadIdx.x, warp = warp_uniform(), lane = tid & 31, wu = warp;  // (measured: with warp = tid >> 5 for the
  // epilogue arithmetic this kernel is 3 us slower, while round 1's one-tile-per-CTA kernel was 30 % FASTER that way)
  const uint32_t sbase = smem_u32(smem);
  // this CTA's tiles, in the order it runs them: a host-built schedule (persist_launch), entry = agent index << 4 | stream << 1 | half
  const int wl_beg = __ldg(wl_off + blockIdx.x), cnt = __ldg(wl_off + blockIdx.x + 1) - wl_beg;
  auto entry = [&](int it) -> int { return __ldg(wl + wl_beg + it); };
  const uint32_t bars = sbase + oCtrl;
  auto bar = [&](int i) -> uint32_t { return bars + 8u * (uint32_t)i; };
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + oCtrl + 8 * N_BARS);

  if (tid == 0) {
    if (sbase & 1023u) __trap();
    mbar_init(bar(FULL_L0), 1); mbar_init(bar(MMA0), 1); mbar_init(bar(MMA1), 1); mbar_init(bar(FULL_BW), 1);
    for (int i = 0; i < kSlots; ++i) { mbar_init(bar(FULL_R + i), 1); mbar_init(bar(EMPTY_R + i), 1); }
    for (int i = 0; i < 4; ++i) mbar_init(bar(HBLK + i), 8);
    mbar_init(bar(ACC0_FREE), kEpiWarps); mbar_init(bar(ACC1_FREE), kEpiWarps);
    mbar_init(bar(EPI0_DONE), kEpiWarps); mbar_init(bar(EPI1_DONE), 1);
    for (int i = 0; i < 4; ++i) { mbar_init(bar(H1_OUT + i), 1); mbar_init(bar(H2_OUT + i), 1); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(sbase + oCtrl + 8 * N_BARS, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  [[maybe_unused]] const bool dbg = P.dbg != nullptr && blockIdx.x == 0;
  // 64-column blocks of a stream's input image (wide: the Q streams read state | action, the others the state alone)
  auto x_blocks = [&](int in) -> int { return kWide ? (in + 63) >> 6 : 1; };
#ifdef IQLB_NO_TIMELINE
#define IQLB_TS(slot) do { } while (0)
#else
#define IQLB_TS(slot) do { if (dbg) P.dbg[(slot)] = (unsigned long long)clock64(); } while (0)
#endif

  auto tile_ptrs = [&](int tile, int& agent, int& s, int& half, const uint8_t*& wnet, const uint8_t*& ximg,
                       const float*& fnet) {
    half = tile & 1;
    s = (tile >> 1) & 7;
    const int r = tile >> 4;
    agent = P.agent0 + (P.rev ? n_agents - 1 - r : r);
    wnet = P.wimg + (int64_t)agent * P.agent_w_bytes + (int64_t)net_of_stream(s) * P.net_bytes;
    ximg = (s == 0 ? P.xnext : P.xsa) + (int64_t)agent * P.x_bytes;
    fnet = L.uses_target(s) ? target + (int64_t)agent * L.PQstride + L.net_off(s)
                            : params + (int64_t)agent * L.Pstride + L.net_off(s);
  };

  auto load_bw = [&](int tile) {  // b1 and W2 (fp32) of a tile into their single buffer
    int agent, s, half;
    const uint8_t *wnet, *ximg;
    const float* fnet;
    tile_ptrs(tile, agent, s, half, wnet, ximg, fnet);
    const int in = L.in_dim(s), nout = L.out_dim(s);
    // W2 [nout][256] and b2 [nout] are adjacent in the parameter vector; network segments are padded to 32 bytes,
    // so rounding b2 up to 16 bytes stays inside the segment
    const uint32_t w2b2 = (uint32_t)nout * 1024u + (((uint32_t)nout * 4u + 15u) & ~15u);
    mbar_expect_tx(bar(FULL_BW), 1024u + w2b2);
    bulk_g2s(sbase + oBW, fnet + L.b1(in), 1024u, bar(FULL_BW));
    bulk_g2s(sbase + oBW + 1024u, fnet + L.w2(in), w2b2, bar(FULL_BW));
  };
  // PDL: the weight images (Adam of the previous step) are older than the immediate predecessor (the gather):
  // the first tile's weights start towards L2 before this grid waits for the gather's input images
  if (wu == 0 && cnt > 0 && elect_one()) {
    int a0, s0, h0;
    const uint8_t *w0p, *x0p;
    const float* f0p;
    tile_ptrs(entry(0), a0, s0, h0, w0p, x0p, f0p);
    bulk_prefetch_l2(w0p, (uint32_t)P.net_bytes);
  }
  pdl_wait();
  pdl_launch();
#ifndef IQLB_NO_TIMELINE
  if (P.dbg != nullptr && tid == 0 && blockIdx.x < 148) {  // (start, end) of every CTA on the global timer: scripts/timeline.py
    unsigned long long gt;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
    P.dbg[7700 + blockIdx.x * 2] = gt;
  }
#endif
  if (wu == 0) {
    // ------------------------------------------------------------------ producer
    if (elect_one()) {
      int c = 0;  // ring entries issued so far (per tile: the W0 K-blocks, then W1 block 0..3)
      for (int it = 0; it < cnt; ++it) {
        const int tile = entry(it);
        int agent, s, half;
        const uint8_t *wnet, *ximg;
        const float* fnet;
        tile_ptrs(tile, agent, s, half, wnet, ximg, fnet);
        const int in = L.in_dim(s);
        if (it + 1 < cnt) {  // next tile's operands start their trip HBM -> L2 one tile ahead
          int a2, s2, h2;
          const uint8_t *w2, *x2;
          const float* f2;
          tile_ptrs(entry(it + 1), a2, s2, h2, w2, x2, f2);
          bulk_prefetch_l2(w2, (uint32_t)P.net_bytes);
          bulk_prefetch_l2(x2 + (int64_t)h2 * 16384, 16384u);
          if (kWide && L.in_dim(s2) > 64) bulk_prefetch_l2(x2 + (int64_t)L.B * 128 + (int64_t)h2 * 16384, 16384u);
        }
        if (it > 0) mbar_wait(bar(MMA0), (uint32_t)((it - 1) & 1));  // layer 0 of the previous tile has consumed X
        IQLB_TS(1024 + it * 8);
        const int nxb = x_blocks(in);
        mbar_expect_tx(bar(FULL_L0), (uint32_t)nxb * 16384u + 1024u);
        for (int kx = 0; kx < nxb; ++kx)
          bulk_g2s(sbase + oX + (uint32_t)kx * 16384u, ximg + (int64_t)kx * L.B * 128 + (int64_t)half * 16384, 16384u, bar(FULL_L0));
        bulk_g2s(sbase + oB0 + (uint32_t)(it & 1) * 1024u, fnet + L.b0(in), 1024u, bar(FULL_L0));
        for (int e = 0; e < nxb + 4; ++e, ++c) {
          const int slot = c % kSlots, use = c / kSlots;
          if (use > 0) mbar_wait(bar(EMPTY_R + slot), (uint32_t)((use - 1) & 1));
          if (!kWide) IQLB_TS(1024 + it * 8 + 1 + e);
          mbar_expect_tx(bar(FULL_R + slot), kSlotBytes);
          bulk_g2s(sbase + oRing + (uint32_t)slot * kSlotBytes,
                   e < nxb ? wnet + (int64_t)e * 32768 : wnet + P.w0_bytes + (int64_t)(e - nxb) * 32768, kSlotBytes,
                   bar(FULL_R + slot));
        }
      }
    }
  } else if (wu == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (elect_one()) {
      int c = 0;
      for (int it = 0; it < cnt; ++it) {
        const int tile = entry(it);
        const uint32_t ph = (uint32_t)(it & 1);
        int agent, s, half;
        const uint8_t *wnet, *ximg;
        const float* fnet;
        tile_ptrs(tile, agent, s, half, wnet, ximg, fnet);
        const int in = L.in_dim(s);
        const int nxb = x_blocks(in);
        for (int kx = 0; kx < nxb; ++kx, ++c) {
          const int slot = c % kSlots, use = c / kSlots;
          if (kx == 0) mbar_wait(bar(FULL_L0), ph);
          mbar_wait(bar(FULL_R + slot), (uint32_t)(use & 1));
          if (kx == 0 && it > 0) mbar_wait(bar(ACC0_FREE), (uint32_t)((it - 1) & 1));
          tc_fence_after();
          if (kx == 0) IQLB_TS(512 + it * 8);
          const int kin = in - 64 * kx;  // input columns left from this block on
          gemm_issue(tmem, operand_k_major(sbase + oX + (uint32_t)kx * 16384u, 128, 0),
                     operand_k_major(sbase + oRing + (uint32_t)slot * kSlotBytes, L.H, 0), 256, kin >= 64 ? 4 : (kin + 15) / 16, kx > 0);
          mma_commit(bar(EMPTY_R + slot));
        }
        mma_commit(bar(MMA0));
        for (int kb = 0; kb < 4; ++kb, ++c) {
          const int slot = c % kSlots, use = c / kSlots;
          mbar_wait(bar(HBLK + kb), ph);
          mbar_wait(bar(FULL_R + slot), (uint32_t)(use & 1));
          if (kb == 0 && it > 0) mbar_wait(bar(ACC1_FREE), (uint32_t)((it - 1) & 1));
          tc_fence_after();
          IQLB_TS(512 + it * 8 + 1 + kb);
          gemm_issue(tmem + 256, operand_k_major(sbase + oH + (uint32_t)kb * 16384u, 128, 0),
                     operand_k_major(sbase + oRing + (uint32_t)slot * kSlotBytes, L.H, 0), 256, 4, kb > 0);
          mma_commit(bar(EMPTY_R + slot));
        }
        mma_commit(bar(MMA1));
      }
    }
  } else if (wu < kAuxWarp) {
    // ------------------------------------------------------------------ epilogue
    // warp -> TMEM lane quarter (warp % 4, the hardware rule), column half of a 64-column block, and
    // block parity: warps of parity p drain blocks p and p + 2, so blocks 0 and 1 finish together and
    // layer 1 starts after half of epilogue 0
    const int et = tid - 64;
    bool h1_store_pending = false;
    int ntrain = 0;  // trainable tiles finished by this CTA (phase of the H1_OUT / H2_OUT barriers)
    const int quarter = warp & 3, grp = (warp - 2) >> 2, chalf = grp & 1, par = grp >> 1;  // par < kPar
    const int row = quarter * 32 + lane;
    const uint32_t lane_base = tmem + ((uint32_t)(quarter * 32) << 16);
    for (int it = 0; it < cnt; ++it) {
      const int tile = entry(it);
      const uint32_t ph = (uint32_t)(it & 1);
      const float* sb1 = reinterpret_cast<const float*>(smem + oBW);
      const float* sw2 = sb1 + 256;
      int agent, s, half;
      const uint8_t *wnet, *ximg;
      const float* fnet;
      tile_ptrs(tile, agent, s, half, wnet, ximg, fnet);
      const int in = L.in_dim(s), nout = L.out_dim(s), t = train_index(s);
      const int brow = half * 128 + row;
      const float* sb0 = reinterpret_cast<const float*>(smem + oB0 + (it & 1) * 1024);
      uint32_t* m1 = t >= 0 ? P.mask1 + (((int64_t)agent * kTrain + t) * L.B + brow) * (L.H / 32) : nullptr;
      uint32_t* m2 = t >= 0 ? P.mask2 + (((int64_t)agent * kTrain + t) * L.B + brow) * (L.H / 32) : nullptr;
      // actor dropout (wide instantiation only): keep mask applied after the ReLU, stored masks = relu' AND keep, so the
      // backward pass needs only the 1 / (1 - p) factor
      const bool drop = kWide && s == 6 && P.drop_thresh != 0;
      const uint64_t drop_seed = drop ? P.drop_rng[2 * agent] : 0ull, drop_step = drop ? (uint64_t)P.drop_it[agent] : 0ull;
      // (b0 landed before layer 0 was issued; FULL_L0 itself may already be a phase ahead, so it is
      //  not waited on here - MMA0 cannot run ahead of this warp because layer 0 waits for ACC0_FREE)
      if (et == 0) IQLB_TS(it * 8);
      mbar_wait(bar(MMA0), ph);
      tc_fence_after();
      if (et == 0) IQLB_TS(it * 8 + 1);
      // ---- epilogue 0: H1 = relu(acc0 + b0), block by block so that layer 1 can start behind us
#pragma unroll 1
      for (int blk = par; blk < 4; blk += kPar) {
        const int c0 = blk * 64 + chalf * 32;
        float v[32];
        tmem_ld32(lane_base + (uint32_t)c0, v);
        if (h1_store_pending) mbar_wait(bar(H2_OUT + blk), (uint32_t)((ntrain - 1) & 1));  // previous tile's H2 block has left
        tmem_ld_wait();
        if (kWide && drop) {
          bias_relu32<false>(v, sb0 + c0);
          dropout32(v, drop_seed, drop_step, 0, brow, L.H, c0, P.drop_thresh, P.drop_scale);
          uint32_t m = 0;
#pragma unroll
          for (int j = 0; j < 32; ++j) m |= (v[j] > 0.f ? 1u : 0u) << j;
          m1[c0 >> 5] = m;  // (the actor stream is trainable: m1 is not null)
        } else if (t >= 0) m1[c0 >> 5] = bias_relu32<true>(v, sb0 + c0);
        else bias_relu32<false>(v, sb0 + c0);
#pragma unroll
        for (int q = 0; q < 4; ++q) st_chunk128(smem + oH, row, c0 + 8 * q, v + 8 * q);
        fence_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(HBLK + blk));
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(ACC0_FREE));
      if (et == 0) IQLB_TS(it * 8 + 2);
      if (lane == 0) mbar_arrive(bar(EPI0_DONE));  // every H1 block of this warp is written and fenced: the store warp takes over
      mbar_wait(bar(FULL_BW), ph);
      if (et == 0) IQLB_TS(it * 8 + 3);
      mbar_wait(bar(MMA1), ph);
      tc_fence_after();
      if (et == 0) IQLB_TS(it * 8 + 4);
      if (et == 0) IQLB_TS(it * 8 + 7);
      // ---- epilogue 1: H2 = relu(acc1 + b1); head = H2 . W2 + b2
      float acc[kOut];
#pragma unroll
      for (int o = 0; o < kOut; ++o) acc[o] = 0.f;
#pragma unroll 1
      for (int blk = par; blk < 4; blk += kPar) {
        const int c0 = blk * 64 + chalf * 32;
        float v[32];
        tmem_ld32(lane_base + 256u + (uint32_t)c0, v);
        if (t >= 0) mbar_wait(bar(H1_OUT + blk), (uint32_t)(ntrain & 1));  // H1 block blk has left: H2 block blk may move in
        tmem_ld_wait();
        uint32_t m = 0;
        if (kWide && drop) {
          bias_relu32<false>(v, sb1 + c0);
          dropout32(v, drop_seed, drop_step, 1, brow, L.H, c0, P.drop_thresh, P.drop_scale);
#pragma unroll
          for (int j = 0; j < 32; ++j) m |= (v[j] > 0.f ? 1u : 0u) << j;
        } else if (t >= 0) m = bias_relu32<true>(v, sb1 + c0);
        else bias_relu32<false>(v, sb1 + c0);
        if (nout == 1) {
          float a = acc[0];
#pragma unroll
          for (int j4 = 0; j4 < 8; ++j4) {
            const float4 w = *reinterpret_cast<const float4*>(sw2 + c0 + 4 * j4);
            a = fmaf(v[4 * j4], w.x, a); a = fmaf(v[4 * j4 + 1], w.y, a);
            a = fmaf(v[4 * j4 + 2], w.z, a); a = fmaf(v[4 * j4 + 3], w.w, a);
          }
          acc[0] = a;
        } else {
#pragma unroll
          for (int o = 0; o < kOut; ++o) {
            if (o < nout) {
              const float* wr = sw2 + o * 256 + c0;
              float a = acc[o];
#pragma unroll
              for (int j4 = 0; j4 < 8; ++j4) {
                const float4 w = *reinterpret_cast<const float4*>(wr + 4 * j4);
                a = fmaf(v[4 * j4], w.x, a); a = fmaf(v[4 * j4 + 1], w.y, a);
                a = fmaf(v[4 * j4 + 2], w.z, a); a = fmaf(v[4 * j4 + 3], w.w, a);
              }
              acc[o] = a;
            }
          }
        }
        if (t >= 0) {  // H2 is staged in the H tile (layer 1 has consumed H1) and leaves as one bulk store
          m2[c0 >> 5] = m;
#pragma unroll
          for (int q = 0; q < 4; ++q) st_chunk128(smem + oH, row, c0 + 8 * q, v + 8 * q);
        }
      }
      if (t >= 0) fence_async_smem();
      if (et == 0) IQLB_TS(it * 8 + 5);
      // The partial head sums of the column groups meet in TENSOR MEMORY: the warps that share a lane quarter own the
      // same 32 rows, so group g parks its eight sums in the first columns of the accumulator block it drained last and
      // group 0 reads them back after the barrier (no shared-memory staging: with 16 epilogue warps it did not fit).
      constexpr int kLastBlk = (4 / kPar - 1) * kPar;
      if (grp > 0) {
#pragma unroll
        for (int o0 = 0; o0 < kOut; o0 += 16) {
          uint32_t pr[16];
#pragma unroll
          for (int o = 0; o < 16; ++o) pr[o] = o0 + o < kOut ? __float_as_uint(acc[o0 + o]) : 0u;
          tmem_st16(lane_base + 256u + (uint32_t)((par + kLastBlk) * 64 + chalf * 32 + o0), pr);
        }
        tmem_st_wait();
      }
      const float b2v = lane < kOut ? sw2[nout * 256 + (lane < nout ? lane : 0)] : 0.f;  // before the bundle is reloaded
      tc_fence_before();
      named_bar_sync(1, kEpiThreads);
      if (et == 0) IQLB_TS(it * 8 + 6);
      if (et == 0) mbar_arrive(bar(EPI1_DONE));  // H2 staged, b1 / W2 / b2 consumed: the store warp stores and reloads
      h1_store_pending = t >= 0;
      if (t >= 0) ++ntrain;
      if (grp > 0) {
        if (lane == 0) mbar_arrive(bar(ACC1_FREE));  // (layer 1 of the next tile still waits for group 0's arrival)
      } else {
        tc_fence_after();
        float ps[kWide ? 1 : (kGroups > 1 ? kGroups - 1 : 1)][16];
        if (kWide) {  // 24 sums per group: one group and 16 sums at a time, added in the same (group) order as below
#pragma unroll
          for (int g = 1; g < kGroups; ++g) {
#pragma unroll
            for (int o0 = 0; o0 < kOut; o0 += 16) {
              tmem_ld16(lane_base + 256u + (uint32_t)(((g >> 1) + kLastBlk) * 64 + (g & 1) * 32 + o0), ps[0]);
              tmem_ld_wait();
#pragma unroll
              for (int o = 0; o < 16; ++o)
                if (o0 + o < kOut) acc[o0 + o] += ps[0][o];
            }
          }
        } else {
#pragma unroll
          for (int g = 1; g < kGroups; ++g)
            tmem_ld16(lane_base + 256u + (uint32_t)(((g >> 1) + kLastBlk) * 64 + (g & 1) * 32), ps[kWide ? 0 : g - 1]);
          tmem_ld_wait();
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(ACC1_FREE));
        float* ho = head_out + (int64_t)agent * L.head_rows() * L.B;
#pragma unroll
        for (int o = 0; o < kOut; ++o) {
          if (o < nout) {
            float z = acc[o];
            if (!kWide) {
#pragma unroll
              for (int g = 0; g < kGroups - 1; ++g) z += ps[g][o];
            }
            z += __shfl_sync(0xffffffffu, b2v, o);
            if (s == 6) ho[(int64_t)(6 + o) * L.B + brow] = tanhf(z);
            else ho[(int64_t)s * L.B + brow] = z;
          }
        }
        if (et == 0) IQLB_TS(1536 + it * 8 + 2);
        // (the parked partial sums were read before this warp's ACC1_FREE arrival: layer 1 of the next tile cannot
        //  overwrite them earlier)
      }
    }
  } else {
    // ------------------------------------------------------------------ store warp
    if (elect_one()) {
      if (cnt > 0) load_bw(entry(0));
      int ntr = 0;
      for (int it = 0; it < cnt; ++it) {
        const int tile = entry(it);
        const uint32_t ph = (uint32_t)(it & 1);
        int agent, s, half;
        const uint8_t *wnet, *ximg;
        const float* fnet;
        tile_ptrs(tile, agent, s, half, wnet, ximg, fnet);
        const int t = train_index(s);
        mbar_wait(bar(EPI0_DONE), ph);
        if (t >= 0) {  // H1: one commit group per 64-column block, released to epilogue 1 block by block
          uint8_t* h1 = P.h1 + ((int64_t)agent * kTrain + t) * P.h_bytes;
          for (int cb = 0; cb < 4; ++cb) {
            bulk_s2g(h1 + (int64_t)cb * L.B * 128 + (int64_t)half * 16384, sbase + oH + (uint32_t)cb * 16384u, 16384u);
            bulk_commit();
          }
          asm volatile("cp.async.bulk.wait_group.read 3;" ::: "memory"); mbar_arrive(bar(H1_OUT + 0));
          asm volatile("cp.async.bulk.wait_group.read 2;" ::: "memory"); mbar_arrive(bar(H1_OUT + 1));
          asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); mbar_arrive(bar(H1_OUT + 2));
          asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); mbar_arrive(bar(H1_OUT + 3));
        }
        mbar_wait(bar(EPI1_DONE), ph);
        if (it + 1 < cnt) load_bw(entry(it + 1));
        if (t >= 0) {
          uint8_t* h2 = P.h2 + ((int64_t)agent * kTrain + t) * P.h_bytes;
          for (int cb = 0; cb < 4; ++cb) {
            bulk_s2g(h2 + (int64_t)cb * L.B * 128 + (int64_t)half * 16384, sbase + oH + (uint32_t)cb * 16384u, 16384u);
            bulk_commit();
          }
          asm volatile("cp.async.bulk.wait_group.read 3;" ::: "memory"); mbar_arrive(bar(H2_OUT + 0));
          asm volatile("cp.async.bulk.wait_group.read 2;" ::: "memory"); mbar_arrive(bar(H2_OUT + 1));
          asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); mbar_arrive(bar(H2_OUT + 2));
          asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); mbar_arrive(bar(H2_OUT + 3));
          ++ntr;
        }
      }
      bulk_wait_read0();  // (read out of shared memory; the global writes complete with the grid)
    }
  }
  tc_fence_before();
  __syncthreads();
#ifndef IQLB_NO_TIMELINE
  if (P.dbg != nullptr && tid == 0 && blockIdx.x < 148) {
    unsigned long long gt;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
    P.dbg[7700 + blockIdx.x * 2 + 1] = gt;
  }
#endif
  if (warp == 1) tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------------------------------------
// Host-built schedule.  Tiles are not equal: per-tile clock64 stamps (scripts/timeline.py) give ~6 k clocks for the three
// streams that keep nothing (next-state V, the two target Qs), ~9.5 k for V / Q1 / Q2 (H1 / H2 tiles and masks leave for the
// backward pass) and ~11.6 k for the actor (8 heads; more with 24 heads and dropout masks).  896 tiles dealt round-robin give
// every CTA six or seven of mixed cost (CTA ends 25-35 us); here they are sorted by cost and each goes to the least-loaded CTA.
// ------------------------------------------------------------------------------------------------
struct FwdSchedule {
  int dev = -1, n_agents = 0, n_sm = 0, A = 0, drop = 0, grid = 0;
  int32_t *d_off = nullptr, *d_list = nullptr;
};

static int fwd_schedule(int n_agents, int n_sm, int A, bool wide, bool drop, const FwdSchedule** out) {
  static std::vector<FwdSchedule> cache;  // (a handful of shapes per process; entries live until exit)
  int dev = 0;
  IQLB_CUDA(cudaGetDevice(&dev));
  for (const FwdSchedule& c : cache)
    if (c.dev == dev && c.n_agents == n_agents && c.n_sm == n_sm && c.A == A && c.drop == (int)drop) { *out = &c; return IQLB_OK; }
  const int n_tiles = n_agents * kStreams * 2;
  const int G = n_tiles < n_sm ? n_tiles : n_sm;
  const double c_keep = 9.5, c_none = 6.0;
  const double c_actor = wide ? 9.5 + 2.0 * ((A + 7) / 8) + (drop ? 10.0 : 0.0) : 11.6;
  auto cost = [&](int s) { return s == 6 ? c_actor : (train_index(s) >= 0 ? c_keep : c_none); };
  std::vector<int32_t> order;
  for (int cls = 0; cls < 3; ++cls)  // actor, then the streams that keep activations, then the rest; natural order inside a class
    for (int r = 0; r < n_agents; ++r)
      for (int s = 0; s < kStreams; ++s) {
        const int c = s == 6 ? 0 : (train_index(s) >= 0 ? 1 : 2);
        if (c != cls) continue;
        for (int h = 0; h < 2; ++h) order.push_back((int32_t)(r << 4 | s << 1 | h));
      }
  std::vector<double> load(G, 0.0);
  std::vector<std::vector<int32_t>> lists(G);
  if (wide) {
    for (int32_t e : order) {
      int b = 0;
      for (int i = 1; i < G; ++i) if (load[i] < load[b]) b = i;
      lists[b].push_back(e);
      load[b] += cost((e >> 1) & 7);
    }
  } else {
    // Narrow shapes: natural (agent, stream, half) order dealt round-robin.  The cost-sorted deal evens the CTA end times
    // (25-35 us -> 29-35 us) without shortening the kernel - it is bound by L2 <-> SM traffic, DESIGN.md section 4 - and reads
    // 110 MB instead of 91 MB from DRAM per launch (profiles/r02j vs r02g): tiles of one agent no longer run together, so
    // fewer of its weight images are found in L2 by the second half batch.
    int i = 0;
    for (int r = 0; r < n_agents; ++r)
      for (int s = 0; s < kStreams; ++s)
        for (int h = 0; h < 2; ++h, ++i) lists[i % G].push_back((int32_t)(r << 4 | s << 1 | h));
  }
  std::vector<int32_t> off(G + 1, 0), flat;
  for (int i = 0; i < G; ++i) {
    off[i + 1] = off[i] + (int32_t)lists[i].size();
    flat.insert(flat.end(), lists[i].begin(), lists[i].end());
  }
  FwdSchedule sc;
  sc.dev = dev; sc.n_agents = n_agents; sc.n_sm = n_sm; sc.A = A; sc.drop = (int)drop; sc.grid = G;
  IQLB_CUDA(cudaMalloc((void**)&sc.d_off, off.size() * sizeof(int32_t)));
  IQLB_CUDA(cudaMalloc((void**)&sc.d_list, flat.size() * sizeof(int32_t)));
  IQLB_CUDA(cudaMemcpy(sc.d_off, off.data(), off.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
  IQLB_CUDA(cudaMemcpy(sc.d_list, flat.data(), flat.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
  cache.push_back(sc);
  *out = &cache.back();
  return IQLB_OK;
}

template <bool kWide>
static int persist_launch(const BfParams& P, const float* params, const float* target, float* head_out, int n_agents,
                          cudaStream_t stream, bool pdl) {
  static int n_sm = 0;
  static bool attr_set = false;
  constexpr uint32_t smem = fwdp::Lay<kWide>::kSmem;
  if (!attr_set) {
    int dev = 0;
    IQLB_CUDA(cudaGetDevice(&dev));
    IQLB_CUDA(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
    IQLB_CUDA(cudaFuncSetAttribute(bf16_fwd_persist_kernel<kWide>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_set = true;
  }
  IQLB_REQUIRE(P.L.B == 256, "forward kernel: batch 256 (two 128-row tiles per stream)");
  const FwdSchedule* sc = nullptr;
  const int rc = fwd_schedule(n_agents, n_sm, P.L.A, kWide, P.drop_thresh != 0, &sc);
  if (rc) return rc;
  IQLB_CUDA(launch_pdl(bf16_fwd_persist_kernel<kWide>, dim3(sc->grid), dim3(fwdp::kThreads), smem, stream, pdl, P, params, target,
                       head_out, n_agents, (const int32_t*)sc->d_off, (const int32_t*)sc->d_list));
  return IQLB_OK;
}

static bool persist_narrow(const BfParams& P) { return P.KXB == 1 && P.L.A <= 8 && P.drop_thresh == 0; }

int bf16_fwd_persist_launch(const BfParams& P, const float* params, const float* target, float* head_out, int n_agents,
                            cudaStream_t stream, bool pdl) {
  // every shape bf16_create accepts (hidden 256, batch 256, state_dim + action_dim <= 125, action_dim <= 24) fits the wide one
  IQLB_REQUIRE(P.KXB <= 2 && P.L.A <= kOutMax && P.L.H == 256, "forward kernel: shape outside what bf16_create accepts");
  return persist_narrow(P) ? persist_launch<false>(P, params, target, head_out, n_agents, stream, pdl)
                           : persist_launch<true>(P, params, target, head_out, n_agents, stream, pdl);
}

}  // namespace iqlb
