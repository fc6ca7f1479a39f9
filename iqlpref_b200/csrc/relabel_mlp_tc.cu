// Tensor-core reward-MLP relabel (split-bf16, fp32-level accuracy) for the shapes the reference uses:
// RewardMLP(d_in <= 128, 1, [w]*depth) with w in {64, 128} and depth <= 3
// (gp_reward-priors/optbnn/bnn/nets/mlp.py:73-93, layers/linear.py:43-47; called from
// /root/reference/algorithms/offline/iql.py:719-724, 982-1011, 1162-1187).
//
// Every product x*W is evaluated as three bf16 MMAs with fp32 accumulation,
//     x_hi*W_hi + x_lo*W_hi + x_hi*W_lo,     v = v_hi + v_lo + O(2^-18 |v|),
// so a layer carries a relative error of ~1e-5 per product (vs 4e-3 for plain bf16): well inside the
// 1e-3 parity bar of the relabelled rewards, at a third of the bf16 tensor rate instead of the FP32 pipe.
//
// One persistent CTA per SM works on pairs of 128-row subtiles.  The activations never touch shared
// memory: the A operand of every layer lives in TENSOR MEMORY (tcgen05.mma with A from TMEM), written
// there by the epilogue warps with tcgen05.st straight from the fp32 accumulator they have just drained
// (bias, activation, hi/lo split), or - for layer 0 - straight from the dataset rows in HBM.
//   TMEM, per subtile p: [ACC w cols fp32 | A_hi AW cols | A_lo AW cols],  AW = max(w, K0)/2
//   SMEM: the hi/lo weight images of all layers of ONE member (K-major UMMA images, 1-D bulk copies),
//         resident for a single model, streamed layer slot by layer slot for ensembles; 2 bias bundles.
//   warp 0: weight / bias producer (+ L2 prefetch of the next tile's rows)      warps 1, 2: MMA issuers of subtile 0, 1
//   warps 4-11: epilogue of subtile 0, warps 12-19: subtile 1 (two warps per TMEM lane quarter, one per column half)
// The two subtiles ping-pong: while the epilogue warps of one drain layer l, the tensor core runs layer l
// (or l+1) of the other.  The width->1 output layer is a dot product in the last epilogue.
// Measured (clock64 timeline, scripts/relabel_timeline.py): a 128x128x16 MMA with A from TMEM retires every ~110
// clocks; TMEM loads/stores of the epilogue warps queue behind the MMAs in flight (a tcgen05.st + wait::st costs
// ~1 k clocks while a 24-MMA batch runs), so variants with fewer, fatter epilogue threads or register prefetch of
// the next rows were slower than this one.
#include <vector>

#include "common.cuh"
#include "umma.cuh"

namespace iqlb {
using namespace umma;

namespace rtc {
constexpr int kMaxDepth = 3;
constexpr int kThreads = 640;  // producer, 2 MMA issuers, spare, 2 x 8 epilogue warps
enum { W_FULL = 0, W_EMPTY = kMaxDepth, B_FULL = 2 * kMaxDepth, B_EMPTY = B_FULL + 2, A_READY = B_EMPTY + 2,
       ACC_READY = A_READY + 2, N_BARS = ACC_READY + 2 };
}  // namespace rtc

struct RtcParams {
  const float* x;          // [rows][ldx]
  int64_t rows;
  int ldx, d_in;
  int w, depth, act;
  int S;                   // members
  int n_ms, ctas_per_ms;   // grid = n_ms * ctas_per_ms; CTA b works on member range b % n_ms, tiles b / n_ms, += ctas_per_ms
  int n_tiles;             // 256-row tiles
  const uint8_t* wimg;     // [S][member_bytes]: per layer [hi image | lo image], then the bias bundle
  int64_t member_bytes;
  uint32_t slot_off[rtc::kMaxDepth];    // byte offset of layer l inside a member block / inside shared memory
  uint32_t slot_bytes[rtc::kMaxDepth];  // 2 * KB_l * w * 128
  uint32_t bundle_off, bundle_bytes;    // floats: b_0 .. b_{depth-1} (w each) | wout (w) | bout, 0, 0, 0
  int ksteps0;             // ceil(d_in / 16)
  int aw;                  // TMEM columns of one A image (hi or lo)
  int skew_clocks;         // start-up delay of the second subtile's MMA issuer
  float scale0, scale_h, scale_out;  // 1/sqrt(n_in) of layer 0, hidden layers, output layer
  float* preds;            // [S][pred_stride]
  unsigned long long* dbg; // optional clock64 timeline of CTA 0 (iqlb_relabel_debug_timeline)
  int64_t pred_stride;
};

// fp32 member parameters (W stored [n_in][n_out]) -> hi/lo bf16 K-major images [n_out rows][n_in cols] + bias bundle
__global__ void __launch_bounds__(256) mlp_tc_prep_kernel(RtcParams P, const float* __restrict__ weights, int64_t pm) {
  const int s = blockIdx.y;
  const float* wm = weights + (int64_t)s * pm;
  uint8_t* dst = const_cast<uint8_t*>(P.wimg) + (int64_t)s * P.member_bytes;
  int64_t off = 0;  // running offset of W_l inside the member's parameters
  int n_in = P.d_in;
  for (int l = 0; l < P.depth; ++l) {
    const int kb = (n_in + 63) / 64;
    const int chunks = kb * 8 * P.w;  // 16-byte chunks of one image
    uint8_t* hi = dst + P.slot_off[l];
    uint8_t* lo = hi + P.slot_bytes[l] / 2;
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < chunks; c += gridDim.x * blockDim.x) {
      const int n = c % P.w, k0 = (c / P.w) * 8;
      uint32_t h[4], q[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float a[2];
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int k = k0 + 2 * j + e;
          a[e] = k < n_in ? __ldg(wm + off + (int64_t)k * P.w + n) : 0.f;
        }
        const __nv_bfloat162 hh = __floats2bfloat162_rn(a[0], a[1]);
        const float2 hf = __bfloat1622float2(hh);
        h[j] = *reinterpret_cast<const uint32_t*>(&hh);
        q[j] = pack_bf16(a[0] - hf.x, a[1] - hf.y);
      }
      const int64_t o = image_off(P.w, n, k0);
      *reinterpret_cast<uint4*>(hi + o) = make_uint4(h[0], h[1], h[2], h[3]);
      *reinterpret_cast<uint4*>(lo + o) = make_uint4(q[0], q[1], q[2], q[3]);
    }
    // biases
    float* bun = reinterpret_cast<float*>(dst + P.bundle_off);
    for (int n = blockIdx.x * blockDim.x + threadIdx.x; n < P.w; n += gridDim.x * blockDim.x)
      bun[l * P.w + n] = __ldg(wm + off + (int64_t)n_in * P.w + n);
    off += (int64_t)n_in * P.w + P.w;
    n_in = P.w;
  }
  float* bun = reinterpret_cast<float*>(dst + P.bundle_off);
  for (int n = blockIdx.x * blockDim.x + threadIdx.x; n < P.w + 4; n += gridDim.x * blockDim.x)
    bun[P.depth * P.w + n] = n <= P.w ? __ldg(wm + off + n) : 0.f;  // wout[w], bout
}

// hi/lo split of 32 fp32 values into two groups of 16 packed bf16x2 words
__device__ __forceinline__ void split32(const float (&v)[32], uint32_t (&hi)[16], uint32_t (&lo)[16]) {
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
    const float2 hf = __bfloat1622float2(h);
    hi[j] = *reinterpret_cast<const uint32_t*>(&h);
    lo[j] = pack_bf16(v[2 * j] - hf.x, v[2 * j + 1] - hf.y);
  }
}

__device__ __forceinline__ void rtc_stamp(const RtcParams& P, int role, int idx) {
  if (P.dbg != nullptr && blockIdx.x == 0 && idx < 512) P.dbg[role * 512 + idx] = (unsigned long long)clock64();
}

template <bool kRelu, int kW>
__global__ void __maxnreg__(96) relabel_mlp_tc_kernel(const RtcParams P) {
  using namespace rtc;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t oBun = P.slot_off[P.depth - 1] + P.slot_bytes[P.depth - 1];
  const uint32_t oBar = oBun + 2 * P.bundle_bytes;
  const uint32_t oTmem = oBar + N_BARS * 8;
  const uint32_t oRed = oTmem + 16;  // 256 partial dot products
  auto bar = [&](int i) { return base + oBar + 8u * (uint32_t)i; };
  const int warp = warp_uniform(), lane = threadIdx.x & 31;
  constexpr int w = kW;
  const int depth = P.depth;

  if (threadIdx.x == 0) {
    for (int l = 0; l < kMaxDepth; ++l) { mbar_init(bar(W_FULL + l), 1); mbar_init(bar(W_EMPTY + l), 2); }
    for (int q = 0; q < 2; ++q) {
      mbar_init(bar(B_FULL + q), 1); mbar_init(bar(B_EMPTY + q), 512);
      mbar_init(bar(A_READY + q), 256); mbar_init(bar(ACC_READY + q), 1);
    }
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(base + oTmem, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<const uint32_t*>(gbase + oTmem);
  const uint32_t PC = (uint32_t)(w + 2 * P.aw);  // TMEM columns per subtile

  // work of this CTA: a fixed member range, every ctas_per_ms-th 256-row tile
  const int ms = blockIdx.x % P.n_ms;
  const int s0 = (int)((int64_t)ms * P.S / P.n_ms), s1 = (int)((int64_t)(ms + 1) * P.S / P.n_ms);
  const int nm = s1 - s0;
  const int tile0 = blockIdx.x / P.n_ms;
  const int my_tiles = tile0 < P.n_tiles ? (P.n_tiles - tile0 + P.ctas_per_ms - 1) / P.ctas_per_ms : 0;
  const int n_it = my_tiles * nm;
  const bool reload = nm > 1;

  if (warp == 0) {
    if (elect_one()) {  // single-thread roles are chosen by elect.sync: the operands of the bulk copies / MMAs stay uniform (umma.cuh)
      const bool x_aligned = (P.ldx % 4 == 0) && ((reinterpret_cast<uintptr_t>(P.x) & 15) == 0);
      for (int i = 0; i < n_it; ++i) {
        const uint8_t* mb = P.wimg + (int64_t)(s0 + i % nm) * P.member_bytes;
        const int q = i & 1;
        if (i % nm == 0 && x_aligned) {  // dataset rows of the NEXT tile: HBM -> L2 while this one computes
          const int64_t r_next = ((int64_t)tile0 + (int64_t)(i / nm + 1) * P.ctas_per_ms) * 256;
          if (r_next < P.rows) {
            const int64_t nr = P.rows - r_next < 256 ? P.rows - r_next : 256;
            const char* src = reinterpret_cast<const char*>(P.x + r_next * P.ldx);
            const int64_t bytes = (nr * P.ldx * 4) & ~15ll;
            for (int64_t o = 0; o < bytes; o += 16384) bulk_prefetch_l2(src + o, (uint32_t)(bytes - o < 16384 ? bytes - o : 16384));
          }
        }
        if (i >= 2) mbar_wait(bar(B_EMPTY + q), (uint32_t)(((i >> 1) - 1) & 1));
        mbar_expect_tx(bar(B_FULL + q), P.bundle_bytes);
        bulk_g2s(base + oBun + q * P.bundle_bytes, mb + P.bundle_off, P.bundle_bytes, bar(B_FULL + q));
        if (reload || i == 0) {
          for (int l = 0; l < depth; ++l) {
            if (i > 0) mbar_wait(bar(W_EMPTY + l), (uint32_t)((i - 1) & 1));
            mbar_expect_tx(bar(W_FULL + l), P.slot_bytes[l]);
            for (uint32_t o = 0; o < P.slot_bytes[l]; o += 32768u) {
              const uint32_t n = P.slot_bytes[l] - o < 32768u ? P.slot_bytes[l] - o : 32768u;
              bulk_g2s(base + P.slot_off[l] + o, mb + P.slot_off[l] + o, n, bar(W_FULL + l));
            }
          }
        }
      }
    }
  } else if (warp == 1 || warp == 2) {
    // One MMA issuer per subtile (scripts/mma_rate.py: with A in tensor memory one elected thread issues a 128 x 128 x 16
    // MMA every 72 clocks = 88 % of the tensor pipe; chosen by `lane == 0` it was 122 clocks, see umma.cuh).
    if (elect_one()) {
      const int p = warp - 1;
      const uint32_t idesc = instr_desc_bf16(w, 0, 0);
      const uint32_t acc = tmem + p * PC, ah = acc + w, al = ah + P.aw;
      uint32_t na = 0;
      if (p == 1 && n_it > 0) {
        // stagger the subtiles by about half a layer, so that one is in its epilogue while the other is in the
        // tensor pipe (started together they stay in lockstep: MMAs overlap MMAs and epilogues overlap epilogues)
        mbar_wait(bar(ACC_READY + 0), 0);
        const long long t0 = clock64();
        while (clock64() - t0 < P.skew_clocks) {
        }
      }
      for (int i = 0; i < n_it; ++i) {
        for (int l = 0; l < depth; ++l) {
          if (reload || i == 0) mbar_wait(bar(W_FULL + l), reload ? (uint32_t)(i & 1) : 0u);
          const int ks = l == 0 ? P.ksteps0 : w / 16;
          const uint64_t bhi = smem_desc(base + P.slot_off[l], 16, 1024);
          const uint64_t blo = smem_desc(base + P.slot_off[l] + P.slot_bytes[l] / 2, 16, 1024);
          mbar_wait(bar(A_READY + p), na & 1);
          ++na;
          tc_fence_after();
          rtc_stamp(P, 0, ((i * 3 + l) * 2 + p) * 2);
          uint32_t accum = 0;
#pragma unroll 1
          for (int term = 0; term < 3; ++term) {
            const uint32_t a = term == 1 ? al : ah;
            const uint64_t b = term == 2 ? blo : bhi;
#pragma unroll 8
            for (int k = 0; k < ks; ++k) {
              // the descriptor's start-address field is (address >> 4): advancing it is one 64-bit add
              mma_bf16_ts(acc, a + 8u * k, b + (uint64_t)(((uint32_t)(k >> 2) * (uint32_t)w * 128u + (uint32_t)(k & 3) * 32u) >> 4),
                          idesc, accum);
              accum = 1;
            }
          }
          mma_commit(bar(ACC_READY + p));
          rtc_stamp(P, 0, ((i * 3 + l) * 2 + p) * 2 + 1);
          if (reload && i + 1 < n_it) mma_commit(bar(W_EMPTY + l));
        }
      }
    }
  } else if (warp == 3) {
    // spare warp: keeps the epilogue warps aligned to the TMEM lane quarters (quarter = warp % 4)
  } else {
    // 16 epilogue warps: subtile p, TMEM lane quarter (= warp % 4, the lanes a warp may touch), column half
    const int ew = warp - 4, p = ew >> 3, quarter = warp & 3, half = (ew & 7) >> 2;
    const uint32_t acc = tmem + ((uint32_t)(quarter * 32) << 16) + p * PC, ah = acc + w, al = ah + P.aw;
    const bool vec = (P.ldx % 4 == 0) && ((reinterpret_cast<uintptr_t>(P.x) & 15) == 0);
    const int k0r32 = (P.d_in + 31) / 32 * 32;
    const int hc = w / 2, col0 = half * hc;  // this thread's accumulator columns [col0, col0 + hc)
    float* red = reinterpret_cast<float*>(gbase + oRed) + p * 128 + quarter * 32 + lane;
    const bool stamper = quarter == 0 && lane == 0;
    uint32_t nacc = 0;
    for (int i = 0; i < n_it; ++i) {
      const int tile = tile0 + (i / nm) * P.ctas_per_ms, s = s0 + i % nm;
      const int64_t row = (int64_t)tile * 256 + p * 128 + quarter * 32 + lane;
      const bool valid = row < P.rows;
      const float* xr = P.x + (valid ? row : 0) * P.ldx;
      if (stamper) rtc_stamp(P, 1 + p + 2 * half, (i * 4) * 2);
      // layer-0 operand: dataset row -> hi/lo bf16 in tensor memory (32-column chunks alternate between the halves)
      for (int c0 = half * 32; c0 < k0r32; c0 += 64) {
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          const int k = c0 + j;
          float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
          if (valid && k < P.d_in) {
            if (vec && k + 4 <= P.ldx) {
              t = __ldg(reinterpret_cast<const float4*>(xr + k));
            } else {
              t.x = __ldg(xr + k);
              if (k + 1 < P.d_in) t.y = __ldg(xr + k + 1);
              if (k + 2 < P.d_in) t.z = __ldg(xr + k + 2);
              if (k + 3 < P.d_in) t.w = __ldg(xr + k + 3);
            }
          }
          v[j] = t.x;
          v[j + 1] = k + 1 < P.d_in ? t.y : 0.f;
          v[j + 2] = k + 2 < P.d_in ? t.z : 0.f;
          v[j + 3] = k + 3 < P.d_in ? t.w : 0.f;
        }
        uint32_t hi[16], lo[16];
        split32(v, hi, lo);
        tmem_st16(ah + (uint32_t)(c0 >> 1), hi);
        tmem_st16(al + (uint32_t)(c0 >> 1), lo);
      }
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(bar(A_READY + p));
      if (stamper) rtc_stamp(P, 1 + p + 2 * half, (i * 4) * 2 + 1);
      const int q = i & 1;
      mbar_wait(bar(B_FULL + q), (uint32_t)((i >> 1) & 1));
      const float* bun = reinterpret_cast<const float*>(gbase + oBun + q * P.bundle_bytes);
      for (int l = 0; l < depth; ++l) {
        mbar_wait(bar(ACC_READY + p), nacc & 1);
        ++nacc;
        tc_fence_after();
        if (stamper) rtc_stamp(P, 1 + p + 2 * half, (i * 4 + l + 1) * 2);
        const float scale = l == 0 ? P.scale0 : P.scale_h;
        const float* b = bun + l * w;
        if (l < depth - 1) {
          for (int c = col0; c < col0 + hc; c += 32) {
            float v[32];
            tmem_ld32(acc + (uint32_t)c, v);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const float4 b4 = *reinterpret_cast<const float4*>(b + c + j);
              v[j] = fmaf(scale, v[j], b4.x); v[j + 1] = fmaf(scale, v[j + 1], b4.y);
              v[j + 2] = fmaf(scale, v[j + 2], b4.z); v[j + 3] = fmaf(scale, v[j + 3], b4.w);
            }
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = kRelu ? fmaxf(v[j], 0.f) : apply_act(P.act, v[j]);
            uint32_t hi[16], lo[16];
            split32(v, hi, lo);
            tmem_st16(ah + (uint32_t)(c >> 1), hi);
            tmem_st16(al + (uint32_t)(c >> 1), lo);
          }
          tmem_st_wait();
          tc_fence_before();
          mbar_arrive(bar(A_READY + p));
          if (stamper) rtc_stamp(P, 1 + p + 2 * half, (i * 4 + l + 1) * 2 + 1);
        } else {
          const float* wo = bun + depth * w;
          float dot = 0.f;
          for (int c = col0; c < col0 + hc; c += 32) {
            float v[32];
            tmem_ld32(acc + (uint32_t)c, v);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const float4 b4 = *reinterpret_cast<const float4*>(b + c + j);
              const float4 w4 = *reinterpret_cast<const float4*>(wo + c + j);
              float h0 = fmaf(scale, v[j], b4.x), h1 = fmaf(scale, v[j + 1], b4.y);
              float h2 = fmaf(scale, v[j + 2], b4.z), h3 = fmaf(scale, v[j + 3], b4.w);
              h0 = kRelu ? fmaxf(h0, 0.f) : apply_act(P.act, h0);
              h1 = kRelu ? fmaxf(h1, 0.f) : apply_act(P.act, h1);
              h2 = kRelu ? fmaxf(h2, 0.f) : apply_act(P.act, h2);
              h3 = kRelu ? fmaxf(h3, 0.f) : apply_act(P.act, h3);
              dot = fmaf(h0, w4.x, dot); dot = fmaf(h1, w4.y, dot);
              dot = fmaf(h2, w4.z, dot); dot = fmaf(h3, w4.w, dot);
            }
          }
          // the two column halves of a row meet in shared memory (named barrier of the subtile's 256 threads)
          if (half == 1) *red = dot;
          asm volatile("bar.sync %0, 256;" ::"r"(1 + p) : "memory");
          if (half == 0 && valid) P.preds[(int64_t)s * P.pred_stride + row] = fmaf(P.scale_out, dot + *red, wo[w]);
          if (stamper) rtc_stamp(P, 1 + p + 2 * half, (i * 4 + l + 1) * 2 + 1);
        }
      }
      mbar_arrive(bar(B_EMPTY + q));
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

// Shapes the tensor-core path covers; everything else stays on the fp32 FFMA path (relabel_mlp.cu).
bool relabel_mlp_tc_supported(int d_in, const int32_t* widths, int depth) {
  if (depth < 1 || depth > rtc::kMaxDepth || d_in < 1 || d_in > 128) return false;
  const int w = widths[0];
  if (w != 64 && w != 128) return false;
  for (int l = 1; l < depth; ++l)
    if (widths[l] != w) return false;
  size_t smem = 0;
  for (int l = 0; l < depth; ++l) smem += 2u * (size_t)(((l == 0 ? d_in : w) + 63) / 64) * w * 128u;
  smem += 2u * (size_t)(depth * w + w + 4) * 4u + rtc::N_BARS * 8 + 16 + 1024 + 1024;
  return smem <= 227u * 1024u;
}

unsigned long long* g_rtc_dbg = nullptr;

struct RelabelTc {
  RtcParams P;
  size_t smem;
  int n_sm;
  bool relu;
  void (*kern)(const RtcParams);
  uint8_t* wimg;
};

// Builds the hi/lo weight images of all members (stream-ordered allocation, freed by relabel_tc_end).
int relabel_tc_begin(RelabelTc** ctx, int ldx, int d_in, const float* weights, int S, const int32_t* widths, int depth,
                     int activation, cudaStream_t st) {
  IQLB_REQUIRE(relabel_mlp_tc_supported(d_in, widths, depth), "shape not covered by the tensor-core relabel");
  RelabelTc* c = new RelabelTc();
  RtcParams& P = c->P;
  P = RtcParams{};
  const int w = widths[0];
  P.ldx = ldx; P.d_in = d_in; P.w = w; P.depth = depth; P.act = activation; P.S = S;
  uint32_t off = 0;
  int64_t pm = 0;
  for (int l = 0; l < depth; ++l) {
    const int n_in = l == 0 ? d_in : w;
    P.slot_off[l] = off;
    P.slot_bytes[l] = 2u * (uint32_t)((n_in + 63) / 64) * (uint32_t)w * 128u;
    off += P.slot_bytes[l];
    pm += (int64_t)n_in * w + w;
  }
  pm += w + 1;
  P.bundle_off = off;
  P.bundle_bytes = (uint32_t)(depth * w + w + 4) * 4u;
  P.member_bytes = ((int64_t)off + P.bundle_bytes + 127) / 128 * 128;
  P.ksteps0 = (d_in + 15) / 16;
  const int k0r32 = (d_in + 31) / 32 * 32;
  P.aw = (w > k0r32 ? w : k0r32) / 2;
  P.scale0 = 1.0f / sqrtf((float)d_in);
  P.scale_h = 1.0f / sqrtf((float)w);
  P.scale_out = P.scale_h;
  P.skew_clocks = getenv("IQLB_RELABEL_SKEW") ? atoi(getenv("IQLB_RELABEL_SKEW")) : 1200;
  c->smem = (size_t)off + 2u * P.bundle_bytes + rtc::N_BARS * 8 + 16 + 1024 + 1024;
  c->relu = activation == IQLB_ACT_RELU;
  c->wimg = nullptr;
  int dev = 0;
  c->n_sm = 148;
  cudaError_t e = cudaGetDevice(&dev);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&c->n_sm, cudaDevAttrMultiProcessorCount, dev);
  c->kern = c->relu ? (w == 128 ? relabel_mlp_tc_kernel<true, 128> : relabel_mlp_tc_kernel<true, 64>)
                    : (w == 128 ? relabel_mlp_tc_kernel<false, 128> : relabel_mlp_tc_kernel<false, 64>);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(c->kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c->smem);
  if (e == cudaSuccess) e = cudaMallocAsync((void**)&c->wimg, (size_t)S * (size_t)P.member_bytes, st);
  if (e == cudaSuccess) {
    P.wimg = c->wimg;
    mlp_tc_prep_kernel<<<dim3(8, S), 256, 0, st>>>(P, weights, pm);
    e = cudaGetLastError();
  }
  if (e != cudaSuccess) {
    if (c->wimg) cudaFreeAsync(c->wimg, st);
    delete c;
    return fail(IQLB_ERR_CUDA, "tensor-core relabel set-up: %s", cudaGetErrorString(e));
  }
  *ctx = c;
  return IQLB_OK;
}

// preds[s * pred_stride + i] = member s on row i of x[rows][ldx]
int relabel_tc_forward(RelabelTc* c, const float* x, int64_t rows, float* preds, int64_t pred_stride, cudaStream_t st) {
  RtcParams P = c->P;
  P.dbg = g_rtc_dbg;
  P.x = x; P.rows = rows; P.preds = preds; P.pred_stride = pred_stride;
  P.n_tiles = (int)((rows + 255) / 256);
  if (P.n_tiles >= c->n_sm) { P.n_ms = 1; P.ctas_per_ms = c->n_sm; }
  else { P.n_ms = c->n_sm / P.n_tiles; if (P.n_ms > P.S) P.n_ms = P.S; P.ctas_per_ms = P.n_tiles; }
  c->kern<<<P.n_ms * P.ctas_per_ms, rtc::kThreads, c->smem, st>>>(P);
  IQLB_LAUNCH_CHECK();
  return IQLB_OK;
}

void relabel_tc_end(RelabelTc* c, cudaStream_t st) {
  if (!c) return;
  if (c->wimg) cudaFreeAsync(c->wimg, st);
  delete c;
}

}  // namespace iqlb

// Debug: CTA 0 of the next tensor-core relabel launches writes clock64 stamps into buf (device, >= 5*512 u64;
// role 0 = MMA issuer, 1..4 = first epilogue thread of subtile 0/1, column half 0/1); NULL switches it off.
extern "C" int iqlb_relabel_debug_timeline(unsigned long long* buf) {
  iqlb::g_rtc_dbg = buf;
  return IQLB_OK;
}
