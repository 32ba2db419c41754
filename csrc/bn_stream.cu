// Streaming BatchNorm kernels, second generation (forward statistics / apply, backward reduce / apply) for NHWC
// bf16 activations whose channel count is a power of two between 8 and 2048 — every BatchNorm of the ResNet
// jobs (model/model_factory.py:47-55). Same maths and the same C entry points as csrc/bn_kernels.cu, which
// keeps serving every other channel count.
//
// Why a second generation. ncu on B200 (profiles/ncu_summary_r1d.md) put the first-generation kernels at 33 %
// (statistics) and 62-66 % (the other three) of DRAM peak with 22-32 % of the warps resident: each thread kept
// 4-8 register-staged 16-byte loads in flight and then stalled on them, so an SM had ~64 KB in flight only in
// bursts. Here a producer warp streams every input tensor with 1-D bulk copies (cp.async.bulk, the TMA engine)
// through a ring of shared-memory stages signalled by mbarriers — 64-96 KB per CTA and 2-3 CTAs per SM are in
// flight ALL the time, independent of registers — while eight consumer warps read the stages with conflict-free
// 16-byte shared loads, do the arithmetic and (apply kernels) write with coalesced 16-byte stores.
//
// A CTA owns a contiguous row range (all channels), so its traffic is one sequential stream per tensor. Thread
// t of the 256 consumers always sees the same 8 channels (vector column t % (C/8)), which keeps the
// per-channel coefficients in registers. Reductions write per-CTA partials that a small second kernel folds in
// a fixed order (compensated fp32 sums): results are bitwise reproducible.
#include "common.cuh"
#include "tc_ptx.cuh"
#include "../include/tiresias_b200.h"

using namespace tb200;
using namespace tb200::ptx;

namespace {

constexpr int kConsumers = 256;
constexpr int kBlock = kConsumers + 32;          // + one producer warp
constexpr int kStages = 4;

__device__ __forceinline__ void unpack8(const uint4& r, float (&f)[8]) {
  f[0] = bf16_lo(r.x); f[1] = bf16_hi(r.x); f[2] = bf16_lo(r.y); f[3] = bf16_hi(r.y);
  f[4] = bf16_lo(r.z); f[5] = bf16_hi(r.z); f[6] = bf16_lo(r.w); f[7] = bf16_hi(r.w);
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint4 o;
  o.x = pack_bf16(f[0], f[1]); o.y = pack_bf16(f[2], f[3]);
  o.z = pack_bf16(f[4], f[5]); o.w = pack_bf16(f[6], f[7]);
  return o;
}
__device__ __forceinline__ float load_param(const void* p, int dtype, int i) {
  return dtype == kBF16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[i])
                        : reinterpret_cast<const float*>(p)[i];
}
__device__ __forceinline__ void store_param(void* p, int dtype, int i, float v) {
  if (dtype == kBF16) reinterpret_cast<__nv_bfloat16*>(p)[i] = __float2bfloat16(v);
  else reinterpret_cast<float*>(p)[i] = v;
}

struct StreamArgs {
  const uint8_t* in[3];       // streamed inputs (bf16 [rows, C])
  uint4* out[2];              // outputs of the apply kernels
  size_t total_vecs;          // rows * C / 8
  size_t vecs_per_cta;        // multiple of the piece size
  int C, vecC;
  int nin;
  // per-channel inputs / outputs
  const float* coef;          // apply: [scale | shift]; bwd apply: [A | mean g | mean g*xhat]
  const float* save_mean;
  const float* save_invstd;
  const void* gamma;
  const void* beta;
  int param_dtype;
  int relu;
  float* partials;            // [gridDim.x][C][2]
  const uint8_t* mask_in;     // backward MASK == 3: one byte per 16-byte vector, bit k = (output element k > 0)
  uint8_t* mask_out;          // forward with residual + ReLU: the same bits, written next to y
};

// The pipeline shared by the four kernels. PIECE_VECS 16-byte vectors per tensor per stage. `body(stage_base,
// first_vec, nvec)` is called by the 256 consumer threads for every piece: tensor i of the stage lives at
// stage_base + i * PIECE_VECS * 16.
// MASKBITS adds a byte stream (a.mask_in, one byte per vector) behind the NIN tensors of every stage.
template <int NIN, int PIECE_VECS, bool MASKBITS = false, typename Body>
__device__ __forceinline__ void stream_pipeline(const StreamArgs& a, uint8_t* smem_raw, Body&& body) {
  constexpr uint32_t kPieceBytes = PIECE_VECS * 16;
  constexpr uint32_t kStageBytes = NIN * kPieceBytes + (MASKBITS ? PIECE_VECS : 0);
  const uint32_t base = smem_u32(smem_raw);
  const uint32_t bars = base + kStages * kStageBytes;
  auto full = [&](int s) { return bars + 8u * s; };
  auto empty = [&](int s) { return bars + 8u * (kStages + s); };
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(full(s), 1); mbar_init(empty(s), kConsumers / 32); }
    mbar_init_fence();
  }
  __syncthreads();
  const size_t v0 = (size_t)blockIdx.x * a.vecs_per_cta;
  size_t v1 = v0 + a.vecs_per_cta;
  if (v1 > a.total_vecs) v1 = a.total_vecs;
  const int pieces = v0 < v1 ? (int)((v1 - v0 + PIECE_VECS - 1) / PIECE_VECS) : 0;
  if (warp == kConsumers / 32) {
    // ---- producer warp: one elected lane issues the bulk copies ----
    int s = 0;
    uint32_t ph = 0;
    for (int i = 0; i < pieces; ++i) {
      mbar_wait(empty(s), ph ^ 1);
      if (elect_one()) {
        const size_t pv = v0 + (size_t)i * PIECE_VECS;
        const uint32_t nv = (uint32_t)((v1 - pv) < (size_t)PIECE_VECS ? (v1 - pv) : (size_t)PIECE_VECS);
        const uint32_t mbytes = MASKBITS ? ((nv + 15u) & ~15u) : 0u;      // bulk copies move multiples of 16 bytes
        mbar_expect_tx(full(s), NIN * nv * 16u + mbytes);
#pragma unroll
        for (int t = 0; t < NIN; ++t)
          bulk_load(base + (uint32_t)s * kStageBytes + (uint32_t)t * kPieceBytes, a.in[t] + pv * 16, nv * 16u, full(s));
        if constexpr (MASKBITS)
          bulk_load(base + (uint32_t)s * kStageBytes + NIN * kPieceBytes, a.mask_in + pv, mbytes, full(s));
      }
      __syncwarp();
      if (++s == kStages) { s = 0; ph ^= 1; }
    }
  } else {
    // ---- consumers ----
    int s = 0;
    uint32_t ph = 0;
    for (int i = 0; i < pieces; ++i) {
      const size_t pv = v0 + (size_t)i * PIECE_VECS;
      const uint32_t nv = (uint32_t)((v1 - pv) < (size_t)PIECE_VECS ? (v1 - pv) : (size_t)PIECE_VECS);
      mbar_wait(full(s), ph);
      body(base + (uint32_t)s * kStageBytes, pv, nv);
      __syncwarp();
      if (lane == 0) mbar_arrive(empty(s));
      if (++s == kStages) { s = 0; ph ^= 1; }
    }
  }
}

// Sum the per-thread 8-channel accumulators of the consumers that share a vector column, in a fixed order,
// and write the CTA's partial row [C][2].
__device__ __forceinline__ void reduce_and_store(const StreamArgs& a, float (&acc)[2][8], float* red /* [2][256][8] */) {
  const int t = threadIdx.x;
  __syncthreads();                              // every warp (producer included) is past the pipeline
  if (t < kConsumers) {
#pragma unroll
    for (int v = 0; v < 2; ++v)
#pragma unroll
      for (int k = 0; k < 8; ++k) red[(v * kConsumers + t) * 8 + k] = acc[v][k];
  }
  __syncthreads();
  const int groups = kConsumers / a.vecC;       // threads per vector column
  float* out = a.partials + (size_t)blockIdx.x * a.C * 2;
  for (int e = t; e < a.C * 2; e += kBlock) {
    const int ch = e >> 1, v = e & 1;
    const int col = ch >> 3, k = ch & 7;
    float s = 0.f;
    for (int g = 0; g < groups; ++g) s += red[(v * kConsumers + col + g * a.vecC) * 8 + k];
    out[e] = s;
  }
}

template <int PIECE_VECS>
__global__ void __launch_bounds__(kBlock, 3) bn_stats_stream_kernel(const StreamArgs a) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  float acc[2][8] = {};
  const int t = threadIdx.x;
  stream_pipeline<1, PIECE_VECS>(a, smem_raw, [&](uint32_t stage, size_t, uint32_t nv) {
#pragma unroll
    for (int k = 0; k < PIECE_VECS / kConsumers; ++k) {
      const uint32_t i = (uint32_t)t + (uint32_t)k * kConsumers;
      if (i < nv) {
        float f[8];
        unpack8(ld_shared_16(stage + i * 16u), f);
#pragma unroll
        for (int j = 0; j < 8; ++j) { acc[0][j] += f[j]; acc[1][j] = fmaf(f[j], f[j], acc[1][j]); }
      }
    }
  });
  reduce_and_store(a, acc, reinterpret_cast<float*>(smem_raw));
}

template <int NIN, int PIECE_VECS, bool WMASK = false>
__global__ void __launch_bounds__(kBlock, 3) bn_apply_stream_kernel(const StreamArgs a) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  const int t = threadIdx.x;
  float sc[8], sh[8];
  if (t < kConsumers) {
    const int col = t % a.vecC;
#pragma unroll
    for (int k = 0; k < 8; ++k) { sc[k] = a.coef[col * 8 + k]; sh[k] = a.coef[a.C + col * 8 + k]; }
  }
  const int relu = a.relu;
  stream_pipeline<NIN, PIECE_VECS>(a, smem_raw, [&](uint32_t stage, size_t pv, uint32_t nv) {
#pragma unroll
    for (int k = 0; k < PIECE_VECS / kConsumers; ++k) {
      const uint32_t i = (uint32_t)t + (uint32_t)k * kConsumers;
      if (i < nv) {
        float f[8], g[8];
        unpack8(ld_shared_16(stage + i * 16u), f);
        if constexpr (NIN == 2) unpack8(ld_shared_16(stage + PIECE_VECS * 16u + i * 16u), g);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float v = fmaf(f[j], sc[j], sh[j]);
          if constexpr (NIN == 2) v += g[j];
          f[j] = relu ? fmaxf(v, 0.f) : v;
        }
        a.out[0][pv + i] = pack8(f);       // default caching: the next layer reads y straight away
        if constexpr (WMASK) {
          // the backward's ReLU mask as one bit per element (1/16 of y): a residual layer then never re-reads y
          uint32_t bits = 0;
#pragma unroll
          for (int j = 0; j < 8; ++j) bits |= (f[j] > 0.f ? 1u : 0u) << j;
          a.mask_out[pv + i] = (uint8_t)bits;
        }
      }
    }
  });
}

// MASK: 0 = no ReLU, 1 = mask from the saved output y (third stream), 2 = mask recomputed from x,
//       3 = mask bits written by the forward (byte stream, 1/16 of a tensor).
template <int MASK, int PIECE_VECS>
__global__ void __launch_bounds__(kBlock, 2) bn_bwd_reduce_stream_kernel(const StreamArgs a) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  constexpr int NIN = MASK == 1 ? 3 : 2;
  const int t = threadIdx.x;
  float mean[8], istd[8], sc[8], sh[8];
  if (t < kConsumers) {
    const int col = t % a.vecC;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int ch = col * 8 + k;
      mean[k] = a.save_mean[ch];
      istd[k] = a.save_invstd[ch];
      sc[k] = sh[k] = 0.f;
      if constexpr (MASK == 2) {
        sc[k] = load_param(a.gamma, a.param_dtype, ch) * istd[k];
        sh[k] = load_param(a.beta, a.param_dtype, ch) - mean[k] * sc[k];
      }
    }
  }
  float acc[2][8] = {};
  stream_pipeline<NIN, PIECE_VECS, MASK == 3>(a, smem_raw, [&](uint32_t stage, size_t, uint32_t nv) {
#pragma unroll
    for (int k = 0; k < PIECE_VECS / kConsumers; ++k) {
      const uint32_t i = (uint32_t)t + (uint32_t)k * kConsumers;
      if (i < nv) {
        float g[8], xv[8], yv[8];
        unpack8(ld_shared_16(stage + i * 16u), g);
        unpack8(ld_shared_16(stage + PIECE_VECS * 16u + i * 16u), xv);
        if constexpr (MASK == 1) unpack8(ld_shared_16(stage + 2 * PIECE_VECS * 16u + i * 16u), yv);
        if constexpr (MASK == 3) {
          const uint32_t bits = smem_raw[(stage - smem_u32(smem_raw)) + 2 * PIECE_VECS * 16u + i];
#pragma unroll
          for (int j = 0; j < 8; ++j) yv[j] = ((bits >> j) & 1u) ? 1.f : 0.f;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          if constexpr (MASK == 2) yv[j] = fmaf(xv[j], sc[j], sh[j]);
          const float gj = (MASK != 0 && !(yv[j] > 0.f)) ? 0.f : g[j];
          acc[0][j] += gj;
          acc[1][j] = fmaf(gj, (xv[j] - mean[j]) * istd[j], acc[1][j]);
        }
      }
    }
  });
  reduce_and_store(a, acc, reinterpret_cast<float*>(smem_raw));
}

template <int MASK, bool DRES, int PIECE_VECS>
__global__ void __launch_bounds__(kBlock, 2) bn_bwd_apply_stream_kernel(const StreamArgs a) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  constexpr int NIN = MASK == 1 ? 3 : 2;
  const int t = threadIdx.x;
  float A[8], mg[8], mgx[8], mean[8], istd[8], sh[8];
  if (t < kConsumers) {
    const int col = t % a.vecC;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int ch = col * 8 + k;
      A[k] = a.coef[ch];                         // gamma * invstd == the forward's scale
      mg[k] = a.coef[a.C + ch];
      mgx[k] = a.coef[2 * a.C + ch];
      mean[k] = a.save_mean[ch];
      istd[k] = a.save_invstd[ch];
      sh[k] = MASK == 2 ? load_param(a.beta, a.param_dtype, ch) - mean[k] * A[k] : 0.f;
    }
  }
  stream_pipeline<NIN, PIECE_VECS, MASK == 3>(a, smem_raw, [&](uint32_t stage, size_t pv, uint32_t nv) {
#pragma unroll
    for (int k = 0; k < PIECE_VECS / kConsumers; ++k) {
      const uint32_t i = (uint32_t)t + (uint32_t)k * kConsumers;
      if (i < nv) {
        float g[8], xv[8], yv[8];
        unpack8(ld_shared_16(stage + i * 16u), g);
        unpack8(ld_shared_16(stage + PIECE_VECS * 16u + i * 16u), xv);
        if constexpr (MASK == 1) unpack8(ld_shared_16(stage + 2 * PIECE_VECS * 16u + i * 16u), yv);
        if constexpr (MASK == 3) {
          const uint32_t bits = smem_raw[(stage - smem_u32(smem_raw)) + 2 * PIECE_VECS * 16u + i];
#pragma unroll
          for (int j = 0; j < 8; ++j) yv[j] = ((bits >> j) & 1u) ? 1.f : 0.f;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          if constexpr (MASK == 2) yv[j] = fmaf(xv[j], A[j], sh[j]);
          const float gj = (MASK != 0 && !(yv[j] > 0.f)) ? 0.f : g[j];
          g[j] = gj;
          xv[j] = A[j] * (gj - mg[j] - (xv[j] - mean[j]) * istd[j] * mgx[j]);
        }
        a.out[0][pv + i] = pack8(xv);
        if constexpr (DRES) a.out[1][pv + i] = pack8(g);
      }
    }
  });
}

// ---- second pass of the reductions: fold the per-CTA partials (fixed order) -------------------------------
struct FoldArgs {
  const float* partials;      // [nparts][C][2]
  int nparts, C;
  double count;
  double inv_count, bessel;   // 1 / count and count / (count - 1): double-precision division costs ~1 us per
                              // channel thread on B200, so the host does it
  const void* gamma;
  const void* beta;
  int param_dtype;
  float* running_mean;
  float* running_var;
  float momentum, eps;
  float* save_mean;
  float* save_invstd;
  float* coef;
  void* dgamma;
  void* dbeta;
  const float* invstd_in;     // backward: the forward's invstd
};

// 64 channels x 16 slices per CTA. Each thread folds every 16th partial of its channel with compensated (Kahan)
// fp32 sums — double-precision adds and float->double conversions run at a small fraction of the fp32 rate on
// B200 and made a first version of this 4-CTA kernel take longer (36 us) than the statistics pass it finishes —
// then the 16 slice results are combined in slice order. Only the last handful of operations per channel
// (mean, E[x^2] - mean^2) use double. The order is a function of nparts only: bitwise reproducible.
constexpr int kFoldSlices = 16;

template <bool BWD>
__global__ void __launch_bounds__(64 * kFoldSlices) bn_fold_kernel(const FoldArgs a) {
  __shared__ float sh[2][kFoldSlices][64];
  const int chl = threadIdx.x % 64, slice = threadIdx.x / 64;
  const int ch = blockIdx.x * 64 + chl;
  float s = 0.f, cs = 0.f, q = 0.f, cq = 0.f;
  if (ch < a.C) {
    const float2* p = reinterpret_cast<const float2*>(a.partials) + ch;
#pragma unroll 4
    for (int part = slice; part < a.nparts; part += kFoldSlices) {
      const float2 v = __ldg(p + (size_t)part * a.C);
      float y = v.x - cs, t = s + y;
      cs = (t - s) - y; s = t;
      y = v.y - cq; t = q + y;
      cq = (t - q) - y; q = t;
    }
  }
  sh[0][slice][chl] = s;
  sh[1][slice][chl] = q;
  __syncthreads();
  if (slice != 0 || ch >= a.C) return;
  double S = 0.0, Q = 0.0;
#pragma unroll
  for (int i = 0; i < kFoldSlices; ++i) { S += (double)sh[0][i][chl]; Q += (double)sh[1][i][chl]; }
  if constexpr (!BWD) {
    const double mean = S * a.inv_count;
    double var = Q * a.inv_count - mean * mean;
    if (var < 0.0) var = 0.0;
    const float invstd = rsqrtf((float)var + a.eps);
    a.save_mean[ch] = (float)mean;
    a.save_invstd[ch] = invstd;
    const float scale = load_param(a.gamma, a.param_dtype, ch) * invstd;
    a.coef[ch] = scale;
    a.coef[a.C + ch] = load_param(a.beta, a.param_dtype, ch) - (float)mean * scale;
    if (a.running_mean) {
      const double unbiased = var * a.bessel;
      a.running_mean[ch] = (1.f - a.momentum) * a.running_mean[ch] + a.momentum * (float)mean;
      a.running_var[ch] = (1.f - a.momentum) * a.running_var[ch] + a.momentum * (float)unbiased;
    }
  } else {
    store_param(a.dbeta, a.param_dtype, ch, (float)S);
    store_param(a.dgamma, a.param_dtype, ch, (float)Q);
    a.coef[ch] = load_param(a.gamma, a.param_dtype, ch) * a.invstd_in[ch];
    a.coef[a.C + ch] = (float)(S * a.inv_count);
    a.coef[2 * a.C + ch] = (float)(Q * a.inv_count);
  }
}

// Two-level variant for the hot path (53 + 53 launches per ResNet-50 step): the single-level kernel above walks
// 444-592 partials per channel with 16 slices, i.e. ~37 dependent global-load round trips per thread, and took
// 9-15 us per call on B200 — latency, not bandwidth. Here grid.y CTAs each fold a 128-part range of one 64-channel
// group (every thread issues its 8 loads at once), write one combined row to `tmp`, and the CTA that arrives last
// at the group's counter folds the grid.y rows in index order and finishes the channel. The order is a function
// of nparts only, whichever CTA arrives last: bitwise reproducible.
constexpr int kFoldRange = kFoldSlices * 8;      // partials per CTA

template <bool BWD>
__global__ void __launch_bounds__(64 * kFoldSlices) bn_fold2_kernel(const FoldArgs a, float* __restrict__ tmp,
                                                                    unsigned* __restrict__ counters) {
  __shared__ float sh[2][kFoldSlices][64];
  __shared__ int s_last;
  const int chl = threadIdx.x % 64, slice = threadIdx.x / 64;
  const int ch = blockIdx.x * 64 + chl;
  const int p0 = blockIdx.y * kFoldRange;
  float s = 0.f, q = 0.f;
  if (ch < a.C) {
    const float2* p = reinterpret_cast<const float2*>(a.partials) + ch;
    float2 v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int part = p0 + slice + u * kFoldSlices;
      v[u] = part < a.nparts ? __ldg(p + (size_t)part * a.C) : make_float2(0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) { s += v[u].x; q += v[u].y; }
  }
  sh[0][slice][chl] = s;
  sh[1][slice][chl] = q;
  __syncthreads();
  if (slice == 0 && ch < a.C) {
    float S = 0.f, Q = 0.f;
#pragma unroll
    for (int i = 0; i < kFoldSlices; ++i) { S += sh[0][i][chl]; Q += sh[1][i][chl]; }
    reinterpret_cast<float2*>(tmp)[(size_t)blockIdx.y * a.C + ch] = make_float2(S, Q);
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned prev = atomicAdd(counters + blockIdx.x, 1u);
    s_last = (prev == gridDim.y - 1);
    if (s_last) counters[blockIdx.x] = 0;          // self-resetting: ready for the next launch on the stream
  }
  __syncthreads();
  if (!s_last || slice != 0 || ch >= a.C) return;
  __threadfence();
  double S = 0.0, Q = 0.0;
  for (unsigned g = 0; g < gridDim.y; ++g) {
    const float2 v = __ldcg(reinterpret_cast<const float2*>(tmp) + (size_t)g * a.C + ch);
    S += (double)v.x;
    Q += (double)v.y;
  }
  if constexpr (!BWD) {
    const double mean = S * a.inv_count;
    double var = Q * a.inv_count - mean * mean;
    if (var < 0.0) var = 0.0;
    const float invstd = rsqrtf((float)var + a.eps);
    a.save_mean[ch] = (float)mean;
    a.save_invstd[ch] = invstd;
    const float scale = load_param(a.gamma, a.param_dtype, ch) * invstd;
    a.coef[ch] = scale;
    a.coef[a.C + ch] = load_param(a.beta, a.param_dtype, ch) - (float)mean * scale;
    if (a.running_mean) {
      const double unbiased = var * a.bessel;
      a.running_mean[ch] = (1.f - a.momentum) * a.running_mean[ch] + a.momentum * (float)mean;
      a.running_var[ch] = (1.f - a.momentum) * a.running_var[ch] + a.momentum * (float)unbiased;
    }
  } else {
    store_param(a.dbeta, a.param_dtype, ch, (float)S);
    store_param(a.dgamma, a.param_dtype, ch, (float)Q);
    a.coef[ch] = load_param(a.gamma, a.param_dtype, ch) * a.invstd_in[ch];
    a.coef[a.C + ch] = (float)(S * a.inv_count);
    a.coef[2 * a.C + ch] = (float)(Q * a.inv_count);
  }
}

template <bool BWD>
int launch_fold(const FoldArgs& f, float* tmp, unsigned* counters, cudaStream_t stream) {
  const dim3 grid((f.C + 63) / 64, (f.nparts + kFoldRange - 1) / kFoldRange);
  bn_fold2_kernel<BWD><<<grid, 64 * kFoldSlices, 0, stream>>>(f, tmp, counters);
  return launched();
}

// ---- host -------------------------------------------------------------------------------------------------
constexpr int kPiece1 = 1024;      // 16 KB per stage (one stream)
constexpr int kPiece2 = 512;       // 8 KB per tensor per stage (two or three streams)

template <int NIN, int PIECE_VECS, bool MASKBITS = false>
constexpr int stream_smem() {
  // the ring, its barriers, and room for the reduction's [2][256][8] floats (reuses the ring)
  constexpr int ring = kStages * (NIN * PIECE_VECS * 16 + (MASKBITS ? PIECE_VECS : 0));
  return (ring > 2 * 256 * 8 * 4 ? ring : 2 * 256 * 8 * 4) + 128;
}

// Row ranges: one wave of CTAs (`per_sm` per SM), each a whole number of pieces.
void plan_stream(StreamArgs& a, size_t rows, int C, int piece_vecs, int per_sm, int* grid) {
  a.C = C;
  a.vecC = C / 8;
  a.total_vecs = rows * (size_t)a.vecC;
  const size_t pieces = (a.total_vecs + piece_vecs - 1) / piece_vecs;
  size_t ctas = (size_t)device_sm_count() * per_sm;
  if (ctas > pieces) ctas = pieces;
  const size_t pieces_per_cta = (pieces + ctas - 1) / ctas;
  a.vecs_per_cta = pieces_per_cta * piece_vecs;
  *grid = (int)((pieces + pieces_per_cta - 1) / pieces_per_cta);
}

template <typename Kern>
int launch_stream(Kern kern, DeviceOnce& once, int smem, int grid, const StreamArgs& a, cudaStream_t stream) {
  if (int rc = ensure_dynamic_smem(once, kern, smem)) return rc;
  kern<<<grid, kBlock, smem, stream>>>(a);
  return launched();
}

}  // namespace

namespace tb200 {

bool bn_stream_eligible(size_t rows, int C) {
  if (C < 8 || C > 2048 || (C & (C - 1))) return false;      // vector column must be fixed per thread
  return rows * (size_t)C >= (size_t)1 << 16;                  // tiny tensors: the old kernels are fine
}

int bn_stream_max_parts() { return device_sm_count() * 3; }

// scratch layout (floats): [0, parts*C*2) partials | then coef (4*C)
int bn_stream_fwd(const void* x, const void* residual, void* y, const void* gamma, const void* beta, int param_dtype,
                  float* running_mean, float* running_var, float momentum, float eps, float* save_mean,
                  float* save_invstd, float* partials, float* coef, const float* conv_partials, int conv_nparts,
                  size_t rows, int C, int relu, cudaStream_t stream, float* fold_tmp, unsigned* counters,
                  uint8_t* mask_out) {
  FoldArgs f{};
  f.C = C; f.count = (double)rows; f.gamma = gamma; f.beta = beta; f.param_dtype = param_dtype;
  f.inv_count = 1.0 / f.count; f.bessel = f.count > 1.0 ? f.count / (f.count - 1.0) : 1.0;
  f.running_mean = running_mean; f.running_var = running_var; f.momentum = momentum; f.eps = eps;
  f.save_mean = save_mean; f.save_invstd = save_invstd; f.coef = coef;
  if (conv_partials) {
    // the producing convolution's epilogue already summed the activation (csrc/conv3x3.cu)
    f.partials = conv_partials;
    f.nparts = conv_nparts;
  } else {
    StreamArgs a{};
    int grid = 0;
    plan_stream(a, rows, C, kPiece1, 3, &grid);
    a.in[0] = reinterpret_cast<const uint8_t*>(x);
    a.nin = 1;
    a.partials = partials;
    static DeviceOnce once;
    if (int rc = launch_stream(bn_stats_stream_kernel<kPiece1>, once, stream_smem<1, kPiece1>(), grid, a, stream)) return rc;
    f.partials = partials;
    f.nparts = grid;
  }
  if (int rc = launch_fold<false>(f, fold_tmp, counters, stream)) return rc;
  StreamArgs ap{};
  int grid = 0;
  ap.in[0] = reinterpret_cast<const uint8_t*>(x);
  ap.in[1] = reinterpret_cast<const uint8_t*>(residual);
  ap.out[0] = reinterpret_cast<uint4*>(y);
  ap.coef = coef;
  ap.relu = relu;
  ap.mask_out = mask_out;
  if (residual) {
    plan_stream(ap, rows, C, kPiece2, 3, &grid);
    if (mask_out && relu) {
      static DeviceOnce once;
      return launch_stream(bn_apply_stream_kernel<2, kPiece2, true>, once, stream_smem<2, kPiece2>(), grid, ap, stream);
    }
    static DeviceOnce once;
    return launch_stream(bn_apply_stream_kernel<2, kPiece2>, once, stream_smem<2, kPiece2>(), grid, ap, stream);
  }
  plan_stream(ap, rows, C, kPiece1, 3, &grid);
  static DeviceOnce once;
  return launch_stream(bn_apply_stream_kernel<1, kPiece1>, once, stream_smem<1, kPiece1>(), grid, ap, stream);
}

int bn_stream_apply(const void* x, const void* residual, void* y, const float* scale_shift, size_t rows, int C, int relu,
                    cudaStream_t stream) {
  StreamArgs ap{};
  int grid = 0;
  ap.in[0] = reinterpret_cast<const uint8_t*>(x);
  ap.in[1] = reinterpret_cast<const uint8_t*>(residual);
  ap.out[0] = reinterpret_cast<uint4*>(y);
  ap.coef = scale_shift;
  ap.relu = relu;
  if (residual) {
    plan_stream(ap, rows, C, kPiece2, 3, &grid);
    static DeviceOnce once;
    return launch_stream(bn_apply_stream_kernel<2, kPiece2>, once, stream_smem<2, kPiece2>(), grid, ap, stream);
  }
  plan_stream(ap, rows, C, kPiece1, 3, &grid);
  static DeviceOnce once;
  return launch_stream(bn_apply_stream_kernel<1, kPiece1>, once, stream_smem<1, kPiece1>(), grid, ap, stream);
}

int bn_stream_bwd(const void* dy, const void* x, const void* y, const void* gamma, const void* beta, int param_dtype,
                  const float* save_mean, const float* save_invstd, void* dx, void* dres, void* dgamma, void* dbeta,
                  float* partials, float* coef, size_t rows, int C, int relu, cudaStream_t stream, float* fold_tmp,
                  unsigned* counters, const uint8_t* mask_in) {
  const int mask = !relu ? 0 : (mask_in ? 3 : (y ? 1 : 2));
  StreamArgs a{};
  int grid = 0;
  plan_stream(a, rows, C, kPiece2, 2, &grid);
  a.in[0] = reinterpret_cast<const uint8_t*>(dy);
  a.in[1] = reinterpret_cast<const uint8_t*>(x);
  a.in[2] = reinterpret_cast<const uint8_t*>(y);
  a.save_mean = save_mean; a.save_invstd = save_invstd;
  a.gamma = gamma; a.beta = beta; a.param_dtype = param_dtype;
  a.partials = partials;
  a.coef = coef;
  a.relu = relu;
  a.mask_in = mask_in;
  int rc;
  if (mask == 3) {
    static DeviceOnce once;
    rc = launch_stream(bn_bwd_reduce_stream_kernel<3, kPiece2>, once, stream_smem<2, kPiece2, true>(), grid, a, stream);
  } else if (mask == 0) {
    static DeviceOnce once;
    rc = launch_stream(bn_bwd_reduce_stream_kernel<0, kPiece2>, once, stream_smem<2, kPiece2>(), grid, a, stream);
  } else if (mask == 1) {
    static DeviceOnce once;
    rc = launch_stream(bn_bwd_reduce_stream_kernel<1, kPiece2>, once, stream_smem<3, kPiece2>(), grid, a, stream);
  } else {
    static DeviceOnce once;
    rc = launch_stream(bn_bwd_reduce_stream_kernel<2, kPiece2>, once, stream_smem<2, kPiece2>(), grid, a, stream);
  }
  if (rc) return rc;
  FoldArgs f{};
  f.partials = partials; f.nparts = grid; f.C = C; f.count = (double)rows;
  f.inv_count = 1.0 / f.count; f.bessel = f.count > 1.0 ? f.count / (f.count - 1.0) : 1.0;
  f.gamma = gamma; f.beta = beta; f.param_dtype = param_dtype; f.coef = coef;
  f.dgamma = dgamma; f.dbeta = dbeta; f.invstd_in = save_invstd;
  if (int rc2 = launch_fold<true>(f, fold_tmp, counters, stream)) return rc2;
  a.out[0] = reinterpret_cast<uint4*>(dx);
  a.out[1] = reinterpret_cast<uint4*>(dres);
#define TB_BWD_APPLY(M, D, NINV)                                                                              \
  {                                                                                                           \
    static DeviceOnce once;                                                                                   \
    return launch_stream(bn_bwd_apply_stream_kernel<M, D, kPiece2>, once, stream_smem<NINV, kPiece2>(), grid, a, \
                         stream);                                                                             \
  }
  if (mask == 3) {
    if (dres) {
      static DeviceOnce once;
      return launch_stream(bn_bwd_apply_stream_kernel<3, true, kPiece2>, once, stream_smem<2, kPiece2, true>(), grid, a, stream);
    }
    static DeviceOnce once;
    return launch_stream(bn_bwd_apply_stream_kernel<3, false, kPiece2>, once, stream_smem<2, kPiece2, true>(), grid, a, stream);
  }
  if (mask == 0) { if (dres) TB_BWD_APPLY(0, true, 2) else TB_BWD_APPLY(0, false, 2) }
  if (mask == 1) { if (dres) TB_BWD_APPLY(1, true, 3) else TB_BWD_APPLY(1, false, 3) }
  if (dres) TB_BWD_APPLY(2, true, 2) else TB_BWD_APPLY(2, false, 2)
#undef TB_BWD_APPLY
}

}  // namespace tb200

extern "C" int tb200_bn_finalize(const float* partials, int nparts, double count, int C, const void* gamma,
                                 const void* beta, int param_dtype, float* running_mean, float* running_var,
                                 float momentum, float eps, float* save_mean, float* save_invstd, float* scale_shift,
                                 cudaStream_t stream) {
  if (!partials || nparts <= 0 || C <= 0 || count <= 0) return fail("tb200_bn_finalize: empty input%s");
  FoldArgs f{};
  f.partials = partials; f.nparts = nparts; f.C = C; f.count = count;
  f.inv_count = 1.0 / count; f.bessel = count > 1.0 ? count / (count - 1.0) : 1.0;
  f.gamma = gamma; f.beta = beta; f.param_dtype = param_dtype;
  f.running_mean = running_mean; f.running_var = running_var; f.momentum = momentum; f.eps = eps;
  f.save_mean = save_mean; f.save_invstd = save_invstd; f.coef = scale_shift;
  bn_fold_kernel<false><<<(C + 63) / 64, 64 * kFoldSlices, 0, stream>>>(f);
  return launched();
}
