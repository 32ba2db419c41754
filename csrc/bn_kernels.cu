// Fused batch normalisation for NHWC bf16 activations (training forward + backward), with the
// ReLU and the residual add of a ResNet block folded in.
//
// Why this exists: one ResNet-50 step of the reference's headline job (model/model_factory.py:47
// "resnet50": 97) spends more than half of its device time in PyTorch's channels-last BatchNorm
// kernels plus the separate add / ReLU passes (profiles/step_launches_r1a.md): ~9.4 of 16.9 ms.
// All of that is pure HBM streaming, so it is restated here as four kernels that each touch the
// activation once:
//
//   forward   bn_stats   : per-channel mean / M2 over the N*H*W rows (fp32, Chan-combined in a
//                          fixed order => bitwise reproducible); the last CTA of a channel group
//                          folds gamma/beta into scale/shift and updates the running statistics
//             bn_apply   : y = relu(x*scale + shift + residual)            (x, residual -> y)
//   backward  bn_bwd_red : sum(g), sum(g*xhat) with g = dy * (y > 0)       (dy, x, y)
//             bn_bwd_app : dx = gamma*invstd*(g - mean(g) - xhat*mean(g*xhat)), dres = g
//
// Tile: a CTA is 8 x 32 threads; thread (tx, ty) owns the 8 channels [64*cg + 8*tx, +8) (one
// 16-byte vector) and rows ty, ty+32, ... of its row chunk. A row of a 64-channel group is one
// 128-byte line, so every warp reads four full lines per step.
#include "common.cuh"
#include "../include/tiresias_b200.h"

#include <stdlib.h>

using namespace tb200;

// csrc/bn_stream.cu: the TMA-streamed second generation (power-of-two channel counts up to 2048)
namespace tb200 {
bool bn_stream_eligible(size_t rows, int C);
int bn_stream_fwd(const void* x, const void* residual, void* y, const void* gamma, const void* beta, int param_dtype,
                  float* running_mean, float* running_var, float momentum, float eps, float* save_mean,
                  float* save_invstd, float* partials, float* coef, const float* conv_partials, int conv_nparts,
                  size_t rows, int C, int relu, cudaStream_t stream, float* fold_tmp, unsigned* counters,
                  uint8_t* mask_out);
int bn_stream_apply(const void* x, const void* residual, void* y, const float* scale_shift, size_t rows, int C, int relu,
                    cudaStream_t stream);
int bn_stream_bwd(const void* dy, const void* x, const void* y, const void* gamma, const void* beta, int param_dtype,
                  const float* save_mean, const float* save_invstd, void* dx, void* dres, void* dgamma, void* dbeta,
                  float* partials, float* coef, size_t rows, int C, int relu, cudaStream_t stream, float* fold_tmp,
                  unsigned* counters, const uint8_t* mask_in);
}  // namespace tb200

namespace {

constexpr int kTX = 8;                 // 16-byte vectors per channel group
constexpr int kTY = 32;                // row lanes
constexpr int kGroup = kTX * 8;        // 64 channels per CTA
constexpr int kCounterWords = 1024;
constexpr int kMaxPartialFloats = 2 * 64 * 1024;
constexpr int kMaxC = 4096;
constexpr int kCoefFloats = 4 * kMaxC;
constexpr int kStreamPartialFloats = 2 * 1024 * 1024;   // [<= 3 CTAs x SMs][C <= 2048][2] of the streamed kernels
static_assert(TB200_BN_SCRATCH_FLOATS == kCounterWords + kMaxPartialFloats + kCoefFloats + kStreamPartialFloats,
              "scratch layout");

// TB200_BN_STREAM=0 keeps every layer on the first-generation kernels (A/B measurements).
bool use_stream(size_t rows, int C) {
  static const bool enabled = [] {
    const char* e = getenv("TB200_BN_STREAM");
    return !(e && e[0] == '0');
  }();
  return enabled && bn_stream_eligible(rows, C) && (size_t)device_sm_count() * 3 * C * 2 <= (size_t)kStreamPartialFloats;
}

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

struct Chunking {
  int nchunks;
  int rows_per_chunk;
};

// Sum the per-thread 8-channel accumulators over the 32 row lanes of the CTA, in a fixed order.
// On return thread t < 64 holds the CTA total of channel t of the group for each of the NV values.
template <int NV>
__device__ __forceinline__ void block_reduce_channels(float (&acc)[NV][8], float (&out)[NV],
                                                      float (*smem)[kTY][kGroup]) {
  const int tx = threadIdx.x, ty = threadIdx.y;
#pragma unroll
  for (int v = 0; v < NV; ++v)
#pragma unroll
    for (int k = 0; k < 8; ++k) smem[v][ty][tx * 8 + k] = acc[v][k];
  __syncthreads();
  const int t = ty * kTX + tx;          // 0..255
  const int ch = t % kGroup, quarter = t / kGroup;   // 4 quarters of 8 row lanes
  float part[NV];
#pragma unroll
  for (int v = 0; v < NV; ++v) {
    float s = 0.f;
#pragma unroll
    for (int r = 0; r < kTY / 4; ++r) s += smem[v][quarter * (kTY / 4) + r][ch];
    part[v] = s;
  }
  __syncthreads();
#pragma unroll
  for (int v = 0; v < NV; ++v) smem[v][quarter][ch] = part[v];
  __syncthreads();
  if (t < kGroup) {
#pragma unroll
    for (int v = 0; v < NV; ++v) out[v] = (smem[v][0][t] + smem[v][1][t]) + (smem[v][2][t] + smem[v][3][t]);
  }
}

// True in exactly one CTA per channel group: the one that arrives last. Partials written before
// the call are visible to it (threadfence + atomic, the classic last-block reduction).
__device__ __forceinline__ bool last_cta_of_group(uint32_t* counter, int nchunks) {
  __shared__ int s_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0 && threadIdx.y == 0) {
    const unsigned prev = atomicAdd(counter, 1u);
    s_last = (prev == (unsigned)nchunks - 1);
    if (s_last) *counter = 0;          // self-resetting: ready for the next launch on the stream
  }
  __syncthreads();
  if (s_last) __threadfence();
  return s_last;
}

struct FwdArgs {
  const uint4* x;
  size_t rows;
  int C, vecC;                 // channels, channels / 8
  Chunking ck;
  const void* gamma;
  const void* beta;
  int param_dtype;
  float* running_mean;
  float* running_var;
  float momentum, eps;
  float* save_mean;
  float* save_invstd;
  uint32_t* counters;
  float* partials;             // [nchunks][C][2] = (mean, M2) of the chunk
  float* coef;                 // [0,C) scale, [C,2C) shift
};

__global__ void __launch_bounds__(kTX * kTY, 2) bn_stats_kernel(const FwdArgs a) {
  __shared__ float smem[2][kTY][kGroup];
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int cg = blockIdx.x, chunk = blockIdx.y;
  const int vcol = cg * kTX + tx;
  const bool col_ok = vcol < a.vecC;
  const size_t r0 = (size_t)chunk * a.ck.rows_per_chunk;
  const size_t r1 = min(a.rows, r0 + (size_t)a.ck.rows_per_chunk);
  float acc[2][8] = {};
  if (col_ok) {
    const uint4* p = a.x + vcol;
    size_t r = r0 + ty;
#define TB_ACC(V)                                                              \
  unpack8(V, f);                                                               \
  _Pragma("unroll") for (int k = 0; k < 8; ++k) { acc[0][k] += f[k]; acc[1][k] += f[k] * f[k]; }
    for (; r + 7 * kTY < r1; r += 8 * kTY) {      // eight independent 16-byte loads in flight per thread
      uint4 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = ld_stream_16(p + (r + u * kTY) * a.vecC);
      float f[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) { TB_ACC(v[u]) }
    }
    for (; r < r1; r += kTY) {
      uint4 v0 = ld_stream_16(p + r * a.vecC);
      float f[8];
      TB_ACC(v0)
    }
#undef TB_ACC
  }
  float tot[2];
  block_reduce_channels<2>(acc, tot, smem);
  const int t = ty * kTX + tx;
  const int ch = cg * kGroup + t;
  if (t < kGroup && ch < a.C) {
    const float n = (float)(r1 - r0);
    const float mean = n > 0 ? tot[0] / n : 0.f;
    float m2 = n > 0 ? tot[1] - tot[0] * mean : 0.f;
    float* out = a.partials + ((size_t)chunk * a.C + ch) * 2;
    out[0] = mean;
    out[1] = fmaxf(m2, 0.f);
  }
  if (!last_cta_of_group(a.counters + cg, a.ck.nchunks)) return;
  // Final combination by the last CTA: 4 lanes per channel each fold every 4th chunk (loads issued
  // 8 at a time), then the 4 lane results are folded in lane order. Chan et al.'s pairwise update;
  // the order is a function of nchunks only, so the result is bitwise reproducible.
  {
    const int fch = t % kGroup, lane = t / kGroup;
    const int gch = cg * kGroup + fch;
    float n = 0.f, mean = 0.f, m2 = 0.f;
    if (gch < a.C) {
      const int nc = a.ck.nchunks;
      for (int k0 = lane; k0 < nc; k0 += 4 * 8) {
        float mk[8], m2k[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int k = k0 + 4 * u;
          if (k < nc) {
            const float2 v = __ldcg(reinterpret_cast<const float2*>(a.partials + ((size_t)k * a.C + gch) * 2));
            mk[u] = v.x; m2k[u] = v.y;
          }
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int k = k0 + 4 * u;
          if (k < nc) {
            const size_t kr0 = (size_t)k * a.ck.rows_per_chunk;
            const float nk = (float)(min(a.rows, kr0 + (size_t)a.ck.rows_per_chunk) - kr0);
            const float delta = mk[u] - mean, ntot = n + nk;
            mean += delta * (nk / ntot);
            m2 += m2k[u] + delta * delta * (n * nk / ntot);
            n = ntot;
          }
        }
      }
    }
    __syncthreads();
    smem[0][lane][fch] = n; smem[0][4 + lane][fch] = mean; smem[0][8 + lane][fch] = m2;
    __syncthreads();
  }
  if (t < kGroup && ch < a.C) {
    float n = 0.f, mean = 0.f, m2 = 0.f;
#pragma unroll
    for (int l = 0; l < 4; ++l) {
      const float nk = smem[0][l][t], mk = smem[0][4 + l][t], m2k = smem[0][8 + l][t];
      if (nk > 0.f) {
        const float delta = mk - mean, ntot = n + nk;
        mean += delta * (nk / ntot);
        m2 += m2k + delta * delta * (n * nk / ntot);
        n = ntot;
      }
    }
    const float var = m2 / n;
    const float invstd = rsqrtf(var + a.eps);
    a.save_mean[ch] = mean;
    a.save_invstd[ch] = invstd;
    const float scale = load_param(a.gamma, a.param_dtype, ch) * invstd;
    a.coef[ch] = scale;
    a.coef[a.C + ch] = load_param(a.beta, a.param_dtype, ch) - mean * scale;
    if (a.running_mean) {
      const float unbiased = n > 1.f ? m2 / (n - 1.f) : var;
      a.running_mean[ch] = (1.f - a.momentum) * a.running_mean[ch] + a.momentum * mean;
      a.running_var[ch] = (1.f - a.momentum) * a.running_var[ch] + a.momentum * unbiased;
    }
  }
}

struct ApplyArgs {
  const uint4* x;
  const uint4* res;        // may be null
  uint4* y;
  const float* coef;       // scale | shift
  size_t rows;
  int C, vecC;
  Chunking ck;
  int relu;
};

__global__ void __launch_bounds__(kTX * kTY) bn_apply_kernel(const ApplyArgs a) {
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int vcol = blockIdx.x * kTX + tx;
  if (vcol >= a.vecC) return;
  float sc[8], sh[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    sc[k] = a.coef[vcol * 8 + k];
    sh[k] = a.coef[a.C + vcol * 8 + k];
  }
  const size_t r0 = (size_t)blockIdx.y * a.ck.rows_per_chunk;
  const size_t r1 = min(a.rows, r0 + (size_t)a.ck.rows_per_chunk);
  const bool has_res = a.res != nullptr;
  auto one = [&](const uint4& xr, const uint4& rr, size_t i) {
    float f[8], g[8];
    unpack8(xr, f);
    if (has_res) unpack8(rr, g);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float v = fmaf(f[k], sc[k], sh[k]);
      if (has_res) v += g[k];
      f[k] = a.relu ? fmaxf(v, 0.f) : v;
    }
    a.y[i] = pack8(f);       // default caching: the next layer reads y straight away
  };
  size_t r = r0 + ty;
  for (; r + 3 * kTY < r1; r += 4 * kTY) {      // 4 (or 8) independent 16-byte loads in flight
    uint4 xr[4], rr[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const size_t i = (r + u * kTY) * a.vecC + vcol;
      xr[u] = ld_stream_16(a.x + i);
      if (has_res) rr[u] = ld_stream_16(a.res + i);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) one(xr[u], rr[u], (r + u * kTY) * a.vecC + vcol);
  }
  for (; r < r1; r += kTY) {
    const size_t i = r * a.vecC + vcol;
    uint4 rr = make_uint4(0, 0, 0, 0);
    if (has_res) rr = ld_stream_16(a.res + i);
    one(ld_stream_16(a.x + i), rr, i);
  }
}

struct BwdArgs {
  const uint4* dy;
  const uint4* x;
  const uint4* y;          // output of the forward (for the ReLU mask); null when relu == 0
  uint4* dx;
  uint4* dres;             // may be null
  size_t rows;
  int C, vecC;
  Chunking ck;
  int relu;
  int recompute_mask;      // relu without a saved output: mask = (x*scale + shift > 0), exact for layers
                           // without a residual input (same fp32 fma as the forward)
  const void* gamma;
  const void* beta;
  int param_dtype;
  const float* save_mean;
  const float* save_invstd;
  void* dgamma;
  void* dbeta;
  uint32_t* counters;
  float* partials;         // [nchunks][C][2] = (sum g, sum g*xhat)
  float* coef;             // [0,C) A = gamma*invstd, [C,2C) mean(g), [2C,3C) mean(g*xhat)
};

// MASK: 0 = no ReLU, 1 = ReLU mask from the saved output y, 2 = ReLU mask recomputed from x.
template <int MASK>
__global__ void __launch_bounds__(kTX * kTY, 2) bn_bwd_reduce_kernel(const BwdArgs a) {
  __shared__ float smem[2][kTY][kGroup];
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int cg = blockIdx.x, chunk = blockIdx.y;
  const int vcol = cg * kTX + tx;
  const bool col_ok = vcol < a.vecC;
  const size_t r0 = (size_t)chunk * a.ck.rows_per_chunk;
  const size_t r1 = min(a.rows, r0 + (size_t)a.ck.rows_per_chunk);
  float acc[2][8] = {};
  if (col_ok) {
    float mean[8], istd[8], sc[8], sh[8];
    constexpr bool saved_mask = MASK == 1;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      mean[k] = a.save_mean[vcol * 8 + k];
      istd[k] = a.save_invstd[vcol * 8 + k];
      sc[k] = sh[k] = 0.f;
      if constexpr (MASK == 2) {
        sc[k] = load_param(a.gamma, a.param_dtype, vcol * 8 + k) * istd[k];
        sh[k] = load_param(a.beta, a.param_dtype, vcol * 8 + k) - mean[k] * sc[k];
      }
    }
    auto one = [&](const uint4& gr, const uint4& xr, const uint4& yr) {
      float g[8], xv[8], yv[8];
      unpack8(gr, g);
      unpack8(xr, xv);
      if (saved_mask) unpack8(yr, yv);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        if constexpr (MASK == 2) yv[k] = fmaf(xv[k], sc[k], sh[k]);
        const float gk = (MASK != 0 && !(yv[k] > 0.f)) ? 0.f : g[k];
        acc[0][k] += gk;
        acc[1][k] += gk * ((xv[k] - mean[k]) * istd[k]);
      }
    };
    size_t r = r0 + ty;
    for (; r + 3 * kTY < r1; r += 4 * kTY) {
      uint4 gr[4], xr[4], yr[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const size_t i = (r + u * kTY) * a.vecC + vcol;
        gr[u] = ld_stream_16(a.dy + i);
        xr[u] = ld_stream_16(a.x + i);
        if (saved_mask) yr[u] = ld_stream_16(a.y + i);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) one(gr[u], xr[u], yr[u]);
    }
    for (; r < r1; r += kTY) {
      const size_t i = r * a.vecC + vcol;
      uint4 yr = make_uint4(0, 0, 0, 0);
      if (saved_mask) yr = ld_stream_16(a.y + i);
      one(ld_stream_16(a.dy + i), ld_stream_16(a.x + i), yr);
    }
  }
  float tot[2];
  block_reduce_channels<2>(acc, tot, smem);
  const int t = ty * kTX + tx;
  const int ch = cg * kGroup + t;
  if (t < kGroup && ch < a.C) {
    float* out = a.partials + ((size_t)chunk * a.C + ch) * 2;
    out[0] = tot[0];
    out[1] = tot[1];
  }
  if (!last_cta_of_group(a.counters + cg, a.ck.nchunks)) return;
  {
    const int fch = t % kGroup, lane = t / kGroup;
    const int gch = cg * kGroup + fch;
    float s1 = 0.f, s2 = 0.f;
    if (gch < a.C) {
      const int nc = a.ck.nchunks;
      for (int k0 = lane; k0 < nc; k0 += 4 * 8) {
        float2 v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int k = k0 + 4 * u;
          v[u] = k < nc ? __ldcg(reinterpret_cast<const float2*>(a.partials + ((size_t)k * a.C + gch) * 2))
                        : make_float2(0.f, 0.f);
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) { s1 += v[u].x; s2 += v[u].y; }
      }
    }
    __syncthreads();
    smem[0][lane][fch] = s1; smem[0][4 + lane][fch] = s2;
    __syncthreads();
  }
  if (t < kGroup && ch < a.C) {
    const float s1 = (smem[0][0][t] + smem[0][1][t]) + (smem[0][2][t] + smem[0][3][t]);
    const float s2 = (smem[0][4][t] + smem[0][5][t]) + (smem[0][6][t] + smem[0][7][t]);
    store_param(a.dbeta, a.param_dtype, ch, s1);
    store_param(a.dgamma, a.param_dtype, ch, s2);
    const float inv_n = 1.f / (float)a.rows;
    a.coef[ch] = load_param(a.gamma, a.param_dtype, ch) * a.save_invstd[ch];
    a.coef[a.C + ch] = s1 * inv_n;
    a.coef[2 * a.C + ch] = s2 * inv_n;
  }
}

template <int MASK>
__global__ void __launch_bounds__(kTX * kTY, 2) bn_bwd_apply_kernel(const BwdArgs a) {
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int vcol = blockIdx.x * kTX + tx;
  if (vcol >= a.vecC) return;
  float A[8], mg[8], mgx[8], mean[8], istd[8], sh[8];
  constexpr bool saved_mask = MASK == 1;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int ch = vcol * 8 + k;
    A[k] = a.coef[ch];                       // gamma * invstd == the forward's scale
    mg[k] = a.coef[a.C + ch];
    mgx[k] = a.coef[2 * a.C + ch];
    mean[k] = a.save_mean[ch];
    istd[k] = a.save_invstd[ch];
    sh[k] = MASK == 2 ? load_param(a.beta, a.param_dtype, ch) - mean[k] * A[k] : 0.f;
  }
  const size_t r0 = (size_t)blockIdx.y * a.ck.rows_per_chunk;
  const size_t r1 = min(a.rows, r0 + (size_t)a.ck.rows_per_chunk);
  auto one = [&](const uint4& gr, const uint4& xr, const uint4& yr, size_t i) {
    float g[8], xv[8], yv[8];
    unpack8(gr, g);
    unpack8(xr, xv);
    if (saved_mask) unpack8(yr, yv);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if constexpr (MASK == 2) yv[k] = fmaf(xv[k], A[k], sh[k]);
      const float gk = (MASK != 0 && !(yv[k] > 0.f)) ? 0.f : g[k];
      g[k] = gk;
      xv[k] = A[k] * (gk - mg[k] - (xv[k] - mean[k]) * istd[k] * mgx[k]);
    }
    a.dx[i] = pack8(xv);
    if (a.dres) a.dres[i] = pack8(g);
  };
  size_t r = r0 + ty;
  for (; r + kTY < r1; r += 2 * kTY) {
    uint4 gr[2], xr[2], yr[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const size_t i = (r + u * kTY) * a.vecC + vcol;
      gr[u] = ld_stream_16(a.dy + i);
      xr[u] = ld_stream_16(a.x + i);
      if (saved_mask) yr[u] = ld_stream_16(a.y + i);
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) one(gr[u], xr[u], yr[u], (r + u * kTY) * a.vecC + vcol);
  }
  for (; r < r1; r += kTY) {
    const size_t i = r * a.vecC + vcol;
    uint4 yr = make_uint4(0, 0, 0, 0);
    if (saved_mask) yr = ld_stream_16(a.y + i);
    one(ld_stream_16(a.dy + i), ld_stream_16(a.x + i), yr, i);
  }
}

int sm_count_bn() { return device_sm_count(); }

// Row chunks so that the grid is ~`per_sm` CTAs per SM, every chunk holds >= one 32-row step, and the
// partials fit the scratch. Reductions use fewer, longer chunks (their last CTA folds one partial
// per chunk), the element-wise passes more.
Chunking plan_chunks(size_t rows, int C, int groups, int per_sm, int max_chunks) {
  int target = sm_count_bn() * per_sm;
  int n = (target + groups - 1) / groups;
  if (n > max_chunks) n = max_chunks;
  const size_t max_by_rows = (rows + kTY - 1) / kTY;
  if ((size_t)n > max_by_rows) n = (int)max_by_rows;
  const int max_by_scratch = kMaxPartialFloats / (2 * C);
  if (n > max_by_scratch) n = max_by_scratch;
  if (n < 1) n = 1;
  Chunking c;
  c.rows_per_chunk = (int)((rows + n - 1) / n);
  c.nchunks = (int)((rows + c.rows_per_chunk - 1) / c.rows_per_chunk);
  return c;
}

int check_shape(const char* who, size_t rows, int C) {
  if (rows == 0 || C <= 0) return fail("%s: empty input", who);
  if (C % 8) return fail("%s: channels (%ld) must be a multiple of 8", who, C);
  if (C > kMaxC) return fail("%s: channels (%ld) above the supported %ld", who, C, kMaxC);
  if (rows > (size_t)1 << 31) return fail("%s: too many rows", who);
  return 0;
}

}  // namespace

extern "C" {

size_t tb200_bn_scratch_floats(void) { return TB200_BN_SCRATCH_FLOATS; }

int tb200_bn_fwd_train(const void* x, const void* residual, void* y, const void* gamma,
                       const void* beta, int param_dtype, float* running_mean, float* running_var,
                       float momentum, float eps, float* save_mean, float* save_invstd,
                       float* scratch, size_t rows, int C, int relu, cudaStream_t stream) {
  if (int rc = check_shape("tb200_bn_fwd_train", rows, C)) return rc;
  if (use_stream(rows, C))
    return bn_stream_fwd(x, residual, y, gamma, beta, param_dtype, running_mean, running_var, momentum, eps, save_mean,
                         save_invstd, scratch + kCounterWords + kMaxPartialFloats + kCoefFloats,
                         scratch + kCounterWords + kMaxPartialFloats, nullptr, 0, rows, C, relu, stream,
                         scratch + kCounterWords, reinterpret_cast<unsigned*>(scratch), nullptr);
  const int groups = (C + kGroup - 1) / kGroup;
  FwdArgs f{};
  f.x = reinterpret_cast<const uint4*>(x);
  f.rows = rows; f.C = C; f.vecC = C / 8;
  f.ck = plan_chunks(rows, C, groups, 6, 296);
  f.gamma = gamma; f.beta = beta; f.param_dtype = param_dtype;
  f.running_mean = running_mean; f.running_var = running_var;
  f.momentum = momentum; f.eps = eps;
  f.save_mean = save_mean; f.save_invstd = save_invstd;
  f.counters = reinterpret_cast<uint32_t*>(scratch);
  f.partials = scratch + kCounterWords;
  f.coef = scratch + kCounterWords + kMaxPartialFloats;
  const dim3 block(kTX, kTY);
  bn_stats_kernel<<<dim3(groups, f.ck.nchunks), block, 0, stream>>>(f);
  if (int rc = launched()) return rc;
  ApplyArgs ap{};
  ap.x = f.x; ap.res = reinterpret_cast<const uint4*>(residual); ap.y = reinterpret_cast<uint4*>(y);
  ap.coef = f.coef; ap.rows = rows; ap.C = C; ap.vecC = C / 8; ap.relu = relu;
  ap.ck = plan_chunks(rows, C, groups, 8, 4096);
  bn_apply_kernel<<<dim3(groups, ap.ck.nchunks), block, 0, stream>>>(ap);
  return launched();
}

// Training forward whose batch statistics come from the producing convolution's epilogue (per-channel
// (sum, sum of squares) partials, csrc/conv3x3.cu / gemm_tcgen05.cu): fold + apply, no statistics pass.
int tb200_bn_fwd_train_stats(const void* x, const void* residual, void* y, const void* gamma, const void* beta,
                             int param_dtype, float* running_mean, float* running_var, float momentum, float eps,
                             float* save_mean, float* save_invstd, float* scratch, const float* conv_partials,
                             int conv_nparts, size_t rows, int C, int relu, cudaStream_t stream) {
  if (int rc = check_shape("tb200_bn_fwd_train_stats", rows, C)) return rc;
  if (!conv_partials || conv_nparts <= 0) return fail("tb200_bn_fwd_train_stats: no statistics partials%s");
  float* coef = scratch + kCounterWords + kMaxPartialFloats;
  if (use_stream(rows, C))
    return bn_stream_fwd(x, residual, y, gamma, beta, param_dtype, running_mean, running_var, momentum, eps, save_mean,
                         save_invstd, nullptr, coef, conv_partials, conv_nparts, rows, C, relu, stream,
                         scratch + kCounterWords, reinterpret_cast<unsigned*>(scratch), nullptr);
  if (int rc = tb200_bn_finalize(conv_partials, conv_nparts, (double)rows, C, gamma, beta, param_dtype, running_mean,
                                 running_var, momentum, eps, save_mean, save_invstd, coef, stream))
    return rc;
  return tb200_bn_apply(x, residual, y, coef, rows, C, relu, stream);
}

// 1 when the streamed kernels serve this shape (power-of-two C <= 2048, a tensor worth streaming): only they
// know the ReLU bit mask of tb200_bn_fwd_train_ex / tb200_bn_bwd_ex.
int tb200_bn_mask_supported(size_t rows, int C) { return use_stream(rows, C) ? 1 : 0; }

// fwd_train with everything optional: statistics from the producing convolution's epilogue (conv_partials, may
// be NULL) and, for a residual + ReLU layer, `relu_mask` (rows * C / 8 bytes + 16, may be NULL): one bit per output
// element, which lets the backward skip re-reading y.
int tb200_bn_fwd_train_ex(const void* x, const void* residual, void* y, const void* gamma, const void* beta,
                          int param_dtype, float* running_mean, float* running_var, float momentum, float eps,
                          float* save_mean, float* save_invstd, float* scratch, const float* conv_partials,
                          int conv_nparts, uint8_t* relu_mask, size_t rows, int C, int relu, cudaStream_t stream) {
  if (int rc = check_shape("tb200_bn_fwd_train_ex", rows, C)) return rc;
  if (relu_mask && !use_stream(rows, C)) return fail("tb200_bn_fwd_train_ex: the ReLU mask needs the streamed kernels%s");
  if (!use_stream(rows, C)) {
    if (conv_partials)
      return tb200_bn_fwd_train_stats(x, residual, y, gamma, beta, param_dtype, running_mean, running_var, momentum, eps,
                                      save_mean, save_invstd, scratch, conv_partials, conv_nparts, rows, C, relu, stream);
    return tb200_bn_fwd_train(x, residual, y, gamma, beta, param_dtype, running_mean, running_var, momentum, eps,
                              save_mean, save_invstd, scratch, rows, C, relu, stream);
  }
  return bn_stream_fwd(x, residual, y, gamma, beta, param_dtype, running_mean, running_var, momentum, eps, save_mean,
                       save_invstd, scratch + kCounterWords + kMaxPartialFloats + kCoefFloats,
                       scratch + kCounterWords + kMaxPartialFloats, conv_partials, conv_nparts, rows, C, relu, stream,
                       scratch + kCounterWords, reinterpret_cast<unsigned*>(scratch), relu_mask);
}

// bwd with the ReLU mask as bits (relu_mask, from tb200_bn_fwd_train_ex) instead of the saved output y.
int tb200_bn_bwd_ex(const void* dy, const void* x, const void* y, const uint8_t* relu_mask, const void* gamma,
                    const void* beta, int param_dtype, const float* save_mean, const float* save_invstd, void* dx,
                    void* dres, void* dgamma, void* dbeta, float* scratch, size_t rows, int C, int relu,
                    cudaStream_t stream) {
  if (!relu_mask)
    return tb200_bn_bwd(dy, x, y, gamma, beta, param_dtype, save_mean, save_invstd, dx, dres, dgamma, dbeta, scratch, rows,
                        C, relu, stream);
  if (int rc = check_shape("tb200_bn_bwd_ex", rows, C)) return rc;
  if (!use_stream(rows, C)) return fail("tb200_bn_bwd_ex: the ReLU mask needs the streamed kernels%s");
  return bn_stream_bwd(dy, x, nullptr, gamma, beta, param_dtype, save_mean, save_invstd, dx, dres, dgamma, dbeta,
                       scratch + kCounterWords + kMaxPartialFloats + kCoefFloats,
                       scratch + kCounterWords + kMaxPartialFloats, rows, C, relu, stream, scratch + kCounterWords,
                       reinterpret_cast<unsigned*>(scratch), relu_mask);
}

int tb200_bn_apply(const void* x, const void* residual, void* y, const float* scale_shift,
                   size_t rows, int C, int relu, cudaStream_t stream) {
  if (int rc = check_shape("tb200_bn_apply", rows, C)) return rc;
  if (use_stream(rows, C)) return bn_stream_apply(x, residual, y, scale_shift, rows, C, relu, stream);
  const int groups = (C + kGroup - 1) / kGroup;
  ApplyArgs ap{};
  ap.x = reinterpret_cast<const uint4*>(x); ap.res = reinterpret_cast<const uint4*>(residual);
  ap.y = reinterpret_cast<uint4*>(y); ap.coef = scale_shift; ap.rows = rows; ap.C = C; ap.vecC = C / 8;
  ap.ck = plan_chunks(rows, C, groups, 8, 4096); ap.relu = relu;
  bn_apply_kernel<<<dim3(groups, ap.ck.nchunks), dim3(kTX, kTY), 0, stream>>>(ap);
  return launched();
}

int tb200_bn_bwd(const void* dy, const void* x, const void* y, const void* gamma, const void* beta,
                 int param_dtype, const float* save_mean, const float* save_invstd, void* dx, void* dres,
                 void* dgamma, void* dbeta, float* scratch, size_t rows, int C, int relu,
                 cudaStream_t stream) {
  if (int rc = check_shape("tb200_bn_bwd", rows, C)) return rc;
  if (relu && !y && !beta) return fail("tb200_bn_bwd: relu needs the forward output, or beta to recompute it%s");
  if (relu && !y && dres) return fail("tb200_bn_bwd: a layer with a residual input must pass its forward output%s");
  if (use_stream(rows, C))
    return bn_stream_bwd(dy, x, y, gamma, beta, param_dtype, save_mean, save_invstd, dx, dres, dgamma, dbeta,
                         scratch + kCounterWords + kMaxPartialFloats + kCoefFloats,
                         scratch + kCounterWords + kMaxPartialFloats, rows, C, relu, stream, scratch + kCounterWords,
                         reinterpret_cast<unsigned*>(scratch), nullptr);
  const int groups = (C + kGroup - 1) / kGroup;
  BwdArgs b{};
  b.dy = reinterpret_cast<const uint4*>(dy); b.x = reinterpret_cast<const uint4*>(x);
  b.y = reinterpret_cast<const uint4*>(y); b.dx = reinterpret_cast<uint4*>(dx);
  b.dres = reinterpret_cast<uint4*>(dres);
  b.rows = rows; b.C = C; b.vecC = C / 8; b.ck = plan_chunks(rows, C, groups, 4, 296); b.relu = relu;
  b.gamma = gamma; b.beta = beta; b.param_dtype = param_dtype; b.save_mean = save_mean; b.save_invstd = save_invstd;
  b.recompute_mask = (relu && !y) ? 1 : 0;
  b.dgamma = dgamma; b.dbeta = dbeta;
  b.counters = reinterpret_cast<uint32_t*>(scratch);
  b.partials = scratch + kCounterWords;
  b.coef = scratch + kCounterWords + kMaxPartialFloats;
  const dim3 block(kTX, kTY);
  const int mask = !relu ? 0 : (b.recompute_mask ? 2 : 1);
  const dim3 rgrid(groups, b.ck.nchunks);
  if (mask == 0) bn_bwd_reduce_kernel<0><<<rgrid, block, 0, stream>>>(b);
  else if (mask == 1) bn_bwd_reduce_kernel<1><<<rgrid, block, 0, stream>>>(b);
  else bn_bwd_reduce_kernel<2><<<rgrid, block, 0, stream>>>(b);
  if (int rc = launched()) return rc;
  b.ck = plan_chunks(rows, C, groups, 8, 4096);
  const dim3 agrid(groups, b.ck.nchunks);
  if (mask == 0) bn_bwd_apply_kernel<0><<<agrid, block, 0, stream>>>(b);
  else if (mask == 1) bn_bwd_apply_kernel<1><<<agrid, block, 0, stream>>>(b);
  else bn_bwd_apply_kernel<2><<<agrid, block, 0, stream>>>(b);
  return launched();
}

}  // extern "C"
