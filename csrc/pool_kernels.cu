// Max pooling for NHWC bf16, forward and backward: 3x3 / stride 2 / pad 1 (the ResNet stem), 2x2 / stride 2
// (VGG) and 3x3 / stride 2 / pad 0 (AlexNet, GoogLeNet, Inception) as instantiations of one pair of kernels.
//
// PyTorch's channels-last kernels for this op run far below memory bandwidth on B200 and the backward
// reads an int64 index per output element (step profile at batch 256: forward 0.75 ms, backward 1.73 ms
// for a 411 MB input). These kernels stream: the forward reads the input once, writes the pooled map and
// one BYTE per output element saying which of the nine window positions held the first maximum (torch's
// tie rule); the backward is input-centric — every input element looks at the (at most four) windows that
// contain it and takes a window's gradient iff that byte names its position — no atomics, deterministic.
#include "common.cuh"
#include "../include/tiresias_b200.h"

using namespace tb200;

namespace {

constexpr int kPoolThreads = 256;

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

struct PoolShape {
  int N, H, W, C, OH, OW, vecC;
};

// One thread per (n, oh, ow, 8 channels). Also records, per channel, which of the K*K window positions
// (kh*K + kw) held the first maximum: one byte per output element, the backward's only bookkeeping.
template <int K, int S, int P>
__global__ void __launch_bounds__(kPoolThreads) maxpool_fwd_kernel(const uint4* __restrict__ x,
                                                                   uint4* __restrict__ out,
                                                                   uint2* __restrict__ argmax, PoolShape s) {
  const size_t total = (size_t)s.N * s.OH * s.OW * s.vecC;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total;
       i += (size_t)gridDim.x * blockDim.x) {
    const int v = (int)(i % s.vecC);
    size_t t = i / s.vecC;
    const int ow = (int)(t % s.OW); t /= s.OW;
    const int oh = (int)(t % s.OH);
    const int n = (int)(t / s.OH);
    float m[8];
    uint32_t pos[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) { m[k] = -INFINITY; pos[k] = 0; }
#pragma unroll
    for (int kh = 0; kh < K; ++kh) {
      const int h = oh * S - P + kh;
      if (h < 0 || h >= s.H) continue;
#pragma unroll
      for (int kw = 0; kw < K; ++kw) {
        const int w = ow * S - P + kw;
        if (w < 0 || w >= s.W) continue;
        float f[8];
        unpack8(ld_stream_16(x + (((size_t)n * s.H + h) * s.W + w) * s.vecC + v), f);
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (f[k] > m[k]) { m[k] = f[k]; pos[k] = kh * K + kw; }     // strict: the first maximum wins
      }
    }
    out[i] = pack8(m);
    if (argmax)
      argmax[i] = make_uint2(pos[0] | (pos[1] << 8) | (pos[2] << 16) | (pos[3] << 24),
                             pos[4] | (pos[5] << 8) | (pos[6] << 16) | (pos[7] << 24));
  }
}

// One thread per (n, h, w, 8 channels) of the INPUT: gathers from the (at most ceil(K/S)^2) windows that
// contain the element. No atomics, deterministic; DRAM traffic = dout + argmax read once, dx written once.
template <int K, int S, int P>
__global__ void __launch_bounds__(kPoolThreads) maxpool_bwd_kernel(const uint2* __restrict__ argmax,
                                                                   const uint4* __restrict__ dout,
                                                                   uint4* __restrict__ dx, PoolShape s) {
  const size_t total = (size_t)s.N * s.H * s.W * s.vecC;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total;
       i += (size_t)gridDim.x * blockDim.x) {
    const int v = (int)(i % s.vecC);
    size_t t = i / s.vecC;
    const int w = (int)(t % s.W); t /= s.W;
    const int h = (int)(t % s.H);
    const int n = (int)(t / s.H);
    float g[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) g[k] = 0.f;
    // windows (oh, ow) with oh*S-P <= h <= oh*S-P+K-1: oh from ceil((h+P-K+1)/S) to floor((h+P)/S)
    const int oh1 = (h + P) / S, ow1 = (w + P) / S;
    const int oh0 = h + P - K + 1 <= 0 ? 0 : (h + P - K + S) / S;
    const int ow0 = w + P - K + 1 <= 0 ? 0 : (w + P - K + S) / S;
    for (int oh = oh0; oh <= oh1; ++oh) {
      if (oh >= s.OH) continue;
      const uint32_t kh = (uint32_t)(h - (oh * S - P));
      for (int ow = ow0; ow <= ow1; ++ow) {
        if (ow >= s.OW) continue;
        const uint32_t me = kh * K + (uint32_t)(w - (ow * S - P));
        const size_t o = (((size_t)n * s.OH + oh) * s.OW + ow) * s.vecC + v;
        const uint2 am = __ldg(argmax + o);
        const uint32_t hit_lo = am.x ^ (me * 0x01010101u), hit_hi = am.y ^ (me * 0x01010101u);
        if ((((hit_lo - 0x01010101u) & ~hit_lo) | ((hit_hi - 0x01010101u) & ~hit_hi)) & 0x80808080u) {
          float go[8];                      // at least one channel's argmax byte equals my position
          unpack8(__ldg(dout + o), go);
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            if (((hit_lo >> (8 * k)) & 0xffu) == 0) g[k] += go[k];
            if (((hit_hi >> (8 * k)) & 0xffu) == 0) g[4 + k] += go[4 + k];
          }
        }
      }
    }
    dx[i] = pack8(g);
  }
}

int check(const char* who, int N, int H, int W, int C) {
  if (N <= 0 || H <= 0 || W <= 0 || C <= 0) return fail("%s: empty input", who);
  if (C % 8) return fail("%s: channels (%ld) must be a multiple of 8", who, C);
  return 0;
}

int grid_for_elems(size_t n) {
  size_t want = (n + kPoolThreads - 1) / kPoolThreads;
  const size_t cap = 148 * 16;
  if (want > cap) want = cap;
  if (want < 1) want = 1;
  return (int)want;
}

template <int K, int S, int P>
int pool_fwd(const void* x, void* out, void* argmax, int N, int H, int W, int C, cudaStream_t stream) {
  if (H + 2 * P < K || W + 2 * P < K) return fail("tb200_maxpool_fwd: input smaller than the window%s");
  PoolShape s{N, H, W, C, (H + 2 * P - K) / S + 1, (W + 2 * P - K) / S + 1, C / 8};
  const size_t total = (size_t)N * s.OH * s.OW * s.vecC;
  maxpool_fwd_kernel<K, S, P><<<grid_for_elems(total), kPoolThreads, 0, stream>>>(
      reinterpret_cast<const uint4*>(x), reinterpret_cast<uint4*>(out), reinterpret_cast<uint2*>(argmax), s);
  return launched();
}

template <int K, int S, int P>
int pool_bwd(const void* argmax, const void* dout, void* dx, int N, int H, int W, int C, cudaStream_t stream) {
  if (H + 2 * P < K || W + 2 * P < K) return fail("tb200_maxpool_bwd: input smaller than the window%s");
  PoolShape s{N, H, W, C, (H + 2 * P - K) / S + 1, (W + 2 * P - K) / S + 1, C / 8};
  const size_t total = (size_t)N * H * W * s.vecC;
  maxpool_bwd_kernel<K, S, P><<<grid_for_elems(total), kPoolThreads, 0, stream>>>(
      reinterpret_cast<const uint2*>(argmax), reinterpret_cast<const uint4*>(dout), reinterpret_cast<uint4*>(dx), s);
  return launched();
}

}  // namespace

extern "C" {

// Geometries served: (k, stride, pad) = (3, 2, 1), (2, 2, 0), (3, 2, 0); floor mode (no ceil_mode).
int tb200_maxpool_supported(int k, int stride, int pad) {
  return (k == 3 && stride == 2 && (pad == 1 || pad == 0)) || (k == 2 && stride == 2 && pad == 0) ? 1 : 0;
}

int tb200_maxpool_fwd(const void* x, void* out, void* argmax, int N, int H, int W, int C, int k, int stride, int pad,
                      cudaStream_t stream) {
  if (int rc = check("tb200_maxpool_fwd", N, H, W, C)) return rc;
  if (k == 3 && stride == 2 && pad == 1) return pool_fwd<3, 2, 1>(x, out, argmax, N, H, W, C, stream);
  if (k == 3 && stride == 2 && pad == 0) return pool_fwd<3, 2, 0>(x, out, argmax, N, H, W, C, stream);
  if (k == 2 && stride == 2 && pad == 0) return pool_fwd<2, 2, 0>(x, out, argmax, N, H, W, C, stream);
  return fail("tb200_maxpool_fwd: geometry not served%s");
}

int tb200_maxpool_bwd(const void* argmax, const void* dout, void* dx, int N, int H, int W, int C, int k, int stride,
                      int pad, cudaStream_t stream) {
  if (int rc = check("tb200_maxpool_bwd", N, H, W, C)) return rc;
  if (!argmax) return fail("tb200_maxpool_bwd: needs the forward's argmax map%s");
  if (k == 3 && stride == 2 && pad == 1) return pool_bwd<3, 2, 1>(argmax, dout, dx, N, H, W, C, stream);
  if (k == 3 && stride == 2 && pad == 0) return pool_bwd<3, 2, 0>(argmax, dout, dx, N, H, W, C, stream);
  if (k == 2 && stride == 2 && pad == 0) return pool_bwd<2, 2, 0>(argmax, dout, dx, N, H, W, C, stream);
  return fail("tb200_maxpool_bwd: geometry not served%s");
}

int tb200_maxpool3x3s2_fwd(const void* x, void* out, void* argmax, int N, int H, int W, int C,
                           cudaStream_t stream) {
  return tb200_maxpool_fwd(x, out, argmax, N, H, W, C, 3, 2, 1, stream);
}

int tb200_maxpool3x3s2_bwd(const void* argmax, const void* dout, void* dx, int N, int H, int W, int C,
                           cudaStream_t stream) {
  return tb200_maxpool_bwd(argmax, dout, dx, N, H, W, C, 3, 2, 1, stream);
}

}  // extern "C"
