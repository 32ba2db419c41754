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

// Forward: a CTA walks output rows (n, oh); its threads cover the row's (ow, 8-channel vector) pairs, so all index
// arithmetic is 32-bit and per row (the first version decoded a 64-bit linear index per element: the divisions, not
// the memory system, bounded it at 43 % of the DRAM roofline). Records, per channel, which of the K*K window positions
// (kh*K + kw) held the first maximum: one byte per output element, the backward's only bookkeeping.
template <int K, int S, int P>
__global__ void __launch_bounds__(kPoolThreads) maxpool_fwd_kernel(const uint4* __restrict__ x,
                                                                   uint4* __restrict__ out,
                                                                   uint2* __restrict__ argmax, PoolShape s) {
  const int rows = s.N * s.OH, per_row = s.OW * s.vecC;
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const int n = row / s.OH, oh = row - n * s.OH;
    const uint4* xin = x + (size_t)n * s.H * s.W * s.vecC;
    for (int e = threadIdx.x; e < per_row; e += kPoolThreads) {
      const int ow = e / s.vecC, v = e - ow * s.vecC;
      float m[8];
      uint32_t pos[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) { m[k] = -INFINITY; pos[k] = 0; }
      // all window loads first (independent), then the comparisons
      uint4 win[K * K];
      bool ok[K * K];
#pragma unroll
      for (int kh = 0; kh < K; ++kh) {
        const int h = oh * S - P + kh;
#pragma unroll
        for (int kw = 0; kw < K; ++kw) {
          const int w = ow * S - P + kw;
          ok[kh * K + kw] = h >= 0 && h < s.H && w >= 0 && w < s.W;
          // (L1-allocating loads: the next output of this row re-reads the overlapping column)
          if (ok[kh * K + kw]) win[kh * K + kw] = __ldg(xin + ((size_t)h * s.W + w) * s.vecC + v);
        }
      }
#pragma unroll
      for (int t = 0; t < K * K; ++t) {
        if (!ok[t]) continue;
        float f[8];
        unpack8(win[t], f);
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (f[k] > m[k]) { m[k] = f[k]; pos[k] = t; }              // strict: the first maximum wins
      }
      const size_t o = (size_t)row * per_row + e;
      out[o] = pack8(m);
      if (argmax)
        argmax[o] = make_uint2(pos[0] | (pos[1] << 8) | (pos[2] << 16) | (pos[3] << 24),
                               pos[4] | (pos[5] << 8) | (pos[6] << 16) | (pos[7] << 24));
    }
  }
}

// Backward: a CTA walks INPUT rows (n, h); every input element gathers from the (at most ceil(K/S)^2) windows that
// contain it and takes a window's gradient iff the window's argmax byte names its position. No atomics,
// deterministic; the argmax words and gradient vectors of all candidate windows are requested before any is used.
template <int K, int S, int P>
__global__ void __launch_bounds__(kPoolThreads) maxpool_bwd_kernel(const uint2* __restrict__ argmax,
                                                                   const uint4* __restrict__ dout,
                                                                   uint4* __restrict__ dx, PoolShape s) {
  constexpr int kMaxWin = (K + S - 1) / S;                  // windows per axis that can contain one element
  const int rows = s.N * s.H, per_row = s.W * s.vecC;
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const int n = row / s.H, h = row - n * s.H;
    // windows (oh, .) with oh*S-P <= h <= oh*S-P+K-1: oh from ceil((h+P-K+1)/S) to floor((h+P)/S)
    const int oh1 = (h + P) / S;
    const int oh0 = h + P - K + 1 <= 0 ? 0 : (h + P - K + S) / S;
    const size_t obase = (size_t)n * s.OH * s.OW * s.vecC;
    for (int e = threadIdx.x; e < per_row; e += kPoolThreads) {
      const int w = e / s.vecC, v = e - w * s.vecC;
      const int ow1 = (w + P) / S;
      const int ow0 = w + P - K + 1 <= 0 ? 0 : (w + P - K + S) / S;
      uint2 am[kMaxWin * kMaxWin];
      uint4 go[kMaxWin * kMaxWin];
      uint32_t me[kMaxWin * kMaxWin];
      bool ok[kMaxWin * kMaxWin];
#pragma unroll
      for (int a = 0; a < kMaxWin; ++a) {
        const int oh = oh0 + a;
#pragma unroll
        for (int b = 0; b < kMaxWin; ++b) {
          const int ow = ow0 + b;
          const int t = a * kMaxWin + b;
          ok[t] = oh <= oh1 && oh < s.OH && ow <= ow1 && ow < s.OW;
          if (ok[t]) {
            const size_t o = obase + ((size_t)oh * s.OW + ow) * s.vecC + v;
            am[t] = __ldg(argmax + o);
            go[t] = __ldg(dout + o);
            me[t] = (uint32_t)(h - (oh * S - P)) * K + (uint32_t)(w - (ow * S - P));
          }
        }
      }
      float g[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) g[k] = 0.f;
#pragma unroll
      for (int t = 0; t < kMaxWin * kMaxWin; ++t) {
        if (!ok[t]) continue;
        const uint32_t hit_lo = am[t].x ^ (me[t] * 0x01010101u), hit_hi = am[t].y ^ (me[t] * 0x01010101u);
        float f[8];
        unpack8(go[t], f);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          if (((hit_lo >> (8 * k)) & 0xffu) == 0) g[k] += f[k];
          if (((hit_hi >> (8 * k)) & 0xffu) == 0) g[4 + k] += f[4 + k];
        }
      }
      dx[(size_t)row * per_row + e] = pack8(g);
    }
  }
}

// 3x3 / stride 2 / pad 1 backward, one thread per 2 x 2 INPUT patch (rows 2a, 2a+1; columns 2b, 2b+1) and channel
// vector: the four inputs lie in windows (a, a+1) x (b, b+1) only - even rows / columns in one of them, odd ones in
// two - so four (argmax, gradient) pairs serve four outputs (the generic kernel fetches up to four pairs per output).
__global__ void __launch_bounds__(kPoolThreads) maxpool321_bwd_patch_kernel(const uint2* __restrict__ argmax,
                                                                            const uint4* __restrict__ dout,
                                                                            uint4* __restrict__ dx, PoolShape s) {
  const int prow = (s.H + 1) / 2, pcol = (s.W + 1) / 2;
  const int rows = s.N * prow, per_row = pcol * s.vecC;
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const int n = row / prow, a = row - n * prow;
    const size_t obase = (size_t)n * s.OH * s.OW * s.vecC;
    const size_t ibase = (size_t)n * s.H * s.W * s.vecC;
    for (int e = threadIdx.x; e < per_row; e += kPoolThreads) {
      const int b = e / s.vecC, v = e - b * s.vecC;
      uint2 am[4];
      uint4 go[4];
      bool ok[4];
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const int oh = a + (t >> 1), ow = b + (t & 1);
        ok[t] = oh < s.OH && ow < s.OW;
        if (ok[t]) {
          const size_t o = obase + ((size_t)oh * s.OW + ow) * s.vecC + v;
          am[t] = __ldg(argmax + o);
          go[t] = __ldg(dout + o);
        }
      }
#pragma unroll
      for (int dy = 0; dy < 2; ++dy) {
        const int h = 2 * a + dy;
        if (h >= s.H) continue;
#pragma unroll
        for (int dxx = 0; dxx < 2; ++dxx) {
          const int w = 2 * b + dxx;
          if (w >= s.W) continue;
          float g[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) g[k] = 0.f;
#pragma unroll
          for (int t = 0; t < 4; ++t) {
            // window (a + t/2, b + t%2) covers rows 2*oh-1 .. 2*oh+1: row 2a+dy lies in window row a always and in
            // a+1 only when dy == 1 (likewise for columns)
            if ((t >> 1) > dy || (t & 1) > dxx || !ok[t]) continue;
            const int oh = a + (t >> 1), ow = b + (t & 1);
            const uint32_t me = (uint32_t)(h - (2 * oh - 1)) * 3u + (uint32_t)(w - (2 * ow - 1));
            const uint32_t hit_lo = am[t].x ^ (me * 0x01010101u), hit_hi = am[t].y ^ (me * 0x01010101u);
            float f[8];
            unpack8(go[t], f);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              if (((hit_lo >> (8 * k)) & 0xffu) == 0) g[k] += f[k];
              if (((hit_hi >> (8 * k)) & 0xffu) == 0) g[4 + k] += f[4 + k];
            }
          }
          dx[ibase + ((size_t)h * s.W + w) * s.vecC + v] = pack8(g);
        }
      }
    }
  }
}

int check(const char* who, int N, int H, int W, int C) {
  if (N <= 0 || H <= 0 || W <= 0 || C <= 0) return fail("%s: empty input", who);
  if (C % 8) return fail("%s: channels (%ld) must be a multiple of 8", who, C);
  return 0;
}

int grid_for_rows(size_t rows) {
  size_t want = rows;
  const size_t cap = (size_t)device_sm_count() * 16;
  if (want > cap) want = cap;
  if (want < 1) want = 1;
  return (int)want;
}

template <int K, int S, int P>
int pool_fwd(const void* x, void* out, void* argmax, int N, int H, int W, int C, cudaStream_t stream) {
  if (H + 2 * P < K || W + 2 * P < K) return fail("tb200_maxpool_fwd: input smaller than the window%s");
  PoolShape s{N, H, W, C, (H + 2 * P - K) / S + 1, (W + 2 * P - K) / S + 1, C / 8};
  maxpool_fwd_kernel<K, S, P><<<grid_for_rows((size_t)N * s.OH), kPoolThreads, 0, stream>>>(
      reinterpret_cast<const uint4*>(x), reinterpret_cast<uint4*>(out), reinterpret_cast<uint2*>(argmax), s);
  return launched();
}

template <int K, int S, int P>
int pool_bwd(const void* argmax, const void* dout, void* dx, int N, int H, int W, int C, cudaStream_t stream) {
  if (H + 2 * P < K || W + 2 * P < K) return fail("tb200_maxpool_bwd: input smaller than the window%s");
  PoolShape s{N, H, W, C, (H + 2 * P - K) / S + 1, (W + 2 * P - K) / S + 1, C / 8};
  maxpool_bwd_kernel<K, S, P><<<grid_for_rows((size_t)N * H), kPoolThreads, 0, stream>>>(
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
  if (k == 3 && stride == 2 && pad == 1) {
    if (H < 3 || W < 3) return pool_bwd<3, 2, 1>(argmax, dout, dx, N, H, W, C, stream);
    PoolShape s{N, H, W, C, (H + 2 - 3) / 2 + 1, (W + 2 - 3) / 2 + 1, C / 8};
    maxpool321_bwd_patch_kernel<<<grid_for_rows((size_t)N * ((H + 1) / 2)), kPoolThreads, 0, stream>>>(
        reinterpret_cast<const uint2*>(argmax), reinterpret_cast<const uint4*>(dout), reinterpret_cast<uint4*>(dx), s);
    return launched();
  }
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
