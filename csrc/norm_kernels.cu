// LayerNorm over the last dimension of bf16 rows [R, D], forward and backward (K5: the Transformer / BERT job
// models of /root/reference model/model_factory.py:19-23 normalise every block input; round 1 left this to torch).
//
//   y  = (x - mean) * rstd * gamma + beta                      mean / rstd per row, fp32, saved for the backward
//   dx = rstd * (g - mean_D(g) - xhat * mean_D(g * xhat))      g = dy * gamma, xhat = (x - mean) * rstd
//   dgamma = sum_rows dy * xhat,   dbeta = sum_rows dy
//
// One warp per row, the whole row in registers (D = 256 * VPL, VPL 16-byte vectors per lane: 256 / 512 / 768 / 1024),
// so every tensor is read once and written once. dgamma / dbeta: each warp keeps its lanes' columns in
// registers over its rows, the CTA's warps are summed through shared memory in warp order into one partial row
// per CTA, and a second kernel sums the CTA rows in CTA order: deterministic, no atomics.
#include "common.cuh"
#include "../include/tiresias_b200.h"

using namespace tb200;

namespace {

constexpr int kLnWarps = 8;
constexpr int kLnThreads = kLnWarps * 32;

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
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float load_param(const void* p, int dtype, int i) {
  return dtype == TB200_F32 ? reinterpret_cast<const float*>(p)[i]
                            : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[i]);
}

// lane l owns vectors l, l + 32, ...: columns (l + 32 j) * 8 .. + 7
template <int VPL>
__global__ void __launch_bounds__(kLnThreads) layernorm_fwd_kernel(const uint4* __restrict__ x, uint4* __restrict__ y,
                                                                   const void* __restrict__ gamma,
                                                                   const void* __restrict__ beta, int param_dtype,
                                                                   float* __restrict__ mean_out,
                                                                   float* __restrict__ rstd_out, size_t rows, float eps) {
  constexpr int D = VPL * 256;
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  float g[VPL][8], b[VPL][8];
#pragma unroll
  for (int j = 0; j < VPL; ++j)
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int c = (lane + 32 * j) * 8 + k;
      g[j][k] = load_param(gamma, param_dtype, c);
      b[j][k] = load_param(beta, param_dtype, c);
    }
  for (size_t r = (size_t)blockIdx.x * kLnWarps + warp; r < rows; r += (size_t)gridDim.x * kLnWarps) {
    float v[VPL][8];
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < VPL; ++j) {
      unpack8(ld_stream_16(x + r * (D / 8) + lane + 32 * j), v[j]);
#pragma unroll
      for (int k = 0; k < 8; ++k) s += v[j][k];
    }
    const float mean = warp_sum(s) * (1.f / D);
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < VPL; ++j)
#pragma unroll
      for (int k = 0; k < 8; ++k) { const float d = v[j][k] - mean; q = fmaf(d, d, q); }
    const float rstd = rsqrtf(warp_sum(q) * (1.f / D) + eps);
#pragma unroll
    for (int j = 0; j < VPL; ++j) {
      float o[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) o[k] = fmaf((v[j][k] - mean) * rstd, g[j][k], b[j][k]);
      y[r * (D / 8) + lane + 32 * j] = pack8(o);
    }
    if (lane == 0) { mean_out[r] = mean; rstd_out[r] = rstd; }
  }
}

template <int VPL>
__global__ void __launch_bounds__(kLnThreads) layernorm_bwd_kernel(const uint4* __restrict__ dy, const uint4* __restrict__ x,
                                                                   const void* __restrict__ gamma, int param_dtype,
                                                                   const float* __restrict__ mean_in,
                                                                   const float* __restrict__ rstd_in,
                                                                   uint4* __restrict__ dx, float* __restrict__ partials,
                                                                   size_t rows) {
  constexpr int D = VPL * 256;
  __shared__ float red[kLnWarps][2][256];                  // one 256-column slice at a time
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  float g[VPL][8], dg[VPL][8], db[VPL][8];
#pragma unroll
  for (int j = 0; j < VPL; ++j)
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      g[j][k] = load_param(gamma, param_dtype, (lane + 32 * j) * 8 + k);
      dg[j][k] = db[j][k] = 0.f;
    }
  for (size_t r = (size_t)blockIdx.x * kLnWarps + warp; r < rows; r += (size_t)gridDim.x * kLnWarps) {
    const float mean = mean_in[r], rstd = rstd_in[r];
    float xh[VPL][8], gy[VPL][8];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int j = 0; j < VPL; ++j) {
      float dv[8];
      unpack8(ld_stream_16(x + r * (D / 8) + lane + 32 * j), xh[j]);
      unpack8(ld_stream_16(dy + r * (D / 8) + lane + 32 * j), dv);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        xh[j][k] = (xh[j][k] - mean) * rstd;
        dg[j][k] = fmaf(dv[k], xh[j][k], dg[j][k]);
        db[j][k] += dv[k];
        gy[j][k] = dv[k] * g[j][k];
        s1 += gy[j][k];
        s2 = fmaf(gy[j][k], xh[j][k], s2);
      }
    }
    const float m1 = warp_sum(s1) * (1.f / D), m2 = warp_sum(s2) * (1.f / D);
#pragma unroll
    for (int j = 0; j < VPL; ++j) {
      float o[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) o[k] = rstd * (gy[j][k] - m1 - xh[j][k] * m2);
      dx[r * (D / 8) + lane + 32 * j] = pack8(o);
    }
  }
  // CTA partial row [2][D]: warps summed in warp order
  float* out = partials + (size_t)blockIdx.x * 2 * D;
#pragma unroll
  for (int j = 0; j < VPL; ++j) {
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      red[warp][0][lane * 8 + k] = dg[j][k];
      red[warp][1][lane * 8 + k] = db[j][k];
    }
    __syncthreads();
    for (int e = threadIdx.x; e < 2 * 256; e += kLnThreads) {
      const int which = e / 256, c = e % 256;
      float s = 0.f;
#pragma unroll
      for (int w = 0; w < kLnWarps; ++w) s += red[w][which][c];
      // slice j holds vectors 32 j .. 32 j + 31 = columns 256 j .. 256 j + 255
      out[which * D + j * 256 + c] = s;
    }
  }
}

// dgamma / dbeta = the CTA rows summed in CTA order; written in the parameters' dtype
__global__ void __launch_bounds__(256) layernorm_fold_kernel(const float* __restrict__ partials, int nparts, int D,
                                                             void* __restrict__ dgamma, void* __restrict__ dbeta,
                                                             int param_dtype) {
  const int e = blockIdx.x * 256 + threadIdx.x;
  if (e >= 2 * D) return;
  float s = 0.f, comp = 0.f;
  for (int p = 0; p < nparts; ++p) {                      // Kahan: hundreds of partials of mixed magnitude
    const float v = partials[(size_t)p * 2 * D + e] - comp;
    const float t = s + v;
    comp = (t - s) - v;
    s = t;
  }
  void* dst = e < D ? dgamma : dbeta;
  const int c = e < D ? e : e - D;
  if (param_dtype == TB200_F32) reinterpret_cast<float*>(dst)[c] = s;
  else reinterpret_cast<__nv_bfloat16*>(dst)[c] = __float2bfloat16(s);
}

int ln_grid(size_t rows) {
  size_t want = (rows + kLnWarps - 1) / kLnWarps;
  const size_t cap = (size_t)device_sm_count() * 4;
  if (want > cap) want = cap;
  return (int)(want < 1 ? 1 : want);
}

int ln_check(const char* who, size_t rows, int D) {
  if (rows == 0 || D <= 0) return fail("%s: empty input", who);
  if (D != 256 && D != 512 && D != 768 && D != 1024) return fail("%s: D (%ld) must be 256, 512, 768 or 1024", who, D);
  return 0;
}

}  // namespace

extern "C" {

int tb200_layernorm_supported(int D) { return (D == 256 || D == 512 || D == 768 || D == 1024) ? 1 : 0; }

// floats of scratch the backward needs (CTA partial rows)
size_t tb200_layernorm_scratch_floats(int D) { return (size_t)device_sm_count() * 4 * 2 * (size_t)D; }

int tb200_layernorm_fwd(const void* x, void* y, const void* gamma, const void* beta, int param_dtype, float* mean,
                        float* rstd, size_t rows, int D, float eps, cudaStream_t stream) {
  if (int rc = ln_check("tb200_layernorm_fwd", rows, D)) return rc;
  const int grid = ln_grid(rows);
  auto X = reinterpret_cast<const uint4*>(x);
  auto Y = reinterpret_cast<uint4*>(y);
  if (D == 256) layernorm_fwd_kernel<1><<<grid, kLnThreads, 0, stream>>>(X, Y, gamma, beta, param_dtype, mean, rstd, rows, eps);
  else if (D == 512) layernorm_fwd_kernel<2><<<grid, kLnThreads, 0, stream>>>(X, Y, gamma, beta, param_dtype, mean, rstd, rows, eps);
  else if (D == 768) layernorm_fwd_kernel<3><<<grid, kLnThreads, 0, stream>>>(X, Y, gamma, beta, param_dtype, mean, rstd, rows, eps);
  else layernorm_fwd_kernel<4><<<grid, kLnThreads, 0, stream>>>(X, Y, gamma, beta, param_dtype, mean, rstd, rows, eps);
  return launched();
}

int tb200_layernorm_bwd(const void* dy, const void* x, const void* gamma, int param_dtype, const float* mean,
                        const float* rstd, void* dx, void* dgamma, void* dbeta, float* scratch, size_t scratch_floats,
                        size_t rows, int D, cudaStream_t stream) {
  if (int rc = ln_check("tb200_layernorm_bwd", rows, D)) return rc;
  const int grid = ln_grid(rows);
  if (!scratch || scratch_floats < (size_t)grid * 2 * D) return fail("tb200_layernorm_bwd: scratch too small%s");
  auto DY = reinterpret_cast<const uint4*>(dy);
  auto X = reinterpret_cast<const uint4*>(x);
  auto DX = reinterpret_cast<uint4*>(dx);
  if (D == 256) layernorm_bwd_kernel<1><<<grid, kLnThreads, 0, stream>>>(DY, X, gamma, param_dtype, mean, rstd, DX, scratch, rows);
  else if (D == 512) layernorm_bwd_kernel<2><<<grid, kLnThreads, 0, stream>>>(DY, X, gamma, param_dtype, mean, rstd, DX, scratch, rows);
  else if (D == 768) layernorm_bwd_kernel<3><<<grid, kLnThreads, 0, stream>>>(DY, X, gamma, param_dtype, mean, rstd, DX, scratch, rows);
  else layernorm_bwd_kernel<4><<<grid, kLnThreads, 0, stream>>>(DY, X, gamma, param_dtype, mean, rstd, DX, scratch, rows);
  if (int rc = launched()) return rc;
  layernorm_fold_kernel<<<(2 * D + 255) / 256, 256, 0, stream>>>(scratch, grid, D, dgamma, dbeta, param_dtype);
  return launched();
}

}  // extern "C"
