// ResNet stem (7x7 / stride 2 / pad 3 convolution over 3 input channels) recast for the tensor cores.
//
// Three input channels cannot feed a 64-channel TMA box, and cuDNN's own 3-channel kernel is the slowest
// convolution of the step (1.5 ms at batch 256). Writing r = 2j + t (j = 0..3, t = 0..1) turns
//     y[oh, ow] = sum_{r,s,c} x[2 oh + r - 3, 2 ow + s - 3, c] * w[r, s, c]
// into a 4 x 1, stride-1 convolution over a packed tensor with exactly 64 "channels":
//     Y[n, q, ow, (t, s, c)] = x[n, 2q + t - 3, 2 ow + s - 3, c]        t < 2, s < 8, c < 4   (zero outside x,
//     y[oh, ow]              = sum_{j < 4} Y[oh + j, ow, :] . Wp[j, :]    for s == 7 and for c == 3)
//     Wp[k, j, (t, s, c)]    = w[k, 2j + t, s, c]                        (zero for 2j + t == 7, s == 7, c == 3)
// which is an ordinary tb200_conv2d_fprop / _wgrad problem (cin = 64, R = 4, S = 1, no padding) on the im2col
// mainloop, BatchNorm-statistics epilogue included. The pack kernel fuses the uint8 -> normalised bf16
// conversion that the input needs anyway. Y is 5.5x the size of the bf16 image (422 MB at batch 256) but is
// written once and read by two GEMMs; the 147-deep reduction becomes 256-deep (57 % useful MACs) on a tensor
// pipe that has cycles to spare at this layer.
//
// The reference only names the model (model/model_factory.py:47 "resnet50": 97); it has no kernels.
#include "common.cuh"
#include "../include/tiresias_b200.h"

using namespace tb200;

namespace {

// one thread per 16-byte unit of Y: unit u of a pixel = (t = u / 4, s = 2 * (u % 4) + {0, 1}) x 4 channels
__global__ void __launch_bounds__(256) stem_pack_kernel(const uint8_t* __restrict__ src, uint4* __restrict__ dst, int N,
                                                        int H, int W, int Q, int OW, float m0, float m1, float m2,
                                                        float s0, float s1, float s2) {
  const size_t total = (size_t)N * Q * OW * 8;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int u = (int)(i & 7);
    size_t pix = i >> 3;
    const int ow = (int)(pix % OW);
    pix /= OW;
    const int q = (int)(pix % Q), n = (int)(pix / Q);
    const int t = u >> 2, sa = (u & 3) * 2;
    const int row = 2 * q + t - 3;
    float f[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (row >= 0 && row < H) {
      const uint8_t* line = src + ((size_t)n * H + row) * (size_t)W * 3;
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const int s = sa + k, col = 2 * ow + s - 3;
        if (s < 7 && col >= 0 && col < W) {
          const uint8_t* px = line + (size_t)col * 3;
          f[4 * k + 0] = (px[0] * (1.0f / 255.0f) - m0) * s0;
          f[4 * k + 1] = (px[1] * (1.0f / 255.0f) - m1) * s1;
          f[4 * k + 2] = (px[2] * (1.0f / 255.0f) - m2) * s2;
        }
      }
    }
    uint4 o;
    o.x = pack_bf16(f[0], f[1]); o.y = pack_bf16(f[2], f[3]);
    o.z = pack_bf16(f[4], f[5]); o.w = pack_bf16(f[6], f[7]);
    dst[i] = o;
  }
}

// w physical [K][7][7][3] (channels_last [K,3,7,7]) -> Wp [K][4][64]
__global__ void stem_weight_pack_kernel(const __nv_bfloat16* __restrict__ w, __nv_bfloat16* __restrict__ wp, int K) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= K * 256) return;
  const int e = i & 63, j = (i >> 6) & 3, k = i >> 8;
  const int t = e >> 5, s = (e >> 2) & 7, c = e & 3, r = 2 * j + t;
  __nv_bfloat16 v = __float2bfloat16(0.f);
  if (r < 7 && s < 7 && c < 3) v = w[((k * 7 + r) * 7 + s) * 3 + c];
  wp[i] = v;
}

// dWp [K][4][64] (fp32) -> dw physical [K][7][7][3] in `dtype`
__global__ void stem_wgrad_unpack_kernel(const float* __restrict__ dwp, void* __restrict__ dw, int K, int dtype) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= K * 147) return;
  const int c = i % 3, s = (i / 3) % 7, r = (i / 21) % 7, k = i / 147;
  const int j = r >> 1, t = r & 1;
  const float v = dwp[(k * 4 + j) * 64 + t * 32 + s * 4 + c];
  if (dtype == kBF16) reinterpret_cast<__nv_bfloat16*>(dw)[i] = __float2bfloat16(v);
  else reinterpret_cast<float*>(dw)[i] = v;
}

}  // namespace

extern "C" {

int tb200_stem7x7_pack(const uint8_t* src, void* dst, int n, int h, int w, const float mean[3], const float inv_std[3],
                       cudaStream_t stream) {
  if (n <= 0 || h <= 0 || w <= 0) return fail("tb200_stem7x7_pack: empty input%s");
  const int oh = (h - 1) / 2 + 1, ow = (w - 1) / 2 + 1, q = oh + 3;
  const size_t units = (size_t)n * q * ow * 8;
  size_t blocks = (units + 255) / 256;
  const size_t cap = (size_t)device_sm_count() * 32;
  if (blocks > cap) blocks = cap;
  stem_pack_kernel<<<(unsigned)blocks, 256, 0, stream>>>(src, reinterpret_cast<uint4*>(dst), n, h, w, q, ow, mean[0], mean[1],
                                                        mean[2], inv_std[0], inv_std[1], inv_std[2]);
  return launched();
}

int tb200_stem7x7_weight_pack(const void* w, void* wp, int cout, cudaStream_t stream) {
  if (cout <= 0) return fail("tb200_stem7x7_weight_pack: empty weight%s");
  stem_weight_pack_kernel<<<(cout * 256 + 255) / 256, 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(w),
                                                                       reinterpret_cast<__nv_bfloat16*>(wp), cout);
  return launched();
}

int tb200_stem7x7_wgrad_unpack(const float* dwp, void* dw, int cout, int dw_dtype, cudaStream_t stream) {
  if (cout <= 0) return fail("tb200_stem7x7_wgrad_unpack: empty weight%s");
  stem_wgrad_unpack_kernel<<<(cout * 147 + 255) / 256, 256, 0, stream>>>(dwp, dw, cout, dw_dtype);
  return launched();
}

}  // extern "C"
