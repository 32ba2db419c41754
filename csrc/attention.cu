// Multi-head attention (head dimension 64) on tcgen05, forward and backward, for the Transformer / BERT job models
// (/root/reference model/model_factory.py:19-23 names them by checkpoint size; round 1 used torch SDPA).
//
// Everything is read in place from the fused projection output qkv [B*S, 3*D] (row = token, columns
// [which*D + head*64, +64)) through one 2-D tensor map, and the gradients are written in place into dqkv of the
// same layout: no head transposes, no concatenation. Tiles are 128 query rows x 128 keys.
//
//   forward  (one CTA per (query block, head, batch)), flash style over key blocks:
//       S = Q K^T              tcgen05, K-major A and B, fp32 in TMEM
//       P = exp2(S*c - m)      four softmax warps, one query row per thread (tcgen05.ld lane = row); running max /
//                              sum in registers; P is written as bf16 into a 128B-swizzled shared-memory tile, i.e.
//                              exactly what a TMA load would have produced, and becomes the A operand of
//       O += P V               V consumed as stored (MN-major B); the product is read back from TMEM and added to
//                              the thread's fp32 output row, which is rescaled whenever the running max moves
//       out: O / l (bf16, [B*S, D]) and lse2 = m + log2(l) per (batch, head, row)
//   backward, deterministic (no atomics): recomputes P from lse2
//       attn_dkv_kernel (one CTA per key block):   dV += P^T dO,  dK += dS^T Q     (P / dS tiles as MN-major A,
//                                                  dO / Q tiles - the very tiles that were K-major A operands of
//                                                  S = Q K^T and dP = dO V^T - as MN-major B)
//       attn_dq_kernel  (one CTA per query block): dQ += dS K                      (K tile as MN-major B)
//       with dS = P * (dP - delta) * scale, delta = rowsum(dO * O) from attn_delta_kernel.
//
// The reference has no native code; nothing here has a counterpart.
#include "common.cuh"
#include "tc_ptx.cuh"
#include "../include/tiresias_b200.h"

#include <cuda.h>
#include <mutex>

using namespace tb200;
using namespace tb200::ptx;

namespace {

constexpr int kThreads = 256;          // warp 0: TMA + MMA issue, warp 2: TMEM, warps 4-7: softmax / epilogue
constexpr int kTile = 16384;           // 128 rows x 128 B
constexpr float kLog2e = 1.4426950408889634f;

struct AttnParams {
  int B, S, H, D;                      // batch, sequence, heads, model width (= H * 64)
  int causal;
  float scale_log2;                    // softmax scale * log2(e)
  float scale;
  __nv_bfloat16* out;                  // forward: O [B*S, D]
  float* lse2;                         // [B, H, S]
  const float* delta;                  // backward: [B, H, S]
  __nv_bfloat16* dqkv;                 // backward: [B*S, 3*D]
};

// MN-major A operand of 128 rows made of two 64-wide slabs kTile bytes apart (the P / dS tiles read transposed)
constexpr uint64_t kDescMNMajorTiles =
    ((uint64_t)(kTile >> 4) << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);

__device__ __forceinline__ void tmem_alloc(uint32_t slot, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_free(uint32_t taddr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}

// bf16 row fragment -> swizzled tile: 32 consecutive columns [c0, c0 + 32) of row r of a [128 x 128] bf16 matrix
// stored as two K-major tiles (columns 0-63, 64-127), rows of 128 bytes, 16-byte units XOR-swizzled by (r & 7)
__device__ __forceinline__ void store_row_chunk(uint32_t tiles, int r, int c0, const float (&v)[32]) {
  const uint32_t tile = tiles + (uint32_t)(c0 >> 6) * kTile + (uint32_t)r * 128u;
  const uint32_t u0 = (uint32_t)(c0 & 63) >> 3, swz = (uint32_t)(r & 7);
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    uint4 o;
    o.x = pack_bf16(v[8 * u], v[8 * u + 1]);
    o.y = pack_bf16(v[8 * u + 2], v[8 * u + 3]);
    o.z = pack_bf16(v[8 * u + 4], v[8 * u + 5]);
    o.w = pack_bf16(v[8 * u + 6], v[8 * u + 7]);
    st_shared_16(tile + (((u0 + (uint32_t)u) ^ swz) << 4), o);
  }
}

// S = A B^T over the 64-wide head dimension: both tiles K-major [128 rows x 64]
__device__ __forceinline__ void mma_qk(uint32_t tmem_d, uint32_t a_tile, uint32_t b_tile) {
  constexpr uint32_t idesc = make_idesc_bf16(128, 128, 0, 0);
  const uint64_t a = kDescKMajor | (uint64_t)desc_units(a_tile), b = kDescKMajor | (uint64_t)desc_units(b_tile);
#pragma unroll
  for (int k = 0; k < 4; ++k) tc_mma_bf16(tmem_d, a + 2u * k, b + 2u * k, idesc, k ? 1u : 0u);
}
// D[128 x 64] (+)= A[128 x 128] B: A = two K-major tiles (reduction columns 0-63, 64-127), B = [128 reduction rows x
// 64] tile as stored (MN-major)
__device__ __forceinline__ void mma_pv(uint32_t tmem_d, uint32_t a_tiles, uint32_t b_tile, uint32_t accumulate) {
  constexpr uint32_t idesc = make_idesc_bf16(128, 64, 0, 1);
  const uint64_t b = kDescMNMajor | (uint64_t)desc_units(b_tile);
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const uint64_t a = kDescKMajor | (uint64_t)(desc_units(a_tiles + (uint32_t)(k >> 2) * kTile) + 2u * (k & 3));
    tc_mma_bf16(tmem_d, a, b + 128u * k, idesc, (k || accumulate) ? 1u : 0u);
  }
}
// D[128 x 64] (+)= A^T B: A^T read out of two K-major tiles [128 reduction rows x 128] (MN-major A, M = the tiles'
// columns), B = [128 reduction rows x 64] tile as stored (MN-major)
__device__ __forceinline__ void mma_tn(uint32_t tmem_d, uint32_t a_tiles, uint32_t b_tile, uint32_t accumulate) {
  constexpr uint32_t idesc = make_idesc_bf16(128, 64, 1, 1);
  const uint64_t a = kDescMNMajorTiles | (uint64_t)desc_units(a_tiles), b = kDescMNMajor | (uint64_t)desc_units(b_tile);
#pragma unroll
  for (int k = 0; k < 8; ++k) tc_mma_bf16(tmem_d, a + 128u * k, b + 128u * k, idesc, (k || accumulate) ? 1u : 0u);
}

// ---------------------------------------------------------------------------------------------------------
// forward
// shared memory: Q | K[2] | V[2] | P (two tiles) | barriers
constexpr int kFwdSmem = kTile * (1 + 2 + 2 + 2) + 256;

__global__ void __launch_bounds__(kThreads, 1) attn_fwd_kernel(const __grid_constant__ CUtensorMap map_qkv,
                                                               const AttnParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = smem_u32(smem_raw);
  const uint32_t q_s = base, k_s = base + kTile, v_s = base + 3 * kTile, p_s = base + 5 * kTile;
  const uint32_t bar = base + 7 * kTile;
  const uint32_t q_full = bar, k_full0 = bar + 8, v_full0 = bar + 24, kv_empty0 = bar + 40, s_full = bar + 56,
                 p_ready = bar + 64, o_full = bar + 72, tmem_slot = bar + 80;
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int qb = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int q0 = qb * 128;
  const int nkb_all = (p.S + 127) / 128;
  const int nkb = p.causal ? (qb + 1 < nkb_all ? qb + 1 : nkb_all) : nkb_all;
  const int row_base = b * p.S;

  if (threadIdx.x == 0) {
    mbar_init(q_full, 1);
    for (int s = 0; s < 2; ++s) { mbar_init(k_full0 + 8 * s, 1); mbar_init(v_full0 + 8 * s, 1); mbar_init(kv_empty0 + 8 * s, 1); }
    mbar_init(s_full, 1);
    mbar_init(p_ready, 4);
    mbar_init(o_full, 1);
    mbar_init_fence();
    prefetch_tensormap(&map_qkv);
  }
  if (warp == 2) tmem_alloc(tmem_slot, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - base));
  const uint32_t tmem_s = tmem_base, tmem_o = tmem_base + 128;

  if (warp == 0) {
    if (elect_one()) {
      auto load_kv = [&](int kb) {
        const int st = kb & 1;
        mbar_expect_tx(k_full0 + 8 * st, kTile);
        tma_load_2d(k_s + st * kTile, &map_qkv, k_full0 + 8 * st, p.D + h * 64, row_base + kb * 128);
        mbar_expect_tx(v_full0 + 8 * st, kTile);
        tma_load_2d(v_s + st * kTile, &map_qkv, v_full0 + 8 * st, 2 * p.D + h * 64, row_base + kb * 128);
      };
      mbar_expect_tx(q_full, kTile);
      tma_load_2d(q_s, &map_qkv, q_full, h * 64, row_base + q0);
      load_kv(0);
      if (nkb > 1) load_kv(1);
      mbar_wait(q_full, 0);
      for (int kb = 0; kb < nkb; ++kb) {
        const int st = kb & 1;
        const uint32_t ph = (uint32_t)(kb >> 1) & 1u;
        mbar_wait(k_full0 + 8 * st, ph);
        tc_fence_after();
        mma_qk(tmem_s, q_s, k_s + st * kTile);
        tc_commit(s_full);
        mbar_wait(p_ready, (uint32_t)kb & 1u);          // P is in shared memory, S has been read
        mbar_wait(v_full0 + 8 * st, ph);
        tc_fence_after();
        mma_pv(tmem_o, p_s, v_s + st * kTile, 0u);
        tc_commit(o_full);
        tc_commit(kv_empty0 + 8 * st);
        if (kb + 2 < nkb) {
          mbar_wait(kv_empty0 + 8 * st, ph);            // the MMAs above have read this stage
          load_kv(kb + 2);
        }
      }
    }
    __syncwarp();
  } else if (warp >= 4) {
    const int ew = warp - 4, r = ew * 32 + lane;          // this thread's query row inside the tile
    const int q_idx = q0 + r;
    const bool q_ok = q_idx < p.S;
    const uint32_t lane_addr = (uint32_t)(ew * 32) << 16;
    float o[64];
#pragma unroll
    for (int i = 0; i < 64; ++i) o[i] = 0.f;
    float m = -INFINITY, l = 0.f;
    for (int kb = 0; kb < nkb; ++kb) {
      mbar_wait(s_full, (uint32_t)kb & 1u);
      tc_fence_after();
      const int key0 = kb * 128;
      // pass 1: row maximum
      float mx = -INFINITY;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        tc_ld_32x32(tmem_s + lane_addr + c * 32, v);
        tc_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int key = key0 + c * 32 + j;
          const bool ok = q_ok && key < p.S && (!p.causal || key <= q_idx);
          if (ok) mx = fmaxf(mx, __uint_as_float(v[j]) * p.scale_log2);
        }
      }
      const float m_new = fmaxf(m, mx);
      const float alpha = (m_new == -INFINITY) ? 1.f : exp2f(m - m_new);
      // pass 2: P, its row sum, P -> shared memory
      float sum = 0.f;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        tc_ld_32x32(tmem_s + lane_addr + c * 32, v);
        tc_ld_wait();
        float pr[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int key = key0 + c * 32 + j;
          const bool ok = q_ok && key < p.S && (!p.causal || key <= q_idx);
          pr[j] = ok ? exp2f(__uint_as_float(v[j]) * p.scale_log2 - m_new) : 0.f;
          sum += pr[j];
        }
        store_row_chunk(p_s, r, c * 32, pr);
      }
      l = l * alpha + sum;
      m = m_new;
#pragma unroll
      for (int i = 0; i < 64; ++i) o[i] *= alpha;
      tc_fence_before();
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_ready);
      // O += P V
      mbar_wait(o_full, (uint32_t)kb & 1u);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t v[32];
        tc_ld_32x32(tmem_o + lane_addr + c * 32, v);
        tc_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) o[c * 32 + j] += __uint_as_float(v[j]);
      }
      tc_fence_before();
    }
    if (q_ok) {
      const float inv = l > 0.f ? 1.f / l : 0.f;
      uint4* dst = reinterpret_cast<uint4*>(p.out + (size_t)(row_base + q_idx) * p.D + h * 64);
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        uint4 w;
        w.x = pack_bf16(o[8 * u] * inv, o[8 * u + 1] * inv);
        w.y = pack_bf16(o[8 * u + 2] * inv, o[8 * u + 3] * inv);
        w.z = pack_bf16(o[8 * u + 4] * inv, o[8 * u + 5] * inv);
        w.w = pack_bf16(o[8 * u + 6] * inv, o[8 * u + 7] * inv);
        dst[u] = w;
      }
      p.lse2[((size_t)b * p.H + h) * p.S + q_idx] = m + log2f(l);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_free(tmem_base, 256);
  }
}

// delta[b, h, row] = sum_d dO[row, h*64 + d] * O[row, h*64 + d]
__global__ void __launch_bounds__(256) attn_delta_kernel(const __nv_bfloat16* __restrict__ dout,
                                                         const __nv_bfloat16* __restrict__ out, float* __restrict__ delta,
                                                         int B, int S, int H, int D) {
  const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= (size_t)B * S * H) return;
  const int h = (int)(i % H);
  const size_t row = i / H;
  const uint4* a = reinterpret_cast<const uint4*>(dout + row * D + h * 64);
  const uint4* c = reinterpret_cast<const uint4*>(out + row * D + h * 64);
  float s = 0.f;
#pragma unroll
  for (int u = 0; u < 8; ++u) {
    const uint4 x = __ldg(a + u), y = __ldg(c + u);
    s += bf16_lo(x.x) * bf16_lo(y.x) + bf16_hi(x.x) * bf16_hi(y.x) + bf16_lo(x.y) * bf16_lo(y.y) + bf16_hi(x.y) * bf16_hi(y.y) +
         bf16_lo(x.z) * bf16_lo(y.z) + bf16_hi(x.z) * bf16_hi(y.z) + bf16_lo(x.w) * bf16_lo(y.w) + bf16_hi(x.w) * bf16_hi(y.w);
  }
  const size_t b = row / S, srow = row % S;
  delta[(b * H + h) * S + srow] = s;
}

// P and dS of one (query block, key block) pair for this thread's query row, 32 keys at a time:
// p = exp2(s*c - lse2), ds = p * (dp - delta) * scale; masked entries are exactly zero.
struct RowCtx {
  int q_idx, key0, S, causal;
  bool q_ok;
  float lse2, delta, scale_log2, scale;
};
__device__ __forceinline__ void p_ds_chunk(const RowCtx& c, int col0, const uint32_t (&s)[32], const uint32_t (&dp)[32],
                                           float (&pr)[32], float (&ds)[32]) {
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    const int key = c.key0 + col0 + j;
    const bool ok = c.q_ok && key < c.S && (!c.causal || key <= c.q_idx);
    const float pv = ok ? exp2f(__uint_as_float(s[j]) * c.scale_log2 - c.lse2) : 0.f;
    pr[j] = pv;
    ds[j] = pv * (__uint_as_float(dp[j]) - c.delta) * c.scale;
  }
}

// ---------------------------------------------------------------------------------------------------------
// backward, key-block kernel: dK, dV
// shared memory: K | V | Q[2] | dO[2] | P (2 tiles) | dS (2 tiles) | barriers
constexpr int kDkvSmem = kTile * (1 + 1 + 2 + 2 + 2 + 2) + 256;

__global__ void __launch_bounds__(kThreads, 1) attn_dkv_kernel(const __grid_constant__ CUtensorMap map_qkv,
                                                               const __grid_constant__ CUtensorMap map_do,
                                                               const AttnParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = smem_u32(smem_raw);
  const uint32_t k_s = base, v_s = base + kTile, q_s = base + 2 * kTile, do_s = base + 4 * kTile, p_s = base + 6 * kTile,
                 ds_s = base + 8 * kTile;
  const uint32_t bar = base + 10 * kTile;
  const uint32_t kv_full = bar, in_full0 = bar + 8, in_empty0 = bar + 24, sdp_full = bar + 40, pds_ready = bar + 48,
                 pds_free = bar + 56, acc_full = bar + 64, tmem_slot = bar + 72;
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int kb = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int key0 = kb * 128;
  const int nqb = (p.S + 127) / 128;
  const int qb0 = p.causal ? kb : 0;
  const int row_base = b * p.S;

  if (threadIdx.x == 0) {
    mbar_init(kv_full, 1);
    for (int s = 0; s < 2; ++s) { mbar_init(in_full0 + 8 * s, 1); mbar_init(in_empty0 + 8 * s, 1); }
    mbar_init(sdp_full, 1);
    mbar_init(pds_ready, 4);
    mbar_init(pds_free, 1);
    mbar_init(acc_full, 1);
    mbar_init_fence();
    prefetch_tensormap(&map_qkv);
    prefetch_tensormap(&map_do);
  }
  if (warp == 2) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - base));
  const uint32_t tmem_s = tmem_base, tmem_dp = tmem_base + 128, tmem_dv = tmem_base + 256, tmem_dk = tmem_base + 320;

  if (warp == 0) {
    if (elect_one()) {
      auto load_in = [&](int i) {                      // i-th query block of this CTA's loop
        const int st = i & 1, qb = qb0 + i;
        mbar_expect_tx(in_full0 + 8 * st, 2 * kTile);
        tma_load_2d(q_s + st * kTile, &map_qkv, in_full0 + 8 * st, h * 64, row_base + qb * 128);
        tma_load_2d(do_s + st * kTile, &map_do, in_full0 + 8 * st, h * 64, row_base + qb * 128);
      };
      const int n = nqb - qb0;
      mbar_expect_tx(kv_full, 2 * kTile);
      tma_load_2d(k_s, &map_qkv, kv_full, p.D + h * 64, row_base + key0);
      tma_load_2d(v_s, &map_qkv, kv_full, 2 * p.D + h * 64, row_base + key0);
      load_in(0);
      if (n > 1) load_in(1);
      mbar_wait(kv_full, 0);
      for (int i = 0; i < n; ++i) {
        const int st = i & 1;
        const uint32_t ph = (uint32_t)(i >> 1) & 1u;
        mbar_wait(in_full0 + 8 * st, ph);
        tc_fence_after();
        mma_qk(tmem_s, q_s + st * kTile, k_s);          // S  = Q_i K^T
        mma_qk(tmem_dp, do_s + st * kTile, v_s);        // dP = dO_i V^T
        tc_commit(sdp_full);
        mbar_wait(pds_ready, (uint32_t)i & 1u);         // P and dS tiles written, S / dP read
        tc_fence_after();
        mma_tn(tmem_dv, p_s, do_s + st * kTile, i ? 1u : 0u);    // dV += P^T dO_i
        mma_tn(tmem_dk, ds_s, q_s + st * kTile, i ? 1u : 0u);    // dK += dS^T Q_i
        tc_commit(pds_free);
        tc_commit(in_empty0 + 8 * st);
        if (i + 2 < n) {
          mbar_wait(in_empty0 + 8 * st, ph);
          load_in(i + 2);
        }
      }
      tc_commit(acc_full);
    }
    __syncwarp();
  } else if (warp >= 4) {
    const int ew = warp - 4, r = ew * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(ew * 32) << 16;
    const int n = nqb - qb0;
    for (int i = 0; i < n; ++i) {
      const int qb = qb0 + i;
      RowCtx c;
      c.q_idx = qb * 128 + r; c.key0 = key0; c.S = p.S; c.causal = p.causal;
      c.q_ok = c.q_idx < p.S;
      c.scale_log2 = p.scale_log2; c.scale = p.scale;
      const size_t srow = ((size_t)b * p.H + h) * p.S + (size_t)(c.q_ok ? c.q_idx : 0);
      c.lse2 = c.q_ok ? p.lse2[srow] : 0.f;
      c.delta = c.q_ok ? p.delta[srow] : 0.f;
      mbar_wait(sdp_full, (uint32_t)i & 1u);
      tc_fence_after();
      if (i > 0) mbar_wait(pds_free, (uint32_t)(i - 1) & 1u);     // the previous pair's MMAs have read the tiles
#pragma unroll 1
      for (int ch = 0; ch < 4; ++ch) {
        uint32_t s[32], dp[32];
        tc_ld_32x32(tmem_s + lane_addr + ch * 32, s);
        tc_ld_32x32(tmem_dp + lane_addr + ch * 32, dp);
        tc_ld_wait();
        float pr[32], ds[32];
        p_ds_chunk(c, ch * 32, s, dp, pr, ds);
        store_row_chunk(p_s, r, ch * 32, pr);
        store_row_chunk(ds_s, r, ch * 32, ds);
      }
      tc_fence_before();
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(pds_ready);
    }
    // epilogue: this thread's key row of dV and dK
    mbar_wait(acc_full, 0);
    tc_fence_after();
    const int key = key0 + r;
#pragma unroll 1
    for (int which = 0; which < 2; ++which) {            // 0: dV -> column block 2, 1: dK -> column block 1
      const uint32_t t = (which == 0 ? tmem_dv : tmem_dk) + lane_addr;
      uint32_t v[64];
      tc_ld_32x32(t, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
      tc_ld_32x32(t + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
      tc_ld_wait();
      if (key < p.S) {
        uint4* dst = reinterpret_cast<uint4*>(p.dqkv + (size_t)(row_base + key) * (3 * p.D) + (which == 0 ? 2 : 1) * p.D + h * 64);
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          uint4 w;
          w.x = pack_bf16(__uint_as_float(v[8 * u]), __uint_as_float(v[8 * u + 1]));
          w.y = pack_bf16(__uint_as_float(v[8 * u + 2]), __uint_as_float(v[8 * u + 3]));
          w.z = pack_bf16(__uint_as_float(v[8 * u + 4]), __uint_as_float(v[8 * u + 5]));
          w.w = pack_bf16(__uint_as_float(v[8 * u + 6]), __uint_as_float(v[8 * u + 7]));
          dst[u] = w;
        }
      }
    }
    tc_fence_before();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_free(tmem_base, 512);
  }
}

// ---------------------------------------------------------------------------------------------------------
// backward, query-block kernel: dQ
// shared memory: Q | dO | K[2] | V[2] | dS (2 tiles) | barriers
constexpr int kDqSmem = kTile * (1 + 1 + 2 + 2 + 2) + 256;

__global__ void __launch_bounds__(kThreads, 1) attn_dq_kernel(const __grid_constant__ CUtensorMap map_qkv,
                                                              const __grid_constant__ CUtensorMap map_do,
                                                              const AttnParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = smem_u32(smem_raw);
  const uint32_t q_s = base, do_s = base + kTile, k_s = base + 2 * kTile, v_s = base + 4 * kTile, ds_s = base + 6 * kTile;
  const uint32_t bar = base + 8 * kTile;
  const uint32_t q_full = bar, kv_full0 = bar + 8, kv_empty0 = bar + 24, sdp_full = bar + 40, ds_ready = bar + 48,
                 ds_free = bar + 56, acc_full = bar + 64, tmem_slot = bar + 72;
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int qb = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int q0 = qb * 128;
  const int nkb_all = (p.S + 127) / 128;
  const int nkb = p.causal ? (qb + 1 < nkb_all ? qb + 1 : nkb_all) : nkb_all;
  const int row_base = b * p.S;

  if (threadIdx.x == 0) {
    mbar_init(q_full, 1);
    for (int s = 0; s < 2; ++s) { mbar_init(kv_full0 + 8 * s, 1); mbar_init(kv_empty0 + 8 * s, 1); }
    mbar_init(sdp_full, 1);
    mbar_init(ds_ready, 4);
    mbar_init(ds_free, 1);
    mbar_init(acc_full, 1);
    mbar_init_fence();
    prefetch_tensormap(&map_qkv);
    prefetch_tensormap(&map_do);
  }
  if (warp == 2) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - base));
  const uint32_t tmem_s = tmem_base, tmem_dp = tmem_base + 128, tmem_dq = tmem_base + 256;

  if (warp == 0) {
    if (elect_one()) {
      auto load_kv = [&](int kb) {
        const int st = kb & 1;
        mbar_expect_tx(kv_full0 + 8 * st, 2 * kTile);
        tma_load_2d(k_s + st * kTile, &map_qkv, kv_full0 + 8 * st, p.D + h * 64, row_base + kb * 128);
        tma_load_2d(v_s + st * kTile, &map_qkv, kv_full0 + 8 * st, 2 * p.D + h * 64, row_base + kb * 128);
      };
      mbar_expect_tx(q_full, 2 * kTile);
      tma_load_2d(q_s, &map_qkv, q_full, h * 64, row_base + q0);
      tma_load_2d(do_s, &map_do, q_full, h * 64, row_base + q0);
      load_kv(0);
      if (nkb > 1) load_kv(1);
      mbar_wait(q_full, 0);
      for (int kb = 0; kb < nkb; ++kb) {
        const int st = kb & 1;
        const uint32_t ph = (uint32_t)(kb >> 1) & 1u;
        mbar_wait(kv_full0 + 8 * st, ph);
        tc_fence_after();
        mma_qk(tmem_s, q_s, k_s + st * kTile);          // S  = Q K_j^T
        mma_qk(tmem_dp, do_s, v_s + st * kTile);        // dP = dO V_j^T
        tc_commit(sdp_full);
        mbar_wait(ds_ready, (uint32_t)kb & 1u);
        tc_fence_after();
        mma_pv(tmem_dq, ds_s, k_s + st * kTile, kb ? 1u : 0u);     // dQ += dS K_j
        tc_commit(ds_free);
        tc_commit(kv_empty0 + 8 * st);
        if (kb + 2 < nkb) {
          mbar_wait(kv_empty0 + 8 * st, ph);
          load_kv(kb + 2);
        }
      }
      tc_commit(acc_full);
    }
    __syncwarp();
  } else if (warp >= 4) {
    const int ew = warp - 4, r = ew * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(ew * 32) << 16;
    RowCtx c;
    c.q_idx = q0 + r; c.S = p.S; c.causal = p.causal;
    c.q_ok = c.q_idx < p.S;
    c.scale_log2 = p.scale_log2; c.scale = p.scale;
    const size_t srow = ((size_t)b * p.H + h) * p.S + (size_t)(c.q_ok ? c.q_idx : 0);
    c.lse2 = c.q_ok ? p.lse2[srow] : 0.f;
    c.delta = c.q_ok ? p.delta[srow] : 0.f;
    for (int kb = 0; kb < nkb; ++kb) {
      c.key0 = kb * 128;
      mbar_wait(sdp_full, (uint32_t)kb & 1u);
      tc_fence_after();
      if (kb > 0) mbar_wait(ds_free, (uint32_t)(kb - 1) & 1u);
#pragma unroll 1
      for (int ch = 0; ch < 4; ++ch) {
        uint32_t s[32], dp[32];
        tc_ld_32x32(tmem_s + lane_addr + ch * 32, s);
        tc_ld_32x32(tmem_dp + lane_addr + ch * 32, dp);
        tc_ld_wait();
        float pr[32], ds[32];
        p_ds_chunk(c, ch * 32, s, dp, pr, ds);
        store_row_chunk(ds_s, r, ch * 32, ds);
      }
      tc_fence_before();
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(ds_ready);
    }
    mbar_wait(acc_full, 0);
    tc_fence_after();
    uint32_t v[64];
    tc_ld_32x32(tmem_dq + lane_addr, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
    tc_ld_32x32(tmem_dq + lane_addr + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
    tc_ld_wait();
    if (c.q_ok) {
      uint4* dst = reinterpret_cast<uint4*>(p.dqkv + (size_t)(row_base + c.q_idx) * (3 * p.D) + h * 64);
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        uint4 w;
        w.x = pack_bf16(__uint_as_float(v[8 * u]), __uint_as_float(v[8 * u + 1]));
        w.y = pack_bf16(__uint_as_float(v[8 * u + 2]), __uint_as_float(v[8 * u + 3]));
        w.z = pack_bf16(__uint_as_float(v[8 * u + 4]), __uint_as_float(v[8 * u + 5]));
        w.w = pack_bf16(__uint_as_float(v[8 * u + 6]), __uint_as_float(v[8 * u + 7]));
        dst[u] = w;
      }
    }
    tc_fence_before();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_free(tmem_base, 512);
  }
}

// ---- host -------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                             const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                             CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeFn attn_encode() {
  static EncodeFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* q = nullptr;
    cudaDriverEntryPointQueryResult r;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &q, cudaEnableDefault, &r) == cudaSuccess &&
        r == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeFn>(q);
  });
  return fn;
}

// row-major bf16 [rows, cols]; box = {64 columns, 128 rows}, 128-byte swizzle
int attn_map(CUtensorMap* out, const void* base, size_t rows, int cols) {
  EncodeFn enc = attn_encode();
  if (!enc) return fail("cuTensorMapEncodeTiled entry point not available%s");
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)cols * 2};
  cuuint32_t box[2] = {64, 128};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail("cuTensorMapEncodeTiled (attention) failed (%s code %ld)", "", (long)r);
  return 0;
}

int attn_check(const char* who, int B, int S, int H) {
  if (B <= 0 || S <= 0 || H <= 0) return fail("%s: empty problem", who);
  if ((long long)B * S >= (1ll << 31) / 4 || S > 65536 || H > 65535 || B > 65535)
    return fail("%s: problem too large", who);
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) {
    if (cudaFree(nullptr) != cudaSuccess) return fail("%s: no CUDA context", who);
    ctx_bound = true;
  }
  return 0;
}

}  // namespace

extern "C" {

// out [B*S, H*64] = softmax(Q K^T / 8 (+ causal mask)) V per (batch, head), all read from qkv [B*S, 3*H*64]
// (columns: Q | K | V, each H heads of 64); lse2 [B, H, S] fp32 is saved for the backward.
int tb200_attention_fwd(const void* qkv, void* out, float* lse2, int B, int S, int H, int causal, cudaStream_t stream) {
  if (int rc = attn_check("tb200_attention_fwd", B, S, H)) return rc;
  const int D = H * 64;
  CUtensorMap mq;
  if (int rc = attn_map(&mq, qkv, (size_t)B * S, 3 * D)) return rc;
  AttnParams p{};
  p.B = B; p.S = S; p.H = H; p.D = D; p.causal = causal ? 1 : 0;
  p.scale = 0.125f; p.scale_log2 = 0.125f * kLog2e;
  p.out = reinterpret_cast<__nv_bfloat16*>(out);
  p.lse2 = lse2;
  static DeviceOnce once;
  if (int rc = ensure_dynamic_smem(once, attn_fwd_kernel, kFwdSmem)) return rc;
  attn_fwd_kernel<<<dim3((S + 127) / 128, H, B), kThreads, kFwdSmem, stream>>>(mq, p);
  return launched();
}

// dqkv [B*S, 3*H*64] from dout [B*S, H*64], the forward's inputs / outputs and lse2; delta: [B, H, S] fp32 scratch.
int tb200_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse2, float* delta, void* dqkv,
                        int B, int S, int H, int causal, cudaStream_t stream) {
  if (int rc = attn_check("tb200_attention_bwd", B, S, H)) return rc;
  const int D = H * 64;
  CUtensorMap mq, md;
  if (int rc = attn_map(&mq, qkv, (size_t)B * S, 3 * D)) return rc;
  if (int rc = attn_map(&md, dout, (size_t)B * S, D)) return rc;
  AttnParams p{};
  p.B = B; p.S = S; p.H = H; p.D = D; p.causal = causal ? 1 : 0;
  p.scale = 0.125f; p.scale_log2 = 0.125f * kLog2e;
  p.lse2 = const_cast<float*>(lse2);
  p.delta = delta;
  p.dqkv = reinterpret_cast<__nv_bfloat16*>(dqkv);
  const size_t n = (size_t)B * S * H;
  attn_delta_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(dout),
                                                                     reinterpret_cast<const __nv_bfloat16*>(out), delta, B, S, H, D);
  if (int rc = launched()) return rc;
  static DeviceOnce once_kv, once_q;
  if (int rc = ensure_dynamic_smem(once_kv, attn_dkv_kernel, kDkvSmem)) return rc;
  if (int rc = ensure_dynamic_smem(once_q, attn_dq_kernel, kDqSmem)) return rc;
  const dim3 grid((S + 127) / 128, H, B);
  attn_dkv_kernel<<<grid, kThreads, kDkvSmem, stream>>>(mq, md, p);
  if (int rc = launched()) return rc;
  attn_dq_kernel<<<grid, kThreads, kDqSmem, stream>>>(mq, md, p);
  return launched();
}

}  // extern "C"
