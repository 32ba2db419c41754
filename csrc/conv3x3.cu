// 3x3 / stride 1 / pad 1 convolution, forward and input gradient, as a halo-tile implicit GEMM on tcgen05
// (K5: "convs as implicit GEMM on the same mainloop"; the trace's model zoo is convolutional —
// /root/reference model/model_factory.py:31-55, per-tensor tables core/models.py:8-20).
//
//   y[n,h,w,k] = sum_{r,s,c} x[n, h+r-1, w+s-1, c] * wt[k,r,s,c]          x, y NHWC bf16, fp32 accumulation
//
// Second generation of csrc/conv_halo.cu. What the first one measured on B200 (profiles/conv_r2.md): the
// numerics were right, but the tensor pipe sat idle 65-79 % of the time because the single MMA-issuing thread
// needed ~640 cycles of integer divisions, descriptor construction and per-tap barrier traffic for every 4
// MMAs. This version
//   * runs the producer and MMA warps converged, with single-thread instructions under elect.sync, so the
//     descriptors live in uniform registers; the nine taps are unrolled with compile-time (r, s); all
//     divisions are multiplications by host-computed reciprocals
//   * groups the weight tiles of three taps behind one barrier pair (BLOCK_N <= 128)
//   * lays positions out with a pitch of W+1 / H+1 (the zero column / row between two image rows / images is
//     shared by both neighbours) instead of a power of two: 3-15 % junk positions at 56x56..14x14 and 31 % at
//     7x7, where the power-of-two layout wasted 18-31 % and 194 %. Tiles are 128 consecutive positions of that
//     flat space; the producer fetches the padded rows a tile touches ONCE per 64-channel block (TMA boxes
//     {64 ch, W+1 px} landing at 128-byte-aligned offsets; out-of-range rows / columns arrive zero-filled =
//     the padding) and every tap is a descriptor start shifted by ((r-1)*(W+1) + (s-1)) rows
//   * can emit per-channel sum / sum of squares of the bf16 outputs from the epilogue (BatchNorm statistics),
//     which removes the statistics pass of the following fused BatchNorm (a full read of the activation)
//   * dgrad mode: the same kernel over dy with the taps flipped and the forward weight read as MN-major
//     B slabs (no transposed copy of the weights).
#include "common.cuh"
#include "tc_ptx.cuh"
#include "../include/tiresias_b200.h"

#include <cuda.h>
#include <mutex>

using namespace tb200;
using namespace tb200::ptx;

namespace {

constexpr int kC3Threads = 256;
constexpr int kC3Acc = 2;          // TMEM accumulator stages
constexpr int kEpiBytes = 4 * 2 * 4096;   // 4 epilogue warps x 2 staging tiles of 32 rows x 128 B

struct C3Params {
  int N, H, W, C, K;               // batch, image, reduction channels, output channels
  int Wp, Hp;                      // pitches of the flat padded space: W+1, H+1 (segment mode: Wp = 130)
  int NR;                          // padded rows fetched per A stage
  int seg;                         // segment mode (wide images): a tile = 128 consecutive columns of ONE image row, its
                                   // stage = three boxes {64 ch, 130 columns} of the rows above / at / below it
  int nseg;                        // segments per image row
  unsigned long long mSeg, mH;     // fastdiv multipliers for nseg and H
  int cblocks;                     // C / 64
  int num_m, num_n;
  int total_pos;                   // N * Hp * Wp
  unsigned long long mWp, mHp;     // fastdiv multipliers
  int a_stage_bytes, a_stages;
  int b_groups;                    // ring depth in groups of taps (1 when resident)
  int epi_off, bar_off;
  int resident;                    // all weight tiles of the (single) n block stay in shared memory
  __nv_bfloat16* y;
  float* stats;                    // [gridDim.x * 4][K][2] per-channel (sum, sum of squares) partials, or null
  const void* bias;                // optional per-output-channel bias (fprop; VGG / AlexNet style layers)
  int bias_dtype;
  int relu;                        // clamp the (biased) output at zero
  int bias_off;                    // shared-memory offset of the epilogue warps' bias rows (4 x 256 floats)
};

template <int BLOCK_N, int TPG, bool DGRAD>
__global__ void __launch_bounds__(kC3Threads, 1)
conv3x3_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w, const C3Params p) {
  constexpr int kTileBytes = BLOCK_N * 128;               // one (tap, 64-channel block) weight tile
  constexpr int kGroups = 9 / TPG;
  constexpr int kTmemCols = kC3Acc * BLOCK_N;
  static_assert(9 % TPG == 0, "taps per group");
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = smem_u32(smem_raw);
  const uint32_t a_smem = base;
  const uint32_t b_smem = base + (uint32_t)(p.a_stages * p.a_stage_bytes);
  const uint32_t epi_smem = base + (uint32_t)p.epi_off;
  const uint32_t bar = base + (uint32_t)p.bar_off;
  auto a_full = [&](int s) { return bar + 8u * s; };
  auto a_empty = [&](int s) { return bar + 8u * (p.a_stages + s); };
  auto b_full = [&](int s) { return bar + 8u * (2 * p.a_stages + s); };
  auto b_empty = [&](int s) { return bar + 8u * (2 * p.a_stages + p.b_groups + s); };
  const uint32_t t_base = bar + 8u * (2 * p.a_stages + 2 * p.b_groups);
  auto t_full = [&](int s) { return t_base + 8u * s; };
  auto t_empty = [&](int s) { return t_base + 8u * (kC3Acc + s); };
  const uint32_t tmem_slot = t_base + 8u * 2 * kC3Acc;

  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&map_x);
    prefetch_tensormap(&map_w);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < p.a_stages; ++s) { mbar_init(a_full(s), 1); mbar_init(a_empty(s), 1); }
    for (int s = 0; s < p.b_groups; ++s) { mbar_init(b_full(s), 1); mbar_init(b_empty(s), 1); }
    for (int s = 0; s < kC3Acc; ++s) { mbar_init(t_full(s), 1); mbar_init(t_empty(s), 4); }
    mbar_init_fence();
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - base));

  // Work items: n block outermost (a CTA's consecutive tiles share the weights).
  const int items = p.num_m * p.num_n;
  const uint32_t row_bytes = (uint32_t)p.Wp * 128u;       // one padded row of 64 channels

  // weight tile of (tap t, reduction block cb, n block): forward = box {64 c, BLOCK_N k} of wt[k][t][c];
  // dgrad = BLOCK_N/64 slabs {64 k_out, 64 c_red} of the forward weight wf[c_red][8-t][k_out] (MN-major)
  auto load_b_tile = [&](uint32_t dst, uint32_t bar_addr, int t, int cb, int n_blk) {
    if constexpr (!DGRAD) {
      tma_load_2d(dst, &map_w, bar_addr, t * p.C + cb * 64, n_blk * BLOCK_N);
    } else {
#pragma unroll
      for (int sl = 0; sl < BLOCK_N / 64; ++sl)
        tma_load_2d(dst + sl * 8192, &map_w, bar_addr, (8 - t) * p.K + n_blk * BLOCK_N + sl * 64, cb * 64);
    }
  };

  if (warp == 0) {
    // ================= producer (converged warp, one elected lane issues) =================
    if (p.resident) {
      if (elect_one()) {
        mbar_expect_tx(b_full(0), (uint32_t)(9 * p.cblocks) * kTileBytes);
        for (int cb = 0; cb < p.cblocks; ++cb)
          for (int t = 0; t < 9; ++t) load_b_tile(b_smem + (uint32_t)(cb * 9 + t) * kTileBytes, b_full(0), t, cb, 0);
      }
      __syncwarp();
    }
    int as = 0, bg = 0;
    uint32_t aph = 0, bph = 0;
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      const int m_tile = it % p.num_m, n_blk = it / p.num_m;
      const uint32_t q0 = (uint32_t)m_tile * 128u;
      // first padded row the tile's windows touch: floor((q0 - Wp - 1) / Wp), can be -1 or -2 for the first tile
      const int rho0 = (int)fastdiv(q0 + (uint32_t)p.Wp - 1u, p.mWp) - 2;
      // segment mode: tile -> (image, row, segment)
      const uint32_t srow = fastdiv((uint32_t)m_tile, p.mSeg), sseg = (uint32_t)m_tile - srow * (uint32_t)p.nseg;
      const uint32_t sn = fastdiv(srow, p.mH), sh = srow - sn * (uint32_t)p.H;
      for (int cb = 0; cb < p.cblocks; ++cb) {
        mbar_wait(a_empty(as), aph ^ 1);
        if (elect_one()) {
          mbar_expect_tx(a_full(as), (uint32_t)p.NR * row_bytes);
          uint32_t dst = a_smem + (uint32_t)as * p.a_stage_bytes;
          if (p.seg) {
            // rows h-1, h, h+1, columns w0-1 .. w0+128 (whatever lies outside the image arrives as zeros)
            for (int i = 0; i < 3; ++i) {
              tma_load_4d(dst, &map_x, a_full(as), cb * 64, (int)sseg * 128 - 1, (int)sh - 1 + i, (int)sn);
              dst += row_bytes;
            }
          } else {
            int n, hp;
            if (rho0 < 0) { n = -1; hp = rho0 + p.Hp; }
            else { n = (int)fastdiv((uint32_t)rho0, p.mHp); hp = rho0 - n * p.Hp; }
            for (int i = 0; i < p.NR; ++i) {
              // whole padded row: columns 0..W (column W, row H and images outside [0, N) arrive as zeros)
              tma_load_4d(dst, &map_x, a_full(as), cb * 64, 0, hp, n);
              dst += row_bytes;
              if (++hp == p.Hp) { hp = 0; ++n; }
            }
          }
        }
        __syncwarp();
        if (++as == p.a_stages) { as = 0; aph ^= 1; }
        if (!p.resident) {
#pragma unroll 1
          for (int g = 0; g < kGroups; ++g) {
            mbar_wait(b_empty(bg), bph ^ 1);
            if (elect_one()) {
              mbar_expect_tx(b_full(bg), TPG * kTileBytes);
#pragma unroll
              for (int tt = 0; tt < TPG; ++tt)
                load_b_tile(b_smem + (uint32_t)(bg * TPG + tt) * kTileBytes, b_full(bg), g * TPG + tt, cb, n_blk);
            }
            __syncwarp();
            if (++bg == p.b_groups) { bg = 0; bph ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer (converged warp, one elected lane issues) =================
    constexpr uint32_t idesc = make_idesc_bf16(128, BLOCK_N, 0, DGRAD ? 1 : 0);
    constexpr uint64_t kBDesc = DGRAD ? kDescMNMajor : kDescKMajor;
    constexpr uint32_t kBStep = DGRAD ? 128u : 2u;        // 16 reduction elements, in 16-byte units
    const int wp8 = p.Wp * 8;                             // one padded row, in 16-byte units
    int as = 0, bg = 0, acc = 0;
    uint32_t aph = 0, bph = 0, accph = 0;
    if (p.resident) {
      mbar_wait(b_full(0), 0);
      tc_fence_after();
    }
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      const int m_tile = it % p.num_m;
      const uint32_t q0 = (uint32_t)m_tile * 128u;
      const int rho0 = (int)fastdiv(q0 + (uint32_t)p.Wp - 1u, p.mWp) - 2;
      // rows between the stage's first position and the tile's first position: in [Wp+1, 2*Wp]
      // (segment mode: always one stage row + one column)
      const uint32_t woff8 = p.seg ? (uint32_t)(p.Wp + 1) * 8u : (uint32_t)((int)q0 - rho0 * p.Wp) * 8u;
      mbar_wait(t_empty(acc), accph ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + (uint32_t)(acc * BLOCK_N);
      uint32_t accumulate = 0;
      for (int cb = 0; cb < p.cblocks; ++cb) {
        mbar_wait(a_full(as), aph);
        tc_fence_after();
        // descriptor (low word) of the window of tap (1, 1): the tile's own positions
        const uint32_t a_mid = desc_units(a_smem + (uint32_t)as * p.a_stage_bytes) + woff8;
#pragma unroll
        for (int g = 0; g < kGroups; ++g) {
          uint32_t b_units;
          if (p.resident) {
            b_units = desc_units(b_smem + (uint32_t)(cb * 9 + g * TPG) * kTileBytes);
          } else {
            mbar_wait(b_full(bg), bph);
            tc_fence_after();
            b_units = desc_units(b_smem + (uint32_t)(bg * TPG) * kTileBytes);
          }
#pragma unroll
          for (int tt = 0; tt < TPG; ++tt) {
            constexpr int kUnitsPerTile = kTileBytes / 16;
            const int t = g * TPG + tt, r = t / 3, s = t % 3;           // compile-time after unrolling
            const uint32_t a_units = a_mid + (uint32_t)((r - 1) * wp8 + (s - 1) * 8);
            const uint64_t adesc = kDescKMajor | (uint64_t)a_units;
            const uint64_t bdesc = kBDesc | (uint64_t)(b_units + (uint32_t)(tt * kUnitsPerTile));
            if (elect_one()) {
#pragma unroll
              for (int k = 0; k < 4; ++k)
                tc_mma_bf16(tmem_d, adesc + 2u * k, bdesc + kBStep * k, idesc, (k == 0) ? accumulate : 1u);
            }
            __syncwarp();
            accumulate = 1;
          }
          if (!p.resident) {
            if (elect_one()) tc_commit(b_empty(bg));
            __syncwarp();
            if (++bg == p.b_groups) { bg = 0; bph ^= 1; }
          }
        }
        if (elect_one()) tc_commit(a_empty(as));
        __syncwarp();
        if (++as == p.a_stages) { as = 0; aph ^= 1; }
      }
      if (elect_one()) tc_commit(t_full(acc));
      __syncwarp();
      if (++acc == kC3Acc) { acc = 0; accph ^= 1; }
    }
  } else if (warp >= 4) {
    // ================= epilogue =================
    // Warp ew owns tile rows [32*ew, +32); lane = row for the TMEM loads. Per 64 output channels:
    // TMEM -> registers -> bf16 -> swizzled staging tile (32 rows x 128 B) -> (statistics from the staged
    // bf16 values) -> coalesced 16-byte global stores of the valid rows (4 rows x 128 B per instruction).
    const int ew = warp - 4;
    int acc = 0;
    uint32_t accph = 0;
    const uint32_t bufs = epi_smem + (uint32_t)(ew * 2) * 4096u;
    const uint32_t swz = (uint32_t)(lane & 7);
    constexpr int kChunks = BLOCK_N / 64;
    float st[kChunks][4];
#pragma unroll
    for (int c = 0; c < kChunks; ++c) st[c][0] = st[c][1] = st[c][2] = st[c][3] = 0.f;
    int cur_n = -1;
    // this warp's copy of the n block's bias row (fp32), refilled when the n block changes
    float* bias_row = reinterpret_cast<float*>(smem_raw + p.bias_off) + ew * 256;
    int bias_n = -1;
    auto flush_stats = [&](int n_blk) {
      const size_t part = (size_t)blockIdx.x * 4 + ew;
#pragma unroll
      for (int c = 0; c < kChunks; ++c) {
        const int k = n_blk * BLOCK_N + c * 64 + 2 * lane;
        if (k < p.K)
          *reinterpret_cast<float4*>(p.stats + (part * p.K + k) * 2) = make_float4(st[c][0], st[c][1], st[c][2], st[c][3]);
        st[c][0] = st[c][1] = st[c][2] = st[c][3] = 0.f;
      }
    };
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      const int m_tile = it % p.num_m, n_blk = it / p.num_m;
      if (p.stats && n_blk != cur_n) {
        if (cur_n >= 0) flush_stats(cur_n);
        cur_n = n_blk;
      }
      if (p.bias && n_blk != bias_n) {
        __syncwarp();
        for (int j = lane; j < BLOCK_N; j += 32) {
          const int k = n_blk * BLOCK_N + j;
          float b = 0.f;
          if (k < p.K)
            b = p.bias_dtype == TB200_F32 ? reinterpret_cast<const float*>(p.bias)[k]
                                          : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p.bias)[k]);
          bias_row[j] = b;
        }
        __syncwarp();
        bias_n = n_blk;
      }
      // this lane's row: flat padded position -> (image, row, column); junk positions get pix = -1
      const uint32_t q = (uint32_t)m_tile * 128u + (uint32_t)(ew * 32 + lane);
      bool valid;
      int pix;
      if (p.seg) {
        const uint32_t srow = fastdiv((uint32_t)m_tile, p.mSeg), sseg = (uint32_t)m_tile - srow * (uint32_t)p.nseg;
        const uint32_t wcol = sseg * 128u + (uint32_t)(ew * 32 + lane);
        valid = wcol < (uint32_t)p.W;
        pix = valid ? (int)(srow * (uint32_t)p.W + wcol) : -1;        // srow = image * H + row
      } else {
        const uint32_t rho = fastdiv(q, p.mWp), j = q - rho * (uint32_t)p.Wp;
        const uint32_t n = fastdiv(rho, p.mHp), h = rho - n * (uint32_t)p.Hp;
        valid = q < (uint32_t)p.total_pos && j < (uint32_t)p.W && h < (uint32_t)p.H;
        pix = valid ? (int)((n * (uint32_t)p.H + h) * (uint32_t)p.W + j) : -1;
      }
      const uint32_t vmask = __ballot_sync(0xffffffffu, valid);
      mbar_wait(t_full(acc), accph);
      tc_fence_after();
      const uint32_t tcol = tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(acc * BLOCK_N);
      int nchunks = (p.K - n_blk * BLOCK_N + 63) / 64;
      if (nchunks > kChunks) nchunks = kChunks;
#pragma unroll
      for (int c = 0; c < kChunks; ++c) {
        if (c < nchunks) {
          const int k0 = n_blk * BLOCK_N + c * 64;
          uint32_t v[64];
          tc_ld_32x32(tcol + c * 64, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
          tc_ld_32x32(tcol + c * 64 + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
          tc_ld_wait();
          if (c == nchunks - 1) {
            // the accumulator has been read: hand the TMEM stage back before the stores
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(t_empty(acc));
          }
          if (p.bias) {
#pragma unroll
            for (int u = 0; u < 16; ++u) {
              const float4 b4 = *reinterpret_cast<const float4*>(bias_row + c * 64 + 4 * u);
              v[4 * u] = __float_as_uint(__uint_as_float(v[4 * u]) + b4.x);
              v[4 * u + 1] = __float_as_uint(__uint_as_float(v[4 * u + 1]) + b4.y);
              v[4 * u + 2] = __float_as_uint(__uint_as_float(v[4 * u + 2]) + b4.z);
              v[4 * u + 3] = __float_as_uint(__uint_as_float(v[4 * u + 3]) + b4.w);
            }
          }
          if (p.relu) {
#pragma unroll
            for (int j = 0; j < 64; ++j) v[j] = __float_as_uint(fmaxf(__uint_as_float(v[j]), 0.f));
          }
          const uint32_t buf = bufs + (uint32_t)(c & 1) * 4096u;
          __syncwarp();                                    // readers of this staging tile (chunk c-2) are done
          const uint32_t row_addr = buf + (uint32_t)lane * 128u;
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            uint4 o;
            o.x = pack_bf16(__uint_as_float(v[8 * u]), __uint_as_float(v[8 * u + 1]));
            o.y = pack_bf16(__uint_as_float(v[8 * u + 2]), __uint_as_float(v[8 * u + 3]));
            o.z = pack_bf16(__uint_as_float(v[8 * u + 4]), __uint_as_float(v[8 * u + 5]));
            o.w = pack_bf16(__uint_as_float(v[8 * u + 6]), __uint_as_float(v[8 * u + 7]));
            st_shared_16(row_addr + (((uint32_t)u ^ swz) << 4), o);
          }
          __syncwarp();
          if (p.stats) {
            // lane l sums channels k0 + 2l, +1 over the valid rows: word l of every staged row
            const uint32_t col = (uint32_t)(lane >> 2), wsel = (uint32_t)(lane & 3) * 4u;
            // packed fp32x2 arithmetic (FADD2 / FFMA2); junk rows are zeroed by a select, not branched around
            float2 sum2 = make_float2(st[c][0], st[c][2]), sq2 = make_float2(st[c][1], st[c][3]);
#pragma unroll
            for (int rr = 0; rr < 32; ++rr) {
              uint32_t w = ld_shared_4(buf + (uint32_t)rr * 128u + ((col ^ (uint32_t)(rr & 7)) << 4) + wsel);
              w = ((vmask >> rr) & 1u) ? w : 0u;
              const float2 v = make_float2(bf16_lo(w), bf16_hi(w));
              sum2 = __fadd2_rn(sum2, v);
              sq2 = __ffma2_rn(v, v, sq2);
            }
            st[c][0] = sum2.x; st[c][2] = sum2.y; st[c][1] = sq2.x; st[c][3] = sq2.y;
          }
          // stores: lane -> (row (lane/8) + 4i, 16-byte unit lane%8)
          const bool unit_ok = k0 + (int)swz * 8 < p.K;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int rr = (lane >> 3) + 4 * i;
            const int pr = __shfl_sync(0xffffffffu, pix, rr);
            const uint4 o = ld_shared_16(buf + (uint32_t)rr * 128u + ((swz ^ (uint32_t)(rr & 7)) << 4));
            if (pr >= 0 && unit_ok)
              *reinterpret_cast<uint4*>(p.y + (size_t)pr * p.K + k0 + swz * 8) = o;
          }
        }
      }
      if (++acc == kC3Acc) { acc = 0; accph ^= 1; }
    }
    if (p.stats && cur_n >= 0) flush_stats(cur_n);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(kTmemCols) : "memory");
  }
}


// =========================================================================================================
// The same convolution on CTA pairs (cta_group::2) for 256 output channels per tile: a pair computes TWO consecutive
// position tiles (UMMA M = 256, each CTA its own 128 positions and its own halo stage) against ONE weight tile of
// which each CTA stages half (128 of the 256 output channels per tap and 64-channel block): the weight traffic
// per SM - the bound of the 256- and 512-channel layers, which re-fetch 2.3 MB of weights per 128-position tile
// (profiles/conv_r2.md) - halves. Plain per-CTA TMA loads; the MMA, its commit and the accumulator hand-back are the
// only cluster-wide operations. Both CTAs place their tile's first position at the SAME stage offset (2 * Wp rows:
// the rows are loaded shifted by up to Wp - 1 rows), because one MMA instruction carries one A descriptor.
template <bool DGRAD>
__global__ void __launch_bounds__(kC3Threads, 1)
conv3x3_pair_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w, const C3Params p) {
  constexpr int BLOCK_N = 256;
  constexpr int kTileBytes = 128 * 128;                   // this CTA's half of one (tap, 64-channel block) weight tile
  constexpr int kTmemCols = kC3Acc * BLOCK_N;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = smem_u32(smem_raw);
  const uint32_t a_smem = base;
  const uint32_t b_smem = base + (uint32_t)(p.a_stages * p.a_stage_bytes);
  const uint32_t epi_smem = base + (uint32_t)p.epi_off;
  const uint32_t bar = base + (uint32_t)p.bar_off;
  auto a_full = [&](int s) { return bar + 8u * s; };                         // leader's are the live ones
  auto a_empty = [&](int s) { return bar + 8u * (p.a_stages + s); };         // per CTA
  auto b_full = [&](int s) { return bar + 8u * (2 * p.a_stages + s); };      // leader's
  auto b_empty = [&](int s) { return bar + 8u * (2 * p.a_stages + p.b_groups + s); };
  const uint32_t t_base = bar + 8u * (2 * p.a_stages + 2 * p.b_groups);
  auto t_full = [&](int s) { return t_base + 8u * s; };                      // per CTA
  auto t_empty = [&](int s) { return t_base + 8u * (kC3Acc + s); };          // leader's
  const uint32_t tmem_slot = t_base + 8u * 2 * kC3Acc;

  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const uint32_t rank = cluster_ctarank();
  const int pair_id = blockIdx.x / 2, num_pairs = gridDim.x / 2;

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&map_x);
    prefetch_tensormap(&map_w);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < p.a_stages; ++s) { mbar_init(a_full(s), 1); mbar_init(a_empty(s), 1); }
    for (int s = 0; s < p.b_groups; ++s) { mbar_init(b_full(s), 1); mbar_init(b_empty(s), 1); }
    for (int s = 0; s < kC3Acc; ++s) { mbar_init(t_full(s), 1); mbar_init(t_empty(s), 8); }
    mbar_init_fence();
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - base));

  const int pm = (p.num_m + 1) / 2;                       // pairs of position tiles
  const int items = pm * p.num_n;
  const uint32_t row_bytes = (uint32_t)p.Wp * 128u;
  const uint32_t fixed_rows = p.seg ? (uint32_t)(p.Wp + 1) : (uint32_t)(2 * p.Wp);   // stage row of the tile's first position

  if (warp == 0) {
    // ================= producer (both CTAs) =================
    int as = 0, bg = 0;
    uint32_t aph = 0, bph = 0;
    for (int it = pair_id; it < items; it += num_pairs) {
      const int m_tile = 2 * (it % pm) + (int)rank, n_blk = it / pm;
      const uint32_t q0 = (uint32_t)m_tile * 128u;
      const int rho0 = (int)fastdiv(q0 + (uint32_t)p.Wp - 1u, p.mWp) - 2;
      const uint32_t woff = (uint32_t)((int)q0 - rho0 * p.Wp);               // in [Wp + 1, 2 * Wp]
      const uint32_t srow = fastdiv((uint32_t)m_tile, p.mSeg), sseg = (uint32_t)m_tile - srow * (uint32_t)p.nseg;
      const uint32_t sn = fastdiv(srow, p.mH), sh = srow - sn * (uint32_t)p.H;
      for (int cb = 0; cb < p.cblocks; ++cb) {
        mbar_wait(a_empty(as), aph ^ 1);
        if (elect_one()) {
          const uint32_t fb = mapa(a_full(as), 0);
          if (rank == 0) mbar_expect_tx(a_full(as), 2u * (uint32_t)p.NR * row_bytes);
          if (p.seg) {
            uint32_t dst = a_smem + (uint32_t)as * p.a_stage_bytes;
            for (int i = 0; i < 3; ++i) {
              tma_load_4d_pair(dst, &map_x, fb, cb * 64, (int)sseg * 128 - 1, (int)sh - 1 + i, (int)sn);
              dst += row_bytes;
            }
          } else {
            int n, hp;
            if (rho0 < 0) { n = -1; hp = rho0 + p.Hp; }
            else { n = (int)fastdiv((uint32_t)rho0, p.mHp); hp = rho0 - n * p.Hp; }
            // shifted so that the tile's first position sits 2 * Wp rows into the stage in BOTH CTAs
            uint32_t dst = a_smem + (uint32_t)as * p.a_stage_bytes + ((uint32_t)(2 * p.Wp) - woff) * 128u;
            for (int i = 0; i < p.NR; ++i) {
              tma_load_4d_pair(dst, &map_x, fb, cb * 64, 0, hp, n);
              dst += row_bytes;
              if (++hp == p.Hp) { hp = 0; ++n; }
            }
          }
        }
        __syncwarp();
        if (++as == p.a_stages) { as = 0; aph ^= 1; }
#pragma unroll 1
        for (int t = 0; t < 9; ++t) {
          mbar_wait(b_empty(bg), bph ^ 1);
          if (elect_one()) {
            const uint32_t fb = mapa(b_full(bg), 0);
            if (rank == 0) mbar_expect_tx(b_full(bg), 2u * kTileBytes);
            const uint32_t dst = b_smem + (uint32_t)bg * kTileBytes;
            if constexpr (!DGRAD) {
              // wt [k_out][9][c_red]: this CTA's 128 of the tile's 256 output channels (box {64 c, 128 k})
              tma_load_2d_pair(dst, &map_w, fb, t * p.C + cb * 64, n_blk * BLOCK_N + (int)rank * 128);
            } else {
              // forward weight wf [c_red][9][k_out]: two MN-major slabs {64 k_out, 64 c_red}
#pragma unroll
              for (int sl = 0; sl < 2; ++sl)
                tma_load_2d_pair(dst + sl * 8192, &map_w, fb, (8 - t) * p.K + n_blk * BLOCK_N + (int)rank * 128 + sl * 64,
                                 cb * 64);
            }
          }
          __syncwarp();
          if (++bg == p.b_groups) { bg = 0; bph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer (leader CTA) =================
    if (rank == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(256, BLOCK_N, 0, DGRAD ? 1 : 0);
      constexpr uint64_t kBDesc = DGRAD ? kDescMNMajor : kDescKMajor;
      constexpr uint32_t kBStep = DGRAD ? 128u : 2u;
      const int wp8 = p.Wp * 8;
      int as = 0, bg = 0, acc = 0;
      uint32_t aph = 0, bph = 0, accph = 0;
      for (int it = pair_id; it < items; it += num_pairs) {
        mbar_wait(t_empty(acc), accph ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)(acc * BLOCK_N);
        uint32_t accumulate = 0;
        for (int cb = 0; cb < p.cblocks; ++cb) {
          mbar_wait(a_full(as), aph);
          tc_fence_after();
          const uint32_t a_mid = desc_units(a_smem + (uint32_t)as * p.a_stage_bytes) + fixed_rows * 8u;
#pragma unroll
          for (int t = 0; t < 9; ++t) {
            mbar_wait(b_full(bg), bph);
            tc_fence_after();
            const int r = t / 3, s2 = t % 3;
            const uint64_t adesc = kDescKMajor | (uint64_t)(a_mid + (uint32_t)((r - 1) * wp8 + (s2 - 1) * 8));
            const uint64_t bdesc = kBDesc | (uint64_t)desc_units(b_smem + (uint32_t)bg * kTileBytes);
            if (elect_one()) {
#pragma unroll
              for (int k = 0; k < 4; ++k)
                tc_mma2_bf16(tmem_d, adesc + 2u * k, bdesc + kBStep * k, idesc, (k == 0) ? accumulate : 1u);
              tc_commit2(b_empty(bg));
            }
            __syncwarp();
            accumulate = 1;
            if (++bg == p.b_groups) { bg = 0; bph ^= 1; }
          }
          if (elect_one()) tc_commit2(a_empty(as));
          __syncwarp();
          if (++as == p.a_stages) { as = 0; aph ^= 1; }
        }
        if (elect_one()) tc_commit2(t_full(acc));
        __syncwarp();
        if (++acc == kC3Acc) { acc = 0; accph ^= 1; }
      }
    }
  } else if (warp >= 4) {
    // ================= epilogue =================
    // Warp ew owns tile rows [32*ew, +32); lane = row for the TMEM loads. Per 64 output channels:
    // TMEM -> registers -> bf16 -> swizzled staging tile (32 rows x 128 B) -> (statistics from the staged
    // bf16 values) -> coalesced 16-byte global stores of the valid rows (4 rows x 128 B per instruction).
    const int ew = warp - 4;
    int acc = 0;
    uint32_t accph = 0;
    const uint32_t bufs = epi_smem + (uint32_t)(ew * 2) * 4096u;
    const uint32_t swz = (uint32_t)(lane & 7);
    constexpr int kChunks = BLOCK_N / 64;
    float st[kChunks][4];
#pragma unroll
    for (int c = 0; c < kChunks; ++c) st[c][0] = st[c][1] = st[c][2] = st[c][3] = 0.f;
    int cur_n = -1;
    // this warp's copy of the n block's bias row (fp32), refilled when the n block changes
    float* bias_row = reinterpret_cast<float*>(smem_raw + p.bias_off) + ew * 256;
    int bias_n = -1;
    auto flush_stats = [&](int n_blk) {
      const size_t part = (size_t)blockIdx.x * 4 + ew;
#pragma unroll
      for (int c = 0; c < kChunks; ++c) {
        const int k = n_blk * BLOCK_N + c * 64 + 2 * lane;
        if (k < p.K)
          *reinterpret_cast<float4*>(p.stats + (part * p.K + k) * 2) = make_float4(st[c][0], st[c][1], st[c][2], st[c][3]);
        st[c][0] = st[c][1] = st[c][2] = st[c][3] = 0.f;
      }
    };
    for (int it = pair_id; it < items; it += num_pairs) {
      const int m_tile = 2 * (it % pm) + (int)rank, n_blk = it / pm;
      if (p.stats && n_blk != cur_n) {
        if (cur_n >= 0) flush_stats(cur_n);
        cur_n = n_blk;
      }
      if (p.bias && n_blk != bias_n) {
        __syncwarp();
        for (int j = lane; j < BLOCK_N; j += 32) {
          const int k = n_blk * BLOCK_N + j;
          float b = 0.f;
          if (k < p.K)
            b = p.bias_dtype == TB200_F32 ? reinterpret_cast<const float*>(p.bias)[k]
                                          : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p.bias)[k]);
          bias_row[j] = b;
        }
        __syncwarp();
        bias_n = n_blk;
      }
      // this lane's row: flat padded position -> (image, row, column); junk positions get pix = -1
      const uint32_t q = (uint32_t)m_tile * 128u + (uint32_t)(ew * 32 + lane);
      bool valid;
      int pix;
      if (p.seg) {
        const uint32_t srow = fastdiv((uint32_t)m_tile, p.mSeg), sseg = (uint32_t)m_tile - srow * (uint32_t)p.nseg;
        const uint32_t wcol = sseg * 128u + (uint32_t)(ew * 32 + lane);
        valid = wcol < (uint32_t)p.W && m_tile < p.num_m;     // (an odd tile count leaves the last pair's second tile empty)
        pix = valid ? (int)(srow * (uint32_t)p.W + wcol) : -1;        // srow = image * H + row
      } else {
        const uint32_t rho = fastdiv(q, p.mWp), j = q - rho * (uint32_t)p.Wp;
        const uint32_t n = fastdiv(rho, p.mHp), h = rho - n * (uint32_t)p.Hp;
        valid = q < (uint32_t)p.total_pos && j < (uint32_t)p.W && h < (uint32_t)p.H;
        pix = valid ? (int)((n * (uint32_t)p.H + h) * (uint32_t)p.W + j) : -1;
      }
      const uint32_t vmask = __ballot_sync(0xffffffffu, valid);
      mbar_wait(t_full(acc), accph);
      tc_fence_after();
      const uint32_t tcol = tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(acc * BLOCK_N);
      int nchunks = (p.K - n_blk * BLOCK_N + 63) / 64;
      if (nchunks > kChunks) nchunks = kChunks;
#pragma unroll
      for (int c = 0; c < kChunks; ++c) {
        if (c < nchunks) {
          const int k0 = n_blk * BLOCK_N + c * 64;
          uint32_t v[64];
          tc_ld_32x32(tcol + c * 64, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
          tc_ld_32x32(tcol + c * 64 + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
          tc_ld_wait();
          if (c == nchunks - 1) {
            // the accumulator has been read: hand the TMEM stage back before the stores
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(mapa(t_empty(acc), 0));
          }
          if (p.bias) {
#pragma unroll
            for (int u = 0; u < 16; ++u) {
              const float4 b4 = *reinterpret_cast<const float4*>(bias_row + c * 64 + 4 * u);
              v[4 * u] = __float_as_uint(__uint_as_float(v[4 * u]) + b4.x);
              v[4 * u + 1] = __float_as_uint(__uint_as_float(v[4 * u + 1]) + b4.y);
              v[4 * u + 2] = __float_as_uint(__uint_as_float(v[4 * u + 2]) + b4.z);
              v[4 * u + 3] = __float_as_uint(__uint_as_float(v[4 * u + 3]) + b4.w);
            }
          }
          if (p.relu) {
#pragma unroll
            for (int j = 0; j < 64; ++j) v[j] = __float_as_uint(fmaxf(__uint_as_float(v[j]), 0.f));
          }
          const uint32_t buf = bufs + (uint32_t)(c & 1) * 4096u;
          __syncwarp();                                    // readers of this staging tile (chunk c-2) are done
          const uint32_t row_addr = buf + (uint32_t)lane * 128u;
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            uint4 o;
            o.x = pack_bf16(__uint_as_float(v[8 * u]), __uint_as_float(v[8 * u + 1]));
            o.y = pack_bf16(__uint_as_float(v[8 * u + 2]), __uint_as_float(v[8 * u + 3]));
            o.z = pack_bf16(__uint_as_float(v[8 * u + 4]), __uint_as_float(v[8 * u + 5]));
            o.w = pack_bf16(__uint_as_float(v[8 * u + 6]), __uint_as_float(v[8 * u + 7]));
            st_shared_16(row_addr + (((uint32_t)u ^ swz) << 4), o);
          }
          __syncwarp();
          if (p.stats) {
            // lane l sums channels k0 + 2l, +1 over the valid rows: word l of every staged row
            const uint32_t col = (uint32_t)(lane >> 2), wsel = (uint32_t)(lane & 3) * 4u;
            // packed fp32x2 arithmetic (FADD2 / FFMA2); junk rows are zeroed by a select, not branched around
            float2 sum2 = make_float2(st[c][0], st[c][2]), sq2 = make_float2(st[c][1], st[c][3]);
#pragma unroll
            for (int rr = 0; rr < 32; ++rr) {
              uint32_t w = ld_shared_4(buf + (uint32_t)rr * 128u + ((col ^ (uint32_t)(rr & 7)) << 4) + wsel);
              w = ((vmask >> rr) & 1u) ? w : 0u;
              const float2 v = make_float2(bf16_lo(w), bf16_hi(w));
              sum2 = __fadd2_rn(sum2, v);
              sq2 = __ffma2_rn(v, v, sq2);
            }
            st[c][0] = sum2.x; st[c][2] = sum2.y; st[c][1] = sq2.x; st[c][3] = sq2.y;
          }
          // stores: lane -> (row (lane/8) + 4i, 16-byte unit lane%8)
          const bool unit_ok = k0 + (int)swz * 8 < p.K;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int rr = (lane >> 3) + 4 * i;
            const int pr = __shfl_sync(0xffffffffu, pix, rr);
            const uint4 o = ld_shared_16(buf + (uint32_t)rr * 128u + ((swz ^ (uint32_t)(rr & 7)) << 4));
            if (pr >= 0 && unit_ok)
              *reinterpret_cast<uint4*>(p.y + (size_t)pr * p.K + k0 + swz * 8) = o;
          }
        }
      }
      if (++acc == kC3Acc) { acc = 0; accph ^= 1; }
    }
    if (p.stats && cur_n >= 0) flush_stats(cur_n);
  }

  __syncwarp();
  tc_fence_before();
  cluster_sync_all();
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(kTmemCols) : "memory");
  }
}

// =========================================================================================================
// Weight gradient of the same convolution, also as a halo-tile implicit GEMM:
//   dW[co, r, s, ci] = sum over positions q of dy[q, co] * x[q + (r-1)*(W+1) + (s-1), ci]
// in the flat padded space used above (dy is zero at the shared pad column / row because its loads are
// zero-filled there too). The reduction runs over POSITIONS, so both operands are MN-major as they land:
// a stage holds, for a super block of 128 positions, the padded rows of dy that cover it (A: M = 64-channel
// slabs of co, the slab stride goes into the descriptor's leading-dimension field) and the padded rows of x that
// cover it plus the halo (B: N = 64 channels of ci); a filter tap is the B descriptor advanced by its shift in
// rows - the same trick as the forward pass, along the reduction dimension this time (probed on B200:
// profiles/probe_shift_r2a.json, "B MN-major"). The im2col-TMA wgrad fetched x once per tap and dy once per
// 128-column tile of the 9*Cin-wide output (L2 -> SM traffic 10-18x the tensors, profiles/conv_r2.md); here a
// work item is (128 output channels, 64 input channels, HALF of the taps - nine 64-column accumulators exceed the
// 512 TMEM columns - , a split of the positions) and reads x (+ halo) and dy once.
// Output: fp32 partial planes [splits][Cout][9*Cin], summed and rounded by wgrad_reduce_kernel.
struct W3Params {
  int N, H, W, Cin, Cout;
  int Wp, Hp, total_pos;
  unsigned long long mWp;
  unsigned long long mHp;
  int NRx, NRy;                    // padded rows per stage of x / dy
  int x_bytes, y_slab_bytes, stage_bytes, stages;
  int m_slabs;                     // 64-channel slabs of dy per stage: 1 (Cout == 64) or 2
  int cblocks, mblocks, splits, num_sb;
  int bar_off;
  float* partials;
};

constexpr int kW3Threads = 256;
constexpr int kW3TmemCols = 512;   // five taps x 64 columns, rounded up to a power of two

__global__ void __launch_bounds__(kW3Threads, 1)
conv3x3_wgrad_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_dy,
                     const W3Params p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = smem_u32(smem_raw);
  const uint32_t bar = base + (uint32_t)p.bar_off;
  auto full = [&](int s) { return bar + 8u * s; };
  auto empty = [&](int s) { return bar + 8u * (p.stages + s); };
  const uint32_t t_full = bar + 8u * (2 * p.stages), t_empty = t_full + 8u;
  const uint32_t tmem_slot = t_empty + 8u;
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&map_x);
    prefetch_tensormap(&map_dy);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < p.stages; ++s) { mbar_init(full(s), 1); mbar_init(empty(s), 1); }
    mbar_init(t_full, 1);
    mbar_init(t_empty, 4);
    mbar_init_fence();
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(kW3TmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - base));

  const int per_split = 2 * p.cblocks * p.mblocks;
  const int items = per_split * p.splits;
  const uint32_t row_bytes = (uint32_t)p.Wp * 128u;
  auto decode = [&](int it, int* tg, int* cb, int* mb, int* sb0, int* sb1) {
    const int ks = it / per_split, r = it - ks * per_split;
    *tg = r & 1;
    *cb = (r >> 1) % p.cblocks;
    *mb = (r >> 1) / p.cblocks;
    *sb0 = (int)((long long)ks * p.num_sb / p.splits);
    *sb1 = (int)((long long)(ks + 1) * p.num_sb / p.splits);
    return ks;
  };

  if (warp == 0) {
    // ================= producer =================
    int st = 0;
    uint32_t ph = 0;
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      int tg, cb, mb, sb0, sb1;
      decode(it, &tg, &cb, &mb, &sb0, &sb1);
      for (int sb = sb0; sb < sb1; ++sb) {
        const uint32_t q0 = (uint32_t)sb * 128u;
        mbar_wait(empty(st), ph ^ 1);
        if (elect_one()) {
          mbar_expect_tx(full(st), (uint32_t)(p.NRx + p.m_slabs * p.NRy) * row_bytes);
          // x: padded rows from the one holding q0 - Wp - 1 on (row -1 / -2 of the first block: zero-filled)
          const int rho0 = (int)fastdiv(q0 + (uint32_t)p.Wp - 1u, p.mWp) - 2;
          int n, hp;
          if (rho0 < 0) { n = -1; hp = rho0 + p.Hp; }
          else { n = (int)fastdiv((uint32_t)rho0, p.mHp); hp = rho0 - n * p.Hp; }
          uint32_t dst = base + (uint32_t)st * p.stage_bytes;
          for (int i = 0; i < p.NRx; ++i) {
            tma_load_4d(dst, &map_x, full(st), cb * 64, 0, hp, n);
            dst += row_bytes;
            if (++hp == p.Hp) { hp = 0; ++n; }
          }
          // dy: padded rows from the one holding q0 on, one slab per 64 output channels
          const int rhoy = (int)fastdiv(q0, p.mWp);
          const int ny0 = (int)fastdiv((uint32_t)rhoy, p.mHp), hy0 = rhoy - ny0 * p.Hp;
          for (int sl = 0; sl < p.m_slabs; ++sl) {
            dst = base + (uint32_t)st * p.stage_bytes + (uint32_t)p.x_bytes + (uint32_t)sl * p.y_slab_bytes;
            n = ny0; hp = hy0;
            for (int i = 0; i < p.NRy; ++i) {
              tma_load_4d(dst, &map_dy, full(st), mb * 128 + sl * 64, 0, hp, n);
              dst += row_bytes;
              if (++hp == p.Hp) { hp = 0; ++n; }
            }
          }
        }
        __syncwarp();
        if (++st == p.stages) { st = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    constexpr uint32_t idesc = make_idesc_bf16(128, 64, 1, 1);
    // A: MN-major, 64-wide slabs y_slab_bytes apart, 8 reduction rows per 1024-byte atom
    const uint64_t a_const = ((uint64_t)((uint32_t)p.y_slab_bytes >> 4) << 16) | ((uint64_t)(1024 >> 4) << 32) |
                             ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
    const int wp8 = p.Wp * 8;
    int st = 0;
    uint32_t ph = 0, tph = 0;
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      int tg, cb, mb, sb0, sb1;
      decode(it, &tg, &cb, &mb, &sb0, &sb1);
      mbar_wait(t_empty, tph ^ 1);
      tc_fence_after();
      uint32_t accumulate = 0;
      for (int sb = sb0; sb < sb1; ++sb) {
        const uint32_t q0 = (uint32_t)sb * 128u;
        const int rho0 = (int)fastdiv(q0 + (uint32_t)p.Wp - 1u, p.mWp) - 2;
        const int rhoy = (int)fastdiv(q0, p.mWp);
        mbar_wait(full(st), ph);
        tc_fence_after();
        const uint32_t stage = base + (uint32_t)st * p.stage_bytes;
        // windows: x at tap (1, 1) = the block's own positions; dy = the block's own positions
        const uint32_t b_mid = desc_units(stage) + (uint32_t)((int)q0 - rho0 * p.Wp) * 8u;
        const uint32_t a_units = desc_units(stage + (uint32_t)p.x_bytes) + (uint32_t)((int)q0 - rhoy * p.Wp) * 8u;
        const uint64_t adesc = a_const | (uint64_t)a_units;
#pragma unroll
        for (int tt = 0; tt < 5; ++tt) {
          if (tt < 5 - tg) {                           // taps 0..4 (tg 0) or 5..8 (tg 1)
            const int t = tg * 5 + tt, r = t / 3, s2 = t - 3 * r;
            const uint64_t bdesc = kDescMNMajor | (uint64_t)(b_mid + (uint32_t)((r - 1) * wp8 + (s2 - 1) * 8));
            if (elect_one()) {
#pragma unroll
              for (int k = 0; k < 8; ++k)              // 8 x 16 positions; 16 rows = two 1024-byte atoms = 128 units
                tc_mma_bf16(tmem_base + (uint32_t)(tt * 64), adesc + 128u * k, bdesc + 128u * k, idesc,
                            (k == 0) ? accumulate : 1u);
            }
            __syncwarp();
          }
        }
        accumulate = 1;
        if (elect_one()) tc_commit(empty(st));
        __syncwarp();
        if (++st == p.stages) { st = 0; ph ^= 1; }
      }
      if (elect_one()) tc_commit(t_full);
      __syncwarp();
      tph ^= 1;
    }
  } else if (warp >= 4) {
    // ================= epilogue: TMEM -> fp32 partial plane of this split =================
    const int ew = warp - 4;
    uint32_t tph = 0;
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      int tg, cb, mb, sb0, sb1;
      const int ks = decode(it, &tg, &cb, &mb, &sb0, &sb1);
      const int co = mb * 128 + ew * 32 + lane;
      mbar_wait(t_full, tph);
      tc_fence_after();
      const uint32_t tcol = tmem_base + ((uint32_t)(ew * 32) << 16);
      float* out_row = p.partials + ((size_t)ks * p.Cout + (size_t)(co < p.Cout ? co : 0)) * (size_t)(9 * p.Cin) + cb * 64;
#pragma unroll 1
      for (int tt = 0; tt < 5 - tg; ++tt) {
        uint32_t v[64];
        tc_ld_32x32(tcol + tt * 64, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
        tc_ld_32x32(tcol + tt * 64 + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
        tc_ld_wait();
        if (co < p.Cout && sb1 > sb0) {
          float4* o = reinterpret_cast<float4*>(out_row + (size_t)(tg * 5 + tt) * p.Cin);
#pragma unroll
          for (int u = 0; u < 16; ++u)
            o[u] = make_float4(__uint_as_float(v[4 * u]), __uint_as_float(v[4 * u + 1]), __uint_as_float(v[4 * u + 2]),
                               __uint_as_float(v[4 * u + 3]));
        } else if (co < p.Cout) {
          float4* o = reinterpret_cast<float4*>(out_row + (size_t)(tg * 5 + tt) * p.Cin);
#pragma unroll
          for (int u = 0; u < 16; ++u) o[u] = make_float4(0.f, 0.f, 0.f, 0.f);   // a split without positions
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(t_empty);
      tph ^= 1;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(kW3TmemCols) : "memory");
  }
}

// dW (bf16, [Cout][9*Cin]) = sum over the split planes, in split order (deterministic)
__global__ void __launch_bounds__(256) wgrad_reduce_kernel(const float* __restrict__ partials, int splits, size_t plane,
                                                          __nv_bfloat16* __restrict__ dw) {
  const size_t i = ((size_t)blockIdx.x * 256 + threadIdx.x) * 4;
  if (i >= plane) return;
  float4 a = *reinterpret_cast<const float4*>(partials + i);
  for (int s = 1; s < splits; ++s) {
    const float4 b = *reinterpret_cast<const float4*>(partials + (size_t)s * plane + i);
    a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
  }
  uint2 o;
  o.x = pack_bf16(a.x, a.y);
  o.y = pack_bf16(a.z, a.w);
  *reinterpret_cast<uint2*>(dw + i) = o;
}

// ---- host ---------------------------------------------------------------------------------------------
typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                             const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                             CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeFn c3_encode() {
  static EncodeFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeFn>(p);
  });
  return fn;
}

// NHWC bf16 tensor [n,h,w,c] as a 4-D map {c, w, h, n}; box = {64 channels, box_w pixels, 1, 1}.
int c3_map_nhwc(CUtensorMap* out, const void* base, int n, int h, int w, int c, int box_w) {
  EncodeFn enc = c3_encode();
  if (!enc) return fail("cuTensorMapEncodeTiled entry point not available%s");
  cuuint64_t dims[4] = {(cuuint64_t)c, (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)n};
  cuuint64_t strides[3] = {(cuuint64_t)c * 2, (cuuint64_t)w * c * 2, (cuuint64_t)h * w * c * 2};
  cuuint32_t box[4] = {64, (cuuint32_t)box_w, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail("cuTensorMapEncodeTiled (NHWC) failed (%s code %ld)", "", (long)r);
  return 0;
}

// row-major bf16 [rows, cols], box = {64 cols, box_rows}
int c3_map_2d(CUtensorMap* out, const void* base, int rows, int cols, int box_rows) {
  EncodeFn enc = c3_encode();
  if (!enc) return fail("cuTensorMapEncodeTiled entry point not available%s");
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)cols * 2};
  cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail("cuTensorMapEncodeTiled (weights) failed (%s code %ld)", "", (long)r);
  return 0;
}

template <int BLOCK_N, int TPG, bool DGRAD>
int launch_c3(const CUtensorMap& mx, const CUtensorMap& mw, C3Params p, int* nparts_out, cudaStream_t stream,
              bool dry_run = false) {
  constexpr int kSmemMax = 227 * 1024;
  constexpr int kTile = BLOCK_N * 128;
  const int row_bytes = p.Wp * 128;
  p.a_stage_bytes = (p.NR * row_bytes + 1023) / 1024 * 1024;
  const int fixed = kEpiBytes + 1024 + 4096;          // staging tiles, barriers, the epilogue warps' bias rows
  const int resident_bytes = 9 * p.cblocks * kTile;
  p.resident = (p.num_n == 1 && resident_bytes + 2 * p.a_stage_bytes + fixed <= kSmemMax) ? 1 : 0;
  int b_bytes;
  if (p.resident) {
    p.b_groups = 1;
    b_bytes = resident_bytes;
    p.a_stages = (kSmemMax - fixed - b_bytes) / p.a_stage_bytes;
  } else {
    // at least two A stages, then as many weight groups as fit (2..4), then a third A stage if there is room
    p.a_stages = 2;
    p.b_groups = (kSmemMax - fixed - 2 * p.a_stage_bytes) / (TPG * kTile);
    const int max_groups = TPG == 1 ? 6 : 3;
    if (p.b_groups > max_groups) p.b_groups = max_groups;
    if (p.b_groups < 2) return fail("tb200_conv3x3: tile does not fit shared memory (W %s%ld)", "", p.W);
    b_bytes = p.b_groups * TPG * kTile;
    p.a_stages = (kSmemMax - fixed - b_bytes) / p.a_stage_bytes;
  }
  if (p.a_stages > 3) p.a_stages = 3;
  if (p.a_stages < 2) return fail("tb200_conv3x3: tile does not fit shared memory (W %s%ld)", "", p.W);
  p.epi_off = p.a_stages * p.a_stage_bytes + b_bytes;
  p.bar_off = p.epi_off + kEpiBytes;
  p.bias_off = p.bar_off + 1024;
  const int total = p.bias_off + 4096;
  if (dry_run) return 0;
  auto kern = conv3x3_kernel<BLOCK_N, TPG, DGRAD>;
  static DeviceOnce configured;
  if (int rc = ensure_dynamic_smem(configured, kern, kSmemMax)) return rc;
  const int sms = device_sm_count();
  const int items = p.num_m * p.num_n;
  const int grid = items < sms ? items : sms;
  if (p.stats) {
    // CTAs that see no tile of some n block leave its slots untouched: start from zero
    TB_CUDA_OK(cudaMemsetAsync(p.stats, 0, (size_t)grid * 4 * p.K * 2 * sizeof(float), stream));
    if (nparts_out) *nparts_out = grid * 4;
  }
  kern<<<grid, kC3Threads, total, stream>>>(mx, mw, p);
  return launched();
}

// SM count, or `dflt` where no device is present (geometry queries on a CPU-only host)
int device_sm_count_or(int dflt) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return dflt; }
  return device_sm_count();
}

// TB200_CONV_PAIR=0 keeps the 256-wide tiles on the one-CTA kernel (A/B runs)
bool conv_pair_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("TB200_CONV_PAIR");
    v = (e && atoi(e) == 0) ? 0 : 1;
  }
  return v == 1;
}

template <bool DGRAD>
int launch_c3_pair(const CUtensorMap& mx, const CUtensorMap& mw, C3Params p, int* nparts_out, cudaStream_t stream,
                   bool dry_run) {
  constexpr int kSmemMax = 227 * 1024;
  constexpr int kTile = 128 * 128;
  const int row_bytes = p.Wp * 128;
  const int rows = p.seg ? 3 : p.NR + 1;                 // flat mode: the rows land shifted by up to Wp - 1 rows
  p.a_stage_bytes = (rows * row_bytes + 1023) / 1024 * 1024;
  const int fixed = kEpiBytes + 1024 + 4096;
  p.resident = 0;
  p.a_stages = 2;
  p.b_groups = (kSmemMax - fixed - 2 * p.a_stage_bytes) / kTile;
  if (p.b_groups > 8) p.b_groups = 8;
  if (p.b_groups < 3) return fail("tb200_conv3x3 (pair): tile does not fit shared memory (W %s%ld)", "", p.W);
  const int b_bytes = p.b_groups * kTile;
  p.a_stages = (kSmemMax - fixed - b_bytes) / p.a_stage_bytes;
  if (p.a_stages > 3) p.a_stages = 3;
  p.epi_off = p.a_stages * p.a_stage_bytes + b_bytes;
  p.bar_off = p.epi_off + kEpiBytes;
  p.bias_off = p.bar_off + 1024;
  const int total = p.bias_off + 4096;
  if (dry_run) return 0;
  auto kern = conv3x3_pair_kernel<DGRAD>;
  static DeviceOnce configured;
  if (int rc = ensure_dynamic_smem(configured, kern, kSmemMax)) return rc;
  const int sms = device_sm_count();
  const int items = ((p.num_m + 1) / 2) * p.num_n;
  const int pairs = items < sms / 2 ? items : sms / 2;
  if (p.stats) {
    TB_CUDA_OK(cudaMemsetAsync(p.stats, 0, (size_t)(2 * pairs) * 4 * p.K * 2 * sizeof(float), stream));
    if (nparts_out) *nparts_out = 2 * pairs * 4;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(2 * pairs);
  cfg.blockDim = dim3(kC3Threads);
  cfg.dynamicSmemBytes = (size_t)total;
  cfg.stream = stream;
  cudaLaunchAttribute attr{};
  attr.id = cudaLaunchAttributeClusterDimension;
  attr.val.clusterDim.x = 2; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = 1;
  cfg.attrs = &attr;
  cfg.numAttrs = 1;
  TB_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, mx, mw, p));
  return launched();
}

int conv3x3_common(const void* x, const void* w, void* y, int n, int h, int wd, int c_red, int k_out, bool dgrad,
                   float* stats, size_t stats_floats, int* nparts_out, cudaStream_t stream, const void* bias = nullptr,
                   int bias_dtype = 0, int relu = 0, bool dry_run = false) {
  const char* who = dgrad ? "tb200_conv3x3_dgrad" : "tb200_conv3x3_fprop";
  if (n <= 0 || h <= 0 || wd <= 0 || c_red <= 0 || k_out <= 0) return fail("%s: empty problem", who);
  if (c_red % 64) return fail("%s: the reduction channel count (%ld) must be a multiple of 64", who, c_red);
  if (k_out % 8) return fail("%s: the output channel count (%ld) must be a multiple of 8", who, k_out);
  static thread_local bool ctx_bound = false;
  if (!ctx_bound && !dry_run) {
    TB_CUDA_OK(cudaFree(nullptr));       // binds the primary context to this thread for the driver-API encode
    ctx_bound = true;
  }
  C3Params p{};
  p.N = n; p.H = h; p.W = wd; p.C = c_red; p.K = k_out;
  p.Wp = wd + 1; p.Hp = h + 1;
  const long long total = (long long)n * p.Hp * p.Wp;
  if (total + 4 * p.Wp >= (1ll << 24) || p.Wp >= 4096 || p.Hp >= 4096)
    return fail("%s: problem too large for the position arithmetic (%ld positions)", who, (long)total);
  p.total_pos = (int)total;
  p.mWp = fastdiv_magic((unsigned)p.Wp);
  p.mHp = fastdiv_magic((unsigned)p.Hp);
  p.NR = (3 * p.Wp + 129 + p.Wp - 1) / p.Wp;
  p.cblocks = c_red / 64;
  p.num_m = (int)((total + 127) / 128);
  p.nseg = 1; p.mSeg = fastdiv_magic(1u); p.mH = fastdiv_magic((unsigned)h);
  // Wide images: whole padded rows no longer fit two stages (W = 112: 72 KB per stage next to the weight ring), so
  // a tile becomes a 128-column segment of one row with a three-row, 130-column stage (50 KB); costs the unused
  // columns of the last segment (12.5 % at W = 112 and 224) instead of the flat layout's 1-3 %.
  if (wd >= 100) {
    p.seg = 1;
    p.nseg = (wd + 127) / 128;
    p.mSeg = fastdiv_magic((unsigned)p.nseg);
    p.Wp = 130;
    p.NR = 3;
    p.mWp = fastdiv_magic(130u);
    const long long tiles = (long long)n * h * p.nseg;
    if (tiles >= (1ll << 24)) return fail("%s: problem too large (%ld tiles)", who, (long)tiles);
    p.num_m = (int)tiles;
  } else if (wd + 1 > 256) {
    return fail("%s: image width %ld exceeds one TMA box (255)", who, wd);
  }
  p.y = reinterpret_cast<__nv_bfloat16*>(y);
  p.stats = stats;
  p.bias = bias; p.bias_dtype = bias_dtype; p.relu = relu;
  if ((bias || relu) && (dgrad || stats)) return fail("%s: bias / ReLU belong to the plain forward pass", who);
  const int block_n = k_out <= 64 ? 64 : (k_out <= 128 ? 128 : 256);
  p.num_n = (k_out + block_n - 1) / block_n;
  if (stats) {
    const int sms = device_sm_count();
    const long long items = (long long)p.num_m * p.num_n;
    const size_t need = (size_t)(items < sms ? items : sms) * 4 * k_out * 2;
    if (stats_floats < need)
      return fail("%s: statistics buffer too small (%ld floats needed)", who, (long)need);
  }
  CUtensorMap mx, mw;
  memset(&mx, 0, sizeof(mx));
  memset(&mw, 0, sizeof(mw));
  if (!dry_run)
    if (int rc = c3_map_nhwc(&mx, x, n, h, wd, c_red, p.Wp)) return rc;      // box: one stage row
  if (!dgrad) {
    // wt [k_out][9][c_red] as [k_out rows, 9*c_red cols]
    const bool pair = block_n == 256 && conv_pair_enabled() && device_sm_count_or(148) >= 2 &&
                      launch_c3_pair<false>(mx, mw, p, nparts_out, stream, true) == 0;
    if (!dry_run)
      if (int rc = c3_map_2d(&mw, w, k_out, 9 * c_red, pair ? 128 : block_n)) return rc;
    if (pair) return launch_c3_pair<false>(mx, mw, p, nparts_out, stream, dry_run);
    if (block_n == 64) return launch_c3<64, 3, false>(mx, mw, p, nparts_out, stream, dry_run);
    if (block_n == 128) {
      // three taps per weight group (48 KB) when two groups fit next to the input stages, else one tap per group
      if (launch_c3<128, 3, false>(mx, mw, p, nparts_out, stream, true) == 0)
        return launch_c3<128, 3, false>(mx, mw, p, nparts_out, stream, dry_run);
      return launch_c3<128, 1, false>(mx, mw, p, nparts_out, stream, dry_run);
    }
    return launch_c3<256, 1, false>(mx, mw, p, nparts_out, stream, dry_run);
  }
  // forward weight wf [c_red][9][k_out] as [c_red rows, 9*k_out cols]; boxes of 64 x 64
  if (!dry_run)
    if (int rc = c3_map_2d(&mw, w, c_red, 9 * k_out, 64)) return rc;
  if (block_n == 256 && conv_pair_enabled() && device_sm_count_or(148) >= 2 &&
      launch_c3_pair<true>(mx, mw, p, nparts_out, stream, true) == 0)
    return launch_c3_pair<true>(mx, mw, p, nparts_out, stream, dry_run);
  if (block_n == 64) return launch_c3<64, 3, true>(mx, mw, p, nparts_out, stream, dry_run);
  if (block_n == 128) {
    if (launch_c3<128, 3, true>(mx, mw, p, nparts_out, stream, true) == 0)
      return launch_c3<128, 3, true>(mx, mw, p, nparts_out, stream, dry_run);
    return launch_c3<128, 1, true>(mx, mw, p, nparts_out, stream, dry_run);
  }
  return launch_c3<256, 1, true>(mx, mw, p, nparts_out, stream, dry_run);
}


// Geometry of the wgrad kernel for a problem: fills p (everything but the pointers), returns 0 / error.
int w3_plan(W3Params& p, int n, int h, int wd, int cin, int cout) {
  const char* who = "tb200_conv3x3_wgrad";
  if (n <= 0 || h <= 0 || wd <= 0 || cin <= 0 || cout <= 0) return fail("%s: empty problem", who);
  if (cin % 64 || cout % 64) return fail("%s: channel counts must be multiples of 64", who);
  if (wd + 1 > 256) return fail("%s: image width %ld exceeds one TMA box (255)", who, wd);
  p.N = n; p.H = h; p.W = wd; p.Cin = cin; p.Cout = cout;
  p.Wp = wd + 1; p.Hp = h + 1;
  const long long total = (long long)n * p.Hp * p.Wp;
  if (total + 4 * p.Wp >= (1ll << 24) || p.Wp >= 4096 || p.Hp >= 4096)
    return fail("%s: problem too large for the position arithmetic (%ld positions)", who, (long)total);
  p.total_pos = (int)total;
  p.mWp = fastdiv_magic((unsigned)p.Wp);
  p.mHp = fastdiv_magic((unsigned)p.Hp);
  p.NRx = (3 * p.Wp + 129 + p.Wp - 1) / p.Wp;
  p.NRy = (p.Wp + 127 + p.Wp - 1) / p.Wp;
  const int row_bytes = p.Wp * 128;
  p.x_bytes = (p.NRx * row_bytes + 1023) / 1024 * 1024;
  p.y_slab_bytes = (p.NRy * row_bytes + 1023) / 1024 * 1024;
  p.m_slabs = cout == 64 ? 1 : 2;
  if (cout % 128 && cout != 64) return fail("%s: Cout (%ld) must be 64 or a multiple of 128", who, cout);
  // Cout == 64: the MMA still spans 128 rows; its second slab aliases the first one 8 rows further (rows 64-127
  // of the accumulator are never read), so the stage reserves those 1024 bytes
  p.stage_bytes = p.x_bytes + p.m_slabs * p.y_slab_bytes + (p.m_slabs == 1 ? 1024 : 0);
  constexpr int kSmemMax = 227 * 1024;
  p.stages = (kSmemMax - 1024) / p.stage_bytes;
  if (p.stages > 4) p.stages = 4;
  if (p.stages < 2) return fail("%s: a stage does not fit shared memory (W %ld)", who, wd);
  p.bar_off = p.stages * p.stage_bytes;
  p.cblocks = cin / 64;
  p.mblocks = (cout + 127) / 128;
  p.num_sb = (int)((total + 127) / 128);
  const int sms = device_sm_count();
  const int per_split = 2 * p.cblocks * p.mblocks;
  p.splits = sms / per_split;
  if (p.splits < 1) p.splits = 1;
  if (p.splits > p.num_sb) p.splits = p.num_sb;
  return 0;
}

}  // namespace

extern "C" {

// fp32 scratch (bytes) tb200_conv3x3_wgrad needs for this problem on the current device; 0 = not eligible
size_t tb200_conv3x3_wgrad_workspace(int n, int h, int wd, int cin, int cout) {
  W3Params p{};
  if (w3_plan(p, n, h, wd, cin, cout)) return 0;
  return (size_t)p.splits * cout * 9 * cin * sizeof(float);
}

// dw[cout][3][3][cin] (bf16) = weight gradient of the 3x3 / stride 1 / pad 1 convolution y = conv(x, w):
// x [n,h,wd,cin], dy [n,h,wd,cout] NHWC bf16. workspace: tb200_conv3x3_wgrad_workspace bytes.
int tb200_conv3x3_wgrad(const void* x, const void* dy, void* dw, int n, int h, int wd, int cin, int cout,
                        float* workspace, size_t workspace_bytes, cudaStream_t stream) {
  W3Params p{};
  if (int rc = w3_plan(p, n, h, wd, cin, cout)) return rc;
  const size_t plane = (size_t)cout * 9 * cin;
  if (!workspace || workspace_bytes < (size_t)p.splits * plane * sizeof(float))
    return fail("tb200_conv3x3_wgrad: workspace too small (%s%ld bytes needed)", "", (long)((size_t)p.splits * plane * 4));
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) {
    TB_CUDA_OK(cudaFree(nullptr));
    ctx_bound = true;
  }
  p.partials = workspace;
  if (p.m_slabs == 1) p.y_slab_bytes = 1024;          // see w3_plan: the unread second slab aliases the first
  CUtensorMap mx, my;
  if (int rc = c3_map_nhwc(&mx, x, n, h, wd, cin, p.Wp)) return rc;
  if (int rc = c3_map_nhwc(&my, dy, n, h, wd, cout, p.Wp)) return rc;
  static DeviceOnce configured;
  if (int rc = ensure_dynamic_smem(configured, conv3x3_wgrad_kernel, 227 * 1024)) return rc;
  const int items = 2 * p.cblocks * p.mblocks * p.splits;
  const int sms = device_sm_count();
  const int grid = items < sms ? items : sms;
  conv3x3_wgrad_kernel<<<grid, kW3Threads, p.bar_off + 1024, stream>>>(mx, my, p);
  if (int rc = launched()) return rc;
  wgrad_reduce_kernel<<<(unsigned)((plane / 4 + 255) / 256), 256, 0, stream>>>(workspace, p.splits, plane,
                                                                             reinterpret_cast<__nv_bfloat16*>(dw));
  return launched();
}

size_t tb200_conv_stats_floats(int cout) { return (size_t)device_sm_count() * 4 * (size_t)cout * 2; }

int tb200_conv3x3_fprop(const void* x, const void* w, void* y, int n, int h, int wd, int cin, int cout, float* stats,
                        size_t stats_floats, int* nparts, cudaStream_t stream) {
  return conv3x3_common(x, w, y, n, h, wd, cin, cout, false, stats, stats_floats, nparts, stream);
}

// The same forward pass with a per-output-channel bias (bias_dtype TB200_F32 / TB200_BF16, may be NULL) and an
// optional ReLU in the epilogue: the convolution layers of VGG / AlexNet.
int tb200_conv3x3_fprop_bias(const void* x, const void* w, const void* bias, int bias_dtype, int relu, void* y, int n,
                             int h, int wd, int cin, int cout, cudaStream_t stream) {
  return conv3x3_common(x, w, y, n, h, wd, cin, cout, false, nullptr, 0, nullptr, stream, bias, bias_dtype, relu ? 1 : 0);
}

// 1 when tb200_conv3x3_fprop / _dgrad serve this geometry (channel multiples, one padded row per TMA box, two
// input stages next to the weight ring in shared memory), without touching the GPU.
int tb200_conv3x3_supported(int n, int h, int wd, int c_red, int k_out, int dgrad) {
  return conv3x3_common(nullptr, nullptr, nullptr, n, h, wd, c_red, k_out, dgrad != 0, nullptr, 0, nullptr, nullptr,
                        nullptr, 0, 0, true) == 0 ? 1 : 0;
}

int tb200_conv3x3_dgrad(const void* dy, const void* w, void* dx, int n, int h, int wd, int cin, int cout,
                        cudaStream_t stream) {
  // dx[n,h,w,ci] = sum_{r,s,co} dy[n, h+1-r, w+1-s, co] * wt[co,r,s,ci]: a forward pass over dy (reduction over
  // cout) with flipped taps whose B tiles are read out of the forward weight as stored
  return conv3x3_common(dy, w, dx, n, h, wd, cout, cin, true, nullptr, 0, nullptr, stream);
}

}  // extern "C"
