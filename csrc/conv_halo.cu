// Halo-tile convolution (EXPERIMENTAL — written after round 1's GPU budget was spent; it compiles for
// sm_100a but has not run yet, so nothing routes to it by default and its tests are opt-in).
//
//   y[n,h,w,k] = sum_{r,s,c} x[n, h+r-pad, w+s-pad, c] * wt[k,r,s,c]      stride 1, pad = (R-1)/2 = (S-1)/2
//   x, y NHWC bf16, wt [K,R,S,C] bf16, fp32 accumulation in TMEM.
//
// Why: the im2col-TMA kernel (gemm_tcgen05.cu) re-fetches the activation tile once per filter tap and is
// bound by L2->SM traffic (profiles/conv_r1g.md: 925 MB over the crossbar for a 51 MB input). Here the
// input rows an output tile needs are fetched ONCE per 64-channel block and the R*S taps are R*S groups of
// tcgen05.mma over the SAME shared-memory tile, each reading a window shifted by (r*Wp + s) pixel rows —
// a descriptor start advanced by that many 128-byte rows. bench/probe_shift.py showed on B200 that such
// windows read correctly with the descriptor's base-offset field at 0 (profiles/probe_shift_r1g.json).
//
// Geometry. Work is laid out in a padded space: every image is Hp = H+R-1 rows of Wp columns, Wp = W+S-1
// rounded up to a multiple of 8 that divides 128 (so a padded row is a whole number of 1024-byte swizzle
// periods and an output tile of 128 positions is TH = 128/Wp whole rows). Output position
// q = (n*Hp + h)*Wp + j is real for h < H, j < W and junk otherwise; its tap (r,s) reads input position
// q + r*Wp + s of the same space, whose image coordinates are (row%Hp - pad, col - pad) — the TMA loads
// fetch whole padded rows {64 ch, Wp, 1, 1} with out-of-range coordinates zero-filled (= the padding), and
// the TMA stores of the epilogue clip junk columns / rows against y's real extent.
//
// Small weight sets (R*S*(C/64) tiles of BLOCK_N x 64) stay resident in shared memory for the CTA's
// lifetime; larger ones stream through a ring, one tile per (channel block, tap).
//
// The PTX wrappers below duplicate those of gemm_tcgen05.cu on purpose: that file is validated on hardware
// and is not touched for an unvalidated kernel.
#include "common.cuh"
#include "../include/tiresias_b200.h"

#include <cuda.h>
#include <mutex>

using namespace tb200;

namespace {

constexpr int kThreadsHalo = 256;
constexpr int kAccStages = 2;
constexpr int kUmmaK = 16;

__device__ __forceinline__ void hb_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void hb_expect(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void hb_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void hb_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
  } while (!done);
}
__device__ __forceinline__ void h_tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
      "l"(map), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void h_tma_load_4d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1,
                                              int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(dst),
      "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void h_tma_store_4d(const CUtensorMap* map, uint32_t src, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(map),
               "r"(src), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}
__device__ __forceinline__ void h_fence_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void h_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void h_store_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void h_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void h_st_shared_16(uint32_t addr, const uint4& v) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void h_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void h_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void h_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void h_mma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void h_ld_32x32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
}
__device__ __forceinline__ void h_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major operand, 128-byte swizzle, 8-row groups 1024 bytes apart; `addr` may start any number of
// 128-byte rows into the tile (base offset stays 0: the swizzle follows the absolute address).
__device__ __forceinline__ uint64_t h_desc(uint32_t addr) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3ffff) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__host__ __device__ constexpr uint32_t h_idesc(int m, int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

struct HaloParams {
  int N, H, W, C, K;          // batch, image, channels in / out
  int R, S, pad;
  int Wp, Hp, TH;             // padded row width (divides 128), padded rows per image, rows per tile
  int a_rows;                 // pixel rows of one A stage that are loaded: TH + R - 1 padded rows
  int cblocks;                // C / 64
  int num_m, num_n;           // tiles over the padded positions / over K
  int resident;               // weights of the CTA's K block stay in shared memory
  int store_w;                // pixels per epilogue TMA store box: min(32, Wp)
};

template <int BLOCK_N>
struct HaloCfg {
  static constexpr int kBTileBytes = BLOCK_N * 128;
  static constexpr int kEpiBytes = 4 * 2 * 4096;               // 4 epilogue warps x 2 staging tiles
  static constexpr int kTmemCols = kAccStages * BLOCK_N;
};

// Shared-memory plan, computed on the host for the actual geometry (all regions 1024-byte aligned).
struct HaloSmem {
  int a_stage_bytes, a_stages;
  int b_bytes, b_stages;       // ring of BLOCK_N x 128 B tiles, or the resident set (b_stages tiles, loaded once)
  int epi_off, bar_off, total;
};

template <int BLOCK_N>
__global__ void __launch_bounds__(kThreadsHalo, 1)
conv_halo_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w,
                 const __grid_constant__ CUtensorMap map_y, const HaloParams p, const HaloSmem sm) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = smem_u32(smem_raw);
  const uint32_t a_smem = base;
  const uint32_t b_smem = base + sm.a_stages * sm.a_stage_bytes;
  const uint32_t epi_smem = base + sm.epi_off;
  const uint32_t bar = base + sm.bar_off;
  // barriers: A full/empty per A stage, B full/empty per B stage (ring) or one "weights landed", TMEM full/empty
  auto a_full = [&](int s) { return bar + 8u * s; };
  auto a_empty = [&](int s) { return bar + 8u * (sm.a_stages + s); };
  auto b_full = [&](int s) { return bar + 8u * (2 * sm.a_stages + s); };
  auto b_empty = [&](int s) { return bar + 8u * (2 * sm.a_stages + sm.b_stages + s); };
  const uint32_t t_base = bar + 8u * (2 * sm.a_stages + 2 * sm.b_stages);
  auto t_full = [&](int s) { return t_base + 8u * s; };
  auto t_empty = [&](int s) { return t_base + 8u * (kAccStages + s); };
  const uint32_t tmem_slot = t_base + 8u * 2 * kAccStages;

  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int taps = p.R * p.S;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_x) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_w) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_y) : "memory");
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < sm.a_stages; ++s) { hb_init(a_full(s), 1); hb_init(a_empty(s), 1); }
    for (int s = 0; s < sm.b_stages; ++s) { hb_init(b_full(s), 1); hb_init(b_empty(s), 1); }
    for (int s = 0; s < kAccStages; ++s) { hb_init(t_full(s), 1); hb_init(t_empty(s), 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot),
                 "n"(HaloCfg<BLOCK_N>::kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  h_fence_before();
  __syncthreads();
  h_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - base));

  // Work items: n block outermost, so that a CTA's consecutive tiles share the weights (and, when they are
  // resident, the whole kernel uses one n block: the host only sets `resident` for num_n == 1).
  const int items = p.num_m * p.num_n;

  if (warp == 0 && lane == 0) {
    // ================= producer =================
    if (p.resident) {
      // all R*S*cblocks weight tiles of n block 0, once: tile (cb, tap) at index cb*taps + tap
      hb_expect(b_full(0), (uint32_t)(taps * p.cblocks) * HaloCfg<BLOCK_N>::kBTileBytes);
      for (int cb = 0; cb < p.cblocks; ++cb)
        for (int t = 0; t < taps; ++t)
          h_tma_load_2d(b_smem + (uint32_t)(cb * taps + t) * HaloCfg<BLOCK_N>::kBTileBytes, &map_w, b_full(0),
                        t * p.C + cb * 64, 0);
    }
    int as = 0, bs = 0;
    uint32_t aph = 0, bph = 0;
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      const int m_tile = it % p.num_m, n_blk = it / p.num_m;
      const int row0 = m_tile * p.TH;                    // first padded row (global: n*Hp + hp) of the tile
      for (int cb = 0; cb < p.cblocks; ++cb) {
        hb_wait(a_empty(as), aph ^ 1);
        hb_expect(a_full(as), (uint32_t)(p.a_rows * p.Wp) * 128u);
        for (int i = 0; i < p.a_rows; ++i) {
          const int rg = row0 + i;
          const int n = rg / p.Hp, hp = rg - n * p.Hp;
          // whole padded row: columns -pad .. Wp-pad-1, image row hp-pad; out of range -> zeros
          h_tma_load_4d(a_smem + (uint32_t)as * sm.a_stage_bytes + (uint32_t)(i * p.Wp) * 128u, &map_x, a_full(as),
                        cb * 64, -p.pad, hp - p.pad, n);
        }
        if (++as == sm.a_stages) { as = 0; aph ^= 1; }
        if (!p.resident) {
          for (int t = 0; t < taps; ++t) {
            hb_wait(b_empty(bs), bph ^ 1);
            hb_expect(b_full(bs), HaloCfg<BLOCK_N>::kBTileBytes);
            h_tma_load_2d(b_smem + (uint32_t)bs * HaloCfg<BLOCK_N>::kBTileBytes, &map_w, b_full(bs),
                          t * p.C + cb * 64, n_blk * BLOCK_N);
            if (++bs == sm.b_stages) { bs = 0; bph ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1 && lane == 0) {
    // ================= MMA issuer =================
    const uint32_t idesc = h_idesc(128, BLOCK_N);
    int as = 0, bs = 0, acc = 0;
    uint32_t aph = 0, bph = 0, accph = 0;
    if (p.resident) {
      hb_wait(b_full(0), 0);
      h_fence_after();
    }
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      hb_wait(t_empty(acc), accph ^ 1);
      h_fence_after();
      const uint32_t tmem_d = tmem_base + acc * BLOCK_N;
      uint32_t first = 1;
      for (int cb = 0; cb < p.cblocks; ++cb) {
        hb_wait(a_full(as), aph);
        h_fence_after();
        const uint32_t a_tile = a_smem + (uint32_t)as * sm.a_stage_bytes;
        for (int t = 0; t < taps; ++t) {
          const int r = t / p.S, s = t - r * p.S;
          uint32_t b_tile;
          if (p.resident) {
            b_tile = b_smem + (uint32_t)(cb * taps + t) * HaloCfg<BLOCK_N>::kBTileBytes;
          } else {
            hb_wait(b_full(bs), bph);
            h_fence_after();
            b_tile = b_smem + (uint32_t)bs * HaloCfg<BLOCK_N>::kBTileBytes;
          }
          // the tap's window: 128 pixel rows starting (r*Wp + s) rows into the halo tile
          const uint64_t adesc = h_desc(a_tile + (uint32_t)(r * p.Wp + s) * 128u);
          const uint64_t bdesc = h_desc(b_tile);
#pragma unroll
          for (int k = 0; k < 64 / kUmmaK; ++k) {
            h_mma(tmem_d, adesc + 2u * k, bdesc + 2u * k, idesc, first ? 0u : 1u);
            first = 0;
          }
          if (!p.resident) {
            h_commit(b_empty(bs));
            if (++bs == sm.b_stages) { bs = 0; bph ^= 1; }
          }
        }
        h_commit(a_empty(as));
        if (++as == sm.a_stages) { as = 0; aph ^= 1; }
      }
      h_commit(t_full(acc));
      if (++acc == kAccStages) { acc = 0; accph ^= 1; }
    }
  } else if (warp >= 4) {
    // ================= epilogue =================
    // Warp ew owns tile rows [32*ew, 32*ew+32) = positions of `32 / store_w` padded-row segments; per 64
    // output channels: TMEM -> registers -> bf16 -> swizzled staging tile -> one TMA store per segment, which
    // clips junk columns (j >= W), junk rows (h >= H) and the batch tail against y's extent.
    const int ew = warp - 4;
    int acc = 0, epi_buf = 0;
    uint32_t accph = 0;
    const uint32_t bufs = epi_smem + (uint32_t)(ew * 2) * 4096u;
    const uint32_t swz = (uint32_t)(lane & 7);
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      const int m_tile = it % p.num_m, n_blk = it / p.num_m;
      hb_wait(t_full(acc), accph);
      h_fence_after();
      const int q0 = m_tile * 128 + ew * 32;              // first padded position of this warp's slice
      const uint32_t tcol = tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(acc * BLOCK_N);
#pragma unroll 1
      for (int c = 0; c < BLOCK_N / 64; ++c) {
        const int k0 = n_blk * BLOCK_N + c * 64;
        if (k0 >= p.K) break;
        uint32_t v[64];
        h_ld_32x32(tcol + c * 64, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
        h_ld_32x32(tcol + c * 64 + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
        h_ld_wait();
        const uint32_t buf = bufs + (uint32_t)epi_buf * 4096u;
        if (lane == 0) h_store_wait_read<1>();             // the stores that last read this staging tile
        __syncwarp();
        const uint32_t row_addr = buf + (uint32_t)lane * 128u;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          uint4 o;
          o.x = pack_bf16(__uint_as_float(v[8 * q]), __uint_as_float(v[8 * q + 1]));
          o.y = pack_bf16(__uint_as_float(v[8 * q + 2]), __uint_as_float(v[8 * q + 3]));
          o.z = pack_bf16(__uint_as_float(v[8 * q + 4]), __uint_as_float(v[8 * q + 5]));
          o.w = pack_bf16(__uint_as_float(v[8 * q + 6]), __uint_as_float(v[8 * q + 7]));
          h_st_shared_16(row_addr + (((uint32_t)q ^ swz) << 4), o);
        }
        h_fence_async();
        __syncwarp();
        if (lane == 0) {
          for (int seg = 0; seg < 32 / p.store_w; ++seg) {
            const int q = q0 + seg * p.store_w;
            const int rg = q / p.Wp, j = q - rg * p.Wp;
            const int n = rg / p.Hp, h = rg - n * p.Hp;
            h_tma_store_4d(&map_y, buf + (uint32_t)(seg * p.store_w) * 128u, k0, j, h, n);
          }
          h_store_commit();
        }
        epi_buf ^= 1;
      }
      h_fence_before();
      __syncwarp();
      if (lane == 0) hb_arrive(t_empty(acc));
      if (++acc == kAccStages) { acc = 0; accph ^= 1; }
    }
    if (lane == 0) h_store_wait_all();
  }

  h_fence_before();
  __syncthreads();
  if (warp == 2) {
    h_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                 "n"(HaloCfg<BLOCK_N>::kTmemCols)
                 : "memory");
  }
}

// ---- host ---------------------------------------------------------------------------------------------
typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                             const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                             CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeFn halo_encode() {
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
int halo_map_nhwc(CUtensorMap* out, const void* base, int n, int h, int w, int c, int box_w) {
  EncodeFn enc = halo_encode();
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

int halo_map_weights(CUtensorMap* out, const void* base, int k, int cols, int box_rows) {
  EncodeFn enc = halo_encode();
  if (!enc) return fail("cuTensorMapEncodeTiled entry point not available%s");
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)k};
  cuuint64_t strides[1] = {(cuuint64_t)cols * 2};
  cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail("cuTensorMapEncodeTiled (weights) failed (%s code %ld)", "", (long)r);
  return 0;
}

template <int BLOCK_N>
int launch_halo(const CUtensorMap& mx, const CUtensorMap& mw, const CUtensorMap& my, HaloParams p, cudaStream_t stream) {
  constexpr int kSmemMax = 227 * 1024;
  HaloSmem sm{};
  const int taps = p.R * p.S;
  // loaded rows plus the (S-1) positions the last tap's window runs past them, rounded to whole 8-row groups
  const int a_rows_alloc = ((p.a_rows * p.Wp + (p.S - 1)) + 7) / 8 * 8;
  sm.a_stage_bytes = a_rows_alloc * 128;
  const int b_tile = HaloCfg<BLOCK_N>::kBTileBytes;
  const int fixed = HaloCfg<BLOCK_N>::kEpiBytes + 1024;                      // staging + barriers
  const int resident_bytes = taps * p.cblocks * b_tile;
  p.resident = (p.num_n == 1 && resident_bytes + 2 * sm.a_stage_bytes + fixed <= kSmemMax) ? 1 : 0;
  if (p.resident) {
    sm.b_stages = 1;                                                          // one "weights landed" barrier pair
    sm.b_bytes = resident_bytes;
  } else {
    sm.b_stages = 4;
    sm.b_bytes = sm.b_stages * b_tile;
  }
  int a_stages = (kSmemMax - fixed - sm.b_bytes) / sm.a_stage_bytes;
  if (a_stages > 4) a_stages = 4;
  if (a_stages < 2) return fail("tb200_conv2d_fprop_halo: tile does not fit shared memory (Wp %s%ld)", "", p.Wp);
  sm.a_stages = a_stages;
  sm.epi_off = sm.a_stages * sm.a_stage_bytes + sm.b_bytes;
  sm.bar_off = sm.epi_off + HaloCfg<BLOCK_N>::kEpiBytes;
  sm.total = sm.bar_off + 1024;
  auto kern = conv_halo_kernel<BLOCK_N>;
  static DeviceOnce configured;      // one per template instantiation
  if (int rc = ensure_dynamic_smem(configured, kern, kSmemMax)) return rc;
  const int sms = device_sm_count();
  const int items = p.num_m * p.num_n;
  kern<<<items < sms ? items : sms, kThreadsHalo, sm.total, stream>>>(mx, mw, my, p, sm);
  return launched();
}


// ---- probe: shifted MN-major operand windows (groundwork for a halo-tile wgrad) -----------------------
// dW[k, (r,s), c] = sum_q dy[q, k] * x_padded[q + r*Wp + s, c]: the reduction runs over pixel rows, so both
// operands are MN-major (rows = reduction index) and a tap is a window of the x tile shifted ALONG the
// reduction dimension. One CTA: D[128, 64] = A[128, 64] . B[shift : shift+64, 0:64] with A K-major (as
// validated) and B stored [256 reduction rows][64 n] bf16, 128-byte swizzled, its MN-major descriptor
// advanced by shift*128 bytes.
__global__ void __launch_bounds__(128, 1)
probe_shift_mn_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                      float* __restrict__ D, int shift) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t a_smem = smem_u32(smem_raw);                 // 128 rows x 128 B (K-major A)
  const uint32_t b_smem = a_smem + 128 * 128;                 // 256 reduction rows x 128 B (MN-major B)
  const uint32_t bar_full = b_smem + 256 * 128;
  const uint32_t bar_done = bar_full + 8;
  const uint32_t tmem_slot = bar_full + 16;
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  if (threadIdx.x == 0) {
    hb_init(bar_full, 1);
    hb_init(bar_done, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(64)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  h_fence_before();
  __syncthreads();
  h_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - a_smem));
  if (threadIdx.x == 0) {
    hb_expect(bar_full, 128 * 128 + 256 * 128);
    h_tma_load_2d(a_smem, &map_a, bar_full, 0, 0);
    h_tma_load_2d(b_smem, &map_b, bar_full, 0, 0);
    h_tma_load_2d(b_smem + 128 * 128, &map_b, bar_full, 0, 128);
    hb_wait(bar_full, 0);
    h_fence_after();
    const uint32_t idesc = h_idesc(128, 64) | (1u << 16);       // B is MN-major
    // MN-major descriptor: 8 reduction rows = one 1024-byte atom (stride byte offset), 64-wide MN slabs 8192
    // bytes apart (leading byte offset; a single slab here)
    uint64_t bdesc = 0;
    bdesc |= (uint64_t)(((b_smem + (uint32_t)shift * 128u) & 0x3ffff) >> 4);
    bdesc |= (uint64_t)(8192 >> 4) << 16;
    bdesc |= (uint64_t)(1024 >> 4) << 32;
    bdesc |= (uint64_t)1 << 46;
    bdesc |= (uint64_t)2 << 61;
    const uint64_t adesc = h_desc(a_smem);
#pragma unroll
    for (int k = 0; k < 64 / kUmmaK; ++k)                       // 16 reduction rows = 2048 bytes = 128 units
      h_mma(tmem_base, adesc + 2u * k, bdesc + 128u * k, idesc, k > 0);
    h_commit(bar_done);
  }
  hb_wait(bar_done, 0);
  h_fence_after();
  const int row = warp * 32 + lane;
#pragma unroll 1
  for (int c = 0; c < 2; ++c) {
    uint32_t r[32];
    h_ld_32x32(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(c * 32), r);
    h_ld_wait();
#pragma unroll
    for (int j = 0; j < 32; ++j) D[(size_t)row * 64 + c * 32 + j] = __uint_as_float(r[j]);
  }
  h_fence_before();
  __syncthreads();
  if (warp == 0) {
    h_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(64) : "memory");
  }
}


// ---- probe: TMA destination that is 128-byte but not 1024-byte aligned -------------------------------
// A halo tile whose padded rows are not multiples of 8 pixels puts each row's TMA box at a shared-memory
// offset that is a multiple of 128 bytes only. This one-CTA kernel loads A[128, 64] (128-byte swizzle) to
// a_smem + shift*128 as `pieces` row boxes and multiplies it from a descriptor at the same address:
// D = A . B^T must come out unchanged if the swizzle of both the TMA write and the MMA read follows the
// absolute shared-memory address.
__global__ void __launch_bounds__(128, 1)
probe_tma_dst_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                     float* __restrict__ D, int shift, int piece_rows) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t a_smem = smem_u32(smem_raw);                 // 256 rows x 128 B
  const uint32_t b_smem = a_smem + 256 * 128;                 // 64 rows x 128 B
  const uint32_t bar_full = b_smem + 64 * 128;
  const uint32_t bar_done = bar_full + 8;
  const uint32_t tmem_slot = bar_full + 16;
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  if (threadIdx.x == 0) {
    hb_init(bar_full, 1);
    hb_init(bar_done, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(64)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  h_fence_before();
  __syncthreads();
  h_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - a_smem));
  if (threadIdx.x == 0) {
    hb_expect(bar_full, 128 * 128 + 64 * 128);
    for (int r0 = 0; r0 < 128; r0 += piece_rows)              // map_a's box is piece_rows rows
      h_tma_load_2d(a_smem + (uint32_t)(shift + r0) * 128u, &map_a, bar_full, 0, r0);
    h_tma_load_2d(b_smem, &map_b, bar_full, 0, 0);
    hb_wait(bar_full, 0);
    h_fence_after();
    const uint32_t idesc = h_idesc(128, 64);
    const uint64_t adesc = h_desc(a_smem + (uint32_t)shift * 128u);
    const uint64_t bdesc = h_desc(b_smem);
#pragma unroll
    for (int k = 0; k < 64 / kUmmaK; ++k) h_mma(tmem_base, adesc + 2u * k, bdesc + 2u * k, idesc, k > 0);
    h_commit(bar_done);
  }
  hb_wait(bar_done, 0);
  h_fence_after();
  const int row = warp * 32 + lane;
#pragma unroll 1
  for (int c = 0; c < 2; ++c) {
    uint32_t r[32];
    h_ld_32x32(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(c * 32), r);
    h_ld_wait();
#pragma unroll
    for (int j = 0; j < 32; ++j) D[(size_t)row * 64 + c * 32 + j] = __uint_as_float(r[j]);
  }
  h_fence_before();
  __syncthreads();
  if (warp == 0) {
    h_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(64) : "memory");
  }
}

}  // namespace

// Hardware probe (see probe_tma_dst_kernel): A bf16 [128, 64], B bf16 [64, 64], D fp32 [128, 64] = A . B^T with A
// landed `shift` 128-byte rows into shared memory, as boxes of `piece_rows` rows (must divide 128).
extern "C" int tb200_probe_tma_dst_shift(const void* A, const void* B, float* D, int shift, int piece_rows,
                                         cudaStream_t stream) {
  if (shift < 0 || shift > 120) return fail("tb200_probe_tma_dst_shift: shift %s(%ld) outside 0..120", "", shift);
  if (piece_rows <= 0 || 128 % piece_rows) return fail("tb200_probe_tma_dst_shift: piece_rows must divide 128%s");
  TB_CUDA_OK(cudaFree(nullptr));
  CUtensorMap ma, mb;
  if (int rc = halo_map_weights(&ma, A, 128, 64, piece_rows)) return rc;
  if (int rc = halo_map_weights(&mb, B, 64, 64, 64)) return rc;
  const int smem = 256 * 128 + 64 * 128 + 64;
  TB_CUDA_OK(cudaFuncSetAttribute(probe_tma_dst_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  probe_tma_dst_kernel<<<1, 128, smem, stream>>>(ma, mb, D, shift, piece_rows);
  return launched();
}

// Hardware probe (see probe_shift_mn_kernel): A bf16 [128, 64] (K-major), B bf16 [256, 64] (reduction rows x n),
// D fp32 [128, 64] = A . B[shift : shift+64, :].
extern "C" int tb200_probe_shifted_mn_operand(const void* A, const void* B, float* D, int shift, cudaStream_t stream) {
  if (shift < 0 || shift > 192) return fail("tb200_probe_shifted_mn_operand: shift %s(%ld) outside 0..192", "", shift);
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) {
    TB_CUDA_OK(cudaFree(nullptr));
    ctx_bound = true;
  }
  CUtensorMap ma, mb;
  if (int rc = halo_map_weights(&ma, A, 128, 64, 128)) return rc;     // [rows, 64 cols], box {64, 128}
  if (int rc = halo_map_weights(&mb, B, 256, 64, 128)) return rc;
  const int smem = 128 * 128 + 256 * 128 + 64;
  static DeviceOnce configured;
  if (int rc = ensure_dynamic_smem(configured, probe_shift_mn_kernel, smem)) return rc;
  probe_shift_mn_kernel<<<1, 128, smem, stream>>>(ma, mb, D, shift);
  return launched();
}

extern "C" int tb200_conv2d_fprop_halo(const void* x, const void* w, void* y, int n, int h, int wd, int cin, int cout,
                                       int R, int S, cudaStream_t stream) {
  if (n <= 0 || h <= 0 || wd <= 0 || cin <= 0 || cout <= 0) return fail("tb200_conv2d_fprop_halo: empty problem%s");
  if (R != S || (R != 3 && R != 1)) return fail("tb200_conv2d_fprop_halo: 3x3 (or 1x1) filters only%s");
  if (cin % 64) return fail("tb200_conv2d_fprop_halo: cin %s(%ld) must be a multiple of 64", "", cin);
  if (cout % 8) return fail("tb200_conv2d_fprop_halo: cout %s(%ld) must be a multiple of 8", "", cout);
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) {
    TB_CUDA_OK(cudaFree(nullptr));
    ctx_bound = true;
  }
  HaloParams p{};
  p.N = n; p.H = h; p.W = wd; p.C = cin; p.K = cout; p.R = R; p.S = S; p.pad = (R - 1) / 2;
  int wp = 8;
  while (wp < wd + S - 1) wp *= 2;                         // 8, 16, 32, 64, 128: divides 128, multiple of 8
  if (wp > 128) return fail("tb200_conv2d_fprop_halo: image width %s(%ld) too large for one 128-position tile row", "", wd);
  p.Wp = wp;
  p.Hp = h + R - 1;
  p.TH = 128 / wp;
  p.a_rows = p.TH + R - 1;
  p.cblocks = cin / 64;
  const long long padded_rows = (long long)n * p.Hp;
  p.num_m = (int)((padded_rows + p.TH - 1) / p.TH);
  p.store_w = wp < 32 ? wp : 32;
  const int block_n = cout <= 64 ? 64 : (cout <= 128 ? 128 : 256);
  p.num_n = (cout + block_n - 1) / block_n;
  CUtensorMap mx, mw, my;
  if (int rc = halo_map_nhwc(&mx, x, n, h, wd, cin, wp)) return rc;
  if (int rc = halo_map_nhwc(&my, y, n, h, wd, cout, p.store_w)) return rc;
  if (int rc = halo_map_weights(&mw, w, cout, R * S * cin, block_n)) return rc;
  if (block_n == 64) return launch_halo<64>(mx, mw, my, p, stream);
  if (block_n == 128) return launch_halo<128>(mx, mw, my, p, stream);
  return launch_halo<256>(mx, mw, my, p, stream);
}
