// 2-CTA (cta_group::2) persistent GEMM: a CTA pair (cluster of two, one TPC) computes a 256 x 256 tile with
// UMMA M = 256. Each CTA of the pair stages only ITS half of both operands - 128 rows of A and 128 rows of B
// per 64-wide k-block, 32 KB per stage instead of the 48 KB of the one-CTA 128 x 256 kernel - and the tensor
// cores of both SMs read the B halves out of both CTAs' shared memory. Per output element that halves the
// L2 -> SM traffic of B and the shared-memory reads per MMA, and leaves room for six stages.
//
// Included by gemm_tcgen05.cu (shares its PTX wrappers, GemmParams and epilogue_tile). Plain GEMMs only
// (K- or MN-major operands, bias / activation / fp32 / split-K epilogues through the TMA-store path); the
// convolution modes, the statistics epilogue and the pull variant stay on the one-CTA kernel.
// Every load here is a plain per-CTA TMA load (no multicast); the only cluster-wide instructions are the MMA
// itself, its commit (which arrives on both CTAs' barriers) and the mbarrier arrivals between the two CTAs.
//
// The reference (matthewygf/Tiresias) has no native code; nothing here has a counterpart.
#pragma once

namespace pair {

constexpr int kStages = 6;
constexpr int kABytes = BLOCK_M * BLOCK_K * 2;            // this CTA's 128 rows of A
constexpr int kBBytes = 128 * BLOCK_K * 2;                // this CTA's 128 rows (of the pair's 256) of B
constexpr int kStageBytes = kABytes + kBBytes;            // 32 KB
constexpr int kTileN = 256;
constexpr int kTmemCols = kAccumStages * kTileN;          // 512
constexpr int kEpiTileBytes = 32 * 128;
constexpr int kEpiBufs = 2;
constexpr int kEpiBytes = 4 * kEpiBufs * kEpiTileBytes;   // 32 KB
constexpr int kEpiBiasBytes = kAccumStages * 256 * 4;
constexpr int kBarrierBytes = (2 * kStages + 2 * kAccumStages) * 8 + 16 + 8 * 8;   // + 2 addend barriers per epilogue warp
constexpr int kSmemBytes = kStages * kStageBytes + kEpiBytes + kEpiBiasBytes + kBarrierBytes;
static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");

__device__ __forceinline__ bool elect_one_sync() {       // one lane of a converged warp
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
// shared::cluster address of `addr` (a shared::cta address of this CTA) in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// per-CTA TMA load whose bytes are counted on the LEADER CTA's barrier (cluster address)
__device__ __forceinline__ void tma_load_2d_pair(uint32_t smem_dst, const CUtensorMap* map, uint32_t cluster_bar, int c0,
                                                 int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_dst),
      "l"(map), "r"(cluster_bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tc_mma2_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrives on the barrier at this shared-memory offset in BOTH CTAs once the pair's MMAs so far have completed
__device__ __forceinline__ void tc_commit2(uint32_t bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
      "h"((uint16_t)3)
      : "memory");
}

// Tile order: groups of kGroupM rows of tiles, column by column inside a group, so the pairs running at one
// time (74) cover a roughly square patch of the output (8 x 9 tiles) and A / B tiles are re-read from L2, not
// from DRAM (m-fastest over the whole matrix re-streams all of A once per wave on a large problem).
constexpr int kGroupM = 8;
__device__ __forceinline__ void tile_coords(const GemmParams& p, int tile, int* m_blk, int* n_blk) {
  const int group_tiles = kGroupM * p.num_n;
  const int g = tile / group_tiles, r = tile - g * group_tiles;
  const int rows = p.num_m - g * kGroupM < kGroupM ? p.num_m - g * kGroupM : kGroupM;
  const int n = r / rows;
  *n_blk = n;
  *m_blk = g * kGroupM + (r - n * rows);
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmThreads, 1)
gemm2_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
             const __grid_constant__ CUtensorMap map_d, const __grid_constant__ PullMaps extra, const GemmParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t smem_base = smem_u32(smem_raw);
  const uint32_t a_smem = smem_base;
  const uint32_t b_smem = a_smem + kStages * kABytes;
  const uint32_t epi_smem = b_smem + kStages * kBBytes;
  const uint32_t epi_bias_smem = epi_smem + kEpiBytes;
  const uint32_t bar_base = epi_bias_smem + kEpiBiasBytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };                       // leader's are the live ones
  auto empty_bar = [&](int s) { return bar_base + 8u * (kStages + s); };          // per CTA
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * kStages + a); };      // per CTA
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * kStages + kAccumStages + a); };   // leader's
  const uint32_t tmem_slot = bar_base + 8u * (2 * kStages + 2 * kAccumStages);
  const uint32_t add_bar_base = tmem_slot + 16u;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_base));

  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const uint32_t rank = cluster_ctarank();
  const int pair_id = blockIdx.x / 2, num_pairs = gridDim.x / 2;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full_bar(s), 1);       // the leader producer's expect_tx arrival; bytes come from both CTAs
      mbar_init(empty_bar(s), 1);      // one multicast commit
    }
    for (int a = 0; a < kAccumStages; ++a) {
      mbar_init(tfull_bar(a), 1);      // one multicast commit
      mbar_init(tempty_bar(a), 8);     // four epilogue warps in each CTA
    }
    for (int i = 0; i < 8; ++i) mbar_init(add_bar_base + 8u * i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  cluster_sync_all();                  // both CTAs' barriers and TMEM exist before anything crosses over
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  const int items = p.num_m * p.num_n * p.splits;

  if (warp == 0) {
    // ================= TMA producer (both CTAs, one elected lane) =================
    if (elect_one_sync()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int it = pair_id; it < items; it += num_pairs) {
        const int ks = it % p.splits;
        const int tile = it / p.splits;
        int m_blk, n_blk;
        tile_coords(p, tile, &m_blk, &n_blk);
        const int kb0 = (int)((long long)ks * p.num_kb / p.splits);
        const int kb1 = (int)((long long)(ks + 1) * p.num_kb / p.splits);
        const int a_row = m_blk * 256 + (int)rank * 128, b_row = n_blk * kTileN + (int)rank * 128;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1);
          const uint32_t fb = mapa(full_bar(stage), 0);
          if (rank == 0) mbar_expect_tx(full_bar(stage), 2 * kStageBytes);
          if (!p.a_mn) {
            tma_load_2d_pair(a_smem + stage * kABytes, &map_a, fb, kb * BLOCK_K, a_row);
          } else {
#pragma unroll
            for (int sl = 0; sl < 2; ++sl)                     // 64-wide slabs of the [K, M] matrix
              tma_load_2d_pair(a_smem + stage * kABytes + sl * 8192, &map_a, fb, a_row + sl * 64, kb * BLOCK_K);
          }
          if (!p.b_mn) {
            tma_load_2d_pair(b_smem + stage * kBBytes, &map_b, fb, kb * BLOCK_K, b_row);
          } else {
#pragma unroll
            for (int sl = 0; sl < 2; ++sl)
              tma_load_2d_pair(b_smem + stage * kBBytes + sl * 8192, &map_b, fb, b_row + sl * 64, kb * BLOCK_K);
          }
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer (leader CTA only) =================
    if (rank == 0 && elect_one_sync()) {
      const uint32_t idesc = make_idesc_bf16(256, kTileN, p.a_mn, p.b_mn);
      const uint32_t a_step = p.a_mn ? 128u : 2u, b_step = p.b_mn ? 128u : 2u;
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      for (int it = pair_id; it < items; it += num_pairs) {
        const int ks = it % p.splits;
        const int kb0 = (int)((long long)ks * p.num_kb / p.splits);
        const int kb1 = (int)((long long)(ks + 1) * p.num_kb / p.splits);
        mbar_wait(tempty_bar(as), aphase ^ 1);       // both CTAs' epilogues have drained this accumulator
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + as * kTileN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint64_t adesc = p.a_mn ? make_mnmajor_sw128_desc(a_smem + stage * kABytes)
                                        : make_kmajor_sw128_desc(a_smem + stage * kABytes);
          const uint64_t bdesc = p.b_mn ? make_mnmajor_sw128_desc(b_smem + stage * kBBytes)
                                        : make_kmajor_sw128_desc(b_smem + stage * kBBytes);
#pragma unroll
          for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
            tc_mma2_bf16(tmem_d, adesc + a_step * k, bdesc + b_step * k, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          tc_commit2(empty_bar(stage));              // both CTAs' stage is reusable once these MMAs have read it
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        tc_commit2(tfull_bar(as));                   // accumulator complete, in both CTAs' TMEM
        if (++as == kAccumStages) { as = 0; aphase ^= 1; }
      }
    }
  } else if (warp >= 4) {
    // ================= epilogue (both CTAs: each its own 128 rows x 256 columns) =================
    const int ew = warp - 4;
    int as = 0;
    uint32_t aphase = 0;
    int epi_buf = 0;
    uint32_t add_count = 0;
    for (int it = pair_id; it < items; it += num_pairs) {
      const int ks = it % p.splits;
      const int tile = it / p.splits;
      int m_blk, n_blk;
      tile_coords(p, tile, &m_blk, &n_blk);
      float* bias_s = reinterpret_cast<float*>(smem_raw + (epi_bias_smem - smem_base)) + as * 256;
      if (p.bias && p.splits == 1) {
        for (int j = threadIdx.x - 128; j < kTileN; j += 128) {
          const int col = n_blk * kTileN + j;
          bias_s[j] = col < p.N ? load_bias(p.bias, p.bias_dtype, col) : 0.f;
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");     // the four epilogue warps only
      }
      mbar_wait(tfull_bar(as), aphase);
      tc_fence_after();
      const int row0 = m_blk * 256 + (int)rank * 128 + ew * 32;
      const uint32_t tcol = tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(as * kTileN);
      const uint32_t bufs = epi_smem + (uint32_t)(ew * kEpiBufs) * kEpiTileBytes;
      const int col_base = n_blk * kTileN;
      const int variant = (p.splits > 1) ? 12 : (p.addend ? 14 : (p.act * 4 + (p.bias ? 2 : 0) + (p.d_dtype == kF32 ? 1 : 0)));
      switch (variant) {
        case 14:
          epilogue_tile<kTileN, kEpiBufs, 0, false, false, false, true>(p, &map_d, tcol, bufs, kEpiTileBytes, epi_buf,
                                                                        bias_s, row0, col_base, ks, lane, nullptr,
                                                                        &extra.addend, add_bar_base + 16u * ew, &add_count);
          break;
#define TB_EPI2(ID, ACT, BIAS, F32)                                                                               \
  case ID:                                                                                                        \
    epilogue_tile<kTileN, kEpiBufs, ACT, BIAS, F32>(p, &map_d, tcol, bufs, kEpiTileBytes, epi_buf, bias_s, row0, \
                                                    col_base, ks, lane);                                          \
    break;
        TB_EPI2(0, 0, false, false) TB_EPI2(1, 0, false, true) TB_EPI2(2, 0, true, false) TB_EPI2(3, 0, true, true)
        TB_EPI2(4, 1, false, false) TB_EPI2(5, 1, false, true) TB_EPI2(6, 1, true, false) TB_EPI2(7, 1, true, true)
        TB_EPI2(8, 2, false, false) TB_EPI2(9, 2, false, true) TB_EPI2(10, 2, true, false) TB_EPI2(11, 2, true, true)
        TB_EPI2(12, 0, false, true)
#undef TB_EPI2
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(mapa(tempty_bar(as), 0));
      if (++as == kAccumStages) { as = 0; aphase ^= 1; }
    }
    if (lane == 0) tma_store_wait_all();     // smem must outlive the last bulk stores
  }

  __syncwarp();
  tc_fence_before();
  cluster_sync_all();                        // neither CTA may leave (or free TMEM) while its peer still uses it
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(kTmemCols) : "memory");
  }
}

}  // namespace pair
