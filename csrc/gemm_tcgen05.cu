// K5 / K2: hand-written tcgen05 GEMM for sm_100a.
//
//   D[M,N] = act(A[M,K] . B[N,K]^T + bias)      A, B bf16 (K contiguous), fp32 accumulate
//
// Structure (one persistent CTA per SM, 256 threads):
//   warp 0 lane 0 : TMA producer  - cp.async.bulk.tensor tiles of A and B into a ring of
//                   128B-swizzled shared-memory stages, completion on "full" mbarriers
//   warp 1 lane 0 : MMA issuer    - tcgen05.mma.cta_group::1.kind::f16, accumulators in TMEM,
//                   tcgen05.commit releases smem stages ("empty") and publishes finished
//                   accumulators ("tmem_full")
//   warp 2        : TMEM allocator (tcgen05.alloc / dealloc, 2 accumulator stages)
//   warp 3        : (pull variant) copy engine: bulk-copies the remote shards of B into the local weight buffer, n tile
//                   by n tile, and publishes a per-tile "ready" flag the producers wait on
//   warps 4-7     : epilogue      - tcgen05.ld TMEM -> registers, bias / activation / addend / BatchNorm statistics,
//                   cast, swizzled staging tile, TMA store (EPI8 instantiations: warps 4-11, two per lane quarter)
// Large plain GEMMs run on the 2-CTA kernel of gemm_2cta.cuh (included below; same parameters and epilogue).
//
// The pull variant (K2) is the load side of the parameter pull that the reference only prices
// as "model_size / bandwidth" (core/network/network_service.py:34-37): B is sharded by rows
// over the parameter-server ranks. Round 1 let every CTA's TMA read its B tile straight out of the owner's
// memory; measured at 4 and 8 GPUs that loses to NCCL all_gather + cuBLAS by 2-17x, because peer memory is
// not cached in the local L2 and every 128-row block of A re-fetched the tile over NVLink
// (profiles/comm_pull_r2.md). Now every remote byte crosses NVLink ONCE: warp 3 of every CTA is a copy engine
// (cp.async.bulk owner -> shared -> local weight buffer, 8 KB pieces through four staging buffers) that walks the n tiles in
// the order the GEMM consumes them and publishes one flag per tile; the producers wait for a tile's flag and
// then load it from LOCAL memory like any other GEMM, so the transfer of tile n+1.. overlaps the MMAs of tile
// n. After the kernel the full gathered matrix is resident locally (backward and the other layers use it).
// FC layer shapes come from core/models.py:9-20 (VGG fc6 25088x4096 etc.).
#include "common.cuh"
#include "../include/tiresias_b200.h"

#include <cuda.h>
#include <mutex>
#include <unordered_map>

using namespace tb200;

namespace {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;   // 64 bf16 = 128 bytes = one swizzle-128B row
constexpr int UMMA_K = 16;
constexpr int kGemmThreads = 256;
constexpr int kAccumStages = 2;

// THIN = short reduction (K <= 256): the kernel is bound by its output stores, not by the mainloop, so it
// trades one mainloop stage for a deeper store pipeline in the epilogue.
constexpr int kCopyPiece = 8 * 1024;    // pull variant: bytes per bulk copy of the copy engine
constexpr int kCopyBufs = 4;            // staging buffers (three loads in flight)

// EPI8 = eight epilogue warps instead of four (two per TMEM lane quarter, alternating 64-column chunks): the
// BatchNorm-statistics epilogue triples the instructions per chunk and made the 1x1 forward convolutions
// epilogue-issue bound (119 us against 88 us for the same GEMM without statistics, profiles/conv_r2.md section 5).
template <int BLOCK_N, bool THIN, bool PULL = false, bool EPI8 = false>
struct Cfg {
  static constexpr int kStages = (BLOCK_N == 256 ? 4 : 6) - (THIN ? 1 : 0) - (PULL ? 1 : 0) - ((EPI8 && !THIN) ? 1 : 0);
  static constexpr int kEpiWarps = EPI8 ? 8 : 4;
  static constexpr int kThreads = 128 + 32 * kEpiWarps;
  static constexpr int kCopyBytes = PULL ? kCopyBufs * kCopyPiece : 0;
  static constexpr int kABytes = BLOCK_M * BLOCK_K * 2;
  static constexpr int kBBytes = BLOCK_N * BLOCK_K * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kTmemCols = kAccumStages * BLOCK_N;  // 512 or 256 (power of two)
  // pipeline barriers, 4 copy-engine barriers, the TMEM slot (16 bytes), 8 addend-tile barriers (2 per epilogue warp)
  static constexpr int kBarrierBytes = (2 * kStages + 2 * kAccumStages + 4) * 8 + 16 + 8 * 8;
  // Epilogue staging: per epilogue warp two 32-row x 128-byte tiles (swizzle-128B), TMA-stored.
  static constexpr int kEpiTileBytes = 32 * 128;
  // kEpiBufs staging tiles per epilogue warp: a TMA store occupies its tile for ~1 us, so with two tiles a
  // warp could not retire more than ~8 GB/s and thin-K (store-bound) shapes topped out near 2.4 TB/s.
  static constexpr int kEpiBufs = (THIN && !EPI8) ? 4 : 2;
  static constexpr int kEpiBytes = kEpiWarps * kEpiBufs * kEpiTileBytes;
  static constexpr int kEpiBiasBytes = kAccumStages * 256 * 4;   // the tile's bias row, per accumulator stage
  static constexpr int kSmemBytes =
      kStages * kStageBytes + kEpiBytes + kEpiBiasBytes + kCopyBytes + kBarrierBytes;
  static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");
};

// ---- PTX wrappers ---------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
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
__device__ __forceinline__ void tma_load_2d(uint32_t smem_dst, const CUtensorMap* map, uint32_t bar,
                                            int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_dst),
      "l"(map), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t smem_dst, const CUtensorMap* map, uint32_t bar,
                                            int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(smem_dst),
      "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
// im2col-mode TMA over an NHWC activation tensor: loads `pixelsPerColumn` consecutive output pixels
// (traversing W, then H, then N inside the map's bounding box, by the map's traversal stride) x
// `channelsPerPixel` channels starting at channel c, for the filter tap (off_h, off_w). (w, h) is the
// input-space position of the first pixel's window corner, i.e. out_pos * stride - pad. Positions that
// fall into the padding or past the last image are zero-filled. The box lands as rows of 128 bytes
// (64 bf16 channels), 128-byte swizzled: the same layout a tiled box {64, rows} produces.
__device__ __forceinline__ void tma_load_im2col_4d(uint32_t smem_dst, const CUtensorMap* map, uint32_t bar,
                                                   int c, int w, int h, int n, int off_w, int off_h) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.im2col.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};" ::"r"(smem_dst),
      "l"(map), "r"(bar), "r"(c), "r"(w), "r"(h), "r"(n), "h"((uint16_t)off_w), "h"((uint16_t)off_h)
      : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, uint32_t smem_src, int c0,
                                             int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(map),
               "r"(smem_src), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* map, uint32_t smem_src, int c0,
                                             int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(map),
               "r"(smem_src), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}
// Make generic-proxy shared-memory writes visible to the async proxy (TMA) before a bulk store.
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
// at most N bulk stores of this thread may still be reading their staging tiles
template <int N>
__device__ __forceinline__ void tma_store_wait_read_n() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void st_shared_16(uint32_t addr, const uint4& v) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void tma_store_wait_n() {   // all but the N latest bulk groups have completed
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void tma_store_wait_all() {
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar)
               : "memory");
}
__device__ __forceinline__ void tc_mma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                            uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// 32 lanes x 32 columns of fp32: thread t of the warp receives row (lane_base + t).
__device__ __forceinline__ void tc_ld_32x32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]),
        "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]),
        "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
}
__device__ __forceinline__ void tc_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// Shared-memory matrix descriptor: K-major operand, 128-byte swizzle, rows of 128 bytes,
// 8-row groups 1024 bytes apart (cute::UMMA::SmemDescriptor, version 1 = Blackwell).
__device__ __forceinline__ uint64_t make_kmajor_sw128_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3ffff) >> 4);        // start address  [0,14)
  d |= (uint64_t)1 << 16;                             // leading byte offset (unused for SW128 K-major)
  d |= (uint64_t)(1024 >> 4) << 32;                   // stride byte offset [32,46)
  d |= (uint64_t)1 << 46;                             // descriptor version
  d |= (uint64_t)2 << 61;                             // SWIZZLE_128B
  return d;
}

// Shared-memory matrix descriptor for an MN-major operand (the non-reduction dimension is the
// contiguous one, e.g. W[N_red, K_out] used as the B of dx = dy . W). The tile is stored as slabs of
// 64 MN-elements x BLOCK_K reduction rows (each slab one TMA box of 128-byte rows, 128-byte swizzle):
// inside a slab 8 reduction rows form one 1024-byte swizzle atom (stride byte offset), slabs are
// 64*128 = 8192 bytes apart (leading byte offset). Canonical form ((8,n),(8,k)):((1,LBO),(8,SBO)) in
// 16-byte units, cute::UMMA::make_umma_desc<Major::MN>.
__device__ __forceinline__ uint64_t make_mnmajor_sw128_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3ffff) >> 4);
  d |= (uint64_t)(8192 >> 4) << 16;                   // leading byte offset: next 64-wide MN slab
  d |= (uint64_t)(1024 >> 4) << 32;                   // stride byte offset: next 8 reduction rows
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

// Instruction descriptor for kind::f16: bf16 x bf16 -> fp32; a_mn / b_mn select MN-major operands.
__host__ __device__ constexpr uint32_t make_idesc_bf16(int umma_m, int umma_n, int a_mn = 0, int b_mn = 0) {
  return (1u << 4)                       // D format  = F32
         | (1u << 7)                     // A format  = BF16
         | (1u << 10)                    // B format  = BF16
         | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16)       // 0 = K-major, 1 = MN-major
         | ((uint32_t)(umma_n >> 3) << 17) | ((uint32_t)(umma_m >> 4) << 24);
}

__device__ __forceinline__ float load_bias(const void* bias, int dtype, size_t i) {
  return dtype == kBF16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(bias)[i])
                        : __ldg(reinterpret_cast<const float*>(bias) + i);
}

__device__ __forceinline__ float apply_act(float x, int act) {
  if (act == 1) return fmaxf(x, 0.f);
  if (act == 2) {
    const float k0 = 0.7978845608028654f, k1 = 0.044715f;
    float t = tanhf(k0 * (x + k1 * x * x * x));
    return 0.5f * x * (1.f + t);
  }
  return x;
}

// Implicit-GEMM convolution (NHWC activations, [Cout, R, S, Cin] weights): the producer warp replaces one
// operand's tiled TMA loads by im2col-mode loads; mainloop, TMEM accumulators and epilogue are the GEMM's.
//   mode 1 (fprop, and dgrad of a stride-1 convolution run as an fprop over dy with flipped taps):
//          A[pixel, (tap, c)] = act[n, oh*stride - pad + r, ow*stride - pad + s, c]   (K-major, 128 pixels)
//   mode 2 (wgrad): B[(tap, c), pixel] — the same box with 64 pixels, used as an MN-major operand whose
//          reduction dimension is the pixel index.
struct ConvGeom {
  int mode;          // 0: plain GEMM
  int OH, OW;        // extent of the pixel grid the GEMM dimension enumerates
  int stride, lower; // traversal stride, lower corner (= -pad) of the im2col bounding box
  int R, S;          // filter taps
  int cblocks;       // 64-channel blocks per tap of the im2col operand
  int flip;          // mode 1: weights are read as tap (R-1-r, S-1-s) through a 3-D map, MN-major (dgrad)
  // n / d as (n * ceil(2^40 / d)) >> 40, exact for n * d < 2^40 (host-computed): the producer is ONE thread, and
  // a hardware-less integer division costs it ~40 dependent instructions — six of them per k-block made the
  // wgrad producer slower than the MMAs it feeds (profiles/conv_r2.md)
  unsigned long long m_plane, m_ow, m_cblocks, m_s;
  int lower_w;       // lower corner along w (== lower except for the parity classes of a strided dgrad)
  // flip mode: window tap (r, s) reads forward-weight tap (wt_rb - wt_rs*r) * wt_S + (wt_sb - wt_ss*s)
  int wt_rb, wt_rs, wt_sb, wt_ss, wt_S;
  // scatter mode (strided dgrad): GEMM row (n, a, b) over the OH x OW class grid is stored to pixel
  // (n, a*sc_stride + sc_ph, b*sc_stride + sc_pw) of an sc_H x sc_W image
  int sc_on, sc_H, sc_W, sc_stride, sc_ph, sc_pw;
};

__device__ __forceinline__ int cdiv40(int n, unsigned long long magic) {
  return (int)(((unsigned long long)(unsigned)n * magic) >> 40);
}

struct GemmParams {
  int M, N, K;
  ConvGeom conv;
  int num_m, num_n, splits, num_kb;
  void* D;            // final output (splits == 1)
  float* partials;    // [splits][M][N] fp32 (splits > 1)
  const void* bias;
  int bias_dtype;
  int d_dtype;
  int act;
  int direct_store;   // 1: per-thread global stores (row stride not 16-byte aligned); 0: TMA store
  int a_mn, b_mn;     // operand is stored [K, M] / [K, N] (non-reduction dimension contiguous)
  const void* addend; // optional: bf16 [M, N] added to the product in the epilogue (residual-gradient fusion)
  float* stats;       // optional: per-channel (sum, sum of squares) partials of the bf16 output,
                      // [gridDim.x * 4][N][2] (BatchNorm statistics from the epilogue; bf16, no bias / activation)
  // pull variant
  int rows_per_rank;
  int world;
  int rank;
  int bucket;
  const uint32_t* epoch;
  const uint32_t* flags;   // local flag pad
  int* status;
  unsigned long long budget;
  // copy engine of the pull variant
  const uint8_t* owner_base[TB200_MAX_WORLD];   // every rank's copy of B (peer mappings), byte address of row 0
  uint8_t* local_base;                          // this rank's copy
  uint32_t* pull_cnt;                           // [num_n] arrivals of the CTAs' copy warps (self-resetting)
  uint32_t* pull_ready;                         // [num_n] epoch at which the tile was last gathered
};

struct PullMaps {
  CUtensorMap b_local;             // this rank's full B: gathered by the copy engine, read by the producers
  CUtensorMap addend;              // addend epilogue: [M, N] bf16 with the output map's box (64 columns x 32 rows)
};

__device__ __forceinline__ void bulk_load_1d(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void bulk_store_1d(void* dst, uint32_t src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void st_release_gpu(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Epilogue of one 32-row slice of an accumulator tile (one epilogue warp): per 128 output bytes per row
// (64 bf16 or 32 fp32 columns) TMEM -> registers, + bias (the tile's bias row sits in shared memory, loaded
// by the epilogue warps while the tile's MMAs were still running; every lane reads it with broadcast
// 16-byte loads), activation, pack, swizzled st.shared, TMA store.
// ADD: the addend tile of a chunk (32 rows x 128 bytes, the output tile's own box) is fetched by TMA straight into
// the staging tile that will hold the chunk's output, one chunk ahead (two barriers per warp, alternating); the lane
// then adds its row in place. A first version read the addend with per-lane global loads (32 different 128-byte
// lines per instruction) and cost as much as the element-wise add it replaced (profiles/conv_r2.md section 5).
template <int BLOCK_N, int EPI_BUFS, int ACT, bool BIAS, bool F32, bool STATS = false, bool ADD = false, int CSTEP = 1>
__device__ __forceinline__ void epilogue_tile(const GemmParams& p, const CUtensorMap* map_d, uint32_t tcol,
                                              uint32_t bufs, uint32_t tile_bytes, int& epi_buf, float* bias_s,
                                              int row0, int col_base, int ks, int lane,
                                              float (*st)[4] = nullptr, const CUtensorMap* map_add = nullptr,
                                              uint32_t add_bars = 0, uint32_t* add_count = nullptr, int c_first = 0) {
  static_assert(CSTEP == 1 || !ADD, "the addend prefetch assumes consecutive chunks");
  constexpr int kCols = F32 ? 32 : 64;
  if (row0 >= p.M) return;
  const uint32_t swz = (uint32_t)(lane & 7);
#pragma unroll 1
  for (int c = c_first; c < BLOCK_N / kCols; c += CSTEP) {
    const int col0 = col_base + c * kCols;
    if (col0 >= p.N) break;
    uint32_t r[kCols];
    if constexpr (ADD) {
      // addend tile of chunk c: issued here for the first chunk of the tile, else one chunk ago; then the next one
      if (lane == 0) {
        if (c == 0) {
          tma_store_wait_read_n<EPI_BUFS - 1>();                    // staging tile epi_buf is free
          const uint32_t bar = add_bars + 8u * (*add_count & 1u);
          mbar_expect_tx(bar, tile_bytes);
          tma_load_2d(bufs + (uint32_t)epi_buf * tile_bytes, map_add, bar, col0, row0);
        }
        if (c + 1 < BLOCK_N / kCols && col0 + kCols < p.N) {
          tma_store_wait_read_n<EPI_BUFS - 2>();                    // so is the next one
          const uint32_t bar = add_bars + 8u * ((*add_count + 1u) & 1u);
          mbar_expect_tx(bar, tile_bytes);
          tma_load_2d(bufs + (uint32_t)((epi_buf + 1) % EPI_BUFS) * tile_bytes, map_add, bar, col0 + kCols, row0);
        }
      }
      __syncwarp();
    }
    tc_ld_32x32(tcol + c * kCols, *reinterpret_cast<uint32_t(*)[32]>(&r[0]));
    if constexpr (!F32) tc_ld_32x32(tcol + c * kCols + 32, *reinterpret_cast<uint32_t(*)[32]>(&r[32]));
    tc_ld_wait();
    if constexpr (ADD) {
      // two loads alternate on two barriers: load k completes phase (k / 2) of barrier (k % 2)
      mbar_wait(add_bars + 8u * (*add_count & 1u), (*add_count >> 1) & 1u);
      const uint32_t arow = bufs + (uint32_t)epi_buf * tile_bytes + (uint32_t)lane * 128u;
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        uint32_t w[4];
        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3])
                     : "r"(arow + (((uint32_t)q ^ swz) << 4)) : "memory");
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          r[8 * q + 2 * e] = __float_as_uint(__uint_as_float(r[8 * q + 2 * e]) + bf16_lo(w[e]));
          r[8 * q + 2 * e + 1] = __float_as_uint(__uint_as_float(r[8 * q + 2 * e + 1]) + bf16_hi(w[e]));
        }
      }
      ++*add_count;
    }
#pragma unroll
    for (int q = 0; q < kCols / 4; ++q) {
      float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
      if constexpr (BIAS) b = *reinterpret_cast<const float4*>(bias_s + c * kCols + 4 * q);
      float x0 = __uint_as_float(r[4 * q]) + b.x, x1 = __uint_as_float(r[4 * q + 1]) + b.y;
      float x2 = __uint_as_float(r[4 * q + 2]) + b.z, x3 = __uint_as_float(r[4 * q + 3]) + b.w;
      if constexpr (ACT == 1) {
        x0 = fmaxf(x0, 0.f); x1 = fmaxf(x1, 0.f); x2 = fmaxf(x2, 0.f); x3 = fmaxf(x3, 0.f);
      } else if constexpr (ACT == 2) {
        x0 = apply_act(x0, 2); x1 = apply_act(x1, 2); x2 = apply_act(x2, 2); x3 = apply_act(x3, 2);
      }
      r[4 * q] = __float_as_uint(x0); r[4 * q + 1] = __float_as_uint(x1);
      r[4 * q + 2] = __float_as_uint(x2); r[4 * q + 3] = __float_as_uint(x3);
    }
    const uint32_t buf = bufs + (uint32_t)epi_buf * tile_bytes;
    if constexpr (!ADD) {
      if (lane == 0) tma_store_wait_read_n<EPI_BUFS - 1>();   // the store that last used this tile is done
    }
    __syncwarp();
    const uint32_t row_addr = buf + (uint32_t)lane * 128u;
#pragma unroll
    for (int q = 0; q < 8; ++q) {                      // eight 16-byte units per 128-byte row
      uint4 o;
      if constexpr (F32) {
        o = make_uint4(r[4 * q], r[4 * q + 1], r[4 * q + 2], r[4 * q + 3]);
      } else {
        o.x = pack_bf16(__uint_as_float(r[8 * q]), __uint_as_float(r[8 * q + 1]));
        o.y = pack_bf16(__uint_as_float(r[8 * q + 2]), __uint_as_float(r[8 * q + 3]));
        o.z = pack_bf16(__uint_as_float(r[8 * q + 4]), __uint_as_float(r[8 * q + 5]));
        o.w = pack_bf16(__uint_as_float(r[8 * q + 6]), __uint_as_float(r[8 * q + 7]));
      }
      st_shared_16(row_addr + (((uint32_t)q ^ swz) << 4), o);
    }
    fence_proxy_async_smem();
    __syncwarp();
    if (lane == 0) {
      if (p.splits > 1) tma_store_3d(map_d, buf, col0, row0, ks);
      else tma_store_2d(map_d, buf, col0, row0);
      tma_store_commit();
    }
    if constexpr (STATS) {
      // BatchNorm statistics of the values just staged (bf16, what the next layer reads): lane l sums
      // channels col0 + 2l, +1 over the tile's 32 rows = word l of every staged row (rows past M were
      // zero-filled by the loads and add nothing). The staging tile is reused only EPI_BUFS chunks later.
      const uint32_t ucol = (uint32_t)(lane >> 2), wsel = (uint32_t)(lane & 3) * 4u;
      // packed fp32x2 arithmetic (FADD2 / FFMA2): one add and one fma per pair of channels
      float2 sum2 = make_float2(0.f, 0.f), sq2 = make_float2(0.f, 0.f);
#pragma unroll
      for (int rr = 0; rr < 32; ++rr) {
        uint32_t w;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w)
                     : "r"(buf + (uint32_t)rr * 128u + ((ucol ^ (uint32_t)(rr & 7)) << 4) + wsel) : "memory");
        const float2 v = make_float2(bf16_lo(w), bf16_hi(w));
        sum2 = __fadd2_rn(sum2, v);
        sq2 = __ffma2_rn(v, v, sq2);
      }
      st[c][0] += sum2.x; st[c][1] += sq2.x; st[c][2] += sum2.y; st[c][3] += sq2.y;
    }
    epi_buf = (epi_buf + 1) % EPI_BUFS;
  }
}

// Pull variant: every rank walks the n tiles starting at its own shard, so the tensor cores work on local tiles
// while the copy engine gathers the first remote ones.
template <bool kPull, int BLOCK_N>
__device__ __forceinline__ int pull_n_order(const GemmParams& p, int t) {
  if constexpr (!kPull) return t;
  const int n = t + (p.rank * (p.rows_per_rank / BLOCK_N)) % p.num_n;
  return n >= p.num_n ? n - p.num_n : n;
}

template <int BLOCK_N, bool kPull, bool THIN, bool EPI8 = false>
__global__ void __launch_bounds__(EPI8 ? 384 : kGemmThreads, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap map_a,
                    const __grid_constant__ CUtensorMap map_b,
                    const __grid_constant__ CUtensorMap map_d,
                    const __grid_constant__ PullMaps pull, const GemmParams p) {
  using C = Cfg<BLOCK_N, THIN, kPull, EPI8>;
  // SWIZZLE_128B tiles need 1024-byte alignment (the kernel has no static shared memory, so the
  // dynamic segment starts at the CTA's 1024-byte aligned shared-memory base).
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t smem_base = smem_u32(smem_raw);
  const uint32_t a_smem = smem_base;
  const uint32_t b_smem = smem_base + C::kStages * C::kABytes;
  const uint32_t epi_smem = smem_base + C::kStages * C::kStageBytes;   // 1024-byte aligned
  const uint32_t epi_bias_smem = epi_smem + C::kEpiBytes;
  const uint32_t copy_smem = epi_bias_smem + C::kEpiBiasBytes;          // pull variant: kCopyBufs x kCopyPiece
  const uint32_t bar_base = copy_smem + C::kCopyBytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (C::kStages + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * C::kStages + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * C::kStages + kAccumStages + s); };
  auto copy_bar = [&](int s) { return bar_base + 8u * (2 * C::kStages + 2 * kAccumStages + s); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * C::kStages + 2 * kAccumStages + 4);
  const uint32_t add_bar_base = tmem_slot + 16u;          // 2 per epilogue warp (addend epilogue)
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x / 32;
  const int lane = threadIdx.x % 32;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_d) : "memory");
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < C::kStages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int s = 0; s < kAccumStages; ++s) {
      mbar_init(tfull_bar(s), 1);
      mbar_init(tempty_bar(s), C::kEpiWarps);  // one arrive per epilogue warp
    }
    for (int i = 0; i < kCopyBufs; ++i) mbar_init(copy_bar(i), 1);
    for (int i = 0; i < 8; ++i) mbar_init(add_bar_base + 8u * i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot),
                 "n"(C::kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  const int items = p.num_m * p.num_n * p.splits;

  if (warp == 0 && lane == 0) {
    // ================= TMA producer =================
    int stage = 0;
    uint32_t phase = 0;
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      const int ks = it % p.splits;
      const int tile = it / p.splits;
      const int m_blk = tile % p.num_m, n_blk = pull_n_order<kPull, BLOCK_N>(p, tile / p.num_m);
      const int kb0 = (int)((long long)ks * p.num_kb / p.splits);
      const int kb1 = (int)((long long)(ks + 1) * p.num_kb / p.splits);
      const CUtensorMap* bmap = &map_b;
      int b_row = n_blk * BLOCK_N;
      if constexpr (kPull) {
        // a tile another rank owns is readable locally once every CTA's copy warp has finished its share of it
        const int owner = b_row / p.rows_per_rank;
        if (owner != p.rank) {
          const uint32_t want = *p.epoch;
          unsigned long long polls = 0;
          while (!epoch_reached(ld_acquire_gpu(p.pull_ready + n_blk), want)) {
            __nanosleep(32);
            if (p.budget && ++polls > p.budget) { atomicExch(p.status, 5); break; }
          }
          asm volatile("fence.proxy.async.global;" ::: "memory");   // the TMA loads below read what bulk stores wrote
        }
      }
      // implicit-GEMM convolution, mode 1: the tile's first output pixel -> (image, window corner)
      int cv_n = 0, cv_h = 0, cv_w = 0;
      if (p.conv.mode == 1) {
        const int pix = m_blk * BLOCK_M, plane = p.conv.OH * p.conv.OW;
        cv_n = cdiv40(pix, p.conv.m_plane);
        const int rem = pix - cv_n * plane, oh = cdiv40(rem, p.conv.m_ow);
        cv_h = oh * p.conv.stride + p.conv.lower;
        cv_w = (rem - oh * p.conv.OW) * p.conv.stride + p.conv.lower_w;
      }
      // mode 1: (tap row, tap column, channel block) of the first k-block, then advanced incrementally
      int t_r = 0, t_s = 0, t_cb = 0;
      if (p.conv.mode == 1) {
        const int tap = cdiv40(kb0, p.conv.m_cblocks);
        t_cb = kb0 - tap * p.conv.cblocks;
        t_r = cdiv40(tap, p.conv.m_s);
        t_s = tap - t_r * p.conv.S;
      }
      // mode 2: the (tap, channel block) of each 64-column slab of this tile does not depend on the k-block
      int sl_c[BLOCK_N / 64], sl_s[BLOCK_N / 64], sl_r[BLOCK_N / 64];
      int w_n = 0, w_oh = 0, w_ow = 0;             // mode 2: image / output row / column of the k-block's first pixel
      if (p.conv.mode == 2) {
#pragma unroll
        for (int sl = 0; sl < BLOCK_N / 64; ++sl) {
          const int blk = (b_row >> 6) + sl;                   // 64-column block index along (tap, c)
          const int tap = cdiv40(blk, p.conv.m_cblocks), cb = blk - tap * p.conv.cblocks;
          const int r = cdiv40(tap, p.conv.m_s), sx = tap - r * p.conv.S;
          // slabs past the last tap (N tail) still have to deliver their bytes: clamp to a valid tap,
          // the epilogue's store clips those columns
          const bool in = tap < p.conv.R * p.conv.S;
          sl_c[sl] = in ? cb * 64 : 0; sl_s[sl] = in ? sx : 0; sl_r[sl] = in ? r : 0;
        }
        const int pix = kb0 * BLOCK_K, plane = p.conv.OH * p.conv.OW;
        w_n = cdiv40(pix, p.conv.m_plane);
        const int rem = pix - w_n * plane;
        w_oh = cdiv40(rem, p.conv.m_ow);
        w_ow = rem - w_oh * p.conv.OW;
      }
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(empty_bar(stage), phase ^ 1);
        mbar_expect_tx(full_bar(stage), C::kStageBytes);
        if (p.conv.mode == 1) {
          // reduction block kb = (tap, 64-channel block); taps in (r, s) order like the weight's K axis
          const int cb = t_cb, r = t_r, sx = t_s;
          if (++t_cb == p.conv.cblocks) { t_cb = 0; if (++t_s == p.conv.S) { t_s = 0; ++t_r; } }
          tma_load_im2col_4d(a_smem + stage * C::kABytes, &map_a, full_bar(stage), cb * 64, cv_w, cv_h, cv_n,
                             sx, r);
          if (!p.conv.flip) {
            tma_load_2d(b_smem + stage * C::kBBytes, bmap, full_bar(stage), kb * BLOCK_K, b_row);
          } else {
            // dgrad: dx[., ci] += dy[. + (pad' - r, pad' - s), co] * W[co, R-1-r, S-1-s, ci]; the weight
            // is a 3-D tensor (ci, tap, co): 64 co rows x 64 ci columns per box, MN-major B slabs
            const int wtap = (p.conv.wt_rb - p.conv.wt_rs * r) * p.conv.wt_S + (p.conv.wt_sb - p.conv.wt_ss * sx);
#pragma unroll
            for (int sl = 0; sl < BLOCK_N / 64; ++sl)
              tma_load_3d(b_smem + stage * C::kBBytes + sl * 8192, bmap, full_bar(stage), b_row + sl * 64,
                          wtap, cb * 64);
          }
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
          continue;
        }
        if (!p.a_mn) {
          tma_load_2d(a_smem + stage * C::kABytes, &map_a, full_bar(stage), kb * BLOCK_K,
                      m_blk * BLOCK_M);
        } else {
#pragma unroll
          for (int sl = 0; sl < BLOCK_M / 64; ++sl)           // 64-wide slabs of the [K, M] matrix
            tma_load_2d(a_smem + stage * C::kABytes + sl * 8192, &map_a, full_bar(stage),
                        m_blk * BLOCK_M + sl * 64, kb * BLOCK_K);
        }
        if (p.conv.mode == 2) {
          // wgrad: reduction block kb = 64 consecutive output pixels; each 64-column slab of the tile is
          // one (tap, 64-channel block) of the im2col matrix
          const int h = w_oh * p.conv.stride + p.conv.lower;
          const int w = w_ow * p.conv.stride + p.conv.lower_w;
#pragma unroll
          for (int sl = 0; sl < BLOCK_N / 64; ++sl)
            tma_load_im2col_4d(b_smem + stage * C::kBBytes + sl * 8192, bmap, full_bar(stage), sl_c[sl], w, h, w_n,
                               sl_s[sl], sl_r[sl]);
          // next k-block starts 64 output pixels further (row-major over ow, oh, n)
          w_ow += BLOCK_K;
          while (w_ow >= p.conv.OW) { w_ow -= p.conv.OW; if (++w_oh == p.conv.OH) { w_oh = 0; ++w_n; } }
        } else if (!p.b_mn) {
          tma_load_2d(b_smem + stage * C::kBBytes, bmap, full_bar(stage), kb * BLOCK_K, b_row);
        } else {
#pragma unroll
          for (int sl = 0; sl < BLOCK_N / 64; ++sl)
            tma_load_2d(b_smem + stage * C::kBBytes + sl * 8192, bmap, full_bar(stage), b_row + sl * 64,
                        kb * BLOCK_K);
        }
        if (++stage == C::kStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1 && lane == 0) {
    // ================= MMA issuer =================
    const uint32_t idesc = make_idesc_bf16(BLOCK_M, BLOCK_N, p.a_mn, p.b_mn);
    // advancing the reduction index by UMMA_K = 16: 32 bytes inside a K-major swizzle atom, two whole
    // 1024-byte atoms for an MN-major operand (descriptor addresses are in 16-byte units)
    const uint32_t a_step = p.a_mn ? 128u : 2u, b_step = p.b_mn ? 128u : 2u;
    int stage = 0;
    uint32_t phase = 0;
    int as = 0;
    uint32_t aphase = 0;
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      const int ks = it % p.splits;
      const int kb0 = (int)((long long)ks * p.num_kb / p.splits);
      const int kb1 = (int)((long long)(ks + 1) * p.num_kb / p.splits);
      mbar_wait(tempty_bar(as), aphase ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + as * BLOCK_N;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(full_bar(stage), phase);
        tc_fence_after();
        const uint64_t adesc = p.a_mn ? make_mnmajor_sw128_desc(a_smem + stage * C::kABytes)
                                      : make_kmajor_sw128_desc(a_smem + stage * C::kABytes);
        const uint64_t bdesc = p.b_mn ? make_mnmajor_sw128_desc(b_smem + stage * C::kBBytes)
                                      : make_kmajor_sw128_desc(b_smem + stage * C::kBBytes);
#pragma unroll
        for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
          tc_mma_bf16(tmem_d, adesc + a_step * k, bdesc + b_step * k, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
        tc_commit(empty_bar(stage));  // smem stage reusable once these MMAs have read it
        if (++stage == C::kStages) { stage = 0; phase ^= 1; }
      }
      tc_commit(tfull_bar(as));       // accumulator complete
      if (++as == kAccumStages) { as = 0; aphase ^= 1; }
    }
  } else if (kPull && warp == 3 && lane == 0) {
    // ================= copy engine (pull variant) =================
    // The remote n tiles (BLOCK_N rows x K, contiguous in the owner's and in the local buffer) in the order the GEMM
    // consumes them, cut into 8 KB pieces numbered across tiles; this CTA moves pieces blockIdx.x,
    // blockIdx.x + gridDim.x, ... through a four-deep staging ring (1-D bulk copies peer -> shared -> local).
    uint32_t owner_ready = 0;                     // owners whose PARAM_DONE flag has been observed
    const size_t tile_bytes = (size_t)BLOCK_N * p.K * 2;
    const int ppt = (int)((tile_bytes + kCopyPiece - 1) / kCopyPiece);      // pieces per tile
    const int own_blks = p.rows_per_rank / BLOCK_N, own_first = p.rank * own_blks;
    const int own_cnt = own_first < p.num_n ? (p.num_n - own_first < own_blks ? p.num_n - own_first : own_blks) : 0;
    const int rot = own_first % p.num_n;          // pull_n_order(): the GEMM starts at this rank's own tiles
    const int remote_tiles = p.num_n - own_cnt;
    const long long total = (long long)remote_tiles * ppt;
    const uint32_t epoch = *p.epoch;
    auto locate = [&](long long g, int* n_blk, size_t* off, uint32_t* sz) {
      const int j = (int)(g / ppt);                // j-th remote tile in consumption order
      int n = rot + own_cnt + j;                   // the tiles after mine, wrapping around
      if (n >= p.num_n) n -= p.num_n;
      *n_blk = n;
      *off = (size_t)(g - (long long)j * ppt) * kCopyPiece;
      *sz = (uint32_t)(tile_bytes - *off < (size_t)kCopyPiece ? tile_bytes - *off : (size_t)kCopyPiece);
    };
    auto issue_load = [&](long long g, int buf) {
      int n_blk; size_t off; uint32_t sz;
      locate(g, &n_blk, &off, &sz);
      const int owner = (n_blk * BLOCK_N) / p.rows_per_rank;
      if (!(owner_ready & (1u << owner))) {
        // shards written during the step that just ended
        if (!wait_epoch(p.flags + TB200_FLAG_PARAM_DONE(p.bucket, owner), epoch - 1, p.budget)) atomicExch(p.status, 5);
        owner_ready |= 1u << owner;
      }
      mbar_expect_tx(copy_bar(buf), sz);
      bulk_load_1d(copy_smem + buf * kCopyPiece, p.owner_base[owner] + (size_t)n_blk * tile_bytes + off, sz, copy_bar(buf));
    };
    // My pieces are g_k = blockIdx.x + k * gridDim.x. Ring of kCopyBufs staging buffers: loads run kCopyBufs - 1
    // pieces ahead of the stores; completed pieces are counted in batches of four behind ONE pair of fences
    // (a fence + counter update per piece made each piece cost ~5 us and held the engine at 400 GB/s of NVLink).
    const long long n_my = total > blockIdx.x ? (total - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    auto piece = [&](long long k) { return (long long)blockIdx.x + k * gridDim.x; };
    auto count_batch = [&](long long k0, long long k1) {      // pieces [k0, k1) of mine are complete in local memory
      if (k0 >= k1) return;
      asm volatile("fence.proxy.async.global;" ::: "memory");
      __threadfence();
      for (long long k = k0; k < k1; ++k) {
        int n_blk; size_t off; uint32_t sz;
        locate(piece(k), &n_blk, &off, &sz);
        const unsigned prev = atomicAdd(p.pull_cnt + n_blk, 1u);
        if (prev == (unsigned)ppt - 1) {
          p.pull_cnt[n_blk] = 0;                     // self-resetting: ready for the next launch
          __threadfence();
          st_release_gpu(p.pull_ready + n_blk, epoch);
        }
      }
    };
    for (long long k = 0; k < kCopyBufs - 1 && k < n_my; ++k) issue_load(piece(k), (int)k);
    long long counted = 0;
    for (long long k = 0; k < n_my; ++k) {
      const int buf = (int)(k % kCopyBufs);
      int n_blk; size_t off; uint32_t sz;
      locate(piece(k), &n_blk, &off, &sz);
      mbar_wait(copy_bar(buf), (uint32_t)(k / kCopyBufs) & 1u);
      fence_proxy_async_smem();
      bulk_store_1d(p.local_base + (size_t)n_blk * tile_bytes + off, copy_smem + buf * kCopyPiece, sz);
      tma_store_commit();
      const long long kn = k + kCopyBufs - 1;        // next load: into the buffer piece k - 1 was stored from
      if (kn < n_my) {
        if (k >= 1) tma_store_wait_read_n<1>();      // every store but the one just committed has read its buffer
        issue_load(piece(kn), (int)(kn % kCopyBufs));
      }
      if ((k & 3) == 3) {
        tma_store_wait_n<1>();                       // every store but the one just committed is complete
        count_batch(counted, k);
        counted = k;
      }
    }
    if (n_my > 0) {
      tma_store_wait_all();
      count_batch(counted, n_my);
    }
  } else if (warp >= 4) {
    // ================= epilogue =================
    const int ew = (warp - 4) & 3;   // == warp % 4: the TMEM lane quarter this warp may touch
    const int half = (warp - 4) >> 2;   // EPI8: this warp takes the 64-column chunks c with c % 2 == half
    int as = 0;
    uint32_t aphase = 0;
    int epi_buf = 0;
    uint32_t add_count = 0;          // addend tiles fetched so far by this warp (addend epilogue)
    // BatchNorm-statistics epilogue: per-lane (sum, sum of squares) of two channels per 64-column chunk, carried
    // across this CTA's tiles of one n block and written as one partial row per (CTA, warp) when it changes
    float st[BLOCK_N / 64][4];
#pragma unroll
    for (int c = 0; c < BLOCK_N / 64; ++c) st[c][0] = st[c][1] = st[c][2] = st[c][3] = 0.f;
    int st_n = -1;
    auto flush_stats = [&](int n_blk) {
      const size_t part = (size_t)blockIdx.x * 4 + ew;         // (EPI8: the two warps of a quarter share the row)
#pragma unroll
      for (int c = 0; c < BLOCK_N / 64; ++c) {
        const int col = n_blk * BLOCK_N + c * 64 + 2 * lane;
        if (col < p.N && (!EPI8 || (c & 1) == half))
          *reinterpret_cast<float4*>(p.stats + (part * p.N + col) * 2) = make_float4(st[c][0], st[c][1], st[c][2], st[c][3]);
        st[c][0] = st[c][1] = st[c][2] = st[c][3] = 0.f;
      }
    };
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      const int ks = it % p.splits;
      const int tile = it / p.splits;
      const int m_blk = tile % p.num_m, n_blk = pull_n_order<kPull, BLOCK_N>(p, tile / p.num_m);
      if (p.stats && n_blk != st_n) {
        if (st_n >= 0) flush_stats(st_n);
        st_n = n_blk;
      }
      if (!EPI8 && p.bias && p.splits == 1 && !p.direct_store) {
        // Stage this tile's bias row (fp32) before waiting for the accumulator: the loads overlap the
        // MMAs. Buffer `as` was last read two tiles ago; the barrier of the tile in between orders it.
        float* bt = reinterpret_cast<float*>(smem_raw + (epi_bias_smem - smem_u32(smem_raw))) + as * 256;
        for (int j = threadIdx.x - 128; j < BLOCK_N; j += 128) {
          const int col = n_blk * BLOCK_N + j;
          bt[j] = col < p.N ? load_bias(p.bias, p.bias_dtype, col) : 0.f;
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");     // the four epilogue warps only
      }
      mbar_wait(tfull_bar(as), aphase);
      tc_fence_after();
      const int row = m_blk * BLOCK_M + ew * 32 + lane;
      const bool row_ok = row < p.M;
      size_t orow = (size_t)(row_ok ? row : 0);      // output row (scatter mode: the pixel of the strided image)
      if (p.conv.sc_on && row_ok) {
        const int plane = p.conv.OH * p.conv.OW;
        const int n = cdiv40(row, p.conv.m_plane), rem = row - n * plane;
        const int a = cdiv40(rem, p.conv.m_ow), b = rem - a * p.conv.OW;
        orow = ((size_t)n * p.conv.sc_H + (size_t)(a * p.conv.sc_stride + p.conv.sc_ph)) * p.conv.sc_W +
               (size_t)(b * p.conv.sc_stride + p.conv.sc_pw);
      }
      if (!p.direct_store) {
        // ---- TMEM -> registers -> swizzled smem tile -> TMA store (coalesced, clipped at M / N) ----
        const int row0 = m_blk * BLOCK_M + ew * 32;
        const uint32_t tcol = tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(as * BLOCK_N);
        const uint32_t bufs = epi_smem + (uint32_t)((warp - 4) * C::kEpiBufs) * C::kEpiTileBytes;

        float* bias_s = reinterpret_cast<float*>(smem_raw + (epi_bias_smem - smem_u32(smem_raw))) + as * 256;
        const int col_base = n_blk * BLOCK_N;
        if constexpr (EPI8) {
          // statistics epilogue only (the host launches this instantiation for nothing else)
          epilogue_tile<BLOCK_N, C::kEpiBufs, 0, false, false, true, false, 2>(p, &map_d, tcol, bufs, C::kEpiTileBytes,
                                                                               epi_buf, bias_s, row0, col_base, ks, lane,
                                                                               st, nullptr, 0, nullptr, half);
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(tempty_bar(as));
          if (++as == kAccumStages) { as = 0; aphase ^= 1; }
          continue;
        }
        // One compact instruction stream per (activation, bias, output type): the element loop must not
        // carry per-element branches (a first version did and was instruction-fetch bound).
        const int variant = (p.splits > 1) ? 12 : (p.stats ? 13 : (p.addend ? 14 : (p.act * 4 + (p.bias ? 2 : 0) + (p.d_dtype == kF32 ? 1 : 0))));
        switch (variant) {
          case 14:
            epilogue_tile<BLOCK_N, C::kEpiBufs, 0, false, false, false, true>(p, &map_d, tcol, bufs, C::kEpiTileBytes,
                                                                              epi_buf, bias_s, row0, col_base, ks, lane,
                                                                              nullptr, &pull.addend,
                                                                              add_bar_base + 16u * ew, &add_count);
            break;
          case 13:
            epilogue_tile<BLOCK_N, C::kEpiBufs, 0, false, false, true>(p, &map_d, tcol, bufs, C::kEpiTileBytes, epi_buf,
                                                                       bias_s, row0, col_base, ks, lane, st);
            break;
#define TB_EPI(ID, ACT, BIAS, F32)                                                                        \
  case ID:                                                                                                \
    epilogue_tile<BLOCK_N, C::kEpiBufs, ACT, BIAS, F32>(p, &map_d, tcol, bufs, C::kEpiTileBytes, epi_buf, bias_s, row0, \
                                           col_base, ks, lane);                                           \
    break;
          TB_EPI(0, 0, false, false) TB_EPI(1, 0, false, true) TB_EPI(2, 0, true, false) TB_EPI(3, 0, true, true)
          TB_EPI(4, 1, false, false) TB_EPI(5, 1, false, true) TB_EPI(6, 1, true, false) TB_EPI(7, 1, true, true)
          TB_EPI(8, 2, false, false) TB_EPI(9, 2, false, true) TB_EPI(10, 2, true, false) TB_EPI(11, 2, true, true)
          TB_EPI(12, 0, false, true)
#undef TB_EPI
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(tempty_bar(as));
        if (++as == kAccumStages) { as = 0; aphase ^= 1; }
        continue;
      }

#pragma unroll 1
      for (int c = 0; c < BLOCK_N / 32; ++c) {
        uint32_t r[32];
        tc_ld_32x32(tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(as * BLOCK_N + c * 32), r);
        tc_ld_wait();
        const int col0 = n_blk * BLOCK_N + c * 32;
        if (!row_ok || col0 >= p.N) continue;
        if (p.splits > 1) {
          float* out = p.partials + ((size_t)ks * p.M + row) * p.N + col0;
          if (col0 + 32 <= p.N && (p.N % 4) == 0) {
#pragma unroll
            for (int j = 0; j < 32; j += 4)
              *reinterpret_cast<float4*>(out + j) =
                  make_float4(__uint_as_float(r[j]), __uint_as_float(r[j + 1]),
                              __uint_as_float(r[j + 2]), __uint_as_float(r[j + 3]));
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (col0 + j < p.N) out[j] = __uint_as_float(r[j]);
          }
          continue;
        }
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          float x = __uint_as_float(r[j]);
          if (p.bias && col0 + j < p.N) x += load_bias(p.bias, p.bias_dtype, col0 + j);
          v[j] = apply_act(x, p.act);
        }
        if (p.d_dtype == kBF16) {
          __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(p.D) + orow * p.N + col0;
          if (col0 + 32 <= p.N && (p.N % 8) == 0) {
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
              uint4 o;
              o.x = pack_bf16(v[j], v[j + 1]); o.y = pack_bf16(v[j + 2], v[j + 3]);
              o.z = pack_bf16(v[j + 4], v[j + 5]); o.w = pack_bf16(v[j + 6], v[j + 7]);
              *reinterpret_cast<uint4*>(out + j) = o;
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (col0 + j < p.N) out[j] = __float2bfloat16(v[j]);
          }
        } else {
          float* out = reinterpret_cast<float*>(p.D) + orow * p.N + col0;
          if (col0 + 32 <= p.N && (p.N % 4) == 0) {
#pragma unroll
            for (int j = 0; j < 32; j += 4)
              *reinterpret_cast<float4*>(out + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (col0 + j < p.N) out[j] = v[j];
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar(as));
      if (++as == kAccumStages) { as = 0; aphase ^= 1; }
    }
    if (p.stats && st_n >= 0) flush_stats(st_n);
    if (lane == 0) tma_store_wait_all();     // smem must outlive the last bulk stores
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                 "n"(C::kTmemCols)
                 : "memory");
  }
}

// ---- probe: shifted K-major operand windows ---------------------------------------------------------
// Groundwork for convolutions that fetch an input halo tile ONCE and form the R x S filter taps as shifted
// windows of it in shared memory (profiles/conv_r1g.md: the im2col-TMA kernel re-fetches the tile per tap
// and is bound by L2->SM traffic). A window that starts `shift` pixel rows (128 bytes each) into a
// 128-byte-swizzled tile is no longer aligned to the 1024-byte swizzle pattern; the descriptor's "matrix
// base offset" field (bits 49-51) exists for that case. This one-CTA kernel measures what the hardware
// does: D[128, 64] = A[shift : shift+128, 0:64] . B[64, 64]^T with the A descriptor advanced by shift*128
// bytes and base offset = `base_offset`. The host compares against the same product in fp32.
__global__ void __launch_bounds__(128, 1)
probe_shift_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                   float* __restrict__ D, int shift, int base_offset) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t a_smem = smem_u32(smem_raw);                 // 256 rows x 128 B
  const uint32_t b_smem = a_smem + 256 * 128;                 // 64 rows x 128 B
  const uint32_t bar_full = b_smem + 64 * 128;
  const uint32_t bar_done = bar_full + 8;
  const uint32_t tmem_slot = bar_full + 16;
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  if (threadIdx.x == 0) {
    mbar_init(bar_full, 1);
    mbar_init(bar_done, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(64)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - a_smem));
  if (threadIdx.x == 0) {
    mbar_expect_tx(bar_full, 256 * 128 + 64 * 128);
    tma_load_2d(a_smem, &map_a, bar_full, 0, 0);
    tma_load_2d(a_smem + 128 * 128, &map_a, bar_full, 0, 128);
    tma_load_2d(b_smem, &map_b, bar_full, 0, 0);
    mbar_wait(bar_full, 0);
    tc_fence_after();
    const uint32_t idesc = make_idesc_bf16(128, 64);
    const uint64_t adesc = make_kmajor_sw128_desc(a_smem + (uint32_t)shift * 128u) |
                           ((uint64_t)(base_offset & 7) << 49);
    const uint64_t bdesc = make_kmajor_sw128_desc(b_smem);
#pragma unroll
    for (int k = 0; k < BLOCK_K / UMMA_K; ++k) tc_mma_bf16(tmem_base, adesc + 2u * k, bdesc + 2u * k, idesc, k > 0);
    tc_commit(bar_done);
  }
  mbar_wait(bar_done, 0);
  tc_fence_after();
  const int row = warp * 32 + lane;
#pragma unroll 1
  for (int c = 0; c < 2; ++c) {
    uint32_t r[32];
    tc_ld_32x32(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(c * 32), r);
    tc_ld_wait();
#pragma unroll
    for (int j = 0; j < 32; ++j) D[(size_t)row * 64 + c * 32 + j] = __uint_as_float(r[j]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(64) : "memory");
  }
}

// Split-K second pass: sum the fp32 partials in split order, then bias/activation/cast.
__global__ void __launch_bounds__(256) splitk_reduce_kernel(const float* __restrict__ partials,
                                                            int splits, size_t mn, int N,
                                                            const void* __restrict__ bias,
                                                            int bias_dtype, int act,
                                                            void* __restrict__ D, int d_dtype) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < mn;
       i += (size_t)gridDim.x * blockDim.x) {
    float acc = 0.f;
    for (int s = 0; s < splits; ++s) acc += partials[(size_t)s * mn + i];
    if (bias) acc += load_bias(bias, bias_dtype, i % N);
    acc = apply_act(acc, act);
    if (d_dtype == kBF16) reinterpret_cast<__nv_bfloat16*>(D)[i] = __float2bfloat16(acc);
    else reinterpret_cast<float*>(D)[i] = acc;
  }
}

// ---- host side --------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const int*, const int*, cuuint32_t, cuuint32_t,
                                   const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeIm2colFn get_encode_im2col() {
  static EncodeIm2colFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeIm2colFn>(p);
  });
  return fn;
}

// im2col-mode map over an NHWC bf16 activation tensor [n, h, w, c] for an R x S filter with the given
// stride and padding: every load delivers `pixels` consecutive output positions x 64 channels of one tap.
// Bounding box of window corners: lower = -pad, upper = pad - (taps - 1) (cuTensorMapEncodeIm2col's
// convention: corners relative to the tensor's first / last element).
int make_im2col_map(CUtensorMap* out, const void* base, int n, int h, int w, int c, int R, int S, int stride,
                    int pad, int pixels, const int* corners = nullptr /* lower_w, lower_h, upper_w, upper_h */) {
  EncodeIm2colFn enc = get_encode_im2col();
  if (!enc) return fail("cuTensorMapEncodeIm2col entry point not available%s");
  cuuint64_t dims[4] = {(cuuint64_t)c, (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)n};
  cuuint64_t strides[3] = {(cuuint64_t)c * 2, (cuuint64_t)w * c * 2, (cuuint64_t)h * w * c * 2};
  int lower[2] = {-pad, -pad};
  int upper[2] = {pad - (S - 1), pad - (R - 1)};
  if (corners) { lower[0] = corners[0]; lower[1] = corners[1]; upper[0] = corners[2]; upper[1] = corners[3]; }
  cuuint32_t estr[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, lower,
                   upper, 64, (cuuint32_t)pixels, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail("cuTensorMapEncodeIm2col failed (%s code %ld)", "", (long)r);
  // Drivers up to 13.1 encode im2col maps of tensors smaller than 128 KiB with a flag that makes the
  // loads fault; CUTLASS (cute/atom/copy_traits_sm90_im2col.hpp) clears the same bit.
  int driver = 0;
  if (cudaDriverGetVersion(&driver) == cudaSuccess && driver <= 13010 &&
      (size_t)n * h * w * c * 2 < 131072)
    reinterpret_cast<uint64_t*>(out)[1] &= ~(1ull << 21);
  return 0;
}

// [cout][taps][cin] bf16 weights as a 3-D tensor for the dgrad's MN-major B slabs: box = 64 cout rows of
// one tap x 64 cin columns.
int make_weight_tap_map(CUtensorMap* out, const void* base, int cout, int taps, int cin) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return fail("cuTensorMapEncodeTiled entry point not available%s");
  cuuint64_t dims[3] = {(cuuint64_t)cin, (cuuint64_t)taps, (cuuint64_t)cout};
  cuuint64_t strides[2] = {(cuuint64_t)cin * 2, (cuuint64_t)taps * cin * 2};
  cuuint32_t box[3] = {64, 1, (cuuint32_t)BLOCK_K};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box,
                   estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail("cuTensorMapEncodeTiled (weight taps) failed (%s code %ld)", "", (long)r);
  return 0;
}

// 2-D bf16 row-major [rows, cols] tensor, box = [box_rows, 64 cols], 128-byte swizzle.
int make_map(CUtensorMap* out, const void* base, int rows, int cols, int box_rows) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return fail("cuTensorMapEncodeTiled entry point not available%s");
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)cols * 2};
  cuuint32_t box[2] = {(cuuint32_t)BLOCK_K, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides,
                   box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail("cuTensorMapEncodeTiled failed (%s code %ld)", "", (long)r);
  return 0;
}

// [k_rows, mn_cols] bf16 row-major matrix of an MN-major operand: box = 64 reduction rows x 64 columns.
int make_map_mn(CUtensorMap* out, const void* base, int k_rows, int mn_cols) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return fail("cuTensorMapEncodeTiled entry point not available%s");
  cuuint64_t dims[2] = {(cuuint64_t)mn_cols, (cuuint64_t)k_rows};
  cuuint64_t strides[1] = {(cuuint64_t)mn_cols * 2};
  cuuint32_t box[2] = {64, (cuuint32_t)BLOCK_K};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box,
                   estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail("cuTensorMapEncodeTiled (MN-major) failed (%s code %ld)", "", (long)r);
  return 0;
}

// Output tensor map for the epilogue's TMA stores: rows of `cols` elements, box = 32 rows x 128 bytes,
// 128-byte swizzle. `planes` > 1 adds a third dimension (split-K partial planes of M x N fp32).
int make_out_map(CUtensorMap* out, const void* base, int rows, int cols, bool f32, int planes) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return fail("cuTensorMapEncodeTiled entry point not available%s");
  const cuuint64_t esz = f32 ? 4 : 2;
  cuuint64_t dims[3] = {(cuuint64_t)cols, (cuuint64_t)rows, (cuuint64_t)planes};
  cuuint64_t strides[2] = {(cuuint64_t)cols * esz, (cuuint64_t)cols * esz * (cuuint64_t)rows};
  cuuint32_t box[3] = {(cuuint32_t)(128 / esz), 32, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  const int rank = planes > 1 ? 3 : 2;
  CUresult r = enc(out, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank,
                   const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail("cuTensorMapEncodeTiled (output) failed (%s code %ld)", "", (long)r);
  return 0;
}

struct Workspace {
  float* ptr = nullptr;
  size_t bytes = 0;
};
std::mutex g_ws_mutex;
// split-K scratch, grown on demand, one per (device, stream): jobs that share a GPU (exec/ packs up to three
// single-GPU jobs per device, each on its own streams) must not share partial sums
std::unordered_map<unsigned long long, Workspace> g_ws;

int sm_count() { return device_sm_count(); }

int get_workspace(size_t bytes, cudaStream_t stream, float** out) {
  int dev = 0;
  TB_CUDA_OK(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lock(g_ws_mutex);
  Workspace& w = g_ws[((unsigned long long)(uintptr_t)stream << 6) ^ (unsigned long long)dev];
  if (w.bytes < bytes) {
    // Growing frees the old block; callers on other streams must not be using it. The
    // Python layer sizes the workspace up-front (tb200_gemm_reserve) before graph capture.
    if (w.ptr) TB_CUDA_OK(cudaFree(w.ptr));
    w.ptr = nullptr;
    w.bytes = 0;
    TB_CUDA_OK(cudaMalloc(&w.ptr, bytes));
    w.bytes = bytes;
  }
  *out = w.ptr;
  return 0;
}

// Tile width and split-K factor of a problem (shared by the launcher and tb200_gemm_plan_splits).
//   block_n: 256 when that still yields at least one full wave of CTAs, else 128.
//   splits : 1 when the tile grid fills the machine. Otherwise the count whose work-item total packs best
//            into whole waves of `sms` CTAs (ties go to fewer splits), bounded by
//              - at least 4 k-blocks (256 reduction elements) per split:  splits <= ceil(K/64) / 4
//              - one split per SM:                                          splits <= sms
//              - the fp32 partial planes may add at most ~10 % to the bf16 input traffic, with a floor of
//                16 (a long-reduction, tiny-output GEMM such as a 1x1 convolution's dW wants one split per
//                SM; a square-ish problem wants few):   splits <= max(16, 0.1 * in_bytes / plane_bytes)
int plan_tiles(int M, int N, int K, bool pull, int rows_per_rank, int sms, int* block_n_out, int* splits_out) {
  const int num_kb = (K + BLOCK_K - 1) / BLOCK_K;
  const int num_m = (M + BLOCK_M - 1) / BLOCK_M;
  int block_n = 256;
  if (N <= 128 || (long long)num_m * ((N + 255) / 256) < sms) block_n = 128;
  if (pull && rows_per_rank % block_n) {
    block_n = 128;
    if (rows_per_rank % 128)
      return fail("tb200_gemm_pull: rows_per_rank %s(%ld) must be a multiple of 128", "", rows_per_rank);
  }
  const int tiles = num_m * ((N + block_n - 1) / block_n);
  int splits = 1;
  if (tiles < sms) {
    int max_splits = num_kb / 4;
    if (max_splits > sms) max_splits = sms;
    const double in_bytes = 2.0 * ((double)M * K + (double)N * K);
    const double plane_bytes = 4.0 * (double)M * N;
    int by_traffic = (int)(0.1 * in_bytes / plane_bytes);
    if (by_traffic < 16) by_traffic = 16;
    if (max_splits > by_traffic) max_splits = by_traffic;
    double best = 0.0;
    for (int s = 1; s <= max_splits; ++s) {
      const int it = tiles * s;
      const int waves = (it + sms - 1) / sms;
      const double eff = (double)it / ((double)waves * sms);
      if (eff > best + 1e-9) { best = eff; splits = s; }
    }
  }
  *block_n_out = block_n;
  *splits_out = splits;
  return 0;
}

}  // namespace
#include "gemm_2cta.cuh"
namespace {

// 0 = never, 1 = whenever the problem is eligible, -1 = eligible and large enough to fill the pairs
// (TB200_GEMM_2CTA or tb200_gemm_set_pair_mode).
int g_pair_mode = -2;
int pair_mode() {
  if (g_pair_mode == -2) {
    const char* e = getenv("TB200_GEMM_2CTA");
    g_pair_mode = e ? atoi(e) : -1;
  }
  return g_pair_mode;
}

// Split-K factor of the 2-CTA kernel: 256 x 256 tiles over sms / 2 pairs, same bounds as plan_tiles().
int plan_pair_splits(int M, int N, int K, int pairs) {
  const int num_kb = (K + BLOCK_K - 1) / BLOCK_K;
  const int tiles = ((M + 255) / 256) * ((N + 255) / 256);
  int splits = 1;
  if (tiles < pairs) {
    int max_splits = num_kb / 4;
    if (max_splits > pairs) max_splits = pairs;
    const double in_bytes = 2.0 * ((double)M * K + (double)N * K);
    const double plane_bytes = 4.0 * (double)M * N;
    int by_traffic = (int)(0.1 * in_bytes / plane_bytes);
    if (by_traffic < 16) by_traffic = 16;
    if (max_splits > by_traffic) max_splits = by_traffic;
    double best = 0.0;
    for (int s = 1; s <= max_splits; ++s) {
      const int it = tiles * s;
      const int waves = (it + pairs - 1) / pairs;
      const double eff = (double)it / ((double)waves * pairs);
      if (eff > best + 1e-9) { best = eff; splits = s; }
    }
  }
  return splits;
}

int launch_gemm_pair(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& md, const PullMaps& extra,
                     const GemmParams& p, int pairs, cudaStream_t stream) {
  static DeviceOnce configured;
  if (int rc = ensure_dynamic_smem(configured, pair::gemm2_kernel, pair::kSmemBytes)) return rc;
  pair::gemm2_kernel<<<2 * pairs, kGemmThreads, pair::kSmemBytes, stream>>>(ma, mb, md, extra, p);
  return launched();
}

template <int BLOCK_N, bool kPull, bool THIN, bool EPI8 = false>
int launch_gemm(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& md, const PullMaps& pull,
                const GemmParams& p, int grid, cudaStream_t stream) {
  using C = Cfg<BLOCK_N, THIN, kPull, EPI8>;
  auto kern = gemm_tcgen05_kernel<BLOCK_N, kPull, THIN, EPI8>;
  static DeviceOnce configured;      // one per template instantiation
  if (int rc = ensure_dynamic_smem(configured, kern, C::kSmemBytes)) return rc;
  kern<<<grid, C::kThreads, C::kSmemBytes, stream>>>(ma, mb, md, pull, p);
  return launched();
}

// TB200_STATS_EPI8=0 keeps the statistics epilogue on four warps (A/B runs)
bool stats_epi8() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("TB200_STATS_EPI8");
    v = (e && atoi(e) == 0) ? 0 : 1;
  }
  return v == 1;
}

// Host-side description of the im2col operand of an implicit-GEMM convolution (see ConvGeom).
struct ConvHost {
  ConvGeom g;
  const void* act;          // NHWC tensor the im2col map walks (x for fprop / wgrad, dy for dgrad)
  int n, h, w, c;           // its extent
  int pad;                  // padding of the walk (dgrad: taps - 1 - pad)
  int wcout, wtaps, wcin;   // weight extent (flip only)
  int has_corners;          // explicit im2col bounding box (strided-dgrad classes)
  int corners[4];
};

int gemm_common(const tb200_world_t* world, int bucket, size_t b_off, int rows_per_rank,
                const void* A, const void* B, void* D, const void* bias, int bias_dtype, int M,
                int N, int K, int d_dtype, int act, const uint32_t* epoch_dev, int* status_dev,
                unsigned long long budget, cudaStream_t stream, float* fixed_scratch = nullptr,
                size_t fixed_scratch_bytes = 0, int a_mn = 0, int b_mn = 0, const ConvHost* conv = nullptr,
                float* stats = nullptr, size_t stats_floats = 0, int* stats_parts = nullptr,
                uint32_t* pull_sync = nullptr, const void* addend = nullptr) {
  const bool pull = world != nullptr;
  if (stats_parts) *stats_parts = 0;
  if (M <= 0 || N <= 0 || K <= 0) return fail("tb200_gemm: empty problem%s");
  // cuTensorMapEncodeTiled is a driver call and needs the primary context bound to THIS thread; a thread
  // that has made no runtime call yet (e.g. autograd's backward thread, when the GEMM is the first thing
  // it runs) has none. cudaFree(0) binds it; once per thread.
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) {
    TB_CUDA_OK(cudaFree(nullptr));
    ctx_bound = true;
  }
  // TMA needs 16-byte row strides: the contiguous dimension of each operand must be a multiple of 8
  if (!a_mn && K % 8) return fail("tb200_gemm: K %s(%ld) must be a multiple of 8", "", K);
  if (!b_mn && K % 8) return fail("tb200_gemm: K %s(%ld) must be a multiple of 8", "", K);
  if (a_mn && M % 8) return fail("tb200_gemm: M %s(%ld) must be a multiple of 8 for an MN-major A", "", M);
  if (b_mn && N % 8) return fail("tb200_gemm: N %s(%ld) must be a multiple of 8 for an MN-major B", "", N);
  if (pull && (a_mn || b_mn)) return fail("tb200_gemm_pull: operands must be K-major%s");
  if (d_dtype != TB200_BF16 && d_dtype != TB200_F32) return fail("tb200_gemm: bad output dtype%s");

  GemmParams p{};
  p.M = M; p.N = N; p.K = K;
  p.D = D; p.bias = bias; p.bias_dtype = bias_dtype; p.d_dtype = d_dtype; p.act = act;
  p.a_mn = a_mn; p.b_mn = b_mn;
  p.addend = addend;
  if (addend && (bias || act || d_dtype != TB200_BF16 || N % 8 || stats || pull || (conv && conv->g.sc_on)))
    return fail("tb200_gemm: the addend epilogue takes a bf16 output with N %% 8 == 0, no bias, no activation%s");
  if (conv) {
    p.conv = conv->g;
    auto magic40 = [](long long d) { return (unsigned long long)(((1ull << 40) + (unsigned long long)d - 1) / (unsigned long long)d); };
    const long long plane = (long long)conv->g.OH * conv->g.OW;
    if ((double)M * (double)plane >= 1.0e12 && conv->g.mode == 1) return fail("tb200_conv2d: problem too large%s");
    if ((double)K * (double)plane >= 1.0e12 && conv->g.mode == 2) return fail("tb200_conv2d: problem too large%s");
    p.conv.m_plane = magic40(plane);
    p.conv.m_ow = magic40(conv->g.OW);
    p.conv.m_cblocks = magic40(conv->g.cblocks);
    p.conv.m_s = magic40(conv->g.S);
    if (!conv->has_corners) p.conv.lower_w = conv->g.lower;
    if (conv->g.flip && !conv->g.wt_S) {          // stride-1 dgrad: plain flip of both tap axes
      p.conv.wt_rb = conv->g.R - 1; p.conv.wt_rs = 1; p.conv.wt_sb = conv->g.S - 1; p.conv.wt_ss = 1; p.conv.wt_S = conv->g.S;
    }
  }
  p.num_kb = (K + BLOCK_K - 1) / BLOCK_K;
  p.num_m = (M + BLOCK_M - 1) / BLOCK_M;

  const int sms = sm_count();
  int block_n = 0, splits = 1;
  // 2-CTA kernel (gemm_2cta.cuh): plain GEMMs with a TMA-storable output; the allocation-free entry point keeps
  // the one-CTA plan its documentation states
  bool use_pair = false;
  if (!pull && !conv && !stats && !fixed_scratch && pair_mode() != 0 && sms >= 2) {
    const int sp = plan_pair_splits(M, N, K, sms / 2);
    const bool f32 = sp > 1 || d_dtype == TB200_F32;
    const bool storable = (N % (f32 ? 4 : 8)) == 0;
    // measured (profiles/gemm_pair_r2.md): the pair tiles win on long reductions (8192^3 1.17x cuBLAS, one-CTA
    // 0.97x) and lose on short ones, which are bound by the epilogue's stores (the one-CTA kernel has the deeper
    // store pipeline); and they must not split K where the one-CTA plan would not (partial planes cost more
    // than the idle pairs)
    bool big = M >= 256 && N >= 256 && K >= 1024;
    if (big && sp > 1) {
      int bn1 = 0, sp1 = 1;
      if (plan_tiles(M, N, K, false, 0, sms, &bn1, &sp1) || sp1 == 1) big = false;
    }
    if (storable && (pair_mode() == 1 || big)) {
      use_pair = true;
      block_n = 256;
      splits = sp;
      p.num_m = (M + 255) / 256;
    }
  }
  if (!use_pair)
    if (int rc = plan_tiles(M, N, K, pull, rows_per_rank, sms, &block_n, &splits)) return rc;
  if (conv && conv->g.sc_on) splits = 1;      // the split-K second pass writes rows in order, not scattered
  if (addend) splits = 1;                     // the addend is applied by the GEMM's own epilogue
  p.num_n = (N + block_n - 1) / block_n;
  const int tiles = p.num_m * p.num_n;
  p.splits = splits;
  if (splits > 1) {
    const size_t need = (size_t)splits * M * N * sizeof(float);
    if (fixed_scratch) {
      // allocation-free entry point: the partial sums must fit the caller's fixed scratch. Every partial
      // the reduction reads is written by this call's GEMM first, so nothing has to be cleared.
      if (need > fixed_scratch_bytes) {
        snprintf(g_last_error, sizeof(g_last_error),
                 "tb200_gemm_bf16_nt_noalloc: M=%d N=%d K=%d needs %zu bytes of split-K scratch "
                 "(%d splits) but the built-in scratch is %zu bytes; use tb200_gemm_bf16_nt",
                 M, N, K, need, splits, fixed_scratch_bytes);
        return 4;
      }
      p.partials = fixed_scratch;
    } else if (int rc = get_workspace(need, stream, &p.partials)) {
      return rc;
    }
  }
  // TMA-store epilogue needs a 16-byte aligned row stride; odd widths fall back to direct stores.
  const bool f32_out = splits > 1 || d_dtype == TB200_F32;
  p.direct_store = (N % (f32_out ? 4 : 8)) != 0 || (conv && conv->g.sc_on);

  CUtensorMap ma, mb, md;
  memset(&md, 0, sizeof(md));
  if (!p.direct_store) {
    if (splits > 1) {
      if (int rc = make_out_map(&md, p.partials, M, N, true, splits)) return rc;
    } else if (int rc = make_out_map(&md, D, M, N, d_dtype == TB200_F32, 1)) {
      return rc;
    }
  }
  PullMaps pm;
  memset(&pm, 0, sizeof(pm));
  if (addend)
    if (int rc = make_out_map(&pm.addend, addend, M, N, false, 1)) return rc;
  if (conv && conv->g.mode == 1) {
    if (int rc = make_im2col_map(&ma, conv->act, conv->n, conv->h, conv->w, conv->c, conv->g.R, conv->g.S,
                                 conv->g.stride, conv->pad, BLOCK_M, conv->has_corners ? conv->corners : nullptr))
      return rc;
  } else if (int rc = a_mn ? make_map_mn(&ma, A, K, M) : make_map(&ma, A, M, K, BLOCK_M)) {
    return rc;
  }
  if (conv && conv->g.mode == 2) {
    if (int rc = make_im2col_map(&mb, conv->act, conv->n, conv->h, conv->w, conv->c, conv->g.R, conv->g.S,
                                 conv->g.stride, conv->pad, BLOCK_K))
      return rc;
  } else if (conv && conv->g.flip) {
    if (int rc = make_weight_tap_map(&mb, B, conv->wcout, conv->wtaps, conv->wcin)) return rc;
  } else if (pull) {
    p.rows_per_rank = rows_per_rank;
    p.world = world->world;
    p.rank = world->rank;
    p.bucket = bucket;
    p.epoch = epoch_dev;
    p.flags = world->flags[world->rank];
    p.status = status_dev;
    p.budget = budget;
    if ((long long)rows_per_rank * world->world < N)
      return fail("tb200_gemm_pull: shards do not cover N%s");
    if (!pull_sync) return fail("tb200_gemm_pull: no synchronisation scratch%s");
    // every rank's copy of B has the same [rows_per_rank * world, K] layout at element offset b_off of its
    // parameter buffer; the copy engine moves whole n tiles between them
    for (int s = 0; s < world->world; ++s)
      p.owner_base[s] = reinterpret_cast<const uint8_t*>(world->param[s]) + b_off * 2;
    p.local_base = reinterpret_cast<uint8_t*>(world->param[world->rank]) + b_off * 2;
    p.pull_cnt = pull_sync;
    p.pull_ready = pull_sync + TB200_PULL_SYNC_WORDS / 2;
    if ((N + block_n - 1) / block_n > TB200_PULL_SYNC_WORDS / 2)
      return fail("tb200_gemm_pull: N %s(%ld) has more column tiles than the synchronisation scratch holds", "", N);
    const __nv_bfloat16* local = reinterpret_cast<const __nv_bfloat16*>(world->param[world->rank]) + b_off;
    if (int rc = make_map(&pm.b_local, local, N, K, block_n)) return rc;
    mb = pm.b_local;
  } else {
    // (the pair kernel's B box is this CTA's 128 rows of the 256-wide tile)
    if (int rc = b_mn ? make_map_mn(&mb, B, K, N) : make_map(&mb, B, N, K, use_pair ? 128 : block_n)) return rc;
  }

  const int items = tiles * splits;
  const int grid = items < sms ? items : sms;
  // BatchNorm-statistics epilogue: only the plain bf16 TMA-store epilogue carries it; a problem that had to be
  // split along K (tiny batch) simply reports no statistics and the caller runs its own pass
  if (stats && splits == 1 && !p.direct_store && d_dtype == TB200_BF16 && !bias && act == 0 && !pull) {
    const size_t need = (size_t)grid * 4 * N * 2;
    if (stats_floats < need) return fail("tb200_gemm: statistics buffer too small (%s%ld floats needed)", "", (long)need);
    TB_CUDA_OK(cudaMemsetAsync(stats, 0, need * sizeof(float), stream));
    p.stats = stats;
    if (stats_parts) *stats_parts = grid * 4;
  }
  int rc;
  const bool thin = !pull && p.num_kb <= 4 && splits == 1;   // store-bound: deeper epilogue pipeline
  if (use_pair) {
    rc = launch_gemm_pair(ma, mb, md, pm, p, items < sms / 2 ? items : sms / 2, stream);
  } else if (p.stats && stats_epi8() && block_n == 256) {
    // eight epilogue warps, two per lane quarter on alternating 64-column chunks: measured on the ResNet-50 forward
    // convolutions (profiles/conv_r2.md section 5) 119 -> 90 us on layer1's 64 -> 256 and 67 -> 55 us on layer2's
    // 128 -> 512; 128-wide tiles have too few chunks per quarter to share and lose the deeper store pipeline
    rc = thin ? launch_gemm<256, false, true, true>(ma, mb, md, pm, p, grid, stream)
              : launch_gemm<256, false, false, true>(ma, mb, md, pm, p, grid, stream);
  } else if (block_n == 256) {
    rc = pull ? launch_gemm<256, true, false>(ma, mb, md, pm, p, grid, stream)
              : (thin ? launch_gemm<256, false, true>(ma, mb, md, pm, p, grid, stream)
                      : launch_gemm<256, false, false>(ma, mb, md, pm, p, grid, stream));
  } else {
    rc = pull ? launch_gemm<128, true, false>(ma, mb, md, pm, p, grid, stream)
              : (thin ? launch_gemm<128, false, true>(ma, mb, md, pm, p, grid, stream)
                      : launch_gemm<128, false, false>(ma, mb, md, pm, p, grid, stream));
  }
  if (rc) return rc;
  if (splits > 1) {
    const size_t mn = (size_t)M * N;
    int rgrid = (int)((mn + 255) / 256);
    if (rgrid > sms * 8) rgrid = sms * 8;
    splitk_reduce_kernel<<<rgrid, 256, 0, stream>>>(p.partials, splits, mn, N, bias, bias_dtype, act, D,
                                                    d_dtype);
    return launched();
  }
  return 0;
}

}  // namespace

extern "C" {

int tb200_gemm_bf16_nt(const void* A, const void* B, void* D, const void* bias, int bias_dtype,
                       int M, int N, int K, int d_dtype, int act, cudaStream_t stream) {
  return gemm_common(nullptr, 0, 0, 0, A, B, D, bias, bias_dtype, M, N, K, d_dtype, act, nullptr,
                     nullptr, 0, stream);
}

// D (bf16) = A . B^T plus, from the epilogue, per-channel (sum, sum of squares) partials of D for a following
// BatchNorm: *nparts rows of [N][2] floats in `stats` (0 rows when the problem had to be split along K).
int tb200_gemm_bf16_nt_stats(const void* A, const void* B, void* D, int M, int N, int K, float* stats,
                             size_t stats_floats, int* nparts, cudaStream_t stream) {
  return gemm_common(nullptr, 0, 0, 0, A, B, D, nullptr, 0, M, N, K, TB200_BF16, 0, nullptr, nullptr, 0, stream, nullptr,
                     0, 0, 0, nullptr, stats, stats_floats, nparts);
}

int tb200_gemm_bf16(const void* A, const void* B, void* D, const void* bias, int bias_dtype, int M, int N,
                    int K, int a_mn_major, int b_mn_major, int d_dtype, int act, cudaStream_t stream) {
  return gemm_common(nullptr, 0, 0, 0, A, B, D, bias, bias_dtype, M, N, K, d_dtype, act, nullptr, nullptr, 0,
                     stream, nullptr, 0, a_mn_major ? 1 : 0, b_mn_major ? 1 : 0);
}

// D = A . B^T + addend (all bf16, addend and D [M, N]; D may alias addend): the residual-gradient add of a
// bottleneck block folded into the dgrad GEMM of its first convolution.
int tb200_gemm_bf16_add(const void* A, const void* B, void* D, const void* addend, int M, int N, int K, int a_mn_major,
                        int b_mn_major, cudaStream_t stream) {
  if (!addend) return fail("tb200_gemm_bf16_add: null addend%s");
  return gemm_common(nullptr, 0, 0, 0, A, B, D, nullptr, 0, M, N, K, TB200_BF16, 0, nullptr, nullptr, 0, stream, nullptr, 0,
                     a_mn_major ? 1 : 0, b_mn_major ? 1 : 0, nullptr, nullptr, 0, nullptr, nullptr, addend);
}

int tb200_gemm_bf16_nt_pull(const tb200_world_t* world, int bucket, size_t b_off,
                            int rows_per_rank, const void* A, void* D, const void* bias,
                            int bias_dtype, int M, int N, int K, int d_dtype, int act,
                            const uint32_t* epoch_dev, int* status_dev,
                            unsigned long long spin_budget, uint32_t* sync, cudaStream_t stream) {
  if (!world) return fail("tb200_gemm_bf16_nt_pull: null world%s");
  return gemm_common(world, bucket, b_off, rows_per_rank, A, nullptr, D, bias, bias_dtype, M, N, K,
                     d_dtype, act, epoch_dev, status_dev, spin_budget, stream, nullptr, 0, 0, 0, nullptr, nullptr, 0,
                     nullptr, sync);
}

// ---- implicit-GEMM convolutions (K5: "convs as implicit GEMM on the same mainloop") ----------------
// x [n, h, w, cin] and y / dy [n, oh, ow, cout] are NHWC bf16 (torch channels_last), weights
// [cout, R, S, cin] bf16 (a channels_last torch weight). Square stride / padding, no dilation, no groups.
static int conv_out(int in, int k, int stride, int pad) { return (in + 2 * pad - k) / stride + 1; }

static int conv_check(const char* who, int n, int h, int w, int cin, int cout, int R, int S, int stride,
                      int pad, int im2col_channels) {
  if (n <= 0 || h <= 0 || w <= 0 || cin <= 0 || cout <= 0) return fail("%s: empty problem", who);
  if (R <= 0 || S <= 0 || R > 16 || S > 16 || stride <= 0 || stride > 8 || pad < 0 || pad >= 16)
    return fail("%s: unsupported filter geometry", who);
  if (conv_out(h, R, stride, pad) <= 0 || conv_out(w, S, stride, pad) <= 0) return fail("%s: empty output", who);
  if (im2col_channels % 64)
    return fail("%s: the gathered tensor's channel count (%ld) must be a multiple of 64", who, im2col_channels);
  return 0;
}

static int conv2d_fprop_impl(const void* x, const void* w, void* y, int n, int h, int wd, int cin, int cout, int R,
                             int S, int stride, int pad, float* stats, size_t stats_floats, int* nparts,
                             cudaStream_t stream);

int tb200_conv2d_fprop(const void* x, const void* w, void* y, int n, int h, int wd, int cin, int cout, int R,
                       int S, int stride, int pad, cudaStream_t stream) {
  return conv2d_fprop_impl(x, w, y, n, h, wd, cin, cout, R, S, stride, pad, nullptr, 0, nullptr, stream);
}

int tb200_conv2d_fprop_stats(const void* x, const void* w, void* y, int n, int h, int wd, int cin, int cout, int R,
                             int S, int stride, int pad, float* stats, size_t stats_floats, int* nparts,
                             cudaStream_t stream) {
  return conv2d_fprop_impl(x, w, y, n, h, wd, cin, cout, R, S, stride, pad, stats, stats_floats, nparts, stream);
}

static int conv2d_fprop_impl(const void* x, const void* w, void* y, int n, int h, int wd, int cin, int cout, int R,
                             int S, int stride, int pad, float* stats, size_t stats_floats, int* nparts,
                             cudaStream_t stream) {
  if (nparts) *nparts = 0;
  if (int rc = conv_check("tb200_conv2d_fprop", n, h, wd, cin, cout, R, S, stride, pad, cin)) return rc;
  const int oh = conv_out(h, R, stride, pad), ow = conv_out(wd, S, stride, pad);
  ConvHost c{};
  c.g = ConvGeom{1, oh, ow, stride, -pad, R, S, cin / 64, 0};
  c.act = x; c.n = n; c.h = h; c.w = wd; c.c = cin; c.pad = pad;
  return gemm_common(nullptr, 0, 0, 0, nullptr, w, y, nullptr, 0, n * oh * ow, cout, R * S * cin, TB200_BF16, 0,
                     nullptr, nullptr, 0, stream, nullptr, 0, 0, 0, &c, stats, stats_floats, nparts);
}

// Strided input gradient as stride x stride parity classes: output pixels with
// (h % stride, w % stride) == (ph, pw) only ever see the taps r = (ph + pad) mod stride, + stride, ..., so each class
// is a stride-1 correlation of dy with a dense R' x S' sub-filter whose rows land on every stride-th pixel of dx.
// One launch per class on the same im2col mainloop; the epilogue scatters its rows to the class's pixels.
static int conv2d_dgrad_strided(const void* dy, const void* w, void* dx, int n, int h, int wd, int cin, int cout, int R,
                                int S, int stride, int pad, cudaStream_t stream) {
  const int oh = conv_out(h, R, stride, pad), ow = conv_out(wd, S, stride, pad);
  struct Axis { int taps, tmax, off, count; };
  auto axis = [&](int ph, int extent, int k) {
    Axis a{};
    a.count = (extent - ph + stride - 1) / stride;             // pixels of this parity
    int tmax = -1;
    for (int t = k - 1; t >= 0; --t)
      if ((ph + pad - t) % stride == 0) { tmax = t; break; }
    a.tmax = tmax;
    a.taps = tmax < 0 ? 0 : tmax / stride + 1;
    a.off = tmax < 0 ? 0 : (ph + pad - tmax) / stride;         // first dy row/column of the window of pixel 0
    return a;
  };
  bool any_empty = false;
  for (int ph = 0; ph < stride; ++ph)
    for (int pw = 0; pw < stride; ++pw) {
      const Axis ah = axis(ph, h, R), aw = axis(pw, wd, S);
      if (ah.count > 0 && aw.count > 0 && (ah.taps == 0 || aw.taps == 0)) any_empty = true;
    }
  // classes no tap reaches (e.g. the odd pixels of a 1x1 / stride-2 convolution) are zero
  if (any_empty) TB_CUDA_OK(cudaMemsetAsync(dx, 0, (size_t)n * h * wd * cin * 2, stream));
  for (int ph = 0; ph < stride; ++ph)
    for (int pw = 0; pw < stride; ++pw) {
      const Axis ah = axis(ph, h, R), aw = axis(pw, wd, S);
      if (ah.count <= 0 || aw.count <= 0 || ah.taps == 0 || aw.taps == 0) continue;
      ConvHost c{};
      c.g = ConvGeom{1, ah.count, aw.count, 1, ah.off, ah.taps, aw.taps, cout / 64, 1};
      c.g.lower_w = aw.off;
      c.g.wt_rb = ah.tmax; c.g.wt_rs = stride; c.g.wt_sb = aw.tmax; c.g.wt_ss = stride; c.g.wt_S = S;
      c.g.sc_on = 1; c.g.sc_H = h; c.g.sc_W = wd; c.g.sc_stride = stride; c.g.sc_ph = ph; c.g.sc_pw = pw;
      c.act = dy; c.n = n; c.h = oh; c.w = ow; c.c = cout; c.pad = 0;
      c.wcout = cout; c.wtaps = R * S; c.wcin = cin;
      // window corners run over [off, off + count): count = extent + upper - lower
      c.has_corners = 1;
      c.corners[0] = aw.off; c.corners[1] = ah.off;
      c.corners[2] = aw.count - ow + aw.off; c.corners[3] = ah.count - oh + ah.off;
      if (int rc = gemm_common(nullptr, 0, 0, 0, nullptr, w, dx, nullptr, 0, n * ah.count * aw.count, cin,
                               ah.taps * aw.taps * cout, TB200_BF16, 0, nullptr, nullptr, 0, stream, nullptr, 0, 0, 1, &c))
        return rc;
    }
  return 0;
}

int tb200_conv2d_dgrad(const void* dy, const void* w, void* dx, int n, int h, int wd, int cin, int cout, int R,
                       int S, int stride, int pad, cudaStream_t stream) {
  if (pad > R - 1 || pad > S - 1) return fail("tb200_conv2d_dgrad: padding larger than the filter%s");
  if (int rc = conv_check("tb200_conv2d_dgrad", n, h, wd, cin, cout, R, S, stride, pad, cout)) return rc;
  if (cin % 8) return fail("tb200_conv2d_dgrad: cin %s(%ld) must be a multiple of 8", "", cin);
  if (stride != 1) return conv2d_dgrad_strided(dy, w, dx, n, h, wd, cin, cout, R, S, stride, pad, stream);
  if (R != S) return fail("tb200_conv2d_dgrad: square filters only%s");
  const int oh = conv_out(h, R, 1, pad), ow = conv_out(wd, S, 1, pad);
  // dx = "full" correlation of dy with the flipped filter: an fprop over dy with padding taps-1-pad whose
  // output grid is x's
  ConvHost c{};
  c.g = ConvGeom{1, h, wd, 1, -(R - 1 - pad), R, S, cout / 64, 1};
  c.act = dy; c.n = n; c.h = oh; c.w = ow; c.c = cout; c.pad = R - 1 - pad;
  c.wcout = cout; c.wtaps = R * S; c.wcin = cin;
  return gemm_common(nullptr, 0, 0, 0, nullptr, w, dx, nullptr, 0, n * h * wd, cin, R * S * cout, TB200_BF16, 0,
                     nullptr, nullptr, 0, stream, nullptr, 0, 0, 1, &c);
}

int tb200_conv2d_wgrad(const void* x, const void* dy, void* dw, int n, int h, int wd, int cin, int cout, int R,
                       int S, int stride, int pad, int dw_dtype, cudaStream_t stream) {
  if (int rc = conv_check("tb200_conv2d_wgrad", n, h, wd, cin, cout, R, S, stride, pad, cin)) return rc;
  if (cout % 8) return fail("tb200_conv2d_wgrad: cout %s(%ld) must be a multiple of 8", "", cout);
  const int oh = conv_out(h, R, stride, pad), ow = conv_out(wd, S, stride, pad);
  // dW[cout, (tap, cin)] = dy[pixel, cout]^T . im2col(x)[pixel, (tap, cin)]: both operands are stored
  // with the reduction (pixel) index outermost, i.e. MN-major
  ConvHost c{};
  c.g = ConvGeom{2, oh, ow, stride, -pad, R, S, cin / 64, 0};
  c.act = x; c.n = n; c.h = h; c.w = wd; c.c = cin; c.pad = pad;
  return gemm_common(nullptr, 0, 0, 0, dy, nullptr, dw, nullptr, 0, cout, R * S * cin, n * oh * ow, dw_dtype, 0,
                     nullptr, nullptr, 0, stream, nullptr, 0, 1, 1, &c);
}

// Hardware probe (see probe_shift_kernel): A bf16 [256, 64], B bf16 [64, 64], D fp32 [128, 64].
int tb200_probe_shifted_operand(const void* A, const void* B, float* D, int shift, int base_offset,
                                cudaStream_t stream) {
  if (shift < 0 || shift > 128) return fail("tb200_probe_shifted_operand: shift %s(%ld) outside 0..128", "", shift);
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) {
    TB_CUDA_OK(cudaFree(nullptr));
    ctx_bound = true;
  }
  CUtensorMap ma, mb;
  if (int rc = make_map(&ma, A, 256, 64, 128)) return rc;
  if (int rc = make_map(&mb, B, 64, 64, 64)) return rc;
  const int smem = 256 * 128 + 64 * 128 + 64;
  static DeviceOnce configured;
  if (int rc = ensure_dynamic_smem(configured, probe_shift_kernel, smem)) return rc;
  probe_shift_kernel<<<1, 128, smem, stream>>>(ma, mb, D, shift, base_offset);
  return launched();
}

// How many split-K partial planes tb200_gemm_bf16_nt / _noalloc would use for this problem on the current
// device (1 = no scratch needed). Scratch bytes = splits * M * N * 4.
// -1 (default): the 2-CTA kernel serves eligible problems of at least 256 x 256 x 256; 0: never; 1: every
// eligible problem (tests). Not part of the reference's surface.
void tb200_gemm_set_pair_mode(int mode) { g_pair_mode = mode; }

int tb200_gemm_plan_splits(int M, int N, int K) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  int block_n = 0, splits = 1;
  if (plan_tiles(M, N, K, false, 0, sm_count(), &block_n, &splits)) return 0;
  return splits;
}

// Pre-size the split-K scratch (call before CUDA-graph capture so no allocation happens
// inside the captured region).
int tb200_gemm_reserve(size_t bytes) {
  float* unused = nullptr;
  return get_workspace(bytes, nullptr, &unused);
}

int tb200_gemm_reserve_stream(size_t bytes, cudaStream_t stream) {
  float* unused = nullptr;
  return get_workspace(bytes, stream, &unused);
}

// Frees the split-K workspace kept for (current device, stream): a long-lived worker process creates new streams
// for every job it runs, and the workspaces of finished jobs would otherwise stay allocated. The caller guarantees
// that nothing is in flight on the stream.
int tb200_gemm_release_stream(cudaStream_t stream) {
  int dev = 0;
  TB_CUDA_OK(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lock(g_ws_mutex);
  auto it = g_ws.find(((unsigned long long)(uintptr_t)stream << 6) ^ (unsigned long long)dev);
  if (it == g_ws.end()) return 0;
  if (it->second.ptr) TB_CUDA_OK(cudaFree(it->second.ptr));
  g_ws.erase(it);
  return 0;
}

}  // extern "C"

// =================================================================================================
// No-allocation entry point (no counterpart in the reference; it has no native code at all).
//
// tb200_gemm_bf16_nt may call cudaMalloc/cudaFree to grow its split-K scratch. Some users of this
// library must never do that once their job is running:
//   * a job that shares its GPU with one or two other jobs (the simulator packs up to three per
//     device): cudaMalloc and cudaFree synchronise the whole device, so one tenant's allocation
//     stalls its neighbours and pollutes the interference that calib/ is trying to measure, and
//     on a GPU whose memory the packed jobs have fully subscribed the allocation can simply fail
//     in the middle of a step;
//   * plain C/C++ hosts that link the C ABI without any allocator of their own and want a call
//     whose memory footprint is fixed when the library is loaded.
// For them the split-K partial sums live in the array below: TB200_GEMM_NOALLOC_SCRATCH_BYTES of
// device memory that every GPU which loads this library holds until the process exits.
// =================================================================================================
static __device__ __align__(256) float tb200_noalloc_splitk_scratch[TB200_GEMM_NOALLOC_SCRATCH_BYTES / sizeof(float)];

namespace {

std::mutex g_noalloc_mutex;
float* g_noalloc_ptr[kMaxDevices] = {};

// In host code the array's name is only a stand-in for the symbol; the kernels need the device
// address, which differs per GPU. Resolved once per device, with that device current.
int noalloc_scratch(float** out) {
  int dev = 0;
  TB_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return fail("tb200_gemm_bf16_nt_noalloc: device %s%ld unsupported", "", dev);
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (!g_noalloc_ptr[dev]) {
    float* p = nullptr;
    TB_CUDA_OK(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), tb200_noalloc_splitk_scratch));
    if (!p) return fail("tb200_gemm_bf16_nt_noalloc: scratch symbol has no device address%s");
    g_noalloc_ptr[dev] = p;
  }
  *out = g_noalloc_ptr[dev];
  return 0;
}

}  // namespace

extern "C" int tb200_gemm_bf16_nt_noalloc(const void* A, const void* B, void* D, const void* bias,
                                          int bias_dtype, int M, int N, int K, int d_dtype, int act,
                                          cudaStream_t stream) {
  // Same planning, kernels and results as tb200_gemm_bf16_nt; only the home of the split-K partial
  // sums differs (the built-in array instead of a cudaMalloc'ed workspace).
  float* scratch = nullptr;
  if (int rc = noalloc_scratch(&scratch)) return rc;
  return gemm_common(nullptr, 0, 0, 0, A, B, D, bias, bias_dtype, M, N, K, d_dtype, act, nullptr, nullptr,
                     0, stream, scratch, (size_t)TB200_GEMM_NOALLOC_SCRATCH_BYTES);
}
