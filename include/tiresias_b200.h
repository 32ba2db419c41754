/*
 * tiresias_b200 native library — C ABI (libtiresias_b200.so, sm_100a only).
 *
 * Every function returns 0 on success; on failure it returns non-zero and
 * tb200_last_error() describes why. All kernels are launched on the stream that is
 * passed in; nothing synchronises the device unless stated.
 *
 * The reference simulator (matthewygf/Tiresias) has no native code.
 * Each entry point below names the piece of the reference's *model* that it turns
 * into an executed sm_100a kernel.
 */
#ifndef TIRESIAS_B200_H_
#define TIRESIAS_B200_H_

#include <stddef.h>
#include <stdint.h>
#include <cuda_runtime_api.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TB200_MAX_WORLD 8          /* one NVSwitch domain: 8 x B200                     */
#define TB200_MAX_BUCKETS 256      /* gradient buckets per job                          */
#define TB200_MAX_RING_CTAS 64     /* independent ring lanes (one per CTA)              */
#define TB200_IPC_HANDLE_BYTES 64

/* dtype codes */
#define TB200_F32 0
#define TB200_BF16 1

/* Layout of one rank's flag pad (uint32 words). Every rank owns one pad in symmetric
 * memory; peers write into it with st.release.sys, the owner polls it locally.          */
#define TB200_FLAG_GRAD_READY(bucket, src) (((bucket) * TB200_MAX_WORLD) + (src))
#define TB200_FLAG_PARAM_DONE(bucket, src) \
  ((TB200_MAX_BUCKETS * TB200_MAX_WORLD) + ((bucket) * TB200_MAX_WORLD) + (src))
#define TB200_FLAG_RING(dir, lane) \
  ((2 * TB200_MAX_BUCKETS * TB200_MAX_WORLD) + ((dir) * TB200_MAX_RING_CTAS) + (lane))
#define TB200_FLAG_WORDS (2 * TB200_MAX_BUCKETS * TB200_MAX_WORLD + 2 * TB200_MAX_RING_CTAS)

/* ---- runtime ------------------------------------------------------------------------ */
int tb200_version(void);
const char* tb200_last_error(void);
/* Number of kernels this library has launched since the last reset (bench "gpu_launches"). */
unsigned long long tb200_launch_count(void);
void tb200_launch_count_reset(void);

/* ---- symmetric memory (replaces the reference's rack/switch network model,
 *      infra/rack.py:10, run_sim.py:59-69) ------------------------------------------- */
int tb200_symm_alloc(size_t bytes, void** out);               /* cudaMalloc + zero      */
int tb200_symm_free(void* ptr);
int tb200_ipc_export(void* ptr, unsigned char handle[TB200_IPC_HANDLE_BYTES]);
int tb200_ipc_open(const unsigned char handle[TB200_IPC_HANDLE_BYTES], void** out);
int tb200_ipc_close(void* ptr);
int tb200_enable_peer_access(int device, int peer);
int tb200_flags_fill(uint32_t* flags, size_t first, size_t count, uint32_t value,
                     cudaStream_t stream);

/* One job's view of the symmetric world. Index [rank] holds the local pointers. */
typedef struct {
  int world;                               /* number of workers == number of PS shards  */
  int rank;
  const void* grad[TB200_MAX_WORLD];       /* flat gradient buffers (grad_dtype)        */
  void* param[TB200_MAX_WORLD];            /* flat parameter buffers (bf16)             */
  uint32_t* flags[TB200_MAX_WORLD];        /* flag pads, TB200_FLAG_WORDS each          */
  const void* grad_mc;                     /* NVLS multicast view of the W gradient buffers, or NULL */
  void* param_mc;                          /* NVLS multicast view of the W parameter buffers, or NULL */
} tb200_world_t;

/* Hyper-parameters live in device memory so a captured CUDA graph picks up changes. */
typedef struct {
  float lr;
  float momentum;
  float weight_decay;
  float grad_scale;                        /* usually 1/world                           */
} tb200_sgd_hyper_t;

/* ---- K4: optimizer step (README.md:7 "Synchronize SGD", core/jobs/job.py:72) --------- */
/* master/momentum fp32, grad in grad_dtype, writes bf16 (or fp32) parameters.           */
int tb200_sgd_momentum(float* master, float* momentum, const void* grad, int grad_dtype,
                       void* param_out, int param_dtype, size_t n,
                       const tb200_sgd_hyper_t* hyper_dev, cudaStream_t stream);

/* ---- K1 + K2 (store side): worker->PS push fused with the SGD update and the PS->worker
 *      broadcast (core/network/network_service.py:30-37 "round trip"; infra/switch.py:97-107
 *      per-node traffic; one PS shard per GPU as in infra/switch.py:100-106).
 *
 * For bucket `bucket`, this rank is the parameter server of elements
 * [shard_off, shard_off + shard_len) of the flat buffers. The kernel
 *   1. posts "my gradients for this bucket are final" to every peer,
 *   2. waits until every worker has posted the same,
 *   3. reads the W gradient slices over NVLink (fixed rank order => bitwise reproducible),
 *      scales, applies SGD+momentum+weight decay to the fp32 master shard,
 *   4. stores the new bf16 parameters into every worker's parameter buffer (broadcast=1),
 *      or only into its own (broadcast=0: the workers pull the shard later, inside
 *      tb200_gemm_bf16_nt_pull),
 *   5. posts "shard done" to every peer.
 * *epoch_dev is the step number (device memory; see tb200_step_end).
 * counter_dev: one zero-initialised uint32 per bucket (scratch owned by the caller).
 * status_dev: set non-zero when a wait ran out of `spin_budget` polls (0 = wait forever).  */
int tb200_ps_push_pull(const tb200_world_t* world, int bucket, size_t shard_off,
                       size_t shard_len, int grad_dtype, float* master, float* momentum,
                       const tb200_sgd_hyper_t* hyper_dev, const uint32_t* epoch_dev,
                       uint32_t* counter_dev, int* status_dev,
                       unsigned long long spin_budget, int max_ctas, int broadcast,
                       cudaStream_t stream);

/* Waits until all PS shards of buckets [0, num_buckets) are done for *epoch_dev, then
 * increments *epoch_dev. Launch once per step after the last push.                       */
int tb200_step_end(const tb200_world_t* world, int num_buckets, uint32_t* epoch_dev,
                   int* status_dev, unsigned long long spin_budget, cudaStream_t stream);

/* ---- K3: all-reduce mode (README.md:6 "All-reduce not yet implemented") ---------------- */
/* In-place sum over `n` elements of the symmetric buffers buf[0..W-1] (same offset on every
 * rank). algo 0 = ring (reduce-scatter + all-gather, 2(W-1) flag-synchronised steps),
 * algo 1 = direct two-shot through NVSwitch, algo 2 = one-shot NVLS (multimem.ld_reduce + multimem.st;
 * bf16, bufs must lie inside world->grad and world->grad_mc be set). seq_dev: device counter bumped
 * per call.                                                                                   */
int tb200_allreduce(const tb200_world_t* world, void* const bufs[TB200_MAX_WORLD], size_t n,
                    int dtype, float scale, int algo, uint32_t* seq_dev, int* status_dev,
                    unsigned long long spin_budget, int num_ctas, cudaStream_t stream);

/* ---- K5: tcgen05 GEMM (model_factory.py:4-55 / core/models.py:9-20 FC layers) ---------- */
/* D[M,N] = act(A[M,K] . B[N,K]^T + bias[N]); A, B bf16 row-major (K contiguous), fp32
 * accumulation in TMEM, D bf16 or fp32; bias fp32 or bf16 (bias_dtype), may be NULL.
 * act: 0 none, 1 relu, 2 gelu(tanh).
 * K must be a multiple of 8 (TMA 16-byte stride rule).                                     */
int tb200_gemm_bf16_nt(const void* A, const void* B, void* D, const void* bias, int bias_dtype,
                       int M, int N, int K, int d_dtype, int act, cudaStream_t stream);

/* Same kernel with selectable operand layouts: D[M,N] = act(op(A) . op(B)^T + bias) where A is stored
 * [M,K] (a_mn_major = 0, K contiguous) or [K,M] (a_mn_major = 1, M contiguous), and B is stored [N,K]
 * (b_mn_major = 0) or [K,N] (b_mn_major = 1). The linear layer's backward needs no transposes:
 *   dx[M,K] = dy[M,N] . W[N,K]      -> A = dy (K-major over N),  B = W with b_mn_major = 1
 *   dW[N,K] = dy[M,N]^T . x[M,K]    -> A = dy with a_mn_major = 1, B = x with b_mn_major = 1
 * The contiguous dimension of every operand must be a multiple of 8.                               */
int tb200_gemm_bf16(const void* A, const void* B, void* D, const void* bias, int bias_dtype, int M, int N,
                    int K, int a_mn_major, int b_mn_major, int d_dtype, int act, cudaStream_t stream);

/* ---- NVLS (NVLink SHARP) multicast: in-switch reduce / broadcast through multimem.* instructions ----
 * tb200_nvls_supported: 1 when `device` can join a multicast object, 0 when not, <0 on error.
 * tb200_nvls_selftest: single-process check and timing of multimem.ld_reduce (sum of all GPUs' copies),
 * multimem.st (store into all copies) and the one-shot all-reduce built from them, over GPUs 0..ndev-1
 * with `bytes` of bf16 per GPU and `ctas` CTAs of 512 threads; `report` receives one JSON object.
 * Returns 0 = all values correct, 2 = NVLS unavailable on this box, 1 = error.                      */
int tb200_nvls_supported(int device);
int tb200_nvls_selftest(int ndev, size_t bytes, int ctas, char* report, size_t report_len);

/* Multi-process symmetric memory that can carry an NVLS view: physical allocations (cuMemCreate) shared
 * between the ranks' processes as POSIX file descriptors (the caller passes them over a unix socket).
 * Order, with the rank's device current: every rank tb200_vmm_round + tb200_vmm_alloc; rank 0
 * tb200_mc_create, the others tb200_mc_import(fd); every rank tb200_mc_add_device; BARRIER; every rank
 * tb200_mc_bind -> multicast pointer; every rank tb200_vmm_import of each peer's fd -> peer pointers.   */
int tb200_vmm_round(int world, size_t bytes, size_t* rounded);
int tb200_vmm_alloc(size_t rounded, unsigned long long* phys, void** ptr, int* fd);
int tb200_vmm_import(int fd, size_t rounded, unsigned long long* phys, void** ptr);
int tb200_vmm_release(unsigned long long phys, void* ptr, size_t rounded);
int tb200_mc_create(int world, size_t rounded, unsigned long long* mc, int* fd);
int tb200_mc_import(int fd, unsigned long long* mc);
int tb200_mc_add_device(unsigned long long mc);
int tb200_mc_bind(unsigned long long mc, unsigned long long phys, size_t rounded, void** mc_ptr);
int tb200_mc_release(unsigned long long mc, void* mc_ptr, size_t rounded);

/* Implicit-GEMM convolutions on the same tcgen05 mainloop (K5): one operand is gathered by im2col-mode
 * TMA loads instead of tiled ones. x [n,h,w,cin], y/dy [n,oh,ow,cout] NHWC bf16 (torch channels_last),
 * w / dw [cout,R,S,cin] (a channels_last weight); square stride and padding, no dilation, no groups.
 *   fprop: y  = conv(x, w)                     cin  % 64 == 0
 *   dgrad: dx = conv_transpose(dy, w)          cout % 64 == 0, cin % 8 == 0, stride 1 only
 *   wgrad: dw = dy^T . im2col(x)               cin  % 64 == 0, cout % 8 == 0; dw bf16 or fp32 (split-K)  */
int tb200_conv2d_fprop(const void* x, const void* w, void* y, int n, int h, int wd, int cin, int cout, int R,
                       int S, int stride, int pad, cudaStream_t stream);
int tb200_conv2d_dgrad(const void* dy, const void* w, void* dx, int n, int h, int wd, int cin, int cout, int R,
                       int S, int stride, int pad, cudaStream_t stream);
int tb200_conv2d_wgrad(const void* x, const void* dy, void* dw, int n, int h, int wd, int cin, int cout, int R,
                       int S, int stride, int pad, int dw_dtype, cudaStream_t stream);

/* EXPERIMENTAL (compiles, not yet run on hardware): 3x3 / stride 1 / pad 1 forward convolution that fetches
 * each input halo tile once and forms the nine taps as shifted shared-memory windows (csrc/conv_halo.cu).
 * Same tensors as tb200_conv2d_fprop; cin % 64 == 0, cout % 8 == 0, image width <= 126.                   */
int tb200_conv2d_fprop_halo(const void* x, const void* w, void* y, int n, int h, int wd, int cin, int cout,
                            int R, int S, cudaStream_t stream);

/* ---- 3x3 / stride 1 / pad 1 convolutions as halo-tile implicit GEMMs (csrc/conv3x3.cu; model zoo of
 * model/model_factory.py:31-55). x [n,h,w,cin], y / dy [n,h,w,cout] NHWC bf16, w [cout,3,3,cin] bf16.
 * cin % 64 == 0 and cout % 8 == 0 for fprop; cout % 64 == 0 and cin % 8 == 0 for dgrad; w <= 255.
 * The input rows a 128-position output tile needs are fetched once per 64-channel block and the nine taps
 * are shifted shared-memory windows of that tile. dgrad reads the forward weight as stored.
 * `stats` (optional, fprop): the epilogue also writes per-channel (sum, sum of squares) of the bf16 outputs
 * as *nparts partial rows [nparts][cout][2] (fp32) for tb200_bn_finalize - the statistics pass of a following
 * BatchNorm comes for free. The buffer must hold tb200_conv_stats_floats(cout) floats.                     */
size_t tb200_conv_stats_floats(int cout);
/* The same statistics from the GEMM / im2col-convolution epilogue (1x1 and strided convolutions): D bf16 [M,N] =
 * A . B^T. *nparts = 0 when the problem had to be split along K (tiny batch): the caller then runs its own pass. */
int tb200_gemm_bf16_nt_stats(const void* A, const void* B, void* D, int M, int N, int K, float* stats,
                             size_t stats_floats, int* nparts, cudaStream_t stream);
int tb200_conv2d_fprop_stats(const void* x, const void* w, void* y, int n, int h, int wd, int cin, int cout, int R,
                             int S, int stride, int pad, float* stats, size_t stats_floats, int* nparts,
                             cudaStream_t stream);
int tb200_conv3x3_fprop(const void* x, const void* w, void* y, int n, int h, int wd, int cin, int cout, float* stats,
                        size_t stats_floats, int* nparts, cudaStream_t stream);
/* tb200_conv3x3_fprop with a per-output-channel bias (TB200_F32 / TB200_BF16; may be NULL) and an optional ReLU in
 * the epilogue (VGG / AlexNet style layers), and a geometry query that runs no GPU work.                      */
int tb200_conv3x3_fprop_bias(const void* x, const void* w, const void* bias, int bias_dtype, int relu, void* y, int n,
                             int h, int wd, int cin, int cout, cudaStream_t stream);
int tb200_conv3x3_supported(int n, int h, int wd, int c_red, int k_out, int dgrad);
int tb200_conv3x3_dgrad(const void* dy, const void* w, void* dx, int n, int h, int wd, int cin, int cout,
                        cudaStream_t stream);
/* Weight gradient of the same convolution on a halo-tile kernel: dw [cout][3][3][cin] bf16 from x [n,h,w,cin] and
 * dy [n,h,w,cout]; cin a multiple of 64, cout 64 or a multiple of 128, w <= 255. `workspace`: fp32 scratch of
 * tb200_conv3x3_wgrad_workspace(...) bytes (0 = shape not served). No reference counterpart.                */
size_t tb200_conv3x3_wgrad_workspace(int n, int h, int wd, int cin, int cout);
int tb200_conv3x3_wgrad(const void* x, const void* dy, void* dw, int n, int h, int wd, int cin, int cout,
                        float* workspace, size_t workspace_bytes, cudaStream_t stream);
/* Per-channel statistics partials [nparts][C][2] over `count` values per channel -> batch mean / invstd,
 * scale_shift = [gamma*invstd | beta - mean*gamma*invstd] for tb200_bn_apply, running statistics (may be NULL).
 * Compensated sums in a fixed order: bitwise reproducible.                                                */
int tb200_bn_finalize(const float* partials, int nparts, double count, int C, const void* gamma, const void* beta,
                      int param_dtype, float* running_mean, float* running_var, float momentum, float eps,
                      float* save_mean, float* save_invstd, float* scale_shift, cudaStream_t stream);

/* Hardware probe for halo-tile convolutions: D[128,64] fp32 = A[shift:shift+128, :] . B^T with A bf16 [256,64],
 * B bf16 [64,64], the A shared-memory descriptor advanced by `shift` 128-byte rows and its matrix-base-offset
 * field set to `base_offset` (bench/probe_shift.py reports which setting reproduces the fp32 product). */
int tb200_probe_shifted_operand(const void* A, const void* B, float* D, int shift, int base_offset,
                                cudaStream_t stream);

/* EXPERIMENTAL probe for a halo-tile wgrad: D[128,64] fp32 = A[128,64] . B[shift:shift+64, :] with B bf16
 * [256 reduction rows, 64] read as an MN-major operand whose descriptor starts `shift` rows into the tile. */
int tb200_probe_shifted_mn_operand(const void* A, const void* B, float* D, int shift, cudaStream_t stream);

/* Pre-size the split-K scratch of the GEMM (call before CUDA-graph capture). The scratch is kept per
 * (device, stream) so that jobs sharing a GPU on different streams never share partial sums; this call sizes
 * the one of the current device's NULL stream, tb200_gemm_reserve_stream the one of `stream`.               */
int tb200_gemm_reserve(size_t bytes);
int tb200_gemm_reserve_stream(size_t bytes, cudaStream_t stream);
/* Frees the workspace kept for (current device, stream); nothing may be in flight on the stream.            */
int tb200_gemm_release_stream(cudaStream_t stream);

/* ---- tb200_gemm_bf16_nt without any allocation (no counterpart in the reference, which has
 *      no native code). Same arguments, same kernels and same results as tb200_gemm_bf16_nt.
 *
 * Who needs it: a job that shares its GPU with other jobs (cudaMalloc/cudaFree synchronise the
 * whole device, so an allocation by one tenant stalls the others, and on a fully subscribed GPU
 * it can fail mid-step), and plain C hosts of this ABI that bring no allocator and want the
 * memory footprint fixed at load time. This entry point never calls cudaMalloc or any
 * allocating routine and frees nothing (the first call a thread makes into the GEMM issues one
 * cudaFree(NULL), which releases nothing: it only binds the CUDA context to that thread): when
 * the problem is split along K, the fp32 partial sums go to a scratch array that is built into
 * the library, TB200_GEMM_NOALLOC_SCRATCH_BYTES per GPU, held by every GPU that loads the
 * library until the process exits.
 *
 *  - Limit: splits * M * N * 4 bytes must fit in the scratch. splits is chosen as in
 *    tb200_gemm_bf16_nt: 1 when the 128 x {128,256} tile grid already fills the GPU, otherwise
 *    up to min(ceil(K/64)/4, SM count, max(16, 0.1 * (M+N)*K*2 / (M*N*4))) - so a long-K,
 *    small-output problem can use one split per SM (148 on a B200), not 16.
 *    tb200_gemm_plan_splits(M, N, K) returns the count for the current device, so a caller can
 *    check the fit in advance. Larger problems are refused with an error that names both sizes;
 *    nothing is launched in that case. Problems that need no split use no scratch.
 *  - The first call on each device looks up the scratch's device address (one
 *    cudaGetSymbolAddress with that device current). Make that first call during set-up,
 *    before any stream capture is started on the device.
 *  - All calls on one GPU share the one scratch array. A call overwrites whatever the previous
 *    call left there and needs nothing cleared beforehand, but two calls that can run at the
 *    same time on the GPU (different streams, no dependency between them) race on it: order
 *    them, or use tb200_gemm_bf16_nt for one of them.                                         */
#define TB200_GEMM_NOALLOC_SCRATCH_BYTES (16u << 20)
int tb200_gemm_bf16_nt_noalloc(const void* A, const void* B, void* D, const void* bias,
                               int bias_dtype, int M, int N, int K, int d_dtype, int act,
                               cudaStream_t stream);

/* Split-K partial planes the one-CTA kernel (always the one behind _noalloc; tb200_gemm_bf16_nt may pick the 2-CTA
 * kernel and its own plan for large problems, whose workspace is internal) uses for an [M,K] x [N,K]^T problem on the
 * current device (1 = no scratch; 0 = empty problem). Scratch bytes = splits * M * N * 4.
 * No counterpart in the reference.                                                           */
int tb200_gemm_plan_splits(int M, int N, int K);

/* D = A . B^T + addend with the operand layouts of tb200_gemm_bf16 (bf16 in and out, N a multiple of 8; D may
 * alias addend). Used for the first convolution's dgrad of a residual block: the skip branch's gradient is added
 * in the epilogue instead of by a separate element-wise kernel. No reference counterpart.                   */
int tb200_gemm_bf16_add(const void* A, const void* B, void* D, const void* addend, int M, int N, int K, int a_mn_major,
                        int b_mn_major, cudaStream_t stream);

/* Plain GEMMs (not the convolutions, the pull variant or the allocation-free entry point) of at least
 * 256 x 256 x 256 run on the 2-CTA kernel (cta_group::2, 256 x 256 pair tiles; csrc/gemm_2cta.cuh).
 * mode: -1 that default, 0 never, 1 every eligible problem. Also TB200_GEMM_2CTA. No reference counterpart. */
void tb200_gemm_set_pair_mode(int mode);

/* ---- K2 (load side): PS->worker pull fused into the first GEMM that consumes the weights.
 * B is sharded by rows across the world: rank s owns rows [s*rows_per_rank, ...) of a [rows_per_rank * world, K]
 * matrix that every rank holds at element offset b_off of its parameter buffer (rows_per_rank a multiple of 128;
 * rows past N are padding). One warp per CTA is a copy engine that bulk-copies the other ranks' rows out of
 * their memory into this rank's copy, column tile by column tile, once the owner's PARAM_DONE flag for
 * `bucket` has reached *epoch_dev - 1 (the step that just ended), and publishes a flag per tile; the GEMM's
 * producers wait for a tile's flag and then TMA-load it locally. Every remote byte crosses NVLink once, the
 * transfer of later tiles overlaps the MMAs of earlier ones, and after the call the gathered matrix is
 * resident locally for the backward pass. `sync`: TB200_PULL_SYNC_WORDS zero-initialised 32-bit words owned by
 * this layer (arrival counters + per-tile epochs; the kernels keep them consistent from launch to launch).  */
#define TB200_PULL_SYNC_WORDS 4096
int tb200_gemm_bf16_nt_pull(const tb200_world_t* world, int bucket, size_t b_off,
                            int rows_per_rank, const void* A, void* D, const void* bias,
                            int bias_dtype, int M, int N, int K, int d_dtype, int act,
                            const uint32_t* epoch_dev, int* status_dev,
                            unsigned long long spin_budget, uint32_t* sync, cudaStream_t stream);

/* ---- small fused helpers ----------------------------------------------------------------- */
/* uint8 NHWC image batch -> normalised bf16 NHWC (mean/std per channel, C == 3).            */
int tb200_u8_to_bf16_norm(const uint8_t* src, void* dst, size_t pixels, const float mean[3],
                          const float inv_std[3], cudaStream_t stream);
/* Same, writing 8 channels per pixel (r, g, b, 0, 0, 0, 0, 0) so the first convolution can use the
 * 8-channel-aligned tensor-core path; dst holds pixels * 8 bf16.                                   */
int tb200_u8_to_bf16_norm_pad8(const uint8_t* src, void* dst, size_t pixels, const float mean[3],
                               const float inv_std[3], cudaStream_t stream);
/* ResNet stem on the tensor cores (csrc/stem.cu): the 7x7 / stride 2 / pad 3 convolution over 3 channels becomes a
 * 4x1 / stride-1 convolution over a packed 64-channel tensor. pack: uint8 NHWC [n,h,w,3] -> normalised bf16
 * Y [n, oh+3, ow, 64] (oh = (h-1)/2+1, ow = (w-1)/2+1); weight_pack: bf16 [cout][7][7][3] -> [cout][4][64];
 * wgrad_unpack: fp32 [cout][4][64] -> [cout][7][7][3] in dw_dtype. Then
 * tb200_conv2d_fprop(Y, Wp, y, n, oh+3, ow, 64, cout, 4, 1, 1, 0) and tb200_conv2d_wgrad with the same geometry. */
int tb200_stem7x7_pack(const uint8_t* src, void* dst, int n, int h, int w, const float mean[3], const float inv_std[3],
                       cudaStream_t stream);
int tb200_stem7x7_weight_pack(const void* w, void* wp, int cout, cudaStream_t stream);
int tb200_stem7x7_wgrad_unpack(const float* dwp, void* dw, int cout, int dw_dtype, cudaStream_t stream);
/* Copy up to any number of bf16 tensors into their destinations with one launch per TB200_PACK_MAX
 * items (gathers per-parameter gradients into the flat symmetric gradient buffer of a bucket).      */
#define TB200_PACK_MAX 64
typedef struct {
  const void* src;
  void* dst;
  unsigned long long n;                    /* elements (bf16) */
} tb200_pack_item_t;
int tb200_pack_bf16(const tb200_pack_item_t* items, int count, cudaStream_t stream);
/* bf16 [R,C] -> bf16 [C,R] tiled transpose (feeds the NT GEMM in backward).                 */
int tb200_transpose_bf16(const void* src, void* dst, int rows, int cols, cudaStream_t stream);
/* out[0] = sum(x^2) over n elements; `partials` is caller-provided scratch of
 * tb200_sumsq_scratch_floats() floats. Deterministic two-pass reduction.                    */
size_t tb200_sumsq_scratch_floats(void);
int tb200_sumsq(const void* x, int dtype, size_t n, float* partials, float* out,
                cudaStream_t stream);

/* ---- fused batch normalisation, NHWC bf16 (the ResNet jobs of model/model_factory.py:47-55).
 * x, residual, y, dy, dx, dres: bf16 [rows, C] with C contiguous (rows = N*H*W), C % 8 == 0,
 * C <= 4096. gamma/beta/dgamma/dbeta in param_dtype (TB200_F32 or TB200_BF16); statistics fp32.
 * `scratch`: TB200_BN_SCRATCH_FLOATS floats, zero-filled ONCE by the caller and then left alone
 * (the kernels keep self-resetting counters in it); one scratch serves every layer as long as
 * the calls are ordered on one stream. Reductions run in a fixed order: results are bitwise
 * reproducible.
 *   fwd_train: batch statistics (biased variance for normalisation, unbiased for running_var,
 *              momentum as in torch), y = relu?(bn(x) + residual?); saves mean / invstd for
 *              the backward. running_mean / running_var may be NULL.
 *   apply:     y = relu?(x*scale[c] + shift[c] + residual?) with scale_shift = [scale | shift]
 *              (inference, or any per-channel affine).
 *   bwd:       g = dy * (y > 0 if relu); dx, dres = g (may be NULL), dgamma, dbeta. For a ReLU layer
 *              WITHOUT residual input, y may be NULL: the mask is then recomputed from x with the
 *              forward's own fp32 fma (needs beta), which saves one activation read per pass.      */
#define TB200_BN_SCRATCH_FLOATS (1024 + 2 * 64 * 1024 + 4 * 4096 + 2 * 1024 * 1024)
size_t tb200_bn_scratch_floats(void);
int tb200_bn_fwd_train(const void* x, const void* residual, void* y, const void* gamma,
                       const void* beta, int param_dtype, float* running_mean, float* running_var,
                       float momentum, float eps, float* save_mean, float* save_invstd,
                       float* scratch, size_t rows, int C, int relu, cudaStream_t stream);
/* fwd_train whose statistics come from the producing convolution's epilogue (tb200_conv3x3_fprop `stats`):
 * no statistics pass over x. conv_partials [conv_nparts][C][2] as written by the convolution.            */
int tb200_bn_fwd_train_stats(const void* x, const void* residual, void* y, const void* gamma, const void* beta,
                             int param_dtype, float* running_mean, float* running_var, float momentum, float eps,
                             float* save_mean, float* save_invstd, float* scratch, const float* conv_partials,
                             int conv_nparts, size_t rows, int C, int relu, cudaStream_t stream);
/* The same two with every option: statistics from a convolution's epilogue (conv_partials, may be NULL) and the
 * ReLU mask as bits. A residual + ReLU layer's backward needs the sign of every output; `relu_mask` (rows*C/8
 * bytes + 16 of slack) carries it as one bit per element, so the backward reads 1/16 of a tensor instead of y in
 * both of its passes. Only where tb200_bn_mask_supported(rows, C) (the TMA-streamed kernels).              */
int tb200_bn_mask_supported(size_t rows, int C);
int tb200_bn_fwd_train_ex(const void* x, const void* residual, void* y, const void* gamma, const void* beta,
                          int param_dtype, float* running_mean, float* running_var, float momentum, float eps,
                          float* save_mean, float* save_invstd, float* scratch, const float* conv_partials,
                          int conv_nparts, uint8_t* relu_mask, size_t rows, int C, int relu, cudaStream_t stream);
int tb200_bn_bwd_ex(const void* dy, const void* x, const void* y, const uint8_t* relu_mask, const void* gamma,
                    const void* beta, int param_dtype, const float* save_mean, const float* save_invstd, void* dx,
                    void* dres, void* dgamma, void* dbeta, float* scratch, size_t rows, int C, int relu,
                    cudaStream_t stream);
int tb200_bn_apply(const void* x, const void* residual, void* y, const float* scale_shift,
                   size_t rows, int C, int relu, cudaStream_t stream);
int tb200_bn_bwd(const void* dy, const void* x, const void* y, const void* gamma, const void* beta,
                 int param_dtype, const float* save_mean, const float* save_invstd, void* dx, void* dres,
                 void* dgamma, void* dbeta, float* scratch, size_t rows, int C, int relu,
                 cudaStream_t stream);

/* ---- 3x3 / stride 2 / pad 1 max pooling, NHWC bf16 (ResNet stem). C % 8 == 0. The forward also writes
 * `argmax` (one byte per output element: kh*3+kw of the window's first maximum, torch's tie rule; may be
 * NULL for inference); the backward gathers with it — deterministic, no atomics. N, H, W, C describe the
 * INPUT of the pooling in both calls.                                                                 */
int tb200_maxpool3x3s2_fwd(const void* x, void* out, void* argmax, int N, int H, int W, int C,
                           cudaStream_t stream);
int tb200_maxpool3x3s2_bwd(const void* argmax, const void* dout, void* dx, int N, int H, int W, int C,
                           cudaStream_t stream);
/* The same pair for the other pooling geometries of the model zoo: (k, stride, pad) = (3,2,1), (2,2,0) (VGG),
 * (3,2,0) (AlexNet / GoogLeNet / Inception); floor mode. argmax: one byte per output element.              */
int tb200_maxpool_supported(int k, int stride, int pad);
int tb200_maxpool_fwd(const void* x, void* out, void* argmax, int N, int H, int W, int C, int k, int stride, int pad,
                      cudaStream_t stream);
int tb200_maxpool_bwd(const void* argmax, const void* dout, void* dx, int N, int H, int W, int C, int k, int stride,
                      int pad, cudaStream_t stream);

/* ---- Multi-head attention, head dimension 64 (csrc/attention.cu): flash-style tcgen05 kernels reading Q, K, V in
 * place from the fused projection output qkv [B*S, 3*H*64] (columns Q | K | V, each H heads of 64) and writing
 * out [B*S, H*64]; softmax scale 1/8, optional causal mask. lse2 [B, H, S] fp32 is saved by the forward; the
 * backward (deterministic: one kernel per key block for dK / dV, one per query block for dQ) writes dqkv in the
 * layout of qkv; delta is a [B, H, S] fp32 scratch. No reference counterpart.                                */
int tb200_attention_fwd(const void* qkv, void* out, float* lse2, int B, int S, int H, int causal, cudaStream_t stream);
int tb200_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse2, float* delta, void* dqkv,
                        int B, int S, int H, int causal, cudaStream_t stream);

/* ---- LayerNorm over the last dimension of bf16 rows [rows, D], D in {256, 512, 768, 1024} (csrc/norm_kernels.cu):
 * one warp per row, the row in registers; mean / rstd (fp32, [rows]) are saved by the forward for the backward;
 * dgamma / dbeta are reduced in a fixed order (scratch: tb200_layernorm_scratch_floats(D) floats).
 * gamma / beta / dgamma / dbeta in `param_dtype` (TB200_F32 or TB200_BF16). No reference counterpart.     */
int tb200_layernorm_supported(int D);
size_t tb200_layernorm_scratch_floats(int D);
int tb200_layernorm_fwd(const void* x, void* y, const void* gamma, const void* beta, int param_dtype, float* mean,
                        float* rstd, size_t rows, int D, float eps, cudaStream_t stream);
int tb200_layernorm_bwd(const void* dy, const void* x, const void* gamma, int param_dtype, const float* mean,
                        const float* rstd, void* dx, void* dgamma, void* dbeta, float* scratch, size_t scratch_floats,
                        size_t rows, int D, cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* TIRESIAS_B200_H_ */
