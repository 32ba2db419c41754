// Parameter-server / worker communication kernels for one NVSwitch domain (sm_100a).
//
// What the reference models and what runs here instead:
//   * worker -> PS gradient push  (network_service.py:34-37 first half of "* 2.0";
//     infra/switch.py:98,104)                       -> ps_push_pull_kernel steps 1-3
//   * synchronous SGD on the PS shard (README.md:7)  -> fused into the same kernel (step 3)
//   * PS -> worker parameter pull (second half of "* 2.0"; infra/switch.py:102)
//                                                    -> store-side broadcast, step 4
//   * all-reduce mode (README.md:6, unimplemented)   -> ring_allreduce_kernel / direct
// All cross-GPU traffic is issued from inside the kernels as loads/stores on peer-mapped
// pointers (NVLink 5 through NVSwitch); there is no NCCL call on these paths.
#include "common.cuh"
#include "../include/tiresias_b200.h"

using namespace tb200;

namespace {

constexpr int kThreads = 256;

struct alignas(16) F8 {
  float v[8];
};

template <typename T>
struct VecIO;

// 8 bf16 values per 16-byte vector.
template <>
struct VecIO<__nv_bfloat16> {
  static constexpr int kElems = 8;
  __device__ static void load_add(const void* base, size_t vec, F8& acc) {
    uint4 r = ld_stream_16(reinterpret_cast<const uint4*>(base) + vec);
    acc.v[0] += bf16_lo(r.x); acc.v[1] += bf16_hi(r.x);
    acc.v[2] += bf16_lo(r.y); acc.v[3] += bf16_hi(r.y);
    acc.v[4] += bf16_lo(r.z); acc.v[5] += bf16_hi(r.z);
    acc.v[6] += bf16_lo(r.w); acc.v[7] += bf16_hi(r.w);
  }
  __device__ static void store(void* base, size_t vec, const F8& x) {
    uint4 o;
    o.x = pack_bf16(x.v[0], x.v[1]); o.y = pack_bf16(x.v[2], x.v[3]);
    o.z = pack_bf16(x.v[4], x.v[5]); o.w = pack_bf16(x.v[6], x.v[7]);
    st_stream_16(reinterpret_cast<uint4*>(base) + vec, o);
  }
};

// fp32: one logical "vector" is still 8 elements = two 16-byte accesses.
template <>
struct VecIO<float> {
  static constexpr int kElems = 8;
  __device__ static void load_add(const void* base, size_t vec, F8& acc) {
    const uint4* p = reinterpret_cast<const uint4*>(base) + 2 * vec;
    uint4 a = ld_stream_16(p), b = ld_stream_16(p + 1);
    acc.v[0] += __uint_as_float(a.x); acc.v[1] += __uint_as_float(a.y);
    acc.v[2] += __uint_as_float(a.z); acc.v[3] += __uint_as_float(a.w);
    acc.v[4] += __uint_as_float(b.x); acc.v[5] += __uint_as_float(b.y);
    acc.v[6] += __uint_as_float(b.z); acc.v[7] += __uint_as_float(b.w);
  }
  __device__ static void store(void* base, size_t vec, const F8& x) {
    uint4 a, b;
    a.x = __float_as_uint(x.v[0]); a.y = __float_as_uint(x.v[1]);
    a.z = __float_as_uint(x.v[2]); a.w = __float_as_uint(x.v[3]);
    b.x = __float_as_uint(x.v[4]); b.y = __float_as_uint(x.v[5]);
    b.z = __float_as_uint(x.v[6]); b.w = __float_as_uint(x.v[7]);
    uint4* p = reinterpret_cast<uint4*>(base) + 2 * vec;
    st_stream_16(p, a);
    st_stream_16(p + 1, b);
  }
};

// Raw (unconverted) 8-element vector: lets a thread put many peer loads in flight before it touches any.
// NVLink needs ~2 MB in flight per GPU to run at link rate (770 GB/s x ~2.5 us); with the 32 CTAs a job
// may use next to its own backward pass that means 8-16 outstanding 16-byte loads per thread.
template <typename T>
struct RawVec {
  static constexpr int kQ = sizeof(T) * 8 / 16;
  uint4 q[kQ];
};
template <typename T>
__device__ __forceinline__ RawVec<T> load_raw(const void* base, size_t vec) {
  RawVec<T> r;
  const uint4* p = reinterpret_cast<const uint4*>(base) + vec * RawVec<T>::kQ;
#pragma unroll
  for (int i = 0; i < RawVec<T>::kQ; ++i) r.q[i] = ld_stream_16(p + i);
  return r;
}
__device__ __forceinline__ void add_raw(const RawVec<__nv_bfloat16>& r, F8& acc) {
  acc.v[0] += bf16_lo(r.q[0].x); acc.v[1] += bf16_hi(r.q[0].x);
  acc.v[2] += bf16_lo(r.q[0].y); acc.v[3] += bf16_hi(r.q[0].y);
  acc.v[4] += bf16_lo(r.q[0].z); acc.v[5] += bf16_hi(r.q[0].z);
  acc.v[6] += bf16_lo(r.q[0].w); acc.v[7] += bf16_hi(r.q[0].w);
}
__device__ __forceinline__ void add_raw(const RawVec<float>& r, F8& acc) {
  acc.v[0] += __uint_as_float(r.q[0].x); acc.v[1] += __uint_as_float(r.q[0].y);
  acc.v[2] += __uint_as_float(r.q[0].z); acc.v[3] += __uint_as_float(r.q[0].w);
  acc.v[4] += __uint_as_float(r.q[1].x); acc.v[5] += __uint_as_float(r.q[1].y);
  acc.v[6] += __uint_as_float(r.q[1].z); acc.v[7] += __uint_as_float(r.q[1].w);
}
// vectors per thread per round, chosen so that (peers x vectors x 16-byte units) stays near 16 loads
template <typename T, int W>
struct CommUnroll {
  static constexpr int kLoads = W * RawVec<T>::kQ;
  static constexpr int value = kLoads >= 16 ? 1 : (kLoads >= 8 ? 2 : (kLoads >= 4 ? 4 : 8));
};
constexpr int kCommThreads = 512;

__device__ __forceinline__ void sgd_update8(F8& g, float* master, float* mom, size_t elem,
                                            const tb200_sgd_hyper_t& h, F8& out) {
  float4* mp = reinterpret_cast<float4*>(master + elem);
  float4* vp = reinterpret_cast<float4*>(mom + elem);
  float4 m0 = mp[0], m1 = mp[1], v0 = vp[0], v1 = vp[1];
  float m[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
  float v[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    float gi = fmaf(g.v[i], h.grad_scale, h.weight_decay * m[i]);
    v[i] = fmaf(h.momentum, v[i], gi);
    m[i] = fmaf(-h.lr, v[i], m[i]);
    out.v[i] = m[i];
  }
  mp[0] = make_float4(m[0], m[1], m[2], m[3]);
  mp[1] = make_float4(m[4], m[5], m[6], m[7]);
  vp[0] = make_float4(v[0], v[1], v[2], v[3]);
  vp[1] = make_float4(v[4], v[5], v[6], v[7]);
}

// ---------------------------------------------------------------------------------------------
// K4: stand-alone optimizer step (the baseline-shaped path and world == 1 jobs).
// ---------------------------------------------------------------------------------------------
template <typename GT, typename PT>
__global__ void __launch_bounds__(kThreads) sgd_momentum_kernel(
    float* __restrict__ master, float* __restrict__ mom, const GT* __restrict__ grad,
    PT* __restrict__ param, size_t nvec, const tb200_sgd_hyper_t* __restrict__ hyper) {
  const tb200_sgd_hyper_t h = *hyper;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < nvec;
       i += (size_t)gridDim.x * blockDim.x) {
    F8 g = {};
    VecIO<GT>::load_add(grad, i, g);
    F8 p;
    sgd_update8(g, master, mom, i * 8, h, p);
    VecIO<PT>::store(param, i, p);
  }
}

// ---------------------------------------------------------------------------------------------
// K1 (+ store side of K2): push + reduce + SGD + broadcast for one bucket shard.
// ---------------------------------------------------------------------------------------------
struct PushArgs {
  tb200_world_t w;
  int bucket;
  size_t shard_off;   // elements
  size_t nvec;        // 8-element vectors in the shard
  float* master;      // shard-local
  float* mom;
  const tb200_sgd_hyper_t* hyper;
  const uint32_t* epoch;
  uint32_t* counter;
  int* status;
  unsigned long long budget;
  int broadcast;      // 1: store new params to every worker; 0: only to this PS's own buffer
};

__device__ __forceinline__ F8 unpack_bf16x8(const uint4& r) {
  F8 f;
  f.v[0] = bf16_lo(r.x); f.v[1] = bf16_hi(r.x); f.v[2] = bf16_lo(r.y); f.v[3] = bf16_hi(r.y);
  f.v[4] = bf16_lo(r.z); f.v[5] = bf16_hi(r.z); f.v[6] = bf16_lo(r.w); f.v[7] = bf16_hi(r.w);
  return f;
}
__device__ __forceinline__ uint4 pack_bf16x8(const F8& x) {
  uint4 o;
  o.x = pack_bf16(x.v[0], x.v[1]); o.y = pack_bf16(x.v[2], x.v[3]);
  o.z = pack_bf16(x.v[4], x.v[5]); o.w = pack_bf16(x.v[6], x.v[7]);
  return o;
}
constexpr int kMcUnroll = 8;     // multicast loads in flight per thread

// MC = NVLS: the W gradient slices are summed inside the NVSwitch by one multimem.ld_reduce per 16 bytes
// (this GPU receives n/W bytes instead of (W-1) n/W) and the new parameters reach all W workers with one
// multimem.st (this GPU sends n/W bytes instead of (W-1) n/W). bf16 gradients only.
template <typename GT, int W, bool MC = false>
__global__ void __launch_bounds__(kCommThreads) ps_push_pull_kernel(const PushArgs a) {
  const int rank = a.w.rank;
  const uint32_t epoch = *a.epoch;
  __shared__ int s_ok;
  if (threadIdx.x == 0) s_ok = 1;
  __syncthreads();

  // 1. My gradients for this bucket are final: the launch is stream-ordered behind the
  //    backward kernels that produced them. One CTA tells every PS.
  if (blockIdx.x == 0 && threadIdx.x < W)
    st_release_sys(a.w.flags[threadIdx.x] + TB200_FLAG_GRAD_READY(a.bucket, rank), epoch);

  // 2. Every worker's gradients are final (each CTA polls for itself: the acquire must be
  //    performed on the SM that will issue the peer loads).
  if (threadIdx.x < W) {
    if (!wait_epoch(a.w.flags[rank] + TB200_FLAG_GRAD_READY(a.bucket, threadIdx.x), epoch,
                    a.budget))
      s_ok = 0;
  }
  __syncthreads();
  if (!s_ok) {
    // Leave the shard untouched; the host watchdog sees status != 0. Still take part in the arrival counter
    // (without publishing PARAM_DONE): CTAs that did get through have incremented it, and a counter left
    // non-zero would let the NEXT launch publish before all of its CTAs have stored their parameters.
    if (threadIdx.x == 0) {
      atomicExch(a.status, 1);
      const unsigned prev = atomicAdd(a.counter, 1u);
      if (prev == gridDim.x - 1) *a.counter = 0;
    }
    return;
  }

  // 3-4. Reduce the W gradient slices in rank order, update, broadcast.
  const tb200_sgd_hyper_t h = *a.hyper;
  const size_t voff = a.shard_off / 8;  // shard_off is a multiple of 8 elements
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  if constexpr (MC) {
    const uint4* gmc = reinterpret_cast<const uint4*>(a.w.grad_mc) + voff;
    uint4* pmc = reinterpret_cast<uint4*>(a.w.param_mc) + voff;
    for (size_t i0 = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i0 < a.nvec; i0 += stride * kMcUnroll) {
      uint4 raw[kMcUnroll];
#pragma unroll
      for (int u = 0; u < kMcUnroll; ++u) {
        const size_t i = i0 + (size_t)u * stride;
        if (i < a.nvec) raw[u] = multimem_ld_reduce_bf16x8(gmc + i);
      }
#pragma unroll
      for (int u = 0; u < kMcUnroll; ++u) {
        const size_t i = i0 + (size_t)u * stride;
        if (i < a.nvec) {
          F8 g = unpack_bf16x8(raw[u]);
          F8 p;
          sgd_update8(g, a.master, a.mom, i * 8, h, p);
          if (a.broadcast) multimem_st_16(pmc + i, pack_bf16x8(p));
          else VecIO<__nv_bfloat16>::store(a.w.param[rank], voff + i, p);
        }
      }
    }
  } else {
  constexpr int U = CommUnroll<GT, W>::value;
  for (size_t i0 = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i0 < a.nvec; i0 += stride * U) {
    RawVec<GT> raw[U][W];
#pragma unroll
    for (int u = 0; u < U; ++u) {                      // all peer loads of this round go out first
      const size_t i = i0 + (size_t)u * stride;
      if (i < a.nvec) {
#pragma unroll
        for (int w = 0; w < W; ++w) raw[u][w] = load_raw<GT>(a.w.grad[w], voff + i);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = i0 + (size_t)u * stride;
      if (i < a.nvec) {
        F8 g = {};
#pragma unroll
        for (int w = 0; w < W; ++w) add_raw(raw[u][w], g);          // fixed rank order: reproducible
        F8 p;
        sgd_update8(g, a.master, a.mom, i * 8, h, p);
#pragma unroll
        for (int w = 0; w < W; ++w)
          if (a.broadcast || w == rank) VecIO<__nv_bfloat16>::store(a.w.param[w], voff + i, p);
      }
    }
  }
  }

  // 5. Publish: all CTAs' stores must be visible system-wide before the flag.
  __threadfence_system();
  __syncthreads();
  __shared__ int s_last;
  if (threadIdx.x == 0) {
    unsigned prev = atomicAdd(a.counter, 1u);
    s_last = (prev == gridDim.x - 1);
    if (s_last) *a.counter = 0;  // ready for the next launch on this stream
  }
  __syncthreads();
  if (s_last) {
    __threadfence_system();
    if (threadIdx.x < W)
      st_release_sys(a.w.flags[threadIdx.x] + TB200_FLAG_PARAM_DONE(a.bucket, rank), epoch);
  }
}

// Waits for all PARAM_DONE flags of this step, then bumps the step counter.
__global__ void step_end_kernel(const tb200_world_t w, int num_buckets, uint32_t* epoch_ptr,
                                int* status, unsigned long long budget) {
  const uint32_t epoch = *epoch_ptr;
  __shared__ int s_ok;
  if (threadIdx.x == 0) s_ok = 1;
  __syncthreads();
  const int total = num_buckets * w.world;
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    int b = i / w.world, s = i % w.world;
    if (!wait_epoch(w.flags[w.rank] + TB200_FLAG_PARAM_DONE(b, s), epoch, budget)) s_ok = 0;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    if (!s_ok) atomicExch(status, 2);
    *epoch_ptr = epoch + 1;
  }
}

// ---------------------------------------------------------------------------------------------
// K3: all-reduce mode.
//
// Ring: the buffer is cut into W chunks; every CTA ("lane") owns the same sub-range of every
// chunk and talks only to the same lane on the neighbouring ranks, so no grid-wide barrier is
// needed. Step s of reduce-scatter: buf[c] += prev.buf[c] with c = (rank - s - 1) mod W, legal
// once prev has finished s steps. All-gather step t copies chunk (rank - t) mod W from prev.
// Each lane publishes its progress to next (data dependency) and to prev (so prev does not
// return and let its caller overwrite a chunk that is still being read).
// ---------------------------------------------------------------------------------------------
struct RingArgs {
  tb200_world_t w;
  void* bufs[TB200_MAX_WORLD];
  size_t n;          // elements (multiple of 8 * W)
  float scale;
  uint32_t* seq;     // per-call sequence number (device)
  int* status;
  unsigned long long budget;
};

template <typename T>
__global__ void __launch_bounds__(kCommThreads) ring_allreduce_kernel(const RingArgs a) {
  const int W = a.w.world, rank = a.w.rank;
  const int prev = (rank + W - 1) % W, next = (rank + 1) % W;
  const int lane = blockIdx.x;
  const uint32_t steps = 2u * (W - 1);
  // progress value p means "lane has completed p - base steps of this call".
  const uint32_t base = (*a.seq) * (steps + 1) + 1;

  const size_t chunk_vec = a.n / 8 / W;
  const size_t per_lane = (chunk_vec + gridDim.x - 1) / gridDim.x;
  const size_t lo = min(chunk_vec, (size_t)lane * per_lane);
  const size_t hi = min(chunk_vec, lo + per_lane);

  uint32_t* my_from_prev = a.w.flags[rank] + TB200_FLAG_RING(0, lane);  // written by prev
  uint32_t* my_from_next = a.w.flags[rank] + TB200_FLAG_RING(1, lane);  // written by next
  uint32_t* tell_next = a.w.flags[next] + TB200_FLAG_RING(0, lane);
  uint32_t* tell_prev = a.w.flags[prev] + TB200_FLAG_RING(1, lane);

  __shared__ int s_ok;
  if (threadIdx.x == 0) {
    s_ok = 1;
    st_release_sys(tell_next, base);
    st_release_sys(tell_prev, base);
  }
  __syncthreads();

  T* mine = reinterpret_cast<T*>(a.bufs[rank]);
  const T* theirs = reinterpret_cast<const T*>(a.bufs[prev]);

  for (uint32_t s = 0; s < steps; ++s) {
    if (threadIdx.x == 0 && !wait_epoch(my_from_prev, base + s, a.budget)) s_ok = 0;
    __syncthreads();
    if (!s_ok) break;
    const bool reduce = s < (uint32_t)(W - 1);
    const int c = reduce ? (rank - (int)s - 1 + 2 * W) % W
                         : (rank - (int)(s - (W - 1)) + 2 * W) % W;
    const size_t cbase = (size_t)c * chunk_vec;
    const bool last_reduce = (s == (uint32_t)(W - 2));
    constexpr int U = 4;
    for (size_t i0 = lo + threadIdx.x; i0 < hi; i0 += (size_t)blockDim.x * U) {
      RawVec<T> rt[U], rm[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {                    // peer (and local) loads first, then the math
        const size_t i = i0 + (size_t)u * blockDim.x;
        if (i < hi) {
          rt[u] = load_raw<T>(theirs, cbase + i);
          if (reduce) rm[u] = load_raw<T>(mine, cbase + i);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const size_t i = i0 + (size_t)u * blockDim.x;
        if (i < hi) {
          F8 acc = {};
          add_raw(rt[u], acc);
          if (reduce) {
            add_raw(rm[u], acc);
            if (last_reduce) {
#pragma unroll
              for (int k = 0; k < 8; ++k) acc.v[k] *= a.scale;
            }
          }
          VecIO<T>::store(mine, cbase + i, acc);
        }
      }
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
      st_release_sys(tell_next, base + s + 1);
      st_release_sys(tell_prev, base + s + 1);
    }
  }
  // Do not return while next may still be reading my buffer.
  if (threadIdx.x == 0 && s_ok && !wait_epoch(my_from_next, base + steps, a.budget)) s_ok = 0;
  __syncthreads();
  if (!s_ok && threadIdx.x == 0) atomicExch(a.status, 3);
  // The host bumps *seq with a trailing 1-thread kernel (bump_seq_kernel) so that every lane
  // of this launch reads the same value.
}

__global__ void bump_seq_kernel(uint32_t* seq) { *seq += 1; }

// Direct two-shot: rank r owns slice r; it reads that slice from all W ranks (NVSwitch gives
// every peer full bandwidth), reduces in rank order, and stores the result to all W ranks.
// Uses the bucket-0 GRAD_READY / PARAM_DONE flag slots of bucket index TB200_MAX_BUCKETS-1.
struct DirectArgs {
  tb200_world_t w;
  void* mc;                     // NVLS multicast view of bufs[0..W-1] (algo 2), else NULL
  void* bufs[TB200_MAX_WORLD];
  size_t n;
  float scale;
  uint32_t* seq;
  uint32_t* counter;
  int* status;
  unsigned long long budget;
};

template <typename T, int W, bool MC = false>
__global__ void __launch_bounds__(kCommThreads) direct_allreduce_kernel(const DirectArgs a) {
  const int rank = a.w.rank;
  const int slot = TB200_MAX_BUCKETS - 1;
  const uint32_t epoch = *a.seq + 1;
  __shared__ int s_ok;
  if (threadIdx.x == 0) s_ok = 1;
  __syncthreads();
  if (blockIdx.x == 0 && threadIdx.x < W)
    st_release_sys(a.w.flags[threadIdx.x] + TB200_FLAG_GRAD_READY(slot, rank), epoch);
  if (threadIdx.x < W &&
      !wait_epoch(a.w.flags[rank] + TB200_FLAG_GRAD_READY(slot, threadIdx.x), epoch, a.budget))
    s_ok = 0;
  __syncthreads();
  if (!s_ok) {
    if (threadIdx.x == 0) {          // keep the arrival counter consistent for the next launch (see ps_push_pull)
      atomicExch(a.status, 4);
      const unsigned prev = atomicAdd(a.counter, 1u);
      if (prev == gridDim.x - 1) *a.counter = 0;
    }
    return;
  }
  const size_t slice_vec = a.n / 8 / W;
  const size_t vbase = (size_t)rank * slice_vec;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  if constexpr (MC) {
    // one-shot NVLS: the switch sums my slice of everybody's buffer; one multicast store returns it
    uint4* mc = reinterpret_cast<uint4*>(a.mc) + vbase;
    for (size_t i0 = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i0 < slice_vec; i0 += stride * kMcUnroll) {
      uint4 raw[kMcUnroll];
#pragma unroll
      for (int u = 0; u < kMcUnroll; ++u) {
        const size_t i = i0 + (size_t)u * stride;
        if (i < slice_vec) raw[u] = multimem_ld_reduce_bf16x8(mc + i);
      }
#pragma unroll
      for (int u = 0; u < kMcUnroll; ++u) {
        const size_t i = i0 + (size_t)u * stride;
        if (i < slice_vec) {
          F8 acc = unpack_bf16x8(raw[u]);
#pragma unroll
          for (int k = 0; k < 8; ++k) acc.v[k] *= a.scale;
          multimem_st_16(mc + i, pack_bf16x8(acc));
        }
      }
    }
  } else {
  constexpr int U = CommUnroll<T, W>::value;
  for (size_t i0 = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i0 < slice_vec; i0 += stride * U) {
    RawVec<T> raw[U][W];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = i0 + (size_t)u * stride;
      if (i < slice_vec) {
#pragma unroll
        for (int w = 0; w < W; ++w) raw[u][w] = load_raw<T>(a.bufs[w], vbase + i);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = i0 + (size_t)u * stride;
      if (i < slice_vec) {
        F8 acc = {};
#pragma unroll
        for (int w = 0; w < W; ++w) add_raw(raw[u][w], acc);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc.v[k] *= a.scale;
#pragma unroll
        for (int w = 0; w < W; ++w) VecIO<T>::store(a.bufs[w], vbase + i, acc);
      }
    }
  }
  }
  __threadfence_system();
  __syncthreads();
  __shared__ int s_last;
  if (threadIdx.x == 0) {
    unsigned prev = atomicAdd(a.counter, 1u);
    s_last = (prev == gridDim.x - 1);
    if (s_last) *a.counter = 0;
  }
  __syncthreads();
  if (s_last) {
    __threadfence_system();
    if (threadIdx.x < W)
      st_release_sys(a.w.flags[threadIdx.x] + TB200_FLAG_PARAM_DONE(slot, rank), epoch);
    __syncthreads();
  }
  // Everyone waits until all slices have landed locally (and, symmetrically, all peers have
  // finished reading this rank's buffer) before the kernel retires.
  if (threadIdx.x < W &&
      !wait_epoch(a.w.flags[rank] + TB200_FLAG_PARAM_DONE(slot, threadIdx.x), epoch, a.budget))
    atomicExch(a.status, 4);
}

// ---------------------------------------------------------------------------------------------
// Small fused helpers
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) u8_to_bf16_norm_kernel(
    const uint8_t* __restrict__ src, __nv_bfloat16* __restrict__ dst, size_t pixels, float m0,
    float m1, float m2, float s0, float s1, float s2) {
  // 4 pixels (12 bytes in, 24 bytes out) per thread keeps both sides vector-aligned.
  const size_t groups = pixels / 4;
  for (size_t g = blockIdx.x * (size_t)blockDim.x + threadIdx.x; g < groups;
       g += (size_t)gridDim.x * blockDim.x) {
    const uint32_t* p = reinterpret_cast<const uint32_t*>(src + g * 12);
    uint32_t a = p[0], b = p[1], c = p[2];
    uint8_t u[12] = {uint8_t(a), uint8_t(a >> 8), uint8_t(a >> 16), uint8_t(a >> 24),
                     uint8_t(b), uint8_t(b >> 8), uint8_t(b >> 16), uint8_t(b >> 24),
                     uint8_t(c), uint8_t(c >> 8), uint8_t(c >> 16), uint8_t(c >> 24)};
    const float mean[3] = {m0, m1, m2}, istd[3] = {s0, s1, s2};
    uint32_t o[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      int e0 = 2 * k, e1 = 2 * k + 1;
      float f0 = (u[e0] * (1.0f / 255.0f) - mean[e0 % 3]) * istd[e0 % 3];
      float f1 = (u[e1] * (1.0f / 255.0f) - mean[e1 % 3]) * istd[e1 % 3];
      o[k] = pack_bf16(f0, f1);
    }
    uint2* q = reinterpret_cast<uint2*>(dst + g * 12);
    q[0] = make_uint2(o[0], o[1]);
    q[1] = make_uint2(o[2], o[3]);
    q[2] = make_uint2(o[4], o[5]);
  }
  // tail pixels (pixels % 4)
  if (blockIdx.x == 0 && threadIdx.x < (pixels % 4) * 3) {
    size_t e = groups * 12 + threadIdx.x;
    const float mean[3] = {m0, m1, m2}, istd[3] = {s0, s1, s2};
    dst[e] = __float2bfloat16((src[e] * (1.0f / 255.0f) - mean[e % 3]) * istd[e % 3]);
  }
}

// 64x64 bf16 tile transpose through padded shared memory.
__global__ void __launch_bounds__(256) transpose_bf16_kernel(const __nv_bfloat16* __restrict__ src,
                                                             __nv_bfloat16* __restrict__ dst,
                                                             int rows, int cols) {
  __shared__ __nv_bfloat16 tile[64][66];
  const int r0 = blockIdx.y * 64, c0 = blockIdx.x * 64;
  const int tx = threadIdx.x % 32, ty = threadIdx.x / 32;  // 32 x 8
  for (int i = ty; i < 64; i += 8) {
    int r = r0 + i;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      int c = c0 + tx * 2 + j;
      tile[i][tx * 2 + j] = (r < rows && c < cols) ? src[(size_t)r * cols + c] : __nv_bfloat16(0.f);
    }
  }
  __syncthreads();
  for (int i = ty; i < 64; i += 8) {
    int c = c0 + i;  // becomes the output row
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      int r = r0 + tx * 2 + j;
      if (c < cols && r < rows) dst[(size_t)c * rows + r] = tile[tx * 2 + j][i];
    }
  }
}

constexpr int kSumsqBlocks = 592;  // 4 per SM on 148 SMs

template <typename T>
__global__ void __launch_bounds__(kThreads) sumsq_partial_kernel(const T* __restrict__ x, size_t n,
                                                                 float* __restrict__ partials) {
  float acc = 0.f;
  const size_t nvec = n / 8;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < nvec;
       i += (size_t)gridDim.x * blockDim.x) {
    F8 v = {};
    VecIO<T>::load_add(x, i, v);
#pragma unroll
    for (int k = 0; k < 8; ++k) acc = fmaf(v.v[k], v.v[k], acc);
  }
  if (blockIdx.x == 0) {
    for (size_t e = nvec * 8 + threadIdx.x; e < n; e += blockDim.x) {
      float f = static_cast<float>(x[e]);
      acc = fmaf(f, f, acc);
    }
  }
  __shared__ float red[kThreads / 32];
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < kThreads / 32 ? red[threadIdx.x] : 0.f;
    for (int o = 4; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (threadIdx.x == 0) partials[blockIdx.x] = v;
  }
}

__global__ void __launch_bounds__(1024) sumsq_final_kernel(const float* __restrict__ partials,
                                                           int count, float* __restrict__ out) {
  __shared__ float red[32];
  float v = 0.f;
  for (int i = threadIdx.x; i < count; i += blockDim.x) v += partials[i];
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x < 32) {
    float t = red[threadIdx.x];
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if (threadIdx.x == 0) out[0] = t;
  }
}

int grid_for(size_t nvec, int per_thread, int max_ctas) {
  size_t want = (nvec + (size_t)kThreads * per_thread - 1) / ((size_t)kThreads * per_thread);
  if (want < 1) want = 1;
  if (want > (size_t)max_ctas) want = max_ctas;
  return (int)want;
}

}  // namespace

// Same normalisation, but every pixel is written as 8 bf16 channels (r, g, b, 0, 0, 0, 0, 0 = one
// 16-byte store): a first convolution with 8 input channels takes cuDNN's sm_100 implicit-GEMM path,
// which needs channel counts that are multiples of 8; with 3 channels it falls back to a legacy kernel
// that is ~4x slower (profiles/step_launches: cutlass__5x_cudnn fprop_optimized 1.47 ms at batch 256).
__global__ void __launch_bounds__(kThreads) u8_to_bf16_norm_pad8_kernel(
    const uint8_t* __restrict__ src, uint4* __restrict__ dst, size_t pixels, float m0, float m1,
    float m2, float s0, float s1, float s2) {
  const size_t groups = pixels / 4;
  const float k = 1.0f / 255.0f;
  for (size_t g = blockIdx.x * (size_t)blockDim.x + threadIdx.x; g < groups;
       g += (size_t)gridDim.x * blockDim.x) {
    const uint32_t* p = reinterpret_cast<const uint32_t*>(src + g * 12);
    const uint32_t w[3] = {p[0], p[1], p[2]};
#pragma unroll
    for (int px = 0; px < 4; ++px) {
      float c[3];
#pragma unroll
      for (int ch = 0; ch < 3; ++ch) {
        const int e = px * 3 + ch;
        c[ch] = (float)((w[e / 4] >> (8 * (e % 4))) & 0xffu) * k;
      }
      dst[g * 4 + px] = make_uint4(pack_bf16((c[0] - m0) * s0, (c[1] - m1) * s1),
                                   pack_bf16((c[2] - m2) * s2, 0.f), 0u, 0u);
    }
  }
  if (blockIdx.x == 0 && threadIdx.x < pixels % 4) {
    const size_t px = groups * 4 + threadIdx.x;
    const uint8_t* q = src + px * 3;
    dst[px] = make_uint4(pack_bf16((q[0] * k - m0) * s0, (q[1] * k - m1) * s1),
                         pack_bf16((q[2] * k - m2) * s2, 0.f), 0u, 0u);
  }
}

// ---------------------------------------------------------------------------------------------
// Gradient packing: up to TB200_PACK_MAX tensors are copied into their slots of the flat, symmetric
// gradient buffer by ONE launch (autograd hands every parameter's gradient over as a separate
// tensor; accumulating each into a flat view costs one add kernel per parameter — 161 launches per
// ResNet-50 step — where one launch per bucket suffices).
// ---------------------------------------------------------------------------------------------
constexpr unsigned kPackElemsPerBlock = 8192;

struct PackArgs {
  int count;
  tb200_pack_item_t items[TB200_PACK_MAX];
  unsigned block_start[TB200_PACK_MAX + 1];
};

__global__ void __launch_bounds__(kThreads) pack_bf16_kernel(const PackArgs a) {
  int it = 0;
  while (it + 1 < a.count && blockIdx.x >= a.block_start[it + 1]) ++it;     // <= 64 items: linear scan
  const tb200_pack_item_t item = a.items[it];
  const size_t begin = (size_t)(blockIdx.x - a.block_start[it]) * kPackElemsPerBlock;
  const size_t end = begin + kPackElemsPerBlock < item.n ? begin + kPackElemsPerBlock : (size_t)item.n;
  const __nv_bfloat16* src = reinterpret_cast<const __nv_bfloat16*>(item.src);
  __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(item.dst);
  if ((((uintptr_t)src | (uintptr_t)dst) & 15) == 0) {
    const size_t v0 = begin / 8, v1 = end / 8;                  // begin is a multiple of 8
    for (size_t v = v0 + threadIdx.x; v < v1; v += kThreads)
      st_stream_16(reinterpret_cast<uint4*>(dst) + v, ld_stream_16(reinterpret_cast<const uint4*>(src) + v));
    for (size_t i = v1 * 8 + threadIdx.x; i < end; i += kThreads) dst[i] = src[i];
  } else {
    for (size_t i = begin + threadIdx.x; i < end; i += kThreads) dst[i] = src[i];
  }
}

extern "C" {

int tb200_sgd_momentum(float* master, float* momentum, const void* grad, int grad_dtype,
                       void* param_out, int param_dtype, size_t n,
                       const tb200_sgd_hyper_t* hyper_dev, cudaStream_t stream) {
  if (n % 8) return fail("tb200_sgd_momentum: n %s(%ld) must be a multiple of 8", "", (long)n);
  if (n == 0) return 0;
  const size_t nvec = n / 8;
  const int grid = grid_for(nvec, 2, 148 * 8);
#define TB_LAUNCH(GT, PT)                                                              \
  sgd_momentum_kernel<GT, PT><<<grid, kThreads, 0, stream>>>(                          \
      master, momentum, reinterpret_cast<const GT*>(grad), reinterpret_cast<PT*>(param_out), \
      nvec, hyper_dev)
  if (grad_dtype == TB200_BF16 && param_dtype == TB200_BF16) TB_LAUNCH(__nv_bfloat16, __nv_bfloat16);
  else if (grad_dtype == TB200_F32 && param_dtype == TB200_BF16) TB_LAUNCH(float, __nv_bfloat16);
  else if (grad_dtype == TB200_BF16 && param_dtype == TB200_F32) TB_LAUNCH(__nv_bfloat16, float);
  else if (grad_dtype == TB200_F32 && param_dtype == TB200_F32) TB_LAUNCH(float, float);
  else return fail("tb200_sgd_momentum: bad dtype%s");
#undef TB_LAUNCH
  return launched();
}

int tb200_ps_push_pull(const tb200_world_t* world, int bucket, size_t shard_off,
                       size_t shard_len, int grad_dtype, float* master, float* momentum,
                       const tb200_sgd_hyper_t* hyper_dev, const uint32_t* epoch_dev,
                       uint32_t* counter_dev, int* status_dev,
                       unsigned long long spin_budget, int max_ctas, int broadcast,
                       cudaStream_t stream) {
  if (!world || world->world < 1 || world->world > TB200_MAX_WORLD)
    return fail("tb200_ps_push_pull: bad world%s");
  if (bucket < 0 || bucket >= TB200_MAX_BUCKETS - 1)
    return fail("tb200_ps_push_pull: bucket %s%ld out of range", "", bucket);
  if ((shard_off % 8) || (shard_len % 8))
    return fail("tb200_ps_push_pull: shard offset/length must be multiples of 8 elements%s");
  PushArgs a;
  a.w = *world;
  a.bucket = bucket;
  a.shard_off = shard_off;
  a.nvec = shard_len / 8;
  a.master = master;
  a.mom = momentum;
  a.hyper = hyper_dev;
  a.epoch = epoch_dev;
  a.counter = counter_dev;
  a.status = status_dev;
  a.budget = spin_budget;
  a.broadcast = broadcast;
  if (max_ctas <= 0) max_ctas = 32;
  const int grid = grid_for(a.nvec * kThreads / kCommThreads, 2, max_ctas);
#define TB_CASE(GT, WW)                                                       \
  case WW:                                                                    \
    ps_push_pull_kernel<GT, WW><<<grid, kCommThreads, 0, stream>>>(a);        \
    break
#define TB_SWITCH(GT)                                                         \
  switch (world->world) {                                                     \
    TB_CASE(GT, 1); TB_CASE(GT, 2); TB_CASE(GT, 3); TB_CASE(GT, 4);           \
    TB_CASE(GT, 5); TB_CASE(GT, 6); TB_CASE(GT, 7); TB_CASE(GT, 8);           \
  }
  if (grad_dtype == TB200_BF16 && world->world > 1 && world->grad_mc && world->param_mc) {
    // NVLS path: in-switch reduction and multicast stores
    switch (world->world) {
#define TB_MC_CASE(WW)                                                                    \
  case WW:                                                                                \
    ps_push_pull_kernel<__nv_bfloat16, WW, true><<<grid, kCommThreads, 0, stream>>>(a);   \
    break
      TB_MC_CASE(2); TB_MC_CASE(3); TB_MC_CASE(4); TB_MC_CASE(5); TB_MC_CASE(6); TB_MC_CASE(7); TB_MC_CASE(8);
#undef TB_MC_CASE
    }
  }
  else if (grad_dtype == TB200_BF16) { TB_SWITCH(__nv_bfloat16) }
  else if (grad_dtype == TB200_F32) { TB_SWITCH(float) }
  else return fail("tb200_ps_push_pull: bad grad dtype%s");
#undef TB_SWITCH
#undef TB_CASE
  return launched();
}

int tb200_step_end(const tb200_world_t* world, int num_buckets, uint32_t* epoch_dev,
                   int* status_dev, unsigned long long spin_budget, cudaStream_t stream) {
  if (!world || num_buckets < 0 || num_buckets >= TB200_MAX_BUCKETS)
    return fail("tb200_step_end: bad arguments%s");
  step_end_kernel<<<1, 256, 0, stream>>>(*world, num_buckets, epoch_dev, status_dev, spin_budget);
  return launched();
}

int tb200_allreduce(const tb200_world_t* world, void* const bufs[TB200_MAX_WORLD], size_t n,
                    int dtype, float scale, int algo, uint32_t* seq_dev, int* status_dev,
                    unsigned long long spin_budget, int num_ctas, cudaStream_t stream) {
  if (!world || world->world < 1 || world->world > TB200_MAX_WORLD)
    return fail("tb200_allreduce: bad world%s");
  const int W = world->world;
  if (n % (8 * (size_t)W))
    return fail("tb200_allreduce: n %s(%ld) must be a multiple of 8*world", "", (long)n);
  if (num_ctas <= 0) num_ctas = 32;
  if (num_ctas > TB200_MAX_RING_CTAS) num_ctas = TB200_MAX_RING_CTAS;
  if (algo == 0) {
    RingArgs a;
    a.w = *world;
    for (int i = 0; i < TB200_MAX_WORLD; ++i) a.bufs[i] = i < W ? bufs[i] : nullptr;
    a.n = n;
    a.scale = scale;
    a.seq = seq_dev;
    a.status = status_dev;
    a.budget = spin_budget;
    if (W == 1) {
      // nothing to exchange; only the scale applies (kept on the GPU for stream ordering)
      num_ctas = 1;
    }
    if (dtype == TB200_BF16) ring_allreduce_kernel<__nv_bfloat16><<<num_ctas, kCommThreads, 0, stream>>>(a);
    else if (dtype == TB200_F32) ring_allreduce_kernel<float><<<num_ctas, kCommThreads, 0, stream>>>(a);
    else return fail("tb200_allreduce: bad dtype%s");
    if (int rc = launched()) return rc;
    bump_seq_kernel<<<1, 1, 0, stream>>>(seq_dev);
    return launched();
  }
  DirectArgs a;
  a.w = *world;
  a.mc = nullptr;
  if (algo == 2) {
    // NVLS: `bufs` must be the gradient buffers of `world` (their multicast view is world->grad_mc)
    if (dtype != TB200_BF16) return fail("tb200_allreduce: the NVLS algorithm reduces bf16 only%s");
    if (W > 1) {
      // `bufs` must point into the gradient buffers of `world` (same offset on every rank): their
      // multicast view is world->grad_mc + that offset
      if (!world->grad_mc) return fail("tb200_allreduce: algo 2 needs world->grad_mc%s");
      const ptrdiff_t off = reinterpret_cast<const char*>(bufs[world->rank]) -
                            reinterpret_cast<const char*>(world->grad[world->rank]);
      if (off < 0 || off % 16) return fail("tb200_allreduce: algo 2: bufs must lie inside world->grad%s");
      a.mc = const_cast<char*>(reinterpret_cast<const char*>(world->grad_mc)) + off;
    }
  }
  for (int i = 0; i < TB200_MAX_WORLD; ++i) a.bufs[i] = i < W ? bufs[i] : nullptr;
  a.n = n;
  a.scale = scale;
  a.seq = seq_dev;
  a.counter = seq_dev + 1;  // seq_dev points at two words: {sequence, arrival counter}
  a.status = status_dev;
  a.budget = spin_budget;
  const int grid = grid_for(n / 8 / W, 2, num_ctas);
#define TB_CASE(T, WW)                                                        \
  case WW:                                                                    \
    direct_allreduce_kernel<T, WW><<<grid, kCommThreads, 0, stream>>>(a);     \
    break
#define TB_SWITCH(T)                                                          \
  switch (W) {                                                                \
    TB_CASE(T, 1); TB_CASE(T, 2); TB_CASE(T, 3); TB_CASE(T, 4);               \
    TB_CASE(T, 5); TB_CASE(T, 6); TB_CASE(T, 7); TB_CASE(T, 8);               \
  }
  if (a.mc && W > 1) {
    switch (W) {
#define TB_MC_CASE(WW)                                                                          \
  case WW:                                                                                      \
    direct_allreduce_kernel<__nv_bfloat16, WW, true><<<grid, kCommThreads, 0, stream>>>(a);     \
    break
      TB_MC_CASE(2); TB_MC_CASE(3); TB_MC_CASE(4); TB_MC_CASE(5); TB_MC_CASE(6); TB_MC_CASE(7); TB_MC_CASE(8);
#undef TB_MC_CASE
    }
  }
  else if (dtype == TB200_BF16) { TB_SWITCH(__nv_bfloat16) }
  else if (dtype == TB200_F32) { TB_SWITCH(float) }
  else return fail("tb200_allreduce: bad dtype%s");
#undef TB_SWITCH
#undef TB_CASE
  if (int rc = launched()) return rc;
  bump_seq_kernel<<<1, 1, 0, stream>>>(seq_dev);
  return launched();
}

int tb200_u8_to_bf16_norm(const uint8_t* src, void* dst, size_t pixels, const float mean[3],
                          const float inv_std[3], cudaStream_t stream) {
  if (pixels == 0) return 0;
  const int grid = grid_for(pixels / 4 + 1, 2, 148 * 8);
  u8_to_bf16_norm_kernel<<<grid, kThreads, 0, stream>>>(
      src, reinterpret_cast<__nv_bfloat16*>(dst), pixels, mean[0], mean[1], mean[2], inv_std[0],
      inv_std[1], inv_std[2]);
  return launched();
}

int tb200_u8_to_bf16_norm_pad8(const uint8_t* src, void* dst, size_t pixels, const float mean[3],
                               const float inv_std[3], cudaStream_t stream) {
  if (pixels == 0) return 0;
  const int grid = grid_for(pixels / 4 + 1, 2, 148 * 8);
  u8_to_bf16_norm_pad8_kernel<<<grid, kThreads, 0, stream>>>(
      src, reinterpret_cast<uint4*>(dst), pixels, mean[0], mean[1], mean[2], inv_std[0], inv_std[1],
      inv_std[2]);
  return launched();
}

int tb200_pack_bf16(const tb200_pack_item_t* items, int count, cudaStream_t stream) {
  if (count <= 0) return 0;
  if (!items) return fail("tb200_pack_bf16: null items%s");
  for (int base = 0; base < count; base += TB200_PACK_MAX) {
    PackArgs a{};
    a.count = count - base < TB200_PACK_MAX ? count - base : TB200_PACK_MAX;
    unsigned blocks = 0;
    for (int i = 0; i < a.count; ++i) {
      a.items[i] = items[base + i];
      a.block_start[i] = blocks;
      blocks += (unsigned)((a.items[i].n + kPackElemsPerBlock - 1) / kPackElemsPerBlock);
    }
    a.block_start[a.count] = blocks;
    if (blocks == 0) continue;
    pack_bf16_kernel<<<blocks, kThreads, 0, stream>>>(a);
    if (int rc = launched()) return rc;
  }
  return 0;
}

int tb200_transpose_bf16(const void* src, void* dst, int rows, int cols, cudaStream_t stream) {
  if (rows <= 0 || cols <= 0) return 0;
  dim3 grid((cols + 63) / 64, (rows + 63) / 64);
  transpose_bf16_kernel<<<grid, 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(src),
                                                  reinterpret_cast<__nv_bfloat16*>(dst), rows, cols);
  return launched();
}

size_t tb200_sumsq_scratch_floats(void) { return kSumsqBlocks; }

int tb200_sumsq(const void* x, int dtype, size_t n, float* partials, float* out,
                cudaStream_t stream) {
  if (!partials || !out) return fail("tb200_sumsq: null scratch/out%s");
  if (dtype == TB200_BF16)
    sumsq_partial_kernel<__nv_bfloat16><<<kSumsqBlocks, kThreads, 0, stream>>>(
        reinterpret_cast<const __nv_bfloat16*>(x), n, partials);
  else if (dtype == TB200_F32)
    sumsq_partial_kernel<float><<<kSumsqBlocks, kThreads, 0, stream>>>(
        reinterpret_cast<const float*>(x), n, partials);
  else return fail("tb200_sumsq: bad dtype%s");
  if (int rc = launched()) return rc;
  sumsq_final_kernel<<<1, 1024, 0, stream>>>(partials, kSumsqBlocks, out);
  return launched();
}

}  // extern "C"
