// Shared device/host helpers for the tiresias_b200 native library (sm_100a only).
//
// Everything here is new code: the reference (matthewygf/Tiresias) has no native
// component at all; these helpers back the kernels that replace the
// reference's *analytic* PS/worker cost model (core/network/network_service.py:3-39).
#pragma once

#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <atomic>

namespace tb200 {

// ---------------------------------------------------------------------------------------------
// Host-side error plumbing. Every extern "C" entry point returns 0 on success and a non-zero
// code on failure; the message is kept per-thread and read with tb200_last_error().
// ---------------------------------------------------------------------------------------------
extern thread_local char g_last_error[512];
extern std::atomic<unsigned long long> g_launch_count;

// `fmt` may hold one %s (it must come first) followed by up to two %ld. A second %s would make snprintf read a
// long as a pointer (it did, in two messages of round 2, and crashed the caller instead of reporting): refuse it.
inline int fail(const char* fmt, const char* a = "", long b = 0, long c = 0) {
  int strings = 0;
  for (const char* p = fmt; *p; ++p)
    if (p[0] == '%' && p[1] == 's') ++strings;
  if (strings > 1) snprintf(g_last_error, sizeof(g_last_error), "(malformed error format) %.400s", fmt);
  else snprintf(g_last_error, sizeof(g_last_error), fmt, a, b, c);
  return 1;
}

#define TB_CUDA_OK(expr)                                                                   \
  do {                                                                                     \
    cudaError_t _e = (expr);                                                               \
    if (_e != cudaSuccess) {                                                               \
      snprintf(tb200::g_last_error, sizeof(tb200::g_last_error), "%s failed: %s (%s:%d)",  \
               #expr, cudaGetErrorString(_e), __FILE__, __LINE__);                         \
      return 2;                                                                            \
    }                                                                                      \
  } while (0)

// One place that counts kernel launches made by this library (bench.py reports it as
// "gpu_launches"). Called right after every <<<>>> / cudaLaunchKernelEx.
inline int launched() {
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    snprintf(g_last_error, sizeof(g_last_error), "kernel launch failed: %s", cudaGetErrorString(e));
    return 3;
  }
  return 0;
}

enum DType : int { kF32 = 0, kBF16 = 1 };

// Per-device one-time set-up. Function attributes (cudaFuncAttributeMaxDynamicSharedMemorySize) and the SM
// count belong to a device, not to the process: a host that drives several GPUs from one process must
// configure every kernel on every GPU it launches on. `DeviceOnce` remembers, per kernel instantiation,
// which devices are done (lock-free; a race only repeats the cheap, idempotent call).
constexpr int kMaxDevices = 64;

inline int current_device() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return 0;
  return dev;
}

struct DeviceOnce {
  std::atomic<bool> done[kMaxDevices];
  bool need(int dev) const { return !done[dev].load(std::memory_order_acquire); }
  void mark(int dev) { done[dev].store(true, std::memory_order_release); }
};

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) once per (kernel, device).
template <typename Kern>
inline int ensure_dynamic_smem(DeviceOnce& once, Kern kern, int bytes) {
  const int dev = current_device();
  if (once.need(dev)) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) {
      snprintf(g_last_error, sizeof(g_last_error), "cudaFuncSetAttribute(max dynamic smem %d) failed: %s", bytes,
               cudaGetErrorString(e));
      return 2;
    }
    once.mark(dev);
  }
  return 0;
}

// SM count of the CURRENT device (cached per device).
inline int device_sm_count() {
  static std::atomic<int> cache[kMaxDevices];
  const int dev = current_device();
  int n = cache[dev].load(std::memory_order_relaxed);
  if (n <= 0) {
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cache[dev].store(n, std::memory_order_relaxed);
  }
  return n;
}

// ---------------------------------------------------------------------------------------------
// Device helpers
// ---------------------------------------------------------------------------------------------
#ifdef __CUDACC__

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// 16-byte streaming global load that bypasses L1 allocation (works for local and peer-mapped
// addresses alike; for peer addresses the request goes out over NVLink).
__device__ __forceinline__ uint4 ld_stream_16(const void* p) {
  uint4 r;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream_16(void* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

// NVLS: loads/stores through a multicast address (csrc/nvls.cu). The load returns the sum of all GPUs'
// copies of 8 bf16 (reduced in the NVSwitch, fp32 accumulation), the store writes all copies.
__device__ __forceinline__ uint4 multimem_ld_reduce_bf16x8(const void* mc_ptr) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(mc_ptr)
               : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_16(void* mc_ptr, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc_ptr), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w)
               : "memory");
}

// System-scope release/acquire on 32-bit flags living in (possibly peer-mapped) global memory.
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Flags only ever grow (they carry the step number), so "reached" is a wrap-safe signed compare.
__device__ __forceinline__ bool epoch_reached(uint32_t seen, uint32_t want) {
  return static_cast<int32_t>(seen - want) >= 0;
}

// Spin until *flag reaches `want`. `budget` bounds the spin (in polls) so that a lost peer turns
// into a reported timeout instead of a hung GPU; 0 means unbounded.
__device__ __forceinline__ bool wait_epoch(const uint32_t* flag, uint32_t want,
                                           unsigned long long budget) {
  unsigned long long polls = 0;
  while (!epoch_reached(ld_acquire_sys(flag), want)) {
    __nanosleep(64);
    if (budget && ++polls > budget) return false;
  }
  return true;
}

__device__ __forceinline__ float bf16_lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}

#endif  // __CUDACC__

}  // namespace tb200
