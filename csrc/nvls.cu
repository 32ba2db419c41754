// NVLS (NVLink SHARP) multicast objects: in-switch reduction and broadcast from inside a kernel.
//
// A multicast object binds one physical allocation per GPU of the NVSwitch domain to a single virtual
// address. `multimem.ld_reduce` on that address returns the SUM of all GPUs' copies (the switch reduces
// in flight, so a GPU receives n/W bytes for its shard instead of (W-1) * n/W), `multimem.st` writes every
// copy with one store. This is the NVLS half of K1/K2/K3: the reference only prices the
// push/pull as model_size / bandwidth * 2 (core/network/network_service.py:30-37).
//
// This file holds the driver-level plumbing (cuMulticast*, cuMem* virtual memory management; entry points
// fetched at run time so the library does not link libcuda) and a single-process self test over the
// first `ndev` GPUs that measures the three primitives the trainers need:
//   reduce      : one GPU sums everybody's buffer         (push onto a parameter-server shard)
//   broadcast   : one GPU stores into everybody's buffer  (pull of an updated shard)
//   all-reduce  : every GPU reduces its 1/W shard and multicasts the result (one-shot NVLS all-reduce)
#include "common.cuh"
#include "../include/tiresias_b200.h"

#include <cuda.h>
#include <cstring>
#include <functional>
#include <string>
#include <vector>

using namespace tb200;

namespace {

// ---- driver entry points ------------------------------------------------------------------------
#define TB_DRV_LIST(X)                                                                                   \
  X(cuDeviceGet) X(cuDeviceGetAttribute) X(cuMulticastCreate) X(cuMulticastAddDevice) X(cuMulticastBindMem) \
  X(cuMulticastGetGranularity) X(cuMulticastUnbind) X(cuMemCreate) X(cuMemRelease) X(cuMemAddressReserve)  \
  X(cuMemAddressFree) X(cuMemMap) X(cuMemUnmap) X(cuMemSetAccess) X(cuMemGetAllocationGranularity)         \
  X(cuMemExportToShareableHandle) X(cuMemImportFromShareableHandle)

struct Drv {
#define X(name) decltype(&name) name##_ = nullptr;
  TB_DRV_LIST(X)
#undef X
  bool ok = false;
  std::string missing;
};

const Drv& drv() {
  static Drv d;
  static bool loaded = false;
  if (!loaded) {
    loaded = true;
    d.ok = true;
    cudaDriverEntryPointQueryResult q;
    void* p = nullptr;
#define X(name)                                                                                         \
  p = nullptr;                                                                                          \
  if (cudaGetDriverEntryPoint(#name, &p, cudaEnableDefault, &q) != cudaSuccess ||                       \
      q != cudaDriverEntryPointSuccess || !p) {                                                         \
    d.ok = false;                                                                                       \
    d.missing += std::string(" ") + #name;                                                              \
  }                                                                                                     \
  d.name##_ = reinterpret_cast<decltype(&name)>(p);
    TB_DRV_LIST(X)
#undef X
  }
  return d;
}

// ---- kernels --------------------------------------------------------------------------------------
__global__ void __launch_bounds__(512) nvls_fill_kernel(uint4* dst, size_t n16, float value) {
  const uint32_t w = pack_bf16(value, value);
  const uint4 v = make_uint4(w, w, w, w);
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x)
    dst[i] = v;
}

// out[i] = sum over GPUs of their copy of element i (read through the multicast address)
__global__ void __launch_bounds__(512) nvls_reduce_kernel(const uint4* mc_src, uint4* out, size_t n16) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x)
    out[i] = multimem_ld_reduce_bf16x8(mc_src + i);
}

// every GPU's copy of element i = src[i] (written through the multicast address)
__global__ void __launch_bounds__(512) nvls_bcast_kernel(uint4* mc_dst, const uint4* src, size_t n16) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x)
    multimem_st_16(mc_dst + i, src[i]);
}

// one-shot all-reduce of this GPU's shard: reduce in the switch, multicast the sum to every GPU's output
__global__ void __launch_bounds__(512) nvls_allreduce_shard_kernel(const uint4* mc_in, uint4* mc_out, size_t first,
                                                                   size_t n16) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x)
    multimem_st_16(mc_out + first + i, multimem_ld_reduce_bf16x8(mc_in + first + i));
}

// ---- one multicast-backed buffer replicated on `ndev` GPUs of this process ---------------------------
struct McBuffer {
  int ndev = 0;
  size_t bytes = 0;
  CUmemGenericAllocationHandle mc = 0;
  std::vector<CUmemGenericAllocationHandle> phys;
  std::vector<CUdeviceptr> uc;       // per-GPU unicast view of that GPU's copy
  CUdeviceptr mc_va = 0;             // the multicast view
  bool mc_mapped = false;
};

#define TB_DRV_STEP(call, what)                                                                  \
  do {                                                                                           \
    CUresult r_ = (call);                                                                        \
    if (r_ != CUDA_SUCCESS) {                                                                    \
      snprintf(g_last_error, sizeof(g_last_error), "nvls: %s failed (CUresult %d)", what, (int)r_); \
      return 1;                                                                                  \
    }                                                                                            \
  } while (0)

int mc_buffer_create(McBuffer* b, int ndev, size_t want_bytes) {
  const Drv& d = drv();
  b->ndev = ndev;
  CUmulticastObjectProp prop{};
  prop.numDevices = (unsigned)ndev;
  prop.handleTypes = 0;        // single process: nothing to export
  prop.flags = 0;
  prop.size = want_bytes;
  size_t gran = 0;
  // (the RECOMMENDED granularity is 512 MiB on B200: fine for one big arena, not for test buffers)
  TB_DRV_STEP(d.cuMulticastGetGranularity_(&gran, &prop, CU_MULTICAST_GRANULARITY_MINIMUM),
              "cuMulticastGetGranularity");
  b->bytes = (want_bytes + gran - 1) / gran * gran;
  prop.size = b->bytes;
  TB_DRV_STEP(d.cuMulticastCreate_(&b->mc, &prop), "cuMulticastCreate");
  for (int i = 0; i < ndev; ++i) {
    CUdevice dev;
    TB_DRV_STEP(d.cuDeviceGet_(&dev, i), "cuDeviceGet");
    TB_DRV_STEP(d.cuMulticastAddDevice_(b->mc, dev), "cuMulticastAddDevice");
  }
  std::vector<CUmemAccessDesc> access(ndev);
  for (int i = 0; i < ndev; ++i) {
    access[i].location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    access[i].location.id = i;
    access[i].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  }
  b->phys.assign(ndev, 0);
  b->uc.assign(ndev, 0);
  for (int i = 0; i < ndev; ++i) {
    TB_CUDA_OK(cudaSetDevice(i));
    CUmemAllocationProp ap{};
    ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    ap.location.id = i;
    ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_NONE;
    TB_DRV_STEP(d.cuMemCreate_(&b->phys[i], b->bytes, &ap, 0), "cuMemCreate");
    TB_DRV_STEP(d.cuMulticastBindMem_(b->mc, 0, b->phys[i], 0, b->bytes, 0), "cuMulticastBindMem");
    TB_DRV_STEP(d.cuMemAddressReserve_(&b->uc[i], b->bytes, gran, 0, 0), "cuMemAddressReserve (unicast)");
    TB_DRV_STEP(d.cuMemMap_(b->uc[i], b->bytes, 0, b->phys[i], 0), "cuMemMap (unicast)");
    TB_DRV_STEP(d.cuMemSetAccess_(b->uc[i], b->bytes, access.data(), (size_t)ndev), "cuMemSetAccess (unicast)");
  }
  TB_DRV_STEP(d.cuMemAddressReserve_(&b->mc_va, b->bytes, gran, 0, 0), "cuMemAddressReserve (multicast)");
  TB_DRV_STEP(d.cuMemMap_(b->mc_va, b->bytes, 0, b->mc, 0), "cuMemMap (multicast)");
  b->mc_mapped = true;
  TB_DRV_STEP(d.cuMemSetAccess_(b->mc_va, b->bytes, access.data(), (size_t)ndev), "cuMemSetAccess (multicast)");
  return 0;
}

void mc_buffer_destroy(McBuffer* b) {
  const Drv& d = drv();
  if (!d.ok) return;
  if (b->mc_va) {
    if (b->mc_mapped) d.cuMemUnmap_(b->mc_va, b->bytes);
    d.cuMemAddressFree_(b->mc_va, b->bytes);
  }
  for (int i = 0; i < (int)b->uc.size(); ++i) {
    if (b->uc[i]) {
      d.cuMemUnmap_(b->uc[i], b->bytes);
      d.cuMemAddressFree_(b->uc[i], b->bytes);
    }
    if (b->phys[i]) {
      CUdevice dev;
      if (b->mc && d.cuDeviceGet_(&dev, i) == CUDA_SUCCESS) d.cuMulticastUnbind_(b->mc, dev, 0, b->bytes);
      d.cuMemRelease_(b->phys[i]);
    }
  }
  if (b->mc) d.cuMemRelease_(b->mc);
  *b = McBuffer{};
}

float time_on(int dev, cudaStream_t s, int iters, const std::function<void()>& fn) {
  cudaSetDevice(dev);
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  for (int i = 0; i < 3; ++i) fn();
  cudaStreamSynchronize(s);
  cudaEventRecord(a, s);
  for (int i = 0; i < iters; ++i) fn();
  cudaEventRecord(b, s);
  cudaEventSynchronize(b);
  float ms = 0.f;
  cudaEventElapsedTime(&ms, a, b);
  cudaEventDestroy(a);
  cudaEventDestroy(b);
  return ms / iters;
}

}  // namespace

// ---- multi-process symmetric memory on the virtual-memory-management API ---------------------------------
// cudaMalloc'ed memory cannot be bound to a multicast object, so buffers that want an NVLS view are
// physical allocations (cuMemCreate) shared between the ranks' processes as POSIX file descriptors
// (passed over a unix socket by the Python layer) instead of CUDA IPC handles.
namespace {

int current_device(CUdevice* dev, int* ordinal) {
  const Drv& d = drv();
  if (!d.ok) {
    snprintf(g_last_error, sizeof(g_last_error), "vmm: driver entry points missing:%s", d.missing.c_str());
    return 2;
  }
  TB_CUDA_OK(cudaFree(nullptr));            // binds the primary context to this thread
  TB_CUDA_OK(cudaGetDevice(ordinal));
  TB_DRV_STEP(d.cuDeviceGet_(dev, *ordinal), "cuDeviceGet");
  return 0;
}

CUmemAllocationProp phys_prop(int ordinal) {
  CUmemAllocationProp ap{};
  ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  ap.location.id = ordinal;
  ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return ap;
}

int map_rw(CUmemGenericAllocationHandle h, size_t bytes, int ordinal, CUdeviceptr* out, const char* what) {
  const Drv& d = drv();
  CUdeviceptr va = 0;
  TB_DRV_STEP(d.cuMemAddressReserve_(&va, bytes, 0, 0, 0), what);
  CUresult r = d.cuMemMap_(va, bytes, 0, h, 0);
  if (r == CUDA_SUCCESS) {
    CUmemAccessDesc acc{};
    acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc.location.id = ordinal;
    acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    r = d.cuMemSetAccess_(va, bytes, &acc, 1);
    if (r != CUDA_SUCCESS) d.cuMemUnmap_(va, bytes);
  }
  if (r != CUDA_SUCCESS) {
    d.cuMemAddressFree_(va, bytes);
    snprintf(g_last_error, sizeof(g_last_error), "vmm: %s: map failed (CUresult %d)", what, (int)r);
    return 1;
  }
  *out = va;
  return 0;
}

}  // namespace

extern "C" {

// Size every rank must use for a symmetric VMM buffer of `bytes` shared by `world` GPUs: a multiple of
// both the allocation granularity and (when the GPU supports it) the multicast granularity.
int tb200_vmm_round(int world, size_t bytes, size_t* rounded) {
  CUdevice dev;
  int ord = 0;
  if (int rc = current_device(&dev, &ord)) return rc;
  const Drv& d = drv();
  CUmemAllocationProp ap = phys_prop(ord);
  size_t gran = 0;
  TB_DRV_STEP(d.cuMemGetAllocationGranularity_(&gran, &ap, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED),
              "cuMemGetAllocationGranularity");
  if (world > 1 && tb200_nvls_supported(ord) == 1) {
    CUmulticastObjectProp prop{};
    prop.numDevices = (unsigned)world;
    prop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    prop.size = bytes;
    size_t mg = 0;
    TB_DRV_STEP(d.cuMulticastGetGranularity_(&mg, &prop, CU_MULTICAST_GRANULARITY_MINIMUM),
                "cuMulticastGetGranularity");
    if (mg > gran) gran = mg;
  }
  if (gran == 0) gran = 2u << 20;
  *rounded = (bytes + gran - 1) / gran * gran;
  return 0;
}

// This rank's copy: physical memory on the current device, mapped read/write, zero-filled, plus a file
// descriptor the other ranks import (close it once they have).
int tb200_vmm_alloc(size_t rounded, unsigned long long* phys, void** ptr, int* fd) {
  CUdevice dev;
  int ord = 0;
  if (int rc = current_device(&dev, &ord)) return rc;
  const Drv& d = drv();
  CUmemAllocationProp ap = phys_prop(ord);
  CUmemGenericAllocationHandle h = 0;
  TB_DRV_STEP(d.cuMemCreate_(&h, rounded, &ap, 0), "cuMemCreate");
  CUdeviceptr va = 0;
  if (int rc = map_rw(h, rounded, ord, &va, "local copy")) {
    d.cuMemRelease_(h);
    return rc;
  }
  int out_fd = -1;
  CUresult r = d.cuMemExportToShareableHandle_(&out_fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) {
    d.cuMemUnmap_(va, rounded);
    d.cuMemAddressFree_(va, rounded);
    d.cuMemRelease_(h);
    snprintf(g_last_error, sizeof(g_last_error), "vmm: cuMemExportToShareableHandle failed (CUresult %d)", (int)r);
    return 1;
  }
  TB_CUDA_OK(cudaMemset(reinterpret_cast<void*>(va), 0, rounded));
  *phys = h;
  *ptr = reinterpret_cast<void*>(va);
  *fd = out_fd;
  return 0;
}

// A peer's copy, from the file descriptor it exported: mapped read/write on the current device.
int tb200_vmm_import(int fd, size_t rounded, unsigned long long* phys, void** ptr) {
  CUdevice dev;
  int ord = 0;
  if (int rc = current_device(&dev, &ord)) return rc;
  const Drv& d = drv();
  CUmemGenericAllocationHandle h = 0;
  TB_DRV_STEP(d.cuMemImportFromShareableHandle_(&h, reinterpret_cast<void*>((uintptr_t)fd),
                                                CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR),
              "cuMemImportFromShareableHandle");
  CUdeviceptr va = 0;
  if (int rc = map_rw(h, rounded, ord, &va, "peer copy")) {
    d.cuMemRelease_(h);
    return rc;
  }
  *phys = h;
  *ptr = reinterpret_cast<void*>(va);
  return 0;
}

int tb200_vmm_release(unsigned long long phys, void* ptr, size_t rounded) {
  const Drv& d = drv();
  if (!d.ok) return 0;
  if (ptr) {
    d.cuMemUnmap_(reinterpret_cast<CUdeviceptr>(ptr), rounded);
    d.cuMemAddressFree_(reinterpret_cast<CUdeviceptr>(ptr), rounded);
  }
  if (phys) d.cuMemRelease_(phys);
  return 0;
}

// Rank 0 creates the multicast object for `world` GPUs and gets a file descriptor for the others.
int tb200_mc_create(int world, size_t rounded, unsigned long long* mc, int* fd) {
  CUdevice dev;
  int ord = 0;
  if (int rc = current_device(&dev, &ord)) return rc;
  const Drv& d = drv();
  CUmulticastObjectProp prop{};
  prop.numDevices = (unsigned)world;
  prop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  prop.size = rounded;
  CUmemGenericAllocationHandle h = 0;
  TB_DRV_STEP(d.cuMulticastCreate_(&h, &prop), "cuMulticastCreate");
  int out_fd = -1;
  CUresult r = d.cuMemExportToShareableHandle_(&out_fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) {
    d.cuMemRelease_(h);
    snprintf(g_last_error, sizeof(g_last_error), "nvls: export of the multicast object failed (CUresult %d)", (int)r);
    return 1;
  }
  *mc = h;
  *fd = out_fd;
  return 0;
}

int tb200_mc_import(int fd, unsigned long long* mc) {
  CUdevice dev;
  int ord = 0;
  if (int rc = current_device(&dev, &ord)) return rc;
  const Drv& d = drv();
  CUmemGenericAllocationHandle h = 0;
  TB_DRV_STEP(d.cuMemImportFromShareableHandle_(&h, reinterpret_cast<void*>((uintptr_t)fd),
                                                CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR),
              "cuMemImportFromShareableHandle (multicast object)");
  *mc = h;
  return 0;
}

// Every rank adds its (current) device; all ranks must have done so before anyone binds memory.
int tb200_mc_add_device(unsigned long long mc) {
  CUdevice dev;
  int ord = 0;
  if (int rc = current_device(&dev, &ord)) return rc;
  TB_DRV_STEP(drv().cuMulticastAddDevice_(mc, dev), "cuMulticastAddDevice");
  return 0;
}

// Bind this rank's copy into the object and map the multicast view on the current device.
int tb200_mc_bind(unsigned long long mc, unsigned long long phys, size_t rounded, void** mc_ptr) {
  CUdevice dev;
  int ord = 0;
  if (int rc = current_device(&dev, &ord)) return rc;
  const Drv& d = drv();
  TB_DRV_STEP(d.cuMulticastBindMem_(mc, 0, phys, 0, rounded, 0), "cuMulticastBindMem");
  CUdeviceptr va = 0;
  if (int rc = map_rw(mc, rounded, ord, &va, "multicast view")) return rc;
  *mc_ptr = reinterpret_cast<void*>(va);
  return 0;
}

int tb200_mc_release(unsigned long long mc, void* mc_ptr, size_t rounded) {
  const Drv& d = drv();
  if (!d.ok) return 0;
  if (mc_ptr) {
    d.cuMemUnmap_(reinterpret_cast<CUdeviceptr>(mc_ptr), rounded);
    d.cuMemAddressFree_(reinterpret_cast<CUdeviceptr>(mc_ptr), rounded);
  }
  if (mc) {
    CUdevice dev;
    int ord = 0;
    if (cudaGetDevice(&ord) == cudaSuccess && d.cuDeviceGet_(&dev, ord) == CUDA_SUCCESS)
      d.cuMulticastUnbind_(mc, dev, 0, rounded);
    d.cuMemRelease_(mc);
  }
  return 0;
}

// 1 when `device` can join a multicast object (NVSwitch fabric with NVLS), 0 when not, <0 on error.
int tb200_nvls_supported(int device) {
  const Drv& d = drv();
  if (!d.ok) return 0;
  static thread_local bool ctx = false;
  if (!ctx) {
    cudaFree(nullptr);
    ctx = true;
  }
  CUdevice dev;
  if (d.cuDeviceGet_(&dev, device) != CUDA_SUCCESS) return -1;
  int v = 0;
  if (d.cuDeviceGetAttribute_(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) != CUDA_SUCCESS) return -1;
  return v ? 1 : 0;
}

// Single-process self test over GPUs 0..ndev-1: checks the three primitives for correctness and times
// them (CUDA events on the launching stream; the all-reduce is the max over GPUs). `report` receives one
// JSON object. Returns 0 when everything checked out, 2 when NVLS is not available, 1 on an error
// (tb200_last_error()).
int tb200_nvls_selftest(int ndev, size_t bytes, int ctas, char* report, size_t report_len) {
  if (report && report_len) report[0] = 0;
  if (ndev < 2 || ndev > TB200_MAX_WORLD) return fail("tb200_nvls_selftest: ndev %s(%ld) must be 2..8", "", ndev);
  const Drv& d = drv();
  if (!d.ok) {
    snprintf(g_last_error, sizeof(g_last_error), "nvls: driver entry points missing:%s", d.missing.c_str());
    return 2;
  }
  int have = 0;
  TB_CUDA_OK(cudaGetDeviceCount(&have));
  if (have < ndev) return fail("tb200_nvls_selftest: %s%ld GPUs visible, %ld wanted", "", have, ndev);
  for (int i = 0; i < ndev; ++i) {
    TB_CUDA_OK(cudaSetDevice(i));
    TB_CUDA_OK(cudaFree(nullptr));
    if (tb200_nvls_supported(i) != 1) {
      snprintf(g_last_error, sizeof(g_last_error), "nvls: GPU %d does not support multicast objects", i);
      return 2;
    }
    for (int j = 0; j < ndev; ++j)
      if (int rc = tb200_enable_peer_access(i, j)) return rc;
  }
  if (ctas <= 0) ctas = 148;
  McBuffer in, out;
  int rc = mc_buffer_create(&in, ndev, bytes);
  if (!rc) rc = mc_buffer_create(&out, ndev, bytes);
  if (rc) {
    mc_buffer_destroy(&in);
    mc_buffer_destroy(&out);
    return rc;
  }
  const size_t n16 = in.bytes / 16;
  std::vector<cudaStream_t> st(ndev);
  std::vector<uint4*> local(ndev);
  for (int i = 0; i < ndev; ++i) {
    cudaSetDevice(i);
    cudaStreamCreate(&st[i]);
    cudaMalloc(&local[i], in.bytes);
    nvls_fill_kernel<<<ctas, 512, 0, st[i]>>>(reinterpret_cast<uint4*>(in.uc[i]), n16, (float)(i + 1));
    nvls_fill_kernel<<<ctas, 512, 0, st[i]>>>(reinterpret_cast<uint4*>(out.uc[i]), n16, 0.f);
    nvls_fill_kernel<<<ctas, 512, 0, st[i]>>>(local[i], n16, 0.f);
  }
  for (int i = 0; i < ndev; ++i) {
    cudaSetDevice(i);
    cudaStreamSynchronize(st[i]);
  }
  auto first_word = [&](const void* p, int dev) {
    uint32_t w = 0;
    cudaSetDevice(dev);
    cudaMemcpy(&w, p, 4, cudaMemcpyDeviceToHost);
    const uint32_t bits = w << 16;           // the buffer's first bf16 element, widened
    float f;
    memcpy(&f, &bits, 4);
    return f;
  };
  const float want_sum = ndev * (ndev + 1) / 2.0f;
  bool ok = true;
  // 1. reduce on GPU 0
  cudaSetDevice(0);
  const uint4* mc_in = reinterpret_cast<const uint4*>(in.mc_va);
  uint4* mc_out = reinterpret_cast<uint4*>(out.mc_va);
  const float reduce_ms = time_on(0, st[0], 10, [&] { nvls_reduce_kernel<<<ctas, 512, 0, st[0]>>>(mc_in, local[0], n16); });
  const float got_first = first_word(local[0], 0);
  const float got_last = first_word(reinterpret_cast<const uint8_t*>(local[0]) + in.bytes - 4, 0);
  ok = ok && got_first == want_sum && got_last == want_sum;
  // 2. broadcast from GPU 0 (its `local` now holds the sums)
  const float bcast_ms = time_on(0, st[0], 10, [&] { nvls_bcast_kernel<<<ctas, 512, 0, st[0]>>>(mc_out, local[0], n16); });
  float seen_min = 1e30f, seen_max = -1e30f;
  for (int i = 0; i < ndev; ++i) {
    const float a = first_word(reinterpret_cast<const void*>(out.uc[i]), i);
    const float b = first_word(reinterpret_cast<const uint8_t*>(out.uc[i]) + in.bytes - 4, i);
    seen_min = fminf(seen_min, fminf(a, b));
    seen_max = fmaxf(seen_max, fmaxf(a, b));
  }
  ok = ok && seen_min == want_sum && seen_max == want_sum;
  // 3. one-shot all-reduce: every GPU its shard, all launched back to back, time = slowest GPU
  for (int i = 0; i < ndev; ++i) {
    cudaSetDevice(i);
    nvls_fill_kernel<<<ctas, 512, 0, st[i]>>>(reinterpret_cast<uint4*>(out.uc[i]), n16, 0.f);
    cudaStreamSynchronize(st[i]);
  }
  const size_t shard16 = n16 / ndev;
  const int iters = 10;
  std::vector<cudaEvent_t> ea(ndev), eb(ndev);
  for (int rep = 0; rep < 2; ++rep) {                      // rep 0 = warm-up
    for (int i = 0; i < ndev; ++i) {
      cudaSetDevice(i);
      if (rep == 1) {
        cudaEventCreate(&ea[i]);
        cudaEventCreate(&eb[i]);
        cudaEventRecord(ea[i], st[i]);
      }
      for (int it = 0; it < (rep ? iters : 3); ++it)
        nvls_allreduce_shard_kernel<<<ctas, 512, 0, st[i]>>>(mc_in, mc_out, (size_t)i * shard16, shard16);
      if (rep == 1) cudaEventRecord(eb[i], st[i]);
    }
    for (int i = 0; i < ndev; ++i) {
      cudaSetDevice(i);
      cudaStreamSynchronize(st[i]);
    }
  }
  float allreduce_ms = 0.f;
  for (int i = 0; i < ndev; ++i) {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ea[i], eb[i]);
    allreduce_ms = fmaxf(allreduce_ms, ms / iters);
    cudaEventDestroy(ea[i]);
    cudaEventDestroy(eb[i]);
  }
  float ar_min = 1e30f, ar_max = -1e30f;
  for (int i = 0; i < ndev; ++i)
    for (int s = 0; s < ndev; ++s) {
      const float a = first_word(reinterpret_cast<const uint8_t*>(out.uc[i]) + (size_t)s * shard16 * 16, i);
      ar_min = fminf(ar_min, a);
      ar_max = fmaxf(ar_max, a);
    }
  ok = ok && ar_min == want_sum && ar_max == want_sum;
  cudaError_t err = cudaSuccess;
  for (int i = 0; i < ndev; ++i) {
    cudaSetDevice(i);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) err = e;
    cudaFree(local[i]);
    cudaStreamDestroy(st[i]);
  }
  const double gb = in.bytes / 1e9;
  if (report && report_len)
    snprintf(report, report_len,
             "{\"ndev\": %d, \"bytes\": %zu, \"ctas\": %d, \"ok\": %s, \"want\": %.1f, \"reduce_first\": %.1f, "
             "\"bcast_min\": %.1f, \"bcast_max\": %.1f, \"allreduce_min\": %.1f, \"allreduce_max\": %.1f, "
             "\"reduce_ms\": %.4f, \"reduce_out_gbs\": %.1f, \"bcast_ms\": %.4f, \"bcast_src_gbs\": %.1f, "
             "\"allreduce_ms\": %.4f, \"allreduce_busbw_gbs\": %.1f, \"cuda_error\": \"%s\"}",
             ndev, in.bytes, ctas, ok ? "true" : "false", want_sum, got_first, seen_min, seen_max, ar_min, ar_max,
             reduce_ms, gb / (reduce_ms * 1e-3), bcast_ms, gb / (bcast_ms * 1e-3), allreduce_ms,
             2.0 * (ndev - 1) / ndev * gb / (allreduce_ms * 1e-3), cudaGetErrorString(err));
  mc_buffer_destroy(&in);
  mc_buffer_destroy(&out);
  cudaSetDevice(0);
  if (err != cudaSuccess) return fail("tb200_nvls_selftest: %s", cudaGetErrorString(err));
  if (!ok) return fail("tb200_nvls_selftest: wrong values%s");
  return 0;
}

}  // extern "C"
