// Runtime plumbing for the native library: error strings, launch accounting, and the
// symmetric-memory layer (per-rank device allocations exported to the other ranks of the box
// as CUDA IPC handles so kernels can dereference peer pointers over NVLink 5 / NVSwitch).
//
// No counterpart in the reference: it only models the network with three numbers
// (bandwidth 1250 MB/s, 15 ms per crossed node, x2 per iteration; run_sim.py:59-69,
// core/network/network_service.py:30-37). This file is what makes the modelled
// worker<->parameter-server traffic executable on one 8xB200 NVSwitch domain.
#include "common.cuh"
#include "../include/tiresias_b200.h"

namespace tb200 {
thread_local char g_last_error[512] = "";
std::atomic<unsigned long long> g_launch_count{0};
}  // namespace tb200

using namespace tb200;

extern "C" {

int tb200_version(void) { return 100; }

const char* tb200_last_error(void) { return g_last_error; }

unsigned long long tb200_launch_count(void) { return g_launch_count.load(); }

void tb200_launch_count_reset(void) { g_launch_count.store(0); }

// ---- symmetric memory -------------------------------------------------------------------------

int tb200_symm_alloc(size_t bytes, void** out) {
  if (!out || bytes == 0) return fail("tb200_symm_alloc: bad arguments%s");
  // IPC-exported allocations must come straight from cudaMalloc (not from a pooling
  // allocator), and we round up so 16-byte vector access at the tail is always in bounds.
  size_t rounded = (bytes + 511) & ~size_t(511);
  TB_CUDA_OK(cudaMalloc(out, rounded));
  TB_CUDA_OK(cudaMemset(*out, 0, rounded));
  return 0;
}

int tb200_symm_free(void* ptr) {
  if (ptr) TB_CUDA_OK(cudaFree(ptr));
  return 0;
}

int tb200_ipc_export(void* ptr, unsigned char handle[TB200_IPC_HANDLE_BYTES]) {
  static_assert(sizeof(cudaIpcMemHandle_t) == TB200_IPC_HANDLE_BYTES, "handle size");
  cudaIpcMemHandle_t h;
  TB_CUDA_OK(cudaIpcGetMemHandle(&h, ptr));
  memcpy(handle, &h, sizeof(h));
  return 0;
}

int tb200_ipc_open(const unsigned char handle[TB200_IPC_HANDLE_BYTES], void** out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  TB_CUDA_OK(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}

int tb200_ipc_close(void* ptr) {
  if (ptr) TB_CUDA_OK(cudaIpcCloseMemHandle(ptr));
  return 0;
}

// Same-process multi-device worlds (tests, single-process executors) use plain peer access.
int tb200_enable_peer_access(int device, int peer) {
  if (device == peer) return 0;
  int can = 0;
  TB_CUDA_OK(cudaDeviceCanAccessPeer(&can, device, peer));
  if (!can) return fail("device %s%ld cannot access peer %ld", "", device, peer);
  int prev = 0;
  TB_CUDA_OK(cudaGetDevice(&prev));
  TB_CUDA_OK(cudaSetDevice(device));
  cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    e = cudaSuccess;
  }
  cudaSetDevice(prev);
  TB_CUDA_OK(e);
  return 0;
}

// Host-side write of a run of flags (used to (re)initialise pads, by single-GPU tests that
// stand in for absent peers, and by the executor's watchdog when it resets a failed job).
int tb200_flags_fill(uint32_t* flags, size_t first, size_t count, uint32_t value,
                     cudaStream_t stream) {
  if (value == 0) {
    TB_CUDA_OK(cudaMemsetAsync(flags + first, 0, count * sizeof(uint32_t), stream));
    return 0;
  }
  // cudaMemsetAsync is byte-wise; use the 32-bit driver-level pattern fill via cudaMemset2D
  // equivalent: a tiny staging copy keeps this dependency-free.
  uint32_t tmp[256];
  if (count > 256) return fail("tb200_flags_fill: count %s%ld > 256", "", (long)count);
  for (size_t i = 0; i < count; ++i) tmp[i] = value;
  TB_CUDA_OK(cudaMemcpyAsync(flags + first, tmp, count * sizeof(uint32_t),
                             cudaMemcpyHostToDevice, stream));
  TB_CUDA_OK(cudaStreamSynchronize(stream));
  return 0;
}

}  // extern "C"
