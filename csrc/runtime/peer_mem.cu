// Peer-memory runtime: the data plane that replaces tf.train.Server's gRPC RecvTensor path
// (run_mnist_distributed.py:96-98). A PS task allocates its shard arena with cudaMalloc, exports
// a 64-byte CUDA IPC handle through the TCP control plane, and every worker maps it into its own
// address space; from then on parameters and gradients move by loads/stores/atomics issued from
// inside kernels over NVLink 5 / NVSwitch. Same-process shards (one process driving several
// GPUs, or PS and worker colocated on one GPU) use the pointer directly after
// cudaDeviceEnablePeerAccess.
#include <chrono>

#include "../kernels/common.cuh"
#include "../../include/dtf_b200.h"

using namespace dtf;

extern "C" int dtf_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

extern "C" int dtf_set_device(int device) {
  DTF_CHECK_CUDA(cudaSetDevice(device));
  return 0;
}

extern "C" int dtf_arena_alloc(void** ptr_out, size_t bytes) {
  DTF_REQUIRE(ptr_out != nullptr && bytes > 0, "arena_alloc: bad arguments");
  void* p = nullptr;
  DTF_CHECK_CUDA(cudaMalloc(&p, bytes));
  // Zero it on a private stream and wait for THAT stream only: a device-wide synchronise here would
  // wait for every kernel on the GPU, including a colocated worker's step that is parked in its
  // device-side flow control / token wait until the very task that is allocating has joined.
  cudaStream_t st = nullptr;
  cudaError_t e = cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaMemsetAsync(p, 0, bytes, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (st != nullptr) cudaStreamDestroy(st);
  if (e != cudaSuccess) {
    cudaFree(p);
    DTF_CHECK_CUDA(e);
  }
  *ptr_out = p;
  return 0;
}

extern "C" int dtf_arena_free(void* ptr) {
  if (ptr == nullptr) return 0;
  DTF_CHECK_CUDA(cudaFree(ptr));
  return 0;
}

extern "C" int dtf_ipc_export(void* ptr, unsigned char* handle_out) {
  static_assert(sizeof(cudaIpcMemHandle_t) == DTF_IPC_HANDLE_BYTES, "IPC handle size");
  cudaIpcMemHandle_t h;
  DTF_CHECK_CUDA(cudaIpcGetMemHandle(&h, ptr));
  memcpy(handle_out, &h, sizeof(h));
  return 0;
}

extern "C" int dtf_ipc_import(const unsigned char* handle, void** ptr_out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  void* p = nullptr;
  DTF_CHECK_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *ptr_out = p;
  return 0;
}

extern "C" int dtf_ipc_close(void* ptr) {
  if (ptr == nullptr) return 0;
  DTF_CHECK_CUDA(cudaIpcCloseMemHandle(ptr));
  return 0;
}

extern "C" int dtf_can_access_peer(int device, int peer_device) {
  int ok = 0;
  if (cudaDeviceCanAccessPeer(&ok, device, peer_device) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return ok;
}

extern "C" int dtf_enable_peer_access(int peer_device) {
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    return 0;
  }
  DTF_CHECK_CUDA(e);
  return 0;
}

extern "C" int dtf_memcpy_d2h(void* dst_host, const void* src_dev, size_t bytes, void* stream) {
  DTF_CHECK_CUDA(cudaMemcpyAsync(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost,
                                 reinterpret_cast<cudaStream_t>(stream)));
  return 0;
}

extern "C" int dtf_memcpy_h2d(void* dst_dev, const void* src_host, size_t bytes, void* stream) {
  DTF_CHECK_CUDA(cudaMemcpyAsync(dst_dev, src_host, bytes, cudaMemcpyHostToDevice,
                                 reinterpret_cast<cudaStream_t>(stream)));
  return 0;
}

extern "C" int dtf_stream_sync(void* stream) {
  DTF_CHECK_CUDA(cudaStreamSynchronize(reinterpret_cast<cudaStream_t>(stream)));
  return 0;
}

extern "C" int dtf_host_alloc_pinned(void** ptr_out, size_t bytes) {
  DTF_CHECK_CUDA(cudaHostAlloc(ptr_out, bytes, cudaHostAllocDefault));
  return 0;
}

extern "C" int dtf_host_free_pinned(void* ptr) {
  if (ptr == nullptr) return 0;
  DTF_CHECK_CUDA(cudaFreeHost(ptr));
  return 0;
}

// Launch an instantiated CUDA graph (cudaGraphExec_t obtained from the framework that captured it)
// on an explicit stream: the per-step hot path of the executor, one call, no Python-side stream
// context juggling.
extern "C" int dtf_graph_launch(void* graph_exec, void* stream) {
  DTF_REQUIRE(graph_exec != nullptr, "graph_launch: null graph");
  DTF_CHECK_CUDA(cudaGraphLaunch(reinterpret_cast<cudaGraphExec_t>(graph_exec),
                                 reinterpret_cast<cudaStream_t>(stream)));
  return 0;
}

// Spin (on the calling host thread, GIL released by ctypes) until a word of pinned host memory
// that a kernel writes equals `expected`. Returns 0 when it does, 1 on timeout.
extern "C" int dtf_host_wait_u32(const unsigned int* word, unsigned int expected, long long timeout_us) {
  const volatile unsigned int* p = word;
  if (*p == expected) return 0;
  const auto t0 = std::chrono::steady_clock::now();
  for (unsigned long long spins = 0;; ++spins) {
    if (*p == expected) return 0;
#if defined(__x86_64__)
    __builtin_ia32_pause();
#endif
    if ((spins & 0xfff) == 0xfff) {
      const auto dt = std::chrono::duration_cast<std::chrono::microseconds>(
                          std::chrono::steady_clock::now() - t0).count();
      if (dt > timeout_us) return 1;
    }
  }
}
