// Host-side helpers shared by all kernels of libdtf_b200_cuda: error reporting through a
// thread-local string (the C ABI returns int status codes), TMA descriptor encoding through the
// driver entry point (no link-time dependency on libcuda, so the library loads on GPU-less boxes).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

namespace dtf {

void set_error(const char* fmt, ...);  // defined in runtime/errors.cu

#define DTF_CHECK_CUDA(expr)                                                              \
  do {                                                                                    \
    cudaError_t _e = (expr);                                                              \
    if (_e != cudaSuccess) {                                                              \
      ::dtf::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__,  \
                       __LINE__);                                                         \
      return -1;                                                                          \
    }                                                                                     \
  } while (0)

#define DTF_REQUIRE(cond, ...)                                                            \
  do {                                                                                    \
    if (!(cond)) {                                                                        \
      ::dtf::set_error(__VA_ARGS__);                                                      \
      return -2;                                                                          \
    }                                                                                     \
  } while (0)

#define DTF_LAUNCH_CHECK()                                                                \
  do {                                                                                    \
    cudaError_t _e = cudaGetLastError();                                                  \
    if (_e != cudaSuccess) {                                                              \
      ::dtf::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e),        \
                       __FILE__, __LINE__);                                               \
      return -1;                                                                          \
    }                                                                                     \
  } while (0)

// Encode a tiled TMA descriptor over a bf16 tensor of rank 2..4 (dims/strides inner -> outer).
// strides_bytes has rank-1 entries (stride of dim 1.., dim 0 is contiguous).
// swizzle_bytes in {0, 32, 64, 128}. Returns 0 on success.
int encode_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                     const uint64_t* strides_bytes, const uint32_t* box, int swizzle_bytes);

// Launch with (optionally) the programmatic-dependent-launch attribute, see ptx.cuh. The switch is
// process-wide (dtf_set_pdl); kernels call griddep_wait() before touching global memory.
extern int g_pdl_enabled;

template <typename... KArgs, typename... Args>
static inline cudaError_t launch_kernel(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                        cudaStream_t stream, Args... args) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = g_pdl_enabled ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
static inline int round_up(int a, int b) { return ceil_div(a, b) * b; }

}  // namespace dtf
