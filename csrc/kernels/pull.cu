// K11 (bulk part): parameter pull, PS shard -> worker, for every tensor that is NOT consumed
// directly from peer memory by a fused GEMM. One launch copies a whole list of tensors with
// 16-byte loads over the NVLink mapping (peer reads bypass the local L2, so each byte crosses the
// link exactly once) and records the global_step observed at pull time, which is what the
// bounded-staleness option compares against. Reference: the per-tensor RecvTensor RPCs issued by
// every mon_sess.run (run_mnist_distributed.py:150,161).
#include "common.cuh"
#include "ptx.cuh"
#include "../../include/dtf_b200.h"

namespace dtf {

constexpr int PULL_THREADS = 256;
constexpr int PULL_UNROLL = 8;
constexpr long long PULL_CHUNK = static_cast<long long>(PULL_THREADS) * PULL_UNROLL * 16;  // 32 KB

struct PullArgs {
  dtf_copy_item items[DTF_MAX_COPY_ITEMS];
  int block_start[DTF_MAX_COPY_ITEMS + 1];
  int n_items;
  const dtf_ps_ctrl* ctrl_step;
  unsigned long long* pulled_step_out;
};

__global__ void __launch_bounds__(PULL_THREADS)
pull_bulk_kernel(const __grid_constant__ PullArgs a) {
  griddep_wait();
  griddep_launch();
  int it = 0;
  while (it + 1 < a.n_items && static_cast<int>(blockIdx.x) >= a.block_start[it + 1]) ++it;
  const dtf_copy_item& c = a.items[it];
  const long long base = static_cast<long long>(blockIdx.x - a.block_start[it]) * PULL_CHUNK;
  const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(c.src) + base);
  uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<char*>(c.dst) + base);
  const long long nvec = min(PULL_CHUNK, c.bytes - base) / 16;
  uint4 v[PULL_UNROLL];
  // all loads first (PULL_UNROLL 16-byte requests in flight per thread), then all stores
#pragma unroll
  for (int u = 0; u < PULL_UNROLL; ++u) {
    const long long i = static_cast<long long>(u) * PULL_THREADS + threadIdx.x;
    if (i < nvec) v[u] = ld_relaxed_u4(src + i);
  }
#pragma unroll
  for (int u = 0; u < PULL_UNROLL; ++u) {
    const long long i = static_cast<long long>(u) * PULL_THREADS + threadIdx.x;
    if (i < nvec) dst[i] = v[u];
  }
  if (blockIdx.x == 0 && threadIdx.x == 0 && a.pulled_step_out != nullptr &&
      a.ctrl_step != nullptr)
    *a.pulled_step_out = ld_acquire_sys_u64(&a.ctrl_step->global_step);
}

}  // namespace dtf

using namespace dtf;

extern "C" int dtf_pull_bulk(const dtf_copy_item* items, int n_items,
                             const dtf_ps_ctrl* ctrl_for_step,
                             unsigned long long* pulled_step_out, void* stream) {
  DTF_REQUIRE(n_items >= 0 && n_items <= DTF_MAX_COPY_ITEMS, "pull_bulk: %d items (max %d)",
              n_items, DTF_MAX_COPY_ITEMS);
  PullArgs a;
  memset(&a, 0, sizeof(a));
  int blocks = 0;
  for (int i = 0; i < n_items; ++i) {
    DTF_REQUIRE(items[i].bytes % 16 == 0, "pull_bulk: item %d has %lld bytes (not /16)", i,
                items[i].bytes);
    DTF_REQUIRE(((reinterpret_cast<uintptr_t>(items[i].src) |
                  reinterpret_cast<uintptr_t>(items[i].dst)) & 15) == 0,
                "pull_bulk: item %d is not 16-byte aligned", i);
    a.items[i] = items[i];
    a.block_start[i] = blocks;
    blocks += static_cast<int>((items[i].bytes + PULL_CHUNK - 1) / PULL_CHUNK);
  }
  a.block_start[n_items] = blocks;
  a.n_items = n_items > 0 ? n_items : 1;
  a.ctrl_step = ctrl_for_step;
  a.pulled_step_out = pulled_step_out;
  if (blocks == 0) {
    if (pulled_step_out == nullptr) return 0;
    blocks = 1;
    a.items[0].bytes = 0;
    a.block_start[1] = 1;
  }
  DTF_CHECK_CUDA(launch_kernel(pull_bulk_kernel, dim3(blocks), dim3(PULL_THREADS), 0,
                               reinterpret_cast<cudaStream_t>(stream), a));
  return 0;
}
