// K9/K10: gradient push fused with the owning shard's optimizer apply.
//
// The worker launches ONE kernel per step. For every (variable slice, gradient slice) pair it
// reads the local fp32 gradient, reads the slice's optimizer state from the owning PS shard
// (local HBM, or a peer GPU's HBM through the NVLink mapping), applies the TF1 update rule and
// writes var / slots back together with the bf16 serving copy that workers pull. Plain SGD can
// instead be pushed with red.global.add.f32 (write-only over NVLink, exact under concurrency).
// The last CTA to retire performs TF's "_finish": beta1_power/beta2_power *= beta, then bumps the
// shared global_step with ONE system-scope 64-bit atomic and publishes the worker's heartbeat.
// Updates are deliberately unlocked (use_locking=False, Hogwild), as in the reference's
// AdamOptimizer(...).minimize (run_mnist_distributed.py:116).
#include "common.cuh"
#include "ptx.cuh"
#include "apply_rules.cuh"
#include "../../include/dtf_b200.h"

namespace dtf {

constexpr int APPLY_THREADS = 256;
constexpr int APPLY_VEC = 4;
constexpr int APPLY_UNROLL = 4;
constexpr int APPLY_CHUNK = APPLY_THREADS * APPLY_VEC * APPLY_UNROLL;  // 4096 elements / CTA

struct ApplyArgs {
  dtf_apply_item items[DTF_MAX_APPLY_ITEMS];
  int block_start[DTF_MAX_APPLY_ITEMS + 1];
  int n_items;
  int total_blocks;
  dtf_apply_hyper h;
  dtf_ps_ctrl* ctrl_powers;
  dtf_ps_ctrl* ctrl_step;
  int finish;
  int worker_id;
  unsigned long long* step_out;
  unsigned int* block_counter;
  dtf_result_sink sink;
};

template <int OPT>
__global__ void __launch_bounds__(APPLY_THREADS)
push_apply_kernel(const __grid_constant__ ApplyArgs a) {
  __shared__ int s_item;
  __shared__ float s_lr_t;
  __shared__ unsigned int s_last;
  griddep_wait();
  griddep_launch();
  if (threadIdx.x == 0) {
    float lr_t = a.h.lr;
    if (OPT == DTF_OPT_ADAM) {
      // beta powers live on the PS (shared by all workers): t counts applies from everyone.
      const float b1p = *reinterpret_cast<volatile float*>(&a.ctrl_powers->beta1_power);
      const float b2p = *reinterpret_cast<volatile float*>(&a.ctrl_powers->beta2_power);
      lr_t = a.h.lr * sqrtf(1.f - b2p) / (1.f - b1p);
    }
    s_lr_t = lr_t;
  }
  // grid-stride over the 4096-element chunks: the grid may be capped (max_ctas) so that a big
  // apply running beside compute kernels holds one slot per SM instead of flooding them
  for (int vb = blockIdx.x; vb < a.total_blocks; vb += gridDim.x) {
  if (threadIdx.x == 0) {
    int it = 0;
    while (it + 1 < a.n_items && vb >= a.block_start[it + 1]) ++it;
    s_item = it;
  }
  __syncthreads();
  const dtf_apply_item& it = a.items[s_item];
  const float lr_t = s_lr_t;
  const long long base = static_cast<long long>(vb - a.block_start[s_item]) * APPLY_CHUNK;
  const bool vec_ok =
      ((reinterpret_cast<uintptr_t>(it.var) | reinterpret_cast<uintptr_t>(it.grad) |
        reinterpret_cast<uintptr_t>(it.slot0) | reinterpret_cast<uintptr_t>(it.slot1)) & 15) == 0;
  __nv_bfloat16* shadow = reinterpret_cast<__nv_bfloat16*>(it.shadow);
  const bool shadow_vec = shadow != nullptr && (it.cols % 4 == 0) && (it.shadow_ld % 4 == 0) &&
                          (it.first % 4 == 0) &&
                          ((reinterpret_cast<uintptr_t>(shadow) & 7) == 0);

  // Phase 1: issue every 16-byte load of this thread's APPLY_UNROLL vectors before using any of
  // them (up to 16 requests in flight per thread; over NVLink each one is a ~2 us round trip).
  const bool rmw = !(OPT == DTF_OPT_SGD && a.h.atomic_sgd);
  float4 ldg_[APPLY_UNROLL], ldv_[APPLY_UNROLL], ld0_[APPLY_UNROLL], ld1_[APPLY_UNROLL];
#pragma unroll
  for (int u = 0; u < APPLY_UNROLL; ++u) {
    const long long i = base + (static_cast<long long>(u) * APPLY_THREADS + threadIdx.x) * APPLY_VEC;
    ldg_[u] = ldv_[u] = ld0_[u] = ld1_[u] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (vec_ok && i + APPLY_VEC <= it.n) {
      ldg_[u] = *reinterpret_cast<const float4*>(it.grad + i);
      if (it.zero_grad)  // self-cleaning accumulator: ready for the next step's atomics
        *reinterpret_cast<float4*>(const_cast<float*>(it.grad) + i) = make_float4(0.f, 0.f, 0.f, 0.f);
      if (rmw) {
        ldv_[u] = *reinterpret_cast<const float4*>(it.var + i);
        if (OPT != DTF_OPT_SGD) ld0_[u] = *reinterpret_cast<const float4*>(it.slot0 + i);
        if (OPT == DTF_OPT_ADAM) ld1_[u] = *reinterpret_cast<const float4*>(it.slot1 + i);
      }
    }
  }
  // Phase 2: update + store.
#pragma unroll
  for (int u = 0; u < APPLY_UNROLL; ++u) {
    const long long i = base + (static_cast<long long>(u) * APPLY_THREADS + threadIdx.x) * APPLY_VEC;
    if (i >= it.n) continue;
    if (vec_ok && i + APPLY_VEC <= it.n) {
      const float4 g4 = ldg_[u];
      float g[4] = {g4.x * a.h.grad_scale, g4.y * a.h.grad_scale, g4.z * a.h.grad_scale,
                    g4.w * a.h.grad_scale};
      if (!rmw) {
        // write-only push: var += -lr * g, one 16-byte reduction atomic at the owner's L2
        red_add_v4_f32_sys(it.var + i, -a.h.lr * g[0], -a.h.lr * g[1], -a.h.lr * g[2], -a.h.lr * g[3]);
        continue;  // shadow is refreshed by the owner in this mode
      }
      const float4 v4 = ldv_[u];
      float v[4] = {v4.x, v4.y, v4.z, v4.w};
      float s0[4] = {ld0_[u].x, ld0_[u].y, ld0_[u].z, ld0_[u].w};
      float s1[4] = {ld1_[u].x, ld1_[u].y, ld1_[u].z, ld1_[u].w};
#pragma unroll
      for (int j = 0; j < 4; ++j) update_one<OPT>(g[j], v[j], s0[j], s1[j], a.h, lr_t);
      *reinterpret_cast<float4*>(it.var + i) = make_float4(v[0], v[1], v[2], v[3]);
      if (OPT != DTF_OPT_SGD)
        *reinterpret_cast<float4*>(it.slot0 + i) = make_float4(s0[0], s0[1], s0[2], s0[3]);
      if (OPT == DTF_OPT_ADAM)
        *reinterpret_cast<float4*>(it.slot1 + i) = make_float4(s1[0], s1[1], s1[2], s1[3]);
      if (shadow != nullptr) {
        const long long f = it.first + i;
        if (shadow_vec) {
          const long long r = f / it.cols, c = f % it.cols;
          uint2 o;
          o.x = pack_bf16x2(v[0], v[1]);
          o.y = pack_bf16x2(v[2], v[3]);
          *reinterpret_cast<uint2*>(shadow + r * it.shadow_ld + c) = o;
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const long long r = (f + j) / it.cols, c = (f + j) % it.cols;
            shadow[r * it.shadow_ld + c] = __float2bfloat16(v[j]);
          }
        }
      }
    } else {
      for (int j = 0; j < APPLY_VEC && i + j < it.n; ++j) {
        const float g = it.grad[i + j] * a.h.grad_scale;
        if (it.zero_grad) const_cast<float*>(it.grad)[i + j] = 0.f;
        if (OPT == DTF_OPT_SGD && a.h.atomic_sgd) {
          red_add_f32_sys(it.var + i + j, -a.h.lr * g);
          continue;
        }
        float v = it.var[i + j];
        float s0 = (OPT != DTF_OPT_SGD) ? it.slot0[i + j] : 0.f;
        float s1 = (OPT == DTF_OPT_ADAM) ? it.slot1[i + j] : 0.f;
        update_one<OPT>(g, v, s0, s1, a.h, lr_t);
        it.var[i + j] = v;
        if (OPT != DTF_OPT_SGD) it.slot0[i + j] = s0;
        if (OPT == DTF_OPT_ADAM) it.slot1[i + j] = s1;
        if (shadow != nullptr) {
          const long long f = it.first + i + j;
          shadow[(f / it.cols) * it.shadow_ld + (f % it.cols)] = __float2bfloat16(v);
        }
      }
    }
  }

  __syncthreads();   // s_item is rewritten by the next chunk
  }  // grid-stride loop

  if (!a.finish) return;
  // ---- elect the last CTA; it performs _finish (K10) ----
  // bar.sync, then ONE thread fences (cumulative: it orders the whole CTA's stores) and takes a
  // ticket. Device scope is enough: nothing is synchronised on global_step across GPUs (updates are
  // Hogwild), and the host only reads a shard after the stream or the result word says so.
  __shared__ float s_loss;
  __shared__ unsigned int s_seq;
  __shared__ unsigned long long s_step;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned int old = atomicInc(a.block_counter, gridDim.x - 1u);
    s_last = (old == gridDim.x - 1u) ? 1u : 0u;
  }
  __syncthreads();
  if (!s_last) return;
  // three independent memory round trips, taken by three warps at the same time
  const bool sink = a.sink.host_dst != nullptr;
  if (threadIdx.x == 0) {
    if (OPT == DTF_OPT_ADAM) {
      volatile float* b1 = &a.ctrl_powers->beta1_power;
      volatile float* b2 = &a.ctrl_powers->beta2_power;
      *b1 = *b1 * a.h.beta1;
      *b2 = *b2 * a.h.beta2;
    }
    if (a.finish != 2) {  // finish == 2: global_step / results were already reported (dtf_step_report)
      const unsigned long long old = atom_add_u64_sys(&a.ctrl_step->global_step, 1ull);
      s_step = old + 1ull;
      if (a.step_out != nullptr) *a.step_out = old + 1ull;
      if (a.worker_id >= 0 && a.worker_id < DTF_MAX_WORKERS)
        *reinterpret_cast<volatile unsigned long long*>(&a.ctrl_step->heartbeat[a.worker_id]) =
            old + 1ull;
      atomicAdd(&a.ctrl_step->apply_count, 1u);
    }
  } else if (threadIdx.x == 32 && sink) {
    s_loss = *reinterpret_cast<const volatile float*>(a.sink.loss_src);
    if (a.sink.clear_loss) *const_cast<volatile float*>(a.sink.loss_src) = 0.f;
  } else if (threadIdx.x == 64 && sink) {
    const unsigned int q = *reinterpret_cast<volatile unsigned int*>(a.sink.seq_dev) + 1u;
    *reinterpret_cast<volatile unsigned int*>(a.sink.seq_dev) = q;
    s_seq = q;
  }
  __syncthreads();
  if (threadIdx.x == 0 && sink) {
    // ONE aligned 16-byte store into pinned host memory = one PCIe write: the host, spinning on
    // the sequence word, can never see it without the loss and the step next to it
    const unsigned long long st = s_step;
    asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(a.sink.host_dst),
                 "r"(__float_as_uint(s_loss)), "r"(s_seq), "r"(static_cast<unsigned int>(st)),
                 "r"(static_cast<unsigned int>(st >> 32))
                 : "memory");
  }
}

// Early step report (see dtf_step_report): one thread bumps / reads global_step and hands the step's
// loss and step number to the host while the rest of the backward pass is still running.
__global__ void step_report_kernel(dtf_ps_ctrl* ctrl_step, int worker_id, int bump,
                                   unsigned long long* step_out, dtf_result_sink sink) {
  griddep_wait();      // launched through launch_kernel(): honour the PDL contract (common.cuh)
  griddep_launch();
  if (threadIdx.x != 0) return;
  unsigned long long step;
  if (bump) {
    step = atom_add_u64_sys(&ctrl_step->global_step, 1ull) + 1ull;
    if (worker_id >= 0 && worker_id < DTF_MAX_WORKERS)
      *reinterpret_cast<volatile unsigned long long*>(&ctrl_step->heartbeat[worker_id]) = step;
    atomicAdd(&ctrl_step->apply_count, 1u);
  } else {
    step = ld_acquire_sys_u64(&ctrl_step->global_step);
  }
  if (step_out != nullptr) *step_out = step;
  if (sink.host_dst != nullptr) {
    const float loss = *reinterpret_cast<const volatile float*>(sink.loss_src);
    if (sink.clear_loss) *const_cast<volatile float*>(sink.loss_src) = 0.f;
    const unsigned int q = *reinterpret_cast<volatile unsigned int*>(sink.seq_dev) + 1u;
    *reinterpret_cast<volatile unsigned int*>(sink.seq_dev) = q;
    asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(sink.host_dst),
                 "r"(__float_as_uint(loss)), "r"(q), "r"(static_cast<unsigned int>(step)),
                 "r"(static_cast<unsigned int>(step >> 32))
                 : "memory");
  }
}

__global__ void refresh_shadow_kernel(const float* __restrict__ var,
                                      __nv_bfloat16* __restrict__ shadow, long long n,
                                      long long first, int cols, int shadow_ld) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long f = first + i;
    shadow[(f / cols) * shadow_ld + (f % cols)] = __float2bfloat16(var[i]);
  }
}

// Test hook: the optimizer rules' division / square root next to the compiler's IEEE ones.
__global__ void div_sqrt_probe_kernel(const float* a, const float* b, float* q_fast, float* q_ieee,
                                      float* s_fast, float* s_ieee, long long n) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    q_fast[i] = div_rn(a[i], b[i]);
    q_ieee[i] = __fdiv_rn(a[i], b[i]);
    s_fast[i] = sqrt_rn(b[i]);
    s_ieee[i] = __fsqrt_rn(b[i]);
  }
}

}  // namespace dtf

using namespace dtf;

extern "C" int dtf_debug_div_sqrt(const float* a, const float* b, float* q_fast, float* q_ieee,
                                  float* s_fast, float* s_ieee, long long n, void* stream) {
  div_sqrt_probe_kernel<<<296, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(a, b, q_fast, q_ieee,
                                                                                   s_fast, s_ieee, n);
  DTF_LAUNCH_CHECK();
  return 0;
}

extern "C" int dtf_push_apply(int opt, const dtf_apply_item* items, int n_items,
                              const dtf_apply_hyper* hyper, dtf_ps_ctrl* ctrl_for_powers,
                              dtf_ps_ctrl* ctrl_for_step, int finish, int worker_id,
                              unsigned long long* step_out, unsigned int* block_counter,
                              void* stream_) {
  return dtf_push_apply_ex(opt, items, n_items, hyper, ctrl_for_powers, ctrl_for_step, finish,
                           worker_id, step_out, block_counter, nullptr, 0, stream_);
}

extern "C" int dtf_push_apply_ex(int opt, const dtf_apply_item* items, int n_items,
                                 const dtf_apply_hyper* hyper, dtf_ps_ctrl* ctrl_for_powers,
                                 dtf_ps_ctrl* ctrl_for_step, int finish, int worker_id,
                                 unsigned long long* step_out, unsigned int* block_counter,
                                 const dtf_result_sink* sink, int max_ctas, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  DTF_REQUIRE(n_items >= 0 && n_items <= DTF_MAX_APPLY_ITEMS, "push_apply: %d items (max %d)",
              n_items, DTF_MAX_APPLY_ITEMS);
  DTF_REQUIRE(!finish || (ctrl_for_step != nullptr && block_counter != nullptr),
              "push_apply: finish needs ctrl_for_step and block_counter");
  DTF_REQUIRE(!(sink != nullptr && finish == 2),
              "push_apply: finish = 2 (results already reported by dtf_step_report) takes no result sink");
  DTF_REQUIRE(opt != DTF_OPT_ADAM || ctrl_for_powers != nullptr,
              "push_apply: Adam needs the control block that owns the beta powers");
  ApplyArgs a;
  memset(&a, 0, sizeof(a));
  int blocks = 0;
  for (int i = 0; i < n_items; ++i) {
    a.items[i] = items[i];
    a.block_start[i] = blocks;
    DTF_REQUIRE(items[i].n >= 0 && items[i].var && items[i].grad, "push_apply: item %d invalid", i);
    DTF_REQUIRE(opt == DTF_OPT_SGD || items[i].slot0, "push_apply: item %d lacks slot0", i);
    DTF_REQUIRE(opt != DTF_OPT_ADAM || items[i].slot1, "push_apply: item %d lacks slot1", i);
    blocks += static_cast<int>((items[i].n + APPLY_CHUNK - 1) / APPLY_CHUNK);
  }
  a.block_start[n_items] = blocks;
  if (blocks == 0) {  // nothing to apply but the step may still need its _finish
    DTF_REQUIRE(n_items > 0 || finish, "push_apply: nothing to do");
    blocks = 1;
    if (n_items == 0) {
      a.items[0].n = 0;
      n_items = 1;
      a.block_start[1] = 1;
    }
  }
  a.n_items = n_items;
  a.total_blocks = blocks;
  a.h = *hyper;
  a.ctrl_powers = ctrl_for_powers;
  a.ctrl_step = ctrl_for_step;
  a.finish = finish;
  a.worker_id = worker_id;
  a.step_out = step_out;
  a.block_counter = block_counter;
  if (sink != nullptr && finish) {
    DTF_REQUIRE(sink->loss_src && sink->host_dst && sink->seq_dev, "push_apply: incomplete result sink");
    a.sink = *sink;
  }
  const int grid = (max_ctas > 0 && blocks > max_ctas) ? max_ctas : blocks;
  switch (opt) {
    case DTF_OPT_SGD: DTF_CHECK_CUDA(launch_kernel(push_apply_kernel<DTF_OPT_SGD>, dim3(grid), dim3(APPLY_THREADS), 0, stream, a)); break;
    case DTF_OPT_MOMENTUM: DTF_CHECK_CUDA(launch_kernel(push_apply_kernel<DTF_OPT_MOMENTUM>, dim3(grid), dim3(APPLY_THREADS), 0, stream, a)); break;
    case DTF_OPT_ADAGRAD: DTF_CHECK_CUDA(launch_kernel(push_apply_kernel<DTF_OPT_ADAGRAD>, dim3(grid), dim3(APPLY_THREADS), 0, stream, a)); break;
    case DTF_OPT_ADAM: DTF_CHECK_CUDA(launch_kernel(push_apply_kernel<DTF_OPT_ADAM>, dim3(grid), dim3(APPLY_THREADS), 0, stream, a)); break;
    default: DTF_REQUIRE(false, "push_apply: unknown optimizer %d", opt);
  }
  DTF_LAUNCH_CHECK();
  return 0;
}

extern "C" int dtf_refresh_shadow(const float* var, void* shadow, long long n, long long first,
                                  int cols, int shadow_ld, void* stream) {
  if (n <= 0) return 0;
  int blocks = static_cast<int>((n + 1023) / 1024);
  if (blocks > 1184) blocks = 1184;
  refresh_shadow_kernel<<<blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      var, reinterpret_cast<__nv_bfloat16*>(shadow), n, first, cols, shadow_ld);
  DTF_LAUNCH_CHECK();
  return 0;
}

// ================================================================================================
// dtf_push_apply_static — the same push+apply, for callers that must not allocate.
//
// dtf_push_apply needs one word of device scratch (the counter that elects the last CTA) and
// makes its caller provide it. Some callers have nowhere to get that word from: a parameter-server
// task that handed the whole HBM budget to its shard arena at start-up, a host thread that must
// not enter cudaMalloc while the workers' kernels are in flight (cudaMalloc synchronises the
// device and takes the driver's allocator lock), or a foreign runtime that reaches this library
// only through the C ABI and owns no allocator we could call. For them the scratch lives inside
// the library image: a file-scope __device__ array of DTF_PUSH_APPLY_STATIC_SCRATCH_BYTES (256)
// bytes, which every GPU that loads the library keeps until the process exits.
//
// Rules that follow from sharing one array:
//   * every call looks up the array's device address (cudaGetSymbolAddress, with the target
//     device current); the first call on a device also loads the module, so make that first call
//     before starting any stream capture on that device;
//   * every call clears the counter on `stream` before its kernel, because the array keeps
//     whatever the previous call (or an aborted launch) left in it;
//   * two calls whose kernels overlap on one GPU — different streams, or different host threads —
//     race on the counter and may elect the wrong "last" CTA: keep calls on one GPU ordered;
//   * requests beyond DTF_MAX_APPLY_ITEMS slices or DTF_PUSH_APPLY_STATIC_MAX_ELEMS elements per
//     call are refused; split them or use dtf_push_apply with caller-owned scratch.
// Nothing here calls cudaMalloc, cudaFree or any allocating library routine.
// ================================================================================================
namespace dtf {

constexpr int kStaticScratchWords = DTF_PUSH_APPLY_STATIC_SCRATCH_BYTES / 4;
constexpr int kStaticMaxDevices = 64;

// In host code this name is only a handle for cudaGetSymbolAddress; it is NOT a device pointer.
static __device__ __align__(256) unsigned int g_push_apply_static_scratch[kStaticScratchWords];

// Resolved on EVERY call: the address belongs to the current context's instance of this module, so
// a pointer remembered per device ordinal would go stale after cudaDeviceReset or under a caller's
// own (non-primary) context. After the first call on a device - which loads the module and therefore
// belongs before any stream capture - this is a host-side table lookup.
static int static_scratch_for_current_device(unsigned int** out) {
  int dev = -1;
  DTF_CHECK_CUDA(cudaGetDevice(&dev));
  DTF_REQUIRE(dev >= 0 && dev < kStaticMaxDevices, "push_apply_static: device %d out of range", dev);
  void* addr = nullptr;
  cudaError_t e = cudaGetSymbolAddress(&addr, g_push_apply_static_scratch);
  if (e != cudaSuccess || addr == nullptr) {
    set_error("push_apply_static: cudaGetSymbolAddress failed on device %d: %s", dev,
              cudaGetErrorString(e));
    return -1;
  }
  *out = static_cast<unsigned int*>(addr);
  return 0;
}

}  // namespace dtf

extern "C" int dtf_push_apply_static(int opt, const dtf_apply_item* items, int n_items,
                                     const dtf_apply_hyper* hyper, dtf_ps_ctrl* ctrl_for_powers,
                                     dtf_ps_ctrl* ctrl_for_step, int finish, int worker_id,
                                     unsigned long long* step_out, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  DTF_REQUIRE(items != nullptr && hyper != nullptr, "push_apply_static: items/hyper is null");
  DTF_REQUIRE(n_items >= 1 && n_items <= DTF_MAX_APPLY_ITEMS,
              "push_apply_static: %d slices in one call, the limit is %d (split the call)", n_items,
              DTF_MAX_APPLY_ITEMS);
  DTF_REQUIRE(!finish || ctrl_for_step != nullptr, "push_apply_static: finish needs ctrl_for_step");
  DTF_REQUIRE(opt != DTF_OPT_ADAM || ctrl_for_powers != nullptr,
              "push_apply_static: Adam needs the control block that owns the beta powers");
  ApplyArgs a;
  memset(&a, 0, sizeof(a));
  long long total = 0;
  int blocks = 0;
  for (int i = 0; i < n_items; ++i) {
    DTF_REQUIRE(items[i].n >= 0 && items[i].var && items[i].grad,
                "push_apply_static: item %d invalid", i);
    DTF_REQUIRE(opt == DTF_OPT_SGD || items[i].slot0, "push_apply_static: item %d lacks slot0", i);
    DTF_REQUIRE(opt != DTF_OPT_ADAM || items[i].slot1, "push_apply_static: item %d lacks slot1", i);
    total += items[i].n;
    DTF_REQUIRE(total <= DTF_PUSH_APPLY_STATIC_MAX_ELEMS,
                "push_apply_static: more than %lld elements in one call (limit of the "
                "no-allocation entry point; split the call or use dtf_push_apply)",
                (long long)DTF_PUSH_APPLY_STATIC_MAX_ELEMS);
    a.items[i] = items[i];
    a.block_start[i] = blocks;
    blocks += static_cast<int>((items[i].n + APPLY_CHUNK - 1) / APPLY_CHUNK);
  }
  a.block_start[n_items] = blocks;
  if (blocks == 0) blocks = 1;  // all slices empty: one CTA still performs the _finish
  a.n_items = n_items;
  a.total_blocks = blocks;
  a.h = *hyper;
  a.ctrl_powers = ctrl_for_powers;
  a.ctrl_step = ctrl_for_step;
  a.finish = finish;
  a.worker_id = worker_id;
  a.step_out = step_out;
  unsigned int* scratch = nullptr;
  if (static_scratch_for_current_device(&scratch) != 0) return -1;
  a.block_counter = scratch;
  // the array keeps the previous call's contents: the election counter must start at zero
  DTF_CHECK_CUDA(cudaMemsetAsync(scratch, 0, sizeof(unsigned int), stream));
  switch (opt) {
    case DTF_OPT_SGD: DTF_CHECK_CUDA(launch_kernel(push_apply_kernel<DTF_OPT_SGD>, dim3(blocks), dim3(APPLY_THREADS), 0, stream, a)); break;
    case DTF_OPT_MOMENTUM: DTF_CHECK_CUDA(launch_kernel(push_apply_kernel<DTF_OPT_MOMENTUM>, dim3(blocks), dim3(APPLY_THREADS), 0, stream, a)); break;
    case DTF_OPT_ADAGRAD: DTF_CHECK_CUDA(launch_kernel(push_apply_kernel<DTF_OPT_ADAGRAD>, dim3(blocks), dim3(APPLY_THREADS), 0, stream, a)); break;
    case DTF_OPT_ADAM: DTF_CHECK_CUDA(launch_kernel(push_apply_kernel<DTF_OPT_ADAM>, dim3(blocks), dim3(APPLY_THREADS), 0, stream, a)); break;
    default: DTF_REQUIRE(false, "push_apply_static: unknown optimizer %d", opt);
  }
  DTF_LAUNCH_CHECK();
  return 0;
}

extern "C" int dtf_step_report(dtf_ps_ctrl* ctrl_for_step, int worker_id, int bump,
                               unsigned long long* step_out, const dtf_result_sink* sink,
                               void* stream) {
  DTF_REQUIRE(ctrl_for_step != nullptr, "step_report: ctrl_for_step is null");
  dtf_result_sink sk;
  memset(&sk, 0, sizeof(sk));
  if (sink != nullptr) {
    DTF_REQUIRE(sink->loss_src && sink->host_dst && sink->seq_dev, "step_report: incomplete result sink");
    sk = *sink;
  }
  DTF_CHECK_CUDA(launch_kernel(dtf::step_report_kernel, dim3(1), dim3(32), 0,
                               reinterpret_cast<cudaStream_t>(stream), ctrl_for_step, worker_id, bump,
                               step_out, sk));
  return 0;
}
