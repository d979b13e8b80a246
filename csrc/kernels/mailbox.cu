// Asynchronous "owner applies" gradient push (SURVEY 5.9 option (b); K9 + K11).
//
// dtf_push_apply lets the WORKER's SMs read-modify-write the owner's var / m / v over NVLink:
// 12 bytes of round-trip reads and 14 bytes of writes per Adam parameter. At 8 GPUs that traffic
// (74 MB per CNN step per worker) is what limits scaling. The reference's PS applies the gradient
// itself after receiving it (the ApplyAdam op is placed on /job:ps, run_mnist_distributed.py:107,
// 116) — this file is that protocol over peer memory:
//
//   worker   mailbox_push_kernel   posted 16-byte stores of the fp32 gradient into its own mailbox
//                                  on every owning shard (4 bytes per parameter, no reads), then
//                                  ONE st.release.sys of a sequence number per owner;
//   owner    mailbox_drain_kernel  part of the owner's own step (side stream): reads var / slots
//                                  once from LOCAL HBM, applies every pending worker's gradient in
//                                  turn, writes var / slots / bf16 serving copy once, publishes
//                                  consumed[w] and bumps global_step by the number applied.
//
// Flow control is a counter pair per (owner, worker): the worker never runs more than `depth`
// pushes ahead of the owner's applies, which bounds staleness (TF1's async mode is unbounded).
#include "common.cuh"
#include "ptx.cuh"
#include "apply_rules.cuh"
#include "../../include/dtf_b200.h"

namespace dtf {

constexpr int MBOX_THREADS = 256;
constexpr int MBOX_UNROLL = 8;
constexpr long long MBOX_CHUNK = static_cast<long long>(MBOX_THREADS) * MBOX_UNROLL * 16;  // 32 KB

struct MboxPushArgs {
  dtf_mbox_push_item items[DTF_MAX_MBOX_ITEMS];
  int block_start[DTF_MAX_MBOX_ITEMS + 1];
  int n_items;
  int total_blocks;
  int depth;
  int worker_id;
  dtf_mbox_flags* owner_flags[DTF_MAX_WORKERS];
  int n_owners;
  int publish;
  unsigned int* seq;
  unsigned int* block_counter;
  unsigned int* min_consumed_out;
  const dtf_ps_ctrl* ctrl_step;
  unsigned long long* step_out;
  dtf_result_sink sink;
  const dtf_mirror_hdr* my_hdr;
  const dtf_mirror_hdr* flow_hdr;   // device-side flow control against this header's consumed[]
  const unsigned int* sync_tag;     // sync mode: local_step of this gradient (device word) or NULL
};

__global__ void __launch_bounds__(MBOX_THREADS)
mailbox_push_kernel(const __grid_constant__ MboxPushArgs a) {
  __shared__ unsigned int s_last;
  __shared__ unsigned int s_min;
  griddep_wait();
  griddep_launch();
  // every CTA reads the sequence number before any CTA can reach the election below, and only the
  // elected (last) CTA changes it
  const unsigned int seq = *reinterpret_cast<volatile unsigned int*>(a.seq);
  const long long buf = seq % static_cast<unsigned int>(a.depth);
  {
    int it = 0;
    while (it + 1 < a.n_items && static_cast<int>(blockIdx.x) >= a.block_start[it + 1]) ++it;
    const dtf_mbox_push_item& c = a.items[it];
    const long long base = static_cast<long long>(blockIdx.x - a.block_start[it]) * MBOX_CHUNK;
    if (a.flow_hdr != nullptr && base < c.bytes && seq >= static_cast<unsigned int>(a.depth)) {
      // Device-side flow control (bounded staleness): buffer seq % depth still holds push
      // seq - depth until its owner has consumed it. The owners' drains keep consumed[] fresh in MY
      // mirror's header, so this polls local memory and the host never blocks on admission. A worker
      // only ever waits for owners that are behind it, so the waits cannot form a cycle.
      if (threadIdx.x == 0) {
        const volatile unsigned int* cons = &a.flow_hdr->consumed[c.owner];
        const unsigned int need = seq - static_cast<unsigned int>(a.depth) + 1u;
        const long long t0 = clock64();
        while (static_cast<int>(*cons - need) < 0) {
          if (*reinterpret_cast<const volatile unsigned int*>(&a.flow_hdr->abort) != 0u) break;
          __nanosleep(200);
          if (clock64() - t0 > 100ll * DTF_WAIT_TIMEOUT_CYCLES) {   // minutes: an owner may be paused for a checkpoint
            printf("dtf mailbox: worker %d waited too long for owner %d to consume push %u\n",
                   a.worker_id, c.owner, need);
            __trap();
          }
        }
        __threadfence();
      }
      __syncthreads();
    }
    if (base < c.bytes) {
      const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(c.src) + base);
      uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<char*>(c.dst0) + buf * c.buf_stride + base);
      const long long nvec = min(MBOX_CHUNK, c.bytes - base) / 16;
      uint4 v[MBOX_UNROLL];
#pragma unroll
      for (int u = 0; u < MBOX_UNROLL; ++u) {
        const long long i = static_cast<long long>(u) * MBOX_THREADS + threadIdx.x;
        if (i < nvec) v[u] = src[i];
      }
      if (c.zero_src) {
        uint4* z = const_cast<uint4*>(src);
#pragma unroll
        for (int u = 0; u < MBOX_UNROLL; ++u) {
          const long long i = static_cast<long long>(u) * MBOX_THREADS + threadIdx.x;
          if (i < nvec) z[i] = make_uint4(0u, 0u, 0u, 0u);
        }
      }
#pragma unroll
      for (int u = 0; u < MBOX_UNROLL; ++u) {
        const long long i = static_cast<long long>(u) * MBOX_THREADS + threadIdx.x;
        if (i < nvec) dst[i] = v[u];
      }
    }
  }
  if (!a.publish) return;
  // ---- the last CTA publishes the push to every owner ----
  // A system-scope fence costs microseconds here, so exactly ONE sits on the critical path: every
  // CTA orders its stores at GPU scope before taking its ticket (bar.sync, then one thread fences:
  // cumulative over the CTA), and the elected CTA's release stores below carry the system-scope
  // fence that makes everything that happened-before them visible to the owners.
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned int old = atomicInc(a.block_counter, static_cast<unsigned int>(a.total_blocks - 1));
    s_last = (old == static_cast<unsigned int>(a.total_blocks - 1)) ? 1u : 0u;
    s_min = 0xffffffffu;
    if (s_last) __threadfence();  // acquire side of the election
  }
  __syncthreads();
  if (!s_last) return;
  if (static_cast<int>(threadIdx.x) < a.n_owners) {
    dtf_mbox_flags* f = a.owner_flags[threadIdx.x];
    if (a.sync_tag != nullptr)   // ordered before produced[] by the release below
      *reinterpret_cast<volatile unsigned int*>(buf == 0 ? &f->tag0[a.worker_id] : &f->tag1[a.worker_id]) =
          *reinterpret_cast<const volatile unsigned int*>(a.sync_tag);
    st_release_sys_u32(&f->produced[a.worker_id], seq + 1u);
    // how far has this owner got with MY gradients (flow control for the next pushes): from my
    // mirror's header when the owners keep one fresh (local read), else over NVLink
    const unsigned int c =
        a.my_hdr != nullptr
            ? *reinterpret_cast<const volatile unsigned int*>(&a.my_hdr->consumed[threadIdx.x])
            : ld_acquire_sys_u32(&f->consumed[a.worker_id]);
    atomicMin(&s_min, c);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    *reinterpret_cast<volatile unsigned int*>(a.seq) = seq + 1u;
    if (a.min_consumed_out != nullptr) *a.min_consumed_out = s_min;
    // what sess.run([train_op, global_step]) reports: the shared counter as this step ends
    unsigned long long gs = 0;
    if (a.my_hdr != nullptr)
      gs = *reinterpret_cast<const volatile unsigned long long*>(&a.my_hdr->global_step);
    else if (a.ctrl_step != nullptr)
      gs = ld_acquire_sys_u64(&a.ctrl_step->global_step);
    if (a.step_out != nullptr && (a.ctrl_step != nullptr || a.my_hdr != nullptr)) *a.step_out = gs;
    if (a.sink.host_dst != nullptr) {
      const float loss = *reinterpret_cast<const volatile float*>(a.sink.loss_src);
      const unsigned int q = *reinterpret_cast<volatile unsigned int*>(a.sink.seq_dev) + 1u;
      *reinterpret_cast<volatile unsigned int*>(a.sink.seq_dev) = q;
      // one aligned 16-byte store = one PCIe write (see apply.cu)
      asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(a.sink.host_dst),
                   "r"(__float_as_uint(loss)), "r"(q), "r"(static_cast<unsigned int>(gs)),
                   "r"(static_cast<unsigned int>(gs >> 32))
                   : "memory");
    }
  }
}

// Claims the shard's drain lock and records which workers have an unconsumed gradient.
// todo[w] = 1 + buffer index of worker w's oldest pending gradient (0 = nothing);
// todo[DTF_MAX_WORKERS] = 1 when this launch holds the lock (the drain kernel releases it).
__global__ void mailbox_snapshot_kernel(dtf_mbox_flags* flags, unsigned int* todo, int n_workers,
                                        int depth) {
  griddep_wait();
  griddep_launch();
  const int w = threadIdx.x;
  unsigned int claimed = 0;
  if (w == 0) {
    if (*reinterpret_cast<volatile unsigned int*>(&flags->paused) == 0)
      claimed = (atomicCAS(&flags->lock, 0u, 1u) == 0u) ? 1u : 0u;
    __threadfence();   // acquire: the previous holder's consumed[] / state precede its unlock
    todo[DTF_MAX_WORKERS] = claimed;
  }
  claimed = __shfl_sync(0xffffffffu, claimed, 0);
  if (w >= DTF_MAX_WORKERS) return;
  unsigned int t = 0;
  if (claimed && w < n_workers) {
    const unsigned int p = ld_acquire_sys_u32(&flags->produced[w]);
    const unsigned int c = *reinterpret_cast<volatile unsigned int*>(&flags->consumed[w]);
    if (static_cast<int>(p - c) > 0) t = c % static_cast<unsigned int>(depth) + 1u;  // oldest first
  }
  todo[w] = t;
}

// Synchronous variant (see dtf_mailbox_drain_sync): classifies every worker's oldest pending
// gradient as fresh (tag == epoch) or stale (tag < epoch). todo[w] = 1 + buffer for the fresh ones
// when at least R are pending (they are averaged and applied once), todo[16] = lock held,
// todo[17] = bit mask of stale gradients to retire unapplied, todo[18] = number of fresh ones applied.
__global__ void mailbox_snapshot_sync_kernel(dtf_mbox_flags* flags, const dtf_ps_ctrl* ctrl,
                                             unsigned int* todo, int n_workers, int R) {
  griddep_wait();
  griddep_launch();
  const int w = threadIdx.x;
  unsigned int claimed = 0;
  if (w == 0) {
    if (*reinterpret_cast<volatile unsigned int*>(&flags->paused) == 0)
      claimed = (atomicCAS(&flags->lock, 0u, 1u) == 0u) ? 1u : 0u;
    __threadfence();
  }
  claimed = __shfl_sync(0xffffffffu, claimed, 0);
  const unsigned int epoch = *reinterpret_cast<const volatile unsigned int*>(&ctrl->epoch);
  bool fresh = false, stale = false;
  unsigned int buf = 0;
  if (claimed && w < n_workers && w < DTF_MAX_WORKERS) {
    const unsigned int p = ld_acquire_sys_u32(&flags->produced[w]);
    const unsigned int c = *reinterpret_cast<volatile unsigned int*>(&flags->consumed[w]);
    if (static_cast<int>(p - c) > 0) {
      buf = c & 1u;
      const unsigned int tag =
          *reinterpret_cast<volatile unsigned int*>(buf == 0 ? &flags->tag0[w] : &flags->tag1[w]);
      fresh = tag == epoch;
      stale = static_cast<int>(tag - epoch) < 0;
    }
  }
  const unsigned int fresh_mask = __ballot_sync(0xffffffffu, fresh);
  const unsigned int stale_mask = __ballot_sync(0xffffffffu, stale);
  const int n_fresh = __popc(fresh_mask);
  const bool apply = n_fresh >= R && n_fresh > 0;
  if (w < DTF_MAX_WORKERS) todo[w] = (fresh && apply) ? buf + 1u : 0u;
  if (w == 0) {
    todo[DTF_MAX_WORKERS] = claimed;
    todo[DTF_MAX_WORKERS + 1] = stale_mask;
    todo[DTF_MAX_WORKERS + 2] = apply ? static_cast<unsigned int>(n_fresh) : 0u;
  }
}

constexpr int DRAIN_THREADS = 256;
constexpr int DRAIN_VEC = 4;                                           // one float4 per access, lane-contiguous
constexpr int DRAIN_UNROLL = 4;                                        // independent vectors per thread
constexpr int DRAIN_CHUNK = DRAIN_THREADS * DRAIN_VEC * DRAIN_UNROLL;  // 4096 elements / CTA pass

struct DrainArgs {
  dtf_apply_item items[DTF_MAX_APPLY_ITEMS];
  dtf_drain_mirror_item mirror[DTF_MAX_APPLY_ITEMS];
  int block_start[DTF_MAX_APPLY_ITEMS + 1];
  int n_items;
  int total_chunks;
  dtf_apply_hyper h;
  dtf_ps_ctrl* ctrl_powers;
  dtf_ps_ctrl* ctrl_step;
  dtf_mbox_flags* flags;
  const unsigned int* todo;
  long long slot_stride;  // bytes
  int depth;
  int n_workers;
  int bcast;              // mirror[] is valid: broadcast into flags->mirror_base[*]
  int shard;
  int mean;               // synchronous round: average the pending gradients, apply once, advance the epoch
  unsigned long long* step_out;
  unsigned int* block_counter;
};

// bf16 store of 4 consecutive elements of a [rows, cols] tensor kept with row pitch ld; f = flat
// index of v[0]. vec: cols, ld, f are multiples of 4 and base is 8-byte aligned (one 8-byte store
// per lane, lane-contiguous: a warp writes 256 contiguous bytes).
__device__ __forceinline__ void put_bf16x4(__nv_bfloat16* base, long long f, int cols, int ld,
                                           const float* v, bool vec) {
  // 32-bit index math whenever it fits (a 64-bit division costs ~10x a 32-bit one)
  long long r;
  int c;
  if (f < 0x7fffffffLL) {
    const unsigned int fu = static_cast<unsigned int>(f), cu = static_cast<unsigned int>(cols);
    const unsigned int ru = fu / cu;
    r = ru;
    c = static_cast<int>(fu - ru * cu);
  } else {
    r = f / cols;
    c = static_cast<int>(f - r * cols);
  }
  if (vec) {
    *reinterpret_cast<uint2*>(base + r * ld + c) =
        make_uint2(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]));
    return;
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    base[r * ld + c] = __float2bfloat16(v[j]);
    if (++c == cols) { c = 0; ++r; }
  }
}

template <int OPT>
__global__ void __launch_bounds__(DRAIN_THREADS)
mailbox_drain_kernel(const __grid_constant__ DrainArgs a) {
  __shared__ int s_n;
  __shared__ long long s_off[DTF_MAX_WORKERS];  // element offset of pending gradient k from item.grad
  __shared__ float s_lr[DTF_MAX_WORKERS];
  __shared__ unsigned long long s_mbase[DTF_MAX_WORKERS];
  __shared__ int s_nm;
  __shared__ unsigned int s_last;
  __shared__ unsigned long long s_step;
  griddep_wait();
  griddep_launch();
  if (threadIdx.x == 0) {
    float b1p = 0.f, b2p = 0.f;
    if (OPT == DTF_OPT_ADAM) {
      b1p = *reinterpret_cast<volatile float*>(&a.ctrl_powers->beta1_power);
      b2p = *reinterpret_cast<volatile float*>(&a.ctrl_powers->beta2_power);
    }
    int n = 0;
    for (int w = 0; w < a.n_workers; ++w) {
      const unsigned int t = a.todo[w];
      if (t == 0) continue;
      s_off[n] = (static_cast<long long>(w) * a.depth + (t - 1)) * (a.slot_stride / 4);
      // the k-th apply of this drain sees the powers the k-1 applies before it left behind
      s_lr[n] = (OPT == DTF_OPT_ADAM) ? a.h.lr * sqrtf(1.f - b2p) / (1.f - b1p) : a.h.lr;
      b1p *= a.h.beta1;
      b2p *= a.h.beta2;
      ++n;
    }
    s_n = n;
    int nm = 0;   // registered mirrors, packed
    if (a.bcast && n > 0) {
      for (int w = 0; w < DTF_MAX_WORKERS; ++w) {
        const unsigned long long b =
            *reinterpret_cast<volatile unsigned long long*>(&a.flags->mirror_base[w]);
        if (b != 0ull) s_mbase[nm++] = b;
      }
    }
    s_nm = nm;
  }
  __syncthreads();
  const int n_pending = s_n;
  const int n_mirror = s_nm;
  // the grid may be capped (a drain that runs beside compute kernels should hold a few CTA slots,
  // not flood the SMs): every CTA strides over the 4096-element chunks
  for (int chunk = blockIdx.x; n_pending > 0 && chunk < a.total_chunks; chunk += gridDim.x) {
  int item = 0;
  while (item + 1 < a.n_items && chunk >= a.block_start[item + 1]) ++item;
  const dtf_apply_item& it = a.items[item];
  const dtf_drain_mirror_item& mi = a.mirror[item];
  if (it.n > 0) {
    const long long base = static_cast<long long>(chunk - a.block_start[item]) * DRAIN_CHUNK;
    const bool vec_ok =
        ((reinterpret_cast<uintptr_t>(it.var) | reinterpret_cast<uintptr_t>(it.grad) |
          reinterpret_cast<uintptr_t>(it.slot0) | reinterpret_cast<uintptr_t>(it.slot1)) & 15) == 0 &&
        (a.slot_stride % 16 == 0);
    __nv_bfloat16* shadow = reinterpret_cast<__nv_bfloat16*>(it.shadow);
    const bool geo4 = (it.cols % 4 == 0) && (it.shadow_ld % 4 == 0);
    const bool shadow_vec = geo4 && (it.first % 4 == 0) && ((reinterpret_cast<uintptr_t>(shadow) & 7) == 0);
    const bool mirror_vec = geo4 && (mi.first % 4 == 0) && (mi.bf16_off % 8 == 0);
    const bool mirror_f32_vec = (mi.first % 4 == 0) && (mi.f32_off % 16 == 0);
    float4 v4[DRAIN_UNROLL], m4[DRAIN_UNROLL], u4[DRAIN_UNROLL], g4[DRAIN_UNROLL];
    bool full[DRAIN_UNROLL];
    // phase 1: every 16-byte load of this thread's vectors (state + the first gradient) is issued
    // before any is used: up to 16 requests in flight per thread
#pragma unroll
    for (int u = 0; u < DRAIN_UNROLL; ++u) {
      const long long i = base + (static_cast<long long>(u) * DRAIN_THREADS + threadIdx.x) * DRAIN_VEC;
      full[u] = vec_ok && i + DRAIN_VEC <= it.n;
      v4[u] = m4[u] = u4[u] = g4[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (full[u]) {
        v4[u] = *reinterpret_cast<const float4*>(it.var + i);
        if (OPT != DTF_OPT_SGD) m4[u] = *reinterpret_cast<const float4*>(it.slot0 + i);
        if (OPT == DTF_OPT_ADAM) u4[u] = *reinterpret_cast<const float4*>(it.slot1 + i);
        // written by a peer GPU since the last drain: read at L2, not through .nc
        const uint4 r = ld_relaxed_u4(reinterpret_cast<const uint4*>(it.grad + s_off[0] + i));
        g4[u] = make_float4(__uint_as_float(r.x), __uint_as_float(r.y), __uint_as_float(r.z),
                            __uint_as_float(r.w));
      }
    }
    // phase 2: apply the pending gradients in worker order (synchronous round: sum them and apply
    // their mean once); gradient k + 1 is in flight while gradient k is consumed
    float4 acc[DRAIN_UNROLL];
#pragma unroll
    for (int u = 0; u < DRAIN_UNROLL; ++u) acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
    for (int k = 0; k < n_pending; ++k) {
      const float lr_t = s_lr[k];
      float4 gn[DRAIN_UNROLL];
      if (k + 1 < n_pending) {
        const float* g = it.grad + s_off[k + 1];
#pragma unroll
        for (int u = 0; u < DRAIN_UNROLL; ++u) {
          const long long i = base + (static_cast<long long>(u) * DRAIN_THREADS + threadIdx.x) * DRAIN_VEC;
          gn[u] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (full[u]) {
            const uint4 r = ld_relaxed_u4(reinterpret_cast<const uint4*>(g + i));
            gn[u] = make_float4(__uint_as_float(r.x), __uint_as_float(r.y), __uint_as_float(r.z),
                                __uint_as_float(r.w));
          }
        }
      }
      const float s = a.h.grad_scale;
      if (a.mean) {
#pragma unroll
        for (int u = 0; u < DRAIN_UNROLL; ++u) {
          acc[u].x += g4[u].x; acc[u].y += g4[u].y; acc[u].z += g4[u].z; acc[u].w += g4[u].w;
        }
        if (k + 1 == n_pending) {
          const float sm = s / static_cast<float>(n_pending);
          const float lr0 = s_lr[0];
#pragma unroll
          for (int u = 0; u < DRAIN_UNROLL; ++u) {
            if (!full[u]) continue;
            update_one<OPT>(acc[u].x * sm, v4[u].x, m4[u].x, u4[u].x, a.h, lr0);
            update_one<OPT>(acc[u].y * sm, v4[u].y, m4[u].y, u4[u].y, a.h, lr0);
            update_one<OPT>(acc[u].z * sm, v4[u].z, m4[u].z, u4[u].z, a.h, lr0);
            update_one<OPT>(acc[u].w * sm, v4[u].w, m4[u].w, u4[u].w, a.h, lr0);
          }
        }
      } else {
#pragma unroll
      for (int u = 0; u < DRAIN_UNROLL; ++u) {
        if (!full[u]) continue;
        update_one<OPT>(g4[u].x * s, v4[u].x, m4[u].x, u4[u].x, a.h, lr_t);
        update_one<OPT>(g4[u].y * s, v4[u].y, m4[u].y, u4[u].y, a.h, lr_t);
        update_one<OPT>(g4[u].z * s, v4[u].z, m4[u].z, u4[u].z, a.h, lr_t);
        update_one<OPT>(g4[u].w * s, v4[u].w, m4[u].w, u4[u].w, a.h, lr_t);
      }
      }
      if (k + 1 < n_pending) {
#pragma unroll
        for (int u = 0; u < DRAIN_UNROLL; ++u) g4[u] = gn[u];
      }
    }
    // phase 3: write back, refresh the serving copy, and push the new values into the mirrors
#pragma unroll
    for (int u = 0; u < DRAIN_UNROLL; ++u) {
      const long long i = base + (static_cast<long long>(u) * DRAIN_THREADS + threadIdx.x) * DRAIN_VEC;
      if (i >= it.n) continue;
      if (full[u]) {
        *reinterpret_cast<float4*>(it.var + i) = v4[u];
        if (OPT != DTF_OPT_SGD) *reinterpret_cast<float4*>(it.slot0 + i) = m4[u];
        if (OPT == DTF_OPT_ADAM) *reinterpret_cast<float4*>(it.slot1 + i) = u4[u];
        const float v[4] = {v4[u].x, v4[u].y, v4[u].z, v4[u].w};
        if (shadow != nullptr) put_bf16x4(shadow, it.first + i, it.cols, it.shadow_ld, v, shadow_vec);
        if (n_mirror > 0) {   // owner pushes: posted stores into every worker's mirror
          const long long f = mi.first + i;
          if (mi.bf16_off >= 0) {
            for (int q = 0; q < n_mirror; ++q)
              put_bf16x4(reinterpret_cast<__nv_bfloat16*>(s_mbase[q] + mi.bf16_off), f, it.cols,
                         it.shadow_ld, v, mirror_vec);
          }
          if (mi.f32_off >= 0) {
            for (int q = 0; q < n_mirror; ++q) {
              float* dst = reinterpret_cast<float*>(s_mbase[q] + mi.f32_off) + f;
              if (mirror_f32_vec) {
                *reinterpret_cast<float4*>(dst) = v4[u];
              } else {
                dst[0] = v[0]; dst[1] = v[1]; dst[2] = v[2]; dst[3] = v[3];
              }
            }
          }
        }
      } else {
        // ragged tail / unaligned slice: element by element
        for (int j = 0; j < DRAIN_VEC && i + j < it.n; ++j) {
          float v = it.var[i + j];
          float s0 = (OPT != DTF_OPT_SGD) ? it.slot0[i + j] : 0.f;
          float s1 = (OPT == DTF_OPT_ADAM) ? it.slot1[i + j] : 0.f;
          if (a.mean) {
            float gs = 0.f;
            for (int k = 0; k < n_pending; ++k)
              gs += *reinterpret_cast<const volatile float*>(it.grad + s_off[k] + i + j);
            update_one<OPT>(gs * (a.h.grad_scale / static_cast<float>(n_pending)), v, s0, s1, a.h, s_lr[0]);
          } else {
            for (int k = 0; k < n_pending; ++k)
              update_one<OPT>(*reinterpret_cast<const volatile float*>(it.grad + s_off[k] + i + j) *
                                  a.h.grad_scale, v, s0, s1, a.h, s_lr[k]);
          }
          it.var[i + j] = v;
          if (OPT != DTF_OPT_SGD) it.slot0[i + j] = s0;
          if (OPT == DTF_OPT_ADAM) it.slot1[i + j] = s1;
          if (shadow != nullptr) {
            const long long f = it.first + i + j;
            shadow[(f / it.cols) * it.shadow_ld + (f % it.cols)] = __float2bfloat16(v);
          }
          const long long f = mi.first + i + j;
          for (int q = 0; q < n_mirror; ++q) {
            if (mi.bf16_off >= 0)
              reinterpret_cast<__nv_bfloat16*>(s_mbase[q] + mi.bf16_off)[(f / it.cols) * it.shadow_ld + (f % it.cols)] =
                  __float2bfloat16(v);
            if (mi.f32_off >= 0) reinterpret_cast<float*>(s_mbase[q] + mi.f32_off)[f] = v;
          }
        }
      }
    }
  }
  }  // chunk loop

  // ---- the last CTA retires the drained gradients ----
  // Fences: system scope costs microseconds, so the CTAs order their stores at GPU scope before the
  // election and the elected CTA issues its system-scope release stores from parallel threads (one
  // per worker): two fence latencies on the critical path instead of 3 + n_workers.
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned int old = atomicInc(a.block_counter, gridDim.x - 1u);
    s_last = (old == gridDim.x - 1u) ? 1u : 0u;
    if (s_last) __threadfence();  // acquire side of the election
  }
  __syncthreads();
  if (!s_last) return;
  const bool holds_lock = a.todo[DTF_MAX_WORKERS] != 0u;
  const unsigned int drop_mask = a.mean ? a.todo[DTF_MAX_WORKERS + 1] : 0u;   // stale: retire unapplied
  if (threadIdx.x == 0) {
    unsigned long long step = 0;
    if (n_pending > 0) {
      const int n_applies = a.mean ? 1 : n_pending;   // a synchronous round is ONE apply of the mean
      if (OPT == DTF_OPT_ADAM) {
        volatile float* b1 = &a.ctrl_powers->beta1_power;
        volatile float* b2 = &a.ctrl_powers->beta2_power;
        float p1 = *b1, p2 = *b2;
        for (int k = 0; k < n_applies; ++k) { p1 *= a.h.beta1; p2 *= a.h.beta2; }
        *b1 = p1;
        *b2 = p2;
      }
      if (a.ctrl_step != nullptr) {
        step = atom_add_u64_sys(&a.ctrl_step->global_step, static_cast<unsigned long long>(n_applies)) +
               static_cast<unsigned long long>(n_applies);
        atomicAdd(&a.ctrl_step->apply_count, static_cast<unsigned int>(n_applies));
      }
      if (a.mean)
        *reinterpret_cast<volatile unsigned int*>(&a.ctrl_powers->epoch) =
            *reinterpret_cast<volatile unsigned int*>(&a.ctrl_powers->epoch) + 1u;
      *reinterpret_cast<volatile unsigned int*>(&a.flags->drains) =
          *reinterpret_cast<volatile unsigned int*>(&a.flags->drains) + 1u;
      *reinterpret_cast<volatile unsigned int*>(&a.flags->applied_total) =
          *reinterpret_cast<volatile unsigned int*>(&a.flags->applied_total) +
          static_cast<unsigned int>(n_pending);
    }
    s_step = step;
    if (a.step_out != nullptr) {
      *a.step_out = (a.ctrl_step != nullptr && n_pending == 0)
                        ? ld_acquire_sys_u64(&a.ctrl_step->global_step)
                        : step;
    }
  }
  __syncthreads();
  // one thread per worker: retire its gradient (the mailbox buffer may be overwritten by worker w
  // from here on) and refresh its mirror's header - what the worker would otherwise read over NVLink
  if ((n_pending > 0 || drop_mask != 0u) && static_cast<int>(threadIdx.x) < DTF_MAX_WORKERS) {
    const int w = threadIdx.x;
    const bool drained = w < a.n_workers && (a.todo[w] != 0u || ((drop_mask >> w) & 1u) != 0u);
    unsigned int c = *reinterpret_cast<volatile unsigned int*>(&a.flags->consumed[w]);
    if (drained) {
      if (a.ctrl_step != nullptr && n_pending > 0)
        *reinterpret_cast<volatile unsigned long long*>(&a.ctrl_step->heartbeat[w]) = s_step;
      c += 1u;
      st_release_sys_u32(&a.flags->consumed[w], c);
    }
    const unsigned long long b =
        a.bcast ? *reinterpret_cast<volatile unsigned long long*>(&a.flags->mirror_base[w]) : 0ull;
    if (b != 0ull) {
      dtf_mirror_hdr* hdr = reinterpret_cast<dtf_mirror_hdr*>(b);
      *reinterpret_cast<volatile unsigned int*>(&hdr->consumed[a.shard]) = c;
      if (n_pending > 0) {
        if (a.ctrl_step != nullptr)
          *reinterpret_cast<volatile unsigned long long*>(&hdr->global_step) = s_step;
        // the release fence orders every CTA's data stores (GPU-scope fences + the election) first;
        // asynchronous: gradients applied so far, synchronous: rounds completed (the worker's token)
        st_release_sys_u32(&hdr->version[a.shard],
                           a.mean ? *reinterpret_cast<volatile unsigned int*>(&a.ctrl_powers->epoch)
                                  : *reinterpret_cast<volatile unsigned int*>(&a.flags->applied_total));
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0 && holds_lock) {
    __threadfence();
    atomicExch(&a.flags->lock, 0u);
  }
}

// One thread: wait for this worker's next token (see dtf_sync_token_wait).
__global__ void sync_token_wait_kernel(const dtf_mirror_hdr* hdr, int n_shards, unsigned int* last_round,
                                       unsigned int* local_step_out, unsigned long long* step_out) {
  griddep_wait();
  griddep_launch();
  if (threadIdx.x != 0) return;
  const unsigned int last = *reinterpret_cast<volatile unsigned int*>(last_round);
  const long long t0 = clock64();
  unsigned int e;
  for (;;) {
    e = 0xffffffffu;
    for (int s = 0; s < n_shards; ++s)
      e = min(e, *reinterpret_cast<const volatile unsigned int*>(&hdr->version[s]));
    if (last == 0xffffffffu || static_cast<int>(e - last) > 0) break;
    if (*reinterpret_cast<const volatile unsigned int*>(&hdr->abort) != 0u) break;
    __nanosleep(200);
    if (clock64() - t0 > 100ll * DTF_WAIT_TIMEOUT_CYCLES) {
      printf("dtf sync: no token after round %u (slowest shard has completed %u)\n", last, e);
      __trap();
    }
  }
  __threadfence();   // the parameters the owners stored before version[] are visible to the step
  *local_step_out = e;
  *last_round = e;
  if (step_out != nullptr)
    *step_out = *reinterpret_cast<const volatile unsigned long long*>(&hdr->global_step);
}

}  // namespace dtf

using namespace dtf;

extern "C" int dtf_mailbox_push(const dtf_mbox_push_item* items, int n_items, int depth,
                                int worker_id, dtf_mbox_flags* const* owner_flags, int n_owners,
                                int publish, unsigned int* seq, unsigned int* block_counter,
                                unsigned int* min_consumed_out, const dtf_ps_ctrl* ctrl_step,
                                unsigned long long* step_out, const dtf_result_sink* sink,
                                const dtf_mirror_hdr* my_hdr, const unsigned int* sync_tag,
                                void* stream) {
  DTF_REQUIRE(n_items >= 0 && n_items <= DTF_MAX_MBOX_ITEMS, "mailbox_push: %d items (max %d)",
              n_items, DTF_MAX_MBOX_ITEMS);
  DTF_REQUIRE(depth >= 1 && seq != nullptr, "mailbox_push: depth %d / seq", depth);
  DTF_REQUIRE(worker_id >= 0 && worker_id < DTF_MAX_WORKERS, "mailbox_push: worker %d", worker_id);
  DTF_REQUIRE(!publish || (n_owners >= 1 && n_owners <= DTF_MAX_WORKERS && owner_flags != nullptr &&
                           block_counter != nullptr),
              "mailbox_push: publish needs 1..%d owners and a block counter", DTF_MAX_WORKERS);
  MboxPushArgs a;
  memset(&a, 0, sizeof(a));
  int blocks = 0;
  for (int i = 0; i < n_items; ++i) {
    DTF_REQUIRE(items[i].bytes >= 0 && items[i].bytes % 16 == 0 && items[i].buf_stride % 16 == 0,
                "mailbox_push: item %d: bytes %lld / stride %lld not multiples of 16", i,
                items[i].bytes, items[i].buf_stride);
    DTF_REQUIRE(((reinterpret_cast<uintptr_t>(items[i].src) |
                  reinterpret_cast<uintptr_t>(items[i].dst0)) & 15) == 0,
                "mailbox_push: item %d is not 16-byte aligned", i);
    a.items[i] = items[i];
    a.block_start[i] = blocks;
    blocks += static_cast<int>((items[i].bytes + MBOX_CHUNK - 1) / MBOX_CHUNK);
  }
  a.block_start[n_items] = blocks;
  if (blocks == 0) {
    if (!publish) return 0;
    blocks = 1;
    a.block_start[n_items > 0 ? n_items : 1] = 1;
  }
  a.n_items = n_items > 0 ? n_items : 1;
  a.total_blocks = blocks;
  a.depth = depth;
  a.worker_id = worker_id;
  a.n_owners = publish ? n_owners : 0;
  for (int i = 0; i < a.n_owners; ++i) {
    DTF_REQUIRE(owner_flags[i] != nullptr, "mailbox_push: owner %d has no flag block", i);
    a.owner_flags[i] = owner_flags[i];
  }
  a.publish = publish;
  a.seq = seq;
  a.block_counter = block_counter;
  a.min_consumed_out = min_consumed_out;
  a.ctrl_step = ctrl_step;
  a.step_out = step_out;
  DTF_REQUIRE(sync_tag == nullptr || depth <= 2, "mailbox_push: tagged (synchronous) pushes need depth <= 2, got %d", depth);
  a.sync_tag = publish ? sync_tag : nullptr;
  a.my_hdr = publish ? my_hdr : nullptr;
  a.flow_hdr = nullptr;
  if (my_hdr != nullptr) {
    bool flow = n_items > 0;
    for (int i = 0; i < n_items; ++i) flow = flow && items[i].owner >= 0 && items[i].owner < DTF_MIRROR_MAX_SHARDS;
    if (flow) a.flow_hdr = my_hdr;
  }
  DTF_REQUIRE(my_hdr == nullptr || n_owners <= DTF_MIRROR_MAX_SHARDS,
              "mailbox_push: %d owners exceed the mirror header (%d)", n_owners, DTF_MIRROR_MAX_SHARDS);
  if (sink != nullptr && publish) {
    DTF_REQUIRE(sink->loss_src && sink->host_dst && sink->seq_dev, "mailbox_push: incomplete result sink");
    a.sink = *sink;
  }
  DTF_CHECK_CUDA(launch_kernel(mailbox_push_kernel, dim3(blocks), dim3(MBOX_THREADS), 0,
                               reinterpret_cast<cudaStream_t>(stream), a));
  return 0;
}

extern "C" int dtf_mailbox_drain(int opt, const dtf_apply_item* items, int n_items,
                                 long long slot_stride, int depth, int n_workers,
                                 const dtf_apply_hyper* hyper, dtf_ps_ctrl* ctrl_powers,
                                 dtf_ps_ctrl* ctrl_step, dtf_mbox_flags* flags, unsigned int* todo,
                                 unsigned int* block_counter, unsigned long long* step_out,
                                 void* stream_) {
  return dtf_mailbox_drain_ex(opt, items, n_items, slot_stride, depth, n_workers, hyper, ctrl_powers,
                              ctrl_step, flags, todo, block_counter, step_out, nullptr, 0, 0, stream_);
}

static int drain_impl(int opt, const dtf_apply_item* items, int n_items, long long slot_stride,
                      int depth, int n_workers, const dtf_apply_hyper* hyper, dtf_ps_ctrl* ctrl_powers,
                      dtf_ps_ctrl* ctrl_step, dtf_mbox_flags* flags, unsigned int* todo,
                      unsigned int* block_counter, unsigned long long* step_out,
                      const dtf_drain_mirror_item* mirror_items, int shard_index, int max_ctas,
                      int sync_R, void* stream_);

extern "C" int dtf_mailbox_drain_ex(int opt, const dtf_apply_item* items, int n_items,
                                    long long slot_stride, int depth, int n_workers,
                                    const dtf_apply_hyper* hyper, dtf_ps_ctrl* ctrl_powers,
                                    dtf_ps_ctrl* ctrl_step, dtf_mbox_flags* flags, unsigned int* todo,
                                    unsigned int* block_counter, unsigned long long* step_out,
                                    const dtf_drain_mirror_item* mirror_items, int shard_index,
                                    int max_ctas, void* stream_) {
  return drain_impl(opt, items, n_items, slot_stride, depth, n_workers, hyper, ctrl_powers, ctrl_step,
                    flags, todo, block_counter, step_out, mirror_items, shard_index, max_ctas, 0, stream_);
}

extern "C" int dtf_mailbox_drain_sync(int opt, const dtf_apply_item* items, int n_items,
                                      long long slot_stride, int n_workers, int replicas_to_aggregate,
                                      const dtf_apply_hyper* hyper, dtf_ps_ctrl* ctrl, dtf_ps_ctrl* ctrl_step,
                                      dtf_mbox_flags* flags, unsigned int* todo, unsigned int* block_counter,
                                      const dtf_drain_mirror_item* mirror_items, int shard_index,
                                      int max_ctas, void* stream_) {
  DTF_REQUIRE(replicas_to_aggregate >= 1 && ctrl != nullptr, "mailbox_drain_sync: needs replicas_to_aggregate >= 1 and the shard's control block");
  return drain_impl(opt, items, n_items, slot_stride, 2, n_workers, hyper, ctrl, ctrl_step, flags, todo,
                    block_counter, nullptr, mirror_items, shard_index, max_ctas, replicas_to_aggregate, stream_);
}

extern "C" int dtf_sync_token_wait(const dtf_mirror_hdr* my_hdr, int n_shards, unsigned int* last_round,
                                   unsigned int* local_step_out, unsigned long long* step_out, void* stream) {
  DTF_REQUIRE(my_hdr && last_round && local_step_out && n_shards >= 1 && n_shards <= DTF_MIRROR_MAX_SHARDS,
              "sync_token_wait: bad arguments (%d shards)", n_shards);
  DTF_CHECK_CUDA(launch_kernel(sync_token_wait_kernel, dim3(1), dim3(32), 0,
                               reinterpret_cast<cudaStream_t>(stream), my_hdr, n_shards, last_round,
                               local_step_out, step_out));
  return 0;
}

static int drain_impl(int opt, const dtf_apply_item* items, int n_items, long long slot_stride,
                      int depth, int n_workers, const dtf_apply_hyper* hyper, dtf_ps_ctrl* ctrl_powers,
                      dtf_ps_ctrl* ctrl_step, dtf_mbox_flags* flags, unsigned int* todo,
                      unsigned int* block_counter, unsigned long long* step_out,
                      const dtf_drain_mirror_item* mirror_items, int shard_index, int max_ctas,
                      int sync_R, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  DTF_REQUIRE(n_items >= 0 && n_items <= DTF_MAX_APPLY_ITEMS, "mailbox_drain: %d items (max %d)",
              n_items, DTF_MAX_APPLY_ITEMS);
  DTF_REQUIRE(depth >= 1 && n_workers >= 1 && n_workers <= DTF_MAX_WORKERS,
              "mailbox_drain: depth %d, %d workers", depth, n_workers);
  DTF_REQUIRE(hyper && flags && todo && block_counter, "mailbox_drain: null argument");
  DTF_REQUIRE(opt != DTF_OPT_ADAM || ctrl_powers != nullptr, "mailbox_drain: Adam needs the beta powers");
  DTF_REQUIRE(slot_stride % 4 == 0, "mailbox_drain: slot stride %lld", slot_stride);
  DTF_REQUIRE(mirror_items == nullptr || (shard_index >= 0 && shard_index < DTF_MIRROR_MAX_SHARDS),
              "mailbox_drain: shard index %d outside the mirror header", shard_index);
  DrainArgs a;
  memset(&a, 0, sizeof(a));
  int blocks = 0;
  for (int i = 0; i < n_items; ++i) {
    DTF_REQUIRE(items[i].n >= 0 && items[i].var && items[i].grad, "mailbox_drain: item %d invalid", i);
    DTF_REQUIRE(opt == DTF_OPT_SGD || items[i].slot0, "mailbox_drain: item %d lacks slot0", i);
    DTF_REQUIRE(opt != DTF_OPT_ADAM || items[i].slot1, "mailbox_drain: item %d lacks slot1", i);
    a.items[i] = items[i];
    if (mirror_items != nullptr) {
      a.mirror[i] = mirror_items[i];
    } else {
      a.mirror[i].bf16_off = a.mirror[i].f32_off = -1;
    }
    a.block_start[i] = blocks;
    blocks += static_cast<int>((items[i].n + DRAIN_CHUNK - 1) / DRAIN_CHUNK);
  }
  a.block_start[n_items] = blocks;
  if (blocks == 0) {  // a shard that only owns global_step still counts the pushes
    blocks = 1;
    if (n_items == 0) {
      n_items = 1;
      a.mirror[0].bf16_off = a.mirror[0].f32_off = -1;
      a.block_start[1] = 1;
    }
  }
  a.n_items = n_items;
  a.total_chunks = blocks;
  if (max_ctas > 0 && blocks > max_ctas) blocks = max_ctas;
  a.h = *hyper;
  a.ctrl_powers = ctrl_powers;
  a.ctrl_step = ctrl_step;
  a.flags = flags;
  a.todo = todo;
  a.slot_stride = slot_stride;
  a.depth = depth;
  a.n_workers = n_workers;
  a.bcast = mirror_items != nullptr ? 1 : 0;
  a.shard = shard_index;
  a.mean = sync_R > 0 ? 1 : 0;
  a.step_out = step_out;
  a.block_counter = block_counter;
  if (sync_R > 0) {
    DTF_CHECK_CUDA(launch_kernel(mailbox_snapshot_sync_kernel, dim3(1), dim3(32), 0, stream, flags,
                                 static_cast<const dtf_ps_ctrl*>(ctrl_powers), todo, n_workers, sync_R));
  } else {
    DTF_CHECK_CUDA(launch_kernel(mailbox_snapshot_kernel, dim3(1), dim3(32), 0, stream, flags, todo,
                                 n_workers, depth));
  }
  switch (opt) {
    case DTF_OPT_SGD: DTF_CHECK_CUDA(launch_kernel(mailbox_drain_kernel<DTF_OPT_SGD>, dim3(blocks), dim3(DRAIN_THREADS), 0, stream, a)); break;
    case DTF_OPT_MOMENTUM: DTF_CHECK_CUDA(launch_kernel(mailbox_drain_kernel<DTF_OPT_MOMENTUM>, dim3(blocks), dim3(DRAIN_THREADS), 0, stream, a)); break;
    case DTF_OPT_ADAGRAD: DTF_CHECK_CUDA(launch_kernel(mailbox_drain_kernel<DTF_OPT_ADAGRAD>, dim3(blocks), dim3(DRAIN_THREADS), 0, stream, a)); break;
    case DTF_OPT_ADAM: DTF_CHECK_CUDA(launch_kernel(mailbox_drain_kernel<DTF_OPT_ADAM>, dim3(blocks), dim3(DRAIN_THREADS), 0, stream, a)); break;
    default: DTF_REQUIRE(false, "mailbox_drain: unknown optimizer %d", opt);
  }
  DTF_LAUNCH_CHECK();
  return 0;
}
