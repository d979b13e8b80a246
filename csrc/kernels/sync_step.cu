// K12: one synchronous training round (SyncReplicasOptimizer semantics,
// templates/00_mnist_replica.py:174-191) as ONE cooperative kernel per rank over NVLink peer
// memory — no NCCL, no host round trips:
//
//   phase 1  reduce-scatter (push): every worker stores its gradient slice for shard s into
//            slot [worker] of shard s's mailbox (P2P 16-byte stores), then publishes
//            arrive[worker] = round on the shard with st.release.sys;
//   phase 2  the shard's owner waits for the round's arrivals, sums the slots in a FIXED order
//            (deterministic, unlike the accumulators' arrival order), takes the mean, applies the
//            TF1 optimizer rule to its fp32 master + slots, and all-gathers: it stores the updated
//            bf16 serving copy (and the fp32 values of vectors) into EVERY worker's mirror;
//            the owner of global_step bumps it once per round; then done[shard] = round on every
//            worker;
//   phase 3  every worker waits for done[*] == round: its mirror now holds the new parameters
//            (this is the token dequeue of the reference), and it reads the new global_step.
//
// A rank is both worker r and shard r (one process per GPU). All ranks launch the kernel for the
// same round; kernels on different GPUs wait on each other through flags in peer memory, every
// wait has a cycle watchdog. The ConditionalAccumulator's "mean of N fresh gradients, applied
// once" and the token barrier are exactly this; stale-drop cannot occur because nobody can start
// round t+1 before all shards finished round t (backup-worker mode stays on the host path).
#include "common.cuh"
#include "ptx.cuh"
#include "apply_rules.cuh"
#include "../../include/dtf_b200.h"

namespace dtf {

constexpr int SYNC_THREADS = 512;

struct SyncArgs {
  const dtf_sync_segment* segs;   // device array (identical on every rank)
  int n_segs;
  int rank, world;
  unsigned int round;             // 1, 2, 3, ... (flags are monotonic)
  int opt;
  dtf_apply_hyper h;
  unsigned long long arena[DTF_MAX_SHARDS];  // base of every rank's arena as mapped HERE
  unsigned int* grid_bar;         // 2 words of local scratch, zero-initialised
  unsigned long long* step_out;   // local u64: global_step after this round
  int step_shard;
};

__device__ __forceinline__ dtf_ps_ctrl* ctrl_of(const SyncArgs& a, int r) {
  return reinterpret_cast<dtf_ps_ctrl*>(a.arena[r]);
}

__device__ void grid_barrier(unsigned int* bar, unsigned int nblocks) {
  __syncthreads();
  if (threadIdx.x == 0) {
    volatile unsigned int* gen = bar + 1;
    const unsigned int my_gen = *gen;
    __threadfence();
    if (atomicAdd(bar, 1u) == nblocks - 1) {
      bar[0] = 0;
      __threadfence();
      atomicAdd(bar + 1, 1u);
    } else {
      long long t0 = clock64();
      while (*gen == my_gen) {
        if (clock64() - t0 > DTF_WAIT_TIMEOUT_CYCLES) {
          printf("dtf sync: grid barrier timed out (block %d)\n", blockIdx.x);
          __trap();
        }
      }
    }
    __threadfence();
  }
  __syncthreads();
}

__device__ void wait_flags(const unsigned int* flags, int n, unsigned int round, const char* what) {
  if (threadIdx.x < n) {
    long long t0 = clock64();
    while (ld_acquire_sys_u32(flags + threadIdx.x) < round) {
      if (clock64() - t0 > DTF_WAIT_TIMEOUT_CYCLES) {
        printf("dtf sync: waiting for %s[%d] >= %u timed out (block %d)\n", what, threadIdx.x, round,
               blockIdx.x);
        __trap();
      }
    }
  }
  __syncthreads();
}

template <int OPT>
__device__ __forceinline__ void sync_update(float g, float& var, float& s0, float& s1,
                                            const dtf_apply_hyper& h, float lr_t) {
  if (OPT == DTF_OPT_SGD) {
    var -= h.lr * g;
  } else if (OPT == DTF_OPT_MOMENTUM) {
    s0 = s0 * h.momentum + g;
    var -= h.use_nesterov ? (g * h.lr + s0 * h.momentum * h.lr) : h.lr * s0;
  } else if (OPT == DTF_OPT_ADAGRAD) {
    s0 += g * g;
    var -= div_rn(h.lr * g, sqrt_rn(s0));
  } else {
    s0 += (g - s0) * (1.f - h.beta1);
    s1 += (g * g - s1) * (1.f - h.beta2);
    var -= div_rn(lr_t * s0, sqrt_rn(s1) + h.epsilon);
  }
}

template <int OPT>
__global__ void __launch_bounds__(SYNC_THREADS, 1)
sync_round_kernel(const __grid_constant__ SyncArgs a) {
  const unsigned int nblocks = gridDim.x;
  const long long tid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long nthreads = static_cast<long long>(nblocks) * blockDim.x;
  dtf_ps_ctrl* my_ctrl = ctrl_of(a, a.rank);

  // ---------------- phase 1: push my gradient slices into the owners' mailboxes ----------------
  for (int si = 0; si < a.n_segs; ++si) {
    const dtf_sync_segment sg = a.segs[si];
    const float* src = reinterpret_cast<const float*>(sg.grad_local);
    float* dst = reinterpret_cast<float*>(a.arena[sg.owner] + sg.mailbox_off +
                                          static_cast<unsigned long long>(a.rank) * sg.mailbox_stride);
    const long long nv = sg.n / 4;
    for (long long i = tid; i < nv; i += nthreads)
      reinterpret_cast<float4*>(dst)[i] = reinterpret_cast<const float4*>(src)[i];
    for (long long i = nv * 4 + tid; i < sg.n; i += nthreads) dst[i] = src[i];
  }
  __threadfence_system();
  grid_barrier(a.grid_bar, nblocks);
  if (blockIdx.x == 0 && threadIdx.x < a.world)
    st_release_sys_u32(&ctrl_of(a, threadIdx.x)->arrive[a.rank], a.round);

  // ---------------- phase 2: owner reduces, applies, all-gathers ---------------------------------
  wait_flags(my_ctrl->arrive, a.world, a.round, "arrive");
  float lr_t = a.h.lr;
  if (OPT == DTF_OPT_ADAM) {
    const float b1p = *reinterpret_cast<volatile float*>(&my_ctrl->beta1_power);
    const float b2p = *reinterpret_cast<volatile float*>(&my_ctrl->beta2_power);
    lr_t = a.h.lr * sqrtf(1.f - b2p) / (1.f - b1p);
  }
  const float inv_w = 1.f / static_cast<float>(a.world);
  for (int si = 0; si < a.n_segs; ++si) {
    const dtf_sync_segment sg = a.segs[si];
    if (sg.owner != a.rank) continue;
    const unsigned long long base = a.arena[a.rank];
    float* var = reinterpret_cast<float*>(base + sg.var_off);
    float* s0p = sg.slot0_off ? reinterpret_cast<float*>(base + sg.slot0_off) : nullptr;
    float* s1p = sg.slot1_off ? reinterpret_cast<float*>(base + sg.slot1_off) : nullptr;
    __nv_bfloat16* shadow = sg.shadow_off ? reinterpret_cast<__nv_bfloat16*>(base + sg.shadow_off) : nullptr;
    // 4 elements per thread (16-byte mailbox / state accesses, 8-byte bf16 stores) when the slice
    // is 4-aligned in every view; scalar otherwise and for the tail
    const bool vec = (sg.n % 4 == 0) && (sg.cols % 4 == 0) && (sg.first % 4 == 0) && (sg.shadow_ld % 4 == 0);
    const long long nvec = vec ? sg.n / 4 : 0;
    for (long long q = tid; q < nvec; q += nthreads) {
      const long long i = q * 4;
      float4 g4 = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int w = 0; w < a.world; ++w) {   // fixed order: bitwise reproducible
        const float4 t = __ldcg(reinterpret_cast<const float4*>(
            base + sg.mailbox_off + static_cast<unsigned long long>(w) * sg.mailbox_stride) + q);
        g4.x += t.x; g4.y += t.y; g4.z += t.z; g4.w += t.w;
      }
      const float sc = inv_w * a.h.grad_scale;
      float g[4] = {g4.x * sc, g4.y * sc, g4.z * sc, g4.w * sc};
      float4 v4 = *reinterpret_cast<float4*>(var + i);
      float4 m4 = s0p ? *reinterpret_cast<float4*>(s0p + i) : make_float4(0.f, 0.f, 0.f, 0.f);
      float4 u4 = s1p ? *reinterpret_cast<float4*>(s1p + i) : make_float4(0.f, 0.f, 0.f, 0.f);
      float v[4] = {v4.x, v4.y, v4.z, v4.w}, s0[4] = {m4.x, m4.y, m4.z, m4.w}, s1[4] = {u4.x, u4.y, u4.z, u4.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) sync_update<OPT>(g[j], v[j], s0[j], s1[j], a.h, lr_t);
      const float4 vo = make_float4(v[0], v[1], v[2], v[3]);
      *reinterpret_cast<float4*>(var + i) = vo;
      if (s0p) *reinterpret_cast<float4*>(s0p + i) = make_float4(s0[0], s0[1], s0[2], s0[3]);
      if (s1p) *reinterpret_cast<float4*>(s1p + i) = make_float4(s1[0], s1[1], s1[2], s1[3]);
      const long long row = i / sg.cols, col = i % sg.cols;
      uint2 vb;
      vb.x = pack_bf16x2(v[0], v[1]);
      vb.y = pack_bf16x2(v[2], v[3]);
      if (shadow) *reinterpret_cast<uint2*>(shadow + row * sg.shadow_ld + col) = vb;
      for (int w = 0; w < a.world; ++w) {   // all-gather into every worker's mirror
        if (sg.mirror_f32_off)
          *reinterpret_cast<float4*>(reinterpret_cast<float*>(a.arena[w] + sg.mirror_f32_off) + sg.first + i) = vo;
        if (sg.mirror_bf16_off)
          *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(a.arena[w] + sg.mirror_bf16_off) +
                                    (sg.first_row + row) * sg.shadow_ld + col) = vb;
      }
    }
    for (long long i = nvec * 4 + tid; i < sg.n; i += nthreads) {
      float g = 0.f;
      for (int w = 0; w < a.world; ++w)   // fixed order: bitwise reproducible
        g += __ldcg(reinterpret_cast<const float*>(base + sg.mailbox_off +
                                                   static_cast<unsigned long long>(w) * sg.mailbox_stride) + i);
      g *= inv_w * a.h.grad_scale;
      float v = var[i];
      float s0 = s0p ? s0p[i] : 0.f, s1 = s1p ? s1p[i] : 0.f;
      sync_update<OPT>(g, v, s0, s1, a.h, lr_t);
      var[i] = v;
      if (s0p) s0p[i] = s0;
      if (s1p) s1p[i] = s1;
      const long long row = i / sg.cols, col = i % sg.cols;
      const __nv_bfloat16 vb = __float2bfloat16(v);
      if (shadow) shadow[row * sg.shadow_ld + col] = vb;
      // all-gather: updated value into every worker's mirror (own included)
      for (int w = 0; w < a.world; ++w) {
        if (sg.mirror_f32_off)
          reinterpret_cast<float*>(a.arena[w] + sg.mirror_f32_off)[sg.first + i] = v;
        if (sg.mirror_bf16_off)
          reinterpret_cast<__nv_bfloat16*>(a.arena[w] + sg.mirror_bf16_off)[(sg.first_row + row) * sg.shadow_ld + col] = vb;
      }
    }
  }
  __threadfence_system();
  grid_barrier(a.grid_bar, nblocks);
  if (blockIdx.x == 0) {
    if (threadIdx.x == 0) {
      if (OPT == DTF_OPT_ADAM) {
        my_ctrl->beta1_power *= a.h.beta1;
        my_ctrl->beta2_power *= a.h.beta2;
      }
      my_ctrl->epoch = a.round;
      if (a.rank == a.step_shard) atom_add_u64_sys(&my_ctrl->global_step, 1ull);
      __threadfence_system();
    }
    __syncthreads();
    if (threadIdx.x < a.world) st_release_sys_u32(&ctrl_of(a, threadIdx.x)->done[a.rank], a.round);
  }

  // ---------------- phase 3: wait until every shard has delivered --------------------------------
  wait_flags(my_ctrl->done, a.world, a.round, "done");
  if (blockIdx.x == 0 && threadIdx.x == 0 && a.step_out != nullptr)
    *a.step_out = ld_acquire_sys_u64(&ctrl_of(a, a.step_shard)->global_step);
}

}  // namespace dtf

using namespace dtf;

extern "C" int dtf_sync_round(int opt, const dtf_sync_segment* segs_dev, int n_segs, int rank, int world,
                              unsigned int round, const dtf_apply_hyper* hyper,
                              const unsigned long long* arena_bases, unsigned int* grid_bar,
                              unsigned long long* step_out, int step_shard, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  DTF_REQUIRE(world >= 1 && world <= DTF_MAX_SHARDS && rank >= 0 && rank < world,
              "sync_round: rank %d / world %d out of range (max %d)", rank, world, DTF_MAX_SHARDS);
  DTF_REQUIRE(round >= 1 && segs_dev != nullptr && grid_bar != nullptr, "sync_round: bad arguments");
  SyncArgs a;
  memset(&a, 0, sizeof(a));
  a.segs = segs_dev;
  a.n_segs = n_segs;
  a.rank = rank;
  a.world = world;
  a.round = round;
  a.opt = opt;
  a.h = *hyper;
  for (int i = 0; i < world; ++i) a.arena[i] = arena_bases[i];
  a.grid_bar = grid_bar;
  a.step_out = step_out;
  a.step_shard = step_shard;

  void (*kern)(SyncArgs) = nullptr;
  switch (opt) {
    case DTF_OPT_SGD: kern = sync_round_kernel<DTF_OPT_SGD>; break;
    case DTF_OPT_MOMENTUM: kern = sync_round_kernel<DTF_OPT_MOMENTUM>; break;
    case DTF_OPT_ADAGRAD: kern = sync_round_kernel<DTF_OPT_ADAGRAD>; break;
    case DTF_OPT_ADAM: kern = sync_round_kernel<DTF_OPT_ADAM>; break;
    default: DTF_REQUIRE(false, "sync_round: unknown optimizer %d", opt);
  }
  int dev = 0, sms = 0, per_sm = 0;
  DTF_CHECK_CUDA(cudaGetDevice(&dev));
  DTF_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  DTF_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, SYNC_THREADS, 0));
  DTF_REQUIRE(per_sm >= 1, "sync_round: kernel does not fit on an SM");
  // all CTAs must be co-resident (they meet at a grid barrier): cooperative launch, one CTA per SM
  void* args[] = {&a};
  DTF_CHECK_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(kern), dim3(sms), dim3(SYNC_THREADS),
                                             args, 0, stream));
  return 0;
}
