/* C ABI of libdtf_b200_cuda.so — the hand-written sm_100a kernels and the peer-memory runtime of
 * the B200-native parameter-server framework. Every entry point returns 0 on success and a
 * negative value on failure (dtf_last_error() gives the reason). Pointers are device pointers
 * unless stated; `stream` is a cudaStream_t passed as void*. Nothing here links against PyTorch;
 * the Python package binds these with ctypes (distributedtensorflow_b200/ops/_lib.py).
 *
 * Reference citations are file:line under the upstream reference tree. */
#ifndef DTF_B200_H_
#define DTF_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* dtf_last_error(void);
int dtf_abi_version(void);

/* ------------------------------------------------------------------------------------------
 * PS shard control block. Lives at the start of every shard arena (peer-mapped into every
 * worker). Replaces the TF resource-manager scalars: global_step
 * (run_mnist_distributed.py:114), Adam's beta1_power/beta2_power (:116), the "variables
 * initialised" readiness poll of MonitoredTrainingSession (:128-132) and SyncReplicasOptimizer's
 * token queue (templates/00_mnist_replica.py:174-191).
 * ------------------------------------------------------------------------------------------ */
#define DTF_MAX_WORKERS 16
#define DTF_MAX_SHARDS 8
typedef struct dtf_ps_ctrl {
  unsigned long long global_step;               /* +0   shared step counter (device atomics) */
  float beta1_power;                            /* +8   */
  float beta2_power;                            /* +12  */
  unsigned int initialized;                     /* +16  chief sets 1 after init / restore */
  unsigned int epoch;                           /* +20  sync mode: completed rounds */
  unsigned int shutdown;                        /* +24  */
  unsigned int apply_count;                     /* +28  applies seen by this shard */
  unsigned long long heartbeat[DTF_MAX_WORKERS];/* +32  last global step each worker published */
  unsigned int arrive[DTF_MAX_WORKERS];         /* +160 sync mode: worker w pushed round r */
  unsigned int done[DTF_MAX_SHARDS];            /* +224 sync mode: shard s finished round r */
} dtf_ps_ctrl;                                  /* 256 bytes */

/* ------------------------------------------------------------------------------------------
 * K1/K2/K3: tcgen05 GEMM  D[M,N] = A * B^T, bf16 in, fp32 accumulate in TMEM.
 *   a_mn = 0: A is stored [M, K] (K contiguous), 1: stored [K, M] (M contiguous)
 *   b_mn = 0: B is stored [N, K] (K contiguous), 1: stored [K, N] (N contiguous)
 *   lda/ldb: row pitch in elements (multiple of 8).
 * (MatMul / BiasAdd / Relu / ReluGrad of run_mnist_distributed.py:67-69 and
 *  templates/00_mnist_replica.py:160-163, plus their autodiff.)
 * ------------------------------------------------------------------------------------------ */
typedef struct dtf_gemm_epilogue {
  void* out;          /* [M, ldo] bf16 (out_f32 = 0) or fp32 (out_f32 = 1); may be peer memory */
  int ldo;            /* row pitch of out in elements */
  int out_f32;
  const float* bias;  /* [N] or NULL: out = acc + bias */
  int relu;           /* out = max(out, 0) */
  const void* mask;   /* bf16 [M, ldmask] or NULL: out = (mask > 0) ? out : 0  (ReLU gradient) */
  int ldmask;
  int reduce;         /* 1: red.global.add.f32 into out (split-K / gradient accumulation) */
  float alpha;        /* out = alpha * acc (applied first) */
  /* split-K with in-kernel fix-up (splits > 1 and reduce = 0): K-slices sum into this fp32
   * workspace and the last slice of each tile runs the epilogue above. Both must be zero on
   * entry; the counters reset themselves, the workspace does not. */
  float* splitk_ws;             /* [M, ldws] fp32, 16-byte aligned, ldws % 4 == 0, ldws >= N; zero on
                                   entry (the caller re-zeroes it; DTF_GEMM_WS_CLEAN=1 makes the
                                   fix-up do it, which measured slower) */
  int ldws;
  unsigned int* tile_counters;  /* ceil(M/128) * ceil(N/block_n) words */
  /* Fused gradient push + optimizer apply (weight-gradient GEMMs, K2 + K9 in one kernel): when
   * apply_enable != 0 the accumulator element (m, n) is the gradient of element m*N + n of a
   * variable [M, N] and the epilogue applies the TF1 update rule to it in place on the owning
   * shard(s) - local or NVLink peer pointers - instead of storing `out` (which is ignored).
   * The variable may be cut into row slices of apply_rows_per_slice rows, slice i living at
   * apply_var[i] etc. (slice-local indexing). Not combinable with split-K. */
  int apply_enable;             /* 0 = off (a zero-initialised struct is a plain GEMM) */
  int apply_opt;                /* DTF_OPT_* */
  int apply_rows_per_slice;
  int apply_shadow_ld;          /* row pitch of the bf16 serving copy, elements */
  float apply_lr, apply_beta1, apply_beta2, apply_epsilon, apply_momentum, apply_grad_scale;
  int apply_use_nesterov;
  int apply_reserved2;
  const struct dtf_ps_ctrl* apply_ctrl_powers;   /* Adam: control block with the beta powers */
  float* apply_var[8];
  float* apply_slot0[8];
  float* apply_slot1[8];
  void* apply_shadow[8];        /* bf16 serving copy per slice, or NULL */
  /* Scatter store (weight-gradient GEMM -> the owners' gradient MAILBOXES, K2 + K11 in one kernel):
   * when scatter_enable != 0 row m of the [M, N] fp32 result belongs to row slice
   * m / scatter_rows_per_slice and is stored at
   *   scatter_base[slice] + (*scatter_seq % scatter_depth) * scatter_buf_stride[slice] (bytes)
   *                       + ((m - slice * rows_per_slice) * N + n) * 4
   * i.e. straight into this worker's current mailbox buffer on the shard that owns the slice (local or
   * NVLink peer memory; posted 16-byte stores, issued tile by tile while the GEMM is still running).
   * `out` is ignored; no gradient buffer and no copy kernel. The step's publishing
   * dtf_mailbox_push (which advances *scatter_seq) must come after this GEMM in stream order. */
  int scatter_enable;
  int scatter_rows_per_slice;
  int scatter_depth;
  int scatter_reserved;
  const unsigned int* scatter_seq;
  float* scatter_base[8];
  long long scatter_buf_stride[8];
  /* Max-pool + ReLU gradient scatter (K2 + K8 in one kernel, the dX GEMM in front of conv2's
   * backward, run_mnist_distributed.py:59-67): with pool_out != NULL the [M, 49*64] result is the
   * gradient of the 7x7x64 pooled map (NHWC). Each element is masked by `mask` (the pooled
   * activation: ReLU gradient) and stored at the position of its 2x2 window that won the forward max
   * (pool_amax u8 [M,7,7,64], value dy*2+dx), zeros at the other three, straight into the
   * zero-bordered bf16 [M,18,18,64] gradient of conv2's output; pool_dbias[64] (fp32, accumulated)
   * receives conv2's bias gradient. `out` is ignored. Works with split-K (fix-up). */
  void* pool_out;
  const unsigned char* pool_amax;
  float* pool_dbias;
} dtf_gemm_epilogue;

int dtf_gemm_bf16(const void* A, int lda, int a_mn, const void* B, int ldb, int b_mn, int M,
                  int N, int K, const dtf_gemm_epilogue* epi, int splits, int block_n,
                  void* stream);

/* ------------------------------------------------------------------------------------------
 * K4/K5: softmax cross-entropy, loss and dlogits in one pass.
 *   sparse: loss = mean_b(lse - logit[label])          (run_mnist_distributed.py:113)
 *   dense : loss = -sum(y * log(clip(softmax,1e-10,1))) (templates/00_mnist_replica.py:163-164)
 * loss_accum (fp32 scalar) and dbias_accum ([C] fp32, may be NULL) are ACCUMULATED with atomics:
 * zero them before the call. dlogits is bf16 [B, ldd] (columns C..ldd-1 are written as 0).
 * ------------------------------------------------------------------------------------------ */
int dtf_softmax_xent_sparse(const float* logits, int ldl, const int* labels, int B, int C,
                            float* loss_accum, void* dlogits, int ldd, float* dbias_accum,
                            void* stream);
int dtf_softmax_xent_dense(const float* logits, int ldl, const float* labels, int ldy, int B,
                           int C, float* loss_accum, void* dlogits, int ldd, float* dbias_accum,
                           void* stream);

/* Fused classifier head for <= 16 classes (both reference models end in a 10-way layer,
 * run_mnist_distributed.py:69, templates/00_mnist_replica.py:163): logits = act W + b, the
 * softmax cross-entropy (sparse/mean or dense/clipped/sum), dlogits, the logits-bias gradient, and
 * — when dact is given — the previous layer's activation gradient (dlogits W^T masked by act > 0)
 * plus its bias gradient, in ONE launch instead of GEMM + xent + GEMM + column-sum. w_shadow is
 * the [H, 16] bf16 serving copy. Pass dlogits = NULL for inference (logits + loss only). */
int dtf_head_fused(const void* act, int lda, int H, const void* w_shadow, int ldw, const float* bias,
                   const int* labels_sparse, const float* labels_dense, int ldy, int B, int C,
                   float* logits, int ldl, float* loss_accum, void* dlogits, int ldd,
                   float* dbias_accum, void* dact, int lddact, float* dprev_bias_accum,
                   float* dw_accum /* [H, C] fp32, zero on entry: += act^T dlogits; or NULL */, void* stream);

/* ------------------------------------------------------------------------------------------
 * K9/K10: gradient push fused with the owning shard's optimizer apply. One launch walks a list
 * of (variable slice, gradient) pairs; var/slots/shadow may be peer (NVLink) pointers. The last
 * CTA to finish performs the "_finish" step: beta powers, heartbeat and the global_step device
 * atomic. use_locking=False semantics (Hogwild): concurrent workers interleave element-wise.
 * TF1 formulas (python/training/{adam,adagrad,momentum,gradient_descent}.py via
 * run_mnist_distributed.py:116, templates/00_between_graph_replication_async_mnist.py:34).
 * ------------------------------------------------------------------------------------------ */
enum { DTF_OPT_SGD = 0, DTF_OPT_MOMENTUM = 1, DTF_OPT_ADAGRAD = 2, DTF_OPT_ADAM = 3 };
#define DTF_MAX_APPLY_ITEMS 24
typedef struct dtf_apply_item {
  float* var;          /* fp32 master slice on the owning shard */
  float* slot0;        /* m / accum / momentum (NULL for SGD) */
  float* slot1;        /* v (Adam only) */
  void* shadow;        /* bf16 serving copy (row-padded), or NULL */
  const float* grad;   /* local fp32 gradient slice */
  long long n;         /* elements in this slice */
  long long first;     /* flat index of the slice's first element inside the variable */
  int cols;            /* variable's innermost dim (for the padded shadow layout) */
  int shadow_ld;       /* shadow row pitch in elements */
  int zero_grad;       /* dtf_push_apply: store 0 back into grad after reading it (the gradient is an
                          atomically accumulated buffer that must be clear for the next step) */
  int reserved;
} dtf_apply_item;
typedef struct dtf_apply_hyper {
  float lr, beta1, beta2, epsilon, momentum, grad_scale;
  int use_nesterov;
  int atomic_sgd;      /* SGD only: red.global.add.f32 of -lr*g (exact under concurrency) */
} dtf_apply_hyper;
/* ctrl_for_powers: control block holding beta powers (shard of the first variable).
 * ctrl_for_step: control block holding global_step. finish: 1 = bump global_step, update beta
 * powers, heartbeat[worker_id]; 2 = update beta powers only (after dtf_step_report). step_out (device u64, local) receives the new global step.
 * block_counter: a zero-initialised local u32 the kernel uses to elect the last CTA. */
/* Where the step's LAST kernel reports the step's results, so no D2H copy nodes trail the step:
 * the finishing thread stores {loss bits, sequence number, global_step} straight into a pinned,
 * device-mapped host buffer (16-byte aligned): word 0 = *loss_src, words 2..3 = global_step (u64),
 * then — after a system fence — word 1 = ++(*seq_dev). The host spins on word 1. */
typedef struct dtf_result_sink {
  const float* loss_src;     /* device: the step's accumulated loss */
  unsigned int* host_dst;    /* pinned host memory, 4 words */
  unsigned int* seq_dev;     /* device counter of steps reported */
  int clear_loss;            /* dtf_push_apply_ex: reset *loss_src to 0 after reading it */
  int reserved;
} dtf_result_sink;
int dtf_push_apply(int opt, const dtf_apply_item* items, int n_items,
                   const dtf_apply_hyper* hyper, dtf_ps_ctrl* ctrl_for_powers,
                   dtf_ps_ctrl* ctrl_for_step, int finish, int worker_id,
                   unsigned long long* step_out, unsigned int* block_counter, void* stream);
/* Same, and the finishing thread also fills `sink` (may be NULL). max_ctas > 0 caps the grid (the
 * kernel strides over its 4096-element chunks): an apply that overlaps compute kernels then holds
 * one CTA slot per SM instead of flooding the SMs. */
int dtf_push_apply_ex(int opt, const dtf_apply_item* items, int n_items,
                      const dtf_apply_hyper* hyper, dtf_ps_ctrl* ctrl_for_powers,
                      dtf_ps_ctrl* ctrl_for_step, int finish, int worker_id,
                      unsigned long long* step_out, unsigned int* block_counter,
                      const dtf_result_sink* sink, int max_ctas, void* stream);

/* Test hook: q_fast = a / b and s_fast = sqrt(b) as the optimizer rules compute them (branch-free,
 * apply_rules.cuh), q_ieee / s_ieee as IEEE round-to-nearest; n elements each. */
int dtf_debug_div_sqrt(const float* a, const float* b, float* q_fast, float* q_ieee, float* s_fast,
                       float* s_ieee, long long n, void* stream);

/* Early step report: bumps (bump = 1) or just reads (bump = 0) global_step and fills `sink` as soon
 * as the loss is known, long before the step's applies have finished, so the host overlaps its
 * per-step work with the rest of the backward pass. The step's last dtf_push_apply is then called
 * with finish = 2 (beta powers only). The next step is stream-ordered behind this one, and host
 * reads of shard memory must synchronise the stream first. */
int dtf_step_report(dtf_ps_ctrl* ctrl_for_step, int worker_id, int bump, unsigned long long* step_out,
                    const dtf_result_sink* sink, void* stream);

/* dtf_push_apply for callers that cannot allocate device memory (no counterpart in the reference:
 * TensorFlow's ApplyAdam kernels allocate through the TF runtime). Same arguments and result as
 * dtf_push_apply minus `block_counter`: the CTA-election counter lives in a 256-byte __device__
 * array inside this library, resident on every GPU that loads it until the process exits.
 * Meant for a PS task whose whole HBM budget is already committed to its shard arena, for a host
 * thread that must not enter cudaMalloc (a device-wide synchronisation) while worker kernels are
 * in flight, and for foreign runtimes that reach this library only through the C ABI.
 *   - Every call looks the array's address up for the current device and context (nothing is
 *     remembered across calls, so cudaDeviceReset or a caller-owned context cannot leave a stale
 *     pointer behind). The FIRST call on a device also loads the module: issue it before any
 *     stream capture on that device.
 *   - Unlike dtf_push_apply, a call with n_items == 0 (a finish-only call) is refused.
 *   - The counter is cleared on `stream` by every call; the array otherwise keeps the previous
 *     call's contents.
 *   - All calls on one GPU share the array: calls whose kernels overlap there race. Keep them
 *     ordered (one stream, or events between streams).
 *   - More than DTF_MAX_APPLY_ITEMS slices or DTF_PUSH_APPLY_STATIC_MAX_ELEMS elements per call
 *     are refused with an error; split the call or use dtf_push_apply.
 * It never calls cudaMalloc / cudaFree or an allocating library routine. */
#define DTF_PUSH_APPLY_STATIC_SCRATCH_BYTES 256
#define DTF_PUSH_APPLY_STATIC_MAX_ELEMS (1LL << 28)   /* 268 435 456 elements (1 GiB of fp32) */
int dtf_push_apply_static(int opt, const dtf_apply_item* items, int n_items,
                          const dtf_apply_hyper* hyper, dtf_ps_ctrl* ctrl_for_powers,
                          dtf_ps_ctrl* ctrl_for_step, int finish, int worker_id,
                          unsigned long long* step_out, void* stream);

/* ------------------------------------------------------------------------------------------
 * K9/K11, asynchronous "owner applies" push (option (b)). The remote read-modify-write
 * of dtf_push_apply moves 26 bytes per Adam parameter over NVLink, half of them as round-trip
 * reads. Here the worker only WRITES its fp32 gradient into a per-worker mailbox inside the
 * owning shard's arena (4 bytes per parameter, posted stores) and publishes a sequence number;
 * the owner's own step stream drains its mailboxes with local HBM traffic, applying each pending
 * gradient separately and in order (worker 0..W-1), exactly as the reference's PS applies each
 * worker's ApplyAdam when it arrives (run_mnist_distributed.py:116,143-150). Mailboxes are
 * `depth` buffers deep per worker: a worker may run at most `depth` pushes ahead of the owner,
 * which is the bounded-staleness knob TF1 lacks.
 *   arena:  flags (dtf_mbox_flags) ... slot(w, b) = slots + (w * depth + b) * slot_stride
 * ------------------------------------------------------------------------------------------ */
typedef struct dtf_mbox_flags {               /* 512 bytes, zero = idle */
  unsigned int produced[DTF_MAX_WORKERS];     /* +0   worker w: gradients published so far (st.release.sys) */
  unsigned int tag0[DTF_MAX_WORKERS];         /* +64  sync mode: local_step of the gradient in worker w's buffer 0 */
  unsigned int consumed[DTF_MAX_WORKERS];     /* +128 owner: gradients of worker w applied (or dropped) so far */
  unsigned int tag1[DTF_MAX_WORKERS];         /* +192 sync mode: ... in buffer 1 (synchronous mailboxes are 2 deep) */
  unsigned int done[DTF_MAX_WORKERS];         /* +256 worker w has finished training */
  unsigned int lock;                          /* +320 a drain is in flight (device-scope CAS: every drain of a
                                                      shard runs on the shard's GPU, whoever launches it) */
  unsigned int paused;                        /* +324 owner: drains are no-ops while set (checkpoint cut) */
  unsigned int drains;                        /* +328 drains that applied at least one gradient */
  unsigned int applied_total;                 /* +332 gradients applied so far (all workers) */
  unsigned int pad2[12];
  unsigned long long mirror_base[DTF_MAX_WORKERS]; /* +384 address of worker w's mirror IN THE OWNER'S PROCESS
                                                      (0 = none): the drain broadcasts fresh parameters there */
} dtf_mbox_flags;

/* ------------------------------------------------------------------------------------------
 * Worker mirrors ("owner pushes, worker never pulls"). The reference fetches every parameter
 * from the PS at the start of every mon_sess.run (run_mnist_distributed.py:150,161) - an NVLink
 * READ round trip at the head of the step. Here every worker owns a mirror of the parameters in
 * its own HBM (bf16 row-padded copies of the matrices its tensor-core kernels read, fp32 copies
 * of the small tensors); the owner's drain, which has just computed the new values, also stores
 * them into every registered mirror with posted 16-byte writes. The worker's step reads only
 * local memory. Reads are unlocked, like TF1's use_locking=False: a step may see a slice that a
 * drain is half-way through rewriting. The header carries what the worker otherwise had to read
 * over NVLink: per owner the count of its own pushes applied (flow control), a version, and the
 * shared global_step.
 * ------------------------------------------------------------------------------------------ */
#define DTF_MIRROR_MAX_SHARDS 32
typedef struct dtf_mirror_hdr {               /* 512 bytes at offset 0 of every worker mirror */
  unsigned long long global_step;             /* +0   last value broadcast by the step-owning shard */
  unsigned int epoch;                         /* +8   sync mode: completed rounds */
  unsigned int abort;                         /* +12  set by the worker's host: kernels of this worker that
                                                      wait on this header give up (teardown / recovery) */
  unsigned int pad0[28];
  unsigned int version[DTF_MIRROR_MAX_SHARDS];   /* +128 shard s: gradients applied (all workers) when it
                                                         last rewrote this mirror; stored after the data */
  unsigned int consumed[DTF_MIRROR_MAX_SHARDS];  /* +256 shard s: how many of THIS worker's pushes it applied */
  unsigned int pad1[32];
} dtf_mirror_hdr;
/* Where a drained slice goes inside every mirror (byte offsets from the mirror base, -1 = the
 * mirror does not hold that form) and the slice's first element inside its variable. */
typedef struct dtf_drain_mirror_item {
  long long bf16_off;     /* bf16 [rows, shadow_ld] copy of the WHOLE variable */
  long long f32_off;      /* fp32 flat copy of the WHOLE variable */
  long long first;        /* flat index of the slice's first element inside the variable */
} dtf_drain_mirror_item;

#define DTF_MAX_MBOX_ITEMS 64
typedef struct dtf_mbox_push_item {
  const void* src;        /* local fp32 gradient slice */
  void* dst0;             /* the slice's place inside MY buffer 0 on the owning shard (peer pointer) */
  long long bytes;        /* multiple of 16 */
  long long buf_stride;   /* bytes from buffer b to buffer b + 1 of the same worker on that shard */
  int zero_src;           /* store 0 back into src after reading it (src is an atomically accumulated
                             buffer that must be clear for the next step: no fill kernel trails the step) */
  int owner;              /* shard that owns the slice (index into dtf_mirror_hdr.consumed), or -1 */
} dtf_mbox_push_item;
/* Copies every item into buffer (*seq % depth) of the caller's mailboxes. With publish = 1 the
 * last CTA then stores produced[worker_id] = *seq + 1 on EVERY owner (owner_flags[0..n_owners)),
 * advances *seq, and writes min over owners of consumed[worker_id] to *min_consumed_out - the
 * host's flow control: push k may only be issued once every owner has consumed push k - depth.
 * With ctrl_step / step_out it also reads the shared global_step into *step_out (local u64).
 * my_hdr (optional): this worker's mirror header; consumed[] and global_step are then read from
 * it (local memory, kept fresh by the owners' drains; owner i of owner_flags is shard i) instead
 * of over NVLink. When every item also names its owner (item.owner >= 0) flow control moves onto
 * the device: before a CTA overwrites buffer (*seq % depth) it waits until
 * my_hdr->consumed[owner] >= *seq - depth + 1 (watchdog-guarded spin on local memory), so the host
 * may enqueue pushes without waiting. sync_tag (optional, device word; needs depth <= 2): the
 * publishing CTA first stores *sync_tag - the local_step this gradient was computed at - into
 * tag{buffer}[worker_id] of every owner (see dtf_mailbox_drain_sync). Several non-publishing
 * calls followed by one publishing call form one push. */
int dtf_mailbox_push(const dtf_mbox_push_item* items, int n_items, int depth, int worker_id,
                     dtf_mbox_flags* const* owner_flags, int n_owners, int publish,
                     unsigned int* seq, unsigned int* block_counter,
                     unsigned int* min_consumed_out, const dtf_ps_ctrl* ctrl_step,
                     unsigned long long* step_out, const dtf_result_sink* sink,
                     const dtf_mirror_hdr* my_hdr, const unsigned int* sync_tag, void* stream);
/* Owner side, two launches: (1) claim the shard's drain lock and snapshot which workers have an
 * unconsumed gradient (at most one per worker per call, the oldest) into `todo` (local scratch,
 * DTF_MAX_WORKERS + 1 words; nothing is claimed while flags->paused or another drain holds the
 * lock); (2) for every item read var / slots once, apply the pending gradients one after another
 * with the beta powers each apply would have seen, write var / slots / bf16 serving copy once. Items
 * are dtf_apply_item whose `grad` points at the slice inside slot (worker 0, buffer 0). The last
 * CTA publishes consumed[w] += 1, multiplies the shard's beta powers, and - when ctrl_step is not
 * NULL (this shard owns global_step) - adds the number of applied gradients to global_step and
 * refreshes the workers' heartbeats; then it releases the lock. step_out (optional, local u64)
 * receives global_step. The drain may be launched by the owner's daemon thread or as a node of a
 * colocated worker's step graph: the lock makes concurrent launches safe. */
int dtf_mailbox_drain(int opt, const dtf_apply_item* items, int n_items, long long slot_stride,
                      int depth, int n_workers, const dtf_apply_hyper* hyper,
                      dtf_ps_ctrl* ctrl_powers, dtf_ps_ctrl* ctrl_step, dtf_mbox_flags* flags,
                      unsigned int* todo, unsigned int* block_counter,
                      unsigned long long* step_out, void* stream);
/* Same, and every updated element is also stored into each mirror registered in
 * flags->mirror_base (mirror_items[i] belongs to items[i]; NULL = no broadcast); afterwards the
 * mirrors' headers receive consumed[shard_index], global_step (step owner only) and, last,
 * version[shard_index] with release semantics. max_ctas > 0 caps the grid (the CTAs stride over
 * the 2048-element chunks): a drain that runs beside a worker's compute kernels then holds a few
 * CTA slots for longer instead of flooding every SM. */
int dtf_mailbox_drain_ex(int opt, const dtf_apply_item* items, int n_items, long long slot_stride,
                         int depth, int n_workers, const dtf_apply_hyper* hyper,
                         dtf_ps_ctrl* ctrl_powers, dtf_ps_ctrl* ctrl_step, dtf_mbox_flags* flags,
                         unsigned int* todo, unsigned int* block_counter,
                         unsigned long long* step_out, const dtf_drain_mirror_item* mirror_items,
                         int shard_index, int max_ctas, void* stream);

/* ------------------------------------------------------------------------------------------
 * Synchronous rounds on the same mailboxes (SyncReplicasOptimizer with backup workers and PS-only
 * tasks; templates/00_mnist_replica.py:60-63,169-191). A worker pushes ONE gradient per token,
 * tagged with its local_step. The owner's drain in sync mode looks at every worker's oldest
 * pending gradient: one whose tag is older than the shard's epoch (= rounds completed) is stale
 * and is retired without being applied (TF's ConditionalAccumulator drops it the same way); when
 * at least `replicas_to_aggregate` fresh gradients (tag == epoch) are pending, ALL fresh ones are
 * averaged and applied ONCE, the epoch advances, global_step (step owner) grows by one, and the
 * new parameters + epoch reach every worker's mirror (version[shard] = epoch). Workers wait for
 * their next token with dtf_sync_token_wait: a spin on their own mirror header.
 * ------------------------------------------------------------------------------------------ */
int dtf_mailbox_drain_sync(int opt, const dtf_apply_item* items, int n_items, long long slot_stride,
                           int n_workers, int replicas_to_aggregate, const dtf_apply_hyper* hyper,
                           dtf_ps_ctrl* ctrl /* this shard's: beta powers + epoch */,
                           dtf_ps_ctrl* ctrl_step /* or NULL */, dtf_mbox_flags* flags,
                           unsigned int* todo /* DTF_MAX_WORKERS + 4 words */,
                           unsigned int* block_counter,
                           const dtf_drain_mirror_item* mirror_items, int shard_index, int max_ctas,
                           void* stream);
/* Token wait: returns (on the stream) once every shard 0..n_shards-1 has completed more rounds than
 * *last_round - i.e. the round this worker last contributed to is over - or at once when
 * *last_round is 0xffffffff (first step). Stores the round to compute next (min over shards of
 * version[]) into *local_step_out, which the step's pushes pass as sync_tag, and into *last_round. */
int dtf_sync_token_wait(const dtf_mirror_hdr* my_hdr, int n_shards, unsigned int* last_round,
                        unsigned int* local_step_out, unsigned long long* step_out, void* stream);


/* Programmatic dependent launch for every kernel launched afterwards by this library (process-wide
 * switch; returns the previous setting). Every kernel calls griddepcontrol.wait before it touches
 * global memory, so a dependent's prologue (TMEM allocation, mbarrier init, descriptor prefetch) may
 * overlap its predecessor's tail. The executor turns it on for the forward chain only. */
int dtf_set_pdl(int enabled);

/* Host-side spin on a pinned word written by a kernel (see dtf_result_sink): 0 = reached
 * `expected`, 1 = timeout. */
int dtf_host_wait_u32(const unsigned int* word, unsigned int expected, long long timeout_us);
/* cudaGraphLaunch of an instantiated graph on an explicit stream (the executor's per-step call). */
int dtf_graph_launch(void* graph_exec, void* stream);

/* The shard owner's apply daemon: one native host thread per shard that polls the mailbox flags
 * (D2H copy, no SM use while idle) and launches dtf_mailbox_drain on its own low-priority stream
 * whenever a worker has published a gradient — the B200 stand-in for the PS task's CPU executing
 * the ApplyAdam ops placed on /job:ps (run_mnist_distributed.py:96-98,107,116). Pointers are
 * device addresses in the OWNER's address space. */
typedef struct dtf_mbox_daemon dtf_mbox_daemon;
/* mirror_items (may be NULL) / shard_index: see dtf_mailbox_drain_ex. backstop_us > 0 turns the
 * daemon into a backstop for a shard whose drains normally run inside a colocated worker's step
 * graph: it then only drains gradients that have been pending for at least that long (an owner
 * whose worker has stopped stepping must still serve the other workers). */
int dtf_mbox_daemon_start(dtf_mbox_daemon** out, int device, int opt, const dtf_apply_item* items,
                          int n_items, long long slot_stride, int depth, int n_workers,
                          const dtf_apply_hyper* hyper, dtf_ps_ctrl* ctrl_powers,
                          dtf_ps_ctrl* ctrl_step, dtf_mbox_flags* flags, int poll_us,
                          const dtf_drain_mirror_item* mirror_items, int shard_index,
                          int backstop_us, int sync_replicas_to_aggregate /* 0 = asynchronous */);
int dtf_mbox_daemon_set_backstop(dtf_mbox_daemon* d, int backstop_us);
int dtf_mbox_daemon_set_max_ctas(dtf_mbox_daemon* d, int max_ctas);   /* grid cap of its drains, 0 = none */
int dtf_mbox_daemon_pause(dtf_mbox_daemon* d, int pause);   /* 1: returns once parked; 0: resume */
int dtf_mbox_daemon_stats(dtf_mbox_daemon* d, unsigned long long* polls,
                          unsigned long long* drains, unsigned long long* applied); /* -1: daemon died */
int dtf_mbox_daemon_stop(dtf_mbox_daemon* d);               /* joins the thread, frees d */

/* ------------------------------------------------------------------------------------------
 * K11: bulk parameter pull, peer -> local (one launch for a list of tensors). Reference: the
 * per-tensor RecvTensor RPCs of every mon_sess.run (run_mnist_distributed.py:150,161).
 * ------------------------------------------------------------------------------------------ */
#define DTF_MAX_COPY_ITEMS 32
typedef struct dtf_copy_item {
  const void* src;
  void* dst;
  long long bytes;     /* multiple of 16 */
} dtf_copy_item;
int dtf_pull_bulk(const dtf_copy_item* items, int n_items, const dtf_ps_ctrl* ctrl_for_step,
                  unsigned long long* pulled_step_out, void* stream);

/* fp32 master -> bf16 row-padded serving copy (used at init / restore). */
int dtf_refresh_shadow(const float* var, void* shadow, long long n, long long first, int cols,
                       int shadow_ld, void* stream);

/* ------------------------------------------------------------------------------------------
 * K13 and small elementwise helpers.
 * ------------------------------------------------------------------------------------------ */
int dtf_u8_to_bf16(const unsigned char* src, void* dst, int rows, int cols, int ldd, float scale,
                   void* stream);                                    /* dataset.py:92-100 */
int dtf_f32_to_bf16(const float* src, int lds, void* dst, int ldd, int rows, int cols,
                    void* stream);
int dtf_colsum_bf16(const void* x, int ldx, int rows, int cols, float* out_accum, void* stream);
int dtf_fill_u32(void* dst, unsigned int value, long long n_words, void* stream);

/* ------------------------------------------------------------------------------------------
 * K6/K7/K8: the reference CNN's conv / pool layers (run_mnist_distributed.py:52-65), NHWC.
 * conv1 (Cin = 1) is a direct CUDA-core kernel fused with bias + ReLU + 2x2 max-pool; conv2
 * (Cin = 32) is an implicit GEMM on tcgen05 fed by 4-D TMA boxes over a zero-padded NHWC
 * activation, fused with bias + ReLU + pool in the TMEM epilogue.
 * ------------------------------------------------------------------------------------------ */
/* x: u8 [B,28,28] (scale applied) -> pool1p bf16 [B,18,18,32] interior (zero border kept by the
 * caller), amax1 u8 [B,14,14,32] (argmax position 0..3). w1: fp32 [5,5,1,32], b1: fp32 [32]. */
int dtf_conv1_fwd(const unsigned char* x_u8, const float* x_f32, float scale, const float* w1,
                  const float* b1, void* pool1p, unsigned char* amax1, int B, void* stream);
/* Same, but x may be PINNED HOST memory (device-mapped): images are fetched with 16-byte loads and,
 * when copy_* is given, mirrored into a device buffer (the backward pass reads that). Replaces the
 * step's H2D copy node. */
int dtf_conv1_fwd_ingest(const unsigned char* x_u8, const float* x_f32, float scale, const float* w1,
                         const float* b1, void* pool1p, unsigned char* amax1, unsigned char* copy_u8,
                         float* copy_f32, const void* y_src, void* y_dst, int y_bytes, int B,
                         void* stream);   /* y_*: labels mirrored host -> device by one CTA (optional) */
/* dpool1 bf16 [B,14,14,32] + amax1 + pooled output -> dW1 fp32 [25,32], db1 fp32 [32]
 * (accumulated with atomics: zero first). */
int dtf_conv1_bwd(const unsigned char* x_u8, const float* x_f32, float scale,
                  const void* dpool1, const void* pool1p, const unsigned char* amax1,
                  float* dw1_accum, float* db1_accum, int B, void* stream);
/* conv2 forward: pool1p [B,18,18,32] (*) w2 bf16 [800,64] -> pool2 bf16 [B,7,7,64],
 * amax2 u8 [B,7,7,64]. */
int dtf_conv2_fwd(const void* pool1p, const void* w2, const float* b2, void* pool2,
                  unsigned char* amax2, int B, void* stream);
/* Same results as dtf_conv2_fwd / dtf_conv2_dgrad from ONE halo tile per image in shared memory: the
 * activation is loaded once (a 5-D TMA box re-lays it as UMMA's un-swizzled K-major operand) and the
 * 25 filter taps are 25 shifted shared-memory descriptors; the whole filter is loaded once per CTA.
 * A quarter of the L2 traffic of the tap-by-tap kernels. */
int dtf_conv2_fwd_halo(const void* pool1p, const void* w2, const float* b2, void* pool2,
                       unsigned char* amax2, int B, void* stream);
int dtf_conv2_dgrad_halo(const void* dconv2p, const void* w2, void* dpool1, int B, void* stream);
/* pool2 gradient + ReLU gradient: dpool2 bf16 [B,7,7,64] -> dconv2p bf16 [B,18,18,64] interior
 * (zero border kept by the caller); db2 fp32 [64] accumulated. */
int dtf_pool2_bwd(const void* dpool2, int ld_dpool2, const void* pool2,
                  const unsigned char* amax2, void* dconv2p, float* db2_accum, int B,
                  void* stream);
/* conv2 weight gradient: dW2 fp32 [800,64] += im2col(pool1p)^T * dconv2 (split-K, atomics). */
int dtf_conv2_wgrad(const void* pool1p, const void* dconv2p, float* dw2_accum, int B,
                    void* stream);
/* conv2 input gradient: dpool1 bf16 [B,14,14,32] = fullcorr(dconv2p, w2). */
int dtf_conv2_dgrad(const void* dconv2p, const void* w2, void* dpool1, int B, void* stream);

/* ------------------------------------------------------------------------------------------
 * K12: synchronous round (SyncReplicasOptimizer, templates/00_mnist_replica.py:174-191):
 * reduce-scatter of the workers' gradients fused with mean + optimizer apply on the owning
 * shard, followed by an all-gather of the updated bf16 serving copy into every worker, all in
 * ONE cooperative kernel per rank over NVLink peer memory (no NCCL).
 * ------------------------------------------------------------------------------------------ */
/* One variable slice as seen by the sync kernel. Offsets are bytes from the base of an arena
 * (0 = absent); every rank's arena has the same layout, so one table serves all ranks.
 * All ranks pass the same table; `grad_local` is the only per-rank field. */
typedef struct dtf_sync_segment {
  int owner;                          /* shard that owns (reduces + applies) this slice */
  int cols;                           /* innermost dim of the variable */
  int shadow_ld;                      /* row pitch of the bf16 copies, elements */
  int reserved;
  long long n;                        /* elements in the slice */
  long long first;                    /* flat index of the slice's first element in the variable */
  long long first_row;                /* first / cols */
  unsigned long long grad_local;      /* device pointer: this rank's fp32 gradient slice */
  unsigned long long mailbox_off;     /* owner arena: slot 0 of this slice's mailbox */
  unsigned long long mailbox_stride;  /* bytes between consecutive workers' slots */
  unsigned long long var_off, slot0_off, slot1_off, shadow_off;   /* owner arena */
  unsigned long long mirror_f32_off;  /* every arena: fp32 mirror of the WHOLE variable */
  unsigned long long mirror_bf16_off; /* every arena: bf16 row-padded mirror of the WHOLE variable */
} dtf_sync_segment;
/* Runs round `round` (1, 2, 3, ... identical on all ranks) on this rank; see sync_step.cu.
 * segs_dev: device copy of the table; arena_bases[world]: every rank's arena as mapped in this
 * process; grid_bar: 2 zero-initialised local words; step_out: local u64 receiving global_step. */
int dtf_sync_round(int opt, const dtf_sync_segment* segs_dev, int n_segs, int rank, int world,
                   unsigned int round, const dtf_apply_hyper* hyper,
                   const unsigned long long* arena_bases, unsigned int* grid_bar,
                   unsigned long long* step_out, int step_shard, void* stream);

/* ------------------------------------------------------------------------------------------
 * Peer-memory runtime (replaces tf.train.Server's gRPC data plane, run_mnist_distributed.py:96).
 * ------------------------------------------------------------------------------------------ */
#define DTF_IPC_HANDLE_BYTES 64
int dtf_arena_alloc(void** ptr_out, size_t bytes);            /* cudaMalloc + zero */
int dtf_arena_free(void* ptr);
int dtf_ipc_export(void* ptr, unsigned char* handle_out);     /* 64-byte cudaIpcMemHandle_t */
int dtf_ipc_import(const unsigned char* handle, void** ptr_out);
int dtf_ipc_close(void* ptr);
int dtf_enable_peer_access(int peer_device);                  /* same-process multi-GPU */
int dtf_can_access_peer(int device, int peer_device);
int dtf_device_count(void);
int dtf_set_device(int device);
int dtf_memcpy_d2h(void* dst_host, const void* src_dev, size_t bytes, void* stream);
int dtf_memcpy_h2d(void* dst_dev, const void* src_host, size_t bytes, void* stream);
int dtf_stream_sync(void* stream);
int dtf_host_alloc_pinned(void** ptr_out, size_t bytes);
int dtf_host_free_pinned(void* ptr);

#ifdef __cplusplus
}
#endif
#endif /* DTF_B200_H_ */
