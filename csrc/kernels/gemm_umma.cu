// Dense GEMM on the 5th-generation tensor cores.
//
//   D[M,N] = A[M,K] * B[N,K]^T      bf16 operands, fp32 accumulation in TMEM
//
// One CTA per 128 x BN output tile (optionally one K-slice of it: split-K):
//   warp 0     TMA producer   cp.async.bulk.tensor -> 128B-swizzled smem ring, mbarrier tx-count
//   warp 1     MMA issuer     one elected thread issues tcgen05.mma (UMMA 128 x BN x 16), the
//                             ring slot is released by tcgen05.commit -> mbarrier
//   warp 2     TMEM allocator tcgen05.alloc / dealloc of BN fp32 columns
//   warps 4-7  epilogue       tcgen05.ld (thread = accumulator row), fused bias / ReLU /
//                             ReLU-grad mask / scale, bf16 or fp32 store, or red.global.add
//
// Both operands may be K-major (the reduction dim is contiguous in memory) or MN-major (the
// M/N dim is contiguous), so forward (X*W), input-gradient (dY*W^T) and weight-gradient (X^T*dY)
// all run without a transpose pass:
//   fwd   Y = X W      A = X  [M,K] K-major      B = W  [K,N] MN-major
//   dX    = dY W^T     A = dY [M,K] K-major      B = W  [N,K] K-major   (W stored [in,out])
//   dW    = X^T dY     A = X  [K,M] MN-major     B = dY [K,N] MN-major
// The B operand of the forward GEMM may live in another GPU's memory (the PS shard's bf16
// serving copy): the TMA descriptor is simply encoded on the peer-mapped address, which fuses the
// parameter pull into the first GEMM (reference: the _Recv of hid_w feeding MatMul,
// templates/00_mnist_replica.py:160 / run_mnist_distributed.py:67).
//
// The MNIST shapes are skinny (M = batch = 128): a 128 x 1024 x 3136 layer has only 8-16 output
// tiles for 148 SMs. Two remedies live here:
//   * the smem ring depth is a launch parameter, so short-K problems take little smem and
//     several CTAs share an SM;
//   * split-K with an in-kernel fix-up: every K-slice CTA adds its partial tile into an fp32
//     workspace with red.global.add.v4.f32, bumps a per-tile counter, and the LAST slice to arrive
//     re-reads the summed tile from L2 and runs the full epilogue (bias / ReLU / mask / bf16), so
//     fused epilogues survive splitting and no second kernel is needed.
#include "common.cuh"
#include "ptx.cuh"
#include "../../include/dtf_b200.h"

namespace dtf {

constexpr int GEMM_BM = 128;
constexpr int GEMM_BK = 64;  // 64 bf16 = 128 B = one swizzle row
constexpr int GEMM_THREADS = 256;
constexpr int GEMM_MAX_STAGES = 6;

template <int BN>
struct GemmCfg {
  static constexpr int A_BYTES = GEMM_BM * GEMM_BK * 2;  // 16 KB
  static constexpr int B_BYTES = BN * GEMM_BK * 2;       // 8 / 16 KB
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int MAX_STAGES = (BN <= 64) ? 6 : 5;
  static constexpr int smem_bytes(int stages) { return stages * STAGE_BYTES + 1024 + 256; }
};

struct GemmParams {
  int M, N, K;
  int kb_per_split;  // k-blocks (of 64) handled by one blockIdx.z
  int num_stages;
  int splits;
  // epilogue
  void* out;
  int ldo;
  int out_f32;
  int n_store;  // columns [0, n_store) are written (N rounded up into the row padding)
  const float* bias;
  int relu;
  const __nv_bfloat16* mask;
  int ldmask;
  int reduce;  // red.global.add.f32 into out instead of a store
  float alpha;
  // split-K fix-up (splits > 1 && !reduce)
  float* ws;                // [M, ldws] fp32, zero on entry
  int ldws;
  unsigned int* counters;   // one per output tile, zero on entry (self-resetting)
  int ws_clean;             // the fix-up leaves the workspace zeroed
  // fused push + optimizer apply (see dtf_gemm_epilogue)
  int apply_opt;
  int apply_rows_per_slice;
  int apply_shadow_ld;
  dtf_apply_hyper ah;
  const dtf_ps_ctrl* apply_ctrl_powers;
  float* avar[8];
  float* aslot0[8];
  float* aslot1[8];
  __nv_bfloat16* ashadow[8];
  // scatter store into gradient mailboxes (see dtf_gemm_epilogue.scatter_*)
  int scatter_enable;
  int scatter_rows_per_slice;
  int scatter_depth;
  const unsigned int* scatter_seq;
  float* scatter_base[8];
  long long scatter_buf_stride[8];
  // max-pool + ReLU gradient scatter (see dtf_gemm_epilogue.pool_*)
  __nv_bfloat16* pool_out;
  const unsigned char* pool_amax;
  float* pool_dbias;
};

__device__ __forceinline__ void gemm_apply_one(int opt, const dtf_apply_hyper& h, float lr_t, float g,
                                               float& var, float& s0, float& s1) {
  if (opt == DTF_OPT_SGD) {
    var -= h.lr * g;
  } else if (opt == DTF_OPT_MOMENTUM) {
    s0 = s0 * h.momentum + g;
    var -= h.use_nesterov ? (g * h.lr + s0 * h.momentum * h.lr) : h.lr * s0;
  } else if (opt == DTF_OPT_ADAGRAD) {
    s0 += g * g;
    var -= h.lr * g * rsqrtf(s0);
  } else {
    s0 += (g - s0) * (1.f - h.beta1);
    s1 += (g * g - s1) * (1.f - h.beta2);
    var -= lr_t * s0 / (sqrtf(s1) + h.epsilon);
  }
}

// Fused epilogue of a weight-gradient GEMM: v[0..31] are the gradients of row m, columns
// nb..nb+31 of a variable [M, N]; read-modify-write var / slots on the owning slice (local or
// peer memory) and refresh its bf16 serving copy. Unlocked, like every PS apply here.
__device__ __forceinline__ void gemm_epilogue_apply(const GemmParams& p, int m, int nb, float (&v)[32],
                                                    float lr_t) {
  const int slice = m / p.apply_rows_per_slice;
  const long long r = m - slice * p.apply_rows_per_slice;
  const long long idx = r * p.N + nb;
  float* var = p.avar[slice] + idx;
  float* s0p = p.aslot0[slice] ? p.aslot0[slice] + idx : nullptr;
  float* s1p = p.aslot1[slice] ? p.aslot1[slice] + idx : nullptr;
  __nv_bfloat16* sh = p.ashadow[slice] ? p.ashadow[slice] + r * p.apply_shadow_ld + nb : nullptr;
  const float gs = p.alpha * p.ah.grad_scale;
  const bool vec = (p.N % 4 == 0) && (nb + 32 <= p.N) && (p.apply_shadow_ld % 4 == 0);
  if (vec) {
#pragma unroll
    for (int g8 = 0; g8 < 4; ++g8) {          // 8 columns at a time keeps the register count flat
      float4 w0 = *reinterpret_cast<float4*>(var + g8 * 8), w1 = *reinterpret_cast<float4*>(var + g8 * 8 + 4);
      float4 a0 = make_float4(0, 0, 0, 0), a1 = a0, b0 = a0, b1 = a0;
      if (s0p) { a0 = *reinterpret_cast<float4*>(s0p + g8 * 8); a1 = *reinterpret_cast<float4*>(s0p + g8 * 8 + 4); }
      if (s1p) { b0 = *reinterpret_cast<float4*>(s1p + g8 * 8); b1 = *reinterpret_cast<float4*>(s1p + g8 * 8 + 4); }
      float w[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
      float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
      for (int j = 0; j < 8; ++j) gemm_apply_one(p.apply_opt, p.ah, lr_t, v[g8 * 8 + j] * gs, w[j], a[j], b[j]);
      *reinterpret_cast<float4*>(var + g8 * 8) = make_float4(w[0], w[1], w[2], w[3]);
      *reinterpret_cast<float4*>(var + g8 * 8 + 4) = make_float4(w[4], w[5], w[6], w[7]);
      if (s0p) {
        *reinterpret_cast<float4*>(s0p + g8 * 8) = make_float4(a[0], a[1], a[2], a[3]);
        *reinterpret_cast<float4*>(s0p + g8 * 8 + 4) = make_float4(a[4], a[5], a[6], a[7]);
      }
      if (s1p) {
        *reinterpret_cast<float4*>(s1p + g8 * 8) = make_float4(b[0], b[1], b[2], b[3]);
        *reinterpret_cast<float4*>(s1p + g8 * 8 + 4) = make_float4(b[4], b[5], b[6], b[7]);
      }
      if (sh) {
        uint4 o;
        o.x = pack_bf16x2(w[0], w[1]); o.y = pack_bf16x2(w[2], w[3]);
        o.z = pack_bf16x2(w[4], w[5]); o.w = pack_bf16x2(w[6], w[7]);
        if ((p.apply_shadow_ld & 7) == 0) {
          *reinterpret_cast<uint4*>(sh + g8 * 8) = o;
        } else {
          *reinterpret_cast<uint2*>(sh + g8 * 8) = make_uint2(o.x, o.y);
          *reinterpret_cast<uint2*>(sh + g8 * 8 + 4) = make_uint2(o.z, o.w);
        }
      }
    }
  } else {
#pragma unroll 1
    for (int j = 0; j < 32; ++j) {
      if (nb + j >= p.N) break;
      float w = var[j];
      float a = s0p ? s0p[j] : 0.f, b = s1p ? s1p[j] : 0.f;
      gemm_apply_one(p.apply_opt, p.ah, lr_t, v[j] * gs, w, a, b);
      var[j] = w;
      if (s0p) s0p[j] = a;
      if (s1p) s1p[j] = b;
      if (sh) sh[j] = __float2bfloat16(w);
    }
  }
}

// Applies the fused epilogue to 32 consecutive accumulator columns of row m and stores them.
// Pool-scatter mode is called by EVERY lane of the warp (valid = this lane's row exists) because its
// bias-gradient column sums use warp shuffles; all other modes are only called with valid = true.
__device__ __forceinline__ void gemm_epilogue_store(const GemmParams& p, int m, int nb, float (&v)[32],
                                                    bool valid = true) {
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] *= p.alpha;
  if (p.pool_out != nullptr) {
    // The GEMM's [B, 49 x 64] result is the gradient of the 7x7x64 pooled map (NHWC): these 32 columns
    // are channels c0..c0+31 of ONE pooled pixel. ReLU gradient (pooled activation > 0), then each
    // value goes to the position of its 2x2 window that won the forward max (amax = dy*2 + dx), zeros
    // to the other three: the [B,18,18,64] zero-bordered conv2 output gradient is written directly, and
    // the conv2 bias gradient is the column sum. Replaces the dpool2 round trip + pool2_bwd kernel.
    const int pixel = nb >> 6, c0 = nb & 63;
    const int py = pixel / 7, px = pixel - py * 7;
    if (!valid) {
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = 0.f;
    } else {
      const __nv_bfloat16* mrow = p.mask + static_cast<size_t>(m) * p.ldmask + nb;
      uint32_t am[8];
      const uint4* ap = reinterpret_cast<const uint4*>(p.pool_amax + (static_cast<size_t>(m) * 49 + pixel) * 64 + c0);
      const uint4 a0 = ap[0], a1 = ap[1];
      am[0] = a0.x; am[1] = a0.y; am[2] = a0.z; am[3] = a0.w;
      am[4] = a1.x; am[5] = a1.y; am[6] = a1.z; am[7] = a1.w;
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        const uint4 mv = *reinterpret_cast<const uint4*>(mrow + g * 8);
        const __nv_bfloat16* mb = reinterpret_cast<const __nv_bfloat16*>(&mv);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (!(__bfloat162float(mb[j]) > 0.f)) v[g * 8 + j] = 0.f;
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        __nv_bfloat16* orow = p.pool_out +
            ((static_cast<size_t>(m) * 18 + 2 + 2 * py + (q >> 1)) * 18 + 2 + 2 * px + (q & 1)) * 64 + c0;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          float t[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int jj = g * 8 + j;
            const uint32_t a = (am[jj >> 2] >> (8 * (jj & 3))) & 0xffu;
            t[j] = (a == static_cast<uint32_t>(q)) ? v[jj] : 0.f;
          }
          *reinterpret_cast<uint4*>(orow + g * 8) =
              make_uint4(pack_bf16x2(t[0], t[1]), pack_bf16x2(t[2], t[3]), pack_bf16x2(t[4], t[5]),
                         pack_bf16x2(t[6], t[7]));
        }
      }
    }
    // column sums over the warp's 32 rows: a transposing butterfly (31 shuffles) leaves column L's
    // sum in lane L
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
      const bool up = (lane & off) != 0;
#pragma unroll
      for (int i = 0; i < off; ++i) {
        const float keep = up ? v[i + off] : v[i];
        const float give = up ? v[i] : v[i + off];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, give, off);
      }
    }
    if (v[0] != 0.f) atomicAdd(p.pool_dbias + c0 + lane, v[0]);
    return;
  }
  if (p.scatter_enable) {
    // weight-gradient rows go straight into this worker's current mailbox buffer on the shard that
    // owns them: posted stores over NVLink (or local), no gradient buffer, no copy kernel
    const int slice = m / p.scatter_rows_per_slice;
    const long long r = m - static_cast<long long>(slice) * p.scatter_rows_per_slice;
    const unsigned int buf = *reinterpret_cast<const volatile unsigned int*>(p.scatter_seq) %
                             static_cast<unsigned int>(p.scatter_depth);
    float* orow = reinterpret_cast<float*>(reinterpret_cast<char*>(p.scatter_base[slice]) +
                                           static_cast<long long>(buf) * p.scatter_buf_stride[slice]) +
                  r * p.N + nb;
    if ((p.N & 3) == 0 && nb + 32 <= p.N) {
#pragma unroll
      for (int g = 0; g < 8; ++g)
        *reinterpret_cast<float4*>(orow + g * 4) =
            make_float4(v[g * 4], v[g * 4 + 1], v[g * 4 + 2], v[g * 4 + 3]);
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (nb + j < p.N) orow[j] = v[j];
    }
    return;
  }
  if (p.bias != nullptr) {
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (nb + j < p.N) v[j] += __ldg(p.bias + nb + j);
  }
  if (p.relu) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
  }
  if (p.mask != nullptr) {
    const __nv_bfloat16* mrow = p.mask + static_cast<size_t>(m) * p.ldmask + nb;
    if ((p.ldmask & 7) == 0 && nb + 32 <= p.ldmask) {
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        uint4 mv = *reinterpret_cast<const uint4*>(mrow + g * 8);
        const __nv_bfloat16* mb = reinterpret_cast<const __nv_bfloat16*>(&mv);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (!(__bfloat162float(mb[j]) > 0.f)) v[g * 8 + j] = 0.f;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (nb + j < p.ldmask && !(__bfloat162float(mrow[j]) > 0.f)) v[j] = 0.f;
    }
  }
  if (p.out_f32) {
    float* orow = reinterpret_cast<float*>(p.out) + static_cast<size_t>(m) * p.ldo + nb;
    if (p.reduce) {
      if ((p.ldo & 3) == 0 && nb + 32 <= p.n_store) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) red_add_v4_f32_sys(orow + j, v[j], v[j + 1], v[j + 2], v[j + 3]);
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (nb + j < p.n_store) red_add_f32_sys(orow + j, v[j]);
      }
    } else if ((p.ldo & 3) == 0 && nb + 32 <= p.n_store) {
#pragma unroll
      for (int g = 0; g < 8; ++g)
        *reinterpret_cast<float4*>(orow + g * 4) =
            make_float4(v[g * 4], v[g * 4 + 1], v[g * 4 + 2], v[g * 4 + 3]);
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (nb + j < p.n_store) orow[j] = v[j];
    }
  } else {
    __nv_bfloat16* orow =
        reinterpret_cast<__nv_bfloat16*>(p.out) + static_cast<size_t>(m) * p.ldo + nb;
    if ((p.ldo & 7) == 0 && nb + 32 <= p.n_store) {
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        uint4 o;
        o.x = pack_bf16x2(v[g * 8 + 0], v[g * 8 + 1]);
        o.y = pack_bf16x2(v[g * 8 + 2], v[g * 8 + 3]);
        o.z = pack_bf16x2(v[g * 8 + 4], v[g * 8 + 5]);
        o.w = pack_bf16x2(v[g * 8 + 6], v[g * 8 + 7]);
        *reinterpret_cast<uint4*>(orow + g * 8) = o;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (nb + j < p.n_store) orow[j] = __float2bfloat16(v[j]);
    }
  }
}

template <int BN, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_umma_kernel(const __grid_constant__ CUtensorMap tmap_a,
                 const __grid_constant__ CUtensorMap tmap_b, const GemmParams p) {
  using Cfg = GemmCfg<BN>;
  const int S = p.num_stages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S * Cfg::STAGE_BYTES);
  uint64_t* empty_bar = full_bar + GEMM_MAX_STAGES;
  uint64_t* tmem_full_bar = empty_bar + GEMM_MAX_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);
  uint32_t* last_flag = tmem_slot + 1;

  const int warp = threadIdx.x >> 5;
  const int n0 = blockIdx.x * BN;
  const int m0 = blockIdx.y * GEMM_BM;
  const int num_kb_total = (p.K + GEMM_BK - 1) / GEMM_BK;
  const int kb_begin = blockIdx.z * p.kb_per_split;
  const int kb_end = min(num_kb_total, kb_begin + p.kb_per_split);
  const int num_kb = max(0, kb_end - kb_begin);

  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
  }
  if (warp == 1 && elect_one()) {
    for (int i = 0; i < S; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    mbar_init(tmem_full_bar, 1);
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, BN);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // PDL: everything above overlapped the previous kernel; its outputs are needed from here on.
  griddep_wait();
  griddep_launch();

  if (warp == 0) {
    // ------------------------------ TMA producer ------------------------------
    if (elect_one()) {
      int s = 0;
      uint32_t ph = 0;
      for (int i = 0; i < num_kb; ++i) {
        mbar_wait(&empty_bar[s], ph ^ 1);
        uint8_t* a_dst = smem + s * Cfg::STAGE_BYTES;
        uint8_t* b_dst = a_dst + Cfg::A_BYTES;
        mbar_arrive_expect_tx(&full_bar[s], Cfg::STAGE_BYTES);
        const int k0 = (kb_begin + i) * GEMM_BK;
        if constexpr (!A_MN) {
          tma_load_2d(a_dst, &tmap_a, &full_bar[s], k0, m0);  // box {64 k, 128 m}
        } else {
#pragma unroll
          for (int c = 0; c < GEMM_BM / 64; ++c)  // box {64 m, 64 k}
            tma_load_2d(a_dst + c * 8192, &tmap_a, &full_bar[s], m0 + c * 64, k0);
        }
        if constexpr (!B_MN) {
          tma_load_2d(b_dst, &tmap_b, &full_bar[s], k0, n0);  // box {64 k, BN n}
        } else {
#pragma unroll
          for (int c = 0; c < BN / 64; ++c)  // box {64 n, 64 k}
            tma_load_2d(b_dst + c * 8192, &tmap_b, &full_bar[s], n0 + c * 64, k0);
        }
        if (++s == S) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer --------------------------------
    if (elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(GEMM_BM, BN, A_MN, B_MN);
      int s = 0;
      uint32_t ph = 0;
      for (int i = 0; i < num_kb; ++i) {
        mbar_wait(&full_bar[s], ph);
        tc_fence_after();
        const uint32_t a_base = smem_u32(smem + s * Cfg::STAGE_BYTES);
        const uint32_t b_base = a_base + Cfg::A_BYTES;
#pragma unroll
        for (int k = 0; k < GEMM_BK / 16; ++k) {
          // K-major, 128B swizzle: rows of 128 B, 8-row groups 1024 B apart; a K=16 slice is 32 B
          // inside the swizzled row. MN-major: 64-element chunks 8192 B apart (LBO), 8 k-rows per
          // 1024 B group (SBO); a K=16 slice is 16 rows = 2048 B.
          const uint64_t adesc = A_MN ? umma_smem_desc(a_base + k * 2048, 8192, 1024, 2)
                                      : umma_smem_desc(a_base + k * 32, 16, 1024, 2);
          const uint64_t bdesc = B_MN ? umma_smem_desc(b_base + k * 2048, 8192, 1024, 2)
                                      : umma_smem_desc(b_base + k * 32, 16, 1024, 2);
          umma_bf16(tmem_base, adesc, bdesc, idesc, (i > 0 || k > 0) ? 1u : 0u);
        }
        umma_commit(&empty_bar[s]);  // frees the smem slot once these MMAs have read it
        if (++s == S) { s = 0; ph ^= 1; }
      }
      umma_commit(tmem_full_bar);  // accumulator complete
    }
  } else if (warp >= 4) {
    // ------------------------------ epilogue ----------------------------------
    const int q = warp - 4;  // TMEM lane quarter == warp_id % 4
    const int lane = threadIdx.x & 31;
    const int m = m0 + q * 32 + lane;
    if (num_kb > 0) {
      mbar_wait(tmem_full_bar, 0);
      tc_fence_after();
    }
    const bool row_ok = m < p.M;
    const bool fixup = p.splits > 1 && !p.reduce;
    float lr_t = p.ah.lr;
    if (p.apply_opt == DTF_OPT_ADAM) {
      // shared beta powers live on the PS; every thread reads them (two cached 4-byte loads)
      const float b1p = *reinterpret_cast<const volatile float*>(&p.apply_ctrl_powers->beta1_power);
      const float b2p = *reinterpret_cast<const volatile float*>(&p.apply_ctrl_powers->beta2_power);
      lr_t = p.ah.lr * sqrtf(1.f - b2p) / (1.f - b1p);
    }
#pragma unroll 1
    for (int c = 0; c < BN / 32; ++c) {
      uint32_t r[32];
      if (num_kb > 0) {
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + c * 32, r);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) r[j] = 0;
      }
      const int nb = n0 + c * 32;
      const bool pool_now = p.pool_out != nullptr && !fixup;   // called warp-wide (shuffles inside)
      if ((!row_ok && !pool_now) || nb >= p.n_store) continue;
      float v[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
      if (p.apply_opt >= 0) {
        gemm_epilogue_apply(p, m, nb, v, lr_t);
      } else if (!fixup) {
        gemm_epilogue_store(p, m, nb, v, row_ok);
      } else {
        // partial tile -> fp32 workspace (ldws is a multiple of 4, rows are 16-byte aligned)
        float* wrow = p.ws + static_cast<size_t>(m) * p.ldws + nb;
#pragma unroll
        for (int j = 0; j < 32; j += 4)
          if (nb + j < p.ldws) red_add_v4_f32_gpu(wrow + j, v[j], v[j + 1], v[j + 2], v[j + 3]);
      }
    }
    if (fixup) {
      // elect the last K-slice of this tile (threadfence-reduction pattern)
      __threadfence();
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (threadIdx.x == 128) {
        const unsigned int tile = blockIdx.y * gridDim.x + blockIdx.x;
        const unsigned int old = atomicInc(p.counters + tile, static_cast<unsigned int>(p.splits - 1));
        *last_flag = (old == static_cast<unsigned int>(p.splits - 1)) ? 1u : 0u;
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (*last_flag) {
        __threadfence();
        if (row_ok || p.pool_out != nullptr) {
#pragma unroll 1
          for (int c = 0; c < BN / 32; ++c) {
            const int nb = n0 + c * 32;
            if (nb >= p.n_store) continue;
            const float* wrow = p.ws + static_cast<size_t>(row_ok ? m : 0) * p.ldws + nb;
            float v[32];
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
              if (row_ok && nb + j < p.ldws) {
                t = __ldcg(reinterpret_cast<const float4*>(wrow + j));
                // self-resetting workspace: the next GEMM's partial sums start from zero again
                if (p.ws_clean)
                  __stcg(reinterpret_cast<float4*>(const_cast<float*>(wrow) + j), make_float4(0.f, 0.f, 0.f, 0.f));
              }
              v[j] = t.x; v[j + 1] = t.y; v[j + 2] = t.z; v[j + 3] = t.w;
            }
            gemm_epilogue_store(p, m, nb, v, row_ok);
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, BN);
  }
}

template <int BN, bool A_MN, bool B_MN>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p,
                       cudaStream_t stream) {
  using Cfg = GemmCfg<BN>;
  auto kern = gemm_umma_kernel<BN, A_MN, B_MN>;
  const int smem = Cfg::smem_bytes(p.num_stages);
  DTF_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      Cfg::smem_bytes(Cfg::MAX_STAGES)));
  dim3 grid(ceil_div(p.N, BN), ceil_div(p.M, GEMM_BM), p.splits);
  DTF_CHECK_CUDA(launch_kernel(kern, grid, dim3(GEMM_THREADS), smem, stream, ta, tb, p));
  return 0;
}


// ------------------------------------------------------------------------------------------------
// Large-GEMM variant (SURVEY K1-K3 at the wide-MLP sizes): persistent CTA PAIRS.
//
// A 128 x 128 single-CTA tile moves 32 KB of operands through L2 and shared memory for every
// 2.1 MFLOP; at tensor peak that is ~128 B/clk per SM in each direction, more than either path
// sustains, and the single-CTA kernel above stops near 60% of the cuBLAS rate. Here two CTAs of a
// cluster (the two SMs of one TPC) issue ONE tcgen05.mma.cta_group::2 of M = 256, N = 256: each
// CTA stages only its own 128 rows of A and its own 128-column half of B (32 KB per k-block for
// 2 x the math), holds its 128 x 256 accumulator half in its own TMEM, and runs its own epilogue.
//   * persistent: grid = one pair per TPC, tiles assigned round-robin in an L2-friendly raster
//     (8 M-tiles wide), so barrier init / TMEM alloc / descriptor prefetch happen once;
//   * TMEM holds TWO accumulator tiles (2 x 256 columns): the epilogue of tile i drains one while
//     the MMAs of tile i+1 fill the other;
//   * every TMA load (from either CTA) credits the LEADER's full barrier; the leader's
//     tcgen05.commit multicasts "slot free" / "accumulator ready" to both CTAs; the peer's
//     epilogue warps release the accumulator with a remote mbarrier arrive.
// ------------------------------------------------------------------------------------------------
constexpr int PAIR_STAGES = 6;
constexpr int PAIR_STAGE_BYTES = 32768;  // per CTA: A 128 x 64 + its half of B 128 x 64, bf16
constexpr int PAIR_TILE = 256;
constexpr int PAIR_RASTER = 8;           // M-tiles per raster group
constexpr int pair_smem_bytes() { return PAIR_STAGES * PAIR_STAGE_BYTES + 1024 + 256; }

__device__ __forceinline__ void pair_tile_origin(int t, int tiles_m, int tiles_n, int& m0, int& n0) {
  const int per_group = PAIR_RASTER * tiles_n;
  const int g = t / per_group, r = t - g * per_group;
  const int gm = min(PAIR_RASTER, tiles_m - g * PAIR_RASTER);
  m0 = (g * PAIR_RASTER + r % gm) * PAIR_TILE;
  n0 = (r / gm) * PAIR_TILE;
}

template <bool A_MN, bool B_MN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS, 1)
gemm_umma_pair_kernel(const __grid_constant__ CUtensorMap tmap_a,
                      const __grid_constant__ CUtensorMap tmap_b, const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + PAIR_STAGES * PAIR_STAGE_BYTES);
  uint64_t* empty_bar = full_bar + PAIR_STAGES;
  uint64_t* tmem_full_bar = empty_bar + PAIR_STAGES;  // [2] accumulator ready   (per CTA)
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;       // [2] accumulator drained (leader's is used)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const uint32_t rank = cluster_ctarank();            // 0 = leader (issues the MMAs)
  const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;
  const int tiles_m = (p.M + PAIR_TILE - 1) / PAIR_TILE, tiles_n = (p.N + PAIR_TILE - 1) / PAIR_TILE;
  const int num_tiles = tiles_m * tiles_n;
  const int num_kb = (p.K + GEMM_BK - 1) / GEMM_BK;

  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
  }
  if (warp == 1 && elect_one()) {
    for (int i = 0; i < PAIR_STAGES; ++i) {
      mbar_init(&full_bar[i], 1);   // the leader producer's arrive.expect_tx (bytes of BOTH CTAs)
      mbar_init(&empty_bar[i], 1);  // the leader's multicast tcgen05.commit
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tmem_full_bar[a], 1);
      mbar_init(&tmem_empty_bar[a], 8);  // 4 epilogue warps x 2 CTAs
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc_pair(tmem_slot, 512);
    tmem_relinquish_pair();
  }
  tc_fence_before();
  cluster_sync_all();  // barriers of BOTH CTAs are initialised before any remote arrive / TMA credit
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  griddep_wait();
  griddep_launch();

  if (warp == 0) {
    // ------------------------------ TMA producer (both CTAs) ------------------------------
    if (elect_one()) {
      int s = 0;
      uint32_t ph = 0;
      for (int t = pair; t < num_tiles; t += num_pairs) {
        int m0, n0;
        pair_tile_origin(t, tiles_m, tiles_n, m0, n0);
        m0 += rank * GEMM_BM;  // my rows of A
        n0 += rank * 128;      // my half of B
        for (int i = 0; i < num_kb; ++i) {
          mbar_wait(&empty_bar[s], ph ^ 1);
          if (rank == 0) mbar_arrive_expect_tx(&full_bar[s], 2 * PAIR_STAGE_BYTES);
          uint8_t* a_dst = smem + s * PAIR_STAGE_BYTES;
          uint8_t* b_dst = a_dst + 16384;
          const int k0 = i * GEMM_BK;
          if constexpr (!A_MN) {
            tma_load_2d_pair(a_dst, &tmap_a, &full_bar[s], k0, m0);
          } else {
            tma_load_2d_pair(a_dst, &tmap_a, &full_bar[s], m0, k0);
            tma_load_2d_pair(a_dst + 8192, &tmap_a, &full_bar[s], m0 + 64, k0);
          }
          if constexpr (!B_MN) {
            tma_load_2d_pair(b_dst, &tmap_b, &full_bar[s], k0, n0);
          } else {
            tma_load_2d_pair(b_dst, &tmap_b, &full_bar[s], n0, k0);
            tma_load_2d_pair(b_dst + 8192, &tmap_b, &full_bar[s], n0 + 64, k0);
          }
          if (++s == PAIR_STAGES) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer (leader only) ------------------------------
    if (rank == 0 && elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(256, 256, A_MN, B_MN);
      int s = 0, it = 0;
      uint32_t ph = 0;
      for (int t = pair; t < num_tiles; t += num_pairs, ++it) {
        const int acc = it & 1;
        const uint32_t acc_ph = (it >> 1) & 1;
        mbar_wait(&tmem_empty_bar[acc], acc_ph ^ 1);  // both epilogues drained this accumulator
        tc_fence_after();
        const uint32_t d = tmem_base + acc * 256;
        for (int i = 0; i < num_kb; ++i) {
          mbar_wait(&full_bar[s], ph);
          tc_fence_after();
          const uint32_t a_base = smem_u32(smem + s * PAIR_STAGE_BYTES);
          const uint32_t b_base = a_base + 16384;
#pragma unroll
          for (int k = 0; k < GEMM_BK / 16; ++k) {
            const uint64_t adesc = A_MN ? umma_smem_desc(a_base + k * 2048, 8192, 1024, 2)
                                        : umma_smem_desc(a_base + k * 32, 16, 1024, 2);
            const uint64_t bdesc = B_MN ? umma_smem_desc(b_base + k * 2048, 8192, 1024, 2)
                                        : umma_smem_desc(b_base + k * 32, 16, 1024, 2);
            umma_bf16_pair(d, adesc, bdesc, idesc, (i > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit_pair(&empty_bar[s], 3);  // slot free in both CTAs
          if (++s == PAIR_STAGES) { s = 0; ph ^= 1; }
        }
        umma_commit_pair(&tmem_full_bar[acc], 3);  // accumulator complete in both CTAs
      }
    }
  } else if (warp >= 4) {
    // ------------------------------ epilogue (both CTAs) ----------------------------------
    const int q = warp - 4;
    const int lane = threadIdx.x & 31;
    const uint32_t leader_empty[2] = {mapa_u32(&tmem_empty_bar[0], 0), mapa_u32(&tmem_empty_bar[1], 0)};
    int it = 0;
    for (int t = pair; t < num_tiles; t += num_pairs, ++it) {
      const int acc = it & 1;
      const uint32_t acc_ph = (it >> 1) & 1;
      int m0, n0;
      pair_tile_origin(t, tiles_m, tiles_n, m0, n0);
      const int m = m0 + rank * GEMM_BM + q * 32 + lane;
      const bool row_ok = m < p.M;
      mbar_wait(&tmem_full_bar[acc], acc_ph);
      tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < PAIR_TILE / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * 256 + c * 32, r);
        tmem_ld_wait();
        const int nb = n0 + c * 32;
        if (!row_ok || nb >= p.n_store) continue;
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
        gemm_epilogue_store(p, m, nb, v);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(leader_empty[acc]);
    }
  }

  tc_fence_before();
  cluster_sync_all();  // the peer's smem / TMEM stay alive until the leader's MMAs have finished
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, 512);
  }
}

template <bool A_MN, bool B_MN>
static int launch_gemm_pair(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p,
                            cudaStream_t stream) {
  auto kern = gemm_umma_pair_kernel<A_MN, B_MN>;
  DTF_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      pair_smem_bytes()));
  static int sm_count = 0;
  if (sm_count == 0) {
    int dev = 0;
    DTF_CHECK_CUDA(cudaGetDevice(&dev));
    DTF_CHECK_CUDA(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
  }
  const int tiles = ceil_div(p.M, PAIR_TILE) * ceil_div(p.N, PAIR_TILE);
  const int pairs = tiles < sm_count / 2 ? tiles : sm_count / 2;
  DTF_CHECK_CUDA(launch_kernel(kern, dim3(2 * pairs), dim3(GEMM_THREADS), pair_smem_bytes(), stream,
                               ta, tb, p));
  return 0;
}

}  // namespace dtf

using namespace dtf;

extern "C" int dtf_gemm_bf16(const void* A, int lda, int a_mn, const void* B, int ldb, int b_mn,
                             int M, int N, int K, const dtf_gemm_epilogue* epi, int splits,
                             int block_n, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  DTF_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: bad shape M=%d N=%d K=%d", M, N, K);
  DTF_REQUIRE(epi != nullptr && (epi->out != nullptr || epi->apply_enable || epi->scatter_enable ||
                                 epi->pool_out != nullptr),
              "gemm: epilogue/out is null");
  DTF_REQUIRE(lda % 8 == 0 && ldb % 8 == 0,
              "gemm: lda=%d ldb=%d must be multiples of 8 elements (TMA 16 B rule)", lda, ldb);
  const int num_kb = ceil_div(K, GEMM_BK);
  if (splits < 1) splits = 1;
  if (splits > num_kb) splits = num_kb;
  // block_n = 256 asks for the persistent CTA-pair kernel; it is also the default for GEMMs with
  // at least one 256 x 256 tile per TPC (it has no split-K and no fused-apply epilogue)
  const bool pair_ok = splits == 1 && !epi->apply_enable && epi->pool_out == nullptr;
  DTF_REQUIRE(block_n != 256 || pair_ok, "gemm: block_n = 256 cannot be combined with split-K or fused apply");
  const bool use_pair = pair_ok && (block_n == 256 || (block_n == 0 && static_cast<long long>(ceil_div(M, 256)) *
                                                                           ceil_div(N, 256) >= 74 && K >= 256));
  const int BN = use_pair ? 128 : ((block_n == 64 || block_n == 128) ? block_n : (N > 64 ? 128 : 64));
  const bool fixup = splits > 1 && !epi->reduce;
  DTF_REQUIRE(splits == 1 || (epi->reduce && epi->out_f32) || (epi->splitk_ws && epi->tile_counters),
              "gemm: split-K needs an fp32 reduce epilogue, or a workspace + tile counters");
  DTF_REQUIRE(!fixup || (epi->ldws % 4 == 0 && epi->ldws >= N &&
                         (reinterpret_cast<uintptr_t>(epi->splitk_ws) & 15) == 0),
              "gemm: split-K workspace needs ldws %% 4 == 0, ldws >= N and a 16-byte aligned base");

  CUtensorMap ta, tb;
  {
    uint64_t dims[2], str[1];
    uint32_t box[2];
    if (!a_mn) {  // memory [M, K]
      dims[0] = K; dims[1] = M; box[0] = 64; box[1] = GEMM_BM;
    } else {      // memory [K, M]
      dims[0] = M; dims[1] = K; box[0] = 64; box[1] = 64;
    }
    str[0] = static_cast<uint64_t>(lda) * 2;
    if (encode_tmap_bf16(&ta, A, 2, dims, str, box, 128) != 0) return -1;
    if (!b_mn) {  // memory [N, K]
      dims[0] = K; dims[1] = N; box[0] = 64; box[1] = BN;
    } else {      // memory [K, N]
      dims[0] = N; dims[1] = K; box[0] = 64; box[1] = 64;
    }
    str[0] = static_cast<uint64_t>(ldb) * 2;
    if (encode_tmap_bf16(&tb, B, 2, dims, str, box, 128) != 0) return -1;
  }

  GemmParams p;
  memset(&p, 0, sizeof(p));
  p.M = M; p.N = N; p.K = K;
  p.kb_per_split = ceil_div(num_kb, splits);
  p.splits = ceil_div(num_kb, p.kb_per_split);
  const int max_stages = BN <= 64 ? GemmCfg<64>::MAX_STAGES : GemmCfg<128>::MAX_STAGES;
  p.num_stages = p.kb_per_split < max_stages ? (p.kb_per_split < 2 ? 2 : p.kb_per_split) : max_stages;
  p.out = epi->out;
  p.ldo = epi->ldo;
  p.out_f32 = epi->out_f32;
  {
    const int pad = epi->out_f32 ? 1 : 8;
    int ns = round_up(N, pad);
    p.n_store = ns < epi->ldo ? ns : epi->ldo;
  }
  p.bias = epi->bias;
  p.relu = epi->relu;
  p.mask = reinterpret_cast<const __nv_bfloat16*>(epi->mask);
  p.ldmask = epi->ldmask;
  p.reduce = epi->reduce;
  p.alpha = epi->alpha;
  p.ws = epi->splitk_ws;
  p.ldws = epi->ldws;
  p.counters = epi->tile_counters;
  {
    static int ws_clean = -1;
    // measured: re-zeroing the tile from the fix-up CTA (load + store to the same lines) costs
    // 6-20 us per GEMM on the MNIST shapes, a separate fill of the workspace is cheaper -> opt-in
    if (ws_clean < 0) ws_clean = getenv("DTF_GEMM_WS_CLEAN") ? 1 : 0;
    p.ws_clean = ws_clean;
  }
  if (epi->pool_out != nullptr) {
    DTF_REQUIRE(N == 49 * 64 && !epi->apply_enable && !epi->scatter_enable && !epi->reduce && !use_pair,
                "gemm: the pool-scatter epilogue is for the [B, 7*7*64] pooled-map gradient (N = %d)", N);
    DTF_REQUIRE(epi->pool_amax != nullptr && epi->pool_dbias != nullptr && epi->mask != nullptr &&
                    epi->ldmask % 8 == 0 && epi->ldmask >= N &&
                    (reinterpret_cast<uintptr_t>(epi->pool_out) & 15) == 0 &&
                    (reinterpret_cast<uintptr_t>(epi->pool_amax) & 15) == 0 &&
                    (reinterpret_cast<uintptr_t>(epi->mask) & 15) == 0,
                "gemm: pool-scatter needs amax, the bias-gradient accumulator and the pooled activation as mask (16-byte aligned)");
    p.pool_out = reinterpret_cast<__nv_bfloat16*>(epi->pool_out);
    p.pool_amax = epi->pool_amax;
    p.pool_dbias = epi->pool_dbias;
    p.n_store = N;
  }
  if (epi->scatter_enable) {
    DTF_REQUIRE(!epi->apply_enable && !epi->reduce && p.splits == 1,
                "gemm: the scatter epilogue cannot be combined with fused apply, reduce or split-K");
    DTF_REQUIRE(epi->scatter_rows_per_slice > 0 && ceil_div(M, epi->scatter_rows_per_slice) <= 8 &&
                    epi->scatter_depth >= 1 && epi->scatter_seq != nullptr,
                "gemm: scatter needs 1..8 row slices, depth >= 1 and the sequence word");
    p.scatter_enable = 1;
    p.scatter_rows_per_slice = epi->scatter_rows_per_slice;
    p.scatter_depth = epi->scatter_depth;
    p.scatter_seq = epi->scatter_seq;
    for (int i = 0; i < ceil_div(M, epi->scatter_rows_per_slice); ++i) {
      DTF_REQUIRE(epi->scatter_base[i] != nullptr &&
                      (reinterpret_cast<uintptr_t>(epi->scatter_base[i]) & 15) == 0 &&
                      epi->scatter_buf_stride[i] % 16 == 0,
                  "gemm: scatter slice %d needs a 16-byte aligned base and stride", i);
      p.scatter_base[i] = epi->scatter_base[i];
      p.scatter_buf_stride[i] = epi->scatter_buf_stride[i];
    }
    p.out_f32 = 1;
    p.n_store = N;
  }
  p.apply_opt = epi->apply_enable ? epi->apply_opt : -1;
  if (epi->apply_enable) {
    DTF_REQUIRE(p.apply_opt >= 0, "gemm: unknown optimizer %d", p.apply_opt);
    DTF_REQUIRE(p.splits == 1, "gemm: the fused apply epilogue cannot be combined with split-K");
    DTF_REQUIRE(epi->apply_rows_per_slice > 0 && ceil_div(M, epi->apply_rows_per_slice) <= 8,
                "gemm: fused apply needs 1..8 row slices (rows_per_slice=%d, M=%d)",
                epi->apply_rows_per_slice, M);
    DTF_REQUIRE(p.apply_opt <= DTF_OPT_ADAM, "gemm: unknown optimizer %d", p.apply_opt);
    DTF_REQUIRE(p.apply_opt != DTF_OPT_ADAM || epi->apply_ctrl_powers != nullptr,
                "gemm: fused Adam needs the control block with the beta powers");
    p.apply_rows_per_slice = epi->apply_rows_per_slice;
    p.apply_shadow_ld = epi->apply_shadow_ld;
    p.ah.lr = epi->apply_lr; p.ah.beta1 = epi->apply_beta1; p.ah.beta2 = epi->apply_beta2;
    p.ah.epsilon = epi->apply_epsilon; p.ah.momentum = epi->apply_momentum;
    p.ah.grad_scale = epi->apply_grad_scale; p.ah.use_nesterov = epi->apply_use_nesterov;
    p.apply_ctrl_powers = epi->apply_ctrl_powers;
    for (int i = 0; i < 8; ++i) {
      p.avar[i] = epi->apply_var[i];
      p.aslot0[i] = epi->apply_slot0[i];
      p.aslot1[i] = epi->apply_slot1[i];
      p.ashadow[i] = reinterpret_cast<__nv_bfloat16*>(epi->apply_shadow[i]);
    }
    p.n_store = N;
    for (int i = 0; i < ceil_div(M, p.apply_rows_per_slice); ++i)
      DTF_REQUIRE(p.avar[i] != nullptr, "gemm: fused apply slice %d has no variable pointer", i);
  }

  if (use_pair) {
    if (!a_mn && !b_mn) return launch_gemm_pair<false, false>(ta, tb, p, stream);
    if (!a_mn && b_mn) return launch_gemm_pair<false, true>(ta, tb, p, stream);
    if (a_mn && !b_mn) return launch_gemm_pair<true, false>(ta, tb, p, stream);
    return launch_gemm_pair<true, true>(ta, tb, p, stream);
  }
#define DTF_DISPATCH(BN_, AM_, BM_) return launch_gemm<BN_, AM_, BM_>(ta, tb, p, stream)
  if (BN == 64) {
    if (!a_mn && !b_mn) DTF_DISPATCH(64, false, false);
    if (!a_mn && b_mn) DTF_DISPATCH(64, false, true);
    if (a_mn && !b_mn) DTF_DISPATCH(64, true, false);
    DTF_DISPATCH(64, true, true);
  } else {
    if (!a_mn && !b_mn) DTF_DISPATCH(128, false, false);
    if (!a_mn && b_mn) DTF_DISPATCH(128, false, true);
    if (a_mn && !b_mn) DTF_DISPATCH(128, true, false);
    DTF_DISPATCH(128, true, true);
  }
#undef DTF_DISPATCH
}
