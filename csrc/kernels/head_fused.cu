// Fused classifier head (K1 + K3 + K4/K5 + K2 of the LAST dense layer) for heads with <= 16 classes.
//
// Both reference models end in a 10-way linear layer (run_mnist_distributed.py:69,
// templates/00_mnist_replica.py:163). As separate kernels that tail is a chain of four launches on
// the step's critical path: logits GEMM (N = 10, split-K), softmax-xent, the dX GEMM with K = 10
// and a column-sum for the previous layer's bias gradient — 22 us of kernels plus three
// kernel-to-kernel gaps for 2.6 MFLOP of math. Here ONE kernel does all of it on CUDA cores
// (N = 10 is a tenth of the narrowest UMMA tile):
//   logits = act * W + b          bf16 operands, fp32 accumulation (same numerics as the GEMM path)
//   loss, dlogits, db_last        same formulas as softmax_xent.cu
//   dact = (dlogits * W^T) masked by act > 0, db_prev = column sums of dact
// A CTA owns HEAD_ROWS batch rows; thread t owns hidden units t, t+256, ... and keeps W's 32-byte
// rows in registers while it walks them, so W is read from L2 once per phase and nothing is staged.
#include "common.cuh"
#include "ptx.cuh"
#include "../../include/dtf_b200.h"

namespace dtf {

constexpr int HEAD_ROWS = 4;
constexpr int HEAD_THREADS = 256;
constexpr int HEAD_HPT = 4;     // hidden units per thread per chunk
constexpr int HEAD_C = 16;  // classes are padded to 16 (the serving copy's row pitch)

struct HeadParams {
  const __nv_bfloat16* act; int lda; int H;
  const __nv_bfloat16* w; int ldw;           // [H, ldw = 16] serving copy of the logits kernel
  const float* bias;                          // [C] or null
  const int* labels; const float* ylabels; int ldy;
  int B, C;
  float* logits; int ldl;
  float* loss_accum;
  __nv_bfloat16* dlogits; int ldd;            // null when not training
  float* dbias_accum;                         // [C] or null
  __nv_bfloat16* dact; int lddact;            // null: no previous layer gradient wanted
  float* dprev_bias_accum;                    // [H] or null
  float* dw_accum;                            // [H, C] fp32 (zero on entry) or null: dW = act^T dlogits
};

__device__ __forceinline__ void unpack16(const uint4& a, const uint4& b, float (&w)[HEAD_C]) {
  const __nv_bfloat162* pa = reinterpret_cast<const __nv_bfloat162*>(&a);
  const __nv_bfloat162* pb = reinterpret_cast<const __nv_bfloat162*>(&b);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 x = __bfloat1622float2(pa[i]), y = __bfloat1622float2(pb[i]);
    w[2 * i] = x.x; w[2 * i + 1] = x.y;
    w[8 + 2 * i] = y.x; w[8 + 2 * i + 1] = y.y;
  }
}

template <bool DENSE>
__global__ void __launch_bounds__(HEAD_THREADS)
head_fused_kernel(const __grid_constant__ HeadParams p) {
  __shared__ float s_part[HEAD_THREADS / 32][HEAD_ROWS * HEAD_C];
  __shared__ float s_logit[HEAD_ROWS][HEAD_C];
  __shared__ float s_dl[HEAD_ROWS][HEAD_C];
  __shared__ float s_loss[HEAD_ROWS];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row0 = blockIdx.x * HEAD_ROWS;
  griddep_wait();
  griddep_launch();

  // ---- phase 1: partial logits of HEAD_ROWS rows over this thread's hidden units ----
  // Chunks of HEAD_THREADS * HEAD_HPT hidden units; every load of a chunk is issued before the first
  // FMA, so a chunk costs one L2 round trip instead of HEAD_HPT dependent ones.
  float acc[HEAD_ROWS * HEAD_C];
#pragma unroll
  for (int i = 0; i < HEAD_ROWS * HEAD_C; ++i) acc[i] = 0.f;
  for (int h0 = 0; h0 < p.H; h0 += HEAD_THREADS * HEAD_HPT) {
    uint4 wa[HEAD_HPT], wb[HEAD_HPT];
    __nv_bfloat16 av[HEAD_HPT][HEAD_ROWS];
#pragma unroll
    for (int j = 0; j < HEAD_HPT; ++j) {
      const int h = h0 + j * HEAD_THREADS + threadIdx.x;
      wa[j] = wb[j] = make_uint4(0, 0, 0, 0);
      if (h < p.H) {
        const uint4* wr = reinterpret_cast<const uint4*>(p.w + static_cast<size_t>(h) * p.ldw);
        wa[j] = wr[0];
        wb[j] = wr[1];
      }
#pragma unroll
      for (int r = 0; r < HEAD_ROWS; ++r)
        av[j][r] = (h < p.H && row0 + r < p.B) ? p.act[static_cast<size_t>(row0 + r) * p.lda + h]
                                                : __float2bfloat16(0.f);
    }
#pragma unroll
    for (int j = 0; j < HEAD_HPT; ++j) {
      float w[HEAD_C];
      unpack16(wa[j], wb[j], w);
#pragma unroll
      for (int r = 0; r < HEAD_ROWS; ++r) {
        const float a = __bfloat162float(av[j][r]);
#pragma unroll
        for (int c = 0; c < HEAD_C; ++c) acc[r * HEAD_C + c] = fmaf(a, w[c], acc[r * HEAD_C + c]);
      }
    }
  }
  // warp reduction by recursive halving: 62 shuffles for the 64 sums instead of 320; lane L ends
  // up with the warp totals of entries 2L and 2L + 1
  {
#pragma unroll
    for (int o = 16, n = HEAD_ROWS * HEAD_C; o >= 1; o >>= 1, n >>= 1) {
      const bool upper = (lane & o) != 0;
#pragma unroll
      for (int i = 0; i < HEAD_ROWS * HEAD_C / 2; ++i) {
        if (i < n / 2) {
          const float send = upper ? acc[i] : acc[i + n / 2];
          const float keep = upper ? acc[i + n / 2] : acc[i];
          acc[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
        }
      }
    }
    s_part[warp][2 * lane] = acc[0];
    s_part[warp][2 * lane + 1] = acc[1];
  }
  __syncthreads();
  if (threadIdx.x < HEAD_ROWS * HEAD_C) {
    float t = 0.f;
#pragma unroll
    for (int wdx = 0; wdx < HEAD_THREADS / 32; ++wdx) t += s_part[wdx][threadIdx.x];
    const int c = threadIdx.x % HEAD_C;
    if (p.bias != nullptr && c < p.C) t += p.bias[c];
    s_logit[threadIdx.x / HEAD_C][c] = t;
  }
  __syncthreads();

  // ---- phase 2: softmax cross-entropy of row `warp` (lane = class) ----
  const bool train = p.dlogits != nullptr;
  if (warp < HEAD_ROWS) {
    const int row = row0 + warp;
    float row_loss = 0.f, g = 0.f;
    const bool in = lane < p.C;
    if (row < p.B) {
      const float z = in ? s_logit[warp][lane & (HEAD_C - 1)] : -INFINITY;
      if (in) p.logits[static_cast<size_t>(row) * p.ldl + lane] = z;
      float mx = z;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      const float e = in ? __expf(z - mx) : 0.f;
      float sum = e;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
      const float prob = e / sum;
      if (!DENSE) {
        const int lab = p.labels[row];
        const float invB = 1.f / static_cast<float>(p.B);
        g = (prob - (lane == lab ? 1.f : 0.f)) * invB;
        float picked = (lane == lab) ? z : 0.f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) picked += __shfl_xor_sync(0xffffffffu, picked, o);
        row_loss = (mx + __logf(sum) - picked) * invB;
      } else {
        const float y = in ? p.ylabels[static_cast<size_t>(row) * p.ldy + lane] : 0.f;
        const float gy = (in && prob >= 1e-10f) ? y : 0.f;
        float gsum = gy, l = in ? -y * __logf(fminf(fmaxf(prob, 1e-10f), 1.f)) : 0.f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          gsum += __shfl_xor_sync(0xffffffffu, gsum, o);
          l += __shfl_xor_sync(0xffffffffu, l, o);
        }
        row_loss = l;
        g = prob * gsum - gy;
      }
      if (!in) g = 0.f;
      if (train) {
        if (lane < p.ldd) p.dlogits[static_cast<size_t>(row) * p.ldd + lane] = __float2bfloat16(g);
        if (p.dbias_accum != nullptr && in) atomicAdd(&p.dbias_accum[lane], g);
      }
    }
    // the dX product below uses the bf16-rounded dlogits, like the GEMM it replaces
    if (lane < HEAD_C) s_dl[warp][lane] = __bfloat162float(__float2bfloat16(g));
    if (lane == 0) s_loss[warp] = row_loss;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
#pragma unroll
    for (int r = 0; r < HEAD_ROWS; ++r) t += s_loss[r];
    atomicAdd(p.loss_accum, t);
  }
  if (!train || (p.dact == nullptr && p.dw_accum == nullptr)) return;

  // ---- phase 3: dact[r, h] = relu'(act) * sum_c dl[r, c] W[h, c];  db_prev[h] = sum_r dact ----
  float dl[HEAD_ROWS][HEAD_C];
#pragma unroll
  for (int r = 0; r < HEAD_ROWS; ++r)
#pragma unroll
    for (int c = 0; c < HEAD_C; ++c) dl[r][c] = s_dl[r][c];
  for (int h0 = 0; h0 < p.H; h0 += HEAD_THREADS * HEAD_HPT) {
    uint4 wa[HEAD_HPT], wb[HEAD_HPT];
    __nv_bfloat16 av[HEAD_HPT][HEAD_ROWS];
#pragma unroll
    for (int j = 0; j < HEAD_HPT; ++j) {     // L1 / L2 hits: phase 1 just read the same lines
      const int h = h0 + j * HEAD_THREADS + threadIdx.x;
      wa[j] = wb[j] = make_uint4(0, 0, 0, 0);
      if (h < p.H) {
        const uint4* wr = reinterpret_cast<const uint4*>(p.w + static_cast<size_t>(h) * p.ldw);
        wa[j] = wr[0];
        wb[j] = wr[1];
      }
#pragma unroll
      for (int r = 0; r < HEAD_ROWS; ++r)
        av[j][r] = (h < p.H && row0 + r < p.B) ? p.act[static_cast<size_t>(row0 + r) * p.lda + h]
                                                : __float2bfloat16(0.f);
    }
#pragma unroll
    for (int j = 0; j < HEAD_HPT; ++j) {
      const int h = h0 + j * HEAD_THREADS + threadIdx.x;
      if (h >= p.H) continue;
      float w[HEAD_C];
      unpack16(wa[j], wb[j], w);
      float colsum = 0.f;
#pragma unroll
      for (int r = 0; r < HEAD_ROWS; ++r) {
        if (row0 + r >= p.B) continue;
        float d = 0.f;
#pragma unroll
        for (int c = 0; c < HEAD_C; ++c) d = fmaf(dl[r][c], w[c], d);
        const __nv_bfloat16 o = __float2bfloat16(__bfloat162float(av[j][r]) > 0.f ? d : 0.f);
        if (p.dact != nullptr) p.dact[static_cast<size_t>(row0 + r) * p.lddact + h] = o;
        colsum += __bfloat162float(o);
      }
      if (p.dprev_bias_accum != nullptr) atomicAdd(&p.dprev_bias_accum[h], colsum);
      if (p.dw_accum != nullptr) {
        // this CTA's rows of dW[h, :] = sum_r act[r, h] * dlogits[r, :]
        float* dwr = p.dw_accum + static_cast<size_t>(h) * p.C;
#pragma unroll
        for (int c = 0; c < HEAD_C; ++c) {
          if (c >= p.C) break;
          float t = 0.f;
#pragma unroll
          for (int r = 0; r < HEAD_ROWS; ++r) t = fmaf(__bfloat162float(av[j][r]), dl[r][c], t);
          red_add_f32_gpu(dwr + c, t);
        }
      }
    }
  }
}

}  // namespace dtf

using namespace dtf;

extern "C" int dtf_head_fused(const void* act, int lda, int H, const void* w_shadow, int ldw,
                              const float* bias, const int* labels_sparse, const float* labels_dense,
                              int ldy, int B, int C, float* logits, int ldl, float* loss_accum,
                              void* dlogits, int ldd, float* dbias_accum, void* dact, int lddact,
                              float* dprev_bias_accum, float* dw_accum, void* stream) {
  DTF_REQUIRE(C >= 1 && C <= HEAD_C && ldw == HEAD_C, "head_fused: needs C <= %d and a %d-wide serving copy (C=%d ldw=%d)",
              HEAD_C, HEAD_C, C, ldw);
  DTF_REQUIRE((reinterpret_cast<uintptr_t>(w_shadow) & 15) == 0, "head_fused: W is not 16-byte aligned");
  DTF_REQUIRE((labels_sparse != nullptr) != (labels_dense != nullptr), "head_fused: exactly one label form");
  DTF_REQUIRE(B > 0 && H > 0 && act && logits && loss_accum, "head_fused: bad arguments");
  DTF_REQUIRE(dlogits == nullptr || ldd <= 32, "head_fused: ldd=%d", ldd);
  HeadParams p;
  p.act = reinterpret_cast<const __nv_bfloat16*>(act); p.lda = lda; p.H = H;
  p.w = reinterpret_cast<const __nv_bfloat16*>(w_shadow); p.ldw = ldw;
  p.bias = bias; p.labels = labels_sparse; p.ylabels = labels_dense; p.ldy = ldy;
  p.B = B; p.C = C; p.logits = logits; p.ldl = ldl; p.loss_accum = loss_accum;
  p.dlogits = reinterpret_cast<__nv_bfloat16*>(dlogits); p.ldd = ldd; p.dbias_accum = dbias_accum;
  p.dact = reinterpret_cast<__nv_bfloat16*>(dact); p.lddact = lddact; p.dprev_bias_accum = dprev_bias_accum;
  p.dw_accum = dw_accum;
  const dim3 grid(ceil_div(B, HEAD_ROWS));
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (labels_dense != nullptr) {
    DTF_CHECK_CUDA(launch_kernel(head_fused_kernel<true>, grid, dim3(HEAD_THREADS), 0, s, p));
  } else {
    DTF_CHECK_CUDA(launch_kernel(head_fused_kernel<false>, grid, dim3(HEAD_THREADS), 0, s, p));
  }
  return 0;
}
