// K4/K5: softmax cross-entropy forward + backward in one pass (one warp per batch row).
//   sparse / mean : tf.losses.sparse_softmax_cross_entropy  (run_mnist_distributed.py:113)
//   dense / sum / clipped: -sum(y_ * log(clip(softmax, 1e-10, 1)))
//                                                (templates/00_mnist_replica.py:163-164)
// Outputs: loss (atomically accumulated scalar), dlogits (bf16, row-padded, feeds the dW / dX
// GEMMs as a TMA operand) and optionally the bias gradient of the logits layer (column sums of
// dlogits, atomically accumulated) so no separate reduction kernel is needed.
#include "common.cuh"
#include "ptx.cuh"
#include "../../include/dtf_b200.h"

namespace dtf {

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

constexpr int XENT_MAX_PER_LANE = 4;  // supports up to 128 classes

template <bool DENSE>
__global__ void __launch_bounds__(256)
softmax_xent_kernel(const float* __restrict__ logits, int ldl, const int* __restrict__ labels,
                    const float* __restrict__ ylabels, int ldy, int B, int C,
                    float* __restrict__ loss_accum, __nv_bfloat16* __restrict__ dlogits, int ldd,
                    float* __restrict__ dbias_accum) {
  __shared__ float s_loss[8];
  __shared__ float s_db[128];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + warp;
  griddep_wait();
  griddep_launch();
  if (dbias_accum != nullptr)
    for (int i = threadIdx.x; i < 128; i += blockDim.x) s_db[i] = 0.f;
  __syncthreads();

  float row_loss = 0.f;
  if (row < B) {
    float z[XENT_MAX_PER_LANE], y[XENT_MAX_PER_LANE];
    float mx = -INFINITY;
#pragma unroll
    for (int j = 0; j < XENT_MAX_PER_LANE; ++j) {
      const int c = lane + 32 * j;
      z[j] = (c < C) ? logits[static_cast<size_t>(row) * ldl + c] : -INFINITY;
      mx = fmaxf(mx, z[j]);
      if (DENSE) y[j] = (c < C) ? ylabels[static_cast<size_t>(row) * ldy + c] : 0.f;
    }
    mx = warp_max(mx);
    float e[XENT_MAX_PER_LANE], sum = 0.f;
#pragma unroll
    for (int j = 0; j < XENT_MAX_PER_LANE; ++j) {
      e[j] = (lane + 32 * j < C) ? __expf(z[j] - mx) : 0.f;
      sum += e[j];
    }
    sum = warp_sum(sum);
    const float inv = 1.f / sum;
    float g[XENT_MAX_PER_LANE];
    if (!DENSE) {
      const int lab = labels[row];
      const float invB = 1.f / static_cast<float>(B);
      float picked = 0.f;
#pragma unroll
      for (int j = 0; j < XENT_MAX_PER_LANE; ++j) {
        const int c = lane + 32 * j;
        const float p = e[j] * inv;
        g[j] = (p - (c == lab ? 1.f : 0.f)) * invB;
        if (c == lab) picked = z[j];
      }
      picked = warp_sum(picked);
      row_loss = (mx + __logf(sum) - picked) * invB;  // mean over the batch
    } else {
      // L = -sum_j y_j log(clip(p_j)); clip passes gradient only where 1e-10 <= p_j <= 1.
      float gy[XENT_MAX_PER_LANE], gsum = 0.f, l = 0.f;
#pragma unroll
      for (int j = 0; j < XENT_MAX_PER_LANE; ++j) {
        const float p = e[j] * inv;
        const bool in = (lane + 32 * j < C);
        const bool pass = in && (p >= 1e-10f);
        gy[j] = pass ? y[j] : 0.f;
        gsum += gy[j];
        if (in) l -= y[j] * __logf(fminf(fmaxf(p, 1e-10f), 1.f));
      }
      gsum = warp_sum(gsum);
      row_loss = warp_sum(l);
#pragma unroll
      for (int j = 0; j < XENT_MAX_PER_LANE; ++j) g[j] = e[j] * inv * gsum - gy[j];
    }
#pragma unroll
    for (int j = 0; j < XENT_MAX_PER_LANE; ++j) {
      const int c = lane + 32 * j;
      if (c < ldd) {
        const float gv = (c < C) ? g[j] : 0.f;
        dlogits[static_cast<size_t>(row) * ldd + c] = __float2bfloat16(gv);
        if (dbias_accum != nullptr && c < C) atomicAdd(&s_db[c], gv);
      }
    }
  }
  if (lane == 0) s_loss[warp] = row_loss;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += s_loss[i];
    atomicAdd(loss_accum, t);
  }
  if (dbias_accum != nullptr)
    for (int i = threadIdx.x; i < C; i += blockDim.x) atomicAdd(&dbias_accum[i], s_db[i]);
}

}  // namespace dtf

using namespace dtf;

extern "C" int dtf_softmax_xent_sparse(const float* logits, int ldl, const int* labels, int B,
                                       int C, float* loss_accum, void* dlogits, int ldd,
                                       float* dbias_accum, void* stream) {
  DTF_REQUIRE(C >= 1 && C <= 32 * XENT_MAX_PER_LANE && ldd <= 32 * XENT_MAX_PER_LANE,
              "softmax_xent: C=%d / ldd=%d exceed %d classes", C, ldd, 32 * XENT_MAX_PER_LANE);
  DTF_CHECK_CUDA(launch_kernel(softmax_xent_kernel<false>, dim3(ceil_div(B, 8)), dim3(256), 0,
                               reinterpret_cast<cudaStream_t>(stream), logits, ldl, labels,
                               static_cast<const float*>(nullptr), 0, B, C, loss_accum,
                               reinterpret_cast<__nv_bfloat16*>(dlogits), ldd, dbias_accum));
  return 0;
}

extern "C" int dtf_softmax_xent_dense(const float* logits, int ldl, const float* labels, int ldy,
                                      int B, int C, float* loss_accum, void* dlogits, int ldd,
                                      float* dbias_accum, void* stream) {
  DTF_REQUIRE(C >= 1 && C <= 32 * XENT_MAX_PER_LANE && ldd <= 32 * XENT_MAX_PER_LANE,
              "softmax_xent: C=%d / ldd=%d exceed %d classes", C, ldd, 32 * XENT_MAX_PER_LANE);
  DTF_CHECK_CUDA(launch_kernel(softmax_xent_kernel<true>, dim3(ceil_div(B, 8)), dim3(256), 0,
                               reinterpret_cast<cudaStream_t>(stream), logits, ldl,
                               static_cast<const int*>(nullptr), labels, ldy, B, C, loss_accum,
                               reinterpret_cast<__nv_bfloat16*>(dlogits), ldd, dbias_accum));
  return 0;
}
