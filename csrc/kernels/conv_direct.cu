// K6/K7/K8 for the reference CNN's FIRST conv block (run_mnist_distributed.py:52-58) and the
// pooling gradients. conv1 has Cin = 1 (implicit-GEMM K = 25): too thin for the tensor cores, so
// it is a direct CUDA-core kernel with the whole image, the 5x5x32 filter bank and the bias in
// shared memory, fused with bias + ReLU + 2x2 max-pool (+argmax) and with the input
// normalisation uint8 -> [0,1] (dataset.py:92-100). Outputs land in the zero-bordered NHWC bf16
// layout that conv2's TMA boxes read.
#include "common.cuh"
#include "ptx.cuh"
#include "../../include/dtf_b200.h"

namespace dtf {

constexpr int C1_IMG = 28, C1_PAD = 32, C1_OUT = 14, C1_CH = 32, C1_HP = 18;
constexpr int C1_PITCH = C1_PAD + 2;  // pitch 34: the four argmax positions {0, 1, P, P+1} hit four different banks

__device__ __forceinline__ void load_image_padded(float (*img)[C1_PITCH], const unsigned char* x_u8,
                                                  const float* x_f32, float scale, int n) {
  for (int i = threadIdx.x; i < C1_PAD * C1_PAD; i += blockDim.x) {
    const int y = i / C1_PAD - 2, x = i % C1_PAD - 2;
    float v = 0.f;
    if (y >= 0 && y < C1_IMG && x >= 0 && x < C1_IMG) {
      const size_t o = static_cast<size_t>(n) * (C1_IMG * C1_IMG) + y * C1_IMG + x;
      v = (x_u8 != nullptr) ? static_cast<float>(x_u8[o]) * scale : x_f32[o] * scale;
    }
    img[i / C1_PAD][i % C1_PAD] = v;
  }
}

// Batch ingest: x may live in PINNED HOST memory (device-mapped). The image is fetched with 16-byte
// loads (49 per u8 image: one PCIe read burst instead of 784 byte loads), optionally mirrored into
// a device buffer for the backward pass, then expanded into the zero-padded float tile. This
// replaces the step's H2D copy node and the copy-engine -> SM hand-off behind it (~12 us).
__device__ __forceinline__ void ingest_image_padded(float (*img)[C1_PITCH], unsigned char* raw,
                                                    const unsigned char* x_u8, const float* x_f32,
                                                    unsigned char* copy_u8, float* copy_f32,
                                                    float scale, int n, bool do_copy) {
  const int bytes = (x_u8 != nullptr) ? C1_IMG * C1_IMG : C1_IMG * C1_IMG * 4;
  const char* src = (x_u8 != nullptr) ? reinterpret_cast<const char*>(x_u8) + static_cast<size_t>(n) * bytes
                                      : reinterpret_cast<const char*>(x_f32) + static_cast<size_t>(n) * bytes;
  char* dst = (copy_u8 != nullptr) ? reinterpret_cast<char*>(copy_u8) + static_cast<size_t>(n) * bytes
              : (copy_f32 != nullptr ? reinterpret_cast<char*>(copy_f32) + static_cast<size_t>(n) * bytes : nullptr);
  for (int i = threadIdx.x; i < bytes / 16; i += blockDim.x) {
    const uint4 v = reinterpret_cast<const uint4*>(src)[i];
    reinterpret_cast<uint4*>(raw)[i] = v;
    if (do_copy && dst != nullptr) reinterpret_cast<uint4*>(dst)[i] = v;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < C1_PAD * C1_PAD; i += blockDim.x) {
    const int y = i / C1_PAD - 2, x = i % C1_PAD - 2;
    float v = 0.f;
    if (y >= 0 && y < C1_IMG && x >= 0 && x < C1_IMG) {
      const int o = y * C1_IMG + x;
      v = (x_u8 != nullptr) ? static_cast<float>(raw[o]) * scale
                            : reinterpret_cast<const float*>(raw)[o] * scale;
    }
    img[i / C1_PAD][i % C1_PAD] = v;
  }
}

// Two CTAs per image (grid B x 2). Work item = (pooled pixel, group of 8 output channels).
__global__ void __launch_bounds__(392)
conv1_fwd_kernel(const unsigned char* __restrict__ x_u8, const float* __restrict__ x_f32,
                 float scale, const float* __restrict__ w1, const float* __restrict__ b1,
                 __nv_bfloat16* __restrict__ pool1p, unsigned char* __restrict__ amax1,
                 unsigned char* __restrict__ copy_u8, float* __restrict__ copy_f32, int ingest,
                 const uint4* __restrict__ y_src, uint4* __restrict__ y_dst, int y_vecs) {
  __shared__ float img[C1_PAD][C1_PITCH];
  __shared__ float w[25][C1_CH];
  __shared__ float bias[C1_CH];
  __shared__ __align__(16) unsigned char raw[C1_IMG * C1_IMG * 4];
  const int n = blockIdx.x;
  griddep_wait();
  griddep_launch();
  if (ingest && y_src != nullptr && blockIdx.x == gridDim.x - 1) {
    // the labels ride along: one extra CTA mirrors them host -> device (16-byte words), so the step
    // needs no copy node and no cross-stream join in front of the loss
    for (int i = threadIdx.x; i < y_vecs; i += blockDim.x) y_dst[i] = y_src[i];
    return;
  }
  // filter bank and bias first: their loads are in flight while the image crosses PCIe
  for (int i = threadIdx.x; i < 25 * C1_CH; i += blockDim.x) w[i / C1_CH][i % C1_CH] = w1[i];
  if (threadIdx.x < C1_CH) bias[threadIdx.x] = b1[threadIdx.x];
  if (ingest)
    ingest_image_padded(img, raw, x_u8, x_f32, copy_u8, copy_f32, scale, n, blockIdx.y == 0);
  else
    load_image_padded(img, x_u8, x_f32, scale, n);
  __syncthreads();

  // gridDim.y CTAs per image (1 when the batch is ingested from host memory, so every image crosses
  // PCIe once; 2 otherwise: upper / lower 7 pooled rows)
  const int part = C1_OUT * C1_OUT / gridDim.y;
  for (int item = threadIdx.x; item < part * 4; item += blockDim.x) {
    const int pp = blockIdx.y * part + (item >> 2), cg = item & 3;
    const int py = pp / C1_OUT, px = pp % C1_OUT;
    float patch[6][6];
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
      for (int j = 0; j < 6; ++j) patch[i][j] = img[2 * py + i][2 * px + j];
    float acc[4][8];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
      for (int c = 0; c < 8; ++c) acc[q][c] = 0.f;
#pragma unroll
    for (int r = 0; r < 5; ++r)
#pragma unroll
      for (int s = 0; s < 5; ++s) {
        float wv[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) wv[c] = w[r * 5 + s][cg * 8 + c];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float xv = patch[(q >> 1) + r][(q & 1) + s];
#pragma unroll
          for (int c = 0; c < 8; ++c) acc[q][c] = fmaf(xv, wv[c], acc[q][c]);
        }
      }
    uint32_t packed[4];
    uint32_t am[2] = {0, 0};
    float o[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      float best = acc[0][c];
      uint32_t bi = 0;
#pragma unroll
      for (int q = 1; q < 4; ++q)
        if (acc[q][c] > best) { best = acc[q][c]; bi = q; }
      o[c] = fmaxf(best + bias[cg * 8 + c], 0.f);
      am[c >> 2] |= bi << (8 * (c & 3));
    }
#pragma unroll
    for (int c = 0; c < 4; ++c) packed[c] = pack_bf16x2(o[2 * c], o[2 * c + 1]);
    const size_t op = ((static_cast<size_t>(n) * C1_HP + (py + 2)) * C1_HP + (px + 2)) * C1_CH + cg * 8;
    *reinterpret_cast<uint4*>(pool1p + op) = make_uint4(packed[0], packed[1], packed[2], packed[3]);
    const size_t oa = ((static_cast<size_t>(n) * C1_OUT + py) * C1_OUT + px) * C1_CH + cg * 8;
    *reinterpret_cast<uint2*>(amax1 + oa) = make_uint2(am[0], am[1]);
  }
}

// conv1 filter / bias gradient. Two CTAs of 16 warps per image (upper / lower 7 pooled rows).
// lane = output channel, warp w walks the pooled pixels w, w+16, ...: for each (pixel, channel) it
// reads the argmax position and the ReLU-masked gradient ONCE and sweeps the 25 taps with 25
// register accumulators (1 smem load per FMA; a thread-per-(tap, channel) layout needs 3). The
// image arrives with 16-byte loads and every staging loop is a single pass, so the kernel is one
// global round trip deep. The 16 warps' partial filters are combined in shared memory (two
// rounds), so a CTA issues 26 coalesced 128-byte reductions.
constexpr int C1B_WARPS = 16;
constexpr int C1_HALF = C1_OUT * C1_OUT / 2;  // 98 pooled pixels per CTA

__global__ void __launch_bounds__(C1B_WARPS * 32)
conv1_bwd_kernel(const unsigned char* __restrict__ x_u8, const float* __restrict__ x_f32,
                 float scale, const __nv_bfloat16* __restrict__ dpool1,
                 const __nv_bfloat16* __restrict__ pool1p, const unsigned char* __restrict__ amax1,
                 float* __restrict__ dw1_accum, float* __restrict__ db1_accum) {
  __shared__ float img[C1_PAD][C1_PITCH];
  __shared__ __align__(16) float g[C1_HALF][C1_CH];                   // 12.5 KB
  __shared__ __align__(16) unsigned char am[C1_HALF][C1_CH];          // 3 KB
  __shared__ __align__(16) float red[C1B_WARPS / 2][26][C1_CH];       // 26.6 KB (also the raw image)
  const int n = blockIdx.x;
  const int pp0 = blockIdx.y * C1_HALF;
  griddep_wait();
  griddep_launch();
  // (pooled pixel, 8 channels) items: issue their global loads first, the image's next
  const int item = threadIdx.x;
  uint4 av = make_uint4(0, 0, 0, 0), dv = av;
  uint2 mv = make_uint2(0, 0);
  const int ipp = item >> 2, icg = item & 3;
  if (item < C1_HALF * 4) {
    const int pp = pp0 + ipp;
    const int py = pp / C1_OUT, px = pp % C1_OUT;
    av = *reinterpret_cast<const uint4*>(
        pool1p + ((static_cast<size_t>(n) * C1_HP + (py + 2)) * C1_HP + (px + 2)) * C1_CH + icg * 8);
    const size_t o = (static_cast<size_t>(n) * (C1_OUT * C1_OUT) + pp) * C1_CH + icg * 8;
    dv = *reinterpret_cast<const uint4*>(dpool1 + o);
    mv = *reinterpret_cast<const uint2*>(amax1 + o);
  }
  ingest_image_padded(img, reinterpret_cast<unsigned char*>(red), x_u8, x_f32, nullptr, nullptr, scale,
                      n, false);
  if (item < C1_HALF * 4) {
    const __nv_bfloat16* a = reinterpret_cast<const __nv_bfloat16*>(&av);
    const __nv_bfloat16* d = reinterpret_cast<const __nv_bfloat16*>(&dv);
    float gv8[8];
#pragma unroll
    for (int j = 0; j < 8; ++j)  // ReLU gradient through the max-pool
      gv8[j] = (__bfloat162float(a[j]) > 0.f) ? __bfloat162float(d[j]) : 0.f;
    // two 16-byte stores (scalar stores here would be 8-way bank conflicts: lanes are 8 words apart)
    *reinterpret_cast<float4*>(&g[ipp][icg * 8]) = make_float4(gv8[0], gv8[1], gv8[2], gv8[3]);
    *reinterpret_cast<float4*>(&g[ipp][icg * 8 + 4]) = make_float4(gv8[4], gv8[5], gv8[6], gv8[7]);
    *reinterpret_cast<uint2*>(&am[ipp][icg * 8]) = mv;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, c = threadIdx.x & 31;
  float acc[25];
#pragma unroll
  for (int t = 0; t < 25; ++t) acc[t] = 0.f;
  float accb = 0.f;
  for (int lp = warp; lp < C1_HALF; lp += C1B_WARPS) {
    const int pp = pp0 + lp;
    const int py = pp / C1_OUT, px = pp - py * C1_OUT;
    const int a = am[lp][c];
    const float gv = g[lp][c];
    const float* base = &img[2 * py + (a >> 1)][2 * px + (a & 1)];
#pragma unroll
    for (int r = 0; r < 5; ++r)
#pragma unroll
      for (int q = 0; q < 5; ++q) acc[r * 5 + q] = fmaf(base[r * C1_PITCH + q], gv, acc[r * 5 + q]);
    accb += gv;
  }
  // cross-warp reduction in two rounds (warps 0-7 store, warps 8-15 add), then 26 x 32 sums
  if (warp < C1B_WARPS / 2) {
#pragma unroll
    for (int t = 0; t < 25; ++t) red[warp][t][c] = acc[t];
    red[warp][25][c] = accb;
  }
  __syncthreads();
  if (warp >= C1B_WARPS / 2) {
#pragma unroll
    for (int t = 0; t < 25; ++t) red[warp - C1B_WARPS / 2][t][c] += acc[t];
    red[warp - C1B_WARPS / 2][25][c] += accb;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 26 * C1_CH; i += blockDim.x) {
    const int t = i / C1_CH, cc = i % C1_CH;
    float v = 0.f;
#pragma unroll
    for (int w = 0; w < C1B_WARPS / 2; ++w) v += red[w][t][cc];
    if (t < 25) atomicAdd(&dw1_accum[t * C1_CH + cc], v);
    else atomicAdd(&db1_accum[cc], v);
  }
}

// pool2 gradient + ReLU gradient, scattered into the zero-bordered [B,18,18,64] layout conv2's
// dgrad / wgrad TMA boxes read; db2 = column sums. Work item = (pooled pixel, 8 channels).
__global__ void __launch_bounds__(256)
pool2_bwd_kernel(const __nv_bfloat16* __restrict__ dpool2, int ld, const __nv_bfloat16* __restrict__ pool2,
                 const unsigned char* __restrict__ amax2, __nv_bfloat16* __restrict__ dconv2p,
                 float* __restrict__ db2_accum, int B) {
  __shared__ float s_db[64];
  if (threadIdx.x < 64) s_db[threadIdx.x] = 0.f;
  __syncthreads();
  griddep_wait();
  griddep_launch();
  const int total = B * 49 * 8;
  const int stride = gridDim.x * blockDim.x;             // multiple of 8: cg is fixed per thread
  const int cg = threadIdx.x & 7;
  float dbs[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  for (int item = blockIdx.x * blockDim.x + threadIdx.x; item < total; item += stride) {
    const int pp = (item >> 3) % 49;
    const int n = (item >> 3) / 49;
    const int py = pp / 7, px = pp % 7;
    const uint4 dv = *reinterpret_cast<const uint4*>(dpool2 + static_cast<size_t>(n) * ld + pp * 64 + cg * 8);
    const uint4 av = *reinterpret_cast<const uint4*>(pool2 + (static_cast<size_t>(n) * 49 + pp) * 64 + cg * 8);
    const uint2 mv = *reinterpret_cast<const uint2*>(amax2 + (static_cast<size_t>(n) * 49 + pp) * 64 + cg * 8);
    const __nv_bfloat16* d = reinterpret_cast<const __nv_bfloat16*>(&dv);
    const __nv_bfloat16* a = reinterpret_cast<const __nv_bfloat16*>(&av);
    const unsigned char* m = reinterpret_cast<const unsigned char*>(&mv);
    __nv_bfloat16 out[4][8];
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const bool on = __bfloat162float(a[c]) > 0.f;
      const __nv_bfloat16 gv = on ? d[c] : __float2bfloat16(0.f);
#pragma unroll
      for (int q = 0; q < 4; ++q) out[q][c] = (m[c] == q) ? gv : __float2bfloat16(0.f);
      if (on) dbs[c] += __bfloat162float(d[c]);
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int y = 2 * py + (q >> 1) + 2, x = 2 * px + (q & 1) + 2;
      *reinterpret_cast<uint4*>(dconv2p + ((static_cast<size_t>(n) * 18 + y) * 18 + x) * 64 + cg * 8) =
          *reinterpret_cast<const uint4*>(out[q]);
    }
  }
  // bias gradient: the 4 lanes of a warp that share a channel group fold with shuffles, then one
  // shared-memory atomic per (warp, channel)
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    float v = dbs[c];
    v += __shfl_xor_sync(0xffffffffu, v, 8);
    v += __shfl_xor_sync(0xffffffffu, v, 16);
    if ((threadIdx.x & 31) < 8) atomicAdd(&s_db[cg * 8 + c], v);
  }
  __syncthreads();
  if (threadIdx.x < 64) atomicAdd(&db2_accum[threadIdx.x], s_db[threadIdx.x]);
}

}  // namespace dtf

using namespace dtf;

extern "C" int dtf_conv1_fwd(const unsigned char* x_u8, const float* x_f32, float scale,
                             const float* w1, const float* b1, void* pool1p,
                             unsigned char* amax1, int B, void* stream) {
  DTF_REQUIRE((x_u8 != nullptr) != (x_f32 != nullptr), "conv1_fwd: pass exactly one of x_u8/x_f32");
  DTF_CHECK_CUDA(launch_kernel(conv1_fwd_kernel, dim3(B, 2), dim3(196), 0,
                               reinterpret_cast<cudaStream_t>(stream), x_u8, x_f32, scale, w1, b1,
                               reinterpret_cast<__nv_bfloat16*>(pool1p), amax1,
                               static_cast<unsigned char*>(nullptr), static_cast<float*>(nullptr), 0,
                               static_cast<const uint4*>(nullptr), static_cast<uint4*>(nullptr), 0));
  return 0;
}

extern "C" int dtf_conv1_fwd_ingest(const unsigned char* x_u8, const float* x_f32, float scale,
                                    const float* w1, const float* b1, void* pool1p,
                                    unsigned char* amax1, unsigned char* copy_u8, float* copy_f32,
                                    const void* y_src, void* y_dst, int y_bytes, int B, void* stream) {
  DTF_REQUIRE(y_src == nullptr || (y_dst != nullptr && y_bytes % 16 == 0 &&
                                   ((reinterpret_cast<uintptr_t>(y_src) | reinterpret_cast<uintptr_t>(y_dst)) & 15) == 0),
              "conv1_fwd_ingest: labels need a destination, 16-byte alignment and a multiple of 16 bytes");
  DTF_REQUIRE((x_u8 != nullptr) != (x_f32 != nullptr), "conv1_fwd_ingest: pass exactly one of x_u8/x_f32");
  const void* x = x_u8 ? static_cast<const void*>(x_u8) : static_cast<const void*>(x_f32);
  DTF_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0, "conv1_fwd_ingest: x must be 16-byte aligned");
  DTF_CHECK_CUDA(launch_kernel(conv1_fwd_kernel, dim3(B + (y_src != nullptr ? 1 : 0), 1), dim3(392), 0,
                               reinterpret_cast<cudaStream_t>(stream), x_u8, x_f32, scale, w1, b1,
                               reinterpret_cast<__nv_bfloat16*>(pool1p), amax1, copy_u8, copy_f32, 1,
                               static_cast<const uint4*>(y_src), static_cast<uint4*>(y_dst), y_bytes / 16));
  return 0;
}

extern "C" int dtf_conv1_bwd(const unsigned char* x_u8, const float* x_f32, float scale,
                             const void* dpool1, const void* pool1p, const unsigned char* amax1,
                             float* dw1_accum, float* db1_accum, int B, void* stream) {
  DTF_REQUIRE((x_u8 != nullptr) != (x_f32 != nullptr), "conv1_bwd: pass exactly one of x_u8/x_f32");
  DTF_CHECK_CUDA(launch_kernel(conv1_bwd_kernel, dim3(B, 2), dim3(C1B_WARPS * 32), 0,
                               reinterpret_cast<cudaStream_t>(stream), x_u8, x_f32, scale,
                               reinterpret_cast<const __nv_bfloat16*>(dpool1),
                               reinterpret_cast<const __nv_bfloat16*>(pool1p),
                               static_cast<const unsigned char*>(amax1), dw1_accum, db1_accum));
  return 0;
}

extern "C" int dtf_pool2_bwd(const void* dpool2, int ld_dpool2, const void* pool2,
                             const unsigned char* amax2, void* dconv2p, float* db2_accum, int B,
                             void* stream) {
  DTF_REQUIRE(ld_dpool2 % 8 == 0, "pool2_bwd: ld_dpool2=%d must be a multiple of 8", ld_dpool2);
  const int total = B * 49 * 8;
  int blocks = ceil_div(total, 256);
  if (blocks > 148 * 4) blocks = 148 * 4;
  DTF_CHECK_CUDA(launch_kernel(pool2_bwd_kernel, dim3(blocks), dim3(256), 0,
                               reinterpret_cast<cudaStream_t>(stream),
                               reinterpret_cast<const __nv_bfloat16*>(dpool2), ld_dpool2,
                               reinterpret_cast<const __nv_bfloat16*>(pool2), amax2,
                               reinterpret_cast<__nv_bfloat16*>(dconv2p), db2_accum, B));
  return 0;
}
