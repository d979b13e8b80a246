// K6/K7/K8 for the reference CNN's second conv block (run_mnist_distributed.py:59-65):
// 5x5 SAME conv 32 -> 64 on 14x14 maps, as IMPLICIT GEMMs on tcgen05.
//
// Activations are NHWC bf16 with a 2-pixel zero border ([B,18,18,C]), so SAME padding needs no
// im2col buffer: for filter tap (r,s) the A tile of a GEMM whose rows are output pixels is one
// 4-D TMA box {C, 16 x, 2 rows, 4 images} at coordinates {0, s, y0+r, n0}. A 128-row tile is
// 4 images x 2 output rows x 16 columns (columns 14,15 are junk and never stored), which maps a
// whole 2x2 pooling window into ONE warp of the TMEM epilogue (lane = row*16 + x), so
// bias + ReLU + max-pool + argmax are fused with warp shuffles.
//
//   forward : D[pixels,64]   = sum_taps A_tap[pixels,32] * W2[tap][32,64]      (+bias,ReLU,pool)
//   dgrad   : D[pixels,32]   = sum_taps dY_tap'[pixels,64] * W2[flip tap][32,64]^T
//   wgrad   : D[4 taps x 32,64] += A_taps^T[.,pixels] * dY[pixels,64]   (split over pixel tiles,
//             red.global.add.f32 into the fp32 gradient)
#include "common.cuh"
#include "ptx.cuh"
#include "../../include/dtf_b200.h"

namespace dtf {

constexpr int CV_THREADS = 256;
constexpr int CV_H = 14, CV_HP = 18, CV_C1 = 32, CV_C2 = 64;
constexpr int CV_YT = 7;  // 2-row tiles per image

struct ConvParams {
  int B;
  int tiles;            // pixel tiles = ceil(B/4) * 7
  int tiles_per_split;  // wgrad only
  const float* bias;
  __nv_bfloat16* out;   // fwd: pool2 [B,7,7,64]; dgrad: dpool1 [B,14,14,32]
  unsigned char* amax;  // fwd: [B,7,7,64]
  float* dw_accum;      // wgrad: [800,64] fp32
};

// ---------------------------------------------------------------------------------------------
// shared prologue / epilogue scaffolding
// ---------------------------------------------------------------------------------------------
struct ConvSmem {
  uint8_t* tiles;
  uint64_t* full_bar;
  uint64_t* empty_bar;
  uint64_t* tmem_full_bar;
  uint32_t* tmem_slot;
};

template <int STAGES, int STAGE_BYTES>
__device__ __forceinline__ ConvSmem conv_carve(uint8_t* smem_raw) {
  ConvSmem s;
  s.tiles = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                       ~static_cast<uintptr_t>(1023));
  s.full_bar = reinterpret_cast<uint64_t*>(s.tiles + STAGES * STAGE_BYTES);
  s.empty_bar = s.full_bar + STAGES;
  s.tmem_full_bar = s.empty_bar + STAGES;
  s.tmem_slot = reinterpret_cast<uint32_t*>(s.tmem_full_bar + 1);
  return s;
}

template <int STAGES>
__device__ __forceinline__ uint32_t conv_setup(const ConvSmem& s, int warp, uint32_t tmem_cols,
                                               const CUtensorMap* ta, const CUtensorMap* tb) {
  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(ta);
    tma_prefetch_desc(tb);
  }
  if (warp == 1 && elect_one()) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&s.full_bar[i], 1);
      mbar_init(&s.empty_bar[i], 1);
    }
    mbar_init(s.tmem_full_bar, 1);
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(s.tmem_slot, tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  // PDL: the prologue above overlapped the previous kernel; its outputs are needed from here on
  griddep_wait();
  griddep_launch();
  return *s.tmem_slot;
}

__device__ __forceinline__ void conv_teardown(int warp, uint32_t tmem_base, uint32_t tmem_cols) {
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, tmem_cols);
  }
}

// ---------------------------------------------------------------------------------------------
// forward: conv + bias + ReLU + 2x2 max-pool
// ---------------------------------------------------------------------------------------------
constexpr int FW_STAGES = 8;
constexpr int FW_A_TAP = 128 * CV_C1 * 2;        // 8 KB per tap
constexpr int FW_A_BYTES = FW_A_TAP;             // one filter tap: 8 KB
constexpr int FW_B_BYTES = 32 * CV_C2 * 2;       // 32 cin rows x 128 B = 4 KB
constexpr int FW_STAGE = FW_A_BYTES + FW_B_BYTES;
constexpr int FW_SMEM = FW_STAGES * FW_STAGE + 1024 + 256;

__global__ void __launch_bounds__(CV_THREADS, 2)
conv2_fwd_kernel(const __grid_constant__ CUtensorMap tmap_x,   // pool1p [B,18,18,32]
                 const __grid_constant__ CUtensorMap tmap_w,   // w2 [800,64]
                 const ConvParams p) {
  extern __shared__ uint8_t smem_raw[];
  const ConvSmem sm = conv_carve<FW_STAGES, FW_STAGE>(smem_raw);
  const int warp = threadIdx.x >> 5;
  const uint32_t tmem_base = conv_setup<FW_STAGES>(sm, warp, 64, &tmap_x, &tmap_w);
  const int n0 = (blockIdx.x / CV_YT) * 4;
  const int y0 = (blockIdx.x % CV_YT) * 2;

  // One ring slot per filter tap (A 8 KB + B 4 KB): 8 slots = 96 KB, so two CTAs share an SM and
  // one CTA's prologue / epilogue overlaps the other's MMAs.
  if (warp == 0) {
    if (elect_one()) {
      for (int t = 0; t < 25; ++t) {
        const int s = t % FW_STAGES;
        const uint32_t ph = (t / FW_STAGES) & 1;
        mbar_wait(&sm.empty_bar[s], ph ^ 1);
        uint8_t* a_dst = sm.tiles + s * FW_STAGE;
        uint8_t* b_dst = a_dst + FW_A_BYTES;
        mbar_arrive_expect_tx(&sm.full_bar[s], FW_STAGE);
        tma_load_4d(a_dst, &tmap_x, &sm.full_bar[s], 0, t % 5, y0 + t / 5, n0);
        tma_load_2d(b_dst, &tmap_w, &sm.full_bar[s], 0, t * 32);
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(128, 64, /*a_mn=*/false, /*b_mn=*/true);
      for (int t = 0; t < 25; ++t) {
        const int s = t % FW_STAGES;
        const uint32_t ph = (t / FW_STAGES) & 1;
        mbar_wait(&sm.full_bar[s], ph);
        tc_fence_after();
        const uint32_t a_base = smem_u32(sm.tiles + s * FW_STAGE);
        const uint32_t b_base = a_base + FW_A_BYTES;
#pragma unroll
        for (int kk = 0; kk < 2; ++kk) {
          // A: K-major, 64 B rows, 64B swizzle: 8-row groups 512 B apart.
          const uint64_t adesc = umma_smem_desc(a_base + kk * 32, 16, 512, 4);
          // B: MN-major (cout contiguous), 128B swizzle, K rows of 128 B, 16 rows per K-slice.
          const uint64_t bdesc = umma_smem_desc(b_base + kk * 2048, 8192, 1024, 2);
          umma_bf16(tmem_base, adesc, bdesc, idesc, (t > 0 || kk > 0) ? 1u : 0u);
        }
        umma_commit(&sm.empty_bar[s]);
      }
      umma_commit(sm.tmem_full_bar);
    }
  } else if (warp >= 4) {
    const int q = warp - 4;
    const int lane = threadIdx.x & 31;
    const int n = n0 + q;
    const int x = lane & 15;
    const bool writer = (lane < 16) && ((x & 1) == 0) && (x < CV_H) && (n < p.B);
    mbar_wait(sm.tmem_full_bar, 0);
    tc_fence_after();
#pragma unroll 1
    for (int c = 0; c < 2; ++c) {
      uint32_t r[32];
      tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + c * 32, r);
      tmem_ld_wait();
      uint32_t packed[16];
      uint32_t am[8];
      float o[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        // 2x2 window of this lane (row yy, column x): x^1 is lane^1, the other row is lane^16
        const float v0 = __uint_as_float(r[j]);
        const float v1 = __shfl_xor_sync(0xffffffffu, v0, 1);
        const float v2 = __shfl_xor_sync(0xffffffffu, v0, 16);
        const float v3 = __shfl_xor_sync(0xffffffffu, v1, 16);
        float best = v0;
        uint32_t bi = 0;
        if (v1 > best) { best = v1; bi = 1; }
        if (v2 > best) { best = v2; bi = 2; }
        if (v3 > best) { best = v3; bi = 3; }
        o[j] = fmaxf(best + __ldg(p.bias + c * 32 + j), 0.f);
        if ((j & 3) == 0) am[j >> 2] = bi;
        else am[j >> 2] |= bi << (8 * (j & 3));
      }
#pragma unroll
      for (int j = 0; j < 16; ++j) packed[j] = pack_bf16x2(o[2 * j], o[2 * j + 1]);
      if (writer) {
        const int py = y0 >> 1, px = x >> 1;
        const size_t o = ((static_cast<size_t>(n) * 7 + py) * 7 + px) * CV_C2 + c * 32;
        uint4* od = reinterpret_cast<uint4*>(p.out + o);
#pragma unroll
        for (int g = 0; g < 4; ++g)
          od[g] = make_uint4(packed[g * 4], packed[g * 4 + 1], packed[g * 4 + 2], packed[g * 4 + 3]);
        uint4* ad = reinterpret_cast<uint4*>(p.amax + o);
        ad[0] = make_uint4(am[0], am[1], am[2], am[3]);
        ad[1] = make_uint4(am[4], am[5], am[6], am[7]);
      }
    }
  }
  conv_teardown(warp, tmem_base, 64);
}

// ---------------------------------------------------------------------------------------------
// dgrad: dpool1[pixels, 32] = sum over taps of dconv2p box * W2[flipped tap]^T
// ---------------------------------------------------------------------------------------------
constexpr int DG_STAGES = 5;
constexpr int DG_A_BYTES = 128 * CV_C2 * 2;  // 16 KB
constexpr int DG_B_BYTES = 32 * CV_C2 * 2;   // 4 KB
constexpr int DG_STAGE = DG_A_BYTES + DG_B_BYTES;
constexpr int DG_SMEM = DG_STAGES * DG_STAGE + 1024 + 256;

__global__ void __launch_bounds__(CV_THREADS, 2)
conv2_dgrad_kernel(const __grid_constant__ CUtensorMap tmap_dy,  // dconv2p [B,18,18,64]
                   const __grid_constant__ CUtensorMap tmap_w,   // w2 [800,64], box {64,32}
                   const ConvParams p) {
  extern __shared__ uint8_t smem_raw[];
  const ConvSmem sm = conv_carve<DG_STAGES, DG_STAGE>(smem_raw);
  const int warp = threadIdx.x >> 5;
  const uint32_t tmem_base = conv_setup<DG_STAGES>(sm, warp, 32, &tmap_dy, &tmap_w);
  const int n0 = (blockIdx.x / CV_YT) * 4;
  const int y0 = (blockIdx.x % CV_YT) * 2;

  if (warp == 0) {
    if (elect_one()) {
      for (int t = 0; t < 25; ++t) {
        const int s = t % DG_STAGES;
        const uint32_t ph = (t / DG_STAGES) & 1;
        mbar_wait(&sm.empty_bar[s], ph ^ 1);
        uint8_t* a_dst = sm.tiles + s * DG_STAGE;
        uint8_t* b_dst = a_dst + DG_A_BYTES;
        mbar_arrive_expect_tx(&sm.full_bar[s], DG_STAGE);
        const int rp = t / 5, sp = t % 5;          // offsets into the padded gradient map
        const int wtap = (4 - rp) * 5 + (4 - sp);  // spatially flipped filter tap
        tma_load_4d(a_dst, &tmap_dy, &sm.full_bar[s], 0, sp, y0 + rp, n0);
        tma_load_2d(b_dst, &tmap_w, &sm.full_bar[s], 0, wtap * 32);
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(128, 32, false, false);
      for (int t = 0; t < 25; ++t) {
        const int s = t % DG_STAGES;
        const uint32_t ph = (t / DG_STAGES) & 1;
        mbar_wait(&sm.full_bar[s], ph);
        tc_fence_after();
        const uint32_t a_base = smem_u32(sm.tiles + s * DG_STAGE);
        const uint32_t b_base = a_base + DG_A_BYTES;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          const uint64_t adesc = umma_smem_desc(a_base + kk * 32, 16, 1024, 2);
          const uint64_t bdesc = umma_smem_desc(b_base + kk * 32, 16, 1024, 2);
          umma_bf16(tmem_base, adesc, bdesc, idesc, (t > 0 || kk > 0) ? 1u : 0u);
        }
        umma_commit(&sm.empty_bar[s]);
      }
      umma_commit(sm.tmem_full_bar);
    }
  } else if (warp >= 4) {
    const int q = warp - 4;
    const int lane = threadIdx.x & 31;
    const int n = n0 + q;
    const int x = lane & 15, y = y0 + (lane >> 4);
    mbar_wait(sm.tmem_full_bar, 0);
    tc_fence_after();
    uint32_t r[32];
    tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16), r);
    tmem_ld_wait();
    if (n < p.B && x < CV_H) {
      uint4* od = reinterpret_cast<uint4*>(
          p.out + ((static_cast<size_t>(n) * CV_H + y) * CV_H + x) * CV_C1);
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        uint4 o;
        o.x = pack_bf16x2(__uint_as_float(r[g * 8 + 0]), __uint_as_float(r[g * 8 + 1]));
        o.y = pack_bf16x2(__uint_as_float(r[g * 8 + 2]), __uint_as_float(r[g * 8 + 3]));
        o.z = pack_bf16x2(__uint_as_float(r[g * 8 + 4]), __uint_as_float(r[g * 8 + 5]));
        o.w = pack_bf16x2(__uint_as_float(r[g * 8 + 6]), __uint_as_float(r[g * 8 + 7]));
        od[g] = o;
      }
    }
  }
  conv_teardown(warp, tmem_base, 32);
}

// ---------------------------------------------------------------------------------------------
// wgrad: dW2[(tap,ci), co] += sum over pixels  pool1p_tap[pixel, ci] * dconv2[pixel, co]
// ---------------------------------------------------------------------------------------------
constexpr int WG_STAGES = 4;
constexpr int WG_A_TAP = 128 * CV_C1 * 2;   // 8 KB per tap
constexpr int WG_A_BYTES = 4 * WG_A_TAP;    // 4 taps = UMMA_M 128
constexpr int WG_B_BYTES = 128 * CV_C2 * 2; // 16 KB
constexpr int WG_STAGE = WG_A_BYTES + WG_B_BYTES;
constexpr int WG_SMEM = WG_STAGES * WG_STAGE + 1024 + 256;

__global__ void __launch_bounds__(CV_THREADS, 1)
conv2_wgrad_kernel(const __grid_constant__ CUtensorMap tmap_x,   // pool1p [B,18,18,32]
                   const __grid_constant__ CUtensorMap tmap_dy,  // dconv2p [B,18,18,64]
                   const ConvParams p) {
  extern __shared__ uint8_t smem_raw[];
  const ConvSmem sm = conv_carve<WG_STAGES, WG_STAGE>(smem_raw);
  const int warp = threadIdx.x >> 5;
  const uint32_t tmem_base = conv_setup<WG_STAGES>(sm, warp, 64, &tmap_x, &tmap_dy);
  const int mt = blockIdx.x;  // M tile: taps 4*mt .. 4*mt+3
  const int tile_begin = blockIdx.y * p.tiles_per_split;
  const int tile_end = min(p.tiles, tile_begin + p.tiles_per_split);
  const int num_it = max(0, tile_end - tile_begin);

  if (warp == 0) {
    if (elect_one()) {
      for (int i = 0; i < num_it; ++i) {
        const int s = i % WG_STAGES;
        const uint32_t ph = (i / WG_STAGES) & 1;
        mbar_wait(&sm.empty_bar[s], ph ^ 1);
        uint8_t* a_dst = sm.tiles + s * WG_STAGE;
        uint8_t* b_dst = a_dst + WG_A_BYTES;
        mbar_arrive_expect_tx(&sm.full_bar[s], WG_STAGE);
        const int tile = tile_begin + i;
        const int n0 = (tile / CV_YT) * 4, y0 = (tile % CV_YT) * 2;
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const int tap = min(mt * 4 + t, 24);  // rows of taps >= 25 are never stored
          tma_load_4d(a_dst + t * WG_A_TAP, &tmap_x, &sm.full_bar[s], 0, tap % 5, y0 + tap / 5, n0);
        }
        tma_load_4d(b_dst, &tmap_dy, &sm.full_bar[s], 0, 2, y0 + 2, n0);
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(128, 64, true, true);
      for (int i = 0; i < num_it; ++i) {
        const int s = i % WG_STAGES;
        const uint32_t ph = (i / WG_STAGES) & 1;
        mbar_wait(&sm.full_bar[s], ph);
        tc_fence_after();
        const uint32_t a_base = smem_u32(sm.tiles + s * WG_STAGE);
        const uint32_t b_base = a_base + WG_A_BYTES;
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          // A: MN-major (channels contiguous, 64 B), 64B swizzle; 4 taps = 4 chunks 8 KB apart.
          const uint64_t adesc = umma_smem_desc(a_base + kk * 1024, 8192, 512, 4);
          // B: MN-major (cout contiguous, 128 B), 128B swizzle.
          const uint64_t bdesc = umma_smem_desc(b_base + kk * 2048, 8192, 1024, 2);
          umma_bf16(tmem_base, adesc, bdesc, idesc, (i > 0 || kk > 0) ? 1u : 0u);
        }
        umma_commit(&sm.empty_bar[s]);
      }
      umma_commit(sm.tmem_full_bar);
    }
  } else if (warp >= 4) {
    const int q = warp - 4;
    const int lane = threadIdx.x & 31;
    const int tap = mt * 4 + q;
    if (num_it > 0) {
      mbar_wait(sm.tmem_full_bar, 0);
      tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < 2; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + c * 32, r);
        tmem_ld_wait();
        if (tap < 25) {
          float* dst = p.dw_accum + (static_cast<size_t>(tap) * 32 + lane) * CV_C2 + c * 32;
#pragma unroll
          for (int j = 0; j < 32; j += 4)
            red_add_v4_f32_gpu(dst + j, __uint_as_float(r[j]), __uint_as_float(r[j + 1]),
                               __uint_as_float(r[j + 2]), __uint_as_float(r[j + 3]));
        }
      }
    }
  }
  conv_teardown(warp, tmem_base, 64);
}

static int make_act_tmap(CUtensorMap* tm, const void* base, int B, int C, int swizzle) {
  uint64_t dims[4] = {static_cast<uint64_t>(C), CV_HP, CV_HP, static_cast<uint64_t>(B)};
  uint64_t str[3] = {static_cast<uint64_t>(C) * 2, static_cast<uint64_t>(CV_HP) * C * 2,
                     static_cast<uint64_t>(CV_HP) * CV_HP * C * 2};
  uint32_t box[4] = {static_cast<uint32_t>(C), 16, 2, 4};
  return encode_tmap_bf16(tm, base, 4, dims, str, box, swizzle);
}

static int make_w2_tmap(CUtensorMap* tm, const void* w2, int box_rows) {
  uint64_t dims[2] = {CV_C2, 800};
  uint64_t str[1] = {CV_C2 * 2};
  uint32_t box[2] = {CV_C2, static_cast<uint32_t>(box_rows)};
  return encode_tmap_bf16(tm, w2, 2, dims, str, box, 128);
}

template <typename K>
static int set_smem(K kern, int bytes) {
  DTF_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  return 0;
}

}  // namespace dtf

using namespace dtf;

extern "C" int dtf_conv2_fwd(const void* pool1p, const void* w2, const float* b2, void* pool2,
                             unsigned char* amax2, int B, void* stream) {
  CUtensorMap tx, tw;
  if (make_act_tmap(&tx, pool1p, B, CV_C1, 64) != 0) return -1;
  if (make_w2_tmap(&tw, w2, 32) != 0) return -1;
  ConvParams p;
  memset(&p, 0, sizeof(p));
  p.B = B;
  p.tiles = ceil_div(B, 4) * CV_YT;
  p.bias = b2;
  p.out = reinterpret_cast<__nv_bfloat16*>(pool2);
  p.amax = amax2;
  if (set_smem(conv2_fwd_kernel, FW_SMEM) != 0) return -1;
  DTF_CHECK_CUDA(launch_kernel(conv2_fwd_kernel, dim3(p.tiles), dim3(CV_THREADS), FW_SMEM,
                               reinterpret_cast<cudaStream_t>(stream), tx, tw, p));
  return 0;
}

extern "C" int dtf_conv2_dgrad(const void* dconv2p, const void* w2, void* dpool1, int B,
                               void* stream) {
  CUtensorMap ty, tw;
  if (make_act_tmap(&ty, dconv2p, B, CV_C2, 128) != 0) return -1;
  if (make_w2_tmap(&tw, w2, 32) != 0) return -1;
  ConvParams p;
  memset(&p, 0, sizeof(p));
  p.B = B;
  p.tiles = ceil_div(B, 4) * CV_YT;
  p.out = reinterpret_cast<__nv_bfloat16*>(dpool1);
  if (set_smem(conv2_dgrad_kernel, DG_SMEM) != 0) return -1;
  DTF_CHECK_CUDA(launch_kernel(conv2_dgrad_kernel, dim3(p.tiles), dim3(CV_THREADS), DG_SMEM,
                               reinterpret_cast<cudaStream_t>(stream), ty, tw, p));
  return 0;
}

extern "C" int dtf_conv2_wgrad(const void* pool1p, const void* dconv2p, float* dw2_accum, int B,
                               void* stream) {
  CUtensorMap tx, ty;
  if (make_act_tmap(&tx, pool1p, B, CV_C1, 64) != 0) return -1;
  if (make_act_tmap(&ty, dconv2p, B, CV_C2, 128) != 0) return -1;
  ConvParams p;
  memset(&p, 0, sizeof(p));
  p.B = B;
  p.tiles = ceil_div(B, 4) * CV_YT;
  int splits = 21;
  if (splits > p.tiles) splits = p.tiles;
  p.tiles_per_split = ceil_div(p.tiles, splits);
  splits = ceil_div(p.tiles, p.tiles_per_split);
  p.dw_accum = dw2_accum;
  if (set_smem(conv2_wgrad_kernel, WG_SMEM) != 0) return -1;
  dim3 grid(7, splits);
  DTF_CHECK_CUDA(launch_kernel(conv2_wgrad_kernel, grid, dim3(CV_THREADS), WG_SMEM,
                               reinterpret_cast<cudaStream_t>(stream), tx, ty, p));
  return 0;
}
