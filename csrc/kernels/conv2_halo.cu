// conv2 forward / input gradient with ONE halo tile per CTA (K6/K7, run_mnist_distributed.py:59-65).
//
// conv2_umma.cu feeds every filter tap with its own TMA box: 25 x (8..16 KB activations + 4 KB
// weights) per 128-pixel tile, 67 MB of L2 -> shared-memory traffic per pass for 2.6 GFLOP - the
// kernels are bound by that traffic (ncu: tensor pipe 16-19% active). Here a CTA owns one image:
//   * ONE 5-D TMA load brings the image's whole zero-bordered activation (20 x 18 pixels; rows
//     18, 19 are out-of-bounds zero fill) into shared memory, re-laid as [channel chunk][y][x][8 ch]
//     (the tensor map's dimension order does the transposition), i.e. UMMA's un-swizzled K-major
//     canonical layout: a core matrix is 8 consecutive pixels x 16 bytes, contiguous;
//   * the 25 taps are 25 SHIFTED descriptors into that tile: start address + (r * 18 + s) pixels,
//     SBO = one image row (288 B), LBO = one channel-chunk plane. An M = 128 tile is 16 rows x 8
//     columns of the image (two tiles per image; rows 14, 15 and columns 14, 15 are junk and never
//     stored), so a 2x2 pooling window still sits inside one warp of the TMEM epilogue;
//   * the whole 5x5x32x64 filter (100 KB) is loaded once per CTA, one TMA box per filter row.
// 128 CTAs x 123 KB = 16 MB of L2 traffic instead of 67 MB, 100 (fwd) / 200 (dgrad) MMAs per CTA
// issued back to back from one resident tile.
#include "common.cuh"
#include "ptx.cuh"
#include "../../include/dtf_b200.h"

namespace dtf {

constexpr int HL_THREADS = 256;
constexpr int HL_ROWS = 20, HL_COLS = 18;                 // halo tile, pixels
constexpr int HL_ROW_B = HL_COLS * 16;                    // 288: one image row of one channel chunk
constexpr int HL_PLANE_B = HL_ROWS * HL_ROW_B;            // 5760: one channel chunk (8 channels)
constexpr int HL_W_BYTES = 800 * 128;                     // all 25 taps: [tap*32 + cin][64 cout] bf16
constexpr int HL_W_ROW_BYTES = 5 * 32 * 128;              // one filter row (5 taps)
constexpr int HL_FWD_A = 4 * HL_PLANE_B;                  // 32 input channels
constexpr int HL_DG_A = 8 * HL_PLANE_B;                   // 64 gradient channels

struct HaloParams {
  int B;
  const float* bias;
  __nv_bfloat16* out;
  unsigned char* amax;
};

__device__ __forceinline__ void tma_load_5d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0,
                                            int c1, int c2, int c3, int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6, %7}], [%2];" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
      : "memory");
}

struct HaloSmem {
  uint8_t* w;        // 1024-byte aligned (128B-swizzled filter)
  uint8_t* a;        // halo tile
  uint64_t* a_bar;
  uint64_t* w_bar;   // [5] one per filter row
  uint64_t* done_bar;
  uint32_t* tmem_slot;
};

template <int A_BYTES>
__device__ __forceinline__ HaloSmem halo_carve(uint8_t* raw) {
  HaloSmem s;
  s.w = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~static_cast<uintptr_t>(1023));
  s.a = s.w + HL_W_BYTES;
  s.a_bar = reinterpret_cast<uint64_t*>(s.a + A_BYTES);
  s.w_bar = s.a_bar + 1;
  s.done_bar = s.w_bar + 5;
  s.tmem_slot = reinterpret_cast<uint32_t*>(s.done_bar + 1);
  return s;
}

template <int A_BYTES>
constexpr int halo_smem_bytes() { return HL_W_BYTES + A_BYTES + 1024 + 128; }

__device__ __forceinline__ uint32_t halo_setup(const HaloSmem& s, int warp, uint32_t tmem_cols,
                                               const CUtensorMap* ta, const CUtensorMap* tw) {
  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(ta);
    tma_prefetch_desc(tw);
  }
  if (warp == 1 && elect_one()) {
    mbar_init(s.a_bar, 1);
    for (int i = 0; i < 5; ++i) mbar_init(&s.w_bar[i], 1);
    mbar_init(s.done_bar, 1);
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(s.tmem_slot, tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  griddep_wait();
  griddep_launch();
  return *s.tmem_slot;
}

__device__ __forceinline__ void halo_teardown(int warp, uint32_t tmem_base, uint32_t tmem_cols) {
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, tmem_cols);
  }
}

// ---------------------------------------------------------------------------------------------
// forward: conv 32 -> 64 + bias + ReLU + 2x2 max-pool (+ argmax)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(HL_THREADS, 1)
conv2_fwd_halo_kernel(const __grid_constant__ CUtensorMap tmap_x,   // pool1p as {8, 18, 18, 4, B}
                      const __grid_constant__ CUtensorMap tmap_w,   // w2 [800, 64], box {64, 160}
                      const HaloParams p) {
  extern __shared__ uint8_t smem_raw[];
  const HaloSmem sm = halo_carve<HL_FWD_A>(smem_raw);
  const int warp = threadIdx.x >> 5;
  const uint32_t tmem_base = halo_setup(sm, warp, 128, &tmap_x, &tmap_w);
  const int n = blockIdx.x;

  if (warp == 0) {
    if (elect_one()) {
      mbar_arrive_expect_tx(sm.a_bar, HL_FWD_A);
      tma_load_5d(sm.a, &tmap_x, sm.a_bar, 0, 0, 0, 0, n);
      for (int r = 0; r < 5; ++r) {
        mbar_arrive_expect_tx(&sm.w_bar[r], HL_W_ROW_BYTES);
        tma_load_2d(sm.w + r * HL_W_ROW_BYTES, &tmap_w, &sm.w_bar[r], 0, r * 160);
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(128, 64, /*a_mn=*/false, /*b_mn=*/true);
      const uint32_t a_base = smem_u32(sm.a);
      const uint32_t w_base = smem_u32(sm.w);
      mbar_wait(sm.a_bar, 0);
      for (int r = 0; r < 5; ++r) {
        mbar_wait(&sm.w_bar[r], 0);
        tc_fence_after();
#pragma unroll 1
        for (int s = 0; s < 5; ++s) {
          const int tap = r * 5 + s;
#pragma unroll
          for (int tile = 0; tile < 2; ++tile) {
#pragma unroll
            for (int kk = 0; kk < 2; ++kk) {
              // A: un-swizzled K-major; core matrix = 8 pixels x 16 B, next 8-row group = next image
              // row (SBO), next 8 channels = next plane (LBO)
              const uint64_t adesc = umma_smem_desc(
                  a_base + (2 * kk) * HL_PLANE_B + (r * HL_COLS + 8 * tile + s) * 16, HL_PLANE_B, HL_ROW_B, 0);
              // B: MN-major (cout contiguous), 128B swizzle, 16 K-rows (cin) per slice
              const uint64_t bdesc = umma_smem_desc(w_base + tap * 4096 + kk * 2048, 8192, 1024, 2);
              umma_bf16(tmem_base + tile * 64, adesc, bdesc, idesc, (tap > 0 || kk > 0) ? 1u : 0u);
            }
          }
        }
      }
      umma_commit(sm.done_bar);
    }
  } else if (warp >= 4) {
    const int q = warp - 4;
    const int lane = threadIdx.x & 31;
    const int xl = lane & 7, yy = lane >> 3;
    const int y = 4 * q + yy;
    mbar_wait(sm.done_bar, 0);
    tc_fence_after();
#pragma unroll 1
    for (int tc = 0; tc < 4; ++tc) {
      const int tile = tc >> 1, c = tc & 1;
      const int x = 8 * tile + xl;
      const bool writer = ((xl & 1) == 0) && ((yy & 1) == 0) && (y < 14) && (x < 14) && (n < p.B);
      uint32_t r[32];
      tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + tile * 64 + c * 32, r);
      tmem_ld_wait();
      uint32_t packed[16];
      uint32_t am[8];
      float o[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        // 2x2 window of this lane: x + 1 is lane ^ 1, the row below is lane ^ 8
        const float v0 = __uint_as_float(r[j]);
        const float v1 = __shfl_xor_sync(0xffffffffu, v0, 1);
        const float v2 = __shfl_xor_sync(0xffffffffu, v0, 8);
        const float v3 = __shfl_xor_sync(0xffffffffu, v1, 8);
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
        const size_t off = ((static_cast<size_t>(n) * 7 + (y >> 1)) * 7 + (x >> 1)) * 64 + c * 32;
        uint4* od = reinterpret_cast<uint4*>(p.out + off);
#pragma unroll
        for (int g = 0; g < 4; ++g)
          od[g] = make_uint4(packed[g * 4], packed[g * 4 + 1], packed[g * 4 + 2], packed[g * 4 + 3]);
        uint4* ad = reinterpret_cast<uint4*>(p.amax + off);
        ad[0] = make_uint4(am[0], am[1], am[2], am[3]);
        ad[1] = make_uint4(am[4], am[5], am[6], am[7]);
      }
    }
  }
  halo_teardown(warp, tmem_base, 128);
}

// ---------------------------------------------------------------------------------------------
// dgrad: dpool1[pixel, 32] = sum over taps of dconv2p window * W2[flipped tap]^T
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(HL_THREADS, 1)
conv2_dgrad_halo_kernel(const __grid_constant__ CUtensorMap tmap_dy,  // dconv2p as {8, 18, 18, 8, B}
                        const __grid_constant__ CUtensorMap tmap_w,   // w2 [800, 64], box {64, 160}
                        const HaloParams p) {
  extern __shared__ uint8_t smem_raw[];
  const HaloSmem sm = halo_carve<HL_DG_A>(smem_raw);
  const int warp = threadIdx.x >> 5;
  const uint32_t tmem_base = halo_setup(sm, warp, 64, &tmap_dy, &tmap_w);
  const int n = blockIdx.x;

  if (warp == 0) {
    if (elect_one()) {
      mbar_arrive_expect_tx(sm.a_bar, HL_DG_A);
      tma_load_5d(sm.a, &tmap_dy, sm.a_bar, 0, 0, 0, 0, n);
      for (int r = 4; r >= 0; --r) {   // the flipped filter is walked from its last row
        mbar_arrive_expect_tx(&sm.w_bar[r], HL_W_ROW_BYTES);
        tma_load_2d(sm.w + r * HL_W_ROW_BYTES, &tmap_w, &sm.w_bar[r], 0, r * 160);
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(128, 32, false, false);
      const uint32_t a_base = smem_u32(sm.a);
      const uint32_t w_base = smem_u32(sm.w);
      mbar_wait(sm.a_bar, 0);
      for (int rp = 0; rp < 5; ++rp) {
        mbar_wait(&sm.w_bar[4 - rp], 0);
        tc_fence_after();
#pragma unroll 1
        for (int sp = 0; sp < 5; ++sp) {
          const int wtap = (4 - rp) * 5 + (4 - sp);   // spatially flipped filter tap
#pragma unroll
          for (int tile = 0; tile < 2; ++tile) {
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
              const uint64_t adesc = umma_smem_desc(
                  a_base + (2 * kk) * HL_PLANE_B + (rp * HL_COLS + 8 * tile + sp) * 16, HL_PLANE_B, HL_ROW_B, 0);
              // B = W2[tap] as [N = 32 cin][K = 64 cout], K-major, 128B swizzle
              const uint64_t bdesc = umma_smem_desc(w_base + wtap * 4096 + kk * 32, 16, 1024, 2);
              umma_bf16(tmem_base + tile * 32, adesc, bdesc, idesc, (rp > 0 || sp > 0 || kk > 0) ? 1u : 0u);
            }
          }
        }
      }
      umma_commit(sm.done_bar);
    }
  } else if (warp >= 4) {
    const int q = warp - 4;
    const int lane = threadIdx.x & 31;
    const int xl = lane & 7, y = 4 * q + (lane >> 3);
    mbar_wait(sm.done_bar, 0);
    tc_fence_after();
#pragma unroll 1
    for (int tile = 0; tile < 2; ++tile) {
      const int x = 8 * tile + xl;
      uint32_t r[32];
      tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + tile * 32, r);
      tmem_ld_wait();
      if (n < p.B && x < 14 && y < 14) {
        uint4* od = reinterpret_cast<uint4*>(p.out + ((static_cast<size_t>(n) * 14 + y) * 14 + x) * 32);
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
  }
  halo_teardown(warp, tmem_base, 64);
}

// activation [B, 18, 18, C] viewed as {8 ch, x, y, C/8 chunks, image}: a box of one image lands in
// shared memory as [chunk][y][x][8 ch] - UMMA's un-swizzled K-major operand layout
static int make_halo_tmap(CUtensorMap* tm, const void* base, int B, int C) {
  uint64_t dims[5] = {8, 18, 18, static_cast<uint64_t>(C / 8), static_cast<uint64_t>(B)};
  uint64_t str[4] = {static_cast<uint64_t>(C) * 2, static_cast<uint64_t>(18) * C * 2, 16,
                     static_cast<uint64_t>(18) * 18 * C * 2};
  uint32_t box[5] = {8, HL_COLS, HL_ROWS, static_cast<uint32_t>(C / 8), 1};
  return encode_tmap_bf16(tm, base, 5, dims, str, box, 0);
}

static int make_w2_row_tmap(CUtensorMap* tm, const void* w2) {
  uint64_t dims[2] = {64, 800};
  uint64_t str[1] = {128};
  uint32_t box[2] = {64, 160};
  return encode_tmap_bf16(tm, w2, 2, dims, str, box, 128);
}

}  // namespace dtf

using namespace dtf;

extern "C" int dtf_conv2_fwd_halo(const void* pool1p, const void* w2, const float* b2, void* pool2,
                                  unsigned char* amax2, int B, void* stream) {
  CUtensorMap tx, tw;
  if (make_halo_tmap(&tx, pool1p, B, 32) != 0) return -1;
  if (make_w2_row_tmap(&tw, w2) != 0) return -1;
  HaloParams p;
  p.B = B;
  p.bias = b2;
  p.out = reinterpret_cast<__nv_bfloat16*>(pool2);
  p.amax = amax2;
  constexpr int smem = halo_smem_bytes<HL_FWD_A>();
  DTF_CHECK_CUDA(cudaFuncSetAttribute(conv2_fwd_halo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  DTF_CHECK_CUDA(launch_kernel(conv2_fwd_halo_kernel, dim3(B), dim3(HL_THREADS), smem,
                               reinterpret_cast<cudaStream_t>(stream), tx, tw, p));
  return 0;
}

extern "C" int dtf_conv2_dgrad_halo(const void* dconv2p, const void* w2, void* dpool1, int B, void* stream) {
  CUtensorMap ty, tw;
  if (make_halo_tmap(&ty, dconv2p, B, 64) != 0) return -1;
  if (make_w2_row_tmap(&tw, w2) != 0) return -1;
  HaloParams p;
  p.B = B;
  p.bias = nullptr;
  p.out = reinterpret_cast<__nv_bfloat16*>(dpool1);
  p.amax = nullptr;
  constexpr int smem = halo_smem_bytes<HL_DG_A>();
  DTF_CHECK_CUDA(cudaFuncSetAttribute(conv2_dgrad_halo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  DTF_CHECK_CUDA(launch_kernel(conv2_dgrad_halo_kernel, dim3(B), dim3(HL_THREADS), smem,
                               reinterpret_cast<cudaStream_t>(stream), ty, tw, p));
  return 0;
}
