// crc32c (Castagnoli polynomial 0x1EDC6F41, reflected 0x82F63B78) with the SSE4.2 crc32
// instruction when available and a slice-by-8 table fallback. Used for TFRecord framing
// (TensorBoard event files), LevelDB-format block trailers and tensor-bundle entry checksums —
// what TensorFlow's core/lib/hash/crc32c.cc provides to the reference.
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#include "../../include/dtf_b200_io.h"

#if defined(__x86_64__) || defined(__i386__)
#include <cpuid.h>
#include <nmmintrin.h>
#define DTF_X86 1
#endif

namespace {

uint32_t g_table[8][256];
bool g_table_ready = false;

void init_table() {
  if (g_table_ready) return;
  for (uint32_t i = 0; i < 256; ++i) {
    uint32_t c = i;
    for (int k = 0; k < 8; ++k) c = (c & 1) ? (c >> 1) ^ 0x82F63B78u : (c >> 1);
    g_table[0][i] = c;
  }
  for (uint32_t i = 0; i < 256; ++i)
    for (int t = 1; t < 8; ++t)
      g_table[t][i] = (g_table[t - 1][i] >> 8) ^ g_table[0][g_table[t - 1][i] & 0xFF];
  g_table_ready = true;
}

uint32_t extend_sw(uint32_t crc, const uint8_t* p, size_t n) {
  init_table();
  uint32_t c = crc ^ 0xFFFFFFFFu;
  while (n >= 8) {
    uint32_t lo, hi;
    memcpy(&lo, p, 4);
    memcpy(&hi, p + 4, 4);
    lo ^= c;
    c = g_table[7][lo & 0xFF] ^ g_table[6][(lo >> 8) & 0xFF] ^ g_table[5][(lo >> 16) & 0xFF] ^
        g_table[4][lo >> 24] ^ g_table[3][hi & 0xFF] ^ g_table[2][(hi >> 8) & 0xFF] ^
        g_table[1][(hi >> 16) & 0xFF] ^ g_table[0][hi >> 24];
    p += 8;
    n -= 8;
  }
  while (n--) c = g_table[0][(c ^ *p++) & 0xFF] ^ (c >> 8);
  return c ^ 0xFFFFFFFFu;
}

#ifdef DTF_X86
bool have_sse42() {
  static int cached = -1;
  if (cached < 0) {
    unsigned a, b, c, d;
    cached = (__get_cpuid(1, &a, &b, &c, &d) && (c & bit_SSE4_2)) ? 1 : 0;
  }
  return cached == 1;
}

__attribute__((target("sse4.2"))) uint32_t extend_hw(uint32_t crc, const uint8_t* p, size_t n) {
  uint64_t c = crc ^ 0xFFFFFFFFu;
  while (n && (reinterpret_cast<uintptr_t>(p) & 7)) {
    c = _mm_crc32_u8(static_cast<uint32_t>(c), *p++);
    --n;
  }
  while (n >= 8) {
    uint64_t v;
    memcpy(&v, p, 8);
    c = _mm_crc32_u64(c, v);
    p += 8;
    n -= 8;
  }
  while (n--) c = _mm_crc32_u8(static_cast<uint32_t>(c), *p++);
  return static_cast<uint32_t>(c) ^ 0xFFFFFFFFu;
}
#endif

}  // namespace

extern "C" uint32_t dtf_crc32c_extend(uint32_t crc, const void* data, size_t n) {
  const uint8_t* p = static_cast<const uint8_t*>(data);
#ifdef DTF_X86
  if (have_sse42()) return extend_hw(crc, p, n);
#endif
  return extend_sw(crc, p, n);
}

extern "C" uint32_t dtf_crc32c(const void* data, size_t n) { return dtf_crc32c_extend(0, data, n); }

// crc32c(A || B) from crc32c(A), crc32c(B) and len(B): multiplication by x^(8 len2) modulo the
// Castagnoli polynomial, done with GF(2) matrix squaring (the scheme zlib uses for crc32_combine).
// Lets every PS task checksum the slice of a tensor it wrote while the chief, which writes the index,
// never touches the bytes.
namespace {
uint32_t gf2_times(const uint32_t* mat, uint32_t vec) {
  uint32_t sum = 0;
  for (int i = 0; vec; vec >>= 1, ++i)
    if (vec & 1) sum ^= mat[i];
  return sum;
}
void gf2_square(uint32_t* square, const uint32_t* mat) {
  for (int n = 0; n < 32; ++n) square[n] = gf2_times(mat, mat[n]);
}
}  // namespace

extern "C" uint32_t dtf_crc32c_combine(uint32_t crc1, uint32_t crc2, uint64_t len2) {
  if (len2 == 0) return crc1;
  uint32_t even[32], odd[32];
  odd[0] = 0x82F63B78u;  // reflected Castagnoli polynomial: the operator for one zero bit
  uint32_t row = 1;
  for (int n = 1; n < 32; ++n) {
    odd[n] = row;
    row <<= 1;
  }
  gf2_square(even, odd);  // two zero bits
  gf2_square(odd, even);  // four zero bits
  do {                    // apply len2 zero BYTES to crc1
    gf2_square(even, odd);
    if (len2 & 1) crc1 = gf2_times(even, crc1);
    len2 >>= 1;
    if (len2 == 0) break;
    gf2_square(odd, even);
    if (len2 & 1) crc1 = gf2_times(odd, crc1);
    len2 >>= 1;
  } while (len2 != 0);
  return crc1 ^ crc2;
}

extern "C" uint32_t dtf_crc32c_mask(uint32_t crc) {
  return ((crc >> 15) | (crc << 17)) + 0xa282ead8u;
}

extern "C" uint32_t dtf_crc32c_unmask(uint32_t masked) {
  uint32_t rot = masked - 0xa282ead8u;
  return (rot >> 17) | (rot << 15);
}

extern "C" int dtf_crc32c_is_accelerated(void) {
#ifdef DTF_X86
  return have_sse42() ? 1 : 0;
#else
  return 0;
#endif
}
