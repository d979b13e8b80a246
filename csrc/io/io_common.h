// Shared helpers of libdtf_b200_host: error string, protobuf wire encoding/decoding, varints.
#pragma once
#include <stdint.h>
#include <string.h>

#include <string>

namespace dtfio {

extern thread_local char g_err[1024];
void set_error(const char* fmt, ...);

inline void put_varint(std::string* s, uint64_t v) {
  while (v >= 0x80) {
    s->push_back(static_cast<char>((v & 0x7F) | 0x80));
    v >>= 7;
  }
  s->push_back(static_cast<char>(v));
}
inline void put_tag(std::string* s, int field, int wt) { put_varint(s, (static_cast<uint64_t>(field) << 3) | wt); }
inline void put_varint_field(std::string* s, int field, uint64_t v) { put_tag(s, field, 0); put_varint(s, v); }
inline void put_bytes(std::string* s, int field, const std::string& v) {
  put_tag(s, field, 2);
  put_varint(s, v.size());
  s->append(v);
}
inline void put_string(std::string* s, int field, const char* v) { put_bytes(s, field, std::string(v)); }
inline void put_fixed32(std::string* s, uint32_t v) { s->append(reinterpret_cast<const char*>(&v), 4); }
inline void put_fixed64(std::string* s, uint64_t v) { s->append(reinterpret_cast<const char*>(&v), 8); }
inline void put_fixed32_field(std::string* s, int field, uint32_t v) { put_tag(s, field, 5); put_fixed32(s, v); }
inline void put_float(std::string* s, int field, float v) {
  uint32_t u;
  memcpy(&u, &v, 4);
  put_fixed32_field(s, field, u);
}
inline void put_double(std::string* s, int field, double v) {
  uint64_t u;
  memcpy(&u, &v, 8);
  put_tag(s, field, 1);
  put_fixed64(s, u);
}

// Returns false on malformed input.
inline bool get_varint(const uint8_t*& p, const uint8_t* end, uint64_t* out) {
  uint64_t v = 0;
  int shift = 0;
  while (p < end && shift <= 63) {
    uint8_t b = *p++;
    v |= static_cast<uint64_t>(b & 0x7F) << shift;
    if (!(b & 0x80)) {
      *out = v;
      return true;
    }
    shift += 7;
  }
  return false;
}

}  // namespace dtfio
