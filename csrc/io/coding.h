// Little helpers shared by the host-side I/O library: varints, fixed-width little-endian integers,
// CRC-32C (Castagnoli) with the TensorFlow/LevelDB mask, and a Snappy block decompressor.
#pragma once
#include <stddef.h>
#include <stdint.h>

#include <string>

namespace ttdio {

// ---- crc32c ----------------------------------------------------------------------------------
uint32_t crc32c_extend(uint32_t crc, const void* data, size_t n);
inline uint32_t crc32c(const void* data, size_t n) { return crc32c_extend(0, data, n); }
inline uint32_t crc_mask(uint32_t crc) { return ((crc >> 15) | (crc << 17)) + 0xa282ead8u; }
inline uint32_t crc_unmask(uint32_t m) {
  const uint32_t rot = m - 0xa282ead8u;
  return (rot >> 17) | (rot << 15);
}

// ---- integer coding ----------------------------------------------------------------------------
inline void put_fixed32(std::string* dst, uint32_t v) {
  char b[4] = {char(v), char(v >> 8), char(v >> 16), char(v >> 24)};
  dst->append(b, 4);
}
inline void put_fixed64(std::string* dst, uint64_t v) {
  put_fixed32(dst, uint32_t(v));
  put_fixed32(dst, uint32_t(v >> 32));
}
inline uint32_t get_fixed32(const char* p) {
  const unsigned char* u = reinterpret_cast<const unsigned char*>(p);
  return uint32_t(u[0]) | (uint32_t(u[1]) << 8) | (uint32_t(u[2]) << 16) | (uint32_t(u[3]) << 24);
}
inline uint64_t get_fixed64(const char* p) { return uint64_t(get_fixed32(p)) | (uint64_t(get_fixed32(p + 4)) << 32); }
inline void put_varint64(std::string* dst, uint64_t v) {
  while (v >= 128) {
    dst->push_back(char((v & 0x7f) | 0x80));
    v >>= 7;
  }
  dst->push_back(char(v));
}
inline void put_varint32(std::string* dst, uint32_t v) { put_varint64(dst, v); }
// returns pointer past the varint, or nullptr on malformed input
inline const char* get_varint64(const char* p, const char* limit, uint64_t* out) {
  uint64_t r = 0;
  for (int shift = 0; shift <= 63 && p < limit; shift += 7) {
    const uint64_t byte = static_cast<unsigned char>(*p++);
    r |= (byte & 0x7f) << shift;
    if (!(byte & 0x80)) {
      *out = r;
      return p;
    }
  }
  return nullptr;
}
inline const char* get_varint32(const char* p, const char* limit, uint32_t* out) {
  uint64_t v = 0;
  p = get_varint64(p, limit, &v);
  *out = static_cast<uint32_t>(v);
  return p;
}

// ---- snappy (decompression only: files written by real TensorFlow may carry compressed blocks) --
bool snappy_uncompress(const char* in, size_t n, std::string* out);

// ---- error string -------------------------------------------------------------------------------
void set_error(const std::string& msg);

}  // namespace ttdio
