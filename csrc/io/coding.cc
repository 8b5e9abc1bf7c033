#include "coding.h"

#include <string.h>

#if defined(__SSE4_2__)
#include <nmmintrin.h>
#endif

namespace ttdio {

static thread_local std::string g_error;
void set_error(const std::string& msg) { g_error = msg; }
const char* last_error() { return g_error.c_str(); }

// ------------------------------------------------------------------------------------------------
// CRC-32C. Hardware path: SSE4.2 crc32 instruction, 8 bytes per step. Software path: byte table.
// ------------------------------------------------------------------------------------------------
static uint32_t g_table[256];
static bool g_table_ready = false;
static void build_table() {
  for (uint32_t i = 0; i < 256; ++i) {
    uint32_t c = i;
    for (int k = 0; k < 8; ++k) c = (c & 1) ? (c >> 1) ^ 0x82f63b78u : (c >> 1);
    g_table[i] = c;
  }
  g_table_ready = true;
}

uint32_t crc32c_extend(uint32_t crc, const void* data, size_t n) {
  const unsigned char* p = static_cast<const unsigned char*>(data);
  uint32_t c = crc ^ 0xffffffffu;
#if defined(__SSE4_2__)
  uint64_t c64 = c;
  while (n >= 8) {
    uint64_t v;
    memcpy(&v, p, 8);
    c64 = _mm_crc32_u64(c64, v);
    p += 8;
    n -= 8;
  }
  c = static_cast<uint32_t>(c64);
  while (n--) c = _mm_crc32_u8(c, *p++);
#else
  if (!g_table_ready) build_table();
  while (n--) c = g_table[(c ^ *p++) & 0xff] ^ (c >> 8);
#endif
  return c ^ 0xffffffffu;
}

// software variant kept callable for the self-test (hardware and table paths must agree)
uint32_t crc32c_software(const void* data, size_t n) {
  if (!g_table_ready) build_table();
  const unsigned char* p = static_cast<const unsigned char*>(data);
  uint32_t c = 0xffffffffu;
  while (n--) c = g_table[(c ^ *p++) & 0xff] ^ (c >> 8);
  return c ^ 0xffffffffu;
}

// ------------------------------------------------------------------------------------------------
// Snappy raw-format decompressor
// ------------------------------------------------------------------------------------------------
bool snappy_uncompress(const char* in, size_t n, std::string* out) {
  const char* p = in;
  const char* end = in + n;
  uint64_t ulen = 0;
  p = get_varint64(p, end, &ulen);
  if (p == nullptr) return false;
  out->clear();
  out->reserve(ulen);
  while (p < end) {
    const unsigned char tag = static_cast<unsigned char>(*p++);
    const int type = tag & 3;
    if (type == 0) {  // literal
      size_t len = (tag >> 2) + 1;
      if (len > 60) {
        const int extra = static_cast<int>(len) - 60;
        if (p + extra > end) return false;
        len = 0;
        for (int i = 0; i < extra; ++i) len |= static_cast<size_t>(static_cast<unsigned char>(p[i])) << (8 * i);
        len += 1;
        p += extra;
      }
      if (p + len > end) return false;
      out->append(p, len);
      p += len;
    } else {
      size_t len, offset;
      if (type == 1) {
        if (p + 1 > end) return false;
        len = ((tag >> 2) & 7) + 4;
        offset = (static_cast<size_t>(tag >> 5) << 8) | static_cast<unsigned char>(*p++);
      } else if (type == 2) {
        if (p + 2 > end) return false;
        len = (tag >> 2) + 1;
        offset = static_cast<unsigned char>(p[0]) | (static_cast<size_t>(static_cast<unsigned char>(p[1])) << 8);
        p += 2;
      } else {
        if (p + 4 > end) return false;
        len = (tag >> 2) + 1;
        offset = get_fixed32(p);
        p += 4;
      }
      if (offset == 0 || offset > out->size()) return false;
      const size_t start = out->size() - offset;
      for (size_t i = 0; i < len; ++i) out->push_back((*out)[start + i]);  // may overlap: byte by byte
    }
  }
  return out->size() == ulen;
}

}  // namespace ttdio

extern "C" const char* ttdio_last_error(void) { return ttdio::last_error(); }
extern "C" uint32_t ttdio_crc32c(const void* data, uint64_t n) { return ttdio::crc32c(data, n); }
namespace ttdio { uint32_t crc32c_software(const void*, size_t); }
extern "C" uint32_t ttdio_crc32c_software(const void* data, uint64_t n) { return ttdio::crc32c_software(data, n); }
extern "C" uint32_t ttdio_crc32c_masked(const void* data, uint64_t n) { return ttdio::crc_mask(ttdio::crc32c(data, n)); }
