// png_out.cpp — RgbImage::save("output/test.png") (main.rs:54; the reference uses the `image` crate's PNG encoder).
// A dependency-free PNG writer: 8-bit RGB, filter 0 on every row, zlib stream of STORED deflate blocks (valid PNG,
// no compression — the image is the output of a render, written once).  Host only.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

#include "../../include/b200rt.h"

namespace {

uint32_t crc_table[256];
bool crc_ready = false;

void crc_init() {
  for (uint32_t n = 0; n < 256; ++n) {
    uint32_t c = n;
    for (int k = 0; k < 8; ++k) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
    crc_table[n] = c;
  }
  crc_ready = true;
}

uint32_t crc32(uint32_t crc, const uint8_t* p, size_t n) {
  if (!crc_ready) crc_init();
  crc = ~crc;
  for (size_t i = 0; i < n; ++i) crc = crc_table[(crc ^ p[i]) & 0xFFu] ^ (crc >> 8);
  return ~crc;
}

void put_be32(std::vector<uint8_t>& v, uint32_t x) {
  v.push_back((uint8_t)(x >> 24)); v.push_back((uint8_t)(x >> 16)); v.push_back((uint8_t)(x >> 8)); v.push_back((uint8_t)x);
}

void chunk(std::vector<uint8_t>& out, const char type[4], const std::vector<uint8_t>& data) {
  put_be32(out, (uint32_t)data.size());
  const size_t at = out.size();
  out.insert(out.end(), type, type + 4);
  out.insert(out.end(), data.begin(), data.end());
  put_be32(out, crc32(0u, out.data() + at, out.size() - at));
}

}  // namespace

extern "C" int b200rt_write_png(const char* path, const uint8_t* rgb8, int32_t width, int32_t height) {
  if (!path || !rgb8 || width <= 0 || height <= 0 || (int64_t)width * height > (1ll << 28)) return B200RT_ERR_INVALID;
  const size_t row = (size_t)width * 3, raw_n = (row + 1) * (size_t)height;
  std::vector<uint8_t> raw(raw_n);
  for (int32_t y = 0; y < height; ++y) {
    raw[(row + 1) * y] = 0;  // filter type 0 (None)
    std::memcpy(&raw[(row + 1) * y + 1], rgb8 + row * y, row);
  }
  // zlib: CMF/FLG, stored blocks of <= 65535 bytes, Adler-32
  std::vector<uint8_t> z;
  z.reserve(raw_n + raw_n / 65535 * 5 + 16);
  z.push_back(0x78); z.push_back(0x01);
  uint32_t a = 1, b = 0;
  size_t pos = 0;
  while (pos < raw_n) {
    const size_t n = raw_n - pos < 65535 ? raw_n - pos : 65535;
    z.push_back(pos + n == raw_n ? 1 : 0);  // BFINAL, BTYPE = 00
    z.push_back((uint8_t)(n & 0xFF)); z.push_back((uint8_t)(n >> 8));
    z.push_back((uint8_t)(~n & 0xFF)); z.push_back((uint8_t)((~n >> 8) & 0xFF));
    z.insert(z.end(), raw.begin() + pos, raw.begin() + pos + n);
    for (size_t i = 0; i < n; ++i) { a = (a + raw[pos + i]) % 65521u; b = (b + a) % 65521u; }
    pos += n;
  }
  put_be32(z, (b << 16) | a);
  std::vector<uint8_t> out;
  static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
  out.insert(out.end(), sig, sig + 8);
  std::vector<uint8_t> ihdr;
  put_be32(ihdr, (uint32_t)width); put_be32(ihdr, (uint32_t)height);
  ihdr.push_back(8); ihdr.push_back(2); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);  // 8-bit RGB, no interlace
  chunk(out, "IHDR", ihdr);
  chunk(out, "IDAT", z);
  chunk(out, "IEND", std::vector<uint8_t>());
  FILE* f = std::fopen(path, "wb");
  if (!f) return B200RT_ERR_IO;
  const size_t w = std::fwrite(out.data(), 1, out.size(), f);
  const int rc = std::fclose(f);
  return (w == out.size() && rc == 0) ? B200RT_OK : B200RT_ERR_IO;
}
