// pm_png.hpp — minimal PNG writer for the RGBA8 framebuffer (the reference calls stbi_write_png,
// ray-tracer/src/hostCode.cu:239-240). 8-bit RGBA, filter 0, zlib stream of stored (uncompressed) blocks.
#pragma once
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

namespace pmh {

inline uint32_t crc32_update(uint32_t crc, const uint8_t* p, size_t n) {
  static uint32_t table[256];
  static bool init = false;
  if (!init) {
    for (uint32_t i = 0; i < 256; i++) {
      uint32_t c = i;
      for (int k = 0; k < 8; k++) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
      table[i] = c;
    }
    init = true;
  }
  crc = ~crc;
  for (size_t i = 0; i < n; i++) crc = table[(crc ^ p[i]) & 0xff] ^ (crc >> 8);
  return ~crc;
}

inline void be32(std::vector<uint8_t>& v, uint32_t x) {
  v.push_back((uint8_t)(x >> 24)); v.push_back((uint8_t)(x >> 16)); v.push_back((uint8_t)(x >> 8)); v.push_back((uint8_t)x);
}

inline void png_chunk(std::vector<uint8_t>& out, const char type[4], const std::vector<uint8_t>& data) {
  be32(out, (uint32_t)data.size());
  const size_t start = out.size();
  out.insert(out.end(), type, type + 4);
  out.insert(out.end(), data.begin(), data.end());
  be32(out, crc32_update(0, out.data() + start, out.size() - start));
}

// pixels: width*height RGBA8 (byte order R,G,B,A = little-endian uint32 of owl::make_rgba), row 0 first.
inline bool write_png_rgba8(const std::string& path, const uint32_t* pixels, int width, int height) {
  std::vector<uint8_t> raw;
  raw.reserve((size_t)height * ((size_t)width * 4 + 1));
  for (int y = 0; y < height; y++) {
    raw.push_back(0);  // filter type None
    const uint8_t* row = reinterpret_cast<const uint8_t*>(pixels + (size_t)y * width);
    raw.insert(raw.end(), row, row + (size_t)width * 4);
  }
  std::vector<uint8_t> z;
  z.push_back(0x78); z.push_back(0x01);
  uint32_t a = 1, b = 0;  // adler32
  size_t off = 0;
  do {
    const size_t n = std::min<size_t>(65535, raw.size() - off);
    z.push_back(off + n >= raw.size() ? 1 : 0);
    z.push_back((uint8_t)(n & 0xff)); z.push_back((uint8_t)(n >> 8));
    z.push_back((uint8_t)(~n & 0xff)); z.push_back((uint8_t)((~n >> 8) & 0xff));
    z.insert(z.end(), raw.begin() + (ptrdiff_t)off, raw.begin() + (ptrdiff_t)(off + n));
    for (size_t i = 0; i < n; i++) { a = (a + raw[off + i]) % 65521u; b = (b + a) % 65521u; }
    off += n;
  } while (off < raw.size());
  be32(z, (b << 16) | a);

  std::vector<uint8_t> out = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
  std::vector<uint8_t> ihdr;
  be32(ihdr, (uint32_t)width); be32(ihdr, (uint32_t)height);
  ihdr.push_back(8); ihdr.push_back(6); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);
  png_chunk(out, "IHDR", ihdr);
  png_chunk(out, "IDAT", z);
  png_chunk(out, "IEND", {});
  FILE* f = fopen(path.c_str(), "wb");
  if (!f) return false;
  const bool ok = fwrite(out.data(), 1, out.size(), f) == out.size();
  return fclose(f) == 0 && ok;
}

}  // namespace pmh
