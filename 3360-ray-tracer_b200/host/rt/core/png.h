// rt::core — a dependency-free PNG writer for the 8-bit image the resolve kernel returns (a binary
// output next to the reference's ASCII "P3", src/core/color.h:18-33).
//
// 8-bit RGB, no interlace, filter 0 on every scanline, the zlib stream made of stored (uncompressed) deflate blocks:
// the file is W·H·3 bytes + 0.01 %, written at memory speed, and every PNG reader accepts it.  Compression is not the
// point — not paying 12 characters of text per pixel is.
#pragma once
#include <array>
#include <cstddef>
#include <cstdint>
#include <stdexcept>
#include <string>

namespace rt::core {

namespace png_detail {

inline const std::array<uint32_t, 256>& CrcTable() {
  static const std::array<uint32_t, 256> table = [] {
    std::array<uint32_t, 256> t{};
    for (uint32_t n = 0; n < 256; ++n) {
      uint32_t c = n;
      for (int k = 0; k < 8; ++k) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
      t[n] = c;
    }
    return t;
  }();
  return table;
}

inline uint32_t Crc(const char* p, size_t n, uint32_t c = 0xFFFFFFFFu) {
  const auto& t = CrcTable();
  for (size_t i = 0; i < n; ++i) c = t[(c ^ static_cast<uint8_t>(p[i])) & 0xFFu] ^ (c >> 8);
  return c;
}

inline void Be32(std::string& s, uint32_t v) {
  s += static_cast<char>(v >> 24);
  s += static_cast<char>(v >> 16);
  s += static_cast<char>(v >> 8);
  s += static_cast<char>(v);
}

// length | type | data | crc(type + data)
inline void Chunk(std::string& out, const char type[4], const std::string& data) {
  Be32(out, static_cast<uint32_t>(data.size()));
  const size_t at = out.size();
  out.append(type, 4);
  out += data;
  Be32(out, Crc(out.data() + at, out.size() - at) ^ 0xFFFFFFFFu);
}

}  // namespace png_detail

// rgb8: W·H·3 bytes, rows top to bottom (the order rtb_resolve writes).
inline std::string EncodePng(const uint8_t* rgb8, int w, int h) {
  using namespace png_detail;
  if (!rgb8 || w <= 0 || h <= 0) throw std::invalid_argument("EncodePng: empty image");
  const size_t row = static_cast<size_t>(w) * 3;
  // the filtered image: one filter byte (0 = None) in front of every scanline
  std::string raw;
  raw.reserve((row + 1) * static_cast<size_t>(h));
  for (int y = 0; y < h; ++y) {
    raw += '\0';
    raw.append(reinterpret_cast<const char*>(rgb8) + row * static_cast<size_t>(y), row);
  }
  // zlib: CMF/FLG, stored blocks of at most 65535 bytes, Adler-32 of the uncompressed bytes
  std::string z;
  z.reserve(raw.size() + raw.size() / 65535 * 5 + 16);
  z += '\x78';
  z += '\x01';
  uint32_t a = 1, b = 0;
  for (size_t at = 0; at < raw.size();) {
    const size_t n = raw.size() - at < 65535 ? raw.size() - at : 65535;
    z += static_cast<char>(at + n == raw.size() ? 1 : 0);  // BFINAL, BTYPE = 00
    z += static_cast<char>(n & 0xFF);
    z += static_cast<char>(n >> 8);
    z += static_cast<char>(~n & 0xFF);
    z += static_cast<char>((~n >> 8) & 0xFF);
    z.append(raw, at, n);
    for (size_t i = 0; i < n;) {  // 5552 bytes is the longest run whose sums cannot overflow 32 bits
      const size_t run = n - i < 5552 ? n - i : 5552;
      for (size_t k = 0; k < run; ++k) {
        a += static_cast<uint8_t>(raw[at + i + k]);
        b += a;
      }
      a %= 65521u;
      b %= 65521u;
      i += run;
    }
    at += n;
  }
  Be32(z, (b << 16) | a);

  std::string out("\x89PNG\r\n\x1a\n", 8);
  std::string ihdr;
  Be32(ihdr, static_cast<uint32_t>(w));
  Be32(ihdr, static_cast<uint32_t>(h));
  ihdr += '\x08';  // bit depth
  ihdr += '\x02';  // colour type: RGB
  ihdr += '\0';    // deflate
  ihdr += '\0';    // adaptive filtering (filter byte per line)
  ihdr += '\0';    // no interlace
  Chunk(out, "IHDR", ihdr);
  Chunk(out, "IDAT", z);
  Chunk(out, "IEND", std::string());
  return out;
}

}  // namespace rt::core
