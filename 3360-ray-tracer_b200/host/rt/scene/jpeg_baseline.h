// rt::scene::jpeg — a baseline JPEG decoder for scene::Image (the reference loads textures through stb_image,
// src/scene/image.cc:16-40, which is not part of this repository).
//
// Scope: what ImageTexture needs — sequential Huffman JPEG (SOF0 / SOF1, 8 bits per sample), 1 or 3 components, sampling
// factors 1 or 2 per direction (4:4:4, 4:2:2, 4:4:0, 4:2:0), restart intervals, interleaved and per-component scans.
// Progressive (SOF2), arithmetic coding, 12-bit and CMYK files are refused with a message.
//
// Numerics follow the published integer pipeline the reference's decoder also implements, so that the byte table
// ImageTexture samples is the reference's, bit for bit (tests/test_host_scene.py::test_image_table_matches_reference_pipeline
// and tests/test_jpeg.py compare with the compiled reference): dequantised coefficients as 16-bit values; the
// Loeffler-Ligtenberg-Moshovitz "islow" 8x8 inverse DCT in 12-bit fixed point (columns kept with 2 extra bits: +512 >> 10;
// rows +65536 + (128 << 17) >> 17, clamped); chroma upsampled with the 3:1 triangle filters; YCbCr -> RGB in 20-bit fixed point
// with the green term's Cb product truncated to its upper 16 bits.
#pragma once
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

namespace rt::scene::jpeg {

namespace detail {

struct Huffman {
  uint8_t fast[512];       // 9-bit lookahead -> symbol index (255 = longer code)
  uint16_t code[256];
  uint8_t values[256], size[257];
  unsigned maxcode[18];
  int delta[17];
  bool Build(const uint8_t* counts) {
    int k = 0;
    for (int i = 0; i < 16; ++i)
      for (int j = 0; j < counts[i]; ++j) {
        if (k >= 256) return false;
        size[k++] = static_cast<uint8_t>(i + 1);
      }
    size[k] = 0;
    unsigned c = 0;
    k = 0;
    for (int j = 1; j <= 16; ++j) {
      delta[j] = k - static_cast<int>(c);
      if (size[k] == j) {
        while (size[k] == j) code[k++] = static_cast<uint16_t>(c++);
        if (c - 1 >= (1u << j)) return false;
      }
      maxcode[j] = c << (16 - j);
      c <<= 1;
    }
    maxcode[17] = 0xffffffffu;
    std::memset(fast, 255, sizeof(fast));
    for (int i = 0; i < k; ++i) {
      const int s = size[i];
      if (s <= 9) {
        const int c0 = code[i] << (9 - s), m = 1 << (9 - s);
        for (int j = 0; j < m; ++j) fast[c0 + j] = static_cast<uint8_t>(i);
      }
    }
    return true;
  }
};

struct Component {
  int id = 0, h = 1, v = 1, tq = 0, hd = 0, ha = 0, dc_pred = 0;
  int x = 0, y = 0, w2 = 0, h2 = 0;  // size in samples, padded size (whole blocks of whole MCUs)
  std::vector<uint8_t> data;
};

struct Decoder {
  const uint8_t* p;
  const uint8_t* end;
  std::string why;
  Huffman huff_dc[4], huff_ac[4];
  bool have_dc[4] = {false, false, false, false}, have_ac[4] = {false, false, false, false};
  uint16_t dequant[4][64];
  int width = 0, height = 0, n_comp = 0, h_max = 1, v_max = 1, mcu_x = 0, mcu_y = 0, restart_interval = 0;
  Component comp[3];
  // entropy-coded segment reader
  uint32_t code_buffer = 0;
  int code_bits = 0;
  bool hit_marker = false;
  uint8_t marker = 0xff;

  bool Fail(const char* m) {
    if (why.empty()) why = m;
    return false;
  }
  int Get8() { return p < end ? *p++ : 0; }
  int Get16() {
    const int a = Get8();
    return (a << 8) | Get8();
  }
  void Grow() {
    do {
      unsigned b = hit_marker ? 0u : static_cast<unsigned>(Get8());
      if (b == 0xff) {
        int c = Get8();
        while (c == 0xff) c = Get8();  // fill bytes
        if (c != 0) {
          marker = static_cast<uint8_t>(c);
          hit_marker = true;
          b = 0;
        }
      }
      code_buffer |= b << (24 - code_bits);
      code_bits += 8;
    } while (code_bits <= 24);
  }
  int DecodeSymbol(const Huffman& h) {
    if (code_bits < 16) Grow();
    const int look = (code_buffer >> 23) & 511;
    int k = h.fast[look];
    if (k < 255) {
      const int s = h.size[k];
      if (s > code_bits) return -1;
      code_buffer <<= s;
      code_bits -= s;
      return h.values[k];
    }
    const unsigned temp = code_buffer >> 16;
    int s = 10;
    while (temp >= h.maxcode[s]) ++s;
    if (s == 17 || s > code_bits) return -1;
    k = static_cast<int>((code_buffer >> (32 - s)) & ((1u << s) - 1)) + h.delta[s];
    if (k < 0 || k > 255) return -1;
    code_buffer <<= s;
    code_bits -= s;
    return h.values[k];
  }
  // n more bits as the signed value JPEG's "extend" gives them
  int Extend(int n) {
    if (code_bits < n) Grow();
    if (code_bits < n) return 0;
    const int sgn = static_cast<int>(code_buffer >> 31);
    const unsigned k = (code_buffer << n) | (code_buffer >> (32 - n));  // rotate left
    const unsigned mask = (1u << n) - 1u;
    code_buffer = k & ~mask;
    code_bits -= n;
    const int bias = sgn ? 0 : (-(1 << n) + 1);
    return static_cast<int>(k & mask) + bias;
  }
  void ResetEntropy() {
    code_bits = 0;
    code_buffer = 0;
    hit_marker = false;
    marker = 0xff;
    for (int i = 0; i < n_comp; ++i) comp[i].dc_pred = 0;
  }

  static const uint8_t* ZigZag() {
    static const uint8_t z[64 + 15] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13,
                                       6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31,
                                       39, 46, 53, 60, 61, 54, 47, 55, 62, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63};
    return z;
  }

  bool DecodeBlock(short data[64], Component& c) {
    const Huffman& hdc = huff_dc[c.hd];
    const Huffman& hac = huff_ac[c.ha];
    const uint16_t* dq = dequant[c.tq];
    if (code_bits < 16) Grow();
    const int t = DecodeSymbol(hdc);
    if (t < 0 || t > 15) return Fail("bad huffman code");
    std::memset(data, 0, 64 * sizeof(short));
    const int diff = t ? Extend(t) : 0;
    c.dc_pred += diff;
    data[0] = static_cast<short>(c.dc_pred * dq[0]);
    const uint8_t* zz = ZigZag();
    int k = 1;
    do {
      const int rs = DecodeSymbol(hac);
      if (rs < 0) return Fail("bad huffman code");
      const int s = rs & 15, r = rs >> 4;
      if (s == 0) {
        if (rs != 0xf0) break;  // end of block
        k += 16;
      } else {
        k += r;
        const unsigned zig = zz[k++];
        data[zig] = static_cast<short>(Extend(s) * dq[zig]);
      }
    } while (k < 64);
    return true;
  }

  static uint8_t Clamp(int x) { return static_cast<unsigned>(x) > 255u ? (x < 0 ? 0 : 255) : static_cast<uint8_t>(x); }

  // one pass of the 8-point inverse DCT (even part / odd part butterflies, constants x 4096)
  struct Pass { int x0, x1, x2, x3, t0, t1, t2, t3; };
  static constexpr int F(double x) { return static_cast<int>(x * 4096 + 0.5); }
  static Pass Idct1D(int s0, int s1, int s2, int s3, int s4, int s5, int s6, int s7) {
    Pass o;
    int p2 = s2, p3 = s6;
    int p1 = (p2 + p3) * F(0.5411961f);
    int t2 = p1 + p3 * F(-1.847759065f) ;
    int t3 = p1 + p2 * F(0.765366865f);
    p2 = s0;
    p3 = s4;
    int t0 = (p2 + p3) * 4096, t1 = (p2 - p3) * 4096;
    o.x0 = t0 + t3; o.x3 = t0 - t3; o.x1 = t1 + t2; o.x2 = t1 - t2;
    t0 = s7; t1 = s5; t2 = s3; t3 = s1;
    p3 = t0 + t2;
    int p4 = t1 + t3;
    p1 = t0 + t3;
    p2 = t1 + t2;
    const int p5 = (p3 + p4) * F(1.175875602f);
    t0 = t0 * F(0.298631336f); t1 = t1 * F(2.053119869f); t2 = t2 * F(3.072711026f); t3 = t3 * F(1.501321110f);
    p1 = p5 + p1 * F(-0.899976223f); p2 = p5 + p2 * F(-2.562915447f);
    p3 = p3 * F(-1.961570560f); p4 = p4 * F(-0.390180644f);
    o.t3 = t3 + p1 + p4; o.t2 = t2 + p2 + p3; o.t1 = t1 + p2 + p4; o.t0 = t0 + p1 + p3;
    return o;
  }
  static void IdctBlock(uint8_t* out, int stride, const short d[64]) {
    int val[64];
    for (int i = 0; i < 8; ++i) {
      const short* c = d + i;
      int* v = val + i;
      if (c[8] == 0 && c[16] == 0 && c[24] == 0 && c[32] == 0 && c[40] == 0 && c[48] == 0 && c[56] == 0) {
        const int dc = c[0] * 4;
        for (int r = 0; r < 8; ++r) v[8 * r] = dc;
      } else {
        Pass a = Idct1D(c[0], c[8], c[16], c[24], c[32], c[40], c[48], c[56]);
        a.x0 += 512; a.x1 += 512; a.x2 += 512; a.x3 += 512;
        v[0] = (a.x0 + a.t3) >> 10; v[56] = (a.x0 - a.t3) >> 10;
        v[8] = (a.x1 + a.t2) >> 10; v[48] = (a.x1 - a.t2) >> 10;
        v[16] = (a.x2 + a.t1) >> 10; v[40] = (a.x2 - a.t1) >> 10;
        v[24] = (a.x3 + a.t0) >> 10; v[32] = (a.x3 - a.t0) >> 10;
      }
    }
    for (int i = 0; i < 8; ++i) {
      const int* v = val + 8 * i;
      uint8_t* o = out + static_cast<size_t>(i) * stride;
      Pass a = Idct1D(v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7]);
      const int bias = 65536 + (128 << 17);
      a.x0 += bias; a.x1 += bias; a.x2 += bias; a.x3 += bias;
      o[0] = Clamp((a.x0 + a.t3) >> 17); o[7] = Clamp((a.x0 - a.t3) >> 17);
      o[1] = Clamp((a.x1 + a.t2) >> 17); o[6] = Clamp((a.x1 - a.t2) >> 17);
      o[2] = Clamp((a.x2 + a.t1) >> 17); o[5] = Clamp((a.x2 - a.t1) >> 17);
      o[3] = Clamp((a.x3 + a.t0) >> 17); o[4] = Clamp((a.x3 - a.t0) >> 17);
    }
  }

  bool ReadHeaderSegment(int m) {
    switch (m) {
      case 0xDD: {
        if (Get16() != 4) return Fail("bad DRI");
        restart_interval = Get16();
        return true;
      }
      case 0xDB: {
        int len = Get16() - 2;
        while (len > 0) {
          const int q = Get8(), prec = q >> 4, t = q & 15;
          if (prec > 1 || t > 3) return Fail("bad DQT");
          for (int i = 0; i < 64; ++i) dequant[t][ZigZag()[i]] = static_cast<uint16_t>(prec ? Get16() : Get8());
          len -= prec ? 129 : 65;
        }
        return len == 0 || Fail("bad DQT length");
      }
      case 0xC4: {
        int len = Get16() - 2;
        while (len > 0) {
          const int q = Get8(), tc = q >> 4, th = q & 15;
          if (tc > 1 || th > 3) return Fail("bad DHT");
          uint8_t counts[16];
          int n = 0;
          for (int i = 0; i < 16; ++i) n += counts[i] = static_cast<uint8_t>(Get8());
          if (n > 256) return Fail("bad DHT");
          Huffman& h = tc ? huff_ac[th] : huff_dc[th];
          if (!h.Build(counts)) return Fail("bad huffman table");
          for (int i = 0; i < n; ++i) h.values[i] = static_cast<uint8_t>(Get8());
          (tc ? have_ac : have_dc)[th] = true;
          len -= 17 + n;
        }
        return len == 0 || Fail("bad DHT length");
      }
      default:
        if ((m >= 0xE0 && m <= 0xEF) || m == 0xFE) {  // APPn, COM
          const int len = Get16();
          if (len < 2 || p + (len - 2) > end) return Fail("bad segment length");
          p += len - 2;
          return true;
        }
        return Fail("unknown marker");
    }
  }

  int NextMarker() {
    if (hit_marker) {
      hit_marker = false;
      const int m = marker;
      marker = 0xff;
      return m;
    }
    int x = Get8();
    while (x != 0xff && p < end) x = Get8();  // tolerate garbage between segments
    while (x == 0xff && p < end) x = Get8();
    return x;
  }

  bool ReadFrame() {
    const int len = Get16();
    if (Get8() != 8) return Fail("only 8-bit JPEG is supported");
    height = Get16();
    width = Get16();
    n_comp = Get8();
    if (width <= 0 || height <= 0) return Fail("empty image");
    if (n_comp != 1 && n_comp != 3) return Fail("only greyscale and YCbCr JPEG are supported");
    if (len != 8 + 3 * n_comp) return Fail("bad SOF length");
    for (int i = 0; i < n_comp; ++i) {
      Component& c = comp[i];
      c.id = Get8();
      const int q = Get8();
      c.h = q >> 4;
      c.v = q & 15;
      c.tq = Get8();
      if (c.h < 1 || c.h > 2 || c.v < 1 || c.v > 2) return Fail("sampling factors above 2 are not supported");
      if (c.tq > 3) return Fail("bad quantisation table index");
      h_max = c.h > h_max ? c.h : h_max;
      v_max = c.v > v_max ? c.v : v_max;
    }
    for (int i = 0; i < n_comp; ++i)
      if (h_max % comp[i].h || v_max % comp[i].v) return Fail("fractional sampling ratios are not supported");
    const int mw = h_max * 8, mh = v_max * 8;
    mcu_x = (width + mw - 1) / mw;
    mcu_y = (height + mh - 1) / mh;
    for (int i = 0; i < n_comp; ++i) {
      Component& c = comp[i];
      c.x = (width * c.h + h_max - 1) / h_max;
      c.y = (height * c.v + v_max - 1) / v_max;
      c.w2 = mcu_x * c.h * 8;
      c.h2 = mcu_y * c.v * 8;
      c.data.assign(static_cast<size_t>(c.w2) * c.h2, 0);
    }
    return true;
  }

  bool ReadScan() {
    Get16();
    const int n = Get8();
    if (n < 1 || n > n_comp) return Fail("bad SOS component count");
    int order[3];
    for (int i = 0; i < n; ++i) {
      const int id = Get8(), q = Get8();
      int which = 0;
      while (which < n_comp && comp[which].id != id) ++which;
      if (which == n_comp) return Fail("bad SOS component id");
      comp[which].hd = q >> 4;
      comp[which].ha = q & 15;
      if (comp[which].hd > 3 || comp[which].ha > 3 || !have_dc[comp[which].hd] || !have_ac[comp[which].ha]) return Fail("missing huffman table");
      order[i] = which;
    }
    const int ss = Get8();
    Get8();
    const int a = Get8();
    if (ss != 0 || a != 0) return Fail("progressive JPEG is not supported");
    ResetEntropy();
    const int interval = restart_interval ? restart_interval : 0x7fffffff;
    int todo = interval;
    short block[64];
    auto restart = [&]() -> bool {
      if (code_bits < 24) Grow();
      if (!(marker >= 0xD0 && marker <= 0xD7)) return false;  // no restart marker: the data ends here
      ResetEntropy();
      return true;
    };
    if (n == 1) {  // one component: its blocks in raster order, only those that cover the image
      Component& c = comp[order[0]];
      const int bw = (c.x + 7) >> 3, bh = (c.y + 7) >> 3;
      for (int j = 0; j < bh; ++j)
        for (int i = 0; i < bw; ++i) {
          if (!DecodeBlock(block, c)) return false;
          IdctBlock(c.data.data() + static_cast<size_t>(c.w2) * j * 8 + i * 8, c.w2, block);
          if (--todo <= 0) {
            if (!restart()) return true;
            todo = interval;
          }
        }
      return true;
    }
    for (int j = 0; j < mcu_y; ++j)
      for (int i = 0; i < mcu_x; ++i) {
        for (int k = 0; k < n; ++k) {
          Component& c = comp[order[k]];
          for (int y = 0; y < c.v; ++y)
            for (int x = 0; x < c.h; ++x) {
              if (!DecodeBlock(block, c)) return false;
              IdctBlock(c.data.data() + static_cast<size_t>(c.w2) * ((j * c.v + y) * 8) + (i * c.h + x) * 8, c.w2, block);
            }
        }
        if (--todo <= 0) {
          if (!restart()) return true;
          todo = interval;
        }
      }
    return true;
  }

  // one output row of a component at full width (the 3:1 triangle filters; `near`/`far` are the two source rows)
  static void Upsample(uint8_t* out, const uint8_t* in_near, const uint8_t* in_far, int w, int hs, int vs) {
    if (hs == 1 && vs == 1) {
      std::memcpy(out, in_near, static_cast<size_t>(w));
    } else if (hs == 1) {  // vertical only
      for (int i = 0; i < w; ++i) out[i] = static_cast<uint8_t>((3 * in_near[i] + in_far[i] + 2) >> 2);
    } else if (vs == 1) {  // horizontal only
      const uint8_t* in = in_near;
      if (w == 1) {
        out[0] = out[1] = in[0];
        return;
      }
      out[0] = in[0];
      out[1] = static_cast<uint8_t>((in[0] * 3 + in[1] + 2) >> 2);
      int i;
      for (i = 1; i < w - 1; ++i) {
        const int n = 3 * in[i] + 2;
        out[i * 2] = static_cast<uint8_t>((n + in[i - 1]) >> 2);
        out[i * 2 + 1] = static_cast<uint8_t>((n + in[i + 1]) >> 2);
      }
      out[i * 2] = static_cast<uint8_t>((in[w - 2] * 3 + in[w - 1] + 2) >> 2);
      out[i * 2 + 1] = in[w - 1];
    } else {  // both
      if (w == 1) {
        out[0] = out[1] = static_cast<uint8_t>((3 * in_near[0] + in_far[0] + 2) >> 2);
        return;
      }
      int t1 = 3 * in_near[0] + in_far[0];
      out[0] = static_cast<uint8_t>((t1 + 2) >> 2);
      for (int i = 1; i < w; ++i) {
        const int t0 = t1;
        t1 = 3 * in_near[i] + in_far[i];
        out[i * 2 - 1] = static_cast<uint8_t>((3 * t0 + t1 + 8) >> 4);
        out[i * 2] = static_cast<uint8_t>((3 * t1 + t0 + 8) >> 4);
      }
      out[w * 2 - 1] = static_cast<uint8_t>((t1 + 2) >> 2);
    }
  }

  void ToRgb(std::vector<uint8_t>& rgb) {
    rgb.assign(static_cast<size_t>(width) * height * 3, 0);
    std::vector<uint8_t> line[3];
    struct Cursor { int hs, vs, ystep, ypos, w_lores; const uint8_t *line0, *line1; } cur[3];
    for (int k = 0; k < n_comp; ++k) {
      Component& c = comp[k];
      line[k].assign(static_cast<size_t>(width) + 3, 0);
      cur[k] = Cursor{h_max / c.h, v_max / c.v, (v_max / c.v) >> 1, 0, (width + h_max / c.h - 1) / (h_max / c.h), c.data.data(), c.data.data()};
    }
    auto fixed = [](float x) { return static_cast<int>(x * 4096.0f + 0.5f) << 8; };
    for (int j = 0; j < height; ++j) {
      const uint8_t* row[3] = {nullptr, nullptr, nullptr};
      for (int k = 0; k < n_comp; ++k) {
        Component& c = comp[k];
        Cursor& r = cur[k];
        const bool y_bot = r.ystep >= (r.vs >> 1);
        Upsample(line[k].data(), y_bot ? r.line1 : r.line0, y_bot ? r.line0 : r.line1, r.w_lores, r.hs, r.vs);
        row[k] = (r.hs == 1 && r.vs == 1) ? (y_bot ? r.line1 : r.line0) : line[k].data();
        if (++r.ystep >= r.vs) {
          r.ystep = 0;
          r.line0 = r.line1;
          if (++r.ypos < c.y) r.line1 += c.w2;
        }
      }
      uint8_t* out = rgb.data() + static_cast<size_t>(j) * width * 3;
      if (n_comp == 1) {
        for (int i = 0; i < width; ++i) out[3 * i] = out[3 * i + 1] = out[3 * i + 2] = row[0][i];
        continue;
      }
      for (int i = 0; i < width; ++i) {
        const int y_fixed = (row[0][i] << 20) + (1 << 19);
        const int cb = row[1][i] - 128, cr = row[2][i] - 128;
        int r = y_fixed + cr * fixed(1.40200f);
        int g = y_fixed + cr * -fixed(0.71414f) + ((cb * -fixed(0.34414f)) & static_cast<int>(0xffff0000u));
        int b = y_fixed + cb * fixed(1.77200f);
        out[3 * i] = Clamp(r >> 20);
        out[3 * i + 1] = Clamp(g >> 20);
        out[3 * i + 2] = Clamp(b >> 20);
      }
    }
  }
};

}  // namespace detail

// 8-bit RGB, row-major, top row first.  false + `why` for files outside the scope above.
inline bool Decode(const uint8_t* data, size_t n, int& w, int& h, std::vector<uint8_t>& rgb, std::string& why) {
  auto d = std::make_unique<detail::Decoder>();
  d->p = data;
  d->end = data + n;
  auto fail = [&](const char* m) {
    why = d->why.empty() ? m : d->why;
    return false;
  };
  if (n < 4 || data[0] != 0xff || data[1] != 0xd8) return fail("not a JPEG file");
  d->p += 2;
  bool have_frame = false, have_scan = false;
  for (;;) {
    const int m = d->NextMarker();
    if (d->p >= d->end && m != 0xD9) break;
    if (m == 0xD9) break;  // EOI
    if (m == 0xC0 || m == 0xC1) {
      if (have_frame) return fail("two frames");
      if (!d->ReadFrame()) return fail("bad frame header");
      have_frame = true;
    } else if (m == 0xC2) {
      return fail("progressive JPEG is not supported (re-save as baseline, or bake a .rgbtab)");
    } else if (m == 0xC9 || m == 0xCA || m == 0xCB) {
      return fail("arithmetic-coded JPEG is not supported");
    } else if (m == 0xDA) {
      if (!have_frame) return fail("scan before frame");
      if (!d->ReadScan()) return fail("bad scan");
      have_scan = true;
    } else if (m >= 0xD0 && m <= 0xD7) {
      continue;  // stray restart marker
    } else if (!d->ReadHeaderSegment(m)) {
      return fail("bad segment");
    }
  }
  if (!have_frame || !have_scan) return fail("no image data");
  w = d->width;
  h = d->height;
  d->ToRgb(rgb);
  return true;
}

}  // namespace rt::scene::jpeg
