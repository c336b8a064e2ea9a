// rt::scene::Image implementation (reference: src/scene/image.cc:16-73).
#include "image.h"

#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>

#include "jpeg_baseline.h"

// Optional: other formats (PNG, progressive JPEG ...) through stb_image when the embedder points RT_STB_INCLUDE at a copy
// (make RT_STB_INCLUDE=dir).  The product does not need it: baseline JPEG - what the reference ships - is decoded by
// jpeg_baseline.h, and nothing is taken from the reference tree at build time.
#ifdef RT_HAVE_STB
#define STB_IMAGE_IMPLEMENTATION
#define STBI_NO_HDR_WRITE
#include <stb_image.h>
#endif

namespace rt::scene {

namespace {

// value*255.999 truncation with saturation, image.cc:69-73
unsigned char QuantiseLinear(float v) {
  if (v <= 0.0f) return 0;
  if (v >= 1.0f) return 255;
  return static_cast<unsigned char>(v * 255.999f);
}

// ".rgbtab": "RGBTAB1\n<w> <h>\n" + w*h*3 bytes — the byte table already in its final form.
bool LoadTable(const std::string& path, int& w, int& h, std::vector<unsigned char>& out) {
  std::ifstream f(path, std::ios::binary);
  if (!f) return false;
  std::string magic;
  f >> magic >> w >> h;
  if (magic != "RGBTAB1" || w <= 0 || h <= 0) return false;
  f.get();  // the single newline after the header
  out.resize(static_cast<size_t>(w) * h * 3);
  f.read(reinterpret_cast<char*>(out.data()), static_cast<std::streamsize>(out.size()));
  return static_cast<size_t>(f.gcount()) == out.size();
}

// The float image stbi_loadf returns for an 8-bit file (the reference's loader, image.cc:31): pow(byte / 255, 2.2) evaluated in
// double on the float quotient, times the scale 1, rounded to float - then the reference's own x 255.999 re-quantisation.
unsigned char LinearisedByte(unsigned char v) {
  const float lin = static_cast<float>(std::pow(static_cast<double>(v / 255.0f), static_cast<double>(2.2f)) * static_cast<double>(1.0f));
  return QuantiseLinear(lin);
}

bool LoadJpeg(const std::string& path, int& w, int& h, std::vector<unsigned char>& out, bool& is_jpeg) {
  is_jpeg = false;
  std::ifstream f(path, std::ios::binary);
  if (!f) return false;
  std::vector<unsigned char> bytes((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
  if (bytes.size() < 4 || bytes[0] != 0xff || bytes[1] != 0xd8) return false;
  is_jpeg = true;
  std::vector<uint8_t> rgb;
  std::string why;
  if (!jpeg::Decode(bytes.data(), bytes.size(), w, h, rgb, why)) {
    std::cerr << "ERROR: '" << path << "': " << why << "\n";
    return false;
  }
  unsigned char table[256];
  for (int v = 0; v < 256; ++v) table[v] = LinearisedByte(static_cast<unsigned char>(v));
  out.resize(rgb.size());
  for (size_t i = 0; i < rgb.size(); ++i) out[i] = table[rgb[i]];
  return true;
}

bool LoadOne(const std::string& path, int& w, int& h, std::vector<unsigned char>& out) {
  if (LoadTable(path + ".rgbtab", w, h, out)) return true;
  if (path.size() > 7 && path.compare(path.size() - 7, 7, ".rgbtab") == 0 && LoadTable(path, w, h, out)) return true;
  bool is_jpeg = false;
  if (LoadJpeg(path, w, h, out, is_jpeg)) return true;
#ifdef RT_HAVE_STB
  int n = 3;
  float* lin = stbi_loadf(path.c_str(), &w, &h, &n, 3);  // gamma-2.2 linearised floats
  if (lin) {
    out.resize(static_cast<size_t>(w) * h * 3);
    for (size_t i = 0; i < out.size(); ++i) out[i] = QuantiseLinear(lin[i]);
    stbi_image_free(lin);
    return true;
  }
#endif
  return false;
}

}  // namespace

Image::Image(const std::string& filename) {
  if (!Load(filename)) std::cerr << "ERROR: Could not load image '" << filename << "'\n";
}

bool Image::Load(const std::string& filename) {
  std::vector<std::string> candidates{filename};
#ifdef IMAGE_DIR
  candidates.push_back(std::string(IMAGE_DIR) + "/" + filename);
#endif
  if (const char* env = std::getenv("RT_IMAGE_DIR")) candidates.push_back(std::string(env) + "/" + filename);
  for (const auto& c : candidates) {
    if (LoadOne(c, width_, height_, bdata_)) {
      path_ = c;
      return true;
    }
  }
  width_ = height_ = 0;
  bdata_.clear();
  return false;
}

const unsigned char* Image::PixelData(int x, int y) const {
  static unsigned char magenta[3] = {255, 0, 255};
  if (bdata_.empty()) return magenta;
  auto clampi = [](int v, int hi) { return v < 0 ? 0 : (v < hi ? v : hi - 1); };
  return &bdata_[(static_cast<size_t>(clampi(y, height_)) * width_ + clampi(x, width_)) * 3];
}

}  // namespace rt::scene
