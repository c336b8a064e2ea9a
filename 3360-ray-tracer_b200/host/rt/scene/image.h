// rt::scene::Image (src/scene/image.h:7-29, image.cc:16-73): loads through stb_image's float
// path (gamma-2.2 linearisation), then re-quantises to bytes with value*255.999 — the byte
// table is what ImageTexture samples.  Search order: the given path, then IMAGE_DIR/.
// Also accepts the pre-decoded ".rgbtab" container (see assets/README.md) so that the byte
// table can be pinned to the reference decoder's output.
#pragma once
#include <string>
#include <vector>
namespace rt::scene {
class Image {
 public:
  Image() = default;
  explicit Image(const std::string& filename);
  bool Load(const std::string& filename);
  int Width() const { return width_; }
  int Height() const { return height_; }
  const unsigned char* PixelData(int x, int y) const;
  const std::vector<unsigned char>& Bytes() const { return bdata_; }
  const std::string& Path() const { return path_; }  // where the data was found ("" if not loaded)
 private:
  int width_ = 0, height_ = 0;
  std::vector<unsigned char> bdata_;
  std::string path_;
};
}  // namespace rt::scene
