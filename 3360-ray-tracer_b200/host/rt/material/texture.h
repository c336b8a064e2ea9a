// rt::material::Texture family (src/material/texture.h:11-77).  Host objects only carry the
// parameters; Value(u,v,p) is evaluated on the device (csrc/shade.cuh) from the RtbTexture
// table the flattener fills through Describe().
#pragma once
#include <memory>
#include <string>
#include "../core/color.h"
#include "../scene/image.h"
namespace rt::material {

class Texture;
class TextureSink {
 public:
  virtual ~TextureSink() = default;
  virtual void OnSolid(const core::Color& albedo) = 0;
  virtual void OnChecker(double scale, double inv_scale, const std::shared_ptr<Texture>& even, const std::shared_ptr<Texture>& odd) = 0;
  virtual void OnImage(const scene::Image& image) = 0;
};

class Texture {
 public:
  virtual ~Texture() = default;
  virtual void Describe(TextureSink& sink) const = 0;
};

class SolidColor : public Texture {
 public:
  SolidColor(const core::Color& albedo) : albedo_(albedo) {}
  SolidColor(double r, double g, double b) : albedo_(r, g, b) {}
  void Describe(TextureSink& sink) const override { sink.OnSolid(albedo_); }
 private:
  core::Color albedo_;
};

class CheckerTexture : public Texture {
 public:
  CheckerTexture(double scale, std::shared_ptr<Texture> even, std::shared_ptr<Texture> odd)
      : scale_(scale), inv_scale_(1.0 / scale), even_(std::move(even)), odd_(std::move(odd)) {}
  CheckerTexture(double scale, const core::Color& c1, const core::Color& c2)
      : CheckerTexture(scale, std::make_shared<SolidColor>(c1), std::make_shared<SolidColor>(c2)) {}
  void Describe(TextureSink& sink) const override { sink.OnChecker(scale_, inv_scale_, even_, odd_); }
 private:
  double scale_, inv_scale_;
  std::shared_ptr<Texture> even_, odd_;
};

class ImageTexture : public Texture {
 public:
  ImageTexture(const char* filename) : image_(filename) {}
  void Describe(TextureSink& sink) const override { sink.OnImage(image_); }
 private:
  scene::Image image_;
};

}  // namespace rt::material
