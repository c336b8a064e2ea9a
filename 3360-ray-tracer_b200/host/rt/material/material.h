// rt::material::Material family (src/material/material.h:13-214, material.cc).
// Same constructors as the reference.  Scatter / Sample / Eval / Pdf / Emitted are evaluated
// on the device per bounce (csrc/shade.cuh, citing material.cc line by line); the host objects
// are parameter carriers that export themselves through Describe().
#pragma once
#include <memory>
#include "../core/color.h"
#include "texture.h"
namespace rt::material {

class MaterialSink {
 public:
  virtual ~MaterialSink() = default;
  virtual void OnLambertian(const std::shared_ptr<Texture>& albedo) = 0;
  virtual void OnMetal(const core::Color& albedo, double fuzz) = 0;
  virtual void OnDielectric(double ref_idx) = 0;
  virtual void OnDiffuseLight(const std::shared_ptr<Texture>& emit) = 0;
};

class Material {
 public:
  virtual ~Material() = default;
  virtual bool IsSpecular() const { return false; }
  virtual void Describe(MaterialSink& sink) const = 0;
};

class Lambertian : public Material {
 public:
  explicit Lambertian(const core::Color& albedo) : tex_(std::make_shared<SolidColor>(albedo)) {}
  explicit Lambertian(std::shared_ptr<Texture> tex) : tex_(std::move(tex)) {}
  void Describe(MaterialSink& sink) const override { sink.OnLambertian(tex_); }
 private:
  std::shared_ptr<Texture> tex_;
};

class Metal : public Material {
 public:
  Metal(const core::Color& albedo, double fuzz) : albedo_(albedo), fuzz_(fuzz < 1.0 ? fuzz : 1.0) {}
  bool IsSpecular() const override { return true; }
  void Describe(MaterialSink& sink) const override { sink.OnMetal(albedo_, fuzz_); }
 private:
  core::Color albedo_;
  double fuzz_;
};

class Dielectric : public Material {
 public:
  explicit Dielectric(double index) : ref_idx_(index) {}
  bool IsSpecular() const override { return true; }
  void Describe(MaterialSink& sink) const override { sink.OnDielectric(ref_idx_); }
 private:
  double ref_idx_;
};

class DiffuseLight : public Material {
 public:
  explicit DiffuseLight(std::shared_ptr<Texture> tex) : emit_(std::move(tex)) {}
  explicit DiffuseLight(const core::Color& c) : emit_(std::make_shared<SolidColor>(c)) {}
  bool IsSpecular() const override { return true; }
  void Describe(MaterialSink& sink) const override { sink.OnDiffuseLight(emit_); }
 private:
  std::shared_ptr<Texture> emit_;
};

}  // namespace rt::material
