// rt::integrator::Sampler family (src/integrator/sampler.h:14-82).
//
// In the reference a Sampler owns the per-pixel spp loop on the CPU (MegaKernel calls SamplePixel once per pixel,
// mega_kernel.h:28-36).  Here that loop is the device's sample scheduler, so a sampler has two faces:
//   * Policy()       what the loop does, as parameters the device runs for every pixel at once (fixed spp, or the
//                    adaptive stop with its minimum / maximum / threshold).  MegaKernel / B200Renderer read this;
//   * SamplePixel()  the reference's own virtual (same signature, sampler.h:14-20), kept so that code written against
//                    the reference compiles and runs: DefaultSampler / AdaptiveSampler implement it as a ONE-PIXEL device
//                    render of pixel (i, j) (a launch per call: a compatibility path, not the fast one).
// A user sampler that only exists as CPU code (it overrides SamplePixel and not Policy) cannot run here - there is no
// CPU path - and is rejected by the renderer with a clear error instead of a compile break.
#pragma once
#include <optional>

#include "../core/color.h"
#include "../scene/camera.h"
#include "../scene/scene.h"

namespace rt::integrator {

struct SamplerPolicy {
  int min_samples = 0;     // adaptive: samples before the stop rule is looked at
  int max_samples = 0;     // fixed: exactly this many
  float threshold = 0.0f;  // adaptive: relative-error bound
  bool adaptive = false;
};

class Sampler {
 public:
  virtual ~Sampler() = default;
  // return number of samples taken (sampler.h:16-19)
  virtual int SamplePixel(core::Color& pixel, const scene::Scene& world, const scene::Camera& cam, int i, int j) const = 0;
  // the device form of this sampler's loop; nullopt = "CPU code only"
  virtual std::optional<SamplerPolicy> Policy() const { return std::nullopt; }
  // convenience views of Policy() (round-1 callers)
  int max_samples() const { return Require().max_samples; }
  int min_samples() const { return Require().adaptive ? Require().min_samples : Require().max_samples; }
  float threshold() const { return Require().threshold; }
  bool adaptive() const { return Require().adaptive; }

 protected:
  SamplerPolicy Require() const;
};

// One pixel of `cam` rendered on the device with `policy`: mean radiance into `pixel`, returns the samples taken.
// Defined in device_pixel.h (included below).
int SamplePixelOnDevice(const SamplerPolicy& policy, core::Color& pixel, const scene::Scene& world, const scene::Camera& cam, int i, int j);

// sampler.h:22-37: exactly num_samples camera rays per pixel, mean of GetPixel
class DefaultSampler : public Sampler {
 public:
  explicit DefaultSampler(int num_samples) : num_samples_(num_samples) {}
  std::optional<SamplerPolicy> Policy() const override { return SamplerPolicy{num_samples_, num_samples_, 0.0f, false}; }
  int SamplePixel(core::Color& pixel, const scene::Scene& world, const scene::Camera& cam, int i, int j) const override {
    return SamplePixelOnDevice(*Policy(), pixel, world, cam, i, j);
  }

 private:
  int num_samples_;
};

// sampler.h:43-82.  The reference's version accumulates the running total into `sum`
// (sampler.h:56-59) and is not used by its main(); here the test it describes - luminance of the
// variance over luminance of the mean - runs per pixel on the device (RTB_ADAPT_LUMINANCE).
class AdaptiveSampler : public Sampler {
 public:
  AdaptiveSampler(int min_samples, int max_samples, float threshold) : min_(min_samples), max_(max_samples), thr_(threshold) {}
  std::optional<SamplerPolicy> Policy() const override { return SamplerPolicy{min_, max_, thr_, true}; }
  int SamplePixel(core::Color& pixel, const scene::Scene& world, const scene::Camera& cam, int i, int j) const override {
    return SamplePixelOnDevice(*Policy(), pixel, world, cam, i, j);
  }

 private:
  int min_, max_;
  float thr_;
};

}  // namespace rt::integrator

#include "device_pixel.h"
