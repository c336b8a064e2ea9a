// rt::integrator::Sampler family (src/integrator/sampler.h:14-82).  In the reference a Sampler
// owns the per-pixel spp loop on the CPU; here the loop is the device's wavefront scheduler, so a
// Sampler is a sampling *policy* the renderer reads.
#pragma once
namespace rt::integrator {
class Sampler {
 public:
  virtual ~Sampler() = default;
  virtual int max_samples() const = 0;
  virtual int min_samples() const { return max_samples(); }
  virtual float threshold() const { return 0.0f; }
  virtual bool adaptive() const { return false; }
};
// sampler.h:22-37: exactly num_samples camera rays per pixel, mean of GetPixel
class DefaultSampler : public Sampler {
 public:
  explicit DefaultSampler(int num_samples) : num_samples_(num_samples) {}
  int max_samples() const override { return num_samples_; }
 private:
  int num_samples_;
};
// sampler.h:43-82.  The reference's version accumulates the running total into `sum`
// (sampler.h:56-59) and is not used by its main(); here the test it describes - luminance of the
// variance over luminance of the mean - runs per pixel on the device (RTB_ADAPT_LUMINANCE).
class AdaptiveSampler : public Sampler {
 public:
  AdaptiveSampler(int min_samples, int max_samples, float threshold) : min_(min_samples), max_(max_samples), thr_(threshold) {}
  int max_samples() const override { return max_; }
  int min_samples() const override { return min_; }
  float threshold() const override { return thr_; }
  bool adaptive() const override { return true; }
 private:
  int min_, max_;
  float thr_;
};
}  // namespace rt::integrator
