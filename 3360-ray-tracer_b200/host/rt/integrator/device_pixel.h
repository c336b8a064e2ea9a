// Sampler::SamplePixel on the device: the reference's per-pixel entry (src/integrator/sampler.h:26-34, 49-76) as a
// 1 x 1 render.  The camera handed to the device is `cam`'s own frame with pixel00 moved to pixel (i, j), so the jittered
// thin-lens rays are those of Camera::GetRay(i, j) (camera.h:134-144) and every sample runs Camera::GetPixel's bounce loop.
// The device scene of the last `world` is cached per host thread (a reference-style caller invokes SamplePixel for every
// pixel of an image); each call still costs a kernel launch - B200Renderer is the fast path.
#pragma once
#include <rtb.h>

#include <memory>
#include <stdexcept>
#include <string>

#include "../b200/flatten.h"
#include "sampler.h"

namespace rt::integrator {

namespace detail {
struct PixelScene {
  const scene::Scene* world = nullptr;
  size_t n_objects = 0;
  RtbScene* scene = nullptr;
  uint32_t calls = 0;  // sample-index offset: successive calls on the same pixel draw new samples, as an RNG-driven CPU loop does
  ~PixelScene() {
    if (scene) rtb_scene_destroy(scene);
  }
};
inline PixelScene& CachedPixelScene(const scene::Scene& world) {
  static thread_local PixelScene c;
  if (c.world != &world || c.n_objects != world.objects_.size() || !c.scene) {
    if (c.scene) rtb_scene_destroy(c.scene);
    c = PixelScene{};
    if (rtb_init(0) != RTB_OK) throw std::runtime_error(std::string("Sampler::SamplePixel: ") + rtb_last_error());
    b200::FlatScene flat = b200::Flattener().Run(world);
    const RtbSceneDesc d = flat.desc();
    if (rtb_scene_create(&d, &c.scene) != RTB_OK) throw std::runtime_error(std::string("Sampler::SamplePixel: ") + rtb_last_error());
    c.world = &world;
    c.n_objects = world.objects_.size();
  }
  return c;
}
}  // namespace detail

inline SamplerPolicy Sampler::Require() const {
  const std::optional<SamplerPolicy> p = Policy();
  if (!p)
    throw std::logic_error(
        "this Sampler only overrides SamplePixel (CPU code): the B200 build has no CPU path - give it a Policy() "
        "(fixed spp or adaptive min / max / threshold) that the device's sample loop can run");
  return *p;
}

inline int SamplePixelOnDevice(const SamplerPolicy& policy, core::Color& pixel, const scene::Scene& world, const scene::Camera& cam, int i, int j) {
  if (policy.max_samples <= 0) throw std::invalid_argument("Sampler::SamplePixel: no samples");
  detail::PixelScene& ps = detail::CachedPixelScene(world);
  RtbCamera dc = cam.ToDevice();  // the caller has run Initialize(), as MegaKernel::Render does before its pixel loop
  if (i < 0 || j < 0 || i >= dc.width || j >= dc.height) throw std::out_of_range("Sampler::SamplePixel: pixel outside the image");
  for (int c = 0; c < 3; ++c) dc.pixel00[c] += static_cast<float>(i) * dc.delta_u[c] + static_cast<float>(j) * dc.delta_v[c];
  dc.width = dc.height = 1;
  RtbRenderParams prm{};
  prm.spp_begin = static_cast<int32_t>((ps.calls++ % 4096u) * 65536u % 0x40000000u);
  prm.spp_end = prm.spp_begin + policy.max_samples;
  prm.max_depth = cam.max_depth_;
  prm.mode = RTB_MODE_A;  // Camera::GetPixel
  // the Philox counter is (pixel, sample, bounce) and the pixel index of a 1 x 1 image is 0: key the stream on (i, j) instead
  prm.seed = 0x3360ull ^ (static_cast<uint64_t>(static_cast<uint32_t>(i)) << 32) ^ (static_cast<uint64_t>(static_cast<uint32_t>(j)) << 12);
  float sum[3] = {0, 0, 0};
  RtbStats st{};
  int n = policy.max_samples;
  if (policy.adaptive) {
    RtbAdaptive ad{};
    ad.rel_threshold = policy.threshold;
    ad.min_samples = policy.min_samples > 0 ? policy.min_samples : 1;
    ad.criterion = RTB_ADAPT_LUMINANCE;
    float sumsq[3] = {0, 0, 0};
    uint32_t count = 0;
    if (rtb_render_adaptive(ps.scene, &dc, &prm, &ad, sum, sumsq, &count, &st) != RTB_OK)
      throw std::runtime_error(std::string("Sampler::SamplePixel: ") + rtb_last_error());
    n = static_cast<int>(count);
  } else if (rtb_render(ps.scene, &dc, &prm, sum, &st) != RTB_OK) {
    throw std::runtime_error(std::string("Sampler::SamplePixel: ") + rtb_last_error());
  }
  const double inv = n > 0 ? 1.0 / n : 0.0;
  pixel = core::Color(sum[0] * inv, sum[1] * inv, sum[2] * inv);
  return n;
}

}  // namespace rt::integrator
