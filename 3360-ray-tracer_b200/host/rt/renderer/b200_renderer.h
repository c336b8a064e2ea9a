// rt::renderer — the drop-in boundary for the render loop (seam 1).
//
// B200Renderer renders scene+camera on one B200 through the C ABI and writes the reference's
// output: ASCII PPM "P3" on stdout (src/renderer/mega_kernel.h:48-51, src/core/color.h:18-33),
// progress on stderr.  MegaKernel and WavefrontRenderer keep the reference's constructor
// signatures (mega_kernel.h:12-13, wavefront.h:24-33) and select estimator A / B.
#pragma once
#include <rtb.h>

#include <cstdint>
#include <fstream>
#include <iostream>
#include <stdexcept>
#include <string>
#include <vector>

#include "../b200/flatten.h"
#include "../core/png.h"
#include "../core/timer.h"
#include "../integrator/ray_integrator.h"
#include "../integrator/sampler.h"
#include "../scene/camera.h"
#include "../scene/scene.h"

namespace rt::renderer {

struct B200Options {
  int mode = RTB_MODE_A;
  int samples_per_pixel = 0;  // 0: camera.samples_per_pixel_
  int max_depth = -1;         // <0: camera.max_depth_
  uint64_t seed = 0x3360;
  int device = 0;
  int gpus = 1;               // > 1: rtb_render_multi over devices 0 .. gpus-1 (spp split, peer reduce); fixed-spp renders
  uint32_t pool_size = 0;
  // adaptive per-pixel stop (rtb_render_adaptive): samples_per_pixel is then the per-pixel maximum
  bool adaptive = false;
  float rel_threshold = 0.05f;  // wavefront.cc:42
  int min_samples = 16;         // wavefront.cc:43
  int criterion = RTB_ADAPT_PIXEL_STATE;
  bool emit_ppm = true;       // write the image to `out`
  bool binary_ppm = false;    // "P6" (3 bytes per pixel) instead of the reference's ASCII "P3": same bytes, 1/4 of the text
  bool png = false;           // PNG (8-bit RGB, core/png.h) instead of a PPM; wins over binary_ppm
  // progressive accumulator: `resume_from` continues a dumped render with its next global sample
  // indices, `dump_to` writes the accumulator after this call.  File: "RTBACC1\n" w h spp | float32 sums.
  std::string resume_from, dump_to;
  std::ostream* out = &std::cout;
  std::ostream* log = &std::clog;
};

class B200Renderer {
 public:
  B200Renderer(const scene::Scene& world, const scene::Camera& cam, B200Options opt = {}) : world_(world), cam_(cam), opt_(opt) {}

  void Render() {
    scene::Camera& cam = const_cast<scene::Camera&>(cam_);
    cam.Initialize();  // MegaKernel::Render does the same first (mega_kernel.h:16)
    const int w = cam.get_image_width(), h = cam.get_image_height();
    const int spp = opt_.samples_per_pixel > 0 ? opt_.samples_per_pixel : cam.samples_per_pixel_;
    const int depth = opt_.max_depth >= 0 ? opt_.max_depth : cam.max_depth_;
    core::Timer clock;
    if (rtb_init(opt_.device) != RTB_OK) throw std::runtime_error(std::string("B200Renderer: ") + rtb_last_error());
    b200::FlatScene flat = b200::Flattener().Run(world_);
    RtbSceneDesc desc = flat.desc();
    struct TreeGuard {  // the device-native tree: host work, once per scene, shared by every device that renders it
      RtbCompiledTree* t = nullptr;
      ~TreeGuard() { rtb_tree_destroy(t); }
    } tree;
    if (rtb_tree_compile(&desc, &tree.t) != RTB_OK) throw std::runtime_error(std::string("B200Renderer: ") + rtb_last_error());
    desc.compiled = tree.t;
    RtbScene* scene = nullptr;
    if (rtb_scene_create(&desc, &scene) != RTB_OK) throw std::runtime_error(std::string("B200Renderer: ") + rtb_last_error());
    const double t_upload = clock.elapsed();
    const RtbCamera dc = cam.ToDevice();
    RtbRenderParams prm{};
    prm.spp_begin = 0;
    prm.spp_end = spp;
    prm.max_depth = depth;
    prm.mode = opt_.mode;
    prm.seed = opt_.seed;
    prm.pool_size = opt_.pool_size;
    sum_.assign(static_cast<size_t>(w) * h * 3, 0.0f);
    int spp_done = 0;
    if (!opt_.resume_from.empty()) {
      if (opt_.adaptive) throw std::runtime_error("B200Renderer: resume is for fixed-spp renders");
      spp_done = LoadAccumulator(opt_.resume_from, w, h, sum_);
      prm.spp_begin = spp_done;  // Philox is keyed on the global sample index: the continuation adds NEW samples
      prm.spp_end = spp_done + spp;
      prm.flags |= RTB_RENDER_KEEP_ACCUM;
    }
    const int spp_total = spp_done + spp;
    std::vector<uint8_t> rgb8(sum_.size());
    int rc2;
    if (opt_.adaptive) {
      // integrator::PixelState per pixel (pixel_state.h:13-19): sum, sum of squares, samples; image = sum / samples
      RtbAdaptive ad{};
      ad.rel_threshold = opt_.rel_threshold;
      ad.min_samples = opt_.min_samples;
      ad.criterion = opt_.criterion;
      sumsq_.assign(sum_.size(), 0.0f);
      count_.assign(static_cast<size_t>(w) * h, 0u);
      const int rc = rtb_render_adaptive(scene, &dc, &prm, &ad, sum_.data(), sumsq_.data(), count_.data(), &stats_);
      rc2 = rc == RTB_OK ? rtb_resolve_counts(sum_.data(), count_.data(), count_.size(), rgb8.data()) : rc;
    } else if (opt_.gpus > 1) {
      if (prm.flags & RTB_RENDER_KEEP_ACCUM) throw std::runtime_error("B200Renderer: resume runs on one GPU");
      const int rc = rtb_render_multi(&desc, opt_.gpus, nullptr, &dc, &prm, sum_.data(), &stats_);
      rc2 = rc == RTB_OK ? rtb_resolve(sum_.data(), static_cast<size_t>(w) * h, spp_total, rgb8.data()) : rc;
    } else {
      const int rc = rtb_render(scene, &dc, &prm, sum_.data(), &stats_);
      rc2 = rc == RTB_OK ? rtb_resolve(sum_.data(), static_cast<size_t>(w) * h, spp_total, rgb8.data()) : rc;
    }
    rtb_scene_destroy(scene);
    if (rc2 != RTB_OK) throw std::runtime_error(std::string("B200Renderer: ") + rtb_last_error());
    width_ = w;
    height_ = h;
    spp_ = spp_total;
    if (!opt_.dump_to.empty()) SaveAccumulator(opt_.dump_to, w, h, spp_total, sum_);
    if (opt_.log) {
      *opt_.log << "B200: " << w << "x" << h << " spp=" << spp << " depth=" << depth << " mode=" << (opt_.mode == RTB_MODE_A ? 'A' : 'B')
                << "  flatten+upload " << t_upload << "s  render " << stats_.ms_total * 1e-3 << "s  "
                << stats_.samples / (stats_.ms_total * 1e3) << " Msamples/s  " << stats_.rays / (stats_.ms_total * 1e3) << " Mrays/s\n";
    }
    if (opt_.emit_ppm && opt_.out && opt_.png) {
      const std::string file = core::EncodePng(rgb8.data(), w, h);
      opt_.out->write(file.data(), static_cast<std::streamsize>(file.size()));
    } else if (opt_.emit_ppm && opt_.out && opt_.binary_ppm) {
      const std::string head = "P6\n" + std::to_string(w) + ' ' + std::to_string(h) + "\n255\n";
      opt_.out->write(head.data(), static_cast<std::streamsize>(head.size()));
      opt_.out->write(reinterpret_cast<const char*>(rgb8.data()), static_cast<std::streamsize>(rgb8.size()));
    } else if (opt_.emit_ppm && opt_.out) {
      std::string text;
      text.reserve(rgb8.size() * 4 + 32);
      text += "P3\n" + std::to_string(w) + ' ' + std::to_string(h) + "\n255\n";
      for (size_t i = 0; i < rgb8.size(); i += 3) {
        text += std::to_string(rgb8[i]);
        text += ' ';
        text += std::to_string(rgb8[i + 1]);
        text += ' ';
        text += std::to_string(rgb8[i + 2]);
        text += '\n';
      }
      opt_.out->write(text.data(), static_cast<std::streamsize>(text.size()));
    }
    if (opt_.log) *opt_.log << "\rDone. Total samples: " << stats_.samples << "Per pixel: " << stats_.samples / (static_cast<uint64_t>(w) * h) << "\n";
  }

  static void SaveAccumulator(const std::string& path, int w, int h, int spp, const std::vector<float>& sum) {
    std::ofstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("B200Renderer: cannot write '" + path + "'");
    f << "RTBACC1\n" << w << ' ' << h << ' ' << spp << '\n';
    f.write(reinterpret_cast<const char*>(sum.data()), static_cast<std::streamsize>(sum.size() * sizeof(float)));
  }
  // returns the samples per pixel already in the file
  static int LoadAccumulator(const std::string& path, int w, int h, std::vector<float>& sum) {
    std::ifstream f(path, std::ios::binary);
    std::string magic;
    int fw = 0, fh = 0, spp = 0;
    if (!f || !(f >> magic >> fw >> fh >> spp) || magic != "RTBACC1") throw std::runtime_error("B200Renderer: '" + path + "' is not an accumulator dump");
    if (fw != w || fh != h || spp < 0) throw std::runtime_error("B200Renderer: accumulator dump is " + std::to_string(fw) + "x" + std::to_string(fh) + ", camera is " + std::to_string(w) + "x" + std::to_string(h));
    f.get();  // the newline after the header
    sum.assign(static_cast<size_t>(w) * h * 3, 0.0f);
    f.read(reinterpret_cast<char*>(sum.data()), static_cast<std::streamsize>(sum.size() * sizeof(float)));
    if (f.gcount() != static_cast<std::streamsize>(sum.size() * sizeof(float))) throw std::runtime_error("B200Renderer: accumulator dump is truncated");
    return spp;
  }

  const std::vector<float>& radiance_sum() const { return sum_; }
  const std::vector<float>& radiance_sumsq() const { return sumsq_; }   // adaptive renders only
  const std::vector<uint32_t>& sample_counts() const { return count_; }  // adaptive renders only
  const RtbStats& stats() const { return stats_; }
  int width() const { return width_; }
  int height() const { return height_; }
  int samples_per_pixel() const { return spp_; }

 private:
  const scene::Scene& world_;
  const scene::Camera& cam_;
  B200Options opt_;
  std::vector<float> sum_, sumsq_;
  std::vector<uint32_t> count_;
  RtbStats stats_{};
  int width_ = 0, height_ = 0, spp_ = 0;
};

// renderer::MegaKernel(scene, camera, sampler).Render()  — estimator A (mega_kernel.h:10-59)
class MegaKernel {
 public:
  MegaKernel(scene::Scene& scene, scene::Camera& camera, integrator::Sampler& sampler) : sampler_(sampler), world_(scene), cam_(camera) {}
  void Render() {
    B200Options o;
    o.mode = RTB_MODE_A;
    o.samples_per_pixel = sampler_.max_samples();
    if (sampler_.adaptive()) {  // AdaptiveSampler(min, max, threshold): its luminance test, per pixel, on the device
      o.adaptive = true;
      o.criterion = RTB_ADAPT_LUMINANCE;
      o.min_samples = sampler_.min_samples();
      o.rel_threshold = sampler_.threshold();
    }
    B200Renderer(world_, cam_, o).Render();
  }
 private:
  integrator::Sampler& sampler_;
  scene::Scene& world_;
  scene::Camera& cam_;
};

// renderer::WavefrontRenderer(world, cam, integrator, max_depth, max_samples, batch).Render()
// — estimator B with its adaptive per-pixel stop (wavefront.cc:40-242).  `integrator` and `batch_size` are accepted for source
// compatibility; intersection runs inside the device wavefront, whose batch is the path pool.
class WavefrontRenderer {
 public:
  WavefrontRenderer(const scene::Scene& world, const scene::Camera& cam, integrator::RayIntegrator& integrator, int max_depth = 10,
                    int max_samples = 128, int batch_size = 8192)
      : world_(world), cam_(cam), integrator_(integrator), max_depth_(max_depth), max_ssp_(max_samples), batch_size_(batch_size) {}
  void Render() {
    B200Options o;
    o.mode = RTB_MODE_B;
    o.samples_per_pixel = max_ssp_;
    o.max_depth = max_depth_;
    o.adaptive = true;  // kRelThresh 0.05, kMinSamples 16, PixelState::IsConverged (wavefront.cc:42-43)
    B200Renderer(world_, cam_, o).Render();
  }
 private:
  const scene::Scene& world_;
  const scene::Camera& cam_;
  integrator::RayIntegrator& integrator_;
  int max_depth_, max_ssp_, batch_size_;
};

}  // namespace rt::renderer
