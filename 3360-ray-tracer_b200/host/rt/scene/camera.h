// rt::scene::Camera / CameraConfig / preset loading (src/scene/camera.h:25-243).
// Initialize() performs the reference's double-precision derivation (camera.h:100-131) and
// ToDevice() narrows the result into the RtbCamera the generate kernel reads.  GetRay/GetPixel
// are the device's generate/shade kernels.
#pragma once
#include <cmath>
#include <string>
#include <unordered_map>
#include <rtb.h>
#include "../core/constants.h"
#include "../core/math_utils.h"
#include "../core/vec3.h"
#include "json_lite.h"

namespace rt::scene {

struct CameraConfig {
  double aspect_ratio = 16 / 9.0;
  int image_width = 400;
  int samples_per_pixel = 50;
  int max_depth = 10;
  double vfov = 90.0;
  core::Vec3 lookfrom = core::Point3(0, 0, 0);
  core::Vec3 lookat = core::Point3(0, 0, -1);
  core::Vec3 vup = core::Point3(0, 1, 0);
  double defocus_angle = 0.0;
  double focus_dist = 10.0;
};

// Same keys and defaults as camera.h:40-56; lookfrom/lookat/vup are mandatory.
inline CameraConfig parseCamera(const json_lite::Value& j) {
  CameraConfig c;
  c.aspect_ratio = j.number_or("aspectRatio", c.aspect_ratio);
  c.image_width = static_cast<int>(j.number_or("imageWidth", c.image_width));
  c.samples_per_pixel = static_cast<int>(j.number_or("samplesPerPixel", c.samples_per_pixel));
  c.max_depth = static_cast<int>(j.number_or("maxDepth", c.max_depth));
  c.vfov = j.number_or("vfov", c.vfov);
  auto vec = [&](const char* key) {
    const json_lite::Value& a = j.at(key);
    return core::Vec3(a.at(0).number(), a.at(1).number(), a.at(2).number());
  };
  c.lookfrom = vec("lookfrom");
  c.lookat = vec("lookat");
  c.vup = vec("vup");
  c.defocus_angle = j.number_or("defocusAngle", c.defocus_angle);
  c.focus_dist = j.number_or("focusDist", c.focus_dist);
  return c;
}

inline std::unordered_map<std::string, CameraConfig> loadCamerasFromText(const std::string& text) {
  std::unordered_map<std::string, CameraConfig> out;
  const json_lite::Value root = json_lite::Parse(text);
  for (const auto& [name, cam] : root.members()) out[name] = parseCamera(cam);
  return out;
}
inline std::unordered_map<std::string, CameraConfig> loadCameras(const std::string& filename) {
  return loadCamerasFromText(json_lite::ReadFile(filename));
}

class Camera {
 public:
  double aspect_ratio_ = 1.0;
  int image_width_ = 100;
  int max_depth_ = 10;
  int samples_per_pixel_ = 10;
  double vfov_ = 90.0;
  core::Vec3 lookfrom_ = core::Point3(0, 0, 0);
  core::Vec3 lookat = core::Point3(0, 0, -1);
  core::Vec3 vup_ = core::Point3(0, 1, 0);
  double defocus_angle_ = 0;
  double focus_dist_ = 10;

  virtual ~Camera() = default;

  void SetFromConfig(const CameraConfig& c) {
    aspect_ratio_ = c.aspect_ratio;
    image_width_ = c.image_width;
    max_depth_ = c.max_depth;
    samples_per_pixel_ = c.samples_per_pixel;
    vfov_ = c.vfov;
    lookfrom_ = c.lookfrom;
    lookat = c.lookat;
    vup_ = c.vup;
    defocus_angle_ = c.defocus_angle;
    focus_dist_ = c.focus_dist;
  }

  void Initialize() {
    image_height_ = int(image_width_ / aspect_ratio_);
    if (image_height_ < 1) image_height_ = 1;
    center_ = lookfrom_;
    const double half_h = std::tan(core::DegreesToRadians(vfov_) / 2);
    const double vp_h = 2 * half_h * focus_dist_;
    const double vp_w = vp_h * (double(image_width_) / image_height_);
    w_ = core::Normalize(lookfrom_ - lookat);
    u_ = core::Normalize(core::Cross(vup_, w_));
    v_ = core::Cross(w_, u_);
    const core::Vec3 across = vp_w * u_;
    const core::Vec3 down = vp_h * -v_;
    pixel_delta_u_ = across / image_width_;
    pixel_delta_v_ = down / image_height_;
    const core::Vec3 upper_left = center_ - (focus_dist_ * w_) - across / 2 - down / 2;
    pixel00_loc_ = upper_left + 0.5 * (pixel_delta_u_ + pixel_delta_v_);
    const double lens_r = focus_dist_ * std::tan(core::DegreesToRadians(defocus_angle_ / 2));
    defocus_disk_u_ = u_ * lens_r;
    defocus_disk_v_ = v_ * lens_r;
  }

  int get_image_height() const { return image_height_; }
  int get_image_width() const { return image_width_; }
  virtual bool IsDepthCamera() const { return false; }

  // Narrow the derived frame to the device struct.  Initialize() must have run.
  RtbCamera ToDevice() const {
    RtbCamera d{};
    d.width = image_width_;
    d.height = image_height_;
    auto put = [](float* dst, const core::Vec3& v) { for (int i = 0; i < 3; ++i) dst[i] = static_cast<float>(v[i]); };
    put(d.center, center_);
    put(d.pixel00, pixel00_loc_);
    put(d.delta_u, pixel_delta_u_);
    put(d.delta_v, pixel_delta_v_);
    put(d.disk_u, defocus_disk_u_);
    put(d.disk_v, defocus_disk_v_);
    d.defocus_angle = static_cast<float>(defocus_angle_);
    d.depth_camera = IsDepthCamera() ? 1 : 0;
    return d;
  }

  const core::Point3& center() const { return center_; }
  const core::Point3& pixel00_loc() const { return pixel00_loc_; }
  const core::Vec3& pixel_delta_u() const { return pixel_delta_u_; }
  const core::Vec3& pixel_delta_v() const { return pixel_delta_v_; }

 private:
  int image_height_ = 0;
  core::Point3 center_, pixel00_loc_;
  core::Vec3 pixel_delta_u_, pixel_delta_v_, u_, v_, w_, defocus_disk_u_, defocus_disk_v_;
};

class ColorCamera : public Camera {};
class DepthCamera : public Camera {  // camera.h:215-241: first-hit distance of the normalised ray in red
 public:
  bool IsDepthCamera() const override { return true; }
};

}  // namespace rt::scene
