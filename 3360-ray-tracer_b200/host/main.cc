// rt_b200 — the reference's main() (src/main.cc:158-197) on the B200 path: load cameras.json,
// pick the preset named by argv[1] ("default" if absent or unknown), build the scene, render,
// PPM on stdout, "Runtime: Xs" on stderr.  The reference selects its scene with a compile-time
// switch (main.cc:178); here it is --scene (default: cornell, the reference's case 4).
//   rt_b200 [preset] [--scene cornell|rtiow|rtiow110|checkered|earth|earth_bubbles|bunny|stress]
//           [--cameras file.json] [--mode A|B] [--spp N] [--depth N] [--seed N] [--assets dir]
//           [--adaptive [threshold [min_samples]]]   per-pixel stop of the reference's WavefrontRenderer
//           [--gpus N] [--width N] [--p6]  binary PPM  [--png]  PNG   [--dump file] [--resume file]  progressive accumulator
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <iomanip>
#include <iostream>
#include <string>

#include "rt/b200/scenes.h"
#include "rt/core/timer.h"
#include "rt/renderer/b200_renderer.h"
#include "rt/scene/camera.h"

using namespace rt;

int main(int argc, char** argv) {
  core::Timer clock;
  clock.reset();
  std::string active = "default", scene_name = "cornell", cameras_file = "cameras.json", assets = "assets";
  renderer::B200Options opt;
  bool preset_given = false;
  int width = 0;  // --width overrides the preset's image_width (tests)
  for (int i = 1; i < argc; ++i) {
    const std::string a = argv[i];
    auto next = [&]() -> std::string { return i + 1 < argc ? argv[++i] : std::string(); };
    if (a == "--scene") scene_name = next();
    else if (a == "--cameras") cameras_file = next();
    else if (a == "--assets") assets = next();
    else if (a == "--mode") opt.mode = (next() == "B") ? RTB_MODE_B : RTB_MODE_A;
    else if (a == "--spp") opt.samples_per_pixel = std::atoi(next().c_str());
    else if (a == "--depth") opt.max_depth = std::atoi(next().c_str());
    else if (a == "--seed") opt.seed = std::strtoull(next().c_str(), nullptr, 0);
    else if (a == "--gpus") opt.gpus = std::max(1, std::atoi(next().c_str()));
    else if (a == "--width") width = std::atoi(next().c_str());
    else if (a == "--p6") opt.binary_ppm = true;
    else if (a == "--png") opt.png = true;
    else if (a == "--dump") opt.dump_to = next();
    else if (a == "--resume") opt.resume_from = next();
    else if (a == "--adaptive") {
      opt.adaptive = true;
      if (i + 1 < argc && std::atof(argv[i + 1]) > 0.0) opt.rel_threshold = static_cast<float>(std::atof(argv[++i]));
      if (i + 1 < argc && std::atoi(argv[i + 1]) > 0) opt.min_samples = std::atoi(argv[++i]);
    }
    else if (!preset_given) { active = a; preset_given = true; }
  }
  try {
    std::unordered_map<std::string, scene::CameraConfig> cameras;
    try {
      cameras = scene::loadCameras(cameras_file);
    } catch (const std::exception&) {
      cameras = scene::loadCameras(assets + "/cameras.json");
    }
    if (!cameras.count(active)) {
      std::cerr << "Camera '" << active << "' not found. Using default.\n";
      active = "default";
    }
    scene::ColorCamera cam;
    if (width > 0) cameras[active].image_width = width;
    cam.SetFromConfig(cameras[active]);
    cam.Initialize();

    b200::BuiltScene built;
    const std::string earth = assets + "/textures/earthmap.jpg";
    if (scene_name == "cornell") b200::cornell_box(built);
    else if (scene_name == "rtiow") b200::rtiow_spheres(built, 11, 1, false, false, earth);
    else if (scene_name == "rtiow110") b200::rtiow_spheres(built, 110, 1, true, true, earth);  // main.cc:72-145 as shipped
    else if (scene_name == "checkered") b200::checkered_spheres(built);
    else if (scene_name == "earth") b200::earth(built, earth);
    else if (scene_name == "earth_bubbles") b200::earth_bubbles(built, earth);
    else if (scene_name == "bunny") b200::bunny(built, assets + "/models/stanford-bunny.obj", 50);
    else if (scene_name == "stress") b200::stress(built, 1415, 1414, 1000, 500, 2);
    else { std::cerr << "unknown scene '" << scene_name << "'\n"; return 2; }

    renderer::B200Renderer(built.world, cam, opt).Render();
  } catch (const std::exception& e) {
    std::cerr << "error: " << e.what() << "\n";
    return 1;
  }
  std::clog << "Runtime: " << std::setprecision(2) << clock.elapsed() << "s" << std::flush;
  return 0;
}
