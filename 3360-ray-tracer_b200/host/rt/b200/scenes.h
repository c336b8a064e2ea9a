// Named scene builders (host glue, SURVEY.md §8a row S1 / §8d "Concrete inputs").
// cornell_box / earth / Spheres / checkered_spheres re-create src/main.cc:23-156 object for
// object; bunny and stress are the BASELINE.json configs the reference's main.cc has no builder
// for.  All scene randomness comes from a host mt19937 with a fixed seed, so that the oracle
// (which is fed the exported primitive list) and the GPU see the same scene.
#pragma once
#include <cmath>
#include <memory>
#include <string>
#include <vector>

#include "../core/math_utils.h"
#include "../core/random.h"
#include "../geom/bvh.h"
#include "../geom/rect.h"
#include "../geom/sphere.h"
#include "../geom/triangle.h"
#include "../material/material.h"
#include "../scene/scene.h"
#include "obj_loader.h"

namespace rt::b200 {

struct BuiltScene {
  scene::Scene world;
  std::vector<std::shared_ptr<geom::Bvh>> bvhs;  // every Bvh created, in creation order (for structural tests)
  std::shared_ptr<geom::Bvh> AddBvh(std::vector<std::shared_ptr<geom::Hittable>>& objs) {
    auto b = std::make_shared<geom::Bvh>(objs);
    bvhs.push_back(b);
    return b;
  }
};

using material::Dielectric;
using material::DiffuseLight;
using material::Lambertian;
using material::Metal;
using core::Color;
using core::Point3;

// src/main.cc:23-61
inline void cornell_box(BuiltScene& out) {
  scene::Scene world;
  auto red = std::make_shared<Lambertian>(Color(.65, .05, .05));
  auto white = std::make_shared<Lambertian>(Color(.73, .73, .73));
  auto green = std::make_shared<Lambertian>(Color(.12, .45, .15));
  auto light = std::make_shared<DiffuseLight>(Color(15, 15, 15));
  const double S = 10.0, eps = 0.01, L0 = 3.0, L1 = 7.0;
  world.Add(std::make_shared<geom::yz_rect>(0, S, 0, S, S, green));
  world.Add(std::make_shared<geom::yz_rect>(0, S, 0, S, 0, red));
  world.Add(std::make_shared<geom::xz_rect>(0, S, 0, S, 0, white));
  world.Add(std::make_shared<geom::xz_rect>(0, S, 0, S, S, white));
  world.Add(std::make_shared<geom::xy_rect>(0, S, 0, S, S, white));
  world.Add(std::make_shared<geom::xz_rect>(L0, L1, L0, L1, S - eps, light));
  auto glass = std::make_shared<Dielectric>(1.5);
  auto metal = std::make_shared<Metal>(Color(0.85, 0.85, 0.95), 0.03);
  auto diffuse = std::make_shared<Lambertian>(Color(0.8, 0.3, 0.1));
  world.Add(std::make_shared<geom::Sphere>(Point3(3.2, 1.0, 7.0), 1.0, diffuse));
  world.Add(std::make_shared<geom::Sphere>(Point3(7.0, 1.0, 4.0), 1.0, metal));
  world.Add(std::make_shared<geom::Sphere>(Point3(5.0, 1.0, 2.5), 1.0, glass));
  out.world.Add(out.AddBvh(world.objects_));
}

// src/main.cc:72-145 ("Spheres"), with the grid half-extent as a parameter: main.cc loops
// a,b in [-110,110); BASELINE config 1 ("~480 spheres") is extent 11.  with_earth keeps the 20 %
// earth-textured branch (main.cc:102-103); intro adds the five "intro" spheres (main.cc:84-88).
inline void rtiow_spheres(BuiltScene& out, int extent, unsigned seed, bool with_earth, bool intro,
                          const std::string& earth_path) {
  core::SeedRng(seed);
  scene::Scene world;
  std::shared_ptr<material::Material> earth_surface;
  if (with_earth) earth_surface = std::make_shared<Lambertian>(std::make_shared<material::ImageTexture>(earth_path.c_str()));
  if (intro) {
    world.Add(std::make_shared<geom::Sphere>(Point3(0.0, -100.5, -1.0), 100.0, std::make_shared<Lambertian>(Color(0.8, 0.8, 0.0))));
    world.Add(std::make_shared<geom::Sphere>(Point3(0.0, 0.0, -1.2), 0.5, std::make_shared<Lambertian>(Color(0.1, 0.2, 0.5))));
    world.Add(std::make_shared<geom::Sphere>(Point3(-1.0, 0.0, -1.0), 0.5, std::make_shared<Dielectric>(1.50)));
    world.Add(std::make_shared<geom::Sphere>(Point3(-1.0, 0.0, -1.0), 0.4, std::make_shared<Dielectric>(1.00 / 1.50)));
    world.Add(std::make_shared<geom::Sphere>(Point3(1.0, 0.0, -1.0), 0.5, std::make_shared<Metal>(Color(0.8, 0.6, 0.2), 1.0)));
  }
  auto checker = std::make_shared<material::CheckerTexture>(0.32, Color(0.2, 0.3, 0.1), Color(.9, .9, .9));
  world.Add(std::make_shared<geom::Sphere>(Point3(0, -1000, 0), 1000, std::make_shared<Lambertian>(checker)));
  for (int a = -extent; a < extent; a++) {
    for (int b = -extent; b < extent; b++) {
      const double choose = core::RandomDouble();
      const Point3 center(a + 0.9 * core::RandomDouble(), 0.2, b + 0.9 * core::RandomDouble());
      if ((center - Point3(4, 0.2, 0)).length() <= 0.9) continue;
      std::shared_ptr<material::Material> m;
      if (with_earth && choose < 0.2) {
        m = earth_surface;
      } else if (choose < 0.8) {
        m = std::make_shared<Lambertian>(core::RandomVec3(0, 1));
      } else if (choose < 0.95) {
        const Color albedo = core::RandomVec3(0.5, 1);
        const double fuzz = core::RandomDouble(0, 0.5);
        m = std::make_shared<Metal>(albedo, fuzz);
      } else {
        m = std::make_shared<Dielectric>(1.5);
      }
      world.Add(std::make_shared<geom::Sphere>(center, 0.2, m));
    }
  }
  world.Add(std::make_shared<geom::Sphere>(Point3(0, 1, 0), 1.0, std::make_shared<Dielectric>(1.5)));
  world.Add(std::make_shared<geom::Sphere>(Point3(-4, 0, 0), 1.0, std::make_shared<Lambertian>(Color(0.4, 0.2, 0.1))));
  world.Add(std::make_shared<geom::Sphere>(Point3(4, 1, 0), 1.0, std::make_shared<Metal>(Color(0.7, 0.6, 0.5), 0.0)));
  out.world.Add(out.AddBvh(world.objects_));
}

// src/main.cc:147-156
inline void checkered_spheres(BuiltScene& out) {
  scene::Scene world;
  auto checker = std::make_shared<material::CheckerTexture>(0.32, Color(.2, .3, .1), Color(.9, .9, .9));
  world.Add(std::make_shared<geom::Sphere>(Point3(0, -10, 0), 10, std::make_shared<Lambertian>(checker)));
  world.Add(std::make_shared<geom::Sphere>(Point3(0, 10, 0), 10, std::make_shared<Lambertian>(checker)));
  out.world.Add(out.AddBvh(world.objects_));
}

// src/main.cc:64-70: the bare globe, not wrapped in a BVH (scene::Scene holds it directly).
inline void earth(BuiltScene& out, const std::string& earth_path) {
  auto tex = std::make_shared<material::ImageTexture>(earth_path.c_str());
  out.world = scene::Scene(std::make_shared<geom::Sphere>(Point3(0, 0, 0), 2, std::make_shared<Lambertian>(tex)));
}

// BASELINE config 4: the globe of main.cc:64-70, the hollow-glass "air bubble" pair of
// main.cc:80-87 (Dielectric 1.5 r=0.5 around Dielectric 1/1.5 r=0.4) in front of it, and a ground.
inline void earth_bubbles(BuiltScene& out, const std::string& earth_path) {
  scene::Scene world;
  auto tex = std::make_shared<material::ImageTexture>(earth_path.c_str());
  world.Add(std::make_shared<geom::Sphere>(Point3(0, 0, 0), 2, std::make_shared<Lambertian>(tex)));
  const Point3 bubble(0.45, 0.95, 2.35);
  world.Add(std::make_shared<geom::Sphere>(bubble, 0.5, std::make_shared<Dielectric>(1.50)));
  world.Add(std::make_shared<geom::Sphere>(bubble, 0.4, std::make_shared<Dielectric>(1.00 / 1.50)));
  world.Add(std::make_shared<geom::Sphere>(Point3(0, -1002, 0), 1000, std::make_shared<Lambertian>(Color(0.5, 0.5, 0.5))));
  out.world.Add(out.AddBvh(world.objects_));
}

// BASELINE config 3: the mesh the commented-out block of main.cc:135-142 loads — centred,
// scaled x50, Lambertian(0.8,0.1,0.1), in its own BVH — plus a ground sphere touching the mesh's
// lowest point, both inside an outer BVH.
inline void bunny(BuiltScene& out, const std::string& obj_path, double scale) {
  auto red = std::make_shared<Lambertian>(Color(0.8, 0.1, 0.1));
  auto mesh = load_obj(obj_path, red, static_cast<float>(scale));
  auto mesh_bvh = out.AddBvh(mesh->tris.objects_);
  const double floor_y = mesh_bvh->BoundingBox().y.min_;
  scene::Scene world;
  world.Add(mesh_bvh);
  world.Add(std::make_shared<geom::Sphere>(Point3(0, floor_y - 1000.0, 0), 1000.0, std::make_shared<Lambertian>(Color(0.5, 0.5, 0.5))));
  out.world.Add(out.AddBvh(world.objects_));
}

namespace detail {
inline double Hash01(int x, int z, unsigned seed) {  // integer lattice hash -> [0,1)
  uint32_t h = static_cast<uint32_t>(x) * 0x8da6b343u ^ static_cast<uint32_t>(z) * 0xd8163841u ^ seed * 0xcb1ab31fu;
  h ^= h >> 16; h *= 0x7feb352du; h ^= h >> 15; h *= 0x846ca68bu; h ^= h >> 16;
  return h * (1.0 / 4294967296.0);
}
inline double ValueNoise(double x, double z, unsigned seed) {  // bilinear, smoothstep-faded
  const double fx = std::floor(x), fz = std::floor(z);
  const int ix = static_cast<int>(fx), iz = static_cast<int>(fz);
  double tx = x - fx, tz = z - fz;
  tx = tx * tx * (3 - 2 * tx);
  tz = tz * tz * (3 - 2 * tz);
  const double a = Hash01(ix, iz, seed), b = Hash01(ix + 1, iz, seed), c = Hash01(ix, iz + 1, seed), d = Hash01(ix + 1, iz + 1, seed);
  return (a + (b - a) * tx) + ((c + (d - c) * tx) - (a + (b - a) * tx)) * tz;
}
}  // namespace detail

// BASELINE config 5: nx x nz quads of a value-noise height field over
// [-half,half]^2 (2 triangles each) in one BVH, plus an ns x ns jittered grid of r=0.3 spheres
// with the RTIOW material mix floating above it in a second BVH.  nx=1415, nz=1414, ns=1000,
// half=500 gives 4 001 620 triangles + 1 000 000 spheres.
inline void stress(BuiltScene& out, int nx, int nz, int ns, double half, unsigned seed) {
  auto ground = std::make_shared<Lambertian>(Color(0.45, 0.5, 0.4));
  std::vector<std::shared_ptr<geom::Hittable>> tris;
  tris.reserve(static_cast<size_t>(nx) * nz * 2);
  auto height = [&](double x, double z) { return 2.0 * detail::ValueNoise(x / 8.0, z / 8.0, 7u); };
  auto vertex = [&](int i, int k) {
    const double x = -half + (2.0 * half) * i / nx, z = -half + (2.0 * half) * k / nz;
    return Point3(x, height(x, z), z);
  };
  for (int k = 0; k < nz; ++k)
    for (int i = 0; i < nx; ++i) {
      const Point3 p00 = vertex(i, k), p10 = vertex(i + 1, k), p01 = vertex(i, k + 1), p11 = vertex(i + 1, k + 1);
      tris.push_back(std::make_shared<geom::Triangle>(p00, p01, p11, ground));
      tris.push_back(std::make_shared<geom::Triangle>(p00, p11, p10, ground));
    }
  core::SeedRng(seed);
  std::vector<std::shared_ptr<geom::Hittable>> balls;
  balls.reserve(static_cast<size_t>(ns) * ns);
  auto glass = std::make_shared<Dielectric>(1.5);
  for (int k = 0; k < ns; ++k)
    for (int i = 0; i < ns; ++i) {
      const double cell = (2.0 * half) / ns;
      const double x = -half + (i + 0.2 + 0.6 * core::RandomDouble()) * cell, z = -half + (k + 0.2 + 0.6 * core::RandomDouble()) * cell;
      const double choose = core::RandomDouble();
      std::shared_ptr<material::Material> m;
      if (choose < 0.8) m = std::make_shared<Lambertian>(core::RandomVec3(0, 1));
      else if (choose < 0.95) { const Color al = core::RandomVec3(0.5, 1); m = std::make_shared<Metal>(al, core::RandomDouble(0, 0.5)); }
      else m = glass;
      balls.push_back(std::make_shared<geom::Sphere>(Point3(x, height(x, z) + 0.3 + 0.5 * core::RandomDouble(), z), 0.3, m));
    }
  scene::Scene world;
  if (!tris.empty()) world.Add(out.AddBvh(tris));
  if (!balls.empty()) world.Add(out.AddBvh(balls));
  out.world.Add(out.AddBvh(world.objects_));
}

}  // namespace rt::b200
