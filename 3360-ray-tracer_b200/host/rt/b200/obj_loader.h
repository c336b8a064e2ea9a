// OBJ -> geom::Mesh.  Stands where the reference's stale load_obj.h stood (src/load_obj.h:10-56,
// written against tinyobjloader, which is an empty submodule in the reference tree): positions are
// parsed to float like tinyobj's real_t, faces with other than 3 vertices are skipped
// (load_obj.h:32-33), the mesh is centred on its vertex centroid and scaled (load_obj.h:43-52).
// Index forms "i", "i/t", "i/t/n", "i//n" and negative (relative) indices are accepted.
#pragma once
#include <array>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>
#include "../geom/mesh.h"
namespace rt::b200 {

struct ObjData {
  std::vector<core::Point3> vertices;
  std::vector<std::array<int, 3>> faces;
};

inline ObjData ParseObj(const std::string& path) {
  std::ifstream in(path);
  if (!in) throw std::runtime_error("cannot open OBJ '" + path + "'");
  ObjData d;
  std::string line;
  while (std::getline(in, line)) {
    const char* s = line.c_str();
    while (*s == ' ' || *s == '\t') ++s;
    if (s[0] == 'v' && (s[1] == ' ' || s[1] == '\t')) {
      char* end = nullptr;
      float xyz[3] = {0, 0, 0};
      const char* p = s + 1;
      for (int i = 0; i < 3; ++i) {
        xyz[i] = static_cast<float>(std::strtod(p, &end));
        p = end;
      }
      d.vertices.emplace_back(xyz[0], xyz[1], xyz[2]);
    } else if (s[0] == 'f' && (s[1] == ' ' || s[1] == '\t')) {
      std::vector<int> idx;
      const char* p = s + 1;
      for (;;) {
        while (*p == ' ' || *p == '\t' || *p == '\r') ++p;
        if (!*p) break;
        char* end = nullptr;
        long v = std::strtol(p, &end, 10);
        if (end == p) break;
        idx.push_back(v > 0 ? static_cast<int>(v - 1) : static_cast<int>(d.vertices.size()) + static_cast<int>(v));
        p = end;
        while (*p && *p != ' ' && *p != '\t') ++p;  // skip /t/n
      }
      if (idx.size() == 3) d.faces.push_back({idx[0], idx[1], idx[2]});
    }
  }
  for (const auto& f : d.faces)
    for (int i : f)
      if (i < 0 || i >= static_cast<int>(d.vertices.size())) throw std::runtime_error("OBJ face index out of range in '" + path + "'");
  return d;
}

inline std::shared_ptr<geom::Mesh> load_obj(const std::string& filename, std::shared_ptr<material::Material> mat, float scale) {
  ObjData d = ParseObj(filename);
  if (d.vertices.empty()) throw std::runtime_error("OBJ '" + filename + "' has no vertices");
  core::Vec3 centroid(0.0, 0.0, 0.0);
  for (const auto& v : d.vertices) centroid += v;
  centroid /= static_cast<double>(d.vertices.size());
  for (auto& v : d.vertices) {
    v = v - centroid;
    v = v * scale;
  }
  return std::make_shared<geom::Mesh>(d.vertices, d.faces, std::move(mat));
}

}  // namespace rt::b200
