// The three scene classes and their loaders, with the reference's names and members:
//   SDFGrid + loadSDFGrid         src/grid_raytracing.hpp:10-22, .cpp:127-134
//   SDFOctree + loadSDFOctree     src/octree_raytracing.hpp:8-49, .cpp:8-16
//   cmesh4::SimpleMesh, LoadMeshFromObj (src/core/mesh.h:15-58, mesh.cpp:178-286), loadAndScale (main.cpp:326-343),
//   BVHBuilder                    src/triangles_raytracing.hpp:38-45
// The host data members stay public as in the reference (size/values, nodes); `upload()` (called by the loaders and by
// BVHBuilder::perform) sends them to the GPU once.  Unlike the reference's loaders, these check the file.
#pragma once
#include <cstdio>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>
#include <string>
#include <tuple>
#include <vector>

#include "raytracing.hpp"

struct SDFGrid final : DeviceScene {
  rtmath::uint3 size;
  std::vector<float> values;
  float sdf(rtmath::uint3 c) const noexcept { return values[(c.x * size.y + c.y) * size.z + c.z]; }
  void upload(int layout = RT_GRID_AUTO) {
    if (values.size() != static_cast<size_t>(size.x) * size.y * size.z) throw std::runtime_error("SDFGrid: size/values mismatch");
    rt_scene *s = nullptr;
    rthost::check(rt_scene_create_grid(values.data(), size.x, size.y, size.z, layout, &s), "SDFGrid::upload");
    adopt(s);
  }
};

inline void loadSDFGrid(SDFGrid &scene, const std::string &path) {
  std::ifstream fs(path, std::ios::binary);
  if (!fs) throw std::runtime_error("loadSDFGrid: cannot open " + path);
  uint32_t dims[3];
  fs.read(reinterpret_cast<char *>(dims), sizeof dims);
  if (!fs) throw std::runtime_error("loadSDFGrid: truncated header in " + path);
  scene.size = {dims[0], dims[1], dims[2]};
  const size_t n = static_cast<size_t>(dims[0]) * dims[1] * dims[2];
  scene.values.resize(n);
  fs.read(reinterpret_cast<char *>(scene.values.data()), static_cast<std::streamsize>(n * sizeof(float)));
  if (!fs) throw std::runtime_error("loadSDFGrid: file shorter than its header says: " + path);
  scene.upload();
}

struct SDFOctreeNode {  // octree_raytracing.hpp:8-18 (36 bytes, read raw from disk)
  float values[8];
  uint32_t childrenOffset = 0;
  bool isLeaf() const noexcept { return childrenOffset == 0; }
};
static_assert(sizeof(SDFOctreeNode) == 36, "SDFOctreeNode must stay 36 bytes");

struct SDFOctree final : DeviceScene {
  std::vector<SDFOctreeNode> nodes;
  void upload() {
    rt_scene *s = nullptr;
    rthost::check(rt_scene_create_octree(nodes.data(), static_cast<uint32_t>(nodes.size()), &s), "SDFOctree::upload");
    adopt(s);
  }
};

inline void loadSDFOctree(SDFOctree &scene, const std::string &path) {
  std::ifstream fs(path, std::ios::binary);
  if (!fs) throw std::runtime_error("loadSDFOctree: cannot open " + path);
  uint32_t n = 0;
  fs.read(reinterpret_cast<char *>(&n), sizeof n);
  if (!fs) throw std::runtime_error("loadSDFOctree: truncated header in " + path);
  scene.nodes.resize(n);
  fs.read(reinterpret_cast<char *>(scene.nodes.data()), static_cast<std::streamsize>(sizeof(SDFOctreeNode) * n));
  if (!fs) throw std::runtime_error("loadSDFOctree: file shorter than its node count says: " + path);
  scene.upload();
}

namespace cmesh4 {
struct SimpleMesh {  // mesh.h:15-55; only positions and indices reach the hot path (triangles_raytracing.cpp:301-309)
  SimpleMesh() {}
  SimpleMesh(size_t nVerts, size_t nIdx) { vPos4f.resize(nVerts); indices.resize(nIdx); }
  size_t VerticesNum() const { return vPos4f.size(); }
  size_t IndicesNum() const { return indices.size(); }
  size_t TrianglesNum() const { return indices.size() / 3; }
  std::vector<rtmath::float4> vPos4f;
  std::vector<unsigned int> indices;
};

// Minimal OBJ reader with LoadMeshFromObj's vertex rule (mesh.cpp:231-267): one vertex per distinct
// (position, normal, texcoord) index triple, numbered in order of first use; faces in file order, polygons as fans.
inline SimpleMesh LoadMeshFromObj(const char *fileName, bool verbose = false) {
  SimpleMesh mesh;
  std::ifstream fs(fileName);
  if (!fs) {
    std::printf("[LoadMeshFromObj::ERROR] Failed to load obj file: %s\n", fileName);
    return mesh;  // the reference returns an empty mesh too (mesh.cpp:193-197)
  }
  std::vector<float> pos;
  int nTex = 0, nNrm = 0;
  std::map<std::tuple<int, int, int>, uint32_t> unique;
  std::string line;
  auto fix = [](int idx, int count) { return idx > 0 ? idx - 1 : (idx < 0 ? count + idx : -1); };
  while (std::getline(fs, line)) {
    std::istringstream ss(line);
    std::string tag;
    ss >> tag;
    if (tag == "v") {
      double x = 0, y = 0, z = 0;
      ss >> x >> y >> z;
      pos.push_back(static_cast<float>(x));
      pos.push_back(static_cast<float>(y));
      pos.push_back(static_cast<float>(z));
    } else if (tag == "vt") {
      ++nTex;
    } else if (tag == "vn") {
      ++nNrm;
    } else if (tag == "f") {
      std::vector<uint32_t> poly;
      std::string item;
      while (ss >> item) {
        int v = 0, vt = 0, vn = 0;
        const char *p = item.c_str();
        v = std::atoi(p);
        if (const char *s1 = std::strchr(p, '/')) {
          if (s1[1] != '/' && s1[1] != '\0') vt = std::atoi(s1 + 1);
          if (const char *s2 = std::strchr(s1 + 1, '/')) vn = std::atoi(s2 + 1);
        }
        const int nPos = static_cast<int>(pos.size() / 3);
        auto key = std::make_tuple(fix(v, nPos), fix(vn, nNrm), fix(vt, nTex));
        if (std::get<0>(key) < 0 || std::get<0>(key) >= nPos) throw std::runtime_error("LoadMeshFromObj: bad vertex index");
        auto it = unique.find(key);
        uint32_t id;
        if (it == unique.end()) {
          id = static_cast<uint32_t>(mesh.vPos4f.size());
          unique.emplace(key, id);
          const float *q = &pos[3 * static_cast<size_t>(std::get<0>(key))];
          mesh.vPos4f.push_back({q[0], q[1], q[2], 1.0f});
        } else {
          id = it->second;
        }
        poly.push_back(id);
      }
      for (size_t k = 1; k + 1 < poly.size(); ++k) {
        mesh.indices.push_back(poly[0]);
        mesh.indices.push_back(poly[k]);
        mesh.indices.push_back(poly[k + 1]);
      }
    }
  }
  if (verbose)
    std::printf("[LoadMeshFromObj::INFO] Loaded obj file %s with %d vertices and %d indices\n", fileName,
                (unsigned)mesh.vPos4f.size(), (unsigned)mesh.indices.size());
  return mesh;
}
}  // namespace cmesh4

// main.cpp:326-343 + calc_bbox (raytracing.hpp:22-37): centre the mesh and scale it into the unit sphere
inline cmesh4::SimpleMesh loadAndScale(const std::string &path) {
  auto mesh = cmesh4::LoadMeshFromObj(path.c_str(), false);
  const float inf = std::numeric_limits<float>::infinity();
  rtmath::float3 lo{inf, inf, inf}, hi{-inf, -inf, -inf};
  for (auto &v : mesh.vPos4f) {
    rtmath::float3 p{v.x / v.w, v.y / v.w, v.z / v.w};
    for (int a = 0; a < 3; ++a) {
      lo[a] = std::min(lo[a], p[a]);
      hi[a] = std::max(hi[a], p[a]);
    }
  }
  rtmath::float3 center = (lo + hi) / 2.0f;
  float scale = rtmath::length(hi - center);
  for (auto &v : mesh.vPos4f) {
    float w = v.w;
    rtmath::float3 p{v.x / w, v.y / w, v.z / w};
    p = p - center;
    p = p / scale;
    v = {p.x * w, p.y * w, p.z * w, 1.0f * w};
  }
  return mesh;
}

class BVHBuilder final : public DeviceScene {  // triangles_raytracing.hpp:38-45
public:
  void perform(cmesh4::SimpleMesh mesh) {
    m_mesh = std::move(mesh);
    rt_scene *s = nullptr;
    rthost::check(rt_scene_create_mesh(&m_mesh.vPos4f[0].x, static_cast<uint32_t>(m_mesh.vPos4f.size()), m_mesh.indices.data(),
                                       static_cast<uint32_t>(m_mesh.indices.size()), RT_MESH_AUTO, &s),
                  "BVHBuilder::perform");
    adopt(s);
  }
  cmesh4::SimpleMesh &&result() { return std::move(m_mesh); }
  float modelMinY() const {  // main.cpp:176-189 puts the ground plane at modelBox.boxMin.y
    float lo[3], hi[3];
    rthost::check(rt_scene_info(handle(), nullptr, lo, hi), "rt_scene_info");
    return lo[1];
  }

private:
  cmesh4::SimpleMesh m_mesh;
};
