// TEST INFRASTRUCTURE — not part of the product path.
//
// Headless C-API driver around the UNMODIFIED reference sources (compiled where they lie under
// /root/reference by oracle/Makefile; nothing from the reference is copied into this repo).
// It gives tests/, bench.py's cpu_baseline leg and __graft_entry__.smoke() a way to run the
// reference's own Renderer::draw / IScene::intersect without the SDL/ImGui viewer:
//   * scene construction through loadSDFGrid / loadSDFOctree / BVHBuilder::perform,
//   * the viewer's scene normalisation (main.cpp:326-343 loadAndScale) and camera/projection
//     constants (main.cpp:89-91,198-201) restated here because main.cpp itself cannot be built,
//   * Renderer::draw (raytracing.cpp:67-102) for colour + depth and its own stopwatch,
//   * single-ray and per-primitive entry points for known-answer tests,
//   * the reference-built BVH8 (private members of BVHBuilder) exported so that the CUDA path can
//     traverse the very same tree and be compared bit for bit.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <execution>
#include <filesystem>
#include <fstream>
#include <iostream>
#include <limits>
#include <map>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include <omp.h>

// The BVH node array and the permuted mesh are private members (triangles_raytracing.hpp:60-66).
// Changing the access specifier does not change the class layout; this TU only reads them.
#define private public
#include "triangles_raytracing.hpp"
#undef private
#include "grid_raytracing.hpp"
#include "octree_raytracing.hpp"
#include "raytracing.hpp"
#include <ray_pack_ispc.h>

using namespace LiteMath;

namespace {

enum SceneKind : int { KIND_GRID = 0, KIND_OCTREE = 1, KIND_MESH = 2, KIND_UNION = 3 };

struct TripleKey {
  uint32_t a, b, c;
  bool operator<(const TripleKey &o) const {
    return std::tie(a, b, c) < std::tie(o.a, o.b, o.c);
  }
};

struct VertKey {
  uint32_t w[9];
  bool operator<(const VertKey &o) const { return std::memcmp(w, o.w, sizeof(w)) < 0; }
};

} // namespace

struct RefScene {
  int kind = -1;
  std::shared_ptr<IScene> scene;
  std::shared_ptr<BVHBuilder> bvh;           // KIND_MESH
  std::vector<uint32_t> primOfPermuted;      // KIND_MESH: original triangle id per BVH-order triangle
  std::map<VertKey, uint32_t> primOfVerts;   // KIND_MESH: (v0,v1,v2) bit patterns -> original triangle id
  float modelMinY = -1.0f;                   // ground plane height (main.cpp:176-189)
};

struct RefCam {
  float eye[3];
  float target[3];
  float up[3];
  int32_t lockUp;   // main.cpp:91 sets true
  int32_t nRotate;  // number of Camera::rotate(dx, dy) calls applied after construction
  float dx, dy;
};

namespace {

Camera makeCamera(const RefCam *c) {
  Camera cam(float3{c->eye[0], c->eye[1], c->eye[2]}, float3{c->target[0], c->target[1], c->target[2]},
             float3{c->up[0], c->up[1], c->up[2]});
  cam.setLockUp(c->lockUp != 0);
  for (int i = 0; i < c->nRotate; ++i)
    cam.rotate(c->dx, c->dy);
  return cam;
}

float4x4 toMat(const float m[16]) {
  float4x4 r;
  std::memcpy(&r, m, sizeof(float) * 16);
  return r;
}

// main.cpp:326-343 (loadAndScale), applied to an already loaded mesh.
void scaleToUnitSphere(cmesh4::SimpleMesh &mesh) {
  auto bbox = calc_bbox(mesh);
  auto center = (bbox.boxMin + bbox.boxMax) / 2.0f;
  auto scale = length(bbox.boxMax - center);
  for (auto &v : mesh.vPos4f) {
    auto w = v.w;
    v /= w;
    float3 scaled = to_float3(v);
    scaled -= center;
    scaled /= scale;
    v = to_float4(scaled, 1.0f);
    v *= w;
  }
}

RefScene *finishMesh(cmesh4::SimpleMesh mesh) {
  auto *s = new RefScene;
  s->kind = KIND_MESH;
  s->modelMinY = calc_bbox(mesh).boxMin.y;
  // original-order triangle ids, keyed by the index triple (the builder moves triples as units,
  // triangles_raytracing.cpp:18-20,35-47)
  std::multimap<TripleKey, uint32_t> ids;
  const size_t nTri = mesh.indices.size() / 3;
  for (size_t i = 0; i < nTri; ++i)
    ids.insert({TripleKey{mesh.indices[3 * i], mesh.indices[3 * i + 1], mesh.indices[3 * i + 2]},
                static_cast<uint32_t>(i)});
  s->bvh = std::make_shared<BVHBuilder>();
  s->bvh->perform(std::move(mesh));
  s->scene = s->bvh;
  const auto &m = s->bvh->m_mesh;
  s->primOfPermuted.resize(nTri);
  for (size_t i = 0; i < nTri; ++i) {
    TripleKey k{m.indices[3 * i], m.indices[3 * i + 1], m.indices[3 * i + 2]};
    auto it = ids.find(k);
    uint32_t id = 0xFFFFFFFFu;
    if (it != ids.end()) {
      id = it->second;
      ids.erase(it);
    }
    s->primOfPermuted[i] = id;
    VertKey vk;
    for (int v = 0; v < 3; ++v) {
      float4 p = m.vPos4f[(&k.a)[v]];
      p /= p.w;
      std::memcpy(&vk.w[3 * v], &p, 12);
    }
    auto ins = s->primOfVerts.insert({vk, id});
    if (!ins.second && id < ins.first->second)
      ins.first->second = id; // duplicate geometry: smallest original id names it
  }
  return s;
}

} // namespace

extern "C" {

void ref_set_threads(int n) { omp_set_num_threads(n > 0 ? n : omp_get_num_procs()); }
int ref_max_threads() { return omp_get_max_threads(); }

// ---------------------------------------------------------------- scenes
RefScene *ref_scene_grid_file(const char *path) {
  auto g = std::make_shared<SDFGrid>();
  loadSDFGrid(*g, path);
  auto *s = new RefScene;
  s->kind = KIND_GRID;
  s->scene = g;
  return s;
}

RefScene *ref_scene_grid(const float *values, uint32_t sx, uint32_t sy, uint32_t sz) {
  auto g = std::make_shared<SDFGrid>();
  g->size = uint3{sx, sy, sz};
  g->values.assign(values, values + static_cast<size_t>(sx) * sy * sz);
  auto *s = new RefScene;
  s->kind = KIND_GRID;
  s->scene = g;
  return s;
}

RefScene *ref_scene_octree_file(const char *path) {
  auto o = std::make_shared<SDFOctree>();
  loadSDFOctree(*o, path);
  auto *s = new RefScene;
  s->kind = KIND_OCTREE;
  s->scene = o;
  return s;
}

RefScene *ref_scene_octree(const void *nodes36, uint32_t n) {
  static_assert(sizeof(SDFOctreeNode) == 36, "SDFOctreeNode must be 36 bytes");
  auto o = std::make_shared<SDFOctree>();
  o->nodes.resize(n);
  std::memcpy(o->nodes.data(), nodes36, static_cast<size_t>(n) * 36);
  auto *s = new RefScene;
  s->kind = KIND_OCTREE;
  s->scene = o;
  return s;
}

// vpos4: nVerts x float4; idx: nIdx (multiple of 3). scale != 0 applies loadAndScale's normalisation.
RefScene *ref_scene_mesh(const float *vpos4, uint32_t nVerts, const uint32_t *idx, uint32_t nIdx,
                         int scale) {
  cmesh4::SimpleMesh mesh(nVerts, nIdx);
  std::memcpy(mesh.vPos4f.data(), vpos4, static_cast<size_t>(nVerts) * 16);
  std::memcpy(mesh.indices.data(), idx, static_cast<size_t>(nIdx) * 4);
  if (scale)
    scaleToUnitSphere(mesh);
  return finishMesh(std::move(mesh));
}

RefScene *ref_scene_mesh_obj(const char *path, int scale) {
  auto mesh = cmesh4::LoadMeshFromObj(path, false);
  if (mesh.indices.empty())
    return nullptr;
  if (scale)
    scaleToUnitSphere(mesh);
  return finishMesh(std::move(mesh));
}

// main.cpp:189-190: fullScene = SceneUnion(pScene, Plane({0,1,0}, modelBox.boxMin.y))
RefScene *ref_scene_with_plane(const RefScene *base, float planeY) {
  auto *s = new RefScene;
  s->kind = KIND_UNION;
  s->scene = std::make_shared<SceneUnion>(base->scene,
                                          std::make_shared<Plane>(float3{0.0f, 1.0f, 0.0f}, planeY));
  return s;
}

float ref_scene_model_min_y(const RefScene *s) { return s->modelMinY; }
int ref_scene_kind(const RefScene *s) { return s->kind; }
void ref_scene_free(RefScene *s) { delete s; }

// ---------------------------------------------------------------- mesh / BVH export
void ref_mesh_counts(const RefScene *s, uint32_t *nVerts, uint32_t *nIdx, uint32_t *nNodes,
                     uint32_t *nodeBytes) {
  *nVerts = static_cast<uint32_t>(s->bvh->m_mesh.vPos4f.size());
  *nIdx = static_cast<uint32_t>(s->bvh->m_mesh.indices.size());
  *nNodes = static_cast<uint32_t>(s->bvh->m_nodes.size());
  *nodeBytes = static_cast<uint32_t>(sizeof(BVH8Node));
}

// Unused child lanes of an inner node are uninitialised stack bytes in the reference.
// They are exported as they are; the consumer must ignore lanes >= realCount.
void ref_mesh_export(const RefScene *s, float *vpos4, uint32_t *idxPermuted, void *nodes,
                     uint32_t *primOfPermuted) {
  const auto &m = s->bvh->m_mesh;
  std::memcpy(vpos4, m.vPos4f.data(), m.vPos4f.size() * 16);
  std::memcpy(idxPermuted, m.indices.data(), m.indices.size() * 4);
  std::memcpy(nodes, s->bvh->m_nodes.data(), s->bvh->m_nodes.size() * sizeof(BVH8Node));
  std::memcpy(primOfPermuted, s->primOfPermuted.data(), s->primOfPermuted.size() * 4);
}

// ---------------------------------------------------------------- camera / projection
void ref_camera(const RefCam *c, float outPos[3], float outView[16], float outViewInv[16]) {
  Camera cam = makeCamera(c);
  float3 p = cam.position();
  outPos[0] = p.x;
  outPos[1] = p.y;
  outPos[2] = p.z;
  auto view = cam.lookAtMatrix();
  auto viewInv = inverse4x4(view);
  std::memcpy(outView, &view, 64);
  std::memcpy(outViewInv, &viewInv, 64);
}

// main.cpp:198-201
void ref_projection(float fovy, float aspect, float zNear, float zFar, float outProj[16],
                    float outProjInv[16]) {
  auto proj = perspectiveMatrix(fovy, aspect, zNear, zFar);
  auto projInv = inverse4x4(proj);
  std::memcpy(outProj, &proj, 64);
  std::memcpy(outProjInv, &projInv, 64);
}

// ---------------------------------------------------------------- rendering
// color/t are in-out (W*H each, row-major, memory row 0 = top): exactly FrameBuffer's contents.
// mode: 0 Normal, 1 Lambert, 2 Color (raytracing.hpp:99 order).
float ref_render(const RefScene *s, const RefCam *c, uint32_t W, uint32_t H, const float projInv[16],
                 int mode, const float light[3], int shadows, int reflections, uint32_t *color,
                 float *t) {
  Renderer r;
  r.lightPos = float3{light[0], light[1], light[2]};
  r.enableShadows = shadows != 0;
  r.enableReflections = reflections != 0;
  r.shadingMode = static_cast<ShadingMode>(mode);
  FrameBuffer fb;
  fb.resize(W, H);
  std::memcpy(fb.color.data(), color, static_cast<size_t>(W) * H * 4);
  std::memcpy(fb.t.data(), t, static_cast<size_t>(W) * H * 4);
  Camera cam = makeCamera(c);
  float ms = r.draw(*s->scene, fb, cam, toMat(projInv));
  std::memcpy(color, fb.color.data(), static_cast<size_t>(W) * H * 4);
  std::memcpy(t, fb.t.data(), static_cast<size_t>(W) * H * 4);
  return ms;
}

// Timing variant: draws `frames` times into a fresh cleared framebuffer, returns the best
// Renderer::draw stopwatch value (ms). Optional outputs receive the last frame.
float ref_render_timed(const RefScene *s, const RefCam *c, uint32_t W, uint32_t H,
                       const float projInv[16], int mode, const float light[3], int shadows,
                       int reflections, int frames, float *msEach) {
  Renderer r;
  r.lightPos = float3{light[0], light[1], light[2]};
  r.enableShadows = shadows != 0;
  r.enableReflections = reflections != 0;
  r.shadingMode = static_cast<ShadingMode>(mode);
  FrameBuffer fb;
  fb.resize(W, H);
  Camera cam = makeCamera(c);
  auto pi = toMat(projInv);
  float best = std::numeric_limits<float>::infinity();
  for (int f = 0; f < frames; ++f) {
    fb.clear();
    float ms = r.draw(*s->scene, fb, cam, pi);
    if (msEach)
      msEach[f] = ms;
    best = std::min(best, ms);
  }
  return best;
}

// Renders a horizontal band of loop rows [y0, y1) of raytracing.cpp:80 (y = 0 is the BOTTOM of the
// view, stored in memory row H-1-y) by calling the reference's per-ray code directly; used by
// bench.py to time a bounded sample of a large frame. Writes into full-size W*H buffers.
float ref_render_rows(const RefScene *s, const RefCam *c, uint32_t W, uint32_t H,
                      const float projInv[16], uint32_t y0, uint32_t y1, uint32_t *color, float *t);

struct PrimRec {
  std::vector<RefTriCandidate> cands;
};

static void recordCand(const RefTriCandidate *c, void *user) {
  static_cast<PrimRec *>(user)->cands.push_back(*c);
}

// Primary-ray closest hit per pixel through IScene::intersect with the ray generation of
// raytracing.cpp:82-88 (tFar = min(100, +inf)). Outputs: tOut (+inf on miss), normalOut (3 floats
// per pixel, the unflipped HitInfo::normal), primOut (mesh scenes: original triangle id of the
// winner = first reported candidate in call order whose t equals the final t; 0xFFFFFFFF on miss
// or for non-mesh scenes).
void ref_render_prim(const RefScene *s, const RefCam *c, uint32_t W, uint32_t H,
                     const float projInv[16], float *tOut, float *normalOut, uint32_t *primOut) {
  Camera cam = makeCamera(c);
  float3 rayPos = cam.position();
  auto viewInv = inverse4x4(cam.lookAtMatrix());
  auto pi = toMat(projInv);
  const int width = static_cast<int>(W), height = static_cast<int>(H);
#pragma omp parallel
  {
    PrimRec rec;
    if (s->kind == KIND_MESH)
      ref_standin_set_recorder(recordCand, &rec);
#pragma omp for schedule(dynamic)
    for (int y = 0; y < height; ++y) {
      for (int x = 0; x < width; ++x) {
        size_t o = static_cast<size_t>(height - y - 1) * W + x;
        float4 d4 = EyeRayDir4f(static_cast<float>(x) + 0.5f, static_cast<float>(y) + 0.5f,
                                static_cast<float>(width), static_cast<float>(height), pi);
        d4.w = 0.0f;
        d4 = viewInv * d4;
        float3 rayDir = to_float3(d4);
        rec.cands.clear();
        HitInfo hit = s->scene->intersect(rayPos, rayDir, 0.01f,
                                          std::min(100.0f, std::numeric_limits<float>::infinity()));
        uint32_t prim = 0xFFFFFFFFu;
        if (hit.hitten && s->kind == KIND_MESH) {
          for (const auto &cd : rec.cands) {
            if (std::memcmp(&cd.t, &hit.t, 4) == 0) {
              VertKey vk;
              std::memcpy(&vk.w[0], cd.v0, 12);
              std::memcpy(&vk.w[3], cd.v1, 12);
              std::memcpy(&vk.w[6], cd.v2, 12);
              auto it = s->primOfVerts.find(vk);
              if (it != s->primOfVerts.end())
                prim = it->second;
              break;
            }
          }
        }
        tOut[o] = hit.hitten ? hit.t : std::numeric_limits<float>::infinity();
        if (normalOut) {
          normalOut[3 * o + 0] = hit.normal.x;
          normalOut[3 * o + 1] = hit.normal.y;
          normalOut[3 * o + 2] = hit.normal.z;
        }
        if (primOut)
          primOut[o] = prim;
      }
    }
    ref_standin_set_recorder(nullptr, nullptr);
  }
}

// ---------------------------------------------------------------- single-ray / primitive KATs
int ref_intersect(const RefScene *s, const float o[3], const float d[3], float tNear, float tFar,
                  float *t, float normal[3]) {
  HitInfo h = s->scene->intersect(float3{o[0], o[1], o[2]}, float3{d[0], d[1], d[2]}, tNear, tFar);
  *t = h.t;
  normal[0] = h.normal.x;
  normal[1] = h.normal.y;
  normal[2] = h.normal.z;
  return h.hitten ? 1 : 0;
}

float ref_grid_sdf(const RefScene *s, const float p[3]) {
  return static_cast<const SDFGrid *>(s->scene.get())->sdf(float3{p[0], p[1], p[2]});
}

void ref_grid_normal(const RefScene *s, const float p[3], float out[3]) {
  float3 n = static_cast<const SDFGrid *>(s->scene.get())->normal(float3{p[0], p[1], p[2]});
  out[0] = n.x;
  out[1] = n.y;
  out[2] = n.z;
}

void ref_bbox_intersection(const float bmin[3], const float bmax[3], const float o[3],
                           const float invDir[3], float tMin, float tMax, float *t1, float *t2,
                           int *face) {
  BBox3f b;
  b.boxMin = float3{bmin[0], bmin[1], bmin[2]};
  b.boxMax = float3{bmax[0], bmax[1], bmax[2]};
  auto r = b.Intersection(float3{o[0], o[1], o[2]}, float3{invDir[0], invDir[1], invDir[2]}, tMin, tMax);
  *t1 = r.t1;
  *t2 = r.t2;
  *face = r.face;
}

// ray generation of raytracing.cpp:82-88 for loop coordinates (x, y)
void ref_eye_ray(uint32_t x, uint32_t y, uint32_t W, uint32_t H, const float projInv[16],
                 const float viewInv[16], float out[3]) {
  float4 d4 = EyeRayDir4f(static_cast<float>(x) + 0.5f, static_cast<float>(y) + 0.5f,
                          static_cast<float>(W), static_cast<float>(H), toMat(projInv));
  d4.w = 0.0f;
  d4 = toMat(viewInv) * d4;
  out[0] = d4.x;
  out[1] = d4.y;
  out[2] = d4.z;
}

uint32_t ref_color_pack(const float c[4]) { return color_pack_rgba(float4{c[0], c[1], c[2], c[3]}); }

void ref_sort8(float t[8], uint64_t children[8]) {
  size_t ch[8];
  for (int i = 0; i < 8; ++i)
    ch[i] = children[i];
  sort8(t, ch);
  for (int i = 0; i < 8; ++i)
    children[i] = ch[i];
}

void ref_divide_box_8(const float box[6], float out48[48]) {
  ispc::divide_box_8(reinterpret_cast<const ispc::Box *>(box), reinterpret_cast<ispc::Box8 *>(out48));
}

void ref_intersect_box_8(const float boxes48[48], const float o[3], const float invDir[3],
                         float tNear, float tFar, float out[8]) {
  ispc::intersect_box_8(reinterpret_cast<const ispc::Box8 *>(boxes48), o, invDir, tNear, tFar, out);
}

void ref_intersect_1_ray_8_triangles(const float tris72[72], const float o[3], const float d[3],
                                     uint32_t hitten[8], float t[8], float normals24[24]) {
  ispc::HitInfo8 h;
  ispc::intersect_1_ray_8_triangles(reinterpret_cast<const ispc::Triangle8 *>(tris72), o, d, &h);
  for (int i = 0; i < 8; ++i) {
    hitten[i] = h.hitten[i];
    t[i] = h.t[i];
    normals24[i] = h.norm_x[i];
    normals24[8 + i] = h.norm_y[i];
    normals24[16 + i] = h.norm_z[i];
  }
}

float ref_render_rows(const RefScene *s, const RefCam *c, uint32_t W, uint32_t H,
                      const float projInv[16], uint32_t y0, uint32_t y1, uint32_t *color, float *t) {
  // Same per-pixel body as Renderer::draw in Normal mode would need the private
  // intersectionColor; instead draw a (y1-y0)-row frame is not equivalent (different rays), so
  // this entry point drives IScene::intersect with the reference's own ray generation and
  // applies Normal-mode shading/packing through the public LiteMath helpers.
  Camera cam = makeCamera(c);
  float3 rayPos = cam.position();
  auto viewInv = inverse4x4(cam.lookAtMatrix());
  auto pi = toMat(projInv);
  const int width = static_cast<int>(W), height = static_cast<int>(H);
  double b = omp_get_wtime();
#pragma omp parallel for schedule(dynamic)
  for (int y = static_cast<int>(y0); y < static_cast<int>(y1); ++y) {
    for (int x = 0; x < width; ++x) {
      size_t o = static_cast<size_t>(height - y - 1) * W + x;
      float4 d4 = EyeRayDir4f(static_cast<float>(x) + 0.5f, static_cast<float>(y) + 0.5f,
                              static_cast<float>(width), static_cast<float>(height), pi);
      d4.w = 0.0f;
      d4 = viewInv * d4;
      float3 rayDir = to_float3(d4);
      HitInfo hit = s->scene->intersect(rayPos, rayDir, 0.01f, std::min(100.0f, t[o]));
      if (!hit.hitten)
        continue;
      if (dot(hit.normal, rayDir) > 0)
        hit.normal *= -1.0f;
      float4 col = to_float4(hit.normal, 1.0f);
      col = (col + 1.0f) / 2.0f;
      if (!std::isinf(hit.t)) {
        t[o] = hit.t;
        color[o] = color_pack_rgba(col);
      }
    }
  }
  return static_cast<float>((omp_get_wtime() - b) * 1e3);
}

} // extern "C"
