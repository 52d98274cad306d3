// TEST INFRASTRUCTURE, not a product path and not a CPU fallback: nothing in the package imports, links or runs this.
//
// The per-ray CUDA source of the product (csrc/rt_device.cuh + csrc/rt_shade.cuh: the text nvcc compiles into the
// render kernels) compiled a second time as plain C++ with `g++ -ffp-contract=off`, so that tests/test_host_emul.py can
// run the SAME functions on the CPU against the oracle, on many more cameras than a GPU box gets time for, and so that a
// change to the arithmetic can be checked for bit-exactness before GPU time is spent on it.  With -fmad=false on the
// device and -ffp-contract=off here every operator is one correctly rounded binary32 operation on both sides.
//
// What is NOT the product's text: the handful of intrinsic shims below, the 30-line pixel loop (the kernels' launch glue:
// tiles, bands, shared memory, TMA stores have no CPU counterpart) and the octree link classification, which restates
// k_oct_split of csrc/ingest.cu (isEmpty octree_raytracing.hpp:12-17, all >= 1e-4 octree_raytracing.cpp:132-135).
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "bvh_builder.h"
#include "rt_scene.h"

// ---- the CUDA intrinsics and keywords the per-ray code uses -----------------------------------------------------------------
template <class T>
static inline T __ldg(const T *p) {
  return *p;
}
static inline uint32_t __float_as_uint(float f) {
  uint32_t u;
  std::memcpy(&u, &f, 4);
  return u;
}
static inline float __uint_as_float(uint32_t u) {
  float f;
  std::memcpy(&f, &u, 4);
  return f;
}
static inline float __int_as_float(int i) {
  float f;
  std::memcpy(&f, &i, 4);
  return f;
}

#ifndef __noinline__
#define __noinline__ __attribute__((noinline))
#endif

namespace rt {
namespace emul {
using std::isfinite;
using std::max;
using std::min;
#include "rt_device.cuh"
#include "rt_shade.cuh"

// shade_pixel of csrc/rt_kernels.cuh without the launch state: memory row `row` (0 = top), column x
template <int KIND, bool SECONDARY, bool COUNT>
static void pixel(const SceneDev &s, const FrameDev &f, const float4 *tris, uint32_t x, uint32_t row, uint32_t *color,
                  float *t, uint32_t *prim, Counters &cnt) {
  const size_t px = (size_t)row * f.W + x;
  const int y = (int)(f.H - 1u - row);
  const V3 o = mk3(f.ray_pos[0], f.ray_pos[1], f.ray_pos[2]);
  const V3 d = eye_ray(f, (int)x, y);
  const float tPrev = f.clear ? __int_as_float(0x7f800000) : t[px];
  C4 col;
  float tNew;
  uint32_t pr;
  if (intersection_color<KIND, COUNT, SECONDARY>(s, f, tris, o, d, tPrev, col, tNew, pr, cnt) && !std::isinf(tNew)) {
    t[px] = tNew;
    color[px] = color_pack_rgba(col);
    if (prim) prim[px] = pr;
  } else if (f.clear) {
    t[px] = __int_as_float(0x7f800000);
    color[px] = 0u;
    if (prim) prim[px] = RT_PRIM_MISS;
  }
}

// count = the kernels' COUNT flavour (workload counters; some paths, e.g. the octree's culled-child handling, differ)
template <int KIND>
static void frame(const SceneDev &s, const FrameDev &f, const float4 *tris, uint32_t *color, float *t, uint32_t *prim,
                  Counters &cnt, bool count) {
  const bool secondary = f.mode == RT_SHADE_LAMBERT && (f.shadows || f.reflections);
  for (uint32_t row = 0; row < f.H; ++row)
    for (uint32_t x = 0; x < f.W; ++x) {
      if (secondary) {
        if (count) pixel<KIND, true, true>(s, f, tris, x, row, color, t, prim, cnt);
        else pixel<KIND, true, false>(s, f, tris, x, row, color, t, prim, cnt);
      } else {
        if (count) pixel<KIND, false, true>(s, f, tris, x, row, color, t, prim, cnt);
        else pixel<KIND, false, false>(s, f, tris, x, row, color, t, prim, cnt);
      }
    }
}

}  // namespace emul
}  // namespace rt

using namespace rt;

struct EmulScene {
  SceneDev dev{};
  std::vector<float> grid;
  std::vector<uint32_t> link;
  std::vector<float4> values;
  HostBvh bvh;
  std::string err;
};

extern "C" {

EmulScene *emul_grid(const float *values, uint32_t sx, uint32_t sy, uint32_t sz) {
  auto *e = new EmulScene;
  e->grid.assign(values, values + (size_t)sx * sy * sz);
  e->grid.resize(e->grid.size() + (size_t)sy * sz + sz + 1, 0.0f);  // the zero padding rt_scene_create_grid uploads
  GridDev &g = e->dev.grid;
  e->dev.kind = RT_SCENE_GRID;
  g.values = e->grid.data();
  g.sx = sx; g.sy = sy; g.sz = sz;
  g.bricks = 0; g.stride_x = sy * sz; g.stride_y = sz;
  g.fx = (float)(sx - 1); g.fy = (float)(sy - 1); g.fz = (float)(sz - 1);
  g.hfx = 0.5f * g.fx; g.hfy = 0.5f * g.fy; g.hfz = 0.5f * g.fz;  // as rt_scene_create_grid (csrc/rt_api.cu)
  return e;
}

// the 8x8x8-sample brick layout (RT_GRID_BRICKS): same rule as k_grid_bricks of csrc/ingest.cu
EmulScene *emul_grid_bricks(const float *values, uint32_t sx, uint32_t sy, uint32_t sz) {
  EmulScene *e = emul_grid(values, sx, sy, sz);
  const uint32_t nbx = (sx + 7) / 8, nby = (sy + 7) / 8, nbz = (sz + 7) / 8;
  std::vector<float> b((size_t)nbx * nby * nbz * 512, 0.0f);
  for (uint32_t x = 0; x < sx; ++x)
    for (uint32_t y = 0; y < sy; ++y)
      for (uint32_t z = 0; z < sz; ++z) {
        const size_t brick = ((size_t)(x >> 3) * nby + (y >> 3)) * nbz + (z >> 3);
        b[brick * 512 + ((x & 7u) << 6 | (y & 7u) << 3 | (z & 7u))] = values[((size_t)x * sy + y) * sz + z];
      }
  e->grid.swap(b);
  GridDev &g = e->dev.grid;
  g.values = e->grid.data();
  g.bricks = 1; g.stride_x = nby * nbz * 512u; g.stride_y = nbz * 512u;  // as rt_scene_create_grid
  return e;
}

EmulScene *emul_octree(const void *nodes36, uint32_t n) {
  auto *e = new EmulScene;
  e->link.resize(n);
  e->values.resize(2 * (size_t)n);
  const unsigned char *raw = static_cast<const unsigned char *>(nodes36);
  for (uint32_t i = 0; i < n; ++i) {
    float v[8];
    uint32_t off;
    std::memcpy(v, raw + (size_t)i * 36, 32);
    std::memcpy(&off, raw + (size_t)i * 36 + 32, 4);
    e->values[2 * (size_t)i] = make_float4(v[0], v[1], v[2], v[3]);
    e->values[2 * (size_t)i + 1] = make_float4(v[4], v[5], v[6], v[7]);
    uint32_t l = off;
    if (off == 0u) {
      bool allBig = true, allZero = true, allPos = true;
      for (int k = 0; k < 8; ++k) {
        if (!(v[k] > 10.0f)) allBig = false;
        if (!(v[k] == 0.0f)) allZero = false;
        if (!(v[k] >= 1e-4f)) allPos = false;
      }
      l = (allBig || allZero || allPos) ? OCT_LEAF_CULLED : OCT_LEAF_SURFACE;
    }
    e->link[i] = l;
  }
  // k_oct_childmask of csrc/ingest.cu: inner link words learn which children are worth a visit
  const bool masks = n < OCT_MASK_MAX_NODES;
  if (masks) {
    std::vector<uint32_t> masked(e->link);
    for (uint32_t i = 0; i < n; ++i) {
      uint32_t off;
      std::memcpy(&off, raw + (size_t)i * 36 + 32, 4);
      if (off == 0u || (unsigned long long)off + 7ull >= (unsigned long long)n) continue;
      uint32_t mask = 0;
      for (uint32_t c = 0; c < 8u; ++c)
        if (e->link[off + c] != OCT_LEAF_CULLED) mask |= 1u << c;
      masked[i] = off | (mask << 24);
    }
    e->link.swap(masked);
  }
  e->dev.kind = RT_SCENE_OCTREE;
  OctreeDev &o = e->dev.oct;
  o.has_mask = masks ? 1u : 0u;
  o.link = e->link.data();
  o.values = e->values.data();
  o.n_nodes = n;
  o.depth = 0;
  std::memcpy(&o.root_inner, raw + 32, 4);
  o.root_inner = o.root_inner != 0u ? 1u : 0u;
  return e;
}

// brute = BRUTE_NONE / BRUTE_BOXED / BRUTE_ROOT_LEAF (rt_scene.h): the product's host-built BVH8 or the no-tree paths
EmulScene *emul_mesh(const float *vpos4, uint32_t n_verts, const uint32_t *indices, uint32_t n_indices, int brute) {
  auto *e = new EmulScene;
  if (!build_bvh8(vpos4, n_verts, indices, n_indices, (uint32_t)brute, e->bvh, e->err)) {
    delete e;
    return nullptr;
  }
  e->dev.kind = RT_SCENE_MESH;
  MeshDev &m = e->dev.mesh;
  m.nodes = e->bvh.nodes.data();
  m.tris = e->bvh.tris.data();
  m.n_nodes = (uint32_t)e->bvh.nodes.size();
  m.n_tris = (uint32_t)e->bvh.tris.size();
  m.brute = e->bvh.brute;
  m.literal_order = 0u;
  return e;
}

// a tree the reference built (rt_scene_create_mesh_bvh8): literal slab test + sort8 visit order
EmulScene *emul_mesh_refbvh(const float *vpos4, uint32_t n_verts, const uint32_t *indices_bvh, uint32_t n_indices,
                            const void *nodes, uint32_t n_nodes, uint32_t node_bytes, const uint32_t *prim_of_permuted) {
  auto *e = new EmulScene;
  if (!import_reference_bvh(vpos4, n_verts, indices_bvh, n_indices, nodes, n_nodes, node_bytes, prim_of_permuted, e->bvh,
                            e->err)) {
    delete e;
    return nullptr;
  }
  e->dev.kind = RT_SCENE_MESH;
  MeshDev &m = e->dev.mesh;
  m.nodes = e->bvh.nodes.data();
  m.tris = e->bvh.tris.data();
  m.n_nodes = (uint32_t)e->bvh.nodes.size();
  m.n_tris = (uint32_t)e->bvh.tris.size();
  m.brute = e->bvh.brute;  // a one-leaf reference tree becomes the BRUTE_ROOT_LEAF scene
  m.literal_order = 1u;
  return e;
}

void emul_free(EmulScene *e) { delete e; }

// IScene::intersect for a batch of rays: the body of rays_kernel (csrc/rt_kernels.cuh)
int emul_intersect(const EmulScene *e, const float *origins, const float *dirs, uint32_t n, float t_near, float t_far,
                   int has_plane, float plane_y, uint8_t *hit, float *t, float *normal, uint32_t *prim) {
  if (!e || !origins || !dirs) return -1;
  const SceneDev &s = e->dev;
  const float4 *tris = reinterpret_cast<const float4 *>(s.mesh.tris);
  for (uint32_t i = 0; i < n; ++i) {
    emul::Counters cnt = {};
    const emul::V3 o = emul::mk3(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]);
    const emul::V3 d = emul::mk3(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]);
    emul::Hit h;
    const bool pl = has_plane != 0;
    if (s.kind == RT_SCENE_GRID && s.grid.bricks)
      h = emul::scene_intersect<emul::KIND_GRID_BRICKS, false>(s, nullptr, pl, plane_y, o, d, t_near, t_far, cnt);
    else if (s.kind == RT_SCENE_GRID)
      h = emul::scene_intersect<emul::KIND_GRID, false>(s, nullptr, pl, plane_y, o, d, t_near, t_far, cnt);
    else if (s.kind == RT_SCENE_OCTREE)
      h = emul::scene_intersect<emul::KIND_OCTREE, false>(s, nullptr, pl, plane_y, o, d, t_near, t_far, cnt);
    else if (s.mesh.brute)
      h = emul::scene_intersect<emul::KIND_BRUTE, false>(s, tris, pl, plane_y, o, d, t_near, t_far, cnt);
    else
      h = emul::scene_intersect<emul::KIND_BVH, false>(s, nullptr, pl, plane_y, o, d, t_near, t_far, cnt);
    if (hit) hit[i] = h.hitten ? 1 : 0;
    if (t) t[i] = h.t;
    if (normal) {
      normal[3 * i] = h.normal.x;
      normal[3 * i + 1] = h.normal.y;
      normal[3 * i + 2] = h.normal.z;
    }
    if (prim) prim[i] = h.prim;
  }
  return 0;
}

// The resumable octree walk (OctWalk: what render_octree_compact_kernel drives lane by lane) run to completion for one
// ray at a time: begin, then step / leaf until a leaf hits or the walk is done.  Octree scenes with an inner root only.
int emul_intersect_walk(const EmulScene *e, const float *origins, const float *dirs, uint32_t n, float t_near,
                        float t_far, uint8_t *hit, float *t, float *normal, unsigned long long *counters3) {
  if (!e || e->dev.kind != RT_SCENE_OCTREE || !e->dev.oct.root_inner) return -1;
  const OctreeDev &oc = e->dev.oct;
  emul::Counters cnt = {};
  for (uint32_t i = 0; i < n; ++i) {
    const emul::V3 o = emul::mk3(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]);
    const emul::V3 d = emul::mk3(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]);
    emul::OctWalk w;
    emul::Hit h = emul::hit_none();
    emul::oct_walk_begin<true>(w, oc, o, d, t_near, t_far, cnt);
    for (;;) {
      const int r = emul::oct_walk_step<true>(w, oc, o, t_near, cnt);
      if (r == emul::OCT_WALK_DONE) break;
      if (r == emul::OCT_WALK_LEAF && emul::oct_walk_leaf<true>(w, oc, o, t_near, h, cnt)) break;
    }
    hit[i] = h.hitten ? 1 : 0;
    t[i] = h.t;
    normal[3 * i] = h.normal.x;
    normal[3 * i + 1] = h.normal.y;
    normal[3 * i + 2] = h.normal.z;
  }
  if (counters3) {
    counters3[0] = cnt.oct_inner;
    counters3[1] = cnt.oct_leaf;
    counters3[2] = cnt.oct_leaf_samples;
  }
  return 0;
}

// the same rays through oct_intersect, with the same three counters
int emul_intersect_octree_counted(const EmulScene *e, const float *origins, const float *dirs, uint32_t n, float t_near,
                                  float t_far, uint8_t *hit, float *t, float *normal, unsigned long long *counters3) {
  if (!e || e->dev.kind != RT_SCENE_OCTREE) return -1;
  emul::Counters cnt = {};
  for (uint32_t i = 0; i < n; ++i) {
    const emul::V3 o = emul::mk3(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]);
    const emul::V3 d = emul::mk3(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]);
    emul::Hit h = emul::oct_intersect<true>(e->dev.oct, o, d, t_near, t_far, cnt);
    hit[i] = h.hitten ? 1 : 0;
    t[i] = h.t;
    normal[3 * i] = h.normal.x;
    normal[3 * i + 1] = h.normal.y;
    normal[3 * i + 2] = h.normal.z;
  }
  if (counters3) {
    counters3[0] = cnt.oct_inner;
    counters3[1] = cnt.oct_leaf;
    counters3[2] = cnt.oct_leaf_samples;
  }
  return 0;
}

// Per-pixel work of a primary-ray frame (analysis of SIMD efficiency under different pixel -> warp mappings,
// tools/warp_efficiency.py): work[px * 4 + {0,1,2,3}] = inner-node visits, leaf visits, SDF samples, triangle tests.
int emul_work_map(const EmulScene *e, const rt_frame *fr, uint32_t *work) {
  if (!e || !fr || !work) return -1;
  rt_frame one = *fr;
  const SceneDev &s = e->dev;
  const float4 *tris = reinterpret_cast<const float4 *>(s.mesh.tris);
  FrameDev f{};
  std::memcpy(f.ray_pos, one.ray_pos, sizeof f.ray_pos);
  std::memcpy(f.view_inv, one.view_inv, sizeof f.view_inv);
  std::memcpy(f.proj_inv, one.proj_inv, sizeof f.proj_inv);
  f.W = one.width; f.H = one.height;
  f.t_near = 0.01f; f.t_far = 100.0f;
  for (uint32_t row = 0; row < f.H; ++row)
    for (uint32_t x = 0; x < f.W; ++x) {
      emul::Counters c = {};
      const emul::V3 o = emul::mk3(f.ray_pos[0], f.ray_pos[1], f.ray_pos[2]);
      const emul::V3 d = emul::eye_ray(f, (int)x, (int)(f.H - 1u - row));
      if (s.kind == RT_SCENE_GRID) emul::scene_intersect<emul::KIND_GRID, true>(s, nullptr, false, 0.f, o, d, f.t_near, f.t_far, c);
      else if (s.kind == RT_SCENE_OCTREE) emul::scene_intersect<emul::KIND_OCTREE, true>(s, nullptr, false, 0.f, o, d, f.t_near, f.t_far, c);
      else if (s.mesh.brute) emul::scene_intersect<emul::KIND_BRUTE, true>(s, tris, false, 0.f, o, d, f.t_near, f.t_far, c);
      else emul::scene_intersect<emul::KIND_BVH, true>(s, nullptr, false, 0.f, o, d, f.t_near, f.t_far, c);
      uint32_t *w = work + ((size_t)row * f.W + x) * 4;
      w[0] = c.oct_inner + c.bvh_inner;
      w[1] = c.oct_leaf + c.bvh_leaf;
      w[2] = c.grid_samples + c.oct_leaf_samples;
      w[3] = c.bvh_tris;
    }
  return 0;
}

// Renderer::draw over the whole frame.  frame = the rt_frame of include/rt_b200.h (band fields ignored).
// counters: 9 x u64 in rt_counters order, or null.
int emul_render(const EmulScene *e, const rt_frame *fr, uint32_t *color, float *t, uint32_t *prim,
                unsigned long long *counters) {
  if (!e || !fr || !color || !t) return -1;
  FrameDev f{};
  std::memcpy(f.ray_pos, fr->ray_pos, sizeof f.ray_pos);
  std::memcpy(f.view_inv, fr->view_inv, sizeof f.view_inv);
  std::memcpy(f.proj_inv, fr->proj_inv, sizeof f.proj_inv);
  f.W = fr->width;
  f.H = fr->height;
  f.band_rows = f.H; f.band_count = 1; f.band_index = 0; f.rows_here = f.H; f.row0 = 0;
  f.mode = fr->shading_mode;
  std::memcpy(f.light, fr->light_pos, sizeof f.light);
  f.shadows = fr->enable_shadows != 0;
  f.reflections = fr->enable_reflections != 0;
  f.has_plane = fr->has_plane != 0;
  f.plane_y = fr->plane_y;
  const bool defaults = fr->t_near == 0.0f && fr->t_far == 0.0f;  // as make_frame (csrc/rt_api.cu)
  f.t_near = defaults ? 0.01f : fr->t_near;
  f.t_far = defaults ? 100.0f : fr->t_far;
  f.clear = fr->clear != 0;
  emul::Counters cnt = {};
  const SceneDev &s = e->dev;
  const float4 *tris = reinterpret_cast<const float4 *>(s.mesh.tris);
  if (s.kind == RT_SCENE_GRID && s.grid.bricks) emul::frame<emul::KIND_GRID_BRICKS>(s, f, nullptr, color, t, prim, cnt, counters != nullptr);
  else if (s.kind == RT_SCENE_GRID) emul::frame<emul::KIND_GRID>(s, f, nullptr, color, t, prim, cnt, counters != nullptr);
  else if (s.kind == RT_SCENE_OCTREE) emul::frame<emul::KIND_OCTREE>(s, f, nullptr, color, t, prim, cnt, counters != nullptr);
  else if (s.mesh.brute) emul::frame<emul::KIND_BRUTE>(s, f, tris, color, t, prim, cnt, counters != nullptr);
  else emul::frame<emul::KIND_BVH>(s, f, nullptr, color, t, prim, cnt, counters != nullptr);
  if (counters) {
    const uint32_t v[9] = {cnt.rays, cnt.hits, cnt.grid_samples, cnt.oct_inner, cnt.oct_leaf, cnt.oct_leaf_samples,
                           cnt.bvh_inner, cnt.bvh_leaf, cnt.bvh_tris};
    for (int i = 0; i < 9; ++i) counters[i] = v[i];
  }
  return 0;
}

}  // extern "C"
