// extern "C" boundary of the B200 primary-ray path (include/rt_b200.h): scene upload, validation, launches.
// No CPU fallback lives here: every entry point needs a CUDA device.
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <chrono>
#include <limits>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "bvh_builder.h"
#include "ingest.h"
#include "rt_scene.h"

using namespace rt;

namespace {

thread_local std::string g_err;
std::atomic<int> g_math{RT_MATH_EXACT};
std::atomic<int> g_sched{RT_SCHED_AUTO};
// Device pointers obtained through rt_ipc_open: another GPU's memory mapped into this process.  (The runtime reports
// the mapping device, not the owner, for such pointers, so the library remembers them itself.)
std::mutex g_peer_mu;
std::vector<const void *> g_peer_ptrs;
std::atomic<uint64_t> g_staged_launches{0};
bool is_peer_pointer(const void *p) {
  std::lock_guard<std::mutex> lock(g_peer_mu);
  for (const void *q : g_peer_ptrs)
    if (q == p) return true;
  return false;
}

int fail(int code, const std::string &msg) {
  g_err = msg;
  return code;
}
int cuda_fail(cudaError_t e, const char *what) {
  g_err = std::string(what) + ": " + cudaGetErrorString(e);
  cudaGetLastError();  // a non-sticky error must not be reported again by the next launch's cudaGetLastError()
  return RT_ERR_CUDA;
}
#define RT_CUDA(call)                                  \
  do {                                                 \
    cudaError_t e__ = (call);                          \
    if (e__ != cudaSuccess) return cuda_fail(e__, #call); \
  } while (0)

struct DeviceGuard {
  int prev = -1;
  bool ok = true;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
    if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

constexpr int N_TILE_COUNTERS = 64;

}  // namespace

// The scene on ONE device.  A scene created after rt_init(devs, n) has one replica per listed device (scene data is
// read-only and small next to 180 GB of HBM: replicated, never sharded); otherwise a single replica on the device
// that was current at creation.
struct Replica {
  int device = 0;
  int sm_count = 148;
  SceneDev dev{};
  std::vector<std::pair<void *, size_t>> allocs;  // device allocations and their sizes (cloned onto the other devices)
  uint32_t *tile_counters = nullptr;              // ring, one per in-flight persistent launch
  std::atomic<uint32_t> next_counter{0};
  unsigned long long *xfer = nullptr;             // running {fetched, stored} bytes of the in/out staged tiles
  // per-device state of rt_render_host / of a fanned-out rt_render_device
  uint32_t *h_color = nullptr;  // device framebuffer of the copy pipeline (allocated on first use)
  float *h_t = nullptr;
  uint32_t *h_prim = nullptr;
  rt_counters *h_cnt = nullptr;
  size_t h_pixels = 0;
  cudaStream_t h_stream = nullptr, h_up = nullptr, h_down = nullptr;  // kernels, H2D copies, D2H copies
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t ev_up[16] = {}, ev_k[16] = {};  // per row chunk: upload done, kernel done
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;  // fan-out of one frame over the replicas (no timing)
  // heavy-first block scheduling (HeavySched, rt_scene.h): what the previous launch on this replica learned
  std::mutex heavy_mu;
  uint32_t *heavy_mem = nullptr;  // [2 x HEAVY_CAP lists | 2 counts | done[tiles] | seen[tiles]]
  size_t heavy_tiles = 0;
  uint64_t heavy_sig = 0;         // frame geometry + kernel shape the lists were built for
  uint32_t heavy_parity = 0, heavy_tag = 0;
};

struct rt_scene {
  int kind = -1;
  std::vector<std::unique_ptr<Replica>> reps;  // reps[0] is where the scene was built
  uint64_t bytes = 0;
  float bbox_min[3] = {-1, -1, -1}, bbox_max[3] = {1, 1, 1};
  float build_ms = 0.0f;  // device time of an on-device BVH build / octree ingestion
  std::mutex host_mu;     // rt_render_host is serialised per scene (it owns the replicas' streams and buffers)
};

namespace {

// rt_init's device list (empty: scenes live on the device current at their creation)
std::mutex g_dev_mu;
std::vector<int> g_devs;

int scene_begin(rt_scene **out, rt_scene *&s) {
  if (!out) return fail(RT_ERR_INVALID, "out pointer is null");
  *out = nullptr;
  int dev = 0;
  {
    std::lock_guard<std::mutex> lock(g_dev_mu);
    if (!g_devs.empty()) dev = g_devs[0];
    else RT_CUDA(cudaGetDevice(&dev));
  }
  cudaDeviceProp p;
  cudaError_t e = cudaGetDeviceProperties(&p, dev);
  if (e != cudaSuccess) return cuda_fail(e, "cudaGetDeviceProperties");
  s = new rt_scene;
  s->reps.emplace_back(new Replica);
  s->reps[0]->device = dev;
  s->reps[0]->sm_count = p.multiProcessorCount;
  return RT_OK;
}

// cudaMalloc + copy on the CURRENT device (the creating entry point holds a DeviceGuard on reps[0]->device)
template <class T>
int upload(rt_scene *s, const T *host, size_t n, const T **devOut) {
  void *d = nullptr;
  size_t bytes = std::max<size_t>(n * sizeof(T), 16);
  RT_CUDA(cudaMalloc(&d, bytes));
  s->reps[0]->allocs.emplace_back(d, bytes);
  s->bytes += n * sizeof(T);
  if (n) RT_CUDA(cudaMemcpy(d, host, n * sizeof(T), cudaMemcpyHostToDevice));
  *devOut = static_cast<const T *>(d);
  return RT_OK;
}

void adopt(rt_scene *s, void *d, size_t bytes) { s->reps[0]->allocs.emplace_back(d, std::max<size_t>(bytes, 16)); }

int finish_replica(Replica &r) {
  void *d = nullptr;
  const size_t bytes = N_TILE_COUNTERS * sizeof(uint32_t) + 2 * sizeof(unsigned long long);
  RT_CUDA(cudaMalloc(&d, bytes));
  r.allocs.emplace_back(d, bytes);
  r.tile_counters = static_cast<uint32_t *>(d);
  r.xfer = reinterpret_cast<unsigned long long *>(r.tile_counters + N_TILE_COUNTERS);
  RT_CUDA(cudaMemset(d, 0, bytes));
  RT_CUDA(cudaDeviceSynchronize());
  return RT_OK;
}

// the device copy of `src` (an allocation of `from`) on the current device
template <class T>
int clone_ptr(const Replica &from, Replica &to, const T *&field) {
  if (!field) return RT_OK;
  for (const auto &a : from.allocs)
    if (a.first == static_cast<const void *>(field)) {
      void *d = nullptr;
      RT_CUDA(cudaMalloc(&d, a.second));
      to.allocs.emplace_back(d, a.second);
      RT_CUDA(cudaMemcpyPeer(d, to.device, a.first, from.device, a.second));
      field = static_cast<const T *>(d);
      return RT_OK;
    }
  return fail(RT_ERR_CUDA, "internal: scene pointer without an allocation record");
}

// Builds the remaining replicas (rt_init's devices 1..n-1) from reps[0] with device-to-device copies: the host data is
// parsed, validated and laid out once, whatever built it (host BVH, device LBVH, device octree ingestion).
int scene_finish(rt_scene *s, rt_scene **out) {
  int rc = finish_replica(*s->reps[0]);
  if (rc) return rc;
  std::vector<int> devs;
  {
    std::lock_guard<std::mutex> lock(g_dev_mu);
    devs = g_devs;
  }
  for (size_t j = 1; j < devs.size(); ++j) {
    DeviceGuard g(devs[j]);
    if (!g.ok) return fail(RT_ERR_CUDA, "cannot select a device of rt_init's list");
    cudaDeviceProp p;
    RT_CUDA(cudaGetDeviceProperties(&p, devs[j]));
    s->reps.emplace_back(new Replica);
    Replica &r = *s->reps.back(), &r0 = *s->reps[0];
    r.device = devs[j];
    r.sm_count = p.multiProcessorCount;
    r.dev = r0.dev;
    if ((rc = clone_ptr(r0, r, r.dev.grid.values))) return rc;
    if ((rc = clone_ptr(r0, r, r.dev.oct.link))) return rc;
    if ((rc = clone_ptr(r0, r, r.dev.oct.values))) return rc;
    if ((rc = clone_ptr(r0, r, r.dev.mesh.nodes))) return rc;
    if ((rc = clone_ptr(r0, r, r.dev.mesh.tris))) return rc;
    if ((rc = finish_replica(r))) return rc;
  }
  *out = s;
  return RT_OK;
}

void replica_free(Replica &r) {
  DeviceGuard g(r.device);
  for (auto &a : r.allocs) cudaFree(a.first);
  if (r.h_color) cudaFree(r.h_color);
  if (r.h_t) cudaFree(r.h_t);
  if (r.h_prim) cudaFree(r.h_prim);
  if (r.h_cnt) cudaFree(r.h_cnt);
  for (auto &e : r.ev)
    if (e) cudaEventDestroy(e);
  for (auto &e : r.ev_up)
    if (e) cudaEventDestroy(e);
  for (auto &e : r.ev_k)
    if (e) cudaEventDestroy(e);
  if (r.ev_fork) cudaEventDestroy(r.ev_fork);
  if (r.ev_join) cudaEventDestroy(r.ev_join);
  if (r.heavy_mem) cudaFree(r.heavy_mem);
  if (r.h_stream) cudaStreamDestroy(r.h_stream);
  if (r.h_up) cudaStreamDestroy(r.h_up);
  if (r.h_down) cudaStreamDestroy(r.h_down);
}

void scene_free(rt_scene *s) {
  if (!s) return;
  for (auto &r : s->reps) replica_free(*r);
  delete s;
}

}  // namespace

extern "C" {

const char *rt_last_error(void) { return g_err.c_str(); }
const char *rt_version(void) { return "rt_b200 0.1 (sm_100a)"; }

int rt_device_count(int *count) {
  if (!count) return fail(RT_ERR_INVALID, "count is null");
  *count = 0;
  RT_CUDA(cudaGetDeviceCount(count));
  return RT_OK;
}

int rt_set_device(int device) {
  RT_CUDA(cudaSetDevice(device));
  return RT_OK;
}

int rt_set_math_mode(int mode) {
  if (mode != RT_MATH_EXACT && mode != RT_MATH_FAST) return fail(RT_ERR_INVALID, "unknown math mode");
  g_math = mode;
  return RT_OK;
}

int rt_set_sched_mode(int mode) {
  if (mode < RT_SCHED_AUTO || mode > RT_SCHED_COMPACT) return fail(RT_ERR_INVALID, "unknown sched mode");
  g_sched = mode;
  return RT_OK;
}

// ---------------------------------------------------------------------------------------------- grid
int rt_scene_create_grid(const float *values, uint32_t sx, uint32_t sy, uint32_t sz, int layout, rt_scene **out) {
  if (!values) return fail(RT_ERR_INVALID, "grid values pointer is null");
  // SDFGrid::sdf maps [-1,1] to [0,size-1] (grid_raytracing.cpp:8-11): every axis needs two samples
  if (sx < 2 || sy < 2 || sz < 2) return fail(RT_ERR_INVALID, "grid needs at least 2 samples per axis");
  // element offsets are 32-bit on the device, also for the brick-padded copy
  const uint64_t px = ((uint64_t)sx + 7) & ~7ull, py = ((uint64_t)sy + 7) & ~7ull, pz = ((uint64_t)sz + 7) & ~7ull;
  if (px >= (1ull << 32) || py >= (1ull << 32) || pz >= (1ull << 32) || px * py >= (1ull << 32) ||
      px * py * pz >= (1ull << 32))
    return fail(RT_ERR_INVALID, "grid too large (element offsets must fit 32 bits)");
  if (layout < RT_GRID_AUTO || layout > RT_GRID_BRICKS) return fail(RT_ERR_INVALID, "unknown grid layout");
  rt_scene *s = nullptr;
  int rc = scene_begin(out, s);
  if (rc) return rc;
  DeviceGuard dg(s->reps[0]->device);
  s->kind = RT_SCENE_GRID;
  s->reps[0]->dev.kind = RT_SCENE_GRID;
  GridDev &g = s->reps[0]->dev.grid;
  g.sx = sx; g.sy = sy; g.sz = sz;
  g.fx = (float)(sx - 1); g.fy = (float)(sy - 1); g.fz = (float)(sz - 1);
  g.hfx = 0.5f * g.fx; g.hfy = 0.5f * g.fy; g.hfz = 0.5f * g.fz;
  const size_t n = (size_t)sx * sy * sz;
  if (layout == RT_GRID_AUTO) layout = RT_GRID_LINEAR;  // chosen by measurement, DESIGN.md "grid layout"
  if (layout == RT_GRID_LINEAR) {
    // the linear-layout sampler multiplies a neighbouring sample by an exact 0 where the reference does not read it
    // (rt_device.cuh: grid_sdf): only right for finite samples, so a grid with infinities / NaNs takes the literal path
    bool finite = true;
    for (size_t i = 0; i < n && finite; ++i) finite = std::isfinite(values[i]);
    if (!finite) layout = RT_GRID_BRICKS;
  }
  if (layout == RT_GRID_BRICKS) {
    // the layout transform runs on the device (ingest.cu): upload the linear samples once, re-tile there
    const uint32_t nby = (sy + 7) / 8, nbz = (sz + 7) / 8;
    float *b = nullptr;
    size_t count = 0;
    std::string ierr;
    const int irc = ingest_grid_bricks_device(values, sx, sy, sz, &b, &count, ierr);
    if (irc != INGEST_OK) {
      scene_free(s);
      return fail(RT_ERR_CUDA, ierr);
    }
    adopt(s, b, count * sizeof(float));
    s->bytes += count * sizeof(float);
    g.values = b;
    g.bricks = 1; g.stride_x = nby * nbz * 512u; g.stride_y = nbz * 512u;
  } else {
    g.bricks = 0; g.stride_x = sy * sz; g.stride_y = sz;
    // readable zeros one slab + one row + one sample past the end (grid_sdf reads c0 + 1 on every axis)
    const size_t padded = n + (size_t)sy * sz + sz + 1;
    void *d = nullptr;
    cudaError_t e = cudaMalloc(&d, padded * sizeof(float));
    if (e == cudaSuccess) e = cudaMemset(static_cast<float *>(d) + n, 0, (padded - n) * sizeof(float));
    if (e == cudaSuccess) e = cudaMemcpy(d, values, n * sizeof(float), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
      if (d) cudaFree(d);
      scene_free(s);
      return cuda_fail(e, "grid upload");
    }
    adopt(s, d, padded * sizeof(float));
    s->bytes += n * sizeof(float);
    g.values = static_cast<const float *>(d);
  }
  if (rc == RT_OK) rc = scene_finish(s, out);
  if (rc != RT_OK) scene_free(s);
  return rc;
}

// -------------------------------------------------------------------------------------------- octree
int rt_scene_create_octree(const void *nodes36, uint32_t n_nodes, rt_scene **out) {
  if (!nodes36 || n_nodes == 0) return fail(RT_ERR_INVALID, "octree needs at least the root node");
  // Split, leaf classification and validation run on the device (ingest.cu): the raw records are uploaded as they are.
  rt_scene *s = nullptr;
  int rc = scene_begin(out, s);
  if (rc) return rc;
  DeviceGuard dg(s->reps[0]->device);
  uint32_t *link = nullptr, depth = 0, has_mask = 0;
  float4 *vals = nullptr;
  std::string ierr;
  const int irc = ingest_octree_device(nodes36, n_nodes, &link, &vals, &depth, &has_mask, &s->build_ms, ierr);
  if (irc != INGEST_OK) {
    scene_free(s);
    return fail(irc == INGEST_INVALID ? RT_ERR_INVALID : RT_ERR_CUDA, ierr);
  }
  adopt(s, link, (size_t)n_nodes * 4);
  adopt(s, vals, (size_t)n_nodes * 32);
  s->bytes += (uint64_t)n_nodes * 36;
  s->kind = RT_SCENE_OCTREE;
  OctreeDev &oc = s->reps[0]->dev.oct;
  s->reps[0]->dev.kind = RT_SCENE_OCTREE;
  oc.n_nodes = n_nodes;
  oc.depth = depth;
  oc.root_inner = depth > 0 ? 1u : 0u;
  oc.has_mask = has_mask;
  oc.link = link;
  oc.values = vals;
  rc = scene_finish(s, out);
  if (rc != RT_OK) scene_free(s);
  return rc;
}

// ---------------------------------------------------------------------------------------------- mesh
static int finish_mesh(HostBvh &h, bool literal_order, rt_scene **out) {
  rt_scene *s = nullptr;
  int rc = scene_begin(out, s);
  if (rc) return rc;
  DeviceGuard dg(s->reps[0]->device);
  s->kind = RT_SCENE_MESH;
  s->reps[0]->dev.kind = RT_SCENE_MESH;
  MeshDev &m = s->reps[0]->dev.mesh;
  m.n_tris = (uint32_t)h.tris.size();
  m.n_nodes = (uint32_t)h.nodes.size();
  m.brute = h.brute;
  m.literal_order = literal_order ? 1u : 0u;
  for (int a = 0; a < 3; ++a) {
    s->bbox_min[a] = h.bbox_min[a];
    s->bbox_max[a] = h.bbox_max[a];
  }
  rc = upload(s, h.tris.data(), h.tris.size(), &m.tris);
  if (rc == RT_OK) rc = upload(s, h.nodes.data(), h.nodes.size(), &m.nodes);
  if (rc == RT_OK) rc = scene_finish(s, out);
  if (rc != RT_OK) scene_free(s);
  return rc;
}

int rt_scene_create_mesh(const float *vpos4, uint32_t n_verts, const uint32_t *indices, uint32_t n_indices, int accel,
                         rt_scene **out) {
  if (accel < RT_MESH_AUTO || accel > RT_MESH_LBVH_GPU) return fail(RT_ERR_INVALID, "unknown mesh accel");
  const uint32_t nTri = n_indices / 3;
  if (accel == RT_MESH_LBVH_GPU && nTri > 32) {
    rt_scene *s = nullptr;
    int rc = scene_begin(out, s);
    if (rc) return rc;
    DeviceGuard dg(s->reps[0]->device);
    DeviceBvh d;
    std::string derr;
    if (!build_bvh8_device(vpos4, n_verts, indices, n_indices, d, derr)) {
      scene_free(s);
      return fail(derr.find("index out of range") != std::string::npos || derr.find("mesh needs") != std::string::npos
                      ? RT_ERR_INVALID
                      : RT_ERR_CUDA,
                  derr);
    }
    adopt(s, d.nodes, (size_t)d.n_nodes * sizeof(BvhNode));
    adopt(s, d.tris, (size_t)d.n_tris * sizeof(TriDev));
    s->bytes += (uint64_t)d.n_nodes * sizeof(BvhNode) + (uint64_t)d.n_tris * sizeof(TriDev);
    s->kind = RT_SCENE_MESH;
    s->reps[0]->dev.kind = RT_SCENE_MESH;
    MeshDev &m = s->reps[0]->dev.mesh;
    m.nodes = d.nodes;
    m.tris = d.tris;
    m.n_nodes = d.n_nodes;
    m.n_tris = d.n_tris;
    m.brute = 0u;
    m.literal_order = 0u;
    s->build_ms = d.build_ms;
    for (int a = 0; a < 3; ++a) {
      s->bbox_min[a] = d.bbox_min[a];
      s->bbox_max[a] = d.bbox_max[a];
    }
    rc = scene_finish(s, out);
    if (rc != RT_OK) scene_free(s);
    return rc;
  }
  if (accel == RT_MESH_LBVH_GPU) accel = RT_MESH_AUTO;  // tiny meshes: nothing to build
  // Up to 8 triangles the reference's own tree is a single leaf whose triangles are tested without any box or range
  // check (triangles_raytracing.cpp:175-186, 292-331): the brute-force kernel in BRUTE_ROOT_LEAF mode IS that, exactly,
  // in every view.  Beyond 8 triangles only a tree reproduces the reference's box culling, so RT_MESH_AUTO builds one;
  // RT_MESH_BRUTE stays available as an explicit choice (BRUTE_BOXED, rt_scene.h) up to 4096 triangles (192 KB of smem).
  uint32_t brute = BRUTE_NONE;
  if (nTri <= ROOT_LEAF_MAX_TRIS && accel != RT_MESH_BVH8) brute = BRUTE_ROOT_LEAF;
  else if (accel == RT_MESH_BRUTE) brute = BRUTE_BOXED;
  if (brute == BRUTE_BOXED && nTri > 4096) return fail(RT_ERR_UNSUPPORTED, "brute force is limited to 4096 triangles");
  HostBvh h;
  std::string err;
  if (!build_bvh8(vpos4, n_verts, indices, n_indices, brute, h, err)) return fail(RT_ERR_INVALID, err);
  return finish_mesh(h, false, out);
}

int rt_scene_create_mesh_bvh8(const float *vpos4, uint32_t n_verts, const uint32_t *indices_bvh, uint32_t n_indices,
                              const void *nodes, uint32_t n_nodes, uint32_t node_bytes,
                              const uint32_t *prim_of_permuted, rt_scene **out) {
  HostBvh h;
  std::string err;
  if (!import_reference_bvh(vpos4, n_verts, indices_bvh, n_indices, nodes, n_nodes, node_bytes, prim_of_permuted, h,
                            err))
    return fail(RT_ERR_INVALID, err);
  return finish_mesh(h, true, out);
}

void rt_scene_destroy(rt_scene *scene) { scene_free(scene); }

int rt_scene_build_ms(const rt_scene *scene, float *device_ms) {
  if (!scene || !device_ms) return fail(RT_ERR_INVALID, "null argument");
  *device_ms = scene->build_ms;
  return RT_OK;
}

int rt_scene_mesh_stats(const rt_scene *scene, uint32_t *n_nodes, uint32_t *n_tris, float *device_build_ms) {
  if (!scene || scene->kind != RT_SCENE_MESH) return fail(RT_ERR_INVALID, "not a mesh scene");
  if (n_nodes) *n_nodes = scene->reps[0]->dev.mesh.n_nodes;
  if (n_tris) *n_tris = scene->reps[0]->dev.mesh.n_tris;
  if (device_build_ms) *device_build_ms = scene->build_ms;
  return RT_OK;
}

int rt_scene_kind_of(const rt_scene *scene) { return scene ? scene->kind : RT_ERR_INVALID; }
int rt_scene_device(const rt_scene *scene) { return scene ? scene->reps[0]->device : RT_ERR_INVALID; }
int rt_scene_replicas(const rt_scene *scene, int *devices, int cap) {
  if (!scene) return RT_ERR_INVALID;
  for (size_t j = 0; j < scene->reps.size() && devices && (int)j < cap; ++j) devices[j] = scene->reps[j]->device;
  return (int)scene->reps.size();
}

int rt_scene_transfer_bytes(const rt_scene *scene, uint64_t *fetched, uint64_t *stored) {
  if (!scene) return fail(RT_ERR_INVALID, "null scene");
  uint64_t f = 0, st = 0;
  for (const auto &rep : scene->reps) {
    DeviceGuard dg(rep->device);
    unsigned long long v[2] = {0, 0};
    RT_CUDA(cudaMemcpy(v, rep->xfer, sizeof(v), cudaMemcpyDeviceToHost));
    f += v[0];
    st += v[1];
  }
  if (fetched) *fetched = f;
  if (stored) *stored = st;
  return RT_OK;
}

int rt_scene_info(const rt_scene *scene, uint64_t *device_bytes, float bbox_min[3], float bbox_max[3]) {
  if (!scene) return fail(RT_ERR_INVALID, "scene is null");
  if (device_bytes) *device_bytes = scene->bytes;
  for (int a = 0; a < 3; ++a) {
    if (bbox_min) bbox_min[a] = scene->bbox_min[a];
    if (bbox_max) bbox_max[a] = scene->bbox_max[a];
  }
  return RT_OK;
}

}  // extern "C"

// =============================================================================================================
// rendering
// =============================================================================================================
namespace {

int make_frame(const rt_frame *f, FrameDev &d) {
  if (!f) return fail(RT_ERR_INVALID, "frame is null");
  if (f->width == 0 || f->height == 0) return fail(RT_ERR_INVALID, "frame has zero size");
  if ((uint64_t)f->width * f->height >= (1ull << 31)) return fail(RT_ERR_INVALID, "frame too large");
  if (f->shading_mode < RT_SHADE_NORMAL || f->shading_mode > RT_SHADE_COLOR)
    return fail(RT_ERR_INVALID, "unknown shading mode");
  std::memcpy(d.ray_pos, f->ray_pos, sizeof d.ray_pos);
  std::memcpy(d.view_inv, f->view_inv, sizeof d.view_inv);
  std::memcpy(d.proj_inv, f->proj_inv, sizeof d.proj_inv);
  d.W = f->width;
  d.H = f->height;
  d.row0 = 0;
  if (f->band_count <= 1) {
    d.band_rows = f->height;
    d.band_count = 1;
    d.band_index = 0;
    d.rows_here = f->height;
  } else {
    if (f->band_rows == 0 || f->band_index >= f->band_count) return fail(RT_ERR_INVALID, "bad band partition");
    d.band_rows = f->band_rows;
    d.band_count = f->band_count;
    d.band_index = f->band_index;
    const uint32_t nBands = (f->height + f->band_rows - 1) / f->band_rows;
    const uint32_t mine = nBands > f->band_index ? (nBands - f->band_index + f->band_count - 1) / f->band_count : 0;
    d.rows_here = mine * f->band_rows;  // the last band may be short: kernels also test row < H
  }
  d.mode = f->shading_mode;
  std::memcpy(d.light, f->light_pos, sizeof d.light);
  d.shadows = f->enable_shadows != 0;
  d.reflections = f->enable_reflections != 0;
  d.has_plane = f->has_plane != 0;
  d.plane_y = f->plane_y;
  const bool defaults = f->t_near == 0.0f && f->t_far == 0.0f;
  d.t_near = defaults ? 0.01f : f->t_near;  // raytracing.cpp:90
  d.t_far = defaults ? 100.0f : f->t_far;
  d.clear = f->clear != 0;
  return RT_OK;
}

// tuning knobs for measurement runs (read once): RT_TILE_BLOCK = threads per block of the tiles kernel,
// RT_HOST_CHUNKS = row chunks of rt_render_host's copy/render pipeline.  Unset = the measured defaults.
int env_int(const char *name) {
  const char *v = std::getenv(name);
  return v ? std::atoi(v) : 0;
}
const int g_env_tile_block = env_int("RT_TILE_BLOCK");
const int g_env_host_chunks = env_int("RT_HOST_CHUNKS");
const int g_env_stage = std::getenv("RT_STAGE_STORES") ? env_int("RT_STAGE_STORES") : -1;  // -1: peer framebuffers only
// RT_HOST_ZEROCOPY: rt_render_host with a pinned host framebuffer; 0 = copy pipeline, 1 = kernel stores straight into the
// host buffers, 2 = the same through the staged 128-byte bulk stores, 3 = 2 but the in/out contract (clear = 0) keeps
// the copies.  Unset = 2, by measurement (profiles/r1_e2e_ab.txt,
// B200, PCIe Gen5): 1080p 0.397 -> 0.346 ms, and 0.58 -> 0.35 ms when the CPU has just rewritten the buffers (the
// reference clears on the host before every draw: dirty cache lines slow the copy engine's writes, not these);
// 4K 1.235 -> 1.295 ms untouched, 1.60 -> 1.31 ms rewritten; sdf_6 octree 4K 1.79 -> 1.32 ms; in/out contract at 1080p
// 0.913 -> 0.458 ms.  Plain 32-byte stores (1) are 1.7x slower than the copies; 256-byte bulk stores measured the same
// as 128-byte ones.
constexpr int HOST_ZEROCOPY_DEFAULT = 2;
const int g_env_zero_copy = std::getenv("RT_HOST_ZEROCOPY") ? env_int("RT_HOST_ZEROCOPY") : HOST_ZEROCOPY_DEFAULT;
const int g_env_oct_refill = env_int("RT_OCT_REFILL");
const int g_env_oct_batch = env_int("RT_OCT_BATCH");
// RT_MULTI=0: scenes with several replicas (rt_init) still render on one device per call (A/B measurements)
const int g_env_multi = std::getenv("RT_MULTI") ? env_int("RT_MULTI") : 1;
// Heavy-first block scheduling (rt_scene.h: HeavySched).  RT_HEAVY_CYCLES: SM cycles above which a block counts as heavy
// and starts first in the next launch (0 = off).  RT_HEAVY_MIN_RAYS / RT_HEAVY_MAX_RAYS: the launch sizes it is used for.
// Measured on one box against the kernels without it (profiles/r2_heavy_ab.txt, kernel ms): one GPU's share of a 4K
// frame over 8 / 4 GPUs, 513^3 grid 0.104 -> 0.088 / 0.138 -> 0.126, bunny 0.080 -> 0.070 / 0.129 -> 0.109; 1080p full
// frame 0.112 -> 0.109 (grid), 0.169 -> 0.160 (default look).  A full 4K frame would lose 4-6 % (the tail it removes is
// small there, the slot blocks and the per-block look-up are not) and frames under ~0.5 Mrays lose more (cube 512^2:
// +20 %): both stay on the plain kernels, which are compiled without any of it.
constexpr int HEAVY_CYCLES_DEFAULT = 35000;
constexpr size_t HEAVY_MIN_GRID_BYTES = 32u << 20;
const int g_env_heavy = std::getenv("RT_HEAVY_CYCLES") ? env_int("RT_HEAVY_CYCLES") : HEAVY_CYCLES_DEFAULT;
const long long g_env_heavy_min_rays = std::getenv("RT_HEAVY_MIN_RAYS") ? std::atoll(std::getenv("RT_HEAVY_MIN_RAYS")) : 600000ll;
const long long g_env_heavy_max_rays = std::getenv("RT_HEAVY_MAX_RAYS") ? std::atoll(std::getenv("RT_HEAVY_MAX_RAYS")) : 3300000ll;

// per-scene-type default scheduling, chosen by measurement (profiles/r1_tune.txt): the octree
// descent has the most uneven per-warp cost and gains from persistent warps pulling tiles; the grid marcher, the BVH
// traversal (with 128-thread blocks) and the tiny brute-force scenes are as fast or faster as plain tiles.
int resolve_sched(const rt_scene *s) {
  int m = g_sched.load();
  if (m != RT_SCHED_AUTO) return m;
  if (s->kind == RT_SCENE_OCTREE) return RT_SCHED_PERSISTENT;
  return RT_SCHED_TILES;
}

// How the tiles of one launch reach the framebuffer.
//   STORE_PLAIN : per-thread 32-byte tile-row stores (local HBM)
//   STORE_AUTO  : STORE_PLAIN, except for a framebuffer opened with rt_ipc_open on a frame split over >= 8 ranks
//   STORE_STAGED: the block's 32 x 4 pixels are parked in shared memory and leave as 128-byte TMA bulk stores
//                 (peer HBM over NVLink, pinned host memory over PCIe); needs the fused clear, W % 8 == 0
//   STORE_INOUT : STORE_STAGED for the reference's in/out framebuffer: the tile's previous contents are first fetched
//                 with bulk loads, only nearer hits replace them
enum StoreMode { STORE_AUTO = -1, STORE_PLAIN = 0, STORE_STAGED = 1, STORE_INOUT = 2 };

// the staged kernels exist for every scene kind, with and without secondary rays and prim ids, under tiles scheduling;
// the workload counters are only in the plain kernels
bool staged_kernel_available(const FrameDev &fd, const void *counters) {
  const int sched_now = g_sched.load();
  return (sched_now == RT_SCHED_AUTO || sched_now == RT_SCHED_TILES) && !counters && fd.W % 8u == 0u;
}

// Points a launch at the replica's heavy-first state (rt_scene.h: HeavySched).  Order only; see the notes there.
int prepare_heavy(Replica &rep, const FrameDev &fd, int kernel_shape, HeavySched &h, cudaStream_t st) {
  h = HeavySched{};
  const long long rays = (long long)fd.W * fd.rows_here;
  if (g_env_heavy <= 0 || rays > g_env_heavy_max_rays || rays < g_env_heavy_min_rays) return RT_OK;
  std::lock_guard<std::mutex> lock(rep.heavy_mu);
  const size_t tiles = (size_t)((fd.W + 7u) / 8u) * ((fd.rows_here + 3u) / 4u);  // the finest tile shape bounds them all
  if (tiles > rep.heavy_tiles || !rep.heavy_mem) {
    if (rep.heavy_mem) RT_CUDA(cudaFree(rep.heavy_mem));  // (synchronises with launches still using it)
    rep.heavy_mem = nullptr;
    rep.heavy_tiles = 0;
    const size_t words = 2 * (size_t)HEAVY_CAP + 2 + 2 * tiles;
    RT_CUDA(cudaMalloc(&rep.heavy_mem, words * 4));
    RT_CUDA(cudaMemset(rep.heavy_mem, 0, words * 4));
    rep.heavy_tiles = tiles;
    rep.heavy_sig = 0;
  }
  uint32_t *lists = rep.heavy_mem, *counts = lists + 2 * HEAVY_CAP, *done = counts + 2, *seen = done + rep.heavy_tiles;
  const uint64_t sig = ((uint64_t)fd.W << 40) ^ ((uint64_t)fd.rows_here << 20) ^ ((uint64_t)fd.band_count << 12) ^
                       ((uint64_t)fd.band_index << 4) ^ ((uint64_t)fd.band_rows << 52) ^ (uint64_t)kernel_shape;
  const uint32_t prev = rep.heavy_parity, next = prev ^ 1u;
  if (sig != rep.heavy_sig) {  // another frame geometry: tile ids of the old list mean nothing
    RT_CUDA(cudaMemsetAsync(counts + prev, 0, 4, st));
    rep.heavy_sig = sig;
  }
  RT_CUDA(cudaMemsetAsync(counts + next, 0, 4, st));
  h.prev = lists + prev * HEAVY_CAP;
  h.prev_count = counts + prev;
  h.next = lists + next * HEAVY_CAP;
  h.next_count = counts + next;
  h.done = done;
  h.seen = seen;
  if (++rep.heavy_tag == 0u) ++rep.heavy_tag;  // never 0 (the arrays start zeroed)
  h.tag = rep.heavy_tag;
  h.cycles = (uint32_t)g_env_heavy;
  rep.heavy_parity = next;
  return RT_OK;
}

int render_on_stream(const rt_scene *scene, Replica &rep, const FrameDev &fd, uint32_t *color, const float *t_in,
                     float *t_out, uint32_t *prim, rt_counters *counters_dev, cudaStream_t st, int store = STORE_AUTO,
                     bool allow_heavy = true) {
  RenderArgs a;
  a.scene = rep.dev;
  a.frame = fd;
  a.color = color;
  a.t_in = t_in;
  a.t_out = t_out;
  a.prim = prim;
  a.counters = reinterpret_cast<unsigned long long *>(counters_dev);
  a.tile_counter = rep.tile_counters + (rep.next_counter.fetch_add(1) % N_TILE_COUNTERS);
  a.tile_block = g_env_tile_block;
  a.stage_stores = 0;
  a.xfer = rep.xfer;
  int sched = resolve_sched(scene);
  if ((fd.clear || store == STORE_INOUT) && staged_kernel_available(fd, counters_dev) &&
      (store >= 0 ? store >= STORE_STAGED : g_env_stage != 0)) {
    bool staged = g_env_stage == 1 || store >= STORE_STAGED;
    // one process per GPU (rt_ipc_open): measured in round 1 (profiles/r1_multi/TABLE.txt) to pay when rank 0's inbound
    // NVLink is saturated, i.e. on 8 GPUs at 4K
    // (RT_STAGE_STORES=2 stages for every peer framebuffer, whatever the frame size and rank count: A/B measurements)
    if (!staged && (g_env_stage == 2 || ((uint64_t)fd.W * fd.H >= (3ull << 20) && fd.band_count >= 8u)))
      staged = is_peer_pointer(color) && is_peer_pointer(t_out);
    a.stage_stores = staged ? (store == STORE_INOUT ? 2 : 1) : 0;
    if (staged && store < 0) g_staged_launches.fetch_add(1);  // counts the one-process-per-GPU path only
    // the staged stores live in the tiles kernel: the octree's default (persistent, ~1 % faster on one GPU) gives way
    if (staged && g_sched.load() == RT_SCHED_AUTO) sched = RT_SCHED_TILES;
  }
  // RT_SCHED_COMPACT thresholds: best point of the sweep in profiles/r1_tune.txt (RT_OCT_REFILL / RT_OCT_BATCH override)
  a.oct_refill_min = g_env_oct_refill > 0 ? g_env_oct_refill : 32;
  a.oct_leaf_batch = g_env_oct_batch > 0 ? g_env_oct_batch : 6;
  a.heavy = HeavySched{};
  // a grid that stays in L2 has no long-latency tail for the order (or the prefetch) to hide: measured with the
  // natural-weights sampler, 65^3 grid, heavy-first on / off: 1080p 0.098 / 0.093 ms, a quarter of 4K 0.095 / 0.087, an
  // eighth 0.056 / 0.056; 257^3 (68 MB) 0.099 / 0.105 and 0.062 / 0.075, 513^3 0.113 / 0.140 and 0.081 / 0.107
  if (scene->kind == RT_SCENE_GRID && scene->bytes < HEAVY_MIN_GRID_BYTES) allow_heavy = false;
  if (allow_heavy && sched == RT_SCHED_TILES && !counters_dev) {
    const int rc = prepare_heavy(rep, fd, (a.stage_stores ? 1 : 0) | (a.tile_block << 2) | (scene->kind << 12), a.heavy, st);
    if (rc) return rc;
  }
  cudaError_t e = (g_math.load() == RT_MATH_FAST) ? fast::launch_render(a, sched, rep.sm_count, st)
                                                  : exact::launch_render(a, sched, rep.sm_count, st);
  if (e != cudaSuccess) return cuda_fail(e, "render launch");
  return RT_OK;
}

// Device-side address of a pinned (page-locked, mapped) host range [p, p + bytes), or null for pageable / device
// memory, a range that is only partly registered, or one the bulk copies cannot address (16-byte alignment).
void *mapped_host_range(const void *p, size_t bytes) {
  cudaPointerAttributes at, last;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  if (at.type != cudaMemoryTypeHost || !at.devicePointer) return nullptr;
  const char *end = static_cast<const char *>(p) + (bytes ? bytes - 1 : 0);
  if (cudaPointerGetAttributes(&last, end) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  if (last.type != cudaMemoryTypeHost ||
      static_cast<char *>(last.devicePointer) != static_cast<char *>(at.devicePointer) + (bytes ? bytes - 1 : 0))
    return nullptr;
  if (reinterpret_cast<uintptr_t>(at.devicePointer) % 16u) return nullptr;
  return at.devicePointer;
}

// streams and events of a replica (current device = rep.device)
int ensure_streams(Replica &r) {
  if (r.h_cnt) return RT_OK;  // created last: a call that failed half-way is completed by the next one
  if (!r.h_stream) RT_CUDA(cudaStreamCreateWithFlags(&r.h_stream, cudaStreamNonBlocking));
  if (!r.h_up) RT_CUDA(cudaStreamCreateWithFlags(&r.h_up, cudaStreamNonBlocking));
  if (!r.h_down) RT_CUDA(cudaStreamCreateWithFlags(&r.h_down, cudaStreamNonBlocking));
  for (auto &e : r.ev)
    if (!e) RT_CUDA(cudaEventCreate(&e));
  for (auto &e : r.ev_up)
    if (!e) RT_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  for (auto &e : r.ev_k)
    if (!e) RT_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  if (!r.ev_fork) RT_CUDA(cudaEventCreateWithFlags(&r.ev_fork, cudaEventDisableTiming));
  if (!r.ev_join) RT_CUDA(cudaEventCreateWithFlags(&r.ev_join, cudaEventDisableTiming));
  RT_CUDA(cudaMalloc(&r.h_cnt, sizeof(rt_counters)));
  return RT_OK;
}

// device framebuffer of the copy pipeline: only the paths that copy allocate it (12 bytes per pixel)
int ensure_device_fb(Replica &r, size_t pixels) {
  if (pixels <= r.h_pixels) return RT_OK;
  if (r.h_color) cudaFree(r.h_color);
  if (r.h_t) cudaFree(r.h_t);
  if (r.h_prim) cudaFree(r.h_prim);
  r.h_color = nullptr; r.h_t = nullptr; r.h_prim = nullptr; r.h_pixels = 0;
  RT_CUDA(cudaMalloc(&r.h_color, pixels * 4));
  RT_CUDA(cudaMalloc(&r.h_t, pixels * 4));
  RT_CUDA(cudaMalloc(&r.h_prim, pixels * 4));
  r.h_pixels = pixels;
  return RT_OK;
}

Replica *replica_on(const rt_scene *s, int device) {
  for (auto &r : s->reps)
    if (r->device == device) return r.get();
  return nullptr;
}

// rows per interleaved band when ONE call fans a frame out over the replicas (a multiple of the 4-row staged block;
// hit density is mid-screen, so contiguous slabs would load-imbalance: SURVEY 8(e))
uint32_t multi_band_rows() {
  static const int v = env_int("RT_MULTI_BAND_ROWS");
  return v > 0 ? (uint32_t)((v + 3) & ~3) : 8u;
}

// frame parameters of replica j's share of a whole frame split over n devices
FrameDev band_share(const FrameDev &whole, uint32_t n, uint32_t j) {
  FrameDev f = whole;
  f.band_rows = multi_band_rows();
  f.band_count = n;
  f.band_index = j;
  const uint32_t nBands = (f.H + f.band_rows - 1) / f.band_rows;
  const uint32_t mine = nBands > j ? (nBands - j + n - 1) / n : 0;
  f.rows_here = mine * f.band_rows;
  return f;
}

// Copies the interleaved bands of replica j between a host frame and the device frame of the same layout, as ONE
// strided 2-D copy per buffer (band k of share j starts at row (k * n + j) * band_rows) plus the frame's ragged tail.
int copy_bands(void *dst, const void *src, const FrameDev &f, cudaMemcpyKind kind, cudaStream_t st) {
  const size_t rowBytes = (size_t)f.W * 4, bandBytes = rowBytes * f.band_rows, pitch = bandBytes * f.band_count;
  const uint32_t nBands = (f.H + f.band_rows - 1) / f.band_rows;
  uint32_t mine = nBands > f.band_index ? (nBands - f.band_index + f.band_count - 1) / f.band_count : 0;
  if (mine == 0) return RT_OK;
  const size_t off = bandBytes * f.band_index;
  const uint32_t lastRow0 = ((mine - 1) * f.band_count + f.band_index) * f.band_rows;  // first row of my last band
  const uint32_t lastRows = std::min(f.band_rows, f.H - lastRow0);
  const uint32_t full = lastRows == f.band_rows ? mine : mine - 1;
  if (full)
    RT_CUDA(cudaMemcpy2DAsync(static_cast<char *>(dst) + off, pitch, static_cast<const char *>(src) + off, pitch,
                              bandBytes, full, kind, st));
  if (full != mine)
    RT_CUDA(cudaMemcpyAsync(static_cast<char *>(dst) + (size_t)lastRow0 * rowBytes,
                            static_cast<const char *>(src) + (size_t)lastRow0 * rowBytes, (size_t)lastRows * rowBytes, kind,
                            st));
  return RT_OK;
}

double wall_ms_since(std::chrono::steady_clock::time_point t0) {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

// ---- Renderer::draw from ONE call over every replica: host framebuffer ----------------------------------------------
// Each device renders its interleaved bands.  Pinned framebuffer: each kernel stores (and, for the in/out contract,
// first loads) its tiles straight in the caller's buffers over ITS OWN PCIe link, so n devices bring n links to a
// transfer that bounds the single-device call (0.35 ms of 1080p frame against a 0.11 ms kernel).  Anything else:
// per-device framebuffers, one strided copy per buffer and direction.
int render_host_multi(rt_scene *s, const FrameDev &fd, uint32_t *color, float *t_inout, uint32_t *prim,
                      float *device_ms, float *total_ms, rt_counters *counters) {
  const auto t0 = std::chrono::steady_clock::now();
  const uint32_t n = (uint32_t)s->reps.size();
  const size_t pixels = (size_t)fd.W * fd.H, bytes = pixels * 4;
  const bool inout = !fd.clear;
  struct Drain {
    rt_scene *s;
    bool armed = true;
    ~Drain() {
      if (!armed) return;
      for (auto &r : s->reps) {
        DeviceGuard g(r->device);
        if (r->h_stream) cudaStreamSynchronize(r->h_stream);
      }
    }
  } drain{s};
  for (auto &r : s->reps) {
    DeviceGuard g(r->device);
    if (!g.ok) return fail(RT_ERR_CUDA, "cannot select a replica's device");
    int rc = ensure_streams(*r);
    if (rc) return rc;
  }
  const bool timing = device_ms != nullptr;
  uint32_t *dc = nullptr;
  float *dt = nullptr;
  uint32_t *dp = nullptr;
  if (g_env_zero_copy != 0 && staged_kernel_available(fd, counters) && (inout ? g_env_zero_copy == 2 : true)) {
    dc = static_cast<uint32_t *>(mapped_host_range(color, bytes));
    dt = static_cast<float *>(mapped_host_range(t_inout, bytes));
    if (prim) dp = static_cast<uint32_t *>(mapped_host_range(prim, bytes));
  }
  const bool direct = dc && dt && (!prim || dp);
  for (uint32_t j = 0; j < n; ++j) {
    Replica &r = *s->reps[j];
    DeviceGuard g(r.device);
    const FrameDev fj = band_share(fd, n, j);
    cudaStream_t st = r.h_stream;
    int rc = RT_OK;
    if (direct) {
      if (timing) RT_CUDA(cudaEventRecord(r.ev[0], st));
      rc = render_on_stream(s, r, fj, dc, inout ? dt : nullptr, dt, dp, nullptr, st,
                            inout ? STORE_INOUT : (g_env_zero_copy >= 2 ? STORE_STAGED : STORE_PLAIN));
      if (rc) return rc;
      if (timing) RT_CUDA(cudaEventRecord(r.ev[3], st));
      continue;
    }
    rc = ensure_device_fb(r, pixels);
    if (rc) return rc;
    if (inout) {
      if ((rc = copy_bands(r.h_t, t_inout, fj, cudaMemcpyHostToDevice, st))) return rc;
      if ((rc = copy_bands(r.h_color, color, fj, cudaMemcpyHostToDevice, st))) return rc;
    }
    if (prim && inout) RT_CUDA(cudaMemsetAsync(r.h_prim, 0xFF, bytes, st));  // pixels not written: MISS
    if (counters) RT_CUDA(cudaMemsetAsync(r.h_cnt, 0, sizeof(rt_counters), st));
    if (timing) RT_CUDA(cudaEventRecord(r.ev[0], st));
    rc = render_on_stream(s, r, fj, r.h_color, r.h_t, r.h_t, prim ? r.h_prim : nullptr, counters ? r.h_cnt : nullptr, st,
                          STORE_PLAIN);
    if (rc) return rc;
    if (timing) RT_CUDA(cudaEventRecord(r.ev[3], st));
  }
  // downloads in a second pass: a copy into pageable memory blocks the host until it is done, and by then every other
  // device's kernel is already running
  for (uint32_t j = 0; j < n && !direct; ++j) {
    Replica &r = *s->reps[j];
    DeviceGuard g(r.device);
    const FrameDev fj = band_share(fd, n, j);
    int rc;
    if ((rc = copy_bands(color, r.h_color, fj, cudaMemcpyDeviceToHost, r.h_stream))) return rc;
    if ((rc = copy_bands(t_inout, r.h_t, fj, cudaMemcpyDeviceToHost, r.h_stream))) return rc;
    if (prim && (rc = copy_bands(prim, r.h_prim, fj, cudaMemcpyDeviceToHost, r.h_stream))) return rc;
  }
  if (counters) std::memset(counters, 0, sizeof *counters);
  float worst = 0.0f;
  for (uint32_t j = 0; j < n; ++j) {
    Replica &r = *s->reps[j];
    DeviceGuard g(r.device);
    RT_CUDA(cudaStreamSynchronize(r.h_stream));
    if (timing) {
      float ms = 0.0f;
      RT_CUDA(cudaEventElapsedTime(&ms, r.ev[0], r.ev[3]));
      worst = std::max(worst, ms);
    }
    if (counters) {
      rt_counters c;
      RT_CUDA(cudaMemcpy(&c, r.h_cnt, sizeof c, cudaMemcpyDeviceToHost));
      const uint64_t *src = reinterpret_cast<const uint64_t *>(&c);
      uint64_t *dst = reinterpret_cast<uint64_t *>(counters);
      for (size_t k = 0; k < sizeof(rt_counters) / sizeof(uint64_t); ++k) dst[k] += src[k];
    }
  }
  drain.armed = false;
  if (device_ms) *device_ms = worst;  // the slowest device's kernel: the frame is done when the last band is
  if (total_ms) *total_ms = (float)wall_ms_since(t0);
  return RT_OK;
}

// ---- the same on a DEVICE-resident framebuffer: tiles gathered over NVLink inside the render kernels ----------------
// The framebuffer lives on one of the replicas' devices (the owner).  The owner renders its bands on the caller's
// stream; every other replica waits for the caller's stream (event fork), renders its bands with the tile staged in
// shared memory and stored into the owner's HBM as 128-byte bulk writes over NVLink (peer access enabled by rt_init),
// and the caller's stream waits for all of them (event join).  Asynchronous like the single-device call.
int render_device_multi(const rt_scene *s, Replica &owner, const FrameDev &fd, uint32_t *color, const float *t_in,
                        float *t_out, uint32_t *prim, cudaStream_t user) {
  const uint32_t n = (uint32_t)s->reps.size();
  // The fork / join events and the replicas' streams belong to the scene: calls from several host threads enqueue their
  // whole record / wait / launch sequence one after the other (a wait captures the event's latest record at call time,
  // so sequences that do not interleave on the host cannot see each other's events).
  std::lock_guard<std::mutex> lock(const_cast<rt_scene *>(s)->host_mu);
  for (auto &r : s->reps) {
    DeviceGuard g(r->device);
    if (!g.ok) return fail(RT_ERR_CUDA, "cannot select a replica's device");
    int rc = ensure_streams(*r);
    if (rc) return rc;
  }
  const bool inout = !fd.clear && t_in != nullptr;
  // Staged 128-byte stores pay when the owner's NVLink ingress is the limit (8 GPUs on a >= 3 Mpixel frame: 4K grid
  // 0.139 -> 0.096 ms in round 1); below that the plain kernel's 32-byte tile rows are faster (2 GPUs, 4K grid: 0.221 vs
  // 0.247 ms, profiles/r2_multi/).  The in/out contract always stages: its bulk loads replace per-thread remote reads.
  const bool worth = g_env_stage == 2 || g_env_stage == 1 || inout || (n >= 8u && (uint64_t)fd.W * fd.H >= (3ull << 20));
  const bool can_stage = worth && staged_kernel_available(fd, nullptr) && (fd.clear || inout) && g_env_stage != 0 &&
                         reinterpret_cast<uintptr_t>(color) % 16u == 0 && reinterpret_cast<uintptr_t>(t_out) % 16u == 0 &&
                         reinterpret_cast<uintptr_t>(prim) % 16u == 0 && (!inout || t_in == t_out);
  {
    DeviceGuard g(owner.device);
    RT_CUDA(cudaEventRecord(owner.ev_fork, user));
  }
  for (uint32_t j = 0; j < n; ++j) {
    Replica &r = *s->reps[j];
    const FrameDev fj = band_share(fd, n, j);
    DeviceGuard g(r.device);
    if (&r == &owner) {
      int rc = render_on_stream(s, r, fj, color, t_in, t_out, prim, nullptr, user, STORE_PLAIN);
      if (rc) return rc;
      continue;
    }
    RT_CUDA(cudaStreamWaitEvent(r.h_stream, owner.ev_fork, 0));
    int rc = render_on_stream(s, r, fj, color, t_in, t_out, prim, nullptr, r.h_stream,
                              can_stage ? (inout ? STORE_INOUT : STORE_STAGED) : STORE_PLAIN);
    if (rc) return rc;
    RT_CUDA(cudaEventRecord(r.ev_join, r.h_stream));
  }
  DeviceGuard g(owner.device);
  for (auto &r : s->reps)
    if (r.get() != &owner) RT_CUDA(cudaStreamWaitEvent(user, r->ev_join, 0));
  return RT_OK;
}

}  // namespace

extern "C" {

// ---- multi-GPU from one process (rt_init(devs, n)) ------------------------------------------------------
int rt_init(const int *devices, int n) {
  if (n < 0 || (n > 0 && !devices)) return fail(RT_ERR_INVALID, "bad device list");
  int count = 0;
  RT_CUDA(cudaGetDeviceCount(&count));
  std::vector<int> devs(devices, devices + n);
  for (int i = 0; i < n; ++i) {
    if (devs[i] < 0 || devs[i] >= count) return fail(RT_ERR_INVALID, "device index out of range");
    for (int k = 0; k < i; ++k)
      if (devs[k] == devs[i]) return fail(RT_ERR_INVALID, "device listed twice");
  }
  // peer access both ways between every pair: the device-resident fan-out stores tiles into the owner's HBM
  for (int i = 0; i < n; ++i) {
    DeviceGuard g(devs[i]);
    if (!g.ok) return fail(RT_ERR_CUDA, "cannot select a listed device");
    for (int k = 0; k < n; ++k) {
      if (k == i) continue;
      int can = 0;
      RT_CUDA(cudaDeviceCanAccessPeer(&can, devs[i], devs[k]));
      if (!can) continue;  // no NVLink / PCIe peer path: rt_render_device then keeps such a pair apart (plain copies)
      cudaError_t e = cudaDeviceEnablePeerAccess(devs[k], 0);
      if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
      else if (e != cudaSuccess) return cuda_fail(e, "cudaDeviceEnablePeerAccess");
    }
  }
  std::lock_guard<std::mutex> lock(g_dev_mu);
  g_devs = devs;
  return RT_OK;
}

int rt_device_list(int *devices, int cap) {
  std::lock_guard<std::mutex> lock(g_dev_mu);
  for (size_t j = 0; j < g_devs.size() && devices && (int)j < cap; ++j) devices[j] = g_devs[j];
  return (int)g_devs.size();
}

int rt_render_device(const rt_scene *scene, const rt_frame *frame, uint32_t *color, const float *t_in, float *t_out,
                     uint32_t *prim, rt_counters *counters_dev, void *stream) {
  if (!scene) return fail(RT_ERR_INVALID, "scene is null");
  if (!color || !t_out) return fail(RT_ERR_INVALID, "color / t_out must not be null");
  FrameDev fd;
  int rc = make_frame(frame, fd);
  if (rc) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (scene->reps.size() > 1) {
    // where does the framebuffer live?  (a pointer opened with rt_ipc_open reports the mapping device: the one-process-
    // per-GPU path, which partitions the frame itself)
    cudaPointerAttributes at;
    Replica *owner = nullptr;
    if (cudaPointerGetAttributes(&at, color) == cudaSuccess && at.type == cudaMemoryTypeDevice)
      owner = replica_on(scene, at.device);
    else
      cudaGetLastError();
    if (owner && fd.band_count <= 1 && !counters_dev && !is_peer_pointer(color) && g_env_multi != 0)
      return render_device_multi(scene, *owner, fd, color, t_in, t_out, prim, st);
    if (!owner) {
      int cur = 0;
      RT_CUDA(cudaGetDevice(&cur));
      owner = replica_on(scene, cur);
    }
    if (!owner) owner = scene->reps[0].get();
    DeviceGuard g(owner->device);
    if (!g.ok) return fail(RT_ERR_CUDA, "cannot select the scene's device");
    return render_on_stream(scene, *owner, fd, color, t_in, t_out, prim, counters_dev, st);
  }
  Replica &r = *scene->reps[0];
  DeviceGuard g(r.device);
  if (!g.ok) return fail(RT_ERR_CUDA, "cannot select the scene's device");
  return render_on_stream(scene, r, fd, color, t_in, t_out, prim, counters_dev, st);
}

int rt_render_host(const rt_scene *scene_c, const rt_frame *frame, uint32_t *color, float *t_inout, uint32_t *prim,
                   float *device_ms, float *total_ms, rt_counters *counters) {
  if (!scene_c) return fail(RT_ERR_INVALID, "scene is null");
  if (!color || !t_inout) return fail(RT_ERR_INVALID, "color / t must not be null");
  FrameDev fd;
  int rc = make_frame(frame, fd);
  if (rc) return rc;
  rt_scene *s = const_cast<rt_scene *>(scene_c);
  std::lock_guard<std::mutex> lock(s->host_mu);
  // one call, every device of rt_init's list (a caller that partitions the frame itself keeps the single-device path)
  if (s->reps.size() > 1 && fd.band_count <= 1 && g_env_multi != 0)
    return render_host_multi(s, fd, color, t_inout, prim, device_ms, total_ms, counters);
  Replica &r = *s->reps[0];
  DeviceGuard g(r.device);
  if (!g.ok) return fail(RT_ERR_CUDA, "cannot select the scene's device");
  const size_t pixels = (size_t)fd.W * fd.H, bytes = pixels * 4;
  rc = ensure_streams(r);
  if (rc) return rc;
  cudaStream_t st = r.h_stream;
  // An error return must not leave copies into the caller's buffers in flight: drain the three streams on the way out.
  struct Drain {
    Replica *r;
    bool armed = true;
    ~Drain() {
      if (!armed) return;
      cudaStreamSynchronize(r->h_up);
      cudaStreamSynchronize(r->h_stream);
      cudaStreamSynchronize(r->h_down);
    }
  } drain{&r};
  // Pinned host framebuffer (colour, depth and, when asked for, prim ids): the kernel works on the caller's buffers over PCIe
  // (no device framebuffer, no copy engine, one launch).  clear = 1: tiles leave as 128-byte bulk stores; clear = 0 (the
  // reference's in/out contract): the tile's previous colour and depth are fetched with bulk loads first.  Banded calls
  // take the same route (the staged kernel walks the call's bands).  Mode 1 = plain 32-byte stores (measurements only).
  const bool inout = !fd.clear;
  const bool staged_ok = staged_kernel_available(fd, counters);
  if (g_env_zero_copy != 0 && (inout ? (g_env_zero_copy == 2 && staged_ok) : (g_env_zero_copy == 1 || staged_ok)) &&
      !counters && fd.W % 8u == 0u && (fd.band_count <= 1 || fd.band_rows % 4u == 0u)) {
    uint32_t *dc = static_cast<uint32_t *>(mapped_host_range(color, bytes));
    float *dt = static_cast<float *>(mapped_host_range(t_inout, bytes));
    uint32_t *dp = prim ? static_cast<uint32_t *>(mapped_host_range(prim, bytes)) : nullptr;
    if (dc && dt && (!prim || dp)) {
      RT_CUDA(cudaEventRecord(r.ev[0], st));
      rc = render_on_stream(s, r, fd, dc, inout ? dt : nullptr, dt, dp, nullptr, st,
                            inout ? STORE_INOUT : (g_env_zero_copy >= 2 && staged_ok ? STORE_STAGED : STORE_PLAIN));
      if (rc) return rc;
      RT_CUDA(cudaEventRecord(r.ev[3], st));
      RT_CUDA(cudaStreamSynchronize(st));
      float ms = 0.0f;
      if (device_ms || total_ms) RT_CUDA(cudaEventElapsedTime(&ms, r.ev[0], r.ev[3]));
      if (device_ms) *device_ms = ms;
      if (total_ms) *total_ms = ms;
      drain.armed = false;
      return RT_OK;
    }
  }
  rc = ensure_device_fb(r, pixels);
  if (rc) return rc;
  // The frame is cut into row chunks that flow through three streams: upload (only when the caller's depth /
  // colour matter: the framebuffer is in/out, raytracing.cpp:17,90-94), render, download.  The PCIe copies of one
  // chunk overlap the kernel of the next, so the call costs ~max(copy time, kernel time), not their sum.
  const bool whole = fd.band_count <= 1;
  const bool need_upload = !(fd.clear && whole);  // a fused clear of the whole frame overwrites every pixel
  // Chunk boundaries grow geometrically: the kernel outruns PCIe (~17 Grays/s vs 56 GB/s / 8 B = 7 Gpixels/s), so a
  // small first chunk gets the first download going after a few microseconds and each later, larger chunk is rendered
  // while the previous one is still on the wire.  RT_HOST_CHUNKS=n forces n equal chunks (measurement sweeps).
  uint32_t bounds[18];
  uint32_t n_chunks = 1;
  bounds[0] = 0;
  bounds[1] = fd.H;
  if (whole && fd.H >= 64) {
    if (g_env_host_chunks > 0) {
      n_chunks = std::min<uint32_t>(16u, std::min<uint32_t>(fd.H, (uint32_t)g_env_host_chunks));
      const uint32_t rows = (fd.H + n_chunks - 1) / n_chunks;
      for (uint32_t k = 1; k <= n_chunks; ++k) bounds[k] = std::min(fd.H, k * rows);
    } else {
      uint32_t rows = std::max<uint32_t>(16u, fd.H / 40u), r0 = 0;
      n_chunks = 0;
      while (r0 < fd.H && n_chunks < 15) {
        r0 = std::min(fd.H, r0 + rows);
        if (fd.H - r0 < rows) r0 = fd.H;  // fold a short remainder into this chunk
        bounds[++n_chunks] = r0;
        rows *= 2;
      }
      if (r0 < fd.H) bounds[++n_chunks] = fd.H;
    }
  }
  RT_CUDA(cudaEventRecord(r.ev[0], st));
  if (prim && need_upload) RT_CUDA(cudaMemsetAsync(r.h_prim, 0xFF, bytes, st));  // rows not rendered: MISS
  if (counters) RT_CUDA(cudaMemsetAsync(r.h_cnt, 0, sizeof(rt_counters), st));
  RT_CUDA(cudaEventRecord(r.ev[1], st));
  RT_CUDA(cudaStreamWaitEvent(r.h_up, r.ev[0], 0));
  for (uint32_t k = 0; k < n_chunks; ++k) {
    const uint32_t r0 = bounds[k], r1 = bounds[k + 1];
    if (r0 >= r1) continue;
    const size_t off = (size_t)r0 * fd.W, cbytes = (size_t)(r1 - r0) * fd.W * 4;
    FrameDev fk = fd;
    if (n_chunks > 1) {  // rows [r0, r1) of the frame
      fk.row0 = r0;
      fk.rows_here = r1 - r0;
    }
    if (need_upload) {
      const size_t uoff = n_chunks > 1 ? off : 0, ubytes = n_chunks > 1 ? cbytes : bytes;
      RT_CUDA(cudaMemcpyAsync(r.h_t + uoff, t_inout + uoff, ubytes, cudaMemcpyHostToDevice, r.h_up));
      RT_CUDA(cudaMemcpyAsync(r.h_color + uoff, color + uoff, ubytes, cudaMemcpyHostToDevice, r.h_up));
      RT_CUDA(cudaEventRecord(r.ev_up[k], r.h_up));
      RT_CUDA(cudaStreamWaitEvent(st, r.ev_up[k], 0));
    }
    rc = render_on_stream(s, r, fk, r.h_color, r.h_t, r.h_t, prim ? r.h_prim : nullptr, counters ? r.h_cnt : nullptr, st,
                          STORE_PLAIN, n_chunks == 1);
    if (rc) return rc;
    RT_CUDA(cudaEventRecord(r.ev_k[k], st));
    RT_CUDA(cudaStreamWaitEvent(r.h_down, r.ev_k[k], 0));
    const size_t doff = n_chunks > 1 ? off : 0, dbytes = n_chunks > 1 ? cbytes : bytes;
    RT_CUDA(cudaMemcpyAsync(color + doff, r.h_color + doff, dbytes, cudaMemcpyDeviceToHost, r.h_down));
    RT_CUDA(cudaMemcpyAsync(t_inout + doff, r.h_t + doff, dbytes, cudaMemcpyDeviceToHost, r.h_down));
    if (prim) RT_CUDA(cudaMemcpyAsync(prim + doff, r.h_prim + doff, dbytes, cudaMemcpyDeviceToHost, r.h_down));
  }
  RT_CUDA(cudaEventRecord(r.ev[2], st));
  if (counters) RT_CUDA(cudaMemcpyAsync(counters, r.h_cnt, sizeof(rt_counters), cudaMemcpyDeviceToHost, st));
  RT_CUDA(cudaStreamSynchronize(st));
  RT_CUDA(cudaEventRecord(r.ev[3], r.h_down));
  RT_CUDA(cudaStreamSynchronize(r.h_down));
  if (device_ms) RT_CUDA(cudaEventElapsedTime(device_ms, r.ev[1], r.ev[2]));
  if (total_ms) RT_CUDA(cudaEventElapsedTime(total_ms, r.ev[0], r.ev[3]));
  drain.armed = false;  // both syncs above succeeded: nothing is pending
  return RT_OK;
}

int rt_intersect_host(const rt_scene *scene, const float *origins, const float *dirs, uint32_t n, float t_near,
                      float t_far, int has_plane, float plane_y, uint8_t *hit, float *t, float *normal,
                      uint32_t *prim) {
  if (!scene) return fail(RT_ERR_INVALID, "scene is null");
  if (n == 0) return RT_OK;
  if (!origins || !dirs) return fail(RT_ERR_INVALID, "origins / dirs must not be null");
  const Replica &r0 = *scene->reps[0];
  DeviceGuard g(r0.device);
  if (!g.ok) return fail(RT_ERR_CUDA, "cannot select the scene's device");
  float *dO = nullptr, *dD = nullptr, *dT = nullptr, *dN = nullptr;
  uint8_t *dH = nullptr;
  uint32_t *dP = nullptr;
  int rc = RT_OK;
  auto cleanup = [&]() {
    cudaFree(dO); cudaFree(dD); cudaFree(dT); cudaFree(dN); cudaFree(dH); cudaFree(dP);
  };
#define RT_TRY(call)                                     \
  do {                                                   \
    cudaError_t e__ = (call);                            \
    if (e__ != cudaSuccess) { rc = cuda_fail(e__, #call); cleanup(); return rc; } \
  } while (0)
  RT_TRY(cudaMalloc(&dO, (size_t)n * 12));
  RT_TRY(cudaMalloc(&dD, (size_t)n * 12));
  RT_TRY(cudaMalloc(&dT, (size_t)n * 4));
  RT_TRY(cudaMalloc(&dN, (size_t)n * 12));
  RT_TRY(cudaMalloc(&dH, (size_t)n));
  RT_TRY(cudaMalloc(&dP, (size_t)n * 4));
  RT_TRY(cudaMemcpy(dO, origins, (size_t)n * 12, cudaMemcpyHostToDevice));
  RT_TRY(cudaMemcpy(dD, dirs, (size_t)n * 12, cudaMemcpyHostToDevice));
  RayArgs a;
  a.scene = r0.dev;
  a.origins = dO; a.dirs = dD; a.n = n;
  a.t_near = t_near; a.t_far = t_far;
  a.has_plane = has_plane != 0; a.plane_y = plane_y;
  a.hit = dH; a.t = dT; a.normal = dN; a.prim = dP;
  RT_TRY((g_math.load() == RT_MATH_FAST) ? fast::launch_rays(a, nullptr) : exact::launch_rays(a, nullptr));
  RT_TRY(cudaDeviceSynchronize());
  if (hit) RT_TRY(cudaMemcpy(hit, dH, (size_t)n, cudaMemcpyDeviceToHost));
  if (t) RT_TRY(cudaMemcpy(t, dT, (size_t)n * 4, cudaMemcpyDeviceToHost));
  if (normal) RT_TRY(cudaMemcpy(normal, dN, (size_t)n * 12, cudaMemcpyDeviceToHost));
  if (prim) RT_TRY(cudaMemcpy(prim, dP, (size_t)n * 4, cudaMemcpyDeviceToHost));
#undef RT_TRY
  cleanup();
  return RT_OK;
}

}  // extern "C"

// =============================================================================================================
// rt_intersect_host_static: rt_intersect_host for callers that must not allocate (no counterpart in the reference).
// rt_intersect_host stages the rays and their results in six device buffers it
// cudaMallocs and frees on every call; here the same staging lives in ONE file-scope device array of fixed size, so the
// call makes no cudaMalloc / cudaFree (nor any allocating library routine).  Every GPU that loads the library holds the
// array (2.8 MiB) until the process exits.  All calls on a device share it and it keeps the last call's contents: the
// result area of the n rays is cleared first, and calls that overlap on one GPU would race (inside one process they are
// serialised by a mutex and the call is synchronous; nothing protects against a caller's own concurrent kernels on it).
// =============================================================================================================
namespace {
constexpr size_t kStaticRays = RT_INTERSECT_STATIC_MAX_RAYS;
// per ray: origin 12 B, direction 12 B, t 4 B, normal 12 B, prim 4 B, hit 1 B; each section starts 256-byte aligned
// (kStaticRays is a multiple of 256)
constexpr size_t kOffO = 0, kOffD = kOffO + kStaticRays * 12, kOffT = kOffD + kStaticRays * 12,
                 kOffN = kOffT + kStaticRays * 4, kOffP = kOffN + kStaticRays * 12, kOffH = kOffP + kStaticRays * 4,
                 kStaticBytes = kOffH + kStaticRays;
static_assert(kStaticRays % 256 == 0, "sections of the static scratch must stay 256-byte aligned");
}  // namespace
// (in host code this name is a host-side stand-in, not a device address: cudaGetSymbolAddress below)
static __device__ __align__(256) unsigned char g_ray_scratch[kStaticBytes];

// clears the result area of the first n rays (device code names the array directly)
static __global__ void ray_scratch_clear_kernel(uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  reinterpret_cast<float *>(g_ray_scratch + kOffT)[i] = 0.0f;
  float *nrm = reinterpret_cast<float *>(g_ray_scratch + kOffN);
  nrm[3 * i] = nrm[3 * i + 1] = nrm[3 * i + 2] = 0.0f;
  reinterpret_cast<uint32_t *>(g_ray_scratch + kOffP)[i] = 0u;
  g_ray_scratch[kOffH + i] = 0;
}

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_ray_scratch_mu;
unsigned char *g_ray_scratch_addr[kMaxDevices] = {};  // per device: the array's DEVICE address, resolved once

// set-up, once per device, with that device current
int ray_scratch_on(int device, unsigned char **out) {
  if (device < 0 || device >= kMaxDevices) return fail(RT_ERR_INVALID, "device index out of range for the static scratch");
  if (!g_ray_scratch_addr[device]) {
    void *p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_ray_scratch);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_ray_scratch)");
    if (!p) return fail(RT_ERR_CUDA, "cudaGetSymbolAddress returned a null address");
    g_ray_scratch_addr[device] = static_cast<unsigned char *>(p);
  }
  *out = g_ray_scratch_addr[device];
  return RT_OK;
}
}  // namespace

extern "C" {

int rt_intersect_host_static(const rt_scene *scene, const float *origins, const float *dirs, uint32_t n, float t_near,
                             float t_far, int has_plane, float plane_y, uint8_t *hit, float *t, float *normal,
                             uint32_t *prim) {
  if (!scene) return fail(RT_ERR_INVALID, "scene is null");
  if (n > kStaticRays)
    return fail(RT_ERR_INVALID, "rt_intersect_host_static: n = " + std::to_string(n) + " rays exceeds the static scratch (at most " +
                                    std::to_string(kStaticRays) + " per call; split the batch or use rt_intersect_host)");
  if (n == 0) return RT_OK;
  if (!origins || !dirs) return fail(RT_ERR_INVALID, "origins / dirs must not be null");
  const Replica &r0 = *scene->reps[0];
  DeviceGuard g(r0.device);
  if (!g.ok) return fail(RT_ERR_CUDA, "cannot select the scene's device");
  std::lock_guard<std::mutex> lock(g_ray_scratch_mu);
  unsigned char *base = nullptr;
  int rc = ray_scratch_on(r0.device, &base);
  if (rc) return rc;
  float *dO = reinterpret_cast<float *>(base + kOffO), *dD = reinterpret_cast<float *>(base + kOffD);
  float *dT = reinterpret_cast<float *>(base + kOffT), *dN = reinterpret_cast<float *>(base + kOffN);
  uint32_t *dP = reinterpret_cast<uint32_t *>(base + kOffP);
  uint8_t *dH = base + kOffH;
  RT_CUDA(cudaMemcpy(dO, origins, (size_t)n * 12, cudaMemcpyHostToDevice));
  RT_CUDA(cudaMemcpy(dD, dirs, (size_t)n * 12, cudaMemcpyHostToDevice));
  ray_scratch_clear_kernel<<<(n + 255u) / 256u, 256>>>(n);  // the array keeps the last call's results
  RT_CUDA(cudaGetLastError());
  RayArgs a;
  a.scene = r0.dev;
  a.origins = dO; a.dirs = dD; a.n = n;
  a.t_near = t_near; a.t_far = t_far;
  a.has_plane = has_plane != 0; a.plane_y = plane_y;
  a.hit = dH; a.t = dT; a.normal = dN; a.prim = dP;
  RT_CUDA((g_math.load() == RT_MATH_FAST) ? fast::launch_rays(a, nullptr) : exact::launch_rays(a, nullptr));
  RT_CUDA(cudaDeviceSynchronize());
  if (hit) RT_CUDA(cudaMemcpy(hit, dH, (size_t)n, cudaMemcpyDeviceToHost));
  if (t) RT_CUDA(cudaMemcpy(t, dT, (size_t)n * 4, cudaMemcpyDeviceToHost));
  if (normal) RT_CUDA(cudaMemcpy(normal, dN, (size_t)n * 12, cudaMemcpyDeviceToHost));
  if (prim) RT_CUDA(cudaMemcpy(prim, dP, (size_t)n * 4, cudaMemcpyDeviceToHost));
  return RT_OK;
}

}  // extern "C"

// =============================================================================================================
// device buffers + CUDA IPC (one process per GPU; tiles are stored into the root's framebuffer over NVLink)
// =============================================================================================================
extern "C" {

int rt_device_malloc(uint64_t bytes, void **dev_ptr) {
  if (!dev_ptr) return fail(RT_ERR_INVALID, "dev_ptr is null");
  *dev_ptr = nullptr;
  RT_CUDA(cudaMalloc(dev_ptr, bytes ? bytes : 1));
  return RT_OK;
}

int rt_device_free(void *dev_ptr) {
  RT_CUDA(cudaFree(dev_ptr));
  return RT_OK;
}

int rt_ipc_export(const void *dev_ptr, unsigned char handle[RT_IPC_HANDLE_BYTES]) {
  static_assert(sizeof(cudaIpcMemHandle_t) == RT_IPC_HANDLE_BYTES, "IPC handle size");
  if (!dev_ptr || !handle) return fail(RT_ERR_INVALID, "null argument");
  cudaIpcMemHandle_t h;
  RT_CUDA(cudaIpcGetMemHandle(&h, const_cast<void *>(dev_ptr)));
  std::memcpy(handle, &h, sizeof h);
  return RT_OK;
}

int rt_ipc_open(const unsigned char handle[RT_IPC_HANDLE_BYTES], void **dev_ptr) {
  if (!dev_ptr || !handle) return fail(RT_ERR_INVALID, "null argument");
  cudaIpcMemHandle_t h;
  std::memcpy(&h, handle, sizeof h);
  *dev_ptr = nullptr;
  RT_CUDA(cudaIpcOpenMemHandle(dev_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  {
    std::lock_guard<std::mutex> lock(g_peer_mu);
    g_peer_ptrs.push_back(*dev_ptr);
  }
  return RT_OK;
}

uint64_t rt_staged_launch_count(void) { return g_staged_launches.load(); }

int rt_ipc_close(void *dev_ptr) {
  {
    std::lock_guard<std::mutex> lock(g_peer_mu);
    for (size_t i = 0; i < g_peer_ptrs.size(); ++i)
      if (g_peer_ptrs[i] == dev_ptr) {
        g_peer_ptrs.erase(g_peer_ptrs.begin() + (long)i);
        break;
      }
  }
  RT_CUDA(cudaIpcCloseMemHandle(dev_ptr));
  return RT_OK;
}

// ---- page-locked host framebuffers (rt_render_host stores straight into them) -------------------------------------
int rt_pinned_alloc(uint64_t bytes, void **host_ptr) {
  if (!host_ptr) return fail(RT_ERR_INVALID, "host_ptr is null");
  *host_ptr = nullptr;
  RT_CUDA(cudaHostAlloc(host_ptr, bytes ? bytes : 1, cudaHostAllocPortable | cudaHostAllocMapped));
  return RT_OK;
}

int rt_pinned_free(void *host_ptr) {
  if (!host_ptr) return RT_OK;
  RT_CUDA(cudaFreeHost(host_ptr));
  return RT_OK;
}

int rt_pinned_register(void *host_ptr, uint64_t bytes) {
  if (!host_ptr || bytes == 0) return fail(RT_ERR_INVALID, "null or empty host range");
  RT_CUDA(cudaHostRegister(host_ptr, bytes, cudaHostRegisterPortable | cudaHostRegisterMapped));
  return RT_OK;
}

int rt_pinned_unregister(void *host_ptr) {
  if (!host_ptr) return RT_OK;
  RT_CUDA(cudaHostUnregister(host_ptr));
  return RT_OK;
}

int rt_memcpy_d2h(void *host, const void *dev_ptr, uint64_t bytes, void *stream) {
  if (bytes == 0) return RT_OK;
  if (!host || !dev_ptr) return fail(RT_ERR_INVALID, "null argument");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  RT_CUDA(cudaMemcpyAsync(host, dev_ptr, bytes, cudaMemcpyDeviceToHost, st));
  RT_CUDA(cudaStreamSynchronize(st));
  return RT_OK;
}

int rt_memcpy_h2d(void *dev_ptr, const void *host, uint64_t bytes, void *stream) {
  if (bytes == 0) return RT_OK;
  if (!host || !dev_ptr) return fail(RT_ERR_INVALID, "null argument");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  RT_CUDA(cudaMemcpyAsync(dev_ptr, host, bytes, cudaMemcpyHostToDevice, st));
  RT_CUDA(cudaStreamSynchronize(st));
  return RT_OK;
}

}  // extern "C"
