// extern "C" boundary of the B200 primary-ray path (include/rt_b200.h): scene upload, validation, launches.
// No CPU fallback lives here: every entry point needs a CUDA device.
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
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

struct rt_scene {
  int kind = -1;
  int device = 0;
  int sm_count = 148;
  SceneDev dev{};
  std::vector<void *> allocs;
  uint64_t bytes = 0;
  float bbox_min[3] = {-1, -1, -1}, bbox_max[3] = {1, 1, 1};
  float build_ms = 0.0f;              // device time of an on-device BVH build
  uint32_t *tile_counters = nullptr;  // ring, one per in-flight persistent launch
  std::atomic<uint32_t> next_counter{0};
  // framebuffer cache of rt_render_host
  std::mutex host_mu;
  uint32_t *h_color = nullptr;
  float *h_t = nullptr;
  uint32_t *h_prim = nullptr;
  rt_counters *h_cnt = nullptr;
  size_t h_pixels = 0;
  cudaStream_t h_stream = nullptr, h_up = nullptr, h_down = nullptr;  // kernels, H2D copies, D2H copies
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t ev_up[16] = {}, ev_k[16] = {};  // per row chunk: upload done, kernel done
};

namespace {

int scene_begin(rt_scene **out, rt_scene *&s) {
  if (!out) return fail(RT_ERR_INVALID, "out pointer is null");
  *out = nullptr;
  int dev = 0;
  RT_CUDA(cudaGetDevice(&dev));
  s = new rt_scene;
  s->device = dev;
  cudaDeviceProp p;
  cudaError_t e = cudaGetDeviceProperties(&p, dev);
  if (e != cudaSuccess) {
    delete s;
    return cuda_fail(e, "cudaGetDeviceProperties");
  }
  s->sm_count = p.multiProcessorCount;
  return RT_OK;
}

template <class T>
int upload(rt_scene *s, const T *host, size_t n, const T **devOut) {
  void *d = nullptr;
  size_t bytes = std::max<size_t>(n * sizeof(T), 16);
  RT_CUDA(cudaMalloc(&d, bytes));
  s->allocs.push_back(d);
  s->bytes += n * sizeof(T);
  if (n) RT_CUDA(cudaMemcpy(d, host, n * sizeof(T), cudaMemcpyHostToDevice));
  *devOut = static_cast<const T *>(d);
  return RT_OK;
}

int scene_finish(rt_scene *s, rt_scene **out) {
  void *d = nullptr;
  RT_CUDA(cudaMalloc(&d, N_TILE_COUNTERS * sizeof(uint32_t)));
  s->allocs.push_back(d);
  s->tile_counters = static_cast<uint32_t *>(d);
  RT_CUDA(cudaMemset(d, 0, N_TILE_COUNTERS * sizeof(uint32_t)));
  RT_CUDA(cudaDeviceSynchronize());
  *out = s;
  return RT_OK;
}

void scene_free(rt_scene *s) {
  if (!s) return;
  DeviceGuard g(s->device);
  for (void *p : s->allocs) cudaFree(p);
  if (s->h_color) cudaFree(s->h_color);
  if (s->h_t) cudaFree(s->h_t);
  if (s->h_prim) cudaFree(s->h_prim);
  if (s->h_cnt) cudaFree(s->h_cnt);
  for (auto &e : s->ev)
    if (e) cudaEventDestroy(e);
  for (auto &e : s->ev_up)
    if (e) cudaEventDestroy(e);
  for (auto &e : s->ev_k)
    if (e) cudaEventDestroy(e);
  if (s->h_stream) cudaStreamDestroy(s->h_stream);
  if (s->h_up) cudaStreamDestroy(s->h_up);
  if (s->h_down) cudaStreamDestroy(s->h_down);
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
  if ((uint64_t)((sx + 7) & ~7u) * ((sy + 7) & ~7u) * ((sz + 7) & ~7u) >= (1ull << 32))
    return fail(RT_ERR_INVALID, "grid too large (element offsets must fit 32 bits)");
  if (layout < RT_GRID_AUTO || layout > RT_GRID_BRICKS) return fail(RT_ERR_INVALID, "unknown grid layout");
  rt_scene *s = nullptr;
  int rc = scene_begin(out, s);
  if (rc) return rc;
  s->kind = RT_SCENE_GRID;
  s->dev.kind = RT_SCENE_GRID;
  GridDev &g = s->dev.grid;
  g.sx = sx; g.sy = sy; g.sz = sz;
  g.fx = (float)(sx - 1); g.fy = (float)(sy - 1); g.fz = (float)(sz - 1);
  g.hfx = 0.5f * g.fx; g.hfy = 0.5f * g.fy; g.hfz = 0.5f * g.fz;
  const size_t n = (size_t)sx * sy * sz;
  if (layout == RT_GRID_AUTO) layout = RT_GRID_LINEAR;  // chosen by measurement, DESIGN.md "grid layout"
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
    s->allocs.push_back(b);
    s->bytes += count * sizeof(float);
    g.values = b;
    g.bricks = 1; g.stride_x = nby * nbz * 512u; g.stride_y = nbz * 512u;
  } else {
    g.bricks = 0; g.stride_x = sy * sz; g.stride_y = sz;
    rc = upload(s, values, n, &g.values);
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
  uint32_t *link = nullptr, depth = 0;
  float4 *vals = nullptr;
  std::string ierr;
  const int irc = ingest_octree_device(nodes36, n_nodes, &link, &vals, &depth, &s->build_ms, ierr);
  if (irc != INGEST_OK) {
    scene_free(s);
    return fail(irc == INGEST_INVALID ? RT_ERR_INVALID : RT_ERR_CUDA, ierr);
  }
  s->allocs.push_back(link);
  s->allocs.push_back(vals);
  s->bytes += (uint64_t)n_nodes * 36;
  s->kind = RT_SCENE_OCTREE;
  s->dev.kind = RT_SCENE_OCTREE;
  s->dev.oct.n_nodes = n_nodes;
  s->dev.oct.depth = depth;
  s->dev.oct.root_inner = depth > 0 ? 1u : 0u;
  s->dev.oct.link = link;
  s->dev.oct.values = vals;
  rc = scene_finish(s, out);
  if (rc != RT_OK) scene_free(s);
  return rc;
}

// ---------------------------------------------------------------------------------------------- mesh
static int finish_mesh(HostBvh &h, bool brute, bool literal_order, rt_scene **out) {
  rt_scene *s = nullptr;
  int rc = scene_begin(out, s);
  if (rc) return rc;
  s->kind = RT_SCENE_MESH;
  s->dev.kind = RT_SCENE_MESH;
  MeshDev &m = s->dev.mesh;
  m.n_tris = (uint32_t)h.tris.size();
  m.n_nodes = (uint32_t)h.nodes.size();
  m.brute = brute ? 1u : 0u;
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
    DeviceBvh d;
    std::string derr;
    if (!build_bvh8_device(vpos4, n_verts, indices, n_indices, d, derr)) {
      scene_free(s);
      return fail(derr.find("index out of range") != std::string::npos || derr.find("mesh needs") != std::string::npos
                      ? RT_ERR_INVALID
                      : RT_ERR_CUDA,
                  derr);
    }
    s->allocs.push_back(d.nodes);
    s->allocs.push_back(d.tris);
    s->bytes += (uint64_t)d.n_nodes * sizeof(BvhNode) + (uint64_t)d.n_tris * sizeof(TriDev);
    s->kind = RT_SCENE_MESH;
    s->dev.kind = RT_SCENE_MESH;
    MeshDev &m = s->dev.mesh;
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
  // brute force keeps the whole mesh in shared memory: 48 B per triangle, up to 4096 triangles (192 KB)
  if (accel == RT_MESH_AUTO) accel = nTri <= 32 ? RT_MESH_BRUTE : RT_MESH_BVH8;
  if (accel == RT_MESH_BRUTE && nTri > 4096) return fail(RT_ERR_UNSUPPORTED, "brute force is limited to 4096 triangles");
  HostBvh h;
  std::string err;
  if (!build_bvh8(vpos4, n_verts, indices, n_indices, accel == RT_MESH_BRUTE, h, err)) return fail(RT_ERR_INVALID, err);
  return finish_mesh(h, accel == RT_MESH_BRUTE, false, out);
}

int rt_scene_create_mesh_bvh8(const float *vpos4, uint32_t n_verts, const uint32_t *indices_bvh, uint32_t n_indices,
                              const void *nodes, uint32_t n_nodes, uint32_t node_bytes,
                              const uint32_t *prim_of_permuted, rt_scene **out) {
  HostBvh h;
  std::string err;
  if (!import_reference_bvh(vpos4, n_verts, indices_bvh, n_indices, nodes, n_nodes, node_bytes, prim_of_permuted, h,
                            err))
    return fail(RT_ERR_INVALID, err);
  return finish_mesh(h, false, true, out);
}

void rt_scene_destroy(rt_scene *scene) { scene_free(scene); }

int rt_scene_build_ms(const rt_scene *scene, float *device_ms) {
  if (!scene || !device_ms) return fail(RT_ERR_INVALID, "null argument");
  *device_ms = scene->build_ms;
  return RT_OK;
}

int rt_scene_mesh_stats(const rt_scene *scene, uint32_t *n_nodes, uint32_t *n_tris, float *device_build_ms) {
  if (!scene || scene->kind != RT_SCENE_MESH) return fail(RT_ERR_INVALID, "not a mesh scene");
  if (n_nodes) *n_nodes = scene->dev.mesh.n_nodes;
  if (n_tris) *n_tris = scene->dev.mesh.n_tris;
  if (device_build_ms) *device_build_ms = scene->build_ms;
  return RT_OK;
}

int rt_scene_kind_of(const rt_scene *scene) { return scene ? scene->kind : RT_ERR_INVALID; }
int rt_scene_device(const rt_scene *scene) { return scene ? scene->device : RT_ERR_INVALID; }

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

// per-scene-type default scheduling, chosen by measurement (profiles/r1_tune.txt): the octree
// descent has the most uneven per-warp cost and gains from persistent warps pulling tiles; the grid marcher, the BVH
// traversal (with 128-thread blocks) and the tiny brute-force scenes are as fast or faster as plain tiles.
int resolve_sched(const rt_scene *s) {
  int m = g_sched.load();
  if (m != RT_SCHED_AUTO) return m;
  if (s->kind == RT_SCENE_OCTREE) return RT_SCHED_PERSISTENT;
  return RT_SCHED_TILES;
}

int render_on_stream(const rt_scene *scene, const FrameDev &fd, uint32_t *color, const float *t_in, float *t_out,
                     uint32_t *prim, rt_counters *counters_dev, cudaStream_t st, int force_stage = -1) {
  RenderArgs a;
  a.scene = scene->dev;
  a.frame = fd;
  a.color = color;
  a.t_in = t_in;
  a.t_out = t_out;
  a.prim = prim;
  a.counters = reinterpret_cast<unsigned long long *>(counters_dev);
  rt_scene *ms = const_cast<rt_scene *>(scene);
  a.tile_counter = scene->tile_counters + (ms->next_counter.fetch_add(1) % N_TILE_COUNTERS);
  a.tile_block = g_env_tile_block;
  // Peer framebuffer (multi-GPU: rank 0's memory mapped over CUDA IPC): stage tiles in shared memory and store them as
  // 128-byte TMA bulk writes.  Needs the fused clear (whole tiles are written) and 32-byte-multiple rows.
  a.stage_stores = 0;
  int sched = resolve_sched(scene);
  // (force_stage, from rt_render_host with a pinned host framebuffer: 0 = plain stores, 1 = staged with the fused clear,
  // 2 = staged in/out, the tile's previous contents fetched by bulk loads)
  if ((fd.clear || force_stage == 2) && !prim && !counters_dev && fd.W % 8u == 0u &&
      (force_stage >= 0 ? force_stage >= 1 : g_env_stage != 0)) {
    bool peer = g_env_stage == 1 || force_stage >= 1;
    // measured (profiles/r1_multi/TABLE.txt): pays when rank 0's inbound NVLink is saturated, i.e. on 8 GPUs at 4K
    // (grid 0.139 -> 0.096 ms, bunny 0.147 -> 0.125 ms); not at 1080p (0.051 -> 0.053 ms), mixed on 4 GPUs (grid 0.131 ->
    // 0.124, bunny 0.139 -> 0.165 ms) and slower on 2 (grid 0.196 -> 0.214 ms): the 32 x 4 block with its barrier is a
    // slower kernel than the 16 x 8 one
    if (!peer && (uint64_t)fd.W * fd.H >= (3ull << 20) && fd.band_count >= 8u)
      peer = is_peer_pointer(color) && is_peer_pointer(t_out);
    a.stage_stores = peer ? (force_stage == 2 ? 2 : 1) : 0;
    if (peer && force_stage < 0) g_staged_launches.fetch_add(1);  // counts the multi-GPU path only
    // the staged stores live in the tiles kernel: for a peer framebuffer the octree's default (persistent, ~1 % faster
    // on one GPU) gives way to it
    if (peer && g_sched.load() == RT_SCHED_AUTO) sched = RT_SCHED_TILES;
  }
  // RT_SCHED_COMPACT thresholds: best point of the sweep in profiles/r1_tune.txt (RT_OCT_REFILL / RT_OCT_BATCH override)
  a.oct_refill_min = g_env_oct_refill > 0 ? g_env_oct_refill : 32;
  a.oct_leaf_batch = g_env_oct_batch > 0 ? g_env_oct_batch : 6;
  cudaError_t e = (g_math.load() == RT_MATH_FAST) ? fast::launch_render(a, sched, scene->sm_count, st)
                                                  : exact::launch_render(a, sched, scene->sm_count, st);
  if (e != cudaSuccess) return cuda_fail(e, "render launch");
  return RT_OK;
}

// Device-side address of a pinned (page-locked, mapped) host buffer, or null for pageable / device memory.
void *mapped_host_pointer(const void *p) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  return at.type == cudaMemoryTypeHost ? at.devicePointer : nullptr;
}

int ensure_host_buffers(rt_scene *s, size_t pixels) {
  if (!s->h_cnt) {  // created last: a call that failed half-way is completed by the next one
    if (!s->h_stream) RT_CUDA(cudaStreamCreateWithFlags(&s->h_stream, cudaStreamNonBlocking));
    if (!s->h_up) RT_CUDA(cudaStreamCreateWithFlags(&s->h_up, cudaStreamNonBlocking));
    if (!s->h_down) RT_CUDA(cudaStreamCreateWithFlags(&s->h_down, cudaStreamNonBlocking));
    for (auto &e : s->ev)
      if (!e) RT_CUDA(cudaEventCreate(&e));
    for (auto &e : s->ev_up)
      if (!e) RT_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (auto &e : s->ev_k)
      if (!e) RT_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    RT_CUDA(cudaMalloc(&s->h_cnt, sizeof(rt_counters)));
  }
  if (pixels > s->h_pixels) {
    if (s->h_color) cudaFree(s->h_color);
    if (s->h_t) cudaFree(s->h_t);
    if (s->h_prim) cudaFree(s->h_prim);
    s->h_color = nullptr; s->h_t = nullptr; s->h_prim = nullptr; s->h_pixels = 0;
    RT_CUDA(cudaMalloc(&s->h_color, pixels * 4));
    RT_CUDA(cudaMalloc(&s->h_t, pixels * 4));
    RT_CUDA(cudaMalloc(&s->h_prim, pixels * 4));
    s->h_pixels = pixels;
  }
  return RT_OK;
}

}  // namespace

extern "C" {

int rt_render_device(const rt_scene *scene, const rt_frame *frame, uint32_t *color, const float *t_in, float *t_out,
                     uint32_t *prim, rt_counters *counters_dev, void *stream) {
  if (!scene) return fail(RT_ERR_INVALID, "scene is null");
  if (!color || !t_out) return fail(RT_ERR_INVALID, "color / t_out must not be null");
  FrameDev fd;
  int rc = make_frame(frame, fd);
  if (rc) return rc;
  DeviceGuard g(scene->device);
  if (!g.ok) return fail(RT_ERR_CUDA, "cannot select the scene's device");
  return render_on_stream(scene, fd, color, t_in, t_out, prim, counters_dev, static_cast<cudaStream_t>(stream));
}

int rt_render_host(const rt_scene *scene_c, const rt_frame *frame, uint32_t *color, float *t_inout, uint32_t *prim,
                   float *device_ms, float *total_ms, rt_counters *counters) {
  if (!scene_c) return fail(RT_ERR_INVALID, "scene is null");
  if (!color || !t_inout) return fail(RT_ERR_INVALID, "color / t must not be null");
  FrameDev fd;
  int rc = make_frame(frame, fd);
  if (rc) return rc;
  rt_scene *s = const_cast<rt_scene *>(scene_c);
  DeviceGuard g(s->device);
  if (!g.ok) return fail(RT_ERR_CUDA, "cannot select the scene's device");
  std::lock_guard<std::mutex> lock(s->host_mu);
  const size_t pixels = (size_t)fd.W * fd.H, bytes = pixels * 4;
  rc = ensure_host_buffers(s, pixels);
  if (rc) return rc;
  cudaStream_t st = s->h_stream;
  // An error return must not leave copies into the caller's buffers in flight: drain the three streams on the way out.
  struct Drain {
    rt_scene *s;
    bool armed = true;
    ~Drain() {
      if (!armed) return;
      cudaStreamSynchronize(s->h_up);
      cudaStreamSynchronize(s->h_stream);
      cudaStreamSynchronize(s->h_down);
    }
  } drain{s};
  // Pinned host framebuffer, whole frame overwritten, nothing else asked for: the kernel stores its tiles straight into
  // the caller's buffers over PCIe (no device framebuffer, no copy engine, one launch).
  // (mode 2 needs the launch to be the staged tiles kernel: primary rays, not the brute-force mesh, tiles scheduling;
  // everything else keeps the copies, which beat plain 32-byte stores over PCIe)
  const bool secondary = fd.mode == RT_SHADE_LAMBERT && (fd.shadows || fd.reflections);
  const int sched_now = g_sched.load();
  const bool staged_ok = !secondary && !(s->kind == RT_SCENE_MESH && s->dev.mesh.brute) &&
                         (sched_now == RT_SCHED_AUTO || sched_now == RT_SCHED_TILES);
  // The in/out contract (clear = 0) takes the same route in mode 2: the kernel fetches each tile's previous colour and
  // depth from the host buffers with bulk loads and sends the merged tile back.
  const bool inout = !fd.clear;
  if (g_env_zero_copy != 0 && (inout ? (g_env_zero_copy == 2 && staged_ok) : (g_env_zero_copy == 1 || staged_ok)) &&
      fd.band_count <= 1 && !prim && !counters && fd.W % 8u == 0u) {
    uint32_t *dc = static_cast<uint32_t *>(mapped_host_pointer(color));
    float *dt = static_cast<float *>(mapped_host_pointer(t_inout));
    if (dc && dt) {
      RT_CUDA(cudaEventRecord(s->ev[0], st));
      rc = render_on_stream(s, fd, dc, inout ? dt : nullptr, dt, nullptr, nullptr, st,
                            inout ? 2 : (g_env_zero_copy >= 2 ? 1 : 0));
      if (rc) return rc;
      RT_CUDA(cudaEventRecord(s->ev[3], st));
      RT_CUDA(cudaStreamSynchronize(st));
      float ms = 0.0f;
      if (device_ms || total_ms) RT_CUDA(cudaEventElapsedTime(&ms, s->ev[0], s->ev[3]));
      if (device_ms) *device_ms = ms;
      if (total_ms) *total_ms = ms;
      drain.armed = false;
      return RT_OK;
    }
  }
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
      uint32_t rows = std::max<uint32_t>(16u, fd.H / 40u), r = 0;
      n_chunks = 0;
      while (r < fd.H && n_chunks < 15) {
        r = std::min(fd.H, r + rows);
        if (fd.H - r < rows) r = fd.H;  // fold a short remainder into this chunk
        bounds[++n_chunks] = r;
        rows *= 2;
      }
      if (r < fd.H) bounds[++n_chunks] = fd.H;
    }
  }
  RT_CUDA(cudaEventRecord(s->ev[0], st));
  if (prim && need_upload) RT_CUDA(cudaMemsetAsync(s->h_prim, 0xFF, bytes, st));  // rows not rendered: MISS
  if (counters) RT_CUDA(cudaMemsetAsync(s->h_cnt, 0, sizeof(rt_counters), st));
  RT_CUDA(cudaEventRecord(s->ev[1], st));
  RT_CUDA(cudaStreamWaitEvent(s->h_up, s->ev[0], 0));
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
      RT_CUDA(cudaMemcpyAsync(s->h_t + uoff, t_inout + uoff, ubytes, cudaMemcpyHostToDevice, s->h_up));
      RT_CUDA(cudaMemcpyAsync(s->h_color + uoff, color + uoff, ubytes, cudaMemcpyHostToDevice, s->h_up));
      RT_CUDA(cudaEventRecord(s->ev_up[k], s->h_up));
      RT_CUDA(cudaStreamWaitEvent(st, s->ev_up[k], 0));
    }
    rc = render_on_stream(s, fk, s->h_color, s->h_t, s->h_t, prim ? s->h_prim : nullptr, counters ? s->h_cnt : nullptr, st);
    if (rc) return rc;
    RT_CUDA(cudaEventRecord(s->ev_k[k], st));
    RT_CUDA(cudaStreamWaitEvent(s->h_down, s->ev_k[k], 0));
    const size_t doff = n_chunks > 1 ? off : 0, dbytes = n_chunks > 1 ? cbytes : bytes;
    RT_CUDA(cudaMemcpyAsync(color + doff, s->h_color + doff, dbytes, cudaMemcpyDeviceToHost, s->h_down));
    RT_CUDA(cudaMemcpyAsync(t_inout + doff, s->h_t + doff, dbytes, cudaMemcpyDeviceToHost, s->h_down));
    if (prim) RT_CUDA(cudaMemcpyAsync(prim + doff, s->h_prim + doff, dbytes, cudaMemcpyDeviceToHost, s->h_down));
  }
  RT_CUDA(cudaEventRecord(s->ev[2], st));
  if (counters) RT_CUDA(cudaMemcpyAsync(counters, s->h_cnt, sizeof(rt_counters), cudaMemcpyDeviceToHost, st));
  RT_CUDA(cudaStreamSynchronize(st));
  RT_CUDA(cudaEventRecord(s->ev[3], s->h_down));
  RT_CUDA(cudaStreamSynchronize(s->h_down));
  if (device_ms) RT_CUDA(cudaEventElapsedTime(device_ms, s->ev[1], s->ev[2]));
  if (total_ms) RT_CUDA(cudaEventElapsedTime(total_ms, s->ev[0], s->ev[3]));
  drain.armed = false;  // both syncs above succeeded: nothing is pending
  return RT_OK;
}

int rt_intersect_host(const rt_scene *scene, const float *origins, const float *dirs, uint32_t n, float t_near,
                      float t_far, int has_plane, float plane_y, uint8_t *hit, float *t, float *normal,
                      uint32_t *prim) {
  if (!scene) return fail(RT_ERR_INVALID, "scene is null");
  if (n == 0) return RT_OK;
  if (!origins || !dirs) return fail(RT_ERR_INVALID, "origins / dirs must not be null");
  DeviceGuard g(scene->device);
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
  a.scene = scene->dev;
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
