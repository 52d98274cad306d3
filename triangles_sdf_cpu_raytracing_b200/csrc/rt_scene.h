// Device-visible scene descriptors and launch parameters (plain structs shared by rt_api.cu and the kernels).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/rt_b200.h"

namespace rt {

// ---- dense SDF grid (replaces SDFGrid, grid_raytracing.hpp:10-21) ---------------------------------------
// LINEAR: values[(x*sy + y)*sz + z], the reference's own order.
// BRICKS: the grid is cut into 8x8x8-sample bricks (bx,by,bz), brick-major with the same x-major/z-fastest
//         order inside a brick: a 2 KB brick = 16 lines of 128 B, so the 8 corners of a trilinear cell touch
//         1-2 lines when they fall inside one brick instead of 4 far-apart lines.
struct GridDev {
  const float *values;
  uint32_t sx, sy, sz;
  uint32_t bricks;              // 0 linear, 1 bricked
  uint32_t stride_x, stride_y;  // linear: sy*sz, sz;  bricked: nby*nbz*512, nbz*512 (elements)
  float fx, fy, fz;             // (float)(size - 1), grid_raytracing.cpp:9-11
  float hfx, hfy, hfz;          // (float)(size - 1) / 2, exact
};

// ---- SDF octree (replaces SDFOctree, octree_raytracing.hpp:8-47) ----------------------------------------
// The 36-byte disk node is split at upload: `link[node]` holds the traversal word, `values[2*node..]` the 8
// corner distances as two 16-byte-aligned float4 (only ever read for surface leaves).
//   link == OCT_LEAF_CULLED : leaf that intersectLeaf rejects from its values alone (isEmpty
//                             octree_raytracing.hpp:12-17, or all >= 1e-4 octree_raytracing.cpp:132-135)
//   link == OCT_LEAF_SURFACE: leaf that must be marched
//   otherwise               : inner node: childrenOffset, and when has_mask (fewer than 2^24 nodes) bits 24..31 say which
//                             of its 8 children are worth a visit (bit c clear: child c is an OCT_LEAF_CULLED leaf; the
//                             traversal counts it as visited without loading its record)
constexpr uint32_t OCT_LEAF_SURFACE = 0u;
constexpr uint32_t OCT_LEAF_CULLED = 0xFFFFFFFFu;
constexpr uint32_t OCT_OFFSET_MASK = 0x00FFFFFFu;
constexpr uint32_t OCT_MASK_MAX_NODES = 1u << 24;  // offsets must leave the top byte free (offset + 7 < n_nodes)
constexpr int OCT_MAX_DEPTH = 21;  // child boxes stay exactly representable; 3 path bits per level in 64
struct OctreeDev {
  const uint32_t *link;
  const float4 *values;
  uint32_t n_nodes;
  uint32_t depth;
  uint32_t root_inner;  // 1: node 0 has children (always, except for a one-node file)
  uint32_t has_mask;    // 1: inner link words carry the kept-children mask in bits 24..31
};

// ---- triangle mesh ------------------------------------------------------------------------------------
// 8-wide BVH node: child boxes SoA as in Box8 (triangles_raytracing.hpp:10-17) followed by 8 child words.
//   child == BVH_EMPTY           : unused lane (the reference leaves garbage there)
//   child &  BVH_LEAF_BIT        : leaf: bits 0..26 first triangle, bits 27..30 triangle count - 1
//   otherwise                    : index of an inner node
// 224 bytes, 32-byte aligned: 14 x LDG.128 per visit.
constexpr uint32_t BVH_EMPTY = 0xFFFFFFFFu;
constexpr uint32_t BVH_LEAF_BIT = 0x80000000u;
constexpr int BVH_STACK = 128;  // scene creation rejects trees that could need more
struct alignas(32) BvhNode {
  float lo[3][8];  // xMin[8], yMin[8], zMin[8]
  float hi[3][8];  // xMax[8], yMax[8], zMax[8]
  uint32_t child[8];
};
static_assert(sizeof(BvhNode) == 224, "BvhNode layout");

// Leaf triangles pre-gathered in BVH order (removes the per-visit gather of triangles_raytracing.cpp:299-320):
// v0.xyz | prim id, v1.xyz | 0, v2.xyz | 0, already divided by w (same IEEE division as :307-309).
struct alignas(16) TriDev {
  float4 v0, v1, v2;
};

// Brute-force scenes (every ray tests every triangle, no tree):
//   BRUTE_ROOT_LEAF: what the reference itself does for a mesh of <= 8 triangles.  createNode makes node 0 a leaf
//                    (triangles_raytracing.cpp:175-186: `end - start <= 24` indices and no split), and traverseNode(0)
//                    on a leaf tests its triangles with NO box test and NO range check (:292-331).  Exact in every view.
//   BRUTE_BOXED    : 9+ triangles, explicit RT_MESH_BRUTE only.  The ray is first culled against the mesh bounding box
//                    with the reference's slab test and a triangle is accepted iff t <= tFar.  Equals the reference
//                    wherever its answer does not depend on its tree (eye outside the mesh, cleared framebuffer), and
//                    on cube.obj in every view; elsewhere it is one of many tree-dependent answers.
constexpr uint32_t BRUTE_NONE = 0u, BRUTE_BOXED = 1u, BRUTE_ROOT_LEAF = 2u;
constexpr uint32_t ROOT_LEAF_MAX_TRIS = 8u;  // triangles_raytracing.cpp:120-122,175-186

struct MeshDev {
  const BvhNode *nodes;
  const TriDev *tris;
  uint32_t n_nodes, n_tris;
  uint32_t brute;  // BRUTE_*: no BVH, all triangles are tested from shared memory
  uint32_t literal_order;  // 1: tree imported from the reference: literal slab test + sort8 visit order
};

struct SceneDev {
  int32_t kind;  // rt_scene_kind
  GridDev grid;
  OctreeDev oct;
  MeshDev mesh;
};

// ---- per-launch parameters ------------------------------------------------------------------------------
struct FrameDev {
  float ray_pos[3];
  float view_inv[16];
  float proj_inv[16];
  uint32_t W, H;
  uint32_t band_rows, band_count, band_index;
  uint32_t rows_here;  // number of memory rows this call renders
  uint32_t row0;       // whole-frame calls (band_count <= 1): first memory row (row chunks of rt_render_host)
  int32_t mode;
  float light[3];
  int32_t shadows, reflections, has_plane;
  float plane_y;
  float t_near, t_far;
  int32_t clear;
};

// Heavy-first block scheduling (tiles kernels).  A frame cannot finish before its longest ray does: one grazing ray
// marching 150 dependent, cache-missing steps takes 20-60 us however many SMs are idle, and launched mid-frame it IS the
// tail.  That fixed cost is what a share of a frame on one of 8 GPUs, or a 1080p frame, mostly consists of (4K 513^3
// grid: 0.366 ms on one GPU, 0.103 ms for an eighth of it).  Successive frames of a viewer are nearly the same picture,
// so every launch records which blocks ran long and the next launch of the same scene / frame geometry starts those
// first: the grid carries extra leading rows of blocks that pick their tile from the previous launch's list; the
// regular blocks skip tiles already taken that way.  Order only: every pixel is still rendered by exactly the same code
// exactly once (both sides claim a tile with one atomic exchange; a stale list only costs the order).
constexpr uint32_t HEAVY_CAP = 2048;
struct HeavySched {
  const uint32_t *prev;        // tile ids (default-order block index) that ran long in the previous launch
  const uint32_t *prev_count;  // how many of them (device memory; <= HEAVY_CAP are used)
  uint32_t *next, *next_count; // this launch's own list
  uint32_t *done;              // [tile] == tag: this launch took the tile from `prev`
  uint32_t *seen;              // [tile] == tag: this launch has already reported the tile
  uint32_t tag;                // unique per launch
  uint32_t cycles;             // a block is heavy when one of its warps ran longer than this; 0 = scheduling off
  uint32_t extra_rows;         // leading grid rows that are heavy slots (set by the launcher)
};

struct RenderArgs {
  SceneDev scene;
  FrameDev frame;
  HeavySched heavy;
  uint32_t *color;
  const float *t_in;
  float *t_out;
  uint32_t *prim;
  unsigned long long *counters;  // rt_counters as 9 x u64, or null
  uint32_t *tile_counter;        // persistent scheduling: next tile to hand out
  int32_t stage_stores;          // tiles: stage the tile in shared memory, store it with TMA (1: fused clear, 2: in/out)
  unsigned long long *xfer;      // in/out staged tiles: running byte counts {fetched, stored} (rt_scene_transfer_bytes), or null
  int32_t tile_block;            // RT_SCHED_TILES: threads per block (32, 64, 128, 256)
  int32_t oct_refill_min, oct_leaf_batch;  // RT_SCHED_COMPACT: idle lanes needed for a refill, leaf lanes for a march
};

struct RayArgs {
  SceneDev scene;
  const float *origins, *dirs;
  uint32_t n;
  float t_near, t_far;
  int32_t has_plane;
  float plane_y;
  uint8_t *hit;
  float *t;
  float *normal;
  uint32_t *prim;
};

// launchers, one set per arithmetic flavour (rt_kernels.cuh compiled twice)
#define RT_DECLARE_LAUNCHERS(NS)                                                         \
  namespace NS {                                                                         \
  cudaError_t launch_render(const RenderArgs &a, int sched, int sm_count, cudaStream_t s); \
  cudaError_t launch_rays(const RayArgs &a, cudaStream_t s);                             \
  }
RT_DECLARE_LAUNCHERS(exact)
RT_DECLARE_LAUNCHERS(fast)

}  // namespace rt
