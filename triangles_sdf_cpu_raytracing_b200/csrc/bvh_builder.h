// Host-side construction of the GPU's 8-wide BVH. Replaces BVHBuilder::perform (triangles_raytracing.cpp:227-258)
// on the product path: closest-hit results do not depend on the tree, so this is our own builder
// (binned SAH binary tree collapsed to 8-wide, leaves folded into their parents), not a restatement of the
// reference's sweep builder.  import_reference_bvh() converts a tree the reference itself built.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "rt_scene.h"

namespace rt {

struct HostBvh {
  std::vector<BvhNode> nodes;  // nodes[0] = root, always an inner node
  std::vector<TriDev> tris;    // leaf triangles in tree order, prim id in v0.w
  float bbox_min[3], bbox_max[3];
  int stack_need = 0;          // worst-case traversal stack entries
  uint32_t brute = BRUTE_NONE; // BRUTE_*: no tree, `tris` in the original order (rt_scene.h)
};

// vpos4: n_verts x (x,y,z,w); indices: 3 per triangle. Triangles are divided by w exactly as
// triangles_raytracing.cpp:307-309 does per visit.  Returns false and sets err on bad input.
// brute = BRUTE_NONE builds the tree; BRUTE_BOXED / BRUTE_ROOT_LEAF only pack the triangles.
bool build_bvh8(const float *vpos4, uint32_t n_verts, const uint32_t *indices, uint32_t n_indices, uint32_t brute,
                HostBvh &out, std::string &err);

// nodes: n_nodes images of the reference's BVH8Node (triangles_raytracing.hpp:10-36): Box8 (192 B) | realCount u32 |
// offset u32 (union with BVHLeafInfo {startIndex, count}) | bool isLeaf at byte 200; node_bytes = 204.
bool import_reference_bvh(const float *vpos4, uint32_t n_verts, const uint32_t *indices_bvh, uint32_t n_indices,
                          const void *nodes, uint32_t n_nodes, uint32_t node_bytes, const uint32_t *prim_of_permuted,
                          HostBvh &out, std::string &err);

// The same tree built on the device (bvh_gpu.cu): linear BVH over 63-bit Morton codes, collapsed to 8-wide.
// nodes / tris are cudaMalloc'ed on the current device and owned by the caller on success.
struct DeviceBvh {
  BvhNode *nodes = nullptr;
  TriDev *tris = nullptr;
  uint32_t n_nodes = 0, n_tris = 0;
  float bbox_min[3] = {0, 0, 0}, bbox_max[3] = {0, 0, 0};
  int stack_need = 0;
  int levels = 0;        // depth of the 8-wide tree
  float build_ms = 0.0f; // device time of the build kernels (excluding the upload of vertices / indices)
};
bool build_bvh8_device(const float *vpos4, uint32_t n_verts, const uint32_t *indices, uint32_t n_indices, DeviceBvh &out,
                       std::string &err);

}  // namespace rt
