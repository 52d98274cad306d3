// On-device resource ingestion (ingest.cu): raw .octree / .grid images -> traversal layouts, validated on the GPU.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "rt_scene.h"

namespace rt {

enum { INGEST_OK = 0, INGEST_INVALID = 1, INGEST_CUDA = 2 };

// nodes36: n raw 36-byte SDFOctreeNode records (host memory).  On success link / values are cudaMalloc'ed arrays owned
// by the caller (rt_scene.h: OctreeDev), depth is the deepest level that holds a node, ms the device time.
// has_mask: inner link words carry the kept-children mask (trees of fewer than 2^24 nodes).
int ingest_octree_device(const void *nodes36, uint32_t n, uint32_t **link, float4 **values, uint32_t *depth,
                         uint32_t *has_mask, float *ms, std::string &err);

// values: the linear grid (host memory) -> brick-major device array of outCount floats (rt_scene.h: GridDev, bricks = 1)
int ingest_grid_bricks_device(const float *values, uint32_t sx, uint32_t sy, uint32_t sz, float **out, size_t *outCount,
                              std::string &err);

}  // namespace rt
