// TEST INFRASTRUCTURE — not part of the product path.
//
// Stand-in for the header the ISPC compiler would generate from the reference's
// src/ray_pack.ispc (ispc is not installed in this image and cannot be).
// It declares, in namespace ispc, the struct layouts the reference's C++ reinterpret_casts
// into and the three exports the reference actually calls:
//   Box8          <- ray_pack.ispc:204-212  (same layout as Box8 in triangles_raytracing.hpp:10-17)
//   Box           <- ray_pack.ispc:214-218  (same layout as LiteMath::BBox3f, 2 x float3)
//   Point8/Triangle8 <- ray_pack.ispc:275-287
//   HitInfo8      <- ray_pack.ispc:81-88
//   divide_box_8                 <- ray_pack.ispc:220-239, called at octree_raytracing.cpp:176
//   intersect_box_8              <- ray_pack.ispc:241-273, called at triangles_raytracing.cpp:274, octree_raytracing.cpp:180
//   intersect_1_ray_8_triangles  <- ray_pack.ispc:289-310, called at triangles_raytracing.cpp:322
// intersect_triangles_8 (ray_pack.ispc:167-202) has no caller in the reference and is not declared.
#pragma once
#include <cstdint>

namespace ispc {

struct Box8 {
  float xMin[8];
  float yMin[8];
  float zMin[8];
  float xMax[8];
  float yMax[8];
  float zMax[8];
};

struct Box {
  float boxMin[3];
  float boxMax[3];
};

struct Point8 {
  float x[8];
  float y[8];
  float z[8];
};

struct Triangle8 {
  Point8 v0;
  Point8 v1;
  Point8 v2;
};

struct HitInfo8 {
  uint32_t hitten[8];
  float t[8];
  float norm_x[8];
  float norm_y[8];
  float norm_z[8];
};

extern "C" {
void divide_box_8(const Box *box, Box8 *result);
void intersect_box_8(const Box8 *pBoxes, const float orig[3],
                     const float dirInverted[3], float tNear, float tFar,
                     float pResults[8]);
void intersect_1_ray_8_triangles(const Triangle8 *pTriangles,
                                 const float orig[3], const float dir[3],
                                 HitInfo8 *pResults);
}

} // namespace ispc

// Test-only hook (not in the reference): when a recorder is installed on the calling thread,
// intersect_1_ray_8_triangles appends every lane it reports as hit, in call order, so that the
// oracle driver can name the winning triangle of a ray (the reference's HitInfo has no primitive
// id). See oracle/ref_driver.cpp: ref_render_prim.
struct RefTriCandidate {
  float t;
  float v0[3], v1[3], v2[3];
};
typedef void (*RefTriRecorder)(const RefTriCandidate *c, void *user);
extern "C" void ref_standin_set_recorder(RefTriRecorder fn, void *user);
