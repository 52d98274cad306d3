// TEST INFRASTRUCTURE — not part of the product path.
//
// Scalar C++ restatement of the three ISPC exports of the reference's src/ray_pack.ispc that
// its C++ calls (the ispc compiler is absent from this image).  Each lane of an ISPC `foreach`
// becomes one iteration of a plain loop; operation order follows the ISPC source token by token
// so that, built with -ffp-contract=off, every lane result is the correctly-rounded IEEE value
// of the expression as written.
//
// ISPC float min/max lower to x86 minps/maxps: min(a,b) = a<b ? a : b and max(a,b) = a>b ? a : b
// (the SECOND operand is returned when either is NaN) — restated as such, SURVEY T11.
#include "ray_pack_ispc.h"

#include <cmath>

namespace {

inline float imin(float a, float b) { return a < b ? a : b; }
inline float imax(float a, float b) { return a > b ? a : b; }

struct V3 {
  float x, y, z;
};
inline V3 sub(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// ray_pack.ispc:12-20: shuffle_yzx(a*b_yzx - a_yzx*b)
inline V3 cross(V3 a, V3 b) {
  V3 tmp = {a.x * b.y - a.y * b.x, a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z};
  return {tmp.y, tmp.z, tmp.x};
}
// ray_pack.ispc:5-10: ray / length  (a division per component, not a multiply by 1/length)
inline V3 normalize(V3 r) {
  float length = std::sqrt(r.x * r.x + r.y * r.y + r.z * r.z);
  return {r.x / length, r.y / length, r.z / length};
}

thread_local RefTriRecorder g_recorder = nullptr;
thread_local void *g_recorder_user = nullptr;

} // namespace

extern "C" void ref_standin_set_recorder(RefTriRecorder fn, void *user) {
  g_recorder = fn;
  g_recorder_user = user;
}

namespace ispc {

// ray_pack.ispc:220-239
extern "C" void divide_box_8(const Box *box, Box8 *result) {
  const float bminx = box->boxMin[0], bminy = box->boxMin[1], bminz = box->boxMin[2];
  const float bmaxx = box->boxMax[0], bmaxy = box->boxMax[1], bmaxz = box->boxMax[2];
  const float cx = (bminx + bmaxx) / 2.0f, cy = (bminy + bmaxy) / 2.0f, cz = (bminz + bmaxz) / 2.0f;
  const float dx = cx - bminx, dy = cy - bminy, dz = cz - bminz;
  for (int id = 0; id < 8; ++id) {
    int x = id >> 2;
    int y = (id & 3) >> 1;
    int z = id & 1;
    result->xMin[id] = (x == 0) ? bminx : cx;
    result->yMin[id] = (y == 0) ? bminy : cy;
    result->zMin[id] = (z == 0) ? bminz : cz;
    result->xMax[id] = result->xMin[id] + dx;
    result->yMax[id] = result->yMin[id] + dy;
    result->zMax[id] = result->zMin[id] + dz;
  }
}

// ray_pack.ispc:241-273
extern "C" void intersect_box_8(const Box8 *pBoxes, const float orig[3],
                                const float dirInverted[3], float tNear, float tFar,
                                float pResults[8]) {
  const float ox = orig[0], oy = orig[1], oz = orig[2];
  const float ix = dirInverted[0], iy = dirInverted[1], iz = dirInverted[2];
  for (int b = 0; b < 8; ++b) {
    float t1x = (pBoxes->xMin[b] - ox) * ix;
    float t1y = (pBoxes->yMin[b] - oy) * iy;
    float t1z = (pBoxes->zMin[b] - oz) * iz;
    float t2x = (pBoxes->xMax[b] - ox) * ix;
    float t2y = (pBoxes->yMax[b] - oy) * iy;
    float t2z = (pBoxes->zMax[b] - oz) * iz;

    float tMin3x = imin(t1x, t2x), tMin3y = imin(t1y, t2y), tMin3z = imin(t1z, t2z);
    float tMax3x = imax(t1x, t2x), tMax3y = imax(t1y, t2y), tMax3z = imax(t1z, t2z);

    float tMin = imax(tMin3x, imax(tMin3y, tMin3z));
    float tMax = imin(tMax3x, imin(tMax3y, tMax3z));

    tMin = imax(tMin, tNear);
    tMax = imin(tMax, tFar);

    if (tMax < 0 || tMin > tMax) {
      pResults[b] = -1.0f;
    } else {
      pResults[b] = tMin;
    }
  }
}

// ray_pack.ispc:289-310 with triangle_intersection (uniform ray x varying triangle) :132-165
extern "C" void intersect_1_ray_8_triangles(const Triangle8 *pTriangles, const float orig[3],
                                            const float dir[3], HitInfo8 *pResults) {
  const V3 rayPos = {orig[0], orig[1], orig[2]};
  const V3 rayDir = {dir[0], dir[1], dir[2]};
  for (int i = 0; i < 8; ++i) {
    V3 v0 = {pTriangles->v0.x[i], pTriangles->v0.y[i], pTriangles->v0.z[i]};
    V3 v1 = {pTriangles->v1.x[i], pTriangles->v1.y[i], pTriangles->v1.z[i]};
    V3 v2 = {pTriangles->v2.x[i], pTriangles->v2.y[i], pTriangles->v2.z[i]};

    uint32_t hitten = 0;
    float t = -1.0f;
    V3 e1 = sub(v1, v0);
    V3 e2 = sub(v2, v0);
    V3 norm = normalize(cross(e1, e2));
    V3 pvec = cross(rayDir, e2);
    float det = dot(e1, pvec);
    do {
      if (det < 1e-8f && det > -1e-8f)
        break;
      float inv_det = 1 / det;
      V3 tvec = sub(rayPos, v0);
      float u = dot(tvec, pvec) * inv_det;
      if (u < 0.0f || u > 1.0f)
        break;
      V3 qvec = cross(tvec, e1);
      float v = dot(rayDir, qvec) * inv_det;
      if (v < 0.0f || u + v > 1.0f)
        break;
      t = dot(e2, qvec) * inv_det;
      hitten = 1;
    } while (false);

    pResults->hitten[i] = hitten;
    pResults->t[i] = t;
    pResults->norm_x[i] = norm.x;
    pResults->norm_y[i] = norm.y;
    pResults->norm_z[i] = norm.z;

    if (hitten && g_recorder) {
      RefTriCandidate c = {t, {v0.x, v0.y, v0.z}, {v1.x, v1.y, v1.z}, {v2.x, v2.y, v2.z}};
      g_recorder(&c, g_recorder_user);
    }
  }
}

} // namespace ispc
