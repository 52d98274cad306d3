/* TEST INFRASTRUCTURE — CPU restatement ("port") of the reference's primary-ray hot path.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may build, load or call this
 * file. It is the checker for the CUDA path, never the thing measured or shipped.
 *
 * Parity status: PINNED against the reference itself — the reference ships no tests or golden
 * vectors, so this port is validated bit for bit (tests/test_oracle_*.py) against
 *   (a) oracle/_ref/libref_oracle.so = the reference's own unmodified C++ compiled by
 *       oracle/Makefile with -ffp-contract=off, where /root/reference exists, and
 *   (b) the committed fixtures under tests/golden/ that (a) generated (tests/golden/make_golden.py).
 *
 * Built with -ffp-contract=off: every expression below is evaluated as written, one correctly
 * rounded IEEE-754 binary32 operation per operator, in the reference's operator order.
 * All file:line citations are into /root/reference.
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

typedef struct { float x, y, z; } v3;

/* std::min / std::max (used by LiteMath: LiteMath.h:793-795): return the FIRST argument unless the
 * comparison says otherwise; ISPC min/max (ray_pack.ispc:258-265) lower to minps/maxps: SECOND
 * argument on NaN. SURVEY T11. */
static inline float std_min(float a, float b) { return (b < a) ? b : a; }
static inline float std_max(float a, float b) { return (a < b) ? b : a; }
static inline float ispc_min(float a, float b) { return a < b ? a : b; }
static inline float ispc_max(float a, float b) { return a > b ? a : b; }

static inline v3 v3_add(v3 a, v3 b) { v3 r = {a.x + b.x, a.y + b.y, a.z + b.z}; return r; }
static inline v3 v3_sub(v3 a, v3 b) { v3 r = {a.x - b.x, a.y - b.y, a.z - b.z}; return r; }
static inline v3 v3_scale(v3 a, float s) { v3 r = {a.x * s, a.y * s, a.z * s}; return r; }
static inline float v3_dot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
/* LiteMath.h:835-840 and ray_pack.ispc:12-20: shuffle_yzx(a*b_yzx - a_yzx*b) */
static inline v3 v3_cross(v3 a, v3 b) {
  v3 tmp = {a.x * b.y - a.y * b.x, a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z};
  v3 r = {tmp.y, tmp.z, tmp.x};
  return r;
}
/* LiteMath.h:814: a * (1/length) */
static inline v3 lm_normalize(v3 a) {
  float lenInv = 1.0f / sqrtf(v3_dot(a, a));
  return v3_scale(a, lenInv);
}
/* ray_pack.ispc:5-10: ray / length */
static inline v3 ispc_normalize(v3 r) {
  float length = sqrtf(r.x * r.x + r.y * r.y + r.z * r.z);
  v3 o = {r.x / length, r.y / length, r.z / length};
  return o;
}

/* ------------------------------------------------------------------ public structs (ctypes mirrors in port_oracle.py) */
typedef struct {
  int32_t kind; /* 0 grid, 1 octree, 2 mesh */
  /* grid: grid_raytracing.hpp:10-15, index (x*sy + y)*sz + z */
  const float *values;
  uint32_t sx, sy, sz;
  /* octree: octree_raytracing.hpp:8-11, n x {float values[8]; uint32 childrenOffset} = 36 B */
  const uint8_t *nodes36;
  uint32_t nOctNodes;
  /* mesh: vPos4f, BVH-ordered indices, BVH8Node array (triangles_raytracing.hpp:10-36, nodeBytes each),
   * original triangle id per BVH-order triangle (may be NULL -> BVH-order id) */
  const float *vpos4;
  const uint32_t *indices;
  const uint8_t *bvhNodes;
  uint32_t nBvhNodes;
  uint32_t bvhNodeBytes;
  const uint32_t *primOfPermuted;
  /* optional checkerboard ground plane y = planeY: SceneUnion(model, Plane({0,1,0}, planeY)), main.cpp:189-190 */
  int32_t hasPlane;
  float planeY;
  /* meshes without a BVH (mesh_brute): bounding box of the triangles' vertices (after the division by w) */
  float bboxMin[3], bboxMax[3];
} PortScene;

typedef struct {
  float rayPos[3];
  float viewInv[16]; /* column-major, LiteMath float4x4::m_col */
  float projInv[16];
  uint32_t W, H;
  int32_t mode; /* 0 Normal, 1 Lambert, 2 Color (raytracing.hpp:99) */
  float light[3];
  int32_t shadows, reflections;
} PortFrame;

/* per-call workload counters (sums over all rays, primary and secondary) */
typedef struct {
  uint64_t rays;       /* IScene::intersect calls on the model */
  uint64_t gridSdf;    /* SDFGrid::sdf(float3) evaluations */
  uint64_t octInner;   /* inner octree nodes visited */
  uint64_t octLeaf;    /* intersectLeaf calls */
  uint64_t octLeafSdf; /* nodeSDF(float3) evaluations */
  uint64_t bvhInner;   /* inner BVH8 nodes visited */
  uint64_t bvhLeaf;    /* BVH leaves visited */
  uint64_t bvhTris;    /* real triangles tested */
  uint64_t hits;
} PortCounters;

typedef struct {
  int hitten;
  float t;
  v3 normal;
  v3 albedo;
  float reflectiveness;
  uint32_t prim;
} Hit;

#define PRIM_MISS 0xFFFFFFFFu
#define PRIM_PLANE 0xFFFFFFFEu

static inline Hit hit_none(void) { /* raytracing.hpp:67-73 defaults */
  Hit h;
  h.hitten = 0;
  h.t = INFINITY;
  h.normal.x = 0.0f; h.normal.y = 1.0f; h.normal.z = 0.0f;
  h.albedo.x = 1.0f; h.albedo.y = 1.0f; h.albedo.z = 1.0f;
  h.reflectiveness = 0.0f;
  h.prim = PRIM_MISS;
  return h;
}

/* ------------------------------------------------------------------ LiteMath::BBox3f::Intersection (LiteMath.h:2005-2067) */
static void bbox_intersection(v3 bmin, v3 bmax, v3 o, v3 inv, float min_t, float max_t, float *t1, float *t2, int *face) {
  const float min_x = inv.x < 0 ? bmax.x : bmin.x;
  const float min_y = inv.y < 0 ? bmax.y : bmin.y;
  const float min_z = inv.z < 0 ? bmax.z : bmin.z;
  const float max_x = inv.x < 0 ? bmin.x : bmax.x;
  const float max_y = inv.y < 0 ? bmin.y : bmax.y;
  const float max_z = inv.z < 0 ? bmin.z : bmax.z;
  const float tmin_x = (min_x - o.x) * inv.x;
  const float tmax_x = (max_x - o.x) * inv.x * 1.00000024f;
  const float tmin_y = (min_y - o.y) * inv.y;
  const float tmax_y = (max_y - o.y) * inv.y * 1.00000024f;
  const float tmin_z = (min_z - o.z) * inv.z;
  const float tmax_z = (max_z - o.z) * inv.z * 1.00000024f;
  float tmax = std_min(tmax_z, std_min(tmax_y, std_min(tmax_x, max_t)));
  float tmin;
  int idx;
  if (tmin_x > tmin_y) {
    if (tmin_z > tmin_x) { idx = 2; tmin = std_max(tmin_z, min_t); }
    else                 { idx = 0; tmin = std_max(tmin_x, min_t); }
  } else {
    if (tmin_z > tmin_y) { idx = 2; tmin = std_max(tmin_z, min_t); }
    else                 { idx = 1; tmin = std_max(tmin_y, min_t); }
  }
  *t1 = tmin; *t2 = tmax; *face = idx;
}

/* ------------------------------------------------------------------ ISPC primitives (ray_pack.ispc) */
/* ray_pack.ispc:241-273, one lane */
static inline float slab_lane(float xMin, float yMin, float zMin, float xMax, float yMax, float zMax,
                              v3 o, v3 inv, float tNear, float tFar) {
  float t1x = (xMin - o.x) * inv.x, t1y = (yMin - o.y) * inv.y, t1z = (zMin - o.z) * inv.z;
  float t2x = (xMax - o.x) * inv.x, t2y = (yMax - o.y) * inv.y, t2z = (zMax - o.z) * inv.z;
  float mnx = ispc_min(t1x, t2x), mny = ispc_min(t1y, t2y), mnz = ispc_min(t1z, t2z);
  float mxx = ispc_max(t1x, t2x), mxy = ispc_max(t1y, t2y), mxz = ispc_max(t1z, t2z);
  float tMin = ispc_max(mnx, ispc_max(mny, mnz));
  float tMax = ispc_min(mxx, ispc_min(mxy, mxz));
  tMin = ispc_max(tMin, tNear);
  tMax = ispc_min(tMax, tFar);
  if (tMax < 0 || tMin > tMax) return -1.0f;
  return tMin;
}

/* raytracing.hpp:188-213: the 19-comparator network, compare with '>' and swap both arrays */
static void sort8(float t[8], uint32_t c[8]) {
  static const uint8_t net[19][2] = {{0,1},{2,3},{4,5},{6,7},{0,2},{1,3},{4,6},{5,7},{1,2},{5,6},
                                     {0,4},{3,7},{1,5},{2,6},{1,4},{3,6},{2,4},{3,5},{3,4}};
  for (int i = 0; i < 19; ++i) {
    int a = net[i][0], b = net[i][1];
    if (t[a] > t[b]) {
      float tf = t[a]; t[a] = t[b]; t[b] = tf;
      uint32_t tc = c[a]; c[a] = c[b]; c[b] = tc;
    }
  }
}

/* ray_pack.ispc:132-165 (uniform ray x varying triangle), without the normal (computed for the winner only:
 * it is a pure function of the triangle, ray_pack.ispc:142) */
static inline int tri_lane(v3 o, v3 d, v3 v0, v3 v1, v3 v2, float *tOut) {
  v3 e1 = v3_sub(v1, v0);
  v3 e2 = v3_sub(v2, v0);
  v3 pvec = v3_cross(d, e2);
  float det = v3_dot(e1, pvec);
  if (det < 1e-8f && det > -1e-8f) return 0;
  float inv_det = 1 / det;
  v3 tvec = v3_sub(o, v0);
  float u = v3_dot(tvec, pvec) * inv_det;
  if (u < 0.0f || u > 1.0f) return 0;
  v3 qvec = v3_cross(tvec, e1);
  float v = v3_dot(d, qvec) * inv_det;
  if (v < 0.0f || u + v > 1.0f) return 0;
  *tOut = v3_dot(e2, qvec) * inv_det;
  return 1;
}

/* ------------------------------------------------------------------ dense grid (grid_raytracing.cpp) */
static inline float grid_at(const PortScene *s, uint32_t x, uint32_t y, uint32_t z) {
  return s->values[((size_t)x * s->sy + y) * s->sz + z]; /* grid_raytracing.hpp:13-15 */
}

/* grid_raytracing.cpp:7-62; *cell receives the linear index of corner c0 */
static float grid_sdf(const PortScene *s, v3 p, uint32_t *cell, PortCounters *cnt) {
  if (cnt) cnt->gridSdf++;
  p.x = (p.x + 1.0f) / 2.0f; p.y = (p.y + 1.0f) / 2.0f; p.z = (p.z + 1.0f) / 2.0f;
  p.x *= (float)(s->sx - 1); p.y *= (float)(s->sy - 1); p.z *= (float)(s->sz - 1);
  float c0fx = floorf(p.x), c0fy = floorf(p.y), c0fz = floorf(p.z);
  float c1fx = ceilf(p.x), c1fy = ceilf(p.y), c1fz = ceilf(p.z);
  uint32_t c0x = (uint32_t)c0fx, c0y = (uint32_t)c0fy, c0z = (uint32_t)c0fz;
  uint32_t c1x = (uint32_t)c1fx, c1y = (uint32_t)c1fy, c1z = (uint32_t)c1fz;
  float px = p.x - c0fx, py = p.y - c0fy, pz = p.z - c0fz;   /* p_c0f */
  float qx = c1fx - p.x, qy = c1fy - p.y, qz = c1fz - p.z;   /* c1f_p */
  if (c1x == c0x) { px = 1.0f; qx = 0.0f; }
  if (c1y == c0y) { py = 1.0f; qy = 0.0f; }
  if (c1z == c0z) { pz = 1.0f; qz = 0.0f; }
  float v[8];
  v[0] = grid_at(s, c0x, c0y, c0z); v[1] = grid_at(s, c0x, c0y, c1z);
  v[2] = grid_at(s, c0x, c1y, c0z); v[3] = grid_at(s, c0x, c1y, c1z);
  v[4] = grid_at(s, c1x, c0y, c0z); v[5] = grid_at(s, c1x, c0y, c1z);
  v[6] = grid_at(s, c1x, c1y, c0z); v[7] = grid_at(s, c1x, c1y, c1z);
  float res = 0.0f;
  res += v[0] * qx * qy * qz;
  res += v[1] * qx * qy * pz;
  res += v[2] * qx * py * qz;
  res += v[3] * qx * py * pz;
  res += v[4] * px * qy * qz;
  res += v[5] * px * qy * pz;
  res += v[6] * px * py * qz;
  res += v[7] * px * py * pz;
  if (cell) *cell = (uint32_t)(((size_t)c0x * s->sy + c0y) * s->sz + c0z);
  return res;
}

/* grid_raytracing.cpp:64-89 */
static v3 grid_normal(const PortScene *s, v3 p, PortCounters *cnt) {
  const float E = 1e-3f;
  float xL = (p.x - E >= -1.0f) ? p.x - E : p.x, xR = (p.x + E <= 1.0f) ? p.x + E : p.x;
  float yL = (p.y - E >= -1.0f) ? p.y - E : p.y, yR = (p.y + E <= 1.0f) ? p.y + E : p.y;
  float zL = (p.z - E >= -1.0f) ? p.z - E : p.z, zR = (p.z + E <= 1.0f) ? p.z + E : p.z;
  v3 a, b;
  a = p; a.x = xR; b = p; b.x = xL; float dx = grid_sdf(s, a, 0, cnt) - grid_sdf(s, b, 0, cnt);
  a = p; a.y = yR; b = p; b.y = yL; float dy = grid_sdf(s, a, 0, cnt) - grid_sdf(s, b, 0, cnt);
  a = p; a.z = zR; b = p; b.z = zL; float dz = grid_sdf(s, a, 0, cnt) - grid_sdf(s, b, 0, cnt);
  v3 g = {dx, dy, dz};
  return lm_normalize(g);
}

/* grid_raytracing.cpp:93-125 */
static Hit grid_intersect(const PortScene *s, v3 o, v3 d, float tNear, float tFar, PortCounters *cnt) {
  Hit r = hit_none();
  v3 inv = {1.0f / d.x, 1.0f / d.y, 1.0f / d.z};
  v3 lo = {-1.0f, -1.0f, -1.0f}, hi = {1.0f, 1.0f, 1.0f};
  float t1, t2; int face;
  bbox_intersection(lo, hi, o, inv, tNear, tFar, &t1, &t2, &face);
  if (t1 > t2) return r;
  float t = t1;
  v3 p = v3_add(o, v3_scale(d, t)); /* rayPos + t * rayDir (t * rayDir componentwise, then add) */
  p.x = std_max(p.x, -1.0f); p.y = std_max(p.y, -1.0f); p.z = std_max(p.z, -1.0f);
  p.x = std_min(p.x, 1.0f);  p.y = std_min(p.y, 1.0f);  p.z = std_min(p.z, 1.0f);
  while (p.x <= 1.0f && p.y <= 1.0f && p.z <= 1.0f && p.x >= -1.0f && p.y >= -1.0f && p.z >= -1.0f) {
    uint32_t cell;
    float sd = grid_sdf(s, p, &cell, cnt);
    if (sd < 1e-3f) {
      r.hitten = 1;
      r.t = t + sd;
      r.normal = grid_normal(s, p, cnt);
      r.prim = cell;
      break;
    }
    t += sd;
    p = v3_add(o, v3_scale(d, t));
  }
  return r;
}

/* ------------------------------------------------------------------ octree (octree_raytracing.cpp) */
static inline const float *oct_values(const PortScene *s, uint32_t node) { return (const float *)(s->nodes36 + (size_t)node * 36); }
static inline uint32_t oct_offset(const PortScene *s, uint32_t node) {
  uint32_t o; memcpy(&o, s->nodes36 + (size_t)node * 36 + 32, 4); return o;
}

/* shared prefix of nodeSDF / nodeNormal: octree_raytracing.cpp:20-35 and :62-79 */
static inline void oct_local(v3 bmin, v3 bmax, v3 p, float pc[3], float cp[3], uint32_t c0[3], uint32_t c1[3]) {
  float q[3] = {(p.x - bmin.x) / (bmax.x - bmin.x), (p.y - bmin.y) / (bmax.y - bmin.y), (p.z - bmin.z) / (bmax.z - bmin.z)};
  for (int i = 0; i < 3; ++i) {
    /* LiteMath clamp(u,a,b) = std::min(std::max(a,u), b), LiteMath.h:97 */
    float u = std_min(std_max(0.0000001f, q[i]), 0.9999999f);
    float f = floorf(u), c = ceilf(u);
    c0[i] = (uint32_t)f; c1[i] = (uint32_t)c;
    pc[i] = u - f; cp[i] = c - u;
  }
}

/* octree_raytracing.cpp:18-57 */
static float oct_node_sdf(const PortScene *s, uint32_t node, v3 bmin, v3 bmax, v3 p, PortCounters *cnt) {
  if (cnt) cnt->octLeafSdf++;
  float pc[3], cp[3]; uint32_t c0[3], c1[3];
  oct_local(bmin, bmax, p, pc, cp, c0, c1);
  const float *val = oct_values(s, node);
#define OV(X, Y, Z) val[((X) << 2) + ((Y) << 1) + (Z)]
  float v[8] = {OV(c0[0], c0[1], c0[2]), OV(c0[0], c0[1], c1[2]), OV(c0[0], c1[1], c0[2]), OV(c0[0], c1[1], c1[2]),
                OV(c1[0], c0[1], c0[2]), OV(c1[0], c0[1], c1[2]), OV(c1[0], c1[1], c0[2]), OV(c1[0], c1[1], c1[2])};
  float res = 0.0f;
  res += v[0] * cp[0] * cp[1] * cp[2];
  res += v[1] * cp[0] * cp[1] * pc[2];
  res += v[2] * cp[0] * pc[1] * cp[2];
  res += v[3] * cp[0] * pc[1] * pc[2];
  res += v[4] * pc[0] * cp[1] * cp[2];
  res += v[5] * pc[0] * cp[1] * pc[2];
  res += v[6] * pc[0] * pc[1] * cp[2];
  res += v[7] * pc[0] * pc[1] * pc[2];
  return res;
}

/* octree_raytracing.cpp:59-118 */
static v3 oct_node_normal(const PortScene *s, uint32_t node, v3 bmin, v3 bmax, v3 p) {
  float pc[3], cp[3]; uint32_t c0[3], c1[3];
  oct_local(bmin, bmax, p, pc, cp, c0, c1);
  const float *val = oct_values(s, node);
  float v[8] = {OV(c0[0], c0[1], c0[2]), OV(c0[0], c0[1], c1[2]), OV(c0[0], c1[1], c0[2]), OV(c0[0], c1[1], c1[2]),
                OV(c1[0], c0[1], c0[2]), OV(c1[0], c0[1], c1[2]), OV(c1[0], c1[1], c0[2]), OV(c1[0], c1[1], c1[2])};
#undef OV
  const float dp = 1.0f, dc = -1.0f; /* dp_c0f, dc1f_p */
  float dfdx = v[0] * dc * cp[1] * cp[2] + v[1] * dc * cp[1] * pc[2] + v[2] * dc * pc[1] * cp[2] + v[3] * dc * pc[1] * pc[2] +
               v[4] * dp * cp[1] * cp[2] + v[5] * dp * cp[1] * pc[2] + v[6] * dp * pc[1] * cp[2] + v[7] * dp * pc[1] * pc[2];
  float dfdy = v[0] * cp[0] * dc * cp[2] + v[1] * cp[0] * dc * pc[2] + v[2] * cp[0] * dp * cp[2] + v[3] * cp[0] * dp * pc[2] +
               v[4] * pc[0] * dc * cp[2] + v[5] * pc[0] * dc * pc[2] + v[6] * pc[0] * dp * cp[2] + v[7] * pc[0] * dp * pc[2];
  float dfdz = v[0] * cp[0] * cp[1] * dc + v[1] * cp[0] * cp[1] * dp + v[2] * cp[0] * pc[1] * dc + v[3] * cp[0] * pc[1] * dp +
               v[4] * pc[0] * cp[1] * dc + v[5] * pc[0] * cp[1] * dp + v[6] * pc[0] * pc[1] * dc + v[7] * pc[0] * pc[1] * dp;
  v3 g = {dfdx, dfdy, dfdz};
  return lm_normalize(g);
}

/* octree_raytracing.cpp:122-164 */
static Hit oct_leaf(const PortScene *s, uint32_t node, v3 bmin, v3 bmax, v3 o, v3 d, float tNear, float tFar, PortCounters *cnt) {
  if (cnt) cnt->octLeaf++;
  Hit r = hit_none();
  const float *val = oct_values(s, node);
  int allBig = 1, allZero = 1, allPos = 1;
  for (int i = 0; i < 8; ++i) {
    if (!(val[i] > 10.0f)) allBig = 0;
    if (!(val[i] == 0.0f)) allZero = 0;
    if (!(val[i] >= 1e-4f)) allPos = 0;
  }
  if (allBig || allZero) return r; /* isEmpty, octree_raytracing.hpp:12-17 */
  if (allPos) return r;            /* octree_raytracing.cpp:132-135 */
  v3 inv = {1.0f / d.x, 1.0f / d.y, 1.0f / d.z};
  float t1, t2; int face;
  bbox_intersection(bmin, bmax, o, inv, tNear, tFar, &t1, &t2, &face);
  if (t1 > t2) return r;
  float t = t1;
  v3 p = v3_add(o, v3_scale(d, t));
  p.x = std_max(p.x, bmin.x); p.y = std_max(p.y, bmin.y); p.z = std_max(p.z, bmin.z);
  p.x = std_min(p.x, bmax.x); p.y = std_min(p.y, bmax.y); p.z = std_min(p.z, bmax.z);
  while (p.x <= bmax.x && p.y <= bmax.y && p.z <= bmax.z && p.x >= bmin.x && p.y >= bmin.y && p.z >= bmin.z) {
    float sd = oct_node_sdf(s, node, bmin, bmax, p, cnt);
    if (sd < 1e-4f) {
      r.hitten = 1;
      r.t = t + sd;
      r.normal = oct_node_normal(s, node, bmin, bmax, p);
      r.prim = node;
      break;
    }
    t += sd;
    p = v3_add(o, v3_scale(d, t));
  }
  return r;
}

/* octree_raytracing.cpp:166-202 with divide_box_8 (ray_pack.ispc:220-239) */
static Hit oct_node(const PortScene *s, uint32_t node, v3 o, v3 d, float tNear, float tFar, v3 bmin, v3 bmax, PortCounters *cnt) {
  uint32_t off = oct_offset(s, node);
  if (off == 0) return oct_leaf(s, node, bmin, bmax, o, d, tNear, tFar, cnt);
  if (cnt) cnt->octInner++;
  v3 c = {(bmin.x + bmax.x) / 2.0f, (bmin.y + bmax.y) / 2.0f, (bmin.z + bmax.z) / 2.0f};
  v3 h = v3_sub(c, bmin);
  float mn[8][3], mx[8][3], ts[8];
  uint32_t ch[8];
  v3 inv = {1.0f / d.x, 1.0f / d.y, 1.0f / d.z};
  for (int id = 0; id < 8; ++id) {
    int x = id >> 2, y = (id & 3) >> 1, z = id & 1;
    mn[id][0] = x == 0 ? bmin.x : c.x; mn[id][1] = y == 0 ? bmin.y : c.y; mn[id][2] = z == 0 ? bmin.z : c.z;
    mx[id][0] = mn[id][0] + h.x; mx[id][1] = mn[id][1] + h.y; mx[id][2] = mn[id][2] + h.z;
    ts[id] = slab_lane(mn[id][0], mn[id][1], mn[id][2], mx[id][0], mx[id][1], mx[id][2], o, inv, tNear, tFar);
    ch[id] = (uint32_t)id;
  }
  sort8(ts, ch);
  Hit r = hit_none();
  for (int i = 0; i < 8; ++i) {
    uint32_t id = ch[i];
    float t = ts[i];
    if (t > 0 && (!r.hitten || r.t > t)) {
      v3 cmin = {mn[id][0], mn[id][1], mn[id][2]}, cmax = {mx[id][0], mx[id][1], mx[id][2]};
      Hit h2 = oct_node(s, off + id, o, d, tNear, tFar, cmin, cmax, cnt);
      if (h2.hitten) { r = h2; break; }
    }
  }
  return r;
}

/* ------------------------------------------------------------------ BVH8 (triangles_raytracing.cpp:266-335) */
typedef struct { int hitten; float t; uint32_t tri; } BvhHit; /* tri = BVH-order triangle index */

static inline v3 mesh_vertex(const PortScene *s, uint32_t idx) { /* v /= v.w, triangles_raytracing.cpp:304-309 */
  const float *p = s->vpos4 + (size_t)idx * 4;
  v3 r = {p[0] / p[3], p[1] / p[3], p[2] / p[3]};
  return r;
}

static BvhHit bvh_node(const PortScene *s, uint32_t index, v3 o, v3 d, float tNear, float tFar, PortCounters *cnt) {
  const uint8_t *node = s->bvhNodes + (size_t)index * s->bvhNodeBytes;
  BvhHit result = {0, INFINITY, 0};
  const int isLeaf = node[200] != 0; /* bool after the 200-byte union */
  if (!isLeaf) {
    if (cnt) cnt->bvhInner++;
    const float *b = (const float *)node; /* Box8: xMin[8] yMin[8] zMin[8] xMax[8] yMax[8] zMax[8] */
    uint32_t realCount, offset;
    memcpy(&realCount, node + 192, 4);
    memcpy(&offset, node + 196, 4);
    float t[8];
    uint32_t ch[8];
    v3 inv = {1.0f / d.x, 1.0f / d.y, 1.0f / d.z};
    for (int i = 0; i < 8; ++i) {
      t[i] = slab_lane(b[i], b[8 + i], b[16 + i], b[24 + i], b[32 + i], b[40 + i], o, inv, tNear, tFar);
      ch[i] = (uint32_t)i;
    }
    sort8(t, ch);
    for (int i = 0; i < 8; ++i) {
      uint32_t id = ch[i];
      if (id >= realCount || (t[i] < 0) || (result.hitten && result.t < t[i])) continue;
      BvhHit cur = bvh_node(s, offset + id, o, d, tNear, tFar, cnt);
      if (cur.hitten && (!result.hitten || result.t > cur.t)) result = cur;
    }
  } else {
    if (cnt) cnt->bvhLeaf++;
    uint32_t start, count;
    memcpy(&start, node, 4);
    memcpy(&count, node + 4, 4);
    uint32_t nTri = count / 3;
    if (nTri > 8) nTri = 8;
    for (uint32_t k = 0; k < nTri; ++k) {
      if (cnt) cnt->bvhTris++;
      v3 v0 = mesh_vertex(s, s->indices[start + 3 * k]);
      v3 v1 = mesh_vertex(s, s->indices[start + 3 * k + 1]);
      v3 v2 = mesh_vertex(s, s->indices[start + 3 * k + 2]);
      float th;
      if (tri_lane(o, d, v0, v1, v2, &th) && (!result.hitten || result.t > th)) {
        result.hitten = 1;
        result.t = th;
        result.tri = start / 3 + k;
      }
    }
  }
  return result;
}

/* No BVH given: every triangle in index order, first closest wins (the loop ray_pack.ispc:167-202 describes,
 * with the leaf acceptance rule of triangles_raytracing.cpp:324-331: no t-range test on the triangle itself).
 * nIdx / 3 is passed in bvhNodeBytes' place.
 *
 * <= 8 triangles: this IS the reference.  BVHBuilder::createNode (triangles_raytracing.cpp:155-186) finds no split
 * for a range of <= 24 indices (tryDivide returns at :120-122) and makes node 0 a leaf; traverseNode(0) on a leaf
 * (:292-331) tests its triangles in index order with no box test and no range check whatsoever, so hits behind the
 * origin, in front of tNear and beyond tFar are all returned.  Pinned bit for bit against the compiled reference on a
 * 2-triangle quad, a tetrahedron and an open 5-triangle fan from outside, grazing and inside, looking at and away,
 * cleared and pre-filled depth, with and without plane + secondary rays (tests/test_small_meshes.py).
 *
 * > 8 triangles: the reference reaches triangles only through child boxes that pass intersect_box_8 against
 * [tNear, tFar] (:271-291), and which triangles share a box is a property of ITS tree.  This loop is then the
 * TREE-INDEPENDENT truth only: it equals the reference wherever the closest hit does not depend on the tree (eye outside
 * the mesh, cleared framebuffer: every benchmark configuration; pinned on 4 meshes x 94 cameras).  Two tree-independent
 * conditions stand in for the box culling elsewhere: the ray must pass the same slab test on the mesh's bounding box
 * (inside which every box of any tree lies), and `th <= tFar`.  On cube.obj that also reproduces the reference from
 * inside and with pre-filled depth (tests/test_host_emul.py); on other meshes those views are tree-dependent and only
 * the reference's own tree (bvh_node above, rt_scene_create_mesh_bvh8 in the product) answers them. */
static BvhHit mesh_brute(const PortScene *s, v3 o, v3 d, float tNear, float tFar, PortCounters *cnt) {
  BvhHit result = {0, INFINITY, 0};
  uint32_t nTri = s->bvhNodeBytes; /* triangle count when bvhNodes == NULL */
  const int rootLeaf = nTri <= 8;
  if (!rootLeaf) {
    v3 inv = {1.0f / d.x, 1.0f / d.y, 1.0f / d.z};
    if (slab_lane(s->bboxMin[0], s->bboxMin[1], s->bboxMin[2], s->bboxMax[0], s->bboxMax[1], s->bboxMax[2], o, inv, tNear,
                  tFar) < 0.0f)
      return result;
  }
  for (uint32_t k = 0; k < nTri; ++k) {
    if (cnt) cnt->bvhTris++;
    v3 v0 = mesh_vertex(s, s->indices[3 * k]);
    v3 v1 = mesh_vertex(s, s->indices[3 * k + 1]);
    v3 v2 = mesh_vertex(s, s->indices[3 * k + 2]);
    float th;
    if (tri_lane(o, d, v0, v1, v2, &th) && (rootLeaf || th <= tFar) && (!result.hitten || result.t > th)) {
      result.hitten = 1;
      result.t = th;
      result.tri = k;
    }
  }
  return result;
}

static Hit mesh_intersect(const PortScene *s, v3 o, v3 d, float tNear, float tFar, PortCounters *cnt) {
  Hit r = hit_none();
  BvhHit b = s->bvhNodes ? bvh_node(s, 0, o, d, tNear, tFar, cnt) : mesh_brute(s, o, d, tNear, tFar, cnt);
  if (!b.hitten) return r;
  r.hitten = 1;
  r.t = b.t;
  v3 v0 = mesh_vertex(s, s->indices[3 * b.tri]);
  v3 v1 = mesh_vertex(s, s->indices[3 * b.tri + 1]);
  v3 v2 = mesh_vertex(s, s->indices[3 * b.tri + 2]);
  r.normal = ispc_normalize(v3_cross(v3_sub(v1, v0), v3_sub(v2, v0)));
  r.prim = s->primOfPermuted ? s->primOfPermuted[b.tri] : b.tri;
  return r;
}

/* ------------------------------------------------------------------ Plane (raytracing.hpp:119-186) for normal (0,1,0), offset planeY */
static Hit plane_intersect(float planeY, v3 o, v3 d, float tNear, float tFar) {
  Hit r = hit_none();
  const v3 n = {0.0f, 1.0f, 0.0f};
  /* recalcBasis(): |n.x| is not the largest component */
  v3 b1raw = {0.0f, n.z, -n.y};
  v3 b1 = lm_normalize(b1raw);
  v3 b2 = lm_normalize(v3_cross(b1, n));
  float dividor = v3_dot(d, n);
  if (fabsf(dividor) < 1e-8f) return r;
  float t = (planeY - v3_dot(o, n)) / dividor;
  if (t < tNear || t > tFar) return r;
  v3 p = v3_add(o, v3_scale(d, t));
  int x = (int)ceilf(v3_dot(p, b1));
  int y = (int)ceilf(v3_dot(p, b2));
  float c = ((x + y) % 2 == 0) ? 0.0f : 1.0f;
  r.hitten = 1;
  r.t = t;
  r.normal = n;
  r.albedo.x = c; r.albedo.y = c; r.albedo.z = c;
  r.reflectiveness = 0.3f;
  r.prim = PRIM_PLANE;
  return r;
}

/* ------------------------------------------------------------------ scene dispatch (IScene::intersect, SceneUnion raytracing.hpp:83-97) */
static Hit model_intersect(const PortScene *s, v3 o, v3 d, float tNear, float tFar, PortCounters *cnt) {
  if (cnt) cnt->rays++;
  switch (s->kind) {
  case 0: return grid_intersect(s, o, d, tNear, tFar, cnt);
  case 1: { v3 lo = {-1.0f, -1.0f, -1.0f}, hi = {1.0f, 1.0f, 1.0f}; return oct_node(s, 0, o, d, tNear, tFar, lo, hi, cnt); }
  default: return mesh_intersect(s, o, d, tNear, tFar, cnt);
  }
}

static Hit scene_intersect(const PortScene *s, v3 o, v3 d, float tNear, float tFar, PortCounters *cnt) {
  Hit a = model_intersect(s, o, d, tNear, tFar, cnt);
  if (!s->hasPlane) return a;
  Hit b = plane_intersect(s->planeY, o, d, tNear, tFar);
  return (a.t < b.t) ? a : b;
}

/* ------------------------------------------------------------------ Renderer (raytracing.cpp) */
typedef struct { float r, g, b, a; } c4;

/* raytracing.cpp:13-65 */
static c4 intersection_color(const PortScene *s, const PortFrame *f, v3 o, v3 d, float tNear, float tFar, float tPrev,
                             int maxDepth, float *tOut, uint32_t *primOut, PortCounters *cnt) {
  Hit hit = scene_intersect(s, o, d, tNear, std_min(tFar, tPrev), cnt);
  c4 color = {0.0f, 0.0f, 0.0f, 1.0f};
  if (!hit.hitten) { *tOut = INFINITY; if (primOut) *primOut = PRIM_MISS; return color; }
  if (cnt) cnt->hits++;
  if (v3_dot(hit.normal, d) > 0) hit.normal = v3_scale(hit.normal, -1.0f);
  if (f->mode == 0) {
    color.r = (hit.normal.x + 1.0f) / 2.0f;
    color.g = (hit.normal.y + 1.0f) / 2.0f;
    color.b = (hit.normal.z + 1.0f) / 2.0f;
    color.a = (1.0f + 1.0f) / 2.0f;
  } else if (f->mode == 2) {
    color.r = hit.albedo.x; color.g = hit.albedo.y; color.b = hit.albedo.z; color.a = 1.0f;
  } else {
    int lightIsVisible = 1;
    v3 light = {f->light[0], f->light[1], f->light[2]};
    v3 point = v3_add(o, v3_scale(d, hit.t)); /* rayPos + hit.t * rayDir */
    if (f->shadows) {
      v3 shadowDir = lm_normalize(v3_sub(light, point));
      Hit sh = scene_intersect(s, v3_add(point, v3_scale(shadowDir, 0.3f)), shadowDir, 0.01f, 100.0f, cnt);
      lightIsVisible = !sh.hitten;
    }
    if (!lightIsVisible) {
      color.r = hit.albedo.x * 0.1f; color.g = hit.albedo.y * 0.1f; color.b = hit.albedo.z * 0.1f; color.a = 1.0f;
    } else {
      v3 ldir = lm_normalize(v3_sub(point, light));
      v3 neg = {-ldir.x, -ldir.y, -ldir.z};
      float lam = std_max(v3_dot(neg, hit.normal), 0.0f); /* Lambert(), raytracing.cpp:8-11: scalar * albedo */
      color.r = std_min(hit.albedo.x * 0.1f + lam * hit.albedo.x, 1.0f);
      color.g = std_min(hit.albedo.y * 0.1f + lam * hit.albedo.y, 1.0f);
      color.b = std_min(hit.albedo.z * 0.1f + lam * hit.albedo.z, 1.0f);
      color.a = 1.0f;
    }
    if (f->reflections && maxDepth > 1 && hit.reflectiveness > 0.0f) {
      /* reflect(dir, n) = n * dot(dir, n) * (-2) + dir, LiteMath.h:1196 */
      float dn = v3_dot(d, hit.normal);
      v3 refl = {hit.normal.x * dn * (-2.0f) + d.x, hit.normal.y * dn * (-2.0f) + d.y, hit.normal.z * dn * (-2.0f) + d.z};
      v3 reflectDir = lm_normalize(refl);
      float tmpT;
      c4 rc = intersection_color(s, f, v3_add(point, v3_scale(reflectDir, 0.02f)), reflectDir, 0.01f, 100.0f, INFINITY,
                                 maxDepth - 1, &tmpT, 0, cnt);
      float k = hit.reflectiveness, ik = 1.0f - hit.reflectiveness;
      color.r = color.r * ik + k * rc.r;
      color.g = color.g * ik + k * rc.g;
      color.b = color.b * ik + k * rc.b;
      color.a = color.a * ik + k * rc.a;
    }
  }
  *tOut = hit.t;
  if (primOut) *primOut = hit.prim;
  return color;
}

/* raytracing.cpp:82-88 + EyeRayDir4f (LiteMath.h:2121-2134) + mat4_colmajor_mul_vec4 (:1371-1377) */
static v3 eye_ray(const PortFrame *f, int x, int y) {
  float fx = (float)x + 0.5f, fy = (float)y + 0.5f;
  float w = (float)f->W, h = (float)f->H;
  float V[4] = {2.0f * (fx + 0.5f) / w - 1.0f, -2.0f * (fy + 0.5f) / h + 1.0f, 0.0f, 1.0f};
  const float *B = f->projInv;
  float pos[4];
  for (int i = 0; i < 4; ++i) pos[i] = V[0] * B[i] + V[1] * B[4 + i] + V[2] * B[8 + i] + V[3] * B[12 + i];
  float pw = pos[3];
  for (int i = 0; i < 4; ++i) pos[i] = pos[i] / pw;
  pos[1] *= (-1.0f);
  float lenInv = 1.0f / sqrtf(pos[0] * pos[0] + pos[1] * pos[1] + pos[2] * pos[2]);
  float D[4] = {pos[0] * lenInv, pos[1] * lenInv, pos[2] * lenInv, 0.0f};
  const float *M = f->viewInv;
  v3 r;
  r.x = D[0] * M[0] + D[1] * M[4] + D[2] * M[8] + D[3] * M[12];
  r.y = D[0] * M[1] + D[1] * M[5] + D[2] * M[9] + D[3] * M[13];
  r.z = D[0] * M[2] + D[1] * M[6] + D[2] * M[10] + D[3] * M[14];
  return r;
}

/* LiteMath.h:1741-1746 + to_uint32 :1153 (C cast = truncation) */
static inline uint32_t color_pack_rgba(c4 c) {
  uint32_t r = (uint32_t)(c.r * 255.0f), g = (uint32_t)(c.g * 255.0f), b = (uint32_t)(c.b * 255.0f), a = (uint32_t)(c.a * 255.0f);
  return (a << 24) | (b << 16) | (g << 8) | r;
}

static void counters_add(PortCounters *dst, const PortCounters *src) {
  uint64_t *d = (uint64_t *)dst; const uint64_t *s = (const uint64_t *)src;
  for (size_t i = 0; i < sizeof(PortCounters) / 8; ++i) d[i] += s[i];
}

/* Renderer::draw (raytracing.cpp:67-102) over loop rows [y0, y1); color/t/prim are full W*H in-out buffers
 * (prim may be NULL; it is written only where color/t are). */
void port_render(const PortScene *s, const PortFrame *f, uint32_t y0, uint32_t y1, uint32_t *color, float *t,
                 uint32_t *prim, PortCounters *counters) {
  const int width = (int)f->W, height = (int)f->H;
  const v3 o = {f->rayPos[0], f->rayPos[1], f->rayPos[2]};
  PortCounters total; memset(&total, 0, sizeof total);
#pragma omp parallel
  {
    PortCounters local; memset(&local, 0, sizeof local);
    PortCounters *cnt = counters ? &local : 0;
#pragma omp for schedule(dynamic)
    for (int y = (int)y0; y < (int)y1; ++y) {
      for (int x = 0; x < width; ++x) {
        size_t px = (size_t)(height - y - 1) * (size_t)width + (size_t)x;
        v3 d = eye_ray(f, x, y);
        float tNew; uint32_t pr;
        c4 col = intersection_color(s, f, o, d, 0.01f, 100.0f, t[px], 2, &tNew, &pr, cnt);
        if (!isinf(tNew)) {
          t[px] = tNew;
          color[px] = color_pack_rgba(col);
          if (prim) prim[px] = pr;
        }
      }
    }
    if (counters) {
#pragma omp critical
      counters_add(&total, &local);
    }
  }
  if (counters) *counters = total;
}

/* single ray through the scene (IScene::intersect of the model or of SceneUnion(model, plane)) */
int port_intersect(const PortScene *s, const float o[3], const float d[3], float tNear, float tFar, float *t,
                   float normal[3], uint32_t *prim) {
  v3 oo = {o[0], o[1], o[2]}, dd = {d[0], d[1], d[2]};
  Hit h = scene_intersect(s, oo, dd, tNear, tFar, 0);
  *t = h.t;
  normal[0] = h.normal.x; normal[1] = h.normal.y; normal[2] = h.normal.z;
  if (prim) *prim = h.prim;
  return h.hitten;
}

/* primitives for KATs */
void port_eye_ray(const PortFrame *f, uint32_t x, uint32_t y, float out[3]) {
  v3 r = eye_ray(f, (int)x, (int)y); out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
float port_grid_sdf(const PortScene *s, const float p[3]) { v3 q = {p[0], p[1], p[2]}; return grid_sdf(s, q, 0, 0); }
void port_grid_normal(const PortScene *s, const float p[3], float out[3]) {
  v3 q = {p[0], p[1], p[2]}; v3 n = grid_normal(s, q, 0); out[0] = n.x; out[1] = n.y; out[2] = n.z;
}
void port_bbox_intersection(const float bmin[3], const float bmax[3], const float o[3], const float inv[3], float tMin,
                            float tMax, float *t1, float *t2, int *face) {
  v3 a = {bmin[0], bmin[1], bmin[2]}, b = {bmax[0], bmax[1], bmax[2]}, oo = {o[0], o[1], o[2]}, ii = {inv[0], inv[1], inv[2]};
  bbox_intersection(a, b, oo, ii, tMin, tMax, t1, t2, face);
}
void port_intersect_box_8(const float boxes48[48], const float o[3], const float inv[3], float tNear, float tFar, float out[8]) {
  v3 oo = {o[0], o[1], o[2]}, ii = {inv[0], inv[1], inv[2]};
  const float *b = boxes48;
  for (int i = 0; i < 8; ++i) out[i] = slab_lane(b[i], b[8 + i], b[16 + i], b[24 + i], b[32 + i], b[40 + i], oo, ii, tNear, tFar);
}
void port_sort8(float t[8], uint32_t c[8]) { sort8(t, c); }
int port_triangle(const float o[3], const float d[3], const float tri9[9], float *t, float normal[3]) {
  v3 oo = {o[0], o[1], o[2]}, dd = {d[0], d[1], d[2]};
  v3 v0 = {tri9[0], tri9[1], tri9[2]}, v1 = {tri9[3], tri9[4], tri9[5]}, v2 = {tri9[6], tri9[7], tri9[8]};
  v3 n = ispc_normalize(v3_cross(v3_sub(v1, v0), v3_sub(v2, v0)));
  normal[0] = n.x; normal[1] = n.y; normal[2] = n.z;
  *t = -1.0f;
  return tri_lane(oo, dd, v0, v1, v2, t);
}
uint32_t port_color_pack(const float c[4]) { c4 k = {c[0], c[1], c[2], c[3]}; return color_pack_rgba(k); }
