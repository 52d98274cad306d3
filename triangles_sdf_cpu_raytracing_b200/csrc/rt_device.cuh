// Device-side restatement of the reference's per-ray arithmetic, written from scratch for one-thread-per-ray
// execution (explicit stacks instead of recursion, per-thread loops instead of 8-lane ISPC calls).
// Included inside `namespace rt::<flavour>`; the exact flavour is compiled with -fmad=false so that every
// operator below is one correctly rounded IEEE-754 binary32 operation in the reference's operator order.
// All file:line citations are into the reference tree.

struct V3 {
  float x, y, z;
};

// std::min/std::max (LiteMath.h:793-795 use them): FIRST argument unless the comparison says otherwise.
// ISPC min/max (ray_pack.ispc:258-265) lower to minps/maxps: SECOND operand when either is NaN. SURVEY T11.
__device__ __forceinline__ float std_min(float a, float b) { return (b < a) ? b : a; }
__device__ __forceinline__ float std_max(float a, float b) { return (a < b) ? b : a; }
__device__ __forceinline__ float ispc_min(float a, float b) { return a < b ? a : b; }
__device__ __forceinline__ float ispc_max(float a, float b) { return a > b ? a : b; }

__device__ __forceinline__ V3 mk3(float x, float y, float z) {
  V3 r;
  r.x = x; r.y = y; r.z = z;
  return r;
}
__device__ __forceinline__ V3 add3(V3 a, V3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 sub3(V3 a, V3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 scale3(V3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ float dot3(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// LiteMath.h:835-840 and ray_pack.ispc:12-20: shuffle_yzx(a*b_yzx - a_yzx*b)
__device__ __forceinline__ V3 cross3(V3 a, V3 b) {
  return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
// LiteMath.h:814: a * (1/length)
__device__ __forceinline__ V3 lm_normalize(V3 a) {
  float lenInv = 1.0f / sqrtf(dot3(a, a));
  return scale3(a, lenInv);
}
// ray_pack.ispc:5-10: ray / length (three divisions)
__device__ __forceinline__ V3 ispc_normalize(V3 r) {
  float length = sqrtf(r.x * r.x + r.y * r.y + r.z * r.z);
  return mk3(r.x / length, r.y / length, r.z / length);
}
// rayPos + t * rayDir as LiteMath evaluates it: componentwise product, then sum
__device__ __forceinline__ V3 along(V3 o, V3 d, float t) { return mk3(o.x + d.x * t, o.y + d.y * t, o.z + d.z * t); }

struct Hit {
  bool hitten;
  float t;
  V3 normal;
  float albedo;          // grey albedo: models (1,1,1), plane checker 0/1 (raytracing.hpp:67-73,160-163)
  float reflectiveness;
  uint32_t prim;
};
__device__ __forceinline__ Hit hit_none() {
  Hit h;
  h.hitten = false;
  h.t = __int_as_float(0x7f800000);
  h.normal = mk3(0.0f, 1.0f, 0.0f);
  h.albedo = 1.0f;
  h.reflectiveness = 0.0f;
  h.prim = RT_PRIM_MISS;
  return h;
}

struct Counters {
  uint32_t grid_samples, oct_inner, oct_leaf, oct_leaf_samples, bvh_inner, bvh_leaf, bvh_tris, rays, hits;
};

// ---- LiteMath::BBox3f::Intersection (LiteMath.h:2005-2067) ------------------------------------------------
__device__ __forceinline__ void bbox_intersection(V3 bmin, V3 bmax, V3 o, V3 inv, float min_t, float max_t, float &t1,
                                                  float &t2) {
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
  t2 = std_min(tmax_z, std_min(tmax_y, std_min(tmax_x, max_t)));
  if (tmin_x > tmin_y) {
    t1 = (tmin_z > tmin_x) ? std_max(tmin_z, min_t) : std_max(tmin_x, min_t);
  } else {
    t1 = (tmin_z > tmin_y) ? std_max(tmin_z, min_t) : std_max(tmin_y, min_t);
  }
}

// ---- ispc::intersect_box_8, one lane (ray_pack.ispc:241-273) ---------------------------------------------
// NANFREE: the caller knows all three 1/dir are finite, so no plane distance can be NaN (NaN needs 0 * inf) and
// ISPC's select-form min/max (second operand on NaN) equals the hardware FMNMX up to the sign of a zero, which
// nothing downstream observes (entry distances are only compared with each other, with hit.t and with 0).
template <bool NANFREE>
__device__ __forceinline__ float mn2(float a, float b) { return NANFREE ? fminf(a, b) : ispc_min(a, b); }
template <bool NANFREE>
__device__ __forceinline__ float mx2(float a, float b) { return NANFREE ? fmaxf(a, b) : ispc_max(a, b); }

template <bool NANFREE>
__device__ __forceinline__ float slab_lane(float xMin, float yMin, float zMin, float xMax, float yMax, float zMax, V3 o,
                                           V3 inv, float tNear, float tFar) {
  float t1x = (xMin - o.x) * inv.x, t1y = (yMin - o.y) * inv.y, t1z = (zMin - o.z) * inv.z;
  float t2x = (xMax - o.x) * inv.x, t2y = (yMax - o.y) * inv.y, t2z = (zMax - o.z) * inv.z;
  float mnx = mn2<NANFREE>(t1x, t2x), mny = mn2<NANFREE>(t1y, t2y), mnz = mn2<NANFREE>(t1z, t2z);
  float mxx = mx2<NANFREE>(t1x, t2x), mxy = mx2<NANFREE>(t1y, t2y), mxz = mx2<NANFREE>(t1z, t2z);
  float tMin = mx2<NANFREE>(mnx, mx2<NANFREE>(mny, mnz));
  float tMax = mn2<NANFREE>(mxx, mn2<NANFREE>(mxy, mxz));
  tMin = mx2<NANFREE>(tMin, tNear);
  tMax = mn2<NANFREE>(tMax, tFar);
  if (tMax < 0 || tMin > tMax) return -1.0f;
  return tMin;
}

// ---- sort8 (raytracing.hpp:188-213): the same 19 comparators in the same order, '>' compare ---------------
#define RT_CSWAP(a, b)                  \
  if (t[a] > t[b]) {                    \
    float tf = t[a]; t[a] = t[b]; t[b] = tf; \
    uint32_t tc = c[a]; c[a] = c[b]; c[b] = tc; \
  }
__device__ __forceinline__ void sort8(float (&t)[8], uint32_t (&c)[8]) {
  RT_CSWAP(0, 1) RT_CSWAP(2, 3) RT_CSWAP(4, 5) RT_CSWAP(6, 7)
  RT_CSWAP(0, 2) RT_CSWAP(1, 3) RT_CSWAP(4, 6) RT_CSWAP(5, 7)
  RT_CSWAP(1, 2) RT_CSWAP(5, 6) RT_CSWAP(0, 4) RT_CSWAP(3, 7)
  RT_CSWAP(1, 5) RT_CSWAP(2, 6)
  RT_CSWAP(1, 4) RT_CSWAP(3, 6)
  RT_CSWAP(2, 4) RT_CSWAP(3, 5)
  RT_CSWAP(3, 4)
}
#undef RT_CSWAP

// ---- Moeller-Trumbore, uniform ray x one triangle (ray_pack.ispc:132-165), without the normal: the normal
// is a pure function of the triangle (:142) and is evaluated for the winner only ---------------------------
__device__ __forceinline__ bool tri_lane(V3 o, V3 d, V3 v0, V3 v1, V3 v2, float &tOut) {
  V3 e1 = sub3(v1, v0);
  V3 e2 = sub3(v2, v0);
  V3 pvec = cross3(d, e2);
  float det = dot3(e1, pvec);
  if (det < 1e-8f && det > -1e-8f) return false;
  float inv_det = 1 / det;
  V3 tvec = sub3(o, v0);
  float u = dot3(tvec, pvec) * inv_det;
  if (u < 0.0f || u > 1.0f) return false;
  V3 qvec = cross3(tvec, e1);
  float v = dot3(d, qvec) * inv_det;
  if (v < 0.0f || u + v > 1.0f) return false;
  tOut = dot3(e2, qvec) * inv_det;
  return true;
}
__device__ __forceinline__ V3 ld3(const float4 &v) { return mk3(v.x, v.y, v.z); }

// =============================================================================================================
// Dense SDF grid (grid_raytracing.cpp)
// =============================================================================================================
// Offsets of sample coordinate c along one axis; the element index is offx(x) + offy(y) + offz(z).
// LINEAR: (x*sy + y)*sz + z (grid_raytracing.hpp:13-15).  BRICKS: 8x8x8 bricks of 512 floats, brick-major.
template <bool BRICKS>
__device__ __forceinline__ uint32_t grid_offx(const GridDev &g, uint32_t x) {
  return BRICKS ? (x >> 3) * g.stride_x + ((x & 7u) << 6) : x * g.stride_x;
}
template <bool BRICKS>
__device__ __forceinline__ uint32_t grid_offy(const GridDev &g, uint32_t y) {
  return BRICKS ? (y >> 3) * g.stride_y + ((y & 7u) << 3) : y * g.stride_y;
}
template <bool BRICKS>
__device__ __forceinline__ uint32_t grid_offz(uint32_t z) {
  return BRICKS ? ((z >> 3) << 9) + (z & 7u) : z;
}

// SDFGrid::sdf(float3), grid_raytracing.cpp:7-62; cell = (c0x, c0y, c0z) packed by the caller when it wants the
// grid's "prim id".
// LINEAR layout: the reference's integer-coordinate special case (:25-38: ceil == floor on an axis -> weights (1, 0),
// both corners the same sample) needs no code.  With c1 = c0 + 1 always, the natural weights on such an axis are
// p - c0 = 0 and c1 - p = 1, both exact: the four corners at c0 carry weight 1 where the reference's four (identical)
// corners at "c1" did, so the same four products come out bit for bit, in the same relative order, and the other four
// terms are (finite sample) x 0 = +-0 where the reference's were +-0 too.  Adding +-0 never changes a non-zero partial
// sum, and a sum that starts at +0 cannot become -0, so the result is the same float.  Needs: finite samples (checked
// at upload; grids with infinities take the literal BRICKS path) and readable memory one slab + one row + one sample
// past the end (the upload pads with zeros).  Saves the three compares and twelve selects per sample, and the z + 1
// loads become immediate offsets of the z loads.
// BRICKS layout: literal form.  ceil(p) is evaluated as floor(p) + (p != floor(p)), the same value for 0 <= p < 2^23.
// (A real branch for the c1 == c0 case instead of the selects was measured and rejected: the case is not rare,
// every ray's first sample sits exactly on the box face it entered through; grid 4K 0.338 -> 0.373 ms.)
template <bool BRICKS, bool COUNT>
__device__ __forceinline__ float grid_sdf(const GridDev &g, V3 p, uint32_t &cell, Counters &cnt) {
  if (COUNT) cnt.grid_samples++;
  // ((p + 1) / 2) * (size - 1): the halving is exact, so the product with hf = (size - 1) / 2 (also exact) rounds the
  // same real number once: the same float as the reference's two multiplications (grid_raytracing.cpp:8-11)
  p.x = (p.x + 1.0f) * g.hfx; p.y = (p.y + 1.0f) * g.hfy; p.z = (p.z + 1.0f) * g.hfz;
  const float c0fx = floorf(p.x), c0fy = floorf(p.y), c0fz = floorf(p.z);
  const uint32_t c0x = (uint32_t)c0fx, c0y = (uint32_t)c0fy, c0z = (uint32_t)c0fz;
  float v0, v1, v2, v3, v4, v5, v6, v7, px, py, pz, qx, qy, qz;
  if (!BRICKS) {
    px = p.x - c0fx; py = p.y - c0fy; pz = p.z - c0fz;                             // p_c0f
    qx = (c0fx + 1.0f) - p.x; qy = (c0fy + 1.0f) - p.y; qz = (c0fz + 1.0f) - p.z;  // c1f_p
    const float *v00 = g.values + ((c0x * g.stride_x + c0y * g.stride_y) + c0z);
    const float *v01 = v00 + g.stride_y, *v10 = v00 + g.stride_x, *v11 = v10 + g.stride_y;
    v0 = __ldg(v00); v1 = __ldg(v00 + 1);
    v2 = __ldg(v01); v3 = __ldg(v01 + 1);
    v4 = __ldg(v10); v5 = __ldg(v10 + 1);
    v6 = __ldg(v11); v7 = __ldg(v11 + 1);
  } else {
    const bool ix = p.x != c0fx, iy = p.y != c0fy, iz = p.z != c0fz;  // c1 != c0 on that axis
    const float c1fx = ix ? c0fx + 1.0f : c0fx, c1fy = iy ? c0fy + 1.0f : c0fy, c1fz = iz ? c0fz + 1.0f : c0fz;
    px = p.x - c0fx; py = p.y - c0fy; pz = p.z - c0fz;  // p_c0f
    qx = c1fx - p.x; qy = c1fy - p.y; qz = c1fz - p.z;  // c1f_p
    if (!ix) { px = 1.0f; qx = 0.0f; }                  // :25-38
    if (!iy) { py = 1.0f; qy = 0.0f; }
    if (!iz) { pz = 1.0f; qz = 0.0f; }
    const uint32_t x0 = grid_offx<BRICKS>(g, c0x), x1 = ix ? grid_offx<BRICKS>(g, c0x + 1u) : x0;
    const uint32_t y0 = grid_offy<BRICKS>(g, c0y), y1 = iy ? grid_offy<BRICKS>(g, c0y + 1u) : y0;
    const uint32_t z0 = grid_offz<BRICKS>(c0z), z1 = iz ? grid_offz<BRICKS>(c0z + 1u) : z0;
    const float *v = g.values;
    const uint32_t a00 = x0 + y0, a01 = x0 + y1, a10 = x1 + y0, a11 = x1 + y1;
    v0 = __ldg(v + (a00 + z0)); v1 = __ldg(v + (a00 + z1));
    v2 = __ldg(v + (a01 + z0)); v3 = __ldg(v + (a01 + z1));
    v4 = __ldg(v + (a10 + z0)); v5 = __ldg(v + (a10 + z1));
    v6 = __ldg(v + (a11 + z0)); v7 = __ldg(v + (a11 + z1));
  }
  float res = 0.0f;
  res += v0 * qx * qy * qz;
  res += v1 * qx * qy * pz;
  res += v2 * qx * py * qz;
  res += v3 * qx * py * pz;
  res += v4 * px * qy * qz;
  res += v5 * px * qy * pz;
  res += v6 * px * py * qz;
  res += v7 * px * py * pz;
  cell = (c0x * g.sy + c0y) * g.sz + c0z;
  return res;
}

// SDFGrid::normal, grid_raytracing.cpp:64-89: clamped +-1e-3 differences, not scaled, then normalize
template <bool BRICKS, bool COUNT>
__device__ __forceinline__ V3 grid_normal(const GridDev &g, V3 p, Counters &cnt) {
  const float E = 1e-3f;
  float xL = (p.x - E >= -1.0f) ? p.x - E : p.x, xR = (p.x + E <= 1.0f) ? p.x + E : p.x;
  float yL = (p.y - E >= -1.0f) ? p.y - E : p.y, yR = (p.y + E <= 1.0f) ? p.y + E : p.y;
  float zL = (p.z - E >= -1.0f) ? p.z - E : p.z, zR = (p.z + E <= 1.0f) ? p.z + E : p.z;
  uint32_t c;
  float dx = grid_sdf<BRICKS, COUNT>(g, mk3(xR, p.y, p.z), c, cnt) - grid_sdf<BRICKS, COUNT>(g, mk3(xL, p.y, p.z), c, cnt);
  float dy = grid_sdf<BRICKS, COUNT>(g, mk3(p.x, yR, p.z), c, cnt) - grid_sdf<BRICKS, COUNT>(g, mk3(p.x, yL, p.z), c, cnt);
  float dz = grid_sdf<BRICKS, COUNT>(g, mk3(p.x, p.y, zR), c, cnt) - grid_sdf<BRICKS, COUNT>(g, mk3(p.x, p.y, zL), c, cnt);
  return lm_normalize(mk3(dx, dy, dz));
}

// samples after which a PREFETCH march starts asking for the lines ahead of it (profiles/r2_prefetch.txt: 8 / 16 / 32)
constexpr uint32_t GRID_PREFETCH_AFTER = 32;
// SDFGrid::intersect, grid_raytracing.cpp:93-125.  The march loop only marches; the hit's normal (6 more samples)
// is evaluated after the loop so that all lanes of the warp that hit do it together instead of one by one at
// the iteration each of them happens to hit.
// PREFETCH: the flavour the small-launch (HEAVY) kernels use.  Measured: one GPU's eighth of the 4K 513^3 frame 0.086 ->
// 0.077 ms; on full frames the extra counter costs the register-tight loop 2-5 %, so they keep the plain flavour.
template <bool BRICKS, bool COUNT, bool PREFETCH = false, bool PIN = false>
__device__ __forceinline__ Hit grid_intersect(const GridDev &g0, V3 o, V3 d, float tNear, float tFar, Counters &cnt) {
  // PIN: the seven launch constants of the march loop (scale, strides, base pointer) are made to look per-thread (a
  // select on a condition the compiler cannot decide), so that they stay in vector registers.  Left alone, ptxas keeps
  // them in uniform registers and copies each one into a vector register again at every use: seven MOVs in a
  // 94-instruction loop.  The kernel still fits its 48 registers.  Measured (profiles/r2_grid_loop.txt): 513^3 grid 4K
  // 0.327 -> 0.303 ms, one GPU's eighth of it 0.081 -> 0.070; the out-of-line copy the secondary rays call loses 4 %
  // with it and keeps the plain form.
#ifdef __CUDA_ARCH__
  constexpr bool PINNED = PIN && !BRICKS;
  GridDev gp;
  if (PINNED) {
    const bool never = d.x != d.x;
    gp = g0;
    gp.hfx = never ? g0.hfy : g0.hfx; gp.hfy = never ? g0.hfz : g0.hfy; gp.hfz = never ? g0.hfx : g0.hfz;
    gp.stride_x = never ? g0.stride_y : g0.stride_x; gp.stride_y = never ? g0.stride_x : g0.stride_y;
    gp.values = never ? g0.values + 1 : g0.values;  // (every alternative is harmless: a NaN direction never samples)
  }
  const GridDev &g = PINNED ? gp : g0;
#else
  const GridDev &g = g0;
#endif
  Hit r = hit_none();
  V3 inv = mk3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
  float t1, t2;
  bbox_intersection(mk3(-1.0f, -1.0f, -1.0f), mk3(1.0f, 1.0f, 1.0f), o, inv, tNear, tFar, t1, t2);
  if (t1 > t2) return r;
  float t = t1;
  V3 p = along(o, d, t);
  p.x = std_max(p.x, -1.0f); p.y = std_max(p.y, -1.0f); p.z = std_max(p.z, -1.0f);
  p.x = std_min(p.x, 1.0f); p.y = std_min(p.y, 1.0f); p.z = std_min(p.z, 1.0f);
  // (one `max(|x|, |y|, |z|) <= 1` and a cell index recomputed after the loop were measured: 1080p 0.112 -> 0.139 ms,
  // 513^3 4K 0.366 -> 0.395 ms; three compares on |x|, |y|, |z| instead of the six: 0.331 -> 0.327 ms)
  uint32_t steps = 0;
  // |x| <= 1 is the reference's pair of tests (-1 <= x && x <= 1, false for a NaN either way) in one compare
  while (fabsf(p.x) <= 1.0f && fabsf(p.y) <= 1.0f && fabsf(p.z) <= 1.0f) {
    uint32_t cell;
    float sd = grid_sdf<BRICKS, COUNT>(g, p, cell, cnt);
    if (sd < 1e-3f) {
      r.hitten = true;
      r.t = t + sd;
      r.prim = cell;
      break;
    }
    t += sd;
    p = along(o, d, t);
    // A march that is still going after GRID_PREFETCH_AFTER samples is one of the few long ones that decide when the launch
    // ends (a dependent chain of 100+ samples, each a cache miss on a grid that does not fit the caches): ask for the
    // lines two steps ahead along the ray while this step's arithmetic runs.  A hint only: no value is read.
    if (PREFETCH && !BRICKS && ++steps >= GRID_PREFETCH_AFTER) {
      const float ta = t + sd + sd;
      const float qx = fminf(fmaxf((o.x + d.x * ta + 1.0f) * g.hfx, 0.0f), g.fx - 1.0f);
      const float qy = fminf(fmaxf((o.y + d.y * ta + 1.0f) * g.hfy, 0.0f), g.fy - 1.0f);
      const float qz = fminf(fmaxf((o.z + d.z * ta + 1.0f) * g.hfz, 0.0f), g.fz - 1.0f);
      const float *q = g.values + ((uint32_t)qx * g.stride_x + (uint32_t)qy * g.stride_y + (uint32_t)qz);
#ifdef __CUDA_ARCH__
      asm volatile("prefetch.global.L1 [%0];" ::"l"(q));
      asm volatile("prefetch.global.L1 [%0];" ::"l"(q + g.stride_y));
      asm volatile("prefetch.global.L1 [%0];" ::"l"(q + g.stride_x));
      asm volatile("prefetch.global.L1 [%0];" ::"l"(q + g.stride_x + g.stride_y));
#else
      (void)q;
#endif
    }
  }
  if (r.hitten) r.normal = grid_normal<BRICKS, COUNT>(g, p, cnt);
  return r;
}

// compare-exchange of two packed sort keys (octree crossing order, BVH child order)
__device__ __forceinline__ void key_cswap(uint32_t &a, uint32_t &b) {
  const uint32_t lo = min(a, b), hi = max(a, b);
  a = lo;
  b = hi;
}

// =============================================================================================================
// SDF octree (octree_raytracing.cpp)
// =============================================================================================================
// SDFOctree::nodeSDF(float3), octree_raytracing.cpp:18-57.  The local coordinate is clamped to
// [1e-7, 0.9999999] so floor/ceil are always 0/1: c0 = 0, c1 = 1, p_c0f = u, c1f_p = 1 - u.
// (p - min) / (max - min): the extent is a power of two, so multiplying by its reciprocal is the same
// correctly rounded result as the reference's division.
__device__ __forceinline__ void oct_local(V3 bmin, float invExt, V3 p, float (&pc)[3], float (&cp)[3]) {
  float q[3] = {(p.x - bmin.x) * invExt, (p.y - bmin.y) * invExt, (p.z - bmin.z) * invExt};
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    float u = std_min(std_max(0.0000001f, q[i]), 0.9999999f);  // LiteMath clamp = min(max(a,u),b), LiteMath.h:97
    pc[i] = u - 0.0f;
    cp[i] = 1.0f - u;
  }
}

__device__ __forceinline__ float oct_node_sdf(const float (&v)[8], V3 bmin, float invExt, V3 p) {
  float pc[3], cp[3];
  oct_local(bmin, invExt, p, pc, cp);
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

// SDFOctree::nodeNormal, octree_raytracing.cpp:59-118: analytic gradient of the same trilinear form
__device__ __forceinline__ V3 oct_node_normal(const float (&v)[8], V3 bmin, float invExt, V3 p) {
  float pc[3], cp[3];
  oct_local(bmin, invExt, p, pc, cp);
  const float dp = 1.0f, dc = -1.0f;  // dp_c0f, dc1f_p
  float dfdx = v[0] * dc * cp[1] * cp[2] + v[1] * dc * cp[1] * pc[2] + v[2] * dc * pc[1] * cp[2] +
               v[3] * dc * pc[1] * pc[2] + v[4] * dp * cp[1] * cp[2] + v[5] * dp * cp[1] * pc[2] +
               v[6] * dp * pc[1] * cp[2] + v[7] * dp * pc[1] * pc[2];
  float dfdy = v[0] * cp[0] * dc * cp[2] + v[1] * cp[0] * dc * pc[2] + v[2] * cp[0] * dp * cp[2] +
               v[3] * cp[0] * dp * pc[2] + v[4] * pc[0] * dc * cp[2] + v[5] * pc[0] * dc * pc[2] +
               v[6] * pc[0] * dp * cp[2] + v[7] * pc[0] * dp * pc[2];
  float dfdz = v[0] * cp[0] * cp[1] * dc + v[1] * cp[0] * cp[1] * dp + v[2] * cp[0] * pc[1] * dc +
               v[3] * cp[0] * pc[1] * dp + v[4] * pc[0] * cp[1] * dc + v[5] * pc[0] * cp[1] * dp +
               v[6] * pc[0] * pc[1] * dc + v[7] * pc[0] * pc[1] * dp;
  return lm_normalize(mk3(dfdx, dfdy, dfdz));
}

// SDFOctree::intersectLeaf, octree_raytracing.cpp:122-164, for a leaf that survived the value culls.
// On a hit the sample point is returned in `pHit`; the caller evaluates the normal once the traversal is over.
// KNOWN_INSIDE: the caller reached this leaf through the closed-form child list (oct_children_fast), i.e. the ISPC
// slab test of the leaf box succeeded with finite, NaN-free plane distances and tNear > 0.  Then
// BBox3f::Intersection cannot reject it (its t2 is the same tMax with every term scaled by 1.00000024 >= the slab
// test's, its t1 the same max of the near-plane distances) and t1 reduces to max(near-plane distances, tNear): the
// same products (plane - o) * inv on the same operands, selected by the sign of 1/dir as LiteMath.h:2011-2016 does.
template <bool COUNT>
__device__ __forceinline__ bool oct_leaf(const OctreeDev &oc, uint32_t node, V3 bmin, float ext, float invExt, V3 o,
                                         V3 d, V3 inv, float tNear, float tFar, bool KNOWN_INSIDE, Hit &r, V3 &pHit,
                                         Counters &cnt) {
  V3 bmax = mk3(bmin.x + ext, bmin.y + ext, bmin.z + ext);
  float t1, t2;
  if (KNOWN_INSIDE) {
    const float nx = inv.x < 0 ? bmax.x : bmin.x, ny = inv.y < 0 ? bmax.y : bmin.y, nz = inv.z < 0 ? bmax.z : bmin.z;
    t1 = fmaxf(fmaxf((nx - o.x) * inv.x, (ny - o.y) * inv.y), fmaxf((nz - o.z) * inv.z, tNear));
  } else {
    bbox_intersection(bmin, bmax, o, inv, tNear, tFar, t1, t2);
    if (t1 > t2) return false;
  }
  const float4 a = __ldg(oc.values + 2 * (size_t)node), b = __ldg(oc.values + 2 * (size_t)node + 1);
  const float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
  float t = t1;
  V3 p = along(o, d, t);
  p.x = std_max(p.x, bmin.x); p.y = std_max(p.y, bmin.y); p.z = std_max(p.z, bmin.z);
  p.x = std_min(p.x, bmax.x); p.y = std_min(p.y, bmax.y); p.z = std_min(p.z, bmax.z);
  while (p.x <= bmax.x && p.y <= bmax.y && p.z <= bmax.z && p.x >= bmin.x && p.y >= bmin.y && p.z >= bmin.z) {
    if (COUNT) cnt.oct_leaf_samples++;
    float sd = oct_node_sdf(v, bmin, invExt, p);
    if (sd < 1e-4f) {
      r.hitten = true;
      r.t = t + sd;
      r.prim = node;
      pHit = p;
      return true;
    }
    t += sd;
    p = along(o, d, t);
  }
  return false;
}

// divide_box_8 + intersect_box_8 for one inner node (ray_pack.ispc:220-273): entry distance of the ray into each of
// the 8 child boxes, -1 on a miss.  The 8 child boxes only have three distinct planes per axis (min, centre, max:
// centre + half is max exactly), so the 48 plane distances are 9 distinct values and the per-axis min/max pairs 12:
// the same operations on the same operands as ray_pack.ispc:255-263, each evaluated once.
// NANFREE: all three 1/dir are finite, so no plane distance can be NaN (NaN needs 0 * inf) and the select-form
// min/max of ISPC (second operand on NaN) equals the hardware FMNMX up to the sign of a zero, which nothing
// downstream can observe (the distances are only compared with each other and with 0).
template <bool NANFREE>
__device__ __forceinline__ void oct_child_entries(V3 bmin, float ext, V3 o, V3 inv, float tNear, float tFar,
                                                  float (&ts)[8], uint32_t (&ch)[8]) {
  const float h = ext * 0.5f;
  const float pl[3][3] = {{bmin.x, bmin.x + h, bmin.x + ext}, {bmin.y, bmin.y + h, bmin.y + ext},
                          {bmin.z, bmin.z + h, bmin.z + ext}};
  const float oo[3] = {o.x, o.y, o.z}, ii[3] = {inv.x, inv.y, inv.z};
  float mn[3][2], mx[3][2];
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const float t0 = (pl[a][0] - oo[a]) * ii[a], t1 = (pl[a][1] - oo[a]) * ii[a], t2 = (pl[a][2] - oo[a]) * ii[a];
    mn[a][0] = mn2<NANFREE>(t0, t1); mx[a][0] = mx2<NANFREE>(t0, t1);
    mn[a][1] = mn2<NANFREE>(t1, t2); mx[a][1] = mx2<NANFREE>(t1, t2);
  }
#pragma unroll
  for (int id = 0; id < 8; ++id) {
    const int cx = (id >> 2) & 1, cy = (id >> 1) & 1, cz = id & 1;
    float tMin = mx2<NANFREE>(mn[0][cx], mx2<NANFREE>(mn[1][cy], mn[2][cz]));
    float tMax = mn2<NANFREE>(mx[0][cx], mn2<NANFREE>(mx[1][cy], mx[2][cz]));
    tMin = mx2<NANFREE>(tMin, tNear);
    tMax = mn2<NANFREE>(tMax, tFar);
    ts[id] = (tMax < 0 || tMin > tMax) ? -1.0f : tMin;
    ch[id] = id;
  }
}

// The same visit list without testing 8 boxes and sorting them.  With no NaN possible, tNear > 0 and no exact
// ties, the validity test and the sort of ray_pack.ispc:264-271 + raytracing.hpp:188-213 are a pure, exact
// function of the 9 plane distances (max/min select, they never round), and that function has a closed form:
// per axis the ray is in the near half on [A, M] and in the far half on [M, B] (A = min, B = max of the outer
// plane distances, M the centre plane's); inside the node's window [T_in, T_out] = [max(tNear, A's), min(tFar, B's)]
// the ray starts in the child whose bits are (M_a < T_in) and moves to a neighbour at every M_a inside the window,
// in increasing M_a.  Those <= 4 children are exactly the boxes whose slab test succeeds, their entry distances are
// strictly increasing (T_in-side first, then the M_a of each crossing), so sort8 would list them in this order.
// Any exact tie (two equal centre distances, or one equal to a window end) makes the closed form ambiguous: the
// caller then falls back to the literal 8-box test + sort8.  Returns false when the fallback is needed.
// (experiment switches, fixed by measurement: profiles/r2_octree_variants.txt)
#ifndef RT_OCT_KEYS
#define RT_OCT_KEYS 1     // crossings ordered through packed 32-bit keys (else a 3-comparator network on (t, axis) pairs)
#endif
#ifndef RT_OCT_MASK
#define RT_OCT_MASK 2     // kept-children mask: 0 unused, 1 tested when a child is taken, 2 applied when the list is built
#endif
#ifndef RT_OCT_STACK64
#define RT_OCT_STACK64 0  // one 64-bit stack entry per level (else two 32-bit arrays); measured slower (+1.3 %)
#endif
#ifndef RT_OCT_CFWORD
#define RT_OCT_CFWORD 0   // 32-bit arrays: "closed-form list" flag kept in bit 31 of the saved list (else a per-level bit mask)
#endif

// keep: bit c set = child c is worth a visit (everything else is a leaf intersectLeaf rejects from its values alone);
// FILTER: drop the others from the list right here (the kernels that count visits keep them and skip them one by one)
template <bool FILTER>
__device__ __forceinline__ bool oct_children_fast(V3 bmin, float ext, V3 o, V3 inv, uint32_t nearMask, float tNear,
                                                  float tFar, uint32_t keep, uint32_t &w) {
  const float h = ext * 0.5f;
  const float ax0 = (bmin.x - o.x) * inv.x, ax1 = (bmin.x + h - o.x) * inv.x, ax2 = (bmin.x + ext - o.x) * inv.x;
  const float ay0 = (bmin.y - o.y) * inv.y, ay1 = (bmin.y + h - o.y) * inv.y, ay2 = (bmin.y + ext - o.y) * inv.y;
  const float az0 = (bmin.z - o.z) * inv.z, az1 = (bmin.z + h - o.z) * inv.z, az2 = (bmin.z + ext - o.z) * inv.z;
  const float tIn = fmaxf(fmaxf(fminf(ax0, ax2), fminf(ay0, ay2)), fmaxf(fminf(az0, az2), tNear));
  const float tOut = fminf(fminf(fmaxf(ax0, ax2), fmaxf(ay0, ay2)), fminf(fmaxf(az0, az2), tFar));
  // crossings strictly inside the window; equal centre distances only matter when both are inside it (two
  // crossings before T_in just mean "far half on both axes", e.g. a camera on a coordinate axis at the root)
  const bool ex = tIn < ax1 && ax1 < tOut, ey = tIn < ay1 && ay1 < tOut, ez = tIn < az1 && az1 < tOut;
#if RT_OCT_KEYS
  // The crossings are ordered through ONE 32-bit key each: the distance's bit pattern (positive: tIn >= tNear > 0, so
  // the patterns order like the floats) with its two low mantissa bits replaced by the axis (0 -> x, 1 -> y, 2 -> z).
  // Two crossings whose distances agree in the upper 30 bits (equal, or within 3 ulp) count as a tie and take the
  // literal path, so the order taken here is always the exact one.  Crossings outside the window sort last.
  uint32_t k0 = ex ? (__float_as_uint(ax1) & ~3u) : 0xFFFFFFFFu;
  uint32_t k1 = ey ? ((__float_as_uint(ay1) & ~3u) | 1u) : 0xFFFFFFFFu;
  uint32_t k2 = ez ? ((__float_as_uint(az1) & ~3u) | 2u) : 0xFFFFFFFFu;
  key_cswap(k0, k1); key_cswap(k1, k2); key_cswap(k0, k1);
  const bool tie = ((k0 ^ k1) < 4u && k1 != 0xFFFFFFFFu) || ((k1 ^ k2) < 4u && k2 != 0xFFFFFFFFu) || ax1 == tIn ||
                   ay1 == tIn || az1 == tIn || ax1 == tOut || ay1 == tOut || az1 == tOut;
  if (tie) return false;
  const bool v0 = k0 != 0xFFFFFFFFu, v1 = k1 != 0xFFFFFFFFu, v2 = k2 != 0xFFFFFFFFu;
  const uint32_t m0 = 4u >> (k0 & 3u), m1 = 4u >> (k1 & 3u), m2 = 4u >> (k2 & 3u);
#else
  const bool tie = (ex && ey && ax1 == ay1) || (ex && ez && ax1 == az1) || (ey && ez && ay1 == az1) || ax1 == tIn ||
                   ay1 == tIn || az1 == tIn || ax1 == tOut || ay1 == tOut || az1 == tOut;
  if (tie) return false;
  const float INF = __int_as_float(0x7f800000);
  float k0 = ex ? ax1 : INF, k1 = ey ? ay1 : INF, k2 = ez ? az1 : INF;
  uint32_t m0 = 4u, m1 = 2u, m2 = 1u;
#define RT_CS3(ka, kb, ma, mb)            \
  if (ka > kb) {                          \
    float tf = ka; ka = kb; kb = tf;      \
    uint32_t tm = ma; ma = mb; mb = tm;   \
  }
  RT_CS3(k0, k1, m0, m1) RT_CS3(k1, k2, m1, m2) RT_CS3(k0, k1, m0, m1)
#undef RT_CS3
  const bool v0 = k0 < INF, v1 = k1 < INF, v2 = k2 < INF;
#endif
  w = 0;
  if (tIn > tOut) return true;  // the ray misses this node's box inside [tNear, tFar]: no child can be hit
  // child at the window start, in ray order (bit set = far half), x -> 4, y -> 2, z -> 1 as in divide_box_8
  uint32_t k = (ax1 < tIn ? 4u : 0u) | (ay1 < tIn ? 2u : 0u) | (az1 < tIn ? 1u : 0u);
  k ^= nearMask;  // ray order -> spatial child id
  if (FILTER) {
    uint32_t sh = 0;
    if ((keep >> k) & 1u) { w = 8u | k; sh = 4u; }
    if (v0) { k ^= m0; if ((keep >> k) & 1u) { w |= (8u | k) << sh; sh += 4u; } }
    if (v1) { k ^= m1; if ((keep >> k) & 1u) { w |= (8u | k) << sh; sh += 4u; } }
    if (v2) { k ^= m2; if ((keep >> k) & 1u) { w |= (8u | k) << sh; } }
  } else {
    w = 8u | k;
    if (v0) { k ^= m0; w |= (8u | k) << 4; }
    if (v1) { k ^= m1; w |= (8u | k) << 8; }
    if (v2) { k ^= m2; w |= (8u | k) << 12; }
  }
  return true;
}

// SDFOctree::intersectNode, octree_raytracing.cpp:166-202, as an explicit-stack ordered descent.
// The reference recursion returns as soon as ANY leaf reports a hit (the first hitting child ends every
// enclosing loop, :188-197), so `hit.t > t` is never evaluated with a hit in hand: the visit rule is just
// t > 0 in sort8 order.  Per level we keep one word: the not-yet-visited children in visit order, 4 bits each
// (bit 3 = valid), and the node's link word.  Child boxes come from integer (level, i, j, k): the dyadic
// subdivision of [-1,1] that divide_box_8 (ray_pack.ispc:220-239) computes is exact.
// ordered list of the children of the inner node with box (bmin, ext) that the ray visits: up to 8 nibbles,
// first to visit in the low nibble, bit 3 = valid, bits 0..2 = child id
template <bool FILTER>
__device__ __forceinline__ uint32_t oct_children(V3 bmin, float ext, V3 o, V3 inv, bool finiteInv, bool fastOk,
                                                 uint32_t nearMask, float tNear, float tFar, uint32_t keep,
                                                 bool &closedForm) {
  uint32_t w = 0;
  // closedForm: the list came from the closed form, so every child in it passed a NaN-free slab test
  closedForm = fastOk && oct_children_fast<FILTER>(bmin, ext, o, inv, nearMask, tNear, tFar, keep, w);
  if (closedForm) return w;
  // literal path: axis-parallel rays, tNear <= 0, exact ties
  float ts[8];
  uint32_t ch[8];
  if (finiteInv) oct_child_entries<true>(bmin, ext, o, inv, tNear, tFar, ts, ch);
  else oct_child_entries<false>(bmin, ext, o, inv, tNear, tFar, ts, ch);
  sort8(ts, ch);
  w = 0;
#pragma unroll
  for (int i = 7; i >= 0; --i)
    if (ts[i] > 0 && (!FILTER || ((keep >> ch[i]) & 1u))) w = (w << 4) | 8u | ch[i];
  return w;
}

template <bool COUNT>
__device__ __forceinline__ Hit oct_intersect(const OctreeDev &oc, V3 o, V3 d, float tNear, float tFar, Counters &cnt) {
  Hit r = hit_none();
  const V3 inv = mk3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
  const bool finiteInv = isfinite(inv.x) && isfinite(inv.y) && isfinite(inv.z);
  // closed-form child list needs: no NaN/inf plane distances, a sign on every axis, and tNear > 0 (so that every
  // box that passes the slab test also passes the reference's `t > 0` visit rule)
  const bool fastOk = finiteInv && inv.x != 0.0f && inv.y != 0.0f && inv.z != 0.0f && tNear > 0.0f && tFar >= tNear;
  const uint32_t nearMask = (inv.x < 0.0f ? 4u : 0u) | (inv.y < 0.0f ? 2u : 0u) | (inv.z < 0.0f ? 1u : 0u);
  V3 hitMin = mk3(0.0f, 0.0f, 0.0f), pHit = hitMin;
  float hitInvExt = 0.0f;
  const uint32_t rootLink = __ldg(oc.link);
  if (rootLink == OCT_LEAF_SURFACE || rootLink == OCT_LEAF_CULLED) {
    if (COUNT) cnt.oct_leaf++;
    hitMin = mk3(-1.0f, -1.0f, -1.0f);
    hitInvExt = 0.5f;
    if (rootLink == OCT_LEAF_SURFACE)
      oct_leaf<COUNT>(oc, 0, hitMin, 2.0f, 0.5f, o, d, inv, tNear, tFar, false, r, pHit, cnt);
  } else {
    // Iterate the children of the CURRENT inner node from registers; only stepping into an inner child saves
    // the parent's state (leaf children, the majority of visits, never touch the stack).
#if RT_OCT_STACK64
    // One 64-bit entry per level: low word = the parent's not-yet-visited children (<= 7 nibbles once a child has been
    // taken) with bit 31 = "that list is a closed-form one", high word = the parent's link word.
    unsigned long long stack[OCT_MAX_DEPTH + 1];
#else
    uint32_t pend[OCT_MAX_DEPTH + 1], base[OCT_MAX_DEPTH + 1];
#if !RT_OCT_CFWORD
    uint32_t closedMask = 0;  // bit l: the child list of the level-l node on the path is a closed-form one
#endif
#endif
    int level = 0;                     // depth of the current inner node
    V3 bmin = mk3(-1.0f, -1.0f, -1.0f);  // its box: dyadic, so parent/child corners are exact sums
    float ext = 2.0f, invExt = 0.5f;   // 1/ext: both powers of two, kept exact by doubling / halving
    unsigned long long path = 0;       // 3 bits per level: child ids on the way down (for stepping back up)
    const bool hasMask = RT_OCT_MASK != 0 && oc.has_mask != 0u;
    const uint32_t offMask = oc.has_mask ? OCT_OFFSET_MASK : 0xFFFFFFFFu;
    // the kernels that count visits take the culled children one by one (each is a leaf visit of the reference)
    constexpr bool FILTER = RT_OCT_MASK == 2 && !COUNT;
    uint32_t lw = rootLink;            // link word of the current node: childrenOffset (| kept-children mask << 24)
    if (COUNT) cnt.oct_inner++;
    bool cf;  // the current node's child list came from the closed form: its children passed a NaN-free slab test
    uint32_t w = oct_children<FILTER>(bmin, ext, o, inv, finiteInv, fastOk, nearMask, tNear, tFar,
                                      hasMask ? lw >> 24 : 0xFFu, cf);
#if !RT_OCT_STACK64 && !RT_OCT_CFWORD
    closedMask = cf ? 1u : 0u;
#endif
    // (A "while-while" form, all lanes walking to their next surface leaf before any of them marches, was measured
    // and rejected: octree6 4K 0.73 -> 0.75 ms, depth-8 stand-in 1.16 -> 1.34 ms; profiles/r1_tune.txt.  So was the
    // milder form of round 2, a lane parking its leaf until n = 4 .. 24 lanes of the warp can march together:
    // 1.058 -> 1.08 .. 1.17 ms, profiles/r2_octree_variants.txt.)
    for (;;) {
      if (w == 0) {
        // ---- back to the parent
        if (level == 0) break;
        --level;
#if RT_OCT_STACK64
        const unsigned long long e = stack[level];
        w = (uint32_t)e;
        lw = (uint32_t)(e >> 32);
        cf = (w >> 31) != 0u;
        w &= 0x7FFFFFFFu;
#else
        w = pend[level];
        lw = base[level];
#if RT_OCT_CFWORD
        cf = (w >> 31) != 0u;
        w &= 0x7FFFFFFFu;
#endif
#endif
        const uint32_t id = (uint32_t)path & 7u;
        path >>= 3;
        bmin = mk3(bmin.x - ((id & 4u) ? ext : 0.0f), bmin.y - ((id & 2u) ? ext : 0.0f), bmin.z - ((id & 1u) ? ext : 0.0f));
        ext *= 2.0f;
        invExt *= 0.5f;
        continue;
      }
      const uint32_t id = w & 7u;
      w >>= 4;
      // a child that is a leaf intersectLeaf rejects from its values alone (octree_raytracing.cpp:128-135) is known from
      // the parent's link word: visited (it counts as a leaf visit) without touching its record
      if (!FILTER && hasMask && !((lw >> (24u + id)) & 1u)) {
        if (COUNT) cnt.oct_leaf++;
        continue;
      }
      const uint32_t child = (lw & offMask) + id;
      const uint32_t link = __ldg(oc.link + child);
      const float h = ext * 0.5f;
      if (link == OCT_LEAF_SURFACE || link == OCT_LEAF_CULLED) {
        if (COUNT) cnt.oct_leaf++;
        if (link == OCT_LEAF_SURFACE) {
          const V3 cmin = mk3(bmin.x + ((id & 4u) ? h : 0.0f), bmin.y + ((id & 2u) ? h : 0.0f), bmin.z + ((id & 1u) ? h : 0.0f));
          const float ih = invExt * 2.0f;
#if RT_OCT_STACK64 || RT_OCT_CFWORD
          const bool known = cf;
#else
          const bool known = ((closedMask >> level) & 1u) != 0u;
#endif
          if (oct_leaf<COUNT>(oc, child, cmin, h, ih, o, d, inv, tNear, tFar, known, r, pHit, cnt)) {
            hitMin = cmin;
            hitInvExt = ih;
            break;  // the first leaf that reports a hit ends the whole descent (octree_raytracing.cpp:188-197)
          }
        }
      } else {
        if (COUNT) cnt.oct_inner++;
#if RT_OCT_STACK64
        stack[level] = ((unsigned long long)lw << 32) | (unsigned long long)(w | (cf ? 0x80000000u : 0u));
#else
        pend[level] = RT_OCT_CFWORD ? (w | (cf ? 0x80000000u : 0u)) : w;
        base[level] = lw;
#endif
        ++level;
        path = (path << 3) | id;
        bmin = mk3(bmin.x + ((id & 4u) ? h : 0.0f), bmin.y + ((id & 2u) ? h : 0.0f), bmin.z + ((id & 1u) ? h : 0.0f));
        ext = h;
        invExt *= 2.0f;
        lw = link;
        w = oct_children<FILTER>(bmin, ext, o, inv, finiteInv, fastOk, nearMask, tNear, tFar,
                                 hasMask ? lw >> 24 : 0xFFu, cf);
#if !RT_OCT_STACK64 && !RT_OCT_CFWORD
        closedMask = (closedMask & ~(1u << level)) | (cf ? (1u << level) : 0u);
#endif
      }
    }
  }
  if (r.hitten) {
    // the hit leaf's normal, evaluated by all lanes that hit together (values are an L1 hit)
    const float4 a = __ldg(oc.values + 2 * (size_t)r.prim), b = __ldg(oc.values + 2 * (size_t)r.prim + 1);
    const float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    r.normal = oct_node_normal(v, hitMin, hitInvExt, pHit);
  }
  return r;
}

// ---- the same descent as a resumable per-lane state machine (render_octree_compact_kernel) ------------------------
// oct_intersect above runs one ray to completion; here the loop body is exposed one child visit at a time so that a
// persistent warp can (a) hand a lane a NEW ray the moment its old one is done and (b) hold lanes that stand on a
// surface leaf until enough of them can march together.  Same functions, same operands, same order per ray.
struct OctWalk {
  V3 d, inv;
  float tFar;
  uint32_t flags;  // bit 0 finiteInv, bit 1 fastOk, bits 4..6 nearMask
  int level;
  V3 bmin;
  float ext, invExt;
  unsigned long long path;
  uint32_t closedMask, cbase, w;
  uint32_t pend[OCT_MAX_DEPTH + 1], base[OCT_MAX_DEPTH + 1];
  // pending surface leaf
  uint32_t leafNode;
  V3 leafMin;
  float leafExt, leafInvExt;
  bool leafClosed;
};
constexpr int OCT_WALK_CONTINUE = 0, OCT_WALK_LEAF = 1, OCT_WALK_DONE = 2;

// the root must be an inner node (the caller checks OctreeDev::root_inner)
template <bool COUNT>
__device__ __forceinline__ void oct_walk_begin(OctWalk &s, const OctreeDev &oc, V3 o, V3 d, float tNear, float tFar,
                                               Counters &cnt) {
  s.d = d;
  s.inv = mk3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
  s.tFar = tFar;
  const bool finiteInv = isfinite(s.inv.x) && isfinite(s.inv.y) && isfinite(s.inv.z);
  const bool fastOk = finiteInv && s.inv.x != 0.0f && s.inv.y != 0.0f && s.inv.z != 0.0f && tNear > 0.0f && tFar >= tNear;
  const uint32_t nearMask = (s.inv.x < 0.0f ? 4u : 0u) | (s.inv.y < 0.0f ? 2u : 0u) | (s.inv.z < 0.0f ? 1u : 0u);
  s.flags = (finiteInv ? 1u : 0u) | (fastOk ? 2u : 0u) | (nearMask << 4);
  s.level = 0;
  s.bmin = mk3(-1.0f, -1.0f, -1.0f);
  s.ext = 2.0f;
  s.invExt = 0.5f;
  s.path = 0;
  s.cbase = __ldg(oc.link);
  if (COUNT) cnt.oct_inner++;
  bool cf;
  s.w = oct_children<false>(s.bmin, s.ext, o, s.inv, finiteInv, fastOk, nearMask, tNear, tFar, 0xFFu, cf);
  s.closedMask = cf ? 1u : 0u;
}

// one iteration of oct_intersect's loop, without the leaf march
template <bool COUNT>
__device__ __forceinline__ int oct_walk_step(OctWalk &s, const OctreeDev &oc, V3 o, float tNear, Counters &cnt) {
  if (s.w == 0) {
    if (s.level == 0) return OCT_WALK_DONE;
    --s.level;
    s.w = s.pend[s.level];
    s.cbase = s.base[s.level];
    const uint32_t id = (uint32_t)s.path & 7u;
    s.path >>= 3;
    s.bmin = mk3(s.bmin.x - ((id & 4u) ? s.ext : 0.0f), s.bmin.y - ((id & 2u) ? s.ext : 0.0f),
                 s.bmin.z - ((id & 1u) ? s.ext : 0.0f));
    s.ext *= 2.0f;
    s.invExt *= 0.5f;
    return OCT_WALK_CONTINUE;
  }
  const uint32_t id = s.w & 7u;
  s.w >>= 4;
  const uint32_t child = (s.cbase & (oc.has_mask ? OCT_OFFSET_MASK : 0xFFFFFFFFu)) + id;
  const uint32_t link = __ldg(oc.link + child);
  const float h = s.ext * 0.5f;
  const V3 cmin = mk3(s.bmin.x + ((id & 4u) ? h : 0.0f), s.bmin.y + ((id & 2u) ? h : 0.0f), s.bmin.z + ((id & 1u) ? h : 0.0f));
  if (link == OCT_LEAF_SURFACE || link == OCT_LEAF_CULLED) {
    if (COUNT) cnt.oct_leaf++;
    if (link == OCT_LEAF_CULLED) return OCT_WALK_CONTINUE;
    s.leafNode = child;
    s.leafMin = cmin;
    s.leafExt = h;
    s.leafInvExt = s.invExt * 2.0f;
    s.leafClosed = ((s.closedMask >> s.level) & 1u) != 0u;
    return OCT_WALK_LEAF;
  }
  if (COUNT) cnt.oct_inner++;
  s.pend[s.level] = s.w;
  s.base[s.level] = s.cbase;
  ++s.level;
  s.path = (s.path << 3) | id;
  s.bmin = cmin;
  s.ext = h;
  s.invExt *= 2.0f;
  s.cbase = link;
  bool cf;
  s.w = oct_children<false>(s.bmin, s.ext, o, s.inv, (s.flags & 1u) != 0u, (s.flags & 2u) != 0u, (s.flags >> 4) & 7u,
                            tNear, s.tFar, 0xFFu, cf);
  s.closedMask = (s.closedMask & ~(1u << s.level)) | (cf ? (1u << s.level) : 0u);
  return OCT_WALK_CONTINUE;
}

// march the pending leaf; on a hit fills r (t, prim, normal) exactly as oct_intersect does
template <bool COUNT>
__device__ __forceinline__ bool oct_walk_leaf(const OctWalk &s, const OctreeDev &oc, V3 o, float tNear, Hit &r,
                                              Counters &cnt) {
  V3 pHit;
  if (!oct_leaf<COUNT>(oc, s.leafNode, s.leafMin, s.leafExt, s.leafInvExt, o, s.d, s.inv, tNear, s.tFar, s.leafClosed, r,
                       pHit, cnt))
    return false;
  const float4 a = __ldg(oc.values + 2 * (size_t)r.prim), b = __ldg(oc.values + 2 * (size_t)r.prim + 1);
  const float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
  r.normal = oct_node_normal(v, s.leafMin, s.leafInvExt, pHit);
  return true;
}

// =============================================================================================================
// Triangle mesh: 8-wide BVH (triangles_raytracing.cpp:266-335) and brute force (ray_pack.ispc:167-202)
// =============================================================================================================
// BVHBuilder::traverseNode as an explicit-stack closest-hit traversal.  Children are pushed far-to-near in
// sort8 order so they pop in the reference's visit order; the cull `hit && hit.t < t_entry` (:282-284) is
// evaluated at pop time against the best hit so far.  (The recursion culls against the best hit of the
// enclosing node only; a subtree it visits and we skip starts behind the current best hit, so it cannot
// change the closest hit — DESIGN.md "BVH culling".)  Ties keep the first hit in visit order (strict '>').
// Fast inner-node path (our own trees, all three 1/dir finite and non-zero, tNear > 0):
//  * the near / far plane of every child on each axis is picked by the SIGN of 1/dir instead of min/max of the two
//    plane distances: lo <= hi and rounding is monotonic, so min((lo-o)*inv, (hi-o)*inv) IS the near-plane product,
//    bit for bit.  The sign is a per-ray constant, so it only changes which 32-byte run of the SoA node is loaded.
//  * children are ordered by a single 32-bit key = entry distance with its 3 low mantissa bits replaced by the
//    child's lane (entry distances are >= tNear > 0, so their bit patterns order like the floats).  Closest-hit
//    results do not depend on the visit order (only on which triangles get tested, and the pop-time cull below uses a
//    key that is <= the exact entry distance, i.e. it culls a subset of what the exact value would); only exact-t
//    ties between triangles can resolve differently, and those already depend on the tree.
// Trees imported from the reference (rt_scene_create_mesh_bvh8) keep the literal slab test + sort8 network so
// that their visit order, and with it every tie, is the reference's own.

#ifndef RT_BVH_DEFER
#define RT_BVH_DEFER 1  // fast path: push (node, lane), fetch the child word at pop time (else fetch all at push time)
#endif
template <bool COUNT>
__device__ __forceinline__ Hit bvh_intersect(const MeshDev &m, V3 o, V3 d, float tNear, float tFar, Counters &cnt) {
  Hit r = hit_none();
  const V3 inv = mk3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
  const bool finiteInv = isfinite(inv.x) && isfinite(inv.y) && isfinite(inv.z);
  const bool fastNode = finiteInv && !m.literal_order && inv.x != 0.0f && inv.y != 0.0f && inv.z != 0.0f &&
                        tNear > 0.0f;
  // float4 index inside the 224-byte node of the near / far plane runs (lo.x 0, lo.y 2, lo.z 4, hi.x 6, hi.y 8, hi.z 10)
  const uint32_t nx4 = inv.x < 0.0f ? 6u : 0u, ny4 = inv.y < 0.0f ? 8u : 2u, nz4 = inv.z < 0.0f ? 10u : 4u;
  uint32_t stackRef[BVH_STACK];
  float stackT[BVH_STACK];
  int sp = 0;
  uint32_t bestTri = 0;
  uint32_t ref = 0;  // root is always an inner node
  for (;;) {
    if (!(ref & BVH_LEAF_BIT)) {
      if (COUNT) cnt.bvh_inner++;
      const float4 *n4 = reinterpret_cast<const float4 *>(m.nodes + ref);
      if (fastNode) {
        uint32_t key[8];
#pragma unroll
        for (int hlf = 0; hlf < 2; ++hlf) {
          const float4 ax = __ldg(n4 + nx4 + hlf), bx = __ldg(n4 + (6u - nx4) + hlf);
          const float4 ay = __ldg(n4 + ny4 + hlf), by = __ldg(n4 + (10u - ny4) + hlf);
          const float4 az = __ldg(n4 + nz4 + hlf), bz = __ldg(n4 + (14u - nz4) + hlf);
          const float nX[4] = {ax.x, ax.y, ax.z, ax.w}, fX[4] = {bx.x, bx.y, bx.z, bx.w};
          const float nY[4] = {ay.x, ay.y, ay.z, ay.w}, fY[4] = {by.x, by.y, by.z, by.w};
          const float nZ[4] = {az.x, az.y, az.z, az.w}, fZ[4] = {bz.x, bz.y, bz.z, bz.w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float tnx = (nX[j] - o.x) * inv.x, tny = (nY[j] - o.y) * inv.y, tnz = (nZ[j] - o.z) * inv.z;
            const float tfx = (fX[j] - o.x) * inv.x, tfy = (fY[j] - o.y) * inv.y, tfz = (fZ[j] - o.z) * inv.z;
            const float tMin = fmaxf(fmaxf(tnx, fmaxf(tny, tnz)), tNear);
            const float tMax = fminf(fminf(tfx, fminf(tfy, tfz)), tFar);
            // unused lanes hold the box (+inf, -inf) and miss here; tMin >= tNear > 0
            key[4 * hlf + j] = (tMax < 0 || tMin > tMax) ? 0xFFFFFFFFu
                                                         : ((__float_as_uint(tMin) & ~7u) | (uint32_t)(4 * hlf + j));
          }
        }
        // the 19-comparator network of raytracing.hpp:188-213 on the packed keys (misses sort last)
        key_cswap(key[0], key[1]); key_cswap(key[2], key[3]); key_cswap(key[4], key[5]); key_cswap(key[6], key[7]);
        key_cswap(key[0], key[2]); key_cswap(key[1], key[3]); key_cswap(key[4], key[6]); key_cswap(key[5], key[7]);
        key_cswap(key[1], key[2]); key_cswap(key[5], key[6]); key_cswap(key[0], key[4]); key_cswap(key[3], key[7]);
        key_cswap(key[1], key[5]); key_cswap(key[2], key[6]);
        key_cswap(key[1], key[4]); key_cswap(key[3], key[6]);
        key_cswap(key[2], key[4]); key_cswap(key[3], key[5]);
        key_cswap(key[3], key[4]);
#if !RT_BVH_DEFER
        const uint32_t *childWord = reinterpret_cast<const uint32_t *>(n4 + 12);
#endif
#pragma unroll
        for (int i = 7; i >= 0; --i) {
          if (key[i] != 0xFFFFFFFFu) {  // far-to-near pushes: the nearest child pops first
#if RT_BVH_DEFER
            stackRef[sp] = (ref << 3) | (key[i] & 7u);  // (node, lane): the child word is fetched when the entry pops
#else
            stackRef[sp] = __ldg(childWord + (key[i] & 7u));
#endif
            stackT[sp] = __uint_as_float(key[i] & ~7u);
            ++sp;
          }
        }
      } else {
      float b[48];
#pragma unroll
      for (int i = 0; i < 12; ++i) {
        float4 q = __ldg(n4 + i);
        b[4 * i] = q.x; b[4 * i + 1] = q.y; b[4 * i + 2] = q.z; b[4 * i + 3] = q.w;
      }
      const uint4 c0 = __ldg(reinterpret_cast<const uint4 *>(n4 + 12)), c1 = __ldg(reinterpret_cast<const uint4 *>(n4 + 13));
      const uint32_t child[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
      float t[8];
      uint32_t ch[8];  // sort8 payload = the child word itself (the reference sorts child indices)
      if (finiteInv) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
          t[i] = slab_lane<true>(b[i], b[8 + i], b[16 + i], b[24 + i], b[32 + i], b[40 + i], o, inv, tNear, tFar);
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i)
          t[i] = slab_lane<false>(b[i], b[8 + i], b[16 + i], b[24 + i], b[32 + i], b[40 + i], o, inv, tNear, tFar);
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (child[i] == BVH_EMPTY) t[i] = -1.0f;  // unused lanes behave as misses
        ch[i] = child[i];
      }
      sort8(t, ch);
#pragma unroll
      for (int i = 7; i >= 0; --i) {
        if (!(t[i] < 0)) {  // :282-284 skips t < 0 (slab_lane never returns NaN: tNear wins every max)
          stackRef[sp] = ch[i];  // depth of the stack is validated against BVH_STACK at scene creation
          stackT[sp] = t[i];
          ++sp;
        }
      }
      }
    } else {
      if (COUNT) cnt.bvh_leaf++;
      const uint32_t first = ref & 0x07FFFFFFu, count = ((ref >> 27) & 15u) + 1u;
      for (uint32_t k = 0; k < count; ++k) {
        if (COUNT) cnt.bvh_tris++;
        const float4 *tp = reinterpret_cast<const float4 *>(m.tris + first + k);
        const float4 a = __ldg(tp), bq = __ldg(tp + 1), cq = __ldg(tp + 2);
        float th;
        if (tri_lane(o, d, ld3(a), ld3(bq), ld3(cq), th) && (!r.hitten || r.t > th)) {
          r.hitten = true;
          r.t = th;
          bestTri = first + k;
        }
      }
    }
    // ---- pop the next subtree that can still hold a closer hit
    for (;;) {
      if (sp == 0) goto done;
      --sp;
      ref = stackRef[sp];
      if (!(r.hitten && r.t < stackT[sp])) break;
    }
#if RT_BVH_DEFER
    if (fastNode) ref = __ldg(reinterpret_cast<const uint32_t *>(reinterpret_cast<const float4 *>(m.nodes + (ref >> 3)) + 12) + (ref & 7u));
#endif
  }
done:
  if (r.hitten) {
    const float4 *tp = reinterpret_cast<const float4 *>(m.tris + bestTri);
    const float4 a = __ldg(tp), bq = __ldg(tp + 1), cq = __ldg(tp + 2);
    r.normal = ispc_normalize(cross3(sub3(ld3(bq), ld3(a)), sub3(ld3(cq), ld3(a))));  // ray_pack.ispc:142
    r.prim = __float_as_uint(a.w);
  }
  return r;
}

// Every triangle against the ray, triangles staged in shared memory by the kernel (smem = n_tris x 3 float4).
template <bool COUNT>
__device__ __forceinline__ Hit brute_intersect(const MeshDev &m, const float4 *smemTris, V3 o, V3 d, float tNear,
                                               float tFar, Counters &cnt) {
  Hit r = hit_none();
  const uint32_t nTris = m.n_tris;
  // BRUTE_ROOT_LEAF (<= 8 triangles): the reference's tree is one leaf, and traverseNode(0) on a leaf tests its
  // triangles with no box test and no range check at all (triangles_raytracing.cpp:292-331): nothing to cull here.
  // BRUTE_BOXED (explicit RT_MESH_BRUTE on 9+ triangles): the reference only reaches triangles through child boxes that
  // pass intersect_box_8 against [tNear, tFar] (:271-291); all of them lie inside the mesh's bounding box, lane 0 of
  // the single node such a scene carries, so a ray that fails the same test on it hits nothing there, and `th <= tFar`
  // stands in for the leaf boxes' own cull.  Tree-independent views only (rt_scene.h).
  const bool boxed = m.brute == BRUTE_BOXED;
  if (boxed) {
    const BvhNode *n = m.nodes;
    const V3 inv = mk3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    if (slab_lane<false>(__ldg(&n->lo[0][0]), __ldg(&n->lo[1][0]), __ldg(&n->lo[2][0]), __ldg(&n->hi[0][0]),
                         __ldg(&n->hi[1][0]), __ldg(&n->hi[2][0]), o, inv, tNear, tFar) < 0.0f)
      return r;
  }
  uint32_t best = 0;
  for (uint32_t k = 0; k < nTris; ++k) {
    if (COUNT) cnt.bvh_tris++;
    const float4 a = smemTris[3 * k], bq = smemTris[3 * k + 1], cq = smemTris[3 * k + 2];
    float th;
    if (tri_lane(o, d, ld3(a), ld3(bq), ld3(cq), th) && (!boxed || th <= tFar) && (!r.hitten || r.t > th)) {
      r.hitten = true;
      r.t = th;
      best = k;
    }
  }
  if (r.hitten) {
    const float4 a = smemTris[3 * best], bq = smemTris[3 * best + 1], cq = smemTris[3 * best + 2];
    r.normal = ispc_normalize(cross3(sub3(ld3(bq), ld3(a)), sub3(ld3(cq), ld3(a))));
    r.prim = __float_as_uint(a.w);
  }
  return r;
}

// =============================================================================================================
// Plane({0,1,0}, planeY) (raytracing.hpp:119-186) and SceneUnion (raytracing.hpp:83-97)
// =============================================================================================================
__device__ __forceinline__ Hit plane_intersect(float planeY, V3 o, V3 d, float tNear, float tFar) {
  Hit r = hit_none();
  const V3 n = mk3(0.0f, 1.0f, 0.0f);
  // recalcBasis() for this normal: b1 = normalize({0, n.z, -n.y}), b2 = normalize(cross(b1, n))
  const V3 b1 = lm_normalize(mk3(0.0f, n.z, -n.y));
  const V3 b2 = lm_normalize(cross3(b1, n));
  float dividor = dot3(d, n);
  if (fabsf(dividor) < 1e-8f) return r;
  float t = (planeY - dot3(o, n)) / dividor;
  if (t < tNear || t > tFar) return r;
  V3 p = along(o, d, t);
  int x = (int)ceilf(dot3(p, b1));
  int y = (int)ceilf(dot3(p, b2));
  r.hitten = true;
  r.t = t;
  r.normal = n;
  r.albedo = ((x + y) % 2 == 0) ? 0.0f : 1.0f;
  r.reflectiveness = 0.3f;
  r.prim = RT_PRIM_PLANE;
  return r;
}
