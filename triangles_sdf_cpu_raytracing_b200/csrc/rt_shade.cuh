// Per-pixel shading shared by every render kernel: IScene::intersect dispatch, SceneUnion with the ground plane, the eye
// ray, Renderer::intersectionColor (raytracing.cpp:13-65) and the colour pack.  Included inside `namespace rt::<flavour>`
// right after rt_device.cuh; nothing here touches threadIdx, shared memory or any other launch state, so the same text
// also compiles as plain C++ (tests/host_emul/, where the parity tests run it on the CPU against the oracle).

constexpr int KIND_GRID = 0, KIND_OCTREE = 1, KIND_BVH = 2, KIND_BRUTE = 3, KIND_GRID_BRICKS = 4;
constexpr int KIND_GRID_PF = 5;  // the linear grid with prefetch along long marches (small launches, rt_device.cuh)

// IScene::intersect of the model, and of SceneUnion(model, plane) (raytracing.hpp:83-97)
// PIN: the inlined primary-ray call (rt_device.cuh, grid_intersect)
template <int KIND, bool COUNT, bool PIN = false>
__device__ __forceinline__ Hit model_intersect(const SceneDev &s, const float4 *smemTris, V3 o, V3 d, float tNear,
                                               float tFar, Counters &cnt) {
  if (COUNT) cnt.rays++;
  if (KIND == KIND_GRID) return grid_intersect<false, COUNT, false, PIN>(s.grid, o, d, tNear, tFar, cnt);
  if (KIND == KIND_GRID_PF) return grid_intersect<false, COUNT, true, PIN>(s.grid, o, d, tNear, tFar, cnt);
  if (KIND == KIND_GRID_BRICKS) return grid_intersect<true, COUNT>(s.grid, o, d, tNear, tFar, cnt);
  if (KIND == KIND_OCTREE) return oct_intersect<COUNT>(s.oct, o, d, tNear, tFar, cnt);
  if (KIND == KIND_BVH) return bvh_intersect<COUNT>(s.mesh, o, d, tNear, tFar, cnt);
  return brute_intersect<COUNT>(s.mesh, smemTris, o, d, tNear, tFar, cnt);
}

template <int KIND, bool COUNT, bool PIN = false>
__device__ __forceinline__ Hit scene_intersect(const SceneDev &s, const float4 *smemTris, bool hasPlane, float planeY,
                                               V3 o, V3 d, float tNear, float tFar, Counters &cnt) {
  Hit a = model_intersect<KIND, COUNT, PIN>(s, smemTris, o, d, tNear, tFar, cnt);
  if (!hasPlane) return a;
  Hit b = plane_intersect(planeY, o, d, tNear, tFar);
  return (a.t < b.t) ? a : b;
}

// the secondary-ray variant calls through one out-of-line copy so the traversal code is not inlined 4 times
template <int KIND, bool COUNT>
__device__ __noinline__ Hit scene_intersect_shared(const SceneDev &s, const float4 *smemTris, bool hasPlane,
                                                   float planeY, V3 o, V3 d, float tNear, float tFar, Counters &cnt) {
  return scene_intersect<KIND, COUNT>(s, smemTris, hasPlane, planeY, o, d, tNear, tFar, cnt);
}

// raytracing.cpp:82-88 + EyeRayDir4f (LiteMath.h:2121-2134) + mat4_colmajor_mul_vec4 (:1371-1377).
// (x, y) are the LOOP coordinates of Renderer::draw: y = 0 is the bottom of the view.
__device__ __forceinline__ V3 eye_ray(const FrameDev &f, int x, int y) {
  float fx = (float)x + 0.5f, fy = (float)y + 0.5f;
  float w = (float)f.W, h = (float)f.H;
  float V0 = 2.0f * (fx + 0.5f) / w - 1.0f, V1 = -2.0f * (fy + 0.5f) / h + 1.0f, V2 = 0.0f, V3_ = 1.0f;
  const float *B = f.proj_inv;
  float pos[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) pos[i] = V0 * B[i] + V1 * B[4 + i] + V2 * B[8 + i] + V3_ * B[12 + i];
  float pw = pos[3];
#pragma unroll
  for (int i = 0; i < 4; ++i) pos[i] = pos[i] / pw;
  pos[1] *= (-1.0f);
  float lenInv = 1.0f / sqrtf(pos[0] * pos[0] + pos[1] * pos[1] + pos[2] * pos[2]);
  float D0 = pos[0] * lenInv, D1 = pos[1] * lenInv, D2 = pos[2] * lenInv, D3 = 0.0f;
  const float *M = f.view_inv;
  return mk3(D0 * M[0] + D1 * M[4] + D2 * M[8] + D3 * M[12], D0 * M[1] + D1 * M[5] + D2 * M[9] + D3 * M[13],
             D0 * M[2] + D1 * M[6] + D2 * M[10] + D3 * M[14]);
}

struct C4 {
  float r, g, b, a;
};
// LiteMath.h:1741-1746 + to_uint32 :1153 (truncation)
__device__ __forceinline__ uint32_t color_pack_rgba(C4 c) {
  uint32_t r = (uint32_t)(c.r * 255.0f), g = (uint32_t)(c.g * 255.0f), b = (uint32_t)(c.b * 255.0f),
           a = (uint32_t)(c.a * 255.0f);
  return (a << 24) | (b << 16) | (g << 8) | r;
}

// Lambert branch of intersectionColor for one hit (raytracing.cpp:33-50); `visible` = shadow-ray result
__device__ __forceinline__ C4 lambert_color(const Hit &hit, V3 point, V3 light, bool visible) {
  C4 c;
  if (!visible) {
    c.r = c.g = c.b = hit.albedo * 0.1f;
  } else {
    V3 ldir = lm_normalize(sub3(point, light));
    float lam = std_max(dot3(mk3(-ldir.x, -ldir.y, -ldir.z), hit.normal), 0.0f);
    c.r = c.g = c.b = std_min(hit.albedo * 0.1f + lam * hit.albedo, 1.0f);
  }
  c.a = 1.0f;
  return c;
}

// Renderer::intersectionColor (raytracing.cpp:13-65) for a primary ray. Returns false on a miss.
template <int KIND, bool COUNT, bool SECONDARY>
__device__ __forceinline__ bool intersection_color(const SceneDev &s, const FrameDev &f, const float4 *smemTris, V3 o,
                                                   V3 d, float tPrev, C4 &color, float &tOut, uint32_t &primOut,
                                                   Counters &cnt) {
  const bool plane = f.has_plane != 0;
  Hit hit = SECONDARY ? scene_intersect_shared<KIND, COUNT>(s, smemTris, plane, f.plane_y, o, d, f.t_near,
                                                           std_min(f.t_far, tPrev), cnt)
                      : scene_intersect<KIND, COUNT, true>(s, smemTris, plane, f.plane_y, o, d, f.t_near,
                                                     std_min(f.t_far, tPrev), cnt);
  if (!hit.hitten) return false;
  if (COUNT) cnt.hits++;
  if (dot3(hit.normal, d) > 0) hit.normal = scale3(hit.normal, -1.0f);
  if (f.mode == RT_SHADE_NORMAL) {
    color.r = (hit.normal.x + 1.0f) / 2.0f;
    color.g = (hit.normal.y + 1.0f) / 2.0f;
    color.b = (hit.normal.z + 1.0f) / 2.0f;
    color.a = (1.0f + 1.0f) / 2.0f;
  } else if (f.mode == RT_SHADE_COLOR) {
    color.r = color.g = color.b = hit.albedo;
    color.a = 1.0f;
  } else {
    const V3 light = mk3(f.light[0], f.light[1], f.light[2]);
    const V3 point = along(o, d, hit.t);
    bool visible = true;
    if (SECONDARY && f.shadows) {
      V3 sdir = lm_normalize(sub3(light, point));
      Hit sh = scene_intersect_shared<KIND, COUNT>(s, smemTris, plane, f.plane_y, along(point, sdir, 0.3f), sdir, 0.01f,
                                                   100.0f, cnt);
      visible = !sh.hitten;
    }
    color = lambert_color(hit, point, light, visible);
    if (SECONDARY && f.reflections && hit.reflectiveness > 0.0f) {
      // reflect(dir, n) = n * dot(dir, n) * (-2) + dir, LiteMath.h:1196; one bounce (maxDepth 2)
      float dn = dot3(d, hit.normal);
      V3 rdir = lm_normalize(mk3(hit.normal.x * dn * (-2.0f) + d.x, hit.normal.y * dn * (-2.0f) + d.y,
                                 hit.normal.z * dn * (-2.0f) + d.z));
      V3 ro = along(point, rdir, 0.02f);
      C4 rc;
      rc.r = rc.g = rc.b = 0.0f;
      rc.a = 1.0f;
      Hit h2 = scene_intersect_shared<KIND, COUNT>(s, smemTris, plane, f.plane_y, ro, rdir, 0.01f,
                                                   std_min(100.0f, __int_as_float(0x7f800000)), cnt);
      if (h2.hitten) {
        if (COUNT) cnt.hits++;
        if (dot3(h2.normal, rdir) > 0) h2.normal = scale3(h2.normal, -1.0f);
        V3 p2 = along(ro, rdir, h2.t);
        bool vis2 = true;
        if (f.shadows) {
          V3 sdir = lm_normalize(sub3(light, p2));
          Hit sh = scene_intersect_shared<KIND, COUNT>(s, smemTris, plane, f.plane_y, along(p2, sdir, 0.3f), sdir,
                                                       0.01f, 100.0f, cnt);
          vis2 = !sh.hitten;
        }
        rc = lambert_color(h2, p2, light, vis2);
      }
      const float k = hit.reflectiveness, ik = 1.0f - hit.reflectiveness;
      color.r = color.r * ik + k * rc.r;
      color.g = color.g * ik + k * rc.g;
      color.b = color.b * ik + k * rc.b;
      color.a = color.a * ik + k * rc.a;
    }
  }
  tOut = hit.t;
  primOut = hit.prim;
  return true;
}

