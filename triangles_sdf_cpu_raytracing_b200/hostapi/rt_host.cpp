// C hooks into the C++ host mirror (include/rt_b200_host.h). Compiled with -ffp-contract=off.
#include <cstdlib>
#include <cstring>

#include "../../include/rt_b200_host.h"
#include "scenes.hpp"

extern "C" {

int rt_host_camera(const float eye[3], const float target[3], const float up[3], int lock_up, int n_rotate, float dx,
                   float dy, float pos[3], float view[16], float view_inv[16]) {
  Camera cam({eye[0], eye[1], eye[2]}, {target[0], target[1], target[2]}, {up[0], up[1], up[2]});
  cam.setLockUp(lock_up != 0);
  for (int i = 0; i < n_rotate; ++i) cam.rotate(dx, dy);
  rtmath::float3 p = cam.position();
  pos[0] = p.x; pos[1] = p.y; pos[2] = p.z;
  rtmath::float4x4 v = cam.lookAtMatrix(), vi = rtmath::inverse4x4(v);
  std::memcpy(view, v.data(), 64);
  std::memcpy(view_inv, vi.data(), 64);
  return 0;
}

int rt_host_projection(float fovy, float aspect, float z_near, float z_far, float proj[16], float proj_inv[16]) {
  rtmath::float4x4 p = rtmath::perspectiveMatrix(fovy, aspect, z_near, z_far), pi = rtmath::inverse4x4(p);
  std::memcpy(proj, p.data(), 64);
  std::memcpy(proj_inv, pi.data(), 64);
  return 0;
}

int rt_host_load_obj(const char *path, int scale, float **vpos4, uint32_t *n_verts, uint32_t **indices,
                     uint32_t *n_indices) {
  try {
    cmesh4::SimpleMesh m = scale ? loadAndScale(path) : cmesh4::LoadMeshFromObj(path, false);
    if (m.indices.empty()) return -1;
    *n_verts = static_cast<uint32_t>(m.vPos4f.size());
    *n_indices = static_cast<uint32_t>(m.indices.size());
    *vpos4 = static_cast<float *>(std::malloc(m.vPos4f.size() * 16));
    *indices = static_cast<uint32_t *>(std::malloc(m.indices.size() * 4));
    std::memcpy(*vpos4, m.vPos4f.data(), m.vPos4f.size() * 16);
    std::memcpy(*indices, m.indices.data(), m.indices.size() * 4);
    return 0;
  } catch (const std::exception &) {
    return -1;
  }
}

void rt_host_free(void *p) { std::free(p); }

}  // extern "C"
