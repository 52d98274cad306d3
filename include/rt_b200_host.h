/* rt_b200_host.h — C hooks into the C++ host mirror (triangles_sdf_cpu_raytracing_b200/hostapi/), exported by the
 * same library so that tests (and bindings that cannot include C++ headers) can reach the host-side pieces of the
 * drop-in: the orbit camera (src/camera.hpp, camera.cpp), the projection set-up of the viewer (src/main.cpp:198-201)
 * and the OBJ loading + normalisation (src/core/mesh.cpp:178-286, src/main.cpp:326-343).  Pure host code, no CUDA. */
#ifndef RT_B200_HOST_H
#define RT_B200_HOST_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* Camera(eye, target, up); setLockUp(lock_up); n_rotate x rotate(dx, dy)  ->  position(), lookAtMatrix(),
 * inverse4x4(lookAtMatrix()) (column-major 16 floats each). */
int rt_host_camera(const float eye[3], const float target[3], const float up[3], int lock_up, int n_rotate, float dx,
                   float dy, float pos[3], float view[16], float view_inv[16]);
/* perspectiveMatrix(fovy, aspect, z_near, z_far) and its inverse4x4 (main.cpp:198-201). */
int rt_host_projection(float fovy, float aspect, float z_near, float z_far, float proj[16], float proj_inv[16]);
/* LoadMeshFromObj (+ loadAndScale when scale != 0). The arrays are owned by the library: rt_host_free them. */
int rt_host_load_obj(const char *path, int scale, float **vpos4, uint32_t *n_verts, uint32_t **indices,
                     uint32_t *n_indices);
void rt_host_free(void *p);

#ifdef __cplusplus
}
#endif
#endif
