/* rt_b200.h — C ABI of the B200 (sm_100a) primary-ray hot path.
 *
 * This is the drop-in boundary for ONE path of iMacsimus/Triangles-SDF-CPU-RayTracing: everything that
 * happens between `Renderer::draw(scene, frameBuffer, camera, projInv)` (src/raytracing.cpp:67-102) and
 * the framebuffer writes, for the three scene types (triangles_raytracing.cpp, grid_raytracing.cpp,
 * octree_raytracing.cpp) plus the ISPC primitives they call (ray_pack.ispc).  All `file:line` citations are
 * into the reference tree.  Plain pointers and sizes only; no C++ / torch types; no exceptions cross it.
 *
 * Conventions
 *   - every entry point returns RT_OK (0) or a negative rt_status; rt_last_error() gives the text
 *     (the reference has no error reporting on this path);
 *   - matrices are 16 floats, column-major, i.e. the memory image of LiteMath::float4x4::m_col
 *     (external/LiteMath/LiteMath.h:1396-1462);
 *   - framebuffers are row-major, index row*W + x with memory row 0 = TOP of the image, exactly
 *     FrameBuffer::color / FrameBuffer::t (src/raytracing.hpp:9-20, Image2d.h:91);
 *   - the depth buffer is IN/OUT: a pixel is overwritten only by a hit closer than the value it holds
 *     (raytracing.cpp:17,90-94); "cleared" means colour 0 and t = +inf (raytracing.hpp:16-19);
 *   - a scene lives on the CUDA device that was current when it was created; it is immutable afterwards
 *     and may be rendered from any host thread.
 *   - there is NO CPU fallback: without a CUDA device every call fails with RT_ERR_CUDA.
 */
#ifndef RT_B200_H
#define RT_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum rt_status {
  RT_OK = 0,
  RT_ERR_INVALID = -1, /* bad argument / malformed scene data (e.g. octree child offset out of range) */
  RT_ERR_CUDA = -2,    /* CUDA runtime failure, including "no device" */
  RT_ERR_UNSUPPORTED = -3
} rt_status;

/* ShadingMode, src/raytracing.hpp:99 (same order). */
typedef enum rt_shading_mode { RT_SHADE_NORMAL = 0, RT_SHADE_LAMBERT = 1, RT_SHADE_COLOR = 2 } rt_shading_mode;

/* Arithmetic flavour of the kernels.
 * RT_MATH_EXACT: every FP32 operation is issued separately, in the reference's operator order (no FMA
 *   contraction), IEEE div/sqrt: bit-identical to the reference built with -ffp-contract=off.
 * RT_MATH_FAST : same algorithm, the compiler may contract a*b+c into FMA (what GCC/ISPC do by default
 *   in the reference's Release build). */
typedef enum rt_math_mode { RT_MATH_EXACT = 0, RT_MATH_FAST = 1 } rt_math_mode;

/* Ray-to-thread scheduling of the render kernels.
 * RT_SCHED_TILES     : one thread per pixel, 8x4-pixel warp tiles, one launch covers the frame.
 * RT_SCHED_PERSISTENT: persistent warps pull 8x4 tiles from an atomic counter, so warps whose rays all end
 *   early pick up new work at once.
 * RT_SCHED_COMPACT   : persistent warps as a pool of 32 ray slots: finished lanes are refilled from the pixel
 *   counter (ballot + popc, one warp-aggregated atomic) and lanes standing on a leaf march in batches.  Octree
 *   scenes without plane / secondary rays; anything else runs as RT_SCHED_PERSISTENT.  Bit-identical output;
 *   measured slower than PERSISTENT on B200, kept selectable.
 * RT_SCHED_AUTO      : the per-scene-type default chosen by measurement. */
typedef enum rt_sched_mode {
  RT_SCHED_AUTO = 0,
  RT_SCHED_TILES = 1,
  RT_SCHED_PERSISTENT = 2,
  RT_SCHED_COMPACT = 3
} rt_sched_mode;

/* Dense-grid storage in HBM. LINEAR = the on-disk order (x-major, z fastest, grid_raytracing.hpp:13-15);
 * BRICKS = 8x8x8-sample bricks so the 8 corners of a trilinear cell share few 128-byte lines. */
typedef enum rt_grid_layout { RT_GRID_AUTO = 0, RT_GRID_LINEAR = 1, RT_GRID_BRICKS = 2 } rt_grid_layout;

/* Triangle-scene traversal. BVH8 = our 8-wide BVH; BRUTE = every ray tests every triangle from shared
 * memory (the loop ray_pack.ispc:167-202 describes; only for small meshes). */
typedef enum rt_mesh_accel {
  RT_MESH_AUTO = 0,
  RT_MESH_BVH8 = 1,     /* binned-SAH tree built on the host, collapsed to 8-wide */
  RT_MESH_BRUTE = 2,
  RT_MESH_LBVH_GPU = 3  /* linear BVH (Morton order) built on the device, collapsed to 8-wide: milliseconds for 10^6 triangles */
} rt_mesh_accel;

typedef enum rt_scene_kind { RT_SCENE_GRID = 0, RT_SCENE_OCTREE = 1, RT_SCENE_MESH = 2 } rt_scene_kind;

#define RT_PRIM_MISS 0xFFFFFFFFu  /* prim id of a pixel that was not written */
#define RT_PRIM_PLANE 0xFFFFFFFEu /* the checkerboard ground plane (src/raytracing.hpp:119-186) */

typedef struct rt_scene rt_scene; /* opaque */

/* Everything Renderer::draw derives per frame (raytracing.cpp:73-75) plus the Renderer's public knobs
 * (raytracing.hpp:103-106) and the part of the frame this call renders. */
typedef struct rt_frame {
  float ray_pos[3];   /* camera.position() */
  float view_inv[16]; /* inverse4x4(camera.lookAtMatrix()) */
  float proj_inv[16]; /* the projInv argument of draw */
  uint32_t width, height; /* full frame */
  /* Screen-space partition (multi-GPU tiles): this call renders the memory rows r with
   * (r / band_rows) % band_count == band_index.  {0,0,0} or band_count <= 1 means the whole frame. */
  uint32_t band_rows, band_count, band_index;
  int32_t shading_mode;     /* rt_shading_mode */
  float light_pos[3];       /* Renderer::lightPos (main.cpp:60: {2,2,2}) */
  int32_t enable_shadows;   /* Renderer::enableShadows   (secondary rays) */
  int32_t enable_reflections; /* Renderer::enableReflections (only the plane reflects) */
  int32_t has_plane;        /* SceneUnion(model, Plane({0,1,0}, plane_y)), main.cpp:189-190 */
  float plane_y;
  float t_near, t_far;      /* raytracing.cpp:90 passes 0.01, 100; 0,0 selects those defaults */
  /* 1: FrameBuffer::clear() (raytracing.hpp:16-19) is fused into the render: the previous depth is taken as +inf
   * and pixels the scene does not cover are written with the clear values (colour 0, t = +inf, prim MISS), so one
   * launch writes every pixel of the call's rows exactly once.  0: the in/out behaviour of Renderer::draw. */
  int32_t clear;
} rt_frame;

/* per-frame workload counters (filled when requested; sums over all rays of the call) */
typedef struct rt_counters {
  uint64_t rays, hits;
  uint64_t grid_samples;                       /* SDFGrid::sdf(float3) evaluations */
  uint64_t oct_inner, oct_leaf, oct_leaf_samples;
  uint64_t bvh_inner, bvh_leaf, bvh_tris;
} rt_counters;

/* ---- library / device ------------------------------------------------------------------------- */
const char *rt_last_error(void);         /* thread-local text of the last failure */
const char *rt_version(void);
int rt_device_count(int *count);         /* number of visible CUDA devices */
int rt_set_device(int device);           /* cudaSetDevice for the calling thread */
/* Multi-GPU behind the reference's single call (src/main.cpp:203,206 -> Renderer::draw, src/raytracing.cpp:67-102: one
 * call from one process renders the frame).  After rt_init(devices, n) every rt_scene_create_* builds the scene once (on
 * devices[0]) and replicates it onto the other listed devices with device-to-device copies, and ONE rt_render_host /
 * rt_render_device call on such a scene fans the frame out over all of them as interleaved 8-row bands:
 *   - rt_render_host, page-locked framebuffer: every GPU stores (clear = 0: first loads) its own bands straight in the
 *     caller's buffers over its OWN PCIe link (128-byte bulk transfers), so n GPUs bring n links to the transfer that
 *     bounds the single-GPU call; pageable buffers: one strided copy per device, buffer and direction;
 *   - rt_render_device, framebuffer in the HBM of one listed device: that device renders its bands on the caller's
 *     stream, the others store theirs into it over NVLink / NVSwitch inside their render kernels (peer access is
 *     enabled here), ordered with the caller's stream by events.  No separate gather, no NCCL.
 * A call whose rt_frame selects bands itself (band_count > 1) runs on one device as before.  n = 0 restores the
 * default: scenes live on the device that is current when they are created.  Call it before creating scenes. */
int rt_init(const int *devices, int n);
int rt_device_list(int *devices, int cap);                           /* -> number of devices of rt_init's list */

/* ---- scene construction (upload once) ---------------------------------------------------------- */
/* replaces SDFGrid + loadSDFGrid (grid_raytracing.hpp:10-22, .cpp:127-134): values[(x*sy + y)*sz + z] */
int rt_scene_create_grid(const float *values, uint32_t sx, uint32_t sy, uint32_t sz, int layout /*rt_grid_layout*/,
                         rt_scene **out);
/* replaces SDFOctree + loadSDFOctree (octree_raytracing.hpp:8-49, .cpp:8-16): n x 36-byte
 * {float values[8]; uint32 childrenOffset}.  Offsets and depth (<= 21 levels) are validated (a bad offset would be an
 * out-of-bounds device read).  Leaves may be shared between parents; an inner node referenced by two parents (a DAG
 * with shared subtrees, which the reference's loader would accept) is rejected with RT_ERR_INVALID, as a cycle would be. */
int rt_scene_create_octree(const void *nodes36, uint32_t n_nodes, rt_scene **out);
/* replaces BVHBuilder::perform (triangles_raytracing.hpp:40, .cpp:227-258): builds our own 8-wide BVH on the
 * host and uploads it with the leaf triangles pre-gathered.  vpos4 = SimpleMesh::vPos4f (x,y,z,w),
 * indices = SimpleMesh::indices (3 per triangle); prim id = triangle index in this order. */
int rt_scene_create_mesh(const float *vpos4, uint32_t n_verts, const uint32_t *indices, uint32_t n_indices,
                         int accel /*rt_mesh_accel*/, rt_scene **out);
/* same, but traverses a BVH8 the reference's own BVHBuilder produced: nodes = n_nodes x node_bytes (204)
 * images of BVH8Node (triangles_raytracing.hpp:10-36), indices in the builder's permuted order, and
 * prim_of_permuted[i] = original id of permuted triangle i (NULL: identity). */
int rt_scene_create_mesh_bvh8(const float *vpos4, uint32_t n_verts, const uint32_t *indices_bvh, uint32_t n_indices,
                              const void *nodes, uint32_t n_nodes, uint32_t node_bytes,
                              const uint32_t *prim_of_permuted, rt_scene **out);
/* mesh scenes: size of the tree and, for RT_MESH_LBVH_GPU, the device time of its construction (0 otherwise).
 * Any output pointer may be NULL. */
int rt_scene_mesh_stats(const rt_scene *scene, uint32_t *n_nodes, uint32_t *n_tris, float *device_build_ms);
/* device time (ms) of the on-device part of scene construction: octree split + validation, RT_MESH_LBVH_GPU build */
int rt_scene_build_ms(const rt_scene *scene, float *device_ms);
void rt_scene_destroy(rt_scene *scene);
int rt_scene_kind_of(const rt_scene *scene);             /* rt_scene_kind */
int rt_scene_device(const rt_scene *scene);               /* the device the scene was built on (replica 0) */
int rt_scene_replicas(const rt_scene *scene, int *devices, int cap); /* -> number of devices holding the scene */
/* bytes of scene data resident in HBM, and the model's bounding box (meshes: calc_bbox, raytracing.hpp:22-37;
 * SDF scenes: [-1,1]^3) */
int rt_scene_info(const rt_scene *scene, uint64_t *device_bytes, float bbox_min[3], float bbox_max[3]);
/* Transfer accounting of the in/out framebuffer contract (rt_render_host on pinned buffers, rt_render_device on a peer
 * framebuffer): running totals, since the scene was created and over all its devices, of the framebuffer bytes the render
 * kernels fetched (previous depth of the 32 x 4-pixel tiles in which a ray hits anything at all; previous colour / prim ids
 * of partly hit tiles) and stored (colour + depth [+ prim ids] of tiles with a nearer hit).  Synchronises with the scene's devices.
 * Any output pointer may be NULL. */
int rt_scene_transfer_bytes(const rt_scene *scene, uint64_t *fetched, uint64_t *stored);

/* ---- options (process-wide defaults; reads are racy by design, set them before rendering) ------- */
int rt_set_math_mode(int mode /*rt_math_mode*/);
int rt_set_sched_mode(int mode /*rt_sched_mode*/);

/* ---- rendering: replaces the body of Renderer::draw (raytracing.cpp:67-102) ---------------------- */
/* Device-resident framebuffer, asynchronous on `stream` (a cudaStream_t passed as void*; NULL = default
 * stream).  color/t/prim are device pointers to W*H elements (they may be peer-mapped memory of another
 * GPU: the kernel then stores its tiles straight into the root framebuffer over NVLink).  t_in may equal
 * t_out (the reference's in/out depth buffer) or be NULL (= cleared, +inf).  prim and counters may be NULL;
 * counters is a DEVICE pointer to an rt_counters that is accumulated into. */
int rt_render_device(const rt_scene *scene, const rt_frame *frame, uint32_t *color, const float *t_in, float *t_out,
                     uint32_t *prim, rt_counters *counters_dev, void *stream);
/* Host framebuffer, synchronous, exactly the reference's call: copies t in, renders, copies colour / t
 * (/ prim) out.  Returns the device time of the render kernels alone in *device_ms (the reference returns
 * its own stopwatch, raytracing.cpp:97-101) and of the whole call incl. copies in *total_ms.  Any of prim,
 * device_ms, total_ms, counters may be NULL.
 * Page-locked framebuffer: when color and t_inout (and prim, if given) are pinned host memory (rt_pinned_alloc /
 * rt_pinned_register below, or any cudaHostAlloc / cudaHostRegister memory) covering all W*H*4 bytes of each buffer and
 * 16-byte aligned (anything else silently takes the copies), W % 8 == 0 (band_rows % 4 == 0 for a banded call) and
 * counters is NULL, the kernel works on the caller's buffers directly over PCIe, for every scene kind and shading mode
 * (secondary rays and the ground plane included): with rt_frame.clear = 1 it stores its tiles as 128-byte bulk stores;
 * with clear = 0 (the reference's in/out framebuffer) a 32 x 4-pixel tile in which no ray hits anything is left alone
 * without being read; of the others the previous depth is fetched with bulk loads (it bounds the rays, raytracing.cpp:17),
 * of partly hit ones the previous colour as well, and the merged tile is stored back (rt_scene_transfer_bytes counts
 * both directions).  No device framebuffer, no copy engine; *device_ms then equals *total_ms.  Pageable
 * buffers and workload counters go through the pipelined copies.  On a scene replicated by rt_init every listed GPU
 * does this for its own bands over its own PCIe link. */
int rt_render_host(const rt_scene *scene, const rt_frame *frame, uint32_t *color, float *t_inout, uint32_t *prim,
                   float *device_ms, float *total_ms, rt_counters *counters);

/* ---- single rays: replaces IScene::intersect (raytracing.hpp:75-81) for n rays at once ------------ */
/* origins/dirs: n x 3 floats (host).  Outputs (host, n elements; normal n x 3; any may be NULL):
 * hit (0/1), t, the unflipped HitInfo::normal, prim id. */
int rt_intersect_host(const rt_scene *scene, const float *origins, const float *dirs, uint32_t n, float t_near,
                      float t_far, int has_plane, float plane_y, uint8_t *hit, float *t, float *normal,
                      uint32_t *prim);

/* The same call for callers that must not allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Identical
 * arguments and results, but the device staging of the rays and their results lives in one fixed file-scope device array
 * (2.8 MiB, held by every GPU that loads the library until the process exits) instead of six buffers allocated and freed
 * per call: the call makes no cudaMalloc / cudaFree and no allocating library call.  n > RT_INTERSECT_STATIC_MAX_RAYS is
 * rejected with RT_ERR_INVALID (split the batch, or use rt_intersect_host).  All calls on a GPU share the array, which keeps
 * the last call's contents (the result area is cleared at the start of each call); calls that overlap on one GPU would
 * race: the library serialises its own calls per process and returns only when the results are on the host. */
#define RT_INTERSECT_STATIC_MAX_RAYS 65536
int rt_intersect_host_static(const rt_scene *scene, const float *origins, const float *dirs, uint32_t n, float t_near,
                             float t_far, int has_plane, float plane_y, uint8_t *hit, float *t, float *normal,
                             uint32_t *prim);

/* ---- device buffers and cross-process sharing (multi-GPU: one process per GPU) ----------------------- */
/* The reference has one framebuffer in one address space.  With one process per GPU the root rank owns the
 * framebuffer (rt_device_malloc on its GPU), exports it (CUDA IPC), and every other rank opens it and passes the
 * mapped pointers to rt_render_device: its tiles are then stored straight into the root framebuffer over
 * NVLink / NVSwitch by the render kernel itself (no separate gather collective).
 * handle = 64 opaque bytes (cudaIpcMemHandle_t). */
#define RT_IPC_HANDLE_BYTES 64
int rt_device_malloc(uint64_t bytes, void **dev_ptr);            /* on the calling thread's current device */
int rt_device_free(void *dev_ptr);
int rt_ipc_export(const void *dev_ptr, unsigned char handle[RT_IPC_HANDLE_BYTES]);
int rt_ipc_open(const unsigned char handle[RT_IPC_HANDLE_BYTES], void **dev_ptr); /* enables peer access lazily */
/* number of render launches so far that stored their tiles into a peer framebuffer (a pointer from rt_ipc_open) as
 * shared-memory staged 128-byte TMA bulk writes (frames of >= 3 Mpixel split over >= 8 ranks, rt_frame.clear = 1, W % 8 == 0) */
uint64_t rt_staged_launch_count(void);
int rt_ipc_close(void *dev_ptr);
/* Page-locked host memory for the framebuffers of rt_render_host.  rt_pinned_register pins an existing range in place:
 * the reference's FrameBuffer keeps its Image2D<uint32_t> color / Image2D<float> t (src/raytracing.hpp:9-15, storage
 * Image2d.h:91) and registers their data() once after FrameBuffer::resize.  Two ranges registered separately must
 * not share a memory page (frame-sized vectors are mmap'd by the allocator and never do; tiny ones may: use
 * rt_pinned_alloc for those). */
int rt_pinned_alloc(uint64_t bytes, void **host_ptr);
int rt_pinned_free(void *host_ptr);
int rt_pinned_register(void *host_ptr, uint64_t bytes);
int rt_pinned_unregister(void *host_ptr);
/* stream-ordered copies between host and device buffers (stream = cudaStream_t as void*, NULL = default stream;
 * the call returns after the copy has completed) */
int rt_memcpy_d2h(void *host, const void *dev_ptr, uint64_t bytes, void *stream);
int rt_memcpy_h2d(void *dev_ptr, const void *host, uint64_t bytes, void *stream);

/* ---- environment knobs (read once when the library is loaded; measurement and A/B runs, not needed in production) ----
 *   RT_HOST_ZEROCOPY   0: rt_render_host always copies; 1: plain stores into pinned host memory; 2 (default): staged
 *                      128-byte bulk transfers; 3: as 2 but the in/out contract copies
 *   RT_STAGE_STORES    0: never stage stores into a peer framebuffer; 1: stage every launch; 2: stage for every peer
 *                      framebuffer; unset: stage at >= 8 GPUs on >= 3 Mpixel
 *   RT_MULTI           0: a scene replicated by rt_init still renders on one device per call
 *   RT_MULTI_BAND_ROWS rows per interleaved band of the one-call fan-out (default 8, multiple of 4)
 *   RT_HEAVY_CYCLES    heavy-first block order: SM cycles above which a block starts first in the next launch of the
 *                      same scene and frame geometry (default 35000; 0 = off); RT_HEAVY_MIN_RAYS / RT_HEAVY_MAX_RAYS:
 *                      launch sizes it applies to (default 600000 .. 3300000 rays).  Order only: output is unchanged
 *   RT_TILE_BLOCK      threads per block of the tiles kernel (32, 64, 128, 256); RT_HOST_CHUNKS: row chunks of the
 *                      copy pipeline; RT_OCT_REFILL / RT_OCT_BATCH: thresholds of RT_SCHED_COMPACT */

#ifdef __cplusplus
}
#endif
#endif /* RT_B200_H */
