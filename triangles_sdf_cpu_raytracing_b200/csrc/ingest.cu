// On-device resource ingestion: the step before the hot path.  The raw file images the reference's
// loaders read (loadSDFGrid grid_raytracing.cpp:127-134, loadSDFOctree octree_raytracing.cpp:8-16) are uploaded as
// they are and turned into the traversal layouts by kernels:
//   octree  k_oct_split     36-byte {float values[8]; uint32 childrenOffset} records -> link[] + 2 x float4 values[],
//                           leaves classified from their values (isEmpty octree_raytracing.hpp:12-17, all >= 1e-4
//                           octree_raytracing.cpp:132-135); records are staged through shared memory so both the
//                           4-byte-aligned reads and the 16-byte writes are coalesced
//           k_oct_frontier  validation as a level-synchronous walk from the root: every inner node's 8 children in
//                           range, every INNER node reached at most once (no cycles; leaves may be shared), depth bounded.
//                           A bad childrenOffset would otherwise be an out-of-bounds device read in the render kernel.
//   grid    k_grid_bricks   linear x-major/z-fastest samples -> 8x8x8-sample bricks (rt_scene.h), one thread per
//                           output sample, coalesced writes
#include <algorithm>
#include <string>

#include "ingest.h"

namespace rt {
namespace {

constexpr int SPLIT_THREADS = 256;

__global__ void __launch_bounds__(SPLIT_THREADS) k_oct_split(const uint32_t *raw, uint32_t n, uint32_t *rawLink,
                                                             uint32_t *link, float4 *values) {
  __shared__ uint32_t rec[SPLIT_THREADS * 9];
  const uint32_t first = blockIdx.x * SPLIT_THREADS;
  const uint32_t here = min((uint32_t)SPLIT_THREADS, n - first);
  const uint32_t words = here * 9u;
  for (uint32_t w = threadIdx.x; w < words; w += SPLIT_THREADS) rec[w] = raw[(size_t)first * 9u + w];
  __syncthreads();
  if (threadIdx.x >= here) return;
  const uint32_t *r = rec + threadIdx.x * 9u;  // stride 9 words: conflict-free across a warp
  float v[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) v[k] = __uint_as_float(r[k]);
  const uint32_t off = r[8];
  const uint32_t i = first + threadIdx.x;
  values[2 * (size_t)i] = make_float4(v[0], v[1], v[2], v[3]);
  values[2 * (size_t)i + 1] = make_float4(v[4], v[5], v[6], v[7]);
  rawLink[i] = off;
  uint32_t l = off;
  if (off == 0u) {
    bool allBig = true, allZero = true, allPos = true;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (!(v[k] > 10.0f)) allBig = false;
      if (!(v[k] == 0.0f)) allZero = false;
      if (!(v[k] >= 1e-4f)) allPos = false;
    }
    l = (allBig || allZero || allPos) ? OCT_LEAF_CULLED : OCT_LEAF_SURFACE;
  }
  link[i] = l;
}

struct WalkState {
  uint32_t error;      // 1: childrenOffset out of range, 2: node referenced twice
  uint32_t count[2];   // frontier sizes (current / next)
};

__global__ void k_oct_frontier(const uint32_t *rawLink, uint32_t n, const uint32_t *frontIn, uint32_t nIn,
                               uint32_t *frontOut, uint32_t outSlot, uint32_t *seen, WalkState *st) {
  const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= nIn) return;
  const uint32_t off = rawLink[frontIn[w]];  // the frontier only holds inner nodes
  if ((unsigned long long)off + 7ull >= (unsigned long long)n) {
    st->error = 1u;
    return;
  }
  for (uint32_t c = 0; c < 8u; ++c) {
    const uint32_t child = off + c;
    // Leaves may be shared between parents (files with de-duplicated empty leaves load and render in the reference,
    // octree_raytracing.cpp:8-16,166-202): a leaf has no box of its own and ends the descent, so sharing one is harmless.
    // An INNER node reached twice is rejected: it is how a cycle would look, and the depth bound below is per node.
    if (rawLink[child] == 0u) continue;
    if (atomicExch(&seen[child], 1u) != 0u) {
      atomicMax(&st->error, 2u);
      continue;
    }
    frontOut[atomicAdd(&st->count[outSlot], 1u)] = child;
  }
}

// Inner nodes learn which of their children are leaves that intersectLeaf rejects from their values alone: the
// traversal then skips those without loading their records.  Runs after the walk above has validated every reachable
// offset; nodes the walk never reached may hold anything, so the range is checked here too.
__global__ void k_oct_childmask(const uint32_t *rawLink, const uint32_t *link, uint32_t n, uint32_t *out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t off = rawLink[i];
  uint32_t l = link[i];
  if (off != 0u && (unsigned long long)off + 7ull < (unsigned long long)n) {
    uint32_t mask = 0;
#pragma unroll
    for (uint32_t c = 0; c < 8u; ++c)
      if (link[off + c] != OCT_LEAF_CULLED) mask |= 1u << c;
    l = off | (mask << 24);
  }
  out[i] = l;
}

__global__ void k_grid_bricks(const float *lin, uint32_t sx, uint32_t sy, uint32_t sz, uint32_t nby, uint32_t nbz,
                              size_t total, float *out) {
  const size_t o = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (o >= total) return;
  const uint32_t in = (uint32_t)(o & 511u);
  const size_t br = o >> 9;
  const uint32_t bz = (uint32_t)(br % nbz), by = (uint32_t)((br / nbz) % nby), bx = (uint32_t)(br / ((size_t)nbz * nby));
  const uint32_t x = (bx << 3) | (in >> 6), y = (by << 3) | ((in >> 3) & 7u), z = (bz << 3) | (in & 7u);
  out[o] = (x < sx && y < sy && z < sz) ? lin[((size_t)x * sy + y) * sz + z] : 0.0f;
}

struct Tmp {
  std::vector<void *> p;
  ~Tmp() {
    for (void *q : p) cudaFree(q);
  }
  template <class T>
  cudaError_t get(T **out, size_t n) {
    void *d = nullptr;
    cudaError_t e = cudaMalloc(&d, std::max<size_t>(n * sizeof(T), 16));
    if (e == cudaSuccess) p.push_back(d);
    *out = static_cast<T *>(d);
    return e;
  }
};

#define IG_CUDA(call)                                            \
  do {                                                           \
    cudaError_t e__ = (call);                                    \
    if (e__ != cudaSuccess) {                                    \
      err = std::string(#call) + ": " + cudaGetErrorString(e__); \
      return INGEST_CUDA;                                        \
    }                                                            \
  } while (0)

}  // namespace

int ingest_octree_device(const void *nodes36, uint32_t n, uint32_t **linkOut, float4 **valuesOut, uint32_t *depthOut,
                         uint32_t *hasMaskOut, float *ms, std::string &err) {
  *hasMaskOut = 0;
  *linkOut = nullptr;
  *valuesOut = nullptr;
  Tmp tmp;
  uint32_t *raw = nullptr, *rawLink = nullptr, *seen = nullptr, *front[2] = {nullptr, nullptr};
  WalkState *st = nullptr;
  IG_CUDA(tmp.get(&raw, (size_t)n * 9));
  IG_CUDA(tmp.get(&rawLink, n));
  IG_CUDA(tmp.get(&seen, n));
  IG_CUDA(tmp.get(&front[0], n));
  IG_CUDA(tmp.get(&front[1], n));
  IG_CUDA(tmp.get(&st, 1));
  uint32_t *link = nullptr;
  float4 *values = nullptr;
  IG_CUDA(cudaMalloc(&link, std::max<size_t>((size_t)n * 4, 16)));
  if (cudaMalloc(&values, (size_t)n * 32) != cudaSuccess) {
    cudaFree(link);
    err = "cudaMalloc of the octree values failed";
    return INGEST_CUDA;
  }
  auto bail = [&](int code, const std::string &why) {
    cudaFree(link);
    cudaFree(values);
    err = why;
    return code;
  };
  struct Events {  // destroyed on every exit path
    cudaEvent_t a = nullptr, b = nullptr;
    Events() {
      cudaEventCreate(&a);
      cudaEventCreate(&b);
    }
    ~Events() {
      cudaEventDestroy(a);
      cudaEventDestroy(b);
    }
  } ev;
  cudaEvent_t e0 = ev.a, e1 = ev.b;
  cudaError_t ce = cudaMemcpy(raw, nodes36, (size_t)n * 36, cudaMemcpyHostToDevice);
  if (ce == cudaSuccess) ce = cudaMemset(seen, 0, (size_t)n * 4);
  WalkState h = {0u, {1u, 0u}};
  if (ce == cudaSuccess) ce = cudaMemcpy(st, &h, sizeof h, cudaMemcpyHostToDevice);
  const uint32_t rootSeen = 1u, rootNode = 0u;
  if (ce == cudaSuccess) ce = cudaMemcpy(seen, &rootSeen, 4, cudaMemcpyHostToDevice);
  if (ce == cudaSuccess) ce = cudaMemcpy(front[0], &rootNode, 4, cudaMemcpyHostToDevice);
  if (ce != cudaSuccess) return bail(INGEST_CUDA, std::string("octree upload: ") + cudaGetErrorString(ce));
  cudaEventRecord(e0);
  k_oct_split<<<(n + SPLIT_THREADS - 1) / SPLIT_THREADS, SPLIT_THREADS>>>(raw, n, rawLink, link, values);
  uint32_t rootLink = 0;
  ce = cudaMemcpy(&rootLink, rawLink, 4, cudaMemcpyDeviceToHost);
  uint32_t depth = 0, nIn = rootLink != 0u ? 1u : 0u;
  int cur = 0;
  while (ce == cudaSuccess && nIn > 0) {
    if (depth + 1 > (uint32_t)OCT_MAX_DEPTH) return bail(INGEST_INVALID, "octree deeper than the supported 21 levels");
    k_oct_frontier<<<(nIn + 127) / 128, 128>>>(rawLink, n, front[cur], nIn, front[cur ^ 1], (uint32_t)(cur ^ 1), seen, st);
    ce = cudaMemcpy(&h, st, sizeof h, cudaMemcpyDeviceToHost);  // also the level's sync point
    if (ce != cudaSuccess) break;
    if (h.error == 1u) return bail(INGEST_INVALID, "octree childrenOffset out of range");
    if (h.error == 2u) return bail(INGEST_INVALID, "octree inner node referenced twice (shared subtrees / cycles are not supported; shared leaves are)");
    ++depth;  // the children of this frontier exist: the tree reaches level `depth`
    nIn = h.count[cur ^ 1];
    const uint32_t zero = 0;
    ce = cudaMemcpy(&st->count[cur], &zero, 4, cudaMemcpyHostToDevice);
    cur ^= 1;
  }
  if (ce == cudaSuccess && n < OCT_MASK_MAX_NODES) {
    uint32_t *masked = nullptr;
    ce = cudaMalloc(&masked, std::max<size_t>((size_t)n * 4, 16));
    if (ce == cudaSuccess) {
      k_oct_childmask<<<(n + 255) / 256, 256>>>(rawLink, link, n, masked);
      cudaFree(link);  // stream-ordered: the kernel above has been enqueued before the free takes effect
      link = masked;
      *hasMaskOut = 1;
    }
  }
  cudaEventRecord(e1);
  if (ce == cudaSuccess) ce = cudaDeviceSynchronize();
  if (ce == cudaSuccess) ce = cudaGetLastError();
  if (ce != cudaSuccess) return bail(INGEST_CUDA, std::string("octree ingestion: ") + cudaGetErrorString(ce));
  if (ms) cudaEventElapsedTime(ms, e0, e1);
  *linkOut = link;
  *valuesOut = values;
  *depthOut = depth;
  return INGEST_OK;
}

int ingest_grid_bricks_device(const float *values, uint32_t sx, uint32_t sy, uint32_t sz, float **out, size_t *outCount,
                              std::string &err) {
  *out = nullptr;
  const uint32_t nbx = (sx + 7) / 8, nby = (sy + 7) / 8, nbz = (sz + 7) / 8;
  const size_t n = (size_t)sx * sy * sz, total = (size_t)nbx * nby * nbz * 512;
  Tmp tmp;
  float *lin = nullptr;
  IG_CUDA(tmp.get(&lin, n));
  IG_CUDA(cudaMemcpy(lin, values, n * 4, cudaMemcpyHostToDevice));
  float *b = nullptr;
  IG_CUDA(cudaMalloc(&b, total * 4));
  k_grid_bricks<<<(unsigned)((total + 255) / 256), 256>>>(lin, sx, sy, sz, nby, nbz, total, b);
  cudaError_t ce = cudaDeviceSynchronize();
  if (ce != cudaSuccess) {
    cudaFree(b);
    err = std::string("grid brick transform: ") + cudaGetErrorString(ce);
    return INGEST_CUDA;
  }
  *out = b;
  *outCount = total;
  return INGEST_OK;
}

}  // namespace rt
