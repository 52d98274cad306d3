// On-device construction of the 8-wide BVH: replaces BVHBuilder::perform
// (triangles_raytracing.cpp:227-258: SAH sweep with OpenMP tasks and a parallel sort, 0.6 s for 69 k triangles on the
// reference's CPU path) with a linear BVH built entirely on the GPU:
//   1. k_setup      one thread per triangle: gather + divide by w exactly as triangles_raytracing.cpp:307-309, triangle
//                   box, scene box (block reduction + ordered-int atomics)
//   2. k_morton     63-bit Morton code of the box centre, cub radix sort of (code, triangle)
//   3. k_radix_tree binary radix tree over the sorted codes, every inner node in parallel (Karras 2012); ties between
//                   equal codes are broken by position so the tree is always a proper binary tree
//   4. k_fit        bottom-up boxes: the second thread to reach a node merges its children
//   5. k_collapse   level by level, one thread per wide node: pull the binary subtree up into <= 8 children, always
//                   opening the child with the largest surface area; subtrees of <= LEAF_TRIS triangles become leaves
//                   (their triangles are contiguous in Morton order)
//   6. k_reorder    triangles written in tree order as the 3 x float4 records the traversal kernel reads
// Closest-hit results do not depend on the tree, so the parity bar is unchanged: depth bit-identical.
#include <cub/device/device_radix_sort.cuh>

#include <algorithm>
#include <cfloat>
#include <cstring>
#include <string>
#include <vector>

#include "bvh_builder.h"

namespace rt {
namespace {

constexpr int LEAF_TRIS = 4;  // same leaf size as the host builder (a leaf word can address 16)

#define GB_CUDA(call)                                                        \
  do {                                                                       \
    cudaError_t e__ = (call);                                                \
    if (e__ != cudaSuccess) {                                                \
      err = std::string(#call) + ": " + cudaGetErrorString(e__);             \
      return false;                                                          \
    }                                                                        \
  } while (0)

// order-preserving float <-> int (for atomicMin / atomicMax on floats)
__device__ __forceinline__ int f2ord(float f) {
  int i = __float_as_int(f);
  return i >= 0 ? i : i ^ 0x7FFFFFFF;
}
__host__ __device__ __forceinline__ float ord2f(int i) {
  int j = i >= 0 ? i : i ^ 0x7FFFFFFF;
#ifdef __CUDA_ARCH__
  return __int_as_float(j);
#else
  float f;
  memcpy(&f, &j, 4);
  return f;
#endif
}

struct BuildState {
  int sceneLo[3], sceneHi[3];  // ordered-int encoded
  uint32_t error;              // 1: index out of range
  uint32_t nWide;              // wide nodes allocated so far
  uint32_t qCount[2];          // work items of the current / next collapse level
};

__global__ void k_setup(const float4 *vpos, uint32_t nVerts, const uint32_t *idx, uint32_t nTri, TriDev *tris,
                        float4 *boxLo, float4 *boxHi, BuildState *st) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  if (t < nTri) {
    const uint32_t i0 = idx[3 * (size_t)t], i1 = idx[3 * (size_t)t + 1], i2 = idx[3 * (size_t)t + 2];
    if (i0 >= nVerts || i1 >= nVerts || i2 >= nVerts) {
      st->error = 1u;
    } else {
      const float4 a = vpos[i0], b = vpos[i1], c = vpos[i2];
      // triangles_raytracing.cpp:307-309: vertex / w, IEEE division
      const float v[3][3] = {{a.x / a.w, a.y / a.w, a.z / a.w}, {b.x / b.w, b.y / b.w, b.z / b.w},
                             {c.x / c.w, c.y / c.w, c.z / c.w}};
      TriDev tr;
      tr.v0 = make_float4(v[0][0], v[0][1], v[0][2], __uint_as_float(t));
      tr.v1 = make_float4(v[1][0], v[1][1], v[1][2], 0.0f);
      tr.v2 = make_float4(v[2][0], v[2][1], v[2][2], 0.0f);
      tris[t] = tr;
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        lo[k] = fminf(v[0][k], fminf(v[1][k], v[2][k]));
        hi[k] = fmaxf(v[0][k], fmaxf(v[1][k], v[2][k]));
      }
      boxLo[t] = make_float4(lo[0], lo[1], lo[2], 0.0f);
      boxHi[t] = make_float4(hi[0], hi[1], hi[2], 0.0f);
    }
  }
  // scene box: warp shuffle reduction, one atomic pair per warp and axis
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    float l = lo[k], h = hi[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      l = fminf(l, __shfl_xor_sync(0xffffffffu, l, o));
      h = fmaxf(h, __shfl_xor_sync(0xffffffffu, h, o));
    }
    if ((threadIdx.x & 31) == 0 && l <= h) {
      atomicMin(&st->sceneLo[k], f2ord(l));
      atomicMax(&st->sceneHi[k], f2ord(h));
    }
  }
}

__device__ __forceinline__ unsigned long long spread21(uint32_t v) {  // 21 bits -> every third bit of 63
  unsigned long long x = v & 0x1FFFFFull;
  x = (x | x << 32) & 0x1F00000000FFFFull;
  x = (x | x << 16) & 0x1F0000FF0000FFull;
  x = (x | x << 8) & 0x100F00F00F00F00Full;
  x = (x | x << 4) & 0x10C30C30C30C30C3ull;
  x = (x | x << 2) & 0x1249249249249249ull;
  return x;
}

__global__ void k_morton(const float4 *boxLo, const float4 *boxHi, uint32_t nTri, const BuildState *st,
                         unsigned long long *keys, uint32_t *vals) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nTri) return;
  const float4 lo = boxLo[t], hi = boxHi[t];
  const float c[3] = {0.5f * (lo.x + hi.x), 0.5f * (lo.y + hi.y), 0.5f * (lo.z + hi.z)};
  uint32_t q[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const float s0 = ord2f(st->sceneLo[k]), s1 = ord2f(st->sceneHi[k]);
    const float ext = s1 - s0;
    float u = ext > 0.0f ? (c[k] - s0) / ext : 0.0f;
    u = fminf(fmaxf(u, 0.0f), 1.0f);
    q[k] = min(2097151u, (uint32_t)(u * 2097152.0f));
  }
  keys[t] = (spread21(q[0]) << 2) | (spread21(q[1]) << 1) | spread21(q[2]);
  vals[t] = t;
}

// length of the common prefix of the keys at sorted positions i and j; equal keys are told apart by position
__device__ __forceinline__ int delta(const unsigned long long *keys, int n, int i, int j) {
  if (j < 0 || j >= n) return -1;
  const unsigned long long a = keys[i], b = keys[j];
  if (a == b) return 64 + __clz((uint32_t)i ^ (uint32_t)j);
  return __clzll((long long)(a ^ b));
}

// node numbering: inner nodes 0 .. n-2, leaf k (sorted position k) = n-1 + k
__global__ void k_radix_tree(const unsigned long long *keys, int n, int2 *children, int2 *range, int *parent) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n - 1) return;
  const int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
  const int dMin = delta(keys, n, i, i - d);
  int lMax = 2;
  while (delta(keys, n, i, i + lMax * d) > dMin) lMax <<= 1;
  int l = 0;
  for (int t = lMax >> 1; t >= 1; t >>= 1)
    if (delta(keys, n, i, i + (l + t) * d) > dMin) l += t;
  const int j = i + l * d;
  const int dNode = delta(keys, n, i, j);
  int s = 0;
  for (int t = (l + 1) >> 1;; t = (t + 1) >> 1) {
    if (delta(keys, n, i, i + (s + t) * d) > dNode) s += t;
    if (t <= 1) break;
  }
  const int gamma = i + s * d + min(d, 0);
  const int first = min(i, j), last = max(i, j);
  const int left = (first == gamma) ? (n - 1 + gamma) : gamma;
  const int right = (last == gamma + 1) ? (n - 1 + gamma + 1) : gamma + 1;
  children[i] = make_int2(left, right);
  range[i] = make_int2(first, last);
  parent[left] = i;
  parent[right] = i;
  if (i == 0) parent[0] = -1;
}

__global__ void k_fit(const uint32_t *sortedTri, const float4 *triLo, const float4 *triHi, int n, const int2 *children,
                      const int *parent, float4 *nodeLo, float4 *nodeHi, uint32_t *visits) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const uint32_t t = sortedTri[k];
  float4 lo = triLo[t], hi = triHi[t];
  int node = n - 1 + k;
  nodeLo[node] = lo;
  nodeHi[node] = hi;
  __threadfence();
  int p = parent[node];
  while (p >= 0) {
    if (atomicAdd(&visits[p], 1u) == 0u) return;  // the sibling subtree is not finished: its thread will carry on
    __threadfence();
    const int2 c = children[p];
    // L2 loads: a neighbouring entry of the same line may sit stale in this SM's L1
    const float4 l0 = __ldcg(nodeLo + c.x), l1 = __ldcg(nodeLo + c.y), h0 = __ldcg(nodeHi + c.x), h1 = __ldcg(nodeHi + c.y);
    lo = make_float4(fminf(l0.x, l1.x), fminf(l0.y, l1.y), fminf(l0.z, l1.z), 0.0f);
    hi = make_float4(fmaxf(h0.x, h1.x), fmaxf(h0.y, h1.y), fmaxf(h0.z, h1.z), 0.0f);
    nodeLo[p] = lo;
    nodeHi[p] = hi;
    __threadfence();
    p = parent[p];
  }
}

__device__ __forceinline__ float box_area(float4 lo, float4 hi) {
  const float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
  return dx * dy + dx * dz + dy * dz;
}

// triangles under binary node b (sorted positions first .. first+count-1)
__device__ __forceinline__ void node_span(int b, int n, const int2 *range, int &first, int &count) {
  if (b >= n - 1) {
    first = b - (n - 1);
    count = 1;
  } else {
    const int2 r = range[b];
    first = r.x;
    count = r.y - r.x + 1;
  }
}

__global__ void k_collapse(const uint2 *qIn, uint32_t nIn, uint2 *qOut, uint32_t outSlot, int n, const int2 *children,
                           const int2 *range, const float4 *nodeLo, const float4 *nodeHi, BvhNode *wide,
                           uint32_t maxWide, BuildState *st) {
  const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= nIn) return;
  const uint2 item = qIn[w];  // x: binary node (inner, more than LEAF_TRIS triangles), y: wide node index
  int kids[8];
  int nk = 2;
  {
    const int2 c = children[item.x];
    kids[0] = c.x;
    kids[1] = c.y;
  }
  while (nk < 8) {
    int pick = -1;
    float best = -1.0f;
    for (int i = 0; i < nk; ++i) {
      int first, count;
      node_span(kids[i], n, range, first, count);
      if (count > LEAF_TRIS) {
        const float a = box_area(nodeLo[kids[i]], nodeHi[kids[i]]);
        if (a > best) {
          best = a;
          pick = i;
        }
      }
    }
    if (pick < 0) break;
    const int2 c = children[kids[pick]];
    kids[pick] = c.x;
    kids[nk++] = c.y;
  }
  BvhNode &out = wide[item.y];
  const float INF = __int_as_float(0x7f800000);
  for (int i = 0; i < 8; ++i) {
    if (i >= nk) {  // unused lanes: the empty box (+inf, -inf) misses every ray
      for (int a = 0; a < 3; ++a) {
        out.lo[a][i] = INF;
        out.hi[a][i] = -INF;
      }
      out.child[i] = BVH_EMPTY;
      continue;
    }
    const float4 lo = nodeLo[kids[i]], hi = nodeHi[kids[i]];
    out.lo[0][i] = lo.x; out.lo[1][i] = lo.y; out.lo[2][i] = lo.z;
    out.hi[0][i] = hi.x; out.hi[1][i] = hi.y; out.hi[2][i] = hi.z;
    int first, count;
    node_span(kids[i], n, range, first, count);
    if (count <= LEAF_TRIS) {
      out.child[i] = BVH_LEAF_BIT | ((uint32_t)(count - 1) << 27) | (uint32_t)first;
    } else {
      const uint32_t idx = atomicAdd(&st->nWide, 1u);
      if (idx >= maxWide) {
        st->error = 2u;
        out.child[i] = BVH_EMPTY;
        continue;
      }
      out.child[i] = idx;
      const uint32_t q = atomicAdd(&st->qCount[outSlot], 1u);
      qOut[q] = make_uint2((uint32_t)kids[i], idx);
    }
  }
}

__global__ void k_reorder(const TriDev *in, const uint32_t *sortedTri, uint32_t nTri, TriDev *out) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < nTri) out[k] = in[sortedTri[k]];
}

struct Scratch {  // frees everything it owns on scope exit
  std::vector<void *> p;
  ~Scratch() {
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

}  // namespace

bool build_bvh8_device(const float *vpos4, uint32_t n_verts, const uint32_t *indices, uint32_t n_indices, DeviceBvh &out,
                       std::string &err) {
  out = DeviceBvh();
  if (!vpos4 || !indices || n_indices == 0 || n_indices % 3 != 0) {
    err = "mesh needs a non-empty index list with 3 indices per triangle";
    return false;
  }
  const uint32_t nTri = n_indices / 3;
  if (nTri >= (1u << 27)) {
    err = "too many triangles (limit 2^27)";
    return false;
  }
  if (nTri <= (uint32_t)LEAF_TRIS) {
    err = "mesh too small for the device builder";
    return false;
  }
  const int n = (int)nTri;
  const uint32_t maxWide = nTri / 2 + 16;
  Scratch tmp;
  float4 *dV = nullptr, *triLo = nullptr, *triHi = nullptr, *nodeLo = nullptr, *nodeHi = nullptr;
  uint32_t *dI = nullptr, *vals = nullptr, *valsSorted = nullptr, *visits = nullptr;
  unsigned long long *keys = nullptr, *keysSorted = nullptr;
  TriDev *trisIn = nullptr;
  int2 *children = nullptr, *range = nullptr;
  int *parent = nullptr;
  uint2 *q[2] = {nullptr, nullptr};
  BuildState *st = nullptr;
  GB_CUDA(tmp.get(&dV, n_verts));
  GB_CUDA(tmp.get(&dI, n_indices));
  GB_CUDA(tmp.get(&trisIn, nTri));
  GB_CUDA(tmp.get(&triLo, nTri));
  GB_CUDA(tmp.get(&triHi, nTri));
  GB_CUDA(tmp.get(&keys, nTri));
  GB_CUDA(tmp.get(&keysSorted, nTri));
  GB_CUDA(tmp.get(&vals, nTri));
  GB_CUDA(tmp.get(&valsSorted, nTri));
  GB_CUDA(tmp.get(&children, nTri));
  GB_CUDA(tmp.get(&range, nTri));
  GB_CUDA(tmp.get(&parent, 2 * (size_t)nTri));
  GB_CUDA(tmp.get(&nodeLo, 2 * (size_t)nTri));
  GB_CUDA(tmp.get(&nodeHi, 2 * (size_t)nTri));
  GB_CUDA(tmp.get(&visits, nTri));
  GB_CUDA(tmp.get(&q[0], maxWide));
  GB_CUDA(tmp.get(&q[1], maxWide));
  GB_CUDA(tmp.get(&st, 1));
  BvhNode *wide = nullptr;
  TriDev *trisOut = nullptr;
  GB_CUDA(cudaMalloc(&wide, (size_t)maxWide * sizeof(BvhNode)));
  if (cudaMalloc(&trisOut, (size_t)nTri * sizeof(TriDev)) != cudaSuccess) {
    cudaFree(wide);
    err = "cudaMalloc of the triangle array failed";
    return false;
  }
  auto bail = [&](const std::string &why) {
    cudaFree(wide);
    cudaFree(trisOut);
    err = why;
    return false;
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
  cudaError_t ce = cudaMemcpy(dV, vpos4, (size_t)n_verts * 16, cudaMemcpyHostToDevice);
  if (ce == cudaSuccess) ce = cudaMemcpy(dI, indices, (size_t)n_indices * 4, cudaMemcpyHostToDevice);
  BuildState h;
  for (int k = 0; k < 3; ++k) {
    h.sceneLo[k] = 0x7FFFFFFF;
    h.sceneHi[k] = (int)0x80000000;
  }
  h.error = 0;
  h.nWide = 1;
  h.qCount[0] = 1;
  h.qCount[1] = 0;
  if (ce == cudaSuccess) ce = cudaMemcpy(st, &h, sizeof h, cudaMemcpyHostToDevice);
  if (ce == cudaSuccess) ce = cudaMemset(visits, 0, (size_t)nTri * 4);
  if (ce != cudaSuccess) return bail(std::string("upload: ") + cudaGetErrorString(ce));
  cudaEventRecord(e0);
  const int B = 256;
  const uint32_t gTri = (nTri + B - 1) / B;
  k_setup<<<gTri, B>>>(dV, n_verts, dI, nTri, trisIn, triLo, triHi, st);
  k_morton<<<gTri, B>>>(triLo, triHi, nTri, st, keys, vals);
  size_t sortBytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, sortBytes, keys, keysSorted, vals, valsSorted, n, 0, 63);
  void *sortTmp = nullptr;
  if (tmp.get((char **)&sortTmp, sortBytes) != cudaSuccess) return bail("cudaMalloc of the sort scratch failed");
  cub::DeviceRadixSort::SortPairs(sortTmp, sortBytes, keys, keysSorted, vals, valsSorted, n, 0, 63);
  k_radix_tree<<<gTri, B>>>(keysSorted, n, children, range, parent);
  k_fit<<<gTri, B>>>(valsSorted, triLo, triHi, n, children, parent, nodeLo, nodeHi, visits);
  k_reorder<<<gTri, B>>>(trisIn, valsSorted, nTri, trisOut);
  // collapse, level by level; the root work item is (binary root 0 -> wide node 0)
  const uint2 rootItem = make_uint2(0u, 0u);
  ce = cudaMemcpyAsync(q[0], &rootItem, sizeof rootItem, cudaMemcpyHostToDevice);
  uint32_t nIn = 1;
  int levels = 0, cur = 0;
  while (ce == cudaSuccess && nIn > 0) {
    ++levels;
    k_collapse<<<(nIn + 127) / 128, 128>>>(q[cur], nIn, q[cur ^ 1], (uint32_t)(cur ^ 1), n, children, range, nodeLo,
                                           nodeHi, wide, maxWide, st);
    uint32_t counts[2];
    ce = cudaMemcpy(counts, st->qCount, sizeof counts, cudaMemcpyDeviceToHost);  // also the level's sync point
    if (ce != cudaSuccess) break;
    nIn = counts[cur ^ 1];
    const uint32_t zero = 0;
    ce = cudaMemcpy(&st->qCount[cur], &zero, 4, cudaMemcpyHostToDevice);
    cur ^= 1;
    if (levels > 64) break;
  }
  cudaEventRecord(e1);
  if (ce == cudaSuccess) ce = cudaDeviceSynchronize();
  if (ce == cudaSuccess) ce = cudaGetLastError();
  if (ce != cudaSuccess) return bail(std::string("device BVH build: ") + cudaGetErrorString(ce));
  cudaMemcpy(&h, st, sizeof h, cudaMemcpyDeviceToHost);
  cudaEventElapsedTime(&out.build_ms, e0, e1);
  if (h.error == 1u) return bail("mesh index out of range");
  if (h.error == 2u) return bail("device BVH build: wide node pool exhausted");
  // the traversal pushes at most 8 children per visited level and pops one before descending
  out.stack_need = 7 * levels + 1;
  if (out.stack_need > BVH_STACK) return bail("BVH too deep for the traversal stack (degenerate Morton order)");
  out.nodes = wide;
  out.tris = trisOut;
  out.n_nodes = h.nWide;
  out.n_tris = nTri;
  out.levels = levels;
  for (int k = 0; k < 3; ++k) {
    out.bbox_min[k] = ord2f(h.sceneLo[k]);
    out.bbox_max[k] = ord2f(h.sceneHi[k]);
  }
  return true;
}

}  // namespace rt
