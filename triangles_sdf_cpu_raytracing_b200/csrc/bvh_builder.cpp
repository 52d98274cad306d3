#include "bvh_builder.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>

namespace rt {
namespace {

constexpr float INF = std::numeric_limits<float>::infinity();
constexpr int LEAF_MAX = 4;    // triangles per leaf of the binary tree (a leaf word can address 16)
constexpr int SAH_BINS = 16;

struct Box {
  float lo[3] = {INF, INF, INF}, hi[3] = {-INF, -INF, -INF};
  void grow(const float *p) {
    for (int a = 0; a < 3; ++a) {
      lo[a] = std::min(lo[a], p[a]);
      hi[a] = std::max(hi[a], p[a]);
    }
  }
  void grow(const Box &b) {
    for (int a = 0; a < 3; ++a) {
      lo[a] = std::min(lo[a], b.lo[a]);
      hi[a] = std::max(hi[a], b.hi[a]);
    }
  }
  float area() const {
    float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    if (!(dx >= 0) || !(dy >= 0) || !(dz >= 0)) return 0.0f;
    return 2.0f * (dx * dy + dx * dz + dy * dz);
  }
};

struct Prim {
  Box box;
  float c[3];
  uint32_t tri;  // index into the input triangle list
};

struct BinNode {
  Box box;
  int left = -1, right = -1;  // inner
  uint32_t first = 0, count = 0;  // leaf when count > 0
};

struct Builder {
  std::vector<Prim> prims;
  std::vector<BinNode> bin;

  // depth: recursion level.  A SAH split can peel one primitive off a skewed range at a time; past MAX_SAH_DEPTH levels
  // the range is halved by count instead, so the recursion (and the traversal stack the tree needs) stays bounded by
  // MAX_SAH_DEPTH + log2(n) whatever the input looks like.
  static constexpr int MAX_SAH_DEPTH = 48;
  int build(uint32_t first, uint32_t count, int depth = 0) {
    int idx = (int)bin.size();
    bin.emplace_back();
    Box box, cbox;
    for (uint32_t i = first; i < first + count; ++i) {
      box.grow(prims[i].box);
      cbox.grow(prims[i].c);
    }
    bin[idx].box = box;
    if (count <= (uint32_t)LEAF_MAX) {
      bin[idx].first = first;
      bin[idx].count = count;
      return idx;
    }
    // binned SAH over the three axes
    int bestAxis = -1, bestBin = -1;
    float bestCost = INF;
    for (int a = 0; a < 3 && depth < MAX_SAH_DEPTH; ++a) {
      float ext = cbox.hi[a] - cbox.lo[a];
      if (!(ext > 0.0f)) continue;
      Box bb[SAH_BINS];
      uint32_t bc[SAH_BINS] = {};
      float k = (float)SAH_BINS / ext;
      for (uint32_t i = first; i < first + count; ++i) {
        int b = std::min(SAH_BINS - 1, std::max(0, (int)((prims[i].c[a] - cbox.lo[a]) * k)));
        bb[b].grow(prims[i].box);
        bc[b]++;
      }
      float rightArea[SAH_BINS];
      uint32_t rightCnt[SAH_BINS];
      Box acc;
      uint32_t n = 0;
      for (int b = SAH_BINS - 1; b > 0; --b) {
        acc.grow(bb[b]);
        n += bc[b];
        rightArea[b] = acc.area();
        rightCnt[b] = n;
      }
      acc = Box();
      n = 0;
      for (int b = 0; b < SAH_BINS - 1; ++b) {
        acc.grow(bb[b]);
        n += bc[b];
        if (n == 0 || rightCnt[b + 1] == 0) continue;
        float cost = acc.area() * (float)n + rightArea[b + 1] * (float)rightCnt[b + 1];
        if (cost < bestCost) {
          bestCost = cost;
          bestAxis = a;
          bestBin = b;
        }
      }
    }
    uint32_t mid;
    if (bestAxis >= 0) {
      float ext = cbox.hi[bestAxis] - cbox.lo[bestAxis];
      float k = (float)SAH_BINS / ext, lo = cbox.lo[bestAxis];
      int a = bestAxis, bsplit = bestBin;
      auto it = std::partition(prims.begin() + first, prims.begin() + first + count, [&](const Prim &p) {
        int b = std::min(SAH_BINS - 1, std::max(0, (int)((p.c[a] - lo) * k)));
        return b <= bsplit;
      });
      mid = (uint32_t)(it - prims.begin());
    } else {
      mid = first + count / 2;  // all centroids coincide: split by count
    }
    if (mid == first || mid == first + count) mid = first + count / 2;
    int l = build(first, mid - first, depth + 1);
    int r = build(mid, first + count - mid, depth + 1);
    bin[idx].left = l;
    bin[idx].right = r;
    return idx;
  }
};

void set_child_box(BvhNode &n, int lane, const Box &b) {
  for (int a = 0; a < 3; ++a) {
    n.lo[a][lane] = b.lo[a];
    n.hi[a][lane] = b.hi[a];
  }
}

void clear_node(BvhNode &n) {
  for (int a = 0; a < 3; ++a)
    for (int l = 0; l < 8; ++l) {
      n.lo[a][l] = INF;  // unused lanes: never consulted (child == BVH_EMPTY forces a miss)
      n.hi[a][l] = -INF;
    }
  for (int l = 0; l < 8; ++l) n.child[l] = BVH_EMPTY;
}

uint32_t leaf_word(uint32_t first, uint32_t count) { return BVH_LEAF_BIT | ((count - 1u) << 27) | first; }

// collapse the binary subtree rooted at b into an 8-wide node; returns its index
uint32_t collapse(const std::vector<BinNode> &bin, int b, std::vector<BvhNode> &out) {
  uint32_t idx = (uint32_t)out.size();
  out.emplace_back();
  clear_node(out[idx]);
  int kids[8];
  int n = 0;
  if (bin[b].count > 0) {
    kids[n++] = b;  // a leaf root: one child
  } else {
    kids[n++] = bin[b].left;
    kids[n++] = bin[b].right;
    while (n < 8) {
      int pick = -1;
      float best = -1.0f;
      for (int i = 0; i < n; ++i)
        if (bin[kids[i]].count == 0) {
          float a = bin[kids[i]].box.area();
          if (a > best) {
            best = a;
            pick = i;
          }
        }
      if (pick < 0) break;
      int k = kids[pick];
      kids[pick] = bin[k].left;
      kids[n++] = bin[k].right;
    }
  }
  for (int i = 0; i < n; ++i) {
    const BinNode &c = bin[kids[i]];
    set_child_box(out[idx], i, c.box);
    uint32_t w = c.count > 0 ? leaf_word(c.first, c.count) : collapse(bin, kids[i], out);
    out[idx].child[i] = w;  // out may have been reallocated by the recursion: index, not reference
  }
  return idx;
}

int stack_need(const std::vector<BvhNode> &nodes, uint32_t idx) {
  int k = 0, worst = 0;
  for (int l = 0; l < 8; ++l)
    if (nodes[idx].child[l] != BVH_EMPTY) ++k;
  for (int l = 0; l < 8; ++l) {
    uint32_t c = nodes[idx].child[l];
    if (c == BVH_EMPTY || (c & BVH_LEAF_BIT)) continue;
    worst = std::max(worst, k - 1 + stack_need(nodes, c));
  }
  return std::max(k, worst);
}

bool load_triangle(const float *vpos4, uint32_t n_verts, const uint32_t *idx3, float v[3][3], std::string &err) {
  for (int k = 0; k < 3; ++k) {
    uint32_t i = idx3[k];
    if (i >= n_verts) {
      err = "mesh index out of range";
      return false;
    }
    const float *p = vpos4 + 4 * (size_t)i;
    v[k][0] = p[0] / p[3];  // triangles_raytracing.cpp:307-309
    v[k][1] = p[1] / p[3];
    v[k][2] = p[2] / p[3];
  }
  return true;
}

TriDev make_tri(const float v[3][3], uint32_t prim) {
  TriDev t;
  float w;
  std::memcpy(&w, &prim, 4);
  t.v0 = make_float4(v[0][0], v[0][1], v[0][2], w);
  t.v1 = make_float4(v[1][0], v[1][1], v[1][2], 0.0f);
  t.v2 = make_float4(v[2][0], v[2][1], v[2][2], 0.0f);
  return t;
}

}  // namespace

bool build_bvh8(const float *vpos4, uint32_t n_verts, const uint32_t *indices, uint32_t n_indices, uint32_t brute,
                HostBvh &out, std::string &err) {
  if (!vpos4 || !indices || n_indices == 0 || n_indices % 3 != 0) {
    err = "mesh needs a non-empty index list with 3 indices per triangle";
    return false;
  }
  const uint32_t nTri = n_indices / 3;
  if (nTri >= (1u << 27)) {
    err = "too many triangles (limit 2^27)";
    return false;
  }
  std::vector<TriDev> all(nTri);
  Builder bld;
  bld.prims.resize(nTri);
  Box total;
  for (uint32_t t = 0; t < nTri; ++t) {
    float v[3][3];
    if (!load_triangle(vpos4, n_verts, indices + 3 * (size_t)t, v, err)) return false;
    all[t] = make_tri(v, t);
    Prim &p = bld.prims[t];
    for (int k = 0; k < 3; ++k) p.box.grow(v[k]);
    for (int a = 0; a < 3; ++a) p.c[a] = 0.5f * (p.box.lo[a] + p.box.hi[a]);
    p.tri = t;
    total.grow(p.box);
  }
  for (int a = 0; a < 3; ++a) {
    out.bbox_min[a] = total.lo[a];
    out.bbox_max[a] = total.hi[a];
  }
  out.nodes.clear();
  out.stack_need = 0;
  out.brute = brute;
  if (brute != BRUTE_NONE) {
    if (brute == BRUTE_ROOT_LEAF && nTri > ROOT_LEAF_MAX_TRIS) {
      err = "root-leaf meshes hold at most 8 triangles";
      return false;
    }
    out.tris = std::move(all);  // original order: ties keep the lowest prim id (the reference's leaf order, :324-331)
    // one node whose lane 0 is the mesh's bounding box: BRUTE_BOXED culls a ray against it the way the reference culls
    // against its child boxes (every one of them lies inside this box); BRUTE_ROOT_LEAF never looks at it
    BvhNode root{};
    for (int l = 0; l < 8; ++l) root.child[l] = BVH_EMPTY;
    for (int a = 0; a < 3; ++a) {
      root.lo[a][0] = total.lo[a];
      root.hi[a][0] = total.hi[a];
    }
    root.child[0] = BVH_LEAF_BIT;
    out.nodes.push_back(root);
    return true;
  }
  bld.bin.reserve(2 * (size_t)nTri / LEAF_MAX + 16);
  bld.build(0, nTri);
  out.tris.resize(nTri);
  for (uint32_t i = 0; i < nTri; ++i) out.tris[i] = all[bld.prims[i].tri];
  out.nodes.reserve(bld.bin.size() / 4 + 4);
  collapse(bld.bin, 0, out.nodes);
  out.stack_need = stack_need(out.nodes, 0);
  if (out.stack_need > BVH_STACK) {
    err = "BVH too deep for the traversal stack";
    return false;
  }
  return true;
}

bool import_reference_bvh(const float *vpos4, uint32_t n_verts, const uint32_t *indices_bvh, uint32_t n_indices,
                          const void *nodes, uint32_t n_nodes, uint32_t node_bytes, const uint32_t *prim_of_permuted,
                          HostBvh &out, std::string &err) {
  if (!vpos4 || !indices_bvh || !nodes || n_nodes == 0 || n_indices == 0 || n_indices % 3 != 0 || node_bytes < 201) {
    err = "bad reference BVH arguments";
    return false;
  }
  const uint32_t nTri = n_indices / 3;
  if (nTri >= (1u << 27)) {
    err = "too many triangles (limit 2^27)";
    return false;
  }
  out.tris.resize(nTri);
  Box total;
  std::vector<Box> triBox(nTri);
  for (uint32_t t = 0; t < nTri; ++t) {
    float v[3][3];
    if (!load_triangle(vpos4, n_verts, indices_bvh + 3 * (size_t)t, v, err)) return false;
    out.tris[t] = make_tri(v, prim_of_permuted ? prim_of_permuted[t] : t);
    for (int k = 0; k < 3; ++k) triBox[t].grow(v[k]);
    total.grow(triBox[t]);
  }
  for (int a = 0; a < 3; ++a) {
    out.bbox_min[a] = total.lo[a];
    out.bbox_max[a] = total.hi[a];
  }
  const uint8_t *raw = static_cast<const uint8_t *>(nodes);
  auto is_leaf = [&](uint32_t i) { return raw[(size_t)i * node_bytes + 200] != 0; };
  auto u32_at = [&](uint32_t i, size_t off) {
    uint32_t v;
    std::memcpy(&v, raw + (size_t)i * node_bytes + off, 4);
    return v;
  };
  auto leaf_ref = [&](uint32_t i, uint32_t &word) {
    uint32_t start = u32_at(i, 0), count = u32_at(i, 4);  // BVHLeafInfo {startIndex, count} in indices
    if (start % 3 || count % 3 || count == 0 || count > 24 || (uint64_t)start + count > n_indices) return false;
    word = leaf_word(start / 3, count / 3);
    return true;
  };
  // inner reference nodes get consecutive new indices in discovery order; leaves fold into their parent
  std::vector<int64_t> newIndex(n_nodes, -1);
  std::vector<uint32_t> order;
  out.nodes.clear();
  out.brute = BRUTE_NONE;
  if (is_leaf(0)) {
    // The reference's whole tree is one leaf (<= 8 triangles): traverseNode(0) tests its triangles with no box test
    // and no range check (triangles_raytracing.cpp:292-331).  That is the BRUTE_ROOT_LEAF scene, on the leaf's range.
    const uint32_t start = u32_at(0, 0), count = u32_at(0, 4);
    if (start % 3 || count % 3 || count == 0 || count > 24 || (uint64_t)start + count > n_indices) {
      err = "reference BVH: bad leaf record";
      return false;
    }
    std::vector<TriDev> leafTris(out.tris.begin() + start / 3, out.tris.begin() + (start + count) / 3);
    out.tris = std::move(leafTris);
    out.nodes.emplace_back();
    clear_node(out.nodes[0]);
    set_child_box(out.nodes[0], 0, total);
    out.nodes[0].child[0] = BVH_LEAF_BIT;
    out.stack_need = 0;
    out.brute = BRUTE_ROOT_LEAF;
    return true;
  }
  newIndex[0] = 0;
  order.push_back(0);
  for (size_t q = 0; q < order.size(); ++q) {
    uint32_t i = order[q];
    uint32_t realCount = u32_at(i, 192), offset = u32_at(i, 196);
    if (realCount == 0 || realCount > 8 || (uint64_t)offset + realCount > n_nodes) {
      err = "reference BVH: child range out of bounds";
      return false;
    }
    for (uint32_t c = 0; c < realCount; ++c) {
      uint32_t ci = offset + c;
      if (!is_leaf(ci)) {
        if (newIndex[ci] >= 0) {
          err = "reference BVH: node referenced twice";
          return false;
        }
        newIndex[ci] = (int64_t)order.size();
        order.push_back(ci);
      }
    }
  }
  out.nodes.resize(order.size());
  for (size_t q = 0; q < order.size(); ++q) {
    uint32_t i = order[q];
    BvhNode &n = out.nodes[q];
    clear_node(n);
    const float *b = reinterpret_cast<const float *>(raw + (size_t)i * node_bytes);
    float box[48];
    std::memcpy(box, b, 192);
    uint32_t realCount = u32_at(i, 192), offset = u32_at(i, 196);
    for (uint32_t c = 0; c < realCount; ++c) {
      for (int a = 0; a < 3; ++a) {
        n.lo[a][c] = box[8 * a + c];
        n.hi[a][c] = box[24 + 8 * a + c];
      }
      uint32_t ci = offset + c, w;
      if (is_leaf(ci)) {
        if (!leaf_ref(ci, w)) {
          err = "reference BVH: bad leaf record";
          return false;
        }
      } else {
        w = (uint32_t)newIndex[ci];
      }
      n.child[c] = w;
    }
  }
  out.stack_need = stack_need(out.nodes, 0);
  if (out.stack_need > BVH_STACK) {
    err = "BVH too deep for the traversal stack";
    return false;
  }
  return true;
}

}  // namespace rt
