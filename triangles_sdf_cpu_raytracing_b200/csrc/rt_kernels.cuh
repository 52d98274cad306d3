// Render kernels: per-pixel eye ray -> scene intersect -> shade -> pack -> depth-tested write, i.e. the body of
// Renderer::draw's pixel loop (raytracing.cpp:80-95) and Renderer::intersectionColor (:13-65), one thread per ray.
// This file is compiled twice (rt_kernels_exact.cu with -fmad=false, rt_kernels_fast.cu with FMA contraction),
// each time into its own namespace RT_NS.
#include <cstdint>
#include <cuda_runtime.h>

#include "rt_scene.h"

namespace rt {
namespace RT_NS {

#include "rt_device.cuh"

#include "rt_shade.cuh"

constexpr int BLOCK_THREADS = 256;
// a warp owns an 8x4 pixel tile; a 256-thread block owns 32x8 pixels (4 warps across, 2 down)
constexpr int TILE_W = 8, TILE_H = 4;

__device__ __forceinline__ void flush_counters(const Counters &c, unsigned long long *g) {
  // order = rt_counters: rays, hits, grid_samples, oct_inner, oct_leaf, oct_leaf_samples, bvh_inner, bvh_leaf, bvh_tris
  const uint32_t v[9] = {c.rays, c.hits, c.grid_samples, c.oct_inner, c.oct_leaf, c.oct_leaf_samples,
                         c.bvh_inner, c.bvh_leaf, c.bvh_tris};
#pragma unroll
  for (int i = 0; i < 9; ++i) {
    uint32_t x = v[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    if ((threadIdx.x & 31) == 0 && x) atomicAdd(g + i, (unsigned long long)x);
  }
}

// Stage the whole triangle list in shared memory (brute-force scenes; 48 B per triangle = 3 x float4).
// The list is one contiguous, 16-byte-aligned run of n_tris * 48 bytes, i.e. tile-regular: one elected thread
// issues a single TMA bulk copy (cp.async.bulk global -> shared, SASS UBLKCP) that completes on an mbarrier; every
// thread of the block then waits on the barrier's phase.  No thread spends load/store instructions on the copy.
__device__ __forceinline__ void stage_triangles(const MeshDev &m, float4 *smemTris) {
  __shared__ __align__(8) unsigned long long mbar;
  const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&mbar);
  const uint32_t dst = (uint32_t)__cvta_generic_to_shared(smemTris);
  const uint32_t bytes = m.n_tris * (uint32_t)sizeof(TriDev);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(1));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(m.tris), "r"(bytes), "r"(bar)
                 : "memory");
  }
  uint32_t done = 0;
  while (!done) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(0)
        : "memory");
  }
}

// one pixel: memory row `row` (0 = top), column x
template <int KIND, bool COUNT, bool SECONDARY>
__device__ __forceinline__ void shade_pixel(const RenderArgs &a, const float4 *smemTris, uint32_t x, uint32_t row,
                                            Counters &cnt) {
  const FrameDev &f = a.frame;
  const size_t px = (size_t)row * f.W + x;
  const int y = (int)(f.H - 1u - row);  // loop coordinate of raytracing.cpp:80-82
  const V3 o = mk3(f.ray_pos[0], f.ray_pos[1], f.ray_pos[2]);
  const V3 d = eye_ray(f, (int)x, y);
  const float tPrev = (a.t_in && !f.clear) ? a.t_in[px] : __int_as_float(0x7f800000);
  C4 col;
  float tNew;
  uint32_t prim;
  if (intersection_color<KIND, COUNT, SECONDARY>(a.scene, f, smemTris, o, d, tPrev, col, tNew, prim, cnt) &&
      !isinf(tNew)) {
    a.t_out[px] = tNew;
    a.color[px] = color_pack_rgba(col);
    if (a.prim) a.prim[px] = prim;
  } else if (f.clear) {  // fused FrameBuffer::clear (raytracing.hpp:16-19)
    a.t_out[px] = __int_as_float(0x7f800000);
    a.color[px] = 0u;
    if (a.prim) a.prim[px] = RT_PRIM_MISS;
  }
}

// virtual row v (0 .. rows_here) of this call -> memory row of the frame (interleaved bands, rt_frame)
__device__ __forceinline__ uint32_t band_row(const FrameDev &f, uint32_t v) {
  if (f.band_count <= 1u) return f.row0 + v;
  const uint32_t b = v / f.band_rows, r = v - b * f.band_rows;
  return (b * f.band_count + f.band_index) * f.band_rows + r;
}

// Rows of tiles are handed out from the middle of the call's rows outwards (0 -> mid, 1 -> mid-1, 2 -> mid+1, ...):
// the model sits mid-screen, so the long marches / deep traversals start first and the cheap box-miss tiles at the
// top and bottom of the frame fill the tail of the launch instead of one late heavy tile stretching it.
__device__ __forceinline__ uint32_t centre_out(uint32_t i, uint32_t n) {
  const uint32_t mid = n >> 1, k = i >> 1;
  return (i & 1u) ? mid - 1u - k : mid + k;
}

// ---- heavy-first block scheduling (HeavySched, rt_scene.h) ----------------------------------------------------
// HEAVY instantiations only (the others compile to exactly the plain kernels: three more live registers were enough to
// cost the 48-register grid marcher 16 %, so the state a block needs at its end lives in shared memory instead).
// The grid is (tilesX, extra_rows + tilesY): the leading extra_rows rows of blocks are slots for the tiles that ran long
// in the previous launch, the rest is the regular grid.  -> the tile position (bx, byi: default-order row, before
// centre_out) this block renders, or false when it has nothing to do.  The answer is the same for every thread of the
// block (one thread looks, the block votes), so a tile is never half rendered.
struct HeavyBlock {
  uint32_t tile;     // this block's tile id
  uint32_t t0[8];    // per warp: clock() when it started
};
__device__ __forceinline__ bool heavy_pick(const HeavySched &h, HeavyBlock &hb, uint32_t tilesX, uint32_t nTiles,
                                           uint32_t &bx, uint32_t &byi) {
  if ((threadIdx.x & 31u) == 0u) hb.t0[threadIdx.x >> 5] = (uint32_t)clock();
  if (blockIdx.y < h.extra_rows) {
    const uint32_t slot = blockIdx.y * tilesX + blockIdx.x;
    if (threadIdx.x == 0) {
      uint32_t tile = slot < min(__ldcg(h.prev_count), HEAVY_CAP) ? __ldcg(h.prev + slot) : 0xFFFFFFFFu;
      // the claim is one atomic exchange on both sides, so a tile is rendered exactly once per launch (an in/out tile
      // rendered twice could fetch its own first, half-landed result as "previous contents")
      if (tile < nTiles && atomicExch(h.done + tile, h.tag) == h.tag) tile = 0xFFFFFFFFu;  // its regular block was faster
      hb.tile = tile;
    }
    __syncthreads();
    const uint32_t tile = hb.tile;
    if (tile >= nTiles) return false;  // empty slot, a list built for another frame geometry, or already taken
    byi = tile / tilesX;
    bx = tile - byi * tilesX;
    return true;
  }
  byi = blockIdx.y - h.extra_rows;
  bx = blockIdx.x;
  const uint32_t tile = byi * tilesX + bx;
  int taken = 0;
  if (threadIdx.x == 0) {
    hb.tile = tile;
    taken = atomicExch(h.done + tile, h.tag) == h.tag;
  }
  return !__syncthreads_or(taken);
}
// called by every warp when its pixels are done
__device__ __forceinline__ void heavy_report(const HeavySched &h, const HeavyBlock &hb) {
  if ((threadIdx.x & 31u) == 0u && (uint32_t)clock() - hb.t0[threadIdx.x >> 5] > h.cycles) {
    const uint32_t tile = hb.tile;
    if (atomicExch(h.seen + tile, h.tag) != h.tag) {
      const uint32_t k = atomicAdd(h.next_count, 1u);
      if (k < HEAVY_CAP) h.next[k] = tile;
    }
  }
}

// ---- RT_SCHED_TILES: one thread per pixel -------------------------------------------------------------------
// A block is WX x WY warp tiles.  An SM releases a block's registers and warp slots only when its LAST warp retires,
// so with long-tailed per-ray costs (one grazing ray marching 100+ steps) big blocks strand resident-warp slots
// behind their slowest warp; the block shape per scene kind is chosen by measurement.
template <int KIND, bool COUNT, bool SECONDARY, int WX, int WY, bool HEAVY>
// (BVH only: capping the traversal at 64 registers buys 8 resident blocks and measured 5 % on both meshes; the same cap
// slows the grid and octree kernels, profiles/r1_tune.txt)
__global__ void __launch_bounds__(32 * WX * WY, KIND == KIND_BVH ? 1024 / (32 * WX * WY) : 0) render_tiles_kernel(const __grid_constant__ RenderArgs a) {
  extern __shared__ float4 smemTris[];
  __shared__ HeavyBlock hb;
  uint32_t bx = blockIdx.x, byi = blockIdx.y, tilesY = gridDim.y;
  if (HEAVY) {
    tilesY = gridDim.y - a.heavy.extra_rows;
    if (!heavy_pick(a.heavy, hb, gridDim.x, gridDim.x * tilesY, bx, byi)) return;
  }
  if (KIND == KIND_BRUTE) stage_triangles(a.scene.mesh, smemTris);
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  const uint32_t x = (bx * WX + warp % WX) * TILE_W + (lane & 7u);
  const uint32_t v = (centre_out(byi, tilesY) * WY + warp / WX) * TILE_H + (lane >> 3);
  Counters cnt = {};
  if (x < a.frame.W && v < a.frame.rows_here) {
    const uint32_t row = band_row(a.frame, v);
    // (small launches are decided by their longest march: the HEAVY flavour of the linear grid prefetches along it)
    constexpr int K = (HEAVY && KIND == KIND_GRID) ? KIND_GRID_PF : KIND;
    if (row < a.frame.H) shade_pixel<K, COUNT, SECONDARY>(a, smemTris, x, row, cnt);
  }
  if (COUNT) flush_counters(cnt, a.counters);
  if (HEAVY) {
    __syncwarp();
    heavy_report(a.heavy, hb);
  }
}

// one pixel of an in/out tile traced with its true previous depth (render_tiles_staged_kernel, INOUT)
template <int K, bool SECONDARY>
__device__ __noinline__ bool retrace_pixel(const SceneDev &scene, const FrameDev &f, const float4 *smemTris, V3 o, V3 d,
                                           float tPrev, uint32_t &color, float &t, uint32_t &prim) {
  Counters cnt = {};
  C4 col;
  float tNew = __int_as_float(0x7f800000);
  uint32_t pr;
  if (!intersection_color<K, false, SECONDARY>(scene, f, smemTris, o, d, tPrev, col, tNew, pr, cnt) || isinf(tNew))
    return false;
  color = color_pack_rgba(col);
  t = tNew;
  prim = pr;
  return true;
}

// ---- RT_SCHED_TILES with the tile staged in shared memory and stored by TMA (peer or pinned-host framebuffer) --------
// When the framebuffer lives on ANOTHER GPU (peer access / CUDA IPC) or in page-locked HOST memory (rt_render_host), the
// plain kernel's stores cross NVLink / PCIe as 32-byte pieces (one 8-pixel tile row per warp).  Here a block owns
// 32 x 4 pixels: every thread parks its pixel in shared memory, and after one barrier four threads issue the block's
// 128-byte row segments (colour, depth and, when asked for, prim ids) as bulk asynchronous stores (cp.async.bulk
// shared -> global, SASS UBLKCP): full 128-byte writes, no per-thread store instructions.  The block only waits until
// the bulk engine has READ its shared memory (wait_group.read), not until the writes have landed on the far side of
// the link: the grid's completion covers that.  W % 8 == 0.
//   INOUT = false: fused clear, every pixel of the tile is produced here.
//   INOUT = true : the reference's in/out framebuffer (raytracing.cpp:90-94).  The tile is traced against a cleared depth
//                  first; if nothing hits it is left alone unread.  Otherwise its previous DEPTH is fetched into shared
//                  memory by bulk loads completing on an mbarrier and the pixels it bounds are traced again with it; a
//                  tile without any nearer hit is then left alone, a fully hit one is stored whole, and only a partly
//                  hit one fetches its previous colour (and prim ids) so that the rows that go back are complete.
//                  (BVH: capped at 64 registers like the plain kernel; bunny 4K in/out call 0.434 -> 0.401 ms.)
template <int KIND, bool INOUT, bool SECONDARY, bool HEAVY>
__global__ void __launch_bounds__(128, KIND == KIND_BVH ? 8 : 0) render_tiles_staged_kernel(const __grid_constant__ RenderArgs a) {
  extern __shared__ float4 smemTris[];
  __shared__ __align__(128) uint32_t sColor[4][32];
  __shared__ __align__(128) float sDepth[4][32];
  __shared__ __align__(128) uint32_t sPrim[4][32];
  __shared__ __align__(8) unsigned long long mbar;
  const FrameDev &f = a.frame;
  __shared__ HeavyBlock hb;
  uint32_t bxi = blockIdx.x, byi = blockIdx.y, tilesY = gridDim.y;
  if (HEAVY) {
    tilesY = gridDim.y - a.heavy.extra_rows;
    if (!heavy_pick(a.heavy, hb, gridDim.x, gridDim.x * tilesY, bxi, byi)) return;
  }
  if (KIND == KIND_BRUTE) stage_triangles(a.scene.mesh, smemTris);
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  const uint32_t x0 = bxi * 32u, v0 = centre_out(byi, tilesY) * 4u;
  const uint32_t lx = warp * TILE_W + (lane & 7u), ly = lane >> 3;
  const uint32_t x = x0 + lx, v = v0 + ly;
  const uint32_t rowBytes = min(32u, f.W - x0) * 4u;  // W % 8 == 0: a multiple of 32 bytes, 16-byte aligned
  const bool wantPrim = a.prim != nullptr;
  const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&mbar);
  // thread 0: bulk-load the tile's rows of one framebuffer plane into shared memory, completing on the mbarrier
  auto load_rows = [&](const void *plane, uint32_t sdst) {
#pragma unroll
    for (uint32_t r = 0; r < 4u; ++r) {
      if (!(v0 + r < f.rows_here)) continue;
      const uint32_t row = band_row(f, v0 + r);
      if (!(row < f.H)) continue;
      const size_t px = (size_t)row * f.W + x0;
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                       sdst + r * 128u),
                   "l"(static_cast<const uint32_t *>(plane) + px), "r"(rowBytes), "r"(bar)
                   : "memory");
    }
  };
  auto wait_phase = [&](uint32_t parity) {
    uint32_t done = 0;
    while (!done) {
      asm volatile(
          "{\n\t.reg .pred p;\n\t"
          "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
          "selp.u32 %0, 1, 0, p;\n\t}"
          : "=r"(done)
          : "r"(bar), "r"(parity)
          : "memory");
    }
  };
  uint32_t rowsHere = 0;
#pragma unroll
  for (uint32_t r = 0; r < 4u; ++r)
    if (v0 + r < f.rows_here && band_row(f, v0 + r) < f.H) ++rowsHere;
  // this thread's pixel and eye ray
  bool inFrame = false;
  const V3 o = mk3(f.ray_pos[0], f.ray_pos[1], f.ray_pos[2]);
  V3 d = mk3(0.0f, 0.0f, 0.0f);
  if (x < f.W && v < f.rows_here) {
    const uint32_t row = band_row(f, v);
    if (row < f.H) {
      inFrame = true;
      d = eye_ray(f, (int)x, (int)(f.H - 1u - row));
    }
  }
  Counters cnt = {};
  bool hit = false;
  uint32_t myColor = 0u, myPrim = RT_PRIM_MISS;
  float myT = __int_as_float(0x7f800000);
  // INOUT: the previous depth only enters the reference's draw as the ray's far bound min(tFar, tPrev)
  // (raytracing.cpp:17), and every intersector only ever REJECTS more with a nearer far bound (the bound clips box
  // windows; no march, triangle test or visit order starts from it).  So the tile is first traced as if its previous
  // depth were +inf, what FrameBuffer::clear leaves there (src/main.cpp:197): if no ray hits then, none can with the
  // true depth, and the tile is left alone before anything has crossed the link.  Otherwise the previous depth is
  // fetched; pixels where it is nearer than tFar (a framebuffer that was not cleared) are traced again with the true
  // bound, exactly as the reference would have (retrace_pixel).
  constexpr int K = (HEAVY && KIND == KIND_GRID) ? KIND_GRID_PF : KIND;
  if (inFrame) {
    C4 col;
    float tNew = __int_as_float(0x7f800000);
    uint32_t prim;
    if (intersection_color<K, false, SECONDARY>(a.scene, f, smemTris, o, d, __int_as_float(0x7f800000), col, tNew, prim, cnt) &&
        !isinf(tNew)) {
      hit = true;
      myColor = color_pack_rgba(col);
      myT = tNew;
      myPrim = prim;
    }
  }
  if (INOUT) {
    if (!__syncthreads_or(hit)) {  // no ray hits anything: nothing to fetch, nothing to store
      if (HEAVY) heavy_report(a.heavy, hb);
      return;
    }
    if (threadIdx.x == 0) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(1));
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(rowsHere * rowBytes) : "memory");
      load_rows(a.t_in, (uint32_t)__cvta_generic_to_shared(&sDepth[0][0]));
      if (a.xfer) atomicAdd(a.xfer, (unsigned long long)(rowsHere * rowBytes));
    }
    wait_phase(0);
    // a previous depth nearer than tFar (a framebuffer that was not cleared): that pixel again, with the true bound
    // (out of line: the common path keeps the registers of a single trace)
    if (inFrame) {
      const float tPrev = sDepth[ly][lx];
      if (tPrev < f.t_far) hit = retrace_pixel<K, SECONDARY>(a.scene, f, smemTris, o, d, tPrev, myColor, myT, myPrim);
    }
  }
  if (INOUT) {
    const int nHit = __syncthreads_count(hit), nIn = __syncthreads_count(inFrame);
    if (nHit == 0) {  // nothing nearer anywhere in the tile: the framebuffer keeps what it holds
      if (HEAVY) heavy_report(a.heavy, hb);
      return;
    }
    if (nHit < nIn) {  // partly hit: the pixels that keep their value need it in the rows that go back
      if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar),
                     "r"(rowsHere * (wantPrim ? 2u : 1u) * rowBytes)
                     : "memory");
        load_rows(a.color, (uint32_t)__cvta_generic_to_shared(&sColor[0][0]));
        if (wantPrim) load_rows(a.prim, (uint32_t)__cvta_generic_to_shared(&sPrim[0][0]));
        if (a.xfer) atomicAdd(a.xfer, (unsigned long long)(rowsHere * (wantPrim ? 2u : 1u) * rowBytes));
      }
      wait_phase(1);
    }
    if (hit) {
      sColor[ly][lx] = myColor;
      sDepth[ly][lx] = myT;
      if (wantPrim) sPrim[ly][lx] = myPrim;
    }
  } else if (inFrame) {  // fused FrameBuffer::clear (raytracing.hpp:16-19): misses carry the clear values
    sColor[ly][lx] = myColor;
    sDepth[ly][lx] = myT;
    if (wantPrim) sPrim[ly][lx] = myPrim;
  }
  // every thread orders ITS generic-proxy writes of the tile before the async proxy (the bulk stores below read the
  // rows through it), then the barrier makes all of them visible to the four issuing threads
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 4u) {
    const uint32_t vr = v0 + threadIdx.x;
    const uint32_t row = vr < f.rows_here ? band_row(f, vr) : f.H;
    if (row < f.H && x0 < f.W) {
      const size_t px = (size_t)row * f.W + x0;
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(a.color + px),
                   "r"((uint32_t)__cvta_generic_to_shared(&sColor[threadIdx.x][0])), "r"(rowBytes)
                   : "memory");
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(a.t_out + px),
                   "r"((uint32_t)__cvta_generic_to_shared(&sDepth[threadIdx.x][0])), "r"(rowBytes)
                   : "memory");
      if (wantPrim)
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(a.prim + px),
                     "r"((uint32_t)__cvta_generic_to_shared(&sPrim[threadIdx.x][0])), "r"(rowBytes)
                     : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      if (INOUT && a.xfer) atomicAdd(a.xfer + 1, (unsigned long long)((wantPrim ? 3u : 2u) * rowBytes));
      // the shared-memory rows must stay valid until the bulk engine has read them; the writes themselves complete
      // before the grid does
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    }
  }
  if (HEAVY) heavy_report(a.heavy, hb);
}

// ---- RT_SCHED_PERSISTENT: persistent warps, dynamic tile fetch ------------------------------------------------
// Each warp repeatedly takes the next 8x4 tile from a global counter, so warps whose rays all terminate early
// (box misses cost a few dozen instructions) immediately pick up new work instead of idling until the slowest
// warp of their block finishes; the grid is sized to the machine (sm_count x resident blocks).
template <int KIND, bool COUNT, bool SECONDARY>
__global__ void __launch_bounds__(BLOCK_THREADS) render_persistent_kernel(const __grid_constant__ RenderArgs a) {
  extern __shared__ float4 smemTris[];
  if (KIND == KIND_BRUTE) stage_triangles(a.scene.mesh, smemTris);
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t tilesX = (a.frame.W + TILE_W - 1) / TILE_W, tilesY = (a.frame.rows_here + TILE_H - 1) / TILE_H;
  const uint32_t nTiles = tilesX * tilesY;
  Counters cnt = {};
  for (;;) {
    uint32_t tile = 0;
    if (lane == 0) tile = atomicAdd(a.tile_counter, 1u);
    tile = __shfl_sync(0xffffffffu, tile, 0);
    if (tile >= nTiles) break;
    const uint32_t tyi = tile / tilesX, tx = tile - tyi * tilesX, ty = centre_out(tyi, tilesY);
    const uint32_t x = tx * TILE_W + (lane & 7u), v = ty * TILE_H + (lane >> 3);
    if (x < a.frame.W && v < a.frame.rows_here) {
      const uint32_t row = band_row(a.frame, v);
      if (row < a.frame.H) shade_pixel<KIND, COUNT, SECONDARY>(a, smemTris, x, row, cnt);
    }
  }
  if (COUNT) flush_counters(cnt, a.counters);
}

// ---- octree, persistent warps with lane refill and batched leaf marches ------------------------------------------
// The north-star scheduling (RT_SCHED_COMPACT) for the most divergent scene type.  MEASURED AND NOT THE DEFAULT: the
// resumable state (OctWalk in local memory, three votes per iteration) costs more than the idle lanes it recovers:
// octree6 4K 0.73 ms persistent tiles vs 1.00 ms here at the best thresholds, depth-8 stand-in 1.16 vs 1.59 ms
// (profiles/r1_tune.txt).  Kept selectable and parity-tested.  A warp is a pool of 32 independent ray slots:
//   * a lane whose ray is finished goes IDLE; when enough lanes are idle (or nothing else is left to do) the idle lanes
//     take the next pixels from a global counter with ONE warp-aggregated atomicAdd (ballot + popc give each lane its
//     rank) and generate their rays together;
//   * lanes that stand on a surface leaf wait (state LEAF) until OCT_LEAF_BATCH of them can march together, or no
//     lane is left walking: the leaf march used to run at 5..13 of 32 lanes (profiles/r1_octree*_lines.txt);
//   * every iteration the warp votes (ballot) for one phase: refill, walk one child, or march.
// Per-ray arithmetic is oct_intersect's, so the frame is bit-identical; primary rays only (no plane / secondary rays).
constexpr int OCT_IDLE = 0, OCT_TRAV = 1, OCT_LEAF = 2;

// primary-ray part of Renderer::intersectionColor (raytracing.cpp:23-50) for a hit without secondary rays
__device__ __forceinline__ C4 shade_primary(const FrameDev &f, V3 o, V3 d, Hit &hit) {
  if (dot3(hit.normal, d) > 0) hit.normal = scale3(hit.normal, -1.0f);
  C4 color;
  if (f.mode == RT_SHADE_NORMAL) {
    color.r = (hit.normal.x + 1.0f) / 2.0f;
    color.g = (hit.normal.y + 1.0f) / 2.0f;
    color.b = (hit.normal.z + 1.0f) / 2.0f;
    color.a = (1.0f + 1.0f) / 2.0f;
  } else if (f.mode == RT_SHADE_COLOR) {
    color.r = color.g = color.b = hit.albedo;
    color.a = 1.0f;
  } else {
    color = lambert_color(hit, along(o, d, hit.t), mk3(f.light[0], f.light[1], f.light[2]), true);
  }
  return color;
}

template <bool COUNT>
__global__ void __launch_bounds__(128) render_octree_compact_kernel(const __grid_constant__ RenderArgs a) {
  const FrameDev &f = a.frame;
  const OctreeDev &oc = a.scene.oct;
  const uint32_t lane = threadIdx.x & 31u, ltMask = (1u << lane) - 1u;
  const uint32_t tilesX = (f.W + TILE_W - 1) / TILE_W, tilesY = (f.rows_here + TILE_H - 1) / TILE_H;
  const uint32_t nSlots = tilesX * tilesY * 32u;  // pixel slots in tile order (some fall outside the frame)
  const V3 o = mk3(f.ray_pos[0], f.ray_pos[1], f.ray_pos[2]);
  const float INF = __int_as_float(0x7f800000);
  Counters cnt = {};
  OctWalk s;
  int state = OCT_IDLE;
  size_t px = 0;
  bool exhausted = false;
  for (;;) {
    const uint32_t idle = __ballot_sync(0xffffffffu, state == OCT_IDLE);
    // ---- refill: idle lanes take the next pixel slots
    if (!exhausted && idle != 0u && (__popc(idle) >= a.oct_refill_min || idle == 0xffffffffu)) {
      const uint32_t n = (uint32_t)__popc(idle), leader = (uint32_t)__ffs(idle) - 1u;
      uint32_t first = 0;
      if (lane == leader) first = atomicAdd(a.tile_counter, n);
      first = __shfl_sync(0xffffffffu, first, leader);
      if (first + n >= nSlots) exhausted = true;
      if (state == OCT_IDLE) {
        const uint32_t slot = first + (uint32_t)__popc(idle & ltMask);
        if (slot < nSlots) {
          const uint32_t tile = slot >> 5, in = slot & 31u;
          const uint32_t tyi = tile / tilesX, tx = tile - tyi * tilesX, ty = centre_out(tyi, tilesY);
          const uint32_t x = tx * TILE_W + (in & 7u), v = ty * TILE_H + (in >> 3);
          const uint32_t row = (x < f.W && v < f.rows_here) ? band_row(f, v) : f.H;
          if (row < f.H) {
            px = (size_t)row * f.W + x;
            const V3 d = eye_ray(f, (int)x, (int)(f.H - 1u - row));
            const float tPrev = (a.t_in && !f.clear) ? a.t_in[px] : INF;
            if (COUNT) cnt.rays++;
            oct_walk_begin<COUNT>(s, oc, o, d, f.t_near, std_min(f.t_far, tPrev), cnt);
            state = OCT_TRAV;
          }
        }
      }
      continue;
    }
    const uint32_t trav = __ballot_sync(0xffffffffu, state == OCT_TRAV);
    const uint32_t leaf = __ballot_sync(0xffffffffu, state == OCT_LEAF);
    if ((trav | leaf) == 0u) {
      if (exhausted) break;
      continue;  // everyone idle: the refill branch fires next iteration
    }
    if (leaf != 0u && (__popc(leaf) >= a.oct_leaf_batch || trav == 0u)) {
      // ---- march phase
      if (state == OCT_LEAF) {
        Hit hit = hit_none();
        if (oct_walk_leaf<COUNT>(s, oc, o, f.t_near, hit, cnt)) {
          if (COUNT) cnt.hits++;
          const C4 col = shade_primary(f, o, s.d, hit);
          a.t_out[px] = hit.t;
          a.color[px] = color_pack_rgba(col);
          if (a.prim) a.prim[px] = hit.prim;
          state = OCT_IDLE;  // the first leaf that reports a hit ends the descent (octree_raytracing.cpp:188-197)
        } else {
          state = OCT_TRAV;
        }
      }
    } else {
      // ---- walk phase: one child visit per walking lane
      if (state == OCT_TRAV) {
        const int r = oct_walk_step<COUNT>(s, oc, o, f.t_near, cnt);
        if (r == OCT_WALK_LEAF) {
          state = OCT_LEAF;
        } else if (r == OCT_WALK_DONE) {
          if (f.clear) {  // fused FrameBuffer::clear (raytracing.hpp:16-19)
            a.t_out[px] = INF;
            a.color[px] = 0u;
            if (a.prim) a.prim[px] = RT_PRIM_MISS;
          }
          state = OCT_IDLE;
        }
      }
    }
  }
  if (COUNT) flush_counters(cnt, a.counters);
}

// ---- IScene::intersect for n arbitrary rays ---------------------------------------------------------------------
template <int KIND>
__global__ void __launch_bounds__(BLOCK_THREADS) rays_kernel(const __grid_constant__ RayArgs a) {
  extern __shared__ float4 smemTris[];
  if (KIND == KIND_BRUTE) stage_triangles(a.scene.mesh, smemTris);
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n) return;
  Counters cnt = {};
  const V3 o = mk3(a.origins[3 * i], a.origins[3 * i + 1], a.origins[3 * i + 2]);
  const V3 d = mk3(a.dirs[3 * i], a.dirs[3 * i + 1], a.dirs[3 * i + 2]);
  Hit h = scene_intersect<KIND, false>(a.scene, smemTris, a.has_plane != 0, a.plane_y, o, d, a.t_near, a.t_far, cnt);
  if (a.hit) a.hit[i] = h.hitten ? 1 : 0;
  if (a.t) a.t[i] = h.t;
  if (a.normal) {
    a.normal[3 * i] = h.normal.x;
    a.normal[3 * i + 1] = h.normal.y;
    a.normal[3 * i + 2] = h.normal.z;
  }
  if (a.prim) a.prim[i] = h.prim;
}

// ---- launchers ------------------------------------------------------------------------------------------------
static inline int kind_of(const SceneDev &s) {
  if (s.kind == RT_SCENE_GRID) return s.grid.bricks ? KIND_GRID_BRICKS : KIND_GRID;
  if (s.kind == RT_SCENE_OCTREE) return KIND_OCTREE;
  return s.mesh.brute ? KIND_BRUTE : KIND_BVH;
}

template <int KIND, bool COUNT, bool SECONDARY, int WX, int WY, bool HEAVY>
static cudaError_t launch_tiles_shape_h(const RenderArgs &a, size_t smem, cudaStream_t st) {
  auto kern = render_tiles_kernel<KIND, COUNT, SECONDARY, WX, WY, HEAVY>;
  if (smem > 40 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const uint32_t tx = (a.frame.W + TILE_W * WX - 1u) / (TILE_W * WX), ty = (a.frame.rows_here + TILE_H * WY - 1u) / (TILE_H * WY);
  RenderArgs b = a;
  b.heavy.extra_rows = HEAVY ? (HEAVY_CAP + tx - 1u) / tx : 0u;
  kern<<<dim3(tx, ty + b.heavy.extra_rows), 32 * WX * WY, smem, st>>>(b);
  return cudaGetLastError();
}
// heavy-first scheduling exists for the default block shape of the kernels that are timed (no counters)
template <int KIND, bool COUNT, bool SECONDARY, int WX, int WY>
static cudaError_t launch_tiles_shape(const RenderArgs &a, size_t smem, cudaStream_t st) {
  if (!COUNT && WX == 2 && WY == 2 && a.heavy.cycles) return launch_tiles_shape_h<KIND, false, SECONDARY, 2, 2, true>(a, smem, st);
  return launch_tiles_shape_h<KIND, COUNT, SECONDARY, WX, WY, false>(a, smem, st);
}

template <int KIND, bool INOUT, bool SECONDARY, bool HEAVY>
static cudaError_t launch_staged_h(const RenderArgs &a, size_t smem, cudaStream_t st) {
  auto kern = render_tiles_staged_kernel<KIND, INOUT, SECONDARY, HEAVY>;
  if (smem > 40 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const uint32_t tx = (a.frame.W + 31u) / 32u, ty = (a.frame.rows_here + 3u) / 4u;
  RenderArgs b = a;
  b.heavy.extra_rows = HEAVY ? (HEAVY_CAP + tx - 1u) / tx : 0u;
  kern<<<dim3(tx, ty + b.heavy.extra_rows), 128, smem, st>>>(b);
  return cudaGetLastError();
}

template <int KIND, bool COUNT, bool SECONDARY>
static cudaError_t launch_render_t(const RenderArgs &a, int sched, int sm_count, cudaStream_t st) {
  const size_t smem = (KIND == KIND_BRUTE) ? (size_t)a.scene.mesh.n_tris * sizeof(TriDev) : 0;
  if (KIND == KIND_OCTREE && !SECONDARY && sched == RT_SCHED_COMPACT && a.scene.oct.root_inner && !a.frame.has_plane) {
    auto kern = render_octree_compact_kernel<COUNT>;
    int perSm = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, kern, 128, 0);
    if (e != cudaSuccess) return e;
    if (perSm < 1) perSm = 1;
    e = cudaMemsetAsync(a.tile_counter, 0, sizeof(uint32_t), st);
    if (e != cudaSuccess) return e;
    kern<<<dim3((unsigned)(sm_count * perSm)), 128, 0, st>>>(a);
    return cudaGetLastError();
  }
  if (sched == RT_SCHED_PERSISTENT || sched == RT_SCHED_COMPACT) {
    auto kern = render_persistent_kernel<KIND, COUNT, SECONDARY>;
    if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    int perSm = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, kern, BLOCK_THREADS, smem);
    if (e != cudaSuccess) return e;
    if (perSm < 1) perSm = 1;
    e = cudaMemsetAsync(a.tile_counter, 0, sizeof(uint32_t), st);
    if (e != cudaSuccess) return e;
    kern<<<dim3((unsigned)(sm_count * perSm)), BLOCK_THREADS, smem, st>>>(a);
  } else {
    // Block shape.  Measured on B200 (profiles/r1_tune.txt): 128 threads (2 x 2 warp tiles) is the fastest shape for
    // the grid, octree and BVH kernels (256 strands warp slots behind a block's slowest warp, 32/64 lose L1 reuse
    // between neighbouring tiles); the brute-force kernel stages its triangles per block and keeps 256.
    // a.tile_block (RT_TILE_BLOCK) overrides it for sweeps; only the plain primary-ray kernels have every shape.
    const int tb = a.tile_block;
    if (KIND == KIND_BRUTE && !(!COUNT && a.stage_stores)) return launch_tiles_shape<KIND, COUNT, SECONDARY, 4, 2>(a, smem, st);
    if (!COUNT && a.stage_stores) {
      if (a.stage_stores == 2)
        return a.heavy.cycles ? launch_staged_h<KIND, true, SECONDARY, true>(a, smem, st)
                              : launch_staged_h<KIND, true, SECONDARY, false>(a, smem, st);
      return a.heavy.cycles ? launch_staged_h<KIND, false, SECONDARY, true>(a, smem, st)
                            : launch_staged_h<KIND, false, SECONDARY, false>(a, smem, st);
    }
    if (!COUNT && !SECONDARY) {
      if (tb == 32) return launch_tiles_shape<KIND, false, false, 1, 1>(a, smem, st);
      if (tb == 64) return launch_tiles_shape<KIND, false, false, 2, 1>(a, smem, st);
      if (tb == 256) return launch_tiles_shape<KIND, false, false, 4, 2>(a, smem, st);
    }
    return launch_tiles_shape<KIND, COUNT, SECONDARY, 2, 2>(a, smem, st);
  }
  return cudaGetLastError();
}

template <int KIND>
static cudaError_t launch_render_k(const RenderArgs &a, int sched, int sm_count, cudaStream_t st) {
  const bool secondary = a.frame.mode == RT_SHADE_LAMBERT && (a.frame.shadows || a.frame.reflections);
  const bool count = a.counters != nullptr;
  if (secondary) return count ? launch_render_t<KIND, true, true>(a, sched, sm_count, st)
                              : launch_render_t<KIND, false, true>(a, sched, sm_count, st);
  return count ? launch_render_t<KIND, true, false>(a, sched, sm_count, st)
               : launch_render_t<KIND, false, false>(a, sched, sm_count, st);
}

cudaError_t launch_render(const RenderArgs &a, int sched, int sm_count, cudaStream_t st) {
  if (a.frame.rows_here == 0 || a.frame.W == 0) return cudaSuccess;
  switch (kind_of(a.scene)) {
    case KIND_GRID: return launch_render_k<KIND_GRID>(a, sched, sm_count, st);
    case KIND_GRID_BRICKS: return launch_render_k<KIND_GRID_BRICKS>(a, sched, sm_count, st);
    case KIND_OCTREE: return launch_render_k<KIND_OCTREE>(a, sched, sm_count, st);
    case KIND_BVH: return launch_render_k<KIND_BVH>(a, sched, sm_count, st);
    default: return launch_render_k<KIND_BRUTE>(a, sched, sm_count, st);
  }
}

template <int KIND>
static cudaError_t launch_rays_k(const RayArgs &a, cudaStream_t st) {
  const size_t smem = (KIND == KIND_BRUTE) ? (size_t)a.scene.mesh.n_tris * sizeof(TriDev) : 0;
  auto kern = rays_kernel<KIND>;
  if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  kern<<<(a.n + BLOCK_THREADS - 1) / BLOCK_THREADS, BLOCK_THREADS, smem, st>>>(a);
  return cudaGetLastError();
}

cudaError_t launch_rays(const RayArgs &a, cudaStream_t st) {
  if (a.n == 0) return cudaSuccess;
  switch (kind_of(a.scene)) {
    case KIND_GRID: return launch_rays_k<KIND_GRID>(a, st);
    case KIND_GRID_BRICKS: return launch_rays_k<KIND_GRID_BRICKS>(a, st);
    case KIND_OCTREE: return launch_rays_k<KIND_OCTREE>(a, st);
    case KIND_BVH: return launch_rays_k<KIND_BVH>(a, st);
    default: return launch_rays_k<KIND_BRUTE>(a, st);
  }
}

}  // namespace RT_NS
}  // namespace rt
