// tools/peaks.cu — the roofs the ray kernels are measured against, MEASURED on the box:
//   * FP32 pipe: independent FFMA streams (2 flops / lane / clock) and independent FADD+FMUL streams (the exact,
//     no-FMA flavour's ceiling: 1 flop / lane / clock);
//   * warp-instruction issue rate (one instruction per SM sub-partition per clock);
//   * 4-byte random gathers served by L1 (per-block 64 KB window), by a 1 MB footprint (the 65^3 grid's size: L1 + L2),
//     by L2 (64 MB footprint) and by HBM (4 GB footprint, sector-granular), in useful GB/s (4 bytes per load) and in
//     giga-loads/s;
//   * dependent-load latency (pointer chase) in the same four regimes: the marchers are dependent chains.
// Prints ONE JSON object.  Built by the Makefile (nvcc, sm_100a); bench.py runs it once per box and caches the result.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include <cuda_runtime.h>

#define CK(call)                                                                              \
  do {                                                                                        \
    cudaError_t e__ = (call);                                                                 \
    if (e__ != cudaSuccess) {                                                                 \
      std::fprintf(stderr, "%s: %s\n", #call, cudaGetErrorString(e__));                       \
      std::exit(1);                                                                           \
    }                                                                                         \
  } while (0)

constexpr int ILP = 8;

// ---- FP32: 8 independent chains per thread -------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_ffma(float *out, int iters, float a, float b) {
  float x[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) x[i] = (float)(threadIdx.x + i);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i) x[i] = __fmaf_rn(x[i], a, b);
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += x[i];
  if (s == 12345.678f) out[0] = s;
}

__global__ void __launch_bounds__(256) k_fmul_fadd(float *out, int iters, float a, float b) {
  float x[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) x[i] = (float)(threadIdx.x + i);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i) x[i] = __fadd_rn(__fmul_rn(x[i], a), b);  // two instructions, never contracted
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += x[i];
  if (s == 12345.678f) out[0] = s;
}

// issue rate: integer ALU stream (one warp instruction per op)
__global__ void __launch_bounds__(256) k_issue(uint32_t *out, int iters, uint32_t a, uint32_t b) {
  uint32_t x[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) x[i] = threadIdx.x * 2654435761u + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i) x[i] = (x[i] ^ a) + b;  // LOP3 + IADD3: 2 instructions
  }
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s ^= x[i];
  if (s == 0x12345678u) out[0] = s;
}

// ---- random 4-byte gathers ------------------------------------------------------------------------------------------
// window = elements each BLOCK gathers from (perBlock != 0: its own window, L1-resident; else one shared footprint)
__global__ void __launch_bounds__(256) k_gather(const float *__restrict__ data, size_t n, size_t window, int perBlock,
                                                int iters, float *out) {
  const float *base = perBlock ? data + ((size_t)blockIdx.x * window) % (n - window + 1) : data;
  const uint32_t mask = (uint32_t)(window - 1);  // window is a power of two
  uint32_t s[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) s[i] = (blockIdx.x * 256u + threadIdx.x) * 747796405u + 2891336453u * (i + 1);
  float acc = 0;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i) {
      s[i] = s[i] * 1664525u + 1013904223u;
      acc += __ldg(base + ((s[i] >> 4) & mask));
    }
  }
  if (acc == 12345.678f) out[0] = acc;
}

// the same with 16-byte loads (octree values and BVH node runs are read as float4)
__global__ void __launch_bounds__(256) k_gather16(const float4 *__restrict__ data, size_t n, size_t window, int perBlock,
                                                  int iters, float *out) {
  const float4 *base = perBlock ? data + ((size_t)blockIdx.x * window) % (n - window + 1) : data;
  const uint32_t mask = (uint32_t)(window - 1);
  uint32_t s[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) s[i] = (blockIdx.x * 256u + threadIdx.x) * 747796405u + 2891336453u * (i + 1);
  float acc = 0;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i) {
      s[i] = s[i] * 1664525u + 1013904223u;
      const float4 v = __ldg(base + ((s[i] >> 4) & mask));
      acc += v.x + v.w;
    }
  }
  if (acc == 12345.678f) out[0] = acc;
}

// dependent chain: next = table[cur]
// (the walk resumes where the previous launch stopped, so a timed launch never revisits lines an earlier one cached)
__global__ void k_chase(const uint32_t *__restrict__ table, int iters, uint32_t *cursor) {
  uint32_t cur = cursor[0];
  for (int it = 0; it < iters; ++it) cur = __ldg(table + cur);
  cursor[0] = cur;
}

template <class F>
static float time_ms(F launch, int reps = 5) {
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a));
  CK(cudaEventCreate(&b));
  launch();  // warm-up
  CK(cudaDeviceSynchronize());
  float best = 1e30f;
  for (int r = 0; r < reps; ++r) {
    CK(cudaEventRecord(a));
    launch();
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    float ms;
    CK(cudaEventElapsedTime(&ms, a, b));
    if (ms < best) best = ms;
  }
  CK(cudaGetLastError());
  cudaEventDestroy(a);
  cudaEventDestroy(b);
  return best;
}

int main() {
  cudaDeviceProp p;
  CK(cudaGetDeviceProperties(&p, 0));
  const int sms = p.multiProcessorCount;
  float *out;
  CK(cudaMalloc(&out, 64));
  const int blocks = sms * 8, threads = 256;
  const double lanes = (double)blocks * threads;

  const int fi = 1 << 14;
  const float ms_ffma = time_ms([&] { k_ffma<<<blocks, threads>>>(out, fi, 1.0001f, 0.5f); });
  const float ms_nofma = time_ms([&] { k_fmul_fadd<<<blocks, threads>>>(out, fi, 1.0001f, 0.5f); });
  const float ms_issue = time_ms([&] { k_issue<<<blocks, threads>>>((uint32_t *)out, fi, 0x9e3779b9u, 0x7f4a7c15u); });
  const double ffma_tflops = lanes * fi * ILP * 2.0 / (ms_ffma * 1e-3) / 1e12;
  const double nofma_tflops = lanes * fi * ILP * 2.0 / (ms_nofma * 1e-3) / 1e12;
  const double issue_ginst = lanes / 32.0 * fi * ILP * 2.0 / (ms_issue * 1e-3) / 1e9;

  // gathers
  const size_t big = (size_t)1 << 30;  // 4 GB of floats
  float *data;
  CK(cudaMalloc(&data, big * 4));
  CK(cudaMemset(data, 0, big * 4));
  struct G {
    const char *name;
    size_t window;
    int perBlock;
    int iters;
  } gs[] = {{"l1_64KB_per_block", (size_t)1 << 14, 1, 2048},
            {"footprint_1MB", (size_t)1 << 18, 0, 2048},
            {"l2_64MB", (size_t)1 << 24, 0, 1024},
            {"hbm_4GB", (size_t)1 << 30, 0, 256}};
  double gload[4], gbs[4], gbs16[4];
  for (int k = 0; k < 4; ++k) {
    const G &g = gs[k];
    const float ms = time_ms([&] { k_gather<<<blocks, threads>>>(data, big, g.window, g.perBlock, g.iters, out); }, 3);
    gload[k] = lanes * g.iters * ILP / (ms * 1e-3) / 1e9;
    gbs[k] = gload[k] * 4.0;
    const float ms16 = time_ms(
        [&] { k_gather16<<<blocks, threads>>>((const float4 *)data, big / 4, g.window / 4, g.perBlock, g.iters, out); }, 3);
    gbs16[k] = lanes * g.iters * ILP / (ms16 * 1e-3) / 1e9 * 16.0;
  }
  // dependent chains: a random cyclic permutation over the footprint (Sattolo), one thread
  double chase_ns[4];
  const size_t chaseN[4] = {(size_t)1 << 13, (size_t)1 << 18, (size_t)1 << 24, (size_t)1 << 26};  // 32 KB, 1 MB, 64 MB, 256 MB
  uint32_t *tab;
  CK(cudaMalloc(&tab, chaseN[3] * 4));
  for (int k = 0; k < 4; ++k) {
    std::vector<uint32_t> h(chaseN[k]);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (uint32_t)i;
    uint64_t rng = 88172645463325252ull;
    for (size_t i = h.size() - 1; i > 0; --i) {
      rng ^= rng << 13; rng ^= rng >> 7; rng ^= rng << 17;
      size_t j = rng % i;
      uint32_t t = h[i]; h[i] = h[j]; h[j] = t;
    }
    CK(cudaMemcpy(tab, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
    const int it = k < 2 ? 200000 : 50000;
    CK(cudaMemset(out, 0, 4));
    // footprints that fit a cache are walked once in full first (warm); the HBM walk must stay cold
    if (k < 3) k_chase<<<1, 1>>>(tab, (int)h.size(), (uint32_t *)out);
    const float ms = time_ms([&] { k_chase<<<1, 1>>>(tab, it, (uint32_t *)out); }, 2);
    chase_ns[k] = ms * 1e6 / it;
  }
  int clk = 0;
  CK(cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0));
  std::printf(
      "{\"gpu\": \"%s\", \"sms\": %d, \"sm_clock_mhz_attr\": %.0f, \"l2_bytes\": %d, "
      "\"fp32_ffma_tflops\": %.3f, \"fp32_nofma_tflops\": %.3f, \"issue_ginst_s\": %.2f, "
      "\"gather_gloads_s\": {\"l1_64KB_per_block\": %.2f, \"footprint_1MB\": %.2f, \"l2_64MB\": %.2f, \"hbm_4GB\": %.2f}, "
      "\"gather_useful_gbs\": {\"l1_64KB_per_block\": %.1f, \"footprint_1MB\": %.1f, \"l2_64MB\": %.1f, \"hbm_4GB\": %.1f}, "
      "\"gather16_useful_gbs\": {\"l1_64KB_per_block\": %.1f, \"footprint_1MB\": %.1f, \"l2_64MB\": %.1f, \"hbm_4GB\": %.1f}, "
      "\"dependent_load_ns\": {\"l1_32KB\": %.1f, \"footprint_1MB\": %.1f, \"l2_64MB\": %.1f, \"hbm_256MB\": %.1f}, "
      "\"how\": \"tools/peaks.cu: %d blocks x %d threads, %d independent chains per thread; best of 3-5 launches, CUDA events\"}\n",
      p.name, sms, clk / 1e3, p.l2CacheSize, ffma_tflops, nofma_tflops, issue_ginst, gload[0], gload[1], gload[2], gload[3],
      gbs[0], gbs[1], gbs[2], gbs[3], gbs16[0], gbs16[1], gbs16[2], gbs16[3], chase_ns[0], chase_ns[1], chase_ns[2], chase_ns[3], blocks, threads, ILP);
  return 0;
}
