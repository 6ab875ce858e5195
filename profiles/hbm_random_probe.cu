// What can HBM deliver for the access shapes of the FM row kernels?  A stand-alone probe (not part of
// the library): 2^24 blocks of 256 bytes (the interleaved [w | h | g] layout at K=16, 4 GiB), touched
// through a list of 2.5 M uniformly random block numbers -- the shape of one cfg2 batch.
//
//   gather80    5 lanes read the first 80 bytes of a block (the scorer's read of w)
//   gather128   8 lanes read the first 128-byte line
//   gather256   16 lanes read the whole block
//   rmw256      16 lanes read the whole block and write it back (the Adagrad update's traffic)
//   red80       5 lanes add 80 bytes into the second line with red.global.add.v4.f32 (the scatter)
//   gather+red  gather80 of line 0 followed by red80 into line 1 of the same block (scorer + scatter)
//
// Each shape runs with 1, 2, 4 and 8 independent blocks in flight per lane group.  Bandwidth is
// reported against the 128-byte lines the shape touches (DRAM traffic is per line, see
// profiles/README.md r1a) and, for comparison, a plain streaming copy of the same 4 GiB.
//
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o /tmp/hbm_random_probe profiles/hbm_random_probe.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

constexpr int kBlockFloats = 64;   // 256 bytes

__device__ __forceinline__ void red4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// MODE 0 gather (LANES x 16 bytes from offset 0), 1 rmw (LANES x 16 bytes), 2 red (LANES x 16 bytes at
// +128), 3 gather from +0 then red at +128
template <int LANES, int UNROLL, int MODE>
__global__ void __launch_bounds__(256) probe(float* __restrict__ store, const int* __restrict__ list, int n,
                                             float* __restrict__ sink) {
  constexpr int GROUPS = 32 / LANES;               // lane groups per warp (trailing lanes idle)
  const int lane = threadIdx.x & 31, g = lane / LANES, c = lane % LANES;
  const bool on = g < GROUPS;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
  float acc = 0.f;
  for (int base = warp * GROUPS * UNROLL; base < n; base += n_warps * GROUPS * UNROLL) {
    float4 v[UNROLL];
    float* p[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int j = base + u * GROUPS + g;
      p[u] = (on && j < n) ? store + (int64_t)list[j] * kBlockFloats + 4 * c : nullptr;
    }
    if (MODE != 2) {
#pragma unroll
      for (int u = 0; u < UNROLL; ++u)
        if (p[u]) v[u] = MODE == 1 ? *reinterpret_cast<float4*>(p[u]) : __ldg(reinterpret_cast<const float4*>(p[u]));
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      if (!p[u]) continue;
      if (MODE == 0) acc += v[u].x + v[u].y + v[u].z + v[u].w;
      if (MODE == 1) { v[u].x += 1.f; *reinterpret_cast<float4*>(p[u]) = v[u]; }
      if (MODE == 2) red4(p[u] + 32, 1.f, 2.f, 3.f, 4.f);
      if (MODE == 3) red4(p[u] + 32, v[u].x, v[u].y, v[u].z, v[u].w);
    }
  }
  if (MODE == 0 && acc == 12345.678f) *sink = acc;
}

__global__ void copy_kernel(const float4* __restrict__ a, float4* __restrict__ b, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) b[i] = a[i];
}

template <int LANES, int UNROLL, int MODE>
void run(const char* name, float lines, float* store, const int* list, int n, float* sink) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  float best = 1e30f;
  int best_blocks = 0;
  for (int per_sm : {2, 4, 8}) {
    const int blocks = 148 * per_sm;
    for (int rep = 0; rep < 4; ++rep) {      // the list walks 2.5 M of 16.8 M blocks; 4 GiB store >> L2
      CK(cudaEventRecord(e0));
      probe<LANES, UNROLL, MODE><<<blocks, 256>>>(store, list + rep * n, n, sink);
      CK(cudaEventRecord(e1));
      CK(cudaEventSynchronize(e1));
      float ms;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      if (rep > 0 && ms < best) { best = ms; best_blocks = per_sm; }
    }
  }
  printf("{\"shape\": \"%s\", \"in_flight_per_group\": %d, \"ms\": %.4f, \"GBps_by_lines\": %.0f, \"blocks_per_sm\": %d}\n",
         name, UNROLL, best, lines * 128.0 * n / best / 1e6, best_blocks);
  CK(cudaGetLastError());
}

template <int LANES, int MODE>
void sweep(const char* name, float lines, float* store, const int* list, int n, float* sink) {
  run<LANES, 1, MODE>(name, lines, store, list, n, sink);
  run<LANES, 2, MODE>(name, lines, store, list, n, sink);
  run<LANES, 4, MODE>(name, lines, store, list, n, sink);
  run<LANES, 8, MODE>(name, lines, store, list, n, sink);
}

int main() {
  const int64_t rows = 1ll << 24;
  const int n = 2555904;           // 65 536 lines x 39 features
  float *store, *sink, *copy_dst;
  int* list;
  CK(cudaMalloc(&store, rows * kBlockFloats * 4));
  CK(cudaMalloc(&copy_dst, rows * kBlockFloats * 4));
  CK(cudaMalloc(&sink, 4));
  CK(cudaMemset(store, 0, rows * kBlockFloats * 4));
  std::vector<int> h(4 * (size_t)n);
  uint64_t s = 0x9E3779B97F4A7C15ull;
  for (auto& x : h) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; x = (int)(s >> 40); }   // 24 random bits
  CK(cudaMalloc(&list, h.size() * 4));
  CK(cudaMemcpy(list, h.data(), h.size() * 4, cudaMemcpyHostToDevice));

  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {
    CK(cudaEventRecord(e0));
    copy_kernel<<<148 * 16, 256>>>((const float4*)store, (float4*)copy_dst, rows * kBlockFloats / 4);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    if (ms < best) best = ms;
  }
  printf("{\"shape\": \"stream copy 4 GiB\", \"ms\": %.4f, \"GBps\": %.0f}\n", best, 2.0 * rows * 256 / best / 1e6);
  CK(cudaFree(copy_dst));

  sweep<5, 0>("gather80 (1 line read)", 1.f, store, list, n, sink);
  sweep<8, 0>("gather128 (1 line read)", 1.f, store, list, n, sink);
  sweep<16, 0>("gather256 (2 lines read)", 2.f, store, list, n, sink);
  sweep<16, 1>("rmw256 (2 lines read + 2 written)", 4.f, store, list, n, sink);
  sweep<5, 2>("red80 (1 line read + written)", 2.f, store, list, n, sink);
  sweep<5, 3>("gather80 + red80 (2 lines read, 1 written)", 3.f, store, list, n, sink);
  return 0;
}
