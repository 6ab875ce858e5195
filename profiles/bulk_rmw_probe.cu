// Can bulk asynchronous copies (cp.async.bulk, the TMA's 1-D form) carry the update's row traffic with few warps?
// A stand-alone probe (not part of the library): 2^24 blocks of 256 bytes (the [w | h] pair layout at K=16, 4 GiB),
// 2.37 M of them -- one cfg2 batch's distinct rows, ascending like the sorted path visits them -- are read
// (160 bytes), changed and written back (160 bytes).
//
//   ldst      the shape of sorted_update_kernel: 5 lanes per row, LDG.128 x 2, STG.128 x 2, rows in registers
//   bulk      lane per row: cp.async.bulk global -> shared (160 B, mbarrier complete_tx), change in shared memory,
//             cp.async.bulk shared -> global (bulk_group); STAGES batches of 32 rows in flight per warp
// Both run at several warps-per-SM settings: the question is how few resident warps keep HBM busy, because the
// warps (and registers) the update does not need are what a parser running beside it gets.
//
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o profiles/bulk_rmw_probe.bin profiles/bulk_rmw_probe.cu
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

constexpr int kBlockFloats = 64;   // 256-byte stride
constexpr int kRowBytes = 160;     // [w | h] at K=16

__global__ void __launch_bounds__(256) ldst_kernel(float* __restrict__ store, const int* __restrict__ list, int n) {
  const int lane = threadIdx.x & 31, g = lane / 5, c = lane % 5;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
  for (int base = warp * 6; base < n; base += n_warps * 6) {
    const int j = base + g;
    if (g >= 6 || j >= n) continue;
    float* p = store + (int64_t)list[j] * kBlockFloats + 4 * c;
    float4 w = *reinterpret_cast<float4*>(p), h = *reinterpret_cast<float4*>(p + 20);
    w.x += 1.f; h.x += 1.f;
    *reinterpret_cast<float4*>(p) = w;
    *reinterpret_cast<float4*>(p + 20) = h;
  }
}

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
  unsigned ok = 0;
  while (!ok)
    asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(unsigned dst, const void* src, unsigned bytes, unsigned bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* dst, unsigned src, unsigned bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
}

// lane per row; a warp keeps STAGES batches of 32 rows in flight
template <int STAGES>
__global__ void bulk_kernel(float* __restrict__ store, const int* __restrict__ list, int n) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
  unsigned char* my = smem + (size_t)warp * STAGES * (32 * kRowBytes);
  unsigned long long* bars = reinterpret_cast<unsigned long long*>(smem + (size_t)warps * STAGES * (32 * kRowBytes)) + warp * STAGES;
  if (lane == 0)
    for (int s = 0; s < STAGES; ++s) mbar_init(smem_u32(bars + s), 1);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncwarp();
  const int gw = blockIdx.x * warps + warp, n_warps = gridDim.x * warps;
  const int n_batches = (n + 31) / 32;
  int my_batches = gw < n_batches ? (n_batches - gw + n_warps - 1) / n_warps : 0;

  auto issue = [&](int it) {            // loads of my it-th batch into stage it % STAGES
    const int s = it % STAGES;
    const int j = (gw + it * n_warps) * 32 + lane;
    const unsigned live = __ballot_sync(0xffffffffu, j < n);
    if (lane == 0) mbar_expect_tx(smem_u32(bars + s), kRowBytes * __popc(live));
    __syncwarp();
    if (j < n)
      bulk_g2s(smem_u32(my + ((size_t)s * 32 + lane) * kRowBytes), store + (int64_t)list[j] * kBlockFloats, kRowBytes,
               smem_u32(bars + s));
  };
  for (int it = 0; it < STAGES && it < my_batches; ++it) issue(it);
  for (int it = 0; it < my_batches; ++it) {
    const int s = it % STAGES;
    mbar_wait(smem_u32(bars + s), (it / STAGES) & 1);
    const int j = (gw + it * n_warps) * 32 + lane;
    if (j < n) {
      float* row = reinterpret_cast<float*>(my + ((size_t)s * 32 + lane) * kRowBytes);
      row[0] += 1.f;
      row[20] += 1.f;
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      bulk_s2g(store + (int64_t)list[j] * kBlockFloats, smem_u32(row), kRowBytes);
    }
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    if (it + STAGES < my_batches) {
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // the stores have read the stage
      __syncwarp();
      issue(it + STAGES);
    }
  }
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

template <class F>
float time_best(F launch, int reps) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  float best = 1e30f;
  for (int r = 0; r < reps; ++r) {
    CK(cudaEventRecord(e0));
    launch(r);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    if (r > 0 && ms < best) best = ms;     // rep 0 warms up
  }
  CK(cudaGetLastError());
  return best;
}

int main() {
  const int64_t n_blocks = 1 << 24;
  const int n = 2371504, reps = 5;
  float* store;
  CK(cudaMalloc(&store, n_blocks * kBlockFloats * sizeof(float)));
  CK(cudaMemset(store, 0, n_blocks * kBlockFloats * sizeof(float)));
  std::vector<int> host((size_t)n * reps);
  std::mt19937_64 rng(1);
  for (int r = 0; r < reps; ++r) {          // distinct rows, ascending, like a sorted batch
    std::vector<int> v(n);
    for (auto& x : v) x = (int)(rng() % n_blocks);
    std::sort(v.begin(), v.end());
    v.erase(std::unique(v.begin(), v.end()), v.end());
    while ((int)v.size() < n) v.push_back(v.back());   // the few duplicates lost: pad (same-row repeats are harmless here)
    std::copy(v.begin(), v.end(), host.begin() + (size_t)r * n);
  }
  int* list;
  CK(cudaMalloc(&list, host.size() * sizeof(int)));
  CK(cudaMemcpy(list, host.data(), host.size() * sizeof(int), cudaMemcpyHostToDevice));
  const double lines_mb = 2.0 * n * 256 / 1e6, bytes_mb = 2.0 * n * kRowBytes / 1e6;

  for (int per_sm : {1, 2, 3, 4, 6, 8}) {
    const float ms = time_best([&](int r) { ldst_kernel<<<148 * per_sm, 256>>>(store, list + (size_t)r * n, n); }, reps);
    printf("{\"kernel\": \"ldst\", \"warps_per_sm\": %d, \"ms\": %.4f, \"gbs_by_lines\": %.0f, \"gbs_by_bytes\": %.0f}\n",
           per_sm * 8, ms, lines_mb / ms, bytes_mb / ms);
  }
#define BULK(STAGES, WARPS)                                                                                          \
  for (int per_sm : {1, 2, 3, 4}) {                                                                                  \
    const size_t smem = (size_t)WARPS * STAGES * 32 * kRowBytes + WARPS * STAGES * 8;                                \
    if (smem * per_sm > 220 * 1024) continue;                                                                        \
    CK(cudaFuncSetAttribute(bulk_kernel<STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));           \
    const float ms = time_best([&](int r) {                                                                          \
      bulk_kernel<STAGES><<<148 * per_sm, WARPS * 32, smem>>>(store, list + (size_t)r * n, n); }, reps);             \
    printf("{\"kernel\": \"bulk\", \"stages\": %d, \"warps_per_sm\": %d, \"smem_per_sm_kb\": %d, \"ms\": %.4f, "     \
           "\"gbs_by_lines\": %.0f, \"gbs_by_bytes\": %.0f}\n", STAGES, per_sm * WARPS, (int)(smem * per_sm / 1024), \
           ms, lines_mb / ms, bytes_mb / ms);                                                                        \
  }
  BULK(2, 4) BULK(4, 4) BULK(8, 4) BULK(4, 2) BULK(8, 2) BULK(4, 8)
  // every touched row was incremented the same number of times by both kernels' runs: spot-check one row
  std::vector<float> row(kBlockFloats);
  CK(cudaMemcpy(row.data(), store + (int64_t)host[0] * kBlockFloats, sizeof(float) * kBlockFloats, cudaMemcpyDeviceToHost));
  printf("{\"check_row\": %d, \"w0\": %.0f, \"h0\": %.0f, \"w1\": %.0f}\n", host[0], row[0], row[20], row[1]);
  return 0;
}
