// fm_common.cuh -- shared host/device helpers for libfm_b200.so
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/fm_b200.h"

#define FM_EXPORT extern "C" __attribute__((visibility("default")))

// host: records a message for fm_last_error() (thread local) and returns `code`
int fm_fail(int code, const char* fmt, ...);

#define FM_REQUIRE(cond, ...)                               \
  do {                                                      \
    if (!(cond)) return fm_fail(FM_E_INVALID, __VA_ARGS__); \
  } while (0)

#define FM_CUDA(call)                                                                        \
  do {                                                                                       \
    cudaError_t e_ = (call);                                                                 \
    if (e_ != cudaSuccess)                                                                   \
      return fm_fail(FM_E_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, \
                     __LINE__);                                                              \
  } while (0)

#define FM_LAUNCH_CHECK() FM_CUDA(cudaGetLastError())

static inline size_t fm_align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// tuning knob read once from the environment (experiments only; the defaults are the measured optima)
int fm_env_int(const char* name, int fallback);

namespace fm {

constexpr int kWarp = 32;
constexpr unsigned kFull = 0xffffffffu;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
  return v;
}
__device__ __forceinline__ int warp_sum(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
  return v;
}

// sum of `n` floats by ONE block, fixed order, double accumulation: used as the
// deterministic second stage of every scalar reduction (reg_score, loss)
template <int kUnused = 0>
__global__ void reduce_partials_kernel(const float* __restrict__ partials, int n, float scale,
                                       float* __restrict__ out) {
  __shared__ double s_w[32];
  double acc = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) acc += (double)partials[i];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += s_w[w];
    *out = (float)(t * (double)scale);
  }
}

}  // namespace fm
