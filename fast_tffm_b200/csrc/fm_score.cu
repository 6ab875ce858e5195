// fm_score.cu -- FmScorer and FmGrad on the GPU.
//
// Replaces cc/fm_scorer_op.h:8-140 (BiasGenerator, FeatureRankGenerator,
// functor::FmScorer: >= 4 Eigen launches with [2,B,K] and [3,B] temporaries,
// one scalar thread per output coefficient, each re-walking the example's rows)
// and cc/fm_grad_op.h:23-163 (FactorSumGenerator, ParamGradGenerator,
// functor::FmGrad) with one kernel each.
//
// One warp per example.  A row (bias + K factors) is spread over LPR lanes
// (lanes-per-row = min(32, pow2 >= K)); the warp walks 32/LPR rows per step, so
// every step is one coalesced read of 32/LPR whole rows.  Per-lane accumulators
// hold sum(x v) and sum((x v)^2) for the lane's factor columns; the second-order
// term uses the identity 0.5 * sum_k[(sum_j x v)^2 - sum_j (x v)^2]
// (fm_scorer_op.h:134-136).  Row-group and column reductions are warp shuffles.
// reg_score is reduced warp -> block -> per-block partial -> one fixed-order
// pass (deterministic; no float atomics on the scalar).
#include <mutex>

#include "fm_common.cuh"

namespace fm {
namespace {

constexpr int kWarps = 8;  // warps (= examples) per block

template <int LPR>
__device__ __forceinline__ float reduce_row_groups(float v) {
  // sum over lanes that hold the same column (lane % LPR) of different rows
#pragma unroll
  for (int o = 16; o >= LPR; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
  return v;
}

// Walks the rows of example [start, end) and accumulates the per-column sums.
// T = factor columns per lane (column k = c + t * LPR).
template <int LPR, int T>
__device__ __forceinline__ void accumulate_example(const int32_t* __restrict__ fids,
                                                   const float* __restrict__ params, int64_t ld,
                                                   const float* __restrict__ vals, int start, int end,
                                                   int K, int lane, float (&S)[T], float (&Q)[T],
                                                   float& bias_dot, float& bias_sq, float& fac_sq) {
  constexpr int RPW = 32 / LPR;  // rows per warp step
  const int r = lane / LPR, c = lane % LPR;
#pragma unroll
  for (int t = 0; t < T; ++t) { S[t] = 0.f; Q[t] = 0.f; }
  bias_dot = bias_sq = fac_sq = 0.f;
  for (int base = start; base < end; base += 32) {
    // one coalesced load of up to 32 (id, value) pairs, then broadcast by shuffle
    const int j = base + lane;
    const int my_fid = j < end ? fids[j] : 0;
    const float my_x = j < end ? vals[j] : 0.f;
    const int rows_here = min(32, end - base);
#pragma unroll 4
    for (int s = 0; s < LPR; ++s) {
      const int jj = s * RPW + r;
      if (s * RPW >= rows_here) break;  // warp-uniform
      const int fid = __shfl_sync(kFull, my_fid, jj);
      const float x = __shfl_sync(kFull, my_x, jj);
      if (jj < rows_here) {
        const float* row = params + (int64_t)fid * ld;
        if (c == 0) {
          const float b = __ldg(row);
          bias_dot = fmaf(x, b, bias_dot);
          bias_sq = fmaf(b, b, bias_sq);
        }
#pragma unroll
        for (int t = 0; t < T; ++t) {
          const int k = c + t * LPR;
          if (k < K) {
            const float v = __ldg(row + 1 + k);
            const float xv = x * v;
            S[t] += xv;
            Q[t] = fmaf(xv, xv, Q[t]);
            fac_sq = fmaf(v, v, fac_sq);
          }
        }
      }
    }
  }
}

template <int LPR, int T>
__global__ void __launch_bounds__(kWarps * 32)
score_kernel(const int32_t* __restrict__ fids, const float* __restrict__ params, int64_t ld,
             const float* __restrict__ vals, const int32_t* __restrict__ poses, int64_t batch, int K,
             float lam_f, float lam_b, float* __restrict__ pred, float* __restrict__ reg_partials) {
  __shared__ float s_reg[kWarps];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t i = (int64_t)blockIdx.x * kWarps + warp;
  float reg = 0.f;
  if (i < batch) {
    // the second-order term is a sum over factor columns, so any K is walked in slabs of LPR * T columns
    // (one slab for every K <= 256; the rows are re-walked per slab beyond that)
    float S[T], Q[T], bias_dot = 0.f, bias_sq = 0.f, fac_sq = 0.f, second = 0.f;
    for (int k0 = 0; k0 < K || k0 == 0; k0 += LPR * T) {
      float bd, bs, fs;
      accumulate_example<LPR, T>(fids, params + k0, ld, vals, poses[i], poses[i + 1], min(K - k0, LPR * T), lane, S, Q,
                                 bd, bs, fs);
      if (k0 == 0) { bias_dot = bd; bias_sq = bs; }   // column k0 of the row is the bias only in the first slab
      fac_sq += fs;
#pragma unroll
      for (int t = 0; t < T; ++t) {
        const float s = reduce_row_groups<LPR>(S[t]);
        const float q = reduce_row_groups<LPR>(Q[t]);
        second += s * s - q;
      }
    }
    if (lane >= LPR) second = 0.f;  // every row group now holds the same totals: count them once
    const float p = warp_sum(bias_dot + 0.5f * second);
    reg = warp_sum(0.5f * lam_f * fac_sq + 0.5f * lam_b * bias_sq);
    if (lane == 0) pred[i] = p;
  }
  if (lane == 0) s_reg[warp] = reg;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) t += s_reg[w];
    reg_partials[blockIdx.x] = t;
  }
}

// FmGrad: pass 1 = factor sums s_k (fm_grad_op.h:37-49), pass 2 = the
// per-occurrence terms (fm_grad_op.h:80-106) scattered with float atomics
// (the reference does the same with atomicAdd, fm_grad_op.h:90-102).
template <int LPR, int T>
__global__ void __launch_bounds__(kWarps * 32)
grad_kernel(const int32_t* __restrict__ fids, const float* __restrict__ params, int64_t ld,
            const float* __restrict__ vals, const int32_t* __restrict__ poses, int64_t batch, int K,
            float lam_f, float lam_b, const float* __restrict__ pred_grad, float reg_grad_host,
            const float* __restrict__ reg_grad_dev, float* __restrict__ grad) {
  constexpr int RPW = 32 / LPR;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int r = lane / LPR, c = lane % LPR;
  const int64_t i = (int64_t)blockIdx.x * kWarps + warp;
  if (i >= batch) return;
  const int start = poses[i], end = poses[i + 1];
  const float pg = pred_grad[i];
  const float reg_grad = reg_grad_dev ? __ldg(reg_grad_dev) : reg_grad_host;
  const float rf = reg_grad * lam_f, rb = reg_grad * lam_b;
  // column k's gradient needs only column k's factor sum: any K is walked in slabs of LPR * T columns
  for (int k0 = 0; k0 < K || k0 == 0; k0 += LPR * T) {
    const int kk = min(K - k0, LPR * T);
    float S[T], Q[T], bias_dot, bias_sq, fac_sq;
    accumulate_example<LPR, T>(fids, params + k0, ld, vals, start, end, kk, lane, S, Q, bias_dot, bias_sq, fac_sq);
#pragma unroll
    for (int t = 0; t < T; ++t) S[t] = reduce_row_groups<LPR>(S[t]);
    for (int base = start; base < end; base += 32) {
      const int j = base + lane;
      const int my_fid = j < end ? fids[j] : 0;
      const float my_x = j < end ? vals[j] : 0.f;
      const int rows_here = min(32, end - base);
#pragma unroll 4
      for (int s = 0; s < LPR; ++s) {
        const int jj = s * RPW + r;
        if (s * RPW >= rows_here) break;
        const int fid = __shfl_sync(kFull, my_fid, jj);
        const float x = __shfl_sync(kFull, my_x, jj);
        if (jj < rows_here) {
          const float* row = params + (int64_t)fid * ld + k0;
          float* grow = grad + (int64_t)fid * ld + k0;
          if (c == 0 && k0 == 0) atomicAdd(grow, fmaf(pg, x, rb * __ldg(row)));          // :88-89
#pragma unroll
          for (int t = 0; t < T; ++t) {
            const int k = c + t * LPR;
            if (k < kk) {
              const float v = __ldg(row + 1 + k);
              // reg_grad*lf*t - pg*x*x*t + pg*x*s_k                           // :96-97
              atomicAdd(grow + 1 + k, rf * v - pg * x * x * v + pg * x * S[t]);
            }
          }
        }
      }
    }
  }
}

// ---- float4 flavour (rows 16-byte aligned, ld a multiple of 4) --------------------------------
// gs lanes share a row (gs = the float4 chunks of a K+1 wide row, at most 32; lane c of a group owns
// chunks c, c + gs, ... T of them), a warp walks 32 / gs rows per step.  Columns past K (row padding)
// are read but masked out.
__device__ __forceinline__ float4 row_chunk(const float* row, int chunk, int K) {
  float4 v = __ldg(reinterpret_cast<const float4*>(row + 4 * chunk));
  const int col = 4 * chunk;
  if (col + 1 > K) v.y = 0.f;
  if (col + 2 > K) v.z = 0.f;
  if (col + 3 > K) v.w = 0.f;
  return v;
}

template <int T>
__device__ __forceinline__ void gather_example_vec(const int32_t* __restrict__ fids, const float* __restrict__ params,
                                                   int64_t ld, const float* __restrict__ vals, int start, int end,
                                                   int K, int lane, int gs, float4 (&S)[T], float& q,
                                                   float& bias_dot, float& bias_sq, float& fac_sq) {
  const int RPW = 32 / gs, r = lane / gs, c = lane - r * gs;
  const bool lane_on = r < RPW;
  const int vc = (K + 4) >> 2;
#pragma unroll
  for (int t = 0; t < T; ++t) S[t] = make_float4(0.f, 0.f, 0.f, 0.f);
  q = bias_dot = bias_sq = fac_sq = 0.f;
  for (int base = start; base < end; base += 32) {
    const int j = base + lane;
    const int my_fid = j < end ? fids[j] : 0;
    const float my_x = j < end ? vals[j] : 0.f;
    const int rows_here = min(32, end - base);
#pragma unroll 4
    for (int s = 0; s * RPW < rows_here; ++s) {  // warp-uniform
      const int jj = s * RPW + r;
      const int fid = __shfl_sync(kFull, my_fid, jj & 31);
      const float x = __shfl_sync(kFull, my_x, jj & 31);
      if (lane_on && jj < rows_here) {
        const float* row = params + (int64_t)fid * ld;
#pragma unroll
        for (int t = 0; t < T; ++t) {
          const int chunk = c + t * gs;
          if (chunk < vc) {
            float4 v = row_chunk(row, chunk, K);
            if (chunk == 0) {  // column 0 is the bias
              bias_dot = fmaf(x, v.x, bias_dot);
              bias_sq = fmaf(v.x, v.x, bias_sq);
              v.x = 0.f;
            }
            const float a0 = x * v.x, a1 = x * v.y, a2 = x * v.z, a3 = x * v.w;
            S[t].x += a0; S[t].y += a1; S[t].z += a2; S[t].w += a3;
            q = fmaf(a0, a0, fmaf(a1, a1, fmaf(a2, a2, fmaf(a3, a3, q))));
            fac_sq = fmaf(v.x, v.x, fmaf(v.y, v.y, fmaf(v.z, v.z, fmaf(v.w, v.w, fac_sq))));
          }
        }
      }
    }
  }
  // column totals over the row groups (lanes c, c + gs, ...): every lane ends up with its columns' totals
#pragma unroll
  for (int t = 0; t < T; ++t) {
    float4 tot = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int g2 = 0; g2 < RPW; ++g2) {
      const int src = c + g2 * gs;
      tot.x += __shfl_sync(kFull, S[t].x, src);
      tot.y += __shfl_sync(kFull, S[t].y, src);
      tot.z += __shfl_sync(kFull, S[t].z, src);
      tot.w += __shfl_sync(kFull, S[t].w, src);
    }
    S[t] = tot;
  }
}

template <int T>
__global__ void __launch_bounds__(kWarps * 32)
score_vec_kernel(const int32_t* __restrict__ fids, const float* __restrict__ params, int64_t ld,
                 const float* __restrict__ vals, const int32_t* __restrict__ poses, int64_t batch, int K,
                 float lam_f, float lam_b, float* __restrict__ pred, float* __restrict__ reg_partials, int gs) {
  __shared__ float s_reg[kWarps];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t i = (int64_t)blockIdx.x * kWarps + warp;
  float reg = 0.f;
  if (i < batch) {
    float4 S[T];
    float q, bias_dot, bias_sq, fac_sq;
    gather_example_vec<T>(fids, params, ld, vals, poses[i], poses[i + 1], K, lane, gs, S, q, bias_dot, bias_sq, fac_sq);
    float second = 0.f;
#pragma unroll
    for (int t = 0; t < T; ++t) second += S[t].x * S[t].x + S[t].y * S[t].y + S[t].z * S[t].z + S[t].w * S[t].w;
    if (lane >= gs) second = 0.f;  // every row group holds the same totals: count them once
    const float p = warp_sum(bias_dot + 0.5f * (second - q));
    reg = warp_sum(0.5f * lam_f * fac_sq + 0.5f * lam_b * bias_sq);
    if (lane == 0) pred[i] = p;
  }
  if (lane == 0) s_reg[warp] = reg;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) t += s_reg[w];
    reg_partials[blockIdx.x] = t;
  }
}

__device__ __forceinline__ void red_add_v4f(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// FmGrad on float4 chunks: the factor sums of pass A, then one 16-byte vector atomic per chunk and occurrence
template <int T>
__global__ void __launch_bounds__(kWarps * 32)
grad_vec_kernel(const int32_t* __restrict__ fids, const float* __restrict__ params, int64_t ld,
                const float* __restrict__ vals, const int32_t* __restrict__ poses, int64_t batch, int K,
                float lam_f, float lam_b, const float* __restrict__ pred_grad, float reg_grad_host,
                const float* __restrict__ reg_grad_dev, float* __restrict__ grad, int gs) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t i = (int64_t)blockIdx.x * kWarps + warp;
  if (i >= batch) return;
  const int RPW = 32 / gs, r = lane / gs, c = lane - r * gs;
  const bool lane_on = r < RPW;
  const int vc = (K + 4) >> 2;
  const int start = poses[i], end = poses[i + 1];
  float4 S[T];
  float q, bias_dot, bias_sq, fac_sq;
  gather_example_vec<T>(fids, params, ld, vals, start, end, K, lane, gs, S, q, bias_dot, bias_sq, fac_sq);
  const float pg = pred_grad[i];
  const float reg_grad = reg_grad_dev ? __ldg(reg_grad_dev) : reg_grad_host;
  const float rf = reg_grad * lam_f, rb = reg_grad * lam_b;
  for (int base = start; base < end; base += 32) {
    const int j = base + lane;
    const int my_fid = j < end ? fids[j] : 0;
    const float my_x = j < end ? vals[j] : 0.f;
    const int rows_here = min(32, end - base);
#pragma unroll 4
    for (int s = 0; s * RPW < rows_here; ++s) {
      const int jj = s * RPW + r;
      const int fid = __shfl_sync(kFull, my_fid, jj & 31);
      const float x = __shfl_sync(kFull, my_x, jj & 31);
      if (lane_on && jj < rows_here) {
        const float* row = params + (int64_t)fid * ld;
        float* grow = grad + (int64_t)fid * ld;
#pragma unroll
        for (int t = 0; t < T; ++t) {
          const int chunk = c + t * gs;
          if (chunk < vc) {
            const float4 v = row_chunk(row, chunk, K);
            const float rv[4] = {v.x, v.y, v.z, v.w};
            const float sv[4] = {S[t].x, S[t].y, S[t].z, S[t].w};
            float o[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int col = 4 * chunk + e;
              if (col == 0) o[e] = fmaf(pg, x, rb * rv[e]);                              // fm_grad_op.h:88-89
              else if (col <= K) o[e] = rf * rv[e] - pg * x * x * rv[e] + pg * x * sv[e];  // :96-97
              else o[e] = 0.f;
            }
            red_add_v4f(grow + 4 * chunk, o[0], o[1], o[2], o[3]);
          }
        }
      }
    }
  }
}

// zero-fill of the first *rows_dev rows when the row count lives on the device
__global__ void __launch_bounds__(256)
zero_rows_kernel(float4* __restrict__ p, int64_t max_vec, int64_t ld, const int64_t* __restrict__ rows_dev) {
  const int64_t n = min((*rows_dev * ld + 3) / 4, max_vec);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    p[i] = make_float4(0.f, 0.f, 0.f, 0.f);
}

template <template <int, int> class Launcher, class... Args>
int dispatch(int64_t K, Args... args) {
  // lanes per row = min(32, pow2 >= K); T = ceil(K / LPR) rounded up to a compiled size
  if (K <= 4) return Launcher<4, 1>::run(args...);
  if (K <= 8) return Launcher<8, 1>::run(args...);
  if (K <= 16) return Launcher<16, 1>::run(args...);
  if (K <= 32) return Launcher<32, 1>::run(args...);
  if (K <= 64) return Launcher<32, 2>::run(args...);
  if (K <= 96) return Launcher<32, 3>::run(args...);
  if (K <= 128) return Launcher<32, 4>::run(args...);
  if (K <= 192) return Launcher<32, 6>::run(args...);
  return Launcher<32, 8>::run(args...);   // K > 256: the kernels walk the columns in slabs of 256
}

template <int LPR, int T>
struct ScoreLaunch {
  static int run(const int32_t* fids, const float* params, int64_t ld, const float* vals, const int32_t* poses,
                 int64_t batch, int K, float lf, float lb, float* pred, float* partials, unsigned blocks,
                 cudaStream_t stream) {
    score_kernel<LPR, T><<<blocks, kWarps * 32, 0, stream>>>(fids, params, ld, vals, poses, batch, K, lf, lb,
                                                             pred, partials);
    return FM_OK;
  }
};
template <int LPR, int T>
struct GradLaunch {
  static int run(const int32_t* fids, const float* params, int64_t ld, const float* vals, const int32_t* poses,
                 int64_t batch, int K, float lf, float lb, const float* pg, float rg, const float* rg_dev,
                 float* grad, unsigned blocks, cudaStream_t stream) {
    grad_kernel<LPR, T><<<blocks, kWarps * 32, 0, stream>>>(fids, params, ld, vals, poses, batch, K, lf, lb,
                                                            pg, rg, rg_dev, grad);
    return FM_OK;
  }
};

// can the rows be walked as float4 chunks?  (ld a multiple of 4 covering the padded row, 16-byte aligned
// bases, at most 3 chunks per lane)
static inline bool vec_shape(int64_t K, int64_t ld, const void* a, const void* b, int* gs, int* t) {
  const int64_t chunks = (K + 4) / 4;
  if (ld % 4 != 0 || ld < 4 * chunks || (((uintptr_t)a | (uintptr_t)b) & 15) != 0) return false;
  *gs = chunks <= 32 ? (int)chunks : 32;
  *t = (int)((chunks + *gs - 1) / *gs);
  return *t <= 3;
}

// the scorer launch shared by fm_score and fm_score_noalloc: float4 chunks when the rows allow it
// (vec_shape), else the scalar kernels; `partials` = one float per block
static int launch_score(const int32_t* feature_ids, const float* params, int64_t params_ld,
                        const float* feature_vals, const int32_t* feature_poses, int64_t batch,
                        int64_t factor_num, float factor_lambda, float bias_lambda, float* pred_score,
                        float* partials, unsigned blocks, cudaStream_t stream) {
  int gs, t;
  if (vec_shape(factor_num, params_ld, params, nullptr, &gs, &t)) {  // rows as 16-byte chunks
#define CALL(T)                                                                                                  \
  score_vec_kernel<T><<<blocks, kWarps * 32, 0, stream>>>(feature_ids, params, params_ld, feature_vals,          \
                                                          feature_poses, batch, (int)factor_num, factor_lambda,  \
                                                          bias_lambda, pred_score, partials, gs)
    if (t == 1) { CALL(1); } else if (t == 2) { CALL(2); } else { CALL(3); }
#undef CALL
  } else {
    const int rc = dispatch<ScoreLaunch>(factor_num, feature_ids, params, params_ld, feature_vals, feature_poses,
                                         batch, (int)factor_num, factor_lambda, bias_lambda, pred_score,
                                         partials, blocks, stream);
    if (rc != FM_OK) return rc;
  }
  FM_LAUNCH_CHECK();
  return FM_OK;
}

}  // namespace
}  // namespace fm

using namespace fm;

FM_EXPORT size_t fm_score_workspace(int64_t batch) {
  const int64_t blocks = batch <= 0 ? 1 : (batch + kWarps - 1) / kWarps;
  return fm_align_up((size_t)blocks * sizeof(float), 256);
}

FM_EXPORT int fm_score(const int32_t* feature_ids, const float* params, int64_t params_ld,
                       const float* feature_vals, const int32_t* feature_poses, int64_t batch,
                       int64_t factor_num, float factor_lambda, float bias_lambda, float* pred_score,
                       float* reg_score, void* workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  FM_REQUIRE(batch >= 0 && factor_num >= 0 && params_ld >= factor_num + 1, "fm_score: bad shape");
  FM_REQUIRE(reg_score && workspace && feature_poses, "fm_score: null pointer");
  FM_REQUIRE(batch == 0 || (pred_score && params), "fm_score: null pointer");
  if (workspace_bytes < fm_score_workspace(batch))
    return fm_fail(FM_E_WORKSPACE, "fm_score: workspace %zu < %zu", workspace_bytes, fm_score_workspace(batch));
  const unsigned blocks = (unsigned)((batch + kWarps - 1) / kWarps);
  float* partials = reinterpret_cast<float*>(workspace);
  if (batch > 0) {
    const int rc = launch_score(feature_ids, params, params_ld, feature_vals, feature_poses, batch, factor_num,
                                factor_lambda, bias_lambda, pred_score, partials, blocks, stream);
    if (rc != FM_OK) return rc;
  }
  reduce_partials_kernel<<<1, 256, 0, stream>>>(partials, (int)blocks, 1.0f, reg_score);
  FM_LAUNCH_CHECK();
  return FM_OK;
}

FM_EXPORT int fm_grad(const int32_t* feature_ids, const float* params, int64_t params_ld,
                      const float* feature_vals, const int32_t* feature_poses, int64_t batch,
                      int64_t factor_num, float factor_lambda, float bias_lambda, const float* pred_grad,
                      float reg_grad, const float* reg_grad_dev, int64_t param_rows,
                      const int64_t* param_rows_dev, float* params_grad, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  FM_REQUIRE(batch >= 0 && factor_num >= 0 && params_ld >= factor_num + 1 && param_rows >= 0, "fm_grad: bad shape");
  FM_REQUIRE(param_rows == 0 || params_grad, "fm_grad: null pointer");
  FM_REQUIRE(batch == 0 || (params && feature_poses && pred_grad), "fm_grad: null pointer");
  if (param_rows > 0) {
    const size_t bytes = (size_t)param_rows * (size_t)params_ld * sizeof(float);
    if (param_rows_dev && ((uintptr_t)params_grad & 15) == 0 && bytes % 16 == 0) {
      // only the rows in use (count on the device) need clearing
      zero_rows_kernel<<<148 * 8, 256, 0, stream>>>(reinterpret_cast<float4*>(params_grad), (int64_t)(bytes / 16),
                                                    params_ld, param_rows_dev);
      FM_LAUNCH_CHECK();
    } else {
      FM_CUDA(cudaMemsetAsync(params_grad, 0, bytes, stream));
    }
  }
  if (batch == 0) return FM_OK;
  const unsigned blocks = (unsigned)((batch + kWarps - 1) / kWarps);
  int gs, t;
  if (vec_shape(factor_num, params_ld, params, params_grad, &gs, &t)) {
#define CALL(T)                                                                                                 \
  grad_vec_kernel<T><<<blocks, kWarps * 32, 0, stream>>>(feature_ids, params, params_ld, feature_vals,          \
                                                         feature_poses, batch, (int)factor_num, factor_lambda,  \
                                                         bias_lambda, pred_grad, reg_grad, reg_grad_dev,        \
                                                         params_grad, gs)
    if (t == 1) { CALL(1); } else if (t == 2) { CALL(2); } else { CALL(3); }
#undef CALL
  } else {
    const int rc = dispatch<GradLaunch>(factor_num, feature_ids, params, params_ld, feature_vals, feature_poses,
                                        batch, (int)factor_num, factor_lambda, bias_lambda, pred_grad, reg_grad,
                                        reg_grad_dev, params_grad, blocks, stream);
    if (rc != FM_OK) return rc;
  }
  FM_LAUNCH_CHECK();
  return FM_OK;
}

// ---------------------------------------------------------------------------
// fm_score_noalloc: fm_score for callers that have no way to hand over a
// workspace.  (No counterpart in the reference.)
//
// fm_score needs one float of scratch per thread block (the per-block partial
// sums of reg_score).  Some hosts cannot give us that memory: a serving process
// whose framework grabbed the whole device for its own pool at start-up (a
// plugin's cudaMalloc then fails, or is simply forbidden by the host's rules),
// or a scoring thread on a latency budget that must not take the device-wide
// lock and implicit synchronisation cudaMalloc / cudaFree bring with them.
// For them the scratch lives in the library image itself: a fixed array in
// device memory that every GPU loading this library carries until the process
// exits (kNoAllocMaxBlocks floats = 512 KiB).
//
// Rules that follow from sharing one array:
//   * batch is limited to FM_SCORE_NOALLOC_MAX_BATCH examples; larger batches
//     are refused with FM_E_LIMIT (use fm_score with your own workspace).
//   * every call on a device uses the same array and leaves its partial sums
//     behind.  Nothing has to start at zero -- each launch overwrites exactly
//     the entries its own reduction reads -- but two calls that overlap on the
//     GPU (different streams, or several host threads) race on it.  Serialise
//     them, or use fm_score.
//   * the first call on each device looks the array's device address up
//     (cudaGetSymbolAddress, with that device current) and caches it.  Make
//     that first call during start-up, before any stream capture begins.
// The call itself allocates nothing, frees nothing and does not synchronise.
// ---------------------------------------------------------------------------
namespace fm {
namespace {

constexpr int64_t kNoAllocMaxBatch = FM_SCORE_NOALLOC_MAX_BATCH;
constexpr int64_t kNoAllocMaxBlocks = (kNoAllocMaxBatch + kWarps - 1) / kWarps;
constexpr int kMaxDevices = 64;

// In host code this name is only a handle for the runtime; the address kernels
// can use comes from cudaGetSymbolAddress below.
static __device__ __align__(256) float g_score_noalloc_scratch[kNoAllocMaxBlocks];

static float* g_noalloc_ptr[kMaxDevices];  // per device, filled on first use
static std::mutex g_noalloc_mu;

static int noalloc_scratch_for_current_device(float** out) {
  int dev = -1;
  FM_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices)
    return fm_fail(FM_E_LIMIT, "fm_score_noalloc: device ordinal %d outside [0, %d)", dev, kMaxDevices);
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (g_noalloc_ptr[dev] == nullptr) {
    void* p = nullptr;
    FM_CUDA(cudaGetSymbolAddress(&p, g_score_noalloc_scratch));
    if (p == nullptr) return fm_fail(FM_E_CUDA, "fm_score_noalloc: cudaGetSymbolAddress returned a null address");
    g_noalloc_ptr[dev] = static_cast<float*>(p);
  }
  *out = g_noalloc_ptr[dev];
  return FM_OK;
}

}  // namespace
}  // namespace fm

FM_EXPORT int fm_score_noalloc(const int32_t* feature_ids, const float* params, int64_t params_ld,
                               const float* feature_vals, const int32_t* feature_poses, int64_t batch,
                               int64_t factor_num, float factor_lambda, float bias_lambda, float* pred_score,
                               float* reg_score, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  FM_REQUIRE(batch >= 0 && factor_num >= 0 && params_ld >= factor_num + 1, "fm_score_noalloc: bad shape");
  if (batch > kNoAllocMaxBatch)
    return fm_fail(FM_E_LIMIT,
                   "fm_score_noalloc: batch %lld exceeds the built-in scratch (FM_SCORE_NOALLOC_MAX_BATCH = %lld); "
                   "call fm_score with a caller-provided workspace instead",
                   (long long)batch, (long long)kNoAllocMaxBatch);
  FM_REQUIRE(reg_score && feature_poses, "fm_score_noalloc: null pointer");
  FM_REQUIRE(batch == 0 || (pred_score && params), "fm_score_noalloc: null pointer");
  float* partials = nullptr;
  const int rc0 = noalloc_scratch_for_current_device(&partials);
  if (rc0 != FM_OK) return rc0;
  const unsigned blocks = (unsigned)((batch + kWarps - 1) / kWarps);
  if (batch > 0) {   // the same kernels fm_score launches for these shapes
    const int rc = launch_score(feature_ids, params, params_ld, feature_vals, feature_poses, batch, factor_num,
                                factor_lambda, bias_lambda, pred_score, partials, blocks, stream);
    if (rc != FM_OK) return rc;
  }
  reduce_partials_kernel<<<1, 256, 0, stream>>>(partials, (int)blocks, 1.0f, reg_score);
  FM_LAUNCH_CHECK();
  return FM_OK;
}
