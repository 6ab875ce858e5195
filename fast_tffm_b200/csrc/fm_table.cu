// fm_table.cu -- the stock-TF ops around the custom ops, on the vocabulary table:
//   fm_lookup          tf.nn.embedding_lookup(vocab_blocks, ori_ids)   tffm/fm_model.py:291
//   fm_loss_grad       weighted logistic / mse mean loss + d/dpred     tffm/fm_model.py:311-326
//   fm_sparse_adagrad  AdagradOptimizer's sparse row update            tffm/fm_model.py:341-348
// Rows are 68..1028 bytes; a row is spread over LPR lanes (LPR = min(32, pow2 >=
// K+1)) so that a warp moves 32/LPR whole rows per step with coalesced accesses.
#include "fm_common.cuh"

namespace fm {
namespace {

__device__ __forceinline__ int64_t dev_count(int64_t n_max, const int64_t* n_dev) {
  return n_dev ? min(*n_dev, n_max) : n_max;
}

template <int LPR>
__global__ void __launch_bounds__(256)
lookup_kernel(const float* __restrict__ table, int64_t table_ld, const int64_t* __restrict__ ori_ids,
              int64_t n_max, const int64_t* __restrict__ n_dev, int width, float* __restrict__ out,
              int64_t out_ld, const unsigned long long* __restrict__ guard) {
  if (guard && *guard != 0xffffffffffffffffull) return;  // ids of a failed parse are not to be trusted
  const int64_t n = dev_count(n_max, n_dev);
  const int c = threadIdx.x % LPR;
  const int64_t rows_per_block = blockDim.x / LPR;
  for (int64_t u = (int64_t)blockIdx.x * rows_per_block + threadIdx.x / LPR; u < n;
       u += (int64_t)gridDim.x * rows_per_block) {
    const float* src = table + ori_ids[u] * table_ld;
    float* dst = out + u * out_ld;
    for (int k = c; k < width; k += LPR) dst[k] = __ldg(src + k);
  }
}

// float4 flavour: thread per (row, 16-byte chunk), no idle lanes
__global__ void __launch_bounds__(256)
lookup_vec_kernel(const float* __restrict__ table, int64_t table_ld, const int64_t* __restrict__ ori_ids,
                  int64_t n_max, const int64_t* __restrict__ n_dev, int chunks, float* __restrict__ out,
                  int64_t out_ld, const unsigned long long* __restrict__ guard) {
  if (guard && *guard != 0xffffffffffffffffull) return;
  const int64_t total = dev_count(n_max, n_dev) * chunks;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t u = t / chunks;
    const int c = (int)(t - u * chunks);
    *reinterpret_cast<float4*>(out + u * out_ld + 4 * c) =
        __ldg(reinterpret_cast<const float4*>(table + ori_ids[u] * table_ld + 4 * c));
  }
}

// acc += g^2 ; w -= lr * g / sqrt(acc)   (TF1 SparseApplyAdagrad: no epsilon; every
// listed row is touched even when g == 0)
template <int LPR>
__global__ void __launch_bounds__(256)
adagrad_kernel(float* __restrict__ table, float* __restrict__ accum, int64_t table_ld,
               const int64_t* __restrict__ ori_ids, int64_t n_max, const int64_t* __restrict__ n_dev,
               int width, const float* __restrict__ grad, int64_t grad_ld, float lr,
               const unsigned long long* __restrict__ guard) {
  if (guard && *guard != 0xffffffffffffffffull) return;  // the batch failed to parse: no update
  const int64_t n = dev_count(n_max, n_dev);
  const int c = threadIdx.x % LPR;
  const int64_t rows_per_block = blockDim.x / LPR;
  for (int64_t u = (int64_t)blockIdx.x * rows_per_block + threadIdx.x / LPR; u < n;
       u += (int64_t)gridDim.x * rows_per_block) {
    const int64_t base = ori_ids[u] * table_ld;
    const float* g = grad + u * grad_ld;
    for (int k = c; k < width; k += LPR) {
      const float gv = __ldg(g + k);
      const float a = fmaf(gv, gv, accum[base + k]);
      accum[base + k] = a;
      table[base + k] -= __fdiv_rn(lr * gv, __fsqrt_rn(a));
    }
  }
}

// float4 flavour: thread per (row, 16-byte chunk); columns >= width (row padding) pass through untouched
__global__ void __launch_bounds__(256)
adagrad_vec_kernel(float* __restrict__ table, float* __restrict__ accum, int64_t table_ld,
                   const int64_t* __restrict__ ori_ids, int64_t n_max, const int64_t* __restrict__ n_dev,
                   int width, int chunks, const float* __restrict__ grad, int64_t grad_ld, float lr,
                   const unsigned long long* __restrict__ guard) {
  if (guard && *guard != 0xffffffffffffffffull) return;
  const int64_t total = dev_count(n_max, n_dev) * chunks;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t u = t / chunks;
    const int c = (int)(t - u * chunks);
    const int64_t at = ori_ids[u] * table_ld + 4 * c;
    const float4 g4 = __ldg(reinterpret_cast<const float4*>(grad + u * grad_ld + 4 * c));
    const float4 a4 = *reinterpret_cast<const float4*>(accum + at);
    const float4 w4 = *reinterpret_cast<const float4*>(table + at);
    const float g[4] = {g4.x, g4.y, g4.z, g4.w};
    float a[4] = {a4.x, a4.y, a4.z, a4.w};
    float w[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      if (4 * c + e < width) {
        a[e] = fmaf(g[e], g[e], a[e]);
        w[e] -= __fdiv_rn(lr * g[e], __fsqrt_rn(a[e]));
      }
    }
    *reinterpret_cast<float4*>(accum + at) = make_float4(a[0], a[1], a[2], a[3]);
    *reinterpret_cast<float4*>(table + at) = make_float4(w[0], w[1], w[2], w[3]);
  }
}

template <template <int> class L, class... Args>
void dispatch_width(int width, Args... args) {
  if (width <= 4) L<4>::run(args...);
  else if (width <= 8) L<8>::run(args...);
  else if (width <= 16) L<16>::run(args...);
  else L<32>::run(args...);
}

static inline unsigned row_blocks(int64_t n, int lpr) {
  const int64_t rows_per_block = 256 / lpr;
  int64_t b = (n + rows_per_block - 1) / rows_per_block;
  if (b > 148 * 32) b = 148 * 32;
  return (unsigned)(b < 1 ? 1 : b);
}

template <int LPR>
struct LookupLaunch {
  static void run(const float* table, int64_t table_ld, const int64_t* ori_ids, int64_t n, const int64_t* n_dev,
                  int width, float* out, int64_t out_ld, const unsigned long long* guard, cudaStream_t stream) {
    lookup_kernel<LPR><<<row_blocks(n, LPR), 256, 0, stream>>>(table, table_ld, ori_ids, n, n_dev, width, out, out_ld,
                                                               guard);
  }
};
template <int LPR>
struct AdagradLaunch {
  static void run(float* table, float* accum, int64_t table_ld, const int64_t* ori_ids, int64_t n,
                  const int64_t* n_dev, int width, const float* grad, int64_t grad_ld, float lr,
                  const unsigned long long* guard, cudaStream_t stream) {
    adagrad_kernel<LPR><<<row_blocks(n, LPR), 256, 0, stream>>>(table, accum, table_ld, ori_ids, n, n_dev, width,
                                                                grad, grad_ld, lr, guard);
  }
};

// loss_i = w_i * l(pred_i, y_i); per-block partial sums; pred_grad_i = w_i * l'(.) / batch
__global__ void __launch_bounds__(256)
loss_kernel(const float* __restrict__ pred, const float* __restrict__ labels, const float* __restrict__ weights,
            int64_t batch, int kind, float* __restrict__ pred_grad, float* __restrict__ partials) {
  __shared__ float s_w[8];
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  float l = 0.f;
  if (i < batch) {
    const float x = pred[i], y = labels[i], w = weights ? weights[i] : 1.0f;
    float d;
    if (kind == FM_LOSS_LOGISTIC) {
      // tf.nn.sigmoid_cross_entropy_with_logits: max(x,0) - x*y + log1p(exp(-|x|))
      l = fmaxf(x, 0.f) - x * y + log1pf(expf(-fabsf(x)));
      d = 1.0f / (1.0f + expf(-x)) - y;
    } else {
      l = (x - y) * (x - y);
      d = 2.0f * (x - y);
    }
    l *= w;
    pred_grad[i] = w * d / (float)batch;
  }
  l = warp_sum(l);
  if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = l;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += s_w[w];
    partials[blockIdx.x] = t;
  }
}

}  // namespace
}  // namespace fm

using namespace fm;

FM_EXPORT int fm_lookup(const float* table, int64_t table_ld, const int64_t* ori_ids, int64_t n_unique,
                        const int64_t* n_unique_dev, int64_t factor_num, float* out, int64_t out_ld,
                        const uint64_t* guard, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  const int64_t width = factor_num + 1;
  FM_REQUIRE(n_unique >= 0 && factor_num >= 0 && table_ld >= width && out_ld >= width, "fm_lookup: bad shape");
  if (n_unique == 0) return FM_OK;
  FM_REQUIRE(table && ori_ids && out, "fm_lookup: null pointer");
  const int64_t chunks = (width + 3) / 4;
  if (table_ld % 4 == 0 && out_ld % 4 == 0 && out_ld >= 4 * chunks && table_ld >= 4 * chunks &&
      (((uintptr_t)table | (uintptr_t)out) & 15) == 0) {
    // rows as 16-byte chunks (the pad columns travel along)
    int64_t blocks = (n_unique * chunks + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    lookup_vec_kernel<<<(unsigned)blocks, 256, 0, stream>>>(table, table_ld, ori_ids, n_unique, n_unique_dev, (int)chunks,
                                                           out, out_ld, reinterpret_cast<const unsigned long long*>(guard));
  } else {
    dispatch_width<LookupLaunch>((int)width, table, table_ld, ori_ids, n_unique, n_unique_dev, (int)width, out,
                                 out_ld, reinterpret_cast<const unsigned long long*>(guard), stream);
  }
  FM_LAUNCH_CHECK();
  return FM_OK;
}

FM_EXPORT int fm_sparse_adagrad(float* table, float* accum, int64_t table_ld, const int64_t* ori_ids,
                                int64_t n_unique, const int64_t* n_unique_dev, int64_t factor_num,
                                const float* grad, int64_t grad_ld, float learning_rate,
                                const uint64_t* guard, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  const int64_t width = factor_num + 1;
  FM_REQUIRE(n_unique >= 0 && factor_num >= 0 && table_ld >= width && grad_ld >= width, "fm_sparse_adagrad: bad shape");
  if (n_unique == 0) return FM_OK;
  FM_REQUIRE(table && accum && ori_ids && grad, "fm_sparse_adagrad: null pointer");
  const int64_t chunks = (width + 3) / 4;
  if (table_ld % 4 == 0 && grad_ld % 4 == 0 && table_ld >= 4 * chunks && grad_ld >= 4 * chunks &&
      (((uintptr_t)table | (uintptr_t)accum | (uintptr_t)grad) & 15) == 0) {
    int64_t blocks = (n_unique * chunks + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    adagrad_vec_kernel<<<(unsigned)blocks, 256, 0, stream>>>(table, accum, table_ld, ori_ids, n_unique, n_unique_dev,
                                                            (int)width, (int)chunks, grad, grad_ld, learning_rate,
                                                            reinterpret_cast<const unsigned long long*>(guard));
  } else {
    dispatch_width<AdagradLaunch>((int)width, table, accum, table_ld, ori_ids, n_unique, n_unique_dev, (int)width,
                                  grad, grad_ld, learning_rate, reinterpret_cast<const unsigned long long*>(guard),
                                  stream);
  }
  FM_LAUNCH_CHECK();
  return FM_OK;
}

FM_EXPORT size_t fm_loss_workspace(int64_t batch) {
  const int64_t blocks = batch <= 0 ? 1 : (batch + 255) / 256;
  return fm_align_up((size_t)blocks * sizeof(float), 256);
}

FM_EXPORT int fm_loss_grad(const float* pred, const float* labels, const float* weights, int64_t batch,
                           int loss_kind, float* loss, float* pred_grad, void* workspace,
                           size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  FM_REQUIRE(batch >= 0, "fm_loss_grad: negative batch");
  FM_REQUIRE(loss_kind == FM_LOSS_LOGISTIC || loss_kind == FM_LOSS_MSE, "fm_loss_grad: unknown loss kind %d", loss_kind);
  FM_REQUIRE(loss && workspace, "fm_loss_grad: null pointer");
  FM_REQUIRE(batch == 0 || (pred && labels && pred_grad), "fm_loss_grad: null pointer");
  if (workspace_bytes < fm_loss_workspace(batch))
    return fm_fail(FM_E_WORKSPACE, "fm_loss_grad: workspace %zu < %zu", workspace_bytes, fm_loss_workspace(batch));
  const unsigned blocks = (unsigned)((batch + 255) / 256);
  float* partials = reinterpret_cast<float*>(workspace);
  if (batch > 0) {
    loss_kernel<<<blocks, 256, 0, stream>>>(pred, labels, weights, batch, loss_kind, pred_grad, partials);
    FM_LAUNCH_CHECK();
  }
  reduce_partials_kernel<<<1, 256, 0, stream>>>(partials, (int)blocks, batch > 0 ? 1.0f / (float)batch : 0.f, loss);
  FM_LAUNCH_CHECK();
  return FM_OK;
}
