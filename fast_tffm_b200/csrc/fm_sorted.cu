// fm_sorted.cu -- FmGrad + the optimizer step as a segment reduce over occurrences sorted by row.
//
// Replaces cc/fm_grad_op.h:80-106 (one atomic float add per occurrence per column into a dense
// [U, K+1] gradient) + TF's SparseApplyAdagrad (tffm/fm_model.py:341-348).  Input: the batch's
// occurrences sorted by table row (fm_sort.cu: keys and (example, x) pairs) and one record
// per example [pg, s_1 .. s_K] (fm_sorted_score).  FmGrad's per-occurrence gradient
//     g_0 = pg x + rg lb b            g_k = rg lf v_k - pg x^2 v_k + pg x s_k
// summed over a row's occurrences is
//     G_0 = C + cnt rg lb b           G_k = v_k (cnt rg lf - D) + B_k
// with C = sum pg x, D = sum pg x^2, B_k = sum pg x s_k, cnt = #occurrences.  One lane group
// (one lane per 16-byte chunk of the row) owns one run of equal ids: it adds the run's terms up in
// registers, in sorted (= batch) order, reads [w | h] of the row once, applies the optimizer and
// writes them back.  No float atomics, no accumulator rows, no third table; the result does not
// depend on scheduling.
//
// Runs are cut at multiples of FM_SORT_CHUNK occurrences (fm_sort.cu), so a hot row under skewed
// ids is many work items, not one long loop: the pieces of a cut run leave partial sums in a small
// scratch array and a second, short kernel adds them up in order and applies the step.
#include "fm_common.cuh"

namespace fm {
namespace {

constexpr unsigned long long kNoError = 0xffffffffffffffffull;
constexpr int kOptAdagrad = 0, kOptSgd = 1;

struct SortedIn {
  const unsigned* keys;      // sorted row keys
  const uint2* pairs;        // (example, x bits) of each sorted occurrence
  int64_t* counts;           // [n, n_unique]: n from the sort, n_unique counted here
  const float* rec;          // [batch, rec_ld] example records [pg, s_1 .. s_K]
  int64_t rec_ld;
};

struct Step {
  float lr, rf, rb;          // learning rate, reg_grad * factor_lambda, reg_grad * bias_lambda
  int opt;                   // kOptAdagrad: acc += g^2, w -= lr g / sqrt(acc);  kOptSgd: w -= lr g
};

__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }

// one occurrence's terms for lane chunk(s) c + t * gs of the row: G[t] = (. | B_k) columns, C, D, cnt
template <int T>
__device__ __forceinline__ void add_occurrence(const SortedIn& in, uint2 pr, int c, int gs, int vc, float4 (&G)[T],
                                               float& C, float& D, float& cnt) {
  const float x = __uint_as_float(pr.y);
  const float* r = in.rec + (int64_t)pr.x * in.rec_ld;
  const float pgx = __ldg(r) * x;
  C += pgx;
  D = fmaf(pgx, x, D);
  cnt += 1.0f;
#pragma unroll
  for (int t = 0; t < T; ++t) {
    const int chunk = c + t * gs;
    if (chunk < vc) {
      const float4 s = __ldg(reinterpret_cast<const float4*>(r + 4 * chunk));
      G[t].x = fmaf(pgx, s.x, G[t].x); G[t].y = fmaf(pgx, s.y, G[t].y);
      G[t].z = fmaf(pgx, s.z, G[t].z); G[t].w = fmaf(pgx, s.w, G[t].w);
    }
  }
}

template <int T>
__device__ __forceinline__ void apply_row(float* __restrict__ table, float* __restrict__ accum, int64_t row, int c,
                                          int gs, int vc, int K, const float4 (&w)[T], const float4 (&h)[T],
                                          const float4 (&G)[T], float D, float cnt, const Step st) {
  const float fac_scale = cnt * st.rf - D, bias_scale = cnt * st.rb;
#pragma unroll
  for (int t = 0; t < T; ++t) {
    const int chunk = c + t * gs;
    if (chunk >= vc) continue;
    const float gv[4] = {G[t].x, G[t].y, G[t].z, G[t].w};
    float wv[4] = {w[t].x, w[t].y, w[t].z, w[t].w};
    float hv[4] = {h[t].x, h[t].y, h[t].z, h[t].w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int col = 4 * chunk + e;
      if (col > K) continue;       // pad columns stay as they are
      const float gr = col == 0 ? fmaf(bias_scale, wv[e], gv[e]) : fmaf(wv[e], fac_scale, gv[e]);
      if (st.opt == kOptAdagrad) {
        hv[e] = fmaf(gr, gr, hv[e]);
        wv[e] = fmaf(-st.lr * gr, rsqrtf(hv[e]), wv[e]);    // TF1 SparseApplyAdagrad: w -= lr g rsqrt(acc)
      } else {
        wv[e] = fmaf(-st.lr, gr, wv[e]);                    // plain SGD
      }
    }
    *reinterpret_cast<float4*>(table + row + 4 * chunk) = make_float4(wv[0], wv[1], wv[2], wv[3]);
    if (st.opt == kOptAdagrad)
      *reinterpret_cast<float4*>(accum + row + 4 * chunk) = make_float4(hv[0], hv[1], hv[2], hv[3]);
  }
}

// One lane group per SORTED POSITION; the positions that open a run of equal keys, or sit on a multiple of
// FM_SORT_CHUNK inside one, do the work (their piece ends at the next such position), the others leave at once.
// gs lanes = the float4 chunks of a K+1 wide row (at most 32; wider rows take T chunks per lane); a warp covers
// 32 / gs positions per step, so keys are read coalesced and a piece's row address is known one L2 round trip
// after the group starts.
template <int T>
__global__ void __launch_bounds__(256)
sorted_update_kernel(float* __restrict__ table, float* __restrict__ accum, int64_t ld, const SortedIn in, int K,
                     const Step st, float* __restrict__ part, int64_t part_ld, unsigned* __restrict__ hot,
                     unsigned* __restrict__ hot_count, const unsigned long long* __restrict__ guard, int gs,
                     unsigned key_mask) {
  if (guard && *guard != kNoError) return;
  const unsigned n = (unsigned)in.counts[0];
  const int lane = threadIdx.x & 31;
  const int per_warp = 32 / gs;
  const int g = lane / gs, c = lane - g * gs;
  const bool lane_on = g < per_warp;
  const int vc = (K + 4) >> 2;
  const int64_t warps_total = (int64_t)gridDim.x * (blockDim.x >> 5);
  const int64_t warp_id = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  unsigned heads = 0;                                  // runs opened by this lane (lane c == 0 of a group counts)
  for (int64_t base = warp_id * per_warp; base < n; base += warps_total * per_warp) {
    const int64_t p64 = base + g;
    if (!(lane_on && p64 < n)) continue;               // no warp-wide operation below
    const unsigned p0 = (unsigned)p64;
    const unsigned key = __ldg(in.keys + p0);
    const bool first = p0 == 0 || __ldg(in.keys + p0 - 1) != key;
    if (!first && (p0 % FM_SORT_CHUNK) != 0) continue; // inside someone else's piece
    unsigned knext = p0 + 1 < n ? __ldg(in.keys + p0 + 1) : ~key;
    uint2 pr = __ldg(in.pairs + p0);
    const int64_t row = (int64_t)(key & key_mask) * ld;
    float4 w[T], h[T], G[T];
    if (first) {                                       // nearly always the whole run: start the row loads now
      heads += c == 0;
#pragma unroll
      for (int t = 0; t < T; ++t) {
        const int chunk = c + t * gs;
        if (chunk < vc) {
          w[t] = ld4(table + row + 4 * chunk);
          h[t] = st.opt == kOptAdagrad ? ld4(accum + row + 4 * chunk) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
    }
#pragma unroll
    for (int t = 0; t < T; ++t) G[t] = make_float4(0.f, 0.f, 0.f, 0.f);
    float C = 0.f, D = 0.f, cnt = 0.f;
    unsigned p1 = p0;
    while (true) {
      add_occurrence<T>(in, pr, c, gs, vc, G, C, D, cnt);
      ++p1;
      if (p1 >= n || (p1 % FM_SORT_CHUNK) == 0 || knext != key) break;
      pr = __ldg(in.pairs + p1);
      knext = p1 + 1 < n ? __ldg(in.keys + p1 + 1) : ~key;
    }
    if (c == 0) G[0].x = C;        // column 0 of a record holds pg, not a factor sum: its column is C itself
    const bool last = p1 >= n || knext != key;         // knext = keys[p1] here
    if (first && last) {
      apply_row<T>(table, accum, row, c, gs, vc, K, w, h, G, D, cnt, st);
      continue;
    }
    // a piece of a cut run: its sums wait in slot 2 * cut (the piece before the cut) or 2 * cut + 1 (after)
    const int64_t slot = !first ? 2 * (int64_t)(p0 / FM_SORT_CHUNK) + 1 : 2 * (int64_t)(p1 / FM_SORT_CHUNK);
    float* dst = part + slot * part_ld;
#pragma unroll
    for (int t = 0; t < T; ++t) {
      const int chunk = c + t * gs;
      if (chunk < vc) *reinterpret_cast<float4*>(dst + 4 * chunk) = G[t];
    }
    if (c == 0) {
      *reinterpret_cast<float4*>(dst + 4 * vc) = make_float4(D, cnt, 0.f, 0.f);
      if (first) hot[atomicAdd(hot_count, 1u)] = p1 / FM_SORT_CHUNK;   // the run's first piece: someone finishes it
    }
  }
  // number of distinct rows (tf.unique's U): one atomic per warp
  heads = (unsigned)warp_sum((int)heads);
  if (lane == 0 && heads) atomicAdd(reinterpret_cast<unsigned long long*>(in.counts + 1), (unsigned long long)heads);
}

// the cut runs: add the pieces' partial sums in order, then the step.  One lane group per run.
template <int T>
__global__ void __launch_bounds__(256)
sorted_finish_kernel(float* __restrict__ table, float* __restrict__ accum, int64_t ld, const SortedIn in, int K,
                     const Step st, const float* __restrict__ part, int64_t part_ld, const unsigned* __restrict__ hot,
                     const unsigned* __restrict__ hot_count, const unsigned long long* __restrict__ guard, int gs,
                     unsigned key_mask) {
  if (guard && *guard != kNoError) return;
  const unsigned n_hot = *hot_count;
  const unsigned n = (unsigned)in.counts[0];
  const int lane = threadIdx.x & 31;
  const int per_warp = 32 / gs;
  const int g = lane / gs, c = lane - g * gs;
  const int vc = (K + 4) >> 2;
  const int64_t warps_total = (int64_t)gridDim.x * (blockDim.x >> 5);
  const int64_t warp_id = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  for (int64_t base = warp_id * per_warp; base < n_hot; base += warps_total * per_warp) {
    const int64_t e = base + g;
    if (!(g < per_warp && e < n_hot)) continue;
    const unsigned cut0 = hot[e];
    const unsigned key = in.keys[cut0 * FM_SORT_CHUNK - 1];
    const int64_t row = (int64_t)(key & key_mask) * ld;
    float4 w[T], h[T], G[T];
#pragma unroll
    for (int t = 0; t < T; ++t) {
      const int chunk = c + t * gs;
      G[t] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (chunk < vc) {
        w[t] = ld4(table + row + 4 * chunk);
        h[t] = st.opt == kOptAdagrad ? ld4(accum + row + 4 * chunk) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
    float D = 0.f, cnt = 0.f;
    auto add = [&](int64_t slot) {
      const float* src = part + slot * part_ld;
#pragma unroll
      for (int t = 0; t < T; ++t) {
        const int chunk = c + t * gs;
        if (chunk < vc) {
          const float4 v = ld4(src + 4 * chunk);
          G[t].x += v.x; G[t].y += v.y; G[t].z += v.z; G[t].w += v.w;
        }
      }
      const float4 tail = ld4(src + 4 * vc);
      D += tail.x; cnt += tail.y;
    };
    add(2 * (int64_t)cut0);
    for (unsigned cut = cut0; (uint64_t)cut * FM_SORT_CHUNK < n && in.keys[cut * FM_SORT_CHUNK] == key; ++cut)
      add(2 * (int64_t)cut + 1);
    apply_row<T>(table, accum, row, c, gs, vc, K, w, h, G, D, cnt, st);
  }
}

static inline int64_t part_ld_of(int64_t factor_num) { return (factor_num + 4) / 4 * 4 + 4; }
static inline int64_t cuts_of(int64_t capacity) { return capacity / FM_SORT_CHUNK + 2; }

}  // namespace
}  // namespace fm

using namespace fm;

FM_EXPORT size_t fm_sorted_workspace(int64_t capacity, int64_t factor_num) {
  if (capacity < 0 || factor_num < 0) return 0;
  return fm_align_up(256 + sizeof(unsigned) * (size_t)cuts_of(capacity), 256) +
         fm_align_up(sizeof(float) * 2 * (size_t)cuts_of(capacity) * (size_t)part_ld_of(factor_num), 256);
}

FM_EXPORT int fm_sorted_update(float* table, float* accum, int64_t table_ld, const uint32_t* sorted_keys,
                               const uint32_t* sorted_pairs, int64_t* counts, int64_t capacity, const float* records, int64_t record_ld, int64_t factor_num,
                               int optimizer, float learning_rate, float factor_lambda, float bias_lambda,
                               float reg_grad, int64_t row_bits, const uint64_t* guard, void* workspace,
                               size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  FM_REQUIRE(capacity >= 0 && factor_num >= 0, "fm_sorted_update: bad shape");
  FM_REQUIRE(table_ld >= factor_num + 1 && table_ld % 4 == 0, "fm_sorted_update: table_ld must be a multiple of 4 and >= K+1");
  FM_REQUIRE(record_ld >= (factor_num + 4) / 4 * 4 && record_ld % 4 == 0, "fm_sorted_update: bad record_ld");
  FM_REQUIRE(optimizer == FM_OPT_ADAGRAD || optimizer == FM_OPT_SGD, "fm_sorted_update: unknown optimizer");
  FM_REQUIRE(row_bits >= 0 && row_bits <= 32, "fm_sorted_update: bad row_bits");
  if (capacity == 0) return FM_OK;
  FM_REQUIRE(table && (accum || optimizer == FM_OPT_SGD) && sorted_keys && sorted_pairs && counts &&
             records && workspace, "fm_sorted_update: null pointer");
  FM_REQUIRE((((uintptr_t)table | (uintptr_t)accum | (uintptr_t)records) & 15) == 0 && ((uintptr_t)sorted_pairs & 7) == 0,
             "fm_sorted_update: table, accum, records 16-byte aligned; sorted_pairs 8-byte aligned");
  if (workspace_bytes < fm_sorted_workspace(capacity, factor_num))
    return fm_fail(FM_E_WORKSPACE, "fm_sorted_update: workspace %zu < %zu", workspace_bytes,
                   fm_sorted_workspace(capacity, factor_num));
  const int64_t chunks = (factor_num + 4) / 4;
  const int gs = chunks <= 32 ? (int)chunks : 32;
  const int t = (int)((chunks + gs - 1) / gs);
  if (t > 3) return fm_fail(FM_E_LIMIT, "fm_sorted_update: factor_num %lld beyond the sorted path (use the claim path)",
                            (long long)factor_num);
  char* ws = static_cast<char*>(workspace);
  unsigned* hot_count = reinterpret_cast<unsigned*>(ws);
  unsigned* hot = reinterpret_cast<unsigned*>(ws + 256);
  float* part = reinterpret_cast<float*>(ws + fm_align_up(256 + sizeof(unsigned) * (size_t)cuts_of(capacity), 256));
  const int64_t part_ld = part_ld_of(factor_num);
  FM_CUDA(cudaMemsetAsync(hot_count, 0, sizeof(unsigned), stream));
  FM_CUDA(cudaMemsetAsync(counts + 1, 0, sizeof(int64_t), stream));
  SortedIn in{sorted_keys, reinterpret_cast<const uint2*>(sorted_pairs), counts, records, record_ld};
  Step st{learning_rate, reg_grad * factor_lambda, reg_grad * bias_lambda, optimizer == FM_OPT_SGD ? kOptSgd : kOptAdagrad};
  // sharded tables sort by owner << row_bits | row: the owner's shard is indexed by the row alone
  const unsigned key_mask = row_bits == 0 || row_bits >= 32 ? 0xffffffffu : ((1u << row_bits) - 1u);
  const unsigned long long* g = reinterpret_cast<const unsigned long long*>(guard);
  const int64_t per_block = 8 * (32 / gs);
  int64_t blocks = (capacity + per_block - 1) / per_block;
  if (blocks > 148 * 16) blocks = 148 * 16;
  int64_t fin_blocks = (cuts_of(capacity) + per_block - 1) / per_block;
  if (fin_blocks > 148 * 4) fin_blocks = 148 * 4;
#define CALL(T)                                                                                                      \
  sorted_update_kernel<T><<<(unsigned)blocks, 256, 0, stream>>>(table, accum, table_ld, in, (int)factor_num, st,    \
                                                                part, part_ld, hot, hot_count, g, gs, key_mask);    \
  FM_LAUNCH_CHECK();                                                                                                 \
  sorted_finish_kernel<T><<<(unsigned)fin_blocks, 256, 0, stream>>>(table, accum, table_ld, in, (int)factor_num, st, \
                                                                    part, part_ld, hot, hot_count, g, gs, key_mask)
  if (t == 1) { CALL(1); } else if (t == 2) { CALL(2); } else { CALL(3); }
#undef CALL
  FM_LAUNCH_CHECK();
  return FM_OK;
}
