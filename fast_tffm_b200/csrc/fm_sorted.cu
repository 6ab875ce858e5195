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
  unsigned heads = 0;                                  // runs opened, counted by lane c == 0 of the group that walks them
  // a warp step covers 32 sorted positions: every lane tests one for "opens a piece", then the lane groups take the
  // pieces found, per_warp at a time (under skewed ids most positions sit inside somebody's piece)
  for (int64_t base = warp_id * 32; base < n; base += warps_total * 32) {
    const int64_t mine = base + lane;
    bool opens = false;
    if (mine < n) {
      const unsigned k0 = __ldg(in.keys + mine);
      opens = mine == 0 || __ldg(in.keys + mine - 1) != k0 || (mine % FM_SORT_CHUNK) == 0;
    }
    const unsigned opening = __ballot_sync(kFull, opens);
    const int n_open = __popc(opening);
    for (int taken = 0; taken < n_open; taken += per_warp) {       // warp-uniform
    if (!(lane_on && taken + g < n_open)) continue;                // no warp-wide operation below
    const unsigned p0 = (unsigned)(base + __fns(opening, 0, taken + g + 1));
    const unsigned key = __ldg(in.keys + p0);
    const bool first = p0 == 0 || __ldg(in.keys + p0 - 1) != key;
    // the piece ends at the next key change, multiple of FM_SORT_CHUNK or n (one more key read in the usual case)
    unsigned p1 = p0 + 1;
    unsigned knext = p1 < n ? __ldg(in.keys + p1) : ~key;
    while (p1 < n && (p1 % FM_SORT_CHUNK) != 0 && knext == key) {
      ++p1;
      knext = p1 < n ? __ldg(in.keys + p1) : ~key;
    }
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
    constexpr int U = 2;                               // occurrences whose gathers are in flight together
    for (unsigned p = p0; p < p1; p += U) {
      uint2 pr[U];
#pragma unroll
      for (int u = 0; u < U; ++u) pr[u] = p + u < p1 ? __ldg(in.pairs + p + u) : make_uint2(0u, 0u);
      float pg[U];
      float4 sv[U][T];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (p + u >= p1) continue;
        const float* r = in.rec + (int64_t)pr[u].x * in.rec_ld;
        pg[u] = __ldg(r);
#pragma unroll
        for (int t = 0; t < T; ++t)
          if (c + t * gs < vc) sv[u][t] = __ldg(reinterpret_cast<const float4*>(r + 4 * (c + t * gs)));
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {                    // added in sorted (= batch) order
        if (p + u >= p1) continue;
        const float x = __uint_as_float(pr[u].y);
        const float pgx = pg[u] * x;
        C += pgx;
        D = fmaf(pgx, x, D);
        cnt += 1.0f;
#pragma unroll
        for (int t = 0; t < T; ++t) {
          if (c + t * gs < vc) {
            G[t].x = fmaf(pgx, sv[u][t].x, G[t].x); G[t].y = fmaf(pgx, sv[u][t].y, G[t].y);
            G[t].z = fmaf(pgx, sv[u][t].z, G[t].z); G[t].w = fmaf(pgx, sv[u][t].w, G[t].w);
          }
        }
      }
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
  }
  // number of distinct rows (tf.unique's U): one atomic per warp
  heads = (unsigned)warp_sum((int)heads);
  if (lane == 0 && heads) atomicAdd(reinterpret_cast<unsigned long long*>(in.counts + 1), (unsigned long long)heads);
}

// the cut runs: add the pieces' partial sums, then the step.  One WARP per run: its lane groups take the pieces
// round-robin (a very hot row has hundreds of them), the groups' sums are combined by shuffles in a fixed order,
// group 0 applies the step.  Same result every time.
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
  const bool lane_on = g < per_warp;
  const int vc = (K + 4) >> 2;
  const int64_t warps_total = (int64_t)gridDim.x * (blockDim.x >> 5);
  const int64_t warp_id = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  for (int64_t e = warp_id; e < n_hot; e += warps_total) {          // warp-uniform
    const unsigned cut0 = hot[e];
    const unsigned key = in.keys[cut0 * FM_SORT_CHUNK - 1];
    const int64_t row = (int64_t)(key & key_mask) * ld;
    float4 w[T], h[T], G[T];
#pragma unroll
    for (int t = 0; t < T; ++t) {
      const int chunk = c + t * gs;
      G[t] = make_float4(0.f, 0.f, 0.f, 0.f);
      w[t] = G[t]; h[t] = G[t];
      if (lane_on && g == 0 && chunk < vc) {
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
    if (lane_on) {
      if (g == 0) add(2 * (int64_t)cut0);                           // the run's first piece
      // the pieces after cut0, cut0 + 1, ...: group g takes every per_warp-th; the keys are sorted, so the first
      // cut that no longer continues the run ends the group's walk
      for (unsigned cut = cut0 + g; (uint64_t)cut * FM_SORT_CHUNK < n && in.keys[cut * FM_SORT_CHUNK] == key;
           cut += per_warp)
        add(2 * (int64_t)cut + 1);
    }
    // group totals, added in group order on every lane
    float Dt = 0.f, ct = 0.f;
    float4 Gt[T];
#pragma unroll
    for (int t = 0; t < T; ++t) Gt[t] = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int g2 = 0; g2 < per_warp; ++g2) {
      const int src = c + g2 * gs;
      Dt += __shfl_sync(kFull, D, src);
      ct += __shfl_sync(kFull, cnt, src);
#pragma unroll
      for (int t = 0; t < T; ++t) {
        Gt[t].x += __shfl_sync(kFull, G[t].x, src); Gt[t].y += __shfl_sync(kFull, G[t].y, src);
        Gt[t].z += __shfl_sync(kFull, G[t].z, src); Gt[t].w += __shfl_sync(kFull, G[t].w, src);
      }
    }
    if (lane_on && g == 0) apply_row<T>(table, accum, row, c, gs, vc, K, w, h, Gt, Dt, ct, st);
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
  // Few resident warps on purpose: read + write-back of rows in ascending order runs fastest at about 24 warps per SM
  // (profiles/r2_bulk_rmw_probe.jsonl: 0.235 ms at 24, 0.29 at 32, 0.55 at 64 for one cfg2 batch), and the
  // warps and registers left over are what the parse of the next batch runs in.
  static const int bpsm = fm_env_int("FM_UPDATE_BLOCKS_PER_SM", 4);
  if (blocks > 148 * bpsm) blocks = 148 * bpsm;
  int64_t fin_blocks = (cuts_of(capacity) + 7) / 8;                  // a warp per cut run
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
