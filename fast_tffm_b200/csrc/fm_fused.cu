// fm_fused.cu -- the training hot path with every table row read once.
//
// The reference runs, per step, embedding_lookup -> FmScorer -> loss -> FmGrad ->
// SparseApplyAdagrad (tffm/fm_model.py:291-348) and each of those walks the
// gathered rows again.  Here:
//
//   fm_fused_score_scatter  one warp per example: gathers the example's rows
//       straight from the vocabulary table with 128-bit loads, computes
//       pred (cc/fm_scorer_op.h:22-36, :60-86, :134-136), the example's loss term
//       and d loss / d pred (tffm/fm_model.py:311-326), then scatters the
//       example's contribution to the gradient of each of its rows.
//       FmGrad's per-occurrence gradient (cc/fm_grad_op.h:88-97)
//           g_0 = pg x + rg lb b            g_k = rg lf v_k - pg x^2 v_k + pg x s_k
//       summed over the occurrences of one table row factorises as
//           G_0 = C + cnt rg lb b           G_k = v_k (cnt rg lf - D) + B_k
//       with C = sum pg x, D = sum pg x^2, B_k = sum pg x s_k, cnt = #occurrences,
//       none of which needs the row itself.  So the kernel adds {C, B_1..B_K, D,
//       cnt} to a compact accumulator row per distinct id (columns 0, 1..K, K+1,
//       K+2) with vector float atomics (red.global.add.v4.f32) and never reads the
//       rows a second time.
//   fm_fused_adagrad  one lane group per distinct id: reads the accumulator row
//       (sequential), the weight and Adagrad rows (random, once), forms G, applies
//       acc += G^2 ; w -= lr G / sqrt(acc)  (TF1 SparseApplyAdagrad) and clears
//       the accumulator row for the next step.
//
// Rows are handled as float4 chunks: row strides (floats) are multiples of 4,
// pointers 16-byte aligned, and the columns K+1 .. 4*ceil((K+1)/4)-1 of a table row
// are zero.  Strides may be larger than the rows (e.g. weight, Adagrad and gradient
// accumulator rows of one id interleaved in one allocation).
#include "fm_common.cuh"
#include "fm_scan.cuh"

namespace fm {
namespace {

constexpr int kWarps = 8;
constexpr unsigned long long kNoError = 0xffffffffffffffffull;

__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }

__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// sys-scope flavour for accumulator rows in a peer GPU's memory (NVLink atomics)
__device__ __forceinline__ void red_add_v4_sys(float* p, float a, float b, float c, float d) {
  asm volatile("red.relaxed.sys.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}
__device__ __forceinline__ void red_or_sys(unsigned* p, unsigned v) {
  asm volatile("red.relaxed.sys.global.or.b32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// The table sharded over the GPUs of one NVLink domain: rank id % n owns row id / n.  Every rank's
// shard is mapped into this process, so a row is one pointer away wherever it lives.
constexpr int kMaxPeers = 16;
struct PeerPtrs {
  const float* table[kMaxPeers];   // shard base of every rank
  float* gtab[kMaxPeers];          // gradient accumulator rows beside them
  unsigned* touched[kMaxPeers];    // bitmap: bit r set <=> row r of that shard got a contribution
  int n;
};
__device__ __forceinline__ void split_id(int64_t id, int n, int* owner, int64_t* row) {
  if ((uint64_t)id >> 32) { *owner = (int)((uint64_t)id % (uint64_t)n); *row = (int64_t)((uint64_t)id / (uint64_t)n); }
  else { const unsigned u = (unsigned)id; *owner = (int)(u % (unsigned)n); *row = (int64_t)(u / (unsigned)n); }
}

template <int LPR>
__device__ __forceinline__ float sum_in_group(float v) {  // over the LPR lanes of one row
#pragma unroll
  for (int o = LPR / 2; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
  return v;
}

// gs lanes share a row (gs = the float4 chunks of a K+3 wide row, at most 32); lane c of a
// group owns chunks c, c + gs, ... (T of them); a warp walks 32 / gs rows per step.
// PEER: rows and accumulator rows are addressed through `peers` (id -> owner shard) instead
// of table / gacc; reads and vector atomics then travel over NVLink where the owner is remote.
template <int T, bool PEER>
__global__ void __launch_bounds__(kWarps * 32)
fused_score_scatter_kernel(const int64_t* __restrict__ ids, const int32_t* __restrict__ local,
                           const float* __restrict__ vals, const int32_t* __restrict__ poses,
                           const float* __restrict__ labels, const float* __restrict__ weights, int64_t batch,
                           int K, const float* __restrict__ table, int64_t ld, float lam_f, float lam_b,
                           int loss_kind, float inv_batch, float* __restrict__ pred, float* __restrict__ gacc, int64_t gacc_ld,
                           float* __restrict__ partials, const unsigned long long* __restrict__ guard,
                           const PeerPtrs peers, int gs, float* __restrict__ emit, int64_t emit_ld) {
  __shared__ float s_loss[kWarps], s_reg[kWarps];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int RPW = 32 / gs;                       // rows per warp step
  const int r = lane / gs, c = lane - r * gs;
  const bool lane_on = r < RPW;                  // 32 % gs trailing lanes idle
  const int vc_table = (K + 4) >> 2, vc_acc = (K + 6) >> 2;  // float4 chunks of a K+1 / K+3 wide row
  const bool active = !(guard && *guard != kNoError);
  float loss_acc = 0.f, reg_acc = 0.f;

  for (int64_t i = (int64_t)blockIdx.x * kWarps + warp; active && i < batch; i += (int64_t)gridDim.x * kWarps) {
    const int start = poses[i], end = poses[i + 1];
    float4 S[T];
#pragma unroll
    for (int t = 0; t < T; ++t) S[t] = make_float4(0.f, 0.f, 0.f, 0.f);
    float q = 0.f, bias_dot = 0.f, bias_sq = 0.f, fac_sq = 0.f;

    // ---- pass A: gather the rows once ------------------------------------------
    for (int base = start; base < end; base += 32) {
      const int j = base + lane;
      unsigned long long my_row = (unsigned long long)table;
      if (j < end) {
        if (PEER) {
          int owner; int64_t row;
          split_id(ids[j], peers.n, &owner, &row);
          my_row = (unsigned long long)(peers.table[owner] + row * ld);
        } else {
          my_row = (unsigned long long)(table + ids[j] * ld);
        }
      }
      const float my_x = j < end ? vals[j] : 0.f;
      const int rows_here = min(32, end - base);
#pragma unroll 4
      for (int s = 0; s * RPW < rows_here; ++s) {  // warp-uniform
        const int jj = s * RPW + r;
        const float* rowp = reinterpret_cast<const float*>(__shfl_sync(kFull, my_row, jj & 31));
        const float x = __shfl_sync(kFull, my_x, jj & 31);
        if (lane_on && jj < rows_here) {
#pragma unroll
          for (int t = 0; t < T; ++t) {
            const int chunk = c + t * gs;
            if (chunk < vc_table) {
              float4 v = ldg4(rowp + 4 * chunk);
              if (chunk == 0) {  // column 0 is the bias
                bias_dot = fmaf(x, v.x, bias_dot);
                bias_sq = fmaf(v.x, v.x, bias_sq);
                v.x = 0.f;
              }
              // pad columns of the table are zero, so they drop out of every sum
              const float a0 = x * v.x, a1 = x * v.y, a2 = x * v.z, a3 = x * v.w;
              S[t].x += a0; S[t].y += a1; S[t].z += a2; S[t].w += a3;
              q = fmaf(a0, a0, fmaf(a1, a1, fmaf(a2, a2, fmaf(a3, a3, q))));
              fac_sq = fmaf(v.x, v.x, fmaf(v.y, v.y, fmaf(v.z, v.z, fmaf(v.w, v.w, fac_sq))));
            }
          }
        }
      }
    }
    float second = 0.f;
#pragma unroll
    for (int t = 0; t < T; ++t) {
      // column totals over the row groups (lanes c, c + gs, ...); every lane ends up with its column's total
      float4 tot = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int g2 = 0; g2 < RPW; ++g2) {
        const int src = c + g2 * gs;
        tot.x += __shfl_sync(kFull, S[t].x, src);
        tot.y += __shfl_sync(kFull, S[t].y, src);
        tot.z += __shfl_sync(kFull, S[t].z, src);
        tot.w += __shfl_sync(kFull, S[t].w, src);
      }
      S[t] = tot;
      second += S[t].x * S[t].x + S[t].y * S[t].y + S[t].z * S[t].z + S[t].w * S[t].w;
    }
    if (r != 0) second = 0.f;  // every row group holds the same column totals: count them once
    const float p = warp_sum(bias_dot + 0.5f * (second - q));
    reg_acc += warp_sum(0.5f * lam_f * fac_sq + 0.5f * lam_b * bias_sq);

    if (labels == nullptr) {  // scoring only: no loss, no scatter
      if (lane == 0) pred[i] = p;
      continue;
    }
    // ---- loss term and d loss / d pred (tffm/fm_model.py:311-326) ------------------
    const float y = labels[i], w = weights ? weights[i] : 1.0f;
    float l, d;
    if (loss_kind == FM_LOSS_LOGISTIC) {
      l = fmaxf(p, 0.f) - p * y + log1pf(expf(-fabsf(p)));
      d = 1.0f / (1.0f + expf(-p)) - y;
    } else {
      l = (p - y) * (p - y);
      d = 2.0f * (p - y);
    }
    const float pg = w * d * inv_batch;
    if (lane == 0) {
      pred[i] = p;
      loss_acc += w * l;
    }

    if (emit != nullptr) {
      // sorted update (fm_sorted.cu): no scatter here.  The example leaves one record [pg, s_1 .. s_K] behind --
      // everything the gradient of its rows needs besides the rows themselves and the occurrence's value
      if (lane_on && r == 0) {
        float* rec = emit + i * emit_ld;
#pragma unroll
        for (int t = 0; t < T; ++t) {
          const int chunk = c + t * gs;
          if (chunk < vc_table) {
            float4 o = S[t];
            if (chunk == 0) o.x = pg;            // column 0 (the bias column, whose total is unused) carries pg
            *reinterpret_cast<float4*>(rec + 4 * chunk) = o;
          }
        }
      }
      continue;
    }
    // ---- pass B: scatter {C, B_k, D, cnt} to the accumulator rows ---------------------
    for (int base = start; base < end; base += 32) {
      const int j = base + lane;
      unsigned long long my_arow = (unsigned long long)gacc;
      if (j < end) {
        if (PEER) {
          int owner; int64_t row;
          split_id(ids[j], peers.n, &owner, &row);
          my_arow = (unsigned long long)(peers.gtab[owner] + row * gacc_ld);
          red_or_sys(peers.touched[owner] + (row >> 5), 1u << (row & 31));
        } else {
          my_arow = (unsigned long long)(gacc + (local ? (int64_t)local[j] : ids[j]) * gacc_ld);
        }
      }
      const float my_x = j < end ? vals[j] : 0.f;
      const int rows_here = min(32, end - base);
#pragma unroll 4
      for (int s = 0; s * RPW < rows_here; ++s) {
        const int jj = s * RPW + r;
        float* arow = reinterpret_cast<float*>(__shfl_sync(kFull, my_arow, jj & 31));
        const float x = __shfl_sync(kFull, my_x, jj & 31);
        if (lane_on && jj < rows_here) {
          const float pgx = pg * x;
#pragma unroll
          for (int t = 0; t < T; ++t) {
            const int chunk = c + t * gs;
            if (chunk < vc_acc) {
              const int col = 4 * chunk;
              float o[4] = {pgx * S[t].x, pgx * S[t].y, pgx * S[t].z, pgx * S[t].w};
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const int cc = col + e;
                if (cc == 0) o[e] = pgx;                 // C
                else if (cc == K + 1) o[e] = pgx * x;    // D
                else if (cc == K + 2) o[e] = 1.0f;       // cnt
                else if (cc > K + 2) o[e] = 0.f;
              }
              if (col <= K + 2) {
                if (PEER) red_add_v4_sys(arow + col, o[0], o[1], o[2], o[3]);
                else red_add_v4(arow + col, o[0], o[1], o[2], o[3]);
              }
            }
          }
        }
      }
    }
  }

  if (lane == 0) { s_loss[warp] = loss_acc; s_reg[warp] = reg_acc; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0.f, b = 0.f;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) { a += s_loss[w]; b += s_reg[w]; }
    partials[blockIdx.x] = a;
    partials[gridDim.x + blockIdx.x] = b;
  }
  if (PEER) __threadfence_system();  // remote atomics are performed before the kernel is seen as done
}

// two fixed-order sums (loss, reg) of `n` per-block partials each
__global__ void __launch_bounds__(256)
reduce2_kernel(const float* __restrict__ partials, int n, float scale0, float* __restrict__ out0,
               float* __restrict__ out1) {
  __shared__ double s_w[2][8];
  double a = 0.0, b = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) { a += (double)partials[i]; b += (double)partials[n + i]; }
  a = warp_sum(a); b = warp_sum(b);
  if ((threadIdx.x & 31) == 0) { s_w[0][threadIdx.x >> 5] = a; s_w[1][threadIdx.x >> 5] = b; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double ta = 0.0, tb = 0.0;
    for (int w = 0; w < 8; ++w) { ta += s_w[0][w]; tb += s_w[1][w]; }
    *out0 = (float)(ta * (double)scale0);
    *out1 = (float)tb;
  }
}

template <int LPR, int T>
__global__ void __launch_bounds__(256)
fused_adagrad_kernel(float* __restrict__ table, float* __restrict__ accum, int64_t ld,
                     const int64_t* __restrict__ ori_ids, int64_t n_max, const int64_t* __restrict__ n_dev, int K,
                     float* __restrict__ gacc, int64_t gacc_ld, float lr, float rf, float rb,
                     const unsigned long long* __restrict__ guard) {
  if (guard && *guard != kNoError) return;
  const int64_t n = n_dev ? min(*n_dev, n_max) : n_max;
  const int c = threadIdx.x % LPR;
  const int64_t groups_per_block = blockDim.x / LPR;
  const int vc_table = (K + 4) >> 2, vc_acc = (K + 6) >> 2;  // float4 chunks of a K+1 / K+3 wide row
  const int64_t n_round = (n + groups_per_block - 1) / groups_per_block * groups_per_block;  // keep warps converged
  for (int64_t u = (int64_t)blockIdx.x * groups_per_block + threadIdx.x / LPR; u < n_round;
       u += (int64_t)gridDim.x * groups_per_block) {
    const bool live = u < n;
    const int64_t row = live ? ori_ids[u] * ld : 0;
    float4 a[T], w[T], h[T];
    float D = 0.f, cnt = 0.f;
#pragma unroll
    for (int t = 0; t < T; ++t) {
      const int chunk = c + t * LPR;
      a[t] = make_float4(0.f, 0.f, 0.f, 0.f);
      w[t] = a[t]; h[t] = a[t];
      if (live && chunk < vc_acc) a[t] = *reinterpret_cast<const float4*>(gacc + u * gacc_ld + 4 * chunk);
      if (live && chunk < vc_table) {
        w[t] = *reinterpret_cast<const float4*>(table + row + 4 * chunk);
        h[t] = *reinterpret_cast<const float4*>(accum + row + 4 * chunk);
      }
      const float av[4] = {a[t].x, a[t].y, a[t].z, a[t].w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        if (4 * chunk + e == K + 1) D += av[e];
        if (4 * chunk + e == K + 2) cnt += av[e];
      }
    }
    D = sum_in_group<LPR>(D);
    cnt = sum_in_group<LPR>(cnt);
    const float fac_scale = cnt * rf - D, bias_scale = cnt * rb;
#pragma unroll
    for (int t = 0; t < T; ++t) {
      const int chunk = c + t * LPR;
      if (live && chunk < vc_table) {
        const float av[4] = {a[t].x, a[t].y, a[t].z, a[t].w};
        float wv[4] = {w[t].x, w[t].y, w[t].z, w[t].w};
        float hv[4] = {h[t].x, h[t].y, h[t].z, h[t].w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int col = 4 * chunk + e;
          if (col > K) continue;   // pad columns stay as they are (an accumulator pad of 0 would give 0 / sqrt(0))
          const float g = col == 0 ? fmaf(bias_scale, wv[e], av[e]) : fmaf(wv[e], fac_scale, av[e]);
          hv[e] = fmaf(g, g, hv[e]);
          wv[e] -= __fdiv_rn(lr * g, __fsqrt_rn(hv[e]));
        }
        *reinterpret_cast<float4*>(table + row + 4 * chunk) = make_float4(wv[0], wv[1], wv[2], wv[3]);
        *reinterpret_cast<float4*>(accum + row + 4 * chunk) = make_float4(hv[0], hv[1], hv[2], hv[3]);
      }
      if (live && chunk < vc_acc)
        *reinterpret_cast<float4*>(gacc + u * gacc_ld + 4 * chunk) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
}

// Adagrad over accumulator rows that live beside the table rows (gtab[row] belongs to
// table[row]): one lane group per OCCURRENCE.  The first group to swap the row's cnt
// column to zero owns the row and applies the update; later occurrences of the same
// row see cnt == 0 and leave.  No de-duplication pass is needed.
// A group is exactly `gs` lanes = the float4 chunks of a K+3 wide row (gs <= 32; wider
// rows take T chunks per lane), so a warp works on 32 / gs rows at once with no idle
// lanes inside a group.
template <int T>
__global__ void __launch_bounds__(256)
fused_adagrad_claim_kernel(float* __restrict__ table, float* __restrict__ accum, int64_t ld,
                           const int64_t* __restrict__ rows, int64_t n_max, const int64_t* __restrict__ n_dev, int K,
                           float* __restrict__ gtab, int64_t gtab_ld, float lr, float rf, float rb,
                           const unsigned long long* __restrict__ guard, int discard_on_guard, int gs) {
  const bool tripped = guard && *guard != kNoError;
  if (tripped && !discard_on_guard) return;
  const int64_t n = n_dev ? min(*n_dev, n_max) : n_max;
  const int lane = threadIdx.x & 31;
  const int rows_per_warp = 32 / gs;
  const int g = lane / gs, c = lane - g * gs;
  const bool lane_on = g < rows_per_warp;
  const int vc_table = (K + 4) >> 2, vc_acc = (K + 6) >> 2;
  const int cnt_chunk = (K + 2) >> 2, cnt_comp = (K + 2) & 3, d_chunk = (K + 1) >> 2, d_comp = (K + 1) & 3;
  const int64_t warps_total = (int64_t)gridDim.x * (blockDim.x >> 5);
  const int64_t warp_id = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  for (int64_t base = warp_id * rows_per_warp; base < n; base += warps_total * rows_per_warp) {  // warp-uniform
    const int64_t j = base + g;
    const bool live = lane_on && j < n;
    const int64_t id = live ? rows[j] : 0;
    const int64_t row = id * ld, grow = id * gtab_ld;
    float4 a[T], w[T], h[T];
#pragma unroll
    for (int t = 0; t < T; ++t) {
      const int chunk = c + t * gs;
      a[t] = make_float4(0.f, 0.f, 0.f, 0.f);
      w[t] = a[t]; h[t] = a[t];
      if (live && chunk < vc_acc) a[t] = *reinterpret_cast<const float4*>(gtab + grow + 4 * chunk);
      if (live && chunk < vc_table) {
        w[t] = *reinterpret_cast<const float4*>(table + row + 4 * chunk);
        h[t] = *reinterpret_cast<const float4*>(accum + row + 4 * chunk);
      }
    }
    // The claim.  It waits for the row's own load (so the atomic hits a line that is already in
    // L2 -- an atomic that misses L2 costs a DRAM round trip inside the atomic unit) and is
    // skipped when the row was visibly consumed already.
    float cnt = 0.f, D = 0.f;
#pragma unroll
    for (int t = 0; t < T; ++t) {
      const int chunk = c + t * gs;
      const float av[4] = {a[t].x, a[t].y, a[t].z, a[t].w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        if (chunk == cnt_chunk && e == cnt_comp && live && av[e] != 0.f)
          cnt = atomicExch(gtab + grow + 4 * cnt_chunk + cnt_comp, 0.f);
        if (chunk == d_chunk && e == d_comp) D = av[e];
      }
    }
    // cnt / D sit in one lane of the group: hand them to the others
    cnt = __shfl_sync(kFull, cnt, g * gs + cnt_chunk % gs);
    D = __shfl_sync(kFull, D, g * gs + d_chunk % gs);
    if (!(live && cnt != 0.f)) continue;  // another occurrence of this row got there first
    const float fac_scale = cnt * rf - D, bias_scale = cnt * rb;
#pragma unroll
    for (int t = 0; t < T; ++t) {
      const int chunk = c + t * gs;
      if (chunk < vc_table && !tripped) {  // tripped: the step is void, only throw the sums away
        const float av[4] = {a[t].x, a[t].y, a[t].z, a[t].w};
        float wv[4] = {w[t].x, w[t].y, w[t].z, w[t].w};
        float hv[4] = {h[t].x, h[t].y, h[t].z, h[t].w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int col = 4 * chunk + e;
          if (col > K) continue;   // pad columns stay as they are (an accumulator pad of 0 would give 0 * inf)
          const float gr = col == 0 ? fmaf(bias_scale, wv[e], av[e]) : fmaf(wv[e], fac_scale, av[e]);
          hv[e] = fmaf(gr, gr, hv[e]);
          wv[e] = fmaf(-lr * gr, rsqrtf(hv[e]), wv[e]);   // w -= lr g / sqrt(acc)  (TF: g * lr * rsqrt(acc))
        }
        *reinterpret_cast<float4*>(table + row + 4 * chunk) = make_float4(wv[0], wv[1], wv[2], wv[3]);
        *reinterpret_cast<float4*>(accum + row + 4 * chunk) = make_float4(hv[0], hv[1], hv[2], hv[3]);
      }
      if (chunk < vc_acc && 4 * chunk <= K + 2)
        *reinterpret_cast<float4*>(gtab + grow + 4 * chunk) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
}

// The same update steered by the bitmap of touched rows (bit r set <=> row r got a contribution this step: the sharded
// scorer sets it, fused_score_scatter_kernel<.., PEER> with any number of ranks).  A warp takes 32 consecutive words =
// 1024 rows, clears them, lists the set bits in its shared-memory slice and its lane groups take the listed rows: every
// touched row exactly once, in ascending order, with no claim atomics, no separate scan, no row list in HBM.
constexpr int kBitmapWarps = 8;
template <int T>
__global__ void __launch_bounds__(kBitmapWarps * 32)
fused_adagrad_bitmap_kernel(float* __restrict__ table, float* __restrict__ accum, int64_t ld,
                            unsigned* __restrict__ bits, int64_t n_words, int K, float* __restrict__ gtab,
                            int64_t gtab_ld, float lr, float rf, float rb,
                            const unsigned long long* __restrict__ guard, int discard_on_guard, int gs,
                            unsigned long long* __restrict__ n_rows_out) {
  __shared__ unsigned short s_list[kBitmapWarps][1024];
  const bool tripped = guard && *guard != kNoError;
  if (tripped && !discard_on_guard) return;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned short* list = s_list[warp];
  const int rows_per_warp = 32 / gs;
  const int g = lane / gs, c = lane - g * gs;
  const bool lane_on = g < rows_per_warp;
  const int vc_table = (K + 4) >> 2, vc_acc = (K + 6) >> 2;
  const int cnt_chunk = (K + 2) >> 2, cnt_comp = (K + 2) & 3, d_chunk = (K + 1) >> 2, d_comp = (K + 1) & 3;
  const int64_t warps_total = (int64_t)gridDim.x * kBitmapWarps;
  const int64_t n_spans = (n_words + 31) >> 5;
  // (a block per span instead of a warp per span, tried: 0.28-0.31 ms against 0.266 on uniform ids, 0.055 against 0.031
  // on skewed ones -- two block barriers per span and one warp listing the bits while seven wait)
  for (int64_t span = (int64_t)blockIdx.x * kBitmapWarps + warp; span < n_spans; span += warps_total) {  // warp-uniform
    const int64_t widx = span * 32 + lane;
    unsigned m = widx < n_words ? bits[widx] : 0u;
    if (!__any_sync(kFull, m != 0u)) continue;
    if (m) bits[widx] = 0u;
    int incl = __popc(m);
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int u = __shfl_up_sync(kFull, incl, o);
      if (lane >= o) incl += u;
    }
    const int total = __shfl_sync(kFull, incl, 31);
    if (n_rows_out != nullptr && lane == 0) atomicAdd(n_rows_out, (unsigned long long)total);
    for (int at = incl - __popc(m); m; m &= m - 1, ++at) list[at] = (unsigned short)(lane * 32 + (__ffs(m) - 1));
    __syncwarp();
    for (int taken = 0; taken < total; taken += rows_per_warp) {          // warp-uniform
      const bool live = lane_on && taken + g < total;
      const int64_t id = span * 1024 + (live ? list[taken + g] : 0);
      const int64_t row = id * ld, grow = id * gtab_ld;
      float4 a[T], w[T], h[T];
#pragma unroll
      for (int t = 0; t < T; ++t) {
        const int chunk = c + t * gs;
        a[t] = make_float4(0.f, 0.f, 0.f, 0.f);
        w[t] = a[t]; h[t] = a[t];
        if (live && chunk < vc_acc) a[t] = *reinterpret_cast<const float4*>(gtab + grow + 4 * chunk);
        if (live && chunk < vc_table && !tripped) {
          w[t] = *reinterpret_cast<const float4*>(table + row + 4 * chunk);
          h[t] = *reinterpret_cast<const float4*>(accum + row + 4 * chunk);
        }
      }
      float cnt = 0.f, D = 0.f;
#pragma unroll
      for (int t = 0; t < T; ++t) {
        const int chunk = c + t * gs;
        const float av[4] = {a[t].x, a[t].y, a[t].z, a[t].w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          if (chunk == cnt_chunk && e == cnt_comp) cnt = av[e];
          if (chunk == d_chunk && e == d_comp) D = av[e];
        }
      }
      cnt = __shfl_sync(kFull, cnt, g * gs + cnt_chunk % gs);
      D = __shfl_sync(kFull, D, g * gs + d_chunk % gs);
      if (!live) continue;
      const float fac_scale = cnt * rf - D, bias_scale = cnt * rb;
#pragma unroll
      for (int t = 0; t < T; ++t) {
        const int chunk = c + t * gs;
        if (chunk < vc_table && !tripped) {  // tripped: the step is void, only throw the sums away
          const float av[4] = {a[t].x, a[t].y, a[t].z, a[t].w};
          float wv[4] = {w[t].x, w[t].y, w[t].z, w[t].w};
          float hv[4] = {h[t].x, h[t].y, h[t].z, h[t].w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int col = 4 * chunk + e;
            if (col > K) continue;
            const float gr = col == 0 ? fmaf(bias_scale, wv[e], av[e]) : fmaf(wv[e], fac_scale, av[e]);
            hv[e] = fmaf(gr, gr, hv[e]);
            wv[e] = fmaf(-lr * gr, rsqrtf(hv[e]), wv[e]);   // as in fused_adagrad_claim_kernel
          }
          *reinterpret_cast<float4*>(table + row + 4 * chunk) = make_float4(wv[0], wv[1], wv[2], wv[3]);
          *reinterpret_cast<float4*>(accum + row + 4 * chunk) = make_float4(hv[0], hv[1], hv[2], hv[3]);
        }
        if (chunk < vc_acc && 4 * chunk <= K + 2)
          *reinterpret_cast<float4*>(gtab + grow + 4 * chunk) = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
    __syncwarp();
  }
}

// ---- rows wider than 96 float4 chunks (factor_num > 378) -------------------------------------
// The same arithmetic with plain loops: one warp per example (scorer) or per row (update), lane
// l owns chunks l, l + 32, ...; an example's column totals sit in shared memory (one float4 per
// chunk per warp) instead of registers.  Correctness first: such widths are rare and every id
// then moves kilobytes, so the walk is bandwidth-bound anyway.
constexpr size_t kWideSmemMax = 200 * 1024;

template <bool PEER>
__global__ void __launch_bounds__(kWarps * 32)
fused_score_scatter_wide_kernel(const int64_t* __restrict__ ids, const int32_t* __restrict__ local,
                                const float* __restrict__ vals, const int32_t* __restrict__ poses,
                                const float* __restrict__ labels, const float* __restrict__ weights, int64_t batch,
                                int K, const float* __restrict__ table, int64_t ld, float lam_f, float lam_b,
                                int loss_kind, float inv_batch, float* __restrict__ pred, float* __restrict__ gacc,
                                int64_t gacc_ld, float* __restrict__ partials,
                                const unsigned long long* __restrict__ guard, const PeerPtrs peers) {
  extern __shared__ float4 s_tot[];   // [warps][vc_acc]
  __shared__ float s_loss[kWarps], s_reg[kWarps];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
  const int vc_table = (K + 4) >> 2, vc_acc = (K + 6) >> 2;
  float4* S = s_tot + (size_t)warp * vc_acc;
  const bool active = !(guard && *guard != kNoError);
  float loss_acc = 0.f, reg_acc = 0.f;

  for (int64_t i = (int64_t)blockIdx.x * warps + warp; active && i < batch; i += (int64_t)gridDim.x * warps) {
    const int start = poses[i], end = poses[i + 1];
    for (int chunk = lane; chunk < vc_acc; chunk += 32) S[chunk] = make_float4(0.f, 0.f, 0.f, 0.f);
    float q = 0.f, bias_dot = 0.f, bias_sq = 0.f, fac_sq = 0.f;
    for (int j = start; j < end; ++j) {
      const float x = vals[j];
      const float* rowp;
      if (PEER) {
        int owner; int64_t row;
        split_id(ids[j], peers.n, &owner, &row);
        rowp = peers.table[owner] + row * ld;
      } else {
        rowp = table + ids[j] * ld;
      }
      for (int chunk = lane; chunk < vc_table; chunk += 32) {
        float4 v = ldg4(rowp + 4 * chunk);
        if (chunk == 0) {
          bias_dot = fmaf(x, v.x, bias_dot);
          bias_sq = fmaf(v.x, v.x, bias_sq);
          v.x = 0.f;
        }
        const float a0 = x * v.x, a1 = x * v.y, a2 = x * v.z, a3 = x * v.w;
        float4 t = S[chunk];
        t.x += a0; t.y += a1; t.z += a2; t.w += a3;
        S[chunk] = t;
        q = fmaf(a0, a0, fmaf(a1, a1, fmaf(a2, a2, fmaf(a3, a3, q))));
        fac_sq = fmaf(v.x, v.x, fmaf(v.y, v.y, fmaf(v.z, v.z, fmaf(v.w, v.w, fac_sq))));
      }
    }
    float second = 0.f;
    for (int chunk = lane; chunk < vc_table; chunk += 32) {
      const float4 t = S[chunk];
      second += t.x * t.x + t.y * t.y + t.z * t.z + t.w * t.w;
    }
    const float p = warp_sum(bias_dot + 0.5f * (second - q));
    reg_acc += warp_sum(0.5f * lam_f * fac_sq + 0.5f * lam_b * bias_sq);
    if (labels == nullptr) {
      if (lane == 0) pred[i] = p;
      continue;
    }
    const float y = labels[i], w = weights ? weights[i] : 1.0f;
    float l, d;
    if (loss_kind == FM_LOSS_LOGISTIC) {
      l = fmaxf(p, 0.f) - p * y + log1pf(expf(-fabsf(p)));
      d = 1.0f / (1.0f + expf(-p)) - y;
    } else {
      l = (p - y) * (p - y);
      d = 2.0f * (p - y);
    }
    const float pg = w * d * inv_batch;
    if (lane == 0) {
      pred[i] = p;
      loss_acc += w * l;
    }
    for (int j = start; j < end; ++j) {
      const float x = vals[j], pgx = pg * x;
      float* arow;
      if (PEER) {
        int owner; int64_t row;
        split_id(ids[j], peers.n, &owner, &row);
        arow = peers.gtab[owner] + row * gacc_ld;
        if (lane == 0) red_or_sys(peers.touched[owner] + (row >> 5), 1u << (row & 31));
      } else {
        arow = gacc + (local ? (int64_t)local[j] : ids[j]) * gacc_ld;
      }
      for (int chunk = lane; chunk < vc_acc; chunk += 32) {
        const int col = 4 * chunk;
        if (col > K + 2) continue;
        const float4 t = S[chunk];   // chunks past the table row hold zeros
        float o[4] = {pgx * t.x, pgx * t.y, pgx * t.z, pgx * t.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int cc = col + e;
          if (cc == 0) o[e] = pgx;                 // C
          else if (cc == K + 1) o[e] = pgx * x;    // D
          else if (cc == K + 2) o[e] = 1.0f;       // cnt
          else if (cc > K + 2) o[e] = 0.f;
        }
        if (PEER) red_add_v4_sys(arow + col, o[0], o[1], o[2], o[3]);
        else red_add_v4(arow + col, o[0], o[1], o[2], o[3]);
      }
    }
  }
  if (lane == 0) { s_loss[warp] = loss_acc; s_reg[warp] = reg_acc; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0.f, b = 0.f;
    for (int w = 0; w < warps; ++w) { a += s_loss[w]; b += s_reg[w]; }
    partials[blockIdx.x] = a;
    partials[gridDim.x + blockIdx.x] = b;
  }
  if (PEER) __threadfence_system();
}

// CLAIM: `rows` are occurrences, accumulator row = g + id * g_ld, first to swap cnt to zero applies it;
// otherwise `rows` are distinct ids and accumulator row u belongs to rows[u]
template <bool CLAIM>
__global__ void __launch_bounds__(256)
fused_adagrad_wide_kernel(float* __restrict__ table, float* __restrict__ accum, int64_t ld,
                          const int64_t* __restrict__ rows, int64_t n_max, const int64_t* __restrict__ n_dev, int K,
                          float* __restrict__ g, int64_t g_ld, float lr, float rf, float rb,
                          const unsigned long long* __restrict__ guard, int discard_on_guard) {
  const bool tripped = guard && *guard != kNoError;
  if (tripped && !(CLAIM && discard_on_guard)) return;
  const int64_t n = n_dev ? min(*n_dev, n_max) : n_max;
  const int lane = threadIdx.x & 31;
  const int vc_table = (K + 4) >> 2, vc_acc = (K + 6) >> 2;
  const int64_t warps_total = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (int64_t u = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); u < n; u += warps_total) {
    const int64_t id = rows[u];
    const int64_t row = id * ld, grow = (CLAIM ? id : u) * g_ld;
    float cnt = 0.f, D = 0.f;
    if (lane == 0) {
      cnt = g[grow + K + 2];
      if (CLAIM && cnt != 0.f) cnt = atomicExch(g + grow + K + 2, 0.f);
      D = g[grow + K + 1];
    }
    cnt = __shfl_sync(kFull, cnt, 0);
    D = __shfl_sync(kFull, D, 0);
    if (CLAIM && cnt == 0.f) continue;   // another occurrence of this row got there first
    const float fac_scale = cnt * rf - D, bias_scale = cnt * rb;
    for (int chunk = lane; chunk < vc_acc; chunk += 32) {
      const float4 a = *reinterpret_cast<const float4*>(g + grow + 4 * chunk);
      if (chunk < vc_table && !tripped) {
        const float4 w = *reinterpret_cast<const float4*>(table + row + 4 * chunk);
        const float4 h = *reinterpret_cast<const float4*>(accum + row + 4 * chunk);
        const float av[4] = {a.x, a.y, a.z, a.w};
        float wv[4] = {w.x, w.y, w.z, w.w};
        float hv[4] = {h.x, h.y, h.z, h.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int col = 4 * chunk + e;
          if (col > K) continue;
          const float gr = col == 0 ? fmaf(bias_scale, wv[e], av[e]) : fmaf(wv[e], fac_scale, av[e]);
          hv[e] = fmaf(gr, gr, hv[e]);
          wv[e] = fmaf(-lr * gr, rsqrtf(hv[e]), wv[e]);
        }
        *reinterpret_cast<float4*>(table + row + 4 * chunk) = make_float4(wv[0], wv[1], wv[2], wv[3]);
        *reinterpret_cast<float4*>(accum + row + 4 * chunk) = make_float4(hv[0], hv[1], hv[2], hv[3]);
      }
      if (4 * chunk <= K + 2) *reinterpret_cast<float4*>(g + grow + 4 * chunk) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
}

// warps per block such that the column totals fit in shared memory; 0 = the row does not fit at all
static inline int wide_warps(int64_t factor_num, size_t* smem) {
  const size_t per_warp = (size_t)((factor_num + 6) / 4) * sizeof(float4);
  if (per_warp > kWideSmemMax) return 0;
  size_t w = (64 * 1024) / per_warp;
  if (w > (size_t)kWarps) w = kWarps;
  if (w < 1) w = 1;
  *smem = w * per_warp;
  return (int)w;
}

// the wide kernels' dynamic shared memory opt-in, once per device instead of once per call
static cudaError_t wide_smem_opt_in(bool peer) {
  static bool done[2][64];
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev >= 0 && dev < 64 && done[peer][dev]) return cudaSuccess;
  e = peer ? cudaFuncSetAttribute(fused_score_scatter_wide_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)kWideSmemMax)
           : cudaFuncSetAttribute(fused_score_scatter_wide_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)kWideSmemMax);
  if (e == cudaSuccess && dev >= 0 && dev < 64) done[peer][dev] = true;   // benign race: the attribute is idempotent
  return e;
}

// bitmap of touched rows -> ascending list of row numbers (and the bitmap is cleared on the way).
// A tile = 256 threads x kTouchedRounds words, read round by round so that a warp reads 32 consecutive words and its
// lanes write consecutive pieces of the list (the first version gave each thread 8 consecutive words: a serial loop
// over ~36 rows per thread and 8-byte stores 256 bytes apart, 0.043-0.051 ms for a 2 MB bitmap).  Tiles take their
// number from a ticket and find their place in the list by decoupled look-back (fm_scan.cuh).
constexpr int kTouchedRounds = 4;
__global__ void __launch_bounds__(kScanThreads)
touched_rows_kernel(unsigned* __restrict__ bits, int64_t n_words, int64_t* __restrict__ rows, int64_t capacity,
                    int64_t* __restrict__ n_out, ScanScratch st, int64_t n_tiles) {
  constexpr int kW = kScanThreads / 32;
  __shared__ unsigned s_tile;
  __shared__ int s_warp[kTouchedRounds][kW];
  __shared__ long long s_prefix;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_tile = atomicAdd(st.ticket, 1u);
  __syncthreads();
  const int64_t tile = s_tile;
  const int64_t base = tile * (int64_t)(kScanThreads * kTouchedRounds);
  unsigned w[kTouchedRounds];
  int incl[kTouchedRounds];
#pragma unroll
  for (int r = 0; r < kTouchedRounds; ++r) {
    const int64_t idx = base + r * kScanThreads + tid;
    w[r] = idx < n_words ? bits[idx] : 0u;
    int v = __popc(w[r]);
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int u = __shfl_up_sync(kFull, v, o);
      if (lane >= o) v += u;
    }
    incl[r] = v;
    if (lane == 31) s_warp[r][warp] = v;
  }
  __syncthreads();
  // list order: round-major, then warp, then lane (= ascending words)
  int before[kTouchedRounds], block_sum = 0;
#pragma unroll
  for (int r = 0; r < kTouchedRounds; ++r) {
    before[r] = block_sum;
#pragma unroll
    for (int x = 0; x < kW; ++x) {
      const int v = s_warp[r][x];
      if (x < warp) before[r] += v;
      block_sum += v;
    }
  }
  if (warp == 0) {
    const long long prefix = scan_look_back(st, tile, block_sum);
    if (lane == 0) s_prefix = prefix;
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < kTouchedRounds; ++r) {
    unsigned m = w[r];
    if (m == 0u) continue;
    const int64_t idx = base + r * kScanThreads + tid;
    bits[idx] = 0u;
    for (long long at = s_prefix + before[r] + incl[r] - __popc(m); m; m &= m - 1, ++at)
      if (at < capacity) rows[at] = idx * 32 + (__ffs(m) - 1);
  }
  if (tile == n_tiles - 1 && tid == 0) *n_out = s_prefix + block_sum;
}

struct Shape { int lpr, t; };
static inline bool pick_shape(int64_t chunks, Shape* s) {
  int lpr = 1;
  while (lpr < 32 && lpr < chunks) lpr <<= 1;
  const int t = (int)((chunks + lpr - 1) / lpr);
  if (t > 3) return false;
  s->lpr = lpr; s->t = t;
  return true;
}

#define FM_DISPATCH_SHAPE(shape, CALL)                     \
  switch ((shape).lpr * 4 + (shape).t) {                   \
    case 1 * 4 + 1: { CALL(1, 1); } break;                 \
    case 2 * 4 + 1: { CALL(2, 1); } break;                 \
    case 4 * 4 + 1: { CALL(4, 1); } break;                 \
    case 8 * 4 + 1: { CALL(8, 1); } break;                 \
    case 16 * 4 + 1: { CALL(16, 1); } break;               \
    case 32 * 4 + 1: { CALL(32, 1); } break;               \
    case 32 * 4 + 2: { CALL(32, 2); } break;               \
    case 32 * 4 + 3: { CALL(32, 3); } break;               \
    default: return fm_fail(FM_E_LIMIT, "unsupported row shape"); \
  }

static inline int fused_blocks(int64_t batch) {
  static const int bpsm = fm_env_int("FM_SCORE_BLOCKS_PER_SM", 8);
  int64_t b = (batch + kWarps - 1) / kWarps;
  if (b > 148 * 8) b = 148 * 8;      // the workspace is sized for this many partials
  if (b > 148 * bpsm && bpsm > 0 && bpsm < 8) b = 148 * bpsm;
  return (int)(b < 1 ? 1 : b);
}

}  // namespace
}  // namespace fm

using namespace fm;

FM_EXPORT int64_t fm_fused_gacc_ld(int64_t factor_num) { return (factor_num + 3 + 3) / 4 * 4; }

FM_EXPORT size_t fm_fused_workspace(int64_t batch) {
  return fm_align_up((size_t)fused_blocks(batch) * 2 * sizeof(float), 256);
}

FM_EXPORT int fm_fused_score_scatter(const int64_t* feature_ids, const int32_t* local_ids, const float* feature_vals,
                                     const int32_t* feature_poses, const float* labels, const float* weights,
                                     int64_t batch, int64_t factor_num, const float* table, int64_t table_ld,
                                     float factor_lambda, float bias_lambda, int loss_kind, int64_t loss_batch,
                                     float* pred_score,
                                     float* loss, float* reg_score, float* gacc, int64_t gacc_ld,
                                     const uint64_t* guard, void* workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  FM_REQUIRE(batch >= 0 && factor_num >= 0, "fm_fused_score_scatter: bad shape");
  FM_REQUIRE(table_ld >= factor_num + 1 && table_ld % 4 == 0, "fm_fused_score_scatter: table_ld must be a multiple of 4 and >= K+1");
  FM_REQUIRE(gacc_ld >= factor_num + 3 && gacc_ld % 4 == 0, "fm_fused_score_scatter: gacc_ld must be a multiple of 4 and >= K+3");
  FM_REQUIRE(loss_kind == FM_LOSS_LOGISTIC || loss_kind == FM_LOSS_MSE, "fm_fused_score_scatter: unknown loss kind");
  FM_REQUIRE(loss && reg_score && workspace, "fm_fused_score_scatter: null pointer");
  FM_REQUIRE(batch == 0 || (feature_ids && feature_vals && feature_poses && table && pred_score),
             "fm_fused_score_scatter: null pointer");
  FM_REQUIRE(batch == 0 || (labels == nullptr) == (gacc == nullptr),
             "fm_fused_score_scatter: labels and gacc go together (both NULL = scoring only)");
  FM_REQUIRE((((uintptr_t)table | (uintptr_t)gacc) & 15) == 0, "fm_fused_score_scatter: table and gacc must be 16-byte aligned");
  if (workspace_bytes < fm_fused_workspace(batch))
    return fm_fail(FM_E_WORKSPACE, "fm_fused_score_scatter: workspace %zu < %zu", workspace_bytes, fm_fused_workspace(batch));
  const int64_t chunks = (factor_num + 6) / 4;
  const int gs = chunks <= 32 ? (int)chunks : 32;
  const int t = (int)((chunks + gs - 1) / gs);
  const int blocks = fused_blocks(batch);
  float* partials = reinterpret_cast<float*>(workspace);
  const int64_t denom = loss_batch > 0 ? loss_batch : batch;
  const float inv_batch = denom > 0 ? 1.0f / (float)denom : 0.f;
  if (t > 3) {  // wider than 96 chunks: the generic kernel
    size_t smem = 0;
    const int warps = wide_warps(factor_num, &smem);
    if (warps == 0) return fm_fail(FM_E_LIMIT, "fm_fused_score_scatter: factor_num %lld too large", (long long)factor_num);
    FM_CUDA(wide_smem_opt_in(false));
    fused_score_scatter_wide_kernel<false><<<blocks, warps * 32, smem, stream>>>(
        feature_ids, local_ids, feature_vals, feature_poses, labels, weights, batch, (int)factor_num, table, table_ld,
        factor_lambda, bias_lambda, loss_kind, inv_batch, pred_score, gacc, gacc_ld, partials,
        reinterpret_cast<const unsigned long long*>(guard), PeerPtrs());
    FM_LAUNCH_CHECK();
    reduce2_kernel<<<1, 256, 0, stream>>>(partials, blocks, inv_batch, loss, reg_score);
    FM_LAUNCH_CHECK();
    return FM_OK;
  }
#define CALL(T)                                                                                                   \
  fused_score_scatter_kernel<T, false><<<blocks, kWarps * 32, 0, stream>>>(                                       \
      feature_ids, local_ids, feature_vals, feature_poses, labels, weights, batch, (int)factor_num, table, table_ld, \
      factor_lambda, bias_lambda, loss_kind, inv_batch, pred_score, gacc, gacc_ld, partials,                      \
      reinterpret_cast<const unsigned long long*>(guard), PeerPtrs(), gs, nullptr, 0)
  if (t == 1) { CALL(1); } else if (t == 2) { CALL(2); } else { CALL(3); }
#undef CALL
  FM_LAUNCH_CHECK();
  reduce2_kernel<<<1, 256, 0, stream>>>(partials, blocks, inv_batch, loss, reg_score);
  FM_LAUNCH_CHECK();
  return FM_OK;
}

FM_EXPORT int64_t fm_sorted_record_ld(int64_t factor_num) { return (factor_num + 4) / 4 * 4; }

// the scorer of the sorted path: fm_fused_score_scatter without the scatter; per example it leaves the record
// [pg, s_1 .. s_K] (pg = d loss / d pred over the batch, s_k = sum_j x_j v_jk) for fm_sorted_update
FM_EXPORT int fm_sorted_score(const int64_t* feature_ids, const float* feature_vals, const int32_t* feature_poses,
                              const float* labels, const float* weights, int64_t batch, int64_t factor_num,
                              const float* table, int64_t table_ld, float factor_lambda, float bias_lambda,
                              int loss_kind, int64_t loss_batch, float* pred_score, float* loss, float* reg_score,
                              float* records, int64_t record_ld, const uint64_t* guard, void* workspace,
                              size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  FM_REQUIRE(batch >= 0 && factor_num >= 0, "fm_sorted_score: bad shape");
  FM_REQUIRE(table_ld >= factor_num + 1 && table_ld % 4 == 0, "fm_sorted_score: table_ld must be a multiple of 4 and >= K+1");
  FM_REQUIRE(record_ld >= fm_sorted_record_ld(factor_num) && record_ld % 4 == 0,
             "fm_sorted_score: record_ld must be a multiple of 4 and >= fm_sorted_record_ld(K)");
  FM_REQUIRE(loss_kind == FM_LOSS_LOGISTIC || loss_kind == FM_LOSS_MSE, "fm_sorted_score: unknown loss kind");
  FM_REQUIRE(loss && reg_score && workspace, "fm_sorted_score: null pointer");
  FM_REQUIRE(batch == 0 || (feature_ids && feature_vals && feature_poses && table && pred_score && labels && records),
             "fm_sorted_score: null pointer");
  FM_REQUIRE((((uintptr_t)table | (uintptr_t)records) & 15) == 0, "fm_sorted_score: table and records must be 16-byte aligned");
  if (workspace_bytes < fm_fused_workspace(batch))
    return fm_fail(FM_E_WORKSPACE, "fm_sorted_score: workspace %zu < %zu", workspace_bytes, fm_fused_workspace(batch));
  const int64_t chunks = (factor_num + 4) / 4;      // a K+1 wide row: no K+3 wide accumulator row on this path
  const int gs = chunks <= 32 ? (int)chunks : 32;
  const int t = (int)((chunks + gs - 1) / gs);
  if (t > 3) return fm_fail(FM_E_LIMIT, "fm_sorted_score: factor_num %lld beyond the sorted path (use the claim path)",
                            (long long)factor_num);
  const int blocks = fused_blocks(batch);
  float* partials = reinterpret_cast<float*>(workspace);
  const int64_t denom = loss_batch > 0 ? loss_batch : batch;
  const float inv_batch = denom > 0 ? 1.0f / (float)denom : 0.f;
#define CALL(T)                                                                                                   \
  fused_score_scatter_kernel<T, false><<<blocks, kWarps * 32, 0, stream>>>(                                       \
      feature_ids, nullptr, feature_vals, feature_poses, labels, weights, batch, (int)factor_num, table, table_ld, \
      factor_lambda, bias_lambda, loss_kind, inv_batch, pred_score, nullptr, 0, partials,                         \
      reinterpret_cast<const unsigned long long*>(guard), PeerPtrs(), gs, records, record_ld)
  if (t == 1) { CALL(1); } else if (t == 2) { CALL(2); } else { CALL(3); }
#undef CALL
  FM_LAUNCH_CHECK();
  reduce2_kernel<<<1, 256, 0, stream>>>(partials, blocks, inv_batch, loss, reg_score);
  FM_LAUNCH_CHECK();
  return FM_OK;
}

FM_EXPORT int fm_fused_adagrad(float* table, float* accum, int64_t table_ld, const int64_t* ori_ids, int64_t n_unique,
                               const int64_t* n_unique_dev, int64_t factor_num, float* gacc, int64_t gacc_ld,
                               float learning_rate, float factor_lambda, float bias_lambda, float reg_grad,
                               const uint64_t* guard, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  FM_REQUIRE(n_unique >= 0 && factor_num >= 0, "fm_fused_adagrad: bad shape");
  FM_REQUIRE(table_ld >= factor_num + 1 && table_ld % 4 == 0, "fm_fused_adagrad: table_ld must be a multiple of 4 and >= K+1");
  FM_REQUIRE(gacc_ld >= factor_num + 3 && gacc_ld % 4 == 0, "fm_fused_adagrad: gacc_ld must be a multiple of 4 and >= K+3");
  if (n_unique == 0) return FM_OK;
  FM_REQUIRE(table && accum && ori_ids && gacc, "fm_fused_adagrad: null pointer");
  FM_REQUIRE((((uintptr_t)table | (uintptr_t)accum | (uintptr_t)gacc) & 15) == 0, "fm_fused_adagrad: 16-byte alignment required");
  Shape sh;
  if (!pick_shape((factor_num + 6) / 4, &sh)) {  // wider than 96 chunks: the generic kernel, a warp per row
    int64_t wide_blocks = (n_unique + 7) / 8;
    if (wide_blocks > 148 * 16) wide_blocks = 148 * 16;
    fused_adagrad_wide_kernel<false><<<(unsigned)wide_blocks, 256, 0, stream>>>(
        table, accum, table_ld, ori_ids, n_unique, n_unique_dev, (int)factor_num, gacc, gacc_ld, learning_rate,
        reg_grad * factor_lambda, reg_grad * bias_lambda, reinterpret_cast<const unsigned long long*>(guard), 0);
    FM_LAUNCH_CHECK();
    return FM_OK;
  }
  const int64_t groups_per_block = 256 / sh.lpr;
  int64_t blocks = (n_unique + groups_per_block - 1) / groups_per_block;
  if (blocks > 148 * 16) blocks = 148 * 16;
#define CALL(LPR, T)                                                                                     \
  fused_adagrad_kernel<LPR, T><<<(unsigned)blocks, 256, 0, stream>>>(                                    \
      table, accum, table_ld, ori_ids, n_unique, n_unique_dev, (int)factor_num, gacc, gacc_ld,           \
      learning_rate, reg_grad * factor_lambda, reg_grad * bias_lambda,                                   \
      reinterpret_cast<const unsigned long long*>(guard))
  FM_DISPATCH_SHAPE(sh, CALL)
#undef CALL
  FM_LAUNCH_CHECK();
  return FM_OK;
}

FM_EXPORT int fm_fused_adagrad_claim(float* table, float* accum, int64_t table_ld, const int64_t* rows, int64_t n,
                                     const int64_t* n_dev, int64_t factor_num, float* gtab, int64_t gtab_ld,
                                     float learning_rate, float factor_lambda, float bias_lambda, float reg_grad,
                                     const uint64_t* guard, int discard_on_guard, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  FM_REQUIRE(n >= 0 && factor_num >= 0, "fm_fused_adagrad_claim: bad shape");
  FM_REQUIRE(table_ld >= factor_num + 1 && table_ld % 4 == 0, "fm_fused_adagrad_claim: table_ld must be a multiple of 4 and >= K+1");
  FM_REQUIRE(gtab_ld >= factor_num + 3 && gtab_ld % 4 == 0, "fm_fused_adagrad_claim: gtab_ld must be a multiple of 4 and >= K+3");
  if (n == 0) return FM_OK;
  FM_REQUIRE(table && accum && rows && gtab, "fm_fused_adagrad_claim: null pointer");
  FM_REQUIRE((((uintptr_t)table | (uintptr_t)accum | (uintptr_t)gtab) & 15) == 0, "fm_fused_adagrad_claim: 16-byte alignment required");
  const int64_t chunks = (factor_num + 6) / 4;
  const int gs = chunks <= 32 ? (int)chunks : 32;
  const int t = (int)((chunks + gs - 1) / gs);
  if (t > 3) {  // wider than 96 chunks: the generic kernel, a warp per occurrence
    int64_t wide_blocks = (n + 7) / 8;
    if (wide_blocks > 148 * 16) wide_blocks = 148 * 16;
    fused_adagrad_wide_kernel<true><<<(unsigned)wide_blocks, 256, 0, stream>>>(
        table, accum, table_ld, rows, n, n_dev, (int)factor_num, gtab, gtab_ld, learning_rate,
        reg_grad * factor_lambda, reg_grad * bias_lambda, reinterpret_cast<const unsigned long long*>(guard),
        discard_on_guard);
    FM_LAUNCH_CHECK();
    return FM_OK;
  }
  const int64_t rows_per_block = 8 * (32 / gs);
  int64_t blocks = (n + rows_per_block - 1) / rows_per_block;
  // grid: many short blocks, not a persistent one (profiles/r2_sweep_claim_blocks.txt: 148 x 6 blocks 0.31 ms, 148 x 16
  // 0.256, 148 x 32 the same alone and 2 % better with the parse of the next batch running beside it)
  static const int bpsm = fm_env_int("FM_CLAIM_BLOCKS_PER_SM", 32);
  if (blocks > 148 * bpsm) blocks = 148 * bpsm;
#define CALL(T)                                                                                           \
  fused_adagrad_claim_kernel<T><<<(unsigned)blocks, 256, 0, stream>>>(                                    \
      table, accum, table_ld, rows, n, n_dev, (int)factor_num, gtab, gtab_ld, learning_rate,              \
      reg_grad * factor_lambda, reg_grad * bias_lambda, reinterpret_cast<const unsigned long long*>(guard), \
      discard_on_guard, gs)
  if (t == 1) { CALL(1); } else if (t == 2) { CALL(2); } else { CALL(3); }
#undef CALL
  FM_LAUNCH_CHECK();
  return FM_OK;
}

FM_EXPORT int fm_fused_adagrad_bitmap(float* table, float* accum, int64_t table_ld, uint32_t* touched, int64_t n_rows,
                                      int64_t factor_num, float* gtab, int64_t gtab_ld, float learning_rate,
                                      float factor_lambda, float bias_lambda, float reg_grad, const uint64_t* guard,
                                      int discard_on_guard, int64_t* rows_seen, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  FM_REQUIRE(n_rows >= 0 && factor_num >= 0, "fm_fused_adagrad_bitmap: bad shape");
  FM_REQUIRE(table_ld >= factor_num + 1 && table_ld % 4 == 0, "fm_fused_adagrad_bitmap: table_ld must be a multiple of 4 and >= K+1");
  FM_REQUIRE(gtab_ld >= factor_num + 3 && gtab_ld % 4 == 0, "fm_fused_adagrad_bitmap: gtab_ld must be a multiple of 4 and >= K+3");
  if (n_rows == 0) return FM_OK;
  FM_REQUIRE(table && accum && touched && gtab, "fm_fused_adagrad_bitmap: null pointer");
  FM_REQUIRE((((uintptr_t)table | (uintptr_t)accum | (uintptr_t)gtab) & 15) == 0, "fm_fused_adagrad_bitmap: 16-byte alignment required");
  const int64_t chunks = (factor_num + 6) / 4;
  const int gs = chunks <= 32 ? (int)chunks : 32;
  const int t = (int)((chunks + gs - 1) / gs);
  if (t > 3) return fm_fail(FM_E_LIMIT, "fm_fused_adagrad_bitmap: factor_num %lld beyond this form (use fm_peer_touched_rows + "
                            "fm_fused_adagrad_claim)", (long long)factor_num);
  const int64_t words = (n_rows + 31) / 32, spans = (words + 31) / 32;
  int64_t blocks = (spans + kBitmapWarps - 1) / kBitmapWarps;
  static const int bpsm = fm_env_int("FM_BITMAP_BLOCKS_PER_SM", 32);
  if (blocks > 148 * (int64_t)bpsm) blocks = 148 * (int64_t)bpsm;
#define CALL(T)                                                                                             \
  fused_adagrad_bitmap_kernel<T><<<(unsigned)blocks, kBitmapWarps * 32, 0, stream>>>(                       \
      table, accum, table_ld, touched, words, (int)factor_num, gtab, gtab_ld, learning_rate,                \
      reg_grad * factor_lambda, reg_grad * bias_lambda, reinterpret_cast<const unsigned long long*>(guard), \
      discard_on_guard, gs, reinterpret_cast<unsigned long long*>(rows_seen))
  if (t == 1) { CALL(1); } else if (t == 2) { CALL(2); } else { CALL(3); }
#undef CALL
  FM_LAUNCH_CHECK();
  return FM_OK;
}

// ---- the same pair over a table sharded across the GPUs of one NVLink domain ----
FM_EXPORT int fm_peer_score_scatter(const int64_t* feature_ids, const float* feature_vals, const int32_t* feature_poses,
                                    const float* labels, const float* weights, int64_t batch, int64_t factor_num,
                                    int64_t n_ranks, const void* const* table_bases, int64_t table_ld,
                                    void* const* gtab_bases, int64_t gtab_ld, void* const* touched_bases,
                                    float factor_lambda, float bias_lambda, int loss_kind, int64_t loss_batch,
                                    float* pred_score, float* loss, float* reg_score, const uint64_t* guard,
                                    void* workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  FM_REQUIRE(batch >= 0 && factor_num >= 0, "fm_peer_score_scatter: bad shape");
  if (n_ranks < 1 || n_ranks > kMaxPeers) return fm_fail(FM_E_LIMIT, "fm_peer_score_scatter: 1..%d ranks", kMaxPeers);
  FM_REQUIRE(table_ld >= factor_num + 1 && table_ld % 4 == 0, "fm_peer_score_scatter: table_ld must be a multiple of 4 and >= K+1");
  FM_REQUIRE(loss_kind == FM_LOSS_LOGISTIC || loss_kind == FM_LOSS_MSE, "fm_peer_score_scatter: unknown loss kind");
  FM_REQUIRE(table_bases && loss && reg_score && workspace, "fm_peer_score_scatter: null pointer");
  FM_REQUIRE(batch == 0 || (feature_ids && feature_vals && feature_poses && pred_score), "fm_peer_score_scatter: null pointer");
  const bool training = labels != nullptr;
  FM_REQUIRE(!training || (gtab_bases && touched_bases && gtab_ld >= factor_num + 3 && gtab_ld % 4 == 0),
             "fm_peer_score_scatter: training needs gtab_bases, touched_bases and gtab_ld >= K+3 (multiple of 4)");
  if (workspace_bytes < fm_fused_workspace(batch))
    return fm_fail(FM_E_WORKSPACE, "fm_peer_score_scatter: workspace %zu < %zu", workspace_bytes, fm_fused_workspace(batch));
  PeerPtrs peers;
  peers.n = (int)n_ranks;
  for (int r = 0; r < kMaxPeers; ++r) {
    const bool in = r < n_ranks;
    peers.table[r] = in ? static_cast<const float*>(table_bases[r]) : nullptr;
    peers.gtab[r] = in && training ? static_cast<float*>(gtab_bases[r]) : nullptr;
    peers.touched[r] = in && training ? static_cast<unsigned*>(touched_bases[r]) : nullptr;
    if (in) {
      FM_REQUIRE(peers.table[r] && ((uintptr_t)peers.table[r] & 15) == 0, "fm_peer_score_scatter: bad table base");
      FM_REQUIRE(!training || (peers.gtab[r] && peers.touched[r] && ((uintptr_t)peers.gtab[r] & 15) == 0),
                 "fm_peer_score_scatter: bad gtab / touched base");
    }
  }
  const int64_t chunks = (factor_num + 6) / 4;
  const int gs = chunks <= 32 ? (int)chunks : 32;
  const int t = (int)((chunks + gs - 1) / gs);
  const int blocks = fused_blocks(batch);
  float* partials = reinterpret_cast<float*>(workspace);
  const int64_t denom = loss_batch > 0 ? loss_batch : batch;
  const float inv_batch = denom > 0 ? 1.0f / (float)denom : 0.f;
  // gacc != NULL tells the kernel to scatter; the real targets come from `peers`
  float* scatter_flag = training ? peers.gtab[0] : nullptr;
  if (t > 3) {  // wider than 96 chunks: the generic kernel
    size_t smem = 0;
    const int warps = wide_warps(factor_num, &smem);
    if (warps == 0) return fm_fail(FM_E_LIMIT, "fm_peer_score_scatter: factor_num %lld too large", (long long)factor_num);
    FM_CUDA(wide_smem_opt_in(true));
    fused_score_scatter_wide_kernel<true><<<blocks, warps * 32, smem, stream>>>(
        feature_ids, nullptr, feature_vals, feature_poses, labels, weights, batch, (int)factor_num, peers.table[0],
        table_ld, factor_lambda, bias_lambda, loss_kind, inv_batch, pred_score, scatter_flag, gtab_ld, partials,
        reinterpret_cast<const unsigned long long*>(guard), peers);
    FM_LAUNCH_CHECK();
    reduce2_kernel<<<1, 256, 0, stream>>>(partials, blocks, inv_batch, loss, reg_score);
    FM_LAUNCH_CHECK();
    return FM_OK;
  }
#define CALL(T)                                                                                                     \
  fused_score_scatter_kernel<T, true><<<blocks, kWarps * 32, 0, stream>>>(                                          \
      feature_ids, nullptr, feature_vals, feature_poses, labels, weights, batch, (int)factor_num, peers.table[0],   \
      table_ld, factor_lambda, bias_lambda, loss_kind, inv_batch, pred_score, scatter_flag, gtab_ld, partials,      \
      reinterpret_cast<const unsigned long long*>(guard), peers, gs, nullptr, 0)
  if (t == 1) { CALL(1); } else if (t == 2) { CALL(2); } else { CALL(3); }
#undef CALL
  FM_LAUNCH_CHECK();
  reduce2_kernel<<<1, 256, 0, stream>>>(partials, blocks, inv_batch, loss, reg_score);
  FM_LAUNCH_CHECK();
  return FM_OK;
}

// the scorer of the record exchange (fm_merge.cu): rows pulled from their owners like fm_peer_score_scatter, nothing
// pushed; per example the record [pg, s_1 .. s_K] stays in local memory for fm_peer_push_records
FM_EXPORT int fm_peer_sorted_score(const int64_t* feature_ids, const float* feature_vals, const int32_t* feature_poses,
                                   const float* labels, const float* weights, int64_t batch, int64_t factor_num,
                                   int64_t n_ranks, const void* const* table_bases, int64_t table_ld,
                                   float factor_lambda, float bias_lambda, int loss_kind, int64_t loss_batch,
                                   float* pred_score, float* loss, float* reg_score, float* records, int64_t record_ld,
                                   const uint64_t* guard, void* workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  FM_REQUIRE(batch >= 0 && factor_num >= 0, "fm_peer_sorted_score: bad shape");
  if (n_ranks < 1 || n_ranks > kMaxPeers) return fm_fail(FM_E_LIMIT, "fm_peer_sorted_score: 1..%d ranks", kMaxPeers);
  FM_REQUIRE(table_ld >= factor_num + 1 && table_ld % 4 == 0, "fm_peer_sorted_score: table_ld must be a multiple of 4 and >= K+1");
  FM_REQUIRE(record_ld >= fm_sorted_record_ld(factor_num) && record_ld % 4 == 0, "fm_peer_sorted_score: bad record_ld");
  FM_REQUIRE(loss_kind == FM_LOSS_LOGISTIC || loss_kind == FM_LOSS_MSE, "fm_peer_sorted_score: unknown loss kind");
  FM_REQUIRE(table_bases && loss && reg_score && workspace, "fm_peer_sorted_score: null pointer");
  FM_REQUIRE(batch == 0 || (feature_ids && feature_vals && feature_poses && pred_score && labels && records),
             "fm_peer_sorted_score: null pointer");
  FM_REQUIRE(((uintptr_t)records & 15) == 0, "fm_peer_sorted_score: records must be 16-byte aligned");
  if (workspace_bytes < fm_fused_workspace(batch))
    return fm_fail(FM_E_WORKSPACE, "fm_peer_sorted_score: workspace %zu < %zu", workspace_bytes, fm_fused_workspace(batch));
  PeerPtrs peers;
  peers.n = (int)n_ranks;
  for (int r = 0; r < kMaxPeers; ++r) {
    peers.table[r] = r < n_ranks ? static_cast<const float*>(table_bases[r]) : nullptr;
    peers.gtab[r] = nullptr;
    peers.touched[r] = nullptr;
    FM_REQUIRE(r >= n_ranks || (peers.table[r] && ((uintptr_t)peers.table[r] & 15) == 0), "fm_peer_sorted_score: bad table base");
  }
  const int64_t chunks = (factor_num + 4) / 4;
  const int gs = chunks <= 32 ? (int)chunks : 32;
  const int t = (int)((chunks + gs - 1) / gs);
  if (t > 3) return fm_fail(FM_E_LIMIT, "fm_peer_sorted_score: factor_num %lld too wide", (long long)factor_num);
  const int blocks = fused_blocks(batch);
  float* partials = reinterpret_cast<float*>(workspace);
  const int64_t denom = loss_batch > 0 ? loss_batch : batch;
  const float inv_batch = denom > 0 ? 1.0f / (float)denom : 0.f;
#define CALL(T)                                                                                                     \
  fused_score_scatter_kernel<T, true><<<blocks, kWarps * 32, 0, stream>>>(                                          \
      feature_ids, nullptr, feature_vals, feature_poses, labels, weights, batch, (int)factor_num, peers.table[0],   \
      table_ld, factor_lambda, bias_lambda, loss_kind, inv_batch, pred_score, nullptr, 0, partials,                 \
      reinterpret_cast<const unsigned long long*>(guard), peers, gs, records, record_ld)
  if (t == 1) { CALL(1); } else if (t == 2) { CALL(2); } else { CALL(3); }
#undef CALL
  FM_LAUNCH_CHECK();
  reduce2_kernel<<<1, 256, 0, stream>>>(partials, blocks, inv_batch, loss, reg_score);
  FM_LAUNCH_CHECK();
  return FM_OK;
}

static inline int64_t touched_tiles(int64_t words) {
  const int64_t per = (int64_t)kScanThreads * kTouchedRounds;
  return words <= 0 ? 1 : (words + per - 1) / per;
}

FM_EXPORT size_t fm_peer_touched_workspace(int64_t shard_rows) {
  return fm_align_up((size_t)touched_tiles((shard_rows + 31) / 32) * 8 + 16, 256);
}

FM_EXPORT int fm_peer_touched_rows(uint32_t* touched, int64_t shard_rows, int64_t* rows, int64_t capacity,
                                   int64_t* n_rows, void* workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  FM_REQUIRE(shard_rows >= 0 && capacity >= 0, "fm_peer_touched_rows: bad size");
  FM_REQUIRE(touched && n_rows && workspace && (capacity == 0 || rows), "fm_peer_touched_rows: null pointer");
  const int64_t words = (shard_rows + 31) / 32;
  const size_t need = fm_peer_touched_workspace(shard_rows);
  if (workspace_bytes < need)
    return fm_fail(FM_E_WORKSPACE, "fm_peer_touched_rows: workspace %zu < %zu", workspace_bytes, need);
  FM_CUDA(cudaMemsetAsync(workspace, 0, need, stream));
  const int64_t tiles = touched_tiles(words);
  touched_rows_kernel<<<(unsigned)tiles, kScanThreads, 0, stream>>>(touched, words, rows, capacity, n_rows,
                                                                    scan_scratch_at(workspace, words, 1), tiles);
  FM_LAUNCH_CHECK();
  return FM_OK;
}
