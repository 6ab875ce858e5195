// fm_sort.cu -- the occurrences of a batch grouped by table row, ahead of the update.
//
// The reference scatters FmGrad's per-occurrence gradients with a CAS-loop float add
// (cc/fm_grad_op.h:3-16, :80-106) and lets TF's unique + SparseApplyAdagrad find the rows
// (tffm/fm_model.py:72, :341-348).  Here the occurrences are SORTED by feature id instead --
// a stable LSD radix sort that needs nothing but the parser's ids, so it can run on a side
// stream next to the parse -- and the update (fm_sorted.cu) walks each row's run of occurrences
// once: no atomics, no accumulator rows, the same summation order every time.
//
//   expand        example index of every occurrence (the inverse of feature_poses)
//   per 8-bit digit, least significant first (3 passes for a 2^24 vocabulary):
//     count       digit histogram of every tile of 3072 keys -> counts[digit][tile]
//     scan        one exclusive scan over that matrix, digit-major: the global position of
//                 (digit, tile)'s first key (fm_scan.cuh, single pass)
//     scatter     per tile: stable ranks by warp-wide match, the tile reordered in shared
//                 memory so that the writes are runs, not single words; the last pass writes
//                 the sorted keys and, gathered through the carried occurrence index, the
//                 (example, value) pair of each occurrence
// No block waits on another block.  Runs of equal keys are not listed: the update gives every
// sorted position a lane group and the positions that open a run (or sit on a multiple of
// FM_SORT_CHUNK inside one) do the work.
//
// Everything is sized by the parser's device-resident nnz; 2.5 M occurrences are 30 MB of keys,
// payload and pairs: the whole sort lives in the 126 MB L2.
#include "fm_common.cuh"
#include "fm_scan.cuh"

namespace fm {
namespace {

constexpr int kThreads = 256;
constexpr int kWarpsPerBlock = kThreads / 32;
constexpr int kItems = 12;                          // per thread
constexpr int kTile = kThreads * kItems;            // 3072 keys per tile
constexpr int kWarpSpan = 32 * kItems;              // keys of one warp, in order: round r, lane l
constexpr int kBins = 256;
constexpr int kMaxPasses = 4;
constexpr int kScanItems = 8;
constexpr unsigned long long kNoError = 0xffffffffffffffffull;

struct SortWs {
  char* scan[kMaxPasses];  // scan scratch of each pass (zeroed per call)
  unsigned* counts;        // [kBins][tiles]: digit counts per tile, then (scanned) global offsets
  unsigned* example;       // [cap] example index of every occurrence
  unsigned* keys[2];       // ping-pong
  unsigned* pay[2];
  size_t zero_bytes;
};

static inline int64_t sort_tiles(int64_t cap) { return cap <= 0 ? 1 : (cap + kTile - 1) / kTile; }

static inline size_t carve(size_t* off, size_t bytes) {
  const size_t at = *off;
  *off = fm_align_up(at + bytes, 256);
  return at;
}

static size_t layout(int64_t cap, char* base, SortWs* ws) {
  size_t off = 0;
  const int64_t cells = sort_tiles(cap) * kBins;
  size_t sc[kMaxPasses];
  for (int p = 0; p < kMaxPasses; ++p) sc[p] = carve(&off, scan_scratch_bytes(cells, kScanItems));
  const size_t zero_end = off;
  const size_t c = carve(&off, sizeof(unsigned) * (size_t)cells);
  const size_t e = carve(&off, sizeof(unsigned) * (size_t)cap);
  const size_t k0 = carve(&off, sizeof(unsigned) * (size_t)cap);
  const size_t k1 = carve(&off, sizeof(unsigned) * (size_t)cap);
  const size_t p0 = carve(&off, sizeof(unsigned) * (size_t)cap);
  const size_t p1 = carve(&off, sizeof(unsigned) * (size_t)cap);
  if (ws) {
    for (int p = 0; p < kMaxPasses; ++p) ws->scan[p] = base + sc[p];
    ws->counts = reinterpret_cast<unsigned*>(base + c);
    ws->example = reinterpret_cast<unsigned*>(base + e);
    ws->keys[0] = reinterpret_cast<unsigned*>(base + k0);
    ws->keys[1] = reinterpret_cast<unsigned*>(base + k1);
    ws->pay[0] = reinterpret_cast<unsigned*>(base + p0);
    ws->pay[1] = reinterpret_cast<unsigned*>(base + p1);
    ws->zero_bytes = zero_end;
  }
  return off;
}

__device__ __forceinline__ int64_t live_count(int64_t cap, const int64_t* n_dev, const unsigned long long* guard) {
  if (guard && *guard != kNoError) return 0;     // the batch did not parse: nothing to sort
  if (!n_dev) return cap;
  const int64_t n = *n_dev;
  return n < 0 ? 0 : (n < cap ? n : cap);
}

// sort key of a feature id.  One GPU: the id itself.  A table sharded over n_ranks GPUs (rank id % n owns row
// id / n): owner << row_bits | row, so that a rank's occurrences come out grouped by owner, rows ascending.
__device__ __forceinline__ unsigned key_of(int64_t id, int n_ranks, int row_bits) {
  const unsigned u = (unsigned)id;
  if (n_ranks <= 1) return u;
  return ((u % (unsigned)n_ranks) << row_bits) | (u / (unsigned)n_ranks);
}

// example[j] = i for poses[i] <= j < poses[i + 1]: eight lanes per example write its range
__global__ void __launch_bounds__(kThreads)
expand_kernel(const int32_t* __restrict__ poses, int64_t batch, unsigned* __restrict__ example, int64_t cap,
              const int64_t* __restrict__ n_dev, const unsigned long long* __restrict__ guard) {
  const int64_t n = live_count(cap, n_dev, guard);
  const int sub = threadIdx.x & 7;
  const int64_t groups = (int64_t)gridDim.x * (kThreads / 8);
  for (int64_t i = (int64_t)blockIdx.x * (kThreads / 8) + (threadIdx.x >> 3); i < batch; i += groups) {
    const int64_t a = __ldg(poses + i), b = min((int64_t)__ldg(poses + i + 1), n);
    for (int64_t j = a + sub; j < b; j += 8) example[j] = (unsigned)i;
  }
}

// digit counts of every tile: counts[digit * n_tiles_cap + tile].  Tiles past the live keys write zeros, so the
// matrix needs no clearing.
template <bool FIRST>
__global__ void __launch_bounds__(kThreads)
count_kernel(const int64_t* __restrict__ ids, const unsigned* __restrict__ keys_in, int64_t cap,
             const int64_t* __restrict__ n_dev, const unsigned long long* __restrict__ guard, int n_ranks, int row_bits,
             int shift, unsigned* __restrict__ counts, int64_t n_tiles_cap) {
  __shared__ unsigned s_hist[kWarpsPerBlock][kBins];     // one histogram per warp: fewer same-address collisions
  const int tid = threadIdx.x, warp = tid >> 5;
  const int64_t n = live_count(cap, n_dev, guard);
  for (int64_t tile = blockIdx.x; tile < n_tiles_cap; tile += gridDim.x) {
    for (int i = tid; i < kWarpsPerBlock * kBins; i += kThreads) (&s_hist[0][0])[i] = 0;
    __syncthreads();
    const int64_t tile_base = tile * kTile;
    unsigned key[kItems];
#pragma unroll
    for (int r = 0; r < kItems; ++r) {
      const int64_t j = tile_base + r * kThreads + tid;
      key[r] = j < n ? (FIRST ? key_of(ids[j], n_ranks, row_bits) : keys_in[j]) : 0u;
    }
#pragma unroll
    for (int r = 0; r < kItems; ++r)
      if (tile_base + r * kThreads + tid < n) atomicAdd(&s_hist[warp][(key[r] >> shift) & 255u], 1u);
    __syncthreads();
    unsigned c = 0;
#pragma unroll
    for (int w = 0; w < kWarpsPerBlock; ++w) c += s_hist[w][tid];
    counts[(int64_t)tid * n_tiles_cap + tile] = c;
    __syncthreads();
  }
}

// the exclusive scan over counts[] in place (digit-major: a cell ends up holding the global position of its first key)
struct CellIn {
  const unsigned* cells; int64_t n;
  __device__ int64_t count() const { return n; }
  __device__ long long operator()(int64_t i) const { return cells[i]; }
};
struct CellOut {
  unsigned* cells;
  __device__ void operator()(int64_t i, long long excl, long long) const { cells[i] = (unsigned)excl; }
  __device__ void total(long long) const {}
};

// One stable 8-bit scatter pass.  FIRST: keys come from the ids, the payload is the occurrence index itself.
// LAST: writes the sorted keys and the (example, value bits) pair of each occurrence instead of the payload.
template <bool FIRST, bool LAST>
__global__ void __launch_bounds__(kThreads)
scatter_kernel(const int64_t* __restrict__ ids, const unsigned* __restrict__ keys_in, const unsigned* __restrict__ pay_in,
               unsigned* __restrict__ keys_out, unsigned* __restrict__ pay_out, uint2* __restrict__ pairs_out,
               const unsigned* __restrict__ example, const float* __restrict__ vals, int64_t cap,
               const int64_t* __restrict__ n_dev, const unsigned long long* __restrict__ guard, int n_ranks, int row_bits,
               int shift, const unsigned* __restrict__ offsets, int64_t n_tiles_cap, int64_t* __restrict__ out_counts) {
  __shared__ unsigned s_cnt[kWarpsPerBlock][kBins];   // per-warp digit counts, then the warp's offset inside its digit
  __shared__ unsigned s_keys[kTile], s_pay[kTile];
  __shared__ unsigned s_tile_off[kBins];              // first slot of each digit in the reordered tile
  __shared__ unsigned s_glob[kBins];                  // global position of slot i of digit d = s_glob[d] + i
  __shared__ unsigned s_scan[kWarpsPerBlock];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t n = live_count(cap, n_dev, guard);
  const int64_t n_tiles = (n + kTile - 1) / kTile;
  if (LAST && blockIdx.x == 0 && tid == 0) out_counts[0] = n;

  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    for (int i = tid; i < kWarpsPerBlock * kBins; i += kThreads) (&s_cnt[0][0])[i] = 0;
    const unsigned first_pos = offsets[(int64_t)tid * n_tiles_cap + tile];   // global position of digit tid's first key
    __syncthreads();
    const int64_t tile_base = tile * kTile;
    const int tile_n = (int)min((int64_t)kTile, n - tile_base);

    // ---- load: warp w owns keys [w * kWarpSpan, (w + 1) * kWarpSpan) of the tile, round r lane l ----
    unsigned key[kItems], pay[kItems], rank[kItems], peers[kItems];
#pragma unroll
    for (int r = 0; r < kItems; ++r) {
      const int local = warp * kWarpSpan + r * 32 + lane;
      const int64_t j = tile_base + local;
      const bool valid = local < tile_n;
      if (FIRST) {
        key[r] = valid ? key_of(ids[j], n_ranks, row_bits) : 0u;
        pay[r] = (unsigned)j;
      } else {
        key[r] = valid ? keys_in[j] : 0u;
        pay[r] = valid ? pay_in[j] : 0u;
      }
    }
    // ---- stable ranks: all the matches first (independent), then the warp's counters round by round ----
#pragma unroll
    for (int r = 0; r < kItems; ++r) {
      const bool valid = warp * kWarpSpan + r * 32 + lane < tile_n;
      peers[r] = __match_any_sync(kFull, valid ? (key[r] >> shift) & 255u : 0xffffffffu);   // lanes with my digit
    }
    volatile unsigned* my_cnt = s_cnt[warp];
#pragma unroll
    for (int r = 0; r < kItems; ++r) {
      const bool valid = warp * kWarpSpan + r * 32 + lane < tile_n;
      const unsigned d = (key[r] >> shift) & 255u;
      const unsigned before = valid ? my_cnt[d] : 0u;        // every lane of the group reads the same word ...
      __syncwarp();
      if (valid && lane == __ffs(peers[r]) - 1) my_cnt[d] = before + __popc(peers[r]);   // ... its first lane bumps it
      __syncwarp();
      rank[r] = before + __popc(peers[r] & ((1u << lane) - 1u));
    }
    __syncthreads();

    // ---- digit d (thread d): offsets of the warps inside the digit, the tile's count, its first slot ----
    unsigned count = 0;
#pragma unroll
    for (int w = 0; w < kWarpsPerBlock; ++w) {
      const unsigned c = s_cnt[w][tid];
      s_cnt[w][tid] = count;
      count += c;
    }
    {
      unsigned incl = count;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned v = __shfl_up_sync(kFull, incl, o);
        if (lane >= o) incl += v;
      }
      if (lane == 31) s_scan[warp] = incl;
      __syncthreads();
      unsigned off = 0;
      for (int w = 0; w < warp; ++w) off += s_scan[w];
      const unsigned tile_off = off + incl - count;
      s_tile_off[tid] = tile_off;
      s_glob[tid] = first_pos - tile_off;            // unsigned wrap is fine: + slot index brings it back
    }
    __syncthreads();

    // ---- reorder the tile in shared memory, then write runs ----
#pragma unroll
    for (int r = 0; r < kItems; ++r) {
      const int local = warp * kWarpSpan + r * 32 + lane;
      if (local < tile_n) {
        const unsigned d = (key[r] >> shift) & 255u;
        const unsigned slot = s_tile_off[d] + s_cnt[warp][d] + rank[r];
        s_keys[slot] = key[r];
        s_pay[slot] = pay[r];
      }
    }
    __syncthreads();
    for (int i = tid; i < tile_n; i += kThreads) {
      const unsigned k = s_keys[i], p = s_pay[i];
      const unsigned pos = s_glob[(k >> shift) & 255u] + (unsigned)i;
      keys_out[pos] = k;
      if (LAST) pairs_out[pos] = make_uint2(__ldg(example + p), __float_as_uint(__ldg(vals + p)));
      else pay_out[pos] = p;
    }
    __syncthreads();
  }
}

static inline int bit_length(uint64_t v) {
  int b = 0;
  while (v) { ++b; v >>= 1; }
  return b;
}

}  // namespace
}  // namespace fm

using namespace fm;

FM_EXPORT size_t fm_sort_workspace(int64_t capacity) {
  if (capacity < 0) return 0;
  return layout(capacity, nullptr, nullptr);
}

FM_EXPORT int fm_sort_by_row(const int64_t* feature_ids, const float* feature_vals, const int32_t* feature_poses,
                             int64_t batch, int64_t capacity, const int64_t* nnz_dev, int64_t vocab_size,
                             int64_t n_ranks, uint32_t* sorted_keys, uint32_t* sorted_pairs,
                             int64_t* counts, const uint64_t* guard, void* workspace, size_t workspace_bytes,
                             int max_blocks, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  FM_REQUIRE(batch >= 0 && capacity >= 0 && vocab_size >= 1 && n_ranks >= 1, "fm_sort_by_row: bad shape");
  if (vocab_size > (int64_t)1 << 32 || capacity >= (int64_t)1 << 30)
    return fm_fail(FM_E_LIMIT, "fm_sort_by_row: vocabulary_size <= 2^32 and capacity < 2^30 (got %lld, %lld)",
                   (long long)vocab_size, (long long)capacity);
  FM_REQUIRE(counts && workspace, "fm_sort_by_row: null pointer");
  FM_REQUIRE(capacity == 0 || (feature_ids && feature_vals && feature_poses && sorted_keys && sorted_pairs),
             "fm_sort_by_row: null pointer");
  FM_REQUIRE(((uintptr_t)sorted_pairs & 7) == 0, "fm_sort_by_row: sorted_pairs must be 8-byte aligned");
  if (workspace_bytes < fm_sort_workspace(capacity))
    return fm_fail(FM_E_WORKSPACE, "fm_sort_by_row: workspace %zu < %zu", workspace_bytes, fm_sort_workspace(capacity));
  SortWs ws;
  layout(capacity, static_cast<char*>(workspace), &ws);
  const unsigned long long* g = reinterpret_cast<const unsigned long long*>(guard);
  // key bits: the id itself, or owner << row_bits | row for a sharded table
  const int64_t rows = (vocab_size + n_ranks - 1) / n_ranks;
  const int row_bits = n_ranks > 1 ? bit_length((uint64_t)(rows > 1 ? rows - 1 : 1)) : 0;   // id / n <= rows - 1
  const int key_bits = n_ranks > 1 ? row_bits + bit_length((uint64_t)(n_ranks - 1))
                                   : bit_length((uint64_t)(vocab_size > 1 ? vocab_size - 1 : 1));
  if (key_bits > 32) return fm_fail(FM_E_LIMIT, "fm_sort_by_row: %d key bits (vocabulary x ranks too large)", key_bits);
  const int passes = (key_bits + 7) / 8;

  FM_CUDA(cudaMemsetAsync(workspace, 0, ws.zero_bytes, stream));
  FM_CUDA(cudaMemsetAsync(counts, 0, 2 * sizeof(int64_t), stream));     // stays 0 when there is nothing to sort
  if (capacity == 0 || batch == 0) return FM_OK;
  const int64_t tiles = sort_tiles(capacity);
  const int64_t cells = tiles * kBins;
  const int cap_blocks = max_blocks > 0 ? max_blocks : 148 * 8;
  const unsigned grid = (unsigned)(tiles < cap_blocks ? tiles : cap_blocks);
  int64_t eb = (batch + kThreads / 8 - 1) / (kThreads / 8);
  if (eb > cap_blocks) eb = cap_blocks;
  expand_kernel<<<(unsigned)eb, kThreads, 0, stream>>>(feature_poses, batch, ws.example, capacity, nnz_dev, g);
  FM_LAUNCH_CHECK();
  const int64_t scan_n_tiles = scan_tiles(cells, kScanItems);
  for (int p = 0; p < passes; ++p) {
    const bool first = p == 0, last = p == passes - 1;
    const unsigned* kin = ws.keys[(p + 1) & 1];
    const unsigned* pin = ws.pay[(p + 1) & 1];
    unsigned* kout = last ? sorted_keys : ws.keys[p & 1];
    unsigned* pout = ws.pay[p & 1];
    if (first)
      count_kernel<true><<<grid, kThreads, 0, stream>>>(feature_ids, kin, capacity, nnz_dev, g, (int)n_ranks, row_bits,
                                                        8 * p, ws.counts, tiles);
    else
      count_kernel<false><<<grid, kThreads, 0, stream>>>(feature_ids, kin, capacity, nnz_dev, g, (int)n_ranks, row_bits,
                                                         8 * p, ws.counts, tiles);
    FM_LAUNCH_CHECK();
    scan_kernel<kScanItems><<<(unsigned)scan_n_tiles, kScanThreads, 0, stream>>>(
        CellIn{ws.counts, cells}, CellOut{ws.counts}, scan_scratch_at(ws.scan[p], cells, kScanItems), scan_n_tiles);
    FM_LAUNCH_CHECK();
#define FM_PASS(F, L)                                                                                                    \
  scatter_kernel<F, L><<<grid, kThreads, 0, stream>>>(feature_ids, kin, pin, kout, pout,                                \
                                                      reinterpret_cast<uint2*>(sorted_pairs), ws.example, feature_vals, \
                                                      capacity, nnz_dev, g, (int)n_ranks, row_bits, 8 * p, ws.counts,   \
                                                      tiles, counts)
    if (first && last) { FM_PASS(true, true); }
    else if (first) { FM_PASS(true, false); }
    else if (last) { FM_PASS(false, true); }
    else { FM_PASS(false, false); }
#undef FM_PASS
    FM_LAUNCH_CHECK();
  }
  return FM_OK;
}
