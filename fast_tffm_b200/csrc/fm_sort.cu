// fm_sort.cu -- the occurrences of a batch grouped by table row, ahead of the update.
//
// The reference scatters FmGrad's per-occurrence gradients with a CAS-loop float add
// (cc/fm_grad_op.h:3-16, :80-106) and lets TF's unique + SparseApplyAdagrad find the rows
// (tffm/fm_model.py:72, :341-348).  Here the occurrences are SORTED by feature id instead --
// a stable LSD radix sort that needs nothing but the parser's ids, so it runs on the side
// stream next to the parse, under the table kernels of the previous batch -- and the update
// (fm_sorted.cu) walks each row's run of occurrences once: no atomics, no accumulator rows,
// the same summation order every time.
//
//   expand   example index of every occurrence (the inverse of feature_poses)
//   hist     digit histograms of all passes in one read of the ids
//   pass x P 8-bit stable counting pass: per-warp match ranking, tile offsets by decoupled
//            look-back (tiles are taken from a ticket, so a tile only waits on started tiles),
//            tile reordered in shared memory so that the scatter writes runs, not single words
//            (last pass: writes the sorted ids and, gathered through the carried occurrence
//            index, the (example, value) pair of each occurrence)
//   chunks   one scan over the sorted ids: the start of every run of equal ids, runs cut at
//            multiples of kSortChunk so that no work item of the update is longer than that
//            (hot rows under skewed ids), the number of chunks and of distinct ids
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
constexpr unsigned long long kNoError = 0xffffffffffffffffull;

struct SortWs {
  unsigned* hist;        // [kMaxPasses][kBins]
  unsigned* tickets;     // [kMaxPasses + 1] tile tickets of the passes
  unsigned* status;      // [kMaxPasses][tiles][kBins] look-back words: flag << 30 | count
  unsigned* example;     // [cap] example index of every occurrence
  unsigned* keys[2];     // ping-pong
  unsigned* pay[2];
  void* scan;            // scratch of the chunk scan
  size_t zero_bytes;     // hist + tickets + status, zeroed per call
};

static inline int64_t sort_tiles(int64_t cap) { return cap <= 0 ? 1 : (cap + kTile - 1) / kTile; }

static inline size_t carve(size_t* off, size_t bytes) {
  const size_t at = *off;
  *off = fm_align_up(at + bytes, 256);
  return at;
}

static size_t layout(int64_t cap, char* base, SortWs* ws) {
  size_t off = 0;
  const size_t h = carve(&off, sizeof(unsigned) * kMaxPasses * kBins);
  const size_t t = carve(&off, sizeof(unsigned) * 64);
  const size_t s = carve(&off, sizeof(unsigned) * kMaxPasses * (size_t)sort_tiles(cap) * kBins);
  const size_t zero_end = off;
  const size_t e = carve(&off, sizeof(unsigned) * (size_t)cap);
  const size_t k0 = carve(&off, sizeof(unsigned) * (size_t)cap);
  const size_t k1 = carve(&off, sizeof(unsigned) * (size_t)cap);
  const size_t p0 = carve(&off, sizeof(unsigned) * (size_t)cap);
  const size_t p1 = carve(&off, sizeof(unsigned) * (size_t)cap);
  const size_t sc = carve(&off, scan_scratch_bytes(cap, 8));
  if (ws) {
    ws->hist = reinterpret_cast<unsigned*>(base + h);
    ws->tickets = reinterpret_cast<unsigned*>(base + t);
    ws->status = reinterpret_cast<unsigned*>(base + s);
    ws->example = reinterpret_cast<unsigned*>(base + e);
    ws->keys[0] = reinterpret_cast<unsigned*>(base + k0);
    ws->keys[1] = reinterpret_cast<unsigned*>(base + k1);
    ws->pay[0] = reinterpret_cast<unsigned*>(base + p0);
    ws->pay[1] = reinterpret_cast<unsigned*>(base + p1);
    ws->scan = base + sc;
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

// example[j] = i for poses[i] <= j < poses[i + 1]
__global__ void __launch_bounds__(kThreads)
expand_kernel(const int32_t* __restrict__ poses, int64_t batch, unsigned* __restrict__ example,
              int64_t cap, const unsigned long long* __restrict__ guard) {
  if (guard && *guard != kNoError) return;
  const int lane = threadIdx.x & 31;
  const int64_t warps = (int64_t)gridDim.x * kWarpsPerBlock;
  for (int64_t i = (int64_t)blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5); i < batch; i += warps) {
    const int64_t a = poses[i], b = min((int64_t)poses[i + 1], cap);
    for (int64_t j = a + lane; j < b; j += 32) example[j] = (unsigned)i;
  }
}

// digit histograms of every pass, one read of the ids
__global__ void __launch_bounds__(kThreads)
hist_kernel(const int64_t* __restrict__ ids, int64_t cap, const int64_t* __restrict__ n_dev,
            const unsigned long long* __restrict__ guard, int n_ranks, int row_bits, int passes,
            unsigned* __restrict__ hist) {
  __shared__ unsigned s_hist[kMaxPasses * kBins];
  const int64_t n = live_count(cap, n_dev, guard);
  for (int i = threadIdx.x; i < kMaxPasses * kBins; i += kThreads) s_hist[i] = 0;
  __syncthreads();
  for (int64_t j = (int64_t)blockIdx.x * kThreads + threadIdx.x; j < n; j += (int64_t)gridDim.x * kThreads) {
    const unsigned key = key_of(ids[j], n_ranks, row_bits);
    for (int p = 0; p < passes; ++p) atomicAdd(&s_hist[p * kBins + ((key >> (8 * p)) & 255u)], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < passes * kBins; i += kThreads)
    if (s_hist[i]) atomicAdd(&hist[i], s_hist[i]);
}

// One stable 8-bit counting pass.  FIRST: keys come from the ids, the payload is the occurrence index itself.
// LAST: writes the sorted keys and the (example, value bits) pair of each occurrence instead of the payload.
template <bool FIRST, bool LAST>
__global__ void __launch_bounds__(kThreads)
pass_kernel(const int64_t* __restrict__ ids, const unsigned* __restrict__ keys_in, const unsigned* __restrict__ pay_in,
            unsigned* __restrict__ keys_out, unsigned* __restrict__ pay_out, uint2* __restrict__ pairs_out,
            const unsigned* __restrict__ example, const float* __restrict__ vals, int64_t cap,
            const int64_t* __restrict__ n_dev, const unsigned long long* __restrict__ guard, int n_ranks, int row_bits,
            int shift, const unsigned* __restrict__ hist, unsigned* __restrict__ status, unsigned* __restrict__ ticket) {
  __shared__ unsigned s_cnt[kWarpsPerBlock][kBins];   // per-warp digit counts, then the warp's offset inside its digit
  __shared__ unsigned s_keys[kTile], s_pay[kTile];
  __shared__ unsigned s_tile_off[kBins];              // first slot of each digit in the reordered tile
  __shared__ unsigned s_glob[kBins];                  // global position of slot i of digit d = s_glob[d] + i
  __shared__ unsigned s_base[kBins];                  // exclusive scan of the pass histogram
  __shared__ unsigned s_scan[kWarpsPerBlock];
  __shared__ unsigned s_tile;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t n = live_count(cap, n_dev, guard);
  if (n == 0) return;
  const int64_t n_tiles = (n + kTile - 1) / kTile;

  {  // s_base: exclusive scan over the 256 digit totals (thread d owns digit d)
    const unsigned c = hist[tid];
    unsigned incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned v = __shfl_up_sync(kFull, incl, o);
      if (lane >= o) incl += v;
    }
    if (lane == 31) s_scan[warp] = incl;
    __syncthreads();
    unsigned off = 0;
    for (int w = 0; w < warp; ++w) off += s_scan[w];
    s_base[tid] = off + incl - c;
    __syncthreads();
  }

  while (true) {
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);
    for (int i = tid; i < kWarpsPerBlock * kBins; i += kThreads) (&s_cnt[0][0])[i] = 0;
    __syncthreads();
    const int64_t tile = s_tile;
    if (tile >= n_tiles) return;
    const int64_t tile_base = tile * kTile;
    const int tile_n = (int)min((int64_t)kTile, n - tile_base);

    // ---- load + rank: warp w owns keys [w * kWarpSpan, (w + 1) * kWarpSpan) of the tile, round r lane l ----
    unsigned key[kItems], pay[kItems], rank[kItems];
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
#pragma unroll
    for (int r = 0; r < kItems; ++r) {
      const int local = warp * kWarpSpan + r * 32 + lane;
      const bool valid = local < tile_n;
      const unsigned d = (key[r] >> shift) & 255u;
      const unsigned peers = __match_any_sync(kFull, valid ? d : 0xffffffffu);   // lanes with my digit (in lane order)
      const int leader = __ffs(peers) - 1;
      unsigned before = 0;
      if (valid && lane == leader) {
        before = s_cnt[warp][d];
        s_cnt[warp][d] = before + __popc(peers);
      }
      before = __shfl_sync(kFull, before, leader);
      rank[r] = before + __popc(peers & ((1u << lane) - 1u));
      __syncwarp();
    }
    __syncthreads();

    // ---- digit d (thread d): offsets of the warps inside the digit, the tile's count ----
    unsigned count = 0;
#pragma unroll
    for (int w = 0; w < kWarpsPerBlock; ++w) {
      const unsigned c = s_cnt[w][tid];
      s_cnt[w][tid] = count;
      count += c;
    }
    // publish the tile's count for digit d, then sum the predecessors' (decoupled look-back per digit)
    unsigned* st = status + (size_t)tile * kBins + tid;
    atomicExch(st, ((tile == 0 ? 2u : 1u) << 30) | count);
    unsigned excl = 0;
    for (int64_t t = tile - 1; t >= 0; --t) {
      const volatile unsigned* sp = status + (size_t)t * kBins + tid;
      unsigned s;
      do { s = *sp; } while ((s >> 30) == 0u);
      excl += s & 0x3fffffffu;
      if ((s >> 30) == 2u) break;
    }
    if (tile != 0) atomicExch(st, (2u << 30) | (excl + count));
    {  // exclusive scan of the tile's digit counts
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
      s_glob[tid] = s_base[tid] + excl - tile_off;     // unsigned wrap is fine: + slot index brings it back
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
      if (LAST) pairs_out[pos] = make_uint2(example[p], __float_as_uint(vals[p]));
      else pay_out[pos] = p;
    }
    __syncthreads();
  }
}

// chunk starts: p is a start if it opens a run of equal keys or is a multiple of kSortChunk
struct ChunkIn {
  const unsigned* keys; int64_t cap; const int64_t* n_dev; const unsigned long long* guard;
  __device__ int64_t count() const { return live_count(cap, n_dev, guard); }
  __device__ long long operator()(int64_t p) const {
    const bool head = p == 0 || keys[p] != keys[p - 1];
    const bool start = head || (p % FM_SORT_CHUNK) == 0;
    return (long long)start | ((long long)head << 32);      // both counts ride one scan
  }
};
struct ChunkOut {
  unsigned* starts; int64_t* counts; int64_t cap; const int64_t* n_dev; const unsigned long long* guard;
  __device__ void operator()(int64_t p, long long excl, long long item) const {
    if (item & 1) starts[excl & 0xffffffffll] = (unsigned)p;
  }
  __device__ void total(long long t) const {
    const int64_t n = live_count(cap, n_dev, guard);
    const long long chunks = t & 0xffffffffll;
    starts[chunks] = (unsigned)n;
    counts[0] = chunks;
    counts[1] = t >> 32;
    counts[2] = n;
  }
};

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
                             int64_t n_ranks, uint32_t* sorted_keys, uint32_t* sorted_pairs, uint32_t* chunk_starts,
                             int64_t* counts, const uint64_t* guard, void* workspace, size_t workspace_bytes,
                             int max_blocks, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  FM_REQUIRE(batch >= 0 && capacity >= 0 && vocab_size >= 1 && n_ranks >= 1, "fm_sort_by_row: bad shape");
  if (vocab_size > (int64_t)1 << 32 || capacity >= (int64_t)1 << 30)
    return fm_fail(FM_E_LIMIT, "fm_sort_by_row: vocabulary_size <= 2^32 and capacity < 2^30 (got %lld, %lld)",
                   (long long)vocab_size, (long long)capacity);
  FM_REQUIRE(counts && chunk_starts && workspace, "fm_sort_by_row: null pointer");
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
  FM_CUDA(cudaMemsetAsync(ws.scan, 0, scan_scratch_bytes(capacity, 8), stream));
  const int64_t tiles = sort_tiles(capacity);
  const int cap_blocks = max_blocks > 0 ? max_blocks : 148 * 4;
  const unsigned grid = (unsigned)(tiles < cap_blocks ? tiles : cap_blocks);
  if (capacity > 0 && batch > 0) {
    int64_t eb = (batch + kWarpsPerBlock - 1) / kWarpsPerBlock;
    if (eb > cap_blocks) eb = cap_blocks;
    expand_kernel<<<(unsigned)eb, kThreads, 0, stream>>>(feature_poses, batch, ws.example, capacity, g);
    FM_LAUNCH_CHECK();
    hist_kernel<<<grid, kThreads, 0, stream>>>(feature_ids, capacity, nnz_dev, g, (int)n_ranks, row_bits, passes, ws.hist);
    FM_LAUNCH_CHECK();
    for (int p = 0; p < passes; ++p) {
      const bool first = p == 0, last = p == passes - 1;
      const unsigned* kin = ws.keys[(p + 1) & 1];
      const unsigned* pin = ws.pay[(p + 1) & 1];
      unsigned* kout = last ? sorted_keys : ws.keys[p & 1];
      unsigned* pout = ws.pay[p & 1];
      unsigned* status = ws.status + (size_t)p * tiles * kBins;
#define FM_PASS(F, L)                                                                                                  \
  pass_kernel<F, L><<<grid, kThreads, 0, stream>>>(feature_ids, kin, pin, kout, pout,                                  \
                                                   reinterpret_cast<uint2*>(sorted_pairs), ws.example, feature_vals,   \
                                                   capacity, nnz_dev, g, (int)n_ranks, row_bits, 8 * p,                \
                                                   ws.hist + p * kBins, status, ws.tickets + p)
      if (first && last) { FM_PASS(true, true); }
      else if (first) { FM_PASS(true, false); }
      else if (last) { FM_PASS(false, true); }
      else { FM_PASS(false, false); }
#undef FM_PASS
      FM_LAUNCH_CHECK();
    }
  }
  ChunkIn in{sorted_keys, capacity, nnz_dev, g};
  ChunkOut out{chunk_starts, counts, capacity, nnz_dev, g};
  const int64_t scan_n_tiles = scan_tiles(capacity, 8);
  // the scan walks live_count() items; its last tile (by capacity) reports the total, so size it by capacity
  scan_kernel<8><<<(unsigned)scan_n_tiles, kScanThreads, 0, stream>>>(in, out, scan_scratch_at(ws.scan, capacity, 8),
                                                                      scan_n_tiles);
  FM_LAUNCH_CHECK();
  return FM_OK;
}
