// fm_scan.cuh -- single-pass exclusive prefix sum (decoupled look-back) over a
// functor-defined sequence.  Used for feature_poses = [0, cumsum(sizes)]
// (tffm/fm_model.py:73), the first-occurrence ranks of tf.unique (:72) and the
// line starts of the text buffer.
//
// One tile = 256 threads x ITEMS items.  Tiles take their index from an atomic
// ticket so a tile only ever waits on tiles that already started.  Scratch: one
// 64-bit status word per tile + one ticket counter, zeroed by the caller's
// stream before the launch (fm::scan_scratch_bytes / cudaMemsetAsync).
#pragma once
#include "fm_common.cuh"

namespace fm {

constexpr int kScanThreads = 256;

struct ScanScratch {
  unsigned long long* status;  // [tiles]  (flag << 62) | value ; flag 1 = tile sum, 2 = inclusive prefix
  unsigned int* ticket;
};

static inline int64_t scan_tiles(int64_t n_items, int items_per_thread) {
  const int64_t per_tile = (int64_t)kScanThreads * items_per_thread;
  return n_items <= 0 ? 1 : (n_items + per_tile - 1) / per_tile;
}
static inline size_t scan_scratch_bytes(int64_t n_items, int items_per_thread) {
  return fm_align_up((size_t)scan_tiles(n_items, items_per_thread) * 8 + 16, 256);
}
static inline ScanScratch scan_scratch_at(void* p, int64_t n_items, int items_per_thread) {
  ScanScratch s;
  s.ticket = reinterpret_cast<unsigned int*>(p);
  s.status = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(p) + 16);
  (void)n_items; (void)items_per_thread;
  return s;
}

// Warp 0 of a tile: publish the tile's sum, then look back over the predecessors (32 tiles at a time) for the
// exclusive prefix of the tile; every lane returns it.
__device__ __forceinline__ long long scan_look_back(const ScanScratch& st, int64_t tile, long long block_sum) {
  const int lane = threadIdx.x & 31;
  if (lane == 0) {
    const unsigned long long flag = tile == 0 ? 2ull : 1ull;
    __threadfence();
    atomicExch(&st.status[tile], (flag << 62) | (unsigned long long)block_sum);
  }
  long long prefix = 0;
  int64_t look = tile - 1;
  while (look >= 0) {  // warp-uniform
    const int64_t idx = look - lane;
    unsigned long long s = 0;
    bool done = false;
    while (!done) {
      s = idx >= 0 ? *reinterpret_cast<volatile unsigned long long*>(&st.status[idx]) : (2ull << 62);
      done = __all_sync(kFull, (s >> 62) != 0);
    }
    const unsigned flags = __ballot_sync(kFull, (s >> 62) == 2ull);
    // lanes up to and including the first inclusive prefix contribute
    const int first_incl = flags ? __ffs(flags) - 1 : 32;
    long long v = (lane <= first_incl) ? (long long)(s & ((1ull << 62) - 1)) : 0;
    if (idx < 0) v = 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
    prefix += v;
    if (flags) break;
    look -= 32;
  }
  if (lane == 0 && tile != 0) {
    __threadfence();
    atomicExch(&st.status[tile], (2ull << 62) | (unsigned long long)(prefix + block_sum));
  }
  return prefix;
}

// In:  int64_t count() const;  long long operator()(int64_t i) const   (i < count())
// Out: void operator()(int64_t i, long long exclusive_prefix, long long item) const;
//      void total(long long t) const    -- called once by the last tile
template <int ITEMS, class In, class Out>
__global__ void __launch_bounds__(kScanThreads) scan_kernel(In in, Out out, ScanScratch st, int64_t n_tiles) {
  __shared__ unsigned int s_tile;
  __shared__ long long s_warp[kScanThreads / 32];
  __shared__ long long s_prefix;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_tile = atomicAdd(st.ticket, 1u);
  __syncthreads();
  const int64_t tile = s_tile;
  const int64_t n = in.count();
  const int64_t base = (tile * kScanThreads + tid) * ITEMS;

  long long item[ITEMS];
  long long tsum = 0;
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    item[k] = (base + k < n) ? in(base + k) : 0;
    tsum += item[k];
  }
  // block-wide exclusive scan of the per-thread sums
  long long incl = tsum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const long long v = __shfl_up_sync(kFull, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  long long warp_off = 0, block_sum = 0;
#pragma unroll
  for (int w = 0; w < kScanThreads / 32; ++w) {
    const long long v = s_warp[w];
    if (w < warp) warp_off += v;
    block_sum += v;
  }
  if (warp == 0) {
    const long long prefix = scan_look_back(st, tile, block_sum);
    if (lane == 0) s_prefix = prefix;
  }
  __syncthreads();
  long long run = s_prefix + warp_off + (incl - tsum);
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    if (base + k < n) out(base + k, run, item[k]);
    run += item[k];
  }
  if (tile == n_tiles - 1 && tid == kScanThreads - 1) out.total(s_prefix + block_sum);
}

}  // namespace fm
