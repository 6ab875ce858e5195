// fm_unique.cu -- tf.unique on the GPU (tffm/fm_model.py:72, also :133, :173).
//
// ori_ids = distinct feature ids in order of FIRST OCCURRENCE, local_ids[j] =
// index of ids[j] in ori_ids.  A sort-unique would give sorted order, so:
//   1. insert: open-addressing hash table keyed by id; every occurrence j does
//      atomicMin(first_pos[slot], j) and remembers its slot;
//   2. scan:   flag[j] = (first_pos[slot[j]] == j); the exclusive prefix sum of
//      the flags is the rank of each first occurrence -> ori_ids[rank] = id,
//      and the rank is stored back into the table (tagged with the top bit);
//   3. finish: local_ids[j] = rank stored at slot[j].
// Scratch (caller's workspace): table keys u64[cap] + first_pos u32[cap]
// (cap = power of two >= n / 0.7), slot i32[n], scan tile status.
#include "fm_common.cuh"
#include "fm_scan.cuh"

namespace fm {
namespace {

constexpr unsigned long long kEmpty = 0xffffffffffffffffull;
constexpr int kScanItems = 8;

static inline int64_t table_capacity(int64_t n) {
  int64_t cap = 1024;
  while (cap * 7 < n * 10) cap <<= 1;
  return cap;
}

struct UniqueWs {
  void* scan;                 // scan scratch (zeroed)
  unsigned long long* keys;   // [cap]  (0xff-filled)
  unsigned int* first_pos;    // [cap]  (0xff-filled)
  int32_t* slot;              // [n]
  size_t scan_bytes, table_bytes, total;
  int64_t cap;
};

static inline UniqueWs layout(void* base, int64_t n) {
  UniqueWs w;
  w.cap = table_capacity(n);
  w.scan_bytes = scan_scratch_bytes(n, kScanItems);
  w.table_bytes = fm_align_up((size_t)w.cap * 12, 256);
  char* p = reinterpret_cast<char*>(base);
  w.scan = p;
  w.keys = reinterpret_cast<unsigned long long*>(p + w.scan_bytes);
  w.first_pos = reinterpret_cast<unsigned int*>(p + w.scan_bytes + (size_t)w.cap * 8);
  w.slot = reinterpret_cast<int32_t*>(p + w.scan_bytes + w.table_bytes);
  w.total = w.scan_bytes + w.table_bytes + fm_align_up((size_t)(n > 0 ? n : 1) * 4, 256);
  return w;
}

__device__ __forceinline__ uint32_t mix(uint64_t k) {
  k *= 0x9E3779B97F4A7C15ull;
  return (uint32_t)(k >> 32) ^ (uint32_t)k;
}

__global__ void __launch_bounds__(256)
unique_insert_kernel(const int64_t* __restrict__ ids, int64_t n_max, const int64_t* __restrict__ n_dev,
                     unsigned long long* keys, unsigned int* first_pos, int32_t* __restrict__ slot,
                     uint32_t mask) {
  const int64_t n = n_dev ? min(*n_dev, n_max) : n_max;
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
    const unsigned long long key = (unsigned long long)ids[j];
    uint32_t h = mix(key) & mask;
    for (;;) {
      unsigned long long k = *reinterpret_cast<volatile unsigned long long*>(&keys[h]);
      if (k == kEmpty) k = atomicCAS(&keys[h], kEmpty, key);
      if (k == kEmpty || k == key) break;
      h = (h + 1) & mask;
    }
    atomicMin(&first_pos[h], (unsigned int)j);
    slot[j] = (int32_t)h;
  }
}

struct FirstFlagIn {
  const int32_t* slot; const unsigned int* first_pos; int64_t n_max; const int64_t* n_dev;
  __device__ int64_t count() const { return n_dev ? min(*n_dev, n_max) : n_max; }
  __device__ long long operator()(int64_t j) const {
    return *reinterpret_cast<const volatile unsigned int*>(&first_pos[slot[j]]) == (unsigned int)j;
  }
};
struct RankOut {
  const int64_t* ids; const int32_t* slot; unsigned int* first_pos; int64_t* ori_ids; int64_t* n_unique;
  __device__ void operator()(int64_t j, long long excl, long long flag) const {
    if (!flag) return;
    ori_ids[excl] = ids[j];
    // only the first occurrence ever writes its slot; the tag keeps readers that still compare
    // first_pos against their own (smaller than 2^31) position from matching
    first_pos[slot[j]] = 0x80000000u | (unsigned int)excl;
  }
  __device__ void total(long long t) const { *n_unique = t; }
};

__global__ void __launch_bounds__(256)
unique_finish_kernel(const int32_t* __restrict__ slot, const unsigned int* __restrict__ first_pos,
                     int64_t n_max, const int64_t* __restrict__ n_dev, int32_t* __restrict__ local_ids) {
  const int64_t n = n_dev ? min(*n_dev, n_max) : n_max;
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x)
    local_ids[j] = (int32_t)(first_pos[slot[j]] & 0x7fffffffu);
}

}  // namespace
}  // namespace fm

using namespace fm;

FM_EXPORT size_t fm_unique_workspace(int64_t n) { return layout(nullptr, n < 0 ? 0 : n).total; }

FM_EXPORT int fm_unique(const int64_t* ids, int64_t n, const int64_t* n_dev, int64_t* ori_ids,
                        int32_t* local_ids, int64_t* n_unique, void* workspace, size_t workspace_bytes,
                        void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  FM_REQUIRE(n >= 0, "fm_unique: negative n");
  if (n >= (1ll << 31)) return fm_fail(FM_E_LIMIT, "fm_unique: n = %lld >= 2^31", (long long)n);
  FM_REQUIRE(n_unique && workspace, "fm_unique: null pointer");
  FM_REQUIRE(n == 0 || (ids && ori_ids && local_ids), "fm_unique: null pointer");
  const UniqueWs w = layout(workspace, n);
  if (workspace_bytes < w.total)
    return fm_fail(FM_E_WORKSPACE, "fm_unique: workspace %zu < %zu", workspace_bytes, w.total);
  FM_CUDA(cudaMemsetAsync(w.scan, 0, w.scan_bytes, stream));
  FM_CUDA(cudaMemsetAsync(w.keys, 0xff, (size_t)w.cap * 12, stream));
  const int threads = 256;
  int64_t nb = (n + threads - 1) / threads;
  if (nb > 148 * 16) nb = 148 * 16;
  const unsigned blocks = (unsigned)nb;
  if (n > 0) {
    unique_insert_kernel<<<blocks, threads, 0, stream>>>(ids, n, n_dev, w.keys, w.first_pos, w.slot,
                                                         (uint32_t)(w.cap - 1));
    FM_LAUNCH_CHECK();
  }
  FirstFlagIn in{w.slot, w.first_pos, n, n_dev};
  RankOut out{ids, w.slot, w.first_pos, ori_ids, n_unique};
  const int64_t tiles = scan_tiles(n, kScanItems);
  scan_kernel<kScanItems><<<(unsigned)tiles, kScanThreads, 0, stream>>>(in, out, scan_scratch_at(w.scan, n, kScanItems), tiles);
  FM_LAUNCH_CHECK();
  if (n > 0) {
    unique_finish_kernel<<<blocks, threads, 0, stream>>>(w.slot, w.first_pos, n, n_dev, local_ids);
    FM_LAUNCH_CHECK();
  }
  return FM_OK;
}
