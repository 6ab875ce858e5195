// fm_shard.cu -- kernels around the all-to-all exchange of the vocab-sharded table.
//
// The reference shards the table over TF parameter servers: vocab_block_i lives on
// ps task i % n_ps, embedding_lookup's "mod" strategy routes id -> block id % nb,
// row id // nb, and sparse gradients travel back the same way
// (tffm/fm_model.py:269-291, run_tffm.py:181-206).  Here rank id % G owns row id // G
// and the routing is explicit:
//   fm_shard_bucket      stable counting sort of the batch's distinct ids by owner:
//                        send_ids (grouped by owner), pos[u] = slot of distinct id u
//                        in that order, counts[G]
//   fm_shard_remap       per occurrence: slot = pos[local_id]   (index into the rows
//                        that come back from the owners, and into the gradient
//                        accumulator rows that go out to them)
//   fm_shard_local_rows  owner side: row = id / G
//   fm_shard_combine     owner side: accumulator rows received from several ranks for
//                        the same id are summed BEFORE Adagrad squares the gradient, so
//                        the update equals the single-process one
#include "fm_common.cuh"

namespace fm {
namespace {

constexpr int kBThreads = 256, kBItems = 4, kBTile = kBThreads * kBItems, kMaxRanks = 64;

__device__ __forceinline__ int owner_of(int64_t id, int g) { return (int)((uint64_t)id % (uint64_t)g); }

__global__ void __launch_bounds__(kBThreads)
bucket_hist_kernel(const int64_t* __restrict__ ids, int64_t n_max, const int64_t* __restrict__ n_dev, int g,
                   int* __restrict__ block_counts) {
  __shared__ int s_cnt[kMaxRanks];
  const int64_t n = n_dev ? min(*n_dev, n_max) : n_max;
  if (threadIdx.x < kMaxRanks) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  const int64_t base = (int64_t)blockIdx.x * kBTile + threadIdx.x * kBItems;
#pragma unroll
  for (int k = 0; k < kBItems; ++k)
    if (base + k < n) atomicAdd(&s_cnt[owner_of(ids[base + k], g)], 1);
  __syncthreads();
  if (threadIdx.x < g) block_counts[(int64_t)blockIdx.x * g + threadIdx.x] = s_cnt[threadIdx.x];
}

// exclusive scan of block_counts in owner-major order: base[blk][o] = sum of counts of
// all smaller owners + same owner in earlier blocks.  One block.
__global__ void __launch_bounds__(256)
bucket_offsets_kernel(const int* __restrict__ block_counts, int64_t n_blocks, int g, int64_t* __restrict__ block_base,
                      int64_t* __restrict__ counts) {
  __shared__ long long s_part[256];
  __shared__ long long s_owner_base[kMaxRanks + 1];
  const int tid = threadIdx.x;
  if (tid == 0) s_owner_base[0] = 0;
  for (int o = 0; o < g; ++o) {
    // per-thread contiguous segment of blocks
    const int64_t per = (n_blocks + 255) / 256, lo = tid * per, hi = min(lo + per, n_blocks);
    long long sum = 0;
    for (int64_t b = lo; b < hi; ++b) sum += block_counts[b * g + o];
    s_part[tid] = sum;
    __syncthreads();
    if (tid == 0) {
      long long run = 0;
      for (int t = 0; t < 256; ++t) { const long long v = s_part[t]; s_part[t] = run; run += v; }
      counts[o] = run;
      s_owner_base[o + 1] = s_owner_base[o] + run;
    }
    __syncthreads();
    long long run = s_owner_base[o] + s_part[tid];
    for (int64_t b = lo; b < hi; ++b) { block_base[b * g + o] = run; run += block_counts[b * g + o]; }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(kBThreads)
bucket_scatter_kernel(const int64_t* __restrict__ ids, int64_t n_max, const int64_t* __restrict__ n_dev, int g,
                      const int64_t* __restrict__ block_base, int64_t* __restrict__ send_ids,
                      int32_t* __restrict__ pos) {
  __shared__ int s_warp[kBThreads / 32];
  const int64_t n = n_dev ? min(*n_dev, n_max) : n_max;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t base = (int64_t)blockIdx.x * kBTile + tid * kBItems;
  int64_t id[kBItems];
  int own[kBItems];
#pragma unroll
  for (int k = 0; k < kBItems; ++k) {
    id[k] = base + k < n ? ids[base + k] : 0;
    own[k] = base + k < n ? owner_of(id[k], g) : -1;
  }
  for (int o = 0; o < g; ++o) {
    int mine = 0;
#pragma unroll
    for (int k = 0; k < kBItems; ++k) mine += own[k] == o;
    int incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int v = __shfl_up_sync(kFull, incl, d);
      if (lane >= d) incl += v;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    int before = incl - mine;
    for (int w = 0; w < warp; ++w) before += s_warp[w];
    int64_t slot = block_base[(int64_t)blockIdx.x * g + o] + before;
#pragma unroll
    for (int k = 0; k < kBItems; ++k)
      if (own[k] == o) {
        send_ids[slot] = id[k];
        pos[base + k] = (int32_t)slot;
        ++slot;
      }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(256)
remap_kernel(const int32_t* __restrict__ local, int64_t n_max, const int64_t* __restrict__ n_dev,
             const int32_t* __restrict__ pos, int64_t* __restrict__ slot64, int32_t* __restrict__ slot32) {
  const int64_t n = n_dev ? min(*n_dev, n_max) : n_max;
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
    const int32_t s = pos[local[j]];
    slot64[j] = s;
    slot32[j] = s;
  }
}

__global__ void __launch_bounds__(256)
local_rows_kernel(const int64_t* __restrict__ ids, int64_t n, int g, int64_t* __restrict__ rows) {
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x)
    rows[j] = (int64_t)((uint64_t)ids[j] / (uint64_t)g);
}

// dst[index[i], :] += src[i, :]  as float4 vector atomics; ld multiple of 4
__global__ void __launch_bounds__(256)
combine_rows_kernel(const float* __restrict__ src, int64_t src_ld, const int64_t* __restrict__ index, int64_t n,
                    int vec_per_row, float* __restrict__ dst, int64_t dst_ld) {
  const int64_t total = n * vec_per_row;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t i = t / vec_per_row;
    const int v = (int)(t - i * vec_per_row);
    const float4 x = __ldg(reinterpret_cast<const float4*>(src + i * src_ld) + v);
    float* d = dst + index[i] * dst_ld + 4 * v;
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(d), "f"(x.x), "f"(x.y), "f"(x.z), "f"(x.w)
                 : "memory");
  }
}

static inline unsigned grid_for(int64_t n, int per_block) {
  int64_t b = (n + per_block - 1) / per_block;
  if (b > 148 * 16) b = 148 * 16;
  return (unsigned)(b < 1 ? 1 : b);
}

}  // namespace
}  // namespace fm

using namespace fm;

FM_EXPORT size_t fm_shard_bucket_workspace(int64_t n, int64_t n_ranks) {
  const int64_t blocks = n <= 0 ? 1 : (n + kBTile - 1) / kBTile;
  return fm_align_up((size_t)blocks * (size_t)n_ranks * (sizeof(int) + sizeof(int64_t)) + 256, 256);
}

FM_EXPORT int fm_shard_bucket(const int64_t* ids, int64_t n, const int64_t* n_dev, int64_t n_ranks, int64_t* send_ids,
                              int32_t* pos, int64_t* counts, void* workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  FM_REQUIRE(n >= 0 && n_ranks >= 1, "fm_shard_bucket: bad size");
  if (n_ranks > kMaxRanks) return fm_fail(FM_E_LIMIT, "fm_shard_bucket: more than %d ranks", kMaxRanks);
  FM_REQUIRE(counts && workspace, "fm_shard_bucket: null pointer");
  FM_REQUIRE(n == 0 || (ids && send_ids && pos), "fm_shard_bucket: null pointer");
  if (workspace_bytes < fm_shard_bucket_workspace(n, n_ranks))
    return fm_fail(FM_E_WORKSPACE, "fm_shard_bucket: workspace too small");
  const int64_t blocks = n <= 0 ? 1 : (n + kBTile - 1) / kBTile;
  int64_t* block_base = reinterpret_cast<int64_t*>(workspace);
  int* block_counts = reinterpret_cast<int*>(block_base + blocks * n_ranks);
  bucket_hist_kernel<<<(unsigned)blocks, kBThreads, 0, stream>>>(ids, n, n_dev, (int)n_ranks, block_counts);
  FM_LAUNCH_CHECK();
  bucket_offsets_kernel<<<1, 256, 0, stream>>>(block_counts, blocks, (int)n_ranks, block_base, counts);
  FM_LAUNCH_CHECK();
  if (n > 0) {
    bucket_scatter_kernel<<<(unsigned)blocks, kBThreads, 0, stream>>>(ids, n, n_dev, (int)n_ranks, block_base, send_ids,
                                                                      pos);
    FM_LAUNCH_CHECK();
  }
  return FM_OK;
}

FM_EXPORT int fm_shard_remap(const int32_t* local_ids, int64_t n, const int64_t* n_dev, const int32_t* pos,
                             int64_t* slot64, int32_t* slot32, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  FM_REQUIRE(n >= 0, "fm_shard_remap: negative n");
  if (n == 0) return FM_OK;
  FM_REQUIRE(local_ids && pos && slot64 && slot32, "fm_shard_remap: null pointer");
  remap_kernel<<<grid_for(n, 256), 256, 0, stream>>>(local_ids, n, n_dev, pos, slot64, slot32);
  FM_LAUNCH_CHECK();
  return FM_OK;
}

FM_EXPORT int fm_shard_local_rows(const int64_t* ids, int64_t n, int64_t n_ranks, int64_t* rows, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  FM_REQUIRE(n >= 0 && n_ranks >= 1, "fm_shard_local_rows: bad size");
  if (n == 0) return FM_OK;
  FM_REQUIRE(ids && rows, "fm_shard_local_rows: null pointer");
  local_rows_kernel<<<grid_for(n, 256), 256, 0, stream>>>(ids, n, (int)n_ranks, rows);
  FM_LAUNCH_CHECK();
  return FM_OK;
}

FM_EXPORT int fm_shard_combine(const float* src_rows, int64_t src_ld, const int64_t* index, int64_t n, int64_t width,
                               float* dst_rows, int64_t dst_ld, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  FM_REQUIRE(n >= 0 && width > 0 && width % 4 == 0 && src_ld % 4 == 0 && dst_ld % 4 == 0 && src_ld >= width &&
                 dst_ld >= width, "fm_shard_combine: width and strides must be positive multiples of 4");
  if (n == 0) return FM_OK;
  FM_REQUIRE(src_rows && index && dst_rows, "fm_shard_combine: null pointer");
  FM_REQUIRE((((uintptr_t)src_rows | (uintptr_t)dst_rows) & 15) == 0, "fm_shard_combine: 16-byte alignment required");
  combine_rows_kernel<<<grid_for(n * (width / 4), 256), 256, 0, stream>>>(src_rows, src_ld, index, n, (int)(width / 4),
                                                                        dst_rows, dst_ld);
  FM_LAUNCH_CHECK();
  return FM_OK;
}

// ---- barrier + tiny all-reduce over peer-mapped memory (the sharded step's two synchronisation points) ----
// Every rank owns FM_PEER_SYNC_BYTES of flag slots inside the allocation that is mapped into all processes:
// slots[parity][sender] = {epoch, word, a, b}.  A call publishes this rank's {word, a, b} into every rank's
// slots[epoch & 1][rank] (payload first, then the epoch with release semantics at system scope), waits until all
// n senders' slots of that parity carry this epoch, and reduces: max of the words, sums of a and b in rank order.
// Ranks move in lock step (nobody can be more than one call ahead of the slowest rank), so two parities suffice.
namespace fm {
namespace {

struct SyncSlot { unsigned long long epoch; long long word; float a, b; unsigned long long pad; };
static_assert(sizeof(SyncSlot) == 32, "slot layout");
constexpr int kSyncMaxRanks = 16;

struct SyncBases { SyncSlot* base[kSyncMaxRanks]; };

__global__ void peer_sync_kernel(const SyncBases bases, int rank, int n, unsigned long long epoch,
                                 const long long* __restrict__ word_in, const float* __restrict__ a_in,
                                 const float* __restrict__ b_in, long long* __restrict__ word_out,
                                 float* __restrict__ sums_out) {
  const int r = threadIdx.x;
  const int parity = (int)(epoch & 1ull);
  long long word = word_in ? *word_in : -1;
  if (word >= 0) word += (long long)rank << 48;           // a parse status word: tag it with the rank that failed
  if (r < n) {
    SyncSlot* dst = bases.base[r] + parity * kSyncMaxRanks + rank;
    dst->word = word;
    dst->a = a_in ? *a_in : 0.f;
    dst->b = b_in ? *b_in : 0.f;
    __threadfence_system();                               // payload (and everything this stream did before) first
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(&dst->epoch), "l"(epoch) : "memory");
  }
  float a = 0.f, b = 0.f;
  word = -1;
  if (r < n) {
    const SyncSlot* src = bases.base[rank] + parity * kSyncMaxRanks + r;
    unsigned long long seen;
    unsigned polls = 0;
    do {
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(seen) : "l"(&src->epoch) : "memory");
    } while (seen != epoch && ++polls < (1u << 26));       // tens of seconds: a rank that died must not hang the box
    word = *reinterpret_cast<const volatile long long*>(&src->word);
    if (seen != epoch) word = (1ll << 62) | ((long long)r << 48);   // "rank r never arrived": voids the step
    a = *reinterpret_cast<const volatile float*>(&src->a);
    b = *reinterpret_cast<const volatile float*>(&src->b);
  }
  // fixed-order reduction over the ranks (lane r holds rank r's contribution)
  long long wmax = -1;
  float asum = 0.f, bsum = 0.f;
  for (int i = 0; i < n; ++i) {
    const long long wi = __shfl_sync(kFull, word, i);
    wmax = wi > wmax ? wi : wmax;
    asum += __shfl_sync(kFull, a, i);
    bsum += __shfl_sync(kFull, b, i);
  }
  if (r == 0) {
    if (word_out) *word_out = wmax;
    if (sums_out) { sums_out[0] = asum; sums_out[1] = bsum; }
  }
  __threadfence_system();
}

}  // namespace
}  // namespace fm

FM_EXPORT size_t fm_peer_sync_bytes(void) { return 2 * kSyncMaxRanks * sizeof(SyncSlot); }

FM_EXPORT int fm_peer_sync(void* const* sync_bases, int64_t rank, int64_t n_ranks, uint64_t epoch,
                           const int64_t* word_in, const float* a_in, const float* b_in, int64_t* word_out,
                           float* sums_out, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (n_ranks < 1 || n_ranks > kSyncMaxRanks) return fm_fail(FM_E_LIMIT, "fm_peer_sync: 1..%d ranks", kSyncMaxRanks);
  FM_REQUIRE(sync_bases && rank >= 0 && rank < n_ranks && epoch > 0, "fm_peer_sync: bad argument (epoch starts at 1)");
  SyncBases bases;
  for (int r = 0; r < kSyncMaxRanks; ++r) {
    bases.base[r] = r < n_ranks ? static_cast<SyncSlot*>(sync_bases[r]) : nullptr;
    FM_REQUIRE(r >= n_ranks || (bases.base[r] && ((uintptr_t)bases.base[r] & 31) == 0), "fm_peer_sync: bad base");
  }
  peer_sync_kernel<<<1, 32, 0, stream>>>(bases, (int)rank, (int)n_ranks, (unsigned long long)epoch,
                                         reinterpret_cast<const long long*>(word_in), a_in, b_in,
                                         reinterpret_cast<long long*>(word_out), sums_out);
  FM_LAUNCH_CHECK();
  return FM_OK;
}
