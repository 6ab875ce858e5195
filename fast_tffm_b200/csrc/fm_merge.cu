// fm_merge.cu -- the sharded step's gradient exchange as sorted records instead of remote atomics.
//
// fm_peer_score_scatter pushes five 16-byte vector atomics and a 4-byte OR per occurrence into the owner's
// memory: 192 bytes on the wire for 84 bytes of payload (profiles/r2_peer_nvlink_train.csv), 58 % of what crosses
// NVLink in a step.  Here a rank first sorts its occurrences by (owner, row) (fm_sort_by_row with n_ranks = G, run
// ahead with the parse), scores by pulling rows as before (fm_peer_sorted_score), and then sends ONE record
// {C, B_1..B_K, D, cnt} per run of equal rows, written as whole 16-byte chunks into the owner's inbox at the run's
// sorted position, beside one 4-byte key word per position (row | piece-start flag).  After a barrier every owner
// merges its G inboxes, each sorted by row, tile by tile: a block takes R consecutive rows of the shard, adds the G
// sub-lists that fall into them -- located through a small directory in shared memory, summed in registers in a
// fixed order -- and applies Adagrad to the rows that got something -- reading and writing the table almost sequentially.  No accumulator table, no touched-row bitmap, no
// global float atomics.
//
//   seg_kernel    (sender) where each owner's segment and each of the owner's row tiles start in the sorted keys;
//                 the tile offsets go to the owner's inbox header
//   push_kernel   (sender) a warp tests 32 sorted positions, writes their key words, and its lane groups walk the
//                 pieces that start there (runs cut at FM_SORT_CHUNK, like fm_sorted_update) and write their records
//   merge_kernel  (owner) the tile merge + Adagrad described above (a directory of record positions per tile in shared
//                 memory; the sums are formed in registers)
#include "fm_common.cuh"

namespace fm {
namespace {

constexpr unsigned long long kNoError = 0xffffffffffffffffull;
constexpr int kMaxRanks = 16;
constexpr unsigned kStartBit = 0x80000000u;   // key word: a piece (a record) starts here
constexpr unsigned kMoreBit = 0x40000000u;    // ... and another piece of the same run follows (runs cut at FM_SORT_CHUNK)
constexpr unsigned kRowMask = 0x3fffffffu;
constexpr int kMergeThreads = 256;
constexpr int kTileCells = 2048;             // directory cells (senders x rows) of a merge tile
constexpr unsigned kNone = 0xffffffffu;

static inline int64_t grad_record_ld(int64_t k) { return (k + 3 + 3) / 4 * 4; }
// Shard rows per merge tile, a power of two in 128..1024: a shard of V / G rows receives about as many records
// whatever G is, so the rows that hold one tile's worth of records shrink with G (sender and owner must agree).
static inline int64_t tile_rows(int64_t n_ranks) {
  int64_t r = 1024;
  while (r > 128 && r * n_ranks > kTileCells) r >>= 1;
  return r;
}
static inline int64_t n_tiles_of(int64_t shard_rows, int64_t n_ranks) {
  return (shard_rows + tile_rows(n_ranks) - 1) / tile_rows(n_ranks);
}

// one sender's part of an owner's inbox
struct InboxLayout {
  size_t tile_off;     // u32[tiles + 1]: first position of each tile's rows in this sender's segment (+ the segment's length)
  size_t keys;         // u32[capacity]: row | kStartBit where a piece (a record) starts
  size_t recs;         // f32[capacity][rl]
  size_t stride;       // bytes per sender
};
static inline InboxLayout inbox_layout(int64_t capacity, int64_t shard_rows, int64_t k, int64_t n_ranks) {
  InboxLayout l;
  l.tile_off = 0;
  l.keys = fm_align_up(sizeof(unsigned) * (size_t)(n_tiles_of(shard_rows, n_ranks) + 1), 256);
  l.recs = l.keys + fm_align_up(sizeof(unsigned) * (size_t)capacity, 256);
  l.stride = l.recs + fm_align_up(sizeof(float) * (size_t)capacity * (size_t)grad_record_ld(k), 256);
  return l;
}

struct InboxBases { char* base[kMaxRanks]; };

__device__ __forceinline__ unsigned lower_bound_u32(const unsigned* __restrict__ a, unsigned n, unsigned key) {
  unsigned lo = 0, hi = n;                   // first index with a[i] >= key
  while (lo < hi) {
    const unsigned mid = (lo + hi) >> 1;
    if (__ldg(a + mid) < key) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// ---- sender side ------------------------------------------------------------------------------------------------
struct PushIn {
  const unsigned* keys;      // sorted (owner << row_bits | row)
  const uint2* pairs;        // (example, x bits) per sorted position
  const int64_t* counts;     // counts[0] = n
  const float* rec;          // [batch, rec_ld]: [pg, s_1 .. s_K] per example
  int64_t rec_ld;
};

// thread (o, t): position of the first key of owner o whose row is >= t * R, relative to the owner's segment;
// t = tiles gives the segment's length.  Written into owner o's inbox (this sender's part); seg_start[] stays local.
__global__ void __launch_bounds__(256)
seg_kernel(const unsigned* __restrict__ keys, const int64_t* __restrict__ counts, int n_ranks, int row_bits,
           int64_t tiles, int log2_rows, InboxBases inbox, size_t sender_off, size_t tile_off_at,
           unsigned* __restrict__ seg_start, const unsigned long long* __restrict__ guard) {
  const unsigned n = (guard && *guard != kNoError) ? 0u : (unsigned)counts[0];
  const int64_t per_owner = tiles + 1;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < per_owner * n_ranks;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int o = (int)(i / per_owner);
    const int64_t t = i - (int64_t)o * per_owner;
    const unsigned long long lo_key = (unsigned long long)o << row_bits, hi_key = (unsigned long long)(o + 1) << row_bits;
    unsigned long long want = lo_key + ((unsigned long long)t << log2_rows);
    if (want > hi_key) want = hi_key;
    const unsigned seg0 = lo_key > 0xffffffffull ? n : lower_bound_u32(keys, n, (unsigned)lo_key);
    const unsigned at = want > 0xffffffffull ? n : lower_bound_u32(keys, n, (unsigned)want);
    unsigned* dst = reinterpret_cast<unsigned*>(inbox.base[o] + sender_off + tile_off_at);
    dst[t] = at - seg0;
    if (t == 0) seg_start[o] = seg0;
    if (t == 0 && o == n_ranks - 1) seg_start[n_ranks] = n;
  }
}

// Every sorted position gets its key word in the owner's inbox; the positions that start a piece (a run of equal
// keys, cut at multiples of FM_SORT_CHUNK) also get the piece's record.  gs lanes = the 16-byte chunks of a K+3 wide
// record (at most 32; wider records take T chunks per lane).
template <int T>
__global__ void __launch_bounds__(256)
push_kernel(const PushIn in, int K, int n_ranks, int row_bits, InboxBases inbox, size_t sender_off, size_t keys_at,
            size_t recs_at, int64_t rl, int64_t inbox_capacity, const unsigned* __restrict__ seg_start,
            fm_parse_status* __restrict__ status, const unsigned long long* __restrict__ guard, int gs) {
  if (guard && *guard != kNoError) return;
  const unsigned n = (unsigned)in.counts[0];
  const int lane = threadIdx.x & 31;
  const int per_warp = 32 / gs;
  const int g = lane / gs, c = lane - g * gs;
  const bool lane_on = g < per_warp;
  const int vc_rec = (K + 4) >> 2, vc_acc = (K + 6) >> 2;       // chunks of an example record / of a gradient record
  const unsigned row_mask = row_bits >= 32 ? 0xffffffffu : ((1u << row_bits) - 1u);
  const int64_t warps_total = (int64_t)gridDim.x * (blockDim.x >> 5);
  const int64_t warp_id = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  for (int64_t base = warp_id * 32; base < n; base += warps_total * 32) {
    const int64_t mine = base + lane;
    bool opens = false;
    uint2 my_pair = make_uint2(0u, 0u);
    if (mine < n) {
      const unsigned k0 = __ldg(in.keys + mine);
      my_pair = __ldg(in.pairs + mine);                 // one coalesced read for the warp's 32 positions
      opens = mine == 0 || __ldg(in.keys + mine - 1) != k0 || (mine % FM_SORT_CHUNK) == 0;
      const int64_t cut = (mine / FM_SORT_CHUNK + 1) * FM_SORT_CHUNK;      // where this piece is cut, if its run goes on
      const bool more = opens && cut < n && __ldg(in.keys + cut) == k0;
      const int o = row_bits >= 32 ? 0 : (int)(k0 >> row_bits);
      const int64_t slot = mine - seg_start[o];
      if (slot < inbox_capacity)
        reinterpret_cast<unsigned*>(inbox.base[o] + sender_off + keys_at)[slot] =
            (k0 & row_mask) | (opens ? kStartBit : 0u) | (more ? kMoreBit : 0u);
      else   // the owner's inbox is smaller than this segment: the step is void (reported like a capacity error)
        atomicMin(reinterpret_cast<unsigned long long*>(&status->first_error), (unsigned long long)FM_PARSE_E_CAPACITY);
    }
    const unsigned opening = __ballot_sync(kFull, opens);
    const int n_open = __popc(opening);
    for (int taken = 0; taken < n_open; taken += per_warp) {       // warp-uniform
      const bool busy = lane_on && taken + g < n_open;
      const int src = busy ? (int)__fns(opening, 0, taken + g + 1) : 0;
      const unsigned first_ex = __shfl_sync(kFull, my_pair.x, src), first_x = __shfl_sync(kFull, my_pair.y, src);
      if (!busy) continue;                                         // no warp-wide operation below
      const unsigned p0 = (unsigned)(base + src);
      const unsigned key = __ldg(in.keys + p0);
      unsigned p1 = p0 + 1;
      while (p1 < n && (p1 % FM_SORT_CHUNK) != 0 && __ldg(in.keys + p1) == key) ++p1;
      float4 G[T];
#pragma unroll
      for (int t = 0; t < T; ++t) G[t] = make_float4(0.f, 0.f, 0.f, 0.f);
      float C = 0.f, D = 0.f, cnt = 0.f;
      for (unsigned p = p0; p < p1; ++p) {
        const uint2 pr = p == p0 ? make_uint2(first_ex, first_x) : __ldg(in.pairs + p);
        const float x = __uint_as_float(pr.y);
        const float* r = in.rec + (int64_t)pr.x * in.rec_ld;
        const float pgx = __ldg(r) * x;
        C += pgx;
        D = fmaf(pgx, x, D);
        cnt += 1.0f;
#pragma unroll
        for (int t = 0; t < T; ++t) {
          const int chunk = c + t * gs;
          if (chunk < vc_rec) {
            const float4 s = __ldg(reinterpret_cast<const float4*>(r + 4 * chunk));
            G[t].x = fmaf(pgx, s.x, G[t].x); G[t].y = fmaf(pgx, s.y, G[t].y);
            G[t].z = fmaf(pgx, s.z, G[t].z); G[t].w = fmaf(pgx, s.w, G[t].w);
          }
        }
      }
      const int o = row_bits >= 32 ? 0 : (int)(key >> row_bits);
      const int64_t slot = (int64_t)p0 - seg_start[o];
      if (slot >= inbox_capacity) continue;
      float* dst = reinterpret_cast<float*>(inbox.base[o] + sender_off + recs_at) + slot * rl;
#pragma unroll
      for (int t = 0; t < T; ++t) {
        const int chunk = c + t * gs;
        if (chunk >= vc_acc) continue;
        float o4[4] = {G[t].x, G[t].y, G[t].z, G[t].w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int col = 4 * chunk + e;
          if (col == 0) o4[e] = C;                       // the record's column 0 held pg: its sum is not a B_k
          else if (col == K + 1) o4[e] = D;
          else if (col == K + 2) o4[e] = cnt;
          else if (col > K + 2 || chunk >= vc_rec) o4[e] = 0.f;
        }
        *reinterpret_cast<float4*>(dst + 4 * chunk) = make_float4(o4[0], o4[1], o4[2], o4[3]);
      }
    }
  }
  __threadfence_system();      // the records are in the owners' memory before this kernel counts as done
}

// ---- owner side -------------------------------------------------------------------------------------------------
// One block per tile of consecutive shard rows.  A thread per inbox position of the tile (all senders' sub-lists in
// one flat range) enters the position of every row's first record into a directory in shared memory
// (dir[sender][row]), marks the sender in the row's mask and, when it is the row's first, appends the row to the
// tile's list.  Then one lane group per LISTED row -- rows that got nothing cost nothing -- gathers the row's
// records (at most one run per sender, its pieces in order, senders in rank order: the same sum every time, no float
// atomics) and applies Adagrad (G_0 = C + cnt rg lb b, G_k = v_k (cnt rg lf - D) + B_k).
template <int T>
__global__ void __launch_bounds__(kMergeThreads)
merge_kernel(float* __restrict__ table, float* __restrict__ accum, int64_t ld, int64_t shard_rows,
             const char* __restrict__ inbox, InboxLayout lay, int64_t rl, int64_t inbox_capacity, int n_ranks, int K,
             int64_t tiles, int log2_rows, float lr, float rf, float rb, const long long* __restrict__ flag, int gs,
             int early) {
  extern __shared__ __align__(16) unsigned dir[];      // [n_ranks][rows]: position of the row's first record, or kNone
  __shared__ unsigned s_a[kMaxRanks], s_b[kMaxRanks];
  __shared__ unsigned s_total, s_listed;
  if (flag && *flag >= 0) return;                      // some rank's batch did not parse: the step is void everywhere
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int rows = 1 << log2_rows;
  unsigned* who = dir + n_ranks * rows;                // [rows]: bit s = sender s has a record for the row
  unsigned short* list = reinterpret_cast<unsigned short*>(who + rows);      // the rows with a record, any order
  const int per_warp = 32 / gs;
  const int g = lane / gs, c = lane - g * gs;
  const bool lane_on = g < per_warp;
  const int vc_table = (K + 4) >> 2, vc_acc = (K + 6) >> 2;
  const int cnt_chunk = (K + 2) >> 2, cnt_comp = (K + 2) & 3, d_chunk = (K + 1) >> 2, d_comp = (K + 1) & 3;
  const unsigned gmask = gs >= 32 ? kFull : (((1u << gs) - 1u) << (g * gs));
  const int groups = (kMergeThreads / 32) * per_warp;
  const int my_group = warp * per_warp + g;
  for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    __syncthreads();                                   // the previous tile's list has been read
    if (tid == 0) { s_total = 0; s_listed = 0; }
    __syncthreads();
    if (tid < n_ranks) {
      const unsigned* off = reinterpret_cast<const unsigned*>(inbox + (size_t)tid * lay.stride + lay.tile_off);
      unsigned a = off[tile], b = off[tile + 1];
      if (b > (unsigned)inbox_capacity) b = (unsigned)inbox_capacity;
      if (a > b) a = b;
      s_a[tid] = a; s_b[tid] = b;
      if (b > a) atomicAdd(&s_total, b - a);
    }
    __syncthreads();
    if (s_total == 0) continue;                        // block-uniform: nothing for these rows this step
    for (int i = tid; i < n_ranks * rows; i += kMergeThreads) dir[i] = kNone;
    for (int i = tid; i < rows; i += kMergeThreads) who[i] = 0u;
    __syncthreads();
    const int64_t row0 = tile << log2_rows;
    for (unsigned i = tid; i < s_total; i += kMergeThreads) {
      unsigned j = i;
      int s = 0;
      while (j >= s_b[s] - s_a[s]) { j -= s_b[s] - s_a[s]; ++s; }       // which sender's sub-list (at most G steps)
      const unsigned p = s_a[s] + j;
      const unsigned kw = __ldg(reinterpret_cast<const unsigned*>(inbox + (size_t)s * lay.stride + lay.keys) + p);
      if (!(kw & kStartBit)) continue;
      const int64_t r = (int64_t)(kw & kRowMask) - row0;
      if (r < 0 || r >= rows || row0 + r >= shard_rows) continue;
      // a cut run has several pieces: the directory keeps the first
      atomicMin(&dir[s * rows + r], (p << 1) | ((kw & kMoreBit) ? 1u : 0u));
      if (atomicOr(&who[r], 1u << s) == 0u) list[atomicAdd(&s_listed, 1u)] = (unsigned short)r;
    }
    __syncthreads();
    const int listed = (int)s_listed;
    for (int i = my_group; lane_on && i < listed; i += groups) {
      const int r = list[i];
      const int64_t row = (row0 + r) * ld;
      float4 w4[T], h4[T];
      auto load_row = [&]() {
#pragma unroll
        for (int t = 0; t < T; ++t) {
          const int chunk = c + t * gs;
          if (chunk < vc_table) {
            w4[t] = *reinterpret_cast<const float4*>(table + row + 4 * chunk);
            h4[t] = *reinterpret_cast<const float4*>(accum + row + 4 * chunk);
          }
        }
      };
      if (early) load_row();                           // in flight while the row's records are gathered
      float4 G[T];
#pragma unroll
      for (int t = 0; t < T; ++t) G[t] = make_float4(0.f, 0.f, 0.f, 0.f);
      auto add = [&](const float4 (&v)[T]) {
#pragma unroll
        for (int t = 0; t < T; ++t)
          if (c + t * gs < vc_acc) { G[t].x += v[t].x; G[t].y += v[t].y; G[t].z += v[t].z; G[t].w += v[t].w; }
      };
      auto fetch = [&](int s, unsigned pos, float4 (&v)[T]) {
        const float* recs = reinterpret_cast<const float*>(inbox + (size_t)s * lay.stride + lay.recs) + (size_t)pos * rl;
#pragma unroll
        for (int t = 0; t < T; ++t)
          if (c + t * gs < vc_acc) v[t] = __ldg(reinterpret_cast<const float4*>(recs + 4 * (c + t * gs)));
      };
      // a cut run's further pieces follow at the next piece starts with the same row
      auto rest = [&](int s, unsigned first) {
        const unsigned* keys = reinterpret_cast<const unsigned*>(inbox + (size_t)s * lay.stride + lay.keys);
        const unsigned want = (unsigned)(row0 + r), end = s_b[s];
        bool more = true;
        for (unsigned q = first + 1; more && q < end; ++q) {
          const unsigned kw = __ldg(keys + q);
          if ((kw & kRowMask) != want) break;
          if (!(kw & kStartBit)) continue;
          float4 x[T];
          fetch(s, q, x);
          add(x);
          more = (kw & kMoreBit) != 0u;
        }
      };
      unsigned senders = who[r];
      while (senders) {                                // two senders' records in flight at a time
        const int s0 = __ffs(senders) - 1;
        senders &= senders - 1u;
        const int s1 = senders ? __ffs(senders) - 1 : -1;
        senders &= senders - 1u;                       // 0 stays 0
        const unsigned e0 = dir[s0 * rows + r], e1 = s1 >= 0 ? dir[s1 * rows + r] : 0u;
        float4 v0[T], v1[T];
        fetch(s0, e0 >> 1, v0);
        if (s1 >= 0) fetch(s1, e1 >> 1, v1);
        add(v0);
        if (e0 & 1u) rest(s0, e0 >> 1);
        if (s1 >= 0) {
          add(v1);
          if (e1 & 1u) rest(s1, e1 >> 1);
        }
      }
      if (!early) load_row();
      // D and cnt sit in one lane of the group: hand them to the others
      float D = 0.f, cnt = 0.f;
#pragma unroll
      for (int t = 0; t < T; ++t) {
        const int chunk = c + t * gs;
        const float gv[4] = {G[t].x, G[t].y, G[t].z, G[t].w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          if (chunk == cnt_chunk && e == cnt_comp) cnt = gv[e];
          if (chunk == d_chunk && e == d_comp) D = gv[e];
        }
      }
      cnt = __shfl_sync(gmask, cnt, g * gs + cnt_chunk % gs);
      D = __shfl_sync(gmask, D, g * gs + d_chunk % gs);
      const float fac_scale = cnt * rf - D, bias_scale = cnt * rb;
#pragma unroll
      for (int t = 0; t < T; ++t) {
        const int chunk = c + t * gs;
        if (chunk >= vc_table) continue;
        const float av[4] = {G[t].x, G[t].y, G[t].z, G[t].w};
        float wv[4] = {w4[t].x, w4[t].y, w4[t].z, w4[t].w};
        float hv[4] = {h4[t].x, h4[t].y, h4[t].z, h4[t].w};
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
    }
  }
}

static inline int ilog2(int64_t v) { int b = 0; while ((1ll << (b + 1)) <= v) ++b; return b; }

}  // namespace
}  // namespace fm

using namespace fm;

FM_EXPORT int64_t fm_merge_tile_rows(int64_t n_ranks) { return n_ranks < 1 ? 0 : tile_rows(n_ranks); }

FM_EXPORT size_t fm_merge_inbox_bytes(int64_t capacity, int64_t shard_rows, int64_t factor_num, int64_t n_ranks) {
  if (capacity < 0 || shard_rows < 0 || factor_num < 0 || n_ranks < 1) return 0;
  return inbox_layout(capacity, shard_rows, factor_num, n_ranks).stride * (size_t)n_ranks;
}

FM_EXPORT size_t fm_merge_push_workspace(int64_t n_ranks) { return fm_align_up(sizeof(unsigned) * (size_t)(n_ranks + 1), 256); }

FM_EXPORT int fm_peer_push_records(const uint32_t* sorted_keys, const uint32_t* sorted_pairs, const int64_t* counts,
                                   int64_t capacity, const float* records, int64_t record_ld, int64_t factor_num,
                                   int64_t rank, int64_t n_ranks, int64_t vocab_size, int64_t shard_rows,
                                   void* const* inbox_bases, int64_t inbox_capacity, fm_parse_status* status,
                                   void* workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  FM_REQUIRE(capacity >= 0 && factor_num >= 0 && vocab_size >= 1 && shard_rows >= 1 && inbox_capacity >= 0,
             "fm_peer_push_records: bad shape");
  if (n_ranks < 1 || n_ranks > kMaxRanks) return fm_fail(FM_E_LIMIT, "fm_peer_push_records: 1..%d ranks", kMaxRanks);
  FM_REQUIRE(rank >= 0 && rank < n_ranks, "fm_peer_push_records: bad rank");
  if (shard_rows > (1ll << 30) || inbox_capacity >= (1ll << 30))     // key words keep two flag bits above the row
    return fm_fail(FM_E_LIMIT, "fm_peer_push_records: at most 2^30 rows per shard and positions per inbox");
  FM_REQUIRE(sorted_keys && sorted_pairs && counts && records && inbox_bases && status && workspace,
             "fm_peer_push_records: null pointer");
  FM_REQUIRE(record_ld >= (factor_num + 4) / 4 * 4 && record_ld % 4 == 0, "fm_peer_push_records: bad record_ld");
  FM_REQUIRE(((uintptr_t)records & 15) == 0 && ((uintptr_t)sorted_pairs & 7) == 0, "fm_peer_push_records: alignment");
  if (workspace_bytes < fm_merge_push_workspace(n_ranks))
    return fm_fail(FM_E_WORKSPACE, "fm_peer_push_records: workspace %zu < %zu", workspace_bytes, fm_merge_push_workspace(n_ranks));
  const int64_t chunks = (factor_num + 6) / 4;
  const int gs = chunks <= 32 ? (int)chunks : 32;
  const int t = (int)((chunks + gs - 1) / gs);
  if (t > 3) return fm_fail(FM_E_LIMIT, "fm_peer_push_records: factor_num %lld too wide", (long long)factor_num);
  // the key layout of fm_sort_by_row(n_ranks): owner << row_bits | row
  const int64_t rows_max = (vocab_size + n_ranks - 1) / n_ranks;
  int row_bits = 0;
  if (n_ranks > 1) { uint64_t v = (uint64_t)(rows_max > 1 ? rows_max - 1 : 1); while (v) { ++row_bits; v >>= 1; } }
  else row_bits = 32;     // one rank: the key is the row
  const InboxLayout lay = inbox_layout(inbox_capacity, shard_rows, factor_num, n_ranks);
  InboxBases bases;
  for (int r = 0; r < kMaxRanks; ++r) {
    bases.base[r] = r < n_ranks ? static_cast<char*>(inbox_bases[r]) : nullptr;
    FM_REQUIRE(r >= n_ranks || (bases.base[r] && ((uintptr_t)bases.base[r] & 255) == 0), "fm_peer_push_records: bad inbox base");
  }
  const int64_t tiles = n_tiles_of(shard_rows, n_ranks);
  const int log2_rows = ilog2(tile_rows(n_ranks));
  const unsigned long long* g = reinterpret_cast<const unsigned long long*>(&status->first_error);
  unsigned* seg_start = static_cast<unsigned*>(workspace);
  const size_t sender_off = lay.stride * (size_t)rank;
  int64_t seg_blocks = ((tiles + 1) * n_ranks + 255) / 256;
  if (seg_blocks > 148 * 8) seg_blocks = 148 * 8;
  seg_kernel<<<(unsigned)seg_blocks, 256, 0, stream>>>(sorted_keys, counts, (int)n_ranks, n_ranks > 1 ? row_bits : 32,
                                                       tiles, log2_rows, bases, sender_off, lay.tile_off, seg_start, g);
  FM_LAUNCH_CHECK();
  if (capacity == 0) return FM_OK;
  PushIn in{sorted_keys, reinterpret_cast<const uint2*>(sorted_pairs), counts, records, record_ld};
  int64_t blocks = (capacity + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
#define CALL(T)                                                                                                        \
  push_kernel<T><<<(unsigned)blocks, 256, 0, stream>>>(in, (int)factor_num, (int)n_ranks, n_ranks > 1 ? row_bits : 32, \
                                                       bases, sender_off, lay.keys, lay.recs, grad_record_ld(factor_num), \
                                                       inbox_capacity, seg_start, status, g, gs)
  if (t == 1) { CALL(1); } else if (t == 2) { CALL(2); } else { CALL(3); }
#undef CALL
  FM_LAUNCH_CHECK();
  return FM_OK;
}

FM_EXPORT int fm_peer_merge_update(float* table, float* accum, int64_t table_ld, int64_t shard_rows, const void* inbox,
                                   int64_t inbox_capacity, int64_t factor_num, int64_t n_ranks, float learning_rate,
                                   float factor_lambda, float bias_lambda, float reg_grad, const int64_t* flag,
                                   void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  FM_REQUIRE(factor_num >= 0 && shard_rows >= 1 && inbox_capacity >= 0, "fm_peer_merge_update: bad shape");
  if (shard_rows > (1ll << 30) || inbox_capacity >= (1ll << 30))
    return fm_fail(FM_E_LIMIT, "fm_peer_merge_update: at most 2^30 rows per shard and positions per inbox");
  if (n_ranks < 1 || n_ranks > kMaxRanks) return fm_fail(FM_E_LIMIT, "fm_peer_merge_update: 1..%d ranks", kMaxRanks);
  FM_REQUIRE(table_ld >= factor_num + 1 && table_ld % 4 == 0, "fm_peer_merge_update: table_ld must be a multiple of 4 and >= K+1");
  FM_REQUIRE(table && accum && inbox, "fm_peer_merge_update: null pointer");
  FM_REQUIRE((((uintptr_t)table | (uintptr_t)accum) & 15) == 0 && ((uintptr_t)inbox & 255) == 0, "fm_peer_merge_update: alignment");
  const int64_t chunks = (factor_num + 6) / 4;
  const int gs = chunks <= 32 ? (int)chunks : 32;
  const int t = (int)((chunks + gs - 1) / gs);
  if (t > 3) return fm_fail(FM_E_LIMIT, "fm_peer_merge_update: factor_num %lld too wide", (long long)factor_num);
  const InboxLayout lay = inbox_layout(inbox_capacity, shard_rows, factor_num, n_ranks);
  const int64_t rows = tile_rows(n_ranks), tiles = n_tiles_of(shard_rows, n_ranks), rl = grad_record_ld(factor_num);
  // the directory (at most 8 KiB), who sent something per row, the list of rows that got something
  const size_t smem = (size_t)rows * ((size_t)n_ranks + 1) * sizeof(unsigned) + (size_t)rows * sizeof(unsigned short);
  static const int early = fm_env_int("FM_MERGE_EARLY_ROWS", 0);     // row loads before (1) / after (0) the gather
  static const int per_sm = fm_env_int("FM_MERGE_BLOCKS_PER_SM", 8);
  int64_t blocks = tiles < 148 * (int64_t)per_sm ? tiles : 148 * (int64_t)per_sm;
#define CALL(T)                                                                                                      \
  {                                                                                                                  \
    merge_kernel<T><<<(unsigned)blocks, kMergeThreads, smem, stream>>>(                                              \
        table, accum, table_ld, shard_rows, static_cast<const char*>(inbox), lay, rl, inbox_capacity, (int)n_ranks,  \
        (int)factor_num, tiles, ilog2(rows), learning_rate, reg_grad * factor_lambda, reg_grad * bias_lambda,        \
        reinterpret_cast<const long long*>(flag), gs, early);                                                        \
  }
  if (t == 1) CALL(1) else if (t == 2) CALL(2) else CALL(3)
#undef CALL
  FM_LAUNCH_CHECK();
  return FM_OK;
}
