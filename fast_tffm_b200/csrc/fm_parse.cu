// fm_parse.cu -- FmParser on the GPU.
//
// Replaces cc/fm_parser_op.cc:29-51 (Compute: a serial loop over the batch) and
// :58-109 (ParseLine) plus the tf.cumsum at tffm/fm_model.py:73.
//
// Structure (one warp per line, text read as 16-byte words, 512 bytes per
// warp iteration):
//   pass 1 (count): count the ' ' bytes of a line and prove it "regular"
//       (no NUL / \t\n\v\f\r, no space at the start, after ':' or after a
//       space).  In a regular line every loop iteration of ParseLine eats
//       exactly one space, so sizes[i] = spaces - (trailing space ? 1 : 0).
//       Any other line is counted by the sequential scanner, which follows
//       ParseLine statement by statement.
//   scan: feature_poses = [0, cumsum(sizes)], single pass.
//   pass 2 (fill): the spaces of a 512-byte chunk are compacted to token
//       starts in shared memory; each lane parses whole tokens (id[:val]).  A
//       lane that meets anything but the plain grammar raises a flag and lane 0
//       re-parses the line sequentially: error reporting and every libc quirk
//       (whitespace skipping, "inf", hex floats ...) live in that one place.
// Numbers are converted by fm_numparse.cuh (bit-exact with glibc).
#include "fm_common.cuh"
#include "fm_numparse.cuh"
#include "fm_scan.cuh"

namespace fm {
namespace {

constexpr int kWarpsPerBlock = 8;
constexpr int kChunk = 512;  // bytes per warp iteration

__device__ __forceinline__ void report_error(fm_parse_status* st, int64_t line, int kind) {
  atomicMin(reinterpret_cast<unsigned long long*>(&st->first_error),
            ((unsigned long long)line << 3) | (unsigned long long)kind);
}

// ParseLine, cc/fm_parser_op.cc:58-109, statement by statement.  ids == nullptr
// => count only.  p[0..n) is the line; at() supplies the terminating NUL.
__device__ __noinline__ int parse_line_seq(const uint8_t* p, int64_t n, int hash, int64_t vocab,
                                           float* label, int64_t* ids, float* vals, int64_t cap,
                                           int32_t* cnt_out) {
  uint32_t bits;
  int64_t c = fmnp::strtof_bits(p, n, &bits);                       // :69
  if (c == 0) return FM_PARSE_E_LABEL;                              // :70
  if (label) *label = __uint_as_float(bits);
  int64_t i = c;
  int32_t cnt = 0;
  for (; fmnp::at(p, n, i) != 0; ++cnt) {                           // :75
    if (fmnp::at(p, n, i) != ' ') return FM_PARSE_E_FORMAT;         // :76
    ++i;
    if (fmnp::at(p, n, i) == 0) break;                              // :79-81
    int64_t id;
    if (hash) {                                                     // :82-84
      int64_t j = i;
      for (uint8_t ch = fmnp::at(p, n, j); ch != ' ' && ch != ':' && ch != 0; ch = fmnp::at(p, n, ++j)) {}
      id = (int64_t)(fmnp::hash64(p + i, j - i) % (uint64_t)vocab);
      i = j;
    } else {                                                        // :86-93
      c = fmnp::strtoll10(p + i, n - i, &id);
      if (c == 0) return FM_PARSE_E_FORMAT;
      if (!(id >= 0 && id < vocab)) return FM_PARSE_E_RANGE;
      i += c;
    }
    float fv = 1.0f;                                                // :103
    if (fmnp::at(p, n, i) == ':') {                                 // :95-101
      ++i;
      c = fmnp::strtof_bits(p + i, n - i, &bits);
      if (c == 0) return FM_PARSE_E_VALUE;
      i += c;
      fv = __uint_as_float(bits);
    }
    if (ids) {
      if (cnt >= cap) return FM_PARSE_E_CAPACITY;
      ids[cnt] = id;                                                // :105-106
      vals[cnt] = fv;
    }
  }
  *cnt_out = cnt;                                                   // :108
  return FM_PARSE_OK;
}

__device__ __forceinline__ uint4 load16(const uint8_t* p) {
  return __ldg(reinterpret_cast<const uint4*>(p));
}

// 16-bit masks over the 16 bytes of v: bytes equal to `c`
__device__ __forceinline__ uint32_t eq_mask(const uint4& v, uint32_t c) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
  uint32_t m = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    // per-byte equality via the XOR + "has zero byte" trick, exact variant
    const uint32_t x = w[k] ^ (c * 0x01010101u);
    const uint32_t z = ~(((x & 0x7f7f7f7fu) + 0x7f7f7f7fu) | x | 0x7f7f7f7fu);  // 0x80 where byte == 0
    m |= (((z >> 7) & 1u) | ((z >> 14) & 2u) | ((z >> 21) & 4u) | ((z >> 28) & 8u)) << (4 * k);
  }
  return m;
}
// bytes in [lo, hi], hi < 128
__device__ __forceinline__ uint32_t range_mask(const uint4& v, uint32_t lo, uint32_t hi) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
  uint32_t m = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const uint32_t byte = (w[k] >> (8 * b)) & 0xffu;
      m |= (uint32_t)(byte - lo <= hi - lo) << (4 * k + b);
    }
  }
  return m;
}

__global__ void __launch_bounds__(kWarpsPerBlock * 32)
parse_count_kernel(const uint8_t* __restrict__ text, const int64_t* __restrict__ off, int64_t n_lines,
                   int64_t vocab, int hash, int32_t* __restrict__ sizes, fm_parse_status* st) {
  const int lane = threadIdx.x & 31;
  const int64_t line = (int64_t)blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
  if (line >= n_lines) return;
  const int64_t s = off[line], e = off[line + 1] - 1;
  int spaces = 0;
  bool irregular = false, trail = false;
  // carry: was the byte just before this lane's first byte a space / ':' (or the line start)?
  uint32_t carry_iter = 1;  // virtual byte before the line start counts as a space
  for (int64_t pos = s & ~(int64_t)15; pos < e; pos += kChunk) {
    const int64_t p = pos + lane * 16;
    uint32_t valid = 0;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (p < e) {
      v = load16(text + p);
      const int64_t lo = s > p ? s - p : 0, hi = e - p < 16 ? e - p : 16;  // valid bytes [lo, hi)
      valid = lo < 16 ? ((0xffffu >> (16 - hi)) & (0xffffu << lo)) & 0xffffu : 0;
    }
    const uint32_t sp = eq_mask(v, ' ') & valid;
    const uint32_t sc = (sp | eq_mask(v, ':')) & valid;
    const uint32_t bad = (eq_mask(v, 0) | range_mask(v, 9, 13)) & valid;
    // previous-byte mask: bit t set when byte t-1 is a space / ':' / the line start
    uint32_t prev_last = __shfl_up_sync(kFull, (sc >> 15) & 1u, 1);
    if (lane == 0) prev_last = carry_iter;
    uint32_t prevm = ((sc << 1) | prev_last) & 0xffffu;
    if (s > p) prevm |= 1u << (s - p);  // line starts inside this word (lane 0 of the first chunk only)
    irregular |= (sp & prevm) != 0 || bad != 0;
    spaces += __popc(sp);
    if (p <= e - 1 && e - 1 < p + 16) trail = (sp >> (e - 1 - p)) & 1u;
    carry_iter = __shfl_sync(kFull, (sc >> 15) & 1u, 31);
  }
  spaces = warp_sum(spaces);
  irregular = __any_sync(kFull, irregular);
  trail = __any_sync(kFull, trail);
  if (lane == 0) {
    int32_t cnt = spaces - (trail ? 1 : 0);
    if (irregular) {
      cnt = 0;
      const int rc = parse_line_seq(text + s, e - s, hash, vocab, nullptr, nullptr, nullptr, 0, &cnt);
      if (rc != FM_PARSE_OK) { report_error(st, line, rc); cnt = 0; }
    }
    sizes[line] = cnt;
  }
}

struct SizesIn {
  const int32_t* sizes; int64_t n;
  __device__ int64_t count() const { return n; }
  __device__ long long operator()(int64_t i) const { return sizes[i]; }
};
struct PosesOut {
  int32_t* poses; int64_t n; fm_parse_status* st;
  __device__ void operator()(int64_t i, long long excl, long long) const { poses[i] = (int32_t)excl; }
  __device__ void total(long long t) const { poses[n] = (int32_t)t; st->nnz = t; }
};

// ---- pass 2 ---------------------------------------------------------------

// [+-]digits[.digits][e[+-]digits] through the cheap exact tiers; anything else
// (or a tier miss) goes through the full glibc-compatible scanner.  Returns the
// bytes consumed (0 = no conversion).  Leading whitespace is NOT accepted here:
// strtof would skip it, which breaks the one-space-per-token invariant.
__device__ __forceinline__ int64_t scan_value(const uint8_t* __restrict__ p, int64_t n, uint32_t* bits) {
  int64_t i = 0;
  uint8_t c = fmnp::at(p, n, 0);
  if (fmnp::is_space(c)) return 0;
  uint32_t sign = 0;
  if (c == '-') { sign = 0x80000000u; c = fmnp::at(p, n, ++i); } else if (c == '+') { c = fmnp::at(p, n, ++i); }
  uint64_t w = 0;
  int nd = 0, frac = 0;
  bool any = false, simple = true;
  for (; fmnp::is_digit(c); c = fmnp::at(p, n, ++i)) { w = w * 10 + (c - '0'); nd += (w != 0); any = true; }
  if (c == '.') {
    c = fmnp::at(p, n, ++i);
    for (; fmnp::is_digit(c); c = fmnp::at(p, n, ++i)) { w = w * 10 + (c - '0'); nd += (w != 0); ++frac; any = true; }
  }
  if (!any || nd > 19) simple = false;
  int64_t ex = 0;
  if (simple && (c | 0x20) == 'e') {
    int64_t k = i + 1;
    uint8_t d = fmnp::at(p, n, k);
    bool eneg = false;
    if (d == '-') { eneg = true; d = fmnp::at(p, n, ++k); } else if (d == '+') { d = fmnp::at(p, n, ++k); }
    if (fmnp::is_digit(d)) {
      for (; fmnp::is_digit(d); d = fmnp::at(p, n, ++k)) if (ex < 100000) ex = ex * 10 + (d - '0');
      if (eneg) ex = -ex;
      i = k;
    }
  }
  uint32_t r;
  if (simple && fmnp::dec_to_f32_fast(w, ex - frac, &r)) { *bits = sign | r; return i; }
  return fmnp::strtof_bits(p, n, bits);
}

__global__ void __launch_bounds__(kWarpsPerBlock * 32)
parse_fill_kernel(const uint8_t* __restrict__ text, const int64_t* __restrict__ off, int64_t n_lines,
                  int64_t vocab, int hash, const int32_t* __restrict__ poses, float* __restrict__ labels,
                  int64_t* __restrict__ ids, float* __restrict__ vals, int64_t capacity, fm_parse_status* st) {
  __shared__ uint16_t s_tok[kWarpsPerBlock][kChunk];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t line = (int64_t)blockIdx.x * kWarpsPerBlock + warp;
  if (line >= n_lines) return;
  const int64_t s = off[line], e = off[line + 1] - 1;
  const int64_t out_base = poses[line];
  const int32_t cnt = poses[line + 1] - poses[line];
  bool fail = out_base + cnt > capacity;
  const bool over_capacity = fail;

  if (lane == 0 && !fail) {  // the label: everything up to the first space
    uint32_t bits = 0;
    const int64_t c = scan_value(text + s, e - s, &bits);
    if (c == 0 || !(s + c == e || text[s + c] == ' ')) fail = true;
    else labels[line] = __uint_as_float(bits);
  }
  fail = __any_sync(kFull, fail);

  int32_t ord_base = 0;
  for (int64_t pos = s & ~(int64_t)15; pos < e && !fail; pos += kChunk) {
    const int64_t p = pos + lane * 16;
    uint32_t sp = 0, nul = 0;
    if (p < e) {
      const uint4 v = load16(text + p);
      const int64_t lo = s > p ? s - p : 0, hi = e - p < 16 ? e - p : 16;
      const uint32_t valid = lo < 16 ? ((0xffffu >> (16 - hi)) & (0xffffu << lo)) & 0xffffu : 0;
      sp = eq_mask(v, ' ') & valid;
      nul = eq_mask(v, 0) & valid;
    }
    // compact the token starts (byte after each space) of this chunk
    const int c = __popc(sp);
    int incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(kFull, incl, o);
      if (lane >= o) incl += t;
    }
    const int tot = __shfl_sync(kFull, incl, 31);
    int w = incl - c;
    for (uint32_t m = sp; m; m &= m - 1) s_tok[warp][w++] = (uint16_t)(lane * 16 + __ffs(m));
    __syncwarp();
    bool bad = nul != 0;
    for (int t = lane; t < tot; t += 32) {
      int64_t q = pos + s_tok[warp][t];
      const int32_t ordinal = ord_base + t;
      if (q >= e) { bad |= ordinal != cnt; continue; }  // trailing space: breaks without counting (:79-81)
      if (ordinal >= cnt) { bad = true; continue; }
      int64_t id = 0;
      if (hash) {
        const int64_t a = q;
        uint8_t ch = text[q];
        while (ch != ':' && ch != ' ') { ++q; if (q >= e) break; ch = text[q]; }
        if (q == a) { bad = true; continue; }  // empty token: leave it to the sequential path
        id = (int64_t)(fmnp::hash64(text + a, q - a) % (uint64_t)vocab);
      } else {
        int nd = 0;
        uint64_t acc = 0;
        for (; q < e && fmnp::is_digit(text[q]) && nd < 18; ++q, ++nd) acc = acc * 10 + (text[q] - '0');
        if (nd == 0 || (q < e && fmnp::is_digit(text[q])) || (int64_t)acc >= vocab) { bad = true; continue; }
        id = (int64_t)acc;
      }
      uint32_t bits = 0x3f800000u;  // 1.0f
      if (q < e && text[q] == ':') {
        ++q;
        const int64_t c2 = scan_value(text + q, e - q, &bits);
        if (c2 == 0) { bad = true; continue; }
        q += c2;
      }
      if (!(q == e || text[q] == ' ')) { bad = true; continue; }
      ids[out_base + ordinal] = id;
      vals[out_base + ordinal] = __uint_as_float(bits);
    }
    ord_base += tot;
    fail = __any_sync(kFull, bad);
    __syncwarp();
  }
  // a line that ended short of its count (cannot happen for a regular line)
  if (!fail && ord_base != cnt && ord_base != cnt + 1) fail = true;

  if (fail && lane == 0) {
    if (over_capacity) {
      report_error(st, line, FM_PARSE_E_CAPACITY);
    } else {
      int32_t got = 0;
      const int rc = parse_line_seq(text + s, e - s, hash, vocab, labels + line, ids + out_base,
                                    vals + out_base, cnt, &got);
      if (rc != FM_PARSE_OK) report_error(st, line, rc);
      else if (got != cnt) report_error(st, line, FM_PARSE_E_FORMAT);
    }
  }
}

// ---- line splitting ---------------------------------------------------------
struct NewlineIn {
  const uint8_t* text; int64_t nbytes;
  __device__ int64_t count() const { return (nbytes + 15) / 16; }
  __device__ uint32_t mask(int64_t g) const {
    const uint4 v = load16(text + g * 16);
    const int64_t left = nbytes - g * 16;
    return eq_mask(v, '\n') & (left >= 16 ? 0xffffu : (0xffffu >> (16 - left)));
  }
  __device__ long long operator()(int64_t g) const { return __popc(mask(g)); }
};
struct LineStartsOut {
  NewlineIn in; int64_t* offsets; int64_t max_lines; fm_parse_status* st;
  __device__ void operator()(int64_t g, long long excl, long long cnt) const {
    if (g == 0) offsets[0] = 0;
    if (cnt == 0) return;
    long long r = excl;
    for (uint32_t m = in.mask(g); m; m &= m - 1, ++r)
      if (r + 1 <= max_lines) offsets[r + 1] = g * 16 + __ffs(m);
  }
  __device__ void total(long long t) const {
    st->n_lines = t;
    if (t > max_lines) report_error(st, max_lines, FM_PARSE_E_CAPACITY);
    else if (in.nbytes > 0 && in.text[in.nbytes - 1] != '\n') report_error(st, t, FM_PARSE_E_FORMAT);
  }
};

}  // namespace
}  // namespace fm

using namespace fm;

FM_EXPORT size_t fm_split_lines_workspace(int64_t text_bytes) {
  return scan_scratch_bytes((text_bytes + 15) / 16, 4);
}

FM_EXPORT int fm_split_lines(const uint8_t* text, int64_t text_bytes, int64_t* line_offsets,
                             int64_t max_lines, fm_parse_status* status, void* workspace,
                             size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  FM_REQUIRE(text_bytes >= 0 && max_lines >= 0, "fm_split_lines: negative size");
  FM_REQUIRE((text || text_bytes == 0) && line_offsets && status && workspace, "fm_split_lines: null pointer");
  FM_REQUIRE(((uintptr_t)text & 15) == 0, "fm_split_lines: text must be 16-byte aligned");
  const int64_t groups = (text_bytes + 15) / 16;
  if (workspace_bytes < scan_scratch_bytes(groups, 4))
    return fm_fail(FM_E_WORKSPACE, "fm_split_lines: workspace %zu < %zu", workspace_bytes, scan_scratch_bytes(groups, 4));
  FM_CUDA(cudaMemsetAsync(workspace, 0, scan_scratch_bytes(groups, 4), stream));
  FM_CUDA(cudaMemsetAsync(&status->first_error, 0xff, sizeof(uint64_t), stream));
  NewlineIn in{text, text_bytes};
  LineStartsOut out{in, line_offsets, max_lines, status};
  const int64_t tiles = scan_tiles(groups, 4);
  scan_kernel<4><<<(unsigned)tiles, kScanThreads, 0, stream>>>(in, out, scan_scratch_at(workspace, groups, 4), tiles);
  FM_LAUNCH_CHECK();
  return FM_OK;
}

FM_EXPORT size_t fm_parse_workspace(int64_t n_lines) { return scan_scratch_bytes(n_lines, 4); }

FM_EXPORT int fm_parse_count(const uint8_t* text, const int64_t* line_offsets, int64_t n_lines,
                             int64_t vocab_size, int hash_feature_id, int32_t* sizes,
                             int32_t* feature_poses, fm_parse_status* status, void* workspace,
                             size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  FM_REQUIRE(n_lines >= 0 && vocab_size > 0, "fm_parse_count: n_lines must be >= 0 and vocab_size > 0");
  FM_REQUIRE(feature_poses && status && workspace, "fm_parse_count: null pointer");
  FM_REQUIRE(n_lines == 0 || (text && line_offsets && sizes), "fm_parse_count: null pointer");
  FM_REQUIRE(((uintptr_t)text & 15) == 0, "fm_parse_count: text must be 16-byte aligned");
  if (workspace_bytes < scan_scratch_bytes(n_lines, 4))
    return fm_fail(FM_E_WORKSPACE, "fm_parse_count: workspace %zu < %zu", workspace_bytes, scan_scratch_bytes(n_lines, 4));
  FM_CUDA(cudaMemsetAsync(workspace, 0, scan_scratch_bytes(n_lines, 4), stream));
  FM_CUDA(cudaMemsetAsync(&status->first_error, 0xff, sizeof(uint64_t), stream));
  if (n_lines > 0) {
    const unsigned blocks = (unsigned)((n_lines + kWarpsPerBlock - 1) / kWarpsPerBlock);
    parse_count_kernel<<<blocks, kWarpsPerBlock * 32, 0, stream>>>(text, line_offsets, n_lines, vocab_size,
                                                                   hash_feature_id, sizes, status);
    FM_LAUNCH_CHECK();
  }
  SizesIn in{sizes, n_lines};
  PosesOut out{feature_poses, n_lines, status};
  const int64_t tiles = scan_tiles(n_lines, 4);
  scan_kernel<4><<<(unsigned)tiles, kScanThreads, 0, stream>>>(in, out, scan_scratch_at(workspace, n_lines, 4), tiles);
  FM_LAUNCH_CHECK();
  return FM_OK;
}

FM_EXPORT int fm_parse_fill(const uint8_t* text, const int64_t* line_offsets, int64_t n_lines,
                            int64_t vocab_size, int hash_feature_id, const int32_t* feature_poses,
                            float* labels, int64_t* feature_ids, float* feature_vals, int64_t capacity,
                            fm_parse_status* status, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  FM_REQUIRE(n_lines >= 0 && vocab_size > 0 && capacity >= 0, "fm_parse_fill: bad size");
  if (n_lines == 0) return FM_OK;
  FM_REQUIRE(text && line_offsets && feature_poses && labels && status, "fm_parse_fill: null pointer");
  FM_REQUIRE(capacity == 0 || (feature_ids && feature_vals), "fm_parse_fill: null output");
  FM_REQUIRE(((uintptr_t)text & 15) == 0, "fm_parse_fill: text must be 16-byte aligned");
  const unsigned blocks = (unsigned)((n_lines + kWarpsPerBlock - 1) / kWarpsPerBlock);
  parse_fill_kernel<<<blocks, kWarpsPerBlock * 32, 0, stream>>>(text, line_offsets, n_lines, vocab_size,
                                                                hash_feature_id, feature_poses, labels,
                                                                feature_ids, feature_vals, capacity, status);
  FM_LAUNCH_CHECK();
  return FM_OK;
}
