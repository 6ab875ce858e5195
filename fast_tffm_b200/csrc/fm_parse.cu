// fm_parse.cu -- FmParser on the GPU.
//
// Replaces cc/fm_parser_op.cc:29-51 (Compute: a serial loop over the batch) and
// :58-109 (ParseLine) plus the tf.cumsum at tffm/fm_model.py:73.
//
// Structure (8 lanes per line, 4 lines per warp; text read as 16-byte words):
//   pass 1 (count): count the ' ' bytes of a line and prove it "regular"
//       (no control bytes, no space at the start, after ':' or after a space)
//       with word-wide byte flags.  In a regular line every loop iteration of
//       ParseLine eats exactly one space, so sizes[i] = spaces - (trailing
//       space ? 1 : 0).  Any other line goes on a list and is counted by a
//       second, tiny kernel that follows ParseLine statement by statement.
//   scan: feature_poses = [0, cumsum(sizes)], single pass.
//   pass 2 (fill): the token starts of a 128-byte chunk are compacted into
//       shared memory; each lane parses whole tokens (label, id[:val]) with
//       the cheap exact number tiers.  A line in which any lane meets anything
//       else goes on a second list; a tiny kernel re-parses those lines
//       sequentially: error reporting and every libc quirk (whitespace
//       skipping, "inf", hex floats, long mantissas ...) live in that one place,
//       out of the hot kernels' register budget.
// Numbers are converted by fm_numparse.cuh (bit-exact with glibc).
#include "fm_common.cuh"
#include "fm_numparse.cuh"
#include "fm_scan.cuh"

namespace fm {
namespace {

constexpr int kWarpsPerBlock = 8;
constexpr int kLPL = 8;                    // lanes per line
constexpr int kLinesPerWarp = 32 / kLPL;
constexpr int kLinesPerBlock = kWarpsPerBlock * kLinesPerWarp;
constexpr int kChunk = kLPL * 16;          // bytes per group iteration (count pass)
constexpr int kFillChunk = kLPL * 32;      // fill pass: 32 bytes per lane and chunk; the tokens found are parsed
                                           // kLPL at a time in full rounds, left-overs wait for the next chunk's
constexpr int kSlowThreads = 128, kSlowBlocks = 148;

struct ParseWs {
  void* scan;        // scan scratch (zeroed per call)
  int* slow_count;   // [2]: lines queued for the sequential count / fill kernels (zeroed per call)
  int32_t* slow_a;   // [n_lines] count pass
  int32_t* slow_b;   // [n_lines] fill pass
  size_t scan_bytes, zero_bytes, total;
};
static inline ParseWs parse_ws(void* base, int64_t n_lines) {
  ParseWs w;
  const size_t n = (size_t)(n_lines > 0 ? n_lines : 1);
  w.scan_bytes = scan_scratch_bytes(n_lines, 4);
  char* p = reinterpret_cast<char*>(base);
  w.scan = p;
  w.slow_count = reinterpret_cast<int*>(p + w.scan_bytes);
  w.zero_bytes = w.scan_bytes + 256;
  w.slow_a = reinterpret_cast<int32_t*>(p + w.zero_bytes);
  w.slow_b = w.slow_a + n;
  w.total = w.zero_bytes + fm_align_up(2 * n * sizeof(int32_t), 256);
  return w;
}

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

// ---- byte flags on 32-bit words: bit 7 of byte k set <=> byte k has the property --------
__device__ __forceinline__ uint32_t flag_eq(uint32_t w, uint32_t c) {   // byte == c (exact)
  const uint32_t x = w ^ (c * 0x01010101u);
  return ~(((x & 0x7f7f7f7fu) + 0x7f7f7f7fu) | x | 0x7f7f7f7fu);
}
__device__ __forceinline__ uint32_t flag_ctrl(uint32_t w) {             // byte < 0x20
  return ~(((w & 0x7f7f7f7fu) + 0x60606060u) | w | 0x7f7f7f7fu);
}
// the four byte flags (bit 7 of each byte) of a word gathered into bits 0..3
__device__ __forceinline__ uint32_t flag_bits(uint32_t f) { return (((f >> 7) * 0x00204081u) >> 21) & 0xfu; }
__device__ __forceinline__ uint32_t flag_first_bytes(int n) {           // bytes 0..n-1 (n clipped to 0..4)
  return n <= 0 ? 0u : n >= 4 ? 0x80808080u : (0x80808080u >> (8 * (4 - n)));
}
// flags of the bytes of word k (bytes 4k..4k+3 of a 16-byte group) that lie in [lo, hi)
__device__ __forceinline__ uint32_t flag_valid(int k, int lo, int hi) {
  return flag_first_bytes(hi - 4 * k) & ~flag_first_bytes(lo - 4 * k);
}

__device__ __forceinline__ unsigned group_mask(int lane) {
  return kLPL == 32 ? kFull : (((1u << kLPL) - 1u) << ((lane / kLPL) * kLPL));
}
template <class T>
__device__ __forceinline__ T group_sum(T v, unsigned gm) {
#pragma unroll
  for (int o = kLPL / 2; o > 0; o >>= 1) v += __shfl_xor_sync(gm, v, o);
  return v;
}

__global__ void __launch_bounds__(kWarpsPerBlock * 32)
parse_count_kernel(const uint8_t* __restrict__ text, const int64_t* __restrict__ begins,
                   const int64_t* __restrict__ ends, int64_t end_adjust, int64_t n_lines,
                   int32_t* __restrict__ sizes, int32_t* __restrict__ slow_list, int* __restrict__ slow_count) {
  const int lane = threadIdx.x & 31, c = lane % kLPL;
  const unsigned gm = group_mask(lane);
  const int64_t line = ((int64_t)blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5)) * kLinesPerWarp + lane / kLPL;
  if (line >= n_lines) return;  // the whole lane group leaves together
  const int64_t s = begins[line], e = ends[line] + end_adjust;
  int spaces = 0;
  uint32_t irregular = 0, trail = 0;
  uint32_t carry = 0x80u;  // "the byte before the line start is a space": a leading space is irregular
  for (int64_t pos = s & ~(int64_t)15; pos < e; pos += kChunk) {
    const int64_t p = pos + c * 16;
    uint32_t w[4] = {0, 0, 0, 0};
    if (p < e) { const uint4 v = load16(text + p); w[0] = v.x; w[1] = v.y; w[2] = v.z; w[3] = v.w; }
    const int lo = s > p ? (int)(s - p) : 0;
    const int hi = e - p >= 16 ? 16 : (e - p > 0 ? (int)(e - p) : 0);
    const bool full = lo == 0 && hi == 16;
    uint32_t sp[4], sc[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      sp[k] = flag_eq(w[k], ' ');
      sc[k] = sp[k] | flag_eq(w[k], ':');
      uint32_t bad = flag_ctrl(w[k]);
      if (!full) {
        const uint32_t v = flag_valid(k, lo, hi);
        sp[k] &= v; sc[k] &= v; bad &= v;
      }
      irregular |= bad;
      spaces += __popc(sp[k]);
    }
    // previous-byte flags: byte t inherits the space/colon flag of byte t-1
    uint32_t prev_in = __shfl_up_sync(gm, sc[3] >> 24, 1, kLPL);
    if (c == 0) prev_in = carry;
    uint32_t pv[4];
    pv[0] = (sc[0] << 8) | prev_in;
#pragma unroll
    for (int k = 1; k < 4; ++k) pv[k] = (sc[k] << 8) | (sc[k - 1] >> 24);
    if (s > p) pv[(s - p) >> 2] |= 0x80u << (8 * ((s - p) & 3));  // the line starts inside this word
#pragma unroll
    for (int k = 0; k < 4; ++k) irregular |= sp[k] & pv[k];
    if (p <= e - 1 && e - 1 < p + 16) trail = (sp[(e - 1 - p) >> 2] >> (8 * ((e - 1 - p) & 3) + 7)) & 1u;
    carry = __shfl_sync(gm, sc[3] >> 24, kLPL - 1, kLPL);
  }
  spaces = group_sum(spaces, gm);
  irregular = group_sum(irregular != 0 ? 1u : 0u, gm);
  trail = group_sum(trail, gm);
  if (c == 0) {
    if (irregular) {
      slow_list[atomicAdd(slow_count, 1)] = (int32_t)line;
      sizes[line] = 0;
    } else {
      sizes[line] = spaces - (trail ? 1 : 0);
    }
  }
}

// the lines the fast counter could not prove regular: sequential ParseLine, count only
__global__ void __launch_bounds__(kSlowThreads)
parse_count_slow_kernel(const uint8_t* __restrict__ text, const int64_t* __restrict__ begins,
                        const int64_t* __restrict__ ends, int64_t end_adjust, int64_t vocab, int hash,
                        int32_t* __restrict__ sizes, const int32_t* __restrict__ slow_list,
                        const int* __restrict__ slow_count, fm_parse_status* st) {
  const int n = *slow_count;
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) {
    const int64_t line = slow_list[t];
    const int64_t s = begins[line], e = ends[line] + end_adjust;
    int32_t cnt = 0;
    const int rc = parse_line_seq(text + s, e - s, hash, vocab, nullptr, nullptr, nullptr, 0, &cnt);
    if (rc != FM_PARSE_OK) { report_error(st, line, rc); cnt = 0; }
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

// [+-]digits[.digits][e[+-]digits] through the cheap exact tiers only.  getc(i) yields byte i of
// the candidate text (0 at and past its end).  Returns the bytes consumed, or 0 when the text is
// anything else (leading whitespace, inf/nan, hex, a mantissa or exponent beyond the tiers): the
// caller then sends the line to the sequential kernel, which knows the whole of strtof.
template <class GetC>
__device__ __forceinline__ int scan_value_fast(GetC getc, uint32_t* bits) {
  int i = 0;
  uint32_t c = getc(0);
  uint32_t sign = 0;
  if (c == '-') { sign = 0x80000000u; c = getc(++i); }
  else if (c == '+') { c = getc(++i); }
  uint64_t w = 0;
  int nd = 0, frac = 0;
  bool any = false;
  for (; c - '0' < 10u; c = getc(++i)) { w = w * 10 + (c - '0'); nd += (w != 0); any = true; }
  if (c == '.') {
    c = getc(++i);
    for (; c - '0' < 10u; c = getc(++i)) { w = w * 10 + (c - '0'); nd += (w != 0); ++frac; any = true; }
  }
  if (!any || nd > 19) return 0;
  int ex = 0;
  if ((c | 0x20) == 'e') {
    int k = i + 1;
    uint32_t d = getc(k);
    bool eneg = false;
    if (d == '-') { eneg = true; d = getc(++k); } else if (d == '+') { d = getc(++k); }
    if (d - '0' < 10u) {
      for (; d - '0' < 10u; d = getc(++k)) if (ex < 10000) ex = ex * 10 + (int)(d - '0');
      if (eneg) ex = -ex;
      i = k;
    }
  }
  uint32_t r;
  if (!fmnp::dec_to_f32_fast(w, (int64_t)ex - frac, &r)) return 0;
  *bits = sign | r;
  return i;
}

struct MemBytes {  // bytes straight from memory (fallback for tokens that outgrow the register window)
  const uint8_t* p; int n;
  __device__ __forceinline__ uint32_t operator()(int i) const { return i < n ? p[i] : 0u; }
};

// ---- tokens from a 32-byte register window ------------------------------------------
// 0x80 in every byte of x that is zero (exact)
__device__ __forceinline__ uint64_t zero_bytes(uint64_t x) {
  const uint64_t k = 0x7f7f7f7f7f7f7f7full;
  return ~(((x & k) + k) | x | k);
}
__device__ __forceinline__ uint64_t keep_low_bytes(int n) {  // mask of the n lowest bytes, n clipped to 0..8
  return n >= 8 ? ~0ull : n <= 0 ? 0ull : ((1ull << (8 * n)) - 1ull);
}
struct Window {   // 32 bytes of text from the token start (a line's last token: bytes at and past its end are 0)
  uint64_t w0, w1, w2, w3;
  __device__ __forceinline__ uint64_t word(int j) const { return j == 0 ? w0 : j == 1 ? w1 : j == 2 ? w2 : j == 3 ? w3 : 0ull; }
  __device__ __forceinline__ uint32_t byte(int i) const { return (uint32_t)(word(i >> 3) >> (8 * (i & 7))) & 0xffu; }
  // 8 bytes starting at byte `off` (zeros past byte 31)
  __device__ __forceinline__ uint64_t at(int off) const {
    const int j = off >> 3, sh = (off & 7) * 8;
    const uint64_t lo = word(j);
    return sh ? ((lo >> sh) | (word(j + 1) << (64 - sh))) : lo;
  }
};
__device__ __forceinline__ Window load_window(const uint8_t* __restrict__ text, int64_t q, int64_t limit, int len,
                                              bool mask_tail) {
  const int64_t a = q & ~(int64_t)7;
  const int span = ((int)q & 7) + len;        // bytes from a to the token's end: only the words that hold them are read
  uint64_t v[5];
  const bool inside = a + 40 <= limit;        // nearly always: no per-word bound checks
#pragma unroll
  for (int k = 0; k < 5; ++k)
    v[k] = (8 * k < span && (inside || a + 8 * k + 8 <= limit)) ? __ldg(reinterpret_cast<const uint64_t*>(text + a + 8 * k))
                                                               : 0ull;
  const int sh = ((int)q & 7) * 8;
  Window w;
  if (sh) {
    w.w0 = (v[0] >> sh) | (v[1] << (64 - sh)); w.w1 = (v[1] >> sh) | (v[2] << (64 - sh));
    w.w2 = (v[2] >> sh) | (v[3] << (64 - sh)); w.w3 = (v[3] >> sh) | (v[4] << (64 - sh));
  } else {
    w.w0 = v[0]; w.w1 = v[1]; w.w2 = v[2]; w.w3 = v[3];
  }
  if (mask_tail) {                            // a line's last token: no space behind it, whatever follows reads as 0
    w.w0 &= keep_low_bytes(len); w.w1 &= keep_low_bytes(len - 8);
    w.w2 &= keep_low_bytes(len - 16); w.w3 &= keep_low_bytes(len - 24);
  }
  return w;
}
// index of the first ':' in the first len (< 32) bytes of the window; len if none
__device__ __forceinline__ int find_colon(const Window& w, int len) {
  const uint64_t co = 0x3a3a3a3a3a3a3a3aull;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    if (8 * j >= len) break;
    const uint64_t f = zero_bytes(w.word(j) ^ co);
    if (f) { const int at = 8 * j + ((__ffsll((long long)f) >> 3) - 1); return at < len ? at : len; }
  }
  return len;
}
struct RegBytes {  // 16 bytes held in two registers: a value of up to 15 bytes and what ends it (a space or 0)
  uint64_t lo, hi;
  __device__ __forceinline__ uint32_t operator()(int i) const {
    return i < 16 ? (uint32_t)((i < 8 ? lo : hi) >> (8 * (i & 7))) & 0xffu : 0u;
  }
};

enum { kTokOk = 0, kTokFallback = 1, kTokBad = 2 };

// four decimal digits held as byte values 0..9, the most significant in the lowest byte -> their number
__device__ __forceinline__ uint32_t four_digits(uint32_t t) {
  t = (t * 10u + (t >> 8)) & 0x00ff00ffu;         // bytes 0 and 2: 10 d0 + d1, 10 d2 + d3
  return (t * 100u + (t >> 16)) & 0xffffu;
}

// one token "id[:val]" (or the label when is_label) in [q, e), from registers: e is the token's end -- the tokenizer
// already knows it --, line_end says that it is the line's end too (else text[e] is a space).
// kTokFallback: the token does not fit the window, use the byte-wise parser.
__device__ __forceinline__ int parse_token_regs(const uint8_t* __restrict__ text, int64_t q, int64_t e, int64_t limit,
                                                int hash, int64_t vocab, bool is_label, bool line_end,
                                                int64_t* id_out, uint32_t* bits_out) {
  if (e - q >= 32) return kTokFallback;
  const int len = (int)(e - q);
  if (len == 0) return kTokBad;    // empty token: the sequential path knows what the reference does with it
  const Window w = load_window(text, q, limit, len, line_end);
  int vstart = 0;
  if (!is_label) {
    const int n = find_colon(w, len);
    if (n == 0) return kTokBad;
    int64_t id;
    if (hash) {
      const uint64_t m = 0xc6a4a7935bd1e995ull;
      uint64_t h = 0xDECAFCAFFEull ^ ((uint64_t)n * m);
#pragma unroll
      for (int j = 0; j < 4; ++j) {             // n < 32: words 0..3 with static indices, the tail inside the loop
        if (8 * (j + 1) <= n) { uint64_t k = w.word(j); k *= m; k ^= k >> 47; k *= m; h ^= k; h *= m; }
        else if (8 * j < n) { h ^= w.word(j) & ((1ull << (8 * (n - 8 * j))) - 1ull); h *= m; }
      }
      h ^= h >> 47; h *= m; h ^= h >> 47;
      id = (int64_t)((vocab & (vocab - 1)) == 0 ? (h & (uint64_t)(vocab - 1)) : (h % (uint64_t)vocab));
    } else {
      if (n > 18) return kTokBad;
      uint64_t acc = 0;
      if (n <= 8) {                              // the usual id: up to eight digits, checked and folded in registers
        uint64_t t = (w.w0 & keep_low_bytes(n)) ^ (0x3030303030303030ull & keep_low_bytes(n));
        if ((((t + 0x7676767676767676ull) | t) & 0x8080808080808080ull) != 0) return kTokBad;
        t <<= 8 * (8 - n);
        acc = (uint64_t)four_digits((uint32_t)t) * 10000u + four_digits((uint32_t)(t >> 32));
      } else {
        for (int i = 0; i < n; ++i) {
          const uint32_t d = w.byte(i) - '0';
          if (d >= 10u) return kTokBad;
          acc = acc * 10 + d;
        }
      }
      if ((int64_t)acc >= vocab) return kTokBad;
      id = (int64_t)acc;
    }
    *id_out = id;
    if (n == len) { *bits_out = 0x3f800000u; return kTokOk; }  // no colon: value 1
    vstart = n + 1;
  }
  // the value (or label): the rest of the token
  const int vlen = len - vstart;
  if (vlen >= 16) return kTokFallback;
  if (vlen == 0) return kTokBad;
  RegBytes v;
  v.lo = w.at(vstart);
  v.hi = w.at(vstart + 8);
  if (vlen <= 10) {
    // [+-]digits[.digits] with at most 8 digits, no exponent -- what a feature value nearly always is ("0.6931",
    // "-0.123456") -- without a loop: sign and dot are taken out, the digits are checked and converted in
    // registers.  Anything else falls through to the general scanner; both end in the same dec_to_f32_fast(w, -frac).
    const uint32_t c0 = (uint32_t)v.lo & 0xffu;
    const bool neg = c0 == '-';
    const int sgn = (neg || c0 == '+') ? 1 : 0;
    const int blen = vlen - sgn;                  // the body: digits and at most one dot, 9 bytes at most to be fast
    const uint64_t lo = sgn ? (v.lo >> 8) | (v.hi << 56) : v.lo;              // body bytes 0..7
    const uint64_t b8 = (sgn ? v.hi >> 8 : v.hi) & 0xffull;                   // body byte 8
    const uint64_t x = lo & keep_low_bytes(blen);
    const uint64_t dotf = zero_bytes(x ^ 0x2e2e2e2e2e2e2e2eull);
    const bool dot8 = !dotf && blen == 9 && b8 == '.';                        // "12345678."
    const bool has_dot = dotf != 0 || dot8;
    const int dp = dotf ? (__ffsll((long long)dotf) >> 3) - 1 : dot8 ? 8 : blen;
    const int ndig = blen - (has_dot ? 1 : 0);
    if (ndig >= 1 && ndig <= 8) {
      const uint64_t below = keep_low_bytes(dp);
      const uint64_t up = dotf ? (x >> 8) | (blen == 9 ? b8 << 56 : 0ull) : x;   // the bytes behind the dot move down
      uint64_t t = ((x & below) | (up & ~below)) ^ (0x3030303030303030ull & keep_low_bytes(ndig));
      if ((((t + 0x7676767676767676ull) | t) & 0x8080808080808080ull) == 0) {
        t <<= 8 * (8 - ndig);                     // leading zeros in front (the low bytes): eight digits, d0 first
        const uint64_t wv = (uint64_t)four_digits((uint32_t)t) * 10000u + four_digits((uint32_t)(t >> 32));
        uint32_t r;
        if (!fmnp::dec_to_f32_fast(wv, -(int64_t)(ndig - dp), &r)) return kTokBad;
        *bits_out = (neg ? 0x80000000u : 0u) | r;
        return kTokOk;
      }
    }
  }
  const int used = scan_value_fast(v, bits_out);
  return used == vlen ? kTokOk : kTokBad;
}

// the same token, byte by byte from memory (any length)
__device__ __noinline__ int parse_token_bytes(const uint8_t* __restrict__ text, int64_t q, int64_t e, int hash,
                                              int64_t vocab, bool is_label, int64_t* id_out, uint32_t* bits_out) {
  if (!is_label) {
    if (hash) {
      const int64_t a = q;
      uint8_t ch = text[q];
      while (ch != ':' && ch != ' ') { ++q; if (q >= e) break; ch = text[q]; }
      if (q == a) return kTokBad;
      const uint64_t hsh = fmnp::hash64(text + a, q - a);
      *id_out = (int64_t)((vocab & (vocab - 1)) == 0 ? (hsh & (uint64_t)(vocab - 1)) : (hsh % (uint64_t)vocab));
    } else {
      int nd = 0;
      uint64_t acc = 0;
      for (; q < e && fmnp::is_digit(text[q]) && nd < 18; ++q, ++nd) acc = acc * 10 + (text[q] - '0');
      if (nd == 0 || (q < e && fmnp::is_digit(text[q])) || (int64_t)acc >= vocab) return kTokBad;
      *id_out = (int64_t)acc;
    }
    if (!(q < e && text[q] == ':')) {
      *bits_out = 0x3f800000u;
      return (q == e || text[q] == ' ') ? kTokOk : kTokBad;
    }
    ++q;
  }
  MemBytes mb{text + q, (int)min((int64_t)64, e - q)};
  const int used = scan_value_fast(mb, bits_out);
  if (used == 0) return kTokBad;
  q += used;
  return (q == e || text[q] == ' ') ? kTokOk : kTokBad;
}

__global__ void __launch_bounds__(kWarpsPerBlock * 32)
parse_fill_kernel(const uint8_t* __restrict__ text, int64_t limit, const int64_t* __restrict__ begins,
                  const int64_t* __restrict__ ends, int64_t end_adjust, int64_t n_lines,
                  int64_t vocab, int hash, const int32_t* __restrict__ poses, float* __restrict__ labels,
                  int64_t* __restrict__ ids, float* __restrict__ vals, int64_t capacity,
                  int32_t* __restrict__ slow_list, int* __restrict__ slow_count) {
  // token starts of a line, as byte offsets from (pos - kFillChunk): a chunk's tokens, behind the few (< kLPL) the
  // previous chunk left over -- tokens are parsed kLPL at a time, in full rounds, whatever chunk they began in
  __shared__ uint16_t s_tok[kWarpsPerBlock][kLinesPerWarp][kFillChunk + kLPL + 6];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane / kLPL, c = lane % kLPL;
  const unsigned gm = group_mask(lane);
  for (int64_t line = ((int64_t)blockIdx.x * kWarpsPerBlock + warp) * kLinesPerWarp + g; line < n_lines;
       line += (int64_t)gridDim.x * kLinesPerBlock) {
  const int64_t s = begins[line], e = ends[line] + end_adjust;
  const int64_t out_base = poses[line];
  const int32_t cnt = poses[line + 1] - poses[line];
  uint16_t* tok = s_tok[warp][g];
  bool bad = out_base + cnt > capacity || e <= s;
  int32_t ord_base = -1;  // token -1 is the label
  int pending = 0;        // tokens carried over from the previous chunk (group-uniform)
  for (int64_t pos = s & ~(int64_t)15; pos < e && !bad; pos += kFillChunk) {
    const int64_t p = pos + c * 32;
    uint32_t w[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (p < e) { const uint4 v = load16(text + p); w[0] = v.x; w[1] = v.y; w[2] = v.z; w[3] = v.w; }
    if (p + 16 < e) { const uint4 v = load16(text + p + 16); w[4] = v.x; w[5] = v.y; w[6] = v.z; w[7] = v.w; }
    const int lo = s > p ? (int)(s - p) : 0;      // < 16: only the first chunk starts before the line
    const int hi = e - p >= 32 ? 32 : (e - p > 0 ? (int)(e - p) : 0);
    // one bit per byte of the lane's 32: spaces and NULs inside [lo, hi)
    const uint32_t below_hi = hi >= 32 ? 0xffffffffu : ((1u << hi) - 1u);
    const uint32_t in_line = hi > lo ? (below_hi & ~((1u << lo) - 1u)) : 0u;
    uint32_t sp32 = 0, zero_somewhere = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      sp32 |= flag_bits(flag_eq(w[k], ' ')) << (4 * k);
      zero_somewhere |= (w[k] - 0x01010101u) & ~w[k];         // bit 7 of some byte set <=> the word has a zero byte
    }
    sp32 &= in_line;
    if (zero_somewhere & 0x80808080u) {           // a NUL ends the line early: sequential path.  Exact where: only
      uint32_t nul32 = 0;                         // the lanes at a line's two ends get here for nothing
#pragma unroll
      for (int k = 0; k < 8; ++k) nul32 |= flag_bits(flag_eq(w[k], 0)) << (4 * k);
      bad |= (nul32 & in_line) != 0;
    }
    int n_here = __popc(sp32);
    const bool first = pos <= s;                  // the chunk that holds the line start also lists the label
    if (first && c == 0) n_here += 1;
    int incl = n_here;
#pragma unroll
    for (int o = 1; o < kLPL; o <<= 1) {
      const int t = __shfl_up_sync(gm, incl, o, kLPL);
      if (c >= o) incl += t;
    }
    const int total = pending + __shfl_sync(gm, incl, kLPL - 1, kLPL);
    int wpos = pending + incl - n_here;
    if (first && c == 0) tok[wpos++] = (uint16_t)(kFillChunk + (s - pos));
    for (uint32_t f = sp32; f; f &= f - 1) tok[wpos++] = (uint16_t)(kFillChunk + c * 32 + __ffs(f));  // the byte after the space
    // full rounds of kLPL tokens, each with its successor already listed (a token ends one byte before the next
    // begins); the line's last chunk takes what is left: its last token ends with the line
    const bool last = pos + kFillChunk >= e;
    const int n_proc = last ? total : total > 0 ? ((total - 1) & ~(kLPL - 1)) : 0;
    if (last && c == 0) tok[total] = (uint16_t)(kFillChunk + (e - pos) + 1);
    __syncwarp(gm);
    for (int t = c; t < n_proc; t += kLPL) {
      const int64_t q = pos - kFillChunk + tok[t];
      const int64_t q_end = pos - kFillChunk + tok[t + 1] - 1;
      const int32_t ordinal = ord_base + t;
      const bool is_label = ordinal < 0;
      if (!is_label) {
        if (q >= e) { bad |= ordinal != cnt; continue; }  // trailing space: breaks without counting (:79-81)
        if (ordinal >= cnt) { bad = true; continue; }
      }
      int64_t id = 0;
      uint32_t bits = 0x3f800000u;
      int rc = parse_token_regs(text, q, q_end, limit, hash, vocab, is_label, q_end >= e, &id, &bits);
      if (rc == kTokFallback) rc = parse_token_bytes(text, q, q_end, hash, vocab, is_label, &id, &bits);
      if (rc != kTokOk) { bad = true; continue; }
      if (is_label) {
        labels[line] = __uint_as_float(bits);
      } else {
        ids[out_base + ordinal] = id;
        vals[out_base + ordinal] = __uint_as_float(bits);
      }
    }
    ord_base += n_proc;
    // the tokens left over move to the front, re-based for the next chunk (at most kLPL: one per lane)
    pending = total - n_proc;
    const int carried = c < pending ? (int)tok[n_proc + c] - kFillChunk : 0;
    bad |= carried < 0;                           // a token that spans a whole chunk: not a regular line
    bad = __any_sync(gm, bad);
    __syncwarp(gm);                               // every read of tok[] above is done
    if (c < pending) tok[c] = (uint16_t)carried;
    __syncwarp(gm);
  }
  bad = __any_sync(gm, bad);
  // a line that ended short of its count (cannot happen for a regular line)
  if (!bad && ord_base != cnt && ord_base != cnt + 1) bad = true;
  if (bad && c == 0) slow_list[atomicAdd(slow_count, 1)] = (int32_t)line;
  }
}

// the lines the fast path gave up on: sequential ParseLine with outputs and error reporting
__global__ void __launch_bounds__(kSlowThreads)
parse_fill_slow_kernel(const uint8_t* __restrict__ text, const int64_t* __restrict__ begins,
                       const int64_t* __restrict__ ends, int64_t end_adjust, int64_t vocab, int hash,
                       const int32_t* __restrict__ poses, float* __restrict__ labels, int64_t* __restrict__ ids,
                       float* __restrict__ vals, int64_t capacity, const int32_t* __restrict__ slow_list,
                       const int* __restrict__ slow_count, fm_parse_status* st) {
  const int n = *slow_count;
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) {
    const int64_t line = slow_list[t];
    const int64_t s = begins[line], e = ends[line] + end_adjust;
    const int64_t out_base = poses[line];
    const int32_t cnt = poses[line + 1] - poses[line];
    if (out_base + cnt > capacity) { report_error(st, line, FM_PARSE_E_CAPACITY); continue; }
    int32_t got = 0;
    const int rc = parse_line_seq(text + s, e - s, hash, vocab, labels + line, ids + out_base, vals + out_base, cnt,
                                  &got);
    if (rc != FM_PARSE_OK) report_error(st, line, rc);
    else if (got != cnt) report_error(st, line, FM_PARSE_E_FORMAT);
  }
}

// ---- line splitting ---------------------------------------------------------
struct NewlineIn {
  const uint8_t* text; int64_t nbytes;
  __device__ int64_t count() const { return (nbytes + 15) / 16; }
  __device__ void flags(int64_t g, uint32_t (&f)[4]) const {
    const uint4 v = load16(text + g * 16);
    const int64_t left = nbytes - g * 16;
    const int hi = left >= 16 ? 16 : (int)left;
    f[0] = flag_eq(v.x, '\n') & flag_valid(0, 0, hi);
    f[1] = flag_eq(v.y, '\n') & flag_valid(1, 0, hi);
    f[2] = flag_eq(v.z, '\n') & flag_valid(2, 0, hi);
    f[3] = flag_eq(v.w, '\n') & flag_valid(3, 0, hi);
  }
  __device__ long long operator()(int64_t g) const {
    uint32_t f[4];
    flags(g, f);
    return __popc(f[0]) + __popc(f[1]) + __popc(f[2]) + __popc(f[3]);
  }
};
struct LineStartsOut {
  NewlineIn in; int64_t* offsets; int64_t max_lines; fm_parse_status* st;
  __device__ void operator()(int64_t g, long long excl, long long cnt) const {
    if (g == 0) offsets[0] = 0;
    if (cnt == 0) return;
    uint32_t f[4];
    in.flags(g, f);
    long long r = excl;
#pragma unroll
    for (int k = 0; k < 4; ++k)
      for (uint32_t m = f[k]; m; m &= m - 1, ++r)
        if (r + 1 <= max_lines) offsets[r + 1] = g * 16 + 4 * k + (__ffs(m) >> 3);
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

FM_EXPORT size_t fm_parse_workspace(int64_t n_lines) { return parse_ws(nullptr, n_lines < 0 ? 0 : n_lines).total; }

FM_EXPORT int fm_parse_count(const uint8_t* text, int64_t text_bytes, const int64_t* line_begins,
                             const int64_t* line_ends, int64_t end_adjust, int64_t n_lines,
                             int64_t vocab_size, int hash_feature_id, int32_t* sizes,
                             int32_t* feature_poses, fm_parse_status* status, void* workspace,
                             size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  FM_REQUIRE(n_lines >= 0 && vocab_size > 0, "fm_parse_count: n_lines must be >= 0 and vocab_size > 0");
  if (n_lines >= (1ll << 31)) return fm_fail(FM_E_LIMIT, "fm_parse_count: n_lines >= 2^31");
  FM_REQUIRE(feature_poses && status && workspace, "fm_parse_count: null pointer");
  FM_REQUIRE(n_lines == 0 || (text && line_begins && line_ends && sizes), "fm_parse_count: null pointer");
  FM_REQUIRE(text_bytes >= 0, "fm_parse_count: negative text_bytes");
  FM_REQUIRE(((uintptr_t)text & 15) == 0, "fm_parse_count: text must be 16-byte aligned");
  const ParseWs ws = parse_ws(workspace, n_lines);
  if (workspace_bytes < ws.total)
    return fm_fail(FM_E_WORKSPACE, "fm_parse_count: workspace %zu < %zu", workspace_bytes, ws.total);
  FM_CUDA(cudaMemsetAsync(workspace, 0, ws.zero_bytes, stream));
  FM_CUDA(cudaMemsetAsync(&status->first_error, 0xff, sizeof(uint64_t), stream));
  if (n_lines > 0) {
    const unsigned blocks = (unsigned)((n_lines + kLinesPerBlock - 1) / kLinesPerBlock);
    parse_count_kernel<<<blocks, kWarpsPerBlock * 32, 0, stream>>>(text, line_begins, line_ends, end_adjust, n_lines,
                                                                   sizes, ws.slow_a, ws.slow_count);
    FM_LAUNCH_CHECK();
    parse_count_slow_kernel<<<kSlowBlocks, kSlowThreads, 0, stream>>>(text, line_begins, line_ends, end_adjust,
                                                                      vocab_size, hash_feature_id, sizes, ws.slow_a,
                                                                      ws.slow_count, status);
    FM_LAUNCH_CHECK();
  }
  SizesIn in{sizes, n_lines};
  PosesOut out{feature_poses, n_lines, status};
  const int64_t tiles = scan_tiles(n_lines, 4);
  scan_kernel<4><<<(unsigned)tiles, kScanThreads, 0, stream>>>(in, out, scan_scratch_at(ws.scan, n_lines, 4), tiles);
  FM_LAUNCH_CHECK();
  return FM_OK;
}

FM_EXPORT int fm_parse_fill(const uint8_t* text, int64_t text_bytes, const int64_t* line_begins,
                            const int64_t* line_ends, int64_t end_adjust, int64_t n_lines,
                            int64_t vocab_size, int hash_feature_id, const int32_t* feature_poses,
                            float* labels, int64_t* feature_ids, float* feature_vals, int64_t capacity,
                            fm_parse_status* status, void* workspace, size_t workspace_bytes, int max_blocks,
                            void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  FM_REQUIRE(n_lines >= 0 && vocab_size > 0 && capacity >= 0, "fm_parse_fill: bad size");
  if (n_lines == 0) return FM_OK;
  FM_REQUIRE(text && line_begins && line_ends && feature_poses && labels && status && workspace,
             "fm_parse_fill: null pointer");
  FM_REQUIRE(text_bytes >= 0, "fm_parse_fill: negative text_bytes");
  FM_REQUIRE(capacity == 0 || (feature_ids && feature_vals), "fm_parse_fill: null output");
  FM_REQUIRE(((uintptr_t)text & 15) == 0, "fm_parse_fill: text must be 16-byte aligned");
  const ParseWs ws = parse_ws(workspace, n_lines);
  if (workspace_bytes < ws.total)
    return fm_fail(FM_E_WORKSPACE, "fm_parse_fill: workspace %zu < %zu", workspace_bytes, ws.total);
  unsigned blocks = (unsigned)((n_lines + kLinesPerBlock - 1) / kLinesPerBlock);
  if (max_blocks > 0 && blocks > (unsigned)max_blocks) blocks = (unsigned)max_blocks;
  const int64_t limit = (text_bytes + 15) & ~(int64_t)15;  // readable extent of the text (header contract)
  parse_fill_kernel<<<blocks, kWarpsPerBlock * 32, 0, stream>>>(text, limit, line_begins, line_ends, end_adjust, n_lines,
                                                                vocab_size, hash_feature_id, feature_poses, labels,
                                                                feature_ids, feature_vals, capacity, ws.slow_b,
                                                                ws.slow_count + 1);
  FM_LAUNCH_CHECK();
  parse_fill_slow_kernel<<<kSlowBlocks, kSlowThreads, 0, stream>>>(text, line_begins, line_ends, end_adjust, vocab_size,
                                                                   hash_feature_id, feature_poses, labels, feature_ids,
                                                                   feature_vals, capacity, ws.slow_b, ws.slow_count + 1,
                                                                   status);
  FM_LAUNCH_CHECK();
  return FM_OK;
}
