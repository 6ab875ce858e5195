// fm_numparse.cuh -- text -> number conversion for the GPU line parser.
//
// Replaces the libc calls the reference parser makes per token
// (strtof at cc/fm_parser_op.cc:69 and :98, strtoll at :86) with functions
// that return the same bits and consume the same bytes as glibc does in the
// "C" locale: leading isspace() skipping, sign, decimal and hexadecimal
// floats with correct round-to-nearest-even (including subnormals, overflow to
// infinity and underflow to zero), "inf"/"infinity", "nan" and "nan(payload)".
//
// Every function is __host__ __device__ so that the same source is fuzzed
// against glibc on the CPU (tests/native/numparse_check.cpp) before it runs on
// the GPU.
//
// Decimal -> binary32 is done in three tiers, each exact:
//   1. w <= 2^24, |q| <= 10: one IEEE float multiply/divide (both operands exact).
//   2. w <= 2^53, |q| <= 22: one IEEE double multiply/divide, accepted unless the
//      double lies exactly on a binary32 rounding boundary (then double rounding
//      could bite) -> tier 3.
//   3. exact big-integer long division on up to 128 significant digits + sticky
//      (a binary32 midpoint has < 128 significant decimal digits, so later
//      digits can only break a tie upward).
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define FMNP_HD __host__ __device__
#define FMNP_NOINLINE __noinline__
#else
#define FMNP_HD
#define FMNP_NOINLINE __attribute__((noinline))
#endif

namespace fmnp {

FMNP_HD inline bool is_space(uint8_t c) { return c == ' ' || (uint8_t)(c - 9) < 5; }
FMNP_HD inline bool is_digit(uint8_t c) { return (uint8_t)(c - '0') < 10; }
FMNP_HD inline uint8_t lower(uint8_t c) { return (uint8_t)(c - 'A') < 26 ? (uint8_t)(c + 32) : c; }
FMNP_HD inline bool is_alnum(uint8_t c) { return is_digit(c) || (uint8_t)(lower(c) - 'a') < 26; }
FMNP_HD inline int hex_val(uint8_t c) {
  if (is_digit(c)) return c - '0';
  c = lower(c);
  return (uint8_t)(c - 'a') < 6 ? c - 'a' + 10 : -1;
}
// byte i of a line of n bytes; reading at or past the end yields the NUL that
// terminates std::string::c_str() in the reference
FMNP_HD inline uint8_t at(const uint8_t* p, int64_t n, int64_t i) { return i < n ? p[i] : (uint8_t)0; }

FMNP_HD inline int bit_length64(uint64_t x) {
#if defined(__CUDA_ARCH__)
  return 64 - __clzll((long long)x);
#else
  return x ? 64 - __builtin_clzll(x) : 0;
#endif
}

FMNP_HD inline uint32_t f32_bits(float f) {
#if defined(__CUDA_ARCH__)
  return __float_as_uint(f);
#else
  union { float f; uint32_t u; } c; c.f = f; return c.u;
#endif
}
FMNP_HD inline uint64_t f64_bits(double d) {
#if defined(__CUDA_ARCH__)
  return (uint64_t)__double_as_longlong(d);
#else
  union { double d; uint64_t u; } c; c.d = d; return c.u;
#endif
}

// Round the positive value (q + sticky) * 2^lsb_exp to binary32, nearest-even.
FMNP_HD inline uint32_t round_pack(uint64_t q, int64_t lsb_exp, bool sticky) {
  if (q == 0) return 0u;  // callers only pass sticky with q != 0
  const int nb = bit_length64(q);
  const int64_t e2 = nb - 1 + lsb_exp;
  if (e2 > 127) return 0x7f800000u;
  if (e2 < -200) return 0u;
  int64_t sh = nb - 24;
  const bool subnormal = e2 < -126;
  if (subnormal) sh = -149 - lsb_exp;
  uint64_t mant;
  if (sh <= 0) {
    mant = q << (-sh);  // exact (sticky is never set with so few bits)
  } else if (sh > 64) {
    return 0u;
  } else {
    const uint64_t half = 1ull << (sh - 1);
    const uint64_t rem = sh == 64 ? q : (q & ((half << 1) - 1));
    mant = sh == 64 ? 0 : (q >> sh);
    if (rem > half || (rem == half && (sticky || (mant & 1)))) mant += 1;
  }
  if (subnormal) return (uint32_t)mant;  // a carry into 2^23 is the smallest normal
  const uint64_t bits = ((uint64_t)(e2 + 126) << 23) + mant;
  return bits >= 0x7f800000ull ? 0x7f800000u : (uint32_t)bits;
}

// ---- tiers 1 and 2 ---------------------------------------------------------
#define FMNP_P10F {1e0f, 1e1f, 1e2f, 1e3f, 1e4f, 1e5f, 1e6f, 1e7f, 1e8f, 1e9f, 1e10f}
#define FMNP_P10D {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11, \
                   1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22}
#if defined(__CUDACC__)
static __device__ __constant__ float kDevP10F[11] = FMNP_P10F;
static __device__ __constant__ double kDevP10D[23] = FMNP_P10D;
#endif
static const float kHostP10F[11] = FMNP_P10F;
static const double kHostP10D[23] = FMNP_P10D;

FMNP_HD inline float pow10_f32(int i) {
#if defined(__CUDA_ARCH__)
  return kDevP10F[i];
#else
  return kHostP10F[i];
#endif
}
FMNP_HD inline double pow10_f64(int i) {
#if defined(__CUDA_ARCH__)
  return kDevP10D[i];
#else
  return kHostP10D[i];
#endif
}

// w * 10^q -> binary32 bits when one of the cheap exact tiers applies
FMNP_HD inline bool dec_to_f32_fast(uint64_t w, int64_t q, uint32_t* bits) {
  if (w == 0) { *bits = 0; return true; }
  if (w <= (1ull << 24) && q >= -10 && q <= 10) {
    const float f = (float)(uint32_t)w;
#if defined(__CUDA_ARCH__)
    *bits = f32_bits(q < 0 ? __fdiv_rn(f, pow10_f32((int)-q)) : __fmul_rn(f, pow10_f32((int)q)));
#else
    volatile float r = q < 0 ? f / pow10_f32((int)-q) : f * pow10_f32((int)q);
    *bits = f32_bits(r);
#endif
    return true;
  }
  if (w <= (1ull << 53) && q >= -22 && q <= 22) {
#if defined(__CUDA_ARCH__)
    const double d = q < 0 ? __ddiv_rn((double)(long long)w, pow10_f64((int)-q))
                           : __dmul_rn((double)(long long)w, pow10_f64((int)q));
    const float f = __double2float_rn(d);
#else
    volatile double d = q < 0 ? (double)w / pow10_f64((int)-q) : (double)w * pow10_f64((int)q);
    const float f = (float)d;
#endif
    // 1 <= w <= 2^53 and |q| <= 22 keep d inside binary32's normal range, where a
    // rounding boundary is a double whose low 29 mantissa bits are 1000...0
    if ((f64_bits(d) & 0x1fffffffull) == 0x10000000ull) return false;
    *bits = f32_bits(f);
    return true;
  }
  return false;
}

// ---- tier 3: exact big-integer path ---------------------------------------
enum { BIG_WORDS = 16, MAX_SIG_DIGITS = 128 };
struct Big { uint32_t w[BIG_WORDS]; };

FMNP_HD inline void big_zero(Big& a) { for (int i = 0; i < BIG_WORDS; ++i) a.w[i] = 0; }
FMNP_HD inline void big_mul_add(Big& a, uint32_t m, uint32_t add) {
  uint64_t c = add;
  for (int i = 0; i < BIG_WORDS; ++i) {
    c += (uint64_t)a.w[i] * m;
    a.w[i] = (uint32_t)c;
    c >>= 32;
  }
}
FMNP_HD inline void big_mul_pow5(Big& a, int64_t e) {
  while (e >= 13) { big_mul_add(a, 1220703125u, 0); e -= 13; }
  uint32_t m = 1;
  for (; e > 0; --e) m *= 5u;
  if (m != 1) big_mul_add(a, m, 0);
}
FMNP_HD inline int big_bits(const Big& a) {
  for (int i = BIG_WORDS - 1; i >= 0; --i)
    if (a.w[i]) return 32 * i + bit_length64(a.w[i]);
  return 0;
}
FMNP_HD inline void big_shl(Big& a, int s) {
  const int ws = s >> 5, bs = s & 31;
  for (int i = BIG_WORDS - 1; i >= 0; --i) {
    const uint32_t lo = (i - ws >= 0) ? a.w[i - ws] : 0u;
    const uint32_t lo2 = (i - ws - 1 >= 0) ? a.w[i - ws - 1] : 0u;
    a.w[i] = bs ? ((lo << bs) | (lo2 >> (32 - bs))) : lo;
  }
}
FMNP_HD inline void big_shr1(Big& a) {
  for (int i = 0; i < BIG_WORDS; ++i) {
    const uint32_t hi = (i + 1 < BIG_WORDS) ? a.w[i + 1] : 0u;
    a.w[i] = (a.w[i] >> 1) | (hi << 31);
  }
}
FMNP_HD inline bool big_ge(const Big& a, const Big& b) {
  for (int i = BIG_WORDS - 1; i >= 0; --i)
    if (a.w[i] != b.w[i]) return a.w[i] > b.w[i];
  return true;
}
FMNP_HD inline void big_sub(Big& a, const Big& b) {
  uint64_t br = 0;
  for (int i = 0; i < BIG_WORDS; ++i) {
    const uint64_t d = (uint64_t)a.w[i] - b.w[i] - br;
    a.w[i] = (uint32_t)d;
    br = (d >> 32) & 1;
  }
}
FMNP_HD inline bool big_is_zero(const Big& a) {
  uint32_t o = 0;
  for (int i = 0; i < BIG_WORDS; ++i) o |= a.w[i];
  return o == 0;
}

// digits: the text from the first significant digit to the last digit (a '.'
// may sit among them); exp10_last is the power of ten of the LAST digit.
FMNP_HD FMNP_NOINLINE uint32_t dec_to_f32_exact(const uint8_t* digits, int64_t n, int64_t exp10_last) {
  Big num;
  big_zero(num);
  int nd = 0;
  int64_t dropped = 0;
  bool sticky = false;
  uint32_t chunk = 0, chunk_mul = 1;
  for (int64_t i = 0; i < n; ++i) {
    const uint8_t c = digits[i];
    if (c == '.') continue;
    if (nd < MAX_SIG_DIGITS) {
      chunk = chunk * 10u + (uint32_t)(c - '0');
      chunk_mul *= 10u;
      ++nd;
      if (chunk_mul == 1000000000u) { big_mul_add(num, chunk_mul, chunk); chunk = 0; chunk_mul = 1; }
    } else {
      sticky |= (c != '0');
      ++dropped;
    }
  }
  if (chunk_mul != 1) big_mul_add(num, chunk_mul, chunk);
  const int64_t e = exp10_last + dropped;  // value = (num + sticky) * 10^e, 10^(nd-1) <= num < 10^nd
  if (nd + e <= -46) return 0u;            // < 1e-46 < 2^-150: rounds to zero
  if (nd - 1 + e >= 39) return 0x7f800000u;  // >= 1e39 > FLT_MAX + half ulp
  Big den;
  big_zero(den);
  den.w[0] = 1;
  if (e >= 0) big_mul_pow5(num, e); else big_mul_pow5(den, -e);
  // value = num / den * 2^e.  Scale so the quotient has 27 or 28 bits.
  const int s = 27 + big_bits(den) - big_bits(num);
  if (s >= 0) big_shl(num, s); else big_shl(den, -s);
  big_shl(den, 27);
  uint64_t q = 0;
  for (int b = 27; b >= 0; --b) {
    if (big_ge(num, den)) { big_sub(num, den); q |= 1ull << b; }
    big_shr1(den);
  }
  sticky |= !big_is_zero(num);
  return round_pack(q, e - s, sticky);
}

// glibc strtoull(base 0) on an alnum/underscore run, for the nan(payload) form
FMNP_HD inline bool nan_payload(const uint8_t* p, int64_t n, uint64_t* out) {
  int64_t i = 0;
  int base = 10;
  if (n >= 1 && p[0] == '0') {
    if (n >= 3 && lower(p[1]) == 'x' && hex_val(p[2]) >= 0) { base = 16; i = 2; }
    else base = 8;
  }
  uint64_t v = 0;
  bool any = false, ovf = false;
  for (; i < n; ++i) {
    const int d = hex_val(p[i]);
    if (d < 0 || d >= base) break;
    if (v > (0xffffffffffffffffull - (uint64_t)d) / (uint64_t)base) ovf = true;
    v = v * (uint64_t)base + (uint64_t)d;
    any = true;
  }
  if (ovf) v = 0xffffffffffffffffull;
  *out = v;
  return any && i == n;
}

// glibc strtof in the "C" locale.  Returns the number of bytes consumed
// (0 = no conversion) and the result's bit pattern.
FMNP_HD FMNP_NOINLINE int64_t strtof_bits(const uint8_t* p, int64_t n, uint32_t* bits) {
  int64_t i = 0;
  *bits = 0;
  while (is_space(at(p, n, i))) ++i;
  uint32_t sign = 0;
  if (at(p, n, i) == '-') { sign = 0x80000000u; ++i; }
  else if (at(p, n, i) == '+') ++i;
  const uint8_t c0 = at(p, n, i);

  if (lower(c0) == 'i') {
    const char* w = "infinity";
    int m = 0;
    while (m < 8 && lower(at(p, n, i + m)) == (uint8_t)w[m]) ++m;
    if (m < 3) return 0;
    *bits = sign | 0x7f800000u;
    return i + (m == 8 ? 8 : 3);
  }
  if (lower(c0) == 'n') {
    if (lower(at(p, n, i + 1)) != 'a' || lower(at(p, n, i + 2)) != 'n') return 0;
    i += 3;
    uint32_t v = 0x7fc00000u;
    if (at(p, n, i) == '(') {
      int64_t j = i + 1;
      while (is_alnum(at(p, n, j)) || at(p, n, j) == '_') ++j;
      if (at(p, n, j) == ')') {
        uint64_t payload;
        if (nan_payload(p + i + 1, j - i - 1, &payload)) v = 0x7fc00000u | (uint32_t)(payload & 0x3fffffull);
        i = j + 1;
      }
    }
    *bits = sign | v;
    return i;
  }

  // hexadecimal: 0x h* [. h*] [p [+-] d+], at least one hex digit
  if (c0 == '0' && lower(at(p, n, i + 1)) == 'x') {
    int64_t j = i + 2;
    uint64_t m = 0;
    int64_t e2 = 0;
    bool any = false, sticky = false, point = false;
    for (;; ++j) {
      const uint8_t c = at(p, n, j);
      if (c == '.' && !point) { point = true; continue; }
      const int d = hex_val(c);
      if (d < 0) break;
      any = true;
      if (m >> 60) {  // mantissa full: later digits only matter as sticky
        sticky |= d != 0;
        if (!point) e2 += 4;
      } else {
        m = (m << 4) | (uint64_t)d;
        if (point) e2 -= 4;
      }
    }
    // a lone '.' consumed above must be given back when no digit followed it at all
    if (any) {
      if (lower(at(p, n, j)) == 'p') {
        int64_t k = j + 1;
        bool eneg = false;
        if (at(p, n, k) == '-') { eneg = true; ++k; } else if (at(p, n, k) == '+') ++k;
        if (is_digit(at(p, n, k))) {
          int64_t ex = 0;
          for (; is_digit(at(p, n, k)); ++k) if (ex < (1ll << 40)) ex = ex * 10 + (at(p, n, k) - '0');
          e2 += eneg ? -ex : ex;
          j = k;
        }
      }
      if (e2 > (1ll << 41)) e2 = 1ll << 41;
      if (e2 < -(1ll << 41)) e2 = -(1ll << 41);
      *bits = sign | round_pack(m, e2, sticky);
      return j;
    }
    // "0x" with no hex digit: only the "0" converts
    *bits = sign;
    return i + 1;
  }

  // decimal: d* [. d*] [e [+-] d+], at least one digit
  int64_t j = i;
  int64_t first_sig = -1, last_digit = -1, frac_digits = 0, nd_total = 0;
  bool any = false, point = false;
  uint64_t w = 0;
  bool truncated = false;
  for (;; ++j) {
    const uint8_t c = at(p, n, j);
    if (c == '.' && !point) { point = true; continue; }
    if (!is_digit(c)) break;
    any = true;
    last_digit = j;
    if (point) ++frac_digits;
    if (first_sig < 0) {
      if (c == '0') continue;
      first_sig = j;
    }
    if (nd_total < 19) w = w * 10 + (uint64_t)(c - '0'); else truncated |= c != '0';
    ++nd_total;
  }
  if (!any) return 0;
  // j is one past the last digit or a '.' that followed digits ("5." consumes the point)
  int64_t ex = 0;
  if (lower(at(p, n, j)) == 'e') {
    int64_t k = j + 1;
    bool eneg = false;
    if (at(p, n, k) == '-') { eneg = true; ++k; } else if (at(p, n, k) == '+') ++k;
    if (is_digit(at(p, n, k))) {
      for (; is_digit(at(p, n, k)); ++k) if (ex < (1ll << 40)) ex = ex * 10 + (at(p, n, k) - '0');
      if (eneg) ex = -ex;
      j = k;
    }
  }
  if (first_sig < 0) { *bits = sign; return j; }  // all digits were zero
  int64_t exp10_last = ex - frac_digits;          // power of ten of the last digit
  if (exp10_last > 1000000) exp10_last = 1000000;
  if (exp10_last < -(1ll << 42)) exp10_last = -(1ll << 42);
  uint32_t r;
  const int64_t q = exp10_last + (nd_total > 19 ? nd_total - 19 : 0);
  if (truncated || !dec_to_f32_fast(w, q, &r))
    r = dec_to_f32_exact(p + first_sig, last_digit - first_sig + 1, exp10_last);
  *bits = sign | r;
  return j;
}

// glibc strtoll(base 10): returns bytes consumed (0 = no conversion); clamps
// to INT64_MIN / INT64_MAX on overflow like libc does.
FMNP_HD FMNP_NOINLINE int64_t strtoll10(const uint8_t* p, int64_t n, int64_t* out) {
  int64_t i = 0;
  *out = 0;
  while (is_space(at(p, n, i))) ++i;
  bool neg = false;
  if (at(p, n, i) == '-') { neg = true; ++i; } else if (at(p, n, i) == '+') ++i;
  if (!is_digit(at(p, n, i))) return 0;
  const uint64_t lim = neg ? (1ull << 63) : (1ull << 63) - 1;
  uint64_t v = 0;
  bool ovf = false;
  for (; is_digit(at(p, n, i)); ++i) {
    const uint64_t d = (uint64_t)(at(p, n, i) - '0');
    if (v > (lim - d) / 10) ovf = true;
    if (!ovf) v = v * 10 + d;
  }
  if (ovf) v = lim;
  *out = neg ? (int64_t)(0 - v) : (int64_t)v;
  return i;
}

// tensorflow::Hash64 (tensorflow/core/lib/hash/hash.cc, default seed
// 0xDECAFCAFFE), the hash cc/fm_parser_op.cc:84 applies to a feature token.
FMNP_HD inline uint64_t hash64(const uint8_t* data, int64_t n) {
  const uint64_t m = 0xc6a4a7935bd1e995ull;
  uint64_t h = 0xDECAFCAFFEull ^ ((uint64_t)n * m);
  int64_t i = 0;
  for (; i + 8 <= n; i += 8) {
    uint64_t k = 0;
    for (int b = 7; b >= 0; --b) k = (k << 8) | data[i + b];
    k *= m; k ^= k >> 47; k *= m;
    h ^= k; h *= m;
  }
  if (i < n) {
    uint64_t k = 0;
    for (int64_t b = n - 1; b >= i; --b) k = (k << 8) | data[b];
    h ^= k; h *= m;
  }
  h ^= h >> 47; h *= m; h ^= h >> 47;
  return h;
}

}  // namespace fmnp
