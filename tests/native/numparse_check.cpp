// Host-side fuzz of fast_tffm_b200/csrc/fm_numparse.cuh against glibc, the
// arithmetic the reference parser uses (cc/fm_parser_op.cc:69,86,98).
// usage: numparse_check <seed> <iterations>      exit 0 = no mismatch
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>

#include "../../fast_tffm_b200/csrc/fm_numparse.cuh"

static long g_bad = 0, g_n = 0, g_quirk = 0;

static void check_f(const std::string& s) {
  char* e;
  float ref = strtof(s.c_str(), &e);
  uint32_t rb;
  memcpy(&rb, &ref, 4);
  long rn = e - s.c_str();
  uint32_t gb = 0;
  // hand the scanner the bytes without a trailing NUL: it must not read past n
  long gn = (long)fmnp::strtof_bits((const uint8_t*)s.data(), (int64_t)strlen(s.c_str()), &gb);
  ++g_n;
  if (gn == rn && gb != rb && (gb & 0x7fffffffu) <= 0x00800000u && (rb & 0x7fffffffu) <= 0x00800000u) {
    // glibc 2.39's hex-float path drops the round bit of a 25-bit mantissa in the subnormal
    // range; a hex string is exact in double, so (float)strtod is the IEEE answer to hold to.
    const char* t = s.c_str();
    while (*t == ' ' || (*t >= 9 && *t <= 13) || *t == '+' || *t == '-') ++t;
    if (t[0] == '0' && (t[1] == 'x' || t[1] == 'X')) {
      float alt = (float)strtod(s.c_str(), 0);
      uint32_t ab;
      memcpy(&ab, &alt, 4);
      if (ab == gb) { ++g_quirk; return; }
    }
  }
  if (gn != rn || (rn > 0 && gb != rb)) {
    if (g_bad < 30) fprintf(stderr, "strtof MISMATCH [%s] glibc bits=%08x n=%ld  ours bits=%08x n=%ld\n", s.c_str(), rb, rn, gb, gn);
    ++g_bad;
  }
}

static void check_ll(const std::string& s) {
  char* e;
  long long ref = strtoll(s.c_str(), &e, 10);
  long rn = e - s.c_str();
  int64_t g = 0;
  long gn = (long)fmnp::strtoll10((const uint8_t*)s.data(), (int64_t)strlen(s.c_str()), &g);
  ++g_n;
  if (gn != rn || (rn > 0 && g != ref)) {
    if (g_bad < 30) fprintf(stderr, "strtoll MISMATCH [%s] glibc=%lld n=%ld ours=%lld n=%ld\n", s.c_str(), ref, rn, (long long)g, gn);
    ++g_bad;
  }
}

int main(int argc, char** argv) {
  unsigned long seed = argc > 1 ? strtoul(argv[1], 0, 10) : 1;
  long iters = argc > 2 ? atol(argv[2]) : 200000;
  std::mt19937_64 rng(seed);
  char buf[1024];
  auto rnd = [&](uint64_t n) { return (uint64_t)(rng() % n); };
  auto rand_float = [&]() {
    uint32_t b = (uint32_t)rng();
    if (((b >> 23) & 0xff) == 0xff) b &= 0xff7fffffu;
    float f;
    memcpy(&f, &b, 4);
    return f;
  };
  for (long it = 0; it < iters; ++it) {
    switch (it % 10) {
      case 0: {  // shortest-ish prints of random floats
        float f = rand_float();
        snprintf(buf, sizeof buf, "%.*g", (int)(1 + rnd(17)), (double)f);
        check_f(buf);
        snprintf(buf, sizeof buf, "%.*e", (int)rnd(12), (double)f);
        check_f(buf);
        break;
      }
      case 1: {  // exact rounding boundaries (midpoints of adjacent floats) and neighbours
        float f = fabsf(rand_float());
        if (rnd(4) == 0) { uint32_t b = (uint32_t)rnd(1u << 24); memcpy(&f, &b, 4); }  // subnormals
        float g = nextafterf(f, INFINITY);
        if (std::isinf(g)) g = f;
        double mid = ((double)f + (double)g) * 0.5;
        snprintf(buf, sizeof buf, "%.200g", mid);
        std::string s = buf;
        check_f(s);
        if (s.find('e') == std::string::npos) {
          if (s.find('.') == std::string::npos) s += ".";
          check_f(s + "0000000000000000000000000000000000001");
          check_f(s + "000");
          std::string t = s;
          for (int i = (int)t.size() - 1; i >= 0; --i) {  // decrement: just below the boundary
            if (t[i] == '.') continue;
            if (t[i] > '0') { t[i]--; break; }
            t[i] = '9';
          }
          check_f(t + "99999999999999999999999999999");
        }
        snprintf(buf, sizeof buf, "%.*e", (int)(8 + rnd(30)), mid);
        check_f(buf);
        break;
      }
      case 2: {  // random digit strings, random point, random exponent
        std::string s;
        if (rnd(3) == 0) s += rnd(2) ? "-" : "+";
        int nd = 1 + (int)rnd(rnd(8) == 0 ? 300 : 25);
        int pt = (int)rnd(nd + 2) - 1;
        for (int i = 0; i < nd; ++i) {
          if (i == pt) s += '.';
          s += (char)('0' + (rnd(5) == 0 ? 0 : rnd(10)));
        }
        if (rnd(2)) {
          s += rnd(2) ? "e" : "E";
          int ex = (int)rnd(120) - 60;
          if (rnd(10) == 0) ex *= 7;
          snprintf(buf, sizeof buf, "%s%d", (ex >= 0 && rnd(2)) ? "+" : "", ex);
          s += buf;
        }
        check_f(s);
        break;
      }
      case 3: {  // typical data values
        snprintf(buf, sizeof buf, "%.4f", (double)rnd(2000000) / 65536.0);
        check_f(buf);
        snprintf(buf, sizeof buf, "%.6g", ((double)rng() / 1.8446744073709552e19 - 0.5) * pow(10.0, (double)rnd(20) - 10));
        check_f(buf);
        snprintf(buf, sizeof buf, "%llu", (unsigned long long)(rng() >> rnd(64)));
        check_f(buf);
        break;
      }
      case 4: {  // around the double-tier boundaries: 15-19 digit mantissas, |q| near 22
        snprintf(buf, sizeof buf, "%llue%d", (unsigned long long)(rng() >> rnd(12)), (int)rnd(60) - 30);
        check_f(buf);
        uint32_t b = (uint32_t)rnd(1u << 25);
        snprintf(buf, sizeof buf, "%u.5e%d", b, (int)rnd(24) - 12);  // 25-bit integers + .5: float midpoints scaled
        check_f(buf);
        snprintf(buf, sizeof buf, "%u", 16777216u + (uint32_t)rnd(64));
        check_f(buf);
        break;
      }
      case 5: {  // overflow / underflow edges
        const char* edges[] = {"3.4028234e38", "3.4028235e38", "3.40282356e38", "3.40282357e38", "3.4028236e38",
                               "340282356779733661637539395458142568447", "340282356779733661637539395458142568448",
                               "340282356779733661637539395458142568449", "1.17549435e-38", "1.1754942e-38",
                               "1.4012984e-45", "7.0064923e-46", "7.00649232162408535461864791644958065640130970938257885878534141944895541342930300743319094181060791015625e-46",
                               "7.00649232162408535461864791644958065640130970938257885878534141944895541342930300743319094181060791015626e-46",
                               "7.00649232162408535461864791644958065640130970938257885878534141944895541342930300743319094181060791015624e-46",
                               "2.1019476964872256e-45", "1e-44", "9e-46", "0.5e-45", "1e38", "9.9999999e37"};
        check_f(edges[rnd(sizeof(edges) / sizeof(*edges))]);
        snprintf(buf, sizeof buf, "%.*ge%d", (int)(1 + rnd(12)), 1.0 + (double)rnd(9000000) / 1e6, -(int)(36 + rnd(12)));
        check_f(buf);
        snprintf(buf, sizeof buf, "%.*ge%d", (int)(1 + rnd(12)), 1.0 + (double)rnd(9000000) / 1e6, (int)(36 + rnd(4)));
        check_f(buf);
        break;
      }
      case 6: {  // hex floats
        std::string s = rnd(4) == 0 ? "-" : "";
        s += rnd(2) ? "0x" : "0X";
        int nd = (int)rnd(rnd(6) == 0 ? 40 : 10);
        int pt = (int)rnd(nd + 2) - 1;
        for (int i = 0; i < nd; ++i) {
          if (i == pt) s += '.';
          s += "0123456789abcdefABCDEF"[rnd(rnd(4) == 0 ? 1 : 22)];
        }
        if (pt >= nd && rnd(2)) s += '.';
        if (rnd(3)) {
          s += rnd(2) ? "p" : "P";
          int ex = (int)rnd(400) - 200;
          snprintf(buf, sizeof buf, "%s%d", (ex >= 0 && rnd(2)) ? "+" : "", ex);
          if (rnd(12)) s += buf;
        }
        check_f(s);
        snprintf(buf, sizeof buf, "0x1.%06xp-%d", (unsigned)rnd(1u << 24) & (rnd(2) ? 0xffffffu : 0xf00001u), (int)(126 + rnd(26)));
        check_f(buf);
        break;
      }
      case 7: {  // garbage from the grammar's alphabet
        const char al[] = "0123456789+-.eExXpPinfatyINFNA()_ \t:";
        std::string s;
        int n = (int)rnd(14);
        for (int i = 0; i < n; ++i) s += al[rnd(sizeof(al) - 1)];
        check_f(s);
        check_ll(s);
        break;
      }
      case 8: {  // special words with suffixes
        const char* w[] = {"inf", "INF", "Infinity", "infinit", "infinityx", "nan", "NaN", "nan(", "nan()", "nan(1)", "nan(0x7)",
                           "nan(017)", "nan(zz)", "nan(1_)", "nan(4194303)", "nan(4194304)", "nan(99999999999999999999999)", "-nan(5)",
                           "+nan", "nan(0x)", "nan(0xg)", "nan(00)", "nan(1", " inf", "\t-inf", "i", "n", "na", "in"};
        std::string s = w[rnd(sizeof(w) / sizeof(*w))];
        if (rnd(3) == 0) s += "0123 x:"[rnd(7)];
        check_f(s);
        break;
      }
      default: {  // strtoll
        std::string s;
        for (int i = (int)rnd(3); i > 0; --i) s += " \t\n"[rnd(3)];
        if (rnd(3) == 0) s += rnd(2) ? "-" : "+";
        int nd = (int)rnd(rnd(5) == 0 ? 25 : 12);
        for (int i = 0; i < nd; ++i) s += (char)('0' + rnd(10));
        if (rnd(4) == 0) s += ":1";
        check_ll(s);
        const char* edges[] = {"9223372036854775806", "9223372036854775807", "9223372036854775808", "-9223372036854775807",
                               "-9223372036854775808", "-9223372036854775809", "18446744073709551616", "-0", "+0"};
        check_ll(edges[rnd(sizeof(edges) / sizeof(*edges))]);
        break;
      }
    }
  }
  printf("checked %ld strings, %ld mismatches, %ld glibc hex-subnormal quirks\n", g_n, g_bad, g_quirk);
  return g_bad ? 1 : 0;
}
