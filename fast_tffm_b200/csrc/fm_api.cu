// fm_api.cu -- library-wide pieces of the C ABI: version and per-thread error text.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>

#include "fm_common.cuh"

static thread_local char g_error[512] = "";

int fm_fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof g_error, fmt, ap);
  va_end(ap);
  return code;
}

int fm_env_int(const char* name, int fallback) {
  const char* v = getenv(name);
  return v && *v ? atoi(v) : fallback;
}

FM_EXPORT int fm_abi_version(void) { return FM_B200_ABI_VERSION; }
FM_EXPORT const char* fm_last_error(void) { return g_error; }
