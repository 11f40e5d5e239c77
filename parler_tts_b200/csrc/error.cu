#include <stdarg.h>
#include <stdlib.h>
#include "common.cuh"

namespace ptts {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
static int64_t g_launches = 0;
void count_launch() { ++g_launches; }
#ifdef PTTS_DIAG
int knob_env(const char* name, int dflt) {
  const char* e = getenv(name);
  return e ? atoi(e) : dflt;
}
#endif
static thread_local L2Hint g_hint = {nullptr, 0};
void set_l2_hint(const void* ptr, long long bytes) {
  g_hint.ptr = ptr;
  g_hint.bytes = ptr ? bytes : 0;
}
L2Hint take_l2_hint() {
  const L2Hint h = g_hint;
  g_hint = {nullptr, 0};
  return h;
}
}  // namespace ptts

extern "C" int32_t ptts_abi_version(void) { return PTTS_ABI_VERSION; }
extern "C" const char* ptts_last_error_string(void) { return ptts::g_err; }
extern "C" int64_t ptts_launch_count(void) { return ptts::g_launches; }
