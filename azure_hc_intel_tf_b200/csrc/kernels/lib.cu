// lib.cu — library-wide pieces: version string and the thread-local last-error message.
#include <cstdarg>

#include "common.cuh"

namespace b200 {
static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
}
}  // namespace b200

extern "C" {
const char* b200_version(void) { return "b200cnn 0.1 (sm_100a)"; }
const char* b200_last_error(void) { return b200::g_err; }
}
