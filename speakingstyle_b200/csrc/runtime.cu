// Status strings, per-thread error message and the launch counter behind the C ABI.
#include <stdarg.h>

#include <atomic>

#include "common.cuh"

namespace ss {
static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
}  // namespace ss

extern "C" {

const char* ss_status_string(int status) {
  switch (status) {
    case SS_OK: return "ok";
    case SS_ERR_INVALID: return "invalid argument";
    case SS_ERR_UNSUPPORTED: return "unsupported configuration";
    case SS_ERR_CUDA: return "CUDA error";
    case SS_ERR_ALIGNMENT: return "misaligned pointer or stride";
    default: return "unknown status";
  }
}

const char* ss_last_error(void) { return ss::g_err; }
int ss_version(void) { return 100; }
int64_t ss_launch_count(void) { return ss::g_launches.load(std::memory_order_relaxed); }
void ss_launch_count_reset(void) { ss::g_launches.store(0, std::memory_order_relaxed); }

}  // extern "C"
