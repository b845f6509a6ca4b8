// Library-level plumbing: last-error string, launch counter, arch string.
#include "common.cuh"
#include "ddl25_kernels.h"
#include <atomic>
#include <cstdarg>

namespace {
thread_local char g_err[512] = "";
std::atomic<int64_t> g_launches{0};
}  // namespace

void ddl_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void ddl_count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

extern "C" {
const char* ddl_last_error(void) { return g_err; }
int64_t ddl_launch_count(void) { return g_launches.load(); }
void ddl_reset_launch_count(void) { g_launches.store(0); }
const char* ddl_arch(void) { return "sm_100a"; }
}
