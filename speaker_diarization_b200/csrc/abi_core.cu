// ABI bookkeeping: version, last-error string, launch counter.
#include <atomic>
#include <stdarg.h>

#include "common.cuh"

namespace eda {
static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
}  // namespace eda

extern "C" {
int eda_abi_version(void) { return EDA_ABI_VERSION; }
const char* eda_last_error_string(void) { return eda::g_err; }
int64_t eda_launch_count(void) { return eda::g_launches.load(std::memory_order_relaxed); }
}
