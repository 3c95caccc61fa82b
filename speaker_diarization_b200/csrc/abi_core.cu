// ABI bookkeeping: version, last-error string, launch counter.
#include <atomic>
#include <map>
#include <mutex>
#include <stdarg.h>
#include <tuple>

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

void* stream_scratch(int tag, cudaStream_t st, size_t bytes) {
  static std::mutex mu;
  static std::map<std::tuple<int, cudaStream_t, int>, std::pair<void*, size_t>> pool;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) { set_error("stream_scratch: cudaGetDevice failed"); return nullptr; }
  std::lock_guard<std::mutex> lock(mu);
  auto& e = pool[std::make_tuple(dev, st, tag)];
  if (bytes > e.second) {
    if (e.first) { cudaStreamSynchronize(st); cudaFree(e.first); e.first = nullptr; e.second = 0; }
    const size_t cap = bytes + bytes / 4;
    cudaError_t rc = cudaMalloc(&e.first, cap);
    if (rc != cudaSuccess) { set_error("stream_scratch: cudaMalloc(%zu) failed: %s", cap, cudaGetErrorString(rc)); e.first = nullptr; return nullptr; }
    e.second = cap;
  }
  return e.first;
}
}  // namespace eda

extern "C" {
int eda_abi_version(void) { return EDA_ABI_VERSION; }
const char* eda_last_error_string(void) { return eda::g_err; }
int64_t eda_launch_count(void) { return eda::g_launches.load(std::memory_order_relaxed); }
}
