// Error plumbing and process-wide counters behind the C ABI.
#include <cuda_runtime.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "../include/ddls_b200.h"

namespace ddls {

static thread_local char g_err[1024] = "";
std::atomic<unsigned long long> g_launches{0};

void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int check_cuda(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return 0;
  set_last_error("%s: %s (%s)", what, cudaGetErrorString(e), cudaGetErrorName(e));
  return static_cast<int>(e) ? static_cast<int>(e) : 1;
}

void count_launch(unsigned n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

}  // namespace ddls

extern "C" const char* ddls_last_error(void) { return ddls::g_err; }
extern "C" int ddls_version(void) { return 100; }
extern "C" unsigned long long ddls_launch_count(void) { return ddls::g_launches.load(); }
