// Error plumbing and process-wide counters behind the C ABI.
#include <cuda_runtime.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>

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

// programmatic dependent launch (common.cuh): on unless DDLS_PDL=0; ddls_set_pdl() overrides at run time
static std::atomic<int> g_pdl{-1};
bool pdl_enabled() {
  int v = g_pdl.load(std::memory_order_relaxed);
  if (v < 0) {
    const char* e = getenv("DDLS_PDL");
    v = (e != nullptr && e[0] == '0') ? 0 : 1;
    g_pdl.store(v, std::memory_order_relaxed);
  }
  return v != 0;
}
void set_pdl(int on) { g_pdl.store(on ? 1 : 0, std::memory_order_relaxed); }

}  // namespace ddls

extern "C" const char* ddls_last_error(void) { return ddls::g_err; }
extern "C" int ddls_version(void) { return 100; }
extern "C" int ddls_set_pdl(int on) { ddls::set_pdl(on); return 0; }
extern "C" unsigned long long ddls_launch_count(void) { return ddls::g_launches.load(); }
