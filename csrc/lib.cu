// hetseq_b200 -- library-level plumbing: error string, launch counter, device attributes.
#include "common.cuh"
#include "host.h"
#include "../include/hetseq_b200.h"

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <mutex>

namespace hs {

static thread_local char g_err[512] = "";
static std::atomic<uint64_t> g_launches{0};

void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int check_cuda(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return 0;
  set_last_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
  return 1;
}

void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

int post_launch(const char* kernel) {
  count_launch(1);
  cudaError_t e = cudaPeekAtLastError();
  if (e != cudaSuccess) {
    set_last_error("launch of %s failed: %s", kernel, cudaGetErrorString(e));
    (void)cudaGetLastError();
    return 1;
  }
  return 0;
}

bool pdl_enabled() {
  static const bool on = []() {
    const char* e = getenv("HETSEQ_PDL");
    return !(e != nullptr && e[0] == '0');
  }();
  return on;
}

int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

}  // namespace hs

extern "C" const char* hs_last_error(void) { return hs::g_err; }
extern "C" int hs_version(void) { return 100; }
extern "C" uint64_t hs_launch_count(void) { return hs::g_launches.load(std::memory_order_relaxed); }
