// Library-wide plumbing: last-error string, launch counter, version stamp.
#include <atomic>
#include <stdarg.h>
#include <stdio.h>

#include "../include/ttd_b200.h"
#include "common.cuh"

namespace ttd {

static thread_local char g_err[512] = "";
std::atomic<uint64_t> g_launch_count{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int check_cuda(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return 0;
  set_error("CUDA error %d (%s) at %s", static_cast<int>(e), cudaGetErrorString(e), what);
  return static_cast<int>(e);
}

}  // namespace ttd

extern "C" const char* ttd_last_error(void) { return ttd::g_err; }
extern "C" const char* ttd_version(void) { return "ttd_b200 sm_100a " __DATE__ " " __TIME__; }
extern "C" uint64_t ttd_launch_count(void) { return ttd::g_launch_count.load(); }
