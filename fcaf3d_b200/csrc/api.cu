// Library-wide state of the C-ABI: last-error string (thread-local) and version / device probes.
#include "common.cuh"
#include <stdarg.h>
#include <string.h>

static thread_local char g_err[512] = "";

void b2_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

#include <atomic>
static std::atomic<long long> g_launches{0};
void b2_count_launches(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

extern "C" {

const char* b2_last_error(void) { return g_err; }

// number of kernels this library has launched in this process so far
long long b2_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int b2_version(void) { return 100; }

// number of SMs of the current device (grid sizing of persistent kernels); <0 on error
int b2_device_sm_count(void) {
  int dev = 0, sms = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return B2_ERR_CUDA;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return B2_ERR_CUDA;
  return sms;
}

}  // extern "C"
