// Version / error reporting / device query for libnfi_b200.so.
#include <mutex>

#include "nfi_common.cuh"

static thread_local char g_err[512] = "";

int nfi_set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

int nfi_sm_count() {
  static int cache[64];
  static std::once_flag once[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  std::call_once(once[dev], [dev]() {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cache[dev] = n;
  });
  return cache[dev];
}

extern "C" int nfi_version(void) { return NFI_VERSION; }
extern "C" const char* nfi_last_error(void) { return g_err; }
