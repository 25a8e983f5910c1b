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

int nfi_set_smem_once(const void* kernel, int bytes, const char* who) {
  struct Entry { const void* k; int dev; int bytes; };
  static std::mutex mu;
  static Entry done[256];
  static int n_done = 0;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return nfi_set_error(NFI_ERR_LAUNCH, "%s: cudaGetDevice: %s", who, cudaGetErrorString(e));
  std::lock_guard<std::mutex> lock(mu);
  for (int i = 0; i < n_done; ++i)
    if (done[i].k == kernel && done[i].dev == dev && done[i].bytes >= bytes) return NFI_OK;
  e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e != cudaSuccess) return nfi_set_error(NFI_ERR_LAUNCH, "%s: cudaFuncSetAttribute: %s", who, cudaGetErrorString(e));
  if (n_done < 256) done[n_done++] = Entry{kernel, dev, bytes};     // table full: fall back to setting it every time
  return NFI_OK;
}

extern "C" int nfi_version(void) { return NFI_VERSION; }
extern "C" const char* nfi_last_error(void) { return g_err; }
