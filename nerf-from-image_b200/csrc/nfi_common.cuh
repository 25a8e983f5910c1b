// Host-side plumbing shared by the .cu translation units: thread-local error text,
// argument checks, launch-error capture.  No global mutable state, no device sync.
#pragma once
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/nfi_b200.h"

int nfi_set_error(int code, const char* fmt, ...);

#define NFI_REQUIRE(cond, code, ...)                         \
  do {                                                       \
    if (!(cond)) return nfi_set_error((code), __VA_ARGS__);  \
  } while (0)

#define NFI_CHECK_LAUNCH(name)                                                        \
  do {                                                                                \
    cudaError_t e__ = cudaGetLastError();                                             \
    if (e__ != cudaSuccess)                                                           \
      return nfi_set_error(NFI_ERR_LAUNCH, "%s: %s", (name), cudaGetErrorString(e__)); \
  } while (0)

static inline bool nfi_aligned16(const void* p) { return (((uintptr_t)p) & 15u) == 0; }
static inline cudaStream_t nfi_cu(nfi_stream_t s) { return (cudaStream_t)s; }

// number of SMs of the *current* device, cached per device id (read-only after first use)
int nfi_sm_count();

// cudaFuncSetAttribute(kernel, MaxDynamicSharedMemorySize, bytes) once per (kernel, device) instead of on
// every launch (per-device set-up is done once).  Thread-safe; returns NFI_OK or NFI_ERR_LAUNCH.
int nfi_set_smem_once(const void* kernel, int bytes, const char* who);
