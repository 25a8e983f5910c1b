// Warp-per-ray building blocks for the scan-shaped stages (A7, A8, A9, A10).
// One warp owns one ray; lane l holds E consecutive samples (blocked layout) in registers;
// the transmittance product and the reverse (suffix) sums are warp scans over __shfl_*_sync.
// Products / prefix sums run in fp64 like ATen's CPU cumprod / cumsum accumulators.
#pragma once
#include "nfi_math.cuh"

#define NFI_FULL 0xffffffffu

__device__ __forceinline__ float nfi_warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(NFI_FULL, v, o);
  return v;
}

// exclusive prefix product over lanes (fp64); returns the product of all lower lanes
__device__ __forceinline__ double nfi_warp_exclusive_prod(double v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const double t = __shfl_up_sync(NFI_FULL, v, o);
    if (lane >= o) v = NFI_DMUL(t, v);
  }
  const double up = __shfl_up_sync(NFI_FULL, v, 1);
  return lane == 0 ? 1.0 : up;
}

// exclusive prefix sum over lanes (fp64)
__device__ __forceinline__ double nfi_warp_exclusive_sum(double v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const double t = __shfl_up_sync(NFI_FULL, v, o);
    if (lane >= o) v = NFI_DADD(t, v);
  }
  const double up = __shfl_up_sync(NFI_FULL, v, 1);
  return lane == 0 ? 0.0 : up;
}

// exclusive SUFFIX sum over lanes (fp32): sum of all higher lanes
__device__ __forceinline__ float nfi_warp_exclusive_suffix_sum(float v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const float t = __shfl_down_sync(NFI_FULL, v, o);
    if (lane + o < 32) v += t;
  }
  const float dn = __shfl_down_sync(NFI_FULL, v, 1);
  return lane == 31 ? 0.f : dn;
}

// load E consecutive values of a row (blocked layout), `fill` beyond n
template <int E>
__device__ __forceinline__ void nfi_load_blocked(const float* __restrict__ row, int n, int stride,
                                                 int lane, float fill, float (&v)[E]) {
  if constexpr (E == 4) {
    if (stride == 1 && (n & 3) == 0 && ((reinterpret_cast<uintptr_t>(row) & 15u) == 0)) {
      if (lane * 4 < n) {
        const float4 t = *reinterpret_cast<const float4*>(row + lane * 4);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
      } else {
#pragma unroll
        for (int e = 0; e < E; ++e) v[e] = fill;
      }
      return;
    }
  }
#pragma unroll
  for (int e = 0; e < E; ++e) {
    const int i = lane * E + e;
    v[e] = i < n ? row[(int64_t)i * stride] : fill;
  }
}

template <int E>
__device__ __forceinline__ void nfi_store_blocked(float* __restrict__ row, int n, int lane,
                                                  const float (&v)[E]) {
#pragma unroll
  for (int e = 0; e < E; ++e) {
    const int i = lane * E + e;
    if (i < n) row[i] = v[e];
  }
}

// Alpha / transmittance / weights of one ray (lib/nerf_utils.py:137-146, 176-181, 20-25).
// z, sigma in blocked layout; samples >= n are padding.  Outputs per element:
//   gap   = z[i+1]-z[i] (0 for the last sample), alpha, keep = 1-alpha+1e-10, T (exclusive
//   product, fp64 accumulate), w = alpha*T.
template <int E>
__device__ __forceinline__ void nfi_warp_weights(const float (&sigma)[E], const float (&z)[E], int n,
                                                 float dnorm, int lane, float (&gap)[E],
                                                 float (&alpha)[E], float (&keep)[E], float (&T)[E],
                                                 float (&w)[E]) {
  const float z_next_lane = __shfl_down_sync(NFI_FULL, z[0], 1);
  double local = 1.0;
  double pre[E];
#pragma unroll
  for (int e = 0; e < E; ++e) {
    const int i = lane * E + e;
    const float zn = (e + 1 < E) ? z[(e + 1 < E) ? e + 1 : e] : z_next_lane;
    gap[e] = (i < n - 1) ? NFI_SUB(zn, z[e]) : 0.f;
    const float dist = NFI_MUL(gap[e], dnorm);
    alpha[e] = (i < n) ? NFI_SUB(1.f, expf(NFI_MUL(-sigma[e], dist))) : 0.f;
    keep[e] = (i < n) ? NFI_ADD(NFI_SUB(1.f, alpha[e]), 1e-10f) : 1.f;
    pre[e] = local;
    local = NFI_DMUL(local, (double)keep[e]);
  }
  const double below = nfi_warp_exclusive_prod(local, lane);
#pragma unroll
  for (int e = 0; e < E; ++e) {
    T[e] = (float)NFI_DMUL(below, pre[e]);
    w[e] = NFI_MUL(alpha[e], T[e]);
  }
}
