// A7 + A8: coarse transmittance weights, EG3D smoothing and inverse-CDF importance sampling
// (lib/nerf_utils.py:166-224, run.py:261-281).  One warp per ray, samples in registers, the CDF in
// a few hundred bytes of shared memory per warp.  HBM-streaming: 8 B read + 4 B written per sample.
//
// Bit-exact index contract: the CDF is built in the *same floating-point order as ATen's CPU
// kernels* (row sum = 8-lane vector partials with 4-way ILP; cumsum = fp64 accumulator), so that
// searchsorted(cdf, u, right=True) returns the reference's indices for identical inputs.
#include "nfi_common.cuh"
#include "nfi_warp.cuh"

namespace {

constexpr int kWarps = 8;
constexpr int kThreads = kWarps * 32;
constexpr int kMaxM = 128;

struct PdfSmem {
  float pw[kMaxM];
  float bins[kMaxM];
  float cdf[kMaxM];
};

// ATen SumKernel order for a contiguous fp32 row of nw < 512 values held in shared memory.
__device__ __forceinline__ float aten_row_sum(const float* pw, int nw, int lane) {
  const int nv = nw >> 3, rows = nv >> 2;
  float part0 = 0.f;
  if (lane < 8) {
    float p[4] = {0.f, 0.f, 0.f, 0.f};
    for (int r = 0; r < rows; ++r) {
#pragma unroll
      for (int k = 0; k < 4; ++k) p[k] = NFI_ADD(p[k], pw[((r * 4 + k) << 3) + lane]);
    }
    for (int i = rows * 4; i < nv; ++i) p[0] = NFI_ADD(p[0], pw[(i << 3) + lane]);
    p[0] = NFI_ADD(p[0], p[1]); p[0] = NFI_ADD(p[0], p[2]); p[0] = NFI_ADD(p[0], p[3]);
    part0 = p[0];
  }
  float acc = 0.f;
  for (int k = nv << 3; k < nw; ++k) acc = NFI_ADD(acc, pw[k]);
#pragma unroll
  for (int l = 0; l < 8; ++l) acc = NFI_ADD(acc, __shfl_sync(NFI_FULL, part0, l));
  return acc;   // identical on every lane
}

// sm.pw[0..M-2] (1e-5 already added) and sm.bins[0..M-1] are filled; writes n samples.
__device__ __forceinline__ void warp_sample_pdf(PdfSmem& sm, int M, int n, const float* __restrict__ u,
                                                float* __restrict__ out, int32_t* __restrict__ inds_out,
                                                int lane) {
  const int nw = M - 1;
  __syncwarp();
  const float total = aten_row_sum(sm.pw, nw, lane);
  // pdf -> fp64 inclusive prefix -> cdf[1..]
  double run = 0.0;
  double inc[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const int i = lane * 4 + e;
    const float pdf = i < nw ? NFI_DIV(sm.pw[i], total) : 0.f;
    run = NFI_DADD(run, (double)pdf);
    inc[e] = run;
  }
  const double below = nfi_warp_exclusive_sum(run, lane);
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const int i = lane * 4 + e;
    if (i < nw) sm.cdf[i + 1] = (float)NFI_DADD(below, inc[e]);
  }
  if (lane == 0) sm.cdf[0] = 0.f;
  __syncwarp();
  const float step = NFI_DIV(1.f, (float)(n - 1));
  for (int j = lane; j < n; j += 32) {
    float uj;
    if (u) uj = u[j];
    else uj = (j < n / 2) ? NFI_MUL(step, (float)j) : NFI_FMA(-step, (float)(n - 1 - j), 1.f);  // torch.linspace
    int lo = 0, hi = M;                                   // searchsorted(right=True)
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (sm.cdf[mid] <= uj) lo = mid + 1; else hi = mid;
    }
    const int b = lo - 1 > 0 ? lo - 1 : 0;
    const int a = lo < M - 1 ? lo : M - 1;
    const float c0 = sm.cdf[b], c1 = sm.cdf[a];
    const float b0 = sm.bins[b], b1 = sm.bins[a];
    float denom = NFI_SUB(c1, c0);
    if (denom < 1e-5f) denom = 1.f;
    const float t = NFI_DIV(NFI_SUB(uj, c0), denom);
    out[j] = NFI_ADD(b0, NFI_MUL(t, NFI_SUB(b1, b0)));
    if (inds_out) inds_out[j] = lo;
  }
  __syncwarp();
}

__global__ void __launch_bounds__(kThreads)
sample_pdf_kernel(const float* __restrict__ bins, const float* __restrict__ weights, int64_t R, int M,
                  int n, const float* __restrict__ u, float* __restrict__ samples,
                  int32_t* __restrict__ inds_out) {
  __shared__ PdfSmem sm[kWarps];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t nwarps = (int64_t)gridDim.x * kWarps;
  for (int64_t r = (int64_t)blockIdx.x * kWarps + warp; r < R; r += nwarps) {
    for (int i = lane; i < M; i += 32) sm[warp].bins[i] = bins[r * M + i];
    for (int i = lane; i < M - 1; i += 32) sm[warp].pw[i] = NFI_ADD(weights[r * (M - 1) + i], 1e-5f);
    warp_sample_pdf(sm[warp], M, n, u ? u + r * n : nullptr, samples + r * n,
                    inds_out ? inds_out + r * n : nullptr, lane);
  }
}

// EG3D smoothing of blocked weights: s_i = (max(w[i-1],w[i]) + max(w[i],w[i+1])) / 2 + 0.01
template <int E>
__device__ __forceinline__ void warp_smooth(const float (&w)[E], int n, int lane, float (&s)[E]) {
  const float up = __shfl_up_sync(NFI_FULL, w[E - 1], 1);
  const float dn = __shfl_down_sync(NFI_FULL, w[0], 1);
#pragma unroll
  for (int e = 0; e < E; ++e) {
    const int i = lane * E + e;
    float prev = (e > 0) ? w[e > 0 ? e - 1 : 0] : up;
    float next = (e + 1 < E) ? w[e + 1 < E ? e + 1 : e] : dn;
    if (i == 0) prev = -INFINITY;
    if (i >= n - 1) next = -INFINITY;
    s[e] = NFI_ADD(NFI_MUL(NFI_ADD(fmaxf(prev, w[e]), fmaxf(w[e], next)), 0.5f), 0.01f);
  }
}

template <int E>
__global__ void __launch_bounds__(kThreads)
weights_only_kernel(const float* __restrict__ sigma, int sigma_stride, const float* __restrict__ dirs,
                    const float* __restrict__ z, int64_t R, int S, int smooth,
                    float* __restrict__ weights) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t nwarps = (int64_t)gridDim.x * kWarps;
  for (int64_t r = (int64_t)blockIdx.x * kWarps + warp; r < R; r += nwarps) {
    float sg[E], zz[E], gap[E], al[E], keep[E], T[E], w[E];
    nfi_load_blocked<E>(sigma + r * S * sigma_stride, S, sigma_stride, lane, 0.f, sg);
    nfi_load_blocked<E>(z + r * S, S, 1, lane, 0.f, zz);
    const float d[3] = {dirs[r * 3], dirs[r * 3 + 1], dirs[r * 3 + 2]};
    nfi_warp_weights<E>(sg, zz, S, nfi_norm3(d), lane, gap, al, keep, T, w);
    if (smooth) {
      float s[E];
      warp_smooth<E>(w, S, lane, s);
      nfi_store_blocked<E>(weights + r * S, S, lane, s);
    } else {
      nfi_store_blocked<E>(weights + r * S, S, lane, w);
    }
  }
}

// run.py:261-281 in one kernel: coarse (sigma, z) -> fine depths.  S <= 64, two samples per lane.
__global__ void __launch_bounds__(kThreads)
importance_kernel(const float* __restrict__ sigma, int sigma_stride, const float* __restrict__ dirs,
                  const float* __restrict__ z_coarse, int64_t R, int S, const float* __restrict__ u,
                  float* __restrict__ z_fine) {
  __shared__ PdfSmem sm[kWarps];
  constexpr int E = 2;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t nwarps = (int64_t)gridDim.x * kWarps;
  for (int64_t r = (int64_t)blockIdx.x * kWarps + warp; r < R; r += nwarps) {
    float sg[E], zz[E], gap[E], al[E], keep[E], T[E], w[E], s[E];
    nfi_load_blocked<E>(sigma + r * S * sigma_stride, S, sigma_stride, lane, 0.f, sg);
    nfi_load_blocked<E>(z_coarse + r * S, S, 1, lane, 0.f, zz);
    const float d[3] = {dirs[r * 3], dirs[r * 3 + 1], dirs[r * 3 + 2]};
    nfi_warp_weights<E>(sg, zz, S, nfi_norm3(d), lane, gap, al, keep, T, w);
    warp_smooth<E>(w, S, lane, s);
    const float z_next_lane = __shfl_down_sync(NFI_FULL, zz[0], 1);
#pragma unroll
    for (int e = 0; e < E; ++e) {
      const int i = lane * E + e;
      if (i >= 1 && i <= S - 2) sm[warp].pw[i - 1] = NFI_ADD(s[e], 1e-5f);       // weights[..., 1:-1]
      const float zn = (e + 1 < E) ? zz[e + 1 < E ? e + 1 : e] : z_next_lane;
      if (i <= S - 2) sm[warp].bins[i] = NFI_MUL(0.5f, NFI_ADD(zn, zz[e]));       // z_vals_mid
    }
    warp_sample_pdf(sm[warp], S - 1, S, u ? u + r * S : nullptr, z_fine + r * S, nullptr, lane);
  }
}

inline int warp_grid(int64_t R) {
  const int64_t blocks = (R + kWarps - 1) / kWarps;
  const int64_t cap = (int64_t)nfi_sm_count() * 8;
  return (int)(blocks < 1 ? 1 : (blocks < cap ? blocks : cap));
}

}  // namespace

extern "C" int nfi_weights_only(const float* sigma, int sigma_stride, const float* dirs, const float* z,
                                int64_t R, int S, int smooth, float* weights, nfi_stream_t stream) {
  NFI_REQUIRE(sigma && dirs && z && weights, NFI_ERR_BAD_ARG, "nfi_weights_only: NULL pointer");
  NFI_REQUIRE(R > 0 && S >= 2 && sigma_stride >= 1, NFI_ERR_BAD_ARG, "nfi_weights_only: bad sizes");
  NFI_REQUIRE(S <= 128, NFI_ERR_UNSUPPORTED, "nfi_weights_only: S = %d > 128 samples per ray", S);
  if (S <= 64)
    weights_only_kernel<2><<<warp_grid(R), kThreads, 0, nfi_cu(stream)>>>(sigma, sigma_stride, dirs, z, R, S, smooth, weights);
  else
    weights_only_kernel<4><<<warp_grid(R), kThreads, 0, nfi_cu(stream)>>>(sigma, sigma_stride, dirs, z, R, S, smooth, weights);
  NFI_CHECK_LAUNCH("weights_only_kernel");
  return NFI_OK;
}

extern "C" int nfi_sample_pdf(const float* bins, const float* weights, int64_t R, int M, int n,
                              const float* u, float* samples, int32_t* inds_out, nfi_stream_t stream) {
  NFI_REQUIRE(bins && weights && samples, NFI_ERR_BAD_ARG, "nfi_sample_pdf: NULL pointer");
  NFI_REQUIRE(R > 0 && M >= 2 && n >= 2, NFI_ERR_BAD_ARG, "nfi_sample_pdf: bad sizes");
  NFI_REQUIRE(M <= kMaxM, NFI_ERR_UNSUPPORTED, "nfi_sample_pdf: M = %d > %d bins", M, kMaxM);
  sample_pdf_kernel<<<warp_grid(R), kThreads, 0, nfi_cu(stream)>>>(bins, weights, R, M, n, u, samples, inds_out);
  NFI_CHECK_LAUNCH("sample_pdf_kernel");
  return NFI_OK;
}

extern "C" int nfi_importance(const float* sigma, int sigma_stride, const float* dirs,
                              const float* z_coarse, int64_t R, int S, const float* u, float* z_fine,
                              nfi_stream_t stream) {
  NFI_REQUIRE(sigma && dirs && z_coarse && z_fine, NFI_ERR_BAD_ARG, "nfi_importance: NULL pointer");
  NFI_REQUIRE(R > 0 && S >= 4 && sigma_stride >= 1, NFI_ERR_BAD_ARG, "nfi_importance: bad sizes");
  NFI_REQUIRE(S <= NFI_MAX_S, NFI_ERR_UNSUPPORTED, "nfi_importance: S = %d > %d", S, NFI_MAX_S);
  importance_kernel<<<warp_grid(R), kThreads, 0, nfi_cu(stream)>>>(sigma, sigma_stride, dirs, z_coarse, R, S, u, z_fine);
  NFI_CHECK_LAUNCH("importance_kernel");
  return NFI_OK;
}
