// A1-A3: ray generation, ray/box near-far, stratified depths, query points (+ backward).
// Replaces lib/nerf_utils.py:28-122, 227-275 (about 100 tiny ATen launches per render) with a
// handful of streaming kernels.  All of these are HBM-bound elementwise work: one thread per
// ray (or per sample), coalesced accesses, grids sized in multiples of the SM count.
#include <mutex>

#include "nfi_common.cuh"
#include "nfi_math.cuh"

namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ------------------------------------------------------------------------------------------
// A1 forward: one thread per ray.  [R,3] outputs are written through shared memory so that the
// global stores are fully coalesced (3 floats per thread would otherwise stride by 12 bytes).
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
ray_bundle_fwd_kernel(const float* __restrict__ cam2world, const float* __restrict__ focal,
                      int focal_stride, const float* __restrict__ center,
                      const float* __restrict__ bbox, int B, int H, int W, int normalize,
                      float* __restrict__ origins, float* __restrict__ dirs) {
  __shared__ float so[kThreads * 3];
  __shared__ float sd[kThreads * 3];
  const int64_t hw = (int64_t)H * W;
  const int64_t R = (int64_t)B * hw;
  for (int64_t base = (int64_t)blockIdx.x * kThreads; base < R; base += (int64_t)gridDim.x * kThreads) {
    const int64_t r = base + threadIdx.x;
    if (r < R) {
      const int b = (int)(r / hw);
      const int pix = (int)(r - (int64_t)b * hw);
      NfiCamera c;
      nfi_load_camera(c, cam2world, focal, focal_stride, center, bbox, b);
      float o[3], d[3];
      nfi_pixel_ray(c, pix / W, pix % W, H, W, o, d);
      if (normalize) nfi_normalize3(d);
#pragma unroll
      for (int k = 0; k < 3; ++k) { so[threadIdx.x * 3 + k] = o[k]; sd[threadIdx.x * 3 + k] = d[k]; }
    }
    __syncthreads();
    const int64_t n = min((int64_t)kThreads, R - base) * 3;
    for (int i = threadIdx.x; i < n; i += kThreads) {
      origins[base * 3 + i] = so[i];
      dirs[base * 3 + i] = sd[i];
    }
    __syncthreads();
  }
}

// A1 backward: each block reduces its rays' contributions to one image's 4x4 matrix + focal.
// grid = (blocks_per_image, B); 18 values: gR[9], gt[3], gt33, gfocal.
__global__ void __launch_bounds__(kThreads)
ray_bundle_bwd_kernel(const float* __restrict__ cam2world, const float* __restrict__ focal,
                      int focal_stride, const float* __restrict__ center,
                      const float* __restrict__ bbox, int H, int W, int normalize,
                      const float* __restrict__ g_origins, const float* __restrict__ g_dirs,
                      float* __restrict__ g_cam, float* __restrict__ g_focal) {
  const int b = blockIdx.y;
  const int hw = H * W;
  NfiCamera c;
  nfi_load_camera(c, cam2world, focal, focal_stride, center, bbox, b);
  float acc[14];
#pragma unroll
  for (int i = 0; i < 14; ++i) acc[i] = 0.f;
  for (int pix = blockIdx.x * kThreads + threadIdx.x; pix < hw; pix += gridDim.x * kThreads) {
    const int64_t r = (int64_t)b * hw + pix;
    float go[3], gd[3];
    if (g_origins) { go[0] = g_origins[r * 3]; go[1] = g_origins[r * 3 + 1]; go[2] = g_origins[r * 3 + 2]; }
    if (g_dirs) { gd[0] = g_dirs[r * 3]; gd[1] = g_dirs[r * 3 + 1]; gd[2] = g_dirs[r * 3 + 2]; }
    nfi_pixel_ray_bwd(c, pix / W, pix % W, H, W, normalize, g_origins ? go : nullptr,
                      g_dirs ? gd : nullptr, acc, acc + 9, acc + 12, acc + 13);
  }
  __shared__ float part[kThreads / 32][14];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < 14; ++i) {
    const float v = warp_sum(acc[i]);
    if (lane == 0) part[warp][i] = v;
  }
  __syncthreads();
  if (threadIdx.x < 14) {
    float v = 0.f;
    for (int w = 0; w < kThreads / 32; ++w) v += part[w][threadIdx.x];
    const int i = threadIdx.x;
    if (i < 9) atomicAdd(&g_cam[b * 16 + (i / 3) * 4 + (i % 3)], v);
    else if (i < 12) atomicAdd(&g_cam[b * 16 + (i - 9) * 4 + 3], v);
    else if (i == 12) atomicAdd(&g_cam[b * 16 + 15], v);
    else atomicAdd(&g_focal[b], v);
  }
}

// ------------------------------------------------------------------------------------------
// A2: pass 1 = per-block NaN-propagating min(near) / max(far) over hitting rays + hit count;
//     pass 2 = every block folds the (<= kMaxBlocks) partials, then finishes its own rays.
// Deterministic (no float atomics); the cross-ray dependency costs one extra tiny pass.
// workspace layout: int32 hits | int32 nblocks | float pmin[kMaxBlocks] | float pmax[..] | int32 pcnt[..]
// ------------------------------------------------------------------------------------------
constexpr int kMaxBlocks = 1024;

__global__ void __launch_bounds__(kThreads)
near_far_reduce_kernel(const float* __restrict__ origins, const float* __restrict__ dirs, int64_t R,
                       float r_box, float* __restrict__ pmin, float* __restrict__ pmax,
                       int* __restrict__ pcnt) {
  float lo = INFINITY, hi = -INFINITY;
  int cnt = 0;
  for (int64_t r = (int64_t)blockIdx.x * kThreads + threadIdx.x; r < R; r += (int64_t)gridDim.x * kThreads) {
    const float o[3] = {origins[r * 3], origins[r * 3 + 1], origins[r * 3 + 2]};
    const float d[3] = {dirs[r * 3], dirs[r * 3 + 1], dirs[r * 3 + 2]};
    float n, f;
    if (nfi_slab(o, d, r_box, n, f)) { lo = nfi_min_nan(lo, n); hi = nfi_max_nan(hi, f); ++cnt; }
  }
  __shared__ float smin[kThreads / 32], smax[kThreads / 32];
  __shared__ int scnt[kThreads / 32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    lo = nfi_min_nan(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = nfi_max_nan(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  }
  if ((threadIdx.x & 31) == 0) { smin[threadIdx.x >> 5] = lo; smax[threadIdx.x >> 5] = hi; scnt[threadIdx.x >> 5] = cnt; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < kThreads / 32; ++w) { lo = nfi_min_nan(lo, smin[w]); hi = nfi_max_nan(hi, smax[w]); cnt += scnt[w]; }
    pmin[blockIdx.x] = lo; pmax[blockIdx.x] = hi; pcnt[blockIdx.x] = cnt;
  }
}

__global__ void __launch_bounds__(kThreads)
near_far_apply_kernel(const float* __restrict__ origins, const float* __restrict__ dirs, int64_t R,
                      float r_box, int nparts, const float* __restrict__ pmin,
                      const float* __restrict__ pmax, const int* __restrict__ pcnt,
                      int* __restrict__ hits_out, float* __restrict__ near, float* __restrict__ far) {
  __shared__ float smin[kThreads / 32], smax[kThreads / 32];
  __shared__ int scnt[kThreads / 32];
  float lo = INFINITY, hi = -INFINITY;
  int cnt = 0;
  for (int i = threadIdx.x; i < nparts; i += kThreads) {
    lo = nfi_min_nan(lo, pmin[i]); hi = nfi_max_nan(hi, pmax[i]); cnt += pcnt[i];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    lo = nfi_min_nan(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = nfi_max_nan(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  }
  if ((threadIdx.x & 31) == 0) { smin[threadIdx.x >> 5] = lo; smax[threadIdx.x >> 5] = hi; scnt[threadIdx.x >> 5] = cnt; }
  __syncthreads();
  lo = smin[0]; hi = smax[0]; cnt = scnt[0];
  for (int w = 1; w < kThreads / 32; ++w) { lo = nfi_min_nan(lo, smin[w]); hi = nfi_max_nan(hi, smax[w]); cnt += scnt[w]; }
  if (blockIdx.x == 0 && threadIdx.x == 0) *hits_out = cnt;
  for (int64_t r = (int64_t)blockIdx.x * kThreads + threadIdx.x; r < R; r += (int64_t)gridDim.x * kThreads) {
    const float o[3] = {origins[r * 3], origins[r * 3 + 1], origins[r * 3 + 2]};
    const float d[3] = {dirs[r * 3], dirs[r * 3 + 1], dirs[r * 3 + 2]};
    float n, f;
    const bool hit = nfi_slab(o, d, r_box, n, f);
    nfi_near_far_finish(hit, lo, hi, n, f);
    near[r] = n; far[r] = f;
  }
}

// ------------------------------------------------------------------------------------------
// A3: one thread per sample; z[r*S + k].
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
stratified_kernel(const float* __restrict__ near, const float* __restrict__ far,
                  const float* __restrict__ u, int64_t R, int S, int randomize, uint64_t seed,
                  uint64_t offset, float* __restrict__ z) {
  const int64_t N = R * S;
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < N; i += (int64_t)gridDim.x * kThreads) {
    const int64_t r = i / S;
    const int k = (int)(i - r * S);
    float uu = 0.f;
    if (randomize) {
      if (u) uu = u[i];
      else { uint32_t rnd[4]; nfi_philox(seed, (uint64_t)i, offset, rnd); uu = nfi_u01(rnd[0]); }
    }
    z[i] = nfi_stratified_z(near[r], far[r], k, S, randomize != 0, uu);
  }
}

// Same, four consecutive depths of one ray per thread (S % 4 == 0, u and z 16-byte aligned): one
// float4 load / store per thread, the ray index and (far - near) / S once per four samples, k / S from
// a per-block table.  Bit-identical to the scalar kernel (same operations per sample).
__global__ void __launch_bounds__(kThreads)
stratified4_kernel(const float* __restrict__ near, const float* __restrict__ far,
                   const float4* __restrict__ u, int64_t R, int S, int randomize, uint64_t seed,
                   uint64_t offset, float4* __restrict__ z) {
  __shared__ float wtab[NFI_MAX_S];
  for (int k = threadIdx.x; k < S; k += kThreads) wtab[k] = NFI_DIV((float)k, (float)S);
  __syncthreads();
  const int64_t N4 = R * S / 4;
  const uint32_t q = (uint32_t)S / 4;                    // float4 groups per ray
  const bool small = N4 < ((int64_t)1 << 31);
  for (int64_t i4 = (int64_t)blockIdx.x * kThreads + threadIdx.x; i4 < N4; i4 += (int64_t)gridDim.x * kThreads) {
    const int64_t r = small ? (int64_t)((uint32_t)i4 / q) : i4 / q;
    const int k0 = (int)(i4 - r * q) * 4;
    const float nr = near[r], fr = far[r];
    const float diff = NFI_SUB(fr, nr);
    float uu[4] = {0.f, 0.f, 0.f, 0.f};
    if (randomize) {
      if (u) { const float4 t = u[i4]; uu[0] = t.x; uu[1] = t.y; uu[2] = t.z; uu[3] = t.w; }
      else {
#pragma unroll
        for (int e = 0; e < 4; ++e) { uint32_t rnd[4]; nfi_philox(seed, (uint64_t)(i4 * 4 + e), offset, rnd); uu[e] = nfi_u01(rnd[0]); }
      }
    }
    const float step = NFI_DIV(diff, (float)S);
    float zz[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float w = wtab[k0 + e];
      float v = (w < 0.5f) ? NFI_FMA(w, diff, nr) : NFI_FMA(NFI_SUB(w, 1.f), diff, fr);
      if (randomize) v = NFI_ADD(v, NFI_MUL(uu[e], step));
      zz[e] = v;
    }
    z[i4] = make_float4(zz[0], zz[1], zz[2], zz[3]);
  }
}

__global__ void __launch_bounds__(kThreads)
points_fwd_kernel(const float* __restrict__ origins, const float* __restrict__ dirs,
                  const float* __restrict__ z, int64_t R, int S, float* __restrict__ pts) {
  const int64_t N = R * S * 3;
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < N; i += (int64_t)gridDim.x * kThreads) {
    const int64_t s = i / 3;
    const int c = (int)(i - s * 3);
    const int64_t r = s / S;
    pts[i] = NFI_ADD(origins[r * 3 + c], NFI_MUL(dirs[r * 3 + c], z[s]));
  }
}

// one warp per ray: g_o = sum_s g, g_d = sum_s g * z
__global__ void __launch_bounds__(kThreads)
points_bwd_kernel(const float* __restrict__ g_pts, const float* __restrict__ z, int64_t R, int S,
                  float* __restrict__ g_o, float* __restrict__ g_d) {
  const int lane = threadIdx.x & 31;
  const int64_t warps = ((int64_t)gridDim.x * kThreads) >> 5;
  for (int64_t r = ((int64_t)blockIdx.x * kThreads + threadIdx.x) >> 5; r < R; r += warps) {
    float a[3] = {0.f, 0.f, 0.f}, b[3] = {0.f, 0.f, 0.f};
    for (int k = lane; k < S; k += 32) {
      const float zz = z[r * S + k];
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float g = g_pts[(r * S + k) * 3 + c];
        a[c] += g; b[c] = fmaf(g, zz, b[c]);
      }
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) { a[c] = warp_sum(a[c]); b[c] = warp_sum(b[c]); }
    if (lane == 0) {
#pragma unroll
      for (int c = 0; c < 3; ++c) { g_o[r * 3 + c] = a[c]; g_d[r * 3 + c] = b[c]; }
    }
  }
}

inline int grid_for(int64_t work_items, int per_block) {
  const int64_t blocks = (work_items + per_block - 1) / per_block;
  const int64_t cap = (int64_t)nfi_sm_count() * 8;   // multiple of the SM count; grid-stride beyond
  return (int)(blocks < 1 ? 1 : (blocks < cap ? blocks : cap));
}

}  // namespace

extern "C" int nfi_ray_bundle_fwd(const float* cam2world, const float* focal, int focal_stride,
                                  const float* center, const float* bbox, int B, int H, int W,
                                  int normalize, float* origins, float* dirs, nfi_stream_t stream) {
  NFI_REQUIRE(cam2world && origins && dirs, NFI_ERR_BAD_ARG, "nfi_ray_bundle_fwd: NULL pointer");
  NFI_REQUIRE(B > 0 && H > 0 && W > 0, NFI_ERR_BAD_ARG, "nfi_ray_bundle_fwd: B,H,W must be positive");
  NFI_REQUIRE(focal_stride == 0 || focal_stride == 1, NFI_ERR_BAD_ARG, "nfi_ray_bundle_fwd: focal_stride must be 0 or 1");
  const int64_t R = (int64_t)B * H * W;
  ray_bundle_fwd_kernel<<<grid_for(R, kThreads), kThreads, 0, nfi_cu(stream)>>>(
      cam2world, focal, focal_stride, center, bbox, B, H, W, normalize, origins, dirs);
  NFI_CHECK_LAUNCH("ray_bundle_fwd_kernel");
  return NFI_OK;
}

extern "C" int nfi_ray_bundle_bwd(const float* cam2world, const float* focal, int focal_stride,
                                  const float* center, const float* bbox, int B, int H, int W,
                                  int normalize, const float* g_origins, const float* g_dirs,
                                  float* g_cam2world, float* g_focal, nfi_stream_t stream) {
  NFI_REQUIRE(cam2world && g_cam2world && g_focal, NFI_ERR_BAD_ARG, "nfi_ray_bundle_bwd: NULL pointer");
  NFI_REQUIRE(B > 0 && H > 0 && W > 0, NFI_ERR_BAD_ARG, "nfi_ray_bundle_bwd: B,H,W must be positive");
  cudaMemsetAsync(g_cam2world, 0, sizeof(float) * 16 * B, nfi_cu(stream));
  cudaMemsetAsync(g_focal, 0, sizeof(float) * B, nfi_cu(stream));
  int per_image = (H * W + kThreads * 4 - 1) / (kThreads * 4);
  per_image = per_image < 1 ? 1 : (per_image > 64 ? 64 : per_image);
  ray_bundle_bwd_kernel<<<dim3(per_image, B), kThreads, 0, nfi_cu(stream)>>>(
      cam2world, focal, focal_stride, center, bbox, H, W, normalize, g_origins, g_dirs, g_cam2world, g_focal);
  NFI_CHECK_LAUNCH("ray_bundle_bwd_kernel");
  return NFI_OK;
}

extern "C" int nfi_near_far(const float* origins, const float* dirs, int64_t R, float scene_range,
                            float* near, float* far, void* workspace, nfi_stream_t stream) {
  NFI_REQUIRE(origins && dirs && near && far && workspace, NFI_ERR_BAD_ARG, "nfi_near_far: NULL pointer");
  NFI_REQUIRE(R > 0, NFI_ERR_BAD_ARG, "nfi_near_far: R must be positive");
  static_assert(8 + kMaxBlocks * 12 <= NFI_NEARFAR_WS_BYTES, "workspace too small");
  int* hits = (int*)workspace;
  float* pmin = (float*)workspace + 2;
  float* pmax = pmin + kMaxBlocks;
  int* pcnt = (int*)(pmax + kMaxBlocks);
  int nb = grid_for(R, kThreads);
  if (nb > kMaxBlocks) nb = kMaxBlocks;
  near_far_reduce_kernel<<<nb, kThreads, 0, nfi_cu(stream)>>>(origins, dirs, R, scene_range, pmin, pmax, pcnt);
  NFI_CHECK_LAUNCH("near_far_reduce_kernel");
  near_far_apply_kernel<<<grid_for(R, kThreads), kThreads, 0, nfi_cu(stream)>>>(
      origins, dirs, R, scene_range, nb, pmin, pmax, pcnt, hits, near, far);
  NFI_CHECK_LAUNCH("near_far_apply_kernel");
  return NFI_OK;
}

// ------------------------------------------------------------------------------------------
// A2 without allocation: the partials of the two-pass reduction live in a file-scope device array
// instead of a caller workspace.  No counterpart in the reference (see include/nfi_b200.h).
// The array's name is only a host-side stand-in in host code: its device address is looked up once
// per device with cudaGetSymbolAddress (that device current) and passed to the kernels.
// ------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_near_far_static_scratch[NFI_NEARFAR_WS_BYTES];

static int near_far_static_scratch(void** out) {
  constexpr int kMaxDev = 64;
  static std::mutex mu;
  static void* addr[kMaxDev];           // zero-initialised: NULL = not looked up yet on that device
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess)
    return nfi_set_error(NFI_ERR_LAUNCH, "nfi_near_far_static: cudaGetDevice: %s", cudaGetErrorString(e));
  NFI_REQUIRE(dev >= 0 && dev < kMaxDev, NFI_ERR_UNSUPPORTED, "nfi_near_far_static: device id %d >= %d", dev, kMaxDev);
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_near_far_static_scratch);
    if (e != cudaSuccess || !p)
      return nfi_set_error(NFI_ERR_LAUNCH, "nfi_near_far_static: cudaGetSymbolAddress on device %d: %s", dev,
                           cudaGetErrorString(e));
    addr[dev] = p;
  }
  *out = addr[dev];
  return NFI_OK;
}

extern "C" int nfi_near_far_static(const float* origins, const float* dirs, int64_t R, float scene_range,
                                   float* near, float* far, int32_t* hits_out, nfi_stream_t stream) {
  NFI_REQUIRE(origins && dirs && near && far, NFI_ERR_BAD_ARG, "nfi_near_far_static: NULL pointer");
  NFI_REQUIRE(R > 0, NFI_ERR_BAD_ARG, "nfi_near_far_static: R must be positive");
  NFI_REQUIRE(R <= (int64_t)NFI_NEARFAR_STATIC_MAX_RAYS, NFI_ERR_BAD_ARG,
              "nfi_near_far_static: R = %lld rays exceeds NFI_NEARFAR_STATIC_MAX_RAYS = %lld; use nfi_near_far "
              "with a caller workspace", (long long)R, (long long)NFI_NEARFAR_STATIC_MAX_RAYS);
  static_assert(8 + kMaxBlocks * 12 <= NFI_NEARFAR_WS_BYTES, "static scratch too small");
  void* scratch = nullptr;
  const int rc = near_far_static_scratch(&scratch);
  if (rc != NFI_OK) return rc;
  // every word the second pass reads ([0, nb) of each partial array) is written by the first pass of
  // THIS call, so nothing left over from an earlier call is ever consumed and nothing needs zeroing
  int* hits = hits_out ? (int*)hits_out : (int*)scratch;
  float* pmin = (float*)scratch + 2;
  float* pmax = pmin + kMaxBlocks;
  int* pcnt = (int*)(pmax + kMaxBlocks);
  int nb = grid_for(R, kThreads);
  if (nb > kMaxBlocks) nb = kMaxBlocks;
  near_far_reduce_kernel<<<nb, kThreads, 0, nfi_cu(stream)>>>(origins, dirs, R, scene_range, pmin, pmax, pcnt);
  NFI_CHECK_LAUNCH("near_far_reduce_kernel (static scratch)");
  near_far_apply_kernel<<<grid_for(R, kThreads), kThreads, 0, nfi_cu(stream)>>>(
      origins, dirs, R, scene_range, nb, pmin, pmax, pcnt, hits, near, far);
  NFI_CHECK_LAUNCH("near_far_apply_kernel (static scratch)");
  return NFI_OK;
}

extern "C" int nfi_stratified(const float* near, const float* far, const float* u, int64_t R, int S,
                              int randomize, uint64_t seed, uint64_t offset, float* z,
                              nfi_stream_t stream) {
  NFI_REQUIRE(near && far && z, NFI_ERR_BAD_ARG, "nfi_stratified: NULL pointer");
  NFI_REQUIRE(R > 0 && S > 0, NFI_ERR_BAD_ARG, "nfi_stratified: R and S must be positive");
  if (S % 4 == 0 && S <= NFI_MAX_S && nfi_aligned16(z) && (!u || nfi_aligned16(u))) {
    stratified4_kernel<<<grid_for(R * S / 4, kThreads), kThreads, 0, nfi_cu(stream)>>>(
        near, far, (const float4*)u, R, S, randomize, seed, offset, (float4*)z);
  } else {
    stratified_kernel<<<grid_for(R * S, kThreads), kThreads, 0, nfi_cu(stream)>>>(
        near, far, u, R, S, randomize, seed, offset, z);
  }
  NFI_CHECK_LAUNCH("stratified_kernel");
  return NFI_OK;
}

extern "C" int nfi_points_fwd(const float* origins, const float* dirs, const float* z, int64_t R,
                              int S, float* points, nfi_stream_t stream) {
  NFI_REQUIRE(origins && dirs && z && points, NFI_ERR_BAD_ARG, "nfi_points_fwd: NULL pointer");
  NFI_REQUIRE(R > 0 && S > 0, NFI_ERR_BAD_ARG, "nfi_points_fwd: R and S must be positive");
  points_fwd_kernel<<<grid_for(R * S * 3, kThreads), kThreads, 0, nfi_cu(stream)>>>(origins, dirs, z, R, S, points);
  NFI_CHECK_LAUNCH("points_fwd_kernel");
  return NFI_OK;
}

extern "C" int nfi_points_bwd(const float* g_points, const float* z, int64_t R, int S,
                              float* g_origins, float* g_dirs, nfi_stream_t stream) {
  NFI_REQUIRE(g_points && z && g_origins && g_dirs, NFI_ERR_BAD_ARG, "nfi_points_bwd: NULL pointer");
  NFI_REQUIRE(R > 0 && S > 0, NFI_ERR_BAD_ARG, "nfi_points_bwd: R and S must be positive");
  points_bwd_kernel<<<grid_for(R * 32, kThreads), kThreads, 0, nfi_cu(stream)>>>(g_points, z, R, S, g_origins, g_dirs);
  NFI_CHECK_LAUNCH("points_bwd_kernel");
  return NFI_OK;
}
