// A9 + A10: depth merge of coarse + fine samples and alpha compositing with cumulative
// transmittance (run.py:283-335, lib/nerf_utils.py:125-163), forward and backward.
// One warp per ray, four samples per lane, transmittance = fp64 warp product scan over
// __shfl_up_sync, backward = warp suffix scan over __shfl_down_sync.  HBM-streaming kernels.
#include "nfi_common.cuh"
#include "nfi_warp.cuh"

namespace {

constexpr int kWarps = 8;
constexpr int kThreads = kWarps * 32;
constexpr int E = 4;   // samples per lane -> up to 128 samples per ray

inline int warp_grid(int64_t R) {
  const int64_t blocks = (R + kWarps - 1) / kWarps;
  const int64_t cap = (int64_t)nfi_sm_count() * 8;
  return (int)(blocks < 1 ? 1 : (blocks < cap ? blocks : cap));
}

// rgb-like [.., S, 3] rows in blocked layout: v[e][c]
__device__ __forceinline__ void load_rgb_blocked(const float* __restrict__ row, int S, int lane,
                                                 float (&v)[E][3]) {
  if ((S & 3) == 0 && ((reinterpret_cast<uintptr_t>(row) & 15u) == 0)) {
    if (lane * 4 < S) {
      const float4* p = reinterpret_cast<const float4*>(row + lane * 12);
      const float4 a = p[0], b = p[1], c = p[2];
      v[0][0] = a.x; v[0][1] = a.y; v[0][2] = a.z; v[1][0] = a.w;
      v[1][1] = b.x; v[1][2] = b.y; v[2][0] = b.z; v[2][1] = b.w;
      v[2][2] = c.x; v[3][0] = c.y; v[3][1] = c.z; v[3][2] = c.w;
    } else {
#pragma unroll
      for (int e = 0; e < E; ++e) v[e][0] = v[e][1] = v[e][2] = 0.f;
    }
    return;
  }
#pragma unroll
  for (int e = 0; e < E; ++e) {
    const int i = lane * E + e;
#pragma unroll
    for (int c = 0; c < 3; ++c) v[e][c] = i < S ? row[i * 3 + c] : 0.f;
  }
}

__device__ __forceinline__ void store_rgb_blocked(float* __restrict__ row, int S, int lane,
                                                  const float (&v)[E][3]) {
  if ((S & 3) == 0 && ((reinterpret_cast<uintptr_t>(row) & 15u) == 0)) {
    if (lane * 4 < S) {
      float4* p = reinterpret_cast<float4*>(row + lane * 12);
      p[0] = make_float4(v[0][0], v[0][1], v[0][2], v[1][0]);
      p[1] = make_float4(v[1][1], v[1][2], v[2][0], v[2][1]);
      p[2] = make_float4(v[2][2], v[3][0], v[3][1], v[3][2]);
    }
    return;
  }
#pragma unroll
  for (int e = 0; e < E; ++e) {
    const int i = lane * E + e;
    if (i < S) {
#pragma unroll
      for (int c = 0; c < 3; ++c) row[i * 3 + c] = v[e][c];
    }
  }
}

// ------------------------------------------------------------------------------------------
// standalone A10
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
composite_fwd_kernel(const float* __restrict__ sigma, const float* __restrict__ rgb,
                     const float* __restrict__ dirs, const float* __restrict__ z,
                     const float* __restrict__ normals, const float* __restrict__ sem, int n_sem,
                     int64_t R, int S, int white, float* __restrict__ rgb_map,
                     float* __restrict__ depth_map, float* __restrict__ mask_out,
                     float* __restrict__ normal_map, float* __restrict__ sem_map) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t nwarps = (int64_t)gridDim.x * kWarps;
  for (int64_t r = (int64_t)blockIdx.x * kWarps + warp; r < R; r += nwarps) {
    float sg[E], zz[E], gap[E], al[E], keep[E], T[E], w[E], col[E][3];
    nfi_load_blocked<E>(sigma + r * S, S, 1, lane, 0.f, sg);
    nfi_load_blocked<E>(z + r * S, S, 1, lane, 0.f, zz);
    load_rgb_blocked(rgb + r * S * 3, S, lane, col);
    const float d[3] = {dirs[r * 3], dirs[r * 3 + 1], dirs[r * 3 + 2]};
    nfi_warp_weights<E>(sg, zz, S, nfi_norm3(d), lane, gap, al, keep, T, w);
    float acc[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int e = 0; e < E; ++e) {
      acc[0] = fmaf(w[e], col[e][0], acc[0]);
      acc[1] = fmaf(w[e], col[e][1], acc[1]);
      acc[2] = fmaf(w[e], col[e][2], acc[2]);
      acc[3] = fmaf(w[e], zz[e], acc[3]);
      acc[4] += w[e];
    }
#pragma unroll
    for (int i = 0; i < 5; ++i) acc[i] = nfi_warp_sum(acc[i]);
    const float bg = white ? 1.f - acc[4] : 0.f;
    if (lane == 0) {
      rgb_map[r * 3 + 0] = acc[0] + bg; rgb_map[r * 3 + 1] = acc[1] + bg; rgb_map[r * 3 + 2] = acc[2] + bg;
      depth_map[r] = acc[3];
      mask_out[r] = acc[4];
    }
    if (normals) {
      load_rgb_blocked(normals + r * S * 3, S, lane, col);
      float na[3] = {0.f, 0.f, 0.f};
#pragma unroll
      for (int e = 0; e < E; ++e)
        for (int c = 0; c < 3; ++c) na[c] = fmaf(w[e], col[e][c], na[c]);
      for (int c = 0; c < 3; ++c) na[c] = nfi_warp_sum(na[c]);
      if (lane == 0)
        for (int c = 0; c < 3; ++c) normal_map[r * 3 + c] = na[c] + bg;
    }
    if (sem) {
      for (int k = 0; k < n_sem; ++k) {
        float a = 0.f;
#pragma unroll
        for (int e = 0; e < E; ++e) {
          const int i = lane * E + e;
          if (i < S) a = fmaf(w[e], sem[(r * S + i) * n_sem + k], a);
        }
        a = nfi_warp_sum(a);
        if (lane == 0) sem_map[r * n_sem + k] = a;
      }
    }
  }
}

// shared by the standalone and the merged backward: given per-sample dL/dw and the forward
// quantities, produce d sigma and the per-ray d|d| term.
__device__ __forceinline__ void weights_bwd(const float (&dLdw)[E], const float (&sg)[E],
                                            const float (&gap)[E], const float (&al)[E],
                                            const float (&keep)[E], const float (&T)[E],
                                            const float (&w)[E], float dnorm, int lane,
                                            float (&dsigma)[E], float& g_dnorm_lane) {
  float suf[E];
  float run = 0.f;
#pragma unroll
  for (int e = E - 1; e >= 0; --e) { suf[e] = run; run = fmaf(dLdw[e], w[e], run); }
  const float above = nfi_warp_exclusive_suffix_sum(run, lane);
  g_dnorm_lane = 0.f;
#pragma unroll
  for (int e = 0; e < E; ++e) {
    const float da = dLdw[e] * T[e] - (above + suf[e]) / keep[e];
    const float one_m_a = 1.f - al[e];
    dsigma[e] = da * (gap[e] * dnorm) * one_m_a;
    g_dnorm_lane = fmaf(da * sg[e] * one_m_a, gap[e], g_dnorm_lane);
  }
}

__global__ void __launch_bounds__(kThreads)
composite_bwd_kernel(const float* __restrict__ sigma, const float* __restrict__ rgb,
                     const float* __restrict__ dirs, const float* __restrict__ z,
                     const float* __restrict__ sem, int n_sem, int64_t R, int S, int white,
                     const float* __restrict__ g_rgb_map, const float* __restrict__ g_mask,
                     const float* __restrict__ g_sem_map, float* __restrict__ g_sigma,
                     float* __restrict__ g_rgb, float* __restrict__ g_dirs,
                     float* __restrict__ g_sem) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t nwarps = (int64_t)gridDim.x * kWarps;
  for (int64_t r = (int64_t)blockIdx.x * kWarps + warp; r < R; r += nwarps) {
    float sg[E], zz[E], gap[E], al[E], keep[E], T[E], w[E], col[E][3];
    nfi_load_blocked<E>(sigma + r * S, S, 1, lane, 0.f, sg);
    nfi_load_blocked<E>(z + r * S, S, 1, lane, 0.f, zz);
    load_rgb_blocked(rgb + r * S * 3, S, lane, col);
    const float d[3] = {dirs[r * 3], dirs[r * 3 + 1], dirs[r * 3 + 2]};
    const float dnorm = nfi_norm3(d);
    nfi_warp_weights<E>(sg, zz, S, dnorm, lane, gap, al, keep, T, w);
    const float G[3] = {g_rgb_map ? g_rgb_map[r * 3] : 0.f, g_rgb_map ? g_rgb_map[r * 3 + 1] : 0.f,
                        g_rgb_map ? g_rgb_map[r * 3 + 2] : 0.f};
    const float Gm = (g_mask ? g_mask[r] : 0.f) - (white ? (G[0] + G[1] + G[2]) : 0.f);
    float dLdw[E];
#pragma unroll
    for (int e = 0; e < E; ++e) dLdw[e] = fmaf(G[0], col[e][0], fmaf(G[1], col[e][1], fmaf(G[2], col[e][2], Gm)));
    if (sem && g_sem_map) {
      for (int k = 0; k < n_sem; ++k) {
        const float gk = g_sem_map[r * n_sem + k];
#pragma unroll
        for (int e = 0; e < E; ++e) {
          const int i = lane * E + e;
          if (i < S) {
            dLdw[e] = fmaf(gk, sem[(r * S + i) * n_sem + k], dLdw[e]);
            if (g_sem) g_sem[(r * S + i) * n_sem + k] = w[e] * gk;
          }
        }
      }
    }
    float ds[E], gdn;
    weights_bwd(dLdw, sg, gap, al, keep, T, w, dnorm, lane, ds, gdn);
    nfi_store_blocked<E>(g_sigma + r * S, S, lane, ds);
    if (g_rgb) {
#pragma unroll
      for (int e = 0; e < E; ++e)
        for (int c = 0; c < 3; ++c) col[e][c] = w[e] * G[c];
      store_rgb_blocked(g_rgb + r * S * 3, S, lane, col);
    }
    if (g_dirs) {
      gdn = nfi_warp_sum(gdn);
      if (lane == 0) {
        const float inv = dnorm > 0.f ? gdn / dnorm : 0.f;
        g_dirs[r * 3] = inv * d[0]; g_dirs[r * 3 + 1] = inv * d[1]; g_dirs[r * 3 + 2] = inv * d[2];
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// A9 + A10 merged.  Rank-sort of the 2S depths in shared memory (stable: ties keep the
// cat(coarse, fine) order), then the same scan as above on the permuted samples.
// ------------------------------------------------------------------------------------------
struct alignas(16) MergeSmem {
  float zcat[128];
  float zs[128];
  uint8_t perm[128];
  float zf_sorted[64];     // fast path scratch: the fine depths in ascending order
  uint8_t chk[64];         // ... and who claimed each of their ranks (tie detection)
};

// number of leading elements of the ascending a[0..63] that satisfy (a[i] <= key) / (a[i] < key)
template <bool LE>
__device__ __forceinline__ int count_leading64(const float* a, float key) {
  int pos = 0;
#pragma unroll
  for (int step = 64; step > 0; step >>= 1) {
    const int q = pos + step;
    const float v = a[q <= 64 ? q - 1 : 63];
    const bool ok = q <= 64 && (LE ? v <= key : v < key);
    pos = ok ? q : pos;
  }
  return pos;
}

// Ranks of the 64 coarse + 64 fine depths of one ray in the stable ascending order of
// cat(coarse, fine) (run.py:283-285), without the 128 x 128 comparison: the coarse depths are
// ascending by construction (stratified bins), so
//   rank(coarse i) = i + #{fine < zc[i]},   rank(fine j) = #{coarse <= zf[j]} + rank of j among the fine,
// the fine-among-fine ranks by a 64 x 64 strict comparison, the cross counts by binary search.
// Returns false -- nothing written that matters -- when the coarse list is not ascending or two
// fine depths are exactly equal; the caller then takes the general path.
__device__ __forceinline__ bool merge_ranks_fast64(MergeSmem& m, int lane) {
  const float2 c = reinterpret_cast<const float2*>(m.zcat)[lane];
  const float2 f = reinterpret_cast<const float2*>(m.zcat + 64)[lane];
  const float cnext = lane < 31 ? m.zcat[2 * lane + 2] : INFINITY;
  if (!__all_sync(NFI_FULL, c.x <= c.y && c.y <= cnext)) return false;
  int r0 = 0, r1 = 0;
#pragma unroll
  for (int k4 = 0; k4 < 16; ++k4) {
    const float4 v = reinterpret_cast<const float4*>(m.zcat + 64)[k4];
    r0 += (v.x < f.x) + (v.y < f.x) + (v.z < f.x) + (v.w < f.x);
    r1 += (v.x < f.y) + (v.y < f.y) + (v.z < f.y) + (v.w < f.y);
  }
  const int j0 = 2 * lane, j1 = j0 + 1;
  m.chk[r0] = (uint8_t)j0; m.zf_sorted[r0] = f.x;
  __syncwarp();
  m.chk[r1] = (uint8_t)j1; m.zf_sorted[r1] = f.y;
  __syncwarp();
  const bool distinct = m.chk[r0] == j0 && m.chk[r1] == j1 && r0 != r1;
  if (!__all_sync(NFI_FULL, distinct)) return false;
  const int rc0 = j0 + count_leading64<false>(m.zf_sorted, c.x);
  const int rc1 = j1 + count_leading64<false>(m.zf_sorted, c.y);
  const int rf0 = r0 + count_leading64<true>(m.zcat, f.x);
  const int rf1 = r1 + count_leading64<true>(m.zcat, f.y);
  m.zs[rc0] = c.x; m.perm[rc0] = (uint8_t)j0;
  m.zs[rc1] = c.y; m.perm[rc1] = (uint8_t)j1;
  m.zs[rf0] = f.x; m.perm[rf0] = (uint8_t)(64 + j0);
  m.zs[rf1] = f.y; m.perm[rf1] = (uint8_t)(64 + j1);
  return true;
}

__global__ void __launch_bounds__(kThreads)
merge_composite_fwd_kernel(const float4* __restrict__ coarse, const float4* __restrict__ fine,
                           const float* __restrict__ z_coarse, const float* __restrict__ z_fine,
                           const float* __restrict__ dirs, int64_t R, int S, int white,
                           const float* __restrict__ cam2world, int64_t rays_per_image,
                           float* __restrict__ rgb_map, float* __restrict__ depth_map,
                           float* __restrict__ mask_out, uint8_t* __restrict__ perm_out,
                           float* __restrict__ z_sorted) {
  __shared__ MergeSmem sm[kWarps];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n = 2 * S;
  const int64_t nwarps = (int64_t)gridDim.x * kWarps;
  for (int64_t r = (int64_t)blockIdx.x * kWarps + warp; r < R; r += nwarps) {
    MergeSmem& m = sm[warp];
    for (int i = lane; i < S; i += 32) { m.zcat[i] = z_coarse[r * S + i]; m.zcat[S + i] = z_fine[r * S + i]; }
    __syncwarp();
    if (!(S == 64 && merge_ranks_fast64(m, lane))) {     // general path: rank by comparison with all 2S
      float mine[E];
      int rank[E];
#pragma unroll
      for (int e = 0; e < E; ++e) { const int j = lane * E + e; mine[e] = j < n ? m.zcat[j] : INFINITY; rank[e] = 0; }
      for (int k = 0; k < n; ++k) {
        const float zk = m.zcat[k];
#pragma unroll
        for (int e = 0; e < E; ++e) {
          const int j = lane * E + e;
          rank[e] += (zk < mine[e] || (zk == mine[e] && k < j)) ? 1 : 0;
        }
      }
#pragma unroll
      for (int e = 0; e < E; ++e) {
        const int j = lane * E + e;
        if (j < n) { m.zs[rank[e]] = mine[e]; m.perm[rank[e]] = (uint8_t)j; }
      }
    }
    __syncwarp();
    float sg[E], zz[E], gap[E], al[E], keep[E], T[E], w[E], col[E][3];
#pragma unroll
    for (int e = 0; e < E; ++e) {
      const int i = lane * E + e;
      if (i < n) {
        const int src = m.perm[i];
        const float4 p = src < S ? coarse[r * S + src] : fine[r * S + (src - S)];
        sg[e] = p.x; col[e][0] = p.y; col[e][1] = p.z; col[e][2] = p.w; zz[e] = m.zs[i];
        if (n != 32 * E) {
          if (perm_out) perm_out[r * n + i] = (uint8_t)src;
          if (z_sorted) z_sorted[r * n + i] = zz[e];
        }
      } else { sg[e] = 0.f; col[e][0] = col[e][1] = col[e][2] = 0.f; zz[e] = 0.f; }
    }
    if (n == 32 * E) {                         // full rows: one 4-byte / 16-byte store per lane
      if (perm_out) reinterpret_cast<uchar4*>(perm_out + r * n)[lane] = reinterpret_cast<const uchar4*>(m.perm)[lane];
      if (z_sorted) reinterpret_cast<float4*>(z_sorted + r * n)[lane] = make_float4(zz[0], zz[1], zz[2], zz[3]);
    }
    const float d[3] = {dirs[r * 3], dirs[r * 3 + 1], dirs[r * 3 + 2]};
    nfi_warp_weights<E>(sg, zz, n, nfi_norm3(d), lane, gap, al, keep, T, w);
    float acc[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int e = 0; e < E; ++e) {
      acc[0] = fmaf(w[e], col[e][0], acc[0]);
      acc[1] = fmaf(w[e], col[e][1], acc[1]);
      acc[2] = fmaf(w[e], col[e][2], acc[2]);
      acc[3] = fmaf(w[e], zz[e], acc[3]);
      acc[4] += w[e];
    }
#pragma unroll
    for (int i = 0; i < 5; ++i) acc[i] = nfi_warp_sum(acc[i]);
    if (lane == 0) {
      const float bg = white ? 1.f - acc[4] : 0.f;
      rgb_map[r * 3 + 0] = acc[0] + bg; rgb_map[r * 3 + 1] = acc[1] + bg; rgb_map[r * 3 + 2] = acc[2] + bg;
      float depth = acc[3];
      if (cam2world) {   // eval_kitti_persp.py:221-228: z-buffer depth = -(R_w2c d)_z * depth
        const float* mtx = cam2world + (r / rays_per_image) * 16;
        const float t33 = mtx[15];
        const float vz = NFI_ADD(NFI_ADD(NFI_MUL(d[0], NFI_DIV(mtx[2], t33)), NFI_MUL(d[1], NFI_DIV(mtx[6], t33))),
                                 NFI_MUL(d[2], NFI_DIV(mtx[10], t33)));
        depth = NFI_MUL(NFI_MUL(vz, depth), -1.f);
      }
      depth_map[r] = depth;
      mask_out[r] = acc[4];
    }
    __syncwarp();
  }
}

__global__ void __launch_bounds__(kThreads)
merge_composite_bwd_kernel(const float4* __restrict__ coarse, const float4* __restrict__ fine,
                           const float* __restrict__ z_sorted, const uint8_t* __restrict__ perm,
                           const float* __restrict__ dirs, int64_t R, int S, int white,
                           const float* __restrict__ g_rgb_map, const float* __restrict__ g_mask,
                           float4* __restrict__ g_coarse, float4* __restrict__ g_fine) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n = 2 * S;
  const int64_t nwarps = (int64_t)gridDim.x * kWarps;
  for (int64_t r = (int64_t)blockIdx.x * kWarps + warp; r < R; r += nwarps) {
    float sg[E], zz[E], gap[E], al[E], keep[E], T[E], w[E], col[E][3];
    int src[E];
    if (n == 32 * E) {                         // full rows: one 4-byte / 16-byte load per lane
      const uchar4 pp = reinterpret_cast<const uchar4*>(perm + r * n)[lane];
      const float4 zv = reinterpret_cast<const float4*>(z_sorted + r * n)[lane];
      src[0] = pp.x; src[1] = pp.y; src[2] = pp.z; src[3] = pp.w;
      zz[0] = zv.x; zz[1] = zv.y; zz[2] = zv.z; zz[3] = zv.w;
    }
#pragma unroll
    for (int e = 0; e < E; ++e) {
      const int i = lane * E + e;
      if (i < n) {
        if (n != 32 * E) { src[e] = perm[r * n + i]; zz[e] = z_sorted[r * n + i]; }
        const float4 p = src[e] < S ? coarse[r * S + src[e]] : fine[r * S + (src[e] - S)];
        sg[e] = p.x; col[e][0] = p.y; col[e][1] = p.z; col[e][2] = p.w;
      } else { src[e] = -1; sg[e] = 0.f; col[e][0] = col[e][1] = col[e][2] = 0.f; zz[e] = 0.f; }
    }
    const float d[3] = {dirs[r * 3], dirs[r * 3 + 1], dirs[r * 3 + 2]};
    const float dnorm = nfi_norm3(d);
    nfi_warp_weights<E>(sg, zz, n, dnorm, lane, gap, al, keep, T, w);
    const float G[3] = {g_rgb_map[r * 3], g_rgb_map[r * 3 + 1], g_rgb_map[r * 3 + 2]};
    const float Gm = (g_mask ? g_mask[r] : 0.f) - (white ? (G[0] + G[1] + G[2]) : 0.f);
    float dLdw[E];
#pragma unroll
    for (int e = 0; e < E; ++e) dLdw[e] = fmaf(G[0], col[e][0], fmaf(G[1], col[e][1], fmaf(G[2], col[e][2], Gm)));
    float ds[E], gdn;
    weights_bwd(dLdw, sg, gap, al, keep, T, w, dnorm, lane, ds, gdn);
#pragma unroll
    for (int e = 0; e < E; ++e) {
      if (src[e] >= 0) {
        const float4 g = make_float4(ds[e], w[e] * G[0], w[e] * G[1], w[e] * G[2]);
        if (src[e] < S) g_coarse[r * S + src[e]] = g; else g_fine[r * S + (src[e] - S)] = g;
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// cumprod_exclusive (lib/nerf_utils.py:20-25): out[i] = prod_{j<i} x[j], out[0] = 1.  One warp per
// row, 128 elements per pass with a carried fp64 running product (ATen's CPU accumulator type).
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
cumprod_exclusive_kernel(const float* __restrict__ x, int64_t R, int n, float* __restrict__ out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t nwarps = (int64_t)gridDim.x * kWarps;
  for (int64_t r = (int64_t)blockIdx.x * kWarps + warp; r < R; r += nwarps) {
    double carry = 1.0;
    for (int base = 0; base < n; base += 32 * E) {
      float v[E];
      double pre[E];
      double local = 1.0;
#pragma unroll
      for (int e = 0; e < E; ++e) {
        const int i = base + lane * E + e;
        v[e] = i < n ? x[r * n + i] : 1.f;
        pre[e] = local;
        local = NFI_DMUL(local, (double)v[e]);
      }
      const double below = NFI_DMUL(carry, nfi_warp_exclusive_prod(local, lane));
#pragma unroll
      for (int e = 0; e < E; ++e) {
        const int i = base + lane * E + e;
        if (i < n) out[r * n + i] = (float)NFI_DMUL(below, pre[e]);
      }
      carry = __shfl_sync(NFI_FULL, NFI_DMUL(below, local), 31);
    }
  }
}

}  // namespace

extern "C" int nfi_cumprod_exclusive(const float* x, int64_t R, int n, float* out, nfi_stream_t stream) {
  NFI_REQUIRE(x && out, NFI_ERR_BAD_ARG, "nfi_cumprod_exclusive: NULL pointer");
  NFI_REQUIRE(R > 0 && n > 0, NFI_ERR_BAD_ARG, "nfi_cumprod_exclusive: bad sizes");
  cumprod_exclusive_kernel<<<warp_grid(R), kThreads, 0, nfi_cu(stream)>>>(x, R, n, out);
  NFI_CHECK_LAUNCH("cumprod_exclusive_kernel");
  return NFI_OK;
}

extern "C" int nfi_composite_fwd(const float* sigma, const float* rgb, const float* dirs, const float* z,
                                 const float* normals, const float* semantics, int n_sem, int64_t R,
                                 int S, int white_background, float* rgb_map, float* depth_map,
                                 float* mask, float* normal_map, float* semantic_map,
                                 nfi_stream_t stream) {
  NFI_REQUIRE(sigma && rgb && dirs && z && rgb_map && depth_map && mask, NFI_ERR_BAD_ARG, "nfi_composite_fwd: NULL pointer");
  NFI_REQUIRE(R > 0 && S >= 1, NFI_ERR_BAD_ARG, "nfi_composite_fwd: bad sizes");
  NFI_REQUIRE(S <= 128, NFI_ERR_UNSUPPORTED, "nfi_composite_fwd: S = %d > 128 samples per ray", S);
  NFI_REQUIRE(!normals || normal_map, NFI_ERR_BAD_ARG, "nfi_composite_fwd: normals without normal_map");
  NFI_REQUIRE(!semantics || (semantic_map && n_sem > 0), NFI_ERR_BAD_ARG, "nfi_composite_fwd: semantics without semantic_map");
  composite_fwd_kernel<<<warp_grid(R), kThreads, 0, nfi_cu(stream)>>>(
      sigma, rgb, dirs, z, normals, semantics, n_sem, R, S, white_background, rgb_map, depth_map, mask,
      normal_map, semantic_map);
  NFI_CHECK_LAUNCH("composite_fwd_kernel");
  return NFI_OK;
}

extern "C" int nfi_composite_bwd(const float* sigma, const float* rgb, const float* dirs, const float* z,
                                 const float* semantics, int n_sem, int64_t R, int S,
                                 int white_background, const float* g_rgb_map, const float* g_mask,
                                 const float* g_semantic_map, float* g_sigma, float* g_rgb,
                                 float* g_dirs, float* g_semantics, nfi_stream_t stream) {
  NFI_REQUIRE(sigma && rgb && dirs && z && g_sigma, NFI_ERR_BAD_ARG, "nfi_composite_bwd: NULL pointer");
  NFI_REQUIRE(R > 0 && S >= 1, NFI_ERR_BAD_ARG, "nfi_composite_bwd: bad sizes");
  NFI_REQUIRE(S <= 128, NFI_ERR_UNSUPPORTED, "nfi_composite_bwd: S = %d > 128 samples per ray", S);
  composite_bwd_kernel<<<warp_grid(R), kThreads, 0, nfi_cu(stream)>>>(
      sigma, rgb, dirs, z, semantics, n_sem, R, S, white_background, g_rgb_map, g_mask, g_semantic_map,
      g_sigma, g_rgb, g_dirs, g_semantics);
  NFI_CHECK_LAUNCH("composite_bwd_kernel");
  return NFI_OK;
}

extern "C" int nfi_merge_composite_fwd(const float* coarse, const float* fine, const float* z_coarse,
                                       const float* z_fine, const float* dirs, int64_t R, int S,
                                       int white_background, const float* cam2world,
                                       int64_t rays_per_image, float* rgb_map, float* depth_map,
                                       float* mask, uint8_t* perm, float* z_sorted,
                                       nfi_stream_t stream) {
  NFI_REQUIRE(coarse && fine && z_coarse && z_fine && dirs && rgb_map && depth_map && mask, NFI_ERR_BAD_ARG,
              "nfi_merge_composite_fwd: NULL pointer");
  NFI_REQUIRE(R > 0 && S >= 1 && (!cam2world || rays_per_image > 0), NFI_ERR_BAD_ARG, "nfi_merge_composite_fwd: bad sizes");
  NFI_REQUIRE(S <= NFI_MAX_S, NFI_ERR_UNSUPPORTED, "nfi_merge_composite_fwd: S = %d > %d", S, NFI_MAX_S);
  NFI_REQUIRE(nfi_aligned16(coarse) && nfi_aligned16(fine), NFI_ERR_ALIGNMENT, "nfi_merge_composite_fwd: float4 buffers must be 16-byte aligned");
  NFI_REQUIRE((!z_sorted || nfi_aligned16(z_sorted)) && (((uintptr_t)perm) & 3u) == 0, NFI_ERR_ALIGNMENT,
              "nfi_merge_composite_fwd: z_sorted must be 16-byte and perm 4-byte aligned");
  merge_composite_fwd_kernel<<<warp_grid(R), kThreads, 0, nfi_cu(stream)>>>(
      (const float4*)coarse, (const float4*)fine, z_coarse, z_fine, dirs, R, S, white_background, cam2world,
      rays_per_image, rgb_map, depth_map, mask, perm, z_sorted);
  NFI_CHECK_LAUNCH("merge_composite_fwd_kernel");
  return NFI_OK;
}

extern "C" int nfi_merge_composite_bwd(const float* coarse, const float* fine, const float* z_sorted,
                                       const uint8_t* perm, const float* dirs, int64_t R, int S,
                                       int white_background, const float* g_rgb_map,
                                       const float* g_mask, float* g_coarse, float* g_fine,
                                       nfi_stream_t stream) {
  NFI_REQUIRE(coarse && fine && z_sorted && perm && dirs && g_rgb_map && g_coarse && g_fine, NFI_ERR_BAD_ARG,
              "nfi_merge_composite_bwd: NULL pointer");
  NFI_REQUIRE(R > 0 && S >= 1, NFI_ERR_BAD_ARG, "nfi_merge_composite_bwd: bad sizes");
  NFI_REQUIRE(S <= NFI_MAX_S, NFI_ERR_UNSUPPORTED, "nfi_merge_composite_bwd: S = %d > %d", S, NFI_MAX_S);
  NFI_REQUIRE(nfi_aligned16(coarse) && nfi_aligned16(fine) && nfi_aligned16(g_coarse) && nfi_aligned16(g_fine),
              NFI_ERR_ALIGNMENT, "nfi_merge_composite_bwd: float4 buffers must be 16-byte aligned");
  NFI_REQUIRE(nfi_aligned16(z_sorted) && (((uintptr_t)perm) & 3u) == 0, NFI_ERR_ALIGNMENT,
              "nfi_merge_composite_bwd: z_sorted must be 16-byte and perm 4-byte aligned");
  merge_composite_bwd_kernel<<<warp_grid(R), kThreads, 0, nfi_cu(stream)>>>(
      (const float4*)coarse, (const float4*)fine, z_sorted, perm, dirs, R, S, white_background, g_rgb_map,
      g_mask, (float4*)g_coarse, (float4*)g_fine);
  NFI_CHECK_LAUNCH("merge_composite_bwd_kernel");
  return NFI_OK;
}
