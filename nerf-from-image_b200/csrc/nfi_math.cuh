// Per-ray / per-sample arithmetic of the render hot path, shared by every kernel.
//
// Everything here is __host__ __device__ so that tests/hostcheck can run the *same* functions
// on the CPU against the oracle before any GPU time is spent (test infrastructure; the product
// only ever calls them from kernels).  Where the reference's results feed an integer decision
// (out-of-box mask, texel index, searchsorted index, sort order) the fp32 operation order of
// the reference/ATen is reproduced exactly -- round-to-nearest intrinsics keep nvcc from
// contracting a*b+c into an FMA where ATen does not, and vice versa.
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define NFI_HD __host__ __device__ __forceinline__
#else
#define NFI_HD inline
#endif

#if defined(__CUDA_ARCH__)
#define NFI_MUL(a, b) __fmul_rn((a), (b))
#define NFI_ADD(a, b) __fadd_rn((a), (b))
#define NFI_SUB(a, b) __fsub_rn((a), (b))
#define NFI_DIV(a, b) __fdiv_rn((a), (b))
#define NFI_FMA(a, b, c) __fmaf_rn((a), (b), (c))
#define NFI_SQRT(a) __fsqrt_rn((a))
#define NFI_DMUL(a, b) __dmul_rn((a), (b))
#define NFI_DADD(a, b) __dadd_rn((a), (b))
#else
// host build: compiled with -ffp-contract=off, so plain operators are single IEEE operations
#define NFI_MUL(a, b) ((a) * (b))
#define NFI_ADD(a, b) ((a) + (b))
#define NFI_SUB(a, b) ((a) - (b))
#define NFI_DIV(a, b) ((a) / (b))
#define NFI_FMA(a, b, c) fmaf((a), (b), (c))
#define NFI_SQRT(a) sqrtf((a))
#define NFI_DMUL(a, b) ((a) * (b))
#define NFI_DADD(a, b) ((a) + (b))
#endif

#define NFI_C 32     // feature channels per plane
#define NFI_HID 64   // decoder hidden width
#define NFI_YPAD 16  // padded decoder output width

// ---------------------------------------------------------------------------------------------
// NaN-propagating min / max / clamp (torch.max/min/clamp semantics, nerf_utils.py:253-265)
// ---------------------------------------------------------------------------------------------
NFI_HD float nfi_max_nan(float a, float b) {
  if (a != a) return a;
  if (b != b) return b;
  return a > b ? a : b;
}
NFI_HD float nfi_min_nan(float a, float b) {
  if (a != a) return a;
  if (b != b) return b;
  return a < b ? a : b;
}

// ---------------------------------------------------------------------------------------------
// A1  pixel -> ray   (lib/nerf_utils.py:28-93)
// ---------------------------------------------------------------------------------------------
struct NfiCamera {
  float R[9];   // cam2world[:3,:3] row-major
  float t[3];   // cam2world[:3,3]
  float t33;    // cam2world[3,3]
  float focal;  // perspective only
  float cx, cy, bx, by, bw, bh;
  int persp, has_center, has_bbox;
};

NFI_HD void nfi_load_camera(NfiCamera& c, const float* cam2world, const float* focal,
                            int focal_stride, const float* center, const float* bbox, int b) {
  const float* m = cam2world + (int64_t)b * 16;
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j) c.R[i * 3 + j] = m[i * 4 + j];
    c.t[i] = m[i * 4 + 3];
  }
  c.t33 = m[15];
  c.persp = focal != nullptr;
  c.focal = c.persp ? focal[(int64_t)b * focal_stride] : 1.f;
  c.has_center = center != nullptr;
  c.has_bbox = bbox != nullptr;
  c.cx = c.has_center ? center[b * 2 + 0] : 0.f;
  c.cy = c.has_center ? center[b * 2 + 1] : 0.f;
  c.bx = c.has_bbox ? bbox[b * 4 + 0] : 0.f;
  c.by = c.has_bbox ? bbox[b * 4 + 1] : 0.f;
  c.bw = c.has_bbox ? bbox[b * 4 + 2] : 0.f;
  c.bh = c.has_bbox ? bbox[b * 4 + 3] : 0.f;
}

// Camera-plane coordinates (u, v) of pixel (h, w): everything before the rotation.
// For the perspective model these are already divided by the focal length.
NFI_HD void nfi_pixel_uv(const NfiCamera& c, int h, int w, int H, int W, float& u, float& v) {
  u = NFI_DIV((float)w, (float)W);
  v = NFI_DIV((float)h, (float)H);
  if (c.persp) {
    if (c.has_center) {  // nerf_utils.py:44-47
      u = NFI_SUB(NFI_SUB(u, NFI_MUL(0.5f, NFI_SUB(NFI_MUL(2.f, c.cx), 1.f))), 0.5f);
      v = NFI_SUB(NFI_SUB(v, NFI_MUL(0.5f, NFI_SUB(NFI_MUL(2.f, c.cy), 1.f))), 0.5f);
    } else {
      u = NFI_SUB(u, 0.5f);
      v = NFI_SUB(v, 0.5f);
    }
    if (c.has_bbox) {    // nerf_utils.py:52-56
      u = NFI_MUL(NFI_ADD(NFI_MUL(c.bw, NFI_ADD(u, 0.5f)), c.bx), 0.5f);
      v = NFI_MUL(-NFI_ADD(NFI_MUL(c.bh, NFI_ADD(-v, 0.5f)), c.by), 0.5f);
    }
    u = NFI_DIV(u, c.focal);
    v = NFI_DIV(v, c.focal);
  } else {
    u = NFI_MUL(NFI_SUB(u, 0.5f), 2.f);
    v = NFI_MUL(NFI_SUB(v, 0.5f), 2.f);
    if (c.has_bbox) {    // nerf_utils.py:73-77
      u = NFI_ADD(NFI_MUL(c.bw, NFI_ADD(NFI_DIV(u, 2.f), 0.5f)), c.bx);
      v = -NFI_ADD(NFI_MUL(c.bh, NFI_ADD(NFI_DIV(-v, 2.f), 0.5f)), c.by);
    }
  }
}

NFI_HD float nfi_dot3_seq(float a0, float a1, float a2, const float* r) {
  return NFI_ADD(NFI_ADD(NFI_MUL(a0, r[0]), NFI_MUL(a1, r[1])), NFI_MUL(a2, r[2]));
}

NFI_HD float nfi_norm3(const float* d) {  // ATen's order: fma chain, then sqrt
  return NFI_SQRT(NFI_FMA(d[2], d[2], NFI_FMA(d[1], d[1], NFI_MUL(d[0], d[0]))));
}

// o, d_raw (un-normalised) for one pixel.
NFI_HD void nfi_pixel_ray(const NfiCamera& c, int h, int w, int H, int W, float* o, float* d) {
  float u, v;
  nfi_pixel_uv(c, h, w, H, W, u, v);
  if (c.persp) {
    for (int k = 0; k < 3; ++k) {
      d[k] = nfi_dot3_seq(u, -v, -1.f, c.R + 3 * k);
      o[k] = c.t[k];
    }
  } else {
    for (int k = 0; k < 3; ++k) {
      o[k] = NFI_ADD(nfi_dot3_seq(u, -v, 0.f, c.R + 3 * k), c.t[k]);
      d[k] = NFI_DIV(nfi_dot3_seq(0.f, 0.f, -1.f, c.R + 3 * k), c.t33);
    }
  }
}

NFI_HD void nfi_normalize3(float* d) {  // F.normalize, eps = 1e-12
  float n = nfi_norm3(d);
  n = n > 1e-12f ? n : 1e-12f;
  d[0] = NFI_DIV(d[0], n);
  d[1] = NFI_DIV(d[1], n);
  d[2] = NFI_DIV(d[2], n);
}

// Backward of nfi_pixel_ray (+ optional normalisation) for one pixel.  Adds into gR[9], gt[3],
// *gt33, *gfocal.  go / gd may be NULL.  center / bbox are data (no gradient).
NFI_HD void nfi_pixel_ray_bwd(const NfiCamera& c, int h, int w, int H, int W, int normalize,
                              const float* go, const float* gd, float* gR, float* gt, float* gt33,
                              float* gfocal) {
  float u, v;
  nfi_pixel_uv(c, h, w, H, W, u, v);
  float graw[3] = {0.f, 0.f, 0.f};
  if (gd) {
    graw[0] = gd[0]; graw[1] = gd[1]; graw[2] = gd[2];
    if (normalize) {
      float o_[3], d[3];
      nfi_pixel_ray(c, h, w, H, W, o_, d);
      float n = nfi_norm3(d);
      if (n > 1e-12f) {
        float dn[3] = {d[0] / n, d[1] / n, d[2] / n};
        float dot = gd[0] * dn[0] + gd[1] * dn[1] + gd[2] * dn[2];
        for (int k = 0; k < 3; ++k) graw[k] = (gd[k] - dn[k] * dot) / n;
      } else {
        for (int k = 0; k < 3; ++k) graw[k] = gd[k] / 1e-12f;
      }
    }
  }
  if (c.persp) {
    const float cam[3] = {u, -v, -1.f};
    float gcam0 = 0.f, gcam1 = 0.f;
    for (int k = 0; k < 3; ++k) {
      for (int j = 0; j < 3; ++j) gR[k * 3 + j] += graw[k] * cam[j];
      gcam0 += graw[k] * c.R[k * 3 + 0];
      gcam1 += graw[k] * c.R[k * 3 + 1];
      if (go) gt[k] += go[k];
    }
    // u = u_n / f, cam1 = -v_n / f
    *gfocal += -(gcam0 * u - gcam1 * v) / c.focal;
  } else {
    const float cam[3] = {u, -v, 0.f};
    for (int k = 0; k < 3; ++k) {
      if (go) {
        for (int j = 0; j < 3; ++j) gR[k * 3 + j] += go[k] * cam[j];
        gt[k] += go[k];
      }
      gR[k * 3 + 2] += -graw[k] / c.t33;
      *gt33 += graw[k] * c.R[k * 3 + 2] / (c.t33 * c.t33);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// A2  slab test against [-r, r]^3   (lib/nerf_utils.py:227-258)
// ---------------------------------------------------------------------------------------------
NFI_HD bool nfi_slab(const float* o, const float* d, float r, float& near, float& far) {
  float tmin[3], tmax[3];
  for (int a = 0; a < 3; ++a) {
    const float inv = NFI_DIV(1.f, d[a]);
    const bool neg = inv < 0.f;
    tmin[a] = NFI_MUL(NFI_SUB(neg ? r : -r, o[a]), inv);
    tmax[a] = NFI_MUL(NFI_SUB(neg ? -r : r, o[a]), inv);
  }
  bool hit = !((tmin[0] > tmax[1]) || (tmin[1] > tmax[0]));
  near = nfi_max_nan(tmin[0], tmin[1]);
  far = nfi_min_nan(tmax[0], tmax[1]);
  hit = hit && !((near > tmax[2]) || (tmin[2] > far));
  near = nfi_max_nan(near, tmin[2]);
  far = nfi_min_nan(far, tmax[2]);
  return hit;
}

// miss-ray substitution, clamps and minimum extent (lib/nerf_utils.py:260-270)
NFI_HD void nfi_near_far_finish(bool hit, float gmin, float gmax, float& near, float& far) {
  if (!hit) { near = gmin; far = gmax; }
  near = (near != near) ? near : (near > 0.1f ? near : 0.1f);
  far = (far != far) ? far : (far > 0.1f ? far : 0.1f);
  if (NFI_SUB(far, near) < 1e-3f) far = NFI_ADD(near, 1e-3f);
}

// ---------------------------------------------------------------------------------------------
// A3  stratified depth k of S   (lib/nerf_utils.py:104-117; torch.lerp's two-branch FMA form)
// ---------------------------------------------------------------------------------------------
NFI_HD float nfi_stratified_z(float near, float far, int k, int S, bool jitter, float u) {
  const float w = NFI_DIV((float)k, (float)S);
  const float diff = NFI_SUB(far, near);
  float z = (w < 0.5f) ? NFI_FMA(w, diff, near) : NFI_FMA(NFI_SUB(w, 1.f), diff, far);
  if (jitter) z = NFI_ADD(z, NFI_MUL(u, NFI_DIV(diff, (float)S)));
  return z;
}

NFI_HD void nfi_point(const float* o, const float* d, float z, float* p) {
  p[0] = NFI_ADD(o[0], NFI_MUL(d[0], z));
  p[1] = NFI_ADD(o[1], NFI_MUL(d[1], z));
  p[2] = NFI_ADD(o[2], NFI_MUL(d[2], z));
}

// Philox4x32-10 (counter-based; used when the caller supplies no uniforms)
NFI_HD void nfi_philox(uint64_t seed, uint64_t ctr_lo, uint64_t ctr_hi, uint32_t out[4]) {
  uint32_t c0 = (uint32_t)ctr_lo, c1 = (uint32_t)(ctr_lo >> 32);
  uint32_t c2 = (uint32_t)ctr_hi, c3 = (uint32_t)(ctr_hi >> 32);
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
  for (int i = 0; i < 10; ++i) {
    const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    c1 = (uint32_t)p1; c3 = (uint32_t)p0; c0 = n0; c2 = n2;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
NFI_HD float nfi_u01(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }  // [0, 1)
