// Triplane field arithmetic: bilinear footprint, decoder MLP, density / colour heads
// (models/generator.py:288-331, 587-681; models/stylegan.py:148-180), forward and backward.
// __host__ __device__, shared with tests/hostcheck (see nfi_math.cuh).
#pragma once
#include "nfi_math.cuh"

// exp on the MUFU pipe in kernels (ex2.approx: relative error ~2^-21, far inside the 1e-4 contract);
// the host emulation build (tests/hostcheck) uses libm.
#if defined(__CUDA_ARCH__)
// flush-to-zero form: no denormal fix-up instructions around the MUFU op
__device__ __forceinline__ float nfi_exp_fast(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x * 1.4426950408889634f));
  return y;
}
#define NFI_EXPF(x) nfi_exp_fast(x)
#else
#define NFI_EXPF(x) expf(x)
#endif

// ---------------------------------------------------------------------------------------------
// Texel order of a channels-last plane in HBM.  When both plane sides are multiples of 8 the plane is
// stored as contiguous 8 x 8 texel blocks (8 KB each, bshift = 3): the 2 x 2 bilinear footprint and the
// ~9 x 17 texel neighbourhood of a pixel-block tile then cover a few contiguous 8 KB runs instead of
// 9 rows one power-of-two stride apart (which alias in L1).  Otherwise row-major (bshift = 0).
//   index(y, x) = nfi_texel_row(y) + nfi_texel_col(x)        (texel units; * 32 floats)
// ---------------------------------------------------------------------------------------------
NFI_HD int nfi_plane_bshift(int Hp, int Wp) { return (Hp % 8 == 0 && Wp % 8 == 0) ? 3 : 0; }
NFI_HD int nfi_texel_row(int y, int bshift, int nbx) {
  return (((y >> bshift) * nbx) << (2 * bshift)) + ((y & ((1 << bshift) - 1)) << bshift);
}
NFI_HD int nfi_texel_col(int x, int bshift) {
  return ((x >> bshift) << (2 * bshift)) + (x & ((1 << bshift) - 1));
}

// ---------------------------------------------------------------------------------------------
// F.grid_sample(bilinear, border, align_corners=True), one axis   (ATen GridSampler.h)
//   i = ((g + 1) / 2) * (size - 1); clamp to [0, size-1]; d i / d g = (size-1)/2 strictly inside
// ---------------------------------------------------------------------------------------------
NFI_HD float nfi_unnormalize(float g, int size, float* dmult) {
  float i = NFI_MUL(NFI_DIV(NFI_ADD(g, 1.f), 2.f), (float)(size - 1));
  const float hi = (float)(size - 1);
  float m = hi * 0.5f;
  if (!(i > 0.f)) { i = 0.f; m = 0.f; }         // also catches NaN like std::max(in, 0)
  else if (i >= hi) { i = hi; m = 0.f; }
  if (dmult) *dmult = m;
  return i;
}

struct NfiFoot {   // bilinear footprint of one sample on one plane
  int x0, y0, x1, y1;        // texel indices (x1/y1 clamped in-bounds; their weight is 0 there)
  float wnw, wne, wsw, wse;  // ATen's weights
  float tx, ty;
};

NFI_HD void nfi_footprint(float ix, float iy, int Wp, int Hp, NfiFoot& f) {
  const float fx = floorf(ix), fy = floorf(iy);
  f.x0 = (int)fx; f.y0 = (int)fy;
  f.x1 = f.x0 + 1 < Wp ? f.x0 + 1 : Wp - 1;
  f.y1 = f.y0 + 1 < Hp ? f.y0 + 1 : Hp - 1;
  const float ex = NFI_SUB(NFI_ADD(fx, 1.f), ix), ey = NFI_SUB(NFI_ADD(fy, 1.f), iy);
  f.tx = NFI_SUB(ix, fx); f.ty = NFI_SUB(iy, fy);
  f.wnw = NFI_MUL(ex, ey); f.wne = NFI_MUL(f.tx, ey);
  f.wsw = NFI_MUL(ex, f.ty); f.wse = NFI_MUL(f.tx, f.ty);
}

// weighted sum of the four texels (one channel) and the mean over the three planes
NFI_HD float nfi_bilerp(float v00, float v01, float v10, float v11, const NfiFoot& f) {
  return fmaf(v11, f.wse, fmaf(v10, f.wsw, fmaf(v01, f.wne, v00 * f.wnw)));
}
NFI_HD float nfi_mean3(float e0, float e1, float e2) {   // (e1 + e2 + e3) / 3, generator.py:328
  return NFI_DIV(NFI_ADD(NFI_ADD(e0, e1), e2), 3.f);
}
// d(bilerp)/d(ix, iy) for one channel, times upstream g; neighbours outside the plane count as 0
// (ATen within_bounds_2d).  vx / vy = 1 if x0+1 / y0+1 is inside.
NFI_HD void nfi_bilerp_grad(float g, float v00, float v01, float v10, float v11, float vx, float vy,
                            const NfiFoot& f, float& gx, float& gy) {
  const float nw = v00, ne = v01 * vx, sw = v10 * vy, se = v11 * vx * vy;
  gx = fmaf(g, (ne - nw) * (1.f - f.ty) + (se - sw) * f.ty, gx);
  gy = fmaf(g, (sw - nw) * (1.f - f.tx) + (se - ne) * f.tx, gy);
}
// d i / d g of nfi_unnormalize, from the clamped coordinate
NFI_HD float nfi_coord_mult(float i, int size) {
  return (i > 0.f && i < (float)(size - 1)) ? 0.5f * (float)(size - 1) : 0.f;
}

// plane p in {xy, xz, yz} reads point coordinates (a -> width index, b -> height index)
NFI_HD int nfi_plane_axis_a(int p) { return p == 2 ? 1 : 0; }
NFI_HD int nfi_plane_axis_b(int p) { return p == 0 ? 1 : 2; }

// ---------------------------------------------------------------------------------------------
// Decoder MLP.  Weight layouts (gain already applied):
//   W0kj [32][64]  (k-major: the 64 outputs of input channel k are contiguous)
//   W0jk [64][32]  (for the backward)
//   W1jo [64][16]  (outputs padded to 16)
// ---------------------------------------------------------------------------------------------
NFI_HD float nfi_softplus(float x) { return x > 20.f ? x : log1pf(NFI_EXPF(x)); }
NFI_HD float nfi_softplus_grad(float x) {
  if (x > 20.f) return 1.f;
  const float e = NFI_EXPF(x);
  return e / (e + 1.f);
}

template <typename F4>
NFI_HD void nfi_mlp_layer0(const float* feat, const F4* W0kj4, const F4* b04, float* hpre) {
#pragma unroll
  for (int j4 = 0; j4 < NFI_HID / 4; ++j4) {
    const F4 b = b04[j4];
    hpre[4 * j4 + 0] = b.x; hpre[4 * j4 + 1] = b.y; hpre[4 * j4 + 2] = b.z; hpre[4 * j4 + 3] = b.w;
  }
#pragma unroll
  for (int k = 0; k < NFI_C; ++k) {
    const float f = feat[k];
#pragma unroll
    for (int j4 = 0; j4 < NFI_HID / 4; ++j4) {
      const F4 w = W0kj4[k * (NFI_HID / 4) + j4];
      hpre[4 * j4 + 0] = fmaf(f, w.x, hpre[4 * j4 + 0]);
      hpre[4 * j4 + 1] = fmaf(f, w.y, hpre[4 * j4 + 1]);
      hpre[4 * j4 + 2] = fmaf(f, w.z, hpre[4 * j4 + 2]);
      hpre[4 * j4 + 3] = fmaf(f, w.w, hpre[4 * j4 + 3]);
    }
  }
}

// y[16] = b1 + softplus(hpre) . W1 ; NO4 = ceil(n_out / 4)
template <typename F4>
NFI_HD void nfi_mlp_layer1(const float* hpre, const F4* W1jo4, const F4* b14, int no4, float* y) {
#pragma unroll
  for (int o4 = 0; o4 < NFI_YPAD / 4; ++o4) {
    const F4 b = b14[o4];
    y[4 * o4 + 0] = b.x; y[4 * o4 + 1] = b.y; y[4 * o4 + 2] = b.z; y[4 * o4 + 3] = b.w;
  }
#pragma unroll
  for (int j = 0; j < NFI_HID; ++j) {
    const float h = nfi_softplus(hpre[j]);
#pragma unroll
    for (int o4 = 0; o4 < NFI_YPAD / 4; ++o4) {
      if (o4 < no4) {
        const F4 w = W1jo4[j * (NFI_YPAD / 4) + o4];
        y[4 * o4 + 0] = fmaf(h, w.x, y[4 * o4 + 0]);
        y[4 * o4 + 1] = fmaf(h, w.y, y[4 * o4 + 1]);
        y[4 * o4 + 2] = fmaf(h, w.z, y[4 * o4 + 2]);
        y[4 * o4 + 3] = fmaf(h, w.w, y[4 * o4 + 3]);
      }
    }
  }
}

// hpre[j] <- d L / d hpre[j] given dy[16] (entries >= n_out must be 0)
template <typename F4>
NFI_HD void nfi_mlp_layer1_bwd(float* hpre, const F4* W1jo4, int no4, const float* dy) {
#pragma unroll
  for (int j = 0; j < NFI_HID; ++j) {
    float dh = 0.f;
#pragma unroll
    for (int o4 = 0; o4 < NFI_YPAD / 4; ++o4) {
      if (o4 < no4) {
        const F4 w = W1jo4[j * (NFI_YPAD / 4) + o4];
        dh = fmaf(dy[4 * o4 + 0], w.x, dh);
        dh = fmaf(dy[4 * o4 + 1], w.y, dh);
        dh = fmaf(dy[4 * o4 + 2], w.z, dh);
        dh = fmaf(dy[4 * o4 + 3], w.w, dh);
      }
    }
    hpre[j] = dh * nfi_softplus_grad(hpre[j]);
  }
}

// dfeat[32] = W0^T . dpre
template <typename F4>
NFI_HD void nfi_mlp_layer0_bwd(const float* dpre, const F4* W0jk4, float* dfeat) {
#pragma unroll
  for (int k = 0; k < NFI_C; ++k) dfeat[k] = 0.f;
#pragma unroll
  for (int j = 0; j < NFI_HID; ++j) {
    const float g = dpre[j];
#pragma unroll
    for (int k4 = 0; k4 < NFI_C / 4; ++k4) {
      const F4 w = W0jk4[j * (NFI_C / 4) + k4];
      dfeat[4 * k4 + 0] = fmaf(g, w.x, dfeat[4 * k4 + 0]);
      dfeat[4 * k4 + 1] = fmaf(g, w.y, dfeat[4 * k4 + 1]);
      dfeat[4 * k4 + 2] = fmaf(g, w.z, dfeat[4 * k4 + 2]);
      dfeat[4 * k4 + 3] = fmaf(g, w.w, dfeat[4 * k4 + 3]);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Heads   (models/generator.py:629-641 density, :661-679 colour)
// ---------------------------------------------------------------------------------------------
struct NfiHeads {
  int A;           // attention values (0 -> plain RGB head)
  int use_sdf;
  float beta, inv_alpha, alpha;
};

// y -> sigma, rgb[3]; probs[15] (softmax, zero-padded) is an output for A > 0.  keep = 1 - out_of_box.
NFI_HD void nfi_heads_fwd(const NfiHeads& hd, const float* y, float keep, const float* palette,
                          float& sigma, float* rgb, float* probs) {
  const float dist = y[0];
  if (hd.use_sdf) {
    const float t = -dist;
    const float e = expf(NFI_DIV(-fabsf(t), hd.beta));   // full precision: sigma feeds exp(-sigma * dist)
    const float sgn = (t > 0.f) ? 1.f : ((t < 0.f) ? -1.f : 0.f);
    const float cdf = 0.5f + 0.5f * sgn * (1.f - e);
    sigma = hd.inv_alpha * (cdf * keep);
  } else {
    sigma = nfi_softplus(dist - 1.f) * keep;
  }
  if (hd.A > 0) {
    // fixed trip counts + predicates keep y/probs in registers (no dynamic indexing)
    float m = y[1];
#pragma unroll
    for (int a = 1; a < NFI_YPAD - 1; ++a) if (a < hd.A) m = fmaxf(m, y[1 + a]);
    float s = 0.f;
#pragma unroll
    for (int a = 0; a < NFI_YPAD - 1; ++a) {
      probs[a] = (a < hd.A) ? NFI_EXPF(y[1 + a] - m) : 0.f;
      s += probs[a];
    }
    const float inv = 1.f / s;
    rgb[0] = rgb[1] = rgb[2] = 0.f;
#pragma unroll
    for (int a = 0; a < NFI_YPAD - 1; ++a) {
      if (a < hd.A) {
        probs[a] *= inv;
        rgb[0] = fmaf(probs[a], palette[a * 3 + 0], rgb[0]);
        rgb[1] = fmaf(probs[a], palette[a * 3 + 1], rgb[1]);
        rgb[2] = fmaf(probs[a], palette[a * 3 + 2], rgb[2]);
      }
    }
  } else {
    for (int c = 0; c < 3; ++c) rgb[c] = (1.f / (1.f + NFI_EXPF(-y[1 + c]))) * 2.004f - 1.002f;
  }
}

// Backward of the heads: (dsigma, drgb[3]) -> dy[16] (zero-padded).  probs from the forward.
// dbeta / dalpha accumulate when non-NULL.  rgb (forward value) is needed for the plain head.
NFI_HD void nfi_heads_bwd(const NfiHeads& hd, const float* y, float keep, const float* palette,
                          const float* probs, const float* rgb, float dsigma, const float* drgb,
                          float* dy, float* dbeta, float* dalpha) {
#pragma unroll
  for (int o = 0; o < NFI_YPAD; ++o) dy[o] = 0.f;
  const float dist = y[0];
  if (hd.use_sdf) {
    const float t = -dist;
    const float at = fabsf(t);
    const float e = expf(NFI_DIV(-at, hd.beta));
    const float sgn = (t > 0.f) ? 1.f : ((t < 0.f) ? -1.f : 0.f);
    // d cdf / d t = 0.5 * sgn^2 * e / beta   (sign() has zero gradient, |t|' = sgn)
    const float dcdf_dt = 0.5f * sgn * sgn * e / hd.beta;
    dy[0] = -dsigma * hd.inv_alpha * keep * dcdf_dt;
    if (dbeta) *dbeta += dsigma * hd.inv_alpha * keep * (0.5f * sgn * (-e) * (at / (hd.beta * hd.beta)));
    if (dalpha) {
      const float cdf = 0.5f + 0.5f * sgn * (1.f - e);
      *dalpha += -dsigma * (cdf * keep) * hd.inv_alpha * hd.inv_alpha;
    }
  } else {
    dy[0] = dsigma * keep * nfi_softplus_grad(dist - 1.f);
  }
  if (hd.A > 0) {
    float dp[NFI_YPAD - 1];
    float dot = 0.f;
#pragma unroll
    for (int a = 0; a < NFI_YPAD - 1; ++a) {
      dp[a] = 0.f;
      if (a < hd.A) {
        dp[a] = drgb[0] * palette[a * 3 + 0] + drgb[1] * palette[a * 3 + 1] + drgb[2] * palette[a * 3 + 2];
        dot = fmaf(probs[a], dp[a], dot);
      }
    }
#pragma unroll
    for (int a = 0; a < NFI_YPAD - 1; ++a) if (a < hd.A) dy[1 + a] = probs[a] * (dp[a] - dot);
  } else {
    for (int c = 0; c < 3; ++c) {
      const float s = (rgb[c] + 1.002f) / 2.004f;
      dy[1 + c] = drgb[c] * 2.004f * s * (1.f - s);
    }
  }
}
