// TEST INFRASTRUCTURE ONLY -- never linked into the product.
// Runs the __host__ __device__ arithmetic of nerf-from-image_b200/csrc/*.cuh in plain CPU loops so
// the per-ray / per-sample math (forward and adjoint) can be compared against the oracle in the
// build container, where no GPU exists.  The kernels' parallel decomposition (shared-memory
// staging, shuffles, atomics) is NOT emulated here; that is what the -m gpu tests verify.
#include <stdint.h>
#include <string.h>

#include <vector>

#include "../../nerf-from-image_b200/csrc/nfi_field_math.cuh"

struct F4 { float x, y, z, w; };

extern "C" {

void hc_rays(const float* cam, const float* focal, int fstride, const float* center, const float* bbox,
             int B, int H, int W, int normalize, float* o, float* d) {
  for (int b = 0; b < B; ++b) {
    NfiCamera c;
    nfi_load_camera(c, cam, focal, fstride, center, bbox, b);
    for (int h = 0; h < H; ++h)
      for (int w = 0; w < W; ++w) {
        const int64_t r = ((int64_t)b * H + h) * W + w;
        nfi_pixel_ray(c, h, w, H, W, o + r * 3, d + r * 3);
        if (normalize) nfi_normalize3(d + r * 3);
      }
  }
}

void hc_rays_bwd(const float* cam, const float* focal, int fstride, const float* center, const float* bbox,
                 int B, int H, int W, int normalize, const float* go, const float* gd, float* gcam,
                 float* gfocal) {
  for (int b = 0; b < B; ++b) {
    NfiCamera c;
    nfi_load_camera(c, cam, focal, fstride, center, bbox, b);
    float acc[14];
    for (int i = 0; i < 14; ++i) acc[i] = 0.f;
    for (int h = 0; h < H; ++h)
      for (int w = 0; w < W; ++w) {
        const int64_t r = ((int64_t)b * H + h) * W + w;
        nfi_pixel_ray_bwd(c, h, w, H, W, normalize, go ? go + r * 3 : nullptr, gd ? gd + r * 3 : nullptr,
                          acc, acc + 9, acc + 12, acc + 13);
      }
    for (int i = 0; i < 16; ++i) gcam[b * 16 + i] = 0.f;
    for (int i = 0; i < 9; ++i) gcam[b * 16 + (i / 3) * 4 + i % 3] = acc[i];
    for (int i = 0; i < 3; ++i) gcam[b * 16 + i * 4 + 3] = acc[9 + i];
    gcam[b * 16 + 15] = acc[12];
    gfocal[b] = acc[13];
  }
}

int hc_near_far(const float* o, const float* d, int64_t R, float r_box, float* near, float* far) {
  float lo = INFINITY, hi = -INFINITY;
  int cnt = 0;
  std::vector<char> hit(R);
  for (int64_t r = 0; r < R; ++r) {
    hit[r] = nfi_slab(o + r * 3, d + r * 3, r_box, near[r], far[r]);
    if (hit[r]) { lo = nfi_min_nan(lo, near[r]); hi = nfi_max_nan(hi, far[r]); ++cnt; }
  }
  for (int64_t r = 0; r < R; ++r) nfi_near_far_finish(hit[r], lo, hi, near[r], far[r]);
  return cnt;
}

void hc_stratified(const float* near, const float* far, const float* u, int64_t R, int S, float* z) {
  for (int64_t r = 0; r < R; ++r)
    for (int k = 0; k < S; ++k)
      z[r * S + k] = nfi_stratified_z(near[r], far[r], k, S, u != nullptr, u ? u[r * S + k] : 0.f);
}

void hc_points(const float* o, const float* d, const float* z, int64_t R, int S, float* pts) {
  for (int64_t r = 0; r < R; ++r)
    for (int k = 0; k < S; ++k) nfi_point(o + r * 3, d + r * 3, z[r * S + k], pts + (r * S + k) * 3);
}

// Field evaluation at explicit points [B, N, 3]; planes channels-last [B,3,Hp,Wp,32].
// mode 0: out[.,4] = (sigma, rgb); mode 1: out[., n_out] raw.  g_out == NULL -> forward only.
void hc_field(const float* W0, const float* b0, const float* W1, const float* b1, float beta, float alpha,
              int n_out, int A, int use_sdf, float scene_range, const float* planes, const float* palette,
              int B, int Hp, int Wp, const float* pts, int64_t N, int mode, float* out, float* probs_out,
              const float* g_out, float* g_planes, float* g_palette, float* g_pts, float* g_beta_alpha) {
  const float g0 = (float)(1.0 / sqrt((double)NFI_C)), g1 = (float)(1.0 / sqrt((double)NFI_HID));
  std::vector<F4> W0kj(NFI_C * NFI_HID / 4), W0jk(NFI_HID * NFI_C / 4), W1jo(NFI_HID * NFI_YPAD / 4), b04(NFI_HID / 4), b14(NFI_YPAD / 4);
  float* w0kj = (float*)W0kj.data(); float* w0jk = (float*)W0jk.data(); float* w1 = (float*)W1jo.data();
  for (int j = 0; j < NFI_HID; ++j)
    for (int k = 0; k < NFI_C; ++k) { w0kj[k * NFI_HID + j] = W0[j * NFI_C + k] * g0; w0jk[j * NFI_C + k] = W0[j * NFI_C + k] * g0; }
  for (int j = 0; j < NFI_HID; ++j)
    for (int o = 0; o < NFI_YPAD; ++o) w1[j * NFI_YPAD + o] = o < n_out ? W1[o * NFI_HID + j] * g1 : 0.f;
  for (int j = 0; j < NFI_HID; ++j) ((float*)b04.data())[j] = b0[j];
  for (int o = 0; o < NFI_YPAD; ++o) ((float*)b14.data())[o] = o < n_out ? b1[o] : 0.f;
  NfiHeads hd; hd.A = A; hd.use_sdf = use_sdf; hd.beta = beta; hd.alpha = alpha; hd.inv_alpha = 1.f / alpha;
  const int no4 = (n_out + 3) >> 2;
  const int64_t plane_sz = (int64_t)Hp * Wp * NFI_C;
  for (int b = 0; b < B; ++b)
    for (int64_t n = 0; n < N; ++n) {
      const int64_t id = (int64_t)b * N + n;
      float x[3];
      for (int c = 0; c < 3; ++c) x[c] = pts[id * 3 + c] / scene_range;
      float keep = 1.f;
      if (mode == 0 && (fabsf(x[0]) > 1.f || fabsf(x[1]) > 1.f || fabsf(x[2]) > 1.f)) keep = 0.f;
      NfiFoot f[3]; float ci[3][2];
      float feat[NFI_C], e[3][NFI_C];
      for (int p = 0; p < 3; ++p) {
        ci[p][0] = nfi_unnormalize(x[nfi_plane_axis_a(p)], Wp, nullptr);
        ci[p][1] = nfi_unnormalize(x[nfi_plane_axis_b(p)], Hp, nullptr);
        nfi_footprint(ci[p][0], ci[p][1], Wp, Hp, f[p]);
        const float* P = planes + ((int64_t)b * 3 + p) * plane_sz;
        for (int c = 0; c < NFI_C; ++c)
          e[p][c] = nfi_bilerp(P[(f[p].y0 * Wp + f[p].x0) * NFI_C + c], P[(f[p].y0 * Wp + f[p].x1) * NFI_C + c],
                               P[(f[p].y1 * Wp + f[p].x0) * NFI_C + c], P[(f[p].y1 * Wp + f[p].x1) * NFI_C + c], f[p]);
      }
      for (int c = 0; c < NFI_C; ++c) feat[c] = nfi_mean3(e[0][c], e[1][c], e[2][c]);
      float hpre[NFI_HID], y[NFI_YPAD], sigma = 0.f, rgb[3] = {0, 0, 0}, probs[NFI_YPAD - 1];
      nfi_mlp_layer0<F4>(feat, W0kj.data(), b04.data(), hpre);
      nfi_mlp_layer1<F4>(hpre, W1jo.data(), b14.data(), no4, y);
      const float* pal = palette ? palette + (int64_t)b * A * 3 : nullptr;
      if (mode == 0) {
        nfi_heads_fwd(hd, y, keep, pal, sigma, rgb, probs);
        if (out) { out[id * 4] = sigma; out[id * 4 + 1] = rgb[0]; out[id * 4 + 2] = rgb[1]; out[id * 4 + 3] = rgb[2]; }
        if (probs_out) for (int a = 0; a < A; ++a) probs_out[id * A + a] = probs[a];
      } else if (out) {
        for (int o = 0; o < n_out; ++o) out[id * n_out + o] = y[o];
      }
      if (!g_out) continue;
      float dy[NFI_YPAD];
      if (mode == 0) {
        const float drgb[3] = {g_out[id * 4 + 1], g_out[id * 4 + 2], g_out[id * 4 + 3]};
        nfi_heads_bwd(hd, y, keep, pal, probs, rgb, g_out[id * 4], drgb, dy,
                      g_beta_alpha ? g_beta_alpha : nullptr, g_beta_alpha ? g_beta_alpha + 1 : nullptr);
        if (g_palette) for (int a = 0; a < A; ++a) for (int c = 0; c < 3; ++c) g_palette[((int64_t)b * A + a) * 3 + c] += probs[a] * drgb[c];
      } else {
        for (int o = 0; o < NFI_YPAD; ++o) dy[o] = o < n_out ? g_out[id * n_out + o] : 0.f;
      }
      nfi_mlp_layer1_bwd<F4>(hpre, W1jo.data(), no4, dy);
      float dfeat[NFI_C];
      nfi_mlp_layer0_bwd<F4>(hpre, W0jk.data(), dfeat);
      float gp[3] = {0, 0, 0};
      for (int p = 0; p < 3; ++p) {
        const float* P = planes + ((int64_t)b * 3 + p) * plane_sz;
        float* G = g_planes ? g_planes + ((int64_t)b * 3 + p) * plane_sz : nullptr;
        const int64_t o00 = (f[p].y0 * Wp + f[p].x0) * NFI_C, o01 = (f[p].y0 * Wp + f[p].x1) * NFI_C;
        const int64_t o10 = (f[p].y1 * Wp + f[p].x0) * NFI_C, o11 = (f[p].y1 * Wp + f[p].x1) * NFI_C;
        const float vx = (f[p].x0 + 1 < Wp) ? 1.f : 0.f, vy = (f[p].y0 + 1 < Hp) ? 1.f : 0.f;
        float gx = 0.f, gy = 0.f;
        for (int c = 0; c < NFI_C; ++c) {
          const float g = dfeat[c] * (1.f / 3.f);
          if (G) { G[o00 + c] += g * f[p].wnw; G[o01 + c] += g * f[p].wne; G[o10 + c] += g * f[p].wsw; G[o11 + c] += g * f[p].wse; }
          nfi_bilerp_grad(g, P[o00 + c], P[o01 + c], P[o10 + c], P[o11 + c], vx, vy, f[p], gx, gy);
        }
        gp[nfi_plane_axis_a(p)] += gx * nfi_coord_mult(ci[p][0], Wp);
        gp[nfi_plane_axis_b(p)] += gy * nfi_coord_mult(ci[p][1], Hp);
      }
      if (g_pts) for (int c = 0; c < 3; ++c) g_pts[id * 3 + c] = gp[c] / scene_range;
    }
}

}  // extern "C"
