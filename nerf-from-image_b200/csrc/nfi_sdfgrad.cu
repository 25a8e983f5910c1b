// the regulariser path of the generator (models/generator.py:505-585, lib/ops.py:58-120) needs the
// gradient of the decoder output with respect to the query coordinates (the eikonal term |grad_x sdf| -> 1) AND the
// gradient of that gradient with respect to the planes and the decoder weights (double backward).  The reference gets
// both from autograd through a double-differentiable restatement of grid_sample; here they are two hand-written kernels:
//
//   decoder_gradx_fwd   gx[n][k]  = sum_o gy[n][o] d y_o / d x_k                      (J^T gy, k = 0..2)
//   decoder_gradx_bwd   given gg[n][k]: gradients of  L2 = sum_n sum_k gg[n][k] gx[n][k]  w.r.t. planes, W0, b0, W1
//
// with  y = W1s softplus(W0s F(x) + b0) + b1,  F = mean over the three planes of the bilinear sample,  s = sigmoid(h):
//   gx_k = sum_j v_j s_j u_kj,      v = W1s^T gy,  u_k = W0s T_k,  T_k = dF / dx_k
//   L2   = sum_j v_j s_j q_j,       q = W0s Tg,    Tg = sum_k gg_k T_k
//   dL2/dh_j = v_j s_j (1 - s_j) q_j = a_j         dL2/dq_j = v_j s_j = r_j       dL2/dv_j = s_j q_j
//   dL2/dW0s[j][c] = a_j F_c + r_j Tg_c            dL2/db0_j = a_j                dL2/dW1s[o][j] = gy_o s_j q_j
//   dL2/dF_c = sum_j a_j W0s[j][c]                 dL2/dTg_c = sum_j r_j W0s[j][c]
// and F, Tg are linear in the texels.  The workload is small (31^3 points per image), so this is a plain SIMT kernel:
// one warp per point, lane = feature channel (a texel is one coalesced 128-byte load), the 64 x 32 matrix-vector
// products by shuffle broadcast against shared-memory copies of W0s.  Coordinates are in box units ([-1, 1]); like
// lib/ops.py::grid_sample2d the coordinate is NOT clamped (d ix / d x = (W - 1) / 2 everywhere), indices are.
#include "nfi_common.cuh"
#include "nfi_field_math.cuh"

namespace {

constexpr int kC = 32, kH = 64, kMaxOut = 16, kThreads = 128, kWarps = kThreads / 32;

struct GxSmem {
  float w0t[kC][kH];        // W0s[j][c] stored [c][j]: lanes read consecutive j
  float w0[kH][kC];         // W0s[j][c] stored [j][c]: lanes read consecutive c
  float w1[kMaxOut][kH];    // W1s[o][j]
  float b0[kH];
};

__device__ __forceinline__ void gx_load_weights(GxSmem& sm, const float* W0, const float* b0, const float* W1, int n_out) {
  const float g0 = rsqrtf((float)kC), g1 = rsqrtf((float)kH);
  for (int i = threadIdx.x; i < kH * kC; i += blockDim.x) {
    const int j = i / kC, c = i - j * kC;
    const float w = W0[i] * g0;
    sm.w0[j][c] = w; sm.w0t[c][j] = w;
  }
  for (int i = threadIdx.x; i < kMaxOut * kH; i += blockDim.x) {
    const int o = i / kH, j = i - o * kH;
    sm.w1[o][j] = o < n_out ? W1[o * kH + j] * g1 : 0.f;
  }
  for (int i = threadIdx.x; i < kH; i += blockDim.x) sm.b0[i] = b0[i];
}

// Bilinear footprint of one coordinate pair on a plane (grid_sample2d, lib/ops.py:66-91)
struct Foot {
  int o00, o01, o10, o11;   // texel indices (clamped)
  float tx, ty;
};
__device__ __forceinline__ Foot gx_foot(float xa, float xb, int Wp, int Hp, int bshift, int nbx) {
  const float ix = ((xa + 1.f) * 0.5f) * (float)(Wp - 1), iy = ((xb + 1.f) * 0.5f) * (float)(Hp - 1);
  const float fx = floorf(ix), fy = floorf(iy);
  const int x0 = min(max((int)fx, 0), Wp - 1), x1 = min(max((int)fx + 1, 0), Wp - 1);
  const int y0 = min(max((int)fy, 0), Hp - 1), y1 = min(max((int)fy + 1, 0), Hp - 1);
  Foot f;
  const int r0 = nfi_texel_row(y0, bshift, nbx), r1 = nfi_texel_row(y1, bshift, nbx);
  const int c0 = nfi_texel_col(x0, bshift), c1 = nfi_texel_col(x1, bshift);
  f.o00 = r0 + c0; f.o01 = r0 + c1; f.o10 = r1 + c0; f.o11 = r1 + c1;
  f.tx = ix - fx; f.ty = iy - fy;
  return f;
}

// Per point, per lane (= channel): F_c, T_kc (k = 0..2) and the footprints.
struct PointFeat {
  float F, T[3];
  Foot ft[3];
};
__device__ __forceinline__ PointFeat gx_features(const float* __restrict__ planes_img, const float x[3], int Hp, int Wp,
                                                 int bshift, int nbx, int lane) {
  PointFeat pf;
  pf.F = 0.f; pf.T[0] = pf.T[1] = pf.T[2] = 0.f;
  const float third = 1.f / 3.f;
  const float mx = 0.5f * (float)(Wp - 1), my = 0.5f * (float)(Hp - 1);
  const int64_t plane = (int64_t)Hp * Wp * kC;
#pragma unroll
  for (int p = 0; p < 3; ++p) {
    const int a = nfi_plane_axis_a(p), b = nfi_plane_axis_b(p);
    const Foot f = gx_foot(x[a], x[b], Wp, Hp, bshift, nbx);
    pf.ft[p] = f;
    const float* P = planes_img + p * plane + lane;
    const float v0 = __ldg(P + (int64_t)f.o00 * kC), v1 = __ldg(P + (int64_t)f.o01 * kC);
    const float v2 = __ldg(P + (int64_t)f.o10 * kC), v3 = __ldg(P + (int64_t)f.o11 * kC);
    const float ex = v1 - v0, ey = v2 - v0, exy = (v3 - v2) - ex;
    const float dix = fmaf(exy, f.ty, ex), diy = fmaf(exy, f.tx, ey);
    pf.F += third * fmaf(f.tx, dix, fmaf(f.ty, ey, v0));
    pf.T[a] = fmaf(third * mx, dix, pf.T[a]);
    pf.T[b] = fmaf(third * my, diy, pf.T[b]);
  }
  return pf;
}

// h_j, u_kj for j = lane, lane + 32 from F_c, T_kc held one channel per lane
__device__ __forceinline__ void gx_matvec(const GxSmem& sm, const PointFeat& pf, int lane, float (&h)[2], float (&u)[3][2]) {
  h[0] = sm.b0[lane]; h[1] = sm.b0[lane + 32];
#pragma unroll
  for (int k = 0; k < 3; ++k) u[k][0] = u[k][1] = 0.f;
#pragma unroll 8
  for (int c = 0; c < kC; ++c) {
    const float w_a = sm.w0t[c][lane], w_b = sm.w0t[c][lane + 32];
    const float Fc = __shfl_sync(0xffffffffu, pf.F, c);
    h[0] = fmaf(w_a, Fc, h[0]); h[1] = fmaf(w_b, Fc, h[1]);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const float Tc = __shfl_sync(0xffffffffu, pf.T[k], c);
      u[k][0] = fmaf(w_a, Tc, u[k][0]); u[k][1] = fmaf(w_b, Tc, u[k][1]);
    }
  }
}

__device__ __forceinline__ float gx_sigmoid(float h) { return 1.f / (1.f + expf(-h)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__global__ void __launch_bounds__(kThreads)
decoder_gradx_fwd_kernel(const float* __restrict__ planes_cl, const float* __restrict__ W0, const float* __restrict__ b0,
                         const float* __restrict__ W1, int n_out, const float* __restrict__ coords,
                         const float* __restrict__ gy, int B, int64_t N, int Hp, int Wp, float* __restrict__ gx) {
  __shared__ GxSmem sm;
  gx_load_weights(sm, W0, b0, W1, n_out);
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int bshift = nfi_plane_bshift(Hp, Wp), nbx = Wp >> bshift;
  const int64_t total = (int64_t)B * N, nwarps = (int64_t)gridDim.x * kWarps;
  for (int64_t n = (int64_t)blockIdx.x * kWarps + warp; n < total; n += nwarps) {
    const int64_t b = n / N;
    const float x[3] = {coords[n * 3], coords[n * 3 + 1], coords[n * 3 + 2]};
    const PointFeat pf = gx_features(planes_cl + b * 3 * Hp * Wp * kC, x, Hp, Wp, bshift, nbx, lane);
    float h[2], u[3][2];
    gx_matvec(sm, pf, lane, h, u);
    float v[2] = {0.f, 0.f};
    for (int o = 0; o < n_out; ++o) {
      const float g = gy[n * n_out + o];
      v[0] = fmaf(g, sm.w1[o][lane], v[0]); v[1] = fmaf(g, sm.w1[o][lane + 32], v[1]);
    }
    const float vs0 = v[0] * gx_sigmoid(h[0]), vs1 = v[1] * gx_sigmoid(h[1]);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const float r = warp_sum(fmaf(vs0, u[k][0], vs1 * u[k][1]));
      if (lane == 0) gx[n * 3 + k] = r;
    }
  }
}

__global__ void __launch_bounds__(kThreads)
decoder_gradx_bwd_kernel(const float* __restrict__ planes_cl, const float* __restrict__ W0, const float* __restrict__ b0,
                         const float* __restrict__ W1, int n_out, const float* __restrict__ coords,
                         const float* __restrict__ gy, const float* __restrict__ gg, int B, int64_t N, int Hp, int Wp,
                         float* __restrict__ g_planes_cl, float* __restrict__ g_W0, float* __restrict__ g_b0,
                         float* __restrict__ g_W1) {
  __shared__ GxSmem sm;
  gx_load_weights(sm, W0, b0, W1, n_out);
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int bshift = nfi_plane_bshift(Hp, Wp), nbx = Wp >> bshift;
  const int64_t total = (int64_t)B * N, nwarps = (int64_t)gridDim.x * kWarps;
  const float third = 1.f / 3.f;
  const float mx = 0.5f * (float)(Wp - 1), my = 0.5f * (float)(Hp - 1);
  const int64_t plane = (int64_t)Hp * Wp * kC;
  float gw0[kH];                     // d W0s[j][lane], accumulated over this warp's points
#pragma unroll
  for (int j = 0; j < kH; ++j) gw0[j] = 0.f;
  float gb[2] = {0.f, 0.f};          // d b0[lane], d b0[lane + 32]
  float gw1[kMaxOut][2];             // d W1s[o][lane], [o][lane + 32]
#pragma unroll
  for (int o = 0; o < kMaxOut; ++o) gw1[o][0] = gw1[o][1] = 0.f;
  for (int64_t n = (int64_t)blockIdx.x * kWarps + warp; n < total; n += nwarps) {
    const int64_t b = n / N;
    const float x[3] = {coords[n * 3], coords[n * 3 + 1], coords[n * 3 + 2]};
    const float g3[3] = {gg[n * 3], gg[n * 3 + 1], gg[n * 3 + 2]};
    const PointFeat pf = gx_features(planes_cl + b * 3 * Hp * Wp * kC, x, Hp, Wp, bshift, nbx, lane);
    float h[2], u[3][2];
    gx_matvec(sm, pf, lane, h, u);
    float v[2] = {0.f, 0.f};
    float gyo[kMaxOut];
#pragma unroll
    for (int o = 0; o < kMaxOut; ++o) {
      gyo[o] = o < n_out ? gy[n * n_out + o] : 0.f;
      v[0] = fmaf(gyo[o], sm.w1[o][lane], v[0]); v[1] = fmaf(gyo[o], sm.w1[o][lane + 32], v[1]);
    }
    float a[2], r[2];
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const float s = gx_sigmoid(h[e]);
      const float q = fmaf(g3[0], u[0][e], fmaf(g3[1], u[1][e], g3[2] * u[2][e]));
      r[e] = v[e] * s;
      a[e] = r[e] * (1.f - s) * q;
      gb[e] += a[e];
      const float sq = s * q;
#pragma unroll
      for (int o = 0; o < kMaxOut; ++o) gw1[o][e] = fmaf(gyo[o], sq, gw1[o][e]);
    }
    const float Tg = fmaf(g3[0], pf.T[0], fmaf(g3[1], pf.T[1], g3[2] * pf.T[2]));
    float dF = 0.f, dTg = 0.f;       // for channel c = lane
#pragma unroll
    for (int j = 0; j < kH; ++j) {
      const float aj = __shfl_sync(0xffffffffu, a[j >> 5], j & 31), rj = __shfl_sync(0xffffffffu, r[j >> 5], j & 31);
      const float w = sm.w0[j][lane];
      dF = fmaf(aj, w, dF); dTg = fmaf(rj, w, dTg);
      gw0[j] = fmaf(aj, pf.F, fmaf(rj, Tg, gw0[j]));
    }
    if (g_planes_cl) {
      float* G = g_planes_cl + b * 3 * plane + lane;
#pragma unroll
      for (int p = 0; p < 3; ++p) {
        const Foot f = pf.ft[p];
        const float ga = g3[nfi_plane_axis_a(p)] * mx, gb_ = g3[nfi_plane_axis_b(p)] * my;
        // F = v0 (1-tx)(1-ty) + v1 tx (1-ty) + v2 (1-tx) ty + v3 tx ty;  dF/dix = (v1-v0)(1-ty) + (v3-v2) ty;  dF/diy likewise
        const float ex = 1.f - f.tx, ey = 1.f - f.ty;
        const float c0 = dF * (ex * ey) + dTg * (-ga * ey - gb_ * ex);
        const float c1 = dF * (f.tx * ey) + dTg * (ga * ey - gb_ * f.tx);
        const float c2 = dF * (ex * f.ty) + dTg * (-ga * f.ty + gb_ * ex);
        const float c3 = dF * (f.tx * f.ty) + dTg * (ga * f.ty + gb_ * f.tx);
        float* P = G + p * plane;
        atomicAdd(P + (int64_t)f.o00 * kC, third * c0);
        atomicAdd(P + (int64_t)f.o01 * kC, third * c1);
        atomicAdd(P + (int64_t)f.o10 * kC, third * c2);
        atomicAdd(P + (int64_t)f.o11 * kC, third * c3);
      }
    }
  }
  // flush the weight gradients (EqualizedLinear gains: d W = g d Ws)
  const float g0 = rsqrtf((float)kC), g1 = rsqrtf((float)kH);
  if (g_W0) {
#pragma unroll
    for (int j = 0; j < kH; ++j) if (gw0[j] != 0.f) atomicAdd(&g_W0[j * kC + lane], gw0[j] * g0);
  }
  if (g_b0) {
    if (gb[0] != 0.f) atomicAdd(&g_b0[lane], gb[0]);
    if (gb[1] != 0.f) atomicAdd(&g_b0[lane + 32], gb[1]);
  }
  if (g_W1) {
#pragma unroll
    for (int o = 0; o < kMaxOut; ++o) {
      if (o < n_out) {
        if (gw1[o][0] != 0.f) atomicAdd(&g_W1[o * kH + lane], gw1[o][0] * g1);
        if (gw1[o][1] != 0.f) atomicAdd(&g_W1[o * kH + lane + 32], gw1[o][1] * g1);
      }
    }
  }
}

inline int gx_grid(int64_t points) {
  const int64_t blocks = (points + kWarps - 1) / kWarps;
  const int64_t cap = (int64_t)nfi_sm_count() * 8;
  return (int)(blocks < 1 ? 1 : (blocks < cap ? blocks : cap));
}

}  // namespace

extern "C" int nfi_decoder_gradx_fwd(const NfiField* field, const float* planes_cl, int B, int Hp, int Wp,
                                     const float* coords, const float* gy, int64_t n_per_image, float* gx,
                                     nfi_stream_t stream) {
  NFI_REQUIRE(field && field->W0 && field->b0 && field->W1 && planes_cl && coords && gy && gx, NFI_ERR_BAD_ARG,
              "nfi_decoder_gradx_fwd: NULL pointer");
  NFI_REQUIRE(B > 0 && Hp >= 2 && Wp >= 2 && n_per_image > 0, NFI_ERR_BAD_ARG, "nfi_decoder_gradx_fwd: bad sizes");
  NFI_REQUIRE(field->n_out >= 1 && field->n_out <= NFI_MAX_OUT, NFI_ERR_UNSUPPORTED, "nfi_decoder_gradx_fwd: n_out = %d", field->n_out);
  NFI_REQUIRE(!field->planes_bf16, NFI_ERR_UNSUPPORTED, "nfi_decoder_gradx_fwd: fp32 planes only");
  decoder_gradx_fwd_kernel<<<gx_grid((int64_t)B * n_per_image), kThreads, 0, nfi_cu(stream)>>>(
      planes_cl, field->W0, field->b0, field->W1, field->n_out, coords, gy, B, n_per_image, Hp, Wp, gx);
  NFI_CHECK_LAUNCH("decoder_gradx_fwd_kernel");
  return NFI_OK;
}

extern "C" int nfi_decoder_gradx_bwd(const NfiField* field, const float* planes_cl, int B, int Hp, int Wp,
                                     const float* coords, const float* gy, const float* gg, int64_t n_per_image,
                                     float* g_planes_cl, float* g_W0, float* g_b0, float* g_W1, nfi_stream_t stream) {
  NFI_REQUIRE(field && field->W0 && field->b0 && field->W1 && planes_cl && coords && gy && gg, NFI_ERR_BAD_ARG,
              "nfi_decoder_gradx_bwd: NULL pointer");
  NFI_REQUIRE(B > 0 && Hp >= 2 && Wp >= 2 && n_per_image > 0, NFI_ERR_BAD_ARG, "nfi_decoder_gradx_bwd: bad sizes");
  NFI_REQUIRE(field->n_out >= 1 && field->n_out <= NFI_MAX_OUT, NFI_ERR_UNSUPPORTED, "nfi_decoder_gradx_bwd: n_out = %d", field->n_out);
  NFI_REQUIRE(!field->planes_bf16, NFI_ERR_UNSUPPORTED, "nfi_decoder_gradx_bwd: fp32 planes only");
  decoder_gradx_bwd_kernel<<<gx_grid((int64_t)B * n_per_image), kThreads, 0, nfi_cu(stream)>>>(
      planes_cl, field->W0, field->b0, field->W1, field->n_out, coords, gy, gg, B, n_per_image, Hp, Wp, g_planes_cl,
      g_W0, g_b0, g_W1);
  NFI_CHECK_LAUNCH("decoder_gradx_bwd_kernel");
  return NFI_OK;
}
