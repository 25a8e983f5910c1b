// N3: the pose parametrisation the inversion loop optimises -- F.normalize on the
// quaternion + pose_utils.pose_to_matrix (run.py:2257-2262, lib/pose_utils.py:30-75) -- forward and
// backward as one kernel each (one thread per image) instead of ~30 ATen launches per step.
//   R[i,:] = e_i + 2 (w (u x e_i) + u x (u x e_i)),   (w, u) = q / max(|q|, 1e-12)
//   perspective: f = 1 + exp(z0), t3 = (t2 / s, f / s), focal = f / 2;   orthographic: t3 = (t2, 1) / s
//   cam2world[:3,:3] = R, cam2world[:3,3] = R t3, flipped camera: columns 1..3 of the top rows negated.
#include "nfi_common.cuh"

namespace {

constexpr int kThreads = 128;

__device__ __forceinline__ void cross3(const float* a, const float* b, float* c) {
  c[0] = a[1] * b[2] - a[2] * b[1];
  c[1] = a[2] * b[0] - a[0] * b[2];
  c[2] = a[0] * b[1] - a[1] * b[0];
}

struct PoseFwd {
  float w, u[3], inv_n;       // unit quaternion, 1 / max(|q|, eps) (1 when not normalising)
  float R[3][3];
  float t3[3], f;
};

__device__ __forceinline__ PoseFwd pose_eval(const float* z0, const float* t2, const float* s, const float* q,
                                             int b, int normalize_q) {
  PoseFwd p;
  float qq[4] = {q[b * 4], q[b * 4 + 1], q[b * 4 + 2], q[b * 4 + 3]};
  p.inv_n = 1.f;
  if (normalize_q) {
    const float n = sqrtf(qq[0] * qq[0] + qq[1] * qq[1] + qq[2] * qq[2] + qq[3] * qq[3]);
    p.inv_n = 1.f / fmaxf(n, 1e-12f);                    // F.normalize: x / max(|x|, eps)
  }
  p.w = qq[0] * p.inv_n;
#pragma unroll
  for (int i = 0; i < 3; ++i) p.u[i] = qq[1 + i] * p.inv_n;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    float e[3] = {0.f, 0.f, 0.f}, a[3], c[3];
    e[i] = 1.f;
    cross3(p.u, e, a);
    cross3(p.u, a, c);
#pragma unroll
    for (int j = 0; j < 3; ++j) p.R[i][j] = e[j] + 2.f * (p.w * a[j] + c[j]);
  }
  const float sb = s[b];
  if (z0) {
    p.f = 1.f + expf(z0[b]);
    p.t3[0] = t2[b * 2] / sb; p.t3[1] = t2[b * 2 + 1] / sb; p.t3[2] = p.f / sb;
  } else {
    p.f = 1.f;
    p.t3[0] = t2[b * 2] / sb; p.t3[1] = t2[b * 2 + 1] / sb; p.t3[2] = 1.f / sb;
  }
  return p;
}

__global__ void __launch_bounds__(kThreads)
pose_fwd_kernel(const float* __restrict__ z0, const float* __restrict__ t2, const float* __restrict__ s,
                const float* __restrict__ q, int B, int flipped, int normalize_q,
                float* __restrict__ cam, float* __restrict__ focal) {
  const int b = blockIdx.x * kThreads + threadIdx.x;
  if (b >= B) return;
  const PoseFwd p = pose_eval(z0, t2, s, q, b, normalize_q);
  const float sg = flipped ? -1.f : 1.f;
  float* m = cam + b * 16;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const float tr = (p.t3[0] * p.R[i][0] + p.t3[1] * p.R[i][1]) + p.t3[2] * p.R[i][2];
    m[i * 4 + 0] = p.R[i][0];
    m[i * 4 + 1] = sg * p.R[i][1];
    m[i * 4 + 2] = sg * p.R[i][2];
    m[i * 4 + 3] = sg * tr;
  }
  m[12] = 0.f; m[13] = 0.f; m[14] = 0.f; m[15] = 1.f;
  if (focal && z0) focal[b] = p.f * 0.5f;
}

__global__ void __launch_bounds__(kThreads)
pose_bwd_kernel(const float* __restrict__ z0, const float* __restrict__ t2, const float* __restrict__ s,
                const float* __restrict__ q, int B, int flipped, int normalize_q,
                const float* __restrict__ g_cam, const float* __restrict__ g_focal,
                float* __restrict__ g_z0, float* __restrict__ g_t2, float* __restrict__ g_s,
                float* __restrict__ g_q) {
  const int b = blockIdx.x * kThreads + threadIdx.x;
  if (b >= B) return;
  const PoseFwd p = pose_eval(z0, t2, s, q, b, normalize_q);
  const float sg = flipped ? -1.f : 1.f;
  const float* G = g_cam + b * 16;
  // gradient wrt the unflipped matrix, then wrt R and t3
  float gR[3][3], gt3[3] = {0.f, 0.f, 0.f};
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const float gtr = sg * G[i * 4 + 3];
    gR[i][0] = G[i * 4 + 0] + gtr * p.t3[0];
    gR[i][1] = sg * G[i * 4 + 1] + gtr * p.t3[1];
    gR[i][2] = sg * G[i * 4 + 2] + gtr * p.t3[2];
#pragma unroll
    for (int j = 0; j < 3; ++j) gt3[j] = fmaf(gtr, p.R[i][j], gt3[j]);
  }
  const float sb = s[b], inv_s = 1.f / sb;
  g_t2[b * 2] = gt3[0] * inv_s;
  g_t2[b * 2 + 1] = gt3[1] * inv_s;
  if (z0) {
    const float gf = gt3[2] * inv_s + (g_focal ? 0.5f * g_focal[b] : 0.f);
    if (g_z0) g_z0[b] = gf * (p.f - 1.f);                                        // d f / d z0 = exp(z0)
    g_s[b] = -(t2[b * 2] * gt3[0] + t2[b * 2 + 1] * gt3[1] + p.f * gt3[2]) * inv_s * inv_s;
  } else {
    g_s[b] = -(t2[b * 2] * gt3[0] + t2[b * 2 + 1] * gt3[1] + gt3[2]) * inv_s * inv_s;
  }
  // rotation: out_i = e_i + 2 (w a_i + c_i), a_i = u x e_i, c_i = u x a_i
  // for c = p x r:  d/dp = r x gc,  d/dr = gc x p
  float gw = 0.f, gu[3] = {0.f, 0.f, 0.f};
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    float e[3] = {0.f, 0.f, 0.f}, a[3];
    e[i] = 1.f;
    cross3(p.u, e, a);
    const float g2[3] = {2.f * gR[i][0], 2.f * gR[i][1], 2.f * gR[i][2]};        // d / d (w a_i + c_i)
    gw += g2[0] * a[0] + g2[1] * a[1] + g2[2] * a[2];
    float t[3], ga[3];
    cross3(a, g2, t);                                    // c_i = u x a_i: d/du = a_i x gc
#pragma unroll
    for (int j = 0; j < 3; ++j) gu[j] += t[j];
    cross3(g2, p.u, ga);                                 // ... d/da_i = gc x u
#pragma unroll
    for (int j = 0; j < 3; ++j) ga[j] = fmaf(p.w, g2[j], ga[j]);
    cross3(e, ga, t);                                    // a_i = u x e_i: d/du = e_i x ga
#pragma unroll
    for (int j = 0; j < 3; ++j) gu[j] += t[j];
  }
  float gq[4] = {gw, gu[0], gu[1], gu[2]};
  if (normalize_q) {                                     // qn = q / n: dq = (g - qn (qn . g)) / n
    const float dot = p.w * gq[0] + p.u[0] * gq[1] + p.u[1] * gq[2] + p.u[2] * gq[3];
    const float qn[4] = {p.w, p.u[0], p.u[1], p.u[2]};
#pragma unroll
    for (int j = 0; j < 4; ++j) gq[j] = (gq[j] - qn[j] * dot) * p.inv_n;
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) g_q[b * 4 + j] = gq[j];
}

}  // namespace

extern "C" int nfi_pose_to_matrix_fwd(const float* z0, const float* t2, const float* s, const float* q, int B,
                                      int camera_flipped, int normalize_q, float* cam2world, float* focal,
                                      nfi_stream_t stream) {
  NFI_REQUIRE(t2 && s && q && cam2world, NFI_ERR_BAD_ARG, "nfi_pose_to_matrix_fwd: NULL pointer");
  NFI_REQUIRE(B > 0, NFI_ERR_BAD_ARG, "nfi_pose_to_matrix_fwd: B must be positive");
  NFI_REQUIRE(!z0 || focal, NFI_ERR_BAD_ARG, "nfi_pose_to_matrix_fwd: perspective poses need the focal output");
  pose_fwd_kernel<<<(B + kThreads - 1) / kThreads, kThreads, 0, nfi_cu(stream)>>>(z0, t2, s, q, B, camera_flipped,
                                                                                normalize_q, cam2world, focal);
  NFI_CHECK_LAUNCH("pose_fwd_kernel");
  return NFI_OK;
}

extern "C" int nfi_pose_to_matrix_bwd(const float* z0, const float* t2, const float* s, const float* q, int B,
                                      int camera_flipped, int normalize_q, const float* g_cam2world,
                                      const float* g_focal, float* g_z0, float* g_t2, float* g_s, float* g_q,
                                      nfi_stream_t stream) {
  NFI_REQUIRE(t2 && s && q && g_cam2world && g_t2 && g_s && g_q, NFI_ERR_BAD_ARG, "nfi_pose_to_matrix_bwd: NULL pointer");
  NFI_REQUIRE(B > 0, NFI_ERR_BAD_ARG, "nfi_pose_to_matrix_bwd: B must be positive");
  pose_bwd_kernel<<<(B + kThreads - 1) / kThreads, kThreads, 0, nfi_cu(stream)>>>(
      z0, t2, s, q, B, camera_flipped, normalize_q, g_cam2world, g_focal, g_z0, g_t2, g_s, g_q);
  NFI_CHECK_LAUNCH("pose_bwd_kernel");
  return NFI_OK;
}
