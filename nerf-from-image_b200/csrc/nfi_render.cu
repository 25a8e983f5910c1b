// A11 / A12 behind ONE pair of C-ABI calls: run.py::render (run.py:176-350) forward and its backward,
// sequencing the stage kernels of this library on the caller's stream inside a caller-owned workspace.
// This is what a non-Python host (or a no_grad forward-only caller) binds; the Python autograd path
// (ops.py / render.py) launches exactly the same kernels stage by stage.
//
// Nothing is allocated here and nothing synchronises: every intermediate (rays, near/far, depths, one
// float4 per sample, the merged order) lives in `workspace`, whose layout is private to this file --
// except its first 4 bytes, which hold the number of rays that hit the scene box once the forward has
// run (stream order): 0 is the case in which the reference raises (lib/nerf_utils.py:260).
#include "nfi_common.cuh"
#include "nfi_math.cuh"

namespace {

constexpr int kThreads = 256;

// u[i] = U[0,1) from Philox4x32-10 (seed, counter = i / 4, offset): the uniforms of
// nerf_utils.sample_pdf (lib/nerf_utils.py:204) when the caller supplies none
__global__ void __launch_bounds__(kThreads)
uniform_fill_kernel(uint64_t seed, uint64_t offset, int64_t n4, float4* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n4; i += (int64_t)gridDim.x * kThreads) {
    uint32_t r[4];
    nfi_philox(seed, (uint64_t)i, offset, r);
    out[i] = make_float4(nfi_u01(r[0]), nfi_u01(r[1]), nfi_u01(r[2]), nfi_u01(r[3]));
  }
}

__global__ void __launch_bounds__(kThreads)
add_inplace_kernel(float* __restrict__ a, const float* __restrict__ b, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kThreads)
    a[i] += b[i];
}

inline int grid_for(int64_t n) {
  const int64_t blocks = (n + kThreads - 1) / kThreads;
  const int64_t cap = (int64_t)nfi_sm_count() * 8;
  return (int)(blocks < 1 ? 1 : (blocks < cap ? blocks : cap));
}

struct Layout {               // byte offsets into the workspace (256-byte aligned)
  size_t nf, origins, dirs, near, far, z_c, coarse, z_f, fine, u_pdf, perm, z_sorted;
  size_t g_coarse, g_fine, g_points, g_o, g_d, g_o2, g_d2;
  size_t fwd_bytes, total_bytes;
};

inline size_t up(size_t x) { return (x + 255) & ~(size_t)255; }

Layout layout_of(const NfiRender& p) {
  const size_t R = (size_t)p.B * p.H * p.W, S = (size_t)p.S, f = sizeof(float);
  Layout L;
  size_t o = 0;
  auto take = [&](size_t bytes) { const size_t at = o; o = up(o + bytes); return at; };
  L.nf = take(NFI_NEARFAR_WS_BYTES);
  L.origins = take(R * 3 * f); L.dirs = take(R * 3 * f);
  L.near = take(R * f); L.far = take(R * f);
  L.z_c = take(R * S * f); L.coarse = take(R * S * 4 * f);
  L.z_f = take(R * S * f); L.fine = take(R * S * 4 * f);
  L.u_pdf = take(R * S * f);
  L.perm = take(R * 2 * S); L.z_sorted = take(R * 2 * S * f);
  L.fwd_bytes = o;
  L.g_coarse = take(R * S * 4 * f); L.g_fine = take(R * S * 4 * f);
  L.g_points = take(R * S * 3 * f);
  L.g_o = take(R * 3 * f); L.g_d = take(R * 3 * f); L.g_o2 = take(R * 3 * f); L.g_d2 = take(R * 3 * f);
  L.total_bytes = o;
  return L;
}

int check(const NfiRender* p, const char* who) {
  NFI_REQUIRE(p, NFI_ERR_BAD_ARG, "%s: NULL parameter block", who);
  NFI_REQUIRE(p->planes_cl && p->cam2world && p->field.W0 && p->field.b0 && p->field.W1 && p->field.b1,
              NFI_ERR_BAD_ARG, "%s: NULL planes / camera / decoder pointer", who);
  NFI_REQUIRE(p->B > 0 && p->H > 0 && p->W > 0 && p->Hp > 0 && p->Wp > 0, NFI_ERR_BAD_ARG, "%s: bad sizes", who);
  NFI_REQUIRE(p->S >= 4 && p->S <= NFI_MAX_S && p->S % 4 == 0, NFI_ERR_UNSUPPORTED,
              "%s: S = %d (need a multiple of 4 in [4, %d])", who, p->S, NFI_MAX_S);
  NFI_REQUIRE(p->field.attention_values == 0 || p->palette, NFI_ERR_BAD_ARG, "%s: attention head without palette", who);
  return NFI_OK;
}

template <typename T>
T* at(void* ws, size_t off) { return reinterpret_cast<T*>(static_cast<char*>(ws) + off); }
template <typename T>
const T* at(const void* ws, size_t off) { return reinterpret_cast<const T*>(static_cast<const char*>(ws) + off); }

}  // namespace

extern "C" size_t nfi_render_workspace_bytes(const NfiRender* p, int with_backward) {
  if (!p || p->B <= 0 || p->H <= 0 || p->W <= 0 || p->S <= 0) return 0;
  const Layout L = layout_of(*p);
  return with_backward ? L.total_bytes : L.fwd_bytes;
}

extern "C" int nfi_render_fwd(const NfiRender* p, void* workspace, float* rgb, float* depth, float* mask,
                              nfi_stream_t stream) {
  int rc = check(p, "nfi_render_fwd");
  if (rc) return rc;
  NFI_REQUIRE(workspace && rgb && depth && mask, NFI_ERR_BAD_ARG, "nfi_render_fwd: NULL workspace / output");
  NFI_REQUIRE((((uintptr_t)workspace) & 255u) == 0, NFI_ERR_ALIGNMENT, "nfi_render_fwd: workspace must be 256-byte aligned");
  const Layout L = layout_of(*p);
  const int64_t R = (int64_t)p->B * p->H * p->W, npi = (int64_t)p->H * p->W;
  const int S = p->S;
  float* o = at<float>(workspace, L.origins); float* d = at<float>(workspace, L.dirs);
  float* z_c = at<float>(workspace, L.z_c);   float* z_f = at<float>(workspace, L.z_f);
  float* coarse = at<float>(workspace, L.coarse); float* fine = at<float>(workspace, L.fine);
  // A1 (+ F.normalize, run.py:196), A2, A3
  rc = nfi_ray_bundle_fwd(p->cam2world, p->focal, p->focal_stride, p->center, p->bbox, p->B, p->H, p->W, 1, o, d, stream);
  if (rc) return rc;
  rc = nfi_near_far(o, d, R, p->field.scene_range, at<float>(workspace, L.near), at<float>(workspace, L.far),
                    at<void>(workspace, L.nf), stream);
  if (rc) return rc;
  rc = nfi_stratified(at<float>(workspace, L.near), at<float>(workspace, L.far), p->u_strat, R, S, p->randomize,
                      p->seed, p->offset, z_c, stream);
  if (rc) return rc;
  // A4-A6 on the coarse samples
  const int img_w = (p->H % 8 == 0 && p->W % 16 == 0) ? p->W : 0;
  rc = nfi_field_fwd(&p->field, p->planes_cl, p->palette, p->B, p->Hp, p->Wp, o, d, z_c, nullptr, npi, S, 0, img_w, 1,
                     coarse, nullptr, stream);
  if (rc) return rc;
  // A7 + A8 (no grad, run.py:261-281)
  const float* u_pdf = nullptr;
  if (p->randomize) {
    u_pdf = p->u_pdf;
    if (!u_pdf) {
      float* gen = at<float>(workspace, L.u_pdf);
      uniform_fill_kernel<<<grid_for(R * S / 4), kThreads, 0, nfi_cu(stream)>>>(
          p->seed ^ 0x9E3779B97F4A7C15ull, p->offset, R * S / 4, reinterpret_cast<float4*>(gen));
      NFI_CHECK_LAUNCH("uniform_fill_kernel");
      u_pdf = gen;
    }
  }
  rc = nfi_importance(coarse, 4, d, z_c, R, S, u_pdf, z_f, stream);
  if (rc) return rc;
  rc = nfi_field_fwd(&p->field, p->planes_cl, p->palette, p->B, p->Hp, p->Wp, o, d, z_f, nullptr, npi, S, 0, img_w, 1,
                     fine, nullptr, stream);
  if (rc) return rc;
  // A9 + A10
  return nfi_merge_composite_fwd(coarse, fine, z_c, z_f, d, R, S, p->white_background,
                                 p->zbuffer_depth ? p->cam2world : nullptr, npi, rgb, depth, mask,
                                 at<uint8_t>(workspace, L.perm), at<float>(workspace, L.z_sorted), stream);
}

extern "C" int nfi_render_bwd(const NfiRender* p, void* workspace, const float* g_rgb, const float* g_mask,
                              float* g_planes_cl, float* g_palette, float* g_cam2world, float* g_focal,
                              float* g_params, int want_weight_grads, nfi_stream_t stream) {
  int rc = check(p, "nfi_render_bwd");
  if (rc) return rc;
  NFI_REQUIRE(workspace && g_rgb, NFI_ERR_BAD_ARG, "nfi_render_bwd: NULL workspace / upstream gradient");
  NFI_REQUIRE((g_cam2world != nullptr) == (g_focal != nullptr), NFI_ERR_BAD_ARG,
              "nfi_render_bwd: g_cam2world and g_focal come together (both or neither)");
  const Layout L = layout_of(*p);
  const int64_t R = (int64_t)p->B * p->H * p->W, npi = (int64_t)p->H * p->W;
  const int S = p->S;
  const float* o = at<float>(workspace, L.origins); const float* d = at<float>(workspace, L.dirs);
  const float* z_c = at<float>(workspace, L.z_c);   const float* z_f = at<float>(workspace, L.z_f);
  const float* coarse = at<float>(workspace, L.coarse); const float* fine = at<float>(workspace, L.fine);
  float* g_coarse = at<float>(workspace, L.g_coarse); float* g_fine = at<float>(workspace, L.g_fine);
  const bool pose = g_cam2world != nullptr;
  float* g_points = pose ? at<float>(workspace, L.g_points) : nullptr;
  const int img_w = (p->H % 8 == 0 && p->W % 16 == 0) ? p->W : 0;
  rc = nfi_merge_composite_bwd(coarse, fine, at<float>(workspace, L.z_sorted), at<uint8_t>(workspace, L.perm), d, R, S,
                               p->white_background, g_rgb, g_mask, g_coarse, g_fine, stream);
  if (rc) return rc;
  // fine pass, then coarse pass: both accumulate into the caller's (zero-filled) plane / palette /
  // parameter gradients; the per-sample point gradients are reduced per ray right after each pass
  for (int pass = 0; pass < 2; ++pass) {
    const float* z = pass == 0 ? z_f : z_c;
    rc = nfi_field_bwd(&p->field, p->planes_cl, p->palette, p->B, p->Hp, p->Wp, o, d, z, nullptr, npi, S, 0, img_w,
                       pass == 0 ? g_fine : g_coarse, g_planes_cl, g_palette, g_points, g_params,
                       want_weight_grads, stream);
    if (rc) return rc;
    if (pose) {
      rc = nfi_points_bwd(g_points, z, R, S, at<float>(workspace, pass == 0 ? L.g_o : L.g_o2),
                          at<float>(workspace, pass == 0 ? L.g_d : L.g_d2), stream);
      if (rc) return rc;
    }
  }
  if (pose) {
    // g_o | g_d and g_o2 | g_d2 are adjacent only up to padding: two small adds
    add_inplace_kernel<<<grid_for(R * 3), kThreads, 0, nfi_cu(stream)>>>(at<float>(workspace, L.g_o), at<float>(workspace, L.g_o2), R * 3);
    add_inplace_kernel<<<grid_for(R * 3), kThreads, 0, nfi_cu(stream)>>>(at<float>(workspace, L.g_d), at<float>(workspace, L.g_d2), R * 3);
    NFI_CHECK_LAUNCH("add_inplace_kernel");
    rc = nfi_ray_bundle_bwd(p->cam2world, p->focal, p->focal_stride, p->center, p->bbox, p->B, p->H, p->W, 1,
                            at<float>(workspace, L.g_o), at<float>(workspace, L.g_d), g_cam2world, g_focal, stream);
    if (rc) return rc;
  }
  return NFI_OK;
}
