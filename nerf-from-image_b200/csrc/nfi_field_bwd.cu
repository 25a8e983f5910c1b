// Backward entry point of the triplane field kernels (kernels are instantiated in
// nfi_field_bwd_heads.cu / nfi_field_bwd_raw.cu so that the translation units build in parallel).
#include "nfi_field_host.cuh"

void nfi_field_bwd_launch_heads(const nfi_field::Args& a, int grid, cudaStream_t st);
void nfi_field_bwd_launch_raw(const nfi_field::Args& a, int grid, cudaStream_t st);

extern "C" int nfi_field_bwd(const NfiField* field, const float* planes_cl, const float* palette, int B,
                             int Hp, int Wp, const float* origins, const float* dirs, const float* z,
                             const float* points, int64_t n_per_image, int S, int mode, int img_w,
                             const float* g_out, float* g_planes_cl, float* g_palette, float* g_points,
                             float* g_params, nfi_stream_t stream) {
  nfi_field::Args a;
  int rc = fill_args(a, "nfi_field_bwd", field, planes_cl, palette, B, Hp, Wp, origins, dirs, z, points,
                     n_per_image, S, mode, img_w);
  if (rc) return rc;
  NFI_REQUIRE(g_out, NFI_ERR_BAD_ARG, "nfi_field_bwd: NULL upstream gradient");
  NFI_REQUIRE(mode == 1 || nfi_aligned16(g_out), NFI_ERR_ALIGNMENT, "nfi_field_bwd: g_out must be 16-byte aligned");
  NFI_REQUIRE(!g_planes_cl || nfi_aligned16(g_planes_cl), NFI_ERR_ALIGNMENT, "nfi_field_bwd: g_planes must be 16-byte aligned");
  a.g_out = g_out; a.g_planes = g_planes_cl; a.g_palette = (mode == 0) ? g_palette : nullptr;
  a.g_points = g_points; a.g_params = g_params;
  const int grid = field_grid(a.n_tiles, 3);
  cudaStream_t st = nfi_cu(stream);
  if (mode == 0) nfi_field_bwd_launch_heads(a, grid, st);
  else nfi_field_bwd_launch_raw(a, grid, st);
  NFI_CHECK_LAUNCH("field_bwd_kernel");
  return NFI_OK;
}
