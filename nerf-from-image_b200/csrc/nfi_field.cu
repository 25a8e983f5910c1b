// Forward entry point of the triplane field kernels + the NCHW <-> channels-last layout passes.
#include "nfi_field_host.cuh"

namespace {

// [B*3][32][HW] -> [B*3][HW][32] (and back).  32 x 32 tiles through padded shared memory: both the
// global reads and the global writes are 128-byte coalesced.
template <bool TO_CL>
__global__ void __launch_bounds__(256)
plane_transpose_kernel(const float* __restrict__ src, float* __restrict__ dst, int HW) {
  __shared__ float tile[32][33];
  const int64_t plane = blockIdx.y;
  const int p0 = blockIdx.x * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8
  const float* s = src + plane * 32 * (int64_t)HW;
  float* d = dst + plane * 32 * (int64_t)HW;
  if (TO_CL) {
#pragma unroll
    for (int c = ty; c < 32; c += 8) tile[c][tx] = (p0 + tx < HW) ? s[(int64_t)c * HW + p0 + tx] : 0.f;
    __syncthreads();
#pragma unroll
    for (int p = ty; p < 32; p += 8) if (p0 + p < HW) d[(int64_t)(p0 + p) * 32 + tx] = tile[tx][p];
  } else {
#pragma unroll
    for (int p = ty; p < 32; p += 8) tile[p][tx] = (p0 + p < HW) ? s[(int64_t)(p0 + p) * 32 + tx] : 0.f;
    __syncthreads();
#pragma unroll
    for (int c = ty; c < 32; c += 8) if (p0 + tx < HW) d[(int64_t)c * HW + p0 + tx] = tile[tx][c];
  }
}

}  // namespace

extern "C" int nfi_planes_to_cl(const float* planes_nchw, int B, int Hp, int Wp, float* planes_cl,
                                nfi_stream_t stream) {
  NFI_REQUIRE(planes_nchw && planes_cl && B > 0 && Hp > 0 && Wp > 0, NFI_ERR_BAD_ARG, "nfi_planes_to_cl: bad argument");
  const int HW = Hp * Wp;
  plane_transpose_kernel<true><<<dim3((HW + 31) / 32, B * 3), 256, 0, nfi_cu(stream)>>>(planes_nchw, planes_cl, HW);
  NFI_CHECK_LAUNCH("plane_transpose_kernel<to_cl>");
  return NFI_OK;
}

extern "C" int nfi_planes_grad_from_cl(const float* g_planes_cl, int B, int Hp, int Wp,
                                       float* g_planes_nchw, nfi_stream_t stream) {
  NFI_REQUIRE(g_planes_cl && g_planes_nchw && B > 0 && Hp > 0 && Wp > 0, NFI_ERR_BAD_ARG, "nfi_planes_grad_from_cl: bad argument");
  const int HW = Hp * Wp;
  plane_transpose_kernel<false><<<dim3((HW + 31) / 32, B * 3), 256, 0, nfi_cu(stream)>>>(g_planes_cl, g_planes_nchw, HW);
  NFI_CHECK_LAUNCH("plane_transpose_kernel<from_cl>");
  return NFI_OK;
}

extern "C" int nfi_field_fwd(const NfiField* field, const float* planes_cl, const float* palette, int B,
                             int Hp, int Wp, const float* origins, const float* dirs, const float* z,
                             const float* points, int64_t n_per_image, int S, int mode, int img_w,
                             int skip_oob, float* out, float* probs_out, nfi_stream_t stream) {
  nfi_field::Args a;
  int rc = fill_args(a, "nfi_field_fwd", field, planes_cl, palette, B, Hp, Wp, origins, dirs, z, points,
                     n_per_image, S, mode, img_w);
  if (rc) return rc;
  NFI_REQUIRE(out, NFI_ERR_BAD_ARG, "nfi_field_fwd: NULL output");
  NFI_REQUIRE(mode == 1 || nfi_aligned16(out), NFI_ERR_ALIGNMENT, "nfi_field_fwd: out must be 16-byte aligned");
  a.out = out; a.probs = probs_out; a.skip_oob = skip_oob;
  const int grid = field_grid(a.n_tiles, 4);
  cudaStream_t st = nfi_cu(stream);
  if (points) {
    if (mode == 0) nfi_field::field_fwd_kernel<true, 0><<<grid, nfi_field::kTile, 0, st>>>(a);
    else nfi_field::field_fwd_kernel<true, 1><<<grid, nfi_field::kTile, 0, st>>>(a);
  } else {
    if (mode == 0) nfi_field::field_fwd_kernel<false, 0><<<grid, nfi_field::kTile, 0, st>>>(a);
    else nfi_field::field_fwd_kernel<false, 1><<<grid, nfi_field::kTile, 0, st>>>(a);
  }
  NFI_CHECK_LAUNCH("field_fwd_kernel");
  return NFI_OK;
}

