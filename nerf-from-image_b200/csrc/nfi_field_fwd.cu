// Forward of the fused triplane field (gather + tcgen05 MLP + heads), see nfi_field_tc.cuh.
#include <stdio.h>
#include <stdlib.h>

#include "nfi_field_fwd_ws.cuh"
#include "nfi_field_tc_host.cuh"

namespace {

// [B*3][32][HW] -> [B*3][HW][32] (and back), the channels-last side in the texel order of
// nfi_field_math.cuh.  32 x 32 tiles through padded shared memory: the NCHW side is 128-byte coalesced,
// the channels-last side moves whole 128-byte texels.
__device__ __forceinline__ int texel_index(int p, int Wp, int bshift) {
  const int y = p / Wp, x = p - y * Wp;
  return nfi_texel_row(y, bshift, Wp >> bshift) + nfi_texel_col(x, bshift);
}

template <bool TO_CL>
__global__ void __launch_bounds__(256)
plane_transpose_kernel(const float* __restrict__ src, float* __restrict__ dst, int HW, int Wp, int bshift) {
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
    for (int p = ty; p < 32; p += 8) if (p0 + p < HW) d[(int64_t)texel_index(p0 + p, Wp, bshift) * 32 + tx] = tile[tx][p];
  } else {
#pragma unroll
    for (int p = ty; p < 32; p += 8) tile[p][tx] = (p0 + p < HW) ? s[(int64_t)texel_index(p0 + p, Wp, bshift) * 32 + tx] : 0.f;
    __syncthreads();
#pragma unroll
    for (int c = ty; c < 32; c += 8) if (p0 + tx < HW) d[(int64_t)c * HW + p0 + tx] = tile[tx][c];
  }
}

// Same transposition with one 16-byte global access per thread on both sides (HW % 128 == 0, Wp % 4 == 0):
// a block moves four 32-texel x 32-channel tiles; NCHW side: thread (c, x4) takes 4 consecutive texels
// of channel c; channels-last side: thread (p, c4) takes 4 consecutive channels of texel p (8 lanes =
// one 128-byte texel).  The padded tile makes all four scalar shared-memory accesses conflict-free.
template <bool TO_CL>
__global__ void __launch_bounds__(256)
plane_transpose4_kernel(const float* __restrict__ src, float* __restrict__ dst, int HW, int Wp, int bshift) {
  __shared__ float tile[32][33];
  const int64_t plane = blockIdx.y;
  const float* s = src + plane * 32 * (int64_t)HW;
  float* d = dst + plane * 32 * (int64_t)HW;
  const int t = threadIdx.x;
  const int c = t >> 3, x4 = (t & 7) * 4;          // NCHW side
  const int p = t >> 3, c4 = (t & 7) * 4;          // channels-last side
#pragma unroll 1
  for (int k = 0; k < 4; ++k) {
    const int p0 = (blockIdx.x * 4 + k) * 32;
    if (TO_CL) {
      const float4 v = *reinterpret_cast<const float4*>(s + (int64_t)c * HW + p0 + x4);
      tile[c][x4] = v.x; tile[c][x4 + 1] = v.y; tile[c][x4 + 2] = v.z; tile[c][x4 + 3] = v.w;
      __syncthreads();
      *reinterpret_cast<float4*>(d + (int64_t)texel_index(p0 + p, Wp, bshift) * 32 + c4) =
          make_float4(tile[c4][p], tile[c4 + 1][p], tile[c4 + 2][p], tile[c4 + 3][p]);
    } else {
      const float4 v = *reinterpret_cast<const float4*>(s + (int64_t)texel_index(p0 + p, Wp, bshift) * 32 + c4);
      tile[c4][p] = v.x; tile[c4 + 1][p] = v.y; tile[c4 + 2][p] = v.z; tile[c4 + 3][p] = v.w;
      __syncthreads();
      *reinterpret_cast<float4*>(d + (int64_t)c * HW + p0 + x4) =
          make_float4(tile[c][x4], tile[c][x4 + 1], tile[c][x4 + 2], tile[c][x4 + 3]);
    }
    __syncthreads();
  }
}

template <bool TO_CL>
void launch_plane_transpose(const float* src, float* dst, int B, int HW, int Wp, int bshift, cudaStream_t st) {
  if (HW % 128 == 0 && Wp % 4 == 0 && nfi_aligned16(src) && nfi_aligned16(dst))
    plane_transpose4_kernel<TO_CL><<<dim3(HW / 128, B * 3), 256, 0, st>>>(src, dst, HW, Wp, bshift);
  else
    plane_transpose_kernel<TO_CL><<<dim3((HW + 31) / 32, B * 3), 256, 0, st>>>(src, dst, HW, Wp, bshift);
}

template <typename K>
int set_smem(K kernel, int bytes, const char* who) {
  return nfi_set_smem_once((const void*)kernel, bytes, who);
}

}  // namespace

extern "C" int nfi_planes_to_cl(const float* planes_nchw, int B, int Hp, int Wp, float* planes_cl,
                                nfi_stream_t stream) {
  NFI_REQUIRE(planes_nchw && planes_cl && B > 0 && Hp > 0 && Wp > 0, NFI_ERR_BAD_ARG, "nfi_planes_to_cl: bad argument");
  const int HW = Hp * Wp;
  launch_plane_transpose<true>(planes_nchw, planes_cl, B, HW, Wp, nfi_plane_bshift(Hp, Wp), nfi_cu(stream));
  NFI_CHECK_LAUNCH("plane_transpose_kernel<to_cl>");
  return NFI_OK;
}

extern "C" int nfi_planes_grad_from_cl(const float* g_planes_cl, int B, int Hp, int Wp,
                                       float* g_planes_nchw, nfi_stream_t stream) {
  NFI_REQUIRE(g_planes_cl && g_planes_nchw && B > 0 && Hp > 0 && Wp > 0, NFI_ERR_BAD_ARG, "nfi_planes_grad_from_cl: bad argument");
  const int HW = Hp * Wp;
  launch_plane_transpose<false>(g_planes_cl, g_planes_nchw, B, HW, Wp, nfi_plane_bshift(Hp, Wp), nfi_cu(stream));
  NFI_CHECK_LAUNCH("plane_transpose_kernel<from_cl>");
  return NFI_OK;
}

extern "C" int nfi_field_fwd(const NfiField* field, const float* planes_cl, const float* palette, int B,
                             int Hp, int Wp, const float* origins, const float* dirs, const float* z,
                             const float* points, int64_t n_per_image, int S, int mode, int img_w,
                             int skip_oob, float* out, float* probs_out, nfi_stream_t stream) {
  using namespace nfi_ftc;
  Args a;
  int rc = fill_args(a, "nfi_field_fwd", field, planes_cl, palette, B, Hp, Wp, origins, dirs, z, points,
                     n_per_image, S, mode, img_w);
  if (rc) return rc;
  NFI_REQUIRE(out, NFI_ERR_BAD_ARG, "nfi_field_fwd: NULL output");
  NFI_REQUIRE(mode == 1 || nfi_aligned16(out), NFI_ERR_ALIGNMENT, "nfi_field_fwd: out must be 16-byte aligned");
  a.out = out; a.probs = probs_out; a.skip_oob = skip_oob;
  cudaStream_t st = nfi_cu(stream);
  {
    const int grid = field_grid_chunks(a);
    const int smem = (int)sizeof(WsFwdSmem);
#define NFI_LAUNCH_WS1(P, M, H)                                                           \
  do {                                                                                    \
    rc = set_smem(field_fwd_ws_kernel<P, M, H>, smem, "nfi_field_fwd");                   \
    if (rc) return rc;                                                                    \
    field_fwd_ws_kernel<P, M, H><<<grid, kWsThreads, smem, st>>>(a);                      \
  } while (0)
#define NFI_LAUNCH_WS(P, M)                                                               \
  do {                                                                                    \
    if (a.f.planes_bf16) NFI_LAUNCH_WS1(P, M, true); else NFI_LAUNCH_WS1(P, M, false);    \
  } while (0)
    if (points) { if (mode == 0) NFI_LAUNCH_WS(true, 0); else NFI_LAUNCH_WS(true, 1); }
    else        { if (mode == 0) NFI_LAUNCH_WS(false, 0); else NFI_LAUNCH_WS(false, 1); }
#undef NFI_LAUNCH_WS
#undef NFI_LAUNCH_WS1
    NFI_CHECK_LAUNCH("field_fwd_ws_kernel");
#ifdef NFI_WS_TRACE
    if (const char* path = getenv("NFI_WS_TRACE_FILE")) {       // developer builds only: blocking dump
      static int launch = 0;
      static long long host[kTrRoles * kTrN * kTrEv];
      cudaStreamSynchronize(st);
      cudaMemcpyFromSymbol(host, g_ws_trace, sizeof(host));
      char name[512];
      snprintf(name, sizeof(name), "%s.%d", path, launch++ % 2);
      if (FILE* fp = fopen(name, "w")) {
        for (int r = 0; r < kTrRoles; ++r)
          for (int t = 0; t < kTrN; ++t) {
            fprintf(fp, "%d %d", r, t);
            for (int e = 0; e < kTrEv; ++e) fprintf(fp, " %lld", host[(r * kTrN + t) * kTrEv + e]);
            fprintf(fp, "\n");
          }
        fclose(fp);
      }
    }
#endif
    return NFI_OK;
  }
}

#ifdef NFI_WAIT_FLAG
// Developer reproducer builds only (see nfi_field_bwd.cu): timed-out bounded waits of THIS translation unit.
extern "C" int nfi_debug_wait_timeouts_fwd(int reset) {
  int v = -1;
  if (cudaDeviceSynchronize() != cudaSuccess) return -2;
  if (cudaMemcpyFromSymbol(&v, nfi_tc::g_nfi_wait_timeouts, sizeof(int)) != cudaSuccess) return -3;
  if (reset) {
    const int zero = 0;
    if (cudaMemcpyToSymbol(nfi_tc::g_nfi_wait_timeouts, &zero, sizeof(int)) != cudaSuccess) return -4;
  }
  return v;
}
#endif
