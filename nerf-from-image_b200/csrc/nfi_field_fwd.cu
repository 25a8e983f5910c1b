// Forward of the fused triplane field (gather + tcgen05 MLP + heads), see nfi_field_tc.cuh.
#include <stdlib.h>

#include "nfi_field_fwd_ws.cuh"
#include "nfi_field_tc_host.cuh"

namespace nfi_ftc {

struct FwdSmem {
  nfi_mlp::Operands op;
  Tables t;
};

template <bool PTS, int MODE>
__global__ void __launch_bounds__(kThreads, 2)
field_fwd_tc_kernel(const Args a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  FwdSmem& sm = *reinterpret_cast<FwdSmem*>(smem_raw);
  nfi_mlp::Operands& op = sm.op;
  Tables& t = sm.t;
  const int tid = threadIdx.x, warp = tid >> 5;
  nfi_mlp::load_weights(op, a.f.W0, a.f.b0, a.f.W1, a.f.b1, a.f.n_out);
  if (warp == 0) nfi_tc::tmem_alloc(&op.tmem_base, nfi_mlp::kTmemCols);
  if (tid == 0) { nfi_tc::mbar_init(&op.bar, 1); nfi_tc::mbar_fence_init(); }
  nfi_tc::fence_proxy_async();
  nfi_tc::fence_before_sync();
  __syncthreads();
  nfi_tc::fence_after_sync();
  const uint32_t tmem = op.tmem_base;
  const uint32_t tmem_lane = tmem + ((uint32_t)((warp & 3) * 32) << 16);
  const NfiHeads hd = make_heads(a.f);
  uint32_t parity = 0;
  const int64_t per_image = a.npi * a.S;
  for (int b = 0; b < a.B; ++b) {
    int64_t lo, hi;
    cta_tile_range(a.tiles_per_image, lo, hi);
    for (int64_t tt = lo; tt < hi; ++tt) {
      int64_t id = -1;
      if (tid < kTile) {
        const int sid = sample_in_image(a, (int)tt, tid);
        if (sid >= 0) id = (int64_t)b * per_image + sid;
        phase1<PTS, MODE, false>(a, t, id, sid, tid);
      }
      const bool mine = tid < kTile && id >= 0 && sample_active(a, t.keep[tid]);
      const int any = __syncthreads_or(mine ? 1 : 0);
      if (!any) {                                  // whole tile outside the box: zeros, no gather, no MMA
        if (MODE == 0 && id >= 0) {
          reinterpret_cast<float4*>(a.out)[id] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (a.probs) for (int q = 0; q < hd.A; ++q) a.probs[id * hd.A + q] = 0.f;
        }
        __syncthreads();                           // tables are rewritten by the next tile
        continue;
      }
      phase2_gather(a, t, op, b);
      nfi_tc::fence_proxy_async();
      nfi_tc::fence_before_sync();
      __syncthreads();
      if (tid == 0) {
        nfi_tc::fence_after_sync();
        nfi_mlp::issue_layer0(op, tmem);
        nfi_tc::mma_commit(&op.bar);
      }
      nfi_tc::mbar_wait(&op.bar, parity); parity ^= 1;
      nfi_tc::fence_after_sync();
      nfi_mlp::epilogue0<2, false>(op, tmem_lane, (warp >> 2) * 32, nullptr);
      nfi_tc::fence_before_sync();
      __syncthreads();
      if (tid == 0) {
        nfi_tc::fence_after_sync();
        nfi_mlp::issue_layer1(op, tmem);
        nfi_tc::mma_commit(&op.bar);
      }
      if (warp < 4) {
        nfi_tc::mbar_wait(&op.bar, parity);
        nfi_tc::fence_after_sync();
        float y[nfi_mlp::kY];
        nfi_mlp::epilogue1(op, tmem_lane, y);
        if (id >= 0) {
          const float keep = t.keep[tid];
          if (MODE == 0) {
            if (!sample_active(a, keep)) {
              reinterpret_cast<float4*>(a.out)[id] = make_float4(0.f, 0.f, 0.f, 0.f);
              if (a.probs) for (int q = 0; q < hd.A; ++q) a.probs[id * hd.A + q] = 0.f;
            } else {
              float sigma, rgb[3], probs[NFI_YPAD - 1];
              nfi_heads_fwd(hd, y, keep, a.palette ? a.palette + (int64_t)b * hd.A * 3 : nullptr, sigma, rgb, probs);
              reinterpret_cast<float4*>(a.out)[id] = make_float4(sigma, rgb[0], rgb[1], rgb[2]);
              if (a.probs) {
#pragma unroll
                for (int q = 0; q < NFI_YPAD - 1; ++q) if (q < hd.A) a.probs[id * hd.A + q] = probs[q];
              }
            }
          } else {
#pragma unroll
            for (int o = 0; o < NFI_YPAD; ++o) if (o < a.f.n_out) a.out[id * a.f.n_out + o] = y[o];
          }
        }
        nfi_tc::fence_before_sync();
      }
      parity ^= 1;
    }
  }
  __syncthreads();
  if (warp == 0) nfi_tc::tmem_dealloc(tmem, nfi_mlp::kTmemCols);
}

}  // namespace nfi_ftc

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

template <typename K>
int set_smem(K kernel, int bytes, const char* who) {
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  NFI_REQUIRE(e == cudaSuccess, NFI_ERR_LAUNCH, "%s: cudaFuncSetAttribute: %s", who, cudaGetErrorString(e));
  return NFI_OK;
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
  using namespace nfi_ftc;
  Args a;
  int rc = fill_args(a, "nfi_field_fwd", field, planes_cl, palette, B, Hp, Wp, origins, dirs, z, points,
                     n_per_image, S, mode, img_w);
  if (rc) return rc;
  NFI_REQUIRE(out, NFI_ERR_BAD_ARG, "nfi_field_fwd: NULL output");
  NFI_REQUIRE(mode == 1 || nfi_aligned16(out), NFI_ERR_ALIGNMENT, "nfi_field_fwd: out must be 16-byte aligned");
  a.out = out; a.probs = probs_out; a.skip_oob = skip_oob;
  cudaStream_t st = nfi_cu(stream);
  static const bool simple = getenv("NFI_FWD_SIMPLE") != nullptr;     // development A/B switch
  if (!simple) {
    const int grid = field_grid(a, 1);
    const int smem = (int)sizeof(WsFwdSmem);
#define NFI_LAUNCH_WS(P, M)                                                               \
  do {                                                                                    \
    rc = set_smem(field_fwd_ws_kernel<P, M>, smem, "nfi_field_fwd");                      \
    if (rc) return rc;                                                                    \
    field_fwd_ws_kernel<P, M><<<grid, kWsThreads, smem, st>>>(a);                         \
  } while (0)
    if (points) { if (mode == 0) NFI_LAUNCH_WS(true, 0); else NFI_LAUNCH_WS(true, 1); }
    else        { if (mode == 0) NFI_LAUNCH_WS(false, 0); else NFI_LAUNCH_WS(false, 1); }
#undef NFI_LAUNCH_WS
    NFI_CHECK_LAUNCH("field_fwd_ws_kernel");
    return NFI_OK;
  }
  const int grid = field_grid(a, 2);
  const int smem = (int)sizeof(FwdSmem);
#define NFI_LAUNCH_FWD(P, M)                                                              \
  do {                                                                                    \
    rc = set_smem(field_fwd_tc_kernel<P, M>, smem, "nfi_field_fwd");                      \
    if (rc) return rc;                                                                    \
    field_fwd_tc_kernel<P, M><<<grid, kThreads, smem, st>>>(a);                           \
  } while (0)
  if (points) { if (mode == 0) NFI_LAUNCH_FWD(true, 0); else NFI_LAUNCH_FWD(true, 1); }
  else        { if (mode == 0) NFI_LAUNCH_FWD(false, 0); else NFI_LAUNCH_FWD(false, 1); }
#undef NFI_LAUNCH_FWD
  NFI_CHECK_LAUNCH("field_fwd_tc_kernel");
  return NFI_OK;
}
