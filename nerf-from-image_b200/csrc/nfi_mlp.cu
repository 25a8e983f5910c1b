// Standalone decoder MLP (models/generator.py:295-299, 329-330) on tcgen05: features [n,32] ->
// y [n, n_out].  Used to validate / profile the tensor-core path in isolation; the fused field
// kernels run the same tile routines (nfi_mlp_tc.cuh) straight after their gather.
#include "nfi_common.cuh"
#include "nfi_mlp_tc.cuh"

namespace {

using namespace nfi_mlp;

__global__ void __launch_bounds__(kTile)
decoder_mlp_fwd_kernel(const float* __restrict__ W0, const float* __restrict__ b0,
                       const float* __restrict__ W1, const float* __restrict__ b1, int n_out,
                       const float* __restrict__ feat, int64_t n, float* __restrict__ y_out,
                       float* __restrict__ hpre_out) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  Operands& s = *reinterpret_cast<Operands*>(smem_raw);
  const int tid = threadIdx.x, warp = tid >> 5;
  load_weights(s, W0, b0, W1, b1, n_out);
  if (warp == 0) nfi_tc::tmem_alloc(&s.tmem_base, kTmemCols);
  if (tid == 0) { nfi_tc::mbar_init(&s.bar, 1); nfi_tc::mbar_fence_init(); }
  nfi_tc::fence_proxy_async();          // weight tiles -> async proxy
  nfi_tc::fence_before_sync();
  __syncthreads();
  nfi_tc::fence_after_sync();
  const uint32_t tmem = s.tmem_base;
  const uint32_t tmem_lane = tmem + ((uint32_t)(warp * 32) << 16);
  uint32_t parity = 0;
  const int64_t n_tiles = (n + kTile - 1) / kTile;
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    // features -> operand tile (8 lanes per row: one 128-byte row per quarter warp)
#pragma unroll 4
    for (int it = 0; it < 8; ++it) {
      const int idx = it * kTile + tid;
      const int row = idx >> 3, c4 = idx & 7;
      const int64_t g = tile * kTile + row;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (g < n) v = __ldg(reinterpret_cast<const float4*>(feat) + g * 8 + c4);
      store_feat4(s, row, c4, v);
    }
    nfi_tc::fence_proxy_async();
    nfi_tc::fence_before_sync();
    __syncthreads();
    if (tid == 0) {
      nfi_tc::fence_after_sync();
      issue_layer0(s, tmem);
      nfi_tc::mma_commit(&s.bar);
    }
    nfi_tc::mbar_wait(&s.bar, parity); parity ^= 1;
    nfi_tc::fence_after_sync();
    float hpre[kH];
    epilogue0<4, true>(s, tmem_lane, 0, hpre);
    nfi_tc::fence_before_sync();
    __syncthreads();
    if (tid == 0) {
      nfi_tc::fence_after_sync();
      issue_layer1(s, tmem);
      nfi_tc::mma_commit(&s.bar);
    }
    nfi_tc::mbar_wait(&s.bar, parity); parity ^= 1;
    nfi_tc::fence_after_sync();
    float y[kY];
    epilogue1(s, tmem_lane, y);
    const int64_t g = tile * kTile + tid;
    if (g < n) {
#pragma unroll
      for (int o = 0; o < kY; ++o) if (o < n_out) y_out[g * n_out + o] = y[o];
      if (hpre_out) {
#pragma unroll
        for (int j = 0; j < kH; ++j) hpre_out[g * kH + j] = hpre[j];
      }
    }
    nfi_tc::fence_before_sync();       // TMEM reads of this tile ordered before the next tile's MMAs
  }
  __syncthreads();
  if (warp == 0) nfi_tc::tmem_dealloc(tmem, kTmemCols);
}

}  // namespace

extern "C" int nfi_decoder_mlp_fwd(const NfiField* field, const float* feat, int64_t n, float* y,
                                   float* hpre, nfi_stream_t stream) {
  NFI_REQUIRE(field && field->W0 && field->b0 && field->W1 && field->b1 && feat && y, NFI_ERR_BAD_ARG,
              "nfi_decoder_mlp_fwd: NULL pointer");
  NFI_REQUIRE(n > 0 && field->n_out >= 1 && field->n_out <= NFI_MAX_OUT, NFI_ERR_BAD_ARG, "nfi_decoder_mlp_fwd: bad sizes");
  NFI_REQUIRE(nfi_aligned16(feat), NFI_ERR_ALIGNMENT, "nfi_decoder_mlp_fwd: feat must be 16-byte aligned");
  const int smem = (int)sizeof(Operands);
  const int rc = nfi_set_smem_once((const void*)decoder_mlp_fwd_kernel, smem, "nfi_decoder_mlp_fwd");
  if (rc) return rc;
  const int64_t n_tiles = (n + kTile - 1) / kTile;
  const int64_t cap = (int64_t)nfi_sm_count() * 2;
  const int grid = (int)(n_tiles < cap ? n_tiles : cap);
  decoder_mlp_fwd_kernel<<<grid, kTile, smem, nfi_cu(stream)>>>(field->W0, field->b0, field->W1, field->b1,
                                                              field->n_out, feat, n, y, hpre);
  NFI_CHECK_LAUNCH("decoder_mlp_fwd_kernel");
  return NFI_OK;
}
