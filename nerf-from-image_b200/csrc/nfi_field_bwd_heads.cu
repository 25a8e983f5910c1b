// field_bwd_kernel instantiations, mode 0 (sigma/rgb heads).
#include "nfi_field_bwd_kernels.cuh"

void nfi_field_bwd_launch_heads(const nfi_field::Args& a, int grid, cudaStream_t st) {
  using namespace nfi_field;
  if (a.points) {
    if (a.g_points) field_bwd_kernel<true, 0, true><<<grid, kTile, 0, st>>>(a);
    else field_bwd_kernel<true, 0, false><<<grid, kTile, 0, st>>>(a);
  } else {
    if (a.g_points) field_bwd_kernel<false, 0, true><<<grid, kTile, 0, st>>>(a);
    else field_bwd_kernel<false, 0, false><<<grid, kTile, 0, st>>>(a);
  }
}
