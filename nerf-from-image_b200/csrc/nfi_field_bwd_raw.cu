// field_bwd_kernel instantiations, mode 1 (raw decoder output).
#include "nfi_field_bwd_kernels.cuh"

void nfi_field_bwd_launch_raw(const nfi_field::Args& a, int grid, cudaStream_t st) {
  using namespace nfi_field;
  if (a.points) {
    if (a.g_points) field_bwd_kernel<true, 1, true><<<grid, kTile, 0, st>>>(a);
    else field_bwd_kernel<true, 1, false><<<grid, kTile, 0, st>>>(a);
  } else {
    if (a.g_points) field_bwd_kernel<false, 1, true><<<grid, kTile, 0, st>>>(a);
    else field_bwd_kernel<false, 1, false><<<grid, kTile, 0, st>>>(a);
  }
}
