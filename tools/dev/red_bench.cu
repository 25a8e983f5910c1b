// Developer microbenchmark: how fast can one SM push 128-byte fp32 reduction lines into L2?
//   A  red.global.add.v4.f32, 8 lanes per line (what field_bwd_ws_kernel's scatter does)
//   B  cp.reduce.async.bulk.global.shared::cta.add.f32, one 128-byte bulk op per thread (TMA engine)
// Random line addresses in a 403 MB buffer (the plane-gradient tensor of config 2).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o red_bench red_bench.cu && ./red_bench
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

__device__ __forceinline__ uint32_t lcg(uint32_t& s) { s = s * 1664525u + 1013904223u; return s >> 8; }

__global__ void __launch_bounds__(128) red_v4_kernel(float* buf, uint32_t nlines, int iters) {
  const int lane = threadIdx.x & 31, c4 = lane & 7;
  uint32_t s = (blockIdx.x * 128 + (threadIdx.x & ~7)) * 2654435761u + 12345u;     // same stream for the 8 lanes of a line
  for (int i = 0; i < iters; ++i) {
    const uint32_t line = lcg(s) % nlines;
    float* p = buf + (size_t)line * 32 + c4 * 4;
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(1.f), "f"(2.f), "f"(3.f), "f"(4.f) : "memory");
  }
}

template <int DEPTH>
__global__ void __launch_bounds__(128) red_bulk_kernel(float* buf, uint32_t nlines, int iters) {
  __shared__ __align__(128) float stage[128 * 36];          // 144-byte stride: conflict-free float4 stores
  float* mine = stage + threadIdx.x * 36;
  for (int i = 0; i < 32; ++i) mine[i] = (float)(i & 3) + 1.f;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  uint32_t s = (blockIdx.x * 128 + threadIdx.x) * 2654435761u + 12345u;
  const uint32_t src = (uint32_t)__cvta_generic_to_shared(mine);
  for (int i = 0; i < iters; ++i) {
    const uint32_t line = lcg(s) % nlines;
    float* p = buf + (size_t)line * 32;
    asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], 128;" ::"l"(p), "r"(src) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(DEPTH) : "memory");
  }
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

template <typename F>
float time_ms(F f) {
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  f(); cudaDeviceSynchronize();
  cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms = 0; cudaEventElapsedTime(&ms, a, b);
  return ms;
}

int run(uint32_t nlines);
int main() {
  printf("== 403 MB of lines (all images: DRAM-resident)\n");
  int rc = run(16u * 3u * 65536u);
  printf("== 25 MB of lines (one image's planes: L2-resident, what the kernel's image-major schedule gives)\n");
  return rc | run(3u * 65536u);
}
int run(uint32_t nlines) {
  float* buf; cudaMalloc(&buf, (size_t)nlines * 128); cudaMemset(buf, 0, (size_t)nlines * 128);
  int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  const int iters = 4096;
  for (int cta_per_sm = 1; cta_per_sm <= 2; cta_per_sm *= 2) {
    const int grid = sms * cta_per_sm;
    // A: 128 threads = 16 lines per iteration per CTA
    float ms = time_ms([&] { red_v4_kernel<<<grid, 128>>>(buf, nlines, iters); });
    double lines = (double)grid * 16 * iters;
    printf("red.v4   ctas/SM %d: %.3f ms  %.2f Glines/s  %.1f cycles per line per SM (at %d MHz)  %.0f GB/s\n", cta_per_sm, ms,
           lines / ms / 1e6, ms * 1e-3 * khz * 1e3 / (lines / sms), khz / 1000, lines * 128 / ms / 1e6);
    ms = time_ms([&] { red_bulk_kernel<8><<<grid, 128>>>(buf, nlines, iters / 8); });
    lines = (double)grid * 128 * (iters / 8);
    printf("bulk d=8 ctas/SM %d: %.3f ms  %.2f Glines/s  %.1f cycles per line per SM  %.0f GB/s\n", cta_per_sm, ms,
           lines / ms / 1e6, ms * 1e-3 * khz * 1e3 / (lines / sms), lines * 128 / ms / 1e6);
    ms = time_ms([&] { red_bulk_kernel<0><<<grid, 128>>>(buf, nlines, iters / 8); });
    printf("bulk d=0 ctas/SM %d: %.3f ms  %.2f Glines/s  %.1f cycles per line per SM\n", cta_per_sm, ms,
           lines / ms / 1e6, ms * 1e-3 * khz * 1e3 / (lines / sms));
  }
  cudaError_t e = cudaGetLastError();
  printf("status: %s\n", cudaGetErrorString(e));
  cudaFree(buf);
  return e != cudaSuccess;
}
