// Developer probe: can tcgen05.mma kind::tf32 read SAMPLE-major tiles as MN-major
// operands, so that the weight gradients  dW0 = dHpre^T F  and  dW1^T = H^T dY  (K = the 128 samples of
// a tile) run on the tensor cores without transposing anything?
//
// The field kernels keep a [128 samples x C] fp32 tile in the canonical K-major no-swizzle layout
//     addr(s, c) = (c / 4) * LBO_A + (s / 8) * 128 + (s % 8) * 16 + (c % 4) * 4
// (nfi_tc.cuh).  Read as an MN-major operand with MN = c and K = s, the canonical no-swizzle layout is
//     addr(mn, k) = (mn % 4) * 4 + (mn / 4) * SBO' + (k % 8) * 16 + (k / 8) * LBO'
// -- the same bytes with SBO' = LBO_A and LBO' = 128.  This program checks that claim and its alternatives on
// the hardware: D[64 x 32] = X^T Y over 128 samples with X [128 x 64], Y [128 x 32] stored sample-major,
// small integer entries (exact in tf32), compared with the host result.
//
// Result on B200 (gpurun, round 1): with the MN-major bits set, the no-swizzle layout (variants 0, 1, 5-7) and
// the 128B-swizzled one (8, 9) leave the accumulator all zero -- not usable.  What works (variant 10, exact):
// plain row-major sample rows of 128 bytes (32 floats; wider tiles as panels of 32 columns), 32-byte chunks
// XOR-swizzled by (row % 4), descriptor layout type 1 (SWIZZLE_128B_BASE32B), LBO = panel stride, SBO = 512
// (a block of 4 rows), start address + 1024 per MMA (K = 8 samples).  M = 64 results sit on lanes
// 32 * (row / 16) + row % 16 (variant 4 is the K-major control for that).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../../nerf-from-image_b200/csrc -o umma_mn_test umma_mn_test.cu && ./umma_mn_test
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "nfi_tc.cuh"

namespace {
constexpr int kS = 128, kJ = 64, kF = 32;
constexpr uint32_t kLboX = kS * 16, kLboY = kS * 16;      // dense: a 16-byte column chunk of all 128 rows

__device__ __host__ inline uint32_t off(int s, int c, uint32_t lbo) {
  return (c / 4) * lbo + (s / 8) * 128 + (s % 8) * 16 + (c % 4) * 4;
}

// kind::tf32, fp32 accumulate, A and B MN-major (bits 15 / 16), M x N tile
__device__ constexpr uint32_t idesc_tf32_mn(int M, int N) {
  return nfi_tc::idesc_tf32(M, N) | (1u << 15) | (1u << 16);
}

struct Smem {
  alignas(128) float x[kS * kJ];
  alignas(128) float y[kS * kF];
  alignas(128) float xt[kJ * kS];     // X^T as a K-major [64 rows j x K = 128 samples] operand (control)
  alignas(128) float yt[kF * kS];     // Y^T as a K-major [32 rows f x K = 128 samples] operand (control)
  alignas(1024) float xs[kS * kJ];    // X row-major, 128-byte rows in two panels of 32 columns, 128B-swizzled
  alignas(1024) float ys[kS * kF];    // Y likewise (one panel)
  alignas(8) uint64_t bar;
  uint32_t tmem_base;
};

// [128 samples x 32 columns] panel: sample row = 128 bytes, 16-byte chunks XOR-swizzled by (row % 8)
__device__ inline uint32_t off_sw(int s, int c) {
  return (c / 32) * (kS * 128) + s * 128 + ((((c % 32) / 4) ^ (s % 8)) * 16) + (c % 4) * 4;
}
// same rows, 32-byte chunks XOR-swizzled by (row % 4): Swizzle<2,5,2>, layout type SWIZZLE_128B_BASE32B
__device__ inline uint32_t off_sw32(int s, int c) {
  return (c / 32) * (kS * 128) + s * 128 + ((((c % 32) / 8) ^ (s % 4)) * 32) + (c % 8) * 4;
}
__device__ inline uint64_t smem_desc_type(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes, int type) {
  return nfi_tc::smem_desc(saddr, lbo_bytes, sbo_bytes) | (static_cast<uint64_t>(type) << 61);
}
__device__ inline uint64_t smem_desc_sw128(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return nfi_tc::smem_desc(saddr, lbo_bytes, sbo_bytes) | (static_cast<uint64_t>(2) << 61);
}

__global__ void __launch_bounds__(128, 1) umma_mn_kernel(const float* X, const float* Y, float* out, int variant) {
  extern __shared__ __align__(1024) unsigned char raw[];
  Smem& sm = *reinterpret_cast<Smem*>(raw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < kS * kJ; i += 128) sm.x[off(i / kJ, i % kJ, kLboX) / 4] = X[i];
  for (int i = tid; i < kS * kF; i += 128) sm.y[off(i / kF, i % kF, kLboY) / 4] = Y[i];
  for (int i = tid; i < kS * kJ; i += 128) sm.xt[off(i % kJ, i / kJ, kJ * 16) / 4] = X[i];    // row j, column s
  for (int i = tid; i < kS * kF; i += 128) sm.yt[off(i % kF, i / kF, kF * 16) / 4] = Y[i];
  for (int i = tid; i < kS * kJ; i += 128) sm.xs[(variant >= 10 ? off_sw32(i / kJ, i % kJ) : off_sw(i / kJ, i % kJ)) / 4] = X[i];
  for (int i = tid; i < kS * kF; i += 128) sm.ys[(variant >= 10 ? off_sw32(i / kF, i % kF) : off_sw(i / kF, i % kF)) / 4] = Y[i];
  if (warp == 0) nfi_tc::tmem_alloc(&sm.tmem_base, 32);
  if (tid == 0) { nfi_tc::mbar_init(&sm.bar, 1); nfi_tc::mbar_fence_init(); }
  nfi_tc::fence_proxy_async();
  nfi_tc::fence_before_sync();
  __syncthreads();
  nfi_tc::fence_after_sync();
  const uint32_t tmem = sm.tmem_base;
  {   // clear the accumulator block so that untouched lanes read as zero
    uint32_t z[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) z[i] = 0u;
    const uint32_t tl = tmem + ((uint32_t)(warp * 32) << 16);
    nfi_tc::tmem_st16(tl, z);
    nfi_tc::tmem_st16(tl + 16, z);
    nfi_tc::tmem_st_wait();
  }
  nfi_tc::fence_before_sync();
  __syncthreads();
  if (tid == 0) {
    nfi_tc::fence_after_sync();
    const uint32_t xa = nfi_tc::smem_u32(sm.x), ya = nfi_tc::smem_u32(sm.y);
    // variants 0, 1, 4..7: D[64 x 32] = X^T Y over K = 128 samples, one MMA per block of 8 samples.
    //   bit 0 (variants 0/1 only): reading of LBO / SBO for MN-major operands
    //   variants 4..7: A MN-major iff (variant & 1), B MN-major iff (variant & 2); the other operand is the
    //   explicitly transposed K-major copy
    const bool mnA = variant < 2 || (variant >= 4 && (variant & 1)), mnB = variant < 2 || (variant >= 4 && (variant & 2));
    const uint32_t xta = nfi_tc::smem_u32(sm.xt), yta = nfi_tc::smem_u32(sm.yt);
    for (int kb = 0; kb < kS / 8 && variant != 2 && variant < 8; ++kb) {
      uint64_t da, db;
      if (mnA) da = variant == 1 ? nfi_tc::smem_desc(xa + kb * 128, kLboX, 128) : nfi_tc::smem_desc(xa + kb * 128, 128, kLboX);
      else     da = nfi_tc::smem_desc(xta + kb * 2 * (kJ * 16), kJ * 16, 128);
      if (mnB) db = variant == 1 ? nfi_tc::smem_desc(ya + kb * 128, kLboY, 128) : nfi_tc::smem_desc(ya + kb * 128, 128, kLboY);
      else     db = nfi_tc::smem_desc(yta + kb * 2 * (kF * 16), kF * 16, 128);
      const uint32_t idesc = nfi_tc::idesc_tf32(kJ, kF) | (mnA ? 1u << 15 : 0u) | (mnB ? 1u << 16 : 0u);
      nfi_tc::mma_tf32_ss(tmem, da, db, idesc, kb > 0 ? 1u : 0u);
    }
    if (variant >= 10) {         // SWIZZLE_128B_BASE32B (descriptor layout type 1): K blocks of 4 rows
      const uint32_t xsa = nfi_tc::smem_u32(sm.xs), ysa = nfi_tc::smem_u32(sm.ys);
      for (int kb = 0; kb < kS / 8; ++kb) {
        const uint32_t panel = kS * 128, kblock = 512;
        const uint64_t da = variant == 10 ? smem_desc_type(xsa + kb * 1024, panel, kblock, 1)
                                          : smem_desc_type(xsa + kb * 1024, kblock, panel, 1);
        const uint64_t db = variant == 10 ? smem_desc_type(ysa + kb * 1024, panel, kblock, 1)
                                          : smem_desc_type(ysa + kb * 1024, kblock, panel, 1);
        nfi_tc::mma_tf32_ss(tmem, da, db, nfi_tc::idesc_tf32(kJ, kF) | (1u << 15) | (1u << 16), kb > 0 ? 1u : 0u);
      }
    } else
    if (variant >= 8) {          // 128B-swizzled row-major tiles read as MN-major operands
      const uint32_t xsa = nfi_tc::smem_u32(sm.xs), ysa = nfi_tc::smem_u32(sm.ys);
      for (int kb = 0; kb < kS / 8; ++kb) {
        const uint32_t panel = kS * 128, kblock = 1024;
        const uint64_t da = variant == 8 ? smem_desc_sw128(xsa + kb * kblock, panel, kblock)
                                         : smem_desc_sw128(xsa + kb * kblock, kblock, panel);
        const uint64_t db = variant == 8 ? smem_desc_sw128(ysa + kb * kblock, panel, kblock)
                                         : smem_desc_sw128(ysa + kb * kblock, kblock, panel);
        nfi_tc::mma_tf32_ss(tmem, da, db, nfi_tc::idesc_tf32(kJ, kF) | (1u << 15) | (1u << 16), kb > 0 ? 1u : 0u);
      }
    }
    if (variant == 2) {          // harness control, K-major as in production: D[s][n] = sum_f Y[s][f] Y[n][f], n < 32
      for (int k = 0; k < kF / 8; ++k) {
        const uint64_t da = nfi_tc::smem_desc(ya + k * 2 * kLboY, kLboY, 128);
        nfi_tc::mma_tf32_ss(tmem, da, da, nfi_tc::idesc_tf32(128, 32), k > 0 ? 1u : 0u);
      }
    }
    nfi_tc::mma_commit(&sm.bar);
  }
  nfi_tc::mbar_wait(&sm.bar, 0);
  nfi_tc::fence_after_sync();
  float v[16];
  const uint32_t tl = tmem + ((uint32_t)(warp * 32) << 16);
  for (int h = 0; h < 2; ++h) {
    nfi_tc::tmem_ld16(tl + 16 * h, v);
    nfi_tc::tmem_ld_wait();
    for (int i = 0; i < 16; ++i) out[(warp * 32 + lane) * 32 + 16 * h + i] = v[i];
  }
  nfi_tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) nfi_tc::tmem_dealloc(tmem, 32);
}
}  // namespace

int main() {
  std::vector<float> X(kS * kJ), Y(kS * kF), ref(kJ * kF, 0.f), out(128 * 32);
  srand(7);
  for (auto& v : X) v = (float)(rand() % 9 - 4);
  for (auto& v : Y) v = (float)(rand() % 9 - 4);
  for (int s = 0; s < kS; ++s)
    for (int j = 0; j < kJ; ++j)
      for (int f = 0; f < kF; ++f) ref[j * kF + f] += X[s * kJ + j] * Y[s * kF + f];
  float *dX, *dY, *dO;
  cudaMalloc(&dX, X.size() * 4); cudaMalloc(&dY, Y.size() * 4); cudaMalloc(&dO, out.size() * 4);
  cudaMemcpy(dX, X.data(), X.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dY, Y.data(), Y.size() * 4, cudaMemcpyHostToDevice);
  cudaFuncSetAttribute(umma_mn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
  for (int variant = 0; variant < 11; ++variant) {      // (variant 11, LBO / SBO swapped for type 1, faults)
    if (variant == 3) continue;
    cudaMemset(dO, 0, out.size() * 4);
    umma_mn_kernel<<<1, 128, sizeof(Smem)>>>(dX, dY, dO, variant);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("variant %d: %s\n", variant, cudaGetErrorString(e)); return 1; }
    cudaMemcpy(out.data(), dO, out.size() * 4, cudaMemcpyDeviceToHost);
    if (variant == 2) {
      int bad = 0;
      for (int r = 0; r < kS; ++r)
        for (int n = 0; n < 32; ++n) {
          float want = 0.f;
          for (int f = 0; f < kF; ++f) want += Y[r * kF + f] * Y[n * kF + f];
          bad += out[r * 32 + n] != want;
        }
      printf("variant 2 (control, K-major M128 N32 K32): mismatches %d of %d\n", bad, kS * 32);
      continue;
    }
    // where may row j of the M = 64 accumulator sit?  (a) lane j, (b) lane (j / 16) * 32 + j % 16
    int bad_a = 0, bad_b = 0, nz = 0;
    for (int j = 0; j < kJ; ++j)
      for (int f = 0; f < kF; ++f) {
        bad_a += out[j * 32 + f] != ref[j * kF + f];
        bad_b += out[((j / 16) * 32 + j % 16) * 32 + f] != ref[j * kF + f];
      }
    for (float v : out) nz += v != 0.f;
    printf("variant %d (%s): mismatches with rows on lanes 0-63: %d, on lanes 16 of every 32: %d (of %d); "
           "non-zero accumulator cells: %d\n", variant,
           variant == 0 ? "A, B MN-major; LBO = K-block stride, SBO = MN-chunk stride" :
           variant == 1 ? "A, B MN-major; LBO / SBO swapped" :
           variant == 4 ? "A, B K-major transposed copies (M = 64 control)" :
           variant == 5 ? "A MN-major, B K-major copy" :
           variant == 6 ? "A K-major copy, B MN-major" :
           variant == 8 ? "A, B MN-major from 128B-swizzled row-major tiles; LBO = panel, SBO = 8-row block" :
           variant == 9 ? "same, LBO / SBO swapped" :
           variant == 10 ? "A, B MN-major, SWIZZLE_128B_BASE32B tiles; LBO = panel, SBO = 4-row block" :
           variant == 11 ? "same, LBO / SBO swapped" : "A, B MN-major (as variant 0)",
           bad_a, bad_b, kJ * kF, nz);
  }
  return 0;
}
