// Backward of the fused triplane field (see nfi_field_tc.cuh for the tiling).  Nothing but the
// 16-byte upstream gradient per sample is read: the forward (gather, MLP, heads) is recomputed per
// tile, the MLP adjoint runs on tcgen05 (nfi_mlp_tc.cuh), and the feature gradient is scattered
// into the channels-last gradient planes with red.global.add.v4.f32 (one 128-byte line per texel
// and 8 lanes).  Summation order into the planes is not deterministic (fp32 atomics).
#include <stdio.h>
#include <stdlib.h>

#include "nfi_field_bwd_ws.cuh"
#include "nfi_field_tc_host.cuh"

namespace nfi_ftc {

constexpr int kDfLd = 36;            // d feature tile row stride (floats): conflict-free float4 rows
constexpr int kNPal = 3 * (NFI_YPAD - 1);

struct BwdSmem {
  nfi_mlp::Operands op;              // op.a_hi / a_lo double as the d feature tile after layer 0
  nfi_mlp::BwdWeights bw;
  Tables t;
  float act[kTile];
};
// decoder weight gradients (generator training, run.py:980-986): row tiles of H, dHpre and dY
constexpr int kHsLd = nfi_mlp::kH + 1, kDysLd = nfi_mlp::kY + 1;
struct BwdSmemW : BwdSmem {
  float hs[kTile * kHsLd];           // H = softplus(pre)
  float dhs[kTile * kHsLd];          // dHpre
  float dys[kTile * kDysLd];         // dY
};

// epilogue 0 / 2 that also leave their row values in shared memory (weight-gradient variant)
__device__ __forceinline__ void epilogue0_keep(const nfi_mlp::Operands& s, uint32_t tmem_lane, int col0, float* hrow) {
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    float v[16];
    nfi_tc::tmem_ld16(tmem_lane + nfi_mlp::kColD0 + col0 + 16 * q, v);
    nfi_tc::tmem_ld_wait();
    uint32_t hi[16], lo[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const float h = nfi_mlp::softplus_f(v[i] + s.b0[col0 + 16 * q + i]);
      hrow[col0 + 16 * q + i] = h;
      hi[i] = nfi_tc::tf32_hi(h);
      lo[i] = __float_as_uint(h - __uint_as_float(hi[i]));
    }
    nfi_tc::tmem_st16(tmem_lane + nfi_mlp::kColHhi + col0 + 16 * q, hi);
    nfi_tc::tmem_st16(tmem_lane + nfi_mlp::kColHlo + col0 + 16 * q, lo);
  }
  nfi_tc::tmem_st_wait();
}
__device__ __forceinline__ void epilogue2_keep(uint32_t tmem_lane, int col0, float* drow) {
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    float dh[16], hh[16], hl[16];
    nfi_tc::tmem_ld16(tmem_lane + nfi_mlp::kColDH + col0 + 16 * q, dh);
    nfi_tc::tmem_ld16(tmem_lane + nfi_mlp::kColHhi + col0 + 16 * q, hh);
    nfi_tc::tmem_ld16(tmem_lane + nfi_mlp::kColHlo + col0 + 16 * q, hl);
    nfi_tc::tmem_ld_wait();
    uint32_t hi[16], lo[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const float sg = 1.f - nfi_tc::ex2_ftz((hh[i] + hl[i]) * -1.4426950408889634f);
      const float g = dh[i] * sg;
      drow[col0 + 16 * q + i] = g;
      hi[i] = nfi_tc::tf32_hi(g);
      lo[i] = __float_as_uint(g - __uint_as_float(hi[i]));
    }
    nfi_tc::tmem_st16(tmem_lane + nfi_mlp::kColHhi + col0 + 16 * q, hi);
    nfi_tc::tmem_st16(tmem_lane + nfi_mlp::kColHlo + col0 + 16 * q, lo);
  }
  nfi_tc::tmem_st_wait();
}
static_assert(kTile * kDfLd <= 2 * 8 * (int)nfi_mlp::kLboA / 4, "d feature tile does not fit the operand tile");

// scatter d feat into the gradient planes; optionally d/d point from the re-read texels
template <bool WANT_PTS>
__device__ __forceinline__ void phase4_scatter(const Args& a, const Tables& t, const float* act,
                                               const float* dft, int b, int64_t tt) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c4 = lane & 7;
  const float third = 1.f / 3.f;
  const int plane_f4 = a.Hp * a.Wp * 8;
#pragma unroll 1
  for (int it = 0; it < 4; ++it) {
    const int s = warp * 16 + it * 4 + (lane >> 3);
    float gp[3] = {0.f, 0.f, 0.f};
    if (act[s] > 0.5f) {
      float4 g = *reinterpret_cast<const float4*>(&dft[s * kDfLd + c4 * 4]);
      g.x *= third; g.y *= third; g.z *= third; g.w *= third;
      const int64_t img_f4 = (int64_t)b * 3 * plane_f4 + c4;
#pragma unroll
      for (int p = 0; p < 3; ++p) {
        const int4 o = t.toff[p][s];
        const float4 w = t.twgt[p][s];
        const int64_t base = img_f4 + (int64_t)p * plane_f4;
        if (a.g_planes) {
          // a clamped neighbour (x1 == x0 at the border) carries weight 0: adding 0 is harmless
          float* G = a.g_planes;
          red_add_v4(G + (base + o.x) * 4, g.x * w.x, g.y * w.x, g.z * w.x, g.w * w.x);
          red_add_v4(G + (base + o.y) * 4, g.x * w.y, g.y * w.y, g.z * w.y, g.w * w.y);
          red_add_v4(G + (base + o.z) * 4, g.x * w.z, g.y * w.z, g.z * w.z, g.w * w.z);
          red_add_v4(G + (base + o.w) * 4, g.x * w.w, g.y * w.w, g.z * w.w, g.w * w.w);
        }
        if (WANT_PTS) {
          const float4* P = reinterpret_cast<const float4*>(a.planes) + base;
          const float4 v00 = __ldg(P + o.x), v01 = __ldg(P + o.y), v10 = __ldg(P + o.z), v11 = __ldg(P + o.w);
          const float4 d = t.tder[p][s];              // tx, ty, d ix / d x, d iy / d x
          // d/d ix = sum_c g_c ((v01 - v00)(1 - ty) + (v11 - v10) ty); d/d iy likewise
          const float ax = (v01.x - v00.x) * g.x + (v01.y - v00.y) * g.y + (v01.z - v00.z) * g.z + (v01.w - v00.w) * g.w;
          const float bx = (v11.x - v10.x) * g.x + (v11.y - v10.y) * g.y + (v11.z - v10.z) * g.z + (v11.w - v10.w) * g.w;
          const float ay = (v10.x - v00.x) * g.x + (v10.y - v00.y) * g.y + (v10.z - v00.z) * g.z + (v10.w - v00.w) * g.w;
          const float by = (v11.x - v01.x) * g.x + (v11.y - v01.y) * g.y + (v11.z - v01.z) * g.z + (v11.w - v01.w) * g.w;
          const float gx = fmaf(bx - ax, d.y, ax), gy = fmaf(by - ay, d.x, ay);
          gp[nfi_plane_axis_a(p)] = fmaf(gx, d.z, gp[nfi_plane_axis_a(p)]);
          gp[nfi_plane_axis_b(p)] = fmaf(gy, d.w, gp[nfi_plane_axis_b(p)]);
        }
      }
    }
    if (WANT_PTS) {
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        gp[c] += __shfl_xor_sync(0xffffffffu, gp[c], 1);
        gp[c] += __shfl_xor_sync(0xffffffffu, gp[c], 2);
        gp[c] += __shfl_xor_sync(0xffffffffu, gp[c], 4);
      }
      if (c4 == 0 && t.keep[s] != -1.f) {
        const int64_t id = (int64_t)b * (a.npi * a.S) + t.sid[s];
        const float inv_r = 1.f / a.f.scene_range;
        a.g_points[id * 3 + 0] = gp[0] * inv_r;
        a.g_points[id * 3 + 1] = gp[1] * inv_r;
        a.g_points[id * 3 + 2] = gp[2] * inv_r;
      }
    }
  }
}

// Non-specialised form (256 threads, __syncthreads between phases), kept for the case that also wants
// the decoder weight gradients: it has the feature tile, H, dHpre and dY of a tile in one place.
//   dW0[j][k] = g0 sum_rows dHpre[j] F[k]   db0[j] = sum_rows dHpre[j]     (g = EqualizedLinear gains,
//   dW1[o][j] = g1 sum_rows dY[o] H[j]      db1[o] = sum_rows dY[o]         models/stylegan.py:170-177)
// Each thread keeps its 8 + 4 (+2) partial sums in registers over all its tiles; one atomicAdd each at
// the end.
template <bool PTS, int MODE, bool WANT_PTS, bool WANT_W>
__global__ void __launch_bounds__(kThreads, WANT_W ? 1 : 2)
field_bwd_tc_kernel(const Args a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  BwdSmemW& sm = *reinterpret_cast<BwdSmemW*>(smem_raw);     // the W tail is only there when WANT_W
  nfi_mlp::Operands& op = sm.op;
  Tables& t = sm.t;
  float* dft = op.a_hi;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  nfi_mlp::load_weights(op, a.f.W0, a.f.b0, a.f.W1, a.f.b1, a.f.n_out);
  nfi_mlp::load_weights_bwd(sm.bw, a.f.W0, a.f.W1, a.f.n_out);
  if (warp == 0) nfi_tc::tmem_alloc(&op.tmem_base, nfi_mlp::kTmemCols);
  if (tid == 0) { nfi_tc::mbar_init(&op.bar, 1); nfi_tc::mbar_fence_init(); }
  nfi_tc::fence_proxy_async();
  nfi_tc::fence_before_sync();
  __syncthreads();
  nfi_tc::fence_after_sync();
  const uint32_t tmem = op.tmem_base;
  const uint32_t tmem_lane = tmem + ((uint32_t)((warp & 3) * 32) << 16);
  const NfiHeads hd = make_heads(a.f);
  const int npal = 3 * hd.A;
  const bool want_pal = MODE == 0 && a.g_palette != nullptr;
  const bool want_par = MODE == 0 && a.g_params != nullptr && hd.use_sdf;
  uint32_t parity = 0;
  const int64_t per_image = a.npi * a.S;
  // weight-gradient partial sums: thread -> (j = tid/4, k0 = 8 (tid%4)) of dW0, (o = tid/16, j0 = 4 (tid%16)) of dW1
  float aw0[8], aw1[4], ab0 = 0.f, ab1 = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) aw0[i] = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) aw1[i] = 0.f;
  for (int b = 0; b < a.B; ++b) {
    int64_t lo, hi;
    cta_tile_range(a.tiles_per_image, lo, hi);
    float dpal[kNPal];               // per-thread partial sums over this CTA's tiles of image b
#pragma unroll
    for (int q = 0; q < kNPal; ++q) dpal[q] = 0.f;
    float dbeta = 0.f, dalpha = 0.f;
    const float* pal = (MODE == 0 && a.palette) ? a.palette + (int64_t)b * npal : nullptr;
    for (int64_t tt = lo; tt < hi; ++tt) {
      int64_t id = -1;
      bool act = false;
      float gy[NFI_YPAD];
#pragma unroll
      for (int o = 0; o < NFI_YPAD; ++o) gy[o] = 0.f;
      if (tid < kTile) {
        const int sid = sample_in_image(a, (int)tt, tid);
        if (sid >= 0) id = (int64_t)b * per_image + sid;
        phase1<PTS, MODE, WANT_PTS>(a, t, id, sid, tid);
        if (id >= 0) {
          if (MODE == 0) {
            const float4 g = reinterpret_cast<const float4*>(a.g_out)[id];
            gy[0] = g.x; gy[1] = g.y; gy[2] = g.z; gy[3] = g.w;
            const bool col = (g.y != 0.f) || (g.z != 0.f) || (g.w != 0.f);
            // outside the box sigma is masked (generator.py:634): only a colour gradient can act;
            // samples with an all-zero upstream gradient are skipped exactly
            act = sample_active(a, t.keep[tid]) && (col || (g.x != 0.f && t.keep[tid] > 0.5f));
          } else {
#pragma unroll
            for (int o = 0; o < NFI_YPAD; ++o) {
              if (o < a.f.n_out) { gy[o] = a.g_out[id * a.f.n_out + o]; act = act || (gy[o] != 0.f); }
            }
          }
        }
        sm.act[tid] = act ? 1.f : 0.f;
        if (!act && t.keep[tid] >= 0.f) t.keep[tid] = -2.f;   // valid sample, nothing to do: no gather
      }
      const int any = __syncthreads_or(act ? 1 : 0);
      if (any) {
        phase2_gather(a, t, op, b);
        nfi_tc::fence_proxy_async();
        nfi_tc::fence_before_sync();
        __syncthreads();
        if (tid == 0) { nfi_tc::fence_after_sync(); nfi_mlp::issue_layer0(op, tmem); nfi_tc::mma_commit(&op.bar); }
        nfi_tc::mbar_wait(&op.bar, parity); parity ^= 1;
        nfi_tc::fence_after_sync();
        if (WANT_W) epilogue0_keep(op, tmem_lane, (warp >> 2) * 32, &sm.hs[((warp & 3) * 32 + lane) * kHsLd]);
        else nfi_mlp::epilogue0<2, false>(op, tmem_lane, (warp >> 2) * 32, nullptr);
        nfi_tc::fence_before_sync();
        __syncthreads();
        if (tid == 0) { nfi_tc::fence_after_sync(); nfi_mlp::issue_layer1(op, tmem); nfi_tc::mma_commit(&op.bar); }
        nfi_tc::mbar_wait(&op.bar, parity); parity ^= 1;
        nfi_tc::fence_after_sync();
        if (warp < 4) {
          float dy[NFI_YPAD];
#pragma unroll
          for (int o = 0; o < NFI_YPAD; ++o) dy[o] = 0.f;
          if (MODE == 0) {
            float y[nfi_mlp::kY];
            nfi_mlp::epilogue1(op, tmem_lane, y);
            if (act) {
              const float keep = t.keep[tid] > 0.5f ? 1.f : 0.f;
              float sigma, rgb[3], probs[NFI_YPAD - 1];
              nfi_heads_fwd(hd, y, keep, pal, sigma, rgb, probs);
              const float drgb[3] = {gy[1], gy[2], gy[3]};
              nfi_heads_bwd(hd, y, keep, pal, probs, rgb, gy[0], drgb, dy, want_par ? &dbeta : nullptr,
                            want_par ? &dalpha : nullptr);
              if (want_pal) {
#pragma unroll
                for (int q = 0; q < NFI_YPAD - 1; ++q) {
                  if (q < hd.A) {
                    dpal[3 * q] = fmaf(probs[q], drgb[0], dpal[3 * q]);
                    dpal[3 * q + 1] = fmaf(probs[q], drgb[1], dpal[3 * q + 1]);
                    dpal[3 * q + 2] = fmaf(probs[q], drgb[2], dpal[3 * q + 2]);
                  }
                }
              }
            }
          } else {
#pragma unroll
            for (int o = 0; o < NFI_YPAD; ++o) dy[o] = gy[o];
          }
          nfi_mlp::store_dy(tmem_lane, dy);
          if (WANT_W) {
#pragma unroll
            for (int o = 0; o < NFI_YPAD; ++o) sm.dys[tid * kDysLd + o] = dy[o];
          }
        }
        nfi_tc::fence_before_sync();
        __syncthreads();
        if (tid == 0) { nfi_tc::fence_after_sync(); nfi_mlp::issue_dh(sm.bw, tmem); nfi_tc::mma_commit(&op.bar); }
        nfi_tc::mbar_wait(&op.bar, parity); parity ^= 1;
        nfi_tc::fence_after_sync();
        if (WANT_W) epilogue2_keep(tmem_lane, (warp >> 2) * 32, &sm.dhs[((warp & 3) * 32 + lane) * kHsLd]);
        else nfi_mlp::epilogue2<2>(tmem_lane, (warp >> 2) * 32);
        nfi_tc::fence_before_sync();
        __syncthreads();
        if (tid == 0) { nfi_tc::fence_after_sync(); nfi_mlp::issue_df(sm.bw, tmem); nfi_tc::mma_commit(&op.bar); }
        if (WANT_W) {      // while the dF MMAs run: rank-128 updates of the weight gradients (idle rows are 0)
          const int j = tid >> 2, c40 = 2 * (tid & 3);
          const int o1 = tid >> 4, j0 = 4 * (tid & 15);
#pragma unroll 4
          for (int r = 0; r < kTile; ++r) {
            const float g = sm.dhs[r * kHsLd + j];
            const int f0 = c40 * (nfi_mlp::kLboA / 4) + r * 4, f1 = f0 + nfi_mlp::kLboA / 4;
            const float4 h0 = *reinterpret_cast<const float4*>(&op.a_hi[f0]), l0 = *reinterpret_cast<const float4*>(&op.a_lo[f0]);
            const float4 h1 = *reinterpret_cast<const float4*>(&op.a_hi[f1]), l1 = *reinterpret_cast<const float4*>(&op.a_lo[f1]);
            aw0[0] = fmaf(g, h0.x + l0.x, aw0[0]); aw0[1] = fmaf(g, h0.y + l0.y, aw0[1]);
            aw0[2] = fmaf(g, h0.z + l0.z, aw0[2]); aw0[3] = fmaf(g, h0.w + l0.w, aw0[3]);
            aw0[4] = fmaf(g, h1.x + l1.x, aw0[4]); aw0[5] = fmaf(g, h1.y + l1.y, aw0[5]);
            aw0[6] = fmaf(g, h1.z + l1.z, aw0[6]); aw0[7] = fmaf(g, h1.w + l1.w, aw0[7]);
            if ((tid & 3) == 0) ab0 += g;
            const float dyv = sm.dys[r * kDysLd + o1];
            const float* hr = &sm.hs[r * kHsLd + j0];
            aw1[0] = fmaf(dyv, hr[0], aw1[0]); aw1[1] = fmaf(dyv, hr[1], aw1[1]);
            aw1[2] = fmaf(dyv, hr[2], aw1[2]); aw1[3] = fmaf(dyv, hr[3], aw1[3]);
            if ((tid & 15) == 0) ab1 += dyv;
          }
          __syncthreads();       // the d feature tile written below overlays the feature tile read above
        }
        nfi_tc::mbar_wait(&op.bar, parity); parity ^= 1;
        nfi_tc::fence_after_sync();
        {   // d feature rows -> shared memory (row-major, padded)
          float v[16];
          const int col0 = (warp >> 2) * 16;
          nfi_tc::tmem_ld16(tmem_lane + nfi_mlp::kColDF + col0, v);
          nfi_tc::tmem_ld_wait();
          float* row = dft + ((warp & 3) * 32 + lane) * kDfLd + col0;
#pragma unroll
          for (int q = 0; q < 4; ++q)
            *reinterpret_cast<float4*>(row + 4 * q) = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
        }
        nfi_tc::fence_before_sync();
        __syncthreads();
      }
      // scatter (+ d/d point); with no active sample this only writes the zero point gradients
      if (any || WANT_PTS) phase4_scatter<WANT_PTS>(a, t, sm.act, dft, b, tt);
      __syncthreads();                 // tables and the d feature tile are rewritten by the next tile
    }
    // palette / beta / alpha partial sums of this CTA for image b: warp reduce, one atomic per warp
    if ((want_pal || want_par) && warp < 4 && hi > lo) {
      if (want_pal) {
#pragma unroll
        for (int q = 0; q < kNPal; ++q) {
          if (q < npal) {
            float v = dpal[q];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (lane == 0 && v != 0.f) atomicAdd(&a.g_palette[(int64_t)b * npal + q], v);
          }
        }
      }
      if (want_par) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          dbeta += __shfl_xor_sync(0xffffffffu, dbeta, o);
          dalpha += __shfl_xor_sync(0xffffffffu, dalpha, o);
        }
        const int base = NFI_HID * NFI_C + NFI_HID + a.f.n_out * NFI_HID + a.f.n_out;
        if (lane == 0) {
          if (dbeta != 0.f) atomicAdd(&a.g_params[base], dbeta);
          if (dalpha != 0.f) atomicAdd(&a.g_params[base + 1], dalpha);
        }
      }
    }
  }
  if (WANT_W && a.g_params) {
    const float g0 = (float)(1.0 / sqrt((double)nfi_mlp::kC)), g1 = (float)(1.0 / sqrt((double)nfi_mlp::kH));
    float* gW0 = a.g_params;
    float* gb0 = gW0 + NFI_HID * NFI_C;
    float* gW1 = gb0 + NFI_HID;
    float* gb1 = gW1 + a.f.n_out * NFI_HID;
    const int j = tid >> 2, k0 = 8 * (tid & 3), o1 = tid >> 4, j0 = 4 * (tid & 15);
#pragma unroll
    for (int i = 0; i < 8; ++i) if (aw0[i] != 0.f) atomicAdd(&gW0[j * NFI_C + k0 + i], aw0[i] * g0);
    if ((tid & 3) == 0 && ab0 != 0.f) atomicAdd(&gb0[j], ab0);
    if (o1 < a.f.n_out) {
#pragma unroll
      for (int i = 0; i < 4; ++i) if (aw1[i] != 0.f) atomicAdd(&gW1[o1 * NFI_HID + j0 + i], aw1[i] * g1);
      if ((tid & 15) == 0 && ab1 != 0.f) atomicAdd(&gb1[o1], ab1);
    }
  }
  __syncthreads();
  if (warp == 0) nfi_tc::tmem_dealloc(tmem, nfi_mlp::kTmemCols);
}

// ---------------------------------------------------------------------------------------------
// Decoder weight gradients of the training path (nfi_field_bwd_train): one pass over the per-sample rows the
// warp-specialised backward left in the workspace,
//   dW0[j][k] = g0 sum_s dHpre[s][j] F[s][k]    db0[j] = sum_s dHpre[s][j]      (g = EqualizedLinear gains,
//   dW1[o][j] = g1 sum_s dY[s][o]   H[s][j]     db1[o] = sum_s dY[s][o]          models/stylegan.py:170-177)
// 64 samples at a time through shared memory; warps 0-3 hold dW0 as 4 x 4 register blocks, warps 4-5 dW1,
// warps 6-7 the biases; one atomicAdd per output and CTA at the end.  Measured at config 5 (4.2 M samples,
// 2.95 GB of rows): 1.24 ms.  Two other forms were measured and were slower: a cp.async ring with packed FFMA2
// (1.41 ms) and one accumulator set per warp with broadcast loads (1.86 ms).
// ---------------------------------------------------------------------------------------------
constexpr int kWgChunk = 64, kWgThreads = 256;
struct WgSmem {
  float dh[kWgChunk][64];
  float h[kWgChunk][64];
  float f[kWgChunk][32];
  float dy[kWgChunk][16];
};

__global__ void __launch_bounds__(kWgThreads)
wgrad_reduce_kernel(const float* __restrict__ st_f, const float* __restrict__ st_h, const float* __restrict__ st_dh,
                    const float* __restrict__ st_dy, const unsigned char* __restrict__ st_ok, int64_t n, int n_out,
                    float* __restrict__ g_params) {
  __shared__ WgSmem sm;
  const int tid = threadIdx.x, warp = tid >> 5;
  float acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = 0.f;
  float bias = 0.f, bias1 = 0.f;
  const int jb = (tid & 127) >> 3, kb = tid & 7;            // dW0 block: rows 4 jb.., cols 4 kb..   (warps 0-3)
  const int ob = (tid - 128) >> 4, j1 = (tid - 128) & 15;   // dW1 block: rows 4 ob.., cols 4 j1..   (warps 4-5)
  const int64_t nchunks = (n + kWgChunk - 1) / kWgChunk;
  for (int64_t ch = blockIdx.x; ch < nchunks; ch += gridDim.x) {
    const int64_t s0 = ch * kWgChunk;
    __syncthreads();
    // cooperative load: rows of samples that took no part (or past the end) are zero
    for (int i = tid; i < kWgChunk * 16; i += kWgThreads) {           // 64-float rows: 16 float4 each
      const int r = i >> 4, q = i & 15;
      const int64_t sidx = s0 + r;
      const bool ok = sidx < n && st_ok[sidx] != 0;
      const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
      reinterpret_cast<float4*>(sm.dh[r])[q] = ok ? reinterpret_cast<const float4*>(st_dh + sidx * 64)[q] : z4;
      reinterpret_cast<float4*>(sm.h[r])[q] = ok ? reinterpret_cast<const float4*>(st_h + sidx * 64)[q] : z4;
      if (q < 8) reinterpret_cast<float4*>(sm.f[r])[q] = ok ? reinterpret_cast<const float4*>(st_f + sidx * 32)[q] : z4;
      if (q < 4) reinterpret_cast<float4*>(sm.dy[r])[q] = ok ? reinterpret_cast<const float4*>(st_dy + sidx * 16)[q] : z4;
    }
    __syncthreads();
    if (warp < 6) {
      const float* gsrc = warp < 4 ? &sm.dh[0][4 * jb] : &sm.dy[0][4 * ob];
      const float* xsrc = warp < 4 ? &sm.f[0][4 * kb] : &sm.h[0][4 * j1];
      const int gld = warp < 4 ? 64 : 16, xld = warp < 4 ? 32 : 64;
#pragma unroll 4
      for (int r = 0; r < kWgChunk; ++r) {
        const float4 g = *reinterpret_cast<const float4*>(gsrc + r * gld);
        const float4 x = *reinterpret_cast<const float4*>(xsrc + r * xld);
        acc[0] = fmaf(g.x, x.x, acc[0]); acc[1] = fmaf(g.x, x.y, acc[1]); acc[2] = fmaf(g.x, x.z, acc[2]); acc[3] = fmaf(g.x, x.w, acc[3]);
        acc[4] = fmaf(g.y, x.x, acc[4]); acc[5] = fmaf(g.y, x.y, acc[5]); acc[6] = fmaf(g.y, x.z, acc[6]); acc[7] = fmaf(g.y, x.w, acc[7]);
        acc[8] = fmaf(g.z, x.x, acc[8]); acc[9] = fmaf(g.z, x.y, acc[9]); acc[10] = fmaf(g.z, x.z, acc[10]); acc[11] = fmaf(g.z, x.w, acc[11]);
        acc[12] = fmaf(g.w, x.x, acc[12]); acc[13] = fmaf(g.w, x.y, acc[13]); acc[14] = fmaf(g.w, x.z, acc[14]); acc[15] = fmaf(g.w, x.w, acc[15]);
      }
    } else {
      const int j = tid - 192;                                          // 0..63
#pragma unroll 8
      for (int r = 0; r < kWgChunk; ++r) {
        bias += sm.dh[r][j];
        if (j < 16) bias1 += sm.dy[r][j];
      }
    }
  }
  const float g0 = (float)(1.0 / sqrt((double)nfi_mlp::kC)), g1 = (float)(1.0 / sqrt((double)nfi_mlp::kH));
  float* gW0 = g_params;
  float* gb0 = gW0 + NFI_HID * NFI_C;
  float* gW1 = gb0 + NFI_HID;
  float* gb1 = gW1 + n_out * NFI_HID;
  if (warp < 4) {
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int c = 0; c < 4; ++c)
        if (acc[4 * a + c] != 0.f) atomicAdd(&gW0[(4 * jb + a) * NFI_C + 4 * kb + c], acc[4 * a + c] * g0);
  } else if (warp < 6) {
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int c = 0; c < 4; ++c)
        if (4 * ob + a < n_out && acc[4 * a + c] != 0.f) atomicAdd(&gW1[(4 * ob + a) * NFI_HID + 4 * j1 + c], acc[4 * a + c] * g1);
  } else {
    const int j = tid - 192;
    if (bias != 0.f) atomicAdd(&gb0[j], bias);
    if (j < n_out && bias1 != 0.f) atomicAdd(&gb1[j], bias1);
  }
}

// The same reduction with the rows brought in by TMA bulk copies (cp.async.bulk, one per array and chunk: 16 + 16 + 8 + 4 KB
// of contiguous rows) into a two-stage shared-memory ring, completion on mbarriers: no load instructions, no shared-memory
// stores and no address arithmetic in the threads, and chunk i + 1 lands while chunk i is reduced.  Rows of samples that
// took no part hold whatever the workspace held: they are skipped by their validity flag (uniform per row).
struct WgTmaSmem {
  WgSmem st[2];
  int ok[2][kWgChunk];
  uint64_t full[2];
};

__device__ __forceinline__ void wg_tma_issue(WgTmaSmem& sm, int s, const float* st_f, const float* st_h, const float* st_dh,
                                             const float* st_dy, int64_t s0, int64_t n) {
  const uint32_t rows = (uint32_t)((n - s0) < kWgChunk ? (n - s0) : kWgChunk);
  nfi_tc::fence_proxy_async();                         // the stage was read through the generic proxy
  nfi_tc::mbar_arrive_expect_tx(&sm.full[s], rows * (64 + 64 + 32 + 16) * 4);
  nfi_tc::bulk_g2s(sm.st[s].dh, st_dh + s0 * 64, rows * 256, &sm.full[s]);
  nfi_tc::bulk_g2s(sm.st[s].h, st_h + s0 * 64, rows * 256, &sm.full[s]);
  nfi_tc::bulk_g2s(sm.st[s].f, st_f + s0 * 32, rows * 128, &sm.full[s]);
  nfi_tc::bulk_g2s(sm.st[s].dy, st_dy + s0 * 16, rows * 64, &sm.full[s]);
}

__global__ void __launch_bounds__(kWgThreads)
wgrad_reduce_tma_kernel(const float* __restrict__ st_f, const float* __restrict__ st_h, const float* __restrict__ st_dh,
                        const float* __restrict__ st_dy, const unsigned char* __restrict__ st_ok, int64_t n, int n_out,
                        float* __restrict__ g_params) {
  extern __shared__ __align__(128) unsigned char wg_raw[];
  WgTmaSmem& sm = *reinterpret_cast<WgTmaSmem*>(wg_raw);
  const int tid = threadIdx.x, warp = tid >> 5;
  if (tid == 0) { nfi_tc::mbar_init(&sm.full[0], 1); nfi_tc::mbar_init(&sm.full[1], 1); nfi_tc::mbar_fence_init(); }
  __syncthreads();
  float acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = 0.f;
  float bias = 0.f, bias1 = 0.f;
  const int jb = (tid & 127) >> 3, kb = tid & 7;
  const int ob = (tid - 128) >> 4, j1 = (tid - 128) & 15;
  const int64_t nchunks = (n + kWgChunk - 1) / kWgChunk;
  const int64_t first = blockIdx.x, stride = gridDim.x;
  if (tid == 0) {
    if (first < nchunks) wg_tma_issue(sm, 0, st_f, st_h, st_dh, st_dy, first * kWgChunk, n);
    if (first + stride < nchunks) wg_tma_issue(sm, 1, st_f, st_h, st_dh, st_dy, (first + stride) * kWgChunk, n);
  }
  int it = 0;
  for (int64_t ch = first; ch < nchunks; ch += stride, ++it) {
    const int s = it & 1;
    const int64_t s0 = ch * kWgChunk;
    if (tid < kWgChunk) sm.ok[s][tid] = (s0 + tid < n && st_ok[s0 + tid] != 0) ? 1 : 0;
    nfi_tc::mbar_wait(&sm.full[s], (it >> 1) & 1);
    __syncthreads();
    const WgSmem& st = sm.st[s];
    const int* ok = sm.ok[s];
    if (warp < 6) {
      const float* gsrc = warp < 4 ? &st.dh[0][4 * jb] : &st.dy[0][4 * ob];
      const float* xsrc = warp < 4 ? &st.f[0][4 * kb] : &st.h[0][4 * j1];
      const int gld = warp < 4 ? 64 : 16, xld = warp < 4 ? 32 : 64;
#pragma unroll 4
      for (int r = 0; r < kWgChunk; ++r) {
        if (!ok[r]) continue;
        const float4 g = *reinterpret_cast<const float4*>(gsrc + r * gld);
        const float4 x = *reinterpret_cast<const float4*>(xsrc + r * xld);
        acc[0] = fmaf(g.x, x.x, acc[0]); acc[1] = fmaf(g.x, x.y, acc[1]); acc[2] = fmaf(g.x, x.z, acc[2]); acc[3] = fmaf(g.x, x.w, acc[3]);
        acc[4] = fmaf(g.y, x.x, acc[4]); acc[5] = fmaf(g.y, x.y, acc[5]); acc[6] = fmaf(g.y, x.z, acc[6]); acc[7] = fmaf(g.y, x.w, acc[7]);
        acc[8] = fmaf(g.z, x.x, acc[8]); acc[9] = fmaf(g.z, x.y, acc[9]); acc[10] = fmaf(g.z, x.z, acc[10]); acc[11] = fmaf(g.z, x.w, acc[11]);
        acc[12] = fmaf(g.w, x.x, acc[12]); acc[13] = fmaf(g.w, x.y, acc[13]); acc[14] = fmaf(g.w, x.z, acc[14]); acc[15] = fmaf(g.w, x.w, acc[15]);
      }
    } else {
      const int j = tid - 192;
#pragma unroll 8
      for (int r = 0; r < kWgChunk; ++r) {
        if (!ok[r]) continue;
        bias += st.dh[r][j];
        if (j < 16) bias1 += st.dy[r][j];
      }
    }
    __syncthreads();                                   // stage s is free: refill it with the chunk two steps ahead
    const int64_t nxt = ch + 2 * stride;
    if (tid == 0 && nxt < nchunks) wg_tma_issue(sm, s, st_f, st_h, st_dh, st_dy, nxt * kWgChunk, n);
  }
  const float g0 = (float)(1.0 / sqrt((double)nfi_mlp::kC)), g1 = (float)(1.0 / sqrt((double)nfi_mlp::kH));
  float* gW0 = g_params;
  float* gb0 = gW0 + NFI_HID * NFI_C;
  float* gW1 = gb0 + NFI_HID;
  float* gb1 = gW1 + n_out * NFI_HID;
  if (warp < 4) {
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int c = 0; c < 4; ++c)
        if (acc[4 * a + c] != 0.f) atomicAdd(&gW0[(4 * jb + a) * NFI_C + 4 * kb + c], acc[4 * a + c] * g0);
  } else if (warp < 6) {
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int c = 0; c < 4; ++c)
        if (4 * ob + a < n_out && acc[4 * a + c] != 0.f) atomicAdd(&gW1[(4 * ob + a) * NFI_HID + 4 * j1 + c], acc[4 * a + c] * g1);
  } else {
    const int j = tid - 192;
    if (bias != 0.f) atomicAdd(&gb0[j], bias);
    if (j < n_out && bias1 != 0.f) atomicAdd(&gb1[j], bias1);
  }
}

static inline size_t wg_align(size_t v) { return (v + 255) & ~(size_t)255; }

}  // namespace nfi_ftc

extern "C" size_t nfi_field_bwd_train_workspace_bytes(int64_t n_samples) {
  if (n_samples <= 0) return 0;
  const size_t n = (size_t)n_samples;
  return nfi_ftc::wg_align(n * 32 * 4) + 2 * nfi_ftc::wg_align(n * 64 * 4) + nfi_ftc::wg_align(n * 16 * 4) + nfi_ftc::wg_align(n);
}

extern "C" int nfi_field_bwd_train(const NfiField* field, const float* planes_cl, const float* palette, int B,
                                   int Hp, int Wp, const float* origins, const float* dirs, const float* z,
                                   int64_t n_per_image, int S, int img_w, const float* g_out, float* g_planes_cl,
                                   float* g_palette, float* g_points, float* g_params, void* workspace,
                                   size_t workspace_bytes, nfi_stream_t stream) {
  using namespace nfi_ftc;
  Args a;
  int rc = fill_args(a, "nfi_field_bwd_train", field, planes_cl, palette, B, Hp, Wp, origins, dirs, z, nullptr,
                     n_per_image, S, 0, img_w);
  if (rc) return rc;
  NFI_REQUIRE(g_out && g_params && workspace, NFI_ERR_BAD_ARG, "nfi_field_bwd_train: NULL g_out / g_params / workspace");
  NFI_REQUIRE(nfi_aligned16(g_out), NFI_ERR_ALIGNMENT, "nfi_field_bwd_train: g_out must be 16-byte aligned");
  NFI_REQUIRE(!g_planes_cl || nfi_aligned16(g_planes_cl), NFI_ERR_ALIGNMENT, "nfi_field_bwd_train: g_planes must be 16-byte aligned");
  NFI_REQUIRE((((uintptr_t)workspace) & 255u) == 0, NFI_ERR_ALIGNMENT, "nfi_field_bwd_train: workspace must be 256-byte aligned");
  const int64_t n = a.n_total;
  NFI_REQUIRE(workspace_bytes >= nfi_field_bwd_train_workspace_bytes(n), NFI_ERR_BAD_ARG,
              "nfi_field_bwd_train: workspace of %zu bytes, %zu needed", workspace_bytes,
              nfi_field_bwd_train_workspace_bytes(n));
  char* w = (char*)workspace;
  a.st_f = (float*)w;  w += wg_align((size_t)n * 32 * 4);
  a.st_h = (float*)w;  w += wg_align((size_t)n * 64 * 4);
  a.st_dh = (float*)w; w += wg_align((size_t)n * 64 * 4);
  a.st_dy = (float*)w; w += wg_align((size_t)n * 16 * 4);
  a.st_ok = (unsigned char*)w;
  a.g_out = g_out; a.g_planes = g_planes_cl; a.g_palette = g_palette; a.g_points = g_points; a.g_params = g_params;
  cudaStream_t st = nfi_cu(stream);
  const int grid = field_grid_chunks(a);
  const int smem = (int)sizeof(BwSmem);
#define NFI_LAUNCH_TRAIN(W, H)                                                                                     \
  do {                                                                                                           \
    rc = nfi_set_smem_once((const void*)field_bwd_ws_kernel<false, 0, W, H, true>, smem, "nfi_field_bwd_train"); \
    if (rc) return rc;                                                                                           \
    field_bwd_ws_kernel<false, 0, W, H, true><<<grid, kBwThreads, smem, st>>>(a);                                \
  } while (0)
  if (a.f.planes_bf16) { if (g_points) NFI_LAUNCH_TRAIN(true, true); else NFI_LAUNCH_TRAIN(false, true); }
  else                 { if (g_points) NFI_LAUNCH_TRAIN(true, false); else NFI_LAUNCH_TRAIN(false, false); }
#undef NFI_LAUNCH_TRAIN
  NFI_CHECK_LAUNCH("field_bwd_ws_kernel (training path)");
  const int64_t nchunks = (n + kWgChunk - 1) / kWgChunk;
  const int64_t cap = (int64_t)nfi_sm_count() * 4;
#ifdef NFI_WGRAD_NO_TMA
  wgrad_reduce_kernel<<<(int)(nchunks < cap ? nchunks : cap), kWgThreads, 0, st>>>(
      a.st_f, a.st_h, a.st_dh, a.st_dy, a.st_ok, n, a.f.n_out, g_params);
#else
  {
    const int wg_smem = (int)sizeof(WgTmaSmem);
    rc = nfi_set_smem_once((const void*)wgrad_reduce_tma_kernel, wg_smem, "nfi_field_bwd_train");
    if (rc) return rc;
    const int64_t cap2 = (int64_t)nfi_sm_count() * 2;            // 2 CTAs of 92 KB per SM
    wgrad_reduce_tma_kernel<<<(int)(nchunks < cap2 ? nchunks : cap2), kWgThreads, wg_smem, st>>>(
        a.st_f, a.st_h, a.st_dh, a.st_dy, a.st_ok, n, a.f.n_out, g_params);
  }
#endif
  NFI_CHECK_LAUNCH("wgrad_reduce_kernel");
  return NFI_OK;
}

extern "C" int nfi_field_bwd(const NfiField* field, const float* planes_cl, const float* palette, int B,
                             int Hp, int Wp, const float* origins, const float* dirs, const float* z,
                             const float* points, int64_t n_per_image, int S, int mode, int img_w,
                             const float* g_out, float* g_planes_cl, float* g_palette, float* g_points,
                             float* g_params, int want_weight_grads, nfi_stream_t stream) {
  using namespace nfi_ftc;
  Args a;
  int rc = fill_args(a, "nfi_field_bwd", field, planes_cl, palette, B, Hp, Wp, origins, dirs, z, points,
                     n_per_image, S, mode, img_w);
  if (rc) return rc;
  NFI_REQUIRE(g_out, NFI_ERR_BAD_ARG, "nfi_field_bwd: NULL upstream gradient");
  NFI_REQUIRE(mode == 1 || nfi_aligned16(g_out), NFI_ERR_ALIGNMENT, "nfi_field_bwd: g_out must be 16-byte aligned");
  NFI_REQUIRE(!g_planes_cl || nfi_aligned16(g_planes_cl), NFI_ERR_ALIGNMENT, "nfi_field_bwd: g_planes must be 16-byte aligned");
  a.g_out = g_out; a.g_planes = g_planes_cl; a.g_palette = (mode == 0) ? g_palette : nullptr;
  a.g_points = g_points; a.g_params = g_params;
  cudaStream_t st = nfi_cu(stream);
  const bool wp = g_points != nullptr;
  const bool want_w = want_weight_grads != 0 && g_params != nullptr;
  NFI_REQUIRE(!(want_w && field->planes_bf16), NFI_ERR_UNSUPPORTED,
              "nfi_field_bwd: decoder weight gradients are not available with a bf16 plane copy");
  if (!want_w) {
    const int grid = field_grid_chunks(a);
    const int smem = (int)sizeof(BwSmem);
#define NFI_LAUNCH_BWS1(P, M, W, H)                                                                     \
  do {                                                                                                  \
    const int rc_ = nfi_set_smem_once((const void*)field_bwd_ws_kernel<P, M, W, H>, smem, "nfi_field_bwd"); \
    if (rc_) return rc_;                                                                                \
    field_bwd_ws_kernel<P, M, W, H><<<grid, kBwThreads, smem, st>>>(a);                                 \
  } while (0)
#define NFI_LAUNCH_BWS(P, M, W)                                                                         \
  do {                                                                                                  \
    if (a.f.planes_bf16) NFI_LAUNCH_BWS1(P, M, W, true); else NFI_LAUNCH_BWS1(P, M, W, false);          \
  } while (0)
    if (points) {
      if (mode == 0) { if (wp) NFI_LAUNCH_BWS(true, 0, true); else NFI_LAUNCH_BWS(true, 0, false); }
      else           { if (wp) NFI_LAUNCH_BWS(true, 1, true); else NFI_LAUNCH_BWS(true, 1, false); }
    } else {
      if (mode == 0) { if (wp) NFI_LAUNCH_BWS(false, 0, true); else NFI_LAUNCH_BWS(false, 0, false); }
      else           { if (wp) NFI_LAUNCH_BWS(false, 1, true); else NFI_LAUNCH_BWS(false, 1, false); }
    }
#undef NFI_LAUNCH_BWS
#undef NFI_LAUNCH_BWS1
    NFI_CHECK_LAUNCH("field_bwd_ws_kernel");
#ifdef NFI_WS_TRACE
    if (const char* path = getenv("NFI_WS_TRACE_FILE")) {       // developer builds only: blocking dump
      static int launch = 0;
      static long long host[kBtRoles * kBtN * kBtEv];
      cudaStreamSynchronize(st);
      cudaMemcpyFromSymbol(host, g_bw_trace, sizeof(host));
      char name[512];
      snprintf(name, sizeof(name), "%s.bwd.%d", path, launch++ % 2);
      if (FILE* fp = fopen(name, "w")) {
        for (int r = 0; r < kBtRoles; ++r)
          for (int t = 0; t < kBtN; ++t) {
            fprintf(fp, "%d %d", r, t);
            for (int e = 0; e < kBtEv; ++e) fprintf(fp, " %lld", host[(r * kBtN + t) * kBtEv + e]);
            fprintf(fp, "\n");
          }
        fclose(fp);
      }
    }
#endif
    return NFI_OK;
  }
  const int grid = field_grid(a, 1);
  const int smem = (int)sizeof(BwdSmemW);
#define NFI_LAUNCH_BWD(P, M, W)                                                                         \
  do {                                                                                                  \
    const int rc_ = nfi_set_smem_once((const void*)field_bwd_tc_kernel<P, M, W, true>, smem, "nfi_field_bwd"); \
    if (rc_) return rc_;                                                                                \
    field_bwd_tc_kernel<P, M, W, true><<<grid, kThreads, smem, st>>>(a);                                \
  } while (0)
  if (points) {
    if (mode == 0) { if (wp) NFI_LAUNCH_BWD(true, 0, true); else NFI_LAUNCH_BWD(true, 0, false); }
    else           { if (wp) NFI_LAUNCH_BWD(true, 1, true); else NFI_LAUNCH_BWD(true, 1, false); }
  } else {
    if (mode == 0) { if (wp) NFI_LAUNCH_BWD(false, 0, true); else NFI_LAUNCH_BWD(false, 0, false); }
    else           { if (wp) NFI_LAUNCH_BWD(false, 1, true); else NFI_LAUNCH_BWD(false, 1, false); }
  }
#undef NFI_LAUNCH_BWD
  NFI_CHECK_LAUNCH("field_bwd_tc_kernel");
  return NFI_OK;
}

#ifdef NFI_WAIT_FLAG
// Developer reproducer builds only (not declared in include/nfi_b200.h, absent from the shipped library):
// number of bounded mbarrier waits of this translation unit that ran out of spins since the last reset.
// Synchronises the device.
extern "C" int nfi_debug_wait_timeouts(int reset) {
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
