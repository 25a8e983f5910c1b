// Backward of the fused gather + MLP + heads kernel (see nfi_field_kernels.cuh for the tiling).
// The forward is recomputed per tile (nothing but 16 B/sample of upstream gradient is read).
#pragma once
#include "nfi_field_kernels.cuh"

namespace nfi_field {

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}

struct BwdSmem {
  float4 W0jk[NFI_HID * NFI_C / 4];   // [64][32]
  float red[kTile / 32][3 * (NFI_YPAD - 1) + 2];
  float act[kTile];                   // 1 = this sample has a non-zero upstream gradient
};

// phase 4: scatter d feat into the gradient planes; optionally d/d point.
template <bool WANT_PTS>
__device__ __forceinline__ void phase4_scatter(const Args& a, TileSmem& t, const BwdSmem& bs,
                                               int64_t tile) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c4 = lane & 7;
  const float third = 1.f / 3.f;
#pragma unroll 1
  for (int it = 0; it < 8; ++it) {
    const int s = warp * 32 + it * 4 + (lane >> 3);
    const bool act = bs.act[s] > 0.5f;
    float gp[3] = {0.f, 0.f, 0.f};
    if (act) {
      float4 g = *reinterpret_cast<const float4*>(&t.feat[s * kFeatLd + c4 * 4]);
      g.x *= third; g.y *= third; g.z *= third; g.w *= third;
      const int64_t plane_sz = (int64_t)a.Hp * a.Wp * NFI_C;
      const int64_t img_off = (int64_t)t.img[s] * 3 * plane_sz + c4 * 4;
#pragma unroll
      for (int p = 0; p < 3; ++p) {
        NfiFoot f;
        const float ix = t.coord[2 * p][s], iy = t.coord[2 * p + 1][s];
        nfi_footprint(ix, iy, a.Wp, a.Hp, f);
        const int64_t o00 = img_off + p * plane_sz + (int64_t)(f.y0 * a.Wp + f.x0) * NFI_C;
        const int64_t o01 = img_off + p * plane_sz + (int64_t)(f.y0 * a.Wp + f.x1) * NFI_C;
        const int64_t o10 = img_off + p * plane_sz + (int64_t)(f.y1 * a.Wp + f.x0) * NFI_C;
        const int64_t o11 = img_off + p * plane_sz + (int64_t)(f.y1 * a.Wp + f.x1) * NFI_C;
        if (a.g_planes) {
          // a clamped neighbour (x1 == x0 at the border) carries weight 0: adding 0 is harmless
          red_add_v4(a.g_planes + o00, g.x * f.wnw, g.y * f.wnw, g.z * f.wnw, g.w * f.wnw);
          red_add_v4(a.g_planes + o01, g.x * f.wne, g.y * f.wne, g.z * f.wne, g.w * f.wne);
          red_add_v4(a.g_planes + o10, g.x * f.wsw, g.y * f.wsw, g.z * f.wsw, g.w * f.wsw);
          red_add_v4(a.g_planes + o11, g.x * f.wse, g.y * f.wse, g.z * f.wse, g.w * f.wse);
        }
        if (WANT_PTS) {
          const float4 v00 = __ldg(reinterpret_cast<const float4*>(a.planes + o00));
          const float4 v01 = __ldg(reinterpret_cast<const float4*>(a.planes + o01));
          const float4 v10 = __ldg(reinterpret_cast<const float4*>(a.planes + o10));
          const float4 v11 = __ldg(reinterpret_cast<const float4*>(a.planes + o11));
          // out-of-range neighbours contribute 0 in ATen (within_bounds_2d)
          const float vx = (f.x0 + 1 < a.Wp) ? 1.f : 0.f, vy = (f.y0 + 1 < a.Hp) ? 1.f : 0.f;
          float gx = 0.f, gy = 0.f;
          nfi_bilerp_grad(g.x, v00.x, v01.x, v10.x, v11.x, vx, vy, f, gx, gy);
          nfi_bilerp_grad(g.y, v00.y, v01.y, v10.y, v11.y, vx, vy, f, gx, gy);
          nfi_bilerp_grad(g.z, v00.z, v01.z, v10.z, v11.z, vx, vy, f, gx, gy);
          nfi_bilerp_grad(g.w, v00.w, v01.w, v10.w, v11.w, vx, vy, f, gx, gy);
          const float mx = nfi_coord_mult(ix, a.Wp), my = nfi_coord_mult(iy, a.Hp);
          gp[nfi_plane_axis_a(p)] += gx * mx;
          gp[nfi_plane_axis_b(p)] += gy * my;
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
        const int64_t id = sample_id(a, tile, s);
        const float inv_r = 1.f / a.f.scene_range;
        a.g_points[id * 3 + 0] = gp[0] * inv_r;
        a.g_points[id * 3 + 1] = gp[1] * inv_r;
        a.g_points[id * 3 + 2] = gp[2] * inv_r;
      }
    }
  }
}

template <bool PTS, int MODE, bool WANT_PTS>
__global__ void __launch_bounds__(kTile)
field_bwd_kernel(const Args a) {
  __shared__ __align__(16) Weights w;
  __shared__ __align__(16) TileSmem t;
  __shared__ __align__(16) BwdSmem bs;
  load_weights(w, a.f, true);
  {
    const float g0 = (float)(1.0 / sqrt((double)NFI_C));
    float* W0jk = reinterpret_cast<float*>(bs.W0jk);
    for (int i = threadIdx.x; i < NFI_HID * NFI_C; i += blockDim.x) W0jk[i] = NFI_MUL(a.f.W0[i], g0);
  }
  const NfiHeads hd = make_heads(a.f);
  const int no4 = (a.f.n_out + 3) >> 2;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int npal = 3 * hd.A;
  for (int64_t tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
    const int64_t id = sample_id(a, tile, tid);
    float pt[3];
    __syncthreads();
    phase1<PTS, MODE>(a, t, id, tid, pt);
    // upstream gradient: samples with an all-zero gradient are skipped exactly
    float gy[NFI_YPAD];
#pragma unroll
    for (int o = 0; o < NFI_YPAD; ++o) gy[o] = 0.f;
    bool act = false;
    if (id >= 0) {
      if (MODE == 0) {
        const float4 g = reinterpret_cast<const float4*>(a.g_out)[id];
        gy[0] = g.x; gy[1] = g.y; gy[2] = g.z; gy[3] = g.w;
        const bool col = (g.y != 0.f) || (g.z != 0.f) || (g.w != 0.f);
        // outside the box sigma is masked (generator.py:634): only a colour gradient can act
        act = col || (g.x != 0.f && t.keep[tid] > 0.5f);
      } else {
#pragma unroll
        for (int o = 0; o < NFI_YPAD; ++o) {
          if (o < a.f.n_out) { gy[o] = a.g_out[id * a.f.n_out + o]; act = act || (gy[o] != 0.f); }
        }
      }
    }
    bs.act[tid] = act ? 1.f : 0.f;
    if (!act && t.keep[tid] >= 0.f) t.keep[tid] = -2.f;   // valid sample, nothing to do: skip the gather
    __syncthreads();
    phase2_gather(a, t);
    __syncthreads();
    float dpal[3 * (NFI_YPAD - 1)];
#pragma unroll
    for (int q = 0; q < 3 * (NFI_YPAD - 1); ++q) dpal[q] = 0.f;
    float dbeta = 0.f, dalpha = 0.f;
    if (act) {
      const float keep = t.keep[tid] > 0.5f ? 1.f : 0.f;
      float feat[NFI_C], hpre[NFI_HID], y[NFI_YPAD], dy[NFI_YPAD];
#pragma unroll
      for (int k4 = 0; k4 < NFI_C / 4; ++k4) {
        const float4 v = *reinterpret_cast<const float4*>(&t.feat[tid * kFeatLd + k4 * 4]);
        feat[4 * k4] = v.x; feat[4 * k4 + 1] = v.y; feat[4 * k4 + 2] = v.z; feat[4 * k4 + 3] = v.w;
      }
      nfi_mlp_layer0<float4>(feat, w.W0kj, w.b0, hpre);
      if (MODE == 0) {
        nfi_mlp_layer1<float4>(hpre, w.W1jo, w.b1, no4, y);
        const float* pal = a.palette ? a.palette + (int64_t)t.img[tid] * npal : nullptr;
        float sigma, rgb[3], probs[NFI_YPAD - 1];
        nfi_heads_fwd(hd, y, keep, pal, sigma, rgb, probs);
        const float drgb[3] = {gy[1], gy[2], gy[3]};
        nfi_heads_bwd(hd, y, keep, pal, probs, rgb, gy[0], drgb, dy,
                      a.g_params ? &dbeta : nullptr, a.g_params ? &dalpha : nullptr);
        if (a.g_palette) {
#pragma unroll
          for (int q = 0; q < NFI_YPAD - 1; ++q) {
            if (q < hd.A) { dpal[3 * q] = probs[q] * drgb[0]; dpal[3 * q + 1] = probs[q] * drgb[1]; dpal[3 * q + 2] = probs[q] * drgb[2]; }
          }
        }
      } else {
#pragma unroll
        for (int o = 0; o < NFI_YPAD; ++o) dy[o] = gy[o];
      }
      nfi_mlp_layer1_bwd<float4>(hpre, w.W1jo, no4, dy);     // hpre <- d hpre
      nfi_mlp_layer0_bwd<float4>(hpre, bs.W0jk, feat);       // feat <- d feat
#pragma unroll
      for (int k4 = 0; k4 < NFI_C / 4; ++k4)
        *reinterpret_cast<float4*>(&t.feat[tid * kFeatLd + k4 * 4]) =
            make_float4(feat[4 * k4], feat[4 * k4 + 1], feat[4 * k4 + 2], feat[4 * k4 + 3]);
    }
    // palette (+ beta, alpha) gradients: warp shuffle reduction, then one atomic per CTA and entry
    if (MODE == 0 && (a.g_palette || a.g_params)) {
      const int b0 = t.img[0];
      const bool uniform = __syncthreads_and((t.keep[tid] == -1.f) || (t.img[tid] == b0));
      if (uniform) {
#pragma unroll
        for (int q = 0; q < 3 * (NFI_YPAD - 1); ++q) {
          if (q < npal) {
            float v = dpal[q];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (lane == 0) bs.red[warp][q] = v;
          }
        }
        float vb = dbeta, va = dalpha;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { vb += __shfl_xor_sync(0xffffffffu, vb, o); va += __shfl_xor_sync(0xffffffffu, va, o); }
        if (lane == 0) { bs.red[warp][3 * (NFI_YPAD - 1)] = vb; bs.red[warp][3 * (NFI_YPAD - 1) + 1] = va; }
        __syncthreads();
        if (tid < npal && a.g_palette) {
          float v = 0.f;
          for (int ww = 0; ww < kTile / 32; ++ww) v += bs.red[ww][tid];
          if (v != 0.f) atomicAdd(&a.g_palette[(int64_t)b0 * npal + tid], v);
        }
        if (a.g_params && tid >= 64 && tid < 66) {
          float v = 0.f;
          for (int ww = 0; ww < kTile / 32; ++ww) v += bs.red[ww][3 * (NFI_YPAD - 1) + (tid - 64)];
          const int base = NFI_HID * NFI_C + NFI_HID + a.f.n_out * NFI_HID + a.f.n_out;
          if (v != 0.f) atomicAdd(&a.g_params[base + (tid - 64)], v);
        }
      } else if (act) {
        if (a.g_palette) {
#pragma unroll
          for (int q = 0; q < 3 * (NFI_YPAD - 1); ++q)
            if (q < npal) atomicAdd(&a.g_palette[(int64_t)t.img[tid] * npal + q], dpal[q]);
        }
        if (a.g_params) {
          const int base = NFI_HID * NFI_C + NFI_HID + a.f.n_out * NFI_HID + a.f.n_out;
          atomicAdd(&a.g_params[base], dbeta); atomicAdd(&a.g_params[base + 1], dalpha);
        }
      }
    }
    __syncthreads();
    phase4_scatter<WANT_PTS>(a, t, bs, tile);
  }
}

}  // namespace nfi_field
