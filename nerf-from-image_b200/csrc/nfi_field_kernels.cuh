// A4-A6 kernels: bilinear triplane gather + decoder MLP + density/colour heads, fused, forward and
// backward (models/generator.py:288-331, 587-681).  Features / hidden activations never touch HBM.
//
// Tile = 128 samples per pass of a 128-thread CTA (persistent CTAs, grid-stride over tiles):
//   phase 1  thread-per-sample : point -> box coordinates -> 3 x (ix, iy) in shared memory
//   phase 2  8 lanes per texel : each lane owns 4 of the 32 channels (one float4 = 16 B), so a
//            texel is one fully-used 128 B line and a warp instruction fetches 4 texels;
//            the weighted sum over 3 planes x 4 texels lands in the shared feature tile
//   phase 3  thread-per-sample : MLP 32 -> 64 -> n_out on FP32 FMA pipes (TF32 is disabled in the
//            reference, run.py:59-60, so the fp32 parity mode cannot use single-pass tensor cores),
//            heads, and in the backward the MLP/heads adjoint -> d feature tile in shared memory
//   phase 4  (backward) 8 lanes per texel: red.global.add.v4.f32 scatter into the channels-last
//            gradient planes + analytic d/d point from the re-read texels.
#pragma once
#include "nfi_common.cuh"
#include "nfi_field_math.cuh"

namespace nfi_field {

constexpr int kTile = 128;
constexpr int kFeatLd = NFI_C + 4;   // 36 floats: conflict-free float4 rows

struct Args {
  NfiField f;
  const float* planes;    // [B,3,Hp,Wp,32]
  const float* palette;   // [B,A,3]
  int B, Hp, Wp;
  const float* origins; const float* dirs; const float* z; const float* points;
  int64_t npi;            // rays (or points) per image
  int S;
  int img_w;              // > 0: pixel-block tile order is allowed
  int skip_oob;           // fused render: out-of-box samples produce (0,0,0,0) without gathering
  int64_t n_total, n_tiles;
  int blocks_x;           // pixel-block order: blocks per image row (0 = linear order)
  float* out; float* probs;
  // backward
  const float* g_out;
  float* g_planes; float* g_palette; float* g_points; float* g_params;
};

struct Weights {          // gain-scaled copies in shared memory
  float4 W0kj[NFI_C * NFI_HID / 4];     // [32][64]
  float4 b0[NFI_HID / 4];
  float4 W1jo[NFI_HID * NFI_YPAD / 4];  // [64][16]
  float4 b1[NFI_YPAD / 4];
};

__device__ __forceinline__ void load_weights(Weights& w, const NfiField& f, bool /*bwd*/) {
  const float g0 = (float)(1.0 / sqrt((double)NFI_C));     // stylegan.py:170: weight * (1/sqrt(in))
  const float g1 = (float)(1.0 / sqrt((double)NFI_HID));
  float* W0 = reinterpret_cast<float*>(w.W0kj);
  float* W1 = reinterpret_cast<float*>(w.W1jo);
  float* b0 = reinterpret_cast<float*>(w.b0);
  float* b1 = reinterpret_cast<float*>(w.b1);
  for (int i = threadIdx.x; i < NFI_C * NFI_HID; i += blockDim.x) {
    const int j = i / NFI_C, k = i % NFI_C;               // source is [64][32]
    W0[k * NFI_HID + j] = NFI_MUL(f.W0[i], g0);
  }
  for (int i = threadIdx.x; i < NFI_HID * NFI_YPAD; i += blockDim.x) {
    const int j = i / NFI_YPAD, o = i % NFI_YPAD;
    W1[i] = o < f.n_out ? NFI_MUL(f.W1[o * NFI_HID + j], g1) : 0.f;
  }
  for (int i = threadIdx.x; i < NFI_HID; i += blockDim.x) b0[i] = f.b0[i];
  for (int i = threadIdx.x; i < NFI_YPAD; i += blockDim.x) b1[i] = i < f.n_out ? f.b1[i] : 0.f;
}

// tile-order position -> linear sample id (ray * S + k), or -1
__device__ __forceinline__ int64_t sample_id(const Args& a, int64_t tile, int tid) {
  if (a.blocks_x == 0) {
    const int64_t id = tile * kTile + tid;
    return id < a.n_total ? id : -1;
  }
  // pixel-block order: 8 x 16 pixels of one image at one depth step
  const int H = (int)(a.npi / a.img_w);
  const int64_t tiles_per_image = (int64_t)(H / 8) * a.blocks_x * a.S;
  const int64_t b = tile / tiles_per_image;
  const int64_t tb = tile - b * tiles_per_image;
  const int64_t blk = tb / a.S;
  const int k = (int)(tb - blk * a.S);
  const int by = (int)(blk / a.blocks_x), bx = (int)(blk - (int64_t)by * a.blocks_x);
  const int py = tid >> 4, px = tid & 15;
  const int64_t ray = b * a.npi + (int64_t)(by * 8 + py) * a.img_w + bx * 16 + px;
  return ray * a.S + k;
}

struct TileSmem {
  float coord[6][kTile];   // clamped (ix, iy) per plane
  float keep[kTile];       // 1 = inside box, 0 = outside, -1 = no sample / skipped
  int img[kTile];
  float feat[kTile * kFeatLd];
};

template <bool PTS, int MODE>
__device__ __forceinline__ void phase1(const Args& a, TileSmem& t, int64_t id, int tid, float* pt) {
  float keep = -1.f;
  int b = 0;
  if (id >= 0) {
    const int64_t ray = id / a.S;
    b = (int)(ray / a.npi);
    if (PTS) { pt[0] = a.points[id * 3]; pt[1] = a.points[id * 3 + 1]; pt[2] = a.points[id * 3 + 2]; }
    else {
      const float o[3] = {a.origins[ray * 3], a.origins[ray * 3 + 1], a.origins[ray * 3 + 2]};
      const float d[3] = {a.dirs[ray * 3], a.dirs[ray * 3 + 1], a.dirs[ray * 3 + 2]};
      nfi_point(o, d, a.z[id], pt);
    }
    float x[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) x[c] = NFI_DIV(pt[c], a.f.scene_range);
    keep = 1.f;
    if (MODE == 0 && (fabsf(x[0]) > 1.f || fabsf(x[1]) > 1.f || fabsf(x[2]) > 1.f)) keep = 0.f;
#pragma unroll
    for (int p = 0; p < 3; ++p) {
      t.coord[2 * p][tid] = nfi_unnormalize(x[nfi_plane_axis_a(p)], a.Wp, nullptr);
      t.coord[2 * p + 1][tid] = nfi_unnormalize(x[nfi_plane_axis_b(p)], a.Hp, nullptr);
    }
  }
  t.keep[tid] = keep;
  t.img[tid] = b;
}

__device__ __forceinline__ bool sample_active(const Args& a, float keep) {
  return keep > 0.5f || (keep == 0.f && !a.skip_oob);
}

// phase 2: the warp gathers its 32 samples, 4 at a time
__device__ __forceinline__ void phase2_gather(const Args& a, TileSmem& t) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c4 = lane & 7;
#pragma unroll 2
  for (int it = 0; it < 8; ++it) {
    const int s = warp * 32 + it * 4 + (lane >> 3);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (sample_active(a, t.keep[s])) {
      const float4* base = reinterpret_cast<const float4*>(a.planes) +
                           (int64_t)t.img[s] * 3 * a.Hp * a.Wp * (NFI_C / 4) + c4;
      NfiFoot f[3];
      float4 v[3][4];
#pragma unroll
      for (int p = 0; p < 3; ++p) {
        nfi_footprint(t.coord[2 * p][s], t.coord[2 * p + 1][s], a.Wp, a.Hp, f[p]);
        const float4* P = base + (int64_t)p * a.Hp * a.Wp * (NFI_C / 4);
        v[p][0] = __ldg(P + (f[p].y0 * a.Wp + f[p].x0) * (NFI_C / 4));
        v[p][1] = __ldg(P + (f[p].y0 * a.Wp + f[p].x1) * (NFI_C / 4));
        v[p][2] = __ldg(P + (f[p].y1 * a.Wp + f[p].x0) * (NFI_C / 4));
        v[p][3] = __ldg(P + (f[p].y1 * a.Wp + f[p].x1) * (NFI_C / 4));
      }
      float4 e[3];
#pragma unroll
      for (int p = 0; p < 3; ++p) {
        e[p].x = nfi_bilerp(v[p][0].x, v[p][1].x, v[p][2].x, v[p][3].x, f[p]);
        e[p].y = nfi_bilerp(v[p][0].y, v[p][1].y, v[p][2].y, v[p][3].y, f[p]);
        e[p].z = nfi_bilerp(v[p][0].z, v[p][1].z, v[p][2].z, v[p][3].z, f[p]);
        e[p].w = nfi_bilerp(v[p][0].w, v[p][1].w, v[p][2].w, v[p][3].w, f[p]);
      }
      acc.x = nfi_mean3(e[0].x, e[1].x, e[2].x);
      acc.y = nfi_mean3(e[0].y, e[1].y, e[2].y);
      acc.z = nfi_mean3(e[0].z, e[1].z, e[2].z);
      acc.w = nfi_mean3(e[0].w, e[1].w, e[2].w);
    }
    *reinterpret_cast<float4*>(&t.feat[s * kFeatLd + c4 * 4]) = acc;
  }
}

__device__ __forceinline__ NfiHeads make_heads(const NfiField& f) {
  NfiHeads hd;
  hd.A = f.attention_values;
  hd.use_sdf = f.use_sdf;
  hd.beta = f.use_sdf ? f.beta[0] : 1.f;
  hd.alpha = f.use_sdf ? f.alpha[0] : 1.f;
  hd.inv_alpha = NFI_DIV(1.f, hd.alpha);
  return hd;
}

// ------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------
template <bool PTS, int MODE>
__global__ void __launch_bounds__(kTile)
field_fwd_kernel(const Args a) {
  __shared__ __align__(16) Weights w;
  __shared__ __align__(16) TileSmem t;
  load_weights(w, a.f, false);
  const NfiHeads hd = make_heads(a.f);
  const int no4 = (a.f.n_out + 3) >> 2;
  const int tid = threadIdx.x;
  for (int64_t tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
    const int64_t id = sample_id(a, tile, tid);
    float pt[3];
    __syncthreads();                       // previous tile done with t (and weights visible)
    phase1<PTS, MODE>(a, t, id, tid, pt);
    __syncthreads();
    phase2_gather(a, t);
    __syncthreads();
    if (id < 0) continue;
    const float keep = t.keep[tid];
    if (MODE == 0 && !sample_active(a, keep)) {
      reinterpret_cast<float4*>(a.out)[id] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (a.probs) for (int q = 0; q < hd.A; ++q) a.probs[id * hd.A + q] = 0.f;
      continue;
    }
    float feat[NFI_C], hpre[NFI_HID], y[NFI_YPAD];
#pragma unroll
    for (int k4 = 0; k4 < NFI_C / 4; ++k4) {
      const float4 v = *reinterpret_cast<const float4*>(&t.feat[tid * kFeatLd + k4 * 4]);
      feat[4 * k4] = v.x; feat[4 * k4 + 1] = v.y; feat[4 * k4 + 2] = v.z; feat[4 * k4 + 3] = v.w;
    }
    nfi_mlp_layer0<float4>(feat, w.W0kj, w.b0, hpre);
    nfi_mlp_layer1<float4>(hpre, w.W1jo, w.b1, no4, y);
    if (MODE == 0) {
      float sigma, rgb[3], probs[NFI_YPAD - 1];
      nfi_heads_fwd(hd, y, keep, a.palette ? a.palette + (int64_t)t.img[tid] * hd.A * 3 : nullptr,
                    sigma, rgb, probs);
      reinterpret_cast<float4*>(a.out)[id] = make_float4(sigma, rgb[0], rgb[1], rgb[2]);
      if (a.probs) {
#pragma unroll
        for (int q = 0; q < NFI_YPAD - 1; ++q) if (q < hd.A) a.probs[id * hd.A + q] = probs[q];
      }
    } else {
#pragma unroll
      for (int o = 0; o < NFI_YPAD; ++o) if (o < a.f.n_out) a.out[id * a.f.n_out + o] = y[o];
    }
  }
}

}  // namespace nfi_field
