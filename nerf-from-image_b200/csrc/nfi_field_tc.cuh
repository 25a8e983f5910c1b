// A4-A6 fused, tensor-core version: bilinear triplane gather -> decoder MLP on tcgen05 (3xTF32,
// accumulators in TMEM) -> density / colour heads (models/generator.py:288-331, 587-681).
// Features, hidden activations and MLP outputs never touch HBM.
//
// CTA = 256 threads, tile = 128 samples (= the 128 TMEM lanes).  Per tile:
//   phase 1  thread-per-sample : point -> box coords -> per plane 4 texel offsets + 4 bilinear weights
//            (shared memory tables, so the 8 lanes of a texel do not redo the index math)
//   phase 2  8 lanes per texel : each lane owns 4 of the 32 channels (one float4 = 16 B): a texel is
//            one fully-used 128-B line; the 12-texel weighted sum lands in the UMMA A-operand tile
//            (hi / lo tf32 parts, K-major core-matrix layout, see nfi_tc.cuh)
//   layer 0  one thread issues 12 tcgen05.mma (M128 N64 K8); completion through an mbarrier
//   epi 0    all 8 warps: tcgen05.ld D0, + b0, softplus, split hi/lo, tcgen05.st (next A operand)
//   layer 1  24 tcgen05.mma (M128 N16 K8, A from TMEM)
//   epi 1    warps 0-3: tcgen05.ld y[16], heads, one float4 store per sample
// Tile order: an 8 x 16 pixel block at one depth step; a CTA walks consecutive depth steps of the
// same block (texel reuse in L1), and all CTAs work on the same image at a time (planes stay in L2).
#pragma once
#include "nfi_common.cuh"
#include "nfi_field_math.cuh"
#include "nfi_mlp_tc.cuh"

namespace nfi_ftc {

constexpr int kTile = 128;
constexpr int kThreads = 256;

struct Args {
  NfiField f;
  const float* planes;    // [B,3,Hp*Wp,32], texel order of nfi_field_math.cuh (nfi_texel_row / _col)
  const float* palette;   // [B,A,3]
  int B, Hp, Wp;
  int bshift, nbx;        // texel order: block shift (3 or 0) and blocks per plane row
  const float* origins; const float* dirs; const float* z; const float* points;
  int64_t npi;            // rays (or points) per image
  int S;
  int img_w;              // > 0: pixel-block tile order is allowed
  int skip_oob;           // fused render: out-of-box samples produce (0,0,0,0) without gathering
  int64_t n_total;        // B * npi * S samples
  int64_t tiles_per_image;
  int blocks_x;           // pixel-block order: blocks per image row (0 = linear order)
  float inv_S, inv_bx;    // 1 / S, 1 / blocks_x for fast_div (tile -> pixel block / depth step without IDIV sequences)
  float* out; float* probs;
  // backward
  const float* g_out;
  float* g_planes; float* g_palette; float* g_points; float* g_params;
  // generator-training path (nfi_field_bwd_train): per-sample rows the backward leaves in a caller workspace for the
  // weight-gradient reduction -- F [n,32], H [n,64], dHpre [n,64], dY [n,16], and a validity byte per sample
  float* st_f; float* st_h; float* st_dh; float* st_dy; unsigned char* st_ok;
};

// One 16-byte feature chunk (4 channels) of a texel.  `off` is the texel's byte offset in the fp32 layout
// (128 bytes per texel); a bf16 plane copy (NfiField.planes_bf16, 64 bytes per texel, same texel order)
// is read with half the offset and an 8-byte load, widened exactly (bf16 -> fp32 is a 16-bit shift).
template <bool BF16>
__device__ __forceinline__ float4 ldg_texel4(const char* base, uint32_t off) {
  if (BF16) {
    const uint2 r = __ldg(reinterpret_cast<const uint2*>(base + (off >> 1)));
    return make_float4(__uint_as_float(r.x << 16), __uint_as_float(r.x & 0xFFFF0000u),
                       __uint_as_float(r.y << 16), __uint_as_float(r.y & 0xFFFF0000u));
  }
  return __ldg(reinterpret_cast<const float4*>(base + off));
}

// n / d for 0 <= n < 2^24, d >= 1, inv = 1.f / d: the float product is within one of the quotient (n is
// exact in fp32), one correction step makes it exact.  Replaces the ~20-instruction integer division
// sequences that ran once per row and tile in every producer warp.
__device__ __forceinline__ int fast_div(int n, int d, float inv) {
  int q = __float2int_rz(__int2float_rn(n) * inv);
  const int r = n - q * d;
  q += (r >= d) ? 1 : 0;
  q -= (r < 0) ? 1 : 0;
  return q;
}

// (tile-in-image, thread) -> sample index WITHIN the image (ray_in_image * S + k), or -1.
// 32-bit arithmetic: fill_args guarantees npi * S < 2^31 (so tile indices are < 2^24).
__device__ __forceinline__ int sample_in_image(const Args& a, int t, int tid) {
  if (a.blocks_x == 0) {
    const int64_t local = (int64_t)t * kTile + tid;
    return local < a.npi * a.S ? (int)local : -1;
  }
  const int blk = fast_div(t, a.S, a.inv_S);   // depth step fastest: consecutive tiles share texels
  const int k = t - blk * a.S;
  const int by = fast_div(blk, a.blocks_x, a.inv_bx), bx = blk - by * a.blocks_x;
  const int py = tid >> 4, px = tid & 15;
  const int ray = (by * 8 + py) * a.img_w + bx * 16 + px;
  return ray * a.S + k;
}

// same mapping, also returning the ray index within the image
__device__ __forceinline__ int sample_and_ray(const Args& a, int t, int tid, int& ray) {
  if (a.blocks_x == 0) {
    const int64_t local = (int64_t)t * kTile + tid;
    if (local >= a.npi * a.S) { ray = 0; return -1; }
    ray = (int)local / a.S;
    return (int)local;
  }
  const int blk = fast_div(t, a.S, a.inv_S);
  const int k = t - blk * a.S;
  const int by = fast_div(blk, a.blocks_x, a.inv_bx), bx = blk - by * a.blocks_x;
  const int py = tid >> 4, px = tid & 15;
  ray = (by * 8 + py) * a.img_w + bx * 16 + px;
  return ray * a.S + k;
}

struct Tables {
  int4 toff[3][kTile];      // texel offsets in float4 units (texel index * 8), per plane: 00, 01, 10, 11
  float4 twgt[3][kTile];    // ATen's weights nw, ne, sw, se
  float4 tder[3][kTile];    // backward only: (tx, ty, d ix / d x_a, d iy / d x_b) -- the last two are 0 where clamped
  float keep[kTile];        // 1 inside box, 0 outside, -1 no sample, -2 valid but nothing to do
  int sid[kTile];           // sample index within the image (-1: none)
};

__device__ __forceinline__ bool sample_active(const Args& a, float keep) {
  return keep > 0.5f || (keep == 0.f && !a.skip_oob);
}

// phase 1.  MODE 0: heads (out-of-box mask applies); MODE 1: raw decoder.  DER: fill tder.
template <bool PTS, int MODE, bool DER>
__device__ __forceinline__ void phase1(const Args& a, Tables& t, int64_t id, int sid, int tid) {
  float keep = -1.f;
  if (id >= 0) {
    float pt[3];
    if (PTS) { pt[0] = a.points[id * 3]; pt[1] = a.points[id * 3 + 1]; pt[2] = a.points[id * 3 + 2]; }
    else {
      const int64_t ray = id / a.S;
      const float o[3] = {a.origins[ray * 3], a.origins[ray * 3 + 1], a.origins[ray * 3 + 2]};
      const float d[3] = {a.dirs[ray * 3], a.dirs[ray * 3 + 1], a.dirs[ray * 3 + 2]};
      nfi_point(o, d, a.z[id], pt);
    }
    float x[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) x[c] = NFI_DIV(pt[c], a.f.scene_range);
    keep = 1.f;
    if (MODE == 0 && (fabsf(x[0]) > 1.f || fabsf(x[1]) > 1.f || fabsf(x[2]) > 1.f)) keep = 0.f;
    if (sample_active(a, keep)) {
#pragma unroll
      for (int p = 0; p < 3; ++p) {
        float mx, my;
        const float ix = nfi_unnormalize(x[nfi_plane_axis_a(p)], a.Wp, &mx);
        const float iy = nfi_unnormalize(x[nfi_plane_axis_b(p)], a.Hp, &my);
        NfiFoot f;
        nfi_footprint(ix, iy, a.Wp, a.Hp, f);
        const int r0 = nfi_texel_row(f.y0, a.bshift, a.nbx), r1 = nfi_texel_row(f.y1, a.bshift, a.nbx);
        const int c0 = nfi_texel_col(f.x0, a.bshift), c1 = nfi_texel_col(f.x1, a.bshift);
        t.toff[p][tid] = make_int4((r0 + c0) * 8, (r0 + c1) * 8, (r1 + c0) * 8, (r1 + c1) * 8);
        t.twgt[p][tid] = make_float4(f.wnw, f.wne, f.wsw, f.wse);
        if (DER) {
          // ATen zeroes out-of-range neighbours (within_bounds_2d) in the coordinate gradient; that
          // only happens at ix == Wp-1 (iy == Hp-1), where the clamp already makes mx (my) zero
          t.tder[p][tid] = make_float4(f.tx, f.ty, mx, my);
        }
      }
    }
  }
  t.keep[tid] = keep;
  t.sid[tid] = sid;
}

// phase 2: 256 threads gather 128 samples, 4 per warp pass, into the A-operand tile.
__device__ __forceinline__ void phase2_gather(const Args& a, const Tables& t, nfi_mlp::Operands& op, int b) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c4 = lane & 7;
  const float third = 1.f / 3.f;
  const int plane_f4 = a.Hp * a.Wp * 8;
  const float4* base = reinterpret_cast<const float4*>(a.planes) + (int64_t)b * 3 * plane_f4 + c4;
#pragma unroll 2
  for (int it = 0; it < 4; ++it) {
    const int s = warp * 16 + it * 4 + (lane >> 3);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (sample_active(a, t.keep[s])) {
      float4 v[3][4];
      float4 w[3];
#pragma unroll
      for (int p = 0; p < 3; ++p) {
        const int4 o = t.toff[p][s];
        w[p] = t.twgt[p][s];
        const float4* P = base + p * plane_f4;
        v[p][0] = __ldg(P + o.x); v[p][1] = __ldg(P + o.y); v[p][2] = __ldg(P + o.z); v[p][3] = __ldg(P + o.w);
      }
#pragma unroll
      for (int p = 0; p < 3; ++p) {
        // ATen's order: nw*w + ne*w + sw*w + se*w accumulated left to right
        float4 e;
        e.x = fmaf(v[p][3].x, w[p].w, fmaf(v[p][2].x, w[p].z, fmaf(v[p][1].x, w[p].y, v[p][0].x * w[p].x)));
        e.y = fmaf(v[p][3].y, w[p].w, fmaf(v[p][2].y, w[p].z, fmaf(v[p][1].y, w[p].y, v[p][0].y * w[p].x)));
        e.z = fmaf(v[p][3].z, w[p].w, fmaf(v[p][2].z, w[p].z, fmaf(v[p][1].z, w[p].y, v[p][0].z * w[p].x)));
        e.w = fmaf(v[p][3].w, w[p].w, fmaf(v[p][2].w, w[p].z, fmaf(v[p][1].w, w[p].y, v[p][0].w * w[p].x)));
        acc.x += e.x; acc.y += e.y; acc.z += e.z; acc.w += e.w;
      }
      acc.x *= third; acc.y *= third; acc.z *= third; acc.w *= third;   // (e1 + e2 + e3) / 3
    }
    nfi_mlp::store_feat4(op, s, c4, acc);
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

// Interleaved static partition of one image's tiles over the grid (warp-specialised kernels): tiles
// are dealt out in chunks of kChunk consecutive tiles (= depth steps of one pixel block: L1 reuse
// inside a chunk), chunk q to CTA q % grid -- every CTA gets pixel blocks from all over the image,
// so empty (out-of-box) regions and dense ones are shared evenly.
constexpr int kChunk = 8;
struct TileMap {
  int ntile;      // tiles of this CTA per image
  int stride;     // grid size in tiles (gridDim.x * kChunk)
  int first;      // first tile of this CTA
  __device__ __forceinline__ int tile(int l) const { return first + (l / kChunk) * stride + (l % kChunk); }
};
__device__ __forceinline__ TileMap cta_tile_map(int64_t tiles) {
  TileMap m;
  const int nchunks = (int)((tiles + kChunk - 1) / kChunk);
  const int c = blockIdx.x, g = gridDim.x;
  const int mine = c < nchunks ? (nchunks - c + g - 1) / g : 0;
  m.stride = g * kChunk;
  m.first = c * kChunk;
  m.ntile = 0;
  if (mine > 0) {
    const int last = c + (mine - 1) * g;                     // only the image's last chunk can be short
    const int64_t left = tiles - (int64_t)last * kChunk;
    m.ntile = (mine - 1) * kChunk + (int)(left < kChunk ? left : kChunk);
  }
  return m;
}

// static partition of one image's tiles over the grid: CTA c gets [lo, hi)
__device__ __forceinline__ void cta_tile_range(int64_t tiles, int64_t& lo, int64_t& hi) {
  lo = tiles * blockIdx.x / gridDim.x;
  hi = tiles * (blockIdx.x + 1) / gridDim.x;
}

}  // namespace nfi_ftc
