// Warp-specialised, software-pipelined backward of the fused triplane field (one persistent CTA per
// SM, 17 warps).  Same tiling as the forward (nfi_field_fwd_ws.cuh): tile = 128 samples = an 8 x 16
// pixel block at one depth index = the 128 TMEM lanes; tile `it` lives in accumulator set it & 1.
//
//   gather / scatter (8 warps)  warp w owns 16 rows of every tile end to end: phase 1 (coords, upstream
//        gradient -> activity, texel offsets, bilinear fractions) -> gather of tile it into operand
//        stage it & 1 -> then, for tile it - 2, reads its rows of dF straight out of TMEM, scatters
//        them into the gradient planes (red.global.add.v4.f32) and forms d/d point from the re-read
//        texels.  The tables of a tile never leave the warp that wrote them.
//   MMA issuer (1 thread)       per-set state machine, polled: layer 0 (needs the operand stage and the
//        set to be free) -> layer 1 -> dH = dY . W1 -> dF = dHpre . W0; tcgen05.commit -> mbarriers.
//   epilogue (2 groups x 4)     group c serves set c, one sample row per thread: softplus (H stays in
//        TMEM) -> heads forward + backward -> dY -> dHpre = dH * sigmoid.  While a group waits for
//        its MMAs the other group computes.
//
// Barriers (c = set, m = meta slot):
//   meta_full[m]  gather -> MMA, epilogue, scatter   keep / id / activity flags of a tile written
//   meta_empty[m] epilogue group + MMA -> gather
//   a_full[c]     gather -> MMA                      operand stage c written
//   a_empty[c]    MMA (commit) -> gather             layer 0 of the tile in stage c has read it
//   acc_full[c]   MMA (commit) -> epilogue group c   next accumulator of the chain is complete
//   acc_ready[c]  epilogue group c -> MMA            next A operand (H, dY, dHpre) is in TMEM
//   df_full[c]    MMA (commit) -> scatter            dF of the tile in set c is complete
//   acc_free[c]   scatter -> MMA                     dF has been read: set c may take a new tile
#pragma once
#include "nfi_field_tc.cuh"

namespace nfi_ftc {

#ifndef NFI_BW_PS
#define NFI_BW_PS 8
#endif
constexpr int kBwPS = NFI_BW_PS;         // gather / scatter warps (8 or 16)
constexpr int kBwRows = kTile / kBwPS;   // rows of a tile owned by one of them
constexpr int kBwE = 8;                  // epilogue warps (2 groups of 4)
constexpr int kBwThreads = (kBwPS + kBwE + 1) * 32;     // 544
#ifndef NFI_BW_EROWS
#define NFI_BW_EROWS 8
#endif
#ifndef NFI_BW_RING
#define NFI_BW_RING 3
#endif
constexpr int kBwNT = NFI_BW_RING;       // table ring: a slot is free once the tile's epilogue group has scattered it
constexpr int kBwER = NFI_BW_EROWS;      // rows of dF an epilogue warp stages at a time for its atomics
constexpr int kBwNM = 8;               // ring of tables + meta: a slot is free once the epilogue group has scattered its tile
constexpr int kBwLag = 2;                // scatter runs this many tiles behind the gather
constexpr int kBwDfLd = 36;              // dF scratch row stride (floats)
constexpr int kBwRed = 32;               // per-image scalar gradients reduced per pass (dbeta, dalpha, palette)

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}

#ifdef NFI_DBG_NOPTS
constexpr bool kDbgNoPts = true;
#else
constexpr bool kDbgNoPts = false;
#endif

struct BwTable {                         // rows are private to the warp that owns them
  uint4 toff[3][kTile];                  // byte offsets of the 4 texels from the image's first plane
  float4 tder[3][kTile];                 // tx, ty, d ix / d x_a, d iy / d x_b (0 where clamped)
  float keep[kTile];                     // 1 / 0 active (in / out of box), -2 valid but idle, -1 no sample
  int sid[kTile];
};
struct BwMeta {                          // its own, deeper ring: the gather warps read the flags of tile
  int flag[kBwPS];                       // it - kBwLag while faster warps already publish tile it + 4
};
struct BwIn {                            // per warp: phase-1 inputs of its 16 rows, fetched with cp.async
  float o[16][4];                        // origin (or the explicit point), [3] = depth
  float d[16][4];                        // direction
  float4 g[16];                          // upstream gradient (heads mode)
};
struct BwSmem {
  nfi_mlp::FwdWeights w;
  nfi_mlp::BwdWeights bw;
  float a[2][2][nfi_mlp::kStageFloats];
  BwTable tab[kBwNT];
  float df[kBwPS][kBwRows * kBwDfLd];
  float dfe[kBwE][kBwER * kBwDfLd];
  BwMeta meta[kBwNM];
  float pal[kBwE][48];
  BwIn in[kBwPS];
  uint64_t a_full[2], a_empty[2], acc_full[2], acc_ready[2], df_full[2], acc_free[2];
  uint64_t meta_full[kBwNM], meta_empty[kBwNM];
  uint32_t tmem_base;
  uint32_t pad;
};

__device__ __forceinline__ int bw_any(const BwMeta& m) {
  int f = 0;
#pragma unroll
  for (int i = 0; i < kBwPS; ++i) f |= m.flag[i];
  return f;
}

// rows of gather/scatter warp w: the 16 TMEM lanes [16 (w >> 2), +16) of lane quarter w & 3, so the
// warp can read its rows of dF with tcgen05.ld (a warp only reaches the quarter warp_id % 4)
__device__ __forceinline__ int bw_row(int w, int i) { return 32 * (w & 3) + kBwRows * (w >> 2) + i; }

__device__ __forceinline__ void cp_async4(void* dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" :: "r"(nfi_tc::smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" :: "r"(nfi_tc::smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// start fetching the phase-1 inputs of one sample (row li of the warp) -- no registers held
template <bool PTS, int MODE>
__device__ __forceinline__ void bw_prefetch(const Args& a, BwIn& in, int b, int t, int row, int li, int per_image) {
  int ray;
  const int sid = sample_and_ray(a, t, row, ray);
  if (sid < 0) return;
  const int64_t id = (int64_t)b * per_image + sid;
  if (PTS) {
#pragma unroll
    for (int i = 0; i < 3; ++i) cp_async4(&in.o[li][i], a.points + id * 3 + i);
  } else {
    const int64_t r3 = ((int64_t)b * a.npi + ray) * 3;
#pragma unroll
    for (int i = 0; i < 3; ++i) { cp_async4(&in.o[li][i], a.origins + r3 + i); cp_async4(&in.d[li][i], a.dirs + r3 + i); }
    cp_async4(&in.o[li][3], a.z + id);
  }
  if (MODE == 0) cp_async16(&in.g[li], reinterpret_cast<const float4*>(a.g_out) + id);
}

// phase 1 of one sample: activity from the upstream gradient, texel offsets, fractions
template <bool PTS, int MODE>
__device__ __forceinline__ bool bw_phase1(const Args& a, BwTable& tb, BwMeta& mt, const BwIn& in, int b, int t,
                                          int row, int li, int per_image) {
  int ray;
  const int sid = sample_and_ray(a, t, row, ray);
  float keep = -1.f;
  bool act = false;
  if (sid >= 0) {
    const int64_t id = (int64_t)b * per_image + sid;
    float pt[3];
    if (PTS) { pt[0] = in.o[li][0]; pt[1] = in.o[li][1]; pt[2] = in.o[li][2]; }
    else {
      const float o[3] = {in.o[li][0], in.o[li][1], in.o[li][2]};
      const float d[3] = {in.d[li][0], in.d[li][1], in.d[li][2]};
      nfi_point(o, d, in.o[li][3], pt);
    }
    float x[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) x[i] = NFI_DIV(pt[i], a.f.scene_range);
    keep = 1.f;
    if (MODE == 0 && (fabsf(x[0]) > 1.f || fabsf(x[1]) > 1.f || fabsf(x[2]) > 1.f)) keep = 0.f;
    if (MODE == 0) {
      const float4 g = in.g[li];
      const bool col = (g.y != 0.f) || (g.z != 0.f) || (g.w != 0.f);
      // outside the box sigma is masked (generator.py:634): only a colour gradient can act;
      // samples with an all-zero upstream gradient are skipped exactly
      act = col || (g.x != 0.f && keep > 0.5f);
    } else {
#pragma unroll
      for (int o = 0; o < NFI_YPAD; ++o) if (o < a.f.n_out) act = act || (a.g_out[id * a.f.n_out + o] != 0.f);
    }
    if (!act) keep = -2.f;
    if (act) {
      const uint32_t plane_texels = (uint32_t)(a.Hp * a.Wp);
#pragma unroll
      for (int p = 0; p < 3; ++p) {
        float mx, my;
        const float ix = nfi_unnormalize(x[nfi_plane_axis_a(p)], a.Wp, &mx);
        const float iy = nfi_unnormalize(x[nfi_plane_axis_b(p)], a.Hp, &my);
        NfiFoot f;
        nfi_footprint(ix, iy, a.Wp, a.Hp, f);
        const uint32_t r0 = p * plane_texels + (uint32_t)nfi_texel_row(f.y0, a.bshift, a.nbx);
        const uint32_t r1 = p * plane_texels + (uint32_t)nfi_texel_row(f.y1, a.bshift, a.nbx);
        const uint32_t c0 = (uint32_t)nfi_texel_col(f.x0, a.bshift), c1 = (uint32_t)nfi_texel_col(f.x1, a.bshift);
        tb.toff[p][row] = make_uint4((r0 + c0) * 128u, (r0 + c1) * 128u, (r1 + c0) * 128u, (r1 + c1) * 128u);
        // ATen zeroes out-of-range neighbours (within_bounds_2d) in the coordinate gradient; that only
        // happens at ix == Wp-1 (iy == Hp-1), where the clamp already makes mx (my) zero
        tb.tder[p][row] = make_float4(f.tx, f.ty, mx, my);
      }
    }
  }
  tb.keep[row] = keep;
  tb.sid[row] = sid;
  return act;
}

__device__ __forceinline__ float4 bw_ldg(const char* base, uint32_t off) {
  return __ldg(reinterpret_cast<const float4*>(base + off));
}
__device__ __forceinline__ void bw_fma4(float4& acc, const float4& v, float w) {
  acc.x = fmaf(v.x, w, acc.x); acc.y = fmaf(v.y, w, acc.y); acc.z = fmaf(v.z, w, acc.z); acc.w = fmaf(v.w, w, acc.w);
}
__device__ __forceinline__ float4 bw_sub4(const float4& a, const float4& b) {
  return make_float4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w);
}
__device__ __forceinline__ float bw_dot4(const float4& v, const float4& g) {
  return fmaf(v.w, g.w, fmaf(v.z, g.z, fmaf(v.y, g.y, v.x * g.x)));
}
// ATen's bilinear weights from the fractions: (fx + 1) - ix and 1 - tx are the same correctly rounded
// number (tx = ix - fx is exact), so these are the forward's weights bit for bit
__device__ __forceinline__ float4 bw_weights(const float4& d) {
  const float ex = 1.f - d.x, ey = 1.f - d.y;
  return make_float4(ex * ey, d.x * ey, ex * d.y, d.x * d.y);
}

// gather of the warp's 16 rows into the operand stage
__device__ __forceinline__ void bw_gather(const Args& a, const BwTable& tb, float* a_hi, float* a_lo, int b,
                                          int w, int lane) {
  const int c4 = lane & 7;
  const float third = 1.f / 3.f;
  const char* base = reinterpret_cast<const char*>(a.planes) + (int64_t)b * 3 * a.Hp * a.Wp * 128 + c4 * 16;
#pragma unroll 2
  for (int it = 0; it < kBwRows / 4; ++it) {
    const int s = bw_row(w, it * 4 + (lane >> 3));
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (tb.keep[s] >= 0.f) {
      float4 v[3][4];
      float4 wt[3];
#pragma unroll
      for (int p = 0; p < 3; ++p) {
        const uint4 o = tb.toff[p][s];
        wt[p] = bw_weights(tb.tder[p][s]);
        v[p][0] = bw_ldg(base, o.x); v[p][1] = bw_ldg(base, o.y); v[p][2] = bw_ldg(base, o.z); v[p][3] = bw_ldg(base, o.w);
      }
#pragma unroll
      for (int p = 0; p < 3; ++p) {
        bw_fma4(acc, v[p][0], wt[p].x * third); bw_fma4(acc, v[p][1], wt[p].y * third);
        bw_fma4(acc, v[p][2], wt[p].z * third); bw_fma4(acc, v[p][3], wt[p].w * third);
      }
    }
    nfi_mlp::store_feat4(a_hi, a_lo, s, c4, acc);
  }
}

// d/d point of the warp's rows of dF (scratch `df`, row-major, kBwDfLd floats per row) from the re-read
// texels, 8 lanes per texel; `any` = the tile went through the MLP (else zeros are written)
__device__ __forceinline__ void bw_dpoint(const Args& a, const BwTable& tb, const float* df, int b, int w,
                                          int lane, int per_image, bool any) {
  const int c4 = lane & 7;
  const float third = 1.f / 3.f;
  const char* pbase = reinterpret_cast<const char*>(a.planes) + (int64_t)b * 3 * a.Hp * a.Wp * 128 + c4 * 16;
#pragma unroll 1
  for (int it = 0; it < kBwRows / 4; ++it) {
    const int sl = it * 4 + (lane >> 3);
    const int s = bw_row(w, sl);
    const float keep = tb.keep[s];
    float gp[3] = {0.f, 0.f, 0.f};
    if (any && keep >= 0.f) {
      float4 g = *reinterpret_cast<const float4*>(&df[sl * kBwDfLd + c4 * 4]);
      g.x *= third; g.y *= third; g.z *= third; g.w *= third;
      float4 d[3];
      float4 v[3][4];
#pragma unroll
      for (int p = 0; p < 3; ++p) {
        const uint4 o = tb.toff[p][s];
        d[p] = tb.tder[p][s];
        v[p][0] = bw_ldg(pbase, o.x); v[p][1] = bw_ldg(pbase, o.y);
        v[p][2] = bw_ldg(pbase, o.z); v[p][3] = bw_ldg(pbase, o.w);
      }
#pragma unroll
      for (int p = 0; p < 3; ++p) {
        // d/d ix = sum_c g_c ((v01 - v00)(1 - ty) + (v11 - v10) ty), d/d iy likewise (texel differences
        // first: dot products of the raw texels cancel to ~1e-4 relative in the pose gradient)
        const float ax = bw_dot4(bw_sub4(v[p][1], v[p][0]), g), bx = bw_dot4(bw_sub4(v[p][3], v[p][2]), g);
        const float ay = bw_dot4(bw_sub4(v[p][2], v[p][0]), g), by = bw_dot4(bw_sub4(v[p][3], v[p][1]), g);
        const float gx = fmaf(bx - ax, d[p].y, ax), gy = fmaf(by - ay, d[p].x, ay);
        gp[nfi_plane_axis_a(p)] = fmaf(gx, d[p].z, gp[nfi_plane_axis_a(p)]);
        gp[nfi_plane_axis_b(p)] = fmaf(gy, d[p].w, gp[nfi_plane_axis_b(p)]);
      }
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      gp[c] += __shfl_xor_sync(0xffffffffu, gp[c], 1);
      gp[c] += __shfl_xor_sync(0xffffffffu, gp[c], 2);
      gp[c] += __shfl_xor_sync(0xffffffffu, gp[c], 4);
    }
    if (c4 == 0 && keep != -1.f) {
      const int64_t id = (int64_t)b * per_image + tb.sid[s];
      const float inv_r = 1.f / a.f.scene_range;
      a.g_points[id * 3 + 0] = gp[0] * inv_r;
      a.g_points[id * 3 + 1] = gp[1] * inv_r;
      a.g_points[id * 3 + 2] = gp[2] * inv_r;
    }
  }
}

// plane-gradient atomics of NR staged rows of dF (scratch `df`, kBwDfLd floats per row; table rows
// row0 ..): 8 lanes per sample, each red.global.add.v4.f32 of a warp covers four whole 128-byte texels
template <int NR>
__device__ __forceinline__ void bw_scatter_rows(const Args& a, const BwTable& tb, const float* df, int b,
                                                int row0, int lane) {
  const int c4 = lane & 7;
  const float third = 1.f / 3.f;
  char* gbase = reinterpret_cast<char*>(a.g_planes) + (int64_t)b * 3 * a.Hp * a.Wp * 128 + c4 * 16;
#pragma unroll 1
  for (int it = 0; it < NR / 4; ++it) {
    const int sl = it * 4 + (lane >> 3);
    const int s = row0 + sl;
    if (tb.keep[s] >= 0.f) {
      float4 g = *reinterpret_cast<const float4*>(&df[sl * kBwDfLd + c4 * 4]);
      g.x *= third; g.y *= third; g.z *= third; g.w *= third;
#pragma unroll
      for (int p = 0; p < 3; ++p) {
        // a clamped neighbour (x1 == x0 at the border) carries weight 0: adding 0 is harmless
        const uint4 o = tb.toff[p][s];
        const float4 wt = bw_weights(tb.tder[p][s]);
        red_add_v4(reinterpret_cast<float*>(gbase + o.x), g.x * wt.x, g.y * wt.x, g.z * wt.x, g.w * wt.x);
        red_add_v4(reinterpret_cast<float*>(gbase + o.y), g.x * wt.y, g.y * wt.y, g.z * wt.y, g.w * wt.y);
        red_add_v4(reinterpret_cast<float*>(gbase + o.z), g.x * wt.z, g.y * wt.z, g.z * wt.z, g.w * wt.z);
        red_add_v4(reinterpret_cast<float*>(gbase + o.w), g.x * wt.w, g.y * wt.w, g.z * wt.w, g.w * wt.w);
      }
    }
  }
}

// sum over the 32 lanes of v[i], delivered to lane i: 31 shuffles instead of 32 x 5
__device__ __forceinline__ float bw_reduce32(float (&v)[32], int lane) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    const bool up = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < off; ++i) {
      const float send = up ? v[i] : v[i + off];
      const float mine = up ? v[i + off] : v[i];
      v[i] = mine + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
  return v[0];
}


// Developer trace (-DNFI_WS_TRACE builds only), see tools/ws_trace.py
#ifdef NFI_WS_TRACE
constexpr int kBtRoles = 5, kBtN = 128, kBtEv = 4, kBtT0 = 256;
__device__ long long g_bw_trace[kBtRoles * kBtN * kBtEv];
#define NFI_BTRACE(role, it, ev)                                                                    \
  do {                                                                                              \
    if (blockIdx.x == 0 && (it) >= (uint32_t)kBtT0 && (it) < (uint32_t)(kBtT0 + kBtN))              \
      g_bw_trace[((role) * kBtN + ((it) - kBtT0)) * kBtEv + (ev)] = clock64();                      \
  } while (0)
#else
#define NFI_BTRACE(role, it, ev) do {} while (0)
#endif

template <bool PTS, int MODE, bool WANT_PTS>
__global__ void __launch_bounds__(kBwThreads, 1)
field_bwd_ws_kernel(const Args a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  BwSmem& sm = *reinterpret_cast<BwSmem*>(smem_raw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  nfi_mlp::load_weights(sm.w, a.f.W0, a.f.b0, a.f.W1, a.f.b1, a.f.n_out);
  nfi_mlp::load_weights_bwd(sm.bw, a.f.W0, a.f.W1, a.f.n_out);
  if (tid == 0) {
    for (int c = 0; c < 2; ++c) {
      nfi_tc::mbar_init(&sm.a_full[c], kBwPS); nfi_tc::mbar_init(&sm.a_empty[c], 1);
      nfi_tc::mbar_init(&sm.acc_full[c], 1);   nfi_tc::mbar_init(&sm.acc_ready[c], kBwE / 2);
      nfi_tc::mbar_init(&sm.df_full[c], 1);    nfi_tc::mbar_init(&sm.acc_free[c], (WANT_PTS ? kBwPS : 0) + kBwE / 2);
    }
    for (int i = 0; i < kBwNM; ++i) { nfi_tc::mbar_init(&sm.meta_full[i], kBwPS); nfi_tc::mbar_init(&sm.meta_empty[i], kBwE / 2 + 1); }
    nfi_tc::mbar_fence_init();
  }
  if (warp == kBwPS + kBwE) nfi_tc::tmem_alloc(&sm.tmem_base, 2 * nfi_mlp::kTmemCols);
  nfi_tc::fence_proxy_async();
  nfi_tc::fence_before_sync();
  __syncthreads();
  nfi_tc::fence_after_sync();
  const uint32_t tmem = sm.tmem_base;
  const int per_image = (int)(a.npi * a.S);
  const TileMap tm = cta_tile_map(a.tiles_per_image);
  const int ntile = tm.ntile;
  const uint32_t total = (uint32_t)a.B * (uint32_t)ntile;

  if (warp < kBwPS) {
    // ------------------------------------------------------------------ gather / scatter
    const int w = warp;
    const int row = bw_row(w, lane & (kBwRows - 1));
    const uint32_t lane_off = (uint32_t)((w & 3) * 32) << 16;
    uint32_t df_ph[2] = {0, 0};
    BwIn& in = sm.in[w];
    if (lane < kBwRows && total > 0) bw_prefetch<PTS, MODE>(a, in, 0, tm.tile(0), row, lane, per_image);
    for (uint32_t it = 0; it < total + kBwLag; ++it) {
      if (it < total) {
        const uint32_t m = it % kBwNM, c = it & 1;
        const int b = (int)(it / (uint32_t)ntile), t = tm.tile((int)(it % (uint32_t)ntile));
        BwTable& tb = sm.tab[it % kBwNT];
        BwMeta& mt = sm.meta[m];
        if (it >= (uint32_t)kBwNT) {                       // table slot: tile it - kBwNT scattered (implies meta slot free)
          const uint32_t r = it - kBwNT;
          nfi_tc::mbar_wait(&sm.meta_empty[r % kBwNM], (r / kBwNM) & 1);
        }
        if (w == 0 && lane == 0) NFI_BTRACE(0, it, 0);
        bool act = false;
        cp_async_wait_all();
        if (lane < kBwRows) act = bw_phase1<PTS, MODE>(a, tb, mt, in, b, t, row, lane, per_image);
        const unsigned wany = __any_sync(0xffffffffu, act);
        if (lane == 0) mt.flag[w] = wany ? 1 : 0;
        __syncwarp();
        if (lane == 0) nfi_tc::mbar_arrive(&sm.meta_full[m]);
        if (lane < kBwRows && it + 1 < total) {            // inputs of the next tile arrive during gather + scatter
          const uint32_t n = it + 1;
          bw_prefetch<PTS, MODE>(a, in, (int)(n / (uint32_t)ntile), tm.tile((int)(n % (uint32_t)ntile)), row, lane, per_image);
        }
        if (w == 0 && lane == 0) NFI_BTRACE(0, it, 1);
        nfi_tc::mbar_wait(&sm.a_empty[c], ((it >> 1) & 1) ^ 1);
        if (w == 0 && lane == 0) NFI_BTRACE(0, it, 2);
        if (wany) bw_gather(a, tb, sm.a[c][0], sm.a[c][1], b, w, lane);
        nfi_tc::fence_proxy_async();
        __syncwarp();
        if (lane == 0) nfi_tc::mbar_arrive(&sm.a_full[c]);
        if (w == 0 && lane == 0) NFI_BTRACE(0, it, 3);
      }
      if (WANT_PTS && it >= (uint32_t)kBwLag) {
        // d/d point of tile it - kBwLag from its dF rows (read straight out of TMEM) and the re-read
        // texels; the plane-gradient atomics of the tile are issued by its epilogue group
        const uint32_t j = it - kBwLag, jm = j % kBwNM, c = j & 1;
        const int b = (int)(j / (uint32_t)ntile);
        const BwTable& tb = sm.tab[j % kBwNT];
        nfi_tc::mbar_wait(&sm.meta_full[jm], (j / kBwNM) & 1);       // completed long ago: flags of all warps
        const bool any = bw_any(sm.meta[jm]) != 0;
        if (w == 0 && lane == 0) NFI_BTRACE(1, j, 0);
        if (any) {
          nfi_tc::mbar_wait(&sm.df_full[c], df_ph[c]); df_ph[c] ^= 1;
          if (w == 0 && lane == 0) NFI_BTRACE(1, j, 1);
          nfi_tc::fence_after_sync();
          float v0[16], v1[16];
          const uint32_t tl = tmem + c * nfi_mlp::kTmemCols + lane_off + nfi_mlp::kColDF;
          nfi_tc::tmem_ld16(tl, v0);
          nfi_tc::tmem_ld16(tl + 16, v1);
          nfi_tc::tmem_ld_wait();
          __syncwarp();                                   // the previous tile's pass has read the scratch
          if (lane / kBwRows == (w >> 2)) {
            float* dst = &sm.df[w][(lane % kBwRows) * kBwDfLd];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              *reinterpret_cast<float4*>(dst + 4 * q) = make_float4(v0[4 * q], v0[4 * q + 1], v0[4 * q + 2], v0[4 * q + 3]);
              *reinterpret_cast<float4*>(dst + 16 + 4 * q) = make_float4(v1[4 * q], v1[4 * q + 1], v1[4 * q + 2], v1[4 * q + 3]);
            }
          }
          nfi_tc::fence_before_sync();
          __syncwarp();
          if (lane == 0) nfi_tc::mbar_arrive(&sm.acc_free[c]);
          if (w == 0 && lane == 0) NFI_BTRACE(1, j, 2);
        }
        bw_dpoint(a, tb, sm.df[w], b, w, lane, per_image, any);
        if (w == 0 && lane == 0) NFI_BTRACE(1, j, 3);
      }
    }
  } else if (warp < kBwPS + kBwE) {
    // ------------------------------------------------------------------ epilogue group c <-> set c
    const int ew = warp - kBwPS, c = ew >> 2;
    const int row = (ew & 3) * 32 + lane;
    const uint32_t tl = tmem + c * nfi_mlp::kTmemCols + ((uint32_t)((ew & 3) * 32) << 16);
    const NfiHeads hd = make_heads(a.f);
    const int npal = 3 * hd.A;
    const bool want_pal = MODE == 0 && a.g_palette != nullptr && hd.A > 0;
    const bool want_par = MODE == 0 && a.g_params != nullptr && hd.use_sdf;
    const bool two_pass = 2 + npal > kBwRed;
    float* pal = sm.pal[ew];
    float acc0 = 0.f, acc1 = 0.f;                  // per-lane sums of reduction slots lane, 32 + lane
    int cur_b = -1;
    uint32_t ph = 0, dfph = 0;
    auto flush = [&]() {
      if (cur_b < 0) return;
      const int pbase = NFI_HID * NFI_C + NFI_HID + a.f.n_out * NFI_HID + a.f.n_out;
      if (lane < 2) { if (want_par && acc0 != 0.f) atomicAdd(&a.g_params[pbase + lane], acc0); }
      else if (want_pal && lane - 2 < npal && acc0 != 0.f) atomicAdd(&a.g_palette[(int64_t)cur_b * npal + lane - 2], acc0);
      if (want_pal && kBwRed - 2 + lane < npal && acc1 != 0.f) atomicAdd(&a.g_palette[(int64_t)cur_b * npal + kBwRed - 2 + lane], acc1);
      acc0 = acc1 = 0.f;
    };
    for (uint32_t it = c; it < total; it += 2) {
      const uint32_t m = it % kBwNM;
      const BwMeta& mt = sm.meta[m];
      const int b = (int)(it / (uint32_t)ntile);
      nfi_tc::mbar_wait(&sm.meta_full[m], (it / kBwNM) & 1);
      const BwTable& tb = sm.tab[it % kBwNT];
      const int sid = tb.sid[row];
      const float keep = tb.keep[row];
      const bool any = bw_any(mt) != 0;
      if (any) {
        if (b != cur_b) {
          flush();
          cur_b = b;
          if (MODE == 0 && a.palette) {
            __syncwarp();
            for (int i = lane; i < npal; i += 32) pal[i] = a.palette[(int64_t)b * npal + i];
            __syncwarp();
          }
        }
        const bool act = keep >= 0.f;
        const int64_t id = (int64_t)b * per_image + sid;
        float gy[NFI_YPAD];
#pragma unroll
        for (int o = 0; o < NFI_YPAD; ++o) gy[o] = 0.f;
        if (act) {
          if (MODE == 0) {
            const float4 g = __ldg(reinterpret_cast<const float4*>(a.g_out) + id);
            gy[0] = g.x; gy[1] = g.y; gy[2] = g.z; gy[3] = g.w;
          } else {
#pragma unroll
            for (int o = 0; o < NFI_YPAD; ++o) if (o < a.f.n_out) gy[o] = a.g_out[id * a.f.n_out + o];
          }
        }
        // ---- layer 0 -> H
        nfi_tc::mbar_wait(&sm.acc_full[c], ph); ph ^= 1;
        if ((ew & 3) == 0 && lane == 0) NFI_BTRACE(2, it, 0);
        nfi_tc::fence_after_sync();
        nfi_mlp::epilogue0_at<4>(sm.w.b0, tl, 0);
        nfi_tc::fence_before_sync();
        __syncwarp();
        if (lane == 0) nfi_tc::mbar_arrive(&sm.acc_ready[c]);
        if ((ew & 3) == 0 && lane == 0) NFI_BTRACE(2, it, 1);
        // ---- layer 1 -> heads forward + backward -> dY
        nfi_tc::mbar_wait(&sm.acc_full[c], ph); ph ^= 1;
        if ((ew & 3) == 0 && lane == 0) NFI_BTRACE(2, it, 2);
        nfi_tc::fence_after_sync();
        float dy[NFI_YPAD];
#pragma unroll
        for (int o = 0; o < NFI_YPAD; ++o) dy[o] = 0.f;
        float probs[NFI_YPAD - 1];                  // zero for idle rows: they add nothing below
#pragma unroll
        for (int q = 0; q < NFI_YPAD - 1; ++q) probs[q] = 0.f;
        const float drgb[3] = {gy[1], gy[2], gy[3]};
        float dbeta = 0.f, dalpha = 0.f;
        if (MODE == 0) {
          float y[nfi_mlp::kY];
          nfi_mlp::epilogue1_at(sm.w.b1, tl, y);
          if (act) {
            const float k01 = keep > 0.5f ? 1.f : 0.f;
            float sigma, rgb[3];
            nfi_heads_fwd(hd, y, k01, a.palette ? pal : nullptr, sigma, rgb, probs);
            nfi_heads_bwd(hd, y, k01, a.palette ? pal : nullptr, probs, rgb, gy[0], drgb, dy,
                          want_par ? &dbeta : nullptr, want_par ? &dalpha : nullptr);
          }
        } else {
#pragma unroll
          for (int o = 0; o < NFI_YPAD; ++o) dy[o] = gy[o];
        }
        nfi_mlp::store_dy(tl, dy);
        nfi_tc::fence_before_sync();
        __syncwarp();
        if (lane == 0) nfi_tc::mbar_arrive(&sm.acc_ready[c]);
        if ((ew & 3) == 0 && lane == 0) NFI_BTRACE(2, it, 3);
        if (want_pal || want_par) {                 // while the dH MMAs run
          // reduction slots: 0 dbeta, 1 dalpha, 2 + 3 q + k  d palette[q][k] = probs[q] * drgb[k]
          float red[kBwRed];
          red[0] = dbeta; red[1] = dalpha;
#pragma unroll
          for (int i = 2; i < kBwRed; ++i) red[i] = probs[(i - 2) / 3] * drgb[(i - 2) % 3];
          acc0 += bw_reduce32(red, lane);
          if (two_pass && want_pal) {
#pragma unroll
            for (int i = 0; i < kBwRed; ++i) {
              const int s = kBwRed - 2 + i;
              red[i] = s < 3 * (NFI_YPAD - 1) ? probs[s / 3 < NFI_YPAD - 1 ? s / 3 : 0] * drgb[s % 3] : 0.f;
            }
            acc1 += bw_reduce32(red, lane);
          }
        }
        // ---- dH -> dHpre
        nfi_tc::mbar_wait(&sm.acc_full[c], ph); ph ^= 1;
        if ((ew & 3) == 0 && lane == 0) NFI_BTRACE(3, it, 0);
        nfi_tc::fence_after_sync();
        nfi_mlp::epilogue2<4>(tl, 0);
        nfi_tc::fence_before_sync();
        __syncwarp();
        if (lane == 0) nfi_tc::mbar_arrive(&sm.acc_ready[c]);
        if ((ew & 3) == 0 && lane == 0) NFI_BTRACE(3, it, 1);
        // ---- dF -> plane gradients: this thread's row of dF, 8 x 16-byte atomics per texel
        nfi_tc::mbar_wait(&sm.df_full[c], dfph); dfph ^= 1;
        nfi_tc::fence_after_sync();
        float g0[16], g1[16];
        nfi_tc::tmem_ld16(tl + nfi_mlp::kColDF, g0);
        nfi_tc::tmem_ld16(tl + nfi_mlp::kColDF + 16, g1);
        nfi_tc::tmem_ld_wait();
        nfi_tc::fence_before_sync();
        __syncwarp();
        if (lane == 0) nfi_tc::mbar_arrive(&sm.acc_free[c]);      // the set may take its next tile
        if ((ew & 3) == 0 && lane == 0) NFI_BTRACE(3, it, 2);
        if (a.g_planes) {
          float* sc = sm.dfe[ew];
#pragma unroll 1
          for (int h = 0; h < 32 / kBwER; ++h) {          // kBwER rows through the warp's scratch at a time
            __syncwarp();
            if (lane / kBwER == h) {
              float* dst = sc + (lane % kBwER) * kBwDfLd;
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                *reinterpret_cast<float4*>(dst + 4 * q) = make_float4(g0[4 * q], g0[4 * q + 1], g0[4 * q + 2], g0[4 * q + 3]);
                *reinterpret_cast<float4*>(dst + 16 + 4 * q) = make_float4(g1[4 * q], g1[4 * q + 1], g1[4 * q + 2], g1[4 * q + 3]);
              }
            }
            __syncwarp();
            bw_scatter_rows<kBwER>(a, tb, sc, b, (ew & 3) * 32 + h * kBwER, lane);
          }
        }
        if ((ew & 3) == 0 && lane == 0) NFI_BTRACE(3, it, 3);
      }
      __syncwarp();
      if (lane == 0) nfi_tc::mbar_arrive(&sm.meta_empty[m]);
    }
    flush();
  } else {
    // ------------------------------------------------------------------ MMA issuer: per-set state machine
    // the whole warp polls (converged, uniform state); one elected lane issues MMAs, commits, arrives
    uint32_t it0 = 0, it1 = 1, stage0 = 0, stage1 = 0, rph0 = 0, rph1 = 0, nact0 = 0, nact1 = 0;
    uint32_t idle = 0;
    auto step = [&](const int c, uint32_t& it, uint32_t& stage, uint32_t& rph, uint32_t& nact) -> bool {
      if (it >= total) return false;
      const uint32_t tacc = tmem + c * nfi_mlp::kTmemCols;
      if (stage == 0) {
        const uint32_t m = it % kBwNM;
        if (!nfi_tc::mbar_test(&sm.meta_full[m], (it / kBwNM) & 1)) return false;
        if (!nfi_tc::mbar_test(&sm.a_full[c], (it >> 1) & 1)) return false;
        const int any = bw_any(sm.meta[m]);
        if (any && nact > 0 && !nfi_tc::mbar_test(&sm.acc_free[c], (nact - 1) & 1)) return false;
        __syncwarp();
        if (!any) {
          if (nfi_tc::elect_one()) { nfi_tc::mbar_arrive(&sm.meta_empty[m]); nfi_tc::mbar_arrive(&sm.a_empty[c]); }
          it += 2;
          return true;
        }
        nfi_tc::fence_after_sync();
        if (nfi_tc::elect_one()) {
          nfi_tc::mbar_arrive(&sm.meta_empty[m]);
          nfi_mlp::issue_layer0_at(sm.w, sm.a[c][0], sm.a[c][1], tacc);
          nfi_tc::mma_commit(&sm.a_empty[c]);
          nfi_tc::mma_commit(&sm.acc_full[c]);
        }
        __syncwarp();
        if (lane == 0) NFI_BTRACE(4, it, 0);
        ++nact; stage = 1;
        return true;
      }
      if (!nfi_tc::mbar_test(&sm.acc_ready[c], rph)) return false;
      rph ^= 1;
      nfi_tc::fence_after_sync();
      if (stage == 1) {
        if (nfi_tc::elect_one()) { nfi_mlp::issue_layer1_at(sm.w, tacc); nfi_tc::mma_commit(&sm.acc_full[c]); }
        __syncwarp();
        if (lane == 0) NFI_BTRACE(4, it, 1);
        stage = 2;
      } else if (stage == 2) {
        if (nfi_tc::elect_one()) { nfi_mlp::issue_dh(sm.bw, tacc); nfi_tc::mma_commit(&sm.acc_full[c]); }
        __syncwarp();
        if (lane == 0) NFI_BTRACE(4, it, 2);
        stage = 3;
      } else {
        if (nfi_tc::elect_one()) { nfi_mlp::issue_df(sm.bw, tacc); nfi_tc::mma_commit(&sm.df_full[c]); }
        __syncwarp();
        if (lane == 0) NFI_BTRACE(4, it, 3);
        stage = 0; it += 2;
      }
      return true;
    };
    while (it0 < total || it1 < total) {
      const bool p0 = step(0, it0, stage0, rph0, nact0);
      const bool p1 = step(1, it1, stage1, rph1, nact1);
      if (p0 || p1) idle = 0;
      else if (++idle > (1u << 26)) __trap();
    }
  }
  nfi_tc::fence_before_sync();
  __syncthreads();
  if (warp == kBwPS + kBwE) nfi_tc::tmem_dealloc(tmem, 2 * nfi_mlp::kTmemCols);
}

}  // namespace nfi_ftc
