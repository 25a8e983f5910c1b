// Warp-specialised, software-pipelined backward of the fused triplane field (one persistent CTA per
// SM, 17 warps).  Same tiling as the forward (nfi_field_fwd_ws.cuh): tile = 128 samples = an 8 x 16
// pixel block at one depth index = the 128 TMEM lanes.
//
// Every texel line is fetched from L2 exactly once per sample: the gather that recomputes the features
// F also forms, from the same registers, the tangents T_k = dF / dx_k (k = 0..2, 32 channels each) and
// parks them in TMEM until the feature gradient dF of the tile comes out of the MLP chain; the point
// gradient is then three 32-term dot products per sample, taken row-per-thread straight from TMEM.
//
//   gather (8 warps)     warp w owns 16 rows of every tile: phase 1 (coords, upstream gradient ->
//        activity, texel offsets, bilinear fractions) -> gather -> F into operand stage it & 1 (shared
//        memory, hi / lo), T through a per-quarter transposition scratch into TMEM buffer it & 1.
//   MMA issuer (1 warp, one elected lane)  layer 0 -> layer 1 -> dH = dY . W1 -> dF = dHpre . W0, each
//        after the epilogue stage before it; tcgen05.commit -> mbarriers.  One accumulator set.
//   epilogue (4 warps)   one sample row per thread: softplus (H stays in TMEM) -> heads forward +
//        backward -> dY -> dHpre = dH * sigmoid; per-image palette / beta / alpha sums.
//   scatter (4 warps)    one sample row per thread: dF and T from TMEM -> d/d point; then the rows of
//        dF go through shared memory 8 at a time so that 8 lanes cover one texel line of the gradient
//        planes with red.global.add.v4.f32.  Runs one tile behind the epilogue warps.
//
// TMEM (512 columns): [0,240) the chain (nfi_mlp_tc.cuh) | [240,432) T, two buffers of 96 |
//                     [432,496) dF, two buffers of 32
//
// Barriers (c = it & 1, m = flag slot):
//   meta_full[m]  gather -> everyone                 tables + activity flags of a tile written
//   meta_empty[m] epilogue + scatter + MMA -> gather tile finished: its table slot may be rewritten
//   a_full[c]     gather -> MMA                      F stage c and T buffer c written (or tile idle)
//   a_empty       MMA (commit) -> gather             layer 0 has read the F stage
//   acc_full      MMA (commit) -> epilogue           next accumulator of the chain is complete
//   acc_ready     epilogue -> MMA                    next A operand (H, dY, dHpre) is in TMEM
//   df_full[c]    MMA (commit) -> scatter            dF buffer c is complete
//   s_free[c]     scatter -> gather                  T and dF buffers c have been read (every tile)
//
// Phase discipline (audited in round 2 after the round-1 launch failure).  A parity wait is sound only
// while the waiter is at most one phase behind the barrier, so for every barrier the arrival that
// completes phase k + 1 must be ordered after every waiter's test of phase k:
//   barrier       arrives (per tile)   waits                     why phase k + 1 cannot complete before the test of k
//   meta_full[m]  gather, every tile   all roles, every tile     meta_full(it + 8) needs meta_empty(it + 5): every role past it
//   meta_empty[m] E + S + MMA, every   gather, every tile >= 3   phase of tile r + 8 needs the gather's own meta_full(r + 8)
//   a_full[c]     gather, every tile   MMA and scatter, EVERY    gather(it + 2) first waits a_empty(it) [MMA past its test]
//                                      tile                      and s_free(it) [all scatter warps past their test]
//   a_empty[c]    MMA, every tile      gather, every tile        a_empty(it) needs a_full(it), posted after the gather's test
//   s_free[c]     scatter, every tile  gather, EVERY tile >= 2   scatter(it) arrives after a_full(it), posted after the test
//   df_full[c]    MMA, active tiles    scatter, active tiles     self-counted; dF(it + 2k) needs a_full <- s_free(it) or meta_empty
//   acc_full / acc_ready               strict ping-pong MMA <-> epilogue, self-counted phases
// Round 1 waited for a_full[c] (scatter) and s_free[c] (gather) on ACTIVE tiles only.  A gather running
// through an idle tile it + 2 then skipped s_free(it) and could complete phase (it >> 1) + 1 of a_full[c]
// before a delayed scatter warp had tested phase it >> 1 for the active tile it; that warp then waited for a
// parity only tile it + 4 could produce, and tile it + 4 needs meta_empty(it + 1) from the same warp:
// deadlock -> bounded wait -> trap ("unspecified launch failure" after ~50 s).  Making only
// the gather's wait unconditional would open the mirror-image race (scatter arriving on s_free for an
// idle tile before a slow gather warp has tested the previous phase), so BOTH waits are unconditional.
#pragma once
#include "nfi_f2.cuh"
#include "nfi_field_tc.cuh"

namespace nfi_ftc {

constexpr int kBwPS = 8;                 // gather warps
constexpr int kBwRows = kTile / kBwPS;   // rows of a tile owned by one of them
constexpr int kBwE = 4;                  // epilogue warps (warp_id % 4 = TMEM lane quarter)
constexpr int kBwS = 4;                  // scatter warps (likewise)
constexpr int kBwThreads = (kBwPS + kBwE + kBwS + 1) * 32;     // 544
#ifndef NFI_BW_EROWS
#define NFI_BW_EROWS 8
#endif
#ifndef NFI_BW_RING
#define NFI_BW_RING 3
#endif
constexpr int kBwNT = NFI_BW_RING;       // table ring: a slot is free once the tile has been scattered
constexpr int kBwER = NFI_BW_EROWS;      // rows of dF a scatter warp stages at a time for its atomics
#ifndef NFI_BW_ASTAGES
#define NFI_BW_ASTAGES 2
#endif
constexpr int kBwNA = NFI_BW_ASTAGES;    // F operand stages.  Measured (round 1k): one stage would free 33 KB of shared
                                         // memory for L1, but couples the gather to the MMA chain (layer 0 of tile it is
                                         // only issued after the chain of tile it - 1): 7.8 ms per launch against 6.9
constexpr int kBwNM = 8;                 // ring of activity flags (deeper: read by warps that lag)
constexpr int kBwDfLd = 36;              // dF scratch row stride (floats)
constexpr int kBwTLd = 100;              // T scratch row stride (floats): 96 + 4, conflict-free both ways
constexpr int kBwRed = 32;               // per-image scalar gradients reduced per pass (dbeta, dalpha, palette)
constexpr int kBwColT = 240, kBwColDF = 432, kBwTmemCols = 512;

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
#ifdef NFI_EXP_STSCATTER   // developer A/B (tools/ab_build.py): plain stores, wrong results, timing only
  asm volatile("st.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
#else
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
#endif
}

struct BwTable {                         // rows are written by the gather warp that owns them
  uint4 toff[3][kTile];                  // byte offsets of the 4 texels from the image's first plane
  float4 tder[3][kTile];                 // tx, ty, d ix / d x_a, d iy / d x_b (0 where clamped)
  float keep[kTile];                     // 1 / 0 active (in / out of box), -2 valid but idle, -1 no sample
  int sid[kTile];
};
struct BwMeta {
  int flag[kBwPS];
};
struct BwIn {                            // per warp: phase-1 inputs of its 16 rows, fetched with cp.async
  float o[16][4];                        // origin (or the explicit point), [3] = depth
  float d[16][4];                        // direction
  float4 g[16];                          // upstream gradient (heads mode)
};
struct BwSmem {
  nfi_mlp::FwdWeights w;
  nfi_mlp::BwdWeights bw;
  float a[kBwNA][2][nfi_mlp::kStageFloats];
  BwTable tab[kBwNT];
  float tq[4][32 * kBwTLd];              // per lane quarter: T rows of the tile being gathered
  float dfe[kBwS][kBwER * kBwDfLd];
  BwMeta meta[kBwNM];
  float pal[kBwE][48];
  BwIn in[kBwPS];
  uint64_t a_full[2], a_empty[kBwNA], df_full[2], s_free[2], acc_full, acc_ready;
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

// rows of gather warp w: the 16 TMEM lanes [16 (w >> 2), +16) of lane quarter w & 3 (warps w and
// w + 4 share a quarter: together they can write its 32 lanes with tcgen05.st)
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
__device__ __forceinline__ bool bw_phase1(const Args& a, BwTable& tb, const BwIn& in, int b, int t,
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

template <bool BF16>
__device__ __forceinline__ float4 bw_ldg(const char* base, uint32_t off) {
#ifdef NFI_EXP_NOGATHER    // developer A/B: every sample reads texel 0 (L1 hits), timing only
  off = 0;
#endif
  return ldg_texel4<BF16>(base, off);
}
__device__ __forceinline__ void bw_fma4(float4& acc, const float4& v, float w) {
  acc.x = fmaf(v.x, w, acc.x); acc.y = fmaf(v.y, w, acc.y); acc.z = fmaf(v.z, w, acc.z); acc.w = fmaf(v.w, w, acc.w);
}
__device__ __forceinline__ float4 bw_sub4(const float4& a, const float4& b) {
  return make_float4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w);
}
// a + (b - a) t
__device__ __forceinline__ float4 bw_lerp4(const float4& a, const float4& b, float t) {
  return make_float4(fmaf(b.x - a.x, t, a.x), fmaf(b.y - a.y, t, a.y), fmaf(b.z - a.z, t, a.z), fmaf(b.w - a.w, t, a.w));
}
// ATen's bilinear weights from the fractions: (fx + 1) - ix and 1 - tx are the same correctly rounded
// number (tx = ix - fx is exact), so these are the forward's weights bit for bit
__device__ __forceinline__ float4 bw_weights(const float4& d) {
  const float ex = 1.f - d.x, ey = 1.f - d.y;
  return make_float4(ex * ey, d.x * ey, ex * d.y, d.x * d.y);
}

// gather of the warp's 16 rows: features into the operand stage and, from the same texels, the
// tangents T_k = dF / dx_k (in units of the normalised coordinate) into the quarter's scratch `tq`
// (row = TMEM lane within the quarter, kBwTLd floats per row, T_k at [32 k, 32 k + 32)).
// Texel order: 0 (x0,y0)  1 (x1,y0)  2 (x0,y1)  3 (x1,y1).
//   dF/d ix = (v1 - v0) + ((v3 - v2) - (v1 - v0)) ty      dF/d iy = (v2 - v0) + ((v3 - v1) - (v2 - v0)) tx
// (texel differences first: sums of the raw texels cancel to ~1e-4 relative in the pose gradient)
// Round-2 A/B (config 2, ms per launch, same box): this scalar form 6.78; the same operations as packed
// FADD2 / FFMA2 pairs (nfi_f2.cuh) 7.18; the factored form F = v0 + tx dF/dix + ty (v2 - v0), packed or
// scalar, 7.24 / 7.47 -- both need more live registers, and at the 96-register cap of this 17-warp CTA
// ptxas then spills inside this loop (320 instead of 144 bytes of spill loads).  The packed form is kept
// where registers allow it: the scatter below and the forward gather (126 registers).
template <bool WANT_PTS, bool BF16, bool STASH>
__device__ __forceinline__ void bw_gather(const Args& a, const BwTable& tb, float* a_hi, float* a_lo, float* tq,
                                          int b, int w, int lane, float* st_f) {
  const int c4 = lane & 7;
  const float third = 1.f / 3.f;
  const char* base = reinterpret_cast<const char*>(a.planes) +
                     (int64_t)b * 3 * a.Hp * a.Wp * (BF16 ? 64 : 128) + c4 * (BF16 ? 8 : 16);
#pragma unroll 1
  for (int it = 0; it < kBwRows / 4; ++it) {
    const int li = it * 4 + (lane >> 3);
    const int s = bw_row(w, li);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (tb.keep[s] >= 0.f) {
      float4 v[3][4];
      float4 d[3];
#pragma unroll
      for (int p = 0; p < 3; ++p) {
        const uint4 o = tb.toff[p][s];
        d[p] = tb.tder[p][s];
        v[p][0] = bw_ldg<BF16>(base, o.x); v[p][1] = bw_ldg<BF16>(base, o.y);
        v[p][2] = bw_ldg<BF16>(base, o.z); v[p][3] = bw_ldg<BF16>(base, o.w);
      }
      float4 t[3];
#pragma unroll
      for (int k = 0; k < 3; ++k) t[k] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int p = 0; p < 3; ++p) {
        const float4 wt = bw_weights(d[p]);
        bw_fma4(acc, v[p][0], wt.x * third); bw_fma4(acc, v[p][1], wt.y * third);
        bw_fma4(acc, v[p][2], wt.z * third); bw_fma4(acc, v[p][3], wt.w * third);
        if (WANT_PTS) {
          const float4 gx = bw_lerp4(bw_sub4(v[p][1], v[p][0]), bw_sub4(v[p][3], v[p][2]), d[p].y);
          const float4 gy = bw_lerp4(bw_sub4(v[p][2], v[p][0]), bw_sub4(v[p][3], v[p][1]), d[p].x);
          bw_fma4(t[nfi_plane_axis_a(p)], gx, d[p].z * third);
          bw_fma4(t[nfi_plane_axis_b(p)], gy, d[p].w * third);
        }
      }
      if (WANT_PTS) {
        float* dst = tq + (kBwRows * (w >> 2) + li) * kBwTLd + c4 * 4;
#pragma unroll
        for (int k = 0; k < 3; ++k) *reinterpret_cast<float4*>(dst + 32 * k) = t[k];
      }
    }
    nfi_mlp::store_feat4(a_hi, a_lo, s, c4, acc);
    // training path: the feature row stays in the caller's workspace for the weight-gradient reduction
    if (STASH && tb.keep[s] >= 0.f) *reinterpret_cast<float4*>(st_f + (int64_t)tb.sid[s] * 32 + c4 * 4) = acc;
  }
}

// T rows of a lane quarter: scratch -> TMEM.  The two warps of the quarter take 48 columns each of all
// 32 lanes (thread i <-> lane i of the quarter); rows that were not gathered carry stale values that
// no one uses.
__device__ __forceinline__ void bw_store_t(const float* tq, uint32_t tmem_t, int half, int lane) {
  const float* src = tq + lane * kBwTLd + 48 * half;
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    uint32_t r[16];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float4 v = *reinterpret_cast<const float4*>(src + 16 * j + 4 * q);
      r[4 * q] = __float_as_uint(v.x); r[4 * q + 1] = __float_as_uint(v.y);
      r[4 * q + 2] = __float_as_uint(v.z); r[4 * q + 3] = __float_as_uint(v.w);
    }
    nfi_tc::tmem_st16(tmem_t + 48 * half + 16 * j, r);
  }
  nfi_tc::tmem_st_wait();
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
      using namespace nfi_f2;
      const f4 g = mul(from4(*reinterpret_cast<const float4*>(&df[sl * kBwDfLd + c4 * 4])), third);
#pragma unroll
      for (int p = 0; p < 3; ++p) {
        // a clamped neighbour (x1 == x0 at the border) carries weight 0: adding 0 is harmless
        const uint4 o = tb.toff[p][s];
        const float4 wt = bw_weights(tb.tder[p][s]);
        const float4 r0 = to4(mul(g, wt.x)), r1 = to4(mul(g, wt.y)), r2 = to4(mul(g, wt.z)), r3 = to4(mul(g, wt.w));
        red_add_v4(reinterpret_cast<float*>(gbase + o.x), r0.x, r0.y, r0.z, r0.w);
        red_add_v4(reinterpret_cast<float*>(gbase + o.y), r1.x, r1.y, r1.z, r1.w);
        red_add_v4(reinterpret_cast<float*>(gbase + o.z), r2.x, r2.y, r2.z, r2.w);
        red_add_v4(reinterpret_cast<float*>(gbase + o.w), r3.x, r3.y, r3.z, r3.w);
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

template <bool PTS, int MODE, bool WANT_PTS, bool BF16, bool STASH = false>
__global__ void __launch_bounds__(kBwThreads, 1)
field_bwd_ws_kernel(const Args a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  BwSmem& sm = *reinterpret_cast<BwSmem*>(smem_raw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  nfi_mlp::load_weights(sm.w, a.f.W0, a.f.b0, a.f.W1, a.f.b1, a.f.n_out);
  nfi_mlp::load_weights_bwd(sm.bw, a.f.W0, a.f.W1, a.f.n_out);
  if (tid == 0) {
    for (int c = 0; c < 2; ++c) {
      nfi_tc::mbar_init(&sm.a_full[c], kBwPS);  if (c < kBwNA) nfi_tc::mbar_init(&sm.a_empty[c], 1);
      nfi_tc::mbar_init(&sm.df_full[c], 1);     nfi_tc::mbar_init(&sm.s_free[c], kBwS);
    }
    nfi_tc::mbar_init(&sm.acc_full, 1);
    nfi_tc::mbar_init(&sm.acc_ready, kBwE);
    for (int i = 0; i < kBwNM; ++i) {
      nfi_tc::mbar_init(&sm.meta_full[i], kBwPS);
      nfi_tc::mbar_init(&sm.meta_empty[i], kBwE + kBwS + 1);
    }
    nfi_tc::mbar_fence_init();
  }
  constexpr int kMmaWarp = kBwPS + kBwE + kBwS;
  if (warp == kMmaWarp) nfi_tc::tmem_alloc(&sm.tmem_base, kBwTmemCols);
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
    // ------------------------------------------------------------------ gather
    const int w = warp, q = w & 3, half = w >> 2;
    const int row = bw_row(w, lane & (kBwRows - 1));
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    BwIn& in = sm.in[w];
    if (lane < kBwRows && total > 0) bw_prefetch<PTS, MODE>(a, in, 0, tm.tile(0), row, lane, per_image);
    for (uint32_t it = 0; it < total; ++it) {
      const uint32_t m = it % kBwNM, c = it & 1;
      const int b = (int)(it / (uint32_t)ntile), t = tm.tile((int)(it % (uint32_t)ntile));
      BwTable& tb = sm.tab[it % kBwNT];
      if (it >= (uint32_t)kBwNT) {                       // table slot: tile it - kBwNT is finished
        const uint32_t r = it - kBwNT;
        nfi_tc::mbar_wait(&sm.meta_empty[r % kBwNM], (r / kBwNM) & 1);
      }
      if (w == 0 && lane == 0) NFI_BTRACE(0, it, 0);
      bool act = false;
      cp_async_wait_all();
      if (lane < kBwRows) act = bw_phase1<PTS, MODE>(a, tb, in, b, t, row, lane, per_image);
      const unsigned wany = __any_sync(0xffffffffu, act);
      if (lane == 0) sm.meta[m].flag[w] = wany ? 1 : 0;
      __syncwarp();
      if (lane == 0) nfi_tc::mbar_arrive(&sm.meta_full[m]);
      if (lane < kBwRows && it + 1 < total) {            // inputs of the next tile arrive during the gather
        const uint32_t n = it + 1;
        bw_prefetch<PTS, MODE>(a, in, (int)(n / (uint32_t)ntile), tm.tile((int)(n % (uint32_t)ntile)), row, lane, per_image);
      }
      if (w == 0 && lane == 0) NFI_BTRACE(0, it, 1);
      nfi_tc::mbar_wait(&sm.meta_full[m], (it / kBwNM) & 1);          // the flags of all eight warps
      // layer 0 of the previous user of this F stage has read it.  Waited for on idle tiles too: every tile
      // completes one phase of the barrier, so the parity test is only sound one phase behind
      const uint32_t ca = kBwNA == 2 ? c : 0u;
      nfi_tc::mbar_wait(&sm.a_empty[ca], (((kBwNA == 2 ? it >> 1 : it)) & 1) ^ 1);
      // T / dF buffers c read by the scatter warps (tile it - 2).  Waited for on idle tiles too, like
      // a_empty -- see "Phase discipline" at the top of the file
#ifdef NFI_BW_OLD_PROTOCOL    // developer reproducer of the round-1 protocol (tests/test_ws_protocol.py)
      if (bw_any(sm.meta[m]))
#endif
      if (it >= 2) nfi_tc::mbar_wait(&sm.s_free[c], ((it >> 1) - 1) & 1);
      if (bw_any(sm.meta[m])) {
        if (w == 0 && lane == 0) NFI_BTRACE(0, it, 2);
        bw_gather<WANT_PTS, BF16, STASH>(a, tb, sm.a[ca][0], sm.a[ca][1], sm.tq[q], b, w, lane,
                                         STASH ? a.st_f + (int64_t)b * per_image * 32 : nullptr);
        if (WANT_PTS) {
          nfi_tc::named_bar_sync(1 + q, 64);             // both halves of the quarter are in the scratch
          bw_store_t(sm.tq[q], tmem + lane_off + kBwColT + 96 * c, half, lane);
          nfi_tc::named_bar_sync(1 + q, 64);             // ... and have been read: the scratch is free
          nfi_tc::fence_before_sync();
        }
        nfi_tc::fence_proxy_async();
      }
      __syncwarp();
      if (lane == 0) nfi_tc::mbar_arrive(&sm.a_full[c]);
      if (w == 0 && lane == 0) NFI_BTRACE(0, it, 3);
    }
  } else if (warp < kBwPS + kBwE) {
    // ------------------------------------------------------------------ epilogue: the MLP chain
    const int ew = warp - kBwPS;
    const int row = ew * 32 + lane;
    const uint32_t tl = tmem + ((uint32_t)(ew * 32) << 16);
    const NfiHeads hd = make_heads(a.f);
    const int npal = 3 * hd.A;
    const bool want_pal = MODE == 0 && a.g_palette != nullptr && hd.A > 0;
    const bool want_par = MODE == 0 && a.g_params != nullptr && hd.use_sdf;
    const bool two_pass = 2 + npal > kBwRed;
    float* pal = sm.pal[ew];
    float acc0 = 0.f, acc1 = 0.f;                  // per-lane sums of reduction slots lane, 32 + lane
    int cur_b = -1;
    uint32_t ph = 0;
    auto flush = [&]() {
      if (cur_b < 0) return;
      const int pbase = NFI_HID * NFI_C + NFI_HID + a.f.n_out * NFI_HID + a.f.n_out;
      if (lane < 2) { if (want_par && acc0 != 0.f) atomicAdd(&a.g_params[pbase + lane], acc0); }
      else if (want_pal && lane - 2 < npal && acc0 != 0.f) atomicAdd(&a.g_palette[(int64_t)cur_b * npal + lane - 2], acc0);
      if (want_pal && kBwRed - 2 + lane < npal && acc1 != 0.f) atomicAdd(&a.g_palette[(int64_t)cur_b * npal + kBwRed - 2 + lane], acc1);
      acc0 = acc1 = 0.f;
    };
    for (uint32_t it = 0; it < total; ++it) {
      const uint32_t m = it % kBwNM;
      const int b = (int)(it / (uint32_t)ntile);
      nfi_tc::mbar_wait(&sm.meta_full[m], (it / kBwNM) & 1);
      const BwTable& tb = sm.tab[it % kBwNT];
      const int sid = tb.sid[row];
      const float keep = tb.keep[row];
      if (STASH && sid >= 0) a.st_ok[(int64_t)b * per_image + sid] = (keep >= 0.f) ? 1 : 0;   // idle tiles too
      if (bw_any(sm.meta[m])) {
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
        nfi_tc::mbar_wait(&sm.acc_full, ph); ph ^= 1;
        if (ew == 0 && lane == 0) NFI_BTRACE(2, it, 0);
        nfi_tc::fence_after_sync();
        if (STASH) nfi_mlp::epilogue0_at_keep<4>(sm.w.b0, tl, 0, act ? a.st_h + id * 64 : nullptr);
        else nfi_mlp::epilogue0_at<4>(sm.w.b0, tl, 0);
        nfi_tc::fence_before_sync();
        __syncwarp();
        if (lane == 0) nfi_tc::mbar_arrive(&sm.acc_ready);
        if (ew == 0 && lane == 0) NFI_BTRACE(2, it, 1);
        // ---- layer 1 -> heads forward + backward -> dY
        nfi_tc::mbar_wait(&sm.acc_full, ph); ph ^= 1;
        if (ew == 0 && lane == 0) NFI_BTRACE(2, it, 2);
        nfi_tc::fence_after_sync();
        float dy[NFI_YPAD];
#pragma unroll
        for (int o = 0; o < NFI_YPAD; ++o) dy[o] = 0.f;
        float probs[NFI_YPAD - 1];                  // zero for idle rows: they add nothing below
#pragma unroll
        for (int i = 0; i < NFI_YPAD - 1; ++i) probs[i] = 0.f;
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
        if (STASH && act) {
#pragma unroll
          for (int i = 0; i < 4; ++i)
            *reinterpret_cast<float4*>(a.st_dy + id * 16 + 4 * i) = make_float4(dy[4 * i], dy[4 * i + 1], dy[4 * i + 2], dy[4 * i + 3]);
        }
        nfi_tc::fence_before_sync();
        __syncwarp();
        if (lane == 0) nfi_tc::mbar_arrive(&sm.acc_ready);
        if (ew == 0 && lane == 0) NFI_BTRACE(2, it, 3);
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
        nfi_tc::mbar_wait(&sm.acc_full, ph); ph ^= 1;
        if (ew == 0 && lane == 0) NFI_BTRACE(3, it, 0);
        nfi_tc::fence_after_sync();
        if (STASH) nfi_mlp::epilogue2_keep<4>(tl, 0, act ? a.st_dh + id * 64 : nullptr);
        else nfi_mlp::epilogue2<4>(tl, 0);
        nfi_tc::fence_before_sync();
        __syncwarp();
        if (lane == 0) nfi_tc::mbar_arrive(&sm.acc_ready);
        if (ew == 0 && lane == 0) NFI_BTRACE(3, it, 1);
      }
      __syncwarp();
      if (lane == 0) nfi_tc::mbar_arrive(&sm.meta_empty[m]);
    }
    flush();
  } else if (warp < kMmaWarp) {
    // ------------------------------------------------------------------ scatter: dF (+ T) -> gradients
    const int sw = warp - kBwPS - kBwE;
    const int row = sw * 32 + lane;
    const uint32_t tl = tmem + ((uint32_t)(sw * 32) << 16);
    const float inv_r = 1.f / a.f.scene_range;
    float* sc = sm.dfe[sw];
    uint32_t dfph[2] = {0, 0};
    for (uint32_t it = 0; it < total; ++it) {
      const uint32_t m = it % kBwNM, c = it & 1;
      const int b = (int)(it / (uint32_t)ntile);
      nfi_tc::mbar_wait(&sm.meta_full[m], (it / kBwNM) & 1);
      const BwTable& tb = sm.tab[it % kBwNT];
      const int sid = tb.sid[row];
      const float keep = tb.keep[row];
      const bool any = bw_any(sm.meta[m]) != 0;
      float gp[3] = {0.f, 0.f, 0.f};
      float g0[16], g1[16];
#ifdef NFI_BW_INJECT_SCATTER_DELAY   // developer reproducer: one scatter warp falls behind now and then
      if (sw == 3 && (it % 5u) == 1u) for (int d = 0; d < NFI_BW_INJECT_SCATTER_DELAY; ++d) __nanosleep(1000);
#endif
#ifdef NFI_BW_OLD_PROTOCOL
      if (any)
#endif
      nfi_tc::mbar_wait(&sm.a_full[c], (it >> 1) & 1);            // T buffer c written; every tile (phase discipline)
      if (any) {
        nfi_tc::mbar_wait(&sm.df_full[c], dfph[c]); dfph[c] ^= 1;
        if (sw == 0 && lane == 0) NFI_BTRACE(1, it, 0);
        nfi_tc::fence_after_sync();
        nfi_tc::tmem_ld16(tl + kBwColDF + 32 * c, g0);
        nfi_tc::tmem_ld16(tl + kBwColDF + 32 * c + 16, g1);
        nfi_tc::tmem_ld_wait();
        if (WANT_PTS) {
#pragma unroll
          for (int k = 0; k < 3; ++k) {
            float t0[16], t1[16];
            nfi_tc::tmem_ld16(tl + kBwColT + 96 * c + 32 * k, t0);
            nfi_tc::tmem_ld16(tl + kBwColT + 96 * c + 32 * k + 16, t1);
            nfi_tc::tmem_ld_wait();
            float s0 = 0.f, s1 = 0.f;
#pragma unroll
            for (int i = 0; i < 16; ++i) { s0 = fmaf(g0[i], t0[i], s0); s1 = fmaf(g1[i], t1[i], s1); }
            gp[k] = keep >= 0.f ? (s0 + s1) * inv_r : 0.f;      // idle rows carry stale T
          }
        }
        nfi_tc::fence_before_sync();
      }
      __syncwarp();
      if (lane == 0) nfi_tc::mbar_arrive(&sm.s_free[c]);
      if (sw == 0 && lane == 0) NFI_BTRACE(1, it, 1);
      if (WANT_PTS && keep != -1.f) {
        float* dst = a.g_points + ((int64_t)b * per_image + sid) * 3;
        dst[0] = gp[0]; dst[1] = gp[1]; dst[2] = gp[2];
      }
#ifdef NFI_EXP_NOSCATTER
      if (false) {
#else
      if (any && a.g_planes) {
#endif
#pragma unroll 1
        for (int h = 0; h < 32 / kBwER; ++h) {          // kBwER rows through the warp's scratch at a time
          __syncwarp();
          if (lane / kBwER == h) {
            float* dst = sc + (lane % kBwER) * kBwDfLd;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              *reinterpret_cast<float4*>(dst + 4 * i) = make_float4(g0[4 * i], g0[4 * i + 1], g0[4 * i + 2], g0[4 * i + 3]);
              *reinterpret_cast<float4*>(dst + 16 + 4 * i) = make_float4(g1[4 * i], g1[4 * i + 1], g1[4 * i + 2], g1[4 * i + 3]);
            }
          }
          __syncwarp();
          bw_scatter_rows<kBwER>(a, tb, sc, b, sw * 32 + h * kBwER, lane);
        }
      }
      if (sw == 0 && lane == 0) NFI_BTRACE(1, it, 2);
      __syncwarp();
      if (lane == 0) nfi_tc::mbar_arrive(&sm.meta_empty[m]);
    }
  } else {
    // ------------------------------------------------------------------ MMA issuer
    // the whole warp waits (converged); one elected lane issues the MMAs, commits and arrives
    uint32_t rph = 0;
    for (uint32_t it = 0; it < total; ++it) {
      const uint32_t m = it % kBwNM, c = it & 1;
      nfi_tc::mbar_wait(&sm.meta_full[m], (it / kBwNM) & 1);
      nfi_tc::mbar_wait(&sm.a_full[c], (it >> 1) & 1);
      const int any = bw_any(sm.meta[m]);
      __syncwarp();
      if (!any) {
        if (nfi_tc::elect_one()) { nfi_tc::mbar_arrive(&sm.meta_empty[m]); nfi_tc::mbar_arrive(&sm.a_empty[kBwNA == 2 ? c : 0u]); }
        __syncwarp();
        continue;
      }
      nfi_tc::fence_after_sync();
      if (nfi_tc::elect_one()) {
        nfi_tc::mbar_arrive(&sm.meta_empty[m]);
        const uint32_t ca = kBwNA == 2 ? c : 0u;
        nfi_mlp::issue_layer0_at(sm.w, sm.a[ca][0], sm.a[ca][1], tmem);
        nfi_tc::mma_commit(&sm.a_empty[ca]);
        nfi_tc::mma_commit(&sm.acc_full);
      }
      __syncwarp();
      if (lane == 0) NFI_BTRACE(4, it, 0);
      nfi_tc::mbar_wait(&sm.acc_ready, rph); rph ^= 1;
      nfi_tc::fence_after_sync();
      if (nfi_tc::elect_one()) { nfi_mlp::issue_layer1_at(sm.w, tmem); nfi_tc::mma_commit(&sm.acc_full); }
      __syncwarp();
      if (lane == 0) NFI_BTRACE(4, it, 1);
      nfi_tc::mbar_wait(&sm.acc_ready, rph); rph ^= 1;
      nfi_tc::fence_after_sync();
      if (nfi_tc::elect_one()) { nfi_mlp::issue_dh(sm.bw, tmem); nfi_tc::mma_commit(&sm.acc_full); }
      __syncwarp();
      if (lane == 0) NFI_BTRACE(4, it, 2);
      nfi_tc::mbar_wait(&sm.acc_ready, rph); rph ^= 1;
      nfi_tc::fence_after_sync();
      if (nfi_tc::elect_one()) { nfi_mlp::issue_df(sm.bw, tmem, kBwColDF + 32 * c); nfi_tc::mma_commit(&sm.df_full[c]); }
      __syncwarp();
      if (lane == 0) NFI_BTRACE(4, it, 3);
    }
  }
  nfi_tc::fence_before_sync();
  __syncthreads();
  if (warp == kMmaWarp) nfi_tc::tmem_dealloc(tmem, kBwTmemCols);
}

}  // namespace nfi_ftc
