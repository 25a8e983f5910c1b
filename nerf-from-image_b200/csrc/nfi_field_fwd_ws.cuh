// Warp-specialised, software-pipelined forward of the fused triplane field (one persistent CTA per
// SM, 13 warps).  Roles talk through mbarriers only; no __syncthreads inside the tile loop.
//
//   producers  (2 x 4 warps) phase 1 (coords -> texel offsets / weights) + phase 2 (gather) of tiles
//                         i+1, i+2 (one per group) into a ring of UMMA A-operand tiles, while ...
//   MMA thread (1 warp)   ... issues tcgen05.mma for whichever tile has its operands ready
//                         (layer 0 of tile i+1 goes ahead of layer 1 of tile i), and ...
//   epilogue   (4 warps)  ... runs softplus / hi-lo split / heads of tile i out of TMEM
//                         (two accumulator sets, so tile i+1's layer 0 never waits for tile i's heads).
//
// Barrier protocol per tile (ring indices advance only on tiles with at least one active sample):
//   meta_full[m]  P -> MMA, E   per-sample keep / id table + tile flag written
//   a_empty[s]    MMA(commit) -> P      stage s operands consumed by layer 0
//   a_full[s]     P -> MMA              stage s operands written (fence.proxy.async before arrive)
//   d0_full[c]    MMA(commit) -> E      layer 0 accumulators of set c complete
//   h_full[c]     E -> MMA              H (hi, lo) of set c written to TMEM, D0 of set c consumed
//   d1_full[c]    MMA(commit) -> E      layer 1 accumulators complete (also: H of set c is free again)
//   meta_empty[m] E -> P                table m no longer needed
//
// Phase discipline (audited in round 2; a parity wait is sound only while the waiter is at most one phase behind):
//   barrier        arrives                      waits                         why phase k + 1 cannot complete before the test of k
//   meta_full[m]   group of tile it, every tile MMA + epilogue, every tile    meta_full(it + 8) needs meta_empty(it): MMA and all
//                                                                             epilogue warps past tile it
//   meta_empty[m]  MMA at once; epilogue when    producers, every tile        the next phase needs the producers' own meta_full(it + 8)
//                  the tile's heads are done
//   a_full[s]      group of tile it, every tile MMA, every tile               group(it + 3) first waits a_empty(it) <- MMA past its test
//   a_empty[s]     MMA, every tile              producers, every tile         a_empty(it) needs a_full(it), posted after the test
//   d0_full / h_full / d1_full [c]              self-counted over ACTIVE tiles; MMA(k + 2) waits h_full(k), epilogue(k + 2) d0_full(k + 2)
// Liveness (the round-2 bug): the heads of active tile i run after the softplus of the next active tile, so that the
// layer-1 MMAs are never waited for.  Deferring them to the next ACTIVE tile is not enough: meta_empty(i) is only
// released by those heads, and after kNM - 1 idle tiles the producers wait for it while the epilogue waits for the
// producers.  Both the MMA warp and the epilogue warps therefore complete the pending tile on the first idle tile.
#pragma once
#include "nfi_f2.cuh"
#include "nfi_field_tc.cuh"

namespace nfi_ftc {

#ifndef NFI_WS_GROUPS
#define NFI_WS_GROUPS 2
#endif
#ifndef NFI_WS_GROUP_WARPS
#define NFI_WS_GROUP_WARPS 4
#endif
constexpr int kWsGroups = NFI_WS_GROUPS;          // producer groups: group g prepares tiles it % kWsGroups == g
constexpr int kWsGroupWarps = NFI_WS_GROUP_WARPS; // warps per group
constexpr int kWsRows = kTile / kWsGroupWarps;    // rows of the tile owned by one producer warp (16 or 32)
// 13 warps leave ptxas 126 registers per thread; with 8 epilogue warps (17 warps, 96 registers) the
// gather cannot keep its 12 LDG.128 in flight and the launch is 13 % slower (2.82 vs 2.49 ms, config 2).
#ifndef NFI_WS_EPI_WARPS
#define NFI_WS_EPI_WARPS 4
#endif
constexpr int kWsProducerWarps = kWsGroups * kWsGroupWarps, kWsEpilogueWarps = NFI_WS_EPI_WARPS;   // 4 or 8
static_assert(kWsEpilogueWarps == 4 || kWsEpilogueWarps == 8, "one or two epilogue warps per TMEM lane quarter");
constexpr int kWsEpiQ = 4 / (kWsEpilogueWarps / 4);      // 16-column groups of the hidden layer per epilogue warp
constexpr int kWsThreads = (kWsProducerWarps + kWsEpilogueWarps + 1) * 32;   // 416
#ifndef NFI_WS_STAGES
#define NFI_WS_STAGES 3
#endif
constexpr int kNS = NFI_WS_STAGES;      // operand stages
constexpr int kNM = 8;      // meta stages (a producer is never more than kNM tiles ahead of the heads)
constexpr int kNA = 2;      // accumulator sets
constexpr int kAccCols = 256;

// Developer trace (-DNFI_WS_TRACE builds only): SM clock at pipeline events of CTA 0, tiles
// [kTrT0, kTrT0 + kTrN) of the launch; dumped by the host after the launch (NFI_WS_TRACE_FILE).
#ifdef NFI_WS_TRACE
constexpr int kTrRoles = 5, kTrN = 128, kTrEv = 4, kTrT0 = 256;
__device__ long long g_ws_trace[kTrRoles * kTrN * kTrEv];
#define NFI_TRACE(role, it, ev)                                                                     \
  do {                                                                                              \
    if (blockIdx.x == 0 && (it) >= (uint32_t)kTrT0 && (it) < (uint32_t)(kTrT0 + kTrN))              \
      g_ws_trace[((role) * kTrN + ((it) - kTrT0)) * kTrEv + (ev)] = clock64();                      \
  } while (0)
#else
#define NFI_TRACE(role, it, ev) do {} while (0)
#endif

struct WsMeta {
  float keep[kTile];
  int sid[kTile];
  int flag[kWsGroupWarps];      // per producer warp: any of its 16 samples active
};
__device__ __forceinline__ int ws_any(const WsMeta& m) {
  int f = 0;
#pragma unroll
  for (int i = 0; i < kWsGroupWarps; ++i) f |= m.flag[i];
  return f;
}

struct WsTables {            // rows 16w .. 16w+15 are private to producer warp w of the group
  uint4 toff[3][kTile];      // byte offsets of the 4 texels from the image's first plane (plane included)
  float4 twgt[3][kTile];     // bilinear weights, pre-scaled by 1/3 (the mean over the planes)
};

struct WsFwdSmem {
  nfi_mlp::FwdWeights w;
  float a[kNS][2][nfi_mlp::kStageFloats];
  WsTables tab[kWsGroups];
  WsMeta meta[kNM];
  uint64_t a_full[kNS], a_empty[kNS], meta_full[kNM], meta_empty[kNM];
  uint64_t d0_full[kNA], h_full[kNA], d1_full[kNA];
  uint32_t tmem_base;
  uint32_t pad;
};

// inputs of phase 1 for one sample, fetched one tile ahead of their use
struct WsCoord {
  float o[3], d[3], z;       // PTS: o is the point itself
  int sid;
};

template <bool PTS>
__device__ __forceinline__ WsCoord ws_load_coord(const Args& a, int b, int t, int row, int per_image) {
  WsCoord c;
  int ray;
  c.sid = sample_and_ray(a, t, row, ray);
  if (c.sid >= 0) {
    const int64_t id = (int64_t)b * per_image + c.sid;
    if (PTS) {
      c.o[0] = __ldg(a.points + id * 3); c.o[1] = __ldg(a.points + id * 3 + 1); c.o[2] = __ldg(a.points + id * 3 + 2);
    } else {
      const int64_t r3 = ((int64_t)b * a.npi + ray) * 3;
#pragma unroll
      for (int i = 0; i < 3; ++i) { c.o[i] = __ldg(a.origins + r3 + i); c.d[i] = __ldg(a.dirs + r3 + i); }
      c.z = __ldg(a.z + id);
    }
  }
  return c;
}

// phase 1 for the pipelined kernels: writes the producer table + the meta entry
template <bool PTS, int MODE>
__device__ __forceinline__ bool ws_phase1(const Args& a, WsTables& tb, WsMeta& mt, const WsCoord& c, int row) {
  float keep = -1.f;
  bool active = false;
  if (c.sid >= 0) {
    float pt[3];
    if (PTS) { pt[0] = c.o[0]; pt[1] = c.o[1]; pt[2] = c.o[2]; }
    else nfi_point(c.o, c.d, c.z, pt);
    float x[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) x[i] = NFI_DIV(pt[i], a.f.scene_range);
    keep = 1.f;
    if (MODE == 0 && (fabsf(x[0]) > 1.f || fabsf(x[1]) > 1.f || fabsf(x[2]) > 1.f)) keep = 0.f;
    active = sample_active(a, keep);
    if (active) {
      const float third = 1.f / 3.f;
      const uint32_t plane_texels = (uint32_t)(a.Hp * a.Wp);
#pragma unroll
      for (int p = 0; p < 3; ++p) {
        const float ix = nfi_unnormalize(x[nfi_plane_axis_a(p)], a.Wp, nullptr);
        const float iy = nfi_unnormalize(x[nfi_plane_axis_b(p)], a.Hp, nullptr);
        NfiFoot f;
        nfi_footprint(ix, iy, a.Wp, a.Hp, f);
        const uint32_t r0 = p * plane_texels + (uint32_t)nfi_texel_row(f.y0, a.bshift, a.nbx);
        const uint32_t r1 = p * plane_texels + (uint32_t)nfi_texel_row(f.y1, a.bshift, a.nbx);
        const uint32_t c0 = (uint32_t)nfi_texel_col(f.x0, a.bshift), c1 = (uint32_t)nfi_texel_col(f.x1, a.bshift);
        tb.toff[p][row] = make_uint4((r0 + c0) * 128u, (r0 + c1) * 128u, (r1 + c0) * 128u, (r1 + c1) * 128u);
        tb.twgt[p][row] = make_float4(f.wnw * third, f.wne * third, f.wsw * third, f.wse * third);
      }
    }
  }
  mt.keep[row] = keep;
  mt.sid[row] = c.sid;
  return active;
}

__device__ __forceinline__ void ws_fma4(float4& acc, const float4& v, float w) {
  acc.x = fmaf(v.x, w, acc.x); acc.y = fmaf(v.y, w, acc.y); acc.z = fmaf(v.z, w, acc.z); acc.w = fmaf(v.w, w, acc.w);
}

// phase 2 for the pipelined kernels: warp w of a group gathers rows 16w .. 16w+15, 8 lanes per texel
template <bool BF16>
__device__ __forceinline__ void ws_gather(const Args& a, const WsTables& tb, const WsMeta& mt, float* a_hi,
                                          float* a_lo, int b, int gw, int lane) {
  const int c4 = lane & 7;
  const char* base = reinterpret_cast<const char*>(a.planes) +
                     (int64_t)b * 3 * a.Hp * a.Wp * (BF16 ? 64 : 128) + c4 * (BF16 ? 8 : 16);
#pragma unroll 2
  for (int it = 0; it < kWsRows / 4; ++it) {
    using namespace nfi_f2;
    const int s = gw * kWsRows + it * 4 + (lane >> 3);
    f4 acc = zero4();
    if (sample_active(a, mt.keep[s])) {
      float4 v[3][4];
      float4 w[3];
#pragma unroll
      for (int p = 0; p < 3; ++p) {
        const uint4 o = tb.toff[p][s];
        w[p] = tb.twgt[p][s];
        v[p][0] = ldg_texel4<BF16>(base, o.x); v[p][1] = ldg_texel4<BF16>(base, o.y);
        v[p][2] = ldg_texel4<BF16>(base, o.z); v[p][3] = ldg_texel4<BF16>(base, o.w);
      }
#pragma unroll
      for (int p = 0; p < 3; ++p) {       // same order and rounding as the scalar form, two channels per FFMA2
        acc = fma(from4(v[p][0]), w[p].x, acc); acc = fma(from4(v[p][1]), w[p].y, acc);
        acc = fma(from4(v[p][2]), w[p].z, acc); acc = fma(from4(v[p][3]), w[p].w, acc);
      }
    }
    nfi_mlp::store_feat4(a_hi, a_lo, s, c4, to4(acc));
  }
}

}  // namespace nfi_ftc

namespace nfi_ftc {

template <bool PTS, int MODE, bool BF16>
__global__ void __launch_bounds__(kWsThreads, 1)
field_fwd_ws_kernel(const Args a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  WsFwdSmem& sm = *reinterpret_cast<WsFwdSmem*>(smem_raw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  nfi_mlp::load_weights(sm.w, a.f.W0, a.f.b0, a.f.W1, a.f.b1, a.f.n_out);
  if (tid == 0) {
    for (int i = 0; i < kNS; ++i) { nfi_tc::mbar_init(&sm.a_full[i], kWsGroupWarps); nfi_tc::mbar_init(&sm.a_empty[i], 1); }
    for (int i = 0; i < kNM; ++i) { nfi_tc::mbar_init(&sm.meta_full[i], kWsGroupWarps); nfi_tc::mbar_init(&sm.meta_empty[i], kWsEpilogueWarps + 1); }
    for (int i = 0; i < kNA; ++i) { nfi_tc::mbar_init(&sm.d0_full[i], 1); nfi_tc::mbar_init(&sm.h_full[i], kWsEpilogueWarps); nfi_tc::mbar_init(&sm.d1_full[i], 1); }
    nfi_tc::mbar_fence_init();
  }
  if (warp == kWsProducerWarps + kWsEpilogueWarps) nfi_tc::tmem_alloc(&sm.tmem_base, kNA * kAccCols);
  nfi_tc::fence_proxy_async();
  nfi_tc::fence_before_sync();
  __syncthreads();
  nfi_tc::fence_after_sync();
  const uint32_t tmem = sm.tmem_base;
  const int64_t per_image = a.npi * a.S;
  const TileMap tm = cta_tile_map(a.tiles_per_image);
  const int ntile = tm.ntile;

  if (warp < kWsProducerWarps) {
    // ------------------------------------------------------------------ producers
    // every warp owns rows 16w .. 16w+15 of its group's tiles end to end (coords by lanes 0-15,
    // gather by all 32 lanes): no cross-warp barrier, warps drift apart and cover each other's
    // load latency; the inputs of the next tile's phase 1 are fetched before this tile's gather
    const int g = warp / kWsGroupWarps, gw = warp % kWsGroupWarps;
    const int row = gw * kWsRows + (lane & (kWsRows - 1));
    const uint32_t total = (uint32_t)a.B * (uint32_t)ntile;
    WsTables& tab = sm.tab[g];
    int b = 0, tl = g;
    while (tl >= ntile) { tl -= ntile; ++b; }
    WsCoord cur;
    cur.sid = -1;
    if (lane < kWsRows && (uint32_t)g < total) cur = ws_load_coord<PTS>(a, b, tm.tile(tl), row, (int)per_image);
    for (uint32_t it = g; it < total; it += kWsGroups) {
      const uint32_t m = it % kNM, s = it % kNS;
      WsMeta& mt = sm.meta[m];
      nfi_tc::mbar_wait(&sm.meta_empty[m], ((it / kNM) & 1) ^ 1);
      if (gw == 0 && lane == 0 && g < 2) NFI_TRACE(g, it, 0);
      bool active = false;
      if (lane < kWsRows) active = ws_phase1<PTS, MODE>(a, tab, mt, cur, row);
      const unsigned wany = __any_sync(0xffffffffu, active);
      if (lane == 0) { mt.flag[gw] = wany ? 1 : 0; }
      __syncwarp();
      if (lane == 0) nfi_tc::mbar_arrive(&sm.meta_full[m]);
      const int bcur = b;
      tl += kWsGroups;
      while (tl >= ntile) { tl -= ntile; ++b; }
      if (lane < kWsRows && it + kWsGroups < total) cur = ws_load_coord<PTS>(a, b, tm.tile(tl), row, (int)per_image);
      if (gw == 0 && lane == 0 && g < 2) NFI_TRACE(g, it, 1);
      nfi_tc::mbar_wait(&sm.a_empty[s], ((it / kNS) & 1) ^ 1);
      if (gw == 0 && lane == 0 && g < 2) NFI_TRACE(g, it, 2);
      if (wany) ws_gather<BF16>(a, tab, mt, sm.a[s][0], sm.a[s][1], bcur, gw, lane);
      nfi_tc::fence_proxy_async();
      __syncwarp();
      if (lane == 0) nfi_tc::mbar_arrive(&sm.a_full[s]);
      if (gw == 0 && lane == 0 && g < 2) NFI_TRACE(g, it, 3);
    }
  } else if (warp < kWsProducerWarps + kWsEpilogueWarps) {
    // ------------------------------------------------------------------ epilogue
    // order per active tile i: epilogue 0 of tile i, THEN heads of tile i-1 -- the layer-1 MMAs of
    // tile i-1 run while this warp is busy with tile i's softplus, so their latency is never exposed
    const int ew = warp - kWsProducerWarps;
    const int row = (ew & 3) * 32 + lane;                       // sample row = TMEM lane
    const uint32_t lane_off = (uint32_t)((ew & 3) * 32) << 16;
    const NfiHeads hd = make_heads(a.f);
    struct { bool valid; uint32_t c, cph, m, it; int64_t id; float keep; int b; } pend;
    pend.valid = false;
    auto finish = [&]() {
      const uint32_t tl = tmem + pend.c * kAccCols + lane_off;
      nfi_tc::mbar_wait(&sm.d1_full[pend.c], pend.cph);
      if (ew == 0 && lane == 0) NFI_TRACE(2, pend.it, 2);
      nfi_tc::fence_after_sync();
      if (ew < 4) {
        float y[nfi_mlp::kY];
        nfi_mlp::epilogue1_at(sm.w.b1, tl, y);
        const int64_t id = pend.id;
        if (id >= 0) {
          if (MODE == 0) {
            if (!sample_active(a, pend.keep)) {
              reinterpret_cast<float4*>(a.out)[id] = make_float4(0.f, 0.f, 0.f, 0.f);
              if (a.probs) for (int q = 0; q < hd.A; ++q) a.probs[id * hd.A + q] = 0.f;
            } else {
              float sigma, rgb[3], probs[NFI_YPAD - 1];
              nfi_heads_fwd(hd, y, pend.keep, a.palette ? a.palette + (int64_t)pend.b * hd.A * 3 : nullptr, sigma, rgb, probs);
              reinterpret_cast<float4*>(a.out)[id] = make_float4(sigma, rgb[0], rgb[1], rgb[2]);
              if (a.probs) {
#pragma unroll
                for (int q = 0; q < NFI_YPAD - 1; ++q) if (q < hd.A) a.probs[id * hd.A + q] = probs[q];
              }
            }
          } else {
#pragma unroll
            for (int o = 0; o < NFI_YPAD; ++o) if (o < a.f.n_out) a.out[id * a.f.n_out + o] = y[o];
          }
        }
      }
      nfi_tc::fence_before_sync();
      __syncwarp();
      if (lane == 0) nfi_tc::mbar_arrive(&sm.meta_empty[pend.m]);
      if (ew == 0 && lane == 0) NFI_TRACE(2, pend.it, 3);
      pend.valid = false;
    };
    uint32_t it = 0, act = 0;
    for (int b = 0; b < a.B; ++b) {
      for (int tl = 0; tl < ntile; ++tl, ++it) {
        const uint32_t m = it % kNM;
        const WsMeta& mt = sm.meta[m];
        nfi_tc::mbar_wait(&sm.meta_full[m], (it / kNM) & 1);
        const int sid = mt.sid[row];
        const float keep = mt.keep[row];
        const int64_t id = sid >= 0 ? (int64_t)b * per_image + sid : -1;
        if (ws_any(mt)) {
          const uint32_t c = act % kNA, cph = (act / kNA) & 1;
          nfi_tc::mbar_wait(&sm.d0_full[c], cph);
          if (ew == 0 && lane == 0) NFI_TRACE(2, it, 0);
          nfi_tc::fence_after_sync();
          nfi_mlp::epilogue0_at<kWsEpiQ>(sm.w.b0, tmem + c * kAccCols + lane_off, (ew >> 2) * 32);
          nfi_tc::fence_before_sync();
          __syncwarp();
          if (lane == 0) nfi_tc::mbar_arrive(&sm.h_full[c]);
          if (ew == 0 && lane == 0) NFI_TRACE(2, it, 1);
          if (pend.valid) finish();
          pend.valid = true; pend.it = it; pend.c = c; pend.cph = cph; pend.m = m; pend.id = id; pend.keep = keep; pend.b = b;
          ++act;
        } else {
          // An idle tile must not leave the previous active tile pending: its meta slot is only released by
          // finish(), and a producer that wraps the ring of kNM meta slots (an active tile followed by >= kNM - 1
          // idle ones: a pixel block that hits the box followed by blocks that miss it) would wait for that
          // release while this warp waits for the producer's next meta_full -- the deadlock behind the launch
          // failures of BASELINE config 3 (round 2; tests/test_ws_protocol.py::test_forward_idle_run_after_active_tile)
#ifndef NFI_FWD_OLD_DEFER       // developer reproducer of the round-1 behaviour
          if (pend.valid) finish();
#endif
          if (ew < 4 && MODE == 0 && id >= 0) {             // whole tile outside the box
            reinterpret_cast<float4*>(a.out)[id] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (a.probs) for (int q = 0; q < hd.A; ++q) a.probs[id * hd.A + q] = 0.f;
          }
          __syncwarp();
          if (lane == 0) nfi_tc::mbar_arrive(&sm.meta_empty[m]);
        }
      }
    }
    if (pend.valid) finish();
  } else {
    // ------------------------------------------------------------------ MMA issuer
    // the whole warp walks the loop (converged); one elected lane issues MMAs, commits and arrives
    uint32_t it = 0, act = 0;
    bool pend = false;
    uint32_t pc = 0, pph = 0;
    for (int b = 0; b < a.B; ++b) {
      for (int tl = 0; tl < ntile; ++tl, ++it) {
        const uint32_t m = it % kNM;
        nfi_tc::mbar_wait(&sm.meta_full[m], (it / kNM) & 1);
        const int any = ws_any(sm.meta[m]);
        __syncwarp();
        if (nfi_tc::elect_one()) nfi_tc::mbar_arrive(&sm.meta_empty[m]);
        const uint32_t s = it % kNS, c = act % kNA;
        nfi_tc::mbar_wait(&sm.a_full[s], (it / kNS) & 1);
        if (lane == 0) NFI_TRACE(3, it, 0);
        if (!any) {
          if (nfi_tc::elect_one()) nfi_tc::mbar_arrive(&sm.a_empty[s]);
#ifndef NFI_FWD_OLD_DEFER
          if (pend) {                                       // layer 1 of the last active tile: do not defer it past idle tiles
            nfi_tc::mbar_wait(&sm.h_full[pc], pph);
            nfi_tc::fence_after_sync();
            if (nfi_tc::elect_one()) {
              nfi_mlp::issue_layer1_at(sm.w, tmem + pc * kAccCols);
              nfi_tc::mma_commit(&sm.d1_full[pc]);
            }
            __syncwarp();
            pend = false;
          }
#endif
          continue;
        }
        nfi_tc::fence_after_sync();
        if (nfi_tc::elect_one()) {
          nfi_mlp::issue_layer0_at(sm.w, sm.a[s][0], sm.a[s][1], tmem + c * kAccCols);
          nfi_tc::mma_commit(&sm.a_empty[s]);
          nfi_tc::mma_commit(&sm.d0_full[c]);
        }
        __syncwarp();
        if (lane == 0) NFI_TRACE(3, it, 1);
        if (pend) {
          nfi_tc::mbar_wait(&sm.h_full[pc], pph);
          if (lane == 0) NFI_TRACE(3, it, 2);
          nfi_tc::fence_after_sync();
          if (nfi_tc::elect_one()) {
            nfi_mlp::issue_layer1_at(sm.w, tmem + pc * kAccCols);
            nfi_tc::mma_commit(&sm.d1_full[pc]);
          }
          __syncwarp();
          if (lane == 0) NFI_TRACE(3, it, 3);
        }
        pend = true; pc = c; pph = (act / kNA) & 1;
        ++act;
      }
    }
    if (pend) {
      nfi_tc::mbar_wait(&sm.h_full[pc], pph);
      nfi_tc::fence_after_sync();
      if (nfi_tc::elect_one()) {
        nfi_mlp::issue_layer1_at(sm.w, tmem + pc * kAccCols);
        nfi_tc::mma_commit(&sm.d1_full[pc]);
      }
      __syncwarp();
    }
  }
  nfi_tc::fence_before_sync();
  __syncthreads();
  if (warp == kWsProducerWarps + kWsEpilogueWarps) nfi_tc::tmem_dealloc(tmem, kNA * kAccCols);
}

}  // namespace nfi_ftc
