// Warp-specialised, software-pipelined forward of the fused triplane field (one persistent CTA per
// SM, 17 warps).  Roles talk through mbarriers only; no __syncthreads inside the tile loop.
//
//   producers  (8 warps)  phase 1 (coords -> texel offsets / weights) + phase 2 (gather) of tile i+1
//                         into a 2-stage ring of UMMA A-operand tiles, while ...
//   MMA thread (1 warp)   ... issues tcgen05.mma for whichever tile has its operands ready
//                         (layer 0 of tile i+1 goes ahead of layer 1 of tile i), and ...
//   epilogue   (8 warps)  ... runs softplus / hi-lo split / heads of tile i out of TMEM
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
#pragma once
#include "nfi_field_tc.cuh"

namespace nfi_ftc {

constexpr int kWsProducerWarps = 8, kWsEpilogueWarps = 8;
constexpr int kWsThreads = (kWsProducerWarps + kWsEpilogueWarps + 1) * 32;   // 544
constexpr int kNS = 3;      // operand stages
constexpr int kNM = 4;      // meta stages
constexpr int kNA = 2;      // accumulator sets
constexpr int kAccCols = 256;

struct WsMeta {
  float keep[kTile];
  int sid[kTile];
  int flag[8];      // per producer warp: any of its 16 samples active
};
__device__ __forceinline__ int ws_any(const WsMeta& m) {
  return m.flag[0] | m.flag[1] | m.flag[2] | m.flag[3] | m.flag[4] | m.flag[5] | m.flag[6] | m.flag[7];
}

struct WsTables {            // rows 16w .. 16w+15 are private to producer warp w
  int4 toff[3][kTile];
  float4 twgt[3][kTile];
};

struct WsFwdSmem {
  nfi_mlp::FwdWeights w;
  float a[kNS][2][nfi_mlp::kStageFloats];
  WsTables tab;
  WsMeta meta[kNM];
  uint64_t a_full[kNS], a_empty[kNS], meta_full[kNM], meta_empty[kNM];
  uint64_t d0_full[kNA], h_full[kNA], d1_full[kNA];
  uint32_t tmem_base;
  uint32_t pad;
};

// phase 1 for the pipelined kernels: writes the producer table + the meta entry
template <bool PTS, int MODE>
__device__ __forceinline__ bool ws_phase1(const Args& a, WsTables& tb, WsMeta& mt, int64_t id, int sid, int tid) {
  float keep = -1.f;
  bool active = false;
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
    active = sample_active(a, keep);
    if (active) {
#pragma unroll
      for (int p = 0; p < 3; ++p) {
        const float ix = nfi_unnormalize(x[nfi_plane_axis_a(p)], a.Wp, nullptr);
        const float iy = nfi_unnormalize(x[nfi_plane_axis_b(p)], a.Hp, nullptr);
        NfiFoot f;
        nfi_footprint(ix, iy, a.Wp, a.Hp, f);
        tb.toff[p][tid] = make_int4((f.y0 * a.Wp + f.x0) * 8, (f.y0 * a.Wp + f.x1) * 8,
                                    (f.y1 * a.Wp + f.x0) * 8, (f.y1 * a.Wp + f.x1) * 8);
        tb.twgt[p][tid] = make_float4(f.wnw, f.wne, f.wsw, f.wse);
      }
    }
  }
  mt.keep[tid] = keep;
  mt.sid[tid] = sid;
  return active;
}

// phase 2 for the pipelined kernels (256 producer threads, ptid = 0..255)
__device__ __forceinline__ void ws_gather(const Args& a, const WsTables& tb, const WsMeta& mt, float* a_hi,
                                          float* a_lo, int b, int ptid) {
  const int lane = ptid & 31, warp = ptid >> 5;
  const int c4 = lane & 7;
  const float third = 1.f / 3.f;
  const int plane_f4 = a.Hp * a.Wp * 8;
  const float4* base = reinterpret_cast<const float4*>(a.planes) + (int64_t)b * 3 * plane_f4 + c4;
#pragma unroll 2
  for (int it = 0; it < 4; ++it) {
    const int s = warp * 16 + it * 4 + (lane >> 3);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (sample_active(a, mt.keep[s])) {
      float4 v[3][4];
      float4 w[3];
#pragma unroll
      for (int p = 0; p < 3; ++p) {
        const int4 o = tb.toff[p][s];
        w[p] = tb.twgt[p][s];
        const float4* P = base + p * plane_f4;
        v[p][0] = __ldg(P + o.x); v[p][1] = __ldg(P + o.y); v[p][2] = __ldg(P + o.z); v[p][3] = __ldg(P + o.w);
      }
#pragma unroll
      for (int p = 0; p < 3; ++p) {
        float4 e;
        e.x = fmaf(v[p][3].x, w[p].w, fmaf(v[p][2].x, w[p].z, fmaf(v[p][1].x, w[p].y, v[p][0].x * w[p].x)));
        e.y = fmaf(v[p][3].y, w[p].w, fmaf(v[p][2].y, w[p].z, fmaf(v[p][1].y, w[p].y, v[p][0].y * w[p].x)));
        e.z = fmaf(v[p][3].z, w[p].w, fmaf(v[p][2].z, w[p].z, fmaf(v[p][1].z, w[p].y, v[p][0].z * w[p].x)));
        e.w = fmaf(v[p][3].w, w[p].w, fmaf(v[p][2].w, w[p].z, fmaf(v[p][1].w, w[p].y, v[p][0].w * w[p].x)));
        acc.x += e.x; acc.y += e.y; acc.z += e.z; acc.w += e.w;
      }
      acc.x *= third; acc.y *= third; acc.z *= third; acc.w *= third;
    }
    nfi_mlp::store_feat4(a_hi, a_lo, s, c4, acc);
  }
}

}  // namespace nfi_ftc

namespace nfi_ftc {

template <bool PTS, int MODE>
__global__ void __launch_bounds__(kWsThreads, 1)
field_fwd_ws_kernel(const Args a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  WsFwdSmem& sm = *reinterpret_cast<WsFwdSmem*>(smem_raw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  nfi_mlp::load_weights(sm.w, a.f.W0, a.f.b0, a.f.W1, a.f.b1, a.f.n_out);
  if (tid == 0) {
    for (int i = 0; i < kNS; ++i) { nfi_tc::mbar_init(&sm.a_full[i], kWsProducerWarps); nfi_tc::mbar_init(&sm.a_empty[i], 1); }
    for (int i = 0; i < kNM; ++i) { nfi_tc::mbar_init(&sm.meta_full[i], kWsProducerWarps); nfi_tc::mbar_init(&sm.meta_empty[i], kWsEpilogueWarps + 1); }
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
  int64_t lo, hi;
  cta_tile_range(a.tiles_per_image, lo, hi);

  if (warp < kWsProducerWarps) {
    // ------------------------------------------------------------------ producers
    // every warp owns rows 16w .. 16w+15 of each tile end to end (coords by lanes 0-15, gather by
    // all 32 lanes): no cross-warp barrier, warps drift apart and cover each other's load latency
    const int ptid = tid;
    const int row = warp * 16 + (lane & 15);
    uint32_t it = 0;
    for (int b = 0; b < a.B; ++b) {
      for (int64_t tt = lo; tt < hi; ++tt, ++it) {
        const uint32_t m = it % kNM, s = it % kNS;
        WsMeta& mt = sm.meta[m];
        nfi_tc::mbar_wait(&sm.meta_empty[m], ((it / kNM) & 1) ^ 1);
        bool active = false;
        if (lane < 16) {
          const int sid = sample_in_image(a, (int)tt, row);
          const int64_t id = sid >= 0 ? (int64_t)b * per_image + sid : -1;
          active = ws_phase1<PTS, MODE>(a, sm.tab, mt, id, sid, row);
        }
        const unsigned wany = __any_sync(0xffffffffu, active);
        if (lane == 0) { mt.flag[warp] = wany ? 1 : 0; }
        __syncwarp();
        if (lane == 0) nfi_tc::mbar_arrive(&sm.meta_full[m]);
        nfi_tc::mbar_wait(&sm.a_empty[s], ((it / kNS) & 1) ^ 1);
        if (wany) ws_gather(a, sm.tab, mt, sm.a[s][0], sm.a[s][1], b, ptid);
        nfi_tc::fence_proxy_async();
        __syncwarp();
        if (lane == 0) nfi_tc::mbar_arrive(&sm.a_full[s]);
      }
    }
  } else if (warp < kWsProducerWarps + kWsEpilogueWarps) {
    // ------------------------------------------------------------------ epilogue
    const int ew = warp - kWsProducerWarps;
    const int row = (ew & 3) * 32 + lane;                       // sample row = TMEM lane
    const uint32_t lane_off = (uint32_t)((ew & 3) * 32) << 16;
    const NfiHeads hd = make_heads(a.f);
    uint32_t it = 0, act = 0;
    for (int b = 0; b < a.B; ++b) {
      for (int64_t tt = lo; tt < hi; ++tt, ++it) {
        const uint32_t m = it % kNM;
        const WsMeta& mt = sm.meta[m];
        nfi_tc::mbar_wait(&sm.meta_full[m], (it / kNM) & 1);
        const int sid = mt.sid[row];
        const float keep = mt.keep[row];
        const int64_t id = sid >= 0 ? (int64_t)b * per_image + sid : -1;
        if (ws_any(mt)) {
          const uint32_t c = act % kNA, cph = (act / kNA) & 1;
          const uint32_t tl = tmem + c * kAccCols + lane_off;
          nfi_tc::mbar_wait(&sm.d0_full[c], cph);
          nfi_tc::fence_after_sync();
          nfi_mlp::epilogue0_at<2>(sm.w.b0, tl, (ew >> 2) * 32);
          nfi_tc::fence_before_sync();
          __syncwarp();
          if (lane == 0) nfi_tc::mbar_arrive(&sm.h_full[c]);
          nfi_tc::mbar_wait(&sm.d1_full[c], cph);
          nfi_tc::fence_after_sync();
          if (ew < 4) {
            float y[nfi_mlp::kY];
            nfi_mlp::epilogue1_at(sm.w.b1, tl, y);
            if (id >= 0) {
              if (MODE == 0) {
                if (!sample_active(a, keep)) {
                  reinterpret_cast<float4*>(a.out)[id] = make_float4(0.f, 0.f, 0.f, 0.f);
                  if (a.probs) for (int q = 0; q < hd.A; ++q) a.probs[id * hd.A + q] = 0.f;
                } else {
                  float sigma, rgb[3], probs[NFI_YPAD - 1];
                  nfi_heads_fwd(hd, y, keep, a.palette ? a.palette + (int64_t)b * hd.A * 3 : nullptr, sigma, rgb, probs);
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
          ++act;
        } else if (ew < 4 && MODE == 0 && id >= 0) {      // whole tile outside the box
          reinterpret_cast<float4*>(a.out)[id] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (a.probs) for (int q = 0; q < hd.A; ++q) a.probs[id * hd.A + q] = 0.f;
        }
        __syncwarp();
        if (lane == 0) nfi_tc::mbar_arrive(&sm.meta_empty[m]);
      }
    }
  } else if (lane == 0) {
    // ------------------------------------------------------------------ MMA issuer (one thread)
    uint32_t it = 0, act = 0;
    bool pend = false;
    uint32_t pc = 0, pph = 0;
    for (int b = 0; b < a.B; ++b) {
      for (int64_t tt = lo; tt < hi; ++tt, ++it) {
        const uint32_t m = it % kNM;
        nfi_tc::mbar_wait(&sm.meta_full[m], (it / kNM) & 1);
        const int any = ws_any(sm.meta[m]);
        nfi_tc::mbar_arrive(&sm.meta_empty[m]);
        const uint32_t s = it % kNS, c = act % kNA;
        nfi_tc::mbar_wait(&sm.a_full[s], (it / kNS) & 1);
        if (!any) { nfi_tc::mbar_arrive(&sm.a_empty[s]); continue; }
        nfi_tc::fence_after_sync();
        nfi_mlp::issue_layer0_at(sm.w, sm.a[s][0], sm.a[s][1], tmem + c * kAccCols);
        nfi_tc::mma_commit(&sm.a_empty[s]);
        nfi_tc::mma_commit(&sm.d0_full[c]);
        if (pend) {
          nfi_tc::mbar_wait(&sm.h_full[pc], pph);
          nfi_tc::fence_after_sync();
          nfi_mlp::issue_layer1_at(sm.w, tmem + pc * kAccCols);
          nfi_tc::mma_commit(&sm.d1_full[pc]);
        }
        pend = true; pc = c; pph = (act / kNA) & 1;
        ++act;
      }
    }
    if (pend) {
      nfi_tc::mbar_wait(&sm.h_full[pc], pph);
      nfi_tc::fence_after_sync();
      nfi_mlp::issue_layer1_at(sm.w, tmem + pc * kAccCols);
      nfi_tc::mma_commit(&sm.d1_full[pc]);
    }
  }
  nfi_tc::fence_before_sync();
  __syncthreads();
  if (warp == kWsProducerWarps + kWsEpilogueWarps) nfi_tc::tmem_dealloc(tmem, kNA * kAccCols);
}

}  // namespace nfi_ftc
