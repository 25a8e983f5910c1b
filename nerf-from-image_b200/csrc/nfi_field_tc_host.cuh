// Host-side argument validation shared by the tensor-core field translation units.
#pragma once
#include "nfi_field_tc.cuh"

namespace nfi_ftc {

static inline int check_field(const NfiField* f, const char* who) {
  NFI_REQUIRE(f && f->W0 && f->b0 && f->W1 && f->b1, NFI_ERR_BAD_ARG, "%s: NULL decoder weights", who);
  NFI_REQUIRE(f->n_out >= 2 && f->n_out <= NFI_MAX_OUT, NFI_ERR_UNSUPPORTED, "%s: n_out = %d not in [2, %d]", who, f->n_out, NFI_MAX_OUT);
  NFI_REQUIRE(f->attention_values == 0 ? f->n_out == 4 : f->n_out == 1 + f->attention_values, NFI_ERR_BAD_ARG,
              "%s: n_out = %d does not match attention_values = %d", who, f->n_out, f->attention_values);
  NFI_REQUIRE(!f->use_sdf || (f->beta && f->alpha), NFI_ERR_BAD_ARG, "%s: use_sdf needs beta and alpha", who);
  NFI_REQUIRE(f->scene_range > 0.f, NFI_ERR_BAD_ARG, "%s: scene_range must be positive", who);
  return NFI_OK;
}

static inline int fill_args(Args& a, const char* who, const NfiField* field, const float* planes_cl,
                            const float* palette, int B, int Hp, int Wp, const float* origins,
                            const float* dirs, const float* z, const float* points, int64_t n_per_image,
                            int S, int mode, int img_w) {
  int rc = check_field(field, who);
  if (rc) return rc;
  NFI_REQUIRE(planes_cl, NFI_ERR_BAD_ARG, "%s: NULL planes", who);
  NFI_REQUIRE(nfi_aligned16(planes_cl), NFI_ERR_ALIGNMENT, "%s: planes must be 16-byte aligned", who);
  NFI_REQUIRE(B > 0 && Hp >= 2 && Wp >= 2 && n_per_image > 0 && S >= 1, NFI_ERR_BAD_ARG, "%s: bad sizes", who);
  NFI_REQUIRE((int64_t)Hp * Wp * 3 * 128 < (int64_t)1 << 32, NFI_ERR_UNSUPPORTED, "%s: plane too large", who);
  NFI_REQUIRE(n_per_image * S < ((int64_t)1 << 31) - 2 * kTile * kChunk, NFI_ERR_UNSUPPORTED, "%s: more than 2^31 samples per image", who);
  NFI_REQUIRE(mode == 0 || mode == 1, NFI_ERR_BAD_ARG, "%s: mode must be 0 or 1", who);
  NFI_REQUIRE(points || (origins && dirs && z), NFI_ERR_BAD_ARG, "%s: need points or origins/dirs/z", who);
  NFI_REQUIRE(!points || S == 1, NFI_ERR_BAD_ARG, "%s: explicit points need S = 1", who);
  NFI_REQUIRE(mode == 1 || field->attention_values == 0 || palette, NFI_ERR_BAD_ARG, "%s: attention head needs a palette", who);
  a.f = *field;
  a.planes = planes_cl; a.palette = palette; a.B = B; a.Hp = Hp; a.Wp = Wp;
  a.bshift = nfi_plane_bshift(Hp, Wp); a.nbx = Wp >> a.bshift;
  a.origins = origins; a.dirs = dirs; a.z = z; a.points = points;
  a.npi = n_per_image; a.S = S; a.img_w = img_w; a.skip_oob = 0;
  a.n_total = (int64_t)B * n_per_image * S;
  a.blocks_x = 0;
  const int64_t per_image = n_per_image * S;
  a.tiles_per_image = (per_image + kTile - 1) / kTile;
  if (!points && img_w > 0 && n_per_image % img_w == 0 && img_w % 16 == 0 && (n_per_image / img_w) % 8 == 0) {
    a.blocks_x = img_w / 16;                      // per_image is then a multiple of 128
  }
  a.inv_S = 1.f / (float)S;
  a.inv_bx = a.blocks_x > 0 ? 1.f / (float)a.blocks_x : 1.f;
  a.out = nullptr; a.probs = nullptr; a.g_out = nullptr;
  a.g_planes = a.g_palette = a.g_points = a.g_params = nullptr;
  a.st_f = a.st_h = a.st_dh = a.st_dy = nullptr; a.st_ok = nullptr;
  return NFI_OK;
}

// grid of the warp-specialised kernels: one CTA per SM, never more CTAs than tile chunks of one image
static inline int field_grid_chunks(const Args& a) {
  const int64_t chunks = (a.tiles_per_image + kChunk - 1) / kChunk;
  const int64_t cap = nfi_sm_count();
  const int64_t g = chunks < cap ? chunks : cap;
  return (int)(g < 1 ? 1 : g);
}

// persistent grid: ctas_per_sm CTAs on every SM, never more CTAs than tiles of one image
static inline int field_grid(const Args& a, int ctas_per_sm) {
  const int64_t cap = (int64_t)nfi_sm_count() * ctas_per_sm;
  const int64_t g = a.tiles_per_image < cap ? a.tiles_per_image : cap;
  return (int)(g < 1 ? 1 : g);
}

}  // namespace nfi_ftc
