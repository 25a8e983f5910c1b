/*
 * nfi_b200 -- C ABI of the B200-native (sm_100a) volume-rendering hot path of nerf-from-image.
 *
 * The reference has no FFI: its boundary is the set of Python call signatures in
 * lib/nerf_utils.py, models/generator.py (TriplanarDecoder.forward, the `sampler` closure) and
 * run.py::render.  The Python package `nfi_b200` keeps those signatures and binds the entry points
 * below with ctypes (INTEGRATION.md shows the stub).  Each entry point names the reference code it
 * replaces (file:line relative to the reference checkout).
 *
 * Conventions
 *   - every pointer is a DEVICE pointer to fp32 data unless stated; the caller owns every buffer,
 *     including outputs and workspaces; the library allocates nothing and keeps no pointer;
 *   - `stream` is a cudaStream_t (passed as void*); kernels are launched on it and nothing
 *     synchronises -- safe under nn.DataParallel (several host threads, one device each);
 *   - return value 0 = success, negative = NFI_ERR_*; nfi_last_error() gives a thread-local message;
 *   - rays are indexed (b*H + h)*W + w; samples of a ray are contiguous ([R,S] row-major).
 */
#ifndef NFI_B200_H
#define NFI_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NFI_VERSION 100

#define NFI_OK 0
#define NFI_ERR_BAD_ARG (-1)      /* NULL pointer, non-positive size, unsupported combination      */
#define NFI_ERR_UNSUPPORTED (-2)  /* feat_ch != 32, hidden != 64, n_out > 16, S > 64 ...           */
#define NFI_ERR_ALIGNMENT (-3)    /* a vector-accessed buffer is not 16-byte aligned               */
#define NFI_ERR_LAUNCH (-4)       /* cudaGetLastError() after a launch                             */

#define NFI_FEAT_CH 32            /* channels per plane      (models/generator.py:384)             */
#define NFI_HIDDEN 64             /* decoder hidden width    (models/generator.py:294)             */
#define NFI_MAX_OUT 16            /* 1 + attention_values (<= 15) or 1 + 3                         */
#define NFI_MAX_S 64              /* samples per ray per pass (run.py:512 uses 64)                 */
#define NFI_NEARFAR_WS_BYTES 16384

typedef void* nfi_stream_t;

/* Decoder + heads description.  Weight pointers are the RAW EqualizedLinear tensors; the
 * 1/sqrt(fan_in) gain (models/stylegan.py:170-177) is applied inside the kernels. */
typedef struct NfiField {
  const float* W0;   /* [64, 32]  decoder.net[0].weight */
  const float* b0;   /* [64]                            */
  const float* W1;   /* [n_out, 64] decoder.net[2].weight */
  const float* b1;   /* [n_out]                         */
  const float* beta;  /* [1] (use_sdf) or NULL           */
  const float* alpha; /* [1] (use_sdf) or NULL           */
  int32_t n_out;            /* 1 + attention_values, or 1 + 3 for the plain RGB head */
  int32_t attention_values; /* A; 0 -> wide-sigmoid RGB head (models/generator.py:665-666) */
  int32_t use_sdf;          /* 1: Laplace-CDF density (generator.py:629-636); 0: softplus(d-1) */
  float scene_range;        /* points are divided by this (generator.py:604); 1.0 for raw decoder coords */
  int32_t planes_bf16;      /* bf16 mode (1e-2 contract): planes_cl points to a bfloat16 copy of the
                             * channels-last planes (same texel order, 64 bytes per texel); plane gradients
                             * stay fp32.  nfi_field_fwd, nfi_field_bwd without weight gradients, and
                             * nfi_field_bwd_train (weight gradients). */
} NfiField;

int nfi_version(void);
const char* nfi_last_error(void);

/* ---- A1: nerf_utils.get_ray_bundle (lib/nerf_utils.py:28-93) + F.normalize (run.py:196) -------
 * cam2world [B,16] row-major 4x4.  focal: NULL = orthographic; focal_stride 1 = [B], 0 = one shared
 * value (the 0-dim focal of eval_kitti_persp.py:948).  center [B,2] or NULL; bbox [B,2,2] or NULL.
 * normalize != 0 applies x / max(|x|, 1e-12) to the directions.  Outputs [B*H*W, 3]. */
int nfi_ray_bundle_fwd(const float* cam2world, const float* focal, int focal_stride,
                       const float* center, const float* bbox, int B, int H, int W, int normalize,
                       float* origins, float* dirs, nfi_stream_t stream);
/* Backward: g_origins/g_dirs [R,3] (either may be NULL) -> g_cam2world [B,16], g_focal [B]
 * (per image even when the focal is shared; both are OVERWRITTEN). */
int nfi_ray_bundle_bwd(const float* cam2world, const float* focal, int focal_stride,
                       const float* center, const float* bbox, int B, int H, int W, int normalize,
                       const float* g_origins, const float* g_dirs, float* g_cam2world,
                       float* g_focal, nfi_stream_t stream);

/* ---- A2: nerf_utils.compute_near_far_planes (lib/nerf_utils.py:227-275) ------------------------
 * workspace: NFI_NEARFAR_WS_BYTES bytes; on return (stream order) ((int32_t*)workspace)[0] holds the
 * number of rays that hit the box -- 0 is the case in which the reference raises (nerf_utils.py:260). */
int nfi_near_far(const float* origins, const float* dirs, int64_t R, float scene_range,
                 float* near, float* far, void* workspace, nfi_stream_t stream);

/* ---- A2 for callers that cannot allocate (CUDA-graph capture, tight serving loops, harnesses that
 *      forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 *      Same kernels and same results as nfi_near_far, but the reduction
 *      scratch is a file-scope `static __device__` array inside the library (NFI_NEARFAR_WS_BYTES =
 *      16 KiB, held by every GPU that loads the library until the process exits) instead of a caller
 *      workspace; the call performs no cudaMalloc / cudaFree and calls no allocating routine.
 *      hits_out: DEVICE int32 [1] receiving the number of rays that hit the box, or NULL.
 *      Limit: R <= NFI_NEARFAR_STATIC_MAX_RAYS, larger R is rejected with NFI_ERR_BAD_ARG.
 *      All calls on one device share the one array: calls that may overlap on the GPU (different
 *      streams without an ordering between them) RACE on it -- serialise them or use nfi_near_far.
 *      Nothing in the array has to start at zero (each call writes every word it later reads). ---- */
#define NFI_NEARFAR_STATIC_MAX_RAYS (1 << 24)
int nfi_near_far_static(const float* origins, const float* dirs, int64_t R, float scene_range,
                        float* near, float* far, int32_t* hits_out, nfi_stream_t stream);

/* ---- A3: nerf_utils.compute_query_points_from_rays (lib/nerf_utils.py:96-122) ------------------
 * u [R,S] uniforms or NULL; randomize with u == NULL draws Philox4x32-10 numbers from (seed, offset). */
int nfi_stratified(const float* near, const float* far, const float* u, int64_t R, int S,
                   int randomize, uint64_t seed, uint64_t offset, float* z, nfi_stream_t stream);
int nfi_points_fwd(const float* origins, const float* dirs, const float* z, int64_t R, int S,
                   float* points, nfi_stream_t stream);
/* g_points [R,S,3] -> g_origins [R,3] = sum_s g, g_dirs [R,3] = sum_s g * z (overwritten). */
int nfi_points_bwd(const float* g_points, const float* z, int64_t R, int S, float* g_origins,
                   float* g_dirs, nfi_stream_t stream);

/* ---- layout boundary: StyleGAN2 planes [B, 3*32, Hp, Wp] NCHW (models/generator.py:475-477)
 *      <-> channels-last [B, 3, Hp*Wp, 32] used by the gather/scatter kernels.  The texel order
 *      inside a plane is the library's own (8 x 8 texel blocks when Hp and Wp are multiples of 8,
 *      else row-major): `planes_cl` buffers are opaque -- only these two calls produce / consume
 *      them, and every `planes_cl` / `g_planes_cl` argument below expects this layout. -------------- */
int nfi_planes_to_cl(const float* planes_nchw, int B, int Hp, int Wp, float* planes_cl,
                     nfi_stream_t stream);
int nfi_planes_grad_from_cl(const float* g_planes_cl, int B, int Hp, int Wp, float* g_planes_nchw,
                            nfi_stream_t stream);

/* ---- A4-A6: TriplanarDecoder.forward (models/generator.py:301-331) + sampler heads (:587-681) ----
 * Sample points are either explicit (`points` [B, n_per_image, 3], S = 1) or implicit
 * o + d * z (`points` NULL; origins/dirs [B*n_per_image, 3], z [B*n_per_image, S]).
 * mode 0: heads -> out float4 (sigma, r, g, b) per sample; probs_out optional [.., A] (semantics);
 * mode 1: raw decoder -> out [.., n_out] (density_or_distance, features), no out-of-box mask.
 * img_w > 0 (ray mode, image width; n_per_image = H*W with H % 8 == 0, W % 16 == 0) lets a tile be
 * an 8x16 pixel block at one depth step (texel reuse in L1) instead of 128 consecutive samples.
 * skip_oob != 0 (fused render): out-of-box samples are written as (0,0,0,0) without any gather --
 * their sigma is 0, so their colour never reaches the image (lib/nerf_utils.py:143-146). */
int nfi_field_fwd(const NfiField* field, const float* planes_cl, const float* palette, int B, int Hp,
                  int Wp, const float* origins, const float* dirs, const float* z,
                  const float* points, int64_t n_per_image, int S, int mode, int img_w,
                  int skip_oob, float* out, float* probs_out, nfi_stream_t stream);
/* Backward.  g_out as `out`.  Accumulating outputs (caller zero-fills): g_planes_cl [B,3,Hp,Wp,32],
 * g_palette [B,A,3], g_params [64*32 + 64 + n_out*64 + n_out + 2] (W0,b0,W1,b1,beta,alpha).
 * g_points [.., 3] per sample is overwritten.  Any of them may be NULL (not computed).
 * want_weight_grads != 0 (generator training, run.py:980-986) also fills the W0,b0,W1,b1 part of
 * g_params; with 0 only beta / alpha are written (inversion: the decoder is frozen, run.py:630-632). */
int nfi_field_bwd(const NfiField* field, const float* planes_cl, const float* palette, int B, int Hp,
                  int Wp, const float* origins, const float* dirs, const float* z,
                  const float* points, int64_t n_per_image, int S, int mode, int img_w,
                  const float* g_out, float* g_planes_cl, float* g_palette, float* g_points,
                  float* g_params, int want_weight_grads, nfi_stream_t stream);

/* Generator-training gradient set (run.py:980-986, models/stylegan.py:170-177): nfi_field_bwd for ray samples in heads
 * mode (points == NULL, mode 0) with want_weight_grads = 1, on the warp-specialised kernel: the backward leaves F, H,
 * dHpre, dY of every sample in `workspace` (704 bytes per sample + 1) and one reduction kernel forms dW0, db0, dW1, db1
 * from them.  g_params as in nfi_field_bwd (accumulating, caller zero-fills; ALL of W0,b0,W1,b1,beta,alpha are
 * written).  workspace: 256-byte aligned, >= nfi_field_bwd_train_workspace_bytes(B * n_per_image * S); callers bound
 * it by splitting the batch by image (gradients accumulate). */
size_t nfi_field_bwd_train_workspace_bytes(int64_t n_samples);
int nfi_field_bwd_train(const NfiField* field, const float* planes_cl, const float* palette, int B, int Hp,
                        int Wp, const float* origins, const float* dirs, const float* z, int64_t n_per_image,
                        int S, int img_w, const float* g_out, float* g_planes_cl, float* g_palette,
                        float* g_points, float* g_params, void* workspace, size_t workspace_bytes,
                        nfi_stream_t stream);

/* ---- decoder MLP alone (models/generator.py:295-299, 329-330; EqualizedLinear gains applied
 *      inside): feat [n,32] -> y [n, n_out] on tcgen05 tensor cores (3xTF32 split, fp32
 *      accumulate in TMEM).  hpre [n,64] (optional) receives the hidden pre-activations. --------- */
int nfi_decoder_mlp_fwd(const NfiField* field, const float* feat, int64_t n, float* y, float* hpre,
                        nfi_stream_t stream);

/* ---- N2: the regulariser path (models/generator.py:505-585; lib/ops.py:58-120 `grid_sample2d`, the
 *      double-differentiable sampler behind TriplanarDecoder.forward(requires_double_backward=True)).
 * coords [B, n_per_image, 3] in box units ([-1, 1], NOT clamped: lib/ops.py:69-70), gy [B, n_per_image, n_out] the
 * upstream gradient of the raw decoder output -> gx [B, n_per_image, 3] = J^T gy, the gradient w.r.t. the coordinates
 * (the reference's torch.autograd.grad(sdf.sum(), bins_in, create_graph=True), generator.py:533-536).
 * _bwd is its backward (the double backward): gg [B, n_per_image, 3] -> ACCUMULATING g_planes_cl, g_W0 [64,32],
 * g_b0 [64], g_W1 [n_out,64] (any may be NULL; caller zero-fills).  fp32 planes only. */
int nfi_decoder_gradx_fwd(const NfiField* field, const float* planes_cl, int B, int Hp, int Wp,
                          const float* coords, const float* gy, int64_t n_per_image, float* gx,
                          nfi_stream_t stream);
int nfi_decoder_gradx_bwd(const NfiField* field, const float* planes_cl, int B, int Hp, int Wp,
                          const float* coords, const float* gy, const float* gg, int64_t n_per_image,
                          float* g_planes_cl, float* g_W0, float* g_b0, float* g_W1, nfi_stream_t stream);

/* ---- A7: render_volume_density_weights_only (lib/nerf_utils.py:166-182) + EG3D smoothing
 *      (run.py:266-272, smooth != 0).  sigma is read with stride `sigma_stride` floats. ----------- */
int nfi_weights_only(const float* sigma, int sigma_stride, const float* dirs, const float* z,
                     int64_t R, int S, int smooth, float* weights, nfi_stream_t stream);
/* ---- A8: nerf_utils.sample_pdf (lib/nerf_utils.py:185-224): bins [R,M], weights [R,M-1] ->
 *      samples [R,n]; u [R,n] or NULL (deterministic linspace); inds_out optional int32 [R,n]. ---- */
int nfi_sample_pdf(const float* bins, const float* weights, int64_t R, int M, int n, const float* u,
                   float* samples, int32_t* inds_out, nfi_stream_t stream);
/* A7 + smoothing + A8 in one pass over the coarse samples (run.py:261-281). */
int nfi_importance(const float* sigma, int sigma_stride, const float* dirs, const float* z_coarse,
                   int64_t R, int S, const float* u, float* z_fine, nfi_stream_t stream);

/* ---- nerf_utils.cumprod_exclusive (lib/nerf_utils.py:20-25): x [R,n] -> out [R,n], forward only
 *      (the reference only uses it inside render_volume_density*, which have their own kernels). */
int nfi_cumprod_exclusive(const float* x, int64_t R, int n, float* out, nfi_stream_t stream);

/* ---- A10: nerf_utils.render_volume_density (lib/nerf_utils.py:125-163) -------------------------
 * sigma [R,S], rgb [R,S,3], z [R,S], dirs [R,3]; optional normals [R,S,3], semantics [R,S,n_sem]. */
int nfi_composite_fwd(const float* sigma, const float* rgb, const float* dirs, const float* z,
                      const float* normals, const float* semantics, int n_sem, int64_t R, int S,
                      int white_background, float* rgb_map, float* depth_map, float* mask,
                      float* normal_map, float* semantic_map, nfi_stream_t stream);
int nfi_composite_bwd(const float* sigma, const float* rgb, const float* dirs, const float* z,
                      const float* semantics, int n_sem, int64_t R, int S, int white_background,
                      const float* g_rgb_map, const float* g_mask, const float* g_semantic_map,
                      float* g_sigma, float* g_rgb, float* g_dirs, float* g_semantics,
                      nfi_stream_t stream);

/* ---- A9 + A10 fused: sort/merge of coarse + fine samples (run.py:283-335) and compositing.
 * coarse/fine: float4 (sigma, r, g, b) [R,S]; perm [R,2S] uint8 (index into cat(coarse, fine)) and
 * z_sorted [R,2S] are outputs kept for the backward.  cam2world != NULL converts the depth to a
 * z-buffer depth (eval_kitti_persp.py:221-228); rays_per_image is then needed. ------------------- */
int nfi_merge_composite_fwd(const float* coarse, const float* fine, const float* z_coarse,
                            const float* z_fine, const float* dirs, int64_t R, int S,
                            int white_background, const float* cam2world, int64_t rays_per_image,
                            float* rgb_map, float* depth_map, float* mask, uint8_t* perm,
                            float* z_sorted, nfi_stream_t stream);
int nfi_merge_composite_bwd(const float* coarse, const float* fine, const float* z_sorted,
                            const uint8_t* perm, const float* dirs, int64_t R, int S,
                            int white_background, const float* g_rgb_map, const float* g_mask,
                            float* g_coarse, float* g_fine, nfi_stream_t stream);

/* ---- N3: F.normalize(q) + pose_utils.pose_to_matrix (run.py:2257-2262, lib/pose_utils.py:30-75) ----
 * z0 [B] or NULL (orthographic), t2 [B,2], s [B], q [B,4] = (w, x, y, z) -> cam2world [B,16] and, for
 * perspective poses, focal [B] = (1 + exp(z0)) / 2.  normalize_q != 0 folds q / max(|q|, 1e-12) in.
 * Backward: g_cam2world [B,16], g_focal [B] or NULL -> g_z0 [B] (perspective; may be NULL), g_t2 [B,2],
 * g_s [B], g_q [B,4] (all overwritten).  The orthographic branch divides per image (the evident intent
 * of pose_utils.py:69, which the reference only gets right for B = 1). */
int nfi_pose_to_matrix_fwd(const float* z0, const float* t2, const float* s, const float* q, int B,
                           int camera_flipped, int normalize_q, float* cam2world, float* focal,
                           nfi_stream_t stream);
int nfi_pose_to_matrix_bwd(const float* z0, const float* t2, const float* s, const float* q, int B,
                           int camera_flipped, int normalize_q, const float* g_cam2world,
                           const float* g_focal, float* g_z0, float* g_t2, float* g_s, float* g_q,
                           nfi_stream_t stream);

/* ---- A11 / A12: the whole render (run.py:176-350) and its backward as ONE pair of calls ----------
 * Sequences A1-A10 above on `stream` inside a caller-owned workspace (nothing is allocated, nothing
 * synchronises).  Ray r = (b*H + h)*W + w.  planes_cl: see nfi_planes_to_cl.  focal NULL = orthographic,
 * focal_stride 0 = one shared focal (eval_kitti_persp.py:948).  randomize != 0: stratified jitter and
 * random inverse-CDF uniforms, read from u_strat [R,S] / u_pdf [R,S] when given, else drawn in-kernel
 * with Philox4x32-10 from (seed, offset); randomize == 0: the reference's deterministic linspace.
 * Fine sampling is always on (arguments.py:124-127 default); S % 4 == 0, S <= NFI_MAX_S.
 * After the forward (stream order) ((int32_t*)workspace)[0] = number of rays that hit the scene box;
 * 0 is the case in which the reference raises (lib/nerf_utils.py:260). */
typedef struct NfiRender {
  NfiField field;
  const float* planes_cl;   /* [B,3,Hp*Wp,32] */
  const float* palette;     /* [B,A,3] attention values, NULL when field.attention_values == 0 */
  const float* cam2world;   /* [B,16] */
  const float* focal;       /* [B] / [1] / NULL */
  const float* center;      /* [B,2] or NULL */
  const float* bbox;        /* [B,2,2] or NULL */
  const float* u_strat;     /* [R,S] or NULL */
  const float* u_pdf;       /* [R,S] or NULL */
  uint64_t seed, offset;
  int32_t B, H, W, S, Hp, Wp;
  int32_t focal_stride;
  int32_t randomize;
  int32_t white_background; /* dataset_config['white_background'] (lib/nerf_utils.py:158-161) */
  int32_t zbuffer_depth;    /* eval_kitti_persp.py:221-228 */
} NfiRender;

/* bytes of workspace nfi_render_fwd needs (with_backward == 0) or fwd + bwd share (!= 0); 256-byte
 * aligned base required.  0 on a malformed parameter block. */
size_t nfi_render_workspace_bytes(const NfiRender* p, int with_backward);
/* rgb [R,3], depth [R], mask [R]. */
int nfi_render_fwd(const NfiRender* p, void* workspace, float* rgb, float* depth, float* mask,
                   nfi_stream_t stream);
/* `workspace` as left by nfi_render_fwd with the same `p`.  g_rgb [R,3], g_mask [R] or NULL.
 * ACCUMULATING outputs (caller zero-fills; NULL = not wanted): g_planes_cl, g_palette [B,A,3],
 * g_params (layout of nfi_field_bwd).  OVERWRITTEN: g_cam2world [B,16] and g_focal [B] (zeros for
 * orthographic cameras) -- pass both for pose gradients, both NULL to skip them (force_no_cam_grad,
 * run.py:211-214). */
int nfi_render_bwd(const NfiRender* p, void* workspace, const float* g_rgb, const float* g_mask,
                   float* g_planes_cl, float* g_palette, float* g_cam2world, float* g_focal,
                   float* g_params, int want_weight_grads, nfi_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* NFI_B200_H */
