/*
 * nope_nerf_b200.h -- C ABI of the B200-native NoPe-NeRF hot path (libnope_nerf_b200.so).
 *
 * The reference (js0n-lai/my-nope-nerf) is pure Python/PyTorch and has NO FFI of its own
 * (SURVEY.md section 8b); the entry points below are the ones a maintainer would bind from
 * the reference's Python call sites, one per stage and direction.  Each cites the reference
 * code it replaces (paths relative to the reference root).
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in `_host`;
 *   - all floating-point data is fp32, row-major, contiguous; indices are int64; masks are uint8;
 *   - the library never allocates or frees device memory: the caller owns every buffer, including
 *     the workspace whose size the *_workspace_bytes functions report;
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*); no hidden syncs;
 *   - return value: 0 = ok, >0 = cudaError_t, <0 = library error (nnb_error_string decodes both);
 *   - gradient outputs named d_* are OVERWRITTEN unless documented as "accumulated (+=)".
 */
#ifndef NOPE_NERF_B200_H
#define NOPE_NERF_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NNB_ERR_BAD_ARG (-1)
#define NNB_ERR_WORKSPACE (-2)
#define NNB_ERR_UNSUPPORTED (-3)
#define NNB_ERR_NO_DEVICE (-4)
#define NNB_ERR_TOO_LARGE (-5) /* problem exceeds the fixed capacity of an allocation-free entry point */

#define NNB_NUM_PARAMS 24 /* 12 Linear layers x {weight, bias}, reference state_dict order */

/* Precision policy of the MLP contraction. */
#define NNB_PREC_FP32 0 /* CUDA-core fp32 (tight-tolerance parity path)            */
#define NNB_PREC_BF16 1 /* tcgen05 / TMEM tensor-core path, bf16 operands, fp32 acc */

int nnb_version(void);
/* kernels launched by this library in this process so far */
unsigned long long nnb_launch_count(void);
const char *nnb_error_string(int code);
/* sm count, compute capability of the current device. */
int nnb_device_info(int *sm_count_host, int *cc_major_host, int *cc_minor_host);

/* ------------------------------------------------------------------------------------------
 * Stage 1 -- pose application and ray generation
 * ---------------------------------------------------------------------------------------- */

/* LearnPose.forward: c2w = [Exp(r[cam]) | t[cam]; 0 0 0 1] @ init_c2w[cam].
 * Replaces model/poses.py:23-31, model/common.py:277-330 (vec2skew, Exp, make_c2w, convert3x4_4x4).
 * r, t: (n_cams,3); init_c2w: (n_cams,4,4) or NULL; c2w: (4,4). */
int nnb_pose_fwd(const float *r, const float *t, const float *init_c2w, int cam_id, float *c2w, void *stream);
/* d_r_row, d_t_row: (3,) gradients for row cam_id (overwritten). */
int nnb_pose_bwd(const float *r, const float *t, const float *init_c2w, int cam_id, const float *d_c2w,
                 float *d_r_row, float *d_t_row, void *stream);

/* 4x4 inverse (world_mat = inverse(c2w), model/training.py:257; torch.inverse call sites
 * model/common.py:139-141,206-208) and its backward dM = -inv^T dInv inv^T. */
int nnb_mat4_inv_fwd(const float *m, float *inv, void *stream);
int nnb_mat4_inv_bwd(const float *inv, const float *d_inv, float *d_m, void *stream);

/* Relative pose between the frame and its reference: c = a @ b on row-major 4x4 matrices (model/training.py:329
 * `ref_Rt @ torch.inverse(world_mat)`, :344 `world_mat @ torch.inverse(ref_Rt)`).  Backward: d_a = d_c b^T, d_b = a^T d_c
 * (either may be NULL). */
int nnb_mat4_mul_fwd(const float *a, const float *b, float *c, void *stream);
int nnb_mat4_mul_bwd(const float *a, const float *b, const float *d_c, float *d_a, float *d_b, void *stream);

/* Learn_Distortion.forward: model/distortions.py:19-27.  out = {scale, shift}.
 * scale below 0.01 -> constant 0.01 (gradient cut); last camera pinned to 1 when fix_scaleN. */
int nnb_distortion_fwd(const float *scales, const float *shifts, int cam_id, int n_cams, int fix_scaleN,
                       float *out2, void *stream);
int nnb_distortion_bwd(const float *scales, int cam_id, int n_cams, int fix_scaleN, const float *d_out2,
                       float *d_scale_row, float *d_shift_row, void *stream);

/* Ray choice (model/training.py:277 `torch.randperm(h*w)[:n_points]`): n_pick DISTINCT indices of [0, n_total), the head of
 * a keyed pseudo-random permutation (6-round Feistel network with cycle walking), without sorting n_total keys.
 * keys: 6 int64 on the device (low 32 bits used), drawn by the caller from its generator; out (n_pick,) int64. */
int nnb_ray_choice(const int64_t *keys, long long n_total, int n_pick, int64_t *out, void *stream);

/* Batch staging for the captured step: n <= NNB_COPY_MAX device-to-device copies (HOST arrays of device pointers and
 * byte counts, passed to the kernel by value) and up to two int64 scalars (the frame indices of model/training.py:262,323;
 * scalar_dst[q] NULL = skip) in ONE launch instead of one memcpy node per tensor of the batch dict. */
#define NNB_COPY_MAX 16
int nnb_copy_many(const void *const *src, void *const *dst, const unsigned long long *bytes, int n,
                  int64_t *const *scalar_dst, const int64_t *scalar, void *stream);

/* arange_pixels (model/common.py:13-40): loc (h*w,2) int64 [x,y], scaled (h*w,2) fp32 in [-1,1]. */
int nnb_pixel_grid(int h, int w, int64_t *loc, float *scaled, void *stream);

/* Ray selection: gathers what a training step needs for `ray_idx` without building the full
 * pixel grid.  Replaces model/training.py:264,285-288 and model/network.py:21-24:
 *   pixels[i]  = arange_pixels(h,w)[ray_idx[i]]
 *   rgb_gt[i]  = img[:, ray_idx[i]]                      (img is (3,h,w))
 *   depth[i]   = depth_img[ray_idx[i]] * scale + shift   (scale_shift = {scale, shift} on device, or NULL)
 * Any output pointer may be NULL. */
int nnb_gather_rays_fwd(const int64_t *ray_idx, int n_rays, int h, int w, const float *img,
                        const float *depth_img, const float *scale_shift, float *pixels, float *rgb_gt,
                        float *depth, void *stream);
/* d_scale_shift (2,) overwritten. */
int nnb_gather_rays_bwd(const int64_t *ray_idx, int n_rays, const float *depth_img, const float *d_depth,
                        float *d_scale_shift, void *stream);

/* Ray generation.  Replaces model/common.py:112-237 (transform_to_world, origin_to_world,
 * image_points_to_world: 9 torch.inverse calls) and model/rendering.py:53-80.
 * camera_mat, world_mat, scale_mat: general (4,4).  pixels (R,2), depth (R,).
 * Outputs: A (4,4) = S^-1 W^-1 K^-1 (kept for backward), origin (3,), dirs (R,3), dir_norm (R,),
 * d_gt (R,) supervision distance, mask (R,) uint8 = finite(d_gt) & d_gt != 0. */
int nnb_raygen_fwd(const float *pixels, const float *depth, const float *camera_mat, const float *world_mat,
                   const float *scale_mat, int n_rays, int normalise_ray, float *A, float *origin, float *dirs,
                   float *dir_norm, float *d_gt, uint8_t *mask, void *stream);
/* d_dirs (R,3), d_origin (3,), d_d_gt (R,) may each be NULL (= zero).
 * Outputs: d_depth (R,), d_camera_mat / d_world_mat / d_scale_mat (4,4) (each may be NULL).
 * scratch: 16 floats of device memory. */
int nnb_raygen_bwd(const float *pixels, const float *depth, const float *camera_mat, const float *world_mat,
                   const float *scale_mat, const float *A, int n_rays, int normalise_ray, const float *d_dirs,
                   const float *d_origin, const float *d_d_gt, float *d_depth, float *d_camera_mat,
                   float *d_world_mat, float *d_scale_mat, float *scratch16, void *stream);

/* NDC ray warp, get_ndc_rays_fxfy (model/common.py:632-675) as called by Renderer.sample_ndc
 * (model/rendering.py:169-181): focal = {camera_mat[0][0], camera_mat[1][1]}, near plane `near_` (1.0 there).
 * origin (3,) shared camera centre, dirs (R,3) -> ndc_o (R,3), ndc_d (R,3). */
int nnb_ndc_rays_fwd(const float *camera_mat, float near_, const float *origin, const float *dirs, int n_rays,
                     float *ndc_o, float *ndc_d, void *stream);
/* d_ndc_o / d_ndc_d (R,3) (either may be NULL = zero) -> d_origin_rows (R,3) per-ray gradient of the shared
 * origin (the caller sums the rows), d_dirs (R,3), d_fxfy (2,) overwritten (gradient of the two focal entries). */
int nnb_ndc_rays_bwd(const float *camera_mat, float near_, const float *origin, const float *dirs, int n_rays,
                     const float *d_ndc_o, const float *d_ndc_d, float *d_origin_rows, float *d_dirs,
                     float *d_fxfy, void *stream);

/* ------------------------------------------------------------------------------------------
 * Stages 2+3 -- stratified sampling, positional encoding, NeRF MLP, alpha compositing
 * ---------------------------------------------------------------------------------------- */

typedef struct {
  int n_rays;
  int n_samples;    /* 128 (rendering.num_points - outside_steps)                  */
  float near_;      /* rendering.depth_range[0]                                    */
  float far_;       /* rendering.depth_range[1]                                    */
  int occ_softplus; /* 1: softplus, 0: relu      (model.occ_activation)            */
  int dist_alpha;   /* rendering.dist_alpha                                         */
  int white_bkgd;   /* rendering.white_background                                   */
  int use_dir;      /* rendering.use_ray_dir                                        */
  int precision;    /* NNB_PREC_*                                                   */
} nnb_render_cfg;

/* Workspace size.  save_for_backward != 0 keeps the activations a later nnb_render_bwd needs. */
size_t nnb_render_workspace_bytes(const nnb_render_cfg *cfg, int save_for_backward);

/* Renderer.nope_nerf forward, `uniform` sampling branch: model/rendering.py:89-126,139-141,183-198
 * + OfficialStaticNerf.forward model/official_nerf.py:60-119.
 * params_host: HOST array of NNB_NUM_PARAMS device pointers in this order:
 *   layers0.{0,2,4,6}.{weight,bias}, layers1.{0,2,4,6}.{weight,bias}, fc_density.{w,b},
 *   fc_feature.{w,b}, rgb_layers.0.{w,b}, fc_rgb.{w,b}        (PyTorch (out,in) row-major fp32)
 * packed: output of nnb_pack_weights (NNB_PREC_BF16 only, else NULL).
 * origin (3,), dirs (R,3) [already normalised if normalise_ray], noise (R,S) in [0,1) or NULL.
 * Outputs: rgb (R,3), dist (R,), weights (R,S), alpha (R,S), z_vals (R,S)  (last three may be NULL
 * when save_for_backward == 0). */
int nnb_render_fwd(const nnb_render_cfg *cfg, const float *const *params_host, const void *packed,
                   const float *origin, const float *dirs, const float *noise, float *rgb, float *dist,
                   float *weights, float *alpha, float *z_vals, void *workspace, size_t workspace_bytes,
                   int save_for_backward, void *stream);

/* Backward of the above.  Needs the workspace of a forward run with save_for_backward != 0 plus that
 * forward's weights/alpha/z_vals.  d_rgb (R,3), d_dist (R,), d_alpha_out (R,S) or NULL (gradient w.r.t. the
 * returned `alpha`), d_weights_out (R,S) or NULL.
 * d_params_host: HOST array of NNB_NUM_PARAMS device pointers, accumulated (+=).
 * d_origin (3,) overwritten, d_dirs (R,3) overwritten. */
int nnb_render_bwd(const nnb_render_cfg *cfg, const float *const *params_host, const void *packed,
                   const float *origin, const float *dirs, const float *weights, const float *alpha,
                   const float *z_vals, const float *d_rgb, const float *d_dist, const float *d_alpha_out,
                   const float *d_weights_out, float *const *d_params_host, float *d_origin, float *d_dirs,
                   void *workspace, size_t workspace_bytes, void *stream);

/* Per-ray extras for the sampling branches whose rays do not share one origin (`sample_ndc`,
 * model/rendering.py:169-181 + get_ndc_rays_fxfy model/common.py:632-675): the sample points are
 * ray_origins[r] + dirs[r] * z while the view-direction encoding still sees -view_dirs[r].
 * NULL members fall back to the shared `origin` / to `dirs`.  The d_* members are backward outputs
 * ((R,3), overwritten) and are required exactly when their input is given.
 * wgrad_stream (bf16 path, backward only; NULL = `stream`): the weight-gradient kernel is launched on this stream, ordered
 * after the data-gradient kernel by an event.  It keeps one small CTA per SM (64 registers x 192 threads), so what follows the
 * render backward on `stream` -- the ray-generation and pose backward, which need d_origin / d_dirs only -- runs beside it.
 * The caller joins wgrad_stream before it reads d_params or reuses the workspace. */
typedef struct {
  const float *ray_origins; /* (R,3) or NULL */
  const float *view_dirs;   /* (R,3) or NULL */
  float *d_ray_origins;     /* (R,3) or NULL */
  float *d_view_dirs;       /* (R,3) or NULL */
  void *wgrad_stream;       /* cudaStream_t or NULL */
} nnb_ray_extras;

/* nnb_render_fwd / nnb_render_bwd with the extras above (ex == NULL: identical to the plain calls). */
int nnb_render_fwd_ex(const nnb_render_cfg *cfg, const float *const *params_host, const void *packed,
                      const float *origin, const float *dirs, const nnb_ray_extras *ex, const float *noise,
                      float *rgb, float *dist, float *weights, float *alpha, float *z_vals, void *workspace,
                      size_t workspace_bytes, int save_for_backward, void *stream);
int nnb_render_bwd_ex(const nnb_render_cfg *cfg, const float *const *params_host, const void *packed,
                      const float *origin, const float *dirs, const nnb_ray_extras *ex, const float *weights,
                      const float *alpha, const float *z_vals, const float *d_rgb, const float *d_dist,
                      const float *d_alpha_out, const float *d_weights_out, float *const *d_params_host,
                      float *d_origin, float *d_dirs, void *workspace, size_t workspace_bytes, void *stream);

/* Stand-alone stage 3 (model/rendering.py:116-126,139-141), used by tests and by callers that bring
 * their own per-sample rgb/alpha.  alpha (R,S) [sigma if dist_alpha], rgb_s (R,S,3), z (R,S). */
int nnb_composite_fwd(const float *alpha_in, const float *rgb_s, const float *z, int n_rays, int n_samples,
                      int dist_alpha, int white_bkgd, float *rgb, float *dist, float *weights, float *alpha_out,
                      void *stream);
int nnb_composite_bwd(const float *alpha_in, const float *rgb_s, const float *z, const float *weights,
                      int n_rays, int n_samples, int dist_alpha, int white_bkgd, const float *d_rgb,
                      const float *d_dist, const float *d_weights_out, float *d_alpha_in, float *d_rgb_s,
                      void *stream);

/* OfficialStaticNerf.forward at explicit points (model/official_nerf.py:69-96), forward only.
 * pts (M,3), viewdirs (M,3) or NULL (then rgb may be NULL: density only).  rgb (M,3), dens (M,) =
 * alpha (or sigma if dist_alpha).  workspace: nnb_mlp_points_workspace_bytes(M). */
size_t nnb_mlp_points_workspace_bytes(long long n_points);
int nnb_mlp_points_fwd(const float *const *params_host, const float *pts, const float *viewdirs, long long n_points,
                       int occ_softplus, int dist_alpha, float *rgb, float *dens, void *workspace,
                       size_t workspace_bytes, void *stream);

/* Repacks the 24 fp32 tensors into the bf16 stage images the tcgen05 kernels stream with bulk TMA
 * (forward images followed by the transposed images of the data-gradient chain). */
size_t nnb_pack_weights_bytes(void);
int nnb_pack_weights(const float *const *params_host, void *packed, void *stream);

/* OfficialStaticNerf.gradient (model/official_nerf.py:46-58): grad (M,3) = -d raw_density / d p at explicit points (the
 * surface normals of the phong renderer, model/rendering.py:255-260, and of normal_loss, :127-135), fp32 path.
 * raw_out (M,) optional: fc_density output before the activation.  Forward only w.r.t. the parameters (no second order). */
size_t nnb_mlp_points_grad_workspace_bytes(long long n_points);
int nnb_mlp_points_density_grad(const float *const *params_host, const float *pts, long long n_points, float *grad,
                                float *raw_out, void *workspace, size_t workspace_bytes, void *stream);

/* Debug aid: device buffer (>= 64 int64) that receives per-layer clock64 stamps of CTA 0 / first ray pair of
 * the following tensor-core forward launches; NULL switches it off. */
int nnb_debug_set_trace(void *buf);
/* Measurement hook: on != 0 makes every later tensor-core nnb_render_bwd record CUDA events around its two kernels (data-gradient
 * chain, weight gradient) on the call's stream; ms2 != NULL (HOST pointer, 2 floats) first reads the last recorded pair
 * {dgrad ms, wgrad ms} (synchronises on that call).  Do not enable while capturing a CUDA graph. */
int nnb_debug_bwd_timing(int on, float *ms2);

/* ------------------------------------------------------------------------------------------
 * Stage 4 -- rgb / depth / reprojection losses
 * ---------------------------------------------------------------------------------------- */

/* Loss.get_rgb_full_loss + the logging mse: model/losses.py:28-33,200.
 * out3 = { sum|d| / n_rays, sum d^2 / n_rays, mean d^2 }.  pred, gt (n_rays,3). */
int nnb_rgb_loss_fwd(const float *pred, const float *gt, int n_rays, float *out3, void *stream);
/* d_pred = g_out3[0]*dL1 + g_out3[1]*dL2 + g_out3[2]*dMSE  (g_out3 on device). */
int nnb_rgb_loss_bwd(const float *pred, const float *gt, int n_rays, const float *g_out3, float *d_pred,
                     void *stream);

/* Loss.get_depth_loss 'l1': model/losses.py:60-63, on DENSE tensors + mask (avoids the boolean-index
 * sync of model/rendering.py:151-153).  out2 = { sum_mask |p-g| / count, count }.  mask may be NULL. */
int nnb_depth_l1_fwd(const float *pred, const float *gt, const uint8_t *mask, int n, float *out2, void *stream);
int nnb_depth_l1_bwd(const float *pred, const float *gt, const uint8_t *mask, int n, const float *out2,
                     const float *g_loss, float *d_pred, float *d_gt, void *stream);

/* Exact k-th smallest (k = (count-1)/2: torch.median's lower median) of the masked values, with the
 * index of its first occurrence.  model/losses.py:43,46.  out_val (1,), out_idx (1,) int64. */
int nnb_median_select(const float *x, const uint8_t *mask, int n, float *out_val, int64_t *out_idx, void *stream);

/* Loss.depth_loss_dpt ('invariant'): model/losses.py:35-58.  stats8 (8 floats) is kept for backward. */
int nnb_depth_invariant_fwd(const float *pred, const float *gt, const uint8_t *mask, int n, float *loss,
                            float *stats8, int64_t *med_idx2, void *stream);
int nnb_depth_invariant_bwd(const float *pred, const float *gt, const uint8_t *mask, int n, const float *stats8,
                            const int64_t *med_idx2, const float *g_loss, float *d_pred, float *d_gt,
                            void *stream);

/* Point-cloud back-projection: model/training.py:354-362 (nearest resize to (h4,w4), clamp below
 * `nearest_limit`, transform_to_world with identity world/scale).  depth_img (h,w) raw;
 * scale_shift {scale,shift} on device or NULL.  pc (h4*w4,3); clamped (h4*w4,) uint8 kept for backward. */
int nnb_backproject_fwd(const float *depth_img, int h, int w, int h4, int w4, const float *scale_shift,
                        const float *camera_mat, float nearest_limit, float *pc, uint8_t *clamped, void *stream);
int nnb_backproject_bwd(const float *depth_img, int h, int w, int h4, int w4, const float *camera_mat,
                        const uint8_t *clamped, const float *d_pc, float *d_scale_shift, void *stream);

/* F.interpolate(img, (h4,w4), mode='bilinear') as called at model/training.py:365-366. img (3,h,w). */
int nnb_resize_bilinear(const float *img, int c, int h, int w, int h4, int w4, float *out, void *stream);

/* Reprojection sampling: model/training.py:367-382, model/common.py:75-109,436-457.
 * img1r, img2r (3,h4,w4) resized images; pc1 (P,3); R_rel (3,3), t_rel (3,); camera_mat (4,4).
 * Outputs rgb_pc1 (P,3), rgb_pc1_proj (P,3), valid (P,) uint8. */
int nnb_reproj_fwd(const float *img1r, const float *img2r, int h4, int w4, const float *pc1, const float *R_rel,
                   const float *t_rel, const float *camera_mat, float nearest_limit, float *rgb_pc1,
                   float *rgb_pc1_proj, uint8_t *valid, void *stream);
/* d_pc1 (P,3), d_R_rel (9,), d_t_rel (3,) overwritten. */
int nnb_reproj_bwd(const float *img2r, int h4, int w4, const float *pc1, const float *R_rel, const float *t_rel,
                   const float *camera_mat, float nearest_limit, const float *d_rgb_pc1_proj, float *d_pc1,
                   float *d_R_rel, float *d_t_rel, void *stream);

/* Loss.get_rgb_s_loss + mean_on_mask: model/losses.py:79-87,152-159.
 * out2 = { sum_valid clamp(|a-b|,0,1) / (3*count), count }; 0 when count == 0. */
int nnb_photometric_fwd(const float *rgb1, const float *rgb2, const uint8_t *valid, int n_points, float *out2,
                        void *stream);
int nnb_photometric_bwd(const float *rgb1, const float *rgb2, const uint8_t *valid, int n_points,
                        const float *out2, const float *g_loss, float *d_rgb1, float *d_rgb2, void *stream);

/* ------------------------------------------------------------------------------------------
 * Stage 5 -- Chamfer point-cloud loss
 * ---------------------------------------------------------------------------------------- */

/* Loss.comp_closest_pts_idx_with_split: model/losses.py:129-144.  src (n_src,3), dst (n_dst,3).
 * idx[i] = first argmin_j of sqrt(sum_c (src_ic - dst_jc)^2) evaluated in fp32 like the reference.
 * dist (n_src,) = that minimum distance (may be NULL).  workspace: nnb_chamfer_workspace_bytes(n_src, n_dst)
 * (bounds + the bf16 operand images of the tensor-core screening pass, 64 B per point). */
size_t nnb_chamfer_workspace_bytes(int n_src, int n_dst);
/* test hook: 0 = both sweeps on tcgen05 (default), 1 = both on the CUDA cores, 2 = tcgen05 screen + CUDA-core exact sweep
 * (returns the previous setting) */
int nnb_debug_chamfer_pass1(int cuda_cores);
int nnb_chamfer_nn(const float *src, int n_src, const float *dst, int n_dst, int64_t *idx, float *dist, void *workspace,
                   size_t workspace_bytes, void *stream);
/* nnb_chamfer_nn without a caller-supplied workspace.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  Same
 * inputs, outputs and results as nnb_chamfer_nn; the scratch is a file-scope `static __device__` array of the library
 * (4.7 MB, held by every GPU that loads the library until exit), so n_src and n_dst are limited to
 * NNB_CHAMFER_STATIC_MAX_POINTS each -- beyond that it returns NNB_ERR_TOO_LARGE.  It calls no cudaMalloc / cudaFree.  All
 * calls share the one array: calls that overlap on the GPU (two streams, two threads) race; serialise them. */
#define NNB_CHAMFER_STATIC_MAX_POINTS 32768
int nnb_chamfer_nn_static(const float *src, int n_src, const float *dst, int n_dst, int64_t *idx, float *dist, void *stream);
/* Loss.comp_point_point_error: model/losses.py:145-150.  loss = mean_i ||X_i - Y[idx_i]||. */
int nnb_chamfer_err_fwd(const float *X, int n, const float *Y, const int64_t *idx, float *loss, void *stream);
/* d_X (n,3) overwritten; d_Y (m,3) accumulated (+=) by scatter-add. */
int nnb_chamfer_err_bwd(const float *X, int n, const float *Y, const int64_t *idx, const float *g_loss,
                        float *d_X, float *d_Y, void *stream);

/* The clouds the Chamfer term compares: model/training.py:394-399.  X = (pc1 / s) R^T + t, Y = pc2 / s with R, t the
 * rotation and translation of the row-major 4x4 Rt (relative pose) and s = scale[0] (scale == NULL: no division).
 * Backward overwrites d_pc1 (n1,3), d_pc2 (n2,3), d_Rt (16; last row zero) and d_scale (1). */
int nnb_align_clouds_fwd(const float *pc1, int n1, const float *pc2, int n2, const float *Rt, const float *scale,
                         float *X, float *Y, void *stream);
int nnb_align_clouds_bwd(const float *pc1, int n1, const float *pc2, int n2, const float *Rt, const float *scale,
                         const float *dX, const float *dY, float *d_pc1, float *d_pc2, float *d_Rt, float *d_scale, void *stream);

/* ------------------------------------------------------------------------------------------
 * Step glue
 * ---------------------------------------------------------------------------------------- */

/* Fused Adam over one flat fp32 buffer (torch.optim.Adam semantics, no amsgrad, weight_decay 0):
 * model/training.py:93-99.  step is the 1-based step count. */
int nnb_adam_step(float *param, const float *grad, float *exp_avg, float *exp_avg_sq, size_t n, float lr,
                  float beta1, float beta2, float eps, int step, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* NOPE_NERF_B200_H */
