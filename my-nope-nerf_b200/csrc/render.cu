// Stages 2+3 host orchestration behind nnb_render_fwd / nnb_render_bwd.
//   NNB_PREC_FP32: layer-by-layer CUDA-core fp32 path (this file + mlp_fp32.cuh)
//   NNB_PREC_BF16: fused tcgen05/TMEM path (mlp_tc.cu)
#include "mlp_fp32.cuh"
#include "render_fp32_kernels.cuh"
#include "mlp_points.cuh"

int nnb_composite_fwd_launch(const float *, const float *, const float *, int, int, int, int, float *, float *, float *, float *, cudaStream_t);
int nnb_composite_bwd_launch(const float *, const float *, const float *, int, int, int, int, const float *, const float *, const float *,
                             const float *, float *, float *, cudaStream_t);
// tensor-core path (mlp_tc.cu)
size_t nnb_tc_workspace_bytes(const nnb_render_cfg *cfg, int save);
int nnb_tc_render_fwd(const nnb_render_cfg *cfg, const float *const *params, const void *packed, const float *origin,
                      const float *dirs, const nnb_ray_extras *ex, const float *noise, float *rgb, float *dist, float *weights, float *alpha, float *z_vals,
                      void *ws, size_t ws_bytes, int save, cudaStream_t s);
int nnb_tc_render_bwd(const nnb_render_cfg *cfg, const float *const *params, const void *packed, const float *origin,
                      const float *dirs, const nnb_ray_extras *ex, const float *weights, const float *alpha, const float *z_vals, const float *d_rgb,
                      const float *d_dist, const float *d_alpha_out, const float *d_weights_out, float *const *d_params,
                      float *d_origin, float *d_dirs, void *ws, size_t ws_bytes, cudaStream_t s);

// parameter indices (reference state_dict order)
enum { P_L0 = 0, P_L1 = 2, P_L2 = 4, P_L3 = 6, P_L4 = 8, P_L5 = 10, P_L6 = 12, P_L7 = 14, P_DEN = 16, P_FEAT = 18, P_RGB0 = 20, P_RGB = 22 };

#define FP32_RENDER_CHUNK 4096  // rays per pass when nothing is saved

namespace {

struct Fp32Buffers {
  float *X0, *H0, *H1, *H2, *Hs, *H3, *H4, *H5, *H8, *raw, *Fc, *Hc, *rgb_s, *alpha_mlp, *z, *w, *G0, *G1, *Genc, *Gdir, *d_rgb_s, *d_alpha;
};

// floats per sample
constexpr size_t kSaveFloats = 64 + 3 * 256 + 320 + 3 * 256 + 256 + 1 + 288 + 128 + 3 + 1;
constexpr size_t kBwdFloats = 256 + 256 + 64 + 32 + 3 + 1;
constexpr size_t kNoSaveFloats = 64 + 320 + 2 * 256 + 1 + 288 + 128 + 3 + 1 + 1 + 1;

void carve_save(float *p, size_t M, Fp32Buffers &b) {
  b.X0 = p; p += M * 64;
  b.H0 = p; p += M * 256; b.H1 = p; p += M * 256; b.H2 = p; p += M * 256;
  b.Hs = p; p += M * 320;
  b.H3 = p; p += M * 256; b.H4 = p; p += M * 256; b.H5 = p; p += M * 256;
  b.H8 = p; p += M * 256;
  b.raw = p; p += M;
  b.Fc = p; p += M * 288;
  b.Hc = p; p += M * 128;
  b.rgb_s = p; p += M * 3;
  b.alpha_mlp = p; p += M;
  b.G0 = p; p += M * 256; b.G1 = p; p += M * 256;
  b.Genc = p; p += M * 64; b.Gdir = p; p += M * 32;
  b.d_rgb_s = p; p += M * 3;
  b.d_alpha = p; p += M;
  b.z = nullptr; b.w = nullptr;
}

void carve_nosave(float *p, size_t M, Fp32Buffers &b) {
  b.X0 = p; p += M * 64;
  b.Hs = p; p += M * 320;
  float *A = p; p += M * 256;
  float *B = p; p += M * 256;
  b.H0 = A; b.H1 = B; b.H2 = A; b.H3 = A; b.H4 = B; b.H5 = A; b.H8 = B;
  b.raw = p; p += M;
  b.Fc = p; p += M * 288;
  b.Hc = p; p += M * 128;
  b.rgb_s = p; p += M * 3;
  b.alpha_mlp = p; p += M;
  b.z = p; p += M;
  b.w = p; p += M;
  b.G0 = b.G1 = b.Genc = b.Gdir = b.d_rgb_s = b.d_alpha = nullptr;
}

inline int ew_grid(long long n) {
  long long g = (n + 255) / 256;
  const long long cap = (long long)nnb_sm_count() * 16;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

#define RET(x) do { int e__ = (x); if (e__) return e__; } while (0)

int linear_fwd(long long M, int N, int K, const float *X, int ldx, const float *W, const float *b, float *Y, int ldy, int relu,
               cudaStream_t s) {
  SgemmEpilogue ep{b, relu, nullptr, 0, 0, 0};
  return sgemm_launch<true, true>(M, N, K, X, ldx, W, K, Y, ldy, ep, 1, s);
}

// dW[N,K] += dY^T X  (split-K over the sample dimension, atomic accumulation)
int wgrad(long long M, int N, int K, const float *dY, int lddy, const float *X, int ldx, float *dW, int ldw, cudaStream_t s) {
  int splits = (int)((M + 4095) / 4096);
  if (splits < 2) splits = 2;  // forces the atomic (+=) epilogue
  SgemmEpilogue we{nullptr, 0, nullptr, 0, 0, 1};
  return sgemm_launch<false, false>(N, K, (int)M, dY, lddy, X, ldx, dW, ldw, we, splits, s);
}
// db[N] += column sums of dY
int bgrad(long long M, int N, const float *dY, int lddy, float *db, cudaStream_t s) {
  int g = (int)((M + 511) / 512);
  const int cap = nnb_sm_count() * 4;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  colsum_kernel<<<g, 256, 0, s>>>(dY, M, N, lddy, db);
  NNB_CHECK_LAUNCH();
  return 0;
}
// dX[M, ncols] (=|+=) dY[M,N] W[N, ncols] (W row stride ldw), optional ReLU mask applied to the (accumulated) result
int dgrad(long long M, int N, int ncols, const float *dY, int lddy, const float *W, int ldw, float *dX, int lddx, const float *mask,
          int ldmask, int accumulate, cudaStream_t s) {
  SgemmEpilogue de{nullptr, 0, mask, ldmask, accumulate, 0};
  return sgemm_launch<true, false>(M, ncols, N, dY, lddy, W, ldw, dX, lddx, de, 1, s);
}

int fp32_forward_chunk(const nnb_render_cfg *cfg, const float *const *P, const float *origin, const float *dirs, const float *ray_o,
                       const float *view, const float *noise,
                       int R, Fp32Buffers &b, float *rgb, float *dist, float *weights, float *alpha, float *z_vals, cudaStream_t s) {
  const int S = cfg->n_samples;
  const long long M = (long long)R * S;
  float *z = z_vals ? z_vals : b.z;
  encode_fwd_kernel<<<(unsigned)((M + 63) / 64), 256, 0, s>>>(origin, dirs, ray_o, view, noise, M, S, cfg->near_, cfg->far_, cfg->use_dir, b.X0, b.Hs,
                                                            b.Fc, z);
  NNB_CHECK_LAUNCH();
  RET(linear_fwd(M, 256, 63, b.X0, 64, P[P_L0], P[P_L0 + 1], b.H0, 256, 1, s));
  RET(linear_fwd(M, 256, 256, b.H0, 256, P[P_L1], P[P_L1 + 1], b.H1, 256, 1, s));
  RET(linear_fwd(M, 256, 256, b.H1, 256, P[P_L2], P[P_L2 + 1], b.H2, 256, 1, s));
  RET(linear_fwd(M, 256, 256, b.H2, 256, P[P_L3], P[P_L3 + 1], b.Hs, 320, 1, s));
  RET(linear_fwd(M, 256, 319, b.Hs, 320, P[P_L4], P[P_L4 + 1], b.H3, 256, 1, s));
  RET(linear_fwd(M, 256, 256, b.H3, 256, P[P_L5], P[P_L5 + 1], b.H4, 256, 1, s));
  RET(linear_fwd(M, 256, 256, b.H4, 256, P[P_L6], P[P_L6 + 1], b.H5, 256, 1, s));
  RET(linear_fwd(M, 256, 256, b.H5, 256, P[P_L7], P[P_L7 + 1], b.H8, 256, 1, s));
  head_density_fwd_kernel<<<ew_grid(M * 32), 256, 0, s>>>(b.H8, M, P[P_DEN], P[P_DEN + 1], cfg->occ_softplus, cfg->dist_alpha, b.raw,
                                                         b.alpha_mlp);
  NNB_CHECK_LAUNCH();
  RET(linear_fwd(M, 256, 256, b.H8, 256, P[P_FEAT], P[P_FEAT + 1], b.Fc, 288, 0, s));
  RET(linear_fwd(M, 128, 283, b.Fc, 288, P[P_RGB0], P[P_RGB0 + 1], b.Hc, 128, 1, s));
  head_rgb_fwd_kernel<<<ew_grid(M * 32), 256, 0, s>>>(b.Hc, M, P[P_RGB], P[P_RGB + 1], b.rgb_s);
  NNB_CHECK_LAUNCH();
  return nnb_composite_fwd_launch(b.alpha_mlp, b.rgb_s, z, R, S, cfg->dist_alpha, cfg->white_bkgd, rgb, dist,
                                  weights ? weights : b.w, alpha, s);
}

}  // namespace

extern "C" {

size_t nnb_render_workspace_bytes(const nnb_render_cfg *cfg, int save_for_backward) {
  if (!cfg || cfg->n_rays <= 0 || cfg->n_samples <= 0) return 0;
  if (cfg->precision == NNB_PREC_BF16) return nnb_tc_workspace_bytes(cfg, save_for_backward);
  if (save_for_backward) return (size_t)cfg->n_rays * cfg->n_samples * (kSaveFloats + kBwdFloats) * sizeof(float);
  const size_t rc = cfg->n_rays < FP32_RENDER_CHUNK ? cfg->n_rays : FP32_RENDER_CHUNK;
  return rc * cfg->n_samples * kNoSaveFloats * sizeof(float);
}

int nnb_render_fwd(const nnb_render_cfg *cfg, const float *const *params_host, const void *packed, const float *origin,
                   const float *dirs, const float *noise, float *rgb, float *dist, float *weights, float *alpha, float *z_vals,
                   void *workspace, size_t workspace_bytes, int save_for_backward, void *stream) {
  return nnb_render_fwd_ex(cfg, params_host, packed, origin, dirs, nullptr, noise, rgb, dist, weights, alpha, z_vals, workspace,
                           workspace_bytes, save_for_backward, stream);
}

int nnb_render_fwd_ex(const nnb_render_cfg *cfg, const float *const *params_host, const void *packed, const float *origin,
                      const float *dirs, const nnb_ray_extras *ex, const float *noise, float *rgb, float *dist, float *weights,
                      float *alpha, float *z_vals, void *workspace, size_t workspace_bytes, int save_for_backward, void *stream) {
  const float *ray_o = ex ? ex->ray_origins : nullptr, *view = ex ? ex->view_dirs : nullptr;
  if (!cfg || !params_host || !origin || !dirs || !rgb || !dist || !workspace) return NNB_ERR_BAD_ARG;
  if (cfg->n_rays <= 0 || cfg->n_samples < 2 || cfg->n_samples > 256) return NNB_ERR_BAD_ARG;
  if (workspace_bytes < nnb_render_workspace_bytes(cfg, save_for_backward)) return NNB_ERR_WORKSPACE;
  cudaStream_t s = nnb_stream(stream);
  if (cfg->precision == NNB_PREC_BF16)
    return nnb_tc_render_fwd(cfg, params_host, packed, origin, dirs, ex, noise, rgb, dist, weights, alpha, z_vals, workspace,
                             workspace_bytes, save_for_backward, s);
  if (cfg->precision != NNB_PREC_FP32) return NNB_ERR_UNSUPPORTED;
  const int S = cfg->n_samples;
  Fp32Buffers b;
  if (save_for_backward) {
    if (!weights || !alpha || !z_vals) return NNB_ERR_BAD_ARG;
    carve_save(static_cast<float *>(workspace), (size_t)cfg->n_rays * S, b);
    return fp32_forward_chunk(cfg, params_host, origin, dirs, ray_o, view, noise, cfg->n_rays, b, rgb, dist, weights, alpha, z_vals, s);
  }
  const int rc = cfg->n_rays < FP32_RENDER_CHUNK ? cfg->n_rays : FP32_RENDER_CHUNK;
  carve_nosave(static_cast<float *>(workspace), (size_t)rc * S, b);
  for (int r0 = 0; r0 < cfg->n_rays; r0 += rc) {
    const int r = cfg->n_rays - r0 < rc ? cfg->n_rays - r0 : rc;
    const size_t o = (size_t)r0 * S;
    RET(fp32_forward_chunk(cfg, params_host, origin, dirs + 3 * (size_t)r0, ray_o ? ray_o + 3 * (size_t)r0 : nullptr,
                           view ? view + 3 * (size_t)r0 : nullptr, noise ? noise + o : nullptr, r, b, rgb + 3 * (size_t)r0,
                           dist + r0, weights ? weights + o : nullptr, alpha ? alpha + o : nullptr, z_vals ? z_vals + o : nullptr, s));
  }
  return 0;
}

int nnb_render_bwd(const nnb_render_cfg *cfg, const float *const *params_host, const void *packed, const float *origin,
                   const float *dirs, const float *weights, const float *alpha, const float *z_vals, const float *d_rgb,
                   const float *d_dist, const float *d_alpha_out, const float *d_weights_out, float *const *d_params_host,
                   float *d_origin, float *d_dirs, void *workspace, size_t workspace_bytes, void *stream) {
  return nnb_render_bwd_ex(cfg, params_host, packed, origin, dirs, nullptr, weights, alpha, z_vals, d_rgb, d_dist, d_alpha_out,
                           d_weights_out, d_params_host, d_origin, d_dirs, workspace, workspace_bytes, stream);
}

int nnb_render_bwd_ex(const nnb_render_cfg *cfg, const float *const *params_host, const void *packed, const float *origin,
                      const float *dirs, const nnb_ray_extras *ex, const float *weights, const float *alpha, const float *z_vals,
                      const float *d_rgb, const float *d_dist, const float *d_alpha_out, const float *d_weights_out,
                      float *const *d_params_host, float *d_origin, float *d_dirs, void *workspace, size_t workspace_bytes,
                      void *stream) {
  if (ex && ((ex->ray_origins && !ex->d_ray_origins) || (ex->view_dirs && !ex->d_view_dirs))) return NNB_ERR_BAD_ARG;
  if (!cfg || !params_host || !d_params_host || !origin || !dirs || !z_vals || !d_origin || !d_dirs || !workspace) return NNB_ERR_BAD_ARG;
  if (workspace_bytes < nnb_render_workspace_bytes(cfg, 1)) return NNB_ERR_WORKSPACE;
  cudaStream_t s = nnb_stream(stream);
  if (cfg->precision == NNB_PREC_BF16)
    return nnb_tc_render_bwd(cfg, params_host, packed, origin, dirs, ex, weights, alpha, z_vals, d_rgb, d_dist, d_alpha_out,
                             d_weights_out, d_params_host, d_origin, d_dirs, workspace, workspace_bytes, s);
  if (cfg->precision != NNB_PREC_FP32) return NNB_ERR_UNSUPPORTED;
  const int R = cfg->n_rays, S = cfg->n_samples;
  const long long M = (long long)R * S;
  const float *const *P = params_host;
  float *const *D = d_params_host;
  Fp32Buffers b;
  carve_save(static_cast<float *>(workspace), (size_t)M, b);
  NNB_CUDA(cudaMemsetAsync(d_origin, 0, 3 * sizeof(float), s));
  // stage 3 backward -> d_alpha [M], d_rgb_s [M,3]
  RET(nnb_composite_bwd_launch(b.alpha_mlp, b.rgb_s, z_vals, R, S, cfg->dist_alpha, cfg->white_bkgd, d_rgb, d_dist, d_weights_out,
                               d_alpha_out, b.d_alpha, b.d_rgb_s, s));
  // fc_rgb (128 -> 3, sigmoid)
  sigmoid_bwd_kernel<<<ew_grid(M * 3), 256, 0, s>>>(b.rgb_s, M * 3, b.d_rgb_s);
  NNB_CHECK_LAUNCH();
  RET(wgrad(M, 3, 128, b.d_rgb_s, 3, b.Hc, 128, D[P_RGB], 128, s));
  RET(bgrad(M, 3, b.d_rgb_s, 3, D[P_RGB + 1], s));
  RET(dgrad(M, 3, 128, b.d_rgb_s, 3, P[P_RGB], 128, b.G0, 128, b.Hc, 128, 0, s));
  // rgb_layers.0 ([feature | dir enc] 283 -> 128, relu)
  RET(wgrad(M, 128, 283, b.G0, 128, b.Fc, 288, D[P_RGB0], 283, s));
  RET(bgrad(M, 128, b.G0, 128, D[P_RGB0 + 1], s));
  RET(dgrad(M, 128, 256, b.G0, 128, P[P_RGB0], 283, b.G1, 256, nullptr, 0, 0, s));
  RET(dgrad(M, 128, 27, b.G0, 128, P[P_RGB0] + 256, 283, b.Gdir, 32, nullptr, 0, 0, s));
  // fc_density (256 -> 1): d_raw in place, G0 = d_raw (x) w
  head_density_bwd_kernel<<<ew_grid(M), 256, 0, s>>>(b.raw, M, cfg->occ_softplus, cfg->dist_alpha, b.d_alpha);
  NNB_CHECK_LAUNCH();
  RET(wgrad(M, 1, 256, b.d_alpha, 1, b.H8, 256, D[P_DEN], 256, s));
  RET(bgrad(M, 1, b.d_alpha, 1, D[P_DEN + 1], s));
  RET(dgrad(M, 1, 256, b.d_alpha, 1, P[P_DEN], 256, b.G0, 256, nullptr, 0, 0, s));
  // fc_feature (256 -> 256, linear): G0 = relu'(H8) * (G0 + G1 Wf)
  RET(wgrad(M, 256, 256, b.G1, 256, b.H8, 256, D[P_FEAT], 256, s));
  RET(bgrad(M, 256, b.G1, 256, D[P_FEAT + 1], s));
  RET(dgrad(M, 256, 256, b.G1, 256, P[P_FEAT], 256, b.G0, 256, b.H8, 256, 1, s));
  // layers1.6 / .4 / .2
  RET(wgrad(M, 256, 256, b.G0, 256, b.H5, 256, D[P_L7], 256, s));
  RET(bgrad(M, 256, b.G0, 256, D[P_L7 + 1], s));
  RET(dgrad(M, 256, 256, b.G0, 256, P[P_L7], 256, b.G1, 256, b.H5, 256, 0, s));
  RET(wgrad(M, 256, 256, b.G1, 256, b.H4, 256, D[P_L6], 256, s));
  RET(bgrad(M, 256, b.G1, 256, D[P_L6 + 1], s));
  RET(dgrad(M, 256, 256, b.G1, 256, P[P_L6], 256, b.G0, 256, b.H4, 256, 0, s));
  RET(wgrad(M, 256, 256, b.G0, 256, b.H3, 256, D[P_L5], 256, s));
  RET(bgrad(M, 256, b.G0, 256, D[P_L5 + 1], s));
  RET(dgrad(M, 256, 256, b.G0, 256, P[P_L5], 256, b.G1, 256, b.H3, 256, 0, s));
  // layers1.0 (skip: [h | enc] 319 -> 256)
  RET(wgrad(M, 256, 319, b.G1, 256, b.Hs, 320, D[P_L4], 319, s));
  RET(bgrad(M, 256, b.G1, 256, D[P_L4 + 1], s));
  RET(dgrad(M, 256, 256, b.G1, 256, P[P_L4], 319, b.G0, 256, b.Hs, 320, 0, s));
  RET(dgrad(M, 256, 63, b.G1, 256, P[P_L4] + 256, 319, b.Genc, 64, nullptr, 0, 0, s));
  // layers0.6 / .4 / .2
  RET(wgrad(M, 256, 256, b.G0, 256, b.H2, 256, D[P_L3], 256, s));
  RET(bgrad(M, 256, b.G0, 256, D[P_L3 + 1], s));
  RET(dgrad(M, 256, 256, b.G0, 256, P[P_L3], 256, b.G1, 256, b.H2, 256, 0, s));
  RET(wgrad(M, 256, 256, b.G1, 256, b.H1, 256, D[P_L2], 256, s));
  RET(bgrad(M, 256, b.G1, 256, D[P_L2 + 1], s));
  RET(dgrad(M, 256, 256, b.G1, 256, P[P_L2], 256, b.G0, 256, b.H1, 256, 0, s));
  RET(wgrad(M, 256, 256, b.G0, 256, b.H0, 256, D[P_L1], 256, s));
  RET(bgrad(M, 256, b.G0, 256, D[P_L1 + 1], s));
  RET(dgrad(M, 256, 256, b.G0, 256, P[P_L1], 256, b.G1, 256, b.H0, 256, 0, s));
  // layers0.0 (enc 63 -> 256): input gradient accumulates onto the skip path's
  RET(wgrad(M, 256, 63, b.G1, 256, b.X0, 64, D[P_L0], 63, s));
  RET(bgrad(M, 256, b.G1, 256, D[P_L0 + 1], s));
  RET(dgrad(M, 256, 63, b.G1, 256, P[P_L0], 63, b.Genc, 64, nullptr, 0, 1, s));
  // encoding + sampling backward -> d_origin, d_dirs
  const int threads = ((S + 31) / 32) * 32;
  encode_bwd_kernel<<<R, threads, 0, s>>>(b.X0, b.Genc, 64, b.Gdir, 32, dirs, ex ? ex->view_dirs : nullptr, z_vals, S, cfg->use_dir, d_origin,
                                          d_dirs, (ex && ex->ray_origins) ? ex->d_ray_origins : nullptr, ex ? ex->d_view_dirs : nullptr);
  NNB_CHECK_LAUNCH();
  (void)weights; (void)alpha; (void)origin;
  return 0;
}

#define POINTS_CHUNK (1 << 19)
size_t nnb_mlp_points_workspace_bytes(long long n_points) {
  const size_t m = n_points < POINTS_CHUNK ? (n_points > 0 ? n_points : 1) : POINTS_CHUNK;
  return m * kNoSaveFloats * sizeof(float);
}

int nnb_mlp_points_fwd(const float *const *P, const float *pts, const float *viewdirs, long long n_points, int occ_softplus,
                       int dist_alpha, float *rgb, float *dens, void *workspace, size_t workspace_bytes, void *stream) {
  if (!P || !pts || !dens || !workspace || n_points <= 0) return NNB_ERR_BAD_ARG;
  if (rgb && !viewdirs) return NNB_ERR_BAD_ARG;
  if (workspace_bytes < nnb_mlp_points_workspace_bytes(n_points)) return NNB_ERR_WORKSPACE;
  cudaStream_t s = nnb_stream(stream);
  const long long chunk = n_points < POINTS_CHUNK ? n_points : POINTS_CHUNK;
  Fp32Buffers b;
  carve_nosave(static_cast<float *>(workspace), (size_t)chunk, b);
  for (long long m0 = 0; m0 < n_points; m0 += chunk) {
    const long long M = n_points - m0 < chunk ? n_points - m0 : chunk;
    encode_points_fwd_kernel<<<ew_grid(M * 3), 256, 0, s>>>(pts + 3 * m0, viewdirs ? viewdirs + 3 * m0 : nullptr, M, b.X0, b.Hs, b.Fc);
    NNB_CHECK_LAUNCH();
    RET(linear_fwd(M, 256, 63, b.X0, 64, P[P_L0], P[P_L0 + 1], b.H0, 256, 1, s));
    RET(linear_fwd(M, 256, 256, b.H0, 256, P[P_L1], P[P_L1 + 1], b.H1, 256, 1, s));
    RET(linear_fwd(M, 256, 256, b.H1, 256, P[P_L2], P[P_L2 + 1], b.H2, 256, 1, s));
    RET(linear_fwd(M, 256, 256, b.H2, 256, P[P_L3], P[P_L3 + 1], b.Hs, 320, 1, s));
    RET(linear_fwd(M, 256, 319, b.Hs, 320, P[P_L4], P[P_L4 + 1], b.H3, 256, 1, s));
    RET(linear_fwd(M, 256, 256, b.H3, 256, P[P_L5], P[P_L5 + 1], b.H4, 256, 1, s));
    RET(linear_fwd(M, 256, 256, b.H4, 256, P[P_L6], P[P_L6 + 1], b.H5, 256, 1, s));
    RET(linear_fwd(M, 256, 256, b.H5, 256, P[P_L7], P[P_L7 + 1], b.H8, 256, 1, s));
    head_density_fwd_kernel<<<ew_grid(M * 32), 256, 0, s>>>(b.H8, M, P[P_DEN], P[P_DEN + 1], occ_softplus, dist_alpha, b.raw, dens + m0);
    NNB_CHECK_LAUNCH();
    if (rgb) {
      RET(linear_fwd(M, 256, 256, b.H8, 256, P[P_FEAT], P[P_FEAT + 1], b.Fc, 288, 0, s));
      RET(linear_fwd(M, 128, 283, b.Fc, 288, P[P_RGB0], P[P_RGB0 + 1], b.Hc, 128, 1, s));
      head_rgb_fwd_kernel<<<ew_grid(M * 32), 256, 0, s>>>(b.Hc, M, P[P_RGB], P[P_RGB + 1], rgb + 3 * m0);
      NNB_CHECK_LAUNCH();
    }
  }
  return 0;
}

// Workspace of nnb_mlp_points_density_grad: saved activations + two gradient buffers for one chunk of points.
#define POINTS_GRAD_CHUNK (1 << 17)
size_t nnb_mlp_points_grad_workspace_bytes(long long n_points) {
  const size_t m = n_points < POINTS_GRAD_CHUNK ? (n_points > 0 ? n_points : 1) : POINTS_GRAD_CHUNK;
  return m * (kSaveFloats + kBwdFloats) * sizeof(float);
}

// OfficialStaticNerf.gradient (model/official_nerf.py:46-58): grad (M,3) = -d fc_density(trunk(p)) / d p  (raw density, before
// the activation; grad_outputs = ones), optionally also the raw density.  fp32 CUDA-core path: trunk forward with saved
// activations, then the data-gradient chain only (no weight gradients).
int nnb_mlp_points_density_grad(const float *const *P, const float *pts, long long n_points, float *grad, float *raw_out,
                                void *workspace, size_t workspace_bytes, void *stream) {
  if (!P || !pts || !grad || !workspace || n_points <= 0) return NNB_ERR_BAD_ARG;
  if (workspace_bytes < nnb_mlp_points_grad_workspace_bytes(n_points)) return NNB_ERR_WORKSPACE;
  cudaStream_t s = nnb_stream(stream);
  const long long chunk = n_points < POINTS_GRAD_CHUNK ? n_points : POINTS_GRAD_CHUNK;
  Fp32Buffers b;
  carve_save(static_cast<float *>(workspace), (size_t)chunk, b);
  for (long long m0 = 0; m0 < n_points; m0 += chunk) {
    const long long M = n_points - m0 < chunk ? n_points - m0 : chunk;
    encode_points_fwd_kernel<<<ew_grid(M * 3), 256, 0, s>>>(pts + 3 * m0, nullptr, M, b.X0, b.Hs, b.Fc);
    NNB_CHECK_LAUNCH();
    RET(linear_fwd(M, 256, 63, b.X0, 64, P[P_L0], P[P_L0 + 1], b.H0, 256, 1, s));
    RET(linear_fwd(M, 256, 256, b.H0, 256, P[P_L1], P[P_L1 + 1], b.H1, 256, 1, s));
    RET(linear_fwd(M, 256, 256, b.H1, 256, P[P_L2], P[P_L2 + 1], b.H2, 256, 1, s));
    RET(linear_fwd(M, 256, 256, b.H2, 256, P[P_L3], P[P_L3 + 1], b.Hs, 320, 1, s));
    RET(linear_fwd(M, 256, 319, b.Hs, 320, P[P_L4], P[P_L4 + 1], b.H3, 256, 1, s));
    RET(linear_fwd(M, 256, 256, b.H3, 256, P[P_L5], P[P_L5 + 1], b.H4, 256, 1, s));
    RET(linear_fwd(M, 256, 256, b.H4, 256, P[P_L6], P[P_L6 + 1], b.H5, 256, 1, s));
    RET(linear_fwd(M, 256, 256, b.H5, 256, P[P_L7], P[P_L7 + 1], b.H8, 256, 1, s));
    if (raw_out) {
      head_density_fwd_kernel<<<ew_grid(M * 32), 256, 0, s>>>(b.H8, M, P[P_DEN], P[P_DEN + 1], 1, 1, raw_out + m0, b.alpha_mlp);
      NNB_CHECK_LAUNCH();
    }
    // d_raw = 1 -> G0 = relu'(H8) * w_density
    fill_kernel<<<ew_grid(M), 256, 0, s>>>(b.d_alpha, M, 1.0f);
    NNB_CHECK_LAUNCH();
    RET(dgrad(M, 1, 256, b.d_alpha, 1, P[P_DEN], 256, b.G0, 256, b.H8, 256, 0, s));
    RET(dgrad(M, 256, 256, b.G0, 256, P[P_L7], 256, b.G1, 256, b.H5, 256, 0, s));
    RET(dgrad(M, 256, 256, b.G1, 256, P[P_L6], 256, b.G0, 256, b.H4, 256, 0, s));
    RET(dgrad(M, 256, 256, b.G0, 256, P[P_L5], 256, b.G1, 256, b.H3, 256, 0, s));
    RET(dgrad(M, 256, 256, b.G1, 256, P[P_L4], 319, b.G0, 256, b.Hs, 320, 0, s));
    RET(dgrad(M, 256, 63, b.G1, 256, P[P_L4] + 256, 319, b.Genc, 64, nullptr, 0, 0, s));
    RET(dgrad(M, 256, 256, b.G0, 256, P[P_L3], 256, b.G1, 256, b.H2, 256, 0, s));
    RET(dgrad(M, 256, 256, b.G1, 256, P[P_L2], 256, b.G0, 256, b.H1, 256, 0, s));
    RET(dgrad(M, 256, 256, b.G0, 256, P[P_L1], 256, b.G1, 256, b.H0, 256, 0, s));
    RET(dgrad(M, 256, 63, b.G1, 256, P[P_L0], 63, b.Genc, 64, nullptr, 0, 1, s));
    encode_points_bwd_kernel<<<ew_grid(M * 3), 256, 0, s>>>(b.X0, b.Genc, M, -1.0f, grad + 3 * m0);
    NNB_CHECK_LAUNCH();
  }
  return 0;
}

}  // extern "C"
