// Stage 1a: pose / 4x4 / distortion single-thread kernels.
// The reference spends ~15 tiny launches per LearnPose.forward; each of these
// is one launch of one thread, latency-bound by construction.
#include "pose_math.cuh"

__global__ void pose_fwd_kernel(const float *__restrict__ r, const float *__restrict__ t,
                                const float *__restrict__ init, int cam, float *__restrict__ c2w) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  double rr[3], tt[3], in[16], out[16];
  for (int i = 0; i < 3; ++i) { rr[i] = r[cam * 3 + i]; tt[i] = t[cam * 3 + i]; }
  if (init) for (int i = 0; i < 16; ++i) in[i] = init[cam * 16 + i];
  pose_forward(rr, tt, init ? in : nullptr, out);
  for (int i = 0; i < 16; ++i) c2w[i] = (float)out[i];
}

__global__ void pose_bwd_kernel(const float *__restrict__ r, const float *__restrict__ t,
                                const float *__restrict__ init, int cam, const float *__restrict__ dc2w,
                                float *__restrict__ dr, float *__restrict__ dt) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  double rr[3], tt[3], in[16], dout[16], gr[3], gt[3];
  for (int i = 0; i < 3; ++i) { rr[i] = r[cam * 3 + i]; tt[i] = t[cam * 3 + i]; }
  if (init) for (int i = 0; i < 16; ++i) in[i] = init[cam * 16 + i];
  for (int i = 0; i < 16; ++i) dout[i] = dc2w[i];
  pose_backward(rr, tt, init ? in : nullptr, dout, gr, gt);
  for (int i = 0; i < 3; ++i) { dr[i] = (float)gr[i]; dt[i] = (float)gt[i]; }
}

__global__ void mat4_inv_fwd_kernel(const float *__restrict__ m, float *__restrict__ o) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  double a[16], b[16];
  for (int i = 0; i < 16; ++i) a[i] = m[i];
  m4_inv(a, b);
  for (int i = 0; i < 16; ++i) o[i] = (float)b[i];
}

__global__ void mat4_inv_bwd_kernel(const float *__restrict__ inv, const float *__restrict__ dinv, float *__restrict__ dm) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  double a[16], b[16], c[16];
  for (int i = 0; i < 16; ++i) { a[i] = inv[i]; b[i] = dinv[i]; }
  m4_inv_bwd(a, b, c);
  for (int i = 0; i < 16; ++i) dm[i] = (float)c[i];
}

// C = A B on 4x4 row-major matrices (the relative pose ref_Rt @ inverse(world_mat), model/training.py:329,344), one thread per
// element, no shared memory: unlike the library GEMM it replaces it can share an SM with the persistent render kernels.
__global__ void mat4_mul_fwd_kernel(const float *__restrict__ a, const float *__restrict__ b, float *__restrict__ c) {
  const int i = threadIdx.x >> 2, j = threadIdx.x & 3;
  if (threadIdx.x >= 16 || blockIdx.x != 0) return;
  float acc = 0.f;
#pragma unroll
  for (int k = 0; k < 4; ++k) acc = fmaf(a[4 * i + k], b[4 * k + j], acc);
  c[threadIdx.x] = acc;
}

// dA = dC B^T (threads 0-15), dB = A^T dC (threads 16-31)
__global__ void mat4_mul_bwd_kernel(const float *__restrict__ a, const float *__restrict__ b, const float *__restrict__ dc,
                                    float *__restrict__ da, float *__restrict__ db) {
  if (blockIdx.x != 0 || threadIdx.x >= 32) return;
  const int e = threadIdx.x & 15, i = e >> 2, j = e & 3;
  float acc = 0.f;
  if (threadIdx.x < 16) {
#pragma unroll
    for (int k = 0; k < 4; ++k) acc = fmaf(dc[4 * i + k], b[4 * j + k], acc);
    if (da) da[e] = acc;
  } else {
#pragma unroll
    for (int k = 0; k < 4; ++k) acc = fmaf(a[4 * k + i], dc[4 * k + j], acc);
    if (db) db[e] = acc;
  }
}

// model/distortions.py:19-27
__global__ void distortion_fwd_kernel(const float *__restrict__ scales, const float *__restrict__ shifts, int cam,
                                      int n, int fix, float *__restrict__ out) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  float s = scales[cam];
  if (s < 0.01f) s = 0.01f;
  if (fix && cam == n - 1) s = 1.0f;
  out[0] = s;
  out[1] = shifts[cam];
}

__global__ void distortion_bwd_kernel(const float *__restrict__ scales, int cam, int n, int fix,
                                      const float *__restrict__ dout, float *__restrict__ ds, float *__restrict__ dsh) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const bool cut = (scales[cam] < 0.01f) || (fix && cam == n - 1);
  ds[0] = cut ? 0.f : dout[0];
  dsh[0] = dout[1];
}

extern "C" {

int nnb_pose_fwd(const float *r, const float *t, const float *init_c2w, int cam_id, float *c2w, void *stream) {
  if (!r || !t || !c2w || cam_id < 0) return NNB_ERR_BAD_ARG;
  pose_fwd_kernel<<<1, 32, 0, nnb_stream(stream)>>>(r, t, init_c2w, cam_id, c2w);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_pose_bwd(const float *r, const float *t, const float *init_c2w, int cam_id, const float *d_c2w,
                 float *d_r_row, float *d_t_row, void *stream) {
  if (!r || !t || !d_c2w || !d_r_row || !d_t_row || cam_id < 0) return NNB_ERR_BAD_ARG;
  pose_bwd_kernel<<<1, 32, 0, nnb_stream(stream)>>>(r, t, init_c2w, cam_id, d_c2w, d_r_row, d_t_row);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_mat4_inv_fwd(const float *m, float *inv, void *stream) {
  if (!m || !inv) return NNB_ERR_BAD_ARG;
  mat4_inv_fwd_kernel<<<1, 32, 0, nnb_stream(stream)>>>(m, inv);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_mat4_inv_bwd(const float *inv, const float *d_inv, float *d_m, void *stream) {
  if (!inv || !d_inv || !d_m) return NNB_ERR_BAD_ARG;
  mat4_inv_bwd_kernel<<<1, 32, 0, nnb_stream(stream)>>>(inv, d_inv, d_m);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_mat4_mul_fwd(const float *a, const float *b, float *c, void *stream) {
  if (!a || !b || !c) return NNB_ERR_BAD_ARG;
  mat4_mul_fwd_kernel<<<1, 32, 0, nnb_stream(stream)>>>(a, b, c);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_mat4_mul_bwd(const float *a, const float *b, const float *d_c, float *d_a, float *d_b, void *stream) {
  if (!a || !b || !d_c || (!d_a && !d_b)) return NNB_ERR_BAD_ARG;
  mat4_mul_bwd_kernel<<<1, 32, 0, nnb_stream(stream)>>>(a, b, d_c, d_a, d_b);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_distortion_fwd(const float *scales, const float *shifts, int cam_id, int n_cams, int fix_scaleN,
                       float *out2, void *stream) {
  if (!scales || !shifts || !out2 || cam_id < 0 || cam_id >= n_cams) return NNB_ERR_BAD_ARG;
  distortion_fwd_kernel<<<1, 32, 0, nnb_stream(stream)>>>(scales, shifts, cam_id, n_cams, fix_scaleN, out2);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_distortion_bwd(const float *scales, int cam_id, int n_cams, int fix_scaleN, const float *d_out2,
                       float *d_scale_row, float *d_shift_row, void *stream) {
  if (!scales || !d_out2 || !d_scale_row || !d_shift_row || cam_id < 0 || cam_id >= n_cams) return NNB_ERR_BAD_ARG;
  distortion_bwd_kernel<<<1, 32, 0, nnb_stream(stream)>>>(scales, cam_id, n_cams, fix_scaleN, d_out2, d_scale_row, d_shift_row);
  NNB_CHECK_LAUNCH();
  return 0;
}

}  // extern "C"
