// Stage 4b: surface-based reprojection sampling (model/training.py:364-382).
#include "nnb_common.cuh"

static inline int pt_grid(int n, int threads) {
  int g = nnb_cdiv(n, threads);
  const int cap = nnb_sm_count() * 8;
  return g < 1 ? 1 : (g > cap ? cap : g);
}

// F.interpolate(mode='bilinear', align_corners=False, antialias=False): model/training.py:365-366
__global__ void resize_bilinear_kernel(const float *__restrict__ img, int c, int h, int w, int h4, int w4, float *__restrict__ out) {
  const float sy = (float)h / (float)h4, sx = (float)w / (float)w4;
  const int n = c * h4 * w4;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int x4 = i % w4, y4 = (i / w4) % h4, ch = i / (w4 * h4);
    float fy = sy * ((float)y4 + 0.5f) - 0.5f; if (fy < 0.f) fy = 0.f;
    float fx = sx * ((float)x4 + 0.5f) - 0.5f; if (fx < 0.f) fx = 0.f;
    const int y0 = (int)fy, x0 = (int)fx;
    const int y1 = y0 + (y0 < h - 1 ? 1 : 0), x1 = x0 + (x0 < w - 1 ? 1 : 0);
    const float ly = fy - (float)y0, lx = fx - (float)x0;
    const float *p = img + (size_t)ch * h * w;
    const float top = (1.f - lx) * p[y0 * w + x0] + lx * p[y0 * w + x1];
    const float bot = (1.f - lx) * p[y1 * w + x0] + lx * p[y1 * w + x1];
    out[i] = (1.f - ly) * top + ly * bot;
  }
}

// grid_sample(bilinear, padding zeros, align_corners=True) of a (3,h,w) image at normalised (x,y).
// Returns the three channels; optionally d/dx, d/dy (in normalised coordinates) of sum_c g_c * value_c.
__device__ __forceinline__ void sample3(const float *__restrict__ img, int h, int w, float x, float y, float out[3],
                                        const float *g, float *dx, float *dy) {
  const float ix = ((x + 1.f) * 0.5f) * (float)(w - 1);
  const float iy = ((y + 1.f) * 0.5f) * (float)(h - 1);
  const float fx0 = floorf(ix), fy0 = floorf(iy);
  const int x0 = (int)fx0, y0 = (int)fy0, x1 = x0 + 1, y1 = y0 + 1;
  const float tx = ix - fx0, ty = iy - fy0;
  const float w00 = (1.f - tx) * (1.f - ty), w01 = tx * (1.f - ty), w10 = (1.f - tx) * ty, w11 = tx * ty;
  const bool vx0 = x0 >= 0 && x0 < w, vx1 = x1 >= 0 && x1 < w, vy0 = y0 >= 0 && y0 < h, vy1 = y1 >= 0 && y1 < h;
  float gx = 0.f, gy = 0.f;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float *p = img + (size_t)c * h * w;
    const float v00 = (vx0 && vy0) ? p[y0 * w + x0] : 0.f;
    const float v01 = (vx1 && vy0) ? p[y0 * w + x1] : 0.f;
    const float v10 = (vx0 && vy1) ? p[y1 * w + x0] : 0.f;
    const float v11 = (vx1 && vy1) ? p[y1 * w + x1] : 0.f;
    out[c] = v00 * w00 + v01 * w01 + v10 * w10 + v11 * w11;
    if (g) {
      gx += g[c] * ((v01 - v00) * (1.f - ty) + (v11 - v10) * ty);
      gy += g[c] * ((v10 - v00) * (1.f - tx) + (v11 - v01) * tx);
    }
  }
  if (dx) *dx = gx * 0.5f * (float)(w - 1);
  if (dy) *dy = gy * 0.5f * (float)(h - 1);
}

__device__ __forceinline__ float grid_px(int x, int w) { return __fsub_rn(__fdiv_rn(2.0f * (float)x, (float)(w - 1)), 1.0f); }

__global__ void reproj_fwd_kernel(const float *__restrict__ img1, const float *__restrict__ img2, int h4, int w4,
                                  const float *__restrict__ pc1, const float *__restrict__ Rm, const float *__restrict__ tv,
                                  const float *__restrict__ K, float nl, float *__restrict__ rgb1, float *__restrict__ rgb2,
                                  uint8_t *__restrict__ valid) {
  __shared__ float sR[9], sT[3], sK[12];
  if (threadIdx.x < 9) sR[threadIdx.x] = Rm[threadIdx.x];
  if (threadIdx.x < 3) sT[threadIdx.x] = tv[threadIdx.x];
  if (threadIdx.x < 12) sK[threadIdx.x] = K[threadIdx.x];
  __syncthreads();
  const int n = h4 * w4;
  for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
    const int y4 = j / w4, x4 = j - y4 * w4;
    float o[3];
    if (rgb1) {
      sample3(img1, h4, w4, grid_px(x4, w4), grid_px(y4, h4), o, nullptr, nullptr, nullptr);
      rgb1[3 * j] = o[0]; rgb1[3 * j + 1] = o[1]; rgb1[3 * j + 2] = o[2];
    }
    const float p0 = pc1[3 * j], p1 = pc1[3 * j + 1], p2 = pc1[3 * j + 2];
    float r[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) r[c] = p0 * sR[c * 3] + p1 * sR[c * 3 + 1] + p2 * sR[c * 3 + 2] + sT[c];
    if (-r[2] < nl) { r[0] = nl; r[1] = nl; r[2] = nl; }  // model/training.py:373-374 (mask expanded over xyz)
    float q[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) q[c] = sK[c * 4] * r[0] + sK[c * 4 + 1] * r[1] + sK[c * 4 + 2] * r[2] + sK[c * 4 + 3];
    const float x = q[0] / q[2], y = q[1] / q[2];
    valid[j] = (fmaxf(fabsf(x), fabsf(y)) <= 1.f) ? 1 : 0;
    sample3(img2, h4, w4, x, y, o, nullptr, nullptr, nullptr);
    rgb2[3 * j] = o[0]; rgb2[3 * j + 1] = o[1]; rgb2[3 * j + 2] = o[2];
  }
}

// acc12 (device, zeroed by the host wrapper) = { dR (9), dt (3) }
__global__ void reproj_bwd_kernel(const float *__restrict__ img2, int h4, int w4, const float *__restrict__ pc1,
                                  const float *__restrict__ Rm, const float *__restrict__ tv, const float *__restrict__ K,
                                  float nl, const float *__restrict__ d_rgb2, float *__restrict__ d_pc1, float *__restrict__ acc12) {
  __shared__ float sR[9], sT[3], sK[12];
  __shared__ float red[32];
  if (threadIdx.x < 9) sR[threadIdx.x] = Rm[threadIdx.x];
  if (threadIdx.x < 3) sT[threadIdx.x] = tv[threadIdx.x];
  if (threadIdx.x < 12) sK[threadIdx.x] = K[threadIdx.x];
  __syncthreads();
  const int n = h4 * w4;
  float acc[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) acc[k] = 0.f;
  for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
    const float p[3] = {pc1[3 * j], pc1[3 * j + 1], pc1[3 * j + 2]};
    float r[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) r[c] = p[0] * sR[c * 3] + p[1] * sR[c * 3 + 1] + p[2] * sR[c * 3 + 2] + sT[c];
    const bool behind = -r[2] < nl;
    if (behind) { r[0] = nl; r[1] = nl; r[2] = nl; }
    float q[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) q[c] = sK[c * 4] * r[0] + sK[c * 4 + 1] * r[1] + sK[c * 4 + 2] * r[2] + sK[c * 4 + 3];
    const float x = q[0] / q[2], y = q[1] / q[2];
    float o[3], dx, dy;
    const float g[3] = {d_rgb2[3 * j], d_rgb2[3 * j + 1], d_rgb2[3 * j + 2]};
    sample3(img2, h4, w4, x, y, o, g, &dx, &dy);
    float dr[3] = {0.f, 0.f, 0.f};
    if (!behind) {
      const float iq = 1.f / q[2];
      const float dq[3] = {dx * iq, dy * iq, -(dx * x + dy * y) * iq};
#pragma unroll
      for (int k = 0; k < 3; ++k) dr[k] = sK[0 * 4 + k] * dq[0] + sK[1 * 4 + k] * dq[1] + sK[2 * 4 + k] * dq[2];
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) d_pc1[3 * j + k] = sR[0 * 3 + k] * dr[0] + sR[1 * 3 + k] * dr[1] + sR[2 * 3 + k] * dr[2];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
#pragma unroll
      for (int k = 0; k < 3; ++k) acc[c * 3 + k] += dr[c] * p[k];
      acc[9 + c] += dr[c];
    }
  }
#pragma unroll
  for (int k = 0; k < 12; ++k) {
    const float s = block_sum(acc[k], red);
    if (threadIdx.x == 0) atomicAdd(acc12 + k, s);
  }
}

extern "C" {

int nnb_resize_bilinear(const float *img, int c, int h, int w, int h4, int w4, float *out, void *stream) {
  if (!img || !out || c < 1 || h < 1 || w < 1 || h4 < 1 || w4 < 1) return NNB_ERR_BAD_ARG;
  resize_bilinear_kernel<<<pt_grid(c * h4 * w4, 256), 256, 0, nnb_stream(stream)>>>(img, c, h, w, h4, w4, out);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_reproj_fwd(const float *img1r, const float *img2r, int h4, int w4, const float *pc1, const float *R_rel,
                   const float *t_rel, const float *camera_mat, float nearest_limit, float *rgb_pc1, float *rgb_pc1_proj,
                   uint8_t *valid, void *stream) {
  if (!img2r || !pc1 || !R_rel || !t_rel || !camera_mat || !rgb_pc1_proj || !valid || h4 < 2 || w4 < 2) return NNB_ERR_BAD_ARG;
  if (rgb_pc1 && !img1r) return NNB_ERR_BAD_ARG;
  reproj_fwd_kernel<<<pt_grid(h4 * w4, 128), 128, 0, nnb_stream(stream)>>>(img1r, img2r, h4, w4, pc1, R_rel, t_rel, camera_mat,
                                                                          nearest_limit, rgb_pc1, rgb_pc1_proj, valid);
  NNB_CHECK_LAUNCH();
  return 0;
}

// d_R_rel doubles as the 12-float accumulator when d_t_rel == d_R_rel + 9; otherwise a two-step copy is used.
int nnb_reproj_bwd(const float *img2r, int h4, int w4, const float *pc1, const float *R_rel, const float *t_rel,
                   const float *camera_mat, float nearest_limit, const float *d_rgb_pc1_proj, float *d_pc1, float *d_R_rel,
                   float *d_t_rel, void *stream) {
  if (!img2r || !pc1 || !R_rel || !t_rel || !camera_mat || !d_rgb_pc1_proj || !d_pc1 || !d_R_rel || !d_t_rel) return NNB_ERR_BAD_ARG;
  if (d_t_rel != d_R_rel + 9) return NNB_ERR_BAD_ARG;  // caller passes one 12-float buffer: [dR(9) | dt(3)]
  cudaStream_t s = nnb_stream(stream);
  NNB_CUDA(cudaMemsetAsync(d_R_rel, 0, 12 * sizeof(float), s));
  reproj_bwd_kernel<<<pt_grid(h4 * w4, 128), 128, 0, s>>>(img2r, h4, w4, pc1, R_rel, t_rel, camera_mat, nearest_limit,
                                                         d_rgb_pc1_proj, d_pc1, d_R_rel);
  NNB_CHECK_LAUNCH();
  return 0;
}

}  // extern "C"
