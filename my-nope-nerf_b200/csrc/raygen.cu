// Stage 1b: pixel grid, ray selection, ray generation, point-cloud back-projection.
// HBM/latency-bound elementwise kernels: one thread per ray/point, coalesced SoA-ish access,
// block-level shuffle reductions for the handful of matrix/scalar gradients.
#include "nnb_common.cuh"
#include <string.h>

// ---------------------------------------------------------------------------------------------
// arange_pixels: model/common.py:13-40
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float scaled_px(int x, int w) { return __fsub_rn(__fdiv_rn(2.0f * (float)x, (float)(w - 1)), 1.0f); }

__global__ void pixel_grid_kernel(int h, int w, int64_t *__restrict__ loc, float *__restrict__ scaled) {
  const int n = h * w;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int y = i / w, x = i - y * w;
    if (loc) { loc[2 * i] = x; loc[2 * i + 1] = y; }
    if (scaled) { scaled[2 * i] = scaled_px(x, w); scaled[2 * i + 1] = scaled_px(y, h); }
  }
}

// ---------------------------------------------------------------------------------------------
// ray selection: model/training.py:264,285-288, model/network.py:21-24
// ---------------------------------------------------------------------------------------------
__global__ void gather_rays_fwd_kernel(const int64_t *__restrict__ ray_idx, int n, int h, int w,
                                       const float *__restrict__ img, const float *__restrict__ depth_img,
                                       const float *__restrict__ ss, float *__restrict__ pixels,
                                       float *__restrict__ rgb_gt, float *__restrict__ depth) {
  const float sc = ss ? ss[0] : 1.f, sh = ss ? ss[1] : 0.f;
  const int hw = h * w;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int idx = (int)ray_idx[i];
    const int y = idx / w, x = idx - y * w;
    if (pixels) { pixels[2 * i] = scaled_px(x, w); pixels[2 * i + 1] = scaled_px(y, h); }
    if (rgb_gt && img) {
      rgb_gt[3 * i + 0] = img[idx];
      rgb_gt[3 * i + 1] = img[hw + idx];
      rgb_gt[3 * i + 2] = img[2 * hw + idx];
    }
    if (depth && depth_img) {
      float d = depth_img[idx];
      if (ss) d = __fadd_rn(__fmul_rn(d, sc), sh);  // two roundings, like the reference's mul then add
      depth[i] = d;
    }
  }
}

__global__ void gather_rays_bwd_kernel(const int64_t *__restrict__ ray_idx, int n, const float *__restrict__ depth_img,
                                       const float *__restrict__ d_depth, float *__restrict__ d_ss) {
  __shared__ float red[32];
  float a = 0.f, b = 0.f;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float g = d_depth[i];
    a += g * depth_img[ray_idx[i]];
    b += g;
  }
  a = block_sum(a, red);
  b = block_sum(b, red);
  if (threadIdx.x == 0) { atomicAdd(d_ss, a); atomicAdd(d_ss + 1, b); }
}

// ---------------------------------------------------------------------------------------------
// ray generation: model/common.py:112-237, model/rendering.py:53-80
// ---------------------------------------------------------------------------------------------
// A = S^-1 W^-1 K^-1 (left-to-right, as the reference's `scale_mat @ world_mat @ camera_mat`)
__global__ void raygen_mats_kernel(const float *__restrict__ K, const float *__restrict__ W, const float *__restrict__ S,
                                   float *__restrict__ A) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  double k[16], w[16], s[16], ki[16], wi[16], si[16], t[16], a[16];
  for (int i = 0; i < 16; ++i) { k[i] = K[i]; w[i] = W[i]; s[i] = S[i]; }
  m4_inv(k, ki); m4_inv(w, wi); m4_inv(s, si);
  m4_mul(si, wi, t);
  m4_mul(t, ki, a);
  for (int i = 0; i < 16; ++i) A[i] = (float)a[i];
}

__global__ void raygen_fwd_kernel(const float *__restrict__ pixels, const float *__restrict__ depth,
                                  const float *__restrict__ A, int n, int normalise, float *__restrict__ origin,
                                  float *__restrict__ dirs, float *__restrict__ dir_norm, float *__restrict__ d_gt,
                                  uint8_t *__restrict__ mask) {
  __shared__ float a[12];
  if (threadIdx.x < 12) a[threadIdx.x] = A[threadIdx.x];
  __syncthreads();
  if (blockIdx.x == 0 && threadIdx.x < 3 && origin) origin[threadIdx.x] = a[threadIdx.x * 4 + 3];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float px = pixels[2 * i], py = pixels[2 * i + 1];
    const float d = depth ? depth[i] : 1.f;
    float r[3], v[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      r[c] = a[c * 4 + 0] * px + a[c * 4 + 1] * py + a[c * 4 + 2];
      v[c] = a[c * 4 + 0] * (px * d) + a[c * 4 + 1] * (py * d) + a[c * 4 + 2] * d;
    }
    const float nrm = sqrtf(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
    float g = sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    if (normalise) {
      const float inv = 1.0f / nrm;
      r[0] *= inv; r[1] *= inv; r[2] *= inv;
    } else {
      g = g / nrm;
    }
    dirs[3 * i] = r[0]; dirs[3 * i + 1] = r[1]; dirs[3 * i + 2] = r[2];
    if (dir_norm) dir_norm[i] = nrm;
    if (d_gt) d_gt[i] = g;
    if (mask) mask[i] = (isfinite(g) && g != 0.f) ? 1 : 0;
  }
}

// scratch[0..8] += sum_rays dA3; one thread per ray
__global__ void raygen_bwd_kernel(const float *__restrict__ pixels, const float *__restrict__ depth,
                                  const float *__restrict__ A, int n, int normalise, const float *__restrict__ d_dirs,
                                  const float *__restrict__ d_dgt, float *__restrict__ d_depth, float *__restrict__ scratch) {
  __shared__ float a[12];
  __shared__ float red[32];
  if (threadIdx.x < 12) a[threadIdx.x] = A[threadIdx.x];
  __syncthreads();
  float acc[9];
#pragma unroll
  for (int k = 0; k < 9; ++k) acc[k] = 0.f;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float px = pixels[2 * i], py = pixels[2 * i + 1];
    const float d = depth ? depth[i] : 1.f;
    const float q[3] = {px, py, 1.f};
    float r[3], v[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      r[c] = a[c * 4 + 0] * px + a[c * 4 + 1] * py + a[c * 4 + 2];
      v[c] = r[c] * d;
    }
    const float nrm = sqrtf(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
    const float g = sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    const float inv = 1.0f / nrm;
    float dr[3] = {0.f, 0.f, 0.f}, dv[3] = {0.f, 0.f, 0.f};
    if (d_dirs) {
      const float gd[3] = {d_dirs[3 * i], d_dirs[3 * i + 1], d_dirs[3 * i + 2]};
      if (normalise) {
        const float u[3] = {r[0] * inv, r[1] * inv, r[2] * inv};
        const float dot = u[0] * gd[0] + u[1] * gd[1] + u[2] * gd[2];
#pragma unroll
        for (int c = 0; c < 3; ++c) dr[c] += (gd[c] - u[c] * dot) * inv;
      } else {
#pragma unroll
        for (int c = 0; c < 3; ++c) dr[c] += gd[c];
      }
    }
    if (d_dgt) {
      const float gg = d_dgt[i];
      float dg = gg;
      if (!normalise) {
        dg = gg * inv;
        const float dn = -gg * g * inv * inv;
#pragma unroll
        for (int c = 0; c < 3; ++c) dr[c] += dn * r[c] * inv;
      }
      if (g > 0.f) {
        const float ig = dg / g;
#pragma unroll
        for (int c = 0; c < 3; ++c) dv[c] = ig * v[c];
      }
    }
    if (d_depth) d_depth[i] = depth ? (dv[0] * r[0] + dv[1] * r[1] + dv[2] * r[2]) : 0.f;
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
      for (int k = 0; k < 3; ++k) acc[c * 3 + k] += dr[c] * q[k] + dv[c] * q[k] * d;
  }
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    const float s = block_sum(acc[k], red);
    if (threadIdx.x == 0) atomicAdd(scratch + k, s);
  }
}

// chain dA -> dK, dW, dS with A = Si Wi Ki and dX = -Xi^T dXi Xi^T
__global__ void raygen_mats_bwd_kernel(const float *__restrict__ K, const float *__restrict__ W, const float *__restrict__ S,
                                       const float *__restrict__ scratch, const float *__restrict__ d_origin,
                                       float *__restrict__ dK, float *__restrict__ dW, float *__restrict__ dS) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  double k[16], w[16], s[16], ki[16], wi[16], si[16], dA[16];
  for (int i = 0; i < 16; ++i) { k[i] = K[i]; w[i] = W[i]; s[i] = S[i]; dA[i] = 0.0; }
  for (int c = 0; c < 3; ++c) {
    for (int j = 0; j < 3; ++j) dA[c * 4 + j] = scratch[c * 3 + j];
    dA[c * 4 + 3] = d_origin ? (double)d_origin[c] : 0.0;
  }
  m4_inv(k, ki); m4_inv(w, wi); m4_inv(s, si);
  double sw[16], wk[16], t1[16], t2[16], dKi[16], dWi[16], dSi[16], out[16];
  m4_mul(si, wi, sw);
  m4_mul(wi, ki, wk);
  // dKi = (Si Wi)^T dA
  m4_transpose(sw, t1); m4_mul(t1, dA, dKi);
  // dWi = Si^T dA Ki^T
  m4_transpose(si, t1); m4_mul(t1, dA, t2); m4_transpose(ki, t1); m4_mul(t2, t1, dWi);
  // dSi = dA (Wi Ki)^T
  m4_transpose(wk, t1); m4_mul(dA, t1, dSi);
  if (dK) { m4_inv_bwd(ki, dKi, out); for (int i = 0; i < 16; ++i) dK[i] = (float)out[i]; }
  if (dW) { m4_inv_bwd(wi, dWi, out); for (int i = 0; i < 16; ++i) dW[i] = (float)out[i]; }
  if (dS) { m4_inv_bwd(si, dSi, out); for (int i = 0; i < 16; ++i) dS[i] = (float)out[i]; }
}

// ---------------------------------------------------------------------------------------------
// point-cloud back-projection: model/training.py:354-362
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ int nearest_src(int dst, float scale, int in) {
  const int s = (int)floorf((float)dst * scale);
  return s < in - 1 ? s : in - 1;
}

__global__ void backproject_fwd_kernel(const float *__restrict__ depth_img, int h, int w, int h4, int w4,
                                       const float *__restrict__ ss, const float *__restrict__ K, float nl,
                                       float *__restrict__ pc, uint8_t *__restrict__ clamped) {
  __shared__ float ki[16];
  if (threadIdx.x == 0) {
    double k[16], o[16];
    for (int i = 0; i < 16; ++i) k[i] = K[i];
    m4_inv(k, o);
    for (int i = 0; i < 16; ++i) ki[i] = (float)o[i];
  }
  __syncthreads();
  const float sc = ss ? ss[0] : 1.f, sh = ss ? ss[1] : 0.f;
  const float sy = (float)h / (float)h4, sx = (float)w / (float)w4;
  const int n = h4 * w4;
  for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
    const int y4 = j / w4, x4 = j - y4 * w4;
    const int src = nearest_src(y4, sy, h) * w + nearest_src(x4, sx, w);
    float d = depth_img[src];
    if (ss) d = __fadd_rn(__fmul_rn(d, sc), sh);
    const bool cl = d < nl;
    if (cl) d = nl;
    const float px = scaled_px(x4, w4), py = scaled_px(y4, h4);
#pragma unroll
    for (int c = 0; c < 3; ++c) pc[3 * j + c] = ki[c * 4 + 0] * (px * d) + ki[c * 4 + 1] * (py * d) + ki[c * 4 + 2] * d + ki[c * 4 + 3];
    if (clamped) clamped[j] = cl ? 1 : 0;
  }
}

__global__ void backproject_bwd_kernel(const float *__restrict__ depth_img, int h, int w, int h4, int w4,
                                       const float *__restrict__ K, const uint8_t *__restrict__ clamped,
                                       const float *__restrict__ d_pc, float *__restrict__ d_ss) {
  __shared__ float ki[16];
  __shared__ float red[32];
  if (threadIdx.x == 0) {
    double k[16], o[16];
    for (int i = 0; i < 16; ++i) k[i] = K[i];
    m4_inv(k, o);
    for (int i = 0; i < 16; ++i) ki[i] = (float)o[i];
  }
  __syncthreads();
  const float sy = (float)h / (float)h4, sx = (float)w / (float)w4;
  const int n = h4 * w4;
  float a = 0.f, b = 0.f;
  for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
    if (clamped[j]) continue;
    const int y4 = j / w4, x4 = j - y4 * w4;
    const int src = nearest_src(y4, sy, h) * w + nearest_src(x4, sx, w);
    const float px = scaled_px(x4, w4), py = scaled_px(y4, h4);
    float dd = 0.f;
#pragma unroll
    for (int c = 0; c < 3; ++c) dd += d_pc[3 * j + c] * (ki[c * 4 + 0] * px + ki[c * 4 + 1] * py + ki[c * 4 + 2]);
    a += dd * depth_img[src];
    b += dd;
  }
  a = block_sum(a, red);
  b = block_sum(b, red);
  if (threadIdx.x == 0) { atomicAdd(d_ss, a); atomicAdd(d_ss + 1, b); }
}

// ---------------------------------------------------------------------------------------------------------------------
// NDC ray warp (model/common.py:632-675), one thread per ray; op order of the reference, each op rounded.
//   t = -(near + oz) / dz;  p = o + t d;  A = px / pz;  B = py / pz;  gx = -1 / (1 / fx)
//   o' = (gx A, gy B, 1 + 2 near / pz);  d' = (gx (dx/dz - A), gy (dy/dz - B), 1 - o'z)
// ---------------------------------------------------------------------------------------------------------------------
__global__ void ndc_rays_fwd_kernel(const float *__restrict__ K, float near_, const float *__restrict__ origin,
                                    const float *__restrict__ dirs, int n, float *__restrict__ no, float *__restrict__ nd) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float gx = __fdiv_rn(-1.f, __fdiv_rn(1.f, K[0])), gy = __fdiv_rn(-1.f, __fdiv_rn(1.f, K[5]));
  const float ox = origin[0], oy = origin[1], oz = origin[2];
  const float dx = dirs[3 * i], dy = dirs[3 * i + 1], dz = dirs[3 * i + 2];
  const float t = __fdiv_rn(-__fadd_rn(near_, oz), dz);
  const float px = __fadd_rn(ox, __fmul_rn(t, dx)), py = __fadd_rn(oy, __fmul_rn(t, dy)), pz = __fadd_rn(oz, __fmul_rn(t, dz));
  const float A = __fdiv_rn(px, pz), B = __fdiv_rn(py, pz);
  const float o2 = __fadd_rn(1.f, __fdiv_rn(__fmul_rn(2.f, near_), pz));
  no[3 * i] = __fmul_rn(gx, A);
  no[3 * i + 1] = __fmul_rn(gy, B);
  no[3 * i + 2] = o2;
  nd[3 * i] = __fmul_rn(gx, __fsub_rn(__fdiv_rn(dx, dz), A));
  nd[3 * i + 1] = __fmul_rn(gy, __fsub_rn(__fdiv_rn(dy, dz), B));
  nd[3 * i + 2] = __fsub_rn(1.f, o2);
}

__global__ void ndc_rays_bwd_kernel(const float *__restrict__ K, float near_, const float *__restrict__ origin,
                                    const float *__restrict__ dirs, int n, const float *__restrict__ g_no,
                                    const float *__restrict__ g_nd, float *__restrict__ d_o_rows, float *__restrict__ d_dirs,
                                    float *__restrict__ d_fxfy) {
  __shared__ float red[32];
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  float gfx = 0.f, gfy = 0.f;
  if (i < n) {
    const float gx = __fdiv_rn(-1.f, __fdiv_rn(1.f, K[0])), gy = __fdiv_rn(-1.f, __fdiv_rn(1.f, K[5]));
    const float ox = origin[0], oy = origin[1], oz = origin[2];
    const float dx = dirs[3 * i], dy = dirs[3 * i + 1], dz = dirs[3 * i + 2];
    const float t = -(near_ + oz) / dz;
    const float px = ox + t * dx, py = oy + t * dy, pz = oz + t * dz;
    const float A = px / pz, B = py / pz, rx = dx / dz, ry = dy / dz;
    const float go0 = g_no ? g_no[3 * i] : 0.f, go1 = g_no ? g_no[3 * i + 1] : 0.f, go2 = g_no ? g_no[3 * i + 2] : 0.f;
    const float gd0 = g_nd ? g_nd[3 * i] : 0.f, gd1 = g_nd ? g_nd[3 * i + 1] : 0.f, gd2 = g_nd ? g_nd[3 * i + 2] : 0.f;
    const float gA = gx * (go0 - gd0), gB = gy * (go1 - gd1), grx = gx * gd0, gry = gy * gd1;
    // d gx / d fx = -1 (gx = -1/(1/fx))
    gfx = -(go0 * A + gd0 * (rx - A));
    gfy = -(go1 * B + gd1 * (ry - B));
    const float go2t = go2 - gd2;
    const float ipz = 1.f / pz;
    const float gpx = gA * ipz, gpy = gB * ipz;
    const float gpz = -(go2t * 2.f * near_ + gA * px + gB * py) * ipz * ipz;
    const float gt = gpx * dx + gpy * dy + gpz * dz;
    const float idz = 1.f / dz;
    d_o_rows[3 * i] = gpx;
    d_o_rows[3 * i + 1] = gpy;
    d_o_rows[3 * i + 2] = gpz - gt * idz;
    d_dirs[3 * i] = t * gpx + grx * idz;
    d_dirs[3 * i + 1] = t * gpy + gry * idz;
    d_dirs[3 * i + 2] = t * gpz - gt * t * idz - (grx * dx + gry * dy) * idz * idz;
  }
  gfx = block_sum(gfx, red);
  gfy = block_sum(gfy, red);
  if (threadIdx.x == 0) { atomicAdd(d_fxfy, gfx); atomicAdd(d_fxfy + 1, gfy); }
}

static inline int grid_for(int n, int threads) {
  int g = nnb_cdiv(n, threads);
  const int cap = nnb_sm_count() * 8;
  return g < 1 ? 1 : (g > cap ? cap : g);
}

// ---- ray choice: the first n_pick entries of a keyed pseudo-random permutation of [0, n_total) -------------------------------------
// Replaces torch.randperm(h*w)[:n_points] (model/training.py:277), which sorts all h*w keys (five radix passes, ~240 us on the
// critical path before the render) to keep 4,096 of them.  A balanced Feistel network over 2^(2 half) >= n_total values is a
// bijection for any round function; values that land outside [0, n_total) walk the cycle until they return inside, which keeps
// the map a bijection of [0, n_total).  Six rounds keyed by `keys` (drawn by the caller from its generator).
__device__ __forceinline__ uint32_t feistel_round(uint32_t x, uint32_t k) {
  x ^= k;
  x *= 0x9E3779B1u; x ^= x >> 15;
  x *= 0x85EBCA77u; x ^= x >> 13;
  x *= 0xC2B2AE3Du; x ^= x >> 16;
  return x;
}

__global__ void ray_choice_kernel(const int64_t *__restrict__ keys, long long n_total, int n_pick, int half, int64_t *__restrict__ out) {
  const uint32_t mask = (1u << half) - 1u;
  uint32_t k[6];
#pragma unroll
  for (int r = 0; r < 6; ++r) k[r] = (uint32_t)keys[r];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_pick; i += gridDim.x * blockDim.x) {
    unsigned long long x = (unsigned long long)i;
    do {
      uint32_t l = (uint32_t)(x >> half) & mask, r = (uint32_t)x & mask;
#pragma unroll
      for (int q = 0; q < 6; ++q) {
        const uint32_t t = l ^ (feistel_round(r, k[q]) & mask);
        l = r; r = t;
      }
      x = ((unsigned long long)l << half) | r;
    } while (x >= (unsigned long long)n_total);
    out[i] = (int64_t)x;
  }
}

// ---- batch staging: up to NNB_COPY_MAX device-to-device copies and two index scalars in one launch ---------------------------------
struct CopyJobs {
  const void *src[NNB_COPY_MAX];
  void *dst[NNB_COPY_MAX];
  unsigned long long bytes[NNB_COPY_MAX];
  int n;
  int64_t *scalar_dst[2];
  int64_t scalar[2];
};

__global__ void __launch_bounds__(256) copy_many_kernel(const CopyJobs J) {
  if (blockIdx.x == 0 && threadIdx.x < 2 && J.scalar_dst[threadIdx.x]) *J.scalar_dst[threadIdx.x] = J.scalar[threadIdx.x];
  const unsigned long long t0 = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x, nt = (unsigned long long)gridDim.x * blockDim.x;
  for (int j = 0; j < J.n; ++j) {
    const unsigned long long nb = J.bytes[j];
    const bool vec = (((uintptr_t)J.src[j] | (uintptr_t)J.dst[j]) & 15) == 0;
    const unsigned long long nv = vec ? nb / 16 : 0;
    const uint4 *s4 = static_cast<const uint4 *>(J.src[j]);
    uint4 *d4 = static_cast<uint4 *>(J.dst[j]);
    for (unsigned long long e = t0; e < nv; e += nt) d4[e] = __ldcs(s4 + e);
    const unsigned char *s1 = static_cast<const unsigned char *>(J.src[j]);
    unsigned char *d1 = static_cast<unsigned char *>(J.dst[j]);
    for (unsigned long long e = nv * 16 + t0; e < nb; e += nt) d1[e] = s1[e];
  }
}

extern "C" {

int nnb_pixel_grid(int h, int w, int64_t *loc, float *scaled, void *stream) {
  if (h < 2 || w < 2) return NNB_ERR_BAD_ARG;
  pixel_grid_kernel<<<grid_for(h * w, 256), 256, 0, nnb_stream(stream)>>>(h, w, loc, scaled);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_gather_rays_fwd(const int64_t *ray_idx, int n_rays, int h, int w, const float *img, const float *depth_img,
                        const float *scale_shift, float *pixels, float *rgb_gt, float *depth, void *stream) {
  if (!ray_idx || n_rays < 0 || h < 2 || w < 2) return NNB_ERR_BAD_ARG;
  if (n_rays == 0) return 0;
  gather_rays_fwd_kernel<<<grid_for(n_rays, 256), 256, 0, nnb_stream(stream)>>>(ray_idx, n_rays, h, w, img, depth_img,
                                                                               scale_shift, pixels, rgb_gt, depth);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_gather_rays_bwd(const int64_t *ray_idx, int n_rays, const float *depth_img, const float *d_depth,
                        float *d_scale_shift, void *stream) {
  if (!ray_idx || !depth_img || !d_depth || !d_scale_shift) return NNB_ERR_BAD_ARG;
  NNB_CUDA(cudaMemsetAsync(d_scale_shift, 0, 2 * sizeof(float), nnb_stream(stream)));
  if (n_rays == 0) return 0;
  gather_rays_bwd_kernel<<<grid_for(n_rays, 256), 256, 0, nnb_stream(stream)>>>(ray_idx, n_rays, depth_img, d_depth, d_scale_shift);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_ray_choice(const int64_t *keys, long long n_total, int n_pick, int64_t *out, void *stream) {
  if (!keys || !out || n_total < 1 || n_pick < 0 || n_pick > n_total || n_total > (1ll << 40)) return NNB_ERR_BAD_ARG;
  if (n_pick == 0) return 0;
  int half = 1;
  while ((1ull << (2 * half)) < (unsigned long long)n_total) ++half;
  ray_choice_kernel<<<grid_for(n_pick, 128), 128, 0, nnb_stream(stream)>>>(keys, n_total, n_pick, half, out);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_copy_many(const void *const *src, void *const *dst, const unsigned long long *bytes, int n, int64_t *const *scalar_dst,
                  const int64_t *scalar, void *stream) {
  if (n < 0 || n > NNB_COPY_MAX || (n > 0 && (!src || !dst || !bytes))) return NNB_ERR_BAD_ARG;
  CopyJobs J;
  memset(&J, 0, sizeof(J));
  unsigned long long total = 0;
  for (int j = 0; j < n; ++j) {
    if (bytes[j] && (!src[j] || !dst[j])) return NNB_ERR_BAD_ARG;
    J.src[j] = src[j]; J.dst[j] = dst[j]; J.bytes[j] = bytes[j];
    total += bytes[j];
  }
  J.n = n;
  for (int q = 0; q < 2; ++q) {
    J.scalar_dst[q] = scalar_dst ? scalar_dst[q] : nullptr;
    J.scalar[q] = scalar ? scalar[q] : 0;
  }
  if (total == 0 && !J.scalar_dst[0] && !J.scalar_dst[1]) return 0;
  const unsigned long long want = (total / 16 + 255) / 256;
  const int sms = nnb_sm_count();
  const int grid = (int)(want < 1 ? 1 : (want > (unsigned long long)(8 * sms) ? (unsigned long long)(8 * sms) : want));
  copy_many_kernel<<<grid, 256, 0, nnb_stream(stream)>>>(J);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_raygen_fwd(const float *pixels, const float *depth, const float *camera_mat, const float *world_mat,
                   const float *scale_mat, int n_rays, int normalise_ray, float *A, float *origin, float *dirs,
                   float *dir_norm, float *d_gt, uint8_t *mask, void *stream) {
  if (!pixels || !camera_mat || !world_mat || !scale_mat || !A || !dirs || n_rays < 0) return NNB_ERR_BAD_ARG;
  cudaStream_t s = nnb_stream(stream);
  raygen_mats_kernel<<<1, 32, 0, s>>>(camera_mat, world_mat, scale_mat, A);
  NNB_CHECK_LAUNCH();
  raygen_fwd_kernel<<<grid_for(n_rays > 0 ? n_rays : 1, 256), 256, 0, s>>>(pixels, depth, A, n_rays, normalise_ray, origin, dirs,
                                                                           dir_norm, d_gt, mask);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_raygen_bwd(const float *pixels, const float *depth, const float *camera_mat, const float *world_mat,
                   const float *scale_mat, const float *A, int n_rays, int normalise_ray, const float *d_dirs,
                   const float *d_origin, const float *d_d_gt, float *d_depth, float *d_camera_mat,
                   float *d_world_mat, float *d_scale_mat, float *scratch16, void *stream) {
  if (!pixels || !camera_mat || !world_mat || !scale_mat || !A || !scratch16 || n_rays < 0) return NNB_ERR_BAD_ARG;
  cudaStream_t s = nnb_stream(stream);
  NNB_CUDA(cudaMemsetAsync(scratch16, 0, 16 * sizeof(float), s));
  if (n_rays > 0) {
    raygen_bwd_kernel<<<grid_for(n_rays, 256), 256, 0, s>>>(pixels, depth, A, n_rays, normalise_ray, d_dirs, d_d_gt, d_depth, scratch16);
    NNB_CHECK_LAUNCH();
  }
  if (d_camera_mat || d_world_mat || d_scale_mat) {
    raygen_mats_bwd_kernel<<<1, 32, 0, s>>>(camera_mat, world_mat, scale_mat, scratch16, d_origin, d_camera_mat, d_world_mat, d_scale_mat);
    NNB_CHECK_LAUNCH();
  }
  return 0;
}

int nnb_backproject_fwd(const float *depth_img, int h, int w, int h4, int w4, const float *scale_shift,
                        const float *camera_mat, float nearest_limit, float *pc, uint8_t *clamped, void *stream) {
  if (!depth_img || !camera_mat || !pc || h4 < 2 || w4 < 2) return NNB_ERR_BAD_ARG;
  backproject_fwd_kernel<<<grid_for(h4 * w4, 128), 128, 0, nnb_stream(stream)>>>(depth_img, h, w, h4, w4, scale_shift, camera_mat,
                                                                                nearest_limit, pc, clamped);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_backproject_bwd(const float *depth_img, int h, int w, int h4, int w4, const float *camera_mat,
                        const uint8_t *clamped, const float *d_pc, float *d_scale_shift, void *stream) {
  if (!depth_img || !camera_mat || !clamped || !d_pc || !d_scale_shift) return NNB_ERR_BAD_ARG;
  NNB_CUDA(cudaMemsetAsync(d_scale_shift, 0, 2 * sizeof(float), nnb_stream(stream)));
  backproject_bwd_kernel<<<grid_for(h4 * w4, 256), 256, 0, nnb_stream(stream)>>>(depth_img, h, w, h4, w4, camera_mat, clamped, d_pc, d_scale_shift);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_ndc_rays_fwd(const float *camera_mat, float near_, const float *origin, const float *dirs, int n_rays, float *ndc_o,
                     float *ndc_d, void *stream) {
  if (!camera_mat || !origin || !dirs || !ndc_o || !ndc_d || n_rays < 0) return NNB_ERR_BAD_ARG;
  if (n_rays == 0) return 0;
  ndc_rays_fwd_kernel<<<nnb_cdiv(n_rays, 256), 256, 0, nnb_stream(stream)>>>(camera_mat, near_, origin, dirs, n_rays, ndc_o, ndc_d);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_ndc_rays_bwd(const float *camera_mat, float near_, const float *origin, const float *dirs, int n_rays, const float *d_ndc_o,
                     const float *d_ndc_d, float *d_origin_rows, float *d_dirs, float *d_fxfy, void *stream) {
  if (!camera_mat || !origin || !dirs || !d_origin_rows || !d_dirs || !d_fxfy || n_rays < 0) return NNB_ERR_BAD_ARG;
  NNB_CUDA(cudaMemsetAsync(d_fxfy, 0, 2 * sizeof(float), nnb_stream(stream)));
  if (n_rays == 0) return 0;
  ndc_rays_bwd_kernel<<<nnb_cdiv(n_rays, 256), 256, 0, nnb_stream(stream)>>>(camera_mat, near_, origin, dirs, n_rays, d_ndc_o, d_ndc_d,
                                                                            d_origin_rows, d_dirs, d_fxfy);
  NNB_CHECK_LAUNCH();
  return 0;
}

}  // extern "C"
