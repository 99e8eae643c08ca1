// Stage 4: rgb / depth losses as vectorised warp-shuffle reductions (HBM/latency-bound).
#include "nnb_common.cuh"

static inline int red_grid(int n, int threads) {
  int g = nnb_cdiv(n, threads * 4);
  const int cap = nnb_sm_count() * 8;
  return g < 1 ? 1 : (g > cap ? cap : g);
}

// ---------------------------------------------------------------------------------------------
// rgb: model/losses.py:28-33,200
// ---------------------------------------------------------------------------------------------
__global__ void rgb_loss_fwd_kernel(const float *__restrict__ pred, const float *__restrict__ gt, int n3, float inv_r,
                                    float *__restrict__ out3) {
  __shared__ float red[32];
  float s1 = 0.f, s2 = 0.f;
  const int tid = blockIdx.x * blockDim.x + threadIdx.x, nthr = gridDim.x * blockDim.x;
  // 16-byte loads where both tensors allow it (two independent loads in flight per iteration), scalar tail
  const bool vec = ((reinterpret_cast<uintptr_t>(pred) | reinterpret_cast<uintptr_t>(gt)) & 15) == 0;
  const int n4 = vec ? n3 >> 2 : 0;
  const float4 *p4 = reinterpret_cast<const float4 *>(pred), *g4 = reinterpret_cast<const float4 *>(gt);
  for (int i = tid; i < n4; i += nthr) {
    const float4 a = __ldg(p4 + i), b = __ldg(g4 + i);
    const float d0 = a.x - b.x, d1 = a.y - b.y, d2 = a.z - b.z, d3 = a.w - b.w;
    s1 += (fabsf(d0) + fabsf(d1)) + (fabsf(d2) + fabsf(d3));
    s2 += (d0 * d0 + d1 * d1) + (d2 * d2 + d3 * d3);
  }
  for (int i = 4 * n4 + tid; i < n3; i += nthr) {
    const float d = pred[i] - gt[i];
    s1 += fabsf(d);
    s2 += d * d;
  }
  s1 = block_sum(s1, red);
  s2 = block_sum(s2, red);
  if (threadIdx.x == 0) {
    atomicAdd(out3 + 0, s1 * inv_r);
    atomicAdd(out3 + 1, s2 * inv_r);
    atomicAdd(out3 + 2, s2 * inv_r * (1.0f / 3.0f));
  }
}

__global__ void rgb_loss_bwd_kernel(const float *__restrict__ pred, const float *__restrict__ gt, int n3, float inv_r,
                                    const float *__restrict__ g3, float *__restrict__ d_pred) {
  const float g1 = g3[0] * inv_r, g2 = 2.f * inv_r * (g3[1] + g3[2] * (1.0f / 3.0f));
  const int tid = blockIdx.x * blockDim.x + threadIdx.x, nthr = gridDim.x * blockDim.x;
  auto grad = [&](float d) { return g1 * (d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f)) + g2 * d; };
  const bool vec = ((reinterpret_cast<uintptr_t>(pred) | reinterpret_cast<uintptr_t>(gt) | reinterpret_cast<uintptr_t>(d_pred)) & 15) == 0;
  const int n4 = vec ? n3 >> 2 : 0;
  const float4 *p4 = reinterpret_cast<const float4 *>(pred), *g4 = reinterpret_cast<const float4 *>(gt);
  for (int i = tid; i < n4; i += nthr) {
    const float4 a = __ldg(p4 + i), b = __ldg(g4 + i);
    reinterpret_cast<float4 *>(d_pred)[i] = make_float4(grad(a.x - b.x), grad(a.y - b.y), grad(a.z - b.z), grad(a.w - b.w));
  }
  for (int i = 4 * n4 + tid; i < n3; i += nthr) d_pred[i] = grad(pred[i] - gt[i]);
}

// ---------------------------------------------------------------------------------------------
// depth l1 on dense tensors + mask: model/losses.py:60-63 (+ model/rendering.py:151-153 mask)
// ---------------------------------------------------------------------------------------------
__global__ void depth_l1_fwd_kernel(const float *__restrict__ p, const float *__restrict__ g, const uint8_t *__restrict__ m,
                                    int n, float *__restrict__ out2) {
  __shared__ float red[32];
  float s = 0.f, c = 0.f;
  const int tid = blockIdx.x * blockDim.x + threadIdx.x, nthr = gridDim.x * blockDim.x;
  const bool vec = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g)) & 15) == 0 && (!m || (reinterpret_cast<uintptr_t>(m) & 3) == 0);
  const int n4 = vec ? n >> 2 : 0;
  for (int i = tid; i < n4; i += nthr) {
    const float4 a = __ldg(reinterpret_cast<const float4 *>(p) + i), b = __ldg(reinterpret_cast<const float4 *>(g) + i);
    const uint32_t mm = m ? __ldg(reinterpret_cast<const uint32_t *>(m) + i) : 0x01010101u;   // four bool bytes
    if (mm & 0x000000ffu) { s += fabsf(a.x - b.x); c += 1.f; }
    if (mm & 0x0000ff00u) { s += fabsf(a.y - b.y); c += 1.f; }
    if (mm & 0x00ff0000u) { s += fabsf(a.z - b.z); c += 1.f; }
    if (mm & 0xff000000u) { s += fabsf(a.w - b.w); c += 1.f; }
  }
  for (int i = 4 * n4 + tid; i < n; i += nthr) {
    if (!m || m[i]) { s += fabsf(p[i] - g[i]); c += 1.f; }
  }
  s = block_sum(s, red);
  c = block_sum(c, red);
  if (threadIdx.x == 0) { atomicAdd(out2, s); atomicAdd(out2 + 1, c); }
}
__global__ void depth_l1_finish_kernel(float *out2) {
  if (threadIdx.x == 0 && blockIdx.x == 0) out2[0] = out2[1] > 0.f ? out2[0] / out2[1] : 0.f;
}
__global__ void depth_l1_bwd_kernel(const float *__restrict__ p, const float *__restrict__ g, const uint8_t *__restrict__ m,
                                    int n, const float *__restrict__ out2, const float *__restrict__ gl,
                                    float *__restrict__ dp, float *__restrict__ dg) {
  const float cnt = out2[1];
  const float k = cnt > 0.f ? gl[0] / cnt : 0.f;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    float v = 0.f;
    if (!m || m[i]) {
      const float d = p[i] - g[i];
      v = d > 0.f ? k : (d < 0.f ? -k : 0.f);
    }
    if (dp) dp[i] = v;
    if (dg) dg[i] = -v;
  }
}

// ---------------------------------------------------------------------------------------------
// exact lower median by MSB-first radix select (single block): model/losses.py:43,46
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t f2key(float f) {
  uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key2f(uint32_t k) {
  uint32_t u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
  return __uint_as_float(u);
}

// Returns (value, first index) of the k-th smallest masked element, k = (count-1)/2.  All threads get the result.
// smem: hist[256] + misc[4]
__device__ void block_median(const float *__restrict__ x, const uint8_t *__restrict__ m, int n, uint32_t *hist,
                             uint32_t *misc, float *val_out, int *idx_out, int *count_out) {
  // count
  if (threadIdx.x == 0) misc[0] = 0;
  __syncthreads();
  uint32_t c = 0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) c += (!m || m[i]) ? 1u : 0u;
  atomicAdd(&misc[0], c);
  __syncthreads();
  const int count = (int)misc[0];
  __syncthreads();
  if (count == 0) { *val_out = 0.f; *idx_out = -1; *count_out = 0; return; }
  uint32_t k = (uint32_t)((count - 1) / 2);
  uint32_t prefix = 0, pmask = 0;
  for (int shift = 24; shift >= 0; shift -= 8) {
    for (int b = threadIdx.x; b < 256; b += blockDim.x) hist[b] = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      if (m && !m[i]) continue;
      const uint32_t key = f2key(x[i]);
      if ((key & pmask) == prefix) atomicAdd(&hist[(key >> shift) & 255u], 1u);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t acc = 0;
      int b = 0;
      for (; b < 256; ++b) {
        if (acc + hist[b] > k) break;
        acc += hist[b];
      }
      misc[1] = (uint32_t)b;
      misc[2] = k - acc;
    }
    __syncthreads();
    prefix |= misc[1] << shift;
    pmask |= 255u << shift;
    k = misc[2];
    __syncthreads();
  }
  const float v = key2f(prefix);
  if (threadIdx.x == 0) misc[3] = 0x7fffffffu;
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    if (m && !m[i]) continue;
    if (f2key(x[i]) == prefix) atomicMin(&misc[3], (uint32_t)i);
  }
  __syncthreads();
  *val_out = v;
  *idx_out = (int)misc[3];
  *count_out = count;
  __syncthreads();
}

__global__ void median_select_kernel(const float *__restrict__ x, const uint8_t *__restrict__ m, int n,
                                     float *__restrict__ out_val, int64_t *__restrict__ out_idx) {
  __shared__ uint32_t hist[256];
  __shared__ uint32_t misc[4];
  float v; int idx, cnt;
  block_median(x, m, n, hist, misc, &v, &idx, &cnt);
  if (threadIdx.x == 0) { out_val[0] = v; out_idx[0] = idx; }
}

// ---------------------------------------------------------------------------------------------
// scale/shift-invariant depth loss: model/losses.py:35-58 (single block: n <= a few 100k)
// stats8 = { t_pred, s_pred, t_gt, s_gt, count, loss, 0, 0 }
// ---------------------------------------------------------------------------------------------
__global__ void depth_invariant_fwd_kernel(const float *__restrict__ p, const float *__restrict__ g,
                                           const uint8_t *__restrict__ m, int n, float *__restrict__ loss,
                                           float *__restrict__ stats, int64_t *__restrict__ med_idx) {
  __shared__ uint32_t hist[256];
  __shared__ uint32_t misc[4];
  __shared__ float red[32];
  __shared__ float bc[2];
  float tp, tg; int ip, ig, cnt;
  block_median(p, m, n, hist, misc, &tp, &ip, &cnt);
  block_median(g, m, n, hist, misc, &tg, &ig, &cnt);
  float ap = 0.f, ag = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    if (m && !m[i]) continue;
    ap += fabsf(p[i] - tp);
    ag += fabsf(g[i] - tg);
  }
  ap = block_sum(ap, red);
  ag = block_sum(ag, red);
  if (threadIdx.x == 0) { bc[0] = ap / (float)cnt; bc[1] = ag / (float)cnt; }
  __syncthreads();
  const float sp = bc[0], sg = bc[1];
  float acc = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    if (m && !m[i]) continue;
    const float d = (p[i] - tp) / sp - (g[i] - tg) / sg;
    acc += d * d;
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) {
    const float l = cnt > 0 ? acc / (float)cnt : 0.f;
    loss[0] = l;
    stats[0] = tp; stats[1] = sp; stats[2] = tg; stats[3] = sg; stats[4] = (float)cnt; stats[5] = l; stats[6] = 0.f; stats[7] = 0.f;
    med_idx[0] = ip; med_idx[1] = ig;
  }
}

// With u = x - t, s = mean|u|, xn = u/s, L = mean (pn - gn)^2, q_i = 2 (pn_i - gn_i)/n * gL:
//   ds_p = -sum q_i up_i / sp^2,  dup_i = q_i/sp + ds_p sign(up_i)/n,  dt_p = -sum dup_i,  dp_i = dup_i + [i==med_p] dt_p
//   (same with q -> -q for gt)
__global__ void depth_invariant_bwd_kernel(const float *__restrict__ p, const float *__restrict__ g,
                                           const uint8_t *__restrict__ m, int n, const float *__restrict__ stats,
                                           const int64_t *__restrict__ med_idx, const float *__restrict__ gl,
                                           float *__restrict__ dp, float *__restrict__ dg) {
  __shared__ float red[32];
  __shared__ float bc[5];
  const float tp = stats[0], sp = stats[1], tg = stats[2], sg = stats[3], cnt = stats[4];
  const float gL = gl[0];
  const float k = cnt > 0.f ? 2.f * gL / cnt : 0.f;
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f, a4 = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    if (m && !m[i]) continue;
    const float up = p[i] - tp, ug = g[i] - tg;
    const float q = k * (up / sp - ug / sg);
    a0 += q * up; a1 += q * ug; a2 += q;
    a3 += up > 0.f ? 1.f : (up < 0.f ? -1.f : 0.f);
    a4 += ug > 0.f ? 1.f : (ug < 0.f ? -1.f : 0.f);
  }
  a0 = block_sum(a0, red); if (threadIdx.x == 0) bc[0] = a0;
  a1 = block_sum(a1, red); if (threadIdx.x == 0) bc[1] = a1;
  a2 = block_sum(a2, red); if (threadIdx.x == 0) bc[2] = a2;
  a3 = block_sum(a3, red); if (threadIdx.x == 0) bc[3] = a3;
  a4 = block_sum(a4, red); if (threadIdx.x == 0) bc[4] = a4;
  __syncthreads();
  const float dsp = -bc[0] / (sp * sp), dsg = bc[1] / (sg * sg);
  const float dtp = -(bc[2] / sp + dsp * bc[3] / cnt);
  const float dtg = -(-bc[2] / sg + dsg * bc[4] / cnt);
  const int ip = (int)med_idx[0], ig = (int)med_idx[1];
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    float vp = 0.f, vg = 0.f;
    if (!m || m[i]) {
      const float up = p[i] - tp, ug = g[i] - tg;
      const float q = k * (up / sp - ug / sg);
      vp = q / sp + dsp * (up > 0.f ? 1.f : (up < 0.f ? -1.f : 0.f)) / cnt;
      vg = -q / sg + dsg * (ug > 0.f ? 1.f : (ug < 0.f ? -1.f : 0.f)) / cnt;
      if (i == ip) vp += dtp;
      if (i == ig) vg += dtg;
    }
    if (dp) dp[i] = vp;
    if (dg) dg[i] = vg;
  }
}

// ---------------------------------------------------------------------------------------------
// photometric (reprojection) loss: model/losses.py:79-87,152-159
// ---------------------------------------------------------------------------------------------
__global__ void photometric_fwd_kernel(const float *__restrict__ a, const float *__restrict__ b, const uint8_t *__restrict__ v,
                                       int n, float *__restrict__ out2) {
  __shared__ float red[32];
  float s = 0.f, c = 0.f;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    if (!v[i]) continue;
    c += 1.f;
#pragma unroll
    for (int k = 0; k < 3; ++k) s += fminf(fabsf(a[3 * i + k] - b[3 * i + k]), 1.f);
  }
  s = block_sum(s, red);
  c = block_sum(c, red);
  if (threadIdx.x == 0) { atomicAdd(out2, s); atomicAdd(out2 + 1, c); }
}
__global__ void photometric_finish_kernel(float *out2) {
  if (threadIdx.x == 0 && blockIdx.x == 0) out2[0] = out2[1] > 0.f ? out2[0] / (3.f * out2[1]) : 0.f;
}
__global__ void photometric_bwd_kernel(const float *__restrict__ a, const float *__restrict__ b, const uint8_t *__restrict__ v,
                                       int n, const float *__restrict__ out2, const float *__restrict__ gl,
                                       float *__restrict__ da, float *__restrict__ db) {
  const float cnt = out2[1];
  const float k = cnt > 0.f ? gl[0] / (3.f * cnt) : 0.f;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      float g = 0.f;
      if (v[i]) {
        const float d = a[3 * i + c] - b[3 * i + c];
        // clamp(|d|,0,1): gradient passes where 0 <= |d| <= 1 (torch clamp passes at the boundary)
        if (fabsf(d) <= 1.f) g = d > 0.f ? k : (d < 0.f ? -k : 0.f);
      }
      if (da) da[3 * i + c] = g;
      if (db) db[3 * i + c] = -g;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// fused Adam (torch.optim.Adam, amsgrad False, weight_decay 0): model/training.py:93-99
// ---------------------------------------------------------------------------------------------
__global__ void adam_kernel(float *__restrict__ p, const float *__restrict__ g, float *__restrict__ m, float *__restrict__ v,
                            size_t n, float lr, float b1, float b2, float eps, float bc1, float bc2_sqrt) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const float gi = g[i];
    const float mi = m[i] + (gi - m[i]) * (1.f - b1);  // lerp form used by torch
    const float vi = v[i] * b2 + (1.f - b2) * gi * gi;
    m[i] = mi; v[i] = vi;
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    p[i] -= (lr / bc1) * (mi / denom);
  }
}

extern "C" {

int nnb_rgb_loss_fwd(const float *pred, const float *gt, int n_rays, float *out3, void *stream) {
  if (!pred || !gt || !out3 || n_rays <= 0) return NNB_ERR_BAD_ARG;
  cudaStream_t s = nnb_stream(stream);
  NNB_CUDA(cudaMemsetAsync(out3, 0, 3 * sizeof(float), s));
  rgb_loss_fwd_kernel<<<red_grid(n_rays * 3, 256), 256, 0, s>>>(pred, gt, n_rays * 3, 1.0f / (float)n_rays, out3);
  NNB_CHECK_LAUNCH();
  return 0;
}
int nnb_rgb_loss_bwd(const float *pred, const float *gt, int n_rays, const float *g_out3, float *d_pred, void *stream) {
  if (!pred || !gt || !g_out3 || !d_pred || n_rays <= 0) return NNB_ERR_BAD_ARG;
  rgb_loss_bwd_kernel<<<red_grid(n_rays * 3, 256), 256, 0, nnb_stream(stream)>>>(pred, gt, n_rays * 3, 1.0f / (float)n_rays, g_out3, d_pred);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_depth_l1_fwd(const float *pred, const float *gt, const uint8_t *mask, int n, float *out2, void *stream) {
  if (!pred || !gt || !out2 || n < 0) return NNB_ERR_BAD_ARG;
  cudaStream_t s = nnb_stream(stream);
  NNB_CUDA(cudaMemsetAsync(out2, 0, 2 * sizeof(float), s));
  if (n > 0) { depth_l1_fwd_kernel<<<red_grid(n, 256), 256, 0, s>>>(pred, gt, mask, n, out2); NNB_CHECK_LAUNCH(); }
  depth_l1_finish_kernel<<<1, 32, 0, s>>>(out2);
  NNB_CHECK_LAUNCH();
  return 0;
}
int nnb_depth_l1_bwd(const float *pred, const float *gt, const uint8_t *mask, int n, const float *out2, const float *g_loss,
                     float *d_pred, float *d_gt, void *stream) {
  if (!pred || !gt || !out2 || !g_loss || n < 0) return NNB_ERR_BAD_ARG;
  if (n == 0) return 0;
  depth_l1_bwd_kernel<<<red_grid(n, 256), 256, 0, nnb_stream(stream)>>>(pred, gt, mask, n, out2, g_loss, d_pred, d_gt);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_median_select(const float *x, const uint8_t *mask, int n, float *out_val, int64_t *out_idx, void *stream) {
  if (!x || !out_val || !out_idx || n <= 0) return NNB_ERR_BAD_ARG;
  median_select_kernel<<<1, 1024, 0, nnb_stream(stream)>>>(x, mask, n, out_val, out_idx);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_depth_invariant_fwd(const float *pred, const float *gt, const uint8_t *mask, int n, float *loss, float *stats8,
                            int64_t *med_idx2, void *stream) {
  if (!pred || !gt || !loss || !stats8 || !med_idx2 || n <= 0) return NNB_ERR_BAD_ARG;
  depth_invariant_fwd_kernel<<<1, 1024, 0, nnb_stream(stream)>>>(pred, gt, mask, n, loss, stats8, med_idx2);
  NNB_CHECK_LAUNCH();
  return 0;
}
int nnb_depth_invariant_bwd(const float *pred, const float *gt, const uint8_t *mask, int n, const float *stats8,
                            const int64_t *med_idx2, const float *g_loss, float *d_pred, float *d_gt, void *stream) {
  if (!pred || !gt || !stats8 || !med_idx2 || !g_loss || n <= 0) return NNB_ERR_BAD_ARG;
  depth_invariant_bwd_kernel<<<1, 1024, 0, nnb_stream(stream)>>>(pred, gt, mask, n, stats8, med_idx2, g_loss, d_pred, d_gt);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_photometric_fwd(const float *rgb1, const float *rgb2, const uint8_t *valid, int n_points, float *out2, void *stream) {
  if (!rgb1 || !rgb2 || !valid || !out2 || n_points <= 0) return NNB_ERR_BAD_ARG;
  cudaStream_t s = nnb_stream(stream);
  NNB_CUDA(cudaMemsetAsync(out2, 0, 2 * sizeof(float), s));
  photometric_fwd_kernel<<<red_grid(n_points, 256), 256, 0, s>>>(rgb1, rgb2, valid, n_points, out2);
  NNB_CHECK_LAUNCH();
  photometric_finish_kernel<<<1, 32, 0, s>>>(out2);
  NNB_CHECK_LAUNCH();
  return 0;
}
int nnb_photometric_bwd(const float *rgb1, const float *rgb2, const uint8_t *valid, int n_points, const float *out2,
                        const float *g_loss, float *d_rgb1, float *d_rgb2, void *stream) {
  if (!rgb1 || !rgb2 || !valid || !out2 || !g_loss || n_points <= 0) return NNB_ERR_BAD_ARG;
  photometric_bwd_kernel<<<red_grid(n_points, 256), 256, 0, nnb_stream(stream)>>>(rgb1, rgb2, valid, n_points, out2, g_loss, d_rgb1, d_rgb2);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_adam_step(float *param, const float *grad, float *exp_avg, float *exp_avg_sq, size_t n, float lr, float beta1,
                  float beta2, float eps, int step, void *stream) {
  if (!param || !grad || !exp_avg || !exp_avg_sq || step < 1) return NNB_ERR_BAD_ARG;
  if (n == 0) return 0;
  const float bc1 = 1.f - powf(beta1, (float)step);
  const float bc2 = sqrtf(1.f - powf(beta2, (float)step));
  int g = nnb_cdiv((long long)n, 256 * 4);
  const int cap = nnb_sm_count() * 8;
  if (g > cap) g = cap;
  adam_kernel<<<g, 256, 0, nnb_stream(stream)>>>(param, grad, exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps, bc1, bc2);
  NNB_CHECK_LAUNCH();
  return 0;
}

}  // extern "C"
