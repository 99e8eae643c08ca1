// Stage 3: alpha compositing -- transmittance scan, weights, rgb, expected distance; hand-written
// reverse-scan backward.  One warp per ray, each lane owns S/32 consecutive samples, warp-shuffle scans.
// HBM-bound: forward 3,088 B/ray, backward 4,624 B/ray at S = 128.
// Replaces model/rendering.py:113-126,139-141 (torch.cumprod + autograd's cumprod backward).
#include "nnb_common.cuh"

#define CP_MAX_PER_LANE 8  // S <= 256
#define CP_EPS 1e-6f       // model/rendering.py:9

__device__ __forceinline__ float warp_excl_prod(float v, int lane) {
  // exclusive prefix product across lanes
  float inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const float t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc *= t;
  }
  float ex = __shfl_up_sync(0xffffffffu, inc, 1);
  return lane == 0 ? 1.f : ex;
}
__device__ __forceinline__ float warp_excl_suffix_sum(float v, int lane) {
  float inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const float t = __shfl_down_sync(0xffffffffu, inc, o);
    if (lane + o < 32) inc += t;
  }
  float ex = __shfl_down_sync(0xffffffffu, inc, 1);
  return lane == 31 ? 0.f : ex;
}

__device__ __forceinline__ float eff_alpha(const float *__restrict__ a_in, const float *__restrict__ z, int s, int S, int dist_alpha) {
  float a = a_in[s];
  if (dist_alpha) {
    if (s == S - 1) return 1.f;
    a = 1.f - expf(-a * (z[s + 1] - z[s]));
  }
  return a;
}

__global__ void composite_fwd_kernel(const float *__restrict__ alpha_in, const float *__restrict__ rgb_s, const float *__restrict__ z,
                                     int R, int S, int dist_alpha, int white, float *__restrict__ rgb, float *__restrict__ dist,
                                     float *__restrict__ weights, float *__restrict__ alpha_out) {
  const int lane = threadIdx.x & 31;
  const int warps_per_block = blockDim.x >> 5;
  const int per = (S + 31) / 32;
  for (int ray = blockIdx.x * warps_per_block + (threadIdx.x >> 5); ray < R; ray += gridDim.x * warps_per_block) {
    const float *a_in = alpha_in + (size_t)ray * S;
    const float *zz = z + (size_t)ray * S;
    const float *cc = rgb_s + (size_t)ray * S * 3;
    float a[CP_MAX_PER_LANE], beta_prod = 1.f;
#pragma unroll
    for (int k = 0; k < CP_MAX_PER_LANE; ++k) {
      const int s = lane * per + k;
      a[k] = 0.f;
      if (k < per && s < S) {
        a[k] = eff_alpha(a_in, zz, s, S, dist_alpha);
        beta_prod *= (1.f - a[k] + CP_EPS);
      }
    }
    float T = warp_excl_prod(beta_prod, lane);
    float r0 = 0.f, r1 = 0.f, r2 = 0.f, dd = 0.f, ws = 0.f;
#pragma unroll
    for (int k = 0; k < CP_MAX_PER_LANE; ++k) {
      const int s = lane * per + k;
      if (k < per && s < S) {
        const float w = a[k] * T;
        T *= (1.f - a[k] + CP_EPS);
        r0 += w * cc[3 * s]; r1 += w * cc[3 * s + 1]; r2 += w * cc[3 * s + 2];
        dd += w * zz[s];
        ws += w;
        if (weights) weights[(size_t)ray * S + s] = w;
        if (alpha_out) alpha_out[(size_t)ray * S + s] = a[k];
      }
    }
    r0 = warp_sum(r0); r1 = warp_sum(r1); r2 = warp_sum(r2); dd = warp_sum(dd);
    if (white) { ws = warp_sum(ws); r0 += 1.f - ws; r1 += 1.f - ws; r2 += 1.f - ws; }
    if (lane == 0) {
      rgb[3 * ray] = r0; rgb[3 * ray + 1] = r1; rgb[3 * ray + 2] = r2;
      dist[ray] = dd;
    }
  }
}

// d_alpha_in[i] = dw_i T_i - (sum_{k>i} dw_k w_k) / beta_i, chained through 1-exp(-sigma*delta) when dist_alpha.
__global__ void composite_bwd_kernel(const float *__restrict__ alpha_in, const float *__restrict__ rgb_s, const float *__restrict__ z,
                                     int R, int S, int dist_alpha, int white, const float *__restrict__ d_rgb,
                                     const float *__restrict__ d_dist, const float *__restrict__ d_wout,
                                     const float *__restrict__ d_aout, float *__restrict__ d_alpha_in, float *__restrict__ d_rgb_s) {
  const int lane = threadIdx.x & 31;
  const int warps_per_block = blockDim.x >> 5;
  const int per = (S + 31) / 32;
  for (int ray = blockIdx.x * warps_per_block + (threadIdx.x >> 5); ray < R; ray += gridDim.x * warps_per_block) {
    const float *a_in = alpha_in + (size_t)ray * S;
    const float *zz = z + (size_t)ray * S;
    const float *cc = rgb_s + (size_t)ray * S * 3;
    const float g0 = d_rgb ? d_rgb[3 * ray] : 0.f, g1 = d_rgb ? d_rgb[3 * ray + 1] : 0.f, g2 = d_rgb ? d_rgb[3 * ray + 2] : 0.f;
    const float gd = d_dist ? d_dist[ray] : 0.f;
    const float gw_white = white ? -(g0 + g1 + g2) : 0.f;
    float a[CP_MAX_PER_LANE], beta_prod = 1.f;
#pragma unroll
    for (int k = 0; k < CP_MAX_PER_LANE; ++k) {
      const int s = lane * per + k;
      a[k] = 0.f;
      if (k < per && s < S) {
        a[k] = eff_alpha(a_in, zz, s, S, dist_alpha);
        beta_prod *= (1.f - a[k] + CP_EPS);
      }
    }
    const float T0 = warp_excl_prod(beta_prod, lane);
    float T = T0;
    float dw[CP_MAX_PER_LANE], Tk[CP_MAX_PER_LANE], local = 0.f;
#pragma unroll
    for (int k = 0; k < CP_MAX_PER_LANE; ++k) {
      const int s = lane * per + k;
      dw[k] = 0.f; Tk[k] = 0.f;
      if (k < per && s < S) {
        const float w = a[k] * T;
        Tk[k] = T;
        T *= (1.f - a[k] + CP_EPS);
        float g = g0 * cc[3 * s] + g1 * cc[3 * s + 1] + g2 * cc[3 * s + 2] + gd * zz[s] + gw_white;
        if (d_wout) g += d_wout[(size_t)ray * S + s];
        dw[k] = g;
        local += g * w;
        if (d_rgb_s) {
          float *o = d_rgb_s + ((size_t)ray * S + s) * 3;
          o[0] = w * g0; o[1] = w * g1; o[2] = w * g2;
        }
      }
    }
    // suffix: G_i = sum_{k>i} dw_k w_k
    float G = warp_excl_suffix_sum(local, lane);
#pragma unroll
    for (int k = CP_MAX_PER_LANE - 1; k >= 0; --k) {
      const int s = lane * per + k;
      if (k < per && s < S) {
        const float beta = 1.f - a[k] + CP_EPS;
        float da = dw[k] * Tk[k] - G / beta;
        G += dw[k] * a[k] * Tk[k];
        if (d_aout) da += d_aout[(size_t)ray * S + s];
        if (dist_alpha) {
          if (s == S - 1) da = 0.f;
          else {
            const float delta = zz[s + 1] - zz[s];
            da = da * delta * expf(-a_in[s] * delta);
          }
        }
        d_alpha_in[(size_t)ray * S + s] = da;
      }
    }
  }
}


// ---------------------------------------------------------------------------------------------------------------------
// S == 128 fast path: a lane's four consecutive samples are one 16-byte chunk of alpha / z / weights and three of rgb_s, so
// every global access is a float4 (one fully used 512-byte line per warp instruction instead of four quarter-used ones).
// The arithmetic and its order are those of the generic kernels above, so the results are bit-identical.
// ---------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float4 ldg4(const float *p) { return __ldg(reinterpret_cast<const float4 *>(p)); }
__device__ __forceinline__ void stcs4(float *p, float a, float b, float c, float d) { __stcs(reinterpret_cast<float4 *>(p), make_float4(a, b, c, d)); }

// effective alpha of the lane's four samples (model/rendering.py:116-122 when dist_alpha)
__device__ __forceinline__ void eff_alpha4(const float4 ai, const float4 zv, int lane, int dist_alpha, float (&a)[4]) {
  a[0] = ai.x; a[1] = ai.y; a[2] = ai.z; a[3] = ai.w;
  if (dist_alpha) {
    const float znext = __shfl_down_sync(0xffffffffu, zv.x, 1);   // z of the next lane's first sample
    a[0] = 1.f - expf(-ai.x * (zv.y - zv.x));
    a[1] = 1.f - expf(-ai.y * (zv.z - zv.y));
    a[2] = 1.f - expf(-ai.z * (zv.w - zv.z));
    a[3] = lane == 31 ? 1.f : 1.f - expf(-ai.w * (znext - zv.w));
  }
}

__global__ void __launch_bounds__(256) composite_fwd128_kernel(const float *__restrict__ alpha_in, const float *__restrict__ rgb_s,
                                                              const float *__restrict__ z, int R, int dist_alpha, int white,
                                                              float *__restrict__ rgb, float *__restrict__ dist, float *__restrict__ weights,
                                                              float *__restrict__ alpha_out) {
  const int lane = threadIdx.x & 31;
  const int warps_per_block = blockDim.x >> 5;
  for (int ray = blockIdx.x * warps_per_block + (threadIdx.x >> 5); ray < R; ray += gridDim.x * warps_per_block) {
    const size_t base = (size_t)ray * 128 + lane * 4;
    const float4 ai = ldg4(alpha_in + base), zv = ldg4(z + base);
    const float4 c0 = ldg4(rgb_s + 3 * base), c1 = ldg4(rgb_s + 3 * base + 4), c2 = ldg4(rgb_s + 3 * base + 8);
    const float cr[4] = {c0.x, c0.w, c1.z, c2.y}, cg[4] = {c0.y, c1.x, c1.w, c2.z}, cb[4] = {c0.z, c1.y, c2.x, c2.w};
    const float zz[4] = {zv.x, zv.y, zv.z, zv.w};
    float a[4], beta_prod = 1.f;
    eff_alpha4(ai, zv, lane, dist_alpha, a);
#pragma unroll
    for (int k = 0; k < 4; ++k) beta_prod *= (1.f - a[k] + CP_EPS);
    float T = warp_excl_prod(beta_prod, lane);
    float r0 = 0.f, r1 = 0.f, r2 = 0.f, dd = 0.f, ws = 0.f, w[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      w[k] = a[k] * T;
      T *= (1.f - a[k] + CP_EPS);
      r0 += w[k] * cr[k]; r1 += w[k] * cg[k]; r2 += w[k] * cb[k];
      dd += w[k] * zz[k];
      ws += w[k];
    }
    if (weights) stcs4(weights + base, w[0], w[1], w[2], w[3]);
    if (alpha_out) stcs4(alpha_out + base, a[0], a[1], a[2], a[3]);
    r0 = warp_sum(r0); r1 = warp_sum(r1); r2 = warp_sum(r2); dd = warp_sum(dd);
    if (white) { ws = warp_sum(ws); r0 += 1.f - ws; r1 += 1.f - ws; r2 += 1.f - ws; }
    if (lane == 0) {
      rgb[3 * ray] = r0; rgb[3 * ray + 1] = r1; rgb[3 * ray + 2] = r2;
      dist[ray] = dd;
    }
  }
}

__global__ void __launch_bounds__(256) composite_bwd128_kernel(const float *__restrict__ alpha_in, const float *__restrict__ rgb_s,
                                                              const float *__restrict__ z, int R, int dist_alpha, int white,
                                                              const float *__restrict__ d_rgb, const float *__restrict__ d_dist,
                                                              const float *__restrict__ d_wout, const float *__restrict__ d_aout,
                                                              float *__restrict__ d_alpha_in, float *__restrict__ d_rgb_s) {
  const int lane = threadIdx.x & 31;
  const int warps_per_block = blockDim.x >> 5;
  for (int ray = blockIdx.x * warps_per_block + (threadIdx.x >> 5); ray < R; ray += gridDim.x * warps_per_block) {
    const size_t base = (size_t)ray * 128 + lane * 4;
    const float4 ai = ldg4(alpha_in + base), zv = ldg4(z + base);
    const float4 c0 = ldg4(rgb_s + 3 * base), c1 = ldg4(rgb_s + 3 * base + 4), c2 = ldg4(rgb_s + 3 * base + 8);
    const float cr[4] = {c0.x, c0.w, c1.z, c2.y}, cg[4] = {c0.y, c1.x, c1.w, c2.z}, cb[4] = {c0.z, c1.y, c2.x, c2.w};
    const float zz[4] = {zv.x, zv.y, zv.z, zv.w}, ain[4] = {ai.x, ai.y, ai.z, ai.w};
    const float g0 = d_rgb ? d_rgb[3 * ray] : 0.f, g1 = d_rgb ? d_rgb[3 * ray + 1] : 0.f, g2 = d_rgb ? d_rgb[3 * ray + 2] : 0.f;
    const float gd = d_dist ? d_dist[ray] : 0.f;
    const float gw_white = white ? -(g0 + g1 + g2) : 0.f;
    float4 dwo = make_float4(0.f, 0.f, 0.f, 0.f), dao = dwo;
    if (d_wout) dwo = ldg4(d_wout + base);
    if (d_aout) dao = ldg4(d_aout + base);
    const float dwout[4] = {dwo.x, dwo.y, dwo.z, dwo.w}, daout[4] = {dao.x, dao.y, dao.z, dao.w};
    const float znext = __shfl_down_sync(0xffffffffu, zv.x, 1);
    float a[4], beta_prod = 1.f;
    eff_alpha4(ai, zv, lane, dist_alpha, a);
#pragma unroll
    for (int k = 0; k < 4; ++k) beta_prod *= (1.f - a[k] + CP_EPS);
    float T = warp_excl_prod(beta_prod, lane);
    float dw[4], Tk[4], w[4], local = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      w[k] = a[k] * T;
      Tk[k] = T;
      T *= (1.f - a[k] + CP_EPS);
      float g = g0 * cr[k] + g1 * cg[k] + g2 * cb[k] + gd * zz[k] + gw_white;
      if (d_wout) g += dwout[k];
      dw[k] = g;
      local += g * w[k];
    }
    if (d_rgb_s) {
      float *o = d_rgb_s + 3 * base;   // 12 consecutive floats: (w0 g0, w0 g1, w0 g2, w1 g0, ...)
      stcs4(o, w[0] * g0, w[0] * g1, w[0] * g2, w[1] * g0);
      stcs4(o + 4, w[1] * g1, w[1] * g2, w[2] * g0, w[2] * g1);
      stcs4(o + 8, w[2] * g2, w[3] * g0, w[3] * g1, w[3] * g2);
    }
    float G = warp_excl_suffix_sum(local, lane);
    float da[4];
#pragma unroll
    for (int k = 3; k >= 0; --k) {
      const float beta = 1.f - a[k] + CP_EPS;
      float v = dw[k] * Tk[k] - G / beta;
      G += dw[k] * a[k] * Tk[k];
      if (d_aout) v += daout[k];
      if (dist_alpha) {
        if (lane == 31 && k == 3) v = 0.f;
        else {
          const float delta = (k == 3 ? znext : zz[k + 1]) - zz[k];
          v = v * delta * expf(-ain[k] * delta);
        }
      }
      da[k] = v;
    }
    stcs4(d_alpha_in + base, da[0], da[1], da[2], da[3]);
  }
}

static inline bool cp_aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

static inline int cp_grid(int R, int threads) {
  int g = nnb_cdiv(R, threads / 32);
  const int cap = nnb_sm_count() * 8;
  return g < 1 ? 1 : (g > cap ? cap : g);
}

// internal (used by render.cu)
int nnb_composite_fwd_launch(const float *alpha_in, const float *rgb_s, const float *z, int R, int S, int dist_alpha, int white,
                             float *rgb, float *dist, float *weights, float *alpha_out, cudaStream_t s) {
  if (S == 128 && cp_aligned16(alpha_in) && cp_aligned16(rgb_s) && cp_aligned16(z) && cp_aligned16(weights) && cp_aligned16(alpha_out))
    composite_fwd128_kernel<<<cp_grid(R, 256), 256, 0, s>>>(alpha_in, rgb_s, z, R, dist_alpha, white, rgb, dist, weights, alpha_out);
  else
    composite_fwd_kernel<<<cp_grid(R, 128), 128, 0, s>>>(alpha_in, rgb_s, z, R, S, dist_alpha, white, rgb, dist, weights, alpha_out);
  NNB_CHECK_LAUNCH();
  return 0;
}
int nnb_composite_bwd_launch(const float *alpha_in, const float *rgb_s, const float *z, int R, int S, int dist_alpha, int white,
                             const float *d_rgb, const float *d_dist, const float *d_wout, const float *d_aout,
                             float *d_alpha_in, float *d_rgb_s, cudaStream_t s) {
  if (S == 128 && cp_aligned16(alpha_in) && cp_aligned16(rgb_s) && cp_aligned16(z) && cp_aligned16(d_wout) && cp_aligned16(d_aout) &&
      cp_aligned16(d_alpha_in) && cp_aligned16(d_rgb_s))
    composite_bwd128_kernel<<<cp_grid(R, 256), 256, 0, s>>>(alpha_in, rgb_s, z, R, dist_alpha, white, d_rgb, d_dist, d_wout, d_aout,
                                                            d_alpha_in, d_rgb_s);
  else
    composite_bwd_kernel<<<cp_grid(R, 128), 128, 0, s>>>(alpha_in, rgb_s, z, R, S, dist_alpha, white, d_rgb, d_dist, d_wout, d_aout,
                                                        d_alpha_in, d_rgb_s);
  NNB_CHECK_LAUNCH();
  return 0;
}

extern "C" {

int nnb_composite_fwd(const float *alpha_in, const float *rgb_s, const float *z, int n_rays, int n_samples, int dist_alpha,
                      int white_bkgd, float *rgb, float *dist, float *weights, float *alpha_out, void *stream) {
  if (!alpha_in || !rgb_s || !z || !rgb || !dist || n_rays <= 0 || n_samples <= 0 || n_samples > 32 * CP_MAX_PER_LANE)
    return NNB_ERR_BAD_ARG;
  return nnb_composite_fwd_launch(alpha_in, rgb_s, z, n_rays, n_samples, dist_alpha, white_bkgd, rgb, dist, weights, alpha_out,
                                  nnb_stream(stream));
}

int nnb_composite_bwd(const float *alpha_in, const float *rgb_s, const float *z, const float *weights, int n_rays, int n_samples,
                      int dist_alpha, int white_bkgd, const float *d_rgb, const float *d_dist, const float *d_weights_out,
                      float *d_alpha_in, float *d_rgb_s, void *stream) {
  (void)weights;  // recomputed in-kernel: cheaper than re-reading 512 B/ray
  if (!alpha_in || !rgb_s || !z || !d_alpha_in || n_rays <= 0 || n_samples <= 0 || n_samples > 32 * CP_MAX_PER_LANE)
    return NNB_ERR_BAD_ARG;
  return nnb_composite_bwd_launch(alpha_in, rgb_s, z, n_rays, n_samples, dist_alpha, white_bkgd, d_rgb, d_dist, d_weights_out,
                                  nullptr, d_alpha_in, d_rgb_s, nnb_stream(stream));
}

}  // extern "C"
