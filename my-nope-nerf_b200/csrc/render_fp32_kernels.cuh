// Elementwise / small kernels around the fp32 MLP: stratified sampling + positional encoding, heads,
// bias-gradient column sums, encoding backward.  model/rendering.py:89-90,183-198 and
// model/official_nerf.py:60-119.
#pragma once
#include "nnb_common.cuh"
#include "sampling.cuh"

#define ENC_POS 63   // (2*10+1)*3
#define ENC_DIR 27   // (2*4+1)*3
#define LD_X0 64
#define LD_HS 320
#define LD_FC 288

// One block = 64 samples.  Phase 1 computes the 63-wide encoding into shared memory, phase 2 writes rows
// coalesced to X0 (ld 64) and to the skip columns of Hs (ld 320); also z and the per-sample copy of the
// view-direction encoding into Fc[:, 256:288].
__global__ void __launch_bounds__(256) encode_fwd_kernel(const float *__restrict__ origin, const float *__restrict__ dirs,
                                                         const float *__restrict__ ray_o, const float *__restrict__ view,
                                                         const float *__restrict__ noise, long long M, int S, float near_,
                                                         float far_, int use_dir, float *__restrict__ X0,
                                                         float *__restrict__ Hs, float *__restrict__ Fc,
                                                         float *__restrict__ zout) {
  __shared__ float enc[64][65];
  __shared__ float denc[64][33];
  const long long base = (long long)blockIdx.x * 64;
  const int tid = threadIdx.x;
  if (tid < 192) {
    const int ls = tid / 3, c = tid % 3;
    const long long m = base + ls;
    if (m < M) {
      const long long ray = m / S;
      const int s = (int)(m - ray * S);
      const float z = z_sample(s, S, near_, far_, noise ? noise + ray * S : nullptr);
      if (c == 0 && zout) zout[m] = z;
      const float d = dirs[3 * ray + c];
      const float x = __fadd_rn(ray_o ? ray_o[3 * ray + c] : origin[c], __fmul_rn(d, z));
      enc[ls][c] = x;
      float f = 1.0f;
#pragma unroll
      for (int l = 0; l < 10; ++l) {
        float sn, cs;
        sincosf(f * x, &sn, &cs);
        enc[ls][3 + 6 * l + c] = sn;
        enc[ls][3 + 6 * l + 3 + c] = cs;
        f *= 2.0f;
      }
      if (c == 0) enc[ls][63] = 0.f;
      const float v = use_dir ? -(view ? view[3 * ray + c] : d) : 1.0f;  // view direction = -ray (model/rendering.py:195-198)
      denc[ls][c] = v;
      f = 1.0f;
#pragma unroll
      for (int l = 0; l < 4; ++l) {
        float sn, cs;
        sincosf(f * v, &sn, &cs);
        denc[ls][3 + 6 * l + c] = sn;
        denc[ls][3 + 6 * l + 3 + c] = cs;
        f *= 2.0f;
      }
      if (c == 0) { for (int k = ENC_DIR; k < 32; ++k) denc[ls][k] = 0.f; }
    }
  }
  __syncthreads();
  for (int e = tid; e < 64 * 64; e += 256) {
    const int ls = e >> 6, k = e & 63;
    const long long m = base + ls;
    if (m < M) {
      const float v = enc[ls][k];
      X0[m * LD_X0 + k] = v;
      Hs[m * LD_HS + 256 + k] = v;
    }
  }
  for (int e = tid; e < 64 * 32; e += 256) {
    const int ls = e >> 5, k = e & 31;
    const long long m = base + ls;
    if (m < M) Fc[m * LD_FC + 256 + k] = denc[ls][k];
  }
}

// density head: raw = x . w + b ; alpha = 1 - exp(-softplus(raw))  (or relu; or sigma itself if dist_alpha)
// warp per sample.  model/official_nerf.py:66,77-83
__global__ void head_density_fwd_kernel(const float *__restrict__ X, long long M, const float *__restrict__ w,
                                        const float *__restrict__ b, int softplus, int dist_alpha, float *__restrict__ raw,
                                        float *__restrict__ alpha) {
  const int lane = threadIdx.x & 31;
  const long long wid = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long nw = ((long long)gridDim.x * blockDim.x) >> 5;
  float wr[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) wr[k] = w[lane + 32 * k];
  for (long long m = wid; m < M; m += nw) {
    const float *x = X + m * 256;
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) acc = fmaf(x[lane + 32 * k], wr[k], acc);
    acc = warp_sum(acc);
    if (lane == 0) {
      const float r = acc + b[0];
      float d = softplus ? (r > 20.f ? r : log1pf(expf(r))) : fmaxf(r, 0.f);
      if (!dist_alpha) d = 1.f - expf(-d);
      raw[m] = r;
      alpha[m] = d;
    }
  }
}

// d_raw (in place over d_alpha) = d_alpha * d(alpha)/d(raw)
__global__ void head_density_bwd_kernel(const float *__restrict__ raw, long long M, int softplus, int dist_alpha,
                                        float *__restrict__ d_alpha_inout) {
  for (long long m = (long long)blockIdx.x * blockDim.x + threadIdx.x; m < M; m += (long long)gridDim.x * blockDim.x) {
    const float r = raw[m];
    float g = d_alpha_inout[m];
    float dact;  // d(activation)/d(raw)
    float act;
    if (softplus) {
      act = r > 20.f ? r : log1pf(expf(r));
      dact = r > 20.f ? 1.f : 1.f / (1.f + expf(-r));
    } else {
      act = fmaxf(r, 0.f);
      dact = r > 0.f ? 1.f : 0.f;
    }
    if (!dist_alpha) g *= expf(-act);
    d_alpha_inout[m] = g * dact;
  }
}

// rgb head: sigmoid(h . W^T + b), W (3,128).  warp per sample.  model/official_nerf.py:91-92
__global__ void head_rgb_fwd_kernel(const float *__restrict__ H, long long M, const float *__restrict__ W,
                                    const float *__restrict__ b, float *__restrict__ rgb) {
  const int lane = threadIdx.x & 31;
  const long long wid = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long nw = ((long long)gridDim.x * blockDim.x) >> 5;
  float wr[3][4];
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int k = 0; k < 4; ++k) wr[c][k] = W[c * 128 + lane + 32 * k];
  for (long long m = wid; m < M; m += nw) {
    const float *h = H + m * 128;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float v = h[lane + 32 * k];
      a0 = fmaf(v, wr[0][k], a0); a1 = fmaf(v, wr[1][k], a1); a2 = fmaf(v, wr[2][k], a2);
    }
    a0 = warp_sum(a0); a1 = warp_sum(a1); a2 = warp_sum(a2);
    if (lane == 0) {
      rgb[3 * m] = 1.f / (1.f + expf(-(a0 + b[0])));
      rgb[3 * m + 1] = 1.f / (1.f + expf(-(a1 + b[1])));
      rgb[3 * m + 2] = 1.f / (1.f + expf(-(a2 + b[2])));
    }
  }
}

// in place: g *= s (1 - s)
__global__ void sigmoid_bwd_kernel(const float *__restrict__ s, long long n, float *__restrict__ g) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float v = s[i];
    g[i] *= v * (1.f - v);
  }
}

// db[n] += sum_m dY[m*ld + n], n < N <= blockDim.x
__global__ void colsum_kernel(const float *__restrict__ dY, long long M, int N, int ld, float *__restrict__ db) {
  const int n = threadIdx.x;
  if (n >= N) return;
  const long long rows_per = (M + gridDim.x - 1) / gridDim.x;
  const long long r0 = (long long)blockIdx.x * rows_per;
  const long long r1 = r0 + rows_per < M ? r0 + rows_per : M;
  float acc = 0.f;
  for (long long m = r0; m < r1; ++m) acc += dY[m * ld + n];
  if (r1 > r0) atomicAdd(db + n, acc);
}

// Backward of the position encoding + sampling.  One block per ray (blockDim = S rounded up to a warp
// multiple).  d_enc row = Genc[m*ldg + 0..62]; enc row (saved) = X0[m*64 + ...].
//   d_x_c = g[c] + sum_l 2^l (cos_l g[3+6l+c] - sin_l g[3+6l+3+c]);  d_origin += d_x; d_dir[ray] += d_x * z
// Also the view-direction encoding: gdir = sum over the ray's samples of Gdir[m*ldgd + 0..26].
__global__ void encode_bwd_kernel(const float *__restrict__ X0, const float *__restrict__ Genc, int ldg,
                                  const float *__restrict__ Gdir, int ldgd, const float *__restrict__ dirs,
                                  const float *__restrict__ view, const float *__restrict__ z, int S, int use_dir,
                                  float *__restrict__ d_origin, float *__restrict__ d_dirs, float *__restrict__ d_ray_o,
                                  float *__restrict__ d_view) {
  __shared__ float red[32];
  __shared__ float gd[32];
  const long long ray = blockIdx.x;
  const int s = threadIdx.x;
  float dx[3] = {0.f, 0.f, 0.f};
  float zz = 0.f;
  if (s < S) {
    const long long m = ray * S + s;
    const float *e = X0 + m * LD_X0;
    const float *g = Genc + m * (long long)ldg;
    zz = z[m];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      float acc = g[c];
      float f = 1.0f;
#pragma unroll
      for (int l = 0; l < 10; ++l) {
        acc += f * (e[3 + 6 * l + 3 + c] * g[3 + 6 * l + c] - e[3 + 6 * l + c] * g[3 + 6 * l + 3 + c]);
        f *= 2.0f;
      }
      dx[c] = acc;
    }
  }
  // view-direction encoding gradient: column sums over the ray's samples
  if (threadIdx.x < 32) gd[threadIdx.x] = 0.f;
  __syncthreads();
  if (use_dir && Gdir) {
    for (int k = threadIdx.x; k < ENC_DIR * 4; k += blockDim.x) {
      // 4 partial sums per column to spread the serial loop
      const int col = k % ENC_DIR, part = k / ENC_DIR;
      float acc = 0.f;
      for (int ss = part; ss < S; ss += 4) acc += Gdir[(ray * S + ss) * (long long)ldgd + col];
      atomicAdd(&gd[col], acc);
    }
  }
  float o[3], dd[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    o[c] = block_sum(dx[c], red);
    dd[c] = block_sum(dx[c] * zz, red);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      float gdir = 0.f;
      if (use_dir && Gdir) {
        const float v = -(view ? view[3 * ray + c] : dirs[3 * ray + c]);
        float acc = gd[c];
        float f = 1.0f;
#pragma unroll
        for (int l = 0; l < 4; ++l) {
          float sn, cs;
          sincosf(f * v, &sn, &cs);
          acc += f * (cs * gd[3 + 6 * l + c] - sn * gd[3 + 6 * l + 3 + c]);
          f *= 2.0f;
        }
        gdir = -acc;  // v = -d
      }
      if (view) { d_dirs[3 * ray + c] = dd[c]; if (d_view) d_view[3 * ray + c] = gdir; }
      else d_dirs[3 * ray + c] = dd[c] + gdir;
      if (d_ray_o) d_ray_o[3 * ray + c] = o[c];
      else atomicAdd(d_origin + c, o[c]);
    }
  }
}
