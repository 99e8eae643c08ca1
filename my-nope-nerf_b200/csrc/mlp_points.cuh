// Positional encoding of explicit query points (point-query entry nnb_mlp_points_fwd).
#pragma once
#include "render_fp32_kernels.cuh"

// thread per (point, channel): writes X0 (ld 64), Hs[:,256:320] (ld 320), Fc[:,256:288] (ld 288)
__global__ void encode_points_fwd_kernel(const float *__restrict__ pts, const float *__restrict__ viewdirs, long long M,
                                         float *__restrict__ X0, float *__restrict__ Hs, float *__restrict__ Fc) {
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < M * 3; e += (long long)gridDim.x * blockDim.x) {
    const long long m = e / 3;
    const int c = (int)(e - m * 3);
    const float x = pts[e];
    float *x0 = X0 + m * LD_X0, *hs = Hs + m * LD_HS + 256;
    x0[c] = x; hs[c] = x;
    float f = 1.0f;
    for (int l = 0; l < 10; ++l) {
      float sn, cs;
      sincosf(f * x, &sn, &cs);
      x0[3 + 6 * l + c] = sn; hs[3 + 6 * l + c] = sn;
      x0[3 + 6 * l + 3 + c] = cs; hs[3 + 6 * l + 3 + c] = cs;
      f *= 2.0f;
    }
    if (c == 0) { x0[63] = 0.f; hs[63] = 0.f; }
    if (viewdirs) {
      const float v = viewdirs[e];
      float *fc = Fc + m * LD_FC + 256;
      fc[c] = v;
      f = 1.0f;
      for (int l = 0; l < 4; ++l) {
        float sn, cs;
        sincosf(f * v, &sn, &cs);
        fc[3 + 6 * l + c] = sn;
        fc[3 + 6 * l + 3 + c] = cs;
        f *= 2.0f;
      }
      if (c == 0) for (int k = ENC_DIR; k < 32; ++k) fc[k] = 0.f;
    }
  }
}

// Input gradient of the position encoding at explicit points: d_x_c = g[c] + sum_l 2^l (cos_l g[3+6l+c] - sin_l g[3+6l+3+c]),
// enc row (saved by the forward) = X0[m*64 + ...], gradient row = Genc[m*64 + ...].  out (M,3) = sign * d_x.
__global__ void encode_points_bwd_kernel(const float *__restrict__ X0, const float *__restrict__ Genc, long long M, float sign,
                                         float *__restrict__ out) {
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < M * 3; e += (long long)gridDim.x * blockDim.x) {
    const long long m = e / 3;
    const int c = (int)(e - m * 3);
    const float *x0 = X0 + m * LD_X0, *g = Genc + m * LD_X0;
    float acc = g[c], f = 1.0f;
#pragma unroll
    for (int l = 0; l < 10; ++l) {
      acc += f * (x0[3 + 6 * l + 3 + c] * g[3 + 6 * l + c] - x0[3 + 6 * l + c] * g[3 + 6 * l + 3 + c]);
      f *= 2.0f;
    }
    out[e] = sign * acc;
  }
}

__global__ void fill_kernel(float *__restrict__ p, long long n, float v) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) p[i] = v;
}
