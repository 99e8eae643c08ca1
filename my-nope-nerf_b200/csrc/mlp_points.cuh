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
