// Stratified depth sampling shared by the fp32 and the tensor-core render paths (model/rendering.py:89-90,183-198).
#pragma once
#include "nnb_common.cuh"

// z of sample s along a ray, bit-compatible with the reference's fp32 op sequence:
// linspace(0,1,S) -> near*(1-u) + far*u -> optional stratified jitter (each op rounded separately).
__device__ __forceinline__ float linspace01(int i, int S) {
  // torch.linspace fp32: step = 1/(S-1); first half start + step*i, second half end - step*(S-1-i).
  // The second half is a fused multiply-add in torch's kernel (verified bit-for-bit against torch CPU).
  const float step = __fdiv_rn(1.0f, (float)(S - 1));
  return (i < S / 2) ? __fmul_rn(step, (float)i) : __fmaf_rn(-step, (float)(S - 1 - i), 1.0f);
}
__device__ __forceinline__ float z_plain(int i, int S, float near_, float far_) {
  const float u = linspace01(i, S);
  return __fadd_rn(__fmul_rn(near_, __fsub_rn(1.0f, u)), __fmul_rn(far_, u));
}
__device__ __forceinline__ float z_sample(int i, int S, float near_, float far_, const float *noise_row) {
  const float zi = z_plain(i, S, near_, far_);
  if (!noise_row) return zi;
  const float lo = (i == 0) ? zi : __fmul_rn(0.5f, __fadd_rn(zi, z_plain(i - 1, S, near_, far_)));
  const float hi = (i == S - 1) ? zi : __fmul_rn(0.5f, __fadd_rn(z_plain(i + 1, S, near_, far_), zi));
  return __fadd_rn(lo, __fmul_rn(__fsub_rn(hi, lo), noise_row[i]));
}

