// SE(3) exp-map pose math, single-thread, double precision internally.
// Restates model/common.py:277-330 (vec2skew, Exp, make_c2w) and model/poses.py:23-31.
// __host__ __device__ so that tests/native/ can exercise it on the CPU.
#pragma once
#include "nnb_common.cuh"

// R = I + a K + b K^2, a = sin(th)/th, b = (1-cos th)/th^2, th = |r| + 1e-15 (the reference's guard;
// no Taylor branch: at r = 0 this gives a = 1, b = 0 exactly, as the reference does in fp32).
struct PoseCoef {
  double th, a, b, da, db;  // da = da/dth, db = db/dth
};

__host__ __device__ inline PoseCoef pose_coef(const double r[3]) {
  PoseCoef c;
  double n = sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
  if (n == 0.0) {
    // fp32 semantics of the reference at r == 0: sin(1e-15)/1e-15 = 1, (1-cos(1e-15))/1e-30 = 0;
    // d|r|/dr = 0 there (torch masks the 0/0), so da, db never contribute.
    c.th = 1e-15; c.a = 1.0; c.b = 0.0; c.da = 0.0; c.db = 0.0;
    return c;
  }
  double th = n + 1e-15;
  double s = sin(th), co = cos(th);
  c.th = th;
  c.a = s / th;
  // 1 - cos th = 2 sin^2(th/2): avoids the cancellation the fp32 reference suffers
  double sh = sin(0.5 * th);
  double omc = 2.0 * sh * sh;
  c.b = omc / (th * th);
  c.da = (th * co - s) / (th * th);
  c.db = (th * s - 2.0 * omc) / (th * th * th);
  return c;
}

__host__ __device__ inline void skew(const double r[3], double K[9]) {
  K[0] = 0; K[1] = -r[2]; K[2] = r[1];
  K[3] = r[2]; K[4] = 0; K[5] = -r[0];
  K[6] = -r[1]; K[7] = r[0]; K[8] = 0;
}

__host__ __device__ inline void m3_mul(const double *a, const double *b, double *c) {
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) c[i * 3 + j] = a[i * 3 + 0] * b[0 * 3 + j] + a[i * 3 + 1] * b[1 * 3 + j] + a[i * 3 + 2] * b[2 * 3 + j];
}

// c2w (row-major 4x4) = [Exp(r) | t; 0 0 0 1] @ init (init may be nullptr)
__host__ __device__ inline void pose_forward(const double r[3], const double t[3], const double *init, double out[16]) {
  PoseCoef c = pose_coef(r);
  double K[9], K2[9], M[16];
  skew(r, K);
  m3_mul(K, K, K2);
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j) M[i * 4 + j] = (i == j ? 1.0 : 0.0) + c.a * K[i * 3 + j] + c.b * K2[i * 3 + j];
    M[i * 4 + 3] = t[i];
  }
  M[12] = 0; M[13] = 0; M[14] = 0; M[15] = 1;
  if (init) m4_mul(M, init, out);
  else for (int i = 0; i < 16; ++i) out[i] = M[i];
}

// Given dL/d(c2w) (row-major 4x4), gradients for r and t.
__host__ __device__ inline void pose_backward(const double r[3], const double * /*t*/, const double *init, const double dout[16],
                                              double dr[3], double dt[3]) {
  // out = M @ init  =>  dM = dout @ init^T
  double dM[16];
  if (init) {
    double it[16];
    m4_transpose(init, it);
    m4_mul(dout, it, dM);
  } else {
    for (int i = 0; i < 16; ++i) dM[i] = dout[i];
  }
  for (int i = 0; i < 3; ++i) dt[i] = dM[i * 4 + 3];
  double dR[9];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) dR[i * 3 + j] = dM[i * 4 + j];
  PoseCoef c = pose_coef(r);
  double K[9], K2[9];
  skew(r, K);
  m3_mul(K, K, K2);
  // scalar paths: dL/da = <dR, K>, dL/db = <dR, K2>
  double dLa = 0, dLb = 0;
  for (int i = 0; i < 9; ++i) { dLa += dR[i] * K[i]; dLb += dR[i] * K2[i]; }
  // matrix path: dK = a dR + b (dR K^T + K^T dR)
  double Kt[9], t1[9], t2[9], dK[9];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) Kt[i * 3 + j] = K[j * 3 + i];
  m3_mul(dR, Kt, t1);
  m3_mul(Kt, dR, t2);
  for (int i = 0; i < 9; ++i) dK[i] = c.a * dR[i] + c.b * (t1[i] + t2[i]);
  // K entries -> r
  dr[0] = dK[7] - dK[5];
  dr[1] = dK[2] - dK[6];
  dr[2] = dK[3] - dK[1];
  double n = sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
  if (n > 0.0) {
    double dth = dLa * c.da + dLb * c.db;
    for (int i = 0; i < 3; ++i) dr[i] += dth * r[i] / n;
  }
}
