// fp32 CUDA-core building blocks of the NeRF MLP (the tight-tolerance parity path, NNB_PREC_FP32).
// Layer-by-layer with activations in HBM, like the reference, but with the concatenations removed
// (skip / view-direction inputs are written in place into wider row buffers) and bias/ReLU/mask fused
// into the GEMM epilogue.  The tensor-core product path lives in mlp_tc.cu.
#pragma once
#include "nnb_common.cuh"

#define SG_BM 128
#define SG_BN 128
#define SG_BK 16
#define SG_PAD 4

struct SgemmEpilogue {
  const float *bias;   // [N] or null
  int relu;            // max(0, .)
  const float *mask;   // zero the output where mask[m*ldmask+n] <= 0 (ReLU backward), or null
  int ldmask;
  int accumulate;      // C += val (non-atomic)
  int atomic;          // atomicAdd into C (split-K)
};

// C[M,N] (+)= op(A)[M,K] * op(B)[K,N]
//   A_KC: A(m,k) = A[m*lda + k]  else A(m,k) = A[k*lda + m]
//   B_KC: B(k,n) = B[n*ldb + k]  else B(k,n) = B[k*ldb + n]
template <bool A_KC, bool B_KC>
__global__ void __launch_bounds__(256) sgemm_kernel(int M, int N, int K, const float *__restrict__ A, int lda,
                                                    const float *__restrict__ B, int ldb, float *__restrict__ C, int ldc,
                                                    SgemmEpilogue ep, int kchunk) {
  __shared__ float As[SG_BK][SG_BM + SG_PAD];
  __shared__ float Bs[SG_BK][SG_BN + SG_PAD];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const long long m0 = (long long)blockIdx.x * SG_BM;
  const int n0 = blockIdx.y * SG_BN;
  const int kbeg = blockIdx.z * kchunk;
  const int kend = min(K, kbeg + kchunk);
  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

  for (int k0 = kbeg; k0 < kend; k0 += SG_BK) {
#pragma unroll
    for (int i = 0; i < (SG_BM * SG_BK) / 256; ++i) {
      const int e = tid + i * 256;
      int m, k;
      if (A_KC) { k = e % SG_BK; m = e / SG_BK; } else { m = e % SG_BM; k = e / SG_BM; }
      const long long gm = m0 + m;
      const int gk = k0 + k;
      float v = 0.f;
      if (gm < M && gk < kend) v = A_KC ? A[gm * lda + gk] : A[(long long)gk * lda + gm];
      As[k][m] = v;
    }
#pragma unroll
    for (int i = 0; i < (SG_BN * SG_BK) / 256; ++i) {
      const int e = tid + i * 256;
      int n, k;
      if (B_KC) { k = e % SG_BK; n = e / SG_BK; } else { n = e % SG_BN; k = e / SG_BN; }
      const int gn = n0 + n, gk = k0 + k;
      float v = 0.f;
      if (gn < N && gk < kend) v = B_KC ? B[(long long)gn * ldb + gk] : B[(long long)gk * ldb + gn];
      Bs[k][n] = v;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < SG_BK; ++k) {
      const float4 a0 = *reinterpret_cast<const float4 *>(&As[k][ty * 4]);
      const float4 a1 = *reinterpret_cast<const float4 *>(&As[k][64 + ty * 4]);
      const float4 b0 = *reinterpret_cast<const float4 *>(&Bs[k][tx * 4]);
      const float4 b1 = *reinterpret_cast<const float4 *>(&Bs[k][64 + tx * 4]);
      const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const long long gm = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
    if (gm >= M) continue;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int gn = n0 + (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
      if (gn >= N) continue;
      float v = acc[i][j];
      if (ep.bias) v += ep.bias[gn];
      if (ep.relu) v = fmaxf(v, 0.f);
      float *c = C + gm * ldc + gn;
      if (ep.accumulate && !ep.atomic) v += *c;  // the mask below applies to the accumulated value
      if (ep.mask && !(ep.mask[gm * ep.ldmask + gn] > 0.f)) v = 0.f;
      if (ep.atomic) atomicAdd(c, v);
      else *c = v;
    }
  }
}

template <bool A_KC, bool B_KC>
static inline int sgemm_launch(long long M, int N, int K, const float *A, int lda, const float *B, int ldb, float *C, int ldc,
                               SgemmEpilogue ep, int splits, cudaStream_t s) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  int kchunk = K;
  if (splits > 1) {
    kchunk = nnb_cdiv(nnb_cdiv(K, splits), SG_BK) * SG_BK;
    splits = nnb_cdiv(K, kchunk);
    ep.atomic = 1;
  } else {
    splits = 1;
  }
  dim3 grid(nnb_cdiv(M, SG_BM), nnb_cdiv(N, SG_BN), splits);
  sgemm_kernel<A_KC, B_KC><<<grid, 256, 0, s>>>((int)M, N, K, A, lda, B, ldb, C, ldc, ep, kchunk);
  NNB_CHECK_LAUNCH();
  return 0;
}
