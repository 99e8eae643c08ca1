// Stage 5: Chamfer nearest-neighbour search and point-to-point error.
//
// Replaces Loss.comp_closest_pts_idx_with_split (model/losses.py:129-144), which materialises a
// (3,S,D) difference tensor (10 GB at P = 28,830) and argmins over sqrt(sum sq).  Here every source
// point streams all destination points through shared memory; nothing is materialised.
//
// Index parity: the reference takes the FIRST index of the minimum of the fp32 value
//     s_j = sqrt_rn( (dx*dx + dy*dy) + dz*dz )          (each op rounded; verified against torch CPU)
// sqrt_rn maps up to a few adjacent d2 values to the same s, so comparing d2 alone would break ties the
// reference does not break: a candidate replaces the incumbent only if its ROUNDED distance s is strictly
// smaller.  Candidates are visited in ascending j, and partial results from different destination ranges are
// merged with an atomicMin on (bits(s) << 32 | j), so the first index wins exactly as in torch.argmin.
#include "nnb_common.cuh"

// chamfer_tc.cu
size_t nnb_chamfer_tc_workspace_bytes(int n_src, int n_dst);
size_t nnb_chamfer_tc_norm2_offset(int n_src, int n_dst);
int nnb_chamfer_tc_bound(const float *src, int n_src, const float *dst, int n_dst, unsigned int *bound, unsigned int *stats, void *workspace,
                         unsigned long long *best_init, unsigned long long *best, int use_f16, cudaStream_t s);

#define CH_THREADS 128
#define CH_SPT 4      // sources per thread
#define CH_TILE 1024  // destination points per shared-memory tile

// Two passes over the (source, destination) pairs, both with the expanded form  a = |q|^2 - 2 s.q :
//   pass 1: bound[s] ~ min_j d2(s, j).  Production: tcgen05 GEMM with a min epilogue (chamfer_tc.cu, error
//           <= 8e-6 (|s-c|^2 + 2 max|q-c|^2)).  EXACT = false below is the CUDA-core version of the same pass, kept as the
//           probe the tensor-core pass is tested against (tests/test_gpu_losses.py).
//   pass 2 (EXACT = true): three FMAs per pair (|q|^2 rides in tile[].w), branch-free; pairs with
//           a < (bound - |s|^2) + margin are re-evaluated exactly (direct differences in the reference's op order).
//           margin = 1e-6 (|s|^2 + 2 max|q|^2 + bound) [this pass's own screen: <= 4 eps (|s|^2 + 2|q|^2), eps = 6e-8, plus
//           sqrt ties] + the pass-1 error term, so no pair the reference could select is screened out and indices stay
//           bit-exact.  Typically a handful of pairs per source survive, so the exact path is cold.
template <bool EXACT>
__global__ void __launch_bounds__(CH_THREADS) chamfer_nn_kernel(const float *__restrict__ src, int n_src,
                                                                const float *__restrict__ dst, int n_dst, int dst_per_block,
                                                                unsigned int *__restrict__ bound, unsigned int *__restrict__ qqmax_g,
                                                                unsigned long long *__restrict__ best,
                                                                const float *__restrict__ tc_norm2, const unsigned int *__restrict__ tc_stats) {
  __shared__ float4 tile[CH_TILE];
  __shared__ float qqmax_s[CH_THREADS / 32];
  const int s0 = (blockIdx.x * CH_THREADS + threadIdx.x) * CH_SPT;
  float sx[CH_SPT], sy[CH_SPT], sz[CH_SPT], ss[CH_SPT], m2x[CH_SPT], m2y[CH_SPT], m2z[CH_SPT], thr[CH_SPT], smin[CH_SPT], amin[CH_SPT];
  int jbest[CH_SPT];
#pragma unroll
  for (int k = 0; k < CH_SPT; ++k) {
    const int s = min(s0 + k, n_src - 1);
    sx[k] = src[3 * s]; sy[k] = src[3 * s + 1]; sz[k] = src[3 * s + 2];
    ss[k] = sx[k] * sx[k] + sy[k] * sy[k] + sz[k] * sz[k];
    m2x[k] = -2.f * sx[k]; m2y[k] = -2.f * sy[k]; m2z[k] = -2.f * sz[k];
    smin[k] = INFINITY; amin[k] = INFINITY; jbest[k] = 0x7fffffff;
    if (EXACT) {
      const float b = __uint_as_float(bound[s]);
      float margin = 1e-6f * (ss[k] + 2.f * __uint_as_float(*qqmax_g) + b);
      if (tc_norm2) margin += 8e-6f * (tc_norm2[s] + 2.f * __uint_as_float(tc_stats[0]) + b);
      thr[k] = (b - ss[k]) + margin;
    } else {
      thr[k] = 0.f;
    }
  }
  const int j_begin = blockIdx.y * dst_per_block;
  const int j_end = min(j_begin + dst_per_block, n_dst);
  float qm = 0.f;
  for (int base = j_begin; base < j_end; base += CH_TILE) {
    const int cnt = min(CH_TILE, j_end - base);
    __syncthreads();
    for (int t = threadIdx.x; t < CH_TILE; t += CH_THREADS) {
      float4 v = make_float4(0.f, 0.f, 0.f, INFINITY);  // padding never wins
      if (t < cnt) {
        const float *p = dst + 3 * (size_t)(base + t);
        v.x = p[0]; v.y = p[1]; v.z = p[2];
        v.w = v.x * v.x + v.y * v.y + v.z * v.z;
        qm = fmaxf(qm, v.w);
      }
      tile[t] = v;
    }
    __syncthreads();
    for (int t = 0; t < cnt; t += 4) {
      float a[CH_SPT][4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const float4 q = tile[t + u];
#pragma unroll
        for (int k = 0; k < CH_SPT; ++k) a[k][u] = fmaf(m2x[k], q.x, fmaf(m2y[k], q.y, fmaf(m2z[k], q.z, q.w)));
      }
      if (!EXACT) {
#pragma unroll
        for (int k = 0; k < CH_SPT; ++k) amin[k] = fminf(fminf(amin[k], fminf(a[k][0], a[k][1])), fminf(a[k][2], a[k][3]));
      } else {
        float worst = INFINITY;
#pragma unroll
        for (int k = 0; k < CH_SPT; ++k) worst = fminf(worst, fminf(fminf(a[k][0], a[k][1]), fminf(a[k][2], a[k][3])) - thr[k]);
        if (worst < 0.f) {  // cold path
#pragma unroll 1
          for (int u = 0; u < 4; ++u) {
            const float4 q = tile[t + u];
#pragma unroll
            for (int k = 0; k < CH_SPT; ++k) {
              const float ak = fmaf(m2x[k], q.x, fmaf(m2y[k], q.y, fmaf(m2z[k], q.z, q.w)));
              if (ak < thr[k]) {
                const float dx = __fsub_rn(sx[k], q.x), dy = __fsub_rn(sy[k], q.y), dz = __fsub_rn(sz[k], q.z);
                const float d2 = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
                const float s = __fsqrt_rn(d2);
                if (s < smin[k]) {  // strictly smaller rounded distance: first index wins ties, as torch.argmin
                  smin[k] = s;
                  jbest[k] = base + t + u;
                }
              }
            }
          }
        }
      }
    }
  }
  if (!EXACT) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) qm = fmaxf(qm, __shfl_xor_sync(0xffffffffu, qm, o));
    if ((threadIdx.x & 31) == 0) qqmax_s[threadIdx.x >> 5] = qm;
    __syncthreads();
    if (threadIdx.x == 0) {
      float q = 0.f;
      for (int w = 0; w < CH_THREADS / 32; ++w) q = fmaxf(q, qqmax_s[w]);
      atomicMax(qqmax_g, __float_as_uint(q));
    }
#pragma unroll
    for (int k = 0; k < CH_SPT; ++k)
      if (s0 + k < n_src) atomicMin(bound + s0 + k, __float_as_uint(fmaxf(amin[k] + ss[k], 0.f)));
  } else {
#pragma unroll
    for (int k = 0; k < CH_SPT; ++k) {
      const int s = s0 + k;
      if (s < n_src && jbest[k] != 0x7fffffff) {
        const unsigned long long key = ((unsigned long long)__float_as_uint(smin[k]) << 32) | (unsigned int)jbest[k];
        atomicMin(best + s, key);
      }
    }
  }
}

// Also the safety net of the screen: a source for which pass 2 found no pair (its bound was violated; never seen, would
// need a pass-1 error beyond the margin) is searched exhaustively here instead of returning garbage.
__global__ void chamfer_unpack_kernel(const unsigned long long *best, int n, int64_t *idx, float *__restrict__ dist,
                                      const float *__restrict__ src, const float *__restrict__ dst, int n_dst) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    unsigned long long k = best[i];
    if (k == ~0ull) {
      const float sx = src[3 * i], sy = src[3 * i + 1], sz = src[3 * i + 2];
      float smin = INFINITY;
      unsigned int jb = 0;
      for (int j = 0; j < n_dst; ++j) {
        const float dx = __fsub_rn(sx, dst[3 * j]), dy = __fsub_rn(sy, dst[3 * j + 1]), dz = __fsub_rn(sz, dst[3 * j + 2]);
        const float d = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
        if (d < smin) { smin = d; jb = j; }
      }
      k = ((unsigned long long)__float_as_uint(smin) << 32) | jb;
    }
    if (dist) dist[i] = __uint_as_float((unsigned int)(k >> 32));
    idx[i] = (int64_t)(unsigned int)(k & 0xffffffffull);  // written last: idx may alias best
  }
}

// model/losses.py:145-150
__global__ void chamfer_err_fwd_kernel(const float *__restrict__ X, int n, const float *__restrict__ Y,
                                       const int64_t *__restrict__ idx, float inv_n, float *__restrict__ loss) {
  __shared__ float red[32];
  float acc = 0.f;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int64_t j = idx[i];
    const float dx = X[3 * i] - Y[3 * j], dy = X[3 * i + 1] - Y[3 * j + 1], dz = X[3 * i + 2] - Y[3 * j + 2];
    acc += sqrtf(dx * dx + dy * dy + dz * dz);
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) atomicAdd(loss, acc * inv_n);
}

__global__ void chamfer_err_bwd_kernel(const float *__restrict__ X, int n, const float *__restrict__ Y,
                                       const int64_t *__restrict__ idx, float inv_n, const float *__restrict__ gl,
                                       float *__restrict__ dX, float *__restrict__ dY) {
  const float g = gl[0] * inv_n;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int64_t j = idx[i];
    const float dx = X[3 * i] - Y[3 * j], dy = X[3 * i + 1] - Y[3 * j + 1], dz = X[3 * i + 2] - Y[3 * j + 2];
    const float nr = sqrtf(dx * dx + dy * dy + dz * dz);
    const float k = nr > 0.f ? g / nr : 0.f;  // zero subgradient at coincidence, like torch
    const float gx = k * dx, gy = k * dy, gz = k * dz;
    if (dX) { dX[3 * i] = gx; dX[3 * i + 1] = gy; dX[3 * i + 2] = gz; }
    if (dY) { atomicAdd(dY + 3 * j, -gx); atomicAdd(dY + 3 * j + 1, -gy); atomicAdd(dY + 3 * j + 2, -gz); }
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// The two clouds as the Chamfer term sees them (model/training.py:394-399): X = (pc1 / s) R^T + t, Y = pc2 / s.
// One launch forward; one launch backward that also reduces dR, dt and ds (block partial sums, then 13 atomics per block).
// ---------------------------------------------------------------------------------------------------------------------
__global__ void align_clouds_fwd_kernel(const float *__restrict__ pc1, int n1, const float *__restrict__ pc2, int n2,
                                        const float *__restrict__ Rt, const float *__restrict__ scale, float *__restrict__ X,
                                        float *__restrict__ Y) {
  const float s = scale ? scale[0] : 1.0f;
  float R[9], t[3];
#pragma unroll
  for (int r = 0; r < 3; ++r) {
#pragma unroll
    for (int c = 0; c < 3; ++c) R[3 * r + c] = Rt[4 * r + c];
    t[r] = Rt[4 * r + 3];
  }
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n1 + n2; i += gridDim.x * blockDim.x) {
    if (i < n1) {
      // same operation order as the reference: divide, then the (1x3)(3x3) product accumulated left to right, then + t
      const float q0 = pc1[3 * i] / s, q1 = pc1[3 * i + 1] / s, q2 = pc1[3 * i + 2] / s;
#pragma unroll
      for (int r = 0; r < 3; ++r) X[3 * i + r] = fmaf(q2, R[3 * r + 2], fmaf(q1, R[3 * r + 1], q0 * R[3 * r])) + t[r];
    } else {
      const int j = i - n1;
#pragma unroll
      for (int c = 0; c < 3; ++c) Y[3 * j + c] = pc2[3 * j + c] / s;
    }
  }
}

// acc[0..8] = dR (row-major 3x3), acc[9..11] = dt, acc[12] = ds
__global__ void __launch_bounds__(256) align_clouds_bwd_kernel(const float *__restrict__ pc1, int n1, const float *__restrict__ pc2,
                                                               int n2, const float *__restrict__ Rt, const float *__restrict__ scale,
                                                               const float *__restrict__ dX, const float *__restrict__ dY,
                                                               float *__restrict__ d_pc1, float *__restrict__ d_pc2,
                                                               float *__restrict__ d_Rt, float *__restrict__ d_scale) {
  const float s = scale ? scale[0] : 1.0f, inv = 1.0f / s;
  float R[9];
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int c = 0; c < 3; ++c) R[3 * r + c] = Rt[4 * r + c];
  float acc[13];
#pragma unroll
  for (int k = 0; k < 13; ++k) acc[k] = 0.0f;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n1 + n2; i += gridDim.x * blockDim.x) {
    if (i < n1) {
      const float p[3] = {pc1[3 * i], pc1[3 * i + 1], pc1[3 * i + 2]};
      const float g[3] = {dX[3 * i], dX[3 * i + 1], dX[3 * i + 2]};
      float dq[3];
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        dq[c] = g[0] * R[c] + g[1] * R[3 + c] + g[2] * R[6 + c];   // R^T dX_i
        d_pc1[3 * i + c] = dq[c] * inv;
      }
#pragma unroll
      for (int r = 0; r < 3; ++r) {
#pragma unroll
        for (int c = 0; c < 3; ++c) acc[3 * r + c] += g[r] * (p[c] * inv);
        acc[9 + r] += g[r];
      }
      acc[12] -= (dq[0] * p[0] + dq[1] * p[1] + dq[2] * p[2]) * inv * inv;
    } else {
      const int j = i - n1;
      float dot = 0.0f;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float g = dY[3 * j + c];
        d_pc2[3 * j + c] = g * inv;
        dot += g * pc2[3 * j + c];
      }
      acc[12] -= dot * inv * inv;
    }
  }
  __shared__ float part[8][13];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 13; ++k) {
    float v = acc[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) part[warp][k] = v;
  }
  __syncthreads();
  if (threadIdx.x < 13) {
    float v = 0.0f;
#pragma unroll
    for (int w = 0; w < 8; ++w) v += part[w][threadIdx.x];
    const int k = threadIdx.x;
    if (k < 9) atomicAdd(d_Rt + 4 * (k / 3) + (k % 3), v);
    else if (k < 12) atomicAdd(d_Rt + 4 * (k - 9) + 3, v);
    else if (d_scale) atomicAdd(d_scale, v);
  }
}

extern "C" {

size_t nnb_chamfer_workspace_bytes(int n_src, int n_dst) {
  if (n_src < 0) n_src = 0;
  if (n_dst < 0) n_dst = 0;
  return (((size_t)n_src + 16) * sizeof(unsigned int) + 255) / 256 * 256 + nnb_chamfer_tc_workspace_bytes(n_src, n_dst);
}

static int g_chamfer_pass1_cuda_cores = 0;
// test hook: 0 = both passes on tcgen05 (production); 1 = both on the CUDA cores; 2 = tcgen05 screen + CUDA-core exact pass
int nnb_debug_chamfer_pass1(int cuda_cores) {
  const int old = g_chamfer_pass1_cuda_cores;
  g_chamfer_pass1_cuda_cores = cuda_cores;
  return old;
}

// `idx` (n_src int64) doubles as the 64-bit packed-key scratch; `workspace` holds the pass-1 bounds and max |q|^2.
int nnb_chamfer_nn(const float *src, int n_src, const float *dst, int n_dst, int64_t *idx, float *dist, void *workspace,
                   size_t workspace_bytes, void *stream) {
  if (!src || !dst || !idx || !workspace || n_src <= 0 || n_dst <= 0) return NNB_ERR_BAD_ARG;
  if (workspace_bytes < nnb_chamfer_workspace_bytes(n_src, n_dst)) return NNB_ERR_WORKSPACE;
  cudaStream_t s = nnb_stream(stream);
  unsigned long long *best = reinterpret_cast<unsigned long long *>(idx);
  unsigned int *bound = static_cast<unsigned int *>(workspace), *qqmax = bound + n_src, *tc_stats = qqmax + 1;
  uint8_t *tc_ws = static_cast<uint8_t *>(workspace) + (((size_t)n_src + 16) * sizeof(unsigned int) + 255) / 256 * 256;
  const float *tc_norm2 = reinterpret_cast<const float *>(tc_ws + nnb_chamfer_tc_norm2_offset(n_src, n_dst));
  if (g_chamfer_pass1_cuda_cores == 1) {
    NNB_CUDA(cudaMemsetAsync(bound, 0x7f, (size_t)n_src * sizeof(unsigned int), s));
    NNB_CUDA(cudaMemsetAsync(qqmax, 0, sizeof(unsigned int), s));
    NNB_CUDA(cudaMemsetAsync(best, 0xff, (size_t)n_src * sizeof(unsigned long long), s));
  }
  const int gx = nnb_cdiv(n_src, CH_THREADS * CH_SPT);
  // split the destination range so that the grid covers the SMs a few times over
  int want = nnb_cdiv(nnb_sm_count() * 6, gx);
  int max_split = nnb_cdiv(n_dst, CH_TILE);
  int gy = want < 1 ? 1 : (want > max_split ? max_split : want);
  int per = nnb_cdiv(nnb_cdiv(n_dst, gy), CH_TILE) * CH_TILE;
  gy = nnb_cdiv(n_dst, per);
  if (g_chamfer_pass1_cuda_cores == 1) {
    chamfer_nn_kernel<false><<<dim3(gx, gy), CH_THREADS, 0, s>>>(src, n_src, dst, n_dst, per, bound, qqmax, best, nullptr, nullptr);
    NNB_CHECK_LAUNCH();
    chamfer_nn_kernel<true><<<dim3(gx, gy), CH_THREADS, 0, s>>>(src, n_src, dst, n_dst, per, bound, qqmax, best, nullptr, nullptr);
  } else if (g_chamfer_pass1_cuda_cores == 0) {
    const int rc = nnb_chamfer_tc_bound(src, n_src, dst, n_dst, bound, tc_stats, tc_ws, best, best, 1, s);
    if (rc) return rc;
  } else {
    const int rc = nnb_chamfer_tc_bound(src, n_src, dst, n_dst, bound, tc_stats, tc_ws, best, nullptr, 0, s);
    if (rc) return rc;
    chamfer_nn_kernel<true><<<dim3(gx, gy), CH_THREADS, 0, s>>>(src, n_src, dst, n_dst, per, bound, tc_stats + 1, best, tc_norm2, tc_stats);
  }
  NNB_CHECK_LAUNCH();
  int g = nnb_cdiv(n_src, 256);
  chamfer_unpack_kernel<<<g, 256, 0, s>>>(best, n_src, idx, dist, src, dst, n_dst);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_chamfer_err_fwd(const float *X, int n, const float *Y, const int64_t *idx, float *loss, void *stream) {
  if (!X || !Y || !idx || !loss || n <= 0) return NNB_ERR_BAD_ARG;
  cudaStream_t s = nnb_stream(stream);
  NNB_CUDA(cudaMemsetAsync(loss, 0, sizeof(float), s));
  int g = nnb_cdiv(n, 256 * 2);
  const int cap = nnb_sm_count() * 4;
  if (g > cap) g = cap;
  chamfer_err_fwd_kernel<<<g, 256, 0, s>>>(X, n, Y, idx, 1.0f / (float)n, loss);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_chamfer_err_bwd(const float *X, int n, const float *Y, const int64_t *idx, const float *g_loss, float *d_X,
                        float *d_Y, void *stream) {
  if (!X || !Y || !idx || !g_loss || n <= 0) return NNB_ERR_BAD_ARG;
  int g = nnb_cdiv(n, 256);
  const int cap = nnb_sm_count() * 8;
  if (g > cap) g = cap;
  chamfer_err_bwd_kernel<<<g, 256, 0, nnb_stream(stream)>>>(X, n, Y, idx, 1.0f / (float)n, g_loss, d_X, d_Y);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_align_clouds_fwd(const float *pc1, int n1, const float *pc2, int n2, const float *Rt, const float *scale, float *X, float *Y,
                         void *stream) {
  if (!pc1 || !pc2 || !Rt || !X || !Y || n1 <= 0 || n2 <= 0) return NNB_ERR_BAD_ARG;
  int g = nnb_cdiv(n1 + n2, 256);
  const int cap = nnb_sm_count() * 8;
  if (g > cap) g = cap;
  align_clouds_fwd_kernel<<<g, 256, 0, nnb_stream(stream)>>>(pc1, n1, pc2, n2, Rt, scale, X, Y);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_align_clouds_bwd(const float *pc1, int n1, const float *pc2, int n2, const float *Rt, const float *scale, const float *dX,
                         const float *dY, float *d_pc1, float *d_pc2, float *d_Rt, float *d_scale, void *stream) {
  if (!pc1 || !pc2 || !Rt || !dX || !dY || !d_pc1 || !d_pc2 || !d_Rt || n1 <= 0 || n2 <= 0 || (scale && !d_scale)) return NNB_ERR_BAD_ARG;
  cudaStream_t s = nnb_stream(stream);
  NNB_CUDA(cudaMemsetAsync(d_Rt, 0, 16 * sizeof(float), s));
  if (d_scale) NNB_CUDA(cudaMemsetAsync(d_scale, 0, sizeof(float), s));
  int g = nnb_cdiv(n1 + n2, 256 * 4);
  const int cap = nnb_sm_count();
  if (g > cap) g = cap;
  align_clouds_bwd_kernel<<<g, 256, 0, s>>>(pc1, n1, pc2, n2, Rt, scale, dX, dY, d_pc1, d_pc2, d_Rt, d_scale);
  NNB_CHECK_LAUNCH();
  return 0;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------------
// nnb_chamfer_nn_static: the same nearest-neighbour search for callers that cannot allocate (CUDA graph capture, serving
// loops, harnesses that forbid cudaMalloc).  The workspace nnb_chamfer_nn takes from its caller lives here, in a file-scope
// device array that every GPU loading this library holds until exit.  No counterpart in the
// reference.
//   * limit: n_src, n_dst <= NNB_CHAMFER_STATIC_MAX_POINTS (32,768 each: the training clouds are 28,830 points);
//     larger problems return NNB_ERR_TOO_LARGE and must use nnb_chamfer_nn with a caller-owned workspace.
//   * in host code the array's NAME is a host-side stand-in, not a device address: the device address is fetched once per
//     device with cudaGetSymbolAddress (that device current) and that pointer is what the kernels receive.
//   * every call shares the array; it keeps the previous call's contents.  nnb_chamfer_nn initialises every word of the
//     workspace it reads (bounds, statistics, operand images) before reading it, so nothing has to be cleared between calls --
//     but two calls that overlap on the GPU (different streams, or two host threads) RACE on it.  Serialise them.
//   * calls no cudaMalloc / cudaFree / allocating library routine.
// ---------------------------------------------------------------------------------------------------------------------
namespace {
constexpr int CH_STATIC_MAX_POINTS = NNB_CHAMFER_STATIC_MAX_POINTS;
constexpr size_t CH_STATIC_BYTES = 4724480 + 256;   // nnb_chamfer_workspace_bytes(32768, 32768) = 4,724,480 (checked at run time below)
constexpr int CH_STATIC_MAX_DEVICES = 64;
}  // namespace
static __device__ __align__(256) unsigned char g_chamfer_static_scratch[CH_STATIC_BYTES];
static void *g_chamfer_static_ptr[CH_STATIC_MAX_DEVICES] = {nullptr};

extern "C" int nnb_chamfer_nn_static(const float *src, int n_src, const float *dst, int n_dst, int64_t *idx, float *dist,
                                     void *stream) {
  if (!src || !dst || !idx || n_src <= 0 || n_dst <= 0) return NNB_ERR_BAD_ARG;
  if (n_src > CH_STATIC_MAX_POINTS || n_dst > CH_STATIC_MAX_POINTS) return NNB_ERR_TOO_LARGE;
  if (nnb_chamfer_workspace_bytes(n_src, n_dst) > CH_STATIC_BYTES) return NNB_ERR_TOO_LARGE;
  int dev = 0;
  NNB_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= CH_STATIC_MAX_DEVICES) return NNB_ERR_UNSUPPORTED;
  if (!g_chamfer_static_ptr[dev]) {
    void *p = nullptr;
    NNB_CUDA(cudaGetSymbolAddress(&p, g_chamfer_static_scratch));   // this device's copy of the array (device `dev` is current)
    if (!p) return NNB_ERR_UNSUPPORTED;
    g_chamfer_static_ptr[dev] = p;
  }
  return nnb_chamfer_nn(src, n_src, dst, n_dst, idx, dist, g_chamfer_static_ptr[dev], CH_STATIC_BYTES, stream);
}
