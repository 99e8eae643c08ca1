// Stage 5, screening pass on the tensor cores: bound[i] = min_j ( |q_j|^2 - 2 s_i.q_j ) + |s_i|^2  ~  min_j d2(i, j).
//
// The -2 s.q + |q|^2 form is one GEMM row x column product.  To keep ~fp32 accuracy on bf16 tensor cores every
// coordinate is split three ways (x = h + m + l, bf16 each, 24 bits together) after centring both clouds on dst[0]
// (shrinks the magnitudes that drive cancellation).  Per coordinate the six largest cross terms are kept
//     u q ~ uh qh + uh qm + um qh + uh ql + ul qh + um qm        (dropped terms <= 2^-24 |u||q|)
// and |q|^2 rides in three more K slots against constant ones: K = 3*6 + 3 = 21 <= 32 = two tcgen05.mma k-steps.
// One CTA owns 128 sources (A image, thread = source row) and streams [256 x 32] destination images through a bulk-TMA
// ring; accumulators are double-buffered in TMEM (2 x 256 columns) so the min-epilogue of one stage (tcgen05.ld +
// 3-input min) overlaps the MMA of the next.  The result only BOUNDS the search: the exact pass in chamfer.cu
// re-evaluates every pair within the error margin with the reference's op order, so indices stay bit-exact.
#include "nnb_common.cuh"
#include "tc_common.cuh"
#include <cuda_fp16.h>

using namespace tc;

namespace {

constexpr int CT_RING = 8;
// Coherent clouds (back-projected depth maps: 256 consecutive destinations = a thin row segment) let the sweeps skip most
// stages.  The fp16 path therefore runs: (1) a screen over every CT_SUB-th destination -> a loose upper bound per source;
// (2) chamfer_tc_filter_kernel: per source block, the list of stages whose bounding box comes within that bound of any of
// its sources; (3) the screen over the listed stages -> the true minimum; (4) the exact sweep over the same lists.
// Clouds in random order flag every stage and pay (1) + (2) on top (about 1/CT_SUB of a sweep).
constexpr int CT_SUB = 8;
constexpr uint32_t CT_STAGE = 16384;                 // 256 destinations x 32 K x bf16
constexpr uint32_t CS_A = 0;                         // 2 x 8192 (128 sources x 32 K)
constexpr uint32_t CS_RING = 16384;
constexpr uint32_t CS_RED = CS_RING + CT_RING * CT_STAGE;  // 128 floats: column-half combine
constexpr uint32_t CS_BARS = CS_RED + 512;
constexpr uint32_t CS_TOTAL = CS_BARS + 32 * 8 + 16;

__device__ __forceinline__ uint32_t pack2(__nv_bfloat16 a, __nv_bfloat16 b) {
  return (uint32_t)__bfloat16_as_ushort(a) | ((uint32_t)__bfloat16_as_ushort(b) << 16);
}

__device__ __forceinline__ void split3(float x, __nv_bfloat16 &h, __nv_bfloat16 &m, __nv_bfloat16 &l) {
  h = __float2bfloat16_rn(x);
  const float r1 = x - __bfloat162float(h);
  m = __float2bfloat16_rn(r1);
  l = __float2bfloat16_rn(r1 - __bfloat162float(m));
}

// one thread per point: A images for the sources (threads [0, ns_pad)), B images for the destinations (the rest);
// also resets bound[] / best[] for the sweeps
__global__ void chamfer_tc_prep_kernel(const float *__restrict__ src, int n_src, int ns_pad, const float *__restrict__ dst, int n_dst,
                                       int nd_pad, uint8_t *__restrict__ a_img, uint8_t *__restrict__ b_img, float *__restrict__ norm2,
                                       unsigned int *__restrict__ stats, unsigned int *__restrict__ bound, unsigned long long *__restrict__ best) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= ns_pad + nd_pad) return;
  const bool is_src = i < ns_pad;
  if (!is_src) i -= ns_pad;
  const float *pts = is_src ? src : dst;
  const int n = is_src ? n_src : n_dst;
  const int rows = is_src ? 128 : 256;
  uint8_t *base = (is_src ? a_img : b_img) + (size_t)(i / rows) * (rows * 64);
  const int r = i % rows;
  __nv_bfloat16 e[32];
  const __nv_bfloat16 zero = __float2bfloat16_rn(0.f), one = __float2bfloat16_rn(1.f);
#pragma unroll
  for (int k = 0; k < 32; ++k) e[k] = zero;
  float nn = 0.f, raw2 = 0.f;
  if (i < n) {
    float p[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float x = pts[3 * (size_t)i + c];
      raw2 += x * x;
      p[c] = x - dst[c];   // centre = first destination point
      nn += p[c] * p[c];
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      __nv_bfloat16 h, m, l;
      split3(is_src ? -2.f * p[c] : p[c], h, m, l);
      if (is_src) { e[6 * c] = h; e[6 * c + 1] = h; e[6 * c + 2] = m; e[6 * c + 3] = h; e[6 * c + 4] = l; e[6 * c + 5] = m; }
      else        { e[6 * c] = h; e[6 * c + 1] = m; e[6 * c + 2] = h; e[6 * c + 3] = l; e[6 * c + 4] = h; e[6 * c + 5] = m; }
    }
    if (is_src) { e[18] = one; e[19] = one; e[20] = one; }
    else { split3(nn, e[18], e[19], e[20]); }
  } else if (!is_src) {
    e[18] = __float2bfloat16_rn(1e30f);   // padding destinations never win (sources see it through their ones)
  } else {
    e[18] = one;
  }
  // row r of the image: four 16-byte core-matrix rows, 128 B apart
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    uint4 v;
    v.x = pack2(e[8 * g], e[8 * g + 1]); v.y = pack2(e[8 * g + 2], e[8 * g + 3]);
    v.z = pack2(e[8 * g + 4], e[8 * g + 5]); v.w = pack2(e[8 * g + 6], e[8 * g + 7]);
    *reinterpret_cast<uint4 *>(base + img_off(r, 8 * g, 32)) = v;
  }
  if (is_src) {
    if (i < n) {
      norm2[i] = nn;
      bound[i] = 0x7f7f7f7fu;
      if (best) best[i] = ~0ull;
    }
  } else if (i < n) {
    atomicMax(stats, __float_as_uint(nn));       // max |q - c|^2 (error scale of the tensor-core screen)
    atomicMax(stats + 1, __float_as_uint(raw2)); // max |q|^2     (error scale of the CUDA-core screen)
  }
}

// ---- fp16 two-way split (production screen): one K = 16 MMA per stage ------------------------------------------------
// Both clouds are centred on the centroid of the destinations (chamfer_tc_centre_kernel; the first destination -- a corner pixel
// of a depth cloud -- made every scaled norm, and with it every error margin, several times larger) and multiplied by a power
// of two (exact) that brings the largest |coordinate| of either cloud into [16, 32): fp16's range (nn' < 3,072, |-2 s'| < 64)
// and its normal range for the low halves.
// x = h + m (fp16 each, 22 bits).  Per coordinate all four cross terms ride in the K slots: source [h, h, m, m] against
// destination [h, m, h, m]; |q'|^2 split three ways against ones in slots 12..14: K = 15 of 16.
// Error of a' = |q'|^2 - 2 s'.q' against exact arithmetic on the scaled points:
//   representation  <= 2^-21 (|s'|^2 + |q'|^2)     (2^-22 relative per factor)
//   fp16 subnormals <= 2^-25 per low half -> <= 8.5e-6 absolute
//   fp32 accumulation of 15 exact products <= 2^-20 (|s'|^2 + 2 |q'|^2)
// i.e. err(s, j) <= 1.5e-6 |s'|^2 + 2.4e-6 |q_j'|^2 + 8.5e-6.  The sixteenth K slot carries the destination's own share of the margin,
// e_j = 8e-6 |q_j'|^2 (fp16, rounded up): the screening sweeps run with +1 in the source's slot 15 and minimise a_j + e_j, the exact
// sweep runs with -1 (a second source image) and tests a_j - e_j < min_k (a_k + e_k) + 8e-6 (|s'|^2 + bound) + 2e-5.  The true
// nearest neighbour j* passes: a*_j <= a*_k for all k (exact values), so a_j* - e_j* <= a_k + e_k + [err(s,j*) - e_j*] + [err(s,k) - e_k]
// and each bracket is below half the source's margin.  A destination near the centre is thus tested against ITS error scale, not
// the cloud's: with max |q'|^2 in the threshold (as before) a cloud whose bulk sits in a small cap around the centre while a few
// points set the extent -- sparse depth maps back-projected with the nearest_limit clamp -- sent hundreds of pairs per source to
// the exact re-evaluation.
__device__ __forceinline__ uint32_t pack2h(__half a, __half b) {
  return (uint32_t)__half_as_ushort(a) | ((uint32_t)__half_as_ushort(b) << 16);
}
__device__ __forceinline__ void split2h(float x, __half &h, __half &m) {
  h = __float2half_rn(x);
  m = __float2half_rn(x - __half2float(h));
}

// stats[6..8] += sum of the finite destinations, stats[9] += their number (float): the centre of the scaled frame.  The sums are
// atomic (order-dependent in the last bits): the centre only shapes the screen, the exact sweep decides indices and distances.
__global__ void chamfer_tc_centre_kernel(const float *__restrict__ dst, int n_dst, unsigned int *__restrict__ stats) {
  float a[4] = {0.f, 0.f, 0.f, 0.f};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_dst; i += gridDim.x * blockDim.x) {
    const float x = dst[3 * (size_t)i], y = dst[3 * (size_t)i + 1], z = dst[3 * (size_t)i + 2];
    if (isfinite(x) && isfinite(y) && isfinite(z)) { a[0] += x; a[1] += y; a[2] += z; a[3] += 1.f; }
  }
#pragma unroll
  for (int c = 0; c < 4; ++c) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) a[c] += __shfl_xor_sync(0xffffffffu, a[c], o);
  }
  if ((threadIdx.x & 31) == 0 && a[3] > 0.f) {
#pragma unroll
    for (int c = 0; c < 4; ++c) atomicAdd(reinterpret_cast<float *>(stats + 6 + c), a[c]);
  }
}
__device__ __forceinline__ void tc_centre(const unsigned int *__restrict__ stats, const float *__restrict__ dst, float (&c)[3]) {
  const float n = __uint_as_float(stats[9]);
#pragma unroll
  for (int k = 0; k < 3; ++k) c[k] = n > 0.f ? __uint_as_float(stats[6 + k]) / n : dst[k];
}

// stats[2] = float bits of max |p - centre|_inf over both clouds
__global__ void chamfer_tc_extent_kernel(const float *__restrict__ src, int n_src, const float *__restrict__ dst, int n_dst,
                                         unsigned int *__restrict__ stats) {
  __shared__ float red[32];
  float m = 0.f;
  float ctr[3];
  tc_centre(stats, dst, ctr);
  const float cx = ctr[0], cy = ctr[1], cz = ctr[2];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_src + n_dst; i += gridDim.x * blockDim.x) {
    const float *p = i < n_src ? src + 3 * (size_t)i : dst + 3 * (size_t)(i - n_src);
    m = fmaxf(m, fmaxf(fabsf(p[0] - cx), fmaxf(fabsf(p[1] - cy), fabsf(p[2] - cz))));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x < 32) {
    m = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (threadIdx.x == 0) atomicMax(stats + 2, __float_as_uint(m));
  }
}

__device__ __forceinline__ float extent_scale(unsigned int bits) {
  const float m = __uint_as_float(bits);
  if (!(m > 0.f) || !isfinite(m)) return 1.f;
  int e;
  frexpf(m, &e);            // m = f * 2^e, f in [0.5, 1)
  return ldexpf(1.f, 5 - e);  // scaled extent in [16, 32)
}

// one thread per point: [128 x 16] fp16 A images for the sources, [256 x 16] B images for the destinations (scaled units)
// Threads [ns_pad + nd_pad, + nsub_pad) build a THIRD set of images, every CT_SUB-th destination: the screening sweep runs
// over those only (1/CT_SUB of the pairs); its minimum is an upper bound of the true one, which is all the exact sweep needs.
__global__ void chamfer_tc_prep16_kernel(const float *__restrict__ src, int n_src, int ns_pad, const float *__restrict__ dst, int n_dst,
                                         int nd_pad, int nsub_pad, int sub, uint8_t *__restrict__ a_img, uint8_t *__restrict__ a_img_x,
                                         uint8_t *__restrict__ b_img, uint8_t *__restrict__ sub_img, float *__restrict__ norm2,
                                         unsigned int *__restrict__ stats, unsigned int *__restrict__ bound,
                                         unsigned long long *__restrict__ best) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= ns_pad + nd_pad + nsub_pad) return;
  const bool is_src = i < ns_pad;
  const bool is_sub = i >= ns_pad + nd_pad;
  if (!is_src) i -= ns_pad;
  if (is_sub) i -= nd_pad;
  const float *pts = is_src ? src : dst;
  const int n = is_src ? n_src : (is_sub ? (n_dst + sub - 1) / sub : n_dst);
  const int rows = is_src ? 128 : 256;
  uint8_t *base = (is_src ? a_img : (is_sub ? sub_img : b_img)) + (size_t)(i / rows) * (rows * 32);
  const int r = i % rows;
  const size_t pi = is_sub ? (size_t)i * sub : (size_t)i;   // point index
  const float scale = extent_scale(stats[2]);
  float ctr[3];
  tc_centre(stats, dst, ctr);
  __half e[16];
  const __half zero = __float2half_rn(0.f), one = __float2half_rn(1.f);
#pragma unroll
  for (int k = 0; k < 16; ++k) e[k] = zero;
  float nn = 0.f;
  if (i < n) {
    float p[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      p[c] = (pts[3 * pi + c] - ctr[c]) * scale;   // centre = centroid of the destinations; power-of-two scale: exact
      nn += p[c] * p[c];
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      __half h, m;
      split2h(is_src ? -2.f * p[c] : p[c], h, m);
      if (is_src) { e[4 * c] = h; e[4 * c + 1] = h; e[4 * c + 2] = m; e[4 * c + 3] = m; }
      else        { e[4 * c] = h; e[4 * c + 1] = m; e[4 * c + 2] = h; e[4 * c + 3] = m; }
    }
    if (is_src) { e[12] = one; e[13] = one; e[14] = one; e[15] = one; }
    else {
      e[12] = __float2half_rn(nn);
      const float r1 = nn - __half2float(e[12]);
      e[13] = __float2half_rn(r1);
      e[14] = __float2half_rn(r1 - __half2float(e[13]));
      e[15] = __float2half_ru(8e-6f * nn);     // this destination's share of the error margin (see above)
    }
  } else if (!is_src) {
    e[12] = __float2half_rn(60000.f);   // padding destinations never win (real a' < 10^4)
  } else {
    e[12] = one;
  }
#pragma unroll
  for (int g = 0; g < 2; ++g) {
    uint4 v;
    v.x = pack2h(e[8 * g], e[8 * g + 1]); v.y = pack2h(e[8 * g + 2], e[8 * g + 3]);
    v.z = pack2h(e[8 * g + 4], e[8 * g + 5]); v.w = pack2h(e[8 * g + 6], e[8 * g + 7]);
    *reinterpret_cast<uint4 *>(base + img_off(r, 8 * g, 16)) = v;
    if (is_src) {   // the exact sweep's source image: slot 15 negated
      if (g == 1) v.w = pack2h(e[14], __hneg(e[15]));
      *reinterpret_cast<uint4 *>(a_img_x + (size_t)(i / rows) * (rows * 32) + img_off(r, 8 * g, 16)) = v;
    }
  }
  if (is_src) {
    if (i < n) {
      norm2[i] = nn;
      bound[i] = 0x7f7f7f7fu;
      if (best) best[i] = ~0ull;
    }
  } else if (i < n && !is_sub) {
    atomicMax(stats, __float_as_uint(nn));   // max |q'|^2 in scaled units
  }
}

// bounding box of every 256-destination stage, raw coordinates: box[st] = {min xyz, max xyz} (8 floats, 2 unused)
// The last block to finish also decides whether stage lists can pay: stats[3] = 1 iff the stage boxes are selective, i.e. their
// mean volume is below 5 % of the cloud's bounding box (axes along which the whole cloud is flat do not count).  Raster-order
// depth clouds give ~1e-3; clouds in random order give ~1 and run the two plain sweeps (stats[5] = completion counter).
__global__ void __launch_bounds__(128) chamfer_tc_boxes_kernel(const float *__restrict__ dst, int n_dst, int n_stages, float *__restrict__ box,
                                                               unsigned int *__restrict__ stats) {
  __shared__ float red[6][4];
  __shared__ bool last;
  const int st = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (int j = st * 256 + lane; st < n_stages && j < min(n_dst, st * 256 + 256); j += 32) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float v = dst[3 * (size_t)j + c];
      lo[c] = fminf(lo[c], v);
      hi[c] = fmaxf(hi[c], v);
    }
  }
#pragma unroll
  for (int c = 0; c < 3; ++c) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo[c] = fminf(lo[c], __shfl_xor_sync(0xffffffffu, lo[c], o));
      hi[c] = fmaxf(hi[c], __shfl_xor_sync(0xffffffffu, hi[c], o));
    }
  }
  if (lane == 0 && st < n_stages) {
    float *b = box + 8 * (size_t)st;
    b[0] = lo[0]; b[1] = lo[1]; b[2] = lo[2]; b[3] = hi[0]; b[4] = hi[1]; b[5] = hi[2];
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) last = atomicAdd(stats + 5, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!last) return;
  __threadfence();
  // global bounding box, then the mean volume fraction of the stage boxes
  const int t = threadIdx.x, w = t >> 5;
  float g[6] = {INFINITY, INFINITY, INFINITY, -INFINITY, -INFINITY, -INFINITY};
  for (int k = t; k < n_stages; k += blockDim.x) {
    const volatile float *b = box + 8 * (size_t)k;
#pragma unroll
    for (int c = 0; c < 3; ++c) { g[c] = fminf(g[c], b[c]); g[3 + c] = fmaxf(g[3 + c], b[3 + c]); }
  }
#pragma unroll
  for (int c = 0; c < 6; ++c) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float v = __shfl_xor_sync(0xffffffffu, g[c], o);
      g[c] = c < 3 ? fminf(g[c], v) : fmaxf(g[c], v);
    }
    if (lane == 0) red[c][w] = g[c];
  }
  __syncthreads();
  float G[3];
#pragma unroll
  for (int c = 0; c < 3; ++c)
    G[c] = fmaxf(fmaxf(red[3 + c][0], red[3 + c][1]), fmaxf(red[3 + c][2], red[3 + c][3])) -
           fminf(fminf(red[c][0], red[c][1]), fminf(red[c][2], red[c][3]));
  float acc = 0.f;
  for (int k = t; k < n_stages; k += blockDim.x) {
    const volatile float *b = box + 8 * (size_t)k;
    float f = 1.f;
#pragma unroll
    for (int c = 0; c < 3; ++c)
      if (G[c] > 0.f) f *= fminf(fmaxf(b[3 + c] - b[c], 1e-3f * G[c]) / G[c], 1.f);
    acc += f;
  }
  __syncthreads();
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) red[0][w] = acc;
  __syncthreads();
  if (t == 0) stats[3] = ((red[0][0] + red[0][1] + red[0][2] + red[0][3]) / (float)n_stages < 0.05f) ? 1u : 0u;
}

// One block per source block (128 sources, one per thread).  Stage st is listed iff some source s of the block has
//   dist^2(s, box[st]) * scale^2 * (1 - 1e-5) <= bound[s] + margin(s)          (scaled units, margin as in the exact sweep)
// i.e. the stage may hold a destination within the source's current upper bound.  The true nearest neighbour (and every
// tie) always is: d'^2 <= bound + margin / 2.  Lists are ascending; list_cnt[sb] entries.
__global__ void __launch_bounds__(128) chamfer_tc_filter_kernel(const float *__restrict__ src, int n_src, int n_stages,
                                                                const float *__restrict__ box, const unsigned int *__restrict__ bound,
                                                                const float *__restrict__ norm2, const unsigned int *__restrict__ stats,
                                                                int *__restrict__ list, int *__restrict__ list_cnt) {
  __shared__ float4 sp[128];
  __shared__ int warp_cnt[4], base_s;
  if (stats[3] == 0) return;   // boxes not selective: the sweeps ignore the lists
  const int sb = blockIdx.x, t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int s = sb * 128 + t;
  const float scale = extent_scale(stats[2]);
  {
    float4 v = make_float4(0.f, 0.f, 0.f, -1.f);   // r2 < 0: padding row, never matches
    if (s < n_src) {
      const float b = __uint_as_float(bound[s]), nn = norm2[s];
      const float margin = 8e-6f * (nn + 2.f * __uint_as_float(stats[0]) + b) + 2e-5f;
      v = make_float4(src[3 * (size_t)s], src[3 * (size_t)s + 1], src[3 * (size_t)s + 2], (b + margin) / (scale * scale * (1.f - 1e-5f)));
    }
    sp[t] = v;
  }
  if (t == 0) base_s = 0;
  __syncthreads();
  int *out = list + (size_t)sb * n_stages;
  for (int st0 = 0; st0 < n_stages; st0 += 128) {
    const int st = st0 + t;
    bool hit = false;
    if (st < n_stages) {
      const float *b = box + 8 * (size_t)st;
      const float lx = b[0], ly = b[1], lz = b[2], hx = b[3], hy = b[4], hz = b[5];
      for (int k = 0; k < 128 && !hit; ++k) {
        const float4 p = sp[k];
        const float dx = fmaxf(fmaxf(lx - p.x, p.x - hx), 0.f), dy = fmaxf(fmaxf(ly - p.y, p.y - hy), 0.f), dz = fmaxf(fmaxf(lz - p.z, p.z - hz), 0.f);
        hit = dx * dx + dy * dy + dz * dz <= p.w;
      }
    }
    const unsigned m = __ballot_sync(0xffffffffu, hit);
    if (lane == 0) warp_cnt[warp] = __popc(m);
    __syncthreads();
    int off = base_s;
    for (int w = 0; w < warp; ++w) off += warp_cnt[w];
    if (hit) out[off + __popc(m & ((1u << lane) - 1))] = st;
    __syncthreads();
    if (t == 0) base_s += warp_cnt[0] + warp_cnt[1] + warp_cnt[2] + warp_cnt[3];
    __syncthreads();
  }
  if (t == 0) list_cnt[sb] = base_s;
}

struct CtArgs {
  const uint8_t *a_img, *b_img;
  const float *norm2;
  unsigned int *bound;
  int n_src, n_sblocks, n_stages, n_splits, stages_per_split;
  // exact pass
  const float *src, *dst;
  const unsigned int *stats;
  unsigned long long *best;
  int n_dst;
  // optional stage lists (chamfer_tc_filter_kernel): source block sb sweeps list[sb * n_stages + 0 .. list_cnt[sb]) only
  const int *list, *list_cnt;
  // stats[3] (chamfer_tc_boxes_kernel): lists are honoured only when it is 1; gate_skip: this launch (the sub-sampled screen)
  // is needed only then and returns at once otherwise
  const unsigned int *gate;
  int gate_skip;
};

// EXACT = false: bound[s] = min_j a(s, j) + |s - c|^2.
// EXACT = true : same sweep (same MMA shape; F16: the source image with slot 15 negated, see the error model above); every pair
//                with a < (bound - |s - c|^2) + margin is re-evaluated with the reference's arithmetic (direct differences,
//                each op rounded, sqrt_rn) and merged with atomicMin on (bits(dist) << 32 | j): first index of the minimum.
// F16: the fp16 two-way-split images above (one K = 16 MMA per 128 x 256 stage, 8 KB stages); else the bf16 three-way split
// (two MMAs, 16 KB stages).
template <bool EXACT, bool F16>
__global__ void __launch_bounds__(320, 1) chamfer_tc_kernel(const CtArgs A) {
  constexpr uint32_t STAGE = F16 ? 8192u : CT_STAGE, ABYTES = F16 ? 4096u : 8192u, GS = F16 ? 256u : 512u;
  const bool gate_on = A.gate && *A.gate != 0;
  if (A.gate_skip && !gate_on) return;
  const int *const list = gate_on ? A.list : nullptr;
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t sbase = smem_u32(smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float *red = reinterpret_cast<float *>(smem + CS_RED);
  const uint32_t bar_full = sbase + CS_BARS, bar_empty = bar_full + CT_RING * 8, bar_a = bar_empty + CT_RING * 8,
                 bar_acc_full = bar_a + 16, bar_acc_free = bar_acc_full + 16, bar_a_free = bar_acc_free + 16;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + CS_BARS + 32 * 8);
  if (threadIdx.x == 0) {
    for (int i = 0; i < CT_RING; ++i) { mbar_init(bar_full + 8 * i, 1); mbar_init(bar_empty + 8 * i, 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(bar_a + 8 * i, 1); mbar_init(bar_acc_full + 8 * i, 1); mbar_init(bar_acc_free + 8 * i, 256); mbar_init(bar_a_free + 8 * i, 1); }
    mbar_fence_init();
  }
  if (warp == 9) tmem_alloc(smem_u32(tmem_slot), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const int n_items = A.n_sblocks * A.n_splits;
  // positions [k0, k1) of item's source block in its stage list (the whole range 0..n_stages without lists)
  auto item_range = [&](int item, int &sb, int &k0, int &k1) {
    sb = item / A.n_splits;
    const int sp = item % A.n_splits;
    const int cnt = list ? A.list_cnt[sb] : A.n_stages;
    k0 = sp * A.stages_per_split;
    k1 = min(cnt, k0 + A.stages_per_split);
    return k0 < k1;
  };
  auto stage_at = [&](int sb, int k) { return list ? list[(size_t)sb * A.n_stages + k] : k; };

  if (warp == 8) {
    if (elect_one()) {
      uint32_t slot = 0, phase = 0;
      int it = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        int sb, k0, k1;
        if (!item_range(item, sb, k0, k1)) continue;
        int st_next = stage_at(sb, k0);
        for (int k = k0; k < k1; ++k) {
          const int st = st_next;
          if (k + 1 < k1) st_next = stage_at(sb, k + 1);   // list entry fetched one stage ahead of its use
          mbar_wait(bar_empty + 8 * slot, phase ^ 1);
          if (k == k0) {
            // the A buffer of item it-2 is free once that item's last MMA has completed (the issuer commits to bar_a_free)
            if (it >= 2) mbar_wait(bar_a_free + 8 * (it & 1), ((it >> 1) - 1) & 1);
            mbar_arrive_expect_tx(bar_a + 8 * (it & 1), ABYTES);
            bulk_g2s(sbase + CS_A + (it & 1) * 8192, A.a_img + (size_t)sb * ABYTES, ABYTES, bar_a + 8 * (it & 1));
          }
          mbar_arrive_expect_tx(bar_full + 8 * slot, STAGE);
          bulk_g2s(sbase + CS_RING + slot * CT_STAGE, A.b_img + (size_t)st * STAGE, STAGE, bar_full + 8 * slot);
          if (++slot == CT_RING) { slot = 0; phase ^= 1; }
        }
        ++it;
      }
    }
    __syncwarp();
  } else if (warp == 9) {
    if (elect_one()) {
      uint32_t slot = 0, phase = 0, buf = 0, free_par = 0;
      uint32_t n_issued = 0;
      // kind::f16 instruction descriptor: A/B format bits 7..9 / 10..12 = 0 (fp16) or 1 (bf16)
      const uint32_t idesc = F16 ? (idesc_bf16(128, 256, 0, 0) & ~((1u << 7) | (1u << 10))) : idesc_bf16(128, 256, 0, 0);
      int it = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        int sb, k0, k1;
        if (!item_range(item, sb, k0, k1)) continue;
        mbar_wait(bar_a + 8 * (it & 1), (it >> 1) & 1);
        const uint32_t a0 = sbase + CS_A + (it & 1) * 8192;
        for (int k = k0; k < k1; ++k) {
          if (n_issued >= 2) { mbar_wait(bar_acc_free + 8 * buf, (free_par >> buf) & 1); free_par ^= 1u << buf; }
          mbar_wait(bar_full + 8 * slot, phase);
          tc_fence_after();
          const uint32_t b0 = sbase + CS_RING + slot * CT_STAGE;
          umma(tmem + buf * 256, desc_kmajor(a0, GS), desc_kmajor(b0, GS), idesc, 0u);
          if (!F16) umma(tmem + buf * 256, desc_kmajor(a0 + 256, GS), desc_kmajor(b0 + 256, GS), idesc, 1u);
          umma_commit(bar_empty + 8 * slot);
          umma_commit(bar_acc_full + 8 * buf);
          ++n_issued;
          buf ^= 1;
          if (++slot == CT_RING) { slot = 0; phase ^= 1; }
        }
        umma_commit(bar_a_free + 8 * (it & 1));
        ++it;
      }
    }
    __syncwarp();
  } else {
    // 8 epilogue warps: quadrant = warp & 3 (rows), column half = warp >> 2
    const int q = warp & 3, half = warp >> 2, row = q * 32 + lane;
    uint32_t buf = 0, full_par = 0;   // bit b = parity of accumulator buffer b
    for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
      int sb, k0, k1;
      if (!item_range(item, sb, k0, k1)) continue;
      float mrun[4] = {INFINITY, INFINITY, INFINITY, INFINITY};
      const int s_row = sb * 128 + row;
      float thr = -INFINITY, sx = 0.f, sy = 0.f, sz = 0.f, smin = INFINITY;
      unsigned int jbest = 0xffffffffu;
      if (EXACT && s_row < A.n_src) {
        const float b = __uint_as_float(A.bound[s_row]), nn = A.norm2[s_row];
        thr = (b - nn) + (F16 ? 8e-6f * (nn + b) + 2e-5f                                  // destinations carry their own share (slot 15)
                              : 8e-6f * (nn + 2.f * __uint_as_float(A.stats[0]) + b));
        sx = A.src[3 * s_row]; sy = A.src[3 * s_row + 1]; sz = A.src[3 * s_row + 2];
      }
      // one 64-column chunk: running minimum (screen) or threshold test + exact re-evaluation of the few survivors
      auto process = [&](const uint32_t (&v0)[32], const uint32_t (&v1)[32], int jb) {
        float g[4];
#pragma unroll
        for (int h = 0; h < 4; ++h) {
          const uint32_t *v = (h < 2) ? v0 : v1;
          const int o = (h & 1) * 16;
          float m = EXACT ? INFINITY : mrun[h];
#pragma unroll
          for (int k = 0; k < 16; k += 2) m = fminf(m, fminf(__uint_as_float(v[o + k]), __uint_as_float(v[o + k + 1])));
          g[h] = m;
          if (!EXACT) mrun[h] = m;
        }
        if (EXACT && fminf(fminf(g[0], g[1]), fminf(g[2], g[3])) < thr) {  // cold: a few pairs per source
#pragma unroll
          for (int h = 0; h < 4; ++h) {
            if (g[h] < thr) {
              const uint32_t *v = (h < 2) ? v0 : v1;
              const int o = (h & 1) * 16;
              uint32_t bits = 0;
#pragma unroll
              for (int k = 0; k < 16; ++k) bits |= (__uint_as_float(v[o + k]) < thr ? 1u : 0u) << k;
              while (bits) {
                const int j = jb + h * 16 + __ffs(bits) - 1;
                bits &= bits - 1;
                if (j < A.n_dst) {
                  const float dx = __fsub_rn(sx, A.dst[3 * j]), dy = __fsub_rn(sy, A.dst[3 * j + 1]), dz = __fsub_rn(sz, A.dst[3 * j + 2]);
                  const float d = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
                  if (d < smin) { smin = d; jbest = (unsigned int)j; }  // ascending j: the first index keeps a tie
                }
              }
            }
          }
        }
      };
      // software pipeline: the tcgen05.ld of the next chunk is in flight while the current one is reduced
      uint32_t va0[32], va1[32], vb0[32], vb1[32];
      const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16) + half * 128;
      mbar_wait(bar_acc_full + 8 * buf, (full_par >> buf) & 1);
      full_par ^= 1u << buf;
      tc_fence_after();
      tmem_ld32(lane_base + buf * 256, va0);
      tmem_ld32(lane_base + buf * 256 + 32, va1);
      tmem_ld_wait();
      for (int k = k0; k < k1; ++k) {
        const int jb = stage_at(sb, k) * 256 + half * 128;
        tmem_ld32(lane_base + buf * 256 + 64, vb0);
        tmem_ld32(lane_base + buf * 256 + 96, vb1);
        process(va0, va1, jb);
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive(bar_acc_free + 8 * buf);   // both chunks of this stage are in registers
        buf ^= 1;
        if (k + 1 < k1) {
          mbar_wait(bar_acc_full + 8 * buf, (full_par >> buf) & 1);
          full_par ^= 1u << buf;
          tc_fence_after();
          tmem_ld32(lane_base + buf * 256, va0);
          tmem_ld32(lane_base + buf * 256 + 32, va1);
        }
        process(vb0, vb1, jb + 64);
        tmem_ld_wait();
      }
      if (EXACT) {
        if (jbest != 0xffffffffu) atomicMin(A.best + s_row, ((unsigned long long)__float_as_uint(smin) << 32) | jbest);
      } else {
        const float m = fminf(fminf(mrun[0], mrun[1]), fminf(mrun[2], mrun[3]));
        // combine the two column halves, then one atomicMin per source
        if (half == 1) red[row] = m;
        asm volatile("bar.sync 1, 256;" ::: "memory");
        if (half == 0 && s_row < A.n_src) atomicMin(A.bound + s_row, __float_as_uint(fmaxf(fminf(m, red[row]) + A.norm2[s_row], 0.f)));
        asm volatile("bar.sync 1, 256;" ::: "memory");
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc(tmem, 512);
}

}  // namespace

size_t nnb_chamfer_tc_workspace_bytes(int n_src, int n_dst) {
  const size_t ns = (size_t)nnb_cdiv(n_src, 128) * 128, nd = (size_t)nnb_cdiv(n_dst, 256) * 256;
  const size_t nsub = (size_t)nnb_cdiv(nnb_cdiv(n_dst, CT_SUB), 256) * 256;
  const size_t n_sblocks = ns / 128, n_stages = nd / 256;
  // images | norm2 | sub-sampled images | stage boxes | stage lists + counts
  return ns * 64 + nd * 64 + ns * 4 + 256 + nsub * 32 + n_stages * 32 + (n_sblocks * n_stages + n_sblocks) * 4 + 256;
}

size_t nnb_chamfer_tc_norm2_offset(int n_src, int n_dst) {
  return (size_t)nnb_cdiv(n_src, 128) * 8192 + (size_t)nnb_cdiv(n_dst, 256) * CT_STAGE;
}

// bound (n_src uint32 = float bits, pre-filled with a large value) <- min over destinations; stats (10 x uint32): [0] = max |q-c|^2,
// [1] = max |q|^2 (float bits), [2] extent, [3] list gate, [5] counter, [6..9] centre sums.  workspace: nnb_chamfer_tc_workspace_bytes.
// Resets bound[] and best_init[] (to ~0) itself.  best != nullptr: also run the exact sweep, best[s] = (bits(dist) << 32 | first argmin j).
int nnb_chamfer_tc_bound(const float *src, int n_src, const float *dst, int n_dst, unsigned int *bound, unsigned int *stats, void *workspace,
                         unsigned long long *best_init, unsigned long long *best, int use_f16, cudaStream_t s) {
  static bool attr_set = false;
  if (!attr_set) {
    NNB_CUDA(cudaFuncSetAttribute(chamfer_tc_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CS_TOTAL));
    NNB_CUDA(cudaFuncSetAttribute(chamfer_tc_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CS_TOTAL));
    NNB_CUDA(cudaFuncSetAttribute(chamfer_tc_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CS_TOTAL));
    NNB_CUDA(cudaFuncSetAttribute(chamfer_tc_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CS_TOTAL));
    attr_set = true;
  }
  const int n_sblocks = nnb_cdiv(n_src, 128), n_stages = nnb_cdiv(n_dst, 256);
  uint8_t *a_img = static_cast<uint8_t *>(workspace);
  uint8_t *b_img = a_img + (size_t)n_sblocks * 8192;
  // the fp16 source images take 4 KB per block of the 8 KB the bf16 ones need: the exact sweep's copy lives in the other half
  uint8_t *a_img_x = a_img + (size_t)n_sblocks * 4096;
  float *norm2 = reinterpret_cast<float *>(b_img + (size_t)n_stages * CT_STAGE);
  NNB_CUDA(cudaMemsetAsync(stats, 0, 10 * sizeof(unsigned int), s));
  const int ns_pad = n_sblocks * 128, nd_pad = n_stages * 256;
  // stage lists pay off when there are enough stages to skip; small problems run the two plain sweeps
  const bool lists = use_f16 && n_stages >= 16;
  const int sub = lists ? CT_SUB : 1;
  const int n_sub_stages = lists ? nnb_cdiv(nnb_cdiv(n_dst, sub), 256) : 0, nsub_pad = n_sub_stages * 256;
  uint8_t *sub_img = reinterpret_cast<uint8_t *>(norm2) + (((size_t)ns_pad * 4 + 255) / 256) * 256;
  float *box = reinterpret_cast<float *>(sub_img + (size_t)nnb_cdiv(nnb_cdiv(n_dst, CT_SUB), 256) * 8192);
  int *list = reinterpret_cast<int *>(box + (size_t)n_stages * 8);
  int *list_cnt = list + (size_t)n_sblocks * n_stages;
  if (use_f16) {
    int cg = nnb_cdiv(n_dst, 256 * 8);
    if (cg > nnb_sm_count()) cg = nnb_sm_count();
    chamfer_tc_centre_kernel<<<cg, 256, 0, s>>>(dst, n_dst, stats);
    NNB_CHECK_LAUNCH();
    int eg = nnb_cdiv(n_src + n_dst, 256 * 8);
    if (eg > nnb_sm_count()) eg = nnb_sm_count();
    chamfer_tc_extent_kernel<<<eg, 256, 0, s>>>(src, n_src, dst, n_dst, stats);
    NNB_CHECK_LAUNCH();
    chamfer_tc_prep16_kernel<<<nnb_cdiv(ns_pad + nd_pad + nsub_pad, 128), 128, 0, s>>>(src, n_src, ns_pad, dst, n_dst, nd_pad, nsub_pad, sub, a_img,
                                                                                     a_img_x, b_img, sub_img, norm2, stats, bound, best_init);
    if (lists) {
      NNB_CHECK_LAUNCH();
      chamfer_tc_boxes_kernel<<<nnb_cdiv(n_stages, 4), 128, 0, s>>>(dst, n_dst, n_stages, box, stats);
    }
  } else {
    chamfer_tc_prep_kernel<<<nnb_cdiv(ns_pad + nd_pad, 128), 128, 0, s>>>(src, n_src, ns_pad, dst, n_dst, nd_pad, a_img, b_img, norm2, stats,
                                                                        bound, best_init);
  }
  NNB_CHECK_LAUNCH();
  // work items = source blocks x splits of the stage range, about `per_sm` per SM, at least `min_stages` stages each
  auto plan = [&](CtArgs &P, int stages, int per_sm, int min_stages) {
    int n_splits = nnb_cdiv(nnb_sm_count() * per_sm, n_sblocks);
    int max_splits = stages / min_stages;
    if (max_splits < 1) max_splits = 1;
    if (n_splits > max_splits) n_splits = max_splits;
    if (n_splits < 1) n_splits = 1;
    const int per = nnb_cdiv(stages, n_splits);
    P.n_splits = nnb_cdiv(stages, per);
    P.stages_per_split = per;
    const int n_items = n_sblocks * P.n_splits;
    return n_items < nnb_sm_count() ? n_items : nnb_sm_count();
  };
  CtArgs A{a_img, b_img, norm2, bound, n_src, n_sblocks, n_stages, 1, n_stages, src, dst, stats, best, n_dst, nullptr, nullptr, nullptr, 0};
  if (lists) {
    // (1) loose bound from every CT_SUB-th destination
    CtArgs S = A;
    S.b_img = sub_img; S.n_stages = n_sub_stages; S.gate = stats + 3; S.gate_skip = 1;
    const int gs = plan(S, n_sub_stages, 2, 4);
    chamfer_tc_kernel<false, true><<<gs, 320, CS_TOTAL, s>>>(S);
    NNB_CHECK_LAUNCH();
    // (2) stage lists
    chamfer_tc_filter_kernel<<<n_sblocks, 128, 0, s>>>(src, n_src, n_stages, box, bound, norm2, stats, list, list_cnt);
    NNB_CHECK_LAUNCH();
    A.list = list; A.list_cnt = list_cnt; A.gate = stats + 3;
  }
  const int grid = plan(A, n_stages, 4, CT_RING);
  if (use_f16) chamfer_tc_kernel<false, true><<<grid, 320, CS_TOTAL, s>>>(A);
  else chamfer_tc_kernel<false, false><<<grid, 320, CS_TOTAL, s>>>(A);
  NNB_CHECK_LAUNCH();
  if (best) {
    if (use_f16) {
      A.a_img = a_img_x;
      chamfer_tc_kernel<true, true><<<grid, 320, CS_TOTAL, s>>>(A);
    } else chamfer_tc_kernel<true, false><<<grid, 320, CS_TOTAL, s>>>(A);
    NNB_CHECK_LAUNCH();
  }
  return 0;
}
