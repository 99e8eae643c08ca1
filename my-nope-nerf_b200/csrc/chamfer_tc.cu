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

using namespace tc;

namespace {

constexpr int CT_RING = 8;
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
};

// EXACT = false: bound[s] = min_j a(s, j) + |s - c|^2.
// EXACT = true : same sweep (bit-identical a values: same operands, same MMA shape); every pair with
//                a < (bound - |s - c|^2) + margin is re-evaluated with the reference's arithmetic (direct differences,
//                each op rounded, sqrt_rn) and merged with atomicMin on (bits(dist) << 32 | j): first index of the minimum.
template <bool EXACT>
__global__ void __launch_bounds__(320, 1) chamfer_tc_kernel(const CtArgs A) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t sbase = smem_u32(smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float *red = reinterpret_cast<float *>(smem + CS_RED);
  const uint32_t bar_full = sbase + CS_BARS, bar_empty = bar_full + CT_RING * 8, bar_a = bar_empty + CT_RING * 8,
                 bar_acc_full = bar_a + 16, bar_acc_free = bar_acc_full + 16;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + CS_BARS + 32 * 8);
  if (threadIdx.x == 0) {
    for (int i = 0; i < CT_RING; ++i) { mbar_init(bar_full + 8 * i, 1); mbar_init(bar_empty + 8 * i, 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(bar_a + 8 * i, 1); mbar_init(bar_acc_full + 8 * i, 1); mbar_init(bar_acc_free + 8 * i, 256); }
    mbar_fence_init();
  }
  if (warp == 9) tmem_alloc(smem_u32(tmem_slot), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const int n_items = A.n_sblocks * A.n_splits;

  if (warp == 8) {
    if (elect_one()) {
      uint32_t slot = 0, phase = 0;
      int it = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
        const int sb = item / A.n_splits, sp = item % A.n_splits;
        const int st0 = sp * A.stages_per_split, st1 = min(A.n_stages, st0 + A.stages_per_split);
        for (int st = st0; st < st1; ++st) {
          mbar_wait(bar_empty + 8 * slot, phase ^ 1);
          if (st == st0) {
            // the A buffer of item it-2 is free: every MMA older than the slot just waited for has completed
            mbar_arrive_expect_tx(bar_a + 8 * (it & 1), 8192);
            bulk_g2s(sbase + CS_A + (it & 1) * 8192, A.a_img + (size_t)sb * 8192, 8192, bar_a + 8 * (it & 1));
          }
          mbar_arrive_expect_tx(bar_full + 8 * slot, CT_STAGE);
          bulk_g2s(sbase + CS_RING + slot * CT_STAGE, A.b_img + (size_t)st * CT_STAGE, CT_STAGE, bar_full + 8 * slot);
          if (++slot == CT_RING) { slot = 0; phase ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 9) {
    if (elect_one()) {
      uint32_t slot = 0, phase = 0, buf = 0, free_par = 0;
      uint32_t n_issued = 0;
      const uint32_t idesc = idesc_bf16(128, 256, 0, 0);
      int it = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
        const int sp = item % A.n_splits;
        const int st0 = sp * A.stages_per_split, st1 = min(A.n_stages, st0 + A.stages_per_split);
        mbar_wait(bar_a + 8 * (it & 1), (it >> 1) & 1);
        const uint32_t a0 = sbase + CS_A + (it & 1) * 8192;
        for (int st = st0; st < st1; ++st) {
          if (n_issued >= 2) { mbar_wait(bar_acc_free + 8 * buf, (free_par >> buf) & 1); free_par ^= 1u << buf; }
          mbar_wait(bar_full + 8 * slot, phase);
          tc_fence_after();
          const uint32_t b0 = sbase + CS_RING + slot * CT_STAGE;
          umma(tmem + buf * 256, desc_kmajor(a0, 512), desc_kmajor(b0, 512), idesc, 0u);
          umma(tmem + buf * 256, desc_kmajor(a0 + 256, 512), desc_kmajor(b0 + 256, 512), idesc, 1u);
          umma_commit(bar_empty + 8 * slot);
          umma_commit(bar_acc_full + 8 * buf);
          ++n_issued;
          buf ^= 1;
          if (++slot == CT_RING) { slot = 0; phase ^= 1; }
        }
      }
    }
    __syncwarp();
  } else {
    // 8 epilogue warps: quadrant = warp & 3 (rows), column half = warp >> 2
    const int q = warp & 3, half = warp >> 2, row = q * 32 + lane;
    uint32_t buf = 0, full_par = 0;   // bit b = parity of accumulator buffer b
    for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
      const int sb = item / A.n_splits, sp = item % A.n_splits;
      const int st0 = sp * A.stages_per_split, st1 = min(A.n_stages, st0 + A.stages_per_split);
      float mrun[4] = {INFINITY, INFINITY, INFINITY, INFINITY};
      const int s_row = sb * 128 + row;
      float thr = -INFINITY, sx = 0.f, sy = 0.f, sz = 0.f, smin = INFINITY;
      unsigned int jbest = 0xffffffffu;
      if (EXACT && s_row < A.n_src) {
        const float b = __uint_as_float(A.bound[s_row]), nn = A.norm2[s_row];
        thr = (b - nn) + 8e-6f * (nn + 2.f * __uint_as_float(A.stats[0]) + b);
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
      for (int st = st0; st < st1; ++st) {
        const int jb = st * 256 + half * 128;
        tmem_ld32(lane_base + buf * 256 + 64, vb0);
        tmem_ld32(lane_base + buf * 256 + 96, vb1);
        process(va0, va1, jb);
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive(bar_acc_free + 8 * buf);   // both chunks of this stage are in registers
        buf ^= 1;
        if (st + 1 < st1) {
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
  return ns * 64 + nd * 64 + ns * 4 + 256;
}

size_t nnb_chamfer_tc_norm2_offset(int n_src, int n_dst) {
  return (size_t)nnb_cdiv(n_src, 128) * 8192 + (size_t)nnb_cdiv(n_dst, 256) * CT_STAGE;
}

// bound (n_src uint32 = float bits, pre-filled with a large value) <- min over destinations; stats[0] = max |q-c|^2,
// stats[1] = max |q|^2 (float bits).  workspace: nnb_chamfer_tc_workspace_bytes.
// Resets bound[] and best_init[] (to ~0) itself.  best != nullptr: also run the exact sweep, best[s] = (bits(dist) << 32 | first argmin j).
int nnb_chamfer_tc_bound(const float *src, int n_src, const float *dst, int n_dst, unsigned int *bound, unsigned int *stats, void *workspace,
                         unsigned long long *best_init, unsigned long long *best, cudaStream_t s) {
  static bool attr_set = false;
  if (!attr_set) {
    NNB_CUDA(cudaFuncSetAttribute(chamfer_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CS_TOTAL));
    NNB_CUDA(cudaFuncSetAttribute(chamfer_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CS_TOTAL));
    attr_set = true;
  }
  const int n_sblocks = nnb_cdiv(n_src, 128), n_stages = nnb_cdiv(n_dst, 256);
  uint8_t *a_img = static_cast<uint8_t *>(workspace);
  uint8_t *b_img = a_img + (size_t)n_sblocks * 8192;
  float *norm2 = reinterpret_cast<float *>(b_img + (size_t)n_stages * CT_STAGE);
  NNB_CUDA(cudaMemsetAsync(stats, 0, 2 * sizeof(unsigned int), s));
  const int ns_pad = n_sblocks * 128, nd_pad = n_stages * 256;
  chamfer_tc_prep_kernel<<<nnb_cdiv(ns_pad + nd_pad, 128), 128, 0, s>>>(src, n_src, ns_pad, dst, n_dst, nd_pad, a_img, b_img, norm2, stats,
                                                                      bound, best_init);
  NNB_CHECK_LAUNCH();
  // work items = source blocks x destination splits, about 4 per SM, at least CT_RING stages each
  int n_splits = nnb_cdiv(nnb_sm_count() * 4, n_sblocks);
  int max_splits = n_stages / CT_RING;
  if (max_splits < 1) max_splits = 1;
  if (n_splits > max_splits) n_splits = max_splits;
  if (n_splits < 1) n_splits = 1;
  const int per = nnb_cdiv(n_stages, n_splits);
  n_splits = nnb_cdiv(n_stages, per);
  CtArgs A{a_img, b_img, norm2, bound, n_src, n_sblocks, n_stages, n_splits, per, src, dst, stats, best, n_dst};
  const int n_items = n_sblocks * n_splits;
  const int grid = n_items < nnb_sm_count() ? n_items : nnb_sm_count();
  chamfer_tc_kernel<false><<<grid, 320, CS_TOTAL, s>>>(A);
  NNB_CHECK_LAUNCH();
  if (best) {
    chamfer_tc_kernel<true><<<grid, 320, CS_TOTAL, s>>>(A);
    NNB_CHECK_LAUNCH();
  }
  return 0;
}
