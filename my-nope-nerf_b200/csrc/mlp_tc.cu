// Stages 2+3 on the tensor cores (NNB_PREC_BF16): stratified sampling + positional encoding + the 12-layer NeRF MLP
// + alpha compositing in ONE persistent tcgen05/TMEM kernel.  model/rendering.py:89-126, model/official_nerf.py:60-119.
//
// Work unit: a tile = 128 samples = one ray (S = 128).  A CTA owns one SM and walks over ray PAIRS: the two tiles of a
// pair advance through the layers in lockstep so that every weight stage fetched from L2 feeds two MMAs.
//   warps 0-3 / 4-7 : "row owners" of tile 0 / 1 (thread = sample): prologue (sampling + encoding -> bf16 operand images
//                     in shared memory), per-layer epilogue (TMEM -> relu -> bf16 -> next layer's A operand, in place),
//                     heads (density, rgb) and the compositing scan across the 128 threads of the ray
//   warp 8          : weight producer: 1-D bulk TMA of pre-packed [N x 16] bf16 stage images into a ring
//   warp 9          : MMA issuer: tcgen05.mma 128 x N x 16, fp32 accumulators in TMEM (2 tiles x 256 columns)
// Biases ride in the MMA: operand image `aux` holds the view-direction encoding (27) followed by two constant-one
// columns; every layer ends with the aux k-step whose weight rows carry bias_hi / bias_lo (bf16 split, ~fp32 exact).
// The skip and view-direction concatenations become extra k-steps on the `enc` / `aux` images (no torch.cat).
#include "nnb_common.cuh"
#include "sampling.cuh"
#include "tc_common.cuh"
#include <vector>
#include <mutex>

using namespace tc;

namespace {

// parameter indices (reference state_dict order)
enum { P_L0 = 0, P_DEN = 16, P_FEAT = 18, P_RGB0 = 20, P_RGB = 22 };

constexpr int TC_MAX_STAGES = 512;
constexpr int TC_RING = 5;               // weight ring slots
constexpr uint32_t TC_SLOT = 8192;       // bytes per slot (256 rows x 16 k x bf16)

enum ASrc : uint8_t { A_ACT = 0, A_ENC = 1, A_AUX = 2 };

// One k-step (K = 16) of one layer.
struct TcStage {
  uint32_t w_off;   // byte offset of the [n x 16] weight image in the packed buffer
  uint16_t n;       // MMA N (output features)
  uint8_t a_src;    // ASrc
  uint8_t a_k;      // k-step inside that operand image
  uint8_t first;    // first k-step of its layer (accumulator overwrite; wait for the A operand)
  uint8_t last;     // last k-step of its layer (commit -> epilogue)
  uint8_t layer;
  uint8_t pad;
};
// How the pack kernel fills a stage image: value(n, kk) = W[n*ld + col0 + kk] for col0 + kk < col_end (else 0), or the
// transposed read W[(col0 + kk)*ld + row0 + n]; kk == 11 / 12 carry bias_hi / bias_lo when bias >= 0.
struct TcPack {
  int16_t wparam;   // index into the 24 parameter pointers, -1 = none
  int16_t bias;     // parameter index of the bias, -1 = none
  int32_t ld;
  int32_t col0, col_end;
  int32_t row0;
  int32_t transposed;
};

__constant__ TcStage c_fwd_stages[TC_MAX_STAGES];
__constant__ TcPack c_fwd_pack[TC_MAX_STAGES];

struct Schedule {
  std::vector<TcStage> st;
  std::vector<TcPack> pk;
  uint32_t bytes = 0;
  void add(int layer, int n, ASrc src, int ak, bool first, bool last, TcPack p) {
    TcStage s{};
    s.w_off = bytes; s.n = (uint16_t)n; s.a_src = src; s.a_k = (uint8_t)ak; s.first = first; s.last = last; s.layer = (uint8_t)layer;
    st.push_back(s);
    pk.push_back(p);
    bytes += (uint32_t)n * 32u;
  }
};

// forward schedule: layers 0..7 trunk, 8 = fc_feature, 9 = rgb_layers.0
Schedule build_fwd_schedule() {
  Schedule S;
  for (int L = 0; L < 8; ++L) {
    const int wp = P_L0 + 2 * L, ld = (L == 0) ? 63 : (L == 4 ? 319 : 256);
    bool first = true;
    if (L != 0)
      for (int j = 0; j < 16; ++j) { S.add(L, 256, A_ACT, j, first, false, TcPack{(int16_t)wp, -1, ld, 16 * j, 256, 0, 0}); first = false; }
    if (L == 0 || L == 4) {
      const int base = (L == 0) ? 0 : 256;
      for (int j = 0; j < 4; ++j) { S.add(L, 256, A_ENC, j, first, false, TcPack{(int16_t)wp, -1, ld, base + 16 * j, base + 63, 0, 0}); first = false; }
    }
    S.add(L, 256, A_AUX, 1, false, true, TcPack{-1, (int16_t)(wp + 1), 0, 0, 0, 0, 0});
  }
  for (int j = 0; j < 16; ++j) S.add(8, 256, A_ACT, j, j == 0, false, TcPack{P_FEAT, -1, 256, 16 * j, 256, 0, 0});
  S.add(8, 256, A_AUX, 1, false, true, TcPack{-1, P_FEAT + 1, 0, 0, 0, 0, 0});
  for (int j = 0; j < 16; ++j) S.add(9, 128, A_ACT, j, j == 0, false, TcPack{P_RGB0, -1, 283, 16 * j, 256, 0, 0});
  S.add(9, 128, A_AUX, 0, false, false, TcPack{P_RGB0, -1, 283, 256, 283, 0, 0});
  S.add(9, 128, A_AUX, 1, false, true, TcPack{P_RGB0, P_RGB0 + 1, 283, 272, 283, 0, 0});
  return S;
}

const Schedule &fwd_schedule() {
  static Schedule s = build_fwd_schedule();
  return s;
}

struct PackPtrs { const float *p[NNB_NUM_PARAMS]; };

// grid = stages, block = 256 threads: thread n fills row n of the [n x 16] image
__global__ void tc_pack_kernel(PackPtrs P, uint8_t *out, int n_stages, int which) {
  const int i = blockIdx.x;
  if (i >= n_stages) return;
  const TcStage st = c_fwd_stages[i];
  const TcPack pk = c_fwd_pack[i];
  (void)which;
  for (int n = threadIdx.x; n < st.n; n += blockDim.x) {
    __nv_bfloat16 row[16];
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      float v = 0.f;
      if (pk.wparam >= 0 && pk.col0 + kk < pk.col_end) {
        const float *W = P.p[pk.wparam];
        v = pk.transposed ? W[(size_t)(pk.col0 + kk) * pk.ld + pk.row0 + n] : W[(size_t)n * pk.ld + pk.col0 + kk];
      }
      row[kk] = __float2bfloat16_rn(v);
    }
    if (pk.bias >= 0) {
      const float b = P.p[pk.bias][n];
      const __nv_bfloat16 hi = __float2bfloat16_rn(b);
      row[11] = hi;
      row[12] = __float2bfloat16_rn(b - __bfloat162float(hi));
    }
    uint8_t *dst = out + st.w_off;
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) *reinterpret_cast<__nv_bfloat16 *>(dst + img_off(n, kk, 16)) = row[kk];
  }
}

}  // namespace

// =====================================================================================================================
// forward kernel
// =====================================================================================================================
namespace {

constexpr uint32_t SM_ACT = 0;                        // 2 x 65536  (128 rows x 256 cols bf16)
constexpr uint32_t SM_ENC = 131072;                   // 2 x 16384  (128 x 64)
constexpr uint32_t SM_AUX = SM_ENC + 32768;           // 2 x 8192   (128 x 32)
constexpr uint32_t SM_RING = SM_AUX + 16384;          // TC_RING x 8192
constexpr uint32_t SM_SMALL = SM_RING + TC_RING * TC_SLOT;  // fp32: w_den[256] w_rgb[384] b_den b_rgb[3] (+pad) = 648 floats
constexpr uint32_t SM_SCAN = SM_SMALL + 648 * 4;      // 2 tiles x 4 warps x 8 floats
constexpr uint32_t SM_BARS = SM_SCAN + 2 * 4 * 8 * 4; // mbarriers
constexpr uint32_t SM_TOTAL = SM_BARS + 16 * 8 + 16;
static_assert(SM_TOTAL <= 232448, "shared memory budget");

constexpr uint32_t GS_ACT = 4096, GS_ENC = 1024, GS_AUX = 512;

struct FwdArgs {
  const uint8_t *wimg;
  const float *origin, *dirs, *noise;
  const float *w_den, *b_den, *w_rgb, *b_rgb;
  float *rgb, *dist, *weights, *alpha, *z_vals;
  int n_rays, S, n_stages;
  float near_, far_;
  int softplus, white, use_dir;
};

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory"); }

__device__ __forceinline__ float warp_incl_prod(float v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const float t = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v *= t;
  }
  return v;
}

// Sampling + positional encoding of one sample -> its rows of the enc (64 cols) and aux (32 cols) operand images.
// Column order follows encode_position (model/official_nerf.py:99-119): [x, sin(2^l x), cos(2^l x)]_l, 3 wide each.
__device__ __forceinline__ void write_encodings(uint32_t enc_row, uint32_t aux_row, const float p[3], const float v[3]) {
  float e[64];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    e[c] = p[c];
    float sn, cs;
    sincosf(p[c], &sn, &cs);
#pragma unroll
    for (int l = 0; l < 10; ++l) {
      e[3 + 6 * l + c] = sn;
      e[3 + 6 * l + 3 + c] = cs;
      const float s2 = 2.f * sn * cs, c2 = 1.f - 2.f * sn * sn;  // double angle
      sn = s2; cs = c2;
    }
  }
  e[63] = 0.f;
#pragma unroll
  for (int g = 0; g < 8; ++g)
    st_shared_v4(enc_row + g * 128, pack_bf16(e[8 * g], e[8 * g + 1]), pack_bf16(e[8 * g + 2], e[8 * g + 3]),
                 pack_bf16(e[8 * g + 4], e[8 * g + 5]), pack_bf16(e[8 * g + 6], e[8 * g + 7]));
  float a[32];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    a[c] = v[c];
    float sn, cs;
    sincosf(v[c], &sn, &cs);
#pragma unroll
    for (int l = 0; l < 4; ++l) {
      a[3 + 6 * l + c] = sn;
      a[3 + 6 * l + 3 + c] = cs;
      const float s2 = 2.f * sn * cs, c2 = 1.f - 2.f * sn * sn;
      sn = s2; cs = c2;
    }
  }
  a[27] = 1.f; a[28] = 1.f; a[29] = 0.f; a[30] = 0.f; a[31] = 0.f;
#pragma unroll
  for (int g = 0; g < 4; ++g)
    st_shared_v4(aux_row + g * 128, pack_bf16(a[8 * g], a[8 * g + 1]), pack_bf16(a[8 * g + 2], a[8 * g + 3]),
                 pack_bf16(a[8 * g + 4], a[8 * g + 5]), pack_bf16(a[8 * g + 6], a[8 * g + 7]));
}

__global__ void __launch_bounds__(320, 1) tc_fwd_kernel(const FwdArgs A) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t sbase = smem_u32(smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float *sm_small = reinterpret_cast<float *>(smem + SM_SMALL);
  float *sm_scan = reinterpret_cast<float *>(smem + SM_SCAN);
  const uint32_t bar_full = sbase + SM_BARS, bar_empty = bar_full + TC_RING * 8, bar_acc = bar_empty + TC_RING * 8,
                 bar_act = bar_acc + 8;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + SM_BARS + 16 * 8);

  if (threadIdx.x == 0) {
    for (int i = 0; i < TC_RING; ++i) { mbar_init(bar_full + 8 * i, 1); mbar_init(bar_empty + 8 * i, 1); }
    mbar_init(bar_acc, 1);
    mbar_init(bar_act, 256);
    mbar_fence_init();
  }
  for (int i = threadIdx.x; i < 256; i += blockDim.x) sm_small[i] = A.w_den[i];
  for (int i = threadIdx.x; i < 384; i += blockDim.x) sm_small[256 + i] = A.w_rgb[i];
  if (threadIdx.x == 0) { sm_small[640] = A.b_den[0]; sm_small[641] = A.b_rgb[0]; sm_small[642] = A.b_rgb[1]; sm_small[643] = A.b_rgb[2]; }
  if (warp == 9) tmem_alloc(smem_u32(tmem_slot), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const int n_pairs = (A.n_rays + 1) >> 1;

  if (warp == 8) {
    // ------------------------------------------------ weight producer
    if (elect_one()) {
      uint32_t slot = 0, phase = 0;
      for (int pair = blockIdx.x; pair < n_pairs; pair += gridDim.x)
        for (int i = 0; i < A.n_stages; ++i) {
          const TcStage st = c_fwd_stages[i];
          mbar_wait(bar_empty + 8 * slot, phase ^ 1);
          const uint32_t bytes = (uint32_t)st.n * 32u;
          mbar_arrive_expect_tx(bar_full + 8 * slot, bytes);
          bulk_g2s(sbase + SM_RING + slot * TC_SLOT, A.wimg + st.w_off, bytes, bar_full + 8 * slot);
          if (++slot == TC_RING) { slot = 0; phase ^= 1; }
        }
    }
    __syncwarp();
  } else if (warp == 9) {
    // ------------------------------------------------ MMA issuer
    if (elect_one()) {
      uint32_t slot = 0, phase = 0, act_par = 0;
      for (int pair = blockIdx.x; pair < n_pairs; pair += gridDim.x)
        for (int i = 0; i < A.n_stages; ++i) {
          const TcStage st = c_fwd_stages[i];
          if (st.first) { mbar_wait(bar_act, act_par); act_par ^= 1; }
          mbar_wait(bar_full + 8 * slot, phase);
          tc_fence_after();
          const uint32_t a_off = (st.a_src == A_ACT) ? SM_ACT : (st.a_src == A_ENC ? SM_ENC : SM_AUX);
          const uint32_t a_sz = (st.a_src == A_ACT) ? 65536u : (st.a_src == A_ENC ? 16384u : 8192u);
          const uint32_t a_gs = (st.a_src == A_ACT) ? GS_ACT : (st.a_src == A_ENC ? GS_ENC : GS_AUX);
          const uint64_t bd = desc_kmajor(sbase + SM_RING + slot * TC_SLOT, 256);
          const uint32_t idesc = idesc_bf16(128, st.n, 0, 0);
#pragma unroll
          for (int t = 0; t < 2; ++t) {
            const uint64_t ad = desc_kmajor(sbase + a_off + t * a_sz + (uint32_t)st.a_k * 256u, a_gs);
            umma(tmem + t * 256, ad, bd, idesc, st.first ? 0u : 1u);
          }
          umma_commit(bar_empty + 8 * slot);
          if (st.last) umma_commit(bar_acc);
          if (++slot == TC_RING) { slot = 0; phase ^= 1; }
        }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------ row owners: prologue, epilogues, heads, compositing
    const int t = warp >> 2, q = warp & 3, row = q * 32 + lane;
    const uint32_t act_row = sbase + SM_ACT + t * 65536u + (row >> 3) * GS_ACT + (row & 7) * 16;
    const uint32_t enc_row = sbase + SM_ENC + t * 16384u + (row >> 3) * GS_ENC + (row & 7) * 16;
    const uint32_t aux_row = sbase + SM_AUX + t * 8192u + (row >> 3) * GS_AUX + (row & 7) * 16;
    const uint32_t tacc = tmem + ((uint32_t)(q * 32) << 16) + t * 256;
    float *scan = sm_scan + t * 32;
    uint32_t acc_par = 0;
    for (int pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
      const int ray_raw = 2 * pair + t;
      const bool valid = ray_raw < A.n_rays;
      const int ray = valid ? ray_raw : A.n_rays - 1;
      // ---- prologue: stratified depth, sample position, encodings
      const float z = z_sample(row, A.S, A.near_, A.far_, A.noise ? A.noise + (size_t)ray * A.S : nullptr);
      {
        float p[3], v[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          const float d = A.dirs[3 * ray + c];
          p[c] = __fadd_rn(A.origin[c], __fmul_rn(d, z));
          v[c] = A.use_dir ? -d : 1.0f;
        }
        write_encodings(enc_row, aux_row, p, v);
      }
      fence_async_smem();
      tc_fence_before();
      mbar_arrive(bar_act);

      float alpha = 0.f, r_s = 0.f, g_s = 0.f, b_s = 0.f;
      for (int L = 0; L < 10; ++L) {
        mbar_wait(bar_acc, acc_par);
        acc_par ^= 1;
        tc_fence_after();
        if (L < 9) {
          float dsum = 0.f;
#pragma unroll 1
          for (int c0 = 0; c0 < 256; c0 += 64) {
            uint32_t v0[32], v1[32];
            tmem_ld32(tacc + c0, v0);
            tmem_ld32(tacc + c0 + 32, v1);
            tmem_ld_wait();
            if (L == 7) {
              const float4 *wd = reinterpret_cast<const float4 *>(sm_small + c0);
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const float4 w = wd[j];
                dsum = fmaf(fmaxf(__uint_as_float(v0[4 * j]), 0.f), w.x, dsum);
                dsum = fmaf(fmaxf(__uint_as_float(v0[4 * j + 1]), 0.f), w.y, dsum);
                dsum = fmaf(fmaxf(__uint_as_float(v0[4 * j + 2]), 0.f), w.z, dsum);
                dsum = fmaf(fmaxf(__uint_as_float(v0[4 * j + 3]), 0.f), w.w, dsum);
              }
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const float4 w = wd[8 + j];
                dsum = fmaf(fmaxf(__uint_as_float(v1[4 * j]), 0.f), w.x, dsum);
                dsum = fmaf(fmaxf(__uint_as_float(v1[4 * j + 1]), 0.f), w.y, dsum);
                dsum = fmaf(fmaxf(__uint_as_float(v1[4 * j + 2]), 0.f), w.z, dsum);
                dsum = fmaf(fmaxf(__uint_as_float(v1[4 * j + 3]), 0.f), w.w, dsum);
              }
            }
            const uint32_t dst = act_row + (c0 >> 3) * 128;
            if (L != 8) {
#pragma unroll
              for (int g = 0; g < 4; ++g)
                st_shared_v4(dst + g * 128, pack_bf16_relu(__uint_as_float(v0[8 * g]), __uint_as_float(v0[8 * g + 1])),
                             pack_bf16_relu(__uint_as_float(v0[8 * g + 2]), __uint_as_float(v0[8 * g + 3])),
                             pack_bf16_relu(__uint_as_float(v0[8 * g + 4]), __uint_as_float(v0[8 * g + 5])),
                             pack_bf16_relu(__uint_as_float(v0[8 * g + 6]), __uint_as_float(v0[8 * g + 7])));
#pragma unroll
              for (int g = 0; g < 4; ++g)
                st_shared_v4(dst + (4 + g) * 128, pack_bf16_relu(__uint_as_float(v1[8 * g]), __uint_as_float(v1[8 * g + 1])),
                             pack_bf16_relu(__uint_as_float(v1[8 * g + 2]), __uint_as_float(v1[8 * g + 3])),
                             pack_bf16_relu(__uint_as_float(v1[8 * g + 4]), __uint_as_float(v1[8 * g + 5])),
                             pack_bf16_relu(__uint_as_float(v1[8 * g + 6]), __uint_as_float(v1[8 * g + 7])));
            } else {
#pragma unroll
              for (int g = 0; g < 4; ++g)
                st_shared_v4(dst + g * 128, pack_bf16(__uint_as_float(v0[8 * g]), __uint_as_float(v0[8 * g + 1])),
                             pack_bf16(__uint_as_float(v0[8 * g + 2]), __uint_as_float(v0[8 * g + 3])),
                             pack_bf16(__uint_as_float(v0[8 * g + 4]), __uint_as_float(v0[8 * g + 5])),
                             pack_bf16(__uint_as_float(v0[8 * g + 6]), __uint_as_float(v0[8 * g + 7])));
#pragma unroll
              for (int g = 0; g < 4; ++g)
                st_shared_v4(dst + (4 + g) * 128, pack_bf16(__uint_as_float(v1[8 * g]), __uint_as_float(v1[8 * g + 1])),
                             pack_bf16(__uint_as_float(v1[8 * g + 2]), __uint_as_float(v1[8 * g + 3])),
                             pack_bf16(__uint_as_float(v1[8 * g + 4]), __uint_as_float(v1[8 * g + 5])),
                             pack_bf16(__uint_as_float(v1[8 * g + 6]), __uint_as_float(v1[8 * g + 7])));
            }
          }
          if (L == 7) {
            // density head (model/official_nerf.py:66,77-83): softplus(beta 1, threshold 20) or relu, then 1 - exp(-d)
            const float raw = dsum + sm_small[640];
            const float d = A.softplus ? (raw > 20.f ? raw : log1pf(expf(raw))) : fmaxf(raw, 0.f);
            alpha = 1.f - expf(-d);
          }
          fence_async_smem();
          tc_fence_before();
          mbar_arrive(bar_act);
        } else {
          // rgb head: relu(rgb_layers.0) (128 wide, TMEM cols 0..127) . fc_rgb^T + b -> sigmoid
          float s0 = sm_small[641], s1 = sm_small[642], s2 = sm_small[643];
#pragma unroll 1
          for (int c0 = 0; c0 < 128; c0 += 32) {
            uint32_t v0[32];
            tmem_ld32(tacc + c0, v0);
            tmem_ld_wait();
            const float *w0 = sm_small + 256 + c0, *w1 = w0 + 128, *w2 = w0 + 256;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const float h = fmaxf(__uint_as_float(v0[j]), 0.f);
              s0 = fmaf(h, w0[j], s0);
              s1 = fmaf(h, w1[j], s1);
              s2 = fmaf(h, w2[j], s2);
            }
          }
          r_s = 1.f / (1.f + expf(-s0));
          g_s = 1.f / (1.f + expf(-s1));
          b_s = 1.f / (1.f + expf(-s2));
        }
      }
      // ---- compositing over the ray (model/rendering.py:124-126): T = exclusive cumprod(1 - alpha + 1e-6)
      const float beta = 1.f - alpha + 1e-6f;
      const float inc = warp_incl_prod(beta, lane);
      if (lane == 31) scan[q] = inc;
      named_bar_sync(1 + t, 128);
      float T = __shfl_up_sync(0xffffffffu, inc, 1);
      if (lane == 0) T = 1.f;
      for (int k = 0; k < q; ++k) T *= scan[k];
      const float w = alpha * T;
      float s_r = warp_sum(w * r_s), s_g = warp_sum(w * g_s), s_b = warp_sum(w * b_s), s_d = warp_sum(w * z), s_w = warp_sum(w);
      if (lane == 0) { scan[4 + q] = s_r; scan[8 + q] = s_g; scan[12 + q] = s_b; scan[16 + q] = s_d; scan[20 + q] = s_w; }
      named_bar_sync(1 + t, 128);
      if (valid) {
        if (row == 0) {
          float o_r = scan[4] + scan[5] + scan[6] + scan[7], o_g = scan[8] + scan[9] + scan[10] + scan[11];
          float o_b = scan[12] + scan[13] + scan[14] + scan[15], o_d = scan[16] + scan[17] + scan[18] + scan[19];
          if (A.white) { const float ws = scan[20] + scan[21] + scan[22] + scan[23]; o_r += 1.f - ws; o_g += 1.f - ws; o_b += 1.f - ws; }
          A.rgb[3 * ray] = o_r; A.rgb[3 * ray + 1] = o_g; A.rgb[3 * ray + 2] = o_b;
          A.dist[ray] = o_d;
        }
        const size_t m = (size_t)ray * A.S + row;
        if (A.weights) A.weights[m] = w;
        if (A.alpha) A.alpha[m] = alpha;
        if (A.z_vals) A.z_vals[m] = z;
      }
      named_bar_sync(1 + t, 128);  // scan scratch is reused by the next pair
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc(tmem, 512);
}

}  // namespace

// =====================================================================================================================
// host side
// =====================================================================================================================
namespace {

std::once_flag g_tables_once;
int g_tables_err = 0;

void upload_tables() {
  const Schedule &S = fwd_schedule();
  if ((int)S.st.size() > TC_MAX_STAGES) { g_tables_err = NNB_ERR_UNSUPPORTED; return; }
  cudaError_t e = cudaMemcpyToSymbol(c_fwd_stages, S.st.data(), S.st.size() * sizeof(TcStage));
  if (e == cudaSuccess) e = cudaMemcpyToSymbol(c_fwd_pack, S.pk.data(), S.pk.size() * sizeof(TcPack));
  if (e == cudaSuccess) e = cudaFuncSetAttribute(tc_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SM_TOTAL);
  g_tables_err = (int)e;
}

int ensure_tables() {
  std::call_once(g_tables_once, upload_tables);
  return g_tables_err;
}

}  // namespace

size_t nnb_tc_workspace_bytes(const nnb_render_cfg *, int) { return 16; }

int nnb_tc_render_fwd(const nnb_render_cfg *cfg, const float *const *params, const void *packed, const float *origin,
                      const float *dirs, const float *noise, float *rgb, float *dist, float *weights, float *alpha, float *z_vals,
                      void *ws, size_t ws_bytes, int save, cudaStream_t s) {
  (void)ws; (void)ws_bytes;
  if (!packed || !params || !origin || !dirs || !rgb || !dist) return NNB_ERR_BAD_ARG;
  if (cfg->n_samples != 128 || cfg->dist_alpha) return NNB_ERR_UNSUPPORTED;  // tile = one ray of 128 samples
  if (save) return NNB_ERR_UNSUPPORTED;
  if (cfg->n_rays <= 0) return 0;
  int e = ensure_tables();
  if (e) return e;
  FwdArgs A{};
  A.wimg = static_cast<const uint8_t *>(packed);
  A.origin = origin; A.dirs = dirs; A.noise = noise;
  A.w_den = params[P_DEN]; A.b_den = params[P_DEN + 1]; A.w_rgb = params[P_RGB]; A.b_rgb = params[P_RGB + 1];
  A.rgb = rgb; A.dist = dist; A.weights = weights; A.alpha = alpha; A.z_vals = z_vals;
  A.n_rays = cfg->n_rays; A.S = cfg->n_samples; A.n_stages = (int)fwd_schedule().st.size();
  A.near_ = cfg->near_; A.far_ = cfg->far_;
  A.softplus = cfg->occ_softplus; A.white = cfg->white_bkgd; A.use_dir = cfg->use_dir;
  const int n_pairs = (cfg->n_rays + 1) / 2;
  const int grid = n_pairs < nnb_sm_count() ? n_pairs : nnb_sm_count();
  tc_fwd_kernel<<<grid, 320, SM_TOTAL, s>>>(A);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_tc_render_bwd(const nnb_render_cfg *, const float *const *, const void *, const float *, const float *, const float *,
                      const float *, const float *, const float *, const float *, const float *, const float *, float *const *,
                      float *, float *, void *, size_t, cudaStream_t) {
  return NNB_ERR_UNSUPPORTED;
}

extern "C" {
size_t nnb_pack_weights_bytes(void) { return (size_t)fwd_schedule().bytes; }

int nnb_pack_weights(const float *const *params_host, void *packed, void *stream) {
  if (!params_host || !packed) return NNB_ERR_BAD_ARG;
  int e = ensure_tables();
  if (e) return e;
  PackPtrs P;
  for (int i = 0; i < NNB_NUM_PARAMS; ++i) P.p[i] = params_host[i];
  const int n = (int)fwd_schedule().st.size();
  tc_pack_kernel<<<n, 256, 0, nnb_stream(stream)>>>(P, static_cast<uint8_t *>(packed), n, 0);
  NNB_CHECK_LAUNCH();
  return 0;
}
}
