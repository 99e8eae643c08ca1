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
constexpr int TC_RING = 3;               // forward weight ring slots
constexpr uint32_t TC_SLOT = 16384;      // bytes per slot: one group = up to 2 k-steps of [256 rows x 16 k] bf16
constexpr int TC_MAX_GROUPS = 192;

enum ASrc : uint8_t { A_ACT = 0, A_ENC = 1, A_AUX = 2, A_ACT128 = 3, A_DSMALL = 4 };
enum Epi : uint8_t { EPI_DIR = 0, EPI_FEAT = 1, EPI_MASK = 2, EPI_ENC = 3 };

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
  int32_t n_valid;  // rows n >= n_valid are zero (0 = all rows valid)
  // CTA-pair layout (forward): gnk > 0 -> this stage is k-step gk of a group of gnk stages starting at w_off; the group is stored
  // as [rows 0..n/2 of every k-step | rows n/2..n of every k-step] so that each CTA of a pair fetches its half with ONE bulk copy
  int16_t gk, gnk;
};

__constant__ TcStage c_fwd_stages[TC_MAX_STAGES];
__constant__ TcPack c_fwd_pack[TC_MAX_STAGES];
__constant__ TcStage c_bwd_stages[TC_MAX_STAGES];
// Issue groups: consecutive k-steps of one layer on one A image (<= 2), one bulk copy + one barrier round trip each.
//   x = byte offset of the weight images   y = A offset in smem (18 bits) | nk << 18 | first << 20 | last << 21
//   z = n | a_gs << 16                        w = A stride between the two tiles
__constant__ uint4 c_fwd_groups[TC_MAX_GROUPS];
__constant__ uint4 c_bwd_groups[TC_MAX_GROUPS];
__constant__ TcPack c_bwd_pack[TC_MAX_STAGES];

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

// backward (data-gradient) schedule: B operand = W^T stage images; `layer` = Epi kind, `pad` = saved-activation slot
Schedule build_bwd_schedule() {
  Schedule S;
  auto add_step = [&](int n, ASrc src, int ksteps, int wparam, int ld, int row0, int n_valid, int epi, int slot, bool den_extra) {
    const size_t first = S.st.size();
    for (int j = 0; j < ksteps; ++j)
      S.add(epi, n, src, j, j == 0, (j == ksteps - 1) && !den_extra, TcPack{(int16_t)wparam, -1, ld, 16 * j, 16 * ksteps, row0, 1, n_valid});
    if (den_extra) S.add(epi, n, A_DSMALL, 0, false, true, TcPack{P_DEN, -1, 256, -3, 1, 0, 1, 0});
    for (size_t i = first; i < S.st.size(); ++i) S.st[i].pad = (uint8_t)slot;
  };
  add_step(32, A_ACT128, 8, P_RGB0, 283, 256, 27, EPI_DIR, 0, false);
  add_step(256, A_ACT128, 8, P_RGB0, 283, 0, 0, EPI_FEAT, 0, false);
  add_step(256, A_ACT, 16, P_FEAT, 256, 0, 0, EPI_MASK, 7, true);
  for (int L = 7; L >= 5; --L) add_step(256, A_ACT, 16, P_L0 + 2 * L, 256, 0, 0, EPI_MASK, L - 1, false);
  add_step(64, A_ACT, 16, P_L0 + 8, 319, 256, 63, EPI_ENC, 0, false);
  add_step(256, A_ACT, 16, P_L0 + 8, 319, 0, 0, EPI_MASK, 3, false);
  for (int L = 3; L >= 1; --L) add_step(256, A_ACT, 16, P_L0 + 2 * L, 256, 0, 0, EPI_MASK, L - 1, false);
  add_step(64, A_ACT, 16, P_L0, 63, 0, 63, EPI_ENC, 1, false);
  return S;
}

Schedule build_bwd_schedule_grouped();
const Schedule &bwd_schedule() {
  static Schedule s = build_bwd_schedule_grouped();
  return s;
}

Schedule build_fwd_schedule_grouped();
const Schedule &fwd_schedule() {
  static Schedule s = build_fwd_schedule_grouped();
  return s;
}

struct PackPtrs { const float *p[NNB_NUM_PARAMS]; };

// grid = stages, block = 256 threads: thread n fills row n of the [n x 16] image
__global__ void tc_pack_kernel(PackPtrs P, uint8_t *out, int n_stages, int which) {
  const int i = blockIdx.x;
  if (i >= n_stages) return;
  const TcStage st = which ? c_bwd_stages[i] : c_fwd_stages[i];
  const TcPack pk = which ? c_bwd_pack[i] : c_fwd_pack[i];
  for (int n = threadIdx.x; n < st.n; n += blockDim.x) {
    __nv_bfloat16 row[16];
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      float v = 0.f;
      if (pk.wparam >= 0 && pk.col0 + kk >= 0 && pk.col0 + kk < pk.col_end && (pk.n_valid == 0 || n < pk.n_valid)) {
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
    int nrow = n;
    if (pk.gnk > 0) {
      const int half = st.n >> 1, h = n >= half;
      dst += (size_t)h * pk.gnk * st.n * 16 + (size_t)pk.gk * st.n * 16;
      nrow = n - h * half;
    }
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) *reinterpret_cast<__nv_bfloat16 *>(dst + img_off(nrow, kk, 16)) = row[kk];
  }
}

}  // namespace


// ---------------------------------------------------------------------------------------------------------------------
// weight producer and MMA issuer loops shared by the forward and backward kernels (one elected thread each)
// ---------------------------------------------------------------------------------------------------------------------
namespace {

template <int RING>
__device__ __forceinline__ void weight_producer(const uint4 *groups, int n_groups, int n_pairs, const uint8_t *wimg, uint32_t ring,
                                                uint32_t bar_full, uint32_t bar_empty) {
  uint32_t slot = 0, phase = 0;
  for (int pair = blockIdx.x; pair < n_pairs; pair += gridDim.x)
    for (int i = 0; i < n_groups; ++i) {
      const uint4 g = groups[i];
      const uint32_t bytes = (g.z & 0xFFFFu) * 32u * ((g.y >> 18) & 3u);
      mbar_wait(bar_empty + 8 * slot, phase ^ 1);
      mbar_arrive_expect_tx(bar_full + 8 * slot, bytes);
      bulk_g2s(ring + slot * TC_SLOT, wimg + g.x, bytes, bar_full + 8 * slot);
      if (++slot == RING) { slot = 0; phase ^= 1; }
    }
}

// D(tile t) at tmem + t*256.  The next group's record is fetched before the barrier waits so that its latency hides there.
template <int RING>
__device__ __forceinline__ void mma_issuer(const uint4 *groups, int n_groups, int n_pairs, uint32_t sbase, uint32_t ring,
                                           uint32_t bar_full, uint32_t bar_empty, uint32_t bar_acc, uint32_t bar_act, uint32_t tmem,
                                           long long *trace) {
  uint32_t slot = 0, phase = 0, act_par = 0;
  uint4 nxt = groups[0];
  for (int pair = blockIdx.x; pair < n_pairs; pair += gridDim.x)
    for (int i = 0; i < n_groups; ++i) {
      const uint4 g = nxt;
      nxt = groups[(i + 1 == n_groups) ? 0 : i + 1];
      const uint32_t nk = (g.y >> 18) & 3u, first = (g.y >> 20) & 1u, last = (g.y >> 21) & 1u;
      const uint32_t n = g.z & 0xFFFFu, a_gs = g.z >> 16;
      if (first) {
        mbar_wait(bar_act, act_par);
        act_par ^= 1;
        if (trace && blockIdx.x == 0 && pair == 0) trace[2 * i] = clock64();
      }
      mbar_wait(bar_full + 8 * slot, phase);
      tc_fence_after();
      const uint32_t idesc = idesc_bf16(128, (int)n, 0, 0);
      const uint32_t a0 = sbase + (g.y & 0x3FFFFu), b0 = ring + slot * TC_SLOT;
#pragma unroll
      for (uint32_t k = 0; k < 2; ++k)
        if (k < nk) {
          const uint64_t bd = desc_kmajor(b0 + k * n * 32u, 256);
          umma(tmem, desc_kmajor(a0 + k * 256u, a_gs), bd, idesc, (first && k == 0) ? 0u : 1u);
          umma(tmem + 256, desc_kmajor(a0 + g.w + k * 256u, a_gs), bd, idesc, (first && k == 0) ? 0u : 1u);
        }
      umma_commit(bar_empty + 8 * slot);
      if (last) {
        umma_commit(bar_acc);
        if (trace && blockIdx.x == 0 && pair == 0) trace[2 * i + 1] = clock64();
      }
      if (++slot == RING) { slot = 0; phase ^= 1; }
    }
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------------------
// CTA-pair flavour (2-CTA clusters, tcgen05.mma.cta_group::2), used by the forward kernel.
// One MMA spans the pair: M = 256 = tile t of CTA 0 (128 rows) + tile t of CTA 1, and every CTA stages only HALF of the
// weight rows (N / 2) in its ring -- half the L2->SM and shared-memory bytes per FLOP of the single-CTA form.  That pays
// for running the two tiles of a CTA OUT of phase (layer L on tile 0, then layer L on tile 1, each consuming the layer's
// weight stages once): the epilogue of one tile overlaps the MMAs of the other instead of idling the tensor pipe.
//   producer (one warp per CTA)  : per group, lanes 0..nk-1 copy this CTA's half of each k-step image; full[slot] (tx)
//   forwarder (CTA 1, one thread): full[slot] observed -> remote arrive on the leader's peer_full[slot]
//   issuer (leader, one thread)  : act[t] (256 arrivals: the row owners of tile t in BOTH CTAs) at a layer's first group,
//                                  full[slot] + peer_full[slot], MMAs, multicast commit -> empty[slot] in both CTAs and,
//                                  after a layer's last group, -> acc[t] in both CTAs
// n_iters (ray pairs per CTA) is the same in both CTAs of a cluster.
// ---------------------------------------------------------------------------------------------------------------------
namespace {

constexpr int TC2_RING = 4;               // slots of TC2_SLOT bytes (this CTA's half of a group of up to TC2_GK k-steps of [256 x 16] images)
constexpr uint32_t TC2_GK = 4;
constexpr uint32_t TC2_SLOT = TC2_GK * 4096;
// pair-flavour group records: y = A offset (18 bits) | nk << 18 (3 bits) | first << 21 | last << 22

__device__ __forceinline__ int layer_end(const uint4 *groups, int i) {
  while (!((groups[i].y >> 22) & 1u)) ++i;
  return i + 1;
}

// whole warp
__device__ __forceinline__ void weight_producer_pair(const uint4 *groups, int n_groups, int n_iters, const uint8_t *wimg, uint32_t ring,
                                                     uint32_t bar_full, uint32_t bar_empty, uint32_t rank, int lane) {
  uint32_t slot = 0, phase = 0;
  const uint64_t keep = l2_policy_evict_last();
  for (int it = 0; it < n_iters; ++it)
    for (int i = 0; i < n_groups;) {
      const int j = layer_end(groups, i);
      for (int t = 0; t < 2; ++t)
        for (int gi = i; gi < j; ++gi) {
          const uint4 g = groups[gi];
          const uint32_t bytes = (g.z & 0xFFFFu) * 16u * ((g.y >> 18) & 7u);   // this CTA's half of the group: N/2 rows x nk k-steps
          if (lane == 0) {
            mbar_wait(bar_empty + 8 * slot, phase ^ 1);
            mbar_arrive_expect_tx(bar_full + 8 * slot, bytes);
            bulk_g2s_hint(ring + slot * TC2_SLOT, wimg + g.x + rank * bytes, bytes, bar_full + 8 * slot, keep);
          }
          __syncwarp();
          if (++slot == TC2_RING) { slot = 0; phase ^= 1; }
        }
      i = j;
    }
}

// CTA 1, lanes 0..TC2_RING-1 of one warp: lane l forwards slot l (every TC2_RING-th group), so that TC2_RING remote arrivals
// (several hundred cycles each) are in flight at once
__device__ __forceinline__ void full_forwarder_pair(const uint4 *groups, int n_groups, int n_iters, uint32_t bar_full, int lane) {
  if (lane >= TC2_RING) return;
  long long total = 0;
  for (int i = 0; i < n_groups;) {
    const int j = layer_end(groups, i);
    total += 2 * (j - i);
    i = j;
  }
  total *= n_iters;
  uint32_t phase = 0;
  for (long long g = lane; g < total; g += TC2_RING) {
    mbar_wait(bar_full + 8 * lane, phase);
    mbar_arrive_remote_relaxed(bar_full + 8 * lane, 0);   // same offset in the leader: its full[slot] counts this arrival
    phase ^= 1;
  }
}

// leader CTA, one thread.  bar_acc / bar_act: one barrier per tile (+0 / +8)
__device__ __forceinline__ void mma_issuer_pair(const uint4 *groups, int n_groups, int n_iters, uint32_t sbase, uint32_t ring,
                                                uint32_t bar_full, uint32_t bar_empty, uint32_t bar_acc, uint32_t bar_act, uint32_t tmem,
                                                long long *trace) {
  uint32_t slot = 0, phase = 0, act_par = 0;
  long long w_act = 0, w_full = 0, w_peer = 0, t_begin = clock64();   // w_peer: unused (merged into w_full)
  for (int it = 0; it < n_iters; ++it)
    for (int i = 0; i < n_groups;) {
      const int j = layer_end(groups, i);
      for (uint32_t t = 0; t < 2; ++t) {
        uint4 nxt = groups[i];
        for (int gi = i; gi < j; ++gi) {
          const uint4 g = nxt;
          if (gi + 1 < j) nxt = groups[gi + 1];
          const uint32_t nk = (g.y >> 18) & 7u, first = (g.y >> 21) & 1u, last = (g.y >> 22) & 1u;
          const uint32_t n = g.z & 0xFFFFu, a_gs = g.z >> 16;
          long long c0 = 0, c1 = 0, c2 = 0;
          if (trace) c0 = clock64();
          if (first) {
            mbar_wait_cluster(bar_act + 8 * t, (act_par >> t) & 1u);
            act_par ^= 1u << t;
          }
          if (trace) c1 = clock64();
          mbar_wait_cluster(bar_full + 8 * slot, phase);
          if (trace) { c2 = clock64(); w_act += c1 - c0; w_full += c2 - c1; }
          tc_fence_after();
          const uint32_t idesc = idesc_bf16(256, (int)n, 0, 0);
          const uint32_t a0 = sbase + (g.y & 0x3FFFFu) + t * g.w, b0 = ring + slot * TC2_SLOT, d = tmem + t * 256;
          // descriptors advance by plain adds on the 14-bit start-address field (256 B per k-step of A, N/2 rows x 32 B of B)
          uint64_t ad = desc_kmajor(a0, a_gs), bd = desc_kmajor(b0, 256);
          const uint64_t bstep = (uint64_t)(n * 16u >> 4);
#pragma unroll
          for (uint32_t k = 0; k < 4; ++k)
            if (k < nk) {
              umma2(d, ad, bd, idesc, (first && k == 0) ? 0u : 1u);
              ad += 16;
              bd += bstep;
            }
          umma2_commit_mc(bar_empty + 8 * slot, (uint16_t)3);
          if (last) umma2_commit_mc(bar_acc + 8 * t, (uint16_t)3);
          if (++slot == TC2_RING) { slot = 0; phase ^= 1; }
        }
      }
      i = j;
    }
  if (trace && blockIdx.x == 0) { trace[100] = w_act; trace[101] = w_full; trace[102] = w_peer; trace[103] = clock64() - t_begin; trace[104] = n_iters; }
}

}  // namespace

// =====================================================================================================================
// forward kernel
// =====================================================================================================================
namespace {

constexpr uint32_t SM_ACT = 0;                        // 2 x 65536  (128 rows x 256 cols bf16)
constexpr uint32_t SM_ENC = 131072;                   // 2 x 16384  (128 x 64)
constexpr uint32_t SM_AUX = SM_ENC + 32768;           // 2 x 512: 8 identical rows x 32 cols (one ray per tile), group stride 0
constexpr uint32_t SM_RING = SM_AUX + 1024;          // TC_RING x TC_SLOT
constexpr uint32_t SM_SMALL = SM_RING + 4 * 16384;   // ring: TC2_RING x TC2_SLOT  // fp32: w_den[256] w_rgb[384] b_den b_rgb[3] (+pad) = 648 floats
constexpr uint32_t SM_SCAN = SM_SMALL + 8 * 4;       // SM_SMALL: b_den, b_rgb[3]; the head weights are read through L1 (__ldg)      // 2 tiles x 4 warps x 8 floats
constexpr uint32_t SM_BARS = SM_SCAN + 2 * 4 * 8 * 4; // mbarriers
constexpr uint32_t SM_TOTAL = SM_BARS + 24 * 8 + 16;
static_assert(SM_TOTAL <= 232448, "shared memory budget");

constexpr uint32_t GS_ACT = 4096, GS_ENC = 1024, GS_AUX = 512;  // GS_AUX: stride of the SAVED aux image (global)

struct FwdArgs {
  const uint8_t *wimg;
  const float *origin, *dirs, *noise;
  const float *ray_o, *view;   // optional per-ray origins (R,3) / view directions (R,3) (NDC sampling); NULL: origin / dirs
  const float *w_den, *b_den, *w_rgb, *b_rgb;
  float *rgb, *dist, *weights, *alpha, *z_vals;
  int n_rays, S, n_stages;
  float near_, far_;
  int softplus, white, use_dir, dist_alpha;
  uint8_t *ws_fwd;   // SAVE: per-ray activation images (FWD_TILE_BYTES each)
  float4 *ws_s;      // SAVE: per-sample {rgb_s, raw density}
  long long *trace;  // debug: per-layer clock64 stamps of CTA 0 (nnb_debug_set_trace), else NULL
};

// saved-activation images of one ray (bf16, operand-image layout), written by the forward when SAVE, read by the
// backward (relu masks) and by the weight-gradient kernel (X operands)
constexpr uint32_t FW_H = 0;                 // H0..H7: 8 x 65536 (layer l at l*65536)
constexpr uint32_t FW_FEAT = 8 * 65536;      // fc_feature output (no activation)
constexpr uint32_t FW_HC = 9 * 65536;        // relu(rgb_layers.0), 128 cols (gs 2048)
constexpr uint32_t FW_ENC = FW_HC + 32768;   // positional encoding, 64 cols
constexpr uint32_t FW_AUX = FW_ENC + 16384;  // view-direction encoding + ones, 32 cols
constexpr uint32_t FW_MASK = FW_AUX + 8192;   // ReLU sign bitmasks: 9 layers (H0..H7, Hc) x 128 rows x 8 words
constexpr uint32_t FWD_TILE_BYTES = FW_MASK + 9 * 4096;
// gradient images written by the backward, read by the weight-gradient kernel
constexpr uint32_t BW_DY = 0;                // dY0..dY7; each 64 KB image is SAVED as two 128-column images (columns 0..127 | 128..255,
                                             // gs 2048) so that the weight-gradient kernel fetches its half of a stage with one bulk copy
constexpr uint32_t BW_DYFEAT = 8 * 65536;
constexpr uint32_t BW_DYRGB0 = 9 * 65536;    // 128 cols
constexpr uint32_t BW_DSMALL = BW_DYRGB0 + 32768;  // 16 cols: d_raw rgb (3), d_raw density (1)
constexpr uint32_t BWD_TILE_BYTES = BW_DSMALL + 4096;

__device__ __forceinline__ void st_global_v4(void *p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  // streaming (evict-first): the saved images are read back only after several GB of other traffic
  asm volatile("st.global.cs.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
// one 16-byte chunk (8 bf16) of an operand image: shared-memory copy (MMA operand) and optional global copy
template <bool SAVE>
__device__ __forceinline__ void put_chunk(uint32_t saddr, uint8_t *gaddr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  st_shared_v4(saddr, a, b, c, d);
  if (SAVE) st_global_v4(gaddr, a, b, c, d);
}

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
template <bool SAVE>
__device__ __forceinline__ void write_encodings(uint32_t enc_row, uint32_t aux_row, bool aux_smem, uint8_t *genc, uint8_t *gaux,
                                                const float p[3], const float v[3]) {
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
    put_chunk<SAVE>(enc_row + g * 128, genc + g * 128, pack_bf16(e[8 * g], e[8 * g + 1]), pack_bf16(e[8 * g + 2], e[8 * g + 3]),
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
  {
    const uint32_t w0 = pack_bf16(a[8 * g], a[8 * g + 1]), w1 = pack_bf16(a[8 * g + 2], a[8 * g + 3]),
                   w2 = pack_bf16(a[8 * g + 4], a[8 * g + 5]), w3 = pack_bf16(a[8 * g + 6], a[8 * g + 7]);
    if (aux_smem) st_shared_v4(aux_row + g * 128, w0, w1, w2, w3);
    if (SAVE) st_global_v4(gaux + g * 128, w0, w1, w2, w3);
  }
}

// 1 where the bf16 half is > 0, as 0xFFFF per half
__device__ __forceinline__ uint32_t gt0_mask2(uint32_t h) {
  uint32_t m;
  asm("set.gt.u32.bf16x2 %0, %1, %2;" : "=r"(m) : "r"(h), "r"(0u));
  return m;
}
// 32 fp32 accumulator columns -> 4 chunks of 8 bf16 (optionally relu) into the operand image (+ global copy).
// Returns the ReLU bitmask of the block: pair p (columns 2p, 2p+1) -> bit p (even column) and bit 16+p (odd column).
template <bool RELU, bool SAVE, bool GLOBAL = SAVE>
__device__ __forceinline__ uint32_t store32(uint32_t saddr, uint8_t *gaddr, const uint32_t (&v)[32]) {
  uint32_t bits = 0;
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    uint32_t w[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float a = __uint_as_float(v[8 * g + 2 * k]), b = __uint_as_float(v[8 * g + 2 * k + 1]);
      w[k] = RELU ? pack_bf16_relu(a, b) : pack_bf16(a, b);
      if (RELU && SAVE) bits |= gt0_mask2(w[k]) & ((1u << (4 * g + k)) | (1u << (16 + 4 * g + k)));
    }
    put_chunk<GLOBAL>(saddr + g * 128, gaddr + g * 128, w[0], w[1], w[2], w[3]);
  }
  return bits;
}

// row owner -> leader's issuer: "my row of tile t is published / my accumulator row is drained"
// (CTA 1: the row's shared-memory stores are already fenced towards the async proxy of its own SM -- the only reader, the
// SM's tensor core -- so the notification itself carries no release: the release.cluster form costs ~900 cycles per layer.)
__device__ __forceinline__ void act_arrive(uint32_t bar, uint32_t rank) {
  if (rank == 0) mbar_arrive(bar);
  else mbar_arrive_remote_relaxed(bar, 0);
}

// FWD_BULK_SAVE: the 64 KB activation images H0..H7 and `feature` leave shared memory as ONE cp.async.bulk shared -> global each
// (issued by a lane of warp 10 once the tile's 128 row owners have published the image, tests/cuda/bulk_store_bw.cu: 43 GB/s per
// SM) instead of 32 st.global.v4 per row owner; the next epilogue of the tile waits until the copy has read the buffer.
// Measured on B200 (4,096 rays): parity green, forward+save 0.665 -> 0.778 ms -- while the tensor core streams its operands
// out of the same shared memory the bulk copy drains too slowly and the row owners end up waiting for the buffer.  Off.
#ifndef FWD_BULK_SAVE
#define FWD_BULK_SAVE 0
#endif
constexpr int FWD_THREADS_SAVE = FWD_BULK_SAVE ? 352 : 320;

template <bool SAVE>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(352, 1) tc_fwd_kernel(const FwdArgs A) {
  constexpr bool BULK = SAVE && FWD_BULK_SAVE;
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t sbase = smem_u32(smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float *sm_small = reinterpret_cast<float *>(smem + SM_SMALL);
  float *sm_scan = reinterpret_cast<float *>(smem + SM_SCAN);
  // full / empty: TC2_RING each; acc / act: one per tile (+0 / +8)
  const uint32_t bar_full = sbase + SM_BARS, bar_empty = bar_full + TC2_RING * 8, bar_acc = bar_empty + TC2_RING * 8,
                 bar_act = bar_acc + 16, bar_actl = bar_act + 16, bar_sdone = bar_actl + 16;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + SM_BARS + 24 * 8);
  const uint32_t rank = cluster_ctarank();

  if (threadIdx.x == 0) {
    // leader: full[slot] completes on its own copies (arrive + tx bytes) AND the peer's forwarded "my half landed" arrival
    for (int i = 0; i < TC2_RING; ++i) { mbar_init(bar_full + 8 * i, rank == 0 ? 2 : 1); mbar_init(bar_empty + 8 * i, 1); }
    mbar_init(bar_acc, 1);
    mbar_init(bar_acc + 8, 1);
    mbar_init(bar_act, 256);       // used in the leader CTA: 128 local + 128 remote arrivals per tile
    mbar_init(bar_act + 8, 256);
    mbar_init(bar_actl, 128);      // BULK: this CTA's 128 row owners of a tile have published an image to be saved
    mbar_init(bar_actl + 8, 128);
    mbar_init(bar_sdone, 1);       // BULK: the bulk store has read the image, the buffer may be overwritten
    mbar_init(bar_sdone + 8, 1);
    mbar_fence_init();
  }
  if (threadIdx.x == 0) { sm_small[0] = A.b_den[0]; sm_small[1] = A.b_rgb[0]; sm_small[2] = A.b_rgb[1]; sm_small[3] = A.b_rgb[2]; }
  if (warp == 9) tmem_alloc2(smem_u32(tmem_slot), 512);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();   // the peer's barriers exist before anything signals them
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const int n_pairs = (A.n_rays + 1) >> 1;
  // both CTAs of a cluster run the same number of ray pairs (they share every MMA); a pair index >= n_pairs is a dummy
  // pass (both tiles invalid: computed, never stored)
  const int n_iters = (n_pairs - (int)(blockIdx.x & ~1u) + (int)gridDim.x - 1) / (int)gridDim.x;

  if (warp == 8) {
    // ------------------------------------------------ weight producer
    weight_producer_pair(c_fwd_groups, A.n_stages, n_iters, A.wimg, sbase + SM_RING, bar_full, bar_empty, rank, lane);
  } else if (warp == 9) {
    // ------------------------------------------------ MMA issuer
    if (rank == 0) {
      if (elect_one())
        mma_issuer_pair(c_fwd_groups, A.n_stages, n_iters, sbase, sbase + SM_RING, bar_full, bar_empty, bar_acc, bar_act, tmem, A.trace);
    } else {
      full_forwarder_pair(c_fwd_groups, A.n_stages, n_iters, bar_full, lane);
    }
    __syncwarp();
  } else if (warp == 10) {
    // ------------------------------------------------ BULK: image saver, lane t serves tile t
    if (BULK && lane < 2) {
      const int t = lane;
      uint32_t par = 0;
      for (int it = 0; it < n_iters; ++it) {
        const int ray = 2 * (blockIdx.x + it * gridDim.x) + t;
        const bool valid = ray < A.n_rays;
        uint8_t *gtile = A.ws_fwd + (size_t)(valid ? ray : 0) * FWD_TILE_BYTES;
        for (int L = 0; L < 9; ++L) {
          mbar_wait(bar_actl + 8 * t, par);
          par ^= 1;
          if (valid) {
            bulk_s2g(gtile + (uint32_t)L * 65536u, sbase + SM_ACT + t * 65536u, 65536u);
            bulk_commit();
            bulk_wait_read0();
          }
          mbar_arrive(bar_sdone + 8 * t);
        }
      }
      bulk_wait0();
    }
    __syncwarp();
  } else {
    // ------------------------------------------------ row owners: prologue, epilogues, heads, compositing
    const int t = warp >> 2, q = warp & 3, row = q * 32 + lane;
    uint32_t n_img = 0;   // BULK: images of this tile handed to the saver so far
    const uint32_t act_row = sbase + SM_ACT + t * 65536u + (row >> 3) * GS_ACT + (row & 7) * 16;
    const uint32_t enc_row = sbase + SM_ENC + t * 16384u + (row >> 3) * GS_ENC + (row & 7) * 16;
    const uint32_t aux_row = sbase + SM_AUX + t * 512u + (row & 7) * 16;  // rows 0..7 write it (same ray => same values)
    const uint32_t tacc = tmem + ((uint32_t)(q * 32) << 16) + t * 256;
    float *scan = sm_scan + t * 32;
    uint32_t acc_par = 0;
    const uint32_t bar_acc_t = bar_acc + 8 * t, bar_act_t = bar_act + 8 * t;
    for (int it = 0; it < n_iters; ++it) {
      const int pair = blockIdx.x + it * gridDim.x;
      const int ray_raw = 2 * pair + t;
      const bool valid = ray_raw < A.n_rays;
      const int ray = valid ? ray_raw : A.n_rays - 1;
      uint8_t *gtile = SAVE ? A.ws_fwd + (size_t)ray * FWD_TILE_BYTES : nullptr;
      const bool sv = SAVE && valid;
      // ---- prologue: stratified depth, sample position, encodings
      const float z = z_sample(row, A.S, A.near_, A.far_, A.noise ? A.noise + (size_t)ray * A.S : nullptr);
      {
        float p[3], v[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          const float d = A.dirs[3 * ray + c];
          p[c] = __fadd_rn(A.ray_o ? A.ray_o[3 * ray + c] : A.origin[c], __fmul_rn(d, z));
          v[c] = A.use_dir ? -(A.view ? A.view[3 * ray + c] : d) : 1.0f;
        }
        if (SAVE && valid) write_encodings<true>(enc_row, aux_row, row < 8, gtile + FW_ENC + (row >> 3) * GS_ENC + (row & 7) * 16,
                                                 gtile + FW_AUX + (row >> 3) * GS_AUX + (row & 7) * 16, p, v);
        else write_encodings<false>(enc_row, aux_row, row < 8, nullptr, nullptr, p, v);
      }
      fence_async_smem();
      tc_fence_before();
      act_arrive(bar_act_t, rank);

      float alpha = 0.f, r_s = 0.f, g_s = 0.f, b_s = 0.f, raw_den = 0.f;
      for (int L = 0; L < 10; ++L) {
        mbar_wait(bar_acc_t, acc_par);
        acc_par ^= 1;
        tc_fence_after();
        if (A.trace && blockIdx.x == 0 && pair == 0 && threadIdx.x == 0) A.trace[L * 4 + 0] = clock64();
        if (L < 9) {
          float ds0 = 0.f, ds1 = 0.f, ds2 = 0.f, ds3 = 0.f;
          uint32_t mw[8];
          if (BULK && n_img > 0) mbar_wait(bar_sdone + 8 * t, (n_img - 1) & 1);   // the previous image of this buffer has been read
#pragma unroll
          for (int c0 = 0; c0 < 256; c0 += 64) {
            uint32_t v0[32], v1[32];
            tmem_ld32(tacc + c0, v0);
            tmem_ld32(tacc + c0 + 32, v1);
            tmem_ld_wait();
            if (L == 7) {
              const float4 *wd = reinterpret_cast<const float4 *>(A.w_den + c0);   // same address in every lane: L1 broadcast
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const float4 w = __ldg(wd + j);
                ds0 = fmaf(fmaxf(__uint_as_float(v0[4 * j]), 0.f), w.x, ds0);
                ds1 = fmaf(fmaxf(__uint_as_float(v0[4 * j + 1]), 0.f), w.y, ds1);
                ds2 = fmaf(fmaxf(__uint_as_float(v0[4 * j + 2]), 0.f), w.z, ds2);
                ds3 = fmaf(fmaxf(__uint_as_float(v0[4 * j + 3]), 0.f), w.w, ds3);
              }
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const float4 w = __ldg(wd + 8 + j);
                ds0 = fmaf(fmaxf(__uint_as_float(v1[4 * j]), 0.f), w.x, ds0);
                ds1 = fmaf(fmaxf(__uint_as_float(v1[4 * j + 1]), 0.f), w.y, ds1);
                ds2 = fmaf(fmaxf(__uint_as_float(v1[4 * j + 2]), 0.f), w.z, ds2);
                ds3 = fmaf(fmaxf(__uint_as_float(v1[4 * j + 3]), 0.f), w.w, ds3);
              }
            }
            const uint32_t dst = act_row + (c0 >> 3) * 128;
            uint8_t *gdst = sv ? gtile + (uint32_t)L * 65536u + (row >> 3) * GS_ACT + (row & 7) * 16 + (c0 >> 3) * 128 : nullptr;
            if (L != 8) {
              if (sv) { mw[c0 >> 5] = store32<true, true, !BULK>(dst, gdst, v0); mw[(c0 >> 5) + 1] = store32<true, true, !BULK>(dst + 512, gdst + 512, v1); }
              else { store32<true, false>(dst, nullptr, v0); store32<true, false>(dst + 512, nullptr, v1); }
            } else {
              if (sv) { store32<false, true, !BULK>(dst, gdst, v0); store32<false, true, !BULK>(dst + 512, gdst + 512, v1); }
              else { store32<false, false>(dst, nullptr, v0); store32<false, false>(dst + 512, nullptr, v1); }
            }
          }
          if (sv && L != 8) {
            uint8_t *gm = gtile + FW_MASK + (uint32_t)L * 4096u + row * 32;
            st_global_v4(gm, mw[0], mw[1], mw[2], mw[3]);
            st_global_v4(gm + 16, mw[4], mw[5], mw[6], mw[7]);
          }
          if (L == 7) {
            // density head (model/official_nerf.py:66,77-83): softplus(beta 1, threshold 20) or relu, then 1 - exp(-d)
            const float raw = (ds0 + ds1) + (ds2 + ds3) + sm_small[0];
            raw_den = raw;
            const float d = A.softplus ? (raw > 20.f ? raw : log1pf(expf(raw))) : fmaxf(raw, 0.f);
            if (!A.dist_alpha) {
              alpha = 1.f - expf(-d);
            } else if (row == A.S - 1) {
              alpha = 1.f;   // model/rendering.py:122
            } else {
              // model/rendering.py:116-121: alpha = 1 - exp(-sigma * (z[i+1] - z[i]))
              const float zn = z_sample(row + 1, A.S, A.near_, A.far_, A.noise ? A.noise + (size_t)ray * A.S : nullptr);
              alpha = 1.f - expf(__fmul_rn(-d, __fsub_rn(zn, z)));
            }
          }
          fence_async_smem();
          tc_fence_before();
          act_arrive(bar_act_t, rank);
          if (BULK) { mbar_arrive(bar_actl + 8 * t); ++n_img; }
          if (A.trace && blockIdx.x == 0 && pair == 0 && threadIdx.x == 0) A.trace[L * 4 + 1] = clock64();
        } else {
          // rgb head: relu(rgb_layers.0) (128 wide, TMEM cols 0..127) . fc_rgb^T + b -> sigmoid
          float s0 = sm_small[1], s1 = sm_small[2], s2 = sm_small[3];
          uint32_t mc[4];
#pragma unroll
          for (int c0 = 0; c0 < 128; c0 += 32) {
            uint32_t v0[32];
            tmem_ld32(tacc + c0, v0);
            tmem_ld_wait();
            const float *w0 = A.w_rgb + c0, *w1 = w0 + 128, *w2 = w0 + 256;
            if (sv) {
              uint8_t *ghc = gtile + FW_HC + (row >> 3) * 2048 + (row & 7) * 16 + (c0 >> 3) * 128;
              uint32_t bits = 0;
#pragma unroll
              for (int g = 0; g < 4; ++g) {
                uint32_t w[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  w[k] = pack_bf16_relu(__uint_as_float(v0[8 * g + 2 * k]), __uint_as_float(v0[8 * g + 2 * k + 1]));
                  bits |= gt0_mask2(w[k]) & ((1u << (4 * g + k)) | (1u << (16 + 4 * g + k)));
                }
                st_global_v4(ghc + g * 128, w[0], w[1], w[2], w[3]);
              }
              mc[c0 >> 5] = bits;
            }
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const float h = fmaxf(__uint_as_float(v0[j]), 0.f);
              s0 = fmaf(h, __ldg(w0 + j), s0);
              s1 = fmaf(h, __ldg(w1 + j), s1);
              s2 = fmaf(h, __ldg(w2 + j), s2);
            }
          }
          r_s = 1.f / (1.f + expf(-s0));
          g_s = 1.f / (1.f + expf(-s1));
          b_s = 1.f / (1.f + expf(-s2));
          if (sv) {
            A.ws_s[(size_t)ray * 128 + row] = make_float4(r_s, g_s, b_s, raw_den);
            st_global_v4(gtile + FW_MASK + 8 * 4096u + row * 32, mc[0], mc[1], mc[2], mc[3]);
          }
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
  cluster_sync_all();   // neither CTA leaves (or frees TMEM) while pair-MMAs or remote arrivals may still be in flight
  if (warp == 9) tmem_dealloc2(tmem, 512);
}

}  // namespace

// =====================================================================================================================
// backward (data-gradient chain) kernel
// =====================================================================================================================
namespace {

constexpr int BW_RING = 4;
constexpr uint32_t SB_ACT = 0;                         // 2 x 65536: dY operand image, updated in place layer by layer
constexpr uint32_t SB_DSMALL = 131072;                 // 2 x 4096 : [128 x 16] d_raw {r,g,b,density}
constexpr uint32_t SB_RING = SB_DSMALL + 8192;         // BW_RING x TC_SLOT
constexpr uint32_t SB_SMALL = SB_RING + BW_RING * TC_SLOT;  // fp32 w_rgb[384]
constexpr uint32_t SB_SCAN = SB_SMALL + 384 * 4;       // 2 tiles x 64 floats
constexpr uint32_t SB_BARS = SB_SCAN + 2 * 64 * 4;
constexpr uint32_t SB_TOTAL = SB_BARS + 24 * 8 + 16;
static_assert(SB_TOTAL <= 232448, "shared memory budget");

struct BwdArgs {
  const uint8_t *wimg;       // backward stage images
  const uint8_t *ws_fwd;     // saved activations
  uint8_t *ws_bwd;           // dY images (output)
  const float4 *ws_s;        // per-sample {rgb_s, raw}
  const float *origin, *dirs, *alpha, *z_vals;
  const float *ray_o, *view;   // as in FwdArgs
  const float *d_rgb, *d_dist, *d_alpha_out, *d_weights_out;
  const float *w_rgb;
  float *d_origin, *d_dirs, *d_b_den, *d_b_rgb;   // d_origin / d_b_*: atomically accumulated
  float *d_ray_o, *d_view;     // per-ray gradients when ray_o / view are given
  int n_rays, n_stages;
  int softplus, white, use_dir, dist_alpha;
};

__device__ __forceinline__ uint4 ld_global_v4(const void *p) {
  uint4 r;
  asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
// keep the halves of the packed bf16x2 word `d` = columns (2p, 2p+1) of a 32-column block whose ReLU was active:
// `bits` is the block's mask word written by the forward (bit p = even column, bit 16+p = odd column).
template <int P>
__device__ __forceinline__ uint32_t relu_mask_bits(uint32_t d, uint32_t bits) {
  uint32_t m;
  const uint32_t t = bits << (15 - P);  // bit p -> sign of byte 1, bit 16+p -> sign of byte 3
  asm("prmt.b32 %0, %1, %1, 0xBB99;" : "=r"(m) : "r"(t));  // replicate those sign bits over the low / high half
  return d & m;
}
__device__ __forceinline__ float warp_excl_suffix(float v, int lane, float &total) {
  float inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const float t = __shfl_down_sync(0xffffffffu, inc, o);
    if (lane + o < 32) inc += t;
  }
  total = __shfl_sync(0xffffffffu, inc, 0);
  const float ex = __shfl_down_sync(0xffffffffu, inc, 1);
  return lane == 31 ? 0.f : ex;
}
// gradient of the positional encoding: x_bar = g[c] + sum_l 2^l (cos(2^l x) g_sin[l] - sin(2^l x) g_cos[l])
template <int LEVELS>
__device__ __forceinline__ void enc_backward(const float x[3], const uint32_t *g, float out[3]) {
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    float acc = __uint_as_float(g[c]);
    float sn, cs, f = 1.f;
    sincosf(x[c], &sn, &cs);
#pragma unroll
    for (int l = 0; l < LEVELS; ++l) {
      acc = fmaf(f, cs * __uint_as_float(g[3 + 6 * l + c]) - sn * __uint_as_float(g[3 + 6 * l + 3 + c]), acc);
      const float s2 = 2.f * sn * cs, c2 = 1.f - 2.f * sn * sn;
      sn = s2; cs = c2; f *= 2.f;
    }
    out[c] += acc;
  }
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(320, 1) tc_bwd_kernel(const BwdArgs A) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t sbase = smem_u32(smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float *sm_wrgb = reinterpret_cast<float *>(smem + SB_SMALL);
  float *sm_scan = reinterpret_cast<float *>(smem + SB_SCAN);
  // CTA-pair flavour, as the forward (see weight_producer_pair): full / empty: TC2_RING each; acc / act: one per tile (+0 / +8)
  const uint32_t bar_full = sbase + SB_BARS, bar_empty = bar_full + TC2_RING * 8, bar_acc = bar_empty + TC2_RING * 8,
                 bar_act = bar_acc + 16;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + SB_BARS + 24 * 8);
  const uint32_t rank = cluster_ctarank();
  if (threadIdx.x == 0) {
    for (int i = 0; i < TC2_RING; ++i) { mbar_init(bar_full + 8 * i, rank == 0 ? 2 : 1); mbar_init(bar_empty + 8 * i, 1); }
    mbar_init(bar_acc, 1);
    mbar_init(bar_acc + 8, 1);
    mbar_init(bar_act, 256);
    mbar_init(bar_act + 8, 256);
    mbar_fence_init();
  }
  for (int i = threadIdx.x; i < 384; i += blockDim.x) sm_wrgb[i] = A.w_rgb[i];
  if (warp == 9) tmem_alloc2(smem_u32(tmem_slot), 512);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const int n_pairs = (A.n_rays + 1) >> 1;
  const int n_iters = (n_pairs - (int)(blockIdx.x & ~1u) + (int)gridDim.x - 1) / (int)gridDim.x;   // same in both CTAs of a pair

  if (warp == 8) {
    weight_producer_pair(c_bwd_groups, A.n_stages, n_iters, A.wimg, sbase + SB_RING, bar_full, bar_empty, rank, lane);
  } else if (warp == 9) {
    if (rank == 0) {
      if (elect_one())
        mma_issuer_pair(c_bwd_groups, A.n_stages, n_iters, sbase, sbase + SB_RING, bar_full, bar_empty, bar_acc, bar_act, tmem, nullptr);
    } else {
      full_forwarder_pair(c_bwd_groups, A.n_stages, n_iters, bar_full, lane);
    }
    __syncwarp();
  } else {
    const int t = warp >> 2, q = warp & 3, row = q * 32 + lane;
    const uint32_t rowoff256 = (row >> 3) * 4096u + (row & 7) * 16u;
    const uint32_t rowoff128 = (row >> 3) * 2048u + (row & 7) * 16u;
    const uint32_t act_row = sbase + SB_ACT + t * 65536u + rowoff256;
    const uint32_t tacc = tmem + ((uint32_t)(q * 32) << 16) + t * 256;
    float *scan = sm_scan + t * 64;
    uint32_t acc_par = 0;
    const uint32_t bar_acc_t = bar_acc + 8 * t, bar_act_t = bar_act + 8 * t;
    float acc_origin[3] = {0.f, 0.f, 0.f}, acc_bias[4] = {0.f, 0.f, 0.f, 0.f};  // meaningful on row 0 of each tile
    for (int it = 0; it < n_iters; ++it) {
      const int pair = blockIdx.x + it * gridDim.x;
      const int ray_raw = 2 * pair + t;
      const bool valid = ray_raw < A.n_rays;
      const int ray = valid ? ray_raw : A.n_rays - 1;
      const size_t m = (size_t)ray * 128 + row;
      const uint8_t *ftile = A.ws_fwd + (size_t)ray * FWD_TILE_BYTES;
      uint8_t *btile = A.ws_bwd + (size_t)ray * BWD_TILE_BYTES;
      // ---- compositing backward (model/rendering.py:124-126) -> d_raw of the two heads
      const float alpha = A.alpha[m], z = A.z_vals[m];
      const float4 sv = A.ws_s[m];
      float g0 = 0.f, g1 = 0.f, g2 = 0.f, gd = 0.f;
      if (valid) {
        if (A.d_rgb) { g0 = A.d_rgb[3 * ray]; g1 = A.d_rgb[3 * ray + 1]; g2 = A.d_rgb[3 * ray + 2]; }
        if (A.d_dist) gd = A.d_dist[ray];
      }
      const float beta = 1.f - alpha + 1e-6f;
      const float inc = warp_incl_prod(beta, lane);
      if (lane == 31) scan[q] = inc;
      named_bar_sync(1 + t, 128);
      float T = __shfl_up_sync(0xffffffffu, inc, 1);
      if (lane == 0) T = 1.f;
      for (int k = 0; k < q; ++k) T *= scan[k];
      const float w = alpha * T;
      float g = g0 * sv.x + g1 * sv.y + g2 * sv.z + gd * z + (A.white ? -(g0 + g1 + g2) : 0.f);
      if (A.d_weights_out && valid) g += A.d_weights_out[m];
      float wtot;
      float G = warp_excl_suffix(g * w, lane, wtot);
      if (lane == 0) scan[4 + q] = wtot;
      named_bar_sync(1 + t, 128);
      for (int k = q + 1; k < 4; ++k) G += scan[4 + k];
      float dalpha = g * T - G / beta;
      if (A.d_alpha_out && valid) dalpha += A.d_alpha_out[m];
      const float dact = A.softplus ? (sv.w > 20.f ? 1.f : 1.f / (1.f + expf(-sv.w))) : (sv.w > 0.f ? 1.f : 0.f);
      float dr_den = dalpha * (1.f - alpha) * dact;
      if (A.dist_alpha) dr_den = (row == 127) ? 0.f : dr_den * __fsub_rn(A.z_vals[m + 1], z);   // d alpha / d sigma = delta * exp(-sigma delta)
      const float dr0 = w * g0 * sv.x * (1.f - sv.x), dr1 = w * g1 * sv.y * (1.f - sv.y), dr2 = w * g2 * sv.z * (1.f - sv.z);
      {
        const uint32_t ds = sbase + SB_DSMALL + t * 4096u + (row >> 3) * 256u + (row & 7) * 16u;
        const uint32_t p0 = pack_bf16(dr0, dr1), p1 = pack_bf16(dr2, dr_den);
        st_shared_v4(ds, p0, p1, 0u, 0u);
        st_shared_v4(ds + 128, 0u, 0u, 0u, 0u);
        if (valid) {
          uint8_t *gds = btile + BW_DSMALL + (row >> 3) * 256u + (row & 7) * 16u;
          st_global_v4(gds, p0, p1, 0u, 0u);
          st_global_v4(gds + 128, 0u, 0u, 0u, 0u);
        }
      }
      // ---- dY of rgb_layers.0 = (fc_rgb^T d_raw_rgb) masked by relu(Hc) > 0  -> 128-col operand image
      {
        const uint4 hcm = ld_global_v4(ftile + FW_MASK + 8 * 4096u + row * 32);
        const uint32_t hb[4] = {hcm.x, hcm.y, hcm.z, hcm.w};
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const float *w0 = sm_wrgb + 8 * j, *w1 = w0 + 128, *w2 = w0 + 256;
          float v[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) v[k] = w0[k] * dr0 + w1[k] * dr1 + w2[k] * dr2;
          const uint32_t bits = hb[j >> 2];
          uint32_t o0 = pack_bf16(v[0], v[1]), o1 = pack_bf16(v[2], v[3]), o2 = pack_bf16(v[4], v[5]), o3 = pack_bf16(v[6], v[7]);
          switch (j & 3) {
            case 0: o0 = relu_mask_bits<0>(o0, bits); o1 = relu_mask_bits<1>(o1, bits); o2 = relu_mask_bits<2>(o2, bits); o3 = relu_mask_bits<3>(o3, bits); break;
            case 1: o0 = relu_mask_bits<4>(o0, bits); o1 = relu_mask_bits<5>(o1, bits); o2 = relu_mask_bits<6>(o2, bits); o3 = relu_mask_bits<7>(o3, bits); break;
            case 2: o0 = relu_mask_bits<8>(o0, bits); o1 = relu_mask_bits<9>(o1, bits); o2 = relu_mask_bits<10>(o2, bits); o3 = relu_mask_bits<11>(o3, bits); break;
            default: o0 = relu_mask_bits<12>(o0, bits); o1 = relu_mask_bits<13>(o1, bits); o2 = relu_mask_bits<14>(o2, bits); o3 = relu_mask_bits<15>(o3, bits); break;
          }
          st_shared_v4(sbase + SB_ACT + t * 65536u + rowoff128 + j * 128, o0, o1, o2, o3);
          if (valid) st_global_v4(btile + BW_DYRGB0 + rowoff128 + j * 128, o0, o1, o2, o3);
        }
      }
      fence_async_smem();
      tc_fence_before();
      act_arrive(bar_act_t, rank);

      float xp[3], xv[3], d_p[3] = {0.f, 0.f, 0.f}, d_v[3] = {0.f, 0.f, 0.f};
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float d = A.dirs[3 * ray + c];
        xp[c] = __fadd_rn(A.ray_o ? A.ray_o[3 * ray + c] : A.origin[c], __fmul_rn(d, z));
        xv[c] = -(A.view ? A.view[3 * ray + c] : d);
      }
      // ---- the chain: 12 accumulator epilogues
      const uint8_t kind[12] = {EPI_DIR, EPI_FEAT, EPI_MASK, EPI_MASK, EPI_MASK, EPI_MASK, EPI_ENC, EPI_MASK, EPI_MASK, EPI_MASK, EPI_MASK, EPI_ENC};
      const int8_t slot_of[12] = {0, 0, 7, 6, 5, 4, 0, 3, 2, 1, 0, 0};
#pragma unroll 1
      for (int e = 0; e < 12; ++e) {
        const int ek = kind[e];
        uint32_t mw[8];
        if (ek == EPI_MASK) {
          const uint8_t *gm = ftile + FW_MASK + (uint32_t)slot_of[e] * 4096u + row * 32;
          const uint4 m0 = ld_global_v4(gm), m1 = ld_global_v4(gm + 16);
          mw[0] = m0.x; mw[1] = m0.y; mw[2] = m0.z; mw[3] = m0.w; mw[4] = m1.x; mw[5] = m1.y; mw[6] = m1.z; mw[7] = m1.w;
        }
        mbar_wait(bar_acc_t, acc_par);
        acc_par ^= 1;
        tc_fence_after();
        if (ek == EPI_DIR) {
          uint32_t v0[32];
          tmem_ld32(tacc, v0);
          tmem_ld_wait();
          if (A.use_dir) enc_backward<4>(xv, v0, d_v);
        } else if (ek == EPI_ENC) {
          uint32_t v0[64];
          tmem_ld32(tacc, *reinterpret_cast<uint32_t(*)[32]>(&v0[0]));
          tmem_ld32(tacc + 32, *reinterpret_cast<uint32_t(*)[32]>(&v0[32]));
          tmem_ld_wait();
          enc_backward<10>(xp, v0, d_p);
        } else {
          uint8_t *grow = btile + (ek == EPI_FEAT ? BW_DYFEAT : BW_DY + (uint32_t)slot_of[e] * 65536u) + rowoff128;   // split-halves layout
          const bool masked = (ek == EPI_MASK);
#pragma unroll
          for (int c0 = 0; c0 < 256; c0 += 32) {
            uint32_t v[32];
            tmem_ld32(tacc + c0, v);
            tmem_ld_wait();
            const uint32_t bits = masked ? mw[c0 >> 5] : 0xFFFFFFFFu;
            uint32_t o[16];
#pragma unroll
            for (int p = 0; p < 16; ++p) o[p] = pack_bf16(__uint_as_float(v[2 * p]), __uint_as_float(v[2 * p + 1]));
            o[0] = relu_mask_bits<0>(o[0], bits); o[1] = relu_mask_bits<1>(o[1], bits); o[2] = relu_mask_bits<2>(o[2], bits); o[3] = relu_mask_bits<3>(o[3], bits);
            o[4] = relu_mask_bits<4>(o[4], bits); o[5] = relu_mask_bits<5>(o[5], bits); o[6] = relu_mask_bits<6>(o[6], bits); o[7] = relu_mask_bits<7>(o[7], bits);
            o[8] = relu_mask_bits<8>(o[8], bits); o[9] = relu_mask_bits<9>(o[9], bits); o[10] = relu_mask_bits<10>(o[10], bits); o[11] = relu_mask_bits<11>(o[11], bits);
            o[12] = relu_mask_bits<12>(o[12], bits); o[13] = relu_mask_bits<13>(o[13], bits); o[14] = relu_mask_bits<14>(o[14], bits); o[15] = relu_mask_bits<15>(o[15], bits);
#pragma unroll
            for (int gch = 0; gch < 4; ++gch) {
              const uint32_t off = ((c0 >> 3) + gch) * 128;
              st_shared_v4(act_row + off, o[4 * gch], o[4 * gch + 1], o[4 * gch + 2], o[4 * gch + 3]);
              if (valid) st_global_v4(grow + (off >> 11) * 32768u + (off & 2047u), o[4 * gch], o[4 * gch + 1], o[4 * gch + 2], o[4 * gch + 3]);
            }
          }
        }
        if (e < 11) {
          fence_async_smem();
          tc_fence_before();
          act_arrive(bar_act_t, rank);
        }
      }
      // ---- per-ray reductions: d_dirs, and the CTA-level accumulators (d_origin, head biases)
      float r[10];
      r[0] = d_p[0]; r[1] = d_p[1]; r[2] = d_p[2];
      r[3] = z * d_p[0]; r[4] = z * d_p[1]; r[5] = z * d_p[2];
      if (!A.view) { r[3] -= d_v[0]; r[4] -= d_v[1]; r[5] -= d_v[2]; }
      r[6] = dr_den; r[7] = dr0; r[8] = dr1; r[9] = dr2;
#pragma unroll
      for (int k = 0; k < 10; ++k) {
        r[k] = warp_sum(r[k]);
        if (lane == 0) scan[8 + 4 * k + q] = r[k];
      }
      if (A.view) {   // separate view directions (NDC sampling): their gradient is its own output
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          const float sv_ = warp_sum(-d_v[k]);
          if (lane == 0) scan[48 + 4 * k + q] = sv_;
        }
      }
      named_bar_sync(1 + t, 128);
      if (row == 0 && valid) {
        float sres[10];
#pragma unroll
        for (int k = 0; k < 10; ++k) sres[k] = scan[8 + 4 * k] + scan[9 + 4 * k] + scan[10 + 4 * k] + scan[11 + 4 * k];
        if (A.d_ray_o) { A.d_ray_o[3 * ray] = sres[0]; A.d_ray_o[3 * ray + 1] = sres[1]; A.d_ray_o[3 * ray + 2] = sres[2]; }
        else { acc_origin[0] += sres[0]; acc_origin[1] += sres[1]; acc_origin[2] += sres[2]; }
        if (A.view && A.d_view) {
#pragma unroll
          for (int k = 0; k < 3; ++k) A.d_view[3 * ray + k] = scan[48 + 4 * k] + scan[49 + 4 * k] + scan[50 + 4 * k] + scan[51 + 4 * k];
        }
        A.d_dirs[3 * ray] = sres[3]; A.d_dirs[3 * ray + 1] = sres[4]; A.d_dirs[3 * ray + 2] = sres[5];
        acc_bias[0] += sres[6]; acc_bias[1] += sres[7]; acc_bias[2] += sres[8]; acc_bias[3] += sres[9];
      }
      named_bar_sync(1 + t, 128);
    }
    if (row == 0) {
      if (A.d_origin) { atomicAdd(A.d_origin, acc_origin[0]); atomicAdd(A.d_origin + 1, acc_origin[1]); atomicAdd(A.d_origin + 2, acc_origin[2]); }
      if (A.d_b_den) atomicAdd(A.d_b_den, acc_bias[0]);
      if (A.d_b_rgb) { atomicAdd(A.d_b_rgb, acc_bias[1]); atomicAdd(A.d_b_rgb + 1, acc_bias[2]); atomicAdd(A.d_b_rgb + 2, acc_bias[3]); }
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 9) tmem_dealloc2(tmem, 512);
}

}  // namespace

// =====================================================================================================================
// weight-gradient kernel: D[128 x N] += P^T Q over samples, P/Q = saved operand images viewed MN-major (K = samples)
// =====================================================================================================================
namespace {

constexpr int WG_MAX_JOBS = 32;
constexpr int WG_MAX_RING = 8;        // ring depth = min(8, ring bytes / stage bytes of the job): 3 for the 256-wide layers, 8 for the heads
constexpr int WG_SUB = 2;             // stages per 128-sample tile
constexpr int WG_GROUPS = 8;          // 8-sample row groups per stage (64 samples, 4 k-steps)
constexpr uint32_t WG_RING_BYTES = 184320;   // a stage is P 16 KB + Q parts up to 44 KB
constexpr uint32_t SW_RING = 0;
constexpr uint32_t SW_BARS = WG_RING_BYTES;
constexpr uint32_t SW_ONES = SW_BARS + 24 * 8 + 64;   // 256 B of bf16 1.0: a constant [16 samples x 8] operand (any layout of all-ones is all-ones)
constexpr uint32_t SW_TOTAL = SW_ONES + 256;
static_assert(SW_TOTAL <= 232448, "shared memory budget");

struct WgPart { uint32_t slot_off, gs; uint16_t n, tm_col; uint8_t ws, pad[3]; };
struct WgSeg { int32_t param, offset, stride_m, stride_n; uint16_t tm_col, count; uint8_t m0, m_n, pad[2]; };   // rows [m0, m0 + m_n) are flushed (m_n = 0: all 128)
struct WgJob {
  uint32_t p_slot_off, p_gs, p_col0;
  uint8_t p_ws, n_parts, n_segs, shared;   // shared: the two CTAs of a cluster stream the same Q images (multicast halves)
  uint16_t splits, cl0;    // clusters [cl0, cl0 + splits) work on this job pair; entry 2p / 2p+1 = cluster rank 0 / 1
  WgPart q[3];
  WgSeg seg[4];
  // optional second product on the same stage: D2[m][n] = sum_s P2[s][m] * Q[p2_q][s][p2_qcol0 + n].  P2 is a narrow image
  // (16 columns of per-sample head gradients); the MMA still has M = 128, rows past the image's width are never flushed.
  uint32_t p2_slot_off, p2_gs;
  uint16_t p2_n, p2_tm_col, p2_qcol0;
  uint8_t p2_on, p2_ws, p2_q;
  // bias gradient db[m] = sum_s P[s][m] as a product with a CONSTANT all-ones operand kept in shared memory (column ones_tm_col of
  // TMEM): the 256-wide layers used to stream the 8 KB `aux` image of every ray just for its ones column (6 % of the kernel's reads)
  uint8_t ones_on;
  uint16_t ones_tm_col;
};
__constant__ WgJob c_wg_jobs[WG_MAX_JOBS];

struct WgArgs {
  const uint8_t *ws_fwd, *ws_bwd;
  float *d_params[NNB_NUM_PARAMS];
  int n_tiles, n_jobs;
  long long *trace;  // debug (nnb_debug_set_trace): globaltimer stamps per CTA at [1024 + 4*cta]: start, streaming done, flushed
};

__device__ __forceinline__ long long globaltimer_ns() { long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(192, 1) tc_wgrad_kernel(const WgArgs A) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t sbase = smem_u32(smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t bar_full = sbase + SW_BARS, bar_empty = bar_full + WG_MAX_RING * 8, bar_done = bar_empty + WG_MAX_RING * 8;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + SW_BARS + 24 * 8);
  // which job pair / split: cluster c runs entry 2p + rank of pair p
  const uint32_t rank = cluster_ctarank();
  const int c = blockIdx.x >> 1;
  int p = 0;
  for (; 2 * p < A.n_jobs; ++p)
    if (c < c_wg_jobs[2 * p].cl0 + c_wg_jobs[2 * p].splits) break;
  const bool have_job = 2 * p < A.n_jobs;
  const WgJob &J = c_wg_jobs[have_job ? 2 * p + rank : 0];
  const bool shared = J.shared != 0;
  int t0 = 0, t1 = 0;
  if (have_job) {
    // shared pairs walk the same rays; independent entries split the range once more between the two ranks
    const int split = shared ? c - J.cl0 : 2 * (c - J.cl0) + (int)rank, nsplit = shared ? J.splits : 2 * J.splits;
    t0 = (int)((long long)A.n_tiles * split / nsplit);
    t1 = (int)((long long)A.n_tiles * (split + 1) / nsplit);
  }
  const bool active = t1 > t0;
  const uint16_t mc_mask = shared ? (uint16_t)3 : (uint16_t)(1u << rank);
  if (A.trace && threadIdx.x == 0) { A.trace[1024 + 4 * blockIdx.x] = globaltimer_ns(); A.trace[1024 + 4 * blockIdx.x + 3] = p; }
  if (threadIdx.x == 0) {
    for (int i = 0; i < WG_MAX_RING; ++i) { mbar_init(bar_full + 8 * i, 1); mbar_init(bar_empty + 8 * i, shared ? 2 : 1); }
    mbar_init(bar_done, 1);
    mbar_fence_init();
  }
  if (threadIdx.x < 64) reinterpret_cast<uint32_t *>(smem + SW_ONES)[threadIdx.x] = 0x3F803F80u;   // bf16 {1.0, 1.0}
  fence_async_smem();   // generic-proxy stores -> visible to the tensor core's operand reads
  if (warp == 5 && active) tmem_alloc(smem_u32(tmem_slot), 512);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();   // the peer's barriers exist before anything is multicast to them
  tc_fence_after();
  const uint32_t tmem = active ? *tmem_slot : 0;
  // stage = P (8 row groups x 2 KB) followed by the job's Q parts; small stages get a deeper ring (latency, not bandwidth,
  // bounds a CTA whose ring holds few bytes).  Both CTAs of a shared pair have the same geometry.
  uint32_t q_off[3];
  uint32_t bytes = WG_GROUPS * 2048;
  for (int q = 0; q < J.n_parts; ++q) { q_off[q] = bytes; bytes += WG_GROUPS * J.q[q].gs; }
  const uint32_t p2_off = bytes;
  if (J.p2_on) bytes += WG_GROUPS * J.p2_gs;
  const uint32_t n_slots = min((uint32_t)WG_MAX_RING, WG_RING_BYTES / bytes);
  if (!active) {
    // nothing to do (fewer rays than splits): fall through to the closing cluster barrier
  } else if (warp == 4) {
    // producer: lane 0 arms the barrier, then lanes 0..(copies-1) each issue one bulk copy.  In a shared pair each rank
    // fetches half of the row groups of every Q part and multicasts them to both CTAs.
    uint32_t slot = 0, phase = 0;
    const uint8_t *pbase = (J.p_ws ? A.ws_bwd : A.ws_fwd) + J.p_slot_off + (J.p_col0 >> 3) * 128;
    const size_t pstride = J.p_ws ? BWD_TILE_BYTES : FWD_TILE_BYTES;
    const bool p_contig = (J.p_gs == 2048);
    const int n_pcopies = p_contig ? 1 : WG_GROUPS;
    const uint32_t qg0 = shared ? rank * (WG_GROUPS / 2) : 0, qgn = shared ? WG_GROUPS / 2 : WG_GROUPS;
    for (int t = t0; t < t1; ++t)
      for (int sub = 0; sub < WG_SUB; ++sub) {
        if (lane == 0) {
          mbar_wait(bar_empty + 8 * slot, phase ^ 1);
          mbar_arrive_expect_tx(bar_full + 8 * slot, bytes);
        }
        __syncwarp();
        const uint32_t dst = sbase + SW_RING + slot * bytes;
        if (lane < n_pcopies) {
          if (p_contig) bulk_g2s(dst, pbase + (size_t)t * pstride + (size_t)(WG_GROUPS * sub) * 2048, WG_GROUPS * 2048, bar_full + 8 * slot);
          else bulk_g2s(dst + lane * 2048, pbase + (size_t)t * pstride + (size_t)(WG_GROUPS * sub + lane) * J.p_gs, 2048, bar_full + 8 * slot);
        } else if (lane < n_pcopies + J.n_parts) {
          const int q = lane - n_pcopies;
          const uint8_t *qb = (J.q[q].ws ? A.ws_bwd + (size_t)t * BWD_TILE_BYTES : A.ws_fwd + (size_t)t * FWD_TILE_BYTES) + J.q[q].slot_off;
          const uint8_t *srcp = qb + (size_t)(WG_GROUPS * sub + qg0) * J.q[q].gs;
          if (shared) bulk_g2s_mc(dst + q_off[q] + qg0 * J.q[q].gs, srcp, qgn * J.q[q].gs, bar_full + 8 * slot, mc_mask);
          else bulk_g2s(dst + q_off[q], srcp, qgn * J.q[q].gs, bar_full + 8 * slot);
        } else if (lane == n_pcopies + J.n_parts && J.p2_on) {
          const uint8_t *pb = (J.p2_ws ? A.ws_bwd + (size_t)t * BWD_TILE_BYTES : A.ws_fwd + (size_t)t * FWD_TILE_BYTES) + J.p2_slot_off;
          bulk_g2s(dst + p2_off, pb + (size_t)(WG_GROUPS * sub) * J.p2_gs, WG_GROUPS * J.p2_gs, bar_full + 8 * slot);
        }
        __syncwarp();
        if (++slot == n_slots) { slot = 0; phase ^= 1; }
      }
  } else if (warp == 5) {
    if (elect_one()) {
      uint32_t slot = 0, phase = 0, accum = 0;
      uint32_t qgs[3], qn_idesc[3], qtm[3];
      for (int q = 0; q < 3; ++q) {
        qgs[q] = q < J.n_parts ? J.q[q].gs : 0;
        qn_idesc[q] = idesc_bf16(128, q < J.n_parts ? J.q[q].n : 16, 1, 1);
        qtm[q] = tmem + (q < J.n_parts ? J.q[q].tm_col : 0);
      }
      const int n_parts = J.n_parts;
      const bool p2_on = J.p2_on != 0;
      const uint32_t p2_gs = J.p2_gs, p2_tm = tmem + J.p2_tm_col, p2_idesc = idesc_bf16(128, p2_on ? J.p2_n : 16, 1, 1);
      const uint32_t p2_b = p2_on ? q_off[J.p2_q] + (J.p2_qcol0 >> 3) * 128u : 0, p2_bgs = p2_on ? J.q[J.p2_q].gs : 0;
      const bool ones_on = J.ones_on != 0;
      const uint32_t ones_tm = tmem + J.ones_tm_col, ones_idesc = idesc_bf16(128, 8, 1, 1);
      const uint64_t ones_desc = desc_mnmajor(sbase + SW_ONES, 128);   // 2 x 128 B: the two 8-sample groups of a k-step
      for (int t = t0; t < t1; ++t)
        for (int sub = 0; sub < WG_SUB; ++sub) {
          mbar_wait(bar_full + 8 * slot, phase);
          tc_fence_after();
          const uint32_t src = sbase + SW_RING + slot * bytes;
#pragma unroll
          for (int ks = 0; ks < WG_GROUPS / 2; ++ks) {
            const uint64_t pd = desc_mnmajor(src + ks * 4096, 2048);
#pragma unroll
            for (int q = 0; q < 3; ++q)
              if (q < n_parts) umma(qtm[q], pd, desc_mnmajor(src + q_off[q] + ks * 2 * qgs[q], qgs[q]), qn_idesc[q], accum);
            if (p2_on) umma(p2_tm, desc_mnmajor(src + p2_off + ks * 2 * p2_gs, p2_gs), desc_mnmajor(src + p2_b + ks * 2 * p2_bgs, p2_bgs), p2_idesc, accum);
            if (ones_on) umma(ones_tm, pd, ones_desc, ones_idesc, accum);
            accum = 1;
          }
          if (shared) umma_commit_mc(bar_empty + 8 * slot, mc_mask);   // the slot is free once BOTH consumers are done
          else umma_commit(bar_empty + 8 * slot);
          if (++slot == n_slots) { slot = 0; phase ^= 1; }
        }
      umma_commit(bar_done);
    }
    __syncwarp();
  } else {
    // flush: thread = output row m (lane of TMEM), red.add into the fp32 gradient tensors
    mbar_wait(bar_done, 0);
    tc_fence_after();
    if (A.trace && threadIdx.x == 0) A.trace[1024 + 4 * blockIdx.x + 1] = globaltimer_ns();
    const int m = warp * 32 + lane;
    for (int sgi = 0; sgi < J.n_segs; ++sgi) {
      const WgSeg sg = J.seg[sgi];
      if (sg.m_n && (warp * 32 + 32 <= sg.m0 || warp * 32 >= sg.m0 + sg.m_n)) continue;   // no row of this warp (warp-uniform)
      const bool row_on = !sg.m_n || (m >= sg.m0 && m < sg.m0 + sg.m_n);   // the TMEM loads below stay warp-wide
      float *dst = A.d_params[sg.param] + sg.offset + (size_t)m * sg.stride_m;
      // rows of 256-wide weights are 16-byte aligned: four columns per reduction instead of one
      const bool vec4 = sg.stride_n == 1 && (sg.count & 31) == 0 && (sg.stride_m & 3) == 0 && (reinterpret_cast<uintptr_t>(dst) & 15) == 0;
      for (int c0 = 0; c0 < sg.count; c0 += 32) {
        uint32_t v[32];
        tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + sg.tm_col + c0, v);
        tmem_ld_wait();
        if (!row_on) continue;
        if (vec4) {
#pragma unroll
          for (int k = 0; k < 32; k += 4)
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + c0 + k), "f"(__uint_as_float(v[k])),
                         "f"(__uint_as_float(v[k + 1])), "f"(__uint_as_float(v[k + 2])), "f"(__uint_as_float(v[k + 3]))
                         : "memory");
        } else {
#pragma unroll
          for (int k = 0; k < 32; ++k)
            if (c0 + k < sg.count) atomicAdd(dst + (size_t)(c0 + k) * sg.stride_n, __uint_as_float(v[k]));
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (A.trace && threadIdx.x == 0) A.trace[1024 + 4 * blockIdx.x + 2] = globaltimer_ns();
  if (warp == 5 && active) tmem_dealloc(tmem, 512);
  cluster_sync_all();   // neither CTA leaves while the other may still signal its barriers
}

struct WgPlan { std::vector<WgJob> jobs; int n_ctas = 0; };

// Job pairs (entry 2p = cluster rank 0, 2p+1 = rank 1).  The kernel keeps one CTA per SM (512 TMEM columns, 184 KB of
// shared memory), so 74 two-CTA clusters are resident at once: the cluster counts sum to 74 and are proportional to the
// bytes a pair streams per ray (the kernel is memory-bound), so that every cluster starts at t = 0 and ends together:
//   7 x 8 (256-wide layers; fc_feature, which also carries fc_density's product) + 8 (layers1.0) + 4 (layers0.0)
//   + 4 (rgb_layers.0) + 2 (fc_rgb).  A CTA's share of a saturated memory system follows the bytes its ring keeps in flight,
//   hence the counts follow (bytes per ray / ring bytes), not bytes alone (tools/trace_wgrad.py prints the finish times).
WgPlan build_wg_plan() {
  WgPlan P;
  int n_clusters = 0;
  auto part = [](int ws, uint32_t off, uint32_t gs, int n, int tm) { WgPart q{}; q.ws = (uint8_t)ws; q.slot_off = off; q.gs = gs; q.n = (uint16_t)n; q.tm_col = (uint16_t)tm; return q; };
  auto seg = [](int param, int offset, int sm, int sn, int tm, int count) { WgSeg s{}; s.param = param; s.offset = offset; s.stride_m = sm; s.stride_n = sn; s.tm_col = (uint16_t)tm; s.count = (uint16_t)count; return s; };
  auto push_pair = [&](WgJob J0, WgJob J1, int shared, int clusters) {
    J0.shared = J1.shared = (uint8_t)shared;
    J0.splits = J1.splits = (uint16_t)clusters;
    J0.cl0 = J1.cl0 = (uint16_t)n_clusters;
    n_clusters += clusters;
    P.jobs.push_back(J0);
    P.jobs.push_back(J1);
  };
  int big_i = 0;
  const int big_clusters[7] = {8, 8, 8, 8, 8, 8, 8};
  for (int L = 0; L < 8; ++L) {
    WgJob H[2];
    for (int h = 0; h < 2; ++h) {
      WgJob J{};
      const int wp = P_L0 + 2 * L;
      J.p_ws = 1; J.p_slot_off = BW_DY + (uint32_t)L * 65536u + (uint32_t)h * 32768u; J.p_gs = 2048; J.p_col0 = 0;
      // the bias gradient comes from the constant all-ones operand (ones_on), not from the `aux` image's ones column
      if (L == 0) {
        J.n_parts = 1;
        J.q[0] = part(0, FW_ENC, 1024, 64, 0);
        J.ones_on = 1; J.ones_tm_col = 64;
        J.n_segs = 2;
        J.seg[0] = seg(wp, 128 * h * 63, 63, 1, 0, 63);
        J.seg[1] = seg(wp + 1, 128 * h, 1, 1, 64, 1);
      } else if (L == 4) {
        J.n_parts = 2;
        J.q[0] = part(0, FW_H + 3 * 65536u, 4096, 256, 0);
        J.q[1] = part(0, FW_ENC, 1024, 64, 256);
        J.ones_on = 1; J.ones_tm_col = 320;
        J.n_segs = 3;
        J.seg[0] = seg(wp, 128 * h * 319, 319, 1, 0, 256);
        J.seg[1] = seg(wp, 128 * h * 319 + 256, 319, 1, 256, 63);
        J.seg[2] = seg(wp + 1, 128 * h, 1, 1, 320, 1);
      } else {
        J.n_parts = 1;
        J.q[0] = part(0, FW_H + (uint32_t)(L - 1) * 65536u, 4096, 256, 0);
        J.ones_on = 1; J.ones_tm_col = 256;
        J.n_segs = 2;
        J.seg[0] = seg(wp, 128 * h * 256, 256, 1, 0, 256);
        J.seg[1] = seg(wp + 1, 128 * h, 1, 1, 256, 1);
      }
      H[h] = J;
    }
    push_pair(H[0], H[1], 1, L == 0 ? 4 : (L == 4 ? 8 : big_clusters[big_i++]));
  }
  {  // fc_feature
    WgJob H[2];
    for (int h = 0; h < 2; ++h) {
      WgJob J{};
      J.p_ws = 1; J.p_slot_off = BW_DYFEAT + (uint32_t)h * 32768u; J.p_gs = 2048; J.p_col0 = 0;
      J.n_parts = 1;
      J.q[0] = part(0, FW_H + 7 * 65536u, 4096, 256, 0);
      J.ones_on = 1; J.ones_tm_col = 256;
      J.n_segs = 2;
      J.seg[0] = seg(P_FEAT, 128 * h * 256, 256, 1, 0, 256);
      J.seg[1] = seg(P_FEAT + 1, 128 * h, 1, 1, 256, 1);
      // fc_density weight rides along: sum_s d_raw_den[s] * H7[s][128h + n] (row 3 of the 16-column head-gradient image)
      J.p2_on = 1; J.p2_ws = 1; J.p2_slot_off = BW_DSMALL; J.p2_gs = 256;
      J.p2_q = 0; J.p2_qcol0 = (uint16_t)(128 * h); J.p2_n = 128; J.p2_tm_col = 288;
      J.n_segs = 3;
      J.seg[2] = seg(P_DEN, 128 * h, 0, 1, 288, 128);
      J.seg[2].m0 = 3; J.seg[2].m_n = 1;
      H[h] = J;
    }
    push_pair(H[0], H[1], 1, big_clusters[big_i++]);
  }
  {  // rgb_layers.0: input [feat(256), dir_enc(27)]; one 128-row block, the two ranks split the rays
    WgJob J{};
    J.p_ws = 1; J.p_slot_off = BW_DYRGB0; J.p_gs = 2048; J.p_col0 = 0;
    J.n_parts = 2;
    J.q[0] = part(0, FW_FEAT, 4096, 256, 0);
    J.q[1] = part(0, FW_AUX, 512, 32, 256);
    J.n_segs = 3;
    J.seg[0] = seg(P_RGB0, 0, 283, 1, 0, 256);
    J.seg[1] = seg(P_RGB0, 256, 283, 1, 256, 27);
    J.seg[2] = seg(P_RGB0 + 1, 0, 1, 1, 256 + 27, 1);
    push_pair(J, J, 0, 4);
  }
  {  // fc_rgb weight (3,128): sum_s Hc[s][m] * d_raw_rgb[s][c]
    WgJob J{};
    J.p_ws = 0; J.p_slot_off = FW_HC; J.p_gs = 2048; J.p_col0 = 0;
    J.n_parts = 1;
    J.q[0] = part(1, BW_DSMALL, 256, 16, 0);
    J.n_segs = 1;
    J.seg[0] = seg(P_RGB, 0, 1, 128, 0, 3);
    push_pair(J, J, 0, 2);
  }
  P.n_ctas = 2 * n_clusters;
  return P;
}
const WgPlan &wg_plan() {
  static WgPlan p = build_wg_plan();
  return p;
}

}  // namespace

// =====================================================================================================================
// host side
// =====================================================================================================================
namespace {


// k-steps -> issue groups (see c_fwd_groups).  a_off/a_stride: shared-memory placement of each A image kind.
struct ALayout { uint32_t off[5], tile_stride[5], gs[5]; };
std::vector<uint4> build_groups(const Schedule &S, const ALayout &L) {
  std::vector<uint4> G;
  size_t i = 0;
  while (i < S.st.size()) {
    const TcStage &a = S.st[i];
    uint32_t nk = 1;
    if (i + 1 < S.st.size()) {
      const TcStage &b = S.st[i + 1];
      if (!a.last && !b.first && b.a_src == a.a_src && b.a_k == a.a_k + 1 && b.n == a.n && b.w_off == a.w_off + a.n * 32u) nk = 2;
    }
    const TcStage &z = S.st[i + nk - 1];
    uint4 g;
    g.x = a.w_off;
    g.y = (L.off[a.a_src] + (uint32_t)a.a_k * 256u) | (nk << 18) | ((uint32_t)a.first << 20) | ((uint32_t)z.last << 21);
    g.z = (uint32_t)a.n | (L.gs[a.a_src] << 16);
    g.w = L.tile_stride[a.a_src];
    G.push_back(g);
    i += nk;
  }
  return G;
}
// pair flavour: up to 4 consecutive k-steps per group (see TC2_SLOT)
std::vector<uint4> build_groups_pair(const Schedule &S, const ALayout &L) {
  std::vector<uint4> G;
  size_t i = 0;
  while (i < S.st.size()) {
    const TcStage &a = S.st[i];
    uint32_t nk = 1;
    while (nk < TC2_GK && i + nk < S.st.size()) {
      const TcStage &p = S.st[i + nk - 1], &b = S.st[i + nk];
      if (p.last || b.first || b.a_src != a.a_src || b.a_k != p.a_k + 1 || b.n != a.n || b.w_off != p.w_off + p.n * 32u) break;
      ++nk;
    }
    const TcStage &z = S.st[i + nk - 1];
    uint4 g;
    g.x = a.w_off;
    g.y = (L.off[a.a_src] + (uint32_t)a.a_k * 256u) | (nk << 18) | ((uint32_t)a.first << 21) | ((uint32_t)z.last << 22);
    g.z = (uint32_t)a.n | (L.gs[a.a_src] << 16);
    g.w = L.tile_stride[a.a_src];
    G.push_back(g);
    i += nk;
  }
  return G;
}
const ALayout kFwdLayout{{SM_ACT, SM_ENC, SM_AUX, 0, 0}, {65536u, 16384u, 512u, 0, 0}, {GS_ACT, GS_ENC, 0u, 0, 0}};
// the forward schedule with every stage told where it sits inside its issue group (TcPack::gk / gnk, w_off = group start)
Schedule build_fwd_schedule_grouped() {
  Schedule S = build_fwd_schedule();
  const std::vector<uint4> G = build_groups_pair(S, kFwdLayout);
  size_t i = 0;
  for (const uint4 &g : G) {
    const int nk = (int)((g.y >> 18) & 7u);
    for (int k = 0; k < nk; ++k) {
      S.st[i + k].w_off = g.x;
      S.pk[i + k].gk = (int16_t)k;
      S.pk[i + k].gnk = (int16_t)nk;
    }
    i += nk;
  }
  return S;
}
const ALayout kBwdLayout{{SB_ACT, 0, 0, SB_ACT, SB_DSMALL}, {65536u, 0, 0, 65536u, 4096u}, {4096u, 0, 0, 2048u, 256u}};
Schedule group_schedule(Schedule S, const ALayout &L) {
  const std::vector<uint4> G = build_groups_pair(S, L);
  size_t i = 0;
  for (const uint4 &g : G) {
    const int nk = (int)((g.y >> 18) & 7u);
    for (int k = 0; k < nk; ++k) {
      S.st[i + k].w_off = g.x;
      S.pk[i + k].gk = (int16_t)k;
      S.pk[i + k].gnk = (int16_t)nk;
    }
    i += nk;
  }
  return S;
}
Schedule build_bwd_schedule_grouped() { return group_schedule(build_bwd_schedule(), kBwdLayout); }
std::vector<uint4> build_fwd_groups() {
  // group boundaries from the plain schedule (the grouped one has w_off rewritten)
  return build_groups_pair(build_fwd_schedule(), kFwdLayout);
}
const std::vector<uint4> &fwd_groups() {
  static std::vector<uint4> g = build_fwd_groups();
  return g;
}
const std::vector<uint4> &bwd_groups() {
  static std::vector<uint4> g = build_groups_pair(build_bwd_schedule(), kBwdLayout);
  return g;
}

long long *g_trace = nullptr;
// debug timing of the two backward kernels (nnb_debug_bwd_timing): events around tc_bwd_kernel and tc_wgrad_kernel
int g_bwd_timing = 0;
cudaEvent_t g_bwd_ev[3] = {nullptr, nullptr, nullptr};
std::once_flag g_tables_once;
int g_tables_err = 0;

void upload_tables() {
  const Schedule &F = fwd_schedule(), &B = bwd_schedule();
  const WgPlan &W = wg_plan();
  if ((int)F.st.size() > TC_MAX_STAGES || (int)B.st.size() > TC_MAX_STAGES || (int)W.jobs.size() > WG_MAX_JOBS) {
    g_tables_err = NNB_ERR_UNSUPPORTED;
    return;
  }
  cudaError_t e = cudaMemcpyToSymbol(c_fwd_stages, F.st.data(), F.st.size() * sizeof(TcStage));
  if (e == cudaSuccess) e = cudaMemcpyToSymbol(c_fwd_pack, F.pk.data(), F.pk.size() * sizeof(TcPack));
  if (e == cudaSuccess) e = cudaMemcpyToSymbol(c_bwd_stages, B.st.data(), B.st.size() * sizeof(TcStage));
  if (e == cudaSuccess) e = cudaMemcpyToSymbol(c_bwd_pack, B.pk.data(), B.pk.size() * sizeof(TcPack));
  if (e == cudaSuccess && (fwd_groups().size() > TC_MAX_GROUPS || bwd_groups().size() > TC_MAX_GROUPS)) { g_tables_err = NNB_ERR_UNSUPPORTED; return; }
  if (e == cudaSuccess) e = cudaMemcpyToSymbol(c_fwd_groups, fwd_groups().data(), fwd_groups().size() * sizeof(uint4));
  if (e == cudaSuccess) e = cudaMemcpyToSymbol(c_bwd_groups, bwd_groups().data(), bwd_groups().size() * sizeof(uint4));
  if (e == cudaSuccess) e = cudaMemcpyToSymbol(c_wg_jobs, W.jobs.data(), W.jobs.size() * sizeof(WgJob));
  if (e == cudaSuccess) e = cudaFuncSetAttribute(tc_fwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SM_TOTAL);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(tc_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SM_TOTAL);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(tc_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SB_TOTAL);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(tc_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SW_TOTAL);
  g_tables_err = (int)e;
}

int ensure_tables() {
  std::call_once(g_tables_once, upload_tables);
  return g_tables_err;
}

inline size_t ws_fwd_bytes(int R) { return (size_t)R * FWD_TILE_BYTES; }
inline size_t ws_bwd_bytes(int R) { return (size_t)R * BWD_TILE_BYTES; }
inline size_t ws_s_bytes(int R) { return (size_t)R * 128 * sizeof(float4); }

}  // namespace

// workspace = [forward activation images | backward dY images | per-sample {rgb_s, raw}]
size_t nnb_tc_workspace_bytes(const nnb_render_cfg *cfg, int save) {
  if (!save) return 16;
  return ws_fwd_bytes(cfg->n_rays) + ws_bwd_bytes(cfg->n_rays) + ws_s_bytes(cfg->n_rays);
}

int nnb_tc_render_fwd(const nnb_render_cfg *cfg, const float *const *params, const void *packed, const float *origin,
                      const float *dirs, const nnb_ray_extras *ex, const float *noise, float *rgb, float *dist, float *weights, float *alpha, float *z_vals,
                      void *ws, size_t ws_bytes, int save, cudaStream_t s) {
  if (!packed || !params || !origin || !dirs || !rgb || !dist) return NNB_ERR_BAD_ARG;
  if (cfg->n_samples != 128) return NNB_ERR_UNSUPPORTED;  // tile = one ray of 128 samples
  if (cfg->n_rays <= 0) return 0;
  if (save) {
    if (!ws || ws_bytes < nnb_tc_workspace_bytes(cfg, 1)) return NNB_ERR_WORKSPACE;
    if (!weights || !alpha || !z_vals) return NNB_ERR_BAD_ARG;  // the backward reads them
  }
  int e = ensure_tables();
  if (e) return e;
  FwdArgs A{};
  A.wimg = static_cast<const uint8_t *>(packed);
  A.origin = origin; A.dirs = dirs; A.noise = noise;
  A.ray_o = ex ? ex->ray_origins : nullptr; A.view = ex ? ex->view_dirs : nullptr;
  A.w_den = params[P_DEN]; A.b_den = params[P_DEN + 1]; A.w_rgb = params[P_RGB]; A.b_rgb = params[P_RGB + 1];
  A.rgb = rgb; A.dist = dist; A.weights = weights; A.alpha = alpha; A.z_vals = z_vals;
  A.n_rays = cfg->n_rays; A.S = cfg->n_samples; A.n_stages = (int)fwd_groups().size();
  A.near_ = cfg->near_; A.far_ = cfg->far_;
  A.softplus = cfg->occ_softplus; A.white = cfg->white_bkgd; A.use_dir = cfg->use_dir; A.dist_alpha = cfg->dist_alpha;
  A.trace = g_trace;
  if (save) {
    A.ws_fwd = static_cast<uint8_t *>(ws);
    A.ws_s = reinterpret_cast<float4 *>(static_cast<uint8_t *>(ws) + ws_fwd_bytes(cfg->n_rays) + ws_bwd_bytes(cfg->n_rays));
  }
  const int n_pairs = (cfg->n_rays + 1) / 2;
  int grid = n_pairs < nnb_sm_count() ? n_pairs : nnb_sm_count();
  grid = (grid + 1) & ~1;                       // 2-CTA clusters
  if (grid > (nnb_sm_count() & ~1)) grid = nnb_sm_count() & ~1;
  if (save) tc_fwd_kernel<true><<<grid, FWD_THREADS_SAVE, SM_TOTAL, s>>>(A);
  else tc_fwd_kernel<false><<<grid, 320, SM_TOTAL, s>>>(A);
  NNB_CHECK_LAUNCH();
  return 0;
}

int nnb_tc_render_bwd(const nnb_render_cfg *cfg, const float *const *params, const void *packed, const float *origin,
                      const float *dirs, const nnb_ray_extras *ex, const float *weights, const float *alpha, const float *z_vals, const float *d_rgb,
                      const float *d_dist, const float *d_alpha_out, const float *d_weights_out, float *const *d_params,
                      float *d_origin, float *d_dirs, void *ws, size_t ws_bytes, cudaStream_t s) {
  (void)weights;
  if (!packed || !params || !origin || !dirs || !alpha || !z_vals || !d_params || !d_origin || !d_dirs) return NNB_ERR_BAD_ARG;
  if (cfg->n_samples != 128) return NNB_ERR_UNSUPPORTED;
  if (ex && ((ex->ray_origins && !ex->d_ray_origins) || (ex->view_dirs && !ex->d_view_dirs))) return NNB_ERR_BAD_ARG;
  if (!ws || ws_bytes < nnb_tc_workspace_bytes(cfg, 1)) return NNB_ERR_WORKSPACE;
  int e = ensure_tables();
  if (e) return e;
  NNB_CUDA(cudaMemsetAsync(d_origin, 0, 3 * sizeof(float), s));
  if (cfg->n_rays <= 0) return 0;
  const int R = cfg->n_rays;
  uint8_t *w8 = static_cast<uint8_t *>(ws);
  BwdArgs B{};
  B.wimg = static_cast<const uint8_t *>(packed) + fwd_schedule().bytes;
  B.ws_fwd = w8; B.ws_bwd = w8 + ws_fwd_bytes(R);
  B.ws_s = reinterpret_cast<const float4 *>(w8 + ws_fwd_bytes(R) + ws_bwd_bytes(R));
  B.origin = origin; B.dirs = dirs; B.alpha = alpha; B.z_vals = z_vals;
  if (ex) { B.ray_o = ex->ray_origins; B.view = ex->view_dirs; B.d_ray_o = ex->ray_origins ? ex->d_ray_origins : nullptr; B.d_view = ex->d_view_dirs; }
  B.d_rgb = d_rgb; B.d_dist = d_dist; B.d_alpha_out = d_alpha_out; B.d_weights_out = d_weights_out;
  B.w_rgb = params[P_RGB];
  B.d_origin = d_origin; B.d_dirs = d_dirs; B.d_b_den = d_params[P_DEN + 1]; B.d_b_rgb = d_params[P_RGB + 1];
  B.n_rays = R; B.n_stages = (int)bwd_groups().size();
  B.softplus = cfg->occ_softplus; B.white = cfg->white_bkgd; B.use_dir = cfg->use_dir; B.dist_alpha = cfg->dist_alpha;
  const int n_pairs = (R + 1) / 2;
  int grid = n_pairs < nnb_sm_count() ? n_pairs : nnb_sm_count();
  grid = (grid + 1) & ~1;                       // 2-CTA clusters
  if (grid > (nnb_sm_count() & ~1)) grid = nnb_sm_count() & ~1;
  if (g_bwd_timing) {
    for (int i = 0; i < 3; ++i)
      if (!g_bwd_ev[i]) NNB_CUDA(cudaEventCreate(&g_bwd_ev[i]));
    NNB_CUDA(cudaEventRecord(g_bwd_ev[0], s));
  }
  tc_bwd_kernel<<<grid, 320, SB_TOTAL, s>>>(B);
  NNB_CHECK_LAUNCH();
  if (g_bwd_timing) NNB_CUDA(cudaEventRecord(g_bwd_ev[1], s));
  WgArgs W{};
  W.ws_fwd = B.ws_fwd; W.ws_bwd = B.ws_bwd;
  W.trace = g_trace;
  for (int i = 0; i < NNB_NUM_PARAMS; ++i) W.d_params[i] = d_params[i];
  W.n_tiles = R; W.n_jobs = (int)wg_plan().jobs.size();
  cudaStream_t sw = (ex && ex->wgrad_stream) ? nnb_stream(ex->wgrad_stream) : s;
  if (sw != s) {   // fork: the weight gradient follows the data-gradient kernel on its own stream; the caller joins
    static cudaEvent_t fork_ev = nullptr;
    if (!fork_ev) NNB_CUDA(cudaEventCreateWithFlags(&fork_ev, cudaEventDisableTiming));
    NNB_CUDA(cudaEventRecord(fork_ev, s));
    NNB_CUDA(cudaStreamWaitEvent(sw, fork_ev, 0));
  }
  tc_wgrad_kernel<<<wg_plan().n_ctas, 192, SW_TOTAL, sw>>>(W);
  NNB_CHECK_LAUNCH();
  if (g_bwd_timing) NNB_CUDA(cudaEventRecord(g_bwd_ev[2], sw));
  return 0;
}

extern "C" {
/* debug: device buffer of >= 64 int64 receiving clock64 stamps of CTA 0 / pair 0 of the next forward launches */
int nnb_debug_set_trace(void *buf) { g_trace = static_cast<long long *>(buf); return 0; }

int nnb_debug_bwd_timing(int on, float *ms2) {
  if (ms2 && g_bwd_ev[2]) {   // read the last timed backward (synchronises on its last event)
    NNB_CUDA(cudaEventSynchronize(g_bwd_ev[2]));
    NNB_CUDA(cudaEventElapsedTime(ms2, g_bwd_ev[0], g_bwd_ev[1]));
    NNB_CUDA(cudaEventElapsedTime(ms2 + 1, g_bwd_ev[1], g_bwd_ev[2]));
  }
  g_bwd_timing = on;
  return 0;
}

size_t nnb_pack_weights_bytes(void) { return (size_t)fwd_schedule().bytes + (size_t)bwd_schedule().bytes; }

int nnb_pack_weights(const float *const *params_host, void *packed, void *stream) {
  if (!params_host || !packed) return NNB_ERR_BAD_ARG;
  int e = ensure_tables();
  if (e) return e;
  PackPtrs P;
  for (int i = 0; i < NNB_NUM_PARAMS; ++i) P.p[i] = params_host[i];
  const int nf = (int)fwd_schedule().st.size(), nb = (int)bwd_schedule().st.size();
  tc_pack_kernel<<<nf, 256, 0, nnb_stream(stream)>>>(P, static_cast<uint8_t *>(packed), nf, 0);
  NNB_CHECK_LAUNCH();
  tc_pack_kernel<<<nb, 256, 0, nnb_stream(stream)>>>(P, static_cast<uint8_t *>(packed) + fwd_schedule().bytes, nb, 1);
  NNB_CHECK_LAUNCH();
  return 0;
}
}
