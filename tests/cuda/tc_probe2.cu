// Bring-up probe for CTA pairs (tcgen05.mma.cta_group::2): D[256 x 256] = A[256 x 64] * B[256 x 64]^T on a 2-CTA cluster.
// CTA r holds A rows [128 r, 128 r + 128) (K-major image) and B rows (= D columns) [128 r, 128 r + 128) as per-k-step stage
// images [128 x 16]; the leader issues the MMAs; CTA 1 forwards "my half landed" to the leader with a remote mbarrier arrive;
// the accumulators come back through a multicast commit.  Also times back-to-back pair-MMAs.
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cstring>
#include <vector>
#include "../../my-nope-nerf_b200/csrc/tc_common.cuh"
using namespace tc;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(2); } } while (0)
static uint16_t f2bf(float f) { uint32_t u; memcpy(&u, &f, 4); u += 0x7FFF + ((u >> 16) & 1); return (uint16_t)(u >> 16); }
static float bf2f(uint16_t h) { uint32_t u = (uint32_t)h << 16; float f; memcpy(&f, &u, 4); return f; }

constexpr int K = 64;
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(192, 1) probe2(const uint8_t *Aimg, const uint8_t *Bimg, float *D, long long *cycles, int reps) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bars[4];     // 0: local full, 1: peer full (leader), 2: acc
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  uint8_t *sA = smem, *sB = smem + 16384;
  const uint32_t bar_full = smem_u32(&bars[0]), bar_peer = smem_u32(&bars[1]), bar_acc = smem_u32(&bars[2]);
  if (threadIdx.x == 0) { mbar_init(bar_full, 1); mbar_init(bar_peer, 1); mbar_init(bar_acc, 1); mbar_fence_init(); }
  if (warp == 4) tmem_alloc2(smem_u32(&tmem_slot), 256);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (warp == 5 && elect_one()) {
    // each CTA loads its A rows and its B half
    mbar_arrive_expect_tx(bar_full, 128 * K * 2 * 2);
    bulk_g2s(smem_u32(sA), Aimg + (size_t)rank * 128 * K * 2, 128 * K * 2, bar_full);
    bulk_g2s(smem_u32(sB), Bimg + (size_t)rank * 128 * K * 2, 128 * K * 2, bar_full);
    mbar_wait(bar_full, 0);
    if (rank == 1) {
      mbar_arrive_remote(bar_peer, 0);          // tell the leader that CTA 1's operands are in place
    } else {
      mbar_wait_cluster(bar_peer, 0);
      tc_fence_after();
      const uint32_t idesc = idesc_bf16(256, 256, 0, 0);
      long long t0 = clock64();
      for (int rep = 0; rep < reps; ++rep)
        for (int ks = 0; ks < K / 16; ++ks)
          umma2(tmem, desc_kmajor(smem_u32(sA) + ks * 256, (K / 8) * 128), desc_kmajor(smem_u32(sB) + ks * 4096, 256), idesc, ks > 0 ? 1u : 0u);
      umma2_commit_mc(bar_acc, 3);
      mbar_wait(bar_acc, 0);
      cycles[0] = clock64() - t0;
    }
  }
  __syncwarp();
  if (warp < 4) {
    mbar_wait(bar_acc, 0);
    tc_fence_after();
    const int row = warp * 32 + lane;
    for (int c0 = 0; c0 < 256; c0 += 32) {
      uint32_t v[32];
      tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c0, v);
      tmem_ld_wait();
      for (int j = 0; j < 32; ++j) D[(size_t)(rank * 128 + row) * 256 + c0 + j] = __uint_as_float(v[j]);
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 4) tmem_dealloc2(tmem, 256);
}

int main() {
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  printf("device %s sm_%d%d\n", prop.name, prop.major, prop.minor);
  CK(cudaFuncSetAttribute(probe2, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
  std::vector<float> A(256 * K), B(256 * K);
  srand(3);
  for (auto &x : A) x = bf2f(f2bf((rand() % 2001 - 1000) / 1000.f));
  for (auto &x : B) x = bf2f(f2bf((rand() % 2001 - 1000) / 1000.f));
  std::vector<uint8_t> ia(256 * K * 2), ib(256 * K * 2);
  for (int m = 0; m < 256; ++m)
    for (int k = 0; k < K; ++k) *(uint16_t *)&ia[(m / 128) * 128 * K * 2 + img_off(m % 128, k, K)] = f2bf(A[m * K + k]);
  for (int n = 0; n < 256; ++n)   // per CTA: K/16 stage images of [128 x 16]
    for (int k = 0; k < K; ++k) *(uint16_t *)&ib[(n / 128) * 128 * K * 2 + (k / 16) * 4096 + img_off(n % 128, k % 16, 16)] = f2bf(B[n * K + k]);
  uint8_t *dA, *dB; float *dD; long long *dC;
  CK(cudaMalloc(&dA, ia.size())); CK(cudaMalloc(&dB, ib.size())); CK(cudaMalloc(&dD, 256 * 256 * 4)); CK(cudaMalloc(&dC, 16));
  CK(cudaMemcpy(dA, ia.data(), ia.size(), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dB, ib.data(), ib.size(), cudaMemcpyHostToDevice));
  CK(cudaMemset(dD, 0, 256 * 256 * 4));
  probe2<<<2, 192, 65536>>>(dA, dB, dD, dC, 1);
  CK(cudaGetLastError());
  CK(cudaDeviceSynchronize());
  std::vector<float> D(256 * 256);
  CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
  double maxerr = 0;
  int bad = 0;
  for (int m = 0; m < 256; ++m)
    for (int n = 0; n < 256; ++n) {
      double s = 0;
      for (int k = 0; k < K; ++k) s += (double)A[m * K + k] * B[n * K + k];
      const double e = fabs(s - D[m * 256 + n]);
      if (e > 1e-3 && bad < 6) { printf("  mismatch D[%d][%d] = %.4f want %.4f\n", m, n, D[m * 256 + n], s); ++bad; }
      maxerr = fmax(maxerr, e);
    }
  printf("cta_group::2 M=256 N=256 K=%d max abs err %.3e  %s\n", K, maxerr, maxerr < 1e-3 ? "OK" : "FAIL");
  long long cyc;
  const int reps = 64;
  probe2<<<2, 192, 65536>>>(dA, dB, dD, dC, reps);
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpy(&cyc, dC, 8, cudaMemcpyDeviceToHost));
  printf("  %d pair-MMAs (256x256x16) in %lld cycles = %.1f cyc/MMA\n", reps * (K / 16), cyc, (double)cyc / (reps * (K / 16)));
  printf(maxerr < 1e-3 ? "PROBE2 OK\n" : "PROBE2 FAILED\n");
  return maxerr < 1e-3 ? 0 : 1;
}
