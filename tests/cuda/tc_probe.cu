// Bring-up probe for the tcgen05 primitives in csrc/tc_common.cuh (run on a B200: tests/test_gpu_tc_probe.py).
//   test 1: D[128x256] = A[128x64] * B[256x64]^T, both operands K-major no-swizzle images, B via bulk TMA stages
//   test 2: D[128x256] = P^T Q, P [128 samples x 128 feats], Q [128 samples x 256 feats]  (both MN-major views)
//   bench : issue rate of 128x256x16 MMAs and TMEM->register read rate
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include "../../my-nope-nerf_b200/csrc/tc_common.cuh"

using namespace tc;

#define CK(x)                                                                          \
  do {                                                                                 \
    cudaError_t e = (x);                                                               \
    if (e != cudaSuccess) {                                                            \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__);   \
      exit(2);                                                                         \
    }                                                                                  \
  } while (0)

static uint16_t f2bf(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  u += 0x7FFF + ((u >> 16) & 1);
  return (uint16_t)(u >> 16);
}
static float bf2f(uint16_t h) {
  uint32_t u = (uint32_t)h << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}

// mode 0: K-major A (128 x K) image + K-major B stages (each 256 x 16) ; mode 1: MN-major P (K x 128), Q (K x 256)
__global__ void __launch_bounds__(192, 1) probe_kernel(const uint8_t *Aimg, const uint8_t *Bimg, float *D, int mode, int K,
                                                       long long *cycles, int reps) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bars[4];
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t *sA = smem;                 // up to 64 KB
  uint8_t *sB = smem + 65536;         // up to 64 KB
  const uint32_t a_bytes = (mode == 0) ? 128 * K * 2 : K * 128 * 2;
  const uint32_t b_bytes = 256 * K * 2;
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bars[0]), 1);
    mbar_init(smem_u32(&bars[1]), 1);
    mbar_fence_init();
  }
  if (warp == 4) tmem_alloc(smem_u32(&tmem_slot), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (warp == 5) {
    if (elect_one()) {
      // operands by bulk TMA
      mbar_arrive_expect_tx(smem_u32(&bars[0]), a_bytes + b_bytes);
      bulk_g2s(smem_u32(sA), Aimg, a_bytes, smem_u32(&bars[0]));
      for (uint32_t off = 0; off < b_bytes; off += 8192) bulk_g2s(smem_u32(sB) + off, Bimg + off, 8192, smem_u32(&bars[0]));
      mbar_wait(smem_u32(&bars[0]), 0);
      tc_fence_after();
      long long t0 = clock64();
      for (int rep = 0; rep < reps; ++rep) {
        if (mode == 0) {
          const uint32_t idesc = idesc_bf16(128, 256, 0, 0);
          const uint32_t gsA = (K / 8) * 128;
          for (int ks = 0; ks < K / 16; ++ks) {
            uint64_t ad = desc_kmajor(smem_u32(sA) + ks * 256, gsA);
            uint64_t bd = desc_kmajor(smem_u32(sB) + ks * 8192, 256);  // stage image: 256 rows x 16 cols
            umma(tmem, ad, bd, idesc, (ks > 0) ? 1u : 0u);
          }
        } else {
          const uint32_t idesc = idesc_bf16(128, 256, 1, 1);
          // P image: K rows x 128 cols (gs = 2048); Q image: K rows x 256 cols (gs = 4096); one MMA = 16 rows (2 groups)
          for (int ks = 0; ks < K / 16; ++ks) {
            uint64_t ad = desc_mnmajor(smem_u32(sA) + ks * 2 * 2048, 2048);
            uint64_t bd = desc_mnmajor(smem_u32(sB) + ks * 2 * 4096, 4096);
            umma(tmem, ad, bd, idesc, (ks > 0) ? 1u : 0u);
          }
        }
      }
      umma_commit(smem_u32(&bars[1]));
      mbar_wait(smem_u32(&bars[1]), 0);
      long long t1 = clock64();
      cycles[0] = t1 - t0;
    }
    __syncwarp();
  }
  if (warp < 4) {
    mbar_wait(smem_u32(&bars[1]), 0);
    tc_fence_after();
    long long t0 = clock64();
    const int row = warp * 32 + lane;
    float acc = 0.f;
    for (int rep = 0; rep < (reps > 1 ? 16 : 1); ++rep)
      for (int c0 = 0; c0 < 256; c0 += 32) {
        uint32_t v[32];
        tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c0, v);
        tmem_ld_wait();
        if (rep == 0)
          for (int j = 0; j < 32; ++j) D[row * 256 + c0 + j] = __uint_as_float(v[j]);
        else
          for (int j = 0; j < 32; ++j) acc += __uint_as_float(v[j]);
      }
    long long t1 = clock64();
    if (threadIdx.x == 0) cycles[1] = t1 - t0;
    if (acc == 123.456f) D[0] = acc;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc(tmem, 512);
}

int main() {
  int dev = 0;
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, dev));
  printf("device %s sm_%d%d SMs %d\n", prop.name, prop.major, prop.minor, prop.multiProcessorCount);
  CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 131072));
  uint8_t *dA, *dB;
  float *dD;
  long long *dC;
  CK(cudaMalloc(&dA, 65536));
  CK(cudaMalloc(&dB, 65536));
  CK(cudaMalloc(&dD, 128 * 256 * 4));
  CK(cudaMalloc(&dC, 16));
  int fails = 0;
  for (int mode = 0; mode < 2; ++mode) {
    const int K = (mode == 0) ? 64 : 128;
    const int Ma = 128, Nb = 256;
    std::vector<float> A(Ma * K), B(Nb * K);
    srand(1 + mode);
    for (auto &x : A) x = bf2f(f2bf((rand() % 2001 - 1000) / 1000.f));
    for (auto &x : B) x = bf2f(f2bf((rand() % 2001 - 1000) / 1000.f));
    std::vector<uint8_t> ia(65536, 0), ib(65536, 0);
    if (mode == 0) {
      // A(m,k): image 128 rows x K cols.  B(n,k): K/16 stage images of 256 rows x 16 cols
      for (int m = 0; m < Ma; ++m)
        for (int k = 0; k < K; ++k) *(uint16_t *)&ia[img_off(m, k, K)] = f2bf(A[m * K + k]);
      for (int n = 0; n < Nb; ++n)
        for (int k = 0; k < K; ++k) *(uint16_t *)&ib[(k / 16) * 8192 + img_off(n, k % 16, 16)] = f2bf(B[n * K + k]);
    } else {
      // P(k,m): image K rows x 128 cols; Q(k,n): image K rows x 256 cols;  A(m,k) = P[k][m], B(n,k) = Q[k][n]
      for (int m = 0; m < Ma; ++m)
        for (int k = 0; k < K; ++k) *(uint16_t *)&ia[img_off(k, m, 128)] = f2bf(A[m * K + k]);
      for (int n = 0; n < Nb; ++n)
        for (int k = 0; k < K; ++k) *(uint16_t *)&ib[img_off(k, n, 256)] = f2bf(B[n * K + k]);
    }
    CK(cudaMemcpy(dA, ia.data(), 65536, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, ib.data(), 65536, cudaMemcpyHostToDevice));
    CK(cudaMemset(dD, 0, 128 * 256 * 4));
    probe_kernel<<<1, 192, 131072>>>(dA, dB, dD, mode, K, dC, 1);
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    std::vector<float> D(128 * 256);
    CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
    double maxerr = 0;
    for (int m = 0; m < Ma; ++m)
      for (int n = 0; n < Nb; ++n) {
        double s = 0;
        for (int k = 0; k < K; ++k) s += (double)A[m * K + k] * B[n * K + k];
        maxerr = fmax(maxerr, fabs(s - D[m * 256 + n]));
      }
    printf("mode %d (%s) K=%d max abs err %.3e  %s\n", mode, mode == 0 ? "K-major" : "MN-major", K, maxerr,
           maxerr < 1e-3 ? "OK" : "FAIL");
    if (!(maxerr < 1e-3)) {
      ++fails;
      printf("  D[0][0..7] =");
      for (int j = 0; j < 8; ++j) printf(" %.4f", D[j]);
      printf("\n");
    }
    // issue-rate bench: reps x (K/16) MMAs back to back
    long long cyc[2];
    const int reps = 64;
    probe_kernel<<<1, 192, 131072>>>(dA, dB, dD, mode, K, dC, reps);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(cyc, dC, 16, cudaMemcpyDeviceToHost));
    printf("  bench: %d MMAs (128x256x16) in %lld cycles = %.1f cyc/MMA; TMEM read 16 x 128x256 fp32 by 4 warps: %lld cycles = %.1f B/cyc\n",
           reps * (K / 16), cyc[0], (double)cyc[0] / (reps * (K / 16)), cyc[1], 16.0 * 128 * 256 * 4 / (double)cyc[1]);
  }
  printf(fails ? "PROBE FAILED\n" : "PROBE OK\n");
  return fails ? 1 : 0;
}
