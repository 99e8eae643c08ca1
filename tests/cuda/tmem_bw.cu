// Microbenchmark: TMEM -> register read throughput (tcgen05.ld 32x32b.x32) vs loads in flight and warps per quadrant.
#include <cstdio>
#include <cstdlib>
#include "../../my-nope-nerf_b200/csrc/tc_common.cuh"
using namespace tc;

template <int INFLIGHT>
__global__ void __launch_bounds__(512, 1) k(long long *out, int reps, int nwarps) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc(smem_u32(&slot), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = slot;
  uint32_t acc = 0;
  long long t0 = clock64();
  if (warp < nwarps) {
    const uint32_t base = tmem + ((uint32_t)((warp & 3) * 32) << 16) + ((warp >> 2) * 128) % 512;
    for (int r = 0; r < reps; ++r) {
      uint32_t v[INFLIGHT][32];
#pragma unroll
      for (int i = 0; i < INFLIGHT; ++i) tmem_ld32(base + (i * 32) % 128, v[i]);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < INFLIGHT; ++i)
#pragma unroll
        for (int j = 0; j < 32; j += 8) acc ^= v[i][j];
    }
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) out[0] = t1 - t0;
  if (acc == 0x12345) out[1] = acc;
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

int main() {
  long long *d, h[2];
  cudaMalloc(&d, 16);
  const int reps = 2000;
  for (int nw : {4, 8, 16}) {
    for (int inf : {1, 2, 4}) {
      if (inf == 1) k<1><<<1, 512>>>(d, reps, nw);
      if (inf == 2) k<2><<<1, 512>>>(d, reps, nw);
      if (inf == 4) k<4><<<1, 512>>>(d, reps, nw);
      cudaDeviceSynchronize();
      cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
      double bytes = (double)reps * inf * nw * 32 * 32 * 4;
      printf("warps %2d inflight %d: %lld cycles, %.1f B/cyc/SM, %.0f cycles per x32 load per warp\n", nw, inf, h[0], bytes / h[0],
             (double)h[0] / (reps * inf));
    }
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
