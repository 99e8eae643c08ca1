// Probe: per-SM throughput of cp.async.bulk shared -> global (TMA store) for the saved-image pattern (tools, not a test):
// every CTA (1 per SM) stores 64 KB images from two alternating shared-memory buffers; one elected thread issues, waits with
// wait_group.read (buffer reusable) before re-issuing from the same buffer.  depth = copies in flight per CTA (1 or 2),
// chunk = bytes per bulk copy (65536 = one copy per image, 16384 = four).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../my-nope-nerf_b200/csrc/tc_common.cuh"
using namespace tc;

__global__ void __launch_bounds__(128, 1) k(uint8_t *base, int n_img_per_cta, int chunk, int depth) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t s0 = smem_u32(smem);
  for (int i = threadIdx.x; i < 2 * 65536 / 16; i += blockDim.x) reinterpret_cast<uint4 *>(smem)[i] = make_uint4(i, blockIdx.x, 1, 2);
  fence_async_smem();
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int i = 0; i < n_img_per_cta; ++i) {
      uint8_t *dst = base + ((size_t)blockIdx.x * n_img_per_cta + i) * 65536;
      const uint32_t src = s0 + (i & 1) * 65536;
      for (int o = 0; o < 65536; o += chunk) bulk_s2g(dst + o, src + o, chunk);
      bulk_commit();
      if (depth == 1) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      else asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
    }
    bulk_wait0();
  }
}

int main() {
  const int n_img = 4096 * 9, per = n_img / 148;
  uint8_t *buf;
  cudaMalloc(&buf, (size_t)148 * per * 65536);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 65536);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int depth = 1; depth <= 2; ++depth)
    for (int chunk = 65536; chunk >= 4096; chunk /= 4) {
      for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(e0);
        k<<<148, 128, 2 * 65536>>>(buf, per, chunk, depth);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (rep) printf("bulk s2g depth %d chunk %6d: %.3f ms, %.0f GB/s (%.1f GB/s per SM)\n", depth, chunk, ms, 148.0 * per * 65536 / ms / 1e6,
                        per * 65536.0 / ms / 1e6);
      }
    }
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
