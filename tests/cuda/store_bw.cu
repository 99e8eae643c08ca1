// Probe: per-SM global store rate for the saved-image pattern of the tensor-core forward (tools, not a test).
//   each CTA (2 tiles x 128 threads) writes 64 KB images, thread = row, 32 x 16 B chunks 128 B apart
//   variant 0: ray-major images (stride 684,032 B between concurrently written images)
//   variant 1: layer-major (images of one layer contiguous: stride 65,536 B)
//   variant 2: linear (thread t writes 16 B at t*16 + k*4096)
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(256, 1) store_kernel(uint8_t *base, int n_pairs, int n_layers, size_t ray_stride, size_t layer_stride, int variant) {
  const int t = threadIdx.x >> 7, row = threadIdx.x & 127;
  for (int pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
    const int ray = 2 * pair + t;
    for (int L = 0; L < n_layers; ++L) {
      uint8_t *img = base + (size_t)ray * ray_stride + (size_t)L * layer_stride;
      uint4 v = make_uint4(ray, L, row, 7);
      if (variant < 2) {
        uint8_t *g = img + (row >> 3) * 4096 + (row & 7) * 16;
#pragma unroll
        for (int c = 0; c < 32; ++c) *reinterpret_cast<uint4 *>(g + c * 128) = v;
      } else {
        uint8_t *g = img + row * 16;
#pragma unroll
        for (int c = 0; c < 32; ++c) *reinterpret_cast<uint4 *>(g + c * 2048) = v;
      }
    }
  }
}
int main() {
  const int n_rays = 4096, n_layers = 9;
  const size_t tile = 684032;
  uint8_t *buf;
  cudaMalloc(&buf, (size_t)n_rays * tile);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int variant = 0; variant < 3; ++variant) {
    const size_t rs = (variant == 1) ? 65536 : tile, ls = (variant == 1) ? (size_t)n_rays * 65536 : 65536;
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(e0);
      store_kernel<<<148, 256>>>(buf, n_rays / 2, n_layers, rs, ls, variant);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      if (rep) printf("variant %d: %.3f ms, %.0f GB/s (%.1f GB/s per SM)\n", variant, ms, n_rays * 9.0 * 65536 / ms / 1e6, n_rays * 9.0 * 65536 / ms / 1e6 / 148);
    }
  }
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
