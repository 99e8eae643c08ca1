// Microbenchmark: L2 -> shared bulk-TMA (cp.async.bulk) throughput per SM for the weight-streaming pattern:
// every CTA streams the same `span` bytes in `chunk`-byte copies through a ring of `slots` buffers.
#include <cstdio>
#include <cstdlib>
#include "../../my-nope-nerf_b200/csrc/tc_common.cuh"
using namespace tc;

__global__ void __launch_bounds__(128, 1) k(const uint8_t *src, size_t span, size_t cta_stride, int chunk, int slots, int iters, long long *out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bars[32];
  if (threadIdx.x == 0) {
    for (int i = 0; i < slots; ++i) mbar_init(smem_u32(&bars[i]), 1);
    mbar_fence_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint8_t *base = src + (size_t)blockIdx.x * cta_stride;
    long long t0 = clock64();
    size_t off = 0;
    // prime
    for (int i = 0; i < slots; ++i) {
      mbar_arrive_expect_tx(smem_u32(&bars[i]), chunk);
      bulk_g2s(smem_u32(smem) + i * chunk, base + off, chunk, smem_u32(&bars[i]));
      off += chunk; if (off >= span) off = 0;
    }
    uint32_t phase = 0; int slot = 0;
    for (int it = 0; it < iters; ++it) {
      mbar_wait(smem_u32(&bars[slot]), phase);
      mbar_arrive_expect_tx(smem_u32(&bars[slot]), chunk);
      bulk_g2s(smem_u32(smem) + slot * chunk, base + off, chunk, smem_u32(&bars[slot]));
      off += chunk; if (off >= span) off = 0;
      if (++slot == slots) { slot = 0; phase ^= 1; }
    }
    for (int i = 0; i < slots; ++i) { mbar_wait(smem_u32(&bars[slot]), phase); if (++slot == slots) { slot = 0; phase ^= 1; } }
    long long t1 = clock64();
    out[blockIdx.x] = t1 - t0;
  }
}

int main() {
  const size_t span = 1261568;  // forward weight images
  uint8_t *d; long long *o, h[148];
  cudaMalloc(&d, span * 148);
  cudaMemset(d, 1, span * 148);
  cudaMalloc(&o, 148 * 8);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const int iters = 4000;
  struct C { int grid, chunk, slots; size_t stride; const char *name; } cs[] = {
      {148, 8192, 5, 0, "all SMs, shared 1.2 MB, 8 KB x5"}, {148, 8192, 12, 0, "all SMs, shared, 8 KB x12"},
      {148, 8192, 24, 0, "all SMs, shared, 8 KB x24"},      {148, 4096, 24, 0, "all SMs, shared, 4 KB x24"},
      {148, 16384, 12, 0, "all SMs, shared, 16 KB x12"},    {148, 8192, 12, span, "all SMs, private 1.2 MB each, 8 KB x12"},
      {1, 8192, 12, 0, "one SM, 8 KB x12"},                 {1, 8192, 24, 0, "one SM, 8 KB x24"},
      {74, 8192, 12, 0, "74 SMs, shared, 8 KB x12"}};
  for (auto &c : cs) {
    k<<<c.grid, 128, c.chunk * c.slots>>>(d, span, c.stride, c.chunk, c.slots, iters, o);
    cudaDeviceSynchronize();
    cudaMemcpy(h, o, c.grid * 8, cudaMemcpyDeviceToHost);
    long long mx = 0; for (int i = 0; i < c.grid; ++i) mx = h[i] > mx ? h[i] : mx;
    double bpc = (double)(iters + c.slots) * c.chunk / mx;
    printf("%-42s: %.1f B/cyc/SM, %.0f cycles per chunk, chip %.0f B/cyc  (%s)\n", c.name, bpc, (double)mx / (iters + c.slots), bpc * c.grid,
           cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
