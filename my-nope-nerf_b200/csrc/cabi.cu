// Library-level entry points: version, error strings, device info.
#include "nnb_common.cuh"
#include <atomic>

unsigned long long g_nnb_launches = 0;

int nnb_sm_count() {
  static std::atomic<int> cached{0};
  int v = cached.load();
  if (v > 0) return v;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  int n = 0;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) return 148;
  cached.store(n);
  return n;
}

extern "C" {

int nnb_version(void) { return 100; }

unsigned long long nnb_launch_count(void) { return g_nnb_launches; }

const char *nnb_error_string(int code) {
  if (code == 0) return "ok";
  if (code > 0) return cudaGetErrorString(static_cast<cudaError_t>(code));
  switch (code) {
    case NNB_ERR_BAD_ARG: return "nope_nerf_b200: bad argument";
    case NNB_ERR_WORKSPACE: return "nope_nerf_b200: workspace too small";
    case NNB_ERR_UNSUPPORTED: return "nope_nerf_b200: unsupported configuration";
    case NNB_ERR_NO_DEVICE: return "nope_nerf_b200: no sm_100 device";
    case NNB_ERR_TOO_LARGE: return "nope_nerf_b200: problem size exceeds the fixed capacity of this allocation-free entry point";
    default: return "nope_nerf_b200: unknown error";
  }
}

int nnb_device_info(int *sm_count_host, int *cc_major_host, int *cc_minor_host) {
  int dev = 0;
  NNB_CUDA(cudaGetDevice(&dev));
  int sm = 0, maj = 0, min = 0;
  NNB_CUDA(cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev));
  NNB_CUDA(cudaDeviceGetAttribute(&maj, cudaDevAttrComputeCapabilityMajor, dev));
  NNB_CUDA(cudaDeviceGetAttribute(&min, cudaDevAttrComputeCapabilityMinor, dev));
  if (sm_count_host) *sm_count_host = sm;
  if (cc_major_host) *cc_major_host = maj;
  if (cc_minor_host) *cc_minor_host = min;
  return 0;
}

}  // extern "C"
