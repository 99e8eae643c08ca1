// tcgen05 / TMEM tensor-core path of the NeRF MLP (NNB_PREC_BF16).  Placeholder until the fused kernels land:
// every entry reports NNB_ERR_UNSUPPORTED so that callers fail loudly instead of silently falling back.
#include "nnb_common.cuh"

size_t nnb_tc_workspace_bytes(const nnb_render_cfg *, int) { return 16; }
int nnb_tc_render_fwd(const nnb_render_cfg *, const float *const *, const void *, const float *, const float *, const float *,
                      float *, float *, float *, float *, float *, void *, size_t, int, cudaStream_t) {
  return NNB_ERR_UNSUPPORTED;
}
int nnb_tc_render_bwd(const nnb_render_cfg *, const float *const *, const void *, const float *, const float *, const float *,
                      const float *, const float *, const float *, const float *, const float *, const float *, float *const *,
                      float *, float *, void *, size_t, cudaStream_t) {
  return NNB_ERR_UNSUPPORTED;
}

extern "C" {
size_t nnb_pack_weights_bytes(void) { return 16; }
int nnb_pack_weights(const float *const *, void *, void *) { return NNB_ERR_UNSUPPORTED; }
}
