// K3 (TF32 path): tcgen05 / TMA GEMMs.  Placeholder until the tensor-core kernels land: every
// entry point fails loudly (there is no silent fallback to the FP32 path).
#include "ctr_common.cuh"

int ctr_tc_mlp_fwd(ctr_handle* h, cudaStream_t, int, int, int, const float*, int64_t, const float*, const float*, int,
                   float, uint64_t, float*, int64_t) {
    return ctr_fail(h, CTR_E_UNSUPPORTED, "TF32 tcgen05 path not built yet");
}
int ctr_tc_mlp_bwd_data(ctr_handle* h, cudaStream_t, int, int, int, const float*, int64_t, const float*, const float*,
                        int64_t, float, float*, int64_t) {
    return ctr_fail(h, CTR_E_UNSUPPORTED, "TF32 tcgen05 path not built yet");
}
int64_t ctr_tc_bwd_weight_ws_bytes(ctr_handle*, int, int, int) { return 256; }
int ctr_tc_mlp_bwd_weight(ctr_handle* h, cudaStream_t, int, int, int, const float*, int64_t, const float*, int64_t,
                          float*, void*, int64_t) {
    return ctr_fail(h, CTR_E_UNSUPPORTED, "TF32 tcgen05 path not built yet");
}
