// Kernel (1), tensor-core arm, backward -- under construction: reports "unsupported" so that callers needing
// gradients stay on the fp32 arm (there is no CPU fallback either way).
#include "common.cuh"
bool tc_bwd_supported(const nerfb_mlp_desc&) { return false; }
int64_t tc_bwd_workspace_bytes(const nerfb_mlp_desc&, int64_t) { return 0; }
int tc_mlp_bwd(const nerfb_mlp_desc&, int, const float*, const float*, const float*, const float*, int64_t, int,
               const float*, const float*, float*, float*, void*, int64_t, void*) { return NERFB_E_UNSUPPORTED; }
