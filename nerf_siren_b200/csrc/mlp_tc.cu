// placeholder until the tcgen05 kernels land (replaced in the next commit)
#include "common.cuh"
int64_t tc_workspace_bytes(const nerfb_mlp_desc&, int64_t, int, int) { return NERFB_E_UNSUPPORTED; }
bool tc_supported(const nerfb_mlp_desc&, int) { return false; }
int tc_mlp_fwd(const nerfb_mlp_desc&, int, const float*, const float*, const float*, const float*, int64_t, int, int, int,
               float*, void*, int64_t, void*) { return NERFB_E_UNSUPPORTED; }
int tc_mlp_bwd(const nerfb_mlp_desc&, int, const float*, const float*, const float*, const float*, int64_t, int,
               const float*, const float*, float*, float*, void*, int64_t, void*) { return NERFB_E_UNSUPPORTED; }
