// C-ABI glue of libnerf_b200.so: version / diagnostics and the MLP entry points, which dispatch on the
// arithmetic (fp32 CUDA-core path in mlp_simt.cu, bf16 tcgen05 path in mlp_tc*.cu).  No CPU fallback
// exists anywhere in this library: an unsupported configuration returns NERFB_E_UNSUPPORTED.
#include "common.cuh"

std::atomic<long long> g_nerfb_launches{0};

int64_t simt_workspace_bytes(const nerfb_mlp_desc& d, int64_t P, int sigma_only, int save);
int simt_mlp_fwd(const nerfb_mlp_desc& d, int in_mode, const float* params, const float* film, const float* in0,
                 const float* in1, int64_t n_rows, int n_samples, int sigma_only, int save, float* out, float* ws,
                 int64_t ws_bytes, void* stream);
int simt_mlp_bwd(const nerfb_mlp_desc& d, int in_mode, const float* params, const float* film, const float* in0,
                 const float* in1, int64_t n_rows, int n_samples, const float* out, const float* d_out, float* grad,
                 float* d_in, float* ws, int64_t ws_bytes, void* stream);
int64_t tc_workspace_bytes(const nerfb_mlp_desc& d, int64_t P, int sigma_only, int save);
bool tc_supported(const nerfb_mlp_desc& d, int in_mode);
struct TcComposite;
int tc_mlp_fwd(const nerfb_mlp_desc& d, int in_mode, const float* params, const float* film, const float* in0,
               const float* in1, int64_t n_rows, int n_samples, int sigma_only, int save, float* out, void* ws,
               int64_t ws_bytes, void* stream, const TcComposite* comp);
int tc_mlp_bwd(const nerfb_mlp_desc& d, int in_mode, const float* params, const float* film, const float* in0,
               const float* in1, int64_t n_rows, int n_samples, const float* out, const float* d_out, float* grad,
               float* d_in, void* ws, int64_t ws_bytes, void* stream);

static bool desc_ok(const nerfb_mlp_desc* d) {
  if (!d) return false;
  if (d->kind != NERFB_KIND_NERF && d->kind != NERFB_KIND_SIREN) return false;
  if (d->D < 1 || d->D > 16 || d->W < 8 || d->W > 1024 || (d->W & 1)) return false;
  if (d->n_freq_xyz < 0 || d->n_freq_xyz > 16 || d->n_freq_dir < 0 || d->n_freq_dir > 16) return false;
  if (d->n_cls < 0 || d->n_cls > 16) return false;
  if (d->skip_mask & 1) return false;          // layer 0 cannot be a skip layer
  if (d->kind == NERFB_KIND_SIREN && !(d->box > 0.f)) return false;
  return true;
}

extern "C" int nerfb_version(void) { return 100; }
extern "C" int nerfb_arch(void) { return 100; }
extern "C" int64_t nerfb_launch_count(void) { return (int64_t)g_nerfb_launches.load(); }

extern "C" const char* nerfb_error_string(int status) {
  switch (status) {
    case NERFB_OK: return "ok";
    case NERFB_E_BADARG: return "nerf_b200: bad argument (null pointer, negative size or malformed descriptor)";
    case NERFB_E_UNSUPPORTED: return "nerf_b200: configuration not implemented by the sm_100a kernels";
    case NERFB_E_WORKSPACE: return "nerf_b200: workspace too small (see nerfb_mlp_workspace_bytes)";
    case NERFB_E_ARCH: return "nerf_b200: device is not sm_100";
    default: return status > 0 ? cudaGetErrorString((cudaError_t)status) : "nerf_b200: unknown status";
  }
}

extern "C" int64_t nerfb_param_count(const nerfb_mlp_desc* d) { return desc_ok(d) ? nerf_offsets(*d).total : -1; }
extern "C" int nerfb_out_channels(const nerfb_mlp_desc* d) { return desc_ok(d) ? 4 + (d->kind == NERFB_KIND_NERF ? d->n_cls : 0) : -1; }

extern "C" int64_t nerfb_mlp_workspace_bytes(const nerfb_mlp_desc* d, int precision, int64_t n_points, int sigma_only,
                                             int save_for_backward) {
  if (!desc_ok(d) || n_points < 0) return NERFB_E_BADARG;
  if (precision == NERFB_PREC_FP32) return simt_workspace_bytes(*d, n_points, sigma_only, save_for_backward);
  if (precision == NERFB_PREC_BF16) return tc_workspace_bytes(*d, n_points, sigma_only, save_for_backward);
  return NERFB_E_BADARG;
}

extern "C" int nerfb_mlp_fwd(const nerfb_mlp_desc* d, int precision, int in_mode, const float* params, const float* film,
                             const float* in0, const float* in1, int64_t n_rows, int n_samples, int sigma_only,
                             int save_for_backward, float* out, void* workspace, int64_t workspace_bytes, void* stream) {
  NERFB_CHECK_ARG(desc_ok(d) && params && in0 && out && n_rows >= 0 && n_samples >= 1);
  NERFB_CHECK_ARG(in_mode == NERFB_IN_EMBEDDED || in_mode == NERFB_IN_GRID || (in_mode == NERFB_IN_RAYS && in1));
  if (in_mode == NERFB_IN_GRID && precision != NERFB_PREC_BF16) return NERFB_E_UNSUPPORTED;
  NERFB_CHECK_ARG(workspace || workspace_bytes == 0);
  if (n_rows == 0) return NERFB_OK;
  if (precision == NERFB_PREC_FP32)
    return simt_mlp_fwd(*d, in_mode, params, film, in0, in1, n_rows, n_samples, sigma_only, save_for_backward, out,
                        (float*)workspace, workspace_bytes, stream);
  if (precision == NERFB_PREC_BF16) {
    if (!tc_supported(*d, in_mode)) return NERFB_E_UNSUPPORTED;
    return tc_mlp_fwd(*d, in_mode, params, film, in0, in1, n_rows, n_samples, sigma_only, save_for_backward, out,
                      workspace, workspace_bytes, stream, nullptr);
  }
  return NERFB_E_BADARG;
}

extern "C" int nerfb_mlp_bwd(const nerfb_mlp_desc* d, int precision, int in_mode, const float* params, const float* film,
                             const float* in0, const float* in1, int64_t n_rows, int n_samples, const float* out,
                             const float* d_out, float* grad_params, float* d_in, void* workspace, int64_t workspace_bytes,
                             void* stream) {
  NERFB_CHECK_ARG(desc_ok(d) && params && in0 && out && d_out && grad_params && workspace && n_rows >= 0 && n_samples >= 1);
  NERFB_CHECK_ARG(in_mode == NERFB_IN_EMBEDDED || (in_mode == NERFB_IN_RAYS && in1 && !d_in));
  if (n_rows == 0) return NERFB_OK;
  if (precision == NERFB_PREC_FP32)
    return simt_mlp_bwd(*d, in_mode, params, film, in0, in1, n_rows, n_samples, out, d_out, grad_params, d_in,
                        (float*)workspace, workspace_bytes, stream);
  if (precision == NERFB_PREC_BF16) {
    if (!tc_supported(*d, in_mode) || d_in) return NERFB_E_UNSUPPORTED;
    return tc_mlp_bwd(*d, in_mode, params, film, in0, in1, n_rows, n_samples, out, d_out, grad_params, d_in, workspace,
                      workspace_bytes, stream);
  }
  return NERFB_E_BADARG;
}
