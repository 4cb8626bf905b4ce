// nerfb_render_fused_fwd (SURVEY 8b proposed export, 7 step 4; row N1): the whole inference render_rays of
// models/rendering.py:193-262 behind ONE call -- stratified sampling -> coarse embedding+MLP -> compositing ->
// importance resampling -> fine embedding+MLP -> compositing -- queued back to back on the caller's stream without
// returning to the host language in between.  On the tcgen05 arm, when a 128-sample tile holds whole rays (S in
// {32, 64, 128}) and there are no class channels, compositing runs in the MLP kernel's epilogue (mlp_tc.cu,
// epilogue_composite): per-sample `raw` (rgb, sigma) never reaches HBM and a pass is ONE launch; otherwise the pass is
// the MLP kernel followed by kernel (4).  Inference only (nothing is saved for a backward).
#include "common.cuh"
#include "tc_common.cuh"

bool tc_supported(const nerfb_mlp_desc& d, int in_mode);
bool tc_composite_supported(const nerfb_mlp_desc& d, int in_mode, int n_samples, int save);
int tc_mlp_fwd(const nerfb_mlp_desc& d, int in_mode, const float* params, const float* film, const float* in0,
               const float* in1, int64_t n_rows, int n_samples, int sigma_only, int save, float* out, void* ws,
               int64_t ws_bytes, void* stream, const TcComposite* comp);

namespace {

struct FusedRenderPlan {
  bool fuse_c, fuse_f;               // compositing in the MLP epilogue for the coarse / fine pass
  int Sc, Sf, Cc, Cf;                // samples and MLP output channels per pass
  int64_t off_z, off_wc, off_zf, off_wf, off_rawc, off_rawf, off_mlp, mlp_bytes, total;
};

bool plan(const nerfb_mlp_desc* dc, const nerfb_mlp_desc* df, const nerfb_render_cfg& cfg, int64_t n, FusedRenderPlan& pl) {
  const bool fine = cfg.n_importance > 0;
  pl.Sc = cfg.n_samples; pl.Sf = cfg.n_samples + cfg.n_importance;
  pl.Cc = cfg.test_time ? 1 : nerfb_out_channels(dc);
  pl.Cf = fine ? nerfb_out_channels(df) : 0;
  if (pl.Cc < 0 || pl.Cf < 0) return false;
  const bool bf16 = cfg.precision == NERFB_PREC_BF16;
  pl.fuse_c = bf16 && tc_composite_supported(*dc, NERFB_IN_RAYS, pl.Sc, 0);
  pl.fuse_f = fine && bf16 && tc_composite_supported(*df, NERFB_IN_RAYS, pl.Sf, 0);
  const int64_t mc = nerfb_mlp_workspace_bytes(dc, cfg.precision, n * pl.Sc, cfg.test_time, 0);
  const int64_t mf = fine ? nerfb_mlp_workspace_bytes(df, cfg.precision, n * pl.Sf, 0, 0) : 0;
  if (mc < 0 || mf < 0) return false;
  pl.mlp_bytes = mc > mf ? mc : mf;
  int64_t off = 0;
  auto take = [&](int64_t bytes) { const int64_t o = off; off += round_up64(bytes, 1024); return o; };
  pl.off_z = take(n * pl.Sc * 4);
  pl.off_wc = take(n * pl.Sc * 4);
  pl.off_zf = take(fine ? n * pl.Sf * 4 : 0);
  pl.off_wf = take(fine && !pl.fuse_f ? n * pl.Sf * 4 : 0);
  pl.off_rawc = take(pl.fuse_c ? 0 : n * pl.Sc * pl.Cc * 4);
  pl.off_rawf = take(!fine || pl.fuse_f ? 0 : n * pl.Sf * pl.Cf * 4);
  pl.off_mlp = take(pl.mlp_bytes);
  pl.total = off;
  return true;
}

}  // namespace

extern "C" int64_t nerfb_render_fused_workspace_bytes(const nerfb_mlp_desc* coarse, const nerfb_mlp_desc* fine,
                                                      const nerfb_render_cfg* cfg, int64_t n_rays) {
  if (!coarse || !cfg || n_rays < 0 || cfg->n_samples < 1 || cfg->n_importance < 0 || (cfg->n_importance > 0 && !fine))
    return NERFB_E_BADARG;
  FusedRenderPlan pl;
  if (!plan(coarse, fine, *cfg, n_rays, pl)) return NERFB_E_UNSUPPORTED;
  return pl.total;
}

extern "C" int nerfb_render_fused_fwd(const nerfb_mlp_desc* coarse, const nerfb_mlp_desc* fine, const nerfb_render_cfg* cfg,
                                      const float* params_coarse, const float* film_coarse, const float* params_fine,
                                      const float* film_fine, const nerfb_render_io* io, int64_t n_rays, void* workspace,
                                      int64_t workspace_bytes, void* stream) {
  NERFB_CHECK_ARG(coarse && cfg && io && params_coarse && io->rays && io->opacity_coarse && n_rays >= 0);
  NERFB_CHECK_ARG(cfg->n_samples >= 1 && cfg->n_importance >= 0);
  const bool fine_pass = cfg->n_importance > 0;
  NERFB_CHECK_ARG(!fine_pass || (fine && params_fine && io->rgb_fine && io->depth_fine && io->opacity_fine && cfg->n_samples >= 3));
  NERFB_CHECK_ARG(cfg->test_time || (io->rgb_coarse && io->depth_coarse));
  NERFB_CHECK_ARG(!(cfg->perturb > 0.f) || (io->perturb_rand && (!fine_pass || io->u)));
  NERFB_CHECK_ARG(cfg->noise_std == 0.f || (io->noise_coarse && (!fine_pass || io->noise_fine)));
  if (n_rays == 0) return NERFB_OK;
  FusedRenderPlan pl;
  if (!plan(coarse, fine, *cfg, n_rays, pl)) return NERFB_E_UNSUPPORTED;
  if (workspace_bytes < pl.total) return NERFB_E_WORKSPACE;
  NERFB_CHECK_ARG(workspace && (reinterpret_cast<uintptr_t>(workspace) & 1023) == 0);
  uint8_t* base = static_cast<uint8_t*>(workspace);
  float* z = reinterpret_cast<float*>(base + pl.off_z);
  float* wc = reinterpret_cast<float*>(base + pl.off_wc);
  float* zf = reinterpret_cast<float*>(base + pl.off_zf);
  float* wf = reinterpret_cast<float*>(base + pl.off_wf);
  void* mlp_ws = base + pl.off_mlp;
  const int64_t n = n_rays;
  const float* nz_c = cfg->noise_std != 0.f ? io->noise_coarse : nullptr;
  const float* nz_f = cfg->noise_std != 0.f ? io->noise_fine : nullptr;
  int st;
#define NERFB_TRY(expr) do { st = (expr); if (st != NERFB_OK) return st; } while (0)
  // (2) stratified depths, models/rendering.py:206-222
  NERFB_TRY(nerfb_sample_coarse(io->rays, cfg->perturb > 0.f ? io->perturb_rand : nullptr, cfg->perturb, cfg->use_disp, n, pl.Sc, z, stream));
  // coarse pass (:227-238): weights only under test_time
  if (pl.fuse_c) {
    TcComposite cp{1, pl.Sc, cfg->white_back, cfg->noise_std, nz_c, wc, cfg->test_time ? nullptr : io->rgb_coarse,
                   cfg->test_time ? nullptr : io->depth_coarse, io->opacity_coarse};
    NERFB_TRY(tc_mlp_fwd(*coarse, NERFB_IN_RAYS, params_coarse, film_coarse, io->rays, z, n, pl.Sc, cfg->test_time, 0, nullptr,
                         mlp_ws, pl.mlp_bytes, stream, &cp));
  } else {
    float* raw = reinterpret_cast<float*>(base + pl.off_rawc);
    NERFB_TRY(nerfb_mlp_fwd(coarse, cfg->precision, NERFB_IN_RAYS, params_coarse, film_coarse, io->rays, z, n, pl.Sc, cfg->test_time, 0,
                            raw, mlp_ws, pl.mlp_bytes, stream));
    NERFB_TRY(nerfb_composite_fwd(raw, pl.Cc, z, io->rays, nz_c, cfg->noise_std, cfg->white_back, n, pl.Sc, wc,
                                  cfg->test_time ? nullptr : io->rgb_coarse, cfg->test_time ? nullptr : io->depth_coarse,
                                  io->opacity_coarse, (!cfg->test_time && pl.Cc > 4) ? io->cls_coarse : nullptr, stream));
  }
  if (!fine_pass) return NERFB_OK;
  // (3) importance resampling, :241-247
  NERFB_TRY(nerfb_resample(z, wc, cfg->perturb > 0.f ? io->u : nullptr, cfg->perturb > 0.f ? 0 : 1, n, pl.Sc, cfg->n_importance, zf, stream));
  // fine pass, :249-260
  if (pl.fuse_f) {
    TcComposite cp{1, pl.Sf, cfg->white_back, cfg->noise_std, nz_f, nullptr, io->rgb_fine, io->depth_fine, io->opacity_fine};
    NERFB_TRY(tc_mlp_fwd(*fine, NERFB_IN_RAYS, params_fine, film_fine, io->rays, zf, n, pl.Sf, 0, 0, nullptr, mlp_ws, pl.mlp_bytes,
                         stream, &cp));
  } else {
    float* raw = reinterpret_cast<float*>(base + pl.off_rawf);
    NERFB_TRY(nerfb_mlp_fwd(fine, cfg->precision, NERFB_IN_RAYS, params_fine, film_fine, io->rays, zf, n, pl.Sf, 0, 0, raw, mlp_ws,
                            pl.mlp_bytes, stream));
    NERFB_TRY(nerfb_composite_fwd(raw, pl.Cf, zf, io->rays, nz_f, cfg->noise_std, cfg->white_back, n, pl.Sf, wf, io->rgb_fine,
                                  io->depth_fine, io->opacity_fine, pl.Cf > 4 ? io->cls_fine : nullptr, stream));
  }
#undef NERFB_TRY
  return NERFB_OK;
}
