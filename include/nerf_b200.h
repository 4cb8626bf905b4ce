/* nerf_b200.h -- C ABI of libnerf_b200.so: the render_rays hot path on B200 (sm_100a).
 *
 * Boundary contract:
 *   - every entry point is `extern "C"`, takes raw DEVICE pointers + sizes + a cudaStream_t (passed
 *     as void*), returns int: 0 = ok, > 0 = a cudaError_t from the launch, < 0 = NERFB_E_* below;
 *   - never throws, never allocates, never synchronises, never touches the legacy default stream
 *     unless the caller passes it; tensors are dense row-major fp32 unless stated otherwise;
 *   - inputs are never written (the reference never mutates its inputs, models/rendering.py:37).
 * The reference has no FFI for this path -- it is a Python call API over torch ops; each function
 * names the reference code (file:line under /root/reference) it replaces.  The native precedent for
 * the checks-then-launch shape is torchsearchsorted/src/cuda/searchsorted_cuda_wrapper.cpp:5-20.
 */
#ifndef NERF_B200_H
#define NERF_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NERFB_OK            0
#define NERFB_E_BADARG     (-1)   /* null pointer / negative size / unsupported shape */
#define NERFB_E_UNSUPPORTED (-2)  /* configuration outside what the kernels implement */
#define NERFB_E_WORKSPACE  (-3)   /* workspace too small */
#define NERFB_E_ARCH       (-4)   /* device is not sm_100 */

/* Model family.  Parameter blobs are flat fp32 in the reference's state_dict order. */
#define NERFB_KIND_NERF   0       /* models/nerf.py:41-124 NeRF; with n_cls>0: models/nerf_cls.py:12-112 NeRF_3D */
#define NERFB_KIND_SIREN  1       /* models/nerf.py:159-216 SemanticNeRF (FiLM sine layers) */

/* Arithmetic of the MLP contraction. */
#define NERFB_PREC_FP32   0       /* fp32 operands, fp32 accumulate, CUDA cores (tight-parity path) */
#define NERFB_PREC_BF16   1       /* bf16 operands on tcgen05 tensor cores, fp32 accumulate in TMEM */

/* Where the MLP input comes from. */
#define NERFB_IN_RAYS     0       /* rays (N,8) + z (N,S): x = o + d*z, encodings computed in-kernel
                                     (replaces models/rendering.py:224,131-150 + models/nerf.py:33-38) */
#define NERFB_IN_EMBEDDED 1       /* x (P, in_xyz [+ in_dir]) already embedded -- NeRF.forward(x) called
                                     directly, e.g. extract_color_mesh.py:136-139 */
#define NERFB_IN_GRID     2       /* dense occupancy query of extract_color_mesh.py:117-144: in0 = SIX float64 values on the
                                     device [xmin xmax ymin ymax zmin zmax], n_samples = N, n_rows = N^3; point p is row p of
                                     np.stack(np.meshgrid(x, y, z), -1).reshape(-1, 3) with np.linspace axes, generated and
                                     encoded in-kernel; bf16 arm.  sigma_only: out (N^3) = max(sigma, 0) (:141); else out
                                     (N^3, 4) = [rgb sigma] with view direction 0 (extract_mesh.ipynb cell 4) */

typedef struct nerfb_mlp_desc {
  int32_t kind;        /* NERFB_KIND_* */
  int32_t D;           /* trunk depth (8) */
  int32_t W;           /* width (256) */
  int32_t n_freq_xyz;  /* 10 -> in_xyz = 3*(1+2*10) = 63; 0 = identity (SIREN) */
  int32_t n_freq_dir;  /* 4  -> in_dir = 27 */
  int32_t skip_mask;   /* bit i: trunk layer i (0-based) consumes cat([xyz_enc, h]) (models/nerf.py:108-109) */
  int32_t n_cls;       /* NeRF_3D class channels (0 = plain NeRF) */
  int32_t logscale;    /* Embedding(logscale=...) models/nerf.py:16-19 */
  float   box;         /* SIREN UniformBoxWarp side length (models/nerf.py:193: 51) */
  int32_t in_xyz;      /* NERFB_IN_EMBEDDED only: columns of x that are the xyz encoding */
  int32_t in_dir;      /* NERFB_IN_EMBEDDED only: columns of x that are the dir encoding */
} nerfb_mlp_desc;

int         nerfb_version(void);                 /* 100*major + minor */
int         nerfb_arch(void);                    /* 100: built for sm_100a only */
const char* nerfb_error_string(int status);
int64_t     nerfb_param_count(const nerfb_mlp_desc* d);   /* floats in the flat parameter blob */
int         nerfb_out_channels(const nerfb_mlp_desc* d);  /* 4 (+ n_cls) */

/* ---- (2) stratified ray sampling: models/rendering.py:206-222 -------------------------------
 * rays (N,8) [o d near far]; perturb_rand (N,S) U[0,1) or NULL when perturb == 0; z_out (N,S). */
int nerfb_sample_coarse(const float* rays, const float* perturb_rand, float perturb, int use_disp,
                        int64_t n_rays, int n_samples, float* z_out, void* stream);

/* ---- (1a) positional encoding alone: models/nerf.py:21-38 Embedding.forward -----------------
 * x (B,C) -> out (B, C*(1+2*n_freqs)), channel order [x, sin f0 x, cos f0 x, sin f1 x, ...]. */
int nerfb_embed(const float* x, int64_t rows, int channels, int n_freqs, int logscale, float* out,
                void* stream);

/* ---- (1b) embedding + MLP, forward: models/rendering.py:129-159 + models/nerf.py:83-124 ------
 * in_mode RAYS: in0 = rays (N,8), in1 = z (N,S), n_rows = N, n_samples = S, P = N*S.
 * in_mode EMBEDDED: in0 = x (P, in_xyz(+in_dir)), in1 = NULL, n_rows = P, n_samples = 1.
 * SIREN: film = [15*frequencies+30 ((D+1)*W) | phase_shifts ((D+1)*W)]: the two FiLM vectors of
 *        forward_with_frequencies_phase_shifts (models/nerf.py:201) AFTER its 15*f+30 map (:202), which
 *        the host applies; slice l*W..(l+1)*W drives trunk layer l, the last slice the colour layer.
 *        NULL for NeRF.
 * out (P, sigma_only ? 1 : 4+n_cls): [r g b sigma cls...], sigma raw (models/nerf.py:112,122).
 * workspace: >= nerfb_mlp_workspace_bytes(..) bytes, 1024-aligned; with save_for_backward != 0 it
 * holds the activations nerfb_mlp_bwd needs and must be kept alive unchanged until then. */
int64_t nerfb_mlp_workspace_bytes(const nerfb_mlp_desc* d, int precision, int64_t n_points,
                                  int sigma_only, int save_for_backward);
int nerfb_mlp_fwd(const nerfb_mlp_desc* d, int precision, int in_mode, const float* params,
                  const float* film, const float* in0, const float* in1, int64_t n_rows,
                  int n_samples, int sigma_only, int save_for_backward, float* out,
                  void* workspace, int64_t workspace_bytes, void* stream);

/* ---- (1c) embedding + MLP, backward (autograd of the above) -----------------------------------
 * d_out (P, 4+n_cls) upstream gradient; grad_params (param_count) is ACCUMULATED into (+=), the
 * caller zeroes it; d_in (P, in_xyz+in_dir) or NULL (only NERFB_IN_EMBEDDED may ask for it). */
int nerfb_mlp_bwd(const nerfb_mlp_desc* d, int precision, int in_mode, const float* params,
                  const float* film, const float* in0, const float* in1, int64_t n_rows,
                  int n_samples, const float* out, const float* d_out, float* grad_params,
                  float* d_in, void* workspace, int64_t workspace_bytes, void* stream);

/* ---- (4) alpha compositing: models/rendering.py:161-190 (+ :379 class channels) --------------
 * raw (N,S,C) with C = 1 (sigma only -> weights only) or >= 4 ([rgb sigma extra...]); z (N,S);
 * rays (N,8) (direction norm, :168); noise (N,S) N(0,1) or NULL (:170).
 * Outputs (any may be NULL except weights): weights (N,S); rgb (N,3); depth (N); opacity (N);
 * extra (N,C-4). */
int nerfb_composite_fwd(const float* raw, int channels, const float* z, const float* rays,
                        const float* noise, float noise_std, int white_back, int64_t n_rays,
                        int n_samples, float* weights, float* rgb, float* depth, float* opacity,
                        float* extra, void* stream);
/* Backward: upstream d_rgb (N,3), d_depth (N), d_opacity (N), d_weights (N,S), d_extra (N,C-4), any
 * NULL = zero; writes d_raw (N,S,C) in full (no gradient reaches z: models/rendering.py:244). */
int nerfb_composite_bwd(const float* raw, int channels, const float* z, const float* rays,
                        const float* noise, float noise_std, int white_back, int64_t n_rays,
                        int n_samples, const float* d_rgb, const float* d_depth,
                        const float* d_opacity, const float* d_weights, const float* d_extra,
                        float* d_raw, void* stream);

/* ---- (3) importance resampling ------------------------------------------------------------------
 * nerfb_searchsorted: the torchsearchsorted contract (src/torchsearchsorted/searchsorted.py:20-53,
 *   src/cuda/searchsorted_cuda_kernel.cu:83-107): a (rows_a,A) sorted rows, v (rows_v,V), rows_a ==
 *   rows_v or either == 1; out int64 (max(rows),V) = numpy.searchsorted(a[r], v[r], side).
 *   With side_left = 0 and a = cdf, v = u this is the index step of sample_pdf
 *   (models/rendering.py:54), bit-exact for identical (cdf,u).
 * nerfb_sample_pdf: models/rendering.py:22-67 on explicit bins (N,nb), weights (N,nb-1); u (N,n_imp)
 *   or NULL when det (u = linspace(0,1,n_imp), :44); optional cdf_out (N,nb), inds_out (N,n_imp).
 * nerfb_resample: the render_rays call site :241-247 fused: z_mid, weights[:,1:-1], sample_pdf,
 *   cat, sort -> z_fine (N, S_c+S_i) ascending. */
int nerfb_searchsorted(const float* a, int64_t rows_a, int64_t len_a, const float* v, int64_t rows_v,
                       int64_t len_v, int side_left, int64_t* out, void* stream);
/* The same contract for every scalar type the reference's kernel is dispatched on (AT_DISPATCH_ALL_TYPES,
 * src/cuda/searchsorted_cuda_kernel.cu:134): a and v share `dtype`, out stays int64. */
#define NERFB_DT_F32 0
#define NERFB_DT_F64 1
#define NERFB_DT_I64 2
#define NERFB_DT_I32 3
#define NERFB_DT_I16 4
#define NERFB_DT_I8 5
#define NERFB_DT_U8 6
int nerfb_searchsorted_typed(const void* a, int64_t rows_a, int64_t len_a, const void* v, int64_t rows_v,
                             int64_t len_v, int dtype, int side_left, int64_t* out, void* stream);
int nerfb_sample_pdf(const float* bins, const float* weights, const float* u, int det, float eps,
                     int64_t n_rays, int n_bins, int n_importance, float* samples, float* cdf_out,
                     int64_t* inds_out, void* stream);
int nerfb_resample(const float* z_coarse, const float* weights_coarse, const float* u, int det,
                   int64_t n_rays, int n_coarse, int n_importance, float* z_fine, void* stream);

/* ---- (5) ray-sharded data-parallel step, device side ---------------------------------------------
 * The reference's step boundary is Lightning DDP + torch.optim.Adam (train.py:41-66, utils/__init__.py:11-31,
 * system.py:257-275).  Here the host all-reduces ONE flat gradient buffer over NCCL (SUM) and this entry applies
 * the Adam update (torch.optim.Adam semantics: lr, betas, eps, no weight decay, no amsgrad) to the flat
 * parameter buffer in one launch; grad_scale (= 1/world_size) turns the SUM into the DDP mean.  step >= 1 is
 * the 1-based update count (bias correction).  lr / betas / eps are doubles, as torch.optim.Adam holds them: 1 - beta
 * and the bias corrections are formed in double and rounded once.  All four buffers hold n floats, 16-byte aligned. */
int nerfb_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n,
                    double lr, double beta1, double beta2, double eps, int64_t step, float grad_scale,
                    void* stream);

/* ---- (6) EG3D tri-plane renderer --------------------------------------------
 * nerfb_triplane_relayout: planes (3, 32, H, W) as the reference holds them (eg3d_training/triplane.py:65,
 *   volumetric_rendering/renderer.py:57-61) -> channels-last (3, H, W, 32) so a texel is one 128-byte line.
 * nerfb_triplane_field: ImportanceRenderer.run_model (renderer.py:144-151) = sample_from_planes (:57-68: projections
 *   (x,y),(x,z),(z,x), bilinear, zeros padding, align_corners=False, coordinates * 2/box_warp) + OSGDecoder
 *   (triplane.py:144-167).  decoder = W1 (64x32) | b1 (64) | W2 (4x64) | b2 (4), the raw parameters of the two
 *   FullyConnectedLayers (networks_stylegan2.py:95-122; gains lr_mul/sqrt(fan_in) applied in-kernel).
 *   coords (P,3) -> rgb (P,3), sigma (P).
 * nerfb_mip_march: MipRayMarcher2.run_forward (ray_marcher.py:25-57) on colors (R,S,3), densities (R,S), depths
 *   (R,S) -> rgb (R,3), depth (R) [clamped to the call-global depth range], weights (R,S-1) (may be null).
 *   workspace: >= 16 bytes of device memory.
 * nerfb_triplane_render: ImportanceRenderer.forward (renderer.py:94-142) fused, one warp per ray: stratified depths
 *   (:172-195; numeric ray_start/ray_end, or per-ray arrays as produced by math_utils.get_ray_limits_box), coarse
 *   field + march, smoothed-weight importance sampling (:197-256), fine field, sorted merge (:160-170), final march.
 *   rand_coarse (R,S) / u (R,Si) are the reference's torch.rand_like / torch.rand draws; pass null to draw from the
 *   in-kernel counter-based generator seeded by `seed`.  Outputs: rgb (R,3), depth (R), opacity = sum of weights (R),
 *   coarse and final; depths_fine_out (R,Si) optional.  workspace: >= 16 bytes. */
int nerfb_triplane_relayout(const float* planes, int channels, int H, int W, float* planes_cl, void* stream);
int nerfb_triplane_field(const float* planes_cl, int H, int W, const float* decoder, float lr_mul,
                         const float* coords, int64_t n_points, float box_warp, float* rgb, float* sigma,
                         void* stream);
int nerfb_mip_march(const float* colors, const float* densities, const float* depths, int64_t n_rays,
                    int n_samples, int white_back, float* rgb, float* depth, float* weights, void* workspace,
                    void* stream);
int nerfb_triplane_render(const float* planes_cl, int H, int W, const float* decoder, float lr_mul,
                          const float* rays_o, const float* rays_d, int64_t n_rays, int n_samples,
                          int n_importance, double ray_start, double ray_end, const float* ray_start_arr,
                          const float* ray_end_arr, int disparity, float box_warp, int white_back,
                          const float* rand_coarse, const float* u, uint64_t seed, float* rgb_coarse,
                          float* depth_coarse, float* opacity_coarse, float* rgb_final, float* depth_final,
                          float* opacity_final, float* depths_fine_out, void* workspace, void* stream);

/* ---- (6b) EG3D training: backward of the field and of the marcher ------------------------------------------------
 * The reference trains through F.grid_sample + OSGDecoder + MipRayMarcher2 with torch autograd (system.py:22,75,111-120 ->
 * eg3d_training/eg3d_renderer.py:47-60 -> volumetric_rendering/renderer.py:94-142); importance depths are detached there
 * (renderer.py:215), so a training step is field(all samples) -> sort by depth -> marcher and these two entries are its
 * backward.  Both recompute their forward (nothing is saved) and ACCUMULATE into their outputs where noted.
 * nerfb_triplane_field_bwd: d_rgb (P,3), d_sigma (P) -> d_planes_cl (3,H,W,32) += (fp32 atomics; zero it first),
 *   d_decoder (2372 floats, layout of `decoder`: W1 | b1 | W2 | b2, raw parameters) +=.
 * nerfb_mip_march_bwd: d_rgb (R,3) / d_depth (R) / d_weights (R,S-1), each may be null -> d_colors (R,S,3), d_densities
 *   (R,S) (overwritten).  n_samples <= 512.  Depths are not differentiated.
 * nerfb_triplane_render_train: nerfb_triplane_render (same arguments, same six outputs, same single launch) that also
 *   writes samples_out (R, 5, S + Si) = depth | sigma | r | g | b of every sample, the S stratified samples first, then the
 *   Si importance samples, unsorted -- all a training step has to keep between its forward and its backward (coarse
 *   marcher backward on [0, S), sort by depth + final marcher backward on all, nerfb_triplane_field_bwd on the points
 *   rays_o + depth * rays_d).
 * nerfb_triplane_samples_bwd: exactly that middle part in one launch, one warp per ray: from samples (R, 5, S + Si) and the
 *   cotangents of the six outputs (rgb (R,3), depth (R), opacity (R), coarse and final; each may be null) -> per sample in
 *   the kept order: coords (R*(S+Si), 3) = rays_o + depth * rays_d (the forward's rounding), d_rgb (R*(S+Si), 3), d_sigma
 *   (R*(S+Si)): the inputs of nerfb_triplane_field_bwd.  The union is sorted by (depth, index) as the forward does. */
int nerfb_triplane_field_bwd(const float* planes_cl, int H, int W, const float* decoder, float lr_mul,
                             const float* coords, int64_t n_points, float box_warp, const float* d_rgb,
                             const float* d_sigma, float* d_planes_cl, float* d_decoder, void* stream);
int nerfb_mip_march_bwd(const float* colors, const float* densities, const float* depths, int64_t n_rays,
                        int n_samples, int white_back, const float* d_rgb, const float* d_depth,
                        const float* d_weights, float* d_colors, float* d_densities, void* stream);
int nerfb_triplane_samples_bwd(const float* samples, const float* rays_o, const float* rays_d, int64_t n_rays, int n_samples,
                               int n_importance, int white_back, const float* d_rgb_coarse, const float* d_depth_coarse,
                               const float* d_opacity_coarse, const float* d_rgb_final, const float* d_depth_final,
                               const float* d_opacity_final, float* coords, float* d_rgb, float* d_sigma, void* stream);
int nerfb_triplane_render_train(const float* planes_cl, int H, int W, const float* decoder, float lr_mul,
                                const float* rays_o, const float* rays_d, int64_t n_rays, int n_samples,
                                int n_importance, double ray_start, double ray_end, const float* ray_start_arr,
                                const float* ray_end_arr, int disparity, float box_warp, int white_back,
                                const float* rand_coarse, const float* u, uint64_t seed, float* rgb_coarse,
                                float* depth_coarse, float* opacity_coarse, float* rgb_final, float* depth_final,
                                float* opacity_final, float* samples_out, void* workspace, void* stream);

/* ---- the whole inference render behind ONE call (SURVEY 8b `nerfb_render_fused_fwd`, 7 step 4) ----------------
 * Replaces models/rendering.py:193-262 for inference (eval.py:85-96 with test_time=True, system.py:207-216 under no_grad):
 * sampling (:206-222) -> coarse embedding + MLP (:129-159) -> compositing (:161-190) -> sample_pdf + sort (:241-247) ->
 * fine embedding + MLP -> compositing, queued back to back on `stream`.  On the bf16 tensor-core arm, when n_samples
 * (resp. n_samples + n_importance) is 32, 64 or 128 and the model has no class head, compositing runs in the MLP kernel's
 * epilogue: per-sample [rgb sigma] never reaches HBM and the pass is one launch; other shapes use kernels (1) + (4).
 * cfg.test_time: the coarse pass is sigma-only and only opacity_coarse is produced (:230-231).  Random inputs (perturb_rand,
 * u when perturb > 0; noise_* when noise_std != 0) are supplied by the caller in the reference's draw order.
 * Outputs that the configuration does not produce may be NULL.  Nothing is kept for a backward pass. */
typedef struct nerfb_render_cfg {
  int32_t precision;         /* NERFB_PREC_* */
  int32_t n_samples, n_importance;
  int32_t use_disp, white_back, test_time;
  float perturb, noise_std;
} nerfb_render_cfg;
typedef struct nerfb_render_io {
  const float* rays;          /* (N,8) [o d near far] */
  const float* perturb_rand;  /* (N,S)      U[0,1), perturb > 0 */
  const float* noise_coarse;  /* (N,S)      N(0,1), noise_std != 0 */
  const float* u;             /* (N,S_i)    U[0,1), perturb > 0 and n_importance > 0 */
  const float* noise_fine;    /* (N,S+S_i)  N(0,1), noise_std != 0 and n_importance > 0 */
  float* rgb_coarse; float* depth_coarse; float* opacity_coarse; float* cls_coarse;   /* (N,3) (N) (N) (N,CLS) */
  float* rgb_fine; float* depth_fine; float* opacity_fine; float* cls_fine;
} nerfb_render_io;
int64_t nerfb_render_fused_workspace_bytes(const nerfb_mlp_desc* coarse, const nerfb_mlp_desc* fine,
                                           const nerfb_render_cfg* cfg, int64_t n_rays);
int nerfb_render_fused_fwd(const nerfb_mlp_desc* coarse, const nerfb_mlp_desc* fine, const nerfb_render_cfg* cfg,
                           const float* params_coarse, const float* film_coarse, const float* params_fine,
                           const float* film_fine, const nerfb_render_io* io, int64_t n_rays, void* workspace,
                           int64_t workspace_bytes, void* stream);

/* ---- the steps either side of the path ----------------------------------
 * nerfb_generate_rays: datasets/ray_utils.py:5-50 (get_ray_directions, get_rays) [+ :53-94 get_ndc_rays when ndc != 0,
 *   near plane 1.0 and near/far = 0/1 as datasets/llff.py:234-239] -> rays (H*W, 8) = [o d near far], row-major pixels
 *   (datasets/blender.py:103-108).  c2w: 12 floats in HOST memory (row-major 3x4 camera-to-world; passed by value to the
 *   kernel, the only host pointer of this ABI).
 * nerfb_finish_image: eval.py:163-193 on the device: rgb (H*W,3) -> uint8 (H*W,3) [(img*255).astype(uint8)], depth (H*W)
 *   -> nan_to_num + rows flipped = the PFM payload of datasets/depth_utils.py:43-69, and sq_err = sum (rgb - gt)^2 (one
 *   float64; metrics.py:4-14 psnr = -10 log10(sq_err / (3 H W))).  Any of rgb_u8 / depth+depth_pfm / gt+sq_err may be null. */
int nerfb_generate_rays(const float* c2w_host, int H, int W, double focal, float near, float far, int ndc,
                        float* rays, void* stream);
int nerfb_finish_image(const float* rgb, const float* depth, const float* gt, int H, int W, uint8_t* rgb_u8,
                       float* depth_pfm, double* sq_err, void* stream);

/* ---- the sparse `.vol` volume of extract_mesh.ipynb cell 7 ---------------------
 * rgbsigma (n_voxels, 4) fp32 = the dense-grid output of nerfb_mlp_fwd (NERFB_IN_GRID, sigma_only = 0), 16-byte aligned;
 * neg_step = (float)(-(xmax - xmin) / N).  Writes uint32 pairs (voxel index, r<<24 | g<<16 | b<<8 | alpha) for every voxel
 * with alpha = 1 - exp(neg_step * max(sigma, 0)) > 0, in index order, up to `capacity` pairs; *n_valid (device int64) =
 * the number of such voxels (call with capacity 0 to size the buffer).  n_voxels < 2^32 (the file format's index width). */
int64_t nerfb_pack_vol_workspace_bytes(int64_t n_voxels);
int nerfb_pack_vol(const float* rgbsigma, int64_t n_voxels, float neg_step, uint32_t* pairs, int64_t capacity,
                   int64_t* n_valid, void* workspace, int64_t workspace_bytes, void* stream);

/* ---- nerfb_pack_vol for callers that cannot allocate ------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same inputs, outputs and results as
 * nerfb_pack_vol, but the per-chunk counts / offsets live in a file-scope `static __device__` array of the library
 * (192 KiB, held by every GPU that loads the library until exit) instead of a caller-provided workspace; the entry
 * allocates nothing (no cudaMalloc / cudaFree / allocating library call).  Limit: n_voxels <=
 * nerfb_pack_vol_static_max_voxels() = 2^24 (a 256^3 grid); larger inputs return NERFB_E_WORKSPACE with a message on
 * stderr -- use nerfb_pack_vol then.  All calls on a device share the array (it keeps the last call's contents; nothing
 * in it has to be cleared between calls): calls that OVERLAP on the GPU (different streams without ordering) race. */
int64_t nerfb_pack_vol_static_max_voxels(void);
int nerfb_pack_vol_static(const float* rgbsigma, int64_t n_voxels, float neg_step, uint32_t* pairs, int64_t capacity,
                          int64_t* n_valid, void* stream);

/* ---- diagnostics: how many kernels this library has launched in this process ------------------- */
int64_t nerfb_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* NERF_B200_H */
