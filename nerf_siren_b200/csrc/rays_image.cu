// The steps either side of the render_rays path:
//   nerfb_generate_rays -- datasets/ray_utils.py:5-50 (get_ray_directions + get_rays), :53-94 (get_ndc_rays) and the
//     (H*W, 8) ray rows assembled in datasets/blender.py:103-108 / datasets/llff.py:234-248,294-306, computed on the
//     device so a full-image render needs no host ray tensor (and no kornia.create_meshgrid).
//   nerfb_finish_image  -- eval.py:163-193: uint8 quantisation of the rendered rgb, nan_to_num + vertical flip of the
//     depth map into the PFM payload order (datasets/depth_utils.py:43-69 writes np.flipud(image)), and the squared-error
//     sum behind metrics.psnr (metrics.py:4-14).
// Both are pure streaming kernels (HBM-bound): 32 B per ray out; 12 + 4 (+12) B in, 3 + 4 B out per pixel.
#include "common.cuh"

namespace {

struct RayGenArgs {
  float c2w[12];              // row-major 3 x 4 camera-to-world
  int H, W; float focal, half_w, half_h;
  float near, far; int ndc;
  float ndc_near; float sx, sy;        // NDC: near plane depth, -1/(W/(2 focal)), -1/(H/(2 focal))
};

__global__ void __launch_bounds__(256) generate_rays_kernel(const RayGenArgs a, float* __restrict__ rays) {
  const int64_t total = (int64_t)a.H * a.W;
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < total; p += (int64_t)gridDim.x * blockDim.x) {
    const int j = (int)(p / a.W), i = (int)(p - (int64_t)j * a.W);       // row j, column i (create_meshgrid: x = column)
    // ray_utils.py:21-22: ((i - W/2)/focal, -(j - H/2)/focal, -1), no half-pixel centring
    const float dx = __fdiv_rn(__fsub_rn((float)i, a.half_w), a.focal);
    const float dy = -__fdiv_rn(__fsub_rn((float)j, a.half_h), a.focal);
    const float dz = -1.f;
    // ray_utils.py:42: directions @ c2w[:, :3].T, then / norm (:43)
    float d[3];
#pragma unroll
    for (int k = 0; k < 3; ++k)
      d[k] = __fmaf_rn(dz, a.c2w[4 * k + 2], __fmaf_rn(dy, a.c2w[4 * k + 1], __fmul_rn(dx, a.c2w[4 * k])));
    const float nrm = __fsqrt_rn(__fmaf_rn(d[2], d[2], __fmaf_rn(d[1], d[1], __fmul_rn(d[0], d[0]))));
#pragma unroll
    for (int k = 0; k < 3; ++k) d[k] = __fdiv_rn(d[k], nrm);
    float o[3] = {a.c2w[3], a.c2w[7], a.c2w[11]};
    float near = a.near, far = a.far;
    if (a.ndc) {                                                        // ray_utils.py:74-94
      const float t = -__fdiv_rn(__fadd_rn(a.ndc_near, o[2]), d[2]);
#pragma unroll
      for (int k = 0; k < 3; ++k) o[k] = __fadd_rn(o[k], __fmul_rn(t, d[k]));
      const float ox_oz = __fdiv_rn(o[0], o[2]), oy_oz = __fdiv_rn(o[1], o[2]);
      const float o0 = __fmul_rn(a.sx, ox_oz), o1 = __fmul_rn(a.sy, oy_oz);
      const float o2 = __fadd_rn(1.f, __fdiv_rn(__fmul_rn(2.f, a.ndc_near), o[2]));
      const float d0 = __fmul_rn(a.sx, __fsub_rn(__fdiv_rn(d[0], d[2]), ox_oz));
      const float d1 = __fmul_rn(a.sy, __fsub_rn(__fdiv_rn(d[1], d[2]), oy_oz));
      const float d2 = __fsub_rn(1.f, o2);
      o[0] = o0; o[1] = o1; o[2] = o2; d[0] = d0; d[1] = d1; d[2] = d2;
    }
    float4* dst = reinterpret_cast<float4*>(rays + p * 8);
    dst[0] = make_float4(o[0], o[1], o[2], d[0]);
    dst[1] = make_float4(d[1], d[2], near, far);
  }
}

// rgb (N,3) -> uint8 (N,3) as numpy's (img * 255).astype(np.uint8) (truncation; eval.py:187); depth (H,W) -> nan_to_num,
// rows flipped (PFM payload); sum of squared rgb error against gt (metrics.py:4-10) accumulated in double.
__global__ void __launch_bounds__(256) finish_image_kernel(const float* __restrict__ rgb, const float* __restrict__ depth,
                                                           const float* __restrict__ gt, int H, int W, uint8_t* __restrict__ rgb_u8,
                                                           float* __restrict__ depth_pfm, double* __restrict__ sq_err) {
  const int64_t n = (int64_t)H * W;
  double acc = 0.0;
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += (int64_t)gridDim.x * blockDim.x) {
    const float r = __ldg(rgb + p * 3), g = __ldg(rgb + p * 3 + 1), b = __ldg(rgb + p * 3 + 2);
    if (rgb_u8) {
      // float -> uint8 of a value outside [0, 256) is undefined in C / numpy; the renderer's rgb is in [0, 1]: clamp
      rgb_u8[p * 3] = (uint8_t)fminf(fmaxf(__fmul_rn(r, 255.f), 0.f), 255.f);
      rgb_u8[p * 3 + 1] = (uint8_t)fminf(fmaxf(__fmul_rn(g, 255.f), 0.f), 255.f);
      rgb_u8[p * 3 + 2] = (uint8_t)fminf(fmaxf(__fmul_rn(b, 255.f), 0.f), 255.f);
    }
    if (depth && depth_pfm) {
      float v = __ldg(depth + p);
      if (v != v) v = 0.f;                                              // np.nan_to_num: nan -> 0, +-inf -> +-max
      else if (isinf(v)) v = v > 0.f ? 3.4028234663852886e38f : -3.4028234663852886e38f;
      const int64_t row = p / W, col = p - row * W;
      depth_pfm[(int64_t)(H - 1 - row) * W + col] = v;
    }
    if (gt) {
      const float e0 = r - __ldg(gt + p * 3), e1 = g - __ldg(gt + p * 3 + 1), e2 = b - __ldg(gt + p * 3 + 2);
      acc += (double)(e0 * e0) + (double)(e1 * e1) + (double)(e2 * e2);
    }
  }
  if (gt && sq_err) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(sq_err, acc);
  }
}

__global__ void zero_double_kernel(double* p) { *p = 0.0; }

}  // namespace

extern "C" int nerfb_generate_rays(const float* c2w_host, int H, int W, double focal, float near, float far, int ndc,
                                   float* rays, void* stream) {
  NERFB_CHECK_ARG(c2w_host && rays && H >= 1 && W >= 1 && focal > 0.0);
  RayGenArgs a{};
  for (int i = 0; i < 12; ++i) a.c2w[i] = c2w_host[i];
  a.H = H; a.W = W; a.focal = (float)focal; a.half_w = (float)(W / 2.0); a.half_h = (float)(H / 2.0);
  a.near = near; a.far = far; a.ndc = ndc; a.ndc_near = 1.0f;          // llff.py:237: near plane always at 1.0
  a.sx = (float)(-1.0 / (W / (2.0 * focal))); a.sy = (float)(-1.0 / (H / (2.0 * focal)));
  if (ndc) { a.near = 0.f; a.far = 1.f; }                              // llff.py:236: NDC near / far are always 0 and 1
  const int64_t want = ceil_div64((int64_t)H * W, 256);
  const int grid = (int)(want < NERFB_SM_COUNT * 8 ? want : NERFB_SM_COUNT * 8);
  NERFB_LAUNCH(generate_rays_kernel, grid, 256, 0, stream, a, rays);
  return NERFB_OK;
}

extern "C" int nerfb_finish_image(const float* rgb, const float* depth, const float* gt, int H, int W, uint8_t* rgb_u8,
                                  float* depth_pfm, double* sq_err, void* stream) {
  NERFB_CHECK_ARG(rgb && H >= 1 && W >= 1 && (!gt || sq_err) && (!depth_pfm || depth));
  if (gt) NERFB_LAUNCH(zero_double_kernel, 1, 1, 0, stream, sq_err);
  const int64_t want = ceil_div64((int64_t)H * W, 256);
  const int grid = (int)(want < NERFB_SM_COUNT * 8 ? want : NERFB_SM_COUNT * 8);
  NERFB_LAUNCH(finish_image_kernel, grid, 256, 0, stream, rgb, depth, gt, H, W, rgb_u8, depth_pfm, sq_err);
  return NERFB_OK;
}
