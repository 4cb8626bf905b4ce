// The steps either side of the render_rays path:
//   nerfb_generate_rays -- datasets/ray_utils.py:5-50 (get_ray_directions + get_rays), :53-94 (get_ndc_rays) and the
//     (H*W, 8) ray rows assembled in datasets/blender.py:103-108 / datasets/llff.py:234-248,294-306, computed on the
//     device so a full-image render needs no host ray tensor (and no kornia.create_meshgrid).
//   nerfb_finish_image  -- eval.py:163-193: uint8 quantisation of the rendered rgb, nan_to_num + vertical flip of the
//     depth map into the PFM payload order (datasets/depth_utils.py:43-69 writes np.flipud(image)), and the squared-error
//     sum behind metrics.psnr (metrics.py:4-14).
//   nerfb_pack_vol      -- extract_mesh.ipynb cell 7: the sparse `.vol` volume Unity renders, 8 bytes per
//     voxel with alpha > 0, from the dense (N^3, 4) [rgb sigma] query of kernel (1).
// All are pure streaming kernels (HBM-bound): 32 B per ray out; 12 + 4 (+12) B in, 3 + 4 B out per pixel; 16 B in per voxel.
#include <cstdio>
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

// ---- .vol packing (extract_mesh.ipynb cell 7) --------------------------------------------------------------------
//   a = 1 - np.exp(-(xmax - xmin) / N * sigma)       float32 throughout (the Python scalar is weak: cast to float32)
//   rgb = (rgbsigma[:, :3] * 255).astype(np.uint32);  i = np.where(a > 0)[0]
//   s = rgb . [1<<24, 1<<16, 1<<8] + (a * 255).astype(np.uint32);  file = uint32 pairs (i, s) in index order
// sigma is clamped at 0 first (cell 4: np.maximum(sigma, 0)).  exp is evaluated in double and rounded once (numpy's float32
// exp is within 1 ulp of that; the alpha byte can differ by one where a * 255 sits on an integer).
constexpr int VOL_BLOCK = 256, VOL_PER_THREAD = 4, VOL_CHUNK = VOL_BLOCK * VOL_PER_THREAD;

__device__ __forceinline__ bool vol_voxel(const float4 v, float k, uint32_t* packed) {
  const float sg = fmaxf(v.w, 0.f);
  const float a = __fsub_rn(1.f, (float)exp((double)__fmul_rn(k, sg)));
  const uint32_t r = (uint32_t)__fmul_rn(v.x, 255.f), g = (uint32_t)__fmul_rn(v.y, 255.f), b = (uint32_t)__fmul_rn(v.z, 255.f);
  *packed = (r << 24) + (g << 16) + (b << 8) + (uint32_t)__fmul_rn(a, 255.f);
  return a > 0.f;
}

// pass 1: number of voxels with alpha > 0 in each chunk of 1024
__global__ void __launch_bounds__(VOL_BLOCK) vol_count_kernel(const float4* __restrict__ rgbsigma, int64_t n, float k,
                                                              uint32_t* __restrict__ chunk_count) {
  __shared__ uint32_t warp_cnt[VOL_BLOCK / 32];
  for (int64_t c = blockIdx.x; c * VOL_CHUNK < n; c += gridDim.x) {
    uint32_t cnt = 0;
#pragma unroll
    for (int t = 0; t < VOL_PER_THREAD; ++t) {
      const int64_t i = c * VOL_CHUNK + t * VOL_BLOCK + threadIdx.x;
      uint32_t pk;
      if (i < n && vol_voxel(__ldg(rgbsigma + i), k, &pk)) ++cnt;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    if ((threadIdx.x & 31) == 0) warp_cnt[threadIdx.x >> 5] = cnt;
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t tot = 0;
#pragma unroll
      for (int w = 0; w < VOL_BLOCK / 32; ++w) tot += warp_cnt[w];
      chunk_count[c] = tot;
    }
    __syncthreads();
  }
}

// pass 2 (one CTA): exclusive prefix sum of the chunk counts into int64 offsets; total -> *n_valid
__global__ void __launch_bounds__(1024) vol_scan_kernel(const uint32_t* __restrict__ chunk_count, int64_t n_chunks,
                                                        int64_t* __restrict__ chunk_off, int64_t* __restrict__ n_valid) {
  __shared__ int64_t warp_sum[32];
  __shared__ int64_t carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int64_t base = 0; base < n_chunks; base += 1024) {
    const int64_t i = base + threadIdx.x;
    const int64_t v = i < n_chunks ? (int64_t)chunk_count[i] : 0;
    int64_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int64_t t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
    if (lane == 31) warp_sum[w] = inc;
    __syncthreads();
    if (w == 0) {
      int64_t ws = warp_sum[lane], wi = ws;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const int64_t t = __shfl_up_sync(0xffffffffu, wi, o); if (lane >= o) wi += t; }
      warp_sum[lane] = wi - ws;                     // exclusive over warps
    }
    __syncthreads();
    const int64_t c0 = carry;
    if (i < n_chunks) chunk_off[i] = c0 + warp_sum[w] + inc - v;
    __syncthreads();
    if (threadIdx.x == 1023) carry = c0 + warp_sum[w] + inc;
    __syncthreads();
  }
  if (threadIdx.x == 0) *n_valid = carry;
}

// pass 3: ordered write of the (index, packed rgba) pairs
__global__ void __launch_bounds__(VOL_BLOCK) vol_write_kernel(const float4* __restrict__ rgbsigma, int64_t n, float k,
                                                              const int64_t* __restrict__ chunk_off, int64_t capacity,
                                                              uint2* __restrict__ pairs) {
  __shared__ uint32_t warp_cnt[VOL_PER_THREAD][VOL_BLOCK / 32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int64_t c = blockIdx.x; c * VOL_CHUNK < n; c += gridDim.x) {
    uint32_t pk[VOL_PER_THREAD], before[VOL_PER_THREAD];
    bool ok[VOL_PER_THREAD];
#pragma unroll
    for (int t = 0; t < VOL_PER_THREAD; ++t) {
      const int64_t i = c * VOL_CHUNK + t * VOL_BLOCK + threadIdx.x;
      ok[t] = i < n && vol_voxel(__ldg(rgbsigma + i), k, &pk[t]);
      const uint32_t m = __ballot_sync(0xffffffffu, ok[t]);
      before[t] = __popc(m & ((1u << lane) - 1u));
      if (lane == 0) warp_cnt[t][w] = __popc(m);
    }
    __syncthreads();
    int64_t off = chunk_off[c];
#pragma unroll
    for (int t = 0; t < VOL_PER_THREAD; ++t) {
      uint32_t pre = 0, row = 0;
#pragma unroll
      for (int x = 0; x < VOL_BLOCK / 32; ++x) { const uint32_t v = warp_cnt[t][x]; if (x < w) pre += v; row += v; }
      const int64_t dst = off + pre + before[t];
      if (ok[t] && dst < capacity) pairs[dst] = make_uint2((uint32_t)(c * VOL_CHUNK + t * VOL_BLOCK + threadIdx.x), pk[t]);
      off += row;
    }
    __syncthreads();
  }
}

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

extern "C" int64_t nerfb_pack_vol_workspace_bytes(int64_t n_voxels) {
  const int64_t n_chunks = ceil_div64(n_voxels > 0 ? n_voxels : 1, VOL_CHUNK);
  return ((n_chunks * 4 + 255) & ~(int64_t)255) + n_chunks * 8;
}

extern "C" int nerfb_pack_vol(const float* rgbsigma, int64_t n_voxels, float neg_step, uint32_t* pairs, int64_t capacity,
                              int64_t* n_valid, void* workspace, int64_t workspace_bytes, void* stream) {
  NERFB_CHECK_ARG(rgbsigma && n_valid && n_voxels >= 0 && n_voxels < ((int64_t)1 << 32) && capacity >= 0 && (pairs || capacity == 0) &&
                  workspace && workspace_bytes >= nerfb_pack_vol_workspace_bytes(n_voxels));
  NERFB_CHECK_ARG((reinterpret_cast<uintptr_t>(rgbsigma) & 15) == 0 && (reinterpret_cast<uintptr_t>(workspace) & 255) == 0 &&
                  (reinterpret_cast<uintptr_t>(pairs) & 7) == 0);
  // an empty volume has ZERO chunks: the scan then writes *n_valid = 0 instead of summing a count no kernel of this call wrote
  const int64_t n_chunks = ceil_div64(n_voxels, VOL_CHUNK);
  const int64_t n_chunks_ws = n_chunks > 0 ? n_chunks : 1;      // the workspace layout is sized for >= 1 chunk
  uint32_t* cnt = static_cast<uint32_t*>(workspace);
  int64_t* off = reinterpret_cast<int64_t*>(static_cast<char*>(workspace) + ((n_chunks_ws * 4 + 255) & ~(int64_t)255));
  const float4* v = reinterpret_cast<const float4*>(rgbsigma);
  const int grid = (int)(n_chunks < NERFB_SM_COUNT * 8 ? n_chunks : NERFB_SM_COUNT * 8);
  if (n_chunks > 0) NERFB_LAUNCH(vol_count_kernel, grid, VOL_BLOCK, 0, stream, v, n_voxels, neg_step, cnt);
  NERFB_LAUNCH(vol_scan_kernel, 1, 1024, 0, stream, cnt, n_chunks, off, n_valid);
  if (capacity > 0 && n_chunks > 0) NERFB_LAUNCH(vol_write_kernel, grid, VOL_BLOCK, 0, stream, v, n_voxels, neg_step, off, capacity, reinterpret_cast<uint2*>(pairs));
  return NERFB_OK;
}

// ---- nerfb_pack_vol without a caller-provided workspace (no counterpart in the reference; added
// for callers that cannot allocate: CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc) ----
// The per-chunk counts and offsets live in a file-scope device array of fixed size: 2^24 voxels (a 256^3 grid) = 16,384 chunks
// of 1024 voxels = 64 KiB of counts + 128 KiB of offsets = 192 KiB, held by every GPU that loads the library until exit.
// In host code `g_vol_scratch` is only a host-side stand-in for the symbol: its DEVICE address is looked up once per device
// (with that device current) by cudaGetSymbolAddress and that pointer is what the kernels get.  Every call on a device shares
// the array and finds the previous call's contents in it; nothing has to start at zero (every word a kernel reads is written
// earlier in the same call: vol_count_kernel writes all n_chunks counts, vol_scan_kernel all n_chunks offsets), but two calls
// that overlap on the GPU (different streams, no ordering between them) race on it.
constexpr int64_t VOL_STATIC_MAX_VOXELS = (int64_t)1 << 24;
constexpr int64_t VOL_STATIC_MAX_CHUNKS = VOL_STATIC_MAX_VOXELS / VOL_CHUNK;
constexpr int64_t VOL_STATIC_CNT_BYTES = (VOL_STATIC_MAX_CHUNKS * 4 + 255) & ~(int64_t)255;
constexpr int64_t VOL_STATIC_BYTES = VOL_STATIC_CNT_BYTES + VOL_STATIC_MAX_CHUNKS * 8;
static __device__ __align__(256) unsigned char g_vol_scratch[VOL_STATIC_BYTES];

extern "C" int64_t nerfb_pack_vol_static_max_voxels(void) { return VOL_STATIC_MAX_VOXELS; }

extern "C" int nerfb_pack_vol_static(const float* rgbsigma, int64_t n_voxels, float neg_step, uint32_t* pairs, int64_t capacity,
                                     int64_t* n_valid, void* stream) {
  NERFB_CHECK_ARG(rgbsigma && n_valid && n_voxels >= 0 && capacity >= 0 && (pairs || capacity == 0));
  if (n_voxels > VOL_STATIC_MAX_VOXELS) {
    fprintf(stderr, "nerf_b200: nerfb_pack_vol_static: n_voxels = %lld exceeds the built-in scratch (max %lld voxels); use nerfb_pack_vol "
            "with a caller-provided workspace\n", (long long)n_voxels, (long long)VOL_STATIC_MAX_VOXELS);
    return NERFB_E_WORKSPACE;
  }
  NERFB_CHECK_ARG((reinterpret_cast<uintptr_t>(rgbsigma) & 15) == 0 && (reinterpret_cast<uintptr_t>(pairs) & 7) == 0);
  int dev = 0;
  NERFB_CUDA(cudaGetDevice(&dev));
  static unsigned char* scratch_dev[64] = {};          // device address of g_vol_scratch, per device
  unsigned char* scratch = (dev >= 0 && dev < 64) ? scratch_dev[dev] : nullptr;
  if (!scratch) {
    void* ptr = nullptr;
    NERFB_CUDA(cudaGetSymbolAddress(&ptr, g_vol_scratch));      // the current device's copy of the symbol
    if (!ptr) return (int)cudaErrorInvalidSymbol;
    scratch = static_cast<unsigned char*>(ptr);
    if (dev >= 0 && dev < 64) scratch_dev[dev] = scratch;
  }
  // an empty volume has ZERO chunks here: the scan then writes *n_valid = 0 without reading a count nobody wrote in this call
  // (the shared scratch still holds the previous call's counts)
  const int64_t n_chunks = ceil_div64(n_voxels, VOL_CHUNK);
  uint32_t* cnt = reinterpret_cast<uint32_t*>(scratch);
  int64_t* off = reinterpret_cast<int64_t*>(scratch + VOL_STATIC_CNT_BYTES);
  const float4* v = reinterpret_cast<const float4*>(rgbsigma);
  const int grid = (int)(n_chunks < NERFB_SM_COUNT * 8 ? n_chunks : NERFB_SM_COUNT * 8);
  if (n_chunks > 0) NERFB_LAUNCH(vol_count_kernel, grid, VOL_BLOCK, 0, stream, v, n_voxels, neg_step, cnt);
  NERFB_LAUNCH(vol_scan_kernel, 1, 1024, 0, stream, cnt, n_chunks, off, n_valid);
  if (capacity > 0 && n_chunks > 0) NERFB_LAUNCH(vol_write_kernel, grid, VOL_BLOCK, 0, stream, v, n_voxels, neg_step, off, capacity, reinterpret_cast<uint2*>(pairs));
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
