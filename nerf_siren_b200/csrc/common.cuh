// Shared helpers for the sm_100a kernels behind include/nerf_b200.h.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <atomic>
#include "../../include/nerf_b200.h"

#ifndef NERFB_SM_COUNT
#define NERFB_SM_COUNT 148          // B200: 2 dies x 74 SMs; grids are sized in multiples of this
#endif

extern std::atomic<long long> g_nerfb_launches;

// Every launch goes through this: counts it (bench.py's gpu_launches) and returns the launch status.
#define NERFB_LAUNCH(kernel, grid, block, smem, stream, ...)                              \
  do {                                                                                    \
    kernel<<<(grid), (block), (smem), (cudaStream_t)(stream)>>>(__VA_ARGS__);             \
    g_nerfb_launches.fetch_add(1, std::memory_order_relaxed);                             \
    cudaError_t _e = cudaPeekAtLastError();                                               \
    if (_e != cudaSuccess) { cudaGetLastError(); return (int)_e; }                        \
  } while (0)

#define NERFB_CHECK_ARG(cond) do { if (!(cond)) return NERFB_E_BADARG; } while (0)
#define NERFB_CUDA(call) do { cudaError_t _e = (call); if (_e != cudaSuccess) return (int)_e; } while (0)

static inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline int64_t round_up64(int64_t a, int64_t b) { return ceil_div64(a, b) * b; }

// torch.linspace(0, 1, n)[i] in fp32 (verified bit-exact against torch CPU): first half
// i*step, second half a single-rounding fma from the right end.
__host__ __device__ __forceinline__ float linspace01(int i, int n) {
  if (n <= 1) return 0.f;
  const float step = 1.0f / (float)(n - 1);
#ifdef __CUDA_ARCH__
  return (i < n / 2) ? __fmul_rn((float)i, step) : __fmaf_rn(-(float)(n - 1 - i), step, 1.0f);
#else
  return (i < n / 2) ? (float)i * step : fmaf(-(float)(n - 1 - i), step, 1.0f);
#endif
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Layout of the flat fp32 parameter blob == the reference's state_dict order.
struct NerfOffsets {
  int64_t w[16], b[16];   // trunk layers
  int in[16];             // fan-in of trunk layer i
  int64_t w_final, b_final, w_dir, b_dir, w_sigma, b_sigma, w_rgb, b_rgb, w_p0, b_p0, w_p1, b_p1;
  int64_t total;
  int in_xyz, in_dir;
};

__host__ __device__ inline int nerfb_in_xyz(const nerfb_mlp_desc& d) {
  return d.in_xyz > 0 ? d.in_xyz : 3 * (1 + 2 * d.n_freq_xyz);
}
__host__ __device__ inline int nerfb_in_dir(const nerfb_mlp_desc& d) {
  return d.in_dir > 0 ? d.in_dir : 3 * (1 + 2 * d.n_freq_dir);
}

inline NerfOffsets nerf_offsets(const nerfb_mlp_desc& d) {
  NerfOffsets o{};
  const int W = d.W;
  o.in_xyz = nerfb_in_xyz(d);
  o.in_dir = nerfb_in_dir(d);
  int64_t p = 0;
  if (d.kind == NERFB_KIND_NERF) {
    for (int i = 0; i < d.D; ++i) {
      o.in[i] = (i == 0) ? o.in_xyz : (((d.skip_mask >> i) & 1) ? W + o.in_xyz : W);
      o.w[i] = p; p += (int64_t)W * o.in[i];
      o.b[i] = p; p += W;
    }
    o.w_final = p; p += (int64_t)W * W;           o.b_final = p; p += W;
    o.w_dir = p;   p += (int64_t)(W / 2) * (W + o.in_dir); o.b_dir = p; p += W / 2;
    o.w_sigma = p; p += W;                         o.b_sigma = p; p += 1;
    o.w_rgb = p;   p += 3 * (W / 2);               o.b_rgb = p; p += 3;
    if (d.n_cls > 0) {
      o.w_p0 = p; p += (int64_t)2 * W * W;         o.b_p0 = p; p += 2 * W;
      o.w_p1 = p; p += (int64_t)d.n_cls * 2 * W;   o.b_p1 = p; p += d.n_cls;
    }
  } else {  // SIREN: network.{0..D-1}.layer, final_layer, color_layer_sine.layer, color_layer_linear.0
    for (int i = 0; i < d.D; ++i) {
      o.in[i] = (i == 0) ? 3 : W;
      o.w[i] = p; p += (int64_t)W * o.in[i];
      o.b[i] = p; p += W;
    }
    o.w_sigma = p; p += W;                 o.b_sigma = p; p += 1;
    o.w_dir = p;   p += (int64_t)W * (W + 3); o.b_dir = p; p += W;     // color_layer_sine: in = [dir(3) | h(W)]
    o.w_rgb = p;   p += 3 * W;             o.b_rgb = p; p += 3;
  }
  o.total = p;
  return o;
}
