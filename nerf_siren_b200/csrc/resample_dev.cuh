// Warp-level inverse-CDF helpers shared by kernel (3) (resample.cu) and the EG3D renderer (triplane.cu).
#pragma once
#include "common.cuh"

// ---- inverse-CDF sampling for one ray, executed by one warp -------------------------------------------
// s_bins[nb], s_w[nb-1] filled by the caller; s_cdf[nb] produced here.  Returns samples through `emit`.
__device__ __forceinline__ void warp_build_cdf(const float* s_w, float* s_cdf, int nb, float eps, int lane) {
  const int nw = nb - 1;
  float part = 0.f;
  for (int i = lane; i < nw; i += 32) part += s_w[i] + eps;                      // rendering.py:37
  const float total = warp_sum(part);
  // inclusive scan of pdf in blocks of 32 with a running carry (rendering.py:38-41)
  float carry = 0.f;
  if (lane == 0) s_cdf[0] = 0.f;
  for (int base = 0; base < nw; base += 32) {
    const int i = base + lane;
    float p = (i < nw) ? (s_w[i] + eps) / total : 0.f;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const float q = __shfl_up_sync(0xffffffffu, p, o);
      if (lane >= o) p += q;
    }
    if (i < nw) s_cdf[i + 1] = carry + p;
    carry += __shfl_sync(0xffffffffu, p, 31);
  }
  __syncwarp();
}

__device__ __forceinline__ float invert_cdf(const float* s_bins, const float* s_cdf, int nb, float u, float eps,
                                            int* ind_out) {
  // searchsorted(cdf, u, right=True) = #(cdf <= u).  The cdf is non-decreasing (a cumulative sum of positive terms), so the count is
  // the largest lo with cdf[lo - 1] <= u: built bit by bit, branch-free, ceil(log2(nb + 1)) probes for every lane (the bisection
  // loop had data-dependent trip counts and two compares per probe).
  int lo = 0;
  for (int step = 1 << (31 - __clz(nb)); step > 0; step >>= 1) {      // warp-uniform trip count; nb >= 1
    const int probe = lo + step;
    if (probe <= nb && s_cdf[probe - 1] <= u) lo = probe;
  }
  if (ind_out) *ind_out = lo;
  const int below = max(lo - 1, 0);                      // rendering.py:55-56
  const int above = min(lo, nb - 1);
  const float c0 = s_cdf[below], c1 = s_cdf[above];
  const float b0 = s_bins[below], b1 = s_bins[above];
  float denom = __fsub_rn(c1, c0);
  if (denom < eps) denom = 1.0f;                         // rendering.py:63
  // bins_g0 + (u - cdf_g0) / denom * (bins_g1 - bins_g0), left to right (rendering.py:66)
  return __fadd_rn(b0, __fmul_rn(__fdiv_rn(__fsub_rn(u, c0), denom), __fsub_rn(b1, b0)));
}

