// Kernel (4): alpha compositing, forward and fused backward.
//   replaces models/rendering.py:161-190 (deltas, noise, alpha, cumprod transmittance, weights, rgb /
//   depth / opacity sums, white background) and the class-score compositing of :379.
// One warp per ray.  Samples are visited in chunks of 32 (lane = sample within the chunk, so every
// load/store of the (N,S[,C]) tensors is coalesced); the transmittance is a warp-shuffle product scan
// with a running carry across chunks.  HBM-bound: each input element is read once, each output written
// once (24 B/sample forward, 40 B/sample backward for C = 4 -- SURVEY 8d).
#include "common.cuh"

#define COMP_WARPS 4
#define COMP_MAX_EXTRA 28            // C <= 32

struct SampleTerms { float alpha, t, e, delta; bool on; };

// alpha_i and friends for sample i of ray r (models/rendering.py:162-175).
__device__ __forceinline__ SampleTerms sample_terms(float sigma, float noise, float noise_std, float z_i, float z_next,
                                                   bool last, float dnorm) {
  SampleTerms s;
  const float d = last ? 1e10f : __fsub_rn(z_next, z_i);
  s.delta = __fmul_rn(d, dnorm);
  const float pre = __fadd_rn(sigma, __fmul_rn(noise, noise_std));
  s.on = pre > 0.f;
  const float a = s.on ? pre : 0.f;
  s.e = expf(-__fmul_rn(s.delta, a));
  s.alpha = __fsub_rn(1.0f, s.e);
  s.t = __fadd_rn(__fsub_rn(1.0f, s.alpha), 1e-10f);
  return s;
}

__device__ __forceinline__ float warp_incl_prod(float p, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const float q = __shfl_up_sync(0xffffffffu, p, o);
    if (lane >= o) p *= q;
  }
  return p;
}

template <int CT>   // CT = 1 (sigma only), 4 (rgb sigma), 0 (generic C >= 4 with extra channels)
__global__ void __launch_bounds__(COMP_WARPS * 32)
composite_fwd_kernel(const float* __restrict__ raw, int C, const float* __restrict__ z, const float* __restrict__ rays,
                     const float* __restrict__ noise, float noise_std, int white_back, int64_t N, int S,
                     float* __restrict__ weights, float* __restrict__ rgb, float* __restrict__ depth,
                     float* __restrict__ opacity, float* __restrict__ extra) {
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = (int64_t)blockIdx.x * COMP_WARPS + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)gridDim.x * COMP_WARPS;
  const int sig_col = (CT == 1) ? 0 : 3;
  const int n_extra = (CT == 0) ? C - 4 : 0;
  for (int64_t r = warp0; r < N; r += nwarps) {
    const float dx = __ldg(rays + r * 8 + 3), dy = __ldg(rays + r * 8 + 4), dz = __ldg(rays + r * 8 + 5);
    const float dnorm = sqrtf(dx * dx + dy * dy + dz * dz);                       // rendering.py:168
    float carry = 1.0f, acc_r = 0.f, acc_g = 0.f, acc_b = 0.f, acc_d = 0.f, acc_w = 0.f;
    float acc_e[CT == 0 ? COMP_MAX_EXTRA : 1];
#pragma unroll
    for (int e = 0; e < (CT == 0 ? COMP_MAX_EXTRA : 1); ++e) acc_e[e] = 0.f;
    const float* zr = z + r * S;
    for (int base = 0; base < S; base += 32) {
      const int i = base + lane;
      const bool valid = i < S;
      const float zi = valid ? __ldg(zr + i) : 0.f;
      float zn = __shfl_down_sync(0xffffffffu, zi, 1);
      if (lane == 31 && i + 1 < S) zn = __ldg(zr + i + 1);
      float sigma = 0.f, cr = 0.f, cg = 0.f, cb = 0.f, nz = 0.f;
      const float* px = raw + (r * S + i) * (int64_t)C;
      if (valid) {
        if (CT == 4) {
          const float4 v = __ldg(reinterpret_cast<const float4*>(px));
          cr = v.x; cg = v.y; cb = v.z; sigma = v.w;
        } else if (CT == 1) {
          sigma = __ldg(px);
        } else {
          cr = __ldg(px); cg = __ldg(px + 1); cb = __ldg(px + 2); sigma = __ldg(px + 3);
        }
        if (noise) nz = __ldg(noise + r * S + i);
      }
      SampleTerms st = sample_terms(sigma, nz, noise_std, zi, zn, i == S - 1, dnorm);
      const float tt = valid ? st.t : 1.0f;
      const float incl = warp_incl_prod(tt, lane);
      float excl = __shfl_up_sync(0xffffffffu, incl, 1);
      if (lane == 0) excl = 1.0f;
      const float T = carry * excl;                                             // rendering.py:176-177
      carry *= __shfl_sync(0xffffffffu, incl, 31);
      const float w = valid ? st.alpha * T : 0.f;
      if (valid) weights[r * S + i] = w;
      acc_w += w;
      if (CT != 1) {
        acc_r += w * cr; acc_g += w * cg; acc_b += w * cb; acc_d += w * zi;
        if (CT == 0 && valid) {
#pragma unroll
          for (int e = 0; e < COMP_MAX_EXTRA; ++e)
            if (e < n_extra) acc_e[e] += w * __ldg(px + 4 + e);
        }
      }
    }
    acc_w = warp_sum(acc_w);
    if (CT != 1) {
      acc_r = warp_sum(acc_r); acc_g = warp_sum(acc_g); acc_b = warp_sum(acc_b); acc_d = warp_sum(acc_d);
      if (CT == 0) {
#pragma unroll
        for (int e = 0; e < COMP_MAX_EXTRA; ++e)
          if (e < n_extra) acc_e[e] = warp_sum(acc_e[e]);
      }
    }
    if (lane == 0) {
      if (opacity) opacity[r] = acc_w;
      if (CT != 1) {
        if (rgb) {
          const float bg = white_back ? (1.0f - acc_w) : 0.f;                    // rendering.py:187-188
          rgb[r * 3 + 0] = acc_r + bg; rgb[r * 3 + 1] = acc_g + bg; rgb[r * 3 + 2] = acc_b + bg;
        }
        if (depth) depth[r] = acc_d;
        if (CT == 0 && extra) {
#pragma unroll
          for (int e = 0; e < COMP_MAX_EXTRA; ++e)
            if (e < n_extra) extra[r * n_extra + e] = acc_e[e];
        }
      }
    }
  }
}

// CT = 4 with S = 32 * NCH known at compile time: all of the ray's loads are issued before the first scan (the runtime-S loop
// above serialises chunk k + 1's loads behind chunk k's shuffles).  Same arithmetic in the same order as composite_fwd_kernel<4>.
template <int NCH>
__global__ void __launch_bounds__(COMP_WARPS * 32)
composite_fwd_reg_kernel(const float* __restrict__ raw, const float* __restrict__ z, const float* __restrict__ rays,
                         const float* __restrict__ noise, float noise_std, int white_back, int64_t N,
                         float* __restrict__ weights, float* __restrict__ rgb, float* __restrict__ depth, float* __restrict__ opacity) {
  constexpr int S = 32 * NCH;
  const int lane = threadIdx.x & 31;
  const int64_t nwarps = (int64_t)gridDim.x * COMP_WARPS;
  for (int64_t r = (int64_t)blockIdx.x * COMP_WARPS + (threadIdx.x >> 5); r < N; r += nwarps) {
    float4 rv[NCH];
    float zv[NCH], nz[NCH];
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      const int64_t i = r * S + 32 * c + lane;
      rv[c] = __ldg(reinterpret_cast<const float4*>(raw) + i);
      zv[c] = __ldg(z + i);
      nz[c] = noise ? __ldg(noise + i) : 0.f;
    }
    const float dx = __ldg(rays + r * 8 + 3), dy = __ldg(rays + r * 8 + 4), dz = __ldg(rays + r * 8 + 5);
    const float dnorm = sqrtf(dx * dx + dy * dy + dz * dz);
    float carry = 1.0f, acc_r = 0.f, acc_g = 0.f, acc_b = 0.f, acc_d = 0.f, acc_w = 0.f;
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      float zn = __shfl_down_sync(0xffffffffu, zv[c], 1);
      const float z0 = __shfl_sync(0xffffffffu, zv[c + 1 < NCH ? c + 1 : c], 0);
      if (lane == 31 && c + 1 < NCH) zn = z0;
      const SampleTerms st = sample_terms(rv[c].w, nz[c], noise_std, zv[c], zn, c == NCH - 1 && lane == 31, dnorm);
      const float incl = warp_incl_prod(st.t, lane);
      float excl = __shfl_up_sync(0xffffffffu, incl, 1);
      if (lane == 0) excl = 1.0f;
      const float T = carry * excl;
      carry *= __shfl_sync(0xffffffffu, incl, 31);
      const float w = st.alpha * T;
      weights[r * S + 32 * c + lane] = w;
      acc_w += w;
      acc_r += w * rv[c].x; acc_g += w * rv[c].y; acc_b += w * rv[c].z; acc_d += w * zv[c];
    }
    acc_w = warp_sum(acc_w);
    acc_r = warp_sum(acc_r); acc_g = warp_sum(acc_g); acc_b = warp_sum(acc_b); acc_d = warp_sum(acc_d);
    if (lane == 0) {
      if (opacity) opacity[r] = acc_w;
      if (rgb) {
        const float bg = white_back ? (1.0f - acc_w) : 0.f;
        rgb[r * 3 + 0] = acc_r + bg; rgb[r * 3 + 1] = acc_g + bg; rgb[r * 3 + 2] = acc_b + bg;
      }
      if (depth) depth[r] = acc_d;
    }
  }
}

extern "C" int nerfb_composite_fwd(const float* raw, int channels, const float* z, const float* rays,
                                   const float* noise, float noise_std, int white_back, int64_t n_rays,
                                   int n_samples, float* weights, float* rgb, float* depth, float* opacity,
                                   float* extra, void* stream) {
  NERFB_CHECK_ARG(raw && z && rays && weights && n_rays >= 0 && n_samples >= 1);
  NERFB_CHECK_ARG(channels == 1 || (channels >= 4 && channels <= 4 + COMP_MAX_EXTRA));
  if (n_rays == 0) return NERFB_OK;
  const int64_t want = ceil_div64(n_rays, COMP_WARPS);
  const int grid = (int)(want < NERFB_SM_COUNT * 16 ? want : NERFB_SM_COUNT * 16);
  if (channels == 1)
    NERFB_LAUNCH(composite_fwd_kernel<1>, grid, COMP_WARPS * 32, 0, stream, raw, channels, z, rays, noise, noise_std,
                 white_back, n_rays, n_samples, weights, rgb, depth, opacity, extra);
  else if (channels == 4 && (reinterpret_cast<uintptr_t>(raw) & 15) == 0 && n_samples == 64)
    NERFB_LAUNCH(composite_fwd_reg_kernel<2>, grid, COMP_WARPS * 32, 0, stream, raw, z, rays, noise, noise_std, white_back, n_rays,
                 weights, rgb, depth, opacity);
  else if (channels == 4 && (reinterpret_cast<uintptr_t>(raw) & 15) == 0 && n_samples == 128)
    NERFB_LAUNCH(composite_fwd_reg_kernel<4>, grid, COMP_WARPS * 32, 0, stream, raw, z, rays, noise, noise_std, white_back, n_rays,
                 weights, rgb, depth, opacity);
  else if (channels == 4 && (reinterpret_cast<uintptr_t>(raw) & 15) == 0 && n_samples == 192)
    NERFB_LAUNCH(composite_fwd_reg_kernel<6>, grid, COMP_WARPS * 32, 0, stream, raw, z, rays, noise, noise_std, white_back, n_rays,
                 weights, rgb, depth, opacity);
  else if (channels == 4 && (reinterpret_cast<uintptr_t>(raw) & 15) == 0)
    NERFB_LAUNCH(composite_fwd_kernel<4>, grid, COMP_WARPS * 32, 0, stream, raw, channels, z, rays, noise, noise_std,
                 white_back, n_rays, n_samples, weights, rgb, depth, opacity, extra);
  else
    NERFB_LAUNCH(composite_fwd_kernel<0>, grid, COMP_WARPS * 32, 0, stream, raw, channels, z, rays, noise, noise_std,
                 white_back, n_rays, n_samples, weights, rgb, depth, opacity, extra);
  return NERFB_OK;
}

// ---------------------------------------------------------------------------------------------------
// Backward.  With g_i = dL/dw_i (all paths), t_i = 1 - alpha_i + 1e-10, T_i = prod_{j<i} t_j:
//   dL/dalpha_i = g_i T_i - (sum_{k>i} g_k w_k) / t_i          (cumprod's division form, as autograd)
//   dL/dsigma_i = dL/dalpha_i * delta_i * exp(-delta_i a_i) * [sigma_i + noise_i > 0]
//   dL/dc_i     = w_i * dL/drgb ;  dL/dextra_i = w_i * dL/dextra_map
// Pass 1 walks the ray forward recomputing alpha/T (nothing but the kernel inputs is saved) and parks
// (T, g) per sample in shared memory; pass 2 walks backward with a warp suffix-sum.
__global__ void __launch_bounds__(COMP_WARPS * 32)
composite_bwd_kernel(const float* __restrict__ raw, int C, const float* __restrict__ z, const float* __restrict__ rays,
                     const float* __restrict__ noise, float noise_std, int white_back, int64_t N, int S,
                     const float* __restrict__ d_rgb, const float* __restrict__ d_depth,
                     const float* __restrict__ d_opacity, const float* __restrict__ d_weights,
                     const float* __restrict__ d_extra, float* __restrict__ d_raw) {
  extern __shared__ float smem[];
  const int lane = threadIdx.x & 31;
  const int wib = threadIdx.x >> 5;
  float* sT = smem + (size_t)wib * 2 * S;
  float* sG = sT + S;
  const int64_t nwarps = (int64_t)gridDim.x * COMP_WARPS;
  const int sig_col = (C == 1) ? 0 : 3;
  const int n_extra = (C > 4) ? C - 4 : 0;
  for (int64_t r = (int64_t)blockIdx.x * COMP_WARPS + wib; r < N; r += nwarps) {
    const float dx = __ldg(rays + r * 8 + 3), dy = __ldg(rays + r * 8 + 4), dz = __ldg(rays + r * 8 + 5);
    const float dnorm = sqrtf(dx * dx + dy * dy + dz * dz);
    float gr = 0.f, gg = 0.f, gb = 0.f, gd = 0.f, go = 0.f;
    if (C != 1 && d_rgb) { gr = __ldg(d_rgb + r * 3); gg = __ldg(d_rgb + r * 3 + 1); gb = __ldg(d_rgb + r * 3 + 2); }
    if (C != 1 && d_depth) gd = __ldg(d_depth + r);
    if (d_opacity) go = __ldg(d_opacity + r);
    if (C != 1 && white_back) go -= (gr + gg + gb);               // rgb += 1 - sum(w)
    const float* zr = z + r * S;
    float carry = 1.0f;
    for (int base = 0; base < S; base += 32) {                    // pass 1: forward recompute
      const int i = base + lane;
      const bool valid = i < S;
      const float zi = valid ? __ldg(zr + i) : 0.f;
      float zn = __shfl_down_sync(0xffffffffu, zi, 1);
      if (lane == 31 && i + 1 < S) zn = __ldg(zr + i + 1);
      const float* px = raw + (r * S + i) * (int64_t)C;
      const float sigma = valid ? __ldg(px + sig_col) : 0.f;
      const float nz = (valid && noise) ? __ldg(noise + r * S + i) : 0.f;
      SampleTerms st = sample_terms(sigma, nz, noise_std, zi, zn, i == S - 1, dnorm);
      const float tt = valid ? st.t : 1.0f;
      const float incl = warp_incl_prod(tt, lane);
      float excl = __shfl_up_sync(0xffffffffu, incl, 1);
      if (lane == 0) excl = 1.0f;
      const float T = carry * excl;
      carry *= __shfl_sync(0xffffffffu, incl, 31);
      if (valid) {
        float g = go + (d_weights ? __ldg(d_weights + r * S + i) : 0.f);
        if (C != 1) {
          g += gr * __ldg(px) + gg * __ldg(px + 1) + gb * __ldg(px + 2) + gd * zi;
          for (int e = 0; e < n_extra; ++e)
            if (d_extra) g += __ldg(d_extra + r * n_extra + e) * __ldg(px + 4 + e);
        }
        sT[i] = T;
        sG[i] = g;
      }
    }
    __syncwarp();
    float suffix = 0.f;                                            // sum_{k > chunk end} g_k w_k
    for (int base = ((S - 1) / 32) * 32; base >= 0; base -= 32) {  // pass 2: backward
      const int i = base + lane;
      const bool valid = i < S;
      const float zi = valid ? __ldg(zr + i) : 0.f;
      const float zn = (valid && i + 1 < S) ? __ldg(zr + i + 1) : 0.f;
      const float* px = raw + (r * S + i) * (int64_t)C;
      const float sigma = valid ? __ldg(px + sig_col) : 0.f;
      const float nz = (valid && noise) ? __ldg(noise + r * S + i) : 0.f;
      SampleTerms st = sample_terms(sigma, nz, noise_std, zi, zn, i == S - 1, dnorm);
      const float T = valid ? sT[i] : 0.f;
      const float g = valid ? sG[i] : 0.f;
      const float w = st.alpha * T;
      const float gw = valid ? g * w : 0.f;
      float incl = gw;                                             // inclusive suffix sum over lanes >= me
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const float q = __shfl_down_sync(0xffffffffu, incl, o);
        if (lane + o < 32) incl += q;
      }
      const float after = suffix + (incl - gw);                    // sum_{k>i} g_k w_k
      suffix += __shfl_sync(0xffffffffu, incl, 0);
      if (valid) {
        const float dalpha = g * T - after / st.t;
        const float dsig = st.on ? dalpha * st.delta * st.e : 0.f;
        float* po = d_raw + (r * S + i) * (int64_t)C;
        po[sig_col] = dsig;
        if (C != 1) {
          po[0] = w * gr; po[1] = w * gg; po[2] = w * gb;
          for (int e = 0; e < n_extra; ++e) po[4 + e] = d_extra ? w * __ldg(d_extra + r * n_extra + e) : 0.f;
        }
      }
    }
    __syncwarp();
  }
}

// The same for the common shape (C = 4, S = 32 * NCH, 16-byte aligned raw / d_raw): every input of the ray is read ONCE into
// registers (raw as float4, z, noise, d_weights), pass 2 runs from registers (the generic kernel re-reads raw / z / noise and
// parks T, g in shared memory), the gradient leaves as one float4 per sample.  Same arithmetic in the same order: bit-identical.
template <int NCH>
__global__ void __launch_bounds__(COMP_WARPS * 32)
composite_bwd_reg_kernel(const float* __restrict__ raw, const float* __restrict__ z, const float* __restrict__ rays,
                         const float* __restrict__ noise, float noise_std, int white_back, int64_t N,
                         const float* __restrict__ d_rgb, const float* __restrict__ d_depth, const float* __restrict__ d_opacity,
                         const float* __restrict__ d_weights, float* __restrict__ d_raw) {
  constexpr int S = 32 * NCH;
  const int lane = threadIdx.x & 31;
  const int64_t nwarps = (int64_t)gridDim.x * COMP_WARPS;
  for (int64_t r = (int64_t)blockIdx.x * COMP_WARPS + (threadIdx.x >> 5); r < N; r += nwarps) {
    float4 rv[NCH];
    float zv[NCH], nz[NCH], g[NCH];
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      const int64_t i = r * S + 32 * c + lane;
      rv[c] = __ldg(reinterpret_cast<const float4*>(raw) + i);
      zv[c] = __ldg(z + i);
      nz[c] = noise ? __ldg(noise + i) : 0.f;
      g[c] = d_weights ? __ldg(d_weights + i) : 0.f;
    }
    const float dx = __ldg(rays + r * 8 + 3), dy = __ldg(rays + r * 8 + 4), dz = __ldg(rays + r * 8 + 5);
    const float dnorm = sqrtf(dx * dx + dy * dy + dz * dz);
    float gr = 0.f, gg = 0.f, gb = 0.f, gd = 0.f, go = 0.f;
    if (d_rgb) { gr = __ldg(d_rgb + r * 3); gg = __ldg(d_rgb + r * 3 + 1); gb = __ldg(d_rgb + r * 3 + 2); }
    if (d_depth) gd = __ldg(d_depth + r);
    if (d_opacity) go = __ldg(d_opacity + r);
    if (white_back) go -= (gr + gg + gb);
    float T[NCH], alpha[NCH], tt[NCH], ee[NCH], dl[NCH];
    bool on[NCH];
    float carry = 1.0f;
#pragma unroll
    for (int c = 0; c < NCH; ++c) {                                // pass 1: forward
      float zn = __shfl_down_sync(0xffffffffu, zv[c], 1);
      const float z0 = __shfl_sync(0xffffffffu, zv[c + 1 < NCH ? c + 1 : c], 0);
      if (lane == 31 && c + 1 < NCH) zn = z0;
      const SampleTerms st = sample_terms(rv[c].w, nz[c], noise_std, zv[c], zn, c == NCH - 1 && lane == 31, dnorm);
      const float incl = warp_incl_prod(st.t, lane);
      float excl = __shfl_up_sync(0xffffffffu, incl, 1);
      if (lane == 0) excl = 1.0f;
      T[c] = carry * excl;
      carry *= __shfl_sync(0xffffffffu, incl, 31);
      float gi = go + g[c];
      gi += gr * rv[c].x + gg * rv[c].y + gb * rv[c].z + gd * zv[c];
      g[c] = gi;
      alpha[c] = st.alpha; tt[c] = st.t; ee[c] = st.e; dl[c] = st.delta; on[c] = st.on;
    }
    float suffix = 0.f;
#pragma unroll
    for (int c = NCH - 1; c >= 0; --c) {                           // pass 2: backward
      const float w = alpha[c] * T[c];
      const float gw = g[c] * w;
      float incl = gw;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const float q = __shfl_down_sync(0xffffffffu, incl, o);
        if (lane + o < 32) incl += q;
      }
      const float after = suffix + (incl - gw);
      suffix += __shfl_sync(0xffffffffu, incl, 0);
      const float dalpha = g[c] * T[c] - after / tt[c];
      const float dsig = on[c] ? dalpha * dl[c] * ee[c] : 0.f;
      reinterpret_cast<float4*>(d_raw)[r * S + 32 * c + lane] = make_float4(w * gr, w * gg, w * gb, dsig);
    }
  }
}

extern "C" int nerfb_composite_bwd(const float* raw, int channels, const float* z, const float* rays,
                                   const float* noise, float noise_std, int white_back, int64_t n_rays,
                                   int n_samples, const float* d_rgb, const float* d_depth,
                                   const float* d_opacity, const float* d_weights, const float* d_extra,
                                   float* d_raw, void* stream) {
  NERFB_CHECK_ARG(raw && z && rays && d_raw && n_rays >= 0 && n_samples >= 1);
  NERFB_CHECK_ARG(channels == 1 || (channels >= 4 && channels <= 4 + COMP_MAX_EXTRA));
  if (n_rays == 0) return NERFB_OK;
  const int64_t want_r = ceil_div64(n_rays, COMP_WARPS);
  const int grid_r = (int)(want_r < NERFB_SM_COUNT * 16 ? want_r : NERFB_SM_COUNT * 16);
  if (channels == 4 && !d_extra && ((reinterpret_cast<uintptr_t>(raw) | reinterpret_cast<uintptr_t>(d_raw)) & 15) == 0 &&
      (n_samples == 32 || n_samples == 64 || n_samples == 128)) {
#define NERFB_COMP_BWD_REG(NCH)                                                                                              \
  NERFB_LAUNCH(composite_bwd_reg_kernel<NCH>, grid_r, COMP_WARPS * 32, 0, stream, raw, z, rays, noise, noise_std, white_back, \
               n_rays, d_rgb, d_depth, d_opacity, d_weights, d_raw)
    if (n_samples == 32) NERFB_COMP_BWD_REG(1);
    else if (n_samples == 64) NERFB_COMP_BWD_REG(2);
    else NERFB_COMP_BWD_REG(4);
#undef NERFB_COMP_BWD_REG
    return NERFB_OK;
  }
  const size_t smem = (size_t)COMP_WARPS * 2 * n_samples * sizeof(float);
  if (smem > 200 * 1024) return NERFB_E_UNSUPPORTED;
  if (smem > 48 * 1024)
    NERFB_CUDA(cudaFuncSetAttribute(composite_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int64_t want = ceil_div64(n_rays, COMP_WARPS);
  const int grid = (int)(want < NERFB_SM_COUNT * 16 ? want : NERFB_SM_COUNT * 16);
  // rgb with white background uses white_back only through d_rgb (folded into go)
  NERFB_LAUNCH(composite_bwd_kernel, grid, COMP_WARPS * 32, smem, stream, raw, channels, z, rays, noise, noise_std,
               white_back, n_rays, n_samples, d_rgb, d_depth, d_opacity, d_weights, d_extra, d_raw);
  return NERFB_OK;
}
