// Kernel (3): importance resampling -- replaces models/rendering.py:22-67 (sample_pdf), its call site
// :241-247 (mid-points, weights[:,1:-1], cat, sort) and the legacy torchsearchsorted extension
// (torchsearchsorted/src/cuda/searchsorted_cuda_kernel.cu:3-142).
// One warp per ray; the row's bins / cdf live in shared memory, the searches are warp-parallel over the
// queries, and the final sort is a rank-by-counting merge in shared memory (exact multiset of
// torch.sort's values, no assumption that the inverse-CDF output is monotone to the ulp).
#include "common.cuh"
#include "resample_dev.cuh"

#define RS_WARPS 4

// ---- searchsorted: out = #(a[row] <  v) (left)  or  #(a[row] <= v) (right) ---------------------------
// Instantiated for every scalar type of the reference's AT_DISPATCH_ALL_TYPES (searchsorted_cuda_kernel.cu:134).
template <typename T>
__global__ void __launch_bounds__(256)
searchsorted_kernel(const T* __restrict__ a, int64_t rows_a, int64_t A, const T* __restrict__ v, int64_t rows_v,
                    int64_t V, int side_left, int64_t rows, int64_t* __restrict__ out) {
  const int64_t total = rows * V;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = idx / V, c = idx - r * V;
    const T* ar = a + (rows_a == 1 ? 0 : r) * A;
    const T val = __ldg(v + (rows_v == 1 ? 0 : r) * V + c);
    int64_t lo = 0, hi = A;                                  // first position whose element fails the test
    while (lo < hi) {
      const int64_t mid = (lo + hi) >> 1;
      const T x = __ldg(ar + mid);
      const bool go_right = side_left ? (x < val) : (x <= val);
      if (go_right) lo = mid + 1; else hi = mid;
    }
    out[idx] = lo;
  }
}

template <typename T>
static int searchsorted_launch(const void* a, int64_t rows_a, int64_t len_a, const void* v, int64_t rows_v, int64_t len_v,
                               int side_left, int64_t* out, void* stream) {
  const int64_t rows = rows_a > rows_v ? rows_a : rows_v;
  const int64_t want = ceil_div64(rows * len_v, 256);
  const int grid = (int)(want < NERFB_SM_COUNT * 8 ? want : NERFB_SM_COUNT * 8);
  NERFB_LAUNCH(searchsorted_kernel<T>, grid, 256, 0, stream, static_cast<const T*>(a), rows_a, len_a, static_cast<const T*>(v),
               rows_v, len_v, side_left, rows, out);
  return NERFB_OK;
}

extern "C" int nerfb_searchsorted_typed(const void* a, int64_t rows_a, int64_t len_a, const void* v, int64_t rows_v,
                                        int64_t len_v, int dtype, int side_left, int64_t* out, void* stream) {
  NERFB_CHECK_ARG(a && v && out && rows_a >= 1 && rows_v >= 1 && len_a >= 0 && len_v >= 0);
  NERFB_CHECK_ARG(rows_a == rows_v || rows_a == 1 || rows_v == 1);
  const int64_t rows = rows_a > rows_v ? rows_a : rows_v;
  if (rows * len_v == 0) return NERFB_OK;
  switch (dtype) {
    case NERFB_DT_F32: return searchsorted_launch<float>(a, rows_a, len_a, v, rows_v, len_v, side_left, out, stream);
    case NERFB_DT_F64: return searchsorted_launch<double>(a, rows_a, len_a, v, rows_v, len_v, side_left, out, stream);
    case NERFB_DT_I64: return searchsorted_launch<long long>(a, rows_a, len_a, v, rows_v, len_v, side_left, out, stream);
    case NERFB_DT_I32: return searchsorted_launch<int>(a, rows_a, len_a, v, rows_v, len_v, side_left, out, stream);
    case NERFB_DT_I16: return searchsorted_launch<short>(a, rows_a, len_a, v, rows_v, len_v, side_left, out, stream);
    case NERFB_DT_I8: return searchsorted_launch<signed char>(a, rows_a, len_a, v, rows_v, len_v, side_left, out, stream);
    case NERFB_DT_U8: return searchsorted_launch<unsigned char>(a, rows_a, len_a, v, rows_v, len_v, side_left, out, stream);
    default: return NERFB_E_UNSUPPORTED;
  }
}

extern "C" int nerfb_searchsorted(const float* a, int64_t rows_a, int64_t len_a, const float* v, int64_t rows_v,
                                  int64_t len_v, int side_left, int64_t* out, void* stream) {
  return nerfb_searchsorted_typed(a, rows_a, len_a, v, rows_v, len_v, NERFB_DT_F32, side_left, out, stream);
}

__global__ void __launch_bounds__(RS_WARPS * 32)
sample_pdf_kernel(const float* __restrict__ bins, const float* __restrict__ weights, const float* __restrict__ u, int det,
                  float eps, int64_t N, int nb, int n_imp, float* __restrict__ samples, float* __restrict__ cdf_out,
                  int64_t* __restrict__ inds_out) {
  extern __shared__ float smem[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  float* s_bins = smem + (size_t)wib * 3 * nb;
  float* s_w = s_bins + nb;
  float* s_cdf = s_w + nb;
  for (int64_t r = (int64_t)blockIdx.x * RS_WARPS + wib; r < N; r += (int64_t)gridDim.x * RS_WARPS) {
    for (int i = lane; i < nb; i += 32) s_bins[i] = __ldg(bins + r * nb + i);
    for (int i = lane; i < nb - 1; i += 32) s_w[i] = __ldg(weights + r * (nb - 1) + i);
    __syncwarp();
    warp_build_cdf(s_w, s_cdf, nb, eps, lane);
    if (cdf_out) for (int i = lane; i < nb; i += 32) cdf_out[r * nb + i] = s_cdf[i];
    for (int j = lane; j < n_imp; j += 32) {
      const float uj = det ? linspace01(j, n_imp) : __ldg(u + r * n_imp + j);
      int ind;
      samples[r * n_imp + j] = invert_cdf(s_bins, s_cdf, nb, uj, eps, &ind);
      if (inds_out) inds_out[r * n_imp + j] = ind;
    }
    __syncwarp();
  }
}

extern "C" int nerfb_sample_pdf(const float* bins, const float* weights, const float* u, int det, float eps,
                                int64_t n_rays, int n_bins, int n_importance, float* samples, float* cdf_out,
                                int64_t* inds_out, void* stream) {
  NERFB_CHECK_ARG(bins && weights && samples && n_rays >= 0 && n_bins >= 2 && n_importance >= 1);
  NERFB_CHECK_ARG(det || u);
  if (n_rays == 0) return NERFB_OK;
  const size_t smem = (size_t)RS_WARPS * 3 * n_bins * sizeof(float);
  if (smem > 200 * 1024) return NERFB_E_UNSUPPORTED;
  if (smem > 48 * 1024)
    NERFB_CUDA(cudaFuncSetAttribute(sample_pdf_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int64_t want = ceil_div64(n_rays, RS_WARPS);
  const int grid = (int)(want < NERFB_SM_COUNT * 16 ? want : NERFB_SM_COUNT * 16);
  NERFB_LAUNCH(sample_pdf_kernel, grid, RS_WARPS * 32, smem, stream, bins, weights, u, det, eps, n_rays, n_bins,
               n_importance, samples, cdf_out, inds_out);
  return NERFB_OK;
}

// In-register bitonic network over NPL * 32 values (element q * 32 + lane in v[q]).  SKIP_LOWER: the lower half is already
// ascending, stages k <= half only sort the upper half.  Everything but the lane-dependent direction bits is a compile-time constant.
template <int NPL, bool SKIP_LOWER>
__device__ __forceinline__ void bitonic_network(float (&v)[NPL > 0 ? NPL : 1], int lane) {
  constexpr int HALF = NPL * 16;
#pragma unroll
  for (int k = 2; k <= NPL * 32; k <<= 1) {
#pragma unroll
    for (int j = k >> 1; j > 0; j >>= 1) {
      if (j >= 32) {
        const int dq = j >> 5;
#pragma unroll
        for (int q = 0; q < NPL; ++q) {
          if ((q & dq) == 0) {
            if (SKIP_LOWER && k <= HALF && q < NPL / 2) continue;
            const bool up = (((q * 32) & k) == 0);
            const float a = v[q], b = v[q | dq];
            const bool sw = (a > b) == up;
            v[q] = sw ? b : a; v[q | dq] = sw ? a : b;
          }
        }
      } else {
        // the lower index keeps the min in an ascending block, the max in a descending one; for k < 32 the block direction
        // depends on the lane only, for k >= 32 on q only
        const bool lower = (lane & j) == 0;
        const bool lane_up = (lane & k) == 0;
#pragma unroll
        for (int q = 0; q < NPL; ++q) {
          if (SKIP_LOWER && k <= HALF && q < NPL / 2) continue;
          const float o = __shfl_xor_sync(0xffffffffu, v[q], j);
          const bool up = k < 32 ? lane_up : (((q * 32) & k) == 0);
          v[q] = (lower == up) ? fminf(v[q], o) : fmaxf(v[q], o);
        }
      }
    }
  }
}

// ---- the fused call site: z_mid -> sample_pdf(weights[1:-1]) -> cat -> sort ---------------------------
// NPL: elements per lane of the in-register sort (n_pad / 32), 0 = shared-memory sort for other sizes.  SC / SI: the sample counts
// when known at compile time (64 + 64 and 64 + 128, the reference's configurations), 0 = runtime: the kernel is instruction-bound
// (~1100 warp-instructions per ray), and its many short `for (i = lane; i < Sc; i += 32)` loops unroll only with constants.
template <int NPL, int SC, int SI>
__global__ void __launch_bounds__(RS_WARPS * 32)
resample_kernel(const float* __restrict__ z_coarse, const float* __restrict__ w_coarse, const float* __restrict__ u, int det,
                int64_t N, int Sc_rt, int Si_rt, int n_pad, float* __restrict__ z_fine) {
  const int Sc = SC > 0 ? SC : Sc_rt, Si = SI > 0 ? SI : Si_rt;
  extern __shared__ float smem[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int nb = Sc - 1, St = Sc + Si;
  float* s_bins = smem + (size_t)wib * (3 * Sc + n_pad);
  float* s_w = s_bins + Sc;
  float* s_cdf = s_w + Sc;
  float* s_all = s_cdf + Sc;                               // [0,Sc) coarse z, [Sc,St) new samples
  const float eps = 1e-5f;
  for (int64_t r = (int64_t)blockIdx.x * RS_WARPS + wib; r < N; r += (int64_t)gridDim.x * RS_WARPS) {
    for (int i = lane; i < Sc; i += 32) s_all[i] = __ldg(z_coarse + r * Sc + i);
    for (int i = lane; i < Sc - 2; i += 32) s_w[i] = __ldg(w_coarse + r * Sc + i + 1);      // weights[:, 1:-1]
    __syncwarp();
    for (int i = lane; i < nb; i += 32) s_bins[i] = __fmul_rn(0.5f, __fadd_rn(s_all[i], s_all[i + 1]));  // :242
    __syncwarp();
    warp_build_cdf(s_w, s_cdf, nb, eps, lane);
    for (int j = lane; j < Si; j += 32) {
      const float uj = det ? linspace01(j, Si) : __ldg(u + r * Si + j);
      s_all[Sc + j] = invert_cdf(s_bins, s_cdf, nb, uj, eps, nullptr);
    }
    __syncwarp();
    // warp bitonic sort of the St values in shared memory, padded with +inf to the next power of two (values only, so the
    // result is the exact multiset torch.sort returns; no assumption that the inverse-CDF output is monotone to the ulp).
    // log2(n)(log2(n)+1)/2 compare-exchange sweeps of n/2 pairs: 28 sweeps x 2 pairs per lane at St = 128, against the
    // 128 x 128 / 32 comparisons per lane of a rank-by-counting sort.
    if (NPL > 0) {
      // in-register bitonic network: element idx = q * 32 + lane lives in v[q] of lane `lane`; partners at distance j < 32
      // are one shuffle away, at distance j >= 32 they are another register of the same lane.
      // Layout: lower half of the network = the coarse depths (+inf padded), upper half = the new samples (+inf padded).
      // The coarse depths come out of kernel (2) ascending, so when a warp-wide check confirms that (no assumption: an
      // unsorted row takes the full network) the stages k <= half only have to sort the UPPER half -- which the network
      // leaves descending -- and the last stage is one bitonic merge: 21 x 2 + 7 x 4 = 70 compare-exchanges per lane at
      // 64 + 64 instead of 28 x 4 = 112, 176 instead of 288 at 64 + 128.  Values only, exact multiset of torch.sort.
      constexpr int HALF = NPL * 16;
      const bool split = Sc <= HALF && Si <= HALF;
      float v[NPL > 0 ? NPL : 1];
#pragma unroll
      for (int q = 0; q < NPL; ++q) {
        const int idx = q * 32 + lane;
        if (split) v[q] = idx < HALF ? (idx < Sc ? s_all[idx] : INFINITY) : (idx - HALF < Si ? s_all[Sc + idx - HALF] : INFINITY);
        else v[q] = idx < St ? s_all[idx] : INFINITY;
      }
      bool presorted = false;
      if (split) {
        bool bad = false;
#pragma unroll
        for (int q = 0; q < NPL / 2; ++q) {                 // v[q] <= the next element of the lower half (idx + 1), +inf past its end
          float nxt = __shfl_down_sync(0xffffffffu, v[q], 1);
          const float wrap = __shfl_sync(0xffffffffu, v[q + 1 < NPL / 2 ? q + 1 : q], 0);
          if (lane == 31) nxt = (q + 1 < NPL / 2) ? wrap : INFINITY;
          bad |= !(v[q] <= nxt);
        }
        presorted = !__any_sync(0xffffffffu, bad);
      }
      if (presorted) bitonic_network<NPL, true>(v, lane);     // warp-uniform: two fully unrolled networks, no per-exchange test
      else bitonic_network<NPL, false>(v, lane);
#pragma unroll
      for (int q = 0; q < NPL; ++q) { const int idx = q * 32 + lane; if (idx < St) z_fine[r * St + idx] = v[q]; }
      __syncwarp();
      continue;
    }
    for (int i = St + lane; i < n_pad; i += 32) s_all[i] = INFINITY;
    __syncwarp();
    for (int k = 2; k <= n_pad; k <<= 1) {
      for (int j = k >> 1; j > 0; j >>= 1) {
        for (int t = lane; t < (n_pad >> 1); t += 32) {
          const int lo = ((t & ~(j - 1)) << 1) | (t & (j - 1));      // index with bit j clear
          const int hi = lo | j;
          const float a = s_all[lo], b = s_all[hi];
          const bool up = (lo & k) == 0;                              // ascending block
          if ((a > b) == up) { s_all[lo] = b; s_all[hi] = a; }
        }
        __syncwarp();
      }
    }
    for (int i = lane; i < St; i += 32) z_fine[r * St + i] = s_all[i];
    __syncwarp();
  }
}

extern "C" int nerfb_resample(const float* z_coarse, const float* weights_coarse, const float* u, int det,
                              int64_t n_rays, int n_coarse, int n_importance, float* z_fine, void* stream) {
  NERFB_CHECK_ARG(z_coarse && weights_coarse && z_fine && n_rays >= 0 && n_coarse >= 3 && n_importance >= 1);
  NERFB_CHECK_ARG(det || u);
  if (n_rays == 0) return NERFB_OK;
  int n_pad = 2;
  while (n_pad < n_coarse + n_importance) n_pad <<= 1;
  const size_t smem = (size_t)RS_WARPS * (3 * n_coarse + n_pad) * sizeof(float);
  if (smem > 200 * 1024) return NERFB_E_UNSUPPORTED;
  if (smem > 48 * 1024)
    NERFB_CUDA(cudaFuncSetAttribute(resample_kernel<0, 0, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int64_t want = ceil_div64(n_rays, RS_WARPS);
  const int grid = (int)(want < NERFB_SM_COUNT * 16 ? want : NERFB_SM_COUNT * 16);
#define NERFB_RESAMPLE(NPL, SC, SI)                                                                                          \
  NERFB_LAUNCH((resample_kernel<NPL, SC, SI>), grid, RS_WARPS * 32, smem, stream, z_coarse, weights_coarse, u, det, n_rays,    \
               n_coarse, n_importance, n_pad, z_fine)
  if (n_coarse == 64 && n_importance == 64) NERFB_RESAMPLE(4, 64, 64);
  else if (n_coarse == 64 && n_importance == 128) NERFB_RESAMPLE(8, 64, 128);
  else if (n_pad == 128 && smem <= 48 * 1024) NERFB_RESAMPLE(4, 0, 0);
  else if (n_pad == 256 && smem <= 48 * 1024) NERFB_RESAMPLE(8, 0, 0);
  else NERFB_RESAMPLE(0, 0, 0);
#undef NERFB_RESAMPLE
  return NERFB_OK;
}
