// fp32 path of kernel (1): positional encoding + MLP forward/backward on the CUDA cores.
//   replaces models/rendering.py:129-159 (chunked embed -> cat -> model) and the autograd of
//   models/nerf.py:83-124 (NeRF), models/nerf_cls.py:70-112 (NeRF_3D) and models/nerf.py:201-216 (SIREN/FiLM).
// This is the tight-parity arm (NERFB_PREC_FP32): fp32 operands, fp32 accumulate, one strided SGEMM kernel
// composed layer by layer with fused bias + activation epilogues, skip / direction concatenations done by
// writing producers straight into column windows of wider buffers (no cat kernels).  The tensor-core arm
// (mlp_tc*.cu) is the fast path; this one exists so parity with the fp32 reference can be stated to 1e-5.
#include "common.cuh"

#define ACT_NONE 0
#define ACT_RELU 1
#define ACT_SIGMOID 2
#define ACT_SINE 3

#define TB_M 64
#define TB_N 64
#define TB_K 16

// C[i*ldc + j] (+)= act( sum_k X[i*sxi + k*sxk] * Y[k*syk + j*syj] + bias[j] )
// split-K: gridDim.z slices of K, results combined with atomicAdd (then bias/act must be off).
__global__ void __launch_bounds__(256)
sgemm_strided_kernel(int64_t M, int N, int64_t K, const float* __restrict__ X, int64_t sxi, int64_t sxk,
                     const float* __restrict__ Y, int64_t syk, int64_t syj, float* __restrict__ C, int64_t ldc,
                     const float* __restrict__ bias, int act, const float* __restrict__ freq,
                     const float* __restrict__ phase, float* __restrict__ zpre, int64_t ldz, int accumulate,
                     int64_t k_per_slice) {
  __shared__ float Xs[TB_K][TB_M + 4];
  __shared__ float Ys[TB_K][TB_N + 4];
  const int tid = threadIdx.x;
  const int64_t i0 = (int64_t)blockIdx.x * TB_M;
  const int j0 = blockIdx.y * TB_N;
  const int64_t kbeg = (int64_t)blockIdx.z * k_per_slice;
  const int64_t kend = (kbeg + k_per_slice < K) ? kbeg + k_per_slice : K;
  const int ty = tid / 16, tx = tid % 16;            // 16x16 threads, 4x4 outputs each
  float acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;

  for (int64_t k0 = kbeg; k0 < kend; k0 += TB_K) {
    // load X tile (TB_M x TB_K = 1024 elements, 4 per thread), mapping chosen by the unit stride
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int lin = tid + e * 256;
      int ii, kk;
      if (sxk == 1) { kk = lin % TB_K; ii = lin / TB_K; } else { ii = lin % TB_M; kk = lin / TB_M; }
      const int64_t gi = i0 + ii, gk = k0 + kk;
      Xs[kk][ii] = (gi < M && gk < kend) ? __ldg(X + gi * sxi + gk * sxk) : 0.f;
    }
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int lin = tid + e * 256;
      int jj, kk;
      if (syk == 1) { kk = lin % TB_K; jj = lin / TB_K; } else { jj = lin % TB_N; kk = lin / TB_N; }
      const int64_t gk = k0 + kk;
      const int gj = j0 + jj;
      Ys[kk][jj] = (gj < N && gk < kend) ? __ldg(Y + gk * syk + (int64_t)gj * syj) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < TB_K; ++kk) {
      float xa[4], yb[4];
#pragma unroll
      for (int a = 0; a < 4; ++a) xa[a] = Xs[kk][ty * 4 + a];
#pragma unroll
      for (int b = 0; b < 4; ++b) yb[b] = Ys[kk][tx * 4 + b];
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = fmaf(xa[a], yb[b], acc[a][b]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int a = 0; a < 4; ++a) {
    const int64_t gi = i0 + ty * 4 + a;
    if (gi >= M) continue;
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const int gj = j0 + tx * 4 + b;
      if (gj >= N) continue;
      float v = acc[a][b];
      if (gridDim.z > 1) { atomicAdd(C + gi * ldc + gj, v); continue; }
      if (bias) v += __ldg(bias + gj);
      if (zpre) zpre[gi * ldz + gj] = v;
      if (act == ACT_RELU) v = fmaxf(v, 0.f);
      else if (act == ACT_SIGMOID) v = 1.0f / (1.0f + expf(-v));
      else if (act == ACT_SINE) v = sinf(__ldg(freq + gj) * v + __ldg(phase + gj));
      if (accumulate) v += C[gi * ldc + gj];
      C[gi * ldc + gj] = v;
    }
  }
}

struct Gemm {
  int64_t M; int N; int64_t K;
  const float* X; int64_t sxi, sxk;
  const float* Y; int64_t syk, syj;
  float* C; int64_t ldc;
  const float* bias = nullptr; int act = ACT_NONE; const float* freq = nullptr; const float* phase = nullptr;
  float* zpre = nullptr; int64_t ldz = 0; int accumulate = 0; int split_k = 0;
};

static int run_gemm(const Gemm& g, void* stream) {
  if (g.M <= 0 || g.N <= 0 || g.K <= 0) return NERFB_OK;
  dim3 grid((unsigned)ceil_div64(g.M, TB_M), (unsigned)ceil_div64(g.N, TB_N), 1);
  int64_t kps = g.K;
  if (g.split_k) {     // reduction over the points: enough slices to fill the chip, atomics combine them
    int64_t want = (int64_t)NERFB_SM_COUNT * 4 / ((int64_t)grid.x * grid.y) + 1;
    int64_t max_slices = ceil_div64(g.K, 1024);
    if (want > max_slices) want = max_slices;
    if (want < 1) want = 1;
    kps = round_up64(ceil_div64(g.K, want), TB_K);
    grid.z = (unsigned)ceil_div64(g.K, kps);
    if (grid.z == 1) {  // atomics path is keyed on gridDim.z > 1; a single slice must still accumulate
      NERFB_LAUNCH(sgemm_strided_kernel, grid, 256, 0, stream, g.M, g.N, g.K, g.X, g.sxi, g.sxk, g.Y, g.syk, g.syj,
                   g.C, g.ldc, (const float*)nullptr, ACT_NONE, (const float*)nullptr, (const float*)nullptr,
                   (float*)nullptr, (int64_t)0, 1, kps);
      return NERFB_OK;
    }
  }
  NERFB_LAUNCH(sgemm_strided_kernel, grid, 256, 0, stream, g.M, g.N, g.K, g.X, g.sxi, g.sxk, g.Y, g.syk, g.syj, g.C,
               g.ldc, g.bias, g.act, g.freq, g.phase, g.zpre, g.ldz, g.accumulate, kps);
  return NERFB_OK;
}
#define RUN(g) do { int _s = run_gemm((g), stream); if (_s != NERFB_OK) return _s; } while (0)

// ---- small elementwise helpers -------------------------------------------------------------------------
// d[i,j] *= act'(saved[i,j])   (relu: saved = output; sigmoid: saved = output; sine: saved = pre-activation)
__global__ void __launch_bounds__(256)
act_bwd_kernel(int act, float* __restrict__ d, int64_t ldd, const float* __restrict__ saved, int64_t lds, int64_t M,
               int N, const float* __restrict__ freq, const float* __restrict__ phase) {
  const int64_t total = M * N;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t i = idx / N; const int j = (int)(idx - i * N);
    const float s = saved[i * lds + j];
    float g = d[i * ldd + j];
    if (act == ACT_RELU) g = (s > 0.f) ? g : 0.f;
    else if (act == ACT_SIGMOID) g = g * s * (1.0f - s);
    else if (act == ACT_SINE) { const float f = __ldg(freq + j); g = g * f * cosf(f * s + __ldg(phase + j)); }
    d[i * ldd + j] = g;
  }
}

// out[i, 0:N] = in[i*ldi + col0 + j] * scale(saved)  -- copy of a column window with act' applied
__global__ void __launch_bounds__(256)
window_act_bwd_kernel(int act, const float* __restrict__ src, int64_t lds_, int col0, const float* __restrict__ saved,
                      int64_t ldsv, int colsv, float* __restrict__ dst, int64_t ldd, int64_t M, int N) {
  const int64_t total = M * N;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t i = idx / N; const int j = (int)(idx - i * N);
    float g = src[i * lds_ + col0 + j];
    if (act == ACT_SIGMOID) { const float s = saved[i * ldsv + colsv + j]; g = g * s * (1.0f - s); }
    dst[i * ldd + j] = g;
  }
}

// bias gradient: out[j] += sum_i d[i*ld + j]
__global__ void __launch_bounds__(256)
colsum_kernel(const float* __restrict__ d, int64_t ld, int64_t M, int N, int64_t rows_per_block, float* __restrict__ out) {
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
  const int64_t r1 = (r0 + rows_per_block < M) ? r0 + rows_per_block : M;
  for (int j = threadIdx.x; j < N; j += blockDim.x) {
    float s = 0.f;
    for (int64_t i = r0; i < r1; ++i) s += d[i * ld + j];
    atomicAdd(out + j, s);
  }
}

static int run_colsum(const float* d, int64_t ld, int64_t M, int N, float* out, void* stream) {
  if (M <= 0) return NERFB_OK;
  int64_t blocks = NERFB_SM_COUNT * 4;
  int64_t rpb = ceil_div64(M, blocks);
  if (rpb < 64) rpb = 64;
  blocks = ceil_div64(M, rpb);
  NERFB_LAUNCH(colsum_kernel, (unsigned)blocks, 256, 0, stream, d, ld, M, N, rpb, out);
  return NERFB_OK;
}

static int run_act_bwd(int act, float* d, int64_t ldd, const float* saved, int64_t lds, int64_t M, int N,
                       const float* freq, const float* phase, void* stream) {
  const int64_t want = ceil_div64(M * N, 256);
  const int grid = (int)(want < NERFB_SM_COUNT * 8 ? want : NERFB_SM_COUNT * 8);
  NERFB_LAUNCH(act_bwd_kernel, grid, 256, 0, stream, act, d, ldd, saved, lds, M, N, freq, phase);
  return NERFB_OK;
}

// ---- input preparation: encodings written into their column windows ----------------------------------
// NeRF: xyz encoding -> enc0 (P,in_xyz) and every skip buffer's first in_xyz columns; dir encoding -> dirbuf cols [W, W+in_dir)
// mode RAYS: x = o + d*z computed here (rendering.py:224); mode EMBEDDED: columns copied from x.
struct PrepArgs {
  const float* in0; const float* in1; int in_mode; int64_t P; int S;
  int n_freq_xyz, n_freq_dir, logscale, in_xyz, in_dir;
  float* dst_xyz[4]; int64_t ld_xyz[4]; int n_xyz_dst;
  float* dst_dir; int64_t ld_dir;        // may be null (sigma_only)
  int siren; float scale;                // SIREN: xyz*2/box -> dst_xyz[0] (3 cols); raw dir -> dst_dir (3 cols)
};

__device__ __forceinline__ float band_f(int k, int n_freqs, int logscale) {
  if (logscale) return (float)(1u << k);
  if (n_freqs <= 1) return 1.0f;
  const float end = (float)(1u << (n_freqs - 1));
  const float step = __fdiv_rn(end - 1.0f, (float)(n_freqs - 1));
  return (k < n_freqs / 2) ? __fmaf_rn((float)k, step, 1.0f) : __fmaf_rn(-(float)(n_freqs - 1 - k), step, end);
}

__global__ void __launch_bounds__(256) prep_inputs_kernel(PrepArgs a) {
  const int64_t total = a.P * 3;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t p = idx / 3; const int c = (int)(idx - p * 3);
    if (a.in_mode == NERFB_IN_RAYS) {
      const int64_t r = p / a.S;
      const float o = __ldg(a.in0 + r * 8 + c), d = __ldg(a.in0 + r * 8 + 3 + c);
      const float x = __fadd_rn(o, __fmul_rn(d, __ldg(a.in1 + p)));          // rendering.py:224
      if (a.siren) {
        a.dst_xyz[0][p * a.ld_xyz[0] + c] = x * a.scale;                     // nerf.py:140,204
        if (a.dst_dir) a.dst_dir[p * a.ld_dir + c] = d;
        continue;
      }
      for (int t = 0; t < a.n_xyz_dst; ++t) a.dst_xyz[t][p * a.ld_xyz[t] + c] = x;
      for (int k = 0; k < a.n_freq_xyz; ++k) {
        float s, co; sincosf(__fmul_rn(band_f(k, a.n_freq_xyz, a.logscale), x), &s, &co);
        for (int t = 0; t < a.n_xyz_dst; ++t) {
          a.dst_xyz[t][p * a.ld_xyz[t] + 3 + 6 * k + c] = s;
          a.dst_xyz[t][p * a.ld_xyz[t] + 6 + 6 * k + c] = co;
        }
      }
      if (a.dst_dir) {
        a.dst_dir[p * a.ld_dir + c] = d;
        for (int k = 0; k < a.n_freq_dir; ++k) {
          float s, co; sincosf(__fmul_rn(band_f(k, a.n_freq_dir, a.logscale), d), &s, &co);
          a.dst_dir[p * a.ld_dir + 3 + 6 * k + c] = s;
          a.dst_dir[p * a.ld_dir + 6 + 6 * k + c] = co;
        }
      }
    } else {   // EMBEDDED: plain column copies, thread c strides over the columns
      const int width = a.in_xyz + (a.dst_dir ? a.in_dir : 0);
      const float* src = a.in0 + p * (int64_t)(a.in_xyz + (a.dst_dir ? a.in_dir : 0));
      if (a.siren) {
        if (c < 3) a.dst_xyz[0][p * a.ld_xyz[0] + c] = src[c] * a.scale;
        if (a.dst_dir) a.dst_dir[p * a.ld_dir + c] = src[3 + c];
        continue;
      }
      for (int j = c; j < width; j += 3) {
        const float v = __ldg(src + j);
        if (j < a.in_xyz) { for (int t = 0; t < a.n_xyz_dst; ++t) a.dst_xyz[t][p * a.ld_xyz[t] + j] = v; }
        else a.dst_dir[p * a.ld_dir + (j - a.in_xyz)] = v;
      }
    }
  }
}

// ---- workspace plan ------------------------------------------------------------------------------------
struct SimtPlan {
  // forward activations (floats offsets into the workspace)
  int64_t enc0;                 // (P, in_xyz)                  NeRF;  SIREN: (P,3) warped points
  int64_t hbuf[16]; int64_t hld[16]; int hcol[16];   // output of trunk layer i lives at hbuf[i] + hcol[i], row stride hld[i]
  int64_t zpre[16];             // SIREN pre-activations (P,W) per trunk layer
  int64_t dirbuf, dirld;        // NeRF: (P, W+in_dir) [feat | dir enc];  SIREN: (P, 3+W) [dir | h_last]
  int64_t hd;                   // NeRF: (P, W/2);  SIREN: hc (P, W)
  int64_t zc;                   // SIREN colour pre-activation (P, W)
  int64_t p0;                   // NeRF_3D parse.0 output (P, 2W)
  // backward scratch
  int64_t ga, gb;               // two (P, W + max(in_xyz, in_dir, 3)) gradient buffers
  int64_t gsmall;               // (P, 16) head pre-activation grads
  int64_t gp0;                  // (P, 2W)
  int64_t total;
};

static SimtPlan simt_plan(const nerfb_mlp_desc& d, int64_t P, bool sigma_only, bool bwd) {
  SimtPlan pl{};
  const NerfOffsets o = nerf_offsets(d);
  const int W = d.W;
  int64_t p = 0;
  auto take = [&](int64_t n) { int64_t at = p; p += round_up64(n, 64); return at; };
  const bool siren = d.kind == NERFB_KIND_SIREN;
  const int in_xyz = siren ? 3 : o.in_xyz;
  const int in_dir = siren ? 3 : o.in_dir;
  pl.enc0 = take(P * in_xyz);
  if (!sigma_only) {
    pl.dirld = siren ? (3 + W) : (W + in_dir);
    pl.dirbuf = take(P * pl.dirld);
  }
  for (int i = 0; i < d.D; ++i) {
    const bool next_skip = !siren && (i + 1 < d.D) && ((d.skip_mask >> (i + 1)) & 1);
    const bool last = (i == d.D - 1);
    if (siren && last && !sigma_only) {          // h_last is produced straight into [dir | h] (nerf.py:213)
      pl.hbuf[i] = pl.dirbuf; pl.hld[i] = pl.dirld; pl.hcol[i] = 3;
    } else if (next_skip) {                      // produced straight into [xyz_enc | h] (nerf.py:109)
      pl.hbuf[i] = take(P * (int64_t)(in_xyz + W)); pl.hld[i] = in_xyz + W; pl.hcol[i] = in_xyz;
    } else {
      pl.hbuf[i] = take(P * (int64_t)W); pl.hld[i] = W; pl.hcol[i] = 0;
    }
    if (siren) pl.zpre[i] = take(P * (int64_t)W);
  }
  if (!sigma_only) {
    pl.hd = take(P * (int64_t)(siren ? W : W / 2));
    if (siren) pl.zc = take(P * (int64_t)W);
    if (d.n_cls > 0) pl.p0 = take(P * (int64_t)2 * W);
  }
  if (bwd) {
    const int64_t gw = W + (in_xyz > in_dir ? in_xyz : in_dir);
    pl.ga = take(P * gw); pl.gb = take(P * gw);
    pl.gsmall = take(P * 16);
    if (d.n_cls > 0) pl.gp0 = take(P * (int64_t)2 * W);
  }
  pl.total = p;
  return pl;
}

int64_t simt_workspace_bytes(const nerfb_mlp_desc& d, int64_t P, int sigma_only, int save) {
  return simt_plan(d, P, sigma_only != 0, save != 0).total * (int64_t)sizeof(float);
}

static int simt_prep(const nerfb_mlp_desc& d, const SimtPlan& pl, float* ws, int in_mode, const float* in0,
                     const float* in1, int64_t P, int S, bool sigma_only, void* stream) {
  const NerfOffsets o = nerf_offsets(d);
  const bool siren = d.kind == NERFB_KIND_SIREN;
  PrepArgs a{};
  a.in0 = in0; a.in1 = in1; a.in_mode = in_mode; a.P = P; a.S = S;
  a.n_freq_xyz = d.n_freq_xyz; a.n_freq_dir = d.n_freq_dir; a.logscale = d.logscale;
  a.in_xyz = siren ? 3 : o.in_xyz; a.in_dir = siren ? 3 : o.in_dir;
  a.siren = siren; a.scale = siren ? 2.0f / d.box : 1.0f;
  a.dst_xyz[0] = ws + pl.enc0; a.ld_xyz[0] = a.in_xyz; a.n_xyz_dst = 1;
  if (!siren)
    for (int i = 1; i < d.D; ++i)
      if ((d.skip_mask >> i) & 1) {
        if (a.n_xyz_dst >= 4) return NERFB_E_UNSUPPORTED;
        a.dst_xyz[a.n_xyz_dst] = ws + pl.hbuf[i - 1]; a.ld_xyz[a.n_xyz_dst] = pl.hld[i - 1]; a.n_xyz_dst++;
      }
  if (!sigma_only) { a.dst_dir = ws + pl.dirbuf + (siren ? 0 : d.W); a.ld_dir = pl.dirld; }
  const int64_t want = ceil_div64(P * 3, 256);
  const int grid = (int)(want < NERFB_SM_COUNT * 8 ? want : NERFB_SM_COUNT * 8);
  NERFB_LAUNCH(prep_inputs_kernel, grid, 256, 0, stream, a);
  return NERFB_OK;
}

// ---- forward -------------------------------------------------------------------------------------------
int simt_mlp_fwd(const nerfb_mlp_desc& d, int in_mode, const float* params, const float* film, const float* in0,
                 const float* in1, int64_t n_rows, int n_samples, int sigma_only, int save, float* out, float* ws,
                 int64_t ws_bytes, void* stream) {
  const int64_t P = (in_mode == NERFB_IN_RAYS) ? n_rows * n_samples : n_rows;
  const SimtPlan pl = simt_plan(d, P, sigma_only != 0, save != 0);
  if (ws_bytes < pl.total * (int64_t)sizeof(float)) return NERFB_E_WORKSPACE;
  const NerfOffsets o = nerf_offsets(d);
  const int W = d.W;
  const bool siren = d.kind == NERFB_KIND_SIREN;
  if (siren && !film) return NERFB_E_BADARG;
  const int C = sigma_only ? 1 : 4 + d.n_cls;
  int st = simt_prep(d, pl, ws, in_mode, in0, in1, P, n_samples, sigma_only != 0, stream);
  if (st != NERFB_OK) return st;
  const float* freq = film;                         // (D+1)*W effective frequencies 15 f + 30 (nerf.py:202)
  const float* phase = film ? film + (int64_t)(d.D + 1) * W : nullptr;

  const float* x = ws + pl.enc0; int64_t ldx = siren ? 3 : o.in_xyz;
  for (int i = 0; i < d.D; ++i) {
    const bool skip = !siren && i > 0 && ((d.skip_mask >> i) & 1);
    if (skip) { x = ws + pl.hbuf[i - 1]; ldx = pl.hld[i - 1]; }              // [xyz_enc | h_{i-1}]
    Gemm g{P, W, o.in[i], x, ldx, 1, params + o.w[i], 1, o.in[i], ws + pl.hbuf[i] + pl.hcol[i], pl.hld[i]};
    g.bias = params + o.b[i];
    if (siren) { g.act = ACT_SINE; g.freq = freq + (int64_t)i * W; g.phase = phase + (int64_t)i * W;
                 g.zpre = ws + pl.zpre[i]; g.ldz = W; }
    else g.act = ACT_RELU;
    RUN(g);
    x = ws + pl.hbuf[i] + pl.hcol[i]; ldx = pl.hld[i];
  }
  {   // sigma head (nerf.py:112 / :212) -> out[:, C-1 or 3]
    Gemm g{P, 1, W, x, ldx, 1, params + o.w_sigma, 1, W, out + (sigma_only ? 0 : 3), C};
    g.bias = params + o.b_sigma;
    RUN(g);
  }
  if (sigma_only) return NERFB_OK;
  if (!siren) {
    Gemm gf{P, W, W, x, ldx, 1, params + o.w_final, 1, W, ws + pl.dirbuf, pl.dirld};     // feat -> [feat | dir]
    gf.bias = params + o.b_final; RUN(gf);
    Gemm gd{P, W / 2, W + o.in_dir, ws + pl.dirbuf, pl.dirld, 1, params + o.w_dir, 1, W + o.in_dir, ws + pl.hd, W / 2};
    gd.bias = params + o.b_dir; gd.act = ACT_RELU; RUN(gd);
    Gemm gr{P, 3, W / 2, ws + pl.hd, W / 2, 1, params + o.w_rgb, 1, W / 2, out, C};
    gr.bias = params + o.b_rgb; gr.act = ACT_SIGMOID; RUN(gr);
    if (d.n_cls > 0) {                                                                   // nerf_cls.py:56-61,104
      Gemm g0{P, 2 * W, W, ws + pl.dirbuf, pl.dirld, 1, params + o.w_p0, 1, W, ws + pl.p0, 2 * W};
      g0.bias = params + o.b_p0; RUN(g0);
      Gemm g1{P, d.n_cls, 2 * W, ws + pl.p0, 2 * W, 1, params + o.w_p1, 1, 2 * W, out + 4, C};
      g1.bias = params + o.b_p1; g1.act = ACT_SIGMOID; RUN(g1);
    }
  } else {
    Gemm gc{P, W, W + 3, ws + pl.dirbuf, pl.dirld, 1, params + o.w_dir, 1, W + 3, ws + pl.hd, W};   // nerf.py:213
    gc.bias = params + o.b_dir; gc.act = ACT_SINE; gc.freq = freq + (int64_t)d.D * W; gc.phase = phase + (int64_t)d.D * W;
    gc.zpre = ws + pl.zc; gc.ldz = W; RUN(gc);
    Gemm gr{P, 3, W, ws + pl.hd, W, 1, params + o.w_rgb, 1, W, out, C};
    gr.bias = params + o.b_rgb; gr.act = ACT_SIGMOID; RUN(gr);
  }
  return NERFB_OK;
}

// ---- backward ------------------------------------------------------------------------------------------
// dW (n_out x n_in) += dZ^T (P x n_out) . A (P x n_in) ; db += colsum(dZ)
static int wgrad(const float* dz, int64_t lddz, int n_out, const float* a, int64_t lda, int n_in, int64_t P, float* gw,
                 float* gb, void* stream) {
  Gemm g{n_out, n_in, P, dz, 1, lddz, a, lda, 1, gw, n_in};
  g.split_k = 1;
  RUN(g);
  return run_colsum(dz, lddz, P, n_out, gb, stream);
}
#define TRY(expr) do { int _s = (expr); if (_s != NERFB_OK) return _s; } while (0)

int simt_mlp_bwd(const nerfb_mlp_desc& d, int in_mode, const float* params, const float* film, const float* in0,
                 const float* in1, int64_t n_rows, int n_samples, const float* out, const float* d_out,
                 float* grad, float* d_in, float* ws, int64_t ws_bytes, void* stream) {
  const int64_t P = (in_mode == NERFB_IN_RAYS) ? n_rows * n_samples : n_rows;
  const SimtPlan pl = simt_plan(d, P, false, true);
  if (ws_bytes < pl.total * (int64_t)sizeof(float)) return NERFB_E_WORKSPACE;
  const NerfOffsets o = nerf_offsets(d);
  const int W = d.W;
  const bool siren = d.kind == NERFB_KIND_SIREN;
  if (siren && d_in) return NERFB_E_UNSUPPORTED;      // SIREN input gradients (d xyz, d dir) are not implemented: refuse, never return zeros
  const int C = 4 + d.n_cls;
  const float* freq = film;
  const float* phase = film ? film + (int64_t)(d.D + 1) * W : nullptr;
  float* ga = ws + pl.ga; float* gb = ws + pl.gb; float* gs = ws + pl.gsmall;
  const int in_xyz = siren ? 3 : o.in_xyz, in_dir = siren ? 3 : o.in_dir;
  const int64_t gld = W + (in_xyz > in_dir ? in_xyz : in_dir);
  auto grid_for = [](int64_t n) { int64_t w = ceil_div64(n, 256); return (int)(w < NERFB_SM_COUNT * 8 ? w : NERFB_SM_COUNT * 8); };

  const int last = d.D - 1;
  const float* h_last = ws + pl.hbuf[last] + pl.hcol[last]; const int64_t ld_last = pl.hld[last];
  float* dh = nullptr;      // gradient w.r.t. the last trunk activation, (P, W) inside ga/gb with row stride gld

  // rgb head: dz_rgb = d_out[:, :3] * y (1-y)
  NERFB_LAUNCH(window_act_bwd_kernel, grid_for(P * 3), 256, 0, stream, ACT_SIGMOID, d_out, (int64_t)C, 0, out, (int64_t)C, 0,
               gs, (int64_t)16, P, 3);
  if (!siren) {
    const float* hd = ws + pl.hd;
    TRY(wgrad(gs, 16, 3, hd, W / 2, W / 2, P, grad + o.w_rgb, grad + o.b_rgb, stream));
    Gemm g1{P, W / 2, 3, gs, 16, 1, params + o.w_rgb, W / 2, 1, ga, gld}; RUN(g1);             // d_hd
    TRY(run_act_bwd(ACT_RELU, ga, gld, hd, W / 2, P, W / 2, nullptr, nullptr, stream));        // dz_dir
    TRY(wgrad(ga, gld, W / 2, ws + pl.dirbuf, pl.dirld, W + o.in_dir, P, grad + o.w_dir, grad + o.b_dir, stream));
    const int n_back = d_in ? W + o.in_dir : W;
    Gemm g2{P, n_back, W / 2, ga, gld, 1, params + o.w_dir, W + o.in_dir, 1, gb, gld}; RUN(g2); // d[feat | dir]
    if (d_in) {
      NERFB_LAUNCH(window_act_bwd_kernel, grid_for(P * o.in_dir), 256, 0, stream, ACT_NONE, gb, gld, W, (const float*)nullptr,
                   (int64_t)0, 0, d_in + o.in_xyz, (int64_t)(o.in_xyz + o.in_dir), P, o.in_dir);
    }
    if (d.n_cls > 0) {                                                                          // class head
      NERFB_LAUNCH(window_act_bwd_kernel, grid_for(P * d.n_cls), 256, 0, stream, ACT_SIGMOID, d_out, (int64_t)C, 4, out,
                   (int64_t)C, 4, gs, (int64_t)16, P, d.n_cls);
      if (d.n_cls > 16) return NERFB_E_UNSUPPORTED;
      TRY(wgrad(gs, 16, d.n_cls, ws + pl.p0, 2 * W, 2 * W, P, grad + o.w_p1, grad + o.b_p1, stream));
      float* gp0 = ws + pl.gp0;
      Gemm g3{P, 2 * W, d.n_cls, gs, 16, 1, params + o.w_p1, 2 * W, 1, gp0, 2 * W}; RUN(g3);
      TRY(wgrad(gp0, 2 * W, 2 * W, ws + pl.dirbuf, pl.dirld, W, P, grad + o.w_p0, grad + o.b_p0, stream));
      Gemm g4{P, W, 2 * W, gp0, 2 * W, 1, params + o.w_p0, W, 1, gb, gld}; g4.accumulate = 1; RUN(g4);   // d_feat +=
    }
    // final layer (no activation): gb[:, :W] = d_feat
    TRY(wgrad(gb, gld, W, h_last, ld_last, W, P, grad + o.w_final, grad + o.b_final, stream));
    Gemm g5{P, W, W, gb, gld, 1, params + o.w_final, W, 1, ga, gld}; RUN(g5);                   // d_h_last
    dh = ga;
  } else {
    const float* hc = ws + pl.hd;
    TRY(wgrad(gs, 16, 3, hc, W, W, P, grad + o.w_rgb, grad + o.b_rgb, stream));
    Gemm g1{P, W, 3, gs, 16, 1, params + o.w_rgb, W, 1, ga, gld}; RUN(g1);                      // d_hc
    TRY(run_act_bwd(ACT_SINE, ga, gld, ws + pl.zc, W, P, W, freq + (int64_t)d.D * W, phase + (int64_t)d.D * W, stream));
    TRY(wgrad(ga, gld, W, ws + pl.dirbuf, pl.dirld, W + 3, P, grad + o.w_dir, grad + o.b_dir, stream));
    Gemm g2{P, W, W, ga, gld, 1, params + o.w_dir + 3, W + 3, 1, gb, gld}; RUN(g2);             // d_h_last (cols 3..)
    dh = gb;
  }
  // sigma head: d_h_last += d_sigma (x) w_sigma ; dw_sigma += d_sigma^T h_last
  {
    Gemm g6{P, W, 1, d_out + 3, C, 1, params + o.w_sigma, W, 1, dh, gld}; g6.accumulate = 1; RUN(g6);
    TRY(wgrad(d_out + 3, C, 1, h_last, ld_last, W, P, grad + o.w_sigma, grad + o.b_sigma, stream));
  }
  // trunk, last layer to first
  float* other = (dh == ga) ? gb : ga;
  bool xyz_written = false;   // d_in[:, :in_xyz] already holds a skip layer's contribution
  for (int i = last; i >= 0; --i) {
    const float* h_i = ws + pl.hbuf[i] + pl.hcol[i];
    if (siren) TRY(run_act_bwd(ACT_SINE, dh, gld, ws + pl.zpre[i], W, P, W, freq + (int64_t)i * W, phase + (int64_t)i * W, stream));
    else TRY(run_act_bwd(ACT_RELU, dh, gld, h_i, pl.hld[i], P, W, nullptr, nullptr, stream));
    const bool skip = !siren && i > 0 && ((d.skip_mask >> i) & 1);
    const float* x; int64_t ldx;
    if (i == 0) { x = ws + pl.enc0; ldx = in_xyz; }
    else if (skip) { x = ws + pl.hbuf[i - 1]; ldx = pl.hld[i - 1]; }
    else { x = ws + pl.hbuf[i - 1] + pl.hcol[i - 1]; ldx = pl.hld[i - 1]; }
    TRY(wgrad(dh, gld, W, x, ldx, o.in[i], P, grad + o.w[i], grad + o.b[i], stream));
    if (i == 0) {
      if (d_in && !siren) {
        Gemm gi{P, o.in_xyz, W, dh, gld, 1, params + o.w[0], o.in[0], 1, d_in, (int64_t)(o.in_xyz + o.in_dir)};
        gi.accumulate = xyz_written ? 1 : 0;          // skip layers already deposited their xyz part
        RUN(gi);
      }
      break;
    }
    const int col0 = skip ? o.in_xyz : 0;
    Gemm gn{P, W, W, dh, gld, 1, params + o.w[i] + col0, o.in[i], 1, other, gld}; RUN(gn);      // d_h_{i-1}
    if (skip && d_in) {   // gradient reaching the xyz encoding through the skip connection
      Gemm gx{P, o.in_xyz, W, dh, gld, 1, params + o.w[i], o.in[i], 1, d_in, (int64_t)(o.in_xyz + o.in_dir)};
      gx.accumulate = xyz_written ? 1 : 0;
      RUN(gx);
      xyz_written = true;
    }
    float* t = dh; dh = other; other = t;
  }
  return NERFB_OK;
}
