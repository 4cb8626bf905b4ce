// Kernel (6): the EG3D tri-plane renderer (BASELINE config 5).
//   replaces volumetric_rendering/renderer.py:57-68 (sample_from_planes: 3 x grid_sample), eg3d_training/triplane.py:
//   144-167 (OSGDecoder 32 -> 64 softplus -> 4), volumetric_rendering/ray_marcher.py:25-57 (MipRayMarcher2) and
//   renderer.py:94-142,160-256 (ImportanceRenderer: stratified depths, smoothed-weight importance sampling, merge sort).
//
// One warp per ray does the WHOLE pipeline; nothing per-sample ever reaches HBM (the reference materialises
// coordinates, (1,3,M*S,32) features, colours, densities, three marcher passes and a sort + 3 gathers).
//   * planes are re-laid out once per call to channels-last (3, H, W, 32): a texel is ONE 128-byte line and lane = channel,
//     so each of the 12 bilinear taps of a sample is a single coalesced request served by L2 (the 25 MB of planes stay
//     resident in the 126 MB L2).  Bound: L2 gather bandwidth, 1,536 B per sample.
//   * decoder on CUDA cores with its 2 x 64 x 32 first-layer weights in registers (lane j owns hidden units j, j + 32),
//     features broadcast through shared memory (8 LDS.128); 2,304 MACs per sample is far below the gather cost.
//   * marcher = warp product-scan over the S-1 intervals with a running carry (same scheme as composite.cu).
//   * the chunk-global depth clamp (ray_marcher.py:49-50: torch.min / torch.max over ALL depths) is two ordered-int
//     atomics per warp plus a tiny second kernel.
#include "common.cuh"
#include "resample_dev.cuh"

namespace {

constexpr int TP_WARPS = 8;
constexpr int TP_C = 32;          // feature channels (lane = channel)
constexpr int TP_HID = 64;
constexpr int TP_MAX_S = 256;     // samples per ray after the merge

// ---- channels-last re-layout: (3, C=32, H, W) -> (3, H, W, 32) ----------------------------------------------------
__global__ void __launch_bounds__(256) triplane_relayout_kernel(const float* __restrict__ in, float* __restrict__ out, int64_t HW) {
  __shared__ float tile[32][33];
  const int p = blockIdx.y;
  const int64_t hw0 = (int64_t)blockIdx.x * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;      // 32 x 8
  for (int c = ty; c < 32; c += 8) {
    const int64_t hw = hw0 + tx;
    tile[c][tx] = hw < HW ? __ldg(in + ((int64_t)p * 32 + c) * HW + hw) : 0.f;
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {
    const int64_t hw = hw0 + r;
    if (hw < HW) out[((int64_t)p * HW + hw) * 32 + tx] = tile[tx][r];
  }
}

// ---- decoder weights held per lane --------------------------------------------------------------------------------
struct Dec {
  float w1a[TP_C], w1b[TP_C];    // rows lane and lane + 32 of W1 (already times the FullyConnectedLayer gain)
  float b1a, b1b;
  float w2a[4], w2b[4];          // columns lane and lane + 32 of W2
  float b2;                      // bias of the output this lane finalises (lanes 0 / 8 / 16 / 24)
};

// params: W1 (64 x 32) | b1 (64) | W2 (4 x 64) | b2 (4), raw nn.Parameters; gains as networks_stylegan2.py:109-110
__device__ __forceinline__ void load_decoder(Dec& d, const float* __restrict__ p, float lr_mul, int lane) {
  const float g1 = lr_mul / sqrtf((float)TP_C), g2 = lr_mul / sqrtf((float)TP_HID);
#pragma unroll
  for (int c = 0; c < TP_C; ++c) {
    d.w1a[c] = __fmul_rn(__ldg(p + lane * TP_C + c), g1);
    d.w1b[c] = __fmul_rn(__ldg(p + (lane + 32) * TP_C + c), g1);
  }
  const float* b1 = p + TP_HID * TP_C;
  d.b1a = __ldg(b1 + lane) * lr_mul; d.b1b = __ldg(b1 + lane + 32) * lr_mul;
  const float* w2 = b1 + TP_HID;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    d.w2a[k] = __fmul_rn(__ldg(w2 + k * TP_HID + lane), g2);
    d.w2b[k] = __fmul_rn(__ldg(w2 + k * TP_HID + lane + 32), g2);
  }
  d.b2 = __ldg(w2 + 4 * TP_HID + (lane >> 3)) * lr_mul;
}

// F.softplus (beta 1, threshold 20) in the overflow-safe form max(x,0) + log(1 + exp(-|x|)); the fast exp / log
// intrinsics keep the ABSOLUTE error below 2e-7 (1 + t is formed exactly to fp32, t in (0,1]), far inside the stated
// parity tolerance, at a quarter of the instructions of log1pf(expf(x)).
__device__ __forceinline__ float softplusf(float x) {
  // beyond the threshold the second term is below half an ulp of x (log(1 + e^-20) = 2e-9 against ulp(20) = 1.9e-6): the sum
  // already IS x in fp32, no select needed
  // six instructions: ex2 / lg2 approximations straight (the arguments need none of __expf / __logf's range handling:
  // -|x| log2(e) <= 0 and 1 + t in (1, 2]), the final scale folded into one FMA
  float t, l;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(-fabsf(x) * 1.4426950408889634f));
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l) : "f"(1.f + t));
  return fmaf(l, 0.6931471805599453f, fmaxf(x, 0.f));
}

// One bilinear tap per lane (lanes 0..11 = plane * 4 + corner) at normalised (gx -> W, gy -> H), zeros padding,
// align_corners=False (F.grid_sample semantics, renderer.py:64).  The offset is clamped in-bounds and an out-of-range tap
// gets weight 0, so the loads are unconditional and all 12 of a sample are in flight together; NaN coordinates fail every
// test -> no tap.  Computing the 12 taps on 12 lanes instead of all of them on every lane removes ~85 of the ~110 address
// instructions per sample; the (offset, weight) pairs go through shared memory (1 STS + 6 broadcast LDS.128).
__device__ __forceinline__ void tap_lane(int H, int W, float x, float y, float z, int lane, int& off, float& wt) {
  const int p = min(lane >> 2, 2), corner = lane & 3;
  // plane axes (renderer.py:29-39): projections (x, y), (x, z), (z, x)
  const float gx = p == 2 ? z : x;
  const float gy = p == 0 ? y : (p == 1 ? z : x);
  const float ix = __fmul_rn(__fsub_rn(__fmul_rn(__fadd_rn(gx, 1.f), (float)W), 1.f), 0.5f);
  const float iy = __fmul_rn(__fsub_rn(__fmul_rn(__fadd_rn(gy, 1.f), (float)H), 1.f), 0.5f);
  const float fx = floorf(ix), fy = floorf(iy);
  const float w = ix - fx, e = 1.f - w, n = iy - fy, s = 1.f - n;
  const bool right = corner & 1, down = corner & 2;
  const float cx = right ? fx + 1.f : fx, cy = down ? fy + 1.f : fy;
  const bool valid = cx >= 0.f && cx < (float)W && cy >= 0.f && cy < (float)H;
  wt = valid ? (down ? n : s) * (right ? w : e) : 0.f;
  const int xi = min(max((int)fmaxf(fminf(cx, (float)W), -2.f), 0), W - 1);
  const int yi = min(max((int)fmaxf(fminf(cy, (float)H), -2.f), 0), H - 1);
  off = ((p * H + yi) * W + xi) * TP_C;
}

// ---- the decoder on the tensor cores (mma.sync m16n8k8 TF32, 3-term split: fp32-grade accuracy) -----------------------
// A mat-vec on CUDA cores needs every feature on every lane, and a warp-uniform shared load (or a shuffle) delivers ONE useful
// word per wavefront of the shared-memory data pipe; mma fragments are distributed over the lanes, nothing is broadcast.
// A warp evaluates 16 points at a time: pre^T [16 x 64] = F^T [16 x 32] . W1^T, A from the per-warp feature tile, B from the
// block's copy of W1, C in registers (rows pt = lane / 4, + 8; columns hid = 8 nt + 2 (lane % 4), + 1).
constexpr int TM_PTS = 16;
constexpr int TM_FS = 36;      // features [pt][ch]: stride = 4 (mod 32): (pt over lane / 4, ch over lane % 4) is conflict-free
constexpr int TM_W1 = 36;      // W1 [hid][ch]
constexpr int TM_W1_FLOATS = 2 * TP_HID * TM_W1;                          // W1 split in its two TF32 terms: hi [hid][TM_W1] | lo
constexpr int TM_DEC_FLOATS = TM_W1_FLOATS + 4 * TP_HID + TP_HID;         // W1s (hi | lo) | W2s [k][hid] | B1s
__device__ __forceinline__ void tf32_split(float x, uint32_t& hi, uint32_t& lo) {      // round-to-nearest terms (weights, once per block)
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hi) : "f"(x));
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(lo) : "f"(x - __uint_as_float(hi)));
}
// per-sample operands: hi = x with the 13 low mantissa bits cleared, lo = x - hi (exact; the tensor core ignores ITS low bits):
// two instructions instead of the five of the rounding split (cvt.rna.tf32 is emulated with integer ops), error <= 2^-20 |x|
__device__ __forceinline__ void tf32_split_fast(float x, uint32_t& hi, uint32_t& lo) {
  hi = __float_as_uint(x) & 0xFFFFE000u;
  lo = __float_as_uint(x - __uint_as_float(hi));
}
__device__ __forceinline__ void mma_tf32(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// c += a . b with both operands split in two TF32 terms
__device__ __forceinline__ void mma_3x(float (&c)[4], const uint32_t (&ah)[4], const uint32_t (&al)[4], float b0, float b1) {
  uint32_t b0h, b0l, b1h, b1l;
  tf32_split_fast(b0, b0h, b0l); tf32_split_fast(b1, b1h, b1l);
  mma_tf32(c, al, b0h, b1h); mma_tf32(c, ah, b0l, b1l); mma_tf32(c, ah, b0h, b1h);
}


// the same with B = two entries of the pre-split W1 copy in shared memory (p and p + second; lo term TP_HID * TM_W1 further)
__device__ __forceinline__ void mma_3x_w1(float (&c)[4], const uint32_t (&ah)[4], const uint32_t (&al)[4], const float* p, int second) {
  const uint32_t b0h = __float_as_uint(p[0]), b1h = __float_as_uint(p[second]);
  const uint32_t b0l = __float_as_uint(p[TP_HID * TM_W1]), b1l = __float_as_uint(p[TP_HID * TM_W1 + second]);
  mma_tf32(c, al, b0h, b1h); mma_tf32(c, ah, b0l, b1l); mma_tf32(c, ah, b0h, b1h);
}

// block-wide: the decoder's effective weights (gains of networks_stylegan2.py:109-110) into shared memory
__device__ __forceinline__ void load_decoder_mma(float* W1s, const float* __restrict__ decp, float lr_mul) {
  float* W2s = W1s + TM_W1_FLOATS;
  float* B1s = W2s + 4 * TP_HID;
  const float g1 = lr_mul / sqrtf((float)TP_C), g2 = lr_mul / sqrtf((float)TP_HID);
  const float* b1p = decp + TP_HID * TP_C;
  const float* w2p = b1p + TP_HID;
  for (int i = threadIdx.x; i < TP_HID * TP_C; i += blockDim.x) {
    uint32_t hi, lo;
    tf32_split(__fmul_rn(__ldg(decp + i), g1), hi, lo);
    W1s[(i >> 5) * TM_W1 + (i & 31)] = __uint_as_float(hi);
    W1s[TP_HID * TM_W1 + (i >> 5) * TM_W1 + (i & 31)] = __uint_as_float(lo);
  }
  for (int i = threadIdx.x; i < 4 * TP_HID; i += blockDim.x) W2s[i] = __fmul_rn(__ldg(w2p + i), g2);
  for (int i = threadIdx.x; i < TP_HID; i += blockDim.x) B1s[i] = __ldg(b1p + i) * lr_mul;
}

// The four taps of plane p for a lane that needs them all (tap_lane's arithmetic, with what the corners share computed once:
// two column / two row validity tests and clamps instead of eight).  Corner order as tap_lane: (x0,y0) (x1,y0) (x0,y1) (x1,y1).
__device__ __forceinline__ void taps_plane(int H, int W, float x, float y, float z, int p, int (&off)[4], float (&wt)[4]) {
  const float gx = p == 2 ? z : x;
  const float gy = p == 0 ? y : (p == 1 ? z : x);
  const float ix = __fmul_rn(__fsub_rn(__fmul_rn(__fadd_rn(gx, 1.f), (float)W), 1.f), 0.5f);
  const float iy = __fmul_rn(__fsub_rn(__fmul_rn(__fadd_rn(gy, 1.f), (float)H), 1.f), 0.5f);
  const float fx = floorf(ix), fy = floorf(iy);
  const float w = ix - fx, e = 1.f - w, n = iy - fy, s = 1.f - n;
  const float x1 = fx + 1.f, y1 = fy + 1.f;
  const bool vx0 = fx >= 0.f && fx < (float)W, vx1 = x1 >= 0.f && x1 < (float)W;      // NaN fails every test: no tap
  const bool vy0 = fy >= 0.f && fy < (float)H, vy1 = y1 >= 0.f && y1 < (float)H;
  // offsets only have to be in bounds (an invalid tap has weight 0): one sanitised float -> int per axis, the neighbour is + 1
  const int ix0 = (int)fmaxf(fminf(fx, (float)W), -2.f), iy0 = (int)fmaxf(fminf(fy, (float)H), -2.f);
  const int xi0 = min(max(ix0, 0), W - 1), xi1 = min(max(ix0 + 1, 0), W - 1);
  const int yi0 = min(max(iy0, 0), H - 1), yi1 = min(max(iy0 + 1, 0), H - 1);
  const int r0 = (p * H + yi0) * W, r1 = (p * H + yi1) * W;
  off[0] = (r0 + xi0) * TP_C; off[1] = (r0 + xi1) * TP_C; off[2] = (r1 + xi0) * TP_C; off[3] = (r1 + xi1) * TP_C;
  // (vx && vy) ? row * col : 0 as a product of per-axis masked factors (an invalid axis contributes an exact 0)
  const float cx0 = vx0 ? e : 0.f, cx1 = vx1 ? w : 0.f, ry0 = vy0 ? s : 0.f, ry1 = vy1 ? n : 0.f;
  wt[0] = ry0 * cx0; wt[1] = ry0 * cx1; wt[2] = ry1 * cx0; wt[3] = ry1 * cx1;
}

// gather of ONE point by 8 lanes (sub4 = 4 (lane % 8): four channels each): all 12 taps computed by the lane itself, rows read
// as float4; same tap order and rounding as field_eval except the final mean (times 1/3 instead of the IEEE division by 3: the
// tensor-core decoder is not bit-identical to the CUDA-core one anyway).  Returns sampled_features.mean(1) for the lane's four channels.
__device__ __forceinline__ float4 gather_quad(const float* __restrict__ planes, int H, int W, float x, float y, float z, int sub4) {
  float4 f = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int pl = 0; pl < 3; ++pl) {
    int off[4]; float wt[4];
    taps_plane(H, W, x, y, z, pl, off, wt);
    float2 a01 = make_float2(0.f, 0.f), a23 = make_float2(0.f, 0.f);          // two-wide FMAs: same IEEE results, half the instructions
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const float4 v = __ldg(reinterpret_cast<const float4*>(planes + off[c] + sub4));
      const float2 ww = make_float2(wt[c], wt[c]);
      a01 = __ffma2_rn(make_float2(v.x, v.y), ww, a01); a23 = __ffma2_rn(make_float2(v.z, v.w), ww, a23);
    }
    f.x += a01.x; f.y += a01.y; f.z += a23.x; f.w += a23.y;
  }
  const float third = 1.f / 3.f;
  f.x *= third; f.y *= third; f.z *= third; f.w *= third;
  return f;
}

// run_model for the 16 points whose features sit in Fs[pt][TM_FS] (written by gather_quad, __syncwarp'ed by the caller).
// out[u][0] = sigma, out[u][1..3] = rgb of point (lane / 4) + 8 u, identical on the four lanes of a group.
__device__ __forceinline__ void decode16(const float* Fs, const float* W1s, const float (&b2)[4], int lane, float (&out)[2][4]) {
  const float* W2s = W1s + TM_W1_FLOATS;
  const float* B1s = W2s + 4 * TP_HID;
  const int g = lane >> 2, t = lane & 3;
  float pre[8][4];
#pragma unroll
  for (int nt = 0; nt < 8; ++nt) {
    const float2 b = *reinterpret_cast<const float2*>(B1s + 8 * nt + 2 * t);
    pre[nt][0] = b.x; pre[nt][1] = b.y; pre[nt][2] = b.x; pre[nt][3] = b.y;
  }
#pragma unroll
  for (int kt = 0; kt < 4; ++kt) {
    uint32_t ah[4], al[4];
    tf32_split_fast(Fs[g * TM_FS + 8 * kt + t], ah[0], al[0]);
    tf32_split_fast(Fs[(g + 8) * TM_FS + 8 * kt + t], ah[1], al[1]);
    tf32_split_fast(Fs[g * TM_FS + 8 * kt + t + 4], ah[2], al[2]);
    tf32_split_fast(Fs[(g + 8) * TM_FS + 8 * kt + t + 4], ah[3], al[3]);
#pragma unroll
    for (int nt = 0; nt < 8; ++nt)
      mma_3x_w1(pre[nt], ah, al, W1s + (8 * nt + g) * TM_W1 + 8 * kt + t, 4);
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) { out[0][k] = 0.f; out[1][k] = 0.f; }
#pragma unroll
  for (int nt = 0; nt < 8; ++nt) {
    float h[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = softplusf(pre[nt][i]);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float2 w = *reinterpret_cast<const float2*>(W2s + k * TP_HID + 8 * nt + 2 * t);
      float2 o2 = make_float2(out[0][k], out[1][k]);                         // both points of the lane in one two-wide FMA
      o2 = __ffma2_rn(make_float2(w.y, w.y), make_float2(h[1], h[3]), o2);
      o2 = __ffma2_rn(make_float2(w.x, w.x), make_float2(h[0], h[2]), o2);
      out[0][k] = o2.x; out[1][k] = o2.y;
    }
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) {
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      float v = out[u][k];
      v += __shfl_xor_sync(0xffffffffu, v, 1);
      v += __shfl_xor_sync(0xffffffffu, v, 2);
      v += b2[k];
      if (k > 0) v = __fsub_rn(__fmul_rn(__fdividef(1.f, 1.f + __expf(-v)), 1.002f), 0.001f);     // triplane.py:165
      out[u][k] = v;
    }
  }
}

// run_model for one point (renderer.py:144-151 + triplane.py:155-167).  Returns on lane 0: sigma, lane 8 / 16 / 24: r, g, b.
// s_tap: 12 x (offset, weight) as int2, s_feat: 32 floats; the caller alternates two such buffers between consecutive samples.
__device__ __forceinline__ float field_eval(const float* __restrict__ planes, int H, int W, float x, float y, float z,
                                            const Dec& d, int lane, int2* s_tap, float* s_feat) {
  int off; float wt;
  tap_lane(H, W, x, y, z, lane, off, wt);
  if (lane < 12) s_tap[lane] = make_int2(off, __float_as_int(wt));
  __syncwarp();
  const float* pl = planes + lane;
  float v[12], wts[12];
#pragma unroll
  for (int t = 0; t < 6; ++t) {
    const int4 q = *reinterpret_cast<const int4*>(s_tap + 2 * t);
    v[2 * t] = __ldg(pl + q.x); wts[2 * t] = __int_as_float(q.y);
    v[2 * t + 1] = __ldg(pl + q.z); wts[2 * t + 1] = __int_as_float(q.w);
  }
  float fp[3];
#pragma unroll
  for (int p = 0; p < 3; ++p) {
    float acc = v[4 * p] * wts[4 * p];
    acc = fmaf(v[4 * p + 1], wts[4 * p + 1], acc);
    acc = fmaf(v[4 * p + 2], wts[4 * p + 2], acc);
    fp[p] = fmaf(v[4 * p + 3], wts[4 * p + 3], acc);
  }
  const float f = __fdiv_rn(__fadd_rn(__fadd_rn(fp[0], fp[1]), fp[2]), 3.f);      // sampled_features.mean(1)
  s_feat[lane] = f;
  __syncwarp();
  float ha = d.b1a, hb = d.b1b;
#pragma unroll
  for (int c4 = 0; c4 < TP_C / 4; ++c4) {                   // all 32 features: 8 broadcast LDS.128 instead of 32 shuffles
    const float4 q = *reinterpret_cast<const float4*>(s_feat + 4 * c4);
    ha = fmaf(d.w1a[4 * c4], q.x, ha); hb = fmaf(d.w1b[4 * c4], q.x, hb);
    ha = fmaf(d.w1a[4 * c4 + 1], q.y, ha); hb = fmaf(d.w1b[4 * c4 + 1], q.y, hb);
    ha = fmaf(d.w1a[4 * c4 + 2], q.z, ha); hb = fmaf(d.w1b[4 * c4 + 2], q.z, hb);
    ha = fmaf(d.w1a[4 * c4 + 3], q.w, ha); hb = fmaf(d.w1b[4 * c4 + 3], q.w, hb);
  }
  ha = softplusf(ha); hb = softplusf(hb);
  float p0 = fmaf(d.w2a[0], ha, d.w2b[0] * hb), p1 = fmaf(d.w2a[1], ha, d.w2b[1] * hb);
  float p2 = fmaf(d.w2a[2], ha, d.w2b[2] * hb), p3 = fmaf(d.w2a[3], ha, d.w2b[3] * hb);
  // transposing warp reduction of 4 values: lanes [0,8) end with output 0, [8,16) 1, [16,24) 2, [24,32) 3
  const bool hi16 = lane & 16, hi8 = lane & 8;
  float a = hi16 ? p2 : p0, sa = hi16 ? p0 : p2;
  float b = hi16 ? p3 : p1, sb = hi16 ? p1 : p3;
  a += __shfl_xor_sync(0xffffffffu, sa, 16);
  b += __shfl_xor_sync(0xffffffffu, sb, 16);
  float c = hi8 ? b : a, sc = hi8 ? a : b;
  c += __shfl_xor_sync(0xffffffffu, sc, 8);
  c += __shfl_xor_sync(0xffffffffu, c, 4);
  c += __shfl_xor_sync(0xffffffffu, c, 2);
  c += __shfl_xor_sync(0xffffffffu, c, 1);
  c += d.b2;
  if (lane >= 8) c = __fsub_rn(__fmul_rn(__fdividef(1.f, 1.f + __expf(-c)), 1.002f), 0.001f);     // triplane.py:165
  return c;
}

// ---- MipRayMarcher2.run_forward for one ray held in shared memory (ray_marcher.py:25-57) ---------------------------
// s_d / s_sig / s_rgb[3] hold S sorted samples.  Writes the S-1 weights to s_w (if non-null) and returns, on every
// lane, (rgb, sum w z_mid, sum w).  dmin / dmax receive the ray's depth range for the chunk-global clamp.
struct March { float r, g, b, wz, wsum, dmin, dmax; };
__device__ __forceinline__ March march_ray(const float* s_d, const float* s_sig, const float* s_r, const float* s_g,
                                           const float* s_b, int S, float* s_w, int lane) {
  March m{0.f, 0.f, 0.f, 0.f, 0.f, INFINITY, -INFINITY};
  float carry = 1.f;                                       // prod_{j < base} (1 - alpha_j + 1e-10)
  for (int base = 0; base < S - 1; base += 32) {
    const int i = base + lane;
    const bool on = i < S - 1;
    float alpha = 0.f, zmid = 0.f, cr = 0.f, cg = 0.f, cb = 0.f;
    if (on) {
      const float d0 = s_d[i], d1 = s_d[i + 1];
      const float delta = __fsub_rn(d1, d0);
      zmid = __fmul_rn(__fadd_rn(d0, d1), 0.5f);
      const float dm = softplusf(__fsub_rn(__fmul_rn(__fadd_rn(s_sig[i], s_sig[i + 1]), 0.5f), 1.f));
      alpha = __fsub_rn(1.f, expf(-__fmul_rn(dm, delta)));
      cr = __fmul_rn(__fadd_rn(s_r[i], s_r[i + 1]), 0.5f);
      cg = __fmul_rn(__fadd_rn(s_g[i], s_g[i + 1]), 0.5f);
      cb = __fmul_rn(__fadd_rn(s_b[i], s_b[i + 1]), 0.5f);
      m.dmin = fminf(m.dmin, fminf(d0, d1)); m.dmax = fmaxf(m.dmax, fmaxf(d0, d1));
    }
    float t = on ? __fadd_rn(__fsub_rn(1.f, alpha), 1e-10f) : 1.f;       // inclusive product scan
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const float q = __shfl_up_sync(0xffffffffu, t, o);
      if (lane >= o) t *= q;
    }
    float excl = __shfl_up_sync(0xffffffffu, t, 1);
    if (lane == 0) excl = 1.f;
    const float wgt = on ? __fmul_rn(alpha, carry * excl) : 0.f;
    carry *= __shfl_sync(0xffffffffu, t, 31);
    if (on && s_w) s_w[i] = wgt;
    m.r = fmaf(wgt, cr, m.r); m.g = fmaf(wgt, cg, m.g); m.b = fmaf(wgt, cb, m.b);
    m.wz = fmaf(wgt, zmid, m.wz); m.wsum += wgt;
  }
  m.r = warp_sum(m.r); m.g = warp_sum(m.g); m.b = warp_sum(m.b); m.wz = warp_sum(m.wz); m.wsum = warp_sum(m.wsum);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    m.dmin = fminf(m.dmin, __shfl_xor_sync(0xffffffffu, m.dmin, o));
    m.dmax = fmaxf(m.dmax, __shfl_xor_sync(0xffffffffu, m.dmax, o));
  }
  return m;
}

// order-preserving float <-> uint map so that atomicMin / atomicMax on unsigned give the float min / max
__device__ __forceinline__ uint32_t f2ord(float f) { const uint32_t b = __float_as_uint(f); return (b & 0x80000000u) ? ~b : (b | 0x80000000u); }
__device__ __forceinline__ float ord2f(uint32_t u) { return __uint_as_float((u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u); }

__global__ void minmax_init_kernel(uint32_t* mm) { if (threadIdx.x < 4) mm[threadIdx.x] = (threadIdx.x & 1) ? 0u : 0xFFFFFFFFu; }

// composite_depth = clamp(nan_to_num(sum w z / sum w, inf), min(depths), max(depths))   (ray_marcher.py:46-50)
__global__ void __launch_bounds__(256) depth_clamp_kernel(float* __restrict__ d0, float* __restrict__ d1, int64_t R, const uint32_t* __restrict__ mm) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= R) return;
  if (d0) { float v = d0[i]; if (v != v) v = INFINITY; d0[i] = fminf(fmaxf(v, ord2f(mm[0])), ord2f(mm[1])); }
  if (d1) { float v = d1[i]; if (v != v) v = INFINITY; d1[i] = fminf(fmaxf(v, ord2f(mm[2])), ord2f(mm[3])); }
}

__device__ __forceinline__ void emit_march(const March& m, int white_back, int64_t r, float* rgb, float* depth, float* opac,
                                           uint32_t* mm, int lane) {
  if (lane == 0) {
    // white_back: composite_rgb + 1 - weight_total, left to right
    rgb[r * 3 + 0] = white_back ? __fadd_rn(__fadd_rn(m.r, 1.f), -m.wsum) : m.r;
    rgb[r * 3 + 1] = white_back ? __fadd_rn(__fadd_rn(m.g, 1.f), -m.wsum) : m.g;
    rgb[r * 3 + 2] = white_back ? __fadd_rn(__fadd_rn(m.b, 1.f), -m.wsum) : m.b;
    depth[r] = __fdiv_rn(m.wz, m.wsum);
    if (opac) opac[r] = m.wsum;
    atomicMin(mm, f2ord(m.dmin)); atomicMax(mm + 1, f2ord(m.dmax));
  }
}

// counter-based uniform in [0,1) for callers that pass no random tensors (not torch's Philox stream)
__device__ __forceinline__ float hash_uniform(uint64_t seed, uint64_t idx) {
  uint64_t x = seed * 0x9E3779B97F4A7C15ull + idx + 0x632BE59BD9B4E019ull;
  x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull; x ^= x >> 27; x *= 0x94D049BB133111EBull; x ^= x >> 31;
  return (float)(x >> 40) * (1.0f / 16777216.0f);
}

// Rank of sample k of x[0, St) under the (value, index) order torch.sort's values imply for cat([coarse, fine]) (renderer.py:160-170).
// The S stratified depths come out ascending; when a warp-wide check confirms that (`coarse_sorted`, no assumption: any other input
// takes the full count), a coarse sample only has to count the importance samples below it, and an importance sample does a
// binary search among the coarse ones plus the count among its own kind: ~48 + 6 comparisons instead of 96 at 48 + 48.
__device__ __forceinline__ bool coarse_is_sorted(const float* x, int S, int lane) {
  bool bad = false;
  for (int i = lane; i < S - 1; i += 32) bad |= !(x[i] <= x[i + 1]);
  return !__any_sync(0xffffffffu, bad);
}
__device__ __forceinline__ int rank_of_sample(const float* x, int S, int St, int k, bool coarse_sorted) {
  const float xk = x[k];
  int rank = 0;
  if (!coarse_sorted) {
    for (int m = 0; m < St; ++m) { const float xm = x[m]; rank += (xm < xk) || (xm == xk && m < k); }
    return rank;
  }
  const bool fine = k >= S;
  if (fine) {                                                 // #(coarse <= xk): a coarse sample wins a tie (smaller index)
    int lo = 0, hi = S;
    while (lo < hi) { const int mid = (lo + hi) >> 1; if (x[mid] <= xk) lo = mid + 1; else hi = mid; }
    rank = lo;
  } else {
    rank = k;                                                 // ascending, ties by index
  }
  for (int m = S; m < St; ++m) { const float xm = x[m]; rank += (xm < xk) || (fine && xm == xk && m < k); }
  return rank;
}

struct RenderArgs {
  const float* planes; int H, W;
  const float* dec; float lr_mul;
  const float* o; const float* d; int64_t R;
  int S, Si;
  float ray_start, ray_end, inv_start, inv_end, delta; const float* start_arr; const float* end_arr;
  int disparity; float coord_scale; int white_back;
  const float* rand_coarse; const float* u; uint64_t seed;
  float *rgb_c, *depth_c, *opac_c, *rgb_f, *depth_f, *opac_f;
  float* depths_fine_out;        // optional (R, Si): the importance samples, for stage parity tests
  float* samples_out;            // optional (R, 5, S + Si): depth | sigma | r | g | b of every sample, coarse then importance, unsorted
  uint32_t* mm;
};

// torch.linspace(a, b, n)[i] in fp32: step = (b - a) / (n - 1); first half a + i*step, second half b - (n-1-i)*step
__device__ __forceinline__ float linspace_ab(float a, float b, int i, int n) {
  const float step = __fdiv_rn(__fsub_rn(b, a), (float)(n - 1));
  return (i < n / 2) ? __fmaf_rn((float)i, step, a) : __fmaf_rn(-(float)(n - 1 - i), step, b);
}

// MMA: the decoder of 16 samples at a time on the tensor cores (decode16) instead of one sample at a time on CUDA cores with the
// first-layer weights in 64 registers per lane (field_eval); the NERFB_TP_MMA=0 build keeps the latter.
template <bool MMA>
__global__ void __launch_bounds__(TP_WARPS * 32, MMA ? 3 : 2) triplane_render_kernel(const RenderArgs a) {
  extern __shared__ __align__(16) float smem[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int S = a.S, Si = a.Si, St = S + Si;
  // per warp: two sample sets of 5 arrays (depth, sigma, r, g, b) x St, weights St, bins / cdf 2 x S
  const int scr_off = (11 * St + 2 * S + 3) & ~3;            // 16-byte aligned scratch behind the per-ray arrays
  constexpr int SCR = MMA ? TM_PTS * TM_FS : 128;             // MMA: the feature tile of 16 samples
  float* const W1s = smem;                                    // MMA: the block's decoder weights in front of the warp areas
  float* base = smem + (MMA ? TM_DEC_FLOATS : 0) + (size_t)wib * (scr_off + SCR);
  float* const A = base;               // sample set A: [depth | sigma | r | g | b] x St, coarse [0,S) then fine [S,St)
  float* const B = base + 5 * St;      // sample set B: the merged, sorted samples
  float* s_w = base + 10 * St;
  float* s_bins = s_w + St;
  float* s_cdf = s_bins + S;
  float* s_scr = base + scr_off;               // 2 x [12 x (offset, weight) | 32 features], or the MMA feature tile
  Dec dec;
  float b2[4] = {0.f, 0.f, 0.f, 0.f};
  if (MMA) {
    load_decoder_mma(W1s, a.dec, a.lr_mul);
#pragma unroll
    for (int k = 0; k < 4; ++k) b2[k] = __ldg(a.dec + TP_HID * TP_C + TP_HID + 4 * TP_HID + k) * a.lr_mul;
    __syncthreads();
  } else {
    load_decoder(dec, a.dec, a.lr_mul, lane);
  }
  for (int64_t r = (int64_t)blockIdx.x * TP_WARPS + wib; r < a.R; r += (int64_t)gridDim.x * TP_WARPS) {
    const float ox = __ldg(a.o + r * 3), oy = __ldg(a.o + r * 3 + 1), oz = __ldg(a.o + r * 3 + 2);
    const float dx = __ldg(a.d + r * 3), dy = __ldg(a.d + r * 3 + 1), dz = __ldg(a.d + r * 3 + 2);
    const float rs = a.start_arr ? __ldg(a.start_arr + r) : a.ray_start, re = a.end_arr ? __ldg(a.end_arr + r) : a.ray_end;
    // ---- stratified depths (renderer.py:172-195) ----
    for (int i = lane; i < S; i += 32) {
      const float rnd = a.rand_coarse ? __ldg(a.rand_coarse + r * S + i) : hash_uniform(a.seed, (uint64_t)(r * S + i));
      float dep;
      if (a.disparity) {
        const float t = __fadd_rn(linspace01(i, S), __fmul_rn(rnd, (float)(1.0 / (double)(S - 1))));
        const float is = a.start_arr ? __fdiv_rn(1.f, rs) : a.inv_start, ie = a.end_arr ? __fdiv_rn(1.f, re) : a.inv_end;
        dep = __fdiv_rn(1.f, __fadd_rn(__fmul_rn(is, __fsub_rn(1.f, t)), __fmul_rn(ie, t)));
      } else if (a.start_arr) {       // math_utils.linspace: start + (i / (S-1)) * (stop - start), delta = (stop - start) / (S-1)
        const float st = __fdiv_rn((float)i, (float)(S - 1));
        dep = __fadd_rn(__fadd_rn(rs, __fmul_rn(st, __fsub_rn(re, rs))), __fmul_rn(rnd, __fdiv_rn(__fsub_rn(re, rs), (float)(S - 1))));
      } else {
        dep = __fadd_rn(linspace_ab(rs, re, i, S), __fmul_rn(rnd, a.delta));
      }
      A[i] = dep;
    }
    __syncwarp();
#pragma unroll 1
    for (int pass = 0; pass < 2; ++pass) {
      // ---- field pass over the coarse (pass 0) / importance (pass 1) samples: ONE copy of the decoder code ----
      const int s0 = pass ? S : 0, s1 = pass ? St : S;
      if (MMA) {
#pragma unroll 1
        for (int sb = s0; sb < s1; sb += TM_PTS) {
#pragma unroll 1
          for (int rr = 0; rr < TM_PTS / 4; ++rr) {           // four samples per round, 8 lanes x float4 per tap row
            const int pt = 4 * rr + (lane >> 3);
            const float t = A[min(sb + pt, s1 - 1)];          // a ragged last batch re-reads the last sample
            const float4 f = gather_quad(a.planes, a.H, a.W, __fmul_rn(a.coord_scale, __fadd_rn(ox, __fmul_rn(t, dx))),
                                         __fmul_rn(a.coord_scale, __fadd_rn(oy, __fmul_rn(t, dy))),
                                         __fmul_rn(a.coord_scale, __fadd_rn(oz, __fmul_rn(t, dz))), (lane & 7) * 4);
            *reinterpret_cast<float4*>(s_scr + pt * TM_FS + (lane & 7) * 4) = f;
          }
          __syncwarp();
          float out[2][4];
          decode16(s_scr, W1s, b2, lane, out);
          if ((lane & 3) == 0) {
#pragma unroll
            for (int u = 0; u < 2; ++u) {
              const int s = sb + (lane >> 2) + 8 * u;
              if (s < s1) { A[St + s] = out[u][0]; A[2 * St + s] = out[u][1]; A[3 * St + s] = out[u][2]; A[4 * St + s] = out[u][3]; }
            }
          }
          __syncwarp();
        }
      } else
#pragma unroll 2
      for (int s = s0; s < s1; ++s) {
        const float t = A[s];
        const float v = field_eval(a.planes, a.H, a.W, __fmul_rn(a.coord_scale, __fadd_rn(ox, __fmul_rn(t, dx))),
                                   __fmul_rn(a.coord_scale, __fadd_rn(oy, __fmul_rn(t, dy))),
                                   __fmul_rn(a.coord_scale, __fadd_rn(oz, __fmul_rn(t, dz))), dec, lane,
                                   reinterpret_cast<int2*>(s_scr + (s & 1) * 64), s_scr + (s & 1) * 64 + 32);
        if ((lane & 7) == 0) A[(1 + (lane >> 3)) * St + s] = v;
      }
      __syncwarp();
      if (pass) break;
      const March mc = march_ray(A, A + St, A + 2 * St, A + 3 * St, A + 4 * St, S, s_w, lane);
      emit_march(mc, a.white_back, r, a.rgb_c, a.depth_c, a.opac_c, a.mm, lane);
      __syncwarp();
      // ---- importance sampling (renderer.py:197-215): smooth, + 0.01, inverse CDF over the mid-points ----
      const int nwt = S - 1;                                  // weights from the marcher
      for (int i = lane; i < nwt; i += 32) {                  // max_pool1d(2,1,pad 1) then avg_pool1d(2,1): S-1 values
        const float wm1 = i > 0 ? s_w[i - 1] : -INFINITY, w0 = s_w[i], wp1 = i + 1 < nwt ? s_w[i + 1] : -INFINITY;
        const float m0 = fmaxf(wm1, w0), m1 = fmaxf(w0, wp1);
        B[i] = __fadd_rn(__fmul_rn(__fadd_rn(m0, m1), 0.5f), 0.01f);
      }
      for (int i = lane; i < S - 1; i += 32) s_bins[i] = __fmul_rn(0.5f, __fadd_rn(A[i], A[i + 1]));
      __syncwarp();
      const int nb = S - 2;                                   // cdf length: weights[:, 1:-1] has S-3 entries
      warp_build_cdf(B + 1, s_cdf, nb, 1e-5f, lane);
      for (int j = lane; j < Si; j += 32) {
        const float uj = a.u ? __ldg(a.u + r * Si + j) : hash_uniform(a.seed ^ 0x5851F42D4C957F2Dull, (uint64_t)(r * Si + j));
        const float zf = invert_cdf(s_bins, s_cdf, nb, uj, 1e-5f, nullptr);
        A[S + j] = zf;
        if (a.depths_fine_out) a.depths_fine_out[r * Si + j] = zf;
      }
      __syncwarp();
    }
    if (a.samples_out) {                                     // what the training backward needs (nothing else is saved)
      float* so = a.samples_out + r * 5 * St;
      for (int i = lane; i < 5 * St; i += 32) so[i] = A[i];
    }
    // ---- unify_samples (renderer.py:160-170): rank-by-counting sort of the depths, payload moved along ----
    const bool c_sorted = coarse_is_sorted(A, S, lane);
    for (int k = lane; k < St; k += 32) {
      const int rank = rank_of_sample(A, S, St, k, c_sorted);
#pragma unroll
      for (int q = 0; q < 5; ++q) B[q * St + rank] = A[q * St + k];
    }
    __syncwarp();
    const March mf = march_ray(B, B + St, B + 2 * St, B + 3 * St, B + 4 * St, St, nullptr, lane);
    emit_march(mf, a.white_back, r, a.rgb_f, a.depth_f, a.opac_f, a.mm + 2, lane);
    __syncwarp();
  }
}

// ---- stand-alone pieces (the reference's public sub-APIs) -----------------------------------------------------------
// run_model on explicit coordinates: rgb (P,3), sigma (P)
__global__ void __launch_bounds__(TP_WARPS * 32) triplane_field_kernel(const float* __restrict__ planes, int H, int W,
    const float* __restrict__ decp, float lr_mul, const float* __restrict__ coords, int64_t P, float coord_scale,
    float* __restrict__ rgb, float* __restrict__ sigma) {
  const int lane = threadIdx.x & 31;
  __shared__ __align__(16) float s_scr[TP_WARPS][64];
  float* scr = s_scr[threadIdx.x >> 5];
  Dec dec;
  load_decoder(dec, decp, lr_mul, lane);
  for (int64_t p = (int64_t)blockIdx.x * TP_WARPS + (threadIdx.x >> 5); p < P; p += (int64_t)gridDim.x * TP_WARPS) {
    const float x = __ldg(coords + p * 3), y = __ldg(coords + p * 3 + 1), z = __ldg(coords + p * 3 + 2);
    const float v = field_eval(planes, H, W, __fmul_rn(coord_scale, x), __fmul_rn(coord_scale, y), __fmul_rn(coord_scale, z), dec, lane,
                               reinterpret_cast<int2*>(scr), scr + 32);
    __syncwarp();
    if (lane == 0) sigma[p] = v;
    else if ((lane & 7) == 0) rgb[p * 3 + (lane >> 3) - 1] = v;
  }
}

// the same, 16 points per warp pass with the decoder on the tensor cores
__global__ void __launch_bounds__(TP_WARPS * 32, 3) triplane_field_mma_kernel(const float* __restrict__ planes, int H, int W,
    const float* __restrict__ decp, float lr_mul, const float* __restrict__ coords, int64_t P, float coord_scale,
    float* __restrict__ rgb, float* __restrict__ sigma) {
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  __shared__ __align__(16) float s_dec[TM_DEC_FLOATS];
  __shared__ __align__(16) float s_f[TP_WARPS][TM_PTS * TM_FS];
  float* Fs = s_f[wib];
  load_decoder_mma(s_dec, decp, lr_mul);
  float b2[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) b2[k] = __ldg(decp + TP_HID * TP_C + TP_HID + 4 * TP_HID + k) * lr_mul;
  __syncthreads();
  const int64_t n_batches = (P + TM_PTS - 1) / TM_PTS;
  for (int64_t batch = (int64_t)blockIdx.x * TP_WARPS + wib; batch < n_batches; batch += (int64_t)gridDim.x * TP_WARPS) {
    const int64_t p0 = batch * TM_PTS;
#pragma unroll 1
    for (int rr = 0; rr < TM_PTS / 4; ++rr) {
      const int pt = 4 * rr + (lane >> 3);
      const int64_t q = p0 + pt < P ? p0 + pt : P - 1;
      const float4 f = gather_quad(planes, H, W, __fmul_rn(coord_scale, __ldg(coords + q * 3)), __fmul_rn(coord_scale, __ldg(coords + q * 3 + 1)),
                                   __fmul_rn(coord_scale, __ldg(coords + q * 3 + 2)), (lane & 7) * 4);
      *reinterpret_cast<float4*>(Fs + pt * TM_FS + (lane & 7) * 4) = f;
    }
    __syncwarp();
    float out[2][4];
    decode16(Fs, s_dec, b2, lane, out);
    if ((lane & 3) == 0) {
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int64_t q = p0 + (lane >> 2) + 8 * u;
        if (q < P) { sigma[q] = out[u][0]; rgb[q * 3] = out[u][1]; rgb[q * 3 + 1] = out[u][2]; rgb[q * 3 + 2] = out[u][3]; }
      }
    }
    __syncwarp();
  }
}

// MipRayMarcher2 on explicit (R,S,3) colours, (R,S) densities, (R,S) depths
__global__ void __launch_bounds__(TP_WARPS * 32) mip_march_kernel(const float* __restrict__ colors, const float* __restrict__ dens,
    const float* __restrict__ depths, int64_t R, int S, int white_back, float* __restrict__ rgb, float* __restrict__ depth,
    float* __restrict__ weights, uint32_t* mm) {
  extern __shared__ float smem[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  float* s = smem + (size_t)wib * 6 * S;
  for (int64_t r = (int64_t)blockIdx.x * TP_WARPS + wib; r < R; r += (int64_t)gridDim.x * TP_WARPS) {
    for (int i = lane; i < S; i += 32) {
      s[i] = __ldg(depths + r * S + i); s[S + i] = __ldg(dens + r * S + i);
      s[2 * S + i] = __ldg(colors + (r * S + i) * 3); s[3 * S + i] = __ldg(colors + (r * S + i) * 3 + 1);
      s[4 * S + i] = __ldg(colors + (r * S + i) * 3 + 2);
    }
    __syncwarp();
    const March m = march_ray(s, s + S, s + 2 * S, s + 3 * S, s + 4 * S, S, s + 5 * S, lane);
    emit_march(m, white_back, r, rgb, depth, nullptr, mm, lane);
    __syncwarp();
    if (weights) for (int i = lane; i < S - 1; i += 32) weights[r * (S - 1) + i] = s[5 * S + i];
    __syncwarp();
  }
}


// =====================================================================================================================
// Backward of the stand-alone pieces (EG3D training, system.py:22,75,111-120 of the reference back-propagates through
// grid_sample + OSGDecoder + MipRayMarcher2).  The fused renderer above stays the no_grad render path; training composes
//   field (fwd / bwd)  ->  sort by depth (torch)  ->  marcher (fwd / bwd)
// on the sample depths the fused kernel chose (importance depths are detached in the reference too, renderer.py:215).
// =====================================================================================================================

constexpr int TP_DEC_PARAMS = TP_HID * TP_C + TP_HID + 4 * TP_HID + 4;      // W1 | b1 | W2 | b2

// d(run_model) for explicit coordinates: given dL/drgb (P,3), dL/dsigma (P) accumulate dL/dplanes (channels-last, fp32
// reductions: each bilinear tap is one 128-byte row, issued as 8 lanes x red.v4.f32, four taps per warp instruction) and
// dL/d(decoder parameters).  One warp per point, the forward is recomputed (nothing was saved).
// Register plan: the kernel is latency-bound, so registers buy warps.  The 64 x 32 first-layer matrix lives in shared memory
// TWICE -- [c][j] for the forward (lane = hidden units j, j + 32) and [j][c] for dL/dfeat (lane = channel c).  Its gradient
// dW1 = sum_p dpre_p (x) feat_p is NOT accumulated per warp (64 registers per lane): every warp parks dpre (64) and feat (32)
// of TPB_NB points in a block tile, and after a block barrier each of the 256 threads adds the tile's 8 * TPB_NB outer
// products into the EIGHT entries of dW1 it owns (row j = tid / 4, columns 8 (tid % 4) ..+7).
__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" :: "l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
// predicated on wt != 0 without a branch (twelve of these per point)
__device__ __forceinline__ void red_add_v4_if(float* p, float a, float b, float c, float d, float wt) {
  asm volatile("{ .reg .pred q; setp.neu.f32 q, %5, 0f00000000; @q red.global.add.v4.f32 [%0], {%1, %2, %3, %4}; }"
               :: "l"(p), "f"(a), "f"(b), "f"(c), "f"(d), "f"(wt) : "memory");
}

constexpr int TPB_NB = 4;                        // points per warp between two block barriers
constexpr int TPB_TILE = TP_WARPS * TPB_NB;      // points per tile

__global__ void __launch_bounds__(TP_WARPS * 32, 3) triplane_field_bwd_kernel(const float* __restrict__ planes, int H, int W,
    const float* __restrict__ decp, float lr_mul, const float* __restrict__ coords, int64_t P, float coord_scale,
    const float* __restrict__ d_rgb, const float* __restrict__ d_sigma, float* __restrict__ d_planes, float* __restrict__ d_dec) {
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  __shared__ __align__(16) float2 s_w1[TP_C * 32];           // [c][lane] = (W1[lane][c], W1[lane + 32][c]), times the layer gain
  __shared__ __align__(16) float2 s_w1t[(TP_HID / 2) * TP_C];  // [j / 2][c] = (W1[j][c], W1[j + 1][c]), j even
  __shared__ __align__(16) float s_scr[TP_WARPS][48];        // taps of two points (2 x 12 x int2)
  __shared__ __align__(16) float s_tile[TPB_TILE * (TP_HID + TP_C)];      // per point: dpre (64) | feat (32); reused as the flush buffer
  static_assert(TPB_NB % 2 == 0 && TPB_TILE * (TP_HID + TP_C) >= TP_DEC_PARAMS, "flush buffer aliases the tile");
  const float g1 = lr_mul / sqrtf((float)TP_C), g2 = lr_mul / sqrtf((float)TP_HID);
  for (int i = threadIdx.x; i < TP_HID * TP_C; i += blockDim.x) {
    const float w = __fmul_rn(__ldg(decp + i), g1);           // i = j * 32 + c
    const int j = i >> 5, c = i & (TP_C - 1);
    reinterpret_cast<float*>(s_w1t)[((j >> 1) * TP_C + c) * 2 + (j & 1)] = w;
    reinterpret_cast<float*>(s_w1)[(c * 32 + (j & 31)) * 2 + (j >> 5)] = w;
  }
  int2* s_tap = reinterpret_cast<int2*>(s_scr[wib]);
  const float* b1p = decp + TP_HID * TP_C;
  const float* w2p = b1p + TP_HID;
  const float b1a = __ldg(b1p + lane) * lr_mul, b1b = __ldg(b1p + lane + 32) * lr_mul;
  float w2a[4], w2b[4], b2[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    w2a[k] = __fmul_rn(__ldg(w2p + k * TP_HID + lane), g2);
    w2b[k] = __fmul_rn(__ldg(w2p + k * TP_HID + lane + 32), g2);
    b2[k] = __ldg(w2p + 4 * TP_HID + k) * lr_mul;
  }
  float2 gw1[4];                                              // dL/dW1_eff[own_j][own_c + 0..7]
#pragma unroll
  for (int i = 0; i < 4; ++i) gw1[i] = make_float2(0.f, 0.f);
  const int own_j = threadIdx.x >> 2, own_c = (threadIdx.x & 3) * 8;
  float gb1a = 0.f, gb1b = 0.f, gw2a[4], gw2b[4], gb2[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) { gw2a[k] = 0.f; gw2b[k] = 0.f; gb2[k] = 0.f; }
  __syncthreads();
  const int grp = lane >> 3, sub4 = (lane & 7) * 4;
  // tiles of TPB_TILE consecutive points; warp w takes points w * TPB_NB .. + TPB_NB - 1 of the tile
  const int64_t n_tiles = (P + TPB_TILE - 1) / TPB_TILE;
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int64_t p0 = tile * TPB_TILE + wib * TPB_NB;
    // two points per pass: every first-layer weight fetched from shared memory serves both, and the two dependency chains
    // (gather -> hidden layer -> warp sums -> dL/dfeat) overlap
#pragma unroll 1
    for (int ip = 0; ip < TPB_NB; ip += 2) {
      float* t_dp0 = s_tile + (wib * TPB_NB + ip) * (TP_HID + TP_C);
      float* t_dp1 = t_dp0 + (TP_HID + TP_C);
      float* t_f0 = t_dp0 + TP_HID;
      float* t_f1 = t_dp1 + TP_HID;
      if (p0 + ip >= P) {                                     // tail: the tile pass must see zeros
        t_dp0[lane] = 0.f; t_dp0[lane + 32] = 0.f; t_f0[lane] = 0.f;
        t_dp1[lane] = 0.f; t_dp1[lane + 32] = 0.f; t_f1[lane] = 0.f;
        continue;
      }
      const bool two = p0 + ip + 1 < P;                       // an odd tail re-reads the last point with zero cotangents
      const int64_t pp[2] = {p0 + ip, two ? p0 + ip + 1 : p0 + ip};
      float go[2][4];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const float x = __fmul_rn(coord_scale, __ldg(coords + pp[u] * 3)), y = __fmul_rn(coord_scale, __ldg(coords + pp[u] * 3 + 1)),
                    z = __fmul_rn(coord_scale, __ldg(coords + pp[u] * 3 + 2));
        const bool on = u == 0 || two;
        go[u][0] = on ? __ldg(d_sigma + pp[u]) : 0.f;
#pragma unroll
        for (int k = 0; k < 3; ++k) go[u][k + 1] = on ? __ldg(d_rgb + pp[u] * 3 + k) : 0.f;
        int off; float wt;
        tap_lane(H, W, x, y, z, lane, off, wt);
        if (lane < 12) s_tap[12 * u + lane] = make_int2(off, __float_as_int(wt));
      }
      __syncwarp();
      // ---- forward, keeping the intermediates ----
      float f[2];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        float fsum = 0.f;
#pragma unroll
        for (int pl = 0; pl < 3; ++pl) {
          const int4 q0 = *reinterpret_cast<const int4*>(s_tap + 12 * u + 4 * pl), q1 = *reinterpret_cast<const int4*>(s_tap + 12 * u + 4 * pl + 2);
          float acc = __ldg(planes + q0.x + lane) * __int_as_float(q0.y);
          acc = fmaf(__ldg(planes + q0.z + lane), __int_as_float(q0.w), acc);
          acc = fmaf(__ldg(planes + q1.x + lane), __int_as_float(q1.y), acc);
          acc = fmaf(__ldg(planes + q1.z + lane), __int_as_float(q1.w), acc);
          fsum += acc;
        }
        f[u] = fsum / 3.f;                                    // sampled_features.mean(1)
      }
      t_f0[lane] = f[0]; t_f1[lane] = f[1];
      __syncwarp();
      float2 pre[2] = {make_float2(b1a, b1b), make_float2(b1a, b1b)};       // two-wide FMAs (FFMA2): both hidden units at once
#pragma unroll
      for (int c4 = 0; c4 < TP_C; c4 += 4) {
        const float4 fa = *reinterpret_cast<const float4*>(t_f0 + c4), fb = *reinterpret_cast<const float4*>(t_f1 + c4);
        const float2 wa = s_w1[(c4 + 0) * 32 + lane], wb = s_w1[(c4 + 1) * 32 + lane], wc = s_w1[(c4 + 2) * 32 + lane], wd = s_w1[(c4 + 3) * 32 + lane];
        pre[0] = __ffma2_rn(wa, make_float2(fa.x, fa.x), pre[0]); pre[1] = __ffma2_rn(wa, make_float2(fb.x, fb.x), pre[1]);
        pre[0] = __ffma2_rn(wb, make_float2(fa.y, fa.y), pre[0]); pre[1] = __ffma2_rn(wb, make_float2(fb.y, fb.y), pre[1]);
        pre[0] = __ffma2_rn(wc, make_float2(fa.z, fa.z), pre[0]); pre[1] = __ffma2_rn(wc, make_float2(fb.z, fb.z), pre[1]);
        pre[0] = __ffma2_rn(wd, make_float2(fa.w, fa.w), pre[0]); pre[1] = __ffma2_rn(wd, make_float2(fb.w, fb.w), pre[1]);
      }
      float ha[2], hb[2], o[2][4];
#pragma unroll
      for (int u = 0; u < 2; ++u) { ha[u] = softplusf(pre[u].x); hb[u] = softplusf(pre[u].y); }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float s0 = fmaf(w2a[k], ha[0], w2b[k] * hb[0]), s1 = fmaf(w2a[k], ha[1], w2b[k] * hb[1]);
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) { s0 += __shfl_xor_sync(0xffffffffu, s0, d); s1 += __shfl_xor_sync(0xffffffffu, s1, d); }
        o[0][k] = s0 + b2[k]; o[1][k] = s1 + b2[k];
      }
      // ---- upstream: sigma = o[0]; rgb_k = sigmoid(o[k]) * 1.002 - 0.001 (triplane.py:165) ----
#pragma unroll
      for (int u = 0; u < 2; ++u) {
#pragma unroll
        for (int k = 1; k < 4; ++k) {
          const float sg = 1.f / (1.f + __expf(-o[u][k]));
          go[u][k] *= 1.002f * sg * (1.f - sg);
        }
        float dha = 0.f, dhb = 0.f;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          gw2a[k] = fmaf(go[u][k], ha[u], gw2a[k]); gw2b[k] = fmaf(go[u][k], hb[u], gw2b[k]); gb2[k] += go[u][k];
          dha = fmaf(w2a[k], go[u][k], dha); dhb = fmaf(w2b[k], go[u][k], dhb);
        }
        // softplus'(x) = sigmoid(x) (1 beyond the threshold of 20)
        const float dpa = dha * (pre[u].x > 20.f ? 1.f : 1.f / (1.f + __expf(-pre[u].x)));
        const float dpb = dhb * (pre[u].y > 20.f ? 1.f : 1.f / (1.f + __expf(-pre[u].y)));
        gb1a += dpa; gb1b += dpb;
        float* t_dp = u ? t_dp1 : t_dp0;
        t_dp[lane] = dpa; t_dp[lane + 32] = dpb;
      }
      __syncwarp();
      // ---- lane = channel: dL/dfeat_c = sum_j W1[j][c] dpre_j ----
      float2 dfa[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)}, dfb[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
#pragma unroll
      for (int j4 = 0; j4 < TP_HID; j4 += 4) {
        const float4 da = *reinterpret_cast<const float4*>(t_dp0 + j4), db = *reinterpret_cast<const float4*>(t_dp1 + j4);
        const float2 w01 = s_w1t[(j4 >> 1) * TP_C + lane], w23 = s_w1t[((j4 >> 1) + 1) * TP_C + lane];
        dfa[0] = __ffma2_rn(w01, make_float2(da.x, da.y), dfa[0]); dfb[0] = __ffma2_rn(w23, make_float2(da.z, da.w), dfb[0]);
        dfa[1] = __ffma2_rn(w01, make_float2(db.x, db.y), dfa[1]); dfb[1] = __ffma2_rn(w23, make_float2(db.z, db.w), dfb[1]);
      }
      // 8 lanes x 4 channels cover one tap's 128-byte row; the 4 lane groups take the 4 corners of a plane
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const float df = ((dfa[u].x + dfa[u].y) + (dfb[u].x + dfb[u].y)) * (1.f / 3.f);
        const float v0 = __shfl_sync(0xffffffffu, df, sub4), v1 = __shfl_sync(0xffffffffu, df, sub4 + 1),
                    v2 = __shfl_sync(0xffffffffu, df, sub4 + 2), v3 = __shfl_sync(0xffffffffu, df, sub4 + 3);
        if (u == 0 || two) {
#pragma unroll
          for (int pl = 0; pl < 3; ++pl) {
            const int2 q = s_tap[12 * u + 4 * pl + grp];
            const float w = __int_as_float(q.y);
            if (w != 0.f) red_add_v4(d_planes + q.x + sub4, v0 * w, v1 * w, v2 * w, v3 * w);
          }
        }
      }
      __syncwarp();
    }
    __syncthreads();
    // ---- the tile's outer products into this thread's 8 entries of dW1 ----
#pragma unroll 8
    for (int t = 0; t < TPB_TILE; ++t) {
      const float* tp = s_tile + t * (TP_HID + TP_C);
      const float d = tp[own_j];
      const float4 f0 = *reinterpret_cast<const float4*>(tp + TP_HID + own_c), f1 = *reinterpret_cast<const float4*>(tp + TP_HID + own_c + 4);
      const float2 dd = make_float2(d, d);
      gw1[0] = __ffma2_rn(dd, make_float2(f0.x, f0.y), gw1[0]); gw1[1] = __ffma2_rn(dd, make_float2(f0.z, f0.w), gw1[1]);
      gw1[2] = __ffma2_rn(dd, make_float2(f1.x, f1.y), gw1[2]); gw1[3] = __ffma2_rn(dd, make_float2(f1.z, f1.w), gw1[3]);
    }
    __syncthreads();
  }
  // ---- flush the decoder gradients: effective -> raw parameters (gains of networks_stylegan2.py:109-110); warps combine in
  // shared memory first so a block issues ONE global atomic per parameter ----
  float* s_g = s_tile;
  for (int i = threadIdx.x; i < TP_DEC_PARAMS; i += blockDim.x) s_g[i] = 0.f;
  __syncthreads();
  {
    float* o = s_g + own_j * TP_C + own_c;                    // sole owner of these 8
    o[0] = gw1[0].x * g1; o[1] = gw1[0].y * g1; o[2] = gw1[1].x * g1; o[3] = gw1[1].y * g1;
    o[4] = gw1[2].x * g1; o[5] = gw1[2].y * g1; o[6] = gw1[3].x * g1; o[7] = gw1[3].y * g1;
  }
  float* gb1 = s_g + TP_HID * TP_C;
  atomicAdd(gb1 + lane, gb1a * lr_mul); atomicAdd(gb1 + lane + 32, gb1b * lr_mul);
  float* gw2 = gb1 + TP_HID;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    atomicAdd(gw2 + k * TP_HID + lane, gw2a[k] * g2);
    atomicAdd(gw2 + k * TP_HID + lane + 32, gw2b[k] * g2);
  }
  if (lane == 0) {
#pragma unroll
    for (int k = 0; k < 4; ++k) atomicAdd(gw2 + 4 * TP_HID + k, gb2[k] * lr_mul);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < TP_DEC_PARAMS; i += blockDim.x) atomicAdd(d_dec + i, s_g[i]);
}

// ---- the same backward with the decoder's four small GEMMs on the tensor cores -------------------------------------------
// The CUDA-core kernel above is bound by the shared-memory data pipe: a mat-vec on CUDA cores needs every feature / every
// dpre on every lane, and a warp-uniform shared load delivers ONE useful word per wavefront (232 wavefronts per point
// measured, ~70 % of the pipe).  mma.sync fragments are distributed over the lanes, so nothing is broadcast.  A warp takes
// 16 points at a time (M = 16):
//   GEMM1  pre^T [16 x 64]  = F^T [16 x 32] . W1^T            A from shared (features), B from shared (W1), C in registers
//   GEMM2  dF^T  [16 x 32]  = dpre^T [16 x 64] . W1            A = GEMM1's C registers (k slots permuted), B from shared
//   GEMM3  dW1^T [32 x 64] += F [32 x 16 pts] . dpre [16 x 64] accumulators in registers for the whole kernel
//   GEMM4  dW2^T [64 x  4] += h^T [64 x 16 pts] . go [16 x 4]  (N padded to 8)
// all as m16n8k8 TF32 with the 3-term split (a_lo b_hi + a_hi b_lo + a_hi b_hi, fp32 accumulate): fp32-grade accuracy.
// Gather / scatter: a lane is (point slot = lane / 8, channel quad = lane % 8): 8 lanes x 16 B cover a tap's 128-byte row, a
// warp instruction serves four points, every lane computes its own point's 12 taps (no tap broadcast either).
constexpr int TM_WARPS = 12;     // warps per block (one block per SM): the kernel is latency-bound, registers cap the count
constexpr int TM_FT = 20;      // features [ch][pt]: stride = 4 (mod 16)
constexpr int TM_DP = 72;      // dpre / h [pt][hid]: stride = 8 (mod 32): (pt over lane % 4, hid over lane / 4) is conflict-free
constexpr int TM_WARP_FLOATS = TM_PTS * TM_FS + TP_C * TM_FT + 2 * TM_PTS * TM_DP + TM_PTS * 4;
constexpr int TM_BLOCK_FLOATS = TM_DEC_FLOATS;
constexpr size_t TM_SMEM = (size_t)(TM_BLOCK_FLOATS + TM_WARPS * TM_WARP_FLOATS) * sizeof(float);
static_assert(TM_WARPS * TM_WARP_FLOATS >= TP_DEC_PARAMS, "flush buffer aliases the warp areas");

__global__ void __launch_bounds__(TM_WARPS * 32, 1) triplane_field_bwd_mma_kernel(const float* __restrict__ planes, int H, int W,
    const float* __restrict__ decp, float lr_mul, const float* __restrict__ coords, int64_t P, float coord_scale,
    const float* __restrict__ d_rgb, const float* __restrict__ d_sigma, float* __restrict__ d_planes, float* __restrict__ d_dec) {
  extern __shared__ __align__(16) float tm_smem[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;                     // mma fragment coordinates
  const int ps = lane >> 3, sub4 = (lane & 7) * 4;            // gather / scatter coordinates
  float* W1s = tm_smem;                                       // [hid][TM_W1] hi | lo, times the layer gain
  float* W2s = W1s + TM_W1_FLOATS;                            // [k][hid], times the layer gain
  float* B1s = W2s + 4 * TP_HID;
  float* wbase = B1s + TP_HID + wib * TM_WARP_FLOATS;
  float* Fs = wbase;                                          // [pt][TM_FS]   features, later dL/dfeat
  float* Ft = Fs + TM_PTS * TM_FS;                            // [ch][TM_FT]   features, transposed
  float* DPs = Ft + TP_C * TM_FT;                             // [pt][TM_DP]   dpre
  float* Hs = DPs + TM_PTS * TM_DP;                           // [pt][TM_DP]   hidden activations
  float* GOs = Hs + TM_PTS * TM_DP;                           // [pt][4]       dL/d(output pre-activations)
  const float g1 = lr_mul / sqrtf((float)TP_C), g2 = lr_mul / sqrtf((float)TP_HID);
  const float* w2p = decp + TP_HID * TP_C + TP_HID;
  load_decoder_mma(W1s, decp, lr_mul);
  float b2[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) b2[k] = __ldg(w2p + 4 * TP_HID + k) * lr_mul;
  float gw1[2][8][4];                                         // dW1^T: [m tile: ch 16 mt + g (+8)][n tile: hid 8 nt + 2 t (+1)]
  float gw2[4][4];                                            // dW2^T: [m tile: hid 16 mt + g (+8)][k = 2 t (+1), t < 2 of a padded 8]
  float gb1[8][2], gb2[4];
#pragma unroll
  for (int nt = 0; nt < 8; ++nt) {
#pragma unroll
    for (int i = 0; i < 4; ++i) { gw1[0][nt][i] = 0.f; gw1[1][nt][i] = 0.f; gw2[nt >> 1][i] = 0.f; }
    gb1[nt][0] = 0.f; gb1[nt][1] = 0.f;
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) gb2[k] = 0.f;
  __syncthreads();
  const int64_t n_batches = (P + TM_PTS - 1) / TM_PTS;
  for (int64_t batch = (int64_t)blockIdx.x * TM_WARPS + wib; batch < n_batches; batch += (int64_t)gridDim.x * TM_WARPS) {
    // The 16 points of a batch are n_batches apart, NOT consecutive: callers pass ray-major samples, and neighbours along a ray
    // fall into the same texels -- four of them in one red instruction serialise on the same L2 address (measured: 1.64 ms
    // on a render's samples against 1.34 ms on random points).  Consecutive batches (the block's other warps) still take
    // neighbouring points, so the gathers keep their L1 / L2 locality.
    auto point = [&](int i) { return (int64_t)i * n_batches + batch; };
    // ---- gather: four points per round, 8 lanes x float4 per tap row ----
#pragma unroll 1
    for (int r = 0; r < TM_PTS / 4; ++r) {
      const int pt = 4 * r + ps;
      const int64_t q = point(pt) < P ? point(pt) : P - 1;    // tail points re-read the last point; their cotangents are 0
      const float x = __fmul_rn(coord_scale, __ldg(coords + q * 3)), y = __fmul_rn(coord_scale, __ldg(coords + q * 3 + 1)),
                  z = __fmul_rn(coord_scale, __ldg(coords + q * 3 + 2));
      const float4 f = gather_quad(planes, H, W, x, y, z, sub4);
      *reinterpret_cast<float4*>(Fs + pt * TM_FS + sub4) = f;
      Ft[(sub4 + 0) * TM_FT + pt] = f.x; Ft[(sub4 + 1) * TM_FT + pt] = f.y; Ft[(sub4 + 2) * TM_FT + pt] = f.z; Ft[(sub4 + 3) * TM_FT + pt] = f.w;
    }
    __syncwarp();
    // ---- GEMM1: pre^T[pt][hid]; C fragment: rows pt = g, g + 8; columns hid = 8 nt + 2 t, + 1 ----
    float pre[8][4];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      const float2 b = *reinterpret_cast<const float2*>(B1s + 8 * nt + 2 * t);
      pre[nt][0] = b.x; pre[nt][1] = b.y; pre[nt][2] = b.x; pre[nt][3] = b.y;
    }
#pragma unroll
    for (int kt = 0; kt < 4; ++kt) {
      uint32_t ah[4], al[4];
      tf32_split_fast(Fs[g * TM_FS + 8 * kt + t], ah[0], al[0]);
      tf32_split_fast(Fs[(g + 8) * TM_FS + 8 * kt + t], ah[1], al[1]);
      tf32_split_fast(Fs[g * TM_FS + 8 * kt + t + 4], ah[2], al[2]);
      tf32_split_fast(Fs[(g + 8) * TM_FS + 8 * kt + t + 4], ah[3], al[3]);
#pragma unroll
      for (int nt = 0; nt < 8; ++nt)                          // B[k = ch][n = hid] = W1[hid][ch]
        mma_3x_w1(pre[nt], ah, al, W1s + (8 * nt + g) * TM_W1 + 8 * kt + t, 4);
    }
    // ---- hidden activations, second layer, upstream cotangents ----
    float o[2][4];                                            // [pt g, g + 8][k]
#pragma unroll
    for (int k = 0; k < 4; ++k) { o[0][k] = 0.f; o[1][k] = 0.f; }
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      float h[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) h[i] = softplusf(pre[nt][i]);
      *reinterpret_cast<float2*>(Hs + g * TM_DP + 8 * nt + 2 * t) = make_float2(h[0], h[1]);
      *reinterpret_cast<float2*>(Hs + (g + 8) * TM_DP + 8 * nt + 2 * t) = make_float2(h[2], h[3]);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float2 w = *reinterpret_cast<const float2*>(W2s + k * TP_HID + 8 * nt + 2 * t);
        float2 o2 = make_float2(o[0][k], o[1][k]);
        o2 = __ffma2_rn(make_float2(w.y, w.y), make_float2(h[1], h[3]), o2);
        o2 = __ffma2_rn(make_float2(w.x, w.x), make_float2(h[0], h[2]), o2);
        o[0][k] = o2.x; o[1][k] = o2.y;
      }
      // softplus'(x) = sigmoid(x) = 1 - exp(-softplus(x)) (1 beyond the threshold of 20); kept in place of pre
#pragma unroll
      for (int i = 0; i < 4; ++i) {                                      // (beyond the softplus threshold of 20 this is 1 to fp32)
        float en;
        asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(en) : "f"(h[i] * -1.4426950408889634f));
        pre[nt][i] = 1.f - en;
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        o[u][k] += __shfl_xor_sync(0xffffffffu, o[u][k], 1);
        o[u][k] += __shfl_xor_sync(0xffffffffu, o[u][k], 2);
        o[u][k] += b2[k];
      }
    }
    float go[2][4];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int64_t q = point(g + 8 * u);
      const bool on = q < P;
      go[u][0] = on ? __ldg(d_sigma + q) : 0.f;
#pragma unroll
      for (int k = 1; k < 4; ++k) {
        const float sg = 1.f / (1.f + __expf(-o[u][k]));      // rgb_k = sigmoid(o_k) * 1.002 - 0.001 (triplane.py:165)
        go[u][k] = on ? __ldg(d_rgb + q * 3 + k - 1) * 1.002f * sg * (1.f - sg) : 0.f;
      }
      if (t == 0) *reinterpret_cast<float4*>(GOs + (g + 8 * u) * 4) = make_float4(go[u][0], go[u][1], go[u][2], go[u][3]);
      if (t == 0) {
#pragma unroll
        for (int k = 0; k < 4; ++k) gb2[k] += go[u][k];
      }
    }
    // dpre[pt][hid] = sigmoid(pre) * sum_k W2[k][hid] go[pt][k], in the C-fragment layout of GEMM1 (in place of pre)
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      float2 dh01 = make_float2(0.f, 0.f), dh23 = make_float2(0.f, 0.f);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float2 w = *reinterpret_cast<const float2*>(W2s + k * TP_HID + 8 * nt + 2 * t);
        dh01 = __ffma2_rn(w, make_float2(go[0][k], go[0][k]), dh01);
        dh23 = __ffma2_rn(w, make_float2(go[1][k], go[1][k]), dh23);
      }
      const float dh[4] = {dh01.x, dh01.y, dh23.x, dh23.y};
#pragma unroll
      for (int i = 0; i < 4; ++i) pre[nt][i] *= dh[i];
      gb1[nt][0] += pre[nt][0] + pre[nt][2]; gb1[nt][1] += pre[nt][1] + pre[nt][3];
      *reinterpret_cast<float2*>(DPs + g * TM_DP + 8 * nt + 2 * t) = make_float2(pre[nt][0], pre[nt][1]);
      *reinterpret_cast<float2*>(DPs + (g + 8) * TM_DP + 8 * nt + 2 * t) = make_float2(pre[nt][2], pre[nt][3]);
    }
    // ---- GEMM2: dF^T[pt][ch] = dpre^T . W1; A = the C fragments above with k slot t <-> hid 2 t, slot t + 4 <-> hid 2 t + 1 ----
    float df[4][4];
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) { df[nt][0] = 0.f; df[nt][1] = 0.f; df[nt][2] = 0.f; df[nt][3] = 0.f; }
#pragma unroll
    for (int kt = 0; kt < 8; ++kt) {
      uint32_t ah[4], al[4];
      tf32_split_fast(pre[kt][0], ah[0], al[0]); tf32_split_fast(pre[kt][2], ah[1], al[1]);
      tf32_split_fast(pre[kt][1], ah[2], al[2]); tf32_split_fast(pre[kt][3], ah[3], al[3]);
#pragma unroll
      for (int nt = 0; nt < 4; ++nt)                          // B[k = hid][n = ch] = W1[hid][ch]
        mma_3x_w1(df[nt], ah, al, W1s + (8 * kt + 2 * t) * TM_W1 + 8 * nt + g, TM_W1);
    }
    __syncwarp();                                             // GEMM1's reads of Fs are done: Fs now carries dL/dfeat / 3
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
      *reinterpret_cast<float2*>(Fs + g * TM_FS + 8 * nt + 2 * t) = make_float2(df[nt][0] * (1.f / 3.f), df[nt][1] * (1.f / 3.f));
      *reinterpret_cast<float2*>(Fs + (g + 8) * TM_FS + 8 * nt + 2 * t) = make_float2(df[nt][2] * (1.f / 3.f), df[nt][3] * (1.f / 3.f));
    }
    __syncwarp();
    // ---- scatter: plane gradients as red.v4 rows, four points per round (taps recomputed: no registers held across the GEMMs) ----
#pragma unroll 1
    for (int r = 0; r < TM_PTS / 4; ++r) {
      const int pt = 4 * r + ps;
      if (point(pt) < P) {
        const int64_t q = point(pt);
        const float x = __fmul_rn(coord_scale, __ldg(coords + q * 3)), y = __fmul_rn(coord_scale, __ldg(coords + q * 3 + 1)),
                    z = __fmul_rn(coord_scale, __ldg(coords + q * 3 + 2));
        const float4 v = *reinterpret_cast<const float4*>(Fs + pt * TM_FS + sub4);
#pragma unroll
        for (int pl = 0; pl < 3; ++pl) {
          int off[4]; float wt[4];
          taps_plane(H, W, x, y, z, pl, off, wt);
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const float2 ww = make_float2(wt[c], wt[c]);
            const float2 p01 = __fmul2_rn(make_float2(v.x, v.y), ww), p23 = __fmul2_rn(make_float2(v.z, v.w), ww);
            red_add_v4_if(d_planes + off[c] + sub4, p01.x, p01.y, p23.x, p23.y, wt[c]);
          }
        }
      }
    }
    // ---- GEMM3: dW1^T[ch][hid] += F[ch][pt] . dpre[pt][hid] ----
#pragma unroll
    for (int kt = 0; kt < 2; ++kt) {
      uint32_t ah[2][4], al[2][4];
#pragma unroll
      for (int mt = 0; mt < 2; ++mt) {
        tf32_split_fast(Ft[(16 * mt + g) * TM_FT + 8 * kt + t], ah[mt][0], al[mt][0]);
        tf32_split_fast(Ft[(16 * mt + g + 8) * TM_FT + 8 * kt + t], ah[mt][1], al[mt][1]);
        tf32_split_fast(Ft[(16 * mt + g) * TM_FT + 8 * kt + t + 4], ah[mt][2], al[mt][2]);
        tf32_split_fast(Ft[(16 * mt + g + 8) * TM_FT + 8 * kt + t + 4], ah[mt][3], al[mt][3]);
      }
      // GEMM4: dW2^T[hid][k] += h^T[hid][pt] . go[pt][k]; B = go, columns k >= 4 of the padded tile are zero
      {
        const float q0 = g < 4 ? GOs[(8 * kt + t) * 4 + (g & 3)] : 0.f, q1 = g < 4 ? GOs[(8 * kt + t + 4) * 4 + (g & 3)] : 0.f;
#pragma unroll
        for (int mt = 0; mt < 4; ++mt) {
          uint32_t hh[4], hl[4];
          tf32_split_fast(Hs[(8 * kt + t) * TM_DP + 16 * mt + g], hh[0], hl[0]);
          tf32_split_fast(Hs[(8 * kt + t) * TM_DP + 16 * mt + g + 8], hh[1], hl[1]);
          tf32_split_fast(Hs[(8 * kt + t + 4) * TM_DP + 16 * mt + g], hh[2], hl[2]);
          tf32_split_fast(Hs[(8 * kt + t + 4) * TM_DP + 16 * mt + g + 8], hh[3], hl[3]);
          mma_3x(gw2[mt], hh, hl, q0, q1);
        }
      }
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {                        // B[k = pt][n = hid]
        const float d0 = DPs[(8 * kt + t) * TM_DP + 8 * nt + g], d1 = DPs[(8 * kt + t + 4) * TM_DP + 8 * nt + g];
        uint32_t b0h, b0l, b1h, b1l;
        tf32_split_fast(d0, b0h, b0l); tf32_split_fast(d1, b1h, b1l);
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
          mma_tf32(gw1[mt][nt], al[mt], b0h, b1h); mma_tf32(gw1[mt][nt], ah[mt], b0l, b1l); mma_tf32(gw1[mt][nt], ah[mt], b0h, b1h);
        }
      }
    }
    __syncwarp();                                             // the next batch overwrites the warp's buffers
  }
  // ---- flush: effective -> raw parameters (gains of networks_stylegan2.py:109-110), one global atomic per parameter per block ----
  __syncthreads();
  float* s_g = B1s + TP_HID;                                  // the warp areas
  for (int i = threadIdx.x; i < TP_DEC_PARAMS; i += blockDim.x) s_g[i] = 0.f;
  __syncthreads();
#pragma unroll
  for (int mt = 0; mt < 2; ++mt) {
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {                          // dW1[hid][ch] = dW1^T[ch][hid]
      const int ch = 16 * mt + g, hid = 8 * nt + 2 * t;
      atomicAdd(s_g + hid * TP_C + ch, gw1[mt][nt][0] * g1); atomicAdd(s_g + (hid + 1) * TP_C + ch, gw1[mt][nt][1] * g1);
      atomicAdd(s_g + hid * TP_C + ch + 8, gw1[mt][nt][2] * g1); atomicAdd(s_g + (hid + 1) * TP_C + ch + 8, gw1[mt][nt][3] * g1);
    }
  }
  float* gb1s = s_g + TP_HID * TP_C;
  float* gw2s = gb1s + TP_HID;
#pragma unroll
  for (int nt = 0; nt < 8; ++nt) {
    atomicAdd(gb1s + 8 * nt + 2 * t, gb1[nt][0] * lr_mul); atomicAdd(gb1s + 8 * nt + 2 * t + 1, gb1[nt][1] * lr_mul);
  }
  if (t < 2) {
#pragma unroll
    for (int mt = 0; mt < 4; ++mt) {
      atomicAdd(gw2s + (2 * t) * TP_HID + 16 * mt + g, gw2[mt][0] * g2); atomicAdd(gw2s + (2 * t + 1) * TP_HID + 16 * mt + g, gw2[mt][1] * g2);
      atomicAdd(gw2s + (2 * t) * TP_HID + 16 * mt + g + 8, gw2[mt][2] * g2); atomicAdd(gw2s + (2 * t + 1) * TP_HID + 16 * mt + g + 8, gw2[mt][3] * g2);
    }
  }
  if (t == 0) {
#pragma unroll
    for (int k = 0; k < 4; ++k) atomicAdd(gw2s + 4 * TP_HID + k, gb2[k] * lr_mul);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < TP_DEC_PARAMS; i += blockDim.x) atomicAdd(d_dec + i, s_g[i]);
}

// d(MipRayMarcher2.run_forward) (ray_marcher.py:25-57): one warp per ray, forward recomputed.  With G_i = dL/dw_i collected from
// every path (weights, rgb = sum w c_mid [+ 1 - sum w], depth = sum w z_mid / sum w), t_i = 1 - alpha_i + 1e-10, T_i = prod_{j<i} t_j:
//   dL/dalpha_i = G_i T_i - (sum_{k>i} G_k w_k) / t_i,   dL/ddens_mid_i = dL/dalpha_i * delta_i * exp(-dens_i delta_i) * sigmoid(pre_i),
//   dL/dc_mid_i = w_i dL/drgb;  mid-point values split their gradient half / half onto the two samples of the interval.
// The depth clamp to the call-global [min, max] never binds for a weighted mean of the ray's own mid-points; sum w = 0 (depth = inf ->
// clamped constant) passes no depth gradient.  Depths are not differentiated (the reference detaches the importance samples).
// the per-ray core: S sorted samples in shared memory -> per INTERVAL s_w[i] (weights) and s_dp[i] (dL/d density mid-point, before
// the shift of the softplus); s_T / s_G are scratch.  g_w: per-interval cotangents of this ray (global, may be null) plus the
// scalar g_w_all added to every interval (the cotangent of opacity = sum of the weights).
__device__ __forceinline__ void march_bwd_ray(const float* s_d, const float* s_sig, const float* s_r, const float* s_g, const float* s_b,
                                              int S, int white_back, float gr, float gg, float gb, float gd, const float* __restrict__ g_w,
                                              float g_w_all, float* s_T, float* s_G, float* s_w, float* s_dp, int lane) {
  // pass 1: forward; T_i, w_i per interval; sum w, sum w z_mid
  float carry = 1.f, wsum = 0.f, wz = 0.f;
  for (int base = 0; base < S - 1; base += 32) {
    const int i = base + lane;
    const bool on = i < S - 1;
    float alpha = 0.f, zmid = 0.f;
    if (on) {
      const float d0 = s_d[i], d1 = s_d[i + 1];
      zmid = 0.5f * (d0 + d1);
      const float dm = softplusf(0.5f * (s_sig[i] + s_sig[i + 1]) - 1.f);
      alpha = 1.f - expf(-dm * (d1 - d0));
    }
    float t = on ? (1.f - alpha) + 1e-10f : 1.f;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const float q = __shfl_up_sync(0xffffffffu, t, o); if (lane >= o) t *= q; }
    float excl = __shfl_up_sync(0xffffffffu, t, 1);
    if (lane == 0) excl = 1.f;
    const float T = carry * excl;
    carry *= __shfl_sync(0xffffffffu, t, 31);
    const float w = on ? alpha * T : 0.f;
    if (on) { s_T[i] = T; s_w[i] = w; }
    wsum += w; wz = fmaf(w, zmid, wz);
  }
  wsum = warp_sum(wsum); wz = warp_sum(wz);
  const bool depth_ok = gd != 0.f && wsum > 0.f;
  const float depth = depth_ok ? wz / wsum : 0.f;
  __syncwarp();
  for (int i = lane; i < S - 1; i += 32) {
    const float cr = 0.5f * (s_r[i] + s_r[i + 1]), cg = 0.5f * (s_g[i] + s_g[i + 1]), cb = 0.5f * (s_b[i] + s_b[i + 1]);
    float G = (g_w ? __ldg(g_w + i) : 0.f) + g_w_all + gr * cr + gg * cg + gb * cb;
    if (white_back) G -= gr + gg + gb;
    if (depth_ok) G += gd * (0.5f * (s_d[i] + s_d[i + 1]) - depth) / wsum;
    s_G[i] = G;
  }
  __syncwarp();
  // pass 2: backward over the intervals with a running suffix sum of G_k w_k
  float suffix = 0.f;
  for (int base = ((S - 2) / 32) * 32; base >= 0; base -= 32) {
    const int i = base + lane;
    const bool on = i < S - 1;
    const float gw = on ? s_G[i] * s_w[i] : 0.f;
    float incl = gw;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const float q = __shfl_down_sync(0xffffffffu, incl, o); if (lane + o < 32) incl += q; }
    const float after = suffix + (incl - gw);
    suffix += __shfl_sync(0xffffffffu, incl, 0);
    if (on) {
      const float d0 = s_d[i], d1 = s_d[i + 1], delta = d1 - d0;
      const float pre = 0.5f * (s_sig[i] + s_sig[i + 1]) - 1.f;
      const float dm = softplusf(pre);
      const float e = expf(-dm * delta);                   // 1 - alpha
      const float t = e + 1e-10f;
      const float dalpha = s_G[i] * s_T[i] - after / t;
      const float sg = pre > 20.f ? 1.f : 1.f / (1.f + __expf(-pre));
      s_dp[i] = dalpha * delta * e * sg;
    }
  }
  __syncwarp();
}

__global__ void __launch_bounds__(TP_WARPS * 32) mip_march_bwd_kernel(const float* __restrict__ colors, const float* __restrict__ dens,
    const float* __restrict__ depths, int64_t R, int S, int white_back, const float* __restrict__ g_rgb, const float* __restrict__ g_depth,
    const float* __restrict__ g_w, float* __restrict__ d_colors, float* __restrict__ d_dens) {
  extern __shared__ float smem[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  float* s = smem + (size_t)wib * 9 * S;
  float *s_d = s, *s_sig = s + S, *s_r = s + 2 * S, *s_g = s + 3 * S, *s_b = s + 4 * S;
  float *s_T = s + 5 * S, *s_G = s + 6 * S, *s_w = s + 7 * S, *s_dp = s + 8 * S;       // per interval
  for (int64_t r = (int64_t)blockIdx.x * TP_WARPS + wib; r < R; r += (int64_t)gridDim.x * TP_WARPS) {
    for (int i = lane; i < S; i += 32) {
      s_d[i] = __ldg(depths + r * S + i); s_sig[i] = __ldg(dens + r * S + i);
      s_r[i] = __ldg(colors + (r * S + i) * 3); s_g[i] = __ldg(colors + (r * S + i) * 3 + 1); s_b[i] = __ldg(colors + (r * S + i) * 3 + 2);
    }
    __syncwarp();
    const float gr = g_rgb ? __ldg(g_rgb + r * 3) : 0.f, gg = g_rgb ? __ldg(g_rgb + r * 3 + 1) : 0.f, gb = g_rgb ? __ldg(g_rgb + r * 3 + 2) : 0.f;
    const float gd = g_depth ? __ldg(g_depth + r) : 0.f;
    march_bwd_ray(s_d, s_sig, s_r, s_g, s_b, S, white_back, gr, gg, gb, gd, g_w ? g_w + r * (S - 1) : nullptr, 0.f, s_T, s_G, s_w, s_dp, lane);
    for (int i = lane; i < S; i += 32) {
      const float a = i < S - 1 ? s_dp[i] : 0.f, b = i > 0 ? s_dp[i - 1] : 0.f;
      d_dens[r * S + i] = 0.5f * (a + b);
      const float wc = 0.5f * ((i < S - 1 ? s_w[i] : 0.f) + (i > 0 ? s_w[i - 1] : 0.f));
      d_colors[(r * S + i) * 3 + 0] = wc * gr; d_colors[(r * S + i) * 3 + 1] = wc * gg; d_colors[(r * S + i) * 3 + 2] = wc * gb;
    }
    __syncwarp();
  }
}

// The renderer's backward up to the field (renderer.py:94-142 differentiated by hand): from the samples nerfb_triplane_render_train
// kept (depth | sigma | r | g | b, S stratified then Si importance, unsorted) and the cotangents of the six outputs, one warp per
// ray runs the marcher backward on the S coarse samples, rank-sorts all S + Si by depth exactly as the forward's unify_samples
// (value, then index), runs the marcher backward on the union, carries the result back through the ranks and writes, per sample in
// the ORIGINAL order: the point (rays_o + depth * rays_d, same rounding as the forward), dL/drgb and dL/dsigma -- the inputs of
// nerfb_triplane_field_bwd.  Depths carry no gradient (importance depths are detached, renderer.py:214).
__global__ void __launch_bounds__(TP_WARPS * 32) triplane_samples_bwd_kernel(const float* __restrict__ samples, const float* __restrict__ ro,
    const float* __restrict__ rd, int64_t R, int S, int Si, int white_back, const float* __restrict__ g_rgb_c, const float* __restrict__ g_depth_c,
    const float* __restrict__ g_op_c, const float* __restrict__ g_rgb_f, const float* __restrict__ g_depth_f, const float* __restrict__ g_op_f,
    float* __restrict__ coords, float* __restrict__ d_rgb, float* __restrict__ d_sigma) {
  extern __shared__ float smem[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int St = S + Si;
  float* A = smem + (size_t)wib * 19 * St;                   // samples as kept: [depth | sigma | r | g | b] x St
  float* B = A + 5 * St;                                      // the same, sorted by depth
  float *s_T = B + 5 * St, *s_G = s_T + St, *s_w = s_G + St, *s_dp = s_w + St;
  float* o_sig = s_dp + St;                                   // accumulated per ORIGINAL sample: dL/dsigma, dL/drgb
  float *o_r = o_sig + St, *o_g = o_r + St, *o_b = o_g + St;
  int* s_rank = reinterpret_cast<int*>(o_b + St);
  for (int64_t r = (int64_t)blockIdx.x * TP_WARPS + wib; r < R; r += (int64_t)gridDim.x * TP_WARPS) {
    for (int i = lane; i < 5 * St; i += 32) A[i] = __ldg(samples + r * 5 * St + i);
    const float ox = __ldg(ro + r * 3), oy = __ldg(ro + r * 3 + 1), oz = __ldg(ro + r * 3 + 2);
    const float dx = __ldg(rd + r * 3), dy = __ldg(rd + r * 3 + 1), dz = __ldg(rd + r * 3 + 2);
    __syncwarp();
    for (int i = lane; i < St; i += 32) {
      const float t = A[i];
      float* c = coords + (r * St + i) * 3;
      c[0] = __fadd_rn(ox, __fmul_rn(t, dx)); c[1] = __fadd_rn(oy, __fmul_rn(t, dy)); c[2] = __fadd_rn(oz, __fmul_rn(t, dz));
    }
    // ---- coarse marcher: samples [0, S) are ascending as drawn ----
    {
      const float gr = g_rgb_c ? __ldg(g_rgb_c + r * 3) : 0.f, gg = g_rgb_c ? __ldg(g_rgb_c + r * 3 + 1) : 0.f, gb = g_rgb_c ? __ldg(g_rgb_c + r * 3 + 2) : 0.f;
      march_bwd_ray(A, A + St, A + 2 * St, A + 3 * St, A + 4 * St, S, white_back, gr, gg, gb, g_depth_c ? __ldg(g_depth_c + r) : 0.f, nullptr,
                    g_op_c ? __ldg(g_op_c + r) : 0.f, s_T, s_G, s_w, s_dp, lane);
      for (int i = lane; i < St; i += 32) {
        float ds = 0.f, wc = 0.f;
        if (i < S) {
          ds = 0.5f * ((i < S - 1 ? s_dp[i] : 0.f) + (i > 0 ? s_dp[i - 1] : 0.f));
          wc = 0.5f * ((i < S - 1 ? s_w[i] : 0.f) + (i > 0 ? s_w[i - 1] : 0.f));
        }
        o_sig[i] = ds; o_r[i] = wc * gr; o_g[i] = wc * gg; o_b[i] = wc * gb;
      }
      __syncwarp();
    }
    // ---- unify_samples (renderer.py:160-170): the forward's rank-by-counting sort ----
    const bool c_sorted = coarse_is_sorted(A, S, lane);
    for (int k = lane; k < St; k += 32) {
      const int rank = rank_of_sample(A, S, St, k, c_sorted);
      s_rank[k] = rank;
#pragma unroll
      for (int q = 0; q < 5; ++q) B[q * St + rank] = A[q * St + k];
    }
    __syncwarp();
    {
      const float gr = g_rgb_f ? __ldg(g_rgb_f + r * 3) : 0.f, gg = g_rgb_f ? __ldg(g_rgb_f + r * 3 + 1) : 0.f, gb = g_rgb_f ? __ldg(g_rgb_f + r * 3 + 2) : 0.f;
      march_bwd_ray(B, B + St, B + 2 * St, B + 3 * St, B + 4 * St, St, white_back, gr, gg, gb, g_depth_f ? __ldg(g_depth_f + r) : 0.f, nullptr,
                    g_op_f ? __ldg(g_op_f + r) : 0.f, s_T, s_G, s_w, s_dp, lane);
      for (int k = lane; k < St; k += 32) {
        const int j = s_rank[k];
        const float ds = 0.5f * ((j < St - 1 ? s_dp[j] : 0.f) + (j > 0 ? s_dp[j - 1] : 0.f));
        const float wc = 0.5f * ((j < St - 1 ? s_w[j] : 0.f) + (j > 0 ? s_w[j - 1] : 0.f));
        d_sigma[r * St + k] = o_sig[k] + ds;
        float* c = d_rgb + (r * St + k) * 3;
        c[0] = o_r[k] + wc * gr; c[1] = o_g[k] + wc * gg; c[2] = o_b[k] + wc * gb;
      }
    }
    __syncwarp();
  }
}

// NERFB_TP_MMA=0 selects the CUDA-core decoder in all tri-plane kernels (read per call: tests flip it)
inline bool tp_use_mma() { const char* e = getenv("NERFB_TP_MMA"); return e ? atoi(e) != 0 : true; }

inline int grid_for(int64_t units) {
  const int64_t want = ceil_div64(units, TP_WARPS);
  const int64_t cap = (int64_t)NERFB_SM_COUNT * 8;           // 8 resident 256-thread CTAs per SM
  return (int)(want < cap ? (want < 1 ? 1 : want) : cap);
}

}  // namespace

extern "C" int nerfb_triplane_relayout(const float* planes, int channels, int H, int W, float* planes_cl, void* stream) {
  NERFB_CHECK_ARG(planes && planes_cl && H >= 1 && W >= 1);
  if (channels != TP_C) return NERFB_E_UNSUPPORTED;
  const int64_t HW = (int64_t)H * W;
  NERFB_LAUNCH(triplane_relayout_kernel, dim3((unsigned)ceil_div64(HW, 32), 3), 256, 0, stream, planes, planes_cl, HW);
  return NERFB_OK;
}

extern "C" int nerfb_triplane_field(const float* planes_cl, int H, int W, const float* decoder, float lr_mul,
                                    const float* coords, int64_t n_points, float box_warp, float* rgb, float* sigma,
                                    void* stream) {
  NERFB_CHECK_ARG(planes_cl && decoder && rgb && sigma && n_points >= 0 && H >= 1 && W >= 1 && box_warp != 0.f);
  if (n_points == 0) return NERFB_OK;
  NERFB_CHECK_ARG(coords);
  if (tp_use_mma()) {
    const int64_t want = ceil_div64(ceil_div64(n_points, TM_PTS), TP_WARPS);
    const int64_t cap = (int64_t)NERFB_SM_COUNT * 3;
    NERFB_LAUNCH(triplane_field_mma_kernel, (int)(want < cap ? want : cap), TP_WARPS * 32, 0, stream, planes_cl, H, W, decoder, lr_mul,
                 coords, n_points, (float)(2.0 / (double)box_warp), rgb, sigma);
    return NERFB_OK;
  }
  NERFB_LAUNCH(triplane_field_kernel, grid_for(n_points), TP_WARPS * 32, 0, stream, planes_cl, H, W, decoder, lr_mul, coords,
               n_points, (float)(2.0 / (double)box_warp), rgb, sigma);
  return NERFB_OK;
}

extern "C" int nerfb_mip_march(const float* colors, const float* densities, const float* depths, int64_t n_rays,
                               int n_samples, int white_back, float* rgb, float* depth, float* weights, void* workspace,
                               void* stream) {
  NERFB_CHECK_ARG(rgb && depth && workspace && n_rays >= 0 && n_samples >= 2);
  if (n_samples > 2048) return NERFB_E_UNSUPPORTED;
  if (n_rays == 0) return NERFB_OK;
  NERFB_CHECK_ARG(colors && densities && depths);
  uint32_t* mm = static_cast<uint32_t*>(workspace);
  const size_t smem = (size_t)TP_WARPS * 6 * n_samples * sizeof(float);
  if (smem > 48 * 1024) NERFB_CUDA(cudaFuncSetAttribute(mip_march_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  NERFB_LAUNCH(minmax_init_kernel, 1, 32, 0, stream, mm);
  NERFB_LAUNCH(mip_march_kernel, grid_for(n_rays), TP_WARPS * 32, smem, stream, colors, densities, depths, n_rays, n_samples,
               white_back, rgb, depth, weights, mm);
  NERFB_LAUNCH(depth_clamp_kernel, (unsigned)ceil_div64(n_rays, 256), 256, 0, stream, depth, (float*)nullptr, n_rays, mm);
  return NERFB_OK;
}

static int triplane_render_impl(const float* planes_cl, int H, int W, const float* decoder, float lr_mul,
                                const float* rays_o, const float* rays_d, int64_t n_rays, int n_samples,
                                int n_importance, double ray_start, double ray_end, const float* ray_start_arr,
                                const float* ray_end_arr, int disparity, float box_warp, int white_back,
                                const float* rand_coarse, const float* u, uint64_t seed, float* rgb_coarse,
                                float* depth_coarse, float* opacity_coarse, float* rgb_final, float* depth_final,
                                float* opacity_final, float* depths_fine_out, float* samples_out, void* workspace, void* stream) {
  NERFB_CHECK_ARG(planes_cl && decoder && rgb_coarse && depth_coarse && opacity_coarse && rgb_final && depth_final &&
                  opacity_final && workspace);
  NERFB_CHECK_ARG(n_rays >= 0 && n_samples >= 4 && n_importance >= 1 && H >= 1 && W >= 1 && box_warp != 0.f);
  NERFB_CHECK_ARG((ray_start_arr == nullptr) == (ray_end_arr == nullptr));
  if (n_samples + n_importance > TP_MAX_S) return NERFB_E_UNSUPPORTED;
  if (n_rays == 0) return NERFB_OK;
  NERFB_CHECK_ARG(rays_o && rays_d);
  RenderArgs a{};
  a.planes = planes_cl; a.H = H; a.W = W; a.dec = decoder; a.lr_mul = lr_mul; a.o = rays_o; a.d = rays_d; a.R = n_rays;
  a.S = n_samples; a.Si = n_importance; a.ray_start = (float)ray_start; a.ray_end = (float)ray_end; a.start_arr = ray_start_arr;
  a.inv_start = (float)(1.0 / ray_start); a.inv_end = (float)(1.0 / ray_end); a.delta = (float)((ray_end - ray_start) / (double)(n_samples - 1));
  a.end_arr = ray_end_arr; a.disparity = disparity; a.coord_scale = (float)(2.0 / (double)box_warp); a.white_back = white_back;
  a.rand_coarse = rand_coarse; a.u = u; a.seed = seed;
  a.rgb_c = rgb_coarse; a.depth_c = depth_coarse; a.opac_c = opacity_coarse; a.rgb_f = rgb_final; a.depth_f = depth_final;
  a.opac_f = opacity_final; a.depths_fine_out = depths_fine_out; a.samples_out = samples_out; a.mm = static_cast<uint32_t*>(workspace);
  const bool mma = tp_use_mma();
  const size_t per_warp = ((11 * (n_samples + n_importance) + 2 * n_samples + 3) & ~3) + (mma ? TM_PTS * TM_FS : 128);
  const size_t smem = ((size_t)TP_WARPS * per_warp + (mma ? TM_DEC_FLOATS : 0)) * sizeof(float);
  if (smem > 200 * 1024) return NERFB_E_UNSUPPORTED;
  if (smem > 48 * 1024)
    NERFB_CUDA(mma ? cudaFuncSetAttribute(triplane_render_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)
                   : cudaFuncSetAttribute(triplane_render_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  NERFB_LAUNCH(minmax_init_kernel, 1, 32, 0, stream, a.mm);
  // grid = resident CTAs only, rays strided (the CUDA-core variant is register-heavy: 64 decoder weights per lane)
  const int64_t want = ceil_div64(n_rays, TP_WARPS);
  const int64_t cap = (int64_t)NERFB_SM_COUNT * (mma ? 3 : 2);
  if (mma) NERFB_LAUNCH(triplane_render_kernel<true>, (int)(want < cap ? want : cap), TP_WARPS * 32, smem, stream, a);
  else NERFB_LAUNCH(triplane_render_kernel<false>, (int)(want < cap ? want : cap), TP_WARPS * 32, smem, stream, a);
  NERFB_LAUNCH(depth_clamp_kernel, (unsigned)ceil_div64(n_rays, 256), 256, 0, stream, depth_coarse, depth_final, n_rays, a.mm);
  return NERFB_OK;
}

extern "C" int nerfb_triplane_render(const float* planes_cl, int H, int W, const float* decoder, float lr_mul,
                                     const float* rays_o, const float* rays_d, int64_t n_rays, int n_samples,
                                     int n_importance, double ray_start, double ray_end, const float* ray_start_arr,
                                     const float* ray_end_arr, int disparity, float box_warp, int white_back,
                                     const float* rand_coarse, const float* u, uint64_t seed, float* rgb_coarse,
                                     float* depth_coarse, float* opacity_coarse, float* rgb_final, float* depth_final,
                                     float* opacity_final, float* depths_fine_out, void* workspace, void* stream) {
  return triplane_render_impl(planes_cl, H, W, decoder, lr_mul, rays_o, rays_d, n_rays, n_samples, n_importance, ray_start, ray_end,
                              ray_start_arr, ray_end_arr, disparity, box_warp, white_back, rand_coarse, u, seed, rgb_coarse, depth_coarse,
                              opacity_coarse, rgb_final, depth_final, opacity_final, depths_fine_out, nullptr, workspace, stream);
}

// The same launch, additionally writing every sample's (depth, sigma, r, g, b): the forward of a training step.
extern "C" int nerfb_triplane_render_train(const float* planes_cl, int H, int W, const float* decoder, float lr_mul,
                                           const float* rays_o, const float* rays_d, int64_t n_rays, int n_samples,
                                           int n_importance, double ray_start, double ray_end, const float* ray_start_arr,
                                           const float* ray_end_arr, int disparity, float box_warp, int white_back,
                                           const float* rand_coarse, const float* u, uint64_t seed, float* rgb_coarse,
                                           float* depth_coarse, float* opacity_coarse, float* rgb_final, float* depth_final,
                                           float* opacity_final, float* samples_out, void* workspace, void* stream) {
  NERFB_CHECK_ARG(samples_out != nullptr);
  return triplane_render_impl(planes_cl, H, W, decoder, lr_mul, rays_o, rays_d, n_rays, n_samples, n_importance, ray_start, ray_end,
                              ray_start_arr, ray_end_arr, disparity, box_warp, white_back, rand_coarse, u, seed, rgb_coarse, depth_coarse,
                              opacity_coarse, rgb_final, depth_final, opacity_final, nullptr, samples_out, workspace, stream);
}

// ---- backward entries (EG3D training) ------------------------------------------------------------------------------
extern "C" int nerfb_triplane_field_bwd(const float* planes_cl, int H, int W, const float* decoder, float lr_mul,
                                        const float* coords, int64_t n_points, float box_warp, const float* d_rgb,
                                        const float* d_sigma, float* d_planes_cl, float* d_decoder, void* stream) {
  NERFB_CHECK_ARG(planes_cl && decoder && d_planes_cl && d_decoder && n_points >= 0 && H >= 1 && W >= 1 && box_warp != 0.f);
  if (n_points == 0) return NERFB_OK;
  NERFB_CHECK_ARG(coords && d_rgb && d_sigma);
  if (tp_use_mma()) {                                              // tensor-core decoder (default); 0 selects the CUDA-core kernel
    static bool attr_set[64] = {};
    int dev = 0;
    NERFB_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
      NERFB_CUDA(cudaFuncSetAttribute(triplane_field_bwd_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TM_SMEM));
      attr_set[dev] = true;
    }
    const int64_t want_m = ceil_div64(ceil_div64(n_points, TM_PTS), TM_WARPS);
    const int grid_m = (int)(want_m < NERFB_SM_COUNT ? want_m : NERFB_SM_COUNT);
    NERFB_LAUNCH(triplane_field_bwd_mma_kernel, grid_m, TM_WARPS * 32, TM_SMEM, stream, planes_cl, H, W, decoder, lr_mul, coords,
                 n_points, (float)(2.0 / (double)box_warp), d_rgb, d_sigma, d_planes_cl, d_decoder);
    return NERFB_OK;
  }
  const int64_t want = ceil_div64(n_points, TPB_TILE);
  const int grid = (int)(want < NERFB_SM_COUNT * 3 ? want : NERFB_SM_COUNT * 3);     // one decoder-gradient flush per block
  NERFB_LAUNCH(triplane_field_bwd_kernel, grid, TP_WARPS * 32, 0, stream, planes_cl, H, W, decoder, lr_mul, coords, n_points,
               (float)(2.0 / (double)box_warp), d_rgb, d_sigma, d_planes_cl, d_decoder);
  return NERFB_OK;
}

extern "C" int nerfb_triplane_samples_bwd(const float* samples, const float* rays_o, const float* rays_d, int64_t n_rays, int n_samples,
                                          int n_importance, int white_back, const float* d_rgb_coarse, const float* d_depth_coarse,
                                          const float* d_opacity_coarse, const float* d_rgb_final, const float* d_depth_final,
                                          const float* d_opacity_final, float* coords, float* d_rgb, float* d_sigma, void* stream) {
  NERFB_CHECK_ARG(coords && d_rgb && d_sigma && n_rays >= 0 && n_samples >= 4 && n_importance >= 1);
  if (n_samples + n_importance > TP_MAX_S) return NERFB_E_UNSUPPORTED;
  if (n_rays == 0) return NERFB_OK;
  NERFB_CHECK_ARG(samples && rays_o && rays_d);
  const size_t smem = (size_t)TP_WARPS * 19 * (n_samples + n_importance) * sizeof(float);
  if (smem > 200 * 1024) return NERFB_E_UNSUPPORTED;
  if (smem > 48 * 1024) NERFB_CUDA(cudaFuncSetAttribute(triplane_samples_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  NERFB_LAUNCH(triplane_samples_bwd_kernel, grid_for(n_rays), TP_WARPS * 32, smem, stream, samples, rays_o, rays_d, n_rays, n_samples,
               n_importance, white_back, d_rgb_coarse, d_depth_coarse, d_opacity_coarse, d_rgb_final, d_depth_final, d_opacity_final,
               coords, d_rgb, d_sigma);
  return NERFB_OK;
}

extern "C" int nerfb_mip_march_bwd(const float* colors, const float* densities, const float* depths, int64_t n_rays,
                                   int n_samples, int white_back, const float* d_rgb, const float* d_depth,
                                   const float* d_weights, float* d_colors, float* d_densities, void* stream) {
  NERFB_CHECK_ARG(d_colors && d_densities && n_rays >= 0 && n_samples >= 2);
  if (n_samples > 512) return NERFB_E_UNSUPPORTED;
  if (n_rays == 0) return NERFB_OK;
  NERFB_CHECK_ARG(colors && densities && depths);
  const size_t smem = (size_t)TP_WARPS * 9 * n_samples * sizeof(float);
  if (smem > 48 * 1024) NERFB_CUDA(cudaFuncSetAttribute(mip_march_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  NERFB_LAUNCH(mip_march_bwd_kernel, grid_for(n_rays), TP_WARPS * 32, smem, stream, colors, densities, depths, n_rays, n_samples,
               white_back, d_rgb, d_depth, d_weights, d_colors, d_densities);
  return NERFB_OK;
}
