// Kernel (1), tensor-core arm, forward: fused positional encoding + NeRF MLP on tcgen05 / TMEM.
//   replaces models/rendering.py:129-159 (embed -> cat -> model per 32k-point chunk) and models/nerf.py:100-124.
//
// One persistent CTA per SM walks 128-sample tiles; per tile the activations never leave the SM.  Warp roles:
//   warp 0      (1 lane)  weight producer: streams the pre-packed bf16 weight slabs (already in the swizzled
//                         shared-memory image, so a slab is ONE 32 KB bulk copy) from L2 into a 3-stage ring;
//   warp 1      (1 lane)  MMA issuer: tcgen05.mma 128 x N x 16 (N = 256 / 128), A from the activation / encoding blocks,
//                         B from the ring, fp32 accumulators ping-ponging between TMEM columns [0,256) and [256,512);
//   warp 2      (1 lane)  store lane (training only): bulk-stores every published bf16 block to the tile's record in
//                         HBM for the backward, off the MMA thread's critical path;
//   warps 3-10  epilogue: tcgen05.ld 64 accumulator columns -> + bias (packed f32x2) -> ReLU folded into the bf16
//                         convert -> st.shared.v4 as the next layer's A operand (128B-swizzled K-major blocks).  The two
//                         warp groups own alternate 64-column blocks, so blocks 0 and 1 are published concurrently;
//   warps 11-14 encoder:  build the bf16 encoding tile X [128 x 64] (x = o + d z, exact sincosf + double-angle ladder)
//                         one tile AHEAD into a double buffer, and rewrite it as the direction encoding as soon as the
//                         last xyz reader's MMAs have completed.
// The layer chain is pipelined at K-block granularity: the MMA of layer l+1 starts on block 0 while blocks 1..3 of layer
// l are still in the epilogue, into the other accumulator.  Skip / direction concatenations are extra K-blocks (the
// resident X tile), never a cat.  sigma (256->1) and rgb (128->3, sigmoid) heads are fp32 dot products in the epilogue.
//
// SIREN / FiLM variant (models/nerf.py:142-216): same kernel, template parameter SIREN.  sin(freq (W x + b) + phase) =
// sin((freq W) x + (freq b + phase)): both folds happen when the weights / biases are packed, so the epilogue is a bias
// add and MUFU.SIN; with SAVE it also records sign(cos) as the backward's gate mask (mlp_tc_bwd.cu).  The 3-wide
// first layer (and the 3 direction columns of the colour layer) are carried as 9 K-columns [x_hi x_lo x_hi] against
// [W_hi W_hi W_lo], which restores ~16 mantissa bits where the first-layer frequency scale w0 = 30 amplifies input
// rounding; the colour layer is N = 256 and its 256->3 sigmoid head rides in the epilogue like NeRF's.
#include <cstdlib>
#include "common.cuh"
#include "tc_common.cuh"

using namespace tc;

// -DNERFB_TRACE_BUILD compiles the clock64 timeline probes in (tools/trace_fwd.py); the product build has none
#ifdef NERFB_TRACE_BUILD
#define NERFB_TRACE(...) __VA_ARGS__
#else
#define NERFB_TRACE(...)
#endif

constexpr int N_STAGES = 3;
constexpr int STAGE_BYTES = 256 * 128;          // one [256 x 64] bf16 weight slab
// the SMSP arbiter favours the highest warp id: roles that mostly WAIT get the low ids, the epilogue the high ones
constexpr int WARP_PRODUCER = 0, WARP_STORE = 1, WARP_ENC0 = 2, WARP_EPI0 = 6, WARP_MMA = 22, WARP_END = 22;
constexpr int NUM_THREADS = 23 * 32;
constexpr int EPI_THREADS = 512;
constexpr uint32_t N_EPI_WARPS = EPI_THREADS / 32;
constexpr int MAX_STEPS = 12;
constexpr int SRC_X = 4;

struct TcStep {
  int n_kb;            // K-blocks of 64
  int src[6];          // operand-A source per K-block: 0..3 = activation block, SRC_X = encoding tile
  int N;               // 256 or 128 output columns
  int relu;            // activation in the epilogue
  int out_blocks;      // 64-column bf16 blocks the epilogue writes back as the next A operand (0 = none)
  int bias_off;        // float offset into the shared bias table
  int head;            // 0 none, 1 sigma head on this step's activations, 2 rgb head (direction layer)
  int wait_x;          // the MMA must wait for a freshly built X tile before this step's SRC_X block
  int x_done;          // this step is the last reader of the current X contents (xyz or direction encoding)
  int save_off;        // byte offset of this step's output blocks in the tile's saved record (-1 = not saved)
  int store_only;      // the output blocks feed no later MMA (saved for the backward only): published on s_ready
  int film;            // SIREN: FiLM layer index of this step's frequency / phase vectors
  int mask_off;        // byte offset of this step's ReLU gate bit-mask (128 rows x 32 B) in the tile's record, -1 = none
  int reuse_a;         // the activation blocks were already waited for by the previous step (class head before direction layer)
};

constexpr int MAX_OPS = 64;
// One K-block of MMAs, pre-decoded for the issuing thread (indexed parameter-bank loads cost ~100 cycles each on the
// critical path; the list is copied to shared memory once).  bits: 0 A = X tile, 1 wait before use, 2 first of step,
// 3 last of step, 4 last reader of X, 5-6 activation block, 7 N = 128
constexpr uint32_t OP_X = 1u, OP_WAIT = 2u, OP_FIRST = 4u, OP_LAST = 8u, OP_XDONE = 16u, OP_N128 = 128u, OP_N16 = 256u;
// packed per-step word for the epilogue: N | relu << 9 | out_blocks << 10 | head << 13 | store_only << 15 | saved << 16
struct TcProgram {
  int n_steps;
  TcStep s[MAX_STEPS];
  int n_ops;
  uint32_t ops[MAX_OPS];
  uint32_t step_word[MAX_STEPS];
  int n_bias;                      // floats in the bias table
  int save_x_off, save_dir_off;    // saved-record offsets of the two encoding tiles
  int tile_save_bytes;             // saved bytes per tile (0 = nothing saved)
  int n_freq_xyz, n_freq_dir, in_xyz, in_dir;
  int sigma_only, out_c;
  int n_cls;                       // NeRF_3D class head width (models/nerf_cls.py:56-61), 0 = plain NeRF
  int siren, rgb_in;               // rgb head fan-in: 128 (NeRF direction layer) or 256 (SIREN colour layer)
  float xyz_scale;                 // SIREN: UniformBoxWarp 2 / sidelength (models/nerf.py:134-140)
};

struct TcFwdArgs {
  TcProgram prog;
  TcComposite comp;
  const uint8_t* wblob;            // packed bf16 weight slabs in schedule order
  const float* bias;               // fp32 bias table [n_bias] | w_sigma[256] | b_sigma | w_rgb[3*rgb_in] | b_rgb[3]
  const float* scale;              // SIREN: per-column frequency table [n_bias] (bias table then holds freq*b + phase)
  const float* in0; const float* in1; int in_mode; int S;
  int64_t P; int64_t n_tiles;
  float* out;
  uint8_t* save;                   // saved activations, tile-major
  long long* trace;                // diagnostics: clock64 timeline of block 0 (tools/trace_fwd.py), normally null
};

// shared memory map (dynamic, 1024-aligned)
struct __align__(1024) Smem {
  uint8_t a[4][BLK_BYTES];               // activation tile: 4 K-blocks
  uint8_t x[2][BLK_BYTES];               // encoding tiles (double buffer: tile t uses x[t & 1])
  uint8_t ring[N_STAGES][STAGE_BYTES];   // weight slabs
  float bias[10 * 256 + 128];
  float w_sigma[256];
  float w_rgb[3 * 256];
  float b_heads[4];                      // b_sigma, b_rgb[3]
  float part[4][3][TILE_M];              // head partial sums [column quarter][channel][row]
  float sigma[TILE_M];
  float comp[5][TILE_M];                 // epilogue compositing: w, w r, w g, w b, w z per row
  float comp_tot[4];                     // transmittance product of each 32-row chunk
  uint64_t w_full[N_STAGES], w_empty[N_STAGES];
  uint64_t a_ready[4];                   // block j holds the current step's output (16 epilogue warps arrive)
  uint64_t s_ready[4];                   // training: block j of a store-only step (direction / colour layer activations)
  uint64_t a_free[4];                    // training: the store lane has read block j out
  uint64_t x_ready[2], x_consumed[2], x_stored[2];
  uint64_t acc_full[2];
  uint32_t tmem_base;
  uint4 ops[2][MAX_OPS];                 // per X buffer: {A descriptor low word, barrier address, flags, parity bit}
  uint32_t step_word[MAX_STEPS], step_bias[MAX_STEPS];
  int step_mask[MAX_STEPS];
};

static_assert(sizeof(Smem) + 1024 <= 232448, "shared memory budget (227 KB per CTA)");

// one bf16 row of 64 columns -> the 128B-swizzled block image
__device__ __forceinline__ void store_row64(uint32_t base, int row, const float (&buf)[64]) {
#pragma unroll
  for (int ch = 0; ch < 8; ++ch) {
    const uint32_t addr = base + (uint32_t)row * 128u + ((((uint32_t)ch) ^ ((uint32_t)row & 7u)) << 4);
    sts128(addr, pack_bf16(buf[8 * ch + 0], buf[8 * ch + 1]), pack_bf16(buf[8 * ch + 2], buf[8 * ch + 3]),
           pack_bf16(buf[8 * ch + 4], buf[8 * ch + 5]), pack_bf16(buf[8 * ch + 6], buf[8 * ch + 7]));
  }
}

// xyz / dir -> bf16 encoding row [v(3), sin(2^k v)(3), cos(2^k v)(3), ...] (models/nerf.py:33-38), zero beyond the width.
// Exact sincosf at the base frequency, then the double-angle recurrence per octave (freq_bands are exact powers of 2).
__device__ __forceinline__ void encode_row(float vx, float vy, float vz, int n_freq, uint32_t base, int row) {
  float s[3], c[3];
  const float v[3] = {vx, vy, vz};
#pragma unroll
  for (int j = 0; j < 3; ++j) sincosf(v[j], &s[j], &c[j]);
  float buf[64];
#pragma unroll
  for (int i = 0; i < 64; ++i) buf[i] = 0.f;
#pragma unroll
  for (int j = 0; j < 3; ++j) buf[j] = v[j];
#pragma unroll
  for (int k = 0; k < 10; ++k) {
    const bool on = k < n_freq;
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      buf[3 + 6 * k + j] = on ? s[j] : 0.f;
      buf[6 + 6 * k + j] = on ? c[j] : 0.f;
      const float s2 = 2.f * s[j] * c[j];            // double-angle step to the next octave
      const float c2 = 1.f - 2.f * s[j] * s[j];
      s[j] = s2; c[j] = c2;
    }
  }
  store_row64(base, row, buf);
}

// SIREN input row: v (3 values, fp32) as 9 bf16 columns [hi(3) lo(3) hi(3)], hi = bf16(v), lo = bf16(v - hi)
__device__ __forceinline__ void hilo_row(float vx, float vy, float vz, uint32_t base, int row) {
  float buf[64];
#pragma unroll
  for (int i = 0; i < 64; ++i) buf[i] = 0.f;
  const float v[3] = {vx, vy, vz};
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    const float hi = __bfloat162float(__float2bfloat16_rn(v[j]));
    buf[j] = hi; buf[3 + j] = v[j] - hi; buf[6 + j] = hi;
  }
  store_row64(base, row, buf);
}

// already-embedded input: src[0..width) (zero padded to 64 columns) -> X tile row
__device__ __forceinline__ void copy_row(const float* src, int width, uint32_t base, int row) {
  float buf[64];
#pragma unroll
  for (int i = 0; i < 64; ++i) buf[i] = (src && i < width) ? __ldg(src + i) : 0.f;
  store_row64(base, row, buf);
}

// ---- compositing in the epilogue (the 4 epilogue warps that own column quarter 0; warp q holds rows 32 q .. 32 q + 31) ----
// Same operations in the same order as composite_fwd_kernel (composite.cu): 32-sample chunks, inclusive warp product, running
// carry over the chunks of a ray, per-lane accumulation over chunks, butterfly sum -- so the result is bit-identical to
// running kernel (4) on the raw values this epilogue would have stored.
struct CompTerms { float alpha, t; };
__device__ __forceinline__ CompTerms comp_terms(float sigma, float noise, float noise_std, float z_i, float z_next, bool last, float dnorm) {
  CompTerms s;
  const float d = last ? 1e10f : __fsub_rn(z_next, z_i);                     // models/rendering.py:162-163
  const float delta = __fmul_rn(d, dnorm);                                   // :168
  const float pre = __fadd_rn(sigma, __fmul_rn(noise, noise_std));           // :170-173
  const float a = pre > 0.f ? pre : 0.f;
  s.alpha = __fsub_rn(1.0f, expf(-__fmul_rn(delta, a)));
  s.t = __fadd_rn(__fsub_rn(1.0f, s.alpha), 1e-10f);                         // :175
  return s;
}
template <typename SmemT>
__device__ __forceinline__ void epilogue_composite(SmemT& sm, const TcComposite& cp, const float* __restrict__ rays,
                                                   const float* __restrict__ zbuf, int64_t p, int64_t P, bool live, int q, int lane,
                                                   float cr, float cg, float cb, float sigma, bool has_rgb) {
  const int S = cp.S, row = q * 32 + lane;
  const int i = row % S;
  const int64_t ray = p / S;
  const float zi = live ? __ldg(zbuf + p) : 0.f;
  float zn = __shfl_down_sync(0xffffffffu, zi, 1);
  if (lane == 31 && live && i + 1 < S) zn = __ldg(zbuf + p + 1);
  float dnorm = 0.f, nz = 0.f;
  if (live) {
    const float dx = __ldg(rays + ray * 8 + 3), dy = __ldg(rays + ray * 8 + 4), dz = __ldg(rays + ray * 8 + 5);
    dnorm = sqrtf(dx * dx + dy * dy + dz * dz);
    if (cp.noise) nz = __ldg(cp.noise + p);
  }
  const CompTerms st = comp_terms(sigma, nz, cp.noise_std, zi, zn, i == S - 1, dnorm);
  float incl = live ? st.t : 1.0f;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const float t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl *= t;
  }
  float excl = __shfl_up_sync(0xffffffffu, incl, 1);
  if (lane == 0) excl = 1.0f;
  if (lane == 31) sm.comp_tot[q] = incl;
  named_bar_sync(3, 128);
  float carry = 1.0f;
  for (int k = q - (i >> 5); k < q; ++k) carry *= sm.comp_tot[k];             // chunks of this ray before mine, in order
  const float w = live ? st.alpha * (carry * excl) : 0.f;                    // :176-177
  if (live && cp.weights) cp.weights[p] = w;
  sm.comp[0][row] = w;
  if (has_rgb) { sm.comp[1][row] = w * cr; sm.comp[2][row] = w * cg; sm.comp[3][row] = w * cb; sm.comp[4][row] = w * zi; }
  named_bar_sync(3, 128);
  if (q < TILE_M / S) {                                                      // warp q reduces ray q of this tile
    const int64_t r = (p - row) / S + q;
    float acc[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
    for (int c = 0; c < S; c += 32) {
#pragma unroll
      for (int k = 0; k < 5; ++k)
        if (k == 0 || has_rgb) acc[k] += (lane + c < S) ? sm.comp[k][q * S + c + lane] : 0.f;
    }
#pragma unroll
    for (int k = 0; k < 5; ++k)
      if (k == 0 || has_rgb) acc[k] = warp_sum(acc[k]);
    if (lane == 0 && r * S < P) {
      if (cp.opacity) cp.opacity[r] = acc[0];
      if (has_rgb) {
        const float bg = cp.white_back ? (1.0f - acc[0]) : 0.f;              // :187-188
        cp.rgb[r * 3 + 0] = acc[1] + bg; cp.rgb[r * 3 + 1] = acc[2] + bg; cp.rgb[r * 3 + 2] = acc[3] + bg;
        cp.depth[r] = acc[4];
      }
    }
  }
}

template <bool SIREN, bool CLS, bool SAVE>
__global__ void __launch_bounds__(NUM_THREADS, 1) mlp_tc_fwd_kernel(const __grid_constant__ TcFwdArgs args) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  Smem& sm = *reinterpret_cast<Smem*>(smem_raw);            // keeps the shared address space: native LDS / STS
  if ((smem_u32(smem_raw) & 1023u) != 0u) __trap();         // the swizzled images need a 1024-byte aligned base
  const TcProgram& pg = args.prog;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // the inference instantiation carries none of the save / gate-mask code; the training one keeps the flag a run-time
  // value (ptxas allocates the epilogue better that way: 56 B of spills instead of 184 B)
  const bool saving = SAVE ? (args.save != nullptr) : false;

  // ---- one-time setup -----------------------------------------------------------------------------------
  for (int i = threadIdx.x; i < pg.n_bias; i += NUM_THREADS) sm.bias[i] = __ldg(args.bias + i);
  for (int i = threadIdx.x; i < 256; i += NUM_THREADS) sm.w_sigma[i] = __ldg(args.bias + pg.n_bias + i);
  if (threadIdx.x == 0) sm.b_heads[0] = __ldg(args.bias + pg.n_bias + 256);
  constexpr int RGB_IN = SIREN ? 256 : 128;
  for (int i = threadIdx.x; i < 3 * RGB_IN; i += NUM_THREADS) sm.w_rgb[i] = __ldg(args.bias + pg.n_bias + 257 + i);
  if (threadIdx.x < 3) sm.b_heads[1 + threadIdx.x] = __ldg(args.bias + pg.n_bias + 257 + 3 * RGB_IN + threadIdx.x);
  if (threadIdx.x < 2 * pg.n_ops) {       // decode the MMA op list once, for both X buffers
    const int b = threadIdx.x / pg.n_ops, i = threadIdx.x % pg.n_ops;
    const uint32_t w = pg.ops[i];
    uint4 o;
    if (w & OP_X) {
      o.x = ((smem_u32(sm.x[b]) >> 4) & 0x3FFFu) | (1u << 16);
      o.y = (w & OP_WAIT) ? smem_u32(&sm.x_ready[b]) : 0u;
      o.w = 4u + (uint32_t)b;
    } else {
      const uint32_t src = (w >> 5) & 3u;
      o.x = ((smem_u32(sm.a[src]) >> 4) & 0x3FFFu) | (1u << 16);
      o.y = (w & OP_WAIT) ? smem_u32(&sm.a_ready[src]) : 0u;      // a second reader of the same blocks does not wait again
      o.w = src;
    }
    if (!o.y) o.w = 31u;                  // no wait: flip a spare parity bit
    o.z = w;
    sm.ops[b][i] = o;
  }
  if (threadIdx.x < pg.n_steps) { sm.step_word[threadIdx.x] = pg.step_word[threadIdx.x]; sm.step_bias[threadIdx.x] = (uint32_t)pg.s[threadIdx.x].bias_off; sm.step_mask[threadIdx.x] = pg.s[threadIdx.x].mask_off; }
  if (threadIdx.x == 0) {
    for (int i = 0; i < N_STAGES; ++i) { mbar_init(&sm.w_full[i], 1); mbar_init(&sm.w_empty[i], 1); }
    for (int i = 0; i < 4; ++i) { mbar_init(&sm.a_ready[i], N_EPI_WARPS); mbar_init(&sm.a_free[i], 1); }
    for (int i = 0; i < 4; ++i) mbar_init(&sm.s_ready[i], N_EPI_WARPS);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&sm.x_ready[i], 4); mbar_init(&sm.x_consumed[i], 1); mbar_init(&sm.x_stored[i], 1);
      mbar_init(&sm.acc_full[i], 1);
    }
    fence_barrier_init();
  }
  if (warp == WARP_MMA) tmem_alloc(&sm.tmem_base, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp == WARP_PRODUCER) {
    // ===== weight producer =====
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int64_t tile = blockIdx.x; tile < args.n_tiles; tile += gridDim.x) {
        const uint8_t* src = args.wblob;
        for (int s = 0; s < pg.n_steps; ++s) {
          const uint32_t bytes = (uint32_t)pg.s[s].N * 128u;
          for (int kb = 0; kb < pg.s[s].n_kb; ++kb) {
            mbar_wait_relaxed(&sm.w_empty[stage], phase ^ 1);
            mbar_expect_tx(&sm.w_full[stage], bytes);
            bulk_g2s(sm.ring[stage], src, bytes, &sm.w_full[stage]);
            src += bytes;
            if (++stage == N_STAGES) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == WARP_MMA) {
    // ===== MMA issuer (single thread) =====
    // Everything this thread needs per K-block is pre-decoded in shared memory (one 16-byte load): it must spend as few
    // cycles as possible away from issuing, because the tensor pipe only queues a couple of MMAs.
    if (lane == 0) {
      uint32_t stage = 0, phase = 0, par = 0, gs = 0, it = 0;
      const int n_ops = pg.n_ops;
      const uint32_t ring_lo0 = ((smem_u32(sm.ring[0]) >> 4) & 0x3FFFu) | (1u << 16);
      const uint32_t wf0 = smem_u32(&sm.w_full[0]), we0 = smem_u32(&sm.w_empty[0]);
      const uint32_t af0 = smem_u32(&sm.acc_full[0]), xc0 = smem_u32(&sm.x_consumed[0]);
      const uint32_t idesc256 = umma_idesc_bf16(TILE_M, 256), idesc128 = umma_idesc_bf16(TILE_M, 128), idesc16 = umma_idesc_bf16(TILE_M, 16);
      NERFB_TRACE(int kbi = 0;)
      for (int64_t tile = blockIdx.x; tile < args.n_tiles; tile += gridDim.x, ++it) {
        const uint32_t b = it & 1;
        const uint32_t ops_base = smem_u32(sm.ops[b]);
        uint4 op = lds128u(ops_base);
        for (int i = 0; i < n_ops; ++i) {
          const uint4 cur = op;
          if (i + 1 < n_ops) op = lds128u(ops_base + 16u * (uint32_t)(i + 1));
          NERFB_TRACE(if (cur.z & OP_FIRST) kbi = 0; if (args.trace && blockIdx.x == 0 && gs < 40) args.trace[gs * 32 + kbi * 4 + 0] = clock64();)
          const uint32_t a_parity = (par >> cur.w) & 1u;
          par ^= 1u << cur.w;
          const uint32_t w_bar = wf0 + 8u * stage;
          // both barriers of this K-block are polled back to back so their latencies overlap
          const bool ok_a = cur.y ? mbar_try_wait_addr(cur.y, a_parity) : true;
          const bool ok_w = mbar_try_wait_addr(w_bar, phase);
          if (!ok_a) mbar_wait_addr(cur.y, a_parity);
          NERFB_TRACE(if (args.trace && blockIdx.x == 0 && gs < 40) args.trace[gs * 32 + kbi * 4 + 1] = clock64();)
          if (!ok_w) mbar_wait_addr(w_bar, phase);
          tc_fence_after();
          NERFB_TRACE(if (args.trace && blockIdx.x == 0 && gs < 40) args.trace[gs * 32 + kbi * 4 + 2] = clock64();)
          umma_bf16_k64(tmem + (gs & 1u) * 256u, cur.x, ring_lo0 + stage * (STAGE_BYTES >> 4), (cur.z & OP_N16) ? idesc16 : (cur.z & OP_N128) ? idesc128 : idesc256,
                        (cur.z & OP_FIRST) ? 0u : 1u);
          umma_commit_addr(we0 + 8u * stage);            // slab free once these MMAs have read it
          NERFB_TRACE(if (args.trace && blockIdx.x == 0 && gs < 40) args.trace[gs * 32 + kbi * 4 + 3] = clock64(); ++kbi;)
          if (++stage == N_STAGES) { stage = 0; phase ^= 1; }
          if (cur.z & OP_LAST) {
            umma_commit_addr(af0 + 8u * (gs & 1u));
            if (cur.z & OP_XDONE) umma_commit_addr(xc0 + 8u * b);
            ++gs;
          }
        }
      }
    }
  } else if (warp == WARP_STORE) {
    // ===== store lane: published blocks -> the tile's record in HBM (training only) =====
    // Up to 3 bulk stores stay in flight: after committing store k the lane waits only until store k-2 has been READ out
    // of shared memory and releases that one's buffer (a_free / x_stored), so the ~1 us store latency is pipelined instead
    // of serialising the 40 stores of a tile.  The release order is the issue order, which is also the order in which the
    // epilogue / encoder need the buffers back, so the lag cannot deadlock (a buffer is needed again a whole layer later).
    if (lane == 0 && saving) {
      uint32_t a_par[4] = {0, 0, 0, 0}, s_par[4] = {0, 0, 0, 0}, xr_par[2] = {0, 0};
      uint32_t it = 0;
      uint32_t pend[2] = {0u, 0u};              // barrier addresses of the two stores still unreleased (0 = none)
      auto issued = [&](uint32_t bar) {         // store k committed: release store k-2
        bulk_commit(); bulk_wait_read2();
        if (pend[0]) mbar_arrive_addr(pend[0]);
        pend[0] = pend[1]; pend[1] = bar;
      };
      for (int64_t tile = blockIdx.x; tile < args.n_tiles; tile += gridDim.x, ++it) {
        const int b = it & 1;
        uint8_t* rec = args.save + tile * (int64_t)pg.tile_save_bytes;
        mbar_wait_relaxed(&sm.x_ready[b], xr_par[b]); xr_par[b] ^= 1;
        bulk_s2g(rec + pg.save_x_off, sm.x[b], BLK_BYTES); issued(smem_u32(&sm.x_stored[b]));
        for (int s = 0; s < pg.n_steps; ++s) {
          const TcStep& st = pg.s[s];
          for (int j = 0; j < st.out_blocks; ++j) {
            if (st.store_only) { mbar_wait_relaxed(&sm.s_ready[j], s_par[j]); s_par[j] ^= 1; }
            else { mbar_wait_relaxed(&sm.a_ready[j], a_par[j]); a_par[j] ^= 1; }
            if (st.save_off >= 0) {
              bulk_s2g(rec + st.save_off + j * BLK_BYTES, sm.a[j], BLK_BYTES); issued(smem_u32(&sm.a_free[j]));
            }
          }
        }
        if (!pg.sigma_only) {
          mbar_wait_relaxed(&sm.x_ready[b], xr_par[b]); xr_par[b] ^= 1;
          bulk_s2g(rec + pg.save_dir_off, sm.x[b], BLK_BYTES); issued(smem_u32(&sm.x_stored[b]));
        }
      }
      bulk_wait_read0();
      if (pend[0]) mbar_arrive_addr(pend[0]);
      if (pend[1]) mbar_arrive_addr(pend[1]);
      bulk_wait_all0();
    }
  } else if (warp >= WARP_ENC0 && warp < WARP_EPI0) {
    // ===== encoder warps: one row per thread, one tile ahead =====
    const int row = (warp - WARP_ENC0) * 32 + lane;
    uint32_t xc_par[2] = {0, 0}, xs_par[2] = {0, 0};
    uint32_t it = 0;
    for (int64_t tile = blockIdx.x; tile < args.n_tiles; tile += gridDim.x, ++it) {
      const int b = it & 1;
      const uint32_t x_base = smem_u32(sm.x[b]);
      const int64_t p = tile * TILE_M + row;
      const int64_t pc = p < args.P ? p : args.P - 1;      // padded rows recompute the last point (never stored)
      float px = 0.f, py = 0.f, pz = 0.f, dirx = 0.f, diry = 0.f, dirz = 0.f;
      const float* emb_src = nullptr;
      if (args.in_mode == NERFB_IN_RAYS) {                 // x = o + d z  (models/rendering.py:224)
        const int64_t r = pc / args.S;
        const float* ray = args.in0 + r * 8;
        const float z = __ldg(args.in1 + pc);
        dirx = __ldg(ray + 3); diry = __ldg(ray + 4); dirz = __ldg(ray + 5);
        px = __fadd_rn(__ldg(ray + 0), __fmul_rn(dirx, z));
        py = __fadd_rn(__ldg(ray + 1), __fmul_rn(diry, z));
        pz = __fadd_rn(__ldg(ray + 2), __fmul_rn(dirz, z));
      } else if (args.in_mode == NERFB_IN_GRID) {          // dense query grid (extract_color_mesh.py:117-125): point
        // p = (iy * N + ix) * N + iz of np.meshgrid(x, y, z) ('xy' indexing), axes = np.linspace in float64 -> float32
        const double* rg = reinterpret_cast<const double*>(args.in0);
        const int64_t N = args.S;
        const int iz = (int)(pc % N), ix = (int)((pc / N) % N), iy = (int)(pc / (N * N));
        const double den = (double)(N > 1 ? N - 1 : 1);
        px = (float)(ix == N - 1 && N > 1 ? rg[1] : rg[0] + (double)ix * ((rg[1] - rg[0]) / den));
        py = (float)(iy == N - 1 && N > 1 ? rg[3] : rg[2] + (double)iy * ((rg[3] - rg[2]) / den));
        pz = (float)(iz == N - 1 && N > 1 ? rg[5] : rg[4] + (double)iz * ((rg[5] - rg[4]) / den));
      } else if (SIREN) {                                  // x = [xyz(3) | dir(3)] rows (SirenField.forward)
        const float* src = args.in0 + pc * (int64_t)(pg.sigma_only ? 3 : 6);
        px = __ldg(src); py = __ldg(src + 1); pz = __ldg(src + 2);
        if (!pg.sigma_only) { dirx = __ldg(src + 3); diry = __ldg(src + 4); dirz = __ldg(src + 5); }
      } else {
        emb_src = args.in0 + pc * (int64_t)(pg.in_xyz + (pg.sigma_only ? 0 : pg.in_dir));
      }
      if (it >= 2) {                                       // the buffer's previous contents: multiplied (and stored)
        mbar_wait_relaxed(&sm.x_consumed[b], xc_par[b]); xc_par[b] ^= 1;
        if (saving) { mbar_wait_relaxed(&sm.x_stored[b], xs_par[b]); xs_par[b] ^= 1; }
      }
      if (SIREN) hilo_row(__fmul_rn(px, pg.xyz_scale), __fmul_rn(py, pg.xyz_scale), __fmul_rn(pz, pg.xyz_scale), x_base, row);
      else if (args.in_mode != NERFB_IN_EMBEDDED) encode_row(px, py, pz, pg.n_freq_xyz, x_base, row);
      else copy_row(emb_src, pg.in_xyz, x_base, row);
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm.x_ready[b]);
      if (!pg.sigma_only) {                                // direction encoding, once the last xyz reader is done
        mbar_wait_relaxed(&sm.x_consumed[b], xc_par[b]); xc_par[b] ^= 1;
        if (saving) { mbar_wait_relaxed(&sm.x_stored[b], xs_par[b]); xs_par[b] ^= 1; }
        if (SIREN) hilo_row(dirx, diry, dirz, x_base, row);
        else if (args.in_mode != NERFB_IN_EMBEDDED) encode_row(dirx, diry, dirz, pg.n_freq_dir, x_base, row);   // grid: dir = 0 (extract_mesh.ipynb cell 4)
        else copy_row(emb_src + pg.in_xyz, pg.in_dir, x_base, row);
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.x_ready[b]);
      }
    }
  } else if (warp >= WARP_EPI0 && warp < WARP_END) {
    // ===== epilogue warps (16): every 64-column block is drained by all of them, 32 rows x 16 columns each, so block 0 --
    // the next layer's first K-block -- is published after one short pass and four warps per SM sub-partition hide the
    // TMEM-load / convert / store latencies of each other =====
    const int ew = warp - WARP_EPI0;         // 0..15
    const int q = warp & 3;                  // TMEM lane quarter this warp may access (warp id mod 4)
    const int c4 = ew >> 2;                  // which 16-column quarter of every 64-column block
    const int row = q * 32 + lane;
    const uint32_t row_off = (uint32_t)row * 128u, row_x = (uint32_t)row & 7u;
    const uint32_t t_lane = tmem + ((uint32_t)(q * 32) << 16) + 16u * (uint32_t)c4;
    uint32_t acc_par = 0, af_par = 0, need_free = 0;       // bit a / bit j
    uint32_t gs = 0;
    for (int64_t tile = blockIdx.x; tile < args.n_tiles; tile += gridDim.x) {
      const int64_t p = tile * TILE_M + row;
      const bool live = p < args.P;
      for (int s = 0; s < pg.n_steps; ++s, ++gs) {
        const uint32_t sw = lds32(smem_u32(sm.step_word) + 4u * (uint32_t)s);
        const uint32_t sb = lds32(smem_u32(sm.step_bias) + 4u * (uint32_t)s);
        const int st_relu = (sw >> 9) & 1, st_out = (sw >> 10) & 7, st_head = (sw >> 13) & 3;
        const int nb = (sw & 511u) >> 6;     // 64-column blocks of this step's accumulator
        const bool st_saved = saving && ((sw >> 16) & 1u);
        const uint32_t ready = ((sw >> 15) & 1u) ? smem_u32(sm.s_ready) : smem_u32(sm.a_ready);
        const uint32_t bias_base = smem_u32(sm.bias) + sb * 4u + 64u * (uint32_t)c4;
        const uint32_t a = gs & 1;
        mbar_wait(&sm.acc_full[a], (acc_par >> a) & 1u); acc_par ^= 1u << a;
        tc_fence_after();
        NERFB_TRACE(const bool tr = args.trace && blockIdx.x == 0 && gs < 40 && lane == 0 && q == 0 && c4 == 0; if (tr) args.trace[gs * 32 + 20] = clock64();)
        float head_acc[3] = {0.f, 0.f, 0.f};
        uint32_t gate_lo = 0u, gate_hi = 0u;       // ReLU gates of this thread's 4 x 16 columns, for the dgrad kernel
        uint32_t v[2][16];
        tmem_ld16(t_lane + a * 256u, v[0]);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (j < nb) {
            const uint32_t col0 = 64u * (uint32_t)j + 16u * (uint32_t)c4;
            float4 bb[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) bb[i] = lds128f(bias_base + 256u * (uint32_t)j + 16u * i);
            tmem_ld_wait();
            NERFB_TRACE(if (tr && j == 0) args.trace[gs * 32 + 21] = clock64();)
            if (j + 1 < nb) tmem_ld16(t_lane + a * 256u + 64u * (uint32_t)(j + 1), v[(j + 1) & 1]);
            NERFB_TRACE(if (tr && j == 0) args.trace[gs * 32 + 26] = clock64();)
            const uint32_t (&vv)[16] = v[j & 1];
            float f[16];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              float2 r0, r1;
              r0 = __fadd2_rn(make_float2(__uint_as_float(vv[4 * i]), __uint_as_float(vv[4 * i + 1])), make_float2(bb[i].x, bb[i].y));
              r1 = __fadd2_rn(make_float2(__uint_as_float(vv[4 * i + 2]), __uint_as_float(vv[4 * i + 3])), make_float2(bb[i].z, bb[i].w));
              if (SIREN) {      // sin(freq (W x + b) + phase) = sin((freq W) x + (freq b + phase)): both folds are done when packing (models/nerf.py:149-151)
                if (SAVE) {       // training: d/dz sin(freq z + phase) = freq cos(.) is the dgrad kernel's multiplicative gate.
                  // Only sign(cos) is saved (one bit per element, same mask record as the ReLU gates): the dgrad rebuilds
                  // |cos| = sqrt(1 - sin^2) from the saved sine image it shares with the wgrad and multiplies by freq.
                  float c0, c1, c2, c3;
                  __sincosf(r0.x, &r0.x, &c0); __sincosf(r0.y, &r0.y, &c1); __sincosf(r1.x, &r1.x, &c2); __sincosf(r1.y, &r1.y, &c3);
                  if (j < 2) {
                    gate_lo = __funnelshift_l(__float_as_uint(c0), gate_lo, 1); gate_lo = __funnelshift_l(__float_as_uint(c1), gate_lo, 1);
                    gate_lo = __funnelshift_l(__float_as_uint(c2), gate_lo, 1); gate_lo = __funnelshift_l(__float_as_uint(c3), gate_lo, 1);
                  } else {
                    gate_hi = __funnelshift_l(__float_as_uint(c0), gate_hi, 1); gate_hi = __funnelshift_l(__float_as_uint(c1), gate_hi, 1);
                    gate_hi = __funnelshift_l(__float_as_uint(c2), gate_hi, 1); gate_hi = __funnelshift_l(__float_as_uint(c3), gate_hi, 1);
                  }
                } else {
                  r0.x = __sinf(r0.x); r0.y = __sinf(r0.y); r1.x = __sinf(r1.x); r1.y = __sinf(r1.y);
                }
              }
              f[4 * i] = r0.x; f[4 * i + 1] = r0.y; f[4 * i + 2] = r1.x; f[4 * i + 3] = r1.y;
            }
            if (st_head == 1) {               // sigma = <relu(h_D), w_sigma> + b on the un-rounded activations (nerf.py:112)
              const uint32_t wb = smem_u32(sm.w_sigma) + col0 * 4u;
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const float4 w = lds128f(wb + 16u * i);
                head_acc[0] = fmaf(SIREN ? f[4 * i] : fmaxf(f[4 * i], 0.f), w.x, head_acc[0]);
                head_acc[0] = fmaf(SIREN ? f[4 * i + 1] : fmaxf(f[4 * i + 1], 0.f), w.y, head_acc[0]);
                head_acc[0] = fmaf(SIREN ? f[4 * i + 2] : fmaxf(f[4 * i + 2], 0.f), w.z, head_acc[0]);
                head_acc[0] = fmaf(SIREN ? f[4 * i + 3] : fmaxf(f[4 * i + 3], 0.f), w.w, head_acc[0]);
              }
            } else if (st_head == 2) {        // rgb pre-activations on the un-rounded direction-layer activations
#pragma unroll
              for (int c = 0; c < 3; ++c) {
                const uint32_t wb = smem_u32(sm.w_rgb) + (uint32_t)(c * RGB_IN) * 4u + col0 * 4u;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                  const float4 w = lds128f(wb + 16u * i);
                  head_acc[c] = fmaf(SIREN ? f[4 * i] : fmaxf(f[4 * i], 0.f), w.x, head_acc[c]);
                  head_acc[c] = fmaf(SIREN ? f[4 * i + 1] : fmaxf(f[4 * i + 1], 0.f), w.y, head_acc[c]);
                  head_acc[c] = fmaf(SIREN ? f[4 * i + 2] : fmaxf(f[4 * i + 2], 0.f), w.z, head_acc[c]);
                  head_acc[c] = fmaf(SIREN ? f[4 * i + 3] : fmaxf(f[4 * i + 3], 0.f), w.w, head_acc[c]);
                }
              }
            }
            NERFB_TRACE(if (tr && j == 0) { asm volatile("" :: "f"(f[0]), "f"(f[15])); args.trace[gs * 32 + 27] = clock64(); })
            if (j < st_out) {
              if (saving && ((need_free >> j) & 1u)) {       // the store lane has read the block's previous contents out
                mbar_wait(&sm.a_free[j], (af_par >> j) & 1u); af_par ^= 1u << j; need_free &= ~(1u << j);
              }
              const uint32_t blk = smem_u32(sm.a[0]) + (uint32_t)j * BLK_BYTES + row_off;
              if (st_relu) {
#pragma unroll
                for (int ch = 0; ch < 2; ++ch)
                  sts128(blk + ((((uint32_t)(2 * c4 + ch)) ^ row_x) << 4), pack_bf16_relu(f[8 * ch + 0], f[8 * ch + 1]),
                         pack_bf16_relu(f[8 * ch + 2], f[8 * ch + 3]), pack_bf16_relu(f[8 * ch + 4], f[8 * ch + 5]),
                         pack_bf16_relu(f[8 * ch + 6], f[8 * ch + 7]));
                if (st_saved) {
                  // gate mask for the dgrad: ONE funnel shift per element pulls the pre-activation's sign bit in
                  // (1 = negative = gated off); element i of block j lands at bit 15 - i of its 16-bit field.
#pragma unroll
                  for (int i = 0; i < 16; ++i) {
                    if (j < 2) gate_lo = __funnelshift_l(__float_as_uint(f[i]), gate_lo, 1);
                    else gate_hi = __funnelshift_l(__float_as_uint(f[i]), gate_hi, 1);
                  }
                }
              } else {
#pragma unroll
                for (int ch = 0; ch < 2; ++ch)
                  sts128(blk + ((((uint32_t)(2 * c4 + ch)) ^ row_x) << 4), pack_bf16(f[8 * ch + 0], f[8 * ch + 1]),
                         pack_bf16(f[8 * ch + 2], f[8 * ch + 3]), pack_bf16(f[8 * ch + 4], f[8 * ch + 5]),
                         pack_bf16(f[8 * ch + 6], f[8 * ch + 7]));
              }
              NERFB_TRACE(if (tr && j == 0) args.trace[gs * 32 + 22] = clock64();)
              fence_proxy_async();
              NERFB_TRACE(if (tr && j == 0) args.trace[gs * 32 + 23] = clock64();)
              if (st_saved) need_free |= 1u << j;
            }
            tc_fence_before();
            __syncwarp();
            if (j < st_out && lane == 0) mbar_arrive_addr(ready + 8u * (uint32_t)j);
            NERFB_TRACE(if (tr && j < 2) args.trace[gs * 32 + 24 + j] = clock64();)
          }
        }
        if (st_saved && (st_relu || (SIREN && SAVE))) {
          const int mo = sm.step_mask[s];
          if (mo >= 0)
            *reinterpret_cast<uint2*>(args.save + tile * (int64_t)pg.tile_save_bytes + mo + c4 * (TILE_M * 8) + row * 8) = make_uint2(gate_lo, gate_hi);   // [c4][row]: a warp writes 256 contiguous bytes
        }
        if (CLS && st_head == 3) {       // class head (models/nerf_cls.py:104): sigmoid of the merged parse.1 o parse.0 logits, 16-column step
          tmem_ld_wait();
          if (c4 == 0 && live) {
#pragma unroll
            for (int k = 0; k < 8; ++k)
              if (k < pg.n_cls) args.out[p * pg.out_c + 4 + k] = 1.f / (1.f + __expf(-(__uint_as_float(v[0][k]) + sm.bias[sb + k])));
          }
          tc_fence_before();
        }
        if (st_head == 1) {       // combine the two groups' partial dot products
          sm.part[c4][0][row] = head_acc[0];
          named_bar_sync(1, EPI_THREADS);
          if (c4 == 0) {
            const float sg = sm.part[0][0][row] + sm.part[1][0][row] + sm.part[2][0][row] + sm.part[3][0][row] + sm.b_heads[0];
            sm.sigma[row] = sg;
            if (pg.sigma_only && args.comp.on)      // weights-only coarse pass of test_time (models/rendering.py:230-231) composited in place
              epilogue_composite(sm, args.comp, args.in0, args.in1, p, args.P, live, q, lane, 0.f, 0.f, 0.f, sg, false);
            else if (pg.sigma_only && live) args.out[p] = (args.in_mode == NERFB_IN_GRID) ? fmaxf(sg, 0.f) : sg;   // grid: np.maximum(sigma, 0), extract_color_mesh.py:141
          }
        } else if (st_head == 2) { // rgb = sigmoid(W_rgb h_dir + b)  (models/nerf.py:119-120)
          sm.part[c4][0][row] = head_acc[0]; sm.part[c4][1][row] = head_acc[1]; sm.part[c4][2][row] = head_acc[2];
          named_bar_sync(1, EPI_THREADS);
          if (c4 == 0 && (live || args.comp.on)) {
            float4 o;
            o.x = 1.f / (1.f + __expf(-(sm.part[0][0][row] + sm.part[1][0][row] + sm.part[2][0][row] + sm.part[3][0][row] + sm.b_heads[1])));
            o.y = 1.f / (1.f + __expf(-(sm.part[0][1][row] + sm.part[1][1][row] + sm.part[2][1][row] + sm.part[3][1][row] + sm.b_heads[2])));
            o.z = 1.f / (1.f + __expf(-(sm.part[0][2][row] + sm.part[1][2][row] + sm.part[2][2][row] + sm.part[3][2][row] + sm.b_heads[3])));
            o.w = sm.sigma[row];
            if (!CLS && args.comp.on) epilogue_composite(sm, args.comp, args.in0, args.in1, p, args.P, live, q, lane, o.x, o.y, o.z, o.w, true);
            else if (!CLS) *reinterpret_cast<float4*>(args.out + p * 4) = o;
            else { float* dst = args.out + p * pg.out_c; dst[0] = o.x; dst[1] = o.y; dst[2] = o.z; dst[3] = o.w; }
          }
        }
      }
      named_bar_sync(2, EPI_THREADS);          // head scratch is reused by the next tile
    }
  }

  // ---- teardown ------------------------------------------------------------------------------------------
  tc_fence_before();
  __syncthreads();
  if (warp == WARP_MMA) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}


// =====================================================================================================================
// Tile-pair forward (round 2, inference instantiations): TWO tiles in flight per SM, one accumulator each.
//
// The single-tile kernel above pipelines a layer's MMAs behind the previous layer's epilogue block by block, so a
// tile-layer costs epilogue + MMA tail (~4200 clk measured, the tensor pipe needing 2048).  Here the event sequence is
//   A.s, B.s, A.s+1, B.s+1, ...   (A, B = the two tiles of the pair, s = layer step)
// and every role walks it: while the 16 epilogue warps drain accumulator A (TMEM columns [0,256)) into A's activation
// blocks, the tensor pipe runs B's layer into columns [256,512) and vice versa.  MMA(slot, s+1) starts only when
// epilogue(slot, s) has drained the whole accumulator, so a thread can own 64 CONTIGUOUS columns of its row (warp =
// lane quarter x 64-column block) and publish ONCE per layer: one fence.proxy.async and one mbarrier arrive instead of
// four (tools/ubench/epi2.cu: 1900 clk per 128 x 256 drain under full-rate MMAs against 2760 for the per-block split).
// Shared memory: activations 2 x 4 x 16 KB; ONE 16 KB X block shared by the skip / direction K-blocks of both tiles
// (step 0's encoding is written straight into the tile's own activation block 0, which is free at that point); weight
// ring = 2 x 32 KB slabs (bulk copies below ~32 KB are op-rate-bound: ~350-500 clk per cp.async.bulk whatever its size,
// tools/ubench/l2_stream.cu), and the bias table is read through L1 (ld.global.nc, 10.5 KB) to make room for it.
// =====================================================================================================================
constexpr int P_STAGES = 2;
constexpr uint32_t OP_SX = 1u << 10, OP_SXDONE = 1u << 11, OP_TILE_LAST = 1u << 12;

struct __align__(1024) SmemP {
  uint8_t a[2][4][BLK_BYTES];            // activation tiles of the two slots
  uint8_t x[BLK_BYTES];                  // shared X block: skip / direction encodings, one use at a time
  union {
    uint8_t ring[P_STAGES][STAGE_BYTES];   // weight slabs (single-CTA pair kernel)
    uint8_t ring2[4][128 * 128];           // this CTA's slab halves (CTA-pair kernel, cta_group::2)
  };
  float w_sigma[256];
  float w_rgb[3 * 256];
  float b_heads[4];
  float part[4][3][TILE_M];              // head partial sums [64-column block][channel][row]
  float sigma[2][TILE_M];                // per slot: lives from the sigma-head step to the rgb step
  float comp[5][TILE_M];
  float comp_tot[4];
  uint64_t w_full[4], w_empty[4];
  uint64_t w_peer[4];                    // CTA-pair kernel, leader: the peer's half of the slab in this stage has landed
  uint64_t acc_full[2];                  // MMAs of the slot's current step have completed
  uint64_t epi_done[2];                  // the slot's accumulator has been drained and its activation blocks published
  uint64_t enc_ready[2];                 // step 0's encoding sits in a[slot][0]
  uint64_t a0_free[2];                   // the slot's last step has read its activation blocks
  uint64_t x_ready, x_free;
  uint32_t tmem_base;
  uint4 ops[2][MAX_OPS];                 // per slot: {A descriptor low word, -, flags, -}
  uint32_t step_word[MAX_STEPS], step_bias[MAX_STEPS];
  uint32_t step_woff[MAX_STEPS];         // byte offset of the step's first slab in the packed blob
  uint32_t step_nkb[MAX_STEPS];          // K-blocks | N << 8
  int use_step[4], use_kind[4], n_uses;  // X uses of a tile in step order: kind 0 = xyz encoding, 1 = direction encoding
  int step_mask[MAX_STEPS];
  uint32_t st_cnt[2], sx_cnt;            // training: events the store lane has finished with per slot / shared-X uses it has seen
};
static_assert(sizeof(SmemP) + 1024 <= 232448, "shared memory budget (227 KB per CTA)");

__device__ __forceinline__ void pack_row64(const float (&buf)[64], uint32_t (&pk)[32]) {
#pragma unroll
  for (int i = 0; i < 32; ++i) pk[i] = pack_bf16(buf[2 * i], buf[2 * i + 1]);
}
__device__ __forceinline__ void store_pk64(uint32_t base, int row, const uint32_t (&pk)[32]) {
#pragma unroll
  for (int ch = 0; ch < 8; ++ch)
    sts128(base + (uint32_t)row * 128u + ((((uint32_t)ch) ^ ((uint32_t)row & 7u)) << 4), pk[4 * ch], pk[4 * ch + 1], pk[4 * ch + 2], pk[4 * ch + 3]);
}
// encode_row / hilo_row / copy_row into registers (same values, same rounding): the row is computed BEFORE the encoder waits
// for its buffer and stored right after, so the hand-off costs eight st.shared.v4 instead of the sincos ladder
__device__ __forceinline__ void encode_pk(float vx, float vy, float vz, int n_freq, uint32_t (&pk)[32]) {
  float s[3], c[3];
  const float v[3] = {vx, vy, vz};
#pragma unroll
  for (int j = 0; j < 3; ++j) sincosf(v[j], &s[j], &c[j]);
  float buf[64];
#pragma unroll
  for (int i = 0; i < 64; ++i) buf[i] = 0.f;
#pragma unroll
  for (int j = 0; j < 3; ++j) buf[j] = v[j];
#pragma unroll
  for (int k = 0; k < 10; ++k) {
    const bool on = k < n_freq;
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      buf[3 + 6 * k + j] = on ? s[j] : 0.f;
      buf[6 + 6 * k + j] = on ? c[j] : 0.f;
      const float s2 = 2.f * s[j] * c[j];
      const float c2 = 1.f - 2.f * s[j] * s[j];
      s[j] = s2; c[j] = c2;
    }
  }
  pack_row64(buf, pk);
}
__device__ __forceinline__ void hilo_pk(float vx, float vy, float vz, uint32_t (&pk)[32]) {
  float buf[64];
#pragma unroll
  for (int i = 0; i < 64; ++i) buf[i] = 0.f;
  const float v[3] = {vx, vy, vz};
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    const float hi = __bfloat162float(__float2bfloat16_rn(v[j]));
    buf[j] = hi; buf[3 + j] = v[j] - hi; buf[6 + j] = hi;
  }
  pack_row64(buf, pk);
}
__device__ __forceinline__ void copy_pk(const float* src, int width, uint32_t (&pk)[32]) {
  float buf[64];
#pragma unroll
  for (int i = 0; i < 64; ++i) buf[i] = (src && i < width) ? __ldg(src + i) : 0.f;
  pack_row64(buf, pk);
}

// SAVE (training): a store lane bulk-stores every published activation tile (one 64 KB copy per event) and the encodings to
// the tile's record for the backward, and counts the events it has finished with per slot (st_cnt) / the shared-X uses
// (sx_cnt); the epilogue and the encoder wait on those monotonic counters before they overwrite a buffer -- a whole event of
// slack, so they practically never block.  ReLU gates / sign(cos) go out as the same [column quarter][row] bit-mask record
// the dgrad kernel reads (mlp_tc_bwd.cu), here as four 16-bit stores per thread and layer.
template <bool SIREN, bool CLS, bool SAVE>
__global__ void __launch_bounds__(NUM_THREADS, 1) mlp_tc_fwd_pair_kernel(const __grid_constant__ TcFwdArgs args) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  SmemP& sm = *reinterpret_cast<SmemP*>(smem_raw);
  if ((smem_u32(smem_raw) & 1023u) != 0u) __trap();
  const TcProgram& pg = args.prog;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t n_pairs = (args.n_tiles + 1) >> 1;
  const bool saving = SAVE ? (args.save != nullptr) : false;

  // ---- one-time setup -----------------------------------------------------------------------------------
  for (int i = threadIdx.x; i < 256; i += NUM_THREADS) sm.w_sigma[i] = __ldg(args.bias + pg.n_bias + i);
  if (threadIdx.x == 0) sm.b_heads[0] = __ldg(args.bias + pg.n_bias + 256);
  constexpr int RGB_IN = SIREN ? 256 : 128;
  for (int i = threadIdx.x; i < 3 * RGB_IN; i += NUM_THREADS) sm.w_rgb[i] = __ldg(args.bias + pg.n_bias + 257 + i);
  if (threadIdx.x < 3) sm.b_heads[1 + threadIdx.x] = __ldg(args.bias + pg.n_bias + 257 + 3 * RGB_IN + threadIdx.x);
  if (threadIdx.x < 2) {                  // decode the MMA op list for slot b (sequential: the step index is needed)
    const int b = threadIdx.x;
    int s = 0; bool sx = false;
    for (int i = 0; i < pg.n_ops; ++i) {
      uint32_t w = pg.ops[i] & ~(OP_SX | OP_SXDONE | OP_TILE_LAST);
      uint4 o;
      if (w & OP_X) {
        if (s == 0) o.x = ((smem_u32(sm.a[b][0]) >> 4) & 0x3FFFu) | (1u << 16);
        else { o.x = ((smem_u32(sm.x) >> 4) & 0x3FFFu) | (1u << 16); w |= OP_SX; sx = true; }
      } else {
        o.x = ((smem_u32(sm.a[b][(w >> 5) & 3u]) >> 4) & 0x3FFFu) | (1u << 16);
      }
      if (w & OP_LAST) {
        if (sx) w |= OP_SXDONE;
        if (s == pg.n_steps - 1) w |= OP_TILE_LAST;
        ++s; sx = false;
      }
      o.y = 0u; o.z = w; o.w = 0u;
      sm.ops[b][i] = o;
    }
  }
  if (threadIdx.x == 32) {                // per-step tables for the producer / epilogue, X uses for the encoder
    uint32_t off = 0; int nu = 0; bool dir_now = false;
    for (int s = 0; s < pg.n_steps; ++s) {
      const TcStep& st = pg.s[s];
      sm.step_word[s] = pg.step_word[s]; sm.step_bias[s] = (uint32_t)st.bias_off;
      sm.step_woff[s] = off; sm.step_nkb[s] = (uint32_t)st.n_kb | ((uint32_t)st.N << 8); sm.step_mask[s] = st.mask_off;
      off += (uint32_t)st.n_kb * (uint32_t)st.N * 128u;
      bool uses_x = false;
      for (int kb = 0; kb < st.n_kb; ++kb) uses_x |= (st.src[kb] == SRC_X);
      if (uses_x && nu < 4) { sm.use_step[nu] = s; sm.use_kind[nu] = dir_now ? 1 : 0; ++nu; }
      if (st.x_done) dir_now = true;      // the xyz encoding's last reader: later X contents are the direction encoding
    }
    sm.n_uses = nu;
  }
  if (threadIdx.x == 0) {
    for (int i = 0; i < P_STAGES; ++i) { mbar_init(&sm.w_full[i], 1); mbar_init(&sm.w_empty[i], 1); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&sm.acc_full[i], 1); mbar_init(&sm.epi_done[i], N_EPI_WARPS);
      mbar_init(&sm.enc_ready[i], 4); mbar_init(&sm.a0_free[i], 1);
    }
    mbar_init(&sm.x_ready, 4); mbar_init(&sm.x_free, 1);
    sm.st_cnt[0] = 0u; sm.st_cnt[1] = 0u; sm.sx_cnt = 0u;
    fence_barrier_init();
  }
  if (warp == WARP_MMA) tmem_alloc(&sm.tmem_base, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp == WARP_PRODUCER) {
    // ===== weight producer: slabs in event order (step s of slot 0, step s of slot 1, step s + 1 ...) =====
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int64_t pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
        for (int s = 0; s < pg.n_steps; ++s) {
          const uint32_t nk = sm.step_nkb[s];
          const uint32_t bytes = (nk >> 8) * 128u;
          const int n_kb = (int)(nk & 255u);
          for (int slot = 0; slot < 2; ++slot) {
            const uint8_t* src = args.wblob + sm.step_woff[s];
            for (int kb = 0; kb < n_kb; ++kb) {
              mbar_wait_relaxed(&sm.w_empty[stage], phase ^ 1);
              mbar_expect_tx(&sm.w_full[stage], bytes);
              bulk_g2s(sm.ring[stage], src, bytes, &sm.w_full[stage]);
              src += bytes;
              if (++stage == P_STAGES) { stage = 0; phase ^= 1; }
            }
          }
        }
      }
    }
  } else if (warp == WARP_MMA) {
    // ===== MMA issuer (single thread) =====
    if (lane == 0) {
      uint32_t stage = 0, phase = 0, ed_par = 0, en_par = 0, xr_par = 0, it = 0;
      NERFB_TRACE(int tev = 0;)
      const int n_ops = pg.n_ops;
      const uint32_t ring_lo0 = ((smem_u32(sm.ring[0]) >> 4) & 0x3FFFu) | (1u << 16);
      const uint32_t wf0 = smem_u32(&sm.w_full[0]), we0 = smem_u32(&sm.w_empty[0]);
      const uint32_t af0 = smem_u32(&sm.acc_full[0]), ed0 = smem_u32(&sm.epi_done[0]), en0 = smem_u32(&sm.enc_ready[0]);
      const uint32_t a0f0 = smem_u32(&sm.a0_free[0]), xr = smem_u32(&sm.x_ready), xf = smem_u32(&sm.x_free);
      const uint32_t idesc256 = umma_idesc_bf16(TILE_M, 256), idesc128 = umma_idesc_bf16(TILE_M, 128), idesc16 = umma_idesc_bf16(TILE_M, 16);
      for (int64_t pair = blockIdx.x; pair < n_pairs; pair += gridDim.x, ++it) {
        int i0 = 0;                           // first op of the current step
        while (i0 < n_ops) {
          int i1 = i0;
          for (uint32_t slot = 0; slot < 2; ++slot) {
            const uint32_t ops_base = smem_u32(sm.ops[slot]);
            const uint32_t d_tmem = tmem + slot * 256u;
            int i = i0;
            NERFB_TRACE(const bool tr = args.trace && blockIdx.x == 0 && tev < 64; long long tw = 0; if (tr) args.trace[tev * 16 + 0] = clock64();)
            for (;;) {
              const uint4 cur = lds128u(ops_base + 16u * (uint32_t)i);
              if (cur.z & OP_FIRST) {         // the slot's previous event has been drained (accumulator free, activations published)
                if (!(it == 0 && i == 0)) { mbar_wait_addr(ed0 + 8u * slot, (ed_par >> slot) & 1u); ed_par ^= 1u << slot; }
                NERFB_TRACE(if (tr) args.trace[tev * 16 + 1] = clock64();)
              }
              if (cur.z & OP_X) {
                if (cur.z & OP_SX) { mbar_wait_addr(xr, xr_par); xr_par ^= 1u; }
                else { mbar_wait_addr(en0 + 8u * slot, (en_par >> slot) & 1u); en_par ^= 1u << slot; }
              }
              const uint32_t idesc = (cur.z & OP_N16) ? idesc16 : (cur.z & OP_N128) ? idesc128 : idesc256;
              NERFB_TRACE(const long long tw0 = clock64();)
              mbar_wait_addr(wf0 + 8u * stage, phase);
              NERFB_TRACE(tw += clock64() - tw0;)
              tc_fence_after();
              umma_bf16_k64(d_tmem, cur.x, ring_lo0 + stage * (STAGE_BYTES >> 4), idesc, (cur.z & OP_FIRST) ? 0u : 1u);
              umma_commit_addr(we0 + 8u * stage);            // slab free once these MMAs have read it
              if (++stage == P_STAGES) { stage = 0; phase ^= 1; }
              ++i;
              if (cur.z & OP_LAST) {
                umma_commit_addr(af0 + 8u * slot);
                if (cur.z & OP_SXDONE) umma_commit_addr(xf);
                if (cur.z & OP_TILE_LAST) umma_commit_addr(a0f0 + 8u * slot);
                NERFB_TRACE(if (tr) { args.trace[tev * 16 + 2] = tw; args.trace[tev * 16 + 3] = clock64(); } ++tev;)
                break;
              }
            }
            i1 = i;
          }
          i0 = i1;
        }
      }
    }
  } else if (warp == WARP_STORE) {
    // ===== store lane (training): published tiles / encodings -> the tile's record in HBM =====
    if (lane == 0 && saving) {
      uint32_t ed_par = 0, en_par = 0, xr_par = 0;
      const uint32_t cnt0 = smem_u32(&sm.st_cnt[0]), cntx = smem_u32(&sm.sx_cnt);
      for (int64_t pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
        int u = 0;                                  // next X use of the tile (same for both slots)
        for (int s = 0; s < pg.n_steps; ++s) {
          const TcStep& st = pg.s[s];
          const bool uses_x = u < sm.n_uses && sm.use_step[u] == s;
          const int kind = uses_x ? sm.use_kind[u] : 0;
          for (uint32_t slot = 0; slot < 2; ++slot) {
            const int64_t tile = 2 * pair + slot;
            const bool live_tile = tile < args.n_tiles;
            uint8_t* rec = args.save + tile * (int64_t)pg.tile_save_bytes;
            if (uses_x && s == 0) {                 // xyz encoding of step 0, in the tile's own block 0
              mbar_wait_relaxed(&sm.enc_ready[slot], (en_par >> slot) & 1u); en_par ^= 1u << slot;
              if (live_tile) { bulk_s2g(rec + pg.save_x_off, sm.a[slot][0], BLK_BYTES); bulk_commit(); bulk_wait_read0(); }
              cnt_publish(cnt0 + 4u * slot);
            } else if (uses_x) {                    // shared X block: the direction encoding is part of the record, the skip copy is not
              mbar_wait_relaxed(&sm.x_ready, xr_par); xr_par ^= 1u;
              if (kind == 1 && pg.save_dir_off >= 0 && live_tile) { bulk_s2g(rec + pg.save_dir_off, sm.x, BLK_BYTES); bulk_commit(); bulk_wait_read0(); }
              cnt_publish(cntx);
            }
            mbar_wait_relaxed(&sm.epi_done[slot], (ed_par >> slot) & 1u); ed_par ^= 1u << slot;
            if (st.save_off >= 0 && st.out_blocks > 0 && live_tile) {
              bulk_s2g(rec + st.save_off, sm.a[slot][0], (uint32_t)st.out_blocks * BLK_BYTES); bulk_commit(); bulk_wait_read0();
            }
            cnt_publish(cnt0 + 4u * slot);
          }
          if (uses_x) ++u;
        }
      }
      bulk_wait_all0();
    }
  } else if (warp >= WARP_ENC0 && warp < WARP_EPI0) {
    // ===== encoder warps: one row per thread; a row is computed into registers, then stored as soon as its buffer is free =====
    const int row = (warp - WARP_ENC0) * 32 + lane;
    uint32_t a0_par = 0, xf_par = 0, it = 0, n_shared = 0;
    bool first_shared = true;
    const int n_uses = sm.n_uses;
    for (int64_t pair = blockIdx.x; pair < n_pairs; pair += gridDim.x, ++it) {
      float pos[2][3], dir[2][3];
      const float* emb[2] = {nullptr, nullptr};
#pragma unroll
      for (int slot = 0; slot < 2; ++slot) {
        const int64_t p = (2 * pair + slot) * TILE_M + row;
        const int64_t pc = p < args.P ? p : args.P - 1;      // padded rows recompute the last point (never stored)
        pos[slot][0] = pos[slot][1] = pos[slot][2] = 0.f; dir[slot][0] = dir[slot][1] = dir[slot][2] = 0.f;
        if (args.in_mode == NERFB_IN_RAYS) {                 // x = o + d z  (models/rendering.py:224)
          const int64_t r = pc / args.S;
          const float* ray = args.in0 + r * 8;
          const float z = __ldg(args.in1 + pc);
          dir[slot][0] = __ldg(ray + 3); dir[slot][1] = __ldg(ray + 4); dir[slot][2] = __ldg(ray + 5);
          pos[slot][0] = __fadd_rn(__ldg(ray + 0), __fmul_rn(dir[slot][0], z));
          pos[slot][1] = __fadd_rn(__ldg(ray + 1), __fmul_rn(dir[slot][1], z));
          pos[slot][2] = __fadd_rn(__ldg(ray + 2), __fmul_rn(dir[slot][2], z));
        } else if (args.in_mode == NERFB_IN_GRID) {          // dense query grid (extract_color_mesh.py:117-125), see the single-tile kernel
          const double* rg = reinterpret_cast<const double*>(args.in0);
          const int64_t N = args.S;
          const int iz = (int)(pc % N), ix = (int)((pc / N) % N), iy = (int)(pc / (N * N));
          const double den = (double)(N > 1 ? N - 1 : 1);
          pos[slot][0] = (float)(ix == N - 1 && N > 1 ? rg[1] : rg[0] + (double)ix * ((rg[1] - rg[0]) / den));
          pos[slot][1] = (float)(iy == N - 1 && N > 1 ? rg[3] : rg[2] + (double)iy * ((rg[3] - rg[2]) / den));
          pos[slot][2] = (float)(iz == N - 1 && N > 1 ? rg[5] : rg[4] + (double)iz * ((rg[5] - rg[4]) / den));
        } else if (SIREN) {                                  // x = [xyz(3) | dir(3)] rows (SirenField.forward)
          const float* src = args.in0 + pc * (int64_t)(pg.sigma_only ? 3 : 6);
          pos[slot][0] = __ldg(src); pos[slot][1] = __ldg(src + 1); pos[slot][2] = __ldg(src + 2);
          if (!pg.sigma_only) { dir[slot][0] = __ldg(src + 3); dir[slot][1] = __ldg(src + 4); dir[slot][2] = __ldg(src + 5); }
        } else {
          emb[slot] = args.in0 + pc * (int64_t)(pg.in_xyz + (pg.sigma_only ? 0 : pg.in_dir));
        }
      }
      for (int u = 0; u < n_uses; ++u) {
        const int us = sm.use_step[u], kind = sm.use_kind[u];
#pragma unroll
        for (int slot = 0; slot < 2; ++slot) {
          uint32_t pk[32];
          if (kind == 0) {
            if (SIREN) hilo_pk(__fmul_rn(pos[slot][0], pg.xyz_scale), __fmul_rn(pos[slot][1], pg.xyz_scale), __fmul_rn(pos[slot][2], pg.xyz_scale), pk);
            else if (args.in_mode != NERFB_IN_EMBEDDED) encode_pk(pos[slot][0], pos[slot][1], pos[slot][2], pg.n_freq_xyz, pk);
            else copy_pk(emb[slot], pg.in_xyz, pk);
          } else {
            if (SIREN) hilo_pk(dir[slot][0], dir[slot][1], dir[slot][2], pk);
            else if (args.in_mode != NERFB_IN_EMBEDDED) encode_pk(dir[slot][0], dir[slot][1], dir[slot][2], pg.n_freq_dir, pk);   // grid: dir = 0
            else copy_pk(emb[slot] + pg.in_xyz, pg.in_dir, pk);
          }
          uint32_t base;
          if (us == 0) {                       // the tile's own block 0: free once the slot's previous tile has been multiplied out
            if (it > 0) { mbar_wait_relaxed(&sm.a0_free[slot], (a0_par >> slot) & 1u); a0_par ^= 1u << slot; }
            if (saving) cnt_wait_relaxed(smem_u32(&sm.st_cnt[slot]), it * (uint32_t)(pg.n_steps + 1));   // the previous tile's stores have been read out
            base = smem_u32(sm.a[slot][0]);
          } else {                             // shared X block: free once the previous use's step has completed
            if (!first_shared) { mbar_wait_relaxed(&sm.x_free, xf_par); xf_par ^= 1u; }
            first_shared = false;
            if (saving) cnt_wait_relaxed(smem_u32(&sm.sx_cnt), n_shared);      // the store lane has seen (and stored) every earlier use
            ++n_shared;
            base = smem_u32(sm.x);
          }
          store_pk64(base, row, pk);
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) mbar_arrive(us == 0 ? &sm.enc_ready[slot] : &sm.x_ready);
        }
      }
    }
  } else if (warp >= WARP_EPI0 && warp < WARP_END) {
    // ===== epilogue warps (16): warp = (TMEM lane quarter q, 64-column block g); a thread drains 64 contiguous columns of
    // its row as four pipelined tcgen05.ld.x16, writes its 128-byte row of activation block g, and the warp publishes once =====
    const int ew = warp - WARP_EPI0;         // 0..15
    const int q = warp & 3;                  // TMEM lane quarter this warp may access (warp id mod 4)
    const int g = ew >> 2;                   // 64-column block
    const int row = q * 32 + lane;
    const uint32_t row_off = (uint32_t)row * 128u, row_x = (uint32_t)row & 7u;
    uint32_t acc_par = 0, it = 0;
    NERFB_TRACE(int tev = 0;)
    for (int64_t pair = blockIdx.x; pair < n_pairs; pair += gridDim.x, ++it) {
      for (int s = 0; s < pg.n_steps; ++s) {
        const uint32_t sw = lds32(smem_u32(sm.step_word) + 4u * (uint32_t)s);
        const uint32_t sb = lds32(smem_u32(sm.step_bias) + 4u * (uint32_t)s);
        const int st_relu = (sw >> 9) & 1, st_out = (sw >> 10) & 7, st_head = (sw >> 13) & 3;
        const int nb = (sw & 511u) >> 6;     // 64-column blocks of this step's accumulator
        const bool st_saved = saving && ((sw >> 16) & 1u);
        const int mo = SAVE ? sm.step_mask[s] : -1;
#pragma unroll 1
        for (int slot = 0; slot < 2; ++slot) {
          const int64_t p = (2 * pair + slot) * TILE_M + row;
          const bool live = p < args.P;
          NERFB_TRACE(const bool tr = args.trace && blockIdx.x == 0 && tev < 64 && ew == 0 && lane == 0; if (tr) args.trace[tev * 16 + 8] = clock64();)
          mbar_wait(&sm.acc_full[slot], (acc_par >> slot) & 1u); acc_par ^= 1u << slot;
          tc_fence_after();
          NERFB_TRACE(if (tr) args.trace[tev * 16 + 9] = clock64();)
          float head_acc[3] = {0.f, 0.f, 0.f};
          // training: the store lane has read out everything this slot's blocks held before (X copy + s earlier events of this tile)
          NERFB_TRACE(const long long tc0 = clock64();)
          if (SAVE && saving && g < st_out) cnt_wait(smem_u32(&sm.st_cnt[slot]), it * (uint32_t)(pg.n_steps + 1) + 1u + (uint32_t)s);
          NERFB_TRACE(if (tr) args.trace[tev * 16 + 11] = clock64() - tc0;)
          const uint32_t t_addr = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)slot * 256u + 64u * (uint32_t)g;
          if (g < nb) {
            const float4* bias4 = reinterpret_cast<const float4*>(args.bias + sb) + 16 * g;      // L1-resident table, warp-uniform address
            const uint32_t blk = smem_u32(sm.a[slot][g]) + row_off;
            uint32_t v[2][16];
            tmem_ld16(t_addr, v[0]);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              float4 bb[4];
#pragma unroll
              for (int i = 0; i < 4; ++i) bb[i] = __ldg(bias4 + 4 * c + i);
              tmem_ld_wait();
              if (c + 1 < 4) tmem_ld16(t_addr + 16u * (uint32_t)(c + 1), v[(c + 1) & 1]);
              const uint32_t (&vv)[16] = v[c & 1];
              float f[16];
              uint32_t gate = 0u;          // this chunk's 16 gate bits (element i at bit 15 - i), for the dgrad kernel
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                float2 r0, r1;
                r0 = __fadd2_rn(make_float2(__uint_as_float(vv[4 * i]), __uint_as_float(vv[4 * i + 1])), make_float2(bb[i].x, bb[i].y));
                r1 = __fadd2_rn(make_float2(__uint_as_float(vv[4 * i + 2]), __uint_as_float(vv[4 * i + 3])), make_float2(bb[i].z, bb[i].w));
                if (SIREN) {      // frequency / phase folded when packing
                  if (SAVE) {       // training: sign(cos) is the saved gate (the dgrad rebuilds |cos| from the sine image)
                    float c0, c1, c2, c3;
                    __sincosf(r0.x, &r0.x, &c0); __sincosf(r0.y, &r0.y, &c1); __sincosf(r1.x, &r1.x, &c2); __sincosf(r1.y, &r1.y, &c3);
                    gate = __funnelshift_l(__float_as_uint(c0), gate, 1); gate = __funnelshift_l(__float_as_uint(c1), gate, 1);
                    gate = __funnelshift_l(__float_as_uint(c2), gate, 1); gate = __funnelshift_l(__float_as_uint(c3), gate, 1);
                  } else { r0.x = __sinf(r0.x); r0.y = __sinf(r0.y); r1.x = __sinf(r1.x); r1.y = __sinf(r1.y); }
                }
                f[4 * i] = r0.x; f[4 * i + 1] = r0.y; f[4 * i + 2] = r1.x; f[4 * i + 3] = r1.y;
              }
              const uint32_t col0 = 64u * (uint32_t)g + 16u * (uint32_t)c;
              if (st_head == 1) {               // sigma = <relu(h_D), w_sigma> + b on the un-rounded activations (nerf.py:112)
                const uint32_t wb = smem_u32(sm.w_sigma) + col0 * 4u;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                  const float4 w = lds128f(wb + 16u * i);
                  head_acc[0] = fmaf(SIREN ? f[4 * i] : fmaxf(f[4 * i], 0.f), w.x, head_acc[0]);
                  head_acc[0] = fmaf(SIREN ? f[4 * i + 1] : fmaxf(f[4 * i + 1], 0.f), w.y, head_acc[0]);
                  head_acc[0] = fmaf(SIREN ? f[4 * i + 2] : fmaxf(f[4 * i + 2], 0.f), w.z, head_acc[0]);
                  head_acc[0] = fmaf(SIREN ? f[4 * i + 3] : fmaxf(f[4 * i + 3], 0.f), w.w, head_acc[0]);
                }
              } else if (st_head == 2) {        // rgb pre-activations on the un-rounded direction-layer activations
#pragma unroll
                for (int ch = 0; ch < 3; ++ch) {
                  const uint32_t wb = smem_u32(sm.w_rgb) + (uint32_t)(ch * RGB_IN) * 4u + col0 * 4u;
#pragma unroll
                  for (int i = 0; i < 4; ++i) {
                    const float4 w = lds128f(wb + 16u * i);
                    head_acc[ch] = fmaf(SIREN ? f[4 * i] : fmaxf(f[4 * i], 0.f), w.x, head_acc[ch]);
                    head_acc[ch] = fmaf(SIREN ? f[4 * i + 1] : fmaxf(f[4 * i + 1], 0.f), w.y, head_acc[ch]);
                    head_acc[ch] = fmaf(SIREN ? f[4 * i + 2] : fmaxf(f[4 * i + 2], 0.f), w.z, head_acc[ch]);
                    head_acc[ch] = fmaf(SIREN ? f[4 * i + 3] : fmaxf(f[4 * i + 3], 0.f), w.w, head_acc[ch]);
                  }
                }
              }
              if (g < st_out) {
                if (st_relu) {
#pragma unroll
                  for (int ch = 0; ch < 2; ++ch)
                    sts128(blk + ((((uint32_t)(2 * c + ch)) ^ row_x) << 4), pack_bf16_relu(f[8 * ch + 0], f[8 * ch + 1]),
                           pack_bf16_relu(f[8 * ch + 2], f[8 * ch + 3]), pack_bf16_relu(f[8 * ch + 4], f[8 * ch + 5]),
                           pack_bf16_relu(f[8 * ch + 6], f[8 * ch + 7]));
                  if (SAVE && st_saved) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) gate = __funnelshift_l(__float_as_uint(f[i]), gate, 1);     // sign bit: 1 = gated off
                  }
                } else {
#pragma unroll
                  for (int ch = 0; ch < 2; ++ch)
                    sts128(blk + ((((uint32_t)(2 * c + ch)) ^ row_x) << 4), pack_bf16(f[8 * ch + 0], f[8 * ch + 1]),
                           pack_bf16(f[8 * ch + 2], f[8 * ch + 3]), pack_bf16(f[8 * ch + 4], f[8 * ch + 5]),
                           pack_bf16(f[8 * ch + 6], f[8 * ch + 7]));
                }
                // bit-mask record [column quarter c][row] = {blocks 0|1, blocks 2|3}, block g's 16 bits in the upper (even g) / lower half
                if (SAVE && st_saved && mo >= 0 && (st_relu || SIREN) && (2 * pair + slot) < args.n_tiles)
                  *reinterpret_cast<uint16_t*>(args.save + (2 * pair + slot) * (int64_t)pg.tile_save_bytes + mo + c * (TILE_M * 8) + row * 8 +
                                               (g >> 1) * 4 + ((g & 1) ? 0 : 2)) = (uint16_t)(gate & 0xFFFFu);
              }
            }
            if (g < st_out) fence_proxy_async();
          }
          if (CLS && st_head == 3 && g == 0) {       // class head (models/nerf_cls.py:104): sigmoid of the merged parse.1 o parse.0 logits, 16-column step
            uint32_t v[16];
            tmem_ld16(t_addr, v);
            tmem_ld_wait();
            if (live) {
#pragma unroll
              for (int k = 0; k < 8; ++k)
                if (k < pg.n_cls) args.out[p * pg.out_c + 4 + k] = 1.f / (1.f + __expf(-(__uint_as_float(v[k]) + __ldg(args.bias + sb + k))));
            }
          }
          tc_fence_before();
          __syncwarp();
          NERFB_TRACE(if (tr) args.trace[tev * 16 + 10] = clock64(); ++tev;)
          if (lane == 0) mbar_arrive(&sm.epi_done[slot]);    // accumulator drained, activation blocks published: the slot's next step may be issued
          if (st_head == 1) {       // combine the four blocks' partial dot products
            sm.part[g][0][row] = head_acc[0];
            named_bar_sync(1, EPI_THREADS);
            if (g == 0) {
              const float sg = sm.part[0][0][row] + sm.part[1][0][row] + sm.part[2][0][row] + sm.part[3][0][row] + sm.b_heads[0];
              sm.sigma[slot][row] = sg;
              if (pg.sigma_only && args.comp.on)      // weights-only coarse pass of test_time (models/rendering.py:230-231) composited in place
                epilogue_composite(sm, args.comp, args.in0, args.in1, p, args.P, live, q, lane, 0.f, 0.f, 0.f, sg, false);
              else if (pg.sigma_only && live) args.out[p] = (args.in_mode == NERFB_IN_GRID) ? fmaxf(sg, 0.f) : sg;   // grid: np.maximum(sigma, 0), extract_color_mesh.py:141
            }
            named_bar_sync(2, EPI_THREADS);          // head scratch is reused by the other slot's event right after
          } else if (st_head == 2) { // rgb = sigmoid(W_rgb h_dir + b)  (models/nerf.py:119-120)
            sm.part[g][0][row] = head_acc[0]; sm.part[g][1][row] = head_acc[1]; sm.part[g][2][row] = head_acc[2];
            named_bar_sync(1, EPI_THREADS);
            if (g == 0 && (live || args.comp.on)) {
              float4 o;
              o.x = 1.f / (1.f + __expf(-(sm.part[0][0][row] + sm.part[1][0][row] + sm.part[2][0][row] + sm.part[3][0][row] + sm.b_heads[1])));
              o.y = 1.f / (1.f + __expf(-(sm.part[0][1][row] + sm.part[1][1][row] + sm.part[2][1][row] + sm.part[3][1][row] + sm.b_heads[2])));
              o.z = 1.f / (1.f + __expf(-(sm.part[0][2][row] + sm.part[1][2][row] + sm.part[2][2][row] + sm.part[3][2][row] + sm.b_heads[3])));
              o.w = sm.sigma[slot][row];
              if (!CLS && args.comp.on) epilogue_composite(sm, args.comp, args.in0, args.in1, p, args.P, live, q, lane, o.x, o.y, o.z, o.w, true);
              else if (!CLS) *reinterpret_cast<float4*>(args.out + p * 4) = o;
              else { float* dst = args.out + p * pg.out_c; dst[0] = o.x; dst[1] = o.y; dst[2] = o.z; dst[3] = o.w; }
            }
            named_bar_sync(2, EPI_THREADS);
          }
        }
      }
    }
  }

  // ---- teardown ------------------------------------------------------------------------------------------
  tc_fence_before();
  __syncthreads();
  if (warp == WARP_MMA) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

// ---- the same schedule on a CTA PAIR (cluster of 2, tcgen05 cta_group::2) -----------------------------------------------------
// One UMMA is 256 x N x 16: rows 0..127 are the leader CTA's tile, rows 128..255 the peer's, and each CTA holds HALF of the
// weight slab (N / 2 rows), so an SM streams 16 KB per K-block instead of 32 KB -- the single-CTA pair kernel is bound by
// that stream (ring bytes / L2 latency).  Four tiles are in flight per cluster (two slots x two CTAs).  Only the leader's
// elected thread issues MMAs; it waits on barriers in ITS shared memory on which the peer's roles arrive remotely
// (mapa + mbarrier.arrive.shared::cluster), and its tcgen05.commit multicasts to the same barrier in both CTAs.
__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, uint32_t rank) {
  uint32_t r; asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank)); return r;
}
// arrive on the barrier at `bar` (a shared::cta address valid in both CTAs) in the LEADER's shared memory
__device__ __forceinline__ void mbar_arrive_leader(uint32_t bar, uint32_t rank) {
  // default semantics (release, CTA scope) as CUTLASS's ClusterBarrier::arrive: an explicit .release.cluster costs a
  // cluster-scope fence per arrival (measured: the CTA-pair kernel 0.82 ms with it)
  if (rank == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
  else asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(mapa_u32(bar, 0)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait_cluster(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
               : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {
  if (mbar_try_wait_cluster(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait_cluster(bar, parity)) {
    if (clock64() - t0 > 4000000000LL) {
      printf("nerf_b200: cluster mbarrier timeout (block %d thread %d bar %u parity %u)\n", (int)blockIdx.x, (int)threadIdx.x, bar, parity);
      const long long t1 = clock64();
      while (clock64() - t1 < 400000000LL) { }
      __trap();
    }
  }
}
__device__ __forceinline__ void umma_commit_mc2(uint32_t bar) {     // arrive on `bar` in both CTAs once all prior UMMAs have completed
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void tmem_alloc2(uint32_t* smem_slot, uint32_t cols) {    // one full warp in EACH CTA of the pair
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc2(uint32_t taddr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
// one 64-wide K-block = four K = 16 UMMAs of 256 x N x 16 on the CTA pair (descriptors as umma_bf16_k64; B = this CTA's N / 2 rows)
__device__ __forceinline__ void umma2_bf16_k64(uint32_t d_tmem, uint32_t a_lo, uint32_t b_lo, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, t;\n\t.reg .b64 a0, b0, a1, b1, a2, b2, a3, b3;\n\t"
      "setp.ne.b32 p, %4, 0;\n\tsetp.eq.b32 t, 0, 0;\n\t"
      "mov.b64 a0, {%1, 0x40004040};\n\tmov.b64 b0, {%2, 0x40004040};\n\t"
      "add.u64 a1, a0, 2;\n\tadd.u64 b1, b0, 2;\n\tadd.u64 a2, a0, 4;\n\tadd.u64 b2, b0, 4;\n\tadd.u64 a3, a0, 6;\n\tadd.u64 b3, b0, 6;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], a0, b0, %3, p;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], a1, b1, %3, t;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], a2, b2, %3, t;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], a3, b3, %3, t;\n\t}"
      ::"r"(d_tmem), "r"(a_lo), "r"(b_lo), "r"(idesc), "r"(accumulate) : "memory");
}

constexpr int P2_STAGES = 4;
constexpr int P2_STAGE_BYTES = 128 * 128;       // this CTA's half (N / 2 rows) of a [256 x 64] slab

template <bool SIREN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS, 1) mlp_tc_fwd_pair2_kernel(const __grid_constant__ TcFwdArgs args) {
  constexpr bool CLS = false;
  const uint32_t rank = cluster_ctarank();
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  SmemP& sm = *reinterpret_cast<SmemP*>(smem_raw);
  if ((smem_u32(smem_raw) & 1023u) != 0u) __trap();
  const TcProgram& pg = args.prog;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t n_quads = (args.n_tiles + 3) >> 2;          // a cluster walks quads of tiles: tile = 4 quad + 2 rank + slot
  const int64_t cl0 = blockIdx.x >> 1, n_cl = gridDim.x >> 1;

  // ---- one-time setup -----------------------------------------------------------------------------------
  for (int i = threadIdx.x; i < 256; i += NUM_THREADS) sm.w_sigma[i] = __ldg(args.bias + pg.n_bias + i);
  if (threadIdx.x == 0) sm.b_heads[0] = __ldg(args.bias + pg.n_bias + 256);
  constexpr int RGB_IN = SIREN ? 256 : 128;
  for (int i = threadIdx.x; i < 3 * RGB_IN; i += NUM_THREADS) sm.w_rgb[i] = __ldg(args.bias + pg.n_bias + 257 + i);
  if (threadIdx.x < 3) sm.b_heads[1 + threadIdx.x] = __ldg(args.bias + pg.n_bias + 257 + 3 * RGB_IN + threadIdx.x);
  if (threadIdx.x < 2) {                  // decode the MMA op list for slot b (sequential: the step index is needed)
    const int b = threadIdx.x;
    int s = 0; bool sx = false;
    for (int i = 0; i < pg.n_ops; ++i) {
      uint32_t w = pg.ops[i] & ~(OP_SX | OP_SXDONE | OP_TILE_LAST);
      uint4 o;
      if (w & OP_X) {
        if (s == 0) o.x = ((smem_u32(sm.a[b][0]) >> 4) & 0x3FFFu) | (1u << 16);
        else { o.x = ((smem_u32(sm.x) >> 4) & 0x3FFFu) | (1u << 16); w |= OP_SX; sx = true; }
      } else {
        o.x = ((smem_u32(sm.a[b][(w >> 5) & 3u]) >> 4) & 0x3FFFu) | (1u << 16);
      }
      if (w & OP_LAST) {
        if (sx) w |= OP_SXDONE;
        if (s == pg.n_steps - 1) w |= OP_TILE_LAST;
        ++s; sx = false;
      }
      o.y = 0u; o.z = w; o.w = 0u;
      sm.ops[b][i] = o;
    }
  }
  if (threadIdx.x == 32) {                // per-step tables for the producer / epilogue, X uses for the encoder
    uint32_t off = 0; int nu = 0; bool dir_now = false;
    for (int s = 0; s < pg.n_steps; ++s) {
      const TcStep& st = pg.s[s];
      sm.step_word[s] = pg.step_word[s]; sm.step_bias[s] = (uint32_t)st.bias_off;
      sm.step_woff[s] = off; sm.step_nkb[s] = (uint32_t)st.n_kb | ((uint32_t)st.N << 8); sm.step_mask[s] = st.mask_off;
      off += (uint32_t)st.n_kb * (uint32_t)st.N * 128u;
      bool uses_x = false;
      for (int kb = 0; kb < st.n_kb; ++kb) uses_x |= (st.src[kb] == SRC_X);
      if (uses_x && nu < 4) { sm.use_step[nu] = s; sm.use_kind[nu] = dir_now ? 1 : 0; ++nu; }
      if (st.x_done) dir_now = true;      // the xyz encoding's last reader: later X contents are the direction encoding
    }
    sm.n_uses = nu;
  }
  if (threadIdx.x == 0) {
    for (int i = 0; i < P2_STAGES; ++i) { mbar_init(&sm.w_full[i], 1); mbar_init(&sm.w_empty[i], 1); mbar_init(&sm.w_peer[i], 1); }
    for (int i = 0; i < 2; ++i) {                         // leader-side barriers collect the arrivals of BOTH CTAs
      mbar_init(&sm.acc_full[i], 1); mbar_init(&sm.epi_done[i], 2 * N_EPI_WARPS);
      mbar_init(&sm.enc_ready[i], 8); mbar_init(&sm.a0_free[i], 1);
    }
    mbar_init(&sm.x_ready, 8); mbar_init(&sm.x_free, 1);
    fence_barrier_init();
  }
  if (warp == WARP_MMA) tmem_alloc2(&sm.tmem_base, 512);
  tc_fence_before();
  cluster_sync_all();                     // barriers of both CTAs initialised before any remote arrive / multicast commit
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp == WARP_PRODUCER) {
    // ===== weight producer: slabs in event order (step s of slot 0, step s of slot 1, step s + 1 ...) =====
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int64_t quad = cl0; quad < n_quads; quad += n_cl) {
        for (int s = 0; s < pg.n_steps; ++s) {
          const uint32_t nk = sm.step_nkb[s];
          const uint32_t bytes = (nk >> 8) * 64u;           // rows [rank N/2, (rank + 1) N/2) of every slab: a contiguous half of the swizzled image
          const int n_kb = (int)(nk & 255u);
          for (int slot = 0; slot < 2; ++slot) {
            const uint8_t* src = args.wblob + sm.step_woff[s] + rank * bytes;
            for (int kb = 0; kb < n_kb; ++kb) {
              mbar_wait_relaxed(&sm.w_empty[stage], phase ^ 1);
              mbar_expect_tx(&sm.w_full[stage], bytes);
              bulk_g2s(sm.ring2[stage], src, bytes, &sm.w_full[stage]);
              src += 2 * bytes;
              if (++stage == P2_STAGES) { stage = 0; phase ^= 1; }
            }
          }
        }
      }
    }
  } else if (warp == WARP_STORE) {
    // ===== peer CTA only: forward "my half of the slab has landed" to the leader's w_peer barrier; lane k owns ring stage k
    // so the remote arrivals (a cross-SM round trip each) of consecutive slabs overlap =====
    if (lane < P2_STAGES && rank == 1) {
      int per_quad = 0;
      for (int s = 0; s < pg.n_steps; ++s) per_quad += 2 * (int)(sm.step_nkb[s] & 255u);
      const int64_t my_quads = cl0 < n_quads ? (n_quads - cl0 + n_cl - 1) / n_cl : 0;
      const int64_t total = my_quads * per_quad;
      const uint32_t remote = mapa_u32(smem_u32(&sm.w_peer[lane]), 0);
      uint32_t phase = 0;
      for (int64_t j = lane; j < total; j += P2_STAGES) {
        mbar_wait(&sm.w_full[lane], phase); phase ^= 1;
        asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
      }
    }
  } else if (warp == WARP_MMA) {
    // ===== MMA issuer (single thread of the LEADER CTA) =====
    if (lane == 0 && rank == 0) {
      uint32_t stage = 0, phase = 0, ed_par = 0, en_par = 0, xr_par = 0, it = 0;
      const int n_ops = pg.n_ops;
      const uint32_t ring_lo0 = ((smem_u32(sm.ring2[0]) >> 4) & 0x3FFFu) | (1u << 16);
      const uint32_t wf0 = smem_u32(&sm.w_full[0]), we0 = smem_u32(&sm.w_empty[0]), wp0 = smem_u32(&sm.w_peer[0]);
      const uint32_t af0 = smem_u32(&sm.acc_full[0]), ed0 = smem_u32(&sm.epi_done[0]), en0 = smem_u32(&sm.enc_ready[0]);
      const uint32_t a0f0 = smem_u32(&sm.a0_free[0]), xr = smem_u32(&sm.x_ready), xf = smem_u32(&sm.x_free);
      const uint32_t idesc256 = umma_idesc_bf16(2 * TILE_M, 256), idesc128 = umma_idesc_bf16(2 * TILE_M, 128), idesc16 = umma_idesc_bf16(2 * TILE_M, 16);
      for (int64_t quad = cl0; quad < n_quads; quad += n_cl, ++it) {
        int i0 = 0;                           // first op of the current step
        while (i0 < n_ops) {
          int i1 = i0;
          for (uint32_t slot = 0; slot < 2; ++slot) {
            const uint32_t ops_base = smem_u32(sm.ops[slot]);
            const uint32_t d_tmem = tmem + slot * 256u;
            int i = i0;
            for (;;) {
              const uint4 cur = lds128u(ops_base + 16u * (uint32_t)i);
              if (cur.z & OP_FIRST) {         // the slot's previous event has been drained (accumulator free, activations published)
                if (!(it == 0 && i == 0)) { mbar_wait_cluster(ed0 + 8u * slot, (ed_par >> slot) & 1u); ed_par ^= 1u << slot; }
              }
              if (cur.z & OP_X) {
                if (cur.z & OP_SX) { mbar_wait_cluster(xr, xr_par); xr_par ^= 1u; }
                else { mbar_wait_cluster(en0 + 8u * slot, (en_par >> slot) & 1u); en_par ^= 1u << slot; }
              }
              const uint32_t idesc = (cur.z & OP_N16) ? idesc16 : (cur.z & OP_N128) ? idesc128 : idesc256;
              mbar_wait_addr(wf0 + 8u * stage, phase);        // my half of the slab ...
              mbar_wait_cluster(wp0 + 8u * stage, phase);     // ... and the peer's
              tc_fence_after();
              umma2_bf16_k64(d_tmem, cur.x, ring_lo0 + stage * (P2_STAGE_BYTES >> 4), idesc, (cur.z & OP_FIRST) ? 0u : 1u);
              umma_commit_mc2(we0 + 8u * stage);              // slab halves free in both CTAs once these MMAs have read them
              if (++stage == P2_STAGES) { stage = 0; phase ^= 1; }
              ++i;
              if (cur.z & OP_LAST) {
                umma_commit_mc2(af0 + 8u * slot);
                if (cur.z & OP_SXDONE) umma_commit_mc2(xf);
                if (cur.z & OP_TILE_LAST) umma_commit_mc2(a0f0 + 8u * slot);
                break;
              }
            }
            i1 = i;
          }
          i0 = i1;
        }
      }
    }
  } else if (warp >= WARP_ENC0 && warp < WARP_EPI0) {
    // ===== encoder warps: one row per thread; a row is computed into registers, then stored as soon as its buffer is free =====
    const int row = (warp - WARP_ENC0) * 32 + lane;
    uint32_t a0_par = 0, xf_par = 0, it = 0;
    bool first_shared = true;
    const int n_uses = sm.n_uses;
    for (int64_t quad = cl0; quad < n_quads; quad += n_cl, ++it) {
      const int64_t pair = 2 * quad + rank;
      float pos[2][3], dir[2][3];
      const float* emb[2] = {nullptr, nullptr};
#pragma unroll
      for (int slot = 0; slot < 2; ++slot) {
        const int64_t p = (2 * pair + slot) * TILE_M + row;
        const int64_t pc = p < args.P ? p : args.P - 1;      // padded rows recompute the last point (never stored)
        pos[slot][0] = pos[slot][1] = pos[slot][2] = 0.f; dir[slot][0] = dir[slot][1] = dir[slot][2] = 0.f;
        if (args.in_mode == NERFB_IN_RAYS) {                 // x = o + d z  (models/rendering.py:224)
          const int64_t r = pc / args.S;
          const float* ray = args.in0 + r * 8;
          const float z = __ldg(args.in1 + pc);
          dir[slot][0] = __ldg(ray + 3); dir[slot][1] = __ldg(ray + 4); dir[slot][2] = __ldg(ray + 5);
          pos[slot][0] = __fadd_rn(__ldg(ray + 0), __fmul_rn(dir[slot][0], z));
          pos[slot][1] = __fadd_rn(__ldg(ray + 1), __fmul_rn(dir[slot][1], z));
          pos[slot][2] = __fadd_rn(__ldg(ray + 2), __fmul_rn(dir[slot][2], z));
        } else if (args.in_mode == NERFB_IN_GRID) {          // dense query grid (extract_color_mesh.py:117-125), see the single-tile kernel
          const double* rg = reinterpret_cast<const double*>(args.in0);
          const int64_t N = args.S;
          const int iz = (int)(pc % N), ix = (int)((pc / N) % N), iy = (int)(pc / (N * N));
          const double den = (double)(N > 1 ? N - 1 : 1);
          pos[slot][0] = (float)(ix == N - 1 && N > 1 ? rg[1] : rg[0] + (double)ix * ((rg[1] - rg[0]) / den));
          pos[slot][1] = (float)(iy == N - 1 && N > 1 ? rg[3] : rg[2] + (double)iy * ((rg[3] - rg[2]) / den));
          pos[slot][2] = (float)(iz == N - 1 && N > 1 ? rg[5] : rg[4] + (double)iz * ((rg[5] - rg[4]) / den));
        } else if (SIREN) {                                  // x = [xyz(3) | dir(3)] rows (SirenField.forward)
          const float* src = args.in0 + pc * (int64_t)(pg.sigma_only ? 3 : 6);
          pos[slot][0] = __ldg(src); pos[slot][1] = __ldg(src + 1); pos[slot][2] = __ldg(src + 2);
          if (!pg.sigma_only) { dir[slot][0] = __ldg(src + 3); dir[slot][1] = __ldg(src + 4); dir[slot][2] = __ldg(src + 5); }
        } else {
          emb[slot] = args.in0 + pc * (int64_t)(pg.in_xyz + (pg.sigma_only ? 0 : pg.in_dir));
        }
      }
      for (int u = 0; u < n_uses; ++u) {
        const int us = sm.use_step[u], kind = sm.use_kind[u];
#pragma unroll
        for (int slot = 0; slot < 2; ++slot) {
          uint32_t pk[32];
          if (kind == 0) {
            if (SIREN) hilo_pk(__fmul_rn(pos[slot][0], pg.xyz_scale), __fmul_rn(pos[slot][1], pg.xyz_scale), __fmul_rn(pos[slot][2], pg.xyz_scale), pk);
            else if (args.in_mode != NERFB_IN_EMBEDDED) encode_pk(pos[slot][0], pos[slot][1], pos[slot][2], pg.n_freq_xyz, pk);
            else copy_pk(emb[slot], pg.in_xyz, pk);
          } else {
            if (SIREN) hilo_pk(dir[slot][0], dir[slot][1], dir[slot][2], pk);
            else if (args.in_mode != NERFB_IN_EMBEDDED) encode_pk(dir[slot][0], dir[slot][1], dir[slot][2], pg.n_freq_dir, pk);   // grid: dir = 0
            else copy_pk(emb[slot] + pg.in_xyz, pg.in_dir, pk);
          }
          uint32_t base;
          if (us == 0) {                       // the tile's own block 0: free once the slot's previous tile has been multiplied out
            if (it > 0) { mbar_wait_relaxed(&sm.a0_free[slot], (a0_par >> slot) & 1u); a0_par ^= 1u << slot; }
            base = smem_u32(sm.a[slot][0]);
          } else {                             // shared X block: free once the previous use's step has completed
            if (!first_shared) { mbar_wait_relaxed(&sm.x_free, xf_par); xf_par ^= 1u; }
            first_shared = false;
            base = smem_u32(sm.x);
          }
          store_pk64(base, row, pk);
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) mbar_arrive_leader(smem_u32(us == 0 ? &sm.enc_ready[slot] : &sm.x_ready), rank);
        }
      }
    }
  } else if (warp >= WARP_EPI0 && warp < WARP_END) {
    // ===== epilogue warps (16): warp = (TMEM lane quarter q, 64-column block g); a thread drains 64 contiguous columns of
    // its row as four pipelined tcgen05.ld.x16, writes its 128-byte row of activation block g, and the warp publishes once =====
    const int ew = warp - WARP_EPI0;         // 0..15
    const int q = warp & 3;                  // TMEM lane quarter this warp may access (warp id mod 4)
    const int g = ew >> 2;                   // 64-column block
    const int row = q * 32 + lane;
    const uint32_t row_off = (uint32_t)row * 128u, row_x = (uint32_t)row & 7u;
    uint32_t acc_par = 0;
    for (int64_t quad = cl0; quad < n_quads; quad += n_cl) {
      const int64_t pair = 2 * quad + rank;
      for (int s = 0; s < pg.n_steps; ++s) {
        const uint32_t sw = lds32(smem_u32(sm.step_word) + 4u * (uint32_t)s);
        const uint32_t sb = lds32(smem_u32(sm.step_bias) + 4u * (uint32_t)s);
        const int st_relu = (sw >> 9) & 1, st_out = (sw >> 10) & 7, st_head = (sw >> 13) & 3;
        const int nb = (sw & 511u) >> 6;     // 64-column blocks of this step's accumulator
#pragma unroll 1
        for (int slot = 0; slot < 2; ++slot) {
          const int64_t p = (2 * pair + slot) * TILE_M + row;
          const bool live = p < args.P;
          mbar_wait(&sm.acc_full[slot], (acc_par >> slot) & 1u); acc_par ^= 1u << slot;
          tc_fence_after();
          float head_acc[3] = {0.f, 0.f, 0.f};
          const uint32_t t_addr = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)slot * 256u + 64u * (uint32_t)g;
          if (g < nb) {
            const float4* bias4 = reinterpret_cast<const float4*>(args.bias + sb) + 16 * g;      // L1-resident table, warp-uniform address
            const uint32_t blk = smem_u32(sm.a[slot][g]) + row_off;
            uint32_t v[2][16];
            tmem_ld16(t_addr, v[0]);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              float4 bb[4];
#pragma unroll
              for (int i = 0; i < 4; ++i) bb[i] = __ldg(bias4 + 4 * c + i);
              tmem_ld_wait();
              if (c + 1 < 4) tmem_ld16(t_addr + 16u * (uint32_t)(c + 1), v[(c + 1) & 1]);
              const uint32_t (&vv)[16] = v[c & 1];
              float f[16];
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                float2 r0, r1;
                r0 = __fadd2_rn(make_float2(__uint_as_float(vv[4 * i]), __uint_as_float(vv[4 * i + 1])), make_float2(bb[i].x, bb[i].y));
                r1 = __fadd2_rn(make_float2(__uint_as_float(vv[4 * i + 2]), __uint_as_float(vv[4 * i + 3])), make_float2(bb[i].z, bb[i].w));
                if (SIREN) { r0.x = __sinf(r0.x); r0.y = __sinf(r0.y); r1.x = __sinf(r1.x); r1.y = __sinf(r1.y); }   // frequency / phase folded when packing
                f[4 * i] = r0.x; f[4 * i + 1] = r0.y; f[4 * i + 2] = r1.x; f[4 * i + 3] = r1.y;
              }
              const uint32_t col0 = 64u * (uint32_t)g + 16u * (uint32_t)c;
              if (st_head == 1) {               // sigma = <relu(h_D), w_sigma> + b on the un-rounded activations (nerf.py:112)
                const uint32_t wb = smem_u32(sm.w_sigma) + col0 * 4u;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                  const float4 w = lds128f(wb + 16u * i);
                  head_acc[0] = fmaf(SIREN ? f[4 * i] : fmaxf(f[4 * i], 0.f), w.x, head_acc[0]);
                  head_acc[0] = fmaf(SIREN ? f[4 * i + 1] : fmaxf(f[4 * i + 1], 0.f), w.y, head_acc[0]);
                  head_acc[0] = fmaf(SIREN ? f[4 * i + 2] : fmaxf(f[4 * i + 2], 0.f), w.z, head_acc[0]);
                  head_acc[0] = fmaf(SIREN ? f[4 * i + 3] : fmaxf(f[4 * i + 3], 0.f), w.w, head_acc[0]);
                }
              } else if (st_head == 2) {        // rgb pre-activations on the un-rounded direction-layer activations
#pragma unroll
                for (int ch = 0; ch < 3; ++ch) {
                  const uint32_t wb = smem_u32(sm.w_rgb) + (uint32_t)(ch * RGB_IN) * 4u + col0 * 4u;
#pragma unroll
                  for (int i = 0; i < 4; ++i) {
                    const float4 w = lds128f(wb + 16u * i);
                    head_acc[ch] = fmaf(SIREN ? f[4 * i] : fmaxf(f[4 * i], 0.f), w.x, head_acc[ch]);
                    head_acc[ch] = fmaf(SIREN ? f[4 * i + 1] : fmaxf(f[4 * i + 1], 0.f), w.y, head_acc[ch]);
                    head_acc[ch] = fmaf(SIREN ? f[4 * i + 2] : fmaxf(f[4 * i + 2], 0.f), w.z, head_acc[ch]);
                    head_acc[ch] = fmaf(SIREN ? f[4 * i + 3] : fmaxf(f[4 * i + 3], 0.f), w.w, head_acc[ch]);
                  }
                }
              }
              if (g < st_out) {
                if (st_relu) {
#pragma unroll
                  for (int ch = 0; ch < 2; ++ch)
                    sts128(blk + ((((uint32_t)(2 * c + ch)) ^ row_x) << 4), pack_bf16_relu(f[8 * ch + 0], f[8 * ch + 1]),
                           pack_bf16_relu(f[8 * ch + 2], f[8 * ch + 3]), pack_bf16_relu(f[8 * ch + 4], f[8 * ch + 5]),
                           pack_bf16_relu(f[8 * ch + 6], f[8 * ch + 7]));
                } else {
#pragma unroll
                  for (int ch = 0; ch < 2; ++ch)
                    sts128(blk + ((((uint32_t)(2 * c + ch)) ^ row_x) << 4), pack_bf16(f[8 * ch + 0], f[8 * ch + 1]),
                           pack_bf16(f[8 * ch + 2], f[8 * ch + 3]), pack_bf16(f[8 * ch + 4], f[8 * ch + 5]),
                           pack_bf16(f[8 * ch + 6], f[8 * ch + 7]));
                }
              }
            }
            if (g < st_out) fence_proxy_async();
          }
          if (CLS && st_head == 3 && g == 0) {       // class head (models/nerf_cls.py:104): sigmoid of the merged parse.1 o parse.0 logits, 16-column step
            uint32_t v[16];
            tmem_ld16(t_addr, v);
            tmem_ld_wait();
            if (live) {
#pragma unroll
              for (int k = 0; k < 8; ++k)
                if (k < pg.n_cls) args.out[p * pg.out_c + 4 + k] = 1.f / (1.f + __expf(-(__uint_as_float(v[k]) + __ldg(args.bias + sb + k))));
            }
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_leader(smem_u32(&sm.epi_done[slot]), rank);    // accumulator drained, activation blocks published: the slot's next step may be issued
          if (st_head == 1) {       // combine the four blocks' partial dot products
            sm.part[g][0][row] = head_acc[0];
            named_bar_sync(1, EPI_THREADS);
            if (g == 0) {
              const float sg = sm.part[0][0][row] + sm.part[1][0][row] + sm.part[2][0][row] + sm.part[3][0][row] + sm.b_heads[0];
              sm.sigma[slot][row] = sg;
              if (pg.sigma_only && args.comp.on)      // weights-only coarse pass of test_time (models/rendering.py:230-231) composited in place
                epilogue_composite(sm, args.comp, args.in0, args.in1, p, args.P, live, q, lane, 0.f, 0.f, 0.f, sg, false);
              else if (pg.sigma_only && live) args.out[p] = (args.in_mode == NERFB_IN_GRID) ? fmaxf(sg, 0.f) : sg;   // grid: np.maximum(sigma, 0), extract_color_mesh.py:141
            }
            named_bar_sync(2, EPI_THREADS);          // head scratch is reused by the other slot's event right after
          } else if (st_head == 2) { // rgb = sigmoid(W_rgb h_dir + b)  (models/nerf.py:119-120)
            sm.part[g][0][row] = head_acc[0]; sm.part[g][1][row] = head_acc[1]; sm.part[g][2][row] = head_acc[2];
            named_bar_sync(1, EPI_THREADS);
            if (g == 0 && (live || args.comp.on)) {
              float4 o;
              o.x = 1.f / (1.f + __expf(-(sm.part[0][0][row] + sm.part[1][0][row] + sm.part[2][0][row] + sm.part[3][0][row] + sm.b_heads[1])));
              o.y = 1.f / (1.f + __expf(-(sm.part[0][1][row] + sm.part[1][1][row] + sm.part[2][1][row] + sm.part[3][1][row] + sm.b_heads[2])));
              o.z = 1.f / (1.f + __expf(-(sm.part[0][2][row] + sm.part[1][2][row] + sm.part[2][2][row] + sm.part[3][2][row] + sm.b_heads[3])));
              o.w = sm.sigma[slot][row];
              if (!CLS && args.comp.on) epilogue_composite(sm, args.comp, args.in0, args.in1, p, args.P, live, q, lane, o.x, o.y, o.z, o.w, true);
              else if (!CLS) *reinterpret_cast<float4*>(args.out + p * 4) = o;
              else { float* dst = args.out + p * pg.out_c; dst[0] = o.x; dst[1] = o.y; dst[2] = o.z; dst[3] = o.w; }
            }
            named_bar_sync(2, EPI_THREADS);
          }
        }
      }
    }
  }

  // ---- teardown ------------------------------------------------------------------------------------------
  tc_fence_before();
  cluster_sync_all();                     // no CTA leaves while its partner may still arrive on its barriers or read its shared memory
  if (warp == WARP_MMA) { tc_fence_after(); tmem_dealloc2(tmem, 512); }
}


// ---- weight packing: fp32 nn.Linear weights -> bf16 slabs in the swizzled shared-memory image, schedule order ----
struct PackSlab { int64_t w_off; int ld; int col0; int ncols; int N; int64_t dst_off; int hilo; int aux; int film; };   // aux: read the merged class-head matrix, not params   // W[n*ld + col0 + kk], kk < ncols
// hilo: 3 source columns expanded to 9 K-columns [W_hi W_hi W_lo] (pairs with hilo_row's [x_hi x_lo x_hi])
// film >= 0 (SIREN): row n is scaled by the FiLM frequency of that layer, sin(freq (W x + b) + phase) = sin((freq W) x + (freq b + phase)),
// so the epilogue only adds the folded bias; bf16(freq W) carries the same relative rounding as bf16(W)
struct PackArgs { int n_slabs; PackSlab slab[48]; int sw64; };   // sw64: tile-pair kernel's image, two [N x 32] halves per slab, SWIZZLE_64B

__global__ void __launch_bounds__(256) pack_weights_kernel(const float* __restrict__ params_in, PackArgs pa, uint8_t* __restrict__ blob,
                                                           const float* __restrict__ aux, const float* __restrict__ film) {
  // four blocks per slab (a model has ~36 slabs: one block each left three quarters of the SMs idle and the kernel latency-bound)
  const PackSlab sl = pa.slab[blockIdx.x >> 2];
  const float* params = sl.aux ? aux : params_in;
  for (int idx = (blockIdx.x & 3) * blockDim.x + threadIdx.x; idx < sl.N * 8; idx += 4 * blockDim.x) {   // one 16-byte chunk per iteration
    const int n = idx >> 3, ch = idx & 7;
    const float fq = (sl.film >= 0 && film) ? __ldg(film + sl.film * sl.N + n) : 1.f;
    float f[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int kk = ch * 8 + e;
      if (sl.hilo) {
        f[e] = 0.f;
        if (kk < 9) {
          const float w = __fmul_rn(fq, __ldg(params + sl.w_off + (int64_t)n * sl.ld + sl.col0 + kk % 3));
          const float hi = __bfloat162float(__float2bfloat16_rn(w));
          f[e] = kk < 6 ? hi : w - hi;
        }
      } else
      f[e] = (kk < sl.ncols) ? __fmul_rn(fq, __ldg(params + sl.w_off + (int64_t)n * sl.ld + sl.col0 + kk)) : 0.f;
    }
    uint4 o;
    o.x = pack_bf16(f[0], f[1]); o.y = pack_bf16(f[2], f[3]); o.z = pack_bf16(f[4], f[5]); o.w = pack_bf16(f[6], f[7]);
    if (pa.sw64)      // half (ch >> 2) of the slab, 64-byte rows, 16-byte chunk XOR ((row >> 1) & 3)  (Swizzle<2,4,3>)
      *reinterpret_cast<uint4*>(blob + sl.dst_off + (int64_t)(ch >> 2) * sl.N * 64 + (int64_t)n * 64 + (((ch & 3) ^ ((n >> 1) & 3)) << 4)) = o;
    else
    *reinterpret_cast<uint4*>(blob + sl.dst_off + (int64_t)n * 128 + ((ch ^ (n & 7)) << 4)) = o;
  }
}

// bias table: [step biases ...] | w_sigma | b_sigma | w_rgb | b_rgb, gathered from the flat parameter blob
struct BiasSeg { int64_t src; int n; int dst; int film; int aux; };     // film >= 0: FiLM layer whose freq / phase fold into this bias
struct BiasArgs { int n_segs; BiasSeg seg[20]; int film_stride; };
// film = [freq (L x W) | phase (L x W)]: table = freq * b + phase, scale = freq  (models/nerf.py:149-151)
__global__ void __launch_bounds__(256) gather_bias_kernel(const float* __restrict__ params, BiasArgs ba, float* __restrict__ table,
                                                          const float* __restrict__ film, float* __restrict__ scale,
                                                          const float* __restrict__ aux) {
  const BiasSeg sg = ba.seg[blockIdx.x];
  for (int i = threadIdx.x; i < sg.n; i += blockDim.x) {
    const float b = __ldg((sg.aux ? aux : params) + sg.src + i);
    if (sg.film >= 0 && film) {
      const float fr = __ldg(film + sg.film * sg.n + i), ph = __ldg(film + ba.film_stride + sg.film * sg.n + i);
      table[sg.dst + i] = fmaf(fr, b, ph);
      scale[sg.dst + i] = fr;
    } else {
      table[sg.dst + i] = b;
    }
  }
}

// NeRF_3D class head (models/nerf_cls.py:56-61,104): parse = Linear(256,512) -> Linear(512,CLS) -> Sigmoid with NO activation
// between the two Linears, so logits = W_eff feat + b_eff with W_eff = W_p1 W_p0 (CLS x 256), b_eff = W_p1 b_p0 + b_p1.
// One block per output row k (16 rows, rows >= CLS zero); weff fp32 [16 x 256], beff [16] -> the bias table slot.
__global__ void __launch_bounds__(256) merge_parse_kernel(const float* __restrict__ params, int64_t w_p0, int64_t b_p0, int64_t w_p1,
                                                          int64_t b_p1, int n_cls, float* __restrict__ weff, float* __restrict__ beff) {
  const int k = blockIdx.x, i = threadIdx.x;
  float acc = 0.f, bacc = 0.f;
  if (k < n_cls) {
    for (int j = 0; j < 512; ++j) {
      const float w1 = __ldg(params + w_p1 + (int64_t)k * 512 + j);
      acc = fmaf(w1, __ldg(params + w_p0 + (int64_t)j * 256 + i), acc);
      if (i == 0) bacc = fmaf(w1, __ldg(params + b_p0 + j), bacc);
    }
    if (i == 0) bacc += __ldg(params + b_p1 + k);
  }
  weff[k * 256 + i] = acc;
  if (i == 0) beff[k] = bacc;
}

// ---- host side: program construction, workspace plan, launch ---------------------------------------------------
struct TcPlan {
  TcProgram prog;
  PackArgs pack;
  BiasArgs bias;
  int64_t wblob_bytes, bias_floats;
  int64_t off_wblob, off_bias, off_scale, off_weff, off_save, total;
  int cls_bias_off;                // slot of b_eff in the bias table (-1: no class head)
};

bool tc_supported(const nerfb_mlp_desc& d, int in_mode) {
  (void)in_mode;
  if (d.kind == NERFB_KIND_SIREN) return d.W == 256 && d.D >= 1 && d.D <= 9 && d.box > 0.f;   // backward: see tc_bwd_supported
  if (d.kind != NERFB_KIND_NERF || d.W != 256 || d.n_cls < 0 || d.n_cls > 8) return false;
  if (d.D < 1 || d.D > 9) return false;
  if (!d.logscale && d.in_xyz <= 0) return false;                 // in-kernel encoding is the power-of-two ladder
  if (nerfb_in_xyz(d) > 64 || nerfb_in_dir(d) > 32) return false;
  if (nerfb_in_xyz(d) < 1) return false;
  return true;
}

TcFwdLayout tc_fwd_layout(const nerfb_mlp_desc& d, int64_t P);
static TcPlan tc_plan(const nerfb_mlp_desc& d, int64_t P, bool sigma_only, bool save, bool pair = false) {
  TcPlan pl{};
  const NerfOffsets o = nerf_offsets(d);
  TcProgram& pg = pl.prog;
  pg.n_freq_xyz = d.n_freq_xyz; pg.n_freq_dir = d.n_freq_dir; pg.in_xyz = o.in_xyz; pg.in_dir = o.in_dir;
  pg.sigma_only = sigma_only; pg.out_c = sigma_only ? 1 : 4 + d.n_cls; pg.n_cls = sigma_only ? 0 : d.n_cls;
  pl.cls_bias_off = -1;
  int64_t wb = 0; int nb = 0, ns = 0, nslab = 0, nseg = 0, sv = 0;
  int slab_film = -1;            // SIREN: FiLM layer whose frequencies scale the slabs being added
  auto add_slab = [&](int64_t w_off, int ld, int col0, int ncols, int N, int hilo = 0, int aux = 0) {
    pl.pack.slab[nslab++] = PackSlab{w_off, ld, col0, ncols, N, wb, hilo, aux, slab_film};
    wb += (int64_t)N * 128;
  };
  auto add_bias = [&](int64_t src, int n, int film = -1, int aux = 0) { pl.bias.seg[nseg++] = BiasSeg{src, n, nb, film, aux}; nb += n; return nb - n; };
  const bool siren = d.kind == NERFB_KIND_SIREN;
  pg.siren = siren; pg.rgb_in = siren ? 256 : 128; pg.xyz_scale = siren ? 2.0f / d.box : 1.0f;
  pl.bias.film_stride = (d.D + 1) * d.W;
  pg.save_x_off = -1; pg.save_dir_off = -1;
  if (save) { pg.save_x_off = sv; sv += BLK_BYTES; }
  int last_x_reader = 0;
  if (siren) {      // models/nerf.py:204-214: 8 FiLM layers, sigma head on the last, colour FiLM layer on [dir | h] + rgb head
    for (int i = 0; i < d.D; ++i) {
      TcStep& st = pg.s[ns++];
      st = TcStep{};
      st.N = 256; st.relu = 0; st.out_blocks = 4; st.head = (i == d.D - 1) ? 1 : 0; st.film = i; st.save_off = -1; st.mask_off = -1;
      slab_film = i;
      if (save) { st.save_off = sv; sv += 4 * BLK_BYTES; st.mask_off = sv; sv += TILE_M * 32; }      // activations (sine image) | sign(cos) bit mask
      if (i == 0) { st.src[st.n_kb++] = SRC_X; st.wait_x = 1; add_slab(o.w[0], 3, 0, 3, 256, 1); }
      else for (int j = 0; j < 4; ++j) { st.src[st.n_kb++] = j; add_slab(o.w[i], 256, 64 * j, 64, 256); }
      st.bias_off = add_bias(o.b[i], 256, i);
      if (sigma_only && i == d.D - 1) st.out_blocks = 0;
    }
    if (!sigma_only) {
      TcStep& cl = pg.s[ns++]; cl = TcStep{};
      cl.N = 256; cl.relu = 0; cl.head = 2; cl.out_blocks = save ? 4 : 0; cl.store_only = 1; cl.film = d.D; cl.save_off = -1; cl.mask_off = -1;
      slab_film = d.D;
      if (save) { cl.save_off = sv; sv += 4 * BLK_BYTES; cl.mask_off = sv; sv += TILE_M * 32; pg.save_dir_off = sv; sv += BLK_BYTES; }
      for (int j = 0; j < 4; ++j) { cl.src[cl.n_kb++] = j; add_slab(o.w_dir, 259, 3 + 64 * j, 64, 256); }
      cl.src[cl.n_kb++] = SRC_X; cl.wait_x = 1; cl.x_done = 1; add_slab(o.w_dir, 259, 0, 3, 256, 1);
      cl.bias_off = add_bias(o.b_dir, 256, d.D);
    }
  } else
  for (int i = 0; i < d.D; ++i) {
    TcStep& st = pg.s[ns++];
    st = TcStep{};
    const bool skip = i > 0 && ((d.skip_mask >> i) & 1);
    st.N = 256; st.relu = 1; st.out_blocks = 4; st.head = (i == d.D - 1) ? 1 : 0;
    if (i == 0) { st.src[st.n_kb++] = SRC_X; st.wait_x = 1; add_slab(o.w[0], o.in[0], 0, o.in_xyz, 256); }
    else {
      // skip layer: the tile-pair kernel takes the X K-block LAST (the shared X block is handed over while the other four run)
      if (skip && !pair) { st.src[st.n_kb++] = SRC_X; add_slab(o.w[i], o.in[i], 0, o.in_xyz, 256); last_x_reader = i; }
      const int c0 = skip ? o.in_xyz : 0;
      for (int j = 0; j < 4; ++j) { st.src[st.n_kb++] = j; add_slab(o.w[i], o.in[i], c0 + 64 * j, 64, 256); }
      if (skip && pair) { st.src[st.n_kb++] = SRC_X; add_slab(o.w[i], o.in[i], 0, o.in_xyz, 256); last_x_reader = i; }
    }
    st.bias_off = add_bias(o.b[i], 256);
    st.save_off = -1; st.mask_off = -1;
    if (save) { st.save_off = sv; sv += 4 * BLK_BYTES; st.mask_off = sv; sv += TILE_M * 32; }
    if (sigma_only && i == d.D - 1) { st.out_blocks = 0; st.save_off = -1; st.mask_off = -1; }
  }
  pg.s[last_x_reader].x_done = 1;
  if (!sigma_only && !siren) {
    TcStep& fin = pg.s[ns++]; fin = TcStep{};
    fin.N = 256; fin.relu = 0; fin.out_blocks = 4;
    for (int j = 0; j < 4; ++j) { fin.src[fin.n_kb++] = j; add_slab(o.w_final, 256, 64 * j, 64, 256); }
    fin.bias_off = add_bias(o.b_final, 256);
    fin.save_off = -1; fin.mask_off = -1;
    if (save) { fin.save_off = sv; sv += 4 * BLK_BYTES; }
    if (d.n_cls > 0) {      // class-head step: 16 output columns, A = the un-activated xyz_encoding_final blocks
      TcStep& pc = pg.s[ns++]; pc = TcStep{};
      pc.N = 16; pc.relu = 0; pc.out_blocks = 0; pc.head = 3; pc.save_off = -1; pc.mask_off = -1;
      for (int j = 0; j < 4; ++j) { pc.src[pc.n_kb++] = j; add_slab(0, 256, 64 * j, 64, 16, 0, 1); }
      pc.bias_off = add_bias(16 * 256, 16, -1, 1);                         // b_eff, written by merge_parse_kernel behind W_eff
      pl.cls_bias_off = pc.bias_off;
    }
    TcStep& dr = pg.s[ns++]; dr = TcStep{};
    dr.N = 128; dr.relu = 1; dr.head = 2; dr.out_blocks = save ? 2 : 0; dr.store_only = 1; dr.reuse_a = d.n_cls > 0;
    for (int j = 0; j < 4; ++j) { dr.src[dr.n_kb++] = j; add_slab(o.w_dir, 256 + o.in_dir, 64 * j, 64, 128); }
    dr.src[dr.n_kb++] = SRC_X; dr.wait_x = 1; dr.x_done = 1; add_slab(o.w_dir, 256 + o.in_dir, 256, o.in_dir, 128);
    dr.bias_off = add_bias(o.b_dir, 128);
    dr.save_off = -1; dr.mask_off = -1;
    if (save) { dr.save_off = sv; sv += 2 * BLK_BYTES; dr.mask_off = sv; sv += TILE_M * 32; pg.save_dir_off = sv; sv += BLK_BYTES; }
  }
  pg.n_steps = ns; pg.n_bias = nb;
  pg.n_ops = 0;
  for (int s = 0; s < ns; ++s) {
    const TcStep& st = pg.s[s];
    for (int kb = 0; kb < st.n_kb; ++kb) {
      uint32_t w = 0;
      if (st.src[kb] == SRC_X) w |= OP_X | (st.wait_x ? OP_WAIT : 0u);
      else w |= (st.reuse_a ? 0u : OP_WAIT) | ((uint32_t)st.src[kb] << 5);
      if (kb == 0) w |= OP_FIRST;
      if (kb == st.n_kb - 1) w |= OP_LAST | (st.x_done ? OP_XDONE : 0u);
      if (st.N == 128) w |= OP_N128;
      if (st.N == 16) w |= OP_N16;
      pg.ops[pg.n_ops++] = w;
    }
    pg.step_word[s] = (uint32_t)st.N | ((uint32_t)st.relu << 9) | ((uint32_t)st.out_blocks << 10) | ((uint32_t)st.head << 13) |
                      ((uint32_t)st.store_only << 15) | ((st.save_off >= 0 ? 1u : 0u) << 16);
  }
  add_bias(o.w_sigma, 256); add_bias(o.b_sigma, 1);
  if (!sigma_only) { add_bias(o.w_rgb, 3 * pg.rgb_in); add_bias(o.b_rgb, 3); }
  pl.pack.n_slabs = nslab; pl.bias.n_segs = nseg; pl.pack.sw64 = 0;
  pg.tile_save_bytes = save ? sv : 0;
  pl.wblob_bytes = wb; pl.bias_floats = pg.n_bias + 256 + 1 + 768 + 3;
  const int64_t n_tiles = ceil_div64(P, TILE_M);
  int64_t off = 0;
  pl.off_wblob = off; off += round_up64(wb, 1024);
  pl.off_bias = off;  off += round_up64(pl.bias_floats * 4, 1024);
  pl.off_scale = off; off += siren ? round_up64((int64_t)pg.n_bias * 4, 1024) : 0;
  pl.off_weff = off;  off += d.n_cls > 0 ? 16 * 256 * 4 + 1024 : 0;        // merged class-head matrix (fp32) | b_eff
  pl.off_save = off;  off += save ? n_tiles * (int64_t)sv : 0;
  pl.total = off;
  return pl;
}

#ifndef NERFB_FWD_PAIR_DEFAULT
#define NERFB_FWD_PAIR_DEFAULT 1
#endif
// NERFB_FWD_PAIR (read per call): 0 = single-tile kernels everywhere; 1 (default) = tile-pair kernel for inference (measured on
// the B200: 0.71 -> 0.62-0.66 ms at 4096 x 128, NeRF 800x800 render 4.18 -> 4.51 M rays/s); 2 = inference on a CTA pair
// (cta_group::2; correct, but the per-layer cross-SM hand-offs make it slower: 0.70-0.75 ms); 3 = tile pair for the training
// forward as well (correct -- the whole GPU suite passes with it -- but 1.07 ms against 1.01 ms: that kernel is bound by
// the 2.8 GB of activation stores, not by the epilogue / MMA serialisation the pair schedule removes).
static int fwd_pair_mode() {
  const char* e = getenv("NERFB_FWD_PAIR");
  return e && *e ? atoi(e) : NERFB_FWD_PAIR_DEFAULT;
}
static long long* g_trace = nullptr;
extern "C" void nerfb_debug_trace(long long* device_buf) { g_trace = device_buf; }   // diagnostics only (40 x 16 slots)

int64_t tc_bwd_workspace_bytes(const nerfb_mlp_desc& d, int64_t P);   // mlp_tc_bwd.cu
bool tc_bwd_supported(const nerfb_mlp_desc& d);

int64_t tc_workspace_bytes(const nerfb_mlp_desc& d, int64_t P, int sigma_only, int save) {
  if (!tc_supported(d, 0)) return NERFB_E_UNSUPPORTED;
  if (save && !tc_bwd_supported(d)) return NERFB_E_UNSUPPORTED;
  int64_t total = tc_plan(d, P, sigma_only != 0, save != 0).total;
  if (save) total = round_up64(total, 1024) + tc_bwd_workspace_bytes(d, P);
  return total;
}

bool tc_composite_supported(const nerfb_mlp_desc& d, int in_mode, int n_samples, int save) {
  return tc_supported(d, in_mode) && in_mode == NERFB_IN_RAYS && !save && d.n_cls == 0 &&
         (n_samples == 32 || n_samples == 64 || n_samples == 128);
}

int tc_mlp_fwd(const nerfb_mlp_desc& d, int in_mode, const float* params, const float* film, const float* in0,
               const float* in1, int64_t n_rows, int n_samples, int sigma_only, int save, float* out, void* ws,
               int64_t ws_bytes, void* stream, const TcComposite* comp) {
  const bool siren = d.kind == NERFB_KIND_SIREN;
  if (comp && comp->on && !tc_composite_supported(d, in_mode, n_samples, save)) return NERFB_E_UNSUPPORTED;
  if (siren && !film) return NERFB_E_BADARG;
  if (in_mode == NERFB_IN_GRID && (save || siren || d.in_xyz > 0)) return NERFB_E_UNSUPPORTED;
  const int64_t P = (in_mode == NERFB_IN_RAYS) ? n_rows * n_samples : n_rows;
  int pair_mode = fwd_pair_mode();                      // two tiles in flight per SM (mlp_tc_fwd_pair_kernel / _pair2_kernel)
  if (save) pair_mode = pair_mode == 3 ? 1 : 0;         // training forward: single-tile kernel unless asked for
  else if (pair_mode == 3) pair_mode = 1;
  const bool pair = pair_mode != 0;
  const TcPlan pl = tc_plan(d, P, sigma_only != 0, save != 0, pair);
  if (ws_bytes < pl.total) return NERFB_E_WORKSPACE;
  if ((reinterpret_cast<uintptr_t>(ws) & 1023) != 0) return NERFB_E_BADARG;
  uint8_t* base = static_cast<uint8_t*>(ws);
  float* weff = reinterpret_cast<float*>(base + pl.off_weff);
  if (pl.cls_bias_off >= 0) {
    const NerfOffsets o = nerf_offsets(d);
    NERFB_LAUNCH(merge_parse_kernel, 16, 256, 0, stream, params, o.w_p0, o.b_p0, o.w_p1, o.b_p1, d.n_cls, weff, weff + 16 * 256);
  }
  NERFB_LAUNCH(pack_weights_kernel, pl.pack.n_slabs * 4, 256, 0, stream, params, pl.pack, base + pl.off_wblob, (const float*)weff,
               siren ? film : (const float*)nullptr);
  NERFB_LAUNCH(gather_bias_kernel, pl.bias.n_segs, 256, 0, stream, params, pl.bias,
               reinterpret_cast<float*>(base + pl.off_bias), siren ? film : (const float*)nullptr,
               reinterpret_cast<float*>(base + pl.off_scale), (const float*)weff);
  TcFwdArgs a{};
  a.prog = pl.prog;
  a.wblob = base + pl.off_wblob;
  a.bias = reinterpret_cast<const float*>(base + pl.off_bias);
  a.scale = reinterpret_cast<const float*>(base + pl.off_scale);
  a.in0 = in0; a.in1 = in1; a.in_mode = in_mode; a.S = n_samples;
  a.P = P; a.n_tiles = ceil_div64(P, TILE_M);
  a.out = out;
  if (comp) a.comp = *comp;
  a.save = save ? base + pl.off_save : nullptr;
  a.trace = g_trace;
  const size_t smem = sizeof(Smem) + 1024, smem_p = sizeof(SmemP) + 1024;
  int dev = 0;
  NERFB_CUDA(cudaGetDevice(&dev));
  static bool attr_set[64] = {};          // cudaFuncSetAttribute is per device
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    NERFB_CUDA(cudaFuncSetAttribute((mlp_tc_fwd_pair_kernel<false, false, false>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_p));
    NERFB_CUDA(cudaFuncSetAttribute((mlp_tc_fwd_pair_kernel<true, false, false>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_p));
    NERFB_CUDA(cudaFuncSetAttribute((mlp_tc_fwd_pair_kernel<false, true, false>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_p));
    NERFB_CUDA(cudaFuncSetAttribute((mlp_tc_fwd_pair_kernel<false, false, true>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_p));
    NERFB_CUDA(cudaFuncSetAttribute((mlp_tc_fwd_pair_kernel<true, false, true>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_p));
    NERFB_CUDA(cudaFuncSetAttribute((mlp_tc_fwd_pair_kernel<false, true, true>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_p));
    NERFB_CUDA(cudaFuncSetAttribute((mlp_tc_fwd_pair2_kernel<false>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_p));
    NERFB_CUDA(cudaFuncSetAttribute((mlp_tc_fwd_pair2_kernel<true>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_p));
    NERFB_CUDA(cudaFuncSetAttribute((mlp_tc_fwd_kernel<false, false, false>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    NERFB_CUDA(cudaFuncSetAttribute((mlp_tc_fwd_kernel<false, false, true>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    NERFB_CUDA(cudaFuncSetAttribute((mlp_tc_fwd_kernel<true, false, false>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    NERFB_CUDA(cudaFuncSetAttribute((mlp_tc_fwd_kernel<true, false, true>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    NERFB_CUDA(cudaFuncSetAttribute((mlp_tc_fwd_kernel<false, true, false>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    NERFB_CUDA(cudaFuncSetAttribute((mlp_tc_fwd_kernel<false, true, true>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  if (pair_mode == 2 && pl.prog.n_cls == 0 && a.n_tiles >= 4) {
    const int64_t n_quads = (a.n_tiles + 3) >> 2;
    const int n_cl = (int)(n_quads < NERFB_SM_COUNT / 2 ? n_quads : NERFB_SM_COUNT / 2);
    if (siren) NERFB_LAUNCH((mlp_tc_fwd_pair2_kernel<true>), 2 * n_cl, NUM_THREADS, smem_p, stream, a);
    else NERFB_LAUNCH((mlp_tc_fwd_pair2_kernel<false>), 2 * n_cl, NUM_THREADS, smem_p, stream, a);
    return NERFB_OK;
  }
  if (pair) {
    const int64_t n_pairs = (a.n_tiles + 1) >> 1;
    const int pgrid = (int)(n_pairs < NERFB_SM_COUNT ? n_pairs : NERFB_SM_COUNT);
    const bool svp = a.save != nullptr;
    if (siren && svp) NERFB_LAUNCH((mlp_tc_fwd_pair_kernel<true, false, true>), pgrid, NUM_THREADS, smem_p, stream, a);
    else if (siren) NERFB_LAUNCH((mlp_tc_fwd_pair_kernel<true, false, false>), pgrid, NUM_THREADS, smem_p, stream, a);
    else if (pl.prog.n_cls > 0 && svp) NERFB_LAUNCH((mlp_tc_fwd_pair_kernel<false, true, true>), pgrid, NUM_THREADS, smem_p, stream, a);
    else if (pl.prog.n_cls > 0) NERFB_LAUNCH((mlp_tc_fwd_pair_kernel<false, true, false>), pgrid, NUM_THREADS, smem_p, stream, a);
    else if (svp) NERFB_LAUNCH((mlp_tc_fwd_pair_kernel<false, false, true>), pgrid, NUM_THREADS, smem_p, stream, a);
    else NERFB_LAUNCH((mlp_tc_fwd_pair_kernel<false, false, false>), pgrid, NUM_THREADS, smem_p, stream, a);
    return NERFB_OK;
  }
  const int grid = (int)(a.n_tiles < NERFB_SM_COUNT ? a.n_tiles : NERFB_SM_COUNT);
  const bool sv = a.save != nullptr;
  if (siren && sv) NERFB_LAUNCH((mlp_tc_fwd_kernel<true, false, true>), grid, NUM_THREADS, smem, stream, a);
  else if (siren) NERFB_LAUNCH((mlp_tc_fwd_kernel<true, false, false>), grid, NUM_THREADS, smem, stream, a);
  else if (pl.prog.n_cls > 0 && sv) NERFB_LAUNCH((mlp_tc_fwd_kernel<false, true, true>), grid, NUM_THREADS, smem, stream, a);
  else if (pl.prog.n_cls > 0) NERFB_LAUNCH((mlp_tc_fwd_kernel<false, true, false>), grid, NUM_THREADS, smem, stream, a);
  else if (sv) NERFB_LAUNCH((mlp_tc_fwd_kernel<false, false, true>), grid, NUM_THREADS, smem, stream, a);
  else NERFB_LAUNCH((mlp_tc_fwd_kernel<false, false, false>), grid, NUM_THREADS, smem, stream, a);
  return NERFB_OK;
}

// Layout of the saved-activation records, for the backward kernels (mlp_tc_bwd.cu).
TcFwdLayout tc_fwd_layout(const nerfb_mlp_desc& d, int64_t P) {
  const TcPlan pl = tc_plan(d, P, false, true);
  TcFwdLayout L{};
  L.n_steps = pl.prog.n_steps;
  for (int s = 0; s < pl.prog.n_steps; ++s) { L.step_save_off[s] = pl.prog.s[s].save_off; L.step_mask_off[s] = pl.prog.s[s].mask_off; }
  L.save_x_off = pl.prog.save_x_off; L.save_dir_off = pl.prog.save_dir_off;
  L.tile_save_bytes = pl.prog.tile_save_bytes;
  L.off_save = pl.off_save; L.total = round_up64(pl.total, 1024);
  L.off_weff = pl.off_weff;
  return L;
}
