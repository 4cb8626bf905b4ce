// Kernel (1), tensor-core arm, forward: fused positional encoding + NeRF MLP on tcgen05 / TMEM.
//   replaces models/rendering.py:129-159 (embed -> cat -> model per 32k-point chunk) and models/nerf.py:100-124.
//
// One persistent CTA per SM walks 128-sample tiles.  Per tile the activations never leave the SM:
//   * warps 2..9 (256 threads) build the bf16 encoding tile X [128 x 64] in shared memory (x = o + d z, sin/cos by
//     exact sincosf + double-angle recurrence), and later run every layer's epilogue: tcgen05.ld the fp32 accumulator
//     from TMEM, + bias, ReLU, round to bf16, store as the next layer's A operand (128B-swizzled K-major blocks);
//   * warp 1 (one elected lane) issues tcgen05.mma 128 x N x 16 (N = 256 / 128) with A from those blocks and B from a
//     4-stage ring of weight slabs; accumulators ping-pong between TMEM columns [0,256) and [256,512);
//   * warp 0 (one lane) streams the pre-packed bf16 weight slabs (already in the swizzled shared-memory image, so a
//     slab is ONE cp.async.bulk of 32 KB) from L2 into the ring.
// The layer chain is pipelined at K-block granularity: the epilogue of layer l publishes its output in four 64-column
// blocks (mbarrier each); the MMA of layer l+1 starts on block 0 while blocks 1..3 are still in the epilogue, into the
// other accumulator.  Skip / direction concatenations are extra K-blocks (the resident X tile), never a cat.
// sigma (256->1) and rgb (128->3, sigmoid) heads are fp32 dot products folded into the epilogues.
// With save != 0 every layer's bf16 output block is also bulk-stored to HBM (same swizzled image) for the backward.
#include "common.cuh"
#include "tc_common.cuh"

using namespace tc;

constexpr int N_STAGES = 4;
constexpr int STAGE_BYTES = 256 * 128;          // one [256 x 64] bf16 weight slab
constexpr int NUM_THREADS = 320;                // warp 0 producer, warp 1 MMA, warps 2..9 epilogue
constexpr int EPI_THREADS = 256;
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
  int write_dir;       // epilogue rebuilds X as the direction encoding once this step's MMAs are done
  int save_off;        // byte offset of this step's output blocks in the tile's saved record (-1 = not saved)
};

struct TcProgram {
  int n_steps;
  TcStep s[MAX_STEPS];
  int n_bias;                      // floats in the bias table
  int save_x_off, save_dir_off;    // saved-record offsets of the two encoding tiles
  int tile_save_bytes;             // saved bytes per tile (0 = nothing saved)
  int n_freq_xyz, n_freq_dir, in_xyz, in_dir;
  int sigma_only, out_c;
};

struct TcFwdArgs {
  TcProgram prog;
  const uint8_t* wblob;            // packed bf16 weight slabs in schedule order
  const float* bias;               // fp32 bias table [n_bias] | w_sigma[256] | b_sigma | w_rgb[3*128] | b_rgb[3]
  const float* in0; const float* in1; int in_mode; int S;
  int64_t P; int64_t n_tiles;
  float* out;
  uint8_t* save;                   // saved activations, tile-major
};

// shared memory map (dynamic, 1024-aligned)
struct __align__(1024) Smem {
  uint8_t a[4][BLK_BYTES];               // activation tile: 4 K-blocks
  uint8_t x[BLK_BYTES];                  // encoding tile
  uint8_t ring[N_STAGES][STAGE_BYTES];   // weight slabs
  float bias[10 * 256 + 128];
  float w_sigma[256];
  float w_rgb[3 * 128];
  float b_heads[4];                      // b_sigma, b_rgb[3]
  float part[2][3][TILE_M];              // head partial sums [column half][channel][row]
  float sigma[TILE_M];
  uint64_t w_full[N_STAGES], w_empty[N_STAGES];
  uint64_t a_ready[4];
  uint64_t x_ready;
  uint64_t acc_full[2];
  uint32_t tmem_base;
};

static_assert(sizeof(Smem) + 1024 <= 232448, "shared memory budget (227 KB per CTA)");

// xyz / dir -> bf16 encoding row.  The row is [v(3), sin(2^k v)(3), cos(2^k v)(3), ...]; this thread produces the 32
// columns [C_LO, C_LO+32) (everything is unrolled so the window lives in registers), zero beyond the encoding width.
template <int C_LO>
__device__ __forceinline__ void encode_row(float vx, float vy, float vz, int n_freq, uint32_t row_base, int row) {
  float s[3], c[3];
  const float v[3] = {vx, vy, vz};
#pragma unroll
  for (int j = 0; j < 3; ++j) sincosf(v[j], &s[j], &c[j]);
  float buf[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) buf[i] = 0.f;
#pragma unroll
  for (int j = 0; j < 3; ++j)
    if (j >= C_LO && j < C_LO + 32) buf[(j - C_LO) & 31] = v[j];
#pragma unroll
  for (int k = 0; k < 10; ++k) {
    const bool on = k < n_freq;
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const int cs = 3 + 6 * k + j, cc = 6 + 6 * k + j;
      if (cs >= C_LO && cs < C_LO + 32) buf[(cs - C_LO) & 31] = on ? s[j] : 0.f;
      if (cc >= C_LO && cc < C_LO + 32) buf[(cc - C_LO) & 31] = on ? c[j] : 0.f;
      const float s2 = 2.f * s[j] * c[j];            // double-angle step to the next octave
      const float c2 = 1.f - 2.f * s[j] * s[j];
      s[j] = s2; c[j] = c2;
    }
  }
#pragma unroll
  for (int ch = 0; ch < 4; ++ch) {
    const int col = C_LO + 8 * ch;
    const uint32_t addr = row_base + (uint32_t)row * 128u + ((((uint32_t)col >> 3) ^ ((uint32_t)row & 7u)) << 4);
    sts128(addr, pack_bf16(buf[8 * ch + 0], buf[8 * ch + 1]), pack_bf16(buf[8 * ch + 2], buf[8 * ch + 3]),
           pack_bf16(buf[8 * ch + 4], buf[8 * ch + 5]), pack_bf16(buf[8 * ch + 6], buf[8 * ch + 7]));
  }
}

// already-embedded input: copy 32 columns [c_lo, c_lo+32) of src[0..width) (zero padded) into the X tile row
__device__ __forceinline__ void copy_row(const float* src, int width, int c_lo, uint32_t row_base, int row) {
#pragma unroll
  for (int ch = 0; ch < 4; ++ch) {
    float b[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) { const int col = c_lo + 8 * ch + e; b[e] = (src && col < width) ? __ldg(src + col) : 0.f; }
    const int col = c_lo + 8 * ch;
    const uint32_t addr = row_base + (uint32_t)row * 128u + ((((uint32_t)col >> 3) ^ ((uint32_t)row & 7u)) << 4);
    sts128(addr, pack_bf16(b[0], b[1]), pack_bf16(b[2], b[3]), pack_bf16(b[4], b[5]), pack_bf16(b[6], b[7]));
  }
}

__global__ void __launch_bounds__(NUM_THREADS, 1) mlp_tc_fwd_kernel(const __grid_constant__ TcFwdArgs args) {
  extern __shared__ uint8_t smem_raw[];
  Smem& sm = *reinterpret_cast<Smem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const TcProgram& pg = args.prog;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // ---- one-time setup -----------------------------------------------------------------------------------
  for (int i = threadIdx.x; i < pg.n_bias; i += NUM_THREADS) sm.bias[i] = __ldg(args.bias + i);
  for (int i = threadIdx.x; i < 256; i += NUM_THREADS) sm.w_sigma[i] = __ldg(args.bias + pg.n_bias + i);
  if (threadIdx.x == 0) sm.b_heads[0] = __ldg(args.bias + pg.n_bias + 256);
  for (int i = threadIdx.x; i < 3 * 128; i += NUM_THREADS) sm.w_rgb[i] = __ldg(args.bias + pg.n_bias + 257 + i);
  if (threadIdx.x < 3) sm.b_heads[1 + threadIdx.x] = __ldg(args.bias + pg.n_bias + 257 + 384 + threadIdx.x);
  if (threadIdx.x == 0) {
    for (int i = 0; i < N_STAGES; ++i) { mbar_init(&sm.w_full[i], 1); mbar_init(&sm.w_empty[i], 1); }
    for (int i = 0; i < 4; ++i) mbar_init(&sm.a_ready[i], 8);          // one arrive per epilogue warp
    mbar_init(&sm.x_ready, 8);
    mbar_init(&sm.acc_full[0], 1); mbar_init(&sm.acc_full[1], 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&sm.tmem_base, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp == 0) {
    // ===== weight producer =====
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int64_t tile = blockIdx.x; tile < args.n_tiles; tile += gridDim.x) {
        const uint8_t* src = args.wblob;
        for (int s = 0; s < pg.n_steps; ++s) {
          const uint32_t bytes = (uint32_t)pg.s[s].N * 128u;
          for (int kb = 0; kb < pg.s[s].n_kb; ++kb) {
            mbar_wait(&sm.w_empty[stage], phase ^ 1);
            mbar_expect_tx(&sm.w_full[stage], bytes);
            bulk_g2s(sm.ring[stage], src, bytes, &sm.w_full[stage]);
            src += bytes;
            if (++stage == N_STAGES) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (single thread) =====
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      uint32_t a_phase[4] = {0, 0, 0, 0}, x_phase = 0;
      uint32_t gs = 0;                                   // global step counter -> accumulator ping-pong
      for (int64_t tile = blockIdx.x; tile < args.n_tiles; tile += gridDim.x) {
        uint8_t* rec = args.save ? args.save + tile * (int64_t)pg.tile_save_bytes : nullptr;
        bool x_is_dir = false;
        for (int s = 0; s < pg.n_steps; ++s, ++gs) {
          const TcStep& st = pg.s[s];
          const uint32_t d_tmem = tmem + (gs & 1) * 256;
          const uint32_t idesc = umma_idesc_bf16(TILE_M, st.N);
          bool stored = false;
          for (int kb = 0; kb < st.n_kb; ++kb) {
            const int src = st.src[kb];
            uint32_t a_addr;
            if (src == SRC_X) {
              if (st.wait_x) {
                mbar_wait(&sm.x_ready, x_phase); x_phase ^= 1;
                if (rec) { bulk_s2g(rec + (x_is_dir ? pg.save_dir_off : pg.save_x_off), sm.x, BLK_BYTES); stored = true; }
                x_is_dir = true;
              }
              a_addr = smem_u32(sm.x);
            } else {
              mbar_wait(&sm.a_ready[src], a_phase[src]); a_phase[src] ^= 1;      // produced by step s-1's epilogue
              if (rec && pg.s[s - 1].save_off >= 0) {
                bulk_s2g(rec + pg.s[s - 1].save_off + src * BLK_BYTES, sm.a[src], BLK_BYTES); stored = true;
              }
              a_addr = smem_u32(sm.a[src]);
            }
            mbar_wait(&sm.w_full[stage], phase);
            tc_fence_after();
            const uint32_t b_addr = smem_u32(sm.ring[stage]);
#pragma unroll
            for (int k = 0; k < KBLK / 16; ++k)
              umma_bf16(d_tmem, umma_desc_k128(a_addr + k * 32), umma_desc_k128(b_addr + k * 32), idesc,
                        (kb > 0 || k > 0) ? 1u : 0u);
            umma_commit(&sm.w_empty[stage]);             // slab free once these MMAs have read it
            if (++stage == N_STAGES) { stage = 0; phase ^= 1; }
          }
          if (stored) { bulk_commit(); bulk_wait_read0(); }   // the epilogue may overwrite the blocks after acc_full
          umma_commit(&sm.acc_full[gs & 1]);
        }
        // blocks produced by the last step (direction-layer activations, saved for the backward only)
        const TcStep& last = pg.s[pg.n_steps - 1];
        if (last.out_blocks > 0) {
          for (int j = 0; j < last.out_blocks; ++j) {
            mbar_wait(&sm.a_ready[j], a_phase[j]); a_phase[j] ^= 1;
            if (rec && last.save_off >= 0) bulk_s2g(rec + last.save_off + j * BLK_BYTES, sm.a[j], BLK_BYTES);
          }
          bulk_commit(); bulk_wait_read0();
        }
      }
      bulk_wait_all0();
    }
  } else {
    // ===== encoding builder + epilogue warps =====
    const int ew = warp - 2;                 // 0..7
    const int q = warp & 3;                  // TMEM lane quarter this warp may access (warp id mod 4)
    const int h = ew >> 2;                   // which 32-column half of every 64-column block
    const int row = q * 32 + lane;
    uint32_t acc_phase[2] = {0, 0};
    uint32_t gs = 0;
    const uint32_t x_base = smem_u32(sm.x);
    for (int64_t tile = blockIdx.x; tile < args.n_tiles; tile += gridDim.x) {
      const int64_t p = tile * TILE_M + row;
      const bool live = p < args.P;
      const int64_t pc = live ? p : args.P - 1;            // padded rows recompute the last point (never stored)
      float dirx = 0.f, diry = 0.f, dirz = 0.f;
      const float* emb_src = nullptr;
      // ---- X tile: xyz encoding (models/rendering.py:224 + models/nerf.py:33-38) ----
      if (args.in_mode == NERFB_IN_RAYS) {
        const int64_t r = pc / args.S;
        const float* ray = args.in0 + r * 8;
        const float z = __ldg(args.in1 + pc);
        dirx = __ldg(ray + 3); diry = __ldg(ray + 4); dirz = __ldg(ray + 5);
        const float px = __fadd_rn(__ldg(ray + 0), __fmul_rn(dirx, z));
        const float py = __fadd_rn(__ldg(ray + 1), __fmul_rn(diry, z));
        const float pz = __fadd_rn(__ldg(ray + 2), __fmul_rn(dirz, z));
        if (h == 0) encode_row<0>(px, py, pz, pg.n_freq_xyz, x_base, row);
        else encode_row<32>(px, py, pz, pg.n_freq_xyz, x_base, row);
      } else {
        emb_src = args.in0 + pc * (int64_t)(pg.in_xyz + (pg.sigma_only ? 0 : pg.in_dir));
        copy_row(emb_src, pg.in_xyz, 32 * h, x_base, row);
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm.x_ready);

      float head_acc[3] = {0.f, 0.f, 0.f};
      for (int s = 0; s < pg.n_steps; ++s, ++gs) {
        const TcStep& st = pg.s[s];
        const int a = gs & 1;
        mbar_wait(&sm.acc_full[a], acc_phase[a]); acc_phase[a] ^= 1;
        tc_fence_after();
        if (st.head) { head_acc[0] = head_acc[1] = head_acc[2] = 0.f; }
        const int n_blocks = st.N / 64;
        for (int j = 0; j < n_blocks; ++j) {
          uint32_t v[32];
          const int col0 = 64 * j + 32 * h;
          tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(a * 256 + col0), v);
          tmem_ld_wait();
          float f[32];
          const float4* b4 = reinterpret_cast<const float4*>(sm.bias + st.bias_off + col0);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float4 b = b4[i];
            f[4 * i + 0] = __uint_as_float(v[4 * i + 0]) + b.x; f[4 * i + 1] = __uint_as_float(v[4 * i + 1]) + b.y;
            f[4 * i + 2] = __uint_as_float(v[4 * i + 2]) + b.z; f[4 * i + 3] = __uint_as_float(v[4 * i + 3]) + b.w;
          }
          if (st.relu) {
#pragma unroll
            for (int i = 0; i < 32; ++i) f[i] = fmaxf(f[i], 0.f);
          }
          if (st.head == 1) {
            const float4* w4 = reinterpret_cast<const float4*>(sm.w_sigma + col0);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float4 w = w4[i];
              head_acc[0] = fmaf(f[4 * i], w.x, head_acc[0]); head_acc[0] = fmaf(f[4 * i + 1], w.y, head_acc[0]);
              head_acc[0] = fmaf(f[4 * i + 2], w.z, head_acc[0]); head_acc[0] = fmaf(f[4 * i + 3], w.w, head_acc[0]);
            }
          } else if (st.head == 2) {
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              head_acc[0] = fmaf(f[i], sm.w_rgb[col0 + i], head_acc[0]);
              head_acc[1] = fmaf(f[i], sm.w_rgb[128 + col0 + i], head_acc[1]);
              head_acc[2] = fmaf(f[i], sm.w_rgb[256 + col0 + i], head_acc[2]);
            }
          }
          if (j < st.out_blocks) {
            const uint32_t blk = smem_u32(sm.a[j]);
#pragma unroll
            for (int ch = 0; ch < 4; ++ch) {
              const uint32_t chunk = (uint32_t)(4 * h + ch);
              const uint32_t addr = blk + (uint32_t)row * 128u + ((chunk ^ ((uint32_t)row & 7u)) << 4);
              sts128(addr, pack_bf16(f[8 * ch + 0], f[8 * ch + 1]), pack_bf16(f[8 * ch + 2], f[8 * ch + 3]),
                     pack_bf16(f[8 * ch + 4], f[8 * ch + 5]), pack_bf16(f[8 * ch + 6], f[8 * ch + 7]));
            }
            fence_proxy_async();
          }
          tc_fence_before();
          __syncwarp();
          if (j < st.out_blocks && lane == 0) mbar_arrive(&sm.a_ready[j]);
        }
        if (st.write_dir) {       // X (xyz encoding) is dead once this step's MMAs completed: rebuild it as dir encoding
          if (args.in_mode == NERFB_IN_RAYS) {
            if (h == 0) encode_row<0>(dirx, diry, dirz, pg.n_freq_dir, x_base, row);
            else copy_row(nullptr, 0, 32, x_base, row);
          } else {
            copy_row(emb_src + pg.in_xyz, pg.in_dir, 32 * h, x_base, row);
          }
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) mbar_arrive(&sm.x_ready);
        }
        if (st.head == 1) {       // sigma = <h_D, w_sigma> + b  (models/nerf.py:112), fp32 on the un-rounded activations
          sm.part[h][0][row] = head_acc[0];
          named_bar_sync(1, EPI_THREADS);
          if (h == 0) {
            const float sg = sm.part[0][0][row] + sm.part[1][0][row] + sm.b_heads[0];
            sm.sigma[row] = sg;
            if (pg.sigma_only && live) args.out[p] = sg;
          }
        } else if (st.head == 2) { // rgb = sigmoid(W_rgb h_dir + b)  (models/nerf.py:119-120)
          sm.part[h][0][row] = head_acc[0]; sm.part[h][1][row] = head_acc[1]; sm.part[h][2][row] = head_acc[2];
          named_bar_sync(1, EPI_THREADS);
          if (h == 0 && live) {
            float4 o;
            o.x = 1.f / (1.f + __expf(-(sm.part[0][0][row] + sm.part[1][0][row] + sm.b_heads[1])));
            o.y = 1.f / (1.f + __expf(-(sm.part[0][1][row] + sm.part[1][1][row] + sm.b_heads[2])));
            o.z = 1.f / (1.f + __expf(-(sm.part[0][2][row] + sm.part[1][2][row] + sm.b_heads[3])));
            o.w = sm.sigma[row];
            *reinterpret_cast<float4*>(args.out + p * 4) = o;
          }
        }
      }
      named_bar_sync(2, EPI_THREADS);          // head scratch is reused by the next tile
    }
  }

  // ---- teardown ------------------------------------------------------------------------------------------
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

// ---- weight packing: fp32 nn.Linear weights -> bf16 slabs in the swizzled shared-memory image, schedule order ----
struct PackSlab { int64_t w_off; int ld; int col0; int ncols; int N; int64_t dst_off; };   // W[n*ld + col0 + kk], kk < ncols
struct PackArgs { int n_slabs; PackSlab slab[48]; };

__global__ void __launch_bounds__(256) pack_weights_kernel(const float* __restrict__ params, PackArgs pa, uint8_t* __restrict__ blob) {
  const PackSlab sl = pa.slab[blockIdx.x];
  for (int idx = threadIdx.x; idx < sl.N * 8; idx += blockDim.x) {       // one 16-byte chunk per iteration
    const int n = idx >> 3, ch = idx & 7;
    float f[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int kk = ch * 8 + e;
      f[e] = (kk < sl.ncols) ? __ldg(params + sl.w_off + (int64_t)n * sl.ld + sl.col0 + kk) : 0.f;
    }
    uint4 o;
    o.x = pack_bf16(f[0], f[1]); o.y = pack_bf16(f[2], f[3]); o.z = pack_bf16(f[4], f[5]); o.w = pack_bf16(f[6], f[7]);
    *reinterpret_cast<uint4*>(blob + sl.dst_off + (int64_t)n * 128 + ((ch ^ (n & 7)) << 4)) = o;
  }
}

// bias table: [step biases ...] | w_sigma | b_sigma | w_rgb | b_rgb, gathered from the flat parameter blob
struct BiasSeg { int64_t src; int n; int dst; };
struct BiasArgs { int n_segs; BiasSeg seg[20]; };
__global__ void __launch_bounds__(256) gather_bias_kernel(const float* __restrict__ params, BiasArgs ba, float* __restrict__ table) {
  const BiasSeg sg = ba.seg[blockIdx.x];
  for (int i = threadIdx.x; i < sg.n; i += blockDim.x) table[sg.dst + i] = __ldg(params + sg.src + i);
}

// ---- host side: program construction, workspace plan, launch ---------------------------------------------------
struct TcPlan {
  TcProgram prog;
  PackArgs pack;
  BiasArgs bias;
  int64_t wblob_bytes, bias_floats;
  int64_t off_wblob, off_bias, off_save, total;
};

bool tc_supported(const nerfb_mlp_desc& d, int in_mode) {
  (void)in_mode;
  if (d.kind != NERFB_KIND_NERF || d.W != 256 || d.n_cls != 0) return false;
  if (d.D < 1 || d.D > 9) return false;
  if (!d.logscale && d.in_xyz <= 0) return false;                 // in-kernel encoding is the power-of-two ladder
  if (nerfb_in_xyz(d) > 64 || nerfb_in_dir(d) > 32) return false;
  if (nerfb_in_xyz(d) < 1) return false;
  return true;
}

TcFwdLayout tc_fwd_layout(const nerfb_mlp_desc& d, int64_t P);
static TcPlan tc_plan(const nerfb_mlp_desc& d, int64_t P, bool sigma_only, bool save) {
  TcPlan pl{};
  const NerfOffsets o = nerf_offsets(d);
  TcProgram& pg = pl.prog;
  pg.n_freq_xyz = d.n_freq_xyz; pg.n_freq_dir = d.n_freq_dir; pg.in_xyz = o.in_xyz; pg.in_dir = o.in_dir;
  pg.sigma_only = sigma_only; pg.out_c = sigma_only ? 1 : 4;
  int64_t wb = 0; int nb = 0, ns = 0, nslab = 0, nseg = 0, sv = 0;
  auto add_slab = [&](int64_t w_off, int ld, int col0, int ncols, int N) {
    pl.pack.slab[nslab++] = PackSlab{w_off, ld, col0, ncols, N, wb};
    wb += (int64_t)N * 128;
  };
  auto add_bias = [&](int64_t src, int n) { pl.bias.seg[nseg++] = BiasSeg{src, n, nb}; nb += n; return nb - n; };
  pg.save_x_off = -1; pg.save_dir_off = -1;
  if (save) { pg.save_x_off = sv; sv += BLK_BYTES; }
  int last_x_reader = 0;
  for (int i = 0; i < d.D; ++i) {
    TcStep& st = pg.s[ns++];
    st = TcStep{};
    const bool skip = i > 0 && ((d.skip_mask >> i) & 1);
    st.N = 256; st.relu = 1; st.out_blocks = 4; st.head = (i == d.D - 1) ? 1 : 0;
    if (i == 0) { st.src[st.n_kb++] = SRC_X; st.wait_x = 1; add_slab(o.w[0], o.in[0], 0, o.in_xyz, 256); }
    else {
      if (skip) { st.src[st.n_kb++] = SRC_X; add_slab(o.w[i], o.in[i], 0, o.in_xyz, 256); last_x_reader = i; }
      const int c0 = skip ? o.in_xyz : 0;
      for (int j = 0; j < 4; ++j) { st.src[st.n_kb++] = j; add_slab(o.w[i], o.in[i], c0 + 64 * j, 64, 256); }
    }
    st.bias_off = add_bias(o.b[i], 256);
    st.save_off = -1;
    if (save) { st.save_off = sv; sv += 4 * BLK_BYTES; }
    if (sigma_only && i == d.D - 1) { st.out_blocks = 0; st.save_off = -1; }
  }
  if (!sigma_only) {
    pg.s[last_x_reader].write_dir = 1;
    TcStep& fin = pg.s[ns++]; fin = TcStep{};
    fin.N = 256; fin.relu = 0; fin.out_blocks = 4;
    for (int j = 0; j < 4; ++j) { fin.src[fin.n_kb++] = j; add_slab(o.w_final, 256, 64 * j, 64, 256); }
    fin.bias_off = add_bias(o.b_final, 256);
    fin.save_off = -1;
    if (save) { fin.save_off = sv; sv += 4 * BLK_BYTES; }
    TcStep& dr = pg.s[ns++]; dr = TcStep{};
    dr.N = 128; dr.relu = 1; dr.head = 2; dr.out_blocks = save ? 2 : 0;
    for (int j = 0; j < 4; ++j) { dr.src[dr.n_kb++] = j; add_slab(o.w_dir, 256 + o.in_dir, 64 * j, 64, 128); }
    dr.src[dr.n_kb++] = SRC_X; dr.wait_x = 1; add_slab(o.w_dir, 256 + o.in_dir, 256, o.in_dir, 128);
    dr.bias_off = add_bias(o.b_dir, 128);
    dr.save_off = -1;
    if (save) { dr.save_off = sv; sv += 2 * BLK_BYTES; pg.save_dir_off = sv; sv += BLK_BYTES; }
  }
  pg.n_steps = ns; pg.n_bias = nb;
  add_bias(o.w_sigma, 256); add_bias(o.b_sigma, 1);
  if (!sigma_only) { add_bias(o.w_rgb, 384); add_bias(o.b_rgb, 3); }
  pl.pack.n_slabs = nslab; pl.bias.n_segs = nseg;
  pg.tile_save_bytes = save ? sv : 0;
  pl.wblob_bytes = wb; pl.bias_floats = pg.n_bias + 256 + 1 + 384 + 3;
  const int64_t n_tiles = ceil_div64(P, TILE_M);
  int64_t off = 0;
  pl.off_wblob = off; off += round_up64(wb, 1024);
  pl.off_bias = off;  off += round_up64(pl.bias_floats * 4, 1024);
  pl.off_save = off;  off += save ? n_tiles * (int64_t)sv : 0;
  pl.total = off;
  return pl;
}

int64_t tc_bwd_workspace_bytes(const nerfb_mlp_desc& d, int64_t P);   // mlp_tc_bwd.cu
bool tc_bwd_supported(const nerfb_mlp_desc& d);

int64_t tc_workspace_bytes(const nerfb_mlp_desc& d, int64_t P, int sigma_only, int save) {
  if (!tc_supported(d, 0)) return NERFB_E_UNSUPPORTED;
  if (save && !tc_bwd_supported(d)) return NERFB_E_UNSUPPORTED;
  int64_t total = tc_plan(d, P, sigma_only != 0, save != 0).total;
  if (save) total = round_up64(total, 1024) + tc_bwd_workspace_bytes(d, P);
  return total;
}

int tc_mlp_fwd(const nerfb_mlp_desc& d, int in_mode, const float* params, const float* film, const float* in0,
               const float* in1, int64_t n_rows, int n_samples, int sigma_only, int save, float* out, void* ws,
               int64_t ws_bytes, void* stream) {
  (void)film;
  const int64_t P = (in_mode == NERFB_IN_RAYS) ? n_rows * n_samples : n_rows;
  const TcPlan pl = tc_plan(d, P, sigma_only != 0, save != 0);
  if (ws_bytes < pl.total) return NERFB_E_WORKSPACE;
  if ((reinterpret_cast<uintptr_t>(ws) & 1023) != 0) return NERFB_E_BADARG;
  uint8_t* base = static_cast<uint8_t*>(ws);
  NERFB_LAUNCH(pack_weights_kernel, pl.pack.n_slabs, 256, 0, stream, params, pl.pack, base + pl.off_wblob);
  NERFB_LAUNCH(gather_bias_kernel, pl.bias.n_segs, 256, 0, stream, params, pl.bias,
               reinterpret_cast<float*>(base + pl.off_bias));
  TcFwdArgs a{};
  a.prog = pl.prog;
  a.wblob = base + pl.off_wblob;
  a.bias = reinterpret_cast<const float*>(base + pl.off_bias);
  a.in0 = in0; a.in1 = in1; a.in_mode = in_mode; a.S = n_samples;
  a.P = P; a.n_tiles = ceil_div64(P, TILE_M);
  a.out = out;
  a.save = save ? base + pl.off_save : nullptr;
  const size_t smem = sizeof(Smem) + 1024;
  static bool attr_set = false;
  if (!attr_set) {
    NERFB_CUDA(cudaFuncSetAttribute(mlp_tc_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_set = true;
  }
  const int grid = (int)(a.n_tiles < NERFB_SM_COUNT ? a.n_tiles : NERFB_SM_COUNT);
  NERFB_LAUNCH(mlp_tc_fwd_kernel, grid, NUM_THREADS, smem, stream, a);
  return NERFB_OK;
}

// Layout of the saved-activation records, for the backward kernels (mlp_tc_bwd.cu).
TcFwdLayout tc_fwd_layout(const nerfb_mlp_desc& d, int64_t P) {
  const TcPlan pl = tc_plan(d, P, false, true);
  TcFwdLayout L{};
  L.n_steps = pl.prog.n_steps;
  for (int s = 0; s < pl.prog.n_steps; ++s) L.step_save_off[s] = pl.prog.s[s].save_off;
  L.save_x_off = pl.prog.save_x_off; L.save_dir_off = pl.prog.save_dir_off;
  L.tile_save_bytes = pl.prog.tile_save_bytes;
  L.off_save = pl.off_save; L.total = round_up64(pl.total, 1024);
  return L;
}
