// Kernel (1), tensor-core arm, backward: autograd of the fused encoding + NeRF MLP (models/nerf.py:100-124) on
// tcgen05 / TMEM, from the bf16 activations the forward kernel parked per 128-sample tile.
//
// Three launches per call:
//   A. mlp_tc_dgrad_kernel   -- persistent, one CTA per SM, same warp roles / K-block pipelined layer chain as the
//      forward but walking the network backwards: per tile, dZ_dir (rgb head folded in on the CUDA cores) ->
//      x W_dir^T -> dZ_final -> x W_final^T (+ d_sigma (x) w_sigma) (.) [h8 > 0] -> dZ_8 -> ... -> dZ_1.  Every dZ block
//      is a [128 x 64] bf16 swizzled image that is (a) the next UMMA's K-major A operand and (b) bulk-stored to HBM.
//   B. mlp_tc_wgrad_kernel   -- dW_l = dZ_l^T A_{l-1}: M = out features, N = in features, K = SAMPLES.  The very same
//      [samples x 64] block images are read back as MN-major UMMA operands (no transpose anywhere); each CTA owns a
//      contiguous range of tiles, accumulates one weight matrix at a time in TMEM (2 x 128 lanes x N columns) over
//      its whole range, flushes it once to a per-CTA partial, and spare warps column-sum dZ from shared memory for
//      the bias gradients.  The sigma / rgb heads ride along as two more items on a 4-column "head block".
//   C. reduce_partials_kernel -- grad += sum over CTAs of the partials (the gradient buffer is accumulated into).
#include "common.cuh"
#include "tc_common.cuh"

// -DNERFB_TRACE_BUILD compiles the clock64 timeline probes of the fused backward in (tools/trace_fused.py); the product build has none
#ifdef NERFB_TRACE_BUILD
#define NERFB_TRACE(...) __VA_ARGS__
#else
#define NERFB_TRACE(...)
#endif
using namespace tc;

TcFwdLayout tc_fwd_layout(const nerfb_mlp_desc& d, int64_t P);     // mlp_tc.cu
bool tc_supported(const nerfb_mlp_desc& d, int in_mode);

namespace {

constexpr int N_STAGES = 4;
constexpr int STAGE_BYTES = 256 * 128;
// dgrad warp roles (same plan as the forward kernel: mostly-waiting roles get the low warp ids)
constexpr int D_WARP_PRODUCER = 0, D_WARP_STORE = 1, D_WARP_EPI0 = 2, D_WARP_MMA = 18;
constexpr int NUM_THREADS = 19 * 32;
constexpr int D_EPI_WARPS = 16;
constexpr int MAX_BSTEPS = 12;

// ---- weight packing for the dgrad chain: slab (n, kk) = W[(row0 + kk) * ld + col0 + n]  (W^T, K-major) -----------
struct PackTSlab { int64_t w_off; int ld; int row0; int nrows; int col0; int64_t dst_off; };
struct PackTArgs { int n_slabs; PackTSlab slab[40]; };

__global__ void __launch_bounds__(256) pack_weights_t_kernel(const float* __restrict__ params, PackTArgs pa,
                                                             uint8_t* __restrict__ blob) {
  const PackTSlab sl = pa.slab[blockIdx.x >> 2];                         // four blocks per slab
  for (int idx = (blockIdx.x & 3) * blockDim.x + threadIdx.x; idx < 256 * 8; idx += 4 * blockDim.x) {   // n fastest across threads: coalesced reads
    const int n = idx & 255, ch = idx >> 8;
    float f[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int kk = ch * 8 + e;
      f[e] = (kk < sl.nrows) ? __ldg(params + sl.w_off + (int64_t)(sl.row0 + kk) * sl.ld + sl.col0 + n) : 0.f;
    }
    uint4 o;
    o.x = pack_bf16(f[0], f[1]); o.y = pack_bf16(f[2], f[3]); o.z = pack_bf16(f[4], f[5]); o.w = pack_bf16(f[6], f[7]);
    *reinterpret_cast<uint4*>(blob + sl.dst_off + (int64_t)n * 128 + ((ch ^ (n & 7)) << 4)) = o;
  }
}

// ---- dgrad chain program -------------------------------------------------------------------------------------------
struct BStep {
  int n_kb;          // K-blocks (64 upstream features each) of this step's A operand = blocks 0..n_kb-1
  int mask_off;      // forward-record offset of the gate bit-mask (128 rows x 32 B) of this step's output (-1: no gate)
  int add_sigma;     // output += d_sigma (x) w_sigma before gating (the sigma head reads the same activation)
  int save_off;      // backward-record offset of this step's output dZ (4 blocks)
  int add_cls;       // output += g_cls W_eff before storing (NeRF_3D class head reads xyz_encoding_final, nerf_cls.py:104)
  int act_off;       // SIREN: forward-record offset of the gated layer's saved sine image (4 blocks); film_off: its FiLM row * 256
  int film_off;
};
struct BProgram {
  int n_steps;
  BStep s[MAX_BSTEPS];
  int hd_off;        // forward-record offset of the direction layer's gate bit-mask, gate of dZ_dir
  int dzdir_off;     // backward-record offset of dZ_dir (2 blocks)
  int h_off;         // backward-record offset of the head block: columns 0..3 = [g_r g_g g_b d_sigma] per sample
  int fwd_tile_bytes, bwd_tile_bytes;
  int n_cls, out_c;  // NeRF_3D: class-head width; floats per output row (4 + n_cls)
  int siren;         // SIREN: gates are freq * cos(.) rebuilt from sign(cos) masks + the saved sine images, colour layer is 256 wide
  int n_pro;         // dZ blocks the prologue produces: 2 (NeRF direction layer, 128 wide) or 4 (SIREN colour layer)
  int rgb_in;        // fan-in of the rgb head: 128 / 256
  int hd_act_off, hd_film_off;   // SIREN: sine image / FiLM row of the colour layer (the prologue's gate)
};

struct DgradArgs {
  BProgram prog;
  const uint8_t* wblob;            // transposed bf16 slabs in schedule order
  const float* heads;              // w_sigma[256] | w_rgb[3*128]
  const float* weff;               // NeRF_3D: merged class-head matrix, fp32 [16 x 256] (forward workspace)
  const float* film;               // SIREN: [frequencies (D+1) x 256 | phase shifts], the forward's FiLM vector
  const float* out; const float* d_out;
  const uint8_t* fwd_rec; uint8_t* bwd_rec;
  int64_t P; int64_t n_tiles;
  long long* trace_chain;          // diagnostics (NERFB_TRACE_BUILD): store-lane timeline of chain CTA 0
};

// Fused backward (mlp_tc_bwd_fused_kernel): the dgrad chain CTAs hand every dZ packet (one step's 2 / 4 blocks) to the
// wgrad CTAs of the SAME launch through a small per-chain-CTA ring in global memory that stays resident in the 126 MB L2,
// so dZ is neither written to nor re-read from HBM.  Flags are monotonic counters in global memory (tc_common.cuh).
constexpr int F_HQ = 4;                    // head-block ring slots per chain CTA
constexpr int F_PKT_BYTES = 4 * BLK_BYTES; // one packet slot
constexpr int F_FLAG_STRIDE = 32;          // uint32 words between per-CTA counters (one 128 B line each)
constexpr int F_MAX_Q = 32;
struct FusedCtl {
  int Nc;                          // chain CTAs = blocks [0, Nc); wgrad CTAs = blocks [Nc, Nc + n_units * G)
  int G;                           // nominal wgrad CTAs per unit (NERFB_FUSED_G); the budget n_units * G is split by measured cost:
  int unit_G[16];                  // wgrad CTAs of unit u: CTA j of the unit owns tiles t = j (mod unit_G[u])
  int unit_w0[16];                 // first wgrad CTA (relative to Nc) of unit u
  int q;                           // packet ring slots per chain CTA
  int PK;                          // packets per tile: prologue + n_steps
  int m_head;                      // items that read the head block
  int pk_need[MAX_BSTEPS + 1];     // items that read packet i of a tile
  uint8_t* ring;                   // [Nc][q][F_PKT_BYTES]
  uint8_t* hring;                  // [Nc][F_HQ][2 * BLK_BYTES]
  uint32_t* produced;              // [Nc * F_FLAG_STRIDE]  packets published by chain CTA c
  uint32_t* hprod;                 // [Nc * F_FLAG_STRIDE]  head-block warps (4 per tile) done
  uint32_t* consumed;              // [Nc][q]      cumulative item reads of the slot
  uint32_t* hcons;                 // [Nc][F_HQ]
};

struct __align__(1024) SmemD {
  uint8_t g[4][BLK_BYTES];               // current dZ tile: 4 K-blocks
  uint8_t ring[N_STAGES][STAGE_BYTES];
  float w_sigma[256];
  float w_rgb[3 * 256];
  float w_cls[8 * 256];                  // NeRF_3D: W_eff rows
  uint64_t w_full[N_STAGES], w_empty[N_STAGES];
  uint64_t a_ready[4];                   // block j holds a freshly written dZ block (16 epilogue warps arrive)
  uint64_t g_free[4];                    // the store lane has read block j out to HBM
  uint64_t acc_full[2];
  uint32_t tmem_base;
  uint32_t slot_ok;                 // fused launch: ring packets whose slot the scout lane has seen free
};

}  // namespace

namespace {

// 16 sign bits of the forward pre-activations (element i at bit 15 - i, 1 = negative): zero the gated-off values
__device__ __forceinline__ void apply_gate16(float (&f)[16], uint32_t bits) {
#pragma unroll
  for (int i = 0; i < 16; ++i) f[i] = ((bits >> (15 - i)) & 1u) ? 0.f : f[i];
}
// SIREN gate freq * cos(freq z + phase) (models/nerf.py:149-151) for this thread's 16 columns of block j: sign(cos) from the
// forward's bit mask (bit 15 - i, 1 = negative), |cos| = sqrt(1 - sin^2) from the saved bf16 sine image (the same swizzled
// block the wgrad reads as its operand), freq from the FiLM vector.  The sine's bf16 rounding makes |cos| ill-conditioned
// only where |cos| < ~0.06 (absolute error <= 0.06 there, <= 2e-3 / |cos| elsewhere).
__device__ __forceinline__ void load_sine16(uint4 (&sv)[2], const uint8_t* act_blk, uint32_t row_off, uint32_t row_x, int c4) {
#pragma unroll
  for (int ch = 0; ch < 2; ++ch)
    sv[ch] = __ldg(reinterpret_cast<const uint4*>(act_blk + row_off + ((((uint32_t)(2 * c4 + ch)) ^ row_x) << 4)));
}
__device__ __forceinline__ void apply_sine_gate16(float (&f)[16], const uint4 (&sv)[2], uint32_t bits, const float* freq) {
  uint32_t sgn = bits << 16;                   // element i's bit (15 - i) moves to bit 31 as the loop shifts left
#pragma unroll
  for (int ch = 0; ch < 2; ++ch) {
    const float4 fa = __ldg(reinterpret_cast<const float4*>(freq + 8 * ch)), fb = __ldg(reinterpret_cast<const float4*>(freq + 8 * ch + 4));
    const uint32_t w[4] = {sv[ch].x, sv[ch].y, sv[ch].z, sv[ch].w};
    const float fr[8] = {fa.x, fa.y, fa.z, fa.w, fb.x, fb.y, fb.z, fb.w};
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float sn = __uint_as_float((e & 1) ? (w[e >> 1] & 0xFFFF0000u) : (w[e >> 1] << 16));
      float c;                                 // |sn| <= 1 (a rounded sine), so 1 - sn^2 >= 0
      asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(c) : "f"(fmaf(-sn, sn, 1.f)));
      const float g = fr[e] * c;
      f[8 * ch + e] *= __uint_as_float(__float_as_uint(g) ^ (sgn & 0x80000000u));
      sgn <<= 1;
    }
  }
}
// this thread's 16 columns (chunks 2*c4, 2*c4+1) of row `row` of a [128 x 64] swizzled bf16 block
__device__ __forceinline__ void store_block16(uint32_t blk, uint32_t row_off, uint32_t row_x, int c4, const float (&f)[16]) {
#pragma unroll
  for (int ch = 0; ch < 2; ++ch)
    sts128(blk + row_off + ((((uint32_t)(2 * c4 + ch)) ^ row_x) << 4), pack_bf16(f[8 * ch + 0], f[8 * ch + 1]),
           pack_bf16(f[8 * ch + 2], f[8 * ch + 3]), pack_bf16(f[8 * ch + 4], f[8 * ch + 5]), pack_bf16(f[8 * ch + 6], f[8 * ch + 7]));
}

// v2: 16 epilogue warps (32 rows x 16 columns of every 64-column block each, double-buffered TMEM loads), ReLU gates
// from the forward's bit-masks (8 B per thread per layer instead of the 64 KB activation image), the dZ stores on their
// own lane with lagged buffer release, and an MMA thread that only waits and issues.
// The chain role of one CTA: `cta` of `n_cta` chain CTAs walks tiles cta, cta + n_cta, ...  FUSED = false: the stand-alone
// dgrad kernel (dZ to the backward record in HBM, wgrad is a later launch).  FUSED = true: dZ packets go to this CTA's
// L2-resident ring and are published to the wgrad CTAs of the same grid (FusedCtl).
template <bool SIREN, bool FUSED>
__device__ __forceinline__ void dgrad_role(const DgradArgs& args, const FusedCtl& fc, uint8_t* smem_raw, const int cta, const int n_cta) {
  SmemD& sm = *reinterpret_cast<SmemD*>(smem_raw);
  if ((smem_u32(smem_raw) & 1023u) != 0u) __trap();
  const BProgram& pg = args.prog;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  for (int i = threadIdx.x; i < 256; i += NUM_THREADS) sm.w_sigma[i] = __ldg(args.heads + i);
  constexpr int RGB_IN = SIREN ? 256 : 128;
  for (int i = threadIdx.x; i < 3 * RGB_IN; i += NUM_THREADS) sm.w_rgb[i] = __ldg(args.heads + 256 + i);
  for (int i = threadIdx.x; i < pg.n_cls * 256; i += NUM_THREADS) sm.w_cls[i] = __ldg(args.weff + i);
  if (threadIdx.x == 0) {
    for (int i = 0; i < N_STAGES; ++i) { mbar_init(&sm.w_full[i], 1); mbar_init(&sm.w_empty[i], 1); }
    for (int i = 0; i < 4; ++i) { mbar_init(&sm.a_ready[i], D_EPI_WARPS); mbar_init(&sm.g_free[i], 1); }
    mbar_init(&sm.acc_full[0], 1); mbar_init(&sm.acc_full[1], 1);
    sm.slot_ok = 0u;
    fence_barrier_init();
  }
  if (warp == D_WARP_MMA) tmem_alloc(&sm.tmem_base, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp == D_WARP_PRODUCER) {
    if (lane == 0) {                       // ===== weight producer: W^T slabs, 32 KB each =====
      int stage = 0; uint32_t phase = 0;
      for (int64_t tile = cta; tile < args.n_tiles; tile += n_cta) {
        const uint8_t* src = args.wblob;
        for (int s = 0; s < pg.n_steps; ++s)
          for (int kb = 0; kb < pg.s[s].n_kb; ++kb) {
            mbar_wait_relaxed(&sm.w_empty[stage], phase ^ 1);
            mbar_expect_tx(&sm.w_full[stage], STAGE_BYTES);
            bulk_g2s(sm.ring[stage], src, STAGE_BYTES, &sm.w_full[stage]);
            src += STAGE_BYTES;
            if (++stage == N_STAGES) { stage = 0; phase ^= 1; }
          }
      }
    } else if (lane == 1 && FUSED) {       // ===== slot scout (fused): polls the consumers' counters AHEAD of the store lane =====
      // the L2 round trip + acquire of "slot n % q has been read out" costs the store lane ~1400 clk per packet on its critical
      // path (tools/trace_fused.py); here it overlaps the block stores and the store lane only polls shared memory
      uint32_t expect[F_MAX_Q];
      for (int i = 0; i < F_MAX_Q; ++i) expect[i] = 0;
      const uint32_t* const cons = fc.consumed + (size_t)cta * fc.q;
      const uint32_t ok = smem_u32(&sm.slot_ok);
      uint32_t slot = 0;
      for (int64_t tile = cta; tile < args.n_tiles; tile += n_cta)
        for (int s = -1; s < pg.n_steps; ++s) {
          gflag_wait_ge(cons + slot, expect[slot]);
          expect[slot] += (uint32_t)fc.pk_need[s + 1];
          cnt_publish(ok);
          if (++slot == (uint32_t)fc.q) slot = 0;
        }
    }
  } else if (warp == D_WARP_MMA) {
    if (lane == 0) {                       // ===== MMA issuer =====
      uint32_t stage = 0, phase = 0, a_par = 0, gs = 0;
      const uint32_t idesc = umma_idesc_bf16(TILE_M, 256);
      const uint32_t g_lo0 = ((smem_u32(sm.g[0]) >> 4) & 0x3FFFu) | (1u << 16);
      const uint32_t ring_lo0 = ((smem_u32(sm.ring[0]) >> 4) & 0x3FFFu) | (1u << 16);
      const uint32_t ar0 = smem_u32(&sm.a_ready[0]), wf0 = smem_u32(&sm.w_full[0]), we0 = smem_u32(&sm.w_empty[0]);
      const uint32_t af0 = smem_u32(&sm.acc_full[0]);
      for (int64_t tile = cta; tile < args.n_tiles; tile += n_cta) {
        for (int s = 0; s < pg.n_steps; ++s, ++gs) {
          const int n_kb = pg.s[s].n_kb;
          for (int kb = 0; kb < n_kb; ++kb) {
            const bool ok_a = mbar_try_wait_addr(ar0 + 8u * kb, (a_par >> kb) & 1u);
            const bool ok_w = mbar_try_wait_addr(wf0 + 8u * stage, phase);
            if (!ok_a) mbar_wait_addr(ar0 + 8u * kb, (a_par >> kb) & 1u);
            a_par ^= 1u << kb;
            if (!ok_w) mbar_wait_addr(wf0 + 8u * stage, phase);
            tc_fence_after();
            umma_bf16_k64(tmem + (gs & 1u) * 256u, g_lo0 + (uint32_t)kb * (BLK_BYTES >> 4), ring_lo0 + stage * (STAGE_BYTES >> 4), idesc,
                          kb > 0 ? 1u : 0u);
            umma_commit_addr(we0 + 8u * stage);
            if (++stage == N_STAGES) { stage = 0; phase ^= 1; }
          }
          umma_commit_addr(af0 + 8u * (gs & 1u));
        }
        for (int j = 0; j < 4; ++j) {           // the last step's output blocks feed no MMA: keep the phase count in step
          mbar_wait_addr(ar0 + 8u * j, (a_par >> j) & 1u); a_par ^= 1u << j;
        }
      }
    }
  } else if (warp == D_WARP_STORE) {
    if (lane == 0 && !FUSED) {             // ===== store lane: every published dZ block -> HBM for the wgrad =====
      uint32_t a_par = 0, pend = 0;
      auto issued = [&](uint32_t bar) {       // store k committed: release the buffer of store k-1 once it has been read
        bulk_commit(); bulk_wait_read1();
        if (pend) mbar_arrive_addr(pend);
        pend = bar;
      };
      for (int64_t tile = cta; tile < args.n_tiles; tile += n_cta) {
        uint8_t* rec = args.bwd_rec + tile * (int64_t)pg.bwd_tile_bytes;
        for (int s = -1; s < pg.n_steps; ++s) {
          const int nb = s < 0 ? pg.n_pro : 4;
          const int off = s < 0 ? pg.dzdir_off : pg.s[s].save_off;
          for (int j = 0; j < nb; ++j) {
            mbar_wait_relaxed(&sm.a_ready[j], (a_par >> j) & 1u); a_par ^= 1u << j;
            bulk_s2g(rec + off + j * BLK_BYTES, sm.g[j], BLK_BYTES);
            issued(smem_u32(&sm.g_free[j]));
          }
        }
      }
      bulk_wait_read0();
      if (pend) mbar_arrive_addr(pend);
      bulk_wait_all0();
    }
    if (lane == 0 && FUSED) {
      // ===== store lane, fused: packet n of this CTA (tile-local order: prologue, step 0, ...) -> ring slot n % q.  Before
      // the first block of a packet the slot's previous occupant (packet n - q) must have been read by all its items;
      // a packet is PUBLISHED (produced[cta] = n + 1, release) once its stores have completed, which is checked lazily --
      // F_LAG bulk groups later -- so the lane never sits on the ~1 us store latency.  q >= 3 keeps the wait chain
      // well-founded: the packet whose slot is awaited was published at least one packet ago.
      constexpr uint32_t F_LAG = 7;            // 3 -> 7: with 3 the lane waited ~2300 clk per packet for the store issued 3 blocks earlier
      uint32_t a_par = 0, pend = 0;
      uint32_t n = 0, grp = 0, pub = 0;          // packets fully issued, bulk groups committed, packets published
      uint32_t last_grp[8] = {0, 0, 0, 0, 0, 0, 0, 0};       // index of the last bulk group of packet (n & 7)
      uint32_t expect[F_MAX_Q];
      for (int i = 0; i < F_MAX_Q; ++i) expect[i] = 0;
      uint32_t* const prod = fc.produced + (size_t)cta * F_FLAG_STRIDE;
      const uint32_t* const cons = fc.consumed + (size_t)cta * fc.q;
      uint8_t* const ring = fc.ring + (size_t)cta * fc.q * F_PKT_BYTES;
      uint32_t slot = 0;
      for (int64_t tile = cta; tile < args.n_tiles; tile += n_cta) {
        for (int s = -1; s < pg.n_steps; ++s) {
          const int nb = s < 0 ? pg.n_pro : 4;
          NERFB_TRACE(long long* trc = args.trace_chain; const bool tr = trc && cta == 0 && n < 150; long long tpub = 0, twa = 0; if (tr) trc[n * 8 + 0] = clock64();)
          if ((int32_t)(cnt_load(smem_u32(&sm.slot_ok)) - (n + 1)) < 0) {
            // About to block on a ring slot.  The lagged publication below would leave the packets already stored UNPUBLISHED
            // for the whole stall, and the wgrad CTA waiting for one of them is the very one whose progress frees slots
            // elsewhere: flush and publish first.
            if (pub < n) { bulk_wait_all0(); fence_proxy_async_all(); pub = n; st_release_gpu(prod, pub); }
            cnt_wait(smem_u32(&sm.slot_ok), n + 1);
          }
          NERFB_TRACE(if (tr) trc[n * 8 + 1] = clock64(); const uint32_t n_here = n;)
          uint8_t* dst = ring + (size_t)slot * F_PKT_BYTES;
          for (int j = 0; j < nb; ++j) {
            NERFB_TRACE(const long long ta0 = clock64();)
            mbar_wait_relaxed(&sm.a_ready[j], (a_par >> j) & 1u); a_par ^= 1u << j;
            NERFB_TRACE(twa += clock64() - ta0;)
            bulk_s2g(dst + j * BLK_BYTES, sm.g[j], BLK_BYTES);
            bulk_commit(); ++grp; bulk_wait_read1();
            if (pend) mbar_arrive_addr(pend);
            pend = smem_u32(&sm.g_free[j]);
            if (j == nb - 1) { last_grp[n & 7u] = grp - 1; ++n; }
            if (pub < n && grp - 1 - last_grp[pub & 7u] >= F_LAG) {      // the oldest unpublished packet's stores are >= F_LAG groups back
              NERFB_TRACE(const long long tp0 = clock64();)
              bulk_wait_all7();
              ++pub; st_release_gpu(prod, pub);        // the stores are complete (wait_group): the flag's release orders them; no proxy fence on this path
              NERFB_TRACE(tpub += clock64() - tp0;)
            }
          }
          NERFB_TRACE(if (tr) { trc[n_here * 8 + 2] = twa; trc[n_here * 8 + 3] = tpub; trc[n_here * 8 + 4] = clock64(); })
          if (++slot == (uint32_t)fc.q) slot = 0;
        }
      }
      bulk_wait_read0();
      if (pend) mbar_arrive_addr(pend);
      bulk_wait_all0();
      fence_proxy_async_all();
      st_release_gpu(prod, n);
    }
  } else if (warp >= D_WARP_EPI0 && warp < D_WARP_MMA) {
    // ===== prologue (rgb head + direction-layer gate) and per-step epilogues =====
    const int ew = warp - D_WARP_EPI0, q = warp & 3, c4 = ew >> 2;
    const int row = q * 32 + lane;
    const uint32_t row_off = (uint32_t)row * 128u, row_x = (uint32_t)row & 7u;
    const uint32_t t_lane = tmem + ((uint32_t)(q * 32) << 16) + 16u * (uint32_t)c4;
    const uint32_t g0_addr = smem_u32(sm.g[0]);
    uint32_t acc_par = 0, gf_par = 0, need_free = 0;
    uint32_t gs = 0, tile_k = 0;               // tile_k: this CTA's local tile counter
    for (int64_t tile = cta; tile < args.n_tiles; tile += n_cta, ++tile_k) {
      const int64_t p = tile * TILE_M + row;
      const bool live = p < args.P;
      const uint8_t* frec = args.fwd_rec + tile * (int64_t)pg.fwd_tile_bytes + c4 * (TILE_M * 8) + row * 8;   // mask layout [c4][row]
      float g0 = 0.f, g1 = 0.f, g2 = 0.f, dsig = 0.f;
      float gc[8];                            // class-head logit gradients d_cls * s (1 - s)
#pragma unroll
      for (int k = 0; k < 8; ++k) gc[k] = 0.f;
      if (live) {
        float4 y, dy;
        if (pg.out_c == 4) {
          y = __ldg(reinterpret_cast<const float4*>(args.out + p * 4));
          dy = __ldg(reinterpret_cast<const float4*>(args.d_out + p * 4));
        } else {
          const float* yo = args.out + p * pg.out_c; const float* dyo = args.d_out + p * pg.out_c;
          y = make_float4(__ldg(yo), __ldg(yo + 1), __ldg(yo + 2), __ldg(yo + 3));
          dy = make_float4(__ldg(dyo), __ldg(dyo + 1), __ldg(dyo + 2), __ldg(dyo + 3));
#pragma unroll
          for (int k = 0; k < 8; ++k)
            if (k < pg.n_cls) { const float sv = __ldg(yo + 4 + k); gc[k] = __ldg(dyo + 4 + k) * sv * (1.f - sv); }
        }
        g0 = dy.x * y.x * (1.f - y.x); g1 = dy.y * y.y * (1.f - y.y); g2 = dy.z * y.z * (1.f - y.z);   // sigmoid'
        dsig = dy.w;
      }
      if (c4 == 0) {                          // head block for the wgrad kernel: [g_rgb d_sigma g_cls(8) 0 0 0 0] bf16 per sample
        uint8_t* hrow;
        if (FUSED) {                          // this CTA's head ring: slot k % F_HQ, free once the head items of tile k - F_HQ have read it
          const uint32_t hs = tile_k % F_HQ;
          if (lane == 0) gflag_wait_ge(fc.hcons + (size_t)cta * F_HQ + hs, (uint32_t)fc.m_head * (tile_k / F_HQ));
          __syncwarp();
          hrow = fc.hring + ((size_t)cta * F_HQ + hs) * (2 * BLK_BYTES);
        } else {
          hrow = args.bwd_rec + tile * (int64_t)pg.bwd_tile_bytes + pg.h_off;
        }
        *reinterpret_cast<uint4*>(hrow + swz_off(row, 0)) =
            make_uint4(pack_bf16(g0, g1), pack_bf16(g2, dsig), pack_bf16(gc[0], gc[1]), pack_bf16(gc[2], gc[3]));
        if (pg.n_cls > 0)
          *reinterpret_cast<uint4*>(hrow + swz_off(row, 8)) = make_uint4(pack_bf16(gc[4], gc[5]), pack_bf16(gc[6], gc[7]), 0u, 0u);
        if (FUSED) {                          // generic-proxy global writes of this warp -> visible to the consumers' bulk reads
          __threadfence();
          fence_proxy_async_all();
          __syncwarp();
          if (lane == 0) red_release_gpu_add(fc.hprod + (size_t)cta * F_FLAG_STRIDE, 1u);
        }
      }
      const uint8_t* arec = args.fwd_rec + tile * (int64_t)pg.fwd_tile_bytes;      // SIREN: saved sine images of this tile
      uint2 mk = __ldg(reinterpret_cast<const uint2*>(frec + pg.hd_off));
      uint4 sv[2][2];                         // SIREN: this thread's 16 saved sines of the current / next block (one block ahead)
      if (SIREN) load_sine16(sv[0], arec + pg.hd_act_off, row_off, row_x, c4);
#pragma unroll
      for (int j = 0; j < (SIREN ? 4 : 2); ++j) {   // dZ_dir = (g . W_rgb) (.) gate     (models/nerf.py:119-120 / :213-215)
        float f[16];
        const int col0 = 64 * j + 16 * c4;
        if (SIREN && j + 1 < 4) load_sine16(sv[(j + 1) & 1], arec + pg.hd_act_off + (j + 1) * BLK_BYTES, row_off, row_x, c4);
#pragma unroll
        for (int i = 0; i < 16; ++i)
          f[i] = g0 * sm.w_rgb[col0 + i] + g1 * sm.w_rgb[RGB_IN + col0 + i] + g2 * sm.w_rgb[2 * RGB_IN + col0 + i];
        if (SIREN) apply_sine_gate16(f, sv[j & 1], ((j < 2 ? mk.x : mk.y) >> (16 * (1 - (j & 1)))) & 0xFFFFu, args.film + pg.hd_film_off + col0);
        else apply_gate16(f, (mk.x >> (16 * (1 - j))) & 0xFFFFu);      // block 0 was shifted in first: upper half
        if ((need_free >> j) & 1u) { mbar_wait(&sm.g_free[j], (gf_par >> j) & 1u); gf_par ^= 1u << j; }
        need_free |= 1u << j;
        store_block16(g0_addr + (uint32_t)j * BLK_BYTES, row_off, row_x, c4, f);
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.a_ready[j]);
      }
      for (int s = 0; s < pg.n_steps; ++s, ++gs) {
        const BStep& st = pg.s[s];
        const uint32_t a = gs & 1u;
        const bool gated = st.mask_off >= 0;
        if (gated) mk = __ldg(reinterpret_cast<const uint2*>(frec + st.mask_off));
        if (gated && SIREN) load_sine16(sv[0], arec + st.act_off, row_off, row_x, c4);
        const bool add_sigma = st.add_sigma != 0;
        mbar_wait(&sm.acc_full[a], (acc_par >> a) & 1u); acc_par ^= 1u << a;
        tc_fence_after();
        uint32_t v[2][16];
        tmem_ld16(t_lane + a * 256u, v[0]);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int col0 = 64 * j + 16 * c4;
          tmem_ld_wait();
          if (j + 1 < 4) tmem_ld16(t_lane + a * 256u + 64u * (uint32_t)(j + 1), v[(j + 1) & 1]);
          if (gated && SIREN && j + 1 < 4) load_sine16(sv[(j + 1) & 1], arec + st.act_off + (j + 1) * BLK_BYTES, row_off, row_x, c4);
          float f[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) f[i] = __uint_as_float(v[j & 1][i]);
          if (add_sigma) {                    // sigma = <h8, w_sigma> + b reads the same activation (models/nerf.py:112)
            const uint32_t wb = smem_u32(sm.w_sigma) + (uint32_t)col0 * 4u;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const float4 w = lds128f(wb + 16u * i);
              f[4 * i] = fmaf(dsig, w.x, f[4 * i]); f[4 * i + 1] = fmaf(dsig, w.y, f[4 * i + 1]);
              f[4 * i + 2] = fmaf(dsig, w.z, f[4 * i + 2]); f[4 * i + 3] = fmaf(dsig, w.w, f[4 * i + 3]);
            }
          }
          if (st.add_cls) {                   // d feat += g_cls W_eff
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              if (k < pg.n_cls) {
                const uint32_t wb = smem_u32(sm.w_cls) + (uint32_t)(k * 256 + col0) * 4u;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                  const float4 w = lds128f(wb + 16u * i);
                  f[4 * i] = fmaf(gc[k], w.x, f[4 * i]); f[4 * i + 1] = fmaf(gc[k], w.y, f[4 * i + 1]);
                  f[4 * i + 2] = fmaf(gc[k], w.z, f[4 * i + 2]); f[4 * i + 3] = fmaf(gc[k], w.w, f[4 * i + 3]);
                }
              }
            }
          }
          if (gated && SIREN) apply_sine_gate16(f, sv[j & 1], ((j < 2 ? mk.x : mk.y) >> (16 * (1 - (j & 1)))) & 0xFFFFu, args.film + st.film_off + col0);
          else if (gated) apply_gate16(f, ((j < 2 ? mk.x : mk.y) >> (16 * (1 - (j & 1)))) & 0xFFFFu);
          if ((need_free >> j) & 1u) { mbar_wait(&sm.g_free[j], (gf_par >> j) & 1u); gf_par ^= 1u << j; }
          need_free |= 1u << j;
          store_block16(g0_addr + (uint32_t)j * BLK_BYTES, row_off, row_x, c4, f);
          fence_proxy_async();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&sm.a_ready[j]);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == D_WARP_MMA) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

template <bool SIREN>
__global__ void __launch_bounds__(NUM_THREADS, 1) mlp_tc_dgrad_kernel(const __grid_constant__ DgradArgs args) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  dgrad_role<SIREN, false>(args, FusedCtl{}, smem_raw, (int)blockIdx.x, (int)gridDim.x);
}

}  // namespace

// =====================================================================================================================
// B. wgrad: dW = dZ^T A over samples, MN-major operands straight from the saved tile images
// =====================================================================================================================
namespace {

constexpr int W_STAGES = 3;
constexpr int W_HALF = 8192;                 // 64 sample rows of one [128 x 64] block
constexpr int W_STAGE_BYTES = 8 * W_HALF;    // dZ blocks [0,4) | input blocks [4,8)
constexpr int W_THREADS = 192;               // warp 0 producer, warp 1 MMA, warps 2..5 bias sums + accumulator flush
constexpr int MAX_ITEMS = 16;

struct WItem {
  int dz_off, nM;            // backward-record offset of dZ, number of 64-column blocks (4: M = 256, 2: M = 128)
  int a_fwd, a_off, nN;      // input activation: record (1 = forward record, 0 = backward record), offset, blocks
  int N;                     // UMMA N = 64 * nN
  int64_t w_off; int ld, col0, nvalid;   // destination: grad[w_off + out * ld + col0 + n], n < nvalid
  int64_t bias_off;          // destination of colsum(dZ), or -1
  int m_lo, m_hi;            // rows (out features) of the accumulator that are real: [m_lo, m_hi)
  int64_t bias_alt;          // head block only (>= 0): colsum of column 3 goes here (b_sigma), columns 0..2 to bias_off
  int aux;                   // destination is the aux buffer (class-head dW_eff | column sums), not the gradient blob
  int hilo;                  // SIREN 3-wide inputs ride as [x_hi x_lo x_hi]: dW[:, k] = column k + column 3 + k
};
struct WgradArgs {
  int n_items; WItem item[MAX_ITEMS];
  int64_t item_off[MAX_ITEMS + 1];     // prefix sums of cost_i * n_tiles: the linearised (item, tile) work space
  const uint8_t* fwd_rec; const uint8_t* bwd_rec; int fwd_tile_bytes, bwd_tile_bytes;
  int64_t n_tiles; int n_active;
  float* partial;                      // [n_active + n_items slots][256 x 256] dense per-item partials
  float* grad;
  float* aux;                          // NeRF_3D: dW_eff [16 x 256] | colsum of the head block [128]
};
constexpr int64_t W_SLOT = 256 * 256;  // floats per partial slot

// Work split: the (item, tile) space is linearised with each item weighted by the bytes it streams per tile
// (cost = nM + nN blocks) and cut into n_active equal pieces, so a CTA owns ONE long K-range of one weight matrix
// (sometimes the tail of one and the head of the next) instead of a short range of all 16: 148 x 16 accumulator
// flushes and 352 MB of partials become ~160 flushes and 40 MB.  Tiles [t0, t1) of item i that CTA c owns:
__host__ __device__ inline int item_cost(const WItem& im) { return im.nM + im.nN; }
__host__ __device__ inline int64_t seg_tile(int64_t x, int64_t off, int cost, int64_t n_tiles) {
  if (x <= off) return 0;
  const int64_t t = (x - off + cost - 1) / cost;
  return t < n_tiles ? t : n_tiles;
}
__host__ __device__ inline void seg_range(const WgradArgs& a, int c, int i, int64_t& t0, int64_t& t1) {
  const int64_t total = a.item_off[a.n_items];
  const int64_t lo = total * c / a.n_active, hi = total * (c + 1) / a.n_active;
  const int cost = item_cost(a.item[i]);
  t0 = seg_tile(lo, a.item_off[i], cost, a.n_tiles);
  t1 = seg_tile(hi, a.item_off[i], cost, a.n_tiles);
}

struct __align__(1024) SmemW {
  uint8_t stage[W_STAGES][W_STAGE_BYTES];
  uint64_t full[W_STAGES], empty[W_STAGES];
  uint64_t acc_done, acc_free;
  uint32_t tmem_base;
  uint32_t scout_cnt;                // fused launch: (tile, item) pairs whose dZ packet the scout lane has seen published
};

__global__ void __launch_bounds__(W_THREADS, 1) mlp_tc_wgrad_kernel(const __grid_constant__ WgradArgs args) {
  extern __shared__ uint8_t smem_raw[];
  SmemW& sm = *reinterpret_cast<SmemW*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // the items this CTA touches: [it_lo, it_hi]
  int it_lo = 0, it_hi = -1;
  {
    const int64_t total = args.item_off[args.n_items];
    const int64_t lo = total * blockIdx.x / args.n_active, hi = total * (blockIdx.x + 1) / args.n_active;
    while (it_lo + 1 < args.n_items && args.item_off[it_lo + 1] <= lo) ++it_lo;
    it_hi = it_lo;
    while (it_hi + 1 < args.n_items && args.item_off[it_hi + 1] < hi) ++it_hi;
  }

  if (threadIdx.x == 0) {
    for (int i = 0; i < W_STAGES; ++i) { mbar_init(&sm.full[i], 1); mbar_init(&sm.empty[i], 5); }
    mbar_init(&sm.acc_done, 1); mbar_init(&sm.acc_free, 4);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&sm.tmem_base, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp == 0) {
    if (lane == 0) {                         // ===== producer: 64-sample halves of the dZ and input blocks =====
      int st = 0; uint32_t ph = 0;
      for (int it = it_lo; it <= it_hi; ++it) {
        const WItem& im = args.item[it];
        int64_t t0, t1; seg_range(args, blockIdx.x, it, t0, t1);
        const int n_chunks = (int)(t1 - t0) * 2;
        const uint8_t* abase = im.a_fwd ? args.fwd_rec : args.bwd_rec;
        const int64_t astride = im.a_fwd ? args.fwd_tile_bytes : args.bwd_tile_bytes;
        for (int c = 0; c < n_chunks; ++c) {
          const int64_t tile = t0 + (c >> 1); const int half = c & 1;
          mbar_wait(&sm.empty[st], ph ^ 1);
          mbar_expect_tx(&sm.full[st], (uint32_t)(im.nM + im.nN) * W_HALF);
          const uint8_t* dz = args.bwd_rec + tile * (int64_t)args.bwd_tile_bytes + im.dz_off + half * W_HALF;
          for (int j = 0; j < im.nM; ++j) bulk_g2s(sm.stage[st] + j * W_HALF, dz + j * BLK_BYTES, W_HALF, &sm.full[st]);
          const uint8_t* a = abase + tile * astride + im.a_off + half * W_HALF;
          for (int j = 0; j < im.nN; ++j) bulk_g2s(sm.stage[st] + (4 + j) * W_HALF, a + j * BLK_BYTES, W_HALF, &sm.full[st]);
          if (++st == W_STAGES) { st = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {                         // ===== MMA issuer =====
      int st = 0; uint32_t ph = 0, free_ph = 0;
      for (int it = it_lo; it <= it_hi; ++it) {
        const WItem& im = args.item[it];
        int64_t t0, t1; seg_range(args, blockIdx.x, it, t0, t1);
        const int n_chunks = (int)(t1 - t0) * 2;
        const uint32_t idesc = umma_idesc_bf16(128, im.N, 1, 1);
        if (it > it_lo) { mbar_wait(&sm.acc_free, free_ph); free_ph ^= 1; tc_fence_after(); }
        for (int c = 0; c < n_chunks; ++c) {
          mbar_wait(&sm.full[st], ph);
          tc_fence_after();
          const uint32_t base = smem_u32(sm.stage[st]);
#pragma unroll
          for (int ks = 0; ks < 4; ++ks)       // 16 samples per UMMA
            for (int mh = 0; mh < (im.nM >> 1); ++mh)
              umma_bf16(tmem + mh * 256, umma_desc_mn128(base + mh * 2 * W_HALF + ks * 2048, W_HALF),
                        umma_desc_mn128(base + 4 * W_HALF + ks * 2048, W_HALF), idesc, (c > 0 || ks > 0) ? 1u : 0u);
          umma_commit(&sm.empty[st]);
          if (++st == W_STAGES) { st = 0; ph ^= 1; }
        }
        umma_commit(&sm.acc_done);
      }
    }
  } else {
    // ===== bias column sums (from shared memory) + accumulator flush =====
    const int q = warp & 3, bt = (warp - 2) * 32 + lane;
    const int blk = bt >> 5, c2 = bt & 31;
    int st = 0; uint32_t ph = 0, done_ph = 0;
    for (int it = it_lo; it <= it_hi; ++it) {
      const WItem& im = args.item[it];
      int64_t t0, t1; seg_range(args, blockIdx.x, it, t0, t1);
      const int n_chunks = (int)(t1 - t0) * 2;
      float* part = args.partial + (int64_t)(blockIdx.x + it) * W_SLOT;      // slot (c, i) -> c + i is unique
      float b0 = 0.f, b1 = 0.f;
      const bool sum_here = im.bias_off >= 0 && blk < im.nM;
      for (int c = 0; c < n_chunks; ++c) {
        mbar_wait(&sm.full[st], ph);
        if (sum_here) {
          const uint8_t* src = sm.stage[st] + blk * W_HALF + ((c2 & 3) << 2);
#pragma unroll 8
          for (int r = 0; r < 64; ++r) {
            const uint32_t w = *reinterpret_cast<const uint32_t*>(src + r * 128 + ((((uint32_t)c2 >> 2) ^ ((uint32_t)r & 7u)) << 4));
            b0 += __uint_as_float(w << 16); b1 += __uint_as_float(w & 0xFFFF0000u);
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.empty[st]);
        if (++st == W_STAGES) { st = 0; ph ^= 1; }
      }
      if (sum_here && n_chunks > 0) {
        if (im.bias_alt < 0) {
          float* bdst = (im.aux ? args.aux : args.grad) + im.bias_off;
          atomicAdd(bdst + 2 * bt, b0);
          atomicAdd(bdst + 2 * bt + 1, b1);
        } else if (bt < 2) {               // head block: columns [g_r g_g g_b d_sigma]
          atomicAdd(args.grad + im.bias_off + 2 * bt, b0);
          atomicAdd(bt == 0 ? args.grad + im.bias_off + 1 : args.grad + im.bias_alt, b1);
        }
      }
      mbar_wait(&sm.acc_done, done_ph); done_ph ^= 1;
      tc_fence_after();
      for (int mh = 0; mh < (im.nM >> 1); ++mh) {
        const int m = mh * 128 + q * 32 + lane;
        if (mh * 128 + q * 32 >= im.m_hi || mh * 128 + q * 32 + 32 <= im.m_lo) continue;      // warp-uniform
        float* dst = part + (int64_t)m * 256;
        const bool real = m >= im.m_lo && m < im.m_hi && n_chunks > 0;
        for (int c0 = 0; c0 < im.N; c0 += 32) {
          uint32_t v[32];
          tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(mh * 256 + c0), v);
          tmem_ld_wait();
          if (real && c0 < im.nvalid) {
#pragma unroll
            for (int i = 0; i < 32; i += 4)
              *reinterpret_cast<float4*>(dst + c0 + i) = make_float4(__uint_as_float(v[i]), __uint_as_float(v[i + 1]),
                                                                    __uint_as_float(v[i + 2]), __uint_as_float(v[i + 3]));
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm.acc_free);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

// ---- fused backward: the wgrad role ---------------------------------------------------------------------------------
// A "unit" is a set of <= 4 items whose accumulators fit the 512 TMEM columns together (one 256 x 256 matrix, or e.g.
// {direction layer, sigma head}); G CTAs serve a unit, CTA j taking tiles t = j (mod G) in increasing order and, inside a
// tile, its items in the order the chain produces their dZ packets (head block first).  Every wait therefore points at a
// lexicographically smaller (tile, packet), which is what makes the chain <-> wgrad hand-off deadlock-free.
constexpr int F_MAX_UNITS = 16;
constexpr int F_UNIT_ITEMS = 4;
struct FusedWArgs {
  int n_items; WItem item[MAX_ITEMS];
  int item_pk[MAX_ITEMS];                    // packet index inside a tile (0 = prologue, s + 1 = step s), -1 = head block
  int item_col[MAX_ITEMS][2];                // TMEM column of the accumulator half mh
  int item_pslot[MAX_ITEMS];                 // partial slot of wgrad CTA w: item_pslot[i] + w_in_unit * unit_n[u]  (+ unit_pbase[u])
  int item_unit[MAX_ITEMS];
  int n_units; int unit_n[F_MAX_UNITS]; int unit_item[F_MAX_UNITS][F_UNIT_ITEMS]; int unit_pbase[F_MAX_UNITS];
  const uint8_t* fwd_rec; int fwd_tile_bytes;
  int64_t n_tiles;
  float* partial; float* grad; float* aux;
};

__device__ __forceinline__ void wgrad_fused_role(const FusedWArgs& args, const FusedCtl& fc, uint8_t* smem_raw, long long* trace) {
  (void)trace;
  SmemW& sm = *reinterpret_cast<SmemW*>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int w = (int)blockIdx.x - fc.Nc;
  int u = 0;
  while (u + 1 < args.n_units && w >= fc.unit_w0[u + 1]) ++u;
  const int G = fc.unit_G[u], jg = w - fc.unit_w0[u];
  const int n_it = args.unit_n[u];
  const int my_tiles = (int)((args.n_tiles - jg + G - 1) / G);             // tiles jg, jg + G, ... (n_tiles >= G is guaranteed by the host)

  if (threadIdx.x == 0) {
    for (int i = 0; i < W_STAGES; ++i) { mbar_init(&sm.full[i], 1); mbar_init(&sm.empty[i], 5); }
    mbar_init(&sm.acc_done, 1); mbar_init(&sm.acc_free, 4);
    sm.scout_cnt = 0u;
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&sm.tmem_base, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp == 0) {
    {                                        // ===== producer: dZ halves from the chain CTAs' rings (L2), input halves from the forward record (HBM) =====
      // The per-ITEM overhead of this lane, not the MMAs, bounded the fused launch (tools/trace_fused.py: ~900 clk flag round
      // trip + ~870 clk fence + ~2200 clk for 16 sequential copies per item; a unit of four items then needs 4 x 4600 clk
      // per tile).  So: the chain CTA's packet counter is re-read only when the cached value does not already cover the
      // packet (one read usually covers every item of a tile), the proxy fence follows only a fresh read, and lanes 0..7
      // issue one 8 KB copy each (dZ blocks on lanes 0..3, input blocks on 4..7) after lane 0 has posted the byte count.
      // ... and the flag polling has moved to a SCOUT lane (warp 6) that runs ahead through the same (tile, item) sequence and
      // counts the packets it has seen into shared memory; this lane only waits on that counter, so the L2 round trip and
      // the proxy fence of item i + 1 overlap the stage wait and the copies of item i.
      int st = 0; uint32_t ph = 0, n_seen = 0;
      const uint64_t stream_once = l2_policy_evict_first();
      NERFB_TRACE(long long acc_flag = 0, acc_empty = 0; const long long t_begin = clock64();)
      int tc_c = jg % fc.Nc; uint32_t tc_k = (uint32_t)(jg / fc.Nc);
      int it_nM[F_UNIT_ITEMS], it_nN[F_UNIT_ITEMS], it_aoff[F_UNIT_ITEMS], it_pk[F_UNIT_ITEMS];     // the unit's items, out of the parameter bank once
#pragma unroll
      for (int li = 0; li < F_UNIT_ITEMS; ++li) {
        const int it = args.unit_item[u][li < n_it ? li : 0];
        it_nM[li] = args.item[it].nM; it_nN[li] = args.item[it].nN; it_aoff[li] = args.item[it].a_off; it_pk[li] = args.item_pk[it];
      }
      const uint32_t qq = (uint32_t)fc.q, PKn = (uint32_t)fc.PK;
      for (int ti = 0; ti < my_tiles; ++ti) {
        const int64_t tile = jg + (int64_t)ti * G;
        const int c = tc_c; const uint32_t k = tc_k;      // tile = c + k * Nc, tracked incrementally (no 64-bit div / mod per tile)
        tc_c += G; while (tc_c >= fc.Nc) { tc_c -= fc.Nc; ++tc_k; }
        NERFB_TRACE(const bool tr = trace && w == 0 && ti < 120 && lane == 0; if (tr) trace[ti * 8 + 0] = clock64();)
        const uint8_t* const arec = args.fwd_rec + tile * (int64_t)args.fwd_tile_bytes;
#pragma unroll
        for (int li = 0; li < F_UNIT_ITEMS; ++li) {
          if (li >= n_it) break;
          struct { int nM, nN; } im = {it_nM[li], it_nN[li]};
          const int pk = it_pk[li];
          const uint8_t* dzp;
          if (pk >= 0) {
            const uint32_t n = k * PKn + (uint32_t)pk;
            dzp = fc.ring + ((size_t)((uint32_t)c * qq + n % qq)) * F_PKT_BYTES;
          } else {
            dzp = fc.hring + ((size_t)((uint32_t)c * F_HQ + k % F_HQ)) * (2 * BLK_BYTES);
          }
          ++n_seen;
          if (lane == 0) {
            NERFB_TRACE(const long long tf0 = clock64();)
            cnt_wait(smem_u32(&sm.scout_cnt), n_seen);
            NERFB_TRACE(acc_flag += clock64() - tf0;)
          }
          NERFB_TRACE(if (tr && li == 0) { trace[ti * 8 + 1] = clock64(); trace[ti * 8 + 2] = trace[ti * 8 + 1]; })
          const uint8_t* ap = arec + it_aoff[li];
          const bool mine = lane < 4 ? lane < im.nM : (lane < 8 && lane - 4 < im.nN);
          for (int half = 0; half < 2; ++half) {
            if (lane == 0) {
              NERFB_TRACE(const long long te0 = clock64();)
              mbar_wait(&sm.empty[st], ph ^ 1);
              NERFB_TRACE(acc_empty += clock64() - te0;)
              mbar_expect_tx(&sm.full[st], (uint32_t)(im.nM + im.nN) * W_HALF);
            }
            NERFB_TRACE(if (tr && li == 0 && half == 0) trace[ti * 8 + 3] = clock64();)
            __syncwarp();                      // packet seen + byte count posted before any lane's copy
            if (mine) {
              if (lane < 4) bulk_g2s(sm.stage[st] + lane * W_HALF, dzp + lane * BLK_BYTES + half * W_HALF, W_HALF, &sm.full[st]);
              else bulk_g2s_hint(sm.stage[st] + lane * W_HALF, ap + (lane - 4) * BLK_BYTES + half * W_HALF, W_HALF, &sm.full[st], stream_once);
            }
            if (++st == W_STAGES) { st = 0; ph ^= 1; }
          }
          NERFB_TRACE(if (tr && li == 0) trace[ti * 8 + 4] = clock64();)
        }
      }
      // per-CTA totals behind the two timelines: [total, packet waits, stage waits, unit, items, tiles]
      NERFB_TRACE(if (trace && lane == 0) { long long* o = trace + 120 * 8 + 150 * 8 + (size_t)w * 8; o[0] = clock64() - t_begin; o[1] = acc_flag; o[2] = acc_empty; o[3] = u; o[4] = n_it; o[5] = my_tiles; })
    }
  } else if (warp == 6) {
    if (lane == 0) {                         // ===== scout: polls the chain CTAs' packet counters ahead of the copy issuer =====
      const uint32_t cnt = smem_u32(&sm.scout_cnt);
      int tc_c = jg % fc.Nc; uint32_t tc_k = (uint32_t)(jg / fc.Nc);
      for (int ti = 0; ti < my_tiles; ++ti) {
        const int64_t tile = jg + (int64_t)ti * G;
        const int c = tc_c; const uint32_t k = tc_k;      // tile = c + k * Nc, tracked incrementally (no 64-bit div / mod per tile)
        tc_c += G; while (tc_c >= fc.Nc) { tc_c -= fc.Nc; ++tc_k; }
        uint32_t seen = 0u, hseen = 0u;        // last values read of this chain CTA's packet / head counters (monotonic)
        for (int li = 0; li < n_it; ++li) {
          const int pk = args.item_pk[args.unit_item[u][li]];
          bool fresh = false;
          if (pk >= 0) {
            const uint32_t n = k * (uint32_t)fc.PK + (uint32_t)pk;
            if ((int32_t)(seen - (n + 1)) < 0) {
              const uint32_t* flag = fc.produced + (size_t)c * F_FLAG_STRIDE;
              gflag_wait_ge(flag, n + 1);
              seen = ld_acquire_gpu(flag); fresh = true;
            }
          } else if ((int32_t)(hseen - 4u * (k + 1)) < 0) {
            const uint32_t* flag = fc.hprod + (size_t)c * F_FLAG_STRIDE;
            gflag_wait_ge(flag, 4u * (k + 1));
            hseen = ld_acquire_gpu(flag); fresh = true;
          }
          if (fresh) fence_proxy_async_all();
          cnt_publish(cnt);                    // release: the issuer's copies follow its acquire of this counter
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {                         // ===== MMA issuer =====
      int st = 0; uint32_t ph = 0;
      for (int ti = 0; ti < my_tiles; ++ti) {
        NERFB_TRACE(const bool tr = trace && w == 0 && ti < 120;)
        for (int li = 0; li < n_it; ++li) {
          const int it = args.unit_item[u][li];
          const WItem& im = args.item[it];
          const uint32_t idesc = umma_idesc_bf16(128, im.N, 1, 1);
          const uint32_t col0 = (uint32_t)args.item_col[it][0], col1 = (uint32_t)args.item_col[it][1];
          for (int half = 0; half < 2; ++half) {
            NERFB_TRACE(if (tr && li == 0 && half == 0) trace[ti * 8 + 5] = clock64();)
            mbar_wait(&sm.full[st], ph);
            NERFB_TRACE(if (tr && li == 0 && half == 0) trace[ti * 8 + 6] = clock64();)
            tc_fence_after();
            const uint32_t base = smem_u32(sm.stage[st]);
#pragma unroll
            for (int ks = 0; ks < 4; ++ks)       // 16 samples per UMMA
              for (int mh = 0; mh < (im.nM >> 1); ++mh)
                umma_bf16(tmem + (mh ? col1 : col0), umma_desc_mn128(base + mh * 2 * W_HALF + ks * 2048, W_HALF),
                          umma_desc_mn128(base + 4 * W_HALF + ks * 2048, W_HALF), idesc, (ti > 0 || half > 0 || ks > 0) ? 1u : 0u);
            umma_commit(&sm.empty[st]);
            NERFB_TRACE(if (tr && li == 0 && half == 1) trace[ti * 8 + 7] = clock64();)
            if (++st == W_STAGES) { st = 0; ph ^= 1; }
          }
        }
      }
      umma_commit(&sm.acc_done);
    }
  } else if (warp < 6) {
    // ===== bias column sums (from shared memory), slot release, accumulator flush =====
    const int q = warp & 3, bt = (warp - 2) * 32 + lane;
    const int blk = bt >> 5, c2 = bt & 31;
    int st = 0; uint32_t ph = 0;
    float b0[F_UNIT_ITEMS], b1[F_UNIT_ITEMS];
#pragma unroll
    for (int li = 0; li < F_UNIT_ITEMS; ++li) { b0[li] = 0.f; b1[li] = 0.f; }
    int tc_c = jg % fc.Nc; uint32_t tc_k = (uint32_t)(jg / fc.Nc);
      for (int ti = 0; ti < my_tiles; ++ti) {
      const int64_t tile = jg + (int64_t)ti * G;
      const int c = tc_c; const uint32_t k = tc_k;
      tc_c += G; while (tc_c >= fc.Nc) { tc_c -= fc.Nc; ++tc_k; }
#pragma unroll
      for (int li = 0; li < F_UNIT_ITEMS; ++li) {
        if (li < n_it) {
          const int it = args.unit_item[u][li];
          const WItem& im = args.item[it];
          const bool sum_here = im.bias_off >= 0 && blk < im.nM;
          for (int half = 0; half < 2; ++half) {
            mbar_wait(&sm.full[st], ph);
            if (half == 1 && bt == 0) {        // both halves of the packet are in shared memory: hand the ring slot back
              const int pk = args.item_pk[it];
              if (pk >= 0) red_release_gpu_add(fc.consumed + (size_t)c * fc.q + (k * (uint32_t)fc.PK + (uint32_t)pk) % (uint32_t)fc.q, 1u);
              else red_release_gpu_add(fc.hcons + (size_t)c * F_HQ + k % F_HQ, 1u);
            }
            if (sum_here) {
              const uint8_t* src = sm.stage[st] + blk * W_HALF + ((c2 & 3) << 2);
#pragma unroll 8
              for (int r = 0; r < 64; ++r) {
                const uint32_t wd = *reinterpret_cast<const uint32_t*>(src + r * 128 + ((((uint32_t)c2 >> 2) ^ ((uint32_t)r & 7u)) << 4));
                b0[li] += __uint_as_float(wd << 16); b1[li] += __uint_as_float(wd & 0xFFFF0000u);
              }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.empty[st]);
            if (++st == W_STAGES) { st = 0; ph ^= 1; }
          }
        }
      }
    }
    mbar_wait(&sm.acc_done, 0);
    tc_fence_after();
#pragma unroll
    for (int li = 0; li < F_UNIT_ITEMS; ++li) {
      if (li < n_it) {
        const int it = args.unit_item[u][li];
        const WItem& im = args.item[it];
        if (im.bias_off >= 0 && blk < im.nM) {
          if (im.bias_alt < 0) {
            float* bdst = (im.aux ? args.aux : args.grad) + im.bias_off;
            atomicAdd(bdst + 2 * bt, b0[li]);
            atomicAdd(bdst + 2 * bt + 1, b1[li]);
          } else if (bt < 2) {               // head block: columns [g_r g_g g_b d_sigma]
            atomicAdd(args.grad + im.bias_off + 2 * bt, b0[li]);
            atomicAdd(bt == 0 ? args.grad + im.bias_off + 1 : args.grad + im.bias_alt, b1[li]);
          }
        }
        float* part = args.partial + (int64_t)(args.unit_pbase[u] + jg * n_it + args.item_pslot[it]) * W_SLOT;
        for (int mh = 0; mh < (im.nM >> 1); ++mh) {
          const int m = mh * 128 + q * 32 + lane;
          if (mh * 128 + q * 32 >= im.m_hi || mh * 128 + q * 32 + 32 <= im.m_lo) continue;      // warp-uniform
          float* dst = part + (int64_t)m * 256;
          const bool real = m >= im.m_lo && m < im.m_hi;
          for (int c0 = 0; c0 < im.N; c0 += 32) {
            uint32_t v[32];
            tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(args.item_col[it][mh] + c0), v);
            tmem_ld_wait();
            if (real && c0 < im.nvalid) {
#pragma unroll
              for (int i = 0; i < 32; i += 4)
                *reinterpret_cast<float4*>(dst + c0 + i) = make_float4(__uint_as_float(v[i]), __uint_as_float(v[i + 1]),
                                                                      __uint_as_float(v[i + 2]), __uint_as_float(v[i + 3]));
            }
          }
        }
      }
    }
    tc_fence_before();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

struct FusedArgs { DgradArgs d; FusedCtl fc; FusedWArgs w; long long* trace; };

template <bool SIREN>
__global__ void __launch_bounds__(NUM_THREADS, 1) mlp_tc_bwd_fused_kernel(const __grid_constant__ FusedArgs args) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  if ((int)blockIdx.x < args.fc.Nc) dgrad_role<SIREN, true>(args.d, args.fc, smem_raw, (int)blockIdx.x, args.fc.Nc);
  else wgrad_fused_role(args.w, args.fc, smem_raw, args.trace);
}

// grad[w_off + r * ld + col0 + n] += sum over the G partials of the item's unit
__global__ void __launch_bounds__(256) reduce_partials_fused_kernel(const __grid_constant__ FusedWArgs ra, const __grid_constant__ FusedCtl fcr) {
  const int it = blockIdx.x;
  const WItem& im = ra.item[it];
  const int u = ra.item_unit[it];
  const int G = fcr.unit_G[u];
  const int n_it = ra.unit_n[u];
  const float* base = ra.partial + (int64_t)(ra.unit_pbase[u] + ra.item_pslot[it]) * W_SLOT;
  const int total = (im.m_hi - im.m_lo) * im.nvalid;
  if (im.nvalid == 256 && !im.hilo && !im.aux && ((im.ld | im.col0) & 3) == 0 && (im.w_off & 3) == 0 &&
      (reinterpret_cast<uintptr_t>(ra.grad) & 15) == 0) {
    // the common item (a full 256-column block of a hidden layer): four columns per thread, 16-byte accesses, no division;
    // same summation order per element
    for (int e = blockIdx.y * blockDim.x + threadIdx.x; e < total / 4; e += gridDim.y * blockDim.x) {
      const int r = im.m_lo + (e >> 6), n = (e & 63) * 4;
      float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int j = 0; j < G; ++j) {
        const float4 p = *reinterpret_cast<const float4*>(base + (int64_t)j * n_it * W_SLOT + r * 256 + n);
        s.x += p.x; s.y += p.y; s.z += p.z; s.w += p.w;
      }
      float4* g = reinterpret_cast<float4*>(ra.grad + im.w_off + (int64_t)r * im.ld + im.col0 + n);
      float4 o = *g;
      o.x += s.x; o.y += s.y; o.z += s.z; o.w += s.w;
      *g = o;
    }
    return;
  }
  for (int e = blockIdx.y * blockDim.x + threadIdx.x; e < total; e += gridDim.y * blockDim.x) {
    const int r = im.m_lo + e / im.nvalid, n = e % im.nvalid;
    float s = 0.f;
    for (int j = 0; j < G; ++j) {
      const float* pp = base + (int64_t)j * n_it * W_SLOT + r * 256 + n;
      s += im.hilo ? pp[0] + pp[3] : pp[0];
    }
    if (im.aux) ra.aux[im.w_off + (int64_t)r * im.ld + im.col0 + n] = s;
    else ra.grad[im.w_off + (int64_t)r * im.ld + im.col0 + n] += s;
  }
}

// grad[w_off + r * ld + col0 + n] += sum over the active CTAs' partials
__global__ void __launch_bounds__(256) reduce_partials_kernel(const __grid_constant__ WgradArgs ra) {
  const int it = blockIdx.x;
  const WItem& im = ra.item[it];
  __shared__ int s_c[NERFB_SM_COUNT + 1];
  __shared__ int s_flag[NERFB_SM_COUNT + 1];
  __shared__ int s_n;
  if (threadIdx.x < ra.n_active) {           // the CTAs that own a non-empty tile range of this item (one thread per CTA)
    int64_t t0, t1; seg_range(ra, threadIdx.x, it, t0, t1);
    s_flag[threadIdx.x] = t1 > t0;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int n = 0;
    for (int c = 0; c < ra.n_active; ++c) if (s_flag[c]) s_c[n++] = c;
    s_n = n;
  }
  __syncthreads();
  const int total = (im.m_hi - im.m_lo) * im.nvalid;
  for (int e = blockIdx.y * blockDim.x + threadIdx.x; e < total; e += gridDim.y * blockDim.x) {
    const int r = im.m_lo + e / im.nvalid, n = e % im.nvalid;
    float s = 0.f;
    for (int k = 0; k < s_n; ++k) {
      const float* pp = ra.partial + (int64_t)(s_c[k] + it) * W_SLOT + r * 256 + n;
      s += im.hilo ? pp[0] + pp[3] : pp[0];
    }
    if (im.aux) ra.aux[im.w_off + (int64_t)r * im.ld + im.col0 + n] = s;
    else ra.grad[im.w_off + (int64_t)r * im.ld + im.col0 + n] += s;
  }
}

// Class-head chain rule (logits = W_p1 (W_p0 feat + b_p0) + b_p1, models/nerf_cls.py:56-61):  with dWe = sum g feat^T
// (n_cls x 256) and db = sum g (n_cls) from the wgrad:
//   dW_p0[j][i] = sum_k W_p1[k][j] dWe[k][i]      db_p0[j] = sum_k W_p1[k][j] db[k]
//   dW_p1[k][j] = sum_i dWe[k][i] W_p0[j][i] + db[k] b_p0[j]      db_p1[k] = db[k]
// One block per hidden unit j (512), one thread per input i (256).  aux = dWe [16 x 256] (rows 4.. of the head item) | colsums.
__global__ void __launch_bounds__(256) parse_chain_kernel(const float* __restrict__ params, const float* __restrict__ aux, int n_cls,
                                                          int64_t w_p0, int64_t b_p0, int64_t w_p1, int64_t b_p1, float* __restrict__ grad) {
  const int j = blockIdx.x, i = threadIdx.x;
  __shared__ float red[8][8];
  const float* dwe = aux;                       // row k at aux + k * 256
  const float* db = aux + 16 * 256 + 4;         // head-block columns 4 .. 4 + n_cls
  const float w0 = __ldg(params + w_p0 + (int64_t)j * 256 + i);
  float acc = 0.f;
  float part[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    part[k] = 0.f;
    if (k < n_cls) {
      const float d = dwe[k * 256 + i];
      acc = fmaf(__ldg(params + w_p1 + (int64_t)k * 512 + j), d, acc);
      part[k] = d * w0;
    }
  }
  grad[w_p0 + (int64_t)j * 256 + i] += acc;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const float v = warp_sum(part[k]);
    if ((i & 31) == 0) red[k][i >> 5] = v;
  }
  __syncthreads();
  if (i < n_cls) {
    float s = 0.f;
    for (int w = 0; w < 8; ++w) s += red[i][w];
    grad[w_p1 + (int64_t)i * 512 + j] += s + db[i] * __ldg(params + b_p0 + j);
  }
  if (i == 32) {
    float s = 0.f;
    for (int k = 0; k < n_cls; ++k) s = fmaf(__ldg(params + w_p1 + (int64_t)k * 512 + j), db[k], s);
    grad[b_p0 + j] += s;
  }
  if (j == 0 && i >= 64 && i < 64 + n_cls) grad[b_p1 + (i - 64)] += db[i - 64];
}

}  // namespace

// =====================================================================================================================
// head parameter gather + the launch plan
// =====================================================================================================================
namespace {

struct HeadArgs {
  const float* out; const float* d_out; const uint8_t* fwd_rec; int fwd_tile_bytes; int hD_off, hd_off;
  int64_t P, n_tiles; int n_active;
  float* grad; int64_t w_sigma, b_sigma, w_rgb, b_rgb;
};

__global__ void __launch_bounds__(256) gather_heads_kernel(const float* __restrict__ params, int64_t w_sigma, int64_t w_rgb,
                                                           float* __restrict__ heads, int rgb_in) {
  for (int i = threadIdx.x; i < 256 + 3 * rgb_in; i += blockDim.x) heads[i] = __ldg(params + (i < 256 ? w_sigma + i : w_rgb + (i - 256)));
}

struct BwdPlan {
  BProgram prog; PackTArgs pack; WgradArgs wg; HeadArgs head;
  int64_t off_wblob, off_heads, off_rec, off_partial, off_aux, total;
  int n_active; int64_t n_tiles; int64_t param_count;
};

BwdPlan bwd_plan_siren(const nerfb_mlp_desc& d, int64_t P);

BwdPlan bwd_plan(const nerfb_mlp_desc& d, int64_t P) {
  if (d.kind == NERFB_KIND_SIREN) return bwd_plan_siren(d, P);
  BwdPlan pl{};
  const NerfOffsets o = nerf_offsets(d);
  const TcFwdLayout L = tc_fwd_layout(d, P);
  const int D = d.D;
  const int dzdir = 0, dzfin = 2 * BLK_BYTES;
  auto dz_off = [&](int i) { return (6 + 4 * (D - i)) * BLK_BYTES; };          // dZ_i, i = D..1 (grad of trunk step i-1's pre-activation)
  BProgram& pg = pl.prog;
  const int dr_idx = D + 1 + (d.n_cls > 0 ? 1 : 0);            // forward step order: trunk 0..D-1, final, [class head], direction
  pg.hd_off = L.step_mask_off[dr_idx]; pg.dzdir_off = dzdir;
  pg.n_cls = d.n_cls; pg.out_c = 4 + d.n_cls; pg.siren = 0; pg.n_pro = 2; pg.rgb_in = 128;
  const int hblk = (6 + 4 * D) * BLK_BYTES;                     // head block (+ one never-written block behind it)
  pg.h_off = hblk;
  pg.fwd_tile_bytes = L.tile_save_bytes; pg.bwd_tile_bytes = (8 + 4 * D) * BLK_BYTES;
  int ns = 0, nslab = 0; int64_t wb = 0;
  auto slabs = [&](int64_t w_off, int ld, int col0, int n_kb) {
    for (int kb = 0; kb < n_kb; ++kb) { pl.pack.slab[nslab++] = PackTSlab{w_off, ld, kb * 64, 64, col0, wb}; wb += STAGE_BYTES; }
  };
  pg.s[ns++] = BStep{2, -1, 0, dzfin, d.n_cls > 0 ? 1 : 0};                       slabs(o.w_dir, 256 + o.in_dir, 0, 2);
  pg.s[ns++] = BStep{4, L.step_mask_off[D - 1], 1, dz_off(D)}; slabs(o.w_final, 256, 0, 4);
  for (int i = D - 1; i >= 1; --i) {
    const bool skip = (d.skip_mask >> i) & 1;
    pg.s[ns++] = BStep{4, L.step_mask_off[i - 1], 0, dz_off(i)};
    slabs(o.w[i], o.in[i], skip ? o.in_xyz : 0, 4);
  }
  pg.n_steps = ns; pl.pack.n_slabs = nslab;
  // wgrad work list
  WgradArgs& wg = pl.wg; int ni = 0;
  for (int i = 0; i < D; ++i) {
    const bool skip = i > 0 && ((d.skip_mask >> i) & 1);
    if (i == 0) wg.item[ni++] = WItem{dz_off(1), 4, 1, L.save_x_off, 1, 64, o.w[0], o.in[0], 0, o.in_xyz, o.b[0], 0, 256, -1};
    else {
      wg.item[ni++] = WItem{dz_off(i + 1), 4, 1, L.step_save_off[i - 1], 4, 256, o.w[i], o.in[i], skip ? o.in_xyz : 0, 256, o.b[i], 0, 256, -1};
      if (skip) wg.item[ni++] = WItem{dz_off(i + 1), 4, 1, L.save_x_off, 1, 64, o.w[i], o.in[i], 0, o.in_xyz, -1, 0, 256, -1};
    }
  }
  wg.item[ni++] = WItem{dzfin, 4, 1, L.step_save_off[D - 1], 4, 256, o.w_final, 256, 0, 256, o.b_final, 0, 256, -1};
  wg.item[ni++] = WItem{dzdir, 2, 1, L.step_save_off[D], 4, 256, o.w_dir, 256 + o.in_dir, 0, 256, o.b_dir, 0, 128, -1};
  wg.item[ni++] = WItem{dzdir, 2, 1, L.save_dir_off, 1, 64, o.w_dir, 256 + o.in_dir, 256, o.in_dir, -1, 0, 128, -1};
  // heads (models/nerf.py:112,119-120): the head block H holds [g_r g_g g_b d_sigma] in columns 0..3, so rows 0..2 of
  // H^T hd are dW_rgb and row 3 of H^T h_D is dW_sigma; its column sums are the two head biases
  wg.item[ni++] = WItem{hblk, 2, 1, L.step_save_off[dr_idx], 2, 128, o.w_rgb, 128, 0, 128, o.b_rgb, 0, 3, o.b_sigma};
  wg.item[ni++] = WItem{hblk, 2, 1, L.step_save_off[D - 1], 4, 256, o.w_sigma - 3 * 256, 256, 0, 256, -1, 3, 4, -1};
  if (d.n_cls > 0)      // class head: rows 4 .. 4+n_cls of H^T feat are dW_eff; head-block column sums -> aux[4096 ..]
    wg.item[ni++] = WItem{hblk, 2, 1, L.step_save_off[D], 4, 256, -4 * 256, 256, 0, 256, 16 * 256, 4, 4 + d.n_cls, -1, 1};
  wg.n_items = ni;
  pl.n_tiles = ceil_div64(P, TILE_M);
  pl.n_active = (int)(pl.n_tiles < NERFB_SM_COUNT ? pl.n_tiles : NERFB_SM_COUNT);
  pl.param_count = o.total;
  wg.fwd_tile_bytes = pg.fwd_tile_bytes; wg.bwd_tile_bytes = pg.bwd_tile_bytes; wg.n_tiles = pl.n_tiles; wg.n_active = pl.n_active;
  wg.item_off[0] = 0;
  for (int i = 0; i < ni; ++i) wg.item_off[i + 1] = wg.item_off[i] + (int64_t)item_cost(wg.item[i]) * pl.n_tiles;
  HeadArgs& h = pl.head;
  h.fwd_tile_bytes = pg.fwd_tile_bytes; h.hD_off = L.step_save_off[D - 1]; h.hd_off = L.step_save_off[dr_idx];
  h.P = P; h.n_tiles = pl.n_tiles; h.n_active = pl.n_active;
  h.w_sigma = o.w_sigma; h.b_sigma = o.b_sigma; h.w_rgb = o.w_rgb; h.b_rgb = o.b_rgb;
  int64_t off = 0;
  pl.off_wblob = off;   off += round_up64(wb, 1024);
  pl.off_heads = off;   off += 4096;
  pl.off_rec = off;     off += pl.n_tiles * (int64_t)pg.bwd_tile_bytes;
  pl.off_partial = off; off += (int64_t)(pl.n_active + MAX_ITEMS) * W_SLOT * 4;
  pl.off_aux = off;     off += (16 * 256 + 128) * 4;
  pl.total = round_up64(off, 1024);
  return pl;
}

// SIREN / FiLM (models/nerf.py:201-216): forward steps 0..D-1 = FiLM layers (h_l saved at step_save_off[l], freq*cos image at
// step_mask_off[l]), step D = colour FiLM layer on [dir | h_{D-1}] (hc, cos image), heads sigma <- h_{D-1}, rgb <- hc.
// dgrad: prologue dZ_c = (g W_rgb) (.) cos_c (4 blocks); step 0: x W_dir[:, 3:]^T + d_sigma w_sigma, (.) cos_{D-1} -> dZ_{D-1};
// step k: x W_{D-k}^T (.) cos_{D-1-k} -> dZ_{D-1-k}, k = 1..D-1.  Backward record: dZ_c | dZ_{D-1} .. dZ_0 | head block.
BwdPlan bwd_plan_siren(const nerfb_mlp_desc& d, int64_t P) {
  BwdPlan pl{};
  const NerfOffsets o = nerf_offsets(d);
  const TcFwdLayout L = tc_fwd_layout(d, P);
  const int D = d.D;
  auto dz_off = [&](int l) { return (4 + 4 * (D - 1 - l)) * BLK_BYTES; };       // dZ_l, l = D-1 .. 0
  BProgram& pg = pl.prog;
  pg.hd_off = L.step_mask_off[D]; pg.dzdir_off = 0; pg.hd_act_off = L.step_save_off[D]; pg.hd_film_off = D * 256;
  pg.n_cls = 0; pg.out_c = 4; pg.siren = 1; pg.n_pro = 4; pg.rgb_in = 256;
  const int hblk = (4 + 4 * D) * BLK_BYTES;
  pg.h_off = hblk;
  pg.fwd_tile_bytes = L.tile_save_bytes; pg.bwd_tile_bytes = (6 + 4 * D) * BLK_BYTES;
  int ns = 0, nslab = 0; int64_t wb = 0;
  auto slabs = [&](int64_t w_off, int ld, int col0, int n_kb) {
    for (int kb = 0; kb < n_kb; ++kb) { pl.pack.slab[nslab++] = PackTSlab{w_off, ld, kb * 64, 64, col0, wb}; wb += STAGE_BYTES; }
  };
  pg.s[ns++] = BStep{4, L.step_mask_off[D - 1], 1, dz_off(D - 1), 0, L.step_save_off[D - 1], (D - 1) * 256};  slabs(o.w_dir, 259, 3, 4);
  for (int l = D - 1; l >= 1; --l) { pg.s[ns++] = BStep{4, L.step_mask_off[l - 1], 0, dz_off(l - 1), 0, L.step_save_off[l - 1], (l - 1) * 256}; slabs(o.w[l], 256, 0, 4); }
  pg.n_steps = ns; pl.pack.n_slabs = nslab;
  WgradArgs& wg = pl.wg; int ni = 0;
  wg.item[ni++] = WItem{dz_off(0), 4, 1, L.save_x_off, 1, 64, o.w[0], 3, 0, 3, o.b[0], 0, 256, -1, 0, 1};
  for (int l = 1; l < D; ++l)
    wg.item[ni++] = WItem{dz_off(l), 4, 1, L.step_save_off[l - 1], 4, 256, o.w[l], 256, 0, 256, o.b[l], 0, 256, -1, 0, 0};
  wg.item[ni++] = WItem{0, 4, 1, L.step_save_off[D - 1], 4, 256, o.w_dir, 259, 3, 256, o.b_dir, 0, 256, -1, 0, 0};
  wg.item[ni++] = WItem{0, 4, 1, L.save_dir_off, 1, 64, o.w_dir, 259, 0, 3, -1, 0, 256, -1, 0, 1};
  wg.item[ni++] = WItem{hblk, 2, 1, L.step_save_off[D], 4, 256, o.w_rgb, 256, 0, 256, o.b_rgb, 0, 3, o.b_sigma, 0, 0};
  wg.item[ni++] = WItem{hblk, 2, 1, L.step_save_off[D - 1], 4, 256, o.w_sigma - 3 * 256, 256, 0, 256, -1, 3, 4, -1, 0, 0};
  wg.n_items = ni;
  pl.n_tiles = ceil_div64(P, TILE_M);
  pl.n_active = (int)(pl.n_tiles < NERFB_SM_COUNT ? pl.n_tiles : NERFB_SM_COUNT);
  pl.param_count = o.total;
  wg.fwd_tile_bytes = pg.fwd_tile_bytes; wg.bwd_tile_bytes = pg.bwd_tile_bytes; wg.n_tiles = pl.n_tiles; wg.n_active = pl.n_active;
  wg.item_off[0] = 0;
  for (int i = 0; i < ni; ++i) wg.item_off[i + 1] = wg.item_off[i] + (int64_t)item_cost(wg.item[i]) * pl.n_tiles;
  HeadArgs& h = pl.head;
  h.w_sigma = o.w_sigma; h.b_sigma = o.b_sigma; h.w_rgb = o.w_rgb; h.b_rgb = o.b_rgb;
  int64_t off = 0;
  pl.off_wblob = off;   off += round_up64(wb, 1024);
  pl.off_heads = off;   off += 8192;
  pl.off_rec = off;     off += pl.n_tiles * (int64_t)pg.bwd_tile_bytes;
  pl.off_partial = off; off += (int64_t)(pl.n_active + MAX_ITEMS) * W_SLOT * 4;
  pl.off_aux = off;     off += (16 * 256 + 128) * 4;
  pl.total = round_up64(off, 1024);
  return pl;
}

}  // namespace

bool tc_bwd_supported(const nerfb_mlp_desc& d) {
  if (!tc_supported(d, 0)) return false;
  if (d.kind == NERFB_KIND_SIREN) return d.D >= 2 && d.D + 4 <= MAX_ITEMS;
  int skips = 0;
  for (int i = 1; i < d.D; ++i) skips += (d.skip_mask >> i) & 1;
  return d.D + skips + 5 + (d.n_cls > 0 ? 1 : 0) <= MAX_ITEMS && d.D + 1 <= MAX_BSTEPS;
}

namespace {

// ---- fused launch plan: units (bin packing of the wgrad items under the 512-column TMEM budget), SM split, ring sizes -------
int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return (v && *v) ? atoi(v) : dflt;
}
// NERFB_BWD_FUSED: 0 = three launches (dgrad -> HBM record -> wgrad), 1 = fused launch wherever it is expressible, 2 (default) =
// fused launch where it was measured to be at least as fast on the B200: plain NeRF (4096 x 128: 1.88 ms against 1.93 ms, and 5.9 GB less
// HBM traffic per fine pass -> under the power cap the step sustains 4.60 ms instead of 4.81 ms) and NeRF_3D (8192 x 128: 3.87 ms against
// 4.02 ms once the ring holds a whole tile's 12 packets); SIREN keeps the 3-launch path.
#ifndef NERFB_BWD_FUSED_DEFAULT
#define NERFB_BWD_FUSED_DEFAULT 2
#endif
// The fused launch spins on flags written by OTHER CTAs of the same grid: every one of its NERFB_SM_COUNT CTAs must be resident
// at once (one per SM; the kernel's shared memory admits exactly one): the launch is COOPERATIVE (gang-scheduled, refused rather
// than deadlocked when it cannot fit).  Checked per device, once: the device has that many SMs and the kernel fits an SM;
// anything else (a smaller part, a MIG slice) silently keeps the 3-launch path, which has no such requirement.
static bool fused_device_ok() {
  static int cached[64] = {};            // 0 unknown, 1 ok, 2 not ok
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return false;
  if (cached[dev] == 0) {
    int sms = 0;
    int per_sm = 0;
    const size_t smem_f = (sizeof(SmemD) > sizeof(SmemW) ? sizeof(SmemD) : sizeof(SmemW)) + 1024;
    const bool ok = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && sms >= NERFB_SM_COUNT &&
                    cudaFuncSetAttribute(mlp_tc_bwd_fused_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_f) == cudaSuccess &&
                    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, mlp_tc_bwd_fused_kernel<false>, NUM_THREADS, smem_f) == cudaSuccess &&
                    per_sm >= 1;
    if (!ok) cudaGetLastError();
    cached[dev] = ok ? 1 : 2;
  }
  return cached[dev] == 1;
}
struct FusedPlan {
  bool on;
  FusedCtl fc; FusedWArgs w;
  int n_units, grid;
  int64_t off_ring, off_hring, off_flags, flags_bytes, off_partial, off_aux, total;   // relative to the backward workspace part
};

// MMA issue cost of one item per tile in clocks: 8 K-steps x M-halves x N/2 (tcgen05 floor), small-N shapes are bound by the
// shared-memory operand fetch instead (A 4 KB + B N*32 B per instruction at 128 B/clk)
int item_clk(const WItem& im) {
  const int per = im.N / 2 > (4096 + im.N * 32) / 128 ? im.N / 2 : (4096 + im.N * 32) / 128;
  return 8 * (im.nM >> 1) * per;
}

FusedPlan fused_plan(const BwdPlan& pl) {
  FusedPlan fp{};
  fp.on = false;
  const int mode = env_int("NERFB_BWD_FUSED", NERFB_BWD_FUSED_DEFAULT);
  if (mode == 0 || !fused_device_ok()) return fp;
  const BProgram& pg = pl.prog;
  if (mode == 2 && pg.siren) return fp;      // SIREN: the sine-gate chain is the slow side (2.83 ms fused at best against 2.61 ms)
  const WgradArgs& wg = pl.wg;
  FusedWArgs& w = fp.w;
  w.n_items = wg.n_items;
  int order[MAX_ITEMS];
  for (int i = 0; i < wg.n_items; ++i) {
    w.item[i] = wg.item[i];
    const int dz = wg.item[i].dz_off;
    int pk = -2;
    if (dz == pg.h_off) pk = -1;
    else if (dz == pg.dzdir_off) pk = 0;
    else for (int s = 0; s < pg.n_steps; ++s) if (pg.s[s].save_off == dz) pk = s + 1;
    if (pk == -2 || !wg.item[i].a_fwd) return fp;                       // not expressible as packets: keep the 3-launch path
    w.item_pk[i] = pk;
    order[i] = i;
  }
  // first-fit decreasing on cost; capacity: 512 TMEM columns, F_UNIT_ITEMS items, ~1.2 x the cost of a full 256 x 256 item
  for (int a = 0; a < wg.n_items; ++a)
    for (int b = a + 1; b < wg.n_items; ++b)
      if (item_clk(wg.item[order[b]]) > item_clk(wg.item[order[a]])) { const int t = order[a]; order[a] = order[b]; order[b] = t; }
  int ucost[F_MAX_UNITS] = {0}, ucols[F_MAX_UNITS] = {0};
  const int cap = (int)(2048 * 1.2);
  int nu = 0;
  for (int a = 0; a < wg.n_items; ++a) {
    const int i = order[a];
    const WItem& im = wg.item[i];
    const int cols = (im.nM >> 1) * im.N, clk = item_clk(im);
    int u = 0;
    for (; u < nu; ++u)
      if (ucols[u] + cols <= 512 && w.unit_n[u] < F_UNIT_ITEMS && ucost[u] + clk <= cap) break;
    if (u == nu) { if (nu == F_MAX_UNITS) return fp; ++nu; w.unit_n[u] = 0; }
    w.unit_item[u][w.unit_n[u]++] = i;
    w.item_unit[i] = u;
    ucols[u] += cols; ucost[u] += clk;
  }
  w.n_units = nu; fp.n_units = nu;
  // inside a unit: packet order (head block first), then TMEM columns and partial slots in that order
  for (int u = 0; u < nu; ++u) {
    int* li = w.unit_item[u];
    for (int a = 0; a < w.unit_n[u]; ++a)
      for (int b = a + 1; b < w.unit_n[u]; ++b)
        if (w.item_pk[li[b]] < w.item_pk[li[a]]) { const int t = li[a]; li[a] = li[b]; li[b] = t; }
    int col = 0;
    for (int a = 0; a < w.unit_n[u]; ++a) {
      const WItem& im = w.item[li[a]];
      w.item_pslot[li[a]] = a;
      for (int mh = 0; mh < (im.nM >> 1); ++mh) { w.item_col[li[a]][mh] = col; col += im.N; }
    }
  }
  FusedCtl& fc = fp.fc;
  fc.G = env_int("NERFB_FUSED_G", 8);
  fc.q = env_int("NERFB_FUSED_Q", pg.n_steps + 2);      // ring slots = one more than the packets of a tile (prologue + n_steps): NeRF 10 packets -> 11 (10: equal, 9: 3.7 ms instead of 1.9), NeRF_3D 11 -> 12 (11: 4.63 ms, 12: 4.06 ms)
  if (fc.G < 1 || fc.q < 3 || fc.q > F_MAX_Q) return fp;
  const int budget_dflt = nu * fc.G < 84 ? nu * fc.G : 84;      // wgrad CTAs: NeRF 10 units -> 80 (74..82 equal), NeRF_3D 11 units -> 84 (88: 4.06 ms, 84: 3.87 ms, 92: 4.38 ms)
  fc.Nc = NERFB_SM_COUNT - env_int("NERFB_FUSED_BUDGET", budget_dflt);
  fc.PK = pg.n_steps + 1;
  // the fused grid needs every wgrad CTA to own tiles and the chain CTAs several rounds each; small problems keep the 3-launch path
  if (fc.Nc < 16 || pl.n_tiles < 4 * (int64_t)NERFB_SM_COUNT) return fp;
  fc.m_head = 0;
  for (int i = 0; i <= MAX_BSTEPS; ++i) fc.pk_need[i] = 0;
  for (int i = 0; i < w.n_items; ++i) { if (w.item_pk[i] < 0) ++fc.m_head; else ++fc.pk_need[w.item_pk[i]]; }
  int pbase = 0;
  // Split the n_units * G wgrad CTAs by the measured per-tile cost of a unit's producer lane (tools/trace_fused.py on the
  // B200: ~2.8k clk for a one-item unit, ~2.1k clk per item for multi-item units -- the stage / copy-issue hand-shakes,
  // not the MMAs, dominate), largest remainder first, at least one CTA per unit.
  {
    const int budget = env_int("NERFB_FUSED_BUDGET", budget_dflt);
    double cost[F_MAX_UNITS], tot = 0.0;
    const double c1 = env_int("NERFB_FUSED_C1", 2800), c0 = env_int("NERFB_FUSED_C0", 300), cn = env_int("NERFB_FUSED_CN", 2050);   // tuning knobs
    for (int u = 0; u < nu; ++u) { cost[u] = w.unit_n[u] == 1 ? c1 : c0 + cn * w.unit_n[u]; tot += cost[u]; }   // measured busy clk per tile: 2.8k / 4.4k (2 items) / 8.5k (4 items)
    int given = 0;
    double frac[F_MAX_UNITS];
    for (int u = 0; u < nu; ++u) {
      const double share = budget * cost[u] / tot;
      fc.unit_G[u] = (int)share < 1 ? 1 : (int)share;
      frac[u] = share - (int)share; given += fc.unit_G[u];
    }
    while (given < budget) {
      int best = 0;
      for (int u = 1; u < nu; ++u) if (frac[u] > frac[best]) best = u;
      ++fc.unit_G[best]; frac[best] = -1.0; ++given;
    }
    while (given > budget) {           // the >= 1 floor overshot: take from the largest
      int best = 0;
      for (int u = 1; u < nu; ++u) if (fc.unit_G[u] > fc.unit_G[best]) best = u;
      --fc.unit_G[best]; --given;
    }
    int w0 = 0;
    for (int u = 0; u < nu; ++u) { fc.unit_w0[u] = w0; w0 += fc.unit_G[u]; }
  }
  for (int u = 0; u < nu; ++u) { w.unit_pbase[u] = pbase; pbase += fc.unit_G[u] * w.unit_n[u]; }
  fp.grid = NERFB_SM_COUNT;
  int64_t off = 0;
  fp.off_ring = off;    off += (int64_t)fc.Nc * fc.q * F_PKT_BYTES;
  fp.off_hring = off;   off += (int64_t)fc.Nc * F_HQ * 2 * BLK_BYTES;
  fp.off_flags = off;   fp.flags_bytes = ((int64_t)fc.Nc * (2 * F_FLAG_STRIDE + fc.q + F_HQ)) * 4; off += round_up64(fp.flags_bytes, 1024);
  fp.off_partial = off; off += (int64_t)pbase * W_SLOT * 4;
  fp.off_aux = off;     off += (16 * 256 + 128) * 4;
  fp.total = round_up64(off, 1024);
  fp.on = true;
  return fp;
}

}  // namespace

static long long* g_trace_bwd = nullptr;
extern "C" void nerfb_debug_trace_bwd(long long* device_buf) { g_trace_bwd = device_buf; }   // diagnostics only (tools/trace_fused.py)

// The backward workspace follows the forward's: packed W^T slabs | head vectors | then EITHER the 3-launch path's per-tile dZ
// record + partials, OR the fused path's rings + flags + partials (decided by the same plan function in both entry points).
int64_t tc_bwd_workspace_bytes(const nerfb_mlp_desc& d, int64_t P) {
  const BwdPlan pl = bwd_plan(d, P);
  const FusedPlan fp = fused_plan(pl);
  return fp.on ? pl.off_rec + fp.total : pl.total;
}

int tc_mlp_bwd(const nerfb_mlp_desc& d, int in_mode, const float* params, const float* film, const float* in0,
               const float* in1, int64_t n_rows, int n_samples, const float* out, const float* d_out, float* grad,
               float* d_in, void* ws, int64_t ws_bytes, void* stream) {
  (void)in0; (void)in1;
  if (d_in || !tc_bwd_supported(d)) return NERFB_E_UNSUPPORTED;
  if (d.kind == NERFB_KIND_SIREN && !film) return NERFB_E_BADARG;      // the dgrad gates need the FiLM frequencies
  const int64_t P = (in_mode == NERFB_IN_RAYS) ? n_rows * n_samples : n_rows;
  const TcFwdLayout L = tc_fwd_layout(d, P);
  BwdPlan pl = bwd_plan(d, P);
  FusedPlan fp = fused_plan(pl);
  if (ws_bytes < L.total + (fp.on ? pl.off_rec + fp.total : pl.total)) return NERFB_E_WORKSPACE;
  if ((reinterpret_cast<uintptr_t>(ws) & 1023) != 0) return NERFB_E_BADARG;
  uint8_t* base = static_cast<uint8_t*>(ws);
  uint8_t* bb = base + L.total;
  const uint8_t* fwd_rec = base + L.off_save;
  float* heads = reinterpret_cast<float*>(bb + pl.off_heads);
  NERFB_LAUNCH(pack_weights_t_kernel, pl.pack.n_slabs * 4, 256, 0, stream, params, pl.pack, bb + pl.off_wblob);
  NERFB_LAUNCH(gather_heads_kernel, 1, 256, 0, stream, params, pl.head.w_sigma, pl.head.w_rgb, heads, pl.prog.rgb_in);

  int dev = 0;
  NERFB_CUDA(cudaGetDevice(&dev));
  static bool attr_set[64] = {};          // the attribute is per device
  const size_t smem_d = sizeof(SmemD) + 1024, smem_w = sizeof(SmemW) + 1024;
  const size_t smem_f = smem_d > smem_w ? smem_d : smem_w;
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    NERFB_CUDA(cudaFuncSetAttribute(mlp_tc_dgrad_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_d));
    NERFB_CUDA(cudaFuncSetAttribute(mlp_tc_dgrad_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_d));
    NERFB_CUDA(cudaFuncSetAttribute(mlp_tc_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_w));
    NERFB_CUDA(cudaFuncSetAttribute(mlp_tc_bwd_fused_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_f));
    NERFB_CUDA(cudaFuncSetAttribute(mlp_tc_bwd_fused_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_f));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  DgradArgs da{};
  da.prog = pl.prog; da.wblob = bb + pl.off_wblob; da.heads = heads; da.out = out; da.d_out = d_out;
  da.fwd_rec = fwd_rec; da.bwd_rec = bb + pl.off_rec; da.P = P; da.n_tiles = pl.n_tiles;
  da.weff = reinterpret_cast<const float*>(base + L.off_weff);
  da.film = film;

  if (fp.on) {
    // ONE launch: chain CTAs [0, Nc) and wgrad CTAs [Nc, grid) of a co-resident grid (<= 1 CTA per SM) exchange dZ through L2
    uint8_t* fb = bb + pl.off_rec;
    FusedArgs fa{};
    fa.d = da; fa.d.bwd_rec = nullptr;
    fa.fc = fp.fc;
    fa.fc.ring = fb + fp.off_ring; fa.fc.hring = fb + fp.off_hring;
    uint32_t* flags = reinterpret_cast<uint32_t*>(fb + fp.off_flags);
    fa.fc.produced = flags;
    fa.fc.hprod = flags + (size_t)fp.fc.Nc * F_FLAG_STRIDE;
    fa.fc.consumed = flags + (size_t)fp.fc.Nc * 2 * F_FLAG_STRIDE;
    fa.fc.hcons = fa.fc.consumed + (size_t)fp.fc.Nc * fp.fc.q;
    fa.w = fp.w;
    fa.w.fwd_rec = fwd_rec; fa.w.fwd_tile_bytes = pl.prog.fwd_tile_bytes; fa.w.n_tiles = pl.n_tiles;
    fa.w.partial = reinterpret_cast<float*>(fb + fp.off_partial); fa.w.grad = grad;
    fa.w.aux = reinterpret_cast<float*>(fb + fp.off_aux);
    fa.trace = g_trace_bwd;
    fa.d.trace_chain = g_trace_bwd ? g_trace_bwd + 120 * 8 : nullptr;
    NERFB_CUDA(cudaMemsetAsync(flags, 0, (size_t)fp.flags_bytes, (cudaStream_t)stream));
    if (d.n_cls > 0) NERFB_CUDA(cudaMemsetAsync(fa.w.aux + 16 * 256, 0, 128 * 4, (cudaStream_t)stream));
    {
      // COOPERATIVE launch: the CTAs of this grid spin on flags written by one another, so all of them must be resident at
      // once.  A cooperative launch is gang-scheduled (it starts only when the whole grid fits, also next to other work on
      // the device) and is refused with an error -- instead of deadlocking until the bounded waits trap -- when it can never fit.
      cudaLaunchConfig_t lc = {};
      lc.gridDim = dim3((unsigned)fp.grid); lc.blockDim = dim3(NUM_THREADS); lc.dynamicSmemBytes = smem_f; lc.stream = (cudaStream_t)stream;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeCooperative; at[0].val.cooperative = 1;
      lc.attrs = at; lc.numAttrs = 1;
      cudaError_t le = pl.prog.siren ? cudaLaunchKernelEx(&lc, mlp_tc_bwd_fused_kernel<true>, fa)
                                     : cudaLaunchKernelEx(&lc, mlp_tc_bwd_fused_kernel<false>, fa);
      g_nerfb_launches.fetch_add(1, std::memory_order_relaxed);
      if (le != cudaSuccess) { cudaGetLastError(); return (int)le; }
    }
    NERFB_LAUNCH(reduce_partials_fused_kernel, dim3(fa.w.n_items, 32), 256, 0, stream, fa.w, fa.fc);
    if (d.n_cls > 0) {
      const NerfOffsets o = nerf_offsets(d);
      NERFB_LAUNCH(parse_chain_kernel, 512, 256, 0, stream, params, (const float*)fa.w.aux, d.n_cls, o.w_p0, o.b_p0, o.w_p1, o.b_p1, grad);
    }
    return NERFB_OK;
  }

  if (pl.prog.siren) NERFB_LAUNCH(mlp_tc_dgrad_kernel<true>, pl.n_active, NUM_THREADS, smem_d, stream, da);
  else NERFB_LAUNCH(mlp_tc_dgrad_kernel<false>, pl.n_active, NUM_THREADS, smem_d, stream, da);

  pl.wg.fwd_rec = fwd_rec; pl.wg.bwd_rec = bb + pl.off_rec;
  pl.wg.partial = reinterpret_cast<float*>(bb + pl.off_partial); pl.wg.grad = grad;
  pl.wg.aux = reinterpret_cast<float*>(bb + pl.off_aux);
  if (d.n_cls > 0) NERFB_CUDA(cudaMemsetAsync(pl.wg.aux + 16 * 256, 0, 128 * 4, (cudaStream_t)stream));
  NERFB_LAUNCH(mlp_tc_wgrad_kernel, pl.n_active, W_THREADS, smem_w, stream, pl.wg);

  NERFB_LAUNCH(reduce_partials_kernel, dim3(pl.wg.n_items, 32), 256, 0, stream, pl.wg);
  if (d.n_cls > 0) {
    const NerfOffsets o = nerf_offsets(d);
    NERFB_LAUNCH(parse_chain_kernel, 512, 256, 0, stream, params, (const float*)pl.wg.aux, d.n_cls, o.w_p0, o.b_p0, o.w_p1, o.b_p1, grad);
  }
  return NERFB_OK;
}
