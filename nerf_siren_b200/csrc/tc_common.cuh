// sm_100a primitives shared by the tensor-core MLP kernels: mbarrier, bulk async copies (UBLKCP), tcgen05
// (UMMA issue, commit, TMEM alloc / load) and the 128-byte-swizzled K-major tile layout they all agree on.
// Hand-written inline PTX; bit layouts follow the PTX ISA's tcgen05 matrix/instruction descriptor tables.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <cstdio>

// where the forward kernel parked each tile's activations (byte offsets inside one tile record)
struct TcFwdLayout {
  int n_steps; int step_save_off[12]; int step_mask_off[12]; int save_x_off, save_dir_off; int tile_save_bytes;
  int64_t off_save, total;     // record array offset inside the workspace; end of the forward's workspace part
  int64_t off_weff;            // NeRF_3D: merged class-head matrix W_eff (fp32 16 x 256) | b_eff (16) in the forward's workspace
};

// Compositing in the forward kernel's epilogue (models/rendering.py:161-190 without `raw` ever reaching HBM):
// valid when a 128-sample tile holds whole rays (S in {32, 64, 128}), input mode RAYS, no class channels, inference.
// Arithmetic = kernel (4)'s, operation for operation.
struct TcComposite {
  int on, S, white_back; float noise_std;
  const float* noise;              // (N, S) or null
  float* weights;                  // (N, S) or null
  float* rgb; float* depth; float* opacity;   // (N,3) (N) (N); rgb / depth unused for a sigma-only pass
};

namespace tc {

constexpr int TILE_M = 128;                 // samples per tile = TMEM lanes
constexpr int KBLK = 64;                    // bf16 elements per 128-byte swizzle row
constexpr int BLK_BYTES = TILE_M * 128;     // one [128 x 64] bf16 block, swizzle-128B K-major   (16 KB)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// byte offset of element (row, col) inside a [rows x 64] bf16 block stored K-major with the 128B swizzle:
// 16-byte chunk index XOR (row mod 8).  Same image for TMA/UMMA SWIZZLE_128B, so one bulk copy moves a block.
__host__ __device__ __forceinline__ uint32_t swz_off(int row, int col) {
  return (uint32_t)row * 128u + ((((uint32_t)col >> 3) ^ ((uint32_t)row & 7u)) << 4) + (((uint32_t)col & 7u) << 1);
}

// ---- mbarrier --------------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug must trap (the host sees a launch failure) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > 4000000000LL) {   // ~2 s at 2 GHz
      printf("nerf_b200: mbarrier timeout (block %d thread %d bar %u parity %u)\n", (int)blockIdx.x, (int)threadIdx.x,
             smem_u32(bar), parity);
      const long long t1 = clock64();
      while (clock64() - t1 < 400000000LL) { }   // let the other stuck roles report before the grid is torn down
      __trap();
    }
  }
}

// Same, for roles that are not on the critical path (producers, store / encoder lanes): back off between polls so the
// spinning warp does not take issue slots from the epilogue warps of its SM sub-partition.
__device__ __forceinline__ void mbar_wait_relaxed(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    __nanosleep(100);
    if (clock64() - t0 > 4000000000LL) {
      printf("nerf_b200: mbarrier timeout (block %d thread %d bar %u parity %u)\n", (int)blockIdx.x, (int)threadIdx.x,
             smem_u32(bar), parity);
      const long long t1 = clock64();
      while (clock64() - t1 < 400000000LL) { }
      __trap();
    }
  }
}

// raw shared-address forms (no pointer arithmetic on the issuing thread's critical path)
__device__ __forceinline__ bool mbar_try_wait_addr(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait_addr(uint32_t bar, uint32_t parity) {
  if (mbar_try_wait_addr(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait_addr(bar, parity)) {
    if (clock64() - t0 > 4000000000LL) {
      printf("nerf_b200: mbarrier timeout (block %d thread %d bar %u parity %u)\n", (int)blockIdx.x, (int)threadIdx.x, bar, parity);
      const long long t1 = clock64();
      while (clock64() - t1 < 400000000LL) { }
      __trap();
    }
  }
}

// ---- shared-memory publish counters (cheaper to poll than an mbarrier) ----------------------------------------------
// producer: st.shared data -> fence.proxy.async -> __syncwarp -> lane 0: cnt_publish (release, +1 per warp)
// consumer: cnt_wait (acquire) until the monotonic counter reaches its target, then issues the async-proxy reads
__device__ __forceinline__ void cnt_publish(uint32_t cnt_addr) {
  asm volatile("red.release.cta.shared::cta.add.u32 [%0], 1;" ::"r"(cnt_addr) : "memory");
}
__device__ __forceinline__ uint32_t cnt_load(uint32_t cnt_addr) {
  uint32_t v;
  asm volatile("ld.acquire.cta.shared::cta.u32 %0, [%1];" : "=r"(v) : "r"(cnt_addr) : "memory");
  return v;
}
__device__ __forceinline__ void cnt_wait(uint32_t cnt_addr, uint32_t target) {
  if ((int32_t)(cnt_load(cnt_addr) - target) >= 0) return;
  const long long t0 = clock64();
  while ((int32_t)(cnt_load(cnt_addr) - target) < 0) {
    if (clock64() - t0 > 4000000000LL) {
      printf("nerf_b200: counter timeout (block %d thread %d cnt %u target %u)\n", (int)blockIdx.x, (int)threadIdx.x, cnt_addr, target);
      const long long t1 = clock64();
      while (clock64() - t1 < 400000000LL) { }
      __trap();
    }
  }
}
__device__ __forceinline__ void cnt_wait_relaxed(uint32_t cnt_addr, uint32_t target) {
  const long long t0 = clock64();
  while ((int32_t)(cnt_load(cnt_addr) - target) < 0) {
    __nanosleep(100);
    if (clock64() - t0 > 4000000000LL) {
      printf("nerf_b200: counter timeout (block %d thread %d cnt %u target %u)\n", (int)blockIdx.x, (int)threadIdx.x, cnt_addr, target);
      const long long t1 = clock64();
      while (clock64() - t1 < 400000000LL) { }
      __trap();
    }
  }
}

// ---- bulk async copies (TMA engine, no tensor map: contiguous images) -------------------------------------------
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* gdst, const void* smem_src, uint32_t bytes) {
#ifdef NERFB_EXP_NOSTORE
  return;
#endif
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
               ::"l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read2() { asm volatile("cp.async.bulk.wait_group.read 2;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// generic-proxy smem writes -> visible to the async proxy (UMMA operand fetch, bulk stores)
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// same copy with an L2 eviction-priority hint (streamed-once operands must not push the L2-resident ring out)
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ void bulk_g2s_hint(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar, uint64_t policy) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
               ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "l"(policy) : "memory");
}
__device__ __forceinline__ void bulk_wait_all3() { asm volatile("cp.async.bulk.wait_group 3;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all7() { asm volatile("cp.async.bulk.wait_group 7;" ::: "memory"); }
__device__ __forceinline__ uint4 ld_relaxed_gpu_v4(const uint32_t* p) {
  uint4 v;
  asm volatile("ld.relaxed.gpu.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void fence_acq_rel_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }
// generic <-> async proxy ordering for ALL state spaces (global data written / read by bulk copies and by ld/st)
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

// ---- cross-CTA flags in global memory (co-resident CTAs of ONE persistent grid: producer / consumer hand-off through L2) ----
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void red_release_gpu_add(uint32_t* p, uint32_t v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// bounded spin until *p >= target (monotonic counters; wrap-safe compare); traps instead of hanging the GPU
__device__ __forceinline__ void gflag_wait_ge(const uint32_t* p, uint32_t target) {
  if ((int32_t)(ld_acquire_gpu(p) - target) >= 0) return;
  const long long t0 = clock64();
  while ((int32_t)(ld_acquire_gpu(p) - target) < 0) {
    __nanosleep(64);
    if (clock64() - t0 > 4000000000LL) {
      printf("nerf_b200: global flag timeout (block %d thread %d flag %p target %u have %u)\n", (int)blockIdx.x, (int)threadIdx.x,
             (const void*)p, target, ld_acquire_gpu(p));
      const long long t1 = clock64();
      while (clock64() - t1 < 400000000LL) { }
      __trap();
    }
  }
}

// ---- tcgen05 ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tmem_alloc(uint32_t* smem_slot, uint32_t cols) {    // one full warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)), "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {       // same warp
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}

// Shared-memory matrix descriptor, K-major, SWIZZLE_128B: start>>4 | LBO(=1, ignored) | SBO = 1024 B (8 rows x 128 B)
// | version 1 (Blackwell) | layout type 2.
__device__ __forceinline__ uint64_t umma_desc_k128(uint32_t smem_addr) {
  return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) |
         ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}
// MN-major, SWIZZLE_128B: atom = 64 MN-elements (128 B) x 8 K-rows; LBO = byte stride between 64-element MN blocks,
// SBO = byte stride between 8-row K groups (1024 B for a dense [K x 64] image).
__device__ __forceinline__ uint64_t umma_desc_mn128(uint32_t smem_addr, uint32_t lbo_bytes) {
  return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16) |
         ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}
// Instruction descriptor, kind::f16: D = f32 (bits 4-5 = 1), A = B = bf16 (bits 7-9, 10-12 = 1), a_major bit 15,
// b_major bit 16 (0 = K-major, 1 = MN-major), N>>3 at bits 17-22, M>>4 at bits 24-28.
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int M, int N, int a_mn = 0, int b_mn = 0) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// D[tmem] (+)= A[smem] * B[smem]^T, issued by ONE thread on behalf of the CTA.
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
// One 64-wide K-block = four K=16 UMMAs on consecutive 32-byte slices of the swizzled rows; descriptors are
// {low word, 0x40004040}: SBO = 1024 B, version 1, SWIZZLE_128B (see umma_desc_k128); a 32-byte step is +2 in the low word.
__device__ __forceinline__ void umma_bf16_k64(uint32_t d_tmem, uint32_t a_lo, uint32_t b_lo, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, t;\n\t.reg .b64 a0, b0, a1, b1, a2, b2, a3, b3;\n\t"
      "setp.ne.b32 p, %4, 0;\n\tsetp.eq.b32 t, 0, 0;\n\t"
      "mov.b64 a0, {%1, 0x40004040};\n\tmov.b64 b0, {%2, 0x40004040};\n\t"
      "add.u64 a1, a0, 2;\n\tadd.u64 b1, b0, 2;\n\tadd.u64 a2, a0, 4;\n\tadd.u64 b2, b0, 4;\n\tadd.u64 a3, a0, 6;\n\tadd.u64 b3, b0, 6;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], a0, b0, %3, p;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], a1, b1, %3, t;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], a2, b2, %3, t;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], a3, b3, %3, t;\n\t}"
      ::"r"(d_tmem), "r"(a_lo), "r"(b_lo), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit_addr(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// mbarrier arrive once every previously issued UMMA of this thread has completed (implies fence::before_thread_sync)
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// TMEM -> registers: this warp's 32 lanes x 32 consecutive fp32 columns
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr) : "memory");
}
// TMEM -> registers: this warp's 32 lanes x 16 consecutive fp32 columns
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// two fp32 -> packed bf16x2 (lo = a, hi = b), optionally through ReLU, round-to-nearest-even
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
  return r;
}
__device__ __forceinline__ uint32_t pack_bf16_relu(float a, float b) {
  uint32_t r;
  asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
  return r;
}
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
  uint32_t r;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(r) : "r"(addr));
  return r;
}
__device__ __forceinline__ void mbar_arrive_addr(uint32_t bar_addr) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar_addr) : "memory");
}
__device__ __forceinline__ uint4 lds128u(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr));
  return r;
}
__device__ __forceinline__ float4 lds128f(uint32_t addr) {
  float4 r;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(addr));
  return r;
}
__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

}  // namespace tc
