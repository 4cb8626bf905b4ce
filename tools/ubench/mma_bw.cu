// Microbenchmark: tcgen05.mma issue rate from shared memory (M=128, N=256/128, K=16, bf16) alone and with 8 epilogue
// warps concurrently draining TMEM -> relu-pack -> st.shared, and with bulk copies refilling a ring: the realistic
// per-SM ceiling for the fused MLP chain.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../nerf_siren_b200/csrc/tc_common.cuh"
using namespace tc;

struct __align__(1024) Sm {
  uint8_t a[4][BLK_BYTES];       // 64 KB activations
  uint8_t b[3][256 * 128];       // 96 KB weights
  uint8_t out[2][BLK_BYTES];     // epilogue target
  float bias[256];
  uint64_t done, cp, cp2;
  uint32_t tbase;
};

// mode bit0: epilogue warps active; bit1: bulk copies active; bits 2-3: epilogue variant (0 plain, 1 +bias FADD2, 2 +bias FADD)
__global__ void __launch_bounds__(320, 1) k(int iters, int N, int mode, int fast, const uint8_t* gsrc, long long* cycles, uint32_t* sink) {
  extern __shared__ uint8_t dyn[];
  Sm& sm = *reinterpret_cast<Sm*>(((uintptr_t)dyn + 1023) & ~(uintptr_t)1023);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { mbar_init(&sm.done, 1); mbar_init(&sm.cp, 1); mbar_init(&sm.cp2, 1); fence_barrier_init(); }
  if (warp == 1) tmem_alloc(&sm.tbase, 512);
  for (int i = threadIdx.x; i < (int)(sizeof(sm.a) + sizeof(sm.b)) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(sm.a)[i] = 0;
  fence_proxy_async();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tmem = sm.tbase;
  __shared__ volatile int stop;
  if (threadIdx.x == 0) stop = 0;
  __syncthreads();
  if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(128, N);
      const long long t0 = clock64();
      uint32_t ph = 0;
      for (int it = 0; it < iters; ++it) {
        // one "layer": K = 256 -> 16 MMAs, into alternating accumulators
        const uint32_t d = tmem + (it & 1) * 256;
        if (fast) {
          for (int kb = 0; kb < 4; ++kb) {
            const uint64_t ad = umma_desc_k128(smem_u32(sm.a[kb])), bd = umma_desc_k128(smem_u32(sm.b[kb % 3]));
            asm volatile(
                "{\n\t.reg .pred p, t;\n\t.reg .b64 a1, b1, a2, b2, a3, b3;\n\t"
                "setp.ne.b32 p, %4, 0;\n\tsetp.eq.b32 t, 0, 0;\n\t"
                "add.u64 a1, %1, 2;\n\tadd.u64 b1, %2, 2;\n\tadd.u64 a2, %1, 4;\n\tadd.u64 b2, %2, 4;\n\tadd.u64 a3, %1, 6;\n\tadd.u64 b3, %2, 6;\n\t"
                "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
                "tcgen05.mma.cta_group::1.kind::f16 [%0], a1, b1, %3, t;\n\t"
                "tcgen05.mma.cta_group::1.kind::f16 [%0], a2, b2, %3, t;\n\t"
                "tcgen05.mma.cta_group::1.kind::f16 [%0], a3, b3, %3, t;\n\t}"
                ::"r"(d), "l"(ad), "l"(bd), "r"(idesc), "r"(kb ? 1u : 0u) : "memory");
          }
        } else
        for (int kb = 0; kb < 4; ++kb)
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            umma_bf16(d, umma_desc_k128(smem_u32(sm.a[kb]) + kk * 32), umma_desc_k128(smem_u32(sm.b[kb % 3]) + kk * 32), idesc, (kb | kk) ? 1u : 0u);
        if ((it & 7) == 7) { umma_commit(&sm.done); mbar_wait(&sm.done, ph); ph ^= 1; }
      }
      const long long t1 = clock64();
      cycles[blockIdx.x] = t1 - t0;
      stop = 1;
    }
  } else if (warp == 0) {
    if (lane == 0 && (mode & 2)) {            // weight-ring refill at full rate: 32 KB slabs, 2 in flight, into the B stages
      uint32_t ph = 0; int n = 0;
      mbar_expect_tx(&sm.cp, 32768);
      bulk_g2s(sm.b[0], gsrc, 32768, &sm.cp);
      while (!stop) {
        ++n;
        mbar_expect_tx(&sm.cp2, 32768);
        bulk_g2s(sm.b[1], gsrc + 32768, 32768, &sm.cp2);
        mbar_wait(&sm.cp, ph);
        mbar_expect_tx(&sm.cp, 32768);
        bulk_g2s(sm.b[2], gsrc + 65536, 32768, &sm.cp);
        mbar_wait(&sm.cp2, ph);
        ph ^= 1; ++n;
      }
      mbar_wait(&sm.cp, ph);
      sink[1400 + blockIdx.x] = n;
    }
  } else if (mode & 1) {
    const int q = warp & 3, h = (warp - 2) >> 2, row = q * 32 + lane;
    uint32_t acc = 0;
    while (!stop) {
      for (int c = h * 128; c < (h + 1) * 128; c += 32) {
        uint32_t v[32];
        tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)c, v);
        const int var = (mode >> 2) & 3;
        if (var) {
          float4 bb[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) bb[i] = lds128f(smem_u32(sm.bias) + c * 4 + 16 * i);
          tmem_ld_wait();
          if (var == 1) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float2 r0 = __fadd2_rn(make_float2(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1])), make_float2(bb[i].x, bb[i].y));
              const float2 r1 = __fadd2_rn(make_float2(__uint_as_float(v[4 * i + 2]), __uint_as_float(v[4 * i + 3])), make_float2(bb[i].z, bb[i].w));
              v[4 * i] = __float_as_uint(r0.x); v[4 * i + 1] = __float_as_uint(r0.y); v[4 * i + 2] = __float_as_uint(r1.x); v[4 * i + 3] = __float_as_uint(r1.y);
            }
          } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              v[4 * i] = __float_as_uint(__uint_as_float(v[4 * i]) + bb[i].x); v[4 * i + 1] = __float_as_uint(__uint_as_float(v[4 * i + 1]) + bb[i].y);
              v[4 * i + 2] = __float_as_uint(__uint_as_float(v[4 * i + 2]) + bb[i].z); v[4 * i + 3] = __float_as_uint(__uint_as_float(v[4 * i + 3]) + bb[i].w);
            }
          }
        } else {
          tmem_ld_wait();
        }
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const int col = (c & 63) + 8 * g;
          const uint32_t addr = smem_u32(sm.out[0]) + (uint32_t)row * 128u + ((((uint32_t)col >> 3) ^ ((uint32_t)row & 7u)) << 4);
          sts128(addr, pack_bf16_relu(__uint_as_float(v[8 * g]), __uint_as_float(v[8 * g + 1])),
                 pack_bf16_relu(__uint_as_float(v[8 * g + 2]), __uint_as_float(v[8 * g + 3])),
                 pack_bf16_relu(__uint_as_float(v[8 * g + 4]), __uint_as_float(v[8 * g + 5])),
                 pack_bf16_relu(__uint_as_float(v[8 * g + 6]), __uint_as_float(v[8 * g + 7])));
        }
        acc++;
      }
      fence_proxy_async();
    }
    if (acc == 0xFFFFFFFFu) sink[threadIdx.x] = acc;
    if (lane == 0 && q == 0 && h == 0) sink[1024 + blockIdx.x] = acc;   // drains per warp: epilogue rate
  }
  tc_fence_before(); __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

int main() {
  long long* cyc; uint32_t* sink; uint8_t* g;
  const int grid = 148;
  cudaMalloc(&cyc, 8 * grid); cudaMalloc(&sink, 4 * 2048); cudaMalloc(&g, 1 << 20); cudaMemset(g, 0, 1 << 20);
  const size_t smem = sizeof(Sm) + 1024;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const int iters = 4000;
  for (int N : {256, 128, 32})
    for (int fast : {0, 1}) for (int mode : {0, 7}) {
      cudaMemset(sink, 0, 4 * 2048);
      k<<<grid, 320, smem>>>(16, N, mode, fast, g, cyc, sink);
      k<<<grid, 320, smem>>>(iters, N, mode, fast, g, cyc, sink);
      cudaError_t e = cudaDeviceSynchronize();
      long long hc; uint32_t dr;
      cudaMemcpy(&hc, cyc, 8, cudaMemcpyDeviceToHost);
      cudaMemcpy(&dr, sink + 1024, 4, cudaMemcpyDeviceToHost);
      uint32_t ncp; cudaMemcpy(&ncp, sink + 1400, 4, cudaMemcpyDeviceToHost);
      printf("fast %d bulk fill %.1f B/cycle  ", fast, hc ? (double)ncp * 32768 / hc : 0.0);
      printf("N=%3d mode=%d (epi %d, bulk %d, variant %d): %s  %.1f cycles per 128xNx16 MMA (floor %d); epilogue warp: %.0f cycles per 32x32 block\n", N, mode,
             mode & 1, (mode >> 1) & 1, mode >> 2, cudaGetErrorString(e), (double)hc / iters / 16, N / 2, dr ? (double)hc / dr : 0.0);
    }
  return 0;
}
