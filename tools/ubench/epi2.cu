// Microbenchmark for the tile-pair forward (round 2): cost of ONE accumulator drain (128 rows x 256 fp32 columns ->
// + bias -> ReLU -> bf16 -> 128B-swizzled shared-memory blocks) by 16 epilogue warps, for two work splits:
//   OLD  warp = (lane quarter q, 16-column quarter c4): every 64-column block is drained by all 16 warps and published
//        on its own (4 x {tcgen05.ld.x16, fence.proxy.async, arrive} per thread and layer) -- mlp_tc.cu today;
//   NEW  warp = (lane quarter q, block j): a thread drains 64 contiguous columns of its row (4 x ld.x16 pipelined or
//        2 x ld.x32), ONE fence and ONE arrive per layer -- possible once the next layer's MMAs no longer start per block.
// Optionally a background warp keeps the tensor pipe busy with 128 x 256 x 16 MMAs into the other TMEM half (shared-memory
// bandwidth contention as in the real kernel).  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o epi2 epi2.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../nerf_siren_b200/csrc/tc_common.cuh"
using namespace tc;

struct __align__(1024) Sm {
  uint8_t a[4][BLK_BYTES];
  uint8_t a2[4][BLK_BYTES];     // A operand of the background MMAs
  uint8_t w[2][256 * 128];      // B operand slabs of the background MMAs
  float bias[256];
  uint64_t ready[4];
  uint64_t mma_done[2];
  uint32_t tmem_base;
  volatile int stop;
  unsigned long long n_mma;
};

template <int MODE>   // 0 OLD, 1 NEW 4 x ld16, 2 NEW 2 x ld32
__global__ void __launch_bounds__(18 * 32, 1) k(int iters, int bg, long long* cycles, unsigned long long* mma_count) {
  extern __shared__ __align__(1024) uint8_t raw[];
  Sm& sm = *reinterpret_cast<Sm*>(raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 4; ++i) mbar_init(&sm.ready[i], 16);
    mbar_init(&sm.mma_done[0], 1); mbar_init(&sm.mma_done[1], 1);
    sm.stop = 0; sm.n_mma = 0;
    fence_barrier_init();
  }
  for (int i = threadIdx.x; i < 256; i += blockDim.x) sm.bias[i] = 0.001f * i;
  for (int i = threadIdx.x; i < (int)(sizeof(sm.a2) + sizeof(sm.w)) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(sm.a2)[i] = 0x3c003c00u;
  if (warp == 16) tmem_alloc(&sm.tmem_base, 512);
  fence_proxy_async();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tmem = sm.tmem_base;
  if (warp == 17) {
    if (lane == 0 && bg) {          // background MMAs into columns [256, 512): batches of 16 (one layer's worth), commit, wait
      const uint32_t idesc = umma_idesc_bf16(128, 256);
      uint32_t ph[2] = {0, 0}; unsigned long long n = 0; uint32_t b = 0;
      while (!sm.stop) {                // two batches in flight: the pipe never drains between batches
        for (int kb = 0; kb < 4; ++kb)
          umma_bf16_k64(tmem + 256u, ((smem_u32(sm.a2[kb]) >> 4) & 0x3FFFu) | (1u << 16), ((smem_u32(sm.w[kb & 1]) >> 4) & 0x3FFFu) | (1u << 16), idesc, kb ? 1u : 0u);
        umma_commit(&sm.mma_done[b]);
        n += 16;
        b ^= 1;
        if (n > 16) { mbar_wait(&sm.mma_done[b], ph[b]); ph[b] ^= 1; }
      }
      b ^= 1; mbar_wait(&sm.mma_done[b], ph[b]);
      sm.n_mma = n;
    }
  } else if (warp < 16) {
    const int q = warp & 3, g = warp >> 2;
    const int row = q * 32 + lane;
    const uint32_t row_off = (uint32_t)row * 128u, row_x = (uint32_t)row & 7u;
    asm volatile("bar.sync 1, 512;");
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      if (MODE == 0) {
        const uint32_t t_lane = tmem + ((uint32_t)(q * 32) << 16) + 16u * (uint32_t)g;
        const uint32_t bias_base = smem_u32(sm.bias) + 64u * (uint32_t)g;
        uint32_t v[2][16];
        tmem_ld16(t_lane, v[0]);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float4 bb[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) bb[i] = lds128f(bias_base + 256u * (uint32_t)j + 16u * i);
          tmem_ld_wait();
          if (j + 1 < 4) tmem_ld16(t_lane + 64u * (uint32_t)(j + 1), v[(j + 1) & 1]);
          const uint32_t (&vv)[16] = v[j & 1];
          float f[16];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float2 r0 = __fadd2_rn(make_float2(__uint_as_float(vv[4 * i]), __uint_as_float(vv[4 * i + 1])), make_float2(bb[i].x, bb[i].y));
            const float2 r1 = __fadd2_rn(make_float2(__uint_as_float(vv[4 * i + 2]), __uint_as_float(vv[4 * i + 3])), make_float2(bb[i].z, bb[i].w));
            f[4 * i] = r0.x; f[4 * i + 1] = r0.y; f[4 * i + 2] = r1.x; f[4 * i + 3] = r1.y;
          }
          const uint32_t blk = smem_u32(sm.a[0]) + (uint32_t)j * BLK_BYTES + row_off;
#pragma unroll
          for (int ch = 0; ch < 2; ++ch)
            sts128(blk + ((((uint32_t)(2 * g + ch)) ^ row_x) << 4), pack_bf16_relu(f[8 * ch + 0], f[8 * ch + 1]), pack_bf16_relu(f[8 * ch + 2], f[8 * ch + 3]),
                   pack_bf16_relu(f[8 * ch + 4], f[8 * ch + 5]), pack_bf16_relu(f[8 * ch + 6], f[8 * ch + 7]));
          fence_proxy_async();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&sm.ready[j]);
        }
      } else {
        constexpr int X = (MODE == 1) ? 16 : 32;
        const uint32_t t_lane = tmem + ((uint32_t)(q * 32) << 16) + 64u * (uint32_t)g;
        const uint32_t bias_base = smem_u32(sm.bias) + 256u * (uint32_t)g;
        const uint32_t blk = smem_u32(sm.a[0]) + (uint32_t)g * BLK_BYTES + row_off;
        uint32_t v[2][X];
        if (X == 16) tmem_ld16(t_lane, reinterpret_cast<uint32_t(&)[16]>(v[0])); else tmem_ld32(t_lane, reinterpret_cast<uint32_t(&)[32]>(v[0]));
#pragma unroll
        for (int c = 0; c < 64 / X; ++c) {
          float4 bb[X / 4];
#pragma unroll
          for (int i = 0; i < X / 4; ++i) bb[i] = lds128f(bias_base + 4u * X * (uint32_t)c + 16u * i);
          tmem_ld_wait();
          if (c + 1 < 64 / X) {
            if (X == 16) tmem_ld16(t_lane + X * (c + 1), reinterpret_cast<uint32_t(&)[16]>(v[(c + 1) & 1]));
            else tmem_ld32(t_lane + X * (c + 1), reinterpret_cast<uint32_t(&)[32]>(v[(c + 1) & 1]));
          }
          const uint32_t* vv = v[c & 1];
          float f[X];
#pragma unroll
          for (int i = 0; i < X / 4; ++i) {
            const float2 r0 = __fadd2_rn(make_float2(__uint_as_float(vv[4 * i]), __uint_as_float(vv[4 * i + 1])), make_float2(bb[i].x, bb[i].y));
            const float2 r1 = __fadd2_rn(make_float2(__uint_as_float(vv[4 * i + 2]), __uint_as_float(vv[4 * i + 3])), make_float2(bb[i].z, bb[i].w));
            f[4 * i] = r0.x; f[4 * i + 1] = r0.y; f[4 * i + 2] = r1.x; f[4 * i + 3] = r1.y;
          }
#pragma unroll
          for (int ch = 0; ch < X / 8; ++ch)
            sts128(blk + ((((uint32_t)(c * (X / 8) + ch)) ^ row_x) << 4), pack_bf16_relu(f[8 * ch + 0], f[8 * ch + 1]), pack_bf16_relu(f[8 * ch + 2], f[8 * ch + 3]),
                   pack_bf16_relu(f[8 * ch + 4], f[8 * ch + 5]), pack_bf16_relu(f[8 * ch + 6], f[8 * ch + 7]));
        }
        fence_proxy_async();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.ready[0]);
      }
    }
    asm volatile("bar.sync 1, 512;");
    const long long t1 = clock64();
    if (threadIdx.x == 0) { cycles[blockIdx.x] = t1 - t0; sm.stop = 1; }
  }
  tc_fence_before(); __syncthreads();
  if (threadIdx.x == 0) mma_count[blockIdx.x] = sm.n_mma;
  if (warp == 16) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

template <int MODE> void run(const char* name, int grid, int bg) {
  long long* cyc; unsigned long long* cnt;
  cudaMalloc(&cyc, 8 * grid); cudaMalloc(&cnt, 8 * grid);
  const int iters = 4000;
  const size_t smem = sizeof(Sm) + 1024;
  cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  k<MODE><<<grid, 18 * 32, smem>>>(10, bg, cyc, cnt);
  k<MODE><<<grid, 18 * 32, smem>>>(iters, bg, cyc, cnt);
  cudaError_t e = cudaDeviceSynchronize();
  long long hc; unsigned long long hn;
  cudaMemcpy(&hc, cyc, 8, cudaMemcpyDeviceToHost); cudaMemcpy(&hn, cnt, 8, cudaMemcpyDeviceToHost);
  printf("%-22s grid %3d bg %d: %s  %.0f cycles per 128x256 drain", name, grid, bg, cudaGetErrorString(e), (double)hc / iters);
  if (bg) printf(", background MMAs: %.1f cycles per 128x256x16 MMA", hn ? (double)hc / (double)hn : 0.0);
  printf("\n");
}

int main() {
  for (int bg = 0; bg < 2; ++bg) {
    run<0>("OLD (q, c4) 4 fences", 1, bg);
    run<1>("NEW (q, j) 4 x ld16", 1, bg);
    run<2>("NEW (q, j) 2 x ld32", 1, bg);
  }
  run<0>("OLD (q, c4) 4 fences", 148, 1);
  run<1>("NEW (q, j) 4 x ld16", 148, 1);
  run<2>("NEW (q, j) 2 x ld32", 148, 1);
  return 0;
}
