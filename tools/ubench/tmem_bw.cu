// Microbenchmark: TMEM -> register read throughput (tcgen05.ld 32x32b.x32 / .x64 / .x128) per SM, and the
// epilogue pattern TMEM -> cvt.relu.bf16x2 -> st.shared.v4.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../nerf_siren_b200/csrc/tc_common.cuh"
using namespace tc;

template <int X> __device__ __forceinline__ void ld(uint32_t taddr, uint32_t* v);
template <> __device__ __forceinline__ void ld<32>(uint32_t taddr, uint32_t* v) {
  uint32_t (&r)[32] = *reinterpret_cast<uint32_t(*)[32]>(v);
  tmem_ld32(taddr, r);
}
template <> __device__ __forceinline__ void ld<64>(uint32_t taddr, uint32_t* v) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x64.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, "
      "%32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, %48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63}, [%64];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
        "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31]),
        "=r"(v[32]), "=r"(v[33]), "=r"(v[34]), "=r"(v[35]), "=r"(v[36]), "=r"(v[37]), "=r"(v[38]), "=r"(v[39]), "=r"(v[40]), "=r"(v[41]), "=r"(v[42]), "=r"(v[43]), "=r"(v[44]), "=r"(v[45]), "=r"(v[46]), "=r"(v[47]),
        "=r"(v[48]), "=r"(v[49]), "=r"(v[50]), "=r"(v[51]), "=r"(v[52]), "=r"(v[53]), "=r"(v[54]), "=r"(v[55]), "=r"(v[56]), "=r"(v[57]), "=r"(v[58]), "=r"(v[59]), "=r"(v[60]), "=r"(v[61]), "=r"(v[62]), "=r"(v[63])
      : "r"(taddr) : "memory");
}

// mode 0: ld only (xor-reduce); mode 1: ld + relu-pack + st.shared (epilogue pattern)
template <int X, int MODE>
__global__ void __launch_bounds__(256, 1) k(int iters, int cols_total, long long* cycles, uint32_t* sink) {
  __shared__ uint32_t tbase;
  extern __shared__ uint8_t dyn[];
  uint8_t* blk = (uint8_t*)(((uintptr_t)dyn + 1023) & ~(uintptr_t)1023);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) tmem_alloc(&tbase, 512);
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tmem = tbase;
  const int q = warp & 3, h = warp >> 2;            // 8 warps: lane quarter q, column half h
  const int row = q * 32 + lane;
  uint32_t acc = 0;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    // one "accumulator drain": 128 lanes x cols_total columns, split in two column halves across warp groups
    for (int c = h * (cols_total / 2); c < (h + 1) * (cols_total / 2); c += X) {
      uint32_t v[X];
      ld<X>(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)c, v);
      tmem_ld_wait();
      if (MODE == 0) {
#pragma unroll
        for (int i = 0; i < X; ++i) acc ^= v[i];
      } else {
#pragma unroll
        for (int g = 0; g < X / 8; ++g) {
          const int col = (c & 63) + 8 * g;           // position inside a 64-column block
          const uint32_t addr = smem_u32(blk) + (uint32_t)(c >> 6) * BLK_BYTES + (uint32_t)row * 128u +
                                ((((uint32_t)col >> 3) ^ ((uint32_t)row & 7u)) << 4);
          sts128(addr, pack_bf16_relu(__uint_as_float(v[8 * g]), __uint_as_float(v[8 * g + 1])),
                 pack_bf16_relu(__uint_as_float(v[8 * g + 2]), __uint_as_float(v[8 * g + 3])),
                 pack_bf16_relu(__uint_as_float(v[8 * g + 4]), __uint_as_float(v[8 * g + 5])),
                 pack_bf16_relu(__uint_as_float(v[8 * g + 6]), __uint_as_float(v[8 * g + 7])));
        }
      }
    }
    if (MODE == 1) { fence_proxy_async(); __syncwarp(); }
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  if (acc == 0x12345678u) sink[threadIdx.x] = acc;
  tc_fence_before(); __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

template <int X, int MODE> void run(const char* name, int grid) {
  long long* cyc; uint32_t* sink;
  cudaMalloc(&cyc, 8 * grid); cudaMalloc(&sink, 4096);
  const int iters = 2000, cols = 256;
  const size_t smem = 4 * BLK_BYTES + 1024;
  cudaFuncSetAttribute(k<X, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  k<X, MODE><<<grid, 256, smem>>>(10, cols, cyc, sink);
  k<X, MODE><<<grid, 256, smem>>>(iters, cols, cyc, sink);
  cudaError_t e = cudaDeviceSynchronize();
  long long hc; cudaMemcpy(&hc, cyc, 8, cudaMemcpyDeviceToHost);
  const double bytes = (double)iters * 128 * cols * 4;
  printf("%-28s grid %3d: %s  %.1f cycles per 128x%d drain, %.1f B/cycle/SM (TMEM read)\n", name, grid,
         cudaGetErrorString(e), (double)hc / iters, cols, bytes / hc);
}

int main() {
  run<32, 0>("ld.x32 only", 1);
  run<64, 0>("ld.x64 only", 1);
  run<32, 1>("ld.x32 + relu-pack + sts", 1);
  run<64, 1>("ld.x64 + relu-pack + sts", 1);
  run<32, 1>("ld.x32 + relu-pack + sts", 148);
  return 0;
}
