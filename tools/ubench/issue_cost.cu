// Microbenchmark: cycles one thread spends per K-block of the fused-MLP issue loop: 4 UMMAs (N = 32 so the tensor pipe
// is not the limit), + tcgen05.commit, + 2 mbarrier.try_wait on completed barriers, + a shared-memory op fetch.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../nerf_siren_b200/csrc/tc_common.cuh"
using namespace tc;

struct __align__(1024) Sm { uint8_t a[BLK_BYTES]; uint8_t b[BLK_BYTES]; uint64_t bar[8]; uint4 ops[64]; uint32_t tbase; };

template <int variant>
__global__ void __launch_bounds__(64, 1) k(int iters, int N, int fullwarp, long long* cycles) {
  extern __shared__ __align__(1024) uint8_t dyn[];
  Sm& sm = *reinterpret_cast<Sm*>(dyn);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { for (int i = 0; i < 8; ++i) mbar_init(&sm.bar[i], 1); fence_barrier_init(); }
  if (warp == 1) tmem_alloc(&sm.tbase, 512);
  for (int i = threadIdx.x; i < 2 * BLK_BYTES / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(sm.a)[i] = 0;
  for (int i = threadIdx.x; i < 64; i += blockDim.x) sm.ops[i] = make_uint4(i, i, i, i);
  fence_proxy_async();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tmem = sm.tbase;
  if (threadIdx.x == 0) { mbar_arrive(&sm.bar[0]); mbar_arrive(&sm.bar[1]); }   // phase 0 of bars 0,1 complete: try_wait(parity 0) succeeds
  __syncthreads();
  if (warp == 1 && (lane == 0 || fullwarp)) {
    const uint32_t idesc = umma_idesc_bf16(128, N);
    const uint32_t a_lo = ((smem_u32(sm.a) >> 4) & 0x3FFFu) | (1u << 16), b_lo = ((smem_u32(sm.b) >> 4) & 0x3FFFu) | (1u << 16);
    const uint32_t bar0 = smem_u32(&sm.bar[0]), bar1 = smem_u32(&sm.bar[1]), cbar = smem_u32(&sm.bar[2]), ops = smem_u32(sm.ops);
    uint32_t acc = 0;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      if (variant & 4) { const uint4 o = lds128u(ops + 16u * (it & 63)); acc += o.x; }
      if (variant & 2) {
        const bool o1 = mbar_try_wait_addr(bar0, 0), o2 = mbar_try_wait_addr(bar1, 0);
        if (!o1) mbar_wait_addr(bar0, 0);
        if (!o2) mbar_wait_addr(bar1, 0);
        tc_fence_after();
      }
      if (lane == 0) {
        if (variant & 8) umma_bf16_k64(tmem, a_lo, b_lo, idesc, it ? 1u : 0u);
        if (variant & 1) umma_commit_addr(cbar);
      }
      if (fullwarp) __syncwarp();
    }
    const long long t1 = clock64();
    if (lane == 0) { cycles[0] = t1 - t0; cycles[1] = acc; umma_commit_addr(smem_u32(&sm.bar[3])); }
    mbar_wait(&sm.bar[3], 0);
  }
  tc_fence_before(); __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

int main() {
  long long* cyc; cudaMalloc(&cyc, 64);
  const int iters = 4000;
  const int N = 32;
  for (int fullwarp : {0, 1})
    for (int variant : {0, 8, 9, 10, 15, 1, 2, 4}) {
      auto launch = [&](auto kern) { cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Sm) + 1024); kern<<<1, 64, sizeof(Sm) + 1024>>>(iters, N, fullwarp, cyc); };
      switch (variant) { case 0: launch(k<0>); break; case 8: launch(k<8>); break; case 9: launch(k<9>); break; case 10: launch(k<10>); break;
        case 15: launch(k<15>); break; case 1: launch(k<1>); break; case 2: launch(k<2>); break; default: launch(k<4>); }
      cudaError_t e = cudaDeviceSynchronize();
      long long h[2]; cudaMemcpy(h, cyc, 16, cudaMemcpyDeviceToHost);
      printf("fullwarp %d N=%3d  mma %d commit %d try_wait x2 %d lds %d: %s  %.1f cycles per K-block\n", fullwarp, N, (variant >> 3) & 1, variant & 1, (variant >> 1) & 1,
             (variant >> 2) & 1, cudaGetErrorString(e), (double)h[0] / iters);
    }
  return 0;
}
