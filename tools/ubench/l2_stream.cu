// Microbenchmark: every SM streams the same L2-resident weight blob (1.19 MB, like the fused MLP) into a shared-memory
// ring with bulk copies (UBLKCP); rate per SM vs ring depth / slab size, unicast vs cluster multicast (each CTA of the
// cluster fetches 1/csz of every slab and multicasts it to all).
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
#include <cooperative_groups.h>
#include "../../nerf_siren_b200/csrc/tc_common.cuh"
using namespace tc;
namespace cg = cooperative_groups;
// non-suspending poll: mbarrier.test_wait returns immediately (try_wait may park the thread for a HW-defined time)
__device__ __forceinline__ void mbar_spin(uint64_t* bar, uint32_t parity) {
  uint32_t ok = 0;
  while (!ok) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  }
}
#define WAIT(bar, par) do { if (n_prod < 0) mbar_spin(bar, par); else mbar_wait(bar, par); } while (0)

__device__ __forceinline__ void bulk_g2s_mc(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar, uint16_t mask) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;"
               ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "h"(mask) : "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint64_t* bar, uint32_t cta) {   // arrive on the same barrier in CTA `cta`
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(bar)), "r"(cta));
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
}

__global__ void __launch_bounds__(256, 1) k(const uint8_t* blob, int blob_bytes, int slab, int stages, int passes, int csz,
                                            long long* cycles, int n_prod) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem);            // [stages]
  uint64_t* empty = full + 32;                                    // [stages]  (<= 32 stages)
  uint8_t* ring = smem + 1024;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t rank = 0;
  if (csz > 1) rank = cg::this_cluster().block_rank();
  if (threadIdx.x == 0) {
    for (int i = 0; i < stages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], csz); }
    fence_barrier_init();
  }
  if (csz > 1) cg::this_cluster().sync(); else __syncthreads();
  const int n_slabs = blob_bytes / slab;
  const long long t0 = clock64();
  if (n_prod > 1 && n_prod < 100 && warp >= 2 && warp < 2 + n_prod && lane == 0) {     // several producer warps: warp 2 + p issues slabs g = p (mod n_prod)
    const int p0 = warp - 2;
    long long g = 0;
    for (int p = 0; p < passes; ++p)
      for (int s = 0; s < n_slabs; ++s, ++g) {
        if ((int)(g % n_prod) != p0) continue;
        const int st = (int)(g % stages); const uint32_t ph = (uint32_t)((g / stages) & 1);
        mbar_wait(&empty[st], ph ^ 1);
        mbar_expect_tx(&full[st], slab);
        bulk_g2s(ring + (size_t)st * slab, blob + (size_t)s * slab, slab, &full[st]);
      }
  } else if (n_prod >= 100 && warp == 0) {      // one slab = several smaller bulk copies on the same barrier: one lane sequentially (1xx) or one lane each (2xx)
    const bool par = n_prod >= 200;
    const int ns = par ? n_prod - 200 : n_prod - 100;
    const uint32_t piece = slab / ns;
    int st = 0; uint32_t ph = 0;
    for (int p = 0; p < passes; ++p)
      for (int s = 0; s < n_slabs; ++s) {
        if (lane == 0) { mbar_wait(&empty[st], ph ^ 1); mbar_expect_tx(&full[st], slab); }
        __syncwarp();
        if (par) { if (lane < ns) bulk_g2s(ring + (size_t)st * slab + lane * piece, blob + (size_t)s * slab + lane * piece, piece, &full[st]); }
        else if (lane == 0) for (int k = 0; k < ns; ++k) bulk_g2s(ring + (size_t)st * slab + k * piece, blob + (size_t)s * slab + k * piece, piece, &full[st]);
        __syncwarp();
        if (++st == stages) { st = 0; ph ^= 1; }
      }
  } else if (n_prod <= 1 && warp == 0 && lane == 0) {            // producer
    int st = 0; uint32_t ph = 0;
    const uint32_t part = slab / csz;
    for (int p = 0; p < passes; ++p)
      for (int s = 0; s < n_slabs; ++s) {
        WAIT(&empty[st], ph ^ 1);
        mbar_expect_tx(&full[st], slab);
        if (csz == 1) bulk_g2s(ring + (size_t)st * slab, blob + (size_t)s * slab, slab, &full[st]);
        else bulk_g2s_mc(ring + (size_t)st * slab + rank * part, blob + (size_t)s * slab + rank * part, part, &full[st], (uint16_t)((1u << csz) - 1));
        if (++st == stages) { st = 0; ph ^= 1; }
      }
  } else if (warp == 1 && lane == 0) {     // consumer: frees the slab right away (in every CTA of the cluster)
    int st = 0; uint32_t ph = 0;
    for (int p = 0; p < passes; ++p)
      for (int s = 0; s < n_slabs; ++s) {
        WAIT(&full[st], ph);
        if (csz == 1) mbar_arrive(&empty[st]);
        else for (int c = 0; c < csz; ++c) mbar_arrive_cluster(&empty[st], c);
        if (++st == stages) { st = 0; ph ^= 1; }
      }
    cycles[blockIdx.x] = clock64() - t0;      // the consumer's clock: the last slab has landed
  }
  __syncthreads();
  const long long t1 = clock64();
  (void)t1;
  if (csz > 1) cg::this_cluster().sync();
}

int main(int argc, char** argv) {
  const int blob_bytes = 1179648;   // 36 slabs of 32 KB
  uint8_t* blob; long long* cyc;
  cudaMalloc(&blob, blob_bytes); cudaMemset(blob, 1, blob_bytes); cudaMalloc(&cyc, 8 * 148);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaFuncSetAttribute(k, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  struct Cfg { int slab, stages, csz, n_prod; };
  const Cfg cfgs[] = {{32768, 2, 1}, {32768, 3, 1}, {32768, 4, 1}, {32768, 6, 1}, {16384, 6, 1}, {16384, 12, 1}, {8192, 12, 1},
                      {32768, 3, 2}, {32768, 4, 2}, {32768, 3, 4}, {32768, 4, 4}, {32768, 3, 1}};
  // custom sweep: l2_stream <grid> slab,stages[,csz] ...   (one CTA per SM is forced by a 200 KB shared-memory request)
  Cfg custom[32]; int n_custom = 0, grid_override = 0;
  if (argc > 2) { grid_override = atoi(argv[1]); for (int i = 2; i < argc && n_custom < 32; ++i) { Cfg c{0, 0, 1, 1}; sscanf(argv[i], "%d,%d,%d,%d", &c.slab, &c.stages, &c.csz, &c.n_prod); if (c.stages > 32) c.stages = 32; custom[n_custom++] = c; } }
  const Cfg* list = n_custom ? custom : cfgs; const int n_list = n_custom ? n_custom : (int)(sizeof(cfgs) / sizeof(cfgs[0]));
  for (int ci = 0; ci < n_list; ++ci) {
    const Cfg& c = list[ci];
    const int passes = 40;
    const size_t smem = n_custom ? 200 * 1024 : 1024 + (size_t)c.slab * c.stages;
    cudaLaunchConfig_t lc = {};
    const int grid = grid_override ? grid_override : (148 / c.csz) * c.csz;
    lc.gridDim = dim3(grid); lc.blockDim = dim3(256); lc.dynamicSmemBytes = smem; lc.stream = 0;
    cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim = {(unsigned)c.csz, 1, 1};
    lc.attrs = at; lc.numAttrs = 1;
    int ncl = -1;
    cudaOccupancyMaxActiveClusters(&ncl, k, &lc);
    for (int rep = 0; rep < 2; ++rep) cudaLaunchKernelEx(&lc, k, (const uint8_t*)blob, blob_bytes, c.slab, c.stages, passes, c.csz, cyc, c.n_prod != 0 ? c.n_prod : 1);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, cyc, 8 * grid, cudaMemcpyDeviceToHost);
    long long mx = 0; for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
    printf("producers %d slab %5d stages %2d cluster %d (max active clusters %3d, grid %d): %s  %.1f B/cycle/SM delivered, chip %.0f B/cycle\n", c.n_prod > 0 ? c.n_prod : 1, c.slab, c.stages,
           c.csz, ncl, grid, cudaGetErrorString(e), (double)blob_bytes * passes / mx, (double)blob_bytes * passes / mx * grid);
  }
  return 0;
}
