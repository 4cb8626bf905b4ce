// Calibration: clock64 cycles per dependent FFMA (expected ~4) and per empty volatile-asm loop iteration, single lane.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(int iters, int mode, float* out, long long* cyc) {
  if (threadIdx.x != 32) return;
  float x = out[0];
  const long long t0 = clock64();
  if (mode == 0) { for (int i = 0; i < iters; ++i) x = fmaf(x, 1.0001f, 0.5f); }
  else if (mode == 1) { for (int i = 0; i < iters; ++i) asm volatile("" ::: "memory"); }
  else { for (int i = 0; i < iters; ++i) { asm volatile("" ::: "memory"); if (out[1] == 123.f) out[2] = i; } }
  const long long t1 = clock64();
  out[3] = x; cyc[0] = t1 - t0;
}
int main() {
  float* o; long long* c; cudaMalloc(&o, 64); cudaMemset(o, 0, 64); cudaMalloc(&c, 8);
  for (int mode = 0; mode < 3; ++mode) {
    k<<<1, 64>>>(100000, mode, o, c); cudaDeviceSynchronize();
    long long h; cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
    printf("mode %d: %.2f cycles per iteration\n", mode, (double)h / 100000);
  }
}
