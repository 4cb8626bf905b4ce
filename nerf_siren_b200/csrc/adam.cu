// Step (5), device side: the optimizer update of the ray-sharded data-parallel step as ONE launch over the flat
// parameter / gradient buffers (replaces utils/__init__.py:20 torch.optim.Adam driven by Lightning after the DDP
// all-reduce, train.py:41-66).  Pure streaming kernel: 4 fp32 reads + 3 fp32 writes per parameter, float4
// vectorised, grid sized in multiples of the SM count.  `grad_scale` folds the 1/world_size of the mean
// all-reduce into the update so the collective can be a plain SUM.
#include "common.cuh"

__global__ void __launch_bounds__(256)
adam_step_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                 int64_t n, float grad_scale, float beta1, float beta2, float omb1, float omb2, float eps, float step_size,
                 float inv_sqrt_bc2) {
  const int64_t n4 = n >> 2;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 pp = reinterpret_cast<float4*>(p)[i];
    const float4 gg = reinterpret_cast<const float4*>(g)[i];
    float4 mm = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    float* pe = &pp.x; const float* ge = &gg.x; float* me = &mm.x; float* ve = &vv.x;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float gk = ge[k] * grad_scale;
      me[k] = me[k] + omb1 * (gk - me[k]);                            // torch.optim.Adam: exp_avg.lerp_(grad, 1-beta1), weight < 0.5 form
      ve[k] = beta2 * ve[k] + omb2 * (gk * gk);                       // exp_avg_sq.mul_(beta2).addcmul_(g, g, 1-beta2)
      const float denom = sqrtf(ve[k]) * inv_sqrt_bc2 + eps;          // (sqrt(v) / sqrt(1-beta2^t)) + eps
      pe[k] = pe[k] - step_size * (me[k] / denom);                    // step_size = lr / (1-beta1^t)
    }
    reinterpret_cast<float4*>(p)[i] = pp;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
  }
  // tail (n not a multiple of 4)
  for (int64_t i = (n4 << 2) + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float gk = g[i] * grad_scale;
    const float mk = m[i] + omb1 * (gk - m[i]);
    const float vk = beta2 * v[i] + omb2 * (gk * gk);
    m[i] = mk; v[i] = vk;
    p[i] = p[i] - step_size * (mk / (sqrtf(vk) * inv_sqrt_bc2 + eps));
  }
}

extern "C" int nerfb_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n,
                               double lr, double beta1, double beta2, double eps, int64_t step, float grad_scale,
                               void* stream) {
  NERFB_CHECK_ARG(params && grads && exp_avg && exp_avg_sq && n >= 0 && step >= 1);
  NERFB_CHECK_ARG((reinterpret_cast<uintptr_t>(params) & 15) == 0 && (reinterpret_cast<uintptr_t>(grads) & 15) == 0 &&
                  (reinterpret_cast<uintptr_t>(exp_avg) & 15) == 0 && (reinterpret_cast<uintptr_t>(exp_avg_sq) & 15) == 0);
  if (n == 0) return NERFB_OK;
  const double bc1 = 1.0 - pow(beta1, (double)step);
  const double bc2 = 1.0 - pow(beta2, (double)step);
  const float step_size = (float)(lr / bc1);
  const float inv_sqrt_bc2 = (float)(1.0 / sqrt(bc2));
  const int64_t want = ceil_div64((n >> 2) + 1, 256);
  const int grid = (int)(want < NERFB_SM_COUNT * 8 ? want : NERFB_SM_COUNT * 8);
  NERFB_LAUNCH(adam_step_kernel, grid, 256, 0, stream, params, grads, exp_avg, exp_avg_sq, n, grad_scale, (float)beta1,
               (float)beta2, (float)(1.0 - beta1), (float)(1.0 - beta2), (float)eps, step_size, inv_sqrt_bc2);
  return NERFB_OK;
}
