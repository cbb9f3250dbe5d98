// Fused optimizer updates on flat parameter buckets + device-resident LR schedule.
//
// Replaces tf.contrib.optimizer_v2.AdamOptimizer + tf.train.exponential_decay
// (reference: model.py:457-467) and adds the momentum-SGD the BASELINE configs name.
// One launch updates a whole bucket: reads the (fp32 or bf16) gradient, applies the 1/N tower
// scale, updates fp32 optimizer state and fp32 master weights, writes the bf16 compute copy and
// clears the gradient for the next step -- a single pass over memory instead of one kernel per
// variable per tower.  The learning rate and Adam bias correction live in device memory so a
// captured CUDA graph replays with a moving schedule.
#include "tdl_common.cuh"

namespace tdl {
namespace optim {

constexpr int kThreads = 256;

// lr_state[0] = scheduled lr, lr_state[1] = Adam lr_t (bias corrected), global_step += 1.
__global__ void lr_schedule_kernel(long long* global_step, float base_lr, float decay_rate, float decay_steps,
                                   float beta1, float beta2, float* lr_state) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const long long step = *global_step;  // value before this step's increment (tf.train.exponential_decay)
  // One thread, once per step: do it in double.  (The translation unit is built with --use_fast_math, under which
  // powf / sqrtf / division turn into approximate intrinsics: 1 - __powf(0.999f, t) loses ~1e-4 relative at small t,
  // and a float cast of the step stops counting at 2^24.  Double-precision pow is not affected by that flag.)
  double lr = (double)base_lr;
  if (decay_steps > 0.f) lr *= pow((double)decay_rate, (double)step / (double)decay_steps);
  const double t = (double)(step + 1);
  lr_state[0] = (float)lr;
  lr_state[1] = (float)(lr * sqrt(1.0 - pow((double)beta2, t)) / (1.0 - pow((double)beta1, t)));
  *global_step = step + 1;
}

template <typename G>
__device__ __forceinline__ float load_grad(const G* g, int64_t i);
template <>
__device__ __forceinline__ float load_grad<float>(const float* g, int64_t i) { return g[i]; }
template <>
__device__ __forceinline__ float load_grad<bf16>(const bf16* g, int64_t i) { return __bfloat162float(g[i]); }

template <typename G>
__global__ void __launch_bounds__(kThreads)
sgd_momentum_kernel(float* __restrict__ p, G* __restrict__ g, float* __restrict__ mom, bf16* __restrict__ shadow,
                    int64_t n, const float* __restrict__ lr_state, float momentum, float weight_decay,
                    float grad_scale, int zero_grad) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const float lr = lr_state[0];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float w = p[i];
    float gr = load_grad<G>(g, i) * grad_scale + weight_decay * w;
    float m = momentum * mom[i] + gr;  // TF MomentumOptimizer: accum = momentum*accum + grad
    w -= lr * m;
    mom[i] = m;
    p[i] = w;
    if (shadow) shadow[i] = __float2bfloat16(w);
    if (zero_grad) g[i] = G(0.f);
  }
}

template <typename G>
__global__ void __launch_bounds__(kThreads)
adam_kernel(float* __restrict__ p, G* __restrict__ g, float* __restrict__ m1, float* __restrict__ m2,
            bf16* __restrict__ shadow, int64_t n, const float* __restrict__ lr_state, float beta1, float beta2,
            float eps, float grad_scale, int zero_grad) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const float lr_t = lr_state[1];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float gr = load_grad<G>(g, i) * grad_scale;
    const float m = beta1 * m1[i] + (1.f - beta1) * gr;
    const float v = beta2 * m2[i] + (1.f - beta2) * gr * gr;
    const float w = p[i] - lr_t * m / (sqrtf(v) + eps);  // TF Adam: epsilon outside the corrected root
    m1[i] = m;
    m2[i] = v;
    p[i] = w;
    if (shadow) shadow[i] = __float2bfloat16(w);
    if (zero_grad) g[i] = G(0.f);
  }
}

__global__ void cast_f32_to_bf16_kernel(const float* __restrict__ src, bf16* __restrict__ dst, int64_t n) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = __float2bfloat16(src[i]);
}

static int blocks_for(int64_t n) {
  int64_t b = cdiv(n, kThreads * 4);
  const int64_t cap = (int64_t)num_sms() * 8;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

}  // namespace optim
}  // namespace tdl

using namespace tdl;
using namespace tdl::optim;

TDL_API int tdl_lr_schedule_step(void* global_step_i64, float base_lr, float decay_rate, float decay_steps,
                                 float beta1, float beta2, float* lr_state, void* stream) {
  {
    cudaError_t _le = tdl::launch_pdl(lr_schedule_kernel, dim3(1), dim3(1), 0, (cudaStream_t)stream,
        (long long*)global_step_i64, base_lr, decay_rate, decay_steps, beta1, beta2, lr_state);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "lr_schedule_kernel");
  }
  TDL_LAUNCH_OK("lr_schedule_kernel");
  return 0;
}

TDL_API int tdl_sgd_momentum_flat(float* param, void* grad, float* mom, void* shadow_bf16, int64_t n,
                                  const float* lr_state, float momentum, float weight_decay, float grad_scale,
                                  int grad_is_bf16, int zero_grad, void* stream) {
  if (n <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  if (grad_is_bf16)
    {
      cudaError_t _le = tdl::launch_pdl(sgd_momentum_kernel<bf16>, dim3(blocks_for(n)), dim3(kThreads), 0, st, param,
          (bf16*)grad, mom, (bf16*)shadow_bf16, n, lr_state, momentum, weight_decay, grad_scale, zero_grad);
      if (_le != cudaSuccess) return tdl::cuda_fail(_le, "sgd_momentum_kernel");
    }
  else
    {
      cudaError_t _le = tdl::launch_pdl(sgd_momentum_kernel<float>, dim3(blocks_for(n)), dim3(kThreads), 0, st, param,
          (float*)grad, mom, (bf16*)shadow_bf16, n, lr_state, momentum, weight_decay, grad_scale, zero_grad);
      if (_le != cudaSuccess) return tdl::cuda_fail(_le, "sgd_momentum_kernel");
    }
  TDL_LAUNCH_OK("sgd_momentum_kernel");
  return 0;
}

TDL_API int tdl_adam_flat(float* param, void* grad, float* m1, float* m2, void* shadow_bf16, int64_t n,
                          const float* lr_state, float beta1, float beta2, float eps, float grad_scale,
                          int grad_is_bf16, int zero_grad, void* stream) {
  if (n <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  if (grad_is_bf16)
    {
      cudaError_t _le = tdl::launch_pdl(adam_kernel<bf16>, dim3(blocks_for(n)), dim3(kThreads), 0, st, param,
          (bf16*)grad, m1, m2, (bf16*)shadow_bf16, n, lr_state, beta1, beta2, eps, grad_scale, zero_grad);
      if (_le != cudaSuccess) return tdl::cuda_fail(_le, "adam_kernel");
    }
  else
    {
      cudaError_t _le = tdl::launch_pdl(adam_kernel<float>, dim3(blocks_for(n)), dim3(kThreads), 0, st, param,
          (float*)grad, m1, m2, (bf16*)shadow_bf16, n, lr_state, beta1, beta2, eps, grad_scale, zero_grad);
      if (_le != cudaSuccess) return tdl::cuda_fail(_le, "adam_kernel");
    }
  TDL_LAUNCH_OK("adam_kernel");
  return 0;
}

TDL_API int tdl_cast_f32_to_bf16(const float* src, void* dst, int64_t n, void* stream) {
  if (n <= 0) return 0;
  {
    cudaError_t _le = tdl::launch_pdl(cast_f32_to_bf16_kernel, dim3(blocks_for(n)), dim3(kThreads), 0,
        (cudaStream_t)stream, src, (bf16*)dst, n);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "cast_f32_to_bf16_kernel");
  }
  TDL_LAUNCH_OK("cast_f32_to_bf16_kernel");
  return 0;
}
