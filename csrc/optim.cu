// K10: fused optimizer steps over the flat parameter arena.
// One launch updates every parameter of the model; the reference's torch.optim.SGD walks a Python
// list of 10..900 tensors (src/main.py:99-100,151).
#include "common.cuh"
#include "../include/fedb200.h"

namespace {

// kStep: the "end of step" form used by the zoo engine - the gradient is cleared as it is consumed (the next step's
// backward pass accumulates into zeros without a separate fill) and the BatchNorm step counters (num_batches_tracked, the
// whole int64 segment of the arena) advance by one here instead of by one tiny launch per BatchNorm layer.
template <bool kShadow, bool kStep = false>
__global__ void __launch_bounds__(256) sgd_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                 float* __restrict__ v, uint16_t* __restrict__ shadow,
                                                 long long n4, float lr, float mu, float wd,
                                                 long long* __restrict__ counters = nullptr, long long n_counters = 0) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  if constexpr (kStep) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_counters; i += stride) counters[i] += 1;
  }
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 pp = reinterpret_cast<float4*>(p)[i];
    const float4 gg = reinterpret_cast<const float4*>(g)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    // torch order: g' = g + wd*p ; v = mu*v + g' ; p = p - lr*v   (each op rounded, no fma contraction)
#define FEDB_SGD1(c)                                                     \
  {                                                                      \
    const float g1 = __fadd_rn(gg.c, __fmul_rn(wd, pp.c));               \
    vv.c = __fadd_rn(__fmul_rn(mu, vv.c), g1);                           \
    pp.c = __fadd_rn(pp.c, __fmul_rn(-lr, vv.c));                        \
  }
    FEDB_SGD1(x) FEDB_SGD1(y) FEDB_SGD1(z) FEDB_SGD1(w)
#undef FEDB_SGD1
    reinterpret_cast<float4*>(p)[i] = pp;
    reinterpret_cast<float4*>(v)[i] = vv;
    if constexpr (kStep) reinterpret_cast<float4*>(const_cast<float*>(g))[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if constexpr (kShadow) {
      __nv_bfloat162 lo = __floats2bfloat162_rn(pp.x, pp.y), hi = __floats2bfloat162_rn(pp.z, pp.w);
      uint2 r;
      r.x = *reinterpret_cast<uint32_t*>(&lo);
      r.y = *reinterpret_cast<uint32_t*>(&hi);
      reinterpret_cast<uint2*>(shadow)[i] = r;
    }
  }
}

__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                  float* __restrict__ m, float* __restrict__ v, long long n,
                                                  float lr, float b1, float b2, float eps, float wd,
                                                  const int32_t* __restrict__ step_dev) {
  const float t = (float)(*step_dev);
  const float c1 = 1.f - powf(b1, t), c2 = 1.f - powf(b2, t);
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    float gi = g[i] + wd * p[i];
    float mi = b1 * m[i] + (1.f - b1) * gi;
    float vi = b2 * v[i] + (1.f - b2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    p[i] -= lr * (mi / c1) / (sqrtf(vi / c2) + eps);
  }
}

}  // namespace

extern "C" int fedb_sgd_step(float* param, const float* grad, float* momentum, uint16_t* shadow_bf16, int64_t n,
                             float lr, float mu, float weight_decay, void* stream) {
  FEDB_REQUIRE(n % 4 == 0, "fedb_sgd_step: n must be a multiple of 4 (arena segments are padded)");
  if (n == 0) return 0;
  int blocks = fedb_ceil_div(n / 4, 256);
  if (blocks > fedb_num_sms() * 8) blocks = fedb_num_sms() * 8;
  if (shadow_bf16)
    sgd_kernel<true><<<blocks, 256, 0, (cudaStream_t)stream>>>(param, grad, momentum, shadow_bf16, n / 4, lr, mu,
                                                               weight_decay);
  else
    sgd_kernel<false><<<blocks, 256, 0, (cudaStream_t)stream>>>(param, grad, momentum, nullptr, n / 4, lr, mu,
                                                                weight_decay);
  FEDB_LAUNCH_OK();
  return 0;
}

extern "C" int fedb_sgd_step_end(float* param, float* grad, float* momentum, int64_t n, float lr, float mu,
                                 float weight_decay, int64_t* counters, int64_t n_counters, void* stream) {
  FEDB_REQUIRE(n % 4 == 0, "fedb_sgd_step_end: n must be a multiple of 4 (arena segments are padded)");
  FEDB_REQUIRE(n_counters == 0 || counters != nullptr, "fedb_sgd_step_end: null counter segment");
  if (n == 0) return 0;
  int blocks = fedb_ceil_div(n / 4, 256);
  if (blocks > fedb_num_sms() * 8) blocks = fedb_num_sms() * 8;
  sgd_kernel<false, true><<<blocks, 256, 0, (cudaStream_t)stream>>>(param, grad, momentum, nullptr, n / 4, lr, mu, weight_decay,
                                                                    (long long*)counters, (long long)n_counters);
  FEDB_LAUNCH_OK();
  return 0;
}

extern "C" int fedb_adam_step(float* param, const float* grad, float* m, float* v, int64_t n, float lr, float beta1,
                              float beta2, float eps, float weight_decay, const int32_t* step_dev, void* stream) {
  if (n == 0) return 0;
  int blocks = fedb_ceil_div(n, 256);
  if (blocks > fedb_num_sms() * 8) blocks = fedb_num_sms() * 8;
  adam_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(param, grad, m, v, n, lr, beta1, beta2, eps, weight_decay,
                                                        step_dev);
  FEDB_LAUNCH_OK();
  return 0;
}
