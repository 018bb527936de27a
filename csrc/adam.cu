// Fused Adam over one flat fp32 parameter arena (F6 in SURVEY §8).
//
// Replaces the reference's unfused `optim.Adam` (experiments/OGB/main.py:114, OGB-LSC/Trainer.py:78-80,
// GraphCast/train_graphcast.py:82): every parameter of the (<= ~1 M parameter) models lives in one contiguous
// buffer, gradients in another, so the whole optimizer step is ONE float4-vectorised kernel that also applies the
// data-parallel gradient scale, L2 weight decay, and optionally refreshes the bf16 working copy of the weights.
#include "common.cuh"

namespace dgb {

struct AdamArgs {
  float lr, beta1, beta2, eps, weight_decay, grad_scale, bias_c1, bias_c2;  // bias_c = 1 - beta^t
};

__global__ void __launch_bounds__(256) adam_flat_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                        float* __restrict__ m, float* __restrict__ v,
                                                        __nv_bfloat16* __restrict__ p_bf16, int64_t n, AdamArgs a,
                                                        const float* __restrict__ dev_state) {
  const int64_t i4 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i4 >= n) return;
  if (dev_state) {  // capturable mode: learning rate and bias corrections live in device memory (see adam_tick_kernel)
    a.lr = dev_state[0];
    a.bias_c1 = dev_state[1];
    a.bias_c2 = dev_state[2];
  }
  const float inv_c1 = 1.f / a.bias_c1;
  const float inv_sqrt_c2 = rsqrtf(a.bias_c2);
  if (i4 + 4 <= n) {
    float4 pv = *reinterpret_cast<float4*>(p + i4);
    const float4 gv = *reinterpret_cast<const float4*>(g + i4);
    float4 mv = *reinterpret_cast<float4*>(m + i4);
    float4 vv = *reinterpret_cast<float4*>(v + i4);
    float pp[4] = {pv.x, pv.y, pv.z, pv.w}, gg[4] = {gv.x, gv.y, gv.z, gv.w};
    float mm[4] = {mv.x, mv.y, mv.z, mv.w}, vw[4] = {vv.x, vv.y, vv.z, vv.w};
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const float gr = gg[t] * a.grad_scale + a.weight_decay * pp[t];
      mm[t] = a.beta1 * mm[t] + (1.f - a.beta1) * gr;
      vw[t] = a.beta2 * vw[t] + (1.f - a.beta2) * gr * gr;
      const float denom = sqrtf(vw[t]) * inv_sqrt_c2 + a.eps;
      pp[t] -= a.lr * (mm[t] * inv_c1) / denom;
    }
    *reinterpret_cast<float4*>(p + i4) = make_float4(pp[0], pp[1], pp[2], pp[3]);
    *reinterpret_cast<float4*>(m + i4) = make_float4(mm[0], mm[1], mm[2], mm[3]);
    *reinterpret_cast<float4*>(v + i4) = make_float4(vw[0], vw[1], vw[2], vw[3]);
    if (p_bf16) {
      uint2 o;
      o.x = pack_bf2(pp[0], pp[1]);
      o.y = pack_bf2(pp[2], pp[3]);
      *reinterpret_cast<uint2*>(p_bf16 + i4) = o;
    }
  } else {
    for (int64_t i = i4; i < n; ++i) {
      const float gr = g[i] * a.grad_scale + a.weight_decay * p[i];
      const float mi = a.beta1 * m[i] + (1.f - a.beta1) * gr;
      const float vi = a.beta2 * v[i] + (1.f - a.beta2) * gr * gr;
      m[i] = mi; v[i] = vi;
      const float pn = p[i] - a.lr * (mi * inv_c1) / (sqrtf(vi) * inv_sqrt_c2 + a.eps);
      p[i] = pn;
      if (p_bf16) p_bf16[i] = __float2bfloat16_rn(pn);
    }
  }
}

// Capturable mode: advances the device-resident step counter and refreshes the bias corrections, so a CUDA graph that
// contains the optimizer step replays with the right `t` (kernel arguments are frozen at capture time).
__global__ void adam_tick_kernel(int* __restrict__ step, float* __restrict__ dev_state, float beta1, float beta2) {
  const int t = *step + 1;
  *step = t;
  dev_state[1] = 1.f - powf(beta1, (float)t);
  dev_state[2] = 1.f - powf(beta2, (float)t);
}

}  // namespace dgb

// One Adam step over flat fp32 arrays of n elements (16-byte aligned).  `step` >= 1.  p_bf16 may be null.
DGB_API int dgb_adam_flat(float* p, const float* g, float* m, float* v, void* p_bf16, int64_t n, float lr, float beta1,
                          float beta2, float eps, float weight_decay, float grad_scale, int step, void* stream) {
  if (n == 0) return 0;
  if (step < 1) return (int)cudaErrorInvalidValue;
  dgb::AdamArgs a{lr, beta1, beta2, eps, weight_decay, grad_scale, 1.f - powf(beta1, (float)step),
                  1.f - powf(beta2, (float)step)};
  const int64_t threads = (n + 3) / 4;
  dgb::adam_flat_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      p, g, m, v, (__nv_bfloat16*)p_bf16, n, a, nullptr);
  DGB_CHECK_LAUNCH();
  return 0;
}

// Same step with the step counter (`step_dev`, int32) and `dev_state` = {lr, 1-beta1^t, 1-beta2^t, unused} in device memory:
// two launches (tick + update), no host-side value that changes from step to step -> safe to capture once and replay.
// The caller writes dev_state[0] (learning rate) with a device-side fill outside the captured region.
DGB_API int dgb_adam_flat_capturable(float* p, const float* g, float* m, float* v, void* p_bf16, int64_t n, float beta1,
                                     float beta2, float eps, float weight_decay, float grad_scale, int* step_dev,
                                     float* dev_state, void* stream) {
  if (n == 0) return 0;
  if (!step_dev || !dev_state) return (int)cudaErrorInvalidValue;
  dgb::adam_tick_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(step_dev, dev_state, beta1, beta2);
  DGB_CHECK_LAUNCH();
  dgb::AdamArgs a{0.f, beta1, beta2, eps, weight_decay, grad_scale, 1.f, 1.f};
  const int64_t threads = (n + 3) / 4;
  dgb::adam_flat_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      p, g, m, v, (__nv_bfloat16*)p_bf16, n, a, dev_state);
  DGB_CHECK_LAUNCH();
  return 0;
}
