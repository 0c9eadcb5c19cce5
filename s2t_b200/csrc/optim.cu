// Fused optimizer step over the FLAT parameter buffer: gradient scaling, global-norm clipping
// and Adam with fp32 master weights in two passes over the gradients.
//
// Replaces, for the flat bf16 parameter / gradient buffers of params.ParamStore, the sequence the reference trainer runs
// per update (fairseq/trainer.py:729-759): optimizer.multiply_grads(c) -> clip_grad_norm(max_norm)
// (fairseq/utils.py:328-369: total_norm = ||g||_2 in fp32, clip_coef = min(1, max_norm / (total_norm + 1e-6))) ->
// Adam.step (fairseq/optim/adam.py:145-231: exp_avg / exp_avg_sq updates, denom = sqrt(v) + eps, step_size =
// lr * sqrt(1 - beta2^t) / (1 - beta1^t), weight decay p -= wd * lr * p BEFORE the Adam update, fp32 math on a float
// copy of low-precision parameters) -> copy back to the bf16 parameters (FP16Optimizer / MemoryEfficientFP16Optimizer,
// fairseq/optim/fp16_optimizer.py).  HBM-bound: pass 1 reads the gradients (2 B / element), pass 2 reads gradient +
// master + two moments and writes master + moments + bf16 parameter (2 + 12 + 12 + 2 B / element).
#include "common.cuh"

#include "../../include/s2t_b200.h"

namespace s2t {

constexpr int kNormThreads = 256;

// partial sums of squares: one fp32 per CTA (fixed grid -> deterministic summation order)
__global__ void __launch_bounds__(kNormThreads)
grad_sqsum_partial_kernel(const __nv_bfloat16* __restrict__ g, long long n, float* __restrict__ partials) {
  pdl_entry();
  __shared__ float s_part[kNormThreads / 32];
  float acc = 0.f;
  const long long nv = n >> 3;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < nv; i += stride) {
    const uint4 v = reinterpret_cast<const uint4*>(g)[i];
    const float f[8] = {bf16_lo(v.x), bf16_hi(v.x), bf16_lo(v.y), bf16_hi(v.y),
                        bf16_lo(v.z), bf16_hi(v.z), bf16_lo(v.w), bf16_hi(v.w)};
#pragma unroll
    for (int j = 0; j < 8; ++j) acc += f[j] * f[j];
  }
  if (blockIdx.x == 0)
    for (long long i = (nv << 3) + threadIdx.x; i < n; i += blockDim.x) {
      const float f = __bfloat162float(g[i]);
      acc += f * f;
    }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < kNormThreads / 32; ++w) t += s_part[w];
    partials[blockIdx.x] = t;
  }
}

// total_norm = mult * sqrt(sum of the partials), summed in double in index order
__global__ void grad_norm_finish_kernel(const float* __restrict__ partials, int n_partials, float mult,
                                        float* __restrict__ total_norm) {
  pdl_entry();
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    double t = 0.0;
    for (int i = 0; i < n_partials; ++i) t += static_cast<double>(partials[i]);
    *total_norm = mult * static_cast<float>(sqrt(t));
  }
}

struct AdamParams {
  float lr, beta1, beta2, eps, weight_decay, grad_mult, max_norm;
  float one_minus_beta1, one_minus_beta2;     // float(1 - beta) with the subtraction in DOUBLE, as the reference's
                                              // Python scalars `1 - beta1`, `1 - beta2` (adam.py:207-208): 1.f - 0.999f is
                                              // 1.3e-5 away from 0.001f, which showed up in exp_avg_sq on hardware
  float bias_correction1, bias_correction2;   // used when step_dev == nullptr
  const float* lr_dev;                         // optional device scalars (a captured graph replays with fresh values)
  const long long* step_dev;
  const float* total_norm;                     // device scalar from s2t_grad_norm (nullptr: no clipping)
};

__global__ void __launch_bounds__(256)
fused_adam_kernel(__nv_bfloat16* __restrict__ param, float* __restrict__ master, const __nv_bfloat16* __restrict__ grad,
                  float* __restrict__ exp_avg, float* __restrict__ exp_avg_sq, long long n, const AdamParams a) {
  pdl_entry();
  float lr = a.lr_dev != nullptr ? *a.lr_dev : a.lr;
  float bc1 = a.bias_correction1, bc2 = a.bias_correction2;
  if (a.step_dev != nullptr) {
    const long long step = *a.step_dev;
    if (step < 1) return;          // steps count from 1: step 0 would make the bias correction 0 and the step size infinite
    const float t = static_cast<float>(step);
    bc1 = 1.f - powf(a.beta1, t);
    bc2 = 1.f - powf(a.beta2, t);
  }
  float coef = a.grad_mult;
  if (a.total_norm != nullptr) {
    // Overflowed gradients (NaN / Inf norm): the whole update is skipped — master weights, moments and bf16 parameters
    // stay as they were — as the reference trainer skips it by raising FloatingPointError before optimizer.step
    // (fairseq/trainer.py:751-753).  The host sees the same condition in total_norm[0].
    const float tn = *a.total_norm;
    if (!isfinite(tn)) return;
    if (a.max_norm > 0.f) coef *= fminf(1.f, a.max_norm / (tn + 1e-6f));
  }
  const float step_size = lr * sqrtf(bc2) / bc1;
  const float decay = a.weight_decay * lr;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float g = __bfloat162float(grad[i]) * coef;
    const float m = a.beta1 * exp_avg[i] + a.one_minus_beta1 * g;
    const float v = a.beta2 * exp_avg_sq[i] + a.one_minus_beta2 * g * g;
    float p = master[i];
    if (a.weight_decay != 0.f) p -= decay * p;
    p -= step_size * (m / (sqrtf(v) + a.eps));
    exp_avg[i] = m;
    exp_avg_sq[i] = v;
    master[i] = p;
    param[i] = __float2bfloat16(p);
  }
}

}  // namespace s2t

using namespace s2t;

extern "C" int s2t_grad_norm_partials(long long n) {
  long long g = (n / 8 + kNormThreads - 1) / kNormThreads;
  if (g > 592) g = 592;       // 4 CTAs per SM on 148 SMs
  if (g < 1) g = 1;
  return static_cast<int>(g);
}

extern "C" int s2t_grad_norm(const void* grad, long long n, float mult, float* partials, float* total_norm, void* stream) {
  S2T_REQUIRE(grad && partials && total_norm, "s2t_grad_norm: null pointer");
  S2T_REQUIRE((reinterpret_cast<uintptr_t>(grad) & 15) == 0, "s2t_grad_norm: gradient buffer must be 16-byte aligned");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int np = s2t_grad_norm_partials(n);
  S2T_LAUNCH(grad_sqsum_partial_kernel, np, kNormThreads, 0, st, reinterpret_cast<const __nv_bfloat16*>(grad), n, partials);
  S2T_LAUNCH_CHECK();
  S2T_LAUNCH(grad_norm_finish_kernel, 1, 32, 0, st, static_cast<const float*>(partials), np, mult, total_norm);
  S2T_LAUNCH_CHECK();
  return 0;
}

extern "C" int s2t_fused_adam(void* param, float* master, const void* grad, float* exp_avg, float* exp_avg_sq,
                              long long n, float lr, double beta1, double beta2, float eps, float weight_decay,
                              long long step, float grad_mult, const float* total_norm, float max_norm,
                              const float* lr_dev, const long long* step_dev, void* stream) {
  S2T_REQUIRE(param && master && grad && exp_avg && exp_avg_sq, "s2t_fused_adam: null pointer");
  S2T_REQUIRE(step >= 1 || step_dev != nullptr, "s2t_fused_adam: step counts from 1");
  S2T_REQUIRE(beta1 >= 0.0 && beta1 < 1.0 && beta2 >= 0.0 && beta2 < 1.0, "s2t_fused_adam: betas must be in [0, 1)");
  if (n <= 0) return 0;
  AdamParams a;
  a.lr = lr; a.beta1 = static_cast<float>(beta1); a.beta2 = static_cast<float>(beta2); a.eps = eps;
  a.one_minus_beta1 = static_cast<float>(1.0 - beta1); a.one_minus_beta2 = static_cast<float>(1.0 - beta2); a.weight_decay = weight_decay;
  a.grad_mult = grad_mult; a.max_norm = max_norm;
  a.bias_correction1 = static_cast<float>(1.0 - pow(beta1, static_cast<double>(step)));
  a.bias_correction2 = static_cast<float>(1.0 - pow(beta2, static_cast<double>(step)));
  a.lr_dev = lr_dev; a.step_dev = step_dev; a.total_norm = total_norm;
  long long g = (n + 255) / 256;
  if (g > 148 * 8) g = 148 * 8;
  S2T_LAUNCH(fused_adam_kernel, static_cast<int>(g), 256, 0, reinterpret_cast<cudaStream_t>(stream),
             reinterpret_cast<__nv_bfloat16*>(param), master, reinterpret_cast<const __nv_bfloat16*>(grad), exp_avg,
             exp_avg_sq, n, a);
  S2T_LAUNCH_CHECK();
  return 0;
}
