// act.cuh — activation functions shared by the GEMM epilogue (gemm_sm100.cu) and the training-side element-wise
// kernels (train_ops.cu), so that a separately launched activation is bit-identical to the fused one.
#pragma once
#include "common.cuh"

namespace mm {

__device__ __forceinline__ float fast_ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float fast_rcp(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float fast_tanh(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// x * sigmoid(k*x) = h + h*tanh(k*h), h = x/2: ONE MUFU op (tanh.approx, rel. error 2^-11 — below the bf16 output
// rounding) instead of ex2 + rcp.  The epilogue's 64 activations per lane and chunk are MUFU-throughput bound
// (4 lanes/clk/scheduler), so this halves the dominant cost of the fc1 epilogue.
__device__ __forceinline__ float x_sigmoid(float x, float k) {
  const float h = 0.5f * x;
  return fmaf(h, fast_tanh(k * h), h);
}

template <int ACT>
__device__ __forceinline__ float apply_act(float x) {
  if (ACT == MM_ACT_QUICK_GELU) {  // x * sigmoid(1.702 x)
    return x_sigmoid(x, 1.702f);
  } else if (ACT == MM_ACT_GELU_TANH) {  // 0.5x(1+tanh(u)), u = sqrt(2/pi)(x + 0.044715x^3)
    const float h = 0.5f * x;
    const float u = x * fmaf(x * x, 0.035677408136300125f, 0.7978845608028654f);
    return fmaf(h, fast_tanh(u), h);
  } else if (ACT == MM_ACT_GELU_ERF) {  // exact GELU; erf by Abramowitz-Stegun 7.1.26 (|err| <= 1.5e-7)
    const float ax = fabsf(x) * 0.70710678118654752440f;
    const float t = fast_rcp(fmaf(0.3275911f, ax, 1.0f));
    float poly = fmaf(t, 1.061405429f, -1.453152027f);
    poly = fmaf(poly, t, 1.421413741f);
    poly = fmaf(poly, t, -0.284496736f);
    poly = fmaf(poly, t, 0.254829592f);
    poly *= t;
    const float erf_abs = 1.0f - poly * fast_ex2(-1.4426950408889634f * ax * ax);
    return 0.5f * x * (1.0f + copysignf(erf_abs, x));
  }
  return x;
}

}  // namespace mm
