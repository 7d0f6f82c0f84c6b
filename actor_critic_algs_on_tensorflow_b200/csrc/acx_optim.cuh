// Arithmetic shared by the optimiser kernels (acx_optim.cu) and the fused cross-GPU exchange + Adam step (acx_xgpu.cu):
// one definition, so the two paths cannot drift apart.
#pragma once
#include "acx_common.cuh"

#define ACX_BETA1 0.9f
#define ACX_BETA2 0.999f
#define ACX_ADAM_EPS 1e-8f

#ifdef __CUDACC__
__device__ inline void acx_adam_one(float& th, float g, float& m, float& v, float alpha) {
  // TF ApplyAdam: m += (g - m)(1 - b1); v += (g*g - v)(1 - b2); var -= (m * alpha) / (sqrt(v) + eps)
  m = __fadd_rn(m, __fmul_rn(__fsub_rn(g, m), __fsub_rn(1.0f, ACX_BETA1)));
  v = __fadd_rn(v, __fmul_rn(__fsub_rn(__fmul_rn(g, g), v), __fsub_rn(1.0f, ACX_BETA2)));
  th = __fsub_rn(th, __fdiv_rn(__fmul_rn(m, alpha), __fadd_rn(__fsqrt_rn(v), ACX_ADAM_EPS)));
}
// lr * sqrt(1 - b2^t) / (1 - b1^t) from the beta-power slots of the opt block
__device__ inline float acx_adam_corr(const float* opt) {
  return __fdiv_rn(__fsqrt_rn(__fsub_rn(1.0f, opt[ACX_OPT_BETA2_POW])), __fsub_rn(1.0f, opt[ACX_OPT_BETA1_POW]));
}
// the KL-driven learning-rate rule (A3C/process.py:262-270): the actor lr the NEXT Adam step uses
__device__ inline float acx_lr_rule(float lr_now, float kl_f, const float* opt) {
  const double kl = (double)kl_f, want = (double)opt[ACX_OPT_DESIRED_KL], lr = (double)lr_now;
  if (kl < want / 4.0) return (float)fmin((double)opt[ACX_OPT_MAX_LR], lr * 1.5);
  if (kl > want * 4.0) return (float)fmax((double)opt[ACX_OPT_MIN_LR], lr / 1.5);
  return lr_now;
}
__device__ inline float4 acx_clip4(float4 g, float clip) {
  g.x = fminf(fmaxf(g.x, -clip), clip); g.y = fminf(fmaxf(g.y, -clip), clip);
  g.z = fminf(fmaxf(g.z, -clip), clip); g.w = fminf(fmaxf(g.w, -clip), clip);
  return g;
}
#endif
