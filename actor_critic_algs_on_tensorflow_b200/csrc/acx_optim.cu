// K5 -- Adam apply for both networks, the KL-driven learning-rate rule and the beta/gamma
// schedules, all on device (no host sync), plus the library's error/info plumbing.
//
// Replaces tf.train.AdamOptimizer.apply_gradients (A3C/policies.py:82,88-89,155,160-161; TF
// 1.3 ApplyAdam arithmetic), Actor.get_kl + the lr rule (A3C/policies.py:106-108,
// A3C/process.py:262-270) and the LinearSchedule updates (A3C/util.py:34-42,
// A3C/process.py:271-278).  HBM-bound: 28 B per parameter (read g, theta, m, v; write theta, m, v).
#include <stdarg.h>

#include "acx_optim.cuh"

// ------------------------------------------------------------------ error / info plumbing
static thread_local char g_err[512] = "";

void acx_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
extern "C" const char* acx_last_error(void) { return g_err; }
extern "C" int acx_abi_version(void) { return ACX_ABI_VERSION; }

int acx_num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev >= 0 && dev < 64 && cached[dev] > 0) return cached[dev];
  int sms = 0;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) sms = 148;
  if (dev >= 0 && dev < 64) cached[dev] = sms;
  return sms;
}

extern "C" int acx_params_layout(int net, int ob_dim, int num_ac, int cont, int* offsets, int* sizes) {
  ACX_REQUIRE(net == 0 || net == 1, "acx_params_layout: net must be 0 (actor) or 1 (critic), got %d", net);
  ACX_REQUIRE(ob_dim >= 1 && ob_dim <= 24, "acx_params_layout: ob_dim %d outside [1,24]", ob_dim);
  ACX_REQUIRE(net == 1 || (num_ac >= 1 && num_ac <= 4), "acx_params_layout: num_ac %d outside [1,4]", num_ac);
  const AcxNetLayout L = net == 0 ? acx_actor_layout(ob_dim, num_ac, cont) : acx_critic_layout(ob_dim);
  if (offsets) {
    const int o[9] = {L.W1, L.b1, L.W2, L.b2, L.W3, L.b3, L.Wh, L.bh, L.log_std};
    memcpy(offsets, o, sizeof(o));
  }
  if (sizes) {
    const int s[9] = {L.ob * L.H1, L.H1, L.H1 * L.H2, L.H2, L.H2 * L.H3, L.H3, L.H3 * L.nout, L.nout,
                      L.log_std >= 0 ? L.nout : 0};
    memcpy(sizes, s, sizeof(s));
  }
  return L.size;
}

// ------------------------------------------------------------------ kernels
__global__ void acx_opt_init_kernel(float* opt, float actor_lr, float critic_lr, float beta_kl, float gamma_ent,
                                    float clip, float desired_kl, float min_lr, float max_lr, int schedule_on) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    for (int i = 0; i < ACX_OPT_SIZE; ++i) opt[i] = 0.f;
    opt[ACX_OPT_ACTOR_LR] = actor_lr;
    opt[ACX_OPT_CRITIC_LR] = critic_lr;
    opt[ACX_OPT_BETA_KL] = beta_kl;
    opt[ACX_OPT_GAMMA_ENT] = gamma_ent;
    opt[ACX_OPT_BETA1_POW] = ACX_BETA1;      // TF initialises beta_power slots to beta
    opt[ACX_OPT_BETA2_POW] = ACX_BETA2;
    opt[ACX_OPT_CLIP] = clip;
    opt[ACX_OPT_DESIRED_KL] = desired_kl;
    opt[ACX_OPT_MIN_LR] = min_lr;
    opt[ACX_OPT_MAX_LR] = max_lr;
    opt[ACX_OPT_ITER] = 0.f;
    opt[ACX_OPT_SCHEDULE] = schedule_on ? 1.f : 0.f;
    opt[ACX_OPT_ADAM_T] = 0.f;
  }
}

__global__ void __launch_bounds__(256) acx_adam_kernel(float* __restrict__ actor, const float* __restrict__ ga,
                                                       float* __restrict__ ma, float* __restrict__ va, int na,
                                                       float* __restrict__ critic, const float* __restrict__ gc,
                                                       float* __restrict__ mc, float* __restrict__ vc, int nc,
                                                       const float* __restrict__ opt, float scale) {
  const float corr = acx_adam_corr(opt);
  const float alpha_a = __fmul_rn(opt[ACX_OPT_ACTOR_LR], corr);      // lr * sqrt(1-b2^t) / (1-b1^t)
  const float alpha_c = __fmul_rn(opt[ACX_OPT_CRITIC_LR], corr);
  const float clip = opt[ACX_OPT_CLIP];
  const int n4a = na >> 2, n4c = nc >> 2;   // layouts are padded to multiples of 4 floats
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n4a + n4c; i += gridDim.x * blockDim.x) {
    const bool is_a = i < n4a;
    const int j = is_a ? i : i - n4a;
    float4* th = reinterpret_cast<float4*>(is_a ? actor : critic) + j;
    float4* m = reinterpret_cast<float4*>(is_a ? ma : mc) + j;
    float4* v = reinterpret_cast<float4*>(is_a ? va : vc) + j;
    float4 g = reinterpret_cast<const float4*>(is_a ? ga : gc)[j];
    g.x *= scale; g.y *= scale; g.z *= scale; g.w *= scale;
    if (is_a && clip > 0.f) g = acx_clip4(g, clip);
    float4 t = *th, mm = *m, vv = *v;
    const float alpha = is_a ? alpha_a : alpha_c;
    acx_adam_one(t.x, g.x, mm.x, vv.x, alpha); acx_adam_one(t.y, g.y, mm.y, vv.y, alpha);
    acx_adam_one(t.z, g.z, mm.z, vv.z, alpha); acx_adam_one(t.w, g.w, mm.w, vv.w, alpha);
    *th = t; *m = mm; *v = vv;
  }
}

__global__ void acx_adam_tick_kernel(float* opt) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    opt[ACX_OPT_BETA1_POW] = __fmul_rn(opt[ACX_OPT_BETA1_POW], ACX_BETA1);
    opt[ACX_OPT_BETA2_POW] = __fmul_rn(opt[ACX_OPT_BETA2_POW], ACX_BETA2);
    opt[ACX_OPT_ADAM_T] += 1.f;
  }
}

__global__ void acx_counter_add_kernel(uint32_t* ctr, uint32_t inc) {
  if (threadIdx.x == 0 && blockIdx.x == 0) *ctr += inc;
}

__global__ void acx_kl_finalize_kernel(const double* __restrict__ part, int nb, double count, float* stats) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    double s = 0.0;
    for (int i = 0; i < nb; ++i) s += part[i];
    stats[ACX_STAT_KL_POST] = (float)(s / count);
  }
}

__device__ inline double lin_sched(double t, double init_t, double end_t, double init_v, double end_v) {
  if (t < init_t) return init_v;                 // A3C/util.py:34-39
  if (t > end_t) return end_v;
  return ((t - init_t) * end_v + (end_t - t) * init_v) / (end_t - init_t);
}

// flags & 1: the lr rule; flags & 2: the schedules + the iteration counter.  kl_sum != NULL: the KL the rule sees is
// *kl_sum * kl_scale (a cross-rank sum that rode along with the gradient all-reduce) and is left in stats[KL_POST].
__global__ void acx_kl_lr_kernel(float* opt, float* stats, const float* kl_sum, float kl_scale, int flags) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    if (flags & 1) {
      if (kl_sum != nullptr) stats[ACX_STAT_KL_POST] = __fmul_rn(*kl_sum, kl_scale);
      const float lr = opt[ACX_OPT_ACTOR_LR];
      stats[ACX_STAT_ACTOR_LR_USED] = lr;
      opt[ACX_OPT_ACTOR_LR] = acx_lr_rule(lr, stats[ACX_STAT_KL_POST], opt);
    }
    if (flags & 2) {
      const int it = (int)opt[ACX_OPT_ITER];
      if (opt[ACX_OPT_SCHEDULE] != 0.f && it % 100 == 0) {          // A3C/process.py:173-174,271-278
        opt[ACX_OPT_GAMMA_ENT] = (float)pow(10.0, lin_sched((double)it, 100.0, 3000.0, -2.0, -8.0));
        opt[ACX_OPT_BETA_KL] = (float)pow(10.0, lin_sched((double)it, 100.0, 3000.0, 0.0, -4.0));
      }
      opt[ACX_OPT_ITER] = (float)(it + 1);
    }
  }
}

// ------------------------------------------------------------------ entry points
extern "C" int acx_opt_init(float* opt, float actor_lr, float critic_lr, float beta_kl, float gamma_ent, float clip,
                            float desired_kl, float min_lr, float max_lr, int schedule_on, void* stream) {
  ACX_REQUIRE(opt, "acx_opt_init: opt is null");
  acx_opt_init_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(opt, actor_lr, critic_lr, beta_kl, gamma_ent, clip,
                                                         desired_kl, min_lr, max_lr, schedule_on);
  ACX_LAUNCH_OK("acx_opt_init_kernel");
  return ACX_OK;
}

extern "C" int acx_adam_apply(float* actor, const float* grad_actor, float* m_actor, float* v_actor, int n_actor,
                              float* critic, const float* grad_critic, float* m_critic, float* v_critic, int n_critic,
                              const float* opt, float grad_scale, void* stream) {
  ACX_REQUIRE(opt, "acx_adam_apply: opt is null");
  ACX_REQUIRE(n_actor >= 0 && n_critic >= 0, "acx_adam_apply: negative parameter count");
  ACX_REQUIRE((n_actor % 4) == 0 && (n_critic % 4) == 0,
              "acx_adam_apply: parameter counts must be the padded layout sizes (multiples of 4): %d, %d", n_actor,
              n_critic);
  if (n_actor > 0)
    ACX_REQUIRE(actor && grad_actor && m_actor && v_actor && acx_aligned16(actor) && acx_aligned16(grad_actor) &&
                    acx_aligned16(m_actor) && acx_aligned16(v_actor),
                "acx_adam_apply: actor buffers must be non-null and 16-byte aligned");
  if (n_critic > 0)
    ACX_REQUIRE(critic && grad_critic && m_critic && v_critic && acx_aligned16(critic) &&
                    acx_aligned16(grad_critic) && acx_aligned16(m_critic) && acx_aligned16(v_critic),
                "acx_adam_apply: critic buffers must be non-null and 16-byte aligned");
  const int n4 = (n_actor + n_critic) / 4;
  if (n4 == 0) return ACX_OK;
  int grid = (n4 + 255) / 256;
  const int cap = acx_num_sms() * 8;
  if (grid > cap) grid = cap;
  acx_adam_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(actor, grad_actor, m_actor, v_actor, n_actor, critic,
                                                         grad_critic, m_critic, v_critic, n_critic, opt, grad_scale);
  ACX_LAUNCH_OK("acx_adam_kernel");
  return ACX_OK;
}

extern "C" int acx_adam_tick(float* opt, void* stream) {
  ACX_REQUIRE(opt, "acx_adam_tick: opt is null");
  acx_adam_tick_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(opt);
  ACX_LAUNCH_OK("acx_adam_tick_kernel");
  return ACX_OK;
}

extern "C" int acx_counter_add(uint32_t* counter, uint32_t inc, void* stream) {
  ACX_REQUIRE(counter, "acx_counter_add: counter is null");
  acx_counter_add_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(counter, inc);
  ACX_LAUNCH_OK("acx_counter_add_kernel");
  return ACX_OK;
}

extern "C" int acx_kl_finalize(const double* kl_partial, int kl_blocks, double count, float* stats, void* stream) {
  ACX_REQUIRE(kl_partial && stats, "acx_kl_finalize: null buffer");
  ACX_REQUIRE(kl_blocks >= 0 && count > 0, "acx_kl_finalize: bad sizes (blocks=%d, count=%g)", kl_blocks, count);
  acx_kl_finalize_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(kl_partial, kl_blocks, count, stats);
  ACX_LAUNCH_OK("acx_kl_finalize_kernel");
  return ACX_OK;
}

extern "C" int acx_kl_lr_update(float* opt, float* stats, void* stream) {
  ACX_REQUIRE(opt && stats, "acx_kl_lr_update: null buffer");
  acx_kl_lr_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(opt, stats, nullptr, 1.f, 3);
  ACX_LAUNCH_OK("acx_kl_lr_kernel");
  return ACX_OK;
}

extern "C" int acx_kl_lr_update_split(float* opt, float* stats, const float* kl_sum, float kl_scale, int flags,
                                      void* stream) {
  ACX_REQUIRE(opt && stats, "acx_kl_lr_update_split: null buffer");
  ACX_REQUIRE(flags >= 1 && flags <= 3, "acx_kl_lr_update_split: flags must be 1 (lr rule), 2 (schedules) or 3 (got %d)",
              flags);
  acx_kl_lr_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(opt, stats, kl_sum, kl_scale, flags);
  ACX_LAUNCH_OK("acx_kl_lr_kernel");
  return ACX_OK;
}
