// Per-env device functions: dynamics / reset / observe for one environment.
// Shared by the stand-alone env-step kernel (acx_env.cu) and the fused rollout step
// (acx_mlp.cu).  Formulas: gym classic_control as restated in SURVEY.md Appendix C and
// oracle/envs.py (third-party, parity unpinned).
#pragma once
#include "acx_common.cuh"

#define ACX_PI_F 3.14159265358979323846f
#define ACX_TWO_PI_F 6.28318530717958647692f

template <int ENV> struct AcxEnv;

// One transition + the observation that follows it, the sequence both env-stepping kernels run:
//   acx_env_transition<ENV>(state, action, reward, terminated, obs)   then, if the episode ended,   reset + observe.
// Environments with FUSED_OBS produce the observation inside the transition (Acrobot: the sines / cosines of the new
// angles also decide termination); for the others `obs` is written by acx_env_observe_after() once the reset is known.
template <int ENV>
__device__ inline void acx_env_transition(float (&s)[AcxEnv<ENV>::S], float a, float& rew, bool& term,
                                          float (&o)[AcxEnv<ENV>::OB]);
template <int ENV>
__device__ inline void acx_env_observe_after(const float (&s)[AcxEnv<ENV>::S], bool was_reset, float (&o)[AcxEnv<ENV>::OB]);

// numpy float32 `%` (floored modulo): fmodf is exact, then fix the sign
__device__ inline float acx_floored_mod(float a, float b) {
  float r = fmodf(a, b);
  if (r != 0.0f) {
    if (r < 0.0f) r = __fadd_rn(r, b);
  } else {
    r = 0.0f;
  }
  return r;
}

template <> struct AcxEnv<ACX_ENV_PENDULUM> {
  static constexpr int S = 2, OB = 3, MAX_STEPS = 200, NACT = 0;
  static constexpr bool CONT = true, FUSED_OBS = false;
  __device__ static inline void dynamics(float (&s)[S], float a, float& rew, bool& term) {
    const float th = s[0], thdot = s[1];
    const float u = fminf(fmaxf(a, -2.0f), 2.0f);
    const float an = __fsub_rn(acx_floored_mod(__fadd_rn(th, ACX_PI_F), ACX_TWO_PI_F), ACX_PI_F);
    float cost = __fadd_rn(__fmul_rn(an, an), __fmul_rn(0.1f, __fmul_rn(thdot, thdot)));
    cost = __fadd_rn(cost, __fmul_rn(0.001f, __fmul_rn(u, u)));
    const float acc = __fadd_rn(__fmul_rn(-15.0f, sinf(__fadd_rn(th, ACX_PI_F))), __fmul_rn(3.0f, u));
    float nthdot = __fadd_rn(thdot, __fmul_rn(acc, 0.05f));
    const float nth = __fadd_rn(th, __fmul_rn(nthdot, 0.05f));
    nthdot = fminf(fmaxf(nthdot, -8.0f), 8.0f);
    s[0] = nth; s[1] = nthdot;
    rew = -cost;
    term = false;
  }
  __device__ static inline void reset(const float (&u)[S], float (&s)[S]) {
    s[0] = __fadd_rn(-ACX_PI_F, __fmul_rn(ACX_TWO_PI_F, u[0]));
    s[1] = __fadd_rn(-1.0f, __fmul_rn(2.0f, u[1]));
  }
  __device__ static inline void observe(const float (&s)[S], float (&o)[OB]) {
    o[0] = cosf(s[0]); o[1] = sinf(s[0]); o[2] = s[1];
  }
};

template <> struct AcxEnv<ACX_ENV_CARTPOLE> {
  static constexpr int S = 4, OB = 4, MAX_STEPS = 200, NACT = 2;
  static constexpr bool CONT = false, FUSED_OBS = false;
  __device__ static inline void dynamics(float (&s)[S], float a, float& rew, bool& term) {
    const float gravity = 9.8f, masspole = 0.1f, total_mass = 1.1f, length = 0.5f, pml = 0.05f, tau = 0.02f;
    float x = s[0], x_dot = s[1], theta = s[2], theta_dot = s[3];
    const float force = (a == 1.0f) ? 10.0f : -10.0f;
    const float ct = cosf(theta), sn = sinf(theta);
    const float temp =
        __fdiv_rn(__fadd_rn(force, __fmul_rn(__fmul_rn(__fmul_rn(pml, theta_dot), theta_dot), sn)), total_mass);
    const float den = __fmul_rn(
        length, __fsub_rn(__fdiv_rn(4.0f, 3.0f), __fdiv_rn(__fmul_rn(__fmul_rn(masspole, ct), ct), total_mass)));
    const float thetaacc = __fdiv_rn(__fsub_rn(__fmul_rn(gravity, sn), __fmul_rn(ct, temp)), den);
    const float xacc = __fsub_rn(temp, __fdiv_rn(__fmul_rn(__fmul_rn(pml, thetaacc), ct), total_mass));
    x = __fadd_rn(x, __fmul_rn(tau, x_dot));
    x_dot = __fadd_rn(x_dot, __fmul_rn(tau, xacc));
    theta = __fadd_rn(theta, __fmul_rn(tau, theta_dot));
    theta_dot = __fadd_rn(theta_dot, __fmul_rn(tau, thetaacc));
    const float thr = 0.20943951023931953f;   // 12 * 2 * pi / 360
    term = (x < -2.4f) || (x > 2.4f) || (theta < -thr) || (theta > thr);
    s[0] = x; s[1] = x_dot; s[2] = theta; s[3] = theta_dot;
    rew = 1.0f;
  }
  __device__ static inline void reset(const float (&u)[S], float (&s)[S]) {
#pragma unroll
    for (int k = 0; k < S; ++k) s[k] = __fadd_rn(-0.05f, __fmul_rn(0.1f, u[k]));
  }
  __device__ static inline void observe(const float (&s)[S], float (&o)[OB]) {
#pragma unroll
    for (int k = 0; k < S; ++k) o[k] = s[k];
  }
};

// sin and cos of a BOUNDED angle (|x| < ~100: Acrobot's angles are wrapped to [-pi, pi] and its velocities clipped, so an
// RK4 stage never leaves |x| < 12): two-constant Cody-Waite reduction to [-pi/4, pi/4] and the single-precision minimax
// polynomials of the Cephes library, 9.2e-8 absolute against float64 over |x| < 40.  sincosf() carries a slow path for
// huge arguments and costs about twice the instructions; the RK4 step calls this ten times and is bound by its arithmetic.
__device__ inline void acx_sincos_bounded(float x, float& sn, float& cs) {
  const float k = rintf(x * 0.636619772367581343f);
  float r = fmaf(k, -1.57079625129699707031f, x);
  r = fmaf(k, -7.54978941586159635335e-8f, r);
  const float z = r * r;
  const float sp = fmaf(r * z, fmaf(z, fmaf(z, -1.9515295891e-4f, 8.3321608736e-3f), -1.6666654611e-1f), r);
  const float cp = fmaf(z * z, fmaf(z, fmaf(z, 2.443315711809948e-5f, -1.388731625493765e-3f), 4.166664568298827e-2f),
                        fmaf(z, -0.5f, 1.0f));
  const int q = (int)k;
  const float s0 = (q & 1) ? cp : sp, c0 = (q & 1) ? sp : cp;
  sn = (q & 2) ? -s0 : s0;
  cs = ((q + 1) & 2) ? -c0 : c0;
}

template <> struct AcxEnv<ACX_ENV_ACROBOT> {
  static constexpr int S = 4, OB = 6, MAX_STEPS = 500, NACT = 3;
  static constexpr bool CONT = false, FUSED_OBS = true;      // step_observe(): the transition also yields the observation
  // "book" dynamics, all link constants 1 except lc = 0.5, g = 9.8.  gym's cos(th1 - pi/2) and cos(th1 + th2 - pi/2)
  // are sin(th1) and sin(th1 + th2) = s1 c2 + c1 s2: two sincosf per evaluation instead of four trigonometric calls
  // (RK4 makes this kernel ALU-bound), and the quotients by d1 share one reciprocal.  Differences from the literal
  // expression are of rounding size (tests: 1e-5 against the fp32 restatement, 1e-4 against float64).
  __device__ static inline void dsdt(const float (&s)[4], float a, float (&d)[4]) {
    const float g = 9.8f;
    const float dth1 = s[2], dth2 = s[3];
    float s1, c1, s2, c2;
    acx_sincos_bounded(s[0], s1, c1);
    acx_sincos_bounded(s[1], s2, c2);
    const float d1 = 0.25f + (1.0f + 0.25f + c2) + 1.0f + 1.0f;
    const float d2 = (0.25f + 0.5f * c2) + 1.0f;
    const float r1 = __frcp_rn(d1);
    const float phi2 = 0.5f * g * (s1 * c2 + c1 * s2);
    const float phi1 = -0.5f * dth2 * dth2 * s2 - 2.0f * 0.5f * dth2 * dth1 * s2 + 1.5f * g * s1 + phi2;
    const float q = d2 * r1;
    const float ddth2 = (a + q * phi1 - 0.5f * dth1 * dth1 * s2 - phi2) * __frcp_rn(0.25f + 1.0f - d2 * q);
    const float ddth1 = -(d2 * ddth2 + phi1) * r1;
    d[0] = dth1; d[1] = dth2; d[2] = ddth1; d[3] = ddth2;
  }
  __device__ static inline float wrap(float x, float lo, float hi) {
    const float diff = hi - lo;
    // gym's `while x > M: x -= diff`; one pass at most for in-range states -- kept as real loops (not unrolled)
#pragma unroll 1
    for (int i = 0; i < 64 && x > hi; ++i) x -= diff;
#pragma unroll 1
    for (int i = 0; i < 64 && x < lo; ++i) x += diff;
    return x;
  }
  // one RK4 transition; also returns the observation of the new state (its sines / cosines decide `term` as well:
  // cos(y0 + y1) = c0 c1 - s0 s1)
  __device__ static inline void step_observe(float (&s)[S], float a, float& rew, bool& term, float (&o)[OB]) {
    const float dt = 0.2f, torque = a - 1.0f;
    float k1[4], k2[4], k3[4], k4[4], y[4];
    dsdt(s, torque, k1);
#pragma unroll
    for (int i = 0; i < 4; ++i) y[i] = s[i] + 0.5f * dt * k1[i];
    dsdt(y, torque, k2);
#pragma unroll
    for (int i = 0; i < 4; ++i) y[i] = s[i] + 0.5f * dt * k2[i];
    dsdt(y, torque, k3);
#pragma unroll
    for (int i = 0; i < 4; ++i) y[i] = s[i] + dt * k3[i];
    dsdt(y, torque, k4);
#pragma unroll
    for (int i = 0; i < 4; ++i) y[i] = s[i] + dt / 6.0f * (k1[i] + 2.0f * k2[i] + 2.0f * k3[i] + k4[i]);
    y[0] = wrap(y[0], -ACX_PI_F, ACX_PI_F);
    y[1] = wrap(y[1], -ACX_PI_F, ACX_PI_F);
    y[2] = fminf(fmaxf(y[2], -4.0f * ACX_PI_F), 4.0f * ACX_PI_F);
    y[3] = fminf(fmaxf(y[3], -9.0f * ACX_PI_F), 9.0f * ACX_PI_F);
    acx_sincos_bounded(y[0], o[1], o[0]);
    acx_sincos_bounded(y[1], o[3], o[2]);
    o[4] = y[2]; o[5] = y[3];
    term = (-o[0] - (o[0] * o[2] - o[1] * o[3])) > 1.0f;
    rew = term ? 0.0f : -1.0f;
#pragma unroll
    for (int i = 0; i < 4; ++i) s[i] = y[i];
  }
  __device__ static inline void dynamics(float (&s)[S], float a, float& rew, bool& term) {
    float o[OB];
    step_observe(s, a, rew, term, o);
  }
  __device__ static inline void reset(const float (&u)[S], float (&s)[S]) {
#pragma unroll
    for (int k = 0; k < S; ++k) s[k] = __fadd_rn(-0.1f, __fmul_rn(0.2f, u[k]));
  }
  __device__ static inline void observe(const float (&s)[S], float (&o)[OB]) {
    acx_sincos_bounded(s[0], o[1], o[0]);
    acx_sincos_bounded(s[1], o[3], o[2]);
    o[4] = s[2]; o[5] = s[3];
  }
};

template <int ENV>
__device__ inline void acx_env_transition(float (&s)[AcxEnv<ENV>::S], float a, float& rew, bool& term,
                                          float (&o)[AcxEnv<ENV>::OB]) {
  if constexpr (AcxEnv<ENV>::FUSED_OBS) AcxEnv<ENV>::step_observe(s, a, rew, term, o);
  else AcxEnv<ENV>::dynamics(s, a, rew, term);
}
template <int ENV>
__device__ inline void acx_env_observe_after(const float (&s)[AcxEnv<ENV>::S], bool was_reset, float (&o)[AcxEnv<ENV>::OB]) {
  if constexpr (AcxEnv<ENV>::FUSED_OBS) {
    if (was_reset) AcxEnv<ENV>::observe(s, o);
  } else {
    AcxEnv<ENV>::observe(s, o);
  }
}
