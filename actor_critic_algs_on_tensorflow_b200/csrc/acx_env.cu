// K1 -- batched classic-control env step (Pendulum-v0, CartPole-v0, Acrobot-v1).
//
// Replaces the reference's per-step `env.step(ac)` / `env.reset()` calls
// (A3C/process.py:79,91; gym classic_control) for N envs at once.
// Layout: structure-of-arrays state[k][n] in HBM, one env per lane slot, four envs per
// thread so every array is moved with 16-byte loads/stores.  HBM-bound: 45 B (Pendulum),
// 65 B (CartPole), 73 B (Acrobot) of algorithmic traffic per env-step.
//
// Pendulum and CartPole arithmetic is written with non-contracting intrinsics in gym's own
// expression order so that the fp32 oracle (oracle/envs.py) differs only by sin/cos rounding.
#include "acx_common.cuh"
#include "acx_env_dev.cuh"

template <int ENV, int VEC>
__global__ void __launch_bounds__(256) acx_env_step_kernel(int n, float* __restrict__ state, int* __restrict__ elapsed,
                                                           const void* __restrict__ action, float* __restrict__ obs,
                                                           float* __restrict__ rew, unsigned char* __restrict__ done,
                                                           const float* __restrict__ reset_u, uint64_t seed,
                                                           uint64_t index0, uint32_t step, const uint32_t* __restrict__ step_base,
                                                           int max_steps) {
  if (step_base) step += *step_base;
  constexpr int S = AcxEnv<ENV>::S, OB = AcxEnv<ENV>::OB;
  const int stride = gridDim.x * blockDim.x * VEC;
  for (int base = (blockIdx.x * blockDim.x + threadIdx.x) * VEC; base < n; base += stride) {
    float s[S][VEC], o[OB][VEC], r[VEC], act[VEC];
    int el[VEC];
    unsigned char d[VEC];
    if (VEC == 4) {
#pragma unroll
      for (int k = 0; k < S; ++k) {
        float4 v = *reinterpret_cast<const float4*>(state + (size_t)k * n + base);
        s[k][0] = v.x; s[k][1] = v.y; s[k][2] = v.z; s[k][3] = v.w;
      }
      int4 e = *reinterpret_cast<const int4*>(elapsed + base);
      el[0] = e.x; el[1] = e.y; el[2] = e.z; el[3] = e.w;
      if (AcxEnv<ENV>::CONT) {
        float4 a = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(action) + base);
        act[0] = a.x; act[1] = a.y; act[2] = a.z; act[3] = a.w;
      } else {
        int4 a = *reinterpret_cast<const int4*>(reinterpret_cast<const int*>(action) + base);
        act[0] = (float)a.x; act[1] = (float)a.y; act[2] = (float)a.z; act[3] = (float)a.w;
      }
    } else {
#pragma unroll
      for (int k = 0; k < S; ++k) s[k][0] = state[(size_t)k * n + base];
      el[0] = elapsed[base];
      act[0] = AcxEnv<ENV>::CONT ? reinterpret_cast<const float*>(action)[base]
                                 : (float)reinterpret_cast<const int*>(action)[base];
    }
#pragma unroll
    for (int v = 0; v < VEC; ++v) {
      float sv[S], ov[OB], ru[S];
#pragma unroll
      for (int k = 0; k < S; ++k) sv[k] = s[k][v];
      bool term;
      acx_env_transition<ENV>(sv, act[v], r[v], term, ov);
      int e = el[v] + 1;
      bool dn = term || (e >= max_steps);
      if (dn) {
        if (reset_u != nullptr) {
#pragma unroll
          for (int k = 0; k < S; ++k) ru[k] = reset_u[(size_t)k * n + base + v];
        } else {
          AcxRng4 g = acx_philox(index0 + (uint64_t)(base + v), step, ACX_STREAM_RESET, seed);
          float u4[4] = {acx_u01(g.x), acx_u01(g.y), acx_u01(g.z), acx_u01(g.w)};
#pragma unroll
          for (int k = 0; k < S; ++k) ru[k] = u4[k];
        }
        AcxEnv<ENV>::reset(ru, sv);
        e = 0;
      }
      acx_env_observe_after<ENV>(sv, dn, ov);
#pragma unroll
      for (int k = 0; k < S; ++k) s[k][v] = sv[k];
#pragma unroll
      for (int k = 0; k < OB; ++k) o[k][v] = ov[k];
      el[v] = e;
      d[v] = dn ? 1 : 0;
    }
    if (VEC == 4) {
#pragma unroll
      for (int k = 0; k < S; ++k)
        *reinterpret_cast<float4*>(state + (size_t)k * n + base) = make_float4(s[k][0], s[k][1], s[k][2], s[k][3]);
#pragma unroll
      for (int k = 0; k < OB; ++k)
        *reinterpret_cast<float4*>(obs + (size_t)k * n + base) = make_float4(o[k][0], o[k][1], o[k][2], o[k][3]);
      *reinterpret_cast<int4*>(elapsed + base) = make_int4(el[0], el[1], el[2], el[3]);
      *reinterpret_cast<float4*>(rew + base) = make_float4(r[0], r[1], r[2], r[3]);
      *reinterpret_cast<uchar4*>(done + base) = make_uchar4(d[0], d[1], d[2], d[3]);
    } else {
#pragma unroll
      for (int k = 0; k < S; ++k) state[(size_t)k * n + base] = s[k][0];
#pragma unroll
      for (int k = 0; k < OB; ++k) obs[(size_t)k * n + base] = o[k][0];
      elapsed[base] = el[0];
      rew[base] = r[0];
      done[base] = d[0];
    }
  }
}

template <int ENV>
__global__ void __launch_bounds__(256) acx_env_reset_kernel(int n, float* __restrict__ state, int* __restrict__ elapsed,
                                                            float* __restrict__ obs, const float* __restrict__ reset_u,
                                                            uint64_t seed, uint64_t index0, uint32_t step) {
  constexpr int S = AcxEnv<ENV>::S, OB = AcxEnv<ENV>::OB;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    float ru[S], sv[S], ov[OB];
    if (reset_u != nullptr) {
#pragma unroll
      for (int k = 0; k < S; ++k) ru[k] = reset_u[(size_t)k * n + i];
    } else {
      AcxRng4 g = acx_philox(index0 + (uint64_t)i, step, ACX_STREAM_RESET, seed);
      float u4[4] = {acx_u01(g.x), acx_u01(g.y), acx_u01(g.z), acx_u01(g.w)};
#pragma unroll
      for (int k = 0; k < S; ++k) ru[k] = u4[k];
    }
    AcxEnv<ENV>::reset(ru, sv);
    AcxEnv<ENV>::observe(sv, ov);
#pragma unroll
    for (int k = 0; k < S; ++k) state[(size_t)k * n + i] = sv[k];
#pragma unroll
    for (int k = 0; k < OB; ++k) obs[(size_t)k * n + i] = ov[k];
    elapsed[i] = 0;
  }
}

static int env_grid(int n, int vec) {
  int blocks = (n + 256 * vec - 1) / (256 * vec);
  int cap = acx_num_sms() * 8;      // 8 resident 256-thread CTAs per SM; grid-stride beyond that
  if (blocks > cap) blocks = cap;
  return blocks < 1 ? 1 : blocks;
}

template <int ENV>
static int env_step_impl(int n, float* state, int* elapsed, const void* action, float* obs, float* rew,
                         unsigned char* done, const float* reset_u, uint64_t seed, uint64_t index0, uint32_t step,
                         const uint32_t* step_base, int max_steps, cudaStream_t st) {
  if (max_steps <= 0) max_steps = AcxEnv<ENV>::MAX_STEPS;
  bool vec = (n % 4 == 0) && acx_aligned16(state) && acx_aligned16(elapsed) && acx_aligned16(action) &&
             acx_aligned16(obs) && acx_aligned16(rew) && ((reinterpret_cast<uintptr_t>(done) & 3u) == 0);
  // Acrobot is bound by its RK4 arithmetic, not by HBM: one env per thread (4x the warps in flight, 40 instead of 74
  // registers) measured 28.9 us against 40.7 us for 1M envs and 181 against 220 us for 8M
  if (ENV == ACX_ENV_ACROBOT) vec = false;
  if (vec)
    acx_env_step_kernel<ENV, 4><<<env_grid(n, 4), 256, 0, st>>>(n, state, elapsed, action, obs, rew, done, reset_u,
                                                               seed, index0, step, step_base, max_steps);
  else
    acx_env_step_kernel<ENV, 1><<<env_grid(n, 1), 256, 0, st>>>(n, state, elapsed, action, obs, rew, done, reset_u,
                                                               seed, index0, step, step_base, max_steps);
  ACX_LAUNCH_OK("acx_env_step_kernel");
  return ACX_OK;
}

extern "C" int acx_env_step(int env, int n, float* state, int* elapsed, const void* action, float* obs_out,
                            float* rew_out, unsigned char* done_out, const float* reset_u, uint64_t seed,
                            uint64_t index0, uint32_t step, const uint32_t* step_base, int max_steps, void* stream) {
  ACX_REQUIRE(n >= 0, "acx_env_step: n must be >= 0 (got %d)", n);
  if (n == 0) return ACX_OK;
  ACX_REQUIRE(state && elapsed && action && obs_out && rew_out && done_out, "acx_env_step: null buffer");
  cudaStream_t st = (cudaStream_t)stream;
  switch (env) {
    case ACX_ENV_PENDULUM:
      return env_step_impl<ACX_ENV_PENDULUM>(n, state, elapsed, action, obs_out, rew_out, done_out, reset_u, seed,
                                             index0, step, step_base, max_steps, st);
    case ACX_ENV_CARTPOLE:
      return env_step_impl<ACX_ENV_CARTPOLE>(n, state, elapsed, action, obs_out, rew_out, done_out, reset_u, seed,
                                             index0, step, step_base, max_steps, st);
    case ACX_ENV_ACROBOT:
      return env_step_impl<ACX_ENV_ACROBOT>(n, state, elapsed, action, obs_out, rew_out, done_out, reset_u, seed,
                                            index0, step, step_base, max_steps, st);
  }
  acx_set_error("acx_env_step: unknown env id %d", env);
  return ACX_ERR_INVALID;
}

extern "C" int acx_env_step_pendulum(int n, float* state, int* elapsed, const float* action, float* obs_out,
                                     float* rew_out, unsigned char* done_out, const float* reset_u, uint64_t seed,
                                     uint64_t index0, uint32_t step, const uint32_t* step_base, int max_steps, void* stream) {
  return acx_env_step(ACX_ENV_PENDULUM, n, state, elapsed, action, obs_out, rew_out, done_out, reset_u, seed, index0,
                      step, step_base, max_steps, stream);
}
extern "C" int acx_env_step_cartpole(int n, float* state, int* elapsed, const int* action, float* obs_out,
                                     float* rew_out, unsigned char* done_out, const float* reset_u, uint64_t seed,
                                     uint64_t index0, uint32_t step, const uint32_t* step_base, int max_steps, void* stream) {
  return acx_env_step(ACX_ENV_CARTPOLE, n, state, elapsed, action, obs_out, rew_out, done_out, reset_u, seed, index0,
                      step, step_base, max_steps, stream);
}
extern "C" int acx_env_step_acrobot(int n, float* state, int* elapsed, const int* action, float* obs_out,
                                    float* rew_out, unsigned char* done_out, const float* reset_u, uint64_t seed,
                                    uint64_t index0, uint32_t step, const uint32_t* step_base, int max_steps, void* stream) {
  return acx_env_step(ACX_ENV_ACROBOT, n, state, elapsed, action, obs_out, rew_out, done_out, reset_u, seed, index0,
                      step, step_base, max_steps, stream);
}

extern "C" int acx_env_reset(int env, int n, float* state, int* elapsed, float* obs_out, const float* reset_u,
                             uint64_t seed, uint64_t index0, uint32_t step, void* stream) {
  ACX_REQUIRE(n >= 0, "acx_env_reset: n must be >= 0 (got %d)", n);
  if (n == 0) return ACX_OK;
  ACX_REQUIRE(state && elapsed && obs_out, "acx_env_reset: null buffer");
  cudaStream_t st = (cudaStream_t)stream;
  int grid = env_grid(n, 1);
  switch (env) {
    case ACX_ENV_PENDULUM:
      acx_env_reset_kernel<ACX_ENV_PENDULUM><<<grid, 256, 0, st>>>(n, state, elapsed, obs_out, reset_u, seed, index0, step);
      break;
    case ACX_ENV_CARTPOLE:
      acx_env_reset_kernel<ACX_ENV_CARTPOLE><<<grid, 256, 0, st>>>(n, state, elapsed, obs_out, reset_u, seed, index0, step);
      break;
    case ACX_ENV_ACROBOT:
      acx_env_reset_kernel<ACX_ENV_ACROBOT><<<grid, 256, 0, st>>>(n, state, elapsed, obs_out, reset_u, seed, index0, step);
      break;
    default:
      acx_set_error("acx_env_reset: unknown env id %d", env);
      return ACX_ERR_INVALID;
  }
  ACX_LAUNCH_OK("acx_env_reset_kernel");
  return ACX_OK;
}

extern "C" int acx_env_dims(int env, int* state_dim, int* obs_dim, int* n_actions, int* max_episode_steps) {
  int s, o, a, m;
  switch (env) {
    case ACX_ENV_PENDULUM: s = 2; o = 3; a = 0; m = 200; break;
    case ACX_ENV_CARTPOLE: s = 4; o = 4; a = 2; m = 200; break;
    case ACX_ENV_ACROBOT: s = 4; o = 6; a = 3; m = 500; break;
    default: acx_set_error("acx_env_dims: unknown env id %d", env); return ACX_ERR_INVALID;
  }
  if (state_dim) *state_dim = s;
  if (obs_dim) *obs_dim = o;
  if (n_actions) *n_actions = a;
  if (max_episode_steps) *max_episode_steps = m;
  return ACX_OK;
}

// ------------------------------------------------------------------ frame stacking (Framer, A3C/process.py:25-43)
// out[f][i][e] = the observation f - (frames-1) steps back from `obs`, within the episode: where the transition that
// produced `obs` ended an episode (done) -- or there is no history at all (prev == NULL) -- every frame is `obs`
// (the reference pads a fresh episode with copies of its first observation).  Frame 0 is the oldest.
__global__ void __launch_bounds__(256) acx_frame_push_kernel(int n, int ob, int frames, const float* __restrict__ prev,
                                                             const float* __restrict__ obs,
                                                             const unsigned char* __restrict__ done,
                                                             float* __restrict__ out) {
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < n; e += gridDim.x * blockDim.x) {
    const bool fresh = prev == nullptr || (done != nullptr && done[e] != 0);
    for (int i = 0; i < ob; ++i) {
      const float cur = obs[(size_t)i * n + e];
      for (int f = 0; f < frames; ++f) {
        const float v = (fresh || f == frames - 1) ? cur : prev[((size_t)(f + 1) * ob + i) * n + e];
        out[((size_t)f * ob + i) * n + e] = v;
      }
    }
  }
}

extern "C" int acx_frame_push(int n, int ob_dim, int frames, const float* prev, const float* obs,
                              const unsigned char* done, float* out, void* stream) {
  ACX_REQUIRE(n >= 0, "acx_frame_push: n must be >= 0 (got %d)", n);
  ACX_REQUIRE(ob_dim >= 1 && frames >= 1, "acx_frame_push: ob_dim %d / frames %d must be >= 1", ob_dim, frames);
  if (n == 0) return ACX_OK;
  ACX_REQUIRE(obs && out, "acx_frame_push: null obs / out");
  ACX_REQUIRE(out != prev, "acx_frame_push: out must not alias prev (frames are shifted)");
  acx_frame_push_kernel<<<env_grid(n, 1), 256, 0, (cudaStream_t)stream>>>(n, ob_dim, frames, prev, obs, done, out);
  ACX_LAUNCH_OK("acx_frame_push_kernel");
  return ACX_OK;
}
