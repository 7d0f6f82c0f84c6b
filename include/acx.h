/* libacx -- C ABI of the B200-native actor-critic hot path.
 *
 * The reference (renly/Actor-Critic-Algs-on-Tensorflow) has no FFI of its own: its hot
 * path is Python calling TensorFlow 1.3 sessions and gym.  This header is the boundary a
 * maintainer binds instead (INTEGRATION.md shows the ctypes stub); every entry point cites
 * the reference code it replaces.  Host code above this ABI lives in
 * actor_critic_algs_on_tensorflow_b200/ and mirrors the reference's Python interface.
 *
 * Conventions
 *   - plain C types only; every pointer is a DEVICE pointer unless stated otherwise;
 *   - nothing is allocated, freed or synchronised inside: the caller owns every buffer,
 *     including workspaces whose size comes from the matching *_workspace_bytes();
 *   - `stream` is a cudaStream_t passed as void*; all work is stream-ordered on it;
 *   - return value: ACX_OK (0) or a negative ACX_ERR_*; acx_last_error() (thread-local)
 *     gives the message;
 *   - batches are structure-of-arrays: obs[rows][ob_dim][n], actions[rows][ac_dim][n] (float)
 *     or [rows][n] (int32), per-sample scalars [rows][n]; rollout buffers are time-major
 *     [T][n];
 *   - network parameters are ONE flat fp32 buffer per network, 16-byte aligned, laid out by
 *     acx_params_layout() (kernels [in][out] row-major as in the TF checkpoint);
 *   - device RNG: Philox4x32-10, counter (index0 + column, step0 + row, stream id), key = seed;
 *     `step_base` (nullable DEVICE pointer to one uint32) is added to the step, so that a
 *     captured CUDA graph can be replayed with fresh randomness (acx_counter_add advances it).
 */
#ifndef ACX_H_
#define ACX_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ACX_ABI_VERSION 1

#define ACX_OK 0
#define ACX_ERR_INVALID (-1)   /* bad argument (null pointer, size, alignment) */
#define ACX_ERR_CUDA (-2)      /* a CUDA runtime call or launch failed */
#define ACX_ERR_UNSUPPORTED (-3)

#define ACX_ENV_PENDULUM 0     /* Pendulum-v0 */
#define ACX_ENV_CARTPOLE 1     /* CartPole-v0 */
#define ACX_ENV_ACROBOT 2      /* Acrobot-v1 */

#define ACX_SCAN_AUTO 0
#define ACX_SCAN_COLS 1        /* one thread per env column (wide batches) */
#define ACX_SCAN_WARP 2        /* warp-level segmented scan over time (few long columns) */

/* device "opt" block: float[ACX_OPT_SIZE] -- the non-trainable variables of
 * A3C/policies.py:77-79,151 plus the Adam bookkeeping TF keeps in slot variables */
#define ACX_OPT_SIZE 16
#define ACX_OPT_ACTOR_LR 0
#define ACX_OPT_CRITIC_LR 1
#define ACX_OPT_BETA_KL 2
#define ACX_OPT_GAMMA_ENT 3
#define ACX_OPT_BETA1_POW 4    /* beta1^t, updated by repeated fp32 multiplication like TF */
#define ACX_OPT_BETA2_POW 5
#define ACX_OPT_CLIP 6         /* elementwise actor gradient clip (0.1 A3C, 1.0 Basic_AC); <=0: off */
#define ACX_OPT_DESIRED_KL 7
#define ACX_OPT_MIN_LR 8
#define ACX_OPT_MAX_LR 9
#define ACX_OPT_ITER 10        /* local iteration counter i (float, exact below 2^24) */
#define ACX_OPT_SCHEDULE 11    /* != 0: apply the beta/gamma LinearSchedule on device */
#define ACX_OPT_ADAM_T 12

/* device "stats" block: float[ACX_STATS_SIZE] */
#define ACX_STATS_SIZE 16
#define ACX_STAT_ADV_MEAN 0
#define ACX_STAT_ADV_INV_STD 1
#define ACX_STAT_ADV_STD 2
#define ACX_STAT_ACTOR_LOSS 3
#define ACX_STAT_CRITIC_LOSS 4
#define ACX_STAT_KL_PRE 5
#define ACX_STAT_ENT_MEAN 6
#define ACX_STAT_KL_POST 7
#define ACX_STAT_REW_LOSS 8
#define ACX_STAT_ACTOR_LR_USED 9
#define ACX_STAT_EV_BEFORE 10     /* var_accounted_for(target = G, pred = V[:T]) -- A3C/util.py:5-13, process.py:111-116 */
#define ACX_STAT_EV_AFTER 11      /* the same after the update (acx_explained_variance on fresh values, process.py:261) */
#define ACX_STAT_EP_REW_MEAN 12   /* mean return of the episodes that ended in the batch (avg_rew, process.py:243) */
#define ACX_STAT_EP_COUNT 13      /* how many ended */

int acx_abi_version(void);
const char* acx_last_error(void);

/* Flat parameter layout.  net: 0 = actor (A3C/policies.py:44-52,63), 1 = critic (:146-150).
 * offsets[9] = {W1,b1,W2,b2,W3,b3,Whead,bhead,log_std(-1 if none)}, sizes[9] likewise (element
 * counts); returns the padded total (floats) or a negative error.  Every tensor starts on a
 * 4-float boundary.  HOST pointers. */
int acx_params_layout(int net, int ob_dim, int num_ac, int cont, int* offsets, int* sizes);

/* ---------------------------------------------------------------- K1: env step
 * Replaces env.reset() / env.step(ac) -- A3C/process.py:79,91 (gym classic_control) -- for n
 * envs.  state[S][n], elapsed[n] int32 are updated in place; obs_out[ob][n], rew_out[n],
 * done_out[n] (uint8) are written.  Where done (termination or TimeLimit: elapsed >=
 * max_steps, max_steps <= 0 -> gym's default) the env is reset at once and obs_out is the
 * new episode's first observation.  reset_u[S][n] uniforms in [0,1) or NULL (Philox). */
int acx_env_dims(int env, int* state_dim, int* obs_dim, int* n_actions, int* max_episode_steps);
int acx_env_reset(int env, int n, float* state, int* elapsed, float* obs_out, const float* reset_u,
                  uint64_t seed, uint64_t index0, uint32_t step, void* stream);
int acx_env_step(int env, int n, float* state, int* elapsed, const void* action, float* obs_out,
                 float* rew_out, unsigned char* done_out, const float* reset_u, uint64_t seed,
                 uint64_t index0, uint32_t step, const uint32_t* step_base, int max_steps, void* stream);
int acx_env_step_pendulum(int n, float* state, int* elapsed, const float* action, float* obs_out,
                          float* rew_out, unsigned char* done_out, const float* reset_u, uint64_t seed,
                          uint64_t index0, uint32_t step, const uint32_t* step_base, int max_steps,
                          void* stream);
int acx_env_step_cartpole(int n, float* state, int* elapsed, const int* action, float* obs_out,
                          float* rew_out, unsigned char* done_out, const float* reset_u, uint64_t seed,
                          uint64_t index0, uint32_t step, const uint32_t* step_base, int max_steps,
                          void* stream);
int acx_env_step_acrobot(int n, float* state, int* elapsed, const int* action, float* obs_out,
                         float* rew_out, unsigned char* done_out, const float* reset_u, uint64_t seed,
                         uint64_t index0, uint32_t step, const uint32_t* step_base, int max_steps,
                          void* stream);

/* Framer (A3C/process.py:25-43, `--frames k`): the network input is the last `frames` observations of the episode,
 * oldest first, a fresh episode padded with copies of its first observation.  prev[frames*ob_dim][n] is the stacked
 * input one step earlier (NULL: no history, e.g. right after acx_env_reset), obs[ob_dim][n] the new observation,
 * done[n] the done flags of the transition that produced it (NULL: none ended); out[frames*ob_dim][n] must not
 * alias prev.  The MLP entry points then run with ob_dim = frames * ob_dim: at most 24 on the FFMA entry points (three
 * frames of Acrobot's six observations: stack_frames = 3 is the default of A3C/process.py:156-157), at most 8 on the
 * *_tc ones, whose shared memory belongs to the weights. */
int acx_frame_push(int n, int ob_dim, int frames, const float* prev, const float* obs, const unsigned char* done,
                   float* out, void* stream);

/* ---------------------------------------------------------------- K2: policy / value forward
 * Actor.act, Gaussian head -- A3C/policies.py:44-58,110-115.  eps[rows][ac_dim][n] standard
 * normals or NULL (Philox; ac_dim <= 2).  Outputs: ac_out[rows][ac_dim][n], logp_out,
 * ent_out, mu_out (the last three nullable). */
int acx_policy_forward_sample_gauss(const float* actor, int ob_dim, int ac_dim, float ac_scale, int rows, int n,
                                    const float* obs, const float* eps, uint64_t seed, uint64_t index0,
                                    uint32_t step0, const uint32_t* step_base, float* ac_out, float* logp_out, float* ent_out,
                                    float* mu_out, void* stream);
/* Actor.act, categorical head -- A3C/policies.py:63-67.  u[rows][n] uniforms in [0,1) or NULL.
 * tf.multinomial is restated as an inverse CDF in double (parity unpinned). */
int acx_policy_forward_sample_cat(const float* actor, int ob_dim, int n_act, int rows, int n, const float* obs,
                                  const float* u, uint64_t seed, uint64_t index0, uint32_t step0,
                                  const uint32_t* step_base, int* ac_out,
                                  float* logp_out, float* ent_out, float* logits_out, void* stream);
/* log-prob of stored actions under the current weights (A3C/policies.py:60,69) and, when
 * logp_old is given, the per-block partial sums of (logp_old - logp)^2 for Actor.get_kl
 * (A3C/policies.py:75,106-108): kl_partial[*kl_blocks] doubles (acx_kl_workspace_bytes()). */
size_t acx_kl_workspace_bytes(void);
int acx_policy_logp_gauss(const float* actor, int ob_dim, int ac_dim, float ac_scale, int rows, int n,
                          const float* obs, const float* acs, float* logp_out, float* ent_out, float* mu_out,
                          const float* logp_old, double* kl_partial, int* kl_blocks /*host*/, void* stream);
int acx_policy_logp_cat(const float* actor, int ob_dim, int n_act, int rows, int n, const float* obs,
                        const int* acs, float* logp_out, float* ent_out, float* logits_out,
                        const float* logp_old, double* kl_partial, int* kl_blocks /*host*/, void* stream);
/* Critic.value -- A3C/policies.py:146-150,176-177; basic_head != 0 reproduces
 * Basic_AC/policies.py:134-135 (value head reads the second layer). */
int acx_value_forward(const float* critic, int ob_dim, int rows, int n, const float* obs, float* v_out,
                      int basic_head, void* stream);

/* Activation cache.  The reference makes ONE update per batch, on the weights that produced the batch
 * (A3C/process.py:221-258), so the hidden activations the rollout and value passes compute are exactly what
 * the update's forward pass would recompute.  The *_cached forwards also write them -- x2_out (second hidden
 * layer, 128 features) and x3_out (third, 64 actor | 128 critic) -- and acx_loss_backward_*_cached reads them
 * instead of re-running those two GEMMs (650 k -> 500 k FLOP per env-step).  Results are
 * bit-identical to the recomputing entry points.
 * Layout: tile-blocked, [rows][ceil(n / 128)][features][128] fp32 -- the 128-sample tile the kernels work on is one
 * contiguous features x 512-byte block (one DRAM page / TLB entry per tile instead of `features` rows 4 n bytes
 * apart); samples beyond n are written as zeros.  Size: acx_activation_cache_floats(rows, features, n) floats,
 * 16-byte aligned.  A forward call with rows = 1 on row r of a [R]-row cache writes at
 * cache + r * acx_activation_cache_floats(1, features, n). */
size_t acx_activation_cache_floats(int rows, int features, int n);
int acx_policy_forward_sample_gauss_cached(const float* actor, int ob_dim, int ac_dim, float ac_scale, int rows,
                                           int n, const float* obs, const float* eps, uint64_t seed,
                                           uint64_t index0, uint32_t step0, const uint32_t* step_base,
                                           float* ac_out, float* logp_out, float* ent_out, float* mu_out,
                                           float* x2_out, float* x3_out, void* stream);
int acx_policy_forward_sample_cat_cached(const float* actor, int ob_dim, int n_act, int rows, int n,
                                         const float* obs, const float* u, uint64_t seed, uint64_t index0,
                                         uint32_t step0, const uint32_t* step_base, int* ac_out, float* logp_out,
                                         float* ent_out, float* logits_out, float* x2_out, float* x3_out,
                                         void* stream);
int acx_value_forward_cached(const float* critic, int ob_dim, int rows, int n, const float* obs, float* v_out,
                             int basic_head, float* x2_out, float* x3_out, void* stream);

/* ---------------------------------------------------------------- K2 on the tensor cores (tcgen05, 3xTF32)
 * Same contracts as the five forward entry points above (same reference lines), with the two 128-wide
 * hidden GEMMs issued as tcgen05.mma kind::tf32 on error-compensated operand splits (x = hi + lo;
 * hi.hi + hi.lo + lo.hi accumulated in fp32 in tensor memory) -- fp32-class accuracy, checked to the
 * same 1e-5 tolerance.  They take, besides the flat parameters (layer 1, biases and head are read from
 * there), the hidden-layer kernels pre-split and pre-arranged for the tensor cores:
 * acx_tc_pack_weights() fills `packed` (acx_tc_packed_floats() floats, 16-byte aligned) and must be
 * re-run whenever the parameters change.  x2_out / x3_out (both or neither, nullable) fill the activation
 * cache like the *_cached forwards.  The FFMA entry points stay the bit-stable reference. */
size_t acx_tc_packed_floats(int net, int ob_dim, int num_ac, int cont);
int acx_tc_pack_weights(int net, const float* params, int ob_dim, int num_ac, int cont, float* packed,
                        void* stream);
int acx_policy_forward_sample_gauss_tc(const float* actor, const float* packed, int ob_dim, int ac_dim,
                                       float ac_scale, int rows, int n, const float* obs, const float* eps,
                                       uint64_t seed, uint64_t index0, uint32_t step0,
                                       const uint32_t* step_base, float* ac_out, float* logp_out,
                                       float* ent_out, float* mu_out, float* x2_out, float* x3_out,
                                       void* stream);
int acx_policy_forward_sample_cat_tc(const float* actor, const float* packed, int ob_dim, int n_act, int rows,
                                     int n, const float* obs, const float* u, uint64_t seed, uint64_t index0,
                                     uint32_t step0, const uint32_t* step_base, int* ac_out, float* logp_out,
                                     float* ent_out, float* logits_out, float* x2_out, float* x3_out,
                                     void* stream);
/* The whole rollout of A3C/process.py:73-98 for n environments and T steps in ONE launch: per step the policy forward +
 * sampling of acx_policy_forward_sample_*_tc and the transition of acx_env_step (same arithmetic, same Philox
 * counters: action noise and reset uniforms of step t use step = *step_base + t), with each environment's state in
 * registers and the next observation handed to layer 1 through shared memory.  A CTA keeps its 128-env tiles for all T
 * steps (environments are independent).  env: ACX_ENV_*; the head follows the env (Pendulum: Gaussian with ac_scale,
 * CartPole / Acrobot: categorical).  obs[T+1][ob][n]: row 0 is the current observation (input), rows 1..T are written;
 * state[S][n] / elapsed[n] are updated in place; noise: eps[T][1][n] (Gaussian) or u[T][n] (categorical) or NULL
 * (Philox); reset_u[T][S][n] or NULL; outputs ac_out[T][1][n] float | [T][n] int32, logp_out / ent_out [T][n]
 * (nullable), rew_out[T][n], done_out[T][n]; x2_out / x3_out: activation cache rows 0..T-1 (both or neither).
 * Results are identical, bit for bit, to T x {acx_policy_forward_sample_*_tc; acx_env_step}. */
int acx_rollout_tc(int env, const float* actor, const float* packed, float ac_scale, int T, int n, float* obs,
                   float* state, int* elapsed, const float* noise, const float* reset_u, uint64_t seed,
                   uint64_t index0, const uint32_t* step_base, int max_steps, void* ac_out, float* logp_out,
                   float* ent_out, float* rew_out, unsigned char* done_out, float* x2_out, float* x3_out, unsigned int* mask_out,
                   void* stream);
int acx_policy_logp_gauss_tc(const float* actor, const float* packed, int ob_dim, int ac_dim, float ac_scale,
                             int rows, int n, const float* obs, const float* acs, float* logp_out,
                             float* ent_out, float* mu_out, const float* logp_old, double* kl_partial,
                             int* kl_blocks /*host*/, void* stream);
int acx_policy_logp_cat_tc(const float* actor, const float* packed, int ob_dim, int n_act, int rows, int n,
                           const float* obs, const int* acs, float* logp_out, float* ent_out,
                           float* logits_out, const float* logp_old, double* kl_partial,
                           int* kl_blocks /*host*/, void* stream);
int acx_value_forward_tc(const float* critic, const float* packed, int ob_dim, int rows, int n,
                         const float* obs, float* v_out, int basic_head, float* x2_out, float* x3_out, unsigned int* mask_out,
                         void* stream);

/* ---------------------------------------------------------------- K3: returns / advantages
 * PathAdv.__call__ -- A3C/process.py:54-66 -- over time-major [T][n] buffers with done masks
 * (each column may hold several episode segments; the tail is cut by the horizon and
 * bootstraps from V[T]); plus the batch moments of A3C/process.py:247:
 * stats_out[0] = mean(adv), [1] = 1/(1e-8 + std(adv)), [2] = std(adv) (population), and the
 * critic diagnostic the reference logs as ev_before (A3C/process.py:111-116, util.py:5-13):
 * stats_out[ACX_STAT_EV_BEFORE] = mean(standardised G * standardised V[:T]).
 * V is [T+1][n].  workspace: 16-byte aligned, first 16 bytes zero before the first call. */
size_t acx_returns_adv_workspace_bytes(int T, int n);
int acx_returns_adv_scan(int T, int n, const float* rew, const unsigned char* done, const float* V,
                         double gamma, int look_ahead, float* G_out, float* adv_out, float* stats_out,
                         void* workspace, size_t workspace_bytes, int mode, void* stream);
/* The same scan for callers that have no way to hand over a workspace.  (No counterpart in the
 * reference: this entry point exists only for embedding.)  Think of a rollout worker living inside
 * another framework's process, where that framework's pool has already claimed the GPU's memory and
 * a fresh cudaMalloc fails; of a control loop that cannot afford the device-wide synchronisation an
 * allocation implies; or of a sandboxed plugin thread that is given input and output buffers and
 * nothing else.  Inputs and outputs are still the caller's; only the few KB of partial-moment
 * scratch come from a fixed array inside the library's device image (about 28 KB that every GPU
 * loading the library keeps until the process exits).
 *   - Limit: n <= ACX_SCAN_STATIC_MAX_COLUMNS env columns per call (T is unlimited); wider batches
 *     are refused with ACX_ERR_INVALID and a message -- use acx_returns_adv_scan for those.
 *   - Nothing is allocated or freed, here or in any library it calls.
 *   - The first call on each device looks up the array's device address (that device must be
 *     current; the look-up is serialised between host threads); make that first call before any
 *     stream capture begins.  cudaDeviceReset() invalidates the cached address: do not call this
 *     entry point again on that device afterwards.
 *   - All calls on a device share the one array and it keeps the previous call's contents; the
 *     entry point re-zeroes the counter it depends on (stream-ordered) before launching.  Two calls
 *     whose kernels can overlap on the GPU (different streams, no dependency) race on it: keep
 *     calls to this entry point ordered with respect to each other. */
#define ACX_SCAN_STATIC_MAX_COLUMNS 131072
int acx_returns_adv_scan_static(int T, int n, const float* rew, const unsigned char* done, const float* V,
                                double gamma, int look_ahead, float* G_out, float* adv_out,
                                float* stats_out, int mode, void* stream);

/* ---------------------------------------------------------------- diagnostics of the training log
 * The reference prints avg_rew, ev_before and ev_after every iteration (A3C/process.py:243,261,280-283).  These keep
 * them on the device: workspace = acx_diag_workspace_bytes() bytes, 16-byte aligned, first 16 bytes zero before the
 * first call (each entry point may have its own or share one, stream-ordered).
 * acx_episode_returns: rew/done [T][n] of one rollout; ep_ret[n] (in/out) is the running return of each column's
 * unfinished episode (a NaN marks an episode whose start was not seen: dropped when it ends).  Writes
 * stats[ACX_STAT_EP_COUNT] = episodes that ended in the batch and, when that is > 0,
 * stats[ACX_STAT_EP_REW_MEAN] = their mean return (tot_rews / rolls, process.py:234,243).
 * acx_explained_variance: *out = mean(standardised target * standardised pred) over count samples
 * (var_accounted_for, A3C/util.py:5-13). */
size_t acx_diag_workspace_bytes(void);
int acx_episode_returns(int T, int n, const float* rew, const unsigned char* done, float* ep_ret, float* stats,
                        void* workspace, size_t workspace_bytes, void* stream);
int acx_explained_variance(long long count, const float* target, const float* pred, float* out, void* workspace,
                           size_t workspace_bytes, void* stream);
/* One log row per iteration without a host sync per iteration (the reference logs every iteration,
 * A3C/process.py:280-281): copies stats[16] ‖ opt[16] into row (opt[ITER] - 1) mod rows of ring[rows][32]; call it
 * after acx_kl_lr_update and read the ring back at leisure. */
int acx_log_push(float* ring, int rows, const float* stats, const float* opt, void* stream);
/* Spread the TimeLimit counters of n freshly reset environments over [0, max_steps) (Philox stream 3 of
 * (seed, index0 + i)), so that a batched rollout sees every phase of the episode in every update, as the reference's
 * whole-episode batches do (A3C/process.py:221-235).  max_steps <= 0: the env's default. */
int acx_env_stagger(int env, int n, int* elapsed, uint64_t seed, uint64_t index0, int max_steps, void* stream);

/* ---------------------------------------------------------------- K4: loss + backward
 * Actor loss  -mean(adv*lp) + beta*mean((lp_old-lp)^2) - gamma*mean(entropy) and its gradient
 * (A3C/policies.py:74-84, unclipped here: the clip of :85 is applied in acx_adam_apply after
 * the cross-rank reduction) and critic loss mean((v-target)^2) with gradient (:152-158), in
 * one call.  adv is raw; adv_stats (device {mean, 1/(1e-8+std)}, e.g. K3's stats_out) applies
 * the normalisation of A3C/process.py:247 on the fly (NULL: use adv as given).  beta/gamma are
 * read from the device opt block.  Either network may be NULL to skip it.  Writes
 * grad_actor / grad_critic (flat, same layout as the parameters) and stats[ACX_STAT_*]. */
size_t acx_loss_backward_workspace_bytes(int ob_dim, int num_ac, int cont);
int acx_loss_backward_gauss(const float* actor, const float* critic, int ob_dim, int ac_dim, float ac_scale,
                            int rows, int n, const float* obs, const float* acs, const float* adv,
                            const float* adv_stats, const float* logp_old, const float* targets,
                            const float* opt, float* grad_actor, float* grad_critic, float* stats,
                            void* workspace, size_t workspace_bytes, void* stream);
int acx_loss_backward_cat(const float* actor, const float* critic, int ob_dim, int n_act, int rows, int n,
                          const float* obs, const int* acs, const float* adv, const float* adv_stats,
                          const float* logp_old, const float* targets, const float* opt, float* grad_actor,
                          float* grad_critic, float* stats, void* workspace, size_t workspace_bytes,
                          void* stream);

/* K4 reading the activation cache (see the *_cached forwards); a null x2/x3 pair recomputes that network. */
int acx_loss_backward_gauss_cached(const float* actor, const float* critic, int ob_dim, int ac_dim, float ac_scale,
                                   int rows, int n, const float* obs, const float* acs, const float* adv,
                                   const float* adv_stats, const float* logp_old, const float* targets,
                                   const float* opt, float* grad_actor, float* grad_critic, float* stats,
                                   const float* actor_x2, const float* actor_x3, const float* critic_x2,
                                   const float* critic_x3, void* workspace, size_t workspace_bytes, void* stream);
int acx_loss_backward_cat_cached(const float* actor, const float* critic, int ob_dim, int n_act, int rows, int n,
                                 const float* obs, const int* acs, const float* adv, const float* adv_stats,
                                 const float* logp_old, const float* targets, const float* opt, float* grad_actor,
                                 float* grad_critic, float* stats, const float* actor_x2, const float* actor_x3,
                                 const float* critic_x2, const float* critic_x3, void* workspace,
                                 size_t workspace_bytes, void* stream);

/* Same contract again (A3C/policies.py:74-85,152-158 on the cached activations), with the two back-propagation
 * chain products  dX2 = dPre3 . W3^T  and  dX1 = dPre2 . W2^T  on the tcgen05 tensor cores (3xTF32 split, fp32
 * accumulation, ~1e-6 relative per product) while the weight-gradient products run on the CUDA cores of the same
 * CTA.  *_packed_bwd: acx_tc_packed_bwd_floats() floats each, 16-byte aligned, filled by
 * acx_tc_pack_weights_bwd() and re-filled whenever the parameters change.  The activation cache is required.
 * Gradients agree with acx_loss_backward_*_cached to 3xTF32 accuracy, not bit for bit.
 * Workspace: with at least acx_loss_backward_workspace_bytes() the weight-gradient products run as fp32 FFMA
 * inside the same kernel; with at least acx_loss_backward_tc_workspace_bytes(ob_dim, num_ac, cont, rows, n) (room for dPre3 and dPre2 of
 * one network, tile-blocked like the cache) and n % 4 == 0 they run on the tensor cores too, in a second kernel
 * per network (acx_wgrad_kernel: TMA tensor loads, 128-byte swizzle, accumulators in tensor memory). */
size_t acx_loss_backward_tc_workspace_bytes(int ob_dim, int num_ac, int cont, int rows, int n);
/* Activation sign masks (optional, nullable everywhere): acx_value_forward_tc / acx_rollout_tc can store, per 128-sample
 * tile, one 32-bit word per 32 features and sample of "X1 > 0" and "X2 > 0" (acx_act_mask_words(net, rows, n) uint32
 * words, 16-byte aligned; net 0 = actor, 1 = critic); acx_loss_backward_*_tc given the same array (actor_masks /
 * critic_masks) reads act'(X1) / act'(X2) from it instead of re-deriving them from the X2 cache (64 loads per thread)
 * and the observations.  Same results. */
size_t acx_act_mask_words(int net, int rows, int n);
size_t acx_tc_packed_bwd_floats(int net, int ob_dim, int num_ac, int cont);
int acx_tc_pack_weights_bwd(int net, const float* params, int ob_dim, int num_ac, int cont, float* packed,
                            void* stream);
int acx_loss_backward_gauss_tc(const float* actor, const float* critic, const float* actor_packed_bwd,
                               const float* critic_packed_bwd, int ob_dim, int ac_dim, float ac_scale, int rows,
                               int n, const float* obs, const float* acs, const float* adv,
                               const float* adv_stats, const float* logp_old, const float* targets,
                               const float* opt, float* grad_actor, float* grad_critic, float* stats,
                               const float* actor_x2, const float* actor_x3, const float* critic_x2,
                               const float* critic_x3, const unsigned int* actor_masks,
                               const unsigned int* critic_masks, void* workspace, size_t workspace_bytes, void* stream);
int acx_loss_backward_cat_tc(const float* actor, const float* critic, const float* actor_packed_bwd,
                             const float* critic_packed_bwd, int ob_dim, int n_act, int rows, int n,
                             const float* obs, const int* acs, const float* adv, const float* adv_stats,
                             const float* logp_old, const float* targets, const float* opt, float* grad_actor,
                             float* grad_critic, float* stats, const float* actor_x2, const float* actor_x3,
                             const float* critic_x2, const float* critic_x3, const unsigned int* actor_masks,
                               const unsigned int* critic_masks, void* workspace,
                             size_t workspace_bytes, void* stream);

/* Critic loss + gradient for the Basic_AC network, whose value head reads the SECOND layer
 * (Basic_AC/policies.py:134-135): the third layer gets a zero gradient, exactly as TF reports None
 * for it.  Workspace as for acx_loss_backward_gauss(ob_dim, 1, 1). */
int acx_loss_backward_basic_critic(const float* critic, int ob_dim, int rows, int n, const float* obs,
                                   const float* targets, float* grad_critic, float* stats, void* workspace,
                                   size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------- K5: optimiser + lr control
 * acx_opt_init fills a device opt block (HOST values in, one small kernel). */
int acx_opt_init(float* opt, float actor_lr, float critic_lr, float beta_kl, float gamma_ent, float clip,
                 float desired_kl, float min_lr, float max_lr, int schedule_on, void* stream);
/* tf.train.AdamOptimizer.apply_gradients (A3C/policies.py:82,88-89,155,160-161), TF ApplyAdam
 * arithmetic, for both networks in one launch.  grad *= grad_scale first (1/world after an
 * all-reduce sum), actor grads are then clipped to +-opt[CLIP] (policies.py:85).  Reads the
 * beta powers from opt; acx_adam_tick advances them (and the step count) afterwards. */
int acx_adam_apply(float* actor, const float* grad_actor, float* m_actor, float* v_actor, int n_actor,
                   float* critic, const float* grad_critic, float* m_critic, float* v_critic, int n_critic,
                   const float* opt, float grad_scale, void* stream);
int acx_adam_tick(float* opt, void* stream);
/* *counter += inc (one uint32 on device): advances the RNG step base inside a captured graph. */
int acx_counter_add(uint32_t* counter, uint32_t inc, void* stream);
/* Actor.get_kl tail + the lr rule of A3C/process.py:262-270 + the LinearSchedule updates of
 * :271-278, all on device: acx_kl_finalize folds kl_partial into stats[ACX_STAT_KL_POST]
 * (mean over count); acx_kl_lr_update applies lr*1.5 / lr/1.5 with the clamp, the schedules
 * (when opt[SCHEDULE] != 0) and increments opt[ITER].  Split so that a cross-rank average of
 * stats[KL_POST] can sit between them. */
int acx_kl_finalize(const double* kl_partial, int kl_blocks, double count, float* stats, void* stream);
int acx_kl_lr_update(float* opt, float* stats, void* stream);
/* The two halves of acx_kl_lr_update on their own, for data-parallel ranks that exchange ONE buffer per update (the
 * reference's workers push and pull once, A3C/policies.py:88-98): flags = 1 applies the lr rule of
 * A3C/process.py:262-270 only, flags = 2 the LinearSchedule updates of :271-278 and the opt[ITER] bump only, 3 both.
 * kl_sum (nullable DEVICE pointer): the rule then uses *kl_sum * kl_scale -- the post-update KL summed over ranks, which
 * travelled as one extra float behind the flat gradient of the NEXT update's all-reduce; the lr is first consumed by
 * that update's acx_adam_apply, so applying the rule there changes nothing -- and leaves it in stats[ACX_STAT_KL_POST]. */
int acx_kl_lr_update_split(float* opt, float* stats, const float* kl_sum, float kl_scale, int flags, void* stream);

/* ---- the exchange of a synchronous multi-GPU update, fused with the optimiser step (csrc/acx_xgpu.cu) ----
 * Replaces the workers' gradient pushes onto the ps variables and the pull (A3C/policies.py:88-98,
 * A3C/process.py:257-260) on the GPUs of ONE node, without a collective library: every rank keeps
 * [flat gradient (actor ‖ critic) ‖ tail ‖ flags] in an exchange buffer that all ranks have mapped into their address
 * space (the caller maps it -- torch symmetric memory, CUDA IPC, cuMemMap; nothing is allocated here), and ONE kernel
 * per rank (a) waits until every rank's gradient is complete, (b) adds the buffers of ranks 0 .. world-1 in that fixed
 * order with plain loads over NVLink (bit-identical sums on every rank), (c) applies the lr rule of
 * A3C/process.py:262-270 to tail[0] * scale when apply_lr_rule != 0 (the previous update's KL, see
 * acx_kl_lr_update_split), (d) runs acx_adam_apply's arithmetic (x scale, clip of the actor's gradient, Adam) on its
 * elements, and (e) returns only when every peer has read this rank's buffer, so the caller may overwrite it at once.
 * It does NOT write the opt block: follow it with acx_kl_lr_update_split(flags = 1, kl_sum = grad_sum + n_actor +
 * n_critic, kl_scale = scale) when apply_lr_rule != 0 (records the same lr) and with acx_adam_tick.
 *   peer_bufs   HOST array of world device pointers: rank r's exchange buffer as mapped on this device
 *               (peer_bufs[rank] is this rank's own); each acx_xgpu_buffer_bytes(n_actor + n_critic + n_tail) long,
 *               flags (from acx_xgpu_flag_offset_bytes on) zeroed on every rank before the first call
 *   grad_sum    local, n_actor + n_critic + n_tail floats: receives the summed block (nullable)
 *   generation  local, ACX_XGPU_MAX_BLOCKS words, zero before the first call, owned by this entry point afterwards
 * All ranks must make the same sequence of calls (it is a collective).  Graph-capturable.  A rank that never arrives:
 * the others wait 10 s, print which rank was missing and trap (launch error), they do not hang.
 * world <= ACX_XGPU_MAX_RANKS; counts are multiples of 4 floats and at most 1024 * ACX_XGPU_MAX_BLOCKS in total. */
#define ACX_XGPU_MAX_RANKS 8
#define ACX_XGPU_MAX_BLOCKS 128
size_t acx_xgpu_flag_offset_bytes(int n_floats);
size_t acx_xgpu_buffer_bytes(int n_floats);
int acx_allreduce_adam(const void* const* peer_bufs, int world, int rank, int n_actor, int n_critic, int n_tail,
                       float* actor, float* m_actor, float* v_actor, float* critic, float* m_critic, float* v_critic,
                       const float* opt, float* grad_sum, unsigned int* generation, float scale, int apply_lr_rule,
                       void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ACX_H_ */
