"""CPU restatement of one synchronous rollout+update iteration, vectorised over N
environments -- the same sequence the CUDA trainer runs, built from the oracle pieces.
TEST INFRASTRUCTURE and the timed CPU baseline of bench.py (see ``oracle/__init__``).

It follows the reference's outer loop body (``A3C/process.py:217-285``):
rollout with the current actor (:221-222,86-96) -> critic values on every visited
observation (:229) -> PathAdv returns/advantages (:230) -> batch advantage normalisation
(:247) -> one critic and one actor Adam step (:257-258) -> post-update KL on the same
batch (:262) -> lr rule (:263-270) -> beta/gamma schedules every 100 iterations (:271-278).
The batch is N envs x T steps with auto-reset instead of ~6 whole episodes.
"""
import numpy as np

from . import envs as E
from . import philox as R
from . import policies as P
from .pathadv import linear_schedule_val, path_adv_columns, var_accounted_for


class A2CState:
    def __init__(self, env, n_envs, seed=12321, rank=0, world=1, frames=1, dtype=np.float32,
                 actor=None, critic=None, actor_lr=0.005, critic_lr=0.001, beta=1.0, gamma_ent=0.01, stagger=False,
                 max_steps=None, basic=False):
        self.env, self.N, self.seed, self.rank, self.world = env, n_envs, seed, rank, world
        self.f = dtype
        self.cont = env == E.PENDULUM
        self.basic = bool(basic)                       # Basic_AC critic: value head on the second layer (policies.py:134-135)
        self.frames = int(frames)                      # Framer (A3C/process.py:25-43): network input = last k observations
        self.env_ob = E.OBS_DIM[env]
        self.ob_dim = self.env_ob * self.frames
        self.num_ac = 1 if self.cont else E.N_ACTIONS[env]
        self.ac_scale = 2.0
        self.alay = P.actor_layout(self.ob_dim, self.num_ac, self.cont)
        self.clay = P.critic_layout(self.ob_dim)
        rng = np.random.default_rng(seed)
        self.actor = P.glorot_init(self.alay, rng) if actor is None else np.array(actor, dtype=np.float32)
        self.critic = P.glorot_init(self.clay, rng) if critic is None else np.array(critic, dtype=np.float32)
        self.am = np.zeros_like(self.actor)
        self.av = np.zeros_like(self.actor)
        self.cm = np.zeros_like(self.critic)
        self.cv = np.zeros_like(self.critic)
        self.t_adam = 0
        self.actor_lr, self.critic_lr, self.beta, self.gamma_ent = actor_lr, critic_lr, beta, gamma_ent
        self.it = 0
        self.gstep = 0        # global env-step counter = Philox "step"
        self.idx = (np.arange(n_envs, dtype=np.uint64) + np.uint64(rank * n_envs))
        u = R.uniforms(self.idx, 0xFFFFFFFF, R.STREAM_RESET, seed).T            # [4, N]
        self.state = E.reset_from_uniforms(env, u[:E.STATE_DIM[env]], dtype)
        self.elapsed = np.zeros(n_envs, dtype=np.int32)
        self.ep_ret = np.zeros(n_envs, dtype=np.float32)       # running return of each column's unfinished episode
        if stagger:     # the batched trainer's phase stagger (acx_env_stagger): time-limit counters over [0, max_steps)
            limit = int(max_steps or E.MAX_EPISODE_STEPS[env])
            self.elapsed = (R.philox4x32(self.idx, 0, R.STREAM_PHASE, seed)[:, 0] % np.uint32(limit)).astype(np.int32)
            self.ep_ret[self.elapsed > 0] = np.nan              # started mid-way: not counted when it ends
        self.obs = np.tile(E.observe(env, self.state, dtype), (self.frames, 1))     # fresh episode: k copies


def rollout(S, T, max_steps=None):
    """T steps of act -> env.step over all envs.  Returns dict of [T,...] buffers."""
    f = S.f
    obs_buf = np.zeros((T + 1, S.ob_dim, S.N), dtype=f)
    acs = np.zeros((T, S.num_ac if S.cont else 1, S.N), dtype=f if S.cont else np.int32)
    logps = np.zeros((T, S.N), dtype=f)
    rew = np.zeros((T, S.N), dtype=f)
    done = np.zeros((T, S.N), dtype=np.uint8)
    ent_sum = 0.0
    ep_sum, ep_count = 0.0, 0
    for t in range(T):
        obs_buf[t] = S.obs
        if S.cont:
            noise = R.normals(S.idx, S.gstep, R.STREAM_ACTION, S.seed)[:, :S.num_ac]
        else:
            noise = R.uniforms(S.idx, S.gstep, R.STREAM_ACTION, S.seed)[:, 0]
        ac, lp, ent, _ = P.actor_act(S.actor, S.alay, S.obs.T, noise, S.cont, S.ac_scale, dtype=f)
        ru = R.uniforms(S.idx, S.gstep, R.STREAM_RESET, S.seed).T[:E.STATE_DIM[S.env]]
        env_ac = ac.T if S.cont else ac
        S.state, S.elapsed, raw, rew[t], done[t] = E.step(S.env, S.state, S.elapsed, env_ac, ru, max_steps, f)
        if S.frames == 1:
            S.obs = raw
        else:       # drop the oldest frame, append the newest; a finished episode restarts with k copies (auto-reset)
            shifted = np.concatenate([S.obs[S.env_ob:], raw], axis=0)
            S.obs = np.where(done[t][None, :] != 0, np.tile(raw, (S.frames, 1)), shifted).astype(f)
        S.ep_ret = (S.ep_ret + rew[t]).astype(np.float32)      # tot_rews of A3C/process.py:234, per finished episode
        ended = done[t] != 0
        counted = ended & ~np.isnan(S.ep_ret)
        ep_sum += float(np.sum(S.ep_ret[counted], dtype=np.float64))
        ep_count += int(counted.sum())
        S.ep_ret[ended] = 0.0
        acs[t] = ac.T if S.cont else ac[None, :]
        logps[t] = lp
        ent_sum += float(np.sum(ent, dtype=np.float64))
        S.gstep += 1
    obs_buf[T] = S.obs
    return dict(obs=obs_buf, acs=acs, logps=logps, rew=rew, done=done, ent_sum=ent_sum, ep_sum=ep_sum,
                ep_count=ep_count)


def update(S, ro, gamma=0.98, look_ahead=40, desired_kl=0.002, min_lr=1e-6, max_lr=0.1, clip=0.1,
           allreduce=None, grad_dtype=np.float32, diagnostics=False):
    """One update from a rollout; ``allreduce(x)`` (optional) averages an array over ranks."""
    f = S.f
    T = ro["rew"].shape[0]
    B = T * S.N
    obs_all = ro["obs"].transpose(0, 2, 1).reshape((T + 1) * S.N, S.ob_dim)
    V = P.critic_value(S.critic, S.clay, obs_all, basic_head=S.basic, dtype=f).reshape(T + 1, S.N)
    G, adv = path_adv_columns(ro["rew"], ro["done"], V, gamma, look_ahead, dtype=np.float64)
    advn = P.normalize_adv(adv.reshape(-1)).astype(f)
    obs = obs_all[:B]
    acs = ro["acs"].transpose(0, 2, 1).reshape(B, -1) if S.cont else ro["acs"].reshape(B)
    lp_old = ro["logps"].reshape(B)
    cg = P.critic_loss_grads(S.critic, S.clay, obs, G.reshape(-1).astype(f), basic_head=S.basic, dtype=grad_dtype)
    ag = P.actor_loss_grads(S.actor, S.alay, obs, acs, advn, lp_old, S.beta, S.gamma_ent, S.cont,
                            S.ac_scale, clip=None, dtype=grad_dtype)
    g_c, g_a = cg["grad"].astype(np.float32), ag["grad_raw"].astype(np.float32)
    if allreduce is not None:
        g_c, g_a = allreduce(g_c), allreduce(g_a)
    g_a = np.clip(g_a, -np.float32(clip), np.float32(clip))
    S.t_adam += 1
    S.critic, S.cm, S.cv = P.adam_step(S.critic, g_c, S.cm, S.cv, S.t_adam, S.critic_lr)
    S.actor, S.am, S.av = P.adam_step(S.actor, g_a, S.am, S.av, S.t_adam, S.actor_lr)
    lp_new = P.actor_logp(S.actor, S.alay, obs, acs, S.cont, S.ac_scale, dtype=f)
    kl = float(np.mean((lp_old.astype(np.float64) - lp_new.astype(np.float64)) ** 2))
    ev_before = float(var_accounted_for(target=G.reshape(-1), pred=V[:T].reshape(-1).astype(np.float64)))
    ev_after = float(var_accounted_for(target=G.reshape(-1), pred=P.critic_value(
        S.critic, S.clay, obs, basic_head=S.basic, dtype=f).astype(np.float64))) if diagnostics else None
    if allreduce is not None:
        kl = float(allreduce(np.array([kl]))[0])
    stats = dict(actor_loss=float(ag["loss"]), critic_loss=float(cg["loss"]), kl=kl, actor_lr=S.actor_lr,
                 avg_ent=ro["ent_sum"] / B, mean_rew=float(ro["rew"].mean()), G=G, adv=adv, V=V, rollout=ro,
                 grad_actor=g_a, grad_critic=g_c, ev_before=ev_before, ev_after=ev_after,
                 avg_rew=(ro["ep_sum"] / ro["ep_count"]) if ro.get("ep_count") else None, episodes=ro.get("ep_count", 0))
    S.actor_lr = float(np.float32(P.kl_lr_rule(S.actor_lr, kl, desired_kl, min_lr, max_lr)))
    if S.it % 100 == 0:                                  # A3C/process.py:271-278
        S.gamma_ent = float(10.0 ** linear_schedule_val(S.it, 100, 3000, -2, -8))
        S.beta = float(10.0 ** linear_schedule_val(S.it, 100, 3000, 0, -4))
    S.it += 1
    return stats


def iteration(S, T=20, **kw):
    ro = rollout(S, T, kw.pop("max_steps", None))
    return update(S, ro, **kw)
