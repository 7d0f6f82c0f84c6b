"""Training / evaluation drivers with the reference's entry points (``A3C/process.py``):
``Framer`` (:14-43), ``PathAdv`` (:46-71), ``rollout`` (:73-98), ``get_roll_params`` (:100-109),
``train_ciritic`` / ``train_actor`` (:111-123), ``test_process`` (:125-153), ``process_fn`` (:156-288).

Two shapes of the same algorithm live here:

* the **reference-shaped** path -- one env, batch-1 ``act`` per step, per-episode ``PathAdv`` -- runs the
  reference's own call sequence on top of the CUDA-backed ``Actor`` / ``Critic`` / ``GymEnv`` objects
  (used by ``test_process`` and by ``Basic_AC``-style training; every arithmetic step is a kernel call);
* the **fused** path -- ``process_fn`` drives ``trainer.A2CTrainer``: N envs x T steps per update, CUDA
  graphs, NCCL all-reduce across ranks.  ``cluster`` / ``job`` / ``task_id`` keep their meaning as far as
  one NVSwitch box needs: ``job == 'ps'`` returns at once (parameters are replicated, there is no
  parameter-server process), ``task_id`` is the rank, ``len(cluster['worker'])`` the world size.
"""
import os

import numpy as np
import torch

from . import _lib as L
from . import envs as envs_mod
from . import policies as pol
from . import util as U
from .tf_bundle import read_bundle, write_bundle, write_checkpoint_state

LOG_ROUND = 10
MAX_ITERS = 1e7
MAX_LR, MIN_LR = .1, 1e-6


class Framer(object):
    """Builds network inputs by concatenating the ``frame_num`` most recent observations (the first one is
    repeated as left padding).  ``last`` gives the feature for the newest step, ``full`` for every step."""

    def __init__(self, frame_num):
        self.frame_num = frame_num

    def _extend(self, obs):
        obs = list(obs)
        return [obs[0]] * (self.frame_num - 1) + obs

    def last(self, obs):
        ext = self._extend(obs)
        return np.concatenate(ext[len(ext) - self.frame_num:])

    def full(self, obs):
        ext = self._extend(obs)
        k = self.frame_num
        return [np.concatenate(ext[i:i + k]) for i in range(len(ext) - k + 1)]


class PathAdv(object):
    """Windowed discounted returns and advantages of one episode, computed by the scan kernel (K3).
    ``__call__(rews, vals, terminal) -> (targets, advs)`` as lists, like the reference."""

    def __init__(self, gamma=0.98, look_ahead=30):
        self.reset(gamma, look_ahead)

    def reset(self, gamma, look_ahead):
        self.gamma, self.look_ahead = gamma, look_ahead
        self.kern = [np.power(gamma, k) for k in range(look_ahead)]

    def __call__(self, rews, vals, terminal):
        n = len(rews)
        assert len(vals) == n + 1
        if n == 0:
            return [], []
        dev = torch.device("cuda", torch.cuda.current_device())
        r = torch.as_tensor(np.asarray(rews, dtype=np.float32).reshape(n, 1), device=dev)
        v = torch.as_tensor(np.asarray(vals, dtype=np.float32).reshape(n + 1, 1), device=dev)
        d = torch.zeros((n, 1), device=dev, dtype=torch.uint8)
        if terminal:
            d[n - 1, 0] = 1
        G, A = torch.empty((n, 1), device=dev), torch.empty((n, 1), device=dev)
        stats = torch.zeros(L.STATS_SIZE, device=dev)
        nbytes = L.load().acx_returns_adv_workspace_bytes(n, 1)
        ws = torch.zeros(nbytes, device=dev, dtype=torch.uint8)
        L.call("acx_returns_adv_scan", n, 1, L.ptr(r), L.ptr(d), L.ptr(v), float(self.gamma), int(self.look_ahead),
               L.ptr(G), L.ptr(A), L.ptr(stats), L.ptr(ws), nbytes, L.SCAN_AUTO, L.cur_stream())
        return list(G[:, 0].cpu().numpy().astype(np.float64)), list(A[:, 0].cpu().numpy().astype(np.float64))


def rollout(env, sess, policy, framer, max_path_length=100, render=False):
    """One episode: reset, then act -> step until ``done`` or ``max_path_length`` (reference loop shape)."""
    t = 0
    ob = env.reset()
    obs, logps, rews, acs = [ob], [], [], []
    sum_ents = 0
    done = False
    while t < max_path_length and not done:
        if render:
            env.render()
        t += 1
        ac, logp, ent = policy(framer.last(obs), sess=sess)
        ob, rew, done, _ = env.step(ac)
        obs.append(ob)
        rews.append(rew)
        acs.append(ac)
        sum_ents += ent
        logps.append(logp)
    return {'rews': rews, 'obs': obs, 'acs': acs, 'terminated': done, 'logps': logps, 'entropy': sum_ents}


def get_roll_params(env_id, basic=False):
    """Environment + rollout sizes (A3C/process.py:100-109; basic=True: Basic_AC/run_AC.py:124-137)."""
    env = envs_mod.make(env_id)
    max_path_length, ep_length_stop = 1200, 3000
    if basic and env_id == "Pendulum-v0":
        max_path_length, ep_length_stop = 400, 1400
    elif env.spec.max_episode_steps is not None:
        max_path_length = env.spec.max_episode_steps
        ep_length_stop = min(max_path_length * (4 if basic else 6), 3000)
    return env, max_path_length, ep_length_stop


def train_ciritic(critic, sess, obs, targets):
    assert len(obs) == len(targets)
    pre_preds = critic.value(obs, sess=sess)
    ev_before = U.var_accounted_for(pred=pre_preds, target=targets)
    loss, _ = critic.optimize(obs=obs, targets=targets, sess=sess)
    return loss, ev_before


def train_actor(actor, sess, obs, advs, logps, acs, rolls=None):
    assert len(obs) == len(advs)
    assert len(advs) == len(acs)
    loss, _ = actor.optimize(sess=sess, obs=obs, acs=acs, advs=advs, logps=logps)
    return loss


def _space_params(env):
    if isinstance(env.action_space, envs_mod.Discrete):
        return 'disc', env.action_space.n, None
    return 'cont', env.action_space.shape[0], np.maximum(env.action_space.high, np.abs(env.action_space.low))


def load_actor(model_path, ob_dim, ac_dim, act_type, ac_scale, scope='global_actor'):
    """``tf.train.Saver().restore`` for the actor variables of a TF V2 checkpoint bundle."""
    actor = pol.Actor(name=scope, num_ob_feat=ob_dim, num_ac=ac_dim, act_type=act_type, ac_scale=ac_scale)
    actor.load_checkpoint_tensors(read_bundle(model_path), scope)
    return actor


def test_process(env_id, random_seed, stack_frames, model_path, num_episodes, animate=True):
    """Restore ``global_actor/*`` from ``model_path`` and roll out ``num_episodes`` episodes with SAMPLED actions,
    printing what the reference prints (A3C/process.py:125-153; ``random_seed`` is unused there too)."""
    env, MAX_PATH_LENGTH, _ = get_roll_params(env_id)
    framer = Framer(frame_num=stack_frames)
    ob_dim = env.observation_space.shape[0] * stack_frames
    act_type, ac_dim, ac_scale = _space_params(env)
    actor = load_actor(model_path, ob_dim, ac_dim, act_type, ac_scale)
    avg_rew = 0
    for i in range(num_episodes):
        path = rollout(env=env, sess=None, policy=actor.act, max_path_length=MAX_PATH_LENGTH, framer=framer,
                       render=animate)
        rew = np.sum(path['rews'])
        print("Iteration {}".format(i))
        print("Reward {}".format(rew))
        print("Episode Length {}\n".format(len(path['rews'])))
        avg_rew += rew / float(num_episodes)
    print('Average reward over {} was {}'.format(num_episodes, avg_rew))
    return avg_rew


def _evaluate_chunk_rollout(vec, params, num_ac, cont, greedy, seed, start, block=20):
    """One chunk of evaluate_batched on the persistent rollout kernel (acx_rollout_tc): ``block`` steps of
    {policy forward on the tensor cores, action, env transition} per launch, env state in registers; the first
    episode's return is the reward sum up to the first ``done`` of each column."""
    n, dev, s = vec.n, vec.dev, L.cur_stream()
    ob, T, horizon = vec.obs_dim, int(block), vec.max_episode_steps
    packed = torch.zeros(L.load().acx_tc_packed_floats(0, ob, num_ac, int(cont)), device=dev)
    L.call("acx_tc_pack_weights", 0, L.ptr(params), ob, num_ac, int(cont), L.ptr(packed), s)
    obs = torch.empty((T + 1, ob, n), device=dev)
    obs[0].copy_(vec.reset())
    noise = torch.zeros((T, 1, n), device=dev) if (greedy and cont) else None       # eps = 0: action = mu
    acs = torch.empty((T, 1, n), device=dev) if cont else torch.empty((T, n), device=dev, dtype=torch.int32)
    rew = torch.empty((T, n), device=dev)
    done = torch.empty((T, n), device=dev, dtype=torch.uint8)
    ctr = torch.zeros(1, device=dev, dtype=torch.int32)
    ret = torch.zeros(n, device=dev)
    alive = torch.ones(n, device=dev)
    for t0 in range(0, horizon, T):
        tc = min(T, horizon - t0)
        L.call("acx_rollout_tc", vec.env, L.ptr(params), L.ptr(packed), 2.0, tc, n, L.ptr(obs), L.ptr(vec.state),
               L.ptr(vec.elapsed), L.ptr(noise), None, seed, start, L.ptr(ctr), 0, L.ptr(acs), None, None, L.ptr(rew),
               L.ptr(done), None, None, None, s)
        L.call("acx_counter_add", L.ptr(ctr), tc, s)
        d = done[:tc].to(torch.float32)
        before = torch.cumsum(d, 0) - d                     # finished episodes before step t of this block
        ret += (rew[:tc] * (before == 0)).sum(0) * alive
        alive = alive * (d.sum(0) == 0)
        obs[0].copy_(obs[tc])
    return ret


def evaluate_batched(env_id, model_path, num_episodes, greedy=True, seed=12321, actor_flat=None, chunk=1 << 20,
                     tensor_cores=False):
    """Config 5: evaluate a checkpoint over ``num_episodes`` episodes stepped in lock-step on the GPU.
    ``greedy=True`` uses action = mu (the derived deterministic variant); ``False`` samples like the
    reference's ``test_process``.  Returns the per-episode returns (numpy, first episode of each env).
    ``tensor_cores=True`` runs each chunk on the persistent rollout kernel (3xTF32 forward, 20 steps per launch)
    instead of one FFMA forward + one env-step launch per step."""
    env_code = envs_mod.ENV_IDS[env_id]
    dev = torch.device("cuda", torch.cuda.current_device())
    out = []
    for start in range(0, num_episodes, chunk):
        n = min(chunk, num_episodes - start)
        vec = envs_mod.VecEnv(env_id, n, seed, index0=start)
        cont = vec.n_actions == 0
        num_ac = 1 if cont else vec.n_actions
        if actor_flat is None:
            actor = load_actor(model_path, vec.obs_dim, num_ac, 'cont' if cont else 'disc', 2.0)
            params = actor.net.params
        else:
            params = torch.as_tensor(np.asarray(actor_flat, dtype=np.float32), device=dev)
        if tensor_cores:
            out.append(_evaluate_chunk_rollout(vec, params, num_ac, cont, greedy, seed, start).cpu().numpy())
            continue
        obs = vec.reset().clone()
        ret = torch.zeros(n, device=dev)
        alive = torch.ones(n, device=dev)
        ac_f = torch.empty((num_ac, n), device=dev)
        ac_i = torch.empty(n, device=dev, dtype=torch.int32)
        lp = torch.empty(n, device=dev)
        zeros = torch.zeros((num_ac, n), device=dev)
        s = L.cur_stream()
        for t in range(vec.max_episode_steps):
            if cont:
                L.call("acx_policy_forward_sample_gauss", L.ptr(params), vec.obs_dim, num_ac, 2.0, 1, n, L.ptr(obs),
                       L.ptr(zeros) if greedy else None, seed, start, t, None, L.ptr(ac_f), L.ptr(lp), None, None, s)
                o, r, d = vec.step(ac_f)
            else:
                L.call("acx_policy_forward_sample_cat", L.ptr(params), vec.obs_dim, num_ac, 1, n, L.ptr(obs), None, seed,
                       start, t, None, L.ptr(ac_i), L.ptr(lp), None, None, s)
                o, r, d = vec.step(ac_i)
            ret += r * alive
            alive = alive * (1.0 - d.to(torch.float32))
            obs.copy_(o)
        out.append(ret.cpu().numpy())
    return np.concatenate(out)


_KEPT = {}      # Saver(max_to_keep) ring per checkpoint directory


def save_checkpoint(save_path, basename, gstep, actor_tensors, critic_tensors, keep=3):
    """What the chief's CheckpointSaverHook + Saver(max_to_keep=3) leave on disk (A3C/process.py:197,211-214):
    ``<save_path>/<basename>-<gstep>.{index,data-00000-of-00001}`` holding the 22 global tensors, plus the
    ``checkpoint`` state file.  The max_to_keep ring is kept per directory."""
    os.makedirs(save_path, exist_ok=True)
    prefix = os.path.join(save_path, "%s-%d" % (basename, gstep))
    t = dict(actor_tensors)
    t.update(critic_tensors)
    write_bundle(prefix, t)
    _kept = _KEPT.setdefault(os.path.abspath(save_path), [])
    if prefix in _kept:
        _kept.remove(prefix)
    _kept.append(prefix)
    while len(_kept) > keep:
        old = _kept.pop(0)
        for ext in (".index", ".data-00000-of-00001"):
            if os.path.exists(old + ext):
                os.remove(old + ext)
    write_checkpoint_state(save_path, os.path.basename(prefix), [os.path.basename(p) for p in _kept])
    return prefix


def trainer_checkpoint_tensors(tr, actor_scope="global_actor", critic_scope="global_critic"):
    """The 22 tensors of the reference's Saver, named as in SURVEY.md Appendix A, from a trainer's flat buffers
    (Basic_AC's variable scopes are ``Actor`` / ``Critic``, Basic_AC/policies.py:40,129)."""
    head = "mu_layer" if tr.cont else "logits"
    a = tr.actor.cpu().numpy()
    c = tr.critic.cpu().numpy()
    la, _ = L.params_layout(0, tr.ob, tr.num_ac, tr.cont)
    lc, _ = L.params_layout(1, tr.ob)
    op = tr.opt.cpu().numpy()
    ha, hc = (128, 128, 64), (256, 128, 128)

    def pick(flat, lay, name, shape):
        off, sz = lay[name]
        return flat[off:off + sz].reshape(shape).copy()
    ga, gc = actor_scope + "/", critic_scope + "/"
    at = {ga + "Variable": np.float32(op[L.OPT_ACTOR_LR]), ga + "Variable_1": np.float32(op[L.OPT_GAMMA_ENT]),
          ga + "Variable_2": np.float32(op[L.OPT_BETA_KL])}
    dims = [tr.ob] + list(ha) + [tr.num_ac]
    for i, (layer, w, b) in enumerate((("first_layer", "W1", "b1"), ("second_layer", "W2", "b2"),
                                       ("third_layer", "W3", "b3"), (head, "Wh", "bh"))):
        at[ga + "%s/kernel" % layer] = pick(a, la, w, (dims[i], dims[i + 1]))
        at[ga + "%s/bias" % layer] = pick(a, la, b, (dims[i + 1],))
    if tr.cont:
        at[ga + "log_std"] = pick(a, la, "log_std", (tr.num_ac,))
    ct = {gc + "Variable": np.float32(op[L.OPT_CRITIC_LR]), gc + "Variable_1": np.float32(op[L.OPT_CRITIC_LR])}
    dims = [tr.ob] + list(hc) + [1]
    for i, (layer, w, b) in enumerate((("first_layer", "W1", "b1"), ("second_layer", "W2", "b2"),
                                       ("third_layer", "W3", "b3"), ("value", "Wh", "bh"))):
        ct[gc + "%s/kernel" % layer] = pick(c, lc, w, (dims[i], dims[i + 1]))
        ct[gc + "%s/bias" % layer] = pick(c, lc, b, (dims[i + 1],))
    return at, ct


def _emit_log_rows(tr, logger, task_id, first, count, stdout_freq, rows=None):
    """One logger row per iteration (A3C/process.py:280-281) from the trainer's device ring: a single read-back for
    ``count`` iterations.  ``task_id`` None: the Basic_AC row, which has no worker_id (Basic_AC/run_AC.py:277-278)."""
    if logger is None or count <= 0:
        return
    extra = {} if task_id is None else {"worker_id": task_id}
    for j, row in enumerate(rows if rows is not None else tr.read_log(first, count), start=first):
        logger(j, act_loss=row['act_loss'], act_lr=row['act_lr'], kl_dist=row['kl_dist'],
               circ_loss=np.sqrt(max(row['crit_loss'], 0.0)), avg_rew=row['avg_rew'], ev_before=row['ev_before'],
               ev_after=row['ev_after'], print_tog=(j % max(1, stdout_freq)) == 0, avg_ent=row['avg_ent'], **extra)
        if j % 100 == 50:
            logger.flush()


def process_fn(cluster, task_id, job, env_id, logger, save_path, stdout_freq, random_seed=12321, gamma=0.98,
               look_ahead=40, stack_frames=3, animate=False, save_every=600, desired_kl=0.002, TB_log=False,
               run_mode='train', checkpoint_basename='model', n_envs=4096, horizon=20, max_iters=None,
               mode='sync', use_graph=True, diagnostics=True):
    """Worker body on the fused GPU path (reference: A3C/process.py:156-288).  One call = one rank.
    Extra keywords (``n_envs``, ``horizon``, ``max_iters``, ``mode``) size the batched rollout; everything
    else keeps the reference's meaning.  ``random_seed`` arrives as seed + task (A3C/train.py:40).
    Every iteration leaves one log row (losses, lr, KL, avg_rew = mean return of the episodes that ended in the
    batch, ev_before, ev_after, entropy) in a device ring; the host reads the ring back once per ``stdout_freq``
    iterations (at most every 128) and feeds the rows to ``logger`` in order, so the log file has the reference's one
    row per iteration without a host synchronisation per iteration.  ``diagnostics=False`` drops ev_after (one extra
    critic pass per update; logged as -1)."""
    from .trainer import A2CTrainer
    if job == 'ps':
        print("ps %d: parameters are replicated on every GPU and reduced over NCCL; nothing to serve" % task_id)
        return None
    world = len(cluster['worker'])
    is_chief = (task_id == 0)
    if world > 1:
        import torch.distributed as dist
        if not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", str(29500 + int(cluster['worker'][0].split(':')[-1]) % 1000))
            torch.cuda.set_device(task_id % torch.cuda.device_count())
            dist.init_process_group("nccl", rank=task_id, world_size=world)
    tr = A2CTrainer(env_id, n_envs=n_envs, horizon=horizon, seed=random_seed - task_id, gamma=gamma,
                    look_ahead=look_ahead, desired_kl=desired_kl, min_lr=MIN_LR, max_lr=MAX_LR, rank=task_id,
                    world=world, mode=mode, use_graph=use_graph, frames=stack_frames, diagnostics=diagnostics)
    if is_chief:
        print('Initilizing chief. Envirnoment action type {}.'.format('cont' if tr.cont else 'disc'))
    print('\n\nREACHING THE MAIN LOOP OF WORKER %d\n' % task_id)
    i, gstep, last_saved, logged = 0, 0, -1, 0
    limit = MAX_ITERS if max_iters is None else max_iters
    window = max(1, min(int(stdout_freq) if stdout_freq else 1, tr.log_rows))
    while gstep < limit:
        tr.iteration()
        gstep += world            # the reference bumps global_step once per worker update (policies.py:89)
        i += 1
        if i - logged >= window or gstep >= limit:
            _emit_log_rows(tr, logger, task_id, logged, i - logged, stdout_freq)      # the only host sync of the loop
            logged = i
        if is_chief and save_path and gstep // save_every > last_saved:
            last_saved = gstep // save_every
            at, ct = trainer_checkpoint_tensors(tr)
            save_checkpoint(save_path, checkpoint_basename, gstep, at, ct)
    if logger is not None:
        logger.flush()
    return tr
