"""The reference's own loop SHAPE on the host (B-ref of BASELINE.md section 3) -- TEST INFRASTRUCTURE / timed CPU baseline.

``oracle/a2c.py`` batches N environments (the shape the GPU path runs, and faster on a CPU than the reference ever
was).  This module keeps the reference's shape instead: ONE environment, a batch-1 ``act`` per step, whole episodes,
``PathAdv`` per episode, one update once enough steps are collected --

  * A3C worker  (``A3C/process.py:217-285``): episodes until ``len(ep_rews) >= EP_LENGTH_STOP`` (:221; Pendulum and
    CartPole: min(200 * 6, 3000) = 1,200 steps, :105-109), then critic + actor update, KL, lr rule;
  * Basic_AC    (``Basic_AC/run_AC.py:208-284``): ``... and rolls < 7`` (:212), EP_LENGTH_STOP = min(200 * 4, 3000) = 800
    for CartPole (:129-134), value head on the second layer (``Basic_AC/policies.py:134-135``), clip +-1 (:81).

TensorFlow's per-call dispatch cost (one ``sess.run`` per env step, the reference's dominant cost) is NOT modelled: the
numbers this gives are an upper bound on the real reference.  ``run_workers`` runs k independent workers in k
processes -- ``A3C/runner.sh:3-4`` starts 1 ps + 3 workers, the README's example 1 ps + 2 (the ps does no arithmetic
worth timing): both are timed.
"""
import time

import numpy as np

from . import envs as E
from . import policies as P
from .pathadv import path_adv_episode


class RefWorker:
    def __init__(self, env_id="Pendulum-v0", seed=12321, basic=False, gamma=0.98, look_ahead=40):
        self.env = E.ENV_IDS[env_id]
        self.basic = bool(basic)
        self.cont = self.env == E.PENDULUM
        self.ob = E.OBS_DIM[self.env]
        self.num_ac = 1 if self.cont else E.N_ACTIONS[self.env]
        self.alay, self.clay = P.actor_layout(self.ob, self.num_ac, self.cont), P.critic_layout(self.ob)
        self.rng = np.random.default_rng(seed)
        self.actor, self.critic = P.glorot_init(self.alay, self.rng), P.glorot_init(self.clay, self.rng)
        self.am, self.av = np.zeros_like(self.actor), np.zeros_like(self.actor)
        self.cm, self.cv = np.zeros_like(self.critic), np.zeros_like(self.critic)
        self.t = 0
        self.lr, self.beta, self.gamma_ent = 0.005, 1.0, 0.01
        self.gamma, self.look_ahead = gamma, look_ahead
        limit = E.MAX_EPISODE_STEPS[self.env]
        if basic and self.cont:
            self.max_path, self.stop = 400, 1400                      # Basic_AC/run_AC.py:130-131
        else:
            self.max_path, self.stop = limit, min(limit * (4 if basic else 6), 3000)
        self.limit = limit
        self.max_rolls = 7 if basic else 1 << 30
        self.clip, self.max_lr = (1.0, 1.0) if basic else (0.1, 0.1)

    def rollout(self):
        """A3C/process.py:73-98: one episode, batch-1 act per step, gym float64 dynamics."""
        s = E.reset_from_uniforms(self.env, self.rng.random((E.STATE_DIM[self.env], 1)), np.float64)
        obs, acs, lps, rews, ent = [E.observe(self.env, s, np.float64)[:, 0]], [], [], [], 0.0
        done, t = False, 0
        while t < self.max_path and not done:
            t += 1
            noise = self.rng.standard_normal((1, 1)) if self.cont else self.rng.random(1)
            ac, lp, en, _ = P.actor_act(self.actor, self.alay, obs[-1][None, :].astype(np.float32), noise, self.cont, 2.0)
            s, r, term = E.dynamics(self.env, s, ac.T if self.cont else ac, np.float64)
            done = bool(term[0]) or t >= self.limit                  # gym's TimeLimit wrapper
            obs.append(E.observe(self.env, s, np.float64)[:, 0])
            acs.append(ac[0]); lps.append(float(lp[0])); rews.append(float(r[0])); ent += float(np.sum(en))
        return dict(obs=obs, acs=acs, logps=lps, rews=rews, terminated=done, entropy=ent)

    def update(self):
        """One outer-loop body; returns the number of env steps it consumed."""
        f = np.float32
        ep_obs, ep_adv, ep_lp, ep_tgt, ep_ac, n, rolls = [], [], [], [], [], 0, 0
        while n < self.stop and rolls < self.max_rolls:
            path = self.rollout()
            o = np.asarray(path["obs"], dtype=f)
            vals = P.critic_value(self.critic, self.clay, o, basic_head=self.basic)
            tgt, adv = path_adv_episode(path["rews"], vals, path["terminated"], self.gamma, self.look_ahead)
            ep_obs.append(o[:-1]); ep_adv.append(adv); ep_tgt.append(tgt); ep_lp += path["logps"]; ep_ac += path["acs"]
            n += len(path["rews"]); rolls += 1
        obs = np.concatenate(ep_obs)
        adv = P.normalize_adv(np.concatenate(ep_adv)).astype(f)
        tgt = np.concatenate(ep_tgt).astype(f)
        lp_old = np.asarray(ep_lp, dtype=f)
        acs = np.asarray(ep_ac, dtype=f).reshape(n, -1) if self.cont else np.asarray(ep_ac).reshape(n)
        cg = P.critic_loss_grads(self.critic, self.clay, obs, tgt, basic_head=self.basic, dtype=f)
        ag = P.actor_loss_grads(self.actor, self.alay, obs, acs, adv, lp_old, self.beta, self.gamma_ent, self.cont, 2.0,
                                clip=self.clip, dtype=f)
        self.t += 1
        self.critic, self.cm, self.cv = P.adam_step(self.critic, cg["grad"], self.cm, self.cv, self.t, 0.001)
        self.actor, self.am, self.av = P.adam_step(self.actor, ag["grad"], self.am, self.av, self.t, self.lr)
        lp_new = P.actor_logp(self.actor, self.alay, obs, acs, self.cont, 2.0)
        kl = float(np.mean((lp_old.astype(np.float64) - lp_new) ** 2))
        self.lr = P.kl_lr_rule(self.lr, kl, 0.002, 1e-6, self.max_lr)
        return n


def _timed(args):
    env_id, seed, basic, seconds = args
    w = RefWorker(env_id, seed, basic)
    w.update()                                           # warm-up (BLAS threads, allocator)
    steps, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < seconds:
        steps += w.update()
    return steps, time.perf_counter() - t0


def run_workers(env_id, workers=1, basic=False, seconds=4.0, seed=12321):
    """env-steps/s of ``workers`` independent reference-shaped workers, one process each (seed + task, A3C/train.py:40;
    ``runner.sh`` starts them as separate interpreters, and so does this)."""
    jobs = [(env_id, seed + k, basic, seconds) for k in range(workers)]
    if workers == 1:
        res = [_timed(jobs[0])]
    else:
        import json
        import os
        import subprocess
        import sys
        root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
        code = "import json, sys; from oracle import ref_loop as R; print(json.dumps(R._timed(tuple(json.loads(sys.argv[1])))))"
        env = dict(os.environ)                         # the host's cores are shared between the workers
        for k in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
            env[k] = str(max(1, (os.cpu_count() or 1) // workers))
        procs = [subprocess.Popen([sys.executable, "-c", code, json.dumps(j)], cwd=root, stdout=subprocess.PIPE, text=True,
                                  env=env) for j in jobs]
        res = [tuple(json.loads(p.communicate(timeout=600)[0].strip().splitlines()[-1])) for p in procs]
    return sum(s / dt for s, dt in res), sum(s for s, _ in res)
