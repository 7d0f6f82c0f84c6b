"""The reference-shaped Python interface (policies.Actor/Critic, process.PathAdv/rollout/test_process, envs)
on the GPU: same calls as the reference makes, checked against golden vectors from the reference's own
PathAdv, SURVEY Appendix B (shipped checkpoint), and the oracle."""
import os

import numpy as np
import pytest
import torch

from oracle import envs as E, policies as P
from actor_critic_algs_on_tensorflow_b200 import envs, policies as pol, process, tf_bundle, util as U
from gpu_util import GOLDEN, demo_weights

pytestmark = pytest.mark.gpu
MODEL = os.path.join(GOLDEN, "model-Pendulum_a3c")


def _flat_of(net, olay):
    """facade net -> flat buffer in the oracle layout (identical layouts; checked in test_abi)."""
    return net.params.cpu().numpy()


def test_path_adv_facade_matches_reference_golden():
    z = np.load(os.path.join(GOLDEN, "pathadv.npz"))
    for c in range(int(z["n_cases"])):
        pa = process.PathAdv(gamma=float(z["gamma_%d" % c]), look_ahead=int(z["L_%d" % c]))
        G, adv = pa(rews=list(z["rews_%d" % c]), vals=z["vals_%d" % c], terminal=bool(z["term_%d" % c]))
        scale = max(1.0, np.abs(z["G_%d" % c]).max())
        np.testing.assert_allclose(G, z["G_%d" % c], rtol=1e-5, atol=2e-6 * scale)
        np.testing.assert_allclose(adv, z["adv_%d" % c], rtol=1e-5, atol=4e-6 * scale)
        assert isinstance(G, list) and len(G) == len(z["rews_%d" % c])
    assert process.PathAdv()([], [0.0], True) == ([], [])


def test_actor_act_and_critic_value_on_the_shipped_checkpoint():
    actor = process.load_actor(MODEL, 3, 1, 'cont', np.array([2.0]))
    assert actor.get_opt_param(None) == [np.float32(0.005), np.float32(1.0), np.float32(0.01)]
    ob = np.array([np.cos(0.5), np.sin(0.5), 1.0])
    ac, logp, ent = actor.act(ob, sess=None, eps=np.zeros((1, 1)))          # 1-D observation, row 0 returned
    np.testing.assert_allclose(ac, [-1.982684587], rtol=2e-5)               # Appendix B: mu
    np.testing.assert_allclose(ent, 1.3884116139142493, rtol=1e-6)
    assert ac.shape == (1,) and np.ndim(logp) == 0
    acs, lps, ents = actor.act_batch(np.tile(ob, (5, 1)), eps=np.arange(5.0).reshape(5, 1))
    std = np.exp(-0.03052692)
    np.testing.assert_allclose(acs[:, 0], -1.982684587 + std * np.arange(5.0), rtol=2e-5)
    np.testing.assert_allclose(actor.logp(ob[None], [[0.5]]), [-4.164295090913109], rtol=1e-5)
    np.testing.assert_allclose(actor.get_kl(None, logp_feeds=[-4.164295090913109 + 0.1], obs=ob[None], acs=[[0.5]]), 0.01,
                               rtol=1e-3)
    critic = pol.Critic(name='global_critic', num_ob_feat=3)
    critic.load_checkpoint_tensors(__import__("actor_critic_algs_on_tensorflow_b200.tf_bundle", fromlist=["x"]).read_bundle(MODEL))
    v = critic.value(np.array([[1, 0, 0], [-1, 0, 0], [0, 1, 0]]), sess=None)
    np.testing.assert_allclose(v, [-6.202648083e-03, -248.6037307, -133.2513682], rtol=2e-5, atol=2e-6)
    assert v.shape == (3,) and actor.optimize is None and critic.optimize is None     # no global twin, like the reference


@pytest.mark.parametrize("act_type,ob,nac", [("cont", 3, 1), ("disc", 4, 2)])
def test_local_global_optimize_and_sync(act_type, ob, nac):
    cont = act_type == "cont"
    g_actor = pol.Actor(name='global_actor', num_ob_feat=ob, num_ac=nac, act_type=act_type, seed=1)
    g_critic = pol.Critic(name='global_critic', num_ob_feat=ob, seed=2)
    step = type("GS", (), {"value": 0})()
    l_actor = pol.Actor(name='local_actor_0', num_ob_feat=ob, num_ac=nac, act_type=act_type, global_actor=g_actor,
                        global_step=step, seed=3)
    l_critic = pol.Critic(name='local_critic_0', num_ob_feat=ob, global_critic=g_critic, seed=4)
    alay, clay = P.actor_layout(ob, nac, cont), P.critic_layout(ob)
    rng = np.random.default_rng(0)
    B = 300
    obs = rng.normal(size=(B, ob)).astype(np.float32)
    acs = rng.normal(size=(B, nac)).astype(np.float32) if cont else rng.integers(0, nac, B)
    advs, tg = rng.normal(size=B), rng.normal(size=B)
    la0, ga0 = _flat_of(l_actor.net, alay), _flat_of(g_actor.net, alay)
    lc0, gc0 = _flat_of(l_critic.net, clay), _flat_of(g_critic.net, clay)
    lp_old = P.actor_logp(la0, alay, obs, acs, cont) + 0.05
    loss_c, none = l_critic.optimize(obs=obs, targets=tg, sess=None)
    loss_a, _ = l_actor.optimize(acs=acs, obs=obs, advs=advs, logps=lp_old, sess=None)
    assert none is None and step.value == 1
    # gradients of the LOCAL loss, applied to the GLOBAL variables with the local Adam moments (policies.py:88-89)
    wa = P.actor_loss_grads(la0, alay, obs, acs, advs, lp_old, 1.0, 0.01, cont, 2.0, clip=0.1)
    wc = P.critic_loss_grads(lc0, clay, obs, tg)
    np.testing.assert_allclose(loss_a, wa["loss"], rtol=2e-5, atol=1e-6)
    np.testing.assert_allclose(loss_c, wc["loss"], rtol=2e-5)
    ga1, _, _ = P.adam_step(ga0, wa["grad"].astype(np.float32), 0 * ga0, 0 * ga0, 1, 0.005)
    gc1, _, _ = P.adam_step(gc0, wc["grad"].astype(np.float32), 0 * gc0, 0 * gc0, 1, 0.001)
    for got, want, before in ((_flat_of(g_actor.net, alay), ga1, ga0), (_flat_of(g_critic.net, clay), gc1, gc0)):
        assert np.mean(np.abs(got - want) > 1e-5) < 0.01          # lr*sign(g) everywhere but at rounding-noise gradients
        assert np.abs(got - before).max() > 1e-4
    np.testing.assert_array_equal(_flat_of(l_actor.net, alay), la0)           # local weights untouched until the pull
    l_actor.sync_w_global(None)
    l_critic.sync_w_global(None)
    np.testing.assert_array_equal(_flat_of(l_actor.net, alay), _flat_of(g_actor.net, alay))
    np.testing.assert_array_equal(_flat_of(l_critic.net, clay), _flat_of(g_critic.net, clay))
    lr, beta, gamma = l_actor.set_opt_param(sess=None, new_lr=0.0075, new_gamma=1e-3)
    assert (lr, beta, gamma) == (np.float32(0.0075), np.float32(1.0), np.float32(1e-3))
    assert l_critic.set_opt_param(0.5, None) == 0.5                           # reference quirk: no effect


def test_reference_shaped_rollout_and_gym_like_env():
    env, max_len, stop = process.get_roll_params("Pendulum-v0")
    assert (max_len, stop) == (200, 1200) and env.action_space.shape == (1,) and env.observation_space.shape == (3,)
    env2, max_len2, stop2 = process.get_roll_params("Pendulum-v0", basic=True)
    assert (max_len2, stop2) == (400, 1400)
    actor = process.load_actor(MODEL, 3, 1, 'cont', 2.0)
    path = process.rollout(env=env, sess=None, policy=actor.act, framer=process.Framer(1), max_path_length=max_len)
    assert len(path['rews']) == 200 and len(path['obs']) == 201 and path['terminated'] is True   # TimeLimit
    assert -2000 < sum(path['rews']) < 0 and len(path['acs']) == 200 and len(path['logps']) == 200
    short = process.rollout(env=env, sess=None, policy=actor.act, framer=process.Framer(1), max_path_length=7)
    assert len(short['rews']) == 7 and short['terminated'] is False                                # length cut
    cp = envs.make("CartPole-v0")
    assert isinstance(cp.action_space, envs.Discrete) and cp.action_space.n == 2 and cp.spec.max_episode_steps == 200
    ob = cp.reset()
    assert ob.shape == (4,) and np.abs(ob).max() <= 0.05
    n = 0
    done = False
    while not done:
        ob, r, done, _ = cp.step(1)         # always push right: falls over within a few dozen steps
        n += 1
        assert r == 1.0
    assert 5 < n < 60
    with pytest.raises(Exception, match="unsupported env"):
        envs.make("MountainCar-v0")


def test_test_process_and_batched_greedy_evaluation(capsys):
    avg = process.test_process("Pendulum-v0", 12321, 1, MODEL, num_episodes=2, animate=False)
    out = capsys.readouterr().out
    assert "Iteration 1" in out and "Episode Length 200" in out and "Average reward over 2 was" in out
    assert -1500 < avg < 0
    # greedy, batched (BASELINE configs[4]): against the oracle's float64 greedy episodes from the same starts
    n = 4096
    ret = process.evaluate_batched("Pendulum-v0", MODEL, n, greedy=True, seed=4)
    from oracle import philox as R
    u = R.uniforms(np.arange(n, dtype=np.uint64), 0xFFFFFFFF, R.STREAM_RESET, 4).T[:2]
    s = E.reset_from_uniforms(E.PENDULUM, u.astype(np.float64))
    actor, _ = demo_weights()
    lay = P.actor_layout(3, 1, True)
    want = np.zeros(n)
    for t in range(200):
        _, _, _, mu = P.actor_act(actor, lay, E.observe(E.PENDULUM, s).T, np.zeros((n, 1)), True, 2.0, dtype=np.float64)
        s, r, _ = E.dynamics(E.PENDULUM, s, mu.T)
        want += r
    rel = np.abs(ret - want) / np.maximum(np.abs(want), 1.0)
    assert np.median(rel) < 1e-4 and np.quantile(rel, 0.95) < 1e-2          # fp32 drift on the chaotic few
    np.testing.assert_allclose(ret.mean(), want.mean(), rtol=2e-3)
    assert -190 < ret.mean() < -150                                          # Appendix B: -171.1 +- 107.4 over random starts
    sampled = process.evaluate_batched("Pendulum-v0", MODEL, n, greedy=False, seed=4)
    assert sampled.mean() < ret.mean() and -230 < sampled.mean() < -170      # Appendix B: -200.9 sampled
    # the same evaluation on the persistent rollout kernel (3xTF32 forward, 20 steps per launch): same bar
    ret_tc = process.evaluate_batched("Pendulum-v0", MODEL, n, greedy=True, seed=4, tensor_cores=True)
    rel = np.abs(ret_tc - want) / np.maximum(np.abs(want), 1.0)
    assert np.median(rel) < 1e-4 and np.quantile(rel, 0.95) < 1e-2
    np.testing.assert_allclose(ret_tc.mean(), want.mean(), rtol=2e-3)
    sampled_tc = process.evaluate_batched("Pendulum-v0", MODEL, n, greedy=False, seed=4, tensor_cores=True)
    np.testing.assert_allclose(sampled_tc.mean(), sampled.mean(), rtol=2e-2)   # same Philox draws, chaotic episodes


def test_batched_evaluation_with_early_termination_agrees_between_engines():
    """CartPole episodes end early: the return is the reward sum up to the first done.  Random-init policy, sampled
    actions from the same Philox counters on both engines: the per-step launches (FFMA forward) and the persistent
    rollout kernel must agree episode by episode except where a 3xTF32-sized logit difference flips a draw."""
    from actor_critic_algs_on_tensorflow_b200.trainer import glorot_flat
    n = 8192
    flat = glorot_flat(0, 4, 2, False, np.random.default_rng(3))
    a = process.evaluate_batched("CartPole-v0", None, n, greedy=False, seed=9, actor_flat=flat)
    b = process.evaluate_batched("CartPole-v0", None, n, greedy=False, seed=9, actor_flat=flat, tensor_cores=True)
    assert a.min() >= 8 and a.max() <= 200 and 15 < a.mean() < 40          # random policy: ~22 steps
    assert (a == b).mean() > 0.995
    np.testing.assert_allclose(b.mean(), a.mean(), rtol=5e-3)


def test_process_fn_and_run_ac_smoke(tmp_path, capsys):
    from actor_critic_algs_on_tensorflow_b200 import run_AC, train
    os.chdir(str(tmp_path))
    assert train.main(["ps", "0", "--worker_num", "1", "--ps_num", "1"]) is None            # no ps on one box
    tr = train.main(["worker", "0", "--worker_num", "1", "--ps_num", "1", "--env", "CartPole-v0", "--n_envs", "256",
                     "--horizon", "8", "--max_iters", "4", "--save_every", "2", "--stdout_freq", "2",
                     "--checkpoint_dir", "ck", "--outdir", "logs"])
    assert tr.iterations == 4
    files = sorted(os.listdir("ck"))
    assert "checkpoint" in files and any(f.startswith("model-CartPole-") and f.endswith(".index") for f in files)
    log = open(os.path.join("logs", "worker_0.log")).read().splitlines()
    assert log[0] == "step avg_rew ev_before ev_after act_loss crit_loss kl_dist avg_ent" and len(log) == 5    # one row per iteration
    assert "Iteration 0" in capsys.readouterr().out
    tr = run_AC.main(["--env", "CartPole-v0", "--iters", "3", "--outdir", "log.txt", "--checkpoint_dir",
                      os.path.join("tmp", "checkpoints"), "--save_every", "1", "--n_envs", "64", "--horizon", "100"])
    out = capsys.readouterr().out
    assert "Discrete Action Space" in out and "Iteration 0" in out and "EV After" in out
    assert any(f.startswith("checkpoints-CartPole-") and f.endswith(".index") for f in os.listdir("tmp"))
    assert tr.basic and tr.iterations == 3 and np.isfinite(tr.params.cpu().numpy()).all()
    t = tf_bundle.read_bundle(os.path.join("tmp", sorted(f for f in os.listdir("tmp") if f.endswith(".index"))[0][:-6]))
    assert "Actor/logits/kernel" in t and "Critic/value/kernel" in t and t["Critic/third_layer/kernel"].shape == (128, 128)
    assert len(open("log.txt").read().splitlines()) == 4
