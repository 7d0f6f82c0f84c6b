"""CPU tests: the oracle against (a) golden vectors produced by the REFERENCE'S OWN classes
(tests/golden/make_golden.py), (b) SURVEY.md Appendix-B known answers on the reference's shipped
checkpoint, (c) torch.autograd for the analytic gradients, (d) its own fp32 vs fp64 modes."""
import os

import numpy as np
import pytest
import torch

from oracle import a2c, envs as E, pathadv as PA, philox as R, policies as P
from gpu_util import demo_weights, random_weights


def test_path_adv_episode_matches_reference_golden(golden_dir):
    z = np.load(os.path.join(golden_dir, "pathadv.npz"))
    for c in range(int(z["n_cases"])):
        G, adv = PA.path_adv_episode(z["rews_%d" % c], z["vals_%d" % c], bool(z["term_%d" % c]), float(z["gamma_%d" % c]),
                                     int(z["L_%d" % c]))
        np.testing.assert_allclose(G, z["G_%d" % c], rtol=1e-13, atol=1e-13)
        np.testing.assert_allclose(adv, z["adv_%d" % c], rtol=1e-13, atol=1e-12)


def test_path_adv_columns_matches_reference_golden(golden_dir):
    z = np.load(os.path.join(golden_dir, "pathadv_cols.npz"))
    for c in range(int(z["n_cases"])):
        G, adv = PA.path_adv_columns(z["rew_%d" % c], z["done_%d" % c], z["V_%d" % c], float(z["gamma_%d" % c]),
                                     int(z["L_%d" % c]))
        np.testing.assert_allclose(G, z["G_%d" % c], rtol=1e-12, atol=1e-12)
        np.testing.assert_allclose(adv, z["adv_%d" % c], rtol=1e-12, atol=1e-11)


def test_path_adv_equals_plain_reverse_scan_iff_short(golden_dir):
    rng = np.random.default_rng(0)
    for n, same in ((20, True), (40, True), (41, False), (200, False)):
        r, v = rng.normal(size=n), rng.normal(size=n + 1)
        G, _ = PA.path_adv_episode(r, v, False, 0.98, 40)
        plain = np.zeros(n)
        nxt = v[n]
        for i in range(n - 1, -1, -1):
            nxt = plain[i] = r[i] + 0.98 * nxt
        assert np.allclose(G, plain, rtol=1e-12) == same


def test_framer_schedule_vaf_advnorm_match_reference_golden(golden_dir):
    z = np.load(os.path.join(golden_dir, "framer.npz"))
    obs = list(z["obs"])
    for k in (1, 2, 3):
        np.testing.assert_array_equal(np.array(PA.framer_full(obs, k)), z["full_%d" % k])
        np.testing.assert_array_equal(PA.framer_last(obs, k), z["last_%d" % k])
        np.testing.assert_array_equal(PA.framer_last(obs[:3], k), z["last3_%d" % k])
    z = np.load(os.path.join(golden_dir, "schedule.npz"))
    for t, g, b in zip(z["ts"], z["gamma"], z["beta"]):
        assert PA.linear_schedule_val(int(t), 100, 3000, -2, -8) == g
        assert PA.linear_schedule_val(int(t), 100, 3000, 0, -4) == b
    z = np.load(os.path.join(golden_dir, "vaf.npz"))
    np.testing.assert_allclose(PA.var_accounted_for(z["target"], z["pred"]), z["vaf"], rtol=1e-14)
    z = np.load(os.path.join(golden_dir, "advnorm.npz"))
    np.testing.assert_allclose(P.normalize_adv(z["advs"]), z["norm"], rtol=1e-14, atol=1e-15)


APPENDIX_B = [((1, 0, 0), -0.086827753, -6.202648083e-03), ((-1, 0, 0), 0.151770045, -248.6037307),
              ((0, 1, 0), -1.819866438, -133.2513682), ((0, -1, 0), 1.847806339, -125.0141224),
              ((np.cos(0.5), np.sin(0.5), 1.0), -1.982684587, -66.72053824),
              ((np.cos(3.0), np.sin(3.0), -8.0), -1.961357115, -76.33828604),
              ((np.cos(-2.0), np.sin(-2.0), 4.5), 1.847943787, -102.1846655)]


@pytest.mark.parametrize("dtype,tol", [(np.float64, 1e-8), (np.float32, 2e-6)])
def test_appendix_b_known_answers(dtype, tol):
    actor, critic = demo_weights()
    lay, clay = P.actor_layout(3, 1, True), P.critic_layout(3)
    obs = np.array([o for o, _, _ in APPENDIX_B])
    _, lp, ent, mu = P.actor_act(actor, lay, obs, np.zeros((len(obs), 1)), True, 2.0, dtype=dtype)
    v = P.critic_value(critic, clay, obs, dtype=dtype)
    np.testing.assert_allclose(mu[:, 0], [m for _, m, _ in APPENDIX_B], rtol=max(tol, 2e-9) * 5)
    np.testing.assert_allclose(v, [x for _, _, x in APPENDIX_B], rtol=tol * 5, atol=tol)
    np.testing.assert_allclose(ent, 1.3884116139142493, rtol=1e-7)
    lp5 = P.actor_logp(actor, lay, obs[4:5], np.array([[0.5]]), True, 2.0, dtype=dtype)
    np.testing.assert_allclose(lp5, -4.164295090913109, rtol=max(tol, 1e-7) * 5)


def test_greedy_pendulum_returns_known_answers():
    """greedy 200-step returns of the shipped checkpoint under the restated dynamics."""
    actor, _ = demo_weights()
    lay = P.actor_layout(3, 1, True)
    starts = np.array([[np.pi, 0.0], [0.0, 0.0], [1.0, 0.5], [-2.5, -1.0], [3.0, 1.0]]).T
    want = [-429.0372, -0.017548, -125.5242, -236.8588, -354.4552]
    s, el, ret = starts.copy(), np.zeros(5, np.int32), np.zeros(5)
    for t in range(200):
        obs = E.observe(E.PENDULUM, s)
        _, _, _, mu = P.actor_act(actor, lay, obs.T, np.zeros((5, 1)), True, 2.0, dtype=np.float64)
        s, _, r, _ = (lambda ns, rw, tm: (ns, None, rw, tm))(*E.dynamics(E.PENDULUM, s, np.clip(mu.T, -2, 2)))
        ret += r
    np.testing.assert_allclose(ret, want, rtol=2e-5, atol=2e-5)


def _torch_actor_loss(flat, lay, obs, acs, advs, lp_old, beta, gam, cont, ac_scale):
    T = {k: torch.tensor(np.asarray(flat[o:o + int(np.prod(s))], dtype=np.float64).reshape(s), requires_grad=True)
         for k, o, s in P.items(lay)}
    lrelu = lambda x: 0.8 * torch.relu(x) + 0.2 * x
    x = lrelu(torch.tensor(obs, dtype=torch.float64) @ T["W1"] + T["b1"])
    x = lrelu(x @ T["W2"] + T["b2"])
    x = lrelu(x @ T["W3"] + T["b3"])
    if cont:
        mu = torch.tanh(x @ T["Wh"] + T["bh"]) * ac_scale
        ls = torch.clamp(T["log_std"], -2.5, 2.5)
        dist = torch.distributions.Normal(mu, torch.exp(ls) * torch.ones_like(mu))
        lp = dist.log_prob(torch.tensor(acs, dtype=torch.float64)).sum(1)
        ent = dist.entropy().sum(1)
    else:
        logits = x @ T["Wh"] + T["bh"]
        logps = torch.log_softmax(logits + 1e-8, dim=1)
        ent = -(torch.softmax(logits, dim=1) * logps).sum(1)
        lp = logps[torch.arange(len(obs)), torch.tensor(acs, dtype=torch.long)]
    loss = -(torch.tensor(advs) * lp).mean() + beta * ((torch.tensor(lp_old) - lp) ** 2).mean() - gam * ent.mean()
    loss.backward()
    return loss.item(), {k: v.grad.numpy() for k, v in T.items()}


@pytest.mark.parametrize("cont,ob,nac", [(True, 3, 1), (True, 3, 2), (False, 4, 2), (False, 6, 3)])
def test_actor_analytic_gradients_match_autograd(cont, ob, nac):
    actor, _ = random_weights(ob, nac, cont, 42, bias_scale=0.3)
    lay = P.actor_layout(ob, nac, cont)
    rng = np.random.default_rng(1)
    B = 97
    obs = rng.normal(size=(B, ob))
    acs = rng.normal(size=(B, nac)) if cont else rng.integers(0, nac, B)
    advs, lp_old = rng.normal(size=B), rng.normal(size=B) - 1
    out = P.actor_loss_grads(actor, lay, obs, acs, advs, lp_old, 0.7, 0.05, cont, 2.0, clip=None, dtype=np.float64)
    loss, g = _torch_actor_loss(actor, lay, obs, acs, advs, lp_old, 0.7, 0.05, cont, 2.0)
    np.testing.assert_allclose(out["loss"], loss, rtol=1e-12)
    for k, o, s in P.items(lay):
        np.testing.assert_allclose(out["grad_raw"][o:o + int(np.prod(s))].reshape(s), g[k], rtol=1e-9, atol=1e-12,
                                   err_msg=k)
    clipped = P.actor_loss_grads(actor, lay, obs, acs, advs, lp_old, 0.7, 0.05, cont, 2.0, clip=0.01)["grad"]
    assert np.abs(clipped).max() <= 0.01 + 1e-15


def test_log_std_clip_blocks_gradient():
    actor, _ = random_weights(3, 1, True, 1)
    lay = P.actor_layout(3, 1, True)
    actor[lay["log_std"][0]] = 3.0          # outside [-2.5, 2.5] -> clip_by_value passes no gradient
    rng = np.random.default_rng(0)
    out = P.actor_loss_grads(actor, lay, rng.normal(size=(9, 3)), rng.normal(size=(9, 1)), rng.normal(size=9),
                             rng.normal(size=9), 1.0, 0.01, True)
    assert out["grad_raw"][lay["log_std"][0]] == 0.0


@pytest.mark.parametrize("basic", [False, True])
def test_critic_analytic_gradients_match_autograd(basic):
    _, critic = random_weights(4, 2, False, 43, bias_scale=0.3)
    lay = P.critic_layout(4)
    rng = np.random.default_rng(2)
    obs, tg = rng.normal(size=(50, 4)), rng.normal(size=50)
    out = P.critic_loss_grads(critic, lay, obs, tg, basic_head=basic)
    T = {k: torch.tensor(np.asarray(critic[o:o + int(np.prod(s))], dtype=np.float64).reshape(s), requires_grad=True)
         for k, o, s in P.items(lay)}
    x = torch.relu(torch.tensor(obs) @ T["W1"] + T["b1"])
    x1 = torch.relu(x @ T["W2"] + T["b2"])
    x2 = torch.relu(x1 @ T["W3"] + T["b3"])
    v = ((x1 if basic else x2) @ T["Wv"] + T["bv"]).reshape(-1)
    loss = ((v - torch.tensor(tg)) ** 2).mean()
    loss.backward()
    np.testing.assert_allclose(out["loss"], loss.item(), rtol=1e-12)
    for k, o, s in P.items(lay):
        want = np.zeros(s) if T[k].grad is None else T[k].grad.numpy()
        np.testing.assert_allclose(out["grad"][o:o + int(np.prod(s))].reshape(s), want, rtol=1e-9, atol=1e-12, err_msg=k)


def test_adam_matches_torch_adam_semantics():
    """TF ApplyAdam (epsilon outside the bias correction) differs from torch.optim.Adam only in
    where eps sits; with eps -> 0 they coincide.  Checks the update direction / bias correction."""
    rng = np.random.default_rng(0)
    th0 = rng.normal(size=100)
    p = torch.tensor(th0.copy(), requires_grad=True)
    opt = torch.optim.Adam([p], lr=0.01, eps=1e-30)
    th, m, v = th0.copy(), np.zeros(100), np.zeros(100)
    for t in range(1, 6):
        g = rng.normal(size=100)
        p.grad = torch.tensor(g.copy())
        opt.step()
        th, m, v = P.adam_step(th, g, m, v, t, 0.01, eps=1e-30, dtype=np.float64)
        np.testing.assert_allclose(th, p.detach().numpy(), rtol=1e-9)


def test_kl_lr_rule():
    assert P.kl_lr_rule(0.005, 0.0004, 0.002) == pytest.approx(0.0075)
    assert P.kl_lr_rule(0.005, 0.009, 0.002) == pytest.approx(0.005 / 1.5)
    assert P.kl_lr_rule(0.005, 0.002, 0.002) == 0.005
    assert P.kl_lr_rule(0.09, 0.0, 0.002) == 0.1 and P.kl_lr_rule(0.9, 0.0, 0.002, max_lr=1.0) == 1.0
    assert P.kl_lr_rule(1.2e-6, 1.0, 0.002) == 1e-6


def test_cat_sample_inverse_cdf():
    logits = np.log(np.array([[0.2, 0.3, 0.5]]))
    for u, a in ((0.0, 0), (0.19, 0), (0.21, 1), (0.49, 1), (0.51, 2), (0.999999, 2)):
        assert P.cat_sample(logits, np.array([u]))[0] == a
    rng = np.random.default_rng(0)
    acs = P.cat_sample(np.repeat(logits, 200000, 0), rng.random(200000))
    np.testing.assert_allclose(np.bincount(acs) / 200000, [0.2, 0.3, 0.5], atol=5e-3)


def test_philox_known_answer_and_distributions():
    # Random123 known-answer test: philox4x32-10, ctr = key = 0 -> 6627e8d5 e169c58d bc57ac4c 9b00dbd8
    x = R.philox4x32(np.array([0], dtype=np.uint64), 0, 0, 0)[0]
    assert [hex(int(v)) for v in x] == ["0x6627e8d5", "0xe169c58d", "0xbc57ac4c", "0x9b00dbd8"]
    # ctr = key = all ones -> 408f276d 41c83b0e a20bc7c6 6d5451fd
    y = R.philox4x32(np.array([0xFFFFFFFFFFFFFFFF], dtype=np.uint64), 0xFFFFFFFF, 0xFFFFFFFF, 0xFFFFFFFFFFFFFFFF)[0]
    assert [hex(int(v)) for v in y] == ["0x408f276d", "0x41c83b0e", "0xa20bc7c6", "0x6d5451fd"]
    u = R.uniforms(np.arange(100000, dtype=np.uint64), 3, 1, 7)
    assert 0 <= u.min() and u.max() < 1 and abs(u.mean() - 0.5) < 5e-3
    z = R.normals(np.arange(100000, dtype=np.uint64), 3, 2, 7)
    assert abs(z.mean()) < 1e-2 and abs(z.std() - 1) < 1e-2 and np.isfinite(z).all()


@pytest.mark.parametrize("env", [E.PENDULUM, E.CARTPOLE, E.ACROBOT])
def test_env_fp32_mode_tracks_fp64(env):
    rng = np.random.default_rng(env)
    n = 2000
    u = rng.random((E.STATE_DIM[env], n))
    s = E.reset_from_uniforms(env, u)
    a = rng.uniform(-2, 2, (1, n)) if env == E.PENDULUM else rng.integers(0, E.N_ACTIONS[env], n)
    out64 = E.step(env, s, np.zeros(n, np.int32), a, u, dtype=np.float64)
    out32 = E.step(env, s.astype(np.float32), np.zeros(n, np.int32), a, u.astype(np.float32), dtype=np.float32)
    for x, y in zip(out64, out32):
        np.testing.assert_allclose(y, x, rtol=1e-5, atol=1e-5)
    assert out32[2].dtype == np.float32


def test_env_semantics():
    # Pendulum: reward uses the pre-step state, never terminates, TimeLimit at 200 resets the env
    s = np.array([[np.pi / 2], [1.0]])
    ns, el, obs, r, d = E.step(E.PENDULUM, s, np.array([198]), np.array([[5.0]]), np.array([[0.5], [0.5]]))
    assert r[0] == pytest.approx(-((np.pi / 2) ** 2 + 0.1 + 0.001 * 4)) and d[0] == 0 and el[0] == 199
    ns, el, obs, r, d = E.step(E.PENDULUM, ns, el, np.array([[0.0]]), np.array([[0.5], [0.5]]))
    assert d[0] == 1 and el[0] == 0 and ns[0, 0] == pytest.approx(0.0) and ns[1, 0] == pytest.approx(0.0)
    np.testing.assert_allclose(obs[:, 0], [1.0, 0.0, 0.0], atol=1e-12)
    # CartPole: reward 1 including the terminating step; leaving the track terminates
    s = np.array([[2.39], [5.0], [0.0], [0.0]])
    ns, el, obs, r, d = E.step(E.CARTPOLE, s, np.array([3]), np.array([1]), np.full((4, 1), 0.5))
    assert d[0] == 1 and r[0] == 1.0 and el[0] == 0 and np.allclose(ns, 0.0)
    # Acrobot: -1 per step until the tip rises above the bar
    s = np.array([[np.pi], [0.0], [0.0], [0.0]])
    ns, el, obs, r, d = E.step(E.ACROBOT, s, np.array([0]), np.array([1]), np.full((4, 1), 0.5))
    assert d[0] == 1 and r[0] == 0.0
    s = np.zeros((4, 1))
    ns, el, obs, r, d = E.step(E.ACROBOT, s, np.array([0]), np.array([2]), np.full((4, 1), 0.5))
    assert d[0] == 0 and r[0] == -1.0 and obs.shape == (6, 1)


def test_a2c_iteration_runs_and_is_reproducible():
    outs = []
    for _ in range(2):
        S = a2c.A2CState(E.PENDULUM, 64, seed=5)
        st = [a2c.iteration(S, T=8) for _ in range(2)]
        outs.append((S.actor.copy(), S.critic.copy(), st[-1]["kl"], S.actor_lr))
    np.testing.assert_array_equal(outs[0][0], outs[1][0])
    np.testing.assert_array_equal(outs[0][1], outs[1][1])
    assert np.isfinite(outs[0][0]).all() and outs[0][2] >= 0
    S = a2c.A2CState(E.CARTPOLE, 32, seed=5)
    st = a2c.iteration(S, T=8)
    assert np.isfinite(st["actor_loss"]) and S.it == 1


def test_a2c_rollout_stacks_frames_like_the_reference_framer():
    """oracle/a2c.py with frames=2: every stacked observation equals Framer.full over the episode it belongs to."""
    from oracle import a2c, envs as E
    S = a2c.A2CState(E.CARTPOLE, 16, seed=5, frames=2)
    assert S.ob_dim == 8 and S.obs.shape == (8, 16)
    ro = a2c.rollout(S, 12, max_steps=4)
    obs, done = ro["obs"], ro["done"]
    assert done.sum() > 0
    for e in range(16):
        start = 0
        for t in range(13):
            if t > 0 and done[t - 1, e]:
                start = t
            raw = [obs[u, 4:, e] for u in range(start, t + 1)]          # newest frame = the raw observation
            np.testing.assert_array_equal(obs[t, :, e], PA.framer_full(raw, 2)[-1])
