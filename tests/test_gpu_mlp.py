"""K2 parity: fused policy / value forward + sampling through the C ABI vs the oracle.

Tolerance (north_star): log-probs / values within 1e-5 relative in fp32; categorical action
indices bit-exact given the same uniforms (checked on the kernel's own logits, and end to end).
Known answers: SURVEY.md Appendix B (self-derived on the reference's shipped checkpoint).
"""
import numpy as np
import pytest
import torch

from oracle import philox as R
from oracle import policies as P
from gpu_util import demo_weights, dev_t, random_weights, rel_err, soa

pytestmark = pytest.mark.gpu
TOL = 1e-5


def gauss_sample(lib, dev, actor, obs_rows, eps_rows, ac_scale=2.0, seed=0, index0=0, step0=0):
    """obs_rows [rows, n, ob]; eps_rows [rows, n, ac] or None."""
    rows, n, ob = obs_rows.shape
    ac_dim = 1 if eps_rows is None else eps_rows.shape[2]
    a = dev_t(actor, dev)
    o = dev_t(np.stack([soa(x) for x in obs_rows]), dev)
    e = None if eps_rows is None else dev_t(np.stack([soa(x) for x in eps_rows]), dev)
    ac = torch.empty((rows, ac_dim, n), device=dev)
    lp = torch.empty((rows, n), device=dev)
    ent = torch.empty((rows, n), device=dev)
    mu = torch.empty((rows, ac_dim, n), device=dev)
    lib.call("acx_policy_forward_sample_gauss", lib.ptr(a), ob, ac_dim, ac_scale, rows, n, lib.ptr(o), lib.ptr(e), seed,
             index0, step0, None, lib.ptr(ac), lib.ptr(lp), lib.ptr(ent), lib.ptr(mu), lib.cur_stream())
    torch.cuda.synchronize()
    return ac.cpu().numpy(), lp.cpu().numpy(), ent.cpu().numpy(), mu.cpu().numpy()


def value_fwd(lib, dev, critic, obs_rows, basic=0):
    rows, n, ob = obs_rows.shape
    c = dev_t(critic, dev)
    o = dev_t(np.stack([soa(x) for x in obs_rows]), dev)
    v = torch.empty((rows, n), device=dev)
    lib.call("acx_value_forward", lib.ptr(c), ob, rows, n, lib.ptr(o), lib.ptr(v), basic, lib.cur_stream())
    torch.cuda.synchronize()
    return v.cpu().numpy()


APPENDIX_B = [  # obs, mu, V  
    ((1, 0, 0), -0.086827753, -6.202648083e-03), ((-1, 0, 0), 0.151770045, -248.6037307),
    ((0, 1, 0), -1.819866438, -133.2513682), ((0, -1, 0), 1.847806339, -125.0141224),
    ((np.cos(0.5), np.sin(0.5), 1.0), -1.982684587, -66.72053824),
    ((np.cos(3.0), np.sin(3.0), -8.0), -1.961357115, -76.33828604),
    ((np.cos(-2.0), np.sin(-2.0), 4.5), 1.847943787, -102.1846655)]


def test_demo_checkpoint_known_answers(lib, dev):
    actor, critic = demo_weights()
    obs = np.array([o for o, _, _ in APPENDIX_B], dtype=np.float32)[None]
    eps = np.zeros((1, len(APPENDIX_B), 1), dtype=np.float32)
    ac, lp, ent, mu = gauss_sample(lib, dev, actor, obs, eps)
    v = value_fwd(lib, dev, critic, obs)
    np.testing.assert_allclose(mu[0, 0], [m for _, m, _ in APPENDIX_B], rtol=2e-5)
    np.testing.assert_allclose(v[0], [x for _, _, x in APPENDIX_B], rtol=2e-5, atol=2e-6)
    np.testing.assert_allclose(ent[0], 1.3884116139142493, rtol=1e-6)
    np.testing.assert_array_equal(ac, mu)          # eps = 0 -> action = mu
    # logp(a = 0.5 | obs row 5) = -4.164295090913109
    a = dev_t(actor, dev)
    o = dev_t(soa(obs[0, 4:5]), dev)
    acs = torch.full((1, 1, 1), 0.5, device=dev)
    out = torch.empty((1, 1), device=dev)
    lib.call("acx_policy_logp_gauss", lib.ptr(a), 3, 1, 2.0, 1, 1, lib.ptr(o), lib.ptr(acs), lib.ptr(out), None, None,
             None, None, None, lib.cur_stream())
    torch.cuda.synchronize()
    np.testing.assert_allclose(out.item(), -4.164295090913109, rtol=1e-5)


@pytest.mark.parametrize("rows,n", [(1, 1), (1, 63), (1, 64), (3, 65), (2, 1000), (1, 20000)])
@pytest.mark.parametrize("weights", ["demo", "random"])
def test_gauss_forward_sample_matches_oracle(lib, dev, rows, n, weights):
    actor, critic = demo_weights() if weights == "demo" else random_weights(3, 1, True, 7)
    rng = np.random.default_rng(rows * 1000 + n)
    th = rng.uniform(-np.pi, np.pi, (rows, n))
    obs = np.stack([np.cos(th), np.sin(th), rng.uniform(-8, 8, (rows, n))], axis=-1).astype(np.float32)
    eps = rng.normal(size=(rows, n, 1)).astype(np.float32)
    ac, lp, ent, mu = gauss_sample(lib, dev, actor, obs, eps)
    v = value_fwd(lib, dev, critic, obs)
    lay, clay = P.actor_layout(3, 1, True), P.critic_layout(3)
    for r in range(rows):
        for dt, tol in ((np.float32, TOL), (np.float64, 3e-5)):
            wac, wlp, went, wmu = P.actor_act(actor, lay, obs[r], eps[r], True, 2.0, dtype=dt)
            assert rel_err(mu[r].T, wmu) < tol
            assert rel_err(ac[r].T, wac) < tol
            np.testing.assert_allclose(lp[r], wlp, rtol=10 * tol, atol=10 * tol)   # logp amplifies mu error by z/std
            assert rel_err(ent[r], went) < tol
            wv = P.critic_value(critic, clay, obs[r], dtype=dt)
            assert rel_err(v[r], wv) < tol


def test_gauss_device_rng_matches_philox_oracle(lib, dev):
    actor, _ = demo_weights()
    n, seed, index0, step0 = 4096, 12321, 70000, 123
    rng = np.random.default_rng(2)
    obs = rng.normal(size=(2, n, 3)).astype(np.float32)
    ac, lp, ent, mu = gauss_sample(lib, dev, actor, obs, None, seed=seed, index0=index0, step0=step0)
    std = np.exp(np.float32(P.unpack(actor, P.actor_layout(3, 1, True))["log_std"][0]))
    for r in range(2):
        eps = R.normals(np.arange(n, dtype=np.uint64) + np.uint64(index0), step0 + r, R.STREAM_ACTION, seed)[:, 0]
        np.testing.assert_allclose((ac[r, 0] - mu[r, 0]) / std, eps, rtol=1e-4, atol=2e-5)
    assert abs(np.mean((ac - mu) / std)) < 0.05 and abs(np.std((ac - mu) / std) - 1) < 0.05


def cat_sample(lib, dev, actor, ob, n_act, obs_rows, u_rows, seed=0, index0=0, step0=0):
    rows, n, _ = obs_rows.shape
    a = dev_t(actor, dev)
    o = dev_t(np.stack([soa(x) for x in obs_rows]), dev)
    u = None if u_rows is None else dev_t(u_rows.astype(np.float32), dev)
    ac = torch.empty((rows, n), device=dev, dtype=torch.int32)
    lp = torch.empty((rows, n), device=dev)
    ent = torch.empty((rows, n), device=dev)
    lg = torch.empty((rows, n_act, n), device=dev)
    lib.call("acx_policy_forward_sample_cat", lib.ptr(a), ob, n_act, rows, n, lib.ptr(o), lib.ptr(u), seed, index0,
             step0, None, lib.ptr(ac), lib.ptr(lp), lib.ptr(ent), lib.ptr(lg), lib.cur_stream())
    torch.cuda.synchronize()
    return ac.cpu().numpy(), lp.cpu().numpy(), ent.cpu().numpy(), lg.cpu().numpy()


@pytest.mark.parametrize("ob,n_act", [(4, 2), (6, 3)])
@pytest.mark.parametrize("rows,n", [(1, 5), (2, 777), (1, 30000)])
def test_cat_forward_sample_matches_oracle(lib, dev, ob, n_act, rows, n):
    actor, critic = random_weights(ob, n_act, False, 11, bias_scale=0.5)
    lay = P.actor_layout(ob, n_act, False)
    o, s = lay["Wh"]
    actor[o:o + s[0] * s[1]] *= 3          # spread the logits so the softmax is not ~uniform
    rng = np.random.default_rng(n + ob)
    obs = rng.normal(size=(rows, n, ob)).astype(np.float32)
    u = rng.random((rows, n)).astype(np.float32)
    ac, lp, ent, lg = cat_sample(lib, dev, actor, ob, n_act, obs, u)
    v = value_fwd(lib, dev, critic, obs)
    for r in range(rows):
        wac, wlp, went, wlg = P.actor_act(actor, lay, obs[r], u[r], False, dtype=np.float32)
        assert rel_err(lg[r].T, wlg) < TOL
        # bit-exact action indices: (a) on the kernel's own logits, (b) end to end
        np.testing.assert_array_equal(ac[r], P.cat_sample(lg[r].T, u[r]))
        np.testing.assert_array_equal(ac[r], wac)
        np.testing.assert_allclose(lp[r], wlp, rtol=TOL, atol=2e-6)
        np.testing.assert_allclose(ent[r], went, rtol=TOL, atol=2e-6)
        assert rel_err(v[r], P.critic_value(critic, P.critic_layout(ob), obs[r], dtype=np.float32)) < TOL
    assert len(np.unique(ac)) >= 2 or n < 10


def test_cat_device_rng_and_logp_eval(lib, dev):
    ob, n_act, n, seed = 4, 2, 5000, 99
    actor, _ = random_weights(ob, n_act, False, 3)
    rng = np.random.default_rng(1)
    obs = rng.normal(size=(1, n, ob)).astype(np.float32)
    ac, lp, ent, lg = cat_sample(lib, dev, actor, ob, n_act, obs, None, seed=seed, index0=10, step0=4)
    u = R.uniforms(np.arange(n, dtype=np.uint64) + np.uint64(10), 4, R.STREAM_ACTION, seed)[:, 0]
    np.testing.assert_array_equal(ac[0], P.cat_sample(lg[0].T, u))
    # evaluating the stored actions reproduces logp, and the KL partials give mean((lp_old - lp)^2)
    a, o = dev_t(actor, dev), dev_t(soa(obs[0])[None], dev)
    acs = dev_t(ac, dev)
    lp_old = dev_t((lp + 0.25).astype(np.float32), dev)
    out = torch.empty((1, n), device=dev)
    part = torch.zeros(1024, device=dev, dtype=torch.float64)
    import ctypes
    nb = ctypes.c_int(0)
    lib.call("acx_policy_logp_cat", lib.ptr(a), ob, n_act, 1, n, lib.ptr(o), lib.ptr(acs), lib.ptr(out), None, None,
             lib.ptr(lp_old), lib.ptr(part), ctypes.byref(nb), lib.cur_stream())
    stats = torch.zeros(16, device=dev)
    lib.call("acx_kl_finalize", lib.ptr(part), nb.value, float(n), lib.ptr(stats), lib.cur_stream())
    torch.cuda.synchronize()
    np.testing.assert_allclose(out.cpu().numpy()[0], lp[0], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(stats[7].item(), 0.0625, rtol=1e-5)


def test_value_basic_head(lib, dev):
    _, critic = random_weights(4, 2, False, 5)
    obs = np.random.default_rng(0).normal(size=(1, 300, 4)).astype(np.float32)
    v = value_fwd(lib, dev, critic, obs, basic=1)
    want = P.critic_value(critic, P.critic_layout(4), obs[0], basic_head=True, dtype=np.float32)
    assert rel_err(v[0], want) < TOL


def test_forward_rejects_bad_arguments(lib, dev):
    t = torch.zeros(70000, device=dev)
    f = lib.load().acx_value_forward
    assert f(t.data_ptr() + 4, 3, 1, 8, t.data_ptr(), t.data_ptr(), 0, None) == -1      # misaligned params
    assert b"aligned" in lib.load().acx_last_error()
    assert f(t.data_ptr(), 25, 1, 8, t.data_ptr(), t.data_ptr(), 0, None) == -1         # ob_dim too large (FFMA engine: 24)
    assert f(t.data_ptr(), 3, 0, 8, t.data_ptr(), t.data_ptr(), 0, None) == 0           # empty batch: no-op
