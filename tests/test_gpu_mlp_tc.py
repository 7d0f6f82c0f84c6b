"""K2 on the tensor cores (tcgen05, 3xTF32): the *_tc entry points against the oracle and against the
FFMA entry points, same inputs, same tolerances as tests/test_gpu_mlp.py (1e-5 relative)."""
import ctypes

import numpy as np
import pytest
import torch

from oracle import policies as P
from gpu_util import act_cache, demo_weights, dev_t, random_weights, rel_err, soa

pytestmark = pytest.mark.gpu
TOL = 1e-5


def pack(lib, dev, net, params_dev, ob, nac, cont):
    n = lib.load().acx_tc_packed_floats(net, ob, nac, int(cont))
    packed = torch.zeros(n, device=dev)
    lib.call("acx_tc_pack_weights", net, lib.ptr(params_dev), ob, nac, int(cont), lib.ptr(packed), lib.cur_stream())
    return packed


@pytest.mark.parametrize("rows,n", [(1, 1), (1, 127), (1, 128), (3, 129), (2, 1000), (21, 4096)])
@pytest.mark.parametrize("weights", ["demo", "random"])
def test_gauss_and_value_tc_match_oracle(lib, dev, rows, n, weights):
    actor, critic = demo_weights() if weights == "demo" else random_weights(3, 1, True, 7)
    rng = np.random.default_rng(rows * 1000 + n)
    th = rng.uniform(-np.pi, np.pi, (rows, n))
    obs = np.stack([np.cos(th), np.sin(th), rng.uniform(-8, 8, (rows, n))], axis=-1).astype(np.float32)
    eps = rng.normal(size=(rows, n, 1)).astype(np.float32)
    a, c = dev_t(actor, dev), dev_t(critic, dev)
    pa, pc = pack(lib, dev, 0, a, 3, 1, True), pack(lib, dev, 1, c, 3, 1, True)
    o = dev_t(np.stack([soa(x) for x in obs]), dev)
    e = dev_t(np.stack([soa(x) for x in eps]), dev)
    ac, lp, ent, mu = (torch.empty((rows, 1, n), device=dev), torch.empty((rows, n), device=dev),
                       torch.empty((rows, n), device=dev), torch.empty((rows, 1, n), device=dev))
    v = torch.empty((rows, n), device=dev)
    lib.call("acx_policy_forward_sample_gauss_tc", lib.ptr(a), lib.ptr(pa), 3, 1, 2.0, rows, n, lib.ptr(o), lib.ptr(e), 0, 0,
             0, None, lib.ptr(ac), lib.ptr(lp), lib.ptr(ent), lib.ptr(mu), None, None, lib.cur_stream())
    lib.call("acx_value_forward_tc", lib.ptr(c), lib.ptr(pc), 3, rows, n, lib.ptr(o), lib.ptr(v), 0, None, None, None, lib.cur_stream())
    # log-prob of the sampled actions must reproduce lp; KL partials against lp + 0.5
    lp2 = torch.empty((rows, n), device=dev)
    old = lp + 0.5
    part = torch.zeros(1024, device=dev, dtype=torch.float64)
    nb = ctypes.c_int(0)
    lib.call("acx_policy_logp_gauss_tc", lib.ptr(a), lib.ptr(pa), 3, 1, 2.0, rows, n, lib.ptr(o), lib.ptr(ac), lib.ptr(lp2),
             None, None, lib.ptr(old), lib.ptr(part), ctypes.byref(nb), lib.cur_stream())
    stats = torch.zeros(16, device=dev)
    lib.call("acx_kl_finalize", lib.ptr(part), nb.value, float(rows * n), lib.ptr(stats), lib.cur_stream())
    torch.cuda.synchronize()
    ac, lp, ent, mu, v, lp2 = (x.cpu().numpy() for x in (ac, lp, ent, mu, v, lp2))
    lay, clay = P.actor_layout(3, 1, True), P.critic_layout(3)
    for r in range(rows):
        wac, wlp, went, wmu = P.actor_act(actor, lay, obs[r], eps[r], True, 2.0, dtype=np.float64)
        assert rel_err(mu[r].T, wmu) < 2 * TOL
        assert rel_err(ac[r].T, wac) < 2 * TOL
        np.testing.assert_allclose(lp[r], wlp, rtol=20 * TOL, atol=20 * TOL)
        assert rel_err(ent[r], went) < TOL
        assert rel_err(v[r], P.critic_value(critic, clay, obs[r], dtype=np.float64)) < 2 * TOL
    np.testing.assert_allclose(lp2, lp, rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(stats[7].item(), 0.25, rtol=1e-4)


@pytest.mark.parametrize("ob,n_act", [(4, 2), (6, 3)])
def test_cat_tc_matches_oracle_and_ffma(lib, dev, ob, n_act):
    actor, critic = random_weights(ob, n_act, False, 11, bias_scale=0.5)
    lay = P.actor_layout(ob, n_act, False)
    off, shp = lay["Wh"]
    actor[off:off + shp[0] * shp[1]] *= 3
    rows, n = 2, 3000
    rng = np.random.default_rng(n + ob)
    obs = rng.normal(size=(rows, n, ob)).astype(np.float32)
    u = rng.random((rows, n)).astype(np.float32)
    a, c = dev_t(actor, dev), dev_t(critic, dev)
    pa, pc = pack(lib, dev, 0, a, ob, n_act, False), pack(lib, dev, 1, c, ob, 1, False)
    o, uu = dev_t(np.stack([soa(x) for x in obs]), dev), dev_t(u, dev)
    out = {}
    for tag in ("tc", "ffma"):
        ac = torch.empty((rows, n), device=dev, dtype=torch.int32)
        lp, ent, lg = torch.empty((rows, n), device=dev), torch.empty((rows, n), device=dev), torch.empty((rows, n_act, n), device=dev)
        v = torch.empty((rows, n), device=dev)
        if tag == "tc":
            lib.call("acx_policy_forward_sample_cat_tc", lib.ptr(a), lib.ptr(pa), ob, n_act, rows, n, lib.ptr(o), lib.ptr(uu), 0,
                     0, 0, None, lib.ptr(ac), lib.ptr(lp), lib.ptr(ent), lib.ptr(lg), None, None, lib.cur_stream())
            lib.call("acx_value_forward_tc", lib.ptr(c), lib.ptr(pc), ob, rows, n, lib.ptr(o), lib.ptr(v), 0, None, None, None, lib.cur_stream())
        else:
            lib.call("acx_policy_forward_sample_cat", lib.ptr(a), ob, n_act, rows, n, lib.ptr(o), lib.ptr(uu), 0, 0, 0, None,
                     lib.ptr(ac), lib.ptr(lp), lib.ptr(ent), lib.ptr(lg), lib.cur_stream())
            lib.call("acx_value_forward", lib.ptr(c), ob, rows, n, lib.ptr(o), lib.ptr(v), 0, lib.cur_stream())
        torch.cuda.synchronize()
        out[tag] = [x.cpu().numpy() for x in (ac, lp, ent, lg, v)]
    ac, lp, ent, lg, v = out["tc"]
    for r in range(rows):
        wac, wlp, went, wlg = P.actor_act(actor, lay, obs[r], u[r], False, dtype=np.float64)
        assert rel_err(lg[r].T, wlg) < 2 * TOL
        np.testing.assert_array_equal(ac[r], P.cat_sample(lg[r].T, u[r]))      # bit-exact on the kernel's own logits
        assert np.mean(ac[r] != wac) < 1e-3                                     # end to end: only threshold ties may flip
        np.testing.assert_allclose(lp[r], wlp, rtol=2 * TOL, atol=4e-6)
        np.testing.assert_allclose(ent[r], went, rtol=2 * TOL, atol=4e-6)
    assert rel_err(lg, out["ffma"][3]) < 2 * TOL and rel_err(v, out["ffma"][4]) < 2 * TOL
    assert np.mean(ac != out["ffma"][0]) < 1e-3


def test_tc_basic_head_and_argument_checks(lib, dev):
    _, critic = random_weights(4, 2, False, 5)
    c = dev_t(critic, dev)
    pc = pack(lib, dev, 1, c, 4, 1, False)
    obs = np.random.default_rng(0).normal(size=(1, 300, 4)).astype(np.float32)
    o = dev_t(soa(obs[0])[None], dev)
    v = torch.empty((1, 300), device=dev)
    lib.call("acx_value_forward_tc", lib.ptr(c), lib.ptr(pc), 4, 1, 300, lib.ptr(o), lib.ptr(v), 1, None, None, None, lib.cur_stream())
    torch.cuda.synchronize()
    want = P.critic_value(critic, P.critic_layout(4), obs[0], basic_head=True, dtype=np.float64)
    assert rel_err(v.cpu().numpy()[0], want) < 2 * TOL
    f = lib.load().acx_value_forward_tc
    assert f(c.data_ptr(), None, 4, 1, 8, o.data_ptr(), v.data_ptr(), 0, None, None, None, None) == -1
    assert b"null" in lib.load().acx_last_error()
    assert f(c.data_ptr(), pc.data_ptr() + 4, 4, 1, 8, o.data_ptr(), v.data_ptr(), 0, None, None, None, None) == -1
    assert f(c.data_ptr(), pc.data_ptr(), 4, 0, 8, o.data_ptr(), v.data_ptr(), 0, None, None, None, None) == 0


def test_tc_forward_fills_the_activation_cache_like_the_ffma_forward(lib, dev):
    actor, critic = random_weights(3, 1, True, 9)
    rows, n = 2, 777
    rng = np.random.default_rng(3)
    obs = rng.normal(size=(rows, n, 3)).astype(np.float32)
    eps = rng.normal(size=(rows, n, 1)).astype(np.float32)
    a, c = dev_t(actor, dev), dev_t(critic, dev)
    pa, pc = pack(lib, dev, 0, a, 3, 1, True), pack(lib, dev, 1, c, 3, 1, True)
    o = dev_t(np.stack([soa(x) for x in obs]), dev)
    e = dev_t(np.stack([soa(x) for x in eps]), dev)
    res = {}
    for tag in ("tc", "ffma"):
        ac, lp = torch.empty((rows, 1, n), device=dev), torch.empty((rows, n), device=dev)
        v = torch.empty((rows, n), device=dev)
        ax2, ax3 = act_cache(lib, dev, rows, 128, n), act_cache(lib, dev, rows, 64, n)
        cx2, cx3 = act_cache(lib, dev, rows, 128, n), act_cache(lib, dev, rows, 128, n)
        if tag == "tc":
            lib.call("acx_policy_forward_sample_gauss_tc", lib.ptr(a), lib.ptr(pa), 3, 1, 2.0, rows, n, lib.ptr(o), lib.ptr(e),
                     0, 0, 0, None, lib.ptr(ac), lib.ptr(lp), None, None, lib.ptr(ax2), lib.ptr(ax3), lib.cur_stream())
            lib.call("acx_value_forward_tc", lib.ptr(c), lib.ptr(pc), 3, rows, n, lib.ptr(o), lib.ptr(v), 0, lib.ptr(cx2),
                     lib.ptr(cx3), None, lib.cur_stream())
        else:
            lib.call("acx_policy_forward_sample_gauss_cached", lib.ptr(a), 3, 1, 2.0, rows, n, lib.ptr(o), lib.ptr(e), 0, 0,
                     0, None, lib.ptr(ac), lib.ptr(lp), None, None, lib.ptr(ax2), lib.ptr(ax3), lib.cur_stream())
            lib.call("acx_value_forward_cached", lib.ptr(c), 3, rows, n, lib.ptr(o), lib.ptr(v), 0, lib.ptr(cx2), lib.ptr(cx3),
                     lib.cur_stream())
        torch.cuda.synchronize()
        res[tag] = [x.cpu().numpy() for x in (ax2, ax3, cx2, cx3, ac, v)]
    for x, y in zip(res["tc"], res["ffma"]):
        assert rel_err(x, y) < 2 * TOL
    assert np.abs(res["tc"][0]).max() > 0
    # layout: tile-blocked [rows][tiles][features][128], zeros beyond n; values = the critic's second hidden layer
    lay = P.critic_layout(3)
    W = {k: critic[o:o + int(np.prod(sh))].reshape(sh).astype(np.float64) for k, o, sh in P.items(lay)}
    x1 = np.maximum(obs.astype(np.float64) @ W["W1"] + W["b1"], 0.0)
    x2 = np.maximum(x1 @ W["W2"] + W["b2"], 0.0)                              # [rows, n, 128]
    for tag in ("tc", "ffma"):
        blk = res[tag][2]                                                     # [rows, tiles, 128, 128]
        flat = blk.transpose(0, 2, 1, 3).reshape(rows, 128, -1)               # [rows, feature, tiles * 128]
        np.testing.assert_allclose(flat[:, :, :n], x2.transpose(0, 2, 1), rtol=2e-5, atol=2e-5)
        assert not flat[:, :, n:].any()
