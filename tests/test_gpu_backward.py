"""K4 / K5 parity: fused loss + backward, Adam apply, KL-lr rule through the C ABI vs the oracle's
analytic gradients (oracle/policies.py, themselves cross-checked against torch.autograd in
tests/test_oracle.py).  Tolerance 1e-5 relative per tensor (max-abs-diff / max-abs), fp32."""
import numpy as np
import pytest
import torch

from oracle import policies as P
from gpu_util import demo_weights, dev_t, random_weights, redraw_near_kinks, rel_err, soa

pytestmark = pytest.mark.gpu
TOL = 1e-5


def make_opt(lib, dev, actor_lr=0.005, critic_lr=0.001, beta=1.0, gamma_ent=0.01, clip=0.1, desired_kl=0.002,
             min_lr=1e-6, max_lr=0.1, schedule=0):
    opt = torch.zeros(16, device=dev)
    lib.call("acx_opt_init", lib.ptr(opt), actor_lr, critic_lr, beta, gamma_ent, clip, desired_kl, min_lr, max_lr,
             schedule, lib.cur_stream())
    return opt


def backward_gpu(lib, dev, cont, actor, critic, ob, num_ac, obs_rows, acs_rows, adv, lp_old, targets, beta, gamma_ent,
                 adv_stats=None):
    rows, n, _ = obs_rows.shape
    a, c = dev_t(actor, dev), dev_t(critic, dev)
    o = dev_t(np.stack([soa(x) for x in obs_rows]), dev)
    if cont:
        acs = dev_t(np.stack([soa(x) for x in acs_rows]), dev)
    else:
        acs = dev_t(acs_rows.astype(np.int32), dev)
    opt = make_opt(lib, dev, beta=beta, gamma_ent=gamma_ent)
    ga, gc = torch.full_like(a, 7.0), torch.full_like(c, 7.0)
    stats = torch.zeros(16, device=dev)
    st = None if adv_stats is None else dev_t(np.asarray(adv_stats, dtype=np.float32), dev)
    nbytes = lib.load().acx_loss_backward_workspace_bytes(ob, num_ac, int(cont))
    ws = torch.empty(nbytes, device=dev, dtype=torch.uint8)
    # keep every device tensor referenced until the kernels ran: lib.ptr() only returns an address
    adv_d, lp_d, tg_d = (dev_t(x.astype(np.float32), dev) for x in (adv, lp_old, targets))
    common = (lib.ptr(o), lib.ptr(acs), lib.ptr(adv_d), lib.ptr(st), lib.ptr(lp_d), lib.ptr(tg_d),
              lib.ptr(opt), lib.ptr(ga), lib.ptr(gc), lib.ptr(stats), lib.ptr(ws), nbytes, lib.cur_stream())
    for _ in range(2):   # twice into the same workspace: partials must not leak between calls
        if cont:
            lib.call("acx_loss_backward_gauss", lib.ptr(a), lib.ptr(c), ob, num_ac, 2.0, rows, n, *common)
        else:
            lib.call("acx_loss_backward_cat", lib.ptr(a), lib.ptr(c), ob, num_ac, rows, n, *common)
    torch.cuda.synchronize()
    return ga.cpu().numpy(), gc.cpu().numpy(), stats.cpu().numpy()


def check_grads(got, want, lay, tol, want_abs=None):
    """Per tensor: max-abs-diff / max-abs < tol.  The mean gradient of a large batch with random-sign advantages is a
    heavily cancelling sum (|result| ~ |terms| / sqrt(B)): where the fp64 oracle's sums over |per-sample term|
    (want_abs, ``abs_sums=True``) are given, the bar is scaled by the summation's condition number
    kappa = max_j sum_i |t_ij| / max_j |sum_i t_ij| -- the forward error bound of any summation in working precision."""
    for k, o, s in P.items(lay):
        sz = int(np.prod(s))
        g, w = got[o:o + sz], want[o:o + sz]
        kappa = 1.0 if want_abs is None else max(1.0, float(np.abs(want_abs[o:o + sz]).max() / max(np.abs(w).max(), 1e-30)))
        bar = tol * kappa
        assert rel_err(g, w) < bar, "%s: rel err %.3g (bar %.3g, kappa %.1f)" % (k, rel_err(g, w), bar, kappa)
    used = np.zeros(lay["_size"], bool)
    for k, o, s in P.items(lay):
        used[o:o + int(np.prod(s))] = True
    assert not got[~used].any()          # alignment padding carries zero gradient


@pytest.mark.parametrize("rows,n", [(1, 64), (1, 50), (2, 333), (20, 512), (3, 20000)])
@pytest.mark.parametrize("weights", ["demo", "random"])
def test_gauss_loss_backward_matches_oracle(lib, dev, rows, n, weights):
    actor, critic = demo_weights() if weights == "demo" else random_weights(3, 1, True, 21)
    if weights == "demo":
        critic = critic * np.float32(0.05)      # the shipped critic outputs ~-250: keep the targets' scale sane
    lay, clay = P.actor_layout(3, 1, True), P.critic_layout(3)
    rng = np.random.default_rng(rows + n)
    B = rows * n
    th = rng.uniform(-np.pi, np.pi, (rows, n))
    obs = np.stack([np.cos(th), np.sin(th), rng.uniform(-8, 8, (rows, n))], axis=-1).astype(np.float32)

    def draw(k):
        t2 = rng.uniform(-np.pi, np.pi, k)
        return np.stack([np.cos(t2), np.sin(t2), rng.uniform(-8, 8, k)], axis=-1)
    obs = redraw_near_kinks(actor, lay, critic, clay, obs.reshape(B, 3), draw).reshape(rows, n, 3)
    acs = rng.uniform(-2.5, 2.5, (rows, n, 1)).astype(np.float32)
    adv = rng.normal(size=(rows, n)).astype(np.float32)
    lp_old = (P.actor_logp(actor, lay, obs.reshape(B, 3), acs.reshape(B, 1), True) +
              rng.normal(size=B).astype(np.float32) * 0.1).reshape(rows, n)
    targets = rng.normal(size=(rows, n)).astype(np.float32) * 3
    beta, gam = 0.7, 0.03
    ga, gc, stats = backward_gpu(lib, dev, True, actor, critic, 3, 1, obs, acs, adv, lp_old, targets, beta, gam)
    wa = P.actor_loss_grads(actor, lay, obs.reshape(B, 3), acs.reshape(B, 1), adv.reshape(B), lp_old.reshape(B), beta,
                            gam, True, 2.0, clip=None, dtype=np.float64, abs_sums=True)
    wc = P.critic_loss_grads(critic, clay, obs.reshape(B, 3), targets.reshape(B), dtype=np.float64, abs_sums=True)
    check_grads(ga, wa["grad_raw"], lay, 2 * TOL, wa["grad_abs"])
    check_grads(gc, wc["grad"], clay, 2 * TOL, wc["grad_abs"])
    np.testing.assert_allclose(stats[3], wa["loss"], rtol=2 * TOL, atol=1e-6)
    np.testing.assert_allclose(stats[4], wc["loss"], rtol=2 * TOL)
    np.testing.assert_allclose(stats[5], wa["kl"], rtol=2 * TOL)
    np.testing.assert_allclose(stats[6], wa["ent"], rtol=2 * TOL)


def test_gauss_backward_applies_advantage_normalisation(lib, dev):
    actor, critic = random_weights(3, 1, True, 4)
    lay = P.actor_layout(3, 1, True)
    rng = np.random.default_rng(0)
    rows, n = 4, 300
    B = rows * n
    obs = rng.normal(size=(rows, n, 3)).astype(np.float32)
    acs = rng.normal(size=(rows, n, 1)).astype(np.float32)
    adv = (rng.normal(size=(rows, n)) * 5 + 2).astype(np.float32)
    lp_old = rng.normal(size=(rows, n)).astype(np.float32) - 2
    mean, inv = adv.mean(dtype=np.float64), 1.0 / (1e-8 + adv.std(dtype=np.float64))
    ga, _, _ = backward_gpu(lib, dev, True, actor, critic, 3, 1, obs, acs, adv, lp_old, adv * 0, 1.0, 0.01,
                            adv_stats=[mean, inv])
    advn = P.normalize_adv(adv.reshape(B))
    wa = P.actor_loss_grads(actor, lay, obs.reshape(B, 3), acs.reshape(B, 1), advn, lp_old.reshape(B), 1.0, 0.01, True,
                            2.0, clip=None, dtype=np.float64)
    check_grads(ga, wa["grad_raw"], lay, 3 * TOL)


@pytest.mark.parametrize("ob,n_act", [(4, 2), (6, 3)])
@pytest.mark.parametrize("rows,n", [(1, 100), (5, 1000)])
def test_cat_loss_backward_matches_oracle(lib, dev, ob, n_act, rows, n):
    actor, critic = random_weights(ob, n_act, False, 31, bias_scale=0.3)
    lay, clay = P.actor_layout(ob, n_act, False), P.critic_layout(ob)
    o, s = lay["Wh"]
    actor[o:o + s[0] * s[1]] *= 5
    rng = np.random.default_rng(rows * n + ob)
    B = rows * n
    obs = rng.normal(size=(rows, n, ob)).astype(np.float32)
    obs = redraw_near_kinks(actor, lay, critic, clay, obs.reshape(B, ob), lambda k: rng.normal(size=(k, ob))).reshape(rows, n, ob)
    acs = rng.integers(0, n_act, (rows, n)).astype(np.int32)
    adv = rng.normal(size=(rows, n)).astype(np.float32)
    lp_old = (P.actor_logp(actor, lay, obs.reshape(B, ob), acs.reshape(B), False) +
              rng.normal(size=B).astype(np.float32) * 0.1).reshape(rows, n)
    targets = rng.normal(size=(rows, n)).astype(np.float32)
    ga, gc, stats = backward_gpu(lib, dev, False, actor, critic, ob, n_act, obs, acs, adv, lp_old, targets, 1.0, 0.01)
    wa = P.actor_loss_grads(actor, lay, obs.reshape(B, ob), acs.reshape(B), adv.reshape(B), lp_old.reshape(B), 1.0,
                            0.01, False, clip=None, dtype=np.float64)
    wc = P.critic_loss_grads(critic, clay, obs.reshape(B, ob), targets.reshape(B), dtype=np.float64)
    check_grads(ga, wa["grad_raw"], lay, 2 * TOL)
    check_grads(gc, wc["grad"], clay, 2 * TOL)
    np.testing.assert_allclose(stats[3], wa["loss"], rtol=2 * TOL, atol=1e-6)
    np.testing.assert_allclose(stats[4], wc["loss"], rtol=2 * TOL)


def test_backward_is_deterministic(lib, dev):
    actor, critic = random_weights(3, 1, True, 8)
    rng = np.random.default_rng(9)
    rows, n = 20, 4096
    obs = rng.normal(size=(rows, n, 3)).astype(np.float32)
    acs = rng.normal(size=(rows, n, 1)).astype(np.float32)
    adv, lp_old, tg = rng.normal(size=(3, rows, n)).astype(np.float32)
    r1 = backward_gpu(lib, dev, True, actor, critic, 3, 1, obs, acs, adv, lp_old, tg, 1.0, 0.01)
    r2 = backward_gpu(lib, dev, True, actor, critic, 3, 1, obs, acs, adv, lp_old, tg, 1.0, 0.01)
    np.testing.assert_array_equal(r1[0], r2[0])
    np.testing.assert_array_equal(r1[1], r2[1])


def test_adam_apply_clip_and_bias_correction(lib, dev):
    rng = np.random.default_rng(0)
    na, nc = 25352, 50564
    th_a, th_c = rng.normal(size=na).astype(np.float32), rng.normal(size=nc).astype(np.float32)
    ma, va, mc, vc = np.zeros(na, np.float32), np.zeros(na, np.float32), np.zeros(nc, np.float32), np.zeros(nc, np.float32)
    A, C = dev_t(th_a, dev), dev_t(th_c, dev)
    MA, VA, MC, VC = (dev_t(x, dev) for x in (ma, va, mc, vc))
    opt = make_opt(lib, dev, actor_lr=0.005, critic_lr=0.001, clip=0.1)
    for t in range(1, 4):
        ga = (rng.normal(size=na) * 0.2).astype(np.float32)
        gc = (rng.normal(size=nc) * 3).astype(np.float32)
        world = 2.0
        GA, GC = dev_t(ga * world, dev), dev_t(gc * world, dev)
        lib.call("acx_adam_apply", lib.ptr(A), lib.ptr(GA), lib.ptr(MA), lib.ptr(VA), na, lib.ptr(C),
                 lib.ptr(GC), lib.ptr(MC), lib.ptr(VC), nc, lib.ptr(opt), 1.0 / world, lib.cur_stream())
        lib.call("acx_adam_tick", lib.ptr(opt), lib.cur_stream())
        th_a, ma, va = P.adam_step(th_a, np.clip(ga, -0.1, 0.1), ma, va, t, 0.005)
        th_c, mc, vc = P.adam_step(th_c, gc, mc, vc, t, 0.001)
        torch.cuda.synchronize()
        np.testing.assert_allclose(A.cpu().numpy(), th_a, rtol=1e-5, atol=1e-7)
        np.testing.assert_allclose(C.cpu().numpy(), th_c, rtol=1e-5, atol=1e-7)
        np.testing.assert_allclose(MA.cpu().numpy(), ma, rtol=1e-6, atol=1e-9)
        np.testing.assert_allclose(VC.cpu().numpy(), vc, rtol=1e-6, atol=1e-9)
    assert opt[12].item() == 3.0


@pytest.mark.parametrize("kl,lr,want", [(0.0001, 0.005, 0.0075), (0.1, 0.005, 0.005 / 1.5), (0.002, 0.005, 0.005),
                                        (0.0, 0.09, 0.1), (1.0, 1.2e-6, 1e-6)])
def test_kl_lr_rule_and_schedule(lib, dev, kl, lr, want):
    opt = make_opt(lib, dev, actor_lr=lr, schedule=1)
    stats = torch.zeros(16, device=dev)
    stats[7] = kl
    lib.call("acx_kl_lr_update", lib.ptr(opt), lib.ptr(stats), lib.cur_stream())
    torch.cuda.synchronize()
    o = opt.cpu().numpy()
    np.testing.assert_allclose(o[0], np.float32(P.kl_lr_rule(lr, kl, 0.002)), rtol=1e-6)
    np.testing.assert_allclose(o[0], want, rtol=1e-5)
    assert o[10] == 1.0
    np.testing.assert_allclose(o[3], 1e-2, rtol=1e-6)          # iteration 0: gamma = 10^-2, beta = 10^0
    np.testing.assert_allclose(o[2], 1.0, rtol=1e-6)
    # iteration 1500: schedules are mid-way (A3C/process.py:173-174)
    from oracle.pathadv import linear_schedule_val
    opt[10] = 1500.0
    lib.call("acx_kl_lr_update", lib.ptr(opt), lib.ptr(stats), lib.cur_stream())
    torch.cuda.synchronize()
    o = opt.cpu().numpy()
    np.testing.assert_allclose(o[3], 10.0 ** linear_schedule_val(1500, 100, 3000, -2, -8), rtol=1e-5)
    np.testing.assert_allclose(o[2], 10.0 ** linear_schedule_val(1500, 100, 3000, 0, -4), rtol=1e-5)


def test_basic_ac_critic_head_backward(lib, dev):
    """Basic_AC/policies.py:134-135: the value head reads layer 2; layer 3 gets no gradient."""
    _, critic = random_weights(4, 2, False, 77, bias_scale=0.3)
    clay = P.critic_layout(4)
    rng = np.random.default_rng(5)
    rows, n = 3, 700
    B = rows * n
    obs = rng.normal(size=(rows, n, 4)).astype(np.float32)
    obs = redraw_near_kinks(None, None, critic, clay, obs.reshape(B, 4), lambda k: rng.normal(size=(k, 4))).reshape(rows, n, 4)
    tg = rng.normal(size=(rows, n)).astype(np.float32)
    c = dev_t(critic, dev)
    o = dev_t(np.stack([soa(x) for x in obs]), dev)
    t = dev_t(tg, dev)
    g = torch.full_like(c, 3.0)
    stats = torch.zeros(16, device=dev)
    nbytes = lib.load().acx_loss_backward_workspace_bytes(4, 1, 1)
    ws = torch.empty(nbytes, device=dev, dtype=torch.uint8)
    lib.call("acx_loss_backward_basic_critic", lib.ptr(c), 4, rows, n, lib.ptr(o), lib.ptr(t), lib.ptr(g), lib.ptr(stats),
             lib.ptr(ws), nbytes, lib.cur_stream())
    torch.cuda.synchronize()
    want = P.critic_loss_grads(critic, clay, obs.reshape(B, 4), tg.reshape(B), basic_head=True, dtype=np.float64, abs_sums=True)
    got = g.cpu().numpy()
    for name in ("W3", "b3"):
        o3, s3 = clay[name]
        assert not got[o3:o3 + int(np.prod(s3))].any()
    for k, off, shp in P.items(clay):
        if k in ("W3", "b3"):
            continue
        sz = int(np.prod(shp))
        w = want["grad"][off:off + sz]
        kappa = max(1.0, float(np.abs(want["grad_abs"][off:off + sz]).max() / np.abs(w).max()))
        assert rel_err(got[off:off + sz], w) < 2 * TOL * kappa, (k, kappa)
    np.testing.assert_allclose(stats[4].item(), want["loss"], rtol=2 * TOL)
