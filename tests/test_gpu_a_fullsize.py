"""Parity at the sizes the benchmark runs (BASELINE.json configs[2] and configs[3]): the tcgen05 forward passes (K2) and
the tcgen05 loss+backward (K4: chain kernel + weight-gradient kernels) through the C ABI against the **fp64 oracle**
(oracle/policies.py, the restatement of A3C/policies.py:44-85,142-158), evaluated in chunks of 65,536 samples -- the
loss and its gradient are means over the batch, hence sums of per-chunk means weighted by chunk size.

Bars (north_star: 1e-5 relative, fp32):
  * K2 values / log-probs: |got - want| <= 1e-5 * (|want| + rms(want)) per element;
  * K4 gradients, two batches per size:
      "one-signed"  advantages with the sign of (action - mean) (categorical: of the chosen category) and targets above
                    every value, so that the per-sample terms of a gradient element do not cancel: max|got - want| <= 1e-5 * max|want| per parameter tensor -- the well-conditioned case,
                    where the bar measures the kernels;
      "zero-mean"   advantages ~ N(0, 1) as after the batch normalisation of A3C/process.py:247: every gradient element is
                    the residual of 1.3 M (5.2 M) cancelling terms, ~1/sqrt(B) of their magnitude.  No fp32 evaluation
                    meets 1e-5 of THAT (the numpy-fp32 restatement of the reference's own math sits at 1e-4..4e-4: both
                    the rounding of the terms and the handful of samples whose pre-activation lies within rounding of the
                    ReLU / leaky-ReLU kink, where the derivative jumps, cost about one per-sample term each).  The bar
                    is the summation's own condition number kappa = max_j sum_i |t_ij| / max_j |sum_i t_ij| (from the
                    fp64 oracle, printed): max|got - want| <= 1e-5 * kappa * max|want|; the numpy-fp32 oracle's own
                    error is printed beside it (numpy sums pairwise, the kernels' per-thread partial sums are
                    sequential: on the actor's head gradients they sit 2-5x above numpy's, on everything else within 3x);
    plus the absolute rounding floor 3e-6 / sqrt(B) that matters only for the one-element tensors (see
    tests/test_gpu_backward_tc.py::check_grads);
  * the tensor-core accumulate truncates: the weight-gradient kernel flushes its tensor-memory accumulators into fp32
    scratch every 2 tiles; its product is checked against an fp64 product of the very operands it read, on
    cancelling AND on one-signed data (where the truncation bias adds up instead of averaging out)
    (was tests/tc/run_wgrad_accuracy.py) so that a longer flush interval cannot slip through.
The measured errors are printed (pytest -s / the GPU log).

This file sorts before the long-running GPU tests on purpose."""
import time

import numpy as np
import pytest
import torch

from oracle import policies as P
from gpu_util import dev_t, rel_err
import test_gpu_backward_tc as TB

pytestmark = pytest.mark.gpu
CHUNK = 65536
BAR = 1e-5


def chunked_oracle(k, dtype, abs_sums=False):
    """fp64 (or fp32) oracle loss + gradients of Case k, accumulated over 65,536-sample chunks.  abs_sums: the same
    sums over |per-sample term| ride along in sums["abs_a"], sums["abs_c"]."""
    B, ob = k.B, k.ob
    obs = k.obs.reshape(B, ob)
    acs = k.acs.reshape(B, k.num_ac) if k.cont else k.acs.reshape(B)
    adv, lp_old, tg = k.adv.reshape(B), k.lp_old.reshape(B), k.targets.reshape(B)
    ga = np.zeros(k.lay["_size"], np.float64)
    gc = np.zeros(k.clay["_size"], np.float64)
    sums = dict(loss_a=0.0, kl=0.0, ent=0.0, loss_c=0.0, rew=0.0, abs_a=np.zeros_like(ga), abs_c=np.zeros_like(gc))
    lps, vs = [], []
    for lo in range(0, B, CHUNK):
        sl = slice(lo, min(B, lo + CHUNK))
        w = (sl.stop - sl.start) / B
        a = P.actor_loss_grads(k.actor, k.lay, obs[sl], acs[sl], adv[sl], lp_old[sl], k.beta, k.gam, k.cont, 2.0,
                               clip=None, dtype=dtype, abs_sums=abs_sums)
        c = P.critic_loss_grads(k.critic, k.clay, obs[sl], tg[sl], dtype=dtype, abs_sums=abs_sums)
        ga += w * a["grad_raw"].astype(np.float64)
        gc += w * c["grad"].astype(np.float64)
        if abs_sums:
            sums["abs_a"] += w * a["grad_abs"].astype(np.float64)
            sums["abs_c"] += w * c["grad_abs"].astype(np.float64)
        sums["loss_a"] += w * float(a["loss"]); sums["kl"] += w * float(a["kl"]); sums["ent"] += w * float(a["ent"])
        sums["rew"] += w * float(a["rew_loss"]); sums["loss_c"] += w * float(c["loss"])
        lps.append(np.asarray(a["lp"], dtype=np.float64)); vs.append(np.asarray(c["v"], dtype=np.float64))
    return ga, gc, sums, np.concatenate(lps), np.concatenate(vs)


def fill_caches_with_the_tensor_core_forwards(k):
    """What the trainer does: the rollout's policy forward and the value pass on tcgen05 write the activation caches
    (and the critic's sign masks) the backward reads.  Returns (V, logp) of those passes."""
    lib, dev, s = k.lib, k.dev, k.lib.cur_stream()
    pf_a = torch.zeros(lib.load().acx_tc_packed_floats(0, k.ob, k.num_ac, int(k.cont)), device=dev)
    lib.call("acx_tc_pack_weights", 0, lib.ptr(k.a), k.ob, k.num_ac, int(k.cont), lib.ptr(pf_a), s)
    k.critic_forward_tc_with_masks()
    pf_c = torch.zeros(lib.load().acx_tc_packed_floats(1, k.ob, 1, 0), device=dev)
    lib.call("acx_tc_pack_weights", 1, lib.ptr(k.c), k.ob, 1, 0, lib.ptr(pf_c), s)
    v = torch.empty((k.rows, k.n), device=dev)
    lib.call("acx_value_forward_tc", lib.ptr(k.c), lib.ptr(pf_c), k.ob, k.rows, k.n, lib.ptr(k.o), lib.ptr(v), 0,
             lib.ptr(k.cx2), lib.ptr(k.cx3), lib.ptr(k.cmask), s)
    # log-prob of the stored actions under the same weights (fills nothing, K2's third entry point)
    lp = torch.empty((k.rows, k.n), device=dev)
    kl_part = torch.zeros(lib.load().acx_kl_workspace_bytes() // 8, device=dev, dtype=torch.float64)
    import ctypes
    nb = ctypes.c_int(0)
    if k.cont:
        lib.call("acx_policy_logp_gauss_tc", lib.ptr(k.a), lib.ptr(pf_a), k.ob, k.num_ac, 2.0, k.rows, k.n, lib.ptr(k.o),
                 lib.ptr(k.acs_d), lib.ptr(lp), None, None, lib.ptr(k.lp_d), lib.ptr(kl_part), ctypes.byref(nb), s)
        # the policy forward on the tensor cores, given the stored actions' noise, rewrites the actor's caches
        eps = torch.zeros((k.rows, k.num_ac, k.n), device=dev)
        ac2, lp2 = torch.empty_like(k.acs_d), torch.empty((k.rows, k.n), device=dev)
        lib.call("acx_policy_forward_sample_gauss_tc", lib.ptr(k.a), lib.ptr(pf_a), k.ob, k.num_ac, 2.0, k.rows, k.n,
                 lib.ptr(k.o), lib.ptr(eps), 0, 0, 0, None, lib.ptr(ac2), lib.ptr(lp2), None, None, lib.ptr(k.ax2),
                 lib.ptr(k.ax3), s)
    else:
        lib.call("acx_policy_logp_cat_tc", lib.ptr(k.a), lib.ptr(pf_a), k.ob, k.num_ac, k.rows, k.n, lib.ptr(k.o),
                 lib.ptr(k.acs_d), lib.ptr(lp), None, None, lib.ptr(k.lp_d), lib.ptr(kl_part), ctypes.byref(nb), s)
        u = torch.full((k.rows, k.n), 0.5, device=dev)
        ac2, lp2 = torch.empty_like(k.acs_d), torch.empty((k.rows, k.n), device=dev)
        lib.call("acx_policy_forward_sample_cat_tc", lib.ptr(k.a), lib.ptr(pf_a), k.ob, k.num_ac, k.rows, k.n,
                 lib.ptr(k.o), lib.ptr(u), 0, 0, 0, None, lib.ptr(ac2), lib.ptr(lp2), None, None, lib.ptr(k.ax2),
                 lib.ptr(k.ax3), s)
    torch.cuda.synchronize()
    # +-1 per sample: the side of the policy's mean (Gaussian: eps = 0 above made ac2 = mu) or the category the stored
    # action lies on -- an advantage of this sign makes every sample push the head the same way
    if k.cont:
        side = np.sign(k.acs_d.cpu().numpy() - ac2.cpu().numpy())[:, 0, :]
    else:
        side = np.where(k.acs_d.cpu().numpy() == 0, 1.0, -1.0)
    return v.cpu().numpy().reshape(-1), lp.cpu().numpy().reshape(-1), side.astype(np.float32)


def per_tensor_errors(got, want, lay):
    out = {}
    for name, o, s in P.items(lay):
        sz = int(np.prod(s))
        g, w = got[o:o + sz].astype(np.float64), want[o:o + sz]
        out[name] = (np.abs(g - w).max(), np.abs(w).max())
    return out


@pytest.mark.parametrize("label,cont,ob,num_ac,rows,n", [
    ("configs[2] Pendulum 20 x 65,536", True, 3, 1, 20, 65536),
    ("configs[3] CartPole 20 x 262,144", False, 4, 2, 20, 262144)])
def test_tcgen05_forward_and_backward_at_the_bench_size_match_the_fp64_oracle(lib, dev, label, cont, ob, num_ac, rows, n):
    k = TB.Case(lib, dev, cont, ob, num_ac, rows, n, 77, beta=1.0, gam=0.01)
    v_got, lp_got, side = fill_caches_with_the_tensor_core_forwards(k)
    floor = 3e-6 / np.sqrt(k.B)
    base_adv, base_tg = k.adv.copy(), k.targets.copy()
    for regime in ("one-signed", "zero-mean"):
        if regime == "one-signed":
            k.adv, k.targets = (np.abs(base_adv) + np.float32(0.1)) * side, base_tg + np.float32(12.0)
        else:
            k.adv, k.targets = base_adv, base_tg
        k.adv_d, k.tg_d = dev_t(k.adv, dev), dev_t(k.targets, dev)
        ga, gc, stats = k.run("tc", critic_masks=True)
        t0 = time.time()
        wa, wc, sums, lp_want, v_want = chunked_oracle(k, np.float64, abs_sums=True)
        t64 = time.time() - t0
        if regime == "one-signed":          # ---- K2 (the forward passes do not depend on the regime)
            for name, got, want in (("V (acx_value_forward_tc)", v_got, v_want),
                                    ("logp (acx_policy_logp_*_tc)", lp_got, lp_want)):
                rms = float(np.sqrt(np.mean(want * want)))
                err = np.abs(got.astype(np.float64) - want) / (np.abs(want) + rms)
                print("%s | %s: max err %.2e of |want| + rms (%d samples)" % (label, name, err.max(), err.size))
                assert err.max() <= BAR, (name, err.max())
            assert (v_want < k.targets.reshape(-1)).all()           # every v - target has the same sign
        else:
            # the numpy-fp32 restatement's own error, for the log (Pendulum size only: one more pass over 5.2 M samples
            # would add a minute to the suite)
            wa32, wc32 = (chunked_oracle(k, np.float32)[:2]) if k.B <= 2_000_000 else (None, None)
        # ---- K4
        failures, worst = [], 0.0
        for net, got, want, lay in (("actor", ga, wa, k.lay), ("critic", gc, wc, k.clay)):
            e = per_tensor_errors(got, want, lay)
            eabs = per_tensor_errors(sums["abs_a"] if net == "actor" else sums["abs_c"], 0 * want, lay)
            if regime == "zero-mean":
                e32 = per_tensor_errors(wa32 if net == "actor" else wc32, want, lay) if wa32 is not None else None
            for name in e:
                err, scale = e[name]
                kappa = eabs[name][0] / scale
                if regime == "one-signed":
                    # log_std: d lp / d log_std = z^2 - 1 has zero mean over the batch whatever the advantages are --
                    # that one element cancels in every regime and gets the condition-number bar
                    strict = name != "log_std"
                    print("%s | %s | %s.%s: cuda %.2e of max |grad| = %.3e; kappa = %.1f" % (
                        label, regime, net, name, err / scale, scale, kappa))
                    ok = err <= BAR * (1.0 if strict else kappa) * scale + floor
                else:
                    print("%s | %s | %s.%s: cuda %.2e, numpy-fp32 oracle %.2e of max |grad| = %.3e; kappa = %.0f" % (
                        label, regime, net, name, err / scale, e32[name][0] / scale if e32 else float("nan"), scale, kappa))
                    ok = err <= BAR * kappa * scale + floor
                if not ok:
                    failures.append((regime, net, name, err / scale))
                worst = max(worst, (err - floor) / scale)
        print("%s | %s | worst K4 tensor error %.2e of its max; fp64 oracle took %.1f s on the host" % (
            label, regime, worst, t64))
        assert not failures, failures
        np.testing.assert_allclose(stats[3], sums["loss_a"], rtol=BAR, atol=1e-6)
        np.testing.assert_allclose(stats[4], sums["loss_c"], rtol=BAR)
        np.testing.assert_allclose(stats[5], sums["kl"], rtol=BAR)
        np.testing.assert_allclose(stats[6], sums["ent"], rtol=BAR)


@pytest.mark.parametrize("one_signed", [False, True])
def test_weight_gradient_kernel_against_fp64_of_its_own_operands(lib, dev, one_signed):
    """acx_wgrad_kernel at 20 x 65,536: dW3 = X2 . dPre3^T and db3 against an fp64 product of the very arrays the
    kernel read (the X2 cache and the dPre3 stream the chain kernel left in the workspace).  The tensor core's fp32
    accumulate truncates once per MMA: with cancelling terms the error was 7.7e-5 without a flush of the tensor-memory
    accumulators and 8.4e-6 with one every 8 tiles; with one-signed terms (targets above every value) the bias adds up:
    2.2e-5 at 8 tiles, 1.1e-5 at 4, 5.4e-6 at the shipped 2.  Bar 1e-5 -- a change of the flush interval has to pass
    here."""
    rows, n = 20, 65536
    k = TB.Case(lib, dev, True, 3, 1, rows, n, 77)
    if one_signed:
        k.targets = k.targets + np.float32(12.0)
        k.tg_d = dev_t(k.targets, dev)
    _, gc, _ = k.run("tc", actor=False)
    _, fc, _ = k.run("ffma", actor=False)
    base = (lib.load().acx_loss_backward_workspace_bytes(3, 1, 1) + 255) & ~255
    split = k.ws[base:].view(torch.float32)
    d3 = split[:rows * n * 128].view(rows, n // 128, 128, 128).double()        # tile-blocked [rows][tiles][feature][128]
    ref_w3 = torch.einsum("rtks,rtjs->kj", k.cx2.double(), d3)                 # dPre already carries 1/B
    ref_b3 = d3.sum(dim=(0, 1, 3))
    o3, ob3 = k.clay["W3"][0], k.clay["b3"][0]
    for name, g, bar in (("tcgen05", gc, 1e-5), ("ffma", fc, 1e-5)):
        w3 = torch.tensor(g[o3:o3 + 128 * 128].reshape(128, 128), device=dev).double()
        b3 = torch.tensor(g[ob3:ob3 + 128], device=dev).double()
        ew = ((w3 - ref_w3).abs().max() / ref_w3.abs().max()).item()
        eb = ((b3 - ref_b3).abs().max() / ref_b3.abs().max()).item()
        print("%s (%s): dW3 %.2e, db3 %.2e of max, against fp64 of the same operands" % (
            name, "one-signed" if one_signed else "cancelling", ew, eb))
        assert ew <= bar and eb <= bar, (name, ew, eb)
