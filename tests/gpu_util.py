"""Shared helpers for the GPU parity tests (test infrastructure)."""
import os

import numpy as np
import torch

from oracle import policies as P

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def demo_weights():
    """Flat actor/critic buffers (oracle layout) from the reference's shipped checkpoint fixture."""
    from actor_critic_algs_on_tensorflow_b200.tf_bundle import read_bundle
    t = read_bundle(os.path.join(GOLDEN, "model-Pendulum_a3c"))
    al, cl = P.actor_layout(3, 1, True), P.critic_layout(3)
    a = {"W1": t["global_actor/first_layer/kernel"], "b1": t["global_actor/first_layer/bias"],
         "W2": t["global_actor/second_layer/kernel"], "b2": t["global_actor/second_layer/bias"],
         "W3": t["global_actor/third_layer/kernel"], "b3": t["global_actor/third_layer/bias"],
         "Wh": t["global_actor/mu_layer/kernel"], "bh": t["global_actor/mu_layer/bias"],
         "log_std": t["global_actor/log_std"]}
    c = {"W1": t["global_critic/first_layer/kernel"], "b1": t["global_critic/first_layer/bias"],
         "W2": t["global_critic/second_layer/kernel"], "b2": t["global_critic/second_layer/bias"],
         "W3": t["global_critic/third_layer/kernel"], "b3": t["global_critic/third_layer/bias"],
         "Wh": t["global_critic/value/kernel"], "bh": t["global_critic/value/bias"]}
    c["Wv"], c["bv"] = c.pop("Wh"), c.pop("bh")
    return P.pack(a, al), P.pack(c, cl)


def random_weights(ob_dim, num_ac, cont, seed, bias_scale=0.1):
    rng = np.random.default_rng(seed)
    al, cl = P.actor_layout(ob_dim, num_ac, cont), P.critic_layout(ob_dim)
    a, c = P.glorot_init(al, rng), P.glorot_init(cl, rng)
    # non-zero biases / log_std so that every gradient path is exercised
    for lay, flat in ((al, a), (cl, c)):
        for k, o, s in P.items(lay):
            if len(s) == 1:
                flat[o:o + s[0]] = rng.normal(size=s) * bias_scale
    return a, c


def soa(x):
    """[B, d] -> [d, B] contiguous float32 (one 'row' of the [rows][d][n] batch layout)."""
    return np.ascontiguousarray(np.asarray(x, dtype=np.float32).T)


def dev_t(x, dev, dtype=None):
    t = torch.as_tensor(np.ascontiguousarray(x), device=dev)
    return t if dtype is None else t.to(dtype)


def rel_err(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-30)


def act_cache(lib, dev, rows, features, n):
    """Zeroed activation-cache array in the library's tile-blocked layout, viewed [rows][tiles][features][128]."""
    total = lib.load().acx_activation_cache_floats(rows, features, n)
    tiles = (n + 127) // 128
    assert total == rows * tiles * features * 128
    return torch.zeros(total, device=dev).view(rows, tiles, features, 128)


def near_kink(actor, alay, critic, clay, obs2d, margin=2e-5):
    """Samples with a hidden pre-activation of either network within `margin` of zero (fp64 forward).  There the
    derivative of ReLU / leaky ReLU jumps, and which side an fp32 forward lands on is decided by rounding: the gradient
    of such a sample is not determined at fp32 precision (a flip moves a bias gradient by ~1 / batch of a term).  The
    gradient parity tests draw their batches without such samples, so that their bars measure arithmetic, not luck."""
    bad = np.zeros(len(obs2d), bool)
    for flat, lay, slope in ((actor, alay, 0.2), (critic, clay, 0.0)):
        if flat is None:
            continue
        W = {k: v.astype(np.float64) for k, v in P.unpack(flat, lay).items()}
        x = obs2d.astype(np.float64)
        for w, b in (("W1", "b1"), ("W2", "b2"), ("W3", "b3")):
            pre = x @ W[w] + W[b]
            bad |= (np.abs(pre) < margin).any(axis=1)
            x = np.where(pre > 0, pre, slope * pre)
    return bad


def redraw_near_kinks(actor, alay, critic, clay, obs2d, draw, margin=2e-5):
    """obs2d with every near-kink sample (see near_kink) replaced by fresh draws: draw(k) -> [k, ob] float32."""
    obs2d = np.array(obs2d, dtype=np.float32)
    for _ in range(50):
        bad = near_kink(actor, alay, critic, clay, obs2d, margin)
        if not bad.any():
            return obs2d
        obs2d[bad] = draw(int(bad.sum())).astype(np.float32)
    raise AssertionError("could not draw a batch away from the activation kinks")
