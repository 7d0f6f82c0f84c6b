"""numpy restatement of the reference's policy / value networks, losses, gradients,
Adam update and KL-driven learning-rate rule.  TEST INFRASTRUCTURE (see
``oracle/__init__``).  **Parity unpinned**: the arithmetic lives in TensorFlow 1.3
(``README.md:6``; un-vendored, not installable here) and the reference has no tests;
this follows the graph the reference builds, line by line.

References (``/root/reference``):
  actor graph            A3C/policies.py:42-72   (Basic_AC/policies.py:41-70)
  actor loss + clip      A3C/policies.py:74-85   (Basic_AC: clip +-1, policies.py:81)
  critic graph + loss    A3C/policies.py:142-158 (Basic_AC head reads x1: policies.py:134-135)
  Adam                   tf.train.AdamOptimizer defaults, A3C/policies.py:82,155
  post-update KL + lr    A3C/policies.py:106-108, A3C/process.py:262-270

Flat parameter layout (floats; every tensor starts on a 4-float boundary so the CUDA
side can move it with 16-byte bulk copies).  Kernels are ``[in, out]`` row-major as in
the TF checkpoint.
"""
import numpy as np

A_H1, A_H2, A_H3 = 128, 128, 64          # A3C/policies.py:44-46
C_H1, C_H2, C_H3 = 256, 128, 128         # A3C/policies.py:146-148
LOG_STD_MIN, LOG_STD_MAX = -2.5, 2.5     # A3C/policies.py:53
HALF_LOG_2PI = 0.9189385332046727


def _al4(x):
    return (x + 3) // 4 * 4


def actor_layout(ob_dim, num_ac, cont):
    """name -> (offset, shape); plus '_size'."""
    lay, off = {}, 0
    for name, shape in (("W1", (ob_dim, A_H1)), ("b1", (A_H1,)), ("W2", (A_H1, A_H2)), ("b2", (A_H2,)),
                        ("W3", (A_H2, A_H3)), ("b3", (A_H3,)), ("Wh", (A_H3, num_ac)), ("bh", (num_ac,))) + \
            ((("log_std", (num_ac,)),) if cont else ()):
        lay[name] = (off, shape)
        off = _al4(off + int(np.prod(shape)))
    lay["_size"] = off
    return lay


def critic_layout(ob_dim):
    lay, off = {}, 0
    for name, shape in (("W1", (ob_dim, C_H1)), ("b1", (C_H1,)), ("W2", (C_H1, C_H2)), ("b2", (C_H2,)),
                        ("W3", (C_H2, C_H3)), ("b3", (C_H3,)), ("Wv", (C_H3, 1)), ("bv", (1,))):
        lay[name] = (off, shape)
        off = _al4(off + int(np.prod(shape)))
    lay["_size"] = off
    return lay


def items(lay):
    """(name, offset, shape) of every tensor of a layout."""
    return [(k, v[0], v[1]) for k, v in lay.items() if k != "_size"]


def unpack(flat, lay):
    return {k: np.asarray(flat[o:o + int(np.prod(s))]).reshape(s) for k, o, s in items(lay)}


def pack(tensors, lay, dtype=np.float32):
    flat = np.zeros(lay["_size"], dtype=dtype)
    for k, o, s in items(lay):
        flat[o:o + int(np.prod(s))] = np.asarray(tensors[k], dtype=dtype).reshape(-1)
    return flat


def glorot_init(lay, rng, dtype=np.float32):
    """tf.contrib.layers.xavier_initializer (uniform, limit sqrt(6/(fan_in+fan_out))),
    zero biases, log_std = 0 (A3C/policies.py:7,44-52).  Used for synthetic weights."""
    t = {}
    for k, o, s in items(lay):
        if len(s) == 2:
            lim = np.sqrt(6.0 / (s[0] + s[1]))
            t[k] = rng.uniform(-lim, lim, size=s)
        else:
            t[k] = np.zeros(s)
    return pack(t, lay, dtype)


# ------------------------------------------------------------------ forward
def lrelu(x, f):
    # A3C/policies.py:9-10: (1-alpha)*relu(x) + alpha*x, evaluated in that form
    return (f(0.8) * np.maximum(x, f(0.0)) + f(0.2) * x).astype(f)


def actor_hidden(P, obs, f):
    x1 = lrelu(obs.astype(f) @ P["W1"].astype(f) + P["b1"].astype(f), f)
    x2 = lrelu(x1 @ P["W2"].astype(f) + P["b2"].astype(f), f)
    x3 = lrelu(x2 @ P["W3"].astype(f) + P["b3"].astype(f), f)
    return x1, x2, x3


def gauss_head(P, x3, ac_scale, f):
    """mu = ac_scale * tanh(dense) (policies.py:50); log_std clipped (:53)."""
    pre = x3 @ P["Wh"].astype(f) + P["bh"].astype(f)
    mu = (np.tanh(pre) * f(ac_scale)).astype(f)
    log_std = np.clip(P["log_std"].astype(f), f(LOG_STD_MIN), f(LOG_STD_MAX))
    return mu, log_std, pre


def gauss_logp(ac, mu, log_std, f):
    """sum_d Normal(mu, std).log_prob(ac) (policies.py:57,60): -0.5 z^2 - log std - 0.5 log 2pi."""
    std = np.exp(log_std)
    z = (ac.astype(f) - mu) / std
    return np.sum(f(-0.5) * z * z - log_std - f(HALF_LOG_2PI), axis=1).astype(f)


def gauss_entropy(log_std, n, f):
    """sum_d (0.5 + 0.5 log 2pi + log std) (policies.py:58), same for every row."""
    return np.full(n, np.sum(f(0.5) + f(HALF_LOG_2PI) + log_std), dtype=f)


def cat_head(P, x3, f):
    logits = (x3 @ P["Wh"].astype(f) + P["bh"].astype(f)).astype(f)     # policies.py:63
    l = logits + f(1e-8)                                                # policies.py:65
    m = l.max(axis=1, keepdims=True)
    logps = (l - (m + np.log(np.sum(np.exp(l - m), axis=1, keepdims=True)))).astype(f)
    m2 = logits.max(axis=1, keepdims=True)
    e = np.exp(logits - m2)
    p = (e / e.sum(axis=1, keepdims=True)).astype(f)                    # tf.nn.softmax(logits)
    ent = (-np.sum(p * logps, axis=1)).astype(f)                        # policies.py:66
    return logits, logps, p, ent


def cat_sample(logits, u):
    """tf.multinomial (TF 1.x CPU kernel): inverse CDF in double over exp(logit - max);
    action = #{j : cdf_j <= u * cdf_last}.  u in [0,1).  int32 (policies.py:64)."""
    l = logits.astype(np.float64)
    e = np.exp(l - l.max(axis=1, keepdims=True))
    cdf = np.cumsum(e, axis=1)
    thr = np.asarray(u, dtype=np.float64)[:, None] * cdf[:, -1:]
    a = np.sum(cdf <= thr, axis=1)
    return np.minimum(a, logits.shape[1] - 1).astype(np.int32)


def actor_act(flat, lay, obs, noise, cont, ac_scale=2.0, dtype=np.float32):
    """Actor.act on a batch (policies.py:110-115): returns (ac, logp, entropy, head).
    noise = eps[B, ac] standard normals (cont) or u[B] uniforms (disc)."""
    f = dtype
    P = unpack(flat, lay)
    _, _, x3 = actor_hidden(P, obs, f)
    if cont:
        mu, log_std, _ = gauss_head(P, x3, ac_scale, f)
        ac = (mu + np.exp(log_std) * np.asarray(noise).astype(f)).astype(f)      # Normal.sample
        return ac, gauss_logp(ac, mu, log_std, f), gauss_entropy(log_std, len(obs), f), mu
    logits, logps, _, ent = cat_head(P, x3, f)
    ac = cat_sample(logits, noise)
    return ac, logps[np.arange(len(ac)), ac], ent, logits


def actor_logp(flat, lay, obs, acs, cont, ac_scale=2.0, dtype=np.float32):
    """log-prob of stored actions under the current weights (policies.py:60 / :69)."""
    f = dtype
    P = unpack(flat, lay)
    _, _, x3 = actor_hidden(P, obs, f)
    if cont:
        mu, log_std, _ = gauss_head(P, x3, ac_scale, f)
        return gauss_logp(np.asarray(acs).reshape(len(obs), -1), mu, log_std, f)
    _, logps, _, _ = cat_head(P, x3, f)
    return logps[np.arange(len(obs)), np.asarray(acs).astype(np.int64)]


def critic_value(flat, lay, obs, basic_head=False, dtype=np.float32):
    """Critic.value (policies.py:146-150,176-177).  basic_head=True reproduces
    Basic_AC/policies.py:134-135 where the value head reads x1 (second layer)."""
    f = dtype
    P = unpack(flat, lay)
    x = np.maximum(obs.astype(f) @ P["W1"].astype(f) + P["b1"].astype(f), f(0))
    x1 = np.maximum(x @ P["W2"].astype(f) + P["b2"].astype(f), f(0))
    x2 = np.maximum(x1 @ P["W3"].astype(f) + P["b3"].astype(f), f(0))
    h = x1 if basic_head else x2
    return (h @ P["Wv"].astype(f) + P["bv"].astype(f)).reshape(-1).astype(f)


# ------------------------------------------------------------------ losses + analytic gradients
def actor_loss_grads(flat, lay, obs, acs, advs, logp_old, beta, gamma_ent, cont,
                     ac_scale=2.0, clip=0.1, dtype=np.float64, abs_sums=False):
    """loss = -mean(adv*lp) + beta*mean((lp_old-lp)^2) - gamma*mean(entropy)  (policies.py:74-80)
    and its gradient wrt every actor tensor, clipped elementwise to +-clip (policies.py:85).
    Returns dict(loss, rew_loss, kl, ent, grad (flat, clipped), grad_raw (flat)).
    abs_sums=True adds ``grad_abs``: every gradient element with its per-sample terms summed as absolute values
    (sum_i |t_ij| instead of sum_i t_ij) -- the condition number of the batch sum is grad_abs / |grad_raw|."""
    f = dtype
    P = {k: v.astype(f) for k, v in unpack(flat, lay).items()}
    B = len(obs)
    obs = obs.astype(f)
    advs = np.asarray(advs).astype(f)
    logp_old = np.asarray(logp_old).astype(f)
    x1, x2, x3 = actor_hidden(P, obs, f)
    G = {}
    if cont:
        acs = np.asarray(acs).astype(f).reshape(B, -1)
        mu, log_std, pre = gauss_head(P, x3, ac_scale, f)
        lp = gauss_logp(acs, mu, log_std, f)
        ent = gauss_entropy(log_std, B, f)
    else:
        acs = np.asarray(acs).astype(np.int64)
        logits, logps, p, ent = cat_head(P, x3, f)
        lp = logps[np.arange(B), acs]
    diff = logp_old - lp
    rew_loss = -np.mean(advs * lp)
    kl = np.mean(diff * diff)
    loss = rew_loss + f(beta) * kl - f(gamma_ent) * np.mean(ent)
    dlp = (-advs - f(2.0) * f(beta) * diff) / f(B)                # dL/dlp_i
    dent = np.full(B, -f(gamma_ent) / f(B))                       # dL/dent_i
    if cont:
        std = np.exp(log_std)
        z = (acs - mu) / std
        dmu = dlp[:, None] * z / std                              # d lp / d mu = (a-mu)/std^2
        inside = (P["log_std"] >= f(LOG_STD_MIN)) & (P["log_std"] <= f(LOG_STD_MAX))   # clip_by_value grad
        dls = np.sum(dlp[:, None] * (z * z - f(1.0)) + dent[:, None], axis=0)
        G["log_std"] = np.where(inside, dls, f(0.0))
        t = mu / f(ac_scale)
        dhead = dmu * f(ac_scale) * (f(1.0) - t * t)              # through ac_scale * tanh
    else:
        onehot = np.zeros_like(logps)
        onehot[np.arange(B), acs] = 1
        pl = np.exp(logps)                                        # softmax(logits + 1e-8)
        dlogps = dlp[:, None] * onehot - dent[:, None] * p        # ent = -sum p*logps
        dlogits = dlogps - pl * dlogps.sum(axis=1, keepdims=True)  # through log_softmax
        dp = -dent[:, None] * logps                               # d ent / d p = -logps
        dlogits = dlogits + p * (dp - np.sum(dp * p, axis=1, keepdims=True))  # through softmax
        dhead = dlogits
    G["Wh"] = x3.T @ dhead
    G["bh"] = dhead.sum(axis=0)
    A = {}
    if abs_sums:
        A["Wh"], A["bh"] = np.abs(x3).T @ np.abs(dhead), np.abs(dhead).sum(axis=0)
        if cont:
            A["log_std"] = np.sum(np.abs(dlp[:, None] * (z * z - f(1.0))) + np.abs(dent[:, None]), axis=0)

    def back(dout, x_out, x_in, W, name=None):
        dpre = dout * np.where(x_out > 0, f(1.0), f(0.2))         # lrelu': 0.8*[x>0] + 0.2
        if abs_sums:
            A["W" + name], A["b" + name] = np.abs(x_in).T @ np.abs(dpre), np.abs(dpre).sum(axis=0)
        return x_in.T @ dpre, dpre.sum(axis=0), dpre @ W.T

    dx3 = dhead @ P["Wh"].T
    G["W3"], G["b3"], dx2 = back(dx3, x3, x2, P["W3"], "3")
    G["W2"], G["b2"], dx1 = back(dx2, x2, x1, P["W2"], "2")
    G["W1"], G["b1"], _ = back(dx1, x1, obs, P["W1"], "1")
    raw = pack(G, lay, dtype=f)
    out = dict(loss=loss, rew_loss=rew_loss, kl=kl, ent=np.mean(ent), lp=lp,
               grad_raw=raw, grad=np.clip(raw, -f(clip), f(clip)) if clip is not None else raw)
    if abs_sums:
        out["grad_abs"] = pack(A, lay, dtype=f)
    return out


def critic_loss_grads(flat, lay, obs, targets, basic_head=False, dtype=np.float64, abs_sums=False):
    """loss = mean((v - target)^2), gradients unclipped (policies.py:152-158).  abs_sums: see actor_loss_grads."""
    f = dtype
    P = {k: v.astype(f) for k, v in unpack(flat, lay).items()}
    B = len(obs)
    obs = obs.astype(f)
    tg = np.asarray(targets).astype(f)
    x = np.maximum(obs @ P["W1"] + P["b1"], 0)
    x1 = np.maximum(x @ P["W2"] + P["b2"], 0)
    x2 = np.maximum(x1 @ P["W3"] + P["b3"], 0)
    h = x1 if basic_head else x2
    v = (h @ P["Wv"] + P["bv"]).reshape(-1)
    err = v - tg
    loss = np.mean(err * err)
    dv = (f(2.0) * err / f(B))[:, None]
    G = {"Wv": h.T @ dv, "bv": dv.sum(axis=0)}
    A = {"Wv": np.abs(h).T @ np.abs(dv), "bv": np.abs(dv).sum(axis=0)} if abs_sums else {}
    dh = dv @ P["Wv"].T

    def back(dout, x_out, x_in, W, name=None):
        dpre = dout * (x_out > 0)
        if abs_sums:
            A["W" + name], A["b" + name] = np.abs(x_in).T @ np.abs(dpre), np.abs(dpre).sum(axis=0)
        return x_in.T @ dpre, dpre.sum(axis=0), dpre @ W.T

    if basic_head:
        G["W3"], G["b3"] = np.zeros_like(P["W3"]), np.zeros_like(P["b3"])   # dead third layer
        if abs_sums:
            A["W3"], A["b3"] = np.zeros_like(P["W3"]), np.zeros_like(P["b3"])
        dx1 = dh
    else:
        G["W3"], G["b3"], dx1 = back(dh, x2, x1, P["W3"], "3")
    G["W2"], G["b2"], dx = back(dx1, x1, x, P["W2"], "2")
    G["W1"], G["b1"], _ = back(dx, x, obs, P["W1"], "1")
    out = dict(loss=loss, v=v, grad=pack(G, lay, dtype=f))
    if abs_sums:
        out["grad_abs"] = pack(A, lay, dtype=f)
    return out


# ------------------------------------------------------------------ optimiser + lr rule
def adam_step(theta, g, m, v, t, lr, beta1=0.9, beta2=0.999, eps=1e-8, dtype=np.float32):
    """TF 1.x ApplyAdam: lr_t = lr*sqrt(1-b2^t)/(1-b1^t); m += (g-m)(1-b1); v += (g^2-v)(1-b2);
    theta -= lr_t*m/(sqrt(v)+eps).  t counts from 1."""
    f = dtype
    theta, g, m, v = (np.asarray(a).astype(f) for a in (theta, g, m, v))
    b1p, b2p = f(1.0), f(1.0)
    for _ in range(int(t)):          # TF keeps beta^t in slot variables, multiplied once per apply
        b1p, b2p = f(b1p * f(beta1)), f(b2p * f(beta2))
    lr_t = f(f(lr) * f(np.sqrt(f(1.0) - b2p) / (f(1.0) - b1p)))
    m = m + (g - m) * (f(1.0) - f(beta1))
    v = v + (g * g - v) * (f(1.0) - f(beta2))
    theta = theta - (m * lr_t) / (np.sqrt(v) + f(eps))
    return theta.astype(f), m.astype(f), v.astype(f)


def kl_lr_rule(lr, kl, desired_kl, min_lr=1e-6, max_lr=0.1):
    """A3C/process.py:262-270 (max_lr=.1, :12) / Basic_AC/run_AC.py:261-266 (max_lr=1., :178)."""
    if kl < desired_kl / 4:
        return min(max_lr, lr * 1.5)
    if kl > desired_kl * 4:
        return max(min_lr, lr / 1.5)
    return lr


def normalize_adv(advs, dtype=np.float64):
    """A3C/process.py:247: (adv - mean)/(1e-8 + std), population std, whole update batch."""
    a = np.asarray(advs).astype(dtype)
    return (a - np.mean(a)) / (dtype(1e-8) + np.std(a))
