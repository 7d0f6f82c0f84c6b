"""``Actor`` / ``Critic`` with the reference's interface (``A3C/policies.py:34-182``; the
``Basic_AC`` constructor spellings are accepted too), backed by the CUDA library.

What is kept: constructor arguments and defaults, method names, argument meaning, return shapes
(``act`` returns row 0 only, ``optimize`` returns ``[loss, None]``, ``value`` returns a 1-D array,
``set_opt_param`` returns ``[lr, beta, gamma]``), and the global/local pairing of A3C
(``global_actor=...`` makes ``optimize`` apply the *local* gradient to the *global* parameters with the
local Adam moments, ``sync_w_global`` copies global -> local).  The ``sess`` argument is accepted and
ignored: there is no TensorFlow session; state lives in device buffers owned by the object.

What is different on purpose: arrays go to the GPU in one batch, so ``act`` / ``value`` / ``optimize``
accept any batch size; ``act_batch`` returns all rows.  Action noise comes from ``Actor.rng`` (numpy
``default_rng``; the reference never seeds the TF graph) unless ``eps``/``u`` is passed.
No CPU fallback.
"""
import ctypes

import numpy as np
import torch

from . import _lib as L

ID_FN = lambda x: x   # noqa: E731  (the reference's name)


def _dev():
    if not torch.cuda.is_available():
        raise L.AcxError("policies need a CUDA device: the hot path has no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def _soa(x, dev):
    """[B, d] host/device array -> [1][d][B] contiguous float32 device tensor (one batch 'row')."""
    t = torch.as_tensor(np.asarray(x, dtype=np.float32) if not torch.is_tensor(x) else x, device=dev)
    return t.to(torch.float32).t().contiguous()


def normalized_columns(shape, gen, norm=0.1):
    """A3C ``normalized_column_initializer`` (policies.py:29-32): N(0,1) columns scaled to L2 norm 0.1."""
    u = gen.normal(size=shape)
    return u * norm / np.sqrt(np.square(u).sum(axis=0, keepdims=True))


class _Net(object):
    """Flat device parameters + Adam moments + layout."""

    def __init__(self, net, ob_dim, num_ac, cont, gen, head_init, log_std_init=0.0):
        self.dev = _dev()
        self.net, self.ob_dim, self.num_ac, self.cont = net, ob_dim, num_ac, cont
        self.layout, self.size = L.params_layout(net, ob_dim, num_ac, cont)
        h = (128, 128, 64) if net == 0 else (256, 128, 128)
        shapes = {"W1": (ob_dim, h[0]), "W2": (h[0], h[1]), "W3": (h[1], h[2]), "Wh": (h[2], num_ac if net == 0 else 1)}
        flat = np.zeros(self.size, dtype=np.float32)
        for k, shp in shapes.items():
            off, sz = self.layout[k]
            if k == "Wh" and head_init is not None:
                w = head_init(shp, gen)
            else:
                lim = np.sqrt(6.0 / (shp[0] + shp[1]))      # tf xavier_initializer / glorot_uniform default
                w = gen.uniform(-lim, lim, size=shp)
            flat[off:off + sz] = np.asarray(w, dtype=np.float32).reshape(-1)
        if "log_std" in self.layout:
            off, sz = self.layout["log_std"]
            flat[off:off + sz] = log_std_init
        self.shapes = shapes
        self.params = torch.from_numpy(flat).to(self.dev)
        self.m = torch.zeros_like(self.params)
        self.v = torch.zeros_like(self.params)
        self.grad = torch.zeros_like(self.params)

    def tensor(self, name):
        off, sz = self.layout[name]
        return self.params[off:off + sz]

    def get(self, name):
        shp = self.shapes.get(name)
        a = self.tensor(name).cpu().numpy()
        return a.reshape(shp) if shp is not None else a

    def set(self, name, value):
        off, sz = self.layout[name]
        v = np.asarray(value, dtype=np.float32).reshape(-1)
        assert v.size == sz, "%s: expected %d values, got %d" % (name, sz, v.size)
        self.params[off:off + sz] = torch.from_numpy(v).to(self.dev)


class Actor(object):
    """MLP actor with entropy and old/new-policy regularisation; Gaussian (``act_type='cont'``) or
    categorical (``'disc'``) head.  Mirrors ``A3C/policies.py:34-134``."""

    CLIP = 0.1            # A3C/policies.py:85 (Basic_AC/policies.py:81 uses 1.0: pass grad_clip=1.0)

    def __init__(self, name="Actor", num_ob_feat=None, num_ac=None, act_type="cont", init_lr=0.005, init_beta=1,
                 init_gamma=0.01, ac_scale=2., ob_scaler=ID_FN, global_actor=None, global_step=None, ac_dim=None,
                 seed=None, grad_clip=None, head_init="a3c"):
        if num_ac is None:
            num_ac = ac_dim                                   # Basic_AC spelling (policies.py:38)
        assert num_ob_feat is not None and num_ac is not None
        self.name, self.act_type = name, act_type
        self.cont = act_type == "cont"
        self.ac_scale = float(np.max(ac_scale)) if ac_scale is not None else 1.0
        self.ob_scaler = ob_scaler
        self.rng = np.random.default_rng(seed)
        if self.cont:
            # A3C: default glorot on mu_layer (policies.py:50); Basic_AC: 0.1 x xavier (policies.py:5-6,48)
            hi = (lambda s, g: 0.1 * g.uniform(-np.sqrt(6.0 / sum(s)), np.sqrt(6.0 / sum(s)), size=s)) \
                if head_init == "basic" else None
        else:
            hi = normalized_columns if head_init == "a3c" else None      # policies.py:63 / Basic :61
        self.net = _Net(0, num_ob_feat, num_ac, self.cont, self.rng, hi)
        self.dev = self.net.dev
        self.opt = torch.zeros(L.OPT_SIZE, device=self.dev)
        self.stats = torch.zeros(L.STATS_SIZE, device=self.dev)
        clip = self.CLIP if grad_clip is None else grad_clip
        L.call("acx_opt_init", L.ptr(self.opt), float(init_lr), 0.0, float(init_beta), float(init_gamma), float(clip),
               0.002, 1e-6, 0.1, 0, L.cur_stream())
        self.global_actor, self.global_step = global_actor, global_step
        self.my_vars = self.net.params
        self._ws = None
        self._kl = torch.zeros(L.load().acx_kl_workspace_bytes() // 8, device=self.dev, dtype=torch.float64)
        if global_actor is None and name not in ("Actor",):
            # A3C: a net without a global twin has no optimize/sync (policies.py:87-98)
            self.optimize = None
            self.sync_w_global = None

    # -- forward -----------------------------------------------------------------------------------
    def act_batch(self, ob, sess=None, eps=None, u=None):
        """All rows: (ac [B, num_ac] float32 | [B] int32, logp [B], entropy [B]) as numpy arrays."""
        ob = np.asarray(self.ob_scaler(np.asarray(ob, dtype=np.float32)))
        if ob.ndim != 2:
            ob = ob[None]
        B, n = ob.shape[0], self.net
        o = _soa(ob, self.dev)
        lp = torch.empty(B, device=self.dev)
        ent = torch.empty(B, device=self.dev)
        s = L.cur_stream()
        if self.cont:
            noise = self.rng.standard_normal((B, n.num_ac)) if eps is None else np.asarray(eps).reshape(B, n.num_ac)
            e = _soa(noise, self.dev)
            ac = torch.empty((n.num_ac, B), device=self.dev)
            L.call("acx_policy_forward_sample_gauss", L.ptr(n.params), n.ob_dim, n.num_ac, self.ac_scale, 1, B, L.ptr(o),
                   L.ptr(e), 0, 0, 0, None, L.ptr(ac), L.ptr(lp), L.ptr(ent), None, s)
            return ac.t().cpu().numpy(), lp.cpu().numpy(), ent.cpu().numpy()
        noise = self.rng.random(B) if u is None else np.asarray(u).reshape(B)
        uu = torch.as_tensor(noise.astype(np.float32), device=self.dev)
        ac = torch.empty(B, device=self.dev, dtype=torch.int32)
        L.call("acx_policy_forward_sample_cat", L.ptr(n.params), n.ob_dim, n.num_ac, 1, B, L.ptr(o), L.ptr(uu), 0, 0, 0,
               None, L.ptr(ac), L.ptr(lp), L.ptr(ent), None, s)
        return ac.cpu().numpy(), lp.cpu().numpy(), ent.cpu().numpy()

    def act(self, ob, sess=None, **kw):
        """Reference semantics (policies.py:110-115): accepts a 1-D or 2-D observation, returns row 0."""
        ac, logp, ent = self.act_batch(ob, sess, **kw)
        return ac[0], logp[0], ent[0]

    def _acs_dev(self, acs, B):
        if self.cont:
            return _soa(np.asarray(acs, dtype=np.float32).reshape(B, self.net.num_ac), self.dev)
        return torch.as_tensor(np.asarray(acs).reshape(B).astype(np.int32), device=self.dev)

    def logp(self, obs, acs, logp_feeds=None):
        """log-prob of stored actions under the current weights; with logp_feeds also the mean squared
        difference (the reference's ``oldnew_kl``)."""
        obs = np.asarray(self.ob_scaler(np.asarray(obs, dtype=np.float32)))
        B, n = obs.shape[0], self.net
        o, a = _soa(obs, self.dev), self._acs_dev(acs, B)
        out = torch.empty(B, device=self.dev)
        old = None if logp_feeds is None else torch.as_tensor(np.asarray(logp_feeds, dtype=np.float32).reshape(B),
                                                              device=self.dev)
        nb = ctypes.c_int(0)
        s = L.cur_stream()
        kl = None if old is None else self._kl
        if self.cont:
            L.call("acx_policy_logp_gauss", L.ptr(n.params), n.ob_dim, n.num_ac, self.ac_scale, 1, B, L.ptr(o), L.ptr(a),
                   L.ptr(out), None, None, L.ptr(old), L.ptr(kl), ctypes.byref(nb), s)
        else:
            L.call("acx_policy_logp_cat", L.ptr(n.params), n.ob_dim, n.num_ac, 1, B, L.ptr(o), L.ptr(a), L.ptr(out), None,
                   None, L.ptr(old), L.ptr(kl), ctypes.byref(nb), s)
        if old is None:
            return out.cpu().numpy()
        L.call("acx_kl_finalize", L.ptr(self._kl), nb.value, float(B), L.ptr(self.stats), s)
        return out.cpu().numpy(), float(self.stats[L.STAT_KL_POST].item())

    def get_kl(self, sess=None, logp_feeds=None, obs=None, acs=None):
        """mean((logp_feed - logp_new)^2) on the batch (policies.py:75,106-108)."""
        return np.float32(self.logp(obs, acs, logp_feeds)[1])

    # -- training ----------------------------------------------------------------------------------
    def optimize(self, acs, obs, advs, logps, sess=None):
        """One Adam step on loss = -mean(adv*lp) + beta*mean((lp_old-lp)^2) - gamma*mean(entropy) with the
        elementwise gradient clip (policies.py:74-92).  With a ``global_actor`` the clipped local gradient
        is applied to the GLOBAL parameters using the LOCAL moments and lr (policies.py:88-89).
        Returns ``[loss, None]`` like ``sess.run([loss, opt_op])``."""
        obs = np.asarray(self.ob_scaler(np.asarray(obs, dtype=np.float32)))
        B, n = obs.shape[0], self.net
        o, a = _soa(obs, self.dev), self._acs_dev(acs, B)
        adv = torch.as_tensor(np.asarray(advs, dtype=np.float32).reshape(B), device=self.dev)
        old = torch.as_tensor(np.asarray(logps, dtype=np.float32).reshape(B), device=self.dev)
        nbytes = L.load().acx_loss_backward_workspace_bytes(n.ob_dim, n.num_ac, int(self.cont))
        if self._ws is None or self._ws.numel() < nbytes:
            self._ws = torch.empty(nbytes, device=self.dev, dtype=torch.uint8)
        s = L.cur_stream()
        if self.cont:
            L.call("acx_loss_backward_gauss", L.ptr(n.params), None, n.ob_dim, n.num_ac, self.ac_scale, 1, B, L.ptr(o),
                   L.ptr(a), L.ptr(adv), None, L.ptr(old), None, L.ptr(self.opt), L.ptr(n.grad), None,
                   L.ptr(self.stats), L.ptr(self._ws), nbytes, s)
        else:
            L.call("acx_loss_backward_cat", L.ptr(n.params), None, n.ob_dim, n.num_ac, 1, B, L.ptr(o), L.ptr(a),
                   L.ptr(adv), None, L.ptr(old), None, L.ptr(self.opt), L.ptr(n.grad), None, L.ptr(self.stats),
                   L.ptr(self._ws), nbytes, s)
        target = n if self.global_actor is None else self.global_actor.net
        L.call("acx_adam_apply", L.ptr(target.params), L.ptr(n.grad), L.ptr(n.m), L.ptr(n.v), n.size, None, None, None,
               None, 0, L.ptr(self.opt), 1.0, s)
        L.call("acx_adam_tick", L.ptr(self.opt), s)
        if self.global_step is not None:
            self.global_step.value += 1                        # apply_gradients(global_step=...) (policies.py:89)
        return [np.float32(self.stats[L.STAT_ACTOR_LOSS].item()), None]

    def sync_w_global(self, sess=None):
        """local <- global copy of every trainable tensor (policies.py:94,117-118)."""
        self.net.params.copy_(self.global_actor.net.params)
        return None

    def set_opt_param(self, sess=None, new_lr=None, new_beta=None, new_gamma=None):
        if new_beta is not None:
            self.opt[L.OPT_BETA_KL] = float(new_beta)
        if new_lr is not None:
            self.opt[L.OPT_ACTOR_LR] = float(new_lr)
        if new_gamma is not None:
            self.opt[L.OPT_GAMMA_ENT] = float(new_gamma)
        return self.get_opt_param(sess)

    def get_opt_param(self, sess=None):
        o = self.opt.cpu().numpy()
        return [o[L.OPT_ACTOR_LR], o[L.OPT_BETA_KL], o[L.OPT_GAMMA_ENT]]

    @property
    def lr(self):
        return self.get_opt_param()[0]

    @property
    def beta(self):
        return self.get_opt_param()[1]

    @property
    def gamma(self):
        return self.get_opt_param()[2]

    def printoo(self, obs, sess=None):
        """Debug printer (policies.py:100-104,133-134): mean of every variable."""
        print("Actor Variable data", [float(self.net.tensor(k).mean().item()) for k in self.net.layout])
        return 0.0

    # -- checkpoint names -----------------------------------------------------
    def checkpoint_tensors(self, scope=None):
        scope = scope or self.name
        head = "mu_layer" if self.cont else "logits"
        lr, beta, gamma = self.get_opt_param()
        out = {scope + "/Variable": np.float32(lr), scope + "/Variable_1": np.float32(gamma),
               scope + "/Variable_2": np.float32(beta)}
        for layer, (w, b) in (("first_layer", ("W1", "b1")), ("second_layer", ("W2", "b2")),
                              ("third_layer", ("W3", "b3")), (head, ("Wh", "bh"))):
            out["%s/%s/kernel" % (scope, layer)] = self.net.get(w)
            out["%s/%s/bias" % (scope, layer)] = self.net.get(b)
        if self.cont:
            out[scope + "/log_std"] = self.net.get("log_std")
        return out

    def load_checkpoint_tensors(self, tensors, scope=None):
        scope = scope or self.name
        head = "mu_layer" if self.cont else "logits"
        for layer, (w, b) in (("first_layer", ("W1", "b1")), ("second_layer", ("W2", "b2")),
                              ("third_layer", ("W3", "b3")), (head, ("Wh", "bh"))):
            self.net.set(w, tensors["%s/%s/kernel" % (scope, layer)])
            self.net.set(b, tensors["%s/%s/bias" % (scope, layer)])
        if self.cont:
            self.net.set("log_std", tensors[scope + "/log_std"])
        if scope + "/Variable" in tensors:
            self.set_opt_param(new_lr=float(tensors[scope + "/Variable"]), new_gamma=float(tensors[scope + "/Variable_1"]),
                               new_beta=float(tensors[scope + "/Variable_2"]))


class Critic(object):
    """ReLU MLP estimating the discounted return.  Mirrors ``A3C/policies.py:137-182``; ``basic_head=True``
    reproduces ``Basic_AC/policies.py:134-135`` (the value head reads the second layer)."""

    def __init__(self, name="Critic", num_ob_feat=None, init_lr=0.001, ob_scaler=ID_FN, global_critic=None, ob_scale=1.,
                 seed=None, basic_head=False):
        assert num_ob_feat is not None
        self.name, self.ob_scaler, self.ob_scale = name, ob_scaler, ob_scale
        self.basic_head = bool(basic_head)
        self.rng = np.random.default_rng(seed)
        self.net = _Net(1, num_ob_feat, 1, False, self.rng, None)
        self.dev = self.net.dev
        self.opt = torch.zeros(L.OPT_SIZE, device=self.dev)
        self.stats = torch.zeros(L.STATS_SIZE, device=self.dev)
        L.call("acx_opt_init", L.ptr(self.opt), 0.0, float(init_lr), 0.0, 0.0, 0.0, 0.002, 1e-6, 0.1, 0, L.cur_stream())
        self.global_critic = global_critic
        self.my_vars = self.net.params
        self._ws = None
        if global_critic is None and name not in ("Critic",):
            self.optimize = None
            self.sync_w_global = None

    def _obs(self, obs):
        obs = np.asarray(self.ob_scaler(np.asarray(obs, dtype=np.float32))) * np.float32(self.ob_scale)
        return obs if obs.ndim == 2 else obs[None]

    def value(self, obs, sess=None):
        obs = self._obs(obs)
        B, n = obs.shape[0], self.net
        o = _soa(obs, self.dev)
        v = torch.empty(B, device=self.dev)
        L.call("acx_value_forward", L.ptr(n.params), n.ob_dim, 1, B, L.ptr(o), L.ptr(v), int(self.basic_head),
               L.cur_stream())
        return v.cpu().numpy()

    def optimize(self, obs, targets, sess=None):
        """One Adam step on mean((v - target)^2), unclipped (policies.py:152-164); with a ``global_critic``
        the local gradient is applied to the global parameters.  Returns ``[loss, None]``."""
        obs = self._obs(obs)
        B, n = obs.shape[0], self.net
        o = _soa(obs, self.dev)
        tg = torch.as_tensor(np.asarray(targets, dtype=np.float32).reshape(B), device=self.dev)
        nbytes = L.load().acx_loss_backward_workspace_bytes(n.ob_dim, 1, 1)
        if self._ws is None or self._ws.numel() < nbytes:
            self._ws = torch.empty(nbytes, device=self.dev, dtype=torch.uint8)
        s = L.cur_stream()
        if self.basic_head:
            L.call("acx_loss_backward_basic_critic", L.ptr(n.params), n.ob_dim, 1, B, L.ptr(o), L.ptr(tg), L.ptr(n.grad),
                   L.ptr(self.stats), L.ptr(self._ws), nbytes, s)
        else:
            L.call("acx_loss_backward_gauss", None, L.ptr(n.params), n.ob_dim, 1, 1.0, 1, B, L.ptr(o), None, None, None,
                   None, L.ptr(tg), L.ptr(self.opt), None, L.ptr(n.grad), L.ptr(self.stats), L.ptr(self._ws), nbytes, s)
        target = n if self.global_critic is None else self.global_critic.net
        L.call("acx_adam_apply", None, None, None, None, 0, L.ptr(target.params), L.ptr(n.grad), L.ptr(n.m), L.ptr(n.v),
               n.size, L.ptr(self.opt), 1.0, s)
        L.call("acx_adam_tick", L.ptr(self.opt), s)
        return [np.float32(self.stats[L.STAT_CRITIC_LOSS].item()), None]

    def sync_w_global(self, sess=None):
        self.net.params.copy_(self.global_critic.net.params)
        return None

    def set_opt_param(self, new_lr, sess=None):
        """The reference feeds instead of assigning, so the critic lr never changes (policies.py:179-180,
        SURVEY.md Appendix D.4); kept: returns the value it was given, changes nothing."""
        return new_lr

    def printoo(self, obs, sess=None):
        print("Critic Variable data", [float(self.net.tensor(k).mean().item()) for k in self.net.layout])
        return [0.0]

    def checkpoint_tensors(self, scope=None):
        scope = scope or self.name
        lr = np.float32(self.opt[L.OPT_CRITIC_LR].item())
        out = {scope + "/Variable": lr, scope + "/Variable_1": lr}          # self.lr is created twice (:151,154)
        for layer, (w, b) in (("first_layer", ("W1", "b1")), ("second_layer", ("W2", "b2")),
                              ("third_layer", ("W3", "b3")), ("value", ("Wh", "bh"))):
            out["%s/%s/kernel" % (scope, layer)] = self.net.get(w)
            out["%s/%s/bias" % (scope, layer)] = self.net.get(b)
        return out

    def load_checkpoint_tensors(self, tensors, scope=None):
        scope = scope or self.name
        for layer, (w, b) in (("first_layer", ("W1", "b1")), ("second_layer", ("W2", "b2")),
                              ("third_layer", ("W3", "b3")), ("value", ("Wh", "bh"))):
            self.net.set(w, tensors["%s/%s/kernel" % (scope, layer)])
            self.net.set(b, tensors["%s/%s/bias" % (scope, layer)])
