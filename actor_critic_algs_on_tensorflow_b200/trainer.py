"""Fused rollout + update loop on one B200 (one process per GPU), the body of the reference's
outer ``while`` loop (``A3C/process.py:217-285``; ``Basic_AC/run_AC.py:208-284``) for N
environments x T steps per update instead of ~6 whole episodes:

    rollout      T x { acx_policy_forward_sample_*  ->  acx_env_step }     process.py:86-96
    values       acx_value_forward on the T+1 observation rows             process.py:229
    returns      acx_returns_adv_scan (+ batch moments)                    process.py:230,247
    gradients    acx_loss_backward_*                                       process.py:257-258
    [NCCL all-reduce of the flat gradient, world > 1]                      replaces ps push/pull
    update       acx_adam_apply, acx_adam_tick                             policies.py:88-89,160-161
    lr control   acx_policy_logp_* + acx_kl_finalize + acx_kl_lr_update    process.py:262-278

Everything stays on the device; the whole iteration is captured into CUDA graphs (the RNG step
counter lives in device memory so replays draw fresh noise).  PyTorch only provides device
memory, streams and torch.distributed.  No CPU fallback.

Multi-GPU (``world > 1``): environments are sharded by rank (index0 = rank * N keeps the Philox
streams disjoint, mirroring seed+task of ``A3C/train.py:40``); parameters, Adam moments and the
opt block are replicated.  ``mode='sync'`` is synchronous A2C: all-reduce(sum) of the flat
gradient, x 1/world, clip, one identical Adam step on every rank.  ONE collective per update: the
post-update KL of update i (which only feeds the lr rule, and the lr is first consumed by update
i+1's Adam step) rides as one extra float behind update i+1's gradient.  ``exchange='p2p'`` (the default when the
ranks' gradient buffers can be mapped into each other's address space -- torch symmetric memory on one NVSwitch node):
the sum over ranks, the lr rule, the clip and the Adam step are ONE kernel reading the peers' buffers over NVLink
(``acx_allreduce_adam``), inside the same CUDA graph as everything else: a step is one graph launch and no NCCL call.
``exchange='nccl'``: graph -> ``all_reduce`` -> graph.
``mode='async'`` mirrors the reference's parameter-server semantics (per-worker clip, per-worker
Adam moments and lr applied to shared parameters, then pull -- ``A3C/policies.py:85-98``,
``A3C/process.py:257-260``): every rank applies its own clipped gradient with its own moments to
the iteration-start parameters (stale by one push, as on the ps) and the parameter *deltas* are
summed over ranks, i.e. every worker's push lands before the pull.
"""
import ctypes
import os
import sys

import numpy as np
import torch

from . import _lib as L

ENV_IDS = {"Pendulum-v0": L.ENV_PENDULUM, "CartPole-v0": L.ENV_CARTPOLE, "Acrobot-v1": L.ENV_ACROBOT}


def env_dims(env):
    s, o, a, m = (ctypes.c_int() for _ in range(4))
    L.call("acx_env_dims", env, ctypes.byref(s), ctypes.byref(o), ctypes.byref(a), ctypes.byref(m))
    return s.value, o.value, a.value, m.value


def glorot_flat(net, ob_dim, num_ac, cont, gen):
    """xavier-uniform kernels, zero biases, log_std = 0 (A3C/policies.py:7,44-52) in the flat layout."""
    lay, total = L.params_layout(net, ob_dim, num_ac, cont)
    flat = np.zeros(total, dtype=np.float32)
    h = (128, 128, 64) if net == 0 else (256, 128, 128)
    shapes = {"W1": (ob_dim, h[0]), "W2": (h[0], h[1]), "W3": (h[1], h[2]), "Wh": (h[2], num_ac if net == 0 else 1)}
    for k, (fi, fo) in shapes.items():
        lim = np.sqrt(6.0 / (fi + fo))
        off, size = lay[k]
        flat[off:off + size] = gen.uniform(-lim, lim, size=size)
    return flat


XGPU_MAX_RANKS = 8     # ACX_XGPU_MAX_RANKS of include/acx.h
GRAD_TAIL = 4      # floats behind the flat gradient in the exchange buffer: [KL of the previous update, 3 x pad]


def sync_exchange(gradbuf, pg=None):
    """The one collective of a synchronous update: SUM over ranks of [flat gradient ‖ previous post-update KL]
    (replaces the workers' gradient pushes onto the ps variables, A3C/policies.py:88-89)."""
    import torch.distributed as dist
    dist.all_reduce(gradbuf, op=dist.ReduceOp.SUM, group=pg)
    return gradbuf


def async_push_pull(params, old_params, pg=None):
    """The one collective of a stale-gradient update: every rank has applied its own Adam step to the
    iteration-start parameters ``old_params``; the per-rank deltas are summed onto them (all pushes land,
    A3C/policies.py:88-89) and every rank continues from the sum (the pull, A3C/process.py:259-260)."""
    import torch.distributed as dist
    params.sub_(old_params)
    dist.all_reduce(params, op=dist.ReduceOp.SUM, group=pg)
    params.add_(old_params)
    return params


class A2CTrainer:
    def __init__(self, env_id="Pendulum-v0", n_envs=65536, horizon=20, seed=12321, gamma=0.98, look_ahead=40,
                 desired_kl=0.002, min_lr=1e-6, max_lr=0.1, clip=0.1, actor_lr=0.005, critic_lr=0.001, beta=1.0,
                 gamma_ent=0.01, schedule=True, rank=0, world=1, mode="sync", device=None, actor_params=None,
                 critic_params=None, use_graph=True, max_episode_steps=0, process_group=None, tensor_cores=True,
                 cache_activations=True, tensor_core_backward=True, frames=1, fused_rollout=True, stagger=True,
                 diagnostics=False, log_rows=128, basic=False, overlap_nets=True, exchange="auto"):
        if not torch.cuda.is_available():
            raise L.AcxError("A2CTrainer needs a CUDA device: the hot path has no CPU fallback")
        L.load()
        self.env = ENV_IDS[env_id] if isinstance(env_id, str) else int(env_id)
        self.S, self.env_ob, self.n_act, default_max = env_dims(self.env)
        # Framer (A3C/process.py:25-43, --frames k): the networks see the last k observations of the episode
        self.frames = int(frames)
        self.ob = self.env_ob * self.frames
        if self.frames < 1 or self.ob > 24:
            raise L.AcxError("frames=%d: the MLP kernels take at most 24 input features (%d observation components x "
                             "frames)" % (self.frames, self.env_ob))
        # more than 8 network inputs (Pendulum / CartPole from 3 frames, Acrobot from 2): the tensor-core kernels keep
        # shared memory for the weights and take 8; the FFMA engine takes 24 and runs the whole path then
        if self.ob > 8:
            tensor_cores = False
        # Basic_AC's critic: the value head reads the SECOND hidden layer, the third layer is dead weight and never gets a
        # gradient (Basic_AC/policies.py:134-135).  Its other differences are plain arguments: clip=1.0 (:81), max_lr=1.0
        # (Basic_AC/run_AC.py:178).  The actor is the same network and runs on the same kernels.
        self.basic = bool(basic)
        self.cont = self.n_act == 0
        self.num_ac = 1 if self.cont else self.n_act
        self.ac_scale = 2.0                       # max(|low|, high) of Pendulum (A3C/process.py:186)
        self.N, self.T = int(n_envs), int(horizon)
        self.seed, self.rank, self.world, self.mode = int(seed), int(rank), int(world), mode
        assert mode in ("sync", "async")
        self.gamma, self.look_ahead = float(gamma), int(look_ahead)
        self.max_steps = int(max_episode_steps) or default_max
        self.dev = torch.device(device if device is not None else "cuda:%d" % torch.cuda.current_device())
        self.pg = process_group
        self.use_graph = bool(use_graph)
        # True: forward passes (rollout sampling, values, post-update log-prob/KL) on tcgen05 with 3xTF32 operand
        # splits (fp32-class accuracy, same test tolerances, ~1.9x faster than the FFMA forward -- profiles/
        # r01_tc_forward.md); False keeps every GEMM on the bit-stable FFMA engine
        self.tc = bool(tensor_cores)
        # keep the hidden activations of the rollout / value passes in HBM and let the update read them instead of
        # recomputing layers 2 and 3 (same weights => bit-identical; 2.4 GB at 65,536 envs x 20 steps)
        self.cache = bool(cache_activations)
        # back-propagation chain products on tcgen05 too (needs both of the above); the weight-gradient products stay
        # fp32 FFMA in the same kernel
        self.tc_bwd = bool(tensor_core_backward) and self.tc and self.cache
        # the T-step rollout as ONE persistent kernel (policy forward + sampling + env step per step, env state in
        # registers); bit-identical to the T x {sample, env step} launches it replaces
        self.fused_rollout = bool(fused_rollout) and self.tc and int(frames) == 1
        d, f32 = self.dev, torch.float32
        N, T = self.N, self.T

        _, self.Pa = L.params_layout(0, self.ob, self.num_ac, self.cont)
        _, self.Pc = L.params_layout(1, self.ob)
        gen = np.random.default_rng(self.seed)
        a0 = glorot_flat(0, self.ob, self.num_ac, self.cont, gen) if actor_params is None else np.asarray(actor_params)
        c0 = glorot_flat(1, self.ob, 1, False, gen) if critic_params is None else np.asarray(critic_params)
        assert a0.size == self.Pa and c0.size == self.Pc, "flat parameter size mismatch"
        P = self.Pa + self.Pc
        self.params = torch.empty(P, device=d, dtype=f32)
        self.params[:self.Pa] = torch.from_numpy(a0.astype(np.float32)).to(d)
        self.params[self.Pa:] = torch.from_numpy(c0.astype(np.float32)).to(d)
        self.actor, self.critic = self.params[:self.Pa], self.params[self.Pa:]
        self.m = torch.zeros(P, device=d, dtype=f32)
        self.v = torch.zeros(P, device=d, dtype=f32)
        # exchange buffer of the multi-rank update: flat gradient (actor ‖ critic) + GRAD_TAIL floats
        self._gradbuf = torch.zeros(P + GRAD_TAIL, device=d, dtype=f32)
        self.grads = self._gradbuf[:P]
        self.old_params = torch.empty(P, device=d, dtype=f32) if mode == "async" else None
        # how a synchronous multi-rank update exchanges gradients: 'p2p' = acx_allreduce_adam over peer-mapped buffers
        # (one kernel: sum over ranks + lr rule + clip + Adam), 'nccl' = all_reduce between two graphs; 'auto' takes p2p
        # when the buffers can be mapped and says so on stderr when they cannot.  ACX_EXCHANGE overrides the argument.
        self.exchange, self.grad_sum = "nccl", None
        want = os.environ.get("ACX_EXCHANGE", exchange)
        assert want in ("auto", "p2p", "nccl"), want
        if self.world > 1 and mode == "sync" and want != "nccl":
            err = None
            try:
                self._setup_p2p(P)
            except Exception as e:       # noqa: BLE001 -- any failure to map peer memory means the NCCL exchange
                err = e
            # the ranks must agree (one that could not map its peers would sit in all_reduce while the others spin in
            # acx_allreduce_adam); the same collective is the barrier after which every rank's flag words are zero
            import torch.distributed as dist
            agreed = torch.tensor([1 if self.exchange == "p2p" else 0], device=d, dtype=torch.int32)
            dist.all_reduce(agreed, op=dist.ReduceOp.MIN, group=self.pg)
            if int(agreed.item()) == 0:
                why = ("%s: %s" % (type(err).__name__, err)) if err is not None else "another rank could not map the buffers"
                if want == "p2p":
                    raise L.AcxError("exchange='p2p' unavailable on rank %d (%s)" % (self.rank, why)) from err
                print("[acx] rank %d: peer-mapped exchange unavailable (%s); using the NCCL all-reduce" % (self.rank, why),
                      file=sys.stderr, flush=True)
                self.exchange = "nccl"          # a mapped exchange buffer stays where it is: NCCL reduces it as well

        self.state = torch.empty((self.S, N), device=d, dtype=f32)
        self.elapsed = torch.zeros(N, device=d, dtype=torch.int32)
        self.obs_buf = torch.empty((T + 1, self.ob, N), device=d, dtype=f32)
        self.raw_obs = torch.empty((self.env_ob, N), device=d, dtype=f32) if self.frames > 1 else None
        self.acs = (torch.empty((T, self.num_ac, N), device=d, dtype=f32) if self.cont
                    else torch.empty((T, N), device=d, dtype=torch.int32))
        self.logps = torch.empty((T, N), device=d, dtype=f32)
        self.ent = torch.empty((T, N), device=d, dtype=f32)
        self.rew = torch.empty((T, N), device=d, dtype=f32)
        self.done = torch.empty((T, N), device=d, dtype=torch.uint8)
        self.V = torch.empty((T + 1, N), device=d, dtype=f32)
        self.G = torch.empty((T, N), device=d, dtype=f32)
        self.adv = torch.empty((T, N), device=d, dtype=f32)
        self.stats = torch.zeros(L.STATS_SIZE, device=d, dtype=f32)
        self.opt = torch.zeros(L.OPT_SIZE, device=d, dtype=f32)
        self.step_ctr = torch.zeros(1, device=d, dtype=torch.int32)
        # training-log diagnostics kept on the device (avg_rew, ev_before, ev_after, one ring row per iteration)
        self.diagnostics = bool(diagnostics)
        self.ep_ret = torch.zeros(N, device=d, dtype=f32)
        self.ws_diag_bytes = L.load().acx_diag_workspace_bytes()
        self.ws_diag = torch.zeros(self.ws_diag_bytes, device=d, dtype=torch.uint8)
        self.V_after = torch.empty((T, N), device=d, dtype=f32) if self.diagnostics else None
        self.log_rows = int(log_rows)
        self.log_ring = torch.zeros((self.log_rows, 32), device=d, dtype=f32)
        self.ws_scan_bytes = L.load().acx_returns_adv_workspace_bytes(T, N)
        self.ws_scan = torch.zeros(self.ws_scan_bytes, device=d, dtype=torch.uint8)
        self.ws_bwd_bytes = L.load().acx_loss_backward_workspace_bytes(self.ob, self.num_ac, int(self.cont))
        if self.tc_bwd and N % 4 == 0:       # room for dPre3 / dPre2 / X1: weight gradients on the tensor cores too
            self.ws_bwd_bytes = L.load().acx_loss_backward_tc_workspace_bytes(self.ob, self.num_ac, int(self.cont), T, N)
        self.ws_bwd = torch.empty(self.ws_bwd_bytes, device=d, dtype=torch.uint8)
        # the two networks' backward passes are independent: the critic's runs on a second stream (a parallel branch of
        # the CUDA graph) with a workspace of its own.  Every kernel of the pair is persistent and fills the GPU, so the
        # two do not run side by side -- but the one's CTAs start on the SMs the other's last wave leaves idle, and the
        # small fixed-order reduce kernels hide behind the other network's work (step 5.15 -> 5.07 ms).  Branching the
        # operand packs and the episode statistics off the same way measured no further gain and was dropped.
        self.overlap_nets = bool(overlap_nets) and self.tc_bwd and N % 4 == 0 and not self.basic
        self._side = torch.cuda.Stream(device=d) if self.overlap_nets else None
        self.ws_bwd2 = torch.empty(self.ws_bwd_bytes, device=d, dtype=torch.uint8) if self.overlap_nets else None
        self.kl_part = torch.zeros(L.load().acx_kl_workspace_bytes() // 8, device=d, dtype=torch.float64)
        self.kl_blocks = ctypes.c_int(0)
        if self.cache:      # tile-blocked [rows][ceil(N/128)][features][128]; one row of the 2-D tensor = one time step
            per = lambda feat: L.load().acx_activation_cache_floats(1, feat, N)
            self.ax2 = torch.empty((T, per(128)), device=d, dtype=f32)
            self.ax3 = torch.empty((T, per(64)), device=d, dtype=f32)
            self.cx2 = torch.empty((T + 1, per(128)), device=d, dtype=f32)
            self.cx3 = torch.empty((T + 1, per(128)), device=d, dtype=f32)
        # activation sign masks for the tensor-core backward (written by the value pass and the fused rollout)
        self.amask = self.cmask = None
        if self.tc_bwd:
            self.cmask = torch.empty(L.load().acx_act_mask_words(1, T + 1, N), device=d, dtype=torch.int32)
            if self.fused_rollout:
                self.amask = torch.empty(L.load().acx_act_mask_words(0, T, N), device=d, dtype=torch.int32)
        self.packed_a = torch.zeros(L.load().acx_tc_packed_floats(0, self.ob, self.num_ac, int(self.cont)), device=d)
        self.packed_c = torch.zeros(L.load().acx_tc_packed_floats(1, self.ob, 1, 0), device=d)
        self.packed_bwd_a = torch.zeros(L.load().acx_tc_packed_bwd_floats(0, self.ob, self.num_ac, int(self.cont)), device=d)
        self.packed_bwd_c = torch.zeros(L.load().acx_tc_packed_bwd_floats(1, self.ob, 1, 0), device=d)
        # optional caller-supplied noise (parity / end-to-end mode), static so graphs can replay on them
        self.noise_buf = None
        self.reset_u_buf = None
        # double-buffered input staging for prefetch(): [noise, reset_u, ready event] per buffer set
        self._stage, self._stage_cur, self._stage_next_ready = [None, None], 0, 0
        self._copy_stream, self._ev_iter, self._own = None, None, None
        self.index0 = self.rank * N
        self.iterations = 0
        self._graphs = {}
        self.launches_per_iteration = 0

        s = L.cur_stream()
        L.call("acx_opt_init", L.ptr(self.opt), actor_lr, critic_lr, beta, gamma_ent, clip, desired_kl, min_lr, max_lr,
               int(bool(schedule)), s)
        L.call("acx_env_reset", self.env, N, L.ptr(self.state), L.ptr(self.elapsed),
               L.ptr(self.obs_buf[0] if self.frames == 1 else self.raw_obs), None, self.seed, self.index0, 0xFFFFFFFF, s)
        if self.frames > 1:
            L.call("acx_frame_push", N, self.env_ob, self.frames, None, L.ptr(self.raw_obs), None, L.ptr(self.obs_buf[0]), s)
        # N environments reset together would hit the TimeLimit together: every 20-step batch would hold one phase of
        # the episode and (Pendulum) all done flags of a run would fall into every tenth update.  The reference batches
        # whole episodes; spreading the time-limit counters over [0, max_steps) gives every update every phase.
        self.stagger = bool(stagger)
        if self.stagger:
            L.call("acx_env_stagger", self.env, N, L.ptr(self.elapsed), self.seed, self.index0, self.max_steps, s)
            # an episode that starts with part of its time limit spent is shorter than a real one: its return is not
            # counted when it ends (NaN marker, acx_episode_returns)
            self.ep_ret.copy_(torch.where(self.elapsed > 0, float("nan"), 0.0))
        self._pack()

    # ------------------------------------------------------------------ pieces (all enqueue on the current stream)
    def _row(self, t, i):      # pointer to row i of a [rows, ...] tensor
        return t.data_ptr() + i * t.stride(0) * t.element_size()

    def _pack(self):
        """Refresh the tensor-core operand copies of the hidden-layer kernels (after every parameter change)."""
        if not self.tc:
            return 0
        s = L.cur_stream()
        L.call("acx_tc_pack_weights", 0, L.ptr(self.actor), self.ob, self.num_ac, int(self.cont), L.ptr(self.packed_a), s)
        L.call("acx_tc_pack_weights", 1, L.ptr(self.critic), self.ob, 1, 0, L.ptr(self.packed_c), s)
        if not self.tc_bwd:
            return 2
        L.call("acx_tc_pack_weights_bwd", 0, L.ptr(self.actor), self.ob, self.num_ac, int(self.cont),
               L.ptr(self.packed_bwd_a), s)
        L.call("acx_tc_pack_weights_bwd", 1, L.ptr(self.critic), self.ob, 1, 0, L.ptr(self.packed_bwd_c), s)
        return 4

    def _rollout(self, buffered_noise):
        s, N = L.cur_stream(), self.N
        if self.fused_rollout:
            L.call("acx_rollout_tc", self.env, L.ptr(self.actor), L.ptr(self.packed_a), self.ac_scale, self.T, N,
                   L.ptr(self.obs_buf), L.ptr(self.state), L.ptr(self.elapsed),
                   L.ptr(self.noise_buf) if buffered_noise else None, L.ptr(self.reset_u_buf) if buffered_noise else None,
                   self.seed, self.index0, L.ptr(self.step_ctr), self.max_steps, L.ptr(self.acs), L.ptr(self.logps),
                   L.ptr(self.ent), L.ptr(self.rew), L.ptr(self.done), L.ptr(self.ax2) if self.cache else None,
                   L.ptr(self.ax3) if self.cache else None, L.ptr(self.amask), s)
            L.call("acx_counter_add", L.ptr(self.step_ctr), self.T, s)
            return 2
        n = 0
        for t in range(self.T):
            noise = None if not buffered_noise else self._row(self.noise_buf, t)
            ru = None if not buffered_noise else self._row(self.reset_u_buf, t)
            tail = (self._row(self.obs_buf, t), noise, self.seed, self.index0, t, L.ptr(self.step_ctr),
                    self._row(self.acs, t), self._row(self.logps, t), self._row(self.ent, t), None, s)
            cache = (self._row(self.ax2, t), self._row(self.ax3, t)) if self.cache else (None, None)
            if self.cont and self.tc:
                L.call("acx_policy_forward_sample_gauss_tc", L.ptr(self.actor), L.ptr(self.packed_a), self.ob,
                       self.num_ac, self.ac_scale, 1, N, *tail[:-1], *cache, s)
            elif self.cont and self.cache:
                L.call("acx_policy_forward_sample_gauss_cached", L.ptr(self.actor), self.ob, self.num_ac, self.ac_scale,
                       1, N, *tail[:-1], self._row(self.ax2, t), self._row(self.ax3, t), s)
            elif self.cont:
                L.call("acx_policy_forward_sample_gauss", L.ptr(self.actor), self.ob, self.num_ac, self.ac_scale, 1, N,
                       *tail)
            elif self.tc:
                L.call("acx_policy_forward_sample_cat_tc", L.ptr(self.actor), L.ptr(self.packed_a), self.ob, self.n_act,
                       1, N, *tail[:-1], *cache, s)
            elif self.cache:
                L.call("acx_policy_forward_sample_cat_cached", L.ptr(self.actor), self.ob, self.n_act, 1, N, *tail[:-1],
                       self._row(self.ax2, t), self._row(self.ax3, t), s)
            else:
                L.call("acx_policy_forward_sample_cat", L.ptr(self.actor), self.ob, self.n_act, 1, N, *tail)
            L.call("acx_env_step", self.env, N, L.ptr(self.state), L.ptr(self.elapsed), self._row(self.acs, t),
                   self._row(self.obs_buf, t + 1) if self.frames == 1 else L.ptr(self.raw_obs), self._row(self.rew, t),
                   self._row(self.done, t), ru, self.seed, self.index0, t, L.ptr(self.step_ctr), self.max_steps, s)
            n += 2
            if self.frames > 1:
                L.call("acx_frame_push", N, self.env_ob, self.frames, self._row(self.obs_buf, t), L.ptr(self.raw_obs),
                       self._row(self.done, t), self._row(self.obs_buf, t + 1), s)
                n += 1
        L.call("acx_counter_add", L.ptr(self.step_ctr), self.T, s)
        return n + 1

    def _episode_stats(self):
        L.call("acx_episode_returns", self.T, self.N, L.ptr(self.rew), L.ptr(self.done), L.ptr(self.ep_ret),
               L.ptr(self.stats), L.ptr(self.ws_diag), self.ws_diag_bytes, L.cur_stream())
        return 1

    def _gradients(self):
        s, N, T = L.cur_stream(), self.N, self.T
        if self.tc:
            L.call("acx_value_forward_tc", L.ptr(self.critic), L.ptr(self.packed_c), self.ob, T + 1, N,
                   L.ptr(self.obs_buf), L.ptr(self.V), int(self.basic), L.ptr(self.cx2) if self.cache else None,
                   L.ptr(self.cx3) if self.cache else None, L.ptr(self.cmask), s)
        elif self.cache:
            L.call("acx_value_forward_cached", L.ptr(self.critic), self.ob, T + 1, N, L.ptr(self.obs_buf), L.ptr(self.V),
                   int(self.basic), L.ptr(self.cx2), L.ptr(self.cx3), s)
        else:
            L.call("acx_value_forward", L.ptr(self.critic), self.ob, T + 1, N, L.ptr(self.obs_buf), L.ptr(self.V),
                   int(self.basic), s)
        L.call("acx_returns_adv_scan", T, N, L.ptr(self.rew), L.ptr(self.done), L.ptr(self.V), self.gamma,
               self.look_ahead, L.ptr(self.G), L.ptr(self.adv), L.ptr(self.stats), L.ptr(self.ws_scan),
               self.ws_scan_bytes, L.SCAN_AUTO, s)
        ga, gc = self.grads[:self.Pa], self.grads[self.Pa:]
        common = (L.ptr(self.obs_buf), L.ptr(self.acs), L.ptr(self.adv), L.ptr(self.stats), L.ptr(self.logps),
                  L.ptr(self.G), L.ptr(self.opt), L.ptr(ga), L.ptr(gc), L.ptr(self.stats), L.ptr(self.ws_bwd),
                  self.ws_bwd_bytes, s)
        critic_ptr, extra = L.ptr(self.critic), 0
        if self.basic:
            # the Basic_AC critic first (FFMA kernel with the head on layer 2; the third layer's gradient is zero, as TF
            # reports None for it), then the actor alone through the engine's usual entry point
            L.call("acx_loss_backward_basic_critic", L.ptr(self.critic), self.ob, T, N, L.ptr(self.obs_buf), L.ptr(self.G),
                   L.ptr(gc), L.ptr(self.stats), L.ptr(self.ws_bwd), self.ws_bwd_bytes, s)
            critic_ptr, extra = None, 3
        if self.tc_bwd:
            cached = (common[:-3] + (L.ptr(self.ax2), L.ptr(self.ax3), L.ptr(self.cx2), L.ptr(self.cx3), L.ptr(self.amask),
                                     L.ptr(self.cmask)) + common[-3:])
            def launch(actor_p, critic_p, ws, stream):
                head = (actor_p, critic_p, L.ptr(self.packed_bwd_a), L.ptr(self.packed_bwd_c), self.ob)
                tail = cached[:-3] + (L.ptr(ws), self.ws_bwd_bytes, stream)
                if self.cont:
                    L.call("acx_loss_backward_gauss_tc", *head, self.num_ac, self.ac_scale, T, N, *tail)
                else:
                    L.call("acx_loss_backward_cat_tc", *head, self.n_act, T, N, *tail)
            if self.overlap_nets:
                cur = torch.cuda.current_stream()
                self._side.wait_stream(cur)
                with torch.cuda.stream(self._side):
                    launch(None, L.ptr(self.critic), self.ws_bwd2, L.cur_stream())
                launch(L.ptr(self.actor), None, self.ws_bwd, s)
                cur.wait_stream(self._side)
            else:
                launch(L.ptr(self.actor), critic_ptr, self.ws_bwd, s)
            split = self.N % 4 == 0
            per_net = 4 if split else 2         # chain, dW3 kernel, dW2 kernel, grad reduce | train, reduce
            return 2 + extra + per_net * (1 if self.basic else 2)      # value, scan, ...
        if self.cache:
            cached = common[:-3] + (L.ptr(self.ax2), L.ptr(self.ax3), L.ptr(self.cx2), L.ptr(self.cx3)) + common[-3:]
            if self.cont:
                L.call("acx_loss_backward_gauss_cached", L.ptr(self.actor), critic_ptr, self.ob, self.num_ac,
                       self.ac_scale, T, N, *cached)
            else:
                L.call("acx_loss_backward_cat_cached", L.ptr(self.actor), critic_ptr, self.ob, self.n_act, T, N,
                       *cached)
        elif self.cont:
            L.call("acx_loss_backward_gauss", L.ptr(self.actor), critic_ptr, self.ob, self.num_ac,
                   self.ac_scale, T, N, *common)
        else:
            L.call("acx_loss_backward_cat", L.ptr(self.actor), critic_ptr, self.ob, self.n_act, T, N, *common)
        return 2 + extra + (3 if self.basic else 6)      # value, scan, per net (transpose, mlp train, grad reduce)

    def _apply(self, grad_scale):
        s, Pa, Pc = L.cur_stream(), self.Pa, self.Pc
        L.call("acx_adam_apply", L.ptr(self.actor), L.ptr(self.grads[:Pa]), L.ptr(self.m[:Pa]), L.ptr(self.v[:Pa]), Pa,
               L.ptr(self.critic), L.ptr(self.grads[Pa:]), L.ptr(self.m[Pa:]), L.ptr(self.v[Pa:]), Pc, L.ptr(self.opt),
               grad_scale, s)
        L.call("acx_adam_tick", L.ptr(self.opt), s)
        return 2 + (self._pack() if self.world == 1 or self.mode == "sync" else 0)

    def _post_kl(self):
        s, N, T = L.cur_stream(), self.N, self.T
        tail = (T, N, L.ptr(self.obs_buf), L.ptr(self.acs), None, None, None, L.ptr(self.logps), L.ptr(self.kl_part),
                ctypes.byref(self.kl_blocks), s)
        if self.cont and self.tc:
            L.call("acx_policy_logp_gauss_tc", L.ptr(self.actor), L.ptr(self.packed_a), self.ob, self.num_ac,
                   self.ac_scale, *tail)
        elif self.cont:
            L.call("acx_policy_logp_gauss", L.ptr(self.actor), self.ob, self.num_ac, self.ac_scale, *tail)
        elif self.tc:
            L.call("acx_policy_logp_cat_tc", L.ptr(self.actor), L.ptr(self.packed_a), self.ob, self.n_act, *tail)
        else:
            L.call("acx_policy_logp_cat", L.ptr(self.actor), self.ob, self.n_act, *tail)
        L.call("acx_kl_finalize", L.ptr(self.kl_part), self.kl_blocks.value, float(T) * float(N), L.ptr(self.stats), s)
        return 2

    def _lr_update(self, flags=3, kl_sum=None, kl_scale=1.0):
        """flags 3: the lr rule on stats[KL_POST] + schedules + iteration bump (acx_kl_lr_update); 1 / 2: one half."""
        if flags == 3 and kl_sum is None:
            L.call("acx_kl_lr_update", L.ptr(self.opt), L.ptr(self.stats), L.cur_stream())
        else:
            L.call("acx_kl_lr_update_split", L.ptr(self.opt), L.ptr(self.stats), kl_sum, kl_scale, flags, L.cur_stream())
        return 1

    def _ev_after(self):
        """ev_after of the reference's log row (A3C/process.py:261): the updated critic on the batch's observations
        against the targets the update was fitted to.  One more value pass -- only with diagnostics=True."""
        if not self.diagnostics:
            return 0
        s, N, T = L.cur_stream(), self.N, self.T
        if self.tc:
            L.call("acx_value_forward_tc", L.ptr(self.critic), L.ptr(self.packed_c), self.ob, T, N, L.ptr(self.obs_buf),
                   L.ptr(self.V_after), int(self.basic), None, None, None, s)
        else:
            L.call("acx_value_forward", L.ptr(self.critic), self.ob, T, N, L.ptr(self.obs_buf), L.ptr(self.V_after),
                   int(self.basic), s)
        L.call("acx_explained_variance", T * N, L.ptr(self.G), L.ptr(self.V_after),
               self.stats.data_ptr() + 4 * L.STAT_EV_AFTER, L.ptr(self.ws_diag), self.ws_diag_bytes, s)
        return 2

    def _log_push(self):
        L.call("acx_log_push", L.ptr(self.log_ring), self.log_rows, L.ptr(self.stats), L.ptr(self.opt), L.cur_stream())
        return 1

    # ------------------------------------------------------------------ segments + graphs
    def _seg_a(self, buffered):
        self.obs_buf[0].copy_(self.obs_buf[self.T])      # the last observation of the previous rollout
        return self._rollout(buffered) + self._episode_stats() + self._gradients() + 1

    def _seg_b_single(self):
        return self._apply(1.0) + self._post_kl() + self._ev_after() + self._lr_update() + self._log_push()

    def _seg_b_sync(self, first):
        """After the all-reduce of [gradient ‖ previous KL]: the previous update's lr rule on the rank-averaged KL
        (not in the very first update), Adam on the averaged gradient, this update's local KL, which is left behind
        the gradient for the next exchange, then the schedules and the iteration bump."""
        P = self.Pa + self.Pc
        n = 0
        if not first:
            n += self._lr_update(1, self._gradbuf.data_ptr() + 4 * P, 1.0 / self.world)
        n += self._apply(1.0 / self.world) + self._post_kl() + self._ev_after()
        self._gradbuf[P:P + 1].copy_(self.stats[L.STAT_KL_POST:L.STAT_KL_POST + 1])
        return n + 1 + self._lr_update(2) + self._log_push()

    def _setup_p2p(self, P):
        """Re-home the exchange buffer [gradient ‖ tail ‖ flags] into memory every rank of the group has mapped (torch
        symmetric memory: plumbing only -- the kernel that reads the peers is acx_allreduce_adam)."""
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        lib = L.load()
        if self.world > XGPU_MAX_RANKS:
            raise L.AcxError("acx_allreduce_adam takes at most %d ranks" % XGPU_MAX_RANKS)
        group = self.pg if self.pg is not None else dist.group.WORLD
        nfl = lib.acx_xgpu_buffer_bytes(P + GRAD_TAIL) // 4
        buf = symm.empty(nfl, dtype=torch.float32, device=self.dev)
        buf.zero_()
        torch.cuda.synchronize()
        hdl = symm.rendezvous(buf, group)
        ptrs = [int(x) for x in hdl.buffer_ptrs]
        if len(ptrs) != self.world or hdl.rank != self.rank or ptrs[self.rank] != buf.data_ptr():
            raise L.AcxError("symmetric-memory group does not match rank %d / world %d" % (self.rank, self.world))
        self._xbuf, self._xhdl = buf, hdl
        self._peers = (ctypes.c_void_p * self.world)(*ptrs)
        self._gradbuf = buf[:P + GRAD_TAIL]
        self.grads = self._gradbuf[:P]
        self.grad_sum = torch.zeros(P + GRAD_TAIL, device=self.dev, dtype=torch.float32)
        self._xgen = torch.zeros(128, device=self.dev, dtype=torch.int32)
        torch.cuda.synchronize()         # flags zeroed; the caller's agreement all_reduce is the barrier before first use
        self.exchange = "p2p"

    def _seg_b_p2p(self, first):
        """_seg_b_sync with the exchange inside: one kernel sums the ranks' [gradient ‖ previous KL] over NVLink,
        applies the previous update's lr rule (not in the very first update), the clip and Adam."""
        Pa, Pc, s = self.Pa, self.Pc, L.cur_stream()
        P, scale = Pa + Pc, 1.0 / self.world
        L.call("acx_allreduce_adam", self._peers, self.world, self.rank, Pa, Pc, GRAD_TAIL, L.ptr(self.actor),
               L.ptr(self.m[:Pa]), L.ptr(self.v[:Pa]), L.ptr(self.critic), L.ptr(self.m[Pa:]), L.ptr(self.v[Pa:]),
               L.ptr(self.opt), L.ptr(self.grad_sum), L.ptr(self._xgen), scale, int(not first), s)
        n = 1
        if not first:                    # records the lr the kernel has just used (same rule, same words)
            n += self._lr_update(1, self.grad_sum.data_ptr() + 4 * P, scale)
        L.call("acx_adam_tick", L.ptr(self.opt), s)
        n += 1 + self._pack() + self._post_kl() + self._ev_after()
        self._gradbuf[P:P + 1].copy_(self.stats[L.STAT_KL_POST:L.STAT_KL_POST + 1])
        return n + 1 + self._lr_update(2) + self._log_push()

    def _seg_c_async(self):
        return self._pack() + self._post_kl() + self._ev_after() + self._lr_update() + self._log_push()

    def _run(self, key, fn):
        """Run fn() on the current stream, through a CUDA graph after the first (eager, warm-up) call."""
        if not self.use_graph:
            return fn()
        ent = self._graphs.get(key)
        if ent is None:
            n = fn()                                   # eager warm-up (also sets lazily configured attributes)
            self._graphs[key] = [n, 1, None]
            return n
        if ent[2] is None:
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                fn()
            ent[2] = g
        ent[2].replay()
        return ent[0]

    iterations_started = False

    def prefetch(self, noise, reset_u):
        """Start the host->device copy of the NEXT iteration's action noise / reset uniforms (pinned host tensors)
        on a copy stream, into the staging set the running iteration does not read; ``iteration(prefetched=True)``
        then consumes it.  The copy overlaps the kernels of the iteration enqueued before this call -- the loader
        running one batch ahead, as a queue-fed ``sess.run`` would."""
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=self.dev)
        nxt = 1 - self._stage_cur
        if self._stage[nxt] is None:
            self._stage[nxt] = [torch.empty(noise.shape, device=self.dev, dtype=noise.dtype),
                                torch.empty((self.T, self.S, self.N), device=self.dev, dtype=torch.float32),
                                torch.cuda.Event()]
        nb, rb, ev = self._stage[nxt]
        cs = self._copy_stream
        # the set was last read by the rollout of the iteration before the latest one: everything enqueued before the
        # latest iteration's start marker has to be through
        if self._ev_iter is not None:
            cs.wait_event(self._ev_iter)
        else:
            cs.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(cs):
            nb.copy_(noise, non_blocking=True)
            rb.copy_(reset_u, non_blocking=True)
            ev.record(cs)
        self._stage_next_ready = nxt

    def wait_prefetch(self):
        """Make the current stream wait for the copy the last prefetch() started (iteration(prefetched=True) does so
        itself; this is for callers that want the copy to end inside an interval they time)."""
        torch.cuda.current_stream().wait_event(self._stage[self._stage_next_ready][2])

    def iteration(self, noise=None, reset_u=None, prefetched=False):
        """One rollout + update.  ``noise``/``reset_u``: optional tensors [T, ac, N] (normals) or [T, N]
        (uniforms) and [T, S, N], copied to the device here; ``prefetched=True`` takes the set a preceding
        ``prefetch()`` staged; default is the device Philox stream."""
        if self._ev_iter is None:
            self._ev_iter = torch.cuda.Event()
        self._ev_iter.record()
        buffered = noise is not None or prefetched
        tag = "buf" if buffered else "rng"
        if prefetched:
            self._stage_cur = self._stage_next_ready
            self.noise_buf, self.reset_u_buf, ev = self._stage[self._stage_cur]
            torch.cuda.current_stream().wait_event(ev)
            tag = "pre%d" % self._stage_cur            # one CUDA graph per staging set
        elif buffered:
            if self._own is None:
                self._own = (torch.empty_like(noise, device=self.dev),
                             torch.empty((self.T, self.S, self.N), device=self.dev, dtype=torch.float32))
            self.noise_buf, self.reset_u_buf = self._own       # the "a_buf" graph is captured on this set
            self.noise_buf.copy_(noise, non_blocking=True)
            self.reset_u_buf.copy_(reset_u, non_blocking=True)
        first = not self.iterations_started
        p2p = self.world > 1 and self.mode == "sync" and self.exchange == "p2p"
        if first:
            # first iteration: obs_buf[0] already holds the reset observations
            n = self._rollout(buffered) + self._episode_stats() + self._gradients()
            self.iterations_started = True
        elif p2p:                        # the whole update, exchange included, is one graph
            n = self._run("ab_" + tag, lambda: self._seg_a(buffered) + self._seg_b_p2p(False))
        else:
            n = self._run("a_" + tag, lambda: self._seg_a(buffered))
        if self.world == 1:
            n += self._run("b1", self._seg_b_single)
        elif p2p:
            if first:
                n += self._seg_b_p2p(True)
        elif self.mode == "sync":
            sync_exchange(self._gradbuf, self.pg)
            n += self._seg_b_sync(True) if first else self._run("b", lambda: self._seg_b_sync(False))
        else:   # async: stale-gradient pushes onto shared parameters, then pull
            self.old_params.copy_(self.params)
            n += self._run("b", lambda: self._apply(1.0))
            async_push_pull(self.params, self.old_params, self.pg)
            n += self._run("c", self._seg_c_async)
        self.iterations += 1
        self.launches_per_iteration = n
        return n

    # ------------------------------------------------------------------ host-side views (synchronise)
    def _row_dict(self, st, op):
        # sync multi-rank: the lr rule of update i runs at the start of update i+1, so opt[ACTOR_LR] of the row is the
        # lr update i used; otherwise the rule has already run and stats[ACTOR_LR_USED] holds it
        deferred = self.world > 1 and self.mode == "sync"
        return dict(act_loss=float(st[L.STAT_ACTOR_LOSS]), crit_loss=float(st[L.STAT_CRITIC_LOSS]),
                    kl_dist=float(st[L.STAT_KL_POST]), kl_pre=float(st[L.STAT_KL_PRE]),
                    avg_ent=float(st[L.STAT_ENT_MEAN]),
                    act_lr=float(op[L.OPT_ACTOR_LR] if deferred else st[L.STAT_ACTOR_LR_USED]),
                    next_lr=float(op[L.OPT_ACTOR_LR]), beta=float(op[L.OPT_BETA_KL]),
                    gamma_ent=float(op[L.OPT_GAMMA_ENT]), adv_mean=float(st[L.STAT_ADV_MEAN]),
                    adv_std=float(st[L.STAT_ADV_STD]), iteration=int(op[L.OPT_ITER]), adam_t=int(op[L.OPT_ADAM_T]),
                    ev_before=float(st[L.STAT_EV_BEFORE]),
                    ev_after=float(st[L.STAT_EV_AFTER]) if self.diagnostics else -1.0,
                    avg_rew=float(st[L.STAT_EP_REW_MEAN]), episodes=int(st[L.STAT_EP_COUNT]))

    def read_stats(self):
        """Device -> host copy of the stats/opt blocks as a dict (the values the reference logs)."""
        return self._row_dict(self.stats.cpu().numpy(), self.opt.cpu().numpy())

    def read_log(self, first, count):
        """The log rows of iterations first .. first+count-1 (count <= log_rows, all already run) from the device
        ring: ONE device -> host copy however many iterations it covers."""
        assert 0 < count <= self.log_rows and first + count <= self.iterations
        ring = self.log_ring.cpu().numpy()
        return [self._row_dict(ring[i % self.log_rows, :16], ring[i % self.log_rows, 16:]) for i in
                range(first, first + count)]

    def env_steps_per_iteration(self):
        return self.N * self.T
