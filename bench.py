#!/usr/bin/env python
"""Benchmark of the actor-critic hot path (BASELINE.json metric: env-steps/sec, rollout+update).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A step = one rollout + update iteration over the synthetic workload of BASELINE.json configs[2]:
Pendulum-v0 A2C, 65,536 envs x 20-step horizon PER GPU (weak scaling; at N > 1 the flat gradient
is all-reduced over NCCL).  Prints ONE JSON line on rank 0 (contract in the task statement).

  value     device-resident throughput: env state, weights and RNG live in HBM; CUDA-event time per
            step, L2 flushed between timed steps, max over ranks.
  e2e       same metric through the public trainer API with HOST buffers: every step copies one
            step's action noise and reset uniforms from pinned host memory (H2D; the loader runs one
            step ahead on a copy stream, trainer.prefetch) and reads the stats block and the updated
            parameters back (D2H), all inside the timed region.
  roofline  for the dominant kernel (fused loss+backward of the critic MLP) plus a per-kernel table
            (`kernels`): algorithmic FLOPs or bytes per launch / CUDA-event duration of that kernel
            timed alone in this run.
  cpu_baseline / --impl reference
            the numpy restatement of the reference's math (oracle/a2c.py; TF 1.3 + gym cannot be
            installed here -- BASELINE.md) on the host cores.  cpu_baseline of the default line: a
            bounded sample (2,048 envs).  --impl reference: FULL-SIZE iterations (65,536 envs x 20
            steps, the GPU arm's config), up to --steps of them within a 120 s timed region.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

ENV_ID, N_ENVS, HORIZON, SEED = "Pendulum-v0", 65536, 20, 12321
WORKLOAD = "Pendulum-v0 A2C, 65536 envs x 20-step horizon per GPU, fused rollout+update (BASELINE.json configs[2])"
FLOP_ACTOR_FWD, FLOP_CRITIC_FWD = 50048.0, 100096.0       # per sample, 2 FLOP/MAC
# --workload: the default line is configs[2]; the other two are extra, separately reported measurements
WORKLOADS = {
    "pendulum": dict(env_id="Pendulum-v0", n_envs=65536, horizon=20, name=WORKLOAD, flop_actor=50048.0,
                     flop_critic=100096.0),
    "cartpole": dict(env_id="CartPole-v0", n_envs=262144, horizon=20,
                     name="CartPole-v0 A2C, 262144 envs x 20-step horizon per GPU, fused rollout+update, NCCL "
                          "flat-gradient all-reduce (BASELINE.json configs[3]; the horizon is not given there: 20, as "
                          "in configs[2])", flop_actor=50432.0, flop_critic=100608.0),
}
FP32_PEAK_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12         # FFMA peak at max boost clock (computed)


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), float(d.get("bf16_tflops", 1590.0)), "measured"
    return 6650.0, 1590.0, "fallback"


class ClockSampler(object):
    """nvidia-smi clocks / throttle reasons streamed (-lms 100) during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc, self.thread = index, [], None, None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
            return

        def pump():
            for line in self.proc.stdout:
                self.rows.append((time.perf_counter(), [x.strip() for x in line.strip().split(",")]))
        self.thread = threading.Thread(target=pump, daemon=True)
        self.thread.start()
        time.sleep(0.35)            # let the first samples arrive before the timed region starts

    def stop(self):
        if self.proc is not None:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()

    def summary(self, t0=None, t1=None):
        rows = [r for t, r in self.rows if (t0 is None or t >= t0 - 0.1) and (t1 is None or t <= t1 + 0.1)] or \
               [r for _, r in self.rows]
        sm = sorted(int(r[0]) for r in rows if len(r) >= 7 and r[0].isdigit())
        mx = [int(r[1]) for r in rows if len(r) >= 7 and r[1].isdigit()]
        pw = [float(r[2]) for r in rows if len(r) >= 7 and r[2].replace(".", "", 1).isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 7 and r[3 + i].lower() == "active" for r in rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm), "power_w_max": max(pw) if pw else None}


def host_threads():
    """Give the numpy/BLAS arithmetic of the CPU arms every host core, explicitly: torchrun exports OMP_NUM_THREADS=1
    to its children, which would quietly make the reference arm single-threaded.  Must run before numpy is imported."""
    cores = os.cpu_count() or 1
    for k in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[k] = str(cores)
    return cores


def blas_threads():
    try:
        from threadpoolctl import threadpool_info
        return max([int(x.get("num_threads", 1)) for x in threadpool_info()] or [1])
    except Exception:
        return None


def cpu_baseline(steps=2, n_envs=2048, warmup=1):
    """Oracle A2C iteration (numpy, all host threads through BLAS) on n_envs x HORIZON env-steps per iteration."""
    from oracle import a2c, envs as E
    S = a2c.A2CState(E.ENV_IDS[ENV_ID], n_envs, seed=SEED)
    for _ in range(warmup):
        a2c.iteration(S, T=HORIZON)
    t0 = time.perf_counter()
    for _ in range(steps):
        a2c.iteration(S, T=HORIZON)
    dt = time.perf_counter() - t0
    return n_envs * HORIZON * steps / dt, dt / steps, "%d envs x %d steps x %d iterations of the same workload" % (
        n_envs, HORIZON, steps)


def b_ref_baselines(seconds=3.0):
    """The reference's OWN loop shape on the host cores (oracle/ref_loop.py; BASELINE.md section 3 "B-ref"): batch-1 act
    per env step, whole episodes, PathAdv per episode, one update per >= 1,200 (A3C) / >= 800 or 7 episodes (Basic_AC)
    steps -- without TensorFlow's per-call dispatch, i.e. an upper bound on the real reference."""
    from oracle import ref_loop as RL
    out = {}
    for key, env, workers, basic in (("a3c_pendulum_1_worker", "Pendulum-v0", 1, False),
                                     ("a3c_pendulum_readme_2_workers", "Pendulum-v0", 2, False),
                                     ("a3c_pendulum_runner_sh_3_workers", "Pendulum-v0", 3, False),
                                     ("basic_ac_cartpole_1_process", "CartPole-v0", 1, True)):
        v, n = RL.run_workers(env, workers, basic, seconds=seconds, seed=SEED)
        out[key] = {"value": v, "unit": "env-steps/s", "processes": workers, "env_steps_timed": n,
                    "config": "BASELINE.json configs[%d]" % (0 if basic else 1)}
    return out


def ref_shape_gpu(iters=60):
    """BASELINE.json configs[0] / configs[1] on the GPU path at the REFERENCE's batch shape (about 800 / 1,200 env steps
    per update), next to cpu_baseline.b_ref: a batch this small is launch-latency-bound on a B200 -- the numbers say what
    the drop-in costs a user who changes nothing else, not what the hardware can do."""
    import torch
    from actor_critic_algs_on_tensorflow_b200.trainer import A2CTrainer
    out = {}
    for key, kw, steps in (("basic_ac_cartpole_8_envs_x_100_steps", dict(env_id="CartPole-v0", n_envs=8, horizon=100, basic=True,
                                                                           clip=1.0, max_lr=1.0), 800),
                           ("a3c_pendulum_8_envs_x_150_steps", dict(env_id="Pendulum-v0", n_envs=8, horizon=150), 1200)):
        tr = A2CTrainer(seed=SEED, **kw)
        for _ in range(4):
            tr.iteration()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            tr.iteration()
        b.record()
        torch.cuda.synchronize()
        sec = a.elapsed_time(b) * 1e-3 / iters
        out[key] = {"value": steps / sec, "unit": "env-steps/s", "ms_per_update": sec * 1e3,
                    "launches_per_update": tr.launches_per_iteration}
        del tr
    return out


def run_reference(args, rank):
    """--impl reference: the numpy restatement of the reference's math on the box's host cores, on THIS arm's config
    (65,536 envs x 20 steps per iteration, full size, 10-15 s each on 16 cores).  Warm-up: --warmup iterations on a
    4,096-env sample (BLAS threads, allocator; the state they leave is discarded).  Timed: --steps full-size
    iterations, stopping early (never before 2) once the next one would carry the timed region past 120 s so that the
    whole run stays within a few minutes: "steps" of the line is the number actually timed, "steps_requested" is
    --steps."""
    if rank != 0:
        return
    from oracle import a2c, envs as E
    cores = os.cpu_count()
    cpu_baseline(steps=max(1, args.warmup), n_envs=4096, warmup=0)
    S = a2c.A2CState(E.ENV_IDS[ENV_ID], N_ENVS, seed=SEED)
    budget, times = 120.0, []
    want = max(1, args.steps)
    while len(times) < want:
        t0 = time.perf_counter()
        a2c.iteration(S, T=HORIZON)
        times.append(time.perf_counter() - t0)
        if len(times) >= 2 and sum(times) + times[-1] > budget:
            break
    k, sec = len(times), sum(times) / len(times)
    v = N_ENVS * HORIZON / sec
    sample = "%d envs x %d steps x %d iterations: the full workload" % (N_ENVS, HORIZON, k)
    bref = b_ref_baselines()
    line = {"impl": "reference", "metric": "env-steps/sec (rollout+update)", "value": v, "unit": "env-steps/s",
            "n_gpus": args.gpus, "steps": k, "warmup": args.warmup, "ms_per_step": sec * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "envs_per_gpu": N_ENVS, "horizon": HORIZON, "mode": args.mode},
            "same_config": True, "steps_requested": args.steps,
            "cpu_baseline": {"value": v, "unit": "env-steps/s", "cores": cores, "blas_threads": blas_threads(),
                             "kind": "port", "sample": sample, "warmup_sample": "4096 envs x %d steps" % HORIZON,
                             "b_ref": bref},
            "e2e": {"value": v, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "numpy restatement of the reference's math (oracle/a2c.py), vectorised over the env batch: TF 1.3 and "
                    "gym are not installable here (BASELINE.md section 2); faster than the real reference, whose own "
                    "shape (batch-1 sess.run per env step) is timed as cpu_baseline.b_ref"}
    print(json.dumps(line), flush=True)


def time_kernel(fn, flush, reps=5):
    import torch
    fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(reps):
        flush()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        tot += a.elapsed_time(b)
    return tot / reps * 1e-3


def kernel_table(tr, flush, hbm_peak):
    """Per-kernel achieved FLOP/s or GB/s, each kernel timed alone on the trainer's own buffers."""
    from actor_critic_algs_on_tensorflow_b200 import _lib as L
    N, T, B = tr.N, tr.T, tr.N * tr.T
    s = L.cur_stream()
    Pa = tr.Pa
    ga, gc = tr.grads[:Pa], tr.grads[Pa:]
    common = (L.ptr(tr.obs_buf), L.ptr(tr.acs), L.ptr(tr.adv), L.ptr(tr.stats), L.ptr(tr.logps), L.ptr(tr.G),
              L.ptr(tr.opt), L.ptr(ga), L.ptr(gc), L.ptr(tr.stats), L.ptr(tr.ws_bwd), tr.ws_bwd_bytes, s)
    ks = {}

    def bwd(actor, critic):
        return lambda: L.call("acx_loss_backward_gauss", actor, critic, tr.ob, tr.num_ac, tr.ac_scale, T, N, *common)
    ks["loss_backward_critic (K4)"] = (bwd(None, L.ptr(tr.critic)), 3 * FLOP_CRITIC_FWD * B, None)
    ks["loss_backward_actor (K4)"] = (bwd(L.ptr(tr.actor), None), 3 * FLOP_ACTOR_FWD * B, None)
    ks["value_forward (K2)"] = (lambda: L.call("acx_value_forward", L.ptr(tr.critic), tr.ob, T + 1, N,
                                               L.ptr(tr.obs_buf), L.ptr(tr.V), 0, s), FLOP_CRITIC_FWD * (T + 1) * N, None)
    ks["policy_forward_sample (K2)"] = (lambda: L.call(
        "acx_policy_forward_sample_gauss", L.ptr(tr.actor), tr.ob, tr.num_ac, tr.ac_scale, 1, N, L.ptr(tr.obs_buf), None,
        tr.seed, 0, 0, None, L.ptr(tr.acs), L.ptr(tr.logps), L.ptr(tr.ent), None, s), FLOP_ACTOR_FWD * N, None)
    ks["policy_logp_kl (K2/K5)"] = (lambda: L.call(
        "acx_policy_logp_gauss", L.ptr(tr.actor), tr.ob, tr.num_ac, tr.ac_scale, T, N, L.ptr(tr.obs_buf), L.ptr(tr.acs),
        None, None, None, L.ptr(tr.logps), L.ptr(tr.kl_part), __import__("ctypes").byref(tr.kl_blocks), s),
        FLOP_ACTOR_FWD * B, None)
    if getattr(tr, "cache", False):
        cached = common[:-3] + (L.ptr(tr.ax2), L.ptr(tr.ax3), L.ptr(tr.cx2), L.ptr(tr.cx3)) + common[-3:]

        def bwd_c(actor, critic):
            return lambda: L.call("acx_loss_backward_gauss_cached", actor, critic, tr.ob, tr.num_ac, tr.ac_scale, T, N,
                                  *cached)
        # algorithmic FLOPs of the cached variant: backward only (2x forward) + layer 1 / head
        ks["loss_backward_critic, cached activations (K4)"] = (bwd_c(None, L.ptr(tr.critic)), 2 * FLOP_CRITIC_FWD * B, None)
        ks["loss_backward_actor, cached activations (K4)"] = (bwd_c(L.ptr(tr.actor), None), 2 * FLOP_ACTOR_FWD * B, None)
        ks["value_forward + cache write (K2)"] = (lambda: L.call(
            "acx_value_forward_cached", L.ptr(tr.critic), tr.ob, T + 1, N, L.ptr(tr.obs_buf), L.ptr(tr.V), 0,
            L.ptr(tr.cx2), L.ptr(tr.cx3), s), FLOP_CRITIC_FWD * (T + 1) * N, None)
        ks["policy_forward_sample + cache write (K2)"] = (lambda: L.call(
            "acx_policy_forward_sample_gauss_cached", L.ptr(tr.actor), tr.ob, tr.num_ac, tr.ac_scale, 1, N,
            L.ptr(tr.obs_buf), None, tr.seed, 0, 0, None, L.ptr(tr.acs), L.ptr(tr.logps), L.ptr(tr.ent), None,
            L.ptr(tr.ax2), L.ptr(tr.ax3), s), FLOP_ACTOR_FWD * N, None)
    ks["value_forward_tc (K2, tcgen05 3xTF32)"] = (lambda: L.call(
        "acx_value_forward_tc", L.ptr(tr.critic), L.ptr(tr.packed_c), tr.ob, T + 1, N, L.ptr(tr.obs_buf), L.ptr(tr.V), 0,
        L.ptr(tr.cx2) if tr.cache else None, L.ptr(tr.cx3) if tr.cache else None, L.ptr(getattr(tr, "cmask", None)), s),
        FLOP_CRITIC_FWD * (T + 1) * N, "tc")
    ks["policy_forward_sample_tc (K2, tcgen05 3xTF32)"] = (lambda: L.call(
        "acx_policy_forward_sample_gauss_tc", L.ptr(tr.actor), L.ptr(tr.packed_a), tr.ob, tr.num_ac, tr.ac_scale, 1, N,
        L.ptr(tr.obs_buf), None, tr.seed, 0, 0, None, L.ptr(tr.acs), L.ptr(tr.logps), L.ptr(tr.ent), None,
        L.ptr(tr.ax2) if tr.cache else None, L.ptr(tr.ax3) if tr.cache else None, s),
        FLOP_ACTOR_FWD * N, "tc")
    if getattr(tr, "fused_rollout", False):
        st3, el3 = tr.state.clone(), tr.elapsed.clone()
        ks["rollout_tc, %d steps in one launch (K1 + K2, tcgen05 3xTF32)" % T] = (lambda: L.call(
            "acx_rollout_tc", tr.env, L.ptr(tr.actor), L.ptr(tr.packed_a), tr.ac_scale, T, N, L.ptr(tr.obs_buf), L.ptr(st3),
            L.ptr(el3), None, None, tr.seed, 0, None, tr.max_steps, L.ptr(tr.acs), L.ptr(tr.logps), L.ptr(tr.ent),
            L.ptr(tr.rew), L.ptr(tr.done), L.ptr(tr.ax2) if tr.cache else None, L.ptr(tr.ax3) if tr.cache else None,
            L.ptr(getattr(tr, "amask", None)), s), FLOP_ACTOR_FWD * B, "tc")
    ks["policy_logp_kl_tc (K2/K5, tcgen05 3xTF32)"] = (lambda: L.call(
        "acx_policy_logp_gauss_tc", L.ptr(tr.actor), L.ptr(tr.packed_a), tr.ob, tr.num_ac, tr.ac_scale, T, N,
        L.ptr(tr.obs_buf), L.ptr(tr.acs), None, None, None, L.ptr(tr.logps), L.ptr(tr.kl_part),
        __import__("ctypes").byref(tr.kl_blocks), s), FLOP_ACTOR_FWD * B, "tc")
    if getattr(tr, "tc_bwd", False):
        base_bytes = L.load().acx_loss_backward_workspace_bytes(tr.ob, tr.num_ac, int(tr.cont))

        def bwd_t(actor, critic, nbytes):
            args = cached[:-3] + (L.ptr(getattr(tr, "amask", None)), L.ptr(getattr(tr, "cmask", None))) + (cached[-3], nbytes, s)
            return lambda: L.call("acx_loss_backward_gauss_tc", actor, critic, L.ptr(tr.packed_bwd_a),
                                  L.ptr(tr.packed_bwd_c), tr.ob, tr.num_ac, tr.ac_scale, T, N, *args)
        # one kernel: chain products on tcgen05, weight-gradient products FFMA (small workspace)
        ks["loss_backward_critic_tc, one kernel (K4, tcgen05 dX + FFMA dW)"] = (
            bwd_t(None, L.ptr(tr.critic), base_bytes), 2 * FLOP_CRITIC_FWD * B, None)
        ks["loss_backward_actor_tc, one kernel (K4, tcgen05 dX + FFMA dW)"] = (
            bwd_t(L.ptr(tr.actor), None, base_bytes), 2 * FLOP_ACTOR_FWD * B, None)
        if tr.ws_bwd_bytes > base_bytes:
            # two kernels: chain kernel + acx_wgrad_kernel, every product on tcgen05 (3xTF32)
            ks["loss_backward_critic_tc, chain + wgrad kernels (K4, tcgen05 3xTF32)"] = (
                bwd_t(None, L.ptr(tr.critic), tr.ws_bwd_bytes), 2 * FLOP_CRITIC_FWD * B, "tc")
            ks["loss_backward_actor_tc, chain + wgrad kernels (K4, tcgen05 3xTF32)"] = (
                bwd_t(L.ptr(tr.actor), None, tr.ws_bwd_bytes), 2 * FLOP_ACTOR_FWD * B, "tc")
    import torch
    st2, el2 = tr.state.clone(), tr.elapsed.clone()
    ks["env_step pendulum (K1)"] = (lambda: L.call(
        "acx_env_step", tr.env, N, L.ptr(st2), L.ptr(el2), L.ptr(tr.acs), L.ptr(tr.obs_buf[1]), L.ptr(tr.rew),
        L.ptr(tr.done), None, tr.seed, 0, 0, None, tr.max_steps, s), None, 45.0 * N)
    G2, A2 = torch.empty_like(tr.G), torch.empty_like(tr.adv)
    ks["returns_adv_scan (K3)"] = (lambda: L.call(
        "acx_returns_adv_scan", T, N, L.ptr(tr.rew), L.ptr(tr.done), L.ptr(tr.V), tr.gamma, tr.look_ahead, L.ptr(G2),
        L.ptr(A2), L.ptr(tr.stats), L.ptr(tr.ws_scan), tr.ws_scan_bytes, 0, s), None, 17.0 * B + 4.0 * N)
    m2, v2, p2 = tr.m.clone(), tr.v.clone(), tr.params.clone()
    ks["adam_apply (K5)"] = (lambda: L.call(
        "acx_adam_apply", L.ptr(p2[:Pa]), L.ptr(ga), L.ptr(m2[:Pa]), L.ptr(v2[:Pa]), Pa, L.ptr(p2[Pa:]), L.ptr(gc),
        L.ptr(m2[Pa:]), L.ptr(v2[Pa:]), tr.Pc, L.ptr(tr.opt), 1.0, s), None, 28.0 * (tr.Pa + tr.Pc))
    # the HBM-bound kernels move only 3-22 MB at 65,536 envs (launch-latency bound): also time them at the
    # 1,048,576-env size of BASELINE configs[4] (env kernels) / a 1M-column scan
    NB = 1 << 20
    from actor_critic_algs_on_tensorflow_b200.envs import VecEnv
    # A 1M-env launch lasts 10-15 us, of which a lone launch between two events spends ~5 us waiting for the launch
    # itself: the 1M rows therefore time EIGHT launches back to back on eight separate sets of buffers (0.4-0.6 GB in
    # all, so every launch still finds its data in HBM, not in the 126 MB L2) and report the time per launch.
    inner = {}
    for env_name, nb, nenv, sets in (("Pendulum-v0", 45.0, NB, 8), ("CartPole-v0", 65.0, NB, 8), ("Acrobot-v1", 73.0, NB, 8),
                                     ("Pendulum-v0", 45.0, 8 * NB, 1), ("CartPole-v0", 65.0, 8 * NB, 1),
                                     ("Acrobot-v1", 73.0, 8 * NB, 1)):
        ves = [VecEnv(env_name, nenv, seed=1 + i) for i in range(sets)]
        for ve in ves:
            ve.reset()
        act = (torch.zeros(nenv, device="cuda") if ves[0].n_actions == 0 else torch.ones(nenv, device="cuda", dtype=torch.int32))
        key = "env_step %s, %dM envs (K1)" % (env_name, nenv >> 20)
        if sets > 1:          # one CUDA graph of the eight launches: the host (ctypes from Python) is slower than the kernels
            for ve in ves:
                ve.step(act)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                for ve in ves:
                    ve.step(act)
            ks[key] = ((lambda g=g, keep=(ves, act): g.replay()), None, nb * nenv)      # the graph owns no buffer
        else:
            ks[key] = ((lambda ve=ves[0], act=act: ve.step(act)), None, nb * nenv)
        inner[key] = sets
    rB, dB = torch.randn((T, NB), device="cuda"), torch.zeros((T, NB), device="cuda", dtype=torch.uint8)
    vB, gB, aB = torch.randn((T + 1, NB), device="cuda"), torch.empty((T, NB), device="cuda"), torch.empty((T, NB), device="cuda")
    wsB = torch.zeros(L.load().acx_returns_adv_workspace_bytes(T, NB), device="cuda", dtype=torch.uint8)
    ks["returns_adv_scan, 1M columns (K3)"] = (lambda: L.call(
        "acx_returns_adv_scan", T, NB, L.ptr(rB), L.ptr(dB), L.ptr(vB), tr.gamma, tr.look_ahead, L.ptr(gB), L.ptr(aB),
        L.ptr(tr.stats), L.ptr(wsB), wsB.numel(), 0, s), None, 17.0 * T * NB + 4.0 * NB)
    NB8 = 8 << 20
    r8, d8 = torch.randn((T, NB8), device="cuda"), torch.zeros((T, NB8), device="cuda", dtype=torch.uint8)
    v8, g8, a8 = torch.randn((T + 1, NB8), device="cuda"), torch.empty((T, NB8), device="cuda"), torch.empty((T, NB8), device="cuda")
    ws8 = torch.zeros(L.load().acx_returns_adv_workspace_bytes(T, NB8), device="cuda", dtype=torch.uint8)
    ks["returns_adv_scan, 8M columns (K3)"] = (lambda: L.call(
        "acx_returns_adv_scan", T, NB8, L.ptr(r8), L.ptr(d8), L.ptr(v8), tr.gamma, tr.look_ahead, L.ptr(g8), L.ptr(a8),
        L.ptr(tr.stats), L.ptr(ws8), ws8.numel(), 0, s), None, 17.0 * T * NB8 + 4.0 * NB8)
    out = {}
    for name, (fn, flops, nbytes) in ks.items():
        sec = time_kernel(fn, flush) / inner.get(name, 1)
        if nbytes == "tc":
            # algorithmic FLOPs (one product per MAC) against TF32 dense peak / 3: every product costs three MMAs
            ach = flops / sec / 1e12
            peak = kernel_table.bf16_peak / 2.0 / 3.0
            out[name] = {"bound": "tensor", "ms": sec * 1e3, "achieved": ach, "peak": peak, "unit": "TFLOP/s",
                         "frac": ach / peak, "peak_note": "measured bf16 peak / 2 (tf32) / 3 (3xTF32 split)",
                         "frac_of_fp32_ffma_peak": ach / FP32_PEAK_TFLOPS}
        elif flops is not None:
            ach = flops / sec / 1e12
            out[name] = {"bound": "fp32", "ms": sec * 1e3, "achieved": ach, "peak": FP32_PEAK_TFLOPS, "unit": "TFLOP/s",
                         "frac": ach / FP32_PEAK_TFLOPS}
        else:
            ach = nbytes / sec / 1e9
            out[name] = {"bound": "hbm", "ms": sec * 1e3, "achieved": ach, "peak": hbm_peak, "unit": "GB/s",
                         "frac": ach / hbm_peak, "algorithmic_bytes": nbytes}
            if inner.get(name, 1) > 1:
                out[name]["timing"] = "%d launches back to back (one CUDA graph) on separate buffer sets, time per launch" % inner[name]
    return out


def critic_pair_only(tr, flush, bf16_peak):
    """Non-default workloads: only the dominant kernel pair (critic chain + weight-gradient kernels), timed alone."""
    from actor_critic_algs_on_tensorflow_b200 import _lib as L
    N, T = tr.N, tr.T
    ga, gc = tr.grads[:tr.Pa], tr.grads[tr.Pa:]
    tail = (T, N, L.ptr(tr.obs_buf), L.ptr(tr.acs), L.ptr(tr.adv), L.ptr(tr.stats), L.ptr(tr.logps), L.ptr(tr.G),
            L.ptr(tr.opt), L.ptr(ga), L.ptr(gc), L.ptr(tr.stats), L.ptr(tr.ax2), L.ptr(tr.ax3), L.ptr(tr.cx2),
            L.ptr(tr.cx3), L.ptr(tr.amask), L.ptr(tr.cmask), L.ptr(tr.ws_bwd), tr.ws_bwd_bytes, L.cur_stream())
    head = (None, L.ptr(tr.critic), L.ptr(tr.packed_bwd_a), L.ptr(tr.packed_bwd_c), tr.ob)
    if tr.cont:
        fn = lambda: L.call("acx_loss_backward_gauss_tc", *head, tr.num_ac, tr.ac_scale, *tail)
    else:
        fn = lambda: L.call("acx_loss_backward_cat_tc", *head, tr.n_act, *tail)
    sec = time_kernel(fn, flush)
    ach = 2 * FLOP_CRITIC_FWD * N * T / sec / 1e12
    peak = bf16_peak / 2.0 / 3.0
    return {"loss_backward_critic_tc, chain + wgrad kernels (K4, tcgen05 3xTF32)": {
        "bound": "tensor", "ms": sec * 1e3, "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
        "peak_note": "measured bf16 peak / 2 (tf32) / 3 (3xTF32 split)", "frac_of_fp32_ffma_peak": ach / FP32_PEAK_TFLOPS}}


def run_ours(args, rank, local_rank, world):
    import torch
    import torch.distributed as dist
    from actor_critic_algs_on_tensorflow_b200.trainer import A2CTrainer

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    hbm_peak, bf16_peak, peaks_src = measured_peaks()
    tr = A2CTrainer(ENV_ID, n_envs=N_ENVS, horizon=HORIZON, seed=SEED, rank=rank, world=world, mode=args.mode,
                    use_graph=not args.no_graph, tensor_cores=not args.ffma_only, exchange=args.exchange)
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")       # > 126 MB L2
    clean_buf = torch.zeros(64 << 20, dtype=torch.float32, device="cuda")      # 256 MiB, only ever read

    def flush():
        # write 256 MiB (evicts everything), then read 256 MiB so that the lines left in L2 are clean: otherwise
        # the timed kernel also pays for writing back up to 126 MB of the flush buffer's dirty lines
        flush_buf.fill_(1)
        clean_buf.sum()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(3, args.warmup)):          # >= 3 warm-up steps (also builds the CUDA graphs)
        tr.iteration()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    evs = []
    barrier()
    t_begin = time.perf_counter()
    for _ in range(args.steps):
        flush()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        launches = tr.iteration()
        b.record()
        evs.append((a, b))
    barrier()
    t_end = time.perf_counter()
    if rank == 0:
        sampler.stop()
    sec = sum(a.elapsed_time(b) for a, b in evs) * 1e-3
    t = torch.tensor([sec], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    sec = float(t.item())
    steps_per_it = N_ENVS * HORIZON
    value = world * steps_per_it * args.steps / sec
    stats = tr.read_stats()

    # ---- end to end through the public API with host buffers
    eps_h = (torch.randn((HORIZON, 1, N_ENVS), dtype=torch.float32) if tr.cont
             else torch.rand((HORIZON, N_ENVS), dtype=torch.float32)).pin_memory()
    ru_h = torch.rand((HORIZON, tr.S, N_ENVS), dtype=torch.float32).pin_memory()
    stats_h = torch.empty(32, dtype=torch.float32).pin_memory()
    params_h = torch.empty(tr.Pa + tr.Pc, dtype=torch.float32).pin_memory()

    def e2e_step():
        # the loader runs one batch ahead: this step's inputs were staged (H2D from pinned memory, on the trainer's
        # copy stream, double-buffered) while the previous step computed; the copy of the NEXT step's inputs starts
        # here and has to end before this step's closing event -- one full H2D copy inside every timed interval
        tr.iteration(prefetched=True)
        tr.prefetch(eps_h, ru_h)
        stats_h[:16].copy_(tr.stats, non_blocking=True)            # D2H: losses / KL / lr ...
        stats_h[16:].copy_(tr.opt, non_blocking=True)
        params_h.copy_(tr.params, non_blocking=True)               # ... and the updated weights
        tr.wait_prefetch()
    tr.prefetch(eps_h, ru_h)
    for _ in range(5):                                             # one eager + one capture per staging set
        e2e_step()
    barrier()
    e_evs = []
    for _ in range(args.steps):
        flush()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        e2e_step()
        b.record()
        e_evs.append((a, b))
    barrier()
    e_sec = sum(a.elapsed_time(b) for a, b in e_evs) * 1e-3
    t = torch.tensor([e_sec], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e_sec = float(t.item())
    h2d = eps_h.numel() * 4 + ru_h.numel() * 4
    d2h = stats_h.numel() * 4 + params_h.numel() * 4

    if rank == 0:
        kernel_table.bf16_peak = bf16_peak
        kernels = kernel_table(tr, flush, hbm_peak) if args.workload == "pendulum" else critic_pair_only(tr, flush, bf16_peak)
        base_bytes_ = __import__("actor_critic_algs_on_tensorflow_b200._lib", fromlist=["x"]).load() \
            .acx_loss_backward_workspace_bytes(tr.ob, tr.num_ac, int(tr.cont))
        if getattr(tr, "tc_bwd", False) and tr.ws_bwd_bytes > base_bytes_:
            dom_name = "loss_backward_critic_tc, chain + wgrad kernels (K4, tcgen05 3xTF32)"
            dom_kernel = ("acx_mlp_tc_chain_kernel<critic> + acx_wgrad_kernel<critic, dW3> + <critic, dW2> (fused loss+backward on "
                          "cached activations, K4)")
            note = "every product of the pair runs on tcgen05 as 3xTF32"
        elif getattr(tr, "tc_bwd", False):
            dom_name = "loss_backward_critic_tc, one kernel (K4, tcgen05 dX + FFMA dW)"
            dom_kernel = "acx_mlp_tc_train_kernel<critic> (fused loss+backward on cached activations, K4)"
            note = ("the weight-gradient half of the kernel's products is fp32 FFMA and bounds it; the chain half runs "
                    "concurrently on tcgen05 (3xTF32)")
        elif tr.cache:
            dom_name = "loss_backward_critic, cached activations (K4)"
            dom_kernel = "acx_mlp_kernel<critic, train, cached> (fused loss+backward, K4)"
            note = "the kernel is fp32 FFMA, not a tensor-core kernel"
        else:
            dom_name = "loss_backward_critic (K4)"
            dom_kernel = "acx_mlp_kernel<critic, train> (fused loss+backward, K4)"
            note = "the kernel is fp32 FFMA, not a tensor-core kernel"
        dom = kernels[dom_name]
        step_ms = sec / args.steps * 1e3
        if dom["bound"] == "tensor":
            roof = {"kernel": dom_kernel, "bound": "tensor", "achieved": dom["achieved"], "peak": dom["peak"],
                    "unit": "TFLOP/s", "frac": dom["frac"],
                    "peak_source": "measured bf16 dense peak (%s: %.0f TFLOP/s) / 2 (tf32) / 3 (three MMAs per product: "
                                   "3xTF32 split for fp32-class accuracy); achieved = algorithmic FLOPs (2 per MAC of "
                                   "the backward pass) per second; %s" % (peaks_src, bf16_peak, note),
                    "frac_of_fp32_ffma_peak": dom["achieved"] / FP32_PEAK_TFLOPS,
                    "share_of_step": dom["ms"] / step_ms,
                    # dram__bytes_read.sum + dram__bytes_write.sum of the group's three kernels per launch, from the ncu
                    # --set full captures of this workload taken in round 2 (profiles/r02_ncu_step.md: chain 0.756 + 1.289 GB,
                    # dW3 kernel 1.342 + 0.006 GB, dW2 kernel 0.687 + 0.004 GB) -- NOT measured in this run
                    "traffic": 4.084e9 if (N_ENVS, HORIZON) == (65536, 20) else None, "traffic_unit": "bytes per launch",
                    "traffic_source": "profiles/r02_ncu_step.md (ncu --set full, 2026-11-07; not re-measured in this run)",
                    "algorithmic_bytes": 4.09e9 if (N_ENVS, HORIZON) == (65536, 20) else None,
                    # the same captures: sm__pipe_tc_cycles_active (the MMA instruction rate measured by
                    # tests/tc/run_mma_rate.py, 2,007 MAC/clk/SM, is above measured-bf16 / 2, so `frac` flatters the pipe)
                    "tensor_pipe_active_ncu": {"chain": 0.393, "wgrad_dW3": 0.555, "wgrad_dW2": 0.562},
                    "hbm_peak_gbs": hbm_peak, "hbm_peak_source": peaks_src}
        else:
            roof = {"kernel": dom_kernel, "bound": "fp32",
                    "achieved": dom["achieved"], "peak": FP32_PEAK_TFLOPS, "unit": "TFLOP/s", "frac": dom["frac"],
                    "peak_source": "computed FFMA peak 148 SM x 128 lanes x 2 x 1.965 GHz; %s (bf16 tensor peak %s: %.0f "
                                   "TFLOP/s)" % (note, peaks_src, bf16_peak),
                    "share_of_step": dom["ms"] / step_ms, "traffic": None,
                    "hbm_peak_gbs": hbm_peak, "hbm_peak_source": peaks_src}
        total_flops = (FLOP_ACTOR_FWD * (1 + 3 + 1) + FLOP_CRITIC_FWD * (3 + (HORIZON + 1) / HORIZON)) * steps_per_it
        cb_v, cb_sec, cb_sample = (cpu_baseline() if world == 1 else (None, None, None))
        bref = b_ref_baselines(seconds=2.0) if world == 1 else None
        if bref is not None and args.workload == "pendulum":
            bref["gpu_same_shape"] = ref_shape_gpu()
        if world == 1:
            parallelism = "dp1"
        elif args.mode == "async":
            parallelism = "dp%d (envs sharded by rank; async: NCCL all-reduce of the parameter deltas)" % world
        elif tr.exchange == "p2p":
            parallelism = ("dp%d (envs sharded by rank; gradient exchange + lr rule + Adam as ONE kernel over NVLink peer "
                           "memory, acx_allreduce_adam, inside the step's CUDA graph)" % world)
        else:
            parallelism = "dp%d (envs sharded by rank; NCCL all-reduce of the flat gradient)" % world
        line = {"metric": "env-steps/sec (rollout+update)", "value": value, "unit": "env-steps/s", "n_gpus": world,
                "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": step_ms, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
                "config": {"workload": WORKLOAD, "envs_per_gpu": N_ENVS, "horizon": HORIZON, "mode": args.mode,
                           "l2": "flushed between timed steps (256 MiB write, then 256 MiB read to leave clean lines)", "weights": "glorot-uniform, seed 12321",
                           "rng": "device Philox4x32-10", "cuda_graphs": not args.no_graph,
                           "forward_engine": "tcgen05 3xTF32" if tr.tc else "fp32 FFMA", "backward_engine": ("tcgen05 3xTF32" if getattr(tr, "tc_bwd", False) and tr.N % 4 == 0
                                               else "tcgen05 3xTF32 (dX) + fp32 FFMA (dW)" if getattr(tr, "tc_bwd", False)
                                               else "fp32 FFMA"),
                           "parallelism": parallelism},
                "e2e": {"value": world * steps_per_it * args.steps / e_sec, "unit": "env-steps/s",
                        "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": e_sec / args.steps * 1e3,
                        "h2d": "pinned host buffers; the copy of step k+1's inputs overlaps step k's kernels on a copy "
                               "stream (double-buffered); every copy starts and ends inside a timed interval"},
                "gpu_launches": int(launches) * args.steps,
                "clocks": sampler.summary(t_begin, t_end),
                "roofline": roof, "kernels": kernels,
                "step_tflops": total_flops / (sec / args.steps) / 1e12,
                "train_stats": stats}
        if cb_v is not None:
            line["cpu_baseline"] = {"value": cb_v, "unit": "env-steps/s", "cores": os.cpu_count(),
                                    "blas_threads": blas_threads(), "kind": "port", "sample": cb_sample,
                                    "sec_per_iteration": cb_sec, "b_ref": bref}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def run_eval(args, rank, local_rank, world):
    """BASELINE configs[4]: greedy evaluation of the shipped Pendulum checkpoint, 1,048,576 episodes x 200 steps per GPU
    (episodes are independent: sharded by rank, no collective).  A step = one whole evaluation."""
    n_ep, horizon = 1 << 20, 200
    model = os.path.join(ROOT, "tests", "golden", "model-Pendulum_a3c")
    name = ("test_model.py greedy evaluation of demos/model-Pendulum_a3c, %d episodes x %d steps per GPU "
            "(BASELINE.json configs[4])" % (n_ep, horizon))
    if args.impl == "reference":
        if rank != 0:
            return
        import numpy as np
        from oracle import envs as E, policies as P, philox as R
        from oracle.demo_ckpt import read_demo_checkpoint
        t = read_demo_checkpoint(model)
        lay = P.actor_layout(3, 1, True)
        g = "global_actor/"
        actor = P.pack({"W1": t[g + "first_layer/kernel"], "b1": t[g + "first_layer/bias"],
                        "W2": t[g + "second_layer/kernel"], "b2": t[g + "second_layer/bias"],
                        "W3": t[g + "third_layer/kernel"], "b3": t[g + "third_layer/bias"],
                        "Wh": t[g + "mu_layer/kernel"], "bh": t[g + "mu_layer/bias"], "log_std": t[g + "log_std"]}, lay)
        n = 2048
        u = R.uniforms(np.arange(n, dtype=np.uint64), 0xFFFFFFFF, R.STREAM_RESET, SEED).T[:2]

        def one():
            s_ = E.reset_from_uniforms(E.PENDULUM, u.astype(np.float32), dtype=np.float32)
            tot = np.zeros(n, dtype=np.float32)
            for _ in range(horizon):
                _, _, _, mu = P.actor_act(actor, lay, E.observe(E.PENDULUM, s_, np.float32).T, np.zeros((n, 1), np.float32),
                                          True, 2.0, dtype=np.float32)
                s_, r, _ = E.dynamics(E.PENDULUM, s_, mu.T, dtype=np.float32)
                tot += r
            return tot
        one()
        t0 = time.perf_counter()
        k = max(1, min(args.steps, 5))
        for _ in range(k):
            ret = one()
        dt = time.perf_counter() - t0
        v = n * horizon * k / dt
        sample = "%d episodes x %d steps x %d evaluations" % (n, horizon, k)
        print(json.dumps({"impl": "reference", "metric": "env-steps/sec (greedy evaluation)", "value": v,
                          "unit": "env-steps/s", "n_gpus": args.gpus, "steps": k, "warmup": 1, "ms_per_step": dt / k * 1e3,
                          "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp32",
                          "data": "synthetic", "config": {"workload": name, "sample": sample},
                          "cpu_baseline": {"value": v, "unit": "env-steps/s", "cores": os.cpu_count(), "kind": "port",
                                           "sample": sample},
                          "e2e": {"value": v, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                          "gpu_launches": 0, "mean_return": float(ret.mean())}), flush=True)
        return
    import torch
    import torch.distributed as dist
    from actor_critic_algs_on_tensorflow_b200 import process, envs as envs_mod
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    actor = process.load_actor(model, 3, 1, 'cont', 2.0)
    params = actor.net.params
    start = rank * n_ep

    def resident():          # parameters resident; returns stay on the device
        vec = envs_mod.VecEnv("Pendulum-v0", n_ep, SEED, index0=start)
        return process._evaluate_chunk_rollout(vec, params, 1, True, True, SEED, start)

    def e2e():               # the public call: checkpoint file -> device, returns -> host
        return process.evaluate_batched("Pendulum-v0", model, n_ep, greedy=True, seed=SEED, tensor_cores=True)
    out = {}
    for key, fn in (("value", resident), ("e2e", e2e)):
        for _ in range(max(1, min(args.warmup, 3))):
            ret = fn()
        barrier()
        tot = 0.0
        steps = max(1, min(args.steps, 10))
        for _ in range(steps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            ret = fn()
            b.record()
            torch.cuda.synchronize()
            tot += a.elapsed_time(b) * 1e-3
        barrier()
        t = torch.tensor([tot], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out[key] = (world * n_ep * horizon * steps / float(t.item()), float(t.item()) / steps, ret)
    if rank == 0:
        ret = out["e2e"][2]
        steps = max(1, min(args.steps, 10))
        launches = (horizon // 20) * 2 + 3
        print(json.dumps({
            "metric": "env-steps/sec (greedy evaluation)", "value": out["value"][0], "unit": "env-steps/s",
            "n_gpus": world, "steps": steps, "warmup": max(1, min(args.warmup, 3)), "ms_per_step": out["value"][1] * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "config": {"workload": name, "episodes_per_gpu": n_ep, "steps_per_episode": horizon,
                       "engine": "acx_rollout_tc (tcgen05 3xTF32 forward + env step, 20 steps per launch)",
                       "l2": "working set per launch (obs/reward rows of 1M envs x 20 steps, 0.4 GB) exceeds L2"},
            "e2e": {"value": out["e2e"][0], "unit": "env-steps/s", "ms_per_step": out["e2e"][1] * 1e3,
                    "h2d_bytes_per_step": 25352 * 4, "d2h_bytes_per_step": n_ep * 4},
            "gpu_launches": launches * steps, "episodes_per_s": out["value"][0] / horizon,
            "mean_return": float(ret.mean()), "std_return": float(ret.std())}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--mode", choices=["sync", "async"], default="sync")
    ap.add_argument("--exchange", choices=["auto", "p2p", "nccl"], default="auto",
                    help="sync multi-GPU gradient exchange: p2p = fused peer-memory kernel, nccl = all_reduce")
    ap.add_argument("--no_graph", action="store_true", help="launch kernels eagerly (for ncu launch lists)")
    ap.add_argument("--ffma_only", action="store_true", help="keep the forward passes on the FFMA engine as well")
    ap.add_argument("--workload", choices=["pendulum", "cartpole", "eval"], default="pendulum",
                    help="pendulum = BASELINE configs[2] (the bench line); cartpole = configs[3]; eval = configs[4]")
    args = ap.parse_args()
    if args.workload in WORKLOADS:
        global ENV_ID, N_ENVS, HORIZON, WORKLOAD, FLOP_ACTOR_FWD, FLOP_CRITIC_FWD
        w = WORKLOADS[args.workload]
        ENV_ID, N_ENVS, HORIZON, WORKLOAD = w["env_id"], w["n_envs"], w["horizon"], w["name"]
        FLOP_ACTOR_FWD, FLOP_CRITIC_FWD = w["flop_actor"], w["flop_critic"]
    host_threads()                     # before numpy / torch load their BLAS: the CPU arms use every host core
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.gpus > 1 and world == 1 and args.impl == "ours":
        # convenience: re-launch ourselves under torchrun, one rank per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus),
               "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29517"), __file__] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    if args.workload == "eval":
        run_eval(args, rank, local_rank, world)
    elif args.impl == "reference":
        run_reference(args, rank)
    else:
        run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
