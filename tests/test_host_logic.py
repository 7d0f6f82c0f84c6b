"""Host-side mirrors of the reference's utilities against golden output of the reference's own classes,
CLI flag sets, and the world-size-2 (gloo) logic of the multi-rank update."""
import io
import os
import contextlib

import numpy as np
import pytest

from actor_critic_algs_on_tensorflow_b200 import util as U
from actor_critic_algs_on_tensorflow_b200.process import Framer


def test_logger_file_and_stdout_match_reference(golden_dir, tmp_path):
    path = str(tmp_path / "logs" / "worker_1.log")
    lg = U.Logger(logfile=path)
    with contextlib.redirect_stdout(io.StringIO()) as so:
        for t in range(5):
            lg(t, act_loss=0.1 * t - 0.3, circ_loss=1.5 + t, kl_dist=0.002 * t, avg_rew=-1000.0 + 13.37 * t,
               print_tog=(t % 2 == 0), act_lr=0.005, avg_ent=1.4189 - 0.01 * t, worker_id=1, ev_before=0.1 * t,
               ev_after=0.2 * t)
        lg.flush()
        for t in range(5, 8):
            lg(t, act_loss=0.5, circ_loss=2.5, kl_dist=0.01, avg_rew=-900.0, print_tog=False, act_lr=0.0075, avg_ent=1.3,
               worker_id=1)
        lg.flush()
    lg.f.close()
    assert open(path).read() == open(os.path.join(golden_dir, "logger_golden.log")).read()
    assert so.getvalue() == open(os.path.join(golden_dir, "logger_golden.stdout")).read()


def test_schedule_vaf_framer_match_reference(golden_dir):
    z = np.load(os.path.join(golden_dir, "schedule.npz"))
    g = U.LinearSchedule(init_t=100, end_t=3000, init_val=-2, end_val=-8, update_every_t=100)
    b = U.LinearSchedule(init_t=100, end_t=3000, init_val=0, end_val=-4, update_every_t=100)
    for t, gv, bv, up in zip(z["ts"], z["gamma"], z["beta"], z["upd"]):
        assert g.val(int(t)) == gv and b.val(int(t)) == bv and g.update_time(int(t)) == bool(up)
    z = np.load(os.path.join(golden_dir, "vaf.npz"))
    np.testing.assert_allclose(U.var_accounted_for(target=z["target"], pred=z["pred"]), z["vaf"], rtol=1e-14)
    z = np.load(os.path.join(golden_dir, "framer.npz"))
    obs = list(z["obs"])
    for k in (1, 2, 3):
        fr = Framer(frame_num=k)
        np.testing.assert_array_equal(np.array(fr.full(obs)), z["full_%d" % k])
        np.testing.assert_array_equal(fr.last(obs), z["last_%d" % k])
        np.testing.assert_array_equal(fr.last(obs[:3]), z["last3_%d" % k])
    a, b2 = U.make_np([1, 2], [[3.0]])
    assert a.tolist() == [1, 2] and b2.shape == (1, 1)


def test_cli_flag_sets_match_reference():
    from actor_critic_algs_on_tensorflow_b200 import run_AC, train
    a = train.build_parser().parse_args(["worker", "2"])
    assert (a.env, a.seed, a.worker_num, a.ps_num, a.initport, a.stdout_freq, a.save_every, a.frames, a.mode,
            a.desired_kl) == ("Pendulum-v0", 12321, 4, 2, 2849, 20, 600, 1, "train", 0.002)
    assert a.outdir == os.path.join("tmp", "logs") and a.checkpoint_dir == os.path.join("tmp", "checkpoints")
    a = train.build_parser().parse_args(["ps", "0", "--env", "CartPole-v0", "--desired_kl", "0.01", "--save_every", "5",
                                         "--checkpoint_dir", "/x", "--animate"])
    assert (a.job, a.env, a.desired_kl, a.save_every, a.checkpoint_dir, a.animate) == ("ps", "CartPole-v0", 0.01, 5, "/x", True)
    b = run_AC.build_parser().parse_args([])
    assert (b.env, b.seed, b.save_every, b.outdir, b.frames, b.mode, b.desired_kl) == (
        "Pendulum-v0", 12321, 600, "log.txt", 1, "train", 0.002)
    with pytest.raises(SystemExit):
        train.build_parser().parse_args(["boss", "0"])


def _gloo_worker(rank, world, port, q):
    """The cross-rank part of trainer.A2CTrainer.iteration -- the trainer's own exchange functions (sync_exchange,
    async_push_pull, the GRAD_TAIL slot that carries the previous update's KL) -- on host tensors over gloo; the
    oracle's adam_step stands in for the acx_adam_apply launch between them."""
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import policies as P
    from actor_critic_algs_on_tensorflow_b200 import trainer as TR
    n = 64
    rng = np.random.default_rng(100 + rank)
    theta0 = np.random.default_rng(7).normal(size=n).astype(np.float32)     # replicated parameters
    g = rng.normal(size=n).astype(np.float32) * 0.3                          # this rank's local gradient
    zeros = np.zeros(n, np.float32)
    # --- sync: ONE all-reduce of [gradient | previous update's local KL | pad], x 1/world, clip, identical Adam step
    buf = torch.zeros(n + TR.GRAD_TAIL)
    buf[:n] = torch.from_numpy(g)
    buf[n] = 0.001 * (rank + 1)                                              # this rank's post-update KL, one update ago
    TR.sync_exchange(buf)
    gm = np.clip(buf[:n].numpy() / world, -0.1, 0.1)
    kl_avg = float(buf[n].item() / world)
    th_sync, _, _ = P.adam_step(theta0, gm, zeros, zeros, 1, P.kl_lr_rule(0.005, kl_avg, 0.002))
    # --- async: own clip + own moments on the stale parameters, then the trainer's push / pull
    th_loc, _, _ = P.adam_step(theta0, np.clip(g, -0.1, 0.1), zeros, zeros, 1, 0.005)
    params, old = torch.from_numpy(th_loc.copy()), torch.from_numpy(theta0.copy())
    TR.async_push_pull(params, old)
    q.put((rank, th_sync, params.numpy().copy(), kl_avg, g))
    dist.destroy_process_group()


def test_world_size_2_update_logic_gloo():
    import torch.multiprocessing as mp
    from oracle import policies as P
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + os.getpid() % 300
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = sorted([q.get(timeout=120) for _ in range(2)], key=lambda x: x[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (_, s0, a0, kl0, g0), (_, s1, a1, kl1, g1) = out
    np.testing.assert_array_equal(s0, s1)                 # replicas stay identical in both modes
    np.testing.assert_array_equal(a0, a1)
    assert kl0 == kl1 == pytest.approx(0.0015)
    theta0 = np.random.default_rng(7).normal(size=64).astype(np.float32)
    zeros = np.zeros(64, np.float32)
    # the averaged KL 0.0015 lies inside [d/4, 4d] of desired_kl = 0.002: the rule leaves lr at 0.005
    want, _, _ = P.adam_step(theta0, np.clip((g0 + g1) / 2, -0.1, 0.1), zeros, zeros, 1, 0.005)
    np.testing.assert_allclose(s0, want, rtol=1e-6, atol=1e-7)
    # async: theta0 + the sum of both workers' own Adam deltas (every push lands, then the pull)
    d = [P.adam_step(theta0, np.clip(g, -0.1, 0.1), zeros, zeros, 1, 0.005)[0] - theta0 for g in (g0, g1)]
    np.testing.assert_allclose(a0, theta0 + d[0] + d[1], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(a0 - theta0, -0.005 * (np.sign(g0) + np.sign(g1)), rtol=1e-3, atol=1e-6)


def test_bench_workload_flop_constants_match_the_network_shapes():
    """bench.py quotes algorithmic FLOPs per sample (2 per MAC) for its workloads: they must follow from the layer shapes
    of A3C/policies.py:44-50,63,146-150 at each env's observation / action sizes."""
    import importlib.util
    import os
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(os.path.dirname(__file__), "..", "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    shapes = {"pendulum": (3, 1), "cartpole": (4, 2)}
    for name, (ob, nout) in shapes.items():
        w = bench.WORKLOADS[name]
        assert w["flop_actor"] == 2 * (ob * 128 + 128 * 128 + 128 * 64 + 64 * nout)
        assert w["flop_critic"] == 2 * (ob * 256 + 256 * 128 + 128 * 128 + 128 * 1)
        assert w["horizon"] == 20
    assert bench.WORKLOADS["pendulum"]["n_envs"] == 65536 and bench.WORKLOADS["cartpole"]["n_envs"] == 262144
    assert (bench.FLOP_ACTOR_FWD, bench.FLOP_CRITIC_FWD) == (50048.0, 100096.0)
