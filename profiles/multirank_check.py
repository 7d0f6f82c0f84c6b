#!/usr/bin/env python
"""Multi-rank semantics on hardware (one process per GPU), both modes of trainer.A2CTrainer; the sync checks run twice, with the
fused peer-memory exchange (acx_allreduce_adam) and with the NCCL all-reduce:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 \
        profiles/multirank_check.py > profiles/r02_multirank_check.txt

  1. sync: after 10 iterations parameters, Adam moments and the opt block are bit-identical on every rank;
  2. sync: the all-reduced gradient of the first update = sum of the gradients single-rank trainers compute on the
     same shards (fp32 sum order is the only difference), and the update equals oracle.policies.adam_step on its
     1/world mean with the actor clip applied after the reduction;
  3. async: parameters after one update = theta0 + sum over ranks of the Adam deltas oracle.policies.adam_step gives
     for each rank's own clipped gradient (every push lands on the shared parameters, then the pull --
     A3C/policies.py:85-98, A3C/process.py:257-260);
  4. async: replicas bit-identical after 10 iterations; per-rank lr may differ (each worker keeps its own).
The oracle is used as the checker only.  Prints one line per check and exits non-zero on a failure."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from actor_critic_algs_on_tensorflow_b200.trainer import A2CTrainer  # noqa: E402
from oracle import policies as P  # noqa: E402


def gather(t):
    out = [torch.empty_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t.contiguous())
    return [o.cpu().numpy() for o in out]


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ok = True

    def report(name, passed, detail=""):
        nonlocal ok
        ok = ok and bool(passed)
        if rank == 0:
            print("%-78s %s %s" % (name, "PASS" if passed else "FAIL", detail), flush=True)

    N, T, seed = 4096, 20, 12321
    for env_id in ("Pendulum-v0", "CartPole-v0"):
        kw = dict(n_envs=N, horizon=T, seed=seed, max_episode_steps=50)
        for exch in ("auto", "nccl"):      # 'auto' = the fused peer-memory kernel where the buffers can be mapped
            # ---- 2: first update, sync
            multi = A2CTrainer(env_id, rank=rank, world=world, mode="sync", exchange=exch, **kw)
            tag = "%s sync[%s]" % (env_id, multi.exchange)
            single = A2CTrainer(env_id, rank=rank, world=1, **kw)               # same shard (index0 = rank * N), no exchange
            single.index0 = rank * N
            theta0 = multi.params.cpu().numpy().copy()
            P_ = multi.Pa + multi.Pc
            # single-rank gradient of this shard: rollout + gradients only
            single._rollout(False); single._episode_stats(); single._gradients()
            g_local = single.grads.clone()
            multi.iteration()
            torch.cuda.synchronize()
            gs = gather(g_local)
            g_sum = np.sum(np.stack(gs).astype(np.float64), axis=0)
            got = (multi.grad_sum if multi.exchange == "p2p" else multi._gradbuf)[:P_].cpu().numpy().astype(np.float64)
            err = np.abs(got - g_sum).max() / np.abs(g_sum).max()
            report("%s: summed gradient = sum of the %d single-rank gradients" % (tag, world), err <= 2e-6,
                   "(%.1e of max)" % err)
            gm = (g_sum / world).astype(np.float32)
            gm[:multi.Pa] = np.clip(gm[:multi.Pa], -0.1, 0.1)
            z = np.zeros(P_, np.float32)
            want = np.concatenate([P.adam_step(theta0[:multi.Pa], gm[:multi.Pa], z[:multi.Pa], z[:multi.Pa], 1, 0.005)[0],
                                   P.adam_step(theta0[multi.Pa:], gm[multi.Pa:], z[multi.Pa:], z[multi.Pa:], 1, 0.001)[0]])
            got = multi.params.cpu().numpy()
            # first Adam step = lr * g / (|g| + eps'): elements whose averaged gradient is rounding noise may flip sign
            frac = float(np.mean(np.abs(got - want) > 1e-6))
            report("%s: update = oracle adam_step(mean gradient, clip after the reduction)" % tag, frac < 2e-3,
                   "(%.2e of the weights differ by > 1e-6)" % frac)
            # ---- 1: replicas identical after 10 iterations
            for _ in range(9):
                multi.iteration()
            torch.cuda.synchronize()
            same = all(all(np.array_equal(x, xs[0]) for x in xs) for xs in
                       (gather(multi.params), gather(multi.m), gather(multi.v), gather(multi.opt)))
            st = multi.read_stats()
            report("%s: params / Adam moments / opt block bit-identical on %d ranks after 10 iterations" % (tag, world),
                   same, "(iteration %d, lr %.5f, kl %.2e)" % (st["iteration"], st["next_lr"], st["kl_dist"]))
            del multi, single
        # ---- 3: first update, async
        am = A2CTrainer(env_id, rank=rank, world=world, mode="async", **kw)
        am.iteration()
        torch.cuda.synchronize()
        deltas = []
        for g in gs:                                   # each rank's own gradient (same shards, same weights as above)
            g = g.copy()
            g[:am.Pa] = np.clip(g[:am.Pa], -0.1, 0.1)
            th = np.concatenate([P.adam_step(theta0[:am.Pa], g[:am.Pa], z[:am.Pa], z[:am.Pa], 1, 0.005)[0],
                                 P.adam_step(theta0[am.Pa:], g[am.Pa:], z[am.Pa:], z[am.Pa:], 1, 0.001)[0]])
            deltas.append(th.astype(np.float64) - theta0)
        want = theta0 + np.sum(deltas, axis=0)
        got = am.params.cpu().numpy()
        frac = float(np.mean(np.abs(got - want) > 1e-6))
        report("%s async: params = theta0 + sum of per-rank oracle Adam deltas (push, push, pull)" % env_id, frac < 2e-3,
               "(%.2e of the weights differ by > 1e-6)" % frac)
        for _ in range(9):
            am.iteration()
        torch.cuda.synchronize()
        same = all(np.array_equal(x, xs[0]) for xs in (gather(am.params),) for x in xs)
        lrs = [float(o[0]) for o in gather(am.opt)]
        report("%s async: shared parameters bit-identical on %d ranks after 10 iterations" % (env_id, world), same,
               "(per-rank lr %s)" % ", ".join("%.5f" % x for x in lrs))
        del am
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
