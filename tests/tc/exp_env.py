"""Dev: env-step kernel timings (L2 flushed) for grid-shape experiments: python tests/tc/exp_env.py <lib.so>"""
import os, sys
sys.path.insert(0, os.getcwd())
from actor_critic_algs_on_tensorflow_b200 import _lib as L
L.LIB_PATH = os.path.abspath(sys.argv[1])
import torch
from actor_critic_algs_on_tensorflow_b200.envs import VecEnv
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda"); clean = torch.zeros(64 << 20, device="cuda")
for name, nb in (("Pendulum-v0", 45.0), ("CartPole-v0", 65.0), ("Acrobot-v1", 73.0)):
    for n in (1 << 20, 2 << 20, 8 << 20):
        ve = VecEnv(name, n, seed=1); ve.reset()
        act = torch.zeros(n, device="cuda") if ve.n_actions == 0 else torch.ones(n, device="cuda", dtype=torch.int32)
        ve.step(act); torch.cuda.synchronize()
        tot = 0.0
        for _ in range(10):
            flush.fill_(1); clean.sum()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); ve.step(act); b.record(); torch.cuda.synchronize(); tot += a.elapsed_time(b)
        us = tot / 10 * 1e3
        print("%s n=%dM %.1f us %.0f GB/s" % (name, n >> 20, us, nb * n / us / 1e3))
