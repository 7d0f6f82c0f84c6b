"""Dev: A/B timing of two builds of libacx.so on the bench workload (python tests/tc/exp_bench.py <lib.so>): ms per
rollout+update step, CUDA graphs, L2 flushed between steps."""
import sys, os, json
sys.path.insert(0, os.getcwd())
from actor_critic_algs_on_tensorflow_b200 import _lib as L
L.LIB_PATH = os.path.abspath(sys.argv[1])
import torch
from actor_critic_algs_on_tensorflow_b200.trainer import A2CTrainer
tr = A2CTrainer("Pendulum-v0", n_envs=65536, horizon=20, seed=12321)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda"); clean = torch.zeros(64 << 20, device="cuda")
for _ in range(4): tr.iteration()
torch.cuda.synchronize()
tot = 0.0
for _ in range(20):
    flush.fill_(1); clean.sum()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); tr.iteration(); b.record(); torch.cuda.synchronize(); tot += a.elapsed_time(b)
print(sys.argv[1], "ms/step", tot / 20)
