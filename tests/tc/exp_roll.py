"""Dev: time acx_rollout_tc alone (n envs x 20 steps, L2 flushed) for a given build: python tests/tc/exp_roll.py <lib.so> [n]"""
import os, sys
sys.path.insert(0, os.getcwd())
from actor_critic_algs_on_tensorflow_b200 import _lib as L
L.LIB_PATH = os.path.abspath(sys.argv[1])
import torch
from actor_critic_algs_on_tensorflow_b200.trainer import A2CTrainer
n = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
tr = A2CTrainer("Pendulum-v0", n_envs=n, horizon=20, seed=12321, use_graph=False)
tr.iteration(); torch.cuda.synchronize()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda"); clean = torch.zeros(64 << 20, device="cuda")
s = L.cur_stream()
st3, el3 = tr.state.clone(), tr.elapsed.clone()
def roll():
    L.call("acx_rollout_tc", tr.env, L.ptr(tr.actor), L.ptr(tr.packed_a), tr.ac_scale, tr.T, tr.N, L.ptr(tr.obs_buf), L.ptr(st3),
           L.ptr(el3), None, None, tr.seed, 0, None, tr.max_steps, L.ptr(tr.acs), L.ptr(tr.logps), L.ptr(tr.ent),
           L.ptr(tr.rew), L.ptr(tr.done), L.ptr(tr.ax2), L.ptr(tr.ax3), L.ptr(tr.amask), s)
roll(); torch.cuda.synchronize()
tot = 0.0
for _ in range(10):
    flush.fill_(1); clean.sum()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); roll(); b.record(); torch.cuda.synchronize(); tot += a.elapsed_time(b)
print(sys.argv[1], "n", n, "rollout ms", tot / 10)
