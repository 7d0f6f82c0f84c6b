"""Two eager (no CUDA graph) rollout+update iterations at the bench size, for ncu launch lists / captures.
Launch order per iteration: 20 x (acx_mlp_kernel<actor,sample>, acx_env_step_kernel), counter_add,
acx_mlp_kernel<critic,eval> (values), scan, acx_mlp_kernel<actor,train>, grad_reduce, acx_mlp_kernel<critic,train>,
grad_reduce, adam, tick, acx_mlp_kernel<actor,eval> (log-prob/KL), kl_finalize, kl_lr."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from actor_critic_algs_on_tensorflow_b200.trainer import A2CTrainer  # noqa: E402

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 2
tr = A2CTrainer("Pendulum-v0", n_envs=65536, horizon=20, seed=12321, use_graph=False)
for _ in range(iters):
    tr.iteration()
torch.cuda.synchronize()
print("ok", tr.read_stats()["kl_dist"])
