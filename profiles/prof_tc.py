"""One eager rollout+update iteration at the bench size (tensor-core forward engine on), for ncu captures."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from actor_critic_algs_on_tensorflow_b200.trainer import A2CTrainer  # noqa: E402

tr = A2CTrainer("Pendulum-v0", n_envs=65536, horizon=20, seed=12321, use_graph=False, tensor_cores=True)
for _ in range(2):
    tr.iteration()
torch.cuda.synchronize()
print("ok", tr.read_stats()["kl_dist"])
