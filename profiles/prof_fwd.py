"""Value pass + log-prob pass at the bench size (65,536 x 21 / x 20 samples), three launches each, for ncu captures of
the tensor-core forward kernels (ACX_TC_ONE_TILE=1: the one-tile kernel; default: two tiles in flight)."""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from actor_critic_algs_on_tensorflow_b200 import _lib as L  # noqa: E402
from actor_critic_algs_on_tensorflow_b200.trainer import A2CTrainer  # noqa: E402

tr = A2CTrainer("Pendulum-v0", n_envs=65536, horizon=20, seed=12321, use_graph=False)
tr.iteration()
s = L.cur_stream()
N, T = tr.N, tr.T
for _ in range(3):
    L.call("acx_value_forward_tc", L.ptr(tr.critic), L.ptr(tr.packed_c), tr.ob, T + 1, N, L.ptr(tr.obs_buf), L.ptr(tr.V), 0,
           L.ptr(tr.cx2), L.ptr(tr.cx3), L.ptr(tr.cmask), s)
    L.call("acx_policy_logp_gauss_tc", L.ptr(tr.actor), L.ptr(tr.packed_a), tr.ob, tr.num_ac, tr.ac_scale, T, N,
           L.ptr(tr.obs_buf), L.ptr(tr.acs), None, None, None, L.ptr(tr.logps), L.ptr(tr.kl_part),
           ctypes.byref(tr.kl_blocks), s)
torch.cuda.synchronize()
print("ok")
