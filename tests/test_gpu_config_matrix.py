"""Every configuration switch of the fused trainer runs two iterations at ragged sizes (env counts that are not a
multiple of the 128-sample tile or of 4) and leaves finite statistics: all three environments, both engines (tcgen05 and
FFMA), frame stacking (A3C/process.py:25-43), the Basic_AC critic (Basic_AC/policies.py:134-135), per-step launches,
no activation cache, diagnostics.  The oracle comparisons of these paths are in test_gpu_trainer.py; this is the cheap
sweep that catches a switch combination nobody compared."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

CASES = [
    dict(env_id="Pendulum-v0", n_envs=300, horizon=5),
    dict(env_id="Pendulum-v0", n_envs=257, horizon=4),                                   # N % 4 != 0: FFMA weight gradients
    dict(env_id="CartPole-v0", n_envs=384, horizon=6, diagnostics=True),
    dict(env_id="Acrobot-v1", n_envs=200, horizon=5),
    dict(env_id="Pendulum-v0", n_envs=260, horizon=5, tensor_cores=False),
    dict(env_id="CartPole-v0", n_envs=130, horizon=5, frames=3),                         # 12 inputs: FFMA engine
    dict(env_id="Acrobot-v1", n_envs=130, horizon=4, frames=3),                          # 18 inputs
    dict(env_id="CartPole-v0", n_envs=256, horizon=8, basic=True, clip=1.0, max_lr=1.0),
    dict(env_id="Pendulum-v0", n_envs=256, horizon=5, fused_rollout=False),
    dict(env_id="Pendulum-v0", n_envs=256, horizon=5, cache_activations=False),
    dict(env_id="Pendulum-v0", n_envs=256, horizon=5, stagger=False, overlap_nets=False),
]


@pytest.mark.parametrize("kw", CASES, ids=lambda kw: "-".join("%s" % v for v in kw.values()))
@pytest.mark.parametrize("graph", [False, True])
def test_two_iterations_finite(kw, graph):
    from actor_critic_algs_on_tensorflow_b200.trainer import A2CTrainer
    tr = A2CTrainer(seed=7, use_graph=graph, **kw)
    for _ in range(4 if graph else 2):           # graphs: eager, eager warm-up of the second segment, capture, replay
        tr.iteration()
    torch.cuda.synchronize()
    st = tr.read_stats()
    for k in ("act_loss", "crit_loss", "kl_dist", "avg_ent", "adv_std", "ev_before"):
        assert math.isfinite(st[k]), (k, st)
    assert st["iteration"] == (4 if graph else 2)
    assert torch.isfinite(tr.params).all()
