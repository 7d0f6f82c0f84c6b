"""B200-native actor-critic hot path (rollout/update loop of renly/Actor-Critic-Algs-on-Tensorflow).

Modules mirror the reference's: ``policies`` (Actor, Critic), ``process`` (Framer, PathAdv, rollout,
process_fn, test_process), ``util`` (Logger, LinearSchedule, ...), the CLIs ``train`` / ``run_AC`` /
``test_model``; plus ``trainer`` (the fused N-env x T-step loop), ``envs`` (batched envs), ``tf_bundle``
(TF V2 checkpoints) and ``_lib`` (ctypes binding of libacx.so, the C ABI of include/acx.h).
Importing the package does not need a GPU; using it does, and there is no CPU fallback.
"""
__all__ = ["policies", "process", "util", "trainer", "envs", "tf_bundle"]
