"""CPU oracle for the actor-critic hot path -- TEST INFRASTRUCTURE, NOT PRODUCT.

This package restates, in numpy, the arithmetic of the reference's rollout /
update loop (``/root/reference/A3C/{process,policies,util}.py`` and the
``Basic_AC`` twins) plus the third-party arithmetic the reference calls
(gym classic-control dynamics, TF 1.3 ``Normal`` / ``multinomial`` /
``AdamOptimizer``).  Each function cites the reference ``file:line`` it follows.

Who may import it: ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs -- as the *checker* or the timed CPU
baseline, never as the thing shipped.  The product package
``actor_critic_algs_on_tensorflow_b200`` never imports it and has no CPU
fallback: it fails loudly when the CUDA library is missing.

Parity status
-------------
* ``PathAdv``, ``Framer``, ``LinearSchedule``, ``var_accounted_for``, ``Logger``:
  PINNED -- checked against the reference's own classes, imported from
  ``/root/reference`` under ``tensorflow``/``gym`` stub modules
  (``tests/golden/make_golden.py`` generated ``tests/golden/*.npz`` from them).
* TF V2 checkpoint bundle: PINNED on the shipped ``demos/model-Pendulum_a3c.*``
  (every CRC verified).
* Network forward / loss / gradients / Adam, TF ``multinomial`` and the gym
  dynamics: **parity unpinned** -- TensorFlow 1.3 and gym are third-party,
  un-vendored, not installable here (no network, Python 3.12), and the
  reference has no tests or golden vectors.  These are restated from the
  reference's call sites (policies.py) and the published algorithms; analytic
  gradients are cross-checked against ``torch.autograd``.
"""
