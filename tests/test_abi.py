"""The C-ABI library loads without a GPU and exports every symbol include/acx.h declares; no compute calls."""
import ctypes
import os
import re

import pytest

from actor_critic_algs_on_tensorflow_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared():
    hdr = open(os.path.join(ROOT, "include", "acx.h")).read()
    return sorted(set(re.findall(r"\b(acx_[a-z0-9_]+)\s*\(", hdr)))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    names = declared()
    assert len(names) >= 26
    for n in names:
        assert getattr(lib, n) is not None, n
    assert sorted(_lib.SIGNATURES) == names               # the ctypes binding covers the header, nothing more


def test_host_side_queries_work_without_a_gpu():
    lib = _lib.load()
    assert lib.acx_abi_version() == 1
    lay, total = _lib.params_layout(0, 3, 1, True)
    assert total == 25352 and lay["W2"] == (512, 16384) and lay["log_std"] == (25348, 1)
    lay, total = _lib.params_layout(1, 3)
    assert total == 50564 and lay["Wh"] == (50432, 128)
    lay, total = _lib.params_layout(0, 4, 2, False)
    assert "log_std" not in lay and lay["Wh"][1] == 128
    s, o, a, m = (ctypes.c_int() for _ in range(4))
    for env, want in ((0, (2, 3, 0, 200)), (1, (4, 4, 2, 200)), (2, (4, 6, 3, 500))):
        assert lib.acx_env_dims(env, ctypes.byref(s), ctypes.byref(o), ctypes.byref(a), ctypes.byref(m)) == 0
        assert (s.value, o.value, a.value, m.value) == want
    assert lib.acx_env_dims(7, None, None, None, None) == -1 and b"unknown env" in lib.acx_last_error()
    assert lib.acx_params_layout(2, 3, 1, 1, None, None) == -1
    with pytest.raises(_lib.AcxError, match="ob_dim"):
        _lib.params_layout(0, 99, 1, True)


def test_layout_matches_the_oracle_layout():
    from oracle import policies as P
    for ob, nac, cont in ((3, 1, True), (4, 2, False), (6, 3, False), (3, 2, True)):
        lay, total = _lib.params_layout(0, ob, nac, cont)
        olay = P.actor_layout(ob, nac, cont)
        assert total == olay["_size"]
        for k, o, s in P.items(olay):
            assert lay[k] == (o, int(__import__("numpy").prod(s)))
        lay, total = _lib.params_layout(1, ob)
        olay = P.critic_layout(ob)
        assert total == olay["_size"]
        assert lay["Wh"][0] == olay["Wv"][0] and lay["bh"][0] == olay["bv"][0]


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from actor_critic_algs_on_tensorflow_b200 import envs, policies, trainer
    with pytest.raises(_lib.AcxError, match="no CPU fallback"):
        trainer.A2CTrainer("Pendulum-v0", n_envs=8, horizon=2)
    with pytest.raises(_lib.AcxError, match="no CPU fallback"):
        policies.Actor(name="global_actor", num_ob_feat=3, num_ac=1)
    with pytest.raises(_lib.AcxError, match="no CPU fallback"):
        envs.make("Pendulum-v0")


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "actor_critic_algs_on_tensorflow_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), fn
