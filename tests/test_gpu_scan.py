"""K3 parity: returns/advantage scan through the C ABI vs (a) golden vectors produced by the
reference's own PathAdv (tests/golden/pathadv_cols.npz) and (b) the oracle on fresh inputs.
Tolerance 1e-5 relative (fp32 outputs of an fp64 recurrence)."""
import os

import numpy as np
import pytest
import torch

from oracle import pathadv as PA

pytestmark = pytest.mark.gpu


def scan_gpu(lib, dev, rew, done, V, gamma, L, mode):
    T, N = rew.shape
    r = torch.tensor(rew.astype(np.float32), device=dev)
    d = torch.tensor(done.astype(np.uint8), device=dev)
    v = torch.tensor(V.astype(np.float32), device=dev)
    G = torch.empty((T, N), device=dev)
    A = torch.empty((T, N), device=dev)
    stats = torch.zeros(16, device=dev)
    nbytes = lib.load().acx_returns_adv_workspace_bytes(T, N)
    ws = torch.zeros(nbytes, device=dev, dtype=torch.uint8)
    for _ in range(2):      # twice: the workspace ticket must reset itself
        lib.call("acx_returns_adv_scan", T, N, lib.ptr(r), lib.ptr(d), lib.ptr(v), float(gamma), int(L), lib.ptr(G),
                 lib.ptr(A), lib.ptr(stats), lib.ptr(ws), nbytes, mode, lib.cur_stream())
    torch.cuda.synchronize()
    return G.cpu().numpy(), A.cpu().numpy(), stats.cpu().numpy()


@pytest.mark.parametrize("mode", [1, 2])
def test_scan_matches_reference_golden(lib, dev, golden_dir, mode):
    z = np.load(os.path.join(golden_dir, "pathadv_cols.npz"))
    for c in range(int(z["n_cases"])):
        rew, V, done = z["rew_%d" % c], z["V_%d" % c], z["done_%d" % c]
        # the kernel consumes fp32 rewards: feed the reference's result for the same rounded inputs
        G, A, stats = scan_gpu(lib, dev, rew, done, V, float(z["gamma_%d" % c]), int(z["L_%d" % c]), mode)
        scale = np.abs(z["G_%d" % c]).max()
        np.testing.assert_allclose(G, z["G_%d" % c], rtol=1e-5, atol=2e-6 * scale, err_msg="case %d" % c)
        np.testing.assert_allclose(A, z["adv_%d" % c], rtol=1e-5, atol=4e-6 * scale, err_msg="case %d" % c)
        adv = z["adv_%d" % c]
        np.testing.assert_allclose(stats[0], adv.mean(), rtol=1e-5, atol=1e-5)
        np.testing.assert_allclose(stats[1], 1.0 / (1e-8 + adv.std()), rtol=1e-5)


@pytest.mark.parametrize("T,N,L,pdone,mode", [(20, 65536, 40, 0.01, 1), (20, 1000, 40, 0.0, 1), (200, 33, 40, 0.01, 2),
                                              (1200, 1, 40, 0.005, 2), (1200, 3, 40, 0.005, 1), (50, 257, 7, 0.1, 0),
                                              (1, 5, 40, 0.5, 1), (64, 2, 64, 0.05, 2), (65, 2, 64, 0.05, 2)])
def test_scan_matches_oracle(lib, dev, T, N, L, pdone, mode):
    rng = np.random.default_rng(T * 131 + N)
    rew = (rng.normal(size=(T, N)) * 2).astype(np.float32)
    V = (rng.normal(size=(T + 1, N)) * 5).astype(np.float32)
    done = rng.random((T, N)) < pdone
    G, A, stats = scan_gpu(lib, dev, rew, done, V, 0.98, L, mode)
    Gw, Aw = PA.path_adv_columns(rew, done, V, 0.98, L)
    scale = np.abs(Gw).max()
    np.testing.assert_allclose(G, Gw, rtol=1e-5, atol=2e-6 * scale)
    np.testing.assert_allclose(A, Aw, rtol=1e-5, atol=4e-6 * scale)
    np.testing.assert_allclose(stats[0], Aw.mean(), rtol=1e-4, atol=1e-5)
    if T * N > 1:
        np.testing.assert_allclose(stats[1], 1.0 / (1e-8 + Aw.std()), rtol=1e-5)


def test_scan_linearity_at_full_size(lib, dev):
    """Size-independent property at config-3 size: G is linear in (rew, V) for fixed done masks."""
    T, N = 20, 65536
    rng = np.random.default_rng(0)
    done = rng.random((T, N)) < 0.02
    r1, r2 = rng.normal(size=(2, T, N)).astype(np.float32)
    v1, v2 = rng.normal(size=(2, T + 1, N)).astype(np.float32)
    g1, _, _ = scan_gpu(lib, dev, r1, done, v1, 0.98, 40, 1)
    g2, _, _ = scan_gpu(lib, dev, r2, done, v2, 0.98, 40, 1)
    g12, _, _ = scan_gpu(lib, dev, r1 + 2 * r2, done, v1 + 2 * v2, 0.98, 40, 1)
    np.testing.assert_allclose(g12, g1 + 2 * g2, rtol=1e-4, atol=1e-4)


def test_scan_rejects_bad_arguments(lib, dev):
    t = torch.zeros(64, device=dev)
    f = lib.load().acx_returns_adv_scan
    assert f(4, 4, t.data_ptr(), t.data_ptr(), t.data_ptr(), 0.98, 0, t.data_ptr(), t.data_ptr(), t.data_ptr(),
             t.data_ptr(), 1 << 20, 0, None) == -1                       # look_ahead < 1
    assert f(4, 4, t.data_ptr(), t.data_ptr(), t.data_ptr(), 0.98, 40, t.data_ptr(), t.data_ptr(), t.data_ptr(),
             t.data_ptr(), 8, 0, None) == -1                             # workspace too small
    assert b"workspace" in lib.load().acx_last_error()
    assert f(0, 4, None, None, None, 0.98, 40, None, None, None, None, 0, 0, None) == 0     # empty: no-op
